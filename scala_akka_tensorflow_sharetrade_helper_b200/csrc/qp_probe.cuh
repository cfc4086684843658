// qp_probe.cuh — a one-CTA test bench for single tcgen05.mma instructions (kind::tf32).
//
// The persistent kernels rest on operand layouts that no library builds for them (a price-window buffer whose rows,
// from the row of window element 0 on, are the K-major A operand of the backward contraction; N = 104 per CTA; an
// M = 64 accumulator read back lane by lane).  This kernel lets a test state such a layout as a plain byte image of
// shared memory plus the two 64-bit matrix descriptors (or A as a block of tensor-memory columns), issues the MMAs,
// and hands back all 128 lanes of the accumulator, so that every layout fact a kernel relies on is pinned by a test
// against a numpy product (tests/test_gpu_umma_layouts.py; the explorations that led there: profiles/umma_*.py).
#pragma once
#include "qp_tc.cuh"

namespace qp {

struct UmmaProbeArgs {
    const uint8_t* a_img; uint32_t a_bytes;    // byte images copied to shared memory: A at offset 0 ...
    const uint8_t* b_img; uint32_t b_bytes;    // ... B at the next 1024-byte boundary
    uint64_t desc_a, desc_b;                   // matrix descriptors whose start-address field is relative to the image
    uint32_t idesc;
    int n_mma;                                 // MMAs issued (the first one overwrites, the others accumulate)
    uint32_t a_step, b_step;                   // added to the start-address fields per MMA (16-byte units)
    uint32_t d_col, d_lane;                    // TMEM column / lane offset of the accumulator
    int n_cols;                                // columns read back (multiple of 16, <= 512)
    int a_tmem;                                // 1: A comes from tensor memory: a_img is [128 rows][a_k] fp32 row-major, written to
    int a_k;                                   //    columns [a_col, a_col + a_k) of lane = row; a_step then counts COLUMNS per MMA
    uint32_t a_col;
    float fill;                                // every TMEM word read back is preset to this
    float* out;                                // [128][n_cols]
};

#ifdef __CUDACC__

__global__ void __launch_bounds__(128, 1) umma_probe_kernel(const UmmaProbeArgs g) {
    extern __shared__ __align__(1024) uint8_t probe_smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    uint8_t* a_s = probe_smem;
    uint8_t* b_s = probe_smem + (((size_t)g.a_bytes + 1023) & ~(size_t)1023);
    for (uint32_t i = tid; i < g.a_bytes / 4; i += 128) reinterpret_cast<uint32_t*>(a_s)[i] = reinterpret_cast<const uint32_t*>(g.a_img)[i];
    for (uint32_t i = tid; i < g.b_bytes / 4; i += 128) reinterpret_cast<uint32_t*>(b_s)[i] = reinterpret_cast<const uint32_t*>(g.b_img)[i];
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    // preset the words that will be read back
    for (int c0 = 0; c0 < g.n_cols; c0 += 16) {
        const uint32_t v = __float_as_uint(g.fill);
        asm volatile(
            "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};\n" ::"r"(
                tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0), "r"(v) : "memory");
    }
    if (g.a_tmem) {
        const float* arow = reinterpret_cast<const float*>(g.a_img) + (size_t)(warp * 32 + lane) * g.a_k;
        for (int c0 = 0; c0 < g.a_k; c0 += 8) {
            uint32_t v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __float_as_uint(arow[c0 + j]);
            asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};\n" ::"r"(
                             tmem_base + ((uint32_t)(warp * 32) << 16) + g.a_col + (uint32_t)c0),
                         "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
        }
    }
    asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (tid == 0 && g.a_tmem) {
        const uint64_t db0 = g.desc_b + (uint64_t)(((uint32_t)__cvta_generic_to_shared(b_s) >> 4) & 0x3FFF);
        for (int i = 0; i < g.n_mma; ++i) {
            const uint32_t acc = i > 0 ? 1u : 0u;
            asm volatile(
                "{\n\t"
                ".reg .pred p;\n\t"
                "setp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t"
                "}\n" ::"r"(tmem_base + (g.d_lane << 16) + g.d_col),
                "r"(tmem_base + g.a_col + (uint32_t)i * g.a_step), "l"(db0 + (uint64_t)i * g.b_step), "r"(g.idesc), "r"(acc), "r"(0u)
                : "memory");
        }
        tc_commit(&bar);
    } else if (tid == 0) {
        const uint64_t da0 = g.desc_a + (uint64_t)(((uint32_t)__cvta_generic_to_shared(a_s) >> 4) & 0x3FFF);
        const uint64_t db0 = g.desc_b + (uint64_t)(((uint32_t)__cvta_generic_to_shared(b_s) >> 4) & 0x3FFF);
        for (int i = 0; i < g.n_mma; ++i)
            tc_mma_tf32(tmem_base + (g.d_lane << 16) + g.d_col, da0 + (uint64_t)i * g.a_step, db0 + (uint64_t)i * g.b_step, g.idesc, i > 0 ? 1u : 0u);
        tc_commit(&bar);
    }
    tc_wait(&bar, 0);
    tc_fence_after();
    for (int c0 = 0; c0 < g.n_cols; c0 += 16) {
        float v[16];
        uint32_t r[16];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
              "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
            : "r"(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0));
        asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
#pragma unroll
        for (int j = 0; j < 16; ++j) g.out[(size_t)(warp * 32 + lane) * g.n_cols + c0 + j] = v[j];
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(512u) : "memory");
}

#endif  // __CUDACC__
}  // namespace qp
