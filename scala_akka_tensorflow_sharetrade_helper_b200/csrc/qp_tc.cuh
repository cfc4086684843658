// qp_tc.cuh — tcgen05 (5th-generation tensor core) GEMM for the GEMM-organised path.
//
//   C[M, N] (+split) = A[M, K] · B[N, K]^T,   A and B K-major fp32 in HBM, accumulated in fp32 in TMEM
//
// Precision: kind::tf32 reads the top 19 bits of each fp32 operand.  To stay inside the fp32
// tolerance of the parity tests every operand is kept as two arrays, hi = tf32(x) and lo = x - hi,
// and each k-block issues three MMAs:  A_hi·B_hi + A_lo·B_hi + A_hi·B_lo  ("3xTF32"; the dropped
// lo·lo term is ~2^-22 relative).  passes = 1 runs plain TF32 (own, looser tolerance).
//
// Structure (one 128 x 128 output tile per CTA, 256 threads):
//   warp 0, one lane : TMA producer — cp.async.bulk.tensor.2d of the four 128 x 32 operand tiles of a
//                      k-block (SWIZZLE_128B) into a 3-stage shared-memory ring, mbarrier complete_tx
//   warp 1, one lane : MMA issuer — tcgen05.mma.cta_group::1.kind::tf32, M = N = 128, K = 8 per
//                      instruction, shared-memory descriptors advanced 32 B per k-substep;
//                      tcgen05.commit releases the stage and, at the end, signals the epilogue
//   warp 2           : tcgen05.alloc / dealloc of 128 TMEM columns
//   warps 4..7       : epilogue — tcgen05.ld 32x32b of their 32 TMEM lanes, fp32 stores to C
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "qp_kernels.cuh"

namespace qp {

constexpr int kTcBM = 128, kTcBN = 128, kTcBK = 32, kTcStages = 3;
constexpr int kTcTileBytes = kTcBM * kTcBK * 4;                 // 16 KB: one operand tile of one k-block
constexpr int kTcStageBytes = 4 * kTcTileBytes;                 // A_hi, A_lo, B_hi, B_lo
constexpr size_t kTcSmemBytes = (size_t)kTcStages * kTcStageBytes + 1024 /*align*/ + 256 /*barriers*/;

struct TcGemmArgs {
    int a_k0, a_row0;       // TMA coordinates of A at k-block 0 / tile row 0: (inner = k, outer = row)
    int b_k0, b_row0;
    int nkb;                // k-blocks of 32 per split (blockIdx.z)
    int passes;             // 3 = 3xTF32, 1 = TF32
    int a_lo_zero;          // 1: every A_lo element is exactly 0 (A is tf32-exact): its tile and MMA are skipped
    float* C; int64_t ldc; int64_t c_split_stride;
    int m_valid, n_valid;   // rows / columns of C that exist
};

#ifdef __CUDACC__

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* tmap, int c_inner, int c_outer, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n" ::"r"(
            (uint32_t)__cvta_generic_to_shared(smem_dst)),
        "l"(tmap), "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(c_inner), "r"(c_outer)
        : "memory");
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start >> 4,
// LBO = 1 (unused for swizzled K-major), SBO = 1024 B (8 rows x 128 B), version 1, layout type 2.
__device__ __forceinline__ uint64_t tc_smem_desc(const void* smem_tile) {
    const uint32_t addr = (uint32_t)__cvta_generic_to_shared(smem_tile);
    return (uint64_t)((addr >> 4) & 0x3FFF) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

// Instruction descriptor (cute::UMMA::InstrDescriptor): fp32 accumulate, TF32 x TF32, both K-major.
__device__ __forceinline__ uint32_t tc_instr_desc(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_c, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t"
        "}\n" ::"r"(tmem_c),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}

__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(bar))
                 : "memory");
}

// bounded wait: a missing arrival traps after ~4 s instead of hanging the GPU
__device__ __forceinline__ void tc_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = (uint32_t)__cvta_generic_to_shared(bar);
    unsigned long long t0 = 0;
    for (uint32_t spins = 0;; ++spins) {
        uint32_t done;
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) return;
        if ((spins & 1023u) == 1023u) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t0 == 0) t0 = t;
            else if (t - t0 > 4000000000ull) __trap();
        }
    }
}

__global__ void __launch_bounds__(256, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tA_hi, const __grid_constant__ CUtensorMap tA_lo,
               const __grid_constant__ CUtensorMap tB_hi, const __grid_constant__ CUtensorMap tB_lo,
               const TcGemmArgs g) {
    extern __shared__ uint8_t tc_smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(tc_smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full = reinterpret_cast<uint64_t*>(base + (size_t)kTcStages * kTcStageBytes);
    uint64_t* empty = full + kTcStages;
    uint64_t* tmem_full = empty + kTcStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

    if (threadIdx.x == 0) {
        for (int s = 0; s < kTcStages; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
        mbar_init(tmem_full, 1);
        mbar_fence_init();
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                         (uint32_t)__cvta_generic_to_shared(tmem_slot)),
                     "r"((uint32_t)kTcBN)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // N tiles vary fastest in launch order: the CTAs that share an A tile run back to back, so A is
    // fetched from HBM once and the (small) B matrix stays L2-resident
    const int m_tile = blockIdx.y, n_tile = blockIdx.x;
    const int kbase = blockIdx.z * g.nkb * kTcBK;
    const int a_row = g.a_row0 + m_tile * kTcBM;
    const int b_row = g.b_row0 + n_tile * kTcBN;
    const bool three = g.passes == 3;
    const bool use_a_lo = three && !g.a_lo_zero;

    if (warp == 0 && lane == 0) {
        // ---------------- TMA producer ----------------
        for (int kb = 0; kb < g.nkb; ++kb) {
            const int s = kb % kTcStages;
            const uint32_t ph = (uint32_t)(kb / kTcStages) & 1u;
            tc_wait(empty + s, ph ^ 1u);
            uint8_t* st = base + (size_t)s * kTcStageBytes;
            mbar_arrive_expect_tx(full + s, (uint32_t)(three ? (use_a_lo ? 4 : 3) : 2) * kTcTileBytes);
            const int ak = g.a_k0 + kbase + kb * kTcBK, bk = g.b_k0 + kbase + kb * kTcBK;
            tma_load_2d(st, &tA_hi, ak, a_row, full + s);
            tma_load_2d(st + 2 * kTcTileBytes, &tB_hi, bk, b_row, full + s);
            if (use_a_lo) tma_load_2d(st + kTcTileBytes, &tA_lo, ak, a_row, full + s);
            if (three) tma_load_2d(st + 3 * kTcTileBytes, &tB_lo, bk, b_row, full + s);
        }
    } else if (warp == 1 && lane == 0) {
        // ---------------- MMA issuer ----------------
        const uint32_t idesc = tc_instr_desc(kTcBM, kTcBN);
        for (int kb = 0; kb < g.nkb; ++kb) {
            const int s = kb % kTcStages;
            const uint32_t ph = (uint32_t)(kb / kTcStages) & 1u;
            tc_wait(full + s, ph);
            tc_fence_after();
            uint8_t* st = base + (size_t)s * kTcStageBytes;
            const uint64_t da_hi = tc_smem_desc(st), da_lo = tc_smem_desc(st + kTcTileBytes);
            const uint64_t db_hi = tc_smem_desc(st + 2 * kTcTileBytes), db_lo = tc_smem_desc(st + 3 * kTcTileBytes);
#pragma unroll
            for (int k4 = 0; k4 < kTcBK / 8; ++k4) {
                const uint64_t adv = (uint64_t)(k4 * 2);          // 8 tf32 = 32 B = 2 x 16 B along K
                tc_mma_tf32(tmem_base, da_hi + adv, db_hi + adv, idesc, (kb > 0 || k4 > 0) ? 1u : 0u);
                if (use_a_lo) tc_mma_tf32(tmem_base, da_lo + adv, db_hi + adv, idesc, 1u);
                if (three) tc_mma_tf32(tmem_base, da_hi + adv, db_lo + adv, idesc, 1u);
            }
            tc_commit(empty + s);        // the stage is free once these MMAs have read it
        }
        tc_commit(tmem_full);            // accumulator complete
    } else if (warp >= 4) {
        // ---------------- epilogue ----------------
        tc_wait(tmem_full, 0);
        tc_fence_after();
        const int q = warp & 3;                                     // TMEM lane quarter this warp may read
        const int64_t m = (int64_t)m_tile * kTcBM + q * 32 + lane;
        float* crow = g.C + (size_t)blockIdx.z * g.c_split_stride + m * g.ldc + (int64_t)n_tile * kTcBN;
#pragma unroll 1
        for (int c0 = 0; c0 < kTcBN; c0 += 16) {
            uint32_t r[16];
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
            if (m < g.m_valid) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int n = n_tile * kTcBN + c0 + 4 * j;
                    if (n < g.n_valid)
                        *reinterpret_cast<float4*>(crow + c0 + 4 * j) =
                            make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                        __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"((uint32_t)kTcBN) : "memory");
    }
}

// hi = tf32(x) (round to nearest), lo = x - hi (exact in fp32).  With row_len > 0 the copy is shifted
// left by `shift` elements inside each row (dst[r][x] = src[r][x + shift], zero past the end): TMA
// needs the first byte of a box to be 16-byte aligned, so the sliding price window is read from the
// copy whose shift makes its start coordinate a multiple of four elements.
__global__ void tc_split_kernel(const float* __restrict__ src, float* __restrict__ hi, float* __restrict__ lo, int64_t n,
                                int64_t row_len = 0, int shift = 0) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float x;
    if (row_len > 0) {
        const int64_t col = i % row_len;
        x = (col + shift < row_len) ? src[i + shift] : 0.0f;
    } else {
        x = src[i];
    }
    uint32_t t;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(t) : "f"(x));
    const float h = __uint_as_float(t);
    hi[i] = h;
    lo[i] = x - h;
}

// flag[0] = 1 if any element is non-zero (used to learn whether a lo array is identically zero)
__global__ void tc_any_nonzero_kernel(const float* __restrict__ x, int64_t n, int* flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && x[i] != 0.0f) atomicOr(flag, 1);
}

// flag[0] = 1 if some element is not exactly representable in tf32
__global__ void tc_inexact_kernel(const float* __restrict__ x, int64_t n, int* flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t t;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(t) : "f"(x[i]));
    if (__uint_as_float(t) != x[i]) atomicOr(flag, 1);
}

// src [R, C] row-major  ->  hiT / loT [C, ldt] (transposed, K-major for the tensor core), zero padding
// of columns R..ldt-1 is the caller's (buffers are zero-initialised once).
__global__ void tc_transpose_split_kernel(const float* __restrict__ src, int64_t R, int64_t C, int64_t ld_src,
                                          float* __restrict__ hiT, float* __restrict__ loT, int64_t ldt) {
    __shared__ float tile[32][33];
    const int64_t c0 = (int64_t)blockIdx.x * 32, r0 = (int64_t)blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int64_t r = r0 + j, c = c0 + threadIdx.x;
        tile[j][threadIdx.x] = (r < R && c < C) ? src[r * ld_src + c] : 0.0f;
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int64_t c = c0 + j, r = r0 + threadIdx.x;
        if (c < C && r < R) {
            const float x = tile[threadIdx.x][j];
            uint32_t t;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(t) : "f"(x));
            const float h = __uint_as_float(t);
            hiT[c * ldt + r] = h;
            loT[c * ldt + r] = x - h;
        }
    }
}

#endif  // __CUDACC__
}  // namespace qp
