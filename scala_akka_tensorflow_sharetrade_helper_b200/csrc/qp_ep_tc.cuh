// qp_ep_tc.cuh — the persistent episode kernel with its two in-CTA contractions on the tcgen05
// tensor cores, for the reference shape (window 201, hidden 200) and its neighbours.  DESIGN.md §4 (i-tc).
//
// Life cycle as episode_kernel (qp_kernels.cuh): one cooperative launch runs every requested env-step;
// CTA c owns episodes [32c, 32c+32) (or, with more tiles than CTAs, tiles c, c+G, ... in waves).
//   * state     the price window of the tile lives in shared memory in the layout of the backward A operand
//               (X^T, K-major SWIZZLE_128B) as a ring of rows: sliding it is writing one row per step
//   * forward   Z[64, Np] = X[64, Kp] · W1[Kp, Np]: rows 0..31 = state s, rows 32..63 = state s'; issued as M = 128
//               MMAs (rows 64..127 read whatever follows in shared memory and land in TMEM lanes nobody reads).
//               A = X is written by the CTA; B = W1^T streams through a 2-stage TMA ring from the tf32 hi/lo
//               copies that phase B of the previous step left in global memory
//   * glue      raw z out of TMEM into shared-memory tiles; Q(s), select, book-keeping, Q(s'), TD target and dZ1
//               on all warps; dZ1 goes straight into the backward B operand as a tf32 hi/lo pair
//   * backward  dW1[256, Np] = X^T[256, 32] · dZ1[32, Np], two M = 128 halves, K = the 32 episodes; the partial
//               leaves TMEM through swizzled boxes and TMA tensor stores (reduce-add from the second wave on)
//   * update    grid barrier; every CTA owns a share of the parameters: the tiles' partial segments arrive as
//               bulk copies and are summed in shared memory; with several GPUs the shares are exchanged over
//               NVLink peer memory ({value, tag} stores, no fence); AdaGrad; tf32 hi/lo of the new W1 published;
//               split barrier (publishers count up, warp 0 of each CTA waits at the top of the next step)
// Precision: prices must be tf32-exact (checked at qp_env_create; the synthetic quarter-tick walk is) so that
// X needs no lo part; W1 and dZ1 are hi/lo pairs -> A·B_hi + A·B_lo, fp32-grade.  The budget/shares columns
// of the state (not tf32-exact) stay on the FFMA pipe, as in episode_kernel.
#pragma once
#include "qp_kernels.cuh"
#include "qp_tc.cuh"

namespace qp {

constexpr int kEtThreads = 384;    // 12 warps; 16 were measured slower (47.3 K vs 45.3 K cycles per step: the per-warp redundant select / book-keeping grows)
constexpr int kEtWarps = kEtThreads / 32;
constexpr int kEtParts = kEtWarps / 4;     // warps per TMEM lane quarter (a warp can only read quarter warp % 4)

struct EpTcArgs {
    EpisodeArgs ea;
    float* w1T_hi; float* w1T_lo;   // [H, Kp] K-major tf32 hi/lo copies of W1[0:wn, :] (rows k >= wn zero)
    int Kp;                          // wn rounded up to 32
    int nkb;                         // Kp / 32
    int Np;                          // hidden rounded up to 16 (MMA N)
    int halves;                      // ceil(wn / 128): M = 128 tiles of the backward contraction
    unsigned long long* published;   // CTAs that have finished phase B, summed over the steps of the launch
    // several GPUs (ea.world > 1): every rank's exchange block, mapped into this process (CUDA IPC)
    unsigned long long* pdata[8];    // [parity][src rank][p_pad] (step tag << 32 | value bits) of the ranks' shares of the gradient
    int64_t p_pad;
    long long* cta_ns;               // tuning aid ("phase_clocks" = 2): [n_steps][gridDim][4] globaltimer stamps of every CTA
};

#ifdef __CUDACC__

// byte offset of element (row, col) of a K-major SWIZZLE_128B tile with 128-byte rows (32 fp32)
__device__ __forceinline__ uint32_t et_swz(int row, int col) {
    return (uint32_t)row * 128u + (((uint32_t)(col >> 2) ^ ((uint32_t)row & 7u)) << 4) + ((uint32_t)(col & 3) << 2);
}
__device__ __forceinline__ float et_tf32(float x) {
    uint32_t t;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(t) : "f"(x));
    return __uint_as_float(t);
}
// h = relu(z_window + budget*W1[wn] + shares*W1[wn+1] + 0.1): one definition, so that every phase
// that re-forms h from the stored window part agrees bit for bit
__device__ __forceinline__ float et_hidden(float zwin, float xb, float xs, float wb, float ws, float bias) {
    const float z = fmaf(xs, ws, fmaf(xb, wb, zwin)) + bias;
    return z > 0.0f ? z : 0.0f;
}
__device__ __forceinline__ void et_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

__device__ __forceinline__ void et_tmem_ld16(uint32_t taddr, float* v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
}

__device__ __forceinline__ void et_tmem_ld32_nowait(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
// TMA tensor store (shared -> global) of one box; completion is tracked by the issuing thread's bulk groups
__device__ __forceinline__ void et_tma_store_3d(const CUtensorMap* tmap, const void* smem_src, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];\n" ::"l"(tmap),
                 "r"((uint32_t)__cvta_generic_to_shared(smem_src)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
// the same box ADDED to what global memory holds (fp32 add performed by the memory system).  Used only by the
// CTA that owns the destination, one wave after the other: the order of the additions is fixed.
__device__ __forceinline__ void et_tma_reduce_add_3d(const CUtensorMap* tmap, const void* smem_src, int c0, int c1, int c2) {
    asm volatile("cp.reduce.async.bulk.tensor.3d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4}], [%1];\n" ::"l"(tmap),
                 "r"((uint32_t)__cvta_generic_to_shared(smem_src)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void et_bulk_commit() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void et_bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;\n" ::: "memory"); }
__device__ __forceinline__ void et_bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory"); }
__device__ __forceinline__ void et_bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory"); }
// Dataflow synchronisation between the CTAs of the (co-resident) grid.  Bounded: a flag that never
// arrives traps the kernel instead of hanging the GPU.
__device__ __forceinline__ void et_wait_count(const unsigned long long* c, unsigned long long want) {
    unsigned long long v;
    for (unsigned int spin = 0;; ++spin) {
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(c) : "memory");
        if (v >= want) return;
        if (spin > 64u) __nanosleep(32);
        if (spin > (1u << 24)) __trap();
    }
}
__device__ __forceinline__ void et_tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }


// W1[0:wn, :] (row-major [k][n]) -> K-major hi/lo copies [n][Kp], for parameters p in a float4 group
__device__ __forceinline__ void et_publish_w(const EpTcArgs& x, int64_t p, const float4& w, int H, int wn) {
    const int k = (int)(p / H), n = (int)(p - (int64_t)k * H);
    if (k >= wn) return;
    const float ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float hi = et_tf32(ww[q]);
        x.w1T_hi[(size_t)(n + q) * x.Kp + k] = hi;
        x.w1T_lo[(size_t)(n + q) * x.Kp + k] = ww[q] - hi;
    }
}

__global__ void __launch_bounds__(kEtThreads, 1)
episode_tc_kernel(const __grid_constant__ CUtensorMap tW_hi, const __grid_constant__ CUtensorMap tW_lo,
                  const __grid_constant__ CUtensorMap tP, const EpTcArgs x) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(1024) uint8_t et_smem_raw[];
    const EpisodeArgs& ea = x.ea;
    const StepArgs& a = ea.s;
    const Hyper& hp = a.hp;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nthreads = blockDim.x, nwarps = nthreads >> 5;
    const int D = hp.D, H = hp.H, wn = hp.wn;
    const int Np = x.Np, nkb = x.nkb, halves = x.halves;
    const uint32_t TB = (uint32_t)Np * 128u;                   // bytes of one [Np x 32] operand tile

    // Every pointer below is the __shared__ symbol plus an offset: an integer round trip (the usual
    // "(addr + 1023) & ~1023") makes the compiler lose the state space and emit generic LD/ST for all
    // shared-memory traffic of the kernel.  The 1024-byte alignment comes from the declaration.
    uint8_t* base = et_smem_raw;
    if (tid == 0 && ((uint32_t)__cvta_generic_to_shared(base) & 1023u) != 0u) __trap();
    const size_t off_ring = (size_t)nkb * 8192;
    const size_t off_xs = off_ring + 4 * (size_t)TB;
    const size_t off_w2s = off_xs + (size_t)halves * 16384;
    uint8_t* Afwd = base;                                       // nkb tiles of [64 x 32] (8 KB each)
    uint8_t* ring = base + off_ring;                            // 2 stages x (hi, lo) of [Np x 32]
    uint8_t* Bbwd_hi = ring;                                    // after the forward: dZ1 hi / lo, [Np x 32 episodes]
    uint8_t* Bbwd_lo = Bbwd_hi + TB;
    float* HS = reinterpret_cast<float*>(Afwd);                 // after the forward: window part of z1(s), then h(s), [Np][33]
    float* ZN = HS + (size_t)Np * kXStride;                     // after the forward: window part of z1(s'), [Np][33]
    // master window, kept as the A operand of the backward contraction: X^T [R = halves*128 rows][32 episodes],
    // K-major SWIZZLE_128B.  The rows are a ring: at step sidx of the launch window element k is row
    // (k + sidx) mod R, so sliding the window is writing ONE row (the next price); the rows outside the
    // window hold stale prices, whose products land in TMEM lanes that are never stored.
    uint8_t* XT = base + off_xs;
    const int Rm = halves * 128 - 1;
    float* W2s = reinterpret_cast<float*>(base + off_w2s);
    float* Wb = W2s + (size_t)H * 4;
    float* Ws = Wb + H;
    float* XE = Ws + H;                                         // [2][32]
    float* qpart = XE + 64;                                     // [warps][Q(s), Q(s') after Buy / Sell / Hold][3][32]
    float* dz2s = qpart + kEtWarps * 4 * 3 * 32;                // [32]
    int* kns = reinterpret_cast<int*>(dz2s + 32);               // [32]
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + ((off_w2s + (size_t)H * 24 + (64 + kEtWarps * 4 * 3 * 32 + 64) * 4 + 7) & ~(size_t)7));
    uint64_t* full = bars;          // [2]
    uint64_t* empty = bars + 2;     // [2]
    uint64_t* tf_fwd = bars + 4;
    uint64_t* tf_bwd = bars + 5;
    uint64_t* pb_bar = bars + 6;    // [4] phase B: a group of 32 partial segments has landed
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);


    const int64_t B = a.B;
    const int64_t n_tiles = (B + kTileEnvs - 1) / kTileEnvs;
    const bool has_tile = (int64_t)blockIdx.x < n_tiles;
    const int64_t e = (int64_t)blockIdx.x * kTileEnvs + lane;
    const bool valid = has_tile && e < B;
    const int64_t ec = has_tile ? (valid ? e : B - 1) : 0;
    const float bias = hp.bias;
    const float* W1g = a.params;
    const float* W2g = a.params + (size_t)D * H;
    const int q4 = warp & 3, part = warp >> 2;                  // TMEM lane quarter / column part of this warp
    const int n_ld = Np >> 4;                                   // 16-column TMEM loads per row
    const int ld0 = part * n_ld / kEtParts, ld1 = (part + 1) * n_ld / kEtParts;

    if (tid == 0) {
        mbar_init(full, 1); mbar_init(full + 1, 1); mbar_init(empty, 1); mbar_init(empty + 1, 1);
        mbar_init(tf_fwd, 1); mbar_init(tf_bwd, 1);
        for (int g = 0; g < 4; ++g) mbar_init(pb_bar + g, 1);
        mbar_fence_init();
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                         (uint32_t)__cvta_generic_to_shared(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    // zero everything the MMAs may read before it is first written (padding rows / columns)
    for (size_t i = tid; i < ((size_t)nkb * 8192 + 4 * (size_t)TB) / 16; i += nthreads)
        reinterpret_cast<float4*>(base)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    double budget = a.budget[ec], share_value = a.share_value[ec];
    int32_t shares = a.shares[ec];
    const float* my_row = a.prices + (int64_t)a.series_of_env[ec] * a.T;       // (per wave when there are several)
    for (int idx = tid; idx < halves * 4096; idx += nthreads) reinterpret_cast<float*>(XT)[idx] = 0.0f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (has_tile) {
        const int segs = (wn + 1 + 31) >> 5, items = kTileEnvs * segs;
        for (int it = warp; it < items; it += nwarps) {
            const int el = it & 31, k = ((it >> 5) << 5) + lane;
            const float* src = reinterpret_cast<const float*>(
                __shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(my_row), el)) + ea.step0;
            if (k <= wn) cp_async4(XT + et_swz(k, el), src + k);
        }
        cp_async_commit();
        cp_async_wait_pending(0);
    }
    // publish the K-major tf32 hi/lo copies of the current W1 (phase B keeps them current afterwards)
    {
        const int64_t P1 = (int64_t)D * H;
        for (int64_t p = ((int64_t)blockIdx.x * nthreads + tid) * 4; p < P1; p += (int64_t)gridDim.x * nthreads * 4)
            et_publish_w(x, p, __ldcg(reinterpret_cast<const float4*>(W1g + p)), H, wn);
    }
    // phase B's division of the parameters over the CTAs (see there)
    const int64_t P1 = (int64_t)D * H, P = P1 + 3 * (int64_t)H;
    const int n1 = (int)min(n_tiles, (int64_t)gridDim.x);      // one partial per CTA that has tiles
    const int n_w1_ctas = max(1, (int)gridDim.x - 2);
    const int chunk = (int)(((P1 + n_w1_ctas - 1) / n_w1_ctas + 3) & ~(int64_t)3);
    const int chunk2 = ((3 * H + 1) / 2 + 3) & ~3;
    const int pitch = max(chunk, chunk2);
    const int rows_cap = min(128, (int)(off_xs / ((size_t)pitch * 4)) & ~3);   // a multiple of 4: see the running sums
    // flow: the grid barrier after phase B is split.  A CTA that has published its share of the new
    // weights adds one to a counter and moves on; only warp 0 of each CTA waits for the counter (before it
    // starts the W1^T stream) while the other warps already build the A operand of the next step.
    const bool flow = gridDim.x >= 8 && pitch <= nthreads && rows_cap >= 4 && x.published != nullptr;
    const int n_publishers = (int)((P1 + chunk - 1) / chunk) + 2;
    if (flow && blockIdx.x == 0 && tid == 0) *x.published = 0ull;
    __threadfence();
    grid.sync();

    const uint32_t idesc = tc_instr_desc(128, Np);
    uint32_t ring_it = 0;                                       // k-blocks pushed through the ring so far
    uint32_t pb_phase = 0;
    // every CTA streams the same W1^T tiles: start each at a different k-block so that the L2 requests
    // of the grid spread over the whole matrix instead of all landing on one tile at a time
    const int kb_first = (int)(blockIdx.x % (unsigned)nkb);
    // More tiles than CTAs: every step runs in waves.  CTA c does tiles c, c + G, c + 2G, ...; the window and
    // the carry of a tile are then reloaded from global memory for every wave (no ring, no registers across
    // steps); the first wave stores its dW1 partial, the later ones add theirs to it (TMA reduce-add by the CTA
    // that owns the partial, in wave order), so the gradient sum over the CTAs is shared by all waves.  (TMEM has
    // no room to accumulate over the waves: forward 208 + backward 416 columns > 512.)
    const int n_waves = (int)((n_tiles + gridDim.x - 1) / gridDim.x);
    const bool multi = n_waves > 1;
    uint32_t mma_it = 0;                                        // tiles this CTA has pushed through the MMAs

    for (int32_t sidx = 0; sidx < ea.n_steps; ++sidx) {
        const int32_t istep = ea.step0 + sidx;
        const int roff = multi ? 0 : sidx;                              // ring offset of the window
        float p_next = 0.0f;
        float acc5[5] = {0.f, 0.f, 0.f, 0.f, 0.f};                      // this thread's column of dW2 and of the budget / shares rows
        float* out = ea.part1 + (size_t)blockIdx.x * ((size_t)D * H);
        float* out2 = a.dw2_part + (size_t)blockIdx.x * (3 * H);
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 0] = clock64();
        auto cta_stamp = [&](int i) {
            if (x.cta_ns && tid == 0) {
                unsigned long long t;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                x.cta_ns[((size_t)sidx * gridDim.x + blockIdx.x) * 4 + i] = (long long)t;
            }
        };
        cta_stamp(0);
      for (int wave = 0; wave < n_waves; ++wave) {
        const int64_t tile = (int64_t)wave * gridDim.x + blockIdx.x;
        const bool has_tile = tile < n_tiles;
        const int64_t e = tile * kTileEnvs + lane;
        const bool valid = has_tile && e < B;
        const uint32_t par = mma_it & 1u;
        if (multi) {
            const int64_t ecw = has_tile ? (valid ? e : B - 1) : 0;
            budget = a.budget[ecw]; share_value = a.share_value[ecw]; shares = a.shares[ecw];
            my_row = a.prices + (int64_t)a.series_of_env[ecw] * a.T;
        }
        float q_s[3] = {0.f, 0.f, 0.f}, q_n[3] = {0.f, 0.f, 0.f};
        int action = 0, explored = 0, a_eff = 2, kn = 0;
        float r32 = 0.0f, target = 0.0f, dz2 = 0.0f;
        double nb = budget, p_new = 0.0, cur = 0.0;
        int32_t ns = shares;
        uint32_t rnd[4] = {0u, 0u, 0u, 0u};
        // producer side of the W1^T ring: k-block number kk of this step into its stage
        auto push_w_tiles = [&](int kk) {
            const int kb = (kk + kb_first) % nkb;
            const uint32_t it = ring_it + (uint32_t)kk, s = it & 1u, ph = (it >> 1) & 1u;
            tc_wait(empty + s, ph ^ 1u);
            uint8_t* st = ring + (size_t)s * 2 * TB;
            mbar_arrive_expect_tx(full + s, 2 * TB);
            tma_load_2d(st, &tW_hi, kb * 32, 0, full + s);
            tma_load_2d(st + TB, &tW_lo, kb * 32, 0, full + s);
        };
        if (tid == 0) {
            // the new weights are complete once every publisher of the previous step is through
            if (flow && sidx > 0 && wave == 0) et_wait_count(x.published, (unsigned long long)sidx * (unsigned long long)n_publishers);
            // both stages are free: start the stream now, so that the first tiles travel while the warps
            // build the A operand
            fence_proxy_async();                                        // phase-B writes of W1^T -> TMA reads
            if (has_tile)
                for (int kk = 0; kk < min(2, nkb); ++kk) push_w_tiles(kk);
        }
        if (ea.dbg && blockIdx.x == 0 && tid == 0 && wave == n_waves - 1) ea.dbg[(size_t)sidx * kDbgSlots + 11] = clock64();
        if (multi) {
            // this wave's window: element k of episode el at row k (the previous wave's MMAs are through)
            if (has_tile) {
                const int segs = (wn + 1 + 31) >> 5, items = kTileEnvs * segs;
                for (int it = warp; it < items; it += nwarps) {
                    const int el = it & 31, k = ((it >> 5) << 5) + lane;
                    const float* src = reinterpret_cast<const float*>(
                        __shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(my_row), el)) + istep;
                    if (k <= wn) cp_async4(XT + et_swz(k, el), src + k);
                }
                cp_async_commit();
                cp_async_wait_pending(0);
            }
            __syncthreads();
        }
        if (ea.dbg && blockIdx.x == 0 && tid == 0 && wave == n_waves - 1) ea.dbg[(size_t)sidx * kDbgSlots + 13] = clock64();
        if (has_tile) {
            // ---- S0: the A operand of the forward (K-major, swizzled) from the master window — it does not
            // depend on the update in flight — then the small tables, once their publishers are through
            const int w_first = flow ? 1 : 0;                           // (flow: warp 0 is busy feeding the ring)
            if (warp >= w_first) {
#pragma unroll 2
                for (int it = warp - w_first; it < 16 * nkb; it += nwarps - w_first) {   // (k-block, 16-byte chunk, state): lanes = episodes
                    const int half = it & 1, chunk16 = (it >> 1) & 7, kb = it >> 4;
                    const int kk = kb * 32 + chunk16 * 4;               // columns kk..kk+3; state s' = window shifted by one
                    float v[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        v[j] = kk + j < wn ? *reinterpret_cast<const float*>(XT + et_swz((kk + j + half + roff) & Rm, lane)) : 0.0f;
                    const int row = half * 32 + lane;
                    *reinterpret_cast<float4*>(Afwd + (size_t)kb * 8192 + (size_t)row * 128 + ((chunk16 ^ (row & 7)) << 4)) =
                        make_float4(v[0], v[1], v[2], v[3]);
                }
            }
        }
        if (wave == 0) {                                                // the small tables serve every wave of the step
            if (flow) __syncthreads();                                  // warp 0 has seen the counter: the tables are final
            if (has_tile) {
                for (int j = tid; j < H; j += nthreads) {
                    reinterpret_cast<float4*>(W2s)[j] = make_float4(__ldcg(W2g + j * 3), __ldcg(W2g + j * 3 + 1), __ldcg(W2g + j * 3 + 2), 0.f);
                    Wb[j] = __ldcg(W1g + (size_t)wn * H + j);
                    Ws[j] = __ldcg(W1g + (size_t)(wn + 1) * H + j);
                }
            }
        }
        if (has_tile) {
            if (!multi && warp == 0 && (int64_t)istep + wn + 1 < a.T) p_next = __ldg(my_row + istep + wn + 1);
            et_fence_async_smem();
        }
        __syncthreads();
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 1] = clock64();
        if (has_tile) {
            // ---- S1: forward contraction on the tensor cores
            if (warp == 0 && lane == 0) {
                for (int kk = 2; kk < nkb; ++kk) push_w_tiles(kk);      // (the first two went out before S0)
            } else if (warp == 1 && lane == 0) {
                for (int kk = 0; kk < nkb; ++kk) {
                    const int kb = (kk + kb_first) % nkb;
                    const uint32_t it = ring_it + (uint32_t)kk, s = it & 1u, ph = (it >> 1) & 1u;
                    tc_wait(full + s, ph);
                    tc_fence_after();
                    uint8_t* st = ring + (size_t)s * 2 * TB;
                    const uint64_t da = tc_smem_desc(Afwd + (size_t)kb * 8192);
                    const uint64_t db_hi = tc_smem_desc(st), db_lo = tc_smem_desc(st + TB);
#pragma unroll
                    for (int k4 = 0; k4 < 4; ++k4) {
                        tc_mma_tf32(tmem_base, da + 2 * k4, db_hi + 2 * k4, idesc, (kk > 0 || k4 > 0) ? 1u : 0u);
                        tc_mma_tf32(tmem_base, da + 2 * k4, db_lo + 2 * k4, idesc, 1u);
                    }
                    tc_commit(empty + s);
                }
                tc_commit(tf_fwd);
            }
            ring_it += (uint32_t)nkb;
            // while the tensor core works: everything of the decision that does not depend on Q
            {
                const uint64_t env_id = (uint64_t)(a.env_id_base + e);
                philox4x32_10((uint32_t)env_id, (uint32_t)(env_id >> 32), (uint32_t)istep, 1u,
                              (uint32_t)hp.seed, (uint32_t)(hp.seed >> 32), rnd);
                p_new = (double)*reinterpret_cast<const float*>(XT + et_swz((wn + roff) & Rm, lane));
                cur = __dadd_rn(budget, __dmul_rn((double)shares, share_value));
            }
            tc_wait(tf_fwd, par);
            tc_fence_after();

            if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 2] = clock64();
            cta_stamp(1);
            // ---- S2: out of TMEM into shared memory: h(s) (lane quarter 0) and the window part of z1(s')
            // (lane quarter 1); the element-wise work after this runs on all twelve warps
            if (q4 < 2) {
                float* dstz = q4 == 0 ? HS : ZN;         // raw window part of z1(s) / z1(s'); h is formed where it is used
#pragma unroll 1
                for (int l = ld0; l < ld1; ++l) {
                    float v[16];
                    et_tmem_ld16(tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(l * 16), v);
#pragma unroll
                    for (int j = 0; j < 16; ++j) dstz[(l * 16 + j) * kXStride + lane] = v[j];
                }
            }
            tc_fence_before();
        }
        __syncthreads();
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 3] = clock64();

        // the three states s' the decision can lead to (index = effective action: 0 Buy, 1 Sell, 2 Hold; an
        // action that is not affordable is demoted to Hold, TrainerChildActor.scala:118-123) — known before Q is
        const double base_b = hp.bookkeeping == 0 ? a.B0 : budget;
        const int32_t base_n = hp.bookkeeping == 0 ? a.N0 : shares;
        const bool can_buy = base_b >= p_new, can_sell = base_n > 0;
        const double nbc[3] = {can_buy ? __dsub_rn(base_b, p_new) : base_b, can_sell ? __dadd_rn(base_b, p_new) : base_b, base_b};
        const int32_t nsc[3] = {can_buy ? base_n + 1 : base_n, can_sell ? base_n - 1 : base_n, base_n};
        if (has_tile) {
            // ---- S3: ONE pass over the hidden columns (dealt round-robin to the warps, fixed order) forms h(s) and
            // the partial sums of Q(s) and of Q(s') for all three outcomes; then SelectionAction, decision, reward
            // (every warp keeps its own copy) pick the outcome.  (Two passes with the decision in between cost a
            // barrier and a third more shared-memory traffic.)
            {
                float pq[4][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
                const float xb = __double2float_rn(budget), xs = (float)shares;
                const float xnb[3] = {__double2float_rn(nbc[0]), __double2float_rn(nbc[1]), __double2float_rn(nbc[2])};
                const float xns[3] = {(float)nsc[0], (float)nsc[1], (float)nsc[2]};
#pragma unroll 2
                for (int c = warp; c < H; c += nwarps) {
                    const float wb = Wb[c], ws = Ws[c], zn = ZN[c * kXStride + lane];
                    const float4 w2 = reinterpret_cast<const float4*>(W2s)[c];
                    const float h = et_hidden(HS[c * kXStride + lane], xb, xs, wb, ws, bias);
                    HS[c * kXStride + lane] = h;                        // from here on the tile holds h(s)
                    pq[0][0] = fmaf(h, w2.x, pq[0][0]); pq[0][1] = fmaf(h, w2.y, pq[0][1]); pq[0][2] = fmaf(h, w2.z, pq[0][2]);
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        const float hn = et_hidden(zn, xnb[j], xns[j], wb, ws, bias);
                        pq[1 + j][0] = fmaf(hn, w2.x, pq[1 + j][0]); pq[1 + j][1] = fmaf(hn, w2.y, pq[1 + j][1]);
                        pq[1 + j][2] = fmaf(hn, w2.z, pq[1 + j][2]);
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int t = 0; t < 3; ++t) qpart[((warp * 4 + j) * 3 + t) * 32 + lane] = pq[j][t];
            }
        }
        __syncthreads();
        if (has_tile) {
#pragma unroll
            for (int t = 0; t < 3; ++t) {
                float pw[kEtWarps];
#pragma unroll
                for (int w = 0; w < kEtWarps; ++w) pw[w] = w < nwarps ? qpart[((w * 4 + 0) * 3 + t) * 32 + lane] : 0.0f;
                float s = 0.0f;
#pragma unroll
                for (int w = 0; w < kEtWarps; ++w) s += pw[w];
                s += bias;
                q_s[t] = s > 0.0f ? s : 0.0f;
            }
            const float u = (float)(rnd[0] >> 8) * (1.0f / 16777216.0f);
            float thr = __fdiv_rn((float)istep, hp.step_div);
            if (hp.epsilon < thr) thr = hp.epsilon;
            if (u < thr) action = argmax3(q_s);
            else { explored = 1; action = (int)(((uint64_t)rnd[1] * 3u) >> 32); }
            if (action == 0 && can_buy) a_eff = 0;
            else if (action == 1 && can_sell) a_eff = 1;
            nb = a_eff == 0 ? nbc[0] : (a_eff == 1 ? nbc[1] : nbc[2]);
            ns = a_eff == 0 ? nsc[0] : (a_eff == 1 ? nsc[1] : nsc[2]);
            const double newp = __dadd_rn(nb, __dmul_rn((double)ns, p_new));
            r32 = __double2float_rn(__dsub_rn(newp, cur));
            if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 4] = clock64();

            // ---- S4: Q(s') of the outcome taken, TD target; dZ1 as a K-major (episodes contiguous) tf32 hi/lo pair = B
            // operand of the backward
#pragma unroll
            for (int t = 0; t < 3; ++t) {
                float pw[kEtWarps];
#pragma unroll
                for (int w = 0; w < kEtWarps; ++w) pw[w] = w < nwarps ? qpart[((w * 4 + 1 + a_eff) * 3 + t) * 32 + lane] : 0.0f;
                float s = 0.0f;
#pragma unroll
                for (int w = 0; w < kEtWarps; ++w) s += pw[w];
                s += bias;
                q_n[t] = s > 0.0f ? s : 0.0f;
            }
            kn = argmax3(q_n);
            const float qn_k = kn == 0 ? q_n[0] : (kn == 1 ? q_n[1] : q_n[2]);
            const float qs_k = kn == 0 ? q_s[0] : (kn == 1 ? q_s[1] : q_s[2]);
            target = __fadd_rn(r32, __fmul_rn(hp.gamma, qn_k));
            const float dq = -__fmul_rn(2.0f, __fsub_rn(target, qs_k));
            dz2 = (valid && qs_k > 0.0f) ? dq : 0.0f;
#pragma unroll 4
            for (int c = warp; c < H; c += nwarps) {
                const float d1 = HS[c * kXStride + lane] > 0.0f ? __fmul_rn(dz2, W2s[c * 4 + kn]) : 0.0f;
                const float hi = et_tf32(d1);
                const uint32_t off = et_swz(c, lane);
                *reinterpret_cast<float*>(Bbwd_hi + off) = hi;
                *reinterpret_cast<float*>(Bbwd_lo + off) = d1 - hi;
            }
            for (int c = H + warp; c < Np; c += nwarps) {               // padding columns of the MMA
                const uint32_t off = et_swz(c, lane);
                *reinterpret_cast<float*>(Bbwd_hi + off) = 0.0f;
                *reinterpret_cast<float*>(Bbwd_lo + off) = 0.0f;
            }
            if (warp == 0) {
                XE[lane] = __double2float_rn(budget);
                XE[32 + lane] = (float)shares;
                kns[lane] = kn;
                dz2s[lane] = dz2;
                if (valid) {
                    const TraceDev& t = a.trace;
                    const int64_t ti = (int64_t)(istep - a.trace_step0) * B + e;
                    if (t.action_sel) t.action_sel[ti] = action;
                    if (t.action_eff) t.action_eff[ti] = a_eff;
                    if (t.explored) t.explored[ti] = explored;
                    if (t.k_next) t.k_next[ti] = kn;
                    if (t.reward) t.reward[ti] = r32;
                    if (t.target) t.target[ti] = target;
                    if (t.budget) t.budget[ti] = nb;
                    if (t.shares) t.shares[ti] = ns;
                    if (t.q_s) { t.q_s[ti * 3] = q_s[0]; t.q_s[ti * 3 + 1] = q_s[1]; t.q_s[ti * 3 + 2] = q_s[2]; }
                    if (t.q_n) { t.q_n[ti * 3] = q_n[0]; t.q_n[ti * 3 + 1] = q_n[1]; t.q_n[ti * 3 + 2] = q_n[2]; }
                }
            }
            budget = nb; shares = ns; share_value = p_new;             // carry (TrainerChildActor.scala:99-102)
            if (multi && warp == 0 && valid) { a.budget[e] = budget; a.shares[e] = shares; a.share_value[e] = share_value; }
            et_fence_async_smem();
        }
        __syncthreads();
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 5] = clock64();

        if (has_tile) {
            // ---- S5: backward contraction on the tensor cores; dW2 and the two extra dW1 rows meanwhile
            if (warp == 1 && lane == 0) {
                tc_fence_after();
                const uint64_t db_hi = tc_smem_desc(Bbwd_hi), db_lo = tc_smem_desc(Bbwd_lo);
                for (int hf = 0; hf < halves; ++hf) {
                    const uint64_t da = tc_smem_desc(XT + (size_t)hf * 16384);
#pragma unroll
                    for (int k4 = 0; k4 < 4; ++k4) {
                        tc_mma_tf32(tmem_base + (uint32_t)(hf * Np), da + 2 * k4, db_hi + 2 * k4, idesc, k4 > 0 ? 1u : 0u);
                        tc_mma_tf32(tmem_base + (uint32_t)(hf * Np), da + 2 * k4, db_lo + 2 * k4, idesc, 1u);
                    }
                }
                tc_commit(tf_bwd);
            }
            // dW2[c][t] = sum_e [kn_e == t] h_e[c] dz2_e and the budget / shares rows of dW1: one thread per
            // hidden column (H <= blockDim), episodes summed sequentially out of shared memory (fixed order, no
            // shuffles), wave after wave into the same five registers
            if (tid < H) {
                const int c = tid;
                const float4 w2 = reinterpret_cast<const float4*>(W2s)[c];
                float s0 = acc5[0], s1 = acc5[1], s2 = acc5[2], x0 = acc5[3], x1 = acc5[4];
#pragma unroll 8
                for (int el = 0; el < kTileEnvs; ++el) {
                    const float h = HS[c * kXStride + el];
                    const float dz = dz2s[el];
                    const int k = kns[el];
                    const float hv = __fmul_rn(h, dz);
                    s0 += k == 0 ? hv : 0.0f; s1 += k == 1 ? hv : 0.0f; s2 += k == 2 ? hv : 0.0f;
                    const float w2k = k == 0 ? w2.x : (k == 1 ? w2.y : w2.z);
                    const float d1 = h > 0.0f ? __fmul_rn(dz, w2k) : 0.0f;
                    x0 = fmaf(XE[el], d1, x0);
                    x1 = fmaf(XE[32 + el], d1, x1);
                }
                acc5[0] = s0; acc5[1] = s1; acc5[2] = s2; acc5[3] = x0; acc5[4] = x1;
            }
            tc_wait(tf_bwd, par);
            tc_fence_after();
            ++mma_it;
            if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 6] = clock64();
            // ---- S6: dW1 partial out of TMEM.  A TMEM load hands every lane one row of the tile; stored
            // straight to global memory that is 32 different lines per instruction.  Instead each warp puts
            // [32 rows x 32 columns] boxes into shared memory (128-byte swizzle, conflict-free) and lane 0
            // hands them to the TMA engine, which writes whole row segments while the warp fetches the next
            // box.  Row p of the tile is window element (p - sidx) mod R; the tensor map ends at row wn, so
            // the rows outside the window are clipped; the one block that wraps stores its wrapped rows itself.
            {
                const int ncb = (H + 31) >> 5;
                // two boxes per warp in the operand ring (dZ1 is dead by now) as far as it reaches; the last
                // warps get one box each behind h(s) in the A-operand region
                const int n_ring = min(kEtWarps, (int)(4 * (size_t)TB / 8192));
                const int nbuf = warp < n_ring ? 2 : 1;
                uint8_t* stg = warp < n_ring ? ring + (size_t)warp * 8192
                                             : base + (((size_t)Np * kXStride * 4 + 1023) & ~(size_t)1023) + (size_t)(warp - n_ring) * 4096;
                int n_box = 0;
                // the additions below must find the previous wave's store / additions complete (they are, a whole
                // wave later; this makes it a guarantee)
                if (wave > 0) et_bulk_wait_all();
                for (int i = part; i < halves * ncb; i += kEtParts) {
                    const int hf = i / ncb, cb = i - hf * ncb;
                    const int k0 = (hf * 128 + q4 * 32 - roff) & Rm;
                    const bool st_a = k0 < wn, st_b = k0 + 32 > Rm + 1;
                    if (!st_a && !st_b) continue;
                    uint32_t r[32];
                    et_tmem_ld32_nowait(tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(hf * Np + cb * 32), r);
                    if (n_box >= nbuf) {                                // the buffer's previous box has left it
                        if (lane == 0) { if (nbuf == 2) et_bulk_wait_read1(); else et_bulk_wait_read0(); }
                        __syncwarp();
                    }
                    et_tmem_wait_ld();
                    uint8_t* buf = stg + (size_t)(n_box & (nbuf - 1)) * 4096;
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<uint4*>(buf + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                            make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
                    et_fence_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        if (st_a) {                                     // later waves add to what the first one stored
                            if (wave == 0) et_tma_store_3d(&tP, buf, cb * 32, k0, (int)blockIdx.x);
                            else et_tma_reduce_add_3d(&tP, buf, cb * 32, k0, (int)blockIdx.x);
                        }
                        et_bulk_commit();
                    }
                    ++n_box;
                    if (st_b) {
                        // the rows that wrapped around the ring (TMA store coordinates cannot be negative):
                        // one 32-row block per step, stored row by row from the registers
                        const int k = k0 + lane - (Rm + 1);
                        if (k >= 0 && k < wn) {
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                if (cb * 32 + 4 * j < H)
                                    *reinterpret_cast<uint4*>(out + (size_t)k * H + cb * 32 + 4 * j) =
                                        make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
                        }
                    }
                }
            }
            // the next wave's W1^T tiles land on the staging buffers: the TMA engine must have read them
            if (wave + 1 < n_waves) et_bulk_wait_read0();
            tc_fence_before();
        }
        __syncthreads();
      }   // waves
        if (has_tile && tid < H) {                                      // (has_tile: the CTA's first tile; it has at least that one)
            out2[tid * 3 + 0] = acc5[0]; out2[tid * 3 + 1] = acc5[1]; out2[tid * 3 + 2] = acc5[2];
            out[(size_t)wn * H + tid] = acc5[3];
            out[(size_t)(wn + 1) * H + tid] = acc5[4];
        }
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 7] = clock64();
        if (!ea.apply_in_kernel) { et_bulk_wait_all(); break; }
        // ---- S7: slide the window: one new row of the ring
        if (!multi && has_tile && warp == 0) *reinterpret_cast<float*>(XT + et_swz((wn + 1 + sidx) & Rm, lane)) = p_next;
        et_bulk_wait_all();                                             // the partial is in global memory
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 8] = clock64();
        cta_stamp(2);
        __threadfence();
        grid.sync();
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 9] = clock64();

        // ---- phase B: gradient sum over the tiles + AdaGrad (as episode_kernel) + tf32 hi/lo of the new W1.
        // W1: CTA c owns `chunk` consecutive parameters; their n_tiles partial segments come in as bulk
        // copies (all in flight at once, one mbarrier per group of 32), then one thread per parameter sums them out of
        // shared memory in the order of the FFMA kernel: four interleaved running sums, (0+1)+(2+3).
        // W2 (3*H parameters, its own partial buffer) is the last two CTAs' share, handled the same way.
        if (flow) {
            const bool w2_cta = (int)blockIdx.x >= (int)gridDim.x - 2;
            const float* src = w2_cta ? a.dw2_part : ea.part1;
            const int64_t stride = w2_cta ? 3 * (int64_t)H : P1;
            const int64_t off = w2_cta ? (int64_t)((int)blockIdx.x - ((int)gridDim.x - 2)) * chunk2 : (int64_t)blockIdx.x * chunk;
            const int64_t p0 = w2_cta ? P1 + off : off;
            const int len = (int)max((int64_t)0, min((int64_t)(w2_cta ? chunk2 : chunk), stride - off));
            if (len > 0) {
                float* stage = reinterpret_cast<float*>(base);          // [rows][pitch], over the operand regions
                const int64_t p = p0 + tid;
                float av = 0.0f, wv = 0.0f;
                if (tid < len) { av = ea.acc[p]; wv = ea.params_rw[p]; }
                float s4[4] = {0.f, 0.f, 0.f, 0.f};
                for (int r0 = 0; r0 < n1; r0 += rows_cap) {
                    const int rows = min(rows_cap, n1 - r0), ngrp = (rows + 31) >> 5;
                    if (tid < ngrp) mbar_arrive_expect_tx(pb_bar + tid, (uint32_t)min(32, rows - tid * 32) * (uint32_t)len * 4u);
                    __syncthreads();
                    if (tid < rows) {
                        fence_proxy_async();
                        bulk_g2s(stage + (size_t)tid * pitch, src + (size_t)(r0 + tid) * stride + off, (uint32_t)len * 4u,
                                 pb_bar + (tid >> 5));
                    }
                    for (int g = 0; g < ngrp; ++g) {
                        mbar_wait(pb_bar + g, (pb_phase >> g) & 1u);
                        if (tid < len) {
                            // r0 and the group starts are multiples of 4, so j mod 4 is the tile index mod 4
                            const int j1 = min(rows, g * 32 + 32);
                            int j = g * 32;
#pragma unroll 2
                            for (; j + 3 < j1; j += 4) {
                                s4[0] += stage[(size_t)(j + 0) * pitch + tid];
                                s4[1] += stage[(size_t)(j + 1) * pitch + tid];
                                s4[2] += stage[(size_t)(j + 2) * pitch + tid];
                                s4[3] += stage[(size_t)(j + 3) * pitch + tid];
                            }
                            for (; j < j1; ++j) s4[j & 3] += stage[(size_t)j * pitch + tid];
                        }
                    }
                    pb_phase ^= (1u << ngrp) - 1u;
                    if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 14] = clock64();
                    __syncthreads();                                    // the stage is free for the next pass
                }
                if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 15] = clock64();
                float g = (s4[0] + s4[1]) + (s4[2] + s4[3]);
                if (ea.world > 1 && tid < len) {
                    // ---- gradient all-reduce over NVLink peer memory, share by share and without a fence:
                    // every value travels as one 8-byte store {value, step tag} into the slot this rank owns in
                    // every peer's exchange block (posted writes, one NVLink crossing); the receiver spins on its
                    // LOCAL slots until the tag is this step's, then adds the values in rank order — the same
                    // bits on every rank.  Two buffers by (global) step parity: a rank cannot send step t+2 before it has
                    // received this rank's step t+1, which is sent after step t has been read.
                    const unsigned int want = ea.flag_base + (unsigned int)sidx + 1u;
                    const size_t slot = (size_t)(want & 1u) * (size_t)ea.world;     // (global step parity: alternates across launches too)
                    // ONE 64-bit scalar store per value (tag in the upper half): single-copy atomic by the memory model,
                    // which a two-element vector store is not
                    const unsigned long long mine = ((unsigned long long)want << 32) | (unsigned long long)__float_as_uint(g);
#pragma unroll
                    for (int r = 0; r < 8; ++r)
                        if (r < ea.world && r != ea.rank)
                            asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(x.pdata[r] + (slot + (size_t)ea.rank) * (size_t)x.p_pad + p),
                                         "l"(mine) : "memory");
                    const unsigned long long* local = x.pdata[ea.rank] + slot * (size_t)x.p_pad + p;
                    float pv[8];
#pragma unroll
                    for (int r = 0; r < 8; ++r) pv[r] = 0.0f;
                    pv[ea.rank & 7] = g;
                    unsigned int pending = ((1u << ea.world) - 1u) & ~(1u << ea.rank);
                    unsigned long long t0 = 0;
                    for (unsigned int spins = 0; pending != 0u; ++spins) {
                        // all outstanding slots are read in one go (independent loads), then inspected
                        unsigned long long v[8];
#pragma unroll
                        for (int r = 0; r < 8; ++r)
                            if ((pending >> r) & 1u)
                                asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v[r])
                                             : "l"(local + (size_t)r * (size_t)x.p_pad) : "memory");
#pragma unroll
                        for (int r = 0; r < 8; ++r)
                            if (((pending >> r) & 1u) && (unsigned int)(v[r] >> 32) == want) {
                                pv[r] = __uint_as_float((unsigned int)v[r]);
                                pending &= ~(1u << r);
                            }
                        if ((spins & 4095u) == 4095u) {
                            unsigned long long t;
                            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                            if (t0 == 0) t0 = t;
                            else if (t - t0 > 20000000000ull) __trap();   // a rank never arrived: fail instead of hanging
                        }
                    }
                    g = 0.0f;
#pragma unroll
                    for (int r = 0; r < 8; ++r) g += pv[r];
                }
                if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 12] = clock64();
                if (tid < len) {
                    ea.grad[p] = g;
                    av = __fadd_rn(av, __fmul_rn(g, g));
                    wv = __fsub_rn(wv, __fmul_rn(__fmul_rn(g, hp.lr), __fdiv_rn(1.0f, __fsqrt_rn(av))));
                    ea.acc[p] = av;
                    ea.params_rw[p] = wv;
                    const int k = (int)(p / H), n = (int)(p - (int64_t)k * H);
                    if (k < wn) {                                       // (W2 and the budget / shares rows: k >= wn)
                        const float hi = et_tf32(wv);
                        x.w1T_hi[(size_t)n * x.Kp + k] = hi;
                        x.w1T_lo[(size_t)n * x.Kp + k] = wv - hi;
                    }
                }
                // tell the consumers
                __syncthreads();
                if (tid == 0)     // release at GPU scope: cumulative over the CTA's stores the barrier above has ordered
                    asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(x.published), "l"(1ull) : "memory");
            }
        } else {
            // shapes the bulk scheme does not cover: four lanes per float4 group, straight from global memory
            const int cta_lo = 0;
            const int64_t total_threads = (int64_t)gridDim.x * nthreads;
            const int64_t task_lo = 0;
            // whole warps stay in the loop (shuffles below): the bound is on the warp's first task
            for (int64_t task = task_lo + (int64_t)((int)blockIdx.x - cta_lo) * nthreads + tid; task - lane < P; task += total_threads) {
                const int64_t p = (task >> 2) << 2;
                const int sub = (int)(task & 3);
                const bool in_range = p < P;
                float4 gsum = make_float4(0.f, 0.f, 0.f, 0.f);
                if (in_range) {
                    const float* src; int64_t stride, off;
                    if (p < P1) { src = ea.part1; stride = P1; off = p; }
                    else { src = a.dw2_part; stride = P - P1; off = p - P1; }
                    for (int j0 = sub; j0 < n1; j0 += 32) {
                        float4 v[8];
#pragma unroll
                        for (int uu = 0; uu < 8; ++uu) {
                            const int j = j0 + 4 * uu;
                            v[uu] = (j < n1) ? __ldcg(reinterpret_cast<const float4*>(src + (int64_t)j * stride + off))
                                             : make_float4(0.f, 0.f, 0.f, 0.f);
                        }
#pragma unroll
                        for (int uu = 0; uu < 8; ++uu) { gsum.x += v[uu].x; gsum.y += v[uu].y; gsum.z += v[uu].z; gsum.w += v[uu].w; }
                    }
                }
#pragma unroll
                for (int m = 1; m <= 2; m <<= 1) {
                    gsum.x += __shfl_xor_sync(0xffffffffu, gsum.x, m);
                    gsum.y += __shfl_xor_sync(0xffffffffu, gsum.y, m);
                    gsum.z += __shfl_xor_sync(0xffffffffu, gsum.z, m);
                    gsum.w += __shfl_xor_sync(0xffffffffu, gsum.w, m);
                }
                if (in_range && sub == 0) {
                    *reinterpret_cast<float4*>(ea.grad + p) = gsum;
                    float4 av = *reinterpret_cast<float4*>(ea.acc + p);
                    float4 wv = *reinterpret_cast<float4*>(ea.params_rw + p);
                    const float gg[4] = {gsum.x, gsum.y, gsum.z, gsum.w};
                    float aa[4] = {av.x, av.y, av.z, av.w};
                    float ww[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
                    for (int qq = 0; qq < 4; ++qq) {
                        aa[qq] = __fadd_rn(aa[qq], __fmul_rn(gg[qq], gg[qq]));
                        ww[qq] = __fsub_rn(ww[qq], __fmul_rn(__fmul_rn(gg[qq], hp.lr), __fdiv_rn(1.0f, __fsqrt_rn(aa[qq]))));
                    }
                    const float4 wnew = make_float4(ww[0], ww[1], ww[2], ww[3]);
                    *reinterpret_cast<float4*>(ea.acc + p) = make_float4(aa[0], aa[1], aa[2], aa[3]);
                    *reinterpret_cast<float4*>(ea.params_rw + p) = wnew;
                    if (p < P1) et_publish_w(x, p, wnew, H, wn);
                }
            }
        }
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 10] = clock64();
        cta_stamp(3);
        if (!flow) {
            __threadfence();
            grid.sync();
        }
    }

    if (!multi && has_tile && warp == 0 && valid) {
        a.budget[e] = budget;
        a.shares[e] = shares;
        a.share_value[e] = share_value;
    }
    if (ea.apply_in_kernel && blockIdx.x == 0 && tid == 0) *ea.d_step_rw = ea.step0 + ea.n_steps;
    tc_fence_before();
    __syncthreads();
    if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(512u) : "memory");
}

#endif  // __CUDACC__
}  // namespace qp
