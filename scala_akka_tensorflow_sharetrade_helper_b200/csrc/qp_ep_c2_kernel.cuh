// qp_ep_c2_kernel.cuh — body of episode_c2_kernel (see qp_ep_c2.cuh for the design).
#pragma once

namespace qp {
#ifdef __CUDACC__

// kXchg: 0 one GPU, 1 several GPUs with the all-to-all share exchange in phase B, 2 with reduce-scatter + all-gather.  A
// template parameter because the kernel sits at the register limit of 384 threads per SM (168): each instantiation
// carries only the scheme it runs, and every extra live value costs spills.
// kWaves: more tiles than pairs.  Pair i then takes the tiles i, i + pairs, ... one after the other in every env-step: the
// window of a tile is gathered from the price matrix per (tile, step) and its episode record lives in global memory between
// steps; the small tables and the wait for the new weights happen once per step; the backward MMAs of all tiles of a step
// accumulate into the SAME TMEM accumulator (and dW2 / the budget and shares rows into the same shared-memory sums), so the
// pair still stores ONE partial per step, after its last tile.
template <int kXchg, bool kSplit, bool kWaves>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kC2Threads, 1)
episode_c2_kernel(const __grid_constant__ CUtensorMap tW_hi, const __grid_constant__ CUtensorMap tW_lo,
                  const __grid_constant__ CUtensorMap tP, const __grid_constant__ CUtensorMap tX, const EpC2Args x) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(1024) uint8_t c2_smem_raw[];
    const EpisodeArgs& ea = x.ea;
    const StepArgs& a = ea.s;
    const Hyper& hp = a.hp;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int nthreads = kC2Threads;
    const int D = hp.D, H = hp.H, wn = hp.wn;
    const int Hh = x.Hh, Nh = x.Nh, nkb = x.nkb;
    const int halves = (wn + 127) >> 7;
    const C2Smem sm = c2_smem(Hh, Nh);
    const uint32_t TBh = (uint32_t)sm.TBh;

    // every pointer is the __shared__ symbol plus an offset (keeps the shared state space: LDS / STS)
    uint8_t* base = c2_smem_raw;
    if (tid == 0 && ((uint32_t)__cvta_generic_to_shared(base) & 1023u) != 0u) __trap();
    uint8_t* XT = base + sm.xt;                                 // [2 blocks][256 rows][32 episodes], SWIZZLE_128B by address
    uint8_t* regA = base + sm.op;                               // the operand region: forward A ring ...
    uint8_t* regW = base + sm.wring;                            // ... and W1^T ring; then (see c2_smem):
    float* HS = reinterpret_cast<float*>(regA);                 // [Hh][65]: window part of z1(s), then h(s)
    float* ZN = HS + (size_t)Hh * kC2EStride;                   // [Hh][65]: window part of z1(s')
    uint8_t* Bbwd_hi = base + sm.bwd;                           // [2 episode blocks][Nh x 32]
    uint8_t* Bbwd_lo = Bbwd_hi + 2 * (size_t)TBh;
    float* W2s = reinterpret_cast<float*>(base + sm.w2s);       // [Hh] float4
    float* Wb = reinterpret_cast<float*>(base + sm.wb);
    float* Ws = reinterpret_cast<float*>(base + sm.ws);
    float* XE = reinterpret_cast<float*>(base + sm.xe);         // [2][64]
    float* dz2s = reinterpret_cast<float*>(base + sm.dz2s);
    int* kns = reinterpret_cast<int*>(base + sm.kns);
    float* qpart = reinterpret_cast<float*>(base + sm.bwd);     // [6 warps of a lane group][12][64] (consumed before dZ1 is written)
    float* qsum = reinterpret_cast<float*>(base + sm.qsum);     // [2 CTAs of the pair][12][64]
    float* part5 = reinterpret_cast<float*>(base + sm.part5);   // [2 lane groups][5][Hh]
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + sm.bars);
    uint64_t* w_full = bars;            // [kC2WStages <= 4]
    uint64_t* w_empty = bars + 4;       // [kC2WStages]
    uint64_t* a_full = bars + 8;        // [kC2AStages <= 4]
    uint64_t* a_empty = bars + 12;      // [kC2AStages]
    uint64_t* tf_fwd = bars + 16;
    uint64_t* tf_bwd = bars + 17;       // [2]: one per 128-row half of the backward accumulator
    uint64_t* qx_bar = bars + 19;       // the pair's exchange of the Q partial sums: one arrival per CTA and step
    uint64_t* xw_full = bars + 20;      // kWaves: the tile's window has arrived (TMA from the transposed price matrix)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 23);

    const uint32_t crank = c2_cluster_rank();                   // which half of the hidden columns
    const int c0 = (int)crank * Hh;                             // first hidden column of this CTA
    const int64_t cidx = (int64_t)(blockIdx.x >> 1), n_clusters = (int64_t)(gridDim.x >> 1);
    int64_t pair = cidx;                                        // (kWaves: the tile in hand)
    const int64_t B = a.B;
    // kSplit: prices that are not tf32-exact.  A pair then owns 32 episodes, whose windows lie in the buffer twice: block 0
    // holds tf32(x), block 1 the remainder x - tf32(x) as the windows of 32 VIRTUAL episodes.  The window part of layer 1
    // is linear in x, so the same MMAs give z(x_hi) and z(x_lo) in separate rows (added after the TMEM load), and the
    // backward contraction over [32 real | 32 virtual] rows with dZ1 written to both blocks is X_hi^T dZ1 + X_lo^T dZ1.
    constexpr int kEpp = kSplit ? 32 : kC2Envs;                 // real episodes per pair
    const int64_t n_pairs = (B + kEpp - 1) / kEpp;
    const bool has_tile = pair < n_pairs;                       // (kWaves: whether this pair has any tile)
    const int g = warp & 1, wsub = warp >> 1;                   // element-wise phases: lane group / column subset
    const int el = g * 32 + lane;                               // episode of this thread within the pair's 64
    int64_t e = pair * kEpp + (kSplit ? lane : el);
    bool valid = has_tile && e < B;
    int64_t ec = has_tile ? (valid ? e : B - 1) : 0;
    const float bias = hp.bias;
    const float* W1g = a.params;
    const float* W2g = a.params + (size_t)D * H;
    const int q4 = warp & 3, part = warp >> 2;                  // TMEM lane quarter / column part of this warp

    if (tid == 0) {
        for (int s = 0; s < kC2WStages; ++s) { mbar_init(w_full + s, 1); mbar_init(w_empty + s, 1); }
        for (int s = 0; s < kC2AStages; ++s) { mbar_init(a_full + s, kC2Builders); mbar_init(a_empty + s, 1); }
        mbar_init(tf_fwd, 1); mbar_init(tf_bwd, 1); mbar_init(tf_bwd + 1, 1);
        mbar_init(qx_bar, 2);
        mbar_init(xw_full, 1);
        mbar_fence_init();
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                         (uint32_t)__cvta_generic_to_shared(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    for (size_t i = tid; i < (sm.op + sm.op_bytes) / 16; i += nthreads) reinterpret_cast<float4*>(base)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    double budget = a.budget[ec], share_value = a.share_value[ec];
    int32_t shares = a.shares[ec];
    const float* my_row = a.prices + (int64_t)a.series_of_env[ec] * a.T;      // (kWaves: set again per tile)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // the window of every episode of the tile: element k of episode (g, l) at row k of block g
    auto load_window = [&](int32_t first_step) {
        const int segs = (wn + 1 + 31) >> 5, items = 32 * segs;
        for (int it = wsub; it < items; it += kC2Sub) {
            const int l = it & 31, k = ((it >> 5) << 5) + lane;
            const float* src = reinterpret_cast<const float*>(
                __shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(my_row), l)) + first_step;
            if constexpr (kSplit) {
                if (k <= wn) {
                    const float v = __ldg(src + k), hi = et_tf32(v);
                    *reinterpret_cast<float*>(XT + (size_t)g * 32768 + et_swz(k, l)) = g == 0 ? hi : v - hi;
                }
            } else {
                if (k <= wn) cp_async4(XT + (size_t)g * 32768 + et_swz(k, l), src + k);
            }
        }
        cp_async_commit();
        cp_async_wait_pending(0);
        et_fence_async_smem();                                  // (the backward MMA reads the buffer through the async proxy)
    };
    if (has_tile && !kWaves) load_window(ea.step0);
    const int64_t P1 = (int64_t)D * H, P = P1 + 3 * (int64_t)H;
    {   // K-major tf32 hi/lo copies of the current W1 (phase B keeps them current afterwards)
        for (int64_t p = ((int64_t)blockIdx.x * nthreads + tid) * 4; p < P1; p += (int64_t)gridDim.x * nthreads * 4) {
            const float4 w = __ldcg(reinterpret_cast<const float4*>(W1g + p));
            const int k = (int)(p / H), n = (int)(p - (int64_t)k * H);
            if (k < wn) {
                const float ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float hi = et_tf32(ww[q]);
                    x.w1T_hi[(size_t)(n + q) * x.Kp + k] = hi;
                    x.w1T_lo[(size_t)(n + q) * x.Kp + k] = ww[q] - hi;
                }
            }
        }
    }
    // phase B's division of the parameters over the CTAs (as episode_tc_kernel; the host has checked `flow`)
    const int n1 = (int)min(n_pairs, n_clusters);               // one partial per pair
    const int Kk = c2_kk(D);
    const int64_t P1T = (int64_t)H * Kk;                        // the transposed W1 index space (one pad element per column)
    const int n_w1_ctas = max(1, (int)gridDim.x - 2);
    const int chunk = (int)(((P1T + n_w1_ctas - 1) / n_w1_ctas + 3) & ~(int64_t)3);
    const int chunk2 = ((3 * H + 1) / 2 + 3) & ~3;
    const int pitch = max(chunk, chunk2);
    if (!(gridDim.x >= 8 && pitch <= nthreads && x.published != nullptr)) __trap();        // (the host has checked)
    const int n_publishers = (int)((P1T + chunk - 1) / chunk) + 2;
    // this CTA's share of the parameters and this thread's parameter (phase B)
    const bool w2_cta = (int)blockIdx.x >= (int)gridDim.x - 2;
    const float* pb_src = w2_cta ? a.dw2_part : ea.part1;
    const int64_t pb_stride = w2_cta ? 3 * (int64_t)H : P1T;
    const int64_t pb_off = w2_cta ? (int64_t)((int)blockIdx.x - ((int)gridDim.x - 2)) * chunk2 : (int64_t)blockIdx.x * chunk;
    const int64_t pb_p0 = w2_cta ? P1T + pb_off : pb_off;       // (index in the exchange blocks: the same on every rank)
    // the parameter thread t of this CTA owns, or -1 for a pad element: W1[k][n] for the transposed index n * Kk + k
    auto pb_param = [&](int t) -> int64_t {
        if (w2_cta) return P1 + pb_off + t;
        const int64_t q = pb_off + t;
        const int n = (int)(q / Kk), k = (int)(q - (int64_t)n * Kk);
        return k < D ? (int64_t)k * H + n : (int64_t)-1;
    };
    const int pb_len = (int)max((int64_t)0, min((int64_t)(w2_cta ? chunk2 : chunk), pb_stride - pb_off));
    // (column, row) of this thread's element of the transposed index space, packed: computed once, used every step
    unsigned int pb_nk = 0u;
    if (!w2_cta) {
        const unsigned int qq = (unsigned int)pb_off + (unsigned int)tid, nq0 = qq / (unsigned int)Kk;
        pb_nk = (nq0 << 16) | (qq - nq0 * (unsigned int)Kk);
    }
    float pb_w = 0.0f, pb_a = 0.0f, pb_g = 0.0f;
    if (tid < pb_len && pb_param(tid) >= 0) { pb_a = ea.acc[pb_param(tid)]; pb_w = ea.params_rw[pb_param(tid)]; }
    if (blockIdx.x == 0 && tid == 0) { x.published[0] = 0ull; x.published[1] = 0ull; }
    __threadfence();
    grid.sync();

    const uint32_t idesc = tc_instr_desc(128, Nh);
    uint32_t ring_it = 0, a_it = 0, mma_it = 0, xw_it = 0;
    const int kb_first = (int)(blockIdx.x % (unsigned)nkb);
    float* out = ea.part1 + (size_t)cidx * (size_t)P1T;        // [H][Kk]
    float* out2 = a.dw2_part + (size_t)cidx * (3 * H);

    for (int32_t sidx = 0; sidx < ea.n_steps; ++sidx) {
        const int32_t istep = ea.step0 + sidx;
        const int roff = kWaves ? 0 : sidx % kC2Slide;          // row of window element 0
        auto stamp = [&](int i) { if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + i] = clock64(); };
        auto cta_stamp = [&](int i) {
            if (x.cta_ns && tid == 0) {
                unsigned long long t;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                x.cta_ns[((size_t)sidx * gridDim.x + blockIdx.x) * 8 + i] = (long long)t;
            }
        };
        cta_stamp(0);
        stamp(0);
        auto fine = [&](int i) { if (x.fine && blockIdx.x == 0) x.fine[(size_t)sidx * 64 + i] = clock64(); };   // tuning aid
        // the W1^T ring: (hi, lo) tiles of this CTA's rows for k-block kk of the tile in hand
        auto push_w_tiles = [&](int kk) {
            const int kb = (kk + kb_first) % nkb;
            const uint32_t it = ring_it + (uint32_t)kk, s = it % kC2WStages, ph = (it / kC2WStages) & 1u;
            tc_wait(w_empty + s, ph ^ 1u);
            fine(kk);
            uint8_t* st = regW + (size_t)s * 2 * TBh;
            mbar_arrive_expect_tx(w_full + s, 2 * TBh);
            tma_load_2d(st, &tW_hi, kb * 32, c0, w_full + s);
            tma_load_2d(st + TBh, &tW_lo, kb * 32, c0, w_full + s);
        };
        if constexpr (!kWaves) {
            constexpr bool first_tile = true, last_tile = true;
            if (has_tile) {
#include "qp_ep_c2_phaseA.inl"
            }
        } else {
            for (int64_t tile = cidx; tile < n_pairs; tile += n_clusters) {
                const bool first_tile = tile == cidx, last_tile = tile + n_clusters >= n_pairs;
                pair = tile;
                e = pair * kEpp + el;
                valid = e < B;
                ec = valid ? e : B - 1;
                if (last_tile) stamp(11);
                budget = a.budget[ec]; share_value = a.share_value[ec]; shares = a.shares[ec];
                my_row = a.prices + (int64_t)a.series_of_env[ec] * a.T;
                // (the previous tile's backward MMAs have completed: its S6 waits for them)
                if (x.win_tma) {
                    // episode e is series e: the window of the tile is two boxes [wn + 1 rows x 32 episodes] of the transposed
                    // price matrix, fetched by the TMA engine straight into the buffer's swizzled layout; the first W1^T tiles
                    // of the tile are queued behind them
                    if (warp == 0 && lane == 0) {
                        mbar_arrive_expect_tx(xw_full, 2u * (uint32_t)(wn + 1) * 128u);
                        tma_load_2d(XT, &tX, (int)(pair * kC2Envs), istep, xw_full);
                        tma_load_2d(XT + 32768, &tX, (int)(pair * kC2Envs) + 32, istep, xw_full);
                        if (first_tile) {
                            if (sidx > 0) et_wait_count(x.published, (unsigned long long)sidx * (unsigned long long)n_publishers);
                            fence_proxy_async();                // phase-B writes of W1^T -> TMA reads (once per step)
                        }
                        for (int kk = 0; kk < min(kC2WStages, nkb); ++kk) push_w_tiles(kk);
                    }
                    if (lane == 0) tc_wait(xw_full, xw_it & 1u);
                    __syncwarp();
                    ++xw_it;
                } else {
                    // any episode -> series map: gathered with 4-byte copies (the W1^T stream starts afterwards: started
                    // before, it slowed the gather down four-fold)
                    load_window(istep);
                    __syncthreads();
                    if (warp == 0 && lane == 0) {
                        if (first_tile) {
                            if (sidx > 0) et_wait_count(x.published, (unsigned long long)sidx * (unsigned long long)n_publishers);
                            fence_proxy_async();
                        }
                        for (int kk = 0; kk < min(kC2WStages, nkb); ++kk) push_w_tiles(kk);
                    }
                }
                if (last_tile) stamp(13);
#include "qp_ep_c2_phaseA.inl"
                if (crank == 0 && wsub == 0 && valid) {         // the episode record goes home between steps
                    a.budget[e] = budget;
                    a.shares[e] = shares;
                    a.share_value[e] = share_value;
                }
            }
        }
        stamp(7);
        if (!ea.apply_in_kernel) { et_bulk_wait_all(); break; }
        et_bulk_wait_all();                                     // the partial is in global memory
        stamp(8);
        cta_stamp(4);
        // ---- grid barrier: every partial is complete.  One release-add per CTA on a counter that only grows, one
        // acquiring poller per CTA (cooperative_groups' grid.sync took 1.8 us from the last arrival; this takes about 1).
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(x.published + 1), "l"(1ull) : "memory");
            et_wait_count(x.published + 1, (unsigned long long)(sidx + 1) * (unsigned long long)gridDim.x);
        }
        __syncthreads();
        stamp(9);
        cta_stamp(5);
#include "qp_ep_c2_phaseB.inl"
        stamp(10);
        cta_stamp(7);
    }

    if (ea.apply_in_kernel && tid < pb_len && pb_param(tid) >= 0) {   // weights, accumulators and the last gradient sum go home
        const int64_t p = pb_param(tid);
        ea.params_rw[p] = pb_w;
        ea.acc[p] = pb_a;
        ea.grad[p] = pb_g;
    }
    if (!kWaves && has_tile && crank == 0 && wsub == 0 && valid && (!kSplit || g == 0)) {
        a.budget[e] = budget;
        a.shares[e] = shares;
        a.share_value[e] = share_value;
    }
    if (ea.apply_in_kernel && blockIdx.x == 0 && tid == 0) *ea.d_step_rw = ea.step0 + ea.n_steps;
    tc_fence_before();
    __syncthreads();
    c2_cluster_sync();                                          // nobody leaves while its partner may still write to it
    if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "r"(512u) : "memory");
}

#endif  // __CUDACC__
}  // namespace qp
