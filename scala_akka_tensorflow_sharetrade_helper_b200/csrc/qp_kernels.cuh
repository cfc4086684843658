// qp_kernels.cuh — sm_100a kernels of the Q-learning decision policy hot path.
//
// One env-step of every episode is three launches on one stream:
//   step_kernel   (K1)  state build from HBM-resident price series, forward of s and s',
//                       Philox epsilon-greedy select, portfolio/reward book-keeping, TD target,
//                       output-layer backward; writes dz1 [B,H] and per-CTA dW2 partials
//   dw1_kernel    (K2)  dW1 = X^T · dZ1 as a split-over-episodes register-tiled contraction;
//                       writes per-CTA partials
//   apply_kernel  (K3)  fixed-order reduction of the partials + AdaGrad
// Reference lines each piece replaces are cited at the piece.  No float atomics anywhere: every
// sum has a fixed order, so a run is bit-reproducible.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#ifdef __CUDACC__
#include <cooperative_groups.h>
#endif

namespace qp {

constexpr int kColsPerWarp = 20;   // hidden columns owned by one warp of the step kernel
constexpr int kTileEnvs = 32;      // episodes per tile = lanes of a warp
constexpr int kXStride = 33;       // padded row stride of the transposed state tile (bank-conflict free)
constexpr int kMaxStepWarps = 12;  // hidden <= 12 * 20 = 240 on this path
constexpr int kDw1Rows = 8;        // register tile of the dW1 contraction: 8 input rows
constexpr int kDw1Cols = 4;        //   x 4 hidden columns (one float4 group)
constexpr int kDw1MaxThreads = 384;
constexpr int kDw1ChunkEnvs = 32;  // episodes per pipeline stage of the dW1 kernel
constexpr int kDw1Stages = 2;
constexpr int kDw1TileCols = 52;   // target width of one column tile of dW1 (grid.y = ceil(H / 52))
constexpr int kDbgSlots = 16;      // clock64 stamps per env-step (qp_debug_phase_clocks)
constexpr int kApplyLanes = 8;     // lanes cooperating on one parameter in the apply kernel
constexpr int kEpRows = 16;        // persistent kernel: register tile of its dW1 contraction, 16 rows
constexpr int kEpCols = 8;         //   x 8 columns (float4 groups ct and ct + n_ct)
constexpr int kEpMaxThreads = 384;

enum StepMode { kEnvStep = 0, kMsgSelect = 1, kMsgUpdate = 2 };

struct Hyper {
    int wn;        // price-window length (hist + 1)
    int D;         // wn + 2
    int H;         // hidden
    int Hs;        // hidden padded to a multiple of kColsPerWarp (smem row stride of W1)
    float epsilon, gamma, lr, bias, step_div;
    int bookkeeping;
    uint64_t seed;
};

struct TraceDev {
    int32_t* action_sel; int32_t* action_eff; int32_t* explored; int32_t* k_next;
    float* reward; float* q_s; float* q_n; float* target; double* budget; int32_t* shares;
};

struct StepArgs {
    Hyper hp;
    const float* params;            // [W1 | W2]
    // episode mode
    const float* prices; int64_t T; const int32_t* series_of_env;
    double* budget; int32_t* shares; double* share_value;
    double B0; int32_t N0;
    int64_t env_id_base;
    const int32_t* d_step;
    // message mode
    const float* states; const float* next_states; const float* rewards; const float* steps;
    const int64_t* env_ids; int32_t* actions_out; float* q_out;
    // common
    int64_t B;                      // episodes / rows
    int64_t Bpad;
    float* dz1;                     // [Bpad, H]
    float* xextra;                  // [2, Bpad]: (float)budget, (float)shares of state s
    float* dw2_part;                // [grid, 3H]
    TraceDev trace; int32_t trace_step0;
};

struct Dw1Args {
    Hyper hp;
    const float* prices; int64_t T; const int32_t* series_of_env; const int32_t* d_step;
    const float* states;            // message mode: [B, D]
    const float* xextra; int64_t Bpad;
    const float* dz1;
    int64_t B;
    float* part;                    // [grid.x, D*H]
    int n_rt;                       // row tiles of 8:  ceil(D / 8)
    int cg_max;                     // float4 column groups per column tile (max over tiles)
};

struct EpisodeArgs {
    StepArgs s;                     // episode-mode fields of the step kernel
    int32_t step0, n_steps;         // env-steps [step0, step0 + n_steps) in this launch
    float* acc; float* grad;        // AdaGrad accumulators, summed gradient (kept for qp_get_last_grad)
    float* params_rw;               // same buffer as s.params, writable
    float* part1;                   // [n_tiles, D*H] dW1 partials (one per CTA that owns a tile)
    int32_t* d_step_rw;
    int n_rt, n_ct;                 // 16-row tiles, pairs of float4 column groups
    int nwf;                        // forward warps (one per 20 hidden columns)
    long long* dbg;                 // optional [n_steps][8] clock64 stamps of CTA 0 (phase timing aid)
    int apply_in_kernel;            // 1: phase B (sum + AdaGrad) runs in the kernel; 0: exactly one env-step, the
                                    // partials are left for apply_kernel (multi-GPU: an all-reduce sits in between)
    // in-kernel gradient all-reduce over NVLink peer memory (world > 1): every rank publishes its
    // local gradient in its own exchange buffer, raises a flag in every peer's memory, then sums all
    // ranks' buffers in rank order (identical on every rank) and applies AdaGrad
    int world, rank;
    float* xg[8];                   // exchange buffers [2][P] of every rank (peer-mapped); xg[rank] is local
    unsigned int* flags[8];         // flag arrays [8] of every rank (peer-mapped): flags[r][q] = steps published by q, seen at r
    unsigned int flag_base;         // value of the flags before this launch
};

struct ApplyArgs {
    Hyper hp;
    float* params; float* acc; float* grad;
    const float* part1; int n1;     // dW1 partials [n1, D*H]
    const float* part2; int n2;     // dW2 partials [n2, 3H]
    int64_t P, P1;
    int do_reduce;                  // 1: grad = sum of partials; 0: grad already holds the (all-reduced) sum
    int do_adagrad;
    int32_t* d_step; int step_inc;
};

// ---------------------------------------------------------------------------------------------
// Philox4x32-10, counter (id_lo, id_hi, t, stream), key = seed.  Replaces the unseeded global
// scala.util.Random of QDecisionPolicyActor.scala:58, :62 with a reproducible counter-based draw.
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                       uint32_t k0, uint32_t k1, uint32_t out[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

#ifdef __CUDACC__

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gmem_src) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
// wait until at most n groups of this thread are still in flight
__device__ __forceinline__ void cp_async_wait_pending(int n) {
    switch (n) {
        case 0: asm volatile("cp.async.wait_group 0;\n" ::); break;
        case 1: asm volatile("cp.async.wait_group 1;\n" ::); break;
        case 2: asm volatile("cp.async.wait_group 2;\n" ::); break;
        case 3: asm volatile("cp.async.wait_group 3;\n" ::); break;
        case 4: asm volatile("cp.async.wait_group 4;\n" ::); break;
        case 5: asm volatile("cp.async.wait_group 5;\n" ::); break;
        case 6: asm volatile("cp.async.wait_group 6;\n" ::); break;
        case 7: asm volatile("cp.async.wait_group 7;\n" ::); break;
        default: asm volatile("cp.async.wait_group 8;\n" ::); break;
    }
}

// ---- TMA bulk copy (cp.async.bulk, SASS UBLKCP) completing on an mbarrier -----------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"((uint32_t)__cvta_generic_to_shared(bar))
                 : "memory");
}
// Waits for phase `parity` of the barrier.  Bounded: a copy that never lands traps the kernel
// instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = (uint32_t)__cvta_generic_to_shared(bar);
    for (uint32_t spins = 0;; ++spins) {
        uint32_t done;
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) return;
        if (spins > (1u << 24)) __trap();
    }
}

// first-max-wins argmax, as Tensor.argmax at QDecisionPolicyActor.scala:60, :69
__device__ __forceinline__ int argmax3(const float q[3]) {
    int k = 0;
    float m = q[0];
    if (q[1] > m) { k = 1; m = q[1]; }
    if (q[2] > m) { k = 2; }
    return k;
}

__device__ __forceinline__ float warp_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v;
}

// =============================================================================================
// K1 — step kernel.  lane <-> episode (32 per tile), warp <-> 20 hidden columns.
//   kEnvStep  : TrainerChildActor.scala:87-102 for env-step *d_step of every episode, with the
//               policy's SelectionAction (QDecisionPolicyActor.scala:56-62) and the forward /
//               TD-target part of UpdateQ (:64-71) fused in.
//   kMsgSelect: SelectionAction on explicit state rows.
//   kMsgUpdate: UpdateQ (forward of s and s', TD target, output-layer backward) on explicit rows.
// Shared memory: W1 (whole, D x Hs), W2 (Hs x float4), the transposed state tile(s), the
// cross-warp Q partials and the CTA's dW2 accumulator.
// =============================================================================================
template <int MODE>
__global__ void __launch_bounds__(kMaxStepWarps * 32, 1) step_kernel(const StepArgs a) {
    extern __shared__ __align__(16) float smem[];
    const Hyper& hp = a.hp;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int nthreads = blockDim.x;
    const int D = hp.D, H = hp.H, Hs = hp.Hs, wn = hp.wn;
    const int rows_x = (MODE == kEnvStep) ? (wn + 1) : D;

    float* W1s = smem;
    float* W2s = W1s + (size_t)D * Hs;
    float* Xs = W2s + (size_t)Hs * 4;
    float* Xn = Xs + (size_t)rows_x * kXStride;  // kMsgUpdate only
    float* qpart = (MODE == kMsgUpdate) ? (Xn + (size_t)rows_x * kXStride) : Xn;
    float* dw2acc = qpart + (size_t)nwarps * 2 * 3 * 32;
    const int n_groups = (D + 31) >> 5;          // W1 arrives in groups of 32 rows, one mbarrier each
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (((size_t)(dw2acc - smem) + (size_t)Hs * 3 + 3) & ~(size_t)3));

    // ---- stage W1 with TMA bulk copies (one per row, issued by warp 0), W2 / padding directly
    const float* W1g = a.params;
    const float* W2g = a.params + (size_t)D * H;
    if (tid == 0) {
        for (int g = 0; g < n_groups; ++g) mbar_init(bars + g, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (warp == 0) {
        if (lane == 0)
            for (int g = 0; g < n_groups; ++g)
                mbar_arrive_expect_tx(bars + g, (uint32_t)(min(32, D - (g << 5)) * H * (int)sizeof(float)));
        __syncwarp();
        if (Hs == H) {   // rows are contiguous in shared memory too: one copy per 32-row group
            if (lane < n_groups) {
                const int r0 = lane << 5;
                bulk_g2s(W1s + (size_t)r0 * Hs, W1g + (size_t)r0 * H,
                         (uint32_t)(min(32, D - r0) * H * (int)sizeof(float)), bars + lane);
            }
            for (int g = 32 + lane; g < n_groups; g += 32) {
                const int r0 = g << 5;
                bulk_g2s(W1s + (size_t)r0 * Hs, W1g + (size_t)r0 * H,
                         (uint32_t)(min(32, D - r0) * H * (int)sizeof(float)), bars + g);
            }
        } else {
            for (int r = lane; r < D; r += 32)
                bulk_g2s(W1s + (size_t)r * Hs, W1g + (size_t)r * H, (uint32_t)(H * sizeof(float)), bars + (r >> 5));
        }
    }
    if (Hs > H) {
        const int padc = Hs - H;
        for (int idx = tid; idx < D * padc; idx += nthreads) {
            const int r = idx / padc, c = H + (idx - r * padc);
            W1s[(size_t)r * Hs + c] = 0.0f;
        }
    }
    for (int j = tid; j < Hs; j += nthreads) {
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        if (j < H) { w.x = W2g[j * 3 + 0]; w.y = W2g[j * 3 + 1]; w.z = W2g[j * 3 + 2]; }
        reinterpret_cast<float4*>(W2s)[j] = w;
    }
    for (int j = tid; j < Hs * 3; j += nthreads) dw2acc[j] = 0.0f;

    const int32_t istep = (MODE == kEnvStep) ? *a.d_step : 0;
    const int64_t B = a.B;
    const int64_t n_tiles = (B + kTileEnvs - 1) / kTileEnvs;
    const int c0 = warp * kColsPerWarp;
    const float bias = hp.bias;
    bool first = true;

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t e = tile * kTileEnvs + lane;
        const bool valid = e < B;
        const int64_t ec = valid ? e : (B - 1);

        // ---- per-lane episode record (every warp keeps its own copy: no broadcast needed later)
        double budget = 0.0, share_value = 0.0;
        int32_t shares = 0;
        if (MODE == kEnvStep) {
            budget = a.budget[ec];
            shares = a.shares[ec];
            share_value = a.share_value[ec];
        }

        // ---- state tile, transposed: Xs[k][episode].  Coalesced row reads from HBM/L2.
        // TrainerChildActor.scala:89-91: prices.slice(i, i + historyDim + 1); one extra element
        // (index wn) is both the last input of s' (:144) and the new share value (:94).
        // Work items = (episode, 32-element segment); each lane knows its own episode's row start
        // and hands it to the warp by shuffle, so the loads of several items are in flight at once.
        {
            const float* my_row;
            if (MODE == kEnvStep) my_row = a.prices + (int64_t)a.series_of_env[ec] * a.T + istep;
            else my_row = a.states + ec * D;
            const float* my_row2 = (MODE == kMsgUpdate) ? a.next_states + ec * D : nullptr;
            const int segs = (rows_x + 31) >> 5;
            const int items = kTileEnvs * segs;
            for (int it = warp; it < items; it += nwarps) {
                const int el = it & 31, k = ((it >> 5) << 5) + lane;
                const float* src = reinterpret_cast<const float*>(
                    __shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(my_row), el));
                if (k < rows_x) cp_async4(Xs + k * kXStride + el, src + k);
                if (MODE == kMsgUpdate) {
                    const float* src2 = reinterpret_cast<const float*>(
                        __shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(my_row2), el));
                    if (k < rows_x) cp_async4(Xn + k * kXStride + el, src2 + k);
                }
            }
            cp_async_commit();
            cp_async_wait_pending(0);
        }
        __syncthreads();

        // ---- hidden layer: z1 = x · W1 for s and s' (QDecisionPolicyActor.scala:43), k ascending
        float acc_s[kColsPerWarp], acc_n[kColsPerWarp];
#pragma unroll
        for (int c = 0; c < kColsPerWarp; ++c) { acc_s[c] = 0.0f; acc_n[c] = 0.0f; }
        const int kmax = (MODE == kEnvStep) ? wn : D;
        float x_carry = Xs[lane];
#pragma unroll 1
        for (int kb = 0; kb < kmax; kb += 32) {
            if (first) mbar_wait(bars + (kb >> 5), 0);  // W1 row group kb/32 must have landed
            const int kend = min(kb + 32, kmax);
#pragma unroll 4
            for (int k = kb; k < kend; ++k) {
                float x_s, x_n = 0.0f;
                if (MODE == kEnvStep) {
                    x_s = x_carry;
                    x_n = Xs[(k + 1) * kXStride + lane];
                    x_carry = x_n;
                } else {
                    x_s = Xs[k * kXStride + lane];
                    if (MODE == kMsgUpdate) x_n = Xn[k * kXStride + lane];
                }
                const float4* w4 = reinterpret_cast<const float4*>(W1s + (size_t)k * Hs + c0);
#pragma unroll
                for (int j = 0; j < kColsPerWarp / 4; ++j) {
                    const float4 w = w4[j];
                    acc_s[4 * j + 0] = fmaf(x_s, w.x, acc_s[4 * j + 0]);
                    acc_s[4 * j + 1] = fmaf(x_s, w.y, acc_s[4 * j + 1]);
                    acc_s[4 * j + 2] = fmaf(x_s, w.z, acc_s[4 * j + 2]);
                    acc_s[4 * j + 3] = fmaf(x_s, w.w, acc_s[4 * j + 3]);
                    if (MODE != kMsgSelect) {
                        acc_n[4 * j + 0] = fmaf(x_n, w.x, acc_n[4 * j + 0]);
                        acc_n[4 * j + 1] = fmaf(x_n, w.y, acc_n[4 * j + 1]);
                        acc_n[4 * j + 2] = fmaf(x_n, w.z, acc_n[4 * j + 2]);
                        acc_n[4 * j + 3] = fmaf(x_n, w.w, acc_n[4 * j + 3]);
                    }
                }
            }
        }
        if (first && MODE == kEnvStep) mbar_wait(bars + n_groups - 1, 0);  // budget/shares rows: last group
        first = false;

        const float* w_budget = W1s + (size_t)wn * Hs + c0;
        const float* w_shares = W1s + (size_t)(wn + 1) * Hs + c0;
        if (MODE == kEnvStep) {
            // last two inputs of s: (float)budget, (float)shares (TrainerChildActor.scala:90)
            const float xb = __double2float_rn(budget);
            const float xs = (float)shares;
#pragma unroll
            for (int c = 0; c < kColsPerWarp; ++c) {
                acc_s[c] = fmaf(xb, w_budget[c], acc_s[c]);
                acc_s[c] = fmaf(xs, w_shares[c], acc_s[c]);
            }
        }
        // h = relu(z1 + 0.1) (:43); partial of q = h · W2 over this warp's columns (:47)
        float pq[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < kColsPerWarp; ++c) {
            const float v = acc_s[c] + bias;
            acc_s[c] = v > 0.0f ? v : 0.0f;
            const float4 w2 = reinterpret_cast<const float4*>(W2s)[c0 + c];
            pq[0] = fmaf(acc_s[c], w2.x, pq[0]);
            pq[1] = fmaf(acc_s[c], w2.y, pq[1]);
            pq[2] = fmaf(acc_s[c], w2.z, pq[2]);
        }
#pragma unroll
        for (int x = 0; x < 3; ++x) qpart[((warp * 2 + 0) * 3 + x) * 32 + lane] = pq[x];
        __syncthreads();

        float q_s[3];
#pragma unroll
        for (int x = 0; x < 3; ++x) {
            float s = 0.0f;
            for (int w = 0; w < nwarps; ++w) s += qpart[((w * 2 + 0) * 3 + x) * 32 + lane];
            s += bias;
            q_s[x] = s > 0.0f ? s : 0.0f;
        }

        // ---- SelectionAction (QDecisionPolicyActor.scala:56-62)
        int action = 0, explored = 0;
        if (MODE != kMsgUpdate) {
            const float step_f = (MODE == kEnvStep) ? (float)istep : a.steps[ec];
            const uint64_t env_id = (MODE == kEnvStep) ? (uint64_t)(a.env_id_base + e)
                                                       : (uint64_t)(a.env_ids ? a.env_ids[ec] : e);
            const uint32_t tctr = (MODE == kEnvStep) ? (uint32_t)istep : (uint32_t)step_f;
            uint32_t r[4];
            philox4x32_10((uint32_t)env_id, (uint32_t)(env_id >> 32), tctr, 1u,
                          (uint32_t)hp.seed, (uint32_t)(hp.seed >> 32), r);
            const float u = (float)(r[0] >> 8) * (1.0f / 16777216.0f);   // nextFloat(): 24 bits
            float thr = __fdiv_rn(step_f, hp.step_div);
            if (hp.epsilon < thr) thr = hp.epsilon;                       // Seq(epsilon, step/1000f).min
            if (u < thr) {
                action = argmax3(q_s);
            } else {
                explored = 1;
                action = (int)(((uint64_t)r[1] * 3u) >> 32);             // nextInt(3)
            }
        }
        if (MODE == kMsgSelect) {
            if (warp == 0 && valid) {
                a.actions_out[e] = action;
                if (a.q_out) {
                    a.q_out[e * 3 + 0] = q_s[0]; a.q_out[e * 3 + 1] = q_s[1]; a.q_out[e * 3 + 2] = q_s[2];
                }
            }
            __syncthreads();
            continue;
        }

        // ---- decision, reward, next state (TrainerChildActor.scala:92-96, :118-123, :136-146)
        float r32 = 0.0f;
        double nb = 0.0, p_new = 0.0;
        int32_t ns = 0;
        int a_eff = 2;
        if (MODE == kEnvStep) {
            p_new = (double)Xs[wn * kXStride + lane];                               // :94
            const double cur = __dadd_rn(budget, __dmul_rn((double)shares, share_value));  // :92
            const double base_b = hp.bookkeeping == 0 ? a.B0 : budget;
            const int32_t base_n = hp.bookkeeping == 0 ? a.N0 : shares;
            nb = base_b; ns = base_n;
            if (action == 0 && base_b >= p_new) { nb = __dsub_rn(base_b, p_new); ns = base_n + 1; a_eff = 0; }
            else if (action == 1 && base_n > 0) { nb = __dadd_rn(base_b, p_new); ns = base_n - 1; a_eff = 1; }
            const double newp = __dadd_rn(nb, __dmul_rn((double)ns, p_new));        // :141
            r32 = __double2float_rn(__dsub_rn(newp, cur));                          // :142, :158
            const float xnb = __double2float_rn(nb);
            const float xns = (float)ns;
#pragma unroll
            for (int c = 0; c < kColsPerWarp; ++c) {
                acc_n[c] = fmaf(xnb, w_budget[c], acc_n[c]);
                acc_n[c] = fmaf(xns, w_shares[c], acc_n[c]);
            }
        } else {
            r32 = a.rewards[ec];
        }
        pq[0] = pq[1] = pq[2] = 0.0f;
#pragma unroll
        for (int c = 0; c < kColsPerWarp; ++c) {
            float v = acc_n[c] + bias;
            v = v > 0.0f ? v : 0.0f;
            const float4 w2 = reinterpret_cast<const float4*>(W2s)[c0 + c];
            pq[0] = fmaf(v, w2.x, pq[0]);
            pq[1] = fmaf(v, w2.y, pq[1]);
            pq[2] = fmaf(v, w2.z, pq[2]);
        }
#pragma unroll
        for (int x = 0; x < 3; ++x) qpart[((warp * 2 + 1) * 3 + x) * 32 + lane] = pq[x];
        __syncthreads();

        // ---- UpdateQ target and output-layer backward (QDecisionPolicyActor.scala:67-73, :49)
        float q_n[3];
#pragma unroll
        for (int x = 0; x < 3; ++x) {
            float s = 0.0f;
            for (int w = 0; w < nwarps; ++w) s += qpart[((w * 2 + 1) * 3 + x) * 32 + lane];
            s += bias;
            q_n[x] = s > 0.0f ? s : 0.0f;
        }
        const int kn = argmax3(q_n);                                                 // :69
        const float qn_k = kn == 0 ? q_n[0] : (kn == 1 ? q_n[1] : q_n[2]);
        const float qs_k = kn == 0 ? q_s[0] : (kn == 1 ? q_s[1] : q_s[2]);
        const float target = __fadd_rn(r32, __fmul_rn(hp.gamma, qn_k));              // :70
        // d/dq of square(y - q) with y[k] = target and y[j] = q[j] elsewhere; ReluGrad on q
        const float dq = -__fmul_rn(2.0f, __fsub_rn(target, qs_k));
        const float dz2 = (valid && qs_k > 0.0f) ? dq : 0.0f;

        float dz1v[kColsPerWarp];
#pragma unroll
        for (int c = 0; c < kColsPerWarp; ++c) {
            const float w2k = W2s[(c0 + c) * 4 + kn];
            dz1v[c] = acc_s[c] > 0.0f ? __fmul_rn(dz2, w2k) : 0.0f;
        }
        if (valid) {
            float4* dst = reinterpret_cast<float4*>(a.dz1 + e * H + c0);
#pragma unroll
            for (int j = 0; j < kColsPerWarp / 4; ++j)
                if (c0 + 4 * j < H) dst[j] = make_float4(dz1v[4 * j], dz1v[4 * j + 1], dz1v[4 * j + 2], dz1v[4 * j + 3]);
        }
        // dW2[:, k] += h * dz2, reduced over the tile's episodes with a fixed shuffle tree
#pragma unroll
        for (int c = 0; c < kColsPerWarp; ++c) {
            const float hv = __fmul_rn(acc_s[c], dz2);
#pragma unroll
            for (int x = 0; x < 3; ++x) {
                const float s = warp_sum(kn == x ? hv : 0.0f);
                if (lane == 0) dw2acc[(c0 + c) * 3 + x] += s;
            }
        }

        if (warp == 0 && valid) {
            if (MODE == kEnvStep) {
                a.xextra[e] = __double2float_rn(budget);
                a.xextra[a.Bpad + e] = (float)shares;
                a.budget[e] = nb;                     // carry (TrainerChildActor.scala:99-102)
                a.shares[e] = ns;
                a.share_value[e] = p_new;
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
            } else if (a.q_out) {  // kMsgUpdate: optional q(s) for inspection
                a.q_out[e * 3 + 0] = q_s[0]; a.q_out[e * 3 + 1] = q_s[1]; a.q_out[e * 3 + 2] = q_s[2];
            }
        }
        __syncthreads();
    }

    __syncthreads();
    if (MODE != kMsgSelect) {
        float* dst = a.dw2_part + (size_t)blockIdx.x * (3 * H);
        for (int j = tid; j < 3 * H; j += nthreads) dst[j] = dw2acc[j];
    }
}

// =============================================================================================
// K2 — dW1 = X^T · dZ1 (the MatMul gradient of QDecisionPolicyActor.scala:43 w.r.t. weight1) on a
// 2-D grid: blockIdx.x = split s over episodes [B*s/S, B*(s+1)/S), blockIdx.y = column tile t of
// about 52 hidden columns.  One D x (tile) partial per CTA, so only S (= #SM / T) partials exist —
// the 2-D split trades a T-fold re-read of the L2-resident price windows for a T-fold cut of the
// partial traffic.  Thread (rt, ct) keeps an 8 x 4 register tile; per episode it reads 2 + 1
// float4 from shared memory for 32 FMAs.  Episodes stream through a 2-stage cp.async pipeline.
// =============================================================================================
template <int MODE>
__global__ void __launch_bounds__(kDw1MaxThreads, 1) dw1_kernel(const Dw1Args a) {
    extern __shared__ __align__(16) float smem[];
    const Hyper& hp = a.hp;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int nthreads = blockDim.x;
    const int D = hp.D, H = hp.H, wn = hp.wn;
    const int H4 = H >> 2;
    const int T = gridDim.y, t = blockIdx.y;
    const int cg0 = (H4 * t) / T, width = (H4 * (t + 1)) / T - cg0;     // this tile's float4 column groups
    const int Dp = a.n_rt * kDw1Rows, Wp = a.cg_max * kDw1Cols;
    const size_t stage_floats = (size_t)kDw1ChunkEnvs * (Dp + Wp);
    const int32_t istep = (MODE == kEnvStep) ? *a.d_step : 0;

    const int64_t S = gridDim.x;
    const int64_t e0 = a.B * (int64_t)blockIdx.x / S, e1 = a.B * ((int64_t)blockIdx.x + 1) / S;
    const int rt = tid / a.cg_max, ct = tid - rt * a.cg_max;
    const bool active = rt < a.n_rt && ct < width;

    float acc[kDw1Rows][kDw1Cols];
#pragma unroll
    for (int r = 0; r < kDw1Rows; ++r)
#pragma unroll
        for (int c = 0; c < kDw1Cols; ++c) acc[r][c] = 0.0f;

    // the states fed with trainOp (QDecisionPolicyActor.scala:72): price window + budget + shares
    auto prefetch = [&](int64_t eb, int ne, float* Xc, float* Zc) {
        for (int el = warp; el < ne; el += nwarps) {
            const int64_t ee = eb + el;
            float* dst = Xc + el * Dp;
            if (MODE == kEnvStep) {
                const float* src = a.prices + (int64_t)a.series_of_env[ee] * a.T + istep;
                for (int k = lane; k < Dp; k += 32) {
                    if (k < wn) cp_async4(dst + k, src + k);       // rows are only 4-byte aligned
                    else dst[k] = (k == wn) ? a.xextra[ee] : ((k == wn + 1) ? a.xextra[a.Bpad + ee] : 0.0f);
                }
            } else {
                const float* src = a.states + ee * D;
                for (int k = lane; k < Dp; k += 32) {
                    if (k < D) cp_async4(dst + k, src + k);
                    else dst[k] = 0.0f;
                }
            }
        }
        for (int idx = tid; idx < ne * a.cg_max; idx += nthreads) {
            const int el = idx / a.cg_max, c = idx - el * a.cg_max;
            float* dst = Zc + el * Wp + c * 4;
            if (c < width) cp_async16(dst, a.dz1 + (eb + el) * H + (cg0 + c) * 4);
            else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    };

    const int n_chunks = (int)((e1 - e0 + kDw1ChunkEnvs - 1) / kDw1ChunkEnvs);
    if (n_chunks > 0) prefetch(e0, (int)min((int64_t)kDw1ChunkEnvs, e1 - e0), smem, smem + (size_t)kDw1ChunkEnvs * Dp);
    cp_async_commit();
    for (int c = 0; c < n_chunks; ++c) {
        const int64_t eb = e0 + (int64_t)c * kDw1ChunkEnvs;
        const int ne = (int)min((int64_t)kDw1ChunkEnvs, e1 - eb);
        float* Xc = smem + (size_t)(c & 1) * stage_floats;
        float* Zc = Xc + (size_t)kDw1ChunkEnvs * Dp;
        if (c + 1 < n_chunks) {
            float* Xn = smem + (size_t)((c + 1) & 1) * stage_floats;
            prefetch(eb + kDw1ChunkEnvs, (int)min((int64_t)kDw1ChunkEnvs, e1 - eb - kDw1ChunkEnvs), Xn,
                     Xn + (size_t)kDw1ChunkEnvs * Dp);
        }
        cp_async_commit();
        cp_async_wait_pending(1);      // everything but the group just committed has landed
        __syncthreads();
        if (active) {
            const float4* xp = reinterpret_cast<const float4*>(Xc + rt * kDw1Rows);
            const float4* zp = reinterpret_cast<const float4*>(Zc) + ct;
            const int xs4 = Dp >> 2, zs4 = Wp >> 2;
#pragma unroll 4
            for (int el = 0; el < ne; ++el) {
                const float4 x0 = xp[el * xs4], x1 = xp[el * xs4 + 1], z = zp[el * zs4];
                const float x[kDw1Rows] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
                const float zz[kDw1Cols] = {z.x, z.y, z.z, z.w};
#pragma unroll
                for (int r = 0; r < kDw1Rows; ++r)
#pragma unroll
                    for (int cc = 0; cc < kDw1Cols; ++cc) acc[r][cc] = fmaf(x[r], zz[cc], acc[r][cc]);
            }
        }
        __syncthreads();               // the stage just read is refilled by the next iteration's prefetch
    }
    cp_async_wait_pending(0);

    if (active) {
        float* out = a.part + (size_t)blockIdx.x * ((size_t)D * H);
#pragma unroll
        for (int r = 0; r < kDw1Rows; ++r) {
            const int row = rt * kDw1Rows + r;
            if (row < D)
                reinterpret_cast<float4*>(out + (size_t)row * H)[cg0 + ct] = make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
        }
    }
}

// =============================================================================================
// K2, large-batch form — used when every CTA owns at least a few hundred episodes (configs[2]:
// 131 072 episodes per GPU).  One split per SM and the whole D x H partial per CTA: with that many
// episodes per split the partial traffic is negligible, and the bigger 16 x 8 register tile needs
// only 24 shared-memory wavefronts per 128 FMAs (the 8 x 4 tile of the 2-D grid needs ~9 per 32).
// Same 2-stage cp.async pipeline as above.
// =============================================================================================
constexpr int kDw1BigRows = 16, kDw1BigCols = 8;

template <int MODE>
__global__ void __launch_bounds__(kDw1MaxThreads, 1) dw1_big_kernel(const Dw1Args a) {
    extern __shared__ __align__(16) float smem[];
    const Hyper& hp = a.hp;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int nthreads = blockDim.x;
    const int D = hp.D, H = hp.H, wn = hp.wn;
    const int H4 = H >> 2;
    const int n_rt = a.n_rt, n_ct = a.cg_max;                 // here: 16-row tiles, pairs of float4 column groups
    const int Dp = n_rt * kDw1BigRows, Hp = n_ct * kDw1BigCols;
    const size_t stage_floats = (size_t)kDw1ChunkEnvs * (Dp + Hp);
    const int32_t istep = (MODE == kEnvStep) ? *a.d_step : 0;
    const int64_t S = gridDim.x;
    const int64_t e0 = a.B * (int64_t)blockIdx.x / S, e1 = a.B * ((int64_t)blockIdx.x + 1) / S;
    const int rt = tid / n_ct, ct = tid - rt * n_ct;
    const bool active = rt < n_rt;

    float acc[kDw1BigRows][kDw1BigCols];
#pragma unroll
    for (int r = 0; r < kDw1BigRows; ++r)
#pragma unroll
        for (int c = 0; c < kDw1BigCols; ++c) acc[r][c] = 0.0f;

    auto prefetch = [&](int64_t eb, int ne, float* Xc, float* Zc) {
        for (int el = warp; el < ne; el += nwarps) {
            const int64_t ee = eb + el;
            float* dst = Xc + el * Dp;
            if (MODE == kEnvStep) {
                const float* src = a.prices + (int64_t)a.series_of_env[ee] * a.T + istep;
                for (int k = lane; k < Dp; k += 32) {
                    if (k < wn) cp_async4(dst + k, src + k);
                    else dst[k] = (k == wn) ? a.xextra[ee] : ((k == wn + 1) ? a.xextra[a.Bpad + ee] : 0.0f);
                }
            } else {
                const float* src = a.states + ee * D;
                for (int k = lane; k < Dp; k += 32) {
                    if (k < D) cp_async4(dst + k, src + k);
                    else dst[k] = 0.0f;
                }
            }
        }
        const int Hp4 = Hp >> 2;
        for (int idx = tid; idx < ne * Hp4; idx += nthreads) {
            const int el = idx / Hp4, c4 = idx - el * Hp4;
            float* dst = Zc + el * Hp + c4 * 4;
            if (c4 < H4) cp_async16(dst, a.dz1 + (eb + el) * H + c4 * 4);
            else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    };

    const int n_chunks = (int)((e1 - e0 + kDw1ChunkEnvs - 1) / kDw1ChunkEnvs);
    if (n_chunks > 0) prefetch(e0, (int)min((int64_t)kDw1ChunkEnvs, e1 - e0), smem, smem + (size_t)kDw1ChunkEnvs * Dp);
    cp_async_commit();
    for (int c = 0; c < n_chunks; ++c) {
        const int64_t eb = e0 + (int64_t)c * kDw1ChunkEnvs;
        const int ne = (int)min((int64_t)kDw1ChunkEnvs, e1 - eb);
        float* Xc = smem + (size_t)(c & 1) * stage_floats;
        float* Zc = Xc + (size_t)kDw1ChunkEnvs * Dp;
        if (c + 1 < n_chunks) {
            float* Xn = smem + (size_t)((c + 1) & 1) * stage_floats;
            prefetch(eb + kDw1ChunkEnvs, (int)min((int64_t)kDw1ChunkEnvs, e1 - eb - kDw1ChunkEnvs), Xn,
                     Xn + (size_t)kDw1ChunkEnvs * Dp);
        }
        cp_async_commit();
        cp_async_wait_pending(1);
        __syncthreads();
        if (active) {
            const float4* xp = reinterpret_cast<const float4*>(Xc + rt * kDw1BigRows);
            const float4* zp0 = reinterpret_cast<const float4*>(Zc) + ct;
            const float4* zp1 = reinterpret_cast<const float4*>(Zc) + ct + n_ct;
            const int xs4 = Dp >> 2, zs4 = Hp >> 2;
#pragma unroll 2
            for (int el = 0; el < ne; ++el) {
                float x[kDw1BigRows];
#pragma unroll
                for (int j = 0; j < kDw1BigRows / 4; ++j) {
                    const float4 v = xp[el * xs4 + j];
                    x[4 * j] = v.x; x[4 * j + 1] = v.y; x[4 * j + 2] = v.z; x[4 * j + 3] = v.w;
                }
                const float4 za = zp0[el * zs4], zb = zp1[el * zs4];
                const float z[kDw1BigCols] = {za.x, za.y, za.z, za.w, zb.x, zb.y, zb.z, zb.w};
#pragma unroll
                for (int r = 0; r < kDw1BigRows; ++r)
#pragma unroll
                    for (int cc = 0; cc < kDw1BigCols; ++cc) acc[r][cc] = fmaf(x[r], z[cc], acc[r][cc]);
            }
        }
        __syncthreads();
    }
    cp_async_wait_pending(0);

    if (active) {
        float* out = a.part + (size_t)blockIdx.x * ((size_t)D * H);
#pragma unroll
        for (int r = 0; r < kDw1BigRows; ++r) {
            const int row = rt * kDw1BigRows + r;
            if (row < D) {
                if (ct < H4)
                    reinterpret_cast<float4*>(out + (size_t)row * H)[ct] = make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
                if (ct + n_ct < H4)
                    reinterpret_cast<float4*>(out + (size_t)row * H)[ct + n_ct] = make_float4(acc[r][4], acc[r][5], acc[r][6], acc[r][7]);
            }
        }
    }
}

// =============================================================================================
// K3 — gradient reduction + AdaGrad.  Eight lanes per parameter sum the partials (interleaved,
// then a fixed xor tree); the leader applies ResourceApplyAdagrad as trainOp does at
// QDecisionPolicyActor.scala:50, :73:   accum += g*g;  var -= (g*lr) * (1/sqrt(accum)).
// With do_reduce = 0 it applies a gradient that is already summed (after the all-reduce).
// =============================================================================================
__global__ void __launch_bounds__(256) apply_kernel(const ApplyArgs a) {
    const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t p = gid / kApplyLanes;
    const int sub = (int)(gid - p * kApplyLanes);
    if (gid == 0 && a.d_step && a.step_inc) *a.d_step += a.step_inc;
    const bool in_range = p < a.P;  // whole 8-lane groups share p, so shuffles below stay converged per group
    float g = 0.0f;
    if (a.do_reduce) {
        if (in_range) {
            const float* src; int n; int64_t stride, off;
            if (p < a.P1) { src = a.part1; n = a.n1; stride = a.P1; off = p; }
            else { src = a.part2; n = a.n2; stride = a.P - a.P1; off = p - a.P1; }
            float s = 0.0f;
            for (int j = sub; j < n; j += kApplyLanes) s += src[(int64_t)j * stride + off];
            g = s;
        }
        g += __shfl_xor_sync(0xffffffffu, g, 1);
        g += __shfl_xor_sync(0xffffffffu, g, 2);
        g += __shfl_xor_sync(0xffffffffu, g, 4);
        if (in_range && sub == 0) a.grad[p] = g;
    } else if (in_range && sub == 0) {
        g = a.grad[p];
    }
    if (a.do_adagrad && in_range && sub == 0) {
        const float acc = __fadd_rn(a.acc[p], __fmul_rn(g, g));
        a.acc[p] = acc;
        const float upd = __fmul_rn(__fmul_rn(g, a.hp.lr), __fdiv_rn(1.0f, __fsqrt_rn(acc)));
        a.params[p] = __fsub_rn(a.params[p], upd);
    }
}

// budget + shares * shareValue (TrainerChildActor.scala:68)
__global__ void portfolio_kernel(const double* budget, const int32_t* shares, const double* sv,
                                 int64_t n, double* out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < n) out[e] = __dadd_rn(budget[e], __dmul_rn((double)shares[e], sv[e]));
}

// Deterministic fp64 block sums of (x - shift)^pow, pow in {1, 2}; finished on the host side of
// the library over <= 1024 block results (TrainerRouterActor.scala:149-151).
__global__ void __launch_bounds__(256) moment_kernel(const double* x, int64_t n, double shift, int pw,
                                                     double* block_out) {
    __shared__ double sh[256];
    double s = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double d = x[i] - shift;
        s += (pw == 2) ? d * d : d;
    }
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) block_out[blockIdx.x] = sh[0];
}

__global__ void fill_kernel(float* p, int64_t n, float v) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

__global__ void env_init_kernel(double* budget, int32_t* shares, double* sv, int64_t n, double B0, int32_t N0) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < n) { budget[e] = B0; shares[e] = N0; sv[e] = 0.0; }   // TrainerChildActor.scala:84-86
}

// =============================================================================================
// Persistent episode kernel — the latency path for up to 32 x #SM episodes (configs 1, 2, 5).
// One cooperative launch runs n_steps env-steps: CTA c owns episodes [32c, 32c+32) for the whole
// launch and keeps their price window in shared memory, shifting it by one row per step (one new
// price per episode per step instead of re-staging 202), and their carry (budget, shares, share
// value) in registers.  Per step:
//   phase A (per CTA)  TMA bulk load of the current W1 into shared memory; forward of s and s',
//                      select, book-keeping, TD target (as step_kernel<kEnvStep>); then the dW1
//                      contraction of its 32 episodes straight from the shared-memory window
//                      (16 x 8 register tiles) -> one D x H partial per CTA
//   grid.sync
//   phase B (all CTAs) fixed-order sum of the partials + AdaGrad on a slice of the parameters
//   grid.sync
// Arithmetic and summation orders inside a tile are those of the three-kernel path; the only
// difference is the split of the gradient sum (per 32-episode tile instead of per S-split).
// =============================================================================================
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;\n" ::: "memory"); }

__global__ void __launch_bounds__(kEpMaxThreads, 1) episode_kernel(const EpisodeArgs ea) {
    namespace cg = cooperative_groups;
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) float smem[];
    const StepArgs& a = ea.s;
    const Hyper& hp = a.hp;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nthreads = blockDim.x;
    const int D = hp.D, H = hp.H, Hs = hp.Hs, wn = hp.wn, nwf = ea.nwf;
    const int H4 = H >> 2;
    const int Dp = ea.n_rt * kEpRows;            // rows of the window tile incl. padding (>= wn + 1)
    const int Hp = ea.n_ct * kEpCols;            // row stride of the dz1 tile

    float* W1s = smem;                            // [D][Hs]; after the forward: Z tile [32][Hp]
    float* W2s = W1s + (size_t)D * Hs;
    float* Xs = W2s + (size_t)Hs * 4;             // [Dp][33] transposed price window, row k = p[i + k]
    float* XE = Xs + (size_t)Dp * kXStride;       // [2][32] (float)budget, (float)shares of state s
    float* qpart = XE + 64;
    float* dw2acc = qpart + (size_t)nwf * 2 * 3 * 32;
    const int n_groups = (D + 31) >> 5;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (((size_t)(dw2acc - smem) + (size_t)Hs * 3 + 3) & ~(size_t)3));
    float* Zs = W1s;

    const int64_t B = a.B;
    const int64_t n_tiles = (B + kTileEnvs - 1) / kTileEnvs;
    const bool has_tile = (int64_t)blockIdx.x < n_tiles;
    const int64_t e = (int64_t)blockIdx.x * kTileEnvs + lane;
    const bool valid = has_tile && e < B;
    const int64_t ec = has_tile ? (valid ? e : B - 1) : 0;
    const bool fwd_warp = warp < nwf;
    const int c0 = warp * kColsPerWarp;
    const float bias = hp.bias;
    const float* W1g = a.params;
    const float* W2g = a.params + (size_t)D * H;

    if (tid == 0) {
        for (int g = 0; g < n_groups; ++g) mbar_init(bars + g, 1);
        mbar_fence_init();
    }
    // carry of the fold (TrainerChildActor.scala:84-86, :99-102), one copy per warp
    double budget = a.budget[ec], share_value = a.share_value[ec];
    int32_t shares = a.shares[ec];
    const float* my_row = a.prices + (int64_t)a.series_of_env[ec] * a.T;
    // initial window: rows 0..wn of step0; rows beyond stay zero
    for (int idx = tid; idx < Dp * kXStride; idx += nthreads) Xs[idx] = 0.0f;
    __syncthreads();
    if (has_tile) {
        const int segs = (wn + 1 + 31) >> 5, items = kTileEnvs * segs, nw = nthreads >> 5;
        for (int it = warp; it < items; it += nw) {
            const int el = it & 31, k = ((it >> 5) << 5) + lane;
            const float* src = reinterpret_cast<const float*>(
                __shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(my_row), el)) + ea.step0;
            if (k <= wn) cp_async4(Xs + k * kXStride + el, src + k);
        }
        cp_async_commit();
        cp_async_wait_pending(0);
    }
    __syncthreads();

    for (int32_t sidx = 0; sidx < ea.n_steps; ++sidx) {
        const int32_t istep = ea.step0 + sidx;
        const uint32_t parity = (uint32_t)sidx & 1u;
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 0] = clock64();
        // ---- current weights into shared memory (written by phase B of the previous step)
        if (has_tile) {
            if (warp == 0) {
                fence_proxy_async();   // generic-proxy writes (phase B, Z tile) before async-proxy copies
                if (lane == 0)
                    for (int g = 0; g < n_groups; ++g)
                        mbar_arrive_expect_tx(bars + g, (uint32_t)(min(32, D - (g << 5)) * H * (int)sizeof(float)));
                __syncwarp();
                if (Hs == H) {
                    for (int g = lane; g < n_groups; g += 32) {
                        const int r0 = g << 5;
                        bulk_g2s(W1s + (size_t)r0 * Hs, W1g + (size_t)r0 * H,
                                 (uint32_t)(min(32, D - r0) * H * (int)sizeof(float)), bars + g);
                    }
                } else {
                    for (int r = lane; r < D; r += 32)
                        bulk_g2s(W1s + (size_t)r * Hs, W1g + (size_t)r * H, (uint32_t)(H * sizeof(float)), bars + (r >> 5));
                }
            }
            if (Hs > H) {
                const int padc = Hs - H;
                for (int idx = tid; idx < D * padc; idx += nthreads) {
                    const int r = idx / padc, c = H + (idx - r * padc);
                    W1s[(size_t)r * Hs + c] = 0.0f;
                }
            }
            for (int j = tid; j < Hs; j += nthreads) {
                float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
                if (j < H) { w.x = __ldcg(W2g + j * 3 + 0); w.y = __ldcg(W2g + j * 3 + 1); w.z = __ldcg(W2g + j * 3 + 2); }
                reinterpret_cast<float4*>(W2s)[j] = w;
            }
            for (int j = tid; j < Hs * 3; j += nthreads) dw2acc[j] = 0.0f;
        }
        // the price that enters the window at the end of this step (row wn + 1 of this step)
        float p_next = 0.0f;
        if (has_tile && warp == 0 && (int64_t)istep + wn + 1 < a.T) p_next = __ldg(my_row + istep + wn + 1);
        __syncthreads();

        float q_s[3] = {0.f, 0.f, 0.f}, q_n[3] = {0.f, 0.f, 0.f};
        float acc_s[kColsPerWarp], acc_n[kColsPerWarp];
        int action = 0, explored = 0, a_eff = 2, kn = 0;
        float r32 = 0.0f, target = 0.0f, dz2 = 0.0f;
        double nb = budget, p_new = 0.0;
        int32_t ns = shares;

        // ---- hidden layer for s and s' (QDecisionPolicyActor.scala:43), k ascending
        if (has_tile && fwd_warp) {
#pragma unroll
            for (int c = 0; c < kColsPerWarp; ++c) { acc_s[c] = 0.0f; acc_n[c] = 0.0f; }
            float x_carry = Xs[lane];
#pragma unroll 1
            for (int kb = 0; kb < wn; kb += 32) {
                mbar_wait(bars + (kb >> 5), parity);
                const int kend = min(kb + 32, wn);
#pragma unroll 4
                for (int k = kb; k < kend; ++k) {
                    const float x_s = x_carry;
                    const float x_n = Xs[(k + 1) * kXStride + lane];
                    x_carry = x_n;
                    const float4* w4 = reinterpret_cast<const float4*>(W1s + (size_t)k * Hs + c0);
#pragma unroll
                    for (int j = 0; j < kColsPerWarp / 4; ++j) {
                        const float4 w = w4[j];
                        acc_s[4 * j + 0] = fmaf(x_s, w.x, acc_s[4 * j + 0]);
                        acc_s[4 * j + 1] = fmaf(x_s, w.y, acc_s[4 * j + 1]);
                        acc_s[4 * j + 2] = fmaf(x_s, w.z, acc_s[4 * j + 2]);
                        acc_s[4 * j + 3] = fmaf(x_s, w.w, acc_s[4 * j + 3]);
                        acc_n[4 * j + 0] = fmaf(x_n, w.x, acc_n[4 * j + 0]);
                        acc_n[4 * j + 1] = fmaf(x_n, w.y, acc_n[4 * j + 1]);
                        acc_n[4 * j + 2] = fmaf(x_n, w.z, acc_n[4 * j + 2]);
                        acc_n[4 * j + 3] = fmaf(x_n, w.w, acc_n[4 * j + 3]);
                    }
                }
            }
            mbar_wait(bars + n_groups - 1, parity);
            const float* w_budget = W1s + (size_t)wn * Hs + c0;
            const float* w_shares = W1s + (size_t)(wn + 1) * Hs + c0;
            const float xb = __double2float_rn(budget);
            const float xs = (float)shares;
            float pq[3] = {0.f, 0.f, 0.f};
#pragma unroll
            for (int c = 0; c < kColsPerWarp; ++c) {
                acc_s[c] = fmaf(xb, w_budget[c], acc_s[c]);
                acc_s[c] = fmaf(xs, w_shares[c], acc_s[c]);
                const float v = acc_s[c] + bias;
                acc_s[c] = v > 0.0f ? v : 0.0f;
                const float4 w2 = reinterpret_cast<const float4*>(W2s)[c0 + c];
                pq[0] = fmaf(acc_s[c], w2.x, pq[0]);
                pq[1] = fmaf(acc_s[c], w2.y, pq[1]);
                pq[2] = fmaf(acc_s[c], w2.z, pq[2]);
            }
#pragma unroll
            for (int x = 0; x < 3; ++x) qpart[((warp * 2 + 0) * 3 + x) * 32 + lane] = pq[x];
        }
        __syncthreads();

        if (has_tile && fwd_warp) {
#pragma unroll
            for (int x = 0; x < 3; ++x) {
                float s = 0.0f;
                for (int w = 0; w < nwf; ++w) s += qpart[((w * 2 + 0) * 3 + x) * 32 + lane];
                s += bias;
                q_s[x] = s > 0.0f ? s : 0.0f;
            }
            // SelectionAction (QDecisionPolicyActor.scala:56-62)
            const uint64_t env_id = (uint64_t)(a.env_id_base + e);
            uint32_t r[4];
            philox4x32_10((uint32_t)env_id, (uint32_t)(env_id >> 32), (uint32_t)istep, 1u,
                          (uint32_t)hp.seed, (uint32_t)(hp.seed >> 32), r);
            const float u = (float)(r[0] >> 8) * (1.0f / 16777216.0f);
            float thr = __fdiv_rn((float)istep, hp.step_div);
            if (hp.epsilon < thr) thr = hp.epsilon;
            if (u < thr) action = argmax3(q_s);
            else { explored = 1; action = (int)(((uint64_t)r[1] * 3u) >> 32); }
            // decision, reward, next state (TrainerChildActor.scala:92-96, :118-123, :136-146)
            p_new = (double)Xs[wn * kXStride + lane];
            const double cur = __dadd_rn(budget, __dmul_rn((double)shares, share_value));
            const double base_b = hp.bookkeeping == 0 ? a.B0 : budget;
            const int32_t base_n = hp.bookkeeping == 0 ? a.N0 : shares;
            nb = base_b; ns = base_n;
            if (action == 0 && base_b >= p_new) { nb = __dsub_rn(base_b, p_new); ns = base_n + 1; a_eff = 0; }
            else if (action == 1 && base_n > 0) { nb = __dadd_rn(base_b, p_new); ns = base_n - 1; a_eff = 1; }
            const double newp = __dadd_rn(nb, __dmul_rn((double)ns, p_new));
            r32 = __double2float_rn(__dsub_rn(newp, cur));
            const float xnb = __double2float_rn(nb);
            const float xns = (float)ns;
            const float* w_budget = W1s + (size_t)wn * Hs + c0;
            const float* w_shares = W1s + (size_t)(wn + 1) * Hs + c0;
            float pq[3] = {0.f, 0.f, 0.f};
#pragma unroll
            for (int c = 0; c < kColsPerWarp; ++c) {
                acc_n[c] = fmaf(xnb, w_budget[c], acc_n[c]);
                acc_n[c] = fmaf(xns, w_shares[c], acc_n[c]);
                float v = acc_n[c] + bias;
                v = v > 0.0f ? v : 0.0f;
                const float4 w2 = reinterpret_cast<const float4*>(W2s)[c0 + c];
                pq[0] = fmaf(v, w2.x, pq[0]);
                pq[1] = fmaf(v, w2.y, pq[1]);
                pq[2] = fmaf(v, w2.z, pq[2]);
            }
#pragma unroll
            for (int x = 0; x < 3; ++x) qpart[((warp * 2 + 1) * 3 + x) * 32 + lane] = pq[x];
        }
        __syncthreads();     // after this barrier nobody reads W1s any more: the Z tile may overwrite it
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 1] = clock64();

        if (has_tile && fwd_warp) {
            // UpdateQ target and output-layer backward (QDecisionPolicyActor.scala:67-73, :49)
#pragma unroll
            for (int x = 0; x < 3; ++x) {
                float s = 0.0f;
                for (int w = 0; w < nwf; ++w) s += qpart[((w * 2 + 1) * 3 + x) * 32 + lane];
                s += bias;
                q_n[x] = s > 0.0f ? s : 0.0f;
            }
            kn = argmax3(q_n);
            const float qn_k = kn == 0 ? q_n[0] : (kn == 1 ? q_n[1] : q_n[2]);
            const float qs_k = kn == 0 ? q_s[0] : (kn == 1 ? q_s[1] : q_s[2]);
            target = __fadd_rn(r32, __fmul_rn(hp.gamma, qn_k));
            const float dq = -__fmul_rn(2.0f, __fsub_rn(target, qs_k));
            dz2 = (valid && qs_k > 0.0f) ? dq : 0.0f;
            float4* zdst = reinterpret_cast<float4*>(Zs + (size_t)lane * Hp + c0);
#pragma unroll
            for (int j = 0; j < kColsPerWarp / 4; ++j) {
                float v[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int c = 4 * j + q;
                    const float w2k = W2s[(c0 + c) * 4 + kn];
                    v[q] = acc_s[c] > 0.0f ? __fmul_rn(dz2, w2k) : 0.0f;
                }
                if (c0 + 4 * j < Hp) zdst[j] = make_float4(v[0], v[1], v[2], v[3]);
            }
#pragma unroll
            for (int c = 0; c < kColsPerWarp; ++c) {
                const float hv = __fmul_rn(acc_s[c], dz2);
#pragma unroll
                for (int x = 0; x < 3; ++x) {
                    const float sred = warp_sum(kn == x ? hv : 0.0f);
                    if (lane == 0) dw2acc[(c0 + c) * 3 + x] += sred;
                }
            }
            if (warp == 0) {
                XE[lane] = __double2float_rn(budget);
                XE[32 + lane] = (float)shares;
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
            budget = nb; shares = ns; share_value = p_new;     // carry (TrainerChildActor.scala:99-102)
        }
        // zero the padding columns of the Z tile (when H/4 is odd) so the contraction reads zeros
        if (has_tile && Hp > H)
            for (int idx = tid; idx < kTileEnvs * (Hp - H); idx += nthreads) {
                const int el = idx / (Hp - H), c = H + (idx - el * (Hp - H));
                Zs[(size_t)el * Hp + c] = 0.0f;
            }
        __syncthreads();

        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 2] = clock64();
        // ---- dW1 partial of this tile: X^T · dZ1 over its 32 episodes, straight from shared memory
        if (has_tile) {
            float* out = ea.part1 + (size_t)blockIdx.x * ((size_t)D * H);
            const int rt = tid / ea.n_ct, ct = tid - rt * ea.n_ct;
            if (rt < ea.n_rt) {
                float acc[kEpRows][kEpCols];
#pragma unroll
                for (int r = 0; r < kEpRows; ++r)
#pragma unroll
                    for (int c = 0; c < kEpCols; ++c) acc[r][c] = 0.0f;
                const float* xp = Xs + (size_t)rt * kEpRows * kXStride;
                const float4* zp0 = reinterpret_cast<const float4*>(Zs) + ct;
                const float4* zp1 = reinterpret_cast<const float4*>(Zs) + ct + ea.n_ct;
                const int zs4 = Hp >> 2;
#pragma unroll 2
                for (int el = 0; el < kTileEnvs; ++el) {
                    float x[kEpRows];
#pragma unroll
                    for (int r = 0; r < kEpRows; ++r) x[r] = xp[r * kXStride + el];
                    const float4 za = zp0[el * zs4], zb = zp1[el * zs4];
                    const float z[kEpCols] = {za.x, za.y, za.z, za.w, zb.x, zb.y, zb.z, zb.w};
#pragma unroll
                    for (int r = 0; r < kEpRows; ++r)
#pragma unroll
                        for (int c = 0; c < kEpCols; ++c) acc[r][c] = fmaf(x[r], z[c], acc[r][c]);
                }
#pragma unroll
                for (int r = 0; r < kEpRows; ++r) {
                    const int row = rt * kEpRows + r;
                    if (row < wn) {     // window rows only; budget/shares rows follow below
                        if (ct < H4)
                            reinterpret_cast<float4*>(out + (size_t)row * H)[ct] = make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
                        if (ct + ea.n_ct < H4)
                            reinterpret_cast<float4*>(out + (size_t)row * H)[ct + ea.n_ct] = make_float4(acc[r][4], acc[r][5], acc[r][6], acc[r][7]);
                    }
                }
            }
            // rows wn (budget) and wn + 1 (shares) of dW1
            for (int idx = tid; idx < 2 * H; idx += nthreads) {
                const int m = idx / H, c = idx - m * H;
                float sacc = 0.0f;
                for (int el = 0; el < kTileEnvs; ++el) sacc = fmaf(XE[m * 32 + el], Zs[(size_t)el * Hp + c], sacc);
                out[(size_t)(wn + m) * H + c] = sacc;
            }
            float* dst2 = a.dw2_part + (size_t)blockIdx.x * (3 * H);
            for (int j = tid; j < 3 * H; j += nthreads) dst2[j] = dw2acc[j];
        }
        __syncthreads();
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 3] = clock64();
        if (!ea.apply_in_kernel) break;     // grid-uniform: single-step launch, the host continues with apply_kernel
        // ---- slide the window: row k <- row k + 1 (k < wn), row wn <- the new price
        if (has_tile) {
            constexpr int kMaxPer = 24;
            float keep[kMaxPer];
            const int total = wn * kTileEnvs;                 // elements (k, el), k in [0, wn)
#pragma unroll
            for (int j = 0; j < kMaxPer; ++j) {
                const int idx = tid + j * nthreads;
                if (idx < total) keep[j] = Xs[((idx >> 5) + 1) * kXStride + (idx & 31)];
            }
            __syncthreads();
#pragma unroll
            for (int j = 0; j < kMaxPer; ++j) {
                const int idx = tid + j * nthreads;
                if (idx < total) Xs[(idx >> 5) * kXStride + (idx & 31)] = keep[j];
            }
            if (warp == 0) Xs[wn * kXStride + lane] = p_next;
        }
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 4] = clock64();
        __threadfence();
        grid.sync();
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 5] = clock64();

        // ---- phase B: gradient sum over the tiles (fixed order) + AdaGrad, all CTAs.
        // Four lanes share one float4 of parameters; each lane loads its partials (j = sub, sub+4, ...)
        // eight at a time before adding them, so that the L2 round trips overlap.
        {
            const int64_t total_threads = (int64_t)gridDim.x * nthreads;
            const int64_t P1 = (int64_t)D * H, P = P1 + 3 * (int64_t)H;
            const int64_t n_tasks = ((P >> 2) * 4 + 31) / 32 * 32;
            const int n1 = (int)n_tiles;
            const bool multi = ea.world > 1;
            float* xg_local = multi ? ea.xg[ea.rank] + (size_t)(sidx & 1) * P : nullptr;
            for (int64_t task = (int64_t)blockIdx.x * nthreads + tid; task < n_tasks; task += total_threads) {
                const int64_t g4 = task >> 2;             // float4 group of parameters
                const int sub = (int)(task & 3);
                const int64_t p = g4 << 2;
                const bool in_range = p < P;
                float4 gsum = make_float4(0.f, 0.f, 0.f, 0.f);
                if (in_range) {
                    const float* src; int64_t stride, off;
                    if (p < P1) { src = ea.part1; stride = P1; off = p; }
                    else { src = a.dw2_part; stride = P - P1; off = p - P1; }
                    for (int j0 = sub; j0 < n1; j0 += 32) {
                        float4 v[8];
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            const int j = j0 + 4 * u;
                            v[u] = (j < n1) ? __ldcg(reinterpret_cast<const float4*>(src + (int64_t)j * stride + off))
                                            : make_float4(0.f, 0.f, 0.f, 0.f);
                        }
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            gsum.x += v[u].x; gsum.y += v[u].y; gsum.z += v[u].z; gsum.w += v[u].w;
                        }
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
                    if (multi) {
                        *reinterpret_cast<float4*>(xg_local + p) = gsum;      // this rank's share of the sum
                    } else {
                        *reinterpret_cast<float4*>(ea.grad + p) = gsum;
                        float4 av = *reinterpret_cast<float4*>(ea.acc + p);
                        float4 wv = *reinterpret_cast<float4*>(ea.params_rw + p);
                        const float gg[4] = {gsum.x, gsum.y, gsum.z, gsum.w};
                        float aa[4] = {av.x, av.y, av.z, av.w};
                        float ww[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            aa[q] = __fadd_rn(aa[q], __fmul_rn(gg[q], gg[q]));
                            ww[q] = __fsub_rn(ww[q], __fmul_rn(__fmul_rn(gg[q], hp.lr), __fdiv_rn(1.0f, __fsqrt_rn(aa[q]))));
                        }
                        *reinterpret_cast<float4*>(ea.acc + p) = make_float4(aa[0], aa[1], aa[2], aa[3]);
                        *reinterpret_cast<float4*>(ea.params_rw + p) = make_float4(ww[0], ww[1], ww[2], ww[3]);
                    }
                }
            }
            if (multi) {
                // ---- gradient all-reduce over NVLink peer memory, fused with the AdaGrad apply
                __threadfence_system();
                grid.sync();                                   // this rank's gradient is complete and visible
                const unsigned int want = ea.flag_base + (unsigned int)sidx + 1u;
                if (blockIdx.x == 0 && tid < ea.world)         // raise my flag in every rank's memory (posted NVLink writes)
                    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(ea.flags[tid] + ea.rank), "r"(want) : "memory");
                if (tid < ea.world) {                          // every CTA waits on the LOCAL copies of all ranks' flags
                    const unsigned int* f = ea.flags[ea.rank] + tid;
                    unsigned long long t0 = 0;
                    for (unsigned int spins = 0;; ++spins) {
                        unsigned int v;
                        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
                        if ((int)(v - want) >= 0) break;
                        if ((spins & 4095u) == 4095u) {
                            unsigned long long t;
                            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                            if (t0 == 0) t0 = t;
                            else if (t - t0 > 20000000000ull) __trap();   // a rank never arrived: fail instead of hanging
                        }
                    }
                }
                __syncthreads();
                // sum the ranks' gradients in rank order (bit-identical on every rank) and apply AdaGrad
                const size_t boff = (size_t)(sidx & 1) * P;
                for (int64_t p = ((int64_t)blockIdx.x * nthreads + tid) * 4; p < P; p += total_threads * 4) {
                    // all peers' loads are issued before the first is consumed: one NVLink round trip, not `world`
                    float4 pv[8];
#pragma unroll
                    for (int r = 0; r < 8; ++r)
                        pv[r] = (r < ea.world) ? __ldcv(reinterpret_cast<const float4*>(ea.xg[r] + boff + p))
                                               : make_float4(0.f, 0.f, 0.f, 0.f);
                    float4 gs = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int r = 0; r < 8; ++r) { gs.x += pv[r].x; gs.y += pv[r].y; gs.z += pv[r].z; gs.w += pv[r].w; }
                    *reinterpret_cast<float4*>(ea.grad + p) = gs;
                    float4 av = *reinterpret_cast<float4*>(ea.acc + p);
                    float4 wv = *reinterpret_cast<float4*>(ea.params_rw + p);
                    const float gg[4] = {gs.x, gs.y, gs.z, gs.w};
                    float aa[4] = {av.x, av.y, av.z, av.w};
                    float ww[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        aa[q] = __fadd_rn(aa[q], __fmul_rn(gg[q], gg[q]));
                        ww[q] = __fsub_rn(ww[q], __fmul_rn(__fmul_rn(gg[q], hp.lr), __fdiv_rn(1.0f, __fsqrt_rn(aa[q]))));
                    }
                    *reinterpret_cast<float4*>(ea.acc + p) = make_float4(aa[0], aa[1], aa[2], aa[3]);
                    *reinterpret_cast<float4*>(ea.params_rw + p) = make_float4(ww[0], ww[1], ww[2], ww[3]);
                }
            }
        }
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 6] = clock64();
        __threadfence();
        grid.sync();
        if (ea.dbg && blockIdx.x == 0 && tid == 0) ea.dbg[(size_t)sidx * kDbgSlots + 7] = clock64();
    }

    if (has_tile && warp == 0 && valid) {
        a.budget[e] = budget;
        a.shares[e] = shares;
        a.share_value[e] = share_value;
    }
    if (ea.apply_in_kernel && blockIdx.x == 0 && tid == 0) *ea.d_step_rw = ea.step0 + ea.n_steps;
}

// Measurement aid: dependency-free FFMA streams, to measure the fp32-FMA roof of this GPU (the
// binding roof of the fp32 path; MEASURED_PEAKS.json carries only HBM and bf16 tensor peaks).
__global__ void __launch_bounds__(256) fma_peak_kernel(float* out, int iters, float a, float b) {
    float v[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = (float)(threadIdx.x + j);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = fmaf(v[j], a, b);
    }
    float s = 0.0f;
#pragma unroll
    for (int j = 0; j < 16; ++j) s += v[j];
    if (s == 123.456f) out[0] = s;  // keeps the loop alive without a store in practice
}

// Measurement aid: the FFMA pattern register-tiled contractions actually issue — an 8 x 8 outer
// product per step, every FFMA with three register sources (acc += a[i] * b[j]).  The dependency-free
// stream above multiplies by kernel-uniform operands, which the hardware can take from the uniform
// datapath at full rate; this one measures the rate of the three-register form.
__global__ void __launch_bounds__(256) fma_outer_kernel(float* out, int iters) {
    float a[8], b[8], acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = 1.0f + 1e-3f * (float)((threadIdx.x + i) & 15); b[i] = 1.0f - 1e-3f * (float)((threadIdx.x * 3 + i) & 15); }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        a[it & 7] += 1e-7f;      // keeps the operands live without adding measurable work
    }
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) s += acc[i][j];
    if (s == 123.456f) out[0] = s;
}

#endif  // __CUDACC__

}  // namespace qp
