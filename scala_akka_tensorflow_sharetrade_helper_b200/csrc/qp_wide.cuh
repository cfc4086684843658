// qp_wide.cuh — the path for policies whose W1 does not fit in shared memory (BASELINE.json
// configs[3]: hist 512, hidden 2048, batch 65 536).  Here the dense layer is a real contraction, so
// an env-step is organised around two GEMMs instead of one fused per-tile kernel:
//   G1  Z1[2B, H]  = Xwin[2B, wn] · W1[0:wn, :]          (rows 0..B-1: state s, rows B..2B-1: s')
//   E1  per episode: + budget/shares rows, bias, ReLU, output layer, select, book-keeping,
//       TD target, dZ1[B, H]                              (QDecisionPolicyActor.scala:43-73)
//   E2  dW2 partials   (column-parallel over episodes)
//   G2  dW1 partials = Xᵀ[D, B] · dZ1[B, H], split over episodes
//   apply_kernel       (fixed-order sum of partials + AdaGrad, shared with the other paths)
// The GEMMs below are the fp32 FFMA versions (register-tiled 128x128x16): exact fp32 semantics,
// same tolerance as the fused path.
#pragma once
#include "qp_kernels.cuh"

namespace qp {

constexpr int kWgBM = 128, kWgBN = 128, kWgBK = 16, kWgThreads = 256;
constexpr int kWgAStride = kWgBM + 4;     // padded: the forward A tile is stored transposed
constexpr int kWideMaxExtraRows = 8;      // rows of dW1 handled outside G2's 128-row tiles on the tensor-core path

struct WideArgs {
    Hyper hp;
    const float* params;
    const float* prices; int64_t T; const int32_t* series_of_env; const int32_t* d_step;
    double* budget; int32_t* shares; double* share_value;
    double B0; int32_t N0; int64_t env_id_base;
    int64_t B, Bpad;
    float* Z1;            // [2B, H]
    float* dZ1;           // [B, H]
    float* xextra;        // [2, Bpad]
    int32_t* kn;          // [B] argmax of q(s')
    float* dz2;           // [B]
    float* part1;         // [S, D*H]
    float* part2;         // [S2, 3H]
    int n_split;          // episode splits of G2
    int n_split2;         // episode splits of E2
    // tensor-core variant: the rows of X^T the 128-row tiles of G2 do not cover (window rows from
    // xe_row0 on, then budget, shares), as a K-major tf32 hi/lo matrix [kWideMaxExtraRows, xe_ld]
    float* xe_hi; float* xe_lo; int64_t xe_ld; int xe_row0;
    TraceDev trace; int32_t trace_step0;
    // message-level calls on explicit rows (SelectionAction / UpdateQ, QDecisionPolicyActor.scala:56-76): msg = 1 / 2.
    // The state rows [B, D] replace the price windows and the episode record; Z1 holds B (select) or 2B (update) rows.
    int msg;
    const float* m_states; const float* m_next; const float* m_rewards; const float* m_steps; const int64_t* m_ids;
    int32_t* m_actions; float* m_q;
};

#ifdef __CUDACC__

// ---------------------------------------------------------------------------------------------
// Register-tiled fp32 GEMM, C[M,N] = sum_k A(m,k) * B[k,N].  MODE 0 (forward): A(m,k) = price window
// element k of row m (m < B: state s of episode m; m >= B: state s' of episode m-B), K = wn.
// MODE 1 (backward): the contraction runs over the episodes of split blockIdx.z; A(k=e, m) = input m
// of episode e (price window, then budget, shares), B = dZ1 rows, C = this split's D x H partial.
// ---------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(kWgThreads, 2) wide_gemm_kernel(const WideArgs a) {
    __shared__ __align__(16) float As[2][kWgBK][kWgAStride];
    __shared__ __align__(16) float Bs[2][kWgBK][kWgBN];
    const Hyper& hp = a.hp;
    const int tid = threadIdx.x;
    const int H = hp.H, wn = hp.wn, D = hp.D;
    const int32_t istep = a.msg ? 0 : *a.d_step;
    const int64_t B = a.B;
    const int64_t M = (MODE == 0) ? (a.msg == 1 ? B : 2 * B) : (int64_t)D;
    const int64_t m0 = (int64_t)blockIdx.x * kWgBM;
    const int n0 = blockIdx.y * kWgBN;
    int64_t k_begin = 0, k_end = wn;
    if (MODE == 1) {
        k_begin = B * (int64_t)blockIdx.z / gridDim.z;
        k_end = B * ((int64_t)blockIdx.z + 1) / gridDim.z;
    }
    const float* Bmat = (MODE == 0) ? a.params : a.dZ1;     // [K, H] row-major in both modes

    // per-thread load coordinates
    // forward A: thread reads 8 elements: kk = tid % 16, rows tid / 16 + 16 j
    // backward A: thread reads 8 elements: m = tid % 128, kk = tid / 128 + 2 j
    float ra[8];
    float4 rb[2];
    const int b_kk = tid >> 5, b_n4 = tid & 31;               // B tile: 16 rows x 32 float4; rows b_kk, b_kk + 8

    auto load_tiles = [&](int64_t kt) {
        if (MODE == 0) {
            const int kk = tid & 15;
            const int64_t k = kt + kk;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int64_t m = m0 + (tid >> 4) + 16 * j;
                float v = 0.0f;
                if (m < M && k < k_end) {
                    const int64_t e = m < B ? m : m - B;
                    if (a.msg) v = __ldg((m < B ? a.m_states : a.m_next) + e * D + k);
                    else v = __ldg(a.prices + (int64_t)a.series_of_env[e] * a.T + istep + (m < B ? 0 : 1) + k);
                }
                ra[j] = v;
            }
        } else {
            const int mloc = tid & 127;
            const int64_t m = m0 + mloc;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int64_t e = kt + (tid >> 7) + 2 * j;
                float v = 0.0f;
                if (e < k_end && a.msg) {
                    if (m < D) v = __ldg(a.m_states + e * D + m);
                } else if (e < k_end) {
                    if (m < wn) v = __ldg(a.prices + (int64_t)a.series_of_env[e] * a.T + istep + m);
                    else if (m == wn) v = a.xextra[e];
                    else if (m == wn + 1) v = a.xextra[a.Bpad + e];
                }
                ra[j] = v;
            }
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int64_t k = kt + b_kk + 8 * j;
            rb[j] = (k < k_end && n0 + b_n4 * 4 < H)
                        ? *reinterpret_cast<const float4*>(Bmat + k * H + n0 + b_n4 * 4)
                        : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    };
    auto store_tiles = [&](int buf) {
        if (MODE == 0) {
            const int kk = tid & 15;
#pragma unroll
            for (int j = 0; j < 8; ++j) As[buf][kk][(tid >> 4) + 16 * j] = ra[j];
        } else {
            const int mloc = tid & 127;
#pragma unroll
            for (int j = 0; j < 8; ++j) As[buf][(tid >> 7) + 2 * j][mloc] = ra[j];
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) *reinterpret_cast<float4*>(&Bs[buf][b_kk + 8 * j][b_n4 * 4]) = rb[j];
    };

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;
    const int ty = tid >> 4, tx = tid & 15;      // rows ty*4 + {0..3} and 64 + ty*4 + {0..3}; cols likewise with tx

    load_tiles(k_begin);
    store_tiles(0);
    __syncthreads();
    int buf = 0;
    for (int64_t kt = k_begin; kt < k_end; kt += kWgBK) {
        const bool more = kt + kWgBK < k_end;
        if (more) load_tiles(kt + kWgBK);
#pragma unroll
        for (int kk = 0; kk < kWgBK; ++kk) {
            const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
            const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        if (more) {
            store_tiles(buf ^ 1);
            __syncthreads();
            buf ^= 1;
        }
    }

    float* C = (MODE == 0) ? a.Z1 : a.part1 + (size_t)blockIdx.z * ((size_t)D * H);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int64_t m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (m >= M) continue;
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
            const int n = n0 + (jj == 0 ? tx * 4 : 64 + tx * 4);
            if (n < H)
                *reinterpret_cast<float4*>(C + m * H + n) =
                    make_float4(acc[i][jj * 4 + 0], acc[i][jj * 4 + 1], acc[i][jj * 4 + 2], acc[i][jj * 4 + 3]);
        }
    }
}

// h = relu(z1 + budget*W1[wn] + shares*W1[wn+1] + 0.1): the one place this is written, so that
// E1 and E2 agree bit for bit on which hidden units are active.
__device__ __forceinline__ float wide_hidden(float zwin, float xb, float xs, float wb, float ws, float bias) {
    const float v = fmaf(xs, ws, fmaf(xb, wb, zwin)) + bias;
    return v > 0.0f ? v : 0.0f;
}

// ---------------------------------------------------------------------------------------------
// E1 — one warp per episode (lanes over hidden columns, float4): everything between the two GEMMs.
// Shared memory: W2 as float4 per hidden unit and the budget/shares rows of W1.
// ---------------------------------------------------------------------------------------------
constexpr int kWideMaxColsPerLane = 64;   // hidden <= 2048 on this path

template <int MSG>      // 0: env-step; 1: SelectionAction rows; 2: UpdateQ rows
__global__ void __launch_bounds__(256) wide_row_kernel(const WideArgs a) {
    extern __shared__ __align__(16) float smem[];
    const Hyper& hp = a.hp;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int H = hp.H, wn = hp.wn, D = hp.D;
    // A lane owns columns c = 128 j + 4 lane + q (so that its global accesses are float4).  The three
    // per-column tables are stored permuted, index p(c) = (4 j + q) * 32 + lane, so that for a fixed
    // (j, q) the 32 lanes read consecutive words: no bank conflicts.
    float* W2s = smem;                       // [H] float4 (w2[c][0..2], 0), permuted
    float* Wb = W2s + (size_t)H * 4;         // [H] budget row of W1, permuted
    float* Ws = Wb + H;                      // [H] shares row of W1, permuted
    const float* W1g = a.params;
    const float* W2g = a.params + (size_t)D * H;
    for (int c = tid; c < H; c += blockDim.x) {
        const int p = (((c >> 7) << 2) + (c & 3)) * 32 + ((c & 127) >> 2);
        reinterpret_cast<float4*>(W2s)[p] = make_float4(W2g[c * 3], W2g[c * 3 + 1], W2g[c * 3 + 2], 0.f);
        Wb[p] = W1g[(size_t)wn * H + c];
        Ws[p] = W1g[(size_t)(wn + 1) * H + c];
    }
    __syncthreads();
    const int32_t istep = MSG ? 0 : *a.d_step;
    const float bias = hp.bias;
    const int nchunk = H >> 7;               // float4 chunks of 128 columns per warp pass

    for (int64_t e = (int64_t)blockIdx.x * nwarps + warp; e < a.B; e += (int64_t)gridDim.x * nwarps) {
        const double budget = MSG ? 0.0 : a.budget[e], share_value = MSG ? 0.0 : a.share_value[e];
        const int32_t shares = MSG ? 0 : a.shares[e];
        const float* prow = MSG ? nullptr : a.prices + (int64_t)a.series_of_env[e] * a.T + istep;
        // (message rows carry their own budget / shares entries: the last two elements of the state)
        const float xb = MSG ? __ldg(a.m_states + e * D + wn) : __double2float_rn(budget);
        const float xs = MSG ? __ldg(a.m_states + e * D + wn + 1) : (float)shares;
        const float4* zs4 = reinterpret_cast<const float4*>(a.Z1 + e * H);
        const float4* zn4 = reinterpret_cast<const float4*>(a.Z1 + (a.B + e) * H);

        // both Z1 rows of this episode are fetched up front (32 independent 16-byte loads per lane):
        // the row of s' does not depend on the action, only its budget/shares terms do
        float4 zsv[kWideMaxColsPerLane / 4], znv[kWideMaxColsPerLane / 4];
#pragma unroll
        for (int j = 0; j < kWideMaxColsPerLane / 4; ++j)
            if (j < nchunk) { zsv[j] = __ldcs(zs4 + j * 32 + lane); if (MSG != 1) znv[j] = __ldcs(zn4 + j * 32 + lane); }

        // hidden layer of s and its Q-values (QDecisionPolicyActor.scala:43, :47)
        float hs[kWideMaxColsPerLane];
        float pq[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < kWideMaxColsPerLane / 4; ++j) {
            if (j < nchunk) {
                const float4 z = zsv[j];
                const float zz[4] = {z.x, z.y, z.z, z.w};
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int p = (j * 4 + q) * 32 + lane;
                    const float h = wide_hidden(zz[q], xb, xs, Wb[p], Ws[p], bias);
                    hs[j * 4 + q] = h;
                    const float4 w2 = reinterpret_cast<const float4*>(W2s)[p];
                    pq[0] = fmaf(h, w2.x, pq[0]); pq[1] = fmaf(h, w2.y, pq[1]); pq[2] = fmaf(h, w2.z, pq[2]);
                }
            }
        }
        float q_s[3];
#pragma unroll
        for (int x = 0; x < 3; ++x) { const float s = warp_sum(pq[x]) + bias; q_s[x] = s > 0.0f ? s : 0.0f; }

        // SelectionAction (:56-62)
        int action = 0, explored = 0;
        if (MSG != 2) {
            const float step_f = MSG ? a.m_steps[e] : (float)istep;
            const uint64_t env_id = MSG ? (uint64_t)(a.m_ids ? a.m_ids[e] : e) : (uint64_t)(a.env_id_base + e);
            uint32_t r[4];
            philox4x32_10((uint32_t)env_id, (uint32_t)(env_id >> 32), MSG ? (uint32_t)step_f : (uint32_t)istep, 1u,
                          (uint32_t)hp.seed, (uint32_t)(hp.seed >> 32), r);
            const float u = (float)(r[0] >> 8) * (1.0f / 16777216.0f);
            float thr = __fdiv_rn(step_f, hp.step_div);
            if (hp.epsilon < thr) thr = hp.epsilon;
            if (u < thr) action = argmax3(q_s);
            else { explored = 1; action = (int)(((uint64_t)r[1] * 3u) >> 32); }
        }
        if (MSG == 1) {
            if (lane == 0) {
                a.m_actions[e] = action;
                if (a.m_q) { a.m_q[e * 3] = q_s[0]; a.m_q[e * 3 + 1] = q_s[1]; a.m_q[e * 3 + 2] = q_s[2]; }
            }
            continue;
        }
        // decision, reward, next state (TrainerChildActor.scala:92-96, :118-123, :136-146)
        const double p_new = MSG ? 0.0 : (double)__ldg(prow + wn);
        const double cur = __dadd_rn(budget, __dmul_rn((double)shares, share_value));
        const double base_b = hp.bookkeeping == 0 ? a.B0 : budget;
        const int32_t base_n = hp.bookkeeping == 0 ? a.N0 : shares;
        double nb = base_b; int32_t ns = base_n; int a_eff = 2;
        if (action == 0 && base_b >= p_new) { nb = __dsub_rn(base_b, p_new); ns = base_n + 1; a_eff = 0; }
        else if (action == 1 && base_n > 0) { nb = __dadd_rn(base_b, p_new); ns = base_n - 1; a_eff = 1; }
        const double newp = __dadd_rn(nb, __dmul_rn((double)ns, p_new));
        const float r32 = MSG ? a.m_rewards[e] : __double2float_rn(__dsub_rn(newp, cur));
        const float xnb = MSG ? __ldg(a.m_next + e * D + wn) : __double2float_rn(nb);
        const float xns = MSG ? __ldg(a.m_next + e * D + wn + 1) : (float)ns;

        // Q-values of s'
        pq[0] = pq[1] = pq[2] = 0.0f;
#pragma unroll
        for (int j = 0; j < kWideMaxColsPerLane / 4; ++j) {
            if (j < nchunk) {
                const float4 z = znv[j];
                const float zz[4] = {z.x, z.y, z.z, z.w};
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int p = (j * 4 + q) * 32 + lane;
                    const float h = wide_hidden(zz[q], xnb, xns, Wb[p], Ws[p], bias);
                    const float4 w2 = reinterpret_cast<const float4*>(W2s)[p];
                    pq[0] = fmaf(h, w2.x, pq[0]); pq[1] = fmaf(h, w2.y, pq[1]); pq[2] = fmaf(h, w2.z, pq[2]);
                }
            }
        }
        float q_n[3];
#pragma unroll
        for (int x = 0; x < 3; ++x) { const float s = warp_sum(pq[x]) + bias; q_n[x] = s > 0.0f ? s : 0.0f; }

        // UpdateQ target and output-layer backward (QDecisionPolicyActor.scala:67-73, :49)
        const int kn = argmax3(q_n);
        const float qn_k = kn == 0 ? q_n[0] : (kn == 1 ? q_n[1] : q_n[2]);
        const float qs_k = kn == 0 ? q_s[0] : (kn == 1 ? q_s[1] : q_s[2]);
        const float target = __fadd_rn(r32, __fmul_rn(hp.gamma, qn_k));
        const float dq = -__fmul_rn(2.0f, __fsub_rn(target, qs_k));
        const float dz2 = qs_k > 0.0f ? dq : 0.0f;
        float4* dz4 = reinterpret_cast<float4*>(a.dZ1 + e * H);
#pragma unroll
        for (int j = 0; j < kWideMaxColsPerLane / 4; ++j) {
            if (j < nchunk) {
                const int c = j * 128 + lane * 4;
                float v[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float4 w2 = reinterpret_cast<const float4*>(W2s)[(j * 4 + q) * 32 + lane];
                    const float w2k = kn == 0 ? w2.x : (kn == 1 ? w2.y : w2.z);
                    v[q] = hs[j * 4 + q] > 0.0f ? __fmul_rn(dz2, w2k) : 0.0f;
                }
                dz4[c >> 2] = make_float4(v[0], v[1], v[2], v[3]);
            }
        }
        if (MSG == 0 && a.xe_hi && lane < D - a.xe_row0) {
            const int m = a.xe_row0 + lane;
            const float x = m < wn ? __ldg(prow + m) : (m == wn ? xb : xs);
            uint32_t t;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(t) : "f"(x));
            const float hpart = __uint_as_float(t);
            a.xe_hi[(int64_t)lane * a.xe_ld + e] = hpart;
            a.xe_lo[(int64_t)lane * a.xe_ld + e] = x - hpart;
        }
        if (lane == 0) {
            a.xextra[e] = xb; a.xextra[a.Bpad + e] = xs;
            a.kn[e] = kn; a.dz2[e] = dz2;
            if (MSG) {
                if (a.m_q) { a.m_q[e * 3] = q_s[0]; a.m_q[e * 3 + 1] = q_s[1]; a.m_q[e * 3 + 2] = q_s[2]; }
            } else {
            a.budget[e] = nb; a.shares[e] = ns; a.share_value[e] = p_new;
            const TraceDev& t = a.trace;
            const int64_t ti = (int64_t)(istep - a.trace_step0) * a.B + e;
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
    }
}

// ---------------------------------------------------------------------------------------------
// E2 — dW2[:, k] += h(s) * dz2 summed over the episodes of split blockIdx.y, one thread per hidden
// column (so the sum over episodes is sequential and fixed).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) wide_dw2_kernel(const WideArgs a) {
    const Hyper& hp = a.hp;
    const int H = hp.H, wn = hp.wn;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= H) return;
    const float wb = a.params[(size_t)wn * H + c], ws = a.params[(size_t)(wn + 1) * H + c];
    const int64_t e0 = a.B * (int64_t)blockIdx.y / gridDim.y, e1 = a.B * ((int64_t)blockIdx.y + 1) / gridDim.y;
    float acc[3] = {0.f, 0.f, 0.f};
#pragma unroll 4
    for (int64_t e = e0; e < e1; ++e) {
        const float dz2 = a.dz2[e];
        if (dz2 == 0.0f) continue;                       // uniform across the block; contributes exactly 0
        const float h = wide_hidden(a.Z1[e * H + c], a.xextra[e], a.xextra[a.Bpad + e], wb, ws, hp.bias);
        const float v = __fmul_rn(h, dz2);
        const int kn = a.kn[e];
        acc[0] += kn == 0 ? v : 0.0f; acc[1] += kn == 1 ? v : 0.0f; acc[2] += kn == 2 ? v : 0.0f;
    }
    float* out = a.part2 + (size_t)blockIdx.y * (3 * H) + (size_t)c * 3;
    out[0] = acc[0]; out[1] = acc[1]; out[2] = acc[2];
}

#endif  // __CUDACC__
}  // namespace qp
