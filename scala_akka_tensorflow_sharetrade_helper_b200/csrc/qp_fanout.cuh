// qp_fanout.cuh — the router fan-out: MANY episodes on FEW price series (TrainerRouterActor.scala:44, 60-66, 88 sends
// ONE series to all its children; BASELINE.json configs[4]: 512 tickers x 256 episodes).
//
// The window part of the hidden layer, sum_k p[i+k] W1[k,:], is the same for every episode of a series at a step: the
// episodes differ only in their budget / shares entries.  So per env-step
//   F1  fan_zwin_kernel   zw[t][s|s'][H] = window part of z1 for the state s and the state s' of series t — once per
//                         series instead of once per episode
//   F2  fan_row_kernel    per chunk of up to 256 episodes of one series: (pass 1, thread = episode) hidden layer from zw +
//                         the episode's budget / shares terms, Q(s), select, decision, reward, Q(s') of the three
//                         outcomes, TD target, dz2 (QDecisionPolicyActor.scala:43-73, TrainerChildActor.scala:87-146);
//                         (pass 2, thread = hidden column) the chunk's sums over its episodes, in episode order:
//                         SdZ[c] = sum_e dZ1_e[c], the budget / shares rows of dW1, dW2
//   F3  fan_grad_kernel   dW1[k, c] = sum_chunks p_t(chunk)[i+k] * SdZ[chunk][c] (the backward of the window rows is an
//                         outer product per series), the other rows and dW2 summed over the chunks, in chunk order
//   apply_kernel          AdaGrad (shared with the other paths; several GPUs: NCCL all-reduce of the gradient before it)
// Same arithmetic as the other paths with the sums over episodes re-associated (per series first): inside the stated
// tolerance, bit-reproducible (every sum has a fixed order).  fp32 FFMA throughout: per episode and step this path does
// ~9 K flops instead of 247 K, so no tensor core is needed — the work left is the element-wise part.
#pragma once
#include "qp_wide.cuh"
#include <type_traits>

namespace qp {

constexpr int kFanChunk = 256;          // episodes per chunk = threads of fan_row_kernel
constexpr int kFanSeriesPerCta = 8;     // series per CTA of fan_zwin_kernel
constexpr int kFanKSplit = 8;           // fan_zwin_kernel: window rows over the warps; fan_grad_kernel: ranges of chunks

struct FanArgs {
    Hyper hp;
    const float* params;
    const float* prices; int64_t T; int64_t n_series; const int32_t* d_step;
    double* budget; int32_t* shares; double* share_value;
    double B0; int32_t N0; int64_t env_id_base;
    int64_t B;
    const int4* chunks; int n_chunks;   // (series, first episode, episodes, 0); the episodes of a chunk are consecutive
    float* zw;                          // [n_series][2][H]
    float* part;                        // [n_chunks][6][H]: SdZ, budget row, shares row, dW2 as [3][H]
    float* part1; float* part2;         // [kFanKSplit][D*H], [kFanKSplit][3H]: the gradient per range of chunks
    TraceDev trace; int32_t trace_step0;
};

#ifdef __CUDACC__

// h = relu(zb + budget * W1[wn] + shares * W1[wn+1]) with zb = window part + bias (added once per series and column in
// the table, not once per episode).  The one place this is written: both passes of fan_row_kernel must agree bit for bit
// on which hidden units are active.
__device__ __forceinline__ float fan_hidden(float zb, float xb, float xs, float wb, float ws) {
    const float v = fmaf(xs, ws, fmaf(xb, wb, zb));
    return v > 0.0f ? v : 0.0f;
}

// F1: a CTA takes kFanSeriesPerCta series and 32 hidden columns; warp w takes one eighth of the window rows (all its
// W1 loads independent, 128 bytes per warp and row), the eight partial sums meet in shared memory and are added in warp
// order.  (One thread per column over all 201 rows was a chain of 50 dependent load batches: 66 us.)
__global__ void __launch_bounds__(32 * kFanKSplit) fan_zwin_kernel(const FanArgs a) {
    extern __shared__ __align__(16) float fan_smem[];
    const Hyper& hp = a.hp;
    const int H = hp.H, wn = hp.wn;
    const int32_t istep = *a.d_step;
    const int64_t t0 = (int64_t)blockIdx.x * kFanSeriesPerCta;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int c = blockIdx.y * 32 + lane;
    float* pw = fan_smem;                                       // [kFanSeriesPerCta][wn + 1]
    float* red = pw + kFanSeriesPerCta * (wn + 1);              // [kFanKSplit][2 * kFanSeriesPerCta][32]
    for (int i = threadIdx.x; i < kFanSeriesPerCta * (wn + 1); i += blockDim.x) {
        const int j = i / (wn + 1), k = i - j * (wn + 1);
        const int64_t t = t0 + j;
        float v = 0.0f;
        if (t < a.n_series && (int64_t)istep + k < a.T) v = __ldg(a.prices + t * a.T + istep + k);
        pw[i] = v;
    }
    __syncthreads();
    float zs[kFanSeriesPerCta], zn[kFanSeriesPerCta];
#pragma unroll
    for (int j = 0; j < kFanSeriesPerCta; ++j) { zs[j] = 0.0f; zn[j] = 0.0f; }
    if (c < H) {
        // warp w takes a contiguous range of window rows: the price of row k + 1 (state s') is the price of row k of the
        // next iteration (state s), so each iteration reads one new price per series
        const int kper = (wn + kFanKSplit - 1) / kFanKSplit;
        const int kb = w * kper, ke = min(wn, kb + kper);
        float pk[kFanSeriesPerCta];
#pragma unroll
        for (int j = 0; j < kFanSeriesPerCta; ++j) pk[j] = kb < wn ? pw[j * (wn + 1) + kb] : 0.0f;
#pragma unroll 4
        for (int k = kb; k < ke; ++k) {
            const float wv = __ldg(a.params + (size_t)k * H + c);
#pragma unroll
            for (int j = 0; j < kFanSeriesPerCta; ++j) {
                const float pn = pw[j * (wn + 1) + k + 1];
                zs[j] = fmaf(pk[j], wv, zs[j]);
                zn[j] = fmaf(pn, wv, zn[j]);
                pk[j] = pn;
            }
        }
    }
#pragma unroll
    for (int j = 0; j < kFanSeriesPerCta; ++j) {
        red[(w * 2 * kFanSeriesPerCta + 2 * j) * 32 + lane] = zs[j];
        red[(w * 2 * kFanSeriesPerCta + 2 * j + 1) * 32 + lane] = zn[j];
    }
    __syncthreads();
    // 2 * kFanSeriesPerCta * 32 outputs, one per thread
    for (int o = threadIdx.x; o < 2 * kFanSeriesPerCta * 32; o += blockDim.x) {
        const int row = o >> 5, l = o & 31, j = row >> 1, half = row & 1;
        float s_ = 0.0f;
#pragma unroll
        for (int ww = 0; ww < kFanKSplit; ++ww) s_ += red[(ww * 2 * kFanSeriesPerCta + row) * 32 + l];
        const int64_t t = t0 + j;
        const int cc = blockIdx.y * 32 + l;
        if (t < a.n_series && cc < H) a.zw[(t * 2 + half) * H + cc] = s_;
    }
}

// F2: one chunk (<= 256 consecutive episodes of one series) per CTA.
__global__ void __launch_bounds__(kFanChunk, 4) fan_row_kernel(const FanArgs a) {
    extern __shared__ __align__(16) float fan_smem[];
    const Hyper& hp = a.hp;
    const int H = hp.H, wn = hp.wn, D = hp.D;
    const int tid = threadIdx.x;
    const int4 ch = a.chunks[blockIdx.x];
    const int64_t t = ch.x, e0 = ch.y;
    const int count = ch.z;
    const int32_t istep = *a.d_step;
    const float bias = hp.bias;
    float4* tab = reinterpret_cast<float4*>(fan_smem);          // [H] (zs + bias, zn + bias, wb, ws)
    float4* W2s = tab + H;                                      // [H] (w2[c][0..2], 0)
    float4* epi = W2s + H;                                      // [kFanChunk] (xb, xs, dz2, kn as int bits)
    const float* W1g = a.params;
    const float* W2g = a.params + (size_t)D * H;
    for (int c = tid; c < H; c += kFanChunk) {
        tab[c] = make_float4(a.zw[(t * 2 + 0) * H + c] + bias, a.zw[(t * 2 + 1) * H + c] + bias, W1g[(size_t)wn * H + c], W1g[(size_t)(wn + 1) * H + c]);
        W2s[c] = make_float4(W2g[c * 3], W2g[c * 3 + 1], W2g[c * 3 + 2], 0.0f);
    }
    __syncthreads();

    // ---- pass 1: thread = episode
    // Which outcomes can occur in this chunk at all?  (An episode that cannot buy / sell has its Buy / Sell outcome equal to
    // Hold; with the reference's stateless book-keeping and no shares to start with, nobody can ever sell.)
    bool my_buy = false, my_sell = false;
    if (tid < count) {
        const int64_t e = e0 + tid;
        const double pn = (double)__ldg(a.prices + t * a.T + istep + wn);
        my_buy = (hp.bookkeeping == 0 ? a.B0 : a.budget[e]) >= pn;
        my_sell = (hp.bookkeeping == 0 ? a.N0 : a.shares[e]) > 0;
    }
    const bool any_buy = __syncthreads_or(my_buy) != 0, any_sell = __syncthreads_or(my_sell) != 0;
    if (tid < count) {
        const int64_t e = e0 + tid;
        const double budget = a.budget[e], share_value = a.share_value[e];
        const int32_t shares = a.shares[e];
        const float xb = __double2float_rn(budget), xs = (float)shares;
        const double p_new = (double)__ldg(a.prices + t * a.T + istep + wn);          // TrainerChildActor.scala:94
        const double cur = __dadd_rn(budget, __dmul_rn((double)shares, share_value));  // :92
        // the three states s' the decision can lead to (index = effective action; :118-123)
        const double base_b = hp.bookkeeping == 0 ? a.B0 : budget;
        const int32_t base_n = hp.bookkeeping == 0 ? a.N0 : shares;
        const bool can_buy = base_b >= p_new, can_sell = base_n > 0;
        const double nbc[3] = {can_buy ? __dsub_rn(base_b, p_new) : base_b, can_sell ? __dadd_rn(base_b, p_new) : base_b, base_b};
        const int32_t nsc[3] = {can_buy ? base_n + 1 : base_n, can_sell ? base_n - 1 : base_n, base_n};
        const float xnb[3] = {__double2float_rn(nbc[0]), __double2float_rn(nbc[1]), __double2float_rn(nbc[2])};
        const float xns[3] = {(float)nsc[0], (float)nsc[1], (float)nsc[2]};
        float pq[4][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
        auto sweep = [&](auto buy_c, auto sell_c) {
            constexpr bool kBuy = decltype(buy_c)::value, kSell = decltype(sell_c)::value;
#pragma unroll 2
            for (int c = 0; c < H; ++c) {
                const float4 tc = tab[c], w2 = W2s[c];
                const float h = fan_hidden(tc.x, xb, xs, tc.z, tc.w);
                pq[0][0] = fmaf(h, w2.x, pq[0][0]); pq[0][1] = fmaf(h, w2.y, pq[0][1]); pq[0][2] = fmaf(h, w2.z, pq[0][2]);
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    if ((j == 0 && !kBuy) || (j == 1 && !kSell)) continue;
                    const float hn = fan_hidden(tc.y, xnb[j], xns[j], tc.z, tc.w);
                    pq[1 + j][0] = fmaf(hn, w2.x, pq[1 + j][0]); pq[1 + j][1] = fmaf(hn, w2.y, pq[1 + j][1]);
                    pq[1 + j][2] = fmaf(hn, w2.z, pq[1 + j][2]);
                }
            }
        };
        if (any_buy && any_sell) sweep(std::true_type{}, std::true_type{});
        else if (any_buy) sweep(std::true_type{}, std::false_type{});
        else if (any_sell) sweep(std::false_type{}, std::true_type{});
        else sweep(std::false_type{}, std::false_type{});
        // (an outcome nobody in the chunk can reach is never selected below: a_eff is 0 / 1 only for an episode that can)
        float q_s[3], q_n[3];
#pragma unroll
        for (int x = 0; x < 3; ++x) { const float s = pq[0][x] + bias; q_s[x] = s > 0.0f ? s : 0.0f; }
        // SelectionAction (QDecisionPolicyActor.scala:56-62)
        int action, explored = 0;
        {
            const uint64_t env_id = (uint64_t)(a.env_id_base + e);
            uint32_t r[4];
            philox4x32_10((uint32_t)env_id, (uint32_t)(env_id >> 32), (uint32_t)istep, 1u, (uint32_t)hp.seed, (uint32_t)(hp.seed >> 32), r);
            const float u = (float)(r[0] >> 8) * (1.0f / 16777216.0f);
            float thr = __fdiv_rn((float)istep, hp.step_div);
            if (hp.epsilon < thr) thr = hp.epsilon;
            if (u < thr) action = argmax3(q_s);
            else { explored = 1; action = (int)(((uint64_t)r[1] * 3u) >> 32); }
        }
        int a_eff = 2;
        if (action == 0 && can_buy) a_eff = 0;
        else if (action == 1 && can_sell) a_eff = 1;
        const double nb = a_eff == 0 ? nbc[0] : (a_eff == 1 ? nbc[1] : nbc[2]);
        const int32_t ns = a_eff == 0 ? nsc[0] : (a_eff == 1 ? nsc[1] : nsc[2]);
        const double newp = __dadd_rn(nb, __dmul_rn((double)ns, p_new));
        const float r32 = __double2float_rn(__dsub_rn(newp, cur));                     // :95-96, :136-146
#pragma unroll
        for (int x = 0; x < 3; ++x) {
            const float s = (a_eff == 0 ? pq[1][x] : (a_eff == 1 ? pq[2][x] : pq[3][x])) + bias;
            q_n[x] = s > 0.0f ? s : 0.0f;
        }
        // UpdateQ (QDecisionPolicyActor.scala:64-76)
        const int kn = argmax3(q_n);
        const float qn_k = kn == 0 ? q_n[0] : (kn == 1 ? q_n[1] : q_n[2]);
        const float qs_k = kn == 0 ? q_s[0] : (kn == 1 ? q_s[1] : q_s[2]);
        const float target = __fadd_rn(r32, __fmul_rn(hp.gamma, qn_k));
        const float dq = -__fmul_rn(2.0f, __fsub_rn(target, qs_k));
        const float dz2 = qs_k > 0.0f ? dq : 0.0f;
        epi[tid] = make_float4(xb, xs, dz2, __int_as_float(kn));
        a.budget[e] = nb; a.shares[e] = ns; a.share_value[e] = p_new;                  // carry (:99-102)
        const TraceDev& tr = a.trace;
        const int64_t ti = (int64_t)(istep - a.trace_step0) * a.B + e;
        if (tr.action_sel) tr.action_sel[ti] = action;
        if (tr.action_eff) tr.action_eff[ti] = a_eff;
        if (tr.explored) tr.explored[ti] = explored;
        if (tr.k_next) tr.k_next[ti] = kn;
        if (tr.reward) tr.reward[ti] = r32;
        if (tr.target) tr.target[ti] = target;
        if (tr.budget) tr.budget[ti] = nb;
        if (tr.shares) tr.shares[ti] = ns;
        if (tr.q_s) { tr.q_s[ti * 3] = q_s[0]; tr.q_s[ti * 3 + 1] = q_s[1]; tr.q_s[ti * 3 + 2] = q_s[2]; }
        if (tr.q_n) { tr.q_n[ti * 3] = q_n[0]; tr.q_n[ti * 3 + 1] = q_n[1]; tr.q_n[ti * 3 + 2] = q_n[2]; }
    }
    __syncthreads();

    // ---- pass 2: thread = hidden column; the chunk's episodes one after the other (fixed order).  Per class k = argmax
    // Q(s') of the episode: S_k = sum [h > 0] dz2, X0_k / X1_k the same times budget / shares, and dW2's h * dz2; the
    // factor W2[c][k] is applied once per class at the end.
    for (int c = tid; c < H; c += kFanChunk) {
        const float4 tc = tab[c], w2 = W2s[c];
        float S[3] = {0.f, 0.f, 0.f}, X0[3] = {0.f, 0.f, 0.f}, X1[3] = {0.f, 0.f, 0.f}, s2[3] = {0.f, 0.f, 0.f};
#pragma unroll 4
        for (int l = 0; l < count; ++l) {
            const float4 ep = epi[l];                           // (broadcast read: every branch below is uniform)
            const float dz = ep.z;
            if (dz == 0.0f) continue;                           // Q(s)[k] <= 0: the episode's gradient is exactly zero
            const int k = __float_as_int(ep.w);
            const float h = fan_hidden(tc.x, ep.x, ep.y, tc.z, tc.w);
            const float hv = __fmul_rn(h, dz);
            const float m = h > 0.0f ? dz : 0.0f;
            if (k == 0) { s2[0] += hv; S[0] += m; X0[0] = fmaf(ep.x, m, X0[0]); X1[0] = fmaf(ep.y, m, X1[0]); }
            else if (k == 1) { s2[1] += hv; S[1] += m; X0[1] = fmaf(ep.x, m, X0[1]); X1[1] = fmaf(ep.y, m, X1[1]); }
            else { s2[2] += hv; S[2] += m; X0[2] = fmaf(ep.x, m, X0[2]); X1[2] = fmaf(ep.y, m, X1[2]); }
        }
        float* out = a.part + (size_t)blockIdx.x * 6 * H;
        out[c] = fmaf(w2.z, S[2], fmaf(w2.y, S[1], __fmul_rn(w2.x, S[0])));
        out[H + c] = fmaf(w2.z, X0[2], fmaf(w2.y, X0[1], __fmul_rn(w2.x, X0[0])));
        out[2 * H + c] = fmaf(w2.z, X1[2], fmaf(w2.y, X1[1], __fmul_rn(w2.x, X1[0])));
        out[3 * H + c] = s2[0]; out[4 * H + c] = s2[1]; out[5 * H + c] = s2[2];
    }
}

// F3: the gradient as a small dense GEMM over the chunks, dW1[k, c] = sum_j X[j][k] * SdZ[j][c] with X[j][k] = price
// p_t(j)[i + k] (32 consecutive k of a chunk's series are one 128-byte line).  A CTA takes a 32 x 32 tile of (k, c) and
// one of kFanKSplit contiguous ranges of chunks, 64 chunks at a time through shared memory (all loads coalesced and
// independent); thread (kq, c) owns rows 4 kq .. 4 kq + 3 of the tile.  The row-tile index behind the window tiles takes
// the five rows without a price factor: the budget and shares rows of dW1 and the three columns of dW2.  Output: one
// partial per chunk range ([kFanKSplit][D*H] and [kFanKSplit][3H]); apply_kernel adds them in range order.
// (Earlier forms: one thread per parameter over all chunks 274 us at 512 chunks; per parameter and chunk subset 53 us;
// four rows per thread with the price factors staged by dependent scattered loads 49 us, a few CTAs running 5x longer.)
constexpr int kFanTile = 32;
constexpr int kFanStage = 64;
__global__ void __launch_bounds__(256) fan_grad_kernel(const FanArgs a) {
    __shared__ __align__(16) float Xs[kFanStage][kFanTile];
    __shared__ float Ss[kFanStage][kFanTile];
    const Hyper& hp = a.hp;
    const int H = hp.H, wn = hp.wn, D = hp.D;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int c = blockIdx.x * kFanTile + lane;
    const int n_wtiles = (wn + kFanTile - 1) / kFanTile;
    const bool window = (int)blockIdx.y < n_wtiles;
    const int k0 = blockIdx.y * kFanTile;
    const int ks = blockIdx.z;
    const int per = (a.n_chunks + kFanKSplit - 1) / kFanKSplit;
    const int jb = ks * per, je = min(a.n_chunks, jb + per);
    const int32_t istep = *a.d_step;
    float* out1 = a.part1 + (size_t)ks * ((size_t)D * H);
    if (window) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int j0 = jb; j0 < je; j0 += kFanStage) {
            const int nj = min(kFanStage, je - j0);
            __syncthreads();
            for (int jj = w; jj < nj; jj += 8) {                // a warp per chunk: 32 consecutive window rows, 32 columns
                const int64_t t = a.chunks[j0 + jj].x;
                Xs[jj][lane] = (k0 + lane < wn) ? __ldg(a.prices + t * a.T + istep + k0 + lane) : 0.0f;
                Ss[jj][lane] = (c < H) ? __ldg(a.part + (size_t)(j0 + jj) * 6 * H + c) : 0.0f;
            }
            __syncthreads();
#pragma unroll 8
            for (int jj = 0; jj < nj; ++jj) {
                const float v = Ss[jj][lane];
                const float4 x = *reinterpret_cast<const float4*>(&Xs[jj][4 * w]);
                acc[0] = fmaf(x.x, v, acc[0]); acc[1] = fmaf(x.y, v, acc[1]);
                acc[2] = fmaf(x.z, v, acc[2]); acc[3] = fmaf(x.w, v, acc[3]);
            }
        }
        if (c < H) {
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int k = k0 + 4 * w + r;
                if (k < wn) out1[(size_t)k * H + c] = acc[r];
            }
        }
    } else if (w < 5 && c < H) {                                // warp = one of the five rows without a price factor
        float acc = 0.0f;
#pragma unroll 8
        for (int j = jb; j < je; ++j) acc += __ldg(a.part + (size_t)j * 6 * H + (size_t)(1 + w) * H + c);
        if (w < 2) out1[(size_t)(wn + w) * H + c] = acc;
        else a.part2[(size_t)ks * (3 * H) + (size_t)c * 3 + (w - 2)] = acc;
    }
}

#endif  // __CUDACC__
}  // namespace qp
