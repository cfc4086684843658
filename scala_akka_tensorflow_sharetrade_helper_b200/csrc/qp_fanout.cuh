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

namespace qp {

constexpr int kFanChunk = 256;          // episodes per chunk = threads of fan_row_kernel
constexpr int kFanSeriesPerCta = 4;     // series per CTA of fan_zwin_kernel

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
    float* grad;                        // [P]
    TraceDev trace; int32_t trace_step0;
};

#ifdef __CUDACC__

// F1: thread = hidden column, kFanSeriesPerCta series per CTA; W1 is read once per CTA, coalesced, straight from L2.
__global__ void __launch_bounds__(256) fan_zwin_kernel(const FanArgs a) {
    extern __shared__ __align__(16) float fan_smem[];
    const Hyper& hp = a.hp;
    const int H = hp.H, wn = hp.wn;
    const int32_t istep = *a.d_step;
    const int64_t t0 = (int64_t)blockIdx.x * kFanSeriesPerCta;
    float* pw = fan_smem;                                       // [kFanSeriesPerCta][wn + 1]
    for (int i = threadIdx.x; i < kFanSeriesPerCta * (wn + 1); i += blockDim.x) {
        const int j = i / (wn + 1), k = i - j * (wn + 1);
        const int64_t t = t0 + j;
        float v = 0.0f;
        if (t < a.n_series && (int64_t)istep + k < a.T) v = __ldg(a.prices + t * a.T + istep + k);
        pw[i] = v;
    }
    __syncthreads();
    const int c = threadIdx.x;
    if (c >= H) return;
    float zs[kFanSeriesPerCta], zn[kFanSeriesPerCta];
#pragma unroll
    for (int j = 0; j < kFanSeriesPerCta; ++j) { zs[j] = 0.0f; zn[j] = 0.0f; }
#pragma unroll 4
    for (int k = 0; k < wn; ++k) {
        const float w = __ldg(a.params + (size_t)k * H + c);
#pragma unroll
        for (int j = 0; j < kFanSeriesPerCta; ++j) {
            zs[j] = fmaf(pw[j * (wn + 1) + k], w, zs[j]);
            zn[j] = fmaf(pw[j * (wn + 1) + k + 1], w, zn[j]);
        }
    }
#pragma unroll
    for (int j = 0; j < kFanSeriesPerCta; ++j) {
        const int64_t t = t0 + j;
        if (t < a.n_series) {
            a.zw[(t * 2 + 0) * H + c] = zs[j];
            a.zw[(t * 2 + 1) * H + c] = zn[j];
        }
    }
}

// F2: one chunk (<= 256 consecutive episodes of one series) per CTA.
__global__ void __launch_bounds__(kFanChunk) fan_row_kernel(const FanArgs a) {
    extern __shared__ __align__(16) float fan_smem[];
    const Hyper& hp = a.hp;
    const int H = hp.H, wn = hp.wn, D = hp.D;
    const int tid = threadIdx.x;
    const int4 ch = a.chunks[blockIdx.x];
    const int64_t t = ch.x, e0 = ch.y;
    const int count = ch.z;
    const int32_t istep = *a.d_step;
    const float bias = hp.bias;
    float4* tab = reinterpret_cast<float4*>(fan_smem);          // [H] (zs, zn, wb, ws)
    float4* W2s = tab + H;                                      // [H] (w2[c][0..2], 0)
    float4* epi = W2s + H;                                      // [kFanChunk] (xb, xs, dz2, kn as int bits)
    const float* W1g = a.params;
    const float* W2g = a.params + (size_t)D * H;
    for (int c = tid; c < H; c += kFanChunk) {
        tab[c] = make_float4(a.zw[(t * 2 + 0) * H + c], a.zw[(t * 2 + 1) * H + c], W1g[(size_t)wn * H + c], W1g[(size_t)(wn + 1) * H + c]);
        W2s[c] = make_float4(W2g[c * 3], W2g[c * 3 + 1], W2g[c * 3 + 2], 0.0f);
    }
    __syncthreads();

    // ---- pass 1: thread = episode
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
#pragma unroll 2
        for (int c = 0; c < H; ++c) {
            const float4 tc = tab[c], w2 = W2s[c];
            const float h = wide_hidden(tc.x, xb, xs, tc.z, tc.w, bias);
            pq[0][0] = fmaf(h, w2.x, pq[0][0]); pq[0][1] = fmaf(h, w2.y, pq[0][1]); pq[0][2] = fmaf(h, w2.z, pq[0][2]);
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const float hn = wide_hidden(tc.y, xnb[j], xns[j], tc.z, tc.w, bias);
                pq[1 + j][0] = fmaf(hn, w2.x, pq[1 + j][0]); pq[1 + j][1] = fmaf(hn, w2.y, pq[1 + j][1]);
                pq[1 + j][2] = fmaf(hn, w2.z, pq[1 + j][2]);
            }
        }
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

    // ---- pass 2: thread = hidden column; the chunk's episodes one after the other (fixed order)
    for (int c = tid; c < H; c += kFanChunk) {
        const float4 tc = tab[c], w2 = W2s[c];
        float sdz = 0.0f, x0 = 0.0f, x1 = 0.0f, s0 = 0.0f, s1 = 0.0f, s2 = 0.0f;
#pragma unroll 4
        for (int l = 0; l < count; ++l) {
            const float4 ep = epi[l];                           // (broadcast read)
            const float dz = ep.z;
            const int k = __float_as_int(ep.w);
            const float h = wide_hidden(tc.x, ep.x, ep.y, tc.z, tc.w, bias);
            const float hv = __fmul_rn(h, dz);
            s0 += k == 0 ? hv : 0.0f; s1 += k == 1 ? hv : 0.0f; s2 += k == 2 ? hv : 0.0f;
            const float w2k = k == 0 ? w2.x : (k == 1 ? w2.y : w2.z);
            const float d1 = h > 0.0f ? __fmul_rn(dz, w2k) : 0.0f;
            sdz += d1;
            x0 = fmaf(ep.x, d1, x0);
            x1 = fmaf(ep.y, d1, x1);
        }
        float* out = a.part + (size_t)blockIdx.x * 6 * H;
        out[c] = sdz; out[H + c] = x0; out[2 * H + c] = x1;
        out[3 * H + c] = s0; out[4 * H + c] = s1; out[5 * H + c] = s2;
    }
}

// F3: thread = parameter; the chunks one after the other in four interleaved running sums, (0 + 1) + (2 + 3).
__global__ void __launch_bounds__(256) fan_grad_kernel(const FanArgs a) {
    __shared__ int series_of_chunk[256];
    const Hyper& hp = a.hp;
    const int H = hp.H, wn = hp.wn, D = hp.D;
    const int64_t P1 = (int64_t)D * H, P = P1 + 3 * (int64_t)H;
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int32_t istep = *a.d_step;
    // where this parameter's addend of chunk j lies: part[j * 6H + off], times the price p_t(j)[istep + k] for a window row
    int k = 0, off = 0;
    bool window = false;
    if (p < P1) {
        k = (int)(p / H);
        const int c = (int)(p - (int64_t)k * H);
        window = k < wn;
        off = window ? c : (1 + (k - wn)) * H + c;
    } else if (p < P) {
        const int idx = (int)(p - P1), c = idx / 3, x = idx - 3 * c;
        off = (3 + x) * H + c;
    }
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int j0 = 0; j0 < a.n_chunks; j0 += 256) {
        const int nj = min(256, a.n_chunks - j0);
        __syncthreads();
        if ((int)threadIdx.x < nj) series_of_chunk[threadIdx.x] = a.chunks[j0 + threadIdx.x].x;
        __syncthreads();
        if (p < P) {
            for (int j = 0; j < nj; j += 4) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (j + q < nj) {
                        const float v = __ldg(a.part + (size_t)(j0 + j + q) * 6 * H + off);
                        const float x = window ? __ldg(a.prices + (int64_t)series_of_chunk[j + q] * a.T + istep + k) : 1.0f;
                        acc[q] = window ? fmaf(x, v, acc[q]) : acc[q] + v;
                    }
                }
            }
        }
    }
    if (p < P) a.grad[p] = (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

#endif  // __CUDACC__
}  // namespace qp
