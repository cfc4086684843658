// qp_ep_c2_phaseB.inl — gradient sum over the pairs' partials + AdaGrad + tf32 hi/lo of the new W1.
// CTA c owns `chunk` consecutive elements of W1 in the TRANSPOSED index space [column][row padded to 4] the pairs'
// partials are stored in (c2_kk; the last two CTAs: W2), thread t of it element p0 + t for the WHOLE launch: its weight and
// AdaGrad accumulator stay in registers from the first step to the last (pb_w, pb_a) and go back to global memory once,
// at the end (one element per column is padding and owns nothing); per step only what other CTAs read is written — the tf32 hi / lo copy of
// a window row of W1, or the parameter itself for the budget / shares rows of W1 and for W2 (the small tables).
//   sum     : the chunk's segment of every partial is read straight from global memory (L2) with 16-byte loads, all
//             of a thread's loads in flight at once: thread (sub, group) adds the partials sub, sub + nsub, ... of
//             one float4 group in that order; the nsub sums meet in shared memory and are added in sub order
//             (a fixed order: bit-reproducible).  (Bulk copies into shared memory did this before: the engine moved
//             the 72 KB of segments at about 20 B/clk per SM.)
//   exchange: several GPUs — as episode_tc_kernel, share by share over NVLink peer memory.
{
    if (pb_len > 0) {
        float* red = reinterpret_cast<float*>(regA);            // [nsub][pitch] partial sums
        const int ng4 = pb_len >> 2;                            // float4 groups of the chunk
        const int nsub = min(8, nthreads / ng4);
        const int gi = tid % ng4, sub = tid / ng4;
        if (sub < nsub) {
            constexpr int kBatch = 8;                           // loads in flight per thread; at most 16 partials per thread (host check)
            const float* srcp = pb_src + pb_off + (size_t)gi * 4;
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
            for (int u0 = 0; u0 * nsub + sub < n1; u0 += kBatch) {
                float4 v[kBatch];
#pragma unroll
                for (int u = 0; u < kBatch; ++u) {
                    const int j = sub + (u0 + u) * nsub;
                    v[u] = j < n1 ? __ldcg(reinterpret_cast<const float4*>(srcp + (size_t)j * pb_stride)) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int u = 0; u < kBatch; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
            }
            *reinterpret_cast<float4*>(red + (size_t)sub * pitch + gi * 4) = s;
        }
        __syncthreads();
        stamp(15);
        cta_stamp(6);
        float g = 0.0f;
        if (tid < pb_len) {
            g = red[tid];
            for (int u = 1; u < nsub; ++u) g += red[(size_t)u * pitch + tid];
        }
        // this thread's element of the transposed index space: column nq, row kq (kq = D .. Kk - 1: padding, nobody's)
        const int nq = (int)(pb_nk >> 16), kq = (int)(pb_nk & 0xffffu);
        const bool real = w2_cta || kq < D;
        if constexpr (kXchg != 0) {
            if (tid < pb_len && real) g = c2_share_exchange<kXchg == 2>(x, g, pb_p0 + tid, sidx, (int)blockIdx.x);
        }
        stamp(12);
        if (tid < pb_len && real) {
            pb_g = g;
            pb_a = __fadd_rn(pb_a, __fmul_rn(g, g));
            pb_w = __fsub_rn(pb_w, __fmul_rn(__fmul_rn(g, hp.lr), __fdiv_rn(1.0f, __fsqrt_rn(pb_a))));
            if (!w2_cta && kq < wn) {                           // a window row of W1: what the forward streams (consecutive
                const float hi = et_tf32(pb_w);                 // threads = consecutive k of one column: coalesced stores)
                x.w1T_hi[(size_t)nq * x.Kp + kq] = hi;
                x.w1T_lo[(size_t)nq * x.Kp + kq] = pb_w - hi;
            } else {                                            // budget / shares rows of W1, W2: read as they are
                ea.params_rw[w2_cta ? P1 + pb_off + tid : (int64_t)kq * H + nq] = pb_w;
            }
        }
        // tell the consumers
        __syncthreads();
        if (tid == 0)     // release at GPU scope: cumulative over the CTA's stores the barrier above has ordered
            asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(x.published), "l"(1ull) : "memory");
    }
}
