// qp_ep_c2_phaseB.inl — gradient sum over the pairs' partials + AdaGrad + tf32 hi/lo of the new W1.
// CTA c owns `chunk` consecutive parameters (the last two CTAs: W2), thread t of it parameter p0 + t for the WHOLE
// launch: its weight and AdaGrad accumulator stay in registers from the first step to the last (pb_w, pb_a) and go
// back to global memory once, at the end; per step only what other CTAs read is written — the tf32 hi / lo copy of
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
            constexpr int kMaxRows = 16;                        // partials per thread: n1 <= 128, nsub >= ... see the host check
            const float* srcp = pb_src + pb_off + (size_t)gi * 4;
            float4 v[kMaxRows];
#pragma unroll
            for (int u = 0; u < kMaxRows; ++u) {
                const int j = sub + u * nsub;
                v[u] = j < n1 ? __ldcg(reinterpret_cast<const float4*>(srcp + (size_t)j * pb_stride)) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            float4 s = v[0];
#pragma unroll
            for (int u = 1; u < kMaxRows; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
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
        if (ea.world > 1 && tid < pb_len) {
            // ---- gradient all-reduce over NVLink peer memory, share by share and without a fence: every value travels
            // as one 8-byte store {value, step tag} into the slot this rank owns in every peer's exchange block (posted
            // writes, one NVLink crossing); the receiver spins on its LOCAL slots until the tag is this step's, then adds
            // the values in rank order — the same bits on every rank.  Two buffers by (global) step parity.
            const int64_t p = pb_p0 + tid;
            const unsigned int want = ea.flag_base + (unsigned int)sidx + 1u;
            const size_t slot = (size_t)(want & 1u) * (size_t)ea.world;
            const unsigned long long mine = ((unsigned long long)want << 32) | (unsigned long long)__float_as_uint(g);
            if (x.mc_pdata) {
                // NVLS: one store to the multicast address; the switch delivers it into the slot [parity][this rank][p] of
                // EVERY rank's block (this rank's included): one NVLink crossing of 8 bytes instead of world - 1
                asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" ::"l"(x.mc_pdata + (slot + (size_t)ea.rank) * (size_t)x.p_pad + p),
                             "l"(mine) : "memory");
            } else {
#pragma unroll
                for (int r = 0; r < 8; ++r)
                    if (r < ea.world && r != ea.rank)
                        asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(x.pdata[r] + (slot + (size_t)ea.rank) * (size_t)x.p_pad + p),
                                     "l"(mine) : "memory");
            }
            const unsigned long long* local = x.pdata[ea.rank] + slot * (size_t)x.p_pad + p;
            float pv[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) pv[r] = 0.0f;
            pv[ea.rank & 7] = g;
            unsigned int pending = ((1u << ea.world) - 1u) & ~(1u << ea.rank);
            unsigned long long t0 = 0;
            for (unsigned int spins = 0; pending != 0u; ++spins) {
                unsigned long long vv[8];
#pragma unroll
                for (int r = 0; r < 8; ++r)
                    if ((pending >> r) & 1u)
                        asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(vv[r])
                                     : "l"(local + (size_t)r * (size_t)x.p_pad) : "memory");
#pragma unroll
                for (int r = 0; r < 8; ++r)
                    if (((pending >> r) & 1u) && (unsigned int)(vv[r] >> 32) == want) {
                        pv[r] = __uint_as_float((unsigned int)vv[r]);
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
        stamp(12);
        if (tid < pb_len) {
            const int64_t p = pb_p0 + tid;
            pb_g = g;
            pb_a = __fadd_rn(pb_a, __fmul_rn(g, g));
            pb_w = __fsub_rn(pb_w, __fmul_rn(__fmul_rn(g, hp.lr), __fdiv_rn(1.0f, __fsqrt_rn(pb_a))));
            const int k = (int)(p / H), n = (int)(p - (int64_t)k * H);
            if (k < wn) {                                       // a window row of W1: what the forward streams
                const float hi = et_tf32(pb_w);
                x.w1T_hi[(size_t)n * x.Kp + k] = hi;
                x.w1T_lo[(size_t)n * x.Kp + k] = pb_w - hi;
            } else {                                            // budget / shares rows of W1, W2: read as they are
                ea.params_rw[p] = pb_w;
            }
        }
        // tell the consumers
        __syncthreads();
        if (tid == 0)     // release at GPU scope: cumulative over the CTA's stores the barrier above has ordered
            asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(x.published), "l"(1ull) : "memory");
    }
}
