// qp_ep_c2_phaseB.inl — gradient sum over the pairs' partials + AdaGrad + tf32 hi/lo of the new W1, as phase B of
// episode_tc_kernel (bulk-copy branch): CTA c owns `chunk` consecutive parameters; their n1 partial segments come in
// as bulk copies, one thread per parameter sums them in shared memory (four interleaved running sums), exchanges the
// sum with the other GPUs over NVLink where there are any, applies AdaGrad and publishes the new weight.
{
        const bool w2_cta = (int)blockIdx.x >= (int)gridDim.x - 2;
        const float* src = w2_cta ? a.dw2_part : ea.part1;
        const int64_t stride = w2_cta ? 3 * (int64_t)H : P1;
        const int64_t off = w2_cta ? (int64_t)((int)blockIdx.x - ((int)gridDim.x - 2)) * chunk2 : (int64_t)blockIdx.x * chunk;
        const int64_t p0 = w2_cta ? P1 + off : off;
        const int len = (int)max((int64_t)0, min((int64_t)(w2_cta ? chunk2 : chunk), stride - off));
        if (len > 0) {
            float* stage = reinterpret_cast<float*>(regA);          // [rows][pitch], over regA + regW
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
                stamp(14);
                __syncthreads();                                    // the stage is free for the next pass
            }
            stamp(15);
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
            stamp(12);
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
}
