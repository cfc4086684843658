// qp_ep_c2_phaseA.inl — one env-step of the pair's 64 episodes up to the stored dW1 / dW2 partial.
// Textually included in episode_c2_kernel's step loop (has_tile is true here).
{
    const uint32_t par = mma_it & 1u;
    // ---- slide: every kC2Slide steps the live rows go back to the top of the buffer.  Source and destination overlap:
    // everything is read into registers first.  (kC2Slide is a multiple of 8, so a row keeps its swizzle phase.)
    if (!kWaves && sidx > 0 && roff == 0) {
        constexpr int kMaxCopy = 12;
        const int n_items = (wn + 1) * 16;                      // (block, row, 16-byte chunk)
        float4 v[kMaxCopy];
#pragma unroll
        for (int i = 0; i < kMaxCopy; ++i) {
            const int it = tid + i * nthreads;
            if (it < n_items) {
                const int blk = it / ((wn + 1) * 8), rem = it - blk * (wn + 1) * 8;
                v[i] = *reinterpret_cast<const float4*>(XT + (size_t)blk * 32768 + (size_t)((rem >> 3) + kC2Slide) * 128 + ((rem & 7) << 4));
            }
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < kMaxCopy; ++i) {
            const int it = tid + i * nthreads;
            if (it < n_items) {
                const int blk = it / ((wn + 1) * 8), rem = it - blk * (wn + 1) * 8;
                *reinterpret_cast<float4*>(XT + (size_t)blk * 32768 + (size_t)(rem >> 3) * 128 + ((rem & 7) << 4)) = v[i];
            }
        }
        et_fence_async_smem();
        __syncthreads();
    }
    float q_s[3] = {0.f, 0.f, 0.f}, q_n[3] = {0.f, 0.f, 0.f};
    int action = 0, explored = 0, a_eff = 2, kn = 0;
    float r32 = 0.0f, target = 0.0f, dz2 = 0.0f, p_next = 0.0f;
    uint32_t rnd[4] = {0u, 0u, 0u, 0u};

    // ---- S0 / S1: forward contraction Z[128, Nh] = X[128, Kp] · W1[Kp, this CTA's columns] on the tensor core.
    // Roles: warp 0 feeds the W1^T ring by TMA, warp 1 issues the MMAs, warps 2..9 transpose the A tiles out of the
    // window buffer (they do not depend on the weights and run ahead of the split barrier), warp 10 loads the small tables.
    if (tid == 0) fine(48);
    if (warp == 0) {
        if (lane == 0) {
            if constexpr (!kWaves) {                            // (kWaves: done before the tile's window is gathered)
                // the new weights are complete once every publisher of the previous step is through
                if (sidx > 0) et_wait_count(x.published, (unsigned long long)sidx * (unsigned long long)n_publishers);
                fence_proxy_async();                            // phase-B writes of W1^T -> TMA reads
                for (int kk = 0; kk < min(kC2WStages, nkb); ++kk) push_w_tiles(kk);
            }
            stamp(1);                                           // (the wait for the other CTAs' phase B ends here)
            cta_stamp(1);
            for (int kk = kC2WStages; kk < nkb; ++kk) push_w_tiles(kk);
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            bool first = true;
            for (int kk = 0; kk < nkb; ++kk) {
                const int kb = (kk + kb_first) % nkb;
                const uint32_t it = ring_it + (uint32_t)kk, s = it % kC2WStages, ph = (it / kC2WStages) & 1u;
                const uint32_t ait = a_it + (uint32_t)kk, sa = ait % kC2AStages, pha = (ait / kC2AStages) & 1u;
                tc_wait(a_full + sa, pha);
                tc_wait(w_full + s, ph);
                fine(8 + kk);
                tc_fence_after();
                uint8_t* st = regW + (size_t)s * 2 * TBh;
                const uint64_t da = tc_smem_desc(regA + (size_t)sa * 16384);
                const uint64_t db_hi = tc_smem_desc(st), db_lo = tc_smem_desc(st + TBh);
#pragma unroll
                for (int k4 = 0; k4 < 4; ++k4) {
                    if (kb * 32 + k4 * 8 < wn) {                // (the columns past the window are zero on both sides)
                        tc_mma_tf32(tmem_base, da + 2 * k4, db_hi + 2 * k4, idesc, first ? 0u : 1u);
                        tc_mma_tf32(tmem_base, da + 2 * k4, db_lo + 2 * k4, idesc, 1u);
                        first = false;
                    }
                }
                tc_commit(w_empty + s);
                tc_commit(a_empty + sa);
            }
            tc_commit(tf_fwd);
            fine(40);
        }
        __syncwarp();                                           // (the other lanes park here instead of polling beside the issuer)
    } else if (warp < 2 + kC2Builders) {
        // Each builder warp owns four (s or s', lane group, 16-byte chunk) items of every k-block.  Row r of a block sits
        // at r * 128 with its 16-byte chunks XOR-ed by (r & 7); rows roff + k + half for consecutive k: the byte offset
        // of the first k-block's element is precomputed per item and element, a k-block further is 32 rows = 4096 bytes.
        const int bw = warp - 2;
        uint32_t src[4][4], dsto[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int item = bw + kC2Builders * i;
            const int half = item & 1, eg = (item >> 1) & 1, ch = item >> 2;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                src[i][j] = (uint32_t)__cvta_generic_to_shared(XT + (size_t)eg * 32768) + et_swz(roff + ch * 4 + j + half, lane);
            const int row = half * 64 + eg * 32 + lane;
            dsto[i] = (uint32_t)row * 128u + (uint32_t)((ch ^ (row & 7)) << 4);
        }
        for (int kk = 0; kk < nkb; ++kk) {
            const int kb = (kk + kb_first) % nkb;
            const uint32_t ait = a_it + (uint32_t)kk, sa = ait % kC2AStages, pha = (ait / kC2AStages) & 1u;
            if (lane == 0) tc_wait(a_empty + sa, pha ^ 1u);
            __syncwarp();
            const uint32_t At = (uint32_t)__cvta_generic_to_shared(regA + (size_t)sa * 16384);
            // rows past the window hold old prices (finite) and meet zero columns of W1^T: no masking; the last rows
            // read stay inside the block (roff + 223 + 1 < 256 + ... see kC2Slide)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                uint32_t v0, v1, v2, v3;
                const uint32_t o = (uint32_t)kb * 4096u;
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v0) : "r"(src[i][0] + o));
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v1) : "r"(src[i][1] + o));
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v2) : "r"(src[i][2] + o));
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v3) : "r"(src[i][3] + o));
                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(At + dsto[i]), "r"(v0), "r"(v1), "r"(v2), "r"(v3) : "memory");
            }
            et_fence_async_smem();
            __syncwarp();
            if (lane == 0) c2_mbar_arrive(a_full + sa);
        }
    } else if (warp == 2 + kC2Builders && first_tile) {
        // the small tables of this CTA's columns (W2, the budget / shares rows of W1): they need the new weights too, so
        // this warp watches the counter itself; every load is issued before the first store
        if (lane == 0 && sidx > 0) et_wait_count(x.published, (unsigned long long)sidx * (unsigned long long)n_publishers);
        __syncwarp();
        constexpr int kTabIter = 4;                             // Hh <= 128
        float t0[kTabIter], t1[kTabIter], t2[kTabIter], t3[kTabIter], t4[kTabIter];
#pragma unroll
        for (int i = 0; i < kTabIter; ++i) {
            const int j = lane + 32 * i, c = c0 + (j < Hh ? j : 0);
            t0[i] = __ldcg(W2g + c * 3); t1[i] = __ldcg(W2g + c * 3 + 1); t2[i] = __ldcg(W2g + c * 3 + 2);
            t3[i] = __ldcg(W1g + (size_t)wn * H + c); t4[i] = __ldcg(W1g + (size_t)(wn + 1) * H + c);
        }
#pragma unroll
        for (int i = 0; i < kTabIter; ++i) {
            const int j = lane + 32 * i;
            if (j < Hh) {
                reinterpret_cast<float4*>(W2s)[j] = make_float4(t0[i], t1[i], t2[i], 0.f);
                Wb[j] = t3[i];
                Ws[j] = t4[i];
            }
        }
    }
    ring_it += (uint32_t)nkb;
    a_it += (uint32_t)nkb;
    // everything of the decision that does not depend on Q
    double p_new, cur;
    {
        const uint64_t env_id = (uint64_t)(a.env_id_base + e);
        philox4x32_10((uint32_t)env_id, (uint32_t)(env_id >> 32), (uint32_t)istep, 1u, (uint32_t)hp.seed, (uint32_t)(hp.seed >> 32), rnd);
        if constexpr (kSplit)      // tf32(x) + (x - tf32(x)) = x, exactly
            p_new = (double)__fadd_rn(*reinterpret_cast<const float*>(XT + et_swz(roff + wn, lane)),
                                      *reinterpret_cast<const float*>(XT + 32768 + et_swz(roff + wn, lane)));
        else
            p_new = (double)*reinterpret_cast<const float*>(XT + (size_t)g * 32768 + et_swz(roff + wn, lane));
        cur = __dadd_rn(budget, __dmul_rn((double)shares, share_value));
        if (!kWaves && wsub == 0 && (int64_t)istep + wn + 1 < a.T) p_next = __ldg(my_row + istep + wn + 1);
    }
    if (lane == 0) tc_wait(tf_fwd, par);                      // (one poller per warp)
    __syncwarp();
    tc_fence_after();
    stamp(2);
    cta_stamp(2);
    if (tid == 0) fine(41);
    if (tid == 0) fine(41);

    // ---- S2: raw z out of TMEM into the [column][episode] tiles.  TMEM lane quarter q holds rows 32q..32q+31 of
    // the A operand: q = 0, 1 -> z1(s) of episodes 0..31 / 32..63; q = 2, 3 -> z1(s')
    {
        float* dstz = (q4 >> 1) ? ZN : HS;
        const int eo = (q4 & 1) * 32 + lane;
        const int n_ld = (Hh + 15) >> 4;
        const int ld0 = part * n_ld / 3, ld1 = (part + 1) * n_ld / 3;
#pragma unroll 1
        for (int l = ld0; l < ld1; ++l) {
            float v[16];
            et_tmem_ld16(tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(l * 16), v);
#pragma unroll
            for (int j = 0; j < 16; ++j)
                if (l * 16 + j < Hh) dstz[(l * 16 + j) * kC2EStride + eo] = v[j];
        }
    }
    tc_fence_before();
    __syncthreads();
    stamp(3);
    if (tid == 0) fine(39);

    // the three states s' the decision can lead to (index = effective action; TrainerChildActor.scala:118-123)
    const double base_b = hp.bookkeeping == 0 ? a.B0 : budget;
    const int32_t base_n = hp.bookkeeping == 0 ? a.N0 : shares;
    const bool can_buy = base_b >= p_new, can_sell = base_n > 0;
    const double nbc[3] = {can_buy ? __dsub_rn(base_b, p_new) : base_b, can_sell ? __dadd_rn(base_b, p_new) : base_b, base_b};
    const int32_t nsc[3] = {can_buy ? base_n + 1 : base_n, can_sell ? base_n - 1 : base_n, base_n};
    // ---- S3: one pass over this CTA's hidden columns: h(s) and the partial sums of Q(s) and of Q(s') for all three
    // outcomes (warp = lane group g x column subset wsub)
    {
        float pq[4][3] = {{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
        const float xb = __double2float_rn(budget), xs = (float)shares;
        const float xnb[3] = {__double2float_rn(nbc[0]), __double2float_rn(nbc[1]), __double2float_rn(nbc[2])};
        const float xns[3] = {(float)nsc[0], (float)nsc[1], (float)nsc[2]};
        // (kSplit: the twelve warps share the columns of the 32 real episodes; z = z(x_hi) + z(x_lo); h goes to both blocks)
        // An outcome that no episode of the warp can reach (nobody can afford the share / nobody holds one) is left out of
        // the sweep: its sums stay zero and are never selected.  The six warps of a lane group and the partner CTA hold
        // the same 32 episodes, so they all decide alike.  (The reference's stateless book-keeping from zero shares never sells.)
        const bool any_buy = __any_sync(0xffffffffu, can_buy), any_sell = __any_sync(0xffffffffu, can_sell);
        auto sweep = [&](auto buy_c, auto sell_c) {
            constexpr bool kBuy = decltype(buy_c)::value, kSell = decltype(sell_c)::value;
#pragma unroll 2
            for (int c = kSplit ? 2 * wsub + g : wsub; c < Hh; c += kSplit ? 2 * kC2Sub : kC2Sub) {
                const float wb = Wb[c], ws = Ws[c];
                const float zn = kSplit ? ZN[c * kC2EStride + lane] + ZN[c * kC2EStride + 32 + lane] : ZN[c * kC2EStride + el];
                const float zs = kSplit ? HS[c * kC2EStride + lane] + HS[c * kC2EStride + 32 + lane] : HS[c * kC2EStride + el];
                const float4 w2 = reinterpret_cast<const float4*>(W2s)[c];
                const float h = et_hidden(zs, xb, xs, wb, ws, bias);
                if constexpr (kSplit) { HS[c * kC2EStride + lane] = h; HS[c * kC2EStride + 32 + lane] = h; }
                else HS[c * kC2EStride + el] = h;               // from here on the tile holds h(s)
                pq[0][0] = fmaf(h, w2.x, pq[0][0]); pq[0][1] = fmaf(h, w2.y, pq[0][1]); pq[0][2] = fmaf(h, w2.z, pq[0][2]);
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    if ((j == 0 && !kBuy) || (j == 1 && !kSell)) continue;
                    const float hn = et_hidden(zn, xnb[j], xns[j], wb, ws, bias);
                    pq[1 + j][0] = fmaf(hn, w2.x, pq[1 + j][0]); pq[1 + j][1] = fmaf(hn, w2.y, pq[1 + j][1]);
                    pq[1 + j][2] = fmaf(hn, w2.z, pq[1 + j][2]);
                }
            }
        };
        if (any_buy && any_sell) sweep(std::true_type{}, std::true_type{});
        else if (any_buy) sweep(std::true_type{}, std::false_type{});
        else if (any_sell) sweep(std::false_type{}, std::true_type{});
        else sweep(std::false_type{}, std::false_type{});
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int t = 0; t < 3; ++t) qpart[((wsub * 4 + j) * 3 + t) * kC2Envs + el] = pq[j][t];
    }
    if (tid == 0) fine(42);
    __syncthreads();
    if (tid == 0) fine(43);
    // this CTA's half of the twelve sums per episode, fixed order, to both CTAs of the pair.
    // One thread per CTA then arrives (release, cluster scope) on the exchange barrier of BOTH CTAs: a CTA's barrier
    // completes when both halves of its copy are written — no cluster barrier (which cost 1.1 K cycles here).
    {
        const uint32_t peer = c2_map_peer(qsum, crank ^ 1u);
        for (int idx = tid; idx < 12 * kC2Envs; idx += nthreads) {
            float pw[kC2Sub];
#pragma unroll
            for (int w = 0; w < kC2Sub; ++w) pw[w] = qpart[w * 12 * kC2Envs + idx];
            float s = (pw[0] + pw[1]) + (pw[2] + pw[3]);
#pragma unroll
            for (int w = 4; w < kC2Sub; w += 2) s += pw[w] + pw[w + 1];
            qsum[crank * 12 * kC2Envs + idx] = s;
            c2_st_cluster(peer + (uint32_t)((crank * 12 * kC2Envs + idx) * 4), s);
        }
        const uint32_t mine = (uint32_t)__cvta_generic_to_shared(qx_bar), theirs = c2_map_peer(qx_bar, crank ^ 1u);
        __syncthreads();                                        // (orders every thread's stores before thread 0's release)
        if (tid == 0) {
            asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(mine) : "memory");
            asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(theirs) : "memory");
            fine(44);
        }
        if (lane == 0) {
            for (uint32_t spins = 0;; ++spins) {
                uint32_t done;
                asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
                             : "=r"(done) : "r"(mine), "r"(par) : "memory");
                if (done) break;
                if (spins > (1u << 26)) __trap();
            }
        }
        __syncwarp();
    }
    if (tid == 0) fine(45);
    {
        // SelectionAction (QDecisionPolicyActor.scala:56-62), decision, reward: every warp of the lane group keeps its copy
#pragma unroll
        for (int t = 0; t < 3; ++t) {
            const float s = c2_qsum<kSplit>(qsum, t, el, lane) + bias;
            q_s[t] = s > 0.0f ? s : 0.0f;
        }
        const float u = (float)(rnd[0] >> 8) * (1.0f / 16777216.0f);
        float thr = __fdiv_rn((float)istep, hp.step_div);
        if (hp.epsilon < thr) thr = hp.epsilon;
        if (u < thr) action = argmax3(q_s);
        else { explored = 1; action = (int)(((uint64_t)rnd[1] * 3u) >> 32); }
        if (action == 0 && can_buy) a_eff = 0;
        else if (action == 1 && can_sell) a_eff = 1;
        const double nb = a_eff == 0 ? nbc[0] : (a_eff == 1 ? nbc[1] : nbc[2]);
        const int32_t ns = a_eff == 0 ? nsc[0] : (a_eff == 1 ? nsc[1] : nsc[2]);
        const double newp = __dadd_rn(nb, __dmul_rn((double)ns, p_new));
        r32 = __double2float_rn(__dsub_rn(newp, cur));
        stamp(4);
        // ---- S4: Q(s') of the outcome taken, TD target; dZ1 as the K-major (episodes contiguous) tf32 hi/lo B operand
#pragma unroll
        for (int t = 0; t < 3; ++t) {
            const int jt = (1 + a_eff) * 3 + t;
            const float s = c2_qsum<kSplit>(qsum, jt, el, lane) + bias;
            q_n[t] = s > 0.0f ? s : 0.0f;
        }
        kn = argmax3(q_n);
        const float qn_k = kn == 0 ? q_n[0] : (kn == 1 ? q_n[1] : q_n[2]);
        const float qs_k = kn == 0 ? q_s[0] : (kn == 1 ? q_s[1] : q_s[2]);
        target = __fadd_rn(r32, __fmul_rn(hp.gamma, qn_k));
        const float dq = -__fmul_rn(2.0f, __fsub_rn(target, qs_k));
        dz2 = (valid && qs_k > 0.0f) ? dq : 0.0f;
        uint8_t* bh = Bbwd_hi + (size_t)g * TBh;
        uint8_t* bl = Bbwd_lo + (size_t)g * TBh;
#pragma unroll 4
        for (int c = wsub; c < Hh; c += kC2Sub) {
            const float d1 = HS[c * kC2EStride + el] > 0.0f ? __fmul_rn(dz2, W2s[c * 4 + kn]) : 0.0f;
            const float hi = et_tf32(d1);
            const uint32_t off = et_swz(c, lane);
            *reinterpret_cast<float*>(bh + off) = hi;
            *reinterpret_cast<float*>(bl + off) = d1 - hi;
        }
        if (tid == 0) fine(46);
        for (int c = Hh + wsub; c < Nh; c += kC2Sub) {          // padding rows of the MMA's N
            const uint32_t off = et_swz(c, lane);
            *reinterpret_cast<float*>(bh + off) = 0.0f;
            *reinterpret_cast<float*>(bl + off) = 0.0f;
        }
        if (wsub == 0) {
            XE[el] = __double2float_rn(budget);
            XE[kC2Envs + el] = (float)shares;
            kns[el] = kn;
            dz2s[el] = (kSplit && g == 1) ? 0.0f : dz2;         // (virtual episodes add nothing to dW2 and the budget / shares rows)
            if (valid && crank == 0 && (!kSplit || g == 0)) {
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
        budget = nb; shares = ns; share_value = p_new;          // carry (TrainerChildActor.scala:99-102)
    }
    if (tid == 0) fine(47);
    et_fence_async_smem();
    __syncthreads();
    stamp(5);
    if (tid == 0) fine(51);
    cta_stamp(3);
#include "qp_ep_c2_backward.inl"
}
