// qp_ep_c2_backward.inl — S5 .. S7 of one env-step: backward contraction, dW2 and the budget / shares rows, the
// dW1 partial out of TMEM into global memory, the next price into the window buffer.
{
    // ---- S5: dW1[256, Nh] = X^T[256, 64] · dZ1[64, Nh] on the tensor core: the window buffer IS the K-major A operand
    // (rows = time, starting at the row of window element 0), contraction over the 64 episodes in two 32-wide blocks.
    // The two 128-row halves complete separately, so that the first leaves TMEM while the second is still computed.
    if (warp == 1 && lane == 0) {
        tc_fence_after();
        for (int hf = 0; hf < halves; ++hf) {
            bool first = first_tile;                            // (kWaves: the later tiles of a step add to the first one's sum)
            for (int kb = 0; kb < 2; ++kb) {
                const uint64_t da = tc_smem_desc(XT + (size_t)kb * 32768 + (size_t)(roff + hf * 128) * 128);
                const uint64_t db_hi = tc_smem_desc(Bbwd_hi + (size_t)kb * TBh), db_lo = tc_smem_desc(Bbwd_lo + (size_t)kb * TBh);
#pragma unroll
                for (int k4 = 0; k4 < 4; ++k4) {
                    tc_mma_tf32(tmem_base + 128u + (uint32_t)(hf * 128), da + 2 * k4, db_hi + 2 * k4, idesc, first ? 0u : 1u);
                    tc_mma_tf32(tmem_base + 128u + (uint32_t)(hf * 128), da + 2 * k4, db_lo + 2 * k4, idesc, 1u);
                    first = false;
                }
            }
            tc_commit(tf_bwd + hf);
        }
        if (halves == 1) tc_commit(tf_bwd + 1);                 // (both barriers complete once per step)
    }
    __syncwarp();
    // meanwhile: dW2[c][t] = sum_e [kn_e == t] h_e[c] dz2_e and the budget / shares rows of dW1, two threads per hidden
    // column (one per lane group), episodes summed sequentially out of shared memory (fixed order)
    if (tid < 2 * Hh) {
        const int c = tid % Hh, gg = tid / Hh;
        const float4 w2 = reinterpret_cast<const float4*>(W2s)[c];
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, x0 = 0.f, x1 = 0.f;
#pragma unroll 8
        for (int l = 0; l < 32; ++l) {
            const int ee = gg * 32 + l;
            const float h = HS[c * kC2EStride + ee];
            const float dz = dz2s[ee];
            const int k = kns[ee];
            const float hv = __fmul_rn(h, dz);
            s0 += k == 0 ? hv : 0.0f; s1 += k == 1 ? hv : 0.0f; s2 += k == 2 ? hv : 0.0f;
            const float w2k = k == 0 ? w2.x : (k == 1 ? w2.y : w2.z);
            const float d1 = h > 0.0f ? __fmul_rn(dz, w2k) : 0.0f;
            x0 = fmaf(XE[ee], d1, x0);
            x1 = fmaf(XE[kC2Envs + ee], d1, x1);
        }
        float* p5 = part5 + (size_t)gg * 5 * Hh + c;
        if (first_tile) { p5[0] = s0; p5[Hh] = s1; p5[2 * Hh] = s2; p5[3 * Hh] = x0; p5[4 * Hh] = x1; }
        else { p5[0] += s0; p5[Hh] += s1; p5[2 * Hh] += s2; p5[3 * Hh] += x0; p5[4 * Hh] += x1; }
    }
    // ---- S6: dW1 partial out of TMEM, stored TRANSPOSED ([column][row], see c2_kk).  Row m of the accumulator is window
    // element m (the A operand started at the window's first row).  A TMEM load hands every lane one ROW of the tile; each
    // warp stages a [32 columns x 32 rows] box in shared memory — lane = row writes word (column, lane): 32 consecutive
    // words per store instruction, conflict-free — and lane 0 hands it to the TMA engine, which writes 128-byte row
    // segments; the tensor map ([pair][half][column][row]) ends at row wn and at column Hh, which clips the boxes.  The
    // staging buffers lie where z1(s') was and behind dZ1: neither h(s) (still read above) nor the operands of the
    // running MMAs are touched.  Warp (q4, part) takes the column boxes part, part + 3 of its 32-row block in both halves.
    {
        const int nbox = (Hh + 31) >> 5;
        uint8_t* buf = warp < 6 ? regA + sm.stg_a + (size_t)warp * 4096 : regA + sm.stg_b + (size_t)(warp - 6) * 4096;
        bool pending = false;
        for (int hf = 0; hf < halves; ++hf) {
            const int k0 = hf * 128 + q4 * 32;
            if (lane == 0) tc_wait(tf_bwd + hf, par);
            __syncwarp();
            tc_fence_after();
            if (k0 >= wn || !last_tile) continue;                 // (kWaves: the sum leaves TMEM after the step's last tile)
            // full boxes part, part + 3, ...; the box of the last few columns (clipped) goes to a different part in each half
            const int nfull = Hh >> 5;
            for (int cbi = part; cbi < nfull + 3; cbi += 3) {
                int cb = cbi;
                if (cbi >= nfull) {
                    if (nbox == nfull || part != 2 - hf) break;
                    cb = nfull;
                }
                uint32_t r[32];
                et_tmem_ld32_nowait(tmem_base + ((uint32_t)(q4 * 32) << 16) + 128u + (uint32_t)(hf * 128 + cb * 32), r);
                if (pending) {                                  // the buffer's previous box has left it
                    if (lane == 0) et_bulk_wait_read0();
                    __syncwarp();
                }
                et_tmem_wait_ld();
                const int nj = min(32, Hh - cb * 32);           // (rows past column Hh are clipped by the tensor map: not staged)
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (j < nj) reinterpret_cast<uint32_t*>(buf)[j * 32 + lane] = r[j];
                et_fence_async_smem();
                __syncwarp();
                if (lane == 0) {
                    c2_tma_store_4d(&tP, buf, k0, cb * 32, (int)crank, (int)cidx);
                    et_bulk_commit();
                }
                pending = true;
            }
        }
        ++mma_it;
    }
    // ---- S7: slide the window: the next price of every episode into the row after the window's last
    if (!kWaves && wsub == 0) {
        const float p_hi = kSplit ? et_tf32(p_next) : p_next;
        *reinterpret_cast<float*>(XT + (size_t)g * 32768 + et_swz(roff + wn + 1, lane)) = (kSplit && g == 1) ? p_next - p_hi : p_hi;
    }
    tc_fence_before();
    __syncthreads();
    stamp(6);
    if (last_tile && tid >= nthreads - 128 && tid - (nthreads - 128) < Hh) {    // the two lane groups' sums, in group order
        const int cl = tid - (nthreads - 128);
        const float* pa = part5 + cl;
        const float* pb = part5 + (size_t)5 * Hh + cl;
        const int c = c0 + cl;
        out2[c * 3 + 0] = pa[0] + pb[0]; out2[c * 3 + 1] = pa[Hh] + pb[Hh]; out2[c * 3 + 2] = pa[2 * Hh] + pb[2 * Hh];
        out[(size_t)c * Kk + wn] = pa[3 * Hh] + pb[3 * Hh];     // (the partial is [column][row])
        out[(size_t)c * Kk + wn + 1] = pa[4 * Hh] + pb[4 * Hh];
    }
}
