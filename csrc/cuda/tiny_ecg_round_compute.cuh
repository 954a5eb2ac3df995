// Body of the eight compute warps of tiny_round_kernel (textually included inside the kernel: it shares the
// kernel's locals).  See tiny_ecg_round.cu for the design; the per-sample phases are those of tiny_umma_kernel
// (tiny_ecg_umma.cu) with block-wide barriers replaced by a 256-thread named barrier and MMA issue handed to
// the ninth warp through mbarriers.
{
    const float invL = 1.0f / (float)L, invB = 1.0f / (float)B;
    constexpr int S4 = kScratchStride / 4;                       // 16-byte groups per replica (448)
    const int q0 = t, q1 = t + kRCompute;                        // the two groups this thread owns (q1 valid for t < 192)
    const int head_end = kScrGH + 17 * C;
    uint32_t it = 0;
    float loss_round = 0.0f, last_loss = 0.0f;                   // kept by the thread that owns the loss slot
    int nbad = 0, lastbad = 0;

    // mask of the slots of group q that are real parameters for this class count
    auto valid_mask = [&](int q) -> uint32_t {
        uint32_t m = g_valid_table.v[q];
        const int o = 4 * q;
        if (o >= kScrGH) {
#pragma unroll
            for (int e = 0; e < 4; ++e) if (o + e >= head_end) m &= ~(1u << e);
        }
        return m;
    };
    const uint32_t vm0 = valid_mask(q0), vm1 = q1 < S4 ? valid_mask(q1) : 0u;

    // Wait for grid barrier `step`, then sum the gradient replicas of that step and update the shared-memory
    // parameter copy (every CTA does the same arithmetic on the same numbers).
    auto apply_update = [&](int step, bool more_steps) {
        if (t == 0) {
            const uint32_t target = (uint32_t)(step + 1) * (uint32_t)G;
            const unsigned long long t0 = globaltimer_ns();
            while ((int32_t)(ld_acquire_gpu(counter) - target) < 0) {
                if (globaltimer_ns() - t0 > 2000000000ull) { err |= ERR_GRID; break; }
            }
        }
        csync();
        RSTAMP(20);
        const float4* gb = reinterpret_cast<const float4*>(R.gacc + (size_t)(step % kRoundBufs) * kBufFloats);
        float4 v0[kScratchReps], v1[kScratchReps];
#pragma unroll
        for (int r = 0; r < kScratchReps; ++r) {
            v0[r] = __ldcg(gb + r * S4 + q0);
            v1[r] = q1 < S4 ? __ldcg(gb + r * S4 + q1) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        float4 g0 = v0[0], g1 = v1[0];
#pragma unroll
        for (int r = 1; r < kScratchReps; ++r) {
            g0.x += v0[r].x; g0.y += v0[r].y; g0.z += v0[r].z; g0.w += v0[r].w;
            g1.x += v1[r].x; g1.y += v1[r].y; g1.z += v1[r].z; g1.w += v1[r].w;
        }
        if (q1 == kScrLoss / 4) { loss_round += g1.x; last_loss = g1.x; }
        const float inv = 1.0f / scale;
        g0.x = (vm0 & 1u) ? g0.x * inv : 0.0f; g0.y = (vm0 & 2u) ? g0.y * inv : 0.0f;
        g0.z = (vm0 & 4u) ? g0.z * inv : 0.0f; g0.w = (vm0 & 8u) ? g0.w * inv : 0.0f;
        g1.x = (vm1 & 1u) ? g1.x * inv : 0.0f; g1.y = (vm1 & 2u) ? g1.y * inv : 0.0f;
        g1.z = (vm1 & 4u) ? g1.z * inv : 0.0f; g1.w = (vm1 & 8u) ? g1.w * inv : 0.0f;
        int bad = !(finite_f(g0.x) && finite_f(g0.y) && finite_f(g0.z) && finite_f(g0.w) &&
                    finite_f(g1.x) && finite_f(g1.y) && finite_f(g1.z) && finite_f(g1.w));
        bad = csync_or(bad);
        if (!bad) {
            const float mo = ra.momentum, wd = ra.wd, lr = ra.lr;
            {
                float4 p = *reinterpret_cast<float4*>(pm + 4 * q0), m = *reinterpret_cast<float4*>(mm + 4 * q0);
                m.x = mo * m.x + (g0.x + wd * p.x); p.x -= lr * m.x;
                m.y = mo * m.y + (g0.y + wd * p.y); p.y -= lr * m.y;
                m.z = mo * m.z + (g0.z + wd * p.z); p.z -= lr * m.z;
                m.w = mo * m.w + (g0.w + wd * p.w); p.w -= lr * m.w;
                *reinterpret_cast<float4*>(mm + 4 * q0) = m; *reinterpret_cast<float4*>(pm + 4 * q0) = p;
            }
            if (q1 < S4) {
                float4 p = *reinterpret_cast<float4*>(pm + 4 * q1), m = *reinterpret_cast<float4*>(mm + 4 * q1);
                m.x = mo * m.x + (g1.x + wd * p.x); p.x -= lr * m.x;
                m.y = mo * m.y + (g1.y + wd * p.y); p.y -= lr * m.y;
                m.z = mo * m.z + (g1.z + wd * p.z); p.z -= lr * m.z;
                m.w = mo * m.w + (g1.w + wd * p.w); p.w -= lr * m.w;
                *reinterpret_cast<float4*>(mm + 4 * q1) = m; *reinterpret_cast<float4*>(pm + 4 * q1) = p;
            }
        }
        if (ra.use_scaler) {                                   // torch.cuda.amp.GradScaler.update(), decided identically by every CTA
            if (bad) { scale *= ra.backoff; growth = 0; }
            else if (++growth >= ra.interval) { scale *= ra.growth; growth = 0; }
        }
        nbad += bad; lastbad = bad;
        if (more_steps) {
            if (!bad) {                                        // (a skipped step leaves the tiles as they are)
                if (q0 < 320) repack_w2(q0);
                if (q1 < 320) repack_w2(q1);
            }
            // the buffer that was summed two steps ago is free now; it is used again at the step after the next one
            const int next = step + 1;
            if (next >= 2) {
                float4* z = reinterpret_cast<float4*>(R.gacc + (size_t)((next + 1) % kRoundBufs) * kBufFloats);
                const int n4 = kBufFloats / 4;
                const int lo = (int)(((long long)blockIdx.x * n4) / G), hi = (int)(((long long)(blockIdx.x + 1) * n4) / G);
                for (int i = lo + t; i < hi; i += kRCompute) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        fence_proxy_async_smem();
        csync();
        RSTAMP(21);
    };

    for (int step = 0; step < ra.steps; ++step) {
        float ghead0 = 0.0f, loss_acc = 0.0f;
        bool first = true;
        for (int s = blockIdx.x; s < B; s += G, ++it, first = false) {
            const uint32_t par = it & 1;
            // ---- [A] stage the prefetched window: xs, im2col plane; start the next window's loads ----------------
            int cur_label = label;
            if (warp == 0 && (unsigned)cur_label >= (unsigned)C) { err |= ERR_LABEL; cur_label = 0; }
            if (!first && !mbar_wait(bar_wg, par ^ 1)) err |= ERR_MBAR;   // previous sample's MMAs released the planes
            RSTAMP(0);
#pragma unroll
            for (int ii = 0; ii < TPT; ++ii) { const int l = (half * TPT + ii) * 128 + row; if (l < L) xs[l + 3] = round_half<FP16>(xin[ii]); }
            if (s + G < B) prefetch_sample(step, s + G);
            else if (step + 1 < ra.steps) prefetch_sample(step + 1, blockIdx.x);
            csync();
            float xv[TPT][7];
#pragma unroll
            for (int i = 0; i < TPT; ++i)
#pragma unroll
                for (int k = 0; k < 7; ++k) xv[i][k] = xs[(half * TPT + i) * 128 + row + k];
#pragma unroll
            for (int i = 0; i < TPT; ++i) {
                const int l = (half * TPT + i) * 128 + row;
                if (l < L)
                    *reinterpret_cast<uint4*>(smem + S::P_XCOL + (l + 8) * 16) =
                        make_uint4(pack2<FP16>(xv[i][0], xv[i][1]), pack2<FP16>(xv[i][2], xv[i][3]),
                                   pack2<FP16>(xv[i][4], xv[i][5]), pack2<FP16>(xv[i][6], 1.0f));
            }
            RSTAMP(1);
            // ---- [B] previous step's gradients -> parameters (after the grid barrier) ----------------------------------
            if (first && step > 0) apply_update(step - 1, true);
            RSTAMP(2);
            const float lscale = scale;                          // loss scale of this step (1 without a scaler)
            // ---- [C] conv1 + ReLU on FFMA -> h1 planes -------------------------------------------------------------------
            {
                uint32_t hp[TPT][8];
#pragma unroll
                for (int c = 0; c < 16; c += 2) {
                    float w0[7], w1[7];
#pragma unroll
                    for (int k = 0; k < 7; ++k) { w0[k] = w1s[k * 32 + c]; w1[k] = w1s[k * 32 + c + 1]; }
                    const float bb0 = b1s[c], bb1 = b1s[c + 1];
#pragma unroll
                    for (int i = 0; i < TPT; ++i) {
                        float a0 = bb0, a1 = bb1;
#pragma unroll
                        for (int k = 0; k < 7; ++k) { a0 = fmaf(w0[k], xv[i][k], a0); a1 = fmaf(w1[k], xv[i][k], a1); }
                        hp[i][c >> 1] = pack2<FP16>(fmaxf(a0, 0.0f), fmaxf(a1, 0.0f));
                    }
                }
#pragma unroll
                for (int i = 0; i < TPT; ++i) {
                    const int l = (half * TPT + i) * 128 + row;
                    if (l < L) {
                        const int ro = (l + 8) * 16;
                        *reinterpret_cast<uint4*>(smem + S::P_H1 + ro) = make_uint4(hp[i][0], hp[i][1], hp[i][2], hp[i][3]);
                        *reinterpret_cast<uint4*>(smem + S::P_H1 + S::PLANE + ro) = make_uint4(hp[i][4], hp[i][5], hp[i][6], hp[i][7]);
                    }
                }
            }
            handoff(rdy_fwd, lane);
            RSTAMP(3);
            if (!mbar_wait(bar_fwd, par)) err |= ERR_MBAR;
            tc_fence_after_sync();
            RSTAMP(4);

            // ---- epilogue 1: bias + ReLU + pool partials, ReLU mask kept in registers ------------------------------------
            float pool[16];
#pragma unroll
            for (int c = 0; c < 16; ++c) pool[c] = 0.0f;
            uint32_t mask[TPT];
#pragma unroll
            for (int i = 0; i < TPT; ++i) {
                mask[i] = 0;
                const int ti = half * TPT + i;                 // warp-uniform
                if (ti < tiles) {
                    uint32_t v[16];
                    tmem_ld16(tmem_lane + ti * 16, v);
                    tmem_ld_wait();
                    const bool valid = (ti * 128 + row) < L;
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        const float z = __uint_as_float(v[c]) + b2s[c];
                        if (valid && z > 0.0f) { pool[c] += z; mask[i] |= 1u << c; }
                    }
                }
            }
            {
                const float v = warp_reduce16(pool, lane);
                if ((lane & 1) == 0) red[warp * 16 + (lane >> 1)] = v;
            }
            tc_fence_before_sync();
            csync();
            if (t < 16) pooled[t] = (red[t] + red[16 + t] + red[32 + t] + red[48 + t] + red[64 + t] + red[80 + t] + red[96 + t] + red[112 + t]) * invL;
            csync();
            RSTAMP(5);

            // ---- head + cross entropy ----------------------------------------------------------------------------------------
            if (warp == 0) {
                float lg = -3.0e38f;
                if (lane < C) {
                    lg = bhs[lane];
#pragma unroll
                    for (int j = 0; j < 16; ++j) lg = fmaf(whs[lane * 16 + j], pooled[j], lg);
                }
                float mx = lg;
#pragma unroll
                for (int o = 4; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
                float se = lane < C ? __expf(lg - mx) : 0.0f;
#pragma unroll
                for (int o = 4; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
                const float lse = mx + __logf(se);
                const float lg_label = __shfl_sync(0xffffffffu, lg, cur_label & 7);
                if (lane == 0) loss_acc += (lse - lg_label) * invB;
                if (lane < C) dlog[lane] = (__expf(lg - lse) - (lane == cur_label ? 1.0f : 0.0f)) * invB * lscale;
            }
            csync();
            RSTAMP(6);

            // ---- head grads, dp, dz2 = mask * dp / L -> dz2 planes ---------------------------------------------------------
            if (t < 16 * C) ghead0 = fmaf(dlog[t >> 4], pooled[t & 15], ghead0);
            else if (t < 17 * C) ghead0 += dlog[t - 16 * C];
            if (t < 16) {
                float v = 0.0f;
                for (int c = 0; c < C; ++c) v = fmaf(whs[c * 16 + t], dlog[c], v);
                gj[t] = v * invL;
            }
            csync();
            {
                float g[16];
#pragma unroll
                for (int c = 0; c < 16; ++c) g[c] = gj[c];
#pragma unroll
                for (int i = 0; i < TPT; ++i) {
                    const int l = (half * TPT + i) * 128 + row;
                    if (l < L) {
                        uint32_t q[8];
#pragma unroll
                        for (int c = 0; c < 16; c += 2)
                            q[c >> 1] = pack2<FP16>((mask[i] >> c) & 1u ? g[c] : 0.0f, (mask[i] >> (c + 1)) & 1u ? g[c + 1] : 0.0f);
                        const int ro = (l + 8) * 16;
                        *reinterpret_cast<uint4*>(smem + S::P_DZ2 + ro) = make_uint4(q[0], q[1], q[2], q[3]);
                        *reinterpret_cast<uint4*>(smem + S::P_DZ2 + S::PLANE + ro) = make_uint4(q[4], q[5], q[6], q[7]);
                    }
                }
            }
            handoff(rdy_bwd, lane);
            RSTAMP(7);
            if (!mbar_wait(bar_dg, par)) err |= ERR_MBAR;
            tc_fence_after_sync();
            RSTAMP(8);

            // ---- epilogue 2: dz1 = dh1 * (h1 > 0) -> dz1 planes ------------------------------------------------------------
#pragma unroll
            for (int i = 0; i < TPT; ++i) {
                const int ti = half * TPT + i;
                if (ti < tiles) {
                    uint32_t v[16];
                    tmem_ld16(tmem_lane + ti * 16, v);
                    tmem_ld_wait();
                    const int l = ti * 128 + row;
                    if (l < L) {
                        const int ro = (l + 8) * 16;
                        const uint4 ha = *reinterpret_cast<const uint4*>(smem + S::P_H1 + ro);
                        const uint4 hb = *reinterpret_cast<const uint4*>(smem + S::P_H1 + S::PLANE + ro);
                        const uint32_t hw[8] = {ha.x, ha.y, ha.z, ha.w, hb.x, hb.y, hb.z, hb.w};
                        uint32_t q[8];
#pragma unroll
                        for (int c = 0; c < 16; c += 2) {
                            const float2 h = unpack2<FP16>(hw[c >> 1]);
                            q[c >> 1] = pack2<FP16>(h.x > 0.0f ? __uint_as_float(v[c]) : 0.0f,
                                                    h.y > 0.0f ? __uint_as_float(v[c + 1]) : 0.0f);
                        }
                        *reinterpret_cast<uint4*>(smem + S::P_DZ1 + ro) = make_uint4(q[0], q[1], q[2], q[3]);
                        *reinterpret_cast<uint4*>(smem + S::P_DZ1 + S::PLANE + ro) = make_uint4(q[4], q[5], q[6], q[7]);
                    }
                }
            }
            handoff(rdy_wg1, lane);
            RSTAMP(9);
        }

        // ---- end of step: weight-gradient accumulators TMEM -> this step's scratch buffer (16-byte reductions) -------
        if (!mbar_wait(bar_wg, (it - 1) & 1)) err |= ERR_MBAR;
        tc_fence_after_sync();
        RSTAMP(10);
        {
            const int r = wgrad_row_of_thread(t);
            constexpr int kWarpsW = CS_TINY_WGRAD_M == 128 ? 2 : 3;
            const int wq = warp & 3, n0 = half * 16;
            float* g = R.gacc + (size_t)(step % kRoundBufs) * kBufFloats + (blockIdx.x % kScratchReps) * kScratchStride;
            if (wq < kWarpsW) {
                float v[16];
                load_sum_sets16<2>(tmem_lane + col_w + n0, v);
                if (r >= 0 && r < 40) {
#pragma unroll
                    for (int n = 0; n < 16; n += 4) red_add_v4(g + r * 32 + n0 + n, v[n], v[n + 1], v[n + 2], v[n + 3]);
                }
            }
            if (wq == 0) {
                float v[16];
                load_sum_sets16<4>(tmem_lane + col_b + n0, v);
                if (r >= 0 && r < 8) {
#pragma unroll
                    for (int n = 0; n < 16; n += 4) red_add_v4(g + kScrGB + r * 32 + n0 + n, v[n], v[n + 1], v[n + 2], v[n + 3]);
                }
            }
            if (t < 17 * C) atomicAdd(g + kScrGH + t, ghead0);
            if (t == 0) atomicAdd(g + kScrLoss, loss_acc);
        }
        tc_fence_before_sync();
        csync();
        if (t == 0) {                                            // arrive at grid barrier `step` (the wait is in apply_update)
            __threadfence();
            atomicAdd(counter, 1u);
            if (ra.prof && step == ra.steps / 2 && blockIdx.y == 0) ra.prof[64 + blockIdx.x] = (long long)globaltimer_ns();
        }
        RSTAMP(11);
    }

    // ---- round end: CTA 0 finishes the last update, un-permutes and aggregates / writes back ---------------------------------
    if (blockIdx.x == 0) {
        const int step = ra.steps - 1;
        apply_update(step, false);
        float loc[kRoundNP], mo[kRoundNP];
#pragma unroll
        for (int j = 0; j < kRoundNP; ++j) {
            const int i = t + j * kRCompute;
            loc[j] = mo[j] = 0.0f;
            if (i < P) { const int o = g_ofs_table.v[i]; loc[j] = pm[o]; mo[j] = mm[o]; }
        }
        if (ra.comm != nullptr) {
            if (ra.wire_bf16) round_aggregate<true>(ra.comm[er], ra, R, loc, mo, P, t, err);
            else round_aggregate<false>(ra.comm[er], ra, R, loc, mo, P, t, err);
        } else {
#pragma unroll
            for (int j = 0; j < kRoundNP; ++j) {
                const int i = t + j * kRCompute;
                if (i < P) { R.params[i] = loc[j]; R.mom[i] = mo[j]; }
            }
        }
        {   // every other CTA has left: all three gradient buffers go back to zero for the next round
            float4* z = reinterpret_cast<float4*>(R.gacc);
            for (int i = t; i < kRoundBufs * kBufFloats / 4; i += kRCompute) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (t == 0) {
            *counter = 0u;
            R.ctrl[CTRL_GROWTH] = growth;
            R.ctrl[CTRL_STEPS] += ra.steps;
            R.ctrl[CTRL_SKIPPED] += nbad;
            R.ctrl[CTRL_FOUND_INF] = lastbad;
            if (ra.use_scaler) R.fstate[FS_SCALE] = scale;
            if (ra.sample_mode == SAMPLE_PERM) *reinterpret_cast<long long*>(R.ctrl + CTRL_POS) = pos0 + (long long)ra.steps * B;
        }
        if (q1 == kScrLoss / 4) {
            R.fstate[FS_LOSS_SUM] += loss_round;
            R.fstate[FS_LAST_LOSS] = last_loss;
        }
        RSTAMP(12);
    }
}
