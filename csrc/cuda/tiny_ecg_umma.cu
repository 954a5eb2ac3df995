// Fused TinyECG kernels, tcgen05 / TMEM path (config G1: bf16 or fp16 operands, fp32 accumulate).
//
// One CTA (128 threads = the 128 TMEM lanes) walks its samples; every activation of a sample lives
// in shared memory in the *no-swizzle UMMA core-matrix layout*, written there directly by the
// producing CUDA-core code, so the tensor core consumes it without any repacking:
//
//   plane p of a [rows, 16ch] activation = rows x 8 channels, 16 bytes per row, rows contiguous.
//   * as a K-major A operand (rows = M = positions, channels = K): SBO = 128 B (8-row group),
//     LBO = plane stride.  A conv tap is a descriptor whose start address is shifted by
//     tap*16 bytes — the activation is stored ONCE, no im2col, no per-tap copies.
//   * as an MN-major operand (rows = K = positions, channels = M or N): LBO = 128 B,
//     SBO = plane stride — or SBO = 16 B, which makes M-group g the SAME plane shifted by g rows,
//     i.e. one MMA produces the weight gradient of 8 channels x up to 16 taps at once.
//
//   conv2 fwd   D[l, co]      = sum_k  h1[l+k-2, :] . W2[co, :, k]        5 MMAs / 128 positions
//   conv2 dgrad D[l, ci]      = sum_k dz2[l-k+2, :] . W2[:, ci, k]        5 MMAs / 128 positions
//   conv2 wgrad D[(k,c8), co] = sum_l  h1[l+k-2, 8j+c8] * dz2[l, co]      2 MMAs / 16 positions
//   conv1 wgrad + both bias grads D[k, n] = sum_l xcol[l, k] * [dz1 | dz2][l, n]   (xcol = im2col of
//               x with a ones column: 7 taps + bias)                        1 MMA  / 16 positions
//
// Weight-gradient accumulators stay in TMEM across all samples of the CTA and are read out once.
// conv1 forward (Cin = 1, reduction 7) does not fit the tensor core and runs on FFMA, writing bf16
// straight into the operand planes.  MODE_TRAIN ends with the last-CTA SGD/scaler tail, so a local
// step is ONE launch (ref: train_step_G1, TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:117-132).
#include "tiny_umma_common.cuh"

#include <stdlib.h>

namespace cs {

constexpr int kUThreads = 256;   // 8 warps: warps w and w+4 share TMEM lane quadrant w%4 and split the position tiles
long long* g_tiny_prof = nullptr;
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define TSTAMP(k) do { if (a.prof && blockIdx.x == 0 && threadIdx.x == 0) a.prof[k] = clock64(); } while (0)
#define TAILSTAMP(k) do { if (a.prof && threadIdx.x == 0) a.prof[k] = (long long)gtimer(); } while (0)
// Last-CTA tail of the tcgen05 training kernel.  Phase 1 sums the replicas in scratch order with 16-byte
// loads (fully coalesced) into shared memory (row stride 33 to spread banks); phase 2 walks the
// parameters four at a time (16-byte params / momentum accesses) and gathers its gradients from shared
// memory through a precomputed offset table.  Every global load is issued before its first use.
// params / mom must be allocated to a multiple of 4 floats (the model's flat arena is).
__device__ __forceinline__ int sg_index(int o) { return o + (o >> 5); }

__device__ __forceinline__ void umma_sgd_tail(const TinyArgs& a, int P, float* sg /* smem, 33/32 * kScratchStride floats */) {
    constexpr int S4 = kScratchStride / 4;                              // float4 groups per replica (448)
    constexpr int NS4 = (S4 + kUThreads - 1) / kUThreads;               // 4
    constexpr int P4MAX = (kOfsTableLen + 3) / 4;
    constexpr int NV4 = (P4MAX + kUThreads - 1) / kUThreads;            // 4
    const int t = threadIdx.x;
    const int P4 = (P + 3) >> 2;
    const float4 fs = __ldcg(reinterpret_cast<const float4*>(a.fstate));     // loss_sum, scale, last, loss_step
    const int4 cs4 = __ldcg(reinterpret_cast<const int4*>(a.ctrl));          // ticket, growth, steps, skipped
    float4 v[kScratchReps][NS4], p[NV4], m[NV4];
    uint2 of[NV4];
#pragma unroll
    for (int r = 0; r < kScratchReps; ++r)
#pragma unroll
        for (int j = 0; j < NS4; ++j) {
            const int q = j * kUThreads + t;
            v[r][j] = q < S4 ? __ldcg(reinterpret_cast<const float4*>(a.gacc + r * kScratchStride) + q) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
    for (int j = 0; j < NV4; ++j) {
        const int q = j * kUThreads + t;
        p[j] = m[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        of[j] = make_uint2(0, 0);
        if (q < P4) {
            p[j] = reinterpret_cast<const float4*>(a.params)[q];
            m[j] = reinterpret_cast<const float4*>(a.mom)[q];
            of[j] = reinterpret_cast<const uint2*>(g_ofs_table.v)[q];
        }
    }
    const float scale = a.use_scaler ? fs.y : 1.0f;
    const float inv = 1.0f / scale;
#pragma unroll
    for (int j = 0; j < NS4; ++j) {
        const int q = j * kUThreads + t;
        if (q < S4) {
            float4 acc = v[0][j];
#pragma unroll
            for (int r = 1; r < kScratchReps; ++r) { acc.x += v[r][j].x; acc.y += v[r][j].y; acc.z += v[r][j].z; acc.w += v[r][j].w; }
            const int o = sg_index(4 * q);                              // 4 consecutive offsets stay inside one row
            sg[o] = acc.x; sg[o + 1] = acc.y; sg[o + 2] = acc.z; sg[o + 3] = acc.w;
        }
    }
    TAILSTAMP(47);
    __syncthreads();
    TAILSTAMP(48);
    float4 g[NV4];
    int bad = 0;
#pragma unroll
    for (int j = 0; j < NV4; ++j) {
        const int q = j * kUThreads + t, i = 4 * q;
        g[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (q < P4) {
            g[j].x = sg[sg_index(of[j].x & 0xffff)] * inv;
            if (i + 1 < P) g[j].y = sg[sg_index(of[j].x >> 16)] * inv;
            if (i + 2 < P) g[j].z = sg[sg_index(of[j].y & 0xffff)] * inv;
            if (i + 3 < P) g[j].w = sg[sg_index(of[j].y >> 16)] * inv;
        }
        if (!finite_f(g[j].x) || !finite_f(g[j].y) || !finite_f(g[j].z) || !finite_f(g[j].w)) bad = 1;
    }
    TAILSTAMP(42);
    bad = __syncthreads_or(bad);
    TAILSTAMP(43);
    if (!bad) {
#pragma unroll
        for (int j = 0; j < NV4; ++j) {
            const int q = j * kUThreads + t;
            if (q < P4) {
                float4 mm, pp;
                mm.x = a.momentum * m[j].x + (g[j].x + a.wd * p[j].x); pp.x = p[j].x - a.lr * mm.x;
                mm.y = a.momentum * m[j].y + (g[j].y + a.wd * p[j].y); pp.y = p[j].y - a.lr * mm.y;
                mm.z = a.momentum * m[j].z + (g[j].z + a.wd * p[j].z); pp.z = p[j].z - a.lr * mm.z;
                mm.w = a.momentum * m[j].w + (g[j].w + a.wd * p[j].w); pp.w = p[j].w - a.lr * mm.w;
                reinterpret_cast<float4*>(a.mom)[q] = mm;
                reinterpret_cast<float4*>(a.params)[q] = pp;
            }
        }
    }
    TAILSTAMP(44);
    float4* z = reinterpret_cast<float4*>(a.gacc);
    for (int i = t; i < kScratchReps * kScratchStride / 4; i += kUThreads) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    TAILSTAMP(45);
    if (t == 0) {
        float nscale = fs.y;
        int growth = cs4.y;
        if (a.use_scaler) {
            if (bad) { nscale = scale * a.backoff; growth = 0; }
            else if (++growth >= a.interval) { nscale = scale * a.growth; growth = 0; }
        }
        *reinterpret_cast<float4*>(a.fstate) = make_float4(fs.x + fs.w, nscale, fs.w, 0.0f);
        *reinterpret_cast<int4*>(a.ctrl) = make_int4(0, growth, cs4.z + 1, cs4.w + (bad ? 1 : 0));
        a.ctrl[CTRL_FOUND_INF] = bad;
    }
}

template <int TILES, bool FP16, int MODE>
__global__ void __launch_bounds__(kUThreads, TILES <= 4 ? 2 : 1) tiny_umma_kernel(TinyArgs a) {
    using S = USmem<TILES>;
    constexpr int TPT = TILES / 2;                         // position tiles per thread
    extern __shared__ __align__(128) uint8_t smem[];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int row = t & 127, half = t >> 7;                // TMEM lane of this thread / which half of the tiles it owns
    const int L = a.L, C = a.C;
    const int tiles = (L + 127) >> 7;
    TSTAMP(0);
    if (a.prof && threadIdx.x == 0) a.prof[64 + 2 * blockIdx.x] = (long long)gtimer();

    uint64_t* bar_fwd = reinterpret_cast<uint64_t*>(smem + S::OFF_BAR);
    uint64_t* bar_dg = bar_fwd + 1;
    uint64_t* bar_wg = bar_fwd + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S::OFF_BAR + 24);
    int* is_last = reinterpret_cast<int*>(smem + S::OFF_BAR + 28);
    float* f32 = reinterpret_cast<float*>(smem + S::OFF_F32);
    float* w1s = f32;            float* b1s = w1s + 112;   float* b2s = b1s + 16;   float* whs = b2s + 16;
    float* bhs = whs + 128;      float* pooled = bhs + 8;  float* dlog = pooled + 16; float* gj = dlog + 8;
    float* red = gj + 16;
    uint8_t* wt = smem + S::OFF_WT;
    float* xs = reinterpret_cast<float*>(smem + S::OFF_XS);
    const float* prm = a.params;

    // ---- the first window's loads are issued before anything else: they overlap the whole prologue ----
    float xin[TPT];
    int label = 0;
    auto prefetch_sample = [&](int s) {
        const int64_t drow = a.idx ? a.idx[s] : (int64_t)s;          // dataset row of sample s
#pragma unroll
        for (int ii = 0; ii < TPT; ++ii) {
            const int l = (half * TPT + ii) * 128 + row;
            xin[ii] = 0.0f;
            if (l < L) xin[ii] = a.x_bf16 ? __bfloat162float(((const __nv_bfloat16*)a.x)[drow * L + l])
                                          : __ldg((const float*)a.x + drow * L + l);
        }
        if (MODE == MODE_TRAIN && warp == 0) label = (int)a.y[drow];
    };
    if ((int)blockIdx.x < a.B) prefetch_sample(blockIdx.x);

    // ---- one-time set-up ------------------------------------------------------------------------
    {   // zero every operand plane + xs: halo rows / tail rows / slack stay zero (and finite) forever
        uint4* z = reinterpret_cast<uint4*>(smem + S::OFF_XS);
        const int n16 = (S::TOTAL - S::OFF_XS) / 16;
        for (int i = t; i < n16; i += kUThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
    if (t == 0) {
        mbar_init(bar_fwd, 1); mbar_init(bar_dg, 1); mbar_init(bar_wg, 1);
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc(tmem_slot, S::TMEM_COLS);
    // Programmatic dependent launch: everything above (window prefetch, smem clear, TMEM allocation) may run while the
    // previous step's kernel is still finishing; parameters / scaler state are read only after it has
    // completed and flushed.  (No-op when the launch carries no programmatic dependency.)
    if (MODE == MODE_TRAIN) asm volatile("griddepcontrol.wait;" ::: "memory");
    {   // all parameter loads are issued before any is consumed: ONE L2 round trip for the prologue
        const int P = TinyLayout::total(C);
        constexpr int NV = (TinyLayout::WH + 17 * kTinyMaxC + kUThreads - 1) / kUThreads;
        float pv[NV];
#pragma unroll
        for (int j = 0; j < NV; ++j) { const int i = t + j * kUThreads; pv[j] = i < P ? __ldg(prm + i) : 0.0f; }
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            const int i = t + j * kUThreads;
            if (i < TinyLayout::B1) w1s[i] = pv[j];
            else if (i < TinyLayout::W2) b1s[i - TinyLayout::B1] = pv[j];
            else if (i < TinyLayout::B2) {          // W2 [co][ci][k] -> K-major 16x16 operand tiles per tap
                const int q = i - TinyLayout::W2;
                const int co = q / 80, ci = (q / 5) & 15, k = q % 5;
                const uint16_t h = pack1<FP16>(pv[j]);
                // forward tile k:  B[n = co][kk = ci];  dgrad tile 5+k:  B[n = ci][kk = co]
                const int off_f = (co & 7) * 16 + (co >> 3) * 128 + (ci >> 3) * 256 + (ci & 7) * 2;
                const int off_d = (ci & 7) * 16 + (ci >> 3) * 128 + (co >> 3) * 256 + (co & 7) * 2;
                *reinterpret_cast<uint16_t*>(wt + k * 512 + off_f) = h;
                if (MODE != MODE_FWD) *reinterpret_cast<uint16_t*>(wt + (5 + k) * 512 + off_d) = h;
            } else if (i < TinyLayout::WH) b2s[i - TinyLayout::B2] = pv[j];
            else if (i < TinyLayout::WH + 16 * C) whs[i - TinyLayout::WH] = pv[j];
            else if (i < P) bhs[i - TinyLayout::WH - 16 * C] = pv[j];
        }
    }
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    // let the next step's grid start its own prologue on SMs as they drain
    if (MODE == MODE_TRAIN) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    TSTAMP(1);
    const uint32_t tmem_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    const uint32_t col_w = 16 * TILES, col_b = 16 * TILES + 64;

    const uint32_t s_base = smem_u32(smem);
    const uint32_t s_dz1 = s_base + S::P_DZ1, s_dz2 = s_base + S::P_DZ2, s_h1 = s_base + S::P_H1,
                   s_xcol = s_base + S::P_XCOL, s_wt = s_base + S::OFF_WT;
    constexpr uint32_t FMT = FP16 ? FMT_F16 : FMT_BF16;
    constexpr uint32_t ID_CONV = idesc_f16(FMT, 128, 16, MAJOR_K, MAJOR_K);
    constexpr uint32_t ID_WG = idesc_f16(FMT, CS_TINY_WGRAD_M, 16, MAJOR_MN, MAJOR_MN);
    constexpr uint32_t ID_WB = idesc_f16(FMT, CS_TINY_WGRAD_M, 32, MAJOR_MN, MAJOR_MN);

    const float invL = 1.0f / (float)L, invB = 1.0f / (float)a.B;
    const float lscale = (MODE == MODE_TRAIN && a.use_scaler) ? a.fstate[FS_SCALE] : 1.0f;
    float ghead0 = 0.0f, loss_acc = 0.0f;
    uint32_t it = 0;

    for (int s = blockIdx.x; s < a.B; s += gridDim.x, ++it) {
        const uint32_t par = it & 1;
        // ---- P1: prefetched window -> xs (rounded to the operand type, as autocast would feed conv1) ----
        int cur_label = label;
        if (MODE == MODE_TRAIN && warp == 0 && (unsigned)cur_label >= (unsigned)C) {   // out-of-range label: class 0 + error bit
            if (lane == 0) atomicOr(a.ctrl + CTRL_ERROR, 1);
            cur_label = 0;
        }
        if (MODE != MODE_FWD && it > 0) mbar_wait(bar_wg, par ^ 1);   // previous sample's MMAs released the planes
#pragma unroll
        for (int ii = 0; ii < TPT; ++ii) { const int l = (half * TPT + ii) * 128 + row; if (l < L) xs[l + 3] = round_half<FP16>(xin[ii]); }
        if (s + (int)gridDim.x < a.B) prefetch_sample(s + gridDim.x);   // next window loads during this one
        __syncthreads();
        if (it == 0) TSTAMP(2);

        // ---- P2: conv1 + ReLU on FFMA -> h1 planes; im2col(x)+ones -> xcol plane ------------------
        {
            float xv[TPT][7];
#pragma unroll
            for (int i = 0; i < TPT; ++i)
#pragma unroll
                for (int k = 0; k < 7; ++k) xv[i][k] = xs[(half * TPT + i) * 128 + row + k];
            uint32_t hp[TPT][8];
#pragma unroll
            for (int c = 0; c < 16; c += 2) {
                float w0[7], w1[7];
#pragma unroll
                for (int k = 0; k < 7; ++k) { w0[k] = w1s[c * 7 + k]; w1[k] = w1s[(c + 1) * 7 + k]; }
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
                    if (MODE != MODE_FWD)
                        *reinterpret_cast<uint4*>(smem + S::P_XCOL + ro) =
                            make_uint4(pack2<FP16>(xv[i][0], xv[i][1]), pack2<FP16>(xv[i][2], xv[i][3]),
                                       pack2<FP16>(xv[i][4], xv[i][5]), pack2<FP16>(xv[i][6], 1.0f));
                }
            }
        }
        fence_proxy_async_smem();
        tc_fence_before_sync();
        __syncthreads();

        // ---- conv2 forward on the tensor core ---------------------------------------------------------
        if (it == 0) TSTAMP(3);
        if (t == 0) {
            tc_fence_after_sync();
            // descriptors differ only in the start-address field (16-byte units = plane rows)
            const uint64_t a0 = smem_desc(s_h1 + 6 * 16, S::PLANE, 128);
            const uint64_t b0 = smem_desc(s_wt, 256, 128);
            // tap-outer / tile-inner: back-to-back MMAs never accumulate into the same TMEM tile
#pragma unroll
            for (int k = 0; k < 5; ++k)
                for (int i = 0; i < tiles; ++i)
                    mma_f16(tmem_base + i * 16, a0 + (uint64_t)(i * 128 + k), b0 + (uint64_t)(k * 32), ID_CONV, k > 0);
            mma_commit(bar_fwd);
        }
        if (it == 0) TSTAMP(4);
        mbar_wait(bar_fwd, par);
        tc_fence_after_sync();
        if (it == 0) TSTAMP(5);

        // ---- epilogue 1: bias + ReLU + pool partials, ReLU mask kept in registers --------------------
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
        __syncthreads();
        if (t < 16) pooled[t] = (red[t] + red[16 + t] + red[32 + t] + red[48 + t] + red[64 + t] + red[80 + t] + red[96 + t] + red[112 + t]) * invL;
        __syncthreads();
        if (it == 0) TSTAMP(6);

        // ---- head + cross entropy ------------------------------------------------------------------------
        if (warp == 0) {
            // lane c < C owns class c: logit, then a warp-shuffle softmax over the first C lanes
            float lg = -3.0e38f;
            if (lane < C) {
                lg = bhs[lane];
#pragma unroll
                for (int j = 0; j < 16; ++j) lg = fmaf(whs[lane * 16 + j], pooled[j], lg);
            }
            if (MODE == MODE_FWD) {
                if (lane < C) a.logits[(size_t)s * C + lane] = lg;
            } else if (MODE == MODE_TRAIN) {
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
            } else {
                if (lane < C) dlog[lane] = a.dlogits[(size_t)s * C + lane];
            }
        }
        if (MODE == MODE_FWD) continue;      // next iteration starts with a __syncthreads before xs reuse
        __syncthreads();
        if (it == 0) TSTAMP(7);

        // ---- head grads, dp, dz2 = mask * dp / L -> dz2 planes -----------------------------------------------
        if (t < 16 * C) ghead0 = fmaf(dlog[t >> 4], pooled[t & 15], ghead0);
        else if (t < 17 * C) ghead0 += dlog[t - 16 * C];
        if (t < 16) {
            float v = 0.0f;
            for (int c = 0; c < C; ++c) v = fmaf(whs[c * 16 + t], dlog[c], v);
            gj[t] = v * invL;
        }
        __syncthreads();
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
        fence_proxy_async_smem();
        tc_fence_before_sync();
        __syncthreads();

        // ---- conv2 dgrad (reuses the forward accumulator columns) + conv2 wgrad -------------------------
        if (it == 0) TSTAMP(8);
        if (t == 0) {
            tc_fence_after_sync();
            const uint64_t a0 = smem_desc(s_dz2 + 10 * 16, S::PLANE, 128);
            const uint64_t b0 = smem_desc(s_wt + 5 * 512, 256, 128);
#pragma unroll
            for (int k = 0; k < 5; ++k)
                for (int i = 0; i < tiles; ++i)
                    mma_f16(tmem_base + i * 16, a0 + (uint64_t)(i * 128 - k), b0 + (uint64_t)(k * 32), ID_CONV, k > 0);
            mma_commit(bar_dg);
            // conv2 wgrad: M-group g = plane shifted by g rows (SBO = 16 B): rows (g, c8) = dW2[:, 8j+c8, g]
            const int chunks = tiles * 8;
            const uint64_t wa0 = smem_desc(s_h1 + 6 * 16, 128, 16);
            const uint64_t wa1 = smem_desc(s_h1 + S::PLANE + 6 * 16, 128, 16);
            const uint64_t wb = smem_desc(s_dz2 + 8 * 16, 128, S::PLANE);
            // two rotating accumulator sets (even / odd chunks): 4 independent MMAs between dependent ones
#pragma unroll 4
            for (int c = 0; c < chunks; ++c) {
                const uint32_t acc = (it > 0 || c >= 2) ? 1u : 0u;
                const uint32_t d = tmem_base + col_w + (c & 1) * 32;
                mma_f16(d, wa0 + (uint64_t)(c * 16), wb + (uint64_t)(c * 16), ID_WG, acc);
                mma_f16(d + 16, wa1 + (uint64_t)(c * 16), wb + (uint64_t)(c * 16), ID_WG, acc);
            }
        }
        if (it == 0) TSTAMP(9);
        mbar_wait(bar_dg, par);
        tc_fence_after_sync();
        if (it == 0) TSTAMP(10);

        // ---- epilogue 2: dz1 = dh1 * (h1 > 0) -> dz1 planes ---------------------------------------------------
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
        fence_proxy_async_smem();
        tc_fence_before_sync();
        __syncthreads();

        // ---- conv1 wgrad + bias grads: D[k, n] = sum_l xcol[l, k] * [dz1 | dz2][l, n] -----------------------
        if (it == 0) TSTAMP(11);
        if (t == 0) {
            tc_fence_after_sync();
            const int chunks = tiles * 8;
            const uint64_t xa = smem_desc(s_xcol + 8 * 16, 128, 16);
            const uint64_t xb = smem_desc(s_dz1 + 8 * 16, 128, S::PLANE);
#pragma unroll 4
            for (int c = 0; c < chunks; ++c)       // four rotating accumulator sets
                mma_f16(tmem_base + col_b + (c & 3) * 32, xa + (uint64_t)(c * 16), xb + (uint64_t)(c * 16), ID_WB,
                        (it > 0 || c >= 4) ? 1u : 0u);
            mma_commit(bar_wg);
        }
        if (it == 0) TSTAMP(12);
    }

    if (MODE != MODE_FWD && it > 0) {
        // ---- read the weight-gradient accumulators out of TMEM once, add into the global buffer -----------
        mbar_wait(bar_wg, (it - 1) & 1);
        tc_fence_after_sync();
        TSTAMP(13);
        // Both halves of the CTA read accumulators: half h owns columns [16h, 16h+16) of every 32-column
        // accumulator set, sums the rotating sets in registers and issues the reductions for its columns.
        const int r = wgrad_row_of_thread(t);             // accumulator row held by this thread's TMEM lane
        constexpr int kWarpsW = CS_TINY_WGRAD_M == 128 ? 2 : 3;   // lane quadrants whose lanes hold rows < 40
        const int wq = warp & 3, n0 = half * 16;
        if (MODE == MODE_TRAIN) {
            float* g = a.gacc + (blockIdx.x % kScratchReps) * kScratchStride;
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
        } else {
            float* g = a.gacc;                            // MODE_BWD: plain parameter-order gradient
            if (wq < kWarpsW) {
                float v[16];
                load_sum_sets16<2>(tmem_lane + col_w + n0, v);
                const int k = r >> 3, c8 = r & 7;
                if (r >= 0 && r < 40) {
#pragma unroll
                    for (int co = 0; co < 16; ++co)      // columns n0.. = plane j = half: ci = 8*half + c8
                        atomicAdd(g + TinyLayout::W2 + (co * 16 + half * 8 + c8) * 5 + k, v[co]);
                }
            }
            if (wq == 0) {
                float v[16];
                load_sum_sets16<4>(tmem_lane + col_b + n0, v);
                if (half == 0) {                          // columns 0..15: dz1 channels -> dW1 rows / db1
                    if (r >= 0 && r < 7) {
#pragma unroll
                        for (int c = 0; c < 16; ++c) atomicAdd(g + TinyLayout::W1 + c * 7 + r, v[c]);
                    } else if (r == 7) {
#pragma unroll
                        for (int c = 0; c < 16; ++c) atomicAdd(g + TinyLayout::B1 + c, v[c]);
                    }
                } else if (r == 7) {                      // columns 16..31: dz2 channels -> db2
#pragma unroll
                    for (int c = 0; c < 16; ++c) atomicAdd(g + TinyLayout::B2 + c, v[c]);
                }
            }
            if (t < 17 * C) atomicAdd(g + TinyLayout::WH + t, ghead0);
        }
    }
    TSTAMP(14);

    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, S::TMEM_COLS);
    TSTAMP(15);

    if (MODE == MODE_TRAIN) {
        if (t == 0) atomicAdd(a.fstate + FS_LOSS_STEP, loss_acc);
        // release: this CTA's gradient reductions are performed before its ticket becomes visible
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
        __syncthreads();
        if (t == 0) {
            const int tk = atomicAdd(a.ctrl + CTRL_TICKET, 1);
            *is_last = (tk == (int)gridDim.x - 1);
        }
        __syncthreads();
        TSTAMP(16);
        if (a.prof && threadIdx.x == 0) a.prof[64 + 2 * blockIdx.x + 1] = (long long)gtimer();
        if (*is_last) {
            asm volatile("fence.acq_rel.gpu;" ::: "memory"); TAILSTAMP(46); umma_sgd_tail(a, TinyLayout::total(C), reinterpret_cast<float*>(smem + S::OFF_PLANES));
            if (a.prof && threadIdx.x == 0) { a.prof[40] = (long long)gtimer(); a.prof[41] = blockIdx.x; }
        }
    }
}

// ------------------------------------------------------------------------------------------------------
template <int TILES, bool FP16, int MODE>
static int launch_tiles_mode(const TinyArgs& a, cudaStream_t st) {
    using S = USmem<TILES>;
    const size_t smem = S::TOTAL;
    // 2 CTAs per SM co-reside (smem ~69 KB, 128 TMEM columns each at TILES=4): while one waits on its
    // MMAs the other runs its CUDA-core phases
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    const int per_sm = (TILES <= 4) ? 2 : 1;
    const int grid = a.B < num_sms() * per_sm ? a.B : num_sms() * per_sm;
    static size_t configured[64] = {};          // one record per kernel instantiation
    if (smem > configured[dev & 63]) {
        CS_CUDA_OK(cudaFuncSetAttribute(tiny_umma_kernel<TILES, FP16, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev & 63] = smem;
    }
    static const bool use_pdl = []() { const char* e = getenv("CS_TINY_PDL"); return !(e && e[0] == '0'); }();
    if (MODE == MODE_TRAIN && use_pdl) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kUThreads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        CS_CUDA_OK(cudaLaunchKernelEx(&cfg, tiny_umma_kernel<TILES, FP16, MODE>, a));
    } else {
        tiny_umma_kernel<TILES, FP16, MODE><<<grid, kUThreads, smem, st>>>(a);
        CS_CUDA_OK(cudaGetLastError());
    }
    return CS_OK;
}

template <int TILES, bool FP16>
static int launch_tiles(int mode, const TinyArgs& a, cudaStream_t st) {
    if (mode == MODE_FWD) return launch_tiles_mode<TILES, FP16, MODE_FWD>(a, st);
    if (mode == MODE_BWD) return launch_tiles_mode<TILES, FP16, MODE_BWD>(a, st);
    return launch_tiles_mode<TILES, FP16, MODE_TRAIN>(a, st);
}

static int launch_umma(int mode, const TinyArgs& a, cudaStream_t st) {
    CS_REQUIRE(tiny_umma_supported(a.L, a.C), "tcgen05 TinyECG path: unsupported L=%d C=%d", a.L, a.C);
    CS_REQUIRE(a.B >= 1, "bad batch %d", a.B);
    const bool fp16 = a.amp == CS_AMP_FP16;
    const_cast<TinyArgs&>(a).prof = g_tiny_prof;
    if (a.L <= 512) return fp16 ? launch_tiles<4, true>(mode, a, st) : launch_tiles<4, false>(mode, a, st);
    return fp16 ? launch_tiles<8, true>(mode, a, st) : launch_tiles<8, false>(mode, a, st);
}

int tiny_umma_forward(const TinyArgs& a, cudaStream_t st) { return launch_umma(MODE_FWD, a, st); }
int tiny_umma_backward(const TinyArgs& a, cudaStream_t st) { return launch_umma(MODE_BWD, a, st); }
int tiny_umma_train(const TinyArgs& a, cudaStream_t st) { return launch_umma(MODE_TRAIN, a, st); }
size_t tiny_umma_scratch_floats() { return (size_t)kScratchReps * kScratchStride; }
size_t tiny_umma_smem_bytes(int L, int C) { return L <= 512 ? USmem<4>::TOTAL : USmem<8>::TOTAL; }

bool tiny_umma_supported(int L, int C) {
    static int sm100[64] = {};                 // per device: 0 unknown, 1 yes, 2 no
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return false;
    if (sm100[dev] == 0)
        sm100[dev] = (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) == cudaSuccess && major == 10) ? 1 : 2;
    return sm100[dev] == 1 && L >= 8 && L <= 1024 && C >= 1 && C <= kTinyMaxC;
}

}  // namespace cs
