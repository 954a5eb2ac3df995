// Persistent FL-round kernel, fp32 configuration (G0) on the tensor core: tcgen05.mma kind::tf32.
//
// The reference's G0 step is fp32 autograd (train_step_G0, TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:103-114); on an
// Ampere-or-later GPU cuDNN runs its convolutions in TF32 by default, so this is the like-for-like tensor-core path.
// Same structure as tiny_ecg_round.cu (one cooperative launch per round, grid barrier per step, per-CTA SGD on a
// shared-memory parameter copy in scratch order, rotating gradient scratch, fused FedAvg tail, in-kernel sampler) with
// 32-bit operands:
//   * conv2 forward / data gradient: K-major, no swizzle.  A 16-byte operand row holds 4 channels, so the 16-channel
//     activation is FOUR planes and K = 8 per instruction makes 2 MMAs per tap and tile (channel planes {0,1}, {2,3}).
//     h1 and dz2 share the four planes (the forward MMAs are done before dz2 exists).
//   * weight gradients reduce over positions, i.e. they need MN-major operands, and for 32-bit MN-major operands the only
//     layout the tensor core reads is the 128-byte swizzle with a 32-byte base (descriptor layout type 1; measured with
//     cs_debug_mma_layout_probe: the no-swizzle form returns zeros).  So each position also gets one 128-byte ROW in two
//     swizzled arrays:  A row = [h1(l) 16 ch | x taps 0..6, 1 | 0 x8],  B row = [dz2(l) 16 | dz1(l) 16].
//     One MN atom is a whole row (32 values) and the atom stride (LBO) is ONE ROW, so M atom g is the array shifted by g
//     positions = conv tap g.  Per 8 positions one M128 x N32 MMA (taps 0..3) and one M64 x N32 MMA (tap 4) produce ALL
//     parameter gradients of both convolutions at once:  D[(g, a), n]:  a < 16, n < 16: dW2[n][a][g];  g = 2, a = 16 + k:
//     n >= 16: dW1[n-16][k] (k < 7), db1 (k = 7), n < 16, k = 7: db2.  The swizzle is a function of the absolute shared
//     address (bits 7..8 xor into bits 5..6), so row-shifted start addresses need no base offset.
//   * ONE sample slot per CTA (the swizzled arrays are 67 KB each): grid <= 148, a CTA with two samples per step runs
//     them back to back, the weight-gradient MMAs of one overlapping conv1 + forward of the next.
//   * TMEM: 64 conv accumulator columns + 4 rotating sets x (M128 32 columns + M64 32 columns); warps 0-3 issue forward /
//     data gradient (one position tile each), warps 4-7 the weight-gradient chunks, each into its own accumulators (one
//     thread issues one small MMA per ~50 cycles, bench/_mma_probe.py).
// The gradient scratch keeps the bf16 kernel's size and regions; only the W2 block is ordered (tap*16 + ci) x co.
#include "tiny_round_common.cuh"

#include <stdlib.h>
#include <string.h>

namespace cs {

__host__ __device__ constexpr int scratch_offset32(int i) {   // flat parameter index -> scratch offset (tf32 kernel)
    if (i >= TinyLayout::W2 && i < TinyLayout::B2) {
        const int q = i - TinyLayout::W2, co = q / 80, ci = (q / 5) & 15, k = q % 5;
        return (k * 16 + ci) * 16 + co;
    }
    return scratch_offset(i);
}
constexpr OfsTable make_ofs_table32() {
    OfsTable t{};
    for (int i = 0; i < kOfsTableLen; ++i) t.v[i] = (uint16_t)scratch_offset32(i);
    return t;
}
static __device__ const OfsTable g_ofs_table32 = make_ofs_table32();
struct alignas(16) ValidTable32 { uint8_t v[kScratchStride / 4]; };
constexpr ValidTable32 make_valid_table32() {
    ValidTable32 t{};
    for (int i = 0; i < kOfsTableLen; ++i) { const int o = scratch_offset32(i); t.v[o >> 2] |= (uint8_t)(1u << (o & 3)); }
    return t;
}
static __device__ const ValidTable32 g_valid_table = make_valid_table32();     // (same name as the bf16 kernel's: shared update code)

struct RS32 {
    static constexpr int TILES = 4;
    static constexpr int ROWS = TILES * 128 + 16;          // K-major planes: 8 halo rows on top, >= 8 below
    static constexpr int PLANE = ROWS * 16;
    static constexpr int RROWS = TILES * 128 + 24;         // swizzled row arrays: tap-shifted reads reach row 512 + 18
    static constexpr int RBYTES = RROWS * 128;
    static constexpr int OFF_BAR = 0;
    static constexpr int OFF_PM = 128;
    static constexpr int OFF_MM = OFF_PM + kScratchStride * 4;
    static constexpr int OFF_WT = OFF_MM + kScratchStride * 4;              // 10 weight tiles x 1024 B
    static constexpr int OFF_F32 = OFF_WT + 10 * 1024;                      // pooled[16] dlog[8] gj[16] red[128]
    static constexpr int OFF_XS = OFF_F32 + 768;
    static constexpr int XS_LEN = TILES * 128 + 8;
    static constexpr int P_A = (OFF_XS + XS_LEN * 4 + 1023) & ~1023;       // [h1 | xcol | 0] rows, swizzled
    static constexpr int P_B = P_A + RBYTES;                                // [dz2 | dz1] rows, swizzled
    static constexpr int P_KM = P_B + RBYTES;                               // 4 K-major planes: h1, then dz2
    static constexpr int TOTAL = P_KM + 4 * PLANE + 512;
    static constexpr int TMEM_COLS = 512;
    static constexpr int COL_G = 64, SETS = 4;             // weight-gradient set s: M128 at COL_G + 64 s, M64 at COL_G + 64 s + 32
    static_assert(TOTAL <= 227 * 1024 && P_A % 1024 == 0 && RBYTES % 512 == 0, "tf32 round kernel: shared-memory budget");
};

#define RSTAMP(k) do { if (ra.prof && blockIdx.x == 0 && blockIdx.y == 0 && t == 0 && step == ra.steps / 2) ra.prof[k] = clock64(); } while (0)

__global__ void __launch_bounds__(256, 1) tiny_round_tf32_kernel(const __grid_constant__ RoundArgs ra) {
    using S = RS32;
    constexpr int TILES = S::TILES, TPT = TILES / 2, NC = 256, NT = 256;
    extern __shared__ __align__(128) uint8_t smem[];
    const int er = blockIdx.y;
    const RoundRank& R = ra.r[er];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int slot = 0, tt = t, ws = warp;
    const int row = tt & 127, half = (tt >> 7) & 1;
    const int wq = warp & 3;
    const int L = ra.L, C = ra.C, B = ra.B, G = gridDim.x;
    const int V = G, v = blockIdx.x;
    const int tiles_w = (L + 127) >> 7;
    const int tiles = ((tiles_w + TPT - 1) / TPT) * TPT;   // whole groups of tiles, see tiny_ecg_round.cu
    const int P = TinyLayout::total(C);

    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::OFF_BAR);
    uint64_t* bar_fwd = bars;      uint64_t* bar_dg = bars + 1;   uint64_t* bar_wg = bars + 2;
    uint64_t* rdy_fwd = bars + 3;  uint64_t* rdy_bwd = bars + 4;  uint64_t* rdy_wg = bars + 5;
    uint32_t* tmem_word = reinterpret_cast<uint32_t*>(smem + S::OFF_BAR + 112);
    float* pm = reinterpret_cast<float*>(smem + S::OFF_PM);
    float* mm = reinterpret_cast<float*>(smem + S::OFF_MM);
    uint8_t* wt = smem + S::OFF_WT;
    float* f32 = reinterpret_cast<float*>(smem + S::OFF_F32);
    float* pooled = f32;  float* dlog = pooled + 16;  float* gj = dlog + 8;  float* red = gj + 16;
    float* xs = reinterpret_cast<float*>(smem + S::OFF_XS);
    const float* w1s = pm + kScrGB;
    const float* b1s = pm + kScrGB + 7 * 32;
    const float* b2s = b1s + 16;
    const float* whs = pm + kScrGH;
    const float* bhs = whs + 16 * C;
    uint32_t* counter = reinterpret_cast<uint32_t*>(R.ctrl + CTRL_TICKET);

    // ---- sampling: a cursor over this slot's samples (step, s), advanced one sample per prefetch ------------------------
    // In-kernel permutation mode keeps (epoch, offset inside the epoch) incrementally - no division per sample - and
    // re-derives the four key words only when the epoch changes.
    int cur_step = 0, cur_s = v;
    long long cur_off = 0, cur_epoch = 0, usable = 1;
    PermKey pk = {};
    if (ra.sample_mode == SAMPLE_PERM) {
        const long long pos0 = *reinterpret_cast<const long long*>(R.ctrl + CTRL_POS);
        usable = (R.n_rows / B) * B;
        cur_epoch = (pos0 + v) / usable;
        cur_off = (pos0 + v) - cur_epoch * usable;
        pk = perm_key((unsigned long long)R.n_rows, perm_epoch_key(R.seed, (unsigned long long)cur_epoch));
    }
    float xin[TPT];
    int label = 0;
    auto prefetch_next = [&]() {                           // loads sample (cur_step, cur_s), then moves the cursor
        const long long q = (long long)cur_step * B + cur_s;
        int64_t drow;
        if (ra.sample_mode == SAMPLE_IDX) drow = R.idx[q];
        else if (ra.sample_mode == SAMPLE_ROWS) drow = q;
        else drow = (int64_t)perm_apply((unsigned long long)cur_off, (unsigned long long)R.n_rows, pk);
#pragma unroll
        for (int ii = 0; ii < TPT; ++ii) {
            const int l = (half * TPT + ii) * 128 + row;
            xin[ii] = 0.0f;
            if (l < L) xin[ii] = ra.x_bf16 ? __bfloat162float(((const __nv_bfloat16*)R.x)[drow * L + l])
                                           : __ldg((const float*)R.x + drow * L + l);
        }
        if ((tt >> 5) == 0) label = (int)R.y[drow];
        int delta = V;
        if (cur_s + V < B) cur_s += V;
        else { delta = B - cur_s + v; cur_s = v; ++cur_step; }
        if (ra.sample_mode == SAMPLE_PERM) {
            cur_off += delta;
            if (cur_off >= usable) {
                cur_off -= usable; ++cur_epoch;
                pk = perm_key((unsigned long long)R.n_rows, perm_epoch_key(R.seed, (unsigned long long)cur_epoch));
            }
        }
    };
    const bool has_work = v < B;
    if (has_work && ra.steps > 0) prefetch_next();

    // ---- one-time set-up ------------------------------------------------------------------------------------------
    {
        uint4* z = reinterpret_cast<uint4*>(smem + S::OFF_PM);
        const int n16 = (S::TOTAL - S::OFF_PM) / 16;
        for (int i = t; i < n16; i += NT) z[i] = make_uint4(0, 0, 0, 0);
    }
    if (t == 0) {
        mbar_init(bar_fwd, TILES); mbar_init(bar_dg, TILES); mbar_init(bar_wg, S::SETS); // MMA commits: one per issuing warp
        mbar_init(rdy_fwd, 8); mbar_init(rdy_bwd, 8); mbar_init(rdy_wg, 8);              // operands written: one arrival per warp
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc(tmem_word, S::TMEM_COLS);
    float scale = ra.use_scaler ? R.fstate[FS_SCALE] : 1.0f;
    int growth = R.ctrl[CTRL_GROWTH];
    __syncthreads();
    for (int i = t; i < P; i += NT) {                      // state_dict order -> scratch order
        const int o = g_ofs_table32.v[i];
        pm[o] = R.params[i];
        mm[o] = R.mom[i];
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_word;
    const uint32_t tmem_lane = tmem_base + ((uint32_t)(wq * 32) << 16);
    const uint32_t s_base = smem_u32(smem);
    const uint32_t s_a = s_base + S::P_A, s_b = s_base + S::P_B, s_km = s_base + S::P_KM, s_wt = s_base + S::OFF_WT;
    constexpr uint32_t ID_CONV = idesc_f16(FMT_TF32, 128, 16, MAJOR_K, MAJOR_K);
    constexpr uint32_t ID_G128 = idesc_f16(FMT_TF32, 128, 32, MAJOR_MN, MAJOR_MN);
    constexpr uint32_t ID_G64 = idesc_f16(FMT_TF32, 64, 32, MAJOR_MN, MAJOR_MN);
    // 32 bytes [lo | hi] of logical 32-byte chunk `chunk` of row `r` of a swizzled row array
    auto st_row = [&](uint32_t s_arr, int r, int chunk, const float4& lo, const float4& hi) {
        const uint32_t la = s_arr + (uint32_t)r * 128u + (uint32_t)chunk * 32u;
        float4* p = reinterpret_cast<float4*>(smem + ((la ^ ((la >> 2) & 0x60u)) - s_base));
        p[0] = lo; p[1] = hi;
    };

    // W2 (fp32, scratch order: row = tap*16 + ci, column = co) -> K-major 16x16 tf32 operand tiles of 1 KB:
    // element (n, kk) at (n/8)*128 + (kk/4)*256 + (n%8)*16 + (kk%4)*4.  Forward tile k = B[n=co][kk=ci], data-gradient tile
    // 5+k = B[n=ci][kk=co].  Group q holds co0..co0+3 of one (k, ci).
    auto repack_w2 = [&](int q) {
        const float4 wf = *reinterpret_cast<const float4*>(pm + 4 * q);
        const int rw = q >> 2, co0 = (q & 3) * 4;
        const int k = rw >> 4, ci = rw & 15;
        uint8_t* tf = wt + k * 1024 + (co0 >> 3) * 128 + (ci >> 2) * 256 + (co0 & 7) * 16 + (ci & 3) * 4;
        const float4 w = make_float4(to_tf32(wf.x), to_tf32(wf.y), to_tf32(wf.z), to_tf32(wf.w));
        *reinterpret_cast<float*>(tf) = w.x;
        *reinterpret_cast<float*>(tf + 16) = w.y;
        *reinterpret_cast<float*>(tf + 32) = w.z;
        *reinterpret_cast<float*>(tf + 48) = w.w;
        uint8_t* td = wt + (5 + k) * 1024 + (ci >> 3) * 128 + (co0 >> 2) * 256 + (ci & 7) * 16;
        *reinterpret_cast<float4*>(td) = w;
    };
    for (int q = t; q < 320; q += NT) repack_w2(q);
    fence_proxy_async_smem();
    __syncthreads();

    int err = 0;
    // ---- MMA issue, spread over the eight warps (each into accumulators of its own) ----------------------------------------
    auto issue_fwd = [&](uint32_t par) {
        if (ws >= TILES) return;
        if (!mbar_wait(rdy_fwd, par)) err |= ERR_MBAR;
        tc_fence_after_sync();
        if (lane == 0) {
            if (ws < tiles) {
                const uint64_t b0 = smem_desc(s_wt, 256, 128);
#pragma unroll
                for (int k = 0; k < 5; ++k)
#pragma unroll
                    for (int j = 0; j < 2; ++j)            // channel planes {0,1} then {2,3}: K = 8 each
                        mma_tf32(tmem_base + ws * 16, smem_desc(s_km + 2 * j * S::PLANE + (6 + ws * 128 + k) * 16, S::PLANE, 128),
                                 b0 + (uint64_t)(k * 64 + j * 32), ID_CONV, (k | j) > 0);
                mma_commit(bar_fwd);
            } else {
                mbar_arrive(bar_fwd);
            }
        }
        __syncwarp();
    };
    auto issue_dgrad = [&](uint32_t par) {
        if (ws >= TILES) return;
        if (!mbar_wait(rdy_bwd, par)) err |= ERR_MBAR;
        tc_fence_after_sync();
        if (lane == 0) {
            if (ws < tiles) {
                const uint64_t b0 = smem_desc(s_wt + 5 * 1024, 256, 128);
#pragma unroll
                for (int k = 0; k < 5; ++k)
#pragma unroll
                    for (int j = 0; j < 2; ++j)
                        mma_tf32(tmem_base + ws * 16, smem_desc(s_km + 2 * j * S::PLANE + (10 + ws * 128 - k) * 16, S::PLANE, 128),
                                 b0 + (uint64_t)(k * 64 + j * 32), ID_CONV, (k | j) > 0);
                mma_commit(bar_dg);
            } else {
                mbar_arrive(bar_dg);
            }
        }
        __syncwarp();
    };
    // all parameter gradients of conv1 and conv2: per chunk of 8 positions D128[(g, a), n] += A[l + g - 2, a] * B[l, n] for taps
    // g = 0..3 and D64 the same for taps 4, 5 (5 unused).  Warps 4-7, each its own rotating accumulator set.
    auto issue_wgrad = [&](uint32_t par, bool first) {
        if (ws < 4) return;
        if (!mbar_wait(rdy_wg, par)) err |= ERR_MBAR;
        tc_fence_after_sync();
        if (lane == 0) {
            const int set = ws - 4;
            const int chunks = tiles_w * 16;
            const uint64_t a0 = smem_desc_mn32(s_a + 6 * 128, 128, 512), b0 = smem_desc_mn32(s_b + 8 * 128, 128, 512);
            const uint32_t d128 = tmem_base + S::COL_G + set * 64, d64 = d128 + 32;
#pragma unroll 2
            for (int c = set; c < chunks; c += S::SETS) {
                const uint32_t acc = (!first || c >= S::SETS) ? 1u : 0u;
                mma_tf32(d128, a0 + (uint64_t)(c * 64), b0 + (uint64_t)(c * 64), ID_G128, acc);
                mma_tf32(d64, a0 + (uint64_t)(c * 64 + 32), b0 + (uint64_t)(c * 64), ID_G64, acc);
            }
            mma_commit(bar_wg);
        }
        __syncwarp();
    };
    {
        const float invL = 1.0f / (float)L, invB = 1.0f / (float)B;
        constexpr int S4 = kScratchStride / 4;                   // 16-byte groups per replica (448)
        constexpr int NQ = (S4 + NC - 1) / NC;                   // groups per thread in the update (1 with two slots, 2 with one)
        const int head_end = kScrGH + 17 * C;
        uint32_t it = 0;
        float loss_round = 0.0f, last_loss = 0.0f;               // kept by the thread that owns the loss slot
        int nbad = 0, lastbad = 0;

        // mask of the slots of group q that are real parameters for this class count
        uint32_t vm[NQ];
#pragma unroll
        for (int j = 0; j < NQ; ++j) {
            const int q = t + j * NC;
            uint32_t m = q < S4 ? g_valid_table.v[q] : 0u;
            const int o = 4 * q;
            if (o >= kScrGH) {
#pragma unroll
                for (int e = 0; e < 4; ++e) if (o + e >= head_end) m &= ~(1u << e);
            }
            vm[j] = m;
        }

        // Wait for grid barrier `step`, then sum the gradient replicas of that step and update the shared-memory
        // parameter copy (every CTA does the same arithmetic on the same numbers).  Called by ALL compute threads.
        auto apply_update = [&](int step, bool more_steps) {
            if (t == 0) {
                const uint32_t target = (uint32_t)(step + 1) * (uint32_t)G;
                const unsigned long long t0 = globaltimer_ns();
                while ((int32_t)(ld_acquire_gpu(counter) - target) < 0) {
                    if (globaltimer_ns() - t0 > 2000000000ull) { err |= ERR_GRID; break; }
                }
                if (ra.prof && step == ra.steps / 2 && blockIdx.y == 0) ra.prof[1024 + 4 * blockIdx.x + 2] = (long long)globaltimer_ns();
            }
            csync<NC>();
            RSTAMP(20);
            const float4* gb = reinterpret_cast<const float4*>(R.gacc + (size_t)(step % kRoundBufs) * kBufFloats);
            float4 vv[NQ][kScratchReps];
#pragma unroll
            for (int j = 0; j < NQ; ++j) {
                const int q = t + j * NC;
#pragma unroll
                for (int r = 0; r < kScratchReps; ++r)
                    vv[j][r] = q < S4 ? __ldcg(gb + r * S4 + q) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            const float inv = 1.0f / scale;
            float4 g[NQ];
            int bad = 0;
#pragma unroll
            for (int j = 0; j < NQ; ++j) {
                const int q = t + j * NC;
                g[j] = vv[j][0];
#pragma unroll
                for (int r = 1; r < kScratchReps; ++r) { g[j].x += vv[j][r].x; g[j].y += vv[j][r].y; g[j].z += vv[j][r].z; g[j].w += vv[j][r].w; }
                if (q == kScrLoss / 4) { loss_round += g[j].x; last_loss = g[j].x; }
                g[j].x = (vm[j] & 1u) ? g[j].x * inv : 0.0f; g[j].y = (vm[j] & 2u) ? g[j].y * inv : 0.0f;
                g[j].z = (vm[j] & 4u) ? g[j].z * inv : 0.0f; g[j].w = (vm[j] & 8u) ? g[j].w * inv : 0.0f;
                if (!(finite_f(g[j].x) && finite_f(g[j].y) && finite_f(g[j].z) && finite_f(g[j].w))) bad = 1;
            }
            bad = csync_or<NC>(bad);
            if (!bad) {
                const float mo = ra.momentum, wd = ra.wd, lr = ra.lr;
#pragma unroll
                for (int j = 0; j < NQ; ++j) {
                    const int q = t + j * NC;
                    if (q < S4) {
                        float4 p = *reinterpret_cast<float4*>(pm + 4 * q), m = *reinterpret_cast<float4*>(mm + 4 * q);
                        m.x = mo * m.x + (g[j].x + wd * p.x); p.x -= lr * m.x;
                        m.y = mo * m.y + (g[j].y + wd * p.y); p.y -= lr * m.y;
                        m.z = mo * m.z + (g[j].z + wd * p.z); p.z -= lr * m.z;
                        m.w = mo * m.w + (g[j].w + wd * p.w); p.w -= lr * m.w;
                        *reinterpret_cast<float4*>(mm + 4 * q) = m; *reinterpret_cast<float4*>(pm + 4 * q) = p;
                    }
                }
            }
            if (ra.use_scaler) {                                 // torch.cuda.amp.GradScaler.update(), decided identically by every CTA
                if (bad) { scale *= ra.backoff; growth = 0; }
                else if (++growth >= ra.interval) { scale *= ra.growth; growth = 0; }
            }
            nbad += bad; lastbad = bad;
            if (more_steps) {
                if (!bad) {                                      // (a skipped step leaves the tiles as they are)
#pragma unroll
                    for (int j = 0; j < NQ; ++j) { const int q = t + j * NC; if (q < 320) repack_w2(q); }
                }
                // the buffer that was summed two steps ago is free now; it is used again at the step after the next one
                const int next = step + 1;
                if (next >= 2) {
                    float4* z = reinterpret_cast<float4*>(R.gacc + (size_t)((next + 1) % kRoundBufs) * kBufFloats);
                    const int n4 = kBufFloats / 4;
                    const int lo = (int)(((long long)blockIdx.x * n4) / G), hi = (int)(((long long)(blockIdx.x + 1) * n4) / G);
                    for (int i = lo + t; i < hi; i += NC) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            fence_proxy_async_smem();
            csync<NC>();
            RSTAMP(21);
        };

        // [A] stage the prefetched window into xs (fp32 as is - the tensor core drops the low mantissa bits itself) and load
        // this thread's 7-tap neighbourhoods; then start the loads of the window after it.
        float xv[TPT][7];
        int cur_label = 0;
        auto stage_sample = [&](bool more) {
            cur_label = label;
            if (ws == 0 && (unsigned)cur_label >= (unsigned)C) { err |= ERR_LABEL; cur_label = 0; }
#pragma unroll
            for (int ii = 0; ii < TPT; ++ii) { const int l = (half * TPT + ii) * 128 + row; if (l < L) xs[l + 3] = xin[ii]; }
            if (more) prefetch_next();
            slot_sync(slot);
#pragma unroll
            for (int i = 0; i < TPT; ++i)
#pragma unroll
                for (int k = 0; k < 7; ++k) xv[i][k] = xs[(half * TPT + i) * 128 + row + k];
        };

        for (int step = 0; step < ra.steps; ++step) {
            float ghead0 = 0.0f, loss_acc = 0.0f;
            if (ra.prof && t == 0 && step == ra.steps / 2 && blockIdx.y == 0) ra.prof[1024 + 4 * blockIdx.x] = (long long)globaltimer_ns();
            RSTAMP(0);
            if (has_work) stage_sample(v + V < B || step + 1 < ra.steps);
            RSTAMP(1);
            if (step > 0) apply_update(step - 1, true);
            RSTAMP(2);
            const float lscale = scale;
            bool first = true;
            for (int s = v; s < B; s += V, ++it, first = false) {
                const uint32_t par = it & 1;
                if (!first) stage_sample(s + V < B || step + 1 < ra.steps);
                // ---- conv1 + ReLU on FFMA -> h1 in registers and in the four K-major planes ---------------------------------------
                float hr[TPT][16];
                uint32_t hmask[TPT];
#pragma unroll
                for (int i = 0; i < TPT; ++i) hmask[i] = 0;
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    float w[4][7], bb[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        bb[e] = b1s[4 * p + e];
#pragma unroll
                        for (int k = 0; k < 7; ++k) w[e][k] = w1s[k * 32 + 4 * p + e];
                    }
#pragma unroll
                    for (int i = 0; i < TPT; ++i) {
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            float a = bb[e];
#pragma unroll
                            for (int k = 0; k < 7; ++k) a = fmaf(w[e][k], xv[i][k], a);
                            hr[i][4 * p + e] = to_tf32(fmaxf(a, 0.0f));
                            if (a > 0.0f) hmask[i] |= 1u << (4 * p + e);
                        }
                        const int l = (half * TPT + i) * 128 + row;
                        if (l < L)
                            *reinterpret_cast<float4*>(smem + S::P_KM + p * S::PLANE + (l + 8) * 16) =
                                make_float4(hr[i][4 * p], hr[i][4 * p + 1], hr[i][4 * p + 2], hr[i][4 * p + 3]);
                    }
                }
                handoff(rdy_fwd, lane);
                issue_fwd(par);
                RSTAMP(3);
                // while the forward MMAs run: the previous sample's weight-gradient MMAs release the row arrays, then this
                // sample's A rows [h1 | x taps, 1 | 0] go in
                if (!first) { if (!mbar_wait(bar_wg, par ^ 1)) err |= ERR_MBAR; }
#pragma unroll
                for (int i = 0; i < TPT; ++i) {
                    const int l = (half * TPT + i) * 128 + row;
                    if (l < L) {
                        st_row(s_a, l + 8, 0, make_float4(hr[i][0], hr[i][1], hr[i][2], hr[i][3]), make_float4(hr[i][4], hr[i][5], hr[i][6], hr[i][7]));
                        st_row(s_a, l + 8, 1, make_float4(hr[i][8], hr[i][9], hr[i][10], hr[i][11]), make_float4(hr[i][12], hr[i][13], hr[i][14], hr[i][15]));
                        st_row(s_a, l + 8, 2, make_float4(to_tf32(xv[i][0]), to_tf32(xv[i][1]), to_tf32(xv[i][2]), to_tf32(xv[i][3])),
                               make_float4(to_tf32(xv[i][4]), to_tf32(xv[i][5]), to_tf32(xv[i][6]), 1.0f));
                    }
                }
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
                    const int ti = half * TPT + i;             // warp-uniform
                    if (ti < tiles) {
                        uint32_t vr[16];
                        tmem_ld16(tmem_lane + ti * 16, vr);
                        tmem_ld_wait();
                        const bool valid = (ti * 128 + row) < L;
#pragma unroll
                        for (int c = 0; c < 16; ++c) {
                            const float z = __uint_as_float(vr[c]) + b2s[c];
                            if (valid && z > 0.0f) { pool[c] += z; mask[i] |= 1u << c; }
                        }
                    }
                }
                {
                    const float pv = warp_reduce16(pool, lane);
                    if ((lane & 1) == 0) red[ws * 16 + (lane >> 1)] = pv;
                }
                tc_fence_before_sync();
                slot_sync(slot);
                if (tt < 16) pooled[tt] = (red[tt] + red[16 + tt] + red[32 + tt] + red[48 + tt] + red[64 + tt] + red[80 + tt] + red[96 + tt] + red[112 + tt]) * invL;
                slot_sync(slot);
                RSTAMP(5);

                // ---- head + cross entropy ----------------------------------------------------------------------------------------
                if (ws == 0) {
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
                slot_sync(slot);
                RSTAMP(6);

                // ---- head grads, dp, dz2 = mask * dp / L -> four dz2 planes ----------------------------------------------------
                if (tt < 16 * C) ghead0 = fmaf(dlog[tt >> 4], pooled[tt & 15], ghead0);
                else if (tt < 17 * C) ghead0 += dlog[tt - 16 * C];
                if (tt < 16) {
                    float gv = 0.0f;
                    for (int c = 0; c < C; ++c) gv = fmaf(whs[c * 16 + tt], dlog[c], gv);
                    gj[tt] = gv * invL;
                }
                slot_sync(slot);
                {
                    float gg[16];
#pragma unroll
                    for (int c = 0; c < 16; ++c) gg[c] = to_tf32(gj[c]);
#pragma unroll
                    for (int i = 0; i < TPT; ++i) {
                        const int l = (half * TPT + i) * 128 + row;
                        if (l < L) {
                            float4 q[4];
#pragma unroll
                            for (int p = 0; p < 4; ++p) {
                                q[p] = make_float4((mask[i] >> (4 * p)) & 1u ? gg[4 * p] : 0.0f, (mask[i] >> (4 * p + 1)) & 1u ? gg[4 * p + 1] : 0.0f,
                                                   (mask[i] >> (4 * p + 2)) & 1u ? gg[4 * p + 2] : 0.0f, (mask[i] >> (4 * p + 3)) & 1u ? gg[4 * p + 3] : 0.0f);
                                *reinterpret_cast<float4*>(smem + S::P_KM + p * S::PLANE + (l + 8) * 16) = q[p];   // (h1 is dead: forward done)
                            }
                            st_row(s_b, l + 8, 0, q[0], q[1]);
                            st_row(s_b, l + 8, 1, q[2], q[3]);
                        }
                    }
                }
                handoff(rdy_bwd, lane);
                issue_dgrad(par);
                RSTAMP(7);
                if (!mbar_wait(bar_dg, par)) err |= ERR_MBAR;
                tc_fence_after_sync();
                RSTAMP(8);

                // ---- epilogue 2: dz1 = dh1 * (h1 > 0) -> second half of the B rows ---------------------------------------------------
#pragma unroll
                for (int i = 0; i < TPT; ++i) {
                    const int ti = half * TPT + i;
                    if (ti < tiles) {
                        uint32_t vr[16];
                        tmem_ld16(tmem_lane + ti * 16, vr);
                        tmem_ld_wait();
                        const int l = ti * 128 + row;
                        if (l < L) {
                            float d[16];
#pragma unroll
                            for (int c = 0; c < 16; ++c) d[c] = (hmask[i] >> c) & 1u ? to_tf32(__uint_as_float(vr[c])) : 0.0f;
                            st_row(s_b, l + 8, 2, make_float4(d[0], d[1], d[2], d[3]), make_float4(d[4], d[5], d[6], d[7]));
                            st_row(s_b, l + 8, 3, make_float4(d[8], d[9], d[10], d[11]), make_float4(d[12], d[13], d[14], d[15]));
                        }
                    }
                }
                handoff(rdy_wg, lane);
                issue_wgrad(par, first);
                RSTAMP(9);
            }

            // ---- end of step: weight-gradient accumulators TMEM -> this step's scratch buffer (16-byte reductions) -------
            if (has_work) {
                if (!mbar_wait(bar_wg, (it - 1) & 1)) err |= ERR_MBAR;
                tc_fence_after_sync();
                RSTAMP(10);
                float* g = R.gacc + (size_t)(step % kRoundBufs) * kBufFloats + (v % kScratchReps) * kScratchStride;
                // M128 accumulator: row (tap g, a) in lane 32 g + a -> warp quadrant = tap.  M64 accumulator (tap 4): row a in lane a.
                // The two warps of a quadrant split the work: warps 0-3 the conv2 rows of taps 0..3, warp 4 tap 4, warp 6 the
                // conv1 / bias rows (tap 2, a = 16 + k: lanes 80..87).
                auto sum_sets16 = [&](uint32_t col, float (&o)[16]) {
#pragma unroll
                    for (int n = 0; n < 16; ++n) o[n] = 0.0f;
#pragma unroll
                    for (int set = 0; set < S::SETS; ++set) {
                        uint32_t a[16];
                        tmem_ld16(tmem_lane + S::COL_G + set * 64 + col, a);
                        tmem_ld_wait();
#pragma unroll
                        for (int n = 0; n < 16; ++n) o[n] += __uint_as_float(a[n]);
                    }
                };
                if (half == 0 || wq == 0) {                     // dW2[co][ci][tap]: scratch (tap * 16 + ci) * 16 + co
                    float vr[16];
                    sum_sets16(half == 0 ? 0u : 32u, vr);
                    const int tap = half == 0 ? wq : 4;
                    if (lane < 16) {
#pragma unroll
                        for (int n = 0; n < 16; n += 4) red_add_v4(g + (tap * 16 + lane) * 16 + n, vr[n], vr[n + 1], vr[n + 2], vr[n + 3]);
                    }
                }
                if (half == 1 && wq == 2) {                     // lanes 16..23 of quadrant 2: x tap k = lane - 16 (k = 7: the ones column)
                    float v2[16], v1[16];
                    sum_sets16(0u, v2);                          // x dz2: only k = 7 is a parameter gradient (db2)
                    sum_sets16(16u, v1);                         // x dz1: dW1[n][k], db1[n]
                    if (lane >= 16 && lane < 24) {
                        float* gr = g + kScrGB + (lane - 16) * 32;
#pragma unroll
                        for (int n = 0; n < 16; n += 4) {
                            red_add_v4(gr + n, v1[n], v1[n + 1], v1[n + 2], v1[n + 3]);
                            red_add_v4(gr + 16 + n, v2[n], v2[n + 1], v2[n + 2], v2[n + 3]);
                        }
                    }
                }
                if (tt < 17 * C) atomicAdd(g + kScrGH + tt, ghead0);
                if (tt == 0) atomicAdd(g + kScrLoss, loss_acc);
            }
            tc_fence_before_sync();
            csync<NC>();
            if (t == 0) {                                        // arrive at grid barrier `step` (the wait is in apply_update)
                __threadfence();
                atomicAdd(counter, 1u);
                if (ra.prof && step == ra.steps / 2 && blockIdx.y == 0) ra.prof[1024 + 4 * blockIdx.x + 1] = (long long)globaltimer_ns();
            }
            RSTAMP(11);
        }

        // ---- round end: CTA 0 finishes the last update, un-permutes and aggregates / writes back ---------------------------------
        if (blockIdx.x == 0) {
            const int step = ra.steps - 1;
            apply_update(step, false);
            constexpr int NP = RoundTail<NC>::NP;
            float loc[NP], mo[NP];
#pragma unroll
            for (int j = 0; j < NP; ++j) {
                const int i = t + j * NC;
                loc[j] = mo[j] = 0.0f;
                if (i < P) { const int o = g_ofs_table32.v[i]; loc[j] = pm[o]; mo[j] = mm[o]; }
            }
            if (ra.comm != nullptr) {
                if (ra.wire_bf16) round_aggregate<true, NC>(ra.comm[er], ra, R, loc, mo, P, t, err);
                else round_aggregate<false, NC>(ra.comm[er], ra, R, loc, mo, P, t, err);
            } else {
#pragma unroll
                for (int j = 0; j < NP; ++j) {
                    const int i = t + j * NC;
                    if (i < P) { R.params[i] = loc[j]; R.mom[i] = mo[j]; }
                }
            }
            {   // every other CTA has left: all three gradient buffers go back to zero for the next round
                float4* z = reinterpret_cast<float4*>(R.gacc);
                for (int i = t; i < kRoundBufs * kBufFloats / 4; i += NC) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            if (t == 0) {
                *counter = 0u;
                R.ctrl[CTRL_GROWTH] = growth;
                R.ctrl[CTRL_STEPS] += ra.steps;
                R.ctrl[CTRL_SKIPPED] += nbad;
                R.ctrl[CTRL_FOUND_INF] = lastbad;
                if (ra.use_scaler) R.fstate[FS_SCALE] = scale;
                if (ra.sample_mode == SAMPLE_PERM) *reinterpret_cast<long long*>(R.ctrl + CTRL_POS) += (long long)ra.steps * B;
            }
            if (t == (kScrLoss / 4) % NC) {                      // the thread whose update group holds the loss slot
                R.fstate[FS_LOSS_SUM] += loss_round;
                R.fstate[FS_LAST_LOSS] = last_loss;
            }
            RSTAMP(12);
        }
    }
    if (err && lane == 0) atomicOr(R.ctrl + CTRL_ERROR, err);
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(*tmem_word, S::TMEM_COLS);
}

// One sample slot per CTA: the grid is min(B, SMs / ranks) and a CTA takes samples v, v + grid, ... of every step.
int launch_round_tf32(const RoundArgs& ra, int ranks, cudaStream_t st, int* grid_out) {
    using S = RS32;
    const size_t smem = S::TOTAL;
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    static int ready[64] = {};
    if (ready[dev & 63] == 0) {
        CS_CUDA_OK(cudaFuncSetAttribute(tiny_round_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 0;
        CS_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, tiny_round_tf32_kernel, 256, smem));
        CS_REQUIRE(occ >= 1, "train_round (tf32): the kernel does not fit on this device (%zu B shared memory)", smem);
        ready[dev & 63] = 1;
    }
    const int cap = num_sms() / ranks;
    CS_REQUIRE(cap >= 1, "train_round: %d emulated ranks do not fit on one device", ranks);
    // an even share of the batch per CTA (B = 256 on 148 SMs: 128 CTAs x 2 samples, not 108 x 2 + 40 x 1)
    const int per = (ra.B + cap - 1) / cap;
    const int grid = (ra.B + per - 1) / per;
    if (grid_out) *grid_out = grid;
    void* args[] = {(void*)&ra};
    CS_CUDA_OK(cudaLaunchCooperativeKernel((const void*)tiny_round_tf32_kernel, dim3(grid, ranks), dim3(256), args, smem, st));
    return CS_OK;
}

}  // namespace cs
