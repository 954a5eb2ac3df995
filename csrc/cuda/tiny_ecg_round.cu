// Persistent FL-round kernel for TinyECG (tcgen05 / TMEM path): ONE cooperative launch runs all local
// steps of a federated round and ends in the FedAvg aggregation.
//
// Replaces the reference's host-driven round (TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:186-231:
// Bcast, 50 x {~40 library launches + cuda.synchronize + loss.item()}, Allreduce) and this library's own
// first-generation round (50 launches of tiny_umma_kernel + 1 aggregation launch).
//
// Shape of a CTA.  The runtime's occupancy calculator admits ONE TMEM-using CTA per SM (measured on B200: it
// reports 1 for every kernel that executes tcgen05.alloc, whatever its register / shared-memory footprint), and
// a cooperative launch may not exceed that.  The two samples an SM has to work on concurrently - 256 samples on
// 148 SMs - therefore live in ONE CTA as two "slots": 2 x 8 compute warps, each slot with its own operand
// planes, mbarriers and 256 TMEM columns (16 warps, 145 KB of shared memory, all 512 TMEM columns).  While one slot
// waits for its MMAs the other runs its CUDA-core phases.
//
// What stays resident for the whole round: TMEM, mbarriers, operand planes and a shared-memory master copy
// of all parameters and their momentum (one per CTA, shared by both slots).  Per local step:
//   * each slot runs the fused forward/backward of its sample(s) exactly as the one-step kernel does
//     (tiny_ecg_umma.cu); warps publish finished operand planes on mbarriers and the MMA batches are issued by ALL
//     warps of the slot, each into accumulators of its own (one thread issues only one small MMA per ~50 cycles);
//   * weight-gradient accumulators leave TMEM as 16-byte vector reductions into one of THREE rotating global
//     scratch buffers (TMEM-shaped, 4-way replicated); the step's loss rides in a spare slot;
//   * ONE grid-wide barrier (arrive: block barrier + fence + atomic add by one thread; wait: ld.acquire.gpu
//     spin) - split-phase: between arrive and wait the slots already stage the next step's windows;
//   * after the barrier EVERY CTA sums the replicas itself (coalesced 16-byte L2 loads) and applies
//     SGD-momentum / the GradScaler rules to its own shared-memory copy.  Parameters, momentum and gradients
//     share the scratch ("TMEM-shaped") ordering, so the update is purely elementwise: no last-CTA serial
//     tail, no gather, no global parameter round trip.  All CTAs read the same totals in the same order, so
//     their copies stay bit-identical without any further communication.
//   * a buffer used at step s is cleared (distributed over CTAs) during step s+2 and reused at s+3: no second
//     barrier is ever needed.
// At the end CTA 0 un-permutes its copy and either writes it back or - fused aggregation - pushes the
// weighted delta to every peer's arena over NVLink, raises/awaits the system-scope flags and applies the
// server update (same protocol and arena layout as fedavg_oneshot_kernel, so both are interchangeable).
// Batch indices come from an index array, from contiguous rows (staged blocks) or from an in-kernel
// counter-based permutation (a bijection on [0, N) keyed by seed and epoch; ref semantics:
// Module_3/shard_dataset.py:118-136 per-epoch randperm, ragged tail dropped).
#include "tiny_round_common.cuh"

#include <stdlib.h>
#include <string.h>

namespace cs {

// which slots of a 4-float scratch group hold real parameters (class count kTinyMaxC; the head region is
// clipped to 17*C at run time).  Everything else - padding, the x-tap x dz2 cross terms of the bias MMA,
// the loss slot - must not enter the update or the inf check.
struct alignas(16) ValidTable { uint8_t v[kScratchStride / 4]; };
constexpr ValidTable make_valid_table() {
    ValidTable t{};
    for (int i = 0; i < kOfsTableLen; ++i) { const int o = scratch_offset(i); t.v[o >> 2] |= (uint8_t)(1u << (o & 3)); }
    return t;
}
static __device__ const ValidTable g_valid_table = make_valid_table();

template <int TILES> struct RSmem {
    static constexpr int SLOTS = round_slots(TILES);
    static constexpr int ROWS = TILES * 128 + 16;          // 8 halo rows on top, >= 8 below
    static constexpr int PLANE = ROWS * 16;                // bytes
    // shared by both slots
    static constexpr int OFF_BAR = 0;                      // per slot 6 mbarriers (48 B) at slot * 48; TMEM base word at 112
    static constexpr int OFF_PM = 128;                     // parameters, scratch order
    static constexpr int OFF_MM = OFF_PM + kScratchStride * 4;              // momentum, scratch order
    static constexpr int OFF_WT = OFF_MM + kScratchStride * 4;              // 10 weight tiles x 512 B
    static constexpr int OFF_SLOT = OFF_WT + 10 * 512;
    // per slot (offsets inside the slot's region)
    static constexpr int S_F32 = 0;                        // pooled[16] dlog[8] gj[16] red[128]
    static constexpr int S_XS = 768;
    static constexpr int XS_LEN = TILES * 128 + 8;
    static constexpr int S_PLANES = (S_XS + XS_LEN * 4 + 127) & ~127;
    static constexpr int P_DZ1 = S_PLANES, P_DZ2 = P_DZ1 + 2 * PLANE, P_H1 = P_DZ2 + 2 * PLANE, P_XCOL = P_H1 + 2 * PLANE;
    static constexpr int SLOT_BYTES = (P_XCOL + PLANE + 512 + 127) & ~127;  // slack: shifted operands read past the end
    static constexpr int TOTAL = OFF_SLOT + SLOTS * SLOT_BYTES;
    static constexpr int SLOT_COLS = (16 * TILES + 64 + 128) <= 256 ? 256 : 512;   // TMEM columns of one slot
    static constexpr int TMEM_COLS = SLOTS * SLOT_COLS;
    static_assert(TMEM_COLS <= 512 && OFF_SLOT % 128 == 0 && TOTAL <= 227 * 1024, "round kernel: TMEM / shared-memory budget");
};

#define RSTAMP(k) do { if (ra.prof && blockIdx.x == 0 && blockIdx.y == 0 && t == 0 && step == ra.steps / 2) ra.prof[k] = clock64(); } while (0)

template <int TILES, bool FP16>
__global__ void __launch_bounds__(round_threads(TILES), 1) tiny_round_kernel(const __grid_constant__ RoundArgs ra) {
    using S = RSmem<TILES>;
    constexpr int SLOTS = S::SLOTS;
    constexpr int NC = SLOTS * kSlotThreads;               // threads of the CTA
    constexpr int NT = round_threads(TILES);
    constexpr int TPT = TILES / 2;                         // position tiles per compute thread
    extern __shared__ __align__(128) uint8_t smem[];
    const int er = blockIdx.y;
    const RoundRank& R = ra.r[er];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int slot = t >> 8;
    const int tt = t & 255;                                // thread index inside the slot
    const int row = tt & 127, half = (tt >> 7) & 1;        // TMEM lane of this thread / which half of the tiles it owns
    const int wq = warp & 3;                               // TMEM lane quadrant this warp may read
    const int L = ra.L, C = ra.C, B = ra.B, G = gridDim.x;
    const int V = SLOTS * G;                               // sample slots ("workers") of this client
    const int v = blockIdx.x * SLOTS + slot;               // this slot's first sample of every step
    const int tiles_w = (L + 127) >> 7;                    // position tiles that hold data (the weight-gradient chunks cover these)
    // Forward / data-gradient MMAs and the epilogues' TMEM loads run over whole groups of TPT tiles (the tiles one thread
    // owns): with a partly used group (3 of 4 tiles at L = 360) the epilogues read wrong accumulator values on B200 although
    // every load is guarded by a warp-uniform test; the extra tile costs 10 of ~136 MMAs on zero rows.
    const int tiles = ((tiles_w + TPT - 1) / TPT) * TPT;
    const int P = TinyLayout::total(C);

    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::OFF_BAR) + 6 * slot;
    uint64_t* bar_fwd = bars;                              // MMA -> compute (tcgen05.commit)
    uint64_t* bar_dg = bars + 1;
    uint64_t* bar_wg = bars + 2;
    uint64_t* rdy_fwd = bars + 3;                          // operand planes written (8 warp arrivals)
    uint64_t* rdy_bwd = bars + 4;
    uint64_t* rdy_wg1 = bars + 5;
    uint32_t* tmem_word = reinterpret_cast<uint32_t*>(smem + S::OFF_BAR + 112);
    float* pm = reinterpret_cast<float*>(smem + S::OFF_PM);
    float* mm = reinterpret_cast<float*>(smem + S::OFF_MM);
    uint8_t* wt = smem + S::OFF_WT;
    uint8_t* sl = smem + S::OFF_SLOT + slot * S::SLOT_BYTES;           // this slot's private region
    float* f32 = reinterpret_cast<float*>(sl + S::S_F32);
    float* pooled = f32;  float* dlog = pooled + 16;  float* gj = dlog + 8;  float* red = gj + 16;
    float* xs = reinterpret_cast<float*>(sl + S::S_XS);
    const float* w1s = pm + kScrGB;                        // W1[c][k] at k*32 + c
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
    const bool has_work = v < B;                           // (only the last slot of an odd batch can be idle)
    if (has_work && ra.steps > 0) prefetch_next();

    // ---- one-time set-up ------------------------------------------------------------------------------------------
    {   // zero the parameter copies, every operand plane and xs (halo / tail rows stay zero forever)
        uint4* z = reinterpret_cast<uint4*>(smem + S::OFF_PM);
        const int n16 = (S::TOTAL - S::OFF_PM) / 16;
        for (int i = t; i < n16; i += NT) z[i] = make_uint4(0, 0, 0, 0);
    }
    if (t < SLOTS) {
        uint64_t* b = reinterpret_cast<uint64_t*>(smem + S::OFF_BAR) + 6 * t;
        mbar_init(b, TILES); mbar_init(b + 1, TILES); mbar_init(b + 2, 8);     // MMA commits: one per issuing warp
        mbar_init(b + 3, 8); mbar_init(b + 4, 8); mbar_init(b + 5, 8);         // operand planes ready: one arrival per warp
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc(tmem_word, S::TMEM_COLS);
    float scale = ra.use_scaler ? R.fstate[FS_SCALE] : 1.0f;
    int growth = R.ctrl[CTRL_GROWTH];
    __syncthreads();
    for (int i = t; i < P; i += NT) {                      // state_dict order -> scratch order
        const int o = g_ofs_table.v[i];
        pm[o] = R.params[i];
        mm[o] = R.mom[i];
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_word + (uint32_t)(slot * S::SLOT_COLS);
    const uint32_t tmem_lane = tmem_base + ((uint32_t)(wq * 32) << 16);
    const uint32_t col_w = 16 * TILES, col_b = 16 * TILES + 64;
    const uint32_t s_sl = smem_u32(sl);
    const uint32_t s_dz1 = s_sl + S::P_DZ1, s_dz2 = s_sl + S::P_DZ2, s_h1 = s_sl + S::P_H1, s_xcol = s_sl + S::P_XCOL,
                   s_wt = smem_u32(wt);
    constexpr uint32_t FMT = FP16 ? FMT_F16 : FMT_BF16;
    constexpr uint32_t ID_CONV = idesc_f16(FMT, 128, 16, MAJOR_K, MAJOR_K);
    constexpr uint32_t ID_WG = idesc_f16(FMT, CS_TINY_WGRAD_M, 16, MAJOR_MN, MAJOR_MN);
    constexpr uint32_t ID_WB = idesc_f16(FMT, CS_TINY_WGRAD_M, 32, MAJOR_MN, MAJOR_MN);

    // W2 (fp32, scratch order) -> K-major 16x16 bf16/fp16 operand tiles: forward tile k = B[n=co][kk=ci],
    // dgrad tile 5+k = B[n=ci][kk=co].  Group q holds co0..co0+3 of one (k, ci).
    auto repack_w2 = [&](int q) {
        const float4 w = *reinterpret_cast<const float4*>(pm + 4 * q);
        const int rw = q >> 3, col = (q & 7) * 4;
        const int k = rw >> 3, c8 = rw & 7, j = col >> 4, co0 = col & 15, ci = j * 8 + c8;
        const uint32_t lo = pack2<FP16>(w.x, w.y), hi = pack2<FP16>(w.z, w.w);
        uint8_t* tf = wt + k * 512 + (co0 >> 3) * 128 + (ci >> 3) * 256 + (ci & 7) * 2 + (co0 & 7) * 16;
        *reinterpret_cast<uint16_t*>(tf) = (uint16_t)(lo & 0xffff);
        *reinterpret_cast<uint16_t*>(tf + 16) = (uint16_t)(lo >> 16);
        *reinterpret_cast<uint16_t*>(tf + 32) = (uint16_t)(hi & 0xffff);
        *reinterpret_cast<uint16_t*>(tf + 48) = (uint16_t)(hi >> 16);
        uint8_t* td = wt + (5 + k) * 512 + (ci & 7) * 16 + (ci >> 3) * 128 + (co0 >> 3) * 256 + (co0 & 7) * 2;
        *reinterpret_cast<uint2*>(td) = make_uint2(lo, hi);
    };
    for (int q = t; q < 320; q += NT) repack_w2(q);
    fence_proxy_async_smem();
    __syncthreads();

    int err = 0;
    const int ws = tt >> 5;                                  // warp index inside the slot
    // ---- MMA issue, spread over the slot's warps ------------------------------------------------------------------------
    // Measured on B200 (bench/_mma_probe.py): ONE thread issues a tcgen05.mma every ~50 cycles whatever its shape, while the
    // tensor pipe retires these small shapes in 24 (M64 x N16/N32) to 40 (M128 x N16) cycles - a single issuer, not the pipe,
    // paces a 136-MMA sample.  So every warp issues a share, each into accumulators no other warp touches: forward / data
    // gradient: warp i owns position tile i; conv2 weight gradient: warps 4-7 own (rotating set, channel plane); conv1 weight
    // gradient + biases: warps 0-3 own one of the four rotating sets.  Each issuing warp commits its own MMAs.
    auto issue_fwd = [&](uint32_t par) {
        if (ws >= TILES) return;
        if (!mbar_wait(rdy_fwd, par)) err |= ERR_MBAR;
        tc_fence_after_sync();
        if (lane == 0) {
            if (ws < tiles) {
                const uint64_t a0 = smem_desc(s_h1 + 6 * 16, S::PLANE, 128) + (uint64_t)(ws * 128);
                const uint64_t b0 = smem_desc(s_wt, 256, 128);
#pragma unroll
                for (int k = 0; k < 5; ++k) mma_f16(tmem_base + ws * 16, a0 + (uint64_t)k, b0 + (uint64_t)(k * 32), ID_CONV, k > 0);
                mma_commit(bar_fwd);
            } else {
                mbar_arrive(bar_fwd);
            }
        }
        __syncwarp();
    };
    auto issue_bwd = [&](uint32_t par, bool first) {
        if (!mbar_wait(rdy_bwd, par)) err |= ERR_MBAR;
        tc_fence_after_sync();
        if (lane == 0) {
            if (ws < TILES) {
                if (ws < tiles) {
                    const uint64_t a0 = smem_desc(s_dz2 + 10 * 16, S::PLANE, 128) + (uint64_t)(ws * 128);
                    const uint64_t b0 = smem_desc(s_wt + 5 * 512, 256, 128);
#pragma unroll
                    for (int k = 0; k < 5; ++k) mma_f16(tmem_base + ws * 16, a0 - (uint64_t)k, b0 + (uint64_t)(k * 32), ID_CONV, k > 0);
                    mma_commit(bar_dg);
                } else {
                    mbar_arrive(bar_dg);
                }
            }
            if (ws >= 4) {                                   // conv2 weight gradient: this warp's (set, plane) accumulator
                const int set = (ws - 4) >> 1, plane = (ws - 4) & 1;
                const int chunks = tiles_w * 8;
                const uint64_t wa = smem_desc(s_h1 + plane * S::PLANE + 6 * 16, 128, 16);
                const uint64_t wb = smem_desc(s_dz2 + 8 * 16, 128, S::PLANE);
                const uint32_t d = tmem_base + col_w + set * 32 + plane * 16;
#pragma unroll 4
                for (int c = set; c < chunks; c += 2)
                    mma_f16(d, wa + (uint64_t)(c * 16), wb + (uint64_t)(c * 16), ID_WG, (!first || c >= 2) ? 1u : 0u);
                mma_commit(bar_wg);
            }
        }
        __syncwarp();
    };
    auto issue_wg1 = [&](uint32_t par, bool first) {
        if (ws >= 4) return;
        if (!mbar_wait(rdy_wg1, par)) err |= ERR_MBAR;
        tc_fence_after_sync();
        if (lane == 0) {
            const int chunks = tiles_w * 8;
            const uint64_t xa = smem_desc(s_xcol + 8 * 16, 128, 16);
            const uint64_t xb = smem_desc(s_dz1 + 8 * 16, 128, S::PLANE);
#pragma unroll 4
            for (int c = ws; c < chunks; c += 4)
                mma_f16(tmem_base + col_b + ws * 32, xa + (uint64_t)(c * 16), xb + (uint64_t)(c * 16), ID_WB, (!first || c >= 4) ? 1u : 0u);
            mma_commit(bar_wg);
        }
        __syncwarp();
    };
    {
        // =========================== compute warps (8 per slot) ==========================================================
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

        // [A] stage the prefetched window of the slot's next sample: xs (rounded to the operand type, as autocast would
        // feed conv1) and the im2col plane; then start the loads of the window after it.  Needs no parameters, so for
        // the first sample of a step it runs BEFORE the grid barrier is awaited.
        float xv[TPT][7];
        int cur_label = 0;
        auto stage_sample = [&](bool more) {
            cur_label = label;
            if (ws == 0 && (unsigned)cur_label >= (unsigned)C) { err |= ERR_LABEL; cur_label = 0; }
#pragma unroll
            for (int ii = 0; ii < TPT; ++ii) { const int l = (half * TPT + ii) * 128 + row; if (l < L) xs[l + 3] = round_half<FP16>(xin[ii]); }
            if (more) prefetch_next();
            slot_sync(slot);
#pragma unroll
            for (int i = 0; i < TPT; ++i)
#pragma unroll
                for (int k = 0; k < 7; ++k) xv[i][k] = xs[(half * TPT + i) * 128 + row + k];
#pragma unroll
            for (int i = 0; i < TPT; ++i) {
                const int l = (half * TPT + i) * 128 + row;
                if (l < L)
                    *reinterpret_cast<uint4*>(sl + S::P_XCOL + (l + 8) * 16) =
                        make_uint4(pack2<FP16>(xv[i][0], xv[i][1]), pack2<FP16>(xv[i][2], xv[i][3]),
                                   pack2<FP16>(xv[i][4], xv[i][5]), pack2<FP16>(xv[i][6], 1.0f));
            }
        };

        for (int step = 0; step < ra.steps; ++step) {
            float ghead0 = 0.0f, loss_acc = 0.0f;
            if (ra.prof && t == 0 && step == ra.steps / 2 && blockIdx.y == 0) ra.prof[1024 + 4 * blockIdx.x] = (long long)globaltimer_ns();
            RSTAMP(0);
            // the first window of the step is staged while other CTAs are still arriving at the previous step's barrier
            if (has_work) stage_sample(v + V < B || step + 1 < ra.steps);
            RSTAMP(1);
            if (step > 0) apply_update(step - 1, true);
            RSTAMP(2);
            if (SLOTS > 1 && ra.stagger > 0 && slot == 1) {      // see RoundArgs::stagger
                if (lane == 0) { const long long c0 = clock64(); while (clock64() - c0 < ra.stagger) { } }
                __syncwarp();
            }
            const float lscale = scale;                          // loss scale of this step (1 without a scaler)
            bool first = true;
            for (int s = v; s < B; s += V, ++it, first = false) {
                const uint32_t par = it & 1;
                if (!first) {
                    if (!mbar_wait(bar_wg, par ^ 1)) err |= ERR_MBAR;   // previous sample's MMAs released the planes
                    stage_sample(s + V < B || step + 1 < ra.steps);
                }
                // ---- conv1 + ReLU on FFMA -> h1 planes -------------------------------------------------------------------
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
                            *reinterpret_cast<uint4*>(sl + S::P_H1 + ro) = make_uint4(hp[i][0], hp[i][1], hp[i][2], hp[i][3]);
                            *reinterpret_cast<uint4*>(sl + S::P_H1 + S::PLANE + ro) = make_uint4(hp[i][4], hp[i][5], hp[i][6], hp[i][7]);
                        }
                    }
                }
                handoff(rdy_fwd, lane);
                issue_fwd(par);
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

                // ---- head grads, dp, dz2 = mask * dp / L -> dz2 planes ---------------------------------------------------------
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
                    for (int c = 0; c < 16; ++c) gg[c] = gj[c];
#pragma unroll
                    for (int i = 0; i < TPT; ++i) {
                        const int l = (half * TPT + i) * 128 + row;
                        if (l < L) {
                            uint32_t q[8];
#pragma unroll
                            for (int c = 0; c < 16; c += 2)
                                q[c >> 1] = pack2<FP16>((mask[i] >> c) & 1u ? gg[c] : 0.0f, (mask[i] >> (c + 1)) & 1u ? gg[c + 1] : 0.0f);
                            const int ro = (l + 8) * 16;
                            *reinterpret_cast<uint4*>(sl + S::P_DZ2 + ro) = make_uint4(q[0], q[1], q[2], q[3]);
                            *reinterpret_cast<uint4*>(sl + S::P_DZ2 + S::PLANE + ro) = make_uint4(q[4], q[5], q[6], q[7]);
                        }
                    }
                }
                handoff(rdy_bwd, lane);
                issue_bwd(par, first);
                RSTAMP(7);
                if (!mbar_wait(bar_dg, par)) err |= ERR_MBAR;
                tc_fence_after_sync();
                RSTAMP(8);

                // ---- epilogue 2: dz1 = dh1 * (h1 > 0) -> dz1 planes ------------------------------------------------------------
#pragma unroll
                for (int i = 0; i < TPT; ++i) {
                    const int ti = half * TPT + i;
                    if (ti < tiles) {
                        uint32_t vr[16];
                        tmem_ld16(tmem_lane + ti * 16, vr);
                        tmem_ld_wait();
                        const int l = ti * 128 + row;
                        if (l < L) {
                            const int ro = (l + 8) * 16;
                            const uint4 ha = *reinterpret_cast<const uint4*>(sl + S::P_H1 + ro);
                            const uint4 hb = *reinterpret_cast<const uint4*>(sl + S::P_H1 + S::PLANE + ro);
                            const uint32_t hw[8] = {ha.x, ha.y, ha.z, ha.w, hb.x, hb.y, hb.z, hb.w};
                            uint32_t q[8];
#pragma unroll
                            for (int c = 0; c < 16; c += 2) {
                                const float2 h = unpack2<FP16>(hw[c >> 1]);
                                q[c >> 1] = pack2<FP16>(h.x > 0.0f ? __uint_as_float(vr[c]) : 0.0f,
                                                        h.y > 0.0f ? __uint_as_float(vr[c + 1]) : 0.0f);
                            }
                            *reinterpret_cast<uint4*>(sl + S::P_DZ1 + ro) = make_uint4(q[0], q[1], q[2], q[3]);
                            *reinterpret_cast<uint4*>(sl + S::P_DZ1 + S::PLANE + ro) = make_uint4(q[4], q[5], q[6], q[7]);
                        }
                    }
                }
                handoff(rdy_wg1, lane);
                issue_wg1(par, first);
                RSTAMP(9);
            }

            // ---- end of step: weight-gradient accumulators TMEM -> this step's scratch buffer (16-byte reductions) -------
            if (has_work) {
                if (!mbar_wait(bar_wg, (it - 1) & 1)) err |= ERR_MBAR;
                tc_fence_after_sync();
                RSTAMP(10);
                const int r = wgrad_row_of_thread(tt);
                constexpr int kWarpsW = CS_TINY_WGRAD_M == 128 ? 2 : 3;
                const int n0 = half * 16;
                float* g = R.gacc + (size_t)(step % kRoundBufs) * kBufFloats + (v % kScratchReps) * kScratchStride;
                if (wq < kWarpsW) {
                    float vr[16];
                    load_sum_sets16<2>(tmem_lane + col_w + n0, vr);
                    if (r >= 0 && r < 40) {
#pragma unroll
                        for (int n = 0; n < 16; n += 4) red_add_v4(g + r * 32 + n0 + n, vr[n], vr[n + 1], vr[n + 2], vr[n + 3]);
                    }
                }
                if (wq == 0) {
                    float vr[16];
                    load_sum_sets16<4>(tmem_lane + col_b + n0, vr);
                    if (r >= 0 && r < 8) {
#pragma unroll
                        for (int n = 0; n < 16; n += 4) red_add_v4(g + kScrGB + r * 32 + n0 + n, vr[n], vr[n + 1], vr[n + 2], vr[n + 3]);
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
                if (i < P) { const int o = g_ofs_table.v[i]; loc[j] = pm[o]; mo[j] = mm[o]; }
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

// ------------------------------------------------------------------------------------------------------------------
static int g_last_round_grid = 0;

template <int TILES, bool FP16>
static int launch_round_t(const RoundArgs& ra, int ranks, cudaStream_t st) {
    using S = RSmem<TILES>;
    constexpr int NT = round_threads(TILES);
    const size_t smem = S::TOTAL;
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    static int per_sm_cached[64] = {};                      // one record per instantiation and device
    if (per_sm_cached[dev & 63] == 0) {
        CS_CUDA_OK(cudaFuncSetAttribute(tiny_round_kernel<TILES, FP16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 0;
        CS_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, tiny_round_kernel<TILES, FP16>, NT, smem));
        CS_REQUIRE(occ >= 1, "train_round: the kernel does not fit on this device (%zu B shared memory, %d threads)", smem, NT);
        per_sm_cached[dev & 63] = 1;                        // the CTA owns all 512 TMEM columns: one per SM
    }
    const int cap = num_sms() / ranks;
    CS_REQUIRE(cap >= 1, "train_round: %d emulated ranks do not fit on one device", ranks);
    const int want = (ra.B + S::SLOTS - 1) / S::SLOTS;      // one sample slot per sample, SLOTS slots per CTA
    const int grid = want < cap ? want : cap;               // every CTA must be co-resident: the step barrier is grid wide
    g_last_round_grid = grid;
    void* args[] = {(void*)&ra};
    CS_CUDA_OK(cudaLaunchCooperativeKernel((const void*)tiny_round_kernel<TILES, FP16>, dim3(grid, ranks), dim3(NT), args, smem, st));
    return CS_OK;
}

static int launch_round(RoundArgs& ra, int ranks, int amp, cudaStream_t st) {
    CS_REQUIRE(amp == CS_AMP_BF16 || amp == CS_AMP_FP16 || amp == CS_AMP_FP32, "train_round: unknown amp mode %d", amp);
    CS_REQUIRE(amp != CS_AMP_FP32 || ra.L <= 512, "train_round: the fp32 (tf32 tensor-core) round kernel handles windows up to 512 samples, got %d", ra.L);
    CS_REQUIRE(tiny_umma_supported(ra.L, ra.C), "train_round: unsupported L=%d C=%d", ra.L, ra.C);
    CS_REQUIRE(ra.B >= 1 && ra.steps >= 1, "train_round: bad batch %d / steps %d", ra.B, ra.steps);
    CS_REQUIRE(ra.sample_mode >= SAMPLE_IDX && ra.sample_mode <= SAMPLE_PERM, "train_round: bad sample mode %d", ra.sample_mode);
    for (int r = 0; r < ranks; ++r) {
        const RoundRank& R = ra.r[r];
        CS_REQUIRE(R.x && R.y && R.params && R.mom && R.gacc && R.ctrl && R.fstate, "train_round: null pointer (rank %d)", r);
        CS_REQUIRE(ra.sample_mode != SAMPLE_IDX || R.idx != nullptr, "train_round: sample mode 0 needs an index array");
        CS_REQUIRE(ra.sample_mode != SAMPLE_PERM || R.n_rows >= ra.B, "train_round: %lld rows < batch %d", R.n_rows, ra.B);
        CS_REQUIRE(ra.sample_mode != SAMPLE_ROWS || R.n_rows == 0 || R.n_rows >= (long long)ra.steps * ra.B,
                   "train_round: the staged block holds %lld rows, the round needs %lld", R.n_rows, (long long)ra.steps * ra.B);
        CS_REQUIRE(ra.comm == nullptr || R.prev != nullptr, "train_round: fused aggregation needs the previous global model");
    }
    ra.prof = g_tiny_prof;
    { static const int stagger = []() { const char* e = getenv("CS_ROUND_STAGGER"); return e ? atoi(e) : 1800; }(); ra.stagger = stagger; }
    if (amp == CS_AMP_FP32) return launch_round_tf32(ra, ranks, st, &g_last_round_grid);
    const bool fp16 = amp == CS_AMP_FP16;
    if (ra.L <= 512) return fp16 ? launch_round_t<4, true>(ra, ranks, st) : launch_round_t<4, false>(ra, ranks, st);
    return fp16 ? launch_round_t<8, true>(ra, ranks, st) : launch_round_t<8, false>(ra, ranks, st);
}

size_t agg_arena_need(long long, int, int);

static int fill_comm(RoundArgs& ra, cs_comm* c, const CommDev* ctxs, int world, int n, int wire_bf16, float server_lr, int reset_mom,
                     size_t arena_off, int channel) {
    CS_REQUIRE(n <= 4096, "train_round: fused aggregation handles up to 4096 parameters (one flag block)");
    CS_REQUIRE(channel >= 0 && channel < kChannels, "train_round: channel %d out of range", channel);
    CS_REQUIRE(arena_off % 16 == 0 && arena_off + agg_arena_need(n, wire_bf16, world) <= c->bytes,
               "train_round: arena too small (%zu B needed at offset %zu, have %zu)", agg_arena_need(n, wire_bf16, world), arena_off, c->bytes);
    ra.comm = ctxs; ra.wire_bf16 = wire_bf16; ra.server_lr = server_lr; ra.reset_mom = reset_mom; ra.arena_off = arena_off; ra.channel = channel;
    return CS_OK;
}

}  // namespace cs

using namespace cs;

extern "C" {

/* debug: CTAs per (emulated) client of the most recent round launch */
int cs_debug_last_round_grid(void) { return g_last_round_grid; }

size_t cs_tiny_ecg_round_scratch_floats(int C) { (void)C; return (size_t)kRoundBufs * kBufFloats; }

unsigned long long cs_perm_index(unsigned long long i, unsigned long long n, unsigned long long seed, unsigned long long epoch) {
    return perm_index(i, n, perm_epoch_key(seed, epoch));
}

int cs_tiny_ecg_train_round(const void* x, const int64_t* y, const int64_t* idx, int x_bf16, int64_t n_rows, int B, int L, int C,
                            int steps, int sample_mode, uint64_t seed, float* params, float* mom, float* gacc, int32_t* ctrl,
                            float* fstate, float lr, float momentum, float wd, int amp, int use_scaler, float growth, float backoff,
                            int growth_interval, cs_comm_t* comm, float* prev, float weight, int wire_bf16, float server_lr,
                            int reset_momentum, size_t arena_off, int channel, cs_stream_t stream) {
    RoundArgs ra;
    memset(&ra, 0, sizeof(ra));
    RoundRank& R = ra.r[0];
    R.x = x; R.y = y; R.idx = idx; R.params = params; R.mom = mom; R.gacc = gacc; R.ctrl = ctrl; R.fstate = fstate; R.prev = prev;
    R.weight = weight; R.seed = seed; R.n_rows = n_rows;
    ra.x_bf16 = x_bf16; ra.B = B; ra.L = L; ra.C = C; ra.steps = steps; ra.sample_mode = sample_mode;
    ra.lr = lr; ra.momentum = momentum; ra.wd = wd; ra.use_scaler = use_scaler; ra.growth = growth; ra.backoff = backoff; ra.interval = growth_interval;
    if (comm != nullptr) {
        cs_comm* c = (cs_comm*)comm;
        CS_REQUIRE(!c->emulated, "train_round: use cs_tiny_ecg_train_round_local for emulated groups");
        int rc = fill_comm(ra, c, c->dev, c->world, TinyLayout::total(C), wire_bf16, server_lr, reset_momentum, arena_off, channel);
        if (rc) return rc;
    }
    return launch_round(ra, 1, amp, (cudaStream_t)stream);
}

int cs_tiny_ecg_train_round_local(cs_comm_t** group, int world, const void* const* x, const int64_t* const* y, const int64_t* const* idx,
                                  int x_bf16, const int64_t* n_rows, int B, int L, int C, int steps, int sample_mode, const uint64_t* seeds,
                                  float* const* params, float* const* mom, float* const* gacc, int32_t* const* ctrl, float* const* fstate,
                                  float lr, float momentum, float wd, int amp, int use_scaler, float growth, float backoff,
                                  int growth_interval, float* const* prev, const float* weights, int wire_bf16, float server_lr,
                                  int reset_momentum, size_t arena_off, int channel, cs_stream_t stream) {
    cs_comm** g = (cs_comm**)group;
    CS_REQUIRE(g && g[0] && g[0]->emulated && g[0]->world == world && world <= kMaxRanks, "train_round_local: not an emulated group");
    RoundArgs ra;
    memset(&ra, 0, sizeof(ra));
    for (int r = 0; r < world; ++r) {
        RoundRank& R = ra.r[r];
        R.x = x[r]; R.y = y[r]; R.idx = idx ? idx[r] : nullptr; R.params = params[r]; R.mom = mom[r]; R.gacc = gacc[r]; R.ctrl = ctrl[r];
        R.fstate = fstate[r]; R.prev = prev[r]; R.weight = weights[r]; R.seed = seeds ? seeds[r] : 0; R.n_rows = n_rows ? n_rows[r] : 0;
    }
    ra.x_bf16 = x_bf16; ra.B = B; ra.L = L; ra.C = C; ra.steps = steps; ra.sample_mode = sample_mode;
    ra.lr = lr; ra.momentum = momentum; ra.wd = wd; ra.use_scaler = use_scaler; ra.growth = growth; ra.backoff = backoff; ra.interval = growth_interval;
    int rc = fill_comm(ra, g[0], g[0]->group_dev, world, TinyLayout::total(C), wire_bf16, server_lr, reset_momentum, arena_off, channel);
    if (rc) return rc;
    return launch_round(ra, world, amp, (cudaStream_t)stream);
}

}  // extern "C"
