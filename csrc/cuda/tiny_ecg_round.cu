// Persistent FL-round kernel for TinyECG (tcgen05 / TMEM path): ONE cooperative launch runs all local
// steps of a federated round and ends in the FedAvg aggregation.
//
// Replaces the reference's host-driven round (TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:186-231:
// Bcast, 50 x {~40 library launches + cuda.synchronize + loss.item()}, Allreduce) and this library's own
// first-generation round (50 launches of tiny_umma_kernel + 1 aggregation launch).
//
// What stays resident for the whole round: the CTA's TMEM allocation, its mbarriers, the operand planes
// and - new - a shared-memory master copy of all parameters and their momentum.  Per local step:
//   * every CTA runs the fused forward/backward of its sample(s) exactly as the one-step kernel does
//     (tiny_ecg_umma.cu), with MMA issue moved to a dedicated ninth warp so the eight compute warps never
//     stall behind descriptor arithmetic;
//   * weight-gradient accumulators leave TMEM as 16-byte vector reductions into one of THREE rotating
//     global scratch buffers (TMEM-shaped, 4-way replicated); the step's loss rides in a spare slot;
//   * ONE grid-wide barrier (arrive: bar.sync + fence + atomic add by one thread; wait: ld.acquire.gpu
//     spin) - split-phase: between arrive and wait the CTA already stages the next step's window;
//   * after the barrier EVERY CTA sums the replicas itself (coalesced 16-byte L2 loads) and applies
//     SGD-momentum / the GradScaler rules to its own shared-memory copy.  Parameters, momentum and
//     gradients share the scratch ("TMEM-shaped") ordering, so the update is purely elementwise: no
//     last-CTA serial tail, no gather, no global parameter round trip.  All CTAs read the same totals in
//     the same order, so their copies stay bit-identical without any further communication.
//   * a buffer used at step s is cleared (distributed over CTAs) during step s+2 and reused at s+3:
//     no second barrier is ever needed.
// At the end CTA 0 un-permutes its copy and either writes it back or - fused aggregation - pushes the
// weighted delta to every peer's arena over NVLink, raises/awaits the system-scope flags and applies the
// server update (same protocol and arena layout as fedavg_oneshot_kernel, so both are interchangeable).
// Batch indices come from an index array, from contiguous rows (staged blocks) or from an in-kernel
// counter-based permutation (a bijection on [0, N) keyed by seed and epoch; ref semantics:
// Module_3/shard_dataset.py:118-136 per-epoch randperm, ragged tail dropped).
#include "tiny_umma_common.cuh"
#include "comm_dev.cuh"

#include <string.h>

namespace cs {

constexpr int kRCompute = 256;                       // 8 compute warps (two per TMEM lane quadrant)
constexpr int kRThreads = kRCompute + 32;            // + the MMA-issuer warp
constexpr int kRoundBufs = 3;
constexpr int kScrLoss = kScrGH + 17 * kTinyMaxC;    // spare scratch slot carrying the step's mean loss
constexpr int kBufFloats = kScratchReps * kScratchStride;
static_assert(kScrLoss < kScratchStride && (kScrLoss & 3) == 0, "loss slot must start a 16-byte group inside the scratch");
static_assert(CS_TINY_NOALLOC_SCRATCH_FLOATS >= kScratchReps * kScratchStride,
              "the built-in scratch of cs_tiny_ecg_train_step_noalloc must hold the tcgen05 kernel's replicas");

enum { SAMPLE_IDX = 0, SAMPLE_ROWS = 1, SAMPLE_PERM = 2 };
enum { ERR_LABEL = 1, ERR_MBAR = 2, ERR_GRID = 4, ERR_PEER = 8 };

struct RoundRank {                                   // per (emulated) rank; index 0 in real multi-process runs
    const void* x; const int64_t* y; const int64_t* idx;
    float* params; float* mom; float* gacc; int32_t* ctrl; float* fstate; float* prev;
    float weight;
    unsigned long long seed;
    long long n_rows;
};

struct RoundArgs {
    RoundRank r[kMaxRanks];
    int x_bf16, B, L, C, steps, sample_mode;
    float lr, momentum, wd; int use_scaler; float growth, backoff; int interval;
    const CommDev* comm; int wire_bf16; float server_lr; int reset_mom; int channel; size_t arena_off;
    long long* prof;
};

// ---- counter-based permutation -------------------------------------------------------------------------------
// A bijection on [0, n): four rounds of {multiply by an odd constant, add a key word, xor-shift} on
// ceil(log2 n) bits, cycle-walking values that fall outside [0, n).  Mirrored in Python
// (ops/tiny_ecg.py::perm_index) so tests and oracles can reproduce the batches.
__host__ __device__ inline unsigned long long splitmix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ inline unsigned long long perm_index(unsigned long long i, unsigned long long n, unsigned long long key) {
    if (n <= 1) return 0;
    int bits = 1;
    while (bits < 63 && (1ull << bits) < n) ++bits;
    const unsigned long long mask = (1ull << bits) - 1;
    const int sh = (bits + 1) >> 1;
    const unsigned long long k0 = splitmix64(key), k1 = splitmix64(k0), k2 = splitmix64(k1), k3 = splitmix64(k2);
    unsigned long long x = i;
    do {
        x = (x * 0x9E3779B97F4A7C15ull + k0) & mask; x ^= x >> sh;
        x = (x * 0xD6E8FEB86659FD93ull + k1) & mask; x ^= x >> sh;
        x = (x * 0xCA5A826395121157ull + k2) & mask; x ^= x >> sh;
        x = (x * 0xA24BAED4963EE407ull + k3) & mask; x ^= x >> sh;
    } while (x >= n);
    return x;
}

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

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void csync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }   // the 8 compute warps only
__device__ __forceinline__ int csync_or(int pred) {
    uint32_t r;
    asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 q, %1, 0;\n\tbar.red.or.pred p, 1, 256, q;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(r) : "r"(pred) : "memory");
    return (int)r;
}

template <int TILES> struct RSmem {
    static constexpr int ROWS = TILES * 128 + 16;          // 8 halo rows on top, >= 8 below
    static constexpr int PLANE = ROWS * 16;                // bytes
    static constexpr int OFF_BAR = 0;                      // 6 mbarriers, tmem slot
    static constexpr int OFF_F32 = 64;                     // pooled[16] dlog[8] gj[16] red[128]
    static constexpr int F32_COUNT = 16 + 8 + 16 + 128;
    static constexpr int OFF_PM = (OFF_F32 + F32_COUNT * 4 + 127) & ~127;   // parameters, scratch order
    static constexpr int OFF_MM = OFF_PM + kScratchStride * 4;              // momentum, scratch order
    static constexpr int OFF_WT = OFF_MM + kScratchStride * 4;              // 10 weight tiles x 512 B
    static constexpr int OFF_XS = OFF_WT + 10 * 512;
    static constexpr int XS_LEN = TILES * 128 + 8;
    static constexpr int OFF_PLANES = (OFF_XS + XS_LEN * 4 + 127) & ~127;
    static constexpr int P_DZ1 = OFF_PLANES, P_DZ2 = P_DZ1 + 2 * PLANE, P_H1 = P_DZ2 + 2 * PLANE,
                         P_XCOL = P_H1 + 2 * PLANE;
    static constexpr int TOTAL = P_XCOL + PLANE + 512;
    static constexpr int TMEM_COLS = (16 * TILES + 64 + 128) <= 256 ? 256 : 512;
};

// one lane per warp publishes "this warp's shared-memory writes are done" to the issuer warp
__device__ __forceinline__ void handoff(uint64_t* bar, int lane) {
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncwarp();
    if (lane == 0) mbar_arrive(bar);
}

// ---- fused FedAvg tail (CTA 0 of a rank, the 256 compute threads) ----------------------------------------------------
// Same protocol, arena layout ([parity 2][src rank P][n] wire elements at arena_off) and epoch word as block 0 of
// fedavg_oneshot_kernel, so a round that ends here and a stand-alone cs_fedavg_aggregate can alternate on a channel.
constexpr int kRoundNP = (kOfsTableLen + kRCompute - 1) / kRCompute;     // parameters per thread in the tail

template <bool BF16>
__device__ __forceinline__ void round_aggregate(const CommDev& c, const RoundArgs& ra, const RoundRank& R, const float (&loc)[kRoundNP],
                                                const float (&mo)[kRoundNP], int P, int t, int& err) {
    using W = Wire<BF16>;
    using T = typename W::T;
    const int Pw = c.world;
    uint32_t* ep = c.epochs + (size_t)ra.channel * kMaxBlocks;
    const uint32_t epoch = *ep + 1;
    const int parity = epoch & 1;
    const size_t slot_elems = (size_t)((P + 7) & ~7);
    float prv[kRoundNP];
#pragma unroll
    for (int j = 0; j < kRoundNP; ++j) {
        const int i = t + j * kRCompute;
        prv[j] = 0.0f;
        if (i < P) {
            prv[j] = R.prev[i];
            const T e = W::enc(R.weight * (loc[j] - prv[j]));
            for (int r = 0; r < Pw; ++r) {
                T* dst = reinterpret_cast<T*>(c.arena[(c.rank + r) % Pw] + ra.arena_off) + ((size_t)parity * Pw + c.rank) * slot_elems;
                dst[i] = e;
            }
        }
    }
    csync();
    int late = 0;
    if (t < Pw) {
        const size_t base = (size_t)ra.channel * kMaxBlocks * kMaxRanks;
        __threadfence_system();
        st_release_sys(c.signals[t] + base + c.rank, epoch);
        const uint32_t* mine = c.signals[c.rank] + base + t;
        const unsigned long long t0 = globaltimer_ns();
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
            if (globaltimer_ns() - t0 > comm_timeout_ns(c)) { atomicExch(c.status, 1); late = 1; break; }
        }
    }
    late = csync_or(late);
    if (late) err |= ERR_PEER;
    const T* base = reinterpret_cast<const T*>(c.arena[c.rank] + ra.arena_off) + (size_t)parity * Pw * slot_elems;
#pragma unroll
    for (int j = 0; j < kRoundNP; ++j) {
        const int i = t + j * kRCompute;
        if (i < P) {
            float out = loc[j];                            // a peer never arrived: keep the local model, do not reduce garbage
            if (!late) {
                float acc = 0.0f;
                for (int r = 0; r < Pw; ++r) acc += W::dec(__ldcg(base + (size_t)r * slot_elems + i));
                out = prv[j] + ra.server_lr * acc;
                R.prev[i] = out;
            }
            R.params[i] = out;
            R.mom[i] = ra.reset_mom ? 0.0f : mo[j];
        }
    }
    if (t == 0) *ep = epoch;
}

#define RSTAMP(k) do { if (ra.prof && blockIdx.x == 0 && blockIdx.y == 0 && t == 0 && step == ra.steps / 2) ra.prof[k] = clock64(); } while (0)

template <int TILES, bool FP16>
__global__ void __launch_bounds__(kRThreads, TILES <= 4 ? 2 : 1) tiny_round_kernel(const __grid_constant__ RoundArgs ra) {
    using S = RSmem<TILES>;
    constexpr int TPT = TILES / 2;                         // position tiles per compute thread
    extern __shared__ __align__(128) uint8_t smem[];
    const int er = blockIdx.y;
    const RoundRank& R = ra.r[er];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int row = t & 127, half = (t >> 7) & 1;          // TMEM lane of this thread / which half of the tiles it owns
    const int L = ra.L, C = ra.C, B = ra.B, G = gridDim.x;
    const int tiles = (L + 127) >> 7;
    const int P = TinyLayout::total(C);

    uint64_t* bar_fwd = reinterpret_cast<uint64_t*>(smem + S::OFF_BAR);   // MMA -> compute (tcgen05.commit)
    uint64_t* bar_dg = bar_fwd + 1;
    uint64_t* bar_wg = bar_fwd + 2;
    uint64_t* rdy_fwd = bar_fwd + 3;                                      // compute -> issuer (8 warp arrivals)
    uint64_t* rdy_bwd = bar_fwd + 4;
    uint64_t* rdy_wg1 = bar_fwd + 5;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S::OFF_BAR + 48);
    float* f32 = reinterpret_cast<float*>(smem + S::OFF_F32);
    float* pooled = f32;  float* dlog = pooled + 16;  float* gj = dlog + 8;  float* red = gj + 16;
    float* pm = reinterpret_cast<float*>(smem + S::OFF_PM);
    float* mm = reinterpret_cast<float*>(smem + S::OFF_MM);
    const float* w1s = pm + kScrGB;                        // W1[c][k] at k*32 + c
    const float* b1s = pm + kScrGB + 7 * 32;
    const float* b2s = b1s + 16;
    const float* whs = pm + kScrGH;
    const float* bhs = whs + 16 * C;
    uint8_t* wt = smem + S::OFF_WT;
    float* xs = reinterpret_cast<float*>(smem + S::OFF_XS);
    uint32_t* counter = reinterpret_cast<uint32_t*>(R.ctrl + CTRL_TICKET);

    // ---- sampling ----------------------------------------------------------------------------------------------
    long long pos0 = 0, usable = 1;
    if (ra.sample_mode == SAMPLE_PERM) {
        pos0 = *reinterpret_cast<const long long*>(R.ctrl + CTRL_POS);
        usable = (R.n_rows / B) * B;
    }
    auto sample_row = [&](int step, int s) -> int64_t {
        const long long q = (long long)step * B + s;
        if (ra.sample_mode == SAMPLE_IDX) return R.idx[q];
        if (ra.sample_mode == SAMPLE_ROWS) return q;
        const long long g = pos0 + q;
        const long long epoch = g / usable, off = g - epoch * usable;
        return (int64_t)perm_index((unsigned long long)off, (unsigned long long)R.n_rows, R.seed + 0x632BE59BD9B4E019ull * (unsigned long long)epoch);
    };
    float xin[TPT];
    int label = 0;
    auto prefetch_sample = [&](int step, int s) {
        const int64_t drow = sample_row(step, s);
#pragma unroll
        for (int ii = 0; ii < TPT; ++ii) {
            const int l = (half * TPT + ii) * 128 + row;
            xin[ii] = 0.0f;
            if (l < L) xin[ii] = ra.x_bf16 ? __bfloat162float(((const __nv_bfloat16*)R.x)[drow * L + l])
                                           : __ldg((const float*)R.x + drow * L + l);
        }
        if (warp == 0) label = (int)R.y[drow];
    };
    if (warp < 8 && (int)blockIdx.x < B && ra.steps > 0) prefetch_sample(0, blockIdx.x);

    // ---- one-time set-up ------------------------------------------------------------------------------------------
    {   // zero every operand plane + xs (halo / tail rows stay zero forever) and the parameter copies
        uint4* z = reinterpret_cast<uint4*>(smem + S::OFF_PM);
        const int n16 = (S::TOTAL - S::OFF_PM) / 16;
        for (int i = t; i < n16; i += kRThreads) z[i] = make_uint4(0, 0, 0, 0);
    }
    if (t == 0) {
        mbar_init(bar_fwd, 1); mbar_init(bar_dg, 1); mbar_init(bar_wg, 1);
        mbar_init(rdy_fwd, 8); mbar_init(rdy_bwd, 8); mbar_init(rdy_wg1, 8);
        fence_mbar_init();
    }
    if (warp == 8) tmem_alloc(tmem_slot, S::TMEM_COLS);
    float scale = ra.use_scaler ? R.fstate[FS_SCALE] : 1.0f;
    int growth = R.ctrl[CTRL_GROWTH];
    __syncthreads();
    for (int i = t; i < P; i += kRThreads) {               // state_dict order -> scratch order
        const int o = g_ofs_table.v[i];
        pm[o] = R.params[i];
        mm[o] = R.mom[i];
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    const uint32_t col_w = 16 * TILES, col_b = 16 * TILES + 64;
    const uint32_t s_base = smem_u32(smem);
    const uint32_t s_dz1 = s_base + S::P_DZ1, s_dz2 = s_base + S::P_DZ2, s_h1 = s_base + S::P_H1,
                   s_xcol = s_base + S::P_XCOL, s_wt = s_base + S::OFF_WT;
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
    if (t < 320) repack_w2(t);
    if (t + kRThreads < 320) repack_w2(t + kRThreads);
    fence_proxy_async_smem();
    __syncthreads();

    int err = 0;
    if (warp == 8) {
        // =========================== MMA issuer warp ===============================================================
        uint32_t it = 0;
        for (int step = 0; step < ra.steps; ++step) {
            bool first = true;
            for (int s = blockIdx.x; s < B; s += G, ++it, first = false) {
                const uint32_t par = it & 1;
                if (!mbar_wait(rdy_fwd, par)) err |= ERR_MBAR;
                tc_fence_after_sync();
                if (lane == 0) {
                    const uint64_t a0 = smem_desc(s_h1 + 6 * 16, S::PLANE, 128);
                    const uint64_t b0 = smem_desc(s_wt, 256, 128);
#pragma unroll
                    for (int k = 0; k < 5; ++k)           // tap-outer / tile-inner: neighbours never share an accumulator
                        for (int i = 0; i < tiles; ++i)
                            mma_f16(tmem_base + i * 16, a0 + (uint64_t)(i * 128 + k), b0 + (uint64_t)(k * 32), ID_CONV, k > 0);
                    mma_commit(bar_fwd);
                }
                __syncwarp();
                if (!mbar_wait(rdy_bwd, par)) err |= ERR_MBAR;
                tc_fence_after_sync();
                if (lane == 0) {
                    const uint64_t a0 = smem_desc(s_dz2 + 10 * 16, S::PLANE, 128);
                    const uint64_t b0 = smem_desc(s_wt + 5 * 512, 256, 128);
#pragma unroll
                    for (int k = 0; k < 5; ++k)
                        for (int i = 0; i < tiles; ++i)
                            mma_f16(tmem_base + i * 16, a0 + (uint64_t)(i * 128 - k), b0 + (uint64_t)(k * 32), ID_CONV, k > 0);
                    mma_commit(bar_dg);
                    const int chunks = tiles * 8;
                    const uint64_t wa0 = smem_desc(s_h1 + 6 * 16, 128, 16);
                    const uint64_t wa1 = smem_desc(s_h1 + S::PLANE + 6 * 16, 128, 16);
                    const uint64_t wb = smem_desc(s_dz2 + 8 * 16, 128, S::PLANE);
#pragma unroll 4
                    for (int c = 0; c < chunks; ++c) {
                        const uint32_t acc = (!first || c >= 2) ? 1u : 0u;
                        const uint32_t d = tmem_base + col_w + (c & 1) * 32;
                        mma_f16(d, wa0 + (uint64_t)(c * 16), wb + (uint64_t)(c * 16), ID_WG, acc);
                        mma_f16(d + 16, wa1 + (uint64_t)(c * 16), wb + (uint64_t)(c * 16), ID_WG, acc);
                    }
                }
                __syncwarp();
                if (!mbar_wait(rdy_wg1, par)) err |= ERR_MBAR;
                tc_fence_after_sync();
                if (lane == 0) {
                    const int chunks = tiles * 8;
                    const uint64_t xa = smem_desc(s_xcol + 8 * 16, 128, 16);
                    const uint64_t xb = smem_desc(s_dz1 + 8 * 16, 128, S::PLANE);
#pragma unroll 4
                    for (int c = 0; c < chunks; ++c)
                        mma_f16(tmem_base + col_b + (c & 3) * 32, xa + (uint64_t)(c * 16), xb + (uint64_t)(c * 16), ID_WB,
                                (!first || c >= 4) ? 1u : 0u);
                    mma_commit(bar_wg);
                }
                __syncwarp();
            }
        }
    } else {
#include "tiny_ecg_round_compute.cuh"
    }
    if (err && lane == 0) atomicOr(R.ctrl + CTRL_ERROR, err);
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 8) tmem_dealloc(tmem_base, S::TMEM_COLS);
}

// ------------------------------------------------------------------------------------------------------------------
template <int TILES, bool FP16>
static int launch_round_t(const RoundArgs& ra, int ranks, cudaStream_t st) {
    using S = RSmem<TILES>;
    const size_t smem = S::TOTAL;
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    static int per_sm_cached[64] = {};                      // one record per instantiation and device
    if (per_sm_cached[dev & 63] == 0) {
        CS_CUDA_OK(cudaFuncSetAttribute(tiny_round_kernel<TILES, FP16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 0;
        CS_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, tiny_round_kernel<TILES, FP16>, kRThreads, smem));
        const int want = TILES <= 4 ? 2 : 1;                // TMEM: 256 (512) columns per CTA
        per_sm_cached[dev & 63] = occ < want ? occ : want;
        CS_REQUIRE(per_sm_cached[dev & 63] >= 1, "train_round: the kernel does not fit on this device (%zu B shared memory)", smem);
    }
    const int cap = num_sms() * per_sm_cached[dev & 63] / ranks;
    CS_REQUIRE(cap >= 1, "train_round: %d emulated ranks do not fit on one device", ranks);
    const int grid = ra.B < cap ? ra.B : cap;               // every CTA must be co-resident: the step barrier is grid wide
    void* args[] = {(void*)&ra};
    CS_CUDA_OK(cudaLaunchCooperativeKernel((const void*)tiny_round_kernel<TILES, FP16>, dim3(grid, ranks), dim3(kRThreads), args, smem, st));
    return CS_OK;
}

static int launch_round(RoundArgs& ra, int ranks, int amp, cudaStream_t st) {
    CS_REQUIRE(amp == CS_AMP_BF16 || amp == CS_AMP_FP16, "train_round: the persistent round kernel is the tcgen05 path (amp bf16 / fp16)");
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

size_t cs_tiny_ecg_round_scratch_floats(int C) { (void)C; return (size_t)kRoundBufs * kBufFloats; }

unsigned long long cs_perm_index(unsigned long long i, unsigned long long n, unsigned long long seed, unsigned long long epoch) {
    return perm_index(i, n, seed + 0x632BE59BD9B4E019ull * epoch);
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
