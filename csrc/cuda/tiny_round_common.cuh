// Pieces shared by the persistent FL-round kernels (bf16 / fp16: tiny_ecg_round.cu, tf32: tiny_ecg_round_tf32.cu):
// the argument block, the counter-based batch sampler, block / grid synchronisation helpers and the fused FedAvg tail.
#pragma once
#include "tiny_umma_common.cuh"
#include "comm_dev.cuh"

namespace cs {

constexpr int kSlotThreads = 256;                    // compute threads of one sample slot: 8 warps, two per TMEM lane quadrant
__host__ __device__ constexpr int round_slots(int tiles) { return tiles <= 4 ? 2 : 1; }           // L <= 512: two slots per CTA
__host__ __device__ constexpr int round_threads(int tiles) { return round_slots(tiles) * kSlotThreads; }
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
    // Cycles by which slot 1 starts its per-step work after slot 0.  Two slots in lockstep want the tensor pipe (and the
    // shared-memory port its operand fetch saturates) at the same moments and the CUDA cores at the same moments; half a
    // phase of offset lets one slot's MMAs run under the other's epilogues.  Measured at B=256, L=500 on B200: 13.4 us per
    // step in lockstep, 12.0 us with 1600-2400 cycles, 13.1 us with 4500.
    int stagger;
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
struct PermKey { unsigned long long k0, k1, k2, k3, mask; int sh; };
__host__ __device__ inline PermKey perm_key(unsigned long long n, unsigned long long key) {
    PermKey p;
    int bits = 1;
    while (bits < 63 && (1ull << bits) < n) ++bits;
    p.mask = (1ull << bits) - 1;
    p.sh = (bits + 1) >> 1;
    p.k0 = splitmix64(key); p.k1 = splitmix64(p.k0); p.k2 = splitmix64(p.k1); p.k3 = splitmix64(p.k2);
    return p;
}
__host__ __device__ inline unsigned long long perm_apply(unsigned long long i, unsigned long long n, const PermKey& p) {
    if (n <= 1) return 0;
    unsigned long long x = i;
    do {
        x = (x * 0x9E3779B97F4A7C15ull + p.k0) & p.mask; x ^= x >> p.sh;
        x = (x * 0xD6E8FEB86659FD93ull + p.k1) & p.mask; x ^= x >> p.sh;
        x = (x * 0xCA5A826395121157ull + p.k2) & p.mask; x ^= x >> p.sh;
        x = (x * 0xA24BAED4963EE407ull + p.k3) & p.mask; x ^= x >> p.sh;
    } while (x >= n);
    return x;
}
__host__ __device__ inline unsigned long long perm_epoch_key(unsigned long long seed, unsigned long long epoch) {
    return seed + 0x632BE59BD9B4E019ull * epoch;
}
__host__ __device__ inline unsigned long long perm_index(unsigned long long i, unsigned long long n, unsigned long long key) {
    return perm_apply(i, n, perm_key(n, key));
}

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// named barriers: 0 = whole CTA, 1 + slot = the 256 compute threads of a slot, 3 = all compute threads of the CTA
__device__ __forceinline__ void slot_sync(int slot) { asm volatile("bar.sync %0, 256;" ::"r"(1 + slot) : "memory"); }
template <int NC> __device__ __forceinline__ void csync() { asm volatile("bar.sync 3, %0;" ::"n"(NC) : "memory"); }
template <int NC> __device__ __forceinline__ int csync_or(int pred) {
    uint32_t r;
    asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 q, %1, 0;\n\tbar.red.or.pred p, 3, %2, q;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(r) : "r"(pred), "n"(NC) : "memory");
    return (int)r;
}

// one lane per warp publishes "this warp's shared-memory writes are done" to the warps that issue the slot's MMAs
__device__ __forceinline__ void handoff(uint64_t* bar, int lane) {
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncwarp();
    if (lane == 0) mbar_arrive(bar);
}

// ---- fused FedAvg tail (CTA 0 of a rank, its NC compute threads) --------------------------------------------------------
// Same protocol, arena layout ([parity 2][src rank P][n] wire elements at arena_off) and epoch word as block 0 of
// fedavg_oneshot_kernel, so a round that ends here and a stand-alone cs_fedavg_aggregate can alternate on a channel.
template <int NC> struct RoundTail { static constexpr int NP = (kOfsTableLen + NC - 1) / NC; };   // parameters per thread

template <bool BF16, int NC>
__device__ __forceinline__ void round_aggregate(const CommDev& c, const RoundArgs& ra, const RoundRank& R,
                                                const float (&loc)[RoundTail<NC>::NP], const float (&mo)[RoundTail<NC>::NP], int P,
                                                int t, int& err) {
    using W = Wire<BF16>;
    using T = typename W::T;
    constexpr int NP = RoundTail<NC>::NP;
    const int Pw = c.world;
    uint32_t* ep = c.epochs + (size_t)ra.channel * kMaxBlocks;
    const uint32_t epoch = *ep + 1;
    const int parity = epoch & 1;
    const size_t slot_elems = (size_t)((P + 7) & ~7);
    float prv[NP];
#pragma unroll
    for (int j = 0; j < NP; ++j) {
        const int i = t + j * NC;
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
    csync<NC>();
    int late = 0;
    if (t < Pw) {
        const size_t base = (size_t)ra.channel * kMaxBlocks * kMaxRanks;
        __threadfence_system();
        st_release_sys(c.signals[t] + base + c.rank, epoch);
        const uint32_t* mine = c.signals[c.rank] + base + t;
        const unsigned long long t0 = globaltimer_ns();
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
            if (globaltimer_ns() - t0 > comm_timeout_ns(c)) { *(volatile int*)c.status = 1; late = 1; break; }
        }
    }
    late = csync_or<NC>(late);
    if (late) err |= ERR_PEER;
    const T* base = reinterpret_cast<const T*>(c.arena[c.rank] + ra.arena_off) + (size_t)parity * Pw * slot_elems;
#pragma unroll
    for (int j = 0; j < NP; ++j) {
        const int i = t + j * NC;
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

// fp32 configuration on the tensor core (kind::tf32), tiny_ecg_round_tf32.cu; L <= 512
int launch_round_tf32(const RoundArgs& ra, int ranks, cudaStream_t st, int* grid_out);

}  // namespace cs
