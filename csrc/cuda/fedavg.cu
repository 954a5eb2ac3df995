// In-kernel NVLink collectives over peer-mapped buffers (K7 fused FedAvg, K8 bucket all-reduce,
// K9 device barrier).  No MPI / NCCL: ranks exchange CUDA IPC handles once on the host, after that
// every collective is ONE kernel per rank that issues st.global / ld.global on peer pointers
// (NVLink 5 through NVSwitch) and synchronises with st.release.sys / ld.acquire.sys flags.
//
// Replaces broadcast_model + fedavg_allreduce (TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:75-98:
// per round 24 blocking memcpys + 12 host MPI collectives) and fuses into the same kernel the
// per-client weighting n_i/sum(n), the fp32<->bf16 wire cast (delta encoded against the previous
// global model so bf16 only quantises the round's *change*), the server update
// theta <- theta_prev + server_lr * avg_delta and the optional momentum reset.
//
// One-shot (small n, latency bound): every rank PUSHES its weighted delta into a slot of every
// peer's arena, raises a flag, waits for all flags and reduces the slots locally
// (one NVLink one-way trip + one flag, no round trip).
// Two-shot (large n, bandwidth bound): reduce-scatter by P2P loads of slice `rank` from every
// peer, all-gather by P2P stores of the reduced slice into every peer — (P-1)/P * n elements in and
// out per GPU, the bandwidth-optimal volume without in-switch reduction.
#include "common/cs_common.cuh"

#include <string.h>
#include <vector>

namespace cs {

constexpr int kMaxRanks = CS_MAX_RANKS;
constexpr int kChannels = 8;          // independent barrier channels (streams / buckets in flight)
constexpr int kMaxBlocks = 128;       // blocks per collective kernel (all must be co-resident)
constexpr int kCommThreads = 512;
constexpr size_t kSignalWords = (size_t)kChannels * kMaxBlocks * kMaxRanks;
constexpr int64_t kOneShotMaxElems = 64 * 1024;   // 256 KiB of fp32

struct CommDev {                      // lives in device memory, one per rank
    int rank, world;
    char* arena[kMaxRanks];           // peer-mapped arenas (arena[rank] is local)
    uint32_t* signals[kMaxRanks];     // peer-mapped signal pads
    uint32_t* epochs;                 // local: [kChannels][kMaxBlocks] barrier epochs
    int* status;                      // local: != 0 after a barrier timeout
    size_t arena_bytes;
};

}  // namespace cs

struct cs_comm {
    int rank = 0, world = 1;
    size_t bytes = 0;
    int device = 0;
    bool emulated = false;            // all ranks on one device (tests / single GPU)
    void* arena = nullptr;
    uint32_t* signals = nullptr;
    uint32_t* epochs = nullptr;
    int* status = nullptr;
    void* peer_arena[cs::kMaxRanks] = {};
    uint32_t* peer_signals[cs::kMaxRanks] = {};
    bool opened[cs::kMaxRanks] = {};
    cs::CommDev* dev = nullptr;       // device copy of the context
    cs_comm** group = nullptr;        // emulation: all members (owned by member 0)
    cs::CommDev* group_dev = nullptr; // emulation: contiguous array of `world` contexts
};

namespace cs {

// ------------------------------------------------------------------------------------------------
// device-side primitives
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// Cross-GPU barrier between block `b` of every rank on channel `ch`.  Epochs are kept in device
// memory (not kernel arguments) so that a captured CUDA graph can be replayed.
__device__ __forceinline__ void xgpu_barrier(const CommDev& c, int ch, int b, uint32_t epoch) {
    __syncthreads();
    const int t = threadIdx.x;
    if (t < c.world) {
        const size_t base = ((size_t)ch * kMaxBlocks + b) * kMaxRanks;
        __threadfence_system();
        st_release_sys(c.signals[t] + base + c.rank, epoch);
        const uint32_t* mine = c.signals[c.rank] + base + t;
        const unsigned long long t0 = globaltimer_ns();
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
            if (globaltimer_ns() - t0 > 4000000000ull) { atomicExch(c.status, 1); break; }  // 4 s: never hang the box
        }
    }
    __syncthreads();
}

struct AggCall {                       // per-call arguments; arrays indexed by emulated rank (blockIdx.y)
    float* local[kMaxRanks];           // fp32 parameters / gradients, in place
    float* prev[kMaxRanks];            // previous global model (NULL: aggregate raw values)
    float* mom[kMaxRanks];             // optional buffer zeroed after aggregation
    float weight[kMaxRanks];
    long long n;
    float server_lr;
    float post_scale;                  // multiplies the reduced value (1/world for gradient means)
    int wire_bf16;
    int channel;
    size_t arena_off;
};

template <bool BF16> struct Wire;
template <> struct Wire<false> {
    using T = float;
    __device__ static T enc(float v) { return v; }
    __device__ static float dec(T v) { return v; }
};
template <> struct Wire<true> {
    using T = __nv_bfloat16;
    __device__ static T enc(float v) { return __float2bfloat16(v); }
    __device__ static float dec(T v) { return __bfloat162float(v); }
};

// ------------------------------------------------------------------------------------------------
// one-shot push aggregate.  Arena layout at arena_off: [parity 2][src rank P][n] wire elements.
// ------------------------------------------------------------------------------------------------
template <bool BF16>
__global__ void __launch_bounds__(kCommThreads) fedavg_oneshot_kernel(const CommDev* ctxs, AggCall a) {
    using W = Wire<BF16>;
    using T = typename W::T;
    const int er = blockIdx.y;                  // emulated rank index (0 in real multi-process runs)
    const CommDev& c = ctxs[er];
    const int P = c.world, b = blockIdx.x, nb = gridDim.x;
    uint32_t* ep = c.epochs + (size_t)a.channel * kMaxBlocks + b;
    const uint32_t epoch = *ep + 1;
    const int parity = epoch & 1;
    const long long n = a.n;
    float* local = a.local[er];
    const float* prev = a.prev[er];
    const float w = a.weight[er];
    const size_t slot_elems = (size_t)((n + 7) & ~7ll);
    const long long lo = n * b / nb, hi = n * (b + 1) / nb;      // this block's element range

    // phase 1: push my weighted (delta) contribution into slot [parity][rank] of every peer
    for (long long i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        const float v = w * (prev ? local[i] - prev[i] : local[i]);
        const T e = W::enc(v);
        for (int r = 0; r < P; ++r) {
            T* dst = reinterpret_cast<T*>(c.arena[(c.rank + r) % P] + a.arena_off) +
                     ((size_t)parity * P + c.rank) * slot_elems;
            dst[i] = e;
        }
    }
    xgpu_barrier(c, a.channel, b, epoch);
    // phase 2: reduce the P slots that landed in my arena, apply the server update in place
    const T* base = reinterpret_cast<const T*>(c.arena[c.rank] + a.arena_off) + (size_t)parity * P * slot_elems;
    for (long long i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        float acc = 0.0f;
        for (int r = 0; r < P; ++r) acc += W::dec(__ldcg(base + (size_t)r * slot_elems + i));
        acc *= a.post_scale;
        float out = prev ? prev[i] + a.server_lr * acc : acc;
        local[i] = out;
        if (prev) a.prev[er][i] = out;
        if (a.mom[er]) a.mom[er][i] = 0.0f;
    }
    if (threadIdx.x == 0) *ep = epoch;
}

// ------------------------------------------------------------------------------------------------
// two-shot: IN region [P ranks' own data is local only][n], OUT region [n]; both in every arena.
// Arena layout at arena_off: IN[n] then OUT[n] (wire elements, 16-byte aligned).
// Element i belongs to slice s = owner rank, and inside a slice to block b: the partition is the
// same on every rank, so block b only ever consumes what peer blocks b produced (per-block flags).
// ------------------------------------------------------------------------------------------------
template <bool BF16>
__global__ void __launch_bounds__(kCommThreads) fedavg_twoshot_kernel(const CommDev* ctxs, AggCall a) {
    using W = Wire<BF16>;
    using T = typename W::T;
    constexpr int V = 16 / sizeof(T);            // elements per 16-byte vector
    const int er = blockIdx.y;
    const CommDev& c = ctxs[er];
    const int P = c.world, b = blockIdx.x, nb = gridDim.x;
    uint32_t* ep = c.epochs + (size_t)a.channel * kMaxBlocks + b;
    const uint32_t e0 = *ep;
    const long long n = a.n;
    float* local = a.local[er];
    const float* prev = a.prev[er];
    const float w = a.weight[er];
    const size_t region = (size_t)((n + 63) & ~63ll);
    // vectors: nv = ceil(n / V); slice s covers vectors [s*vps, (s+1)*vps); block b a sub-range of it
    const long long nv = (n + V - 1) / V;
    const long long vps = (nv + P - 1) / P;                  // vectors per slice
    const long long vpb = (vps + nb - 1) / nb;               // vectors per (slice, block)

    T* my_in = reinterpret_cast<T*>(c.arena[c.rank] + a.arena_off);
    // phase 1: IN = enc(w * (local - prev)) for the (slice, block) pieces this block owns
    for (int s = 0; s < P; ++s) {
        const long long v0 = s * vps + (long long)b * vpb;
        long long v1 = v0 + vpb; if (v1 > (s + 1) * vps) v1 = (s + 1) * vps; if (v1 > nv) v1 = nv;
        for (long long v = v0 + threadIdx.x; v < v1; v += blockDim.x) {
            T tmp[V];
#pragma unroll
            for (int j = 0; j < V; ++j) {
                const long long i = v * V + j;
                float x = 0.0f;
                if (i < n) x = w * (prev ? local[i] - prev[i] : local[i]);
                tmp[j] = W::enc(x);
            }
            *reinterpret_cast<uint4*>(my_in + v * V) = *reinterpret_cast<const uint4*>(tmp);
        }
    }
    xgpu_barrier(c, a.channel, b, e0 + 1);
    // phase 2: reduce my slice from every peer's IN (P2P loads), push the result to every OUT
    {
        const int s = c.rank;
        const long long v0 = s * vps + (long long)b * vpb;
        long long v1 = v0 + vpb; if (v1 > (s + 1) * vps) v1 = (s + 1) * vps; if (v1 > nv) v1 = nv;
        for (long long v = v0 + threadIdx.x; v < v1; v += blockDim.x) {
            float acc[V];
#pragma unroll
            for (int j = 0; j < V; ++j) acc[j] = 0.0f;
            uint4 raw[kMaxRanks];
#pragma unroll
            for (int r = 0; r < kMaxRanks; ++r) {
                if (r < P) {
                    const T* src = reinterpret_cast<const T*>(c.arena[(c.rank + r) % P] + a.arena_off);
                    raw[r] = __ldcg(reinterpret_cast<const uint4*>(src + v * V));
                }
            }
#pragma unroll
            for (int r = 0; r < kMaxRanks; ++r) {
                if (r < P) {
                    const T* t = reinterpret_cast<const T*>(&raw[r]);
#pragma unroll
                    for (int j = 0; j < V; ++j) acc[j] += W::dec(t[j]);
                }
            }
            T outv[V];
#pragma unroll
            for (int j = 0; j < V; ++j) outv[j] = W::enc(acc[j] * a.post_scale);
            const uint4 packed = *reinterpret_cast<const uint4*>(outv);
            for (int r = 0; r < P; ++r) {
                T* dst = reinterpret_cast<T*>(c.arena[(c.rank + r) % P] + a.arena_off) + region;
                *reinterpret_cast<uint4*>(dst + v * V) = packed;
            }
        }
    }
    xgpu_barrier(c, a.channel, b, e0 + 2);
    // phase 3: local server update from OUT
    const T* my_out = reinterpret_cast<const T*>(c.arena[c.rank] + a.arena_off) + region;
    for (int s = 0; s < P; ++s) {
        const long long v0 = s * vps + (long long)b * vpb;
        long long v1 = v0 + vpb; if (v1 > (s + 1) * vps) v1 = (s + 1) * vps; if (v1 > nv) v1 = nv;
        for (long long v = v0 + threadIdx.x; v < v1; v += blockDim.x) {
            const uint4 raw = __ldcg(reinterpret_cast<const uint4*>(my_out + v * V));
            const T* t = reinterpret_cast<const T*>(&raw);
#pragma unroll
            for (int j = 0; j < V; ++j) {
                const long long i = v * V + j;
                if (i < n) {
                    const float d = W::dec(t[j]);
                    const float out = prev ? prev[i] + a.server_lr * d : d;
                    local[i] = out;
                    if (prev) a.prev[er][i] = out;
                    if (a.mom[er]) a.mom[er][i] = 0.0f;
                }
            }
        }
    }
    if (threadIdx.x == 0) *ep = e0 + 2;
}

__global__ void barrier_kernel(const CommDev* ctxs, int channel) {
    const CommDev& c = ctxs[blockIdx.y];
    uint32_t* ep = c.epochs + (size_t)channel * kMaxBlocks + blockIdx.x;
    const uint32_t epoch = *ep + 1;
    xgpu_barrier(c, channel, blockIdx.x, epoch);
    if (threadIdx.x == 0) *ep = epoch;
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int alloc_rank(cs_comm* c) {
    CS_CUDA_OK(cudaGetDevice(&c->device));
    CS_CUDA_OK(cudaMalloc(&c->arena, c->bytes));
    CS_CUDA_OK(cudaMemset(c->arena, 0, c->bytes));
    CS_CUDA_OK(cudaMalloc((void**)&c->signals, kSignalWords * sizeof(uint32_t)));
    CS_CUDA_OK(cudaMemset(c->signals, 0, kSignalWords * sizeof(uint32_t)));
    CS_CUDA_OK(cudaMalloc((void**)&c->epochs, kChannels * kMaxBlocks * sizeof(uint32_t)));
    CS_CUDA_OK(cudaMemset(c->epochs, 0, kChannels * kMaxBlocks * sizeof(uint32_t)));
    CS_CUDA_OK(cudaMalloc((void**)&c->status, sizeof(int)));
    CS_CUDA_OK(cudaMemset(c->status, 0, sizeof(int)));
    CS_CUDA_OK(cudaMalloc((void**)&c->dev, sizeof(CommDev)));
    return CS_OK;
}

static int upload_ctx(cs_comm* c, CommDev* dst) {
    CommDev h;
    memset(&h, 0, sizeof(h));
    h.rank = c->rank; h.world = c->world; h.epochs = c->epochs; h.status = c->status; h.arena_bytes = c->bytes;
    for (int r = 0; r < c->world; ++r) { h.arena[r] = (char*)c->peer_arena[r]; h.signals[r] = c->peer_signals[r]; }
    CS_CUDA_OK(cudaMemcpy(dst, &h, sizeof(h), cudaMemcpyHostToDevice));
    return CS_OK;
}

static size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

int comm_create(int rank, int world, size_t bytes, cs_comm** out, void* ah, void* sh) {
    CS_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, "comm: bad rank/world %d/%d", rank, world);
    static_assert(sizeof(cudaIpcMemHandle_t) <= CS_IPC_HANDLE_BYTES, "ipc handle size");
    cs_comm* c = new cs_comm();
    c->rank = rank; c->world = world; c->bytes = round_up(bytes < 4096 ? 4096 : bytes, 4096);
    int rc = alloc_rank(c);
    if (rc) { delete c; return rc; }
    c->peer_arena[rank] = c->arena; c->peer_signals[rank] = c->signals;
    memset(ah, 0, CS_IPC_HANDLE_BYTES); memset(sh, 0, CS_IPC_HANDLE_BYTES);
    if (world > 1) {
        cudaIpcMemHandle_t h;
        CS_CUDA_OK(cudaIpcGetMemHandle(&h, c->arena));   memcpy(ah, &h, sizeof(h));
        CS_CUDA_OK(cudaIpcGetMemHandle(&h, c->signals)); memcpy(sh, &h, sizeof(h));
    } else {
        rc = upload_ctx(c, c->dev);
        if (rc) return rc;
    }
    *out = c;
    return CS_OK;
}

int comm_connect(cs_comm* c, const void* ahs, const void* shs) {
    CS_REQUIRE(c != nullptr, "comm: null");
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char*)ahs + (size_t)r * CS_IPC_HANDLE_BYTES, sizeof(h));
        CS_CUDA_OK(cudaIpcOpenMemHandle(&c->peer_arena[r], h, cudaIpcMemLazyEnablePeerAccess));
        memcpy(&h, (const char*)shs + (size_t)r * CS_IPC_HANDLE_BYTES, sizeof(h));
        CS_CUDA_OK(cudaIpcOpenMemHandle((void**)&c->peer_signals[r], h, cudaIpcMemLazyEnablePeerAccess));
        c->opened[r] = true;
    }
    return upload_ctx(c, c->dev);
}

int comm_create_local(int world, size_t bytes, cs_comm** out) {
    CS_REQUIRE(world >= 1 && world <= kMaxRanks, "comm: bad world %d", world);
    cs_comm** g = new cs_comm*[world];
    for (int r = 0; r < world; ++r) {
        g[r] = new cs_comm();
        g[r]->rank = r; g[r]->world = world; g[r]->emulated = true;
        g[r]->bytes = round_up(bytes < 4096 ? 4096 : bytes, 4096);
        int rc = alloc_rank(g[r]);
        if (rc) return rc;
    }
    CommDev* arr = nullptr;
    CS_CUDA_OK(cudaMalloc((void**)&arr, sizeof(CommDev) * world));
    for (int r = 0; r < world; ++r) {
        for (int q = 0; q < world; ++q) { g[r]->peer_arena[q] = g[q]->arena; g[r]->peer_signals[q] = g[q]->signals; }
        g[r]->group = g; g[r]->group_dev = arr;
        int rc = upload_ctx(g[r], arr + r);
        if (rc) return rc;
        rc = upload_ctx(g[r], g[r]->dev);
        if (rc) return rc;
        out[r] = g[r];
    }
    return CS_OK;
}

int comm_destroy(cs_comm* c) {
    if (!c) return CS_OK;
    if (c->emulated) {
        if (c->rank != 0) return CS_OK;      // member 0 tears the whole group down
        cs_comm** g = c->group; int world = c->world;
        cudaFree(c->group_dev);
        for (int r = 0; r < world; ++r) {
            cudaFree(g[r]->arena); cudaFree(g[r]->signals); cudaFree(g[r]->epochs); cudaFree(g[r]->status); cudaFree(g[r]->dev);
            delete g[r];
        }
        delete[] g;
        return CS_OK;
    }
    for (int r = 0; r < c->world; ++r)
        if (c->opened[r]) { cudaIpcCloseMemHandle(c->peer_arena[r]); cudaIpcCloseMemHandle(c->peer_signals[r]); }
    cudaFree(c->arena); cudaFree(c->signals); cudaFree(c->epochs); cudaFree(c->status); cudaFree(c->dev);
    delete c;
    return CS_OK;
}

static int pick_blocks(long long n, bool oneshot) {
    if (oneshot) { long long b = (n + 4095) / 4096; return (int)(b < 1 ? 1 : (b > 16 ? 16 : b)); }
    int cap = num_sms() < kMaxBlocks ? num_sms() : kMaxBlocks;
    long long b = (n + 16383) / 16384;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

size_t agg_arena_need(long long n, int wire_bf16, int world) {
    const size_t es = wire_bf16 ? 2 : 4;
    if (n <= kOneShotMaxElems) return 2 * (size_t)world * (size_t)((n + 7) & ~7ll) * es;
    return 2 * (size_t)((n + 63) & ~63ll) * es;
}

// Launch for a real rank (ranks = 1 context) or for an emulated group (ranks = world contexts).
static int launch_agg(const CommDev* ctxs, int ranks, int world, size_t arena_bytes, const AggCall& a, cudaStream_t st) {
    CS_REQUIRE(a.n > 0, "aggregate: n must be positive");
    CS_REQUIRE(a.channel >= 0 && a.channel < kChannels, "aggregate: channel %d out of range", a.channel);
    CS_REQUIRE(a.arena_off % 16 == 0, "aggregate: arena offset must be 16-byte aligned");
    CS_REQUIRE(a.arena_off + agg_arena_need(a.n, a.wire_bf16, world) <= arena_bytes,
               "aggregate: arena too small (%zu B needed at offset %zu, have %zu)",
               agg_arena_need(a.n, a.wire_bf16, world), a.arena_off, arena_bytes);
    const bool oneshot = a.n <= kOneShotMaxElems;
    int nb = pick_blocks(a.n, oneshot);
    if (ranks > 1) {   // emulated ranks share one GPU: keep the whole grid co-resident
        int cap = (num_sms() * 2) / ranks;
        if (nb > cap) nb = cap < 1 ? 1 : cap;
    }
    dim3 grid(nb, ranks), block(kCommThreads);
    void* args[] = {(void*)&ctxs, (void*)&a};
    const void* fn;
    if (oneshot) fn = a.wire_bf16 ? (const void*)fedavg_oneshot_kernel<true> : (const void*)fedavg_oneshot_kernel<false>;
    else fn = a.wire_bf16 ? (const void*)fedavg_twoshot_kernel<true> : (const void*)fedavg_twoshot_kernel<false>;
    if (ranks > 1) {   // emulation: every rank's blocks must be co-resident -> cooperative launch
        CS_CUDA_OK(cudaLaunchCooperativeKernel(fn, grid, block, args, 0, st));
    } else {
        CS_CUDA_OK(cudaLaunchKernel(fn, grid, block, args, 0, st));
    }
    return CS_OK;
}

int fedavg_aggregate(cs_comm* c, float* local, float* prev, long long n, float weight, int wire_bf16,
                     float server_lr, float* mom, size_t arena_off, int channel, float post_scale, cudaStream_t st) {
    CS_REQUIRE(c && !c->emulated, "aggregate: use the *_local entry point for emulated groups");
    AggCall a; memset(&a, 0, sizeof(a));
    a.local[0] = local; a.prev[0] = prev; a.mom[0] = mom; a.weight[0] = weight;
    a.n = n; a.server_lr = server_lr; a.post_scale = post_scale; a.wire_bf16 = wire_bf16; a.channel = channel; a.arena_off = arena_off;
    return launch_agg(c->dev, 1, c->world, c->bytes, a, st);
}

int fedavg_aggregate_local(cs_comm** g, int world, float* const* local, float* const* prev, long long n,
                           const float* weights, int wire_bf16, float server_lr, float* const* mom,
                           size_t arena_off, int channel, float post_scale, cudaStream_t st) {
    CS_REQUIRE(g && g[0] && g[0]->emulated && g[0]->world == world, "aggregate_local: not an emulated group");
    AggCall a; memset(&a, 0, sizeof(a));
    for (int r = 0; r < world; ++r) {
        a.local[r] = local[r]; a.prev[r] = prev ? prev[r] : nullptr; a.mom[r] = mom ? mom[r] : nullptr; a.weight[r] = weights[r];
    }
    a.n = n; a.server_lr = server_lr; a.post_scale = post_scale; a.wire_bf16 = wire_bf16; a.channel = channel; a.arena_off = arena_off;
    return launch_agg(g[0]->group_dev, world, world, g[0]->bytes, a, st);
}

int comm_barrier(cs_comm* c, int channel, cudaStream_t st) {
    CS_REQUIRE(c != nullptr, "barrier: null comm");
    CS_REQUIRE(channel >= 0 && channel < kChannels, "barrier: bad channel");
    if (c->emulated) {
        const CommDev* ctxs = c->group_dev; dim3 grid(1, c->world), block(32);
        void* args[] = {(void*)&ctxs, (void*)&channel};
        CS_CUDA_OK(cudaLaunchCooperativeKernel((const void*)barrier_kernel, grid, block, args, 0, st));
    } else {
        barrier_kernel<<<dim3(1, 1), 32, 0, st>>>(c->dev, channel);
        CS_CUDA_OK(cudaGetLastError());
    }
    return CS_OK;
}

int comm_status(cs_comm* c, int* out) {
    CS_CUDA_OK(cudaMemcpy(out, c->status, sizeof(int), cudaMemcpyDeviceToHost));
    return CS_OK;
}

}  // namespace cs

namespace cs {
void* comm_arena(cs_comm* c, int r) { return c ? (r < 0 ? c->arena : c->peer_arena[r]) : nullptr; }
size_t comm_bytes(cs_comm* c) { return c ? c->bytes : 0; }
}  // namespace cs
