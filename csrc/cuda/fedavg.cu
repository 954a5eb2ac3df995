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
#include "comm_dev.cuh"

#include <stdlib.h>
#include <string.h>
#include <vector>

namespace cs {

// Cross-GPU barrier between block `b` of every rank on channel `ch`.  Epochs are kept in device
// memory (not kernel arguments) so that a captured CUDA graph can be replayed.
// Returns false (to every thread of the block) when a peer's flag did not arrive within the communicator's time-out:
// the status word is raised and the caller must NOT consume peer data - it skips its reduction / write-back so that
// the local model stays what local training left, instead of a sum over stale slots.
__device__ __forceinline__ bool xgpu_barrier(const CommDev& c, int ch, int b, uint32_t epoch) {
    __syncthreads();
    const int t = threadIdx.x;
    int late = 0;
    if (t < c.world) {
        const size_t base = ((size_t)ch * kMaxBlocks + b) * kMaxRanks;
        __threadfence_system();
        st_release_sys(c.signals[t] + base + c.rank, epoch);
        const uint32_t* mine = c.signals[c.rank] + base + t;
        const unsigned long long t0 = globaltimer_ns();
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
            if (globaltimer_ns() - t0 > c.timeout_ns) { *(volatile int*)c.status = 1; late = 1; break; }   // never hang the box
        }
    }
    return __syncthreads_or(late) == 0;
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
    if (!xgpu_barrier(c, a.channel, b, epoch)) { if (threadIdx.x == 0) *ep = epoch; return; }   // peer missing: keep the local model
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
// two-shot.  Arena layout at arena_off: IN[n] then OUT[n] (wire elements).  The vector is cut into groups
// of 4 elements; slice s (owner rank s) and, inside a slice, block b get fixed group ranges — the same on
// every rank — so block b only ever consumes what peer blocks b produced (per-block flags suffice and no
// block waits for another block of its own GPU).  Every loop issues kU independent 16-byte loads per
// thread before the first use: ~5 MB in flight per GPU, enough to cover HBM and NVLink latency.
// ------------------------------------------------------------------------------------------------
constexpr int kU = 4;
static int env_knob(const char* name, int dflt) { const char* e = getenv(name); return (e && *e) ? atoi(e) : dflt; }

template <bool BF16> struct Wire4;
template <> struct Wire4<false> {
    static constexpr int BYTES = 16;
    __device__ static void st(char* base, long long g, float4 v) { reinterpret_cast<float4*>(base)[g] = v; }
    __device__ static float4 ld(const char* base, long long g) { return __ldcg(reinterpret_cast<const float4*>(base) + g); }
};
template <> struct Wire4<true> {
    static constexpr int BYTES = 8;
    __device__ static void st(char* base, long long g, float4 v) {
        __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
        reinterpret_cast<uint2*>(base)[g] = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
    }
    __device__ static float4 ld(const char* base, long long g) {
        const uint2 r = __ldcg(reinterpret_cast<const uint2*>(base) + g);
        const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&r.x));
        const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&r.y));
        return make_float4(a.x, a.y, b.x, b.y);
    }
};

// 4 floats of a possibly ragged fp32 vector (the last group may be partial)
__device__ __forceinline__ float4 ld4_tail(const float* p, long long g, long long n) {
    if (4 * g + 3 < n) return *reinterpret_cast<const float4*>(p + 4 * g);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (4 * g < n) v.x = p[4 * g];
    if (4 * g + 1 < n) v.y = p[4 * g + 1];
    if (4 * g + 2 < n) v.z = p[4 * g + 2];
    return v;
}
__device__ __forceinline__ void st4_tail(float* p, long long g, long long n, float4 v) {
    if (4 * g + 3 < n) { *reinterpret_cast<float4*>(p + 4 * g) = v; return; }
    if (4 * g < n) p[4 * g] = v.x;
    if (4 * g + 1 < n) p[4 * g + 1] = v.y;
    if (4 * g + 2 < n) p[4 * g + 2] = v.z;
}

template <bool BF16>
__global__ void __launch_bounds__(kCommThreads) fedavg_twoshot_kernel(const CommDev* ctxs, AggCall a) {
    using W = Wire4<BF16>;
    const int er = blockIdx.y;
    const CommDev& c = ctxs[er];
    const int P = c.world, b = blockIdx.x, nb = gridDim.x, T = blockDim.x, tid = threadIdx.x;
    uint32_t* ep = c.epochs + (size_t)a.channel * kMaxBlocks + b;
    const uint32_t e0 = *ep;
    const long long n = a.n;
    float* local = a.local[er];
    float* prev = a.prev[er];
    float* mom = a.mom[er];
    const float w = a.weight[er];
    const long long n4 = (n + 3) >> 2;                       // groups of 4 elements
    const long long gps = (n4 + P - 1) / P;                  // groups per slice
    const long long gpb = (gps + nb - 1) / nb;               // groups per (slice, block)
    const size_t region = (size_t)((n4 * W::BYTES + 255) & ~255ll);   // bytes of IN (OUT follows)
    char* my_in = c.arena[c.rank] + a.arena_off;
    const char* my_out = my_in + region;

    // phase 1: IN = enc(w * (local - prev)) for every (slice, this block) piece
    for (int s = 0; s < P; ++s) {
        const long long g0 = s * gps + (long long)b * gpb;
        long long g1 = g0 + gpb; if (g1 > (s + 1) * gps) g1 = (s + 1) * gps; if (g1 > n4) g1 = n4;
        for (long long g = g0 + tid; g < g1; g += (long long)kU * T) {
            float4 x[kU], q[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const long long gi = g + (long long)u * T;
                if (gi < g1) { x[u] = ld4_tail(local, gi, n); if (prev) q[u] = ld4_tail(prev, gi, n); }
            }
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const long long gi = g + (long long)u * T;
                if (gi < g1) {
                    float4 v = x[u];
                    if (prev) { v.x -= q[u].x; v.y -= q[u].y; v.z -= q[u].z; v.w -= q[u].w; }
                    W::st(my_in, gi, make_float4(w * v.x, w * v.y, w * v.z, w * v.w));
                }
            }
        }
    }
    if (!xgpu_barrier(c, a.channel, b, e0 + 1)) { if (threadIdx.x == 0) *ep = e0 + 2; return; }
    // phase 2: reduce my slice from every peer's IN (P2P loads), push the result into every peer's OUT
    {
        const int s = c.rank;
        const long long g0 = s * gps + (long long)b * gpb;
        long long g1 = g0 + gpb; if (g1 > (s + 1) * gps) g1 = (s + 1) * gps; if (g1 > n4) g1 = n4;
        constexpr int U2 = 2;
        for (long long g = g0 + tid; g < g1; g += (long long)U2 * T) {
            float4 r[U2][kMaxRanks];
#pragma unroll
            for (int u = 0; u < U2; ++u) {
                const long long gi = g + (long long)u * T;
#pragma unroll
                for (int p = 0; p < kMaxRanks; ++p)
                    if (p < P && gi < g1) r[u][p] = W::ld(c.arena[(c.rank + p) % P] + a.arena_off, gi);
            }
#pragma unroll
            for (int u = 0; u < U2; ++u) {
                const long long gi = g + (long long)u * T;
                if (gi < g1) {
                    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int p = 0; p < kMaxRanks; ++p)
                        if (p < P) { acc.x += r[u][p].x; acc.y += r[u][p].y; acc.z += r[u][p].z; acc.w += r[u][p].w; }
                    acc.x *= a.post_scale; acc.y *= a.post_scale; acc.z *= a.post_scale; acc.w *= a.post_scale;
                    for (int p = 0; p < P; ++p) W::st(c.arena[(c.rank + p) % P] + a.arena_off + region, gi, acc);
                }
            }
        }
    }
    if (!xgpu_barrier(c, a.channel, b, e0 + 2)) { if (threadIdx.x == 0) *ep = e0 + 2; return; }
    // phase 3: local server update from OUT
    for (int s = 0; s < P; ++s) {
        const long long g0 = s * gps + (long long)b * gpb;
        long long g1 = g0 + gpb; if (g1 > (s + 1) * gps) g1 = (s + 1) * gps; if (g1 > n4) g1 = n4;
        for (long long g = g0 + tid; g < g1; g += (long long)kU * T) {
            float4 d[kU], q[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const long long gi = g + (long long)u * T;
                if (gi < g1) { d[u] = W::ld(my_out, gi); if (prev) q[u] = ld4_tail(prev, gi, n); }
            }
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const long long gi = g + (long long)u * T;
                if (gi < g1) {
                    float4 o = d[u];
                    if (prev) {
                        o = make_float4(q[u].x + a.server_lr * o.x, q[u].y + a.server_lr * o.y, q[u].z + a.server_lr * o.z,
                                        q[u].w + a.server_lr * o.w);
                        st4_tail(prev, gi, n, o);
                    }
                    st4_tail(local, gi, n, o);
                    if (mom) st4_tail(mom, gi, n, make_float4(0.f, 0.f, 0.f, 0.f));
                }
            }
        }
    }
    if (threadIdx.x == 0) *ep = e0 + 2;
}

// ------------------------------------------------------------------------------------------------
// in-place two-shot on symmetric parameters: the model's fp32 parameters LIVE in the arena (same offset
// on every rank), so there is no staging at all.  Rank r reads slice r of every peer's parameters over
// NVLink, forms sum_j w_j theta_j (optionally theta_prev + lr * (avg - theta_prev)) and stores the result
// into slice r of EVERY replica.  Slice r is read only by rank r and written only by rank r, so two
// barriers (all local training finished / all pushes landed) are the whole protocol.  Per GPU this moves
// (P-1)/P * n floats in each NVLink direction and touches HBM once — the bandwidth-optimal schedule.
// ------------------------------------------------------------------------------------------------
struct InplaceCall {
    size_t param_off[kMaxRanks];       // byte offset of the parameter vector inside each (emulated) rank's arena
    float* prev[kMaxRanks];            // optional previous global model (slice-local use only), may be NULL
    float* mom[kMaxRanks];             // optional buffer zeroed after aggregation
    float weights[kMaxRanks];          // aggregation weight of every rank (same array on all ranks)
    long long n;
    float server_lr;
    int channel;
};

template <int U2>
__global__ void __launch_bounds__(kCommThreads) fedavg_inplace_kernel(const CommDev* ctxs, InplaceCall a) {
    const int er = blockIdx.y;
    const CommDev& c = ctxs[er];
    const int P = c.world, b = blockIdx.x, nb = gridDim.x, T = blockDim.x, tid = threadIdx.x;
    uint32_t* ep = c.epochs + (size_t)a.channel * kMaxBlocks + b;
    const uint32_t e0 = *ep;
    const long long n4 = a.n >> 2;                           // n is a multiple of 4 (checked on the host)
    const long long gps = (n4 + P - 1) / P, gpb = (gps + nb - 1) / nb;
    float* prev = a.prev[er];
    float* mom = a.mom[er];
    if (!xgpu_barrier(c, a.channel, b, e0 + 1)) { if (threadIdx.x == 0) *ep = e0 + 2; return; }   // every rank's local training is complete
    {
        const int s = c.rank;
        const long long g0 = s * gps + (long long)b * gpb;
        long long g1 = g0 + gpb; if (g1 > (s + 1) * gps) g1 = (s + 1) * gps; if (g1 > n4) g1 = n4;
        
        for (long long g = g0 + tid; g < g1; g += (long long)U2 * T) {
            float4 r[U2][kMaxRanks], q[U2];
#pragma unroll
            for (int u = 0; u < U2; ++u) {
                const long long gi = g + (long long)u * T;
#pragma unroll
                for (int p = 0; p < kMaxRanks; ++p)
                    if (p < P && gi < g1) {
                        const int src = (c.rank + p) % P;
                        r[u][p] = __ldcg(reinterpret_cast<const float4*>(c.arena[src] + a.param_off[er]) + gi);
                    }
                if (prev && gi < g1) q[u] = reinterpret_cast<const float4*>(prev)[gi];
            }
#pragma unroll
            for (int u = 0; u < U2; ++u) {
                const long long gi = g + (long long)u * T;
                if (gi < g1) {
                    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int p = 0; p < kMaxRanks; ++p)
                        if (p < P) {
                            const float w = a.weights[(c.rank + p) % P];
                            acc.x = fmaf(w, r[u][p].x, acc.x); acc.y = fmaf(w, r[u][p].y, acc.y);
                            acc.z = fmaf(w, r[u][p].z, acc.z); acc.w = fmaf(w, r[u][p].w, acc.w);
                        }
                    if (prev) {
                        acc = make_float4(q[u].x + a.server_lr * (acc.x - q[u].x), q[u].y + a.server_lr * (acc.y - q[u].y),
                                          q[u].z + a.server_lr * (acc.z - q[u].z), q[u].w + a.server_lr * (acc.w - q[u].w));
                        reinterpret_cast<float4*>(prev)[gi] = acc;
                    }
                    for (int p = 0; p < P; ++p)
                        reinterpret_cast<float4*>(c.arena[(c.rank + p) % P] + a.param_off[er])[gi] = acc;
                }
            }
        }
    }
    if (mom) {                                               // local momentum reset (all slices)
        for (long long g = (long long)b * T + tid; g < n4; g += (long long)nb * T)
            reinterpret_cast<float4*>(mom)[g] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    xgpu_barrier(c, a.channel, b, e0 + 2);                   // every push has landed in my replica
    if (threadIdx.x == 0) *ep = e0 + 2;
}

// ------------------------------------------------------------------------------------------------
// Bucket-fused FedAvg for large models (K7 "bucketed so it pipelines with the producing kernel").  One launch per
// gradient bucket, issued on a side stream the moment the LAST local step's backward has produced that bucket:
//   phase 0  this rank's SGD-momentum update of the bucket (the optimizer step of the last local step, in place in
//            the symmetric arena) - the separate optimizer launch of that step disappears;
//   barrier  every rank has updated this bucket;
//   phase 1  in-place two-shot on the bucket: rank r reads slice r of every replica over NVLink, forms the weighted
//            mean (+ server update) and stores it into slice r of EVERY replica;
//   barrier  every push has landed;
//   phase 2  bf16 shadow of the bucket (what the conv kernels read) - the separate full-model cast pass disappears.
// While bucket i is on the wire the compute stream is still running backward for bucket i+1, so only the last
// bucket's collective is exposed.  Work is partitioned like the in-place kernel: block b only ever consumes what
// block b of a peer produced, so per-block flags suffice and no block waits on another block of its own GPU.
// ------------------------------------------------------------------------------------------------
struct BucketCall {
    size_t param_off[kMaxRanks];       // byte offset of the bucket's parameters inside each (emulated) rank's arena
    const float* grad[kMaxRanks];      // bucket-local pointers
    float* mom[kMaxRanks];
    float* prev[kMaxRanks];            // optional previous global model of the bucket (server_lr != 1; slice-local use)
    void* shadow[kMaxRanks];           // optional bf16 copy of the bucket
    float weights[kMaxRanks];
    long long n;                       // multiple of 4
    float lr, momentum, wd, server_lr, inv_scale;
    int do_sgd, channel;
};

__global__ void __launch_bounds__(kCommThreads) fedavg_bucket_kernel(const CommDev* ctxs, BucketCall a) {
    const int er = blockIdx.y;
    const CommDev& c = ctxs[er];
    const int P = c.world, b = blockIdx.x, nb = gridDim.x, T = blockDim.x, tid = threadIdx.x;
    uint32_t* ep = c.epochs + (size_t)a.channel * kMaxBlocks + b;
    const uint32_t e0 = *ep;
    const long long n4 = a.n >> 2;
    const long long gps = (n4 + P - 1) / P, gpb = (gps + nb - 1) / nb;
    float4* mine = reinterpret_cast<float4*>(c.arena[c.rank] + a.param_off[er]);
    float* prev = a.prev[er];
    if (a.do_sgd) {                                            // phase 0: local optimizer step on every (slice, this block) piece
        const float4* grad = reinterpret_cast<const float4*>(a.grad[er]);
        float4* mom = reinterpret_cast<float4*>(a.mom[er]);
        const float mu = a.momentum, wd = a.wd, lr = a.lr, is = a.inv_scale;
        for (int s = 0; s < P; ++s) {
            const long long g0 = s * gps + (long long)b * gpb;
            long long g1 = g0 + gpb; if (g1 > (s + 1) * gps) g1 = (s + 1) * gps; if (g1 > n4) g1 = n4;
            for (long long g = g0 + tid; g < g1; g += (long long)kU * T) {
                float4 p[kU], m[kU], d[kU];
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    const long long gi = g + (long long)u * T;
                    if (gi < g1) { p[u] = mine[gi]; m[u] = mom[gi]; d[u] = __ldcs(grad + gi); }
                }
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    const long long gi = g + (long long)u * T;
                    if (gi < g1) {
                        m[u].x = mu * m[u].x + (d[u].x * is + wd * p[u].x); p[u].x -= lr * m[u].x;
                        m[u].y = mu * m[u].y + (d[u].y * is + wd * p[u].y); p[u].y -= lr * m[u].y;
                        m[u].z = mu * m[u].z + (d[u].z * is + wd * p[u].z); p[u].z -= lr * m[u].z;
                        m[u].w = mu * m[u].w + (d[u].w * is + wd * p[u].w); p[u].w -= lr * m[u].w;
                        mom[gi] = m[u]; mine[gi] = p[u];
                    }
                }
            }
        }
    }
    const bool ok = xgpu_barrier(c, a.channel, b, e0 + 1);     // every rank has stepped this bucket
    if (ok) {
        const int s = c.rank;
        const long long g0 = s * gps + (long long)b * gpb;
        long long g1 = g0 + gpb; if (g1 > (s + 1) * gps) g1 = (s + 1) * gps; if (g1 > n4) g1 = n4;
        constexpr int U2 = 2;
        for (long long g = g0 + tid; g < g1; g += (long long)U2 * T) {
            float4 r[U2][kMaxRanks], q[U2];
#pragma unroll
            for (int u = 0; u < U2; ++u) {
                const long long gi = g + (long long)u * T;
#pragma unroll
                for (int p = 0; p < kMaxRanks; ++p)
                    if (p < P && gi < g1)
                        r[u][p] = __ldcg(reinterpret_cast<const float4*>(c.arena[(c.rank + p) % P] + a.param_off[er]) + gi);
                if (prev && gi < g1) q[u] = reinterpret_cast<const float4*>(prev)[gi];
            }
#pragma unroll
            for (int u = 0; u < U2; ++u) {
                const long long gi = g + (long long)u * T;
                if (gi < g1) {
                    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                    for (int p = 0; p < kMaxRanks; ++p)
                        if (p < P) {
                            const float w = a.weights[(c.rank + p) % P];
                            acc.x = fmaf(w, r[u][p].x, acc.x); acc.y = fmaf(w, r[u][p].y, acc.y);
                            acc.z = fmaf(w, r[u][p].z, acc.z); acc.w = fmaf(w, r[u][p].w, acc.w);
                        }
                    if (prev) {
                        acc = make_float4(q[u].x + a.server_lr * (acc.x - q[u].x), q[u].y + a.server_lr * (acc.y - q[u].y),
                                          q[u].z + a.server_lr * (acc.z - q[u].z), q[u].w + a.server_lr * (acc.w - q[u].w));
                        reinterpret_cast<float4*>(prev)[gi] = acc;
                    }
                    for (int p = 0; p < P; ++p)
                        reinterpret_cast<float4*>(c.arena[(c.rank + p) % P] + a.param_off[er])[gi] = acc;
                }
            }
        }
    }
    if (ok) xgpu_barrier(c, a.channel, b, e0 + 2);             // every push has landed in my replica
    if (a.shadow[er]) {                                        // phase 2: bf16 shadow of the whole bucket (local)
        uint2* sh = reinterpret_cast<uint2*>(a.shadow[er]);
        for (int s = 0; s < P; ++s) {
            const long long g0 = s * gps + (long long)b * gpb;
            long long g1 = g0 + gpb; if (g1 > (s + 1) * gps) g1 = (s + 1) * gps; if (g1 > n4) g1 = n4;
            for (long long g = g0 + tid; g < g1; g += (long long)kU * T) {
                float4 p[kU];
#pragma unroll
                for (int u = 0; u < kU; ++u) { const long long gi = g + (long long)u * T; if (gi < g1) p[u] = __ldcg(mine + gi); }
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    const long long gi = g + (long long)u * T;
                    if (gi < g1) {
                        __nv_bfloat162 lo = __floats2bfloat162_rn(p[u].x, p[u].y), hi = __floats2bfloat162_rn(p[u].z, p[u].w);
                        sh[gi] = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
                    }
                }
            }
        }
    }
    if (threadIdx.x == 0) *ep = e0 + 2;
}

// ------------------------------------------------------------------------------------------------
// NVLS variant of the in-place aggregation: the parameters live in a symmetric buffer that is also mapped
// through an NVSwitch MULTICAST address.  Rank r pulls the switch-reduced slice r with
// multimem.ld_reduce (the switch fetches and adds all replicas: N/P floats come in instead of (P-1)N/P),
// applies the server update and broadcasts it to every replica with ONE multimem.st.  Per GPU and
// direction that is ~N floats on the wire instead of ~2N(P-1)/P for the peer-to-peer version.
// Unequal client weights are applied by pre-scaling the local replica (one extra local pass).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 multimem_ld_reduce_add_f32x4(const void* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st_f32x4(void* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

struct NvlsCall {
    float* local;          // this rank's replica (unicast address inside the symmetric buffer)
    char* mc;              // multicast address of the same buffer
    float* prev; float* mom;
    long long n;
    float prescale, post_scale, server_lr;
    int channel;
};

template <int kU>
__global__ void __launch_bounds__(kCommThreads) fedavg_nvls_kernel(const CommDev* ctxs, NvlsCall a) {
    const CommDev& c = ctxs[0];
    const int P = c.world, b = blockIdx.x, nb = gridDim.x, T = blockDim.x, tid = threadIdx.x;
    uint32_t* ep = c.epochs + (size_t)a.channel * kMaxBlocks + b;
    const uint32_t e0 = *ep;
    const long long n4 = a.n >> 2;
    const long long gps = (n4 + P - 1) / P, gpb = (gps + nb - 1) / nb;
    if (a.prescale != 1.0f) {                                   // theta_i *= w_i * P (so that the plain switch sum / P is the weighted mean)
        for (int s = 0; s < P; ++s) {
            const long long g0 = s * gps + (long long)b * gpb;
            long long g1 = g0 + gpb; if (g1 > (s + 1) * gps) g1 = (s + 1) * gps; if (g1 > n4) g1 = n4;
            for (long long g = g0 + tid; g < g1; g += T) {
                float4 v = reinterpret_cast<float4*>(a.local)[g];
                v.x *= a.prescale; v.y *= a.prescale; v.z *= a.prescale; v.w *= a.prescale;
                reinterpret_cast<float4*>(a.local)[g] = v;
            }
        }
    }
    if (!xgpu_barrier(c, a.channel, b, e0 + 1)) {               // peer missing: undo the pre-scaling, touch nothing else
        if (a.prescale != 1.0f) {
            const float inv = 1.0f / a.prescale;
            for (int s = 0; s < P; ++s) {
                const long long g0 = s * gps + (long long)b * gpb;
                long long g1 = g0 + gpb; if (g1 > (s + 1) * gps) g1 = (s + 1) * gps; if (g1 > n4) g1 = n4;
                for (long long g = g0 + tid; g < g1; g += T) {
                    float4 v = reinterpret_cast<float4*>(a.local)[g];
                    v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
                    reinterpret_cast<float4*>(a.local)[g] = v;
                }
            }
        }
        if (threadIdx.x == 0) *ep = e0 + 2;
        return;
    }
    {
        const int s = c.rank;
        const long long g0 = s * gps + (long long)b * gpb;
        long long g1 = g0 + gpb; if (g1 > (s + 1) * gps) g1 = (s + 1) * gps; if (g1 > n4) g1 = n4;
        for (long long g = g0 + tid; g < g1; g += (long long)kU * T) {
            float4 v[kU], q[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const long long gi = g + (long long)u * T;
                if (gi < g1) { v[u] = multimem_ld_reduce_add_f32x4(a.mc + gi * 16); if (a.prev) q[u] = reinterpret_cast<const float4*>(a.prev)[gi]; }
            }
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const long long gi = g + (long long)u * T;
                if (gi < g1) {
                    float4 o = make_float4(v[u].x * a.post_scale, v[u].y * a.post_scale, v[u].z * a.post_scale, v[u].w * a.post_scale);
                    if (a.prev) {
                        o = make_float4(q[u].x + a.server_lr * (o.x - q[u].x), q[u].y + a.server_lr * (o.y - q[u].y),
                                        q[u].z + a.server_lr * (o.z - q[u].z), q[u].w + a.server_lr * (o.w - q[u].w));
                        reinterpret_cast<float4*>(a.prev)[gi] = o;
                    }
                    multimem_st_f32x4(a.mc + gi * 16, o);
                }
            }
        }
    }
    if (a.mom) {
        for (long long g = (long long)b * T + tid; g < n4; g += (long long)nb * T)
            reinterpret_cast<float4*>(a.mom)[g] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    xgpu_barrier(c, a.channel, b, e0 + 2);
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
    CS_CUDA_OK(cudaHostAlloc((void**)&c->status_host, sizeof(int), cudaHostAllocMapped));
    *c->status_host = 0;
    CS_CUDA_OK(cudaHostGetDevicePointer((void**)&c->status, c->status_host, 0));
    if (const char* e = getenv("CS_COMM_TIMEOUT_MS")) { const long long ms = atoll(e); if (ms > 0) c->timeout_ns = (unsigned long long)ms * 1000000ull; }
    CS_CUDA_OK(cudaMalloc((void**)&c->dev, sizeof(CommDev)));
    return CS_OK;
}

static int upload_ctx(cs_comm* c, CommDev* dst) {
    CommDev h;
    memset(&h, 0, sizeof(h));
    h.rank = c->rank; h.world = c->world; h.epochs = c->epochs; h.status = c->status; h.arena_bytes = c->bytes; h.timeout_ns = c->timeout_ns;
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

// Communicator over mappings the caller made itself (symm_vmm.cu: VMM allocations imported from the peers).  arena[r] /
// signals[r]: this process's mapping of rank r's arena and signal pad (r == rank: the local allocation); the signal pad must
// hold kSignalWords zeroed 32-bit words.
int comm_create_external(int rank, int world, size_t bytes, void* const* arena, void* const* signals, cs_comm** out) {
    CS_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, "comm: bad rank/world %d/%d", rank, world);
    for (int r = 0; r < world; ++r) CS_REQUIRE(arena[r] && signals[r], "comm_create_external: null mapping for rank %d", r);
    cs_comm* c = new cs_comm();
    c->rank = rank; c->world = world; c->bytes = bytes; c->external = true;
    CS_CUDA_OK(cudaGetDevice(&c->device));
    c->arena = arena[rank]; c->signals = (uint32_t*)signals[rank];
    CS_CUDA_OK(cudaMalloc((void**)&c->epochs, kChannels * kMaxBlocks * sizeof(uint32_t)));
    CS_CUDA_OK(cudaMemset(c->epochs, 0, kChannels * kMaxBlocks * sizeof(uint32_t)));
    CS_CUDA_OK(cudaHostAlloc((void**)&c->status_host, sizeof(int), cudaHostAllocMapped));
    *c->status_host = 0;
    CS_CUDA_OK(cudaHostGetDevicePointer((void**)&c->status, c->status_host, 0));
    if (const char* e = getenv("CS_COMM_TIMEOUT_MS")) { const long long ms = atoll(e); if (ms > 0) c->timeout_ns = (unsigned long long)ms * 1000000ull; }
    CS_CUDA_OK(cudaMalloc((void**)&c->dev, sizeof(CommDev)));
    for (int r = 0; r < world; ++r) { c->peer_arena[r] = arena[r]; c->peer_signals[r] = (uint32_t*)signals[r]; }
    int rc = upload_ctx(c, c->dev);
    if (rc) return rc;
    *out = c;
    return CS_OK;
}
size_t comm_signal_bytes() { return kSignalWords * sizeof(uint32_t); }

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
            cudaFree(g[r]->arena); cudaFree(g[r]->signals); cudaFree(g[r]->epochs); cudaFreeHost(g[r]->status_host); cudaFree(g[r]->dev);
            delete g[r];
        }
        delete[] g;
        return CS_OK;
    }
    if (c->external) {
        cudaFree(c->epochs); cudaFreeHost(c->status_host); cudaFree(c->dev);
        delete c;
        return CS_OK;
    }
    for (int r = 0; r < c->world; ++r)
        if (c->opened[r]) { cudaIpcCloseMemHandle(c->peer_arena[r]); cudaIpcCloseMemHandle(c->peer_signals[r]); }
    cudaFree(c->arena); cudaFree(c->signals); cudaFree(c->epochs); cudaFreeHost(c->status_host); cudaFree(c->dev);
    delete c;
    return CS_OK;
}

static int pick_blocks(long long n, bool oneshot) {
    if (oneshot) { long long b = (n + 4095) / 4096; return (int)(b < 1 ? 1 : (b > 16 ? 16 : b)); }
    int cap = num_sms() < kMaxBlocks ? num_sms() : kMaxBlocks;
    long long b = (n + 8191) / 8192;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

size_t agg_arena_need(long long n, int wire_bf16, int world) {
    const size_t es = wire_bf16 ? 2 : 4;
    if (n <= kOneShotMaxElems) return 2 * (size_t)world * (size_t)((n + 7) & ~7ll) * es;
    return 2 * (size_t)(((((n + 3) >> 2) * (es * 4)) + 255) & ~255ll);
}

// Launch for a real rank (ranks = 1 context) or for an emulated group (ranks = world contexts).
static int launch_agg(const CommDev* ctxs, int ranks, int world, size_t arena_bytes, const AggCall& a, cudaStream_t st, int max_blocks = 0) {
    CS_REQUIRE(a.n > 0, "aggregate: n must be positive");
    CS_REQUIRE(a.channel >= 0 && a.channel < kChannels, "aggregate: channel %d out of range", a.channel);
    CS_REQUIRE(a.arena_off % 16 == 0, "aggregate: arena offset must be 16-byte aligned");
    CS_REQUIRE(a.arena_off + agg_arena_need(a.n, a.wire_bf16, world) <= arena_bytes,
               "aggregate: arena too small (%zu B needed at offset %zu, have %zu)",
               agg_arena_need(a.n, a.wire_bf16, world), a.arena_off, arena_bytes);
    const bool oneshot = a.n <= kOneShotMaxElems;
    int nb = pick_blocks(a.n, oneshot);
    if (max_blocks > 0 && nb > max_blocks) nb = max_blocks;    // (the partition depends on nb: every rank must use the same cap)
    const void* fn;
    if (oneshot) fn = a.wire_bf16 ? (const void*)fedavg_oneshot_kernel<true> : (const void*)fedavg_oneshot_kernel<false>;
    else fn = a.wire_bf16 ? (const void*)fedavg_twoshot_kernel<true> : (const void*)fedavg_twoshot_kernel<false>;
    if (ranks > 1) {   // emulated ranks share one GPU: the cooperative launch needs the whole grid co-resident
        int per_sm = 1;
        CS_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kCommThreads, 0));
        int cap = (num_sms() * (per_sm < 1 ? 1 : per_sm)) / ranks;
        if (nb > cap) nb = cap < 1 ? 1 : cap;
    }
    dim3 grid(nb, ranks), block(kCommThreads);
    void* args[] = {(void*)&ctxs, (void*)&a};
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

// Test hook: launch ONE member of an emulated group on its own, so that its peers never arrive and the time-out path runs.
int debug_fedavg_aggregate_alone(cs_comm* member, float* local, float* prev, long long n, float weight, float* mom,
                                 cudaStream_t st) {
    CS_REQUIRE(member && member->emulated, "debug_aggregate_alone: needs a member of an emulated group");
    AggCall a; memset(&a, 0, sizeof(a));
    a.local[0] = local; a.prev[0] = prev; a.mom[0] = mom; a.weight[0] = weight;
    a.n = n; a.server_lr = 1.0f; a.post_scale = 1.0f; a.wire_bf16 = 0; a.channel = 0; a.arena_off = 0;
    return launch_agg(member->dev, 1, member->world, member->bytes, a, st);
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

static int launch_inplace(const CommDev* ctxs, int ranks, const InplaceCall& a, cudaStream_t st, int max_blocks = 0) {
    CS_REQUIRE(a.n > 0 && a.n % 4 == 0, "aggregate_inplace: n must be a positive multiple of 4");
    CS_REQUIRE(a.channel >= 0 && a.channel < kChannels, "aggregate_inplace: bad channel");
    int nb = pick_blocks(a.n, false);
    if (max_blocks > 0 && nb > max_blocks) nb = max_blocks;
    // independent 16-byte loads per thread and source rank before the first use (tuning knob, read at every call)
    const int unroll = env_knob("CS_INPLACE_UNROLL", 2);
    const void* fn = unroll == 1 ? (const void*)fedavg_inplace_kernel<1> : unroll == 3 ? (const void*)fedavg_inplace_kernel<3>
                                                                                        : (const void*)fedavg_inplace_kernel<2>;
    { const int cap = env_knob("CS_AGG_BLOCKS", 0); if (cap > 0 && nb > cap) nb = cap; }
    if (ranks > 1) {
        int per_sm = 1;
        CS_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kCommThreads, 0));
        int cap = (num_sms() * (per_sm < 1 ? 1 : per_sm)) / ranks;
        if (nb > cap) nb = cap < 1 ? 1 : cap;
    }
    dim3 grid(nb, ranks), block(kCommThreads);
    void* args[] = {(void*)&ctxs, (void*)&a};
    if (ranks > 1) CS_CUDA_OK(cudaLaunchCooperativeKernel(fn, grid, block, args, 0, st));
    else CS_CUDA_OK(cudaLaunchKernel(fn, grid, block, args, 0, st));
    return CS_OK;
}

int fedavg_aggregate_inplace(cs_comm* c, size_t param_off, float* prev, long long n, const float* weights,
                             float server_lr, float* mom, int channel, cudaStream_t st) {
    CS_REQUIRE(c && !c->emulated, "aggregate_inplace: use the *_local entry point for emulated groups");
    CS_REQUIRE(param_off % 16 == 0 && param_off + (size_t)n * 4 <= c->bytes, "aggregate_inplace: parameters must lie inside the arena");
    InplaceCall a; memset(&a, 0, sizeof(a));
    a.param_off[0] = param_off; a.prev[0] = prev; a.mom[0] = mom;
    for (int r = 0; r < c->world; ++r) a.weights[r] = weights[r];
    a.n = n; a.server_lr = server_lr; a.channel = channel;
    return launch_inplace(c->dev, 1, a, st, c->max_blocks);
}

int fedavg_aggregate_inplace_local(cs_comm** g, int world, size_t param_off, float* const* prev, long long n,
                                   const float* weights, float server_lr, float* const* mom, int channel, cudaStream_t st) {
    CS_REQUIRE(g && g[0] && g[0]->emulated && g[0]->world == world, "aggregate_inplace_local: not an emulated group");
    CS_REQUIRE(param_off % 16 == 0 && param_off + (size_t)n * 4 <= g[0]->bytes, "aggregate_inplace: parameters must lie inside the arena");
    InplaceCall a; memset(&a, 0, sizeof(a));
    for (int r = 0; r < world; ++r) {
        a.param_off[r] = param_off; a.prev[r] = prev ? prev[r] : nullptr; a.mom[r] = mom ? mom[r] : nullptr; a.weights[r] = weights[r];
    }
    a.n = n; a.server_lr = server_lr; a.channel = channel;
    return launch_inplace(g[0]->group_dev, world, a, st);
}

static int launch_bucket(const CommDev* ctxs, int ranks, const BucketCall& a, int max_blocks, cudaStream_t st) {
    CS_REQUIRE(a.n > 0 && a.n % 4 == 0, "fedavg_bucket: n must be a positive multiple of 4");
    CS_REQUIRE(a.channel >= 0 && a.channel < kChannels, "fedavg_bucket: bad channel");
    int nb = pick_blocks(a.n, false);
    if (max_blocks > 0 && nb > max_blocks) nb = max_blocks;
    const void* fn = (const void*)fedavg_bucket_kernel;
    if (ranks > 1) {
        int per_sm = 1;
        CS_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kCommThreads, 0));
        int cap = (num_sms() * (per_sm < 1 ? 1 : per_sm)) / ranks;
        if (nb > cap) nb = cap < 1 ? 1 : cap;
    }
    dim3 grid(nb, ranks), block(kCommThreads);
    void* args[] = {(void*)&ctxs, (void*)&a};
    if (ranks > 1) CS_CUDA_OK(cudaLaunchCooperativeKernel(fn, grid, block, args, 0, st));
    else CS_CUDA_OK(cudaLaunchKernel(fn, grid, block, args, 0, st));
    return CS_OK;
}

int fedavg_bucket(cs_comm* c, size_t param_off, const float* grad, float* mom, float* prev, void* shadow, long long n,
                  const float* weights, float lr, float momentum, float wd, float inv_scale, float server_lr, int do_sgd,
                  int channel, int max_blocks, cudaStream_t st) {
    CS_REQUIRE(c && !c->emulated, "fedavg_bucket: use the *_local entry point for emulated groups");
    CS_REQUIRE(param_off % 16 == 0 && param_off + (size_t)n * 4 <= c->bytes, "fedavg_bucket: the bucket must lie inside the arena");
    CS_REQUIRE(!do_sgd || (grad && mom), "fedavg_bucket: the fused optimizer step needs grad and momentum");
    BucketCall a; memset(&a, 0, sizeof(a));
    a.param_off[0] = param_off; a.grad[0] = grad; a.mom[0] = mom; a.prev[0] = prev; a.shadow[0] = shadow;
    for (int r = 0; r < c->world; ++r) a.weights[r] = weights[r];
    a.n = n; a.lr = lr; a.momentum = momentum; a.wd = wd; a.inv_scale = inv_scale; a.server_lr = server_lr; a.do_sgd = do_sgd; a.channel = channel;
    return launch_bucket(c->dev, 1, a, max_blocks, st);
}

int fedavg_bucket_local(cs_comm** g, int world, size_t param_off, const float* const* grad, float* const* mom, float* const* prev,
                        void* const* shadow, long long n, const float* weights, float lr, float momentum, float wd, float inv_scale,
                        float server_lr, int do_sgd, int channel, cudaStream_t st) {
    CS_REQUIRE(g && g[0] && g[0]->emulated && g[0]->world == world, "fedavg_bucket_local: not an emulated group");
    CS_REQUIRE(param_off % 16 == 0 && param_off + (size_t)n * 4 <= g[0]->bytes, "fedavg_bucket: the bucket must lie inside the arena");
    BucketCall a; memset(&a, 0, sizeof(a));
    for (int r = 0; r < world; ++r) {
        a.param_off[r] = param_off; a.grad[r] = grad ? grad[r] : nullptr; a.mom[r] = mom ? mom[r] : nullptr;
        a.prev[r] = prev ? prev[r] : nullptr; a.shadow[r] = shadow ? shadow[r] : nullptr; a.weights[r] = weights[r];
    }
    a.n = n; a.lr = lr; a.momentum = momentum; a.wd = wd; a.inv_scale = inv_scale; a.server_lr = server_lr; a.do_sgd = do_sgd; a.channel = channel;
    return launch_bucket(g[0]->group_dev, world, a, 0, st);
}

int fedavg_aggregate_nvls(cs_comm* c, float* local, void* mc, float* prev, long long n, float prescale, float post_scale,
                          float server_lr, float* mom, int channel, cudaStream_t st) {
    CS_REQUIRE(c && !c->emulated && c->world > 1, "aggregate_nvls: needs a real multi-GPU communicator");
    CS_REQUIRE(local && mc && n > 0 && n % 4 == 0, "aggregate_nvls: bad buffer / n must be a multiple of 4");
    CS_REQUIRE((((uintptr_t)local | (uintptr_t)mc) & 15) == 0, "aggregate_nvls: buffers must be 16-byte aligned");
    CS_REQUIRE(channel >= 0 && channel < kChannels, "aggregate_nvls: bad channel");
    NvlsCall a;
    a.local = local; a.mc = (char*)mc; a.prev = prev; a.mom = mom; a.n = n; a.prescale = prescale; a.post_scale = post_scale;
    a.server_lr = server_lr; a.channel = channel;
    const CommDev* ctxs = c->dev;
    const int unroll = env_knob("CS_NVLS_UNROLL", 4);
    // The switch, not the SMs, sets the pace of multimem traffic: 8 x B200, 96.6 MB: 0.265 ms with 148 CTAs, 0.240 with 64,
    // 0.232 with 32, 0.229 with 16 (profiles/r2/agg_sweep_n8*.json); powers of two partition the slices evenly.
    int nb = pick_blocks(n, false);
    { const int cap = env_knob("CS_AGG_BLOCKS", 32); if (cap > 0 && nb > cap) nb = cap; }
    if (unroll == 8) fedavg_nvls_kernel<8><<<nb, kCommThreads, 0, st>>>(ctxs, a);
    else if (unroll == 2) fedavg_nvls_kernel<2><<<nb, kCommThreads, 0, st>>>(ctxs, a);
    else fedavg_nvls_kernel<4><<<nb, kCommThreads, 0, st>>>(ctxs, a);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
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

// The status word lives in host-mapped pinned memory: reading it costs nothing and synchronises nothing.  It is
// meaningful for kernels the caller has already synchronised with (every round of FedAvgRunner.run_round does).
int comm_status(cs_comm* c, int* out) {
    CS_REQUIRE(c && out, "comm_status: null");
    *out = *(volatile int*)c->status_host;
    return CS_OK;
}
int comm_set_timeout_ms(cs_comm* c, long long ms) {
    CS_REQUIRE(c && ms > 0, "comm_set_timeout: bad arguments");
    cs_comm** members = c->emulated ? c->group : &c;
    const int n = c->emulated ? c->world : 1;
    for (int r = 0; r < n; ++r) {
        members[r]->timeout_ns = (unsigned long long)ms * 1000000ull;
        const unsigned long long v = members[r]->timeout_ns;
        CS_CUDA_OK(cudaMemcpy(&members[r]->dev->timeout_ns, &v, sizeof(v), cudaMemcpyHostToDevice));
        if (c->emulated) CS_CUDA_OK(cudaMemcpy(&c->group_dev[r].timeout_ns, &v, sizeof(v), cudaMemcpyHostToDevice));
    }
    return CS_OK;
}
int comm_set_max_blocks(cs_comm* c, int n) {
    CS_REQUIRE(c != nullptr && n >= 0 && n <= kMaxBlocks, "comm_set_max_blocks: bad arguments");
    c->max_blocks = n;
    return CS_OK;
}
int comm_clear_status(cs_comm* c) {
    CS_REQUIRE(c != nullptr, "comm_clear_status: null");
    *(volatile int*)c->status_host = 0;
    return CS_OK;
}

}  // namespace cs

namespace cs {
void* comm_arena(cs_comm* c, int r) { return c ? (r < 0 ? c->arena : c->peer_arena[r]) : nullptr; }
size_t comm_bytes(cs_comm* c) { return c ? c->bytes : 0; }
}  // namespace cs
