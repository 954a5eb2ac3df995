// Memory-bound helper kernels: batch gather (K11), cross-entropy + loss scale (K5), flat SGD with
// fused unscale / inf-check / scaler update (K6), z-score staging (K10), BN / ReLU / residual /
// pooling (K4) and layout converters.  All are CUDA-core kernels sized against HBM bandwidth:
// 128-bit accesses, grid-stride loops, one pass over the data where the math allows it.
// Reference call sites they replace: x_gpu[sel] (Module_3/shard_dataset.py:135-136), F.cross_entropy + GradScaler
// (Module_3/TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:121-127), torch.optim.SGD(lr=1e-2, momentum=0.9) (:181), the
// per-window z-score of the LABL loader (Module_1/labl_loader(EXPERIMENTAL).py:65-69); BN / pooling have no counterpart.
#include "common/cs_common.cuh"
#include <stdlib.h>

namespace cs {

// ------------------------------------------------------------------------------------------------
// K11 gather: one warp per row, 128-bit loads when L % 4 == 0
// ------------------------------------------------------------------------------------------------
__global__ void gather_rows_kernel(const float* __restrict__ x, const int64_t* __restrict__ y,
                                   const int64_t* __restrict__ idx, void* __restrict__ xo,
                                   int64_t* __restrict__ yo, int B, int L, int out_bf16) {
    const int warps_per_block = blockDim.x >> 5;
    const int lane = threadIdx.x & 31;
    for (int b = blockIdx.x * warps_per_block + (threadIdx.x >> 5); b < B; b += gridDim.x * warps_per_block) {
        const int64_t r = idx[b];
        const float* src = x + r * (int64_t)L;
        if (lane == 0 && yo != nullptr) yo[b] = y[r];
        if (!out_bf16) {
            float* dst = (float*)xo + (int64_t)b * L;
            if ((L & 3) == 0) {
                const float4* s4 = reinterpret_cast<const float4*>(src);
                float4* d4 = reinterpret_cast<float4*>(dst);
                for (int i = lane; i < (L >> 2); i += 32) d4[i] = __ldg(s4 + i);
            } else {
                for (int i = lane; i < L; i += 32) dst[i] = __ldg(src + i);
            }
        } else {
            __nv_bfloat16* dst = (__nv_bfloat16*)xo + (int64_t)b * L;
            for (int i = lane; i < L; i += 32) dst[i] = __float2bfloat16(__ldg(src + i));
        }
    }
}

int gather_rows(const float* x, const int64_t* y, const int64_t* idx, void* xo, int64_t* yo, int B,
                int L, int out_bf16, cudaStream_t st) {
    CS_REQUIRE(B > 0 && L > 0, "gather: bad shape B=%d L=%d", B, L);
    int blocks = (B + 7) / 8;
    int cap = num_sms() * 8;
    if (blocks > cap) blocks = cap;
    gather_rows_kernel<<<blocks, 256, 0, st>>>(x, y, idx, xo, yo, B, L, out_bf16);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// ------------------------------------------------------------------------------------------------
// K5 cross entropy (mean) forward + backward with the loss scale folded into dlogits
// ------------------------------------------------------------------------------------------------
__global__ void ce_fwd_bwd_kernel(const float* __restrict__ logits, const int64_t* __restrict__ y,
                                  int B, int C, const float* __restrict__ scale_ptr,
                                  float* __restrict__ loss_out, float* __restrict__ dlogits) {
    // single CTA; thread per sample, grid-stride over B; block reduce of the loss
    __shared__ float red[32];
    const float scale = scale_ptr ? *scale_ptr : 1.0f;
    const float invB = 1.0f / (float)B;
    float lsum = 0.0f;
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
        const float* lg = logits + (size_t)b * C;
        float mx = -3.0e38f;
        for (int c = 0; c < C; ++c) mx = fmaxf(mx, lg[c]);
        float se = 0.0f;
        for (int c = 0; c < C; ++c) se += __expf(lg[c] - mx);
        const float lse = mx + __logf(se);
        const int lab = (int)y[b];
        lsum += (lse - lg[lab]) * invB;
        if (dlogits) {
            for (int c = 0; c < C; ++c)
                dlogits[(size_t)b * C + c] = (__expf(lg[c] - lse) - (c == lab ? 1.0f : 0.0f)) * invB * scale;
        }
    }
    lsum = warp_sum(lsum);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = lsum;
    __syncthreads();
    if (threadIdx.x < 32) {
        float v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0f;
        v = warp_sum(v);
        if (threadIdx.x == 0) loss_out[0] = v;
    }
}

int cross_entropy(const float* logits, const int64_t* y, int B, int C, const float* scale_ptr,
                  float* loss_out, float* dlogits, cudaStream_t st) {
    CS_REQUIRE(B > 0 && C > 0 && C <= 64, "cross_entropy: bad shape B=%d C=%d", B, C);
    int threads = B >= 1024 ? 1024 : ((B + 31) / 32) * 32;
    ce_fwd_bwd_kernel<<<1, threads, 0, st>>>(logits, y, B, C, scale_ptr, loss_out, dlogits);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// ------------------------------------------------------------------------------------------------
// K6 flat SGD-momentum, found-inf aware
// ------------------------------------------------------------------------------------------------
__global__ void inf_check_kernel(const float* __restrict__ g, int64_t n, int32_t* __restrict__ ctrl) {
    int bad = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        if (!finite_f(g[i])) bad = 1;
    bad = __syncthreads_or(bad);
    if (bad && threadIdx.x == 0) atomicExch(ctrl + CTRL_FOUND_INF, 1);
}

__global__ void sgd_momentum_kernel(float* __restrict__ p, float* __restrict__ mom,
                                    const float* __restrict__ g, int64_t n, float lr, float momentum,
                                    float wd, int use_scaler, const int32_t* __restrict__ ctrl,
                                    const float* __restrict__ fstate,
                                    __nv_bfloat16* __restrict__ shadow) {
    const int bad = use_scaler ? ctrl[CTRL_FOUND_INF] : 0;
    if (bad) return;                                   // skipped step: nothing is written
    const float inv = use_scaler ? 1.0f / fstate[FS_SCALE] : 1.0f;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t n4 = n >> 2;
    float4* p4 = reinterpret_cast<float4*>(p);
    float4* m4 = reinterpret_cast<float4*>(mom);
    const float4* g4 = reinterpret_cast<const float4*>(g);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 pv = p4[i], mv = m4[i], gv = g4[i];
        mv.x = momentum * mv.x + (gv.x * inv + wd * pv.x); pv.x -= lr * mv.x;
        mv.y = momentum * mv.y + (gv.y * inv + wd * pv.y); pv.y -= lr * mv.y;
        mv.z = momentum * mv.z + (gv.z * inv + wd * pv.z); pv.z -= lr * mv.z;
        mv.w = momentum * mv.w + (gv.w * inv + wd * pv.w); pv.w -= lr * mv.w;
        p4[i] = pv; m4[i] = mv;
        if (shadow) {
            __nv_bfloat162* s2 = reinterpret_cast<__nv_bfloat162*>(shadow) + 2 * i;
            s2[0] = __floats2bfloat162_rn(pv.x, pv.y);
            s2[1] = __floats2bfloat162_rn(pv.z, pv.w);
        }
    }
    for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float pv = p[i];
        float mv = momentum * mom[i] + (g[i] * inv + wd * pv);
        pv -= lr * mv;
        mom[i] = mv; p[i] = pv;
        if (shadow) shadow[i] = __float2bfloat16(pv);
    }
}

__global__ void scaler_update_kernel(int32_t* ctrl, float* fstate, float growth, float backoff, int interval) {
    const int bad = ctrl[CTRL_FOUND_INF];
    const float scale = fstate[FS_SCALE];
    ctrl[CTRL_STEPS] += 1;
    if (bad) { fstate[FS_SCALE] = scale * backoff; ctrl[CTRL_GROWTH] = 0; ctrl[CTRL_SKIPPED] += 1; }
    else {
        int t = ctrl[CTRL_GROWTH] + 1;
        if (t >= interval) { fstate[FS_SCALE] = scale * growth; t = 0; }
        ctrl[CTRL_GROWTH] = t;
    }
    ctrl[CTRL_FOUND_INF] = 0;   // cleared for the next step's inf_check (the update kernel ran before us)
}

int sgd_momentum(float* p, float* mom, const float* g, int64_t n, float lr, float momentum, float wd,
                 int use_scaler, int32_t* ctrl, float* fstate, float growth, float backoff,
                 int interval, void* shadow, cudaStream_t st) {
    CS_REQUIRE(n > 0, "sgd: n must be positive");
    CS_REQUIRE(!use_scaler || (ctrl && fstate), "sgd: scaler state missing");
    CS_REQUIRE((((uintptr_t)p | (uintptr_t)mom | (uintptr_t)g) & 15) == 0, "sgd: buffers must be 16-byte aligned");
    int64_t want = (n / 4 + 255) / 256;
    int blocks = (int)(want < 1 ? 1 : (want > (int64_t)num_sms() * 8 ? (int64_t)num_sms() * 8 : want));
    if (use_scaler) {
        inf_check_kernel<<<blocks, 256, 0, st>>>(g, n, ctrl);
        CS_CUDA_OK(cudaGetLastError());
    }
    sgd_momentum_kernel<<<blocks, 256, 0, st>>>(p, mom, g, n, lr, momentum, wd, use_scaler, ctrl, fstate,
                                                (__nv_bfloat16*)shadow);
    CS_CUDA_OK(cudaGetLastError());
    if (use_scaler) {
        scaler_update_kernel<<<1, 1, 0, st>>>(ctrl, fstate, growth, backoff, interval);
        CS_CUDA_OK(cudaGetLastError());
    }
    return CS_OK;
}

// ------------------------------------------------------------------------------------------------
// K10 staging helpers
// ------------------------------------------------------------------------------------------------
__global__ void zscore_rows_kernel(const float* __restrict__ x, void* __restrict__ y, int B, int L, int out_bf16) {
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    for (int b = blockIdx.x * wpb + (threadIdx.x >> 5); b < B; b += gridDim.x * wpb) {
        const float* src = x + (size_t)b * L;
        float s = 0.0f, ss = 0.0f;
        for (int i = lane; i < L; i += 32) { float v = src[i]; s += v; ss = fmaf(v, v, ss); }
        s = warp_sum(s); ss = warp_sum(ss);
        const float mean = s / L;
        const float var = fmaxf(ss / L - mean * mean, 0.0f);
        const float inv = 1.0f / (sqrtf(var) + 1e-8f);
        if (out_bf16) {
            __nv_bfloat16* dst = (__nv_bfloat16*)y + (size_t)b * L;
            for (int i = lane; i < L; i += 32) dst[i] = __float2bfloat16((src[i] - mean) * inv);
        } else {
            float* dst = (float*)y + (size_t)b * L;
            for (int i = lane; i < L; i += 32) dst[i] = (src[i] - mean) * inv;
        }
    }
}

// The reference's native kernel on the GPU, same contract: single-channel VALID cross-correlation, fp32,
// y[b, i] = sum_k x[b, i + k] * w[k], outL = L - K + 1 (ref: Module_2/conv1d_openmp_simd.c:21-61).  One CTA = 1024 outputs
// of one row: the window (+ K - 1 halo) and the taps are staged in shared memory, each thread produces 4 outputs.
__global__ void __launch_bounds__(256) conv1d_valid_f32_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                               float* __restrict__ y, int L, int K, int outL) {
    extern __shared__ float cv[];
    float* ws = cv;                        // [K]
    float* xs = cv + ((K + 3) & ~3);       // [1024 + K - 1]
    const int b = blockIdx.y, i0 = blockIdx.x * 1024;
    const float* xr = x + (size_t)b * L;
    for (int k = threadIdx.x; k < K; k += 256) ws[k] = w[k];
    for (int i = threadIdx.x; i < 1024 + K - 1; i += 256) xs[i] = (i0 + i < L) ? __ldg(xr + i0 + i) : 0.0f;
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int i = threadIdx.x + r * 256;
        if (i0 + i < outL) {
            float a = 0.0f;
            for (int k = 0; k < K; ++k) a = fmaf(xs[i + k], ws[k], a);
            y[(size_t)b * outL + i0 + i] = a;
        }
    }
}

int conv1d_valid_f32(const float* x, const float* w, float* y, int B, int L, int K, cudaStream_t st) {
    CS_REQUIRE(B >= 1 && K >= 1 && K <= 1024 && L >= K, "conv1d_valid: bad shape B=%d L=%d K=%d", B, L, K);
    const int outL = L - K + 1;
    const size_t smem = (((size_t)K + 3) / 4 * 4 + 1024 + K) * sizeof(float);
    conv1d_valid_f32_kernel<<<dim3((outL + 1023) / 1024, B), 256, smem, st>>>(x, w, y, L, K, outL);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

int zscore_rows(const float* x, void* y, int B, int L, int out_bf16, cudaStream_t st) {
    CS_REQUIRE(B > 0 && L > 0, "zscore: bad shape");
    int blocks = (B + 7) / 8; int cap = num_sms() * 8; if (blocks > cap) blocks = cap;
    zscore_rows_kernel<<<blocks, 256, 0, st>>>(x, y, B, L, out_bf16);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// ------------------------------------------------------------------------------------------------
// K4 BN / ReLU / residual / pool on channels-last bf16 [rows, C], C % 8 == 0.
// Every thread moves 16 bytes (8 channels) per access.  Reductions: a thread owns one 8-channel group and
// strides over rows with 4 independent loads in flight, partial sums meet in shared memory, and each CTA
// finishes with ONE atomic per channel and quantity (<= 148 CTAs, so the atomic tail stays short).
// ------------------------------------------------------------------------------------------------
struct bf16x8 { uint4 v; };
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 t = __bfloat1622float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return u;
}

constexpr int kRedThreads = 256;

// block-level finish of a column reduction: acc[Q][8] per thread -> atomics on out[q*C + channel]
// (called by ALL threads of the block; threads with rl >= rpb hold no partials)
template <int Q>
__device__ __forceinline__ void column_reduce_finish(float (&acc)[Q][8], float* const (&out)[Q], int C, int g, int rl, int rpb,
                                                     float* sm /* [rpb][Q][C] */) {
    if (rl < rpb) {
#pragma unroll
        for (int q = 0; q < Q; ++q)
#pragma unroll
            for (int j = 0; j < 8; ++j) sm[((size_t)rl * Q + q) * C + g * 8 + j] = acc[q][j];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < Q * C; i += blockDim.x) {
        float s = 0.0f;
        for (int r = 0; r < rpb; ++r) s += sm[(size_t)r * Q * C + i];
        atomicAdd(out[i / C] + (i % C), s);
    }
}

// Optional tail of the two column reductions below, run by the LAST block to arrive (ticket): the per-channel totals
// are complete in `acc` (L2 atomics), so that block turns them into the layer's outputs and then CLEARS acc and the
// ticket — the accumulator is a persistent, self-cleaning workspace and no fill / finalize / add launches are needed.
// (BnStatsTail and bn_stats_tail_channels live in common/cs_common.cuh: the conv epilogue shares them)
struct BnBwdTail {
    int32_t* ticket;
    float* dgamma; float* dbeta;          // this batch's reductions
    float* gamma_grad; float* beta_grad;  // optional: += into the parameter gradient arena
    const float* gamma; float* coef;      // optional: dx = a*d + p*x + q per channel, coef = [a | p | q] (3C)
    float inv_rows;
};

__device__ __forceinline__ bool last_block_arrives(int32_t* ticket) {
    __shared__ int is_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = (atomicAdd(ticket, 1) == (int)gridDim.x - 1);
    __syncthreads();
    if (is_last) __threadfence();
    return is_last != 0;
}

__global__ void __launch_bounds__(kRedThreads) bn_stats_kernel(const uint4* __restrict__ x, float* __restrict__ stats, int64_t rows, int C,
                                                               BnStatsTail t) {
    extern __shared__ float sm_red[];
    const int G = C >> 3, rpb = kRedThreads / G;
    const int g = threadIdx.x % G, rl = threadIdx.x / G;
    float acc[2][8] = {};
    if (rl < rpb) {
        const int64_t stride = (int64_t)gridDim.x * rpb;
        for (int64_t r = (int64_t)blockIdx.x * rpb + rl; r < rows; r += 4 * stride) {
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { const int64_t rr = r + u * stride; if (rr < rows) v[u] = __ldg(x + rr * G + g); }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (r + u * stride < rows) {
                    float f[8]; unpack8(v[u], f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) { acc[0][j] += f[j]; acc[1][j] = fmaf(f[j], f[j], acc[1][j]); }
                }
            }
        }
    }
    float* const out[2] = {stats, stats + C};
    column_reduce_finish<2>(acc, out, C, g, rl, rpb, sm_red);
    if (t.ticket == nullptr || !last_block_arrives(t.ticket)) return;
    bn_stats_tail_channels(stats, C, t, threadIdx.x, blockDim.x);
    if (threadIdx.x == 0) *t.ticket = 0;
}

__global__ void bn_apply_kernel(const uint4* __restrict__ x, const uint4* __restrict__ res, uint4* __restrict__ y,
                                const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
                                const float* __restrict__ beta, int64_t rows, int C, int relu) {
    const int G = C >> 3;
    const int64_t total = rows * G;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % G) * 8;
        float f[8], r[8];
        unpack8(__ldg(x + i), f);
        if (res) unpack8(__ldg(res + i), r);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float o = (f[j] - mean[c + j]) * (rstd[c + j] * gamma[c + j]) + beta[c + j];
            if (res) o += r[j];
            f[j] = relu ? fmaxf(o, 0.0f) : o;
        }
        y[i] = pack8(f);
    }
}

__global__ void __launch_bounds__(kRedThreads) bn_bwd_reduce_kernel(const uint4* __restrict__ x, const uint4* __restrict__ y,
                                                                    const uint4* __restrict__ dy, const float* __restrict__ mean,
                                                                    const float* __restrict__ rstd, float* __restrict__ dgamma,
                                                                    float* __restrict__ dbeta, int64_t rows, int C, int relu, BnBwdTail t) {
    extern __shared__ float sm_red[];
    const int G = C >> 3, rpb = kRedThreads / G;
    const int g = threadIdx.x % G, rl = threadIdx.x / G;
    float acc[2][8] = {};
    if (rl < rpb) {
        float m[8], rs[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { m[j] = mean[g * 8 + j]; rs[j] = rstd[g * 8 + j]; }
        const int64_t stride = (int64_t)gridDim.x * rpb;
        for (int64_t r = (int64_t)blockIdx.x * rpb + rl; r < rows; r += 2 * stride) {
            uint4 vx[2], vy[2], vd[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int64_t rr = r + u * stride;
                if (rr < rows) { vx[u] = __ldg(x + rr * G + g); vd[u] = __ldg(dy + rr * G + g); if (relu) vy[u] = __ldg(y + rr * G + g); }
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                if (r + u * stride < rows) {
                    float fx[8], fd[8], fy[8];
                    unpack8(vx[u], fx); unpack8(vd[u], fd);
                    if (relu) unpack8(vy[u], fy);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float d = (relu && fy[j] <= 0.0f) ? 0.0f : fd[j];
                        acc[0][j] = fmaf(d, (fx[j] - m[j]) * rs[j], acc[0][j]);
                        acc[1][j] += d;
                    }
                }
            }
        }
    }
    float* const out[2] = {dgamma, dbeta};
    column_reduce_finish<2>(acc, out, C, g, rl, rpb, sm_red);
    if (t.ticket == nullptr || !last_block_arrives(t.ticket)) return;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {      // here dgamma / dbeta are the persistent accumulators
        const float dg = __ldcg(dgamma + c), db = __ldcg(dbeta + c);
        const float gg = t.gamma_grad ? t.gamma_grad[c] : 0.0f, bg = t.gamma_grad ? t.beta_grad[c] : 0.0f;
        const float ga = t.coef ? t.gamma[c] : 0.0f, rs = rstd[c], mu = mean[c];
        t.dgamma[c] = dg; t.dbeta[c] = db;
        if (t.gamma_grad) { t.gamma_grad[c] = gg + dg; t.beta_grad[c] = bg + db; }
        if (t.coef) {
            // dx = gamma*rstd*(d - (dbeta + xhat*dgamma)/n), xhat = (x - mean)*rstd   ->   a*d + p*x + q
            const float a = ga * rs;
            const float pc = -a * t.inv_rows * dg * rs;
            t.coef[c] = a;
            t.coef[C + c] = pc;
            t.coef[2 * C + c] = -a * t.inv_rows * db - pc * mu;
        }
        dgamma[c] = 0.0f; dbeta[c] = 0.0f;
    }
    if (threadIdx.x == 0) *t.ticket = 0;
}

__global__ void bn_bwd_apply_kernel(const uint4* __restrict__ x, const uint4* __restrict__ y, const uint4* __restrict__ dy,
                                    const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
                                    const float* __restrict__ dgamma, const float* __restrict__ dbeta, uint4* __restrict__ dx,
                                    uint4* __restrict__ dres, int64_t rows, int C, int relu) {
    const int G = C >> 3;
    const int64_t total = rows * G;
    const float invn = 1.0f / (float)rows;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % G) * 8;
        float fx[8], fd[8], fy[8];
        unpack8(__ldg(x + i), fx); unpack8(__ldg(dy + i), fd);
        if (relu) unpack8(__ldg(y + i), fy);
#pragma unroll
        for (int j = 0; j < 8; ++j) if (relu && fy[j] <= 0.0f) fd[j] = 0.0f;
        if (dres) dres[i] = pack8(fd);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float xh = (fx[j] - mean[c + j]) * rstd[c + j];
            fx[j] = gamma[c + j] * rstd[c + j] * (fd[j] - invn * (dbeta[c + j] + xh * dgamma[c + j]));
        }
        dx[i] = pack8(fx);
    }
}

// Coefficient forms of the two elementwise passes: a thread owns one 8-channel group, loads its per-channel
// coefficients ONCE (vector loads) and then streams rows with four independent 16-byte loads in flight.
__device__ __forceinline__ void load8(const float* p, float (&f)[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

__global__ void __launch_bounds__(kRedThreads) bn_apply_coef_kernel(const uint4* __restrict__ x, const uint4* __restrict__ res,
                                                                    uint4* __restrict__ y, const float* __restrict__ coef,
                                                                    int64_t rows, int C, int relu) {
    const int G = C >> 3, rpb = kRedThreads / G;
    const int g = threadIdx.x % G, rl = threadIdx.x / G;
    if (rl >= rpb) return;
    float sc[8], sh[8];
    load8(coef + g * 8, sc); load8(coef + C + g * 8, sh);
    const int64_t stride = (int64_t)gridDim.x * rpb;
    for (int64_t r = (int64_t)blockIdx.x * rpb + rl; r < rows; r += 4 * stride) {
        uint4 v[4], w[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t rr = r + u * stride;
            if (rr < rows) { v[u] = __ldg(x + rr * G + g); if (res) w[u] = __ldg(res + rr * G + g); }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t rr = r + u * stride;
            if (rr < rows) {
                float f[8], q[8];
                unpack8(v[u], f);
                if (res) unpack8(w[u], q);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float o = fmaf(f[j], sc[j], sh[j]);
                    if (res) o += q[j];
                    f[j] = relu ? fmaxf(o, 0.0f) : o;
                }
                y[rr * G + g] = pack8(f);
            }
        }
    }
}

__global__ void __launch_bounds__(kRedThreads) bn_bwd_apply_coef_kernel(const uint4* __restrict__ x, const uint4* __restrict__ y,
                                                                        const uint4* __restrict__ dy, const float* __restrict__ coef,
                                                                        uint4* __restrict__ dx, uint4* __restrict__ dres,
                                                                        int64_t rows, int C, int relu) {
    const int G = C >> 3, rpb = kRedThreads / G;
    const int g = threadIdx.x % G, rl = threadIdx.x / G;
    if (rl >= rpb) return;
    float a[8], p[8], q[8];
    load8(coef + g * 8, a); load8(coef + C + g * 8, p); load8(coef + 2 * C + g * 8, q);
    const int64_t stride = (int64_t)gridDim.x * rpb;
    for (int64_t r = (int64_t)blockIdx.x * rpb + rl; r < rows; r += 2 * stride) {
        uint4 vx[2], vy[2], vd[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int64_t rr = r + u * stride;
            if (rr < rows) { vx[u] = __ldg(x + rr * G + g); vd[u] = __ldg(dy + rr * G + g); if (relu) vy[u] = __ldg(y + rr * G + g); }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int64_t rr = r + u * stride;
            if (rr < rows) {
                float fx[8], fd[8], fy[8];
                unpack8(vx[u], fx); unpack8(vd[u], fd);
                if (relu) {
                    unpack8(vy[u], fy);
#pragma unroll
                    for (int j = 0; j < 8; ++j) if (fy[j] <= 0.0f) fd[j] = 0.0f;
                }
                if (dres) dres[rr * G + g] = pack8(fd);
#pragma unroll
                for (int j = 0; j < 8; ++j) fx[j] = fmaf(a[j], fd[j], fmaf(p[j], fx[j], q[j]));
                dx[rr * G + g] = pack8(fx);
            }
        }
    }
}

static int stream_grid(int64_t rows, int rpb, int unroll) {
    int64_t want = (rows + (int64_t)rpb * unroll - 1) / ((int64_t)rpb * unroll);
    const int64_t cap = (int64_t)num_sms() * 8;
    if (want > cap) want = cap;
    return (int)(want < 1 ? 1 : want);
}

int bn_apply_coef(const void* x, const void* res, void* y, const float* coef, int64_t rows, int C, int relu, cudaStream_t st) {
    CS_REQUIRE(C >= 8 && C % 8 == 0 && C <= 2048, "bn_apply: C=%d must be a multiple of 8 and <= 2048", C);
    bn_apply_coef_kernel<<<stream_grid(rows, kRedThreads / (C / 8), 4), kRedThreads, 0, st>>>((const uint4*)x, (const uint4*)res,
                                                                                             (uint4*)y, coef, rows, C, relu);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

int bn_bwd_apply_coef(const void* x, const void* y, const void* dy, const float* coef, void* dx, void* dres, int64_t rows, int C,
                      int relu, cudaStream_t st) {
    CS_REQUIRE(C >= 8 && C % 8 == 0 && C <= 2048, "bn_bwd: C=%d must be a multiple of 8 and <= 2048", C);
    bn_bwd_apply_coef_kernel<<<stream_grid(rows, kRedThreads / (C / 8), 2), kRedThreads, 0, st>>>(
        (const uint4*)x, (const uint4*)y, (const uint4*)dy, coef, (uint4*)dx, (uint4*)dres, rows, C, relu);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

static int env_red_ctas_per_sm() {
    static int v = 0;
    if (!v) { const char* e = getenv("CS_BN_RED_CTAS"); v = (e && *e) ? atoi(e) : 2; if (v < 1) v = 1; if (v > 8) v = 8; }
    return v;
}

static int red_grid(int64_t rows, int rpb, int unroll) {
    int64_t want = (rows + (int64_t)rpb * unroll * 2 - 1) / ((int64_t)rpb * unroll * 2);   // >= 2 passes per thread
    const int64_t cap = (int64_t)num_sms() * env_red_ctas_per_sm();
    if (want > cap) want = cap;
    return (int)(want < 1 ? 1 : want);
}

static int ew_grid(int64_t total) {
    int64_t want = (total + 255) / 256; int64_t cap = (int64_t)num_sms() * 16;
    return (int)(want < 1 ? 1 : (want > cap ? cap : want));
}

int bn_stats(const void* x, float* stats, int64_t rows, int C, cudaStream_t st) {
    CS_REQUIRE(C >= 8 && C % 8 == 0 && C <= 2048, "bn_stats: C=%d must be a multiple of 8 and <= 2048", C);
    const int rpb = kRedThreads / (C / 8);
    const size_t smem = (size_t)rpb * 2 * C * sizeof(float);
    bn_stats_kernel<<<red_grid(rows, rpb, 4), kRedThreads, smem, st>>>((const uint4*)x, stats, rows, C, BnStatsTail{});
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// statistics + finalize in ONE launch.  `acc` [2C] fp32 and `ticket` [1] int32 form a persistent workspace that must
// be zero before the first call; the kernel leaves both zero again.
int bn_stats_finalize(const void* x, float* acc, int32_t* ticket, float* mean, float* rstd, float* rm, float* rv, int64_t rows, int C,
                      float eps, float momentum, const float* gamma, const float* beta, float* coef, cudaStream_t st) {
    CS_REQUIRE(C >= 8 && C % 8 == 0 && C <= 2048, "bn_stats: C=%d must be a multiple of 8 and <= 2048", C);
    CS_REQUIRE(acc && ticket && mean && rstd, "bn_stats_finalize: null workspace / output");
    const int rpb = kRedThreads / (C / 8);
    const size_t smem = (size_t)rpb * 2 * C * sizeof(float);
    CS_REQUIRE(coef == nullptr || (gamma && beta), "bn_stats_finalize: coef needs gamma and beta");
    BnStatsTail t{ticket, mean, rstd, rm, rv, (float)rows, eps, momentum, gamma, beta, coef};
    bn_stats_kernel<<<red_grid(rows, rpb, 4), kRedThreads, smem, st>>>((const uint4*)x, acc, rows, C, t);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

int bn_apply(const void* x, const void* res, void* y, const float* mean, const float* rstd,
             const float* gamma, const float* beta, int64_t rows, int C, int relu, cudaStream_t st) {
    CS_REQUIRE(C >= 8 && C % 8 == 0, "bn_apply: C must be a multiple of 8");
    bn_apply_kernel<<<ew_grid(rows * (C / 8)), 256, 0, st>>>((const uint4*)x, (const uint4*)res, (uint4*)y, mean, rstd, gamma, beta,
                                                              rows, C, relu);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

int bn_bwd_reduce(const void* x, const void* y, const void* dy, const float* mean, const float* rstd,
                  float* dgamma, float* dbeta, int64_t rows, int C, int relu, cudaStream_t st) {
    CS_REQUIRE(C >= 8 && C % 8 == 0 && C <= 2048, "bn_bwd: C=%d must be a multiple of 8 and <= 2048", C);
    const int rpb = kRedThreads / (C / 8);
    const size_t smem = (size_t)rpb * 2 * C * sizeof(float);
    bn_bwd_reduce_kernel<<<red_grid(rows, rpb, 2), kRedThreads, smem, st>>>((const uint4*)x, (const uint4*)y, (const uint4*)dy, mean,
                                                                            rstd, dgamma, dbeta, rows, C, relu, BnBwdTail{});
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// same reduction through the persistent workspace (`acc` [2C] + `ticket`, zero before the first call, left zero):
// dgamma / dbeta need no zero fill, and with gamma_grad / beta_grad the parameter gradients are accumulated here too
int bn_bwd_reduce_ws(const void* x, const void* y, const void* dy, const float* mean, const float* rstd, float* acc, int32_t* ticket,
                     float* dgamma, float* dbeta, float* gamma_grad, float* beta_grad, const float* gamma, float* coef, int64_t rows,
                     int C, int relu, cudaStream_t st) {
    CS_REQUIRE(C >= 8 && C % 8 == 0 && C <= 2048, "bn_bwd: C=%d must be a multiple of 8 and <= 2048", C);
    CS_REQUIRE(acc && ticket && dgamma && dbeta, "bn_bwd_reduce_ws: null workspace / output");
    CS_REQUIRE((gamma_grad == nullptr) == (beta_grad == nullptr), "bn_bwd_reduce_ws: give both gradient pointers or neither");
    const int rpb = kRedThreads / (C / 8);
    const size_t smem = (size_t)rpb * 2 * C * sizeof(float);
    CS_REQUIRE(coef == nullptr || gamma, "bn_bwd_reduce_ws: coef needs gamma");
    BnBwdTail t{ticket, dgamma, dbeta, gamma_grad, beta_grad, gamma, coef, 1.0f / (float)rows};
    bn_bwd_reduce_kernel<<<red_grid(rows, rpb, 2), kRedThreads, smem, st>>>((const uint4*)x, (const uint4*)y, (const uint4*)dy, mean,
                                                                            rstd, acc, acc + C, rows, C, relu, t);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

int bn_bwd_apply(const void* x, const void* y, const void* dy, const float* mean, const float* rstd,
                 const float* gamma, const float* dgamma, const float* dbeta, void* dx, void* dres,
                 int64_t rows, int C, int relu, cudaStream_t st) {
    CS_REQUIRE(C >= 8 && C % 8 == 0, "bn_bwd: C must be a multiple of 8");
    bn_bwd_apply_kernel<<<ew_grid(rows * (C / 8)), 256, 0, st>>>((const uint4*)x, (const uint4*)y, (const uint4*)dy, mean, rstd, gamma,
                                                                  dgamma, dbeta, (uint4*)dx, (uint4*)dres, rows, C, relu);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// mean / rstd from the accumulated sums, plus the running-statistics update of nn.BatchNorm1d (momentum,
// unbiased variance) — one tiny kernel instead of ~8 framework launches per layer
__global__ void bn_finalize_kernel(const float* __restrict__ stats, float* __restrict__ mean, float* __restrict__ rstd,
                                   float* __restrict__ running_mean, float* __restrict__ running_var, int C, float rows,
                                   float eps, float momentum) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float m = stats[c] / rows;
    const float var = fmaxf(stats[C + c] / rows - m * m, 0.0f);
    mean[c] = m;
    rstd[c] = rsqrtf(var + eps);
    if (running_mean) {
        running_mean[c] += momentum * (m - running_mean[c]);
        const float unbiased = var * (rows / fmaxf(rows - 1.0f, 1.0f));
        running_var[c] += momentum * (unbiased - running_var[c]);
    }
}
int bn_finalize(const float* stats, float* mean, float* rstd, float* rm, float* rv, int C, int64_t rows, float eps, float momentum,
                cudaStream_t st) {
    bn_finalize_kernel<<<(C + 127) / 128, 128, 0, st>>>(stats, mean, rstd, rm, rv, C, (float)rows, eps, momentum);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// 2:1 average pooling along L of a channels-last bf16 tensor (odd tail dropped), and its backward
__global__ void avgpool2_fwd_kernel(const __nv_bfloat162* __restrict__ x, __nv_bfloat162* __restrict__ y, int B, int L, int C) {
    const int C2 = C >> 1, Lo = L >> 1;
    const int64_t total = (int64_t)B * Lo * C2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int cp = (int)(i % C2); const int64_t r = i / C2; const int lo = (int)(r % Lo); const int64_t b = r / Lo;
        const __nv_bfloat162* src = x + ((b * L + 2 * lo) * C2 + cp);
        const float2 a = __bfloat1622float2(src[0]), c = __bfloat1622float2(src[C2]);
        y[i] = __floats2bfloat162_rn(0.5f * (a.x + c.x), 0.5f * (a.y + c.y));
    }
}
__global__ void avgpool2_bwd_kernel(const __nv_bfloat162* __restrict__ dy, __nv_bfloat162* __restrict__ dx, int B, int L, int C) {
    const int C2 = C >> 1, Lo = L >> 1;
    const int64_t total = (int64_t)B * L * C2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int cp = (int)(i % C2); const int64_t r = i / C2; const int l = (int)(r % L); const int64_t b = r / L;
        float2 g = make_float2(0.f, 0.f);
        if ((l >> 1) < Lo) { g = __bfloat1622float2(dy[(b * Lo + (l >> 1)) * C2 + cp]); g.x *= 0.5f; g.y *= 0.5f; }
        dx[i] = __floats2bfloat162_rn(g.x, g.y);
    }
}
int avgpool2_fwd(const void* x, void* y, int B, int L, int C, cudaStream_t st) {
    CS_REQUIRE(C % 2 == 0 && L >= 2, "avgpool2: C must be even and L >= 2");
    avgpool2_fwd_kernel<<<ew_grid((int64_t)B * (L / 2) * (C / 2)), 256, 0, st>>>((const __nv_bfloat162*)x, (__nv_bfloat162*)y, B, L, C);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}
int avgpool2_bwd(const void* dy, void* dx, int B, int L, int C, cudaStream_t st) {
    CS_REQUIRE(C % 2 == 0 && L >= 2, "avgpool2: C must be even and L >= 2");
    avgpool2_bwd_kernel<<<ew_grid((int64_t)B * L * (C / 2)), 256, 0, st>>>((const __nv_bfloat162*)dy, (__nv_bfloat162*)dx, B, L, C);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// global average pool: block per (b, channel-pair group); threads stride over L
__global__ void avgpool_fwd_kernel(const __nv_bfloat162* __restrict__ x, float* __restrict__ out, int B, int L, int C) {
    const int C2 = C >> 1;
    const int b = blockIdx.x;
    const __nv_bfloat162* xb = x + (size_t)b * L * C2;
    for (int cp = threadIdx.x % C2; cp < C2; cp += C2) {
        const int rlane = threadIdx.x / C2, rstep = blockDim.x / C2;
        float s0 = 0, s1 = 0;
        if (rlane < rstep)
            for (int l = rlane; l < L; l += rstep) { float2 v = __bfloat1622float2(xb[(size_t)l * C2 + cp]); s0 += v.x; s1 += v.y; }
        if (rlane < rstep) { atomicAdd(out + (size_t)b * C + 2 * cp, s0 / L); atomicAdd(out + (size_t)b * C + 2 * cp + 1, s1 / L); }
    }
}

__global__ void avgpool_bwd_kernel(const float* __restrict__ dout, __nv_bfloat162* __restrict__ dx, int B, int L, int C) {
    const int C2 = C >> 1;
    const int64_t total = (int64_t)B * L * C2;
    const float invL = 1.0f / L;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int cp = (int)(i % C2); const int64_t b = i / ((int64_t)L * C2);
        dx[i] = __floats2bfloat162_rn(dout[b * C + 2 * cp] * invL, dout[b * C + 2 * cp + 1] * invL);
    }
}

int avgpool_fwd(const void* x, float* out, int B, int L, int C, cudaStream_t st) {
    CS_REQUIRE(C % 2 == 0 && C <= 2048, "avgpool: C must be even and <= 2048");
    CS_CUDA_OK(cudaMemsetAsync(out, 0, (size_t)B * C * sizeof(float), st));
    int threads = C / 2 >= 1024 ? C / 2 : 1024;
    avgpool_fwd_kernel<<<B, threads, 0, st>>>((const __nv_bfloat162*)x, out, B, L, C);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

int avgpool_bwd(const float* dout, void* dx, int B, int L, int C, cudaStream_t st) {
    CS_REQUIRE(C % 2 == 0, "avgpool: C must be even");
    avgpool_bwd_kernel<<<ew_grid((int64_t)B * L * (C / 2)), 256, 0, st>>>(dout, (__nv_bfloat162*)dx, B, L, C);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// layout converters through a 32x32 smem tile transpose
__global__ void ncl_to_nlc_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, int C, int L) {
    __shared__ float tile[32][33];
    const int b = blockIdx.z, c0 = blockIdx.y * 32, l0 = blockIdx.x * 32;
    const float* xb = x + (size_t)b * C * L; __nv_bfloat16* yb = y + (size_t)b * C * L;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        int c = c0 + j, l = l0 + threadIdx.x;
        tile[j][threadIdx.x] = (c < C && l < L) ? xb[(size_t)c * L + l] : 0.0f;
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        int l = l0 + j, c = c0 + threadIdx.x;
        if (c < C && l < L) yb[(size_t)l * C + c] = __float2bfloat16(tile[threadIdx.x][j]);
    }
}
__global__ void nlc_to_ncl_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ y, int C, int L) {
    __shared__ float tile[32][33];
    const int b = blockIdx.z, c0 = blockIdx.y * 32, l0 = blockIdx.x * 32;
    const __nv_bfloat16* xb = x + (size_t)b * C * L; float* yb = y + (size_t)b * C * L;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        int l = l0 + j, c = c0 + threadIdx.x;
        tile[j][threadIdx.x] = (c < C && l < L) ? __bfloat162float(xb[(size_t)l * C + c]) : 0.0f;
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        int c = c0 + j, l = l0 + threadIdx.x;
        if (c < C && l < L) yb[(size_t)c * L + l] = tile[threadIdx.x][j];
    }
}
int ncl_to_nlc(const float* x, void* y, int B, int C, int L, cudaStream_t st) {
    dim3 grid((L + 31) / 32, (C + 31) / 32, B), block(32, 8);
    ncl_to_nlc_kernel<<<grid, block, 0, st>>>(x, (__nv_bfloat16*)y, C, L);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}
int nlc_to_ncl(const void* x, float* y, int B, int C, int L, cudaStream_t st) {
    dim3 grid((L + 31) / 32, (C + 31) / 32, B), block(32, 8);
    nlc_to_ncl_kernel<<<grid, block, 0, st>>>((const __nv_bfloat16*)x, y, C, L);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

}  // namespace cs
