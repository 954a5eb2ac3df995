// Fused TinyECG kernels, fp32 FFMA path (config G0 and the exact-precision oracle on device).
//
// One CTA owns one sample at a time and keeps every activation of that sample in shared memory:
// x -> conv1+ReLU -> conv2+ReLU -> avg-pool -> head -> CE -> full backward, so HBM traffic per
// sample is the 2 KB input row plus the weight-gradient atomics at the end.  In MODE_TRAIN the
// last CTA to finish (ticket counter) performs grad unscale / inf-check / SGD-momentum / scaler
// update, so a whole local step is a single launch (ref: train_step_G0,
// TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:103-114, ~40 launches + sync + .item()).
#include "tiny_ecg.cuh"

namespace cs {

constexpr int kThreads = 256;
constexpr int kMaxC = kTinyMaxC;

__host__ __device__ inline int pitch_of(int L) { int p = L + 4; return (p & 1) ? p : p + 1; }
__host__ __device__ inline int xs_len(int L) { return (L + 6 + 3) & ~3; }

struct SmemPlan {
    int w1, b1, w2f, w2b, b2, wh, bh, pooled, dlog, gj, red, xs, h1, dz2, total;
    __host__ __device__ SmemPlan(int L, int C) {
        int o = 0;
        w1 = o; o += 112; b1 = o; o += 16; w2f = o; o += 1280; w2b = o; o += 1280;
        b2 = o; o += 16; wh = o; o += 16 * kMaxC; bh = o; o += kMaxC; pooled = o; o += 16;
        dlog = o; o += kMaxC; gj = o; o += 16; red = o; o += 8 * 16 + 8;
        o = (o + 3) & ~3;
        xs = o; o += xs_len(L);
        h1 = o; o += 16 * pitch_of(L);
        dz2 = o; o += 16 * pitch_of(L);
        total = o;
    }
};

template <int MODE>
__global__ void __launch_bounds__(kThreads, 2) tiny_ecg_f32_kernel(TinyArgs a) {
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int L = a.L, C = a.C, P = pitch_of(L);
    const SmemPlan sp(L, C);
    float* w1s = sm + sp.w1;  float* b1s = sm + sp.b1;  float* w2f = sm + sp.w2f;
    float* w2b = sm + sp.w2b; float* b2s = sm + sp.b2;  float* whs = sm + sp.wh;
    float* bhs = sm + sp.bh;  float* pooled = sm + sp.pooled; float* dlog = sm + sp.dlog;
    float* gj = sm + sp.gj;   float* red = sm + sp.red; float* xs = sm + sp.xs;
    float* h1 = sm + sp.h1;   float* dz2 = sm + sp.dz2;
    const float* prm = a.params;

    // ---- stage weights ---------------------------------------------------------------------
    for (int i = tid; i < 112; i += kThreads) w1s[i] = prm[TinyLayout::W1 + i];
    if (tid < 16) { b1s[tid] = prm[TinyLayout::B1 + tid]; b2s[tid] = prm[TinyLayout::B2 + tid]; }
    for (int i = tid; i < 1280; i += kThreads) {          // global order [co][ci][k]
        int co = i / 80, ci = (i / 5) % 16, k = i % 5;
        float w = prm[TinyLayout::W2 + i];
        w2f[(ci * 5 + k) * 16 + co] = w;                  // forward: 16 co contiguous
        w2b[(co * 5 + k) * 16 + ci] = w;                  // dgrad:   16 ci contiguous
    }
    for (int i = tid; i < 16 * C; i += kThreads) whs[i] = prm[TinyLayout::WH + i];
    if (tid < C) bhs[tid] = prm[TinyLayout::bh(C) + tid];
    // halos stay zero for the whole kernel (only interiors are rewritten per sample)
    for (int i = tid; i < xs_len(L); i += kThreads) xs[i] = 0.0f;
    for (int i = tid; i < 16 * P; i += kThreads) { h1[i] = 0.0f; dz2[i] = 0.0f; }
    __syncthreads();

    const float invL = 1.0f / (float)L;
    const float invB = 1.0f / (float)a.B;
    const float lscale = (MODE == MODE_TRAIN && a.use_scaler) ? a.fstate[FS_SCALE] : 1.0f;

    // per-thread gradient accumulators that persist over this CTA's samples
    float gw2[5] = {0.f, 0.f, 0.f, 0.f, 0.f};   // thread (co = tid/16, ci = tid%16), taps 0..4
    float gsmall = 0.0f;                        // dW1 (0..111), db1 (112..127), db2 (128..143)
    float ghead = 0.0f;                         // dWh (0..16C-1), dbh (16C..17C-1)
    float loss_acc = 0.0f;

    for (int s = blockIdx.x; s < a.B; s += gridDim.x) {
        const int64_t row = a.idx ? a.idx[s] : (int64_t)s;
        // ---- load the window ----------------------------------------------------------------
        if (a.x_bf16) {
            const __nv_bfloat16* xr = (const __nv_bfloat16*)a.x + row * L;
            for (int l = tid; l < L; l += kThreads) xs[l + 3] = __bfloat162float(xr[l]);
        } else {
            const float* xr = (const float*)a.x + row * L;
            for (int l = tid; l < L; l += kThreads) xs[l + 3] = __ldg(xr + l);
        }
        int label = 0;
        if (MODE == MODE_TRAIN) {
            label = (int)a.y[row];
            // a label outside [0, C) would index the logits out of bounds: train it as class 0 and raise the error bit
            if ((unsigned)label >= (unsigned)C) { if (threadIdx.x == 0) atomicOr(a.ctrl + CTRL_ERROR, 1); label = 0; }
        }
        __syncthreads();

        // ---- conv1 (1->16, k7, pad 3) + ReLU --------------------------------------------------
        for (int l = tid; l < L; l += kThreads) {
            float xv[7];
#pragma unroll
            for (int k = 0; k < 7; ++k) xv[k] = xs[l + k];
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                float acc = b1s[c];
#pragma unroll
                for (int k = 0; k < 7; ++k) acc = fmaf(w1s[c * 7 + k], xv[k], acc);
                h1[c * P + l + 2] = fmaxf(acc, 0.0f);
            }
        }
        __syncthreads();

        // ---- conv2 (16->16, k5, pad 2) + ReLU + pool partials; mask -> dz2 ---------------------
        float pool[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) pool[c] = 0.0f;
        for (int base = 0; base < L; base += 2 * kThreads) {
            const int l0 = base + tid, l1 = l0 + kThreads;
            const bool v0 = l0 < L, v1 = l1 < L;
            const int i0 = v0 ? l0 : 0, i1 = v1 ? l1 : 0;
            float acc0[16], acc1[16];
#pragma unroll
            for (int c = 0; c < 16; ++c) { acc0[c] = b2s[c]; acc1[c] = b2s[c]; }
#pragma unroll 1
            for (int ci = 0; ci < 16; ++ci) {
#pragma unroll
                for (int k = 0; k < 5; ++k) {
                    const float a0 = h1[ci * P + i0 + k], a1 = h1[ci * P + i1 + k];
                    const float4* w4 = reinterpret_cast<const float4*>(w2f + (ci * 5 + k) * 16);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        float4 w = w4[q];
                        acc0[4*q+0] = fmaf(w.x, a0, acc0[4*q+0]); acc1[4*q+0] = fmaf(w.x, a1, acc1[4*q+0]);
                        acc0[4*q+1] = fmaf(w.y, a0, acc0[4*q+1]); acc1[4*q+1] = fmaf(w.y, a1, acc1[4*q+1]);
                        acc0[4*q+2] = fmaf(w.z, a0, acc0[4*q+2]); acc1[4*q+2] = fmaf(w.z, a1, acc1[4*q+2]);
                        acc0[4*q+3] = fmaf(w.w, a0, acc0[4*q+3]); acc1[4*q+3] = fmaf(w.w, a1, acc1[4*q+3]);
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                if (v0) { pool[c] += fmaxf(acc0[c], 0.0f); if (MODE != MODE_FWD) dz2[c * P + l0 + 2] = acc0[c] > 0.0f ? 1.0f : 0.0f; }
                if (v1) { pool[c] += fmaxf(acc1[c], 0.0f); if (MODE != MODE_FWD) dz2[c * P + l1 + 2] = acc1[c] > 0.0f ? 1.0f : 0.0f; }
            }
        }
#pragma unroll
        for (int c = 0; c < 16; ++c) {
            float v = warp_sum(pool[c]);
            if (lane == 0) red[warp * 16 + c] = v;
        }
        __syncthreads();
        if (tid < 16) {
            float v = 0.0f;
#pragma unroll
            for (int w = 0; w < kThreads / 32; ++w) v += red[w * 16 + tid];
            pooled[tid] = v * invL;
        }
        __syncthreads();

        // ---- head + cross entropy --------------------------------------------------------------
        if (tid == 0) {
            float lg[kMaxC];
            float mx = -3.0e38f;
            for (int c = 0; c < C; ++c) {
                float v = bhs[c];
#pragma unroll
                for (int j = 0; j < 16; ++j) v = fmaf(whs[c * 16 + j], pooled[j], v);
                lg[c] = v; mx = fmaxf(mx, v);
            }
            if (MODE == MODE_FWD) {
                for (int c = 0; c < C; ++c) a.logits[(size_t)s * C + c] = lg[c];
            } else if (MODE == MODE_TRAIN) {
                float se = 0.0f;
                for (int c = 0; c < C; ++c) se += __expf(lg[c] - mx);
                const float lse = mx + __logf(se);
                loss_acc += (lse - lg[label]) * invB;
                for (int c = 0; c < C; ++c) {
                    float p = __expf(lg[c] - lse);
                    dlog[c] = (p - (c == label ? 1.0f : 0.0f)) * invB * lscale;
                }
            } else {
                for (int c = 0; c < C; ++c) dlog[c] = a.dlogits[(size_t)s * C + c];
            }
        }
        if (MODE == MODE_FWD) { __syncthreads(); continue; }
        __syncthreads();

        // ---- head grads, dp -> gj = dp / L --------------------------------------------------------
        if (tid < 16 * C) ghead = fmaf(dlog[tid >> 4], pooled[tid & 15], ghead);
        else if (tid < 17 * C) ghead += dlog[tid - 16 * C];
        if (tid < 16) {
            float v = 0.0f;
            for (int c = 0; c < C; ++c) v = fmaf(whs[c * 16 + tid], dlog[c], v);
            gj[tid] = v * invL;
        }
        __syncthreads();
        // dz2 = mask * gj[co]
        for (int i = tid; i < 16 * L; i += kThreads) {
            int co = i / L, l = i - co * L;
            dz2[co * P + l + 2] *= gj[co];
        }
        __syncthreads();

        // ---- wgrad conv2: dW2[co][ci][k] += sum_l dz2[co][l] * h1[ci][l+k-2] ----------------------
        {
            const int co = tid >> 4, ci = tid & 15;
            const float* dr = dz2 + co * P + 2;
            const float* hr = h1 + ci * P;            // hr[l + k] = h1 at position l+k-2
            float h0 = hr[0], h1v = hr[1], h2 = hr[2], h3 = hr[3];
            for (int l = 0; l < L; ++l) {
                const float d = dr[l];
                const float h4 = hr[l + 4];
                gw2[0] = fmaf(d, h0, gw2[0]); gw2[1] = fmaf(d, h1v, gw2[1]); gw2[2] = fmaf(d, h2, gw2[2]);
                gw2[3] = fmaf(d, h3, gw2[3]); gw2[4] = fmaf(d, h4, gw2[4]);
                h0 = h1v; h1v = h2; h2 = h3; h3 = h4;
            }
        }
        __syncthreads();

        // ---- dgrad conv2 + ReLU mask of conv1: dz1 (in place over h1) ----------------------------
        for (int base = 0; base < L; base += 2 * kThreads) {
            const int l0 = base + tid, l1 = l0 + kThreads;
            const bool v0 = l0 < L, v1 = l1 < L;
            const int i0 = v0 ? l0 : 0, i1 = v1 ? l1 : 0;
            float acc0[16], acc1[16];
#pragma unroll
            for (int c = 0; c < 16; ++c) { acc0[c] = 0.0f; acc1[c] = 0.0f; }
#pragma unroll 1
            for (int co = 0; co < 16; ++co) {
#pragma unroll
                for (int k = 0; k < 5; ++k) {
                    // dh1[ci][l] += dz2[co][l - k + 2] * w2[co][ci][k];  dz2 storage index = pos + 2
                    const float a0 = dz2[co * P + i0 + 4 - k], a1 = dz2[co * P + i1 + 4 - k];
                    const float4* w4 = reinterpret_cast<const float4*>(w2b + (co * 5 + k) * 16);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        float4 w = w4[q];
                        acc0[4*q+0] = fmaf(w.x, a0, acc0[4*q+0]); acc1[4*q+0] = fmaf(w.x, a1, acc1[4*q+0]);
                        acc0[4*q+1] = fmaf(w.y, a0, acc0[4*q+1]); acc1[4*q+1] = fmaf(w.y, a1, acc1[4*q+1]);
                        acc0[4*q+2] = fmaf(w.z, a0, acc0[4*q+2]); acc1[4*q+2] = fmaf(w.z, a1, acc1[4*q+2]);
                        acc0[4*q+3] = fmaf(w.w, a0, acc0[4*q+3]); acc1[4*q+3] = fmaf(w.w, a1, acc1[4*q+3]);
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < 16; ++c) {
                if (v0) { float h = h1[c * P + l0 + 2]; h1[c * P + l0 + 2] = h > 0.0f ? acc0[c] : 0.0f; }
                if (v1) { float h = h1[c * P + l1 + 2]; h1[c * P + l1 + 2] = h > 0.0f ? acc1[c] : 0.0f; }
            }
        }
        __syncthreads();

        // ---- small weight grads: dW1, db1, db2 -----------------------------------------------------
        if (tid < 112) {
            const int c = tid / 7, k = tid - c * 7;
            const float* dr = h1 + c * P + 2;       // dz1[c][l]
            float acc = 0.0f;
            for (int l = 0; l < L; ++l) acc = fmaf(dr[l], xs[l + k], acc);
            gsmall += acc;
        } else if (tid < 128) {
            const float* dr = h1 + (tid - 112) * P + 2;
            float acc = 0.0f;
            for (int l = 0; l < L; ++l) acc += dr[l];
            gsmall += acc;
        } else if (tid < 144) {
            const float* dr = dz2 + (tid - 128) * P + 2;
            float acc = 0.0f;
            for (int l = 0; l < L; ++l) acc += dr[l];
            gsmall += acc;
        }
        __syncthreads();
    }

    if (MODE == MODE_FWD) return;

    // ---- flush this CTA's gradient partials ----------------------------------------------------------
    float* g = a.gacc;
    {
        const int co = tid >> 4, ci = tid & 15;
#pragma unroll
        for (int k = 0; k < 5; ++k) atomicAdd(g + TinyLayout::W2 + (co * 16 + ci) * 5 + k, gw2[k]);
    }
    if (tid < 112) atomicAdd(g + TinyLayout::W1 + tid, gsmall);
    else if (tid < 128) atomicAdd(g + TinyLayout::B1 + (tid - 112), gsmall);
    else if (tid < 144) atomicAdd(g + TinyLayout::B2 + (tid - 128), gsmall);
    if (tid < 17 * C) atomicAdd(g + TinyLayout::WH + tid, ghead);   // WH then bh are contiguous

    if (MODE == MODE_TRAIN) {
        if (tid == 0) atomicAdd(a.fstate + FS_LOSS_STEP, loss_acc);
        __threadfence();
        __syncthreads();
        __shared__ int is_last;
        if (tid == 0) {
            int t = atomicAdd(a.ctrl + CTRL_TICKET, 1);
            is_last = (t == (int)gridDim.x - 1);
        }
        __syncthreads();
        if (is_last) { __threadfence(); tiny_sgd_tail<kThreads>(a, TinyLayout::total(C)); }
    }
}

template <int MODE>
static int launch_f32_mode(const TinyArgs& a, cudaStream_t st) {
    CS_REQUIRE(a.C >= 1 && a.C <= kMaxC, "TinyECG fused kernels support 1..%d classes (got %d)", kMaxC, a.C);
    CS_REQUIRE(a.B >= 1 && a.L >= 8, "bad shape B=%d L=%d", a.B, a.L);
    SmemPlan sp(a.L, a.C);
    const size_t smem = (size_t)sp.total * sizeof(float);
    CS_REQUIRE(smem <= 225 * 1024, "window length %d needs %zu B smem (> 225 KB)", a.L, smem);
    const int per_sm = smem <= 110 * 1024 ? 2 : 1;
    const int grid = a.B < num_sms() * per_sm ? a.B : num_sms() * per_sm;
    // opt in to >48 KB dynamic smem; raised only when a longer window needs more (never in the steady
    // state, so CUDA-graph capture sees no attribute calls).  One record per kernel instantiation.
    static size_t configured[64] = {};
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    if (smem > configured[dev & 63]) {
        CS_CUDA_OK(cudaFuncSetAttribute(tiny_ecg_f32_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev & 63] = smem;
    }
    tiny_ecg_f32_kernel<MODE><<<grid, kThreads, smem, st>>>(a);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

static int launch_f32(int mode, const TinyArgs& a, cudaStream_t st) {
    if (mode == MODE_FWD) return launch_f32_mode<MODE_FWD>(a, st);
    if (mode == MODE_BWD) return launch_f32_mode<MODE_BWD>(a, st);
    return launch_f32_mode<MODE_TRAIN>(a, st);
}

int tiny_f32_forward(const TinyArgs& a, cudaStream_t st) { return launch_f32(MODE_FWD, a, st); }
int tiny_f32_backward(const TinyArgs& a, cudaStream_t st) { return launch_f32(MODE_BWD, a, st); }
int tiny_f32_train(const TinyArgs& a, cudaStream_t st) { return launch_f32(MODE_TRAIN, a, st); }
size_t tiny_f32_smem_bytes(int L, int C) { return (size_t)SmemPlan(L, C).total * sizeof(float); }

}  // namespace cs
