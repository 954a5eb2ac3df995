// Argument block shared by the fused TinyECG kernels (fp32 FFMA path and tcgen05 path).
#pragma once
#include "common/cs_common.cuh"

namespace cs {

constexpr int kTinyMaxC = 8;
enum { MODE_FWD = 0, MODE_BWD = 1, MODE_TRAIN = 2 };

struct TinyArgs {
    const void* x; const int64_t* y; const int64_t* idx; int x_bf16;
    int B, L, C;
    float* params; float* mom; float* gacc; int32_t* ctrl; float* fstate;
    float* logits; const float* dlogits;
    float lr, momentum, wd; int use_scaler; float growth, backoff; int interval;
    int amp;
    long long* prof;   // debug: per-phase clock64 stamps of CTA 0 (NULL in production)
};

int tiny_f32_forward(const TinyArgs& a, cudaStream_t st);
int tiny_f32_backward(const TinyArgs& a, cudaStream_t st);
int tiny_f32_train(const TinyArgs& a, cudaStream_t st);
size_t tiny_f32_smem_bytes(int L, int C);

int tiny_umma_forward(const TinyArgs& a, cudaStream_t st);
int tiny_umma_backward(const TinyArgs& a, cudaStream_t st);
int tiny_umma_train(const TinyArgs& a, cudaStream_t st);
size_t tiny_umma_smem_bytes(int L, int C);
bool tiny_umma_supported(int L, int C);
size_t tiny_umma_scratch_floats();
extern long long* g_tiny_prof;   // set by cs_debug_set_tiny_profile

// SGD / scaler tail shared by both paths; run by the last CTA of a training-step kernel.
template <int THREADS>
__device__ __forceinline__ void tiny_sgd_tail(const TinyArgs& a, int P) {
    const int tid = threadIdx.x;
    const float scale = a.use_scaler ? a.fstate[FS_SCALE] : 1.0f;
    const float inv = 1.0f / scale;
    int bad = 0;
    for (int i = tid; i < P; i += THREADS) {
        float g = __ldcg(a.gacc + i) * inv;
        if (!finite_f(g)) bad = 1;
    }
    bad = __syncthreads_or(bad);
    if (!bad) {
        for (int i = tid; i < P; i += THREADS) {
            float p = a.params[i];
            float g = __ldcg(a.gacc + i) * inv + a.wd * p;
            float m = a.momentum * a.mom[i] + g;
            a.mom[i] = m;
            a.params[i] = p - a.lr * m;
        }
    }
    for (int i = tid; i < P; i += THREADS) a.gacc[i] = 0.0f;
    if (tid == 0) {
        float loss = __ldcg(a.fstate + FS_LOSS_STEP);
        a.fstate[FS_LOSS_SUM] += loss;
        a.fstate[FS_LAST_LOSS] = loss;
        a.fstate[FS_LOSS_STEP] = 0.0f;
        a.ctrl[CTRL_TICKET] = 0;
        a.ctrl[CTRL_STEPS] += 1;
        a.ctrl[CTRL_FOUND_INF] = bad;
        if (bad) a.ctrl[CTRL_SKIPPED] += 1;
        if (a.use_scaler) {   // torch.cuda.amp.GradScaler.update()
            if (bad) { a.fstate[FS_SCALE] = scale * a.backoff; a.ctrl[CTRL_GROWTH] = 0; }
            else {
                int t = a.ctrl[CTRL_GROWTH] + 1;
                if (t >= a.interval) { a.fstate[FS_SCALE] = scale * a.growth; t = 0; }
                a.ctrl[CTRL_GROWTH] = t;
            }
        }
    }
}

}  // namespace cs
