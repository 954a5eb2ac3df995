// Shared host/device helpers for the sm_100a kernel library (libcrossscale_cuda.so).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../../include/crossscale.h"

namespace cs {

// ---- error plumbing (thread-local last error string, returned by cs_last_error) ----------
void set_error(const char* fmt, ...);
int  fail(cudaError_t e, const char* what);

#define CS_CUDA_OK(expr)                                             \
    do {                                                             \
        cudaError_t _e = (expr);                                     \
        if (_e != cudaSuccess) return ::cs::fail(_e, #expr);         \
    } while (0)

#define CS_REQUIRE(cond, ...)                                        \
    do {                                                             \
        if (!(cond)) { ::cs::set_error(__VA_ARGS__); return CS_ERR_INVALID; } \
    } while (0)

int num_sms();   // SM count of the current device (cached per device)

// ---- TinyECG flat parameter layout (state_dict order, models/tiny_ecg.py) ------------------
struct TinyLayout {
    static constexpr int W1 = 0;        // [16][1][7]
    static constexpr int B1 = 112;      // [16]
    static constexpr int W2 = 128;      // [16 co][16 ci][5]
    static constexpr int B2 = 1408;     // [16]
    static constexpr int WH = 1424;     // [C][16]
    __host__ __device__ static constexpr int bh(int C) { return WH + 16 * C; }
    __host__ __device__ static constexpr int total(int C) { return WH + 17 * C; }
};

// ctrl / fstate slots of the on-device training state (see include/crossscale.h)
enum { CTRL_TICKET = 0, CTRL_GROWTH = 1, CTRL_STEPS = 2, CTRL_SKIPPED = 3, CTRL_FOUND_INF = 4, CTRL_ERROR = 5, CTRL_POS = 6 /* int64: samples drawn so far (in-kernel permutation sampler) */ };
enum { FS_LOSS_SUM = 0, FS_SCALE = 1, FS_LAST_LOSS = 2, FS_LOSS_STEP = 3 };

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- BatchNorm statistics tail (shared by bn_stats_kernel and the conv forward epilogue) ------------------------
// Run by the LAST block to arrive on `ticket`: the per-channel totals are complete in `stats` (L2 atomics), so that
// block turns them into the layer's outputs and CLEARS the accumulators — a persistent, self-cleaning workspace.
struct BnStatsTail {
    int32_t* ticket;          // nullptr: plain accumulate-only kernel
    float* mean; float* rstd; float* running_mean; float* running_var;
    float rows, eps, momentum;
    const float* gamma; const float* beta; float* coef;   // optional: coef[0:C] = rstd*gamma, coef[C:2C] = beta - mean*rstd*gamma
};

__device__ __forceinline__ void bn_stats_tail_channels(float* stats, int C, const BnStatsTail& t, int tid, int nthreads) {
    for (int c = tid; c < C; c += nthreads) {
        // every load of the tail is issued before the first use: one L2 round trip, not four
        const float s0 = __ldcg(stats + c), s1 = __ldcg(stats + C + c);
        const float rm = t.running_mean ? t.running_mean[c] : 0.0f, rv = t.running_mean ? t.running_var[c] : 0.0f;
        const float ga = t.coef ? t.gamma[c] : 0.0f, be = t.coef ? t.beta[c] : 0.0f;
        const float m = s0 / t.rows;
        const float var = fmaxf(s1 / t.rows - m * m, 0.0f);
        const float rs = rsqrtf(var + t.eps);
        t.mean[c] = m;
        t.rstd[c] = rs;
        if (t.running_mean) {
            t.running_mean[c] = rm + t.momentum * (m - rm);
            t.running_var[c] = rv + t.momentum * (var * (t.rows / fmaxf(t.rows - 1.0f, 1.0f)) - rv);
        }
        if (t.coef) {
            t.coef[c] = rs * ga;
            t.coef[C + c] = be - m * rs * ga;
        }
        stats[c] = 0.0f; stats[C + c] = 0.0f;
    }
}

__device__ __forceinline__ bool finite_f(float v) { return fabsf(v) <= 3.402823466e38f; }

}  // namespace cs
