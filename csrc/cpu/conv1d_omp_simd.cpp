// Batched single-channel "valid" cross-correlation on the CPU:  y[b,i] = sum_k x[b,i+k] * w[k].
//
// Same exported C symbol and argument list as the reference kernel
// (Module_2/conv1d_openmp_simd.c:21-28) so its ctypes benchmarks bind unchanged, but a different
// algorithm: the reference vectorises over the K taps (its AVX2 body never runs for K < 8, SURVEY
// C13); here the SIMD lanes run over *output positions* — taps are broadcast scalars held in
// registers — so K = 3/5/7 are fully vectorised, with AVX-512 / AVX2+FMA / scalar variants picked
// once at run time.  OpenMP splits (row, position-block) pairs so small batches still fill all cores.
#include <immintrin.h>
#include <omp.h>
#include <stddef.h>
#include <stdint.h>

#include <algorithm>

#include "crossscale.h"

namespace {

constexpr int kPosBlock = 2048;   // outputs per work item (8 KB of y; x slice stays in L1)

void row_scalar(const float* x, const float* w, float* y, int n, int K) {
    for (int i = 0; i < n; ++i) {
        float acc = 0.0f;
        for (int k = 0; k < K; ++k) acc += x[i + k] * w[k];
        y[i] = acc;
    }
}

__attribute__((target("avx2,fma"))) void row_avx2(const float* x, const float* w, float* y, int n, int K) {
    int i = 0;
    for (; i + 16 <= n; i += 16) {             // two accumulators hide FMA latency
        __m256 a0 = _mm256_setzero_ps(), a1 = _mm256_setzero_ps();
        for (int k = 0; k < K; ++k) {
            const __m256 wk = _mm256_broadcast_ss(w + k);
            a0 = _mm256_fmadd_ps(_mm256_loadu_ps(x + i + k), wk, a0);
            a1 = _mm256_fmadd_ps(_mm256_loadu_ps(x + i + 8 + k), wk, a1);
        }
        _mm256_storeu_ps(y + i, a0);
        _mm256_storeu_ps(y + i + 8, a1);
    }
    for (; i + 8 <= n; i += 8) {
        __m256 a0 = _mm256_setzero_ps();
        for (int k = 0; k < K; ++k)
            a0 = _mm256_fmadd_ps(_mm256_loadu_ps(x + i + k), _mm256_broadcast_ss(w + k), a0);
        _mm256_storeu_ps(y + i, a0);
    }
    if (i < n) row_scalar(x + i, w, y + i, n - i, K);
}

__attribute__((target("avx512f"))) void row_avx512(const float* x, const float* w, float* y, int n, int K) {
    int i = 0;
    for (; i + 32 <= n; i += 32) {
        __m512 a0 = _mm512_setzero_ps(), a1 = _mm512_setzero_ps();
        for (int k = 0; k < K; ++k) {
            const __m512 wk = _mm512_set1_ps(w[k]);
            a0 = _mm512_fmadd_ps(_mm512_loadu_ps(x + i + k), wk, a0);
            a1 = _mm512_fmadd_ps(_mm512_loadu_ps(x + i + 16 + k), wk, a1);
        }
        _mm512_storeu_ps(y + i, a0);
        _mm512_storeu_ps(y + i + 16, a1);
    }
    if (i < n) {                                // masked tail: no scalar clean-up loop
        for (; i < n; i += 16) {
            const int rem = std::min(16, n - i);
            const __mmask16 m = (__mmask16)((1u << rem) - 1u);
            __m512 a0 = _mm512_setzero_ps();
            for (int k = 0; k < K; ++k)
                a0 = _mm512_fmadd_ps(_mm512_maskz_loadu_ps(m, x + i + k), _mm512_set1_ps(w[k]), a0);
            _mm512_mask_storeu_ps(y + i, m, a0);
        }
    }
}

using RowFn = void (*)(const float*, const float*, float*, int, int);

int g_isa = -1;
RowFn g_row = nullptr;

void pick() {
    if (g_row) return;
    __builtin_cpu_init();
    if (__builtin_cpu_supports("avx512f")) { g_row = row_avx512; g_isa = 5; }
    else if (__builtin_cpu_supports("avx2") && __builtin_cpu_supports("fma")) { g_row = row_avx2; g_isa = 2; }
    else { g_row = row_scalar; g_isa = 0; }
}

}  // namespace

extern "C" {

void conv1d_batch_omp_simd(const float* x, const float* w, float* y, int batch, int L, int K, int nthreads) {
    pick();
    const int outL = L - K + 1;
    if (batch <= 0 || outL <= 0) return;
    const int nblk = (outL + kPosBlock - 1) / kPosBlock;
    const long items = (long)batch * nblk;
    const RowFn row = g_row;
    if (nthreads < 1) nthreads = 1;
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (long it = 0; it < items; ++it) {
        const int b = (int)(it / nblk), blk = (int)(it % nblk);
        const int i0 = blk * kPosBlock;
        const int n = std::min(kPosBlock, outL - i0);
        row(x + (size_t)b * L + i0, w, y + (size_t)b * outL + i0, n, K);
    }
}

int cs_cpu_conv1d_isa(void) { pick(); return g_isa; }

}  // extern "C"
