// Shared host helpers of the conv1d translation units (tensor-map encoding, tuning overrides, cross-file entry points).
#pragma once
#include "common/cs_common.cuh"
#include "common/tcgen05.cuh"

#include <cuda.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

namespace cs {

using namespace tc;

constexpr int kFwdThreads = 192;        // warp 0 TMA producer, warp 1 MMA issuer, warps 2-5 epilogue
constexpr int kWgThreads = 192;

// driver entry point fetched through the runtime: no -lcuda needed
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// integer tuning / diagnostic override from the environment (read at every call: tests flip them per case)
inline int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return (e && *e) ? atoi(e) : dflt;
}

// host-side bundle for the fused output statistics of the forward kernel (see Fwd128Shape::bn_acc)
struct ConvBnArgs {
    float* acc; int32_t* ticket; float* mean; float* rstd; float* running_mean; float* running_var;
    float eps, momentum; const float* gamma; const float* beta; float* coef;
};

// conv1d_sw128.cu — wide layers on 128-byte swizzled rows.  Both return CS_OK after launching, an error code, or -1
// when the shape is outside the variant (the callers in conv1d_planes.cu then use the 16-byte plane kernels).
int conv1d_fwd_sw128(const void* x, const void* w, const float* bias, void* y, int B, int L, int Cin, int Cout, int K, int pad,
                     int relu, int Lout, cudaStream_t st, int wmn = 0, const ConvBnArgs* bn = nullptr);
int conv1d_wgrad_sw128(const void* x, const void* dy, float* dw, int B, int L, int Cin, int Cout, int K, int pad, int Lout,
                       cudaStream_t st);

}  // namespace cs
