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

// conv1d_narrow.cu - second generation for narrow layers (channels < 64): persistent, pipelined kernels on the 16-byte
// plane layout.  Same return convention as the sw128 entry points.
int conv1d_fwd_narrow(const void* x, const void* w, const float* bias, void* y, int B, int L, int Cin, int Cout, int K, int pad,
                      int relu, int Lout, cudaStream_t st);
int conv1d_wgrad_narrow(const void* x, const void* dy, float* dw, int B, int L, int Cin, int Cout, int K, int pad, int Lout,
                        cudaStream_t st);
// the same layers as 64-channel layers on packed positions (f positions per 128-byte row) through the sw128 kernels
int conv1d_fwd_packed(const void* x, const void* w, const float* bias, void* y, int B, int L, int Cin, int Cout, int K, int pad,
                      int relu, int Lout, cudaStream_t st);
int conv1d_wgrad_packed(const void* x, const void* dy, float* dw, int B, int L, int Cin, int Cout, int K, int pad, int Lout,
                        cudaStream_t st);

// planes view of a channels-last bf16 tensor [outer, rows, C]: dims (c8 = 8, row, plane = C/8, outer)
inline int make_plane_map(CUtensorMap* m, const void* base, int outer, int rows, int C, int box_rows, int box_planes) {
    EncodeTiledFn fn = encode_fn();
    CS_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from this driver");
    cuuint64_t dims[4] = {8, (cuuint64_t)rows, (cuuint64_t)(C / 8), (cuuint64_t)outer};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, 16, (cuuint64_t)rows * C * 2};
    cuuint32_t box[4] = {8, (cuuint32_t)box_rows, (cuuint32_t)box_planes, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CS_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) rows=%d C=%d box=%dx%d", (int)r, rows, C, box_rows, box_planes);
    return CS_OK;
}

__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

}  // namespace cs
