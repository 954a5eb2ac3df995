// Micro-probe: what does ONE small tcgen05.mma cost when many are issued back to back?  (debug entry point, no
// counterpart in the reference.)  The fused TinyECG kernels issue M128/M64 x N16/N32 x K16 MMAs - far below the
// tile sizes GEMM kernels use - so their tensor time is set by per-instruction costs, not by the FLOP rate.  This
// kernel issues `count` MMAs of one shape into `rot` rotating accumulators from zeroed shared memory and reports
// the cycles from first issue to the commit's arrival, for no-swizzle K-major (conv forward style) and MN-major
// (weight-gradient style) operands.
#include "common/cs_common.cuh"
#include "common/tcgen05.cuh"

namespace cs {
using namespace tc;

// fully unrolled issue (constant descriptor increments), `issuers` warps issuing concurrently into disjoint TMEM columns
template <int M, int N, bool MN, int ROT>
__global__ void __launch_bounds__(128, 1) mma_probe_kernel(int issuers, int shift_rows, long long* out) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    uint32_t* slot = reinterpret_cast<uint32_t*>(smem + 64);
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    for (int i = t; i < (96 * 1024) / 16; i += 128) reinterpret_cast<uint4*>(smem + 128)[i] = make_uint4(0, 0, 0, 0);
    if (t == 0) { for (int w = 0; w < 4; ++w) mbar_init(bar + w, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc(slot, 512);
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tb = *slot + (uint32_t)warp * 128;              // 128 columns per issuing warp
    const uint32_t sa = smem_u32(smem + 128) + warp * 8192, sb = smem_u32(smem + 128) + 40 * 1024;
    constexpr uint32_t idesc = idesc_f16(FMT_BF16, M, N, MN ? MAJOR_MN : MAJOR_K, MN ? MAJOR_MN : MAJOR_K);
    long long t0 = 0, t1 = 0, t2 = 0;
    constexpr int COUNT = 64;
    if (warp < issuers && lane == 0) {
        const uint64_t a0 = MN ? smem_desc(sa, 128, 16) : smem_desc(sa, 8192, 128);
        const uint64_t b0 = MN ? smem_desc(sb, 128, 8192) : smem_desc(sb, 256, 128);
        const uint64_t sh = (uint64_t)shift_rows;
        t0 = clock64();
#pragma unroll
        for (int c = 0; c < COUNT; ++c)
            mma_f16(tb + (c % ROT) * (N < 128 / ROT ? N : 128 / ROT), a0 + sh * (uint64_t)(c & 15), b0, idesc, c >= ROT ? 1u : 0u);
        t1 = clock64();
        mma_commit(bar + warp);
    }
    if (warp < issuers) {
        mbar_wait(bar + warp, 0);
        tc_fence_after_sync();
        if (lane == 0) { t2 = clock64(); out[2 * warp] = t1 - t0; out[2 * warp + 1] = t2 - t0; }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(*slot, 512);
}

template <int M, int N, bool MN, int ROT>
static int run_probe(int issuers, int shift_rows, long long* d) {
    const size_t smem = 128 + 96 * 1024;
    CS_CUDA_OK(cudaFuncSetAttribute(mma_probe_kernel<M, N, MN, ROT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int rep = 0; rep < 3; ++rep) mma_probe_kernel<M, N, MN, ROT><<<1, 128, smem>>>(issuers, shift_rows, d);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// ---- operand-layout probe: ONE MMA on host-provided shared-memory images, accumulator returned raw --------------------
// Lets a test pin down which shared-memory word feeds which (row, k) of an operand for a given kind / major-ness /
// LBO / SBO instead of trusting a reading of the documentation (tests/test_gpu_umma_layout.py).
struct LayoutProbeArgs {
    const uint8_t* a_img; const uint8_t* b_img; int a_bytes, b_bytes;
    uint32_t idesc; uint32_t lbo_a, sbo_a, lbo_b, sbo_b; int tf32, n_cols; float* out; uint32_t layout_a, layout_b;   // out[128 lanes][n_cols]
};
__global__ void __launch_bounds__(128, 1) mma_layout_probe_kernel(LayoutProbeArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    uint32_t* slot = reinterpret_cast<uint32_t*>(smem + 64);
    uint8_t* sa = smem + 1024; uint8_t* sb = smem + 1024 + 32 * 1024;     // 1024-aligned: swizzles act on address bits
    const int t = threadIdx.x, warp = t >> 5;
    for (int i = t; i < (64 * 1024) / 16; i += 128) reinterpret_cast<uint4*>(smem + 1024)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    for (int i = t; i < a.a_bytes / 4; i += 128) reinterpret_cast<uint32_t*>(sa)[i] = reinterpret_cast<const uint32_t*>(a.a_img)[i];
    for (int i = t; i < a.b_bytes / 4; i += 128) reinterpret_cast<uint32_t*>(sb)[i] = reinterpret_cast<const uint32_t*>(a.b_img)[i];
    if (t == 0) { mbar_init(bar, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc(slot, 64);
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tb = *slot;
    if (t == 0) {
        const uint64_t da = smem_desc(smem_u32(sa), a.lbo_a, a.sbo_a) | ((uint64_t)a.layout_a << 61),
                       db = smem_desc(smem_u32(sb), a.lbo_b, a.sbo_b) | ((uint64_t)a.layout_b << 61);
        if (smem_u32(smem) & 1023u) a.out[0] = -12345.0f;        // (never seen: dynamic shared memory starts 1024-aligned)
        if (a.tf32) mma_tf32(tb, da, db, a.idesc, 0u); else mma_f16(tb, da, db, a.idesc, 0u);
        mma_commit(bar);
    }
    mbar_wait(bar, 0);
    tc_fence_after_sync();
    for (int c0 = 0; c0 < a.n_cols; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tb + ((uint32_t)(warp * 32) << 16) + c0, v);
        tmem_ld_wait();
        for (int j = 0; j < 16; ++j) a.out[(size_t)t * a.n_cols + c0 + j] = __uint_as_float(v[j]);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(*slot, 64);
}

}  // namespace cs

/* debug: 64 MMAs per issuing warp.  config: 0 = M128 N16 K-major, 1 = M64 N16 MN-major, 2 = M64 N32 MN-major, 3 = M128 N64 K-major,
 * 4 = M128 N128 K-major; rot = 1 | 4 rotating accumulators.  cycles_host[2w] = issue loop of warp w, [2w+1] = issue -> complete. */
extern "C" int cs_debug_mma_probe(int config, int rot, int issuers, int shift_rows, long long* cycles_host) {
    using namespace cs;
    CS_REQUIRE(issuers >= 1 && issuers <= 4 && (rot == 1 || rot == 4), "mma_probe: bad arguments");
    long long* d = nullptr;
    CS_CUDA_OK(cudaMalloc(&d, 64));
    CS_CUDA_OK(cudaMemset(d, 0, 64));
    int rc = CS_ERR_INVALID;
    if (rot == 1) {
        if (config == 0) rc = run_probe<128, 16, false, 1>(issuers, shift_rows, d);
        if (config == 1) rc = run_probe<64, 16, true, 1>(issuers, shift_rows, d);
        if (config == 2) rc = run_probe<64, 32, true, 1>(issuers, shift_rows, d);
        if (config == 3) rc = run_probe<128, 64, false, 1>(issuers, shift_rows, d);
        if (config == 4) rc = run_probe<128, 128, false, 1>(issuers, shift_rows, d);
    } else {
        if (config == 0) rc = run_probe<128, 16, false, 4>(issuers, shift_rows, d);
        if (config == 1) rc = run_probe<64, 16, true, 4>(issuers, shift_rows, d);
        if (config == 2) rc = run_probe<64, 32, true, 4>(issuers, shift_rows, d);
        if (config == 3) rc = run_probe<128, 64, false, 4>(issuers, shift_rows, d);
        if (config == 4) rc = run_probe<128, 128, false, 4>(issuers, shift_rows, d);
    }
    if (rc == CS_OK) CS_CUDA_OK(cudaMemcpy(cycles_host, d, 64, cudaMemcpyDeviceToHost));
    cudaFree(d);
    return rc;
}

/* debug: one tcgen05.mma (fmt 0 f16 / 1 bf16 / 2 tf32; M 64|128; N multiple of 16 up to 64) on the given shared-memory
 * images (host pointers, each <= 32 KB); LBO / SBO in bytes; layout_* = descriptor bits [61,64): 0 none, 1 128B with 32 B
 * base (MN-major tf32), 2 128B, 4 64B, 6 32B.  d_host receives the raw
 * accumulator: 128 TMEM lanes x N fp32. */
extern "C" int cs_debug_mma_layout_probe(int fmt, int M, int N, int a_mn_major, int b_mn_major, int lbo_a, int sbo_a, int lbo_b, int sbo_b,
                                         const void* a_img, int a_bytes, const void* b_img, int b_bytes, float* d_host, int layout_a, int layout_b) {
    using namespace cs;
    CS_REQUIRE((M == 64 || M == 128) && N >= 16 && N <= 64 && N % 16 == 0 && fmt >= 0 && fmt <= 2, "layout probe: bad shape");
    CS_REQUIRE(a_bytes > 0 && b_bytes > 0 && a_bytes <= 32 * 1024 && b_bytes <= 32 * 1024 && a_bytes % 4 == 0 && b_bytes % 4 == 0, "layout probe: image size");
    uint8_t *da = nullptr, *db = nullptr; float* dout = nullptr;
    CS_CUDA_OK(cudaMalloc(&da, a_bytes)); CS_CUDA_OK(cudaMalloc(&db, b_bytes)); CS_CUDA_OK(cudaMalloc(&dout, (size_t)128 * N * 4));
    CS_CUDA_OK(cudaMemcpy(da, a_img, a_bytes, cudaMemcpyHostToDevice));
    CS_CUDA_OK(cudaMemcpy(db, b_img, b_bytes, cudaMemcpyHostToDevice));
    CS_CUDA_OK(cudaMemset(dout, 0, (size_t)128 * N * 4));
    LayoutProbeArgs a{da, db, a_bytes, b_bytes, idesc_f16((uint32_t)fmt, (uint32_t)M, (uint32_t)N, a_mn_major ? MAJOR_MN : MAJOR_K, b_mn_major ? MAJOR_MN : MAJOR_K),
                      (uint32_t)lbo_a, (uint32_t)sbo_a, (uint32_t)lbo_b, (uint32_t)sbo_b, fmt == 2, N, dout, (uint32_t)layout_a, (uint32_t)layout_b};
    const size_t smem = 1024 + 64 * 1024;
    CS_CUDA_OK(cudaFuncSetAttribute(mma_layout_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    mma_layout_probe_kernel<<<1, 128, smem>>>(a);
    CS_CUDA_OK(cudaGetLastError());
    CS_CUDA_OK(cudaMemcpy(d_host, dout, (size_t)128 * N * 4, cudaMemcpyDeviceToHost));
    cudaFree(da); cudaFree(db); cudaFree(dout);
    return CS_OK;
}
