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
