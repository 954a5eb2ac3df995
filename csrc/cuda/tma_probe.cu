// Micro-probe: how many bytes per clock does ONE SM pull from cold DRAM through the TMA unit, as a function of the copy shape?
// (debug entry point, no counterpart in the reference.)  A persistent CTA per SM keeps `stages` loads in flight into a
// shared-memory ring and recycles a slot as soon as its mbarrier completes - no consumer, so the number is the load path's own
// ceiling.  Shapes: tiled 3-D boxes with rows of 32 / 64 / 128 bytes (what the conv kernels use), and 1-D bulk copies
// (cp.async.bulk.shared.global) of the same bytes.  The narrow conv forward (conv1d_narrow.cu) sits on this ceiling.
#include "conv1d_common.cuh"

namespace cs {

__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct TmaProbe { const uint8_t* base; long long total_bytes; int row_bytes, rows, stages, mode, chunk_bytes; long long* cycles; };

// mode 0: tensor box (row_bytes x rows) per load; mode 1: one 1-D bulk copy of chunk_bytes per load
__global__ void __launch_bounds__(128, 1) tma_stream_probe_kernel(const __grid_constant__ CUtensorMap map, TmaProbe p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint8_t* ring = smem + 1024;
    const int chunk = p.mode == 0 ? p.row_bytes * p.rows : p.chunk_bytes;
    const int stride = (chunk + 1023) & ~1023;
    const long long n_chunks = p.total_bytes / chunk;
    if (threadIdx.x == 0) {
        for (int i = 0; i < p.stages; ++i) mbar_init(full + i, 1);
        fence_mbar_init();
        tma_prefetch_desc(&map);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const long long t0 = clock64();
        long long issued = 0, done = 0;
        const long long mine = (n_chunks - blockIdx.x + gridDim.x - 1) / gridDim.x;
        while (done < mine) {
            while (issued < mine && issued - done < p.stages) {
                const int sa = (int)(issued % p.stages);
                const long long c = blockIdx.x + issued * gridDim.x;
                mbar_arrive_expect_tx(full + sa, chunk);
                if (p.mode == 0) tma_load_3d(ring + sa * stride, &map, full + sa, 0, (int)((c * p.rows) % (1 << 30)), 0);
                else bulk_load_1d(ring + sa * stride, p.base + c * chunk, chunk, full + sa);
                ++issued;
            }
            const int sd = (int)(done % p.stages);
            mbar_wait(full + sd, (uint32_t)((done / p.stages) & 1));
            ++done;
        }
        p.cycles[blockIdx.x] = clock64() - t0;
    }
}

}  // namespace cs

/* debug: stream `total_bytes` of `buf` (device pointer, >= L2 size so that the reads are cold) through every SM's TMA unit.
 * mode 0: tensor boxes of `rows` rows of `row_bytes` (32 / 64 / 128, matching swizzle); mode 1: 1-D bulk copies of chunk_bytes.
 * Returns the kernel time in ms (CUDA events) in *ms_out and the slowest CTA's cycle count in *max_cycles. */
extern "C" int cs_debug_tma_stream_probe(const void* buf, long long total_bytes, int mode, int row_bytes, int rows, int chunk_bytes,
                                         int stages, float* ms_out, long long* max_cycles) {
    using namespace cs;
    CS_REQUIRE(mode == 0 || mode == 1, "tma probe: mode");
    CS_REQUIRE(stages >= 1 && stages <= 32, "tma probe: stages");
    TmaProbe p;
    p.base = (const uint8_t*)buf; p.total_bytes = total_bytes; p.row_bytes = row_bytes; p.rows = rows; p.stages = stages; p.mode = mode;
    p.chunk_bytes = chunk_bytes;
    CUtensorMap map;
    memset(&map, 0, sizeof(map));
    if (mode == 0) {
        CS_REQUIRE((row_bytes == 32 || row_bytes == 64 || row_bytes == 128 || row_bytes == 16) && rows >= 1 && rows <= 256, "tma probe: box");
        EncodeTiledFn fn = encode_fn();
        CS_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available");
        const long long n_rows = total_bytes / row_bytes;
        cuuint64_t dims[3] = {(cuuint64_t)(row_bytes / 2), (cuuint64_t)n_rows, 1};
        cuuint64_t strides[2] = {(cuuint64_t)row_bytes, (cuuint64_t)n_rows * row_bytes};
        cuuint32_t box[3] = {(cuuint32_t)(row_bytes / 2), (cuuint32_t)rows, 1};
        cuuint32_t estr[3] = {1, 1, 1};
        const CUtensorMapSwizzle sw = row_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : row_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                      : row_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE;
        CUresult r = fn(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(buf), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        CS_REQUIRE(r == CUDA_SUCCESS, "tma probe: cuTensorMapEncodeTiled failed (%d)", (int)r);
    } else {
        CS_REQUIRE(chunk_bytes >= 16 && chunk_bytes % 16 == 0 && chunk_bytes <= 64 * 1024, "tma probe: chunk");
    }
    const int chunk = mode == 0 ? row_bytes * rows : chunk_bytes;
    const size_t smem = 1024 + (size_t)stages * ((chunk + 1023) & ~1023);
    CS_REQUIRE(smem <= 200 * 1024, "tma probe: ring too large (%zu B)", smem);
    CS_CUDA_OK(cudaFuncSetAttribute(tma_stream_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long* cyc = nullptr;
    const int grid = num_sms();
    CS_CUDA_OK(cudaMalloc(&cyc, grid * sizeof(long long)));
    p.cycles = cyc;
    cudaEvent_t e0, e1;
    CS_CUDA_OK(cudaEventCreate(&e0)); CS_CUDA_OK(cudaEventCreate(&e1));
    CS_CUDA_OK(cudaEventRecord(e0));
    tma_stream_probe_kernel<<<grid, 128, smem>>>(map, p);
    CS_CUDA_OK(cudaEventRecord(e1));
    CS_CUDA_OK(cudaEventSynchronize(e1));
    CS_CUDA_OK(cudaGetLastError());
    CS_CUDA_OK(cudaEventElapsedTime(ms_out, e0, e1));
    long long* h = (long long*)malloc(grid * sizeof(long long));
    CS_CUDA_OK(cudaMemcpy(h, cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost));
    long long mx = 0;
    for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    *max_cycles = mx;
    free(h); cudaFree(cyc); cudaEventDestroy(e0); cudaEventDestroy(e1);
    return CS_OK;
}
