// conv1d on tcgen05, second generation for NARROW layers (Cin, Cout < 64 ... 64): persistent, pipelined kernels on the
// 16-byte plane layout of conv1d_planes.cu.
//
// Why: a narrow layer is memory bound (C = 16, K = 3, B = 4096, L = 500 moves 131 MB for 1.6 GFLOP), and the first-generation
// kernel runs ONE (sample, 4 tiles) work item per CTA: TMA -> MMA -> epilogue back to back, nothing of the next item in
// flight, 2.8 us per item.  Here a CTA per SM stays resident and streams work items through a pipeline:
//   forward / data gradient (conv1d_fwd_narrow):
//     warp 0        TMA producer: whole-position rows (32 / 64 bytes, 32B / 64B swizzle) of the next items into an NS-deep
//                   ring; the whole weight tensor is loaded ONCE per CTA and stays in shared memory
//     warps 1-2     MMA issuers (tiles 0,2 / 1,3 of an item; one thread issues one small MMA per ~50 cycles), accumulators
//                   double-buffered in TMEM so item i+1 is multiplied while item i is written out
//     warps 3-10    two epilogue groups that alternate on the accumulator buffers: tcgen05.ld -> bias / ReLU -> bf16 -> global
//   weight gradient (conv1d_wgrad_narrow): dw[k][co][ci] = sum_{b,l} dy[b,l,co] x[b,l+k-pad,ci].  A = x as an MN-major operand
//     with SBO = 16 B: M-group g is the plane shifted by g rows = tap g, so ONE M128 MMA covers 16 taps x 8 channels (the first
//     generation issued one MMA per tap with 7/8 of M idle and re-read the tensors once per 7 taps).  The accumulators
//     (tap group x channel plane x Cout columns) stay in TMEM for the whole kernel; each CTA flushes once.
// Reference counterpart: nn.Conv1d -> cuDNN (Module_3/tiny_ecg_model.py:17-20); grid of Module_2/benchmark_part_2.py:12-19.
#include "conv1d_common.cuh"

namespace cs {

constexpr int kNarrowThreads = 352;      // warp 0 TMA, warps 1-2 MMA, warps 3-10 epilogue (two groups of four)
constexpr int kNarrowMaxStages = 8;

// [outer, rows, C] bf16 viewed as (C, rows, outer) with C * 2 = 32 / 64 / 128 bytes per row: box = one whole row x box_rows,
// written to shared memory with the matching 32 / 64 / 128-byte swizzle
static int make_narrow_row_map(CUtensorMap* m, const void* base, int outer, int rows, int C, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    CS_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from this driver");
    cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)rows, (cuuint64_t)outer};
    cuuint64_t strides[2] = {(cuuint64_t)C * 2, (cuuint64_t)rows * C * 2};
    cuuint32_t box[3] = {(cuuint32_t)C, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    const CUtensorMapSwizzle sw = C == 16 ? CU_TENSOR_MAP_SWIZZLE_32B : (C == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B);
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CS_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (narrow rows) failed (%d) rows=%d C=%d box_rows=%d", (int)r, rows, C, box_rows);
    return CS_OK;
}

// K-major operand whose rows are exactly one swizzle span (32 / 64 / 128 bytes): 8-row groups `sbo` bytes apart
__device__ __forceinline__ uint64_t smem_desc_krow(uint32_t saddr, uint32_t sbo_bytes, uint32_t layout) {
    uint64_t d = (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)layout << 61;                            // 6: 32-byte, 4: 64-byte, 2: 128-byte swizzle
    return d;
}

extern long long* g_tiny_prof;   // cs_debug_set_tiny_profile (tiny_ecg_umma.cu)

struct NarrowFwd {
    long long* prof;   // debug timeline of CTA 0 (null in production)
    int B, L, Lout, Cin, Cout, K, pad, relu;
    int R;             // rows of one activation tile = 128 + K - 1
    int MT;            // 128-position tiles per work item (<= 4); tiles are numbered b * tiles + t, an item is MT consecutive
                       // ones (so short packed samples share an item and both MMA warps have work)
    int units;         // B * tiles
    int items;         // ceil(units / MT)
    int tiles;         // 128-position tiles per sample
    int NS;            // ring depth
    int a_tile_stride; // bytes of one tile in the ring (1024-aligned: the swizzles act on address bits 7..9)
    int w_tap_bytes;   // Cout * Cin * 2, 1024-aligned
    int w_bytes;       // K taps
    int stage_bytes;
};

// tile 0 of work item `item` as a global tile number b * tiles + t, and the number of tiles in the item
__device__ __forceinline__ void narrow_item(const NarrowFwd& s, int item, int& first, int& mt) {
    first = item * s.MT;
    mt = s.units - first < s.MT ? s.units - first : s.MT;
}

// Operand rows are whole positions: [row = position][Cin x bf16 = 32 / 64 / 128 B], written by TMA with the swizzle of that
// width: ONE TMA row per position (the plane layout needs Cin / 8 rows of 16 bytes).  What one SM's TMA unit delivers from
// cold DRAM grows with the bytes per load and does not grow with the ring depth (bench/tma_stream_probe.py, 128-row boxes:
// 0.93 / 1.82 / 3.35 / 4.5 TB/s over the GPU for 16 / 32 / 64 / 128-byte rows; 1-D bulk copies of 4 / 16 KB: 1.9 / 4.5 TB/s),
// and C = 16 at K = 3 (37 us for 65.5 MB in = 1.8 TB/s) sits on the 32-byte-row figure; loading the same rows as 256-row
// boxes was tried and changed nothing.  The channel-packed path below gets 128-byte rows.  A tap is a start-address shift of one
// row; the swizzle is a function of the absolute address, so shifted descriptors stay consistent.  One MMA's accumulate is
// ~130 cycles behind its issue when the next MMA targets the same accumulator (timeline of CTA 0, debug hook `prof`), so
// each issuing warp interleaves its two tiles.
__global__ void __launch_bounds__(kNarrowThreads, 1)
conv1d_fwd_narrow_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                         const float* __restrict__ bias, __nv_bfloat16* __restrict__ y, NarrowFwd s) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    uint64_t* full = bars;                                  // [NS]   TMA -> MMA
    uint64_t* empty = bars + kNarrowMaxStages;              // [NS]   MMA -> TMA (tcgen05.commit, one per issuing warp)
    uint64_t* acc_full = bars + 2 * kNarrowMaxStages;       // [2]    MMA -> epilogue
    uint64_t* acc_empty = acc_full + 2;                     // [2]    epilogue -> MMA (4 warp arrivals)
    uint64_t* w_full = acc_empty + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_full + 1);
    float* bias_s = reinterpret_cast<float*>(smem + 256);      // [64]
    uint8_t* w_smem = smem + 1024;
    uint8_t* a_ring = w_smem + s.w_bytes;
    const int stage_bytes = s.stage_bytes;
    const int buf_cols = s.MT * s.Cout;
    uint32_t tmem_cols = 32;
    while ((int)tmem_cols < 2 * buf_cols) tmem_cols <<= 1;
    const int my_items = (s.items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const int row_bytes = s.Cin * 2;

    if (threadIdx.x == 0) {
        for (int i = 0; i < s.NS; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 2); }
        for (int i = 0; i < 2; ++i) { mbar_init(acc_full + i, 2); mbar_init(acc_empty + i, 4); }
        mbar_init(w_full, 1);
        fence_mbar_init();
        tma_prefetch_desc(&map_w); tma_prefetch_desc(&map_x);
    }
    if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
    if (threadIdx.x >= 64 && threadIdx.x < 128) bias_s[threadIdx.x - 64] = (bias != nullptr && (int)threadIdx.x - 64 < s.Cout) ? bias[threadIdx.x - 64] : 0.0f;
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            mbar_arrive_expect_tx(w_full, s.K * s.Cout * row_bytes);
            for (int k = 0; k < s.K; ++k) tma_load_3d(w_smem + k * s.w_tap_bytes, &map_w, w_full, 0, 0, k);
            for (int it = 0; it < my_items; ++it) {
                int u0, mt;
                narrow_item(s, blockIdx.x + it * gridDim.x, u0, mt);
                const int sa = it % s.NS;
                if (it >= s.NS) mbar_wait(empty + sa, ((it / s.NS) - 1) & 1);
                if (s.prof && blockIdx.x == 0 && it < 60) s.prof[it * 4] = clock64();
                mbar_arrive_expect_tx(full + sa, s.R * row_bytes * mt);
                for (int i = 0; i < mt; ++i) {
                    const int b = (u0 + i) / s.tiles, t = (u0 + i) - b * s.tiles;
                    tma_load_3d(a_ring + sa * stage_bytes + i * s.a_tile_stride, &map_x, full + sa, 0, t * 128 - s.pad, b);
                }
            }
        }
    } else if (warp <= 2) {
        // ===== MMA issuers: warp 1 tiles 0, 2; warp 2 tiles 1, 3 =====
        if (elect_one()) {
            const int m = warp - 1;
            const uint32_t idesc = idesc_f16(FMT_BF16, 128, s.Cout, MAJOR_K, MAJOR_K);
            const uint32_t layout = s.Cin == 16 ? 6u : (s.Cin == 32 ? 4u : 2u);
            const uint32_t sbo = 8 * row_bytes;
            mbar_wait(w_full, 0);
            for (int it = 0; it < my_items; ++it) {
                int u0, mt;
                narrow_item(s, blockIdx.x + it * gridDim.x, u0, mt);
                const int sa = it % s.NS, buf = it & 1;
                mbar_wait(full + sa, (it / s.NS) & 1);
                if (s.prof && blockIdx.x == 0 && it < 60 && m == 0) s.prof[it * 4 + 1] = clock64();
                if (it >= 2) mbar_wait(acc_empty + buf, ((it >> 1) - 1) & 1);
                if (s.prof && blockIdx.x == 0 && it < 60 && m == 0) s.prof[it * 4 + 2] = clock64();
                tc_fence_after_sync();
                const uint32_t a_base = smem_u32(a_ring + sa * stage_bytes), w_base = smem_u32(w_smem);
                // this warp's tiles are independent accumulators: interleave them so that the tensor pipe always has a second
                // chain to work on while one MMA's accumulate is still in flight
                const bool has0 = m < mt, has1 = m + 2 < mt;
                const uint32_t d0 = tmem_base + buf * buf_cols + m * s.Cout, d1 = d0 + 2 * s.Cout;
                const uint64_t a0 = smem_desc_krow(a_base + m * s.a_tile_stride, sbo, layout);
                const uint64_t a1 = smem_desc_krow(a_base + (m + 2) * s.a_tile_stride, sbo, layout);
                const uint64_t b0 = smem_desc_krow(w_base, sbo, layout);
                const int nj = s.Cin / 16;                   // 16 channels = 32 bytes inside the row
#pragma unroll 2
                for (int k = 0; k < s.K; ++k) {
                    const uint32_t ak = (uint32_t)(k * row_bytes) >> 4, bk = (uint32_t)(k * s.w_tap_bytes) >> 4;
#pragma unroll 4
                    for (int j = 0; j < nj; ++j) {
                        const uint32_t acc = (k | j) ? 1u : 0u;
                        if (has0) mma_f16(d0, a0 + (uint64_t)(ak + 2 * j), b0 + (uint64_t)(bk + 2 * j), idesc, acc);
                        if (has1) mma_f16(d1, a1 + (uint64_t)(ak + 2 * j), b0 + (uint64_t)(bk + 2 * j), idesc, acc);
                    }
                }
                mma_commit(empty + sa);                     // (a warp without a tile in this item still signals: counts are fixed)
                mma_commit(acc_full + buf);
            }
        }
    } else {
        // ===== epilogue groups: group e takes the items with (it & 1) == e =====
        const int e = (warp - 3) >> 2, q = warp & 3;
        const int r = q * 32 + lane;
        for (int it = e; it < my_items; it += 2) {
            int u0, mt;
            narrow_item(s, blockIdx.x + it * gridDim.x, u0, mt);
            mbar_wait(acc_full + e, (it >> 1) & 1);
            if (s.prof && blockIdx.x == 0 && it < 60 && q == 0 && lane == 0) s.prof[it * 4 + 3] = clock64();
            tc_fence_after_sync();
            // up to 64 accumulator columns go to registers at once; the TMEM buffer is handed back to the MMA warps as soon as the
            // last group has been loaded, BEFORE the conversion and the stores
            for (int cb = 0; cb < buf_cols; cb += 64) {
                uint32_t v[4][16];
                const int nch = (buf_cols - cb) >= 64 ? 4 : (buf_cols - cb) >> 4;
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (u < nch) tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + e * buf_cols + cb + u * 16, v[u]);
                tmem_ld_wait();
                if (cb + 64 >= buf_cols) {
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(acc_empty + e);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (u < nch) {
                        const int col = cb + u * 16, i = col / s.Cout, c0 = col - i * s.Cout;
                        const int b = (u0 + i) / s.tiles, l = ((u0 + i) - b * s.tiles) * 128 + r;
                        if (i < mt && l < s.Lout) {
                            uint32_t pk[8];
#pragma unroll
                            for (int j = 0; j < 16; j += 2) {
                                float f0 = __uint_as_float(v[u][j]) + bias_s[c0 + j], f1 = __uint_as_float(v[u][j + 1]) + bias_s[c0 + j + 1];
                                if (s.relu) { f0 = fmaxf(f0, 0.0f); f1 = fmaxf(f1, 0.0f); }
                                __nv_bfloat162 h = __floats2bfloat162_rn(f0, f1);
                                pk[j >> 1] = *reinterpret_cast<uint32_t*>(&h);
                            }
                            uint4* d4 = reinterpret_cast<uint4*>(y + ((size_t)b * s.Lout + l) * s.Cout + c0);
                            d4[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                            d4[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                        }
                    }
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

static int fwd_narrow_impl(const void* x, const void* w, const float* bias, void* y, int B, int L, int Cin, int Cout, int K, int pad,
                           int relu, int Lout, cudaStream_t st, bool packed) {
    if (env_int("CS_CONV_NARROW", 1) == 0) return -1;
    // rows of one swizzle span; plain 64-channel layers belong to conv1d_sw128.cu, packed narrow layers come here as 64 x 64
    if (!((Cin == 16 || Cin == 32) || (packed && Cin == 64)) || Cout > 64 || Cout % 16) return -1;
    if (Lout + K - 1 <= 64) return -1;                      // short windows: the sample-packing mode of the first generation
    NarrowFwd s;
    memset(&s, 0, sizeof(s));
    s.prof = g_tiny_prof;
    s.B = B; s.L = L; s.Lout = Lout; s.Cin = Cin; s.Cout = Cout; s.K = K; s.pad = pad; s.relu = relu;
    s.R = 128 + K - 1;
    s.w_tap_bytes = (Cout * Cin * 2 + 1023) & ~1023;
    s.w_bytes = K * s.w_tap_bytes;
    if (s.w_bytes > 72 * 1024) return -1;                   // the whole weight tensor stays resident
    s.tiles = (Lout + 127) / 128;
    s.units = B * s.tiles;
    const size_t budget = 200 * 1024 - 1024 - s.w_bytes;
    const int row_bytes = Cin * 2;
    auto stage_of = [&](int mt) -> size_t { return (size_t)((s.R * row_bytes + 1023) & ~1023) * mt; };
    s.MT = s.units >= 4 ? 4 : s.units;
    // two tiles per MMA warp (two accumulate chains in flight) matter more than ring depth for the many-tap shapes
    while (s.MT > 1 && (2 * s.MT * Cout > 512 || budget / stage_of(s.MT) < 3)) s.MT >>= 1;
    s.a_tile_stride = (s.R * row_bytes + 1023) & ~1023;
    s.items = (s.units + s.MT - 1) / s.MT;
    const size_t stage = stage_of(s.MT);
    s.stage_bytes = (int)stage;
    s.NS = (int)(budget / stage);
    if (s.NS > kNarrowMaxStages) s.NS = kNarrowMaxStages;
    if (s.NS < 3) return -1;
    const size_t smem = 1024 + s.w_bytes + (size_t)s.NS * stage;
    CUtensorMap mw, mx;
    int rc = make_narrow_row_map(&mw, w, K, Cout, Cin, Cout);
    if (rc) return rc;
    rc = make_narrow_row_map(&mx, x, B, L, Cin, s.R);
    if (rc) return rc;
    static size_t configured[64] = {};
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    if (smem > configured[dev & 63]) {
        CS_CUDA_OK(cudaFuncSetAttribute(conv1d_fwd_narrow_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev & 63] = smem;
    }
    int grid = num_sms();
    if (grid > s.items) grid = s.items;
    conv1d_fwd_narrow_kernel<<<grid, kNarrowThreads, smem, st>>>(mx, mw, bias, (__nv_bfloat16*)y, s);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

int conv1d_fwd_narrow(const void* x, const void* w, const float* bias, void* y, int B, int L, int Cin, int Cout, int K, int pad,
                      int relu, int Lout, cudaStream_t st) {
    return fwd_narrow_impl(x, w, bias, y, B, L, Cin, Cout, K, pad, relu, Lout, st, false);
}

// ------------------------------------------------------------------------------------------------------------------
// weight gradient, narrow layers
// ------------------------------------------------------------------------------------------------------------------
constexpr int kNarrowWgThreads = 224;    // warp 0 TMA, warps 1-2 MMA, warps 3-6 flush
constexpr int kNarrowWgStages = 8;

struct NarrowWg {
    int B, L, Lout, Cin, Cout, K, pad;
    int TG;            // tap groups of 16
    int RX;            // rows of one x tile = 128 + 16 * TG - 1 (taps past K read finite junk that is never flushed)
    int tiles;         // 128-position tiles per sample
    int total;         // B * tiles
    int NS;
    int dy_bytes, x_bytes, stage_bytes;
};

__global__ void __launch_bounds__(kNarrowWgThreads, 1)
conv1d_wgrad_narrow_kernel(const __grid_constant__ CUtensorMap map_dy, const __grid_constant__ CUtensorMap map_x, float* __restrict__ dw, NarrowWg s) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    uint64_t* full = bars, *empty = bars + kNarrowWgStages, *acc_full = bars + 2 * kNarrowWgStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);
    uint8_t* ring = smem + 256;
    const int Pi = s.Cin / 8, nacc = s.TG * Pi;
    uint32_t tmem_cols = 32;
    while ((int)tmem_cols < nacc * s.Cout) tmem_cols <<= 1;
    const int t_begin = (int)(((long long)s.total * blockIdx.x) / gridDim.x);
    const int t_end = (int)(((long long)s.total * (blockIdx.x + 1)) / gridDim.x);
    const int n_iter = t_end - t_begin;

    if (threadIdx.x == 0) {
        for (int i = 0; i < s.NS; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 2); }
        mbar_init(acc_full, 2);
        tma_prefetch_desc(&map_dy); tma_prefetch_desc(&map_x);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer: dy tile (128 rows) and x tile (RX rows, starting pad rows earlier) per 128 positions =====
        if (elect_one()) {
            for (int it = 0; it < n_iter; ++it) {
                const int tile = t_begin + it, b = tile / s.tiles, l0 = (tile % s.tiles) * 128;
                const int sg = it % s.NS;
                if (it >= s.NS) mbar_wait(empty + sg, ((it / s.NS) - 1) & 1);
                mbar_arrive_expect_tx(full + sg, s.dy_bytes + s.x_bytes);
                tma_load_4d(ring + sg * s.stage_bytes, &map_dy, full + sg, 0, l0, 0, b);
                tma_load_4d(ring + sg * s.stage_bytes + s.dy_bytes, &map_x, full + sg, 0, l0 - s.pad, 0, b);
            }
        }
    } else if (warp <= 2) {
        // accumulator a = (tap group, x plane): D[(g, ci8), co] += x[l + 16 tg + g - pad, 8 p + ci8] * dy[l, co]; warp 1 takes the
        // even accumulators, warp 2 the odd ones
        if (elect_one()) {
            const int m = warp - 1;
            const uint32_t idesc = idesc_f16(FMT_BF16, 128, s.Cout, MAJOR_MN, MAJOR_MN);
            for (int it = 0; it < n_iter; ++it) {
                const int sg = it % s.NS;
                mbar_wait(full + sg, (it / s.NS) & 1);
                tc_fence_after_sync();
                const uint32_t dy_base = smem_u32(ring + sg * s.stage_bytes), x_base = dy_base + s.dy_bytes;
                for (int a = m; a < nacc; a += 2) {
                    const int tg = a / Pi, p = a % Pi;
                    const uint64_t a0 = smem_desc(x_base + p * s.RX * 16 + tg * 16 * 16, 128, 16);
                    const uint64_t b0 = smem_desc(dy_base, 128, 128 * 16);
#pragma unroll
                    for (int c = 0; c < 8; ++c)              // 8 chunks of 16 positions
                        mma_f16(tmem_base + a * s.Cout, a0 + (uint64_t)(c * 16), b0 + (uint64_t)(c * 16), idesc, (it > 0 || c > 0) ? 1u : 0u);
                }
                mma_commit(empty + sg);
            }
            mma_commit(acc_full);
        }
    } else if (n_iter > 0) {
        const int q = warp & 3;
        const int r = q * 32 + lane, g = r >> 3, c8 = r & 7;
        mbar_wait(acc_full, 0);
        tc_fence_after_sync();
        for (int a0 = 0; a0 < nacc; ++a0) {
            const int a = (a0 + (int)blockIdx.x) % nacc;     // CTAs start at different accumulators: fewer same-address collisions
            const int tg = a / Pi, p = a % Pi, k = tg * 16 + g;
            for (int c0 = 0; c0 < s.Cout; c0 += 16) {
                uint32_t v[16];
                tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + a * s.Cout + c0, v);
                tmem_ld_wait();
                if (k < s.K) {
                    float* dst = dw + ((size_t)k * s.Cout + c0) * s.Cin + p * 8 + c8;
#pragma unroll
                    for (int j = 0; j < 16; ++j) atomicAdd(dst + (size_t)j * s.Cin, __uint_as_float(v[j]));
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

int conv1d_wgrad_narrow(const void* x, const void* dy, float* dw, int B, int L, int Cin, int Cout, int K, int pad, int Lout,
                        cudaStream_t st) {
    if (env_int("CS_CONV_NARROW", 1) == 0) return -1;
    if (Cin >= 64 && Cout >= 64) return -1;
    if (Cin > 64 || Cout > 64 || Cin % 8 || Cout % 16) return -1;
    if (Lout + K - 1 <= 64) return -1;                      // short windows: sample packing of the first generation
    NarrowWg s;
    memset(&s, 0, sizeof(s));
    s.B = B; s.L = L; s.Lout = Lout; s.Cin = Cin; s.Cout = Cout; s.K = K; s.pad = pad;
    s.TG = (K + 15) / 16;
    if (s.TG * (Cin / 8) * Cout > 512) return -1;           // every (tap group, plane) accumulator lives in TMEM
    s.RX = 128 + 16 * s.TG - 1;
    s.tiles = (Lout + 127) / 128; s.total = B * s.tiles;
    s.dy_bytes = 128 * 16 * (Cout / 8);
    s.x_bytes = s.RX * 16 * (Cin / 8);
    s.stage_bytes = (s.dy_bytes + s.x_bytes + 127) & ~127;
    s.NS = (200 * 1024) / s.stage_bytes;
    if (s.NS > kNarrowWgStages) s.NS = kNarrowWgStages;
    if (s.NS < 3) return -1;
    const size_t smem = 256 + (size_t)s.NS * s.stage_bytes + 512;   // slack: the last M-groups of the last plane read past the tile
    CUtensorMap mdy, mx;
    int rc = make_plane_map(&mdy, dy, B, Lout, Cout, 128, Cout / 8);
    if (rc) return rc;
    rc = make_plane_map(&mx, x, B, L, Cin, s.RX, Cin / 8);
    if (rc) return rc;
    static size_t configured[64] = {};
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    if (smem > configured[dev & 63]) {
        CS_CUDA_OK(cudaFuncSetAttribute(conv1d_wgrad_narrow_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev & 63] = smem;
    }
    int grid = num_sms();
    if (grid > s.total) grid = s.total;
    conv1d_wgrad_narrow_kernel<<<grid, kNarrowWgThreads, smem, st>>>(mdy, mx, dw, s);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// channel-packed 128-byte rows: a narrow layer as a 64-channel layer on L / f packed positions
// ------------------------------------------------------------------------------------------------------------------
// x [B, L, C] with C = 16 (32) is the same memory as xp [B, L/4, 64] ([B, L/2, 64]): f consecutive positions per 128-byte
// row.  The convolution becomes  yp[m', (r', co)] = sum_{dm, (r, ci)} Wp[dm][(r', co)][(r, ci)] xp[m' + dm, (r, ci)]  with
// Wp[dm][(r', co)][(r, ci)] = W[k = f dm + r - r' + pad][co][ci] (zero outside 0 <= k < K): K' = O(K / f + 2) packed taps of
// a 64-channel layer, which the 128-byte-row kernels (conv1d_sw128.cu: TMA rows of 128 B instead of 16-byte plane rows,
// K = 64 reduction per MMA instead of 16) run near the copy roofline.  The extra arithmetic on the zero blocks is 4x at
// K = 3 and 1.16x at K = 31 - free for a layer that moves 131 MB per 1.6 GFLOP.  dgrad goes the same way (it is a forward
// convolution on dy); the weight gradient is computed packed and folded back: dW[k] = sum_{r'} dWp[dm(k, r')][r'][r(k, r')].
struct PackGeom { int f, Kp, padp, dm_min; };
static bool pack_geom(int L, int Lout, int Cin, int Cout, int K, int pad, PackGeom& g) {
    if (Cin != 16 && Cin != 32) return false;
    g.f = 64 / Cin;
    if (Cout % 16 || Cout > 64 || L % g.f || Lout % g.f) return false;
    g.dm_min = -((pad + g.f - 1) / g.f);
    const int dm_max = (g.f - 1 + K - 1 - pad) / g.f;
    g.Kp = dm_max - g.dm_min + 1;
    g.padp = -g.dm_min;
    return Lout / g.f <= L / g.f + 2 * g.padp - g.Kp + 1;   // every packed output row is one the packed convolution produces
}

__global__ void pack_weight_kernel(const __nv_bfloat16* __restrict__ w, const float* __restrict__ bias, __nv_bfloat16* __restrict__ wp,
                                   float* __restrict__ bp, int K, int Cout, int Cin, int f, int pad, int dm_min, int Kp) {
    const int Cop = f * Cout, Cip = f * Cin;
    const int total = Kp * Cop * Cip;
    const int i0 = blockIdx.x * blockDim.x + threadIdx.x;
    for (int i = i0; i < total; i += gridDim.x * blockDim.x) {
        const int cip = i % Cip, cop = (i / Cip) % Cop, dmi = i / (Cip * Cop);
        const int r = cip / Cin, ci = cip - r * Cin, rp = cop / Cout, co = cop - rp * Cout;
        const int k = f * (dmi + dm_min) + r - rp + pad;
        wp[i] = (k >= 0 && k < K) ? w[((size_t)k * Cout + co) * Cin + ci] : __float2bfloat16(0.0f);
    }
    if (bp != nullptr && i0 < Cop) bp[i0] = bias[i0 % Cout];
}

// dw[k][co][ci] += sum_{r'} dwp[dm - dm_min][r' Cout + co][r Cin + ci],  f dm + r = k - pad + r'
__global__ void unpack_wgrad_kernel(const float* __restrict__ dwp, float* __restrict__ dw, int K, int Cout, int Cin, int f, int pad, int dm_min) {
    const int Cop = f * Cout, Cip = f * Cin;
    const int total = K * Cout * Cin;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int ci = i % Cin, co = (i / Cin) % Cout, k = i / (Cin * Cout);
        float acc = 0.0f;
        for (int rp = 0; rp < f; ++rp) {
            const int j = k - pad + rp;
            const int dm = j >= 0 ? j / f : -((-j + f - 1) / f);
            const int r = j - f * dm;
            acc += dwp[((size_t)(dm - dm_min) * Cop + rp * Cout + co) * Cip + r * Cin + ci];
        }
        dw[i] += acc;
    }
}

// packed weights / packed weight-gradient scratch: one allocation per device, made at the first call (stream-ordered
// pool allocations cost ~150 us per call when the pool trims itself, measured).  Calls on ONE stream are ordered; layers
// run on different streams at the same time would have to use separate scratch - the engine runs all convolutions of a
// model on the compute stream.
constexpr size_t kPackScratchBytes = 2u << 20;
static uint8_t* pack_scratch() {
    static uint8_t* buf[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    if (!buf[dev & 63]) {
        if (cudaMalloc(&buf[dev & 63], kPackScratchBytes) != cudaSuccess) { (void)cudaGetLastError(); buf[dev & 63] = nullptr; return nullptr; }
    }
    return buf[dev & 63];
}

// Measured (B200, L = 500): packing wins where the tap count, not the row count, sets the time - C = 16 from K = 5
// (K = 31, B = 4096: 56 us vs 119 us), C = 32 from K = 16; at K = 3 the row kernel is ahead (37 vs 44 us), and below ~64 K
// positions the extra weight-packing launch costs more than it saves.
static bool packed_worth_it(int B, int L, int Cin, int K) {
    if (env_int("CS_CONV_PACK_MIN", -1) == 0) return true;                       // tests force the path
    return (long long)B * L >= 65536 && (Cin == 16 ? K >= 5 : K >= 16);
}

int conv1d_fwd_packed(const void* x, const void* w, const float* bias, void* y, int B, int L, int Cin, int Cout, int K, int pad,
                      int relu, int Lout, cudaStream_t st) {
    PackGeom g;
    if (!env_int("CS_CONV_PACK", 1) || !packed_worth_it(B, L, Cin, K) || !pack_geom(L, Lout, Cin, Cout, K, pad, g)) return -1;
    const int Cop = g.f * Cout, Cip = g.f * Cin;
    const size_t w_bytes = ((size_t)g.Kp * Cop * Cip * 2 + 255) & ~(size_t)255;
    uint8_t* buf = pack_scratch();
    if (buf == nullptr || w_bytes + (size_t)Cop * 4 > (1u << 20)) return -1;
    float* bp = bias ? reinterpret_cast<float*>(buf + w_bytes) : nullptr;
    const int total = g.Kp * Cop * Cip;
    pack_weight_kernel<<<(total + 255) / 256, 256, 0, st>>>((const __nv_bfloat16*)w, bias, (__nv_bfloat16*)buf, bp, K, Cout, Cin, g.f, pad, g.dm_min, g.Kp);
    CS_CUDA_OK(cudaGetLastError());
    if (env_int("CS_CONV_PACK_PERSISTENT", 1)) {             // the persistent pipelined kernel on 128-byte rows
        const int rc = fwd_narrow_impl(x, buf, bp, y, B, L / g.f, Cip, Cop, g.Kp, g.padp, relu, Lout / g.f, st, true);
        if (rc >= 0) return rc;
    }
    return conv1d_fwd_sw128(x, buf, bp, y, B, L / g.f, Cip, Cop, g.Kp, g.padp, relu, Lout / g.f, st);
}

int conv1d_wgrad_packed(const void* x, const void* dy, float* dw, int B, int L, int Cin, int Cout, int K, int pad, int Lout, cudaStream_t st) {
    PackGeom g;
    if (!env_int("CS_CONV_PACK", 1) || !packed_worth_it(B, L, Cin, K) || !pack_geom(L, Lout, Cin, Cout, K, pad, g)) return -1;
    const int Cop = g.f * Cout, Cip = g.f * Cin;
    const size_t bytes = (size_t)g.Kp * Cop * Cip * 4;
    uint8_t* buf = pack_scratch();
    if (buf == nullptr || bytes + (1u << 20) > kPackScratchBytes) return -1;
    float* dwp = reinterpret_cast<float*>(buf + (1u << 20));       // (the first MiB holds packed weights of a forward / dgrad call)
    CS_CUDA_OK(cudaMemsetAsync(dwp, 0, bytes, st));
    int rc = conv1d_wgrad_sw128(x, dy, dwp, B, L / g.f, Cip, Cop, g.Kp, g.padp, Lout / g.f, st);
    if (rc == 0) {
        const int total = K * Cout * Cin;
        unpack_wgrad_kernel<<<(total + 255) / 256, 256, 0, st>>>(dwp, dw, K, Cout, Cin, g.f, pad, g.dm_min);
        CS_CUDA_OK(cudaGetLastError());
    }
    return rc;
}

}  // namespace cs
