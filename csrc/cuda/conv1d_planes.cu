// conv1d on tcgen05, first generation: 16-byte "plane" operand rows (K1-K3 for narrow layers), the public
// dispatchers, the Cin = 1 FFMA kernels and small helpers.  The wide-layer kernels are in conv1d_sw128.cu.
//
//   x [B, L, Cin]   w [K, Cout, Cin] (tap-major)   y [B, Lout, Cout]   stride 1, zero padding, channels-last bf16
//
// Operand staging: ONE TMA (cp.async.bulk.tensor, 4-D map) per tile writes the activation into
// shared memory as "planes" — [plane = 8 channels][row = position][8 x bf16 = 16 B] — by declaring the
// tensor-map dimensions in the order (c8, position, plane, batch).  That is the no-swizzle UMMA
// core-matrix layout, in which
//   * a plane pair is a K-major A operand (rows = positions = M, SBO = 128 B, LBO = plane stride) and
//     a convolution TAP IS A START-ADDRESS SHIFT of tap*16 bytes: the tile (128 + K - 1 rows) is loaded
//     once and reused by all K taps; TMA out-of-bounds zero fill provides the padding and the
//     isolation between samples;
//   * the same planes are MN-major operands for the weight gradient (rows = positions = reduction).
// Accumulators live in TMEM; one elected thread issues tcgen05.mma; tcgen05.commit -> mbarrier recycles
// the TMA ring slots; the epilogue warps read TMEM with tcgen05.ld and fuse bias + ReLU + bf16 cast.
// Replaces nn.Conv1d -> cuDNN of the reference (Module_3/tiny_ecg_model.py:17-20) for the wide models.
#include "conv1d_common.cuh"

namespace cs {

// ------------------------------------------------------------------------------------------------
// tensor maps (driver entry point fetched through the runtime: no -lcuda needed)
// ------------------------------------------------------------------------------------------------
// (make_plane_map / tma_load_4d: conv1d_common.cuh, shared with conv1d_narrow.cu)

// ------------------------------------------------------------------------------------------------
// K1 forward (dgrad reuses it with flipped / transposed weights)
// CTA = (n tile, group of MT 128-position tiles, sample).  192 threads: warp 0 TMA, warp 1 MMA, warps 2-5
// epilogue.  Every weight stage (one tap of one 64-channel chunk) is reused by all MT position tiles, which
// cuts the L2->SM weight traffic per FLOP by MT (the single-tile version is L2-bandwidth bound at C = 256).
// ------------------------------------------------------------------------------------------------
constexpr int kAStages = 2, kWStages = 4;

struct FwdShape {
    int B, L, Lout, Cin, Cout, K, pad, relu;
    int CK;          // channels per reduction chunk (16 / 32 / 64)
    int NT;          // output channels per CTA (16 .. 128)
    int R;           // rows per activation tile = 128 + K - 1
    int MT;          // 128-position tiles per CTA (1, 2 or 4): weights are streamed once for all of them
    int slot;        // 0: a tile is 128 positions of one sample.  > 0 (16/32/64): short windows (L + K - 1 <= slot):
                     // a tile packs 128/slot SAMPLES, each in its own `slot` rows (windows never straddle slots)
};

__global__ void __launch_bounds__(kFwdThreads, 1)
conv1d_fwd_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                  const __grid_constant__ CUtensorMap map_xs, const float* __restrict__ bias, __nv_bfloat16* __restrict__ y, FwdShape s) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * s.NT, l0 = blockIdx.y * 128 * s.MT, b = blockIdx.z;
    const int a_tile = s.R * 16 * (s.CK / 8);               // one position tile of a chunk: CK/8 planes of R rows
    const int a_tile_stride = (a_tile + 127) & ~127;
    const int a_bytes = a_tile_stride * s.MT;               // ring slot: MT tiles
    const int w_bytes = s.NT * 16 * (s.CK / 8);             // one tap of weights for that chunk
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    uint64_t* full_a = bars, *empty_a = bars + kAStages, *full_w = bars + 2 * kAStages, *empty_w = full_w + kWStages;
    uint64_t* acc_full = empty_w + kWStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);
    uint8_t* a_ring = smem + 256;
    uint8_t* w_ring = a_ring + kAStages * a_bytes;
    const int a_stride = a_bytes, w_stride = (w_bytes + 127) & ~127;
    const int chunks = s.Cin / s.CK;
    uint32_t tmem_cols = 32;
    while ((int)tmem_cols < s.NT * s.MT) tmem_cols <<= 1;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kAStages; ++i) { mbar_init(full_a + i, 1); mbar_init(empty_a + i, 1); }
        for (int i = 0; i < kWStages; ++i) { mbar_init(full_w + i, 1); mbar_init(empty_w + i, 1); }
        mbar_init(acc_full, 1);
        fence_mbar_init();
        tma_prefetch_desc(&map_x); tma_prefetch_desc(&map_w); tma_prefetch_desc(&map_xs);
    }
    if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const int spt = s.slot ? 128 / s.slot : 1;              // samples per tile in packed mode

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            int wi = 0;
            for (int c = 0; c < chunks; ++c) {
                const int sa = c % kAStages;
                if (c >= kAStages) mbar_wait(empty_a + sa, ((c / kAStages) - 1) & 1);
                if (!s.slot) {
                    mbar_arrive_expect_tx(full_a + sa, a_tile * s.MT);
                    for (int i = 0; i < s.MT; ++i)
                        tma_load_4d(a_ring + sa * a_stride + i * a_tile_stride, &map_x, full_a + sa, 0, l0 + i * 128 - s.pad,
                                    c * (s.CK / 8), b);
                } else {
                    // packed short windows: one (slot rows x 1 plane) box per (sample, plane), dropped at its slot inside the
                    // plane-major tile; samples past the batch are fetched out of bounds (zero fill)
                    const int planes = s.CK / 8;
                    mbar_arrive_expect_tx(full_a + sa, s.MT * 128 * 16 * planes);
                    for (int i = 0; i < s.MT; ++i)
                        for (int j = 0; j < spt; ++j)
                            for (int p = 0; p < planes; ++p)
                                tma_load_4d(a_ring + sa * a_stride + i * a_tile_stride + p * s.R * 16 + j * s.slot * 16, &map_xs,
                                            full_a + sa, 0, -s.pad, c * planes + p, (b * s.MT + i) * spt + j);
                }
                for (int k = 0; k < s.K; ++k, ++wi) {
                    const int sw = wi % kWStages;
                    if (wi >= kWStages) mbar_wait(empty_w + sw, ((wi / kWStages) - 1) & 1);
                    mbar_arrive_expect_tx(full_w + sw, w_bytes);
                    tma_load_4d(w_ring + sw * w_stride, &map_w, full_w + sw, 0, n0, c * (s.CK / 8), k);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (elect_one()) {
            const uint32_t idesc = idesc_f16(FMT_BF16, 128, s.NT, MAJOR_K, MAJOR_K);
            const uint32_t plane_a = s.R * 16, plane_w = s.NT * 16;
            int wi = 0;
            uint32_t first = 1;
            for (int c = 0; c < chunks; ++c) {
                const int sa = c % kAStages;
                mbar_wait(full_a + sa, (c / kAStages) & 1);
                tc_fence_after_sync();
                const uint64_t a0 = smem_desc(smem_u32(a_ring + sa * a_stride), plane_a, 128);
                for (int k = 0; k < s.K; ++k, ++wi) {
                    const int sw = wi % kWStages;
                    mbar_wait(full_w + sw, (wi / kWStages) & 1);
                    tc_fence_after_sync();
                    const uint64_t b0 = smem_desc(smem_u32(w_ring + sw * w_stride), plane_w, 128);
                    for (int j = 0; j < s.CK / 16; ++j) {
                        // 16 channels = planes 2j, 2j+1; tap k = k rows further down the tile
                        const uint64_t bj = b0 + (uint64_t)((2 * j * plane_w) >> 4);
                        for (int i = 0; i < s.MT; ++i)
                            mma_f16(tmem_base + i * s.NT, a0 + (uint64_t)((i * a_tile_stride + 2 * j * plane_a) >> 4) + (uint64_t)k,
                                    bj, idesc, (first && j == 0) ? 0u : 1u);
                    }
                    first = 0;
                    mma_commit(empty_w + sw);            // weights slot free once these MMAs retire
                }
                mma_commit(empty_a + sa);
            }
            mma_commit(acc_full);
        }
    } else {
        // ===== epilogue: TMEM -> registers -> bias/ReLU -> bf16 -> global =====
        const int q = warp & 3;                           // TMEM sub-partition of this warp (warp id mod 4)
        mbar_wait(acc_full, 0);
        tc_fence_after_sync();
        for (int i = 0; i < s.MT; ++i) {
        const int r = q * 32 + lane;                       // accumulator row = TMEM lane
        int l = l0 + i * 128 + r, bb = b;
        if (s.slot) { l = r % s.slot; bb = (b * s.MT + i) * spt + r / s.slot; if (bb >= s.B) l = s.Lout; }
        __nv_bfloat16* dst = y + ((size_t)bb * s.Lout + l) * s.Cout + n0;
        for (int c0 = 0; c0 < s.NT; c0 += 16) {
            uint32_t v[16];
            tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + i * s.NT + c0, v);
            tmem_ld_wait();
            if (l < s.Lout) {
                uint32_t pk[8];
#pragma unroll
                for (int j = 0; j < 16; j += 2) {
                    float f0 = __uint_as_float(v[j]), f1 = __uint_as_float(v[j + 1]);
                    if (bias) { f0 += bias[n0 + c0 + j]; f1 += bias[n0 + c0 + j + 1]; }
                    if (s.relu) { f0 = fmaxf(f0, 0.0f); f1 = fmaxf(f1, 0.0f); }
                    __nv_bfloat162 h = __floats2bfloat162_rn(f0, f1);
                    pk[j >> 1] = *reinterpret_cast<uint32_t*>(&h);
                }
                uint4* d4 = reinterpret_cast<uint4*>(dst + c0);
                d4[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                d4[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            }
        }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

static int pick_ck(int C) { return C % 64 == 0 ? 64 : (C % 32 == 0 ? 32 : 16); }
static int pick_nt(int C) { return C % 128 == 0 ? 128 : (C % 64 == 0 ? 64 : (C % 32 == 0 ? 32 : 16)); }

int conv1d_fwd(const void* x, const void* w, const float* bias, void* y, int B, int L, int Cin, int Cout, int K, int pad,
               int stride, int relu, cudaStream_t st) {
    CS_REQUIRE(stride == 1, "conv1d: the tcgen05 path implements stride 1 (got %d); pool/subsample separately", stride);
    CS_REQUIRE(Cin % 16 == 0 && Cout % 16 == 0 && Cin >= 16 && Cout >= 16, "conv1d: Cin=%d / Cout=%d must be multiples of 16", Cin, Cout);
    CS_REQUIRE(K >= 1 && K <= 64 && pad >= 0 && pad < K, "conv1d: bad K=%d pad=%d", K, pad);
    CS_REQUIRE((((uintptr_t)x | (uintptr_t)w | (uintptr_t)y) & 15) == 0, "conv1d: pointers must be 16-byte aligned");
    FwdShape s;
    s.B = B; s.L = L; s.Cin = Cin; s.Cout = Cout; s.K = K; s.pad = pad; s.relu = relu;
    s.Lout = L + 2 * pad - K + 1;
    CS_REQUIRE(s.Lout >= 1, "conv1d: empty output");
    {
        const int rc128 = conv1d_fwd_sw128(x, w, bias, y, B, L, Cin, Cout, K, pad, relu, s.Lout, st);
        if (rc128 >= 0) return rc128;
        const int rcp = conv1d_fwd_packed(x, w, bias, y, B, L, Cin, Cout, K, pad, relu, s.Lout, st);
        if (rcp >= 0) return rcp;
        const int rcn = conv1d_fwd_narrow(x, w, bias, y, B, L, Cin, Cout, K, pad, relu, s.Lout, st);
        if (rcn >= 0) return rcn;
    }
    s.CK = pick_ck(Cin); s.NT = pick_nt(Cout); s.R = 128 + K - 1;
    CUtensorMap mx, mw;
    int rc = make_plane_map(&mx, x, B, L, Cin, s.R, s.CK / 8);
    if (rc) return rc;
    rc = make_plane_map(&mw, w, K, Cout, Cin, s.NT, s.CK / 8);
    if (rc) return rc;
    const int tiles = (s.Lout + 127) / 128;
    s.MT = tiles >= 4 ? 4 : tiles;                          // TMEM: MT * NT <= 512 columns
    // short windows: pack several samples into one 128-row tile (each in its own power-of-two slot)
    s.slot = 0;
    const int need = s.Lout + K - 1;                        // input rows (incl. padding) one sample's outputs read
    if (need <= 64) {
        s.slot = need <= 16 ? 16 : (need <= 32 ? 32 : 64);
        s.R = (128 + K - 1 + 7) & ~7;                       // per-plane TMA destinations must stay 128-byte aligned
        const int spt = 128 / s.slot, groups = (B + spt - 1) / spt;
        s.MT = groups >= 4 ? 4 : groups;
    }
    const int a_tile = (s.R * 16 * (s.CK / 8) + 127) & ~127, w_bytes = s.NT * 16 * (s.CK / 8);
    size_t smem = 256 + (size_t)kAStages * a_tile * s.MT + (size_t)kWStages * ((w_bytes + 127) & ~127);
    while (smem > 220 * 1024 && s.MT > 1) {                 // very wide taps: fewer tiles per CTA
        s.MT >>= 1;
        smem = 256 + (size_t)kAStages * a_tile * s.MT + (size_t)kWStages * ((w_bytes + 127) & ~127);
    }
    static size_t configured[64] = {};
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    if (smem > configured[dev & 63]) {
        CS_CUDA_OK(cudaFuncSetAttribute(conv1d_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev & 63] = smem;
    }
    CUtensorMap mxs = mx;
    dim3 grid(Cout / s.NT, (tiles + s.MT - 1) / s.MT, B);
    if (s.slot) {
        rc = make_plane_map(&mxs, x, B, L, Cin, s.slot, 1);
        if (rc) return rc;
        const int per_cta = s.MT * (128 / s.slot);
        grid = dim3(Cout / s.NT, 1, (B + per_cta - 1) / per_cta);
    }
    conv1d_fwd_kernel<<<grid, kFwdThreads, smem, st>>>(mx, mw, mxs, bias, (__nv_bfloat16*)y, s);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// w [K][Cout][Cin] -> wt [K][Cin][Cout] with taps flipped: wt[K-1-k][ci][co] = w[k][co][ci]
// (32x32 shared-memory tile transpose per tap: coalesced reads and writes)
__global__ void dgrad_weight_kernel(const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ wt, int K, int Cout, int Cin) {
    __shared__ __nv_bfloat16 tile[32][34];
    const int k = blockIdx.z, co0 = blockIdx.y * 32, ci0 = blockIdx.x * 32;
    const __nv_bfloat16* src = w + (size_t)k * Cout * Cin;
    __nv_bfloat16* dst = wt + (size_t)(K - 1 - k) * Cin * Cout;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int co = co0 + j, ci = ci0 + threadIdx.x;
        if (co < Cout && ci < Cin) tile[j][threadIdx.x] = src[(size_t)co * Cin + ci];
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int ci = ci0 + j, co = co0 + threadIdx.x;
        if (co < Cout && ci < Cin) dst[(size_t)ci * Cout + co] = tile[threadIdx.x][j];
    }
}
int conv1d_prepare_dgrad_weight(const void* w, void* wt, int K, int Cout, int Cin, cudaStream_t st) {
    dim3 grid((Cin + 31) / 32, (Cout + 31) / 32, K), block(32, 8);
    dgrad_weight_kernel<<<grid, block, 0, st>>>((const __nv_bfloat16*)w, (__nv_bfloat16*)wt, K, Cout, Cin);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// ------------------------------------------------------------------------------------------------
// K3 weight gradient:  dw[k, co, ci] += sum_{b,l} dy[b,l,co] * x[b, l+k-pad, ci]     (+ dbias)
// GEMM with M = co (128), N = ci (64), reduction over positions: both operands MN-major planes.
// CTA = (co tile, ci tile, split of the (b, l-tile) list); one TMEM accumulator per tap (<= 7 taps per
// pass); partial sums are added to the fp32 output with vector reductions.
// ------------------------------------------------------------------------------------------------
constexpr int kWgStages = 2;
constexpr int kWgTapsPerPass = 7;
constexpr int kWgNT = 64, kWgMT = 128;

struct WgShape { int B, L, Lout, Cin, Cout, K, pad, k0, ktaps, R, tiles_per_sample, total_tiles, my_cout, my_cin, slot; };

__global__ void __launch_bounds__(kWgThreads, 1)
conv1d_wgrad_kernel(const __grid_constant__ CUtensorMap map_dy, const __grid_constant__ CUtensorMap map_x,
                    const __grid_constant__ CUtensorMap map_dys, const __grid_constant__ CUtensorMap map_xs,
                    float* __restrict__ dw, float* __restrict__ dbias, WgShape s) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int co0 = blockIdx.x * kWgMT, ci0 = blockIdx.y * kWgNT;
    const int planes_dy = s.my_cout / 8, planes_x = s.my_cin / 8;
    const int dy_bytes = 128 * 16 * planes_dy, x_bytes = s.R * 16 * planes_x;
    const int dy_alloc = 128 * 16 * (kWgMT / 8);               // always room for 16 planes: M = 128 reads them all
    const int x_alloc = ((s.R * 16 * (kWgNT / 8)) + 127) & ~127;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    uint64_t* full = bars, *empty = bars + kWgStages, *acc_full = bars + 2 * kWgStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);
    uint8_t* ring = smem + 256;
    const int stage_bytes = dy_alloc + x_alloc;
    const uint32_t tmem_cols = 512;

    // operand planes that TMA never writes (Cout < 128) must hold finite values: clear the ring once
    for (int i = threadIdx.x; i < kWgStages * stage_bytes / 16; i += kWgThreads)
        reinterpret_cast<uint4*>(ring)[i] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x == 0) {
        for (int i = 0; i < kWgStages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
        mbar_init(acc_full, 1);
        fence_mbar_init();
        tma_prefetch_desc(&map_dy); tma_prefetch_desc(&map_x);
    }
    if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    // this CTA's share of the (sample, 128-position tile) list
    const int t_begin = (int)(((long long)s.total_tiles * blockIdx.z) / gridDim.z);
    const int t_end = (int)(((long long)s.total_tiles * (blockIdx.z + 1)) / gridDim.z);
    const int n_iter = t_end - t_begin;

    if (warp == 0) {
        if (elect_one()) {
            for (int it = 0; it < n_iter; ++it) {
                const int tile = t_begin + it, b = tile / s.tiles_per_sample, l0 = (tile % s.tiles_per_sample) * 128;
                const int sg = it % kWgStages;
                if (it >= kWgStages) mbar_wait(empty + sg, ((it / kWgStages) - 1) & 1);
                if (!s.slot) {
                    mbar_arrive_expect_tx(full + sg, dy_bytes + x_bytes);
                    tma_load_4d(ring + sg * stage_bytes, &map_dy, full + sg, 0, l0, co0 / 8, b);
                    tma_load_4d(ring + sg * stage_bytes + dy_alloc, &map_x, full + sg, 0, l0 + s.k0 - s.pad, ci0 / 8, b);
                } else {
                    // packed short windows: `tile` is a group of 128/slot samples, each dropped into its own slot of rows
                    const int spt = 128 / s.slot;
                    mbar_arrive_expect_tx(full + sg, 128 * 16 * (planes_dy + planes_x));
                    for (int j = 0; j < spt; ++j) {
                        for (int p = 0; p < planes_dy; ++p)
                            tma_load_4d(ring + sg * stage_bytes + p * 128 * 16 + j * s.slot * 16, &map_dys, full + sg, 0, 0,
                                        co0 / 8 + p, tile * spt + j);
                        for (int p = 0; p < planes_x; ++p)
                            tma_load_4d(ring + sg * stage_bytes + dy_alloc + p * s.R * 16 + j * s.slot * 16, &map_xs, full + sg, 0,
                                        s.k0 - s.pad, ci0 / 8 + p, tile * spt + j);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            const uint32_t idesc = idesc_f16(FMT_BF16, kWgMT, kWgNT, MAJOR_MN, MAJOR_MN);
            for (int it = 0; it < n_iter; ++it) {
                const int sg = it % kWgStages;
                mbar_wait(full + sg, (it / kWgStages) & 1);
                tc_fence_after_sync();
                // A = dy planes: M-groups (8 channels) one plane apart, K-groups (8 positions) 128 B apart
                const uint64_t a0 = smem_desc(smem_u32(ring + sg * stage_bytes), 128, 128 * 16);
                const uint64_t b0 = smem_desc(smem_u32(ring + sg * stage_bytes + dy_alloc), 128, s.R * 16);
                for (int c = 0; c < 8; ++c)                      // 8 chunks of 16 positions
                    for (int k = 0; k < s.ktaps; ++k)
                        mma_f16(tmem_base + k * kWgNT, a0 + (uint64_t)(c * 16), b0 + (uint64_t)(c * 16 + k), idesc,
                                (it > 0 || c > 0) ? 1u : 0u);
                mma_commit(empty + sg);
            }
            mma_commit(acc_full);
        }
    } else if (n_iter > 0) {
        const int q = warp & 3;
        mbar_wait(acc_full, 0);
        tc_fence_after_sync();
        const int co = co0 + q * 32 + lane;
        for (int k = 0; k < s.ktaps; ++k) {
            for (int c0 = 0; c0 < kWgNT; c0 += 16) {
                uint32_t v[16];
                tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + k * kWgNT + c0, v);
                tmem_ld_wait();
                if (co < s.Cout && ci0 + c0 < s.Cin) {
                    float* dst = dw + ((size_t)(s.k0 + k) * s.Cout + co) * s.Cin + ci0 + c0;
#pragma unroll
                    for (int j = 0; j < 16; j += 4)
                        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j), "f"(__uint_as_float(v[j])),
                                     "f"(__uint_as_float(v[j + 1])), "f"(__uint_as_float(v[j + 2])), "f"(__uint_as_float(v[j + 3])) : "memory");
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
    (void)dbias;
}

// per-channel column sum of a channels-last bf16 tensor (bias gradient)
__global__ void colsum_kernel(const __nv_bfloat162* __restrict__ x, float* __restrict__ out, int64_t rows, int C) {
    const int C2 = C >> 1;
    const int cp = threadIdx.x % C2, rl = threadIdx.x / C2, rs = blockDim.x / C2;
    if (rl >= rs) return;
    float s0 = 0.f, s1 = 0.f;
    for (int64_t r = (int64_t)blockIdx.x * rs + rl; r < rows; r += (int64_t)gridDim.x * rs) {
        float2 v = __bfloat1622float2(x[r * C2 + cp]);
        s0 += v.x; s1 += v.y;
    }
    atomicAdd(out + 2 * cp, s0); atomicAdd(out + 2 * cp + 1, s1);
}

int conv1d_wgrad(const void* x, const void* dy, float* dw, float* dbias, int B, int L, int Cin, int Cout, int K, int pad,
                 int stride, cudaStream_t st) {
    CS_REQUIRE(stride == 1, "conv1d_wgrad: stride %d not supported", stride);
    CS_REQUIRE(Cin % 16 == 0 && Cout % 16 == 0, "conv1d_wgrad: Cin=%d / Cout=%d must be multiples of 16", Cin, Cout);
    CS_REQUIRE(K >= 1 && K <= 64 && pad >= 0 && pad < K, "conv1d_wgrad: bad K=%d pad=%d", K, pad);
    const int Lout = L + 2 * pad - K + 1;
    WgShape s;
    memset(&s, 0, sizeof(s));
    s.B = B; s.L = L; s.Lout = Lout; s.Cin = Cin; s.Cout = Cout; s.K = K; s.pad = pad;
    s.tiles_per_sample = (Lout + 127) / 128; s.total_tiles = B * s.tiles_per_sample;
    s.my_cout = Cout < kWgMT ? Cout : kWgMT; s.my_cin = Cin < kWgNT ? Cin : kWgNT;
    const int gx = (Cout + kWgMT - 1) / kWgMT, gy = (Cin + kWgNT - 1) / kWgNT;
    CS_REQUIRE(Cout % s.my_cout == 0 && Cin % s.my_cin == 0, "conv1d_wgrad: channel tiling needs Cout %% 128 == 0 or Cout < 128 (same for Cin / 64)");
    int splits = (2 * num_sms()) / (gx * gy); if (splits < 1) splits = 1; if (splits > s.total_tiles) splits = s.total_tiles;
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    // short windows: pack 128/slot samples per 128-row reduction tile (every tap of a sample stays inside its slot)
    s.slot = 0;
    if (Lout + K - 1 <= 64) {
        const int need = Lout + K - 1;
        s.slot = need <= 16 ? 16 : (need <= 32 ? 32 : 64);
        s.tiles_per_sample = 1;
        s.total_tiles = (B + 128 / s.slot - 1) / (128 / s.slot);
        if (splits > s.total_tiles) splits = s.total_tiles;
    }
    int rc128 = conv1d_wgrad_sw128(x, dy, dw, B, L, Cin, Cout, K, pad, Lout, st);
    if (rc128 > 0) return rc128;
    if (rc128 < 0) {                                        // narrow layers: persistent tap-group kernel (conv1d_narrow.cu)
        int rcn = env_int("CS_CONV_PACK_WGRAD", 0) ? conv1d_wgrad_packed(x, dy, dw, B, L, Cin, Cout, K, pad, Lout, st) : -1;
        if (rcn < 0) rcn = conv1d_wgrad_narrow(x, dy, dw, B, L, Cin, Cout, K, pad, Lout, st);
        if (rcn > 0) return rcn;
        if (rcn == 0) rc128 = 0;
    }
    for (int k0 = 0; rc128 < 0 && k0 < K; k0 += kWgTapsPerPass) {
        s.k0 = k0; s.ktaps = K - k0 < kWgTapsPerPass ? K - k0 : kWgTapsPerPass;
        s.R = s.slot ? ((128 + s.ktaps - 1 + 7) & ~7) : 128 + s.ktaps - 1;
        CUtensorMap mdy, mx;
        int rc = make_plane_map(&mdy, dy, B, Lout, Cout, 128, s.my_cout / 8);
        if (rc) return rc;
        rc = make_plane_map(&mx, x, B, L, Cin, s.R, s.my_cin / 8);
        if (rc) return rc;
        const size_t stage = (size_t)128 * 16 * (kWgMT / 8) + (((size_t)s.R * 16 * (kWgNT / 8) + 127) & ~(size_t)127);
        const size_t smem = 256 + kWgStages * stage;
        static size_t configured[64] = {};
        if (smem > configured[dev & 63]) {
            CS_CUDA_OK(cudaFuncSetAttribute(conv1d_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            configured[dev & 63] = smem;
        }
        CUtensorMap mdys = mdy, mxs = mx;
        if (s.slot) {
            rc = make_plane_map(&mdys, dy, B, Lout, Cout, s.slot, 1);
            if (rc) return rc;
            rc = make_plane_map(&mxs, x, B, L, Cin, s.slot, 1);
            if (rc) return rc;
        }
        conv1d_wgrad_kernel<<<dim3(gx, gy, splits), kWgThreads, smem, st>>>(mdy, mx, mdys, mxs, dw, dbias, s);
        CS_CUDA_OK(cudaGetLastError());
    }
    if (dbias) {
        const int64_t rows = (int64_t)B * Lout;
        int threads = 1024, rs = threads / (Cout / 2);
        if (rs < 1) { threads = Cout / 2; rs = 1; }
        int64_t want = (rows + 64LL * rs - 1) / (64LL * rs); if (want > num_sms() * 2) want = num_sms() * 2; if (want < 1) want = 1;
        colsum_kernel<<<(int)want, threads, 0, st>>>((const __nv_bfloat162*)dy, dbias, rows, Cout);
        CS_CUDA_OK(cudaGetLastError());
    }
    return CS_OK;
}

// ------------------------------------------------------------------------------------------------
// first layer (Cin = 1): x [B, L] fp32 -> y [B, Lout, Cout] bf16 on FFMA (reduction K does not fit the
// tensor core), and its weight gradient.
// ------------------------------------------------------------------------------------------------
__global__ void cin1_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                                __nv_bfloat16* __restrict__ y, int L, int Lout, int Cout, int K, int pad, int relu) {
    extern __shared__ float sm1[];
    float* ws = sm1;                       // [Cout][K]
    float* xs = sm1 + Cout * K;            // 256 + K - 1 window
    const int b = blockIdx.y, l0 = blockIdx.x * 256;
    for (int i = threadIdx.x; i < Cout * K; i += blockDim.x) ws[i] = w[i];
    for (int i = threadIdx.x; i < 256 + K - 1; i += blockDim.x) {
        const int l = l0 + i - pad;
        xs[i] = (l >= 0 && l < L) ? x[(size_t)b * L + l] : 0.0f;
    }
    __syncthreads();
    // thread = (position, channel pair group): consecutive threads write consecutive channels of one row
    const int C2 = Cout >> 1;
    for (int i = threadIdx.x; i < 256 * C2; i += blockDim.x) {
        const int p = i / C2, c = (i % C2) * 2, l = l0 + p;
        if (l >= Lout) continue;
        float a0 = bias ? bias[c] : 0.0f, a1 = bias ? bias[c + 1] : 0.0f;
        for (int k = 0; k < K; ++k) { const float v = xs[p + k]; a0 = fmaf(ws[c * K + k], v, a0); a1 = fmaf(ws[(c + 1) * K + k], v, a1); }
        if (relu) { a0 = fmaxf(a0, 0.0f); a1 = fmaxf(a1, 0.0f); }
        reinterpret_cast<__nv_bfloat162*>(y + ((size_t)b * Lout + l) * Cout)[c >> 1] = __floats2bfloat162_rn(a0, a1);
    }
}

int conv1d_cin1_fwd(const float* x, const float* w, const float* bias, void* y, int B, int L, int Cout, int K, int pad,
                    int stride, int relu, cudaStream_t st) {
    CS_REQUIRE(stride == 1 && Cout % 2 == 0 && K >= 1 && K <= 64, "conv1d_cin1: unsupported stride/Cout/K");
    const int Lout = L + 2 * pad - K + 1;
    const size_t smem = ((size_t)Cout * K + 256 + K) * sizeof(float);
    CS_REQUIRE(smem <= 48 * 1024, "conv1d_cin1: Cout*K too large");
    cin1_fwd_kernel<<<dim3((Lout + 255) / 256, B), 256, smem, st>>>(x, w, bias, (__nv_bfloat16*)y, L, Lout, Cout, K, pad, relu);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// dw[c][k] += sum_{b,l} dy[b,l,c] * x[b, l+k-pad];  dbias[c] += sum dy.  One CTA per (position block, sample
// group); thread = (channel, tap); block-local accumulation, one atomic per output per CTA.
__global__ void cin1_wgrad_kernel(const float* __restrict__ x, const __nv_bfloat16* __restrict__ dy, float* __restrict__ dw,
                                  float* __restrict__ dbias, int B, int L, int Lout, int Cout, int K, int pad) {
    extern __shared__ float sm2[];
    float* xs = sm2;                                        // 128 + K - 1
    __nv_bfloat16* ds = reinterpret_cast<__nv_bfloat16*>(xs + 128 + K);   // [128][Cout]
    const int tiles = (Lout + 127) / 128;
    const int outs = Cout * (K + 1);                        // K taps + bias column per channel
    float acc[4] = {0.f, 0.f, 0.f, 0.f};                    // up to 4 outputs per thread (outs <= 4 * blockDim)
    for (int t = blockIdx.x; t < B * tiles; t += gridDim.x) {
        const int b = t / tiles, l0 = (t % tiles) * 128;
        __syncthreads();
        for (int i = threadIdx.x; i < 128 + K - 1; i += blockDim.x) {
            const int l = l0 + i - pad;
            xs[i] = (l >= 0 && l < L) ? x[(size_t)b * L + l] : 0.0f;
        }
        for (int i = threadIdx.x; i < 128 * Cout; i += blockDim.x) {
            const int p = i / Cout, c = i % Cout, l = l0 + p;
            ds[i] = l < Lout ? dy[((size_t)b * Lout + l) * Cout + c] : __float2bfloat16(0.0f);
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int o = threadIdx.x + r * blockDim.x;
            if (o < outs) {
                const int c = o / (K + 1), k = o % (K + 1);
                float a = 0.0f;
                if (k < K) { for (int p = 0; p < 128; ++p) a = fmaf(__bfloat162float(ds[p * Cout + c]), xs[p + k], a); }
                else { for (int p = 0; p < 128; ++p) a += __bfloat162float(ds[p * Cout + c]); }
                acc[r] += a;
            }
        }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int o = threadIdx.x + r * blockDim.x;
        if (o < outs) {
            const int c = o / (K + 1), k = o % (K + 1);
            if (k < K) atomicAdd(dw + c * K + k, acc[r]);
            else if (dbias) atomicAdd(dbias + c, acc[r]);
        }
    }
}

int conv1d_cin1_wgrad(const float* x, const void* dy, float* dw, float* dbias, int B, int L, int Cout, int K, int pad,
                      int stride, cudaStream_t st) {
    CS_REQUIRE(stride == 1 && K >= 1 && K <= 64, "conv1d_cin1_wgrad: unsupported stride/K");
    CS_REQUIRE(Cout * (K + 1) <= 4 * 512, "conv1d_cin1_wgrad: Cout*(K+1) = %d too large", Cout * (K + 1));
    const int Lout = L + 2 * pad - K + 1;
    const size_t smem = (128 + K + 4) * sizeof(float) + (size_t)128 * Cout * 2;
    CS_REQUIRE(smem <= 48 * 1024, "conv1d_cin1_wgrad: tile too large");
    int blocks = B * ((Lout + 127) / 128); if (blocks > num_sms() * 2) blocks = num_sms() * 2;
    cin1_wgrad_kernel<<<blocks, 512, smem, st>>>(x, (const __nv_bfloat16*)dy, dw, dbias, B, L, Lout, Cout, K, pad);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

}  // namespace cs
