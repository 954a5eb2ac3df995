// conv1d on tcgen05, second generation: 128-byte swizzled operand rows for wide layers (Cin % 64 == 0) —
// forward / data gradient (single CTAs, two per SM, or cta_group::2 pairs), fused BatchNorm statistics in the
// epilogue, and the weight gradient on MN-major tiles.  See DESIGN.md section 3c for the measurements behind it.
// Replaces nn.Conv1d -> cuDNN of the reference (Module_3/tiny_ecg_model.py:17-20) for the wide models.
#include "conv1d_common.cuh"

namespace cs {

// ------------------------------------------------------------------------------------------------
// K1, wide-channel variant (Cin % 64 == 0, L long enough for 128-position tiles).
// The plane layout (conv1d_planes.cu) costs one TMA "row" per 16 bytes: at C = 256 the copy engine, not the tensor core,
// sets the pace (ncu: tensor pipe 46 % active, L2 sectors 1.7x the useful bytes).  Here both operands are
// staged as 128-byte rows (64 channels) with the 128-byte swizzle, 8x fewer TMA rows for the same bytes:
//   A tile  = R rows (positions) x 128 B, K-major SWIZZLE_128B;  W tile = NT rows (cout) x 128 B.
// A tap is still a start-address shift (tap * 128 B).  That start is not aligned to the 1024-byte swizzle
// pattern, and that is fine: the tensor core applies the XOR to the ABSOLUTE shared-memory address bits
// (the same ones TMA used when it wrote the tile), so row r+tap is found where TMA put it.  Measured on
// B200: with the descriptor's base-offset field left 0 the results match the fp32 oracle; setting it to
// (start >> 7) & 7 double-counts the phase and gives wrong sums.
// The epilogue packs bf16 rows into a swizzled 32-row slab per warp and writes it with one TMA store
// (128-byte rows; rows past Lout are clipped by the tensor map) instead of 16-byte scattered stores.
// ------------------------------------------------------------------------------------------------
struct Fwd128Shape {
    int B, L, Lout, Cin, Cout, K, pad, relu;
    int NT, R;               // cout per CTA, rows per A tile (128 + K - 1)
    int MT;                  // 128-position tiles per CTA (each with its own TMEM accumulator); the tiles of a CTA are
                             // consecutive entries of the flat (sample, tile) list, so a CTA may straddle two samples
    int AS, WS;              // activation ring stages (1 or 2), weight ring stages (2..8)
    int tiles, total;        // tiles per sample, B * tiles
    int tile_stride;         // bytes between the MT A tiles of one stage (multiple of 1024)
    int tma_store;           // epilogue through shared memory + TMA store (NT % 64 == 0)
    int wmn;                 // 1: data gradient straight from the forward weights w[K][red][N] (N contiguous): the weight tile
                             // is an MN-major B operand and the taps are read in reverse — no transposed copy of w
    int slot;                // 0, or 32 / 64: short windows (Lout + K - 1 <= slot), a tile packs 128/slot whole samples, each
                             // in its own `slot` rows (the taps of one sample never reach the next sample's outputs)
    // fused BatchNorm statistics of the OUTPUT (training mode): per-channel sum / sum of squares of the bf16 values
    // this kernel stores, accumulated into bn_acc[2 * Cout]; the last CTA to arrive turns them into mean / rstd /
    // running statistics / affine coefficients and clears the workspace (same contract as cs_bn_stats_finalize)
    float* bn_acc;           // nullptr: no statistics
    int bn_arrivals;         // CTAs that reach the ticket
    BnStatsTail bn;
};

constexpr int kMaxWStages = 8;

static int make_row_map(CUtensorMap* m, const void* base, int outer, int rows, int C, int box_rows) {
    // [outer, rows, C] bf16 viewed as (C, rows, outer); box = 64 channels (one 128-byte swizzled row) x box_rows
    EncodeTiledFn fn = encode_fn();
    CS_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from this driver");
    cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)rows, (cuuint64_t)outer};
    cuuint64_t strides[2] = {(cuuint64_t)C * 2, (cuuint64_t)rows * C * 2};
    cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CS_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (128B rows) failed (%d) rows=%d C=%d box_rows=%d", (int)r, rows, C, box_rows);
    return CS_OK;
}

// Sized so that TWO CTAs fit one SM (<= 113 KB shared, <= 256 TMEM columns each): while one CTA drains its
// accumulators and ramps up its first loads the other keeps the tensor core busy (measured 1.16 -> 1.36
// PFLOP/s at C = 256, K = 7 against one 4-tile CTA per SM).
// PAIR = true: the CTA is one half of a cta_group::2 pair (cluster of two along grid.x).  The pair issues ONE
// M256 x N(NT) x K16 MMA per step: each CTA stages its own position tiles (A rows) and its own NT/2 rows of the
// weight tile, so per MMA a CTA reads half the B bytes from shared memory (the single-CTA M128 x N128 shape sits on
// the 128 B/clk shared-memory port) and streams half the weights from L2.  Only the leader (cluster rank 0) waits on
// the "full" barriers and issues MMAs; both CTAs load, and both drain their own 128 accumulator lanes.
template <bool PAIR>
__device__ __forceinline__ void conv1d_fwd_sw128_body(const CUtensorMap& map_x, const CUtensorMap& map_w, const CUtensorMap& map_y,
                                                      const CUtensorMap& map_xs, const float* __restrict__ bias,
                                                      __nv_bfloat16* __restrict__ y, const Fwd128Shape& s) {
    extern __shared__ uint8_t smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
    // single CTAs: grid = (cout tile, unit y, unit z).  Pairs: grid = (unit, cout tile) — the two CTAs of a
    // cta_group::2 pair must be neighbours along x (a cluster with an odd x extent is rejected at launch)
    const int n0 = (PAIR ? blockIdx.y : blockIdx.x) * s.NT;
    const int t0 = (PAIR ? blockIdx.x : blockIdx.z * gridDim.y + blockIdx.y) * s.MT;
    if (t0 - (int)rank * s.MT >= s.total) return;            // padding CTAs of the unit split (pair-uniform)
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);
    uint64_t* full_a = bars, *empty_a = bars + 2, *full_w = bars + 4, *empty_w = full_w + kMaxWStages;
    uint64_t* acc_full = empty_w + kMaxWStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);
    // operand rings start on a 1024-byte boundary of the shared window (the swizzle pattern is address based)
    uint8_t* a_ring = smem_raw + (((smem_u32(smem_raw) + 256 + 1023) & ~1023u) - smem_u32(smem_raw));
    const int a_bytes = s.tile_stride * s.MT;
    const int w_rows = PAIR ? s.NT / 2 : s.NT;               // weight rows (output channels) this CTA stages
    const int w_bytes = w_rows * 128;
    uint8_t* w_ring = a_ring + s.AS * a_bytes;
    const int chunks = s.Cin / 64;
    uint32_t tmem_cols = 32;
    while ((int)tmem_cols < s.NT * s.MT) tmem_cols <<= 1;

    if (threadIdx.x == 0) {
        for (int i = 0; i < s.AS; ++i) { mbar_init(full_a + i, 1); mbar_init(empty_a + i, 1); }
        for (int i = 0; i < s.WS; ++i) { mbar_init(full_w + i, 1); mbar_init(empty_w + i, 1); }
        mbar_init(acc_full, 1);
        fence_mbar_init();
        tma_prefetch_desc(&map_x); tma_prefetch_desc(&map_w); tma_prefetch_desc(&map_y); tma_prefetch_desc(&map_xs);
    }
    if (PAIR) {
        cluster_sync();                                      // the peer's barriers exist before anything can arrive on them
        if (warp == 1) tmem_alloc_pair(tmem_slot, tmem_cols);
    } else if (warp == 1) {
        tmem_alloc(tmem_slot, tmem_cols);
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    // loads of both CTAs are counted on the LEADER's full barriers
    const uint32_t txm = PAIR ? 2u : 1u;
    auto load3 = [&](void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
        if (PAIR) tma_load_3d_pair(dst, map, mapa_shared(smem_u32(bar), 0), c0, c1, c2);
        else tma_load_3d(dst, map, bar, c0, c1, c2);
    };

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            int wi = 0;
            for (int c = 0; c < chunks; ++c) {
                const int sa = c % s.AS;
                if (c >= s.AS) mbar_wait(empty_a + sa, ((c / s.AS) - 1) & 1);
                if (!s.slot) {
                    if (rank == 0) mbar_arrive_expect_tx(full_a + sa, txm * s.R * 128 * s.MT);
                    for (int i = 0; i < s.MT; ++i) {
                        // tiles past the end of the list are fetched out of bounds (sample index >= B): zero fill
                        const int t = t0 + i, tb = t / s.tiles, tl = (t % s.tiles) * 128;
                        load3(a_ring + sa * a_bytes + i * s.tile_stride, &map_x, full_a + sa, c * 64, tl - s.pad, tb);
                    }
                } else {
                    // one `slot`-row box per sample; rows 128 .. R-1 of the tile are never written and only feed outputs
                    // that the epilogue drops (l >= Lout of the last sample)
                    const int spt = 128 / s.slot;
                    if (rank == 0) mbar_arrive_expect_tx(full_a + sa, txm * 128 * 128 * s.MT);
                    for (int i = 0; i < s.MT; ++i)
                        for (int j = 0; j < spt; ++j)
                            load3(a_ring + sa * a_bytes + i * s.tile_stride + j * s.slot * 128, &map_xs, full_a + sa, c * 64,
                                  -s.pad, (t0 + i) * spt + j);
                }
                for (int k = 0; k < s.K; ++k, ++wi) {
                    const int sw = wi % s.WS;
                    if (wi >= s.WS) mbar_wait(empty_w + sw, ((wi / s.WS) - 1) & 1);
                    if (rank == 0) mbar_arrive_expect_tx(full_w + sw, txm * w_bytes);
                    const int nw = n0 + (int)rank * w_rows;     // this CTA's slice of the output channels
                    if (!s.wmn) {
                        load3(w_ring + sw * w_bytes, &map_w, full_w + sw, c * 64, nw, k);
                    } else {
                        for (int h = 0; h < w_rows / 64; ++h)   // 64 reduction rows x 64 output channels per box
                            load3(w_ring + sw * w_bytes + h * 8192, &map_w, full_w + sw, nw + h * 64, c * 64, s.K - 1 - k);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (pair: the leader only) =====
        if (rank == 0 && elect_one()) {
            const uint32_t idesc = idesc_f16(FMT_BF16, PAIR ? 256 : 128, s.NT, MAJOR_K, s.wmn ? MAJOR_MN : MAJOR_K);
            auto commit = [&](uint64_t* bar) { if (PAIR) mma_commit_pair(bar); else mma_commit(bar); };
            int wi = 0;
            uint32_t first = 1;
            for (int c = 0; c < chunks; ++c) {
                const int sa = c % s.AS;
                mbar_wait(full_a + sa, (c / s.AS) & 1);
                tc_fence_after_sync();
                const uint32_t a_addr = smem_u32(a_ring + sa * a_bytes);
                for (int k = 0; k < s.K; ++k, ++wi) {
                    const int sw = wi % s.WS;
                    mbar_wait(full_w + sw, (wi / s.WS) & 1);
                    tc_fence_after_sync();
                    const uint32_t w_addr = smem_u32(w_ring + sw * w_bytes);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {            // 4 x 16 channels inside the 128-byte row
                        // K-major weights: 16-channel step = 32 B inside the row.  MN-major (data gradient): the 16
                        // reduction channels are 16 rows (2048 B) and the 64-channel output blocks sit 8192 B apart.
                        const uint64_t bd = s.wmn ? smem_desc_sw128_mn(w_addr + j * 2048, 8192, 1024)
                                                  : smem_desc_sw128(w_addr + j * 32, 1024);
                        for (int i = 0; i < s.MT; ++i) {
                            // tap k = k rows (k * 128 B) further down the tile; 16-channel step = 32 B inside the row
                            const uint64_t a = smem_desc_sw128(a_addr + i * s.tile_stride + k * 128 + j * 32, 1024);
                            if (PAIR) mma_f16_pair(tmem_base + i * s.NT, a, bd, idesc, (first && j == 0) ? 0u : 1u);
                            else mma_f16(tmem_base + i * s.NT, a, bd, idesc, (first && j == 0) ? 0u : 1u);
                        }
                    }
                    first = 0;
                    commit(empty_w + sw);
                }
                commit(empty_a + sa);
            }
            commit(acc_full);
        }
    } else {
        // ===== epilogue: TMEM -> registers -> bias/ReLU -> bf16 -> (swizzled slab -> TMA store | global) =====
        const int q = warp & 3;
        mbar_wait(acc_full, 0);                              // every MMA retired: the operand rings are free
        tc_fence_after_sync();
        if (s.tma_store) {
            // two 4 KB slabs per warp inside the (now idle) activation ring; slab row r = accumulator lane
            uint8_t* slabs = a_ring + q * 8192;
            int n_store = 0;
            // BN statistics: this lane's channel pair of each 64-channel block, kept in the (idle) weight ring:
            // [warp][64-channel block][channel][sum, sumsq] — lane-private entries until the warps meet below
            float* red = reinterpret_cast<float*>(w_ring);
            if (s.bn_acc)
                for (int h4 = 0; h4 * 64 < s.NT; ++h4)
                    *reinterpret_cast<float4*>(red + ((q * 4 + h4) * 64 + 2 * lane) * 2) = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int i = 0; i < s.MT; ++i) {
                const int t = t0 + i;
                int tb = t / s.tiles, tl = (t % s.tiles) * 128 + q * 32;
                if (s.slot) { tb = t * (128 / s.slot) + (q * 32) / s.slot; tl = (q * 32) % s.slot; }
                if (t >= s.total || tl >= s.Lout || tb >= s.B) continue;
                for (int hh = 0; hh < s.NT; hh += 64) {
                    uint8_t* slab = slabs + (n_store & 1) * 4096;
                    uint32_t v[64];
                    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + i * s.NT + hh;
                    tmem_ld32(taddr, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
                    tmem_ld32(taddr + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
                    if (n_store >= 2) {                       // the store that last read this slab must be done with it
                        if (lane == 0) tma_store_wait_read<1>();
                        __syncwarp();
                    }
                    tmem_ld_wait();
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        uint32_t pk[4];
#pragma unroll
                        for (int j = 0; j < 8; j += 2) {
                            float f0 = __uint_as_float(v[c * 8 + j]), f1 = __uint_as_float(v[c * 8 + j + 1]);
                            if (bias) { f0 += __ldg(bias + n0 + hh + c * 8 + j); f1 += __ldg(bias + n0 + hh + c * 8 + j + 1); }
                            if (s.relu) { f0 = fmaxf(f0, 0.0f); f1 = fmaxf(f1, 0.0f); }
                            __nv_bfloat162 h2 = __floats2bfloat162_rn(f0, f1);
                            pk[j >> 1] = *reinterpret_cast<uint32_t*>(&h2);
                        }
                        *reinterpret_cast<uint4*>(slab + lane * 128 + ((c ^ (lane & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        tma_store_3d(&map_y, slab, n0 + hh, tl, tb);     // rows past Lout are clipped by the map
                        tma_store_commit();
                    }
                    ++n_store;
                    if (s.bn_acc) {
                        // column sums of the slab just written (the stored bf16 values, rows inside the sample only):
                        // lane = channel pair (2*lane, 2*lane+1) -> 16-byte chunk lane/4, word lane%4 of every row
                        const int valid = s.Lout - tl < 32 ? s.Lout - tl : 32;
                        float a0 = 0.f, a1 = 0.f, q0 = 0.f, q1 = 0.f;
                        for (int r0 = 0; r0 < 32; r0 += 8) {           // 8 independent loads in flight (rows r0 .. r0+7)
                            uint32_t u[8];
#pragma unroll
                            for (int j = 0; j < 8; ++j)                // (r0 + j) & 7 == j: the swizzle phase is the unroll index
                                u[j] = (r0 + j < valid)
                                           ? *reinterpret_cast<const uint32_t*>(slab + (r0 + j) * 128 + (((lane >> 2) ^ j) << 4) + (lane & 3) * 4)
                                           : 0u;
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u[j]));
                                a0 += f.x; q0 = fmaf(f.x, f.x, q0); a1 += f.y; q1 = fmaf(f.y, f.y, q1);
                            }
                        }
                        float4* p = reinterpret_cast<float4*>(red + ((q * 4 + (hh >> 6)) * 64 + 2 * lane) * 2);
                        float4 cur = *p;
                        cur.x += a0; cur.y += q0; cur.z += a1; cur.w += q1;
                        *p = cur;
                    }
                }
            }
            if (lane == 0) tma_store_wait_read<0>();
            __syncwarp();
            if (s.bn_acc) {
                // the four epilogue warps meet in the idle weight ring: [warp][64-channel block][channel][sum, sumsq],
                // then ONE atomic per channel and quantity per CTA; the last CTA to arrive finishes the layer statistics
                const int et = threadIdx.x - 64;             // 0..127 over warps 2..5
                volatile uint32_t* last_flag = tmem_slot + 1;
                asm volatile("bar.sync 1, 128;" ::: "memory");
                for (int o = et; o < s.NT * 2; o += 128) {
                    const int ch = o >> 1, qty = o & 1;
                    float v = 0.0f;
#pragma unroll
                    for (int w = 0; w < 4; ++w) v += red[((w * 4 + (ch >> 6)) * 64 + (ch & 63)) * 2 + qty];
                    atomicAdd(s.bn_acc + qty * s.Cout + n0 + ch, v);
                }
                __threadfence();
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (et == 0) *last_flag = (atomicAdd(s.bn.ticket, 1) == s.bn_arrivals - 1) ? 1u : 0u;
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (*last_flag) {
                    __threadfence();
                    bn_stats_tail_channels(s.bn_acc, s.Cout, s.bn, et, 128);
                    if (et == 0) *s.bn.ticket = 0;
                }
            }
        } else {
            for (int i = 0; i < s.MT; ++i) {
                const int t = t0 + i;
                int tb = t / s.tiles, l = (t % s.tiles) * 128 + q * 32 + lane;
                if (s.slot) { tb = t * (128 / s.slot) + (q * 32 + lane) / s.slot; l = (q * 32 + lane) % s.slot; }
                if (t >= s.total || tb >= s.B) continue;
                __nv_bfloat16* dst = y + ((size_t)tb * s.Lout + l) * s.Cout + n0;
                for (int c0 = 0; c0 < s.NT; c0 += 32) {
                    uint32_t v[32];
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + i * s.NT + c0, v);
                    tmem_ld_wait();
                    if (l < s.Lout) {
#pragma unroll
                        for (int h = 0; h < 32; h += 8) {
                            uint32_t pk[4];
#pragma unroll
                            for (int j = 0; j < 8; j += 2) {
                                float f0 = __uint_as_float(v[h + j]), f1 = __uint_as_float(v[h + j + 1]);
                                if (bias) { f0 += bias[n0 + c0 + h + j]; f1 += bias[n0 + c0 + h + j + 1]; }
                                if (s.relu) { f0 = fmaxf(f0, 0.0f); f1 = fmaxf(f1, 0.0f); }
                                __nv_bfloat162 hh = __floats2bfloat162_rn(f0, f1);
                                pk[j >> 1] = *reinterpret_cast<uint32_t*>(&hh);
                            }
                            *reinterpret_cast<uint4*>(dst + c0 + h) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                        }
                    }
                }
            }
        }
    }
    tc_fence_before_sync();
    if (PAIR) {
        cluster_sync();                                      // both CTAs are done with both shared memories and TMEMs
        if (warp == 1) tmem_dealloc_pair(tmem_base, tmem_cols);
    } else {
        __syncthreads();
        if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
    }
}

// (launch bound 2: two of these CTAs must fit one SM's register file — 168 registers per thread at most)
__global__ void __launch_bounds__(kFwdThreads, 2)
conv1d_fwd_sw128_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                        const __grid_constant__ CUtensorMap map_y, const __grid_constant__ CUtensorMap map_xs,
                        const float* __restrict__ bias, __nv_bfloat16* __restrict__ y, const Fwd128Shape s) {
    conv1d_fwd_sw128_body<false>(map_x, map_w, map_y, map_xs, bias, y, s);
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kFwdThreads, 2)
conv1d_fwd_pair_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                       const __grid_constant__ CUtensorMap map_y, const __grid_constant__ CUtensorMap map_xs,
                       const float* __restrict__ bias, __nv_bfloat16* __restrict__ y, const Fwd128Shape s) {
    conv1d_fwd_sw128_body<true>(map_x, map_w, map_y, map_xs, bias, y, s);
}


// Launches the 128-byte-row variant and returns CS_OK (or an error), or -1 when the shape is not one it covers.
// Tuning overrides for experiments: CS_CONV_SW128=0 (use the plane kernel), CS_CONV_PAIR=0 (no cta_group::2 pairs),
// CS_CONV_MT / CS_CONV_AS / CS_CONV_WS.
// Host-side choice of kernel variant and tile configuration (pure arithmetic: also reachable through cs_conv1d_plan,
// which the CPU tests use to check the resource invariants).  Returns 0, or -1 when the shape is not covered.
static int plan_fwd128(int B, int L, int Cin, int Cout, int K, int pad, int Lout, int wmn, Fwd128Shape& s, int& pair, size_t& smem,
                       dim3& grid, int& units) {
    if (Cin % 64 != 0 || Cout % 32 != 0 || Lout + K - 1 <= 16 || !env_int("CS_CONV_SW128", 1)) return -1;
    if (wmn && Cout % 64 != 0) return -1;
    memset(&s, 0, sizeof(s));
    s.wmn = wmn;
    s.B = B; s.L = L; s.Lout = Lout; s.Cin = Cin; s.Cout = Cout; s.K = K; s.pad = pad; 
    s.R = 128 + K - 1;
    s.tile_stride = (s.R * 128 + 1023) & ~1023;
    s.tiles = (Lout + 127) / 128;
    s.total = B * s.tiles;
    s.slot = 0;
    if (Lout + K - 1 <= 64) {                                // short windows: whole samples packed into the 128-row tile
        s.slot = Lout + K - 1 <= 32 ? 32 : 64;
        s.tiles = 1;
        s.total = (B + 128 / s.slot - 1) / (128 / s.slot);
    }
    const size_t kHalfSm = 113 * 1024, kFullSm = 227 * 1024;
    pair = 0;
    auto need = [&]() {
        return (size_t)2048 + (size_t)s.AS * s.tile_stride * s.MT + (size_t)s.WS * (pair ? s.NT / 2 : s.NT) * 128;
    };
    // --- CTA pairs: N up to 256 per MMA, each CTA stages half of the weight tile.  Worth it once there are enough
    //     tiles to fill the machine with pairs; (wmn needs whole 64-channel blocks per CTA)
    {
        const int nt = Cout % 256 == 0 ? 256 : (Cout % 128 == 0 ? 128 : (Cout % 64 == 0 ? 64 : 0));
        if (nt && (!wmn || nt >= 128) && env_int("CS_CONV_PAIR", 1)) {
            s.NT = nt; s.MT = 256 / nt; s.AS = 2; s.WS = 4;
            pair = 1;
            while (need() > kHalfSm && s.WS > 2) --s.WS;
            if (need() > kHalfSm) { s.AS = 1; s.WS = 4; while (need() > kHalfSm && s.WS > 2) --s.WS; }
            // measured: pairs win 5-8 % with N = 256 per MMA (1.40 vs 1.33 PFLOP/s at C=256 K=7, 1.52 vs 1.40 at K=31) and
            // lose a little with N = 128 / 64 or when there are too few tiles to keep every TPC busy with pairs
            const long long ctas = (long long)(Cout / nt) * ((s.total + s.MT - 1) / s.MT);
            if (need() > kHalfSm || ctas < 4LL * num_sms() || nt != 256) pair = 0;
            if (env_int("CS_CONV_PAIR", 1) == 2 && need() <= kFullSm) pair = 1;      // forced (tests)
        }
    }
    if (!pair) {
        s.NT = Cout % 128 == 0 ? 128 : (Cout % 64 == 0 ? 64 : 32);
        // preferred: two co-resident CTAs per SM (256 TMEM columns and <= 113 KB shared memory each)
        s.MT = 256 / s.NT; if (s.MT > 4) s.MT = 4; if (s.MT > s.total) s.MT = s.total;
        // small problems: fewer tiles per CTA so that every SM gets work (weights are re-streamed from L2, which is cheap here)
        while (s.MT > 1 && (long long)(Cout / s.NT) * ((s.total + s.MT - 1) / s.MT) < 2LL * num_sms()) s.MT >>= 1;
        s.AS = 2; s.WS = 4;
        while (need() > kHalfSm && s.WS > 2) --s.WS;
        if (need() > kHalfSm) {                              // wide taps: one activation stage, the sibling CTA covers the bubble
            s.AS = 1; s.WS = 4;
            while (need() > kHalfSm && s.WS > 2) --s.WS;
        }
        if (need() > kHalfSm) {                              // very wide taps: one CTA per SM, all of TMEM
            s.AS = 2; s.WS = 4;
            s.MT = 512 / s.NT; if (s.MT > 4) s.MT = 4; if (s.MT > s.total) s.MT = s.total;
            while (need() > kFullSm && s.MT > 1) s.MT >>= 1;
            while (need() > kFullSm && s.WS > 2) --s.WS;
            if (need() > kFullSm) return -1;
        }
        { const int v = env_int("CS_CONV_MT", 0); if (v > 0 && v * s.NT <= 512) s.MT = v; }
    }
    { const int v = env_int("CS_CONV_AS", 0); if (v == 1 || v == 2) s.AS = v; }
    { const int v = env_int("CS_CONV_WS", 0); if (v >= 2 && v <= kMaxWStages) s.WS = v; }
    if (need() > kFullSm) return -1;
    s.tma_store = (s.NT % 64 == 0 && (size_t)s.AS * s.tile_stride * s.MT >= 32768 && env_int("CS_CONV_TMA_STORE", 1)) ? 1 : 0;
    smem = need();
    units = (s.total + s.MT - 1) / s.MT;
    if (pair) units = (units + 1) & ~1;                      // whole pairs; the odd one out loads zeros and stores nothing
    const int gy = units < 32768 ? units : 32768;
    grid = dim3(Cout / s.NT, gy, (units + gy - 1) / gy);
    if (pair) grid = dim3(units, Cout / s.NT, 1);            // pairs are neighbours along x
    return 0;
}

int conv1d_fwd_sw128(const void* x, const void* w, const float* bias, void* y, int B, int L, int Cin, int Cout, int K, int pad,
                     int relu, int Lout, cudaStream_t st, int wmn, const ConvBnArgs* bn) {
    Fwd128Shape s;
    int pair = 0, units = 0;
    size_t smem = 0;
    dim3 grid;
    if (plan_fwd128(B, L, Cin, Cout, K, pad, Lout, wmn, s, pair, smem, grid, units) != 0) return -1;
    s.relu = relu;
    const int w_rows = pair ? s.NT / 2 : s.NT;
    CUtensorMap mx, mw;
    int rc = make_row_map(&mx, x, B, L, Cin, s.R);
    if (rc) return rc;
    rc = wmn ? make_row_map(&mw, w, K, Cin, Cout, 64) : make_row_map(&mw, w, K, Cout, Cin, w_rows);
    if (rc) return rc;
    CUtensorMap my = mw, mxs = mx;
    if (s.tma_store) {
        rc = make_row_map(&my, y, B, Lout, Cout, 32);
        if (rc) return rc;
    }
    if (s.slot) {
        rc = make_row_map(&mxs, x, B, L, Cin, s.slot);
        if (rc) return rc;
    }
    static size_t configured[2][64] = {};
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    if (smem > configured[pair][dev & 63]) {
        if (pair) CS_CUDA_OK(cudaFuncSetAttribute(conv1d_fwd_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        else CS_CUDA_OK(cudaFuncSetAttribute(conv1d_fwd_sw128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        // ask for the largest shared-memory carve-out: two ~104 KB CTAs only fit one SM with the full 228 KB
        if (env_int("CS_CONV_CARVEOUT", 1)) {
            if (pair) CS_CUDA_OK(cudaFuncSetAttribute(conv1d_fwd_pair_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
            else CS_CUDA_OK(cudaFuncSetAttribute(conv1d_fwd_sw128_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        }
        configured[pair][dev & 63] = smem;
    }
    if (bn) {
        if (!s.tma_store) return -1;                         // statistics are taken from the shared-memory slabs
        s.bn_acc = bn->acc;
        s.bn_arrivals = (Cout / s.NT) * units;               // every CTA that does not return early
        s.bn = BnStatsTail{bn->ticket, bn->mean, bn->rstd, bn->running_mean, bn->running_var, (float)((double)B * Lout), bn->eps,
                           bn->momentum, bn->gamma, bn->beta, bn->coef};
    }
    if (env_int("CS_CONV_DEBUG", 0)) {
        int occ = -1;
        if (pair) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = grid; cfg.blockDim = dim3(kFwdThreads); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr; cfg.numAttrs = 1;
            cudaOccupancyMaxActiveClusters(&occ, conv1d_fwd_pair_kernel, &cfg);
        } else {
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, conv1d_fwd_sw128_kernel, kFwdThreads, smem);
        }
        fprintf(stderr, "[conv sw128] pair=%d grid=(%u,%u,%u) smem=%zu NT=%d MT=%d AS=%d WS=%d occupancy(%s)=%d\n", pair, grid.x, grid.y,
                grid.z, smem, s.NT, s.MT, s.AS, s.WS, pair ? "clusters on the device" : "CTAs per SM", occ);
    }
    if (!pair) conv1d_fwd_sw128_kernel<<<grid, kFwdThreads, smem, st>>>(mx, mw, my, mxs, bias, (__nv_bfloat16*)y, s);
    else conv1d_fwd_pair_kernel<<<grid, kFwdThreads, smem, st>>>(mx, mw, my, mxs, bias, (__nv_bfloat16*)y, s);   // cluster (2,1,1)
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// conv forward + training-mode BatchNorm statistics of its output in one launch (wide layers only).
// CS_ERR_UNSUPPORTED (nothing launched) when the shape is not on the slab-epilogue path: run cs_conv1d_fwd and
// cs_bn_stats_finalize instead.
int conv1d_fwd_bn(const void* x, const void* w, const float* bias, void* y, int B, int L, int Cin, int Cout, int K, int pad, int relu,
                  float* acc, int32_t* ticket, float* mean, float* rstd, float* rm, float* rv, float eps, float momentum,
                  const float* gamma, const float* beta, float* coef, cudaStream_t st) {
    CS_REQUIRE(K >= 1 && K <= 64 && pad >= 0 && pad < K, "conv1d: bad K=%d pad=%d", K, pad);
    CS_REQUIRE((((uintptr_t)x | (uintptr_t)w | (uintptr_t)y) & 15) == 0, "conv1d: pointers must be 16-byte aligned");
    CS_REQUIRE(acc && ticket && mean && rstd, "conv1d_fwd_bn: null workspace / output");
    CS_REQUIRE(coef == nullptr || (gamma && beta), "conv1d_fwd_bn: coef needs gamma and beta");
    CS_REQUIRE((rm == nullptr) == (rv == nullptr), "conv1d_fwd_bn: give both running statistics or neither");
    const int Lout = L + 2 * pad - K + 1;
    if (Lout < 1 || !env_int("CS_CONV_BN_FUSE", 1)) return CS_ERR_UNSUPPORTED;
    ConvBnArgs bn{acc, ticket, mean, rstd, rm, rv, eps, momentum, gamma, beta, coef};
    const int rc = conv1d_fwd_sw128(x, w, bias, y, B, L, Cin, Cout, K, pad, relu, Lout, st, 0, &bn);
    return rc < 0 ? CS_ERR_UNSUPPORTED : rc;
}

// dx = conv(dy, flipped/transposed w) computed from the forward weights themselves (see Fwd128Shape::wmn).
// CS_ERR_UNSUPPORTED (no error text) when the shape needs the transposed copy + conv1d_fwd instead.
int conv1d_dgrad_direct(const void* dy, const void* w, void* dx, int B, int L, int Cin, int Cout, int K, int pad, cudaStream_t st) {
    CS_REQUIRE(K >= 1 && K <= 64 && pad >= 0 && pad < K, "conv1d_dgrad: bad K=%d pad=%d", K, pad);
    CS_REQUIRE((((uintptr_t)dy | (uintptr_t)w | (uintptr_t)dx) & 15) == 0, "conv1d_dgrad: pointers must be 16-byte aligned");
    const int Lout = L + 2 * pad - K + 1;
    const int rc = conv1d_fwd_sw128(dy, w, nullptr, dx, B, Lout, Cout, Cin, K, K - 1 - pad, 0, L, st, 1);
    return rc < 0 ? CS_ERR_UNSUPPORTED : rc;
}

// ------------------------------------------------------------------------------------------------
// K3, wide-channel variant (Cin % 64 == 0 and Cout % 64 == 0): the same 128-byte swizzled rows as the forward
// kernel.  A channels-last tile (row = position, 64 channels = 128 B) IS the MN-major SWIZZLE_128B operand
// layout with the positions as the reduction dimension, so dy and x are used exactly as TMA delivers them:
//   A = dy tile   [128 positions x 128 cout]  (two 64-channel blocks, LBO apart)      M = cout
//   B = x tile    [R positions   x NC cin ]   (NC / 64 blocks), tap = start + tap * 128 B   N = cin
// One TMEM accumulator [128 x NC] per tap, TP = 512 / NC taps per CTA.  grid = (cout tile, cin tile, tap group x
// split of the (sample, position tile) list): all tap groups run in the same launch.
// ------------------------------------------------------------------------------------------------
constexpr int kWg128Stages = 3;
struct Wg128Shape { int B, L, Lout, Cin, Cout, K, pad, TP, S, R, NC, x_block, tiles_per_sample, total_tiles, my_cout, slot; };

__global__ void __launch_bounds__(kWgThreads, 1)
conv1d_wgrad_sw128_kernel(const __grid_constant__ CUtensorMap map_dy, const __grid_constant__ CUtensorMap map_x,
                          const __grid_constant__ CUtensorMap map_dys, const __grid_constant__ CUtensorMap map_xs,
                          float* __restrict__ dw, Wg128Shape s) {
    extern __shared__ uint8_t smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int co0 = blockIdx.x * 128, ci0 = blockIdx.y * s.NC;
    const int dy_alloc = 2 * 16384;                            // M = 128 always reads two 64-channel blocks
    const int x_alloc = s.x_block * (s.NC / 64);
    const int stage_bytes = dy_alloc + x_alloc;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);
    uint64_t* full = bars, *empty = bars + kWg128Stages, *acc_full = bars + 2 * kWg128Stages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);
    uint8_t* ring = smem_raw + (((smem_u32(smem_raw) + 256 + 1023) & ~1023u) - smem_u32(smem_raw));
    const uint32_t tmem_cols = 512;

    // rows / blocks TMA never writes (Cout == 64; rows 128.. of a packed x tile) must hold finite values that meet
    // zeros on the other operand: clear the ring once
    if (s.my_cout < 128 || s.slot)
        for (int i = threadIdx.x; i < kWg128Stages * stage_bytes / 16; i += kWgThreads)
            reinterpret_cast<uint4*>(ring)[i] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x == 0) {
        for (int i = 0; i < kWg128Stages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
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
    const int split = blockIdx.z % s.S, k0 = (blockIdx.z / s.S) * s.TP;
    const int ktaps = s.K - k0 < s.TP ? s.K - k0 : s.TP;
    const int t_begin = (int)(((long long)s.total_tiles * split) / s.S);
    const int t_end = (int)(((long long)s.total_tiles * (split + 1)) / s.S);
    const int n_iter = t_end - t_begin;

    if (warp == 0) {
        if (elect_one()) {
            const uint32_t tx = (uint32_t)(s.my_cout / 64) * 16384u + (uint32_t)(s.NC / 64) * s.R * 128u;
            for (int it = 0; it < n_iter; ++it) {
                const int tile = t_begin + it, b = tile / s.tiles_per_sample, l0 = (tile % s.tiles_per_sample) * 128;
                const int sg = it % kWg128Stages;
                if (it >= kWg128Stages) mbar_wait(empty + sg, ((it / kWg128Stages) - 1) & 1);
                uint8_t* st = ring + sg * stage_bytes;
                if (!s.slot) {
                    mbar_arrive_expect_tx(full + sg, tx);
                    for (int h = 0; h < s.my_cout / 64; ++h) tma_load_3d(st + h * 16384, &map_dy, full + sg, co0 + h * 64, l0, b);
                    for (int h = 0; h < s.NC / 64; ++h)
                        tma_load_3d(st + dy_alloc + h * s.x_block, &map_x, full + sg, ci0 + h * 64, l0 + k0 - s.pad, b);
                } else {
                    // packed short windows: `tile` is a group of 128/slot samples, one `slot`-row box each (dy rows past
                    // Lout arrive as zeros, which is what keeps a sample's taps out of its neighbour)
                    const int spt = 128 / s.slot;
                    mbar_arrive_expect_tx(full + sg, (uint32_t)(s.my_cout / 64 + s.NC / 64) * 16384u);
                    for (int j = 0; j < spt; ++j) {
                        for (int h = 0; h < s.my_cout / 64; ++h)
                            tma_load_3d(st + h * 16384 + j * s.slot * 128, &map_dys, full + sg, co0 + h * 64, 0, tile * spt + j);
                        for (int h = 0; h < s.NC / 64; ++h)
                            tma_load_3d(st + dy_alloc + h * s.x_block + j * s.slot * 128, &map_xs, full + sg, ci0 + h * 64,
                                        k0 - s.pad, tile * spt + j);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            const uint32_t idesc = idesc_f16(FMT_BF16, 128, s.NC, MAJOR_MN, MAJOR_MN);
            for (int it = 0; it < n_iter; ++it) {
                const int sg = it % kWg128Stages;
                mbar_wait(full + sg, (it / kWg128Stages) & 1);
                tc_fence_after_sync();
                const uint32_t a_addr = smem_u32(ring + sg * stage_bytes), b_addr = a_addr + dy_alloc;
                for (int c = 0; c < 8; ++c) {                   // 8 chunks of 16 positions (16 rows x 128 B = 2048 B)
                    const uint64_t a = smem_desc_sw128_mn(a_addr + c * 2048, 16384, 1024);
                    for (int k = 0; k < ktaps; ++k)
                        mma_f16(tmem_base + k * s.NC, a, smem_desc_sw128_mn(b_addr + c * 2048 + k * 128, s.x_block, 1024), idesc,
                                (it > 0 || c > 0) ? 1u : 0u);
                }
                mma_commit(empty + sg);
            }
            mma_commit(acc_full);
        }
    } else if (n_iter > 0) {
        const int q = warp & 3;
        mbar_wait(acc_full, 0);
        tc_fence_after_sync();
        const int co = co0 + q * 32 + lane;
        for (int k = 0; k < ktaps; ++k) {
            for (int c0 = 0; c0 < s.NC; c0 += 32) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + k * s.NC + c0, v);
                tmem_ld_wait();
                if (co < s.Cout) {
                    float* dst = dw + ((size_t)(k0 + k) * s.Cout + co) * s.Cin + ci0 + c0;
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j), "f"(__uint_as_float(v[j])),
                                     "f"(__uint_as_float(v[j + 1])), "f"(__uint_as_float(v[j + 2])), "f"(__uint_as_float(v[j + 3])) : "memory");
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, tmem_cols);
}

// returns -1 when the shape is outside this variant (the plane kernel handles it)
static int plan_wgrad128(int B, int L, int Cin, int Cout, int K, int pad, int Lout, Wg128Shape& s, int& passes, size_t& smem, dim3& grid) {
    if (Cin % 64 != 0 || Cout % 64 != 0 || (Cout > 64 && Cout % 128 != 0) || Lout + K - 1 <= 16 || !env_int("CS_CONV_SW128", 1)) return -1;
    memset(&s, 0, sizeof(s));
    s.B = B; s.L = L; s.Lout = Lout; s.Cin = Cin; s.Cout = Cout; s.K = K; s.pad = pad;
    s.tiles_per_sample = (Lout + 127) / 128; s.total_tiles = B * s.tiles_per_sample;
    s.my_cout = Cout < 128 ? Cout : 128;
    if (Lout + K - 1 <= 64) {                                // short windows: 128/slot whole samples per reduction tile
        s.slot = Lout + K - 1 <= 32 ? 32 : 64;
        s.tiles_per_sample = 1;
        s.total_tiles = (B + 128 / s.slot - 1) / (128 / s.slot);
    }
    // 128 input channels per CTA halve the shared-memory operand bytes per FLOP (4 taps per pass); very wide filters
    // are better off with 64 channels x 8 taps per pass (fewer passes over dy and x)
    s.NC = (K <= 8 && Cin % 128 == 0) ? 128 : 64;
    { const int v = env_int("CS_WGRAD_NC", 0); if ((v == 64 || v == 128) && Cin % v == 0) s.NC = v; }
    s.TP = 512 / s.NC;
    passes = (K + s.TP - 1) / s.TP;
    const int gx = (Cout + 127) / 128, gy = Cin / s.NC, outs = gx * gy * passes;
    // Split of the reduction (sample, tile) list.  Every split ends with an atomic flush of its [taps x 128 x NC]
    // partial (256 KB at most), and all CTAs flush at about the same time, so the flush costs ~ splits * outs * c_epi
    // while the MMA time falls as tiles / splits: take the minimum of the sum, capped at two CTAs per SM.
    const double t_tile = 8.0 * (K < s.TP ? K : s.TP) * (s.NC / 2) / 1.7e3;     // us per tile (MMA issue bound, 1.7 GHz)
    const double c_epi = 0.17;                                                    // us per flushed CTA (measured ~1.5 TB/s)
    int splits = (int)(sqrt((double)s.total_tiles * t_tile / (outs * c_epi)) + 0.5);
    { const int v = env_int("CS_WGRAD_SPLITS", 0); if (v > 0) splits = v; }
    // whole waves: one CTA per SM is resident (all of TMEM), so outs * splits should sit just under 1x or 2x the SM count
    const int one_wave = num_sms() / outs > 0 ? num_sms() / outs : 1, two_waves = (2 * num_sms()) / outs > 0 ? (2 * num_sms()) / outs : 1;
    if (env_int("CS_WGRAD_SPLITS", 0) <= 0) {
        if (splits >= two_waves) splits = two_waves;
        else if (splits > one_wave) splits = (2 * splits > 3 * one_wave) ? two_waves : one_wave;
    }
    if (splits > s.total_tiles) splits = s.total_tiles;
    if (splits < 1) splits = 1;
    s.S = splits;
    s.R = 128 + (K < s.TP ? K : s.TP) - 1;
    s.x_block = (s.R * 128 + 1023) & ~1023;
    smem = 2048 + (size_t)kWg128Stages * (2 * 16384 + (size_t)s.x_block * (s.NC / 64));
    grid = dim3(gx, gy, passes * splits);
    return 0;
}

int conv1d_wgrad_sw128(const void* x, const void* dy, float* dw, int B, int L, int Cin, int Cout, int K, int pad, int Lout,
                       cudaStream_t st) {
    Wg128Shape s;
    int passes = 0;
    size_t smem = 0;
    dim3 grid;
    if (plan_wgrad128(B, L, Cin, Cout, K, pad, Lout, s, passes, smem, grid) != 0) return -1;
    CUtensorMap mdy, mx;
    int rc = make_row_map(&mdy, dy, B, Lout, Cout, 128);
    if (rc) return rc;
    rc = make_row_map(&mx, x, B, L, Cin, s.R);
    if (rc) return rc;
    CUtensorMap mdys = mdy, mxs = mx;
    if (s.slot) {
        rc = make_row_map(&mdys, dy, B, Lout, Cout, s.slot);
        if (rc) return rc;
        rc = make_row_map(&mxs, x, B, L, Cin, s.slot);
        if (rc) return rc;
    }
    static size_t configured[64] = {};
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    if (smem > configured[dev & 63]) {
        CS_CUDA_OK(cudaFuncSetAttribute(conv1d_wgrad_sw128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev & 63] = smem;
    }
    conv1d_wgrad_sw128_kernel<<<grid, kWgThreads, smem, st>>>(mdy, mx, mdys, mxs, dw, s);
    CS_CUDA_OK(cudaGetLastError());
    return CS_OK;
}

// What the dispatch would do for a shape, without launching anything (diagnostics and CPU-side tests).
// direction: 0 forward, 1 data gradient from the forward weights, 2 weight gradient.  out[16]:
//   forward / dgrad: {variant (0 plane kernel, 1 swizzled rows, 2 CTA pairs), NT, MT, AS, WS, slot, tma_store, smem bytes,
//                     grid x, y, z, TMEM columns per CTA, tiles, units, 0, 0}
//   weight gradient: {variant (0 plane kernel, 1 swizzled rows), NC, taps per CTA, tap groups, splits, slot, 0, smem bytes,
//                     grid x, y, z, 512, tiles, 0, 0, 0}
int conv1d_plan(int B, int L, int Cin, int Cout, int K, int pad, int direction, int* out) {
    CS_REQUIRE(out != nullptr && direction >= 0 && direction <= 2, "conv1d_plan: bad arguments");
    CS_REQUIRE(K >= 1 && K <= 64 && pad >= 0 && pad < K && B >= 1 && L >= 1, "conv1d_plan: bad shape");
    for (int i = 0; i < 16; ++i) out[i] = 0;
    const int Lout = L + 2 * pad - K + 1;
    CS_REQUIRE(Lout >= 1, "conv1d_plan: empty output");
    size_t smem = 0;
    dim3 grid;
    if (direction == 2) {
        Wg128Shape s;
        int passes = 0;
        if (plan_wgrad128(B, L, Cin, Cout, K, pad, Lout, s, passes, smem, grid) != 0) return CS_OK;
        const int v[16] = {1, s.NC, s.TP, passes, s.S, s.slot, 0, (int)smem, (int)grid.x, (int)grid.y, (int)grid.z, 512, s.total_tiles, 0, 0, 0};
        memcpy(out, v, sizeof(v));
        return CS_OK;
    }
    Fwd128Shape s;
    int pair = 0, units = 0;
    // the data gradient is the forward kernel on dy: reduction over Cout, outputs Cin, padding K-1-pad, output length L
    const int rc = direction == 0 ? plan_fwd128(B, L, Cin, Cout, K, pad, Lout, 0, s, pair, smem, grid, units)
                                  : plan_fwd128(B, Lout, Cout, Cin, K, K - 1 - pad, L, 1, s, pair, smem, grid, units);
    if (rc != 0) return CS_OK;
    int cols = 32;
    while (cols < s.NT * s.MT) cols <<= 1;
    const int v[16] = {pair ? 2 : 1, s.NT, s.MT, s.AS, s.WS, s.slot, s.tma_store, (int)smem, (int)grid.x, (int)grid.y, (int)grid.z, cols,
                       s.total, units, 0, 0};
    memcpy(out, v, sizeof(v));
    return CS_OK;
}

}  // namespace cs
