// Pieces shared by the tcgen05 TinyECG kernels (the one-step kernel in tiny_ecg_umma.cu and the persistent
// FL-round kernel in tiny_ecg_round.cu): the TMEM-shaped gradient scratch layout, operand packing helpers,
// the shared-memory carve-up of the per-sample working set and small warp-level reductions.
#pragma once
#include "tiny_ecg.cuh"
#include "common/tcgen05.cuh"

namespace cs {

using namespace tc;

#ifndef CS_TINY_WGRAD_M
#define CS_TINY_WGRAD_M 64       // M of the weight-gradient MMAs: 64 rows = 8 tap groups x 8 channels
#endif
// TMEM lane that holds accumulator row r: M=128 -> lane r; M=64 -> 16 lanes per 32-lane subpartition
__device__ __forceinline__ int wgrad_row_of_thread(int t) {
#if CS_TINY_WGRAD_M == 128
    return t & 127;
#else
    return (t & 31) < 16 ? ((t >> 5) & 3) * 16 + (t & 15) : -1;
#endif
}

// Gradient scratch of the training kernel: TMEM-shaped so each thread adds whole rows with 16-byte
// vector reductions, and replicated kScratchReps times (CTA b uses replica b % reps) so that same-address
// traffic in L2 is spread out.  The SGD tail sums the replicas and un-permutes.
//   [0, 1280)      conv2 weight grad  row (k*8 + c8) x col (j*16 + co)      -> dW2[co][8j+c8][k]
//   [1280, 1536)   bias MMA           row k x col n: n<16 dW1[n][k] (k<7), db1[n] (k=7); n>=16: db2[n-16] (k=7)
//   [1536, 1672)   head weight + bias grads (17*C <= 136)
constexpr int kScratchReps = 4;
constexpr int kScratchStride = 1792;
constexpr int kScrGB = 1280, kScrGH = 1536;
__host__ __device__ constexpr int scratch_offset(int i) {     // flat parameter index -> scratch offset
    if (i < TinyLayout::B1) return kScrGB + (i % 7) * 32 + i / 7;
    if (i < TinyLayout::W2) return kScrGB + 7 * 32 + (i - TinyLayout::B1);
    if (i < TinyLayout::B2) {
        const int q = i - TinyLayout::W2, co = q / 80, ci = (q / 5) & 15, k = q % 5;
        return (k * 8 + (ci & 7)) * 32 + (ci >> 3) * 16 + co;
    }
    if (i < TinyLayout::WH) return kScrGB + 7 * 32 + 16 + (i - TinyLayout::B2);
    return kScrGH + (i - TinyLayout::WH);
}
// parameter index -> scratch offset as a table (the tail gathers 13 values per thread; the divisions in
// scratch_offset() cost more than the whole L2 round trip when done on the critical path)
constexpr int kOfsTableLen = TinyLayout::WH + 17 * kTinyMaxC;
struct alignas(16) OfsTable { uint16_t v[(kOfsTableLen + 7) & ~7]; };
constexpr OfsTable make_ofs_table() {
    OfsTable t{};
    for (int i = 0; i < kOfsTableLen; ++i) t.v[i] = (uint16_t)scratch_offset(i);
    return t;
}
static __device__ const OfsTable g_ofs_table = make_ofs_table();

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// sum of SETS rotating 32-column accumulator sets (warp-collective TMEM loads)
template <int SETS> __device__ __forceinline__ void load_sum_sets(uint32_t taddr, float (&out)[32]) {
#pragma unroll
    for (int n = 0; n < 32; ++n) out[n] = 0.0f;
#pragma unroll
    for (int q = 0; q < SETS; ++q) {
        uint32_t v[32];
        tmem_ld32(taddr + q * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int n = 0; n < 32; ++n) out[n] += __uint_as_float(v[n]);
    }
}

// 16 columns summed over SETS rotating accumulator sets that sit 32 columns apart
template <int SETS> __device__ __forceinline__ void load_sum_sets16(uint32_t taddr, float (&out)[16]) {
#pragma unroll
    for (int n = 0; n < 16; ++n) out[n] = 0.0f;
#pragma unroll
    for (int q = 0; q < SETS; ++q) {
        uint32_t v[16];
        tmem_ld16(taddr + q * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int n = 0; n < 16; ++n) out[n] += __uint_as_float(v[n]);
    }
}

template <int TILES> struct USmem {
    static constexpr int ROWS = TILES * 128 + 16;          // 8 halo rows on top, >= 8 below
    static constexpr int PLANE = ROWS * 16;                // bytes
    static constexpr int OFF_BAR = 0;                      // 3 mbarriers, tmem slot, flags
    static constexpr int OFF_F32 = 64;                     // fp32 scalars (see F32_* below)
    static constexpr int F32_COUNT = 112 + 16 + 16 + 128 + 8 + 16 + 8 + 16 + 128;
    static constexpr int OFF_WT = (OFF_F32 + F32_COUNT * 4 + 127) & ~127;   // 10 weight tiles x 512 B
    static constexpr int OFF_XS = OFF_WT + 10 * 512;
    static constexpr int XS_LEN = TILES * 128 + 8;
    static constexpr int OFF_PLANES = (OFF_XS + XS_LEN * 4 + 127) & ~127;
    // plane order matters: [dz1 p0][dz1 p1][dz2 p0][dz2 p1] are the 4 N-groups of the bias MMA
    static constexpr int P_DZ1 = OFF_PLANES, P_DZ2 = P_DZ1 + 2 * PLANE, P_H1 = P_DZ2 + 2 * PLANE,
                         P_XCOL = P_H1 + 2 * PLANE;
    static constexpr int TOTAL = P_XCOL + PLANE + 512;     // slack: shifted operands read past the end
    // conv accumulators 16*TILES | conv2-wgrad 2 rotating sets x 2 planes x 16 | bias-MMA 4 rotating sets x 32
    static constexpr int TMEM_COLS = (16 * TILES + 64 + 128) <= 256 ? 256 : 512;
};

template <bool FP16> __device__ __forceinline__ uint32_t pack2(float a, float b) {
    if (FP16) { __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
template <bool FP16> __device__ __forceinline__ float round_half(float a) {
    return FP16 ? __half2float(__float2half_rn(a)) : __bfloat162float(__float2bfloat16(a));
}
template <bool FP16> __device__ __forceinline__ float2 unpack2(uint32_t u) {
    if (FP16) return __half22float2(*reinterpret_cast<__half2*>(&u));
    return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u));
}
template <bool FP16> __device__ __forceinline__ uint16_t pack1(float a) {
    if (FP16) { __half h = __float2half_rn(a); return *reinterpret_cast<uint16_t*>(&h); }
    __nv_bfloat16 h = __float2bfloat16(a);
    return *reinterpret_cast<uint16_t*>(&h);
}

// 16 per-lane partial sums -> lane j ends with the warp total of channel (j >> 1)
__device__ __forceinline__ float warp_reduce16(const float (&p)[16], int lane) {
    float a8[8], a4[4], a2[2];
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4, h2 = lane & 2;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float send = h16 ? p[j] : p[j + 8], keep = h16 ? p[j + 8] : p[j];
        a8[j] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float send = h8 ? a8[j] : a8[j + 4], keep = h8 ? a8[j + 4] : a8[j];
        a4[j] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const float send = h4 ? a4[j] : a4[j + 2], keep = h4 ? a4[j + 2] : a4[j];
        a2[j] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    const float send = h2 ? a2[0] : a2[1], keep = h2 ? a2[1] : a2[0];
    float v = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v;
}

}  // namespace cs
