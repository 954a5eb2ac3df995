// Device-visible communicator state and the system-scope flag primitives shared by the in-kernel NVLink
// collectives (fedavg.cu) and by kernels that end in a collective (the persistent FL-round kernel).
#pragma once
#include "common/cs_common.cuh"

namespace cs {

constexpr int kMaxRanks = CS_MAX_RANKS;
constexpr int kChannels = 8;          // independent barrier channels (streams / buckets in flight)
constexpr int kMaxBlocks = 160;       // blocks per collective kernel (a block only ever waits for the same block on peers)
constexpr int kCommThreads = 512;
constexpr size_t kSignalWords = (size_t)kChannels * kMaxBlocks * kMaxRanks;
constexpr int64_t kOneShotMaxElems = 64 * 1024;   // 256 KiB of fp32

struct CommDev {                      // lives in device memory, one per rank
    int rank, world;
    char* arena[kMaxRanks];           // peer-mapped arenas (arena[rank] is local)
    uint32_t* signals[kMaxRanks];     // peer-mapped signal pads
    uint32_t* epochs;                 // local: [kChannels][kMaxBlocks] barrier epochs
    int* status;                      // this rank's status word (host-mapped pinned memory): != 0 after a barrier timeout
    size_t arena_bytes;
    unsigned long long timeout_ns;    // how long a cross-GPU flag wait may spin before it gives up
};

}  // namespace cs

struct cs_comm {
    int rank = 0, world = 1;
    size_t bytes = 0;
    int device = 0;
    bool emulated = false;            // all ranks on one device (tests / single GPU)
    void* arena = nullptr;
    uint32_t* signals = nullptr;
    uint32_t* epochs = nullptr;
    int* status = nullptr;            // device alias of status_host
    int* status_host = nullptr;       // pinned + mapped: the host reads it without a copy or a synchronisation of its own
    unsigned long long timeout_ns = 4000000000ull;
    int max_blocks = 0;               // > 0: cap on the CTAs of every collective launched on this communicator (same on all ranks)
    void* peer_arena[cs::kMaxRanks] = {};
    uint32_t* peer_signals[cs::kMaxRanks] = {};
    bool opened[cs::kMaxRanks] = {};
    bool external = false;            // arena / signal mappings are owned by the caller (VMM allocator): never freed here
    cs::CommDev* dev = nullptr;       // device copy of the context
    cs_comm** group = nullptr;        // emulation: all members (owned by member 0)
    cs::CommDev* group_dev = nullptr; // emulation: contiguous array of `world` contexts
};

namespace cs {

// ------------------------------------------------------------------------------------------------
// device-side primitives
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ unsigned long long comm_timeout_ns(const CommDev& c) { return c.timeout_ns; }

// wire formats of the one-shot collectives
template <bool BF16> struct Wire;
template <> struct Wire<false> {
    using T = float;
    __device__ static T enc(float v) { return v; }
    __device__ static float dec(T v) { return v; }
};
template <> struct Wire<true> {
    using T = __nv_bfloat16;
    __device__ static T enc(float v) { return __float2bfloat16(v); }
    __device__ static float dec(T v) { return __bfloat162float(v); }
};

}  // namespace cs
