// cs_tiny_ecg_train_step_noalloc — the fused TinyECG training step for callers that cannot allocate
// device memory.  (No counterpart in the reference.)
//
// Who cannot allocate?  A federated client of this library is often a guest inside someone else's
// process: an inference server or a training framework that has already handed all device memory to
// its own pool, so a cudaMalloc from a plugin either fails or steals from the host; a monitoring or
// control loop where the device-wide synchronisation that cudaMalloc / cudaFree imply is not
// acceptable between two steps; or a thin C / ctypes driver (the reference's benchmark scripts are of
// that kind) that only knows how to pass input and output buffers and has no allocator at hand.
// cs_tiny_ecg_train_step asks all of them for a zeroed gradient scratch buffer they cannot provide.
//
// This variant keeps that scratch inside the library image instead: a file-scope __device__ array of
// CS_TINY_NOALLOC_SCRATCH_FLOATS floats (28 KiB), present on every GPU that loads the library for as
// long as the process lives.  The caller passes only inputs (windows, labels, indices) and the state it
// owns anyway (parameters, momentum, ctrl, fstate).
//
// Rules that follow from sharing one array:
//   * The first call on each device resolves the array's device address (cudaGetSymbolAddress) with
//     that device current.  Make that first call outside any stream capture.
//   * Every call clears the part of the array it uses on `stream` before launching, because the array
//     still holds whatever the previous call — any caller, any model — left there.
//   * Two calls whose kernels overlap on the same GPU (different streams, or two threads) race on the
//     array and corrupt each other's gradients.  Serialise them on one stream, or give concurrent
//     trainers their own buffer through cs_tiny_ecg_train_step.
//   * Nothing here calls cudaMalloc, cudaFree or any allocating library routine.
//   * This is the slower of the two entry points when steps are issued back to back: the cudaMemsetAsync that
//     clears the array sits between consecutive step kernels, so the programmatic dependent launch that lets
//     cs_tiny_ecg_train_step start its prologue under the previous step's tail never engages here.
//   * The per-device address cache is a table of atomics: two host threads that make their first call on the
//     same device at the same time both look the address up and store the same value.
#include "tiny_ecg.cuh"

#include <string.h>

#include <atomic>

namespace cs {

static __device__ __align__(256) float g_tiny_static_scratch[CS_TINY_NOALLOC_SCRATCH_FLOATS];

// The symbol above is only a host-side handle; kernels need its device address.  One look-up per
// device, done with that device current.
static int static_scratch_ptr(float** out) {
    static std::atomic<float*> ptr[64];
    int dev = 0;
    CS_CUDA_OK(cudaGetDevice(&dev));
    CS_REQUIRE(dev >= 0 && dev < 64, "noalloc: device ordinal %d out of range", dev);
    float* cached = ptr[dev].load(std::memory_order_acquire);
    if (cached == nullptr) {
        void* p = nullptr;
        CS_CUDA_OK(cudaGetSymbolAddress(&p, g_tiny_static_scratch));
        CS_REQUIRE(p != nullptr, "noalloc: cudaGetSymbolAddress returned a null address");
        cached = static_cast<float*>(p);
        ptr[dev].store(cached, std::memory_order_release);
    }
    *out = cached;
    return CS_OK;
}

}  // namespace cs

using namespace cs;

// Test hook: overwrite the built-in array with NaNs on `stream`, as if a previous call had been torn down half way.
// The next cs_tiny_ecg_train_step_noalloc must still produce the right update (it clears what it uses first).
extern "C" int cs_debug_noalloc_poison(cs_stream_t stream) {
    float* scratch = nullptr;
    int rc = static_scratch_ptr(&scratch);
    if (rc != CS_OK) return rc;
    CS_CUDA_OK(cudaMemsetAsync(scratch, 0xFF, sizeof(float) * CS_TINY_NOALLOC_SCRATCH_FLOATS, (cudaStream_t)stream));
    return CS_OK;
}

extern "C" int cs_tiny_ecg_train_step_noalloc(const void* x, const int64_t* y, const int64_t* idx, int x_bf16, int B,
                                              int L, int C, float* params, float* mom, int32_t* ctrl, float* fstate,
                                              float lr, float momentum, float wd, int amp, int use_scaler,
                                              float growth, float backoff, int growth_interval, cs_stream_t stream) {
    CS_REQUIRE(x && y && params && mom && ctrl && fstate, "train_step_noalloc: null pointer argument");
    CS_REQUIRE(C >= 1 && B >= 1 && L >= 8, "train_step_noalloc: bad shape B=%d L=%d C=%d", B, L, C);
    CS_REQUIRE(C <= kTinyMaxC,
               "train_step_noalloc: %d classes exceed the %d the built-in %d-float scratch array is laid out for; "
               "use cs_tiny_ecg_train_step with a caller-provided buffer", C, kTinyMaxC, CS_TINY_NOALLOC_SCRATCH_FLOATS);
    const size_t need = cs_tiny_ecg_scratch_floats(C);
    CS_REQUIRE(need <= (size_t)CS_TINY_NOALLOC_SCRATCH_FLOATS,
               "train_step_noalloc: this configuration needs %zu floats of gradient scratch, the built-in array "
               "holds %d; use cs_tiny_ecg_train_step with a caller-provided buffer", need, CS_TINY_NOALLOC_SCRATCH_FLOATS);
    float* scratch = nullptr;
    int rc = static_scratch_ptr(&scratch);
    if (rc != CS_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // the array remembers the previous call: the gradient accumulators must start from zero
    CS_CUDA_OK(cudaMemsetAsync(scratch, 0, need * sizeof(float), st));

    TinyArgs a;
    memset(&a, 0, sizeof(a));
    a.x = x; a.y = y; a.idx = idx; a.x_bf16 = x_bf16; a.B = B; a.L = L; a.C = C; a.amp = amp;
    a.params = params; a.mom = mom; a.gacc = scratch; a.ctrl = ctrl; a.fstate = fstate;
    a.lr = lr; a.momentum = momentum; a.wd = wd; a.use_scaler = use_scaler;
    a.growth = growth; a.backoff = backoff; a.interval = growth_interval;
    const bool tensor_core_path = amp != CS_AMP_FP32 && tiny_umma_supported(L, C);
    return tensor_core_path ? tiny_umma_train(a, st) : tiny_f32_train(a, st);
}
