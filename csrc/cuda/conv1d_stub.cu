// temporary stubs (replaced by conv1d_umma.cu)
#include "tiny_ecg.cuh"
namespace cs {
#define UNSUP(name) set_error(name ": not built yet"); return CS_ERR_UNSUPPORTED
int conv1d_fwd(const void*, const void*, const float*, void*, int, int, int, int, int, int, int, int, cudaStream_t) { UNSUP("conv1d_fwd"); }
int conv1d_prepare_dgrad_weight(const void*, void*, int, int, int, cudaStream_t) { UNSUP("prepare_dgrad"); }
int conv1d_wgrad(const void*, const void*, float*, float*, int, int, int, int, int, int, int, cudaStream_t) { UNSUP("wgrad"); }
int conv1d_cin1_fwd(const float*, const float*, const float*, void*, int, int, int, int, int, int, int, cudaStream_t) { UNSUP("cin1_fwd"); }
int conv1d_cin1_wgrad(const float*, const void*, float*, float*, int, int, int, int, int, int, cudaStream_t) { UNSUP("cin1_wgrad"); }
}
