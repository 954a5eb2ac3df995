/*
 * crossscale.h — C ABI of the CrossScale-B200 native libraries.
 *
 *   libcrossscale_cuda.so   hand-written sm_100a kernels + launchers      (csrc/cuda)
 *   libcrossscale_cpu.so    OpenMP/AVX2 CPU conv1d kernel                  (csrc/cpu)
 *   libcrossscale_rt.so     C++ runtime: mmap shard reader, pinned ring    (csrc/runtime)
 *
 * Conventions: every cs_* function returns 0 on success and a non-zero code on failure;
 * cs_last_error() then returns a message (thread local).  Pointers are device pointers unless the
 * comment says "host".  Launchers are asynchronous on `stream` (a cudaStream_t passed as void*)
 * and allocate nothing unless stated, so they can be captured into CUDA graphs.
 *
 * "ref:" comments cite the reference component (path:line under /root/reference) each entry
 * point replaces.
 */
#ifndef CROSSSCALE_H_
#define CROSSSCALE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CS_OK 0
#define CS_ERR_INVALID 1
#define CS_ERR_CUDA 2
#define CS_ERR_UNSUPPORTED 3

typedef void* cs_stream_t; /* cudaStream_t */

/* precision modes of the fused TinyECG kernels */
#define CS_AMP_FP32 0 /* G0: fp32 (step kernel: FFMA; round kernel: tf32 MMAs)   */
#define CS_AMP_BF16 1 /* G1: bf16 operands, fp32 accumulate (tcgen05)         */
#define CS_AMP_FP16 2 /* G1 reference-faithful: fp16 operands + loss scaling  */

const char* cs_last_error(void);
int cs_version(void);
/* 1 when the current device is compute capability 10.x (tcgen05 paths usable) */
int cs_device_is_sm100(void);

/* ---------------------------------------------------------------------------------------------
 * K11  on-device batch gather.   ref: Module_3/shard_dataset.py:132-136 (x_gpu[sel], y_gpu[sel])
 * xo[b,:] = x[idx[b],:], yo[b] = y[idx[b]].  x is [N,L] fp32; out_bf16 != 0 writes xo as bf16.
 * ------------------------------------------------------------------------------------------- */
int cs_gather_rows(const float* x, const int64_t* y, const int64_t* idx, void* xo, int64_t* yo,
                   int B, int L, int out_bf16, cs_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K6b  fused TinyECG kernels.    ref: Module_3/tiny_ecg_model.py:16-29 (model),
 *      TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:103-132 (train_step_G0 / train_step_G1)
 *
 * `params` is the flat fp32 arena in state_dict order (1424 + 17*C floats, C <= 8).
 * x is [rows, L]; when idx != NULL sample b reads row idx[b] (fused gather), else row b.
 * x_bf16 != 0: x holds bf16 instead of fp32.
 * ------------------------------------------------------------------------------------------- */

/* bytes of dynamic shared memory the fused kernels need for window length L (0 if unsupported) */
size_t cs_tiny_ecg_smem_bytes(int L, int C, int amp);

/* floats the `gacc` gradient scratch of cs_tiny_ecg_train_step must hold (zeroed once by the caller;
 * every step leaves it zero).  The tcgen05 kernel keeps 4 TMEM-shaped replicas of 1792 floats there. */
size_t cs_tiny_ecg_scratch_floats(int C);

/* logits[B,C] = TinyECG(x) */
int cs_tiny_ecg_forward(const void* x, const int64_t* idx, int x_bf16, int B, int L, int C,
                        const float* params, float* logits, int amp, cs_stream_t stream);

/* grad[0:P] += dLoss/dparams given dlogits[B,C] (forward is recomputed on chip) */
int cs_tiny_ecg_backward(const void* x, const int64_t* idx, int x_bf16, int B, int L, int C,
                         const float* params, const float* dlogits, float* grad, int amp,
                         cs_stream_t stream);

/* One whole local training step in ONE launch: gather + forward + mean cross-entropy (x loss
 * scale) + backward + grad unscale/inf-check + SGD-momentum update + scaler update.
 *   y      labels, indexed like x (y[idx[b]] or y[b])
 *   mom    momentum buffer  [P]        gacc   gradient scratch, cs_tiny_ecg_scratch_floats(C) floats,
 *                                              zero on first use; the kernel leaves it zero again
 *   ctrl   int32[8]  {ticket, growth_tracker, steps_done, skipped_steps, found_inf, error_bits,
 *                      samples_drawn (int64 in slots 6-7)}      error_bits: CS_TINY_ERR_*
 *   fstate fp32[8]   {loss_sum, loss_scale, last_loss, loss_step_scratch, ...}
 * SGD: buf = momentum*buf + (g + wd*p); p -= lr*buf            (torch.optim.SGD, ref :181)
 * scaler (use_scaler != 0): torch.cuda.amp.GradScaler semantics  (ref :121-127)            */
int cs_tiny_ecg_train_step(const void* x, const int64_t* y, const int64_t* idx, int x_bf16,
                           int B, int L, int C, float* params, float* mom, float* gacc,
                           int32_t* ctrl, float* fstate, float lr, float momentum, float wd,
                           int amp, int use_scaler, float growth, float backoff,
                           int growth_interval, cs_stream_t stream);

/* The same training step for callers that cannot allocate device memory (a client library living
 * inside another framework's pre-reserved memory pool, a control loop that must not hit the device
 * synchronisation of cudaMalloc/cudaFree, a bare C/ctypes driver that only passes inputs and outputs).
 * This entry point has no counterpart in the reference.
 * The gradient scratch is a CS_TINY_NOALLOC_SCRATCH_FLOATS-float (28 KiB) __device__ array built into
 * the library, resident on every GPU that loads it until the process exits; C > 8 (more scratch than
 * the array holds) is rejected with an error.  The first call on a device looks the array up — issue
 * it outside any stream capture.  All calls share the array and each clears what it uses first, so
 * calls that overlap on one GPU race: keep them on one stream.  Allocates nothing. */
#define CS_TINY_NOALLOC_SCRATCH_FLOATS 7168
int cs_tiny_ecg_train_step_noalloc(const void* x, const int64_t* y, const int64_t* idx, int x_bf16,
                                   int B, int L, int C, float* params, float* mom, int32_t* ctrl,
                                   float* fstate, float lr, float momentum, float wd, int amp,
                                   int use_scaler, float growth, float backoff, int growth_interval,
                                   cs_stream_t stream);

/* device-side error bits the fused TinyECG kernels OR into ctrl[5] (0 = healthy).  The Python wrappers
 * read the word whenever they read the round's loss and raise. */
#define CS_TINY_ERR_LABEL 1 /* a label outside [0, C): the sample was trained as class 0                  */
#define CS_TINY_ERR_MBAR  2 /* an mbarrier wait inside the CTA timed out (protocol bug / hardware fault)   */
#define CS_TINY_ERR_GRID  4 /* the grid-wide step barrier timed out                                        */
#define CS_TINY_ERR_PEER  8 /* fused aggregation: a peer's flag never arrived; the local model was kept    */

/* K6b persistent: ONE cooperative launch = `steps` local training steps + (optionally) the FedAvg aggregation.
 * ref: the whole round loop body, TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:186-231 (Bcast, 50 x train_step_G1,
 * Allreduce-mean) - here 50 steps x ~40 launches, 50 synchronisations and 36 host-staged collectives become one kernel.
 * tcgen05 path: amp = CS_AMP_BF16 / CS_AMP_FP16 (16-bit operands), or CS_AMP_FP32 = fp32 storage with the conv2 operands fed to
 * the tensor core as TF32 (kind::tf32, windows up to 512 samples; what cuDNN does for train_step_G0's fp32 convolutions).  Per step: fused forward/backward
 * per sample, gradients summed through a rotating global scratch, one grid-wide barrier, then every CTA applies
 * SGD-momentum (and the GradScaler rules) to its own shared-memory copy of the model.
 *   sample_mode 0: sample b of step s reads row idx[s*B + b]            (idx int64 [steps*B])
 *               1: row s*B + b of x (a staged block of steps*B rows; n_rows = rows in the block or 0)
 *               2: in-kernel per-epoch permutation of the n_rows dataset rows (ragged tail dropped, like
 *                  make_gpu_batch_iter, Module_3/shard_dataset.py:118-136): row = cs_perm_index(pos % usable, n_rows,
 *                  seed, pos / usable), usable = n_rows / B * B, pos = ctrl[6..7] (int64), advanced by steps*B per call
 *   gacc   cs_tiny_ecg_round_scratch_floats(C) floats, zero before the first call; left zero
 *   ctrl / fstate / SGD / scaler: as cs_tiny_ecg_train_step (fstate[0] += sum of the steps' mean losses)
 *   comm != NULL: the round ends in the one-shot aggregation of cs_fedavg_aggregate (same arena layout, flags and
 *   epoch word: the two are interchangeable on a channel): theta <- prev + server_lr * sum_i weight_i (theta_i - prev),
 *   written to params and prev; momentum is zeroed when reset_momentum != 0.  comm == NULL: params/mom are written back.
 * Every CTA of the launch must be co-resident (cooperative launch; grid = min(B, 2 x SMs)). */
size_t cs_tiny_ecg_round_scratch_floats(int C);
struct cs_comm;
int cs_tiny_ecg_train_round(const void* x, const int64_t* y, const int64_t* idx, int x_bf16, int64_t n_rows,
                            int B, int L, int C, int steps, int sample_mode, uint64_t seed, float* params,
                            float* mom, float* gacc, int32_t* ctrl, float* fstate, float lr, float momentum,
                            float wd, int amp, int use_scaler, float growth, float backoff, int growth_interval,
                            struct cs_comm* comm, float* prev, float weight, int wire_bf16, float server_lr,
                            int reset_momentum, size_t arena_off, int channel, cs_stream_t stream);
/* the same for `world` emulated clients on ONE device (arrays indexed by client; blockIdx.y = client; all CTAs of all
 * clients co-resident, so B * world <= 2 x SMs): the cross-client flag protocol runs inside one cooperative launch */
int cs_tiny_ecg_train_round_local(struct cs_comm** group, int world, const void* const* x, const int64_t* const* y,
                                  const int64_t* const* idx, int x_bf16, const int64_t* n_rows, int B, int L, int C,
                                  int steps, int sample_mode, const uint64_t* seeds, float* const* params,
                                  float* const* mom, float* const* gacc, int32_t* const* ctrl, float* const* fstate,
                                  float lr, float momentum, float wd, int amp, int use_scaler, float growth,
                                  float backoff, int growth_interval, float* const* prev, const float* weights,
                                  int wire_bf16, float server_lr, int reset_momentum, size_t arena_off, int channel,
                                  cs_stream_t stream);
/* the sampler's bijection on [0, n) (host-callable mirror of the device function) */
unsigned long long cs_perm_index(unsigned long long i, unsigned long long n, unsigned long long seed,
                                 unsigned long long epoch);

/* ---------------------------------------------------------------------------------------------
 * K5   head-less cross entropy + loss scale, and AMP unscale / inf-check / scale update
 *      ref: F.cross_entropy (:107,:123), GradScaler (:125-127)
 * ------------------------------------------------------------------------------------------- */
/* loss_out[0] = mean CE (unscaled); dlogits = scale*(softmax-onehot)/B; scale read from *scale_ptr
 * (NULL = 1.0).  logits fp32 [B,C], C <= 64. */
int cs_cross_entropy_fwd_bwd(const float* logits, const int64_t* y, int B, int C,
                             const float* scale_ptr, float* loss_out, float* dlogits,
                             cs_stream_t stream);

/* K6  flat multi-tensor SGD-momentum with fused unscale + found-inf skip + scaler update.
 * g is a flat fp32 gradient [n]; fstate/ctrl as in cs_tiny_ecg_train_step (may be NULL when
 * use_scaler == 0).  bf16_shadow (optional) receives bf16(p) after the update.
 * Two launches internally when use_scaler != 0 (inf-check reduce, then update).             */
int cs_sgd_momentum(float* p, float* mom, const float* g, int64_t n, float lr, float momentum,
                    float wd, int use_scaler, int32_t* ctrl, float* fstate, float growth,
                    float backoff, int growth_interval, void* bf16_shadow, cs_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K1-K3  conv1d implicit GEMM (channels-last bf16, fp32 accumulate in TMEM, TMA-fed tcgen05)
 *        ref: nn.Conv1d -> cuDNN (tiny_ecg_model.py:17-20); CPU C13 for the microbench
 *   x  [B, L, Cin]  bf16     w  [K, Cout, Cin] bf16 (tap-major)     y [B, Lout, Cout] bf16
 *   stride 1 or 2, zero padding `pad`; Cin, Cout multiples of 16 (Cin<=1024, Cout<=1024).
 *   epilogue: + bias (fp32, may be NULL), optional ReLU.
 * ------------------------------------------------------------------------------------------- */
int cs_conv1d_fwd(const void* x, const void* w, const float* bias, void* y, int B, int L, int Cin,
                  int Cout, int K, int pad, int stride, int relu, cs_stream_t stream);
/* cs_conv1d_fwd + the training-mode BatchNorm statistics of its output in ONE launch: the epilogue sums the
 * bf16 values it stores per channel, and the last CTA to finish turns the totals into mean / rstd / running
 * statistics / affine coefficients exactly like cs_bn_stats_finalize (same persistent workspace contract:
 * acc [2*Cout] fp32 + ticket [1] int32, zero before the first call, left zero).  Wide layers only
 * (Cin % 64 == 0, Cout % 64 == 0): returns CS_ERR_UNSUPPORTED and launches nothing otherwise — then call
 * cs_conv1d_fwd followed by cs_bn_stats_finalize. */
int cs_conv1d_fwd_bn(const void* x, const void* w, const float* bias, void* y, int B, int L, int Cin,
                     int Cout, int K, int pad, int stride, int relu, float* acc, int32_t* ticket,
                     float* mean, float* rstd, float* running_mean, float* running_var, float eps,
                     float momentum, const float* gamma, const float* beta, float* coef,
                     cs_stream_t stream);
/* Which kernel variant and tile configuration the dispatch would use for a shape; launches nothing and needs no
 * GPU (SM count defaults to 148 without one).  direction: 0 forward, 1 data gradient, 2 weight gradient.
 * out16 (forward / data gradient): {variant: 0 plane kernel | 1 swizzled rows | 2 cta_group::2 pairs, NT, MT,
 * activation stages, weight stages, slot, tma_store, shared-memory bytes, grid x, y, z, TMEM columns per CTA,
 * tiles, units, 0, 0}; (weight gradient): {variant 0 | 1, NC, taps per CTA, tap groups, splits, slot, 0,
 * shared-memory bytes, grid x, y, z, 512, tiles, 0, 0, 0}. */
int cs_conv1d_plan(int B, int L, int Cin, int Cout, int K, int pad, int direction, int* out16);
/* dx = conv_transpose(dy, w):  dy [B,Lout,Cout] bf16 -> dx [B,L,Cin] bf16.  wt is the
 * tap-flipped, channel-transposed weight [K, Cin, Cout] (cs_conv1d_prepare_dgrad_weight). */
int cs_conv1d_dgrad(const void* dy, const void* wt, void* dx, int B, int L, int Cin, int Cout,
                    int K, int pad, int stride, cs_stream_t stream);
int cs_conv1d_prepare_dgrad_weight(const void* w, void* wt, int K, int Cout, int Cin,
                                   cs_stream_t stream);
/* Same dx, read straight from the FORWARD weights w [K, Cout, Cin] (the tile is used as an
 * MN-major operand, taps in reverse): no transposed copy.  Covers Cin % 64 == 0 && Cout % 64 == 0
 * with windows longer than 16 rows; returns CS_ERR_UNSUPPORTED (and launches nothing) otherwise —
 * the caller then falls back to cs_conv1d_prepare_dgrad_weight + cs_conv1d_dgrad. */
int cs_conv1d_dgrad_direct(const void* dy, const void* w, void* dx, int B, int L, int Cin,
                           int Cout, int K, int pad, int stride, cs_stream_t stream);
/* dw[K,Cout,Cin] (fp32, accumulated: caller zeroes) += sum_{b,l} dy[b,l,co]*x[b,l*s+k-pad,ci];
 * dbias[Cout] (fp32, optional) += sum dy. */
int cs_conv1d_wgrad(const void* x, const void* dy, float* dw, float* dbias, int B, int L, int Cin,
                    int Cout, int K, int pad, int stride, cs_stream_t stream);
/* first layer (Cin = 1): x [B,L] fp32 -> y [B,Lout,Cout] bf16 on FFMA, + bias, optional ReLU */
int cs_conv1d_cin1_fwd(const float* x, const float* w, const float* bias, void* y, int B, int L,
                       int Cout, int K, int pad, int stride, int relu, cs_stream_t stream);
int cs_conv1d_cin1_wgrad(const float* x, const void* dy, float* dw, float* dbias, int B, int L,
                         int Cout, int K, int pad, int stride, cs_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K4   BN / ReLU / residual / global-average-pool (channels-last bf16 activations)
 * ------------------------------------------------------------------------------------------- */
/* per-channel sum and sum of squares over [rows, C] -> stats[2*C] fp32 (accumulated) */
int cs_bn_stats(const void* x, float* stats, int64_t rows, int C, cs_stream_t stream);
/* y = relu?( (x-mean)*rstd*gamma + beta (+ residual) ); mean/rstd fp32 [C] */
int cs_bn_apply(const void* x, const void* residual, void* y, const float* mean,
                const float* rstd, const float* gamma, const float* beta, int64_t rows, int C,
                int relu, cs_stream_t stream);
/* backward of cs_bn_apply (training-mode batch statistics):
 *   pass 1 (cs_bn_bwd_reduce): dgamma, dbeta accumulated (fp32); relu mask from y
 *   pass 2 (cs_bn_bwd_apply):  dx (bf16), and dres = masked dy when residual was used */
int cs_bn_bwd_reduce(const void* x, const void* y, const void* dy, const float* mean,
                     const float* rstd, float* dgamma, float* dbeta, int64_t rows, int C, int relu,
                     cs_stream_t stream);
int cs_bn_bwd_apply(const void* x, const void* y, const void* dy, const float* mean,
                    const float* rstd, const float* gamma, const float* dgamma,
                    const float* dbeta, void* dx, void* dres, int64_t rows, int C, int relu,
                    cs_stream_t stream);
/* mean/rstd [C] from the stats of cs_bn_stats over `rows` rows; updates running_mean/var (may be NULL) like
 * nn.BatchNorm1d (momentum, unbiased variance) */
int cs_bn_finalize(const float* stats, float* mean, float* rstd, float* running_mean,
                   float* running_var, int C, int64_t rows, float eps, float momentum,
                   cs_stream_t stream);
/* One-launch variants built on a persistent workspace: `acc` [2*C] fp32 + `ticket` [1] int32, zero before the
 * first call and left zero by every call (the last block to finish consumes the totals and clears them), so
 * a layer needs no fill, finalize or gradient-add launches.  One workspace per layer and stream.
 *   cs_bn_stats_finalize = cs_bn_stats + cs_bn_finalize; with gamma/beta/coef it also writes the forward
 *                          coefficients coef[0:C] = rstd*gamma, coef[C:2C] = beta - mean*rstd*gamma
 *   cs_bn_bwd_reduce_ws  = cs_bn_bwd_reduce into dgamma/dbeta (no zero fill needed); with gamma_grad /
 *                          beta_grad the same sums are added into the parameter gradients; with gamma/coef
 *                          it writes the backward coefficients [a | p | q] (3*C): dx = a*d + p*x + q
 *   cs_bn_apply_coef     : y = relu?(x*coef[c] + coef[C+c] (+ residual))
 *   cs_bn_bwd_apply_coef : d = dy masked by y > 0 (if relu); dres = d; dx = a*d + p*x + q */
int cs_bn_stats_finalize(const void* x, float* acc, int32_t* ticket, float* mean, float* rstd,
                         float* running_mean, float* running_var, int64_t rows, int C, float eps,
                         float momentum, const float* gamma, const float* beta, float* coef,
                         cs_stream_t stream);
int cs_bn_bwd_reduce_ws(const void* x, const void* y, const void* dy, const float* mean,
                        const float* rstd, float* acc, int32_t* ticket, float* dgamma, float* dbeta,
                        float* gamma_grad, float* beta_grad, const float* gamma, float* coef,
                        int64_t rows, int C, int relu, cs_stream_t stream);
int cs_bn_apply_coef(const void* x, const void* residual, void* y, const float* coef, int64_t rows,
                     int C, int relu, cs_stream_t stream);
int cs_bn_bwd_apply_coef(const void* x, const void* y, const void* dy, const float* coef, void* dx,
                         void* dres, int64_t rows, int C, int relu, cs_stream_t stream);
/* y[b,j,c] = (x[b,2j,c] + x[b,2j+1,c]) / 2, j < L/2 (bf16 channels-last); backward spreads dy/2 */
int cs_avgpool2_fwd(const void* x, void* y, int B, int L, int C, cs_stream_t stream);
int cs_avgpool2_bwd(const void* dy, void* dx, int B, int L, int C, cs_stream_t stream);
/* out[b,c] = mean_l x[b,l,c]  (fp32 out);   backward: dx[b,l,c] = dout[b,c]/L */
int cs_avgpool_fwd(const void* x, float* out, int B, int L, int C, cs_stream_t stream);
int cs_avgpool_bwd(const float* dout, void* dx, int B, int L, int C, cs_stream_t stream);
/* fp32 [B,C,L] <-> bf16 [B,L,C] layout/precision converters */
int cs_ncl_f32_to_nlc_bf16(const float* x, void* y, int B, int C, int L, cs_stream_t stream);
int cs_nlc_bf16_to_ncl_f32(const void* x, float* y, int B, int C, int L, cs_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K7-K9  in-kernel NVLink collectives over peer-mapped ("symmetric") buffers.
 *        ref: broadcast_model / fedavg_allreduce, part3_fedavg_overlap_mpi_gpu.py:75-98
 * No MPI / NCCL call on these paths.  A communicator is built from buffers every rank
 * allocated identically and exchanged through CUDA IPC handles (host side, once).
 * ------------------------------------------------------------------------------------------- */
#define CS_MAX_RANKS 8
#define CS_IPC_HANDLE_BYTES 64

typedef struct cs_comm cs_comm_t;

/* Allocate this rank's symmetric arena (`bytes`) + signal pad; writes the two IPC handles
 * (host buffers of CS_IPC_HANDLE_BYTES each).  Allocates with cudaMalloc (not capturable). */
int cs_comm_create(int rank, int world, size_t bytes, cs_comm_t** out, void* arena_handle_out,
                   void* signal_handle_out);
/* Map every peer's arena and signal pad.  `arena_handles`/`signal_handles`: host arrays of
 * world * CS_IPC_HANDLE_BYTES gathered from all ranks (any host-side transport). */
int cs_comm_connect(cs_comm_t* c, const void* arena_handles, const void* signal_handles);
/* Communicator over mappings the caller made itself (the VMM allocator below): arena[r] / signals[r] = this process's
 * mapping of rank r's arena (`bytes`) and zeroed signal pad (cs_comm_signal_bytes()); nothing is freed by cs_comm_destroy. */
int cs_comm_create_external(int rank, int world, size_t bytes, void* const* arena, void* const* signals, cs_comm_t** out);
size_t cs_comm_signal_bytes(void);

/* Own symmetric-memory allocator (csrc/cuda/symm_vmm.cu): CUDA VMM allocations exported as POSIX file descriptors (sent to
 * the peers over AF_UNIX sockets by crossscale/parallel/vmm.py), imported + mapped by every rank, optionally bound to an
 * NVSwitch multicast object (NVLS).  cs_vmm_support(): bit 0 = fd-exportable VMM allocations, bit 1 = multicast. */
int cs_vmm_support(void);
int cs_vmm_alloc(size_t bytes, int for_multicast, int world, void** ptr, size_t* alloc_bytes, unsigned long long* handle, int* fd);
int cs_vmm_import(int fd, size_t alloc_bytes, void** ptr, unsigned long long* handle);
int cs_vmm_free(void* ptr, size_t alloc_bytes, unsigned long long handle);
int cs_vmm_mc_create(int world, size_t alloc_bytes, unsigned long long* mc_handle, int* fd);
int cs_vmm_mc_import(int fd, unsigned long long* mc_handle);
int cs_vmm_mc_add_device(unsigned long long mc_handle);
int cs_vmm_mc_bind_map(unsigned long long mc_handle, unsigned long long mem_handle, size_t alloc_bytes, void** mc_ptr);
int cs_vmm_close_fd(int fd);

/* Single-process emulation for tests/1-GPU: all `world` arenas live on the current device. */
int cs_comm_create_local(int world, size_t bytes, cs_comm_t** out /* array of `world` */);
void* cs_comm_arena(cs_comm_t* c);          /* this rank's arena (device pointer)              */
void* cs_comm_peer_arena(cs_comm_t* c, int r);
size_t cs_comm_arena_bytes(cs_comm_t* c);
int cs_comm_destroy(cs_comm_t* c);

/* != 0 after a device-side barrier timed out (a peer never arrived); sticky */
int cs_comm_status(cs_comm_t* c, int* status_out);
/* how long a cross-GPU flag wait may spin before it gives up, raises the status word and SKIPS its reduction / write-back
 * (default 4000 ms, or the environment variable CS_COMM_TIMEOUT_MS at creation) */
int cs_comm_set_timeout_ms(cs_comm_t* c, long long ms);
int cs_comm_clear_status(cs_comm_t* c);
/* cap the CTAs of the aggregation / bucket all-reduce launches on this communicator (0 = size by the vector).  A collective
 * that runs UNDER backward (cs_allreduce_bucket from the gradient hooks) should leave the SMs to backward: its CTAs spin at
 * the cross-GPU barriers.  The work partition depends on the count, so every rank must set the same value. */
int cs_comm_set_max_blocks(cs_comm_t* c, int n);

/* K9: device-side barrier across ranks (one tiny kernel; st.release.sys / ld.acquire.sys flags in
 * the peers' signal pads).  `channel` (0..7) selects an independent flag set so collectives on
 * different streams do not interfere. */
int cs_comm_barrier(cs_comm_t* c, int channel, cs_stream_t stream);

/* arena bytes one aggregation of n elements needs at its offset */
size_t cs_fedavg_arena_bytes(int64_t n, int wire_bf16, int world);

/* K7: fused FedAvg.  Every rank calls it with the same n / arena_off / channel / wire format.
 *   local      this rank's fp32 parameters [n] — read, then overwritten with the new global model
 *   prev       global model at the start of the round [n] (updated in place), or NULL.
 *              With prev the wire carries weight*(local-prev) (delta encoding: bf16 only quantises
 *              the round's change) and theta <- prev + server_lr * sum_i w_i delta_i.
 *              With prev == NULL the wire carries weight*local and theta <- sum_i w_i local_i.
 *   weight     this client's aggregation weight n_i / sum_j n_j
 *   wire_bf16  1: bf16 on NVLink, 0: fp32
 *   mom        optional buffer [n] zeroed after aggregation (NULL keeps it; the reference keeps
 *              the optimizer momentum)
 * One-shot (n <= 65536): every rank pushes its contribution into every peer's arena, one flag
 * round, local reduce.  Two-shot (larger): rank r reduces slice r over P2P loads and pushes the
 * result to every peer.  Both are ONE kernel launch per rank including the cross-GPU barriers. */
int cs_fedavg_aggregate(cs_comm_t* c, float* local, float* prev, int64_t n, float weight,
                        int wire_bf16, float server_lr, float* mom, size_t arena_off, int channel,
                        cs_stream_t stream);
/* Same kernel for an emulated group (cs_comm_create_local): ONE cooperative launch covers all
 * ranks (blockIdx.y = rank); host arrays of `world` device pointers / weights. */
int cs_fedavg_aggregate_local(cs_comm_t** group, int world, float* const* local,
                              float* const* prev, int64_t n, const float* weights, int wire_bf16,
                              float server_lr, float* const* mom, size_t arena_off, int channel,
                              cs_stream_t stream);

/* K7, zero-staging variant for large models whose fp32 parameters LIVE in the arena at `param_off` (same
 * offset on every rank): rank r reads slice r of every peer's parameters over NVLink, forms
 * sum_j weights[j]*theta_j (then prev + server_lr*(avg - prev) when prev != NULL) and stores it into slice r
 * of every replica — (P-1)/P*n floats per NVLink direction, one HBM pass, two cross-GPU barriers, ONE
 * launch per rank.  `weights` is a host array of `world` floats (identical on all ranks); n % 4 == 0. */
int cs_fedavg_aggregate_inplace(cs_comm_t* c, size_t param_off, float* prev, int64_t n,
                                const float* weights, float server_lr, float* mom, int channel,
                                cs_stream_t stream);
int cs_fedavg_aggregate_inplace_local(cs_comm_t** group, int world, size_t param_off,
                                      float* const* prev, int64_t n, const float* weights,
                                      float server_lr, float* const* mom, int channel,
                                      cs_stream_t stream);

/* K7 over NVLS: `local` is this rank's replica inside a symmetric buffer that is also mapped at the NVSwitch
 * multicast address `multicast` (e.g. torch.distributed._symmetric_memory: buffer_ptrs[rank] / multicast_ptr).
 * Rank r pulls the switch-reduced slice r (multimem.ld_reduce), applies post_scale and the optional server
 * update, and broadcasts it to every replica with multimem.st.  prescale != 1 multiplies the local replica
 * first (unequal client weights: prescale = w_i * world, post_scale = 1/world).  The communicator only
 * supplies the cross-GPU barrier flags.  ONE launch per rank; n % 4 == 0. */
int cs_fedavg_aggregate_nvls(cs_comm_t* c, float* local, void* multicast, float* prev, int64_t n,
                             float prescale, float post_scale, float server_lr, float* mom,
                             int channel, cs_stream_t stream);

/* K7 bucketed: optimizer step + in-place FedAvg + bf16 shadow of ONE gradient bucket in one launch, meant to be issued on a
 * side stream as soon as the round's last backward has produced the bucket, so the aggregation of bucket i crosses NVLink
 * while backward still computes bucket i+1.  ref: everything in comm_ms is exposed in the reference
 * (TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:188-190,209-214: Allreduce only after the last step's optimizer.step()).
 * The bucket's parameters live in the symmetric arena at byte offset param_off (same offset on every rank).
 *   do_sgd != 0: buf = momentum*buf + (inv_scale*grad + wd*p); p -= lr*buf   on the whole bucket (grad, mom: bucket-local
 *   pointers), then barrier, then rank r forms sum_j weights[j]*p_j (+ server update against prev when prev != NULL) on
 *   slice r and stores it into slice r of every replica, barrier, then shadow_bf16 (optional) = bf16(bucket).
 * max_blocks > 0 caps the CTAs of the launch (it shares the GPU with backward). */
int cs_fedavg_bucket(cs_comm_t* c, size_t param_off, const float* grad, float* mom, float* prev, void* shadow_bf16,
                     int64_t n, const float* weights /* host, world floats */, float lr, float momentum, float wd,
                     float inv_scale, float server_lr, int do_sgd, int channel, int max_blocks, cs_stream_t stream);
int cs_fedavg_bucket_local(cs_comm_t** group, int world, size_t param_off, const float* const* grad, float* const* mom,
                           float* const* prev, void* const* shadow_bf16, int64_t n, const float* weights, float lr,
                           float momentum, float wd, float inv_scale, float server_lr, int do_sgd, int channel,
                           cs_stream_t stream);

/* K8: gradient all-reduce of one bucket, in place: grad <- scale * sum_ranks grad  (fp32 [n]).
 * `arena_off` / `channel` let several buckets be in flight on different streams. */
int cs_allreduce_bucket(cs_comm_t* c, float* grad, int64_t n, float scale, int wire_bf16,
                        size_t arena_off, int channel, cs_stream_t stream);
int cs_allreduce_bucket_local(cs_comm_t** group, int world, float* const* grads, int64_t n,
                              float scale, int wire_bf16, size_t arena_off, int channel,
                              cs_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * K10  pinned-host -> device staging (cudaMemcpyAsync on the given copy stream) + optional
 *      on-device z-score normalise / fp32->bf16 repack.
 *      ref: dead G1 code part3_mpi_gpu_train.py:222-229; LABL ring labl_loader:30-136
 * ------------------------------------------------------------------------------------------- */
int cs_stage_h2d(void* dst_device, const void* src_pinned_host, size_t bytes, cs_stream_t stream);
int cs_stage_d2h(void* dst_pinned_host, const void* src_device, size_t bytes, cs_stream_t stream);
/* The reference's native CPU kernel on the GPU, same contract (device pointers): single-channel VALID cross-correlation,
 * fp32, y[b,i] = sum_k x[b,i+k]*w[k], y is [batch, L-K+1].  ref: conv1d_batch_omp_simd, Module_2/conv1d_openmp_simd.c:21-61
 * (the GPU arm of the Part-2 benchmark, Module_2/benchmark_part_2.py:12-19). */
int cs_conv1d_batch_valid(const float* x, const float* w, float* y, int batch, int L, int K, cs_stream_t stream);

/* y[b,:] = (x[b,:]-mean_b)/(std_b+1e-8) per window; out_bf16 selects the output type */
int cs_zscore_rows(const float* x, void* y, int B, int L, int out_bf16, cs_stream_t stream);

/* =============================================================================================
 * libcrossscale_cpu.so
 * C13: single-channel valid cross-correlation, batched, fp32 — same symbol and signature as the
 * reference (Module_2/conv1d_openmp_simd.c:21-28):  y[b,i] = sum_k x[b,i+k]*w[k], outL = L-K+1
 * ============================================================================================= */
void conv1d_batch_omp_simd(const float* x, const float* w, float* y, int batch, int L, int K,
                           int nthreads);
/* which ISA path the dispatcher picked: 0 scalar, 2 AVX2+FMA, 5 AVX-512 */
int cs_cpu_conv1d_isa(void);

/* =============================================================================================
 * libcrossscale_rt.so — LABL-style pinned ring prefetcher (C8-C10), native threads.
 * The caller owns the slot buffers (pinned host memory, num_slots x batch x L elements).
 * ============================================================================================= */
typedef struct cs_prefetcher cs_prefetcher_t;
/* out_bf16: slots hold bf16 instead of fp32.  loop != 0: restart from shard 0 at end of data.
 * label_slots (optional, host int64 [num_slots][batch]) are filled from .lbl sidecars (0 if none). */
cs_prefetcher_t* cs_prefetcher_create(const char* const* shard_paths, int num_shards, int batch,
                                      int num_slots, void* const* slot_ptrs,
                                      int64_t* const* label_slots, int normalize, int out_bf16,
                                      int loop);
int cs_prefetcher_window_len(cs_prefetcher_t* p);
int cs_prefetcher_start(cs_prefetcher_t* p);
/* Blocks until a slot is full. Returns 1 and sets slot/n_filled/fill_ms; 0 at end of data;
 * -1 on error (message via cs_rt_last_error). */
int cs_prefetcher_next(cs_prefetcher_t* p, int* slot, int* n_filled, double* fill_ms);
int cs_prefetcher_recycle(cs_prefetcher_t* p, int slot);
void cs_prefetcher_destroy(cs_prefetcher_t* p);
const char* cs_rt_last_error(void);

/* File-system rendezvous (libcrossscale_rt.so): all-gather of a small blob per rank through a shared
 * directory — the one-time exchange of CUDA IPC handles for C callers that do not run under
 * torch.distributed.  ref: MPI.COMM_WORLD bootstrap, part3_fedavg_overlap_mpi_gpu.py:149-151 */
int cs_rendezvous_allgather(const char* dir, const char* tag, int rank, int world, const void* blob,
                            size_t blob_bytes, void* out, double timeout_s);
int cs_rendezvous_cleanup(const char* dir, const char* tag, int rank);
const char* cs_rendezvous_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* CROSSSCALE_H_ */
