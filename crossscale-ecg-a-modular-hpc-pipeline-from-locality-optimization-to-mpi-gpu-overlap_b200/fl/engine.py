"""Local-step engines and the FedAvg round loop.

Parity map:
  * ``train_step_G0`` / ``train_step_G1``      ``TRUE_FL_M3/part3_fedavg_overlap_mpi_gpu.py:103-132``
  * ``run_baseline_gpu`` / ``run_overlap_gpu``  ``Module_3/part3_mpi_gpu_train.py:100-184, 306-412``
  * ``FedAvgRunner.run``                         round loop of ``part3_fedavg...py:186-231``

B200 design: on CUDA a TinyECG local step is ONE kernel (``ops.tiny_ecg``), a round's
``local_steps`` launches are captured once in a CUDA graph and replayed, the loss is accumulated on
the device and read back once per round, and aggregation is one in-kernel NVLink collective
(``parallel.fedavg``).  The reference's per-step ``cuda.synchronize()`` + ``loss.item()`` are gone.
On CPU (no GPU) everything runs through stock PyTorch — that path is the plumbing config of
``BASELINE.json`` and the numerical oracle of the GPU tests.
"""
from __future__ import annotations

import time
from dataclasses import asdict
from typing import Callable, Dict, Iterator, List, Optional, Tuple

import torch
import torch.nn.functional as F

from ..models.tiny_ecg import TinyECG
from ..ops import tiny_ecg as tk
from ..parallel.fedavg import FedAvgAggregator, broadcast_model, client_weight
from .stats import BenchStats, RoundStatsExt

G0_LABEL, G1_LABEL = "G0_baseline_GPU_CACHE", "G1_overlap_amp"


# ---------------------------------------------------------------------------------------------
# reference-shaped step functions (work on any nn.Module, CPU or CUDA)
# ---------------------------------------------------------------------------------------------
def train_step_G0(model, x, y, opt, device) -> float:
    model.train()
    opt.zero_grad(set_to_none=True)
    loss = F.cross_entropy(model(x), y)
    loss.backward()
    opt.step()
    if torch.device(device).type == "cuda":
        torch.cuda.synchronize()
    return float(loss.item())


def train_step_G1(model, x, y, opt, scaler, device, dtype: torch.dtype = torch.float16) -> float:
    model.train()
    opt.zero_grad(set_to_none=True)
    dev_type = torch.device(device).type
    with torch.autocast(device_type=dev_type, dtype=dtype if dev_type == "cuda" else torch.bfloat16):
        loss = F.cross_entropy(model(x), y)
    if scaler is not None:
        scaler.scale(loss).backward()
        scaler.step(opt)
        scaler.update()
    else:
        loss.backward()
        opt.step()
    if dev_type == "cuda":
        torch.cuda.synchronize()
    return float(loss.item())


# ---------------------------------------------------------------------------------------------
# index stream: the permutation-batch order of make_gpu_batch_iter, produced a round at a time
# ---------------------------------------------------------------------------------------------
class IndexStream:
    """Per-epoch ``randperm(N)``, ragged tail dropped, handed out ``steps*B`` indices at a time."""

    def __init__(self, n: int, batch: int, device, generator: Optional[torch.Generator] = None):
        if n < batch:
            raise ValueError(f"dataset has {n} windows < batch {batch}")
        self.n, self.batch, self.device, self.gen = n, batch, device, generator
        self.usable = (n // batch) * batch
        self.perm = None
        self.pos = 0

    def fill(self, out: torch.Tensor) -> None:
        need, done = out.numel(), 0
        while done < need:
            if self.perm is None or self.pos >= self.usable:
                self.perm = torch.randperm(self.n, device=self.device, generator=self.gen)
                self.pos = 0
            take = min(need - done, self.usable - self.pos)
            out[done:done + take].copy_(self.perm[self.pos:self.pos + take])
            done += take
            self.pos += take


# ---------------------------------------------------------------------------------------------
# trainers
# ---------------------------------------------------------------------------------------------
class TorchLocalTrainer:
    """Stock-PyTorch local training (CPU plumbing config, oracle, and non-TinyECG models)."""

    def __init__(self, model: torch.nn.Module, device, lr=1e-2, momentum=0.9, amp: str = "none"):
        self.model, self.device, self.amp = model, torch.device(device), amp
        self.opt = torch.optim.SGD(model.parameters(), lr=lr, momentum=momentum)
        self.scaler = None
        if amp == "fp16" and self.device.type == "cuda":
            self.scaler = torch.amp.GradScaler("cuda")
        self.launches = 0

    def flat_parameters(self) -> torch.Tensor:
        return torch.cat([p.detach().reshape(-1) for p in self.model.parameters()])

    def load_flat(self, flat: torch.Tensor) -> None:
        off = 0
        with torch.no_grad():
            for p in self.model.parameters():
                n = p.numel()
                p.copy_(flat[off:off + n].view_as(p))
                off += n

    def run_steps(self, batch_iter: Iterator, steps: int) -> Tuple[float, int]:
        loss_acc, n = 0.0, 0
        for _ in range(steps):
            x, y = next(batch_iter)
            n += x.size(0)
            if self.amp == "none":
                loss_acc += train_step_G0(self.model, x, y, self.opt, self.device)
            else:
                dt = torch.float16 if self.amp == "fp16" else torch.bfloat16
                loss_acc += train_step_G1(self.model, x, y, self.opt, self.scaler, self.device, dt)
        return loss_acc, n


from ..ops.ce_scaler import ScaledCrossEntropy as _ScaledCE  # noqa: E402


class KernelLocalTrainer:
    """Any model built from this package's ops (e.g. ResNet-1D wide) on CUDA: autograd drives the
    tcgen05 conv / BN kernels, the loss is the CE kernel, the update is ONE flat SGD-momentum kernel
    over the parameter arena (found-inf aware when the fp16-style scaler is on), and with
    ``sync_grads`` the gradient buckets are all-reduced in-kernel over NVLink while backward runs."""

    def __init__(self, model: torch.nn.Module, device, lr=1e-2, momentum=0.9, use_scaler: bool = False,
                 sync_grads: bool = False, bucket_bytes: int = 4 << 20, wire: str = "fp32", use_graph: bool = False):
        from ..models.tiny_ecg import FlatParams
        from ..ops import _lib
        self.model, self.device, self._lib = model, torch.device(device), _lib
        self.flat = FlatParams(model, align=4)
        self.lr, self.momentum, self.use_scaler = float(lr), float(momentum), bool(use_scaler)
        self.mom = torch.zeros_like(self.flat.flat)
        self.ctrl = torch.zeros(8, dtype=torch.int32, device=self.device)
        self.fstate = torch.zeros(8, dtype=torch.float32, device=self.device)
        self.fstate[tk.FS_SCALE] = 65536.0 if use_scaler else 1.0
        self.loss_sum = torch.zeros((), dtype=torch.float32, device=self.device)
        # bf16 shadow of the whole arena, refreshed by the SGD kernel: conv layers read their weights from it
        # (no per-layer cast kernels) and accumulate weight gradients straight into the flat gradient arena
        from ..ops.conv1d import Conv1dB200
        from ..ops.norm_act import BNReLU1d
        self.shadow = self.flat.flat.to(torch.bfloat16)
        for mod in model.modules():
            if isinstance(mod, Conv1dB200) and mod.cin > 1:
                i = next(j for j, q in enumerate(self.flat.params) if q is mod.weight)
                off = self.flat.offsets[i]
                mod.weight_bf16 = self.shadow[off:off + mod.weight.numel()].view_as(mod.weight)
                mod.fused_wgrad = True
            elif isinstance(mod, BNReLU1d):
                mod.fused_grad = True                       # dgamma / dbeta are added into the arena by the reduction kernel
        self.sync = None
        self.bucket_bytes, self.sync_grads = bucket_bytes, bool(sync_grads)
        if sync_grads:
            from ..parallel.grad_sync import GradSync
            self.sync = GradSync(self.flat, bucket_bytes=bucket_bytes, wire=wire)
        self.launches = 0
        # whole-step CUDA graph (forward + backward + bucket all-reduces + SGD): ~600 launches become one replay
        self.use_graph, self._graph, self._warm = bool(use_graph), None, 0
        self._fused = None            # set by enable_fused_aggregation: the round's last step ends in the bucketed FedAvg
        self._graph_f, self._warm_f = None, 0

    import os as _os
    fused_max_blocks = int(_os.environ.get("CS_BUCKET_BLOCKS", "8"))   # CTAs per bucket launch: they share the GPU with backward

    def enable_fused_aggregation(self, agg, prev: Optional[torch.Tensor]) -> bool:
        """Let the LAST local step of a round carry the aggregation: each gradient bucket, as soon as backward has
        produced it, gets ONE launch on a side stream that applies the optimizer step to the bucket, averages it
        across clients in place over NVLink and refreshes its bf16 shadow (``cs_fedavg_bucket``).  Needs the
        parameters in the aggregator's symmetric arena (in-place P2P mode), no loss scaler, no momentum reset and no
        per-step gradient all-reduce.  Returns whether the fused path is active."""
        if (self.use_scaler or self.sync_grads or not getattr(agg, "inplace", False) or agg.nvls is not None
                or agg.reset_momentum):
            return False
        if self.flat.flat.data_ptr() != agg.param_view.data_ptr():
            return False
        from ..parallel.grad_sync import GradSync
        if self.sync is None:
            self.sync = GradSync(self.flat, bucket_bytes=self.bucket_bytes, backend="off")
        self._fused = {"agg": agg, "prev": prev}
        return True

    def _fused_bucket_op(self, j, b, ch, side) -> None:
        f = self._fused
        agg = f["agg"]
        lo, n = b["lo_al"], b["hi_al"] - b["lo_al"]
        use_prev = agg.server_lr != 1.0 and f["prev"] is not None
        self._lib.check(self._lib.cuda_lib().cs_fedavg_bucket(
            agg.comm.handle, lo * 4, self.flat.grad.data_ptr() + lo * 4, self.mom.data_ptr() + lo * 4,
            (f["prev"].data_ptr() + lo * 4) if use_prev else None, self.shadow.data_ptr() + lo * 2, n, agg._weights,
            self.lr, self.momentum, 0.0, 1.0, agg.server_lr, 1, ch, self.fused_max_blocks, side.cuda_stream), "cs_fedavg_bucket")
        agg.launches += 1

    def flat_parameters(self) -> torch.Tensor:
        return self.flat.flat

    def load_flat(self, flat: torch.Tensor) -> None:
        if flat.data_ptr() != self.flat.flat.data_ptr():
            self.flat.flat.copy_(flat)
        self.refresh_shadow()

    def refresh_shadow(self) -> None:
        """after the parameters changed outside the optimizer (aggregation, checkpoint load, arena rebind)"""
        self.shadow.copy_(self.flat.flat)

    def step(self, x, y, fused_agg: bool = False) -> torch.Tensor:
        if fused_agg:
            return self._step_fused_outer(x, y)
        if not self.use_graph:
            return self._step(x, y)
        if self._graph is None:
            if self._warm < 3:                              # eager warm-up: kernel attributes, allocator pools
                self._warm += 1
                return self._step(x, y)
            self._sx, self._sy = x.clone(), y.clone()
            torch.cuda.synchronize()
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph):
                self._sloss = self._step(self._sx, self._sy)
        self._sx.copy_(x)
        self._sy.copy_(y)
        self._graph.replay()
        return self._sloss

    def _step_fused_outer(self, x, y) -> torch.Tensor:
        """the round's last step: backward feeds the bucketed optimizer-step + aggregation launches (own CUDA graph)"""
        assert self._fused is not None
        if not self.use_graph:
            return self._step(x, y, fused_agg=True)
        if self._graph_f is None:
            if self._warm_f < 2:
                self._warm_f += 1
                return self._step(x, y, fused_agg=True)
            self._fx, self._fy = x.clone(), y.clone()
            torch.cuda.synchronize()
            self._graph_f = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph_f):
                self._floss = self._step(self._fx, self._fy, fused_agg=True)
        self._fx.copy_(x)
        self._fy.copy_(y)
        self._graph_f.replay()
        return self._floss

    def _step(self, x, y, fused_agg: bool = False) -> torch.Tensor:
        self.model.train()
        self.flat.grad.zero_()
        if self.sync is not None:
            self.sync.fused_op = self._fused_bucket_op if fused_agg else None
        loss = _ScaledCE.apply(self.model(x), y, self.fstate[tk.FS_SCALE:tk.FS_SCALE + 1] if self.use_scaler else None)
        loss.backward()
        if self.sync is not None:
            self.sync.finish()
            self.sync.fused_op = None
        if fused_agg:                                       # optimizer step, aggregation and shadow refresh already happened per bucket
            self.loss_sum += loss.detach()
            return loss.detach()
        n = self.flat.flat.numel()
        self._lib.check(self._lib.cuda_lib().cs_sgd_momentum(
            self.flat.flat.data_ptr(), self.mom.data_ptr(), self.flat.grad.data_ptr(), n, self.lr, self.momentum, 0.0,
            int(self.use_scaler), self.ctrl.data_ptr(), self.fstate.data_ptr(), 2.0, 0.5, 2000, self.shadow.data_ptr(),
            self._lib.stream_ptr()), "cs_sgd_momentum")
        self.launches += 1
        self.loss_sum += loss.detach()
        return loss.detach()

    def run_steps(self, batch_iter: Iterator, steps: int, fuse_last: bool = False) -> Tuple[float, int]:
        n = 0
        self.loss_sum.zero_()
        for s in range(steps):
            x, y = next(batch_iter)
            n += x.size(0)
            self.step(x, y, fused_agg=(fuse_last and s == steps - 1 and self._fused is not None))
        return float(self.loss_sum.item()), n


class FusedTinyTrainer:
    """TinyECG local training on the fused kernels.

    G1 (bf16 / fp16): the persistent round kernel - ``local_steps`` steps (and, when the caller passes
    ``fused_agg``, the FedAvg aggregation) are ONE cooperative launch; batches come from the kernel's own
    per-epoch permutation, so no ``randperm`` / index copy is on the stream.  G0 (fp32) takes the same path with
    fp32 operands fed to the tensor core as TF32 (what cuDNN does for the reference's fp32 convolutions by default);
    ``CROSSSCALE_G0_TF32=0`` or ``use_round=False``: one launch per step (G0: the fp32 FFMA kernel), a round's launches
    captured once in a CUDA graph."""

    def __init__(self, model: TinyECG, x_gpu: torch.Tensor, y_gpu: torch.Tensor, batch: int,
                 local_steps: int, lr=1e-2, momentum=0.9, amp: str = "none", use_graph: bool = True,
                 generator: Optional[torch.Generator] = None, use_round: Optional[bool] = None,
                 check_labels: bool = True):
        import os
        assert x_gpu.is_cuda and x_gpu.dim() == 2
        self.model, self.batch, self.local_steps = model, batch, local_steps
        if check_labels:
            tk.validate_labels(y_gpu, model.num_classes)        # one sync at set-up; the kernels only flag, never fault
        self.flat = model.flatten_parameters()
        amp_code = tk.AMP_BY_NAME[amp]
        self.state = tk.TinyTrainState(self.flat.flat, model.num_classes, lr=lr, momentum=momentum,
                                       amp=amp_code, scaler=tk.ScalerConfig(enabled=(amp == "fp16")))
        self.x, self.y = x_gpu, y_gpu
        if x_gpu.shape[0] < batch:
            raise ValueError(f"dataset has {x_gpu.shape[0]} windows < batch {batch}")
        if use_round is None:
            use_round = os.environ.get("CROSSSCALE_TINY_ROUND", "1") != "0"
        self.use_round = bool(use_round) and self.state.round_supported(int(x_gpu.shape[1])) and tk.round_kernel_available()
        self.seed = int(generator.initial_seed() if generator is not None else torch.initial_seed()) & ((1 << 63) - 1)
        self.stream_idx = None if self.use_round else IndexStream(x_gpu.shape[0], batch, x_gpu.device, generator)
        self.idx = None if self.use_round else torch.zeros(local_steps * batch, dtype=torch.int64, device=x_gpu.device)
        self.use_graph, self.graph, self._warm = use_graph, None, False
        self.launches = 0

    def _resident_steps(self):
        b = self.batch
        for s in range(self.local_steps):
            self.state.step(self.x, self.y, self.idx[s * b:(s + 1) * b], batch=b)

    def run_round_resident(self, fused_agg: Optional[dict] = None) -> None:
        """``local_steps`` fused steps over the GPU-resident dataset (asynchronous).  ``fused_agg``: keyword
        arguments of :meth:`TinyTrainState.round` that make the launch end in the FedAvg aggregation."""
        if self.use_round:
            self.state.round(self.x, self.y, self.local_steps, self.batch, sample_mode=tk.SAMPLE_PERM, seed=self.seed,
                             **(fused_agg or {}))
            self.launches += 1
            return
        assert fused_agg is None
        self.stream_idx.fill(self.idx)
        if not self.use_graph or not self._warm:
            self._resident_steps()                 # first round runs eagerly (one-time kernel set-up)
            self._warm = True
        else:
            if self.graph is None:
                self.graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self.graph):
                    self._resident_steps()
            self.graph.replay()
        self.launches += self.local_steps

    def run_round_staged(self, xb: torch.Tensor, yb: torch.Tensor, fused_agg: Optional[dict] = None) -> None:
        """``local_steps`` fused steps over a staged device buffer ``[steps*B, L]`` (fp32 or bf16)."""
        b = self.batch
        if self.use_round:
            self.state.round(xb, yb, self.local_steps, b, sample_mode=tk.SAMPLE_ROWS, **(fused_agg or {}))
            self.launches += 1
            return
        assert fused_agg is None
        for s in range(self.local_steps):
            self.state.step(xb[s * b:(s + 1) * b], yb[s * b:(s + 1) * b], None, batch=b)
        self.launches += self.local_steps

    def flat_parameters(self) -> torch.Tensor:
        return self.flat.flat


# ---------------------------------------------------------------------------------------------
# pseudo-federated engines (no parameter exchange): BenchStats rows
# ---------------------------------------------------------------------------------------------
def _bench_row(config, world, rank, batch, steps, data_ms, h2d_ms, compute_ms) -> dict:
    step_ms = (data_ms + h2d_ms + compute_ms) / max(steps, 1)
    return asdict(BenchStats(config=config, world_size=world, rank=rank, batch_size=batch, steps=steps,
                             data_ms=data_ms / max(steps, 1), h2d_ms=h2d_ms / max(steps, 1),
                             compute_ms=compute_ms / max(steps, 1), step_ms=step_ms,
                             samples_per_s=batch / (step_ms / 1e3) if step_ms > 0 else 0.0))


def _run_pseudo(model, batch_iter, device, steps, rank, batch_size, world, amp, label, x_gpu=None, y_gpu=None):
    device = torch.device(device)
    fused = (device.type == "cuda" and isinstance(model, TinyECG) and tk.use_kernels()
             and x_gpu is not None and model.num_classes <= tk.MAX_CLASSES)
    if fused:
        # the steps run as replays of one CUDA graph of `chunk` fused steps (steps is rounded up to whole chunks)
        chunk = min(steps, 50)
        tr = FusedTinyTrainer(model, x_gpu, y_gpu, batch_size, chunk, amp=amp)
        tr.run_round_resident()                       # warm-up (eager steps on the graph path)
        tr.run_round_resident()                       # graph capture + first replay on the graph path
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        done = 0
        while done < steps:
            tr.run_round_resident()
            done += chunk
            if rank == 0:
                print(f"[{label}] step {done}/{steps}", flush=True)
        b.record()
        torch.cuda.synchronize()
        tr.state.check_errors()
        return _bench_row(label, world, rank, batch_size, done, 0.0, 0.0, a.elapsed_time(b))
    tr = TorchLocalTrainer(model, device, amp=amp)
    t0 = time.perf_counter()
    for s in range(steps):
        if rank == 0 and s % 10 == 0:
            print(f"[{label}] step {s}/{steps}", flush=True)
        tr.run_steps(batch_iter, 1)
    if device.type == "cuda":
        torch.cuda.synchronize()
    return _bench_row(label, world, rank, batch_size, steps, 0.0, 0.0, (time.perf_counter() - t0) * 1e3)


def run_baseline_gpu(model, batch_iter, device, steps, rank, batch_size, world_size=1, x_gpu=None, y_gpu=None):
    """G0: fp32 local training, data GPU-resident.  Returns a ``BenchStats`` dict."""
    return _run_pseudo(model, batch_iter, device, steps, rank, batch_size, world_size, "none", G0_LABEL, x_gpu, y_gpu)


def run_overlap_gpu(model, batch_iter, device, steps, rank, batch_size, world_size=1, x_gpu=None, y_gpu=None,
                    amp: str = "bf16"):
    """G1: AMP local training (bf16 by default; ``amp='fp16'`` adds the loss scaler)."""
    return _run_pseudo(model, batch_iter, device, steps, rank, batch_size, world_size, amp, G1_LABEL, x_gpu, y_gpu)


# ---------------------------------------------------------------------------------------------
# FedAvg round loop
# ---------------------------------------------------------------------------------------------
class FedAvgRunner:
    """One client's view of FedAvg training (SPMD: every rank runs one of these)."""

    def __init__(self, model: torch.nn.Module, device, x_gpu: torch.Tensor, y_gpu: torch.Tensor,
                 batch_size: int = 256, local_steps: int = 50, config: str = "G0", amp: str = "auto",
                 rank: int = 0, world: int = 1, lr: float = 1e-2, momentum: float = 0.9,
                 agg_backend: str = "auto", wire: str = "fp32", server_lr: float = 1.0,
                 reset_momentum: bool = False, use_graph: bool = True,
                 generator: Optional[torch.Generator] = None, model_name: str = "tiny_ecg",
                 sync_grads: bool = False, inplace: str = "auto", fuse_bucket_agg: bool = True,
                 bucket_bytes: int = 4 << 20):
        self.device = torch.device(device)
        self.model, self.rank, self.world = model, rank, world
        self.batch, self.local_steps, self.config = batch_size, local_steps, config
        if amp == "auto":
            amp = "none" if config == "G0" else "bf16"
        self.amp, self.model_name = amp, model_name
        self.fused = (self.device.type == "cuda" and isinstance(model, TinyECG) and tk.use_kernels()
                      and model.num_classes <= tk.MAX_CLASSES)
        if self.fused:
            self.trainer = FusedTinyTrainer(model, x_gpu, y_gpu, batch_size, local_steps, lr, momentum,
                                            amp, use_graph, generator)
            self.flat = self.trainer.flat.flat
            self.numel = self.trainer.flat.numel
            self.mom = self.trainer.state.mom
            self.batch_iter = None
        else:
            from ..data.shard_dataset import make_gpu_batch_iter
            self.kernel_generic = (self.device.type == "cuda" and tk.use_kernels() and not isinstance(model, TinyECG))
            if self.kernel_generic:
                self.trainer = KernelLocalTrainer(model, self.device, lr, momentum, use_scaler=(amp == "fp16"),
                                                  sync_grads=sync_grads, use_graph=use_graph, bucket_bytes=bucket_bytes)
                self.mom = self.trainer.mom
            else:
                self.trainer = TorchLocalTrainer(model, self.device, lr, momentum,
                                                 amp if self.device.type == "cuda" else "none")
                self.mom = None
            self.batch_iter = make_gpu_batch_iter(x_gpu, y_gpu, batch_size, generator)
            self.flat = self.trainer.flat_parameters()
            self.numel = self.flat.numel()
        self.weight = client_weight(x_gpu.shape[0], self.device)
        broadcast_model(self.flat)                       # round 0: common initial model
        if not self.fused:
            self.trainer.load_flat(self.flat)
        self.agg = FedAvgAggregator(self.numel, self.device, backend=agg_backend, wire=wire,
                                    server_lr=server_lr, reset_momentum=reset_momentum, inplace=inplace)
        if self.agg.inplace:                             # large model: parameters move into peer-mapped memory
            self.trainer.flat.rebind(self.agg.param_view)
            self.flat = self.trainer.flat.flat
            if hasattr(self.trainer, "refresh_shadow"):
                self.trainer.refresh_shadow()
            self.agg.set_weights(self.weight)
        self.prev = self.flat.clone()
        self.rows: List[dict] = []
        # large kernel-path models: the round's last step carries the bucketed aggregation (see KernelLocalTrainer)
        self.fuse_bucket_agg = bool(not self.fused and getattr(self, "kernel_generic", False) and self.agg.inplace
                                    and fuse_bucket_agg and self.trainer.enable_fused_aggregation(self.agg, self.prev))
        # TinyECG on the persistent round kernel: the aggregation rides in the tail of the same launch
        self.fuse_agg = (self.fused and self.trainer.use_round and self.agg.backend == "kernel" and not self.agg.inplace
                         and self.numel <= 4096)

    def _fused_agg_args(self) -> dict:
        a = self.agg
        return dict(comm=a.comm.handle, prev=self.prev, weight=self.weight, wire_bf16=bool(a.wire_bf16),
                    server_lr=a.server_lr, reset_momentum=a.reset_momentum, arena_off=0, channel=a.channel)

    def check_health(self) -> None:
        """Raise if a device-side time-out or error was flagged (call after a synchronisation point)."""
        comm = getattr(self.agg, "comm", None)
        if comm is not None and self.agg.backend == "kernel" and comm.status() != 0:
            raise RuntimeError("FedAvg aggregation: a peer did not reach the cross-GPU barrier in time; the affected "
                               "blocks kept their local parameters (replicas may differ) - aborting instead of training on")
        if self.fused:
            self.trainer.state.check_errors()

    def _events(self):
        return [torch.cuda.Event(enable_timing=True) for _ in range(3)]

    def run_round(self, round_idx: int, record: bool = True) -> dict:
        cuda = self.device.type == "cuda"
        n_samples = self.local_steps * self.batch
        if cuda:
            e0, e1, e2 = self._events()
            e0.record()
        t0 = time.perf_counter()
        if self.fused:
            self.trainer.state.reset_loss()
            # (timed rounds keep local training and aggregation as two launches so that both columns of the
            # reference's CSV - local_train_ms, comm_ms - are measured; run_round_async fuses them)
            self.trainer.run_round_resident()
        else:
            if self.fuse_bucket_agg:
                loss_acc, n_samples = self.trainer.run_steps(self.batch_iter, self.local_steps, fuse_last=True)
            else:
                loss_acc, n_samples = self.trainer.run_steps(self.batch_iter, self.local_steps)
            self.flat = self.trainer.flat_parameters()   # (a view of the arena for the kernel trainer)
        t1 = time.perf_counter()
        if cuda:
            e1.record()
        if not self.fuse_bucket_agg:                     # (fused: aggregated bucket by bucket under the last backward)
            self.agg.aggregate(self.flat, self.prev, self.weight, self.mom)
            if not self.fused:
                self.trainer.load_flat(self.flat)
        if cuda:
            e2.record()
            torch.cuda.synchronize()
            self.check_health()
            local_ms, comm_ms = e0.elapsed_time(e1), e1.elapsed_time(e2)
        else:
            t2 = time.perf_counter()
            local_ms, comm_ms = (t1 - t0) * 1e3, (t2 - t1) * 1e3
        avg_loss = (self.trainer.state.loss_sum() if self.fused else loss_acc) / self.local_steps
        row = asdict(RoundStatsExt(
            config=self.config, world_size=self.world, rank=self.rank, round_idx=round_idx,
            batch_size=self.batch, local_steps=self.local_steps, local_train_ms=local_ms, comm_ms=comm_ms,
            samples_per_s=n_samples / (local_ms / 1e3) if local_ms > 0 else 0.0, avg_loss=avg_loss,
            round_ms_dev=local_ms + comm_ms, agg_ms_dev=comm_ms, exposed_agg_ms=comm_ms,
            amp=self.amp, model=self.model_name))
        if record:
            self.rows.append(row)
        return row

    def run(self, rounds: int) -> List[dict]:
        for r in range(rounds):
            self.run_round(r)
        return self.rows

    # -- asynchronous round (device-timed benchmarks): nothing here synchronises the host -------
    def run_round_async(self, aggregate: bool = True) -> None:
        """One round, enqueued only.  On the persistent-kernel path local steps AND aggregation are one launch;
        ``aggregate=False`` runs the local steps alone (the no-communication round of the exposed-time measurement)."""
        assert self.fused, "async rounds need the fused CUDA trainer"
        if self.fuse_agg:
            self.trainer.run_round_resident(self._fused_agg_args() if aggregate else None)
            self.agg.launches += 0
            return
        self.trainer.run_round_resident()
        if aggregate:
            self.agg.aggregate(self.flat, self.prev, self.weight, self.mom)

    # -- end-to-end round: pinned-host block -> H2D on the copy stream -> train -> aggregate ->
    #    loss read back to the host (the call a user makes when data does not live on the GPU) ---
    def enable_e2e(self, store, seed: int = 0) -> None:
        from ..data.staging import BlockSampler, H2DStager
        from ..ops import _lib
        assert self.fused, "the staged end-to-end path needs the fused CUDA trainer"
        rows = self.local_steps * self.batch
        self._store, self._sampler = store, BlockSampler(store.n, rows, seed)
        self._stager = H2DStager(rows, store.l, store.x.dtype, self.device)
        # loss read-back ring: round r writes slot r % 4; the host reads every round's loss, `e2e_depth` rounds late
        self._loss_slots = 4
        self.e2e_depth = 2
        self._loss_host = torch.zeros(4 * self._loss_slots, dtype=torch.float32, pin_memory=True)
        self._loss_prev, self._last_loss, self._e2e_rounds, self._e2e_collected = 0.0, 0.0, 0, 0
        self._loss_events = [torch.cuda.Event() for _ in range(self._loss_slots)]
        # the D2H copy itself runs on a side stream from a per-round device snapshot of the running loss sum, so the
        # training stream never waits on the copy engine between rounds
        self._loss_ring = torch.zeros(self._loss_slots, dtype=torch.float32, device=self.device)
        self._loss_snap = [torch.cuda.Event() for _ in range(self._loss_slots)]
        self._rb_stream = torch.cuda.Stream(device=self.device)
        self._graphs = [None, None]
        self._slot = 0
        self._lib = _lib
        self.d2h_bytes = 4
        self._stage_next(0)

    def _stage_next(self, slot: int) -> None:
        hx, hy = self._store.block(self._sampler.next_start(), self.local_steps * self.batch)
        self._stager.stage(slot, hx, hy)

    def run_round_e2e(self, pipelined: bool = False) -> float:
        """One FL round fed from pinned host memory; returns a round's mean loss (host float).

        ``pipelined=False``: the host waits for THIS round's loss before returning.
        ``pipelined=True``: this round is enqueued (H2D of the next block, 50 fused steps, aggregation, D2H of the
        loss) and the host then collects the loss of the round ``e2e_depth`` (2) rounds back, so a host hiccup
        shorter than a round or two never idles the GPU; every round's loss is still read back by the host, in
        order — call :meth:`drain_e2e` for the tail."""
        trace = getattr(self, "e2e_trace", None)        # optional diagnostics: list of per-round dicts (bench --trace-e2e)
        if trace is not None:
            h0 = time.perf_counter()
        slot = self._slot
        xb, yb = self._stager.acquire(slot)
        self._stage_next(slot ^ 1)                  # next block's H2D overlaps this round's compute
        tr = self.trainer
        if trace is not None:
            ev0, ev1, ev2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            ev0.record()
        fused_here = self.fuse_agg
        if tr.use_round:
            tr.run_round_staged(xb, yb, self._fused_agg_args() if fused_here else None)
        elif not tr.use_graph:
            tr.run_round_staged(xb, yb)
        else:
            if self._graphs[slot] is None:
                tr.run_round_staged(xb, yb)         # eager once (also the first round's work)
                g = torch.cuda.CUDAGraph()
                self._graphs[slot] = (g, xb, yb)
            else:
                g = self._graphs[slot][0]
                if not getattr(g, "_cs_captured", False):
                    with torch.cuda.graph(g):
                        tr.run_round_staged(xb, yb)
                    g._cs_captured = True
                g.replay()
                tr.launches += self.local_steps
        if trace is not None:
            ev1.record()
        if not fused_here:
            self.agg.aggregate(self.flat, self.prev, self.weight, self.mom)
        # the slot is handed back AFTER the aggregation: the next block's H2D copy waits on this event, and a 13 MB DMA
        # running under the aggregation kernel slows its system-scope flag traffic (traced: 0.05 ms per round at two
        # GPUs, 0.16-0.21 ms at eight, against 0.012-0.018 ms when nothing else touches the memory system)
        self._stager.release(slot)
        st = torch.cuda.current_stream(self.device)
        if trace is not None:
            ev2.record()
            h1 = time.perf_counter()
        k = self._e2e_rounds % self._loss_slots
        self._loss_ring[k:k + 1].copy_(tr.state.fstate[tk.FS_LOSS_SUM:tk.FS_LOSS_SUM + 1])      # snapshot, in stream order
        self._loss_snap[k].record(st)
        self._rb_stream.wait_event(self._loss_snap[k])
        self._lib.check(self._lib.cuda_lib().cs_stage_d2h(self._loss_host.data_ptr() + 16 * k, self._loss_ring.data_ptr() + 4 * k,
                                                          4, self._rb_stream.cuda_stream), "cs_stage_d2h")
        self._loss_events[k].record(self._rb_stream)
        self._e2e_rounds += 1
        self._slot ^= 1
        depth = min(self.e2e_depth, self._loss_slots - 1) if pipelined else 0
        out = self._collect_losses(self._e2e_rounds - depth)
        if trace is not None:
            trace.append({"ev": (ev0, ev1, ev2), "host_enqueue_ms": (h1 - h0) * 1e3, "host_total_ms": (time.perf_counter() - h0) * 1e3})
        return out

    def _collect_losses(self, upto: int) -> float:
        """host-side read of the per-round losses, in order, for rounds < ``upto``"""
        while self._e2e_collected < upto:
            k = self._e2e_collected % self._loss_slots
            self._loss_events[k].synchronize()
            total = float(self._loss_host[4 * k])
            self._last_loss = (total - self._loss_prev) / self.local_steps
            self._loss_prev = total
            self._e2e_collected += 1
        return self._last_loss

    def drain_e2e(self) -> float:
        """Read the losses of the rounds still in flight (pipelined mode) and surface device-side errors."""
        out = self._collect_losses(self._e2e_rounds)
        self.check_health()
        return out

    def close(self):
        if getattr(self.agg, "inplace", False):          # parameters live in the arena: move them out before it is freed
            torch.cuda.synchronize()
            self.trainer.flat.rebind(self.flat.clone())
            self.flat = self.trainer.flat.flat
        self.agg.close()
