"""CrossScale-B200: Blackwell-native federated ECG training framework.

Layer map (mirrors SURVEY.md section 1, re-designed for sm_100a):

  data/      shard format, shard_prep, shard_dataset, synthetic generators, pinned-ring prefetch
  models/    TinyECG (state_dict compatible with the reference), widened ResNet-1D
  ops/       ctypes bindings to the hand-written sm_100a kernels in csrc/cuda (+ CPU kernel)
  parallel/  symmetric peer memory, in-kernel FedAvg aggregation, bucketed gradient all-reduce
  fl/        step engines (G0/G1), FedAvg round loop, CUDA-graph capture, stats rows
  utils/     timing, CSV append, clocks sampling, seeds, checkpoints
  cli/       flag-compatible entry points (shard_prep, part3_mpi_gpu_train, ...)

Import it as ``crossscale`` (see ``crossscale/__init__.py`` at the repo root).
"""
__version__ = "0.1.0"
