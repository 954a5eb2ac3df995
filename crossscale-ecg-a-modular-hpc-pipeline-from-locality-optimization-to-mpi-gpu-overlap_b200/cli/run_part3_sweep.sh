#!/usr/bin/env bash
# FedAvg scaling sweep on ONE multi-GPU node (replaces the SLURM/srun loop of
# TRUE_FL_M3/run_part3_sweep.sh:20-49): world sizes 1 2 4 8, REPEATS runs each, both configs,
# batch 256, 5 rounds x 50 local steps, max 20000 windows per client.
set -euo pipefail
DATA_ROOT="${1:?usage: run_part3_sweep.sh <DATA_ROOT with ecg_*.bin> [RESULTS_DIR]}"
RESULTS_DIR="${2:-results}"
WORLD_SIZES=(${WORLD_SIZES:-1 2 4 8})
REPEATS="${REPEATS:-5}"
PORT="${MASTER_PORT:-29611}"
for W in "${WORLD_SIZES[@]}"; do
  for R in $(seq 1 "${REPEATS}"); do
    echo "=== world=${W} repeat=${R}"
    python -m torch.distributed.run --nnodes=1 --nproc-per-node "${W}" --master-addr 127.0.0.1 --master-port "${PORT}" \
      -m crossscale.cli.part3_fedavg_overlap_mpi_gpu --data-root "${DATA_ROOT}" --batch-size 256 --rounds 5 \
      --local-steps 50 --config both --max-windows 20000 --results-dir "${RESULTS_DIR}"
  done
done
python -m crossscale.cli.plots part3_fl --results-dir "${RESULTS_DIR}"
