#!/usr/bin/env bash
# Multi-process launcher: one rank per GPU on this host (replaces cifar_example/train.sh, which
# prompts for node count / batch size / iterations and calls `mpirun -n N -npernode 1 ...`).
#   examples/launch.sh <nproc> <script.py> [script args...]
set -euo pipefail
N=${1:?usage: launch.sh <nproc> <script.py> [args...]}
shift
PORT=${MASTER_PORT:-29500}
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" \
     --master-addr 127.0.0.1 --master-port "$PORT" "$@"
