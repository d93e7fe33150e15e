#!/usr/bin/env bash
# First GPU call after a stretch of CPU-only work (one B200 is enough):
#   gpurun --timeout 900 -- 'bash tools/pending_gpu_checks.sh'
# 1. tests of code that has never run on a GPU (marker `wip`): move them to `gpu` once they pass
# 2. the whole GPU suite (host-side changes in csrc/comm*.cu, common.cuh and preprocess.cu since the last run)
# 3. a short flagship bench as a smoke test of the unchanged hot path
# Everything is logged under gpurun_out/ so that a truncated tail does not hide a failure.
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m wip -q > gpurun_out/pending_wip.log 2>&1; echo "wip rc=$?" | tee -a gpurun_out/pending_wip.log
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pending_gpu.log 2>&1; echo "gpu rc=$?" | tee -a gpurun_out/pending_gpu.log
timeout 200 python bench.py --steps 50 --warmup 5 > gpurun_out/pending_bench.json 2> gpurun_out/pending_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/pending_wip.log; tail -3 gpurun_out/pending_gpu.log; cut -c1-300 gpurun_out/pending_bench.json
