set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29601 tests/_dist_gpu_checks.py > gpurun_out/dist8.log 2>&1; tail -2 gpurun_out/dist8.log
timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29602 examples/allreduce_sweep.py --max-mb 1024 > gpurun_out/sweep8.log 2>&1; tail -11 gpurun_out/sweep8.log | cut -c1-420
timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29603 examples/allreduce_sweep.py --max-mb 256 --dtype bf16 > gpurun_out/sweep8_bf16.log 2>&1; tail -3 gpurun_out/sweep8_bf16.log | cut -c1-420
python bench.py --steps 200 --warmup 10 2>/dev/null | tail -1 > gpurun_out/bench_n1.json; cut -c1-200 gpurun_out/bench_n1.json
for n in 2 4 8; do timeout -k 10 200 $TR --nproc-per-node $n --master-port 2961$n bench.py --gpus $n --steps 200 --warmup 10 2>/dev/null | tail -1 > gpurun_out/bench_n$n.json; cut -c1-200 gpurun_out/bench_n$n.json; done
TADNN_DP_OVERLAP=0 timeout -k 10 200 $TR --nproc-per-node 8 --master-port 29619 bench.py --gpus 8 --steps 200 --warmup 10 --no-e2e 2>/dev/null | tail -1 > gpurun_out/bench_n8_nooverlap.json; cut -c1-200 gpurun_out/bench_n8_nooverlap.json
timeout -k 10 200 $TR --nproc-per-node 8 --master-port 29620 bench.py --gpus 8 --steps 100 --warmup 10 --impl baseline 2>/dev/null | tail -1 > gpurun_out/bench_n8_baseline.json; cut -c1-200 gpurun_out/bench_n8_baseline.json
for cfg in wide readme; do for impl in tadnn nccl; do timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29621 examples/mp_mlp_bench.py --config $cfg --impl $impl --phases 2>&1 | tail -1 | cut -c1-900; done; done
