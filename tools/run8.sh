set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29601 tests/_dist_gpu_checks.py > gpurun_out/dist8.log 2>&1; tail -2 gpurun_out/dist8.log
timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29602 examples/allreduce_sweep.py --max-mb 1024 > gpurun_out/sweep8.log 2>&1; tail -12 gpurun_out/sweep8.log | cut -c1-330
for n in 8 4; do timeout -k 10 200 $TR --nproc-per-node $n --master-port 2961$n bench.py --gpus $n --steps 100 --warmup 10 2>/dev/null | tail -1 > gpurun_out/bench_n$n.json; cut -c1-200 gpurun_out/bench_n$n.json; done
timeout -k 10 200 $TR --nproc-per-node 8 --master-port 29620 bench.py --gpus 8 --steps 100 --warmup 10 --impl baseline 2>/dev/null | tail -1 | cut -c1-200
timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29621 examples/mp_mlp_bench.py --config wide --impl tadnn 2>&1 | tail -2
timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29622 examples/mp_mlp_bench.py --config wide --impl nccl 2>&1 | tail -2
timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29623 examples/mp_mlp_bench.py --config readme --impl tadnn 2>&1 | tail -2
timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29624 examples/mp_mlp_bench.py --config readme --impl nccl 2>&1 | tail -2
nvidia-smi topo -m > gpurun_out/topo8.txt 2>&1
