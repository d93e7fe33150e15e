set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29601 tests/_dist_gpu_checks.py > gpurun_out/dist8.log 2>&1; tail -1 gpurun_out/dist8.log
python bench.py 2>/dev/null | tail -1 > gpurun_out/bench_n1.json; cut -c1-180 gpurun_out/bench_n1.json
for n in 2 4 8; do timeout -k 10 200 $TR --nproc-per-node $n --master-port 2961$n bench.py --gpus $n 2>/dev/null | tail -1 > gpurun_out/bench_n$n.json; cut -c1-180 gpurun_out/bench_n$n.json; done
timeout -k 10 200 $TR --nproc-per-node 8 --master-port 29619 bench.py --gpus 8 --conv2 pooled 2>/dev/null | tail -1 > gpurun_out/bench_n8_pooled.json; cut -c1-180 gpurun_out/bench_n8_pooled.json
for cfg in wide readme; do timeout -k 10 300 $TR --nproc-per-node 8 --master-port 29621 examples/mp_mlp_bench.py --config $cfg --impl tadnn --phases 2>&1 | tail -1 | cut -c1-700; done
timeout -k 10 200 $TR --nproc-per-node 8 --master-port 29625 examples/cifar_dp.py -usegpu -data --miniBatch 1000 -iterations 3 -learningRate 0.05 --results-dir gpurun_out 2>&1 | tail -6
