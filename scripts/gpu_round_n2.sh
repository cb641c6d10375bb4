#!/bin/bash
# Two GPUs: sharded tests (incl. the two-process NCCL ones) and the N = 2 bench lines.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_n2.txt
timeout 900 python -m pytest tests/test_gpu_sharded.py -q -m gpu -p no:cacheprovider --durations=5 > gpurun_out/pytest_sharded_n2.log 2>&1; echo "pytest_sharded rc=$?" | tee -a gpurun_out/rc_n2.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_c3_n2.json 2> gpurun_out/bench_c3_n2.err; echo "bench n2 rc=$?" | tee -a gpurun_out/rc_n2.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --config c4 --no-cpu --relax 0 > gpurun_out/bench_c4_n2.json 2> gpurun_out/bench_c4_n2.err; echo "bench c4 n2 rc=$?" | tee -a gpurun_out/rc_n2.txt
tail -8 gpurun_out/pytest_sharded_n2.log; python scripts/bench_short.py < gpurun_out/bench_c3_n2.json; tail -3 gpurun_out/bench_c3_n2.err; python scripts/bench_short.py < gpurun_out/bench_c4_n2.json; tail -3 gpurun_out/bench_c4_n2.err
