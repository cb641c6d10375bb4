#!/bin/bash
# Final visit: full -m gpu tests, smoke, the C3 line (with reference arm), launch list and ncu of the fused kernel.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu -p no:cacheprovider --durations=5 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_all rc=$?" | tee -a gpurun_out/rc.txt
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_c3_n1.json 2> gpurun_out/bench_c3_n1.err; echo "bench rc=$?" | tee -a gpurun_out/rc.txt
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 20 --warmup 5 --no-cpu --no-long-run --e2e-reps 1 > gpurun_out/ncu_launch.log 2>&1; echo "ncu_list rc=$?" | tee -a gpurun_out/rc.txt
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'^k_force_eval$' -s 6 -c 1 -f -o gpurun_out/fused_eval \
   python bench.py --steps 20 --warmup 5 --no-cpu --no-long-run --e2e-reps 1 > gpurun_out/ncu_full.log 2>&1; echo "ncu_full rc=$?" | tee -a gpurun_out/rc.txt
tail -9 gpurun_out/pytest_gpu.log; tail -1 gpurun_out/smoke.log; python scripts/bench_short.py < gpurun_out/bench_c3_n1.json 2>/dev/null | head -2
