#!/bin/bash
# One GPU-box visit: full -m gpu tests, the no-alloc test alone, bench lines of all configs, mask-path bench, ncu.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_all rc=$?" | tee -a gpurun_out/rc.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "pytest_noalloc rc=$?" | tee -a gpurun_out/rc.txt
CRM_B200_TRACE=1 timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_c3_n1.json 2> gpurun_out/bench_c3_n1.err; echo "bench rc=$?" | tee -a gpurun_out/rc.txt
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_c3_reference.json 2> gpurun_out/bench_c3_reference.err; echo "bench reference rc=$?" | tee -a gpurun_out/rc.txt
for c in c2 c4 c5; do
  timeout 600 python bench.py --config $c > gpurun_out/bench_${c}_n1.json 2> gpurun_out/bench_${c}_n1.err; echo "bench $c rc=$?" | tee -a gpurun_out/rc.txt
done
timeout 600 python scripts/bench_raster.py > gpurun_out/raster.jsonl 2> gpurun_out/raster.err; echo "raster rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 20 --warmup 5 --no-cpu --no-long-run --e2e-reps 1 > gpurun_out/ncu_launch.log 2>&1; echo "ncu_list rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'^k_force_eval$' -s 6 -c 1 -f -o gpurun_out/fused_eval \
   python bench.py --steps 20 --warmup 5 --no-cpu --no-long-run --e2e-reps 1 > gpurun_out/ncu_full.log 2>&1; echo "ncu_full rc=$?" | tee -a gpurun_out/rc.txt
tail -12 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_noalloc_alone.log; python scripts/bench_short.py < gpurun_out/bench_c3_n1.json; grep "crm trace" gpurun_out/bench_c3_n1.err | tail -2
for c in c2 c4 c5; do python scripts/bench_short.py < gpurun_out/bench_${c}_n1.json | head -1; done; cat gpurun_out/raster.jsonl | cut -c1-400; tail -3 gpurun_out/raster.err
