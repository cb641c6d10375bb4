#!/bin/bash
# One GPU-box visit: full -m gpu tests, the no-alloc test alone, bench lines, A/B of the list kernel's occupancy, ncu.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest_all rc=$?" | tee -a gpurun_out/rc.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "pytest_noalloc rc=$?" | tee -a gpurun_out/rc.txt
CRM_B200_TRACE=1 timeout 900 python bench.py > gpurun_out/bench_c3_n1.json 2> gpurun_out/bench_c3_n1.err; echo "bench rc=$?" | tee -a gpurun_out/rc.txt
for c in c2 c4; do
  timeout 600 python bench.py --config $c --no-cpu > gpurun_out/bench_${c}_n1.json 2> gpurun_out/bench_${c}_n1.err; echo "bench $c rc=$?" | tee -a gpurun_out/rc.txt
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_force_eval -s 6 -c 1 -f -o gpurun_out/fused_eval \
   python bench.py --steps 20 --warmup 3 --no-cpu --no-long-run --e2e-reps 1 > gpurun_out/ncu_full.log 2>&1; echo "ncu_full rc=$?" | tee -a gpurun_out/rc.txt
# A/B: the same library with the list kernel compiled for 6 resident blocks per SM
if [ -f cr_mech_coli_b200/libcrm_b200_mb6.so ]; then
  cp cr_mech_coli_b200/libcrm_b200.so /tmp/lib_default.so
  cp cr_mech_coli_b200/libcrm_b200_mb6.so cr_mech_coli_b200/libcrm_b200.so
  timeout 600 python bench.py --steps 40 --warmup 5 --no-cpu --no-long-run --e2e-reps 1 > gpurun_out/bench_c3_mb6.json 2> gpurun_out/bench_c3_mb6.err; echo "bench mb6 rc=$?" | tee -a gpurun_out/rc.txt
  cp /tmp/lib_default.so cr_mech_coli_b200/libcrm_b200.so
  timeout 600 python bench.py --steps 40 --warmup 5 --no-cpu --no-long-run --e2e-reps 1 > gpurun_out/bench_c3_mb8.json 2> gpurun_out/bench_c3_mb8.err; echo "bench mb8 rc=$?" | tee -a gpurun_out/rc.txt
fi
tail -12 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_noalloc_alone.log; python scripts/bench_short.py < gpurun_out/bench_c3_n1.json; grep "crm trace" gpurun_out/bench_c3_n1.err | tail -4
echo mb6; python scripts/bench_short.py < gpurun_out/bench_c3_mb6.json; echo mb8; python scripts/bench_short.py < gpurun_out/bench_c3_mb8.json
