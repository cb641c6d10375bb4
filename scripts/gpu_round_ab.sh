#!/bin/bash
# A/B of search-kernel variants (staged rods per pass, resident blocks per SM) on C4 (every step searches) and on
# C3's rebuild step; then the C4 line with the relaxation phase and an ncu capture of the search kernel.
mkdir -p gpurun_out
cp cr_mech_coli_b200/libcrm_b200.so /tmp/lib_default.so
for v in default l8b4 l8b5 l8b6; do
  if [ $v != default ]; then cp cr_mech_coli_b200/libcrm_b200_$v.so cr_mech_coli_b200/libcrm_b200.so; fi
  timeout 300 python bench.py --config c4 --no-cpu --no-long-run --e2e-reps 1 > gpurun_out/ab_c4_$v.json 2> gpurun_out/ab_c4_$v.err; echo "c4 $v rc=$?" | tee -a gpurun_out/rc_ab.txt
  timeout 300 python bench.py --no-cpu --e2e-reps 1 > gpurun_out/ab_c3_$v.json 2> gpurun_out/ab_c3_$v.err; echo "c3 $v rc=$?" | tee -a gpurun_out/rc_ab.txt
  python - <<PY
import json
for c in ("c4","c3"):
    try:
        d=json.load(open("gpurun_out/ab_%s_$v.json"%c))
        print("$v", c, "ms/step %.4f"%d["ms_per_step"], "force %.4f"%d["step_breakdown_ms"]["force"], (d.get("long_run") or {}).get("rebuild_step_ms"), (d.get("long_run") or {}).get("ms_per_step"))
    except Exception as e: print("$v", c, "failed", e)
PY
done
cp /tmp/lib_default.so cr_mech_coli_b200/libcrm_b200.so
timeout 600 python bench.py --config c4 --no-cpu > gpurun_out/bench_c4_n1.json 2> gpurun_out/bench_c4_n1.err; echo "c4 relaxed rc=$?" | tee -a gpurun_out/rc_ab.txt
python -c "
import json; d=json.load(open('gpurun_out/bench_c4_n1.json')); print('c4', d['ms_per_step'], json.dumps(d.get('relaxed')))"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_force<' -s 4 -c 1 -f -o gpurun_out/search_c4 \
   python bench.py --config c4 --steps 10 --warmup 3 --no-cpu --no-long-run --e2e-reps 1 > gpurun_out/ncu_search.log 2>&1; echo "ncu_search rc=$?" | tee -a gpurun_out/rc_ab.txt
timeout 300 python -m pytest tests/test_gpu_parity.py -q -m gpu -p no:cacheprovider -k "batched or staging or fused or sharded or patch" > gpurun_out/pytest_subset.log 2>&1; echo "pytest subset rc=$?" | tee -a gpurun_out/rc_ab.txt; tail -3 gpurun_out/pytest_subset.log
