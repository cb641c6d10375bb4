#!/bin/bash
# chunked staging of the list kernel: parity subset, then A/B of 8 vs 7 resident blocks per SM on C3 (short + long run)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -q -m gpu -p no:cacheprovider -x -k "more_than_16 or fused or persistent or batched or mixed or patch or sharded_persistent or multilayer" > gpurun_out/pytest_subset.log 2>&1; echo "pytest subset rc=$?" | tee -a gpurun_out/rc_ab.txt; tail -4 gpurun_out/pytest_subset.log
cp cr_mech_coli_b200/libcrm_b200.so /tmp/lib_default.so
for v in mb8 mb7; do
  if [ $v != mb8 ]; then cp cr_mech_coli_b200/libcrm_b200_$v.so cr_mech_coli_b200/libcrm_b200.so; fi
  timeout 300 python bench.py --steps 40 --warmup 5 --no-cpu --e2e-reps 1 > gpurun_out/ab_c3_$v.json 2> gpurun_out/ab_c3_$v.err; echo "c3 $v rc=$?" | tee -a gpurun_out/rc_ab.txt
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/ab_c3_$v.json")); L=d["long_run"]
    print("$v", "ms/step %.4f"%d["ms_per_step"], "long %.4f"%L["ms_per_step"], "late %.4f"%L["list_steps_at_end"]["ms_per_step"], "longest", L["list_steps_at_end"]["longest_list"], "rebuild", L["rebuild_step_ms"])
except Exception as e: print("$v failed", e)
PY
done
cp /tmp/lib_default.so cr_mech_coli_b200/libcrm_b200.so
