# Round-2 evidence run (one gpurun call): GPU test log, bench lines, launch list, one full ncu capture of k_step at steady state.
set -x
python -m pytest tests -m gpu -q -s 2>&1 | grep -vE "^\s*$" | tail -120 > gpurun_out/r2_pytest_gpu_final.log
tail -3 gpurun_out/r2_pytest_gpu_final.log
python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference_arm.json 2>/dev/null
python bench.py --ctrl position --envs-per-gpu 4096 --steps 100 --warmup 10 > gpurun_out/r2_bench_config2_position_4096.json 2>/dev/null
python bench.py --biped --steps 100 --warmup 10 > gpurun_out/r2_bench_biped.json 2>/dev/null
python bench.py --actions policy --steps 100 --warmup 10 > gpurun_out/r2_bench_policy_actions.json 2>/dev/null
python bench.py --steps 3 --warmup 3 > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 3 --warmup 3 > gpurun_out/r2_ncu_launches.log 2>&1
python tools/prof_step.py 16384 300 20 torque stagger > gpurun_out/r2_prof_step_steady.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 402 -c 1 -f -o gpurun_out/r2_final python tools/prof_step.py 16384 300 4 torque stagger > gpurun_out/r2_ncu_full.log 2>&1
cut -c1-200 gpurun_out/r2_prof_step_steady.log; tail -2 gpurun_out/r2_ncu_full.log
python - <<'PY'
import json
for f in ("r2_bench_n1","r2_bench_config2_position_4096","r2_bench_biped","r2_bench_policy_actions","r2_bench_reference_arm"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json")); print(f, d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("value"), d.get("clocks"), (d.get("steady_state") or {}))
    except Exception as e: print(f, "ERR", e)
PY
