set -x
python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/r1_pytest_gpu_final.log
python bench.py > gpurun_out/r1_bench_n1_final.json 2> gpurun_out/r1_bench_n1_final.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1_reference_arm_final.json 2>/dev/null
python bench.py --ctrl position --envs-per-gpu 4096 --steps 100 --warmup 10 > gpurun_out/r1_bench_config2_final.json 2>/dev/null
python bench.py --biped --steps 100 --warmup 10 > gpurun_out/r1_bench_biped_final.json 2>/dev/null
python bench.py --steps 3 --warmup 3 > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1_launches_final.csv python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_final.log 2>&1
tail -3 gpurun_out/r1_pytest_gpu_final.log
python - <<'PY'
import json
for f in ("r1_bench_n1_final","r1_bench_config2_final","r1_bench_biped_final","r1_reference_arm_final"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json")); print(f, d.get("value"), d.get("ms_per_step"), (d.get("e2e") or {}).get("value"), d.get("clocks"))
    except Exception as e: print(f, "ERR", e)
PY
