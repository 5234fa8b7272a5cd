#!/bin/bash
# final 1-GPU evidence of round 2 on the final library: full -m gpu suite, smoke(), default contract line, ncu launch list
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r2h_pytest_gpu_final.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2h_pytest_gpu_final.log; tail -n 3 gpurun_out/r2h_pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2h_smoke_final.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2h_smoke_final.log; tail -n 2 gpurun_out/r2h_smoke_final.log
python bench.py > gpurun_out/r2h_bench_1gpu_final.json 2> gpurun_out/r2h_bench_1gpu_final.err; echo "bench rc=$?"
python bench.py --impl reference > gpurun_out/r2h_bench_reference_arm.json 2>> gpurun_out/r2h_bench_1gpu_final.err; echo "reference arm rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r2h_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-parity > gpurun_out/r2h_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r2h_bench_1gpu_final.json").read().strip().splitlines()[-1])
print("value", d["value"], "ms", d["ms_per_step"], "fwd", d["fwd"]["ms"], "bwd", d["bwd"]["ms"], "parity", d["parity"]["ok"], "e2e", d["e2e"]["value"], d["clocks"])
r = json.loads(open("gpurun_out/r2h_bench_reference_arm.json").read().strip().splitlines()[-1])
print("reference arm", r["value"], r["unit"], r["cpu_baseline"]["cores"])
PY
grep -E "rowdot|pad_stats|cast_dq|bwd_kernel|fwd_dbuf" gpurun_out/r2h_launches_bench.csv | tail -n 12 | cut -c1-220
