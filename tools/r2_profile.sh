#!/bin/bash
# round-2 evidence run on one GPU: C2 bench line, launch list of the contract bench, ncu --set full of the dominant kernels
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench.py --workload c2 --steps 20 --warmup 5 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "c2 rc=$?"; tail -3 gpurun_out/bench_c2.err; cut -c1-1500 gpurun_out/bench_c2.json
python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_prof_plain.json 2>&1; echo "plain rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r2_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu --no-parity > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python tools/ncu_target.py > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'fwd_dbuf_kernel|bwd_kernel' -s 4 -c 2 \
    -o gpurun_out/r2_full_c3 -f python tools/ncu_target.py > gpurun_out/ncu_full_c3.log 2>&1; echo "ncu full c3 rc=$?"
python tools/ncu_target_c2.py > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'bwd_tf32_kernel|fwd_kernel' -s 4 -c 2 \
    -o gpurun_out/r2_full_c2 -f python tools/ncu_target_c2.py > gpurun_out/ncu_full_c2.log 2>&1; echo "ncu full c2 rc=$?"
ls -la gpurun_out/*.ncu-rep
