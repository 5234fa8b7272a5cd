#!/bin/bash
# full -m gpu suite + the contract bench line (with its parity gate) on one GPU
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?"
tail -5 gpurun_out/bench1.err; cat gpurun_out/bench1.json
bash tools/run_probe.sh > /dev/null 2>&1; tail -8 gpurun_out/probe.log
python tools/ref_cuda_compare.py > gpurun_out/ref_cuda_compare.json 2> gpurun_out/ref_cuda_compare.err; echo "refcuda rc=$?"; tail -3 gpurun_out/ref_cuda_compare.err
