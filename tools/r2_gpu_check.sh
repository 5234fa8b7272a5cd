#!/bin/bash
# full -m gpu suite + the contract bench line (with its parity gate) on one GPU
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
if [ -n "$1" ]; then python -m pytest $1 -x -q -m gpu 2>&1 | tail -40; fi
python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
python tools/ref_cuda_compare.py > gpurun_out/ref_cuda_compare.json 2> gpurun_out/ref_cuda_compare.err; echo "refcuda rc=$?"; tail -3 gpurun_out/ref_cuda_compare.err
grep -n "_ms\|speedup\|family" gpurun_out/ref_cuda_compare.json
python bench.py --steps 10 --warmup 3 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?"
tail -5 gpurun_out/bench1.err; cut -c1-900 gpurun_out/bench1.json
