#!/bin/bash
# 2-GPU check: the real ring (every transport / forward schedule) against the oracle, then the contract bench at N=2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/ring_worker.py > gpurun_out/ring_worker_$N.log 2>&1; echo "ring_worker rc=$?"
grep -v "^W\|^\[W\|Warning" gpurun_out/ring_worker_$N.log | tail -30
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_$N.json 2> gpurun_out/bench_$N.err; echo "bench rc=$?"
tail -5 gpurun_out/bench_$N.err; cat gpurun_out/bench_$N.json
