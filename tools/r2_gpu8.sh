#!/bin/bash
# 8-GPU evidence run: ring_worker (all transports / schedules vs the oracle) at 8 and 4 ranks, contract bench at 8 and 4,
# C5 and the causal extension at 8
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { # nranks port script args...
  local n=$1 port=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"
}
run 8 29521 tests/ring_worker.py > gpurun_out/ring_worker_8.log 2>&1; echo "ring_worker 8 rc=$?"; grep -c "RING_OK" gpurun_out/ring_worker_8.log
run 8 29522 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_8.json 2> gpurun_out/bench_8.err; echo "bench 8 rc=$?"; tail -3 gpurun_out/bench_8.err; cut -c1-700 gpurun_out/bench_8.json
CUDA_VISIBLE_DEVICES=0,1,2,3 run 4 29523 tests/ring_worker.py > gpurun_out/ring_worker_4.log 2>&1; echo "ring_worker 4 rc=$?"; grep -c "RING_OK" gpurun_out/ring_worker_4.log
CUDA_VISIBLE_DEVICES=0,1,2,3 run 4 29524 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/bench_4.json 2> gpurun_out/bench_4.err; echo "bench 4 rc=$?"; cut -c1-400 gpurun_out/bench_4.json
run 8 29525 bench.py --gpus 8 --steps 3 --warmup 3 --workload c5 > gpurun_out/bench_8_c5.json 2> gpurun_out/bench_8_c5.err; echo "c5 rc=$?"; tail -3 gpurun_out/bench_8_c5.err; cut -c1-400 gpurun_out/bench_8_c5.json
run 8 29526 bench.py --gpus 8 --steps 10 --warmup 3 --causal > gpurun_out/bench_8_causal.json 2> gpurun_out/bench_8_causal.err; echo "causal rc=$?"; cut -c1-400 gpurun_out/bench_8_causal.json
