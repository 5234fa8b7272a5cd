#!/bin/bash
# final multi-GPU evidence of round 2: ring battery at 8 ranks (home backward included), scaling series 8/4/2/1 on one box, C5
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { local n=$1 port=$2; shift 2; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
run 8 29551 tests/ring_worker.py > gpurun_out/ring_worker_8.log 2>&1; echo "ring_worker 8 rc=$?"; grep -c "RING_OK" gpurun_out/ring_worker_8.log
run 8 29552 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_8.json 2> gpurun_out/bench_8.err; echo "bench 8 rc=$?"; cut -c1-300 gpurun_out/bench_8.json
CUDA_VISIBLE_DEVICES=0,1,2,3 run 4 29553 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/bench_4.json 2> gpurun_out/bench_4.err; echo "bench 4 rc=$?"
CUDA_VISIBLE_DEVICES=0,1 run 2 29554 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_2.json 2> gpurun_out/bench_2.err; echo "bench 2 rc=$?"
CUDA_VISIBLE_DEVICES=0 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_1.json 2> gpurun_out/bench_1.err; echo "bench 1 rc=$?"
run 8 29555 bench.py --gpus 8 --steps 3 --warmup 3 --workload c5 > gpurun_out/bench_8_c5.json 2> gpurun_out/bench_8_c5.err; echo "c5 rc=$?"; tail -2 gpurun_out/bench_8_c5.err
CUDA_VISIBLE_DEVICES=0,1,2,3 run 4 29556 tests/ring_worker.py > gpurun_out/ring_worker_4.log 2>&1; echo "ring_worker 4 rc=$?"; grep -c "RING_OK" gpurun_out/ring_worker_4.log
