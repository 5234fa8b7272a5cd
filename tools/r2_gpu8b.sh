#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { local n=$1 port=$2; shift 2; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
run 8 29541 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_8.json 2> gpurun_out/bench_8.err; echo "bench 8 rc=$?"; cut -c1-600 gpurun_out/bench_8.json
run 8 29542 tools/ring_modes.py 1,2,16384 1,4,8192 1,8,8192 1,16,8192 1,4,32768 2>&1 | grep "^P=" | tee gpurun_out/ring_modes_8.txt
CUDA_VISIBLE_DEVICES=0,1,2,3 run 4 29543 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/bench_4.json 2> gpurun_out/bench_4.err; echo "bench 4 rc=$?"; cut -c1-300 gpurun_out/bench_4.json
CUDA_VISIBLE_DEVICES=0,1 run 2 29544 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_2.json 2> gpurun_out/bench_2.err; echo "bench 2 rc=$?"; cut -c1-300 gpurun_out/bench_2.json
CUDA_VISIBLE_DEVICES=0 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_1.json 2> gpurun_out/bench_1.err; echo "bench 1 rc=$?"; cut -c1-300 gpurun_out/bench_1.json
