#!/bin/bash
# Native timing of the single-GPU BASELINE.json configs (no Python start-up; a few seconds of GPU time):
#   C2  FA2 forward B=1 H=8 N=4096 d=64  -- fp32 I/O (tf32 tensor cores, the reference's type) and bf16
#   C3  FA2 forward + backward B=8 H=16 N=8192 d=128 bf16 (non-causal, and the causal extension)
#   gpurun --timeout 120 -- 'bash tools/time_configs.sh'
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
[ -x build/kbench ] || nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o build/kbench tools/kbench.cu -ldl
{
  timeout 60 build/kbench fwd 1 8 4096 64 f32 50 --check
  timeout 60 build/kbench fwd 1 8 4096 64 bf16 50 --check
  timeout 60 build/kbench both 8 16 8192 128 bf16 10
  timeout 60 build/kbench both 8 16 8192 128 bf16 10 --causal
} 2>&1 | tee gpurun_out/time_configs.log
