#!/bin/bash
# GPU check of the experimental backward issue loop (B200_BWD_VARIANT=1, bwd_kernel<128, 1>): gradient parity against the
# generic CUDA-core kernels, then timing beside the shipped kernel at C3.  Native harness, seconds of GPU time.
#   gpurun --timeout 200 -- 'bash tools/try_bwd_variant.sh'
cd "$(dirname "$0")/.."
V=${1:-1}
mkdir -p gpurun_out
out=gpurun_out/try_bwd_variant_$V.log
: > $out
[ -x build/kbench ] || nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o build/kbench tools/kbench.cu -ldl
run() {
    echo "\$ $*" >> $out
    timeout 40 env "$@" >> $out 2>&1
    local rc=$?
    echo "[exit $rc]" >> $out
    if [ $rc -ne 0 ]; then echo "stopped at the first failure" >> $out; cat $out; exit 1; fi
}
run B200_BWD_VARIANT=$V build/kbench bwd 1 1 256 128 bf16 2 --check          # 4 iterations: all in the head loop
run B200_BWD_VARIANT=$V build/kbench bwd 1 2 1024 128 bf16 2 --check         # 16 iterations: one unrolled block + tail
run B200_BWD_VARIANT=$V build/kbench bwd 1 2 1000 128 bf16 2 --check         # ragged
run B200_BWD_VARIANT=$V build/kbench bwd 1 2 1472 128 bf16 2 --check --causal
run B200_BWD_VARIANT=$V build/kbench bwd 2 4 4096 128 bf16 3 --check
run B200_BWD_VARIANT=0 build/kbench bwd 8 16 8192 128 bf16 12
run B200_BWD_VARIANT=$V build/kbench bwd 8 16 8192 128 bf16 12
run B200_BWD_VARIANT=0 build/kbench bwd 8 16 8192 128 bf16 12
run B200_BWD_VARIANT=$V build/kbench bwd 8 16 8192 128 bf16 12
run B200_BWD_VARIANT=0 build/kbench bwd 8 16 8192 128 bf16 12 --causal
run B200_BWD_VARIANT=$V build/kbench bwd 8 16 8192 128 bf16 12 --causal
cat $out
