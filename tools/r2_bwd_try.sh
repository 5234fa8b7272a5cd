#!/bin/bash
# round-2 backward experiments: parity (vs the generic CUDA-core kernels) and timing of B200_BWD_VARIANT=$1 beside variant 0
cd "$(dirname "$0")/.."
V=${1:-2}
mkdir -p gpurun_out
out=gpurun_out/bwd_try_$V.log
: > $out
run() {
    echo "\$ $*" >> $out
    timeout 60 env "$@" >> $out 2>&1
    local rc=$?
    echo "[exit $rc]" >> $out
    if [ $rc -ne 0 ]; then echo "stopped at the first failure" >> $out; cat $out; exit 1; fi
}
run B200_BWD_VARIANT=$V build/kbench bwd 1 1 256 128 bf16 2 --check
run B200_BWD_VARIANT=$V build/kbench bwd 1 2 64 128 bf16 2 --check
run B200_BWD_VARIANT=$V build/kbench bwd 1 2 1000 128 bf16 2 --check
run B200_BWD_VARIANT=$V build/kbench bwd 1 2 1472 128 bf16 2 --check --causal
run B200_BWD_VARIANT=$V build/kbench bwd 2 4 4096 128 bf16 3 --check
run B200_BWD_VARIANT=$V build/kbench bwd 1 3 1000 64 bf16 2 --check
run B200_BWD_VARIANT=$V build/kbench bwd 2 4 2048 64 bf16 2 --check --causal
for rep in 1 2; do
run B200_BWD_VARIANT=0 build/kbench bwd 8 16 8192 128 bf16 12
run B200_BWD_VARIANT=$V build/kbench bwd 8 16 8192 128 bf16 12
done
run B200_BWD_VARIANT=0 build/kbench bwd 8 16 8192 128 bf16 12 --causal
run B200_BWD_VARIANT=$V build/kbench bwd 8 16 8192 128 bf16 12 --causal
run B200_BWD_VARIANT=0 build/kbench bwd 8 16 8192 64 bf16 12
run B200_BWD_VARIANT=$V build/kbench bwd 8 16 8192 64 bf16 12
run B200_BWD_VARIANT=0 build/kbench bwd 8 16 8192 128 bf16 300
run B200_BWD_VARIANT=$V build/kbench bwd 8 16 8192 128 bf16 300
if [ -f cuda_flashattention_b200/libb200attn_trace.so ]; then
run B200ATTN_LIB=cuda_flashattention_b200/libb200attn_trace.so B200_BWD_VARIANT=$V build/kbench bwd 8 16 8192 128 bf16 3 --trace
fi
cat $out
