#!/bin/bash
# A/B of a backward variant in the side library (B200_BWD_VARIANT=$1) against the shipped kernel: parity vs the generic kernels, then C3 timing
cd "$(dirname "$0")/.."
V=${1:-4}
mkdir -p gpurun_out
out=gpurun_out/bwd_ab_$V.log; : > $out
export B200ATTN_LIB=cuda_flashattention_b200/libb200attn_exp.so
run() { echo "\$ $*" >> $out; timeout 60 env "$@" >> $out 2>&1; local rc=$?; echo "[exit $rc]" >> $out; if [ $rc -ne 0 ]; then cat $out; exit 1; fi; }
run B200_BWD_VARIANT=$V build/kbench bwd 1 2 1000 128 bf16 2 --check
run B200_BWD_VARIANT=$V build/kbench bwd 1 2 1472 128 bf16 2 --check --causal
run B200_BWD_VARIANT=$V build/kbench bwd 2 4 4096 128 bf16 3 --check
for rep in 1 2 3; do
run B200_BWD_VARIANT=0 build/kbench bwd 8 16 8192 128 bf16 12
run B200_BWD_VARIANT=$V build/kbench bwd 8 16 8192 128 bf16 12
done
run B200_BWD_VARIANT=0 build/kbench bwd 8 16 8192 128 bf16 12 --causal
run B200_BWD_VARIANT=$V build/kbench bwd 8 16 8192 128 bf16 12 --causal
grep -v "^check\|^\[exit" $out
grep -c PASS $out
