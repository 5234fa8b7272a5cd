#!/bin/bash
# Subtraction profile of the backward kernel (timing only, wrong results): what each component costs at C3.
# Needs libb200attn_exp.so (make -C cuda_flashattention_b200 EXPERIMENTS=1).  B200_BWD_DBG bits: see csrc/bwd_sm100.cu.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/bwd_subtract.log
: > $out
for dbg in 0 1 4 8 5 13 16 32 48 64 128 0; do
    echo "== B200_BWD_DBG=$dbg" >> $out
    B200ATTN_LIB=cuda_flashattention_b200/libb200attn_exp.so B200_BWD_DBG=$dbg timeout 60 build/kbench bwd 8 16 8192 128 bf16 10 >> $out 2>&1
done
cat $out
