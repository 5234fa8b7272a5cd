#!/bin/bash
# A/B of the backward with the paired M1 (S^T of two q tiles as one N = 128 MMA group; product library) against the
# previous kernel (libb200attn_old.so = `make VARIANT=old` at the parent commit): parity against the generic CUDA-core
# kernels on shapes that hit every tail (1 tile, odd tile counts, ragged, causal, d = 64), then C3 timings, alternating.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/r2h_bwd_pair_ab.log; : > $out
NEW=cuda_flashattention_b200/libb200attn.so
OLD=cuda_flashattention_b200/libb200attn_old.so
run() { echo "\$ $*" >> $out; timeout 60 env "$@" >> $out 2>&1; local rc=$?; echo "[exit $rc]" >> $out; if [ $rc -ne 0 ]; then tail -n 30 $out; exit 1; fi; }
for shape in "1 1 64 128" "1 2 100 128" "1 2 192 128" "2 3 1000 128" "1 2 1472 128" "1 2 448 64" "2 2 1000 64"; do
  run B200ATTN_LIB=$NEW build/kbench bwd $shape bf16 2 --check
done
run B200ATTN_LIB=$NEW build/kbench bwd 1 2 1472 128 bf16 2 --check --causal
run B200ATTN_LIB=$NEW build/kbench bwd 1 2 960 128 bf16 2 --check --causal
run B200ATTN_LIB=$NEW build/kbench bwd 2 4 4096 128 bf16 3 --check
if [ -f $OLD ]; then
  for rep in 1 2 3; do
    run B200ATTN_LIB=$OLD build/kbench bwd 8 16 8192 128 bf16 12
    run B200ATTN_LIB=$NEW build/kbench bwd 8 16 8192 128 bf16 12
  done
  run B200ATTN_LIB=$OLD build/kbench bwd 8 16 8192 128 bf16 12 --causal
  run B200ATTN_LIB=$NEW build/kbench bwd 8 16 8192 128 bf16 12 --causal
  run B200ATTN_LIB=$OLD build/kbench bwd 8 16 8192 64 bf16 12
  run B200ATTN_LIB=$NEW build/kbench bwd 8 16 8192 64 bf16 12
else
  run B200ATTN_LIB=$NEW build/kbench bwd 8 16 8192 128 bf16 12
fi
grep -v "^\[exit" $out | tail -n 60
echo "PASS lines: $(grep -c PASS $out)"
