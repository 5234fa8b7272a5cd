#!/bin/bash
# GPU check of an experimental forward variant (B200_FWD_VARIANT=$1; 3 = CTA pairs, 4 = double-buffered 64-key tiles):
# parity against the generic CUDA-core kernels, then timing beside the default kernel at C3.  Native harness only (no
# Python start-up), every step under its own timeout, stops at the first failure: a wrong barrier protocol traps through
# the mbarrier watchdog (ptx.cuh) instead of hanging the box.
#   gpurun --timeout 200 -- 'bash tools/try_fwd_variant.sh 4'
cd "$(dirname "$0")/.."
V=${1:-4}
mkdir -p gpurun_out
out=gpurun_out/try_fwd_variant_$V.log
: > $out
[ -x build/kbench ] || nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o build/kbench tools/kbench.cu -ldl
# (the variants live in the side library only: make -C cuda_flashattention_b200 EXPERIMENTS=1)
export B200ATTN_LIB=${B200ATTN_LIB:-cuda_flashattention_b200/libb200attn_exp.so}
run() {
    echo "\$ $*" >> $out
    timeout 25 env "$@" >> $out 2>&1
    local rc=$?
    echo "[exit $rc]" >> $out
    if [ $rc -ne 0 ]; then echo "stopped at the first failure" >> $out; cat $out; exit 1; fi
}
run B200_FWD_VARIANT=$V build/kbench fwd 1 1 512 128 bf16 2 --check
run B200_FWD_VARIANT=$V build/kbench fwd 1 3 1024 128 bf16 2 --check
if [ "$V" != 3 ]; then          # the pair variant only takes full tiles, non-causal (other shapes fall back to the default)
    run B200_FWD_VARIANT=$V build/kbench fwd 1 2 64 128 bf16 2 --check
    run B200_FWD_VARIANT=$V build/kbench fwd 1 2 1000 128 bf16 2 --check
    run B200_FWD_VARIANT=$V build/kbench fwd 1 2 1024 128 bf16 2 --check --causal
    run B200_FWD_VARIANT=$V build/kbench fwd 1 2 1472 128 bf16 2 --check --causal
    run B200_FWD_VARIANT=$V build/kbench fwd 1 2 512 128 bf16 2 --check --nkv 2048     # cross lengths (ring steps)
    run B200_FWD_VARIANT=$V build/kbench fwd 1 2 1000 128 bf16 2 --check --nkv 200
fi
run B200_FWD_VARIANT=$V build/kbench fwd 2 4 4096 128 bf16 3 --check
run B200_FWD_VARIANT=1 build/kbench fwd 8 16 8192 128 bf16 10
run B200_FWD_VARIANT=$V build/kbench fwd 8 16 8192 128 bf16 10
run B200_FWD_VARIANT=1 build/kbench fwd 8 16 8192 128 bf16 10
run B200_FWD_VARIANT=$V build/kbench fwd 8 16 8192 128 bf16 10
if [ "$V" != 3 ]; then
    run B200_FWD_VARIANT=1 build/kbench fwd 8 16 8192 128 bf16 10 --causal
    run B200_FWD_VARIANT=$V build/kbench fwd 8 16 8192 128 bf16 10 --causal
fi
cat $out
