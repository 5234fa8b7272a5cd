#!/bin/bash
# First GPU run of the CTA-pair forward (B200_FWD_VARIANT=3, csrc/fwd_pair_sm100.cu): parity against the generic
# CUDA-core kernels on small and mid-sized problems, then timing beside the default kernel at C3.  Native harness
# only (no Python start-up), every step under its own timeout, stops at the first failure: a wrong barrier protocol
# traps through the mbarrier watchdog (ptx.cuh) instead of hanging the box.
#   gpurun --timeout 300 -- 'bash tools/try_fwd_pair.sh'
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/try_fwd_pair.log
: > $out
[ -x build/kbench ] || nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o build/kbench tools/kbench.cu -ldl
run() {
    echo "\$ $*" >> $out
    timeout 25 env "$@" >> $out 2>&1
    local rc=$?
    echo "[exit $rc]" >> $out
    if [ $rc -ne 0 ]; then echo "stopped at the first failure" >> $out; cat $out; exit 1; fi
}
run B200_FWD_VARIANT=3 build/kbench fwd 1 1 512 128 bf16 2 --check
run B200_FWD_VARIANT=3 build/kbench fwd 1 3 1024 128 bf16 2 --check
run B200_FWD_VARIANT=3 build/kbench fwd 2 4 4096 128 bf16 3 --check
run B200_FWD_VARIANT=1 build/kbench fwd 8 16 8192 128 bf16 10
run B200_FWD_VARIANT=3 build/kbench fwd 8 16 8192 128 bf16 10
run B200_FWD_VARIANT=1 build/kbench fwd 8 16 8192 128 bf16 10
run B200_FWD_VARIANT=3 build/kbench fwd 8 16 8192 128 bf16 10
cat $out
