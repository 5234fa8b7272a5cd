#!/bin/bash
# Does parking the polling threads help under the power cap?  mbarrier.try_wait takes a suspend-time hint
# (B200_MBAR_SUSPEND_NS in csrc/ptx.cuh, default 0 = plain polling).  Every kernel here keeps a producer thread, one or two
# MMA-issuing threads and idle softmax / compute warps spinning on try_wait; under the 1 kW cap the sustained clock, not the
# burst one, decides bench.py's numbers.  This builds side libraries with the hint set (HERE, before gpurun: the .so files
# travel with the snapshot) and times 300 back-to-back launches of forward and backward with each.
#   bash tools/try_suspend.sh build          # in the build container
#   gpurun --timeout 300 -- 'bash tools/try_suspend.sh run'
cd "$(dirname "$0")/.."
NS_LIST="200 1000 5000"
if [ "$1" = build ]; then
    for ns in $NS_LIST; do make -C cuda_flashattention_b200 -j4 VARIANT=s$ns EXTRA=-DB200_MBAR_SUSPEND_NS=$ns > /dev/null || exit 1; done
    ls -la cuda_flashattention_b200/libb200attn_s*.so
    exit 0
fi
mkdir -p gpurun_out
out=gpurun_out/try_suspend.log
: > $out
for lib in libb200attn.so $(for ns in $NS_LIST; do echo libb200attn_s$ns.so; done); do
    [ -f cuda_flashattention_b200/$lib ] || continue
    echo "== $lib" >> $out
    B200ATTN_LIB=cuda_flashattention_b200/$lib timeout 120 build/kbench both 8 16 8192 128 bf16 300 >> $out 2>&1
done
cat $out
