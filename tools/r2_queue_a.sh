#!/bin/bash
# round-2 first GPU call: the ready-to-run queue DESIGN.md 7 left behind (each step under its own timeout)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/a_smi.txt 2>&1
nproc >> gpurun_out/a_smi.txt
bash tools/try_bwd_variant.sh 1 > /dev/null 2>&1
bash tools/time_configs.sh > /dev/null 2>&1
bash tools/try_suspend.sh run > /dev/null 2>&1
B200_DBUF_POLY=4 B200_DBUF_PEEL=1 bash tools/try_fwd_variant.sh 4 > /dev/null 2>&1; mv gpurun_out/try_fwd_variant_4.log gpurun_out/try_fwd_variant_4_peel.log
B200_DBUF_POLY=4 bash tools/try_fwd_variant.sh 4 > /dev/null 2>&1
bash tools/promote_fwd_variant.sh > gpurun_out/promote.log 2>&1
tail -5 gpurun_out/promote.log
