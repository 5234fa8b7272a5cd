#!/bin/bash
# Gate for promoting an experimental forward kernel to the default: the forward-facing GPU parity
# suite under the variant, then the contract bench and an ncu launch list with the variant selected through the
# environment.  (The backward's B200_BWD_VARIANT=1 goes through tests/test_gpu_kernel_variants.py the same way.)
# One gpurun call, about 10 GPU-minutes:
#   gpurun --timeout 900 -- 'bash tools/promote_fwd_variant.sh'
# If everything is green, make the variant the default in dispatch_fwd (csrc/fwd_sm100.cu: `variant = e ? atoi(e) : 1`)
# and commit the new bench line / launch list under profiles/.
cd "$(dirname "$0")/.."
export B200_FWD_VARIANT=${B200_FWD_VARIANT:-4} B200_DBUF_POLY=${B200_DBUF_POLY:-4}
mkdir -p gpurun_out
set -x
python -m pytest -x -q -m gpu tests/test_gpu_fwd_tc.py tests/test_gpu_causal.py tests/test_gpu_layouts.py tests/test_gpu_host.py \
    tests/test_gpu_ring.py > gpurun_out/promote_pytest.log 2>&1 || { tail -30 gpurun_out/promote_pytest.log; exit 1; }
tail -3 gpurun_out/promote_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/promote_bench.json 2> gpurun_out/promote_bench.err || { tail gpurun_out/promote_bench.err; exit 1; }
cat gpurun_out/promote_bench.json
B200_FWD_VARIANT=1 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/promote_bench_default.json 2>> gpurun_out/promote_bench.err
cat gpurun_out/promote_bench_default.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/promote_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/promote_ncu.log 2>&1
tail -3 gpurun_out/promote_ncu.log
