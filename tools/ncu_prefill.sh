#!/bin/bash
# ncu --set full on the prefill kernels of ONE layer at 4 x 2048 tokens (after the same command ran plain).
mkdir -p gpurun_out
CMD="python tools/prefill_bench.py --layers 1 --cases 4x2048 --iters 2"
timeout 200 $CMD > gpurun_out/ncu_prefill_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_prefill_plain.log; exit 1; }
tail -1 gpurun_out/ncu_prefill_plain.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"gemm_pair|gemm_tcgen05|attn_prefill_tc" --launch-skip 5 --launch-count 5 \
    -o gpurun_out/prefill_full -f $CMD > gpurun_out/ncu_prefill_run.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/prefill_full.ncu-rep --page raw --csv > gpurun_out/ncu_full_prefill_r1_v2_raw.csv 2>/dev/null; wc -c gpurun_out/ncu_full_prefill_r1_v2_raw.csv
