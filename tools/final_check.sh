#!/bin/bash
# Round-end verification on one B200: GPU tests, smoke, both bench arms, prefill / per-kernel timings, ncu launch list.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -x 2>&1 | grep -E "passed|failed|rror" | tail -5
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 400 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; tail -c 1500 gpurun_out/bench_final.json
timeout 400 python bench.py --impl reference --steps 8 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; tail -c 600 gpurun_out/bench_ref.json
timeout 200 python tools/prefill_bench.py > gpurun_out/prefill_noncausal.txt 2>&1; tail -4 gpurun_out/prefill_noncausal.txt
timeout 200 python tools/prefill_bench.py --causal 1 > gpurun_out/prefill_causal.txt 2>&1; tail -4 gpurun_out/prefill_causal.txt
timeout 200 python tools/kernel_bench.py > gpurun_out/kernel_bench.txt 2>&1; grep rows= gpurun_out/kernel_bench.txt
K='regex:gemv|attn|gemm|rmsnorm|argmax|embed|prep_kernel|kv_alloc|load_residual|lastrow|convert'
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" --launch-skip 2000 -c 1500 --csv --log-file gpurun_out/launches_r2_ncu.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_run.log 2>&1; echo "ncu rc=$?"
