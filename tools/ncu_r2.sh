#!/bin/bash
# Round-2 profiler evidence on one B200 (run under gpurun AFTER the plain bench has exited 0): the launch list of the
# default bench command, --set full captures of the decode GEMVs (one layer's four, gate/up first) and of the batched-decode
# kernels (tensor-core attention + streaming GEMMs at 32 rows x 640 keys).
mkdir -p gpurun_out
K='regex:gemv|attn|gemm|rmsnorm|argmax|embed|prep_kernel|kv_alloc|load_residual|lastrow|convert'
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" --launch-skip 3000 -c 1200 --csv \
  --log-file gpurun_out/launches_r2_ncu.csv python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-extras --no-parity > gpurun_out/ncu_run_r2.log 2>&1; echo "launch list rc=$?"
# gemv launches per layer: qkv, o_proj, gate/up, down -> skip a multiple of 4 plus 2 to start at gate/up
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemv_kernel --launch-skip 2002 -c 4 \
  -o gpurun_out/prof_gemv_r2 -f python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-extras --no-parity > gpurun_out/ncu_full_r2.log 2>&1; echo "set full gemv rc=$?"
ncu -i gpurun_out/prof_gemv_r2.ncu-rep --page raw --csv > gpurun_out/ncu_full_gemv_r2_raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gemm_stream_tc_kernel|attn_decode_tc_kernel' --launch-skip 60 -c 10 \
  -o gpurun_out/prof_batch_r2 -f python tools/batch_decode_bench.py --rows 32 --layers 2 --steps 8 --context 640 > gpurun_out/ncu_batch_r2.log 2>&1; echo "set full batch rc=$?"
ncu -i gpurun_out/prof_batch_r2.ncu-rep --page raw --csv > gpurun_out/ncu_full_batch32_r2_raw.csv 2>/dev/null
tail -3 gpurun_out/ncu_batch_r2.log
ls -la gpurun_out | grep r2_raw
