#!/bin/bash
# Multi-GPU confirmation runs.  usage: tools/multi_gpu_check.sh <n_gpus_on_box>
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run() { # name nproc args...
  name=$1; np=$2; shift 2
  timeout 420 $TR --nproc-per-node $np --master-port $((29500 + RANDOM % 400)) bench.py --gpus $np "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  echo "== $name rc=$?"; grep '^{' gpurun_out/$name.json | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print({k: d.get(k) for k in ('value', 'ms_per_step', 'prefill_tok_s', 'prompt_tok_s', 'hop', 'hop_error_flag')}, d['e2e']['value'] if 'e2e' in d else None)" 2>/dev/null || tail -3 gpurun_out/$name.err
}
if [ "$N" = "2" ]; then
  timeout 300 python -m pytest tests/test_gpu_chain.py -q -m gpu -x 2>&1 | grep -E "passed|failed|rror" | tail -3
  run bench_n2 2 --steps 128 --warmup 4
  run bench_n2_c4 2 --streams 8 --rows 2 --prompt-len 2048 --steps 16 --warmup 3
  run bench_n2_c3_batch32 2 --streams 2 --rows 16 --prompt-len 128 --steps 128 --warmup 4
fi
if [ "$N" = "4" ]; then
  run bench_n3_c5_mixed 3 --splits 4,12,16 --mixed --steps 96 --warmup 4
  run bench_n4 4 --steps 128 --warmup 4
  run bench_n4_c3_batch32 4 --streams 4 --rows 8 --prompt-len 128 --steps 128 --warmup 4
fi
if [ "$N" = "8" ]; then
  run bench_n8 8 --steps 128 --warmup 4
  run bench_n8_c3_batch32 8 --streams 8 --rows 4 --prompt-len 128 --steps 128 --warmup 4
  run bench_n8_c4_prefill 8 --streams 8 --rows 2 --prompt-len 2048 --steps 16 --warmup 3
fi
