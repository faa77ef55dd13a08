#!/bin/bash
# diagnosis of the 8-GPU step time: clock sampler on/off, 20 vs 200 steps, fused vs NCCL all-gather
mkdir -p gpurun_out
exec > >(tee gpurun_out/scale2.log) 2>&1
run() { # label, env..., extra args
  label=$1; shift
  echo "== $label"
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus 8 --warmup 3 $EXTRA 2> gpurun_out/scale2_$label.err | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(json.dumps({k: d[k] for k in ('value','ms_per_step','steps','clocks','gpu_launches')} | {'kernel_ms': d['roofline']['kernel_ms'], 'e2e': d['e2e']['value'], 'allgather': d['config']['allgather'][:12]}))
"
}
EXTRA="--steps 20" run sampler_20 A=1
EXTRA="--steps 20" run nosampler_20 HMB200_BENCH_NO_SAMPLER=1
EXTRA="--steps 200" run sampler_200 A=1
EXTRA="--steps 200" run nosampler_200 HMB200_BENCH_NO_SAMPLER=1
EXTRA="--steps 200" run nccl_nosampler_200 HMB200_BENCH_NO_SAMPLER=1 HMB200_NO_FUSED=1
