#!/bin/bash
# final 8-GPU session: the metric at 8 and 1 GPUs (driver settings), configs[4] (N=2M GMRES on 8 GPUs, peer collectives)
mkdir -p gpurun_out
exec > >(tee gpurun_out/scale4.log) 2>&1
tr8() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) bench.py --gpus 8 "$@"; }
echo "== bench 8 GPUs"; timeout 600 bash -c "$(declare -f tr8); tr8 --steps 20 --warmup 3" 2> gpurun_out/s4_8.err | tee gpurun_out/s4_8.json | cut -c1-220
echo "== bench 1 GPU"; timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline 2> gpurun_out/s4_1.err | tee gpurun_out/s4_1.json | cut -c1-220
echo "== N=2M GMRES on 8 GPUs"; HMB200_BENCH_N=2000000 timeout 900 bash -c "$(declare -f tr8); tr8 --steps 20 --warmup 3 --gmres --no-cpu-baseline" 2> gpurun_out/s4_gmres2M.err | tee gpurun_out/s4_gmres2M.json | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['value'], d['ms_per_step'], d['gmres'])
"
grep -i "Traceback\|Error" gpurun_out/s4_*.err | head -5
