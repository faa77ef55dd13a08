#!/bin/bash
# 8-GPU box: the metric at N = 8, 4, 1 GPUs (driver settings), BASELINE configs[4] (N=2M GMRES on 8 GPUs), configs[3] (64 RHS on 8 GPUs)
mkdir -p gpurun_out
exec > >(tee gpurun_out/scale3.log) 2>&1
tr() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) bench.py --gpus $1 "${@:2}"; }
for N in 8 4; do
  echo "== bench $N GPUs"; timeout 600 bash -c "$(declare -f tr); tr $N --steps 20 --warmup 3" 2> gpurun_out/s3_$N.err | tee gpurun_out/s3_$N.json
done
echo "== bench 1 GPU"; timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline 2> gpurun_out/s3_1.err | tee gpurun_out/s3_1.json
echo "== bench 8 GPUs, 200 steps"; timeout 600 bash -c "$(declare -f tr); tr 8 --steps 200 --warmup 3" 2> gpurun_out/s3_8_200.err | tee gpurun_out/s3_8_200.json
echo "== N=2M GMRES on 8 GPUs"; timeout 900 bash -c "$(declare -f tr); tr 8 --steps 20 --warmup 3 --n 2000000 --gmres --no-cpu-baseline" 2> gpurun_out/s3_gmres2M.err | tee gpurun_out/s3_gmres2M.json; grep "gmres" gpurun_out/s3_gmres2M.err | tail -3
echo "== 64 RHS on 8 GPUs"; timeout 600 bash -c "$(declare -f tr); tr 8 --steps 5 --warmup 3 --rhs 64" 2> gpurun_out/s3_rhs64.err | tee gpurun_out/s3_rhs64.json
tail -3 gpurun_out/s3_rhs64.err
