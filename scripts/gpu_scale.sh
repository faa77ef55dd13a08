#!/bin/bash
# scaling session on one 8-GPU box: bench at N = 8, 4 (and 1 for the same-box ratio)
mkdir -p gpurun_out
exec > >(tee gpurun_out/scale.log) 2>&1
nvidia-smi --query-gpu=name --format=csv,noheader | sort | uniq -c
for N in ${SCALE_NS:-8 4 1}; do
  echo "== bench $N GPUs"
  if [ "$N" = 1 ]; then
    timeout 900 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline 2> gpurun_out/scale_$N.err | tee gpurun_out/scale_$N.json
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520+N)) bench.py --gpus $N --steps 20 --warmup 3 2> gpurun_out/scale_$N.err | tee gpurun_out/scale_$N.json
  fi
  grep "^\[bench\]" gpurun_out/scale_$N.err | tail -4
done
