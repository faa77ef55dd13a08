#!/bin/bash
# 2-GPU session: distributed tests (fused all-gather over peer memory, device GMRES on 2 ranks), bench at N=1M on 2 GPUs
mkdir -p gpurun_out
exec > >(tee gpurun_out/r3.log) 2>&1
nvidia-smi --query-gpu=name --format=csv,noheader; nvidia-smi topo -m | head -6
echo "== pytest distributed/fused/graphs"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
echo "== bench 2 GPUs"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 2> gpurun_out/bench_2gpu.err | tee gpurun_out/bench_2gpu.json; grep -v "^W\|^\*" gpurun_out/bench_2gpu.err | tail -12
echo "== bench 2 GPUs gmres"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 --gmres --no-cpu-baseline 2> gpurun_out/bench_2gpu_gmres.err | tee gpurun_out/bench_2gpu_gmres.json; grep -v "^W\|^\*" gpurun_out/bench_2gpu_gmres.err | tail -6
