#!/bin/bash
# BASELINE configs[4]: GMRES on the N=2M Laplace operator, 8 GPUs (device-resident Krylov basis, fused/NCCL collectives)
mkdir -p gpurun_out
export HMB200_BENCH_N=2000000
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29811 bench.py --gpus 8 --steps 20 --warmup 3 --gmres --no-cpu-baseline 2> gpurun_out/gmres2M_8gpu.err | tee gpurun_out/gmres2M_8gpu.json
grep "^\[bench\]" gpurun_out/gmres2M_8gpu.err | tail -6
