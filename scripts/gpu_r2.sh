#!/bin/bash
# GPU session: parity tests, smoke, bench (default N=1M) with the GMRES extra.  Logs under gpurun_out/.
mkdir -p gpurun_out
exec > >(tee gpurun_out/r2.log) 2>&1
echo "== probe"; nproc; nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv
echo "== pytest -m gpu"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
echo "== bench N=200k gmres"; timeout 600 python bench.py --n 200000 --steps 10 --warmup 3 --gmres 2> gpurun_out/bench_200k.err | tee gpurun_out/bench_200k_gmres.json; tail -12 gpurun_out/bench_200k.err
echo "== bench N=1M gmres"; timeout 1500 python bench.py --steps 20 --warmup 3 --gmres 2> gpurun_out/bench_1M.err | tee gpurun_out/bench_1M_gmres.json; tail -12 gpurun_out/bench_1M.err
