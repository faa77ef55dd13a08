#!/bin/bash
# final 1-GPU session: parity tests, smoke, both bench arms, ncu capture of the stream kernels, per-slab scan
mkdir -p gpurun_out
exec > >(tee gpurun_out/final1.log) 2>&1
echo "== pytest -m gpu"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
echo "== bench (b200 arm)"; timeout 900 python bench.py --steps 20 --warmup 3 2> gpurun_out/f1_bench.err | tee gpurun_out/f1_bench.json | cut -c1-200; grep "^\[bench\]" gpurun_out/f1_bench.err | tail -4
echo "== bench (reference arm)"; timeout 900 python bench.py --impl reference --steps 5 --warmup 1 2> gpurun_out/f1_ref.err | tee gpurun_out/f1_ref.json | cut -c1-300
echo "== ncu full (stream kernels, after the plain run above exited 0)"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:hmatrix_stream -s 8 -c 2 -f -o gpurun_out/stream_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/f1_ncu_full.log 2>&1; tail -1 gpurun_out/f1_ncu_full.log | cut -c1-120
echo "== slab scan P=8"; timeout 600 python scripts/slab_scan.py 1000000 8 2> /dev/null | tee gpurun_out/slab_scan_P8.jsonl | cut -c1-420
