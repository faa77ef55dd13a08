#!/bin/bash
mkdir -p gpurun_out
exec > >(tee gpurun_out/r4.log) 2>&1
echo "== pytest -m gpu"; timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
echo "== gmres plain"; timeout 600 python scripts/gmres_profile.py 1000000 45 2> gpurun_out/r4_gmres.err
echo "== ncu launch list (gmres)"; timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_gmres.csv python scripts/gmres_profile.py 1000000 45 > gpurun_out/r4_ncu_gmres.log 2>&1; tail -1 gpurun_out/r4_ncu_gmres.log
echo "== bench N=1M gmres"; timeout 1500 python bench.py --steps 20 --warmup 3 --gmres 2> gpurun_out/r4_bench.err | tee gpurun_out/r4_bench_1M_gmres.json | cut -c1-150; grep "gmres" gpurun_out/r4_bench.err
