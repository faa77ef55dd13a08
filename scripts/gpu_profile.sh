#!/bin/bash
# profiling session (1 GPU): each command first runs WITHOUT ncu, then under ncu
mkdir -p gpurun_out
exec > >(tee gpurun_out/profile.log) 2>&1
echo "== bench plain"; timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/p_bench.json 2> gpurun_out/p_bench.err && cut -c1-200 gpurun_out/p_bench.json
echo "== ncu launch list (bench)"; timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/p_ncu_bench.log 2>&1; tail -1 gpurun_out/p_ncu_bench.log | cut -c1-200
echo "== gmres plain"; timeout 600 python scripts/gmres_profile.py 1000000 45 2> gpurun_out/p_gmres.err
echo "== ncu launch list (gmres)"; timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_gmres.csv python scripts/gmres_profile.py 1000000 45 > gpurun_out/p_ncu_gmres.log 2>&1; tail -1 gpurun_out/p_ncu_gmres.log
echo "== ncu full (krylov kernels late in the cycle)"; timeout 900 ncu --set full --clock-control none --import-source on -k regex:krylov_dots\|krylov_combine -s 60 -c 4 -f -o gpurun_out/krylov_full python scripts/gmres_profile.py 1000000 45 > gpurun_out/p_ncu_krylov.log 2>&1; tail -1 gpurun_out/p_ncu_krylov.log
ls -la gpurun_out | grep -i "ncu-rep\|launches"
