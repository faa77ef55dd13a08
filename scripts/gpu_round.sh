#!/bin/bash
# one GPU session: probe, parity tests, bench at increasing N, launch list.  Everything is logged under gpurun_out/.
mkdir -p gpurun_out
exec > >(tee gpurun_out/round.log) 2>&1
echo "== probe"; nproc; free -g | sed -n 2p; nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,power.limit --format=csv; which mpirun mpiexec
echo "== pytest -m gpu"; timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
echo "== smoke"; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
for N in ${BENCH_NS:-200000 1000000}; do
  echo "== bench N=$N"; timeout ${BENCH_TIMEOUT:-1500} python bench.py --n $N --steps 10 --warmup 3 2> gpurun_out/bench_$N.err | tee gpurun_out/bench_$N.json; tail -8 gpurun_out/bench_$N.err
done
if [ -n "$NCU" ]; then
  echo "== ncu launch list"
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; tail -2 gpurun_out/ncu_launches.log
  echo "== ncu full (stream kernels of the 2nd product)"
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:hmatrix_stream -s 2 -c 2 -f -o gpurun_out/stream_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; tail -2 gpurun_out/ncu_full.log
  ls -la gpurun_out
fi
