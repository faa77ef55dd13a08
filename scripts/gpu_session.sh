#!/bin/bash
# One parameterised GPU session: scripts/gpu_session.sh <tag> <step> [<step> ...]
# Every step runs under its own `timeout`; output goes to gpurun_out/<tag>.log (merged back by gpurun).
# Steps: tests | smoke | bench[:<extra bench.py flags>] | benchn:<ranks>[:<flags>] (torchrun) | mmsweep[:N[:mus]] | py:<script and args> | sh:<command>
tag=$1; shift
mkdir -p gpurun_out
exec > >(tee gpurun_out/${tag}.log) 2>&1
nvidia-smi -L
for step in "$@"; do
  kind=${step%%:*}; rest=""; [[ "$step" == *:* ]] && rest=${step#*:}
  echo "== [$tag] $step  ($(date +%T))"
  case $kind in
    tests)     timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 ;;

    smoke)     timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3 ;;
    bench)     timeout 1500 python bench.py $rest 2> gpurun_out/${tag}_bench.err | tee -a gpurun_out/${tag}_bench.jsonl | cut -c1-400; tail -5 gpurun_out/${tag}_bench.err ;;
    benchn)    ng=${rest%%:*}; fl=""; [[ "$rest" == *:* ]] && fl=${rest#*:}
               timeout 1800 python -m torch.distributed.run --nnodes=1 --nproc-per-node $ng --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $ng $fl 2> gpurun_out/${tag}_bench${ng}.err | tee -a gpurun_out/${tag}_bench.jsonl | cut -c1-400; grep -v "^W\|^$" gpurun_out/${tag}_bench${ng}.err | tail -6 ;;
    mmsweep)   timeout 1200 python scripts/mm_width_sweep.py ${rest//:/ } 2> gpurun_out/${tag}_mmsweep.err ;;
    py)        timeout 1500 python $rest ;;
    sh)        timeout 1500 bash -c "$rest" ;;
    *) echo "unknown step $step" ;;
  esac
done
echo "== [$tag] done ($(date +%T))"
