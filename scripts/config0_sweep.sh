#!/bin/bash
# BASELINE.json configs[0]: N=20k cylinder, 1/(0.1+|x-y|), eps 1e-3, eta 10; leaf size 10 (the reference's
# default maximal_leaf_size), 64, 128; symmetric ('S','L' as example/use_hmatrix.py) and non-symmetric.  One bench line per
# case and arm into gpurun_out/r03_config0_sweep.jsonl
out=gpurun_out/r03_config0_sweep.jsonl
: > $out
for sym in S N; do
  for leaf in 10 64 128; do
    python bench.py --config 0 --leaf $leaf --symmetry $sym --steps 50 --warmup 5 --sustained-steps 0 2>> gpurun_out/r03_config0_sweep.err | tail -1 >> $out
    python bench.py --impl reference --config 0 --leaf $leaf --symmetry $sym --steps 20 --warmup 3 2>> gpurun_out/r03_config0_sweep.err | tail -1 >> $out
  done
done
python - <<'PY'
import json
print("| sym | leaf | arm | matvecs/s | ms/step | e2e matvecs/s | kernel GB/s (frac) | parity |")
for l in open("gpurun_out/r03_config0_sweep.jsonl"):
    try: d = json.loads(l)
    except Exception: print("bad line:", l[:200]); continue
    w = d["config"]["workload"]; sym = "S" if "symmetric" in w else "N"; leaf = w.split("leaf ")[1].split(",")[0]
    r = d.get("roofline") or {}
    print(f"| {sym} | {leaf} | {d.get('impl','b200')} | {d['value']} | {d['ms_per_step']} | {(d.get('e2e') or {}).get('value')} | {r.get('achieved')} ({r.get('frac')}) | {d.get('parity_rel_err_vs_reference')} |")
PY
