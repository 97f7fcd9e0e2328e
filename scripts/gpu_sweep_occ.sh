#!/bin/bash
# k_prune occupancy sweep: columns per lane (4 / 3) x register cap (launch bounds 12 / 14 CTAs of one warp)
mkdir -p gpurun_out
tag=${1:-r4f}
out=gpurun_out/sweep_occ_$tag.jsonl; : > $out
run() { # label, extra bench args
  python bench.py --no-cpu-baseline --no-extras --no-e2e --steps 3 $2 > /tmp/line.json 2>> gpurun_out/sweep_occ_$tag.err
  python - "$1" <<PY >> $out
import json, sys
d = json.loads(open("/tmp/line.json").read().splitlines()[-1])
print(json.dumps({"variant": sys.argv[1], "value": d["value"], "k_prune_ms": d["stage_ms_per_step"]["k_prune"], "fp64_frac": d["roofline"]["frac"],
                  "finite_fraction": d["finite_fraction"]}))
PY
  tail -1 $out
}
run "ns4_bounds12 (HEAD)" ""
run "ns3_bounds12" "--opt sites_per_thread=3"
for m in 14 16; do
  touch linearham_b200/csrc/kernels.cu
  make -C linearham_b200/csrc NVCC_EXTRA="-DLH_PRUNE_MIN_CTAS=$m -DLH_PRUNE_MIN_CTAS3=$m -DLH_PRUNE_MIN_CTAS2=16 -DLH_PRUNE_MIN_CTAS1=24" > /dev/null 2>&1 || echo BUILD FAILED
  grep -A1 "k_pruneILi[34]" build/obj/kernels.ptxas.log | grep -i "spill" | head -2
  run "ns4_bounds$m" ""
  run "ns3_bounds$m" "--opt sites_per_thread=3"
done
python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
