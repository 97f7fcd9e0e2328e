#!/bin/bash
# one full ncu capture of the device Newick parser named in $1 (regex), C4 strings, 4096 trees
mkdir -p gpurun_out
k=${1:-k_newick_warp}; tag=${2:-r5}
python scripts/gpu_newick_probe.py c4 4096 > gpurun_out/newick_probe_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o gpurun_out/prof_${k}_$tag -f python scripts/gpu_newick_probe.py c4 4096 > gpurun_out/ncu_${k}_$tag.log 2>&1
grep '"kernel": [23]' gpurun_out/newick_probe_plain.log | tail -4; ls -la gpurun_out/prof_${k}_$tag.ncu-rep
