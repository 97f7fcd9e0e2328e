#!/bin/bash
# ncu evidence of round 2: launch list of the bench command, full captures of k_prune (c4 at the benchmark's launch
# size, c3), k_pmatrix, k_hmm and k_newick_cta. Every capture only after the same command ran clean without ncu.
mkdir -p gpurun_out
tag=${1:-r4e}
B="python bench.py --no-cpu-baseline --no-extras"
$B --steps 2 > gpurun_out/plain_$tag.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_$tag.csv $B --steps 2 > gpurun_out/ncu_launch_$tag.log 2>&1
C4="$B --steps 1 --warmup 3 --no-e2e"
$C4 > gpurun_out/plain2_$tag.log 2>&1 && {
ncu --set full --clock-control none --import-source on -k regex:k_prune -s 3 -c 1 -o gpurun_out/prof_prune_c4_$tag -f $C4 > gpurun_out/ncu_prune_c4_$tag.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_pmatrix -s 3 -c 1 -o gpurun_out/prof_pmatrix_c4_$tag -f $C4 > gpurun_out/ncu_pmatrix_$tag.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_hmm -s 3 -c 1 -o gpurun_out/prof_hmm_c4_$tag -f $C4 > gpurun_out/ncu_hmm_$tag.log 2>&1
}
C3="$B --config c3 --steps 1 --warmup 3 --no-e2e"
$C3 > gpurun_out/plain3_$tag.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_prune -s 3 -c 1 -o gpurun_out/prof_prune_c3_$tag -f $C3 > gpurun_out/ncu_prune_c3_$tag.log 2>&1
python scripts/gpu_newick_probe.py c4 4096 > gpurun_out/newick_probe_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_newick_cta -s 1 -c 1 -o gpurun_out/prof_newick_$tag -f python scripts/gpu_newick_probe.py c4 4096 > gpurun_out/ncu_newick_$tag.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/ncu_prune_c4_$tag.log
