#!/bin/bash
# ncu evidence for the round's last code: launch list of the bench command, k_sample capture
mkdir -p gpurun_out
B="python bench.py --no-cpu-baseline --no-extras --steps 2"
$B > gpurun_out/plain_r6.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_r6.csv $B > gpurun_out/ncu_launch_r6.log 2>&1
python scripts/gpu_pipe_probe.py 4096 > gpurun_out/pipe_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_sample -s 8 -c 1 -o gpurun_out/prof_k_sample_r6 -f python scripts/gpu_pipe_probe.py 4096 > gpurun_out/ncu_sample_r6.log 2>&1
ls -la gpurun_out/launches_r6.csv gpurun_out/prof_k_sample_r6.ncu-rep; tail -1 gpurun_out/plain_r6.log | cut -c1-200
