#!/bin/bash
# the whole -m gpu suite (what the driver runs at round end)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/gputests_${1:-rX}.log; cat gpurun_out/gputests_${1:-rX}.log
