#!/bin/bash
# the no-allocation entry point's test alone in a fresh process, then the whole -m gpu suite (it runs last there)
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -15 > gpurun_out/gputests_noalloc_alone.log; cat gpurun_out/gputests_noalloc_alone.log
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/gputests_${1:-rX}.log; cat gpurun_out/gputests_${1:-rX}.log
