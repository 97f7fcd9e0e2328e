#!/bin/bash
python -m pytest tests/test_gpu_newick.py -x -q -m gpu -k "2 or lenient or chunks or pinned" 2>&1 | tail -2
python scripts/gpu_newick_probe.py c4 4096 2>&1 | grep '"kernel": 2' | tail -2 | cut -c1-110
python scripts/gpu_newick_probe.py c4 148 2>&1 | grep '"kernel": 2' | tail -1 | cut -c1-110
python scripts/gpu_newick_probe.py c3 16384 2>&1 | grep '"kernel": 2' | tail -1 | cut -c1-110
