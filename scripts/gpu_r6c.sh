#!/bin/bash
python -m pytest tests/test_gpu_sample_paths.py -x -q -m gpu 2>&1 | tail -15
