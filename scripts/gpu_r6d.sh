#!/bin/bash
python scripts/gpu_pipe_check.py 2>&1 | tail -3
