#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_catch_compat.py -x -q -m gpu 2>&1 | tail -15
d=$(mktemp -d); python - <<PY
import numpy as np, os
z=np.load('tests/golden/fixture_catch_files.npz')
for n,t in zip(z['names'],z['texts']):
    p=os.path.join('$d','data',str(n)); os.makedirs(os.path.dirname(p),exist_ok=True); open(p,'w').write(str(t))
PY
(cd $d && /root/repo/build/compat/catch_phylohmm 2>&1 | tail -40) | tee gpurun_out/catch_phylohmm_r5d.txt | tail -40
