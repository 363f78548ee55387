#!/bin/bash
# A/B of two builds of the library on one box: bash scripts/ab_e2e.sh <path/to/other/libtfs_gpu.so>
# (build the other revision with `git archive <rev> rust-tensorflow-serving2_b200/csrc include | tar -x -C _ab/<rev>` + make there;
#  _ab/ is git-ignored but travels with gpurun).  Prints value and e2e of `bench.py --no-extras` for each, twice.
other=$1
run() { python bench.py --no-extras --sustain-s 0 --steps 5 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$1', d['value'], d['e2e']['value'])
"; }
for rep in 1 2; do
  TFS_LIB_PATH=$other run other
  run this
done
