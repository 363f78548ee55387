#!/bin/bash
# GPU box: the suite, then the plan kernel's phase cycles on the flagship batch and on the many-short-strings batch.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -1
TFS_PLAN_TIMING=1 python bench.py --no-extras --steps 3 --warmup 2 > gpurun_out/c2_cycles.log 2>&1
echo "C2:"; grep "k_plan_small cycles" gpurun_out/c2_cycles.log | tail -2
TFS_PLAN_TIMING=1 python scripts/tokens_decode.py > gpurun_out/tok_cycles.log 2>&1
echo "tokens:"; grep -E "cycles|decode_into" gpurun_out/tok_cycles.log | tail -2
python bench.py --no-extras > gpurun_out/b.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/b.json')); print(d['value'], d['ms_per_step'], d['roofline']['step_breakdown_ms'])"
python scripts/c5_decode.py 2>&1 | grep -i "c5" | tail -2
python scripts/c1_batch.py | tail -1
