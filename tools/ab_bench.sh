#!/bin/bash
# same-box A/B of library variants: tools/ab_bench.sh <rounds> <variant.so>...   (prints ms/step and the per-kernel split)
rounds=$1; shift
for r in $(seq 1 $rounds); do for v in "$@"; do
  TSQR_B200_LIB=$PWD/tsqr_gpu_b200/variants/$v timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); o=dict(d['roofline']['other_kernels_ms']); k=d['roofline']['kernel']; key='pc_pass1' if 'read A once' in k else 'pc_pass2'; o[key]=d['roofline']['ms_per_launch']
print('$v', 'step', round(d['ms_per_step'],3), 'pass1', round(o.get('pc_pass1',0),3), 'pass2', round(o.get('pc_pass2',0),3), 'sample', round(o.get('pc_sample',0),3), 'finish', round(o.get('pc_finish',0),3))"
done; done
