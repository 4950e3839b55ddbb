#!/bin/bash
# run bench.py against every keyframe-based-slam_b200/libkfs_var_*.so (kernel-variant experiments) and print stage times
for so in keyframe-based-slam_b200/libkfs_var_*.so; do
  KFS_LIB=$PWD/$so python bench.py --no-cpu-baseline --steps 8 --e2e-steps 2 > /tmp/v.json 2>/tmp/v.err || { echo "$so FAILED"; tail -3 /tmp/v.err; continue; }
  python - "$so" <<'P'
import json,sys
d=json.load(open('/tmp/v.json'))
print(sys.argv[1].split('/')[-1], 'fps', d['value'], 'e2e', d['e2e']['value'], ' '.join('%s=%.4f'%(k['stage'],k['ms_per_batch']) for k in d['kernels']), 'knn=%.4f'%d['roofline_popc']['ms_per_batch'])
P
done
