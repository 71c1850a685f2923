#!/bin/bash
# the same command in N fresh processes: which timeline does the slow mode have?
Q="--steps 300 --warmup 10 --no-cpu-baseline --no-e2e --no-extras --no-strong --windows 2"
for i in $(seq 1 ${1:-8}); do
  env $2 timeout 200 python bench.py $Q 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
ig = d['roofline']['in_graph']
print(round(d['ms_per_step']*1e3,1), 'us | traced span', ig['span_us'], '|', ' '.join('%s %.0f-%.0f' % (l['kernel'].replace('_kernel','').replace('loss_','').replace('assign_','a_').replace('nms_','n_'), l['start_us'], l['end_us']) for l in ig['launches']))
"
done
