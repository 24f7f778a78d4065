#!/bin/bash
# Kernel tuning sweep (run on the GPU box): ingest variants x {fused, split}.
for v in ${VARIANTS:-0 1 2 3 4 5 6}; do
  for m in fused split; do
    GA_INGEST_VARIANT=$v python bench.py --mode $m --steps 20 --no-e2e --no-cpu-baseline 2>&1 | tail -1 | \
      python -c "import sys,json; d=json.loads(sys.stdin.read()); print('variant $v mode $m: value %.3fM audio-s/s  step %.3f ms  ingest %.3f ms  frac %.3f'%(d['value']/1e6,d['ms_per_step'],d['roofline']['ms_per_launch'],d['roofline']['frac']))"
  done
done
