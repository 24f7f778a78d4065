#!/bin/bash
# Kernel comparison on the GPU box: staging {tmap, bulk} x mode {fused, split}.
# (The GA_INGEST_VARIANT slice/stage sweeps recorded in profiles/r1_sweep_history.log were run with a
# temporary variant switch in ga_k_ingest that has since been removed.)
for st in tmap bulk; do
  for m in fused split; do
    python bench.py --staging $st --mode $m --steps 20 --no-e2e --no-cpu-baseline --no-sweep --no-tick 2>&1 | tail -1 | \
      python -c "import sys,json; d=json.loads(sys.stdin.read()); print('staging $st mode $m: value %.3fM audio-s/s  step %.3f ms  ingest %.3f ms  frac %.3f'%(d['value']/1e6,d['ms_per_step'],d['roofline']['ms_per_launch'],d['roofline']['frac']))"
  done
done
