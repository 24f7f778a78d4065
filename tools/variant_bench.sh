#!/bin/bash
# Compare prebuilt kernel variants (loqa_hub_b200/variants/libgpuaudio_<name>.so) on the GPU box:
# each is copied over libgpuaudio.so in the box's scratch copy of the repo and benched.
cp loqa_hub_b200/libgpuaudio.so /tmp/libgpuaudio_orig.so
for so in loqa_hub_b200/variants/libgpuaudio_*.so; do
  name=$(basename $so .so); name=${name#libgpuaudio_}
  cp $so loqa_hub_b200/libgpuaudio.so; touch loqa_hub_b200/libgpuaudio.so
  for m in fused split; do
    python bench.py --mode $m --steps 20 --no-e2e --no-cpu-baseline --no-sweep --no-tick $VARIANT_BENCH_ARGS 2>&1 | tail -1 | \
      python -c "import sys,json; d=json.loads(sys.stdin.read()); print('variant $name mode $m: value %.3fM  step %.3f ms  ingest %.3f ms  frac %.3f'%(d['value']/1e6,d['ms_per_step'],d['roofline']['ms_per_launch'],d['roofline']['frac']))"
  done
done
cp /tmp/libgpuaudio_orig.so loqa_hub_b200/libgpuaudio.so
