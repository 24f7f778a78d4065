#!/bin/bash
# Compare prebuilt library variants (loqa_hub_b200/variants/libgpuaudio_<name>.so) on configs[3] (48 kHz stereo).
cp loqa_hub_b200/libgpuaudio.so /tmp/libgpuaudio_orig.so
for so in loqa_hub_b200/variants/libgpuaudio_*.so; do
  name=$(basename $so .so); name=${name#libgpuaudio_}
  cp $so loqa_hub_b200/libgpuaudio.so; touch loqa_hub_b200/libgpuaudio.so
  python bench.py --workload cfg4 --steps 20 --warmup 3 --no-e2e --no-cpu-baseline --no-sweep --no-tick 2>/dev/null | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('variant $name: value %.3fM  step %.3f ms  sha %.3f ms  decimator %.3f ms (%.3f of HBM)'%(d['value']/1e6,d['ms_per_step'],d['roofline']['ms_per_launch'],d['roofline_decimator']['ms_per_launch'],d['roofline_decimator']['frac']))"
done
cp /tmp/libgpuaudio_orig.so loqa_hub_b200/libgpuaudio.so
