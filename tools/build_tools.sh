#!/bin/bash
# Builds the native harness tools against libgpuaudio.so (C ABI only).  Outputs tools/*.bin (git-ignored).
set -e
cd "$(dirname "$0")/.."
g++ -O2 -std=c++17 -Wall tools/hub_realtime.cpp -o tools/hub_realtime.bin \
    -Lloqa_hub_b200 -lgpuaudio -Wl,-rpath,'$ORIGIN/../loqa_hub_b200' -lpthread -ldl
