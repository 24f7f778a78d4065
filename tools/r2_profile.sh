#!/bin/bash
# Round-2 profiling pass on the GPU box (one GPU): launch list + ncu --set full captures of the
# kernels DESIGN.md cites.  Each capture only after the same command has exited 0 without ncu.
# Usage: tools/r2_profile.sh <tag>      (writes gpurun_out/<tag>_*)
tag=${1:-r2a}
out=gpurun_out
B="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-sweep --no-tick"
EXTRA="sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed.sum,sm__cycles_active.avg,sm__cycles_elapsed.max"
set -x
$B > $out/${tag}_plain_cfg3.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_cfg3.csv $B > $out/${tag}_ncu_l.log 2>&1
ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:ingest_kernel -s 3 -c 1 -f -o $out/${tag}_ingest_fused $B > $out/${tag}_ncu_f.log 2>&1
ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:score_kernel -s 3 -c 1 -f -o $out/${tag}_score $B > $out/${tag}_ncu_s.log 2>&1
$B --workload cfg4 > $out/${tag}_plain_cfg4.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches_cfg4.csv $B --workload cfg4 > $out/${tag}_ncu_l4.log 2>&1
ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:ingest_kernel -s 3 -c 1 -f -o $out/${tag}_ingest_sha48 $B --workload cfg4 > $out/${tag}_ncu_f4.log 2>&1
ncu --set full --metrics $EXTRA --clock-control none --import-source on -k regex:resample48_kernel -s 3 -c 1 -f -o $out/${tag}_resample48 $B --workload cfg4 > $out/${tag}_ncu_r4.log 2>&1
ls -la $out/${tag}_*
