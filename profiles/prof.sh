# Profiling recipe of this round (run under gpurun, one GPU).  Plain run first, then the launch
# list, then one full capture per kernel.  Numbers printed under ncu are never bench values.
set -e
CMD="python bench.py --batch 8 --steps 2 --warmup 3 --no-cpu --no-extras"
$CMD > gpurun_out/prof_plain.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fwd_kernel --launch-skip 3 -c 1 -f -o gpurun_out/prof_fwd $CMD > gpurun_out/ncu_f.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:bwd_ --launch-skip 3 -c 1 -f -o gpurun_out/prof_bwd $CMD > gpurun_out/ncu_b.log 2>&1
