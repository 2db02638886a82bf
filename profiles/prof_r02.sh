# Round-2 profile captures (run on the GPU box through gpurun; every ncu pass only after the plain run exited 0).
set -e
CMD="python bench.py --batch 8 --steps 2 --warmup 3 --no-cpu --no-extras --no-compile"
$CMD > gpurun_out/prof_plain.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:bwd_tma_kernel --launch-skip 3 -c 1 -f -o gpurun_out/prof_bwd $CMD > gpurun_out/ncu_b.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fwd_kernel --launch-skip 3 -c 1 -f -o gpurun_out/prof_fwd $CMD > gpurun_out/ncu_f.log 2>&1
