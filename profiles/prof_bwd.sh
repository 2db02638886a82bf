set -e
CMD="python bench.py --batch 8 --steps 2 --warmup 3 --no-cpu --no-extras"
$CMD > gpurun_out/prof_plain.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:bwd_ --launch-skip 3 -c 1 -f -o gpurun_out/prof_bwd $CMD > gpurun_out/ncu_b.log 2>&1
