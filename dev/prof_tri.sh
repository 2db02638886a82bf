set -e
CMD="python dev/tribench.py"
$CMD > gpurun_out/tri_plain.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:trispace_bwd_kernel --launch-skip 2 -c 1 -f -o gpurun_out/prof_tribwd $CMD > gpurun_out/ncu_tri.log 2>&1
