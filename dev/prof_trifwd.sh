set -e
CMD="python dev/tribench.py"
ncu --set full --clock-control none --import-source on -k regex:trispace_fwd_kernel --launch-skip 2 -c 1 -f -o gpurun_out/prof_trifwd $CMD > gpurun_out/ncu_trifwd.log 2>&1
