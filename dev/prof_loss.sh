set -e
CMD="python dev/lossbench.py"
$CMD > gpurun_out/loss_plain.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:loss_fwd_kernel --launch-skip 2 -c 1 -f -o gpurun_out/prof_lossfwd $CMD > gpurun_out/ncu_lossfwd.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:loss_bwd_kernel --launch-skip 2 -c 1 -f -o gpurun_out/prof_lossbwd $CMD > gpurun_out/ncu_lossbwd.log 2>&1
