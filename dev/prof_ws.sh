set -e
export KB_B=8 KB_REPS=2
CMD="python dev/kbench.py dev/libs/ws_s24t352.so"
$CMD > gpurun_out/prof_ws_plain.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:bwd_tma --launch-skip 3 -c 1 -f -o gpurun_out/prof_ws $CMD > gpurun_out/ncu_ws.log 2>&1
