# Round-2 TriSpace captures (gpurun; ncu only after the plain run exited 0): the inference forward (no saved sigmoids,
# 126 coefficients, the rolled variant) and the image-gradient kernel, 4K x 8.
set -e
export TB_4K=1
TB_SAVE=0 python dev/tribench.py > gpurun_out/tri_plain_nosave.log 2>&1
TB_SAVE=1 python dev/tribench.py > gpurun_out/tri_plain_save.log 2>&1
TB_SAVE=0 ncu --set full --clock-control none --import-source on -k regex:trispace_fwd_kernel --launch-skip 2 -c 1 -f -o gpurun_out/prof_trifwd_nosave python dev/tribench.py > gpurun_out/ncu_trifwd.log 2>&1
TB_SAVE=1 ncu --set full --clock-control none --import-source on -k regex:trispace_gimg_kernel --launch-skip 2 -c 1 -f -o gpurun_out/prof_trigimg python dev/tribench.py > gpurun_out/ncu_trigimg.log 2>&1
