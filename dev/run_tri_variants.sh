# Dev: times TriSpace experiment builds (dev/variants/tri_*.so) with dev/tribench.py, then runs the GPU tests.
set -x
for v in ${TRI_VARIANTS:-head hybrid hybrid_v4}; do
  for s in 1 0; do echo "== $v save=$s"; TB_4K=1 TB_SAVE=$s timeout 300 python dev/tribench.py dev/variants/tri_$v.so 2>&1 | tail -2; done
done
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
