# Dev: times TriSpace builds with dev/tribench.py (TRI_VARIANTS: names under dev/variants/tri_*.so; "product" = the
# shipped library), then runs the GPU tests.
set -x
for v in ${TRI_VARIANTS:-product}; do
  lib=dev/variants/tri_$v.so; [ "$v" = product ] && lib=curl_b200/libcurl_b200.so
  for s in 1 0; do echo "== $v save=$s"; TB_4K=1 TB_SAVE=$s timeout 300 python dev/tribench.py $lib 2>&1 | tail -2; done
done
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
