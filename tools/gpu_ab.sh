#!/bin/bash
# A/B: old library vs new on the same box
mkdir -p gpurun_out
cp spmm_denseblock_b200/libspmmdb.so /tmp/new.so
for which in new old new old; do
  if [ $which = old ]; then cp tools/tmp_old/libspmmdb_old.so spmm_denseblock_b200/libspmmdb.so; else cp /tmp/new.so spmm_denseblock_b200/libspmmdb.so; fi
  echo "== $which"; timeout 600 python tools/time_kernels.py --cases c3d --maths 1 2>/dev/null | grep bsrmm | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['bs'], 'ms_med %.4f ms_min %.4f dense %.1f TF frac %.3f'%(d['ms_med'],d['ms_min'],d['dense_tflops'],d['frac_hbm']))"
done
cp /tmp/new.so spmm_denseblock_b200/libspmmdb.so
