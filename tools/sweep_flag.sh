#!/bin/bash
# A/B of a compile-time flag on the GPU box: sweep_flag.sh "" "-DFLAG" ...
for v in "$@"; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared -fmad=false $v -o data_compression_b200/libdcb200.so data_compression_b200/csrc/dcb_abi.cu data_compression_b200/csrc/dcb_legacy.cu 2>&1 | grep -i error
  echo "== flags: '$v'"
  for f in smooth noisy; do for ct in 5 7; do
    python bench.py --ct $ct --field $f --steps 10 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']; r=d['roofline']
print('$f ct=$ct enc=%.3f ms (kernel %.3f + compact %.3f) dec=%.3f' % (c['compress_ms'], r['encode_kernel_ms'], r['compact_kernel_ms'], c['decompress_ms']))"
  done; done
done
