#!/bin/bash
# build variants of the warp kernel on the GPU box and time them: sweep.sh "NT BLOCKS BLOCKS_CT7" ...
for v in "$@"; do set -- $v; 
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared -fmad=false -DDCB_WNT=$1 -DDCB_WARP_BLOCKS=$2 -DDCB_WARP_BLOCKS_CT7=$3 -o data_compression_b200/libdcb200.so data_compression_b200/csrc/dcb_abi.cu data_compression_b200/csrc/dcb_legacy.cu 2>&1 | grep -i error
  echo "== NT=$1 blocks=$2 blocks_ct7=$3"
  for f in smooth noisy; do for ct in 5 7; do
    DCB_ENCODER=warp python bench.py --ct $ct --field $f --steps 10 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']
print('$f ct=$ct enc=%.3f ms' % (c['compress_ms']))"
  done; done
done
