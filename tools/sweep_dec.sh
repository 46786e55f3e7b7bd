#!/bin/bash
# build variants of the decoder on the GPU box and time them: sweep_dec.sh "NT MINB" ...
for v in "$@"; do set -- $v
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared -fmad=false -DDEC_NT=$1 -DDEC_MINB=$2 -Xptxas -v -o data_compression_b200/libdcb200.so data_compression_b200/csrc/dcb_abi.cu data_compression_b200/csrc/dcb_legacy.cu 2>&1 | grep -A2 "decode_kernelILi7" | grep "Used" | head -1
  echo "== DEC_NT=$1 MINB=$2"
  for f in smooth noisy; do for ct in 5 7; do
    python bench.py --ct $ct --field $f --steps 10 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']
print('$f ct=$ct dec=%.3f ms ok=%s' % (c['decompress_ms'], c['decode_status_ok']))"
  done; done
done
