#!/bin/bash
# usage: cls.sh [extra nvcc flags]  -> op mix of the classification loop of the CT5 encoder
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared -fmad=false "$@" -o /tmp/t2.so /root/repo/data_compression_b200/csrc/dcb_abi.cu -Xptxas -v 2>&1 | grep -A2 "encode_kernelILi5" | grep "Used\|spill"
cuobjdump -sass /tmp/t2.so | awk '/Function : /{p=($3 ~ /encode_kernelILi5/)} p{print}' > /tmp/sass/e5.sass
python - <<'PY'
import re,collections
lines=[l for l in open('/tmp/sass/e5.sass') if re.match(r'\s*/\*[0-9a-f]{4,5}\*/',l)]
ops=[]
for l in lines:
    m=re.match(r'\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);',l); t=m.group(2).split(); op=t[1] if t[0].startswith('@') else t[0]; ops.append((op.split('.')[0],m.group(2)))
first=next(i for i,(o,s) in enumerate(ops) if o=='FMNMX3' and '|' in s)
last=next(i for i,(o,s) in enumerate(ops) if o=='VOTE' and i>first)
c=collections.Counter(o for o,s in ops[first-8:last])
print('total static', len(ops), 'classification', last-first+8, dict(c.most_common()))
PY
