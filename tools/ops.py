import re,sys,collections
f=sys.argv[1]
lo=int(sys.argv[2],16) if len(sys.argv)>2 else 0
hi=int(sys.argv[3],16) if len(sys.argv)>3 else 1<<30
c=collections.Counter()
n=0
for line in open(f):
    m=re.match(r'\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);',line)
    if not m: continue
    a=int(m.group(1),16)
    if a<lo or a>=hi: continue
    ins=m.group(2).split()
    op=ins[0]
    if op.startswith('@'): op=ins[1]
    op=op.split('.')[0]
    c[op]+=1; n+=1
print(n, dict(c.most_common()))
