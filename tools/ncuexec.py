"""ncuexec.py <rep>: instruction totals grouped by execution count class (which path runs how often)"""
import csv,sys,subprocess,os,collections
LAUNCH=int(os.environ.get('LAUNCH','0'))
src=subprocess.run(['ncu','-i',sys.argv[1],'--page','source','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
starts=[i for i,r in enumerate(rows) if r and r[0]=='Address']
a=starts[LAUNCH]; b=(starts[LAUNCH+1]-1) if LAUNCH+1<len(starts) else len(rows)
hdr=rows[a]; data=[r for r in rows[a+1:b] if len(r)>5]
iex=hdr.index('Instructions Executed'); isamp=hdr.index('# Samples')
tot=sum(int(r[iex]) for r in data); ts=sum(int(r[isamp]) for r in data)
cls=collections.OrderedDict()
for r in data:
    e=int(r[iex]); k=None
    for c in cls:
        if abs(c-e)<=0.03*max(c,1): k=c; break
    if k is None: k=e; cls[k]=[0,0,0]
    cls[k][0]+=1; cls[k][1]+=e; cls[k][2]+=int(r[isamp])
print('total',tot,'samples',ts)
for k,(n,e,s) in sorted(cls.items(),key=lambda kv:-kv[1][1])[:25]:
    print('exec~%9d  static=%5d  dyn=%11d (%5.1f%%)  samples=%5.1f%%'%(k,n,e,100*e/tot,100*s/ts))
