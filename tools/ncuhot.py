"""ncuhot.py <rep> <first> <last> [minpct]: per-instruction samples/stalls of a SASS range of an ncu report"""
import csv,sys,subprocess,os
SEL=[]
LAUNCH=int(os.environ.get('LAUNCH','0'))
def section(rows):
    starts=[i for i,r in enumerate(rows) if r and r[0]=='Address']
    a=starts[LAUNCH]; b=(starts[LAUNCH+1]-1) if LAUNCH+1<len(starts) else len(rows)
    return [rows[a-1] if a>0 else []]+rows[a:b]
rep,a,b=sys.argv[1],int(sys.argv[2]),int(sys.argv[3]); minp=float(sys.argv[4]) if len(sys.argv)>4 else 0.0
src=subprocess.run(['ncu','-i',rep,'--page','source','--csv']+SEL,capture_output=True,text=True).stdout
rows=section(list(csv.reader(src.splitlines()))); hdr=rows[1]
isrc=hdr.index('Source'); iex=hdr.index('Instructions Executed'); isamp=hdr.index('# Samples')
stall=[h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
data=[r for r in rows[2:] if len(r)>iex]
ts=sum(int(r[isamp]) for r in data)
for i in range(a,min(b,len(data))):
    r=data[i]; s=int(r[isamp])
    if 100*s/ts<minp: continue
    st=sorted(((int(r[hdr.index(h)] or 0),h[6:]) for h in stall),reverse=True)[:2]
    print('%5d %6.2f%%  %-70s %s'%(i,100*s/ts,r[isrc].strip()[:70],' '.join('%s=%d'%(k,v) for v,k in st if v)))
