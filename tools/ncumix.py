"""ncumix.py <rep> [minexec]: opcode mix (weighted by execution count) of an ncu report's launch $LAUNCH, split by source line"""
import csv,sys,subprocess,os,collections,re
LAUNCH=int(os.environ.get('LAUNCH','0'))
src=subprocess.run(['ncu','-i',sys.argv[1],'--page','source','--csv'],capture_output=True,text=True).stdout
rows=list(csv.reader(src.splitlines()))
starts=[i for i,r in enumerate(rows) if r and r[0]=='Address']
a=starts[LAUNCH]; b=(starts[LAUNCH+1]-1) if LAUNCH+1<len(starts) else len(rows)
hdr=rows[a]; data=[r for r in rows[a+1:b] if len(r)>5]
iex=hdr.index('Instructions Executed'); isrc=hdr.index('Source')
ALU={'FMNMX','FMNMX3','FSETP','FSET','SEL','FSEL','LOP3','SHF','PRMT','IADD3','IADD','ISETP','LEA','VOTE','POPC','FLO','BREV','IMNMX','VIMNMX','VIMNMX3','ICMP','P2R','R2P','PLOP3','IABS','SGXT','BMSK','MOV','CS2R','FCHK','VABSDIFF','IADD32I'}
FMA={'FADD','FMUL','FFMA','IMAD','HFMA2','HADD2','HMUL2'}
tot=collections.Counter(); byop=collections.Counter()
for r in data:
    e=int(r[iex]); s=r[isrc].strip(); t=s.split(); 
    if not t: continue
    op=(t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    pipe='alu' if op in ALU else ('fma' if op in FMA else 'other')
    tot[pipe]+=e; byop[(pipe,op)]+=e
T=sum(tot.values())
print({k:'%.1f%%'%(100*v/T) for k,v in tot.items()}, 'total',T)
for (p,op),v in byop.most_common(40): print('%-6s %-10s %6.2f%%  %.2f/elem'%(p,op,100*v/T, v*32/2**28))
