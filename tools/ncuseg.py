import csv,sys,collections,subprocess,os
SEL=[]
LAUNCH=int(os.environ.get('LAUNCH','0'))
def section(rows):
    starts=[i for i,r in enumerate(rows) if r and r[0]=='Address']
    a=starts[LAUNCH]; b=(starts[LAUNCH+1]-1) if LAUNCH+1<len(starts) else len(rows)
    return [rows[a-1] if a>0 else []]+rows[a:b]
rep=sys.argv[1]
raw=subprocess.run(['ncu','-i',rep,'--page','raw','--csv']+SEL,capture_output=True,text=True).stdout
rows=list(csv.reader(raw.splitlines()))
hdr=rows[0]; units=rows[1]
want=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','launch__registers_per_thread','launch__grid_size','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__inst_executed_pipe_alu.sum','smsp__inst_executed_pipe_fma.sum','smsp__inst_executed_pipe_lsu.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active','sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active','sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active','sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active']
for r in rows[2:]:
    print(r[hdr.index('Kernel Name')])
    for k in want:
        if k in hdr: i=hdr.index(k); print('  ',k, r[i], units[i])
src=subprocess.run(['ncu','-i',rep,'--page','source','--csv']+SEL,capture_output=True,text=True).stdout
rows=section(list(csv.reader(src.splitlines())))
hdr=rows[1]
isrc=hdr.index('Source'); iex=hdr.index('Instructions Executed'); isamp=hdr.index('# Samples')
stall=[h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
data=[r for r in rows[2:] if len(r)>iex]
tot=sum(int(r[iex]) for r in data); ts=sum(int(r[isamp]) for r in data)
print('total inst',tot,'samples',ts)
agg=collections.Counter()
for r in data:
    for h in stall: agg[h]+=int(r[hdr.index(h)] or 0)
print({k:round(100*v/ts,1) for k,v in agg.most_common(8)})
minseg=int(sys.argv[2]) if len(sys.argv)>2 else 0
a=0
for i in range(1,len(data)+1):
    if i==len(data) or abs(int(data[i][iex])-int(data[a][iex]))>0.02*max(int(data[a][iex]),1):
        n=sum(int(r[iex]) for r in data[a:i]); s=sum(int(r[isamp]) for r in data[a:i])
        if 100*n/tot>=0.5 or 100*s/ts>=1.0:
            st=collections.Counter()
            for r in data[a:i]:
                for h in stall: st[h]+=int(r[hdr.index(h)] or 0)
            top=', '.join('%s %.1f'%(k[6:],100*v/ts) for k,v in st.most_common(3) if v)
            print('%5d-%5d n=%4d exec=%8d inst%%=%5.1f samp%%=%5.1f  %-40s | %s'%(a,i,i-a,int(data[a][iex]),100*n/tot,100*s/ts,data[a][isrc].strip()[:40],top))
        a=i
