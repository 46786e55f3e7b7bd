"""profile_summary.py <rep> <out_prefix>: text summaries (one per captured launch) of an `ncu --set full` report:
headline metrics, dram bytes, stall mix and the SASS segments that carry the time.  The summaries are what
profiles/ keeps; the .ncu-rep files stay in gpurun_out/ (scratch)."""
import csv, os, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
        'launch__shared_mem_per_block_static', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'sm__cycles_elapsed.avg.per_second']
for li, r in enumerate(rows[2:]):
    name = r[hdr.index('Kernel Name')]
    short = 'encode' if 'encode' in name else ('decode' if 'decode' in name else 'k%d' % li)
    path = '%s_%s.txt' % (out, short)
    with open(path, 'w') as f:
        f.write('# %s\n# from %s (ncu --set full --clock-control none --import-source on), launch %d of the report\n' % (name, os.path.basename(rep), li))
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                f.write('%-70s %s %s\n' % (k, r[i], units[i]))
        env = dict(os.environ, LAUNCH=str(li * int(os.environ.get('SRC_SECTIONS_PER_LAUNCH', '1'))))   # some reports list every launch twice on the source page
        seg = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), 'ncuseg.py'), rep], capture_output=True, text=True, env=env).stdout
        keep = seg[seg.index('total inst'):] if 'total inst' in seg else seg
        f.write('\n# stall mix (% of warp samples) and SASS segments by execution count: first-last index, #instr, times executed,\n# share of executed instructions, share of samples, first instruction, top stalls\n')
        f.write(keep)
        ex = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), 'ncuexec.py'), rep], capture_output=True, text=True, env=env).stdout
        f.write('\n# executed instructions grouped by how often a path runs\n' + ex)
    print(path)
