"""ncu_by_line.py <rep> <cubin> <mangled kernel>: executed warp instructions and stall samples per SOURCE line
(SASS order of the ncu source page joined with nvdisasm -g line markers; inlined code is attributed to the innermost line)."""
import csv, os, re, subprocess, sys, collections
rep, cubin, kern = sys.argv[1], sys.argv[2], sys.argv[3]
LAUNCH = int(os.environ.get('LAUNCH', '0'))
dis = subprocess.run(['nvdisasm', '-g', '-c', cubin], capture_output=True, text=True).stdout.splitlines()
lines = []; cur = None; on = False
for l in dis:
    if l.startswith('.text.'):
        on = (l.strip() == '.text.%s:' % kern); continue
    if not on: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', l): lines.append(cur)
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
a = starts[LAUNCH]; b = (starts[LAUNCH + 1] - 1) if LAUNCH + 1 < len(starts) else len(rows)
hdr = rows[a]; data = [r for r in rows[a + 1:b] if len(r) > 5]
iex = hdr.index('Instructions Executed'); isamp = hdr.index('# Samples')
print('sass instrs: nvdisasm %d, ncu %d' % (len(lines), len(data)))
agg = collections.defaultdict(lambda: [0, 0, 0])
for ln, r in zip(lines, data):
    k = ln or ('?', 0)
    agg[k][0] += 1; agg[k][1] += int(r[iex]); agg[k][2] += int(r[isamp])
tot = sum(v[1] for v in agg.values()); ts = sum(v[2] for v in agg.values())
srccache = {}
def text(f, n):
    p = os.path.join('/root/repo/data_compression_b200/csrc', f)
    if p not in srccache:
        try: srccache[p] = open(p).read().splitlines()
        except Exception: srccache[p] = []
    L = srccache[p]
    return L[n - 1].strip()[:70] if 0 < n <= len(L) else ''
top = int(sys.argv[4]) if len(sys.argv) > 4 else 45
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print('%-16s %5d  sass=%4d  inst=%5.2f%%  samp=%5.2f%%  %s' % (k[0], k[1], v[0], 100 * v[1] / tot, 100 * v[2] / ts, text(*k)))
