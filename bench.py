#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 float32 bit-wise codec.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--ct 7] [--field smooth]
                  [--log2n 28] [--eb 0.01] [--chunk 4096]

Metric (BASELINE.json): compress + decompress INPUT GB/s.  A step = one pass of the hot path over one batch:
compress the array (min-subtract, [medium/type], encode, look-back, pack) and decompress it again (+min).
N=1 workload = BASELINE.json configs[1]: CT=7 bitmask bit-wise, 2^28-element synthetic smooth float32 field.
N>1: each rank holds its own 2^28-element contiguous slice of the global field (chunk ranges shard with no
data-path collective) -> weak scaling; value = all ranks' input bytes / max-over-ranks time.

One JSON line on rank 0.  `value` is device-resident throughput (inputs in HBM when the clock starts);
`e2e` goes through the host-pointer C ABI (dcb_compress_host / dcb_decompress_host) with pinned host buffers,
H2D and D2H inside the timed region.  `roofline` is the dominant kernel's algorithmic bytes / CUDA-event time
against the measured HBM copy peak.  `cpu_baseline` / `--impl reference` time the reference's own C code
(oracle/_ref, compiled from /root/reference by oracle/build_ref.sh) on the box's host cores on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# Rank 0 prints exactly ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner on fd 1), so
# fd 1 is pointed at stderr for the life of the process and the JSON line goes to a private copy of the original stdout.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

ALL_CPUS = os.sched_getaffinity(0)           # before run_ours narrows the affinity to the GPU's NUMA node
METRIC = "compress+decompress input GB/s"
UNIT = "GB/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=("ours", "reference"))
    ap.add_argument("--ct", type=int, default=7)
    ap.add_argument("--field", default="smooth", choices=("smooth", "noisy", "constant"))
    ap.add_argument("--log2n", type=int, default=28, help="elements per GPU = 2^log2n")
    ap.add_argument("--eb", type=float, default=0.01)
    ap.add_argument("--chunk", type=int, default=4096)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-threads", type=int, default=4, help="host threads making the blocking host-pointer calls")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--sweep", default="full", choices=("full", "quick", "none"),
                    help="extra workloads measured after the headline and reported under \"sweep\" (BASELINE.json configs 1, 3, 5)")
    ap.add_argument("--sweep-steps", type=int, default=5)
    ap.add_argument("--no-nccl", action="store_true", help="skip the pingpong / Himeno NCCL sub-records at --gpus >= 2")
    ap.add_argument("--cpu-log2n", type=int, default=28, help="bounded CPU sample = 2^cpu_log2n elements (default: the whole per-GPU field, ~15 s of CPU work over 16 cores)")
    return ap.parse_args()


def workload_name(a):
    return "CT=%d %s float32 field, 2^%d elements per GPU, absErrorBound=%g, chunk=%d" % (
        a.ct, a.field, a.log2n, a.eb, a.chunk)


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------ clocks sampler
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            f = [v.strip() for v in line.split(",")]
            if len(f) < 6:
                continue
            try:
                clk, mx = float(f[0]), float(f[1])
            except ValueError:
                continue
            if t0 - 0.05 <= ts <= t1 + 0.15:
                sm.append(clk)
                for nm, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------ reference on CPU
def _ref_worker(args):
    """One forked worker: the reference run chunk-by-chunk over its slice (the decoder leaks ~48 B/value,
    SURVEY F10 -- the leak dies with the process)."""
    import oracle_lib as ol
    kind, start, count, chunk, ct, eb, use_ref = args
    from data_compression_b200.fields import make_field
    x = make_field(kind, count, start=start)
    base = 5 if ct in (5, 8) else (7 if ct in (7, 9) else ct)
    t0 = time.perf_counter()
    nbytes = 0
    if use_ref:
        R = ol.RefLib(eb)
        for lo in range(0, count, chunk):
            seg = x[lo:lo + chunk]
            mn, y = R.min_subtract(seg)
            t, m17 = 0, 0
            if base == 7:
                med, t = R.med(y, 0)
                m17 = ol.mask17_of(med)
            s, _ = R.compress(base, y, t, m17)
            if ct in (8, 9):
                R.crc32(s)
            d = R.decompress(base, s, y.size, t, m17, pad=False)
            with np.errstate(invalid="ignore"):            # the reference's CT=7 decoder leaves undefined values in its corners
                d += mn
            nbytes += s.size
    else:
        enc = ol.o_compress_chunked(base, x, chunk, eb, want_crc=ct in (8, 9))
        ol.o_decompress_chunked(base, enc, count, chunk, eb)
        nbytes = enc["stream"].size
    return time.perf_counter() - t0, nbytes


def cpu_reference_run(a, log2n, start=0):
    """-> (input GB/s over all cores, cores, kind, sample description, seconds)"""
    import multiprocessing as mp
    import oracle_lib as ol
    ol.build_oracle()
    use_ref = ol.ref_available(a.eb)
    if use_ref:
        ol.RefLib(a.eb)                      # dlopen oracle/_ref/libdcref_<eb>.so in the parent: the forked workers inherit it
    cores = len(ALL_CPUS) or 1
    n = 1 << log2n
    per = max(a.chunk, (n // cores) // a.chunk * a.chunk)
    jobs = [(a.field, start + lo, min(per, n - lo), a.chunk, a.ct, a.eb, use_ref) for lo in range(0, n, per)]
    ctx = mp.get_context("fork")
    narrowed = os.sched_getaffinity(0)
    os.sched_setaffinity(0, ALL_CPUS)            # the CPU baseline gets every host core (forked workers inherit the mask)
    t0 = time.perf_counter()
    try:
        with ctx.Pool(min(cores, len(jobs))) as pool:
            res = pool.map(_ref_worker, jobs)
    finally:
        os.sched_setaffinity(0, narrowed)
    wall = time.perf_counter() - t0
    busy = max(r[0] for r in res)
    gbs = 4.0 * n / busy / 1e9
    kind = "reference" if use_ref else "port"
    sample = "first 2^%d elements of the same field, chunk-by-chunk (C=%d), %d forked workers, slowest worker %.2f s (wall %.2f s)" % (
        log2n, a.chunk, min(cores, len(jobs)), busy, wall)
    return gbs, min(cores, len(jobs)), kind, sample, busy


# ------------------------------------------------------------------------------------------ reference arm
def run_reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return                                   # rank 0 alone runs and prints it
    times, gbs_list = [], []
    info = None
    log2n = a.cpu_log2n
    for i in range(a.warmup + a.steps):
        gbs, cores, kind, sample, secs = cpu_reference_run(a, log2n)
        info = (cores, kind, sample)
        if i >= a.warmup:
            times.append(secs)
            gbs_list.append(gbs)
    val = float(np.mean(gbs_list))
    cores, kind, sample = info
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * float(np.mean(times)), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32/u32", "data": "synthetic",
        "config": {"workload": workload_name(a), "ct": a.ct, "field": a.field, "n_per_gpu": 1 << a.log2n,
                   "abs_error_bound": a.eb, "chunk_elems": a.chunk, "sample": sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------ sweep
EBS = (1e-1, 1e-2, 1e-3, 1e-4, 1e-5, 1e-6)


def sweep_plan(kind, log2n):
    """(ct, field, n, eb, chunk) of every extra workload.  BASELINE.json: configs[2] = CT 5/6/11 x absErrorBound 1e-1..1e-6
    (the loop of impl/experiment.sh:9-19), configs[4] = CT 8/9 on the noisy field, configs[0] = the pingpong shape
    (131,072 floats, ONE chunk: what impl/pingpong.c:164-169 and :553-566 call), configs[3]'s Himeno face (33,929 floats)."""
    n = 1 << log2n
    plan = []
    if kind == "quick":
        for ct in (5, 7):
            for f in ("smooth", "noisy", "constant"):
                plan.append((ct, f, n, 0.01, 4096))
        plan.append((5, "smooth", 131072, 0.01, 0))
        return plan
    for ct in (5, 6, 7, 11):
        for f in ("smooth", "noisy", "constant"):
            plan.append((ct, f, n, 0.01, 4096))
    for ct in (5, 6, 11):
        for f in ("smooth", "noisy"):
            for eb in EBS:
                if eb != 0.01:
                    plan.append((ct, f, n, eb, 4096))
    for ct in (8, 9):
        for f in ("noisy", "smooth"):
            plan.append((ct, f, n, 0.01, 4096))
    for ct in (5, 7):
        plan.append((ct, "smooth", 1 << 30, 0.01, 4096))
    for ct in (8, 9):
        plan.append((ct, "noisy", 1 << 30, 0.01, 4096))          # BASELINE.json configs[4] at its own size, on one GPU
    for ct in (5, 7, 6, 11):
        plan.append((ct, "smooth", 131072, 0.01, 0))            # pingpong shape, one chunk (legacy call)
    plan.append((5, "noisy", 131072, 0.01, 0))
    plan.append((5, "smooth", 33929, 0.01, 0))                  # Himeno face, size L 2x2x2 (impl/himenoBMTxps.c:644-708)
    return plan


def oracle_check(api, ol, x, comp, out, ct, eb, chunk):
    """Sampled chunks of the finished round trip against the CPU oracle (stream bytes, CRC, decoded floats).  The oracle is
    the checker only; nothing here is timed."""
    n = x.numel()
    base = 5 if ct == 8 else (7 if ct == 9 else ct)
    offs = comp.chunk_offsets.cpu().numpy().astype(np.int64)
    if chunk == 0 or chunk >= n:
        cs, C = [0], n
    else:
        K = offs.size - 1
        cs, C = sorted(set([0, K // 3, (2 * K) // 3, K - 1])), chunk
    for c in cs:
        lo, hi = c * C, min(n, (c + 1) * C)
        xs = x[lo:hi].cpu().numpy()
        enc = ol.o_compress_chunked(base, xs, hi - lo, eb, want_crc=ct in (8, 9))
        got = comp.stream[int(offs[c]):int(offs[c + 1])].cpu().numpy()
        if got.size != enc["stream"].size or not np.array_equal(got, enc["stream"]):
            return False
        if ct in (8, 9) and int(comp.chunk_crc[c].item()) & 0xFFFFFFFF != int(enc["chunk_crc"][0]):
            return False
        dec = ol.o_decompress_chunked(base, enc, hi - lo, hi - lo, eb)
        if not np.array_equal(out[lo:hi].cpu().numpy().view(np.uint32), dec.view(np.uint32)):
            return False
    return True


def run_sweep(a, dev, world, rank, peak, fields):
    """-> list of entries; every rank measures its own slice (weak scaling, like the headline), times are max over ranks."""
    import torch
    import torch.distributed as dist
    from data_compression_b200 import api
    from data_compression_b200.fields import make_field_torch
    import oracle_lib as ol
    ol.build_oracle()
    plan = sweep_plan(a.sweep, a.log2n)
    rows, times = [], []
    ws_cache = {}
    for (ct, field, n, eb, chunk) in plan:
        key = (field, n)
        if key not in fields:
            for k in [k for k in fields if k[1] > (1 << 20)]:                  # one big field at a time
                del fields[k]
            for k in [k for k in ws_cache if k[0] != n and k[0] > (1 << 20)]:
                del ws_cache[k]
            torch.cuda.empty_cache()
            fields[key] = make_field_torch(field, n, dev, start=rank * n)
        x = fields[key]
        if (n, chunk) not in ws_cache:
            ws_cache[(n, chunk)] = (api.Workspace(n, chunk, dev), torch.empty(n, dtype=torch.float32, device=dev),
                                    torch.zeros(1, dtype=torch.int32, device=dev))
        ws, out, status = ws_cache[(n, chunk)]
        steps = a.sweep_steps * (4 if n < (1 << 20) else 1)
        for _ in range(2):
            comp = api.compress(x, ct=ct, abs_error_bound=eb, chunk_elems=chunk, subtract_min=True, workspace=ws)
            api.decompress(comp, out=out, status=status)
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
        torch.cuda.synchronize()
        for i in range(steps):
            ev[i][0].record()
            comp = api.compress(x, ct=ct, abs_error_bound=eb, chunk_elems=chunk, subtract_min=True, workspace=ws)
            ev[i][1].record()
            api.decompress(comp, out=out, status=status)
            ev[i][2].record()
        torch.cuda.synchronize()
        tc = float(np.mean([ev[i][0].elapsed_time(ev[i][1]) for i in range(steps)]))
        td = float(np.mean([ev[i][1].elapsed_time(ev[i][2]) for i in range(steps)]))
        B = float(comp.total_bytes())
        ok = bool(status.item() == 0) and oracle_check(api, ol, x, comp, out, ct, eb, chunk)
        rows.append({"ct": ct, "field": field, "n": n, "eb": eb, "chunk": chunk, "B": B, "ok": ok})
        times.append([tc, td])
    # ---- float64 codecs (the *_double entry points of the k-means / MM / LU callers), 2^(log2n-1) doubles
    del ws_cache
    fields.clear()
    torch.cuda.empty_cache()
    n64 = 1 << (a.log2n - 1)
    f64_plan = [] if a.sweep == "quick" else [(5, "smooth", n64, 4096), (7, "smooth", n64, 4096), (6, "smooth", n64, 4096),
                                              (11, "smooth", n64, 4096), (5, "noisy", n64, 4096), (7, "noisy", n64, 4096),
                                              (5, "smooth", 1 << 20, 0), (5, "noisy", 1 << 20, 0)]   # chunk 0: ONE call (k-means / MM / LU shape)
    for ct, field, n64, ch64 in f64_plan:
        x = make_field_torch(field, n64, dev, start=rank * n64).double()
        ws = api.WorkspaceF64(n64, ch64, dev)
        out = torch.empty(n64, dtype=torch.float64, device=dev)
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        for _ in range(2):
            comp = api.compress_f64(x, ct=ct, abs_error_bound=0.01, chunk_elems=ch64, workspace=ws)
            api.decompress_f64(comp, out=out, status=status)
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(a.sweep_steps)]
        torch.cuda.synchronize()
        for i in range(a.sweep_steps):
            ev[i][0].record()
            comp = api.compress_f64(x, ct=ct, abs_error_bound=0.01, chunk_elems=ch64, workspace=ws)
            ev[i][1].record()
            api.decompress_f64(comp, out=out, status=status)
            ev[i][2].record()
        torch.cuda.synchronize()
        tc = float(np.mean([e[0].elapsed_time(e[1]) for e in ev]))
        td = float(np.mean([e[1].elapsed_time(e[2]) for e in ev]))
        offs = comp.chunk_offsets.cpu().numpy().astype(np.int64)
        ok = bool(status.item() == 0)
        if ch64 == 0:                                                      # one call: the whole array against the oracle
            xs = x.cpu().numpy()
            enc = ol.o_compress_chunked_f64(ct, xs, n64, 0.01)
            ok = ok and np.array_equal(comp.stream[:int(offs[-1])].cpu().numpy(), enc["stream"])
            dec = ol.o_decompress_chunked_f64(ct, enc, n64, n64, 0.01)
            ok = ok and np.array_equal(out.cpu().numpy().view(np.uint64), dec.view(np.uint64))
        for c in (() if ch64 == 0 else (0, (offs.size - 1) // 2, offs.size - 2)):   # sampled chunks against the float64 oracle
            xs = x[c * 4096:(c + 1) * 4096].cpu().numpy()
            enc = ol.o_compress_chunked_f64(ct, xs, 4096, 0.01)
            ok = ok and np.array_equal(comp.stream[int(offs[c]):int(offs[c + 1])].cpu().numpy(), enc["stream"])
            dec = ol.o_decompress_chunked_f64(ct, enc, 4096, 4096, 0.01)
            ok = ok and np.array_equal(out[c * 4096:(c + 1) * 4096].cpu().numpy().view(np.uint64), dec.view(np.uint64))
        rows.append({"ct": ct, "field": field, "n": n64, "eb": 0.01, "chunk": ch64, "B": float(comp.total_bytes()), "ok": ok,
                     "elem": 8})
        times.append([tc, td])
        del x, ws, out
    tt = torch.tensor(times, dtype=torch.float64, device=dev)
    tb = torch.tensor([r["B"] for r in rows], dtype=torch.float64, device=dev)
    okt = torch.tensor([1.0 if r["ok"] else 0.0 for r in rows], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(tb, op=dist.ReduceOp.SUM)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
    tt, tb, okt = tt.cpu().numpy(), tb.cpu().numpy(), okt.cpu().numpy()
    out_rows = []
    for r, (tc, td), Bsum, okv in zip(rows, tt, tb, okt):
        n = r["n"]
        el = float(r.get("elem", 4))                       # bytes per element: 4 (float32) or 8 (the float64 codecs)
        Bg = Bsum / world                                  # stream bytes per GPU
        out_rows.append({
            "ct": r["ct"], "field": r["field"], "dtype": "f64" if el == 8 else "f32", "log2n": round(float(np.log2(n)), 3),
            "n_per_gpu": n, "eb": r["eb"],
            "chunk": r["chunk"], "ratio": el * n / max(Bg, 1.0), "compress_ms": tc, "decompress_ms": td,
            "input_gbs": el * n * world / ((tc + td) * 1e-3) / 1e9,
            "roundtrip_frac_of_hbm": (2 * el * n + 2.0 * Bg) / ((tc + td) * 1e-3) / 1e9 / peak,
            "compress_frac_of_hbm": (el * n + Bg) / (tc * 1e-3) / 1e9 / peak,
            "decompress_frac_of_hbm": (el * n + Bg) / (td * 1e-3) / 1e9 / peak,
            "oracle_check": bool(okv > 0.5)})
    return out_rows


# ------------------------------------------------------------------------------------------ row N4 (CT=1, CT=10)
def n4_records(a, dev, peak):
    """SURVEY.md 8f row N4, measured (rank 0's GPU): the byte-wise codec CT=1 (myCompress / myDecompress,
    impl/dataCompression.c:3977 / :3942) chunk by chunk, and the Hamming SECDED pass of CT=10 (hamming_encode / hamming_decode
    :5742 / :5752 as impl/pingpong.c:296-311, :487-505 call them) over a finished CT=5 stream, blocks of 125,000 bytes.
    Sampled chunks / blocks are compared with the oracle."""
    import torch
    from data_compression_b200 import api
    from data_compression_b200.fields import make_field_torch
    import oracle_lib as ol
    n, chunk, eb, steps = 1 << a.log2n, 4096, 0.01, a.sweep_steps
    recs = []

    def timed(fn):
        for _ in range(2):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            r = fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps, r

    for field in ("smooth", "noisy"):
        x = make_field_torch(field, n, dev)
        enc = api.bytewise_compress(x, eb, chunk)
        out = torch.empty(n, dtype=torch.float32, device=dev)
        tc, enc = timed(lambda: api.bytewise_compress(x, eb, chunk, buffers=enc))
        td, _ = timed(lambda: api.bytewise_decompress(enc, out=out))
        nf = enc["nfloat"].cpu().numpy().astype(np.int64)
        nfl = int(nf.sum())
        B = 4.0 * nfl + 5.0 * (n - nfl)                      # floats + (char, int displacement) per hit, as the reference sends them
        ok = True
        for c in (0, nf.size // 2, nf.size - 1):
            xs = x[c * chunk:(c + 1) * chunk].cpu().numpy()
            af, ac, ad = ol.o_bytewise_compress(xs, eb)
            lo = c * chunk
            ok = ok and int(nf[c]) == af.size
            ok = ok and np.array_equal(enc["array_float"][lo:lo + af.size].cpu().numpy().view(np.uint32), af.view(np.uint32))
            ok = ok and np.array_equal(enc["array_char"][lo:lo + ac.size].cpu().numpy(), ac)
            ok = ok and np.array_equal(enc["displacement"][lo:lo + ad.size].cpu().numpy(), ad)
            dec = ol.o_bytewise_decompress(af, ac, ad, xs.size)
            ok = ok and np.array_equal(out[lo:lo + xs.size].cpu().numpy().view(np.uint32), dec.view(np.uint32))
        recs.append({"what": "bytewise_ct1", "field": field, "n": n, "eb": eb, "chunk": chunk, "ratio": 4.0 * n / B,
                     "compress_ms": tc, "decompress_ms": td, "input_gbs": 4.0 * n / ((tc + td) * 1e-3) / 1e9,
                     "compress_frac_of_hbm": (4.0 * n + B) / (tc * 1e-3) / 1e9 / peak,
                     "decompress_frac_of_hbm": (4.0 * n + B) / (td * 1e-3) / 1e9 / peak, "oracle_check": bool(ok)})
        del enc, out
        if field == "noisy":
            comp = api.compress(x, ct=5, abs_error_bound=eb, chunk_elems=chunk)
            total, bs = comp.total_bytes(), 125000
            te, (chk, r) = timed(lambda: api.hamming_encode(comp.stream, total, bs))
            tdec, et = timed(lambda: api.hamming_decode(comp.stream, total, bs, chk))
            nb = chk.numel()
            ok = bool((et == 0).all().item())
            for b in (0, nb // 2, nb - 1):
                blk = comp.stream[b * bs:min(total, (b + 1) * bs)].cpu().numpy()
                orr, oc = ol.o_hamming_encode(blk)
                got = bytes(ord("0") + ((int(chk[b].item()) >> i) & 1) for i in range(orr + 1))
                ok = ok and int(r[b].item()) == orr and got == oc
            recs.append({"what": "hamming_ct10", "stream_bytes": total, "block_bytes": bs, "blocks": nb, "encode_ms": te,
                         "decode_ms": tdec, "encode_gbs": total / (te * 1e-3) / 1e9, "decode_gbs": total / (tdec * 1e-3) / 1e9,
                         "encode_frac_of_hbm": total / (te * 1e-3) / 1e9 / peak, "oracle_check": bool(ok)})
            del comp
        del x
        torch.cuda.empty_cache()
    return recs


# ------------------------------------------------------------------------------------------ NCCL sub-records (N >= 2)
def nccl_records(a, dev, rank, world):
    """Compressed payloads moving between GPUs with NCCL send/recv (BASELINE.json configs[0] and configs[3]): the pingpong hop
    of impl/pingpong.c:226-611 between rank pairs, and the Himeno solver with compressed k-faces (impl/himenoBMTxps.c:247-299,
    :644-708) on all ranks.  Times are CUDA events, max over ranks."""
    import importlib.util
    import torch
    import torch.distributed as dist
    from data_compression_b200 import api, transport as tr
    from data_compression_b200.fields import make_field_torch
    peer = rank ^ 1
    paired = peer < world
    sender_first = (rank & 1) == 0

    def timed(fn, reps):
        fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize(); dist.barrier()
        t = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    pp = []
    for name, field, n, chunk, reps in (("pingpong 131072 (float_eq shape, one chunk)", "constant", 131072, 0, 50),
                                        ("pingpong 131072 smooth (one chunk)", "smooth", 131072, 0, 50),
                                        ("pingpong 2^24 smooth (chunk 4096)", "smooth", 1 << 24, 4096, 10)):
        x = make_field_torch(field, n, dev, start=0)
        ws = api.Workspace(n, chunk, dev)
        out = torch.empty(n, dtype=torch.float32, device=dev)
        rec = {"case": name, "n": n, "chunk": chunk, "field": field}
        xr = torch.empty_like(x)

        def hop_raw():                                          # CT=0: impl/pingpong.c:233 / :336
            if not paired:
                return
            if sender_first:
                dist.send(x, peer); dist.recv(xr, peer)
            else:
                dist.recv(xr, peer); dist.send(x, peer)
        rec["ct0_round_trip_us"] = 1e3 * timed(hop_raw, reps)
        for ct in (5, 8):
            fixed = tr.use_fixed_frame(n, chunk)
            cws = tr.FrameWorkspace(n, chunk, dev, ct=ct) if fixed else ws        # the codec writes straight into the frame
            comp = api.compress(x, ct=ct, abs_error_bound=0.01, chunk_elems=chunk, subtract_min=True, workspace=cws)
            flags = comp.flags | (4 if ct == 8 else 0)
            got_box = [None]
            rbuf = torch.empty(tr.fixed_frame_bytes(n, chunk), dtype=torch.uint8, device=dev) if fixed else None

            def hop():
                if not paired:
                    return
                if fixed:
                    if sender_first:
                        dist.send(cws.frame, peer)
                        got_box[0] = tr.recv_fixed(n, chunk, peer, dev, ct, 0.01, flags, api.Compressed, buffer=rbuf)
                    else:
                        got_box[0] = tr.recv_fixed(n, chunk, peer, dev, ct, 0.01, flags, api.Compressed, buffer=rbuf)
                        dist.send(cws.frame, peer)
                else:
                    if sender_first:
                        tr.send_compressed(comp, peer); got_box[0] = tr.recv_compressed(n, chunk, peer, dev, api.Compressed)
                    else:
                        got_box[0] = tr.recv_compressed(n, chunk, peer, dev, api.Compressed); tr.send_compressed(comp, peer)
            rec["ct%d_round_trip_us" % ct] = 1e3 * timed(hop, reps)
            rec["ct%d_frame" % ct] = "fixed-capacity, one message" if fixed else "table + stream, two messages"
            rec["ct%d_compress_us" % ct] = 1e3 * timed(lambda: api.compress(x, ct=ct, abs_error_bound=0.01, chunk_elems=chunk,
                                                                              subtract_min=True, workspace=cws), reps)
            ok = True
            if paired:
                got = got_box[0]
                rec["ct%d_decompress_us" % ct] = 1e3 * timed(lambda: api.decompress(got, out=out), reps)
                K = got.chunk_offsets.numel() - 1
                ok = bool(torch.equal(out, api.decompress(comp)))        # both ranks hold the same field: the peer's array decodes to mine
                if ct == 8:                                              # receiver side of impl/pingpong.c:380-410
                    crc = api.crc32_chunks(got.stream, got.chunk_offsets)
                    ok = ok and bool(torch.equal(crc[:K].view(torch.int32), got.chunk_crc[:K].view(torch.int32)))
            else:
                timed(lambda: None, 1)
            okt = torch.tensor([1 if ok else 0], device=dev)
            dist.all_reduce(okt, op=dist.ReduceOp.MIN)
            rec["ct%d_check" % ct] = bool(okt.item())
            rec["ct%d_bytes" % ct] = int(comp.total_bytes())
        pp.append(rec)
        del x, ws, out, xr

    spec = importlib.util.spec_from_file_location("himeno_nccl", os.path.join(ROOT, "examples", "himeno_nccl.py"))
    hm = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(hm)
    m, nn, eb = 257, 20, 1e-4                                   # size L faces: 257 x 257 = 66,049 values
    him = {"grid": "%dx%dx%d on %d ranks (chain along k)" % (m, m, world * (m - 2) + 2, world), "nn": nn, "abs_error_bound": eb}
    base = None
    for label, ct, mode in (("ct0", 0, "fixed"), ("ct5_one_call_per_face", 5, "fixed"), ("ct5_chunked_4096", 5, "fixed_chunked"),
                            ("ct5_two_rounds_r1", 5, "two_round")):
        import argparse as _ap
        args = _ap.Namespace(size=m, eb=eb, nn=nn, mode=mode)
        hm.run(_ap.Namespace(size=m, eb=eb, nn=2, mode=mode), ct, rank, world, dev)
        blk, gosa, sent, ms = hm.run(args, ct, rank, world, dev)
        him[label + "_ms_per_iter"] = ms / nn
        him[label + "_gosa"] = gosa
        him[label + "_halo_bytes_rank0"] = sent
        if ct == 0:
            base = blk.p.clone()
        else:
            d = torch.tensor([float((blk.p - base).abs().max().item())], dtype=torch.float64, device=dev)
            dist.all_reduce(d, op=dist.ReduceOp.MAX)
            him[label + "_max_dev_from_ct0"] = float(d.item())
            him[label + "_within_nn_eb"] = bool(d.item() <= nn * eb)
    return {"pingpong": pp, "himeno": him}


def bind_to_gpu_numa_node(index):
    """CPU affinity of this process = the cores NVML reports next to GPU `index` (pinned host buffers then land on that NUMA
    node, and the copy threads run there).  Returns the number of CPUs, or None when NVML does not say."""
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = int(vis.split(",")[index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else index
        h = pynvml.nvmlDeviceGetHandleByIndex(phys)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


# ------------------------------------------------------------------------------------------ our arm
def run_ours(a):
    import torch
    import torch.distributed as dist
    from data_compression_b200 import api
    from data_compression_b200.fields import make_field_torch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    affinity = bind_to_gpu_numa_node(local)       # before any pinned allocation: first touch places the pages
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    api.lib()

    n = 1 << a.log2n
    x = make_field_torch(a.field, n, dev, start=rank * n)
    ws = api.Workspace(n, a.chunk, dev)
    out = torch.empty(n, dtype=torch.float32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    has_crc = a.ct in (8, 9)

    def step():
        comp = api.compress(x, ct=a.ct, abs_error_bound=a.eb, chunk_elems=a.chunk, subtract_min=True, workspace=ws)
        return comp

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(a.warmup, 3)):
        comp = step()
        api.decompress(comp, out=out, status=status)
    barrier()
    total_bytes = comp.total_bytes()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(a.steps)]
    barrier()
    t0 = time.time()
    for i in range(a.steps):
        ev[i][0].record()
        comp = step()
        ev[i][1].record()
        api.decompress(comp, out=out, status=status)
        ev[i][2].record()
    barrier()
    t1 = time.time()
    clocks = sampler.stop(t0, t1) if rank == 0 else None
    t_c = np.array([ev[i][0].elapsed_time(ev[i][1]) for i in range(a.steps)])     # ms
    t_d = np.array([ev[i][1].elapsed_time(ev[i][2]) for i in range(a.steps)])
    t_total = ev[0][0].elapsed_time(ev[-1][2])                                      # device time of the K steps
    tt = torch.tensor([t_total, float(t_c.mean()), float(t_d.mean())], dtype=torch.float64, device=dev)
    tb = torch.tensor([float(total_bytes)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(tb, op=dist.ReduceOp.SUM)
    t_total, tc_ms, td_ms = [float(v) for v in tt.cpu()]
    all_bytes = float(tb.item())

    # sanity: the round trip must reproduce the oracle's decode on a sampled chunk range (checked outside the clock)
    ok = bool(status.item() == 0)

    # per-kernel durations of the compress call (the slot path launches the encoder kernel, then the compaction kernel): the
    # library records CUDA events on the launching stream around each; a separate pass so that reading them back does not
    # put host synchronisations into the timed region above
    enc_k_ms = cmp_k_ms = None
    if a.chunk == 4096 or n <= 4096:
        import ctypes
        L = api.lib()
        L.dcb_profile_enable(1)
        e_ms, c_ms, es, cs = ctypes.c_float(), ctypes.c_float(), [], []
        for _ in range(a.steps):
            step()
            if L.dcb_profile_last_ms(ctypes.byref(e_ms), ctypes.byref(c_ms)) == 0:
                es.append(e_ms.value)
                cs.append(c_ms.value)
        L.dcb_profile_enable(0)
        if es:
            tk = torch.tensor([float(np.mean(es)), float(np.mean(cs))], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tk, op=dist.ReduceOp.MAX)
            enc_k_ms, cmp_k_ms = [float(v) for v in tk.cpu()]

    # ---- end to end through the host-pointer C ABI (pinned host buffers; H2D + D2H inside the timed region) ----
    e2e = None
    if not a.no_e2e:
        # T host threads, each making the blocking calls a reference driver makes (compress, then decompress) on its own
        # pinned buffers.  The library keeps one CUDA stream and one set of device buffers per host thread, so one
        # thread's host->device copy overlaps the other's device->host copy (PCIe is full duplex) -- the same overlap
        # two MPI ranks sharing a GPU would get.  T=1 is the strictly serial call sequence.
        T = max(1, a.e2e_threads)
        xh = torch.empty(n, dtype=torch.float32).pin_memory()
        xh.copy_(x)
        xnp = xh.numpy()
        ohs = [torch.empty(n, dtype=torch.float32).pin_memory() for _ in range(T)]
        shs = [torch.empty(api.lib().dcb_max_stream_bytes(n, a.chunk), dtype=torch.uint8).pin_memory() for _ in range(T)]
        reps = max(2, min(a.steps, 5))
        encs, errs = [None] * T, []

        # worker threads live across warm-up and timed repetitions (the library caches device buffers per host thread; a
        # thread that exits gives them back), stepped by barriers
        gate = threading.Barrier(T + 1)
        plan = {"count": 0, "stop": False}

        def worker(t):
            try:
                torch.cuda.set_device(dev)                  # the current CUDA device is per host thread
                while True:
                    gate.wait()
                    if plan["stop"]:
                        break
                    count = plan["count"]
                    # odd threads start with the decompress of their warm-up result, so that one thread's host->device copy
                    # runs against the other's device->host copy instead of both pulling in the same direction
                    for _ in range(count):
                        if t % 2 == 1 and encs[t] is not None:
                            api.decompress_host(encs[t], out=ohs[t].numpy())
                        encs[t] = api.compress_host(xnp, ct=a.ct, abs_error_bound=a.eb, chunk_elems=a.chunk,
                                                    out={"stream": shs[t].numpy()})
                        if not (t % 2 == 1 and count == reps):
                            api.decompress_host(encs[t], out=ohs[t].numpy())
                    gate.wait()
            except Exception as e:  # pragma: no cover
                errs.append(e)
                gate.abort()
            finally:
                api.lib().dcb_host_release()

        ths = [threading.Thread(target=worker, args=(t,)) for t in range(T)]
        for th in ths:
            th.start()

        def run(count):
            plan["count"] = count
            gate.wait()                                     # go
            gate.wait()                                     # done
            if errs:
                raise errs[0]

        run(1)                                              # warm-up: device buffers of every thread get allocated
        barrier()
        te0 = time.perf_counter()
        run(reps)
        barrier()
        te = (time.perf_counter() - te0) / (reps * T)
        plan["stop"] = True
        gate.wait()
        for th in ths:
            th.join()
        tmax = torch.tensor([te], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        te = float(tmax.item())
        enc = encs[0]
        assert all(torch.equal(o, ohs[0]) for o in ohs[1:])
        K = enc["offsets"].size - 1
        tables = 8 * (K + 1) + 12 * K
        e2e = {"value": 4.0 * n * world / te / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": int(4 * n + enc["total_bytes"] + tables),
               "d2h_bytes_per_step": int(4 * n + enc["total_bytes"] + tables),
               "ms_per_step": 1e3 * te, "host_threads": T, "cpu_affinity_cpus": affinity,
               "api": "dcb_compress_host + dcb_decompress_host (blocking C-ABI calls, pinned host buffers, "
                      "%d host thread%s per GPU)" % (T, "" if T == 1 else "s")}
        # ---- what ONE caller gets: the strictly serial call sequence of a reference driver (compress, then decompress, one
        #      thread), with pinned buffers and with plain malloc'd (pageable) ones, as impl/pingpong.c would pass them
        def serial(xin, sbuf, obuf, k):
            api.decompress_host(api.compress_host(xin, ct=a.ct, abs_error_bound=a.eb, chunk_elems=a.chunk, out={"stream": sbuf}), out=obuf)
            barrier()
            t0_ = time.perf_counter()
            for _ in range(k):
                api.decompress_host(api.compress_host(xin, ct=a.ct, abs_error_bound=a.eb, chunk_elems=a.chunk, out={"stream": sbuf}), out=obuf)
            barrier()
            tt_ = torch.tensor([(time.perf_counter() - t0_) / k], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tt_, op=dist.ReduceOp.MAX)
            return 4.0 * n * world / float(tt_.item()) / 1e9
        e2e["serial_pinned_gbs"] = serial(xnp, shs[0].numpy(), ohs[0].numpy(), 2)
        xpg = np.empty(n, dtype=np.float32); xpg[:] = xnp
        e2e["serial_pageable_gbs"] = serial(xpg, np.empty(shs[0].numel(), dtype=np.uint8), np.empty(n, dtype=np.float32), 2)
        api.lib().dcb_host_release()
        # ---- the box's ceiling for this traffic: all ranks copy 4n + B bytes host->device and device->host at the same
        #      time (plain cudaMemcpyAsync on two streams, pinned buffers); a round trip moves exactly that much each way
        B_ = int(enc["total_bytes"])
        nb = 4 * n + B_
        hsrc = torch.empty(nb, dtype=torch.uint8).pin_memory()
        hdst = torch.empty(nb, dtype=torch.uint8).pin_memory()
        dsrc = torch.empty(nb, dtype=torch.uint8, device=dev)
        ddst = torch.empty(nb, dtype=torch.uint8, device=dev)
        s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

        def both():
            with torch.cuda.stream(s1):
                ddst.copy_(hsrc, non_blocking=True)
            with torch.cuda.stream(s2):
                hdst.copy_(dsrc, non_blocking=True)
            s1.synchronize(); s2.synchronize()
        both()
        barrier()
        t0_ = time.perf_counter()
        for _ in range(3):
            both()
        barrier()
        tc_ = torch.tensor([(time.perf_counter() - t0_) / 3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tc_, op=dist.ReduceOp.MAX)
        ceil_gbs = 4.0 * n * world / float(tc_.item()) / 1e9
        e2e["ceiling"] = {"input_gbs": ceil_gbs, "per_direction_gbs_per_gpu": nb / float(tc_.item()) / 1e9,
                          "how": "every rank copies 4n+B bytes H2D and D2H concurrently (cudaMemcpyAsync, pinned, two streams)"}
        e2e["frac_of_ceiling"] = e2e["value"] / ceil_gbs
        del hsrc, hdst, dsrc, ddst

    peak, peak_src = measured_peak()
    sweep = None
    if a.sweep != "none":
        del out
        sweep = run_sweep(a, dev, world, rank, peak, {(a.field, n): x})

    n4 = None
    if a.sweep == "full" and rank == 0:
        n4 = n4_records(a, dev, peak)

    nccl = None
    if world > 1 and not a.no_nccl:
        nccl = nccl_records(a, dev, rank, world)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    traffic = None                                  # dram bytes per launch of the dominant kernel, from the committed ncu capture
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        key = "ct%d/%s/2^%d/eb%g/chunk%d" % (a.ct, a.field, a.log2n, a.eb, a.chunk)
        traffic = tj.get(key)
    except Exception:
        pass
    B = float(total_bytes)
    alg_c, alg_d = 4.0 * n + B, B + 4.0 * n
    # one-tile chunks take the slot path: the warp-autonomous encoder kernel plus the
    # compaction kernel; the CUDA-event bracket of `compress` covers both (and the workspace memset), which is what is divided
    slot_path = (a.chunk == 4096) or n <= 4096
    enc_kernel = "dcb_encode_warp_kernel" if slot_path else "dcb_encode_kernel"
    dom = enc_kernel if tc_ms >= td_ms else "dcb_decode_kernel"
    # the encoder kernel alone when its own events are available (it reads 4n and writes B into the slots: the same
    # algorithmic bytes), else the whole compress call
    t_dom_ms = (enc_k_ms if enc_k_ms else tc_ms) if tc_ms >= td_ms else td_ms
    ach = ((alg_c if tc_ms >= td_ms else alg_d) / (t_dom_ms * 1e-3)) / 1e9
    value = 4.0 * n * world * a.steps / (t_total * 1e-3) / 1e9
    launches_per_step = (3 if slot_path else 2) + (1 if has_crc else 0)      # encode [+ compact] + decode [+ crc]
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
        "ms_per_step": t_total / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32/u32", "data": "synthetic",
        "config": {"workload": workload_name(a), "ct": a.ct, "field": a.field, "n_per_gpu": n,
                   "abs_error_bound": a.eb, "chunk_elems": a.chunk, "compression_ratio": 4.0 * n * world / all_bytes,
                   "l2": "inputs larger than L2 (%.0f MiB per GPU vs 126 MB)" % (4.0 * n / 2**20),
                   "compress_ms": tc_ms, "decompress_ms": td_ms,
                   "compress_input_gbs": 4.0 * n / tc_ms / 1e6, "decompress_input_gbs": 4.0 * n / td_ms / 1e6,
                   "roundtrip_frac_of_hbm": ((alg_c + alg_d) / ((tc_ms + td_ms) * 1e-3) / 1e9) / peak,
                   "decode_status_ok": ok},
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                     "frac": ach / peak, "traffic": (traffic or {}).get(dom), "traffic_source": (traffic or {}).get("source"),
                     "kernel_ms": t_dom_ms, "encode_kernel_ms": enc_k_ms, "compact_kernel_ms": cmp_k_ms,
                     "timed": "achieved: CUDA events recorded by the library around the dominant kernel on its launching stream; "
                              "compress / decompress: events around the whole dcb_compress (encoder + compaction kernel) / dcb_decompress call",
                     "peak_source": peak_src,
                     "compress": {"achieved": alg_c / tc_ms / 1e6, "frac": alg_c / tc_ms / 1e6 / peak},
                     "decompress": {"achieved": alg_d / td_ms / 1e6, "frac": alg_d / td_ms / 1e6 / peak}},
        "clocks": clocks, "gpu_launches": launches_per_step * a.steps,
    }
    if e2e:
        line["e2e"] = e2e
    if sweep is not None:
        line["sweep"] = sweep
    if n4 is not None:
        line["n4"] = n4
    if nccl is not None:
        line["pingpong"] = nccl["pingpong"]
        line["himeno"] = nccl["himeno"]
    if world == 1 and not a.no_cpu:
        gbs, cores, kind, sample, _ = cpu_reference_run(a, a.cpu_log2n)
        line["cpu_baseline"] = {"value": gbs, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)
