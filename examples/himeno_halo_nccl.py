#!/usr/bin/env python
"""Himeno halo exchange with compressed faces over NCCL (BASELINE.json configs[3]; impl/himenoBMTxps.c sendp3, :580-650).
Ranks form an open chain along k; every rank keeps its pressure block p[MI][MJ][MK] on its GPU, gathers the two inner
k-faces, compresses each as one reference call (CT=5/6/7), exchanges the frames with batched isend/irecv and writes the
received faces into its halo planes -- no host copy anywhere.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 \
      examples/himeno_halo_nccl.py [--ct 5] [--size 129] [--eb 0.001] [--iters 20] [--check]
"""
import argparse
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from data_compression_b200 import api, halo          # noqa: E402


def block(rank, world, m):
    """Himeno's initial pressure p = k^2 / (kmax-1)^2 in GLOBAL k, plus a small smooth perturbation in i, j."""
    i, j, k = np.meshgrid(np.arange(m), np.arange(m), np.arange(m), indexing="ij")
    kg = k + rank * (m - 2)                                           # neighbouring blocks overlap by the two halo planes
    kmax_g = world * (m - 2) + 2
    return ((kg * kg) / float((kmax_g - 1) ** 2) + 0.01 * np.sin(0.05 * i) * np.cos(0.04 * j)).astype(np.float32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ct", type=int, default=5)
    ap.add_argument("--size", type=int, default=129)
    ap.add_argument("--eb", type=float, default=0.001)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    m = a.size
    p = torch.from_numpy(block(rank, world, m)).to(dev)
    lo = rank - 1 if rank > 0 else None
    hi = rank + 1 if rank < world - 1 else None
    sent = halo.exchange_axis(p, 3, m, m, m, lo, hi, ct=a.ct, eb=a.eb)     # warm-up + the exchange that is checked
    ok = True
    if a.check:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as ol
        pn = p.cpu().numpy()
        for peer, v_src, v_dst in ((hi, 1, m - 1), (lo, m - 2, 0)):
            if peer is None:
                continue
            face = block(peer, world, m)[:, :, v_src].reshape(-1).copy()
            enc = ol.o_compress_chunked(a.ct, face, face.size, a.eb)
            want = ol.o_decompress_chunked(a.ct, enc, face.size, face.size, a.eb)
            ok = ok and np.array_equal(pn[:, :, v_dst].reshape(-1).view(np.uint32), want.view(np.uint32))
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(a.iters):
        sent = halo.exchange_axis(p, 3, m, m, m, lo, hi, ct=a.ct, eb=a.eb)
    torch.cuda.synchronize()
    dist.barrier()
    dt = (time.perf_counter() - t0) / a.iters
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        faces = (1 if lo is not None else 0) + (1 if hi is not None else 0)
        raw = faces * m * m * 4
        print("himeno halo ct=%d %d^3 per rank, %d ranks: %.3f ms per exchange, %d -> %d bytes sent by rank 0 (ratio %.2f)%s" % (
            a.ct, m, world, 1e3 * dt, raw, sent, raw / max(sent, 1),
            "  oracle-check=%s" % ("OK" if flag.item() else "MISMATCH") if a.check else ""), flush=True)
    dist.destroy_process_group()
    if a.check and not flag.item():
        raise SystemExit(1)


if __name__ == "__main__":
    main()
