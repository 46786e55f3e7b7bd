#!/usr/bin/env python
"""2-GPU compressed ping-pong over NCCL (the shape of impl/pingpong.c:226-611 with the MPI hop replaced by
transport.send_compressed / recv_compressed, and the CT=8/9 CRC + ack + resend protocol of :260-289, :363-447).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      examples/pingpong_nccl.py [--ct 5] [--elems 131072] [--eb 0.01] [--hops 100] [--check]

Rank 0 compresses once (as the reference does, impl/pingpong.c:164-169), the compressed frame bounces `hops` times,
the last receiver decompresses.  --check compares the final floats with the CPU oracle (test infrastructure)."""
import argparse
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from data_compression_b200 import api, transport as tr          # noqa: E402
from data_compression_b200.fields import make_field             # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ct", type=int, default=5)
    ap.add_argument("--elems", dest="n", type=int, default=131072)             # float_eq_131072 shape (BASELINE.json configs[0])
    ap.add_argument("--eb", type=float, default=0.01)
    ap.add_argument("--chunk", type=int, default=4096)
    ap.add_argument("--hops", type=int, default=100)
    ap.add_argument("--field", default="noisy")
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    rank, local = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    assert dist.get_world_size() == 2
    peer = 1 - rank
    x = make_field(a.field, a.n)
    crc = a.ct in (8, 9, 10)
    cct = 8 if a.ct == 10 else a.ct                              # CT=10 = the CT=5 stream + CRC + Hamming SECDED blocks (impl/pingpong.c:290-326)
    comp = api.compress(torch.from_numpy(x).to(dev), ct=cct, abs_error_bound=a.eb, chunk_elems=a.chunk) if rank == 0 else None
    hamming = None
    if a.ct == 10:
        def hm_dec(stream, total, bs, chk):
            api.hamming_decode(stream, total, bs, chk)
        hamming = (125000, lambda s_, t_, b_: api.hamming_encode(s_, t_, b_)[0], hm_dec)     # block_size at BER = 1e-6 (:5868)
    resends = 0
    flipped = [False]

    def corrupt(got, attempt):                                   # bit_flip on the first delivery only (impl/pingpong.c:368-377)
        if not flipped[0]:
            api.flip_bits(got.stream, [11])
            flipped[0] = True

    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for hop in range(a.hops):
        sender = (hop % 2) == rank
        if crc:
            got, r = tr.pingpong_hop_crc(comp, peer, sender, a.n, a.chunk, dev, api.Compressed, api.crc32_chunks,
                                         corrupt_fn=corrupt if (hop == 0 and not sender) else None, hamming=hamming)
            resends += r
        elif sender:
            tr.send_compressed(comp, peer)
            got = None
        else:
            got = tr.recv_compressed(a.n, a.chunk, peer, dev, api.Compressed)
        if got is not None:
            comp = got
    torch.cuda.synchronize()
    dist.barrier()
    dt = time.perf_counter() - t0
    holder = a.hops % 2                                          # rank that received the last hop
    if rank == holder:
        out = api.decompress(comp).cpu().numpy()
        total = comp.total_bytes()
        line = "pingpong ct=%d n=%d eb=%g hops=%d: %.3f ms/hop, %d bytes/hop (ratio %.3f), resends=%d" % (
            a.ct, a.n, a.eb, a.hops, 1e3 * dt / a.hops, total, 4.0 * a.n / total, resends)
        if a.check:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import oracle_lib as ol
            base = 5 if a.ct in (5, 8, 10) else (7 if a.ct in (7, 9) else a.ct)
            enc = ol.o_compress_chunked(base, x, a.chunk, a.eb)
            want = ol.o_decompress_chunked(base, enc, a.n, a.chunk, a.eb)
            same = np.array_equal(out.view(np.uint32), want.view(np.uint32)) and \
                np.array_equal(comp.stream[:total].cpu().numpy(), enc["stream"])
            line += "  oracle-check=%s gosa=%.3e" % ("OK" if same else "MISMATCH", float(np.abs(out - x).mean()))
            if not same:
                print(line, flush=True)
                raise SystemExit(1)
        print(line, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
