"""world_size-2 `gloo` tests (CPU) of the N>1 host logic: chunk-range sharding and the send/recv frames of
data_compression_b200/transport.py.  The codec work in these tests is done by the CPU oracle (test infrastructure);
on the GPU box the same functions carry api.Compressed objects produced by the CUDA kernels."""
import os
import socket
from dataclasses import dataclass

import numpy as np
import pytest

import oracle_lib as ol
from data_compression_b200 import transport as tr
from data_compression_b200.fields import make_field

torch = pytest.importorskip("torch")


@dataclass
class HostCompressed:                      # same field order as api.Compressed
    ct: int
    n: int
    chunk_elems: int
    abs_error_bound: float
    flags: int
    fixed_type: int
    fixed_mask17: int
    stream: object
    chunk_offsets: object
    chunk_min: object
    chunk_medium: object
    chunk_type: object
    chunk_crc: object


def oracle_compressed(x, ct, chunk, eb):
    base = 5 if ct in (5, 8) else (7 if ct in (7, 9) else ct)
    enc = ol.o_compress_chunked(base, x, chunk, eb, want_crc=True)
    pad = np.zeros(((enc["stream"].size + 31) // 16) * 16, np.uint8)
    pad[:enc["stream"].size] = enc["stream"]
    return HostCompressed(ct, x.size, chunk, eb, 1, 0, 0, torch.from_numpy(pad),
                          torch.from_numpy(enc["offsets"].astype(np.int64)), torch.from_numpy(enc["chunk_min"]),
                          torch.from_numpy(enc["chunk_medium"]), torch.from_numpy(enc["chunk_type"]),
                          torch.from_numpy(enc["chunk_crc"].view(np.int32).copy())), enc


def _pack_check(c):
    return sum((1 << i) for i, ch in enumerate(c) if ch == ord("1"))


def crc_chunks(stream, offsets):
    s, o = stream.numpy(), offsets.numpy()
    return torch.from_numpy(np.array([ol.o_crc32(s[o[c]:o[c + 1]]) for c in range(o.size - 1)], dtype=np.uint32).view(np.int32))


def test_shard_range_partitions_chunks():
    for n, chunk in ((1 << 20, 4096), (20000, 4096), (4096 * 7 + 5, 4096), (100, 4096), (1 << 16, 0)):
        for world in (1, 2, 4, 8):
            pieces = [tr.shard_range(n, chunk, r, world) for r in range(world)]
            assert pieces[0][0] == 0 and pieces[-1][1] == n
            for (a0, a1), (b0, b1) in zip(pieces, pieces[1:]):
                assert a1 == b0 and a0 <= a1
            if chunk and chunk < n:
                assert all(lo % chunk == 0 for lo, _ in pieces)


def test_sharded_compress_equals_whole():
    """Chunks are independent: the per-rank streams concatenated are the whole-array stream."""
    x = make_field("noisy", 4096 * 13 + 100)
    whole = ol.o_compress_chunked(5, x, 4096, 0.01)
    for world in (2, 4):
        parts = []
        for r in range(world):
            lo, hi = tr.shard_range(x.size, 4096, r, world)
            parts.append(ol.o_compress_chunked(5, x[lo:hi], 4096, 0.01)["stream"])
        assert np.array_equal(np.concatenate(parts), whole["stream"])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        dev = torch.device("cpu")
        n, chunk, eb = 4096 * 5 + 77, 4096, 0.01
        x = make_field("smooth", n, start=1000 * rank)
        peer = 1 - rank
        # ---- plain hop (impl/pingpong.c:250 / :353), both directions -------------------------------------------------
        for ct in (5, 7):
            comp, enc = oracle_compressed(x, ct, chunk, eb)
            if rank == 0:
                tr.send_compressed(comp, peer)
                got = tr.recv_compressed(n, chunk, peer, dev, HostCompressed)
            else:
                got = tr.recv_compressed(n, chunk, peer, dev, HostCompressed)
                tr.send_compressed(comp, peer)
            xp = make_field("smooth", n, start=1000 * peer)
            want, wenc = oracle_compressed(xp, ct, chunk, eb)
            total = int(got.chunk_offsets[-1])
            assert (got.ct, got.n, got.chunk_elems, got.abs_error_bound) == (ct, n, chunk, eb)
            assert torch.equal(got.chunk_offsets, want.chunk_offsets)
            assert np.array_equal(got.stream[:total].numpy(), wenc["stream"])
            assert torch.equal(got.chunk_min.view(torch.int32), want.chunk_min.view(torch.int32))
            assert torch.equal(got.chunk_type, want.chunk_type) and torch.equal(got.chunk_crc, want.chunk_crc)
            dec = ol.o_decompress_chunked(ct, dict(stream=got.stream[:total].numpy(), offsets=got.chunk_offsets.numpy().astype(np.uint64),
                                                   chunk_min=got.chunk_min.numpy(), chunk_medium=got.chunk_medium.numpy(),
                                                   chunk_type=got.chunk_type.numpy()), n, chunk, eb)
            assert np.array_equal(dec.view(np.uint32), ol.o_decompress_chunked(ct, wenc, n, chunk, eb).view(np.uint32))
        # ---- the same hop as ONE fixed-capacity message (no length read back on the host) ---------------------------------
        for ct in (5, 7, 8):
            comp, enc = oracle_compressed(x, ct, chunk, eb)
            flags = 1 | (4 if ct == 8 else 0)
            comp.flags = flags
            assert tr.use_fixed_frame(n, chunk)
            if rank == 0:
                tr.send_fixed(comp, peer)
                got = tr.recv_fixed(n, chunk, peer, dev, ct, eb, flags, HostCompressed)
            else:
                got = tr.recv_fixed(n, chunk, peer, dev, ct, eb, flags, HostCompressed)
                tr.send_fixed(comp, peer)
            xp = make_field("smooth", n, start=1000 * peer)
            want, wenc = oracle_compressed(xp, ct, chunk, eb)
            total = int(got.chunk_offsets[-1])
            assert torch.equal(got.chunk_offsets, want.chunk_offsets)
            assert np.array_equal(got.stream[:total].numpy(), wenc["stream"])
            assert torch.equal(got.chunk_min.view(torch.int32), want.chunk_min.view(torch.int32))
            if ct == 7:
                assert torch.equal(got.chunk_type, want.chunk_type)
                assert torch.equal(got.chunk_medium.view(torch.int32), want.chunk_medium.view(torch.int32))
            else:
                assert got.chunk_medium is None and got.chunk_type is None
            if ct == 8:
                assert torch.equal(got.chunk_crc, want.chunk_crc)
            else:
                assert got.chunk_crc is None
        # ---- CT=8 hop with a corrupted first delivery: one resend (impl/pingpong.c:363-410) -----------------------------
        comp, enc = oracle_compressed(x, 8, chunk, eb)

        def corrupt(got, attempt):
            if attempt == 0:
                got.stream[5] ^= 0x10                         # the reference's bit_flip on the received buffer

        got, resends = tr.pingpong_hop_crc(comp, peer, rank == 0, n, chunk, dev, HostCompressed, crc_chunks,
                                           corrupt_fn=corrupt if rank == 1 else None)
        assert resends == 1
        if rank == 1:
            x0 = make_field("smooth", n, start=0)
            assert np.array_equal(got.stream[:int(got.chunk_offsets[-1])].numpy(), oracle_compressed(x0, 8, chunk, eb)[1]["stream"])
        # ---- CT=10 hop (impl/pingpong.c:290-326, :448-510): the same single-bit fault is rectified by the Hamming blocks ----
        def hm_encode(stream, total, bs):
            sn = stream.numpy()
            return torch.tensor([_pack_check(ol.o_hamming_encode(sn[b:min(total, b + bs)])[1]) for b in range(0, total, bs)], dtype=torch.int32)

        def hm_decode(stream, total, bs, chk):
            sn = stream.numpy()
            for i, b in enumerate(range(0, total, bs)):
                blk = sn[b:min(total, b + bs)]
                r = ol.o_hm_length(8 * blk.size)
                c = bytes(ord("0") + ((int(chk[i]) >> k) & 1) for k in range(r + 1))
                _, fixed, _ = ol.o_hamming_decode(blk, c, r)
                sn[b:b + blk.size] = fixed

        got, resends = tr.pingpong_hop_crc(comp, peer, rank == 0, n, chunk, dev, HostCompressed, crc_chunks,
                                           corrupt_fn=corrupt if rank == 1 else None, hamming=(512, hm_encode, hm_decode))
        assert resends == 0
        if rank == 1:
            assert np.array_equal(got.stream[:int(got.chunk_offsets[-1])].numpy(), oracle_compressed(x0, 8, chunk, eb)[1]["stream"])
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, "FAIL: %s\n%s" % (e, traceback.format_exc())))
    finally:
        dist.destroy_process_group()


def test_frames_over_gloo_world2():
    import torch.multiprocessing as mp
    ol.build_oracle()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res


# ---- Himeno halo exchange protocol (data_compression_b200/halo.py) with the oracle as the codec -------------------------
def _np_face(p, ijk, v, imax, jmax, kmax):
    if ijk == 1:
        return p[v, :jmax, :kmax].reshape(-1).copy()
    if ijk == 2:
        return p[:imax, v, :kmax].reshape(-1).copy()
    return p[:imax, :jmax, v].reshape(-1).copy()


def _halo_field(rank, MI, MJ, MK):
    i, j, k = np.meshgrid(np.arange(MI), np.arange(MJ), np.arange(MK), indexing="ij")
    return (0.01 * i + 0.02 * j + 0.03 * k + 10.0 * rank + 0.001 * np.sin(i * j + k)).astype(np.float32)


def _halo_worker(rank, world, port, q):
    import torch.distributed as dist
    from data_compression_b200 import halo
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        MI, MJ, MK = 10, 12, 14
        imax, jmax, kmax = 9, 11, 13
        eb = 0.001

        def compress_fn(p, ijk, v, im, jm, km, ct, e):
            face = _np_face(p.numpy(), ijk, v, im, jm, km)
            enc = ol.o_compress_chunked(ct, face, face.size, e, want_crc=True)
            pad = np.zeros(((enc["stream"].size + 31) // 16) * 16, np.uint8)
            pad[:enc["stream"].size] = enc["stream"]
            return HostCompressed(ct, face.size, 0, e, 1, 0, 0, torch.from_numpy(pad), torch.from_numpy(enc["offsets"].astype(np.int64)),
                                  torch.from_numpy(enc["chunk_min"]), torch.from_numpy(enc["chunk_medium"]),
                                  torch.from_numpy(enc["chunk_type"]), torch.from_numpy(enc["chunk_crc"].view(np.int32).copy()))

        def write_fn(p, ijk, v, im, jm, km, comp):
            total = int(comp.chunk_offsets[-1])
            enc = dict(stream=comp.stream[:total].numpy(), offsets=comp.chunk_offsets.numpy().astype(np.uint64),
                       chunk_min=comp.chunk_min.numpy(), chunk_medium=comp.chunk_medium.numpy(), chunk_type=comp.chunk_type.numpy())
            dec = ol.o_decompress_chunked(comp.ct, enc, comp.n, comp.n, comp.abs_error_bound)
            a = p.numpy()
            if ijk == 1:
                a[v, :jm, :km] = dec.reshape(jm, km)
            elif ijk == 2:
                a[:im, v, :km] = dec.reshape(im, km)
            else:
                a[:im, :jm, v] = dec.reshape(im, jm)

        def expect(src_rank, ijk, v):
            face = _np_face(_halo_field(src_rank, MI, MJ, MK), ijk, v, imax, jmax, kmax)
            enc = ol.o_compress_chunked(5, face, face.size, eb)
            return ol.o_decompress_chunked(5, enc, face.size, face.size, eb)

        peer = 1 - rank
        # (a) open chain along k: rank 0 has only a hi neighbour, rank 1 only a lo neighbour
        p = torch.from_numpy(_halo_field(rank, MI, MJ, MK))
        before = p.numpy().copy()
        sent = halo.exchange_axis(p, 3, imax, jmax, kmax, None if rank == 0 else 0, 1 if rank == 0 else None, ct=5, eb=eb,
                                  compressed_cls=HostCompressed, compress_fn=compress_fn, write_fn=write_fn)
        assert sent > 0
        a = p.numpy()
        if rank == 0:
            assert np.array_equal(_np_face(a, 3, kmax - 1, imax, jmax, kmax).view(np.uint32), expect(1, 3, 1).view(np.uint32))
            assert np.array_equal(a[:, :, 0], before[:, :, 0])                      # no lo neighbour: plane 0 untouched
        else:
            assert np.array_equal(_np_face(a, 3, 0, imax, jmax, kmax).view(np.uint32), expect(0, 3, kmax - 2).view(np.uint32))
            assert np.array_equal(a[:, :, kmax - 1], before[:, :, kmax - 1])
        # (b) periodic along i with two ranks: both neighbours are the same rank
        p = torch.from_numpy(_halo_field(rank, MI, MJ, MK))
        halo.exchange_axis(p, 1, imax, jmax, kmax, peer, peer, ct=5, eb=eb, compressed_cls=HostCompressed,
                           compress_fn=compress_fn, write_fn=write_fn)
        a = p.numpy()
        assert np.array_equal(_np_face(a, 1, imax - 1, imax, jmax, kmax).view(np.uint32), expect(peer, 1, 1).view(np.uint32))
        assert np.array_equal(_np_face(a, 1, 0, imax, jmax, kmax).view(np.uint32), expect(peer, 1, imax - 2).view(np.uint32))
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, "FAIL: %s\n%s" % (e, traceback.format_exc())))
    finally:
        dist.destroy_process_group()


def test_halo_exchange_over_gloo_world2():
    import torch.multiprocessing as mp
    ol.build_oracle()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_halo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res


# ---- broadcast protocol of the float64 callers: MPI_Bcast_bitwise_double / _crc / _mask_crc ----------------
@dataclass
class HostCompressedF64:                   # same field order as api.CompressedF64
    ct: int
    n: int
    chunk_elems: int
    abs_error_bound: float
    flags: int
    fixed_type: int
    fixed_mask20: int
    stream: object
    chunk_offsets: object
    chunk_min: object
    chunk_mask20: object
    chunk_type: object
    chunk_medium: object
    chunk_crc: object


def oracle_compressed_f64(x, ct, chunk, eb):
    base = 5 if ct in (5, 8) else (7 if ct in (7, 9) else ct)
    enc = ol.o_compress_chunked_f64(base, x, chunk, eb, want_crc=True)
    pad = np.zeros(((enc["stream"].size + 31) // 16) * 16, np.uint8)
    pad[:enc["stream"].size] = enc["stream"]
    return HostCompressedF64(ct, x.size, chunk, eb, 1, 0, 0, torch.from_numpy(pad),
                             torch.from_numpy(enc["offsets"].astype(np.int64)), torch.from_numpy(enc["chunk_min"]),
                             torch.from_numpy(enc["chunk_mask20"].view(np.int32).copy()), torch.from_numpy(enc["chunk_type"]),
                             None, torch.from_numpy(enc["chunk_crc"].view(np.int32).copy())), enc


def _bcast_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        dev = torch.device("cpu")
        root = 0
        # (a) float64, CT=5, no CRC: MPI_Bcast_bitwise_double (impl/dataCompression.c:165-224); single chunk = the legacy shape
        n, eb = 3000, 1e-4
        xd = np.cumsum(np.random.default_rng(1).normal(0, 0.01, n)) + 3.0
        comp, enc = oracle_compressed_f64(xd, 5, n, eb)
        got, resends = tr.bcast_compressed(comp if rank == root else None, root, n, 0, dev, HostCompressedF64, f64=True)
        assert resends == 0
        total = int(got.chunk_offsets[-1])
        assert np.array_equal(got.stream[:total].numpy(), enc["stream"]) and total == enc["stream"].size
        assert torch.equal(got.chunk_min.view(torch.int64), torch.from_numpy(enc["chunk_min"]).view(torch.int64))
        dec = ol.o_decompress_chunked_f64(5, dict(stream=got.stream[:total].numpy(), offsets=got.chunk_offsets.numpy().astype(np.uint64),
                                                  chunk_min=got.chunk_min.numpy(), chunk_mask20=got.chunk_mask20.numpy().view(np.uint32),
                                                  chunk_type=got.chunk_type.numpy()), n, n, eb)
        assert np.array_equal(dec.view(np.uint64), ol.o_decompress_chunked_f64(5, enc, n, n, eb).view(np.uint64))
        assert np.max(np.abs(dec - xd)) < 0.5                                 # buffer[i] = decompressed + min (:218-221)
        # (b) float64, CT=9 (mask + CRC), chunked: MPI_Bcast_bitwise_mask_crc (:968-1089); rank 1's first copy is corrupted
        n, chunk = 4096 * 3 + 50, 4096
        xd = np.abs(np.sin(np.arange(n) * 1e-3)) * 7.0 + 1.0
        comp, enc = oracle_compressed_f64(xd, 9, chunk, eb)

        def crc_chunks64(stream, offsets):
            return crc_chunks(stream, offsets)

        def corrupt(got, attempt):
            got.stream[int(got.chunk_offsets[1]) + 3] ^= 0x04               # the BER fault model (:1018-1027), second chunk

        got, resends = tr.bcast_compressed(comp if rank == root else None, root, n, chunk, dev, HostCompressedF64, f64=True,
                                           crc_fn=crc_chunks64, corrupt_fn=corrupt if rank == 1 else None)
        assert resends == 1                                                  # root: one rank re-sent; rank 1: received it
        total = int(got.chunk_offsets[-1])
        assert np.array_equal(got.stream[:total].numpy(), enc["stream"])
        assert np.array_equal(got.chunk_mask20.numpy().view(np.uint32), enc["chunk_mask20"])
        assert np.array_equal(got.chunk_type.numpy(), enc["chunk_type"]) and np.array_equal(got.chunk_crc.numpy().view(np.uint32), enc["chunk_crc"])
        # (c) float32 CT=8 frame through the same collective, clean delivery
        x = make_field("noisy", 4096 * 2 + 9)
        comp, enc = oracle_compressed(x, 8, 4096, 0.01)
        got, resends = tr.bcast_compressed(comp if rank == root else None, root, x.size, 4096, dev, HostCompressed, crc_fn=crc_chunks)
        assert resends == 0
        assert np.array_equal(got.stream[:int(got.chunk_offsets[-1])].numpy(), enc["stream"])
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, "FAIL: %s\n%s" % (e, traceback.format_exc())))
    finally:
        dist.destroy_process_group()


def test_bcast_protocol_over_gloo_world2():
    import torch.multiprocessing as mp
    ol.build_oracle()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_bcast_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res
