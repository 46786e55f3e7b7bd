"""Moving compressed chunks between GPUs: chunk-range sharding and the send/recv frames that replace the
MPI messages of the reference's drivers.

* shard_range(): SURVEY.md 8(e) -- chunks are independent and byte aligned, so rank g of G takes the contiguous chunk
  range [g*K/G, (g+1)*K/G) and compresses it with no data-path collective.
* send_compressed() / recv_compressed(): the pingpong hop (impl/pingpong.c:250 MPI_Send / :353 MPI_Recv of
  data_bits) and the Himeno halo hop (impl/himenoBMTxps.c:667-693: the byte count first, then min + stream).  NCCL
  needs matching counts on both ends, so a hop is two messages exactly like the reference's: a fixed-size header
  with the chunk table (size known from n and chunk_elems), then the stream bytes the header announced.
* pingpong_hop_crc(): the CT=8/9 protocol of impl/pingpong.c:260-289 (sender) and :363-447 (receiver): stream,
  CRC, one-byte ack 'y'/'n', resend on 'n'.
* bcast_compressed(): the collective wrappers of the float64 callers -- MPI_Bcast_bitwise_double
  (impl/dataCompression.c:165-224: one padded message [int bytes][double min][payload]) and MPI_Bcast_bitwise{,_mask}_crc
  (:968-1089, :1092-1197: bytes, min, payload, medium, type, crc broadcast; 'y'/'n' gathered at the root; the root re-sends
  to every 'n' rank).  Here: the fixed-size table broadcast, the stream broadcast at its announced length (no padding to the
  worst case), acks by all_gather (NCCL has no gather), point-to-point resends.  Frames carry float32 (api.Compressed) or
  float64 (api.CompressedF64: double min, 20-bit masks) chunk tables.

torch.distributed is the plumbing (backend "nccl" over NVLink on the GPU box, "gloo" in the CPU tests); the payload is
whatever device the tensors live on.  Nothing here encodes or decodes.
"""
from dataclasses import dataclass
from typing import Callable, Optional

import numpy as np

HEADER_WORDS = 8          # int64 words: magic, ct, n, chunk_elems, abs_error_bound (bits), flags|fixed_type<<8, fixed_mask17, total_bytes
MAGIC = 0x44434232303042  # "DCB200B"


def shard_range(n, chunk_elems, rank, world):
    """-> (first_element, end_element) of rank's contiguous chunk range; ragged last chunk goes to the last rank."""
    if chunk_elems <= 0 or chunk_elems >= n:
        return (0, n) if rank == 0 else (n, n)          # a single chunk does not shard
    K = (n + chunk_elems - 1) // chunk_elems
    c0, c1 = rank * K // world, (rank + 1) * K // world
    return min(n, c0 * chunk_elems), min(n, c1 * chunk_elems)


def num_chunks(n, chunk_elems):
    if n == 0:
        return 0
    return 1 if (chunk_elems <= 0 or chunk_elems >= n) else (n + chunk_elems - 1) // chunk_elems


F64_FLAG = 1 << 16        # header word 5: the frame carries a float64 chunk table


def table_words(K, f64=False):
    """int64 words of the header message: fixed header + offsets[K+1] + the chunk table --
    float32: (min, medium, type, crc)[K] as 4K x 32 bit; float64: min[K] as 64 bit, then (mask20, type, crc, 0)[K] x 32 bit."""
    return HEADER_WORDS + (K + 1) + (3 * K if f64 else 2 * K)


def _is_f64(comp):
    return hasattr(comp, "chunk_mask20")


@dataclass
class Frame:
    """One compressed array on the wire: header+table message and stream message (both torch tensors)."""
    table: "object"      # int64 [table_words(K)]
    stream: "object"     # uint8 [total_bytes]


def pack_frame(comp):
    """comp: api.Compressed (or anything with the same fields) -> Frame on comp's device."""
    import torch
    K = comp.chunk_offsets.numel() - 1
    dev = comp.stream.device
    f64 = _is_f64(comp)
    t = torch.zeros(table_words(K, f64), dtype=torch.int64, device=dev)
    eb_bits = int(np.array([comp.abs_error_bound], dtype=np.float64).view(np.int64)[0])
    total = int(comp.chunk_offsets[-1].item()) if K else 0
    hdr = [MAGIC, comp.ct, comp.n, comp.chunk_elems, eb_bits,
           (comp.flags & 0xFF) | ((comp.fixed_type & 0xFF) << 8) | (F64_FLAG if f64 else 0),
           comp.fixed_mask20 if f64 else comp.fixed_mask17, total]
    t[:HEADER_WORDS] = torch.tensor(hdr, dtype=torch.int64, device=dev)
    t[HEADER_WORDS:HEADER_WORDS + K + 1] = comp.chunk_offsets
    if K and not f64:
        side = torch.zeros(4 * K, dtype=torch.int32, device=dev)
        for i, f in enumerate((comp.chunk_min, comp.chunk_medium, comp.chunk_type, comp.chunk_crc)):
            if f is not None:
                side[i * K:(i + 1) * K] = f[:K].view(torch.int32)
        t[HEADER_WORDS + K + 1:] = side.view(torch.int64)
    elif K:
        base = HEADER_WORDS + K + 1
        if comp.chunk_min is not None:
            t[base:base + K] = comp.chunk_min[:K].view(torch.int64)
        side = torch.zeros(4 * K, dtype=torch.int32, device=dev)
        for i, f in enumerate((comp.chunk_mask20, comp.chunk_type, comp.chunk_crc)):
            if f is not None:
                side[i * K:(i + 1) * K] = f[:K].view(torch.int32)
        t[base + K:] = side.view(torch.int64)
    return Frame(t, comp.stream[:total])


def unpack_frame(frame, compressed_cls):
    """Inverse of pack_frame -> compressed_cls(ct, n, chunk_elems, eb, flags, fixed_type, fixed_mask17, stream, offsets,
    chunk_min, chunk_medium, chunk_type, chunk_crc)."""
    import torch
    t = frame.table
    hdr = [int(v) for v in t[:HEADER_WORDS].tolist()]
    if hdr[0] != MAGIC:
        raise ValueError("not a dcb200 frame")
    ct, n, chunk = hdr[1], hdr[2], hdr[3]
    eb = float(np.array([hdr[4]], dtype=np.int64).view(np.float64)[0])
    flags, ftype, fmask = hdr[5] & 0xFF, (hdr[5] >> 8) & 0xFF, hdr[6]
    K = num_chunks(n, chunk)
    offs = t[HEADER_WORDS:HEADER_WORDS + K + 1].clone()
    if hdr[5] & F64_FLAG:
        # api.CompressedF64(ct, n, chunk_elems, eb, flags, fixed_type, fixed_mask20, stream, offsets, chunk_min, chunk_mask20,
        #                   chunk_type, chunk_medium, chunk_crc)
        base = HEADER_WORDS + K + 1
        cmin = t[base:base + K].clone().view(torch.float64) if K else None
        side = t[base + K:].contiguous().view(torch.int32) if K else torch.zeros(0, dtype=torch.int32)
        return compressed_cls(ct, n, chunk, eb, flags, ftype, fmask, frame.stream, offs, cmin,
                              side[0:K] if K else None, side[K:2 * K] if K else None, None, side[2 * K:3 * K] if K else None)
    side = t[HEADER_WORDS + K + 1:].contiguous().view(torch.int32) if K else torch.zeros(0, dtype=torch.int32)
    cmin = side[0:K].view(torch.float32) if K else None
    cmed = side[K:2 * K].view(torch.float32) if K else None
    ctype = side[2 * K:3 * K] if K else None
    ccrc = side[3 * K:4 * K] if K else None
    return compressed_cls(ct, n, chunk, eb, flags, ftype, fmask, frame.stream, offs, cmin, cmed, ctype, ccrc)


def send_compressed(comp, dst, group=None):
    """Two messages, as impl/himenoBMTxps.c:667-693 sends `bytes` first: table (fixed size), then the stream."""
    import torch.distributed as dist
    f = pack_frame(comp)
    dist.send(f.table, dst, group=group)
    if f.stream.numel():
        dist.send(f.stream.contiguous(), dst, group=group)
    return f


def recv_compressed(n, chunk_elems, src, device, compressed_cls, group=None, stream_buffer=None, f64=False):
    """Posts the fixed-size table receive, reads total_bytes from it, then receives exactly that many stream bytes.
    f64: the frame carries an api.CompressedF64 (3K table words instead of 2K, codes of up to 64 bits)."""
    import torch
    import torch.distributed as dist
    K = num_chunks(n, chunk_elems)
    t = torch.empty(table_words(K, f64), dtype=torch.int64, device=device)
    dist.recv(t, src, group=group)
    hdr_n, hdr_chunk = int(t[2].item()), int(t[3].item())
    if (hdr_n, hdr_chunk) != (n, chunk_elems):
        raise ValueError("frame is for n=%d chunk=%d, expected n=%d chunk=%d" % (hdr_n, hdr_chunk, n, chunk_elems))
    total = int(t[HEADER_WORDS - 1].item())
    cap = (8 if f64 else 4) * n + 16                            # worst case of the stream (every code <= 32 / 64 bits)
    if total > cap:
        raise ValueError("frame announces %d bytes for %d values" % (total, n))
    buf = stream_buffer if stream_buffer is not None else torch.zeros(((cap + 15) // 16) * 16, dtype=torch.uint8, device=device)
    if total:
        dist.recv(buf[:total], src, group=group)
    return unpack_frame(Frame(t, buf), compressed_cls)


def pingpong_hop_crc(comp, peer, is_sender, n, chunk_elems, device, compressed_cls,
                     crc_fn: Callable, corrupt_fn: Optional[Callable] = None, max_resend=8, group=None, hamming=None):
    """CT=8/9 hop.  Sender (impl/pingpong.c:260-289): send stream + per-chunk CRC (they travel in the table), wait for
    the one-byte ack, resend on 'n'.  Receiver (:363-447): [test hook: corrupt the received stream, the reference's
    bit_flip], recompute the CRC of what arrived, compare, ack.  Returns (Compressed the receiver holds | None, resends).
    CT=10 (impl/pingpong.c:290-326, :448-510): hamming = (block_bytes, encode_fn, decode_fn).  The sender also sends the SECDED
    check words of every block of the stream (encode_fn(stream, total, block_bytes) -> int32[blocks]); the receiver runs
    decode_fn(stream, total, block_bytes, check) on what arrived -- single-bit errors are rectified in place -- BEFORE the CRC
    comparison, so a one-bit fault per block costs no resend."""
    import torch
    import torch.distributed as dist
    resend = 0
    ack = torch.zeros(1, dtype=torch.uint8, device=device)
    while True:
        if is_sender:
            f = send_compressed(comp, peer, group=group)
            if hamming is not None and f.stream.numel():
                chk = hamming[1](f.stream, f.stream.numel(), hamming[0])
                dist.send(chk.contiguous(), peer, group=group)
            dist.recv(ack, peer, group=group)
            if ack.item() == ord("y") or resend >= max_resend:
                return None, resend
            resend += 1
        else:
            got = recv_compressed(n, chunk_elems, peer, device, compressed_cls, group=group)
            if corrupt_fn is not None:
                corrupt_fn(got, resend)
            K = got.chunk_offsets.numel() - 1
            total = int(got.chunk_offsets[-1].item()) if K else 0
            if hamming is not None and total:
                nb = (total + hamming[0] - 1) // hamming[0]
                chk = torch.zeros(nb, dtype=torch.int32, device=device)
                dist.recv(chk, peer, group=group)
                hamming[2](got.stream, total, hamming[0], chk)
            ok = bool((crc_fn(got.stream, got.chunk_offsets)[:K] == got.chunk_crc[:K]).all().item())
            ack[0] = ord("y") if ok else ord("n")
            dist.send(ack, peer, group=group)
            if ok or resend >= max_resend:
                return got, resend
            resend += 1


def bcast_compressed(comp, root, n, chunk_elems, device, compressed_cls, f64=False, crc_fn: Optional[Callable] = None,
                     corrupt_fn: Optional[Callable] = None, group=None):
    """MPI_Bcast_bitwise_double / MPI_Bcast_bitwise{,_mask}_crc (impl/dataCompression.c:165-224, :968-1089, :1092-1197) on
    torch.distributed.  `comp` is the root's compressed array (ignored elsewhere).  Every rank returns (the frame it
    holds, resends): with crc_fn every non-root recomputes the per-chunk CRCs of what arrived [corrupt_fn: the reference's
    BER fault model, called as corrupt_fn(frame, attempt) before the check], the 'y'/'n' acks are gathered, and the root
    re-sends the stream point to point to every 'n' rank (once, as the reference does; the re-sent copy is not checked again)."""
    import torch
    import torch.distributed as dist
    rank = dist.get_rank(group)
    world = dist.get_world_size(group)
    K = num_chunks(n, chunk_elems)
    elem = 8 if f64 else 4
    if rank == root:
        f = pack_frame(comp)
        t, stream = f.table, f.stream.contiguous()
    else:
        t = torch.empty(table_words(K, f64), dtype=torch.int64, device=device)
    dist.broadcast(t, root, group=group)                       # data_bytes, min, medium, type, crc: one fixed-size message
    total = int(t[HEADER_WORDS - 1].item())
    cap = elem * n + 16
    if total > cap:
        raise ValueError("frame announces %d bytes for %d values" % (total, n))
    if rank != root:
        stream = torch.zeros(((total + 15) // 16) * 16 + 16, dtype=torch.uint8, device=device)
    if total:
        dist.broadcast(stream[:total], root, group=group)      # the payload at its real length (:1007)
    got = comp if rank == root else unpack_frame(Frame(t, stream), compressed_cls)
    resends = 0
    if crc_fn is not None:
        ok = True
        if rank != root:
            if corrupt_fn is not None:
                corrupt_fn(got, 0)
            ok = bool((crc_fn(got.stream, got.chunk_offsets)[:K] == got.chunk_crc[:K]).all().item())
        ack = torch.tensor([ord("y") if ok else ord("n")], dtype=torch.uint8, device=device)
        acks = [torch.zeros(1, dtype=torch.uint8, device=device) for _ in range(world)]
        dist.all_gather(acks, ack, group=group)                # MPI_Gather of crc_ok (:1049)
        bad = [r for r in range(world) if r != root and int(acks[r].item()) == ord("n")]
        if rank == root:
            for r in bad:                                      # :1053-1060
                if total:
                    dist.send(stream[:total], r, group=group)
                resends += 1
        elif rank in bad:
            if total:
                dist.recv(got.stream[:total], root, group=group)   # :1062-1065
            resends = 1
    return got, resends


# ---- fixed-capacity frames: one message per hop, no host round trip ------------------------------------------------------
# The two-message hop above reads the announced stream length on the host (a device synchronisation per hop), which is what a
# hop of a 66-136 KB Himeno face or of the 131,072-float pingpong array costs most over NVLink.  Both ends know n and
# chunk_elems, hence the worst-case frame size: [table int64 words][stream capacity 4n + 16 bytes] goes out as ONE buffer of
# that size.  The length travels inside the table (word 7) and stays on the device; the unused tail of
# the stream region is never read by the decoder.  Meant for latency-bound payloads (default: frames up to 4 MiB); a large
# array keeps the two-message hop, whose second message has the compressed length.
FIXED_FRAME_MAX_BYTES = 4 << 20


def _fixed_table_bytes(K):
    return ((8 * table_words(K) + 15) // 16) * 16              # the stream region starts 16-byte aligned (dcb_decompress needs it)


def fixed_frame_bytes(n, chunk_elems):
    K = num_chunks(n, chunk_elems)
    return _fixed_table_bytes(K) + ((4 * n + 16 + 15) // 16) * 16


def use_fixed_frame(n, chunk_elems):
    return fixed_frame_bytes(n, chunk_elems) <= FIXED_FRAME_MAX_BYTES


def pack_fixed(comp):
    """comp (float32 api.Compressed) -> one uint8 tensor of fixed_frame_bytes(n, chunk_elems); no host synchronisation."""
    import torch
    K = comp.chunk_offsets.numel() - 1
    dev = comp.stream.device
    tw = table_words(K)
    buf = torch.empty(fixed_frame_bytes(comp.n, comp.chunk_elems), dtype=torch.uint8, device=dev)
    t = buf[:8 * tw].view(torch.int64)
    eb_bits = int(np.array([comp.abs_error_bound], dtype=np.float64).view(np.int64)[0])
    hdr = [MAGIC, comp.ct, comp.n, comp.chunk_elems, eb_bits, (comp.flags & 0xFF) | ((comp.fixed_type & 0xFF) << 8),
           comp.fixed_mask17, 0]
    t[:HEADER_WORDS] = torch.tensor(hdr, dtype=torch.int64).to(dev, non_blocking=True)
    t[HEADER_WORDS:HEADER_WORDS + K + 1] = comp.chunk_offsets
    t[HEADER_WORDS - 1:HEADER_WORDS] = comp.chunk_offsets[K:K + 1]          # total bytes, device to device
    side = t[HEADER_WORDS + K + 1:].view(torch.int32)
    side.zero_()
    for i, f in enumerate((comp.chunk_min, comp.chunk_medium, comp.chunk_type, comp.chunk_crc)):
        if f is not None:
            side[i * K:(i + 1) * K] = f[:K].view(torch.int32)
    tb = _fixed_table_bytes(K)
    cap = min(buf.numel() - tb, comp.stream.numel())
    buf[tb:tb + cap] = comp.stream[:cap]
    return buf


def unpack_fixed(buf, n, chunk_elems, ct, abs_error_bound, flags, compressed_cls, fixed_type=0, fixed_mask17=0):
    """Views into a received fixed frame.  The codec parameters are the receiver's own (both ends of a pingpong / halo hop
    run the same configuration): nothing is read back to the host."""
    import torch
    K = num_chunks(n, chunk_elems)
    tw = table_words(K)
    t = buf[:8 * tw].view(torch.int64)
    offs = t[HEADER_WORDS:HEADER_WORDS + K + 1]
    side = t[HEADER_WORDS + K + 1:].view(torch.int32)
    per_chunk_mask = ct in (7, 9) and not (flags & 2)
    has_crc = bool(flags & 4) or ct in (8, 9)
    return compressed_cls(ct, n, chunk_elems, abs_error_bound, flags, fixed_type, fixed_mask17, buf[_fixed_table_bytes(K):], offs,
                          side[0:K].view(torch.float32) if (flags & 1) else None,
                          side[K:2 * K].view(torch.float32) if per_chunk_mask else None,
                          side[2 * K:3 * K] if per_chunk_mask else None,
                          side[3 * K:4 * K] if has_crc else None)


def send_fixed(comp, dst, group=None):
    import torch.distributed as dist
    buf = pack_fixed(comp)
    dist.send(buf, dst, group=group)
    return buf


def recv_fixed(n, chunk_elems, src, device, ct, abs_error_bound, flags, compressed_cls, group=None, buffer=None):
    import torch
    import torch.distributed as dist
    buf = buffer if buffer is not None else torch.empty(fixed_frame_bytes(n, chunk_elems), dtype=torch.uint8, device=device)
    dist.recv(buf, src, group=group)
    return unpack_fixed(buf, n, chunk_elems, ct, abs_error_bound, flags, compressed_cls)


class FrameWorkspace:
    """An api.Workspace whose tensors are views of ONE fixed-capacity frame buffer: dcb_compress writes the chunk table and the
    stream where the frame wants them, so a hop is `dist.send(ws.frame)` with no packing step at all (pack_fixed costs about
    ten small device ops).  Same field names as api.Workspace; pass it as `workspace=` to api.compress."""

    def __init__(self, n, chunk_elems, device, ct=5, abs_error_bound=0.01, flags=1):
        import torch
        from . import api
        L = api.lib()
        self.n, self.chunk_elems, self.device = n, chunk_elems, device
        K = self.K = L.dcb_num_chunks(n, chunk_elems)
        tw, tb = table_words(K), _fixed_table_bytes(K)
        self.frame = torch.zeros(fixed_frame_bytes(n, chunk_elems), dtype=torch.uint8, device=device)
        t = self.frame[:8 * tw].view(torch.int64)
        eb_bits = int(np.array([abs_error_bound], dtype=np.float64).view(np.int64)[0])
        t[:HEADER_WORDS] = torch.tensor([MAGIC, ct, n, chunk_elems, eb_bits, flags & 0xFF, 0, 0], dtype=torch.int64).to(device)
        self.offsets = t[HEADER_WORDS:HEADER_WORDS + K + 1]
        side = t[HEADER_WORDS + K + 1:].view(torch.int32)
        self.cmin = side[0:K].view(torch.float32)
        self.cmed = side[K:2 * K].view(torch.float32)
        self.ctype = side[2 * K:3 * K]
        self.ccrc = side[3 * K:4 * K]
        self.stream = self.frame[tb:]
        self.capacity = self.stream.numel()
        assert self.capacity >= L.dcb_max_stream_bytes(n, chunk_elems)
        self.ws_bytes = L.dcb_workspace_bytes(n, chunk_elems)
        self.ws = torch.empty(self.ws_bytes, dtype=torch.uint8, device=device)
        self.status = torch.zeros(1, dtype=torch.int32, device=device)
