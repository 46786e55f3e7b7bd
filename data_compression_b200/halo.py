"""Himeno halo exchange with compressed faces, device resident (BASELINE.json configs[3]; the shape of
impl/himenoBMTxps.c sendp1/2/3, e.g. :580-650 for CT=5 k-faces): gather the two inner faces next to the halo planes,
min-subtract + compress each as ONE reference call (single chunk), move the frames with send/recv (NCCL on the box),
decompress, add the sender's min and write the halo planes.  Absent neighbours (MPI_PROC_NULL) are skipped."""
from . import api, transport as tr

AXIS_DIM = {1: 0, 2: 1, 3: 2}
_FRAMES, _INBUFS = {}, {}            # exchange_axis_fixed: frame workspaces / receive buffers, reused from call to call


def compress_face(p, ijk, v, imax, jmax, kmax, ct, eb):
    """-> api.Compressed of the face (one chunk == one myCompress_bitwise* call after toSmallDataset_float)."""
    face = api.gather_face(p, ijk, v, imax, jmax, kmax)
    return api.compress(face, ct=ct, abs_error_bound=eb, chunk_elems=0, subtract_min=True)


def write_face(p, ijk, v, imax, jmax, kmax, comp):
    vals = api.decompress(comp, check_status=False)         # already decompressed[i] + min (same float add as the driver's loop);
                                                      # no status read-back: a halo hop must not synchronise the host
    api.scatter_face(p, ijk, v, imax, jmax, kmax, vals, 0.0)


def exchange_axis(p, ijk, imax, jmax, kmax, lo_peer, hi_peer, ct=5, eb=0.01, compressed_cls=None,
                  compress_fn=None, write_fn=None):
    """One sendpN: my face at index 1 goes to lo_peer (it becomes THEIR halo plane max-1), my face at max-2 goes to hi_peer
    (their plane 0); hi_peer's face 1 arrives for my plane max-1 and lo_peer's face max-2 for my plane 0.  lo_peer / hi_peer =
    rank or None (no neighbour: nothing is sent, the plane keeps its value).  As in the reference (himenoBMTxps.c:588-626:
    Irecv/Isend of the byte counts + Waitall, then of min + payload + Waitall) the transfer is two batched rounds: the
    fixed-size tables, then the streams whose sizes the tables announce.  Sends are posted (lo, hi), receives (hi, lo): if both
    neighbours are the same rank (two ranks, periodic) my lo-send still pairs with its hi-receive.
    Returns the compressed bytes sent (the driver's `cr` statistic)."""
    import torch
    import torch.distributed as dist
    cls = compressed_cls or api.Compressed
    compress_fn = compress_fn or compress_face
    write_fn = write_fn or write_face
    ext = {1: imax, 2: jmax, 3: kmax}[ijk]
    n = {1: jmax * kmax, 2: imax * kmax, 3: imax * jmax}[ijk]
    dev = p.device
    out_lo = tr.pack_frame(compress_fn(p, ijk, 1, imax, jmax, kmax, ct, eb)) if lo_peer is not None else None
    out_hi = tr.pack_frame(compress_fn(p, ijk, ext - 2, imax, jmax, kmax, ct, eb)) if hi_peer is not None else None
    tw = tr.table_words(1)
    in_hi_t = torch.empty(tw, dtype=torch.int64, device=dev) if hi_peer is not None else None
    in_lo_t = torch.empty(tw, dtype=torch.int64, device=dev) if lo_peer is not None else None
    ops = []
    if out_lo is not None: ops.append(dist.P2POp(dist.isend, out_lo.table, lo_peer))
    if out_hi is not None: ops.append(dist.P2POp(dist.isend, out_hi.table, hi_peer))
    if in_hi_t is not None: ops.append(dist.P2POp(dist.irecv, in_hi_t, hi_peer))
    if in_lo_t is not None: ops.append(dist.P2POp(dist.irecv, in_lo_t, lo_peer))
    if not ops:
        return 0
    for w in dist.batch_isend_irecv(ops):
        w.wait()
    cap = 4 * n + 16

    def stream_buf(t):
        total = int(t[tr.HEADER_WORDS - 1].item())
        if total > cap:
            raise ValueError("frame announces %d bytes for %d values" % (total, n))
        return torch.zeros(((cap + 15) // 16) * 16, dtype=torch.uint8, device=dev), total

    in_hi_s = stream_buf(in_hi_t) if in_hi_t is not None else None
    in_lo_s = stream_buf(in_lo_t) if in_lo_t is not None else None
    ops = []
    if out_lo is not None and out_lo.stream.numel(): ops.append(dist.P2POp(dist.isend, out_lo.stream.contiguous(), lo_peer))
    if out_hi is not None and out_hi.stream.numel(): ops.append(dist.P2POp(dist.isend, out_hi.stream.contiguous(), hi_peer))
    if in_hi_s is not None and in_hi_s[1]: ops.append(dist.P2POp(dist.irecv, in_hi_s[0][:in_hi_s[1]], hi_peer))
    if in_lo_s is not None and in_lo_s[1]: ops.append(dist.P2POp(dist.irecv, in_lo_s[0][:in_lo_s[1]], lo_peer))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    if in_hi_t is not None:
        write_fn(p, ijk, ext - 1, imax, jmax, kmax, tr.unpack_frame(tr.Frame(in_hi_t, in_hi_s[0]), cls))
    if in_lo_t is not None:
        write_fn(p, ijk, 0, imax, jmax, kmax, tr.unpack_frame(tr.Frame(in_lo_t, in_lo_s[0]), cls))
    return (out_lo.stream.numel() if out_lo is not None else 0) + (out_hi.stream.numel() if out_hi is not None else 0)


def exchange_axis_fixed(p, ijk, imax, jmax, kmax, lo_peer, hi_peer, ct=5, eb=0.01, chunk_elems=0, compressed_cls=None):
    """exchange_axis with fixed-capacity frames (transport.pack_fixed): ONE batched isend/irecv round per axis and no host
    round trip -- the announced stream length stays on the device.  chunk_elems = 0 compresses a face as ONE reference call
    (the shape of impl/himenoBMTxps.c:644-708); chunk_elems = 4096 makes every 4096 values of the face a reference call of
    their own, which lets the receiver decode the chunks side by side (a face of predicted codes only is otherwise one dependent
    chain).  Returns the frame bytes sent."""
    import torch
    import torch.distributed as dist
    cls = compressed_cls or api.Compressed
    ext = {1: imax, 2: jmax, 3: kmax}[ijk]
    n = {1: jmax * kmax, 2: imax * kmax, 3: imax * jmax}[ijk]
    dev = p.device
    fb = tr.fixed_frame_bytes(n, chunk_elems)

    def comp_face(v, slot):
        # the codec writes table and stream straight into the frame (transport.FrameWorkspace): nothing to pack
        key = (n, chunk_elems, ct, eb, str(dev), slot)
        ws = _FRAMES.get(key)
        if ws is None:
            ws = _FRAMES[key] = tr.FrameWorkspace(n, chunk_elems, dev, ct=ct, abs_error_bound=eb)
        face = api.gather_face(p, ijk, v, imax, jmax, kmax)
        api.compress(face, ct=ct, abs_error_bound=eb, chunk_elems=chunk_elems, subtract_min=True, workspace=ws)
        return ws.frame

    def in_buf(slot):
        key = (fb, str(dev), slot)
        b = _INBUFS.get(key)
        if b is None:
            b = _INBUFS[key] = torch.empty(fb, dtype=torch.uint8, device=dev)
        return b

    out_lo = comp_face(1, 0) if lo_peer is not None else None
    out_hi = comp_face(ext - 2, 1) if hi_peer is not None else None
    in_hi = in_buf(0) if hi_peer is not None else None
    in_lo = in_buf(1) if lo_peer is not None else None
    ops = []
    if out_lo is not None: ops.append(dist.P2POp(dist.isend, out_lo, lo_peer))
    if out_hi is not None: ops.append(dist.P2POp(dist.isend, out_hi, hi_peer))
    if in_hi is not None: ops.append(dist.P2POp(dist.irecv, in_hi, hi_peer))
    if in_lo is not None: ops.append(dist.P2POp(dist.irecv, in_lo, lo_peer))
    if not ops:
        return 0
    for w in dist.batch_isend_irecv(ops):
        w.wait()
    flags = api.F_SUBTRACT_MIN
    if in_hi is not None:
        write_face(p, ijk, ext - 1, imax, jmax, kmax, tr.unpack_fixed(in_hi, n, chunk_elems, ct, eb, flags, cls))
    if in_lo is not None:
        write_face(p, ijk, 0, imax, jmax, kmax, tr.unpack_fixed(in_lo, n, chunk_elems, ct, eb, flags, cls))
    return fb * ((out_lo is not None) + (out_hi is not None))


def exchange_axis_raw(p, ijk, imax, jmax, kmax, lo_peer, hi_peer):
    """CT=0: the uncompressed faces (the reference's MPI vector types, impl/himenoBMTxps.c sendp1/2/3 default branch) -- one
    batched round of float32 faces.  Returns the bytes sent."""
    import torch
    import torch.distributed as dist
    ext = {1: imax, 2: jmax, 3: kmax}[ijk]
    n = {1: jmax * kmax, 2: imax * kmax, 3: imax * jmax}[ijk]
    dev = p.device
    out_lo = api.gather_face(p, ijk, 1, imax, jmax, kmax) if lo_peer is not None else None
    out_hi = api.gather_face(p, ijk, ext - 2, imax, jmax, kmax) if hi_peer is not None else None
    in_hi = torch.empty(n, dtype=torch.float32, device=dev) if hi_peer is not None else None
    in_lo = torch.empty(n, dtype=torch.float32, device=dev) if lo_peer is not None else None
    ops = []
    if out_lo is not None: ops.append(dist.P2POp(dist.isend, out_lo, lo_peer))
    if out_hi is not None: ops.append(dist.P2POp(dist.isend, out_hi, hi_peer))
    if in_hi is not None: ops.append(dist.P2POp(dist.irecv, in_hi, hi_peer))
    if in_lo is not None: ops.append(dist.P2POp(dist.irecv, in_lo, lo_peer))
    if not ops:
        return 0
    for w in dist.batch_isend_irecv(ops):
        w.wait()
    if in_hi is not None:
        api.scatter_face(p, ijk, ext - 1, imax, jmax, kmax, in_hi, 0.0)
    if in_lo is not None:
        api.scatter_face(p, ijk, 0, imax, jmax, kmax, in_lo, 0.0)
    return 4 * n * ((out_lo is not None) + (out_hi is not None))
