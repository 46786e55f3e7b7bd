"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref, built from /root/reference by
oracle/build_ref.sh).  Run here (the authoring container); the fixtures travel, the reference does not.

    python tests/golden/make_golden.py

Fixtures:
  datasets.npz   the reference's own float datasets parsed the way its tools parse them
                 (fscanf("%f") until feof, num = n-1; tools/mycompress_bitwise_float.c:29-35):
                 testfloat_8_8_128 (8,192) and float_rand_8192, plus the known-answer table of
                 SURVEY.md section 4 (bytes, pos, crc32, sha256 of stream and of decoded floats).
  bitdump.npz    impl/dataset/bit.txt + bitcomp.txt packed: per-element 32-bit word of the min-subtracted
                 testfloat_8_8_128 and the code R(x) (length, bits) compress_bitwise_float emits at eb=1e-6.
  chunked.npz    the three synthetic fields (n=20000, ragged last chunk) run chunk-by-chunk (C=4096)
                 through the reference for CT=5/6/7/11 at eb=0.01 and 1e-4: offsets, mins, medium/type,
                 crc per chunk, stream, decoded floats.
"""
import ctypes as C
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as ol  # noqa: E402
from data_compression_b200.fields import make_field  # noqa: E402

REF_DATA = "/root/reference/impl/dataset"


def parse_like_fscanf(path):
    libc = C.CDLL(None)
    libc.strtof.restype = C.c_float
    libc.strtof.argtypes = [C.c_char_p, C.c_void_p]
    toks = open(path, "rb").read().split()
    return np.array([libc.strtof(t, None) for t in toks], dtype=np.float32)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    out = {}
    ka = []  # known answers
    for name in ("testfloat_8_8_128", "float_rand_8192"):
        x = parse_like_fscanf(os.path.join(REF_DATA, name + ".txt"))
        out[name] = x
        for eb in (1e-6, 0.01):
            R = ol.RefLib(eb)
            mn, y = R.min_subtract(x)
            med, t = R.med(y, 0)
            m17 = ol.mask17_of(med)
            for ct in ol.ALL_CTS:
                s, pos = R.compress(ct, y, t, m17)
                defined = ct != ol.CT_MASK or ol.ref_mask_decode_defined(t, m17, eb)
                d = R.decompress(ct, s, y.size, t, m17) if defined else None
                ka.append((name, eb, ct, float(mn), float(med), t, m17, s.size, pos, R.crc32(s), sha(s),
                           sha(d) if d is not None else ""))
                out["stream_%s_%g_%d" % (name, eb, ct)] = s
                if d is not None:
                    out["decoded_%s_%g_%d" % (name, eb, ct)] = d
    out["known_answers"] = np.array(
        ka, dtype=[("name", "U32"), ("eb", "f8"), ("ct", "i4"), ("min", "f4"), ("medium", "f4"), ("type", "i4"),
                   ("mask17", "u4"), ("bytes", "i8"), ("pos", "i4"), ("crc32", "u4"), ("sha_stream", "U64"),
                   ("sha_decoded", "U64")])
    np.savez_compressed(os.path.join(HERE, "datasets.npz"), **out)

    # golden bit dumps of the reference itself
    words = np.array([int("".join(l.split()), 2) for l in open(os.path.join(REF_DATA, "bit.txt"))], dtype=np.uint32)
    lens, codes = [], []
    for l in open(os.path.join(REF_DATA, "bitcomp.txt")):
        b = "".join(l.split())
        lens.append(len(b))
        codes.append(int(b, 2) if b else 0)
    np.savez_compressed(os.path.join(HERE, "bitdump.npz"), words=words, code_len=np.array(lens, np.int32),
                        code=np.array(codes, np.uint32))

    ch = {}
    n, chunk = 20000, 4096
    for kind in ("smooth", "noisy", "constant"):
        x = make_field(kind, n)
        ch["input_sha_" + kind] = np.array(sha(x))
        for eb in (0.01, 1e-4):
            R = ol.RefLib(eb)
            for ct in ol.ALL_CTS:
                enc = R.compress_chunked(ct, x, chunk, subtract_min=True, want_crc=True)
                key = "%s_%g_%d_" % (kind, eb, ct)
                for k, v in enc.items():
                    ch[key + k] = v
                dec = np.empty(n, np.float32)
                ok = True
                for c in range(len(enc["offsets"]) - 1):
                    lo, hi = c * chunk, min(n, (c + 1) * chunk)
                    t, m17 = int(enc["chunk_type"][c]), ol.mask17_of(enc["chunk_medium"][c])
                    if ct == ol.CT_MASK and not ol.ref_mask_decode_defined(t, m17, eb):
                        ok = False
                        break
                    s = enc["stream"][int(enc["offsets"][c]):int(enc["offsets"][c + 1])]
                    dec[lo:hi] = R.decompress(ct, s, hi - lo, t, m17) + enc["chunk_min"][c]
                if ok:
                    ch[key + "decoded"] = dec
    np.savez_compressed(os.path.join(HERE, "chunked.npz"), **ch)
    for f in ("datasets.npz", "bitdump.npz", "chunked.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
