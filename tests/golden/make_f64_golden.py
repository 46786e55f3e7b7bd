#!/usr/bin/env python
"""Generates tests/golden/f64_golden.npz from the UNMODIFIED reference (oracle/_ref, built from /root/reference by
oracle/build_ref.sh): min-subtracted float64 inputs, the reference's own type/mask, its streams and -- where its decoder is
defined -- its decoded values.  Run here (the GPU box has no /root/reference); the fixture travels."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol          # noqa: E402


def main():
    rng = np.random.default_rng(20261104)
    n = 700
    fields = {
        "walk": np.cumsum(rng.standard_normal(n)) * 0.02 + 20.0,
        "uniform": rng.random(n) * 10.0,
        "sine": 1 + 0.5 * np.sin(np.arange(n) / 40.0),
        "centres": np.sort(rng.random(n) * 1000.0),
    }
    out = {}
    for eb in (0.01, 1e-4, 1e-6):
        R = ol.RefLibF64(eb)
        for name, x in fields.items():
            _, y = R.min_subtract(x)
            med, t = R.med(y, 0)
            m20 = ol.mask20_of(med)
            for ct in ol.ALL_CTS:
                tag = "%s_eb%s_ct%d" % (name, ol.eb_str(eb), ct)
                s, _ = R.compress(ct, y, t, m20)
                out[tag + "_x"], out[tag + "_eb"], out[tag + "_ct"] = y, np.float64(eb), np.int32(ct)
                out[tag + "_type"], out[tag + "_mask20"], out[tag + "_stream"] = np.int32(t), np.int64(m20), s
                if ct != ol.CT_MASK or ol.ref_mask_decode_defined_f64(t, m20, eb):
                    out[tag + "_decoded"] = R.decompress(ct, s, y.size, t, m20)
    np.savez_compressed(os.path.join(HERE, "f64_golden.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
