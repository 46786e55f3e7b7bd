"""CPU-side checks of the C ABI: the library builds for sm_100a, loads, exports every symbol include/*.h declares,
and fails loudly (never silently falls back) without a GPU.  No compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from data_compression_b200 import api, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def L():
    build.build()
    return api.lib()


def declared_symbols():
    names = set()
    for h in os.listdir(os.path.join(ROOT, "include")):
        src = open(os.path.join(ROOT, "include", h)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        for m in re.finditer(r"^[A-Za-z_][\w \*]*?[\s\*]([A-Za-z_]\w*)\s*\(", src, flags=re.M):
            if m.group(1) not in ("defined",):
                names.add(m.group(1))
    return names


def test_exports_every_declared_symbol(L):
    names = declared_symbols()
    assert {"dcb_compress", "dcb_decompress", "dcb_compress_host", "dcb_crc32_chunks"} <= names
    missing = [s for s in sorted(names) if not hasattr(L, s)]
    assert not missing, missing


def test_host_logic(L):
    # to_absErrorBound_binary (dataCompression.c:5512): SURVEY.md 3.5 table
    for eb, n in ((0.1, 4), (0.01, 7), (1e-3, 10), (1e-4, 14), (1e-5, 17), (1e-6, 20), (1.0, 0), (2.0 ** -99, 99)):
        assert L.dcb_n_eb(eb) == n
    assert L.dcb_n_eb(2.0 ** -100) == -1
    assert L.dcb_num_chunks(0, 4096) == 0
    assert L.dcb_num_chunks(20000, 4096) == 5
    assert L.dcb_num_chunks(20000, 0) == 1
    assert L.dcb_num_chunks(100, 4096) == 1
    assert L.dcb_max_stream_bytes(1000, 0) >= 4000
    assert L.dcb_workspace_bytes(1 << 20, 4096) >= 8 * 256


def test_host_logic_of_the_widened_rows(L):
    # hmLength (dataCompression.c:5587): smallest r with 2^r - 1 - r - k >= 0
    assert [L.dcb_hm_length(k) for k in (1, 4, 11, 12, 1000000)] == [2, 3, 4, 5, 20]
    # slot path: one 16 KB slot per chunk on top of the tables; multi-tile chunks keep the small workspace
    n = 1 << 20
    assert L.dcb_workspace_bytes(n, 4096) >= (n // 4096) * 16384
    assert L.dcb_workspace_bytes(n, 8192) < 1 << 16
    assert L.dcb_bytewise_workspace_bytes(n, 4096) >= n


def test_no_silent_cpu_fallback(L):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    assert L.dcb_device_count() == 0
    x = np.zeros(16, np.float32)
    with pytest.raises(api.DcbError, match="ENODEVICE"):
        api.compress_host(x, ct=5, abs_error_bound=0.01)
    with pytest.raises(api.DcbError):
        api.compress(torch.zeros(16), ct=5)
    buf = np.zeros(64, np.uint8)
    chk = np.zeros(1, np.uint32)
    assert L.dcb_hamming_encode(buf.ctypes.data, 64, 64, chk.ctypes.data, None, None) == -4          # DCB_ENODEVICE
    assert L.dcb_bytewise_decompress(buf.ctypes.data, buf.ctypes.data, buf.ctypes.data, chk.ctypes.data, 4, 0, buf.ctypes.data, None) == -4


def test_product_does_not_import_oracle():
    """The oracle is test infrastructure: nothing under data_compression_b200/ may reference it."""
    pkg = os.path.join(ROOT, "data_compression_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".c")):
                src = open(os.path.join(dp, f)).read()
                assert "oracle_lib" not in src and "dc_oracle" not in src and "oracle/" not in src, f
