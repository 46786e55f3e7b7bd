"""The link-time drop-in of INTEGRATION.md section 2, executed: the reference's own drivers impl/pingpong.c and
impl/himenoBMTxps.c are compiled where they lie, the 14 hot symbols of the reference object are localized with objcopy and the
drivers are linked against libdcb200.so (oracle/dropin/build_dropin.sh; a single-process MPI stand-in replaces MPI, which this
image does not have).  CPU part: the link succeeds, the hot symbols bind to the library, and the stand-in is validated by the
UNMODIFIED reference (driver + reference object) printing the known ratios of impl/pingpong.csv:36-38.  GPU part: the drop-in
executable prints the same ratios."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, "oracle", "_ref", "dropin")
HOT = ("myCompress_bitwise myDecompress_bitwise myCompress_bitwise_np myDecompress_bitwise_np myCompress_bitwise_op "
       "myDecompress_bitwise_op myCompress_bitwise_mask myDecompress_bitwise_mask toSmallDataset_float med_dataset_float "
       "do_crc32 floattostr strtofloat to_absErrorBound_binary").split()
KNOWN = {5: "1.392546", 7: "1.418774", 11: "1.112212"}            # impl/pingpong.csv:36-38 (testfloat_8_8_128, eb 1e-6)


@pytest.fixture(scope="module")
def dropin():
    from data_compression_b200 import build
    build.build()
    if os.path.isdir("/root/reference/impl"):
        subprocess.check_call([os.path.join(ROOT, "oracle", "dropin", "build_dropin.sh")], stdout=subprocess.DEVNULL)
    if not os.path.exists(os.path.join(DROPIN, "pingpong_dcb")):
        pytest.skip("oracle/_ref/dropin was not built (no /root/reference on this machine)")
    return DROPIN


def nm(path, *flags):
    out = subprocess.run(["nm", *flags, path], capture_output=True, text=True, check=True).stdout
    return {ln.split()[-1]: ln.split()[-2] for ln in out.splitlines() if len(ln.split()) >= 2}


@pytest.fixture()
def dataset_txt(tmp_path, golden_dir):
    """impl/dataset/testfloat_8_8_128.txt again, from the committed fixture (9 significant digits round-trip a float32)."""
    x = np.load(os.path.join(golden_dir, "datasets.npz"))["testfloat_8_8_128"]
    p = tmp_path / "t.txt"                 # the driver keeps the name in a char[64] (impl/pingpong.c:80-82): short, relative
    p.write_text("".join("%.9g\n" % v for v in x))
    return "t.txt"


def run_driver(exe, ct, dataset, cwd, extra_env=None):
    env = dict(os.environ, DC_ABS_ERROR_BOUND="0.000001", MPISTUB_SIZE="2")
    env.update(extra_env or {})
    r = subprocess.run([exe, str(ct), dataset], capture_output=True, text=True, cwd=cwd, env=env, timeout=300)
    return r


def test_hot_symbols_bind_to_the_library(dropin):
    lib = nm(os.path.join(ROOT, "data_compression_b200", "libdcb200.so"), "-D", "--defined-only")
    cold = nm(os.path.join(dropin, "dataCompression_cold.o"))
    for s in HOT:
        assert lib.get(s) == "T", "libdcb200.so does not export %s" % s
        assert cold.get(s) == "t", "%s is still global in the localized reference object" % s
    for exe, used in (("pingpong_dcb", 12), ("bmt_dcb", 7)):
        und = nm(os.path.join(dropin, exe), "-u")
        bound = [s for s in HOT if s in und]
        assert len(bound) >= used, (exe, bound)
        needed = subprocess.run(["readelf", "-d", os.path.join(dropin, exe)], capture_output=True, text=True).stdout
        assert "libdcb200.so" in needed


def test_mpi_stand_in_is_validated_by_the_unmodified_reference(dropin, dataset_txt, tmp_path):
    for ct, ratio in KNOWN.items():
        r = run_driver(os.path.join(dropin, "pingpong_ref"), ct, dataset_txt, str(tmp_path))
        assert r.returncode == 0, r.stderr
        assert "data_num = 8192" in r.stdout
        assert ("%s" % ratio) in [ln.split(":")[1].split()[0] for ln in r.stdout.splitlines() if ln.startswith("Compression rate")]


@pytest.mark.gpu
def test_drop_in_pingpong_prints_the_reference_ratios(dropin, dataset_txt, tmp_path):
    for ct, ratio in KNOWN.items():
        r = run_driver(os.path.join(dropin, "pingpong_dcb"), ct, dataset_txt, str(tmp_path), {"LD_DEBUG": "bindings"})
        assert r.returncode == 0, (r.stdout[-500:], r.stderr[-500:])
        rates = [ln.split(":")[1].split()[0] for ln in r.stdout.splitlines() if ln.startswith("Compression rate")]
        assert rates == [ratio], (ct, rates)
        assert "gosa = 0.000000" in r.stdout
        # the dynamic linker's own record: the driver's calls went to libdcb200.so
        bind = [ln for ln in r.stderr.splitlines() if "pingpong_dcb" in ln and "libdcb200.so" in ln and "normal symbol" in ln]
        names = {ln.split("`")[1].split("'")[0] for ln in bind}
        assert {"myCompress_bitwise", "myDecompress_bitwise", "toSmallDataset_float", "med_dataset_float"} <= names, names
