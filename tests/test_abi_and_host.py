"""CPU: the C-ABI library loads and exports every symbol include/*.h declares (no compute without a GPU), the
product fails loudly without a device, and the host-compilable pieces of the CUDA path (std::sort emulation) agree
with libstdc++."""
import ctypes as C
import glob
import os
import re
import subprocess
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = []
    for h in sorted(glob.glob(os.path.join(ROOT, "include", "*.h"))):
        txt = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names += re.findall(r"\b((?:orbx|orbm)_[a-z0-9_]+)\s*\(", txt)
    return sorted(set(names))


@pytest.fixture(scope="module")
def native():
    from mam3slam_b200 import _lib
    if not os.path.exists(_lib.SO_PATH):
        import __graft_entry__ as g
        g.build()
    return _lib.lib()


def test_every_declared_symbol_is_exported(native):
    names = _declared()
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(native, n)]
    assert not missing, missing


def test_no_device_fails_loudly(native):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mam3slam_b200 import ORBextractor, NativeError
    assert native.orbx_device_count() == 0
    with pytest.raises(NativeError, match="no CUDA device"):
        ORBextractor(1000, 1.2, 8, 20, 7)


def test_product_does_not_import_oracle():
    for f in glob.glob(os.path.join(ROOT, "mam3slam_b200", "**", "*"), recursive=True):
        if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp")):
            assert "oracle" not in open(f, errors="ignore").read().replace("oracle/", "").replace("the oracle", ""), f


@pytest.fixture(scope="module")
def emul(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("native") / "sort_emul_host.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-o", so,
                           os.path.join(ROOT, "tests", "native", "sort_emul_host.cc")])
    return C.CDLL(so)


def _emul_sort(emul, size, ulx, heap_only=0):
    size = np.ascontiguousarray(size, np.int32)
    ulx = np.ascontiguousarray(ulx, np.int32)
    perm = np.empty(len(size), np.int32)
    emul.emul_sort_nodes(size.ctypes.data_as(C.c_void_p), ulx.ctypes.data_as(C.c_void_p), len(size),
                         perm.ctypes.data_as(C.c_void_p), heap_only)
    return perm


@pytest.mark.parametrize("n", [0, 1, 2, 3, 15, 16, 17, 31, 33, 64, 100, 257, 1000, 5000])
def test_sort_emulation_matches_std_sort(oracle_mod, emul, n):
    """Same permutation as std::sort for heavily tied keys (the tie order is what the quadtree depends on)."""
    for seed in range(4):
        rng = np.random.default_rng(1000 * n + seed)
        size = rng.integers(2, 2 + (3 if seed < 2 else 40), n)
        ulx = rng.integers(0, 6 if seed % 2 == 0 else 700, n)
        ref = oracle_mod.sort_nodes(size, ulx)
        assert np.array_equal(_emul_sort(emul, size, ulx), ref), (n, seed)
        # the round-based schedule (independent ranges, per-leaf insertion) the GPU warp runs, in two work orders
        assert np.array_equal(_emul_sort(emul, size, ulx, 2), ref), (n, seed)
        assert np.array_equal(_emul_sort(emul, size, ulx, 3), ref), (n, seed)


def test_sort_emulation_adversarial(oracle_mod, emul):
    # sorted, reversed, constant and organ-pipe inputs drive introsort towards its depth limit / heap-sort fallback
    n = 3000
    for size in (np.arange(n), np.arange(n)[::-1], np.full(n, 5), np.minimum(np.arange(n), n - 1 - np.arange(n)),
                 np.arange(n) % 2, (np.arange(n) * 7919) % 13):
        ulx = np.zeros(n, np.int32)
        ref = oracle_mod.sort_nodes(size, ulx)
        assert np.array_equal(_emul_sort(emul, size, ulx), ref)
        assert np.array_equal(_emul_sort(emul, size, ulx, 2), ref)
        assert np.array_equal(_emul_sort(emul, size, ulx, 3), ref)


def _build_adapter_demo(tmpdir):
    exe = os.path.join(str(tmpdir), "adapter_demo")
    libdir = os.path.join(ROOT, "mam3slam_b200")
    subprocess.check_call(["g++", "-std=c++14", "-O1", "-Wall", "-Werror", "-o", exe,
                           os.path.join(ROOT, "tests", "native", "adapter_demo.cc"), "-L" + libdir, "-lmam3slam_b200",
                           "-Wl,-rpath," + libdir])
    return exe


def test_cpp_adapters_compile_and_link(native, tmp_path):
    """mam3slam_b200/cpp/ORBextractor.h / ORBmatcher.h (the reference's class surface) build as C++14 against the cv:: shim
    and link against the C-ABI library; without a GPU the constructor must throw (no CPU fallback)."""
    exe = _build_adapter_demo(tmp_path)
    import torch
    if not torch.cuda.is_available():
        img = tmp_path / "im.raw"
        img.write_bytes(bytes(64 * 64))
        r = subprocess.run([exe, str(img), "64", "64"], capture_output=True, text=True)
        assert r.returncode != 0 and "no CUDA device" in (r.stderr + r.stdout)


@pytest.mark.gpu
def test_cpp_adapter_matches_python_mirror(native, tmp_path):
    from mam3slam_b200 import ORBextractor, synth
    exe = _build_adapter_demo(tmp_path)
    im = synth.frame(752, 480, 0)
    p = tmp_path / "im.raw"
    p.write_bytes(im.tobytes())
    out = subprocess.run([exe, str(p), "752", "480"], capture_output=True, text=True, check=True).stdout.split()
    ex = ORBextractor(1000, 1.2, 8, 20, 7)
    mono, kps, desc = ex(im)
    cs = 0
    for k, d in zip(kps, desc):
        cs = (cs * 1315423911 + int(k["x"]) * 31 + int(k["y"]) * 7 + int(k["octave"])) % (1 << 64)
        for b in d:
            cs = (cs * 131 + int(b)) % (1 << 64)
    assert int(out[0]) == mono and int(out[1]) == len(kps) and int(out[2]) == cs
    assert int(out[3]) == int(np.unpackbits(desc[0] ^ desc[1]).sum()) and int(out[4]) == 210
    ex.close()
