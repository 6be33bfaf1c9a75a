"""CPU tests of the drop-in boundary: the C-ABI library loads without a GPU, exports every symbol the headers
declare, and refuses to compute without a device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from tests import util  # noqa: F401

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lara_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(built):
    from lara_b200 import _lib
    lib = C.CDLL(_lib.LIB_PATH)
    names = _declared("lara_b200.h") + _declared("lara_host.h")
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), "symbol %s declared in include/ but not exported" % n
    # and the ctypes table covers exactly the declared set
    assert sorted(_lib.ALL_SYMBOLS) == sorted(set(names))


def test_struct_sizes_match_header(built):
    from lara_b200 import _lib
    assert C.sizeof(_lib.LaraParams) == 4 * 4 + 4 * 4 + 2 * 4 + 25 * 4 + 3 * 4 + 9 * 4
    assert C.sizeof(_lib.LaraPairResult) == 32
    assert C.sizeof(_lib.LaraStats) == 7 * 8 + 10 * 8 + 8 + 16 * 8
    assert _lib.load().lara_gpu_version() == 1


def test_no_cpu_fallback(built):
    """Without a CUDA device lara_gpu_create must fail loudly with LARA_E_NO_DEVICE."""
    import lara_b200 as lb
    from lara_b200 import _lib
    if _lib.load().lara_gpu_visible_devices() > 0:
        pytest.skip("a CUDA device is visible")
    with pytest.raises(lb.LaraError) as e:
        lb.Context(lb.Parameters())
    assert "no CPU fallback" in str(e.value) or "-2" in str(e.value)


def test_unsupported_options_are_refused(built):
    import lara_b200 as lb
    from lara_b200 import _lib
    lib = _lib.load()
    for kw in (dict(matching=0), dict(matching=9), dict(num_iterations=0)):
        p = lb.Parameters()
        p.set(**kw)
        h = C.c_void_p()
        rc = lib.lara_gpu_create(C.byref(h), C.byref(p.c), 0, None)
        assert rc in (-1, -4), (kw, rc)


def test_product_does_not_reference_the_oracle():
    """Nothing under lara_b200/ (the shipped path) may import, link or call oracle/."""
    bad = []
    for d, _, files in os.walk(os.path.join(ROOT, "lara_b200")):
        if "build" in d or "__pycache__" in d:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".sh")):
                if "oracle" in open(os.path.join(d, f), errors="ignore").read():
                    bad.append(os.path.join(d, f))
    assert not bad, bad
