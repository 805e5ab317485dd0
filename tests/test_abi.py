"""The C-ABI library loads and exports every symbol include/spmmdb.h declares; argument
validation follows the reference's conventions.  No compute (runs without a GPU)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def L():
    import __graft_entry__ as g
    import spmm_denseblock_b200 as pkg
    if not os.path.exists(pkg.lib_path()):
        g.build()
    return pkg.lib()


def declared_symbols():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            txt = open(os.path.join(ROOT, "include", fn)).read()
            txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
            names |= set(re.findall(r"\b(spmmdb_[a-z0-9_]+)\s*\(", txt))
    return sorted(names)


def test_header_symbols_are_exported(L):
    syms = declared_symbols()
    assert len(syms) >= 10
    for s in syms:
        assert hasattr(L, s), f"libspmmdb.so does not export {s}"


def test_binding_covers_header(L):
    from spmm_denseblock_b200._lib import SIGNATURES
    assert set(declared_symbols()) <= set(SIGNATURES), set(declared_symbols()) - set(SIGNATURES)


def test_version_and_status(L):
    assert b"spmmdb" in L.spmmdb_version()
    assert L.spmmdb_csrmm_segment(1000) == 64 and L.spmmdb_csrmm_segment(62_000_000) == 512
    # long average rows raise it (4 x 271 -> 2048); the override is validated and process-wide
    assert L.spmmdb_csrmm_segment_for(132534, 35_950_920) == 2048 and L.spmmdb_csrmm_segment_for(169343, 1_166_243) == 64
    assert L.spmmdb_set_csrmm_segment(100) == -1 and L.spmmdb_set_csrmm_segment(1024) == 0
    assert L.spmmdb_get_csrmm_segment() == 1024 and L.spmmdb_set_csrmm_segment(0) == 0


def test_tuning_knobs(L, monkeypatch):
    """spmmdb_set_tuning overrides the environment, INT_MIN clears the override."""
    assert L.spmmdb_get_tuning(b"NO_SUCH_KNOB", 7) == 7
    monkeypatch.setenv("SPMMDB_NO_SUCH_KNOB", "5")
    assert L.spmmdb_get_tuning(b"NO_SUCH_KNOB", 7) == 5
    assert L.spmmdb_set_tuning(b"NO_SUCH_KNOB", 3) == 0 and L.spmmdb_get_tuning(b"NO_SUCH_KNOB", 7) == 3
    assert L.spmmdb_set_tuning(b"NO_SUCH_KNOB", -2**31) == 0 and L.spmmdb_get_tuning(b"NO_SUCH_KNOB", 7) == 5
    assert L.spmmdb_set_tuning(None, 1) == -1
    # the drop-in BSR default arithmetic is the tensor-core path
    assert L.spmmdb_get_bsrmm_math() in (0, 1, 2)


def test_validation_without_device(L):
    z = C.c_void_p(0)
    # negative sizes / bad layout -> INVALID_VALUE before any CUDA call (rocsparse_bsrmm.h:146-148)
    assert L.spmmdb_csrmm_f32(-1, 4, 4, 0, z, z, z, z, 4, 0, z, 4, 0, 1.0, 0.0, z) == -1
    assert L.spmmdb_csrmm_f32(4, 4, 4, 0, z, z, z, z, 4, 7, z, 4, 0, 1.0, 0.0, z) == -1
    # empty problem -> quick success
    assert L.spmmdb_csrmm_f32(0, 4, 4, 0, z, z, z, z, 4, 0, z, 4, 0, 1.0, 0.0, z) == 0
    # bsrmm: bad trans_B -> NOT_SUPPORTED (:139-143), block_dim <= 0 -> INVALID (:146), quick return (:151)
    assert L.spmmdb_bsrmm_f32(0, 2, 1, 1, 1, 1, 2, 1.0, z, z, z, z, 2, 0.0, z, 2, 1, 0, z) == -2
    assert L.spmmdb_bsrmm_f32(0, 0, 1, 1, 1, 1, 0, 1.0, z, z, z, z, 2, 0.0, z, 2, 1, 0, z) == -1
    assert L.spmmdb_bsrmm_f32(0, 0, 0, 1, 1, 1, 2, 1.0, z, z, z, z, 2, 0.0, z, 2, 1, 0, z) == 0
    # null pointers -> INVALID (:156-159)
    assert L.spmmdb_bsrmm_f32(0, 0, 1, 1, 1, 1, 2, 1.0, z, z, z, z, 2, 0.0, z, 2, 1, 0, z) == -1


def test_product_does_not_touch_oracle():
    """The product tree must not import, link or mention the oracle."""
    bad = []
    for base in ("spmm_denseblock_b200", "include", "drivers", "baseline"):
        for dp, _, fns in os.walk(os.path.join(ROOT, base)):
            if "build" in dp.split(os.sep):
                continue
            for fn in fns:
                if fn.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp", "Makefile")):
                    txt = open(os.path.join(dp, fn), errors="ignore").read()
                    if re.search(r"\boracle\b", txt):
                        bad.append(os.path.join(dp, fn))
    assert not bad, bad
