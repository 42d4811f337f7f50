"""The C-ABI library builds, loads, and exports exactly what include/newmap_b200.h declares
(no compute call: there is no GPU where the CPU suite runs)."""
import ctypes
import re

from conftest import ROOT


def _declared():
    text = (ROOT / "include" / "newmap_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(nm_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from newmap_b200 import _lib
    L = _lib.load()
    names = _declared()
    assert len(names) >= 18
    for n in names:
        assert hasattr(L, n), n
    assert sorted(_lib.SIGNATURES) == names
    assert L.nm_abi_version() == 3
    assert L.nm_strerror(-2).decode() == "not a valid .awfmi index"


def test_no_cpu_fallback_without_a_gpu():
    """On a box without a CUDA device opening an index must fail loudly (skipped on GPU boxes)."""
    import pytest
    from newmap_b200 import _lib
    if _lib.load().nm_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(Exception) as e:
        _lib.Index(ROOT / "tests" / "golden" / "genome_s1.awfmi", device=0)
    assert "CUDA" in str(e.value)


def test_product_does_not_import_the_oracle():
    """Only tests/, __graft_entry__.smoke()/build() and bench.py may touch oracle/."""
    for p in (ROOT / "newmap_b200").rglob("*.py"):
        src = p.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), p
        assert "liboracle" not in src and "hostemu" not in src, p
    for p in (ROOT / "newmap_b200" / "csrc").glob("*"):
        if p.suffix in (".cu", ".cuh", ".cpp", ".h"):
            assert "oracle" not in p.read_text().replace("oracle/fm_oracle.c ora_upper_bounds", ""), p
