"""The C-ABI library loads without a GPU and exports every symbol include/orbx.h declares; with no
device every entry point fails loudly (no CPU fallback)."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def declared_symbols():
    src = (ROOT / "include" / "orbx.h").read_text()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(orbx_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(pkg):
    syms = declared_symbols()
    assert len(syms) >= 25
    L = C.CDLL(str(pkg.LIB_PATH))
    for s in syms:
        assert hasattr(L, s), f"liborbx.so lacks {s}"
    assert set(syms) == set(pkg.SIGNATURES), "python binding and header disagree"
    assert pkg.lib().orbx_version().decode().endswith("sm_100a")


def test_no_cpu_fallback(pkg):
    if pkg.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.OrbxError) as e:
        pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)
    with pytest.raises(pkg.OrbxError):
        pkg.ORBmatcher()
    import numpy as np
    with pytest.raises(pkg.OrbxError):
        pkg.blur7_u8(np.zeros((16, 16), np.uint8))


def test_product_never_imports_oracle():
    pk = ROOT / "orb-slam2-_b200"
    for p in list(pk.rglob("*.py")) + list(pk.rglob("*.cu")) + list(pk.rglob("*.cuh")) + list(pk.rglob("*.h")) + list(pk.rglob("*.cc")):
        if "build" in p.parts:
            continue
        txt = p.read_text(errors="ignore")
        assert "oracle_lib" not in txt and "liborb_oracle" not in txt and "orb_oracle.h" not in txt, p
