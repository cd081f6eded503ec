"""CPU checks of the drop-in libraries (no device work): oracle/_ref/libdropin_matcher.so and libdropin_front.so load,
bind liborbx.so, carry the product's ORBmatcher / ORBextractor instead of the reference's, and the one member that is
pure host code (ORBmatcher::DescriptorDistance) answers like the reference's."""
import subprocess

import numpy as np
import pytest

import ref_front_lib as rf

pytestmark = pytest.mark.skipif(not (rf.available() and rf.DROPIN_PATH.exists() and rf.DROPIN_MATCHER_PATH.exists()),
                                reason="oracle/_ref libraries not built")


def _syms(path, flag):
    out = subprocess.run(["nm", "-D", flag, str(path)], capture_output=True, text=True).stdout
    return {line.split()[-1] for line in out.splitlines() if line.strip()}


def test_symbols():
    assert not any(s.startswith("orbx_") for s in _syms(rf.LIB_PATH, "-u"))
    need = _syms(rf.DROPIN_MATCHER_PATH, "-u")
    assert {"orbx_matcher_create", "orbx_match_window", "orbx_match_blocks"} <= need and "orbx_extract" not in need
    need = _syms(rf.DROPIN_PATH, "-u")
    assert {"orbx_matcher_create", "orbx_match_window", "orbx_match_blocks", "orbx_extract", "orbx_stereo_match"} <= need
    assert not any("ExtractorNode" in s for s in _syms(rf.DROPIN_PATH, "--defined-only"))


@pytest.mark.parametrize("path", [rf.DROPIN_MATCHER_PATH, rf.DROPIN_PATH], ids=["matcher", "front"])
def test_loads_and_descriptor_distance(path):
    drop = rf.variant(path)
    L = drop.lib()          # dlopen resolves every orbx_* import against orb-slam2-_b200/liborbx.so
    ref = rf.lib()
    rng = np.random.default_rng(9)
    a, b = rng.integers(0, 256, (200, 32), dtype=np.uint8), rng.integers(0, 256, (200, 32), dtype=np.uint8)
    b[:5] = a[:5]
    b[5] = ~a[5]
    for i in range(200):
        e = int(np.unpackbits(a[i] ^ b[i]).sum())
        assert L.ref_descriptor_distance(rf._p(a[i]), rf._p(b[i])) == ref.ref_descriptor_distance(rf._p(a[i]), rf._p(b[i])) == e
