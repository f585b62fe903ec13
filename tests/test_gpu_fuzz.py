"""Randomized round trips (tools/fuzz_roundtrip.py): random segment mixes, sizes around block / sub-block / unit boundaries,
random preferences; our frames through the oracle and our decoder, the LZ4F_* layer with random chunkings, and frames the
reference library makes (linked / independent / flushed) through our decoder.  900 cases of it ran clean on a B200
(profiles/r1d_fuzz.txt); the suite runs a short slice."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fuzz_roundtrip_slice():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_roundtrip.py"), "60", "7"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "fuzz ok: 60 cases" in r.stdout
