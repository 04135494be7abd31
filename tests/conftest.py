"""Shared pytest configuration.

Markers:
  gpu  -- tests the round-end driver runs on the B200 box (``-m gpu``).  This
          repository has no GPU code path (BASELINE.json
          "TIER MISMATCH"); the tests carrying the marker are the ones that must
          hold on a machine where ``/root/reference`` does not exist -- parity
          against the committed golden fixtures and, when it travelled, against
          the staged copy of the reference under ``oracle/_ref``.  They need no
          GPU and say so.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

os.environ.setdefault("LC_ALL", "C")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: runs on the round-end B200 box (no GPU needed; see conftest)")


def reference_root():
    """Where an unmodified reference tree can be found, or None.

    ``/root/reference`` exists only in the build container; ``oracle/_ref`` is
    the staged copy made by ``oracle/stage_ref.py`` (git-ignored, travels to the
    box with the snapshot)."""
    for cand in ("/root/reference", os.path.join(ROOT, "oracle", "_ref", "reference")):
        if os.path.isfile(os.path.join(cand, "setup.sh")):
            return cand
    return None


@pytest.fixture(scope="session")
def ref_root():
    r = reference_root()
    if r is None:
        pytest.skip("no reference tree available (neither /root/reference nor oracle/_ref)")
    return r
