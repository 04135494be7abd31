"""Shared pytest configuration.

Markers:
  gpu  -- tests the round-end driver runs on the B200 box (``-m gpu``).  This
          repository has no GPU code path (BASELINE.json
          "TIER MISMATCH"); the tests carrying the marker are the ones that must
          hold on a machine where ``/root/reference`` does not exist -- parity
          against the committed golden fixtures and, when it travelled, against
          the staged copy of the reference under ``oracle/_ref``.  They need no
          GPU and say so.

          Which tests carry it: a few are marked in their files (the ``*_on_box``
          twins).  Every other test in this directory holds on the box just as
          well -- none needs a GPU, and the ones that need a reference tree skip
          themselves when there is none -- so a run that *selects by the marker*
          (``-m gpu``) gets the marker added to all of them, except the long
          hypothesis sweeps and the tests that already have a marked twin.  A run
          that deselects by it (``-m "not gpu"``, the build container) is
          unaffected and still runs the whole suite.  Net effect: every SURVEY.md
          §8 row's tests run in both places.
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


@pytest.hookimpl(tryfirst=True)
def pytest_cmdline_main(config):
    """The suite is ~1 500 process-spawning tests: ten minutes on one core, under three on eight.  When
    pytest-xdist is there and nobody chose a worker count (``-n``, ``-p no:xdist``, ``--pdb``), run the whole
    directory on the cores the machine has.  ``TK8S_TEST_WORKERS=0`` keeps it in one process."""
    workers = os.environ.get("TK8S_TEST_WORKERS", "")
    if not config.pluginmanager.hasplugin("xdist") or getattr(config.option, "numprocesses", None) is not None \
            or getattr(config.option, "usepdb", False) or getattr(config.option, "collectonly", False) or workers == "0" \
            or os.environ.get("PYTEST_XDIST_WORKER"):
        return None
    if len(config.args) != 1 or os.path.basename(os.path.normpath(config.args[0])) != "tests":
        return None                                   # a single file or test: not worth the workers' start-up
    count = int(workers) if workers.isdigit() else min(8, os.cpu_count() or 1)
    if count >= 2:
        config.option.numprocesses = count
        config.option.dist = "load"
        config.option.tx = ["popen"] * count
    return None


#: not added to the box run: hypothesis sweeps against the live reference (minutes
#: each), and tests whose ``*_on_box`` twin is marked in the file itself
_NOT_ON_BOX = ("test_property_", "test_replacement_matches_golden[", "test_five_of_five", "test_hypothesis_",
               "test_tools_agree_on_random_scenarios", "test_never_crash")


@pytest.hookimpl(tryfirst=True)
def pytest_collection_modifyitems(config, items):
    expr = (config.getoption("markexpr", "") or "").strip()
    if expr != "gpu":
        return
    for item in items:
        if item.get_closest_marker("gpu") is None and not any(item.name.startswith(p) for p in _NOT_ON_BOX):
            item.add_marker(pytest.mark.gpu)


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
