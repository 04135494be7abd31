"""Interactive parity: on a terminal the operator sees the same transcript
(prompts on stderr, echoed input, output) from the replacement as was recorded
from the reference (tests/golden/tty, oracle/make_golden_tty.py)."""
import json
import os
import shutil
import tempfile

import pytest

from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.sandbox import Sandbox, run_on_pty
from tritonk8ssupervisor_b200.conformance.scenario import Scenario

TTY_DIR = os.path.join(rp.golden_dir(), "tty")
CASES = sorted(f for f in os.listdir(TTY_DIR) if f.endswith(".json"))


def _replay(name):
    with open(os.path.join(TTY_DIR, name), encoding="utf-8") as fh:
        doc = json.load(fh)
    tmp = tempfile.mkdtemp(prefix="tk8s-tty-")
    try:
        skel = rp.build_skeleton(doc, os.path.join(tmp, "skel"))
        with Sandbox(skel, Scenario.from_json(doc["scenario"])) as sbx:
            got = run_on_pty(sbx, rp.REPLACEMENT_ARGV, doc["answers"], rp.REPLACEMENT_ENV)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return doc["transcript"], got


@pytest.mark.parametrize("name", CASES)
def test_transcript_matches_reference(name):
    want, got = _replay(name)
    assert got == want


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_transcript_matches_reference_on_box(name):
    want, got = _replay(name)
    assert got == want


def test_catalogue_and_fixtures_agree():
    from tritonk8ssupervisor_b200.conformance.catalog import TTY_CASES
    assert sorted(sc.name + ".json" for sc, _ in TTY_CASES) == CASES
    for sc, answers in TTY_CASES:
        with open(os.path.join(TTY_DIR, sc.name + ".json"), encoding="utf-8") as fh:
            doc = json.load(fh)
        assert doc["scenario"] == sc.to_json() and doc["answers"] == answers
