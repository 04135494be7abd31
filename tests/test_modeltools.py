"""The in-process tool models must behave exactly like stubs/stubtool.sh: the
replacement run in-process against them reproduces every golden observation
(which was recorded from the reference talking to the real stub processes)."""
import os

import pytest

from tritonk8ssupervisor_b200.conformance import replay as rp

GOLDENS = rp.list_goldens()


@pytest.mark.parametrize("path", GOLDENS, ids=lambda p: os.path.basename(p)[:-5])
def test_in_process_matches_golden(path):
    expected, got = rp.replay_in_process(rp.load_golden(path))
    assert expected.diff(got) == []
