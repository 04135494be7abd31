"""Parity of the replacement with the reference on the committed golden
fixtures: stdout, stderr, exit status, every generated/changed/removed file and
the child-process trace (argv, cwd, exported environment) -- byte for byte.

These run everywhere (no reference tree needed).  The same test is collected a
second time under the ``gpu`` marker so that the round-end box, where
``/root/reference`` does not exist, exercises it too.
"""
import os

import pytest

from tritonk8ssupervisor_b200.conformance import replay as rp

GOLDENS = rp.list_goldens()


def _id(path):
    return os.path.basename(path)[:-5]


def _replay(path):
    doc = rp.load_golden(path)
    why = rp.not_reproducible_here(doc)
    if why:
        pytest.skip(why)
    return rp.replay(doc)


def test_golden_set_is_complete():
    from tritonk8ssupervisor_b200.conformance.catalog import ALL
    assert sorted(_id(p) for p in GOLDENS) == sorted(s.name for s in ALL)


def test_golden_scenarios_match_catalogue():
    """A fixture is only as good as the scenario recorded in it."""
    from tritonk8ssupervisor_b200.conformance.catalog import BY_NAME
    for path in GOLDENS:
        doc = rp.load_golden(path)
        assert doc["scenario"] == BY_NAME[_id(path)].to_json(), "stale fixture %s: rerun oracle/make_golden.py" % path


@pytest.mark.parametrize("path", GOLDENS, ids=_id)
def test_replacement_matches_golden(path):
    expected, got = _replay(path)
    assert expected.diff(got) == []


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDENS, ids=_id)
def test_replacement_matches_golden_on_box(path):
    """Same check, selected by ``-m gpu`` (see conftest: no GPU is used)."""
    expected, got = _replay(path)
    assert expected.diff(got) == []
