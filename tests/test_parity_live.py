"""Parity against the reference *run live* (skipped where no reference tree is
available): the oracle is the unmodified setup.sh in the sandbox, the
replacement runs in an identical sandbox, and the two observations must be
byte-identical.  Also: the committed goldens are what the reference produces
now (they cannot go stale silently), and the reference is deterministic.
"""
import os

import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from oracle.reference_runner import find_reference, reference_sha256, run_reference
from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.catalog import ALL, CONFIGS
from tritonk8ssupervisor_b200.conformance.sandbox import Observation, run_in_sandbox
from tritonk8ssupervisor_b200.conformance.scenario import DOC_NETWORKS, Scenario, triton_table

pytestmark = pytest.mark.skipif(find_reference() is None, reason="no reference tree (not /root/reference, not oracle/_ref)")


def ours(scenario, root):
    return run_in_sandbox(rp.REPLACEMENT_ARGV, root, scenario, extra_env=rp.REPLACEMENT_ENV)


@pytest.mark.parametrize("scenario", ALL, ids=lambda s: s.name)
def test_live_parity(scenario, ref_root):
    assert run_reference(scenario, ref_root).diff(ours(scenario, ref_root)) == []


@pytest.mark.gpu
@pytest.mark.parametrize("scenario", CONFIGS, ids=lambda s: s.name)
def test_live_parity_on_box(scenario, ref_root):
    """The five scored configs against the staged reference on the box."""
    assert run_reference(scenario, ref_root).diff(ours(scenario, ref_root)) == []


@pytest.mark.parametrize("scenario", ALL, ids=lambda s: s.name)
def test_goldens_are_current(scenario, ref_root):
    doc = rp.load_golden(os.path.join(rp.golden_dir(), scenario.name + ".json"))
    assert doc["reference"]["setup_sh_sha256"] == reference_sha256(ref_root)
    assert Observation.from_json(doc["expected"]).diff(run_reference(scenario, ref_root)) == []


def test_reference_is_deterministic(ref_root):
    for sc in CONFIGS:
        a, b = run_reference(sc, ref_root), run_reference(sc, ref_root)
        assert a.diff(b) == []


# --- property tests (SURVEY.md §8d "Seeds / distributions") -----------------
_hostish = st.one_of(
    st.text(st.sampled_from(list("abcXYZ019")), min_size=0, max_size=12),
    st.sampled_from(["-", "_", " ", "1abc", "a", "a b", "a-b", "ab_", "Ab9", "kube master", "é"]))
_indexish = st.one_of(
    st.lists(st.integers(0, 20), min_size=0, max_size=5).map(lambda l: ",".join(map(str, l))),
    st.sampled_from(["007", "1,", ",1", "1,,2", "2 ,3", " 4", "4 2", "x", "1-2", "99", "100", "-1", "*", "1;2", "1.5"]))
_freetext = st.text(st.sampled_from(list("ab 12\"'\\*?[]-ne$~;=.,/\t")), min_size=0, max_size=14)
_countish = st.one_of(st.integers(0, 12).map(str), st.sampled_from(["", "x", "1 ", "01", "-3", "9 9"]))
_yes_no = st.sampled_from(["yes", "yes", "yes", "no", "y", ""])


def _net_table(names):
    rows = [(n, "%08x-0000-4000-8000-%012x" % (i, i)) for i, n in enumerate(names, 1)]
    return triton_table(rows)


_net_names = st.lists(
    st.sampled_from(["kube", "kube-net", "kube-net-2", "Joyent-SDC-Public", "Joyent-SDC-Private", "a", "a-b", "a=b",
                     "zz", "Z", "NAME", "net_1", "net.1", "My-Fabric-Network", "kubernetes"]),
    min_size=1, max_size=12, unique=True)


@settings(max_examples=60, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
@given(name=_freetext, descr=_freetext, host=_hostish, prefix=_hostish, count=_countish,
       mnet=_indexish, nnet=_indexish, pkg=_indexish, confirm=_yes_no, nets=_net_names)
def test_property_random_answers(ref_root, name, descr, host, prefix, count, mnet, nnet, pkg, confirm, nets):
    """Random valid and invalid answers; trailing blank lines let every
    re-prompt fall back to its default so the run always reaches the end."""
    answers = [name, descr, host, "", prefix, "", count, "", mnet, "", nnet, "", pkg, "", confirm, "yes"]
    stdin = "".join(a + "\n" for a in answers)
    sc = Scenario(name="prop", stdin=stdin, networks=_net_table(nets))
    assert run_reference(sc, ref_root).diff(ours(sc, ref_root)) == []
