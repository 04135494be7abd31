"""The in-process tool models must behave exactly like stubs/stubtool.sh: the
replacement run in-process against them reproduces every golden observation
(which was recorded from the reference talking to the real stub processes)."""
import os

import pytest

from tritonk8ssupervisor_b200.conformance import replay as rp

GOLDENS = rp.list_goldens()


@pytest.mark.parametrize("path", GOLDENS, ids=lambda p: os.path.basename(p)[:-5])
def test_in_process_matches_golden(path):
    doc = rp.load_golden(path)
    why = rp.not_reproducible_here(doc)
    if why:
        pytest.skip(why)
    expected, got = rp.replay_in_process(doc)
    assert expected.diff(got) == []


# --- beyond the goldens: the two ways of running the replacement agree on random input ---
from hypothesis import HealthCheck, given, settings, strategies as st  # noqa: E402

from tritonk8ssupervisor_b200.conformance.harness import WORKSPACE_TEMPLATE  # noqa: E402
from tritonk8ssupervisor_b200.conformance.modeltools import ModelSandbox  # noqa: E402
from tritonk8ssupervisor_b200.conformance.sandbox import Sandbox  # noqa: E402
from tritonk8ssupervisor_b200.conformance.scenario import Scenario  # noqa: E402

_answer = st.one_of(st.sampled_from(["", "yes", "no", "2", "9", "10", "1,3", "4,2,2", "km01", "-n x", "a*", 'q"q', "14", "99"]),
                    st.text(st.sampled_from(list("ab 12\"\\*?-ne,")), max_size=8))
_ps = st.sampled_from(["", "1111 img k8s_kubernetes-dashboard.a\n3333 pause k8s_POD.x_kubernetes-dashboard-1\n", "5555 agent rancher-agent\n"])


@settings(max_examples=40, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
@given(answers=st.lists(_answer, min_size=0, max_size=14), clean=st.booleans(),
       curl=st.lists(st.sampled_from(["503 Service Unavailable\n", "starting\n", "kubernetes\n", ""]), max_size=3),
       dates=st.lists(st.sampled_from(["1500000000", "1500000007", "1500000015"]), max_size=3), ps=_ps,
       leftovers=st.sets(st.sampled_from(["config", "terraform/masters.ip", "terraform/hosts.ip", "terraform/rancher.tf", "tmp/x"])))
def test_in_process_equals_cli_on_random_scenarios(answers, clean, curl, dates, ps, leftovers):
    """Same scenario, once through the CLI talking to stub *processes*, once
    in-process talking to the tool *models*: identical observations.  (No reference
    needed, so this also runs where there is none.)"""
    sc = Scenario(name="rand", stdin="".join(a + "\n" for a in answers) + "yes\nyes\n", args=["-c"] if clean else [],
                  curl_seq=curl, date_seq=dates, ssh_ps={"10.0.1.1": ps, "10.0.1.2": ps},
                  pre_files={p: "10.0.0.9\n" for p in leftovers})
    with Sandbox(WORKSPACE_TEMPLATE, sc) as sbx:
        via_cli = sbx.run(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV)
    with ModelSandbox(WORKSPACE_TEMPLATE, sc) as msbx:
        in_process = msbx.run()
    assert via_cli.diff(in_process) == []
