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
from tritonk8ssupervisor_b200.conformance.scenario import Scenario, triton_table

pytestmark = pytest.mark.skipif(find_reference() is None, reason="no reference tree (not /root/reference, not oracle/_ref)")


_REF_CACHE = {}


def reference_observation(scenario, root):
    """The reference is deterministic (test_reference_is_deterministic), so one run per
    catalogue scenario serves both the parity test and the goldens-are-current test."""
    key = (scenario.name, root)
    if key not in _REF_CACHE:
        _REF_CACHE[key] = run_reference(scenario, root)
    return _REF_CACHE[key]


def ours(scenario, root):
    return run_in_sandbox(rp.REPLACEMENT_ARGV, root, scenario, extra_env=rp.REPLACEMENT_ENV)


@pytest.mark.parametrize("scenario", ALL, ids=lambda s: s.name)
def test_live_parity(scenario, ref_root):
    assert reference_observation(scenario, ref_root).diff(ours(scenario, ref_root)) == []


@pytest.mark.gpu
@pytest.mark.parametrize("scenario", CONFIGS, ids=lambda s: s.name)
def test_live_parity_on_box(scenario, ref_root):
    """The five scored configs against the staged reference on the box."""
    assert run_reference(scenario, ref_root).diff(ours(scenario, ref_root)) == []


@pytest.mark.parametrize("scenario", ALL, ids=lambda s: s.name)
def test_goldens_are_current(scenario, ref_root):
    doc = rp.load_golden(os.path.join(rp.golden_dir(), scenario.name + ".json"))
    assert doc["reference"]["setup_sh_sha256"] == reference_sha256(ref_root)
    assert Observation.from_json(doc["expected"]).diff(reference_observation(scenario, ref_root)) == []


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


@settings(max_examples=40, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
@given(name=_freetext, descr=_freetext, host=_hostish, prefix=_hostish, count=_countish,
       mnet=_indexish, nnet=_indexish, pkg=_indexish, confirm=_yes_no, nets=_net_names)
def test_property_random_answers(ref_root, name, descr, host, prefix, count, mnet, nnet, pkg, confirm, nets):
    """Random valid and invalid answers; trailing blank lines let every
    re-prompt fall back to its default so the run always reaches the end."""
    answers = [name, descr, host, "", prefix, "", count, "", mnet, "", nnet, "", pkg, "", confirm, "yes"]
    stdin = "".join(a + "\n" for a in answers)
    sc = Scenario(name="prop", stdin=stdin, networks=_net_table(nets))
    assert run_reference(sc, ref_root).diff(ours(sc, ref_root)) == []


# --- §8f rank 2: dashboard wait loop as a state machine -----------------------
_PS_LINES = st.lists(st.sampled_from([
    "1111aaaa2222  gcr.io/kubernetes-dashboard-amd64:v1  k8s_kubernetes-dashboard.abc",
    "3333bbbb4444  gcr.io/pause-amd64:3.0  k8s_POD.d8dbe16c_kubernetes-dashboard-716739405",
    "5555cccc6666  rancher/agent:v1.2.0  rancher-agent",
    "7777dddd8888  gcr.io/pause-amd64:3.0  k8s_POD.aaaa_kube-dns-1",
    "   ", "CONTAINER ID  IMAGE  NAMES"]), max_size=4).map(lambda ls: "".join(l + "\n" for l in ls))
_BODIES = st.sampled_from(["", "503 Service Unavailable\n", "service unavailable", "502 Bad Gateway\n", "<html>starting</html>\n",
                           "* ? [a]\n", "-n kubernetes\n", "-e x\\nkubernetes\n", "KuBeRnEtEs", "two\nlines\n"])
_EPOCHS = st.sampled_from(["1500000000", "1500000001", "1500000015", "1500000029", "1500000030", "0",
                           # setup.sh:64 evaluates the reading as an arithmetic expression
                           "abc", "", "08", "-15", "-16", "15 30", "0x1e", "0x1f", "16#f", "1.5", "i", "n++", "1/0", "(30", "yn",
                           "99999999999999999999", "DASHBOARD_CONTAINER_COUNT+1", "node", "i[0]", "i[1]", "2**4-1"])


@settings(max_examples=25, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
@given(n=st.integers(1, 3), bodies=st.lists(_BODIES, max_size=5), epochs=st.lists(_EPOCHS, max_size=5),
       ps=st.lists(_PS_LINES, min_size=3, max_size=3), ssh_fail=st.sampled_from([None, 0, 1]))
def test_property_wait_loop(ref_root, n, bodies, epochs, ps, ssh_fail):
    """Random clock readings, dashboard bodies and `docker ps` listings: same
    polls, same ssh calls, same containers stopped, same dots on stdout."""
    ips = ["10.0.1.%d" % i for i in range(1, n + 1)]
    sc = Scenario(name="wait", stdin="\n\n\n\n%d\n\n\n\nyes\n" % n, curl_seq=bodies, date_seq=epochs,
                  curl_default="<title>Kubernetes Dashboard</title>\n",
                  ssh_ps={ip: ps[i] for i, ip in enumerate(ips)},
                  ssh_exit=({ips[ssh_fail % n]: 255} if ssh_fail is not None else {}))
    assert run_reference(sc, ref_root).diff(ours(sc, ref_root)) == []


# --- §8f rank 3: clean lifecycle ------------------------------------------------
_LEFTOVERS = {
    "config": "A=1\n", "terraform/rancher.tf": "# x\n", "terraform/masters.ip": "10.0.0.1\n",
    "terraform/hosts.ip": "10.0.1.1\n10.0.1.2\n", "terraform/terraform.tfstate": "{}\n",
    "terraform/terraform.tfstate.backup": "{}\n", "terraform/.terraform/modules/m/x.tf": "#\n",
    "terraform/.terraform.lock": "l\n", "ansible/hosts": "[MASTER]\n", "ansible/clusterUp.retry": "h\n",
    "ansible/other.retry": "h\n", "ansible/ansible.cfg.tmp": "t\n", "tmp/a": "a\n", "tmp/.hidden": "h\n",
    "tmp/d/e": "e\n", "ansible/tmp/extra.id": "z\n", "config.tmp": "c\n", "terraform/notes.txt": "keep\n",
}


@settings(max_examples=40, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
@given(present=st.sets(st.sampled_from(sorted(_LEFTOVERS))), known_hosts=st.booleans(),
       answers=st.lists(st.sampled_from(["yes", "no", "y", "", "YES", " yes"]), min_size=1, max_size=3),
       cfg_key=st.sampled_from(["~/.ssh/id_rsa", "", "/k/with~tilde", "a&b"]))
def test_property_clean(ref_root, present, known_hosts, answers, cfg_key):
    """Random leftovers of earlier runs, with and without known_hosts: the same
    files go, the same stay (ansible/tmp, dot files under tmp/), same tool calls."""
    pre = {p: _LEFTOVERS[p] for p in present}
    pre["ansible/ansible.cfg"] = "[defaults]\nhost_key_checking = False\nprivate_key_file = %s\n# private_key_file = old\n" % cfg_key
    sc = Scenario(name="clean", args=["-c"], stdin="".join(a + "\n" for a in answers) + "yes\n", pre_files=pre,
                  known_hosts="10.0.0.1 ssh-rsa AAAA\n" if known_hosts else None)
    assert run_reference(sc, ref_root).diff(ours(sc, ref_root)) == []


# --- a2: triton profile + ssh key discovery -------------------------------------
_FILE_NAMES = st.sampled_from(["id_rsa", "id_rsa.pub", "id_ed25519", "id_ed25519.pub", "authorized_keys", "config",
                               "known_hosts", "Zkey", "mypub", "a.pub.pub", "github_rsa", "x-pub", ".hidden"])
_FPS = st.sampled_from(["aa:bb:cc:dd", "11:22:33:44", "", "SHA256:abc"])


@settings(max_examples=25, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
@given(names=st.sets(_FILE_NAMES, max_size=6), key_id=_FPS, data=st.data(), no_E=st.booleans(),
       env_extra=st.sampled_from(["", "export DOCKER_HOST=tcp://1.2.3.4:2376\n", "unset SDC_ACCOUNT\n", "# only a comment\n",
                                  "export SDC_URL='https://single.quoted'\n", "SDC_TESTING=1\nexport SDC_TESTING\n"]))
def test_property_key_discovery(ref_root, names, key_id, data, no_E, env_extra):
    """Random ~/.ssh contents and fingerprints (modern and legacy ssh-keygen):
    same ssh-keygen calls in the same order, same SDC_KEY line or the same
    key-not-found exit."""
    from tritonk8ssupervisor_b200.conformance.scenario import triton_env_text
    keygen, legacy, sha = {}, {}, {}
    for n in sorted(names):
        stem = n[:-4] if n.endswith(".pub") else n
        fp = data.draw(_FPS)
        kind = data.draw(st.sampled_from(["md5", "legacy", "sha", "none"]))
        if kind == "md5" and fp:
            keygen[stem] = "2048 MD5:%s c (RSA)\n" % fp
            sha[stem] = "2048 SHA256:zz c (RSA)\n"
        elif kind == "legacy" and fp:
            legacy[stem] = "2048 %s c (RSA)\n" % fp
        elif kind == "sha":
            sha[stem] = "2048 SHA256:zz c (RSA)\n"
    files = {n: "x\n" for n in names}
    for stem in list(keygen) + list(legacy) + list(sha):
        files.setdefault(stem, "k\n")
    sc = Scenario(name="keys", stdin="\n" * 8 + "yes\nyes\n", ssh_files=files, keygen=keygen, keygen_legacy=legacy,
                  keygen_sha=sha, keygen_no_E=no_E, triton_env=triton_env_text(key_id=key_id) + env_extra)
    assert run_reference(sc, ref_root).diff(ours(sc, ref_root)) == []


# --- the operator's environment ---------------------------------------------------
@pytest.mark.parametrize("scenario,env", [
    (Scenario(name="env_sdc_preset_triton_env_fails", stdin="\n" * 8 + "yes\n", fail={"triton-env": 1}),
     {"SDC_URL": "https://preset", "SDC_ACCOUNT": "presetacct", "SDC_KEY_ID": "aa:bb:cc:dd"}),
    (Scenario(name="env_path_variables_preset", stdin="\n" * 8 + "yes\n"),
     {"KUBERNETES_NAME": "from env", "HOST_PACKAGE": "envpkg", "RANCHER_MASTER_NETWORKS": "envnet",
      "KUBERNETES_NUMBER_OF_NODES": "7", "ANSIBLE_HOST_KEY_CHECKING": "True"}),
    (Scenario(name="env_unrelated_exports", stdin="\n" * 8 + "yes\n"),
     {"SDC_TESTING": "1", "TRITON_PROFILE": "other", "KUBERNETES_EXTRA": "x y"}),
], ids=lambda v: v.name if isinstance(v, Scenario) else "env")
def test_preset_environment(ref_root, scenario, env):
    """Variables already exported by the operator's shell: assignments to them stay
    exported (bash), `triton env` overrides or unsets them, and terraform/ansible see
    exactly the same environment under both implementations (the stubs log it)."""
    from oracle.reference_runner import REFERENCE_ARGV
    from tritonk8ssupervisor_b200.conformance.sandbox import Sandbox
    with Sandbox(ref_root, scenario) as s:
        theirs = s.run(REFERENCE_ARGV, env)
    with Sandbox(ref_root, scenario) as s:
        mine = s.run(rp.REPLACEMENT_ARGV, dict(rp.REPLACEMENT_ENV, **env))
    assert theirs.diff(mine) == []
    assert any(c.startswith("  env:") for c in theirs.calls)
