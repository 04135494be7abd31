"""A rougher generator than the hypothesis strategies of test_parity_live.py:
whole scenarios at random -- answers made of shell metacharacters and echo
options, shuffled and truncated `triton` tables, stale files, scripted tool
failures, tools that are not installed, tools that eat stdin, non-numeric
clock readings, and an operator environment that exports the very names the
script uses for its own variables.  Replacement and reference must agree on
everything observable, the complete environment of every child included.

During development 4 000 seeds of this generator were run against the
reference; what it found is fixed and pinned as catalogue scenarios
(``env_exported_names_of_script_variables``, ``env_shell_level_*``).  The test
runs a slice of it on every run where a reference tree is available.
"""
import random

import pytest

from oracle.reference_runner import find_reference, run_reference
from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.sandbox import run_in_sandbox
from tritonk8ssupervisor_b200.conformance.scenario import DOC_NETWORKS, DOC_PACKAGES, Scenario, triton_env_text, triton_table

pytestmark = pytest.mark.skipif(find_reference() is None, reason="no reference tree (not /root/reference, not oracle/_ref)")

ALPHABET = list("ab 12\"'\\*?[]-ne$~;=.,/\t#&|<>(){}!%^`@:+_") + [
    "-e ", "-n ", "\\\\n", "\\\\0", "\\\\c", "\\\\t", "yes", "no", "é", "\x01", "\x7f", "\udcff", "  ", "0", "9", "10", "1,2", "4,2,2",
    "$HOME", "$(id)", "KUBERNETES_NUMBER_OF_NODES=3", "\\\\055e "]
EXPORTED = [("SHLVL", "7"), ("KUBERNETES_NAME", "from env"), ("i", "5"), ("yn", "q"), ("network", "n"), ("HOST_PACKAGE", "p"),
            ("OLDPWD", "/tmp"), ("PWD", "/x"), ("f", "ff"), ("line", "l"), ("master", "m"), ("_", "u"), ("IFS", ","), ("tmp", "t"),
            ("networks", "nn"), ("packages", "pp"), ("foundKey", "fk"), ("countNetwork", "9"), ("values", "vv"),
            ("gotValidInput", "g"), ("package", "pk"), ("cRetVal", "c"), ("node", "nd"), ("host_key", "hk"), ("theargument", "ta")]
CLOCK = ["1500000000", "1500000001", "abc", "", "08", "-15", "i", "1 2", "0x1e", "yn", "x=1", "1/0", "n[", "i[0]"]
FAILING = ["triton-env", "triton-networks", "triton-packages", "terraform-get", "terraform-apply", "ansible-playbook--i",
           "curl---connect-timeout", "ssh-keygen--E", "ssh-keygen--l", "date-+%s", "sleep-2", "sleep-1", "ssh--o"]


def _text(r, k=6):
    return "".join(r.choice(ALPHABET) for _ in range(r.randint(0, k)))


def _answers(r):
    pools = [lambda: _text(r), lambda: "", lambda: r.choice(["km01", "a", "1x", "abc def", "Ab9"]), lambda: str(r.randint(0, 12)),
             lambda: ",".join(str(r.randint(0, 6)) for _ in range(r.randint(1, 4))), lambda: r.choice(["yes", "no", "y", "YES", " yes "])]
    return "".join(r.choice(pools)() + "\n" for _ in range(r.randint(0, 22))) + r.choice(["yes\n", "yes", "", "no\n", "yes\nyes\nyes\n"])


def _table(r, rows):
    rows = r.sample(rows, r.randint(0, len(rows)))
    rows += [(r.choice(["kube", "kube-net", "a=b", "NAME", "NAMEX", "x y", "*", "k4-highcpu-kvm-7.75G", "Joyent-SDC-Public", "z-kvm-1"]),
              r.choice(["id-%d" % i, "", "a b", "*"])) for i in range(r.randint(0, 3))]
    r.shuffle(rows)
    t = triton_table(rows) if rows else "NAME  ID\n"
    return r.choice([t, t, t.replace("NAME", "NAME ", 1), t.rstrip("\n"), ""])


def make(seed: int) -> Scenario:
    r = random.Random(seed)
    kw = dict(name="fuzz%d" % seed, stdin=_answers(r), networks=_table(r, list(DOC_NETWORKS)), packages=_table(r, list(DOC_PACKAGES)))
    if r.random() < 0.3:
        kw["args"] = r.choice([["-c"], ["-c", "x"], ["x"], ["-C"], [""]])
    if r.random() < 0.3:
        kw["pre_files"] = {p: "x\n" for p in r.sample(
            ["config", "terraform/rancher.tf", "terraform/masters.ip", "terraform/hosts.ip", "ansible/hosts", "tmp/a", "config.tmp",
             "ansible/x.retry", "terraform/terraform.tfstate"], r.randint(1, 4))}
    if r.random() < 0.3:
        kw["date_seq"] = [r.choice(CLOCK) for _ in range(r.randint(1, 4))]
        kw["curl_seq"] = [r.choice(["starting\n", "", "503 Service Unavailable\n", "* ?\n", "-n kubernetes\n"]) for _ in range(r.randint(0, 3))]
    if r.random() < 0.2:
        kw["fail"] = {r.choice(FAILING): r.choice([1, 2, 127, 255])}
    if r.random() < 0.2:
        kw["missing_tools"] = [r.choice(["triton", "ssh-keygen", "terraform", "ansible-playbook", "curl", "ssh"])]
    if r.random() < 0.25:
        kw["env"] = dict(r.sample(EXPORTED, r.randint(1, 4)))
    if r.random() < 0.15:
        kw["triton_env"] = triton_env_text(key_id=r.choice(["aa:bb:cc:dd", "", "x y", "*"])) + r.choice(
            ["", "export A=$HOME\n", "unset HOME\n", "export HOME=/nonexistent\n", "export PATH=$PATH\n", "B=~/x\nexport B\n",
             "export network=zz\n", "export i\n", "unset PWD\n"])
    if r.random() < 0.15:
        kw["ssh_files"] = {r.choice(["id_rsa", "a b", "*", "k.pub", "x;y", "-n", "z&"]): "k\n" for _ in range(r.randint(0, 3))}
    if r.random() < 0.1:
        kw["stdin_eaters"] = {r.choice(["triton-networks", "triton-packages", "terraform-apply", "ssh-keygen--E",
                                        "curl---connect-timeout", "date-+%s", "sleep-2"]): r.randint(1, 2)}
    return Scenario(**kw)


@pytest.mark.parametrize("block", range(6))
def test_property_fuzzed_scenarios(ref_root, block):
    bad = []
    for seed in range(7000 + 20 * block, 7000 + 20 * (block + 1)):
        sc = make(seed)
        theirs = run_reference(sc, ref_root)
        mine = run_in_sandbox(rp.REPLACEMENT_ARGV, ref_root, sc, extra_env=rp.REPLACEMENT_ENV)
        diffs = theirs.diff(mine)
        if diffs:
            bad.append((seed, sc.to_json(), diffs))
    assert bad == []
