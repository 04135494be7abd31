"""Pin the oracle to the only known answers the reference ships.

The reference has no tests, golden vectors or fixtures.  What it
does ship is documentation that quotes its own output -- the walkthrough
transcript docs/detailed.md:125-219, the prompt list README.md:112-121, and two
stale run artefacts checked into git.  The oracle (reference run in the stub
sandbox) must reproduce those, otherwise the sandbox misrepresents the
reference.  Needs a reference tree; skipped elsewhere."""
import json
import os
import re

import pytest

from oracle.reference_runner import find_reference, run_reference
from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.catalog import CONFIGS
from tritonk8ssupervisor_b200.conformance.scenario import DOC_NETWORKS, DOC_PACKAGES

pytestmark = pytest.mark.skipif(find_reference() is None, reason="no reference tree")


def _doc_lines(ref_root, rel):
    with open(os.path.join(ref_root, rel), encoding="utf-8") as fh:
        return fh.read().split("\n")


def test_fixture_tables_are_the_walkthrough_tables(ref_root):
    """docs/detailed.md:156-159 and :178-195 are where DOC_NETWORKS/DOC_PACKAGES come from."""
    lines = _doc_lines(ref_root, "docs/detailed.md")
    rows = [re.fullmatch(r"(\d+)\.\t(\S+)  ([0-9a-f-]{36})", l) for l in lines]
    rows = [(int(m.group(1)), m.group(2), m.group(3)) for m in rows if m]
    nets = [(n, u) for i, n, u in rows[:4]]
    pkgs = [(n, u) for i, n, u in rows[8:]]
    assert nets == DOC_NETWORKS and rows[4:8] == rows[:4]          # the menu is printed twice
    assert pkgs == DOC_PACKAGES and [i for i, _, _ in rows[8:]] == list(range(1, 19))


def test_oracle_reproduces_the_walkthrough_menus(ref_root):
    """Menu rows in the oracle's stdout are the documented lines, verbatim."""
    out = run_reference(CONFIGS[0], ref_root).stdout.split("\n")
    doc = _doc_lines(ref_root, "docs/detailed.md")
    documented = [l for l in doc if re.match(r"\d+\.\t", l)]
    printed = [l for l in out if re.match(r"\d+\.\t", l)]
    assert printed == documented and len(printed) == 4 + 4 + 18


def test_oracle_reproduces_the_verify_block(ref_root):
    """docs/detailed.md:203-218 vs the oracle, line by line; the one known stale
    line (:218, reworded in setup.sh:468) is the only difference."""
    doc = _doc_lines(ref_root, "docs/detailed.md")
    start = doc.index("Verify that the following configuration is correct:")
    block = doc[start:doc.index("Is the above config correct (yes | no)? yes")]
    out = run_reference(CONFIGS[0], ref_root).stdout.split("\n")
    ostart = out.index("Verify that the following configuration is correct:")
    got = out[ostart:ostart + len(block)]
    diff = [(d, g) for d, g in zip(block, got) if d != g]
    assert len(diff) == 1
    assert diff[0][0] == "Make sure that the nodes and master are part of networks that can communicate with each other."
    assert diff[0][1].startswith("WARN: Make sure that the nodes and master are part of networks that can communicate")


def test_oracle_reproduces_the_readme_prompts(ref_root):
    """README.md:112-121: the eight prompts with their defaults, as the terminal shows them."""
    readme = _doc_lines(ref_root, "README.md")
    start = readme.index("Name your Kubernetes environment: (k8s dev)")
    prompts = readme[start:start + 8]
    with open(os.path.join(rp.golden_dir(), "tty", "tty_quickstart.json"), encoding="utf-8") as fh:
        transcript = json.load(fh)["transcript"].split("\n")
    shown = [l.rstrip() for l in transcript]
    for p in prompts:
        assert p.rstrip() in shown, p


def test_generated_vars_file_has_the_checked_in_shape(ref_root):
    """ansible/roles/ranchermaster/vars/vars.yml:1-3 is a leftover of a real run: same three keys, same quoting."""
    stale = _doc_lines(ref_root, "ansible/roles/ranchermaster/vars/vars.yml")
    made = run_reference(CONFIGS[0], ref_root).files["ansible/roles/ranchermaster/vars/vars.yml"].split("\n")
    assert [l.split(":")[0] for l in stale] == [l.split(":")[0] for l in made]
    assert stale[1:] == made[1:]                                   # name/description lines are identical
    assert re.fullmatch(r"master: [0-9.]+", stale[0]) and re.fullmatch(r"master: [0-9.]+", made[0])


def test_banner_uses_the_checked_in_environment_id(ref_root):
    """ansible/tmp/kubernetes_environment.id (`1a7`) survives `-c` and shows up in the banner URLs."""
    with open(os.path.join(ref_root, "ansible/tmp/kubernetes_environment.id")) as fh:
        env_id = fh.read().strip()
    out = run_reference(CONFIGS[0], ref_root).stdout
    assert "/env/%s/infra/containers" % env_id in out and "/r/projects/%s/kubernetes-dashboard:9090/" % env_id in out


def test_oracle_reproduces_the_documented_banner_urls(ref_root):
    """docs/detailed.md:229-249 quotes the end-of-run banner with placeholders; the three URL
    templates must be the ones the oracle prints (the prose around them was reworded since:
    typographic dashes, 'Dashboard' capitalised -- setup.sh:53,88-89 win)."""
    doc = "\n".join(_doc_lines(ref_root, "docs/detailed.md"))
    urls = re.findall(r"http://<ip of master>:8080(/[^\n]*?)(?=\n|$)", doc)
    assert urls == ["/env/<env id>/infra/containers", "/r/projects/<env id>/kubernetes-dashboard:9090/",
                    "/env/<env id>/kubernetes/kubectl"]
    out = run_reference(CONFIGS[0], ref_root).stdout
    for u in urls:
        assert "http://10.0.0.1:8080" + u.replace("<env id>", "1a7") in out
    assert "Congratulations, your Kubernetes cluster setup has been complete." in doc and \
        "Congratulations, your Kubernetes cluster setup has been complete." in out
    assert "CONGRATULATIONS, YOU HAVE CONFIGURED YOUR KUBERNETES ENVIRONMENT!" in doc and \
        "CONGRATULATIONS, YOU HAVE CONFIGURED YOUR KUBERNETES ENVIRONMENT!" in out
