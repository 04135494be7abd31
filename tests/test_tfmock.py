"""`terraform apply` against the mocked provider: the IP-file contract the
path relies on is *derived from the module files* (main.tf:29-31 local-exec),
and agrees with what the shell stub hard-codes -- i.e. with the goldens."""
import os
import shutil

import pytest

from tritonk8ssupervisor_b200 import tfmock
from tritonk8ssupervisor_b200.conformance import harness, replay as rp
from tritonk8ssupervisor_b200.conformance.catalog import CONFIGS


def _sources():
    from conftest import reference_root
    out = [("package", os.path.join(harness.WORKSPACE_TEMPLATE, "terraform"))]
    if reference_root():
        out.append(("reference", os.path.join(reference_root(), "terraform")))
    return out


@pytest.mark.parametrize("label,src", _sources(), ids=[s[0] for s in _sources()])
@pytest.mark.parametrize("scenario", CONFIGS, ids=lambda s: s.name)
def test_mock_apply_reproduces_the_ip_files(tmp_path, scenario, label, src):
    doc = harness.golden_for(scenario)
    files = doc["expected"]["files"]
    tf = tmp_path / "terraform"
    shutil.copytree(src, tf)
    (tf / "rancher.tf").write_text(files["terraform/rancher.tf"])
    machines, report = tfmock.apply(str(tf))
    assert report.ok, [str(f) for f in report.errors]
    assert (tf / "masters.ip").read_text() == files["terraform/masters.ip"]
    assert (tf / "hosts.ip").read_text() == files["terraform/hosts.ip"]
    names = harness.expected_module_names(rp.Observation.from_json(doc["expected"]))
    assert [m.module for m in machines] == names
    for m in machines:
        # what the provider is asked for: the generated inputs + the module defaults
        assert m.args["name"] == m.module and m.args["tags"] == {"name": m.module, "hostname": m.module}
        assert m.args["image"] == "0867ef86-e69d-4aaa-ba3b-8d2aef0c204e"
        assert m.args["package"] in files["terraform/rancher.tf"] and isinstance(m.args["networks"], list)
        assert m.args["root_authorized_keys"] == '${file("~/.ssh/id_rsa")}'
        # the bootstrap the machine would get over ssh (main.tf:13-27)
        (boot,) = m.remote_exec
        assert boot["connection"]["host"] == m.computed["primaryip"] and boot["connection"]["user"] == "ubuntu"
        assert boot["connection"]["agent"] is False
        assert boot["inline"][0] == "sleep 30" and any("python-minimal" in c for c in boot["inline"])
        assert m.local_exec == ["echo %s >> %s" % (m.computed["primaryip"], "masters.ip" if m.rname == "master" else "hosts.ip")]


def test_count_scaling_to_ten_workers(tmp_path):
    """Config 5's literal request, as an extension: 10 workers through generator + mock apply."""
    tf = tmp_path / "terraform"
    shutil.copytree(os.path.join(harness.WORKSPACE_TEMPLATE, "terraform"), tf)
    blocks = [harness.golden_for(CONFIGS[0])["expected"]["files"]["terraform/rancher.tf"]]
    one = blocks[0][blocks[0].index('\nmodule "kubenode1"'):]
    for i in range(2, 11):
        blocks.append(one.replace("kubenode1", "kubenode%d" % i))
    (tf / "rancher.tf").write_text("".join(blocks))
    machines, report = tfmock.apply(str(tf))
    assert report.ok and len(machines) == 11
    assert (tf / "hosts.ip").read_text().split() == ["10.0.1.%d" % i for i in range(1, 11)]


def test_nothing_is_applied_when_validation_fails(tmp_path):
    tf = tmp_path / "terraform"
    shutil.copytree(os.path.join(harness.WORKSPACE_TEMPLATE, "terraform"), tf)
    text = harness.golden_for(CONFIGS[0])["expected"]["files"]["terraform/rancher.tf"]
    (tf / "rancher.tf").write_text(text.replace('hostname = "kubenode1"', 'hostnam = "kubenode1"'))
    machines, report = tfmock.apply(str(tf))
    assert machines == [] and not report.ok and not (tf / "masters.ip").exists()


def test_failing_local_exec_stops_the_apply(tmp_path):
    tf = tmp_path / "terraform"
    shutil.copytree(os.path.join(harness.WORKSPACE_TEMPLATE, "terraform"), tf)
    (tf / "rancher.tf").write_text(harness.golden_for(CONFIGS[1])["expected"]["files"]["terraform/rancher.tf"])
    p = tf / "host" / "machine.tf"
    p.write_text(p.read_text().replace(">> hosts.ip", ">> /nonexistent-dir/hosts.ip"))
    machines, report = tfmock.apply(str(tf))
    assert "local-exec" in report.codes("error") and [m.module for m in machines] == ["kubemaster", "kubenode1"]


@pytest.mark.parametrize("label,src", _sources(), ids=[s[0] for s in _sources()])
def test_bootstrap_makes_machines_usable_by_ansible(tmp_path, label, src):
    """main.tf:21-26: after the remote-exec lines root can log in with the operator's key and python exists."""
    tf = tmp_path / "terraform"
    shutil.copytree(src, tf)
    (tf / "rancher.tf").write_text(harness.golden_for(CONFIGS[1])["expected"]["files"]["terraform/rancher.tf"])
    machines, report = tfmock.apply(str(tf))
    assert report.ok
    for m in machines:
        st = tfmock.bootstrap_state(m)
        assert st["ansible_ready"], st
        assert st["waited_s"] == 30 and st["packages"] == ["python-minimal"] and st["unmodelled"] == []


@pytest.mark.parametrize("old,new,needle", [
    ('"sudo apt-get update",\n', "", "before apt-get update"),
    ("-y", "", "would wait for confirmation"),
    ('user        = "ubuntu"', 'user        = "root"', "only admits 'ubuntu'"),
])
def test_broken_bootstrap_is_noticed(tmp_path, old, new, needle):
    tf = tmp_path / "terraform"
    shutil.copytree(os.path.join(harness.WORKSPACE_TEMPLATE, "terraform"), tf)
    (tf / "rancher.tf").write_text(harness.golden_for(CONFIGS[0])["expected"]["files"]["terraform/rancher.tf"])
    p = tf / "host" / "machine.tf"
    text = p.read_text()
    assert old in text
    p.write_text(text.replace(old, new, 1))
    machines, report = tfmock.apply(str(tf))
    host = [m for m in machines if m.rname == "host"][0]
    st = tfmock.bootstrap_state(host)
    assert not st["ansible_ready"] and any(needle in e for e in st["errors"]), st
    assert tfmock.bootstrap_state([m for m in machines if m.rname == "master"][0])["ansible_ready"]
