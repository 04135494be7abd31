"""Terraform structural validator: accepts what the path generates (against the
package's modules and, where available, the reference's own), and rejects each
class of breakage with its own code."""
import os
import shutil

import pytest

from tritonk8ssupervisor_b200 import tfvalidate
from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.catalog import CONFIGS
from tritonk8ssupervisor_b200.conformance.harness import WORKSPACE_TEMPLATE

GOOD_ROOT = '''provider "triton" {
    account = "fixture"
    key_material = "${file("~/.ssh/id_rsa")}"
    key_id = "aa:bb"
    url = "https://api"
}

module "kubemaster" {
    source = "master"
    hostname = "kubemaster"
    networks = ["n1"]
    root_authorized_keys = "${file("~/.ssh/id_rsa")}"
    package = "p1"
}

module "kubenode1" {
    source = "host"
    hostname = "kubenode1"
    networks = ["n1","n2"]
    root_authorized_keys = "${file("~/.ssh/id_rsa")}"
    package = "p1"
}
'''


@pytest.fixture
def tf_dir(tmp_path):
    shutil.copytree(os.path.join(WORKSPACE_TEMPLATE, "terraform"), tmp_path / "terraform")
    (tmp_path / "terraform" / "rancher.tf").write_text(GOOD_ROOT)
    return tmp_path / "terraform"


def _generated(scenario):
    doc = rp.load_golden(os.path.join(rp.golden_dir(), scenario.name + ".json"))
    return doc["expected"]["files"]["terraform/rancher.tf"]


@pytest.mark.parametrize("scenario", CONFIGS, ids=lambda s: s.name)
def test_generated_config_validates_against_package_modules(tmp_path, scenario):
    shutil.copytree(os.path.join(WORKSPACE_TEMPLATE, "terraform"), tmp_path / "terraform")
    (tmp_path / "terraform" / "rancher.tf").write_text(_generated(scenario))
    rep = tfvalidate.validate(str(tmp_path / "terraform"))
    assert rep.findings == [], [str(f) for f in rep.findings]


@pytest.mark.parametrize("scenario", CONFIGS, ids=lambda s: s.name)
def test_generated_config_validates_against_reference_modules(tmp_path, scenario, ref_root):
    """The reference's own modules pass with exactly the era warnings SURVEY.md
    §7.4 lists: the duplicated ``tags`` and the self-reference by name."""
    shutil.copytree(os.path.join(ref_root, "terraform"), tmp_path / "terraform")
    (tmp_path / "terraform" / "rancher.tf").write_text(_generated(scenario))
    rep = tfvalidate.validate(str(tmp_path / "terraform"))
    assert rep.ok, [str(f) for f in rep.errors]
    assert sorted(set(rep.codes("warning"))) == ["attr-duplicate", "self-reference"]
    assert rep.codes("warning").count("attr-duplicate") == 2       # master + host main.tf:6,33
    assert rep.codes("warning").count("self-reference") == 4       # main.tf:15,30 in both modules


def test_plan_resolves_inputs(tf_dir):
    plan = tfvalidate.plan(str(tf_dir))
    assert [p["address"] for p in plan] == ["module.kubemaster.triton_machine.master", "module.kubenode1.triton_machine.host"]
    node = plan[1]["args"]
    assert node["name"] == "kubenode1" and node["networks"] == ["n1", "n2"] and node["package"] == "p1"
    assert node["image"] == "0867ef86-e69d-4aaa-ba3b-8d2aef0c204e"      # never set by the generator: module default
    assert node["root_authorized_keys"] == '${file("~/.ssh/id_rsa")}'
    assert node["tags"] == {"name": "kubenode1", "hostname": "kubenode1"}


def _mutate(tf_dir, rel, old, new):
    p = tf_dir / rel
    text = p.read_text()
    assert old in text, (rel, old)
    p.write_text(text.replace(old, new, 1))


MUTATIONS = [
    ("rancher.tf", 'source = "host"', 'source = "nosuchdir"', "module-source"),
    ("rancher.tf", '    source = "host"\n', "", "module-source"),
    ("rancher.tf", 'module "kubenode1"', 'module "kubemaster"', "module-dup"),
    ("rancher.tf", 'package = "p1"\n}\n\nmodule "kubenode1"', 'pakage = "p1"\n}\n\nmodule "kubenode1"', "arg-unknown"),
    ("rancher.tf", '    hostname = "kubenode1"\n', "", "arg-missing"),
    ("rancher.tf", 'networks = ["n1","n2"]', 'networks = "n1"', "arg-type"),
    ("rancher.tf", 'hostname = "kubenode1"', 'hostname = ["kubenode1"]', "arg-type"),
    ("rancher.tf", 'networks = ["n1","n2"]', 'networks = ["n1","n2"', "syntax"),
    ("rancher.tf", '${file("~/.ssh/id_rsa")}"\n    package = "p1"\n}\n\nmodule "kubenode1"', '${nofunc("x")}"\n    package = "p1"\n}\n\nmodule "kubenode1"', "func-unknown"),
    ("rancher.tf", 'key_material = "${file("~/.ssh/id_rsa")}"', 'key_material = "${file("a", "b")}"', "func-arity"),
    ("rancher.tf", 'url = "https://api"', 'url = "${var.nope}"', "var-undeclared"),
    ("host/machine.tf", "${var.package}", "${var.pkg}", "var-undeclared"),
    ("host/machine.tf", 'name                 = "${var.hostname}"', 'name = "${triton_machine.other.id}"', "ref-unknown"),
    ("host/machine.tf", 'name                 = "${var.hostname}"', 'name = "${self.id}"', "self-outside"),
    ("host/machine.tf", 'provisioner "local-exec"', 'provisioner "nope-exec"', "provisioner"),
    ("host/machine.tf", "command = ", "comand = ", "provisioner"),
    ("host/machine.tf", "agent       = false", "agnt = false", "provisioner"),
    ("host/inputs.tf", 'type        = "list"', 'type        = "array"', "variable-shape"),
    ("host/inputs.tf", 'description = "image UUID', 'descr = "image UUID', "variable-shape"),
    ("master/inputs.tf", 'variable "image" {', 'variable "package" {', "module-dup"),
]


@pytest.mark.parametrize("rel,old,new,code", MUTATIONS, ids=[m[3] + "-" + str(i) for i, m in enumerate(MUTATIONS)])
def test_breakage_is_rejected(tf_dir, rel, old, new, code):
    assert tfvalidate.validate(str(tf_dir)).ok
    _mutate(tf_dir, rel, old, new)
    rep = tfvalidate.validate(str(tf_dir))
    assert not rep.ok
    assert code in rep.codes("error"), [str(f) for f in rep.findings]


def test_warnings_do_not_fail(tf_dir):
    _mutate(tf_dir, "host/machine.tf", 'image                = "${var.image}"', 'image = "${var.image}"\n  colour = "red"')
    _mutate(tf_dir, "rancher.tf", 'url = "https://api"', 'url = "https://api"\n    colour = "red"')
    rep = tfvalidate.validate(str(tf_dir))
    assert rep.ok and rep.codes("warning").count("attr-unknown") == 2
