"""The whole chain with nothing scripted in between: setup -> rancher.tf ->
`terraform apply` on the mocked provider -> IP files -> inventory/vars ->
`ansible-playbook` dry-run against the Rancher mock -> environment id -> banner.

Run once under the replacement and (where available) once under the reference
script on top of the same simulated tools: identical observations."""
import re

from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.harness import WORKSPACE_TEMPLATE
from tritonk8ssupervisor_b200.conformance.sandbox import Sandbox
from tritonk8ssupervisor_b200.conformance.scenario import Scenario

ANSWERS = "sim env\nsimulated stack\nsimmaster\nsimnode\n3\n2,4\n2\n10\nyes\n"


def _run(workspace_src, argv, env):
    sc = Scenario(name="simulated", stdin=ANSWERS, pre_remove=["ansible/tmp/kubernetes_environment.id"])
    with Sandbox(workspace_src, sc, sim_tools=True) as sbx:
        return sbx.run(argv, env, timeout=120)


def _check(obs):
    assert obs.exit_status == 0, obs.stdout + obs.stderr
    assert obs.files["terraform/masters.ip"] == "10.0.0.1\n"
    assert obs.files["terraform/hosts.ip"] == "10.0.1.1\n10.0.1.2\n10.0.1.3\n"
    assert obs.files["ansible/hosts"] == "[MASTER]\n10.0.0.1\n[HOST]\n10.0.1.1\n10.0.1.2\n10.0.1.3\n"
    # the id was handed out by the mock during the playbook dry run, not planted by a stub
    assert obs.files["ansible/tmp/kubernetes_environment.id"] == "1a7\n"
    assert "----> Kubernetes dashboard is at http://10.0.0.1:8080/r/projects/1a7/kubernetes-dashboard:9090/" in obs.stdout
    assert "module.simmaster.triton_machine.master: Creation complete (primaryip = 10.0.0.1)" in obs.stdout
    assert "module.simnode3.triton_machine.host: Creation complete (primaryip = 10.0.1.3)" in obs.stdout
    assert "Apply complete! Resources: 4 added" in obs.stdout
    assert len(re.findall(r"unreachable=0    failed=0", obs.stdout)) == 4
    assert 'kubernetes_name: "sim env"' in obs.files["ansible/roles/ranchermaster/vars/vars.yml"]
    tools = [c.split("|")[0] for c in obs.calls if not c.startswith("  env:")]
    assert tools.count("terraform") == 2 and tools.count("ansible-playbook") == 1
    assert '  env:KUBERNETES_NAME="sim env"' in obs.calls


def test_replacement_on_simulated_tools():
    _check(_run(WORKSPACE_TEMPLATE, rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV))


def test_reference_on_simulated_tools_gives_the_same_observation(ref_root):
    from oracle.reference_runner import REFERENCE_ARGV
    theirs = _run(ref_root, REFERENCE_ARGV, None)
    _check(theirs)
    ours = _run(ref_root, rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV)
    assert theirs.diff(ours) == []


def test_simulated_terraform_refuses_a_broken_config(tmp_path):
    """A module that does not fit the generated file: `terraform get` fails, setup dies with its status (errexit)."""
    import shutil
    ws = tmp_path / "ws"
    shutil.copytree(WORKSPACE_TEMPLATE, ws)
    p = ws / "terraform/host/inputs.tf"
    p.write_text(p.read_text().replace('variable "package"', 'variable "pkg"'))
    obs = _run(str(ws), rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV)
    assert obs.exit_status == 1 and "arg-unknown" in obs.stdout and "terraform/masters.ip" not in obs.files


def test_bootstrap_that_forgets_root_keys_makes_the_playbook_fail(tmp_path):
    """Drop the line that copies the keys to root: terraform still succeeds, the
    playbook cannot reach the nodes, setup dies with ansible-playbook's status."""
    import shutil
    ws = tmp_path / "ws"
    shutil.copytree(WORKSPACE_TEMPLATE, ws)
    p = ws / "terraform/host/machine.tf"
    text = p.read_text()
    line = [l for l in text.split("\n") if "authorized_keys /root" in l][0]
    p.write_text(text.replace(line + "\n", ""))
    obs = _run(str(ws), rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV)
    assert obs.exit_status == 2
    assert "UNREACHABLE: Permission denied (publickey) for root" in obs.stdout
    assert "Apply complete! Resources: 4 added" in obs.stdout and "Congratulations" not in obs.stdout
