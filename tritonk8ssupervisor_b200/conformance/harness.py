"""The offline conformance harness: BASELINE.json's metric.

    metric = configs passing / 5

A config passes iff, with no network, no Triton credentials and none of the
real tools:

 (i)   parity -- the replacement, run through its CLI in the sandbox on the
       config's answers, reproduces the recorded observation of the reference
       byte for byte (stdout, stderr, exit status, generated/changed/removed
       files, child-process trace);
 (ii)  terraform -- the structural validator accepts the generated
       ``terraform/rancher.tf`` together with the two modules it instantiates,
       the plan holds one ``triton_machine`` per module block with the resolved
       inputs, and ``apply`` against the mocked provider (tfmock: the modules'
       own local-exec lines, interpreted) writes the IP files the run recorded;
 (iii) ansible -- the structural checker accepts ``clusterUp.yml`` + roles with
       the generated vars file, ``ansible.cfg`` is valid, and the inventory
       rendered from the generated ``ansible/hosts`` equals the expected
       ``ansible-inventory --list`` document;
 (iv)  determinism -- a second run gives the identical observation.

The workspace the replacement runs in is the package's own template
(``tritonk8ssupervisor_b200/workspace``); the few static files whose *content*
the path reads come from the golden fixture so the bytes are comparable.
"""
from __future__ import annotations

import dataclasses
import os
import re
import time
from typing import Any, Dict, List, Optional, Sequence

from .. import ansiblecheck, inventory, tfvalidate
from . import replay as rp
from .catalog import CONFIGS
from .sandbox import Observation, Sandbox
from .scenario import Scenario

WORKSPACE_TEMPLATE = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "workspace")


@dataclasses.dataclass
class ConfigResult:
    name: str
    parity: bool = False
    terraform: bool = False
    ansible: bool = False
    inventory: bool = False
    deterministic: bool = False
    details: List[str] = dataclasses.field(default_factory=list)
    n_modules: int = 0
    n_hosts: int = 0
    wall_s: float = 0.0

    @property
    def passed(self) -> bool:
        return self.parity and self.terraform and self.ansible and self.inventory and self.deterministic

    def to_json(self) -> Dict[str, Any]:
        d = dataclasses.asdict(self)
        d["passed"] = self.passed
        return d


def expected_inventory(obs: Observation) -> Dict[str, Any]:
    """The ``ansible-inventory --list`` document implied by what terraform
    recorded: masters.ip -> MASTER, hosts.ip -> HOST (setup.sh:122-126)."""
    masters = obs.files.get("terraform/masters.ip", "").split()
    hosts = obs.files.get("terraform/hosts.ip", "").split()
    doc: Dict[str, Any] = {"_meta": {"hostvars": {h: {} for h in sorted(set(masters + hosts))}},
                           "all": {"children": ["HOST", "MASTER", "ungrouped"]}}
    if hosts:
        doc["HOST"] = {"hosts": sorted(set(hosts))}
    if masters:
        doc["MASTER"] = {"hosts": sorted(set(masters))}
    return doc


def expected_module_names(scenario_obs: Observation) -> List[str]:
    """Module labels in file order, read back from the generated rancher.tf."""
    return re.findall(r'^module "([^"]*)" \{$', scenario_obs.files.get("terraform/rancher.tf", ""), re.M)


def expected_machine_count(obs: Observation) -> int:
    """1 master + the node count the run wrote to ./config (setup.sh:202); where the
    observation carries no ./config (the count-scaling extension), one per IP recorded."""
    m = re.search(r"^KUBERNETES_NUMBER_OF_NODES=([1-9][0-9]*)$", obs.files.get("config", ""), re.M)
    if m:
        return 1 + int(m.group(1))
    return len((obs.files.get("terraform/masters.ip", "") + obs.files.get("terraform/hosts.ip", "")).split())


def check_generated_tree(workspace: str, obs: Observation, res: ConfigResult) -> None:
    """(ii) and (iii) on a workspace that holds the generated files."""
    tf = tfvalidate.validate(os.path.join(workspace, "terraform"))
    res.details += ["terraform: %s" % f for f in tf.errors]
    names = expected_module_names(obs)
    planned = tfvalidate.plan(os.path.join(workspace, "terraform"))
    res.n_modules = len(planned)
    plan_ok = [p["address"].split(".")[1] for p in planned] == names and len(set(names)) == len(names) \
        and all(p["type"] == "triton_machine" and p["args"].get("name") == p["address"].split(".")[1] for p in planned)
    if not plan_ok:
        res.details.append("terraform: plan does not hold one triton_machine per generated module block: %r vs %r"
                           % ([p["address"] for p in planned], names))
    mock_ok = _mock_apply_agrees(workspace, obs, res) if tf.ok else False
    want_machines = expected_machine_count(obs)
    if len(planned) != want_machines:
        res.details.append("terraform: %d machines planned, the run asked for %d (1 master + KUBERNETES_NUMBER_OF_NODES)"
                           % (len(planned), want_machines))
    res.terraform = tf.ok and plan_ok and mock_ok and len(planned) == want_machines

    hosts_file = os.path.join(workspace, "ansible", "hosts")
    try:
        inv = inventory.parse_file(hosts_file)
    except (OSError, inventory.InventoryError) as exc:
        res.details.append("inventory: %s" % exc)
        inv = None
    if inv is not None:
        got, want = inv.to_list_json(), expected_inventory(obs)
        res.n_hosts = len(got["_meta"]["hostvars"])
        res.inventory = got == want
        if not res.inventory:
            res.details.append("inventory: rendered %r, expected %r" % (got, want))
    an = ansiblecheck.check_playbook(os.path.join(workspace, "ansible", "clusterUp.yml"), inv)
    cfg = ansiblecheck.check_ansible_cfg(os.path.join(workspace, "ansible", "ansible.cfg"))
    res.details += ["ansible: %s" % f for f in an.errors + cfg.errors]
    res.ansible = an.ok and cfg.ok


def _mock_apply_agrees(workspace: str, obs: Observation, res: ConfigResult) -> bool:
    """`terraform apply` against the mocked provider (tfmock), in a scratch copy of
    terraform/: interpreting the modules' own local-exec lines must give the IP
    files the run recorded."""
    import shutil
    import tempfile

    from .. import tfmock
    tmp = tempfile.mkdtemp(prefix="tk8s-tfmock-")
    try:
        tf = shutil.copytree(os.path.join(workspace, "terraform"), os.path.join(tmp, "terraform"),
                             ignore=shutil.ignore_patterns("*.ip", "terraform.tfstate*", ".terraform*"))
        machines, rep = tfmock.apply(tf)
        got = {}
        for name in ("masters.ip", "hosts.ip"):
            p = os.path.join(tf, name)
            got["terraform/" + name] = open(p).read() if os.path.exists(p) else None
        want = {k: obs.files.get(k) for k in got}
        if not rep.ok or got != want:
            res.details.append("terraform: mock apply gives %r, the run recorded %r (%s)"
                               % (got, want, "; ".join(str(f) for f in rep.errors)))
            return False
        return True
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def run_config(doc: dict, argv: Optional[Sequence[str]] = None, extra_env: Optional[Dict[str, str]] = None,
               workspace_src: str = WORKSPACE_TEMPLATE, repeat: int = 2) -> ConfigResult:
    """Score one config from its golden fixture ``doc``."""
    scenario = Scenario.from_json(doc["scenario"])
    expected = Observation.from_json(doc["expected"])
    res = ConfigResult(name=scenario.name)
    argv = list(argv or rp.REPLACEMENT_ARGV)
    env = rp.REPLACEMENT_ENV if extra_env is None else extra_env
    t0 = time.perf_counter()
    observations: List[Observation] = []
    for k in range(repeat):
        with Sandbox(workspace_src, scenario, input_overrides=doc["inputs"]) as sbx:
            obs = sbx.run(argv, env)
            observations.append(obs)
            if k == 0:
                diffs = expected.diff(obs)
                res.parity = not diffs
                res.details += ["parity: %s" % d for d in diffs]
                check_generated_tree(sbx.workspace, obs, res)
    res.deterministic = all(not observations[0].diff(o) for o in observations[1:])
    if not res.deterministic:
        res.details.append("determinism: two runs differ")
    res.wall_s = time.perf_counter() - t0
    return res


def golden_for(scenario: Scenario) -> dict:
    return rp.load_golden(os.path.join(rp.golden_dir(), scenario.name + ".json"))


def run_all(argv: Optional[Sequence[str]] = None, extra_env: Optional[Dict[str, str]] = None,
            workspace_src: str = WORKSPACE_TEMPLATE, repeat: int = 2) -> List[ConfigResult]:
    return [run_config(golden_for(sc), argv, extra_env, workspace_src, repeat) for sc in CONFIGS]


def replay_catalogue(docs: Optional[Sequence[dict]] = None) -> Dict[str, Any]:
    """The whole recorded parity surface, in-process (about 10 ms a scenario):
    {"replayed": n, "passed": n, "skipped": [...], "failed": [...]}."""
    if docs is None:
        docs = [rp.load_golden(p) for p in rp.list_goldens()]
    out: Dict[str, Any] = {"replayed": 0, "passed": 0, "skipped": [], "failed": []}
    for doc in docs:
        name = doc["scenario"]["name"]
        if rp.not_reproducible_here(doc):
            out["skipped"].append(name)
            continue
        out["replayed"] += 1
        try:
            expected, got = rp.replay_in_process(doc)
            ok = not expected.diff(got)
        except Exception:                      # a crash is a failure, not a reason to lose the line
            ok = False
        if ok:
            out["passed"] += 1
        else:
            out["failed"].append(name)
    return out


def pass_rate(results: Sequence[ConfigResult]) -> float:
    return sum(1 for r in results if r.passed) / float(len(results)) if results else 0.0


# ---------------------------------------------------------------------------
# config 5b: count scaling beyond the reference's domain (EXTENSION, labelled)
# ---------------------------------------------------------------------------
def extension_count_scaling(n_workers: int = 10, workspace_src: str = WORKSPACE_TEMPLATE) -> Dict[str, Any]:
    """BASELINE config 5 asks for 10 workers; the reference refuses anything but
    1-9 at the prompt (setup.sh:301) and has no ``count`` anywhere, so this is
    *not* reference behaviour and is not part of the parity score.  It drives
    the generator half of the path (setup.sh:138-198, 116-137) directly with
    ``KUBERNETES_NUMBER_OF_NODES=n`` -- the state ``exportVars`` would leave for
    an accepted answer -- and checks that the output scales linearly: n+1 module
    blocks with unique names, n+1 planned machines, an inventory of n+1 hosts,
    and the byte-size law measured at N=1,2."""
    from ..supervisor import Supervisor
    from .modeltools import ModelRuntime, ModelSandbox
    sizes: Dict[int, int] = {}
    out: Dict[str, Any] = {"n_workers": n_workers, "extension": True}
    for n in (1, 2, n_workers):
        sc = Scenario(name="extension_n%d" % n)
        with ModelSandbox(workspace_src, sc, input_overrides=golden_for(CONFIGS[0])["inputs"]) as sbx:
            rt = ModelRuntime(sc, sbx.workspace, sbx.home)
            sup = Supervisor(rt)
            sup.v.update({
                "SDC_URL": '"https://us-east-1.api.joyent.com"', "SDC_ACCOUNT": '"fixture"', "SDC_KEY_ID": '"aa:bb:cc:dd"',
                "SDC_KEY": '"~/.ssh/id_rsa"', "RANCHER_MASTER_NETWORKS": "31428241-4878-47d6-9fba-9a8436b596a4",
                "KUBERNETES_NODE_NETWORKS": "31428241-4878-47d6-9fba-9a8436b596a4", "KUBERNETES_NUMBER_OF_NODES": str(n),
                "KUBERNETES_NAME": '"k8s dev"', "KUBERNETES_DESCRIPTION": '"k8s dev"',
                "RANCHER_MASTER_HOSTNAME": '"kubemaster"', "KUBERNETES_NODE_HOSTNAME_BEGINSWITH": '"kubenode"',
                "HOST_PACKAGE": '"14b6fade-d0f8-11e5-85c5-4ff7918ab5c1"'})
            sup.run_terraform_tasks()
            sup.create_ansible_configs()
            text = rt.read_text("terraform/rancher.tf")
            sizes[n] = len(text.encode("utf-8"))
            if n == n_workers:
                obs = Observation(0, "", "", {"terraform/rancher.tf": text,
                                              "terraform/masters.ip": rt.read_text("terraform/masters.ip"),
                                              "terraform/hosts.ip": rt.read_text("terraform/hosts.ip")}, [], [])
                res = ConfigResult(name=sc.name)
                check_generated_tree(sbx.workspace, obs, res)
                names = expected_module_names(obs)
                out.update({"modules": len(names), "unique_names": len(set(names)) == len(names),
                            "terraform": res.terraform, "ansible": res.ansible, "inventory": res.inventory,
                            "planned_machines": res.n_modules, "inventory_hosts": res.n_hosts, "details": res.details})
    per_worker = sizes[2] - sizes[1]
    out["rancher_tf_bytes"] = sizes[n_workers]
    out["size_law"] = {"base": sizes[1] - per_worker, "per_worker": per_worker,
                       "predicted": sizes[1] + per_worker * (n_workers - 1)}
    # worker names get one character longer from 10 on ("kubenode10"): 2 bytes per block (label + hostname)
    extra = 2 * sum(len(str(i)) - 1 for i in range(1, n_workers + 1))
    out["linear"] = sizes[n_workers] == out["size_law"]["predicted"] + extra
    out["ok"] = bool(out.get("terraform") and out.get("ansible") and out.get("inventory") and out["linear"]
                     and out.get("unique_names") and out.get("modules") == n_workers + 1
                     and out.get("planned_machines") == n_workers + 1 and out.get("inventory_hosts") == n_workers + 1)
    return out
