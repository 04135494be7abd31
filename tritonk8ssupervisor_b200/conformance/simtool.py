"""Simulated ``terraform`` and ``ansible-playbook``: the stub contract with the
behaviour *derived* instead of scripted.

The shell stub hard-codes what the tools below the offline line leave behind
(IP files, an environment id).  Installed instead of it (``Scenario.sim_tools``),
this program produces the same artefacts by actually interpreting the workspace:

* ``terraform get|apply|destroy`` -> :mod:`tritonk8ssupervisor_b200.tfmock`
  (validate, then apply against the mocked Triton provider: the modules' own
  ``local-exec`` lines write ``masters.ip`` / ``hosts.ip``);
* ``ansible-playbook -i hosts clusterUp.yml`` -> a
  :class:`~tritonk8ssupervisor_b200.ranchermock.RancherMock` is started and the
  playbook is dry-run against it with :mod:`tritonk8ssupervisor_b200.roledryrun`
  (real Jinja, real HTTP to the mock); the environment id the mock hands out is
  what ends up in ``tmp/kubernetes_environment.id`` and in setup's banner.

Every call is logged in the stub's format, so observations stay comparable.  Both
the reference script and the replacement can be run on top of it; the chain
setup -> rancher.tf -> mock apply -> inventory -> playbook dry run -> banner then
holds together with nothing scripted in between (tests/test_simulated_stack.py).
"""
from __future__ import annotations

import os
import re
import sys

_ENV_LOGGED = re.compile(r"^(SDC_|KUBERNETES_|RANCHER_|HOST_PACKAGE=|ANSIBLE_|TRITON_)")


def _log(tool: str, args) -> None:
    state = os.environ["TK8S_STUB_STATE"]
    root = os.environ.get("TK8S_STUB_ROOT", os.getcwd())
    cwd = os.getcwd()
    rel = "." if cwd == root else ("./" + os.path.relpath(cwd, root) if cwd.startswith(root + os.sep) else cwd)
    with open(os.path.join(state, "calls.log"), "a", encoding="utf-8") as fh:
        fh.write("%s|%s|%s\n" % (tool, rel, " ".join(args)))
        for line in sorted("%s=%s" % kv for kv in os.environ.items()):
            if _ENV_LOGGED.match(line):
                fh.write("  env:%s\n" % line)


def terraform(args) -> int:
    from .. import tfmock, tfvalidate
    sub = args[0] if args else ""
    if sub == "get":
        rep = tfvalidate.validate(".")
        for f in rep.errors:
            print("Error loading modules: %s" % f)
        print("Get: %d module(s)" % len(tfvalidate.load_module(".", tfvalidate.Report()).modules))
        return 0 if rep.ok else 1
    if sub == "apply":
        machines, rep = tfmock.apply(".")
        import json
        boot = {}
        for m in machines:
            print("%s: Creation complete (primaryip = %s)" % (m.address, m.computed["primaryip"]))
            boot[m.computed["primaryip"]] = tfmock.bootstrap_state(m)
        with open(os.path.join(os.environ["TK8S_STUB_STATE"], "bootstrap.json"), "w", encoding="utf-8") as fh:
            json.dump(boot, fh)
        for f in rep.errors:
            print("Error applying plan: %s" % f)
        if rep.ok:
            print("Apply complete! Resources: %d added, 0 changed, 0 destroyed." % len(machines))
        return 0 if rep.ok else 1
    if sub == "destroy":
        print("Destroy complete!")
        return 0
    sys.stderr.write("terraform (simulated): unsupported: %s\n" % " ".join(args))
    return 2


def ansible_playbook(args) -> int:
    import yaml

    from .. import inventory, ranchermock, roledryrun
    inv_path, playbook = "hosts", None
    it = iter(args)
    for a in it:
        if a == "-i":
            inv_path = next(it, "hosts")
        elif not a.startswith("-"):
            playbook = a
    if playbook is None:
        sys.stderr.write("ansible-playbook (simulated): no playbook given\n")
        return 2
    inv = inventory.parse_file(inv_path)
    with open(os.path.join("roles", "ranchermaster", "vars", "vars.yml"), encoding="utf-8") as fh:
        master = str((yaml.safe_load(fh) or {}).get("master", ""))
    os.makedirs("tmp", exist_ok=True)
    # machines are usable only if terraform's bootstrap made them so (tfmock.bootstrap_state)
    unreachable = {}
    boot_file = os.path.join(os.environ["TK8S_STUB_STATE"], "bootstrap.json")
    if os.path.isfile(boot_file):
        import json
        with open(boot_file, encoding="utf-8") as fh:
            boot = json.load(fh)
        for host in inv.host_order:
            st = boot.get(host)
            if st is None:
                unreachable[host] = "%s: UNREACHABLE: no such machine was created" % host
            elif not st["root_login"]:
                unreachable[host] = "%s: UNREACHABLE: Permission denied (publickey) for root" % host
            elif not st["python"]:
                unreachable[host] = "%s: MODULE FAILURE: /usr/bin/python: not found" % host
    with ranchermock.RancherMock(public_host=master) as mock:
        run = roledryrun.DryRun(playbook, inv, {"%s:8080" % master: mock.address}, unreachable=unreachable).run()
    print("PLAY RECAP *********************************************************************")
    for host, c in run.recap().items():
        print("%-26s : ok=%-4d changed=%-4d unreachable=0    failed=%d" % (host, c["ok"], c["changed"], c["failed"]))
    for why in run.failed_hosts.values():
        print("fatal: %s" % why)
    return 2 if run.failed_hosts else 0


def main() -> int:
    tool, args = sys.argv[1], sys.argv[2:]
    _log(tool, args)
    if tool == "terraform":
        return terraform(args)
    if tool == "ansible-playbook":
        return ansible_playbook(args)
    sys.stderr.write("simtool: no behaviour for %s\n" % tool)
    return 127


if __name__ == "__main__":
    sys.exit(main())
