"""Offline tools around the path, as one command::

    python -m tritonk8ssupervisor_b200.tools <command> [options]

    init DIR               lay down the package's workspace (terraform/, ansible/) in DIR
    validate [DIR]         `terraform validate` stand-in on DIR/terraform        (exit 1 on errors)
    plan [DIR]             resource instances + resolved inputs, as JSON
    apply-mock [DIR]       `terraform apply` against a mocked Triton provider (writes masters.ip/hosts.ip)
    syntax-check [DIR]     `ansible-playbook --syntax-check` stand-in on DIR/ansible/clusterUp.yml
    inventory [DIR]        `ansible-inventory -i DIR/ansible/hosts --list`
    lint [DIR]             advisory modern-dialect report (never fails)
    dry-run [DIR]          run the playbook against simulated machines + the Rancher mock
    conformance            BASELINE.json's metric: the five configs, pass rate

DIR defaults to the current directory (the workspace root, where setup runs).
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import sys
from typing import List, Optional

from . import ansiblecheck, inventory, lintmodern, tfvalidate
from .report import Report

TEMPLATE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "workspace")


def _print_report(rep: Report, out) -> None:
    for f in rep.findings:
        out.write("%s\n" % f)
    out.write("%d error(s), %d warning(s)\n" % (len(rep.errors), len(rep.warnings)))


def _inventory(root: str) -> Optional[inventory.Inventory]:
    p = os.path.join(root, "ansible", "hosts")
    return inventory.parse_file(p) if os.path.isfile(p) else None


def main(argv: Optional[List[str]] = None, out=sys.stdout) -> int:
    ap = argparse.ArgumentParser(prog="tritonk8ssupervisor_b200.tools", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("command", choices=["init", "validate", "plan", "apply-mock", "syntax-check", "inventory", "lint", "dry-run", "conformance"])
    ap.add_argument("dir", nargs="?", default=".")
    args = ap.parse_args(argv)
    root = args.dir

    if args.command == "init":
        os.makedirs(root, exist_ok=True)
        for sub in ("terraform", "ansible"):
            dst = os.path.join(root, sub)
            if os.path.exists(dst):
                out.write("refusing to overwrite %s\n" % dst)
                return 1
            shutil.copytree(os.path.join(TEMPLATE, sub), dst)
        out.write("workspace laid down in %s; run `python -m tritonk8ssupervisor_b200` from there\n" % os.path.abspath(root))
        return 0
    if args.command == "validate":
        rep = tfvalidate.validate(os.path.join(root, "terraform"))
        _print_report(rep, out)
        return 0 if rep.ok else 1
    if args.command == "plan":
        json.dump(tfvalidate.plan(os.path.join(root, "terraform")), out, indent=1)
        out.write("\n")
        return 0
    if args.command == "apply-mock":
        from . import tfmock
        machines, rep = tfmock.apply(os.path.join(root, "terraform"))
        for m in machines:
            out.write("%s: created (primaryip %s)\n" % (m.address, m.computed.get("primaryip")))
        _print_report(rep, out)
        return 0 if rep.ok else 1
    if args.command == "syntax-check":
        try:
            inv = _inventory(root)
        except inventory.InventoryError as exc:
            out.write("error[inventory] %s\n" % exc)
            return 1
        rep = ansiblecheck.check_playbook(os.path.join(root, "ansible", "clusterUp.yml"), inv)
        rep.findings += ansiblecheck.check_ansible_cfg(os.path.join(root, "ansible", "ansible.cfg")).findings
        _print_report(rep, out)
        return 0 if rep.ok else 1
    if args.command == "inventory":
        try:
            inv = _inventory(root)
        except inventory.InventoryError as exc:
            out.write("error[inventory] %s\n" % exc)
            return 1
        if inv is None:
            out.write("error[inventory] %s not found\n" % os.path.join(root, "ansible", "hosts"))
            return 1
        json.dump(inv.to_list_json(), out, indent=4, sort_keys=False)
        out.write("\n")
        return 0
    if args.command == "lint":
        for f in lintmodern.lint_workspace(root).findings:
            out.write("%s\n" % f)
        return 0
    if args.command == "dry-run":
        return _dry_run(root, out)
    from .conformance import harness
    results = harness.run_all()
    for r in results:
        out.write("%-40s %s\n" % (r.name, "PASS" if r.passed else "FAIL  " + "; ".join(r.details)[:300]))
    out.write("pass rate: %d/%d\n" % (sum(r.passed for r in results), len(results)))
    return 0 if all(r.passed for r in results) else 1


def _dry_run(root: str, out) -> int:
    import tempfile

    import yaml

    from . import ranchermock, roledryrun
    inv = _inventory(root)
    if inv is None or not inv.hosts_of("MASTER"):
        out.write("error: %s/ansible/hosts with a [MASTER] group is needed (run setup first)\n" % root)
        return 1
    tmp = tempfile.mkdtemp(prefix="tk8s-dry-")
    try:
        work = shutil.copytree(os.path.join(root, "ansible"), os.path.join(tmp, "ansible"))
        os.makedirs(os.path.join(work, "tmp"), exist_ok=True)
        with open(os.path.join(work, "roles", "ranchermaster", "vars", "vars.yml"), encoding="utf-8") as fh:
            master = str((yaml.safe_load(fh) or {}).get("master", inv.hosts_of("MASTER")[0]))
        with ranchermock.RancherMock(public_host=master) as mock:
            run = roledryrun.DryRun(os.path.join(work, "clusterUp.yml"), inv, {"%s:8080" % master: mock.address}).run()
            for line in mock.wire_log():
                out.write("wire: %s\n" % line)
        for host, counts in run.recap().items():
            out.write("%-16s ok=%d changed=%d skipped=%d failed=%d\n" % (host, counts["ok"], counts["changed"],
                                                                           counts["skipped"], counts["failed"]))
        for host, why in run.failed_hosts.items():
            out.write("FAILED %s\n" % why)
        return 1 if run.failed_hosts else 0
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    sys.exit(main())
