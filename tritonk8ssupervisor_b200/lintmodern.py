"""Advisory lint: what present-day Terraform (>= 0.12/1.x) and Ansible (>= 2.9)
would reject or deprecate in a workspace written for Terraform <= 0.9 and
Ansible ~2.2 (SURVEY.md §8f rank 4, §7.4 "Era dialect").

It never changes what the path generates and never fails a config: "validate
passes" in this repository means "passes the era-correct rules" (tfvalidate,
ansiblecheck); this report tells a reader *why* the qualifier is needed.  The
statements about newer releases come from their upgrade guides / changelogs
(third-party, not verifiable offline -- SURVEY.md §7.4), so every item is
"advice", never "error".

Codes
  TF-SOURCE     module source without ./ (0.12+: local paths must start with ./ or ../)
  TF-TYPE       quoted type keyword ("list") instead of a type expression (list(string))
  TF-INTERP     attribute that is only "${expr}" (0.12+: write the bare expression)
  TF-DUPATTR    attribute given twice in one block (HCL2: error)
  TF-SELFREF    resource refers to itself by name (0.12+: cycle; use self.)
  TF-PROVIDER   no required_providers source for a non-hashicorp provider (0.13+)
  TF-CLI        command-line usage of the path that changed: `terraform get` without
                `init`, unattended `apply` (needs -auto-approve), `destroy -force`
  AN-WHEN       when/until wrapped in {{ }} (deprecated 2.3)
  AN-MODULE     module removed or renamed (docker -> docker_container)
  AN-HEADER     uri HEADER_* arguments (removed 2.9; use headers:)
  AN-KV         key=value argument strings (still accepted; YAML mappings preferred)
  AN-LOOPPKG    package module inside with_items (squashing deprecated 2.7; pass a list)
  AN-PY2        target bootstrap installs Python 2 (python-minimal)
  AN-SHADOW     play var shadowing a magic variable
"""
from __future__ import annotations

import os
import re
from typing import Any

from . import ansiblecheck as ac
from . import hcl1
from .hcl1 import Body, HVar, Literal
from .report import Finding, Report

CLI_ADVICE = [
    ("TF-CLI", "setup.sh:156", "`terraform get` alone no longer prepares a working directory: `terraform init` (0.10+) "
                               "fetches modules and the provider plugin"),
    ("TF-CLI", "setup.sh:157", "`terraform apply` asks for confirmation since 0.11; an unattended run needs -auto-approve"),
    ("TF-CLI", "setup.sh:501", "`terraform destroy -force` was replaced by -auto-approve (0.11.4) and removed in 0.15"),
]


def _advice(report: Report, code: str, where: str, msg: str) -> None:
    report.findings.append(Finding("advice", code, where, msg))


def _interp_only(value: Any) -> bool:
    return isinstance(value, Literal) and value.kind == "string" and value.parts is not None \
        and len(value.parts) == 1 and not isinstance(value.parts[0], str)


def _lint_body(body: Body, where: str, report: Report, resource=None) -> None:
    seen = {}
    for it in body.items:
        at = "%s:%d" % (where, it.line)
        if it.name in seen and it.name not in ("provisioner", "connection", "nic"):
            _advice(report, "TF-DUPATTR", at, "%r is given twice in this block (first at line %d); HCL2 rejects that"
                    % (it.name, seen[it.name]))
        seen.setdefault(it.name, it.line)
        if _interp_only(it.value):
            _advice(report, "TF-INTERP", at, '%s = "${…}" is interpolation-only; 0.12+ wants the bare expression' % it.name)
        if resource is not None and not isinstance(it.value, Body):
            for parts, line in hcl1.templates_in(it.value):
                for node in hcl1.walk_hil(parts):
                    if isinstance(node, HVar) and tuple(node.path[:2]) == resource:
                        _advice(report, "TF-SELFREF", "%s:%d" % (where, line),
                                "${%s} refers to the enclosing resource; 0.12+ reports a cycle, use self.%s"
                                % (".".join(node.path), ".".join(node.path[2:])))
        if isinstance(it.value, Body):
            _lint_body(it.value, where, report, resource)


def lint_terraform(tf_dir: str) -> Report:
    """Root directory and every local module directory below it."""
    report = Report()
    dirs = [tf_dir] + sorted(os.path.join(tf_dir, d) for d in os.listdir(tf_dir)
                             if os.path.isdir(os.path.join(tf_dir, d)) and not d.startswith("."))
    providers, required = set(), False
    for d in dirs:
        for name in sorted(n for n in os.listdir(d) if n.endswith(".tf")):
            path = os.path.join(d, name)
            rel = os.path.relpath(path, tf_dir)
            try:
                body = hcl1.parse_file(path)
            except hcl1.HCLError as exc:
                _advice(report, "TF-SYNTAX", rel, str(exc))
                continue
            for it in body.items:
                at = "%s:%d" % (rel, it.line)
                if it.name == "module" and isinstance(it.value, Body):
                    src = it.value.attr("source")
                    text = "".join(p for p in (src.value.parts or []) if isinstance(p, str)) if src is not None and isinstance(src.value, Literal) else ""
                    if text and not text.startswith(("./", "../", "/")) and "/" not in text and "::" not in text:
                        _advice(report, "TF-SOURCE", at, 'source = "%s": 0.12+ reads a bare name as a registry address; '
                                                         'local modules need "./%s"' % (text, text))
                    _lint_body(it.value, rel, report)
                elif it.name == "variable" and isinstance(it.value, Body):
                    t = it.value.attr("type")
                    if t is not None and isinstance(t.value, Literal) and t.value.kind == "string":
                        new = {"list": "list(string)", "map": "map(string)", "string": "string"}.get(t.value.value, "?")
                        _advice(report, "TF-TYPE", "%s:%d" % (rel, t.line), 'type = "%s" is the 0.11 spelling; 0.12+ wants type = %s'
                                % (t.value.value, new))
                elif it.name == "resource" and isinstance(it.value, Body) and len(it.labels) == 2:
                    _lint_body(it.value, rel, report, resource=(it.labels[0], it.labels[1]))
                    providers.add(it.labels[0].split("_", 1)[0])
                elif it.name == "provider" and isinstance(it.value, Body):
                    providers.add(it.labels[0] if it.labels else "?")
                    _lint_body(it.value, rel, report)
                elif it.name == "terraform" and isinstance(it.value, Body) and it.value.blocks("required_providers"):
                    required = True
    for p in sorted(providers - {"?"}):
        if not required and p in ("triton",):
            _advice(report, "TF-PROVIDER", "terraform", 'provider "%s" is not in the hashicorp namespace: 0.13+ needs '
                                                        'required_providers { %s = { source = "joyent/%s" } }' % (p, p, p))
    for code, where, msg in CLI_ADVICE:
        _advice(report, code, where, msg)
    return report


_PKG_MODULES = {"apt", "yum", "dnf", "package", "pip", "apk", "zypper"}


def lint_ansible(playbook: str) -> Report:
    report = Report()
    base = os.path.dirname(os.path.abspath(playbook))
    loader_report = Report()
    plays = ac.load_yaml(playbook, loader_report) or []
    files = []
    for pidx, play in enumerate(plays if isinstance(plays, list) else [], 1):
        for k in (play.get("vars") or {}):
            if k in ac.MAGIC_VARS:
                _advice(report, "AN-SHADOW", "%s play %d" % (os.path.basename(playbook), pidx),
                        "vars: %s is a magic variable; the play var is ignored" % k)
        for entry in play.get("roles") or []:
            name = ac._role_name(entry)
            f = os.path.join(base, "roles", str(name), "tasks", "main.yml")
            if os.path.isfile(f) and f not in files:
                files.append(f)
    for f in files:
        rel = os.path.relpath(f, base)
        for idx, task in enumerate(ac.load_yaml(f, loader_report) or [], 1):
            if not isinstance(task, dict):
                continue
            at = "%s task %d" % (rel, idx)
            for key in ("when", "until", "changed_when", "failed_when"):
                if isinstance(task.get(key), str) and "{{" in task[key]:
                    _advice(report, "AN-WHEN", at, "%s: %r -- conditionals are expressions already; the braces are deprecated" % (key, task[key]))
            act = ac._task_action(task, at, Report())
            if act is None:
                continue
            module, args, _ = act
            if module == "docker":
                _advice(report, "AN-MODULE", at, "the `docker` module was deprecated in 2.1 and removed in 2.4: use docker_container")
            if module == "uri" and any(k.startswith("HEADER_") for k in args):
                _advice(report, "AN-HEADER", at, "uri HEADER_* arguments were removed in 2.9: use headers: {Name: value}")
            spec = task.get(module, task.get("action", task.get("local_action")))
            if isinstance(spec, str) and module not in ac.FREE_FORM and "=" in spec:
                _advice(report, "AN-KV", at, "%s arguments given as a key=value string; a YAML mapping avoids quoting traps" % module)
            if module in _PKG_MODULES and any(k.startswith("with_") for k in task) and "item" in str(args.get("name", "")):
                _advice(report, "AN-LOOPPKG", at, "looping a package module over items is squashed only until 2.11: pass name: [list]")
    return report


def lint_bootstrap(tf_dir: str) -> Report:
    """The remote-exec bootstrap lines (terraform/*/main.tf:21-26)."""
    report = Report()
    for d, _, names in os.walk(tf_dir):
        for n in sorted(names):
            if n.endswith(".tf"):
                with open(os.path.join(d, n), encoding="utf-8") as fh:
                    for ln, line in enumerate(fh, 1):
                        if re.search(r"\bpython-minimal\b", line):
                            _advice(report, "AN-PY2", "%s:%d" % (os.path.relpath(os.path.join(d, n), tf_dir), ln),
                                    "python-minimal is Python 2 on Ubuntu 16.04; Ansible >= 2.12 needs Python 3 on targets")
    return report


def lint_workspace(root: str) -> Report:
    """Everything, for a workspace root holding terraform/ and ansible/."""
    report = Report()
    tf = os.path.join(root, "terraform")
    if os.path.isdir(tf):
        report.findings += lint_terraform(tf).findings + lint_bootstrap(tf).findings
    pb = os.path.join(root, "ansible", "clusterUp.yml")
    if os.path.isfile(pb):
        report.findings += lint_ansible(pb).findings
    return report
