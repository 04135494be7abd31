"""``terraform apply`` against a mocked Triton provider (BASELINE config 5:
"terraform plan against a mocked triton provider (count scaling)").

The real thing needs the terraform binary, the Go plugin SDK and CloudAPI
(unavailable on this box).  What the *path* depends on is much
smaller -- SURVEY.md §4: "`apply` writes one line to masters.ip and one line per
`source = "host"` module to hosts.ip, which is precisely the contract of
main.tf:29-31" -- and that contract is written in the module files themselves:
a ``local-exec`` provisioner whose command appends the machine's ``primaryip``
to a file.  So instead of hard-coding it (as the shell stub does), this mock
*interprets the modules*: it walks the planned resources in file order, lets a
fake provider assign the computed attributes (``id``, ``primaryip``), resolves
``${self.*}`` / ``${triton_machine.NAME.*}`` / ``${var.*}`` in the connection and
provisioner blocks, records the ``remote-exec`` scripts it would have run over
SSH, and really runs the ``local-exec`` commands (``/bin/sh -c`` in the
terraform directory, as terraform does).

Deterministic where the real tool is not: terraform applies independent modules
concurrently, so the order of lines in hosts.ip is arbitrary in a real run
(SURVEY.md §5 "latent races"); here it is file order.
"""
from __future__ import annotations

import dataclasses
import os
import subprocess
from typing import Any, Callable, Dict, List, Optional, Tuple

from . import hcl1, tfvalidate
from .hcl1 import Body, HCall, HVar, ListVal, Literal
from .report import Report


@dataclasses.dataclass
class Machine:
    address: str                      # module.kubenode1.triton_machine.host
    module: str
    rtype: str
    rname: str
    args: Dict[str, Any]              # resolved arguments sent to the provider
    computed: Dict[str, Any]          # what the provider answered (id, primaryip, ...)
    remote_exec: List[Dict[str, Any]] = dataclasses.field(default_factory=list)   # {"connection": {...}, "inline": [...]}
    local_exec: List[str] = dataclasses.field(default_factory=list)               # commands that were run


def default_allocator() -> Callable[[str, str, Dict[str, Any]], Dict[str, Any]]:
    """The fake provider: masters get 10.0.0.N, everything else 10.0.1.N -- the
    addresses the shell stub uses, so both ways of faking `apply` agree."""
    counters = {"master": 0, "other": 0}

    def alloc(rtype: str, rname: str, args: Dict[str, Any]) -> Dict[str, Any]:
        kind = "master" if rname == "master" else "other"
        counters[kind] += 1
        ip = "10.0.0.%d" % counters[kind] if kind == "master" else "10.0.1.%d" % counters[kind]
        n = counters["master"] + counters["other"]
        return {"id": "00000000-0000-4000-8000-%012d" % n, "primaryip": ip, "ips": [ip], "state": "running",
                "dataset": "ubuntu-certified-16.04", "type": "virtualmachine"}
    return alloc


class _Scope:
    def __init__(self, variables: Dict[str, Any], resources: Dict[Tuple[str, str], Dict[str, Any]],
                 self_attrs: Optional[Dict[str, Any]], module_dir: str, root_dir: str):
        self.variables, self.resources, self.self_attrs = variables, resources, self_attrs
        self.module_dir, self.root_dir = module_dir, root_dir


def _eval(node: Any, sc: _Scope) -> Any:
    if isinstance(node, str):
        return node
    if isinstance(node, hcl1.HLit):
        return node.value
    if isinstance(node, hcl1.HStr):
        return "".join(str(_eval(p, sc)) for p in node.parts)
    if isinstance(node, HVar):
        p = node.path
        if p[0] == "var" and len(p) >= 2 and p[1] in sc.variables:
            return sc.variables[p[1]]
        if p[0] == "self" and sc.self_attrs is not None and len(p) == 2 and p[1] in sc.self_attrs:
            return sc.self_attrs[p[1]]
        if p[0] == "path" and len(p) == 2:
            return {"module": sc.module_dir, "root": sc.root_dir, "cwd": sc.root_dir}.get(p[1], "<computed>")
        if len(p) >= 3 and (p[0], p[1]) in sc.resources and p[2] in sc.resources[(p[0], p[1])]:
            return sc.resources[(p[0], p[1])][p[2]]
        return "<computed>"
    if isinstance(node, HCall):
        args = [_eval(a, sc) for a in node.args]
        if node.func == "file" and len(args) == 1:
            return '${file("%s")}' % args[0]          # key material is never read by the mock
        if node.func == "join" and len(args) == 2 and isinstance(args[1], list):
            return str(args[0]).join(str(x) for x in args[1])
        if node.func == "length" and len(args) == 1 and isinstance(args[0], (list, str, dict)):
            return len(args[0])
        if node.func == "element" and len(args) == 2 and isinstance(args[0], list) and args[0]:
            return args[0][int(args[1]) % len(args[0])]
        if node.func in ("lower", "upper") and len(args) == 1:
            return getattr(str(args[0]), node.func)()
        if node.func == "format" and args:
            try:
                return str(args[0]) % tuple(args[1:])
            except (TypeError, ValueError):
                return "<computed>"
        return "<computed>"
    if isinstance(node, hcl1.HBin):
        l, r = _eval(node.left, sc), _eval(node.right, sc)
        try:
            return {"+": lambda: l + r, "-": lambda: l - r, "*": lambda: l * r, "/": lambda: l / r, "%": lambda: l % r,
                    "==": lambda: l == r, "!=": lambda: l != r, "<": lambda: l < r, ">": lambda: l > r,
                    "<=": lambda: l <= r, ">=": lambda: l >= r, "&&": lambda: bool(l) and bool(r),
                    "||": lambda: bool(l) or bool(r)}[node.op]()
        except Exception:
            return "<computed>"
    if isinstance(node, hcl1.HCond):
        return _eval(node.then if _eval(node.cond, sc) else node.other, sc)
    if isinstance(node, hcl1.HUnary):
        v = _eval(node.operand, sc)
        return (not v) if node.op == "!" else (-v if isinstance(v, (int, float)) else "<computed>")
    if isinstance(node, hcl1.HIndex):
        t, k = _eval(node.target, sc), _eval(node.key, sc)
        try:
            return t[k]
        except Exception:
            return "<computed>"
    return "<computed>"


def _value(value: Any, sc: _Scope) -> Any:
    if isinstance(value, ListVal):
        return [_value(v, sc) for v in value.items]
    if isinstance(value, Body):
        out: Dict[str, Any] = {}
        for it in value.items:
            out[it.name] = _value(it.value, sc)
        return out
    if isinstance(value, Literal):
        if value.kind != "string" or value.parts is None:
            return value.value
        if len(value.parts) == 1 and not isinstance(value.parts[0], str):
            return _eval(value.parts[0], sc)           # "${var.networks}" keeps its list type
        return "".join(str(_eval(p, sc)) for p in value.parts)
    return value


def apply(root_dir: str, allocator: Optional[Callable[[str, str, Dict[str, Any]], Dict[str, Any]]] = None,
          run_local_exec: bool = True) -> Tuple[List[Machine], Report]:
    """Mock ``terraform apply`` in ``root_dir``.  Returns the machines "created"
    and the validation report; nothing is applied if validation has errors."""
    report = tfvalidate.validate(root_dir)
    machines: List[Machine] = []
    if not report.ok:
        return machines, report
    alloc = allocator or default_allocator()
    root = tfvalidate.load_module(root_dir, Report())
    for mname, item in root.modules.items():
        body: Body = item.value
        src = "".join(p for p in (body.attr("source").value.parts or []) if isinstance(p, str))
        mdir = os.path.normpath(os.path.join(root_dir, src))
        child = tfvalidate.load_module(mdir, Report())
        root_scope = _Scope({}, {}, None, root_dir, root_dir)
        variables: Dict[str, Any] = {n: _value(v.default, root_scope) for n, v in child.variables.items() if v.has_default}
        for a in body.items:
            if a.name != "source":
                variables[a.name] = _value(a.value, root_scope)
        created: Dict[Tuple[str, str], Dict[str, Any]] = {}
        for (rtype, rname), res in child.resources.items():
            sc = _Scope(variables, created, None, mdir, root_dir)
            args: Dict[str, Any] = {}
            for a in res.value.items:
                if a.name in ("provisioner", "connection"):
                    continue
                v = _value(a.value, sc)
                args[a.name] = {**args[a.name], **v} if isinstance(v, dict) and isinstance(args.get(a.name), dict) else v
            computed = alloc(rtype, rname, args)
            attrs = {**args, **computed}
            created[(rtype, rname)] = attrs
            m = Machine("module.%s.%s.%s" % (mname, rtype, rname), mname, rtype, rname, args, computed)
            psc = _Scope(variables, created, attrs, mdir, root_dir)
            res_conn = res.value.attr("connection")
            for prov in res.value.attrs("provisioner"):
                ptype = prov.labels[0]
                pbody: Body = prov.value
                conn_item = pbody.attr("connection") or res_conn
                if ptype == "remote-exec":
                    m.remote_exec.append({"connection": _value(conn_item.value, psc) if conn_item is not None else {},
                                          "inline": _value(pbody.attr("inline").value, psc) if pbody.attr("inline") else []})
                elif ptype == "local-exec":
                    cmd = str(_value(pbody.attr("command").value, psc))
                    m.local_exec.append(cmd)
                    if run_local_exec:
                        p = subprocess.run(["/bin/sh", "-c", cmd], cwd=root_dir, capture_output=True, text=True)
                        if p.returncode != 0:
                            report.error("local-exec", m.address, "command %r exited %d: %s" % (cmd, p.returncode, p.stderr.strip()))
                            return machines + [m], report
            machines.append(m)
    return machines, report


# ---------------------------------------------------------------------------
# what the remote-exec bootstrap leaves on the machine
# ---------------------------------------------------------------------------
def bootstrap_state(machine: Machine) -> Dict[str, Any]:
    """Interpret the ``remote-exec`` scripts of a machine against a small model
    of the ``ubuntu-certified-16.04`` image (docs/manual-setup.md:9-40): only the
    user ``ubuntu`` can log in and holds the keys given as ``root_authorized_keys``,
    root has none, there is no python, and the apt lists are empty until
    ``apt-get update``.  The playbook needs the opposite (``remote_user: root``,
    ansible/clusterUp.yml:4,10,20, and python for every module), so this says
    whether the bootstrap lines of main.tf:21-26 achieve it."""
    st: Dict[str, Any] = {"root_login": False, "python": False, "apt_updated": False, "waited_s": 0,
                          "packages": [], "errors": [], "unmodelled": []}
    for script in machine.remote_exec:
        conn = script.get("connection") or {}
        if conn.get("user") != "ubuntu":
            st["errors"].append("connection as %r: the image only admits 'ubuntu' before the bootstrap" % conn.get("user"))
            continue
        if conn.get("host") != machine.computed.get("primaryip"):
            st["errors"].append("connection host %r is not the machine's primaryip" % conn.get("host"))
            continue
        if not conn.get("private_key"):
            st["errors"].append("connection without private_key (agent is off)")
            continue
        for cmd in script.get("inline") or []:
            words = str(cmd).split()
            if words[:1] == ["sudo"]:
                words = words[1:]
            if words[:1] == ["sleep"] and len(words) == 2 and words[1].isdigit():
                st["waited_s"] += int(words[1])
            elif words[:1] in (["cp"], ["install"]):
                paths = [w for w in words[1:] if w.startswith("/")]
                if len(paths) == 2 and paths[0] == "/home/ubuntu/.ssh/authorized_keys" \
                        and paths[1] in ("/root/.ssh/", "/root/.ssh", "/root/.ssh/authorized_keys"):
                    st["root_login"] = True
                else:
                    st["unmodelled"].append(cmd)
            elif words[:2] == ["apt-get", "update"]:
                st["apt_updated"] = True
            elif words[:2] == ["apt-get", "install"]:
                pkgs = [w for w in words[2:] if not w.startswith("-")]
                if "-y" not in words and "--yes" not in words:
                    st["errors"].append("%r would wait for confirmation (no -y)" % cmd)
                elif not st["apt_updated"]:
                    st["errors"].append("%r before apt-get update: package lists are empty on a fresh image" % cmd)
                else:
                    st["packages"] += pkgs
                    if any(p in ("python-minimal", "python", "python2.7") for p in pkgs):
                        st["python"] = True
            else:
                st["unmodelled"].append(cmd)
    st["ansible_ready"] = bool(st["root_login"] and st["python"] and not st["errors"])
    return st
