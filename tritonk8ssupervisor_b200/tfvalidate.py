"""Structural validation of a Terraform (<= 0.9 dialect) root configuration and
the local modules it instantiates -- the offline stand-in for ``terraform
validate`` / ``terraform plan`` on this path.

Real ``terraform`` is unavailable on this box, and current
releases would reject the reference's dialect anyway, so "validate passes" here
means "passes the era-correct rules", stated below.  The Triton provider's
schema is third-party code that the reference neither vendors nor pins; the
attribute tables here restate what the reference itself uses plus the
provider's documented arguments, and anything outside them is a *warning*, never
an error ("parity unpinned" for provider acceptance).

Errors (E) and warnings (W):

  E syntax            any file fails to parse
  E module-source     ``source`` missing, not a string literal, or not a directory
  E module-dup        two modules (or two resources of one type) share a name
  E arg-unknown       a module argument names no ``variable`` of the module
  E arg-missing       a ``variable`` without default is not supplied
  E arg-type          list given to a string variable / string to a list variable
  E var-undeclared    ``${var.X}`` with no ``variable "X"``
  E ref-unknown       ``${type.name.attr}`` names no resource of the module;
                      ``${module.X.out}`` names no module/output
  E self-outside      ``self.*`` used outside provisioner/connection
  E func-unknown      interpolation calls an unknown function
  E func-arity        ``file()`` takes exactly one argument
  E variable-shape    ``variable`` block with unknown key or bad ``type``
  E provisioner       unknown provisioner type / missing required argument
  W attr-unknown      resource/provider attribute outside the known tables
  W attr-duplicate    attribute repeated in one block (the reference's ``tags``)
  W self-reference    resource refers to itself by name inside a provisioner
                      (accepted by <= 0.9, ``self.`` is the supported form)
  W count-index       ``count.index`` in a resource without ``count``
"""
from __future__ import annotations

import dataclasses
import os
from typing import Any, Dict, List, Optional, Tuple

from . import hcl1
from .hcl1 import Body, HCall, HVar, Item, ListVal, Literal
from .report import Report

# interpolation functions of Terraform 0.9 (config/interpolate_funcs.go)
FUNCTIONS = {
    "basename", "base64decode", "base64encode", "base64sha256", "ceil", "chomp", "cidrhost", "cidrnetmask",
    "cidrsubnet", "coalesce", "coalescelist", "compact", "concat", "dirname", "distinct", "element", "file",
    "floor", "format", "formatlist", "index", "join", "jsonencode", "keys", "length", "list", "lookup", "lower",
    "map", "matchkeys", "max", "md5", "merge", "min", "pathexpand", "replace", "sha1", "sha256", "signum",
    "slice", "sort", "split", "substr", "timestamp", "title", "trimspace", "upper", "uuid", "values", "zipmap",
}

PROVIDER_ARGS = {
    "triton": {"account", "key_material", "key_id", "url", "insecure_skip_tls_verify", "user", "alias", "version"},
}

#: triton_machine arguments (what the reference uses + the provider's documented ones) and exported attributes
RESOURCE_ARGS = {
    "triton_machine": {"name", "package", "image", "tags", "networks", "root_authorized_keys", "user_script",
                       "user_data", "firewall_enabled", "nic", "metadata", "administrator_pw", "cloud_config",
                       "affinity", "cns", "locality", "deletion_protection_enabled"},
}
RESOURCE_ATTRS = {
    "triton_machine": RESOURCE_ARGS["triton_machine"] | {"id", "primaryip", "ips", "dataset", "disk", "memory",
                                                         "type", "state", "created", "updated", "domain_names",
                                                         "compute_node"},
}
META_ARGS = {"count", "depends_on", "provider", "lifecycle", "provisioner", "connection"}

PROVISIONERS = {
    "remote-exec": ({"inline", "script", "scripts"}, {"inline", "script", "scripts", "connection"}),
    "local-exec": ({"command"}, {"command", "interpreter", "working_dir", "environment"}),
    "file": ({"destination"}, {"source", "content", "destination", "connection"}),
    "chef": (set(), None),
}
CONNECTION_ARGS = {"type", "host", "user", "password", "private_key", "agent", "port", "timeout", "script_path",
                   "bastion_host", "bastion_port", "bastion_user", "bastion_password", "bastion_private_key",
                   "https", "insecure", "cacert"}
VARIABLE_KEYS = {"type", "default", "description"}
VARIABLE_TYPES = {"string", "list", "map"}


@dataclasses.dataclass
class Variable:
    name: str
    type: str
    has_default: bool
    default: Any
    where: str


@dataclasses.dataclass
class Module:
    """All ``*.tf`` of one directory, merged the way Terraform loads them."""
    path: str
    variables: Dict[str, Variable]
    resources: Dict[Tuple[str, str], Item]
    outputs: Dict[str, Item]
    modules: Dict[str, Item]
    providers: List[Item]
    files: Dict[str, Body]


def load_module(path: str, report: Report) -> Module:
    mod = Module(path, {}, {}, {}, {}, [], {})
    try:
        names = sorted(n for n in os.listdir(path) if n.endswith(".tf"))
    except OSError as exc:
        report.error("module-source", path, "cannot list directory: %s" % exc.strerror)
        return mod
    if not names:
        report.error("module-source", path, "no *.tf files")
    for n in names:
        full = os.path.join(path, n)
        try:
            body = hcl1.parse_file(full)
        except hcl1.HCLError as exc:
            report.error("syntax", full, str(exc))
            continue
        mod.files[full] = body
        for it in body.items:
            where = "%s:%d" % (full, it.line)
            if it.name == "variable":
                _load_variable(it, where, mod, report)
            elif it.name == "resource":
                if len(it.labels) != 2 or not isinstance(it.value, Body):
                    report.error("syntax", where, 'resource needs two labels: resource "TYPE" "NAME" { }')
                    continue
                key = (it.labels[0], it.labels[1])
                if key in mod.resources:
                    report.error("module-dup", where, "resource %s.%s is defined twice" % key)
                mod.resources[key] = it
            elif it.name == "output":
                if len(it.labels) == 1:
                    mod.outputs[it.labels[0]] = it
            elif it.name == "module":
                if len(it.labels) != 1 or not isinstance(it.value, Body):
                    report.error("syntax", where, 'module needs one label: module "NAME" { }')
                    continue
                if it.labels[0] in mod.modules:
                    report.error("module-dup", where, "module %r is defined twice" % it.labels[0])
                mod.modules[it.labels[0]] = it
            elif it.name == "provider":
                mod.providers.append(it)
            elif it.name in ("terraform", "data", "atlas", "locals"):
                pass
            else:
                report.error("syntax", where, "unknown top-level block %r" % it.name)
    return mod


def _load_variable(it: Item, where: str, mod: Module, report: Report) -> None:
    if len(it.labels) != 1 or not isinstance(it.value, Body):
        report.error("variable-shape", where, 'variable needs one label: variable "NAME" { }')
        return
    name = it.labels[0]
    body: Body = it.value
    for a in body.items:
        if a.name not in VARIABLE_KEYS:
            report.error("variable-shape", where, "variable %r: unknown key %r" % (name, a.name))
    default = body.attr("default")
    vtype = None
    t = body.attr("type")
    if t is not None:
        if not isinstance(t.value, Literal) or t.value.kind != "string" or t.value.value not in VARIABLE_TYPES:
            report.error("variable-shape", where, "variable %r: type must be one of %s" % (name, sorted(VARIABLE_TYPES)))
        else:
            vtype = t.value.value
    if vtype is None:
        # type is inferred from the default (config/loader_hcl.go), else string
        if default is not None and isinstance(default.value, ListVal):
            vtype = "list"
        elif default is not None and isinstance(default.value, Body):
            vtype = "map"
        else:
            vtype = "string"
    elif default is not None and _shape(default.value) not in (vtype, "interp"):
        report.error("variable-shape", where, "variable %r: default is a %s but type is %s" % (name, _shape(default.value), vtype))
    if name in mod.variables:
        report.error("module-dup", where, "variable %r is declared twice" % name)
    mod.variables[name] = Variable(name, vtype, default is not None, default.value if default else None, where)


def _shape(value: Any) -> str:
    if isinstance(value, ListVal):
        return "list"
    if isinstance(value, Body):
        return "map"
    if isinstance(value, Literal) and value.kind == "string" and value.parts is not None \
            and any(not isinstance(p, str) for p in value.parts):
        return "interp"  # an interpolation can produce any type
    return "string"


# ---------------------------------------------------------------------------
# reference checking inside one module
# ---------------------------------------------------------------------------
def _check_templates(value: Any, where: str, mod: Module, report: Report, *, in_resource: Optional[Tuple[str, str]],
                     in_provisioner: bool, has_count: bool, child_modules: Dict[str, Module]) -> None:
    for parts, line in hcl1.templates_in(value):
        at = "%s (line %d)" % (where, line)
        for node in hcl1.walk_hil(parts):
            if isinstance(node, HCall):
                if node.func not in FUNCTIONS:
                    report.error("func-unknown", at, "unknown interpolation function %s()" % node.func)
                elif node.func == "file" and len(node.args) != 1:
                    report.error("func-arity", at, "file() takes exactly one argument, got %d" % len(node.args))
            elif isinstance(node, HVar):
                _check_var(node, at, mod, report, in_resource, in_provisioner, has_count, child_modules)


def _check_var(node: HVar, at: str, mod: Module, report: Report, in_resource, in_provisioner, has_count,
               child_modules) -> None:
    p = node.path
    head = p[0]
    if head == "var":
        if len(p) < 2 or p[1] not in mod.variables:
            report.error("var-undeclared", at, "${%s}: no variable %r is declared" % (".".join(p), p[1] if len(p) > 1 else ""))
    elif head == "self":
        if not in_provisioner:
            report.error("self-outside", at, "${%s}: self.* is only valid in provisioner/connection blocks" % ".".join(p))
    elif head == "count":
        if not has_count:
            report.warn("count-index", at, "${%s} in a resource without count" % ".".join(p))
    elif head == "path":
        if len(p) != 2 or p[1] not in ("module", "root", "cwd"):
            report.error("ref-unknown", at, "${%s}: path.module, path.root or path.cwd expected" % ".".join(p))
    elif head == "module":
        if len(p) < 3 or p[1] not in mod.modules:
            report.error("ref-unknown", at, "${%s}: no module %r" % (".".join(p), p[1] if len(p) > 1 else ""))
        else:
            child = child_modules.get(p[1])
            if child is not None and p[2] not in child.outputs:
                report.error("ref-unknown", at, "${%s}: module %r has no output %r" % (".".join(p), p[1], p[2]))
    elif head in ("terraform", "data", "local"):
        pass
    else:
        # TYPE.NAME.ATTR
        if len(p) < 3:
            report.error("ref-unknown", at, "${%s}: expected TYPE.NAME.ATTRIBUTE" % ".".join(p))
            return
        key = (p[0], p[1])
        if key not in mod.resources:
            report.error("ref-unknown", at, "${%s}: no resource %s.%s in this module" % (".".join(p), p[0], p[1]))
            return
        attr = p[2] if p[2] not in ("*",) and not p[2].isdigit() else (p[3] if len(p) > 3 else "")
        known = RESOURCE_ATTRS.get(p[0])
        if known is not None and attr and attr not in known:
            report.warn("attr-unknown", at, "${%s}: %s has no known attribute %r" % (".".join(p), p[0], attr))
        if in_resource == key:
            if in_provisioner:
                report.warn("self-reference", at, "${%s}: resource refers to itself; self.%s is the supported form" % (".".join(p), attr))
            else:
                report.error("ref-unknown", at, "${%s}: resource refers to itself (cycle)" % ".".join(p))


def _check_duplicates(body: Body, where: str, report: Report, repeatable=("provisioner", "connection", "nic")) -> None:
    seen: Dict[str, int] = {}
    for it in body.items:
        if it.name in repeatable:
            continue
        if it.name in seen:
            report.warn("attr-duplicate", "%s (line %d)" % (where, it.line),
                        "attribute %r repeated (first at line %d); HCL1 merges the two" % (it.name, seen[it.name]))
        else:
            seen[it.name] = it.line


def _check_resource(it: Item, mod: Module, report: Report, child_modules: Dict[str, Module]) -> None:
    rtype, rname = it.labels
    where = "%s.%s" % (rtype, rname)
    body: Body = it.value
    has_count = body.attr("count") is not None
    _check_duplicates(body, where, report)
    known = RESOURCE_ARGS.get(rtype)
    for a in body.items:
        if a.name == "provisioner":
            _check_provisioner(a, where, (rtype, rname), mod, report, has_count, child_modules)
            continue
        if a.name == "connection":
            _check_connection(a, where, (rtype, rname), mod, report, has_count, child_modules)
            continue
        if a.name not in META_ARGS and known is not None and a.name not in known:
            report.warn("attr-unknown", "%s (line %d)" % (where, a.line), "%s has no known argument %r" % (rtype, a.name))
        _check_templates(a.value, where, mod, report, in_resource=(rtype, rname), in_provisioner=False,
                         has_count=has_count, child_modules=child_modules)


def _check_provisioner(a: Item, where: str, res, mod, report, has_count, child_modules) -> None:
    at = "%s provisioner (line %d)" % (where, a.line)
    if len(a.labels) != 1 or not isinstance(a.value, Body):
        report.error("provisioner", at, 'provisioner needs one label: provisioner "TYPE" { }')
        return
    ptype = a.labels[0]
    if ptype not in PROVISIONERS:
        report.error("provisioner", at, "unknown provisioner type %r" % ptype)
        return
    need_one_of, allowed = PROVISIONERS[ptype]
    present = {x.name for x in a.value.items}
    if need_one_of and not (present & need_one_of):
        report.error("provisioner", at, "%s needs one of %s" % (ptype, sorted(need_one_of)))
    if allowed is not None:
        for x in a.value.items:
            if x.name not in allowed:
                report.error("provisioner", at, "%s has no argument %r" % (ptype, x.name))
    for x in a.value.items:
        if x.name == "connection":
            _check_connection(x, where, res, mod, report, has_count, child_modules)
        else:
            _check_templates(x.value, at, mod, report, in_resource=res, in_provisioner=True, has_count=has_count,
                             child_modules=child_modules)


def _check_connection(a: Item, where: str, res, mod, report, has_count, child_modules) -> None:
    at = "%s connection (line %d)" % (where, a.line)
    if not isinstance(a.value, Body):
        report.error("provisioner", at, "connection must be a block")
        return
    for x in a.value.items:
        if x.name not in CONNECTION_ARGS:
            report.error("provisioner", at, "connection has no argument %r" % x.name)
        _check_templates(x.value, at, mod, report, in_resource=res, in_provisioner=True, has_count=has_count,
                         child_modules=child_modules)


def _check_module_body(mod: Module, report: Report, child_modules: Dict[str, Module]) -> None:
    for it in mod.resources.values():
        _check_resource(it, mod, report, child_modules)
    for name, it in mod.outputs.items():
        _check_templates(it.value, "output.%s" % name, mod, report, in_resource=None, in_provisioner=False,
                         has_count=False, child_modules=child_modules)
    for it in mod.providers:
        pname = it.labels[0] if it.labels else "?"
        where = "provider.%s" % pname
        if isinstance(it.value, Body):
            _check_duplicates(it.value, where, report)
            known = PROVIDER_ARGS.get(pname)
            for a in it.value.items:
                if known is not None and a.name not in known:
                    report.warn("attr-unknown", "%s (line %d)" % (where, a.line), "provider %s has no known argument %r" % (pname, a.name))
            _check_templates(it.value, where, mod, report, in_resource=None, in_provisioner=False, has_count=False,
                             child_modules=child_modules)


# ---------------------------------------------------------------------------
# root validation
# ---------------------------------------------------------------------------
def validate(root_dir: str) -> Report:
    """Validate the root configuration in ``root_dir`` (``terraform/`` of a
    workspace) and every local module it instantiates."""
    report = Report()
    root = load_module(root_dir, report)
    children: Dict[str, Module] = {}
    loaded: Dict[str, Module] = {}
    for name, it in root.modules.items():
        where = "module.%s" % name
        body: Body = it.value
        _check_duplicates(body, where, report)
        src = body.attr("source")
        if src is None or not isinstance(src.value, Literal) or src.value.kind != "string" \
                or any(not isinstance(p, str) for p in (src.value.parts or [])):
            report.error("module-source", where, "source must be a literal string")
            continue
        src_text = "".join(src.value.parts or [])
        if "://" in src_text or src_text.startswith(("github.com/", "git::", "hg::")):
            report.error("module-source", where, "remote source %r cannot be resolved offline" % src_text)
            continue
        path = os.path.normpath(os.path.join(root_dir, src_text))
        if not os.path.isdir(path):
            report.error("module-source", where, "source %r is not a directory (%s)" % (src_text, path))
            continue
        if path not in loaded:
            sub = Report()
            loaded[path] = load_module(path, sub)
            _check_module_body(loaded[path], sub, {})
            report.findings.extend(sub.findings)
        child = loaded[path]
        children[name] = child
        supplied = {}
        for a in body.items:
            if a.name == "source":
                continue
            if a.name not in child.variables:
                report.error("arg-unknown", "%s (line %d)" % (where, a.line),
                             "module %r (source %r) has no variable %r" % (name, src_text, a.name))
                continue
            supplied[a.name] = a
            var = child.variables[a.name]
            shape = _shape(a.value)
            if shape != "interp" and shape != var.type:
                report.error("arg-type", "%s (line %d)" % (where, a.line),
                             "variable %r is a %s, got a %s" % (a.name, var.type, shape))
        for vname, var in child.variables.items():
            if not var.has_default and vname not in supplied:
                report.error("arg-missing", where, "required variable %r is not set" % vname)
        _check_templates(body, where, root, report, in_resource=None, in_provisioner=False, has_count=False,
                         child_modules=children)
    _check_module_body(root, report, children)
    return report


# ---------------------------------------------------------------------------
# plan: which resources an apply would create, with their resolved arguments
# ---------------------------------------------------------------------------
def _render(value: Any, scope: Dict[str, Any]) -> Any:
    """Evaluate a value with module variables in scope; anything that needs the
    provider or the filesystem stays symbolic (``<computed>`` / ``${file(…)}``)."""
    if isinstance(value, ListVal):
        return [_render(v, scope) for v in value.items]
    if isinstance(value, Body):
        return {it.name: _render(it.value, scope) for it in value.items}
    if isinstance(value, Literal):
        if value.kind != "string" or value.parts is None:
            return value.value
        if len(value.parts) == 1 and isinstance(value.parts[0], HVar):
            return _render_var(value.parts[0], scope)
        out = []
        for p in value.parts:
            if isinstance(p, str):
                out.append(p)
            elif isinstance(p, HVar):
                r = _render_var(p, scope)
                out.append(r if isinstance(r, str) else str(r))
            elif isinstance(p, HCall) and p.func == "file" and len(p.args) == 1 and isinstance(p.args[0], hcl1.HStr):
                out.append('${file("%s")}' % "".join(x for x in p.args[0].parts if isinstance(x, str)))
            else:
                out.append("<computed>")
        return "".join(out)
    return value


def _render_var(node: HVar, scope: Dict[str, Any]) -> Any:
    if node.path[0] == "var" and len(node.path) >= 2 and node.path[1] in scope:
        return scope[node.path[1]]
    return "<computed>"


def plan(root_dir: str) -> List[Dict[str, Any]]:
    """Resource instances an apply of ``root_dir`` would create, in file order:
    ``[{"address": "module.kubenode1.triton_machine.host", "type": …, "args": {…}}, …]``.
    This is the "terraform plan against a mocked provider" of BASELINE config 5,
    reduced to what can be known without a provider: the graph and the inputs."""
    report = Report()
    root = load_module(root_dir, report)
    out: List[Dict[str, Any]] = []
    for name, it in root.modules.items():
        body: Body = it.value
        src = body.attr("source")
        if src is None or not isinstance(src.value, Literal):
            continue
        path = os.path.normpath(os.path.join(root_dir, "".join(p for p in (src.value.parts or []) if isinstance(p, str))))
        child = load_module(path, report)
        scope: Dict[str, Any] = {}
        for vname, var in child.variables.items():
            if var.has_default:
                scope[vname] = _render(var.default, {})
        for a in body.items:
            if a.name != "source":
                scope[a.name] = _render(a.value, {})
        for (rtype, rname), res in child.resources.items():
            args = {}
            for a in res.value.items:
                if a.name in ("provisioner", "connection"):
                    continue
                rendered = _render(a.value, scope)
                if a.name in args and isinstance(args[a.name], dict) and isinstance(rendered, dict):
                    args[a.name] = {**args[a.name], **rendered}  # HCL1 merges repeated map attributes
                else:
                    args[a.name] = rendered
            out.append({"address": "module.%s.%s.%s" % (name, rtype, rname), "type": rtype, "args": args})
    return out
