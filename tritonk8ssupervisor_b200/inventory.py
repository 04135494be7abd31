"""INI inventory -> the JSON document ``ansible-inventory --list`` prints.

The path generates ``ansible/hosts`` (setup.sh:122-126): a ``[MASTER]`` section
with the master's address and a ``[HOST]`` section with one worker address per
line.  ``ansible-inventory`` is unavailable on this box, so the
INI rules and the shape of the listing are restated here from Ansible's
published behaviour (``lib/ansible/plugins/inventory/ini.py`` and
``lib/ansible/cli/inventory.py`` -- third-party, not vendored by the reference,
version unpinned: Ansible 2.2-2.4 era by SURVEY.md §7.4).

Covered INI features: ``[group]``, ``[group:children]``, ``[group:vars]``,
ungrouped hosts before the first section, ``host key=value …`` host variables
(values go through Python-literal evaluation like Ansible does), ``host:port``,
numeric and alphabetic ranges ``web[01:03]``/``db-[a:c]``, ``#``/``;`` comments.
"""
from __future__ import annotations

import ast
import re
import shlex
from typing import Any, Dict, List, Optional, Tuple


class InventoryError(ValueError):
    def __init__(self, msg: str, lineno: int = 0):
        super().__init__("inventory line %d: %s" % (lineno, msg) if lineno else msg)
        self.lineno = lineno


_SECTION = re.compile(r"^\[([^:\]\s]+)(?::(\w+))?\]\s*(?:#.*)?$")
_RANGE = re.compile(r"\[([0-9]+|[A-Za-z]):([0-9]+|[A-Za-z])(?::([0-9]+))?\]")


class Group:
    def __init__(self, name: str):
        self.name = name
        self.hosts: List[str] = []
        self.children: List[str] = []
        self.vars: Dict[str, Any] = {}


class Inventory:
    def __init__(self) -> None:
        self.groups: Dict[str, Group] = {"all": Group("all"), "ungrouped": Group("ungrouped")}
        self.hostvars: Dict[str, Dict[str, Any]] = {}
        self.host_order: List[str] = []

    def group(self, name: str) -> Group:
        if name not in self.groups:
            self.groups[name] = Group(name)
        return self.groups[name]

    def add_host(self, group: str, host: str, variables: Dict[str, Any]) -> None:
        g = self.group(group)
        if host not in g.hosts:
            g.hosts.append(host)
        if host not in self.hostvars:
            self.hostvars[host] = {}
            self.host_order.append(host)
        self.hostvars[host].update(variables)

    # -- queries -----------------------------------------------------------
    def hosts_of(self, group: str) -> List[str]:
        """Hosts of a group including its descendants, in inventory order."""
        if group == "all":
            return list(self.host_order)
        seen: List[str] = []

        def visit(name: str, trail: Tuple[str, ...]) -> None:
            if name in trail or name not in self.groups:
                return
            g = self.groups[name]
            for h in g.hosts:
                if h not in seen:
                    seen.append(h)
            for c in g.children:
                visit(c, trail + (name,))
        visit(group, ())
        return seen

    def effective_hostvars(self, host: str) -> Dict[str, Any]:
        """Host variables with group variables merged in (parents first)."""
        merged: Dict[str, Any] = dict(self.groups["all"].vars)
        for g in self.groups.values():
            if g.name != "all" and host in self.hosts_of(g.name):
                merged.update(g.vars)
        merged.update(self.hostvars.get(host, {}))
        return merged

    # -- ansible-inventory --list -------------------------------------------
    def to_list_json(self) -> Dict[str, Any]:
        """The document ``ansible-inventory -i hosts --list`` prints: ``_meta.hostvars``
        for every host, ``all.children`` with every top-level group (sorted, and
        ``ungrouped`` always listed), one entry per non-empty group with sorted
        ``hosts`` / ``children``."""
        child_names = {c for g in self.groups.values() for c in g.children}
        top = sorted(n for n in self.groups if n != "all" and n not in child_names)
        doc: Dict[str, Any] = {"_meta": {"hostvars": {h: self.effective_hostvars(h) for h in sorted(self.hostvars)}}}
        doc["all"] = {"children": top}
        for name in sorted(self.groups):
            if name == "all":
                continue
            g = self.groups[name]
            entry: Dict[str, Any] = {}
            if g.hosts:
                entry["hosts"] = sorted(g.hosts)
            if g.children:
                entry["children"] = sorted(g.children)
            if entry:
                doc[name] = entry
        return doc


def _expand_ranges(pattern: str, lineno: int) -> List[str]:
    m = _RANGE.search(pattern)
    if not m:
        return [pattern]
    lo, hi, step = m.group(1), m.group(2), int(m.group(3) or 1)
    head, tail = pattern[:m.start()], pattern[m.end():]
    out: List[str] = []
    if lo.isdigit() != hi.isdigit():
        raise InventoryError("range [%s:%s] mixes numbers and letters" % (lo, hi), lineno)
    if lo.isdigit():
        width = len(lo) if lo.startswith("0") and len(lo) > 1 else 0
        if int(lo) > int(hi):
            raise InventoryError("range [%s:%s]: start is greater than end" % (lo, hi), lineno)
        seq = [str(i).zfill(width) for i in range(int(lo), int(hi) + 1, step)]
    else:
        if ord(lo) > ord(hi):
            raise InventoryError("range [%s:%s]: start is greater than end" % (lo, hi), lineno)
        seq = [chr(c) for c in range(ord(lo), ord(hi) + 1, step)]
    for s in seq:
        out.extend(_expand_ranges(head + s + tail, lineno))
    return out


def _value(text: str) -> Any:
    """Host-line values are Python literals when they parse as one, else strings."""
    try:
        v = ast.literal_eval(text)
    except (ValueError, SyntaxError):
        return text
    if isinstance(v, (set, complex)):
        return text
    return v


def parse_ini(text: str) -> Inventory:
    inv = Inventory()
    section, kind = "ungrouped", "hosts"
    pending_children: List[Tuple[str, str, int]] = []
    for lineno, raw in enumerate(text.splitlines(), 1):
        line = raw.strip()
        if not line or line[0] in "#;":
            continue
        if line.startswith("["):
            m = _SECTION.match(line)
            if m and not _RANGE.fullmatch(line):
                section, kind = m.group(1), m.group(2) or "hosts"
                if kind not in ("hosts", "children", "vars"):
                    raise InventoryError("section [%s:%s] has an unknown type" % (section, kind), lineno)
                inv.group(section)
                continue
            if not m and "]" not in line:
                raise InventoryError("section header is not closed: %r" % line, lineno)
        if kind == "vars":
            if "=" not in line:
                raise InventoryError("expected key=value in [%s:vars], got %r" % (section, line), lineno)
            k, _, v = line.partition("=")
            inv.group(section).vars[k.strip()] = _value(v.strip())
            continue
        if kind == "children":
            name = line.split("#", 1)[0].strip()
            pending_children.append((section, name, lineno))
            continue
        try:
            tokens = shlex.split(line, comments=True)
        except ValueError as exc:
            raise InventoryError("cannot split host line: %s" % exc, lineno)
        if not tokens:
            continue
        pattern, variables = tokens[0], {}
        for tok in tokens[1:]:
            if "=" not in tok:
                raise InventoryError("expected key=value after the host name, got %r" % tok, lineno)
            k, _, v = tok.partition("=")
            variables[k] = _value(v)
        port: Optional[int] = None
        mport = re.fullmatch(r"(.+?):([0-9]+)", pattern)
        if mport and pattern.count(":") == 1:
            pattern, port = mport.group(1), int(mport.group(2))
        for host in _expand_ranges(pattern, lineno):
            hv = dict(variables)
            if port is not None:
                hv["ansible_port"] = port
            inv.add_host(section, host, hv)
    for parent, child, lineno in pending_children:
        inv.group(child)
        if child not in inv.groups[parent].children:
            inv.groups[parent].children.append(child)
    # hosts that are in a named group are not "ungrouped"
    grouped = {h for n, g in inv.groups.items() if n not in ("all", "ungrouped") for h in g.hosts}
    inv.groups["ungrouped"].hosts = [h for h in inv.groups["ungrouped"].hosts if h not in grouped]
    return inv


def parse_file(path: str) -> Inventory:
    with open(path, encoding="utf-8") as fh:
        return parse_ini(fh.read())
