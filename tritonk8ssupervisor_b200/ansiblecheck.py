"""Structural check of an Ansible playbook, its roles and the generated
inventory/vars -- the offline stand-in for ``ansible-playbook --syntax-check``
on this path.

``ansible`` is unavailable on this box, and the reference is
written in the Ansible 2.2-2.3 dialect (``when: "{{ … }}"``, the ``docker:``
module, ``HEADER_*`` arguments to ``uri``), so the rules are the era-correct
ones.  Module argument tables restate the documented arguments of the modules
the path uses; the modules themselves are third-party and unpinned, so an
argument outside a table is a *warning* ("parity unpinned").

What ``--syntax-check`` really does is load the playbook, resolve roles and
vars_files, and parse every task into (module, args) without running it; that is
what is done here, plus two things the real tool only finds at run time and that
the generated files can break: every Jinja expression must parse and every
variable it reads must be defined somewhere (vars file, ``register``, loop
``item``, fact or magic variable).

  E yaml               file is not valid YAML / not the expected shape
  E play-shape         play without ``hosts``, unknown play keyword
  E role-missing       role directory or ``tasks/main.yml`` missing
  E task-shape         task is not a mapping / has no action / has two actions
  E task-keyword       unknown task keyword (probably a misspelt module)
  E module-arg-missing required module argument missing
  E jinja-syntax       a ``{{ … }}`` / ``when`` expression does not parse
  E jinja-filter       unknown filter or test
  E var-undefined      expression reads a variable nothing defines
  E vars-file          vars_files entry missing or not a mapping
  W module-arg-unknown argument outside the module's table
  W when-braces        ``when``/``until`` wrapped in ``{{ }}`` (2.3 deprecation)
  W magic-shadowed     play var named like a magic variable (``playbook_dir``)
  W yaml-duplicate-key mapping key given twice
  W hosts-unmatched    play targets a group the inventory does not have
  W register-other-hosts variable registered in a play on other hosts
"""
from __future__ import annotations

import functools
import os
import re
from typing import Any, Dict, Iterable, List, Optional, Set, Tuple

import jinja2
import jinja2.meta
import jinja2.nodes
import yaml

from .inventory import Inventory
from .report import Report

PLAY_KEYWORDS = {
    "name", "hosts", "remote_user", "vars", "vars_files", "vars_prompt", "roles", "tasks", "pre_tasks", "post_tasks",
    "handlers", "gather_facts", "become", "become_user", "become_method", "serial", "max_fail_percentage",
    "any_errors_fatal", "environment", "tags", "connection", "strategy", "sudo", "sudo_user", "port",
    "force_handlers", "ignore_errors", "no_log", "order", "gather_subset", "check_mode", "accelerate",
}
TASK_KEYWORDS = {
    "name", "action", "local_action", "args", "register", "when", "loop", "loop_control", "until", "retries", "delay",
    "run_once", "delegate_to", "delegate_facts", "ignore_errors", "become", "become_user", "become_method", "tags",
    "changed_when", "failed_when", "notify", "vars", "environment", "no_log", "async", "poll", "remote_user",
    "connection", "any_errors_fatal", "check_mode", "always_run", "sudo", "sudo_user", "first_available_file",
    "port", "become_flags",
}
FREE_FORM = {"command", "shell", "raw", "script", "include", "include_tasks", "import_tasks", "include_vars",
             "meta", "set_fact", "add_host", "group_by", "include_role", "import_role"}

# module -> (required, required_one_of, known arguments or None for "anything")
_CMD = {"chdir", "creates", "removes", "executable", "warn", "stdin", "_raw_params", "_uses_shell"}
MODULES: Dict[str, Tuple[Set[str], Set[str], Optional[Set[str]]]] = {
    "command": (set(), set(), _CMD),
    "shell": (set(), set(), _CMD),
    "raw": (set(), set(), {"executable", "_raw_params"}),
    "script": (set(), set(), {"creates", "removes", "chdir", "_raw_params"}),
    "apt": (set(), set(), {"name", "pkg", "package", "state", "update_cache", "cache_valid_time", "upgrade",
                           "install_recommends", "force", "default_release", "deb", "autoremove", "purge",
                           "dpkg_options", "allow_unauthenticated", "only_upgrade", "autoclean", "force_apt_get"}),
    "pip": (set(), {"name", "requirements"}, {"name", "requirements", "version", "state", "virtualenv", "executable",
                                              "extra_args", "editable", "chdir", "umask", "virtualenv_command",
                                              "virtualenv_python", "virtualenv_site_packages"}),
    "apt_key": (set(), {"id", "url", "data", "file", "keyserver"}, {"id", "data", "file", "keyring", "url",
                                                                    "keyserver", "state", "validate_certs"}),
    "apt_repository": ({"repo"}, set(), {"repo", "state", "mode", "update_cache", "validate_certs", "filename",
                                         "codename"}),
    "docker_container": ({"name"}, set(), None),
    "docker": ({"image"}, set(), None),
    "debug": (set(), set(), {"msg", "var", "verbosity"}),
    "pause": (set(), set(), {"seconds", "minutes", "prompt", "echo"}),
    "uri": ({"url"}, set(), {"url", "method", "body", "body_format", "status_code", "return_content", "headers",
                             "user", "password", "url_username", "url_password", "force_basic_auth", "dest",
                             "follow_redirects", "creates", "removes", "timeout", "validate_certs", "client_cert",
                             "client_key", "others", "force", "use_proxy"}),
    "slurp": ({"src"}, set(), {"src", "path"}),
    "set_fact": (set(), set(), None), "copy": (set(), set(), None), "file": (set(), set(), None),
    "template": (set(), set(), None), "service": (set(), set(), None), "systemd": (set(), set(), None),
    "lineinfile": (set(), set(), None), "wait_for": (set(), set(), None), "fail": (set(), set(), None),
    "assert": (set(), set(), None), "include": (set(), set(), None), "include_tasks": (set(), set(), None),
    "import_tasks": (set(), set(), None), "include_vars": (set(), set(), None), "include_role": (set(), set(), None),
    "import_role": (set(), set(), None), "get_url": (set(), set(), None), "stat": (set(), set(), None),
    "user": (set(), set(), None), "group": (set(), set(), None), "package": (set(), set(), None),
    "setup": (set(), set(), None), "ping": (set(), set(), None), "meta": (set(), set(), None),
    "add_host": (set(), set(), None), "group_by": (set(), set(), None), "docker_image": (set(), set(), None),
    "authorized_key": (set(), set(), None), "hostname": (set(), set(), None), "sysctl": (set(), set(), None),
    "unarchive": (set(), set(), None), "git": (set(), set(), None), "cron": (set(), set(), None),
}
# slurp's ``src`` has the alias ``path``
_ALIASES = {"slurp": {"path": "src"}, "apt": {"pkg": "name", "package": "name"}}

MAGIC_VARS = {
    "inventory_hostname", "inventory_hostname_short", "groups", "group_names", "hostvars", "playbook_dir", "role_path",
    "inventory_dir", "inventory_file", "ansible_play_hosts", "play_hosts", "omit", "ansible_check_mode",
    "ansible_version", "environment", "vars", "ansible_play_batch", "ansible_playbook_python", "role_name",
    "lookup", "query", "q", "range", "dict", "lipsum", "cycler", "joiner", "namespace", "true", "false", "none",
    "True", "False", "None",
}
ANSIBLE_FILTERS = {
    "b64decode", "b64encode", "to_json", "to_nice_json", "from_json", "to_yaml", "to_nice_yaml", "from_yaml", "bool",
    "quote", "regex_replace", "regex_search", "regex_findall", "regex_escape", "basename", "dirname", "expanduser",
    "realpath", "relpath", "splitext", "mandatory", "ternary", "combine", "hash", "password_hash", "checksum", "md5",
    "sha1", "ipaddr", "ipv4", "ipv6", "to_uuid", "to_datetime", "comment", "type_debug", "extract", "flatten",
    "union", "intersect", "difference", "symmetric_difference", "unique", "zip", "zip_longest", "shuffle", "random",
    "log", "pow", "root", "human_readable", "human_to_bytes", "json_query", "dict2items", "items2dict", "win_basename",
    "win_dirname", "win_splitdrive", "fileglob", "success", "failed", "changed", "skipped", "min", "max",
}
ANSIBLE_TESTS = {"match", "search", "regex", "success", "succeeded", "failed", "failure", "changed", "change",
                 "skipped", "skip", "version_compare", "version", "file", "directory", "link", "exists", "abs",
                 "subset", "superset", "issubset", "issuperset", "any", "all", "finished", "started"}

_JENV = jinja2.Environment()
KNOWN_FILTERS = set(_JENV.filters) | ANSIBLE_FILTERS
KNOWN_TESTS = set(_JENV.tests) | ANSIBLE_TESTS
# the analysis passes of jinja2 insist that every filter/test name resolves;
# placeholders are enough, nothing is rendered with them here
for _n in ANSIBLE_FILTERS:
    _JENV.filters.setdefault(_n, lambda value, *a, **k: value)
for _n in ANSIBLE_TESTS:
    _JENV.tests.setdefault(_n, lambda value, *a, **k: True)


# ---------------------------------------------------------------------------
# YAML loading with duplicate-key detection
# ---------------------------------------------------------------------------
_BaseLoader = getattr(yaml, "CSafeLoader", yaml.SafeLoader)   # libyaml parser when the wheel has it


class _Loader(_BaseLoader):
    pass


def _construct_mapping(loader: _Loader, node: yaml.MappingNode, deep: bool = False) -> Dict[Any, Any]:
    seen: Dict[Any, int] = {}
    for key_node, _ in node.value:
        try:
            key = loader.construct_object(key_node, deep=True)
            hash(key)
        except Exception:
            continue
        if key in seen:
            loader.duplicates.append((key, key_node.start_mark.line + 1))  # type: ignore[attr-defined]
        seen[key] = key_node.start_mark.line + 1
    return yaml.constructor.SafeConstructor.construct_mapping(loader, node, deep=deep)


_Loader.add_constructor(yaml.resolver.BaseResolver.DEFAULT_MAPPING_TAG, _construct_mapping)


def load_yaml(path: str, report: Report) -> Any:
    try:
        with open(path, encoding="utf-8") as fh:
            loader = _Loader(fh.read())
    except OSError as exc:
        report.error("yaml", path, "cannot read: %s" % exc.strerror)
        return None
    loader.duplicates = []  # type: ignore[attr-defined]
    try:
        data = loader.get_single_data()
    except yaml.YAMLError as exc:
        report.error("yaml", path, "not valid YAML: %s" % str(exc).replace("\n", " "))
        return None
    finally:
        loader.dispose()
    for key, line in loader.duplicates:  # type: ignore[attr-defined]
        report.warn("yaml-duplicate-key", "%s:%d" % (path, line), "key %r given twice in one mapping" % (key,))
    return data


# ---------------------------------------------------------------------------
# key=value argument strings (Ansible's parse_kv / split_args)
# ---------------------------------------------------------------------------
def split_args(text: str) -> List[str]:
    """Split on blanks outside quotes and outside ``{{ }}`` / ``{% %}`` blocks."""
    tokens: List[str] = []
    cur: List[str] = []
    quote: Optional[str] = None
    depth = 0
    i, n = 0, len(text)
    while i < n:
        c = text[i]
        two = text[i:i + 2]
        if quote:
            cur.append(c)
            if c == "\\" and i + 1 < n:
                cur.append(text[i + 1])
                i += 2
                continue
            if c == quote:
                quote = None
        elif two in ("{{", "{%", "{#"):
            depth += 1
            cur.append(two)
            i += 2
            continue
        elif two in ("}}", "%}", "#}") and depth:
            depth -= 1
            cur.append(two)
            i += 2
            continue
        elif c in "\"'":
            quote = c
            cur.append(c)
        elif c in " \t\n" and depth == 0:
            if cur:
                tokens.append("".join(cur))
                cur = []
        else:
            cur.append(c)
        i += 1
    if quote:
        raise ValueError("unbalanced quote in %r" % text)
    if depth:
        raise ValueError("unbalanced jinja2 block in %r" % text)
    if cur:
        tokens.append("".join(cur))
    return tokens


_KEY = re.compile(r"[A-Za-z_][A-Za-z0-9_\-]*\Z")


def _unquote(v: str) -> str:
    if len(v) >= 2 and v[0] == v[-1] and v[0] in "\"'":
        return v[1:-1]
    return v


def parse_kv(text: str, free_form: bool) -> Dict[str, str]:
    """``k=v k2="v 2" rest`` -> dict; tokens without ``=`` collect in
    ``_raw_params`` for free-form modules and are an error otherwise."""
    args: Dict[str, str] = {}
    raw: List[str] = []
    for tok in split_args(text):
        eq = _find_unquoted_eq(tok)
        if eq > 0 and _KEY.match(tok[:eq]) and not (free_form and raw):
            args[tok[:eq]] = _unquote(tok[eq + 1:])
        else:
            raw.append(tok)
    if raw:
        if not free_form:
            raise ValueError("expected key=value, got %r" % raw[0])
        args["_raw_params"] = " ".join(raw)
    return args


def _find_unquoted_eq(tok: str) -> int:
    quote, depth = None, 0
    i = 0
    while i < len(tok):
        c = tok[i]
        if quote:
            if c == "\\":
                i += 1
            elif c == quote:
                quote = None
        elif tok.startswith("{{", i):
            depth += 1
            i += 1
        elif tok.startswith("}}", i) and depth:
            depth -= 1
            i += 1
        elif c in "\"'":
            quote = c
        elif c == "=" and depth == 0:
            return i
        i += 1
    return -1


# ---------------------------------------------------------------------------
# Jinja
# ---------------------------------------------------------------------------
def _strings_in(value: Any) -> Iterable[str]:
    if isinstance(value, str):
        yield value
    elif isinstance(value, dict):
        for k, v in value.items():
            yield from _strings_in(k)
            yield from _strings_in(v)
    elif isinstance(value, (list, tuple)):
        for v in value:
            yield from _strings_in(v)


@functools.lru_cache(maxsize=4096)
def _analyse_template(source: str) -> Tuple[Tuple[Tuple[str, str], ...], frozenset]:
    """((code, message), ...) and the free variables of one template source."""
    if "{{" not in source and "{%" not in source and "{#" not in source:
        return (), frozenset()
    try:
        ast = _JENV.parse(source)
    except jinja2.TemplateSyntaxError as exc:
        return (("jinja-syntax", "%s in %r" % (exc.message, source)),), frozenset()
    problems = []
    for node in ast.find_all(jinja2.nodes.Filter):
        if node.name not in KNOWN_FILTERS:
            problems.append(("jinja-filter", "unknown filter %r in %r" % (node.name, source)))
    for node in ast.find_all(jinja2.nodes.Test):
        if node.name not in KNOWN_TESTS:
            problems.append(("jinja-filter", "unknown test %r in %r" % (node.name, source)))
    try:
        names = frozenset(jinja2.meta.find_undeclared_variables(ast))
    except jinja2.TemplateAssertionError:
        names = frozenset()  # unknown filter/test, already recorded above
    return tuple(problems), names


def jinja_names(source: str, where: str, report: Report) -> Set[str]:
    """Parse a template; report syntax errors and unknown filters/tests; return
    the free variables it reads."""
    problems, names = _analyse_template(source)
    for code, message in problems:
        report.error(code, where, message)
    return set(names)


def expression_names(expr: Any, keyword: str, where: str, report: Report) -> Set[str]:
    """``when``/``until``/``changed_when``/``failed_when``: a bare expression
    (or a list of them)."""
    names: Set[str] = set()
    if isinstance(expr, bool) or expr is None:
        return names
    if isinstance(expr, (list, tuple)):
        for e in expr:
            names |= expression_names(e, keyword, where, report)
        return names
    text = str(expr)
    if "{{" in text or "{%" in text:
        report.warn("when-braces", where, "%s is already an expression; the {{ }} around %r is deprecated since 2.3" % (keyword, text))
        return jinja_names(text, where, report)
    return jinja_names("{{ (%s) }}" % text, where, report)


# ---------------------------------------------------------------------------
# tasks
# ---------------------------------------------------------------------------
class _Scope:
    def __init__(self) -> None:
        self.vars: Set[str] = set()
        #: registered name -> hosts pattern of the play that registered it
        self.registered: Dict[str, str] = {}

    def defined(self, name: str, play_hosts: str) -> Tuple[bool, Optional[str]]:
        if name in MAGIC_VARS or name.startswith("ansible_") or name in self.vars:
            return True, None
        if name in self.registered:
            src = self.registered[name]
            if src == play_hosts or "all" in (src, play_hosts):
                return True, None
            return True, src
        return False, None


def _task_action(task: Dict[str, Any], where: str, report: Report) -> Optional[Tuple[str, Dict[str, Any], bool]]:
    """(module, args, is_local) of a task, or None after reporting."""
    candidates = [k for k in task if k not in TASK_KEYWORDS and not k.startswith("with_")]
    explicit = [k for k in ("action", "local_action") if k in task]
    unknown = [k for k in candidates if k not in MODULES]
    if unknown:
        report.error("task-keyword", where, "%r is neither a task keyword nor a known module" % unknown[0])
        return None
    if len(candidates) + len(explicit) == 0:
        report.error("task-shape", where, "task has no action")
        return None
    if len(candidates) + len(explicit) > 1:
        report.error("task-shape", where, "task has conflicting actions: %s" % ", ".join(candidates + explicit))
        return None
    is_local = False
    if explicit:
        is_local = explicit[0] == "local_action"
        spec = task[explicit[0]]
        if isinstance(spec, dict):
            spec = dict(spec)
            module = spec.pop("module", None)
            if not isinstance(module, str):
                report.error("task-shape", where, "%s mapping needs a 'module' key" % explicit[0])
                return None
            args: Any = spec
        elif isinstance(spec, str):
            parts = spec.strip().split(None, 1)
            if not parts:
                report.error("task-shape", where, "%s is empty" % explicit[0])
                return None
            module, args = parts[0], (parts[1] if len(parts) > 1 else "")
        else:
            report.error("task-shape", where, "%s must be a string or a mapping" % explicit[0])
            return None
        if module not in MODULES:
            report.error("task-keyword", where, "%s names an unknown module %r" % (explicit[0], module))
            return None
    else:
        module, args = candidates[0], task[candidates[0]]
    free = module in FREE_FORM
    if args is None:
        args = {}
    if isinstance(args, str):
        try:
            args = parse_kv(args, free)
        except ValueError as exc:
            report.error("task-shape", where, "%s: %s" % (module, exc))
            return None
    elif not isinstance(args, dict):
        report.error("task-shape", where, "%s: arguments must be a mapping or a key=value string" % module)
        return None
    extra = task.get("args")
    if isinstance(extra, dict):
        args = {**extra, **args}
    return module, args, is_local


def _check_module_args(module: str, args: Dict[str, Any], where: str, report: Report) -> None:
    required, one_of, known = MODULES[module]
    names = {(_ALIASES.get(module, {}).get(k, k)) for k in args}
    for r in sorted(required):
        if r not in names:
            report.error("module-arg-missing", where, "%s: required argument %r is missing" % (module, r))
    if one_of and not (names & one_of):
        report.error("module-arg-missing", where, "%s: one of %s is required" % (module, sorted(one_of)))
    if module in ("command", "shell", "raw", "script") and not args.get("_raw_params") and "cmd" not in args \
            and "argv" not in args:
        report.error("module-arg-missing", where, "%s: no command given" % module)
    if known is not None:
        for k in args:
            if k in known or _ALIASES.get(module, {}).get(k) in known:
                continue
            if module == "uri" and k.startswith("HEADER_"):
                continue  # 2.x way of passing headers (ranchermaster/tasks/main.yml:41-42)
            report.warn("module-arg-unknown", where, "%s has no documented argument %r" % (module, k))


def check_tasks(tasks: Any, path: str, scope: _Scope, play_hosts: str, report: Report) -> None:
    if tasks is None:
        return
    if not isinstance(tasks, list):
        report.error("yaml", path, "a task file must be a list of tasks")
        return
    for idx, task in enumerate(tasks, 1):
        if not isinstance(task, dict):
            report.error("task-shape", "%s task %d" % (path, idx), "task must be a mapping")
            continue
        label = task.get("name") or next((k for k in task if k in MODULES), "task")
        where = "%s task %d (%s)" % (path, idx, label)
        if "block" in task:
            # block / rescue / always (2.0+): task lists sharing the block's keywords
            extra = [k for k in task if k not in TASK_KEYWORDS and k not in ("block", "rescue", "always")
                     and not k.startswith("with_")]
            for k in extra:
                report.error("task-keyword", where, "%r is not valid on a block" % k)
            reads = expression_names(task.get("when"), "when", where, report)
            for name in sorted(reads):
                ok, _ = scope.defined(name, play_hosts)
                if not ok:
                    report.error("var-undefined", where, "variable %r is not defined by any vars file, register, loop or fact" % name)
            for section in ("block", "rescue", "always"):
                if section in task:
                    check_tasks(task[section], "%s %s" % (where, section), scope, play_hosts, report)
            continue
        act = _task_action(task, where, report)
        loop_keys = [k for k in task if k.startswith("with_")] + (["loop"] if "loop" in task else [])
        local: Set[str] = set()
        if loop_keys:
            lc = task.get("loop_control")
            local.add(lc.get("loop_var", "item") if isinstance(lc, dict) else "item")
        if isinstance(task.get("vars"), dict):
            local |= set(task["vars"])
        own_register = task.get("register")
        reads: Set[str] = set()
        if act is not None:
            module, args, _ = act
            _check_module_args(module, args, where, report)
            for s in _strings_in(args):
                reads |= jinja_names(s, where, report)
        for k in loop_keys:
            for s in _strings_in(task[k]):
                reads |= jinja_names(s, where, report)
        for k in ("delegate_to", "environment", "vars", "name", "become_user", "remote_user"):
            for s in _strings_in(task.get(k)):
                reads |= jinja_names(s, where, report)
        reads |= expression_names(task.get("when"), "when", where, report)
        late: Set[str] = set()
        for k in ("until", "changed_when", "failed_when"):
            late |= expression_names(task.get(k), k, where, report)
        if isinstance(own_register, str):
            late.discard(own_register)  # evaluated after the task has produced its result
        for name in sorted((reads | late) - local):
            ok, other = scope.defined(name, play_hosts)
            if not ok:
                report.error("var-undefined", where, "variable %r is not defined by any vars file, register, loop or fact" % name)
            elif other is not None:
                report.warn("register-other-hosts", where, "variable %r was registered in a play on hosts %r, this play runs on %r"
                            % (name, other, play_hosts))
        if isinstance(own_register, str):
            scope.registered[own_register] = play_hosts
        elif own_register is not None:
            report.error("task-shape", where, "register must be a variable name")
        if act is not None and act[0] == "set_fact":
            for k in act[1]:
                if k not in ("cacheable", "_raw_params"):
                    scope.vars.add(k)


# ---------------------------------------------------------------------------
# playbook
# ---------------------------------------------------------------------------
def _role_name(entry: Any) -> Optional[str]:
    if isinstance(entry, str):
        return entry
    if isinstance(entry, dict):
        r = entry.get("role", entry.get("name"))
        return r if isinstance(r, str) else None
    return None


def _hosts_groups(pattern: Any) -> List[str]:
    if isinstance(pattern, list):
        pattern = ":".join(str(p) for p in pattern)
    return [p.lstrip("!&") for p in re.split(r"[:,]", str(pattern)) if p.lstrip("!&")]


def check_playbook(playbook: str, inventory: Optional[Inventory] = None) -> Report:
    """Check ``playbook`` (path to e.g. ``ansible/clusterUp.yml``) with the roles
    next to it; ``inventory`` (parsed ``ansible/hosts``) enables the hosts check."""
    report = Report()
    base = os.path.dirname(os.path.abspath(playbook))
    plays = load_yaml(playbook, report)
    if plays is None:
        return report
    if not isinstance(plays, list):
        report.error("yaml", playbook, "a playbook must be a list of plays")
        return report
    scope = _Scope()
    for pidx, play in enumerate(plays, 1):
        where = "%s play %d" % (playbook, pidx)
        if not isinstance(play, dict):
            report.error("play-shape", where, "play must be a mapping")
            continue
        where = "%s play %d (%s)" % (playbook, pidx, play.get("name", "unnamed"))
        for k in play:
            if k not in PLAY_KEYWORDS:
                report.error("play-shape", where, "unknown play keyword %r" % k)
        if "hosts" not in play:
            report.error("play-shape", where, "play has no 'hosts'")
            continue
        if isinstance(play["hosts"], str):
            hosts = play["hosts"]
        elif isinstance(play["hosts"], list) and all(isinstance(h, str) for h in play["hosts"]):
            hosts = ":".join(play["hosts"])
        else:
            report.error("play-shape", where, "hosts must be a pattern string or a list of them")
            continue
        if inventory is not None:
            for g in _hosts_groups(play["hosts"]):
                if g not in ("all", "*", "localhost") and g not in inventory.groups and g not in inventory.hostvars:
                    report.warn("hosts-unmatched", where, "hosts pattern %r matches no group or host of the inventory" % g)
                elif g in inventory.groups and g != "all" and not inventory.hosts_of(g):
                    report.warn("hosts-unmatched", where, "group %r has no hosts" % g)
        play_vars = play.get("vars") or {}
        if not isinstance(play_vars, dict):
            report.error("play-shape", where, "vars must be a mapping")
            play_vars = {}
        play_scope_vars: Set[str] = set()
        for k in play_vars:
            if k in MAGIC_VARS:
                report.warn("magic-shadowed", where, "play var %r cannot override the magic variable of that name" % k)
            play_scope_vars.add(k)
        vars_files = play.get("vars_files") or []
        if isinstance(vars_files, str):
            vars_files = [vars_files]
        elif not isinstance(vars_files, list):
            report.error("play-shape", where, "vars_files must be a list of paths")
            vars_files = []
        for entry in vars_files:
            paths = entry if isinstance(entry, list) else [entry]
            loaded = False
            for raw in paths:
                try:
                    rendered = _JENV.from_string(str(raw)).render(playbook_dir=base, inventory_dir=base)
                except jinja2.TemplateError as exc:
                    report.error("jinja-syntax", where, "vars_files entry %r: %s" % (raw, exc))
                    continue
                vpath = rendered if os.path.isabs(rendered) else os.path.join(base, rendered)
                if not os.path.isfile(vpath):
                    continue
                sub = Report()
                data = load_yaml(vpath, sub)
                report.findings.extend(sub.findings)
                if data is None and not sub.errors:
                    data = {}
                if not isinstance(data, dict):
                    report.error("vars-file", where, "%s must hold a mapping of variables" % vpath)
                else:
                    play_scope_vars |= {str(k) for k in data}
                    for s in _strings_in(list(data.values())):
                        jinja_names(s, vpath, report)
                loaded = True
                break
            if not loaded:
                report.error("vars-file", where, "vars_files entry %r does not exist" % (entry,))
        scope.vars = set(play_scope_vars)
        for section in ("pre_tasks",):
            check_tasks(play.get(section), "%s %s" % (playbook, section), scope, hosts, report)
        roles = play.get("roles") or []
        if not isinstance(roles, list):
            report.error("play-shape", where, "roles must be a list")
            roles = []
        for entry in roles:
            rname = _role_name(entry)
            if rname is None:
                report.error("play-shape", where, "role entry %r has no name" % (entry,))
                continue
            rdir = os.path.join(base, "roles", rname)
            tasks_file = os.path.join(rdir, "tasks", "main.yml")
            if not os.path.isdir(rdir) or not (os.path.isfile(tasks_file) or os.path.isfile(os.path.join(rdir, "meta", "main.yml"))):
                report.error("role-missing", where, "role %r: %s not found" % (rname, os.path.relpath(tasks_file, base)))
                continue
            role_vars = set(scope.vars)
            for sub_dir in ("defaults", "vars"):
                vfile = os.path.join(rdir, sub_dir, "main.yml")
                if os.path.isfile(vfile):
                    data = load_yaml(vfile, report)
                    if isinstance(data, dict):
                        role_vars |= {str(k) for k in data}
            if isinstance(entry, dict):
                role_vars |= {str(k) for k in entry if k not in ("role", "name", "tags", "when", "become")}
            saved = scope.vars
            scope.vars = role_vars
            if os.path.isfile(tasks_file):
                check_tasks(load_yaml(tasks_file, report), os.path.relpath(tasks_file, base), scope, hosts, report)
            scope.vars = saved | (scope.vars - role_vars)
        for section in ("tasks", "post_tasks", "handlers"):
            check_tasks(play.get(section), "%s %s" % (playbook, section), scope, hosts, report)
    return report


def check_ansible_cfg(path: str) -> Report:
    """``ansible.cfg`` is INI with a ``[defaults]`` section; the path rewrites its
    ``private_key_file`` line (setup.sh:133,511)."""
    import configparser
    report = Report()
    cp = configparser.ConfigParser(interpolation=None, strict=True)
    try:
        with open(path, encoding="utf-8") as fh:
            cp.read_file(fh)
    except OSError as exc:
        report.error("yaml", path, "cannot read: %s" % exc.strerror)
        return report
    except configparser.Error as exc:
        report.error("yaml", path, "not a valid INI file: %s" % str(exc).replace("\n", " "))
        return report
    if not cp.has_section("defaults"):
        report.error("play-shape", path, "no [defaults] section")
    elif not cp.has_option("defaults", "private_key_file"):
        report.warn("module-arg-unknown", path, "[defaults] has no private_key_file line for the path to rewrite")
    return report
