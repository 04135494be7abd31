"""Dry-run the playbook against simulated machines and a Rancher mock
("a local HTTP mock + a tiny task interpreter lets the
Jinja expressions be evaluated for real").

This is a small interpreter for the subset of Ansible the playbook uses: plays
over inventory groups, roles, ``vars_files``, the linear strategy, ``register``,
``when`` (bare or in ``{{ }}``), ``until``/``retries``, ``with_items``,
``run_once``, ``delegate_to``, ``action:``/``local_action:`` in both spellings,
and the modules ``command``, ``shell``, ``apt``, ``pip``, ``apt_key``,
``apt_repository``, ``docker_container``, ``docker``, ``debug``, ``pause``,
``uri``, ``slurp``, ``copy``.  Templating is real Jinja2 with the filters the
roles use.  Nothing touches a real machine: package and container modules act on
:class:`SimHost` state, ``uri`` speaks real HTTP but only to the address map it
is given (the :class:`~tritonk8ssupervisor_b200.ranchermock.RancherMock`), and
``local_action`` file work happens inside the playbook directory, which callers
make a scratch copy.

Ansible itself is third-party and unavailable here; semantics restated from its
documentation for the 2.2-2.4 line (strategy/linear.py, playbook/conditional.py,
the modules' documented return values).  Parity unpinned by the reference.
"""
from __future__ import annotations

import base64
import json
import os
import re
import subprocess
import urllib.error
import urllib.request
from typing import Any, Dict, List, Optional, Tuple

import jinja2

from . import ansiblecheck as ac
from .inventory import Inventory


class TaskFailed(Exception):
    def __init__(self, host: str, task: str, msg: str, result: Optional[Dict[str, Any]] = None):
        super().__init__("%s: task %r failed: %s" % (host, task, msg))
        self.host, self.task, self.msg, self.result = host, task, msg, result or {}


class SimHost:
    """What the modules can observe/change on one machine."""

    def __init__(self, name: str, docker_version: Optional[str] = None):
        self.name = name
        self.docker_version = docker_version          # e.g. "1.12.6" once docker-engine is installed
        self.packages: List[str] = []
        self.pip: List[str] = []
        self.apt_keys: List[str] = []
        self.apt_repos: List[str] = []
        self.dist_upgraded = False
        self.containers: List[Dict[str, Any]] = []    # {"name","image","command",...}
        self.facts = {"ansible_default_ipv4": {"address": name}, "ansible_hostname": name}
        self.logs_calls = 0

    # what `docker ps -a` prints
    def docker_ps(self) -> str:
        rows = ["CONTAINER ID        IMAGE                   COMMAND             NAMES"]
        for i, c in enumerate(self.containers):
            rows.append("%012x        %-22s  %-18r  %s" % (i + 1, c["image"], c.get("command") or "/run", c["name"]))
        return "\n".join(rows)


def _b64decode(v: Any) -> str:
    return base64.b64decode(str(v)).decode("utf-8")


def _b64encode(v: Any) -> str:
    return base64.b64encode(str(v).encode("utf-8")).decode("ascii")


def make_env() -> jinja2.Environment:
    env = jinja2.Environment(undefined=jinja2.StrictUndefined, keep_trailing_newline=True)
    env.filters.update({
        "b64decode": _b64decode, "b64encode": _b64encode, "to_json": json.dumps, "from_json": json.loads,
        "bool": lambda v: str(v).lower() in ("1", "true", "yes", "on") if not isinstance(v, bool) else v,
        "quote": lambda v: "'" + str(v).replace("'", "'\\''") + "'",
        "regex_replace": lambda v, pat, rep="": re.sub(pat, rep, str(v)),
        "basename": lambda v: os.path.basename(str(v)), "dirname": lambda v: os.path.dirname(str(v)),
    })
    return env


_SINGLE = re.compile(r"\{\{((?:(?!\{\{|\}\}).)*)\}\}\Z", re.S)


class Templar:
    def __init__(self) -> None:
        self.env = make_env()

    def render(self, value: Any, variables: Dict[str, Any]) -> Any:
        """Template a task argument; a string that is one ``{{ expr }}`` keeps the
        native type of the expression, as in Ansible."""
        if isinstance(value, str):
            if "{{" not in value and "{%" not in value:
                return value
            m = _SINGLE.match(value)          # exact: "{{ x }}\n" is text, and keeps its newline
            if m:
                out = self.env.compile_expression(m.group(1).strip(), undefined_to_none=False)(**variables)
                if isinstance(out, jinja2.Undefined):
                    raise jinja2.UndefinedError("%r is undefined" % m.group(1).strip())
                return out
            return self.env.from_string(value).render(**variables)
        if isinstance(value, dict):
            return {self.render(k, variables): self.render(v, variables) for k, v in value.items()}
        if isinstance(value, list):
            return [self.render(v, variables) for v in value]
        return value

    def truth(self, expr: Any, variables: Dict[str, Any]) -> bool:
        """``when``-style conditional: Jinja truthiness of the expression."""
        if expr is None:
            return True
        if isinstance(expr, bool):
            return expr
        if isinstance(expr, (list, tuple)):
            return all(self.truth(e, variables) for e in expr)
        text = str(expr).strip()
        m = _SINGLE.match(text)
        if m:
            text = m.group(1).strip()
        elif "{{" in text:
            text = str(self.render(text, variables))
        tpl = self.env.from_string("{% if " + text + " %}True{% else %}False{% endif %}")
        return tpl.render(**variables) == "True"


# ---------------------------------------------------------------------------
# modules
# ---------------------------------------------------------------------------
class Modules:
    """Simulated module implementations.  ``address_map`` maps the authority a
    role uses (``"10.0.0.1:8080"``) to where the mock really listens."""

    def __init__(self, playbook_dir: str, address_map: Dict[str, str], rancher_up_after: int = 2,
                 agent_bootstrap_fails: bool = False):
        self.dir = playbook_dir
        self.address_map = address_map
        self.rancher_up_after = rancher_up_after
        #: the bootstrap container exits without starting the long-lived agent
        #: (bad token, server not ready): `docker ps -a` then shows it alone
        self.agent_bootstrap_fails = agent_bootstrap_fails
        self.paused_s = 0

    def run(self, module: str, args: Dict[str, Any], host: SimHost, local: bool) -> Dict[str, Any]:
        fn = getattr(self, "m_" + module, None)
        if fn is None:
            raise TaskFailed(host.name, module, "module %r is not simulated" % module)
        return fn(args, host, local)

    # -- commands ------------------------------------------------------------
    def m_command(self, args, host, local, shell=False):
        cmd = str(args.get("_raw_params", "")).strip()
        if local:
            p = subprocess.run(cmd if shell else cmd.split(), shell=shell, cwd=self.dir, capture_output=True, text=True)
            return {"changed": True, "rc": p.returncode, "stdout": p.stdout.rstrip("\n"), "stderr": p.stderr.rstrip("\n"),
                    "failed": p.returncode != 0, "cmd": cmd}
        out, rc = "", 0
        if cmd.startswith("docker ps"):
            out = host.docker_ps()
        elif cmd.startswith("docker logs"):
            host.logs_calls += 1
            out = "time=... msg=\"Starting rancher\"" + ("\ntime=... msg=\"Listening on :8080\"" if host.logs_calls >= self.rancher_up_after else "")
        elif "docker --version" in cmd:
            banner = "Docker version %s, build 78d1802" % host.docker_version if host.docker_version else ""
            if "grep -c" in cmd:
                out = "1" if host.docker_version == "1.12.6" else "0"
            elif "grep 1.12.6" in cmd:
                out = banner if host.docker_version == "1.12.6" else ""
            else:
                out = banner
        else:
            rc, out = 127, ""
        return {"changed": True, "rc": rc, "stdout": out, "stdout_lines": out.split("\n") if out else [], "stderr": "",
                "failed": rc != 0, "cmd": cmd}

    def m_shell(self, args, host, local):
        return self.m_command(args, host, local, shell=True)

    # -- packages ------------------------------------------------------------
    def m_apt(self, args, host, local):
        changed = False
        if args.get("upgrade"):
            host.dist_upgraded = changed = True
        name = args.get("name") or args.get("pkg") or args.get("package")
        for n in ([name] if isinstance(name, str) else (name or [])):
            if n not in host.packages:
                host.packages.append(n)
                changed = True
            if n.startswith("docker-engine="):
                host.docker_version = n.split("=", 1)[1].split("-", 1)[0]
        return {"changed": changed}

    def m_pip(self, args, host, local):
        host.pip.append(str(args.get("name")))
        return {"changed": True}

    def m_apt_key(self, args, host, local):
        host.apt_keys.append(str(args.get("id") or args.get("url")))
        return {"changed": True}

    def m_apt_repository(self, args, host, local):
        host.apt_repos.append(str(args["repo"]))
        return {"changed": True}

    # -- containers ----------------------------------------------------------
    def m_docker_container(self, args, host, local):
        """``state: started``: nothing to do for a running container with the same
        settings; one that has exited is started again; one whose settings differ
        (a new registration URL as ``command``) is recreated."""
        name = str(args["name"])
        wanted = {"name": name, "image": args.get("image"), "command": args.get("command"),
                  "privileged": args.get("privileged"), "volumes": args.get("volumes"),
                  "published_ports": args.get("published_ports"), "restart_policy": args.get("restart_policy")}
        one_shot = str(args.get("image", "")).startswith("rancher/agent")     # registers the host, then exits
        for c in host.containers:
            if c["name"] == name:
                same = all(c.get(k) == v for k, v in wanted.items())
                if same and c.get("running", True):
                    return {"changed": False}
                host.containers.remove(c)
                break
        host.containers.append(dict(wanted, running=not one_shot))
        if one_shot and not self.agent_bootstrap_fails and not any(c["name"] == "rancher-agent" for c in host.containers):
            # the bootstrap container starts the long-lived agent
            host.containers.append({"name": "rancher-agent", "image": args.get("image"), "command": "run"})
        return {"changed": True}

    def m_docker(self, args, host, local):
        """The pre-2.1 ``docker`` module (rancherhost/tasks/main.yml:27): unnamed container."""
        args = dict(args)
        args.setdefault("name", "pensive_curie%d" % len(host.containers))     # docker invents a name
        return self.m_docker_container(args, host, local)

    # -- misc ------------------------------------------------------------------
    def m_debug(self, args, host, local):
        return {"changed": False, "msg": args.get("msg", args.get("var"))}

    def m_pause(self, args, host, local):
        self.paused_s += int(args.get("seconds", 0) or 0) + 60 * int(args.get("minutes", 0) or 0)
        return {"changed": False}

    def m_slurp(self, args, host, local):
        src = str(args.get("src") or args.get("path"))
        path = os.path.join(self.dir, src) if local else src
        if not local:
            raise TaskFailed(host.name, "slurp", "remote slurp is not simulated")
        try:
            with open(path, "rb") as fh:
                data = fh.read()
        except OSError as exc:
            return {"failed": True, "msg": "file not found: %s (%s)" % (src, exc.strerror)}
        return {"changed": False, "content": base64.b64encode(data).decode("ascii"), "encoding": "base64", "source": src}

    def m_copy(self, args, host, local):
        if not local or "content" not in args:
            raise TaskFailed(host.name, "copy", "only local copy with content= is simulated")
        path = os.path.join(self.dir, str(args["dest"]))
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "w", encoding="utf-8") as fh:
            fh.write(str(args["content"]))
        return {"changed": True, "dest": str(args["dest"])}

    def m_uri(self, args, host, local):
        url = str(args["url"])
        m = re.match(r"(https?)://([^/]+)(/.*)?$", url)
        if not m:
            return {"failed": True, "msg": "unsupported url %r" % url, "url": url}
        authority = m.group(2)
        if authority not in self.address_map:
            return {"failed": True, "status": -1, "url": url,
                    "msg": "connection refused: %s is not reachable in this dry run" % authority}
        real = "%s://%s%s" % (m.group(1), self.address_map[authority], m.group(3) or "/")
        method = str(args.get("method", "GET")).upper()
        headers = {str(k): str(v) for k, v in (args.get("headers") or {}).items()}
        for k, v in args.items():
            if k.startswith("HEADER_"):
                headers[k[len("HEADER_"):]] = str(v)
        body = args.get("body")
        data = None
        if body is not None:
            if str(args.get("body_format", "raw")) == "json":
                headers.setdefault("Content-Type", "application/json")
                data = (body if isinstance(body, str) else json.dumps(body)).encode("utf-8")
            else:
                data = str(body).encode("utf-8")
        elif method == "POST":
            data = b""
        req = urllib.request.Request(real, data=data, method=method, headers=headers)
        try:
            with urllib.request.urlopen(req, timeout=10) as resp:
                status, content = resp.status, resp.read().decode("utf-8")
        except urllib.error.HTTPError as exc:
            status, content = exc.code, exc.read().decode("utf-8")
        want = args.get("status_code", 200)
        want = [int(x) for x in (want if isinstance(want, list) else str(want).split(","))]
        result: Dict[str, Any] = {"changed": False, "status": status, "url": url, "content": content}
        try:
            result["json"] = json.loads(content)
        except ValueError:
            pass
        if status not in want:
            result.update(failed=True, msg="Status code was %d and not %s" % (status, want))
        return result


# ---------------------------------------------------------------------------
# the play runner (linear strategy)
# ---------------------------------------------------------------------------
class DryRun:
    """Run a playbook over simulated hosts.

    ``events`` is the flat log: ``(play, host, task name, status)`` with status
    in ``ok | changed | skipped | failed``; ``results[host][register_name]``
    holds registered variables."""

    def __init__(self, playbook: str, inventory: Inventory, address_map: Dict[str, str],
                 hosts: Optional[Dict[str, SimHost]] = None, rancher_up_after: int = 2,
                 unreachable: Optional[Dict[str, str]] = None, agent_bootstrap_fails: bool = False):
        self.playbook = os.path.abspath(playbook)
        self.dir = os.path.dirname(self.playbook)
        self.inventory = inventory
        self.hosts: Dict[str, SimHost] = hosts if hosts is not None else {}
        for h in inventory.host_order:
            self.hosts.setdefault(h, SimHost(h))
        self.modules = Modules(self.dir, address_map, rancher_up_after, agent_bootstrap_fails)
        self.templar = Templar()
        self.events: List[Tuple[str, str, str, str]] = []
        self.hostvars: Dict[str, Dict[str, Any]] = {h: {} for h in self.hosts}
        #: hosts Ansible cannot use at all (no root login, no python): host -> reason
        self.failed_hosts: Dict[str, str] = dict(unreachable or {})
        self.slept_s = 0

    # -- variables -----------------------------------------------------------
    def _vars_for(self, host: str, play_vars: Dict[str, Any], extra: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
        v: Dict[str, Any] = {}
        v.update(self.inventory.effective_hostvars(host))
        v.update(play_vars)
        v.update(self.hosts[host].facts)
        v.update(self.hostvars[host])
        v.update({"inventory_hostname": host, "playbook_dir": self.dir, "inventory_dir": self.dir,
                  "groups": {g: self.inventory.hosts_of(g) for g in self.inventory.groups},
                  "hostvars": self.hostvars})
        if extra:
            v.update(extra)
        return v

    # -- one task on one host ------------------------------------------------
    def _execute(self, task: Dict[str, Any], host: str, play_vars: Dict[str, Any], play_name: str) -> Optional[Dict[str, Any]]:
        report = ac.Report()
        label = str(task.get("name") or next((k for k in task if k in ac.MODULES), "task"))
        act = ac._task_action(task, label, report)
        if act is None:
            raise TaskFailed(host, label, "; ".join(f.message for f in report.findings))
        module, raw_args, is_local = act
        variables = self._vars_for(host, play_vars)
        if not self.templar.truth(task.get("when"), variables):
            self.events.append((play_name, host, label, "skipped"))
            return {"skipped": True, "changed": False}
        loop = task.get("with_items", task.get("loop"))
        sim = self.hosts[host]

        def once(extra: Optional[Dict[str, Any]]) -> Dict[str, Any]:
            scope = self._vars_for(host, play_vars, extra)
            args = self.templar.render(raw_args, scope)
            if isinstance(args.get("_raw_params"), str) and "=" in args["_raw_params"] and module not in ("command", "shell"):
                args.update(ac.parse_kv(args.pop("_raw_params"), False))
            return self.modules.run(module, args, sim, is_local)

        attempts = int(task.get("retries", 3)) if "until" in task else 1
        result: Dict[str, Any] = {}
        for attempt in range(1, attempts + 1):
            if loop is not None:
                items = self.templar.render(loop, variables)
                results = [dict(once({"item": it}), item=it) for it in items]
                result = {"results": results, "changed": any(r.get("changed") for r in results),
                          "failed": any(r.get("failed") for r in results)}
            else:
                result = once(None)
            result["attempts"] = attempt
            if "until" not in task:
                break
            reg = task.get("register")
            scope = self._vars_for(host, play_vars, {reg: result} if isinstance(reg, str) else None)
            if self.templar.truth(task["until"], scope):
                break
            if attempt < attempts:
                self.slept_s += int(task.get("delay", 5))
        else:
            result["failed"] = True
            result.setdefault("msg", "until condition never became true after %d attempts" % attempts)
        reg = task.get("register")
        late_scope = self._vars_for(host, play_vars, {reg: result} if isinstance(reg, str) else None)
        if "changed_when" in task:
            result["changed"] = self.templar.truth(task["changed_when"], late_scope)
        if "failed_when" in task:
            result["failed"] = self.templar.truth(task["failed_when"], late_scope)
        if isinstance(reg, str):
            self.hostvars[host][reg] = result
        if result.get("failed") and not task.get("ignore_errors"):
            self.events.append((play_name, host, label, "failed"))
            raise TaskFailed(host, label, str(result.get("msg", "module failed")), result)
        self.events.append((play_name, host, label, "changed" if result.get("changed") else "ok"))
        return result

    # -- plays -----------------------------------------------------------------
    def run(self) -> "DryRun":
        report = ac.Report()
        plays = ac.load_yaml(self.playbook, report)
        if not isinstance(plays, list) or report.errors:
            raise ValueError("cannot load playbook: %s" % [str(f) for f in report.errors])
        for play in plays:
            name = str(play.get("name", "play"))
            pattern = play["hosts"] if isinstance(play["hosts"], str) else ":".join(play["hosts"])
            targets: List[str] = []
            for g in ac._hosts_groups(pattern):
                for h in (self.inventory.hosts_of(g) if g in self.inventory.groups else ([g] if g in self.hosts else [])):
                    if h not in targets:
                        targets.append(h)
            play_vars: Dict[str, Any] = {}
            for k, v in (play.get("vars") or {}).items():
                if k not in ac.MAGIC_VARS:      # magic variables win over play vars
                    play_vars[k] = v
            for entry in play.get("vars_files") or []:
                path = self.templar.render(str(entry), {"playbook_dir": self.dir, "inventory_dir": self.dir})
                path = path if os.path.isabs(path) else os.path.join(self.dir, path)
                data = ac.load_yaml(path, report)
                if report.errors:
                    raise ValueError("cannot load vars file: %s" % [str(f) for f in report.errors])
                play_vars.update(data or {})
            tasks: List[Dict[str, Any]] = []
            for entry in play.get("roles") or []:
                rname = ac._role_name(entry)
                tasks += ac.load_yaml(os.path.join(self.dir, "roles", str(rname), "tasks", "main.yml"), report) or []
            tasks += play.get("tasks") or []
            for task in tasks:
                alive = [h for h in targets if h not in self.failed_hosts]
                if not alive:
                    break
                run_on = alive[:1] if task.get("run_once") else alive
                for host in run_on:
                    try:
                        result = self._execute(task, host, play_vars, name)
                    except TaskFailed as exc:
                        self.failed_hosts[host] = str(exc)
                        if task.get("run_once"):
                            for h in alive:
                                self.failed_hosts.setdefault(h, str(exc))
                        continue
                    if task.get("run_once") and isinstance(task.get("register"), str) and result is not None:
                        for h in alive:
                            self.hostvars[h][task["register"]] = result
        return self

    # -- summaries ---------------------------------------------------------------
    def recap(self) -> Dict[str, Dict[str, int]]:
        out: Dict[str, Dict[str, int]] = {h: {"ok": 0, "changed": 0, "skipped": 0, "failed": 0} for h in self.hosts}
        for _, host, _, status in self.events:
            out[host][status] += 1
            if status == "changed":
                out[host]["ok"] += 1
        return out
