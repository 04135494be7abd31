"""In-process models of the stub tools, and a :class:`Runtime` built on them.

``stubs/stubtool.sh`` is what both the reference and the replacement talk to in
the sandbox (separate processes, the real boundary).  For the resident-input leg
of ``bench.py`` and for fast property tests the replacement can instead run
in-process against :class:`ModelRuntime`: same scenario data, same tool
behaviour, same call log -- but no fork/exec, stdin and stdout are byte buffers.
tests/test_modeltools.py holds it to the golden fixtures, so the two ways of
running the path cannot drift apart.
"""
from __future__ import annotations

import hashlib
import os
import re
import shutil
import tempfile
import time
from typing import Dict, List, Optional, Sequence, Tuple

from .. import bashsem
from ..runtime import CAPTURE, INHERIT, TO_STDOUT, Runtime, child_env
from ..supervisor import Supervisor
from .sandbox import Observation, _snapshot, _write, lay_down_home, lay_down_symlinks
from .scenario import Scenario

_ENV_HIDDEN = re.compile(r"^(TK8S_STUB_|PYTHON)")


class ModelRuntime(Runtime):
    """Workspace on disk, everything else in memory."""

    def __init__(self, scenario: Scenario, root: str, home: str):
        self.sc = scenario
        self.root = root
        sbx = os.path.dirname(home)
        self.env: Dict[str, str] = {"PATH": sbx + "/bin:/usr/bin:/bin", "HOME": home, "LC_ALL": "C",
                                    "TK8S_STUB_STATE": sbx + "/state", "TK8S_STUB_ROOT": root}
        self._sbx = sbx
        for key, value in scenario.env.items():
            self.env[key] = value.replace("<SBX>", sbx)
        for key in scenario.unset_env:
            self.env.pop(key, None)
        self._stdin = scenario.stdin.encode("utf-8", "surrogateescape")
        self._pos = 0
        self.stdout: List[str] = []
        self.stderr: List[str] = []
        self.calls: List[str] = []
        self._seq: Dict[str, int] = {}
        self._n_calls = 0
        self._n_reads = 0
        self.max_calls = 2000

    # -- terminal ----------------------------------------------------------
    def out(self, text: str) -> None:
        self.stdout.append(text)

    def err(self, text: str) -> None:
        self.stderr.append(text)

    def _next_byte(self) -> Optional[int]:
        if self._pos >= len(self._stdin):
            return None
        self._pos += 1
        return self._stdin[self._pos - 1]

    def read(self, prompt: str) -> Tuple[str, bool]:
        self._n_reads += 1
        if self._n_reads > self.max_calls:
            raise RuntimeError("more than %d reads in one run: runaway prompt loop" % self.max_calls)
        return bashsem.read_builtin(self._next_byte)

    def command_exists(self, name: str) -> bool:
        from .sandbox import STUB_TOOLS
        if name in STUB_TOOLS:
            return name not in self.sc.missing_tools
        return super().command_exists(name)

    # -- clock -------------------------------------------------------------
    def sleep(self, seconds: int, where: str = "", job=None) -> int:
        return self.run(["sleep", str(seconds)], where=where)[0]

    def epoch(self, where: str = "") -> Tuple[int, str]:
        return self.run(["date", "+%s"], stdout=CAPTURE, where=where, lone_subst=True)

    # -- tools -------------------------------------------------------------
    def run(self, argv: Sequence[str], cwd: str = ".", stdout: str = INHERIT,
            stderr: str = INHERIT, where: str = "", lone_subst: bool = False, job=None) -> Tuple[int, str]:
        tool, args = argv[0], list(argv[1:])
        self._n_calls += 1
        if self._n_calls > self.max_calls:
            # an implementation that never leaves the wait loop (or a re-prompt loop at EOF) must not hang the caller
            raise RuntimeError("more than %d tool calls in one run: runaway loop" % self.max_calls)
        if tool in self.sc.missing_tools:
            msg = "%s%s: command not found\n" % (where + ": " if where else "", tool)
            if stderr == INHERIT:
                self.stderr.append(msg)
            elif stderr == TO_STDOUT:
                self.stdout.append(msg)
            return 127, ""
        rel = "." if cwd in (".", "") else "./" + cwd
        self.calls.append("%s|%s|%s" % (tool, rel, " ".join(args)))
        # what stubtool.sh records of its environment: `env | grep -v | sort | sed`, line by line
        env = child_env(self.env, "%s/bin/%s" % (self._sbx, tool), lone_subst)
        # the stub is a /bin/sh script, and dash exports a PWD of its own when it inherits none (or a wrong one)
        here = os.path.normpath(os.path.join(self.root, cwd))
        try:
            inherited_ok = env.get("PWD", "").startswith("/") and os.path.samefile(env["PWD"], here)
        except OSError:
            inherited_ok = False
        if not inherited_ok:
            env["PWD"] = os.path.realpath(here)
        dump = "".join("%s=%s\n" % kv for kv in env.items())
        lines = sorted((l for l in dump.split("\n")[:-1] if not _ENV_HIDDEN.match(l)), key=bashsem.c_sort_key)
        lines = [l.replace(self._sbx, "<SBX>") for l in lines]
        if tool in ("terraform", "ansible-playbook"):
            self.calls.extend("  env:" + l for l in lines)
        else:
            digest = hashlib.md5(bashsem.to_bytes("".join(l + "\n" for l in lines))).hexdigest()[:12]
            self.calls.append("  env~:" + digest)
        for _ in range(self.sc.stdin_eaters.get("%s-%s" % (tool, args[0] if args else ""), 0)):
            # a tool that reads the stdin it inherits takes whole lines (stubtool.sh: `read -r`)
            end = self._stdin.find(b"\n", self._pos)
            if end == -1:
                self._pos = len(self._stdin)     # an unterminated tail is consumed, `read` fails, nothing logged
                break
            self.calls.append("  stdin:" + bashsem.from_bytes(self._stdin[self._pos:end]))
            self._pos = end + 1
        status, out, err = self._tool(tool, args, os.path.join(self.root, cwd))
        if stderr == INHERIT:
            self.stderr.append(err)
        elif stderr == TO_STDOUT:
            self.stdout.append(err)
        if stdout == CAPTURE:
            return status, out
        if stdout == INHERIT:
            self.stdout.append(out)
        return status, ""

    def _next(self, name: str, seq: List[str], default: Optional[str]) -> Optional[str]:
        n = self._seq.get(name, 0) + 1
        self._seq[name] = n
        if n <= len(seq):
            return seq[n - 1]
        return default

    def _tool(self, tool: str, args: List[str], cwd: str) -> Tuple[int, str, str]:
        sc = self.sc
        sub = args[0] if args else ""
        key = "%s-%s" % (tool, sub)
        if key in sc.fail:
            return sc.fail[key], "", ""
        if tool == "triton":
            table = {"env": sc.triton_env, "networks": sc.networks, "packages": sc.packages}.get(sub)
            if table is None:
                return 2, "", "triton stub: unsupported: %s\n" % " ".join(args)
            return 0, table, ""
        if tool == "ssh-keygen":
            return self._ssh_keygen(sub, args)
        if tool == "terraform":
            return self._terraform(sub, args, cwd)
        if tool == "ansible-playbook":
            if sc.env_id is not None:
                _write(os.path.join(cwd, "tmp", "kubernetes_environment.id"), sc.env_id + "\n")
            return 0, "PLAY RECAP (stub) %s\n" % " ".join(args), ""
        if tool == "curl":
            body = self._next("curl", sc.curl_seq, sc.curl_default)
            return 0, body if body is not None else "<title>Kubernetes Dashboard</title>\n", ""
        if tool == "ssh":
            ip = args[2][len("root@"):] if len(args) > 2 and args[2].startswith("root@") else (args[2] if len(args) > 2 else "")
            out = sc.ssh_ps.get(ip, "") if len(args) > 3 and args[3] == "docker" else ""
            return sc.ssh_exit.get(ip, 0), out, ""
        if tool == "date":
            v = self._next("date", [d + "\n" for d in sc.date_seq], None)
            return 0, v if v is not None else "1500000000\n", ""
        if tool == "sleep":
            return 0, "", ""
        return 127, "", "stubtool: no behaviour for %s\n" % tool

    def _ssh_keygen(self, sub: str, args: List[str]) -> Tuple[int, str, str]:
        sc = self.sc
        if sub == "-E":
            if sc.keygen_no_E:
                return 1, "", "ssh-keygen: illegal option -- E\n"
            path = args[3] if len(args) > 3 else ""
            if not os.path.isfile(path):
                return 255, "", "ssh-keygen: %s: No such file or directory\n" % path
            line = sc.keygen.get(os.path.basename(path))
            if line is None:
                return 255, "", "%s is not a public key file.\n" % path
            return 0, line, ""
        if sub == "-l":
            path = args[2] if len(args) > 2 else ""
            if not os.path.isfile(path):
                return 255, "", "ssh-keygen: %s: No such file or directory\n" % path
            base = os.path.basename(path)
            line = sc.keygen_legacy.get(base, sc.keygen_sha.get(base))
            if line is None:
                return 255, "", "%s is not a public key file.\n" % path
            return 0, line, ""
        if sub == "-R":
            host = args[1] if len(args) > 1 else ""
            return 0, "# Host %s found: line 1\n" % host, "%s/.ssh/known_hosts updated.\n" % self.env["HOME"]
        return 2, "", "ssh-keygen stub: unsupported: %s\n" % " ".join(args)

    def _terraform(self, sub: str, args: List[str], cwd: str) -> Tuple[int, str, str]:
        sc = self.sc
        rel = os.path.relpath(cwd, self.root)
        rel = "." if rel == "." else "./" + rel
        if sub == "get":
            return 0, "Get: file://%s/master\n" % rel, ""
        if sub == "apply":
            if sc.terraform_no_ips:
                return 0, "Error applying plan: provisioning limit\n", ""
            try:
                with open(os.path.join(cwd, "rancher.tf"), "rb") as fh:
                    text = bashsem.from_bytes(fh.read())
            except OSError:
                text = ""
            sources = [l for l in text.split("\n") if "source = " in l]
            i = 0
            for l in sources:
                fields = l.split()
                src = fields[2] if len(fields) > 2 else ""
                if src == '"master"':
                    with open(os.path.join(cwd, "masters.ip"), "a") as fh:
                        fh.write((sc.master_ip or "10.0.0.1") + "\n")
                elif src == '"host"':
                    i += 1
                    with open(os.path.join(cwd, "hosts.ip"), "a") as fh:
                        fh.write("10.0.1.%d\n" % i)
            return 0, "Apply complete! Resources: %d added, 0 changed, 0 destroyed.\n" % len(sources), ""
        if sub == "destroy":
            return 0, "Destroy complete!\n", ""
        return 2, "", "terraform stub: unsupported: %s\n" % " ".join(args)


class ModelSandbox:
    """Like :class:`Sandbox` but nothing is spawned: the path runs in this process."""

    def __init__(self, workspace_src: str, scenario: Scenario, input_overrides: Optional[Dict[str, str]] = None,
                 base_dir: Optional[str] = None):
        self.scenario = scenario
        self.tmp = tempfile.mkdtemp(prefix="tk8s-msbx-", dir=base_dir)
        self.workspace = os.path.join(self.tmp, "ws")
        self.home = os.path.join(self.tmp, scenario.home_name)
        shutil.copytree(workspace_src, self.workspace, symlinks=True,
                        ignore=shutil.ignore_patterns(".git", "img", "__pycache__"))
        for rel, content in (input_overrides or {}).items():
            _write(os.path.join(self.workspace, rel), content)
        lay_down_home(self.tmp, scenario)
        for rel in scenario.pre_remove:
            p = os.path.join(self.workspace, rel)
            if os.path.isdir(p) and not os.path.islink(p):
                shutil.rmtree(p)
            elif os.path.lexists(p):
                os.remove(p)
        for rel, content in scenario.pre_files.items():
            _write(os.path.join(self.workspace, rel), content)
        lay_down_symlinks(self.workspace, scenario)

    def run(self) -> Observation:
        before = _snapshot(self.workspace)
        rt = ModelRuntime(self.scenario, self.workspace, self.home)
        t0 = time.perf_counter()
        status = Supervisor(rt, program=self.scenario.program_label).main(list(self.scenario.args))
        wall = time.perf_counter() - t0
        after = _snapshot(self.workspace)
        scrub = lambda s: s.replace(self.tmp, "<SBX>")  # noqa: E731
        files = {rel: scrub(data.decode("utf-8", "surrogateescape")) for rel, data in after.items() if before.get(rel) != data}
        removed = sorted(rel for rel in before if rel not in after)
        return Observation(exit_status=status, stdout=scrub("".join(rt.stdout)), stderr=scrub("".join(rt.stderr)),
                           files=files, removed=removed, calls=[scrub(c) for c in rt.calls], wall_s=wall)

    def close(self) -> None:
        shutil.rmtree(self.tmp, ignore_errors=True)

    def __enter__(self) -> "ModelSandbox":
        return self

    def __exit__(self, *exc) -> None:
        self.close()
