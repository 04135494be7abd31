"""Replay a golden fixture against an implementation of the path.

A golden fixture (tests/golden/*.json, written by oracle/make_golden.py) carries
everything needed to reproduce the recorded run without the reference tree: the
scenario, the contents of the few static files the path reads, and the *names*
of the files in the workspace (free-text answers are pathname-expanded against
them).  Replay builds that skeleton, runs the implementation in the sandbox and
returns both observations.
"""
from __future__ import annotations

import json
import os
import shutil
import sys
import tempfile
from typing import Dict, List, Optional, Sequence, Tuple

from .sandbox import Observation, Sandbox
from .scenario import Scenario

PACKAGE_PARENT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

#: how the replacement is invoked inside the sandbox (cwd = workspace root)
REPLACEMENT_ARGV = [sys.executable, "-m", "tritonk8ssupervisor_b200"]
#: TK8S_PROGRAM is what a launcher passes as "$0" (INTEGRATION.md); <PROGRAM> is filled in per scenario by the sandbox
REPLACEMENT_ENV = {"PYTHONPATH": PACKAGE_PARENT, "PYTHONDONTWRITEBYTECODE": "1", "TK8S_PROGRAM": "<PROGRAM>"}


def load_golden(path: str) -> dict:
    with open(path, encoding="utf-8") as fh:
        return json.load(fh)


def golden_dir() -> str:
    return os.path.join(PACKAGE_PARENT, "tests", "golden")


def list_goldens(directory: Optional[str] = None) -> List[str]:
    d = directory or golden_dir()
    return sorted(os.path.join(d, f) for f in os.listdir(d) if f.endswith(".json"))


def build_skeleton(doc: dict, dest: str) -> str:
    """Workspace with the recorded path listing (empty files) and input contents."""
    for rel in doc["workspace_listing"]:
        p = os.path.join(dest, rel)
        os.makedirs(os.path.dirname(p), exist_ok=True)
        with open(p, "w", encoding="utf-8") as fh:
            fh.write(doc["inputs"].get(rel, ""))
    os.makedirs(dest, exist_ok=True)
    return dest


def not_reproducible_here(doc: dict) -> str:
    """Why this fixture cannot be replayed on this machine ('' if it can).  Two
    scenarios look outside the sandbox by construction: with HOME unset ``~`` is
    the password-database home of whoever runs the test, with HOME empty it is
    ``/``; their fixtures were recorded where neither had an ``.ssh`` directory."""
    import pwd
    sc = doc["scenario"]
    if "HOME" in sc.get("unset_env", []):
        try:
            home = pwd.getpwuid(os.getuid()).pw_dir
        except KeyError:
            home = "/"
        if "ls: cannot access '%s/.ssh'" % home not in doc["expected"]["stderr"] or os.path.exists(os.path.join(home, ".ssh")):
            return "recorded for a user whose home has no .ssh; this one's is %s" % home
    if sc.get("env", {}).get("HOME") == "" and os.path.exists("/.ssh"):
        return "/.ssh exists on this machine"
    # a fixture recorded with a locale variable naming a locale that was not installed, or one that was
    import locale
    for key, name in sc.get("env", {}).items():
        if (key.startswith("LC_") or key == "LANG") and name not in ("", "C", "POSIX"):
            saved = locale.setlocale(locale.LC_CTYPE)
            try:
                locale.setlocale(locale.LC_CTYPE, name)
                installed = True
            except locale.Error:
                installed = False
            finally:
                locale.setlocale(locale.LC_CTYPE, saved)
            if installed != (not name.startswith("xx")):
                return "recorded where locale %s was %s" % (name, "missing" if installed else "installed")
    return ""


def replay(doc: dict, argv: Sequence[str] = None, extra_env: Optional[Dict[str, str]] = None
           ) -> Tuple[Observation, Observation]:
    """Returns (expected, got)."""
    scenario = Scenario.from_json(doc["scenario"])
    expected = Observation.from_json(doc["expected"])
    tmp = tempfile.mkdtemp(prefix="tk8s-skel-")
    try:
        skel = build_skeleton(doc, os.path.join(tmp, "skel"))
        with Sandbox(skel, scenario) as sbx:
            got = sbx.run(list(argv or REPLACEMENT_ARGV), extra_env if extra_env is not None else REPLACEMENT_ENV)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return expected, got


def replay_in_process(doc: dict) -> Tuple[Observation, Observation]:
    """Like :func:`replay`, but the replacement runs in this process against the
    in-memory tool models (no fork/exec).  Returns (expected, got)."""
    from .modeltools import ModelSandbox
    scenario = Scenario.from_json(doc["scenario"])
    expected = Observation.from_json(doc["expected"])
    tmp = tempfile.mkdtemp(prefix="tk8s-skel-")
    try:
        skel = build_skeleton(doc, os.path.join(tmp, "skel"))
        with ModelSandbox(skel, scenario) as sbx:
            got = sbx.run()
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return expected, got
