"""Run scenarios through the unmodified reference ``setup.sh`` (TEST INFRASTRUCTURE).

Only tests/, ``__graft_entry__.smoke()`` and bench.py's baseline legs may import
this module; the product package never does.
"""
from __future__ import annotations

import hashlib
import os
from typing import Dict, Optional

from tritonk8ssupervisor_b200.conformance.sandbox import Observation, Sandbox
from tritonk8ssupervisor_b200.conformance.scenario import Scenario

HERE = os.path.dirname(os.path.abspath(__file__))
STAGED = os.path.join(HERE, "_ref", "reference")
CANDIDATES = ("/root/reference", STAGED)

#: how the reference is invoked inside the sandbox (cwd = workspace root) unless the
#: scenario says otherwise: the documented way, README.md:105 `$ ./setup.sh`
REFERENCE_ARGV = ["./setup.sh"]


def find_reference() -> Optional[str]:
    """An unmodified reference tree, or None (e.g. on a box it did not travel to)."""
    for cand in CANDIDATES:
        if os.path.isfile(os.path.join(cand, "setup.sh")):
            return cand
    return None


def reference_sha256(root: Optional[str] = None) -> str:
    root = root or find_reference()
    if root is None:
        raise FileNotFoundError("no reference tree (looked in %s)" % (CANDIDATES,))
    with open(os.path.join(root, "setup.sh"), "rb") as fh:
        return hashlib.sha256(fh.read()).hexdigest()


def run_reference(scenario: Scenario, root: Optional[str] = None,
                  input_overrides: Optional[Dict[str, str]] = None) -> Observation:
    root = root or find_reference()
    if root is None:
        raise FileNotFoundError("no reference tree (looked in %s)" % (CANDIDATES,))
    with Sandbox(root, scenario, input_overrides) as sbx:
        return sbx.run(scenario.reference_argv())
