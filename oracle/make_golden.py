"""Regenerate tests/golden/*.json from the reference (run in the build container).

    python oracle/make_golden.py            # all catalogue scenarios
    python oracle/make_golden.py NAME ...   # some

Each fixture is self-contained: scenario, the non-generated input files the path
reads (ansible.cfg, the stored environment id, the stale vars.yml), the path
listing of the workspace (names only -- pathname expansion of free-text answers
depends on it), the SHA-256 of the reference's setup.sh, and the observation.
"""
from __future__ import annotations

import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle.reference_runner import find_reference, reference_sha256  # noqa: E402
from tritonk8ssupervisor_b200.conformance.catalog import ALL  # noqa: E402
from tritonk8ssupervisor_b200.conformance.sandbox import Sandbox  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def make(scenario, ref_root):
    with Sandbox(ref_root, scenario) as sbx:
        inputs = sbx.read_inputs()
        listing = sorted(sbx.snapshot().keys())
        # files the scenario itself planted are not part of the static tree
        listing = [p for p in listing if p not in scenario.pre_files]
        obs = sbx.run(scenario.reference_argv())
    return {
        "generated_by": "oracle/make_golden.py",
        "reference": {"repo": "cheapRoc/tritonK8ssupervisor", "setup_sh_sha256": reference_sha256(ref_root),
                      "argv": scenario.reference_argv()},
        "scenario": scenario.to_json(),
        "workspace_listing": listing,
        "inputs": inputs,
        "expected": obs.to_json(),
    }


def main(argv):
    ref_root = find_reference()
    if ref_root is None:
        print("no reference tree available; cannot regenerate goldens", file=sys.stderr)
        return 1
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    wanted = set(argv)
    for sc in ALL:
        if wanted and sc.name not in wanted:
            continue
        doc = make(sc, ref_root)
        path = os.path.join(GOLDEN_DIR, sc.name + ".json")
        with open(path, "w", encoding="utf-8") as fh:
            json.dump(doc, fh, indent=1, sort_keys=True, ensure_ascii=True)
            fh.write("\n")
        print("wrote %s (exit %d, %d files, %d calls)" % (os.path.relpath(path, ROOT), doc["expected"]["exit_status"],
                                                           len(doc["expected"]["files"]), len(doc["expected"]["calls"])))
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
