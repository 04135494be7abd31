"""Record interactive (pseudo-terminal) transcripts of the reference into
tests/golden/tty/*.json (run in the build container).

Off a terminal ``read -p`` prints nothing, so the prompt texts the README
documents (README.md:112-121) are only observable this way.
"""
from __future__ import annotations

import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle.reference_runner import find_reference, reference_sha256  # noqa: E402
from tritonk8ssupervisor_b200.conformance.catalog import TTY_CASES  # noqa: E402
from tritonk8ssupervisor_b200.conformance.sandbox import Sandbox, run_on_pty  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "tty")


def main() -> int:
    ref = find_reference()
    if ref is None:
        print("no reference tree available", file=sys.stderr)
        return 1
    os.makedirs(OUT, exist_ok=True)
    for sc, answers in TTY_CASES:
        with Sandbox(ref, sc) as sbx:
            inputs = sbx.read_inputs()
            listing = sorted(p for p in sbx.snapshot() if p not in sc.pre_files)
            transcript = run_on_pty(sbx, sc.reference_argv(), answers)
        doc = {"generated_by": "oracle/make_golden_tty.py", "reference": {"setup_sh_sha256": reference_sha256(ref)},
               "scenario": sc.to_json(), "answers": answers, "inputs": inputs, "workspace_listing": listing,
               "transcript": transcript}
        with open(os.path.join(OUT, sc.name + ".json"), "w", encoding="utf-8") as fh:
            json.dump(doc, fh, indent=1, sort_keys=True)
            fh.write("\n")
        print("wrote tty/%s.json (%d bytes of transcript)" % (sc.name, len(transcript)))
    return 0


if __name__ == "__main__":
    sys.exit(main())
