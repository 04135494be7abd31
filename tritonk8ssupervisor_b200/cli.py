"""Command-line entry point.

Usage (same as the reference, setup.sh:8-12)::

    tk8s-setup        # interactive: prompts, generates, provisions
    tk8s-setup -c     # clean: destroy + reset configuration

Environment:
  TK8S_EXTERNAL_CLOCK=1  take ``sleep``/``date +%s`` from PATH (as the reference
                         does) instead of in-process calls; used by the offline
                         harness to script time.
  TK8S_CACHE_TRITON=1    fetch each `triton` table once instead of once per
                         resolved index (same files, fewer CloudAPI round trips).
"""
from __future__ import annotations

import sys
from typing import Optional, Sequence

from .runtime import HostRuntime
from .supervisor import Supervisor


def main(argv: Optional[Sequence[str]] = None, root: str = ".") -> int:
    args = list(sys.argv[1:] if argv is None else argv)
    try:
        return Supervisor(HostRuntime(root=root), program="setup.sh").main(args)
    except KeyboardInterrupt:
        return 130          # what a shell script reports after SIGINT
    except BrokenPipeError:
        return 141          # ... and after SIGPIPE on a closed stdout
