"""Command-line entry point.

Usage (same as the reference, setup.sh:8-12)::

    tk8s-setup        # interactive: prompts, generates, provisions
    tk8s-setup -c     # clean: destroy + reset configuration

Environment (none of these reaches the tools the program starts):
  TK8S_PROGRAM=<name>     the name the launcher was invoked by; bash labels its own
                          diagnostics with it ("./setup.sh: line 546: export: ...").
                          Default ``./setup.sh``, how the reference is run
                          (README.md:105).
  TK8S_CACHE_TRITON=1     fetch each `triton` table once instead of once per
                          resolved index (same files, fewer CloudAPI round trips,
                          shorter child-process trace).
  TK8S_INPROCESS_CLOCK=1  sleep and read the clock in-process instead of running
                          ``sleep``/``date +%s`` from PATH as the reference does
                          (setup.sh:36,41,46,63,64); shorter child-process trace.
"""
from __future__ import annotations

import os
import signal
import sys
from typing import Optional, Sequence

from .runtime import HostRuntime, SubshellQuit
from .supervisor import Supervisor


def main(argv: Optional[Sequence[str]] = None, root: str = ".") -> int:
    """Run the path; returns the exit status.  Signals end the program the way
    they end a bash script: it *dies of* SIGINT (after a foreground child has
    finished, see :meth:`HostRuntime.run`) and of SIGPIPE, so the caller sees
    "killed by signal", not an exit status of 130 or 141."""
    args = list(sys.argv[1:] if argv is None else argv)
    try:
        signal.signal(signal.SIGPIPE, signal.SIG_DFL)      # Python ignores it by default; bash does not
    except (ValueError, OSError):
        pass                                                # not the main thread: leave it
    rt = HostRuntime(root=root)

    def on_sigquit(signum, frame):
        # a non-interactive bash ignores SIGQUIT; the subshell of a $(...) does not.  A handler rather than SIG_IGN,
        # so that the tools started from here get the default action back on exec
        if rt.subshell_read:
            raise SubshellQuit()
    try:
        signal.signal(signal.SIGQUIT, on_sigquit)
    except (ValueError, OSError):
        pass
    try:
        return Supervisor(rt, program=rt.env.get("TK8S_PROGRAM") or "./setup.sh").main(args)
    except SubshellQuit:
        return 128 + signal.SIGQUIT                         # the substitution's status, which errexit makes the script's
    except KeyboardInterrupt:
        signal.signal(signal.SIGINT, signal.SIG_DFL)
        os.kill(os.getpid(), signal.SIGINT)
        return 130          # not reached
