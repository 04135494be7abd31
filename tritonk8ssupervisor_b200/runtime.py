"""The process boundary of the path: stdin/stdout/stderr, child processes, the
clock and the workspace directory.

The reference script touches the outside world in exactly these ways, and the
replacement does so only through a :class:`Runtime`, so the same path logic runs

* as a command-line program on a real machine (:class:`HostRuntime`, file
  descriptors 0/1/2, tools found on ``PATH``), and
* in-process against in-memory tool models for property tests and for the
  resident-input leg of ``bench.py`` (:class:`tritonk8ssupervisor_b200.conformance.modeltools.ModelRuntime`).

Child processes inherit stdin/stdout/stderr exactly where the reference lets them
(terraform and ansible output interleaves with the script's own lines), and stdin
is consumed one byte at a time so a child that reads stdin sees what it would see
under bash.
"""
from __future__ import annotations

import os
import pwd
import signal
import subprocess
import time
from typing import Dict, Optional, Sequence, Tuple

from . import bashsem

INHERIT, DEVNULL, CAPTURE, TO_STDOUT = "inherit", "devnull", "capture", "to_stdout"


class ScriptExit(Exception):
    """The script ends with this status (bash ``exit N`` or errexit)."""

    def __init__(self, status: int):
        super().__init__(status)
        self.status = status


class SubshellQuit(Exception):
    """SIGQUIT arrived while the script stood in for a ``$(...)`` subshell blocked in ``read``."""


class Runtime:
    """Abstract boundary; see :class:`HostRuntime`."""

    root: str
    env: Dict[str, str]
    #: the prompt being waited at belongs to a command substitution (setup.sh:94-111 is only ever called that way)
    subshell_read = False

    # -- terminal ----------------------------------------------------------
    def out(self, text: str) -> None:
        raise NotImplementedError

    def err(self, text: str) -> None:
        raise NotImplementedError

    def read(self, prompt: str) -> Tuple[str, bool]:
        """bash ``read -p PROMPT var``."""
        raise NotImplementedError

    # -- children ----------------------------------------------------------
    def run(self, argv: Sequence[str], cwd: str = ".", stdout: str = INHERIT,
            stderr: str = INHERIT, where: str = "", lone_subst: bool = False,
            job: Optional[Tuple[str, str]] = None) -> Tuple[int, str]:
        """Run a tool; returns (exit status, captured stdout or '').  ``job`` = (label, command text as the
        script spells it) for a command bash runs in the foreground: when a signal other than SIGINT/SIGPIPE
        kills it, bash reports that (``LABEL: <pid> Killed   terraform apply``).  ``where``
        ("setup.sh: line 157") prefixes the shell's own diagnostics, e.g. when
        the tool is not installed.  ``lone_subst``: the tool is the only command
        of a ``$(...)`` (``$(triton env)``, ``$(curl ...)``, `` `date +%s` ``):
        bash then execs it straight from the substitution's subshell, and takes the
        shell level it had added back out of SHLVL first -- see :func:`child_env`."""
        raise NotImplementedError

    def sleep(self, seconds: int, where: str = "", job: Optional[Tuple[str, str]] = None) -> int:
        """``sleep N``; returns its exit status."""
        raise NotImplementedError

    def epoch(self, where: str = "") -> Tuple[int, str]:
        """Exit status and output (with its newline) of ``date +%s``."""
        raise NotImplementedError

    # -- filesystem (all paths relative to the workspace root) --------------
    def path(self, rel: str) -> str:
        return os.path.join(self.root, rel)

    def exists(self, rel: str) -> bool:
        return os.path.exists(self.path(rel))

    def read_text(self, rel: str) -> str:
        with open(self.path(rel), "rb") as fh:
            return bashsem.from_bytes(fh.read())

    def write_text(self, rel: str, text: str) -> None:
        with open(self.path(rel), "wb") as fh:
            fh.write(bashsem.to_bytes(text))

    def append_text(self, rel: str, text: str) -> None:
        with open(self.path(rel), "ab") as fh:
            fh.write(bashsem.to_bytes(text))

    def utf8_locale(self) -> bool:
        """Is the character set of the locale the *children* (grep, sed, wc ...) run in UTF-8?  They call
        ``setlocale(LC_ALL, "")``: LC_ALL if set, else every category from its LC_* variable or LANG -- and if any
        one of those names a locale that is not installed the call fails as a whole and they run in C."""
        env = self.env
        if env.get("LC_ALL"):
            return self._utf8_codeset(env["LC_ALL"])
        for category in ("LC_COLLATE", "LC_MESSAGES", "LC_MONETARY", "LC_NUMERIC", "LC_TIME", "LC_ADDRESS", "LC_IDENTIFICATION",
                         "LC_MEASUREMENT", "LC_NAME", "LC_PAPER", "LC_TELEPHONE"):
            if self.locale_unavailable(env.get(category) or env.get("LANG") or ""):
                return False
        return self._utf8_codeset(env.get("LC_CTYPE") or env.get("LANG") or "")

    def shell_utf8_locale(self) -> bool:
        """... and the locale *bash* runs in: it sets LANG for all categories first and LC_CTYPE on top, each on
        its own, so a variable that names no installed locale is skipped instead of spoiling the rest."""
        env = self.env
        if env.get("LC_ALL"):
            return self._utf8_codeset(env["LC_ALL"])
        ctype = env.get("LC_CTYPE") or ""
        if ctype and not self.locale_unavailable(ctype):
            return self._utf8_codeset(ctype)
        return self._utf8_codeset(env.get("LANG") or "")

    def _utf8_codeset(self, name: str) -> bool:
        if name in ("", "C", "POSIX"):
            return False
        import locale
        saved = locale.setlocale(locale.LC_CTYPE)
        try:
            locale.setlocale(locale.LC_CTYPE, name)
            return locale.nl_langinfo(locale.CODESET).upper().replace("-", "") == "UTF8"
        except locale.Error:
            return False
        finally:
            locale.setlocale(locale.LC_CTYPE, saved)

    def locale_unavailable(self, name: str) -> bool:
        """Would ``setlocale(LC_ALL, name)`` fail?  (bash warns about that for LC_ALL at start-up and runs in C.)"""
        if name in ("", "C", "POSIX"):
            return False
        import locale
        saved = locale.setlocale(locale.LC_CTYPE)
        try:
            locale.setlocale(locale.LC_CTYPE, name)
            return False
        except locale.Error:
            return True
        finally:
            locale.setlocale(locale.LC_CTYPE, saved)

    def command_exists(self, name: str) -> bool:
        """Would bash find ``name`` on PATH (as an executable, or at least as a file)?"""
        try:
            find_command(name, self.env.get("PATH"), self.root)
        except FileNotFoundError:
            return False
        except PermissionError:
            return True
        return True

    def passwd_home(self) -> str:
        """The user's home directory according to the password database: what
        bash expands ``~`` to when ``HOME`` is unset (setup.sh:216-217)."""
        try:
            return pwd.getpwuid(os.getuid()).pw_dir
        except KeyError:
            return "/"


class HostRuntime(Runtime):
    """The real thing: fds 0/1/2, ``subprocess``, wall clock.

    ``sleep`` and ``date +%s`` are child processes found on ``PATH``, exactly as
    in the reference (setup.sh:36,41,46,63,64): same child-process trace, and a
    scripted ``date``/``sleep`` on ``PATH`` controls time in offline runs.
    ``inprocess_clock`` (``TK8S_INPROCESS_CLOCK=1``) replaces both by in-process
    calls: ~2 fork/execs fewer per second of waiting, at the price of a trace
    that differs from the reference's; like the table cache it is opt-in.
    """

    def __init__(self, root: str = ".", env: Optional[Dict[str, str]] = None,
                 stdin_fd: int = 0, stdout_fd: int = 1, stderr_fd: int = 2,
                 inprocess_clock: Optional[bool] = None):
        self.root = root
        self.env = inherited_environment() if env is None else dict(env)
        self.stdin_fd, self.stdout_fd, self.stderr_fd = stdin_fd, stdout_fd, stderr_fd
        if inprocess_clock is None:
            inprocess_clock = self.env.get("TK8S_INPROCESS_CLOCK", "") == "1"
        self.inprocess_clock = inprocess_clock
        try:
            self._tty = os.isatty(stdin_fd)
        except OSError:
            self._tty = False

    def out(self, text: str) -> None:
        _write_all(self.stdout_fd, bashsem.to_bytes(text))

    def err(self, text: str) -> None:
        try:
            _write_all(self.stderr_fd, bashsem.to_bytes(text))
        except OSError:
            pass                # a diagnostic that cannot be written is lost, as bash's are

    def _next_byte(self) -> Optional[int]:
        try:
            b = os.read(self.stdin_fd, 1)
        except OSError:
            return None
        return b[0] if b else None

    def read(self, prompt: str) -> Tuple[str, bool]:
        if self._tty:
            # bash prints the -p prompt (on stderr) only when stdin is a terminal
            self.err(prompt)
        return bashsem.read_builtin(self._next_byte)

    def run(self, argv: Sequence[str], cwd: str = ".", stdout: str = INHERIT,
            stderr: str = INHERIT, where: str = "", lone_subst: bool = False,
            job: Optional[Tuple[str, str]] = None) -> Tuple[int, str]:
        # A child inherits our descriptors as they are -- closed ones included (`./setup.sh 2>&-`);
        # naming 0/1/2 explicitly would make Popen dup them and fail on a closed one.
        def inherit(fd: int, std: int):
            return None if fd == std else fd
        out_target = {INHERIT: inherit(self.stdout_fd, 1), DEVNULL: subprocess.DEVNULL,
                      CAPTURE: subprocess.PIPE}[stdout]
        err_target = {INHERIT: inherit(self.stderr_fd, 2), DEVNULL: subprocess.DEVNULL,
                      TO_STDOUT: self.stdout_fd}[stderr]
        in_target = inherit(self.stdin_fd, 0)
        # Ctrl-C goes to the whole foreground process group, the child included.  bash
        # does not act on it while the child runs: it waits for the child to finish
        # whatever it does about the signal (terraform saving its state), and only
        # then dies of it itself -- if the child did.  So SIGINT is merely noted from
        # before the child starts until it has been waited for.  (Letting it raise
        # KeyboardInterrupt inside Popen.wait() can lose the child's status:
        # interrupted between waitpid() and the bookkeeping, the retry finds no child
        # and reports 0.)
        interrupted: list = []
        restore = None
        try:
            if signal.getsignal(signal.SIGINT) is signal.default_int_handler:
                signal.signal(signal.SIGINT, lambda signum, frame: interrupted.append(signum))
                restore = signal.default_int_handler
        except ValueError:
            pass                            # not the main thread: signals are not ours to handle
        try:
            try:
                exe = find_command(argv[0], self.env.get("PATH"), self.path(cwd))
                # bash puts the path of the command it starts into the command's environment as `_`
                proc = subprocess.Popen(list(argv), executable=exe, cwd=self.path(cwd),
                                        env=child_env(self.env, exe, lone_subst),
                                        stdin=in_target, stdout=out_target, stderr=err_target)
            except (FileNotFoundError, PermissionError) as exc:
                # what bash says; the message follows the command's own redirections
                missing = isinstance(exc, FileNotFoundError)
                msg = "%s%s: %s\n" % (where + ": " if where else "", argv[0],
                                      "command not found" if missing else "Permission denied")
                if stderr == INHERIT:
                    self.err(msg)
                elif stderr == TO_STDOUT:
                    self.out(msg)
                return (127 if missing else 126), ""
            data, _ = proc.communicate()
        finally:
            if restore is not None:
                signal.signal(signal.SIGINT, restore)
        if interrupted and proc.returncode == -signal.SIGINT:
            raise KeyboardInterrupt         # the child died of it: so does the script
        if lone_subst and proc.returncode == -signal.SIGINT:
            raise KeyboardInterrupt         # the only command of a $(...) died of SIGINT: bash takes that for its own
        status = proc.returncode
        if status < 0:
            status = 128 - status
            killer = -proc.returncode
            if job is not None and killer not in (signal.SIGINT, signal.SIGPIPE):
                # bash's notice about a foreground command that a signal killed
                if killer == signal.SIGTERM:
                    self.err("Terminated\n")
                else:
                    self.err("%s: %5d %-24s%s\n" % (job[0], proc.pid, signal.strsignal(killer) or "Unknown signal %d" % killer, job[1]))
        return status, (bashsem.from_bytes(data) if data is not None else "")

    def sleep(self, seconds: int, where: str = "", job: Optional[Tuple[str, str]] = None) -> int:
        if self.inprocess_clock:
            time.sleep(seconds)
            return 0
        return self.run(["sleep", str(seconds)], where=where, job=job)[0]

    def epoch(self, where: str = "") -> Tuple[int, str]:
        if self.inprocess_clock:
            return 0, "%d\n" % int(time.time())
        return self.run(["date", "+%s"], stdout=CAPTURE, where=where, lone_subst=True)


def inherited_environment() -> Dict[str, str]:
    """The environment this process was *started with*.  CPython may add
    ``LC_CTYPE=C.UTF-8`` to its own environment at start-up (locale coercion,
    when no locale is configured at all); bash does no such thing, so a variable
    that was not there when the process was exec'ed is not passed on, and one that
    was gets the value it had."""
    env = dict(os.environ)
    if "LC_CTYPE" in env:
        try:
            with open("/proc/self/environ", "rb") as fh:
                initial = dict(item.split(b"=", 1) for item in fh.read().split(b"\0") if b"=" in item)
            if b"LC_CTYPE" not in initial:
                del env["LC_CTYPE"]
            else:
                env["LC_CTYPE"] = initial[b"LC_CTYPE"].decode("utf-8", "surrogateescape")   # `C` coerced to `C.UTF-8`
        except OSError:
            pass
    return env


def child_env(env: Dict[str, str], exe: str, lone_subst: bool = False) -> Dict[str, str]:
    """The environment bash hands a command: the exported variables, ``_`` = the
    path the command was found at, and -- for the only command of a command
    substitution -- SHLVL one lower (never below 0)."""
    out = dict(env, _=exe)
    if lone_subst:
        try:
            out["SHLVL"] = str(max(0, int(env.get("SHLVL", "0")) - 1))
        except ValueError:
            out["SHLVL"] = "0"
    return out


#: the search path bash falls back to when PATH is unset
DEFAULT_PATH = "/usr/local/sbin:/usr/local/bin:/usr/sbin:/usr/bin:/sbin:/bin:."


def find_command(name: str, path: Optional[str], cwd: str = ".") -> str:
    """Where bash finds ``name``: as given when it contains a slash, else the first
    executable regular file along PATH (an empty element is the current directory).
    Raises FileNotFoundError, or PermissionError when only non-executable files of
    that name were met (bash: status 126, "Permission denied")."""
    if "/" in name:
        full = name if os.path.isabs(name) else os.path.join(cwd, name)
        if not os.path.exists(full):
            raise FileNotFoundError(name)
        if os.path.isdir(full) or not os.access(full, os.X_OK):
            raise PermissionError(name)
        return name
    denied = False
    for d in (DEFAULT_PATH if path is None else path).split(":"):
        cand = os.path.join(d or ".", name)
        full = cand if os.path.isabs(cand) else os.path.join(cwd, cand)
        if os.path.isfile(full):
            if os.access(full, os.X_OK):
                return cand
            denied = True
    if denied:
        raise PermissionError(name)
    raise FileNotFoundError(name)


def _write_all(fd: int, data: bytes) -> None:
    view = memoryview(data)
    while view:
        n = os.write(fd, view)
        view = view[n:]
