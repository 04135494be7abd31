"""The provisioning path: prompts -> ``./config`` -> ``terraform/rancher.tf`` ->
(terraform) -> ``ansible/hosts``, ``vars.yml``, ``ansible.cfg`` -> (ansible) ->
dashboard wait loop; and ``-c`` (clean).

This is a from-scratch replacement for the reference's ``setup.sh`` behind the
same process/CLI/filesystem boundary: same answers on stdin,
same child processes with the same argv in the same order and environment, same
bytes on stdout and in every generated file, same exit statuses -- including
the ones that look wrong (exit 0 from the previous-run guard, setup.sh:14-18).

Design, as opposed to the reference's: the prompts are data (message, validator,
error lines, reference line) driven by three small loops where the reference
has eight copy-pasted blocks, the Terraform block shapes come from one emitter, and
everything bash does to a string implicitly is an explicit call into
:mod:`bashsem` / :mod:`textops`.  All I/O goes through a :class:`Runtime`, so the
same code runs as a CLI and in-process.

Function-by-function correspondence (reference file:line):

  main                      setup.sh:8-92
  set_var_defaults          setup.sh:240-254   (a1)
  set_config_from_triton    setup.sh:209-239   (a2)
  get_argument              setup.sh:94-110    (a3)
  get_config_from_user      setup.sh:255-451   (a4)
  network_ids / package_id  setup.sh:532-542   (a5, a6)
  verify_config             setup.sh:452-483   (a7)
  set_config_to_file        setup.sh:199-208   (a8)
  export_vars               setup.sh:543-549   (a9)
  terraform blocks          setup.sh:162-198   (a10-a12)
  run_terraform_tasks       setup.sh:138-161   (a13)
  create_ansible_configs    setup.sh:116-137   (a15)
  run_ansible               setup.sh:111-115
  wait_for_dashboard        setup.sh:59-85     (§8f rank 2)
  clean_runner              setup.sh:484-521   (a17, §8f rank 3)
"""
from __future__ import annotations

import contextlib
import os
import re
import shutil
from typing import Callable, Dict, List, Optional, Sequence, Tuple

from . import bashsem as sh
from . import sedprog
from . import textops as tx
from .basharith import ArithAbort, ArithError, ArithFatal, ArithRefused, Evaluator
from .runtime import CAPTURE, DEVNULL, TO_STDOUT, Runtime, ScriptExit
from .tritonenv import (NOT_MODELLED as _NOT_MODELLED, READONLY as _READONLY, SET_AT_STARTUP as _SET_AT_STARTUP, SHELL_BEHAVIOUR,
                        BashDiagnostic, CommandNotFound, TritonEnvError, apply_triton_env)

RULE = "#" * 80
DASHES = "---------------"
HOSTNAME_RE = re.compile(r"[a-zA-Z][0-9a-zA-Z]+")            # setup.sh:276,288 (anchored: fullmatch)
NODE_COUNT_RE = re.compile(r"[1-9]")                         # setup.sh:301
INDEX_LIST_RE = r"^[1-9][0-9]?(,[1-9][0-9]?)*$"              # setup.sh:337,380 (BRE there)
INDEX_RE = r"^[1-9][0-9]*$"                                  # setup.sh:428
_IDENT = re.compile(r"[A-Za-z_][A-Za-z0-9_]*")

#: environment variables that configure this program and nothing else
#: every command word of setup.sh that an imported function could shadow (its own functions are defined later and win)
_COMMAND_WORDS = frozenset((
    "triton ssh-keygen terraform ansible-playbook curl ssh date sleep sed grep sort tr uniq awk tail cat wc ls rm mv "
    "echo read cd export eval let local exit return true false break set [ source .").split())

OWN_ENV = ("TK8S_PROGRAM", "TK8S_CACHE_TRITON", "TK8S_INPROCESS_CLOCK")

TRITON_NETWORKS = ["triton", "networks", "-oname,id"]
TRITON_PACKAGES = ["triton", "packages", "-oname,id"]


#: `export --help` of bash 5.2 (a line of ./config can be exactly that)
_EXPORT_HELP = (
    'export: export [-fn] [name[=value] ...] or export -p\n'
    '    Set export attribute for shell variables.\n'
    '    \n'
    '    Marks each NAME for automatic export to the environment of subsequently\n'
    '    executed commands.  If VALUE is supplied, assign VALUE before exporting.\n'
    '    \n'
    '    Options:\n'
    '      -f\trefer to shell functions\n'
    '      -n\tremove the export property from each NAME\n'
    '      -p\tdisplay a list of all exported variables and functions\n'
    '    \n'
    "    An argument of `--' disables further option processing.\n"
    '    \n'
    '    Exit Status:\n'
    '    Returns success unless an invalid option is given or NAME is invalid.\n'
)


class _ShellVars(dict):
    """Shell variables; assigning to one that is exported updates the
    environment children get, as in bash."""

    def __init__(self, env: Dict[str, str]):
        super().__init__(env)
        self._env = env
        #: names given the export attribute before they had a value (`export NAME`)
        self.export_when_set: set = set()

    def __setitem__(self, key: str, value: str) -> None:
        super().__setitem__(key, value)
        if key in self._env or key in self.export_when_set:
            self.export_when_set.discard(key)
            self._env[key] = value


class Supervisor:
    """One run of the path against a :class:`Runtime`."""

    def __init__(self, rt: Runtime, program: str = "./setup.sh"):
        self.rt = rt
        #: how the script was invoked: bash labels its diagnostics with it
        #: ("./setup.sh: line 546: export: ...", README.md:105)
        self.program = program
        self.cache_triton = rt.env.get("TK8S_CACHE_TRITON", "") == "1"
        # the replacement's own switches are not part of the environment the
        # reference's children see
        for knob in OWN_ENV:
            rt.env.pop(knob, None)
        #: the character set of the locale is UTF-8: grep then takes bytes that are not valid UTF-8 for binary data
        self.utf8 = rt.utf8_locale()
        #: where unquoted words are pathname-expanded (the workspace), with the locale's idea of a character
        #: how `echo -e` writes \\uHHHH: in a locale that is not a UTF-8 one bash goes by the *name* of the locale
        #: (LC_ALL=en_US.UTF-8 on a machine without that locale still gets UTF-8 out of it)
        #: ... and bash itself is in such a locale (its matcher and its quoting then look at characters)
        self.shell_utf8 = rt.shell_utf8_locale()
        locale_name = rt.env.get("LC_ALL") or rt.env.get("LC_CTYPE") or rt.env.get("LANG") or ""
        self.wide = True if self.shell_utf8 else sh.locale_charset(locale_name)
        self.root = sh.Cwd(rt.root, self.shell_utf8, self.wide)
        #: shell variables (the exported ones are mirrored in ``rt.env``)
        self.v: Dict[str, str] = _ShellVars(rt.env)
        self._shell_bookkeeping()
        self._tables: Dict[tuple, str] = {}
        self._lineno = 0
        self._scopes: List[Dict[str, tuple]] = []
        self._evaluator = Evaluator(self._arith_get, self.v.__setitem__, errexit=True)   # setup.sh:3

    # ------------------------------------------------------------------
    # what bash itself exports to every child: SHLVL, PWD, OLDPWD
    # (``_``, the path of the command being started, is the runtime's business)
    # ------------------------------------------------------------------
    def _shell_bookkeeping(self) -> None:
        """The script *is* a shell level: children see SHLVL one higher than the
        caller's, PWD as bash validated it at start-up, and an OLDPWD that does not
        name a directory is dropped."""
        env = self.rt.env
        old = env.get("SHLVL", "")
        level = (int(old) if re.fullmatch(r"[ \t]*[-+]?[0-9]+[ \t]*", old) else 0) + 1
        if level < 0:
            level = 0
        elif level >= 1000:
            self.rt.err("bash: warning: shell level (%d) too high, resetting to 1\n" % level)
            level = 1
        if self.rt.locale_unavailable(env.get("LC_ALL", "")):
            self.rt.err("bash: warning: setlocale: LC_ALL: cannot change locale (%s)\n" % env["LC_ALL"])
        self.export("SHLVL", str(level))
        # inherited values bash overwrites at start-up (the export attribute stays)
        if "OPTERR" in env:
            self.v["OPTERR"] = "1"
        if "OPTIND" in env:
            self.v["OPTIND"] = "1"
        if "PPID" in env:
            self.v["PPID"] = str(os.getppid())
        if "PS4" in env and os.geteuid() == 0:
            self.v["PS4"] = "+ "                      # root does not inherit PS4
        root = os.path.abspath(self.root)
        pwd = env.get("PWD", "")
        try:
            inherited_ok = pwd.startswith("/") and os.path.samefile(pwd, root)
        except OSError:
            inherited_ok = False
        #: bash's own idea of the working directory; `cd` works from it, not from $PWD (which `triton env` may unset)
        self._cwd = pwd if inherited_ok else os.path.realpath(root)
        self.export("PWD", self._cwd)
        if "OLDPWD" not in env or not os.path.isdir(env["OLDPWD"]):
            # no usable OLDPWD: the variable exists with the export attribute and no value until the first cd
            env.pop("OLDPWD", None)
            self.v.pop("OLDPWD", None)
            self.v.export_when_set.add("OLDPWD")

    # ------------------------------------------------------------------
    # `local`: a function's local variable shadows a global of the same name until
    # the function returns, and takes over its export attribute -- so when the
    # operator's shell exports `tmp`, the function's children see the *local* tmp.
    # A `$(...)` subshell is the same thing seen from outside: what it assigns is
    # gone afterwards.
    # ------------------------------------------------------------------
    @contextlib.contextmanager
    def scope(self):
        frame: Dict[str, tuple] = {}
        self._scopes.append(frame)
        try:
            yield
        finally:
            self._scopes.pop()
            env = self.rt.env
            for name, (had_v, old_v, had_env, old_env, was_pending) in frame.items():
                if was_pending:
                    self.v.export_when_set.add(name)
                if had_v:
                    dict.__setitem__(self.v, name, old_v)
                else:
                    self.v.pop(name, None)
                if had_env:
                    env[name] = old_env
                else:
                    env.pop(name, None)

    def _remember(self, name: str) -> None:
        frame = self._scopes[-1]
        if name not in frame:
            env = self.rt.env
            frame[name] = (name in self.v, self.v.get(name), name in env, env.get(name), name in self.v.export_when_set)

    def local(self, name: str, value: Optional[str] = None) -> None:
        """``local name[=value]`` in the innermost scope.  The new variable takes
        over the export attribute of the one it shadows, so children of the
        function see the *local* value -- or nothing while it has none."""
        self._remember(name)
        if value is None:
            # unset for the shell; but bash assembles a child's environment from the
            # *visible* exported variables of every scope, and a value-less local is
            # not visible: children keep seeing the outer exported value until the
            # local is assigned (rt.env is left as it is)
            self.v.pop(name, None)
        else:
            self.v[name] = value

    def subshell_set(self, name: str, value: str) -> None:
        """Assignment inside a ``$(...)``: exported if the variable is, undone when the scope ends."""
        self._remember(name)
        self.v[name] = value

    def export(self, name: str, value: str) -> None:
        self.v[name] = value
        self.rt.env[name] = value

    def cd(self, rel: str) -> None:
        """``cd rel``: no process of ours changes directory (children are started
        with ``cwd=``), but PWD and OLDPWD are part of what they inherit."""
        target = os.path.normpath(os.path.join(self._cwd, rel))
        if self.v.get("CDPATH") and not rel.startswith(("/", "./", "../")) and rel not in (".", ".."):
            target = self._cd_through_cdpath(rel, target)
        # plain assignments: both are exported because bash gave them the attribute at
        # start-up, and stop being so once `triton env` has unset them
        if "PWD" in self.v:
            self.v["OLDPWD"] = self.v["PWD"]
        else:
            # OLDPWD=<value of an unset PWD>: no value for the shell; bash patches its cached
            # environment in place, so children of the next commands find an empty OLDPWD
            exported = "OLDPWD" in self.rt.env or "OLDPWD" in self.v.export_when_set
            self.v.pop("OLDPWD", None)
            if exported:
                self.v.export_when_set.discard("OLDPWD")
                self.rt.env["OLDPWD"] = ""
        self._cwd = target
        self.v["PWD"] = self._cwd

    def _cd_through_cdpath(self, rel: str, plain: str) -> str:
        """CDPATH (inherited, or exported by a line of ./config) is searched before the current directory, and a
        directory found through a non-empty entry is announced on stdout.  An entry that leads back into the
        workspace is followed; one that leads anywhere else would move the whole run and is refused."""
        home = self.v.get("HOME", "")
        for entry in self.v["CDPATH"].split(":"):
            where = entry
            if entry == "~" or entry.startswith("~/"):
                where = home + entry[1:]
            elif entry.startswith("~"):
                continue                              # ~user: the password database is not consulted for this
            found = os.path.normpath(os.path.join(self._cwd, where, rel))
            if not os.path.isdir(found):
                continue
            if entry == "":
                return plain
            if os.path.realpath(found) != os.path.realpath(plain):
                self.rt.err("%s: refusing to go on: CDPATH takes `cd %s` to %s, out of the workspace\n" % (self.program, rel, found))
                raise ScriptExit(2)
            self.write_out(found + "\n")
            return found
        return plain

    # ------------------------------------------------------------------
    # small helpers: the shell idioms the reference repeats
    # ------------------------------------------------------------------
    def echo(self, line: str = "", lineno: int = 0) -> None:
        """``echo "<line>"`` with a double-quoted argument (``lineno``: its line in the reference)."""
        self.write_out(sh.echo_builtin([line]), lineno)

    def write_out(self, text: str, lineno: int = 0) -> None:
        """An ``echo`` to the script's stdout.  One that cannot be written (stdout
        closed, disk full) costs bash's message and, under errexit, the run."""
        try:
            self.rt.out(text)
        except BrokenPipeError:
            raise                               # SIGPIPE has its default action; only an embedding caller gets here
        except OSError as exc:
            self.rt.err("%s: echo: write error: %s\n" % (self.at(lineno) if lineno else self.program, exc.strerror))
            raise ScriptExit(1)

    def echo_var(self, value: str) -> str:
        """``$(echo $VALUE)``: unquoted, so field-split and globbed in the workspace."""
        return sh.cmdsubst(sh.echo_unquoted(value, self.root))

    def unq(self, value: str) -> str:
        """``$(echo $VALUE | sed 's/"//g')`` -- the reference's quote-stripper."""
        return sh.cmdsubst(tx.strip_dquotes(sh.echo_unquoted(value, self.root)))

    def args_of(self, value: str) -> List[str]:
        """Positional parameters produced by an unquoted ``$VALUE`` argument."""
        return sh.expand_unquoted(value, self.root)

    def append_line(self, rel: str, line: str, lineno: int = 0) -> None:
        """``echo "<line>" >> rel``; a failing redirection ends the script (errexit).
        ``lineno`` is the reference line, which bash names in its message."""
        try:
            self.rt.append_text(rel, sh.echo_builtin([line]))
        except OSError as exc:
            self._redirect_failed(rel, exc, lineno)

    def _redirect_failed(self, rel: str, exc: OSError, lineno: int = 0) -> None:
        where = self.at(lineno) if lineno else self.program
        self.rt.err("%s: %s: %s\n" % (where, rel, exc.strerror))
        raise ScriptExit(1)

    def subst(self, output: str, lineno: int) -> str:
        """``$(...)`` around something whose output is data: trailing newlines go, and so
        do NUL bytes -- those with bash's warning, labelled with the reference line."""
        if "\x00" in output:
            self.rt.err("%s: warning: command substitution: ignored null byte in input\n" % self.at(lineno))
        return sh.cmdsubst(output)

    def cat_subst(self, rel: str, lineno: int = 0) -> str:
        """``$(cat rel)``; a missing file costs a message on stderr, not the run."""
        try:
            text = self.rt.read_text(rel)
            return self.subst(text, lineno) if lineno else sh.cmdsubst(text)
        except OSError as exc:
            self.rt.err("cat: %s: %s\n" % (rel, exc.strerror))
            return ""

    def _binary_matches(self, name: str = "(standard input)") -> None:
        """What grep says instead of printing lines of input it takes for a binary file."""
        self.rt.err("grep: %s: binary file matches\n" % name)

    def at(self, line: int) -> str:
        """``setup.sh: line N`` -- how bash labels its own diagnostics."""
        return "%s: line %d" % (self.program, line)

    def must(self, status: int) -> None:
        """errexit: an unguarded child failure is the script's exit status."""
        if status != 0:
            raise ScriptExit(status)

    def home(self) -> str:
        """What ``~`` stands for: ``$HOME``, set-but-empty included; the password
        database entry when HOME is unset."""
        return self.v["HOME"] if "HOME" in self.v else self.rt.passwd_home()

    def tilde(self, path: str) -> str:
        """Tilde expansion of a word starting with ``~/``."""
        return self.home() + path[1:] if path.startswith("~/") or path == "~" else path

    # -- arithmetic on data (basharith) ----------------------------------------
    def _arith_get(self, name: str) -> Optional[str]:
        if name == "LINENO":
            return str(self._lineno)
        if name in self.v:
            return self.v[name]
        return self._bash_variable(name)

    def _bash_variable(self, name: str) -> Optional[str]:
        """The few variables bash provides by itself that have a fixed value at
        the places where the path evaluates data (INTEGRATION.md lists what is
        not modelled: RANDOM, SECONDS, BASHPID and the like)."""
        if name == "HOME":
            return self.rt.passwd_home()     # set by bash at start-up when unset, not exported
        if name in ("OPTIND", "OPTERR"):
            return "1"
        if name == "IFS":
            return " \t\n"
        if name in ("UID", "EUID"):
            return str(os.getuid() if name == "UID" else os.geteuid())
        if name == "PPID":
            return str(os.getppid())
        if name == "BASH_SUBSHELL":
            return "0"
        if name == "FUNCNAME":
            return "main"
        return None

    def arith_eval(self, expr: str, lineno: int, command: str, last_status: int = 0) -> Optional[int]:
        """Value of an arithmetic expression evaluated by ``command`` (``((`` or
        ``[[``) at reference line ``lineno``; None after bash's diagnostic when
        the expression is in error (the command then fails, it does not end the
        script by itself).  ``last_status`` is the status of the command that ran
        just before, which is what the script exits with if the evaluation trips
        over something that is fatal under ``set -e``."""
        self._lineno = lineno
        try:
            return self._evaluator.evaluate(expr)
        except ArithError as exc:
            self.rt.err("%s: %s: %s\n" % (self.at(lineno), command, exc.text))
            return None
        except ArithAbort as exc:
            # an error inside an array subscript: bash drops the whole top-level
            # command, which for this script is `main "$@"`
            self.rt.err("%s: %s\n" % (self.at(lineno), exc.text))
            raise ScriptExit(1)
        except ArithFatal as exc:
            self.rt.err("%s: %s\n" % (self.at(lineno), exc.text))
            raise ScriptExit(last_status or 1)
        except ArithRefused as exc:
            self.rt.err("%s: %s\n" % (self.program, exc))
            raise ScriptExit(2)

    # ------------------------------------------------------------------
    # main  (setup.sh:8-92)
    # ------------------------------------------------------------------
    def main(self, argv: Sequence[str]) -> int:
        try:
            self._main(argv)
        except ScriptExit as e:
            return e.status
        return 0

    def _main(self, argv: Sequence[str]) -> None:
        rt = self.rt
        # a function exported by the caller (`export -f`) is defined in the script's bash before line 1, and one named
        # like a command of the script would be run in its place
        # variables that reconfigure bash itself when it *inherits* them: the reference would run in another mode
        # (posix mode alone changes the outcome of half of all odd inputs), which is not restated
        for name, why in (("POSIXLY_CORRECT", "puts bash into posix mode"), ("POSIX_PEDANTIC", "puts bash into posix mode"),
                          ("SHELLOPTS", "switches shell options on before the script starts"),
                          ("BASHOPTS", "switches shell options on before the script starts")):
            if name in rt.env:
                rt.err("%s: refusing to go on: %s in the environment %s; unset it\n" % (self.program, name, why))
                raise ScriptExit(2)
        startup_file = rt.env.get("BASH_ENV", "")
        if startup_file and (any(c in startup_file for c in "$`~") or os.path.exists(os.path.join(rt.root, startup_file))):
            rt.err("%s: refusing to go on: BASH_ENV names a file bash would run before the script; unset it\n" % self.program)
            raise ScriptExit(2)
        for key, value in rt.env.items():
            if key.startswith("BASH_FUNC_") and key.endswith("%%") and value.startswith("() {") and key[10:-2] in _COMMAND_WORDS:
                rt.err("%s: refusing to go on: the environment exports a shell function named %s, which the script would run "
                       "instead of the command\n" % (self.program, key[10:-2]))
                raise ScriptExit(2)
        if argv and argv[0] == "-c":
            self.clean_runner()
            raise ScriptExit(0)
        if rt.exists("terraform/rancher.tf"):
            self.echo("error: configuration for a previous run has been found", lineno=15)
            self.echo("    clean the configuration (./setup -c)", lineno=16)
            raise ScriptExit(0)  # bare `exit` after a successful echo

        self.set_var_defaults()
        self.set_config_from_triton()
        self.get_config_from_user()
        self.verify_config()
        self.set_config_to_file()
        self.export_vars()

        self.banner("Starting terraform tasks...", 36)
        self.run_terraform_tasks()
        self.banner("Creating ansible configs...", 41)
        self.create_ansible_configs()
        self.banner("Running ansible tasks...", 46)
        self.run_ansible()

        self.echo("", lineno=49)
        self.echo("Congratulations, your Kubernetes cluster setup has been complete.", lineno=50)
        self.echo("", lineno=51)
        self.echo("It will take a few minutes for all the Kubernetes process to start up before you can access Kubernetes Dashboard", lineno=52)
        # the two `cat`s are part of the next line's argument: what they say on stderr comes after the lines above
        master = self.cat_subst("terraform/masters.ip", 53)
        env_id = self.cat_subst("ansible/tmp/kubernetes_environment.id", 53)
        self.echo("----> To check what processes/containers are coming up, go to http://%s:8080/env/%s/infra/containers" % (master, env_id), lineno=53)
        self.echo("    once all these containers are up, you should be able to access Kubernetes by its dashboard or using CLI", lineno=54)
        self.echo("Waiting on Kubernetes dashboard to come up.", lineno=56)
        self.echo("", lineno=57)
        self.wait_for_dashboard()
        self.echo("", lineno=87)
        master = self.cat_subst("terraform/masters.ip", 88)
        env_id = self.cat_subst("ansible/tmp/kubernetes_environment.id", 88)
        self.echo("----> Kubernetes dashboard is at http://%s:8080/r/projects/%s/kubernetes-dashboard:9090/" % (master, env_id), lineno=88)
        master = self.cat_subst("terraform/masters.ip", 89)
        env_id = self.cat_subst("ansible/tmp/kubernetes_environment.id", 89)
        self.echo("----> Kubernetes CLI config is at http://%s:8080/env/%s/kubernetes/kubectl" % (master, env_id), lineno=89)
        self.echo("", lineno=90)
        self.echo("    CONGRATULATIONS, YOU HAVE CONFIGURED YOUR KUBERNETES ENVIRONMENT!", lineno=91)

    def banner(self, text: str, line: int = 0) -> None:
        self.echo(RULE, lineno=line - 3)
        self.echo("### " + text, lineno=line - 2)
        self.echo(RULE, lineno=line - 1)
        self.must(self.rt.sleep(2, self.at(line), job=(self.at(8), "sleep 2")))

    # ------------------------------------------------------------------
    # a1  setVarDefaults  (setup.sh:240-254)
    # ------------------------------------------------------------------
    def set_var_defaults(self) -> None:
        if self.rt.exists("config"):
            self.echo("error: old configuration found", lineno=242)
            self.clean_runner()
        v = self.v
        v["KUBERNETES_NAME"] = "k8s dev"
        v["KUBERNETES_DESCRIPTION"] = v["KUBERNETES_NAME"]
        v["RANCHER_MASTER_HOSTNAME"] = "kubemaster"
        v["KUBERNETES_NODE_HOSTNAME_BEGINSWITH"] = "kubenode"
        v["KUBERNETES_NUMBER_OF_NODES"] = "1"
        v["RANCHER_MASTER_NETWORKS"] = ""
        v["KUBERNETES_NODE_NETWORKS"] = ""
        v["HOST_PACKAGE"] = ""
        self.append_line("config", "ANSIBLE_HOST_KEY_CHECKING=False", 253)

    # ------------------------------------------------------------------
    # a2  setConfigFromTritonENV  (setup.sh:209-239)
    # ------------------------------------------------------------------
    def set_config_from_triton(self) -> None:
        rt, v = self.rt, self.v
        _, text = rt.run(["triton", "env"], stdout=CAPTURE, where=self.at(210), lone_subst=True)  # status is swallowed by eval "$(...)"
        try:
            apply_triton_env(self.subst(text, 210), v, rt.env, home=rt.passwd_home, exists=rt.command_exists,
                             provided=self._bash_variable)
        except CommandNotFound as exc:
            rt.err("%s: %s\n" % (self.at(210 + exc.line), exc))
            raise ScriptExit(127)
        except BashDiagnostic as exc:
            # bash's own verdict, not a refusal: eval numbers the lines of its text from the line it stands on
            for message in exc.messages():
                rt.err("%s: %s\n" % (self.at(210 + exc.line), message))
            raise ScriptExit(1)
        except TritonEnvError as exc:
            rt.err("%s: %s\n" % (self.program, exc))
            raise ScriptExit(2)
        for lineno, name in ((211, "SDC_URL"), (212, "SDC_ACCOUNT"), (213, "SDC_KEY_ID")):
            self.append_line("config", '%s="%s"' % (name, v.get(name, "")), lineno)

        with self.scope():
            self.local("foundKey", "false")           # setup.sh:215
            self._find_key()

    def _find_key(self) -> None:
        rt, v = self.rt, self.v
        key_id = v.get("SDC_KEY_ID", "")
        found = False
        for f in self.args_of(self._ls_ssh_dir()):
            # ~/.ssh/$(echo $f | sed 's/.pub$//') -- unquoted, so it can be several words
            stem = sh.cmdsubst(tx.sed_sub(sh.echo_unquoted(f, self.root), r".pub$", "", utf8=self.utf8))
            v["f"] = f                                 # the loop variable is a global of the script
            # tilde expansion yields quoted text: never split, never globbed
            key_args = sh.expand_word([(sh.quote_for_glob(self.tilde("~/.ssh/")), False), (stem, True)], self.root)
            _, out = rt.run(["ssh-keygen", "-E", "md5", "-lf"] + key_args, stdout=CAPTURE, stderr=DEVNULL)
            fp = self.subst(tx.sed_sub(tx.awk_print_field(out, 2), r"^MD5:", ""), 217)
            matched_at = 218
            if fp == key_id:
                found = True
            else:
                matched_at = 225
                # ssh-keygen too old for -E prints MD5 by default (setup.sh:222-229)
                _, out = rt.run(["ssh-keygen", "-l", "-f"] + key_args, stdout=CAPTURE, stderr=DEVNULL)
                if self.subst(tx.grep(tx.awk_print_field(out, 2), r"^SHA256", binary=self._binary_matches, utf8=self.utf8), 223) == "":
                    _, out = rt.run(["ssh-keygen", "-l", "-f"] + key_args, stdout=CAPTURE, stderr=DEVNULL)
                    if self.subst(tx.awk_print_field(out, 2), 224) == key_id:
                        found = True
            if found:
                self.append_line("config", 'SDC_KEY="~/.ssh/%s"' % stem, matched_at)
                v["foundKey"] = "true"
                break
        if not found:
            self.echo("error: couldn't find the ssh key associated with fingerprint %s in ~/.ssh/ directory..." % key_id, lineno=232)
            self.echo("    Clean the setup and make sure your triton profile is set up.", lineno=233)
            self.echo("    To confirm your profile is set up, try `triton info`.", lineno=234)
            self.clean_runner()
            raise ScriptExit(1)
        self.append_line("config", "", 238)

    def _ls_ssh_dir(self) -> str:
        """``$(ls ~/.ssh)``: names one per line, C collation, dot files hidden."""
        d = self.tilde("~/.ssh")
        if os.path.exists(d) and not os.path.isdir(d):
            return d                                  # `ls FILE` prints the name it was given
        try:
            names = os.listdir(d)
        except OSError as exc:
            self.rt.err("ls: cannot access '%s': %s\n" % (d, exc.strerror))
            return ""
        names = sorted((n for n in names if not n.startswith(".")), key=sh.c_sort_key)
        return sh.cmdsubst(tx.unlines(names))

    # ------------------------------------------------------------------
    # a3  getArgument  (setup.sh:94-110)
    # ------------------------------------------------------------------
    def get_argument(self, message: str, default: str, lineno: int = 0) -> str:
        """``$(getArgument "message" "default")``: empty answer (or EOF) = default.
        Both branches ``echo`` unquoted.  ``lineno`` is the reference line of the
        call, which bash names when it drops a NUL byte from the substitution."""
        # getArgument runs inside $(...): the process waiting at this prompt is a subshell, and a SIGQUIT from the
        # terminal kills *it* (status 131, which errexit makes the script's) while the script's own bash ignores it
        self.rt.subshell_read = True
        try:
            answer, _ok = self.rt.read("%s (%s) " % (message, default))
        finally:
            self.rt.subshell_read = False
        printed = sh.echo_unquoted(default if answer == "" else answer, self.root)
        if "\x00" in printed:
            self.rt.err("%s: line %d: warning: command substitution: ignored null byte in input\n" % (self.program, lineno))
        return sh.cmdsubst(printed)

    # ------------------------------------------------------------------
    # a4  getConfigFromUser  (setup.sh:255-451)
    # ------------------------------------------------------------------
    def _menu_rows(self, argv: List[str], kvm_only: bool) -> str:
        """The table a menu is built from: header dropped, runs of blanks
        squeezed, blanks turned into '=', sorted (setup.sh:257,259)."""
        text = self._triton_table(argv, 259 if kvm_only else 257)
        if kvm_only:
            text = tx.grep(text, r"-kvm-", binary=self._binary_matches, utf8=self.utf8)
        text = tx.grep(text, r"^NAME.*ID$", invert=True, binary=self._binary_matches, utf8=self.utf8)
        return sh.cmdsubst(tx.sort_lines(tx.tr(tx.tr_squeeze_spaces(text), " ", "=")))

    def _triton_table(self, argv: List[str], line: int) -> str:
        """Output of ``triton networks|packages -oname,id``.  The reference asks the
        CLI again for every index it resolves (3 + k ``networks`` and 2-3
        ``packages`` calls per run, each a CloudAPI round trip); with
        ``TK8S_CACHE_TRITON=1`` each table is fetched once.  Off by default: it
        shortens the child-process trace, which the parity tests compare."""
        key = tuple(argv)
        if self.cache_triton and key in self._tables:
            return self._tables[key]
        _, text = self.rt.run(argv, stdout=CAPTURE, where=self.at(line))
        self._tables[key] = text
        return text

    def _print_menu(self, rows: str, default_prefix: Optional[str], loop_var: str) -> Tuple[int, int]:
        """Numbered menu; returns (row count, 1-based position of the default row)."""
        count, default_at = 0, 1
        self.v["tmp"] = "0"
        for row in self.args_of(rows):
            self.v[loop_var] = row
            count += 1
            self.v["tmp"] = str(count)
            shown = sh.cmdsubst(tx.sed_sub(sh.echo_unquoted(row, self.root), r"=", "  ", count=0))
            self.write_out(sh.echo_builtin(["-e", "%d.\\t%s" % (count, shown)], self.wide), {"network": 316, "package": 409}.get(loop_var, 0))
            if default_prefix is not None and row.startswith(default_prefix):
                default_at = count
        return count, default_at

    def _prompt_validated(self, message: str, var: str, ok: Callable[[str], bool], errors: List[str],
                          lineno: int) -> None:
        default = self.unq(self.v[var])  # evaluated once per attempt in the reference; it cannot change
        self.v["gotValidInput"] = "false"
        while True:
            answer = self.v["tmp_ValidatedInput"] = self.get_argument(message, default, lineno)
            if ok(answer):
                self.v["gotValidInput"] = "true"
                self.v[var] = answer
                return
            for line in errors:
                self.echo(line)

    def get_config_from_user(self) -> None:
        with self.scope():
            self._get_config_from_user()

    def _get_config_from_user(self) -> None:
        v = self.v
        # `local networks=$(...)`: the substitution runs before the local exists
        networks = self._menu_rows(TRITON_NETWORKS, kvm_only=False)
        self.local("networks", networks)
        packages = self._menu_rows(TRITON_PACKAGES, kvm_only=True)
        self.local("packages", packages)
        self.local("tmp", "0")
        self.local("gotValidInput", "false")
        self.local("tmp_ValidatedInput")

        self.echo(DASHES, lineno=264)
        v["KUBERNETES_NAME"] = self.get_argument("Name your Kubernetes environment:", self.unq(v["KUBERNETES_NAME"]), 265)
        self.echo(DASHES, lineno=266)
        # the description default is whatever the description already holds
        # ("k8s dev" from the defaults), not the name just entered (setup.sh:267-271)
        blank = v["KUBERNETES_DESCRIPTION"] == ""
        descr_default = v["KUBERNETES_NAME"] if blank else v["KUBERNETES_DESCRIPTION"]
        v["KUBERNETES_DESCRIPTION"] = self.get_argument("Describe this Kubernetes environment:", self.unq(descr_default),
                                                        268 if blank else 270)
        self.echo(DASHES, lineno=272)
        self._prompt_validated(
            "Hostname of the master:", "RANCHER_MASTER_HOSTNAME", lambda s: HOSTNAME_RE.fullmatch(s) is not None,
            ["error: Enter a valid hostname or leave blank to use the default.",
             "    Must start with a letter and can only include letters and numbers"], 275)
        self.echo(DASHES, lineno=284)
        self._prompt_validated(
            "Enter a string to use for appending to hostnames of all the nodes:", "KUBERNETES_NODE_HOSTNAME_BEGINSWITH",
            lambda s: HOSTNAME_RE.fullmatch(s) is not None,
            ["error: Enter a valid value or leave blank to use the default.",
             "    Must start with a letter and can only include letters and numbers"], 287)
        self.echo(DASHES, lineno=296)
        # hard limit 1-9: the setup has no HA (setup.sh:296)
        self._prompt_validated(
            "How many nodes should this Kubernetes cluster have:", "KUBERNETES_NUMBER_OF_NODES",
            lambda s: NODE_COUNT_RE.fullmatch(s) is not None,
            ["error: Enter a valid value (number between 1-9) or leave blank to use the default."], 300)

        public_at = 1
        self.local("publicNetworkLocation", "1")      # setup.sh:311-312
        self.local("countNetwork")
        for var, message, first in (
                ("RANCHER_MASTER_NETWORKS", "What networks should the master be a part of, provide comma separated values:", True),
                ("KUBERNETES_NODE_NETWORKS", "What networks should the nodes be a part of, provide comma separated values:", False)):
            lineno = 332 if first else 375
            self.echo(DASHES, lineno=308 if first else 358)
            self.echo("From the networks below:", lineno=309 if first else 359)
            count, at = self._print_menu(networks, "Joyent-SDC-Public=" if first else None, "network")
            if first:
                public_at = at
                v["publicNetworkLocation"], v["countNetwork"] = str(at), str(count)
            if v[var] == "":
                v[var] = self.network_ids(self.args_of(str(public_at)))
            self._prompt_index_list(var, message, count, lineno)

        self.echo(DASHES, lineno=401)
        self.echo("From the packages below:", lineno=402)
        self.local("packageLocation", "1")            # setup.sh:404-405
        self.local("countPackages")
        count, package_at = self._print_menu(packages, "k4-highcpu-kvm-7.75G=", "package")
        v["packageLocation"], v["countPackages"] = str(package_at), str(count)
        if v["HOST_PACKAGE"] == "":
            v["HOST_PACKAGE"] = self.package_id(self.args_of(str(package_at)))
        self._prompt_package(count)
        v["HOST_PACKAGE"] = self.unq(v["HOST_PACKAGE"])

    def _normalise_list(self, value: str) -> str:
        """``echo $V | tr ',' '\\n' | sort | uniq | tr '\\n' ',' | sed 's/\\(.*\\),$/\\1/'``
        -- lexical sort: ``10,2`` stays ``10,2`` (setup.sh:333-334)."""
        t = tx.uniq(tx.sort_lines(tx.tr(sh.echo_unquoted(value, self.root), ",", "\n")))
        return sh.cmdsubst(tx.sed_sub(tx.tr(t, "\n", ","), r"(.*),$", r"\1", utf8=self.utf8))

    def _out_of_range(self, items: List[str], count: int, var: str, lineno: int) -> int:
        """How many of ``items`` fall outside 1..count by
        ``[[ "$item" -gt "$count" || "$item" -lt 1 ]]`` (setup.sh:340,383,431).
        ``-gt``/``-lt`` inside ``[[ ]]`` evaluate their operands as arithmetic
        expressions; one bash cannot evaluate makes that test false after a
        message, so such an item counts as in range -- twice, once per test."""
        bad = 0
        for item in items:
            self.v[var] = item                      # the loop variable is a global of the script
            n = self.arith_eval(item, lineno, "[[")
            if n is not None and n > count:
                bad += 1
                continue
            n = self.arith_eval(item, lineno, "[[")
            if n is not None and n < 1:
                bad += 1
        return bad

    def _prompt_index_list(self, var: str, message: str, count: int, lineno: int) -> None:
        v = self.v
        errors = ["error: Enter a valid option or leave blank to use the default.",
                  "    Values should be comma separated between 1 and %d." % count]
        v["gotValidInput"] = "false"
        while True:
            answer = self.get_argument(message, self.unq(v[var]), lineno)
            v[var] = self._normalise_list(v[var])
            answer = v["tmp_" + var] = self._normalise_list(answer)
            shown = sh.echo_unquoted(answer, self.root)
            if sh.cmdsubst(tx.grep(shown, INDEX_LIST_RE)) != "":
                v["gotValidInput"] = "true"
                items = self.args_of(sh.cmdsubst(tx.tr(shown, ",", "\n")))
                bad = self._out_of_range(items, count, "network", lineno + 8)
                if bad:
                    v["gotValidInput"] = "false"
                for _ in range(bad):
                    for line in errors:
                        self.echo(line)
                if bad == 0:
                    v[var] = self.network_ids(self.args_of(answer))
                    return
            elif sh.pattern_match(answer, v[var], self.shell_utf8):
                v["gotValidInput"] = "true"
                return  # the default UUID list was left as it is
            else:
                for line in errors:
                    self.echo(line)

    def _prompt_package(self, count: int) -> None:
        v = self.v
        errors = ["error: Enter a valid option or leave blank to use the default.",
                  "    Value should be between 1 and %d." % count]
        v["gotValidInput"] = "false"
        while True:
            answer = v["tmp_HOST_PACKAGE"] = self.get_argument("What KVM package should the master and nodes run on:",
                                                               self.unq(v["HOST_PACKAGE"]), 425)
            shown = sh.echo_unquoted(answer, self.root)
            if sh.cmdsubst(tx.grep(shown, INDEX_RE)) != "":
                v["gotValidInput"] = "true"
                items = self.args_of(sh.cmdsubst(tx.tr(shown, ",", "\n")))
                bad = self._out_of_range(items, count, "package", 431)
                if bad:
                    v["gotValidInput"] = "false"
                for _ in range(bad):
                    for line in errors:
                        self.echo(line)
                if bad == 0:
                    v["HOST_PACKAGE"] = self.package_id(self.args_of(answer))
                    self.echo("entered %s and got %s" % (answer, v["HOST_PACKAGE"]), lineno=440)
                    return
            elif sh.pattern_match(answer, self.unq(v["HOST_PACKAGE"]), self.shell_utf8):
                v["gotValidInput"] = "true"
                return
            else:
                for line in errors:
                    self.echo(line)

    # ------------------------------------------------------------------
    # a5/a6  getNetworkIDs / getPackageID  (setup.sh:532-542)
    # ------------------------------------------------------------------
    def network_ids(self, args: List[str]) -> str:
        """Index list -> UUID list, one ``triton networks`` call per index.  The
        rows are sorted *raw* here but after the blank->'=' rewrite in the menu
        (setup.sh:257 vs :536), which is the ordering hazard of SURVEY.md §7.4."""
        arg1 = args[0] if args else ""
        with self.scope():                            # runs inside $(...): its assignments do not outlive it
            values = sh.cmdsubst(tx.tr(sh.echo_unquoted(arg1, self.root), ",", " "))
            self.subshell_set("values", values)
            self.local("networks")
            acc = ""
            for index in self.args_of(values):
                self.subshell_set("network", index)   # what `triton` finds in its environment if `network` is exported
                text = self._triton_table(TRITON_NETWORKS, 536)
                rows = tx.grep(tx.sort_lines(text), r"^NAME *ID$", invert=True, binary=self._binary_matches, utf8=self.utf8)
                acc = acc + "," + self.subst(tx.awk_last_field_if_many(self._nth_row(rows, index, 536)), 536)
                self.v["networks"] = acc
        t = tx.sed_sub(sh.echo_builtin([acc]), r"^,(.*)$", r"\1", utf8=self.utf8)
        return sh.cmdsubst(tx.sed_sub(t, r"(.*),$", r"\1", utf8=self.utf8))

    def package_id(self, args: List[str]) -> str:
        arg1 = args[0] if args else ""
        text = self._triton_table(TRITON_PACKAGES, 541)
        rows = tx.grep(tx.grep(text, r"-kvm-", binary=self._binary_matches, utf8=self.utf8), r"^NAME.*ID$", invert=True, binary=self._binary_matches, utf8=self.utf8)
        rows = tx.sort_lines(tx.tr_squeeze_spaces(rows))
        picked = self.subst(tx.awk_last_field_if_many(self._nth_row(rows, arg1)), 541)
        return sh.cmdsubst(sh.echo_builtin([picked]))

    def _nth_row(self, rows: str, index: str, lineno: int = 541) -> str:
        """``sed -n "${index}p"``.  The index has been through the script's own check, which lets through numbers
        and -- an answer such as ``-e \\055ne 1`` leaves ``-ne 1`` behind -- words `echo` takes for options:
        `sed -n "-ne"p` is `sed -n -n -e p`."""
        script = index + "p"
        if len(index) > 1 and index.startswith("-"):
            script = None
            for k, flag in enumerate(index[1:], 1):
                if flag in "nE":
                    continue                          # -n again; -E: extended regular expressions, of which there are none
                if flag == "e":
                    script = index[k + 1:] + "p"      # the rest of the word is the script
                    break
                self._refuse_sed(script=index + "p", lineno=lineno)
            if script is None:
                self.rt.err("sed: invalid option -- 'p'\n" + tx.SED_USAGE)
                return ""
        if script == "p":
            return rows                               # `sed -n p` prints every row
        m = re.fullmatch(r"([0-9]+)p", script)
        if m:
            return tx.sed_print_line(rows, int(m.group(1)) & 0xFFFFFFFFFFFFFFFF)  # GNU sed line numbers wrap at 2**64
        head = script.lstrip("0123456789")
        at = len(script) - len(head) + 1
        if head[0] == "n":                            # `n` wants `;`, a newline or the end after it
            self.rt.err("sed: -e expression #1, char %d: extra characters after command\n" % (at + 1))
            return ""
        if head[0] not in "{}=abcdDeFgGhHilnNpPqQrRstTvwWxyz:#/\\$!~,+; \t\n":
            self.rt.err("sed: -e expression #1, char %d: unknown command: `%s'\n" % (at, head[0]))
            return ""
        self._refuse_sed(script, lineno)

    def _refuse_sed(self, script: str, lineno: int):
        self.rt.err("%s: refusing to go on: the index spliced into the sed program at line %d makes it %s, which is not restated here\n"
                    % (self.program, lineno, repr(script)))
        raise ScriptExit(2)

    # ------------------------------------------------------------------
    # a7  verifyConfig  (setup.sh:452-483)
    # ------------------------------------------------------------------
    def verify_config(self) -> None:
        v = self.v
        self.echo(RULE, lineno=453)
        self.echo("Verify that the following configuration is correct:", lineno=454)
        self.echo("", lineno=455)
        self.echo("Name of kubernetes environment: %s" % v["KUBERNETES_NAME"], lineno=456)
        self.echo("Kubernetes environment description: %s" % v["KUBERNETES_DESCRIPTION"], lineno=457)
        self.echo("Master hostname: %s" % v["RANCHER_MASTER_HOSTNAME"], lineno=458)
        self.echo("All node hostnames will start with: %s" % v["KUBERNETES_NODE_HOSTNAME_BEGINSWITH"], lineno=459)
        self.echo("Kubernetes environment will have %s nodes" % v["KUBERNETES_NUMBER_OF_NODES"], lineno=460)
        self.echo("Master server will be part of these networks: %s" % v["RANCHER_MASTER_NETWORKS"], lineno=461)
        self.echo("Kubernetes nodes will be a part of these networks: %s" % v["KUBERNETES_NODE_NETWORKS"], lineno=462)
        self.echo("This package will be used for all the hosts: %s" % v["HOST_PACKAGE"], lineno=463)
        self.echo("", lineno=464)
        self.echo("Make sure the above information is correct before answering:", lineno=465)
        self.echo('    to view list of networks call "triton networks -l"', lineno=466)
        self.echo('    to view list of packages call "triton packages -l"', lineno=467)
        self.echo("WARN: Make sure that the nodes and master are part of networks that can communicate with each other and this system from which the setup is running.", lineno=468)
        while True:
            yn, ok = self.rt.read("Is the above config correct (yes | no)? ")
            v["yn"] = yn
            if not ok:
                raise ScriptExit(1)  # read fails at EOF and errexit ends the run
            if yn == "yes":
                return
            if yn == "no":
                raise ScriptExit(0)  # ./config stays behind
            self.echo("Please answer yes or no.", lineno=480)

    # ------------------------------------------------------------------
    # a8  setConfigToFile  (setup.sh:199-208)
    # ------------------------------------------------------------------
    def set_config_to_file(self) -> None:
        v = self.v
        for k, name in enumerate(("RANCHER_MASTER_NETWORKS", "KUBERNETES_NODE_NETWORKS", "KUBERNETES_NUMBER_OF_NODES")):
            self.append_line("config", "%s=%s" % (name, v[name]), 200 + k)
        for k, name in enumerate(("KUBERNETES_NAME", "KUBERNETES_DESCRIPTION", "RANCHER_MASTER_HOSTNAME",
                                  "KUBERNETES_NODE_HOSTNAME_BEGINSWITH", "HOST_PACKAGE")):
            self.append_line("config", '%s="%s"' % (name, v[name]), 203 + k)

    # ------------------------------------------------------------------
    # a9  exportVars  (setup.sh:543-549)
    # ------------------------------------------------------------------
    def export_vars(self) -> None:
        """Every non-blank line of ./config is exported *verbatim*: values keep
        their literal double quotes from here on, and each line
        goes through ``read`` once more (backslashes processed again)."""
        # grep -v "^$" config > config.tmp: the redirection comes first, then grep's own troubles
        try:
            self.rt.write_text("config.tmp", "")
        except OSError as exc:
            self._redirect_failed("config.tmp", exc, 544)
        try:
            raw = self.rt.read_text("config")
        except OSError as exc:
            self.rt.err("grep: config: %s\n" % exc.strerror)
            raise ScriptExit(2)
        text = tx.grep(raw, r"^$", invert=True, utf8=self.utf8, binary=lambda: self._binary_matches("config"))
        self.rt.write_text("config.tmp", text)  # stays behind if an export fails (errexit)
        if not any(line != "" for line in tx.lines(raw)):
            raise ScriptExit(1)                 # grep selected nothing: status 1, errexit
        # done <config.tmp: what is read back is what the file holds (a link to /dev/null holds nothing)
        data = sh.to_bytes(self.rt.read_text("config.tmp"))
        pos = 0

        def next_byte() -> Optional[int]:
            nonlocal pos
            if pos >= len(data):
                return None
            pos += 1
            return data[pos - 1]

        while True:
            line, ok = sh.read_builtin(next_byte)
            self.v["line"] = line
            if not ok:
                break
            self._export(line)
        os.remove(self.rt.path("config.tmp"))

    def _export(self, word: str) -> None:
        if len(word) > 1 and word.startswith("-"):
            return self._export_options(word)
        name, eq, value = word.partition("=")
        append = name.endswith("+") and bool(eq)
        if append:
            name = name[:-1]
        if not _IDENT.fullmatch(name):
            self.rt.err("%s: line 546: export: `%s': not a valid identifier\n" % (self.program, word))
            raise ScriptExit(1)
        if eq and name in _READONLY:
            self.rt.err("%s: %s: readonly variable\n" % (self.at(546), name))
            raise ScriptExit(1)
        if name in _NOT_MODELLED or (eq and name in SHELL_BEHAVIOUR) or (not eq and name in _SET_AT_STARTUP and name not in self.v):
            self.rt.err("%s: refusing to go on: a line of ./config exports %s, which %s\n"
                        % (self.program, name, "changes how bash reads the rest of the script" if name in SHELL_BEHAVIOUR
                           else "bash computes by itself (its value is not restated here)"))
            raise ScriptExit(2)
        if eq and name == "OPTIND":
            # an integer variable: what is assigned is the value of the text as an arithmetic expression
            number = self.arith_eval(value, 546, "export")
            if number is None:
                raise ScriptExit(1)
            value = str(number + (int(self._arith_get("OPTIND") or 0) if append else 0))
            append = False
        if eq:
            self.v[name] = (self.v.get(name, "") + value) if append else value
        if name in self.v:
            self.rt.env[name] = self.v[name]
        elif self._bash_variable(name) is not None:
            self.rt.env[name] = self._bash_variable(name)       # UID, PPID, IFS ...: bash has a value for it
        else:
            self.v.export_when_set.add(name)    # the attribute waits for a value (`i`, `master` ... get one later)

    def _export_options(self, word: str) -> None:
        """``export "-..."``: a line of ./config that begins with a dash is the *options* of export (one word, so
        there are no names behind them).  ``-p``, ``-n``, ``--`` list what is exported; ``-f``/``-a``/``-A`` list
        the exported functions/arrays, of which there are none; any other letter is a usage error, status 2."""
        if word == "--help":
            self.write_out(_EXPORT_HELP, 546)
            raise ScriptExit(2)
        if word != "--":
            for flag in word[1:]:
                if flag not in "aAfnp":
                    self.rt.err("%s: export: -%s: invalid option\nexport: usage: export [-fn] [name[=value] ...] or export -p\n"
                                % (self.at(546), flag))
                    raise ScriptExit(2)
            if any(flag in "aAf" for flag in word[1:]):
                return
        env, pending = self.rt.env, self.v.export_when_set
        listing = []
        for name in sorted((set(env) | pending) - {"_"}, key=sh.to_bytes):
            if not _IDENT.fullmatch(name):
                continue                            # never became a shell variable
            listing.append("declare -x %s=%s\n" % (name, sh.declare_value(env[name], self.shell_utf8)) if name in env else "declare -x %s\n" % name)
        self.write_out("".join(listing), 546)

    # ------------------------------------------------------------------
    # a10-a13  Terraform root config + run  (setup.sh:138-198)
    # ------------------------------------------------------------------
    def run_terraform_tasks(self) -> None:
        v, rt = self.v, self.rt
        if rt.exists("terraform/rancher.tf"):
            # unreachable from main (guarded earlier) but part of the function
            self.echo("warning: a previous terraform configuration has been found", lineno=141)
            self.echo("    skipping terraform configuration and execution...", lineno=142)
            return
        self.echo("Generating terraform configs for environment...", lineno=144)
        self.update_terraform_config(["provider", "triton"])
        self.echo("    Master hostname: %s" % v.get("RANCHER_MASTER_HOSTNAME", ""))  # quotes and all (setup.sh:146, lineno=146)
        self.update_terraform_config(["master"] + self.args_of(self.unq(v.get("RANCHER_MASTER_HOSTNAME", ""))))
        # for (( i = 1; i <= $KUBERNETES_NUMBER_OF_NODES; i++ )): the count is text
        # from ./config spliced into the test expression, `i` a global of the script
        self.arith_eval("i = 1", 148, "((")
        while True:
            more = self.arith_eval("i <= %s" % v.get("KUBERNETES_NUMBER_OF_NODES", ""), 148, "((")
            if more is None or more == 0:
                break                 # an arithmetic error fails the `for` command, which errexit lets pass

            i = v.get("i", "")
            prefix = v.get("KUBERNETES_NODE_HOSTNAME_BEGINSWITH", "")
            self.echo("    Kubernetes node %s: %s%s" % (i, self.unq(prefix), i), lineno=150)
            self.update_terraform_config(["host"] + self.args_of(self.unq("%s%s" % (prefix, i))))
            self.arith_eval("i++", 148, "((")
        self.cd("terraform")
        self.echo("Starting terraform tasks", lineno=155)
        self.must(rt.run(["terraform", "get"], cwd="terraform", where=self.at(156), job=(self.at(138), "terraform get"))[0])
        self.must(rt.run(["terraform", "apply"], cwd="terraform", where=self.at(157), job=(self.at(138), "terraform apply"))[0])
        self.echo("    terraform tasks completed", lineno=158)
        self.cd("..")

    def update_terraform_config(self, args: List[str]) -> None:
        """One block of ``terraform/rancher.tf``.  ``provider`` truncates the
        file, the module blocks append (setup.sh:162-198)."""
        v = self.v
        kind = args[0] if args else ""
        label = args[1] if len(args) > 1 else ""
        key = self.unq(v.get("SDC_KEY", ""))
        rel = "terraform/rancher.tf"
        if kind in ("master", "host"):
            nets = v.get("RANCHER_MASTER_NETWORKS" if kind == "master" else "KUBERNETES_NODE_NETWORKS", "")
            net_list = sh.cmdsubst(tx.sed_sub(sh.echo_unquoted(nets, self.root), r",", '","', count=0))
            block = [
                "",
                'module "%s" {' % label,
                '    source = "%s"' % kind,
                '    hostname = "%s"' % label,
                '    networks = ["%s"]' % net_list,
                '    root_authorized_keys = "${file("%s")}"' % key,   # the private key path: reference quirk, kept
                '    package = "%s"' % self.unq(v.get("HOST_PACKAGE", "")),
                "}",
            ]
            first = 164 if kind == "master" else 176
            for k, line in enumerate(block):
                # the commented-out image line (setup.sh:170,182) sits between the key and the package line
                self.append_line(rel, line, first + k + (1 if k >= 6 else 0))
            return
        if kind == "provider":
            block = [
                'provider "triton" {',
                '    account = "%s"' % self.unq(v.get("SDC_ACCOUNT", "")),
                '    key_material = "${file("%s")}"' % key,
                '    key_id = "%s"' % self.unq(v.get("SDC_KEY_ID", "")),
                '    url = "%s"' % self.unq(v.get("SDC_URL", "")),
                "}",
            ]
            try:
                self.rt.write_text(rel, sh.echo_builtin([block[0]]))
            except OSError as exc:
                self._redirect_failed(rel, exc, 188)
            for k, line in enumerate(block[1:]):
                self.append_line(rel, line, 189 + k)
            return
        self.echo("error: problem updating terraform configuration...", lineno=196)
        raise ScriptExit(1)

    # ------------------------------------------------------------------
    # a15  createAnsibleConfigs  (setup.sh:116-137)
    # ------------------------------------------------------------------
    def create_ansible_configs(self) -> None:
        rt, v = self.rt, self.v
        if not rt.exists("terraform/masters.ip") or not rt.exists("terraform/hosts.ip"):
            self.echo("Terraform had too many errors. Make sure you haven't reached your provisioning limit.", lineno=118)
            raise ScriptExit(1)
        self.echo("Creating ansible hosts file and variable files", lineno=121)
        hosts = "ansible/hosts"
        if os.path.isdir(rt.path(hosts)) and not os.path.islink(rt.path(hosts)):
            raise ScriptExit(1)  # `rm -f` cannot remove a directory (message hidden by 2>/dev/null), errexit
        if os.path.lexists(rt.path(hosts)):
            os.remove(rt.path(hosts))
        self.append_line(hosts, "[MASTER]", 123)
        self._cat_append("terraform/masters.ip", hosts, 124)
        self.append_line(hosts, "[HOST]", 125)
        self._cat_append("terraform/hosts.ip", hosts, 126)
        self.echo("    created: ansible/hosts", lineno=127)

        vars_yml = "ansible/roles/ranchermaster/vars/vars.yml"
        try:
            last = tx.tail_1(rt.read_text("terraform/masters.ip"))
        except OSError as exc:
            rt.err("tail: error reading 'terraform/masters.ip': %s\n" % exc.strerror)
            raise ScriptExit(1)                 # master=$(tail ...): the assignment has tail's status, errexit
        master = v["master"] = self.subst(last, 128)
        try:
            rt.write_text(vars_yml, sh.echo_builtin(["master: %s" % master]))
        except OSError as exc:
            self._redirect_failed(vars_yml, exc, 129)
        self.append_line(vars_yml, 'kubernetes_name: "%s"' % self.unq(v.get("KUBERNETES_NAME", "")), 130)
        self.append_line(vars_yml, 'kubernetes_description: "%s"' % self.unq(v.get("KUBERNETES_DESCRIPTION", "")), 131)

        self.cd("ansible")
        self._rewrite_private_key_file("ansible/ansible.cfg", "ansible/tmp.cfg", ";", self.unq(v.get("SDC_KEY", "")),
                                       shown_src="ansible.cfg")
        self.cd("..")
        self.echo("    created: ansible/roles/ranchermaster/vars/vars.yml", lineno=136)

    def _cat_append(self, src: str, dst: str, lineno: int) -> None:
        """``cat src >> dst`` under errexit."""
        try:
            self.rt.append_text(dst, "")        # the redirection is set up before cat runs
        except OSError as exc:
            self._redirect_failed(dst, exc, lineno)
        try:
            text = self.rt.read_text(src)
        except OSError as exc:
            self.rt.err("cat: %s: %s\n" % (src, exc.strerror))
            raise ScriptExit(1)
        self.rt.append_text(dst, text)

    def _rewrite_private_key_file(self, cfg: str, tmp: str, delim: str, value: str, shown_src: str,
                                  lineno: int = 133) -> None:
        """``sed "s<d>private_key_file = .*$<d>private_key_file = VALUE<d>g" cfg > tmp && mv tmp cfg``
        (setup.sh:133 with ';', setup.sh:511 with '~').  VALUE is spliced into the sed
        program text, so sed's own replacement syntax applies to it."""
        rt = self.rt
        try:
            rt.write_text(tmp, "")  # the redirection creates/truncates the target first
        except OSError as exc:
            rt.err("%s: line %d: %s: %s\n" % (self.program, lineno, os.path.basename(tmp), exc.strerror))
            return
        try:
            text = rt.read_text(cfg)
            readable = True
        except OSError as exc:
            text, readable, why = "", False, exc.strerror
        res = sedprog.run(text, value, delim, utf8=self.utf8)
        if res.refused:
            rt.err("%s: refusing to go on: the key path spliced into the sed program at line %d makes it %s\n"
                   % (self.program, lineno, res.refused))
            raise ScriptExit(2)
        if not res.ok:
            rt.err(res.error)        # a program sed rejects is reported before it opens any input
            return
        if not readable:
            if os.path.isdir(rt.path(cfg)):
                rt.err("sed: read error on %s: %s\n" % (shown_src, why))     # a directory opens fine and fails on read
            else:
                rt.err("sed: can't read %s: %s\n" % (shown_src, why))
            return
        for name, content in res.wfiles.items():     # a `w` flag smuggled in through VALUE writes next to the input
            rt.write_text(os.path.join(os.path.dirname(tmp), name), content)
        rt.write_text(tmp, res.output)
        # mv tmp cfg: into cfg when that is a directory; the last command of an && list is subject to errexit
        dest, shown_dest = rt.path(cfg), shown_src
        shown_tmp = os.path.basename(tmp) if lineno == 133 else tmp
        try:
            same = os.path.samefile(rt.path(tmp), dest)
        except OSError:
            same = False
        if same:                                     # tmp is a link to cfg: sed has just emptied it through the link
            rt.err("mv: '%s' and '%s' are the same file\n" % (shown_tmp, shown_dest))
            raise ScriptExit(1)
        if os.path.isdir(dest):
            dest = os.path.join(dest, os.path.basename(tmp))
            shown_dest = "%s/%s" % (shown_src, os.path.basename(tmp))
            if os.path.isdir(dest) and not os.path.islink(dest):
                rt.err("mv: cannot overwrite directory '%s' with non-directory\n" % shown_dest)
                raise ScriptExit(1)
        try:
            os.replace(rt.path(tmp), dest)
        except OSError as exc:
            rt.err("mv: cannot move '%s' to '%s': %s\n" % (shown_tmp, shown_dest, exc.strerror))
            raise ScriptExit(1)

    # ------------------------------------------------------------------
    # runAnsible  (setup.sh:111-115)
    # ------------------------------------------------------------------
    def run_ansible(self) -> None:
        if not os.path.isdir(self.rt.path("ansible")):
            self.rt.err("%s: line 112: cd: ansible: No such file or directory\n" % self.program)
            raise ScriptExit(1)
        self.cd("ansible")
        self.must(self.rt.run(["ansible-playbook", "-i", "hosts", "clusterUp.yml"], cwd="ansible", where=self.at(113),
                                   job=(self.at(111), "ansible-playbook -i hosts clusterUp.yml"))[0])
        self.cd("..")

    # ------------------------------------------------------------------
    # dashboard wait loop  (setup.sh:59-85)
    # ------------------------------------------------------------------
    def _dashboard_url_words(self, line: int = 0) -> List[str]:
        """``http://$(cat terraform/masters.ip):8080/r/projects/$(cat …id)/kubernetes-dashboard:9090/``
        as an unquoted word."""
        return sh.expand_word([
            ("http://", False), (self.cat_subst("terraform/masters.ip", line), True),
            (":8080/r/projects/", False), (self.cat_subst("ansible/tmp/kubernetes_environment.id", line), True),
            ("/kubernetes-dashboard:9090/", False)], self.root)

    def _curl_dashboard(self, line: int) -> Tuple[int, str]:
        # cRetVal=$(curl ...) at :66 holds curl alone; at :74 curl heads a pipeline
        return self.rt.run(["curl", "--connect-timeout", "5", "--max-time", "5", "-s"] + self._dashboard_url_words(line),
                           stdout=CAPTURE, where=self.at(line), lone_subst=line == 66)

    def _ssh_docker_ps(self, node: str, line: int) -> str:
        return self.rt.run(["ssh", "-o", "StrictHostKeyChecking=no", "root@%s" % node, "docker", "ps"],
                           stdout=CAPTURE, where=self.at(line))[1]

    def wait_for_dashboard(self) -> None:
        """Poll once a second; every time the epoch is a multiple of 15 ask the
        dashboard URL.  A body mentioning "kubernetes" ends the wait.  Otherwise,
        when the nodes run exactly two dashboard containers and the proxy
        answers "Service Unavailable", stop the stuck pod container on every
        node that has one (upstream-bug workaround, setup.sh:69-82)."""
        rt, v = self.rt, self.v
        v["KUBERNETES_DASHBOARD_UP"], v["DASHBOARD_CONTAINER_COUNT"] = "false", "0"
        up = False
        while not up:
            self.write_out(".", 62)
            self.must(rt.sleep(1, self.at(63), job=(self.at(8), "sleep 1")))
            # if ! ((`date +%s` % 15)): the clock reading is spliced into an arithmetic
            # expression.  Zero -- or an expression bash cannot evaluate -- means "poll now".
            date_status, now = rt.epoch(self.at(64))
            v["_"] = "1"                               # last argument of `sleep 1`
            remainder = self.arith_eval("%s %% 15" % self.subst(now, 64), 64, "((", date_status)
            if remainder is not None and remainder != 0:
                continue
            v["DASHBOARD_CONTAINER_COUNT"] = "0"
            status, body = self._curl_dashboard(66)
            self.must(status)  # cRetVal=$(curl …): a failing curl ends the script
            v["cRetVal"] = self.subst(body, 66)
            shown = sh.echo_unquoted(v["cRetVal"], self.root)
            if tx.wc_l(tx.grep(shown, r"kubernetes", ignore_case=True, binary=self._binary_matches, utf8=self.utf8)) != 0:
                up = True
                v["KUBERNETES_DASHBOARD_UP"] = "true"
                continue
            count = 0
            for node in self.args_of(self.cat_subst("terraform/hosts.ip", 71)):
                v["node"] = node
                ps = tx.grep(self._ssh_docker_ps(node, 72), r"kubernetes-dashboard", binary=self._binary_matches, utf8=self.utf8)
                count += tx.wc_w(tx.awk_print_field(ps, 1), utf8=self.utf8)
                v["DASHBOARD_CONTAINER_COUNT"] = str(count)
            if count != 2:
                continue
            _, body = self._curl_dashboard(74)
            if tx.wc_l(tx.grep(body, r"Service Unavailable", ignore_case=True, binary=self._binary_matches, utf8=self.utf8)) == 0:
                continue
            for node in self.args_of(self.cat_subst("terraform/hosts.ip", 75)):
                v["node"] = node
                ps = tx.grep(self._ssh_docker_ps(node, 76), r"k8s_POD.*dashboard", binary=self._binary_matches, utf8=self.utf8)
                container = v["dashboard_container"] = self.subst(tx.awk_print_field(ps, 1), 76)
                if container != "":
                    status, _ = rt.run(["ssh", "-o", "StrictHostKeyChecking=no", "root@%s" % node,
                                        "docker stop -t0 %s" % container], stdout=DEVNULL, stderr=DEVNULL)
                    self.must(status)
                self.write_out(".", 80)

    # ------------------------------------------------------------------
    # cleanRunner  (setup.sh:484-521)
    # ------------------------------------------------------------------
    def clean_runner(self) -> None:
        rt = self.rt
        self.echo("Clearing settings....", lineno=485)
        while True:
            if rt.exists("terraform/masters.ip"):
                self.echo("WARNING: You are about to destroy the following KVMs associated with Rancher cluster:", lineno=488)
                for rel in ("terraform/masters.ip", "terraform/hosts.ip"):
                    try:
                        rt.out(rt.read_text(rel))
                    except OSError:
                        raise ScriptExit(1)  # `cat … 2>/dev/null` still fails (unreadable, or stdout gone): errexit
                yn, ok = rt.read("Do you wish to destroy the KVMs and reset configuration (yes | no)? ")
            else:
                yn, ok = rt.read("Do you wish to reset configuration (yes | no)? ")
            self.v["yn"] = yn
            if not ok:
                raise ScriptExit(1)
            if yn == "yes":
                self._clean()
                self.echo("    All clear!", lineno=515)
                return
            if yn == "no":
                raise ScriptExit(0)
            self.echo("Please answer yes or no.", lineno=518)

    def _clean(self) -> None:
        rt = self.rt
        if rt.exists("terraform/rancher.tf"):
            self.cd("terraform")
            self.echo("    destroying images...", lineno=500)
            rt.run(["terraform", "destroy", "-force"], cwd="terraform", stderr=DEVNULL,
                   job=(self.at(496), "terraform destroy -force 2> /dev/null"))  # `|| true`
            self.cd("..")
        if (rt.exists("terraform/hosts.ip") and rt.exists("terraform/masters.ip")
                and os.path.exists(self.tilde("~/.ssh/known_hosts"))):
            recorded = ""
            for rel in ("terraform/hosts.ip", "terraform/masters.ip"):      # $(cat a b) in a `for` list: errors only cost a message
                try:
                    recorded += rt.read_text(rel)
                except OSError as exc:
                    rt.err("cat: %s: %s\n" % (rel, exc.strerror))
            recorded = self.subst(recorded, 505)
            for host_key in self.args_of(recorded):
                self.v["host_key"] = host_key
                # `2>&1 >> /dev/null`: stderr joins the script's stdout, stdout is dropped
                status, _ = rt.run(["ssh-keygen", "-R"] + self.args_of(host_key), stdout=DEVNULL, stderr=TO_STDOUT, where=self.at(506),
                                    job=(self.at(505), "ssh-keygen -R $host_key 2>&1 >> /dev/null"))
                self.must(status)
        for rel in ("terraform/hosts.ip", "terraform/masters.ip", "terraform/rancher.tf"):
            _rm_rf(rt.path(rel))
        for pattern in ("terraform/terraform.*", "terraform/.terraform*"):
            for rel in sh.glob_word(pattern, self.root):
                _rm_rf(rt.path(rel))
        self._rewrite_private_key_file("ansible/ansible.cfg", "tmp", "~", "", shown_src="ansible/ansible.cfg", lineno=511)
        # rm -f ansible/hosts ansible/*retry ansible/ansible.cfg.tmp 2>&1 >> /dev/null: rm goes through
        # all its arguments, complains on the script's *stdout* (2>&1 comes first) and fails at the end
        failed = False
        for rel in ["ansible/hosts"] + sh.glob_word("ansible/*retry", self.root) + ["ansible/ansible.cfg.tmp"]:
            p = rt.path(rel)
            if os.path.isdir(p) and not os.path.islink(p):
                try:
                    rt.out("rm: cannot remove '%s': Is a directory\n" % rel)
                except OSError:
                    pass
                failed = True
            elif os.path.lexists(p):
                os.remove(p)
        if failed:
            raise ScriptExit(1)
        _rm_rf(rt.path("config"))
        for rel in sh.glob_word("tmp/*", self.root):
            _rm_rf(rt.path(rel))


def _rm_rf(path: str) -> None:
    if os.path.isdir(path) and not os.path.islink(path):
        shutil.rmtree(path, ignore_errors=True)
    elif os.path.lexists(path):
        try:
            os.remove(path)
        except OSError:
            pass
