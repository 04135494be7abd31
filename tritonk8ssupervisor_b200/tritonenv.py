"""Apply the output of ``triton env`` to the shell state.

The reference does ``eval "$(triton env)"`` (setup.sh:210), i.e. it executes
whatever the CLI prints.  What the CLI prints is a fixed, tiny dialect --
``export NAME="value"``, ``unset NAME`` and ``#`` comments -- and that dialect,
widened by everything that executes nothing (quoting, ``NAME=value``, ``$NAME``
and ``${NAME}`` references, ``~`` in assignments), is what is accepted here.
What would run something -- command substitution, backquotes, arithmetic,
pipes, redirections, commands -- is refused loudly instead of executed.  That is
the one deliberate narrowing of the boundary ("Deviations").
"""
from __future__ import annotations

import pwd
import re
from typing import Callable, Dict, List, Optional, Tuple, Union

_NAME = re.compile(r"[A-Za-z_][A-Za-z0-9_]*\Z")
_NAME_AT = re.compile(r"[A-Za-z_][A-Za-z0-9_]*")


class TritonEnvError(ValueError):
    pass


class CommandNotFound(TritonEnvError):
    """The text names a command that does not exist: nothing would run, bash says
    ``NAME: command not found`` (status 127, which errexit makes the script's)."""

    def __init__(self, name: str, line: int):
        super().__init__("%s: command not found" % ansic_quote(name))
        self.name, self.line = name, line


#: words bash does not look up on PATH: builtins, keywords, and the functions of setup.sh
_SHELL_WORDS = frozenset((
    ". : [ alias bg bind break builtin caller cd command compgen complete compopt continue declare dirs disown echo enable "
    "eval exec exit export false fc fg getopts hash help history jobs kill let local logout mapfile popd printf pushd pwd read "
    "readarray readonly return set shift shopt source suspend test times trap true type typeset ulimit umask unalias unset wait "
    "if then else elif fi case esac for select while until do done in function time { } ! [[ ]] coproc "
    "main getArgument runAnsible createAnsibleConfigs runTerraformTasks updateTerraformConfig setConfigToFile "
    "setConfigFromTritonENV setVarDefaults getConfigFromUser verifyConfig cleanRunner debugVars getNetworkIDs getPackageID "
    "exportVars").split())


def ansic_quote(text: str) -> str:
    """How bash prints a word that holds non-printable bytes: ``$'exp\\377ort'``."""
    data = text.encode("utf-8", "surrogateescape")
    if all(0x20 <= b < 0x7F for b in data):
        return text
    names = {7: "\\a", 8: "\\b", 27: "\\E", 12: "\\f", 10: "\\n", 13: "\\r", 9: "\\t", 11: "\\v", 92: "\\\\", 39: "\\'"}
    return "$'" + "".join(names.get(b) or (chr(b) if 0x20 <= b < 0x7F else "\\%03o" % b) for b in data) + "'"


class BashDiagnostic(TritonEnvError):
    """Not a refusal: what bash itself says to the command without running anything (status 1, which errexit
    makes the script's).  ``line`` is the 0-based line of the text the command started on."""

    def __init__(self, messages: List[str], line: int):
        super().__init__("; ".join(messages))
        self._messages, self.line = list(messages), line

    def messages(self) -> List[str]:
        return list(self._messages)


class InvalidIdentifier(BashDiagnostic):
    """``export WORD`` where WORD is no ``NAME[=value]``: not a refusal -- bash itself
    rejects it (``export: `WORD': not a valid identifier``, status 1, which errexit makes
    the script's).  ``line`` is the 0-based line of the text the command started on."""

    def __init__(self, words: List[str], line: int):
        super().__init__(["export: `%s': not a valid identifier" % w for w in words], line)
        #: every word of the command that was rejected (the valid ones have taken effect)
        self.words = words


#: variables that change how bash reads the *rest of the script* (field splitting,
#: globbing, locale of its pattern matching and of sort/grep/sed).  `eval` would let
#: `triton env` set them; the path here is written for their default values.
SHELL_BEHAVIOUR = frozenset(["IFS", "GLOBIGNORE", "POSIXLY_CORRECT", "BASH_COMPAT", "EXECIGNORE", "FUNCNEST",
                             "LC_ALL", "LC_COLLATE", "LC_CTYPE", "LANG"])


#: assigning to these, or unsetting them, fails: bash made them read-only at start-up
READONLY = frozenset(["UID", "EUID", "PPID", "SHELLOPTS", "BASHOPTS", "BASH_VERSINFO"])
#: integer variables: what is assigned to them is evaluated as an arithmetic expression first
INTEGER = frozenset(["OPTIND"])
#: what bash computes each time it is asked: children would see *that* value, whatever was assigned
NOT_MODELLED = frozenset([
    "RANDOM", "SRANDOM", "SECONDS", "EPOCHSECONDS", "EPOCHREALTIME", "BASHPID", "BASH_COMMAND", "BASH_LINENO", "BASH_SOURCE",
    "BASH_ARGC", "BASH_ARGV", "BASH_ARGV0", "BASH_SUBSHELL", "FUNCNAME", "GROUPS", "DIRSTACK", "PIPESTATUS", "HISTCMD", "LINENO",
    "COMP_WORDBREAKS", "BASH_ALIASES", "BASH_CMDS", "BASH_REMATCH"])
#: ... and what it sets once at start-up from things this program does not know: known again once assigned
SET_AT_STARTUP = frozenset(["SHELLOPTS", "BASHOPTS", "BASH_VERSINFO", "BASH", "BASH_VERSION", "HOSTNAME", "HOSTTYPE", "MACHTYPE",
                            "OSTYPE", "PS4", "SHELL", "TERM", "PATH"])


def _guard(name: str) -> None:
    if name in SHELL_BEHAVIOUR:
        raise TritonEnvError("refusing to let `triton env` output change %s: it alters how the rest of the script is read" % name)
    if name in NOT_MODELLED:
        raise TritonEnvError("refusing to let `triton env` output touch %s: bash computes it by itself (its value is not restated here)" % name)


def _lookup(name: str, shell_vars: Dict[str, str], provided: Optional[Callable[[str], Optional[str]]]) -> str:
    """``$NAME``: a shell variable, else one of the few bash provides by itself that have a known value."""
    if name in shell_vars:
        return shell_vars[name]
    if name in NOT_MODELLED or name in SET_AT_STARTUP:
        raise TritonEnvError("refusing to expand $%s in `triton env` output: bash computes it by itself (its value is not restated here)" % name)
    value = provided(name) if provided is not None else None
    return "" if value is None else value


class _Var:
    """``$NAME`` / ``${NAME}`` waiting for its value; ``quoted``: inside double quotes."""

    def __init__(self, name: str, quoted: bool):
        self.name, self.quoted = name, quoted


#: a word is a sequence of (character, quoted?) pairs and parameter references
_Piece = Union[Tuple[str, bool], _Var]


def _parameter(text: str, i: int, quoted: bool) -> Tuple[Optional[_Var], int]:
    """The expansion starting at the ``$`` at ``text[i]``.  Plain parameter
    references are the only expansions that execute nothing, so they are the
    only ones accepted; a ``$`` that starts no expansion is itself."""
    nxt = text[i + 1] if i + 1 < len(text) else ""
    if nxt == "{":
        m = _NAME_AT.match(text, i + 2)
        if m and m.end() < len(text) and text[m.end()] == "}":
            return _Var(m.group(0), quoted), m.end() + 1
        raise TritonEnvError("refusing to evaluate an expansion in `triton env` output: %r" % text[i:i + 20])
    m = _NAME_AT.match(text, i + 1)
    if m and m.group(0) != "_":
        return _Var(m.group(0), quoted), m.end()
    special = nxt != "" and (nxt in "([0123456789?$!#@*-_" or (nxt in "'\"" and not quoted))
    if special:
        # $(...), positional and special parameters, $'...' and $"..."
        raise TritonEnvError("refusing to evaluate an expansion in `triton env` output: %r" % text[i:i + 20])
    return None, i + 1


def _split_commands(text: str) -> Tuple[List[List[List[_Piece]]], List[int]]:
    """Tokenise into commands (newline or ';' separated), each a list of words,
    honouring single quotes, double quotes and backslashes.  Also returns, per
    command, the 0-based line of the text it starts on."""
    cmds: List[List[List[_Piece]]] = []
    cmd_lines: List[int] = []
    words: List[List[_Piece]] = []
    cur: List[_Piece] = []
    have_word = False
    i, n = 0, len(text)

    def end_word():
        nonlocal have_word, cur
        if have_word:
            words.append(cur)
            cur = []
            have_word = False

    def end_cmd():
        nonlocal words
        end_word()
        if words:
            cmds.append(words)
            cmd_lines.append(text.count("\n", 0, i) - sum(1 for w in words for p in w if not isinstance(p, _Var) and p[0] == "\n"))
        words = []

    while i < n:
        c = text[i]
        if c == "\n" or c == ";":
            end_cmd()
            i += 1
        elif c in " \t":
            end_word()
            i += 1
        elif c == "#" and not have_word:
            while i < n and text[i] != "\n":
                i += 1
        elif c == "'":
            j = text.find("'", i + 1)
            if j == -1:
                raise TritonEnvError("unterminated single quote in `triton env` output")
            cur.extend((ch, True) for ch in text[i + 1:j])
            have_word = True
            i = j + 1
        elif c == '"':
            i += 1
            have_word = True
            while True:
                if i >= n:
                    raise TritonEnvError("unterminated double quote in `triton env` output")
                d = text[i]
                if d == '"':
                    i += 1
                    break
                if d == "\\" and i + 1 < n and text[i + 1] in '"\\$`\n':
                    if text[i + 1] != "\n":
                        cur.append((text[i + 1], True))
                    i += 2
                    continue
                if d == "`":
                    raise TritonEnvError("refusing to evaluate an expansion in `triton env` output: %r" % text[i:i + 20])
                if d == "$":
                    var, i = _parameter(text, i, True)
                    cur.append(var if var is not None else ("$", True))
                    continue
                cur.append((d, True))
                i += 1
        elif c == "\\":
            if i + 1 < n:
                if text[i + 1] != "\n":
                    cur.append((text[i + 1], True))
                    have_word = True
                i += 2
            else:
                i += 1
        elif c == "$":
            var, i = _parameter(text, i, False)
            cur.append(var if var is not None else ("$", False))
            have_word = True
        elif c in "`()<>|&":
            raise TritonEnvError("refusing to evaluate shell syntax in `triton env` output: %r" % text[i:i + 20])
        elif c == "{" and words and not isinstance(words[0][0], _Var) and "".join(
                p[0] for p in words[0] if not isinstance(p, _Var)) in ("export", "unset"):
            # arguments of a command undergo brace expansion ({a,b}, {1..3}); a plain assignment does not
            raise TritonEnvError("refusing to evaluate a brace expression in `triton env` output: %r" % text[i:i + 20])
        else:
            cur.append((c, False))
            have_word = True
            i += 1
    end_cmd()
    return cmds, cmd_lines


def _default_home() -> str:
    import os
    try:
        return pwd.getpwuid(os.getuid()).pw_dir
    except KeyError:
        return "/"


def _assignment(word: List[_Piece], shell_vars: Dict[str, str], home: Callable[[], str],
                provided: Optional[Callable[[str], Optional[str]]] = None) -> Optional[Tuple[str, Optional[str]]]:
    """(name, value) when the word is ``NAME=...`` (value None for a bare ``NAME``);
    None when it is not an assignment word.  The value is expanded the way bash
    expands an assignment: parameter references without field splitting or
    pathname expansion, ``~`` at its start and after each unquoted ``:``."""
    k = 0
    while k < len(word) and not isinstance(word[k], _Var) and not word[k][1] and word[k][0] not in "=":
        k += 1
    name = "".join(p[0] for p in word[:k])       # type: ignore[index]
    if not _NAME.match(name):
        return None
    if k == len(word):
        return name, None
    if isinstance(word[k], _Var) or word[k] != ("=", False):
        return None
    out: List[str] = []
    rest = word[k + 1:]
    i = 0
    at_start = True
    while i < len(rest):
        p = rest[i]
        if isinstance(p, _Var):
            out.append(_lookup(p.name, shell_vars, provided))
            at_start = False
            i += 1
            continue
        ch, quoted = p
        if at_start and ch == "~" and not quoted:
            j = i + 1
            while j < len(rest) and not isinstance(rest[j], _Var) and not rest[j][1] and rest[j][0] not in "/:":
                j += 1
            if j < len(rest) and (isinstance(rest[j], _Var) or rest[j][1]):
                pass                                  # a quoted character inside the tilde-prefix: no expansion
            else:
                user = "".join(q[0] for q in rest[i + 1:j])      # type: ignore[index]
                if user == "":
                    out.append(home())
                    i = j
                    at_start = False
                    continue
                if user in ("+", "-"):
                    raise TritonEnvError("refusing to evaluate ~%s in `triton env` output" % user)
                try:
                    out.append(pwd.getpwnam(user).pw_dir)
                    i = j
                    at_start = False
                    continue
                except KeyError:
                    pass                              # no such user: the text stays as it is
        out.append(ch)
        at_start = ch == ":" and not quoted
        i += 1
    return name, "".join(out)


def apply_triton_env(text: str, shell_vars: Dict[str, str], env: Dict[str, str],
                     home: Optional[Callable[[], str]] = None,
                     exists: Optional[Callable[[str], bool]] = None,
                     provided: Optional[Callable[[str], Optional[str]]] = None) -> None:
    """Mutate ``shell_vars`` (all shell variables) and ``env`` (the exported
    ones) as ``eval`` of ``text`` would for the supported dialect.  ``home()``
    is what ``~`` stands for when ``HOME`` is not among the shell variables;
    ``exists(name)`` says whether a command of that name can be found on PATH;
    ``provided(name)`` is the value of a variable bash sets by itself (UID, PPID, IFS ...), if it is known."""
    def tilde_home() -> str:
        if "HOME" in shell_vars:
            return shell_vars["HOME"]
        return (home or _default_home)()

    def plain(word: List[_Piece]) -> str:
        if any(isinstance(p, _Var) for p in word):
            # outside an assignment the value would be split into fields and globbed
            raise TritonEnvError("refusing to expand a parameter outside an assignment in `triton env` output")
        return "".join(p[0] for p in word)        # type: ignore[index]

    pending = getattr(shell_vars, "export_when_set", None)
    if pending is None:
        pending = set()         # a plain dict: the attribute is remembered for the rest of this text at least
    commands, lines = _split_commands(text)
    for words, line in zip(commands, lines):
        head = plain(words[0]) if not any(isinstance(p, _Var) for p in words[0]) else None
        if head == "export":
            if len(words) == 1:
                # a bare `export` lists every exported variable on the script's stdout
                raise TritonEnvError("refusing to run a bare `export` from `triton env` output")
            said: List[str] = []
            for w in words[1:]:
                if not isinstance(w[0], _Var) and w[0] == ("-", False):
                    raise TritonEnvError("refusing to run `export` with options from `triton env` output: %r" % _show(w))
                parsed = _assignment(w, shell_vars, tilde_home, provided)
                if parsed is None:
                    if any(isinstance(p, _Var) for p in w):
                        raise TritonEnvError("not a valid identifier in `triton env` output: %r" % _show(w))
                    said.append("export: `%s': not a valid identifier" % _show(w))
                    continue
                name, value = parsed
                _guard(name)
                if value is not None and name in INTEGER:
                    raise TritonEnvError("refusing to assign %s from `triton env` output: bash would evaluate the value as arithmetic" % name)
                if value is not None and name in READONLY:
                    # under errexit (setup.sh:3) the first of these is the last thing bash says
                    raise BashDiagnostic(said + ["%s: readonly variable" % name], line)
                if value is not None:
                    shell_vars[name] = value
                if name in shell_vars:
                    env[name] = shell_vars[name]
                elif name in SET_AT_STARTUP:
                    raise TritonEnvError("refusing to export %s from `triton env` output: bash computes it by itself (its value is not restated here)" % name)
                elif provided is not None and provided(name) is not None:
                    env[name] = provided(name)      # UID, PPID, IFS ...: bash has a value for it
                else:
                    pending.add(name)           # `export NAME` before NAME has a value: exported once it gets one
            if said:
                if all(m.startswith("export: `") for m in said):
                    raise InvalidIdentifier([m[len("export: `"):-len("': not a valid identifier")] for m in said], line)
                raise BashDiagnostic(said, line)
        elif head == "unset":
            said = []
            for w in words[1:]:
                name = plain(w)
                if name.startswith("-"):
                    raise TritonEnvError("refusing to run `unset` with options from `triton env` output: %r" % name)
                if not _NAME.fullmatch(name):
                    continue                    # bash (not in posix mode) looks for a function of that name, silently
                _guard(name)
                if name in READONLY:
                    said.append("unset: %s: cannot unset: readonly variable" % name)
                    continue
                shell_vars.pop(name, None)
                env.pop(name, None)
                pending.discard(name)
            if said:
                raise BashDiagnostic(said, line)
        else:
            parsed = _assignment(words[0], shell_vars, tilde_home, provided) if len(words) == 1 else None
            if parsed is None or parsed[1] is None:
                # a command.  One that cannot be found runs nothing, and bash's answer to it can be
                # given without running anything; every other command is refused.
                if (exists is not None and head is not None and head != "" and "/" not in head and "=" not in head
                        and head not in _SHELL_WORDS and all(not isinstance(p, _Var) for w in words for p in w)
                        and not exists(head)):
                    raise CommandNotFound(head, line)
                raise TritonEnvError("refusing to run a command from `triton env` output: %r" % " ".join(_show(w) for w in words))
            name, value = parsed
            _guard(name)
            if name in INTEGER:
                raise TritonEnvError("refusing to assign %s from `triton env` output: bash would evaluate the value as arithmetic" % name)
            if name in READONLY:
                raise BashDiagnostic(["%s: readonly variable" % name], line)
            shell_vars[name] = value
            if name in env or name in pending:
                pending.discard(name)
                env[name] = value


def _show(word: List[_Piece]) -> str:
    return "".join("$" + p.name if isinstance(p, _Var) else p[0] for p in word)
