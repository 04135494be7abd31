"""Apply the output of ``triton env`` to the shell state.

The reference does ``eval "$(triton env)"`` (setup.sh:210), i.e. it executes
whatever the CLI prints.  What the CLI prints is a fixed, tiny dialect --
``export NAME="value"``, ``unset NAME`` and ``#`` comments -- and that dialect
is what is accepted here; anything else is refused loudly instead of executed.
That is the one deliberate narrowing of the boundary ("Deviations").
"""
from __future__ import annotations

import re
from typing import Dict, List

_NAME = re.compile(r"[A-Za-z_][A-Za-z0-9_]*\Z")


class TritonEnvError(ValueError):
    pass


def _split_commands(text: str) -> List[List[str]]:
    """Tokenise into commands (newline or ';' separated), each a list of words,
    honouring single quotes, double quotes and backslashes."""
    cmds: List[List[str]] = []
    words: List[str] = []
    cur: List[str] = []
    have_word = False
    i, n = 0, len(text)

    def end_word():
        nonlocal have_word
        if have_word:
            words.append("".join(cur))
            cur.clear()
            have_word = False

    def end_cmd():
        nonlocal words
        end_word()
        if words:
            cmds.append(words)
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
            cur.append(text[i + 1:j])
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
                        cur.append(text[i + 1])
                    i += 2
                    continue
                if d in "$`":
                    raise TritonEnvError("refusing to evaluate an expansion in `triton env` output: %r" % text[i:i + 20])
                cur.append(d)
                i += 1
        elif c == "\\":
            if i + 1 < n:
                if text[i + 1] != "\n":
                    cur.append(text[i + 1])
                    have_word = True
                i += 2
            else:
                i += 1
        elif c in "$`()<>|&":
            raise TritonEnvError("refusing to evaluate shell syntax in `triton env` output: %r" % text[i:i + 20])
        else:
            cur.append(c)
            have_word = True
            i += 1
    end_cmd()
    return cmds


def apply_triton_env(text: str, shell_vars: Dict[str, str], env: Dict[str, str]) -> None:
    """Mutate ``shell_vars`` (all shell variables) and ``env`` (the exported
    ones) as ``eval`` of ``text`` would for the supported dialect."""
    for words in _split_commands(text):
        head = words[0]
        if head == "export":
            for w in words[1:]:
                name, eq, value = w.partition("=")
                if not _NAME.match(name):
                    raise TritonEnvError("not a valid identifier in `triton env` output: %r" % w)
                if eq:
                    shell_vars[name] = value
                if name in shell_vars:
                    env[name] = shell_vars[name]
        elif head == "unset":
            for w in words[1:]:
                shell_vars.pop(w, None)
                env.pop(w, None)
        elif "=" in head and _NAME.match(head.split("=", 1)[0]) and len(words) == 1:
            name, _, value = head.partition("=")
            shell_vars[name] = value
            if name in env:
                env[name] = value
        else:
            raise TritonEnvError("refusing to run a command from `triton env` output: %r" % " ".join(words))
