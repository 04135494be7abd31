"""The one sed program of the path whose *text* contains data.

``createAnsibleConfigs`` and ``cleanRunner`` rewrite the key path in
``ansible.cfg`` with

    sed "s;private_key_file = .*$;private_key_file = VALUE;g"      (setup.sh:133)
    sed "s~private_key_file = .*$~private_key_file = ~g"           (setup.sh:511)

VALUE is ``~/.ssh/<name of the operator's key file>`` spliced into the program,
so sed's own syntax applies to it: ``&`` is the matched text, ``\\n`` a newline,
and an unescaped ``;`` *ends the replacement* -- what follows is parsed as flags
of ``s`` and then as further sed commands, with the script's own trailing ``;g``
becoming the command ``g`` ("replace the pattern space with the empty hold
space": every line of the file comes out empty).  Far-fetched as a key file
name, but it is what the reference does, so it is restated here (GNU sed 4.8
rules, ``sed/compile.c:mark_subst_opts``; third-party, checked against the real
``sed`` in tests/test_sedprog.py).

Restated: the replacement syntax, every flag of ``s`` but one, and the commands
``g G h H x p d q`` after it.  Refused (the caller stops with status 2) rather
than guessed at: the ``e`` flag, which would execute the pattern space as a shell
command, and a value that goes on with any other sed command, a block or an
address -- all legal sed, all absurd in the name of a key file.
"""
from __future__ import annotations

import re
from typing import Dict, List, Optional, Tuple

REGEX = r"private_key_file = .*$"
PREFIX = "private_key_file = "
#: 1-based position in the program text of the first character of VALUE
VALUE_AT = len("s;" + "private_key_file = .*$" + ";" + PREFIX) + 1


class SedResult:
    def __init__(self, ok: bool, output: str = "", error: str = "", wfiles: Optional[Dict[str, str]] = None,
                 refused: str = ""):
        self.ok, self.output, self.error, self.wfiles = ok, output, error, wfiles or {}
        #: non-empty: the value continues the program with something sed would accept and this module
        #: does not restate (or would not run: the e flag) -- the caller must stop, not guess
        self.refused = refused


def _refuse(what: str) -> SedResult:
    return SedResult(False, refused=what)


#: what may legitimately follow the s command in GNU sed and is not restated here: other commands,
#: blocks, and anything that starts an address
_NOT_RESTATED = set("abcDeFilLnNPQrRstTvwWyz=:{!$/\\0123456789~")
_SIMPLE_ESCAPES = {"a": "\a", "f": "\f", "r": "\r", "t": "\t", "v": "\v"}


def _gnu_escape(text: str, at: int) -> Optional[Tuple[str, int]]:
    """GNU sed's character escapes in a replacement: ``\\a \\f \\r \\t \\v``, ``\\cX``, ``\\dNNN``,
    ``\\oNNN``, ``\\xHH``.  ``text[at]`` is the character after the backslash; returns
    (character, index after the escape) or None when it is not one of these."""
    d = text[at]
    if d in _SIMPLE_ESCAPES:
        return _SIMPLE_ESCAPES[d], at + 1
    if d == "c" and at + 1 < len(text):
        return chr(ord(text[at + 1].upper()) ^ 0x40), at + 2
    for letter, digits, base, width in (("d", "0123456789", 10, 3), ("o", "01234567", 8, 3), ("x", "0123456789abcdefABCDEF", 16, 2)):
        if d == letter:
            j = at + 1
            while j < len(text) and j - at - 1 < width and text[j] in digits:
                j += 1
            if j == at + 1:
                return None
            return bytes([int(text[at + 1:j], base) & 0xFF]).decode("utf-8", "surrogateescape"), j
    return None



def _fail(pos: int, what: str) -> SedResult:
    return SedResult(False, error="sed: -e expression #1, char %d: %s\n" % (pos, what))


def run(text: str, value: str, delim: str, utf8: bool = False) -> SedResult:
    """Run ``s<d>private_key_file = .*$<d>private_key_file = VALUE<d>g`` over ``text``."""
    program_tail = value + delim + "g"
    # --- replacement: up to the first unescaped delimiter -----------------------
    parts: List[Tuple[str, str]] = []
    i, n = 0, len(program_tail)
    closed = False
    bad_reference: Optional[SedResult] = None
    while i < n:
        c = program_tail[i]
        if c == "\\" and i + 1 < n:
            d = program_tail[i + 1]
            if d in "123456789":
                # compile.c reports it once the whole command has been read -- if it can be read
                if bad_reference is None:
                    bad_reference = _fail(VALUE_AT + _end_of_s(program_tail, i, delim),
                                          "invalid reference \\%s on `s' command's RHS" % d)
                i += 2
                continue
            if d in "LUluE":
                return _refuse("a case conversion (\\%s) in the replacement" % d)
            esc = _gnu_escape(program_tail, i + 1)
            if esc is not None:
                parts.append(("lit", esc[0]))
                i = esc[1]
                continue
            parts.append(("lit", "\n") if d == "n" else ("match", "") if d == "0" else ("lit", d))
            i += 2
            continue
        if c == delim:
            closed = True
            i += 1
            break
        if c == "\n":
            return _fail(VALUE_AT + i - 1, "unterminated `s' command")
        parts.append(("match", "") if c == "&" else ("lit", c))
        i += 1
    if not closed:
        return _fail(VALUE_AT + n - 1, "unterminated `s' command")
    if bad_reference is not None:
        return bad_reference

    # --- flags of s -------------------------------------------------------------
    flag_g = flag_p = icase = False
    number = 0
    wfile: Optional[str] = None
    while i < n:
        c = program_tail[i]
        pos = VALUE_AT + i
        if c == "g":
            if flag_g:
                return _fail(pos, "multiple `g' options to `s' command")
            flag_g = True
        elif c == "p":
            if flag_p:
                return _fail(pos, "multiple `p' options to `s' command")
            flag_p = True
        elif c in "iI":
            icase = True
        elif c in "mM":
            pass
        elif c.isdigit():
            j = i
            while j < n and program_tail[j].isdigit():
                j += 1
            if number:
                return _fail(VALUE_AT + j - 1, "multiple number options to `s' command")
            number = int(program_tail[i:j])
            if number == 0:
                return _fail(VALUE_AT + j - 1, "number option to `s' command may not be zero")
            i = j
            continue
        elif c == "w":
            rest = program_tail[i + 1:]
            wfile = rest.split("\n", 1)[0].lstrip(" \t")
            if not wfile:
                return _fail(VALUE_AT + n - 1, "missing filename in r/R/w/W commands")
            # the file name runs to the end of the line; a newline in the value lets the program go on after it
            i = i + 1 + rest.index("\n") if "\n" in rest else n
            break
        elif c == "e":
            return _refuse("the e flag of sed's s command (it would run the pattern space as a shell command)")
        elif c in " \t":
            pass
        elif c in ";\n#}":
            break
        else:
            return _fail(pos, "unknown option to `s'")
        i += 1

    # --- further commands ---------------------------------------------------------
    commands: List[str] = []
    while i < n:
        c = program_tail[i]
        if c in "; \t\n":
            i += 1
        elif c == "#":
            while i < n and program_tail[i] != "\n":
                i += 1
        elif c == "}":
            return _fail(VALUE_AT + i, "unexpected `}'")
        elif c in "gpdhxGHq":
            commands.append(c)
            i += 1
            if c == "q":                      # q takes an optional exit status
                j = i
                while j < n and program_tail[j] in " \t":
                    j += 1
                k = j
                while k < n and program_tail[k].isdigit():
                    k += 1
                if k > j:
                    if int(program_tail[j:k]) != 0:
                        return _refuse("a q command with a non-zero exit status")
                    i = k
            while i < n and program_tail[i] in " \t":
                i += 1
            if i < n and program_tail[i] not in ";\n#}":
                return _fail(VALUE_AT + i, "extra characters after command")
        elif c in _NOT_RESTATED:
            return _refuse("the sed program continues with %r, which is not restated here" % program_tail[i:i + 12])
        else:
            return _fail(VALUE_AT + i, "unknown command: `%s'" % c)

    # --- execute ----------------------------------------------------------------------
    # in a UTF-8 locale `.` does not match a byte that is no character
    rx = re.compile(REGEX.replace(".", "[^\udc80-\udcff\n]") if utf8 else REGEX, re.I if icase else 0)

    def replacement(m: "re.Match[str]") -> str:
        return PREFIX + "".join(m.group(0) if k == "match" else t for k, t in parts)

    lines = text.split("\n")
    final_newline = lines[-1] == ""
    if final_newline:
        lines.pop()
    if text == "":
        lines = []
    out = _Stream()
    wout = _Stream()
    hold, hold_nl = "", True
    for idx, line in enumerate(lines):
        # GNU sed remembers per buffer whether its text came with a newline ("chomped")
        line_nl = final_newline or idx < len(lines) - 1
        made = False
        if number <= 1 or flag_g:
            new, k = rx.subn(replacement, line, count=1)
            made = k > 0 and (number <= 1)
            if made:
                line = new
        if made and flag_p:
            out.write(line, line_nl)
        if made and wfile is not None:
            wout.write(line, line_nl)
        deleted = quit_now = False
        for cmd in commands:
            if cmd == "g":
                line, line_nl = hold, hold_nl
            elif cmd == "G":
                line = line + "\n" + hold
            elif cmd == "h":
                hold, hold_nl = line, line_nl
            elif cmd == "H":
                hold = hold + "\n" + line
            elif cmd == "x":
                line, hold, line_nl, hold_nl = hold, line, hold_nl, line_nl
            elif cmd == "p":
                out.write(line, line_nl)
            elif cmd == "d":
                deleted = True
                break
            elif cmd == "q":
                quit_now = True
                break
        if not deleted:
            out.write(line, line_nl)
        if quit_now:
            if out.missing:             # quitting flushes the pending newline, a normal end of input does not
                out.write("", False)
                out.missing = False
            break
    result = SedResult(True, output=out.text())
    if wfile is not None:
        result.wfiles[wfile] = wout.text()
    return result


class _Stream:
    """An output of sed: a record written without its newline gets it as soon as
    anything else is written to the same stream (``output_missing_newline``)."""

    def __init__(self) -> None:
        self.parts: List[str] = []
        self.missing = False

    def write(self, text: str, newline: bool) -> None:
        if self.missing:
            self.parts.append("\n")
        self.parts.append(text + ("\n" if newline else ""))
        self.missing = not newline

    def text(self) -> str:
        return "".join(self.parts)


def _end_of_s(tail: str, start: int, delim: str) -> int:
    """Offset (into VALUE-relative text) of the last character of the s command."""
    i, n = start, len(tail)
    while i < n:
        if tail[i] == "\\":
            i += 2
            continue
        if tail[i] == delim:
            break
        i += 1
    # flags up to the next separator
    j = i + 1
    while j < n and tail[j] not in ";\n}#":
        j += 1
    if j < n and tail[j] in ";\n":
        return j                      # the separator has been read by then
    return min(j, n) - 1
