"""The handful of bash behaviours the reference's results depend on, restated.

The reference is a bash script, and a surprising amount of what it writes to
disk is decided not by its own logic but by what bash does to a string on the
way through: ``read`` without ``-r`` eats backslashes (setup.sh:99,101,472),
every ``echo $VAR`` is unquoted and so field-splits and pathname-expands its
value (setup.sh:104-106,130-131,265 and ~30 more sites), ``echo`` itself takes a
leading ``-n``/``-e`` in the *data* as an option, ``[[ a == $b ]]`` treats an
unquoted right-hand side as a pattern (setup.sh:345,388,442) and ``$(...)``
drops trailing newlines.  A replacement that wants byte-identical files has to
do the same things to the same strings, so they are written down here once, as
plain functions on ``str``, and checked against the real ``bash`` in
tests/test_bashsem.py.

Locale: everything here is the C locale (``LC_ALL=C``): bytes compare and sort
by value, ``[a-z]`` is ASCII.  Strings are ``str`` decoded with
``surrogateescape`` so arbitrary bytes survive a round trip.
"""
from __future__ import annotations

import os
import re
from typing import Callable, List, Optional, Sequence, Tuple

from . import shmatch

IFS_WS = " \t\n"


def to_bytes(s: str) -> bytes:
    return s.encode("utf-8", "surrogateescape")


def from_bytes(b: bytes) -> str:
    return b.decode("utf-8", "surrogateescape")


def c_sort_key(s: str) -> bytes:
    """Collation of ``sort``/``ls``/glob results under LC_ALL=C."""
    return to_bytes(s)


# Quoted text inside a word that is otherwise subject to pathname expansion
# (the result of tilde expansion, setup.sh:216-217: bash never globs or splits
# it): its pattern characters are carried as private-use code points so the
# globbing code sees ordinary characters, and mapped back on the way out.
_QUOTED_BASE = 0xE000
_QUOTABLE = "*?[]\\"


def quote_for_glob(text: str) -> str:
    """Mark ``text`` as quoted: none of its characters is a pattern character."""
    return "".join(chr(_QUOTED_BASE + ord(c)) if c in _QUOTABLE else c for c in text)


def _unquote(text: str) -> str:
    return "".join(chr(ord(c) - _QUOTED_BASE) if _QUOTED_BASE <= ord(c) < _QUOTED_BASE + 128 else c for c in text)


# ---------------------------------------------------------------------------
# read  (bash builtin, no -r, one variable, default IFS)
# ---------------------------------------------------------------------------
def read_builtin(next_byte: Callable[[], Optional[int]]) -> Tuple[str, bool]:
    """``read var``: returns (value, ok); ok is False when EOF came before a
    newline (bash's ``read`` then returns non-zero but still assigns).

    ``next_byte`` returns the next byte of stdin or None at EOF; it is called
    once per byte so that nothing beyond the newline is consumed -- children
    that inherit stdin see the rest, as they do under bash.
    """
    chars: List[Tuple[int, bool]] = []  # (byte, escaped)
    ok = True
    pending_backslash = False
    while True:
        b = next_byte()
        if b is None:
            ok = False
            break
        if pending_backslash:
            pending_backslash = False
            if b == 0x0A:
                continue  # backslash-newline: line continuation
            if b != 0:
                chars.append((b, True))
            continue
        if b == 0x5C:
            pending_backslash = True
            continue
        if b == 0x0A:
            break
        if b == 0:
            continue  # read drops NUL bytes
        chars.append((b, False))
    # One variable: leading unescaped IFS whitespace goes.  If what is left is a
    # single field, that field is the value (escaped blanks are part of it).  If
    # there are more fields, bash assigns the whole remainder and then strips
    # trailing blanks *whether escaped or not* (strip_trailing_ifs_whitespace).
    ws = (0x20, 0x09, 0x0A)
    start, end = 0, len(chars)
    while start < end and chars[start][0] in ws and not chars[start][1]:
        start += 1
    k = start
    while k < end and not (chars[k][0] in ws and not chars[k][1]):
        k += 1
    first_end = k
    while k < end and chars[k][0] in ws and not chars[k][1]:
        k += 1
    if k >= end:
        end = first_end
    else:
        while end > start and chars[end - 1][0] in ws:
            end -= 1
    if pending_backslash and start == end:
        # bash quirk: a lone backslash cut off by EOF leaves its internal
        # quoting byte (CTLESC, 0x01) in the variable
        return "\x01", ok
    return from_bytes(bytes(c for c, _ in chars[start:end])), ok


# ---------------------------------------------------------------------------
# field splitting and pathname expansion of an unquoted expansion
# ---------------------------------------------------------------------------
def split_fields(s: str) -> List[str]:
    """Field splitting of an unquoted ``$var`` with the default IFS."""
    return [w for w in re.split(r"[ \t\n]+", s) if w]


def _has_glob_chars(word: str) -> bool:
    """bash's unquoted_glob_pattern_p: does this word need pathname expansion?"""
    open_br = False
    i, n = 0, len(word)
    while i < n:
        c = word[i]
        i += 1
        if c in "?*":
            return True
        if c == "[":
            open_br = True
        elif c == "]":
            if open_br:
                return True
        elif c == "/":
            open_br = False
        elif c == "\\":
            if i < n and word[i] != "/":
                i += 1
            elif i >= n:
                return False
    return False


def pattern_match(s: str, pat: str, utf8: bool = False) -> bool:
    """``[[ $s == $pat ]]`` with an unquoted right-hand side (setup.sh:351,394,443):
    a pattern match, with the extended operators bash switches on inside ``[[ ]]``."""
    return shmatch.strmatch(pat, s, shmatch.EXTMATCH | (shmatch.UTF8 if utf8 else 0))


class Cwd(str):
    """A directory to expand pathnames in, which also knows whether the locale's character set is
    UTF-8 (``?`` then matches a multi-byte character as one)."""
    utf8 = False
    #: how `echo -e` writes \\uHHHH: True (UTF-8), False (ASCII only) or the name of a character set
    wide = False

    def __new__(cls, path: str, utf8: bool = False, wide=None):
        obj = super().__new__(cls, path)
        obj.utf8 = utf8
        obj.wide = utf8 if wide is None else wide
        return obj


def _glob_component(pattern: str, utf8: bool = False) -> Callable[[str], bool]:
    """Matcher for one pathname component (no extended operators: extglob is off in a script)."""
    # quoted text (see quote_for_glob) reaches the matcher as backslash-escaped characters
    pattern = "".join("\\" + _unquote(c) if _unquote(c) != c else c for c in pattern)

    def match(name: str) -> bool:
        if name in (".", ".."):
            return False  # globskipdots (default on in bash 5.2)
        return shmatch.strmatch(pattern, name, shmatch.PERIOD | (shmatch.UTF8 if utf8 else 0))
    return match


def _unescape_glob(s: str) -> str:
    return re.sub(r"\\(.)", r"\1", s, flags=re.S)


def glob_word(word: str, cwd: str) -> List[str]:
    """Pathname expansion of one field; the word itself when nothing matches."""
    if not _has_glob_chars(word):
        return [_unquote(word)]
    absolute = word.startswith("/")
    parts = word.split("/")
    # candidates are (display path, filesystem path)
    cands: List[Tuple[str, str]] = [("/" if absolute else "", "/" if absolute else cwd)]
    comps = parts[1:] if absolute else parts
    for idx, comp in enumerate(comps):
        last = idx == len(comps) - 1
        nxt: List[Tuple[str, str]] = []
        if comp == "":
            # "a//b" or trailing slash: keep the slash, require a directory
            for disp, fs in cands:
                if last:
                    if os.path.isdir(fs):
                        nxt.append((disp + "/" if not disp.endswith("/") else disp, fs))
                else:
                    nxt.append((disp + "/" if disp and not disp.endswith("/") else disp, fs))
            cands = nxt
            continue
        if _has_glob_chars(comp):
            m = _glob_component(comp, getattr(cwd, "utf8", False))
            for disp, fs in cands:
                try:
                    names = os.listdir(fs)
                except OSError:
                    continue
                for name in sorted(names, key=c_sort_key):
                    if m(name):
                        p = os.path.join(fs, name)
                        if last or os.path.isdir(p):
                            nxt.append((_join_disp(disp, name), p))
        else:
            name = _unquote(_unescape_glob(comp))
            for disp, fs in cands:
                p = os.path.join(fs, name)
                if os.path.lexists(p) and (last or os.path.isdir(p)):
                    nxt.append((_join_disp(disp, name), p))
        cands = nxt
        if not cands:
            break
    if not cands:
        return [_unquote(word)]
    return sorted((d for d, _ in cands), key=c_sort_key)


def _join_disp(disp: str, name: str) -> str:
    if disp == "" or disp.endswith("/"):
        return disp + name
    return disp + "/" + name


def expand_unquoted(s: str, cwd: str) -> List[str]:
    """What an unquoted ``$var`` becomes: fields, each pathname-expanded."""
    out: List[str] = []
    for w in split_fields(s):
        out.extend(glob_word(w, cwd))
    return out


def expand_word(segments: Sequence[Tuple[str, bool]], cwd: str) -> List[str]:
    """Expand one shell word made of literal text and unquoted substitutions.

    ``segments`` is a list of (text, is_unquoted_expansion).  Literal text
    (pass it through :func:`quote_for_glob` when the shell treats it as quoted)
    glues to the neighbouring fields the way it does in the shell:
    ``http://$(cat f):8080`` with a two-line ``f`` gives two words
    (setup.sh:66)."""
    words: List[str] = []
    cur: Optional[str] = None
    for text, is_exp in segments:
        if not is_exp:
            cur = (cur or "") + text
            continue
        if text == "":
            continue
        fields = split_fields(text)
        if not fields:
            # only IFS whitespace: it still terminates the current field
            if cur is not None:
                words.append(cur)
                cur = None
            continue
        if text[0] in IFS_WS and cur is not None:
            words.append(cur)
            cur = None
        for k, f in enumerate(fields):
            if k == 0:
                cur = (cur or "") + f
            else:
                words.append(cur)  # type: ignore[arg-type]
                cur = f
        if text[-1] in IFS_WS:
            words.append(cur)  # type: ignore[arg-type]
            cur = None
    if cur is not None:
        words.append(cur)
    out: List[str] = []
    for w in words:
        out.extend(glob_word(w, cwd))
    return out


# ---------------------------------------------------------------------------
# echo  (bash builtin, xpg_echo off)
# ---------------------------------------------------------------------------
_ECHO_OPT = re.compile(r"-[neE]+\Z")


def _encode_wide(value: int) -> str:
    """UTF-8 as bash writes it for any 31-bit value (surrogates and what lies beyond U+10FFFF included)."""
    if value < 0x80:
        return chr(value)
    if value < 0x800:
        raw = [0xC0 | value >> 6]
        rest = 1
    elif value < 0x10000:
        raw = [0xE0 | value >> 12]
        rest = 2
    elif value < 0x200000:
        raw = [0xF0 | value >> 18]
        rest = 3
    elif value < 0x4000000:
        raw = [0xF8 | value >> 24]
        rest = 4
    else:
        raw = [0xFC | value >> 30]
        rest = 5
    raw += [0x80 | (value >> (6 * k)) & 0x3F for k in range(rest - 1, -1, -1)]
    return from_bytes(bytes(raw))


def locale_charset(name: str) -> str:
    """The character set bash reads out of the *name* of a locale it could not set (or that is no multi-byte one):
    what follows the last dot, without an @modifier; the whole name when there is no dot; ASCII for no name."""
    if name == "":
        return "ASCII"
    dot = name.rfind(".")
    if dot == -1:
        return name
    return name[dot + 1:].split("@", 1)[0]


def _to_charset(value: int, charset: str) -> Optional[str]:
    """iconv from UTF-8 to ``charset``; None when iconv knows no such character set or it has no such character.
    (Python's codec tables stand in for glibc's: the same for UTF-8, the ISO-8859 family, KOI8, the Windows code
    pages; corner cases of the East Asian multi-byte sets may differ.)"""
    import codecs
    if 0xD800 <= value <= 0xDFFF:
        return None                                   # iconv rejects the UTF-8 of a surrogate
    try:
        codec = codecs.lookup(charset)
    except LookupError:
        return None
    if codec.name == "utf-8":
        return _encode_wide(value)
    if value > 0x10FFFF:
        return None
    try:
        return from_bytes(chr(value).encode(codec.name))
    except (UnicodeEncodeError, LookupError):
        return None


def _echo_escapes(s: str, utf8=False) -> Tuple[str, bool]:
    """Backslash processing of ``echo -e``; returns (text, stop) where stop is
    True after ``\\c``.  ``\\uHHHH`` / ``\\UHHHHHHHH`` come out in the locale's
    character set: UTF-8 when it is that (``utf8``); in the C locale only ASCII
    can, anything else is written back as the escape it was."""
    out: List[str] = []
    i, n = 0, len(s)
    simple = {"a": "\a", "b": "\b", "e": "\x1b", "E": "\x1b", "f": "\f", "n": "\n",
              "r": "\r", "t": "\t", "v": "\v", "\\": "\\"}
    while i < n:
        c = s[i]
        i += 1
        if c != "\\" or i >= n:
            out.append(c)
            continue
        d = s[i]
        i += 1
        if d in simple:
            out.append(simple[d])
        elif d == "c":
            return "".join(out), True
        elif d == "0":
            # echo -e takes octal only in the \0nnn form (up to three digits)
            digits = 3
            val = 0
            while digits and i < n and s[i] in "01234567":
                val = val * 8 + int(s[i])
                i += 1
                digits -= 1
            out.append(from_bytes(bytes([val & 0xFF])))
        elif d == "x":
            j = i
            while j < n and j - i < 2 and s[j] in "0123456789abcdefABCDEF":
                j += 1
            if j == i:
                out.append("\\x")
            else:
                out.append(from_bytes(bytes([int(s[i:j], 16) & 0xFF])))
                i = j
        elif d in "uU":
            width = 4 if d == "u" else 8
            j = i
            while j < n and j - i < width and s[j] in "0123456789abcdefABCDEF":
                j += 1
            if j == i:
                out.append("\\" + d)
            else:
                value = int(s[i:j], 16)
                if value < 0x80:
                    out.append(chr(value))
                elif value > 0x7FFFFFFF:
                    pass                                   # not a code point bash can convert: nothing is written
                elif utf8 is True:
                    out.append(_encode_wide(value))
                else:
                    converted = _to_charset(value, utf8) if isinstance(utf8, str) else None
                    out.append(converted if converted is not None else ("\\u%04X" if value <= 0xFFFF else "\\U%08X") % value)
                i = j
        else:
            out.append("\\" + d)
    return "".join(out), False


def echo_builtin(args: Sequence[str], utf8=False) -> str:
    """Exactly what ``echo ARGS...`` writes (including the trailing newline,
    or not)."""
    newline, escapes = True, False
    k = 0
    while k < len(args) and _ECHO_OPT.match(args[k]):
        for ch in args[k][1:]:
            if ch == "n":
                newline = False
            elif ch == "e":
                escapes = True
            else:
                escapes = False
        k += 1
    text = " ".join(args[k:])
    if escapes:
        text, stop = _echo_escapes(text, utf8)
        if stop:
            return text
    return text + ("\n" if newline else "")


def cmdsubst(output: str) -> str:
    """``$(...)``: trailing newlines dropped, NUL bytes dropped."""
    return output.replace("\x00", "").rstrip("\n")


def echo_unquoted(value: str, cwd: str) -> str:
    """``echo $value`` -- the workhorse of the reference (unquoted on purpose
    there or not, it is what decides the bytes)."""
    return echo_builtin(expand_unquoted(value, cwd), getattr(cwd, "wide", False))


# --------------------------------------------------------------------------
# how `export -p` / `declare -p` print a value (bash: show_var_attributes ->
# ansic_shouldquote / ansic_quote / sh_double_quote)
# --------------------------------------------------------------------------
_ANSIC_NAMES = {0x1B: "\\E", 7: "\\a", 8: "\\b", 12: "\\f", 10: "\\n", 13: "\\r", 9: "\\t", 11: "\\v", 92: "\\\\", 39: "\\'"}


def _wide_printable(c: str) -> bool:
    import unicodedata
    return unicodedata.category(c) not in ("Cc", "Cs", "Cn", "Zl", "Zp")


def declare_value(value: str, utf8: bool = False) -> str:
    """``"value"`` with ``" $ ` \\`` escaped -- or, when the value holds anything the locale cannot print,
    ``$'...'`` with C escapes and three-digit octal for the bytes of what is not printable."""
    def printable(c: str) -> bool:
        if ord(c) < 0x80:
            return 0x20 <= ord(c) < 0x7F
        return utf8 and not "\udc80" <= c <= "\udcff" and _wide_printable(c)
    if all(printable(c) for c in value):
        return '"' + "".join("\\" + c if c in '"$`\\' else c for c in value) + '"'
    out = []
    for c in value:
        if ord(c) in _ANSIC_NAMES:
            out.append(_ANSIC_NAMES[ord(c)])
        elif printable(c):
            out.append(c)
        else:
            out.append("".join("\\%03o" % b for b in to_bytes(c)))
    return "$'" + "".join(out) + "'"
