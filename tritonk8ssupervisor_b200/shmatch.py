"""bash's pattern matcher, restated (C locale).

``[[ $answer == $default ]]`` (setup.sh:351,394,443) has an unquoted right-hand
side, so the *default* -- a UUID list resolved from whatever ``triton`` printed
-- is a pattern, and inside ``[[ ]]`` bash matches it with the extended
operators ``?(..) *(..) +(..) @(..) !(..)`` switched on.  Pathname expansion of
unquoted words uses the same matcher without them.  The corner cases decide
real outcomes: a ``[`` without its ``]`` is an ordinary character -- unless the
scan for the ``]`` runs into a dangling range (``[a-``) or a trailing backslash,
in which case nothing matches at all and the prompt loops for ever.

This follows the structure of bash 5.2's matcher (one routine for the pattern,
one for a bracket expression, one for an extended group, one to find the end of
a group) and is held to the real ``bash`` by tests/test_shmatch.py.
"""
from __future__ import annotations

from functools import lru_cache
from typing import Optional

PATHNAME = 1        # a slash is matched only by a slash
PERIOD = 2          # a leading period is matched only by a period
EXTMATCH = 4        # ?(..) *(..) +(..) @(..) !(..)
UTF8 = 8            # the locale's character set is UTF-8: a multi-byte character is one character

_CLASSES = {
    "alnum": lambda c: c.isascii() and c.isalnum(), "alpha": lambda c: c.isascii() and c.isalpha(),
    "ascii": lambda c: ord(c) < 128, "blank": lambda c: c in " \t", "cntrl": lambda c: ord(c) < 32 or ord(c) == 127,
    "digit": lambda c: c in "0123456789", "graph": lambda c: 33 <= ord(c) <= 126, "lower": lambda c: "a" <= c <= "z",
    "print": lambda c: 32 <= ord(c) <= 126, "punct": lambda c: 33 <= ord(c) <= 126 and not c.isalnum(),
    "space": lambda c: c in " \t\n\r\f\v", "upper": lambda c: "A" <= c <= "Z",
    "word": lambda c: c == "_" or (c.isascii() and c.isalnum()), "xdigit": lambda c: c in "0123456789abcdefABCDEF",
}
#: names POSIX gives to single characters inside [. .]
_COLLSYMS = {"NUL": "\0", "tab": "\t", "newline": "\n", "space": " ", "hyphen": "-", "hyphen-minus": "-", "period": ".",
             "full-stop": ".", "slash": "/", "solidus": "/", "backslash": "\\", "reverse-solidus": "\\", "underscore": "_",
             "low-line": "_", "left-square-bracket": "[", "right-square-bracket": "]", "asterisk": "*", "question-mark": "?",
             "exclamation-mark": "!", "circumflex": "^", "circumflex-accent": "^", "tilde": "~", "vertical-line": "|",
             "left-parenthesis": "(", "right-parenthesis": ")", "plus-sign": "+", "commercial-at": "@", "equals-sign": "=",
             "comma": ",", "colon": ":", "semicolon": ";", "dollar-sign": "$", "number-sign": "#", "percent-sign": "%",
             "ampersand": "&", "apostrophe": "'", "quotation-mark": '"', "less-than-sign": "<", "greater-than-sign": ">",
             "grave-accent": "`", "left-brace": "{", "left-curly-bracket": "{", "right-brace": "}", "right-curly-bracket": "}",
             "zero": "0", "one": "1", "two": "2", "three": "3", "four": "4", "five": "5", "six": "6", "seven": "7", "eight": "8",
             "nine": "9"}


def _wide_class(name: str, c: str) -> bool:
    """Character classes beyond ASCII, as a UTF-8 locale has them (a byte that is no character belongs to none)."""
    if "\udc80" <= c <= "\udcff":
        return False
    import unicodedata
    cat = unicodedata.category(c)
    return {"alnum": c.isalnum(), "alpha": c.isalpha(), "ascii": False, "blank": cat == "Zs", "cntrl": cat == "Cc",
            "digit": False, "graph": cat[0] not in "CZ", "lower": c.islower(), "print": cat[0] != "C" and cat not in ("Zl", "Zp"),
            "punct": cat[0] in "PS", "space": c.isspace(), "upper": c.isupper(), "word": c.isalnum(), "xdigit": False}[name]


def _at(text: str, i: int) -> str:
    """The character at ``i``, or NUL past the end (the C string's terminator)."""
    return text[i] if 0 <= i < len(text) else "\0"


def _collsym(pat: str, p: int) -> tuple:
    """``pat[p]`` is the ``.`` of ``[.``; returns (character or None when invalid, index after ``.]``)."""
    end = pat.find(".]", p + 1)
    if end == -1:
        return None, len(pat)
    name = pat[p + 1:end]
    if len(name) == 1:
        return name, end + 2
    return _COLLSYMS.get(name), end + 2


def _bracket(pat: str, p: int, test: str, flags: int) -> Optional[int]:
    """Match ``test`` against the bracket expression whose ``[`` is at ``pat[p-1]``.
    Returns the index in ``pat`` to continue from, or None for "no match".  A
    ``[`` that never finds its ``]`` matches a literal ``[`` and the pattern
    continues right behind it."""
    savep = p
    literal = savep if test == "[" else None
    negate = _at(pat, p) in "!^" and p < len(pat)
    if negate:
        p += 1
    c = _at(pat, p)
    p += 1
    matched = False
    while True:
        cstart = cend = c
        if c == "[" and _at(pat, p) == "=" and _at(pat, p + 2) == "=" and _at(pat, p + 3) == "]":
            # equivalence class: in the C locale a character is equivalent to itself only
            pc = _at(pat, p + 1)
            p += 4
            if pc == test:
                p += 1              # the skip below starts by stepping back one
                matched = True
                break
            c = _at(pat, p)
            p += 1
            if c == "\0":
                return literal
            if c == "]":
                break
            continue
        if c == "[" and _at(pat, p) == ":":
            close = pat.find(":]", p + 1)
            if close != -1:
                name = pat[p + 1:close]
                ok = name in _CLASSES and (_wide_class(name, test) if (flags & UTF8) and ord(test) > 127 else _CLASSES[name](test))
                p = close + 2
                if ok:
                    p += 1
                    matched = True
                    break
                # this cannot be the first point of a range
                c = _at(pat, p)
                p += 1
                if c == "\0":
                    return literal
                if c == "]":
                    break
                continue
            # no closing ":]": the '[' is an ordinary member
        invalid_start = False
        if c == "[" and _at(pat, p) == ".":
            sym, p = _collsym(pat, p)
            if sym is None:
                invalid_start = True
            else:
                cstart = cend = sym
        if c == "\\":
            if p >= len(pat):
                return None
            cstart = cend = pat[p]
            p += 1
        isrange = False
        if c == "\0":
            return literal
        c = _at(pat, p)
        p += 1
        if c == "\0":
            return literal
        if (flags & PATHNAME) and c == "/":
            return None
        if c == "-" and _at(pat, p) != "]":
            cend = _at(pat, p)
            p += 1
            if cend == "\\":
                cend = _at(pat, p)
                p += 1
            if cend == "\0":
                return None
            invalid_end = False
            if cend == "[" and _at(pat, p) == ".":
                sym, p = _collsym(pat, p)
                if sym is None:
                    invalid_end = True
                else:
                    cend = sym
            c = _at(pat, p)
            p += 1
            if invalid_start or invalid_end or cstart > cend:
                if c == "]":
                    break
                continue
            isrange = True
        if not invalid_start:
            if not isrange and test == cstart:
                matched = True
                break
            if isrange and cstart <= test <= cend:
                matched = True
                break
        if c == "]":
            break
    if not matched:
        return p if negate else None
    # skip the rest of the expression that has already matched
    p -= 1
    c = _at(pat, p)
    depth = 1
    while depth > 0:
        if c == "\0":
            return literal
        c = _at(pat, p)
        p += 1
        if c == "[" and _at(pat, p) in "=:.":
            depth += 1
        elif c == "]":
            depth -= 1
        elif c == "\\":
            if p < len(pat):
                p += 1
            else:
                return None
    return None if negate else p


def _patscan(pat: str, start: int, end: int, delim: str) -> Optional[int]:
    """Index just past the ``)`` closing the group scanned from ``start`` (or past
    a top-level ``|`` when ``delim`` is ``|``); None when the group never closes."""
    pnest = bnest = 0
    cchar = ""
    bfirst = -1
    skip = False
    if start >= end:
        return None
    s = start
    while s < len(pat):
        c = pat[s]
        if s >= end:
            return s
        if skip:
            skip = False
        elif c == "\\":
            skip = True
        elif c == "[":
            if bnest == 0:
                bfirst = s + 1
                if _at(pat, bfirst) in "!^" and bfirst < len(pat):
                    bfirst += 1
                bnest += 1
            elif _at(pat, s + 1) in ":.=" and s + 1 < len(pat):
                cchar = pat[s + 1]
        elif c == "]":
            if bnest:
                if cchar and pat[s - 1] == cchar:
                    cchar = ""
                elif s != bfirst:
                    bnest -= 1
                    bfirst = -1
        elif c == "(":
            if bnest == 0:
                pnest += 1
        elif c == ")":
            if bnest == 0:
                if pnest <= 0:
                    return s + 1
                pnest -= 1
        elif c == "|":
            if bnest == 0 and pnest == 0 and delim == "|":
                return s + 1
        s += 1
    return None


class _Matcher:
    def __init__(self, pat: str, text: str):
        self.pat, self.text = pat, text
        self.gmatch = lru_cache(maxsize=None)(self._gmatch)

    def _leading_period(self, n: int, start: int, flags: int) -> bool:
        return bool(flags & PERIOD) and _at(self.text, n) == "." and (
            n == start or (bool(flags & PATHNAME) and n > 0 and self.text[n - 1] == "/"))

    def _gmatch(self, n: int, se: int, p: int, pe: int, flags: int) -> bool:
        """Does text[n:se] match pat[p:pe]?"""
        pat, text = self.pat, self.text
        start = n
        while p < pe:
            c = pat[p]
            p += 1
            sc = text[n] if n < se else "\0"
            if (flags & EXTMATCH) and p < pe and pat[p] == "(" and c in "+*?@!":
                lflags = flags if n == start else flags & ~PERIOD
                return self._extmatch(c, n, se, p, pe, lflags)
            if c == "?":
                if n >= se:
                    return False
                if (flags & PATHNAME) and sc == "/":
                    return False
                if self._leading_period(n, start, flags):
                    return False
            elif c == "\\":
                if p == pe and sc == "\\" and n + 1 == se:
                    pass                        # a trailing backslash matches a final backslash
                else:
                    if p == pe:
                        return False
                    c = pat[p]
                    p += 1
                    if n >= se or sc != c:
                        return False
            elif c == "*":
                if self._leading_period(n, start, flags):
                    return False
                return self._star(n, se, p, pe, flags)
            elif c == "[":
                if n >= se:
                    return False
                if self._leading_period(n, start, flags):
                    return False
                # the expression ends with the (sub)pattern: what follows pe belongs to somebody else
                nxt = _bracket(pat[:pe], p, sc, flags)
                if nxt is None:
                    return False
                p = nxt
            else:
                if n >= se or c != sc:
                    return False
            n += 1
        return n == se

    def _star(self, n: int, se: int, p: int, pe: int, flags: int) -> bool:
        """Everything after a ``*`` (``p`` is just behind it)."""
        pat, text = self.pat, self.text

        def tail_ok() -> bool:
            return not (flags & PATHNAME) or "/" not in text[n:se]
        if p == pe:
            return True
        # a run of * and ? collapses; every ? still takes one character
        c = pat[p]
        p += 1
        while c in "?*":
            sc = text[n] if n < se else "\0"
            if (flags & PATHNAME) and sc == "/":
                return False
            if (flags & EXTMATCH) and c == "?" and _at(pat[:pe], p) == "(":
                if self._extmatch(c, n, se, p, pe, flags):
                    return True
                newp = _patscan(pat, p + 1, pe, "")       # zero occurrences: skip the group
                p = newp if newp is not None else pe
            elif c == "?":
                if n >= se:
                    return False
                n += 1
            if (flags & EXTMATCH) and c == "*" and _at(pat[:pe], p) == "(":
                for newn in range(n, se):
                    if self._extmatch(c, newn, se, p, pe, flags):
                        return True
                newp = _patscan(pat, p + 1, pe, "")
                p = newp if newp is not None else pe
            if p == pe:
                return True                                # the wildcards were the end of the pattern
            c = pat[p]
            p += 1
        if n == se and (flags & EXTMATCH) and c in "!?" and _at(pat[:pe], p) == "(":
            hit = self._extmatch(c, n, se, p - 1, pe, flags)
            return (not hit) if c == "!" else hit
        p -= 1                                             # back onto the character that ended the run
        endp = se
        if flags & PATHNAME:
            slash = text.find("/", n, se)
            if slash != -1:
                endp = slash
        # bash only tries positions whose character can start the rest of the pattern; for a rest
        # that is a lone trailing backslash that character is the terminator, which nothing equals
        first = None
        is_group = p + 1 < pe and pat[p + 1] == "(" and c in "?*+@!"
        if c != "[" and not ((flags & EXTMATCH) and (is_group or c in "?*+@!")):
            first = _at(pat[:pe], p + 1) if c == "\\" else c
        return any((first is None or text[k] == first) and self.gmatch(k, se, p, pe, flags & ~PERIOD) for k in range(n, endp))

    def _extmatch(self, xc: str, s: int, se: int, p: int, pe: int, flags: int) -> bool:
        """``xc`` is one of ``+*?@!`` and ``pat[p]`` the ``(`` behind it."""
        pat, text = self.pat, self.text
        prest = _patscan(pat, p + 1, pe, "")
        if prest is None:
            # not a well-formed group: the pattern text is compared as a string
            return pat[p - 1:pe] == text[s:se]
        nosubs = flags & ~PERIOD

        def alternatives():
            psub = p + 1
            while True:
                pnext = _patscan(pat, psub, pe, "|")
                if pnext is None:
                    pnext = prest
                yield psub, pnext - 1
                if pnext == prest:
                    return
                psub = pnext

        if xc in "+*":
            if xc == "*" and self.gmatch(s, se, prest, pe, flags):
                return True
            for a, b in alternatives():
                for srest in range(s, se + 1):
                    if self.gmatch(s, srest, a, b, flags):
                        xflags = nosubs if srest > s else flags
                        if self.gmatch(srest, se, prest, pe, xflags) or (
                                s != srest and self.gmatch(srest, se, p - 1, pe, xflags)):
                            return True
            return False
        if xc in "?@":
            if xc == "?" and self.gmatch(s, se, prest, pe, flags):
                return True
            for a, b in alternatives():
                first = se if prest == pe else s
                for srest in range(first, se + 1):
                    xflags = nosubs if srest > s else flags
                    if self.gmatch(s, srest, a, b, flags) and self.gmatch(srest, se, prest, pe, xflags):
                        return True
            return False
        # '!': anything except one of the alternatives
        for srest in range(s, se + 1):
            hit = any(self.gmatch(s, srest, a, b, flags) for a, b in alternatives())
            if not hit and (flags & PERIOD) and _at(text, s) == ".":
                return False
            xflags = nosubs if srest > s else flags
            if not hit and self.gmatch(srest, se, prest, pe, xflags):
                return True
        return False


def _bytes(text: str) -> str:
    """One character per byte (strings here are UTF-8 decoded with surrogateescape)."""
    return text.encode("utf-8", "surrogateescape").decode("latin-1")


def strmatch(pattern: str, text: str, flags: int = 0) -> bool:
    """Does ``text`` match ``pattern``?  (bash ``strmatch``; C locale, i.e. byte by
    byte: a two-byte character is matched by ``??``, not by ``?``; with the ``UTF8`` flag,
    character by character, a byte that is no character counting as one)"""
    if (flags & UTF8) and any("\udc80" <= c <= "\udcff" for c in pattern + text):
        flags &= ~UTF8          # bash cannot convert either to wide characters and falls back on bytes
    if not flags & UTF8:
        pattern, text = _bytes(pattern), _bytes(text)
    m = _Matcher(pattern, text)
    return m.gmatch(0, len(text), 0, len(pattern), flags)
