"""Bash arithmetic evaluation, restated.

Three places on the path hand *data* to bash's arithmetic evaluator:

* ``if ! ((`date +%s` % 15))`` (setup.sh:64) -- whatever ``date`` prints is
  spliced into an arithmetic expression.  A bare word is a variable name (unset
  = 0, so the poll fires), ``-15`` is divisible by 15, ``08`` / nothing at all /
  ``15 30`` are arithmetic *errors*: the ``((`` command fails with a message on
  stderr and ``!`` turns that into "poll now";
* ``for (( i = 1; i <= $KUBERNETES_NUMBER_OF_NODES; i++ ))`` (setup.sh:148) --
  the exported value from ``./config``;
* ``[[ "$network" -gt "$countNetwork" || "$network" -lt 1 ]]`` (setup.sh:340,
  383,431) -- both operands of ``-gt``/``-lt`` inside ``[[ ]]`` are arithmetic
  expressions, and an operand bash cannot evaluate makes the test *false* after
  a message.

So the evaluator is part of the path.  This module follows the behaviour of
bash 5.2's evaluator (precedence climbing with one token of look-ahead, tokens
evaluated as they are read, variables whose values are themselves expressions,
64-bit wrap-around) closely enough to reproduce its diagnostics, including the
``(error token is "...")`` tail, and is held to the real ``bash`` by
tests/test_basharith.py.

Not evaluated, by design: a ``$``, backquote or backslash inside an array
subscript.  bash runs word expansion -- command substitution included -- on the
text between ``[`` and ``]``; here that raises :class:`ArithRefused`
(INTEGRATION.md "Deviations").
"""
from __future__ import annotations

import sys
from typing import Callable, Dict, Optional, Tuple

MASK = 0xFFFFFFFFFFFFFFFF
INT_MIN = -(1 << 63)
MAX_DEPTH = 1024            # bash: MAX_EXPR_RECURSION_LEVEL

_WS = " \t\n"
_ALPHA = frozenset("abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ")
_DIGITS = frozenset("0123456789")
_VAR_START = _ALPHA | {"_"}
_VAR_CHAR = _VAR_START | _DIGITS
_ALNUM = _ALPHA | _DIGITS
_NUM_CHAR = _ALNUM | {"#", "@", "_"}

# token kinds: single-character operators are themselves; the rest are named
END, NUM, STR = "END", "NUM", "STR"
EQEQ, NEQ, LEQ, GEQ, LAND, LOR, LSH, RSH = "==", "!=", "<=", ">=", "&&", "||", "<<", ">>"
OP_ASSIGN, COND, POWER = "op=", "?:", "**"
PREINC, PREDEC, POSTINC, POSTDEC = "++x", "--x", "x++", "x--"
_SINGLE_OPS = frozenset("=><+-*/%!()&|^~?:,")
_MULTI_OPS = frozenset([EQEQ, NEQ, LEQ, GEQ, LAND, LOR, LSH, RSH, OP_ASSIGN, COND, POWER,
                        PREINC, PREDEC, POSTINC, POSTDEC])


def wrap(v: int) -> int:
    """Reduce to a signed 64-bit value (intmax_t arithmetic wraps)."""
    v &= MASK
    return v - (1 << 64) if v >> 63 else v


class ArithError(Exception):
    """The expression is in error: the command that evaluated it fails (status 1)
    after printing ``<shell prefix><text>`` on stderr."""

    def __init__(self, text: str):
        super().__init__(text)
        self.text = text


class ArithAbort(Exception):
    """An error while evaluating an array subscript: bash abandons the whole
    top-level command (for this script: ``main "$@"``, i.e. the run ends with
    status 1).  ``text`` goes to stderr after the plain ``prog: line N: `` prefix
    (no command name)."""

    def __init__(self, text: str):
        super().__init__(text)
        self.text = text


class ArithFatal(Exception):
    """A diagnostic bash prints with ``report_error`` (a bad array subscript):
    harmless by itself -- the element just counts as unset -- but under ``set -e``
    that routine ends the shell on the spot, with the status of the last command
    if that was non-zero, else 1.  ``text`` follows the ``prog: line N: `` prefix."""

    def __init__(self, text: str):
        super().__init__(text)
        self.text = text


class ArithRefused(Exception):
    """Evaluating this would mean running word expansion on data (see module
    docstring)."""


class _Frame:
    """Evaluation state of one (sub)expression."""

    __slots__ = ("s", "tp", "lasttp", "curtok", "lasttok", "tokstr", "tokval", "assigntok", "shown_end", "lval_ind")

    def __init__(self, s: str):
        self.s = s
        self.tp = 0
        self.lasttp: Optional[int] = None
        self.curtok = END
        self.lasttok = END
        self.tokstr: Optional[str] = None
        self.tokval = 0
        self.assigntok = ""
        #: while a number is being converted bash has cut the string off after it
        self.shown_end: Optional[int] = None
        #: element index of the last variable reference evaluated (-1: not an element)
        self.lval_ind = -1


class Evaluator:
    """``get(name)`` returns a variable's value or None when unset; ``put(name,
    value)`` assigns.  ``messages`` collects diagnostics that do not end the
    evaluation (bad subscripts): each is text for after the ``prog: line N: ``
    prefix."""

    def __init__(self, get: Callable[[str], Optional[str]], put: Callable[[str, str], None],
                 arrays: Optional[Dict[str, Dict[int, str]]] = None, errexit: bool = False):
        self._get = get
        self._put = put
        #: the shell runs with ``set -e``: see :class:`ArithFatal`
        self._errexit = errexit
        #: variables that became arrays because an expression assigned an element
        self._arrays: Dict[str, Dict[int, str]] = {} if arrays is None else arrays
        self.messages: list = []
        self._depth = 0
        self._noeval = 0
        self._f = _Frame("")
        self._outer_lasttp_text = ""

    # -- public ------------------------------------------------------------
    def evaluate(self, expr: str) -> int:
        """Value of ``expr``; raises :class:`ArithError`, :class:`ArithAbort`,
        :class:`ArithFatal` or :class:`ArithRefused`."""
        self._depth = 0
        self._noeval = 0
        saved = self._f
        # a variable may name a variable that names a variable ... up to MAX_DEPTH
        # levels, each a couple of dozen Python frames deep per level of parentheses
        limit = sys.getrecursionlimit()
        sys.setrecursionlimit(max(limit, 1000000))
        try:
            return self._subexpr(expr)
        except RecursionError:
            # deeper than any clock reading or config value will ever nest; bash itself
            # would be near the end of its C stack here
            raise ArithError("%s: expression recursion level exceeded (error token is \"%s\")" % (expr.lstrip(" \t"), expr))
        finally:
            sys.setrecursionlimit(limit)
            self._f = saved

    # -- errors ------------------------------------------------------------
    def _report(self, text: str) -> None:
        if self._errexit:
            raise ArithFatal(text)
        self.messages.append(text)

    def _error(self, msg: str) -> None:
        f = self._f
        shown = f.s if f.shown_end is None else f.s[:f.shown_end]
        shown = shown.lstrip(" \t")
        if f.lasttp is None:
            token = self._outer_lasttp_text
        else:
            end = len(f.s) if f.shown_end is None else f.shown_end
            token = f.s[f.lasttp:end]
        raise ArithError('%s: %s (error token is "%s")' % (shown, msg, token))

    # -- expression stack ----------------------------------------------------
    def _subexpr(self, expr: str) -> int:
        if expr.strip(_WS) == "":
            return 0
        if self._depth >= MAX_DEPTH:
            self._error("expression recursion level exceeded")
        outer = self._f
        # a fresh expression inherits the error-token pointer of the one that
        # referred to it until its own first token is read
        if outer.lasttp is not None:
            end = len(outer.s) if outer.shown_end is None else outer.shown_end
            self._outer_lasttp_text = outer.s[outer.lasttp:end]
        self._depth += 1
        self._f = _Frame(expr)
        try:
            self._readtok()
            val = self._expcomma()
            if self._f.curtok != END:
                self._error("syntax error in expression")
        finally:
            self._f = outer
            self._depth -= 1
        return val

    # -- variables -----------------------------------------------------------
    def _split_subscript(self, tok: str) -> Tuple[str, Optional[str]]:
        if tok.endswith("]") and "[" in tok:
            name, _, rest = tok.partition("[")
            return name, rest[:-1]
        return tok, None

    def _index(self, sub: str) -> int:
        """Value of an array subscript.  Its errors abandon the command."""
        if any(c in sub for c in "$`\\"):
            raise ArithRefused("refusing to run word expansion on an array subscript: %r" % sub)
        text = sub.replace('"', "")          # quote removal of the (double-quoted) subscript
        inner = Evaluator(self._get, self._put, self._arrays, self._errexit)
        try:
            val = inner.evaluate(text)
        except ArithError as exc:
            raise ArithAbort(exc.text)
        finally:
            self.messages.extend(inner.messages)
        return val

    def _lookup(self, tok: str) -> Tuple[Optional[str], int]:
        """(value, element index or -1) of ``name`` or ``name[sub]``.  Variables
        are scalars -- element 0 is the value, any other element is unset --
        until an element is assigned from inside an expression."""
        name, sub = self._split_subscript(tok)
        arr = self._arrays.get(name)
        if sub is None:
            return (self._get(name) if arr is None else arr.get(0)), -1
        if sub in ("@", "*"):
            self._report("%s: bad array subscript" % tok)
            return None, -1
        ind = self._index(sub)
        if ind < 0 and arr:
            ind += max(arr) + 1
        if ind < 0:
            self._report("%s: bad array subscript" % name)
            return None, ind
        if arr is not None:
            return arr.get(ind), ind
        return (self._get(name) if ind == 0 else None), ind

    def _streval(self, tok: str) -> int:
        if self._noeval:
            return 0
        f = self._f
        value, ind = self._lookup(tok)
        val = self._subexpr(value) if value else 0
        f.lval_ind = ind if self._split_subscript(tok)[1] is not None else -1
        return val

    def _bind(self, tok: str, value: int, ind: int = -1) -> None:
        """Assign; ``ind`` is the element index remembered from evaluating the
        left-hand side (bash does not evaluate the subscript a second time)."""
        name, sub = self._split_subscript(tok)
        if ind == -1 and sub is not None:
            ind = self._index(sub)
            if ind < 0 and self._arrays.get(name):
                ind += max(self._arrays[name]) + 1
            if ind < 0:
                self._report("%s: bad array subscript" % tok)
                return
        if ind == -1:
            if name in self._arrays:
                self._arrays[name][0] = str(value)
            self._put(name, str(value))
            return
        if name not in self._arrays:
            cur = self._get(name)
            self._arrays[name] = {} if cur is None else {0: cur}
        self._arrays[name][ind] = str(value)
        if ind == 0:
            self._put(name, str(value))

    # -- tokenizer -------------------------------------------------------------
    def _skip_subscript(self, s: str, start: int) -> int:
        """Index of the ``]`` matching the ``[`` at ``start``, or len(s)."""
        depth = 0
        i, n = start, len(s)
        while i < n:
            c = s[i]
            if c == "\\":
                i += 2
                continue
            if c == "[":
                depth += 1
            elif c == "]":
                depth -= 1
                if depth == 0:
                    return i
            elif c in "'\"`":
                j = s.find(c, i + 1)
                if j == -1:
                    return n
                i = j
            i += 1
        return n

    def _readtok(self) -> None:
        f = self._f
        s, n = f.s, len(f.s)
        cp = f.tp
        while cp < n and s[cp] in _WS:
            cp += 1
        if cp >= n:
            f.lasttok = f.curtok
            f.curtok = END
            f.tp = n
            return
        c = s[cp]
        f.lasttp = f.tp = cp
        cp += 1
        if c in _VAR_START:
            while cp < n and s[cp] in _VAR_CHAR:
                cp += 1
            if cp < n and s[cp] == "[":
                e = self._skip_subscript(s, cp)
                if e < n:
                    cp = e + 1
                else:
                    self._error("bad array subscript")
            tokstr = s[f.tp:cp]
            # peek at the next token (nothing is evaluated while peeking)
            saved = (f.curtok, f.lasttok, f.tp, f.lasttp, f.tokval, f.tokstr, f.assigntok)
            saved_noeval = self._noeval
            f.tp = cp
            self._noeval = 1
            f.curtok = STR
            f.tokstr = None
            self._readtok()
            peek = f.curtok
            f.curtok, f.lasttok, f.tp, f.lasttp, f.tokval, _, f.assigntok = saved
            self._noeval = saved_noeval
            f.tokstr = tokstr
            if f.lasttok in (PREINC, PREDEC) or peek != "=":
                f.tokval = self._streval(tokstr)
            else:
                f.tokval = 0
            f.lasttok = f.curtok
            f.curtok = STR
        elif c in _DIGITS:
            while cp < n and s[cp] in _NUM_CHAR:
                cp += 1
            f.shown_end = cp
            f.tokval = self._strlong(s[f.tp:cp])
            f.shown_end = None
            f.lasttok = f.curtok
            f.curtok = NUM
        else:
            c1 = s[cp] if cp < n else ""
            cp += 1                       # past c1 (may be one past the end; every path below fixes it)
            tok = c
            if c == "=" and c1 == "=":
                tok = EQEQ
            elif c == "!" and c1 == "=":
                tok = NEQ
            elif c == ">" and c1 == "=":
                tok = GEQ
            elif c == "<" and c1 == "=":
                tok = LEQ
            elif c == "<" and c1 == "<":
                if cp < n and s[cp] == "=":
                    f.assigntok, tok = LSH, OP_ASSIGN
                    cp += 1
                else:
                    tok = LSH
            elif c == ">" and c1 == ">":
                if cp < n and s[cp] == "=":
                    f.assigntok, tok = RSH, OP_ASSIGN
                    cp += 1
                else:
                    tok = RSH
            elif c == "&" and c1 == "&":
                tok = LAND
            elif c == "|" and c1 == "|":
                tok = LOR
            elif c == "*" and c1 == "*":
                tok = POWER
            elif c in "-+" and c1 == c and f.curtok == STR:
                tok = POSTDEC if c == "-" else POSTINC
            elif c in "-+" and c1 == c and f.curtok == NUM and f.lasttok in (PREINC, PREDEC):
                self._error("%s%s: assignment requires lvalue" % (c, c))
            elif c in "-+" and c1 == c:
                xp = cp
                while xp < n and s[xp] in _WS:
                    xp += 1
                if xp < n and s[xp] in _VAR_START:
                    tok = PREDEC if c == "-" else PREINC
                else:
                    cp -= 1
            elif c1 == "=" and c in "*/%+-&^|":
                f.assigntok, tok = c, OP_ASSIGN
            elif c not in _SINGLE_OPS:
                if f.curtok == END or f.curtok in _SINGLE_OPS or f.curtok in _MULTI_OPS:
                    self._error("syntax error: operand expected")
                self._error("syntax error: invalid arithmetic operator")
            else:
                cp -= 1
            f.lasttok = f.curtok
            f.curtok = tok
        f.tp = min(cp, n)

    def _strlong(self, num: str) -> int:
        i, n = 0, len(num)
        base, foundbase, val = 10, False, 0
        if num[0] == "0":
            i = 1
            if i >= n:
                return 0
            if num[i] in "xX":
                base = 16
                i += 1
            else:
                base = 8
            foundbase = True
        while i < n:
            c = num[i]
            i += 1
            if c == "#":
                if foundbase:
                    self._error("invalid number")
                if val < 2 or val > 64:
                    self._error("invalid arithmetic base")
                base, val, foundbase = val, 0, True
                if not (i < n and num[i] in (_ALNUM | {"_", "@"})):
                    self._error("invalid integer constant")
            else:
                if c in _DIGITS:
                    d = ord(c) - 48
                elif "a" <= c <= "z":
                    d = ord(c) - 97 + 10
                elif "A" <= c <= "Z":
                    d = ord(c) - 65 + (10 if base <= 36 else 36)
                elif c == "@":
                    d = 62
                else:
                    d = 63
                if d >= base:
                    self._error("value too great for base")
                val = wrap(val * base + d)
        return val

    # -- grammar, lowest precedence first ---------------------------------------
    def _expcomma(self) -> int:
        val = self._expassign()
        while self._f.curtok == ",":
            self._readtok()
            val = self._expassign()
        return val

    def _expassign(self) -> int:
        f = self._f
        val = self._expcond()
        if f.curtok in ("=", OP_ASSIGN):
            special = f.curtok == OP_ASSIGN
            if f.lasttok != STR:
                self._error("attempted assignment to non-variable")
            op, lvalue = "", 0
            if special:
                op, lvalue = f.assigntok, val
            if f.tokstr is None:
                self._error("syntax error in variable assignment")
            lhs = f.tokstr
            lind = f.lval_ind if special else -1
            self._readtok()
            val = self._expassign()
            if special:
                if op in "/%" and val == 0:
                    if self._noeval == 0:
                        self._error("division by 0")
                    val = 1
                lvalue = self._binary(op, lvalue, val)
                val = lvalue
            if self._noeval == 0:
                self._bind(lhs, val, lind)
            f.tokstr = None
        return val

    def _expcond(self) -> int:
        f = self._f
        cval = self._explor()
        rval = cval
        if f.curtok == "?":
            skip = cval == 0
            if skip:
                self._noeval += 1
            self._readtok()
            if f.curtok in (END, ":"):
                self._error("expression expected")
            val1 = self._expcomma()
            if skip:
                self._noeval -= 1
            if f.curtok != ":":
                self._error("`:' expected for conditional expression")
            skip = cval != 0
            if skip:
                self._noeval += 1
            self._readtok()
            if f.curtok == END:
                self._error("expression expected")
            val2 = self._expcond()
            if skip:
                self._noeval -= 1
            rval = val1 if cval else val2
            f.lasttok = COND
        return rval

    def _explor(self) -> int:
        f = self._f
        val1 = self._expland()
        while f.curtok == LOR:
            skip = val1 != 0
            if skip:
                self._noeval += 1
            self._readtok()
            val2 = self._expland()
            if skip:
                self._noeval -= 1
            val1 = 1 if (val1 or val2) else 0
            f.lasttok = LOR
        return val1

    def _expland(self) -> int:
        f = self._f
        val1 = self._level(0)
        while f.curtok == LAND:
            skip = val1 == 0
            if skip:
                self._noeval += 1
            self._readtok()
            val2 = self._level(0)
            if skip:
                self._noeval -= 1
            val1 = 1 if (val1 and val2) else 0
            f.lasttok = LAND
        return val1

    #: left-associative binary levels between && and * / %
    _LEVELS = (("|",), ("^",), ("&",), (EQEQ, NEQ), (LEQ, GEQ, "<", ">"), (LSH, RSH), ("+", "-"))

    def _level(self, k: int) -> int:
        if k == len(self._LEVELS):
            return self._expmuldiv()
        f = self._f
        ops = self._LEVELS[k]
        val1 = self._level(k + 1)
        while f.curtok in ops:
            op = f.curtok
            self._readtok()
            val2 = self._level(k + 1)
            val1 = self._binary(op, val1, val2)
            f.lasttok = NUM
        return val1

    def _expmuldiv(self) -> int:
        f = self._f
        val1 = self._exppower()
        while f.curtok in ("*", "/", "%"):
            op = f.curtok
            stp = f.tp
            self._readtok()
            val2 = self._exppower()
            if op in "/%" and val2 == 0:
                if self._noeval == 0:
                    while stp < len(f.s) and f.s[stp] in " \t":
                        stp += 1
                    f.lasttp = stp
                    self._error("division by 0")
                val2 = 1
            val1 = self._binary(op, val1, val2)
            f.lasttok = NUM
        return val1

    def _exppower(self) -> int:
        f = self._f
        val1 = self._exp1()
        while f.curtok == POWER:
            self._readtok()
            val2 = self._exppower()
            f.lasttok = NUM
            if val2 == 0:
                return 1
            if val2 < 0:
                self._error("exponent less than 0")
            val1 = wrap(pow(val1, val2, 1 << 64))
        return val1

    def _exp1(self) -> int:
        f = self._f
        tok = f.curtok
        if tok in ("!", "~", "-", "+"):
            self._readtok()
            v = self._exp1()
            f.lasttok = NUM
            if tok == "!":
                return 0 if v else 1
            if tok == "~":
                return wrap(~v)
            if tok == "-":
                return wrap(-v)
            return v
        return self._exp0()

    def _exp0(self) -> int:
        f = self._f
        if f.curtok in (PREINC, PREDEC):
            stok = f.lasttok = f.curtok
            self._readtok()
            if f.curtok != STR:
                self._error("identifier expected after pre-increment or pre-decrement")
            val = wrap(f.tokval + (1 if stok == PREINC else -1))
            if self._noeval == 0:
                self._bind(f.tokstr or "", val, f.lval_ind)
            f.curtok = NUM
            self._readtok()
            return val
        if f.curtok == "(":
            self._readtok()
            val = self._expcomma()
            if f.curtok != ")":
                self._error("missing `)'")
            self._readtok()
            return val
        if f.curtok in (NUM, STR):
            val = f.tokval
            if f.curtok == STR:
                saved = (f.curtok, f.lasttok, f.tp, f.lasttp, f.tokval, f.tokstr, f.assigntok)
                saved_noeval, saved_ind = self._noeval, f.lval_ind
                self._noeval = 1
                self._readtok()
                stok = f.curtok
                f.lval_ind = saved_ind
                if stok in (POSTINC, POSTDEC):
                    f.tokstr = saved[5]
                    self._noeval = saved_noeval
                    f.lasttok = STR
                    v2 = wrap(val + (1 if stok == POSTINC else -1))
                    if self._noeval == 0:
                        self._bind(f.tokstr or "", v2, saved_ind)
                    f.curtok = NUM
                else:
                    f.curtok, f.lasttok, f.tp, f.lasttp, f.tokval, f.tokstr, f.assigntok = saved
                    self._noeval = saved_noeval
            self._readtok()
            return val
        self._error("syntax error: operand expected")
        return 0

    # -- operators -----------------------------------------------------------
    @staticmethod
    def _binary(op: str, a: int, b: int) -> int:
        if op == "*":
            return wrap(a * b)
        if op in "/%" and len(op) == 1:
            if a == INT_MIN and b == -1:
                return 0 if op == "%" else INT_MIN
            q = abs(a) // abs(b)
            if (a < 0) != (b < 0):
                q = -q
            return wrap(q) if op == "/" else wrap(a - q * b)
        if op == "+":
            return wrap(a + b)
        if op == "-":
            return wrap(a - b)
        if op == LSH:
            return wrap(a << (b & 63))
        if op == RSH:
            return a >> (b & 63)
        if op == "&":
            return a & b
        if op == "|":
            return a | b
        if op == "^":
            return a ^ b
        if op == EQEQ:
            return int(a == b)
        if op == NEQ:
            return int(a != b)
        if op == "<":
            return int(a < b)
        if op == ">":
            return int(a > b)
        if op == LEQ:
            return int(a <= b)
        if op == GEQ:
            return int(a >= b)
        raise AssertionError("unknown operator %r" % op)


def evaluate(expr: str, variables: Dict[str, str]) -> int:
    """Convenience wrapper over a plain dict of shell variables."""
    return Evaluator(variables.get, variables.__setitem__).evaluate(expr)
