"""A parser for HCL1 -- the configuration language of Terraform <= 0.11 -- and
for HIL, the ``${...}`` interpolation language inside its strings.

Why it exists: the path generates ``terraform/rancher.tf`` (setup.sh:162-198)
and the reference's ``terraform validate`` cannot be run offline (no terraform
binary, no network).  The structural validator in
:mod:`tfvalidate` stands in for it, and needs a real syntax tree of both the
generated file and the two modules it instantiates
(terraform/{master,host}/{main,vars}.tf).  The dialect is the Terraform 0.8/0.9
one those files are written in: quoted interpolation-only expressions, bare
``source = "master"``, ``type = "list"``, repeated attributes.

The grammar follows the published HCL1 specification (github.com/hashicorp/hcl,
``hcl/parser`` and ``hcl/scanner``, v1 branch) -- third-party code that is not
vendored in the reference; it is restated here, not copied.

AST
---
``Body``      ordered list of ``Item``
``Item``      ``keys`` (first is the attribute/block type, the rest are labels),
              ``value`` (a scalar ``Literal``, ``ListVal``, or ``Body``),
              ``is_block`` (written ``k "l" { … }`` rather than ``k = …``), ``line``
``Literal``   ``kind`` in {"string","number","bool","heredoc"}, ``value``, ``line``;
              for strings ``parts`` is the parsed template: a list of ``str`` and
              HIL expression nodes
"""
from __future__ import annotations

import dataclasses
import os
import re
from typing import Any, Iterator, List, Optional, Tuple, Union


class HCLError(ValueError):
    def __init__(self, msg: str, filename: str = "<hcl>", line: int = 0, col: int = 0):
        super().__init__("%s:%d:%d: %s" % (filename, line, col, msg))
        self.msg, self.filename, self.line, self.col = msg, filename, line, col


# ---------------------------------------------------------------------------
# AST
# ---------------------------------------------------------------------------
@dataclasses.dataclass
class Literal:
    kind: str
    value: Any
    line: int = 0
    parts: Optional[List[Any]] = None  # template parts for strings


@dataclasses.dataclass
class ListVal:
    items: List[Any]
    line: int = 0


@dataclasses.dataclass
class Item:
    keys: List[str]
    value: Any
    is_block: bool
    line: int = 0

    @property
    def name(self) -> str:
        return self.keys[0]

    @property
    def labels(self) -> List[str]:
        return self.keys[1:]


@dataclasses.dataclass
class Body:
    items: List[Item]
    line: int = 0

    def blocks(self, name: str) -> List[Item]:
        return [it for it in self.items if it.name == name and isinstance(it.value, Body)]

    def attrs(self, name: str) -> List[Item]:
        return [it for it in self.items if it.name == name]

    def attr(self, name: str) -> Optional[Item]:
        found = self.attrs(name)
        return found[-1] if found else None


# --- HIL nodes -------------------------------------------------------------
@dataclasses.dataclass
class HVar:
    """``var.x``, ``triton_machine.master.primaryip``, ``self.id``, ``count.index``"""
    path: List[str]


@dataclasses.dataclass
class HCall:
    func: str
    args: List[Any]


@dataclasses.dataclass
class HLit:
    value: Any


@dataclasses.dataclass
class HStr:
    """A quoted string inside an interpolation; itself a template."""
    parts: List[Any]


@dataclasses.dataclass
class HBin:
    op: str
    left: Any
    right: Any


@dataclasses.dataclass
class HUnary:
    op: str
    operand: Any


@dataclasses.dataclass
class HCond:
    cond: Any
    then: Any
    other: Any


@dataclasses.dataclass
class HIndex:
    target: Any
    key: Any


def walk_hil(node: Any) -> Iterator[Any]:
    """Every HIL node under ``node`` (a node, or a template part list)."""
    if isinstance(node, list):
        for p in node:
            yield from walk_hil(p)
        return
    if isinstance(node, str) or node is None:
        return
    yield node
    if isinstance(node, HCall):
        yield from walk_hil(node.args)
    elif isinstance(node, HStr):
        yield from walk_hil(node.parts)
    elif isinstance(node, HBin):
        yield from walk_hil([node.left, node.right])
    elif isinstance(node, HUnary):
        yield from walk_hil(node.operand)
    elif isinstance(node, HCond):
        yield from walk_hil([node.cond, node.then, node.other])
    elif isinstance(node, HIndex):
        yield from walk_hil([node.target, node.key])


# ---------------------------------------------------------------------------
# HIL: template splitting and expression parsing
# ---------------------------------------------------------------------------
_ESCAPES = {"n": "\n", "t": "\t", "r": "\r", '"': '"', "\\": "\\"}


def parse_template(raw: str, filename: str = "<hcl>", line: int = 0) -> List[Any]:
    """Split the *raw* body of a quoted string into literal text and parsed
    ``${...}`` expressions.  ``$${`` is an escaped ``${``."""
    parts: List[Any] = []
    buf: List[str] = []
    i, n = 0, len(raw)
    while i < n:
        c = raw[i]
        if c == "$" and raw.startswith("$${", i):
            buf.append("${")
            i += 3
            continue
        if c == "$" and raw.startswith("${", i):
            end = _find_interp_end(raw, i + 2, filename, line)
            if buf:
                parts.append("".join(buf))
                buf = []
            parts.append(_HilParser(raw[i + 2:end], filename, line).parse())
            i = end + 1
            continue
        if c == "\\" and i + 1 < n:
            d = raw[i + 1]
            if d in _ESCAPES:
                buf.append(_ESCAPES[d])
                i += 2
                continue
            if d == "u" and re.match(r"[0-9a-fA-F]{4}", raw[i + 2:i + 6] or ""):
                buf.append(chr(int(raw[i + 2:i + 6], 16)))
                i += 6
                continue
            raise HCLError("unknown escape sequence \\%s in string" % d, filename, line)
        buf.append(c)
        i += 1
    if buf:
        parts.append("".join(buf))
    return parts


def _find_interp_end(raw: str, start: int, filename: str, line: int) -> int:
    """Index of the '}' closing the interpolation opened before ``start``;
    nested braces and quoted strings (which may hold ``${`` again) are skipped."""
    depth = 1
    i, n = start, len(raw)
    while i < n:
        c = raw[i]
        if c == '"':
            i = _skip_quoted(raw, i + 1, filename, line)
            continue
        if c == "{":
            depth += 1
        elif c == "}":
            depth -= 1
            if depth == 0:
                return i
        i += 1
    raise HCLError("interpolation ${ is not closed", filename, line)


def _skip_quoted(raw: str, i: int, filename: str, line: int) -> int:
    """``i`` is just after an opening quote; returns the index after the closing one."""
    n = len(raw)
    while i < n:
        c = raw[i]
        if c == "\\":
            i += 2
            continue
        if c == "$" and raw.startswith("${", i):
            i = _find_interp_end(raw, i + 2, filename, line) + 1
            continue
        if c == '"':
            return i + 1
        i += 1
    raise HCLError("string inside interpolation is not closed", filename, line)


_HIL_TOKEN = re.compile(r"""
    (?P<ws>\s+)
  | (?P<num>[0-9]+(?:\.[0-9]+)?)
  | (?P<id>[A-Za-z_][A-Za-z0-9_\-]*(?:\.(?:[A-Za-z_][A-Za-z0-9_\-]*|[0-9]+|\*))*)
  | (?P<op>&&|\|\||==|!=|<=|>=|[-+*/%<>!?:(),\[\]])
""", re.X)


class _HilParser:
    def __init__(self, src: str, filename: str, line: int):
        self.src, self.filename, self.line = src, filename, line
        self.toks: List[Tuple[str, Any]] = []
        self._lex()
        self.pos = 0

    def _lex(self) -> None:
        i, n = 0, len(self.src)
        while i < n:
            if self.src[i] == '"':
                end = _skip_quoted(self.src, i + 1, self.filename, self.line)
                self.toks.append(("str", self.src[i + 1:end - 1]))
                i = end
                continue
            m = _HIL_TOKEN.match(self.src, i)
            if not m:
                raise HCLError("unexpected character %r in interpolation %r" % (self.src[i], self.src), self.filename, self.line)
            i = m.end()
            if m.lastgroup == "ws":
                continue
            self.toks.append((m.lastgroup, m.group(m.lastgroup)))
        self.toks.append(("eof", None))

    def _peek(self) -> Tuple[str, Any]:
        return self.toks[self.pos]

    def _next(self) -> Tuple[str, Any]:
        t = self.toks[self.pos]
        self.pos += 1
        return t

    def _accept(self, op: str) -> bool:
        k, v = self._peek()
        if k == "op" and v == op:
            self.pos += 1
            return True
        return False

    def _expect(self, op: str) -> None:
        if not self._accept(op):
            raise HCLError("expected %r in interpolation %r" % (op, self.src), self.filename, self.line)

    def parse(self) -> Any:
        if self._peek()[0] == "eof":
            raise HCLError("empty interpolation", self.filename, self.line)
        node = self._cond()
        if self._peek()[0] != "eof":
            raise HCLError("unexpected %r in interpolation %r" % (self._peek()[1], self.src), self.filename, self.line)
        return node

    def _cond(self) -> Any:
        node = self._binary(0)
        if self._accept("?"):
            then = self._cond()
            self._expect(":")
            other = self._cond()
            return HCond(node, then, other)
        return node

    _LEVELS = [["||"], ["&&"], ["==", "!="], ["<", ">", "<=", ">="], ["+", "-"], ["*", "/", "%"]]

    def _binary(self, level: int) -> Any:
        if level >= len(self._LEVELS):
            return self._unary()
        node = self._binary(level + 1)
        while True:
            k, v = self._peek()
            if k == "op" and v in self._LEVELS[level]:
                self.pos += 1
                node = HBin(v, node, self._binary(level + 1))
            else:
                return node

    def _unary(self) -> Any:
        if self._accept("-"):
            return HUnary("-", self._unary())
        if self._accept("!"):
            return HUnary("!", self._unary())
        return self._postfix()

    def _postfix(self) -> Any:
        node = self._primary()
        while self._accept("["):
            key = self._cond()
            self._expect("]")
            node = HIndex(node, key)
        return node

    def _primary(self) -> Any:
        k, v = self._next()
        if k == "num":
            return HLit(float(v) if "." in v else int(v))
        if k == "str":
            return HStr(parse_template(v, self.filename, self.line))
        if k == "id":
            if v in ("true", "false"):
                return HLit(v == "true")
            if self._accept("("):
                args: List[Any] = []
                if not self._accept(")"):
                    while True:
                        args.append(self._cond())
                        if self._accept(")"):
                            break
                        self._expect(",")
                return HCall(v, args)
            return HVar(v.split("."))
        if k == "op" and v == "(":
            node = self._cond()
            self._expect(")")
            return node
        raise HCLError("unexpected %r in interpolation %r" % (v, self.src), self.filename, self.line)


# ---------------------------------------------------------------------------
# HCL1 scanner + parser
# ---------------------------------------------------------------------------
_IDENT_START = re.compile(r"[A-Za-z_]")
_IDENT = re.compile(r"[A-Za-z_][A-Za-z0-9_\-.]*")
_NUMBER = re.compile(r"-?(?:0[xX][0-9a-fA-F]+|[0-9]+(?:\.[0-9]+)?(?:[eE][-+]?[0-9]+)?)")


class _Scanner:
    def __init__(self, src: str, filename: str):
        self.src, self.filename = src, filename
        self.i, self.line, self.col0 = 0, 1, 0

    def error(self, msg: str) -> HCLError:
        return HCLError(msg, self.filename, self.line, self.i - self.col0 + 1)

    def _advance(self, k: int) -> None:
        seg = self.src[self.i:self.i + k]
        nl = seg.count("\n")
        if nl:
            self.line += nl
            self.col0 = self.i + seg.rfind("\n") + 1
        self.i += k

    def skip_space(self) -> None:
        n = len(self.src)
        while self.i < n:
            c = self.src[self.i]
            if c in " \t\r\n":
                self._advance(1)
            elif c == "#" or self.src.startswith("//", self.i):
                j = self.src.find("\n", self.i)
                self._advance((j if j != -1 else n) - self.i)
            elif self.src.startswith("/*", self.i):
                j = self.src.find("*/", self.i + 2)
                if j == -1:
                    raise self.error("comment /* is not closed")
                self._advance(j + 2 - self.i)
            else:
                return

    def peek(self) -> str:
        self.skip_space()
        return self.src[self.i] if self.i < len(self.src) else ""

    def take(self, ch: str) -> bool:
        if self.peek() == ch:
            self._advance(1)
            return True
        return False

    def ident(self) -> str:
        self.skip_space()
        m = _IDENT.match(self.src, self.i)
        if not m:
            raise self.error("expected an identifier, found %r" % self.src[self.i:self.i + 10])
        self._advance(m.end() - self.i)
        return m.group(0)

    def number(self) -> Union[int, float]:
        m = _NUMBER.match(self.src, self.i)
        if not m:
            raise self.error("malformed number")
        self._advance(m.end() - self.i)
        t = m.group(0)
        if re.fullmatch(r"-?0[xX][0-9a-fA-F]+", t):
            return int(t, 16)
        return float(t) if re.search(r"[.eE]", t) else int(t)

    def string_raw(self) -> str:
        """Raw body of a quoted string starting at the opening quote.  Newlines
        are not allowed outside interpolations (HCL1 scanner rule)."""
        assert self.src[self.i] == '"'
        start_line = self.line
        j, n = self.i + 1, len(self.src)
        while j < n:
            c = self.src[j]
            if c == "\\":
                j += 2
                continue
            if c == "$" and self.src.startswith("${", j):
                try:
                    j = _find_interp_end(self.src, j + 2, self.filename, start_line) + 1
                except HCLError:
                    raise self.error("interpolation ${ is not closed")
                continue
            if c == "\n":
                raise self.error("literal string not terminated (newline in string)")
            if c == '"':
                raw = self.src[self.i + 1:j]
                self._advance(j + 1 - self.i)
                return raw
            j += 1
        raise self.error("literal string not terminated")

    def heredoc(self) -> str:
        m = re.compile(r"<<(-?)([A-Za-z_][A-Za-z0-9_]*)\r?\n").match(self.src, self.i)
        if not m:
            raise self.error("malformed heredoc anchor")
        indent, tag = m.group(1) == "-", m.group(2)
        body_start = m.end()
        end = re.compile(r"^[ \t]*%s[ \t]*\r?$" % re.escape(tag), re.M).search(self.src, body_start)
        if not end:
            raise self.error("heredoc <<%s is not terminated" % tag)
        body = self.src[body_start:end.start()]
        if indent:
            ls = body.split("\n")
            pad = min((len(l) - len(l.lstrip(" \t")) for l in ls if l.strip()), default=0)
            body = "\n".join(l[pad:] for l in ls)
        self._advance(end.end() - self.i)
        return body


class Parser:
    def __init__(self, src: str, filename: str = "<hcl>"):
        self.s = _Scanner(src, filename)
        self.filename = filename

    def parse(self) -> Body:
        body = self._body(top=True)
        if self.s.peek() != "":
            raise self.s.error("unexpected %r" % self.s.peek())
        return body

    def _body(self, top: bool) -> Body:
        items: List[Item] = []
        line = self.s.line
        while True:
            c = self.s.peek()
            if c == "" or (c == "}" and not top):
                return Body(items, line)
            if c == "}" and top:
                raise self.s.error("unexpected '}'")
            items.append(self._item())
            self.s.take(",")  # object items may be comma separated

    def _key(self) -> str:
        if self.s.peek() == '"':
            raw = self.s.string_raw()
            return "".join(p for p in parse_template(raw, self.filename, self.s.line) if isinstance(p, str))
        return self.s.ident()

    def _item(self) -> Item:
        self.s.skip_space()
        line = self.s.line
        keys = [self._key()]
        while True:
            c = self.s.peek()
            if c == "=":
                if len(keys) != 1:
                    raise self.s.error("nested object expected: '{' after labels, found '='")
                self.s.take("=")
                return Item(keys, self._value(), False, line)
            if c == "{":
                self.s.take("{")
                body = self._body(top=False)
                if not self.s.take("}"):
                    raise self.s.error("object expected closing '}'")
                return Item(keys, body, True, line)
            if c == '"' or (c and _IDENT_START.match(c)):
                keys.append(self._key())
                continue
            raise self.s.error("expected '=' or '{' after %r, found %r" % (keys[-1], c))

    def _value(self) -> Any:
        c = self.s.peek()
        line = self.s.line
        if c == '"':
            raw = self.s.string_raw()
            parts = parse_template(raw, self.filename, line)
            return Literal("string", raw, line, parts)
        if c == "<":
            return Literal("heredoc", self.s.heredoc(), line, None)
        if c == "[":
            self.s.take("[")
            items: List[Any] = []
            while True:
                if self.s.take("]"):
                    return ListVal(items, line)
                items.append(self._value())
                if self.s.take(","):
                    continue
                if self.s.take("]"):
                    return ListVal(items, line)
                raise self.s.error("expected ',' or ']' in list")
        if c == "{":
            self.s.take("{")
            body = self._body(top=False)
            if not self.s.take("}"):
                raise self.s.error("object expected closing '}'")
            return body
        if c and (c.isdigit() or c == "-"):
            return Literal("number", self.s.number(), line)
        if c and _IDENT_START.match(c):
            word = self.s.ident()
            if word in ("true", "false"):
                return Literal("bool", word == "true", line)
            raise self.s.error("unknown token %r: values must be quoted in this dialect" % word)
        raise self.s.error("expected a value, found %r" % c)


def parse(src: str, filename: str = "<hcl>") -> Body:
    return Parser(src, filename).parse()


_FILE_CACHE: dict = {}


def parse_file(path: str) -> Body:
    """Parse a file; results are memoised on (path, size, mtime) -- validate, plan
    and the mock apply all walk the same few module files.  The tree is treated
    as read-only by every consumer."""
    st = os.stat(path)
    key = (path, st.st_size, st.st_mtime_ns)
    hit = _FILE_CACHE.get(key)
    if hit is None:
        with open(path, encoding="utf-8") as fh:
            hit = parse(fh.read(), path)
        if len(_FILE_CACHE) > 256:
            _FILE_CACHE.clear()
        _FILE_CACHE[key] = hit
    return hit


def templates_in(value: Any) -> Iterator[Tuple[List[Any], int]]:
    """Every string template (parts, line) under a value, depth first."""
    if isinstance(value, Literal):
        if value.kind == "string" and value.parts is not None:
            yield value.parts, value.line
    elif isinstance(value, ListVal):
        for v in value.items:
            yield from templates_in(v)
    elif isinstance(value, Body):
        for it in value.items:
            yield from templates_in(it.value)
