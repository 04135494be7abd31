"""Line tools of the path (``grep``, ``sed``, ``tr``, ``sort``, ``uniq``, ``awk``,
``tail``) restated as functions on text, C locale.

Each function takes the bytes a pipeline stage would read (as ``str``) and
returns the bytes it would write, so a pipeline of the reference reads here as a
chain of calls in the same order.  Only the option/flag combinations the
reference uses are covered (cited per function).
"""
from __future__ import annotations

import re
import unicodedata
from typing import Callable, List, Optional

from .bashsem import c_sort_key


def lines(text: str) -> List[str]:
    """Records as a line-oriented tool sees them (a last unterminated line counts)."""
    if text == "":
        return []
    ls = text.split("\n")
    if ls[-1] == "":
        ls.pop()
    return ls


def unlines(ls: List[str]) -> str:
    return "".join(l + "\n" for l in ls)


def grep(text: str, pattern: str, invert: bool = False, ignore_case: bool = False,
         binary: Optional[Callable[[], None]] = None, utf8: bool = False) -> str:
    """``grep [-v] [-i] PATTERN`` with a Python-syntax equivalent of the BRE.

    What grep takes for binary data it does not print.  Input holding a NUL byte is binary as a
    whole: no line comes out (and while matching, each NUL ends a line).  When the locale's character set is UTF-8 (``utf8``), a selected line
    that is not valid UTF-8 is left out, the others are printed.  Either way grep says ``binary
    file matches`` on stderr, once, if it held anything back -- ``binary()`` is called then."""
    rx = re.compile(pattern, re.I if ignore_case else 0)

    def hit(line: str) -> bool:
        if utf8:
            # a byte that is no character is matched by nothing, not even by `.`; Python's `.` stops at a newline
            line = "".join("\n" if "\udc80" <= c <= "\udcff" else c for c in line)
        return rx.search(line) is not None
    held_back = False
    if "\0" in text:
        # binary input: a NUL ends a line just as a newline does while grep looks for matches (so `.` never
        # gets to match one), and nothing is printed
        held_back, selected = any(hit(l) != invert for l in lines(text.replace("\0", "\n"))), []
    else:
        selected = [l for l in lines(text) if hit(l) != invert]
        if utf8:
            clean = [l for l in selected if not any("\udc80" <= c <= "\udcff" for c in l)]
            held_back, selected = len(clean) != len(selected), clean
    if held_back and binary is not None:
        binary()
    return unlines(selected)


def tr_squeeze_spaces(text: str) -> str:
    """``tr -s " "`` (setup.sh:257,259,541)."""
    return re.sub(r" +", " ", text)


def tr(text: str, a: str, b: str) -> str:
    """``tr A B`` for single characters (setup.sh:257,334,533)."""
    return text.replace(a, b)


def sort_lines(text: str) -> str:
    """``sort`` under LC_ALL=C (setup.sh:257,334,536)."""
    return unlines(sorted(lines(text), key=c_sort_key))


def uniq(text: str) -> str:
    """``uniq`` (setup.sh:334)."""
    out: List[str] = []
    for l in lines(text):
        if not out or out[-1] != l:
            out.append(l)
    return unlines(out)


#: what `.` matches when the locale's character set is UTF-8: any character, but not a byte that is none
UTF8_DOT = "[^\udc80-\udcff\n]"


def utf8_pattern(pattern: str) -> str:
    """The same regular expression for a UTF-8 locale (unescaped dots only; the patterns of the path are that simple)."""
    return re.sub(r"(?<!\\)\.", lambda m: UTF8_DOT, pattern)


def sed_sub(text: str, pattern: str, repl, count: int = 1, utf8: bool = False) -> str:
    """``sed 's/PATTERN/REPL/[g]'`` line by line; ``count=0`` is the ``g`` flag.
    GNU sed keeps a missing final newline missing."""
    rx = re.compile(utf8_pattern(pattern) if utf8 else pattern)
    ls = text.split("\n")
    tail_nl = ls[-1] == ""
    if tail_nl:
        ls.pop()
    out = "\n".join(rx.sub(repl, l, count=count) for l in ls)
    if text == "":
        return ""
    return out + ("\n" if tail_nl else "")


def strip_dquotes(text: str) -> str:
    """``sed 's/"//g'`` -- the quote-stripper used at 15+ sites (e.g. setup.sh:130)."""
    return text.replace('"', "")

#: what GNU sed 4.9 prints below "invalid option" (its usage(), status 1): reachable at setup.sh:541 with an index
#: word such as ``-n`` -- `sed -n "-n"p` is `sed -n -np`
SED_USAGE = (
    'Usage: sed [OPTION]... {script-only-if-no-other-script} [input-file]...\n'
    '\n'
    '  -n, --quiet, --silent\n'
    '                 suppress automatic printing of pattern space\n'
    '      --debug\n'
    '                 annotate program execution\n'
    '  -e script, --expression=script\n'
    '                 add the script to the commands to be executed\n'
    '  -f script-file, --file=script-file\n'
    '                 add the contents of script-file to the commands to be executed\n'
    '  --follow-symlinks\n'
    '                 follow symlinks when processing in place\n'
    '  -i[SUFFIX], --in-place[=SUFFIX]\n'
    '                 edit files in place (makes backup if SUFFIX supplied)\n'
    '  -l N, --line-length=N\n'
    "                 specify the desired line-wrap length for the `l' command\n"
    '  --posix\n'
    '                 disable all GNU extensions.\n'
    '  -E, -r, --regexp-extended\n'
    '                 use extended regular expressions in the script\n'
    '                 (for portability use POSIX -E).\n'
    '  -s, --separate\n'
    '                 consider files as separate rather than as a single,\n'
    '                 continuous long stream.\n'
    '      --sandbox\n'
    '                 operate in sandbox mode (disable e/r/w commands).\n'
    '  -u, --unbuffered\n'
    '                 load minimal amounts of data from the input files and flush\n'
    '                 the output buffers more often\n'
    '  -z, --null-data\n'
    '                 separate lines by NUL characters\n'
    '      --help     display this help and exit\n'
    '      --version  output version information and exit\n'
    '\n'
    'If no -e, --expression, -f, or --file option is given, then the first\n'
    'non-option argument is taken as the sed script to interpret.  All\n'
    'remaining arguments are names of input files; if no input files are\n'
    'specified, then the standard input is read.\n'
    '\n'
    'GNU sed home page: <https://www.gnu.org/software/sed/>.\n'
    'General help using GNU software: <https://www.gnu.org/gethelp/>.\n'
)


def sed_print_line(text: str, n: int) -> str:
    """``sed -n "${n}p"`` (setup.sh:536,541)."""
    ls = lines(text)
    if 1 <= n <= len(ls):
        last_unterminated = n == len(ls) and not text.endswith("\n")
        return ls[n - 1] + ("" if last_unterminated else "\n")   # GNU sed keeps a missing final newline missing
    return ""


def _awk_fields(line: str) -> List[str]:
    return [f for f in re.split(r"[ \t]+", line) if f]


def awk_print_field(text: str, idx: int) -> str:
    """``awk '{print $idx}'`` (setup.sh:68,72,217)."""
    out = []
    for l in lines(text):
        f = _awk_fields(l)
        out.append(f[idx - 1] if len(f) >= idx else "")
    return unlines(out)


def awk_last_field_if_many(text: str) -> str:
    """``awk 'NF>1{print $NF}'`` (setup.sh:536,541)."""
    out = []
    for l in lines(text):
        f = _awk_fields(l)
        if len(f) > 1:
            out.append(f[-1])
    return unlines(out)


def tail_1(text: str) -> str:
    """``tail -1`` (setup.sh:127)."""
    if text == "":
        return ""
    body = text[:-1] if text.endswith("\n") else text
    last = body.rsplit("\n", 1)[-1]
    return last + ("\n" if text.endswith("\n") else "")


def wc_l(text: str) -> int:
    """``wc -l``: newline count."""
    return text.count("\n")


#: what separates words for GNU wc in a UTF-8 locale, beyond the ASCII blanks (found by asking it)
_UTF8_BLANKS = frozenset("\xa0\u1680\u2000\u2001\u2002\u2003\u2004\u2005\u2006\u2007\u2008\u2009\u200a\u202f\u205f\u2060\u3000")


def wc_w(text: str, utf8: bool = False) -> int:
    """``wc -w``: a word is a run of non-blank characters holding at least one printable
    one; what is neither (control characters, bytes that are no character) starts no word
    and ends none.  In the C locale only ASCII is printable; in a UTF-8 locale (``utf8``)
    the characters beyond it are classified too."""
    count, in_word = 0, False
    for c in text:
        if c in " \t\n\r\f\v" or (utf8 and ord(c) > 0x7F and c in _UTF8_BLANKS):
            in_word = False
        elif (0x20 < ord(c) < 0x7F or (utf8 and ord(c) > 0x7F and unicodedata.category(c) not in ("Cc", "Cs", "Cn", "Zl", "Zp"))) and not in_word:
            in_word = True
            count += 1
    return count
