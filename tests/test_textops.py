"""textops restates grep/sed/tr/sort/uniq/awk/tail/wc for the option sets the
path uses; here each is checked against the real tool (coreutils/sed/awk are in
the image, LC_ALL=C)."""
import subprocess

import pytest
from hypothesis import given, settings, strategies as st

from tritonk8ssupervisor_b200 import textops as tx
from tritonk8ssupervisor_b200.bashsem import from_bytes, to_bytes

ENV = {"PATH": "/usr/bin:/bin", "LC_ALL": "C"}


def real(argv, text):
    p = subprocess.run(argv, input=to_bytes(text), stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, env=ENV)
    return from_bytes(p.stdout)


_line = st.text(st.sampled_from(list("ab =,-kvmNAMEID01\t\"")), max_size=12)
_text = st.builds(lambda ls, nl: "\n".join(ls) + ("\n" if nl and ls else ""), st.lists(_line, max_size=8), st.booleans())

CASES = [
    (lambda t: tx.sort_lines(t), ["sort"]),
    (lambda t: tx.uniq(t), ["uniq"]),
    (lambda t: tx.tr_squeeze_spaces(t), ["tr", "-s", " "]),
    (lambda t: tx.tr(t, " ", "="), ["tr", " ", "="]),
    (lambda t: tx.tr(t, ",", "\n"), ["tr", ",", "\n"]),
    (lambda t: tx.tr(t, "\n", ","), ["tr", "\n", ","]),
    (lambda t: tx.grep(t, r"-kvm-"), ["grep", r"\-kvm-"]),
    (lambda t: tx.grep(t, r"^NAME.*ID$", invert=True), ["grep", "-v", "^NAME.*ID$"]),
    (lambda t: tx.grep(t, r"^NAME *ID$", invert=True), ["grep", "-v", "^NAME *ID$"]),
    (lambda t: tx.grep(t, r"^$", invert=True), ["grep", "-v", "^$"]),
    (lambda t: tx.grep(t, r"name", ignore_case=True), ["grep", "-i", "name"]),
    (lambda t: tx.grep(t, r"^[1-9][0-9]?(,[1-9][0-9]?)*$"), ["grep", r"^[1-9][0-9]\?\(,[1-9][0-9]\?\)*$"]),
    (lambda t: tx.strip_dquotes(t), ["sed", 's/"//g']),
    (lambda t: tx.sed_sub(t, r"=", "  ", count=0), ["sed", "s/=/  /g"]),
    (lambda t: tx.sed_sub(t, r",", '","', count=0), ["sed", 's/,/","/g']),
    (lambda t: tx.sed_sub(t, r"(.*),$", r"\1"), ["sed", r"s/\(.*\),$/\1/"]),
    (lambda t: tx.sed_sub(t, r"^,(.*)$", r"\1"), ["sed", r"s/^,\(.*\)$/\1/"]),
    (lambda t: tx.sed_sub(t, r".pub$", ""), ["sed", "s/.pub$//"]),
    (lambda t: tx.sed_print_line(t, 2), ["sed", "-n", "2p"]),
    (lambda t: tx.sed_print_line(t, 9), ["sed", "-n", "9p"]),
    (lambda t: tx.awk_print_field(t, 2), ["awk", "{print $2}"]),
    (lambda t: tx.awk_print_field(t, 1), ["awk", "{print $1}"]),
    (lambda t: tx.awk_last_field_if_many(t), ["awk", "NF>1{print $NF}"]),
    (lambda t: tx.tail_1(t), ["tail", "-1"]),
    (lambda t: "%d\n" % tx.wc_l(t), ["wc", "-l"]),
    (lambda t: "%d\n" % tx.wc_w(t), ["wc", "-w"]),
]

FIXED = ["", "\n", "a\n", "a", "b\na\nb\n", "NAME   ID\nk4-kvm-1  u1\nx  u2\n", "2,4,\n", ",a,b\n", "a  b   c\n \n",
         "id_rsa.pub\nmypub\n", "x\n\ny\n", "1,2\n10,2\n007\n", 'say "hi"\n', "\t lead\ttab  \n", "a\n\n"]


@pytest.mark.parametrize("idx", range(len(CASES)), ids=[" ".join(c[1]) for c in CASES])
def test_against_real_tool_fixed(idx):
    ours, argv = CASES[idx]
    for text in FIXED:
        assert ours(text) == real(argv, text), (argv, text)


@settings(max_examples=25, deadline=None, derandomize=True)
@given(text=_text)
def test_against_real_tools_random(text):
    for ours, argv in CASES:
        assert ours(text) == real(argv, text), (argv, text)


def test_grep_in_a_utf8_locale_holds_back_what_is_no_text():
    """LANG=C.UTF-8: a selected line that is not valid UTF-8 is not printed (the others are),
    `.` does not match such a byte, and a NUL byte anywhere silences grep altogether."""
    env = {"PATH": "/usr/bin:/bin", "LANG": "C.UTF-8"}
    cases = [("a=1\n\nb=\udcffx\nc=3\nd=\udcff\ne=5\n", ["grep", "-v", "^$"], dict(pattern=r"^$", invert=True)),
             ("NAME  \udcff ID\nNAME   ID\nx y\n", ["grep", "-v", "^NAME.*ID$"], dict(pattern=r"^NAME.*ID$", invert=True)),
             ("kubernetes \udcff\nkubernetes\n", ["grep", "-i", "kubernetes"], dict(pattern=r"kubernetes", ignore_case=True)),
             ("a\x00b\nc\n", ["grep", "c"], dict(pattern=r"c")),
             ("plain\ntext\n", ["grep", "t"], dict(pattern=r"t"))]
    for text, argv, kw in cases:
        p = subprocess.run(argv, input=to_bytes(text), capture_output=True, env=env)
        said = []
        assert tx.grep(text, utf8=True, binary=lambda: said.append(1), **kw) == from_bytes(p.stdout), argv
        assert bool(said) == (b"binary file matches" in p.stderr), argv


def test_sed_in_a_utf8_locale_does_not_match_a_stray_byte_with_a_dot():
    env = {"PATH": "/usr/bin:/bin", "LANG": "C.UTF-8"}
    cases = [("x\udcffpub\nxapub\n.pub\n", ["sed", "s/.pub$//"], (r".pub$", "")),
             ("1,\udcff,\n2,3,\n", ["sed", r"s/\(.*\),$/\1/"], (r"(.*),$", r"\1")),
             (",a\udcff\n,b\n", ["sed", r"s/^,\(.*\)$/\1/"], (r"^,(.*)$", r"\1"))]
    for text, argv, (pattern, repl) in cases:
        p = subprocess.run(argv, input=to_bytes(text), capture_output=True, env=env)
        assert tx.sed_sub(text, pattern, repl, utf8=True) == from_bytes(p.stdout), argv
        c = subprocess.run(argv, input=to_bytes(text), capture_output=True, env=ENV)
        assert tx.sed_sub(text, pattern, repl) == from_bytes(c.stdout), argv


def test_wc_w_classifies_characters_as_the_locale_does():
    import random
    r = random.Random(2)
    chars = ["\u00e9", "\udcff", "a", " ", "\xa0", "\u2003", "\u200b", "\x01", "\u65e5", "\t", "\x85", "\u2028", "\x7f", "\udc80", "b",
             "\xad", "\ue000", "\U0001F600", "\u3000", "\u1680", "\x1c", "\u2007", "\u180e", "\u2060", "\u2029", "\u205f"]
    for _ in range(300):
        text = "".join(r.choice(chars) for _ in range(r.randint(0, 8))) + "\n"
        for env, utf8 in (({"LC_ALL": "C"}, False), ({"LANG": "C.UTF-8"}, True)):
            real = int(subprocess.run(["wc", "-w"], input=to_bytes(text), capture_output=True, env=dict(env, PATH="/usr/bin:/bin")).stdout)
            assert tx.wc_w(text, utf8=utf8) == real, (ascii(text), utf8)


def test_grep_on_input_with_nul_bytes_matches_the_real_grep():
    """Binary input: nothing is printed, `binary file matches` is said if a line was selected -- and while it
    looks for matches grep lets a NUL end a line, so `^NAME.*ID$` does not match `NAME \\0 ID`."""
    import random
    import subprocess
    from tritonk8ssupervisor_b200 import textops as tx
    r = random.Random(4)
    pieces = ["NAME", "ID", " ", "  ", "\0", "\n", "x-kvm-1", "kubernetes", "Service Unavailable", "k8s_POD.a_dashboard", "a", "SHA256", ""]
    patterns = [("^NAME.*ID$", r"^NAME.*ID$", True, False), ("^NAME *ID$", r"^NAME *ID$", True, False), ("\\-kvm-", r"-kvm-", False, False),
                ("kubernetes", r"kubernetes", False, True), ("k8s_POD.*dashboard", r"k8s_POD.*dashboard", False, False), ("^$", r"^$", True, False),
                ("^SHA256", r"^SHA256", False, False)]
    for _ in range(300):
        text = "".join(r.choice(pieces) for _ in range(r.randint(1, 8))) + r.choice(["\n", "", "\0"])
        bre, py, invert, icase = r.choice(patterns)
        argv = ["grep"] + (["-v"] if invert else []) + (["-i"] if icase else []) + [bre]
        real = subprocess.run(argv, input=text.encode(), capture_output=True, env={"LC_ALL": "C", "PATH": "/usr/bin:/bin"})
        said = []
        mine = tx.grep(text, py, invert=invert, ignore_case=icase, binary=lambda: said.append(1))
        assert (mine.encode(), bool(said)) == (real.stdout, b"binary file matches" in real.stderr), (text, bre)
