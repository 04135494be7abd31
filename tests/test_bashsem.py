"""bashsem restates bash behaviour; here it is checked against the real bash
(5.2 in this image, present on the GPU box as well)."""
import subprocess

import pytest
from hypothesis import given, settings, strategies as st

from tritonk8ssupervisor_b200 import bashsem

ENV = {"PATH": "/usr/bin:/bin", "LC_ALL": "C"}


@pytest.fixture(scope="module")
def globdir(tmp_path_factory):
    d = tmp_path_factory.mktemp("globdir")
    for name in ["README.md", "setup.sh", "config", "a1", "a2", "b", ".hidden", "x y", "[z]", "-n", "s*r"]:
        (d / name).write_text("")
    (d / "terraform").mkdir()
    (d / "terraform" / "rancher.tf").write_text("")
    (d / "terraform" / "master").mkdir()
    (d / "terraform" / "master" / "main.tf").write_text("")
    (d / "ansible").mkdir()
    (d / "ansible" / "hosts").write_text("")
    (d / "ansible" / ".cfg").write_text("")
    return str(d)


def bash_read(data: bytes):
    p = subprocess.run(["bash", "-c", 'read -p "q " v; s=$?; printf "%s|%s" "$s" "$v"'],
                       input=data, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, env=ENV)
    status, _, val = p.stdout.partition(b"|")
    return val.decode("utf-8", "surrogateescape"), status == b"0"


def our_read(data: bytes):
    it = iter(data)
    return bashsem.read_builtin(lambda: next(it, None))


READ_CASES = [
    b"", b"\n", b"abc\n", b"abc", b"  a  b  \n", b"\tx\t\n", b"a\\ b\n", b"a\\\nb\n", b"a\\\\b\n",
    b"trailing\\", b"\\", b"\\\n", b"\\ \n", b"  \\  \n", b"a b \\ \n", b'my "quoted" desc\n',
    b"-n foo\n", b"*\n", b"a\\tb\n", b"x\\", b"caf\xc3\xa9\n", b"\xff\xfe\n", b"a\x00b\n",
    b"line1\nline2\n", b"\\\\\n", b"\\\\\\\n", b"a\\\n\\\nb\n",
]


@pytest.mark.parametrize("data", READ_CASES)
def test_read_cases(data):
    assert our_read(data) == bash_read(data)


@settings(max_examples=150, deadline=None, derandomize=True)
@given(st.binary(max_size=12).map(lambda b: b.replace(b"\x01", b"").replace(b"\x7f", b"")))
def test_read_random(data):
    assert our_read(data) == bash_read(data)


def bash_echo_unquoted(values, cwd):
    """`echo $v` for many values in one bash; outputs separated by \\x1e."""
    script = 'for v in "$@"; do echo $v; printf "\\036"; done'
    p = subprocess.run(["bash", "-c", script, "_"] + values, cwd=cwd, stdout=subprocess.PIPE,
                       stderr=subprocess.DEVNULL, env=ENV)
    outs = p.stdout.split(b"\x1e")[:-1]
    return [o.decode("utf-8", "surrogateescape") for o in outs]


ECHO_CASES = [
    "k8s dev", "  a   b  ", "*", "a*", "?", "??", "[ab]*", "[!a]*", "[z]", "\\[z]", "[", "]", "a[", "*.md",
    "terraform/*", "*/*", "*/*/*", ".*", "ansible/.*", "-n foo", "-e a\\nb", "-e a\\tb", "-E x", "-ne x\\c y",
    "-n", "-e", "- n", "-x", "-nx y", "x -n", 'my "quoted" desc', "a\\b", "a\\*", "\\*", "s*r", "s\\*r", "x*y",
    "-e \\0101\\x41\\101", "-e \\q\\\"", "-e tail\\", "nomatch*zzz", "[a-c]", "[a-c]?", "[[:alpha:]]", "[[:digit:]]*",
    "x?y", "terraform/", "terraform/m*/", "*/", "/dev/nul?", "//", "a//b", "~", "$HOME", "`id`", "a;b", "a|b",
    "tab\there", "nl\nhere", "[]]", "[!]]*", "[a-]", "[]-a]", "*\\", "-e \\x", "-e \\xZ", "-e \\1", "-e \\08", "-e a\\cb",
    "-e \\e[0m", "-", "--", "-nn", "-nE -e \\n",
]


def test_echo_unquoted_cases(globdir):
    got = [bashsem.echo_unquoted(v, globdir) for v in ECHO_CASES]
    want = bash_echo_unquoted(ECHO_CASES, globdir)
    for v, g, w in zip(ECHO_CASES, got, want):
        assert g == w, "value %r: ours %r bash %r" % (v, g, w)


_alphabet = st.sampled_from(list(" \t*?[]!-abnexyz12./\\\"'$~;="))


@settings(max_examples=60, deadline=None, derandomize=True)
@given(st.lists(st.text(_alphabet, max_size=8), min_size=1, max_size=25))
def test_echo_unquoted_random(globdir, values):
    got = [bashsem.echo_unquoted(v, globdir) for v in values]
    want = bash_echo_unquoted(values, globdir)
    for v, g, w in zip(values, got, want):
        assert g == w, "value %r: ours %r bash %r" % (v, g, w)


def bash_match(pairs):
    script = 'while [ $# -gt 0 ]; do if [[ $1 == $2 ]]; then printf y; else printf n; fi; shift 2; done'
    flat = [x for p in pairs for x in p]
    p = subprocess.run(["bash", "-c", script, "_"] + flat, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, env=ENV)
    return [c == "y" for c in p.stdout.decode()]


MATCH_CASES = [
    ("abc", "abc"), ("abc", "a*"), ("abc", "a?c"), ("a/b", "a*b"), ("a/b", "a?b"), ("", ""), ("", "*"),
    ("31428241-4878", "31428241-4878"), ("x", "[x]"), ("x", "[!x]"), ("[", "["), ("a]", "a]"), ("a*", "a\\*"),
    ("ab", "a\\*"), (".x", "*"), ("a b", "a b"), ("a", "[a-"), ("-", "[a-]"), ("\\", "\\\\"), ("\\", "\\"),
    ("a\\", "a\\"), ("k4", "k[[:digit:]]"), ("uuid1,uuid2", "uuid1,uuid2"),
]


def test_pattern_match_cases():
    got = [bashsem.pattern_match(s, p) for s, p in MATCH_CASES]
    assert got == bash_match(MATCH_CASES), list(zip(MATCH_CASES, got, bash_match(MATCH_CASES)))


@settings(max_examples=60, deadline=None, derandomize=True)
@given(st.lists(st.tuples(st.text(st.sampled_from(list("ab*?[]!-\\/.")), max_size=5),
                          st.text(st.sampled_from(list("ab*?[]!-\\/.")), max_size=5)), min_size=1, max_size=30))
def test_pattern_match_random(pairs):
    got = [bashsem.pattern_match(s, p) for s, p in pairs]
    want = bash_match(pairs)
    assert got == want, [(pr, g, w) for pr, g, w in zip(pairs, got, want) if g != w]


def test_expand_word(globdir):
    cases = [
        ("A", ["http://10.0.0.1:8080/x"]),
        ("A\nB", None), (" A ", None), ("", None), ("  ", None), ("A B\n", None), ("*", None),
    ]
    for val, _ in cases:
        ours = bashsem.expand_word([("http://", False), (val, True), (":8080/x", False)], globdir)
        p = subprocess.run(["bash", "-c", 'printf "%s\\036" http://$1:8080/x', "_", val], cwd=globdir,
                           stdout=subprocess.PIPE, env=ENV)
        want = p.stdout.decode().split("\x1e")[:-1]
        if want == [""]:
            want = []  # printf with zero args still prints the format once
        assert ours == want, (val, ours, want)


def test_cmdsubst():
    assert bashsem.cmdsubst("a\n\n") == "a"
    assert bashsem.cmdsubst("a\nb\n") == "a\nb"
    assert bashsem.cmdsubst("\n") == ""


def test_echo_e_escapes_in_both_locales():
    """\\uHHHH / \\UHHHHHHHH are written in the locale's character set: UTF-8 where that is
    the locale's, otherwise only ASCII can be and the rest comes back as the escape it was."""
    import random
    import subprocess
    from tritonk8ssupervisor_b200 import bashsem as sh
    r = random.Random(3)
    tokens = ["\\u", "\\U", "\\x", "\\0", "\\", "\\\\", "4", "1", "e", "9", "F", "0", "Z", " ", "a", "\\n", "\\c", "\\t", "8", "d", "D8", "00"]
    cases = ["\\u00e9", "\\u65e5x", "\\U0001F600", "\\u41", "\\uZZ", "\\u", "\\ud800", "\\U00110000", "\\U7FFFFFFF", "\\UFFFFFFFF",
             "\\U00e9\\u7f\\u80"] + ["".join(r.choice(tokens) for _ in range(r.randint(1, 8))) for _ in range(150)]
    for env, utf8 in (({"LC_ALL": "C"}, False), ({"LANG": "C.UTF-8"}, True)):
        for text in cases:
            real = subprocess.run(["bash", "-c", 'echo -e "$1"', "_", text], capture_output=True, env=dict(env, PATH="/usr/bin:/bin")).stdout
            assert sh.to_bytes(sh.echo_builtin(["-e", text], utf8)) == real, (text, utf8)


def test_declare_value_in_both_locales():
    """`export -p` prints "value" -- or $'...' once the value holds something the locale cannot print."""
    import random
    import subprocess
    from tritonk8ssupervisor_b200 import bashsem as sh
    r = random.Random(1)
    chars = list("ab \"$`\\'!*") + ["\t", "\n", "\x01", "\x7f", "\x1b", "\r", "é", "日", "\udcff", "\udc80", "​", "", "͸",
                                    " ", "\xa0", "\xad", "\U0001f600", "\U000e0001", "￾", "̀"]
    for env, utf8 in (({"LC_ALL": "C"}, False), ({"LANG": "C.UTF-8"}, True)):
        for _ in range(150):
            v = "".join(r.choice(chars) for _ in range(r.randint(0, 6)))
            out = subprocess.run(["bash", "-c", 'V=$1; export V; export -p | grep -a "^declare -x V="', "_", v], capture_output=True,
                                 env=dict(env, PATH="/usr/bin:/bin")).stdout.decode("utf-8", "surrogateescape")
            assert "declare -x V=" + sh.declare_value(v, utf8) == out.split("\n")[0], (ascii(v), utf8)


def test_echo_e_escapes_go_by_the_name_of_a_locale_that_is_not_installed():
    """LC_ALL=xx.UTF-8 cannot be set, bash runs in C -- and still writes \\u00e9 as UTF-8: where wctomb() gives up it
    asks iconv for the character set named after the dot."""
    import subprocess
    from tritonk8ssupervisor_b200 import bashsem as sh
    text = "\\u00e9|\\u20ac|\\u044f|\\u65e5|\\ud800|\\U0001F600|\\U00110000|\\u41"
    for name in ["xx", "xx.UTF-8", "xx.ISO-8859-1", "UTF-8", "en_US", "xx.foo", "xx.utf8@euro", "xx.LATIN1", "xx.KOI8-R", "xx.CP1252",
                 "xx.ASCII", "xx.utf-8", "xx.", ".UTF-8", "xx.UTF8.x"]:
        real = subprocess.run(["bash", "-c", 'echo -e "$1"', "_", text], capture_output=True, env={"LC_ALL": name, "PATH": "/usr/bin:/bin"}).stdout
        assert sh.to_bytes(sh.echo_builtin(["-e", text], sh.locale_charset(name))) == real, name
