"""shmatch against the real bash: ``[[ $s == $p ]]`` (extended operators on) and
pathname expansion of an unquoted word in a directory of awkward names."""
import os
import random
import subprocess

import pytest

from tritonk8ssupervisor_b200 import bashsem, shmatch

ENV = {"LC_ALL": "C", "PATH": "/usr/bin:/bin"}


def _q(text):
    return "$'" + "".join("\\x%02x" % b for b in text.encode("utf-8", "surrogateescape")) + "'" if text else "''"


def bash_matches(pairs, tmp_path):
    script = ["t() { if [[ $1 == $2 ]]; then echo 1; else echo 0; fi; }"] + ["t %s %s" % (_q(s), _q(p)) for s, p in pairs]
    path = tmp_path / "m.sh"
    path.write_text("\n".join(script) + "\n")
    out = subprocess.run(["bash", str(path)], capture_output=True, env=ENV, timeout=120).stdout.decode().split()
    assert len(out) == len(pairs)
    return [o == "1" for o in out]


CORPUS = [
    ("abc", "abc"), ("abc", "a*c"), ("abc", "a?c"), ("abc", "a[b]c"), ("abc", "a[!b]c"), ("a]c", "a[]]c"), ("a-c", "a[a-]c"),
    # a '[' that never closes is a character ... unless the scan trips first
    ("[a", "[a"), ("[a-", "[a-"), ("[-kvm-", "[-kvm-"), ("[-a", "[-a"), ("x[a-", "x[a-"), ("[", "["), ("[]", "[]"), ("[!", "[!"),
    ("[a]", "[a]"), ("a]", "a]"), ("[[:alpha:", "[[:alpha:"), ("[[.", "[[."), ("[a\\", "[a\\"), ("a^", "[@([:alpha:]*"),
    ("a", "[[=a=]"), ("ab", "[[=a=]b"), ("a]", "[[:alpha:]]"), ("-", "[[.hyphen.]]"), ("b", "[a-[.c.]]"), ("b", "[c-a]"), ("c", "[c-ab]"),
    # extended operators
    ("xyz", "@(abc|x?z)"), ("abd", "@(abc|x?z)"), ("aaa", "+(a)"), ("", "+(a)"), ("", "*(a)"), ("ab", "?(a)b"), ("b", "?(a)b"),
    ("c", "!(a|b)"), ("a", "!(a|b)"), ("abab", "*(ab)"), ("aba", "*(ab)"), ("a(b", "a(b"), ("a", "@(a"), ("@(a", "@(a"),
    ("ab", "a*(b|c)"), ("ad", "a*(b|c)"), ("a", "a*!(x)"), ("ax", "a*!(x)"), ("a.b", "*.+([a-z])"), ("x", "@(x|)"), ("", "@(x|)"),
    ("a|b", "a\\|b"), ("a)", "@(a\\))"), ("[", "@([)"), ("]", "@([]])"),
    # backslashes
    ("*", "\\*"), ("a", "\\a"), ("\\", "\\\\"), ("a\\", "a\\"), ("a", "a\\"), ("*a*?\\", "*a*?\\"), ("]-(*\\", "]-(*\\"),
    ("31428241-4878-47d6-9fba-9a8436b596a4", "31428241-4878-47d6-9fba-9a8436b596a4"), ("zqz-9", "z?z-9"), ("qxyzq", "q*q"),
]


def test_corpus_matches_bash(tmp_path):
    want = bash_matches(CORPUS, tmp_path)
    got = [shmatch.strmatch(p, s, shmatch.EXTMATCH) for s, p in CORPUS]
    assert [(c, w, g) for c, w, g in zip(CORPUS, want, got) if w != g] == []


_P = list("ab[]-!^*?\\()|+@.:=") + ["[:alpha:]", "[a-", "[!a]", "[]]", "+(", "@(", "!(", "?(", "*(", "[.a.]", "[=a=]", "[.hyphen.]",
                                     "[a-c]", "[c-a]"]
_S = list("ab[]-!^*?\\()|+@.")


@pytest.mark.parametrize("seed", range(8))
def test_random_patterns_match_bash(tmp_path, seed):
    r = random.Random(seed)
    pairs = [("".join(r.choice(_S) for _ in range(r.randint(0, 5))), "".join(r.choice(_P) for _ in range(r.randint(0, 6))))
             for _ in range(300)]
    pairs += [(p.replace("\\", "").replace("*", "a"), p) for _, p in pairs[:150]] + [(s, s) for s, _ in pairs[:50]]
    want = bash_matches(pairs, tmp_path)
    got = [shmatch.strmatch(p, s, shmatch.EXTMATCH) for s, p in pairs]
    assert [(c, w, g) for c, w, g in zip(pairs, want, got) if w != g] == []


NAMES = ["a", "ab", "abc", ".hidden", ".a", "a.b", "[a", "[a-", "a]", "a*b", "a?b", "-n", "x y", "a\\b", "@(a)", "+(a)", "a|b", "B", "z[1]",
         "d1/in", "d1/.dot", "d2/in", "d2/x[", "Makefile", "é"]
WORDS = ["*", ".*", "a*", "[a]*", "[a-", "[a-*", "[a*", "*]", "a?b", "a\\*b", "a[*]b", "@(a)", "+(a)", "*(a)", "?", "??", "[!a]", "[!a]*",
         "*/in", "*/*", "d?/.*", "*/x[", "d[12]/in", "nomatch*", "[[:upper:]]*", "z\\[1]", "z[[]1]", "*[", "\\-n", "-*", "x*y", "a|b", "*|*",
         "[.hyphen.]n", "a[.]b", "*.b", ".[!.]*", "M*e", "[a-z]", "[B-a]", "d1/", "d*/", "*/", "é", "[é]"]


def test_pathname_expansion_matches_bash(tmp_path):
    root = tmp_path / "w"
    for n in NAMES:
        p = root / n
        p.parent.mkdir(parents=True, exist_ok=True)
        p.write_bytes(b"")
    script = ["cd %s" % _q(str(root))] + ["w=%s; printf '%%s\\036' $w; printf '\\n'" % _q(w) for w in WORDS]
    path = tmp_path / "g.sh"
    path.write_text("\n".join(script) + "\n")
    out = subprocess.run(["bash", str(path)], capture_output=True, env=ENV, timeout=60).stdout.decode("utf-8", "surrogateescape")
    want = [line.split("\x1e")[:-1] for line in out.split("\n")[:-1]]
    got = [bashsem.expand_unquoted(w, str(root)) for w in WORDS]
    assert [(w, a, b) for w, a, b in zip(WORDS, want, got) if a != b] == []


def test_quoted_text_is_never_a_pattern(tmp_path):
    (tmp_path / "h*me").mkdir()
    (tmp_path / "home").mkdir()
    for d in ("h*me", "home"):
        (tmp_path / d / "id_rsa").write_text("k")
    word = [(bashsem.quote_for_glob(str(tmp_path / "h*me") + "/"), False), ("id_*", True)]
    assert bashsem.expand_word(word, str(tmp_path)) == [str(tmp_path / "h*me" / "id_rsa")]
    assert os.path.exists(str(tmp_path / "home" / "id_rsa"))


def test_utf8_locale_matches_bash(tmp_path):
    """LANG=C.UTF-8: a multi-byte character is one character for ``?`` and for bracket
    expressions; with a byte that is no character anywhere, bash falls back on bytes."""
    r = random.Random(5)
    pch = list("ab[]-!*?\\()|+@") + ["é", "日", "\udcff", "[:alpha:]", "[:upper:]", "[:punct:]", "[a-é]", "[é-日]",
                                     "[!é]", "+(", "@(", "[:space:]", "[:print:]"]
    sch = list("ab[-*?") + ["é", "日", "\udcff", "É", "\xa0", "¿", " "]
    pairs = [("".join(r.choice(sch) for _ in range(r.randint(0, 4))), "".join(r.choice(pch) for _ in range(r.randint(0, 5))))
             for _ in range(1200)]
    pairs += [(s, s) for s, _ in pairs[:100]] + [("zéz", "z?z"), ("zéz", "z??z"), ("\udcffÉ", "[!é][!é]")]
    script = ["t() { if [[ $1 == $2 ]]; then echo 1; else echo 0; fi; }"] + ["t %s %s" % (_q(s), _q(p)) for s, p in pairs]
    path = tmp_path / "m8.sh"
    path.write_text("\n".join(script) + "\n")
    out = subprocess.run(["bash", str(path)], capture_output=True, env={"LANG": "C.UTF-8", "PATH": "/usr/bin:/bin"}, timeout=120).stdout.decode().split()
    assert len(out) == len(pairs)
    got = [shmatch.strmatch(p, s, shmatch.EXTMATCH | shmatch.UTF8) for s, p in pairs]
    assert [(ascii(c), w, g) for c, w, g in zip(pairs, [o == "1" for o in out], got) if w != g] == []
