"""basharith against the real bash: values, side effects on variables, exit
status and the exact diagnostics (``... (error token is "...")``), for a fixed
corpus and for random token soup.  bash is present here and on the box; each
expression is its own top-level ``let`` in a script file, because an error
inside an array subscript makes bash drop the whole top-level command.
"""
import os
import random
import re
import subprocess

import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from tritonk8ssupervisor_b200 import basharith as ba

VARS = {"x": "1 +", "y": "x", "z": "y * 2", "w": "w", "q": '"a"', "n": "5", "e": "", "sp": "  ", "m": "-3", "h": "0x10",
        "big": "9223372036854775807", "p": "n + 1", "a1": "b1", "b1": "a1"}
SHOWN = ("n", "x", "nn", "m")


def _sq(text):
    return "'" + text.replace("'", "'\\''") + "'"


def run_bash(exprs, tmp_path, errexit=False):
    """[(stdout, stderr)] per expression."""
    reset = ["unset nn zz " + " ".join(VARS)] + ["%s=%s" % (k, _sq(v)) for k, v in VARS.items()]
    script = []
    for i, e in enumerate(exprs):
        script += reset
        script.append("E=%s" % _sq(e))
        script.append('echo "@@ %d"; echo "@@ %d" >&2' % (i, i))
        if errexit:
            # a subshell per case: under errexit some diagnostics end the shell
            script.append('( set -e; if let "$E"; then echo "st=0 %s"; else echo "st=1 %s"; fi; exit 0 )'
                          % ((" ".join("%s=$%s" % (k, k) for k in SHOWN),) * 2))
            script.append('echo "sub=$?"')
        else:
            script.append('let "$E"')
            script.append('echo "st=$? %s"' % " ".join("%s=$%s" % (k, k) for k in SHOWN))
    path = tmp_path / "case.sh"
    path.write_bytes(("\n".join(script) + "\n").encode("utf-8", "surrogateescape"))
    p = subprocess.run(["bash", str(path)], capture_output=True, env={"LC_ALL": "C", "PATH": "/usr/bin:/bin"}, timeout=120)

    def split(raw):
        text = raw.decode("utf-8", "surrogateescape")
        parts = re.split(r"^@@ (\d+)\n", text, flags=re.M)
        return {int(parts[j]): parts[j + 1] for j in range(1, len(parts), 2)}
    out, err = split(p.stdout), split(p.stderr)
    label = re.compile(r"^%s: line \d+: " % re.escape(str(path)), re.M)
    return [(out.get(i, ""), label.sub("L: ", err.get(i, ""))) for i in range(len(exprs))]


def run_mine(expr, errexit=False):
    v = dict(VARS)
    ev = ba.Evaluator(v.get, v.__setitem__, errexit=errexit)

    def shown(st):
        return "st=%d %s\n" % (st, " ".join("%s=%s" % (k, v.get(k, "")) for k in SHOWN))
    try:
        val = ev.evaluate(expr)
        err = "".join("L: %s\n" % m for m in ev.messages)
        out = shown(1 if val == 0 else 0)
    except ba.ArithError as exc:
        err = "".join("L: %s\n" % m for m in ev.messages) + "L: let: %s\n" % exc.text
        out = shown(1)
    except ba.ArithAbort as exc:
        err = "".join("L: %s\n" % m for m in ev.messages) + "L: %s\n" % exc.text
        out = "" if errexit else shown(1)       # the command is dropped; the next top-level command still runs
    except ba.ArithFatal as exc:
        return "sub=1\n", "L: %s\n" % exc.text
    except ba.ArithRefused:
        return None
    return (out + "sub=0\n" if errexit and out else ("sub=1\n" if errexit else out)), err


CORPUS = [
    'abc', '', '08', '-15', '15 30', '0x1e', '1.5', '$x', 'a b', '"q"', '1+', '(', ')', '16#1e', '2#102', 'nn=3', '1/0', '1%0', 'n++',
    '++', '7 ?', '1 ? 2', 'n[1]', '99999999999999999999', '0x', '65#1', '1#0', '08#1', '3 **', '2 ** -1', '1,', ',1', '1 2', 'a=b=',
    '5=3', 'a+=', "'a'", '\\1', '1;2', '#', '@', '1 # 3', '_x', '  ', '1\n2', '!', '~', '-', '1 <', '1 <<', '1 &&', '1 ||', '1 |',
    '1 ^', '1 &', '1 ==', '1 !=', '1 =', '1 >=', '1 >>=', 'x y', 'x+y z', '3 x', '3x', '0x1g', '1e3', 'a.b', '1 ? 2 :', '(1', '1)',
    '()', 'n++ ++', '++n++', ' --1', '- -1', '++1', '1++', 'n--', '1 1', 'x', 'y', 'z', 'w', 'q', 'n[0]', 'n[0', 'n[1+]', 'n[x]', 'e',
    'sp', 'n[-1]', 'n[n]', 'n[ 0 ]', 'n[0]++', '++n[0]', 'n[0]=4', 'n[@]', 'n[*]', 'n[1]=2', 'p', 'p*p', 'a1', 'big+1', 'big*2',
    '-big-2', '1<<64', '1<<63', '-1>>70', '7/-2', '-7/2', '-7%3', '7%-3', '2**62*4', '2**64', '3**40', 'n+=2', 'n*=n', 'n/=0', 'n%=0',
    'n<<=2', 'n>>=1', 'n&=4', 'n|=8', 'n^=1', 'n=', 'n==5', 'n!=5', 'n<=5', 'n>=6', '!n', '~n', 'n?1:2', '0?1/0:3', '1?2:1/0',
    '0&&1/0', '1||1/0', '0||1/0', '1&&1/0', '(n)', '((n))', '(n', 'n)', '1,2,3', 'nn=n=7', 'nn=1,nn+1', 'n = 5 = 3', '++n', ' --n',
    '++ n', ' -- n', '++nn', 'n ++', 'n++ + ++n', 'n+++n', 'n---n', 'n++++', '++--n', '!!n', '-~n', '64#@_', '64#a', '36#Z', '37#Z',
    '10#08', '8#8', '2#', '2#1#1', '0x#1', '1#', '010', '0X1F', '00', '0_1', '1_000', 'n[0][0]', 'h', 'h+h', 'm*m', ' -m',
    '1 ? 2 : 3 : 4', '1 ? : 2', '? 1 : 2', ': ', '1 :', 'a ? b', '= 1', '== 1', '* 2', '/ 2', 'n /', 'n %', 'n **', 'n ** 2 ** 0',
    '2 ** 3 ** 2', ' -2 ** 2', 'n n', 'n 1', '1 n', 'n (1)', '(1) n', 'n[0] n', '1 ( 2', 'n,', 'n,,1', 'n=1,', 'é', '1 é',
    'n é', '\x05', '1 \x05', 'n\x051', 'n$', '1`', 'n{', 'n}', '1]', '[1]', 'n]', '1;', 'n;', 'n[0] -- < n[n-5]',
    'n[n-5] |= n[0]   ++', 'm[1]=7, m[-1]', 'm[2]=7, m[-3]', 'm[2]=7, m[-4]', '-9223372036854775808 / -1', 'big + 1 % -1',
    '(-big-1) % -1', '(-big-1) / -1', '1 ? nn=4 : 5', '0 ? nn=4 : (nn=6)', '0 && (nn=1)', '1 || nn++',
]


def test_corpus_matches_bash(tmp_path):
    got = run_bash(CORPUS, tmp_path)
    bad = [(e, b, run_mine(e)) for e, b in zip(CORPUS, got) if run_mine(e) is not None and run_mine(e) != b]
    assert bad == []


def test_corpus_matches_bash_under_errexit(tmp_path):
    """`set -e` (setup.sh:3) makes "bad array subscript" fatal; nothing else changes."""
    got = run_bash(CORPUS, tmp_path, errexit=True)
    bad = [(e, b, run_mine(e, True)) for e, b in zip(CORPUS, got) if run_mine(e, True) is not None and run_mine(e, True) != b]
    assert bad == []


TOKENS = ["0", "1", "2", "7", "08", "15", "0x1f", "2#11", "64#_@", "10#9", "n", "nn", "x", "y", "p", "m", "h", "e", "sp", "w", "q", "big",
          "zz", "n[0]", "n[1]", "n[n-5]", "nn[0]", "+", "-", "*", "/", "%", "**", "++", "--", "!", "~", "(", ")", "<<", ">>", "<", ">", "<=",
          ">=", "==", "!=", "&", "|", "^", "&&", "||", "?", ":", "=", "+=", "-=", "*=", "/=", "%=", "<<=", ">>=", "&=", "|=", "^=", ",",
          " ", " ", "  ", ".", "#", "@", "[", "]", "$", "'", '"', ";", "\\", "\t", "\n", "é", "9223372036854775807",
          "-9223372036854775808", "63", "64", "65"]


def _soup(rnd):
    sep = rnd.choice(["", " ", " "])
    e = sep.join(rnd.choice(TOKENS) for _ in range(rnd.randint(1, 7)))
    return " " + e if e.lstrip().startswith("-") else e       # a leading dash would be an option of `let`


@pytest.mark.parametrize("seed", range(6))
def test_token_soup_matches_bash(tmp_path, seed):
    rnd = random.Random(seed)
    exprs = [_soup(rnd) for _ in range(400)]
    got = run_bash(exprs, tmp_path)
    bad = [(e, b, run_mine(e)) for e, b in zip(exprs, got) if run_mine(e) is not None and run_mine(e) != b]
    assert bad == []


@settings(max_examples=30, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
@given(st.lists(st.lists(st.sampled_from(TOKENS), min_size=1, max_size=9).map(" ".join), min_size=20, max_size=20))
def test_hypothesis_matches_bash(tmp_path_factory, exprs):
    exprs = [" " + e if e.lstrip().startswith("-") else e for e in exprs]
    got = run_bash(exprs, tmp_path_factory.mktemp("arith"))
    bad = [(e, b, run_mine(e)) for e, b in zip(exprs, got) if run_mine(e) is not None and run_mine(e) != b]
    assert bad == []


def test_expansions_inside_subscripts_are_refused():
    """bash would run word expansion -- `$(...)` included -- on the subscript."""
    for expr in ["n[$(id)]", "n[`id`]", "n[$x]", "n[\\1]"]:
        with pytest.raises(ba.ArithRefused):
            ba.evaluate(expr, dict(VARS))


def test_recursion_limit_is_bashs():
    chain = {"v%d" % i: "v%d" % (i + 1) for i in range(1100)}
    chain["v1100"] = "7"
    with pytest.raises(ba.ArithError, match="expression recursion level exceeded"):
        ba.evaluate("v0", chain)
    assert ba.evaluate("v100", chain) == 7          # 1001 levels deep: fine


def test_the_three_sites_on_the_path():
    """setup.sh:64, :148 and :340 as the supervisor phrases them."""
    v = {"i": "3", "KUBERNETES_NUMBER_OF_NODES": "2"}
    assert ba.evaluate("1500000000 % 15", v) == 0
    assert ba.evaluate("-15 % 15", v) == 0
    assert ba.evaluate("abc % 15", v) == 0
    assert ba.evaluate("i <= KUBERNETES_NUMBER_OF_NODES", v) == 0
    with pytest.raises(ba.ArithError) as exc:
        ba.evaluate("08 % 15", v)
    assert exc.value.text == '08: value too great for base (error token is "08")'
    with pytest.raises(ba.ArithError) as exc:
        ba.evaluate(" % 15", v)
    assert exc.value.text == '% 15: syntax error: operand expected (error token is "% 15")'
    with pytest.raises(ba.ArithError) as exc:
        ba.evaluate("15 30 % 15", v)
    assert exc.value.text == '15 30 % 15: syntax error in expression (error token is "30 % 15")'
    assert os.path.exists("/usr/bin/bash") or os.path.exists("/bin/bash")
