"""`triton env` output is parsed, not eval-ed (deviation 1).  For
the accepted dialect the parser must leave exactly the variables that bash's
``eval`` leaves; everything else must be refused."""
import subprocess

import pytest

from tritonk8ssupervisor_b200.conformance.scenario import triton_env_text
from tritonk8ssupervisor_b200.tritonenv import TritonEnvError, apply_triton_env

NAMES = ["SDC_URL", "SDC_ACCOUNT", "SDC_KEY_ID", "SDC_USER", "SDC_TESTING", "TRITON_PROFILE", "DOCKER_HOST", "PLAIN", "EMPTY", "A", "B"]
START = {"SDC_USER": "old-user", "SDC_TESTING": "1", "PLAIN": "exported-before", "HOME": "/home/op er*", "GLOBBY": "a  b * c"}


def bash_eval(text):
    """(shell value, exported?) per name after `eval "$text"` in a real bash."""
    script = 'eval "$1"; shift; for n in "$@"; do if [ -n "${!n+x}" ]; then printf "%s=%s|%s\\036" "$n" "${!n}" ' \
             '"$(env | grep -c "^$n=")"; fi; done'
    env = dict(START, PATH="/usr/bin:/bin", LC_ALL="C")
    p = subprocess.run(["bash", "-c", script, "_", text] + NAMES, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env)
    out = {}
    for rec in p.stdout.decode().split("\x1e")[:-1]:
        kv, _, exported = rec.rpartition("|")
        k, _, v = kv.partition("=")
        out[k] = (v, exported != "0")
    return out


def ours(text):
    shell, env = dict(START), dict(START)
    apply_triton_env(text, shell, env)
    return {n: (shell[n], n in env) for n in NAMES if n in shell}


ACCEPTED = [
    triton_env_text(),
    triton_env_text(url="https://eu-ams-1.api.joyent.com", account="acme corp", key_id="SHA256:abc/def+ghi="),
    'export SDC_URL=https://bare.example\nexport SDC_ACCOUNT="q\\"uote"\n',
    "export SDC_URL='single $HOME `x`'\n",
    'export SDC_URL="a" SDC_ACCOUNT="b"; export SDC_KEY_ID=c\n',
    "PLAIN=changed\nEMPTY=\nexport EMPTY\n",
    "unset SDC_USER SDC_TESTING\n# comment\n\n   \n",
    'export DOCKER_HOST=tcp://1.2.3.4:2376 # trailing comment\n',
    'export SDC_URL="with#hash" SDC_ACCOUNT=a#b\n',
    'export SDC_URL=back\\ slash\\"x\n',
    "export SDC_ACCOUNT\n",
    "",
    # parameter references execute nothing: expanded as bash expands an assignment
    # (no field splitting, no pathname expansion), earlier assignments visible to later ones
    'export SDC_URL="https://$HOME/x"\n',
    'export SDC_URL="${HOME}y$SDC_USER" SDC_ACCOUNT=$SDC_USER$UNSET_THING.\n',
    "export SDC_URL=$GLOBBY\nexport SDC_ACCOUNT=\"$GLOBBY\"\nPLAIN=$GLOBBY\n",
    'A=one\nB="$A two"\nexport SDC_URL="$B $A"\nunset A\nexport SDC_ACCOUNT="[$A]"\n',
    'export SDC_URL="cost: 5$ and $ and $. and a$"\nexport SDC_ACCOUNT=x$\n',
    'export SDC_URL="\\$HOME" SDC_ACCOUNT=\\$HOME SDC_KEY_ID=\'$HOME\'\n',
    # tilde: at the start of an assigned value and after an unquoted colon, nowhere else
    "export SDC_URL=~/x\nexport SDC_ACCOUNT=~\nexport SDC_KEY_ID=a~/b\n",
    "export SDC_URL=~/a:~/b:c~:~\nPLAIN=~/p\n",
    'export SDC_URL="~/quoted" SDC_ACCOUNT=\\~/escaped SDC_KEY_ID=~"/half"\n',
    "export SDC_URL=~root/x SDC_ACCOUNT=~nosuchuser_zz/x SDC_KEY_ID=~ro\"ot\"/x\n",
    "export SDC_URL=x=~/y\nexport SDC_ACCOUNT=$HOME:~\n",
]


@pytest.mark.parametrize("text", ACCEPTED, ids=range(len(ACCEPTED)))
def test_same_variables_as_bash_eval(text):
    assert ours(text) == bash_eval(text)


@pytest.mark.parametrize("text", [
    'export SDC_URL="$(curl evil)"\n', "export SDC_URL=`id`\n", 'export SDC_URL="`id`"\n', "rm -rf /\n",
    "export SDC_URL=a | tee x\n", "export 1BAD=x\n", "export SDC_URL=$(id)\n", "echo hi > /tmp/x\n", "export A=(1 2)\n",
    'export SDC_URL="unterminated\n', "export SDC_URL='unterminated\n", "eval foo\n", "A=1 B=2\n",
    'export SDC_URL="$((1+1))"\n', 'export SDC_URL="${HOME:-x}"\n', 'export SDC_URL="${HOME"\n', "export SDC_URL=$1\n",
    'export SDC_URL="$?"\n', "export SDC_URL=$'ansi'\n", "unset $HOME\n", "$HOME\n", "export $HOME=x\n", "export SDC_URL=$_\n",
    "export SDC_URL=~+/x\n", "export SDC_URL=${!HOME}\n", "export SDC_URL=$$\n",
    "export\n", "export IFS=:\n", "IFS=,\n", "unset IFS\n", "export LC_ALL=en_US.UTF-8\n", "GLOBIGNORE=*\n",
])
def test_everything_else_is_refused(text):
    with pytest.raises(TritonEnvError):
        apply_triton_env(text, {}, {})


def test_refusal_end_to_end(tmp_path):
    """A `triton env` that prints a command: the reference would run it; the
    replacement stops with status 2 before anything is generated."""
    from tritonk8ssupervisor_b200.conformance import replay as rp
    from tritonk8ssupervisor_b200.conformance.harness import WORKSPACE_TEMPLATE
    from tritonk8ssupervisor_b200.conformance.sandbox import run_in_sandbox
    from tritonk8ssupervisor_b200.conformance.scenario import Scenario
    marker = tmp_path / "pwned"
    sc = Scenario(name="evil", stdin="\n" * 8 + "yes\n", triton_env='export SDC_URL="x"\ntouch %s\n' % marker)
    obs = run_in_sandbox(rp.REPLACEMENT_ARGV, WORKSPACE_TEMPLATE, sc, extra_env=rp.REPLACEMENT_ENV)
    assert obs.exit_status == 2 and "refusing to run a command" in obs.stderr
    assert not marker.exists()
    assert "terraform/rancher.tf" not in obs.files
    assert [c.split("|")[0] for c in obs.calls if not c.startswith("  env")] == ["triton"]


# --- random texts of the accepted dialect: whatever is not refused must leave what bash's eval leaves ----
_VAL = ["x", "y z", "$A", "${B}", "$PLAIN", "$HOME", "~", "~/", ":", "/p", "a=b", "é", "*", "?", "[", "~root", "-", "$", "\\ ", "#h",
        "'q $A'", '"d $A ~"', '"\\$A"', "\\$A", "=", "+", "{a,b}", "$[1]"]


def _random_text(r):
    out = []
    for _ in range(r.randint(1, 4)):
        k = r.random()
        name = r.choice(["A", "B", "SDC_URL", "PLAIN", "EMPTY"])
        val = "".join(r.choice(_VAL) for _ in range(r.randint(0, 4)))
        if k < 0.5:
            out.append("export %s=%s" % (name, val))
        elif k < 0.7:
            out.append("%s=%s" % (name, val))
        elif k < 0.8:
            out.append("unset %s" % name)
        elif k < 0.9:
            out.append("export %s %s=%s" % (name, r.choice(["A", "B"]), val))
        else:
            out.append("# " + val)
        out.append(r.choice(["\n", ";", " ;\n", "\n\n", " # c\n"]))
    return "".join(out)


@pytest.mark.parametrize("seed", range(4))
def test_random_accepted_texts_match_bash_eval(seed):
    import random
    r = random.Random(seed)
    accepted = 0
    for _ in range(120):
        text = _random_text(r)
        try:
            mine = ours(text)
        except TritonEnvError:
            continue
        accepted += 1
        assert mine == bash_eval(text), text
    assert accepted > 40


@pytest.mark.parametrize("text", ["export A={a,b}\n", "export A=x{1..3}\n", "export A=$[1+1]\n", "unset {A,B}\n"])
def test_expansions_of_command_words_are_refused(text):
    with pytest.raises(TritonEnvError):
        apply_triton_env(text, {}, {})


def test_plain_assignment_has_no_brace_expansion():
    shell, env = {}, {}
    apply_triton_env("A={a,b}\n", shell, env)
    assert shell == {"A": "{a,b}"} and env == {}


def test_a_command_that_does_not_exist_gets_bashs_answer():
    """Nothing would run, so nothing needs refusing: `exp\\xffort X=1` is `command not found`,
    labelled with the line of the eval text it stands on (the eval is at setup.sh:210)."""
    from tritonk8ssupervisor_b200.tritonenv import CommandNotFound, ansic_quote
    with pytest.raises(CommandNotFound) as exc:
        apply_triton_env('export A="1"\n\nnosuchcommand_zz --flag\n', {}, {}, exists=lambda name: False)
    assert str(exc.value) == "nosuchcommand_zz: command not found" and exc.value.line == 2
    assert ansic_quote("exp\udcffort") == "$'exp\\377ort'" and ansic_quote("unse\rt") == "$'unse\\rt'"
    # a command that exists, a builtin, a function of the script, a path: refused, never looked at further
    for text in ("ls\n", "echo hi\n", "cleanRunner\n", "./x\n", "eval x\n"):
        with pytest.raises(TritonEnvError) as exc:
            apply_triton_env(text, {}, {}, exists=lambda name: name == "ls")
        assert not isinstance(exc.value, CommandNotFound), text


def _bash_says(text):
    p = subprocess.run(["bash", "-c", 'set -e; eval "$1"; echo survived', "_", text], capture_output=True,
                       env=dict(START, PATH="/usr/bin:/bin", LC_ALL="C"))
    return p.returncode, [line.split(": ", 2)[2] if line.startswith("_: line") else line for line in p.stderr.decode().splitlines()]


@pytest.mark.parametrize("text", ["export UID=5\n", "UID=5\n", "export A=1 UID=5 EUID=6 B=2\n", "unset UID\n", "unset A PPID B\n",
                                  "export SHELLOPTS=x\n", "BASHOPTS=y\n", "unset 1x UID\n", "export 1x UID=3 2y\n"])
def test_read_only_variables_get_bashs_answer(text):
    """Not refusals: bash answers these itself (status 1), running nothing."""
    from tritonk8ssupervisor_b200.tritonenv import BashDiagnostic
    status, said = _bash_says(text)
    assert status == 1
    with pytest.raises(BashDiagnostic) as exc:
        apply_triton_env(text, dict(START), dict(START))
    assert exc.value.messages() == said


@pytest.mark.parametrize("text", ["export RANDOM=3\n", "SECONDS=1\n", "export LINENO\n", "unset RANDOM\n", "export A=$RANDOM\n", "A=$HOSTNAME\n",
                                  "export HOSTNAME\n", "export BASH_VERSION\n", "export -p\n", "export -n A\n", "unset -v A\n", "unset -f A\n",
                                  "export A=${BASH_VERSION}\n", "export FUNCNAME\n", "export PS4\n", "OPTIND=1+1\n", "export OPTIND=x\n"])
def test_what_bash_computes_by_itself_is_refused(text):
    from tritonk8ssupervisor_b200.tritonenv import BashDiagnostic
    with pytest.raises(TritonEnvError) as exc:
        apply_triton_env(text, dict(START), dict(START))
    assert not isinstance(exc.value, BashDiagnostic) and "refusing" in str(exc.value)


def test_variables_bash_provides_expand_to_their_values():
    import os
    shell, env = dict(START), dict(START)
    provided = {"UID": str(os.getuid()), "IFS": " \t\n", "OPTIND": "1"}.get
    apply_triton_env('export A="$UID/$OPTIND" B=$NOSUCH\nexport UID\nHOSTNAME=h\nexport HOSTNAME\n', shell, env, provided=provided)
    assert env["A"] == "%d/1" % os.getuid() and env["B"] == "" and env["UID"] == str(os.getuid()) and env["HOSTNAME"] == "h"
    out = subprocess.run(["bash", "-c", 'eval "$1"; echo "$A|$B"; env | grep -c "^UID="', "_", 'export A="$UID/$OPTIND" B=$NOSUCH\nexport UID\n'],
                         capture_output=True, env=dict(START, PATH="/usr/bin:/bin")).stdout.decode().split()
    assert out == ["%d/1|" % os.getuid(), "1"]
