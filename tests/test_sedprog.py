"""sedprog restates what GNU sed does with the key path spliced into its program
text (setup.sh:133); checked here against the real sed."""
import os
import subprocess

import pytest

from tritonk8ssupervisor_b200 import sedprog

CFG = "[defaults]\ndeprecation_warnings = False\nPrivate_Key_File = Z\nprivate_key_file = old\n# private_key_file = x private_key_file = y\n"
VALUES = ["~/.ssh/id_rsa", "", "a&b", "a\\&b", "a\\\\b", "a\\nb", "a\\0b", "a\\;b", "x/y z", "a;b", "a;", "a;g", "a;p", "a;2", "a;1",
          "a;I", "a;i", "a; ", "a;;", "a;w", "a;w out.txt", "a;}", "a;#c", "a;gx", ";", "a;b;c", "a;gg", "a;pp", "a;0", "a;2g",
          "a;g;p", "a;p;d", "a;;d", "a;;x", "a;;q", "a;;h;g", "a;;G", "a\\1b", "a;e", "a;M", "a;;gg", "a;;p p", "a;3;p"]


def real_sed(value, text, tmp_path):
    prog = "s;private_key_file = .*$;private_key_file = %s;g" % value
    p = subprocess.run(["sed", prog], input=text.encode(), cwd=str(tmp_path), capture_output=True,
                       env={"PATH": "/usr/bin:/bin", "LC_ALL": "C"})
    return p.returncode, p.stdout.decode(), p.stderr.decode()


@pytest.mark.parametrize("value", VALUES)
@pytest.mark.parametrize("text", [CFG, CFG[:-1], "", "private_key_file = only"], ids=["cfg", "no-final-newline", "empty", "one-line"])
def test_against_real_sed(tmp_path, value, text):
    if value == "a;e":
        pytest.skip("the e flag would execute the pattern space; deliberately refused")
    rc, out, err = real_sed(value, text, tmp_path)
    res = sedprog.run(text, value, ";")
    assert res.ok == (rc == 0), (value, err, res.error)
    if rc == 0:
        assert res.output == out
        written = {f: open(os.path.join(str(tmp_path), f)).read() for f in os.listdir(str(tmp_path))}
        assert res.wfiles == written
    else:
        assert res.error == err


def test_e_flag_is_refused_not_run():
    res = sedprog.run(CFG, "a;e", ";")
    assert not res.ok and "e flag" in res.refused and res.error == ""


# no 'e'/'E' in the alphabet (the e flag, the e command), no '/' (files stay in the scratch directory)
_CH = list("ab;&\\gpwiIdDhHxnNqP{}0123 #~!$=") + ["\\n", "\\&", ";;", "w o", "\\;", "2g", "gp", "pw x", "\\\\", "\n", "\\x41", "\\t"]


@pytest.mark.parametrize("seed", range(4))
def test_random_values_against_real_sed(tmp_path, seed):
    """Whatever the module does not refuse, it must do exactly as sed does: output, files written, messages."""
    import random
    r = random.Random(seed)
    checked = 0
    for _ in range(150):
        value = "".join(r.choice(_CH) for _ in range(r.randint(0, 6)))
        text = r.choice([CFG, CFG[:-1], "", "private_key_file = only", "x\nprivate_key_file = q\n"])
        res = sedprog.run(text, value, ";")
        if res.refused:
            continue
        for f in os.listdir(str(tmp_path)):
            os.remove(os.path.join(str(tmp_path), f))
        rc, out, err = real_sed(value, text, tmp_path)
        checked += 1
        assert res.ok == (rc == 0), (value, err, res.error)
        if rc == 0:
            assert res.output == out, value
            assert res.wfiles == {f: open(os.path.join(str(tmp_path), f)).read() for f in os.listdir(str(tmp_path))}, value
        else:
            assert res.error == err, value
    assert checked > 60


def test_clean_form():
    """setup.sh:511: delimiter ~ and an empty value."""
    res = sedprog.run(CFG, "", "~")
    assert res.ok and "private_key_file = \n" in res.output and "old" not in res.output


def test_unrestated_program_is_refused_end_to_end():
    """A key file called ``k;;=``: the reference's sed would go on with the ``=`` command (print
    line numbers into ansible.cfg); the replacement stops with status 2 and says why."""
    from tritonk8ssupervisor_b200.conformance import replay as rp
    from tritonk8ssupervisor_b200.conformance.harness import WORKSPACE_TEMPLATE
    from tritonk8ssupervisor_b200.conformance.sandbox import run_in_sandbox
    from tritonk8ssupervisor_b200.conformance.scenario import DEFAULT_KEY_ID, Scenario
    sc = Scenario(name="sedmore", stdin="\n" * 8 + "yes\n", ssh_files={"k;;=": "K\n"},
                  keygen={"k;;=": "2048 MD5:%s c (RSA)\n" % DEFAULT_KEY_ID}, keygen_sha={})
    obs = run_in_sandbox(rp.REPLACEMENT_ARGV, WORKSPACE_TEMPLATE, sc, extra_env=rp.REPLACEMENT_ENV)
    assert obs.exit_status == 2
    assert "refusing to go on: the key path spliced into the sed program at line 133" in obs.stderr
    assert "ansible/ansible.cfg" not in obs.files          # untouched


def test_index_word_that_makes_sed_execute_is_refused():
    """An answer that leaves ``-ee 1`` in the variable: `sed -n "-ee"p` is `sed -n -e ep`, and GNU sed's `e`
    command runs ``p`` through sh once per row.  Status 2, no terraform input generated."""
    from tritonk8ssupervisor_b200.conformance import replay as rp
    from tritonk8ssupervisor_b200.conformance.harness import WORKSPACE_TEMPLATE
    from tritonk8ssupervisor_b200.conformance.sandbox import run_in_sandbox
    from tritonk8ssupervisor_b200.conformance.scenario import Scenario
    sc = Scenario(name="sedexec", stdin="\n" * 7 + "-e \\\\055ee 1\n\nyes\n")
    obs = run_in_sandbox(rp.REPLACEMENT_ARGV, WORKSPACE_TEMPLATE, sc, extra_env=rp.REPLACEMENT_ENV)
    assert obs.exit_status == 2
    assert "refusing to go on: the index spliced into the sed program at line 541 makes it 'ep'" in obs.stderr
    assert "terraform/rancher.tf" not in obs.files and not [c for c in obs.calls if c.startswith("terraform")]
