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
    assert not res.ok and "unknown option" in res.error


def test_clean_form():
    """setup.sh:511: delimiter ~ and an empty value."""
    res = sedprog.run(CFG, "", "~")
    assert res.ok and "private_key_file = \n" in res.output and "old" not in res.output
