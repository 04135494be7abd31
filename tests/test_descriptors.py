"""Standard descriptors that are closed or cannot be written: the program must
behave like the bash script, not die with a traceback.  Expected values are
what the reference does (compared live where a reference tree is available)."""
import os
import subprocess

import pytest

from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.catalog import CONFIGS
from tritonk8ssupervisor_b200.conformance.harness import WORKSPACE_TEMPLATE
from tritonk8ssupervisor_b200.conformance.sandbox import Sandbox

CASES = {
    # redirection applied to the program : (exit status, stderr, last line of stdout)
    "stdout closed": (">&-", 1, "./setup.sh: line 264: echo: write error: Bad file descriptor\n", None),
    "stdout on a full device": ("> /dev/full", 1, "./setup.sh: line 264: echo: write error: No space left on device\n", None),
    "stderr closed": ("2>&-", 0, "", "    CONGRATULATIONS, YOU HAVE CONFIGURED YOUR KUBERNETES ENVIRONMENT!"),
    "stdin from /dev/null": ("< /dev/null", 1, "", "WARN: Make sure that the nodes and master are part of networks that can communicate "
                                                     "with each other and this system from which the setup is running."),
}


def _run(argv, env, workspace_src, redirection):
    with Sandbox(workspace_src, CONFIGS[0]) as sbx:
        p = subprocess.run(["sh", "-c", 'exec "$@" ' + redirection, "sh"] + list(argv), cwd=sbx.workspace,
                           input=None if "<" in redirection else CONFIGS[0].stdin.encode(),
                           stdin=subprocess.DEVNULL if "<" in redirection else None,
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=sbx.env(env), timeout=60)
        out = p.stdout.decode().splitlines()
        return p.returncode, p.stderr.decode().replace(sbx.tmp, "<SBX>"), (out[-1] if out else None)


@pytest.mark.parametrize("label", sorted(CASES))
def test_odd_descriptors(label):
    redirection, status, stderr, last = CASES[label]
    got = _run(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, WORKSPACE_TEMPLATE, redirection)
    assert got == (status, stderr, last)
    assert "Traceback" not in got[1]


@pytest.mark.parametrize("label", sorted(CASES))
def test_odd_descriptors_same_as_reference(ref_root, label):
    redirection = CASES[label][0]
    assert _run(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, ref_root, redirection) == _run(["./setup.sh"], None, ref_root, redirection)


@pytest.mark.parametrize("kind", ["file", "append", "offset", "devnull"])
def test_answers_from_a_regular_file(ref_root, kind):
    """`./setup.sh < answers.txt`: same run, and the descriptor is left at the same offset (tools that read the
    stdin they inherit find the file where bash would have left it for them)."""
    import subprocess
    from tritonk8ssupervisor_b200.conformance import replay as rp
    from tritonk8ssupervisor_b200.conformance.sandbox import Sandbox
    from tritonk8ssupervisor_b200.conformance.scenario import Scenario
    sc = Scenario(name="filestdin", stdin="x" * 300 + "\n\\\ncontinued\n" + "\n" * 6 + "yes\nA\nB\n" + "y" * 5000 + "\n",
                  stdin_eaters={"terraform-apply": 1, "sleep-2": 1})

    def run(argv, env_extra):
        with Sandbox(ref_root, sc) as sbx:
            path = os.path.join(sbx.tmp, "answers.txt")
            with open(path, "wb") as fh:
                fh.write(sc.stdin.encode())
            fh = open("/dev/null", "rb") if kind == "devnull" else open(path, "ab+" if kind == "append" else "rb")
            if kind == "offset":
                fh.seek(1)
            before = sbx.snapshot()
            p = subprocess.run(list(argv), cwd=sbx.workspace, stdin=fh, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=sbx.env(env_extra), timeout=60)
            where = os.lseek(fh.fileno(), 0, os.SEEK_CUR)
            fh.close()
            return sbx.observe(before, p.returncode, p.stdout, p.stderr), where
    theirs, their_offset = run(["./setup.sh"], None)
    mine, my_offset = run(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV)
    assert theirs.diff(mine) == [] and their_offset == my_offset


def _scenarios_that_say_something_on_stderr():
    import json
    from tritonk8ssupervisor_b200.conformance.catalog import ALL
    names = []
    for sc in ALL:
        path = os.path.join(rp.golden_dir(), sc.name + ".json")
        if os.path.exists(path) and not sc.name.endswith("__bash_setup_sh") and sc.name != "cfg_line_is_export_p":
            with open(path, encoding="utf-8") as fh:
                if json.load(fh)["expected"]["stderr"]:
                    names.append(sc.name)
    return names


@pytest.mark.parametrize("name", _scenarios_that_say_something_on_stderr())
def test_stdout_and_stderr_on_one_pipe(ref_root, name):
    """`./setup.sh 2>&1 | tee log`: the diagnostics fall between the same lines of output as the reference's
    (each `$(...)` is evaluated when the line that holds it is reached, not before)."""
    from tritonk8ssupervisor_b200.conformance.catalog import BY_NAME
    sc = BY_NAME[name]
    why = rp.not_reproducible_here({"scenario": sc.to_json(), "expected": {"stderr": "ls: cannot access"}}) if "HOME" not in sc.unset_env else "home"
    if why:
        pytest.skip(why)

    def run(argv, env_extra):
        with Sandbox(ref_root, sc) as sbx:
            p = subprocess.run(list(argv) + list(sc.args), cwd=sbx.workspace, input=sc.stdin.encode("utf-8", "surrogateescape"),
                               stdout=subprocess.PIPE, stderr=subprocess.STDOUT, env=sbx.env(env_extra), timeout=60)
            return p.returncode, p.stdout.replace(sbx.tmp.encode(), b"<SBX>")
    assert run(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV) == run(sc.reference_argv(), None)
