"""Ctrl-C while a provisioning tool runs, and a reader that goes away.

bash does not act on SIGINT while a foreground child is running: it waits for
the child, and then dies of the signal itself only if the child did.  A tool
that traps the signal to save its state (terraform does) therefore always gets
to finish before the operator has the prompt back.  Compared live with the
reference where a reference tree is available."""
import os
import signal
import subprocess
import time

import pytest

from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.catalog import CONFIGS
from tritonk8ssupervisor_b200.conformance.harness import WORKSPACE_TEMPLATE
from tritonk8ssupervisor_b200.conformance.sandbox import Sandbox

#: terraform stand-ins: apply announces itself, then sleeps until interrupted
_TRAPS_AND_SAVES = """#!/bin/sh
[ "$1" = apply ] || exit 0
trap 'sleep 0.5; echo "state saved" >> saved.txt; echo 10.0.0.1 >> masters.ip; echo 10.0.1.1 >> hosts.ip; exit %d' INT
touch started
while :; do /bin/sleep 0.05; done
"""
_DIES = """#!/bin/sh
[ "$1" = apply ] || exit 0
touch started
exec /bin/sleep 30
"""


def _interrupt_during_apply(argv, env_extra, workspace_src, stub):
    with Sandbox(workspace_src, CONFIGS[0]) as sbx:
        tf = os.path.join(sbx.bin, "terraform")
        with open(tf, "w") as fh:
            fh.write(stub)
        os.chmod(tf, 0o755)
        proc = subprocess.Popen(list(argv), cwd=sbx.workspace, stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                                env=sbx.env(env_extra), start_new_session=True)
        proc.stdin.write(CONFIGS[0].stdin.encode())
        proc.stdin.flush()
        started = os.path.join(sbx.workspace, "terraform", "started")
        deadline = time.time() + 30
        while not os.path.exists(started) and time.time() < deadline:
            time.sleep(0.02)
        assert os.path.exists(started), proc.stderr.read()
        os.killpg(proc.pid, signal.SIGINT)              # what the terminal does on Ctrl-C
        out, err = proc.communicate(timeout=30)
        saved = os.path.exists(os.path.join(sbx.workspace, "terraform", "saved.txt"))
        return {"status": proc.returncode, "saved_before_exit": saved, "tail": out.decode().splitlines()[-1:],
                "ansible_ran": b"PLAY RECAP" in out}


CASES = [("child traps SIGINT and fails", _TRAPS_AND_SAVES % 1, {"status": 1, "saved_before_exit": True, "ansible_ran": False}),
         ("child traps SIGINT and succeeds", _TRAPS_AND_SAVES % 0, {"status": 0, "saved_before_exit": True, "ansible_ran": True}),
         ("child dies of SIGINT", _DIES, {"status": -signal.SIGINT, "saved_before_exit": False, "ansible_ran": False})]


@pytest.mark.parametrize("label,stub,want", CASES, ids=[c[0] for c in CASES])
def test_sigint_during_terraform_apply(label, stub, want):
    got = _interrupt_during_apply(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, WORKSPACE_TEMPLATE, stub)
    assert {k: got[k] for k in want} == want


@pytest.mark.parametrize("label,stub,want", CASES, ids=[c[0] for c in CASES])
def test_sigint_same_as_reference(ref_root, label, stub, want):
    theirs = _interrupt_during_apply(["./setup.sh"], None, ref_root, stub)
    mine = _interrupt_during_apply(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, ref_root, stub)
    assert mine == theirs


def _reader_goes_away(argv, env_extra, workspace_src):
    """stdout is a pipe whose reader exits after the first line."""
    with Sandbox(workspace_src, CONFIGS[0]) as sbx:
        proc = subprocess.Popen(list(argv), cwd=sbx.workspace, stdin=subprocess.PIPE, stdout=subprocess.PIPE,
                                stderr=subprocess.DEVNULL, env=sbx.env(env_extra))
        proc.stdout.readline()
        proc.stdout.close()
        proc.stdin.write(CONFIGS[0].stdin.encode())
        proc.stdin.close()
        return proc.wait(timeout=30), os.path.exists(os.path.join(sbx.workspace, "terraform", "rancher.tf"))


def test_sigpipe_kills_like_a_shell_script():
    status, wrote = _reader_goes_away(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, WORKSPACE_TEMPLATE)
    assert status == -signal.SIGPIPE and not wrote


def test_sigpipe_same_as_reference(ref_root):
    assert _reader_goes_away(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, ref_root) == _reader_goes_away(["./setup.sh"], None, ref_root)


# --- a tool that a signal kills, with no signal sent to the script itself (the OOM killer, a crash) -----
def _tool_dies_of(argv, env_extra, workspace_src, tool, first, signame):
    import re
    with Sandbox(workspace_src, CONFIGS[0]) as sbx:
        path = os.path.join(sbx.bin, tool)
        os.rename(path, os.path.join(sbx.bin, "stub-of-" + tool))
        with open(path, "w") as fh:
            fh.write('#!/bin/sh\nif [ "$1" = "%s" ]; then kill -%s $$; fi\nexit 0\n' % (first, signame))
        os.chmod(path, 0o755)
        p = subprocess.run(list(argv), cwd=sbx.workspace, input=CONFIGS[0].stdin.encode(), stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                           env=sbx.env(env_extra), timeout=60)
        err = re.sub(r"line (\d+): +\d+ ", r"line \1: <PID> ", p.stderr.decode(errors="replace").replace(sbx.tmp, "<SBX>"))
        return p.returncode, p.stdout.decode(errors="replace")[-60:], err[-300:]


KILLED = [("terraform", "get", "KILL", 137, "./setup.sh: line 138: <PID> Killed                  terraform get\n"),
          ("terraform", "apply", "SEGV", 139, "./setup.sh: line 138: <PID> Segmentation fault      terraform apply\n"),
          ("terraform", "get", "TERM", 143, "Terminated\n"),
          ("ansible-playbook", "-i", "ABRT", 134, "./setup.sh: line 111: <PID> Aborted                 ansible-playbook -i hosts clusterUp.yml\n"),
          ("sleep", "2", "HUP", 129, "./setup.sh: line 8: <PID> Hangup                  sleep 2\n"),
          ("terraform", "get", "INT", 130, ""), ("terraform", "get", "PIPE", 141, ""),
          ("curl", "--connect-timeout", "INT", -signal.SIGINT, ""), ("date", "+%s", "INT", -signal.SIGINT, "")]


@pytest.mark.parametrize("tool,first,signame,status,said", KILLED, ids=["%s-%s-%s" % k[:3] for k in KILLED])
def test_tool_killed_by_a_signal(tool, first, signame, status, said):
    """bash reports a foreground command that a signal killed (not SIGINT, not SIGPIPE; for SIGTERM the bare word),
    and takes the SIGINT death of the only command of a $(...) for its own."""
    got = _tool_dies_of(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, WORKSPACE_TEMPLATE, tool, first, signame)
    assert got[0] == status and got[2].endswith(said) and (said != "" or "<PID>" not in got[2])


@pytest.mark.parametrize("tool,first,signame,status,said", KILLED, ids=["%s-%s-%s" % k[:3] for k in KILLED])
def test_tool_killed_by_a_signal_same_as_reference(ref_root, tool, first, signame, status, said):
    assert _tool_dies_of(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, ref_root, tool, first, signame) == \
        _tool_dies_of(["./setup.sh"], None, ref_root, tool, first, signame)


def _quit_at_first_prompt(argv, env_extra, workspace_src):
    with Sandbox(workspace_src, CONFIGS[0]) as sbx:
        proc = subprocess.Popen(list(argv), cwd=sbx.workspace, stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                                env=sbx.env(env_extra), start_new_session=True)
        log, last, since = os.path.join(sbx.state, "calls.log"), -1, time.time()
        while time.time() - since < 0.6:                # the tools stop being called: the script sits at its first prompt
            size = os.path.getsize(log) if os.path.exists(log) else 0
            if size != last:
                last, since = size, time.time()
            time.sleep(0.02)
        os.killpg(proc.pid, signal.SIGQUIT)             # Ctrl-\\
        out, err = proc.communicate(timeout=30)
        return proc.returncode, out.decode()[-20:], err.decode()


def test_sigquit_at_a_prompt():
    """The prompt of getArgument is read by the subshell of a $(...): SIGQUIT kills that (131, errexit), bash itself ignores it."""
    assert _quit_at_first_prompt(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, WORKSPACE_TEMPLATE) == (131, "---------------\n", "")


def test_sigquit_at_a_prompt_same_as_reference(ref_root):
    assert _quit_at_first_prompt(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, ref_root) == _quit_at_first_prompt(["./setup.sh"], None, ref_root)


def test_tool_killed_by_a_signal_while_cleaning(ref_root):
    """`terraform destroy ... || true` and `ssh-keygen -R` in cleanRunner: same notice, labelled with the line of
    the `case` / `for` they stand in."""
    import re
    from tritonk8ssupervisor_b200.conformance.scenario import Scenario
    sc = Scenario(name="c", args=["-c"], stdin="yes\n", known_hosts="x\n",
                  pre_files={"terraform/rancher.tf": "x\n", "terraform/masters.ip": "10.0.0.1\n", "terraform/hosts.ip": "10.0.1.1\n"})

    def run(argv, env_extra, tool, first, signame):
        with Sandbox(ref_root, sc) as sbx:
            path = os.path.join(sbx.bin, tool)
            os.remove(path)
            with open(path, "w") as fh:
                fh.write('#!/bin/sh\nif [ "$1" = "%s" ]; then kill -%s $$; fi\nexit 0\n' % (first, signame))
            os.chmod(path, 0o755)
            p = subprocess.run(list(argv) + ["-c"], cwd=sbx.workspace, input=b"yes\n", stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                               env=sbx.env(env_extra), timeout=60)
            return p.returncode, p.stdout.decode()[-40:], re.sub(r"line (\d+): +\d+ ", r"line \1: <PID> ", p.stderr.decode())
    for tool, first, signame in (("terraform", "destroy", "KILL"), ("terraform", "destroy", "TERM"), ("ssh-keygen", "-R", "SEGV"), ("ssh-keygen", "-R", "INT")):
        mine = run(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, tool, first, signame)
        assert mine == run(["./setup.sh"], None, tool, first, signame)
        if signame in ("KILL", "SEGV"):
            assert "<PID>" in mine[2]
