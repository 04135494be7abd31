"""Permission bits and link counts of everything the run leaves in the workspace, under several umasks and with an
ansible.cfg that had bits of its own (`sed ... > tmp && mv tmp ansible.cfg` gives it the bits of a new file)."""
import os
import stat
import subprocess

import pytest

from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.sandbox import Sandbox
from tritonk8ssupervisor_b200.conformance.scenario import Scenario

FULL = Scenario(name="modes", stdin="\n" * 8 + "yes\n")
CLEAN = Scenario(name="modes-clean", args=["-c"], stdin="yes\n",
                 pre_files={"terraform/rancher.tf": "x\n", "terraform/masters.ip": "10.0.0.1\n", "terraform/hosts.ip": "10.0.1.1\n", "config": "A=1\n"})


def _modes(argv, env_extra, root, sc, umask, cfg_mode):
    with Sandbox(root, sc) as sbx:
        if cfg_mode is not None:
            os.chmod(os.path.join(sbx.workspace, "ansible", "ansible.cfg"), cfg_mode)
        p = subprocess.run(list(argv) + list(sc.args), cwd=sbx.workspace, input=sc.stdin.encode(), stdout=subprocess.PIPE,
                           stderr=subprocess.PIPE, env=sbx.env(env_extra), timeout=60, preexec_fn=lambda: os.umask(umask))
        found = {}
        for d, _, files in os.walk(sbx.workspace):
            for f in files:
                st = os.lstat(os.path.join(d, f))
                found[os.path.relpath(os.path.join(d, f), sbx.workspace)] = (oct(stat.S_IMODE(st.st_mode)), st.st_nlink)
        return p.returncode, found


@pytest.mark.parametrize("umask", [0o022, 0o077, 0o000])
@pytest.mark.parametrize("cfg_mode", [None, 0o600, 0o755])
@pytest.mark.parametrize("sc", [FULL, CLEAN], ids=["run", "clean"])
def test_modes_same_as_reference(ref_root, sc, cfg_mode, umask):
    theirs = _modes(["./setup.sh"], None, ref_root, sc, umask, cfg_mode)
    mine = _modes(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, ref_root, sc, umask, cfg_mode)
    assert mine == theirs
    want = oct(0o666 & ~umask)
    assert theirs[1]["ansible/ansible.cfg"][0] == want          # whatever bits it had: rewritten through a new file
