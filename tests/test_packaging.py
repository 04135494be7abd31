"""The package installs as a wheel that needs nothing from this checkout: both
commands are console scripts, the workspace template and the stub tools ride
along as package data, and the setup path runs from an unrelated directory
without PYTHONPATH pointing at the repository."""
import os
import subprocess
import sys
import zipfile

import pytest

from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.catalog import CONFIGS
from tritonk8ssupervisor_b200.conformance.harness import golden_for

ROOT = rp.PACKAGE_PARENT


@pytest.fixture(scope="module")
def unpacked_wheel(tmp_path_factory):
    out = tmp_path_factory.mktemp("wheel")
    src = tmp_path_factory.mktemp("src")
    # build from a copy: the build writes build/ and *.egg-info next to pyproject.toml
    subprocess.run(["cp", "-r", os.path.join(ROOT, "tritonk8ssupervisor_b200"), os.path.join(ROOT, "pyproject.toml"),
                    os.path.join(ROOT, "README.md"), str(src)], check=True)
    p = subprocess.run([sys.executable, "-m", "pip", "wheel", "--no-index", "--no-deps", "--no-build-isolation", "-q",
                        "-w", str(out), str(src)], capture_output=True, text=True, timeout=600)
    if p.returncode != 0:
        pytest.skip("cannot build a wheel here: %s" % p.stderr[-500:])
    (whl,) = [f for f in os.listdir(out) if f.endswith(".whl")]
    site = out / "site"
    with zipfile.ZipFile(out / whl) as z:
        names = z.namelist()
        z.extractall(site)
    return str(site), names


def test_wheel_carries_data_and_entry_points(unpacked_wheel):
    site, names = unpacked_wheel
    for rel in ("workspace/terraform/master/machine.tf", "workspace/terraform/host/inputs.tf", "workspace/ansible/clusterUp.yml",
                "workspace/ansible/ansible.cfg", "workspace/ansible/roles/rancherhost/tasks/main.yml",
                "workspace/ansible/roles/ranchermaster/vars/vars.yml", "conformance/stubs/stubtool.sh", "basharith.py"):
        assert "tritonk8ssupervisor_b200/" + rel in names, rel
    (ep,) = [n for n in names if n.endswith("entry_points.txt")]
    text = open(os.path.join(site, ep)).read()
    assert "tk8s-setup = tritonk8ssupervisor_b200.cli:main" in text and "tk8s-tools = tritonk8ssupervisor_b200.tools:main" in text
    assert not any(n.startswith(("tests/", "oracle/")) for n in names)


def test_setup_runs_from_the_installed_copy_in_an_unrelated_cwd(unpacked_wheel, tmp_path):
    """Config 2 replayed with the *installed* package first on sys.path and the checkout nowhere on it."""
    site, _ = unpacked_wheel
    ws = tmp_path / "elsewhere" / "cluster"
    p = subprocess.run([sys.executable, "-m", "tritonk8ssupervisor_b200.tools", "init", str(ws)], cwd="/", capture_output=True,
                       text=True, env={"PATH": os.environ["PATH"], "PYTHONPATH": site, "PYTHONDONTWRITEBYTECODE": "1"})
    assert p.returncode == 0 and (ws / "terraform/master/machine.tf").exists(), p.stderr
    probe = subprocess.run([sys.executable, "-c", "import tritonk8ssupervisor_b200 as m; print(m.__file__)"], cwd="/",
                           capture_output=True, text=True, env={"PATH": os.environ["PATH"], "PYTHONPATH": site})
    assert probe.stdout.startswith(site)
    expected, got = rp.replay(golden_for(CONFIGS[1]), extra_env=dict(rp.REPLACEMENT_ENV, PYTHONPATH=site))
    assert expected.diff(got) == []
