"""The five-config harness (BASELINE.json's metric) and its failure modes."""
import json
import os
import subprocess
import sys

import pytest

from tritonk8ssupervisor_b200.conformance import harness
from tritonk8ssupervisor_b200.conformance.catalog import CONFIGS

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_five_of_five():
    results = harness.run_all()
    assert [r.name for r in results] == [c.name for c in CONFIGS]
    for r in results:
        assert r.passed, (r.name, r.details)
    assert harness.pass_rate(results) == 1.0
    assert [r.n_modules for r in results] == [2, 3, 4, 6, 10]      # 1 master + N workers
    assert [r.n_hosts for r in results] == [2, 3, 4, 6, 10]


@pytest.mark.gpu
def test_five_of_five_on_box():
    """Selected by ``-m gpu``; no GPU is used (see conftest)."""
    assert harness.pass_rate(harness.run_all()) == 1.0


def test_count_scaling_extension():
    ext = harness.extension_count_scaling(10)
    assert ext["ok"], ext
    assert ext["modules"] == 11 and ext["planned_machines"] == 11 and ext["inventory_hosts"] == 11
    assert ext["size_law"]["per_worker"] == 234          # +234 bytes per worker block


def test_machine_count_must_match_the_run(tmp_path):
    """Check (ii) holds the plan to 1 master + KUBERNETES_NUMBER_OF_NODES from the run's
    own ./config, not to a floor: a generator that drops a worker block fails it."""
    import shutil
    from tritonk8ssupervisor_b200.conformance.sandbox import Observation
    doc = harness.golden_for(CONFIGS[1])
    obs = Observation.from_json(doc["expected"])
    assert harness.expected_machine_count(obs) == 3
    ws = tmp_path / "ws"
    shutil.copytree(harness.WORKSPACE_TEMPLATE, ws)
    text = obs.files["terraform/rancher.tf"]
    cut = text[:text.index('\nmodule "kubenode2"')]
    (ws / "terraform" / "rancher.tf").write_text(cut)
    (ws / "ansible" / "hosts").write_text(obs.files["ansible/hosts"])
    (ws / "ansible/roles/ranchermaster/vars/vars.yml").write_text(obs.files["ansible/roles/ranchermaster/vars/vars.yml"])
    res = harness.ConfigResult(name="dropped")
    harness.check_generated_tree(str(ws), obs, res)
    assert not res.terraform and any("2 machines planned, the run asked for 3" in d for d in res.details)


def test_a_wrong_implementation_fails_parity(tmp_path):
    """The harness must be able to say no: an implementation that gets one byte
    wrong scores 0 on parity."""
    bad = tmp_path / "bad.py"
    bad.write_text(
        "import sys, runpy\n"
        "import tritonk8ssupervisor_b200.supervisor as s\n"
        "s.DASHES = '--------------'\n"                       # one dash short
        "sys.argv = ['setup.sh'] + sys.argv[1:]\n"
        "runpy.run_module('tritonk8ssupervisor_b200', run_name='__main__')\n")
    res = harness.run_config(harness.golden_for(CONFIGS[0]), argv=[sys.executable, str(bad)], repeat=1)
    assert not res.parity and not res.passed
    assert res.terraform and res.ansible and res.inventory


def test_structural_failure_is_detected(tmp_path):
    """Break a module of the workspace: parity still holds, terraform check fails."""
    import shutil
    ws = tmp_path / "ws"
    shutil.copytree(harness.WORKSPACE_TEMPLATE, ws)
    p = ws / "terraform" / "host" / "inputs.tf"
    p.write_text(p.read_text().replace('variable "package"', 'variable "packge"'))
    res = harness.run_config(harness.golden_for(CONFIGS[1]), workspace_src=str(ws), repeat=1)
    assert res.parity and not res.terraform and not res.passed
    assert any("arg-unknown" in d for d in res.details)


def test_bench_line_contract():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "0", "--median-runs", "2"],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert p.returncode == 0, p.stderr
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "roofline", "cpu_baseline", "e2e", "gpu_launches", "clocks"):
        assert key in d, key
    assert d["value"] == 1.0 and d["e2e"]["value"] == 1.0
    assert d["gpu_launches"] == 0 and d["roofline"]["frac"] is None and d["vs_baseline"] is None
    assert "workload" in d["config"]
    assert set(d["e2e"]["median_run_ms_per_config"]) == {c.name for c in CONFIGS}
    # the whole recorded parity surface rides along, and no criterion is asserted rather than measured
    from tritonk8ssupervisor_b200.conformance import replay as rp
    n = len(rp.list_goldens())
    assert d["catalogue"]["replayed"] + d["catalogue"]["skipped_not_reproducible_here"] == n
    assert d["catalogue"]["passed"] == d["catalogue"]["replayed"] >= n - 2
    for leg in ("per_config", "per_config_e2e"):
        assert set(d[leg]) == {c.name for c in CONFIGS}
        assert all(v["deterministic"] and v["passed"] for v in d[leg].values())
    assert d["scaling"] == "strong"


def test_bench_reference_arm(ref_root):
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert p.returncode == 0, p.stderr
    d = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][0])
    assert d["impl"] == "reference" and d["value"] == 1.0 and d["cpu_baseline"]["kind"] == "reference"
    assert d["e2e"]["h2d_bytes_per_step"] == 0


def test_bench_two_ranks_gloo():
    """N > 1: independent CPU replicas, gloo for the barrier and the max/min over
    ranks; rank 0 prints the one line."""
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["n_gpus"] == 2 and d["value"] == 1.0 and d["e2e"]["value"] == 1.0
    assert d["cpu_baseline"]["value"] is None      # baseline is timed at N=1 only
    # the ranks divide the work instead of repeating it: rank 0 saw 3 of the 5 configs, the sums cover everything
    from tritonk8ssupervisor_b200.conformance import replay as rp
    assert len(d["per_config"]) == 3
    assert d["catalogue"]["replayed"] + d["catalogue"]["skipped_not_reproducible_here"] == len(rp.list_goldens())
    assert d["catalogue"]["passed"] == d["catalogue"]["replayed"]


def test_triton_table_cache_changes_only_the_trace():
    """TK8S_CACHE_TRITON=1: byte-identical stdout and files, strictly fewer
    `triton` calls (one `networks`, one `packages`), nothing else reordered."""
    from tritonk8ssupervisor_b200.conformance import replay as rp
    for sc in CONFIGS:
        doc = harness.golden_for(sc)
        env = dict(rp.REPLACEMENT_ENV, TK8S_CACHE_TRITON="1")
        want, got = rp.replay(doc, extra_env=env)
        assert (got.stdout, got.stderr, got.exit_status, got.removed) == (want.stdout, want.stderr, want.exit_status, want.removed)
        assert got.files == want.files

        def strip(calls):
            """Drop each `triton` call together with the environment digest recorded under it."""
            out, skip = [], False
            for c in calls:
                if c.startswith("triton|"):
                    skip = True
                    continue
                if skip and c.startswith("  env~:"):
                    continue
                skip = False
                out.append(c)
            return out
        assert strip(got.calls) == strip(want.calls)
        triton = [c for c in got.calls if c.startswith("triton|")]
        assert triton == ["triton|.|env", "triton|.|networks -oname,id", "triton|.|packages -oname,id"]
        assert len([c for c in want.calls if c.startswith("triton|")]) >= 6      # env + 2 menus + 3 default resolutions


def test_bench_reference_arm_under_torchrun(ref_root):
    """N > 1 reference arm: rank 0 alone runs and prints, the others exit 0 (no process group, no hang)."""
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "bench.py"),
                        "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1 and json.loads(lines[0])["impl"] == "reference"
