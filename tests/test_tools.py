"""The offline tools command and the whole user journey: init a workspace, run
setup in it (stub tools on PATH), then validate / plan / syntax-check /
inventory / lint / dry-run the result."""
import io
import json
import os
import subprocess
import sys

from tritonk8ssupervisor_b200 import tools
from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.sandbox import Sandbox
from tritonk8ssupervisor_b200.conformance.scenario import Scenario


def _tool(*argv):
    out = io.StringIO()
    return tools.main(list(argv), out=out), out.getvalue()


def test_user_journey(tmp_path):
    ws = tmp_path / "cluster"
    rc, text = _tool("init", str(ws))
    assert rc == 0 and (ws / "terraform/master/machine.tf").exists() and (ws / "ansible/clusterUp.yml").exists()
    assert _tool("init", str(ws))[0] == 1                       # refuses to overwrite

    # setup, 3 workers, in the freshly initialised workspace
    sc = Scenario(name="journey", stdin="\n\n\n\n3\n\n\n\nyes\n", env_id="1a7")
    with Sandbox(str(ws), sc) as sbx:
        obs = sbx.run(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV)
        assert obs.exit_status == 0 and "CONGRATULATIONS" in obs.stdout
        root = sbx.workspace
        rc, text = _tool("validate", root)
        assert rc == 0 and text.endswith("0 error(s), 0 warning(s)\n")
        rc, text = _tool("plan", root)
        plan = json.loads(text)
        assert [p["address"] for p in plan] == ["module.kubemaster.triton_machine.master"] + \
            ["module.kubenode%d.triton_machine.host" % i for i in (1, 2, 3)]
        rc, text = _tool("syntax-check", root)
        assert rc == 0, text
        rc, text = _tool("inventory", root)
        doc = json.loads(text)
        assert doc["MASTER"]["hosts"] == ["10.0.0.1"] and doc["HOST"]["hosts"] == ["10.0.1.1", "10.0.1.2", "10.0.1.3"]
        rc, text = _tool("lint", root)
        assert rc == 0 and "advice[TF-SOURCE]" in text
        rc, text = _tool("dry-run", root)
        assert rc == 0, text
        assert "wire: POST /v2-beta/projects" in text and text.count("wire: POST /v1/registrationtokens?projectId=1a7") == 3

        # break the generated file: validate says so and exits 1
        p = os.path.join(root, "terraform", "rancher.tf")
        with open(p) as fh:
            broken = fh.read().replace('hostname = "kubenode2"', 'hostnme = "kubenode2"')
        with open(p, "w") as fh:
            fh.write(broken)
        rc, text = _tool("validate", root)
        assert rc == 1 and "error[arg-unknown]" in text and "error[arg-missing]" in text


def test_tools_as_a_process(tmp_path):
    p = subprocess.run([sys.executable, "-m", "tritonk8ssupervisor_b200.tools", "init", str(tmp_path / "w")],
                       capture_output=True, text=True, env=dict(os.environ, PYTHONPATH=rp.PACKAGE_PARENT))
    assert p.returncode == 0 and "workspace laid down" in p.stdout
    p = subprocess.run([sys.executable, "-m", "tritonk8ssupervisor_b200.tools", "inventory", str(tmp_path / "w")],
                       capture_output=True, text=True, env=dict(os.environ, PYTHONPATH=rp.PACKAGE_PARENT))
    assert p.returncode == 1 and "not found" in p.stdout


def test_launcher_script_is_the_same_program():
    """bin/tk8s-setup from any cwd == python -m tritonk8ssupervisor_b200 (config 2 replay)."""
    from tritonk8ssupervisor_b200.conformance.catalog import CONFIGS
    from tritonk8ssupervisor_b200.conformance.harness import golden_for
    launcher = os.path.join(rp.PACKAGE_PARENT, "bin", "tk8s-setup")
    expected, got = rp.replay(golden_for(CONFIGS[1]), argv=[launcher],
                              extra_env={"PYTHONDONTWRITEBYTECODE": "1", "TK8S_PROGRAM": "<PROGRAM>"})
    assert expected.diff(got) == []


def test_inprocess_clock_is_opt_in():
    """Default: `sleep`/`date +%s` are child processes from PATH, as in the reference
    (setup.sh:36,41,46,63,64) -- that is what every golden replays.  TK8S_INPROCESS_CLOCK=1
    removes exactly those entries from the trace and changes nothing else; the switch
    itself does not reach the tools."""
    import time
    from tritonk8ssupervisor_b200.conformance.catalog import CONFIGS
    from tritonk8ssupervisor_b200.conformance.harness import golden_for
    doc = golden_for(CONFIGS[0])
    t0 = time.time()
    want, got = rp.replay(doc, extra_env=dict(rp.REPLACEMENT_ENV, TK8S_INPROCESS_CLOCK="1"))
    assert time.time() - t0 >= 6            # three real `sleep 2`s plus the wait loop's seconds
    assert (got.stdout.replace(".", ""), got.stderr, got.exit_status, got.files) == (want.stdout.replace(".", ""), want.stderr, want.exit_status, want.files)

    def tools(calls):
        out, keep = [], True
        for c in calls:
            if not c.startswith("  "):
                keep = c.split("|")[0] not in ("sleep", "date")
            if keep:
                out.append(c)
        return out
    assert got.calls == tools(got.calls) == tools(want.calls)
    assert len(want.calls) > len(got.calls)
    assert not any("TK8S_INPROCESS_CLOCK" in c for c in got.calls)
