"""the advisory modern-dialect report.  Advisory means
advisory: it never fails a workspace."""
import collections
import os
import shutil

from tritonk8ssupervisor_b200 import lintmodern
from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.harness import WORKSPACE_TEMPLATE


def _workspace(tmp_path, src):
    ws = tmp_path / "ws"
    shutil.copytree(src, ws, ignore=shutil.ignore_patterns(".git", "img"))
    doc = rp.load_golden(os.path.join(rp.golden_dir(), "config2_quickstart_2workers.json"))
    for rel, content in doc["expected"]["files"].items():
        if rel.startswith(("terraform/", "ansible/")):
            (ws / rel).write_text(content)
    return ws


def test_reference_workspace_gets_the_era_list(tmp_path, ref_root):
    """Exactly the items SURVEY.md §7.4 enumerates, at the lines it cites."""
    rep = lintmodern.lint_workspace(str(_workspace(tmp_path, ref_root)))
    assert rep.ok and all(f.severity == "advice" for f in rep.findings)
    n = collections.Counter(f.code for f in rep.findings)
    assert n["TF-SOURCE"] == 3            # setup.sh:166,178 -> one per generated module block
    assert n["TF-TYPE"] == 2              # vars.tf:6 in both modules
    assert n["TF-DUPATTR"] == 2           # main.tf:6-8 vs 33-35
    assert n["TF-SELFREF"] == 4           # main.tf:15,30
    assert n["TF-CLI"] == 3               # setup.sh:156,157,501
    assert n["TF-PROVIDER"] == 1
    assert n["AN-WHEN"] == 8              # ranchermaster:13,20,22,27 + rancherhost:9,18,25,34
    assert n["AN-MODULE"] == 1            # rancherhost:27 `docker:`
    assert n["AN-HEADER"] == 1            # ranchermaster:41-42
    assert n["AN-SHADOW"] == 2            # clusterUp.yml:11-12,21-22
    assert n["AN-PY2"] == 2               # main.tf:25
    where = {f.where for f in rep.findings if f.code == "TF-SELFREF"}
    assert where == {"master/main.tf:15", "master/main.tf:30", "host/main.tf:15", "host/main.tf:30"}


def test_package_workspace_keeps_only_what_the_generated_file_forces(tmp_path):
    """The package's own modules/roles avoid the avoidable items; what remains is
    dictated by the generated rancher.tf (bare source, quoted interpolation), by
    the 0.9 dialect it has to fit (quoted types) and by the target image (python 2)."""
    rep = lintmodern.lint_workspace(str(_workspace(tmp_path, WORKSPACE_TEMPLATE)))
    codes = {f.code for f in rep.findings}
    assert codes == {"TF-SOURCE", "TF-TYPE", "TF-INTERP", "TF-CLI", "TF-PROVIDER", "AN-PY2", "AN-LOOPPKG"}
    assert not codes & {"TF-DUPATTR", "TF-SELFREF", "AN-WHEN", "AN-MODULE", "AN-HEADER", "AN-SHADOW", "AN-KV"}
