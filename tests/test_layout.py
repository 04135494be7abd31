"""Repository invariants that the documents assert."""
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _walk(top, exts):
    for d, dirs, files in os.walk(top):
        dirs[:] = [x for x in dirs if x not in (".git", "_ref", "gpurun_out", "__pycache__")]
        for f in files:
            if f.endswith(exts):
                yield os.path.join(d, f)


def test_no_gpu_code_and_no_c_abi():
    """BASELINE.json: TIER MISMATCH, no GPU graft target.  include/ holds no header
    because the reference has no FFI to mirror (include/README.md)."""
    assert list(_walk(ROOT, (".cu", ".cuh"))) == []
    assert list(_walk(os.path.join(ROOT, "include"), (".h", ".hpp"))) == []
    assert list(_walk(os.path.join(ROOT, "tritonk8ssupervisor_b200"), (".so",))) == []


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: only tests/, __graft_entry__.smoke()/build() and bench.py may use it."""
    for path in _walk(os.path.join(ROOT, "tritonk8ssupervisor_b200"), (".py",)):
        text = open(path, encoding="utf-8").read()
        assert "import oracle" not in text and "from oracle" not in text, path
        assert "/root/reference" not in text, path


def test_nothing_on_the_box_path_reads_root_reference():
    """/root/reference does not exist on the GPU box: bench.py and __graft_entry__ find the reference only
    through oracle.reference_runner.find_reference(), which falls back to the staged copy."""
    for name in ("bench.py",):
        assert "/root/reference" not in open(os.path.join(ROOT, name), encoding="utf-8").read().replace(
            "neither /root/reference nor", "")


def test_no_reference_sources_in_git():
    import subprocess
    tracked = subprocess.run(["git", "ls-files"], cwd=ROOT, capture_output=True, text=True).stdout.split()
    assert not [t for t in tracked if t.startswith("oracle/_ref/")]
    assert "setup.sh" not in [os.path.basename(t) for t in tracked]
