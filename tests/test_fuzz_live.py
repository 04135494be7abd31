"""A rougher generator than the hypothesis strategies of test_parity_live.py:
whole scenarios at random -- answers made of shell metacharacters and echo
options, shuffled and truncated `triton` tables, stale files, scripted tool
failures, tools that are not installed, tools that eat stdin, non-numeric
clock readings, and an operator environment that exports the very names the
script uses for its own variables.  Replacement and reference must agree on
everything observable, the complete environment of every child included.

During development some 25 000 seeds of the generators below were run against the
reference; what it found is fixed and pinned as catalogue scenarios
(``env_exported_names_of_script_variables``, ``env_shell_level_*``).  The test
runs a slice of it on every run where a reference tree is available.
"""
import random

import pytest

from oracle.reference_runner import find_reference, run_reference
from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.sandbox import run_in_sandbox
from tritonk8ssupervisor_b200.conformance.scenario import DOC_NETWORKS, DOC_PACKAGES, Scenario, triton_env_text, triton_table

pytestmark = pytest.mark.skipif(find_reference() is None, reason="no reference tree (not /root/reference, not oracle/_ref)")

ALPHABET = list("ab 12\"'\\*?[]-ne$~;=.,/\t#&|<>(){}!%^`@:+_") + [
    "-e ", "-n ", "\\\\n", "\\\\0", "\\\\c", "\\\\t", "yes", "no", "é", "\x01", "\x7f", "\udcff", "  ", "0", "9", "10", "1,2", "4,2,2",
    "$HOME", "$(id)", "KUBERNETES_NUMBER_OF_NODES=3", "\\\\055e "]
EXPORTED = [("SHLVL", "7"), ("KUBERNETES_NAME", "from env"), ("i", "5"), ("yn", "q"), ("network", "n"), ("HOST_PACKAGE", "p"),
            ("OLDPWD", "/tmp"), ("PWD", "/x"), ("f", "ff"), ("line", "l"), ("master", "m"), ("_", "u"), ("IFS", ","), ("tmp", "t"),
            ("networks", "nn"), ("packages", "pp"), ("foundKey", "fk"), ("countNetwork", "9"), ("values", "vv"),
            ("gotValidInput", "g"), ("package", "pk"), ("cRetVal", "c"), ("node", "nd"), ("host_key", "hk"), ("theargument", "ta")]
CLOCK = ["1500000000", "1500000001", "abc", "", "08", "-15", "i", "1 2", "0x1e", "yn", "x=1", "1/0", "n[", "i[0]"]
FAILING = ["triton-env", "triton-networks", "triton-packages", "terraform-get", "terraform-apply", "ansible-playbook--i",
           "curl---connect-timeout", "ssh-keygen--E", "ssh-keygen--l", "date-+%s", "sleep-2", "sleep-1", "ssh--o"]


def _text(r, k=6):
    return "".join(r.choice(ALPHABET) for _ in range(r.randint(0, k)))


def _answers(r):
    pools = [lambda: _text(r), lambda: "", lambda: r.choice(["km01", "a", "1x", "abc def", "Ab9"]), lambda: str(r.randint(0, 12)),
             lambda: ",".join(str(r.randint(0, 6)) for _ in range(r.randint(1, 4))), lambda: r.choice(["yes", "no", "y", "YES", " yes "])]
    return "".join(r.choice(pools)() + "\n" for _ in range(r.randint(0, 22))) + r.choice(["yes\n", "yes", "", "no\n", "yes\nyes\nyes\n"])


def _table(r, rows):
    rows = r.sample(rows, r.randint(0, len(rows)))
    rows += [(r.choice(["kube", "kube-net", "a=b", "NAME", "NAMEX", "x y", "*", "k4-highcpu-kvm-7.75G", "Joyent-SDC-Public", "z-kvm-1"]),
              r.choice(["id-%d" % i, "", "a b", "*"])) for i in range(r.randint(0, 3))]
    r.shuffle(rows)
    t = triton_table(rows) if rows else "NAME  ID\n"
    return r.choice([t, t, t.replace("NAME", "NAME ", 1), t.rstrip("\n"), ""])


def make(seed: int) -> Scenario:
    r = random.Random(seed)
    kw = dict(name="fuzz%d" % seed, stdin=_answers(r), networks=_table(r, list(DOC_NETWORKS)), packages=_table(r, list(DOC_PACKAGES)))
    if r.random() < 0.3:
        kw["args"] = r.choice([["-c"], ["-c", "x"], ["x"], ["-C"], [""]])
    if r.random() < 0.3:
        kw["pre_files"] = {p: "x\n" for p in r.sample(
            ["config", "terraform/rancher.tf", "terraform/masters.ip", "terraform/hosts.ip", "ansible/hosts", "tmp/a", "config.tmp",
             "ansible/x.retry", "terraform/terraform.tfstate"], r.randint(1, 4))}
    if r.random() < 0.3:
        kw["date_seq"] = [r.choice(CLOCK) for _ in range(r.randint(1, 4))]
        kw["curl_seq"] = [r.choice(["starting\n", "", "503 Service Unavailable\n", "* ?\n", "-n kubernetes\n"]) for _ in range(r.randint(0, 3))]
    if r.random() < 0.2:
        kw["fail"] = {r.choice(FAILING): r.choice([1, 2, 127, 255])}
    if r.random() < 0.2:
        kw["missing_tools"] = [r.choice(["triton", "ssh-keygen", "terraform", "ansible-playbook", "curl", "ssh"])]
    if r.random() < 0.25:
        kw["env"] = dict(r.sample(EXPORTED, r.randint(1, 4)))
    if r.random() < 0.15:
        kw["triton_env"] = triton_env_text(key_id=r.choice(["aa:bb:cc:dd", "", "x y", "*"])) + r.choice(
            ["", "export A=$HOME\n", "unset HOME\n", "export HOME=/nonexistent\n", "export PATH=$PATH\n", "B=~/x\nexport B\n",
             "export network=zz\n", "export i\n", "unset PWD\n"])
    if r.random() < 0.15:
        kw["ssh_files"] = {r.choice(["id_rsa", "a b", "*", "k.pub", "x;y", "-n", "z&"]): "k\n" for _ in range(r.randint(0, 3))}
    if r.random() < 0.1:
        kw["stdin_eaters"] = {r.choice(["triton-networks", "triton-packages", "terraform-apply", "ssh-keygen--E",
                                        "curl---connect-timeout", "date-+%s", "sleep-2"]): r.randint(1, 2)}
    return Scenario(**kw)


@pytest.mark.parametrize("block", range(4))
def test_property_fuzzed_scenarios(ref_root, block):
    bad = []
    for seed in range(7000 + 12 * block, 7000 + 12 * (block + 1)):
        sc = make(seed)
        theirs = run_reference(sc, ref_root)
        mine = run_in_sandbox(rp.REPLACEMENT_ARGV, ref_root, sc, extra_env=rp.REPLACEMENT_ENV)
        diffs = theirs.diff(mine)
        if diffs:
            bad.append((seed, sc.to_json(), diffs))
    assert bad == []


# --- a wider variant: what the tools below the offline line print and leave behind ---------
_PS = ["1111 img k8s_kubernetes-dashboard.a\n3333 pause k8s_POD.x_kubernetes-dashboard-1\n", "5555 agent rancher-agent\n", "",
       "a b kubernetes-dashboard\nc d kubernetes-dashboard\n", "* k8s_POD.dashboard kubernetes-dashboard\n? x kubernetes-dashboard\n"]


def make_wide(seed: int) -> Scenario:
    import dataclasses
    sc = make(seed)
    r = random.Random(seed * 7 + 1)
    kw = {}
    if r.random() < 0.4:
        kw["master_ip"] = r.choice(["10.0.0.1", "1.2.3.4 5.6.7.8", "*", "a b", "", "-n", "ter*"])
    if r.random() < 0.4:
        kw["env_id"] = r.choice(["1a7", "", "x y", "*", "1a7\n1a8", "-e"])
    if r.random() < 0.5:
        kw["ssh_ps"] = {"10.0.1.%d" % i: r.choice(_PS) for i in range(1, 4)}
        kw["curl_seq"] = [r.choice(["starting\n", "", "503 Service Unavailable\n", "service unavailable", "KuBeRnEtEs", "* ?\n"])
                          for _ in range(r.randint(0, 4))]
    if r.random() < 0.3:
        kw["known_hosts"] = "x\n"
    if r.random() < 0.3:
        kw["keygen_no_E"] = True
    if r.random() < 0.3:
        kw["keygen"] = {"id_rsa": r.choice(["2048 MD5:aa:bb:cc:dd c (RSA)\n", "2048 MD5:zz c\n", "", "one\n", "a b c\nd e f\n", "2048 * c\n"])}
        kw["keygen_sha"] = {"id_rsa": r.choice(["2048 SHA256:zz c\n", "2048 aa:bb:cc:dd c\n", "", "x SHA256\n"])}
    if r.random() < 0.3 and not sc.stdin.startswith("yes"):
        kw["stdin"] = "\n" * 4 + r.choice(["", "1", "2", "3", "9"]) + "\n\n\n\nyes\n" + sc.stdin
    if r.random() < 0.2:
        kw["terraform_no_ips"] = True
    if r.random() < 0.15:
        kw["invocation"] = "bash setup.sh"
    if r.random() < 0.1:
        kw["home_name"] = r.choice(["my home", "h*me", "h[o]me"])
    if r.random() < 0.1:
        kw["unset_env"] = [r.choice(["LC_ALL", "SHLVL"])]
    return dataclasses.replace(sc, **kw)


# --- and one for the state of the workspace: files, directories and links in the way ------------
_PATHS = ["config", "config.tmp", "tmp", "terraform/rancher.tf", "terraform/masters.ip", "terraform/hosts.ip", "terraform/terraform.tfstate",
          "terraform/.terraform", "ansible/hosts", "ansible/ansible.cfg", "ansible/tmp.cfg", "ansible/ansible.cfg.tmp", "ansible/a.retry",
          "ansible/roles/ranchermaster/vars/vars.yml", "ansible/tmp/kubernetes_environment.id", "ansible/tmp", "tmp/x"]
#: link targets stay inside the sandbox or point nowhere: the reference's `rm -rf tmp/*` goes *through* a link to a
#: directory, and a target such as /tmp would have it empty that directory for real
_TARGETS = ["/dev/null", "nowhere", "../config", "clusterUp.yml", "../ansible/ansible.cfg", "hosts", "/nonexistent/x", "ansible",
            "../terraform", "roles", "ansible.cfg"]
_CFGS = ["[defaults]\nprivate_key_file = old\n", "private_key_file = a\nprivate_key_file = b", "", "x\n",
         "[defaults]\r\nprivate_key_file = q\r\n", "# private_key_file = c\n private_key_file = d\n"]


def make_workspace(seed: int) -> Scenario:
    r = random.Random(seed)
    kw = dict(name="ws%d" % seed, args=["-c"] if r.random() < 0.4 else [],
              stdin=r.choice(["yes\n", "no\n", "yes\nyes\n", "\n" * 8 + "yes\n", "yes\n" + "\n" * 8 + "yes\n", "\n\n\n\n2\n\n\n\nyes\nyes\n"]))
    pre_files, pre_remove, pre_symlinks = {}, [], {}
    for _ in range(r.randint(0, 4)):
        p, k = r.choice(_PATHS), r.random()
        if k < 0.35:
            pre_files[p] = r.choice(_CFGS) if p.endswith("cfg") else r.choice(["x\n", "", "10.9.9.9\n", "a b\n*\n", "1a9"])
        elif k < 0.5:
            pre_files[p + "/keep"] = "k\n"
        elif k < 0.7:
            pre_remove.append(p)
        else:
            pre_symlinks[p] = r.choice(_TARGETS)
    taken = pre_remove + list(pre_symlinks)
    pre_files = {p: c for p, c in pre_files.items() if not any(p == q or p.startswith(q + "/") or q.startswith(p + "/") for q in taken)}
    pre_files = {p: c for p, c in pre_files.items() if not any(q != p and q.startswith(p + "/") for q in pre_files)}
    pre_symlinks = {p: t for p, t in pre_symlinks.items() if not any(p != q and p.startswith(q + "/") for q in taken)}
    kw.update(pre_files=pre_files, pre_remove=pre_remove, pre_symlinks=pre_symlinks)
    if r.random() < 0.3:
        kw["known_hosts"] = "h\n"
    if r.random() < 0.2:
        kw["env_id"] = "1b2"
    if r.random() < 0.15:
        kw["fail"] = {r.choice(["terraform-destroy", "terraform-apply", "ssh-keygen--R", "ansible-playbook--i"]): 1}
    return Scenario(**kw)


def _sweep(ref_root, generator, seeds):
    bad = []
    for seed in seeds:
        try:
            sc = generator(seed)
            theirs = run_reference(sc, ref_root)
            mine = run_in_sandbox(rp.REPLACEMENT_ARGV, ref_root, sc, extra_env=rp.REPLACEMENT_ENV)
        except OSError:
            continue        # the generator asked for a workspace that cannot be laid down (a file inside a link ...)
        diffs = theirs.diff(mine)
        if diffs:
            bad.append((seed, sc.to_json(), diffs))
    return bad


@pytest.mark.parametrize("block", range(2))
def test_property_fuzzed_tools(ref_root, block):
    assert _sweep(ref_root, make_wide, range(20000 + 12 * block, 20000 + 12 * (block + 1))) == []


@pytest.mark.parametrize("block", range(3))
def test_property_fuzzed_workspaces(ref_root, block):
    assert _sweep(ref_root, make_workspace, range(30000 + 12 * block, 30000 + 12 * (block + 1))) == []


# --- and one for the shape of what `triton networks|packages` prints ---------------------------
_TCH = list("abK-=_ .*?[]\t\r0123()|@+!^:") + ["-kvm-", "  ", "NAME", "ID", "Joyent-SDC-Public", "k4-highcpu-kvm-7.75G", "é", "\\"]


def _odd_table(r):
    def word(k=4):
        return "".join(r.choice(_TCH) for _ in range(r.randint(1, k)))
    lines = [r.choice(["NAME  ID", "NAME ID", "NAME    ID  ", "name id", "", "NAMEID", "NAME\tID"])]
    for _ in range(r.randint(0, 7)):
        shape = r.choice(["{}  {}", "{} {}", "{}\t{}", "{}", " {}  {} ", "{}  {}  extra"])
        lines.append(shape.format(word(), word())[:60])
    if r.random() < 0.2:
        lines.insert(r.randint(0, len(lines)), "")
    return "\n".join(lines) + r.choice(["\n", "", "\n\n"])


def make_tables(seed: int) -> Scenario:
    r = random.Random(seed)
    answers = ["", "", "", "", r.choice(["", "2"]), r.choice(["", "1", "2,1", "3", "1,2,3"]), r.choice(["", "1", "2"]),
               r.choice(["", "1", "2", "3"]), "", "", "", "yes", "yes"]
    return Scenario(name="tb%d" % seed, stdin="".join(a + "\n" for a in answers), networks=_odd_table(r), packages=_odd_table(r))


@pytest.mark.parametrize("block", range(2))
def test_property_fuzzed_tables(ref_root, block):
    """A default that is a pattern matching nothing (an id such as ``[a-``) makes the
    reference re-prompt for ever once stdin is exhausted; then both must spin."""
    import subprocess
    from tritonk8ssupervisor_b200.conformance.sandbox import Sandbox
    bad = []
    for seed in range(3012 + 8 * block, 3012 + 8 * (block + 1)):
        sc = make_tables(seed)
        res = []
        for argv, env in ((sc.reference_argv(), None), (rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV)):
            with Sandbox(ref_root, sc) as sbx:
                try:
                    res.append(sbx.run(argv, env, timeout=10))
                except subprocess.TimeoutExpired:
                    res.append(None)
        if res[0] is None or res[1] is None:
            if (res[0] is None) != (res[1] is None):
                bad.append((seed, "only one of the two never ends", sc.to_json()))
        elif res[0].diff(res[1]):
            bad.append((seed, res[0].diff(res[1]), sc.to_json()))
    assert bad == []


# --- NUL bytes, control characters and stray high bytes in what the tools print ----------------
def _inject(r, text):
    if not text:
        return text
    k = r.randint(0, len(text))
    return text[:k] + r.choice(["\0", "\0\0", "\x01", "\udcff", "\r"]) + text[k:]


def make_binary(seed: int) -> Scenario:
    import dataclasses
    sc = make_wide(seed)
    r = random.Random(seed * 13 + 5)
    kw = {}
    for field in ("networks", "packages", "triton_env"):
        if r.random() < 0.35:
            kw[field] = _inject(r, getattr(sc, field))
    if r.random() < 0.3:
        kw["master_ip"] = _inject(r, sc.master_ip or "10.0.0.1")
    if r.random() < 0.3:
        kw["env_id"] = _inject(r, sc.env_id or "1a7")
    if r.random() < 0.4:
        kw["curl_seq"] = [_inject(r, b) for b in (sc.curl_seq or ["starting\n", "kubernetes\n"])]
    if r.random() < 0.4:
        kw["ssh_ps"] = {k: _inject(r, v) for k, v in (sc.ssh_ps or {"10.0.1.1": _PS[0]}).items()}
    if r.random() < 0.3:
        kw["keygen"] = {k: _inject(r, v) for k, v in sc.keygen.items()}
    if r.random() < 0.2:
        kw["date_seq"] = [_inject(r, d) for d in (sc.date_seq or ["1500000000"])]
    return dataclasses.replace(sc, **kw)


@pytest.mark.parametrize("block", range(2))
def test_property_fuzzed_binary_output(ref_root, block):
    assert _sweep(ref_root, make_binary, range(12 * block, 12 * (block + 1))) == []


# --- the same scenarios in a UTF-8 locale (grep, sed and bash's matcher all look at characters then) ---
def make_utf8(seed: int) -> Scenario:
    import dataclasses
    sc = (make, make_wide, make_binary)[seed % 3](seed)
    return dataclasses.replace(sc, env=dict(sc.env, LANG="C.UTF-8"), unset_env=["LC_ALL"])


@pytest.mark.parametrize("block", range(2))
def test_property_fuzzed_in_a_utf8_locale(ref_root, block):
    assert _sweep(ref_root, make_utf8, range(500 + 12 * block, 500 + 12 * (block + 1))) == []


# --- answers and table names full of echo's escapes, in both locales -----------------------------------
_ESC = ["-e ", "-e ", "-E ", "-ne ", "-en ", "-eE ", "-n ", "- ", "-ex ", "\\\\u00e9", "\\\\u65e5", "\\\\U0001F600", "\\\\u41", "\\\\ud800", "\\\\u",
        "\\\\x41", "\\\\xe9", "\\\\xc3\\\\xa9", "\\\\0101", "\\\\0351", "\\\\101", "\\\\c", "\\\\e", "\\\\E", "\\\\a", "\\\\b", "\\\\f", "\\\\v", "\\\\r",
        "\\\\n", "\\\\t", "\\\\\\\\", "\\\\", "\\", "a", "1", " ", "é", "\udcff", "x", "2", ","]
_ESC_NAMES = ["caf\\u00e9", "a\\x41b", "t\\tb", "n\\nl", "c\\cz", "o\\0101", "b\\\\s", "e\\e[1m", "u\\U0001F600", "bad\\xzz", "\\u41-kvm-x",
              "k\\u00e9-kvm-1"]


def make_escapes(seed: int) -> Scenario:
    import dataclasses
    sc = make(seed)
    r = random.Random(seed * 31 + 7)
    lines = ["".join(r.choice(_ESC) for _ in range(r.randint(1, 6))) if r.random() < 0.6 else r.choice(["", "1", "2", "1,2", "3"])
             for _ in range(r.randint(3, 14))]
    kw = {"stdin": "\n".join(lines) + "\n" + r.choice(["yes\n", "", "yes"])}
    if r.random() < 0.5:
        kw["networks"] = triton_table(list(DOC_NETWORKS)[:2] + [(r.choice(_ESC_NAMES), "id-%d" % i) for i in range(r.randint(1, 3))])
    if r.random() < 0.5:
        kw["packages"] = triton_table(list(DOC_PACKAGES)[:2] + [(r.choice(_ESC_NAMES) + "-kvm-", "pid-%d" % i) for i in range(r.randint(1, 3))])
    if seed % 2:
        kw["env"] = dict(sc.env, LANG="C.UTF-8")
        kw["unset_env"] = ["LC_ALL"]
    return dataclasses.replace(sc, **kw)


@pytest.mark.parametrize("block", range(2))
def test_property_fuzzed_echo_escapes(ref_root, block):
    assert _sweep(ref_root, make_escapes, range(12 * block, 12 * (block + 1))) == []


# --- answers that leave a word echo (and then sed) takes for options in the variable --------------------
_OPTS = ["e", "n", "E", "ne", "en", "eE", "nE", "Ee", "nn", "neE", "Ene", "x", "", "-", "ex", "e "]      # not "ee": refused (sed would execute)
_OPT_WORDS = ["1", "2", "3", "14", "1,2", "2,1,2", "1.5", "/b/", "$", "x", "08", "i", "0", "99", "19", "-1", "1 2", "*", "a b", "", "tmp",
              "kube", "-e", "-n"]
_OPT_SEP = [" ", "\\\\\\\\n", "\\\\\\\\t", ",", "\\\\\\\\c", "\\\\n"]


def make_option_words(seed: int) -> Scenario:
    import dataclasses
    sc = make(seed)
    r = random.Random(seed * 17 + 3)

    def answer():
        k = r.random()
        if k < 0.25:
            return r.choice(["", "", "1", "2", "1,2", "km"])
        words = [r.choice(_OPT_WORDS) for _ in range(r.randint(1, 3))]
        payload = words[0] + "".join(r.choice(_OPT_SEP) + w for w in words[1:])
        if k < 0.85:
            return "-e \\\\055" + r.choice(_OPTS) + " " + payload           # read and the first echo leave "-<opts> <payload>"
        return r.choice(["-e ", "-n ", "-E ", "-ne ", ""]) + payload
    kw = {"stdin": "\n".join(answer() for _ in range(r.randint(6, 14))) + "\n" + r.choice(["yes\n", "", "yes", "yes\nyes\n"])}
    if seed % 3 == 0:
        kw["env"] = dict(sc.env, LANG="C.UTF-8")
        kw["unset_env"] = ["LC_ALL"]
    return dataclasses.replace(sc, **kw)


@pytest.mark.parametrize("block", range(2))
def test_property_fuzzed_option_words(ref_root, block):
    assert _sweep(ref_root, make_option_words, range(12 * block, 12 * (block + 1))) == []


@pytest.mark.parametrize("word", ["-p", "-n X", "-f", "--", "-x", "-n", "-", "-fn", "--help", "--x", "-a", "master", "A B=2"])
def test_config_line_that_export_takes_for_options(ref_root, word):
    """A description of three lines leaves WORD as a line of ./config, and setup.sh:546 runs `export "WORD"`."""
    for stdin in ("-e a\\\\n%s\n\n\n\n\n\n\n\n\nyes\n" % word, "\n-e a\\\\n%s\\\\nQ=\n\n\n\n\n\n\nyes\n" % word):
        sc = Scenario(name="exportword", stdin=stdin)
        theirs = run_reference(sc, ref_root)
        mine = run_in_sandbox(rp.REPLACEMENT_ARGV, ref_root, sc, extra_env=rp.REPLACEMENT_ENV)
        assert theirs.diff(mine) == []


@pytest.mark.parametrize("line,why", [("RANDOM=4", "bash computes by itself"), ("HOSTNAME", "bash computes by itself"),
                                      ("IFS=,", "changes how bash reads the rest of the script"),
                                      ("LC_ALL=C.UTF-8", "changes how bash reads the rest of the script")])
def test_config_line_that_is_refused(ref_root, line, why):
    """Exports whose effect on bash itself (or on what children inherit) is not restated stop the run with status 2."""
    sc = Scenario(name="exportrefused", stdin="\n-e a\\\\n%s\\\\nQ=\n\n\n\n\n\n\nyes\n" % line)
    mine = run_in_sandbox(rp.REPLACEMENT_ARGV, ref_root, sc, extra_env=rp.REPLACEMENT_ENV)
    assert mine.exit_status == 2 and why in mine.stderr and "refusing to go on: a line of ./config exports" in mine.stderr
    assert not [c for c in mine.calls if c.startswith("terraform")]


def test_cdpath_that_leaves_the_workspace_is_refused(ref_root, tmp_path):
    (tmp_path / "elsewhere" / "terraform").mkdir(parents=True)
    sc = Scenario(name="cdpathaway", stdin="\n" * 8 + "yes\n", env={"CDPATH": str(tmp_path / "elsewhere")})
    mine = run_in_sandbox(rp.REPLACEMENT_ARGV, ref_root, sc, extra_env=rp.REPLACEMENT_ENV)
    assert mine.exit_status == 2 and "CDPATH takes `cd terraform` to" in mine.stderr


# --- answers of several lines: each extra line becomes a line of ./config, i.e. an argument of `export` ---
_CFG_PIECES = ["A", "B", "i", "=", "=", "+=", "[0]", "[", "]", " ", "-", "-p", "-n", "-f", "--", "x", "1", "_", "$A", "$(id)", "`id`", "*", "?", "'",
               '"', "\\\\\\\\", "#", ";", "&", "|", "~", "é", "\udcff", "KUBERNETES_NAME", "KUBERNETES_NUMBER_OF_NODES=2", "HOME=/x", "PWD",
               "OLDPWD=", "SHLVL=x", "master", "network=", "tmp", "CDPATH=.", "UID", "OPTIND=", "\\\\\\\\t", "\\\\\\\\x41", "-e", "\\\\\\\\055"]


def make_config_lines(seed: int) -> Scenario:
    import dataclasses
    sc = make(seed)
    r = random.Random(seed * 23 + 11)
    answers = [""] * 9
    for _ in range(r.randint(1, 3)):
        lines = ["".join(r.choice(_CFG_PIECES) for _ in range(r.randint(1, 5))) for _ in range(r.randint(1, 3))]
        answers[r.choice([0, 1, 2, 3])] = "-e a" + "".join("\\\\n" + line for line in lines) + r.choice(["", "\\\\nQ="])
    kw = {"stdin": "\n".join(answers) + "\n" + r.choice(["yes\n", "yes"])}
    if seed % 3 == 0:
        kw["env"] = dict(sc.env, LANG="C.UTF-8")
        kw["unset_env"] = ["LC_ALL"]
    return dataclasses.replace(sc, **kw)


@pytest.mark.parametrize("block", range(2))
def test_property_fuzzed_config_lines(ref_root, block):
    assert _sweep(ref_root, make_config_lines, range(70 + 12 * block, 70 + 12 * (block + 1))) == []


def test_exported_function_named_like_a_command_is_refused(ref_root):
    sc = Scenario(name="funcshadow", stdin="\n" * 8 + "yes\n", env={"BASH_FUNC_sleep%%": "() {  echo zzz\n}", "BASH_FUNC_ml%%": "() {  :\n}"})
    mine = run_in_sandbox(rp.REPLACEMENT_ARGV, ref_root, sc, extra_env=rp.REPLACEMENT_ENV)
    assert mine.exit_status == 2 and "exports a shell function named sleep" in mine.stderr and mine.calls == []
    sc = Scenario(name="funcother", stdin="\n" * 8 + "yes\n", env={"BASH_FUNC_ml%%": "() {  :\n}"})
    theirs = run_reference(sc, ref_root)
    mine = run_in_sandbox(rp.REPLACEMENT_ARGV, ref_root, sc, extra_env=rp.REPLACEMENT_ENV)
    assert theirs.diff(mine) == []


# --- locale variables that disagree, or name locales that are not installed ------------------------------
_LOCALES = [{"LC_CTYPE": "xx", "LANG": "C.UTF-8"}, {"LC_TIME": "xx", "LANG": "C.UTF-8"}, {"LC_ALL": "xx.UTF-8"}, {"LANG": "xx.UTF-8"},
            {"LC_CTYPE": "C.UTF-8", "LANG": "xx"}, {"LC_ALL": "xx_XX.ISO-8859-1"}, {"LC_CTYPE": "C.UTF-8"}, {"LC_COLLATE": "C", "LANG": "C.UTF-8"},
            {"LC_CTYPE": "C", "LANG": "C.UTF-8"}, {"LC_ALL": "C.UTF-8", "LANG": "xx"}]


def make_locales(seed: int) -> Scenario:
    import dataclasses
    gens = (make, make_wide, make_binary, make_escapes, make_option_words, make_config_lines)
    sc = gens[seed % len(gens)](seed)
    loc = _LOCALES[(seed // len(gens)) % len(_LOCALES)]
    env = {k: v for k, v in sc.env.items() if not k.startswith(("LC_", "LANG"))}
    env.update(loc)
    return dataclasses.replace(sc, env=env, unset_env=[] if "LC_ALL" in loc else ["LC_ALL"])


@pytest.mark.parametrize("block", range(3))
def test_property_fuzzed_locale_settings(ref_root, block):
    assert _sweep(ref_root, make_locales, range(20 * block, 20 * (block + 1))) == []


def test_inherited_variables_that_reconfigure_bash(ref_root, tmp_path):
    """Refused: what would put the reference's bash into another mode.  Followed: what makes no difference."""
    (tmp_path / "rc").write_text("echo sourced\n")
    for env in ({"POSIXLY_CORRECT": "1"}, {"POSIXLY_CORRECT": ""}, {"SHELLOPTS": "xtrace"}, {"BASHOPTS": "nullglob"}, {"BASH_ENV": str(tmp_path / "rc")},
                {"BASH_ENV": "~/.bashenv"}):
        sc = Scenario(name="reconf", stdin="\n" * 8 + "yes\n", env=env)
        mine = run_in_sandbox(rp.REPLACEMENT_ARGV, ref_root, sc, extra_env=rp.REPLACEMENT_ENV)
        assert mine.exit_status == 2 and "refusing to go on" in mine.stderr and mine.calls == [] and mine.stdout == "", env
    sc = Scenario(name="noreconf", stdin="\n\n\n\n\n*\n\n\n\nyes\n", env={"GLOBIGNORE": "*.tf:ansible", "BASH_ENV": "/nonexistent", "ENV": "/x", "BASH_COMPAT": "4.4",
                                                                         "TMOUT": "1", "FUNCNEST": "3", "EXECIGNORE": "*/sleep"})
    theirs = run_reference(sc, ref_root)
    mine = run_in_sandbox(rp.REPLACEMENT_ARGV, ref_root, sc, extra_env=rp.REPLACEMENT_ENV)
    assert theirs.diff(mine) == []
