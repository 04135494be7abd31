"""Robustness: whatever bytes they are fed, the parsers either succeed or raise
their own error type -- a checker that dies with a traceback on a malformed
generated file would turn a 'fail' into a crash."""
from hypothesis import given, settings, strategies as st

from tritonk8ssupervisor_b200 import ansiblecheck, hcl1, inventory, sedprog, tritonenv

_hclish = st.text(st.sampled_from(list('ab1 ={}[]"$\\\n#/*,.-_<EOT')), max_size=60)
_iniish = st.text(st.sampled_from(list("ab1 =[]:#;\n'\"-.")), max_size=60)
_kvish = st.text(st.sampled_from(list("ab= \"'{}%|\\")), max_size=40)
_shish = st.text(st.sampled_from(list("exportunsAB=\"'$`;#\n \\()|")), max_size=50)


@settings(max_examples=400, deadline=None, derandomize=True)
@given(_hclish)
def test_hcl(src):
    try:
        hcl1.parse(src)
    except hcl1.HCLError:
        pass


@settings(max_examples=300, deadline=None, derandomize=True)
@given(_hclish)
def test_hil_template(src):
    try:
        hcl1.parse_template(src)
    except hcl1.HCLError:
        pass


@settings(max_examples=300, deadline=None, derandomize=True)
@given(_iniish)
def test_inventory(src):
    try:
        inventory.parse_ini(src).to_list_json()
    except inventory.InventoryError:
        pass


@settings(max_examples=300, deadline=None, derandomize=True)
@given(_kvish, st.booleans())
def test_parse_kv(src, free):
    try:
        ansiblecheck.parse_kv(src, free)
    except ValueError:
        pass


@settings(max_examples=300, deadline=None, derandomize=True)
@given(_shish)
def test_triton_env(src):
    try:
        tritonenv.apply_triton_env(src, {}, {})
    except tritonenv.TritonEnvError:
        pass


@settings(max_examples=300, deadline=None, derandomize=True)
@given(st.text(st.sampled_from(list("ab;&\\gpw123 #}\nIx")), max_size=12), st.sampled_from([";", "~"]))
def test_sedprog(value, delim):
    res = sedprog.run("[defaults]\nprivate_key_file = old\n", value, delim)
    assert res.ok or res.refused or res.error.startswith("sed: -e expression #1, char ")


_scalar = st.one_of(st.none(), st.booleans(), st.integers(-3, 3),
                    st.sampled_from(["all", "MASTER", "x", "{{ a }}", "{{ a", "a=b c", "docker ps", "roles", "yes"]))
_key = st.sampled_from(["name", "hosts", "roles", "tasks", "vars", "vars_files", "when", "register", "command", "uri", "url",
                        "with_items", "block", "action", "local_action", "module", "until", "apt", "debug", "msg", "role", "x"])
_node = st.recursive(_scalar, lambda kids: st.one_of(st.lists(kids, max_size=3), st.dictionaries(_key, kids, max_size=4)),
                     max_leaves=12)


@settings(max_examples=300, deadline=None, derandomize=True)
@given(_node, _node)
def test_check_playbook_on_arbitrary_yaml(tmp_path_factory, playbook, role_tasks):
    import yaml
    d = tmp_path_factory.mktemp("pb")
    (d / "roles" / "x" / "tasks").mkdir(parents=True)
    (d / "roles" / "x" / "tasks" / "main.yml").write_text(yaml.safe_dump(role_tasks))
    (d / "clusterUp.yml").write_text(yaml.safe_dump(playbook))
    inv = inventory.parse_ini("[MASTER]\n10.0.0.1\n")
    report = ansiblecheck.check_playbook(str(d / "clusterUp.yml"), inv)
    assert isinstance(report.ok, bool)
