"""Ansible structural checker and inventory renderer."""
import os
import shutil

import pytest

from tritonk8ssupervisor_b200 import ansiblecheck, inventory
from tritonk8ssupervisor_b200.conformance import replay as rp
from tritonk8ssupervisor_b200.conformance.catalog import CONFIGS
from tritonk8ssupervisor_b200.conformance.harness import WORKSPACE_TEMPLATE

HOSTS = "[MASTER]\n10.0.0.1\n[HOST]\n10.0.1.1\n10.0.1.2\n"


@pytest.fixture
def ans_dir(tmp_path):
    shutil.copytree(os.path.join(WORKSPACE_TEMPLATE, "ansible"), tmp_path / "ansible")
    (tmp_path / "ansible" / "hosts").write_text(HOSTS)
    return tmp_path / "ansible"


def _check(ans_dir):
    inv = inventory.parse_file(str(ans_dir / "hosts"))
    return ansiblecheck.check_playbook(str(ans_dir / "clusterUp.yml"), inv)


def test_package_playbook_is_clean(ans_dir):
    rep = _check(ans_dir)
    assert rep.findings == [], [str(f) for f in rep.findings]
    assert ansiblecheck.check_ansible_cfg(str(ans_dir / "ansible.cfg")).findings == []


@pytest.mark.parametrize("scenario", CONFIGS, ids=lambda s: s.name)
def test_reference_playbook_with_generated_files(tmp_path, scenario, ref_root):
    """The reference's playbook + roles with the generated inventory/vars pass,
    with exactly the era warnings of SURVEY.md §7.4 / §3.4."""
    shutil.copytree(os.path.join(ref_root, "ansible"), tmp_path / "ansible")
    doc = rp.load_golden(os.path.join(rp.golden_dir(), scenario.name + ".json"))
    for rel, content in doc["expected"]["files"].items():
        if rel.startswith("ansible/"):
            (tmp_path / rel).write_text(content)
    rep = _check(tmp_path / "ansible")
    assert rep.ok, [str(f) for f in rep.errors]
    assert sorted(set(rep.codes("warning"))) == ["magic-shadowed", "when-braces"]
    assert rep.codes("warning").count("magic-shadowed") == 2      # clusterUp.yml:11-12,21-22
    assert rep.codes("warning").count("when-braces") == 8         # ranchermaster 4 + rancherhost 4
    assert ansiblecheck.check_ansible_cfg(str(tmp_path / "ansible" / "ansible.cfg")).ok


def _mutate(ans_dir, rel, old, new):
    p = ans_dir / rel
    text = p.read_text()
    assert old in text, (rel, old)
    p.write_text(text.replace(old, new, 1))


MASTER = "roles/ranchermaster/tasks/main.yml"
HOSTR = "roles/rancherhost/tasks/main.yml"
DOCKER = "roles/dockersetup/tasks/main.yml"
MUTATIONS = [
    ("clusterUp.yml", "  hosts: MASTER\n", "", "play-shape"),
    ("clusterUp.yml", "  remote_user: root\n  roles:\n    - dockersetup", "  remote_usr: root\n  roles:\n    - dockersetup", "play-shape"),
    ("clusterUp.yml", "    - rancherhost", "    - nosuchrole", "role-missing"),
    ("clusterUp.yml", "- name: docker-engine", "name: docker-engine", "yaml"),
    ("clusterUp.yml", "roles/ranchermaster/vars/vars.yml\"\n  roles:\n    - ranchermaster", "roles/ranchermaster/vars/missing.yml\"\n  roles:\n    - ranchermaster", "vars-file"),
    (MASTER, "  docker_container:\n    name: master", "  docker_contaner:\n    name: master", "task-keyword"),
    (MASTER, "  register: containers\n", "  registr: containers\n", "task-keyword"),
    (MASTER, "  command: docker ps -a\n", "", "task-shape"),
    (MASTER, "  command: docker ps -a\n", "  command: docker ps -a\n  shell: true\n", "task-shape"),
    (MASTER, "    url: \"http://{{ master }}:8080/v2-beta/projectTemplates?name=kubernetes\"\n", "", "module-arg-missing"),
    (MASTER, "{{ kubernetes_name }}", "{{ kubernetes_name ", "jinja-syntax"),
    (MASTER, "{{ kubernetes_name }}", "{{ kubernetes_nme }}", "var-undefined"),
    (MASTER, "'Listening on' in rancher_logs.stdout", "'Listening on' in rancher_logs.stdout and", "jinja-syntax"),
    (HOSTR, "b64decode", "b46decode", "jinja-filter"),
    (HOSTR, "  register: project_id\n", "", "var-undefined"),
    (HOSTR, "{{ master }}", "{{ item }}", "var-undefined"),
    (DOCKER, "  apt_repository:\n    repo: deb https://apt.dockerproject.org/repo ubuntu-xenial main\n", "  apt_repository:\n", "module-arg-missing"),
    (DOCKER, "  pip:\n    name: docker-py\n", "  pip:\n    state: present\n", "module-arg-missing"),
    (DOCKER, "  shell: docker --version", "  shell: \"docker --version", "yaml"),
]


@pytest.mark.parametrize("rel,old,new,code", MUTATIONS, ids=[m[3] + "-" + str(i) for i, m in enumerate(MUTATIONS)])
def test_breakage_is_rejected(ans_dir, rel, old, new, code):
    _mutate(ans_dir, rel, old, new)
    rep = _check(ans_dir)
    assert not rep.ok
    assert code in rep.codes("error"), [str(f) for f in rep.findings]


def test_generated_vars_file_breakage(ans_dir):
    (ans_dir / "roles/ranchermaster/vars/vars.yml").write_text('master: 10.0.0.1\nkubernetes_name: "a"b"\n')
    assert "yaml" in _check(ans_dir).codes("error")
    (ans_dir / "roles/ranchermaster/vars/vars.yml").write_text("master: 10.0.0.1\n")
    assert "var-undefined" in _check(ans_dir).codes("error")


def test_warnings(ans_dir):
    _mutate(ans_dir, DOCKER, "    upgrade: dist\n", "    upgrade: dist\n    colour: red\n")
    _mutate(ans_dir, "clusterUp.yml", "  hosts: HOST\n", "  hosts: WORKERS\n")
    _mutate(ans_dir, MASTER, "    state: started\n", "    state: started\n    state: present\n")
    rep = _check(ans_dir)
    assert rep.ok
    assert {"module-arg-unknown", "hosts-unmatched", "yaml-duplicate-key"} <= set(rep.codes("warning"))


def test_register_on_other_hosts_is_flagged(ans_dir):
    _mutate(ans_dir, HOSTR, "{{ master }}:8080/v1", "{{ kubernetes_environment_id.json.id }}:8080/v1")
    assert "register-other-hosts" in _check(ans_dir).codes("warning")


def test_parse_kv():
    kv = ansiblecheck.parse_kv('method=POST status_code=201 url="http://{{ m }}:8080/x?projectId={{ p[\'content\'] | b64decode | replace(\'\\n\', \'\') }}" return_content=yes', False)
    assert kv["method"] == "POST" and kv["return_content"] == "yes"
    assert kv["url"].startswith("http://{{ m }}:8080/x?projectId={{ p['content']") and kv["url"].endswith("}}")
    assert ansiblecheck.parse_kv("docker logs master", True) == {"_raw_params": "docker logs master"}
    assert ansiblecheck.parse_kv("echo a=b chdir=/tmp", True) == {"_raw_params": "echo a=b chdir=/tmp"}
    assert ansiblecheck.parse_kv("chdir=/tmp echo hi", True) == {"chdir": "/tmp", "_raw_params": "echo hi"}
    with pytest.raises(ValueError):
        ansiblecheck.parse_kv("name=x stray", False)
    with pytest.raises(ValueError):
        ansiblecheck.parse_kv('name="x', False)


# --- inventory ---------------------------------------------------------------
def test_inventory_render_of_generated_hosts():
    doc = inventory.parse_ini(HOSTS).to_list_json()
    assert doc == {"_meta": {"hostvars": {"10.0.0.1": {}, "10.0.1.1": {}, "10.0.1.2": {}}},
                   "all": {"children": ["HOST", "MASTER", "ungrouped"]},
                   "HOST": {"hosts": ["10.0.1.1", "10.0.1.2"]}, "MASTER": {"hosts": ["10.0.0.1"]}}


def test_inventory_features():
    inv = inventory.parse_ini("""
lonely ansible_user=admin
# comment
[web]
web[01:03].example.com http_port=80 flag=True
alpha:2222 name='quoted value'
[db]
db-[a:b]
[prod:children]
web
db
[prod:vars]
env=prod
[web:vars]
tier = front
""")
    assert inv.hosts_of("web") == ["web01.example.com", "web02.example.com", "web03.example.com", "alpha"]
    assert inv.hosts_of("prod") == inv.hosts_of("web") + ["db-a", "db-b"]
    assert inv.groups["ungrouped"].hosts == ["lonely"]
    hv = inv.effective_hostvars("alpha")
    assert hv == {"env": "prod", "tier": "front", "ansible_port": 2222, "name": "quoted value"}
    assert inv.effective_hostvars("web02.example.com")["http_port"] == 80
    assert inv.effective_hostvars("web02.example.com")["flag"] is True
    doc = inv.to_list_json()
    assert doc["all"]["children"] == ["prod", "ungrouped"]
    assert doc["prod"] == {"children": ["db", "web"]}
    assert doc["ungrouped"] == {"hosts": ["lonely"]}


@pytest.mark.parametrize("text", ["[web\nh1\n", "[g:nope]\nh\n", "[g]\nh1 stray\n", "[g:vars]\nnovalue\n", "[g]\nh[3:1]\n", "[g]\nh[1:b]\n"])
def test_inventory_errors(text):
    with pytest.raises(inventory.InventoryError):
        inventory.parse_ini(text)


def test_blocks_are_walked(ans_dir):
    (ans_dir / "roles/rancherhost/tasks/main.yml").write_text("""---
- name: guarded group
  when: "'x' in master"
  block:
    - name: inside
      command: docker ps -a
      register: inner
    - debug:
        msg: "{{ inner.stdout }} {{ undefined_in_block }}"
  rescue:
    - debug:
        msg: "{{ inner.rc }}"
  always:
    - name: typo
      comand: echo
""")
    rep = _check(ans_dir)
    codes = rep.codes("error")
    assert codes.count("var-undefined") == 1 and "task-keyword" in codes
    assert any("undefined_in_block" in f.message for f in rep.errors)
