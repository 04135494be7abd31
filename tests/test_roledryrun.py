"""the roles dry-run against simulated machines and the
Rancher REST mock.  The package's own roles and (where available) the
reference's roles must put the *same* requests on the wire and leave the same
end state."""
import json
import os
import shutil

import pytest

from tritonk8ssupervisor_b200 import inventory, ranchermock, roledryrun
from tritonk8ssupervisor_b200.conformance.harness import WORKSPACE_TEMPLATE

HOSTS = "[MASTER]\n10.0.0.1\n[HOST]\n10.0.1.1\n10.0.1.2\n"
VARS = 'master: 10.0.0.1\nkubernetes_name: "k8s dev"\nkubernetes_description: "my quoted desc"\n'
WIRE = ["GET /v2-beta/projectTemplates?name=kubernetes", "POST /v2-beta/projects",
        "POST /v1/registrationtokens?projectId=1a7", "POST /v1/registrationtokens?projectId=1a7",
        "GET /v1/registrationtokens/1c1", "GET /v1/registrationtokens/1c2"]


def _tree(tmp_path, src):
    d = tmp_path / "ansible"
    shutil.copytree(src, d)
    (d / "hosts").write_text(HOSTS)
    (d / "roles/ranchermaster/vars/vars.yml").write_text(VARS)
    (d / "tmp").mkdir(exist_ok=True)
    env_id = d / "tmp" / "kubernetes_environment.id"
    if env_id.exists():
        env_id.unlink()
    return d


def _run(d, mock, **kw):
    inv = inventory.parse_file(str(d / "hosts"))
    return roledryrun.DryRun(str(d / "clusterUp.yml"), inv, {"10.0.0.1:8080": mock.address}, **kw).run()


def _sources():
    from conftest import reference_root
    out = [("package", os.path.join(WORKSPACE_TEMPLATE, "ansible"))]
    ref = reference_root()
    if ref:
        out.append(("reference", os.path.join(ref, "ansible")))
    return out


@pytest.mark.parametrize("label,src", _sources(), ids=[s[0] for s in _sources()])
def test_full_bring_up(tmp_path, label, src):
    d = _tree(tmp_path, src)
    with ranchermock.RancherMock() as mock:
        run = _run(d, mock)
        assert run.failed_hosts == {}
        assert mock.wire_log() == WIRE
        # 2. the project body (setup-generated name/description end up here)
        (body,) = mock.projects.values()
        assert body == {"description": "my quoted desc", "name": "k8s dev", "projectTemplateId": "1pt5",
                        "allowSystemRole": False, "members": [], "virtualMachine": False,
                        "servicesPortRange": None, "projectLinks": []}
        post = [r for r in mock.requests if r["path"] == "/v2-beta/projects"][0]
        assert post["headers"]["content-type"] == "application/json" and post["headers"]["accept"] == "application/json"
    # the id is kept on the controller for the node play and for setup's banner (setup.sh:53)
    assert (d / "tmp/kubernetes_environment.id").read_text() == "1a7\n"
    master, n1, n2 = (run.hosts[h] for h in ("10.0.0.1", "10.0.1.1", "10.0.1.2"))
    assert [c["name"] for c in master.containers] == ["master"]
    assert master.containers[0]["image"] == "rancher/server:stable" and master.containers[0]["published_ports"] == "8080:8080"
    # 5. each agent is started with the registration URL of its own token
    assert n1.containers[0]["image"] == "rancher/agent:v1.2.0" and n1.containers[0]["privileged"] in (True, "yes")
    assert n1.containers[0]["command"] == "http://10.0.0.1:8080/v1/scripts/TOKEN1C1:1a7"
    assert n2.containers[0]["command"] == "http://10.0.0.1:8080/v1/scripts/TOKEN1C2:1a7"
    for h in (master, n1, n2):
        assert h.docker_version == "1.12.6" and "docker-py" in h.pip and h.dist_upgraded
        assert h.apt_repos == ["deb https://apt.dockerproject.org/repo ubuntu-xenial main"]
        assert {"linux-headers-generic", "build-essential", "zfs", "python-pip", "apt-transport-https", "ca-certificates",
                "curl", "software-properties-common", "docker-engine=1.12.6-0~ubuntu-xenial"} <= set(h.packages)
    assert run.slept_s == 10 and run.modules.paused_s == 15       # one retry of `docker logs`, then the 15 s pause


@pytest.mark.parametrize("label,src", _sources(), ids=[s[0] for s in _sources()])
def test_second_run_is_idempotent(tmp_path, label, src):
    """docker present, master and agents running: every guarded task is skipped,
    no machine changes -- only the (unguarded) environment creation repeats, as in
    the reference (ranchermaster/tasks/main.yml:29-52 carry no `when`)."""
    d = _tree(tmp_path, src)
    with ranchermock.RancherMock() as mock:
        first = _run(d, mock)
        before = {h: (list(s.packages), [c["name"] for c in s.containers]) for h, s in first.hosts.items()}
        n_req = len(mock.requests)
        second = _run(d, mock, hosts=first.hosts)
        assert second.failed_hosts == {}
        after = {h: (list(s.packages), [c["name"] for c in s.containers]) for h, s in second.hosts.items()}
        assert after == before
        assert mock.wire_log()[n_req:] == ["GET /v2-beta/projectTemplates?name=kubernetes", "POST /v2-beta/projects"]
    skipped = [(h, t) for _, h, t, st in second.events if st == "skipped"]
    assert len([1 for h, _ in skipped if h == "10.0.1.1"]) >= 4 + 6


@pytest.mark.parametrize("label,src", _sources(), ids=[s[0] for s in _sources()])
def test_failed_join_is_retried_by_a_second_run(tmp_path, label, src):
    """The agent's bootstrap container exits without starting `rancher-agent`
    (bad token, server not ready).  The guard on every node task is the substring
    'rancher-agent' in `docker ps -a` (rancherhost/tasks/main.yml:9-34), so a
    second run must still find it open and ask for tokens again -- which it only
    does if the bootstrap container's own name does not contain that substring."""
    d = _tree(tmp_path, src)
    with ranchermock.RancherMock() as mock:
        first = _run(d, mock, agent_bootstrap_fails=True)
        assert first.failed_hosts == {}
        for h in ("10.0.1.1", "10.0.1.2"):
            assert "rancher-agent" not in first.hosts[h].docker_ps()
        n_req = len(mock.requests)
        second = _run(d, mock, hosts=first.hosts)
        assert second.failed_hosts == {}
        retried = [r for r in mock.wire_log()[n_req:] if r.startswith("POST /v1/registrationtokens")]
        assert len(retried) == 2
        for h in ("10.0.1.1", "10.0.1.2"):
            assert "rancher-agent" in second.hosts[h].docker_ps()


@pytest.mark.parametrize("label,src", _sources(), ids=[s[0] for s in _sources()])
def test_project_creation_rejected(tmp_path, label, src):
    """Rancher answers 500 to POST /v2-beta/projects: the task fails on status_code 201,
    run_once takes every host of the play down, nodes never ask for tokens."""
    d = _tree(tmp_path, src)
    with ranchermock.RancherMock(fail={("POST", "/v2-beta/projects"): 500}) as mock:
        run = _run(d, mock)
        assert "10.0.0.1" in run.failed_hosts and "Status code was 500" in run.failed_hosts["10.0.0.1"]
        assert not any(r["path"].startswith("/v1/registrationtokens") and r["method"] == "GET" for r in mock.requests)
    assert not (d / "tmp/kubernetes_environment.id").exists()


def test_rancher_never_comes_up(tmp_path):
    d = _tree(tmp_path, os.path.join(WORKSPACE_TEMPLATE, "ansible"))
    with ranchermock.RancherMock() as mock:
        run = _run(d, mock, rancher_up_after=10 ** 6)
        assert "until condition never became true after 30 attempts" in run.failed_hosts["10.0.0.1"]
        assert run.slept_s == 29 * 10
        assert mock.requests == []


def test_mock_validates_what_it_is_sent():
    import urllib.error
    import urllib.request
    with ranchermock.RancherMock() as mock:
        def post(path, doc):
            req = urllib.request.Request("http://%s%s" % (mock.address, path), data=json.dumps(doc).encode(), method="POST",
                                         headers={"Content-Type": "application/json"})
            try:
                return urllib.request.urlopen(req).status
            except urllib.error.HTTPError as exc:
                return exc.code
        assert post("/v2-beta/projects", {"name": "x"}) == 422                              # members missing
        assert post("/v1/registrationtokens?projectId=nope", {}) == 404
        assert post("/v2-beta/nothing", {}) == 404


def test_templar_matches_ansible_conventions():
    t = roledryrun.Templar()
    v = {"r": {"json": {"data": [{"id": "1pt5"}]}, "stdout": "abc"}, "project_id": {"content": "MWE3Cg=="}}
    assert t.render("{{ r.json.data[0].id }}", v) == "1pt5"
    assert t.render("{{ r.json.data }}", v) == [{"id": "1pt5"}]                           # native type kept
    assert t.render("x{{ project_id['content'] | b64decode | replace('\\n', '') }}y", v) == "x1a7y"
    assert t.truth("{{ 'abc' not in r.stdout }}", v) is False                              # 2.2-style braces
    assert t.truth("r.stdout.find('zzz')", v) is True                                      # -1 is truthy (dockersetup quirk)
    assert t.truth("r.stdout.find('abc')", v) is False                                     # 0 is falsy
    with pytest.raises(Exception):
        t.render("{{ undefined_thing }}", v)
