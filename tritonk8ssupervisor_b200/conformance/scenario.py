"""Scenario = everything outside the program that one run of the provisioning
path can observe: the answers on stdin, the tables the ``triton`` CLI prints,
the operator's ``~/.ssh``, files left behind by an earlier run, and the scripted
behaviour of the tools below the offline line (terraform, ansible, curl, ssh).

A scenario is pure data (JSON-serialisable) so that a golden fixture can carry
the exact scenario it was recorded for (tests/golden/*.json).

The tables in :data:`DOC_NETWORKS` / :data:`DOC_PACKAGES` are the ones printed in
the reference's own walkthrough (docs/detailed.md:156-159 and :178-195); they are
the only "fixture" the reference publishes for this path.
"""
from __future__ import annotations

import dataclasses
import json
from typing import Dict, List, Optional

# docs/detailed.md:156-159 (menu rows are "name  uuid")
DOC_NETWORKS = [
    ("Joyent-SDC-Private", "909c0c0d-1455-404f-85bd-04f48b7b0059"),
    ("Joyent-SDC-Public", "31428241-4878-47d6-9fba-9a8436b596a4"),
    ("My-Fabric-Network", "0882d255-ac1e-41b2-b1d5-e08200ebb380"),
    ("kubernetes", "0b206464-d655-4723-a848-86d0f28764c8"),
]

# docs/detailed.md:178-195
DOC_PACKAGES = [
    ("k4-bigdisk-kvm-15.75G", "7741b8f6-2733-11e6-bdb9-bf11c4147d38"),
    ("k4-bigdisk-kvm-31.75G", "14c01a1a-d0f8-11e5-ad69-1fd27456ad73"),
    ("k4-bigdisk-kvm-63.75G", "14c0992c-d0f8-11e5-bd78-e71dad0f8626"),
    ("k4-fastdisk-kvm-31.75G", "14bd9600-d0f8-11e5-a69c-97be6e961834"),
    ("k4-fastdisk-kvm-63.75G", "14be13c8-d0f8-11e5-b55b-47eb44d4e064"),
    ("k4-general-kvm-15.75G", "14ac8f5e-d0f8-11e5-a0e5-9b622a20595f"),
    ("k4-general-kvm-3.75G", "14aba044-d0f8-11e5-8c88-eb339a5da5d0"),
    ("k4-general-kvm-31.75G", "14ad1a32-d0f8-11e5-a465-8f264489308b"),
    ("k4-general-kvm-7.75G", "14ac17a4-d0f8-11e5-a400-e39503e18b19"),
    ("k4-highcpu-kvm-1.75G", "14b5edc4-d0f8-11e5-b4d2-b3e6e8c05f9d"),
    ("k4-highcpu-kvm-15.75G", "14b783d2-d0f8-11e5-8d93-6ba10192d750"),
    ("k4-highcpu-kvm-250M", "14b4ff36-d0f8-11e5-a8b1-e343c129d7f0"),
    ("k4-highcpu-kvm-3.75G", "14b67ef6-d0f8-11e5-ba19-479de37c6f75"),
    ("k4-highcpu-kvm-7.75G", "14b6fade-d0f8-11e5-85c5-4ff7918ab5c1"),
    ("k4-highcpu-kvm-750M", "14b5760a-d0f8-11e5-9cb1-23c9c232c00e"),
    ("k4-highram-kvm-15.75G", "14ba876c-d0f8-11e5-8a1b-ab02fdd17b07"),
    ("k4-highram-kvm-31.75G", "14bafb20-d0f8-11e5-a5cf-e386b841ed87"),
    ("k4-highram-kvm-63.75G", "14bb84f0-d0f8-11e5-8014-2fb7b19ccb24"),
]

# rows the filters on the path must drop: non-KVM packages (setup.sh:259 keeps
# only names containing "-kvm-")
NON_KVM_PACKAGES = [
    ("g4-highcpu-1G", "2bbe1e9e-4ef8-c26a-e3c8-e3a8d2d0f1a1"),
    ("g4-general-4G", "bb5ef2d3-8a66-6c8c-d1a8-9e7a3a8a7e44"),
]

DEFAULT_KEY_ID = "aa:bb:cc:dd"


def triton_table(rows, header=("NAME", "ID"), width: Optional[int] = None) -> str:
    """Render rows the way ``triton <thing> -oname,id`` does: a header line and
    space-padded columns (the path depends on that shape: setup.sh:257,536)."""
    names = [header[0]] + [r[0] for r in rows]
    w = width if width is not None else max(len(n) for n in names)
    out = ["%s  %s" % (header[0].ljust(w), header[1])]
    for name, ident in rows:
        out.append("%s  %s" % (name.ljust(w), ident))
    return "\n".join(out) + "\n"


def triton_env_text(url="https://us-east-1.api.joyent.com", account="fixture",
                    key_id=DEFAULT_KEY_ID, profile="env") -> str:
    """What ``triton env`` prints for a profile (the path ``eval``s it: setup.sh:210)."""
    return (
        'export TRITON_PROFILE="%s"\n'
        'export SDC_URL="%s"\n'
        'export SDC_ACCOUNT="%s"\n'
        'unset SDC_USER\n'
        'export SDC_KEY_ID="%s"\n'
        'unset SDC_TESTING\n'
        '# Run this command to configure your shell:\n'
        '#     eval "$(triton env)"\n' % (profile, url, account, key_id)
    )


@dataclasses.dataclass
class Scenario:
    """All observable inputs of one run."""

    name: str
    #: bytes fed on stdin (answers, one per line; setup.sh:265-450,472)
    stdin: str = ""
    #: command-line arguments after the program name (``-c`` = clean, setup.sh:9)
    args: List[str] = dataclasses.field(default_factory=list)
    triton_env: str = dataclasses.field(default_factory=triton_env_text)
    networks: str = dataclasses.field(default_factory=lambda: triton_table(DOC_NETWORKS))
    packages: str = dataclasses.field(
        default_factory=lambda: triton_table(DOC_PACKAGES + NON_KVM_PACKAGES))
    #: files in ``$HOME/.ssh`` (name -> content)
    ssh_files: Dict[str, str] = dataclasses.field(
        default_factory=lambda: {"id_rsa": "PRIVATE KEY FIXTURE\n", "id_rsa.pub": "ssh-rsa AAAA fixture\n"})
    #: what ``ssh-keygen -E md5 -lf ~/.ssh/<name>`` prints (name -> line)
    keygen: Dict[str, str] = dataclasses.field(
        default_factory=lambda: {"id_rsa": "2048 MD5:%s fixture@host (RSA)\n" % DEFAULT_KEY_ID})
    #: what legacy ``ssh-keygen -l -f`` prints (name -> line); falls back to ``keygen_sha``
    keygen_legacy: Dict[str, str] = dataclasses.field(default_factory=dict)
    keygen_sha: Dict[str, str] = dataclasses.field(
        default_factory=lambda: {"id_rsa": "2048 SHA256:abcdefghijklmnopqrstuvwxyz0123456789ABCDEFG fixture@host (RSA)\n"})
    #: ssh-keygen too old to know ``-E`` (setup.sh:222 comment)
    keygen_no_E: bool = False
    #: files present in the workspace before the run (relative path -> content)
    pre_files: Dict[str, str] = dataclasses.field(default_factory=dict)
    #: files removed from the workspace before the run
    pre_remove: List[str] = dataclasses.field(default_factory=list)
    #: scripted exit statuses, "<tool>-<first arg>" -> status
    fail: Dict[str, int] = dataclasses.field(default_factory=dict)
    #: terraform apply writes no IP files (provisioning limit hit; setup.sh:117-120)
    terraform_no_ips: bool = False
    master_ip: Optional[str] = None
    #: what the stubbed playbook run leaves in ansible/tmp/kubernetes_environment.id
    env_id: Optional[str] = None
    #: successive ``curl`` bodies (index from 1), then ``curl_default``
    curl_seq: List[str] = dataclasses.field(default_factory=list)
    curl_default: Optional[str] = None
    #: successive ``date +%s`` outputs, then 1500000000 (divisible by 15: setup.sh:64)
    date_seq: List[str] = dataclasses.field(default_factory=list)
    #: ``docker ps`` output per node ip for the stubbed ssh
    ssh_ps: Dict[str, str] = dataclasses.field(default_factory=dict)
    ssh_exit: Dict[str, int] = dataclasses.field(default_factory=dict)
    #: ``$HOME/.ssh/known_hosts`` present (setup.sh:504)
    known_hosts: Optional[str] = None
    #: tools that are not installed at all (no stub on PATH): bash says "command not found"
    missing_tools: List[str] = dataclasses.field(default_factory=list)
    #: how the program is started: "./setup.sh" (README.md:105, the documented way) or
    #: "bash setup.sh"; bash labels its own diagnostics with the name it was given
    invocation: str = "./setup.sh"
    #: extra environment of the run (both implementations), applied over the sandbox's own;
    #: the token <SBX> in a value stands for the sandbox directory
    env: Dict[str, str] = dataclasses.field(default_factory=dict)
    #: variables removed from the environment of the run (e.g. HOME, LC_ALL)
    unset_env: List[str] = dataclasses.field(default_factory=list)
    #: name of the home directory inside the sandbox, and directories created next to it
    #: with the same ~/.ssh content (a HOME of "h*me" beside "home": tilde results are never globbed)
    home_name: str = "home"
    home_decoys: List[str] = dataclasses.field(default_factory=list)
    #: symbolic links present in the workspace before the run (relative path -> target)
    pre_symlinks: Dict[str, str] = dataclasses.field(default_factory=dict)
    #: tools that read the stdin they inherit: "<tool>-<first arg>" -> number of lines taken
    stdin_eaters: Dict[str, int] = dataclasses.field(default_factory=dict)
    note: str = ""

    @property
    def program_label(self) -> str:
        """What bash puts in front of "line N:" for this invocation."""
        return "setup.sh" if self.invocation == "bash setup.sh" else self.invocation

    def reference_argv(self) -> List[str]:
        return self.invocation.split(" ")

    def to_json(self) -> dict:
        return dataclasses.asdict(self)

    @classmethod
    def from_json(cls, d: dict) -> "Scenario":
        known = {f.name for f in dataclasses.fields(cls)}
        return cls(**{k: v for k, v in d.items() if k in known})

    def dumps(self) -> str:
        return json.dumps(self.to_json(), indent=1, sort_keys=True)
