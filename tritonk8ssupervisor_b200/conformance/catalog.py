"""The scenario catalogue: BASELINE.json's five configs plus the edge cases the
path has (SURVEY.md §5 validation table, §7.4 hard parts, §8d negative sub-cases).

``CONFIGS`` are the five scored configurations.  ``EDGE`` are
parity cases: every branch of the prompt validation, the guards, the failure
exits, the clean path and the dashboard wait loop.  The reference has no tests
(SURVEY.md §4), so each edge case cites the script lines it exercises instead of
a reference test.
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List

from .scenario import DEFAULT_KEY_ID, DOC_NETWORKS, DOC_PACKAGES, NON_KVM_PACKAGES, Scenario, triton_env_text, triton_table

YES = "yes\n"
BLANK8 = "\n" * 8


def _answers(*lines: str) -> str:
    return "".join(l + "\n" for l in lines)


# ---------------------------------------------------------------------------
# the five BASELINE.json configs
# ---------------------------------------------------------------------------
CONFIGS: List[Scenario] = [
    Scenario(name="config1_defaults_1worker", stdin=BLANK8 + YES,
             note="BASELINE config 1: every prompt left at its default"),
    Scenario(name="config2_quickstart_2workers", stdin=_answers("", "", "", "", "2", "", "", "", "yes"),
             note="BASELINE config 2: README.md:112-121 answers"),
    Scenario(name="config3_custom_names_multinet",
             stdin=_answers("prod k8s", 'my "quoted" desc', "km01", "wrk", "3", "4,2,2", "1,3", "1", "yes"),
             note="BASELINE config 3: custom hostname + prefix + several networks (sorted, de-duplicated)"),
    Scenario(name="config4_5workers_inventory", stdin=_answers("", "", "", "", "5", "", "", "", "yes"),
             note="BASELINE config 4: 5 workers, inventory render"),
    Scenario(name="config5_10workers_rejected_then_9", stdin=_answers("", "", "", "", "10", "9", "", "", "", "yes"),
             note="BASELINE config 5: 10 is outside the reference's domain (setup.sh:301); 9 is the maximum"),
]

# ---------------------------------------------------------------------------
# edge cases
# ---------------------------------------------------------------------------
_TWELVE_NETWORKS = [("net%02d" % i, "00000000-0000-0000-0000-%012d" % i) for i in range(1, 13)]
_COLLIDING_NETWORKS = [("kube", "aaaaaaaa-0000-0000-0000-000000000001"),
                       ("kube-net", "bbbbbbbb-0000-0000-0000-000000000002"),
                       ("Joyent-SDC-Public", "31428241-4878-47d6-9fba-9a8436b596a4")]
_COLLIDING_PACKAGES = [("k4-kvm-1G", "cccccccc-0000-0000-0000-000000000001"),
                       ("k4-kvm-1G-kvm-x", "dddddddd-0000-0000-0000-000000000002"),
                       ("k4-highcpu-kvm-7.75G", "14b6fade-d0f8-11e5-85c5-4ff7918ab5c1")]

_RAN_FILES: Dict[str, str] = {
    "config": "ANSIBLE_HOST_KEY_CHECKING=False\nSDC_URL=\"x\"\n",
    "terraform/rancher.tf": 'provider "triton" {\n}\n',
    "terraform/masters.ip": "10.0.0.1\n",
    "terraform/hosts.ip": "10.0.1.1\n10.0.1.2\n",
    "terraform/terraform.tfstate": "{}\n",
    "terraform/terraform.tfstate.backup": "{}\n",
    "terraform/.terraform/modules/abc/main.tf": "# fetched module\n",
    "ansible/hosts": "[MASTER]\n10.0.0.1\n[HOST]\n10.0.1.1\n10.0.1.2\n",
    "ansible/clusterUp.retry": "10.0.1.2\n",
    "ansible/ansible.cfg.tmp": "stale\n",
    "tmp/scratch.txt": "scratch\n",
    "tmp/sub/deep.txt": "deep\n",
    "tmp/.hidden": "kept: tmp/* does not match dot files\n",
}

_PS_TWO_DASHBOARDS = (
    "CONTAINER ID        IMAGE                                  COMMAND        NAMES\n"
    "1111aaaa2222        gcr.io/kubernetes-dashboard-amd64:v1   \"/dashboard\"   k8s_kubernetes-dashboard.abc\n"
    "3333bbbb4444        gcr.io/pause-amd64:3.0                 \"/pause\"       k8s_POD.d8dbe16c_kubernetes-dashboard-716739405\n"
    "5555cccc6666        rancher/agent:v1.2.0                   \"/run.sh\"      rancher-agent\n")
_PS_NONE = (
    "CONTAINER ID        IMAGE                                  COMMAND        NAMES\n"
    "5555cccc6666        rancher/agent:v1.2.0                   \"/run.sh\"      rancher-agent\n")

EDGE: List[Scenario] = [
    # --- prompt validation (setup.sh:273-307) --------------------------------
    Scenario(name="bad_hostnames", stdin=_answers("", "", "1bad", "a", "has-dash", "km", "x_y", "9n", "n0", "", "", "", "", "yes"),
             note="hostname/prefix regex ^[a-zA-Z][0-9a-zA-Z]+$ incl. the 1-char rejection"),
    Scenario(name="bad_node_counts", stdin=_answers("", "", "", "", "0", "10", "x", "1 2", "", "", "", "", "yes"),
             note="node count regex ^[1-9]$"),
    # --- network selection (setup.sh:309-400) --------------------------------
    Scenario(name="net_out_of_range", stdin=_answers("", "", "", "", "", "5", "0", "1,9", "007", "2,", ",2", "x", "3", "4,4,4", "", "yes"),
             note="index range checks and the index-list regex"),
    Scenario(name="net_lexical_sort", networks=triton_table(_TWELVE_NETWORKS),
             stdin=_answers("", "", "", "", "", "10,2", "9,10,11", "", "yes"),
             note="sort without -n: 10,2 stays 10,2 (SURVEY §7.4)"),
    Scenario(name="net_with_blanks", stdin=_answers("", "", "", "", "", "2, 4", "4 ,2", " 3 ", "", "yes"),
             note="blanks inside the index list: what echo's field splitting makes of them"),
    Scenario(name="net_prefix_collision", networks=triton_table(_COLLIDING_NETWORKS),
             stdin=_answers("", "", "", "", "", "2", "3", "", "yes"),
             note="menu order != resolver order when one name is a '-' extension of another (SURVEY §7.4)"),
    Scenario(name="net_uuid_typed_in", stdin=_answers("", "", "", "", "", "31428241-4878-47d6-9fba-9a8436b596a4", "deadbeef", "", "", "yes"),
             note="typing the default UUID is accepted by the [[ == ]] branch; any other string is not"),
    Scenario(name="net_header_lookalike",
             networks=triton_table([("NAMEX", "11111111-ID"), ("Joyent-SDC-Public", "31428241-4878-47d6-9fba-9a8436b596a4"),
                                    ("Alonely", ""), ("zeta", "22222222-zeta")]).replace("Alonely            \n", "Alonely\n"),
             stdin=_answers("", "", "", "", "", "1", "3", "", "yes"),
             note="a row that looks like the header is dropped from the menu (^NAME.*ID$) but not by the resolver "
                  "(^NAME *ID$): menu entry 3 (zeta) resolves to NAMEX's id; a row without an id (entry 1) resolves to "
                  "nothing (awk NF>1)"),
    Scenario(name="net_three_digit_indices", stdin=_answers("", "", "", "", "", "100,200", "7,8,9", "", "", "", "yes"),
             note="the regex admits at most two digits per index: '100,200' costs ONE error pair (regex), '7,8,9' three (range)"),
    Scenario(name="net_default_is_a_pattern", networks=triton_table([("Joyent-SDC-Public", "z?z-9")]),
             packages=triton_table([("k4-highcpu-kvm-7.75G", "q*q")]),
             stdin=_answers("", "", "", "", "", "zqz-9", "nomatch", "", "qxyzq", "yes"),
             note="[[ $answer == $default ]] has an unquoted right-hand side: the default is a *pattern* (setup.sh:345,388,442)"),
    Scenario(name="net_default_with_extended_pattern", networks=triton_table([("Joyent-SDC-Public", "@(abc|x?z)+(-)!(q)")]),
             stdin=_answers("", "", "", "", "", "xyz-r", "abd-", "abc--", "", "yes"),
             note="inside [[ ]] the pattern operators ?() *() +() @() !() are on: a default of '@(abc|x?z)+(-)!(q)' accepts "
                  "'xyz-r' and 'abc--', not 'abd-'"),
    Scenario(name="pkg_default_is_a_broken_pattern",
             packages=triton_table([("k4-highcpu-kvm-7.75G", "[a-"), ("k4-highcpu-kvm-1.75G", "14b5edc4-d0f8-11e5-b4d2-b3e6e8c05f9d")]),
             stdin=_answers("", "", "", "", "", "", "", "", "[a-", "1", "yes"),
             note="a '[' that finds no ']' is an ordinary character -- unless the search runs into a dangling range: '[a-' "
                  "matches nothing, not even itself, so the empty answer (= the default) is rejected until an index is typed"),
    Scenario(name="pkg_index_word_is_sed_options_ne", stdin=_answers("", "", "", "", "", "", "", "-e \\\\055ne 1", "", "yes"),
             note="the answer leaves '-ne 1' in the variable: echo takes -ne for its own options (the check sees '1'), and "
                  "`sed -n \"-ne\"p` is `sed -n -n -e p`: every row is printed and HOST_PACKAGE is all 18 ids (setup.sh:427,541)"),
    Scenario(name="pkg_index_word_is_sed_options_n", stdin=_answers("", "", "", "", "", "", "", "-e \\\\055n 1", "", "yes"),
             note="... '-n': `sed -n -np` -- invalid option 'p', sed's usage on stderr, an empty package id"),
    Scenario(name="pkg_index_word_is_sed_options_en", stdin=_answers("", "", "", "", "", "", "", "-e \\\\055en 1", "", "yes"),
             note="... '-en': the script is 'np' -- extra characters after command"),
    Scenario(name="pkg_index_word_is_sed_options_eE", stdin=_answers("", "", "", "", "", "", "", "-e \\\\055eE 1", "", "yes"),
             note="... '-eE': the script is 'Ep' -- unknown command: `E'"),
    Scenario(name="cfg_line_is_export_p", stdin=_answers("", "-e a\\\\n-p\\\\nQ=", "", "", "", "", "", "", "", "yes"),
             note="a description of three lines puts a line '-p' into ./config: `export \"-p\"` lists every exported variable on "
                  "stdout (declare -x ..., the ones without a value by name only) and the run goes on (setup.sh:546)"),
    Scenario(name="cfg_line_is_an_invalid_export_option", stdin=_answers("", "-e a\\\\n-n X\\\\nQ=", "", "", "", "", "", "", "", "yes"),
             note="... '-n X' is one word: the options n, ' ', X -- `export: - : invalid option`, usage, status 2, errexit"),
    Scenario(name="cfg_line_exports_a_name_without_value", stdin=_answers("", "-e a\\\\ni\\\\nQ=", "", "", "", "", "", "", "", "yes"),
             note="... 'i': the loop counter of setup.sh:148 has the export attribute from then on, terraform sees i=2"),
    Scenario(name="cfg_line_assigns_a_readonly_variable", stdin=_answers("", "-e a\\\\nUID=0\\\\nQ=", "", "", "", "", "", "", "", "yes"),
             note="... 'UID=0': `./setup.sh: line 546: UID: readonly variable`, status 1, errexit; config.tmp stays behind"),
    Scenario(name="net_header_line_holds_a_nul", stdin=BLANK8 + YES, networks="NAME \x00 ID\n" + triton_table(DOC_NETWORKS).split("\n", 1)[1],
             note="binary input: while it looks for matches grep lets a NUL end a line, so `^NAME.*ID$` does not match the "
                  "header `NAME \\0 ID`, `grep -v` selects it and says `binary file matches` (setup.sh:257)"),
    Scenario(name="net_no_public_network", networks=triton_table(DOC_NETWORKS[:1] + DOC_NETWORKS[2:]),
             stdin=BLANK8 + YES, note="no Joyent-SDC-Public row: default position 1 (setup.sh:311)"),
    Scenario(name="net_single_row_table", networks=triton_table(DOC_NETWORKS[1:2]), stdin=_answers("", "", "", "", "", "2", "1", "", "", "yes"),
             note="one network: index 2 is out of range"),
    Scenario(name="net_empty_table", networks="NAME  ID\n", stdin=BLANK8 + YES,
             note="no networks at all: empty default, everything still flows"),
    # --- package selection (setup.sh:401-450) --------------------------------
    Scenario(name="pkg_out_of_range", stdin=_answers("", "", "", "", "", "", "", "19", "0", "1,2", "abc", "10", "yes"),
             note="package index checks; prints 'entered 10 and got <uuid>'"),
    Scenario(name="pkg_default_missing", packages=triton_table(DOC_PACKAGES[:13] + NON_KVM_PACKAGES),
             stdin=BLANK8 + YES, note="no k4-highcpu-kvm-7.75G row: default position 1 (setup.sh:404)"),
    Scenario(name="pkg_prefix_collision", packages=triton_table(_COLLIDING_PACKAGES),
             stdin=_answers("", "", "", "", "", "", "", "1", "yes"),
             note="menu sorted after blank->'=', resolver sorted with blanks kept (setup.sh:259 vs :541)"),
    Scenario(name="pkg_huge_index", stdin=_answers("", "", "", "", "", "", "", "99999999999999999999", "18446744073709551617", "yes"),
             note="64-bit wrap-around of bash arithmetic in the range check"),
    # --- free text and quoting ---------------------------------
    Scenario(name="text_quotes_and_blanks", stdin=_answers('  spaced   out  ', 'say "hi"  twice', "", "", "", "", "", "", "yes"),
             note="field splitting collapses blanks; quotes survive in config, not in vars.yml"),
    Scenario(name="text_backslashes", stdin=_answers("back\\\\slash", "tab\\there c:\\\\dir\\\\", "", "", "", "", "", "", "yes"),
             note="read without -r, twice (prompt and exportVars)"),
    Scenario(name="text_glob_chars", stdin=_answers("*", "terraform/* ? [a-z]*", "", "", "", "", "", "", "yes"),
             note="unquoted echo pathname-expands the answer against the workspace"),
    Scenario(name="text_echo_options", stdin=_answers("-n", "-e a\\\\tb", "", "", "", "", "", "", "yes"),
             note="a leading -n/-e in the answer is taken by echo as an option"),
    Scenario(name="text_shell_metachars", stdin=_answers("$HOME `id` $(id)", "a;b|c&d>e<f 'q' ~ #x", "", "", "", "", "", "", "yes"),
             note="nothing is evaluated: the answers are data"),
    Scenario(name="text_utf8", stdin=_answers("café ☕", "naïve — test", "", "", "", "", "", "", "yes"),
             note="non-ASCII bytes pass through untouched under LC_ALL=C"),
    Scenario(name="text_embedded_newline", stdin=_answers("-e a\\\\nb", "", "", "", "", "", "", "", "yes"),
             note="echo -e turns the answer into two lines; exportVars then dies on `b\"` (exit 1, config.tmp left)"),
    Scenario(name="text_nul_and_c_escape", stdin=_answers("-e a\\\\0b", "-e cut\\\\chere", "", "", "", "", "", "", "yes"),
             note="\\0 makes a NUL that $(...) drops with a warning on stderr; \\c cuts the output"),
    Scenario(name="text_line_continuation", stdin="first\\\nsecond\n" + "\n" * 7 + YES,
             note="backslash-newline joins two input lines (read without -r)"),
    Scenario(name="text_crlf_answers", stdin="name\r\n\r\n\r\n\r\n2\r\n\r\n\r\n\r\nyes\r\nyes\n",
             note="CR is data: 'km\\r' style answers fail the regexes, '\\r' is not the empty answer"),
    Scenario(name="idx_echo_options", stdin=_answers("", "", "", "", "", "-n 4", "-e 1\\\\n2", "-n 4", "", "yes"),
             note="-n/-e in an index answer: what grep sees differs from what the resolver gets"),
    # --- confirmation (setup.sh:471-482) --------------------------------------
    Scenario(name="verify_no", stdin=BLANK8 + "no\n", note="'no' exits 0 and leaves ./config behind"),
    Scenario(name="verify_garbage_then_yes", stdin=BLANK8 + _answers("maybe", "YES", "y", "", " yes ", "yes"),
             note="only exact yes/no; ' yes ' is trimmed by read and accepted"),
    Scenario(name="verify_eof", stdin=BLANK8, note="EOF at the confirmation: read fails, errexit, status 1"),
    Scenario(name="stdin_empty", stdin="", note="every prompt at EOF takes its default; dies at the confirmation"),
    Scenario(name="stdin_no_final_newline", stdin=BLANK8 + "yes", note="'yes' without newline: read returns 1 -> exit 1"),
    # --- guards (setup.sh:14-18, 241-244) -------------------------------------
    Scenario(name="guard_previous_run", stdin=BLANK8 + YES, pre_files={"terraform/rancher.tf": "# left over\n"},
             note="previous-run guard: message and exit 0"),
    Scenario(name="guard_stale_config_clean_yes", stdin="yes\n" + BLANK8 + YES, pre_files={"config": "OLD=1\n"},
             note="stale ./config: cleanRunner consumes one extra answer, then the run proceeds"),
    Scenario(name="guard_stale_config_clean_no", stdin="no\n" + BLANK8 + YES, pre_files={"config": "OLD=1\n"},
             note="stale ./config, operator declines: exit 0"),
    # --- key discovery (setup.sh:215-238) --------------------------------------
    Scenario(name="key_not_found_then_clean", stdin="yes\n", keygen={"id_rsa": "2048 MD5:ff:ff:ff:ff other (RSA)\n"},
             note="fingerprint mismatch: message, cleanRunner, exit 1"),
    Scenario(name="key_not_found_decline_clean", stdin="no\n", keygen={"id_rsa": "2048 MD5:ff:ff:ff:ff other (RSA)\n"},
             note="fingerprint mismatch, decline clean: exit 0 with ./config left"),
    Scenario(name="key_legacy_ssh_keygen", stdin=BLANK8 + YES, keygen_no_E=True, keygen={},
             keygen_legacy={"id_rsa": "2048 %s fixture@host (RSA)\n" % DEFAULT_KEY_ID}, keygen_sha={},
             note="ssh-keygen without -E prints MD5 by default (setup.sh:222-229)"),
    Scenario(name="key_among_other_files", stdin=BLANK8 + YES,
             ssh_files={"authorized_keys": "ssh-rsa AAAA\n", "config": "Host *\n", "github_rsa": "K\n", "github_rsa.pub": "P\n",
                        "id_rsa": "K\n", "id_rsa.pub": "P\n", "known_hosts": "h\n"},
             keygen={"github_rsa": "4096 MD5:11:22:33:44 gh (RSA)\n", "id_rsa": "2048 MD5:%s fixture@host (RSA)\n" % DEFAULT_KEY_ID},
             keygen_sha={"github_rsa": "4096 SHA256:zzzz gh (RSA)\n", "id_rsa": "2048 SHA256:yyyy x (RSA)\n"},
             note="~/.ssh walked in ls order; non-key files cost two failed ssh-keygen calls each"),
    Scenario(name="key_pub_suffix_sed", stdin=BLANK8 + YES,
             ssh_files={"mykey.pub": "P\n", "mykey": "K\n"},
             keygen={"mykey": "2048 MD5:%s fixture@host (RSA)\n" % DEFAULT_KEY_ID}, keygen_sha={"mykey": "2048 SHA256:q x (RSA)\n"},
             note="the .pub suffix is stripped before ssh-keygen is asked (setup.sh:217)"),
    Scenario(name="key_dot_matches_any_char", stdin=BLANK8 + YES,
             ssh_files={"akeyxpub": "not a key\n", "id_rsa": "K\n", "id_rsa.pub": "P\n"},
             note="sed 's/.pub$//': the dot is a regex dot, so 'akeyxpub' is asked about as 'akey' (setup.sh:217)"),
    Scenario(name="key_name_with_ampersand", stdin=BLANK8 + YES, ssh_files={"a&b": "K\n"},
             keygen={"a&b": "2048 MD5:%s c (RSA)\n" % DEFAULT_KEY_ID}, keygen_sha={},
             note="the key path is spliced into a sed program (setup.sh:133): '&' there means 'the matched text'"),
    Scenario(name="key_name_with_semicolon", stdin=BLANK8 + YES, ssh_files={"k;": "K\n"},
             keygen={"k;": "2048 MD5:%s c (RSA)\n" % DEFAULT_KEY_ID}, keygen_sha={},
             note="';' ends the sed replacement; the script's own trailing ';g' becomes the command g and blanks every "
                  "line of ansible.cfg"),
    Scenario(name="key_name_breaks_the_sed_program", stdin=BLANK8 + YES, ssh_files={"a;b": "K\n"},
             keygen={"a;b": "2048 MD5:%s c (RSA)\n" % DEFAULT_KEY_ID}, keygen_sha={},
             note="'a;b': sed rejects the program (unknown option to s), ansible.cfg keeps its old line, tmp.cfg stays behind"),
    Scenario(name="key_no_ssh_dir_entries", stdin="yes\n", ssh_files={}, note="empty ~/.ssh: key not found"),
    Scenario(name="triton_env_fails", stdin="yes\n", fail={"triton-env": 1},
             note="triton env failing is swallowed by eval \"$(...)\"; empty SDC_* then matches the first non-key"),
    Scenario(name="stale_ip_files_are_appended_to", stdin=_answers("", "", "", "", "2", "", "", "", "yes"),
             pre_files={"terraform/masters.ip": "192.0.2.1\n", "terraform/hosts.ip": "192.0.2.2\n"},
             curl_seq=["not yet\n", "kubernetes\n"],
             note="IP files of an earlier apply are appended to, not replaced: vars.yml takes the LAST master (tail -1), "
                  "the banner and the curl URL get both lines, ssh visits the stale node too"),
    # --- below the offline line (setup.sh:117-120,154-159,113) ----------------
    Scenario(name="terraform_get_fails", stdin=BLANK8 + YES, fail={"terraform-get": 3}, note="errexit with terraform's status"),
    Scenario(name="terraform_apply_fails", stdin=BLANK8 + YES, fail={"terraform-apply": 1}, note="errexit with terraform's status"),
    Scenario(name="terraform_writes_no_ips", stdin=BLANK8 + YES, terraform_no_ips=True,
             note="provisioning limit: 'Terraform had too many errors', exit 1 (setup.sh:117-120)"),
    Scenario(name="ansible_fails", stdin=BLANK8 + YES, fail={"ansible-playbook--i": 2}, note="errexit with ansible-playbook's status"),
    Scenario(name="ansible_cfg_missing", stdin=BLANK8 + YES, pre_remove=["ansible/ansible.cfg"],
             note="sed cannot read ansible.cfg: '&&' list fails quietly, tmp.cfg is left behind"),
    Scenario(name="env_id_from_playbook", stdin=_answers("", "", "", "", "2", "", "", "", "yes"), env_id="1a9", master_ip="165.225.1.2",
             note="banner URLs come from masters.ip and the id the playbook stored"),
    Scenario(name="env_id_missing", stdin=BLANK8 + YES, pre_remove=["ansible/tmp/kubernetes_environment.id"],
             note="no stored id: cat fails inside $(...), URLs have an empty id"),
    # --- a workspace that is not what setup expects ------------------------------
    Scenario(name="ws_no_terraform_dir", stdin=BLANK8 + YES, pre_remove=["terraform"],
             note="no terraform/: the first write of rancher.tf fails (message names line 188), exit 1"),
    Scenario(name="ws_no_ansible_dir", stdin=BLANK8 + YES, pre_remove=["ansible"],
             note="no ansible/: machines are provisioned, then `echo >> ansible/hosts` fails (line 123), exit 1"),
    Scenario(name="ws_no_vars_dir", stdin=BLANK8 + YES, pre_remove=["ansible/roles/ranchermaster/vars"],
             note="inventory written, vars.yml cannot be (line 129), exit 1"),
    Scenario(name="ws_hosts_is_a_directory", stdin=BLANK8 + YES, pre_files={"ansible/hosts/keep": ""},
             note="`rm -f ansible/hosts` on a directory fails silently (2>/dev/null) and errexit ends the run"),
    Scenario(name="ws_config_is_a_directory", stdin="yes\n" + BLANK8 + YES, pre_files={"config/keep": ""},
             note="a directory named config counts as an old configuration; clean removes it"),
    Scenario(name="ws_clean_tmp_is_a_file", args=["-c"], stdin="yes\n", pre_files={"tmp": "a file\n"},
             note="clean writes the rewritten ansible.cfg through ./tmp: an existing file of that name is overwritten and moved"),
    Scenario(name="ws_clean_without_ansible_dir", args=["-c"], stdin="yes\n", pre_remove=["ansible"],
             note="clean with no ansible/: sed fails quietly, an empty ./tmp is left behind"),
    # --- tools that are not installed (bash: "command not found", status 127) ---
    Scenario(name="missing_triton", stdin="yes\n", missing_tools=["triton"],
             note="no triton CLI: eval of nothing, empty SDC_*, menus empty; the first ssh file 'matches' the empty id"),
    Scenario(name="missing_ssh_keygen", stdin="yes\n", missing_tools=["ssh-keygen"],
             note="no ssh-keygen: messages hidden by 2>/dev/null, key never found, clean, exit 1"),
    Scenario(name="missing_terraform", stdin=BLANK8 + YES, missing_tools=["terraform"],
             note="no terraform: rancher.tf is written, then errexit with 127"),
    Scenario(name="missing_ansible", stdin=BLANK8 + YES, missing_tools=["ansible-playbook"],
             note="no ansible-playbook: machines exist, inventory written, errexit with 127"),
    Scenario(name="missing_terraform_on_clean", args=["-c"], stdin="yes\n", pre_files=_RAN_FILES, missing_tools=["terraform"],
             note="clean without terraform: the failed destroy is swallowed, files are removed anyway"),
    # --- dashboard wait loop (setup.sh:59-85) ---------------------------------
    Scenario(name="wait_not_on_15s_boundary", stdin=BLANK8 + YES, date_seq=["1500000001", "1500000002", "1500000014"],
             note="curl is only tried when epoch % 15 == 0"),
    Scenario(name="wait_not_ready_then_up", stdin=_answers("", "", "", "", "2", "", "", "", "yes"),
             curl_seq=["<html>starting</html>\n", "", "<title>KUBERNETES dashboard</title>\n"],
             ssh_ps={"10.0.1.1": _PS_NONE, "10.0.1.2": _PS_NONE},
             note="not ready: count dashboard containers over ssh on every node, keep polling"),
    Scenario(name="wait_stuck_dashboard_killed", stdin=_answers("", "", "", "", "2", "", "", "", "yes"),
             curl_seq=["503 Service Unavailable\n", "503 service unavailable\n", "<title>Kubernetes Dashboard</title>\n"],
             ssh_ps={"10.0.1.1": _PS_TWO_DASHBOARDS, "10.0.1.2": _PS_NONE},
             note="exactly two dashboard containers + Service Unavailable: docker stop -t0 the pod container"),
    Scenario(name="wait_two_containers_but_other_error", stdin=_answers("", "", "", "", "2", "", "", "", "yes"),
             curl_seq=["502 Bad Gateway\n", "502 Bad Gateway\n", "kubernetes\n"],
             ssh_ps={"10.0.1.1": _PS_TWO_DASHBOARDS, "10.0.1.2": _PS_NONE},
             note="two containers but not 'Service Unavailable': nothing is stopped"),
    Scenario(name="wait_three_containers", stdin=BLANK8 + YES, curl_seq=["503 Service Unavailable\n", "kubernetes\n"],
             ssh_ps={"10.0.1.1": _PS_TWO_DASHBOARDS + "9999eeee0000        gcr.io/x   \"/d\"   k8s_kubernetes-dashboard.second\n"},
             note="three dashboard containers is not 'exactly two': no second curl, nothing stopped (setup.sh:71)"),
    Scenario(name="wait_ssh_fails", stdin=BLANK8 + YES, curl_seq=["nope\n", "kubernetes\n"], ssh_exit={"10.0.1.1": 255},
             note="ssh failures while counting are swallowed by '|| true'"),
    Scenario(name="wait_curl_fails", stdin=BLANK8 + YES, fail={"curl---connect-timeout": 7},
             note="cRetVal=$(curl …) under errexit: a refused connection ends the script with curl's status"),
    Scenario(name="wait_binary_output_from_tools", stdin=_answers("", "", "", "", "2", "", "", "", "yes"),
             curl_seq=["starting\x00\n", "503 Service Unavailable\x00\n", "503 Service Unavailable\n", "kubernetes\n"],
             ssh_ps={"10.0.1.1": _PS_TWO_DASHBOARDS + "\x00\n", "10.0.1.2": "\x01\x1b \x7f k8s_kubernetes-dashboard.z\n" + _PS_TWO_DASHBOARDS},
             note="a NUL byte makes the input a binary file for grep: no lines, 'binary file matches' on stderr instead -- so "
                  "node 1's containers are not counted; $(...) drops NULs with a warning; wc -w counts only words that hold a "
                  "printable character"),
    # --- clean (setup.sh:484-521) ----------------------------------------------
    Scenario(name="clean_fresh_yes", args=["-c"], stdin="yes\n", note="nothing to clean: still rewrites ansible.cfg"),
    Scenario(name="clean_fresh_no", args=["-c"], stdin="no\n", note="decline: exit 0"),
    Scenario(name="clean_garbage_then_yes", args=["-c"], stdin="y\nYes\n\nyes\n", note="re-prompts until yes/no"),
    Scenario(name="clean_eof", args=["-c"], stdin="", note="EOF at the clean prompt: exit 1"),
    Scenario(name="clean_after_run", args=["-c"], stdin="yes\n", pre_files=_RAN_FILES, known_hosts="10.0.0.1 ssh-rsa AAAA\n",
             note="full lifecycle: terraform destroy, known_hosts scrub, every generated file removed; ansible/tmp survives"),
    Scenario(name="clean_after_run_no_known_hosts", args=["-c"], stdin="yes\n", pre_files=_RAN_FILES,
             note="without ~/.ssh/known_hosts no ssh-keygen -R calls"),
    Scenario(name="clean_destroy_fails", args=["-c"], stdin="yes\n", pre_files=_RAN_FILES, fail={"terraform-destroy": 1},
             note="terraform destroy errors are swallowed (|| true)"),
    Scenario(name="clean_masters_without_hosts", args=["-c"], stdin="yes\n", pre_files={"terraform/masters.ip": "10.0.0.1\n"},
             note="masters.ip without hosts.ip: `cat hosts.ip` fails under errexit before the prompt"),
    Scenario(name="clean_keygen_R_fails", args=["-c"], stdin="yes\n", pre_files=_RAN_FILES, known_hosts="x\n", fail={"ssh-keygen--R": 1},
             note="ssh-keygen -R failing ends the clean half-way (errexit)"),
    Scenario(name="clean_other_arg_ignored", args=["--clean"], stdin=BLANK8 + YES, note="only -c means clean; anything else is a normal run"),
]

_STARTING = "<html>starting</html>\n"
_CFG = "[defaults]\nhost_key_checking = False\nprivate_key_file =\n"

EDGE += [
    # --- the clock reading is an arithmetic expression (setup.sh:64) -------------
    Scenario(name="wait_clock_word", stdin=BLANK8 + YES, date_seq=["abc", "1500000001"], curl_seq=[_STARTING],
             note="a bare word is an unset variable = 0: `! ((abc % 15))` fires on the first reading"),
    Scenario(name="wait_clock_negative", stdin=BLANK8 + YES, date_seq=["-15", "-14", "- 30"], curl_seq=[_STARTING],
             note="-15 and `- 30` are divisible by 15, -14 is not"),
    Scenario(name="wait_clock_leading_zero", stdin=BLANK8 + YES, date_seq=["08", "010", "017"], curl_seq=[_STARTING],
             note="08: 'value too great for base', status 1, `!` makes it fire; 010 is octal 8 (no poll), 017 is 15 (poll)"),
    Scenario(name="wait_clock_empty", stdin=BLANK8 + YES, date_seq=["", " "], curl_seq=[_STARTING],
             note="no reading at all: `(( % 15))` is a syntax error (operand expected) and the poll fires"),
    Scenario(name="wait_clock_two_words", stdin=BLANK8 + YES, date_seq=["15 30", "1500000000 x"], curl_seq=[_STARTING],
             note="two words: 'syntax error in expression', fires"),
    Scenario(name="wait_clock_bases_and_operators", stdin=BLANK8 + YES,
             date_seq=["0x1e", "0x1f", "16#1E", "2#102", "65#1", "1/0", "14+1", "2**4-1", "1.5", "99999999999999999999", "7 ?", "(15"],
             curl_seq=[_STARTING] * 9,
             note="hex/base# literals, operators, division by zero, 64-bit wrap-around, and bash's wording for each error"),
    Scenario(name="wait_clock_names_script_variables", stdin=_answers("", "", "", "", "2", "", "", "", "yes"),
             date_seq=["KUBERNETES_NUMBER_OF_NODES", "i", "i*5", "yn", "master", "RANCHER_MASTER_HOSTNAME", "HOST_PACKAGE",
                       "LINENO", "DASHBOARD_CONTAINER_COUNT", "t=45", "t", "t++ - 31", "t"],
             curl_seq=[_STARTING] * 6,
             note="an identifier is looked up among the script's own variables (i is 3 after the node loop, yn is 'yes' = "
                  "unset variable yes, the quoted hostname is an error) and assignments persist between readings"),
    Scenario(name="wait_clock_subscript_error", stdin=BLANK8 + YES, date_seq=["i[1+]"],
             note="an error inside an array subscript makes bash drop the whole top-level command: main ends, status 1"),
    Scenario(name="wait_clock_subscripts", stdin=BLANK8 + YES, date_seq=["i[0]", "i[1]", "i[-1]", "i[@]", "i[0"],
             curl_seq=[_STARTING] * 3,
             note="a scalar is element 0 of itself; other elements are unset; a negative or unterminated subscript is reported"),
    # --- text from ./config reaches the node loop's test expression (setup.sh:148) ---
    Scenario(name="count_expression_via_description",
             stdin=_answers("", "-e x\\\\nKUBERNETES_NUMBER_OF_NODES=1+1\\\\nPAD=", "", "", "", "", "", "", "yes"),
             note="an echo -e newline smuggles a second KUBERNETES_NUMBER_OF_NODES line into ./config; exportVars exports "
                  "it and `i <= 1+1` provisions two workers"),
    Scenario(name="count_invalid_via_name",
             stdin=_answers("-e a\\\\nKUBERNETES_NUMBER_OF_NODES=2", "", "", "", "", "", "", "", "yes"),
             note="same route, but the closing quote of the name lands in the count: `i <= 2\"` is an arithmetic error, "
                  "the for command fails and errexit ends the run after the master block"),
    # --- a word bash cannot evaluate reaches [[ -gt ]] (setup.sh:431) ----------------
    Scenario(name="pkg_arith_operands",
             stdin=_answers("", "", "", "", "", "", "", "-e \\\\055e 3\\\\\\\\nzz", "-e \\\\055e 4\\\\\\\\n077", "-e \\\\055e 5\\\\\\\\n1.5", "yes"),
             note="echo -e '\\055e ...' prints a leading -e, so the *next* echo makes two lines: grep passes on the first, "
                  "and the second word is an operand of [[ -gt ]] (setup.sh:431): an unset name is 0 (below 1), 077 is octal 63 "
                  "(above the count of 18), and 1.5 is an arithmetic error, which makes both tests false: in range"),
    Scenario(name="pkg_arith_octal_operand",
             stdin=_answers("", "", "", "", "", "", "", "-e \\\\055e 6\\\\\\\\n022", "yes"),
             note="022 is octal 18, just inside 1..18 (decimal 22 would not be)"),
    # --- tilde expansion (setup.sh:216-217,504) ---------------------------------------
    Scenario(name="home_glob_chars", stdin=BLANK8 + YES, home_name="h*me", home_decoys=["home"],
             note="HOME=/.../h*me beside /.../home: the result of ~ is never globbed, ssh-keygen gets one path"),
    Scenario(name="home_with_space", stdin=BLANK8 + YES, home_name="my home",
             note="... and never split"),
    Scenario(name="home_unset", stdin="yes\n", unset_env=["HOME"],
             note="HOME unset: ~ is the password-database home of the user, not the empty string"),
    Scenario(name="home_empty", stdin="yes\n", env={"HOME": ""},
             note="HOME set but empty: ~ expands to nothing, `ls /.ssh`"),
    Scenario(name="home_clean_known_hosts_glob", args=["-c"], stdin="yes\n", pre_files=_RAN_FILES, known_hosts="h\n",
             home_name="h[o]me", home_decoys=["home"],
             note="[[ -e ~/.ssh/known_hosts ]] with pattern characters in HOME: still a literal path"),
    # --- `triton env` output that refers to variables (eval at setup.sh:210) ------------
    Scenario(name="triton_env_refers_to_variables", stdin=BLANK8 + YES,
             triton_env='PREFIX=https://$HOME\nexport SDC_URL="${PREFIX}/x"\nexport SDC_ACCOUNT=$UNSET_ONE"acct"\n'
                        'export SDC_KEY_ID="%s"\nexport TRITON_PROFILE=~/profile:~\n' % DEFAULT_KEY_ID,
             note="$NAME, ${NAME} and ~ in assignments execute nothing and are expanded as bash's eval expands them"),
    Scenario(name="triton_env_exports_names_without_values", stdin=_answers("", "", "", "", "2", "1,2", "", "", "yes"),
             triton_env=triton_env_text() + "export i network\nexport master yn\nunset yn\n",
             note="`export NAME` before NAME has a value: it is exported the moment the script assigns it (the loop "
                  "counter i reaches terraform as i=3); unset takes the attribute away again"),
    Scenario(name="triton_env_garbled_command", stdin=BLANK8 + YES,
             triton_env='export SDC_URL="u"\nexp\udcffort SDC_ACCOUNT="a"\nexport SDC_KEY_ID="%s"\n' % DEFAULT_KEY_ID,
             note="a byte flipped in `export`: for eval that is a command name; it does not exist, so nothing runs -- bash's "
                  "`command not found` (the name in $'...' quoting, the line counted from the eval at 210), status 127"),
    Scenario(name="triton_env_invalid_identifier", stdin=BLANK8 + YES,
             triton_env='export SDC_URL="u"\nexport SDC-ACCOUNT="a" 2nd=x\nexport SDC_KEY_ID="%s"\n' % DEFAULT_KEY_ID,
             note="export rejects the word itself: `not a valid identifier`, status 1, and errexit ends the run"),
    Scenario(name="triton_env_unsets_pwd", stdin=BLANK8 + YES, triton_env=triton_env_text() + "unset PWD OLDPWD\n",
             note="unset takes the export attribute with it: `cd` goes on working (from bash's own idea of the directory) and "
                  "sets both variables again, but terraform and ansible no longer find them in their environment"),
    Scenario(name="triton_env_unsets_pwd_only", stdin=BLANK8 + YES, triton_env=triton_env_text() + "unset PWD\n",
             note="... while OLDPWD keeps its attribute: absent after the first cd (it was given the value of an unset PWD), "
                  "exported again from the second cd on; PWD stays unexported"),
    # --- children that read the stdin they inherit ------------------------------------
    Scenario(name="stdin_taken_by_ssh_keygen", stdin="eaten by ssh-keygen\n" + BLANK8 + YES, stdin_eaters={"ssh-keygen--E": 1},
             note="stdin is read a byte at a time, so a child sees exactly what is left: here the first line"),
    Scenario(name="stdin_taken_by_triton", stdin=_answers("menu", "resolver", "", "", "", "", "", "", "", "", "yes") + "tail\n",
             stdin_eaters={"triton-networks": 1},
             note="`triton networks` heads six pipelines; each call takes the line the next prompt would have read"),
    Scenario(name="stdin_taken_by_terraform", stdin=BLANK8 + YES + "approve\nsecond\nthird\n", stdin_eaters={"terraform-apply": 2},
             note="terraform asking for approval reads what follows the confirmation"),
    Scenario(name="stdin_taken_by_ssh_in_wait_loop", stdin=BLANK8 + YES + "one\ntwo\nthree\n", stdin_eaters={"ssh--o": 1},
             curl_seq=[_STARTING], note="ssh without -n in the wait loop drains the script's stdin"),
    Scenario(name="stdin_taken_by_destroy_on_clean", args=["-c"], stdin="yes\nfor terraform\n", pre_files=_RAN_FILES,
             stdin_eaters={"terraform-destroy": 1}, note="clean: terraform destroy inherits stdin too"),
    # --- odd workspaces: symbolic links ---------------------------------------------------
    Scenario(name="ws_rancher_tf_dangling_symlink", stdin=BLANK8 + YES, pre_symlinks={"terraform/rancher.tf": "elsewhere.tf"},
             note="the previous-run guard is [ -e ] (setup.sh:14), false for a dangling link: the run proceeds and writes through it"),
    Scenario(name="ws_hosts_symlink_to_dev_null", stdin=BLANK8 + YES, pre_symlinks={"ansible/hosts": "/dev/null"},
             note="rm -f removes the link, the inventory is a regular file afterwards"),
    Scenario(name="ws_ansible_cfg_is_a_symlink", stdin=BLANK8 + YES, pre_remove=["ansible/ansible.cfg"],
             pre_files={"ansible/real.cfg": _CFG}, pre_symlinks={"ansible/ansible.cfg": "real.cfg"},
             note="sed reads through the link, mv replaces the link itself (setup.sh:133): real.cfg is untouched"),
    Scenario(name="ws_clean_with_symlinks", args=["-c"], stdin="yes\n", pre_files={"ansible/real.cfg": _CFG, "elsewhere/keep": "k\n"},
             pre_remove=["ansible/ansible.cfg"],
             pre_symlinks={"ansible/ansible.cfg": "real.cfg", "terraform/rancher.tf": "../elsewhere/keep", "tmp/link": "../elsewhere",
                           "terraform/.terraform": "../elsewhere"},
             note="rm -rf removes links, never what they point to"),
    # --- odd workspaces: a directory or a dead link where the script expects a file -----------
    Scenario(name="ws_clean_retry_is_a_directory", args=["-c"], stdin="yes\n",
             pre_files={"ansible/a.retry/keep": "k\n", "ansible/b.retry": "h\n", "ansible/ansible.cfg.tmp/keep": "k\n", "config": "x\n"},
             note="rm -f goes through all its arguments, complains on the script's stdout (2>&1 >> /dev/null) and fails at "
                  "the end: errexit, ./config survives"),
    Scenario(name="ws_cfg_is_a_directory", stdin=BLANK8 + YES, pre_remove=["ansible/ansible.cfg"], pre_files={"ansible/ansible.cfg/keep": "k\n"},
             note="sed opens a directory and fails on read ('read error on', not 'can't read'); tmp.cfg stays behind, the run goes on"),
    Scenario(name="ws_clean_cfg_is_a_directory", args=["-c"], stdin="yes\n", pre_remove=["ansible/ansible.cfg"],
             pre_files={"ansible/ansible.cfg/keep": "k\n"}, note="the same at setup.sh:511, where sed is given the longer path"),
    Scenario(name="ws_config_tmp_is_a_dead_link", stdin=BLANK8 + YES, pre_symlinks={"config.tmp": "/nonexistent/x"},
             note="grep ... > config.tmp: the redirection fails first (message names line 544), errexit"),
    Scenario(name="ws_config_tmp_is_a_directory", stdin=BLANK8 + YES, pre_files={"config.tmp/keep": "k\n"},
             note="... and so it does on a directory"),
    Scenario(name="ws_config_tmp_links_to_dev_null", stdin=BLANK8 + YES, pre_symlinks={"config.tmp": "/dev/null"},
             note="exportVars reads back what config.tmp holds: nothing. No variable is exported, values keep the shapes they "
                  "had as plain shell variables (no literal quotes), SDC_KEY is not set at all"),
    Scenario(name="ws_tmp_cfg_links_to_ansible_cfg", stdin=BLANK8 + YES, pre_symlinks={"ansible/tmp.cfg": "ansible.cfg"},
             note="`sed ... ansible.cfg > tmp.cfg` empties ansible.cfg through the link before sed reads it, and mv refuses: "
                  "'are the same file', errexit"),
    Scenario(name="ws_masters_ip_is_a_directory", stdin=BLANK8 + YES, terraform_no_ips=True,
             pre_files={"terraform/masters.ip/keep": "k\n", "terraform/hosts.ip": "10.0.1.1\n"},
             note="[[ -e ]] is satisfied by a directory; `cat terraform/masters.ip >> ansible/hosts` is not: cat's message, errexit"),
    Scenario(name="ws_clean_ip_files_are_directories", args=["-c"], stdin="yes\n", known_hosts="h\n",
             pre_files={"terraform/masters.ip/keep": "k\n", "terraform/hosts.ip/keep": "k\n"},
             note="clean: `cat masters.ip 2>/dev/null` before the prompt fails silently and errexit ends it there"),
    # --- bytes ------------------------------------------------------------------------------
    Scenario(name="text_70kb_line", stdin=_answers("x" * 70000, "", "", "", "", "", "", "", "yes"),
             note="a 70 KB answer: no line-length limit anywhere on the path"),
    Scenario(name="text_raw_bytes", stdin="\udcff\udcfe\x01\x1b[31m\x7f\n" + "\udc80 \x02\t\x03\n" + "\n" * 6 + YES,
             note="0xFF/0xFE, C0 controls and ESC are data; 0x01 and 0x7f are bash's internal quoting bytes"),
    # --- the operator's environment -----------------------------------------------------------
    Scenario(name="env_lang_utf8_no_lc_all", stdin=_answers("caf\u00e9 \u2615", "", "", "", "", "2,1", "", "", "yes"),
             env={"LANG": "C.UTF-8"}, unset_env=["LC_ALL"], note="LANG=C.UTF-8 with LC_ALL unset: same bytes, same order"),
    Scenario(name="env_utf8_bytes_that_are_no_characters", stdin=_answers("ok", "bad \udcff byte", "", "", "2", "", "", "", "yes"),
             env={"LANG": "C.UTF-8"}, unset_env=["LC_ALL"],
             networks=triton_table(DOC_NETWORKS).replace("NAME  ", "NAME \udcff", 1),
             note="in a UTF-8 locale grep holds back lines that are not valid UTF-8 ('binary file matches'): the description "
                  "never reaches config.tmp and is not exported; `.` does not match such a byte, so the header row survives "
                  "grep -v '^NAME.*ID$' only to be held back as well"),
    Scenario(name="env_utf8_key_name_with_a_stray_byte", stdin=BLANK8 + YES, env={"LANG": "C.UTF-8"}, unset_env=["LC_ALL"],
             ssh_files={"k\udcffpub": "K\n", "id_rsa": "K\n"}, keygen={"k\udcffpub": "2048 MD5:%s c (RSA)\n" % DEFAULT_KEY_ID},
             keygen_sha={},
             note="sed 's/.pub$//' in a UTF-8 locale: the dot does not match a byte that is no character, so 'k<ff>pub' keeps "
                  "its name (in the C locale it would be asked about as 'k')"),
    Scenario(name="env_utf8_question_mark_is_a_character", stdin=_answers("?", "??", "", "", "", "", "", "", "yes"),
             env={"LANG": "C.UTF-8"}, unset_env=["LC_ALL"], pre_files={"\u00e9": "", "\u65e5": ""},
             note="pathname expansion of an answer in a UTF-8 locale: `?` matches the files called é and 日 (one character "
                  "each), `??` neither; in the C locale it is the other way round for é (two bytes)"),
    Scenario(name="text_question_marks_in_the_c_locale", stdin=_answers("?", "??", "", "", "", "", "", "", "yes"),
             pre_files={"\u00e9": "", "\u65e5": ""}, note="... as here: `??` is é, `?` is nothing"),
    Scenario(name="env_utf8_unicode_escapes", stdin=_answers("-e caf\\\\u00e9 \\\\U0001F600", "-e \\\\ud800 \\\\u41", "", "", "", "", "", "", "yes"),
             env={"LANG": "C.UTF-8"}, unset_env=["LC_ALL"],
             networks=triton_table([("Joyent-SDC-Public", "31428241-4878-47d6-9fba-9a8436b596a4"), ("caf\\u00e9", "0b206464-d655-4723-a848-86d0f28764c8")]),
             note="echo -e writes \\uHHHH in the locale's character set: UTF-8 here -- in the answers and in the menu, "
                  "whose rows go through echo -e as well"),
    Scenario(name="text_unicode_escapes_in_the_c_locale", stdin=_answers("-e caf\\\\u00e9 \\\\U0001F600", "-e \\\\ud800 \\\\u41", "", "", "", "", "", "", "yes"),
             networks=triton_table([("Joyent-SDC-Public", "31428241-4878-47d6-9fba-9a8436b596a4"), ("caf\\u00e9", "0b206464-d655-4723-a848-86d0f28764c8")]),
             note="... in the C locale only ASCII can be written: \\u00e9 comes back as \\u00E9, \\u41 is A"),
    Scenario(name="env_cdpath_names_the_workspace", stdin=BLANK8 + YES, env={"CDPATH": "/nonexistent:.:<SBX>/ws"},
             note="CDPATH is searched before the current directory; `cd terraform` / `cd ansible` found through the non-empty "
                  "entry '.' announce the new directory on stdout (three extra lines)"),
    Scenario(name="env_opterr_is_reset", stdin=BLANK8 + YES, env={"OPTERR": "0", "OPTIND": "7"},
             note="bash sets OPTERR=1 and OPTIND=1 at start-up whatever it inherits; the export attribute stays, so that is what "
                  "terraform and ansible-playbook find in their environment"),
    Scenario(name="env_lc_all_names_no_installed_locale", stdin=_answers("-e caf\\\\u00e9", "", "", "", "", "", "", "", "yes"),
             env={"LC_ALL": "xx_XX.UTF-8"},
             note="a locale forwarded from elsewhere: bash warns (`setlocale: LC_ALL: cannot change locale`), runs in C -- and "
                  "still writes \\u00e9 as UTF-8, because where wctomb() gives up it asks iconv for the character set in the name"),
    Scenario(name="env_lc_ctype_invalid_lang_utf8", stdin=_answers("-e caf\\\\u00e9", "", "", "", "", "", "", "", "yes"),
             env={"LC_CTYPE": "xx_XX", "LANG": "C.UTF-8"}, unset_env=["LC_ALL"],
             networks=triton_table([("Joyent-SDC-Public", "31428241-4878-47d6-9fba-9a8436b596a4"), ("caf\xe9", "0b206464-d655-4723-a848-86d0f28764c8")]),
             note="bash sets LANG for every category and skips the LC_CTYPE it cannot set: it runs in UTF-8; grep, sed and sort call "
                  "setlocale(LC_ALL, \"\"), which fails as a whole, and run in C"),
    Scenario(name="env_path_without_usr_bin", stdin=BLANK8 + YES, env={"PATH": "<SBX>/bin:/bin"},
             note="PATH without /usr/bin"),
    Scenario(name="env_shell_level_and_directories", stdin=BLANK8 + YES,
             env={"SHLVL": "5", "PWD": "/bogus/elsewhere", "OLDPWD": "/nonexistent"},
             note="what bash exports by itself: SHLVL one up (one down again for the lone command of a $(...)), PWD "
                  "re-established at start-up and following each cd, an OLDPWD that is no directory dropped"),
    Scenario(name="env_shell_level_garbage", stdin=BLANK8 + YES, env={"SHLVL": "abc", "OLDPWD": "/tmp"},
             note="a SHLVL that is no number counts as 0; an OLDPWD that is a directory is passed on until the first cd"),
    Scenario(name="env_shell_level_too_high", stdin=BLANK8 + YES, env={"SHLVL": "999"},
             note="shell level 1000: bash warns and resets to 1"),
    Scenario(name="env_no_locale_at_all", stdin=_answers("caf\u00e9", "", "", "", "", "4,1", "", "", "yes"), unset_env=["LC_ALL"],
             note="no LC_ALL, no LANG: children must not find an LC_CTYPE that the reference's children do not have"),
    Scenario(name="env_exported_names_of_script_variables", stdin=_answers("", "", "", "", "2", "1,3", "", "7", "yes"),
             env={"tmp": "t", "networks": "nn", "packages": "pp", "foundKey": "fk", "countNetwork": "99", "values": "vv",
                  "gotValidInput": "g", "network": "n", "package": "pk", "i": "5", "yn": "q", "line": "l", "master": "m",
                  "f": "ff", "node": "nd", "cRetVal": "c", "tmp_ValidatedInput": "tv", "publicNetworkLocation": "9"},
             curl_seq=[_STARTING],
             note="the operator's shell happens to export names the script uses: assignments to them stay exported, a "
                  "`local` takes over the export attribute (children of the function see the local value), what a $(...) "
                  "assigns is gone after it, and `triton` inside getNetworkIDs' loop finds the current index in `network`"),
    Scenario(name="env_own_switches_do_not_leak", stdin=BLANK8 + YES,
             env={"TK8S_PROGRAM_": "1", "KUBERNETES_TK8S": "kept"},
             note="variables that merely look like the replacement's own switches pass through to terraform/ansible"),
]

#: scenarios whose stderr carries bash's own diagnostics: recorded under both ways
#: of starting the script, since bash labels them with the name it was given
_DIAGNOSTIC = ["text_embedded_newline", "text_nul_and_c_escape", "clean_after_run", "missing_triton", "missing_ssh_keygen",
               "missing_terraform", "missing_ansible", "missing_terraform_on_clean", "ws_no_terraform_dir", "ws_no_ansible_dir",
               "ws_no_vars_dir", "ws_clean_without_ansible_dir", "key_name_breaks_the_sed_program", "wait_clock_leading_zero",
               "count_invalid_via_name", "pkg_arith_operands", "config1_defaults_1worker"]
EDGE += [dataclasses.replace(sc, name=sc.name + "__bash_setup_sh", invocation="bash setup.sh",
                             note="started as `bash setup.sh` instead of `./setup.sh`: " + sc.note)
         for sc in CONFIGS + EDGE if sc.name in _DIAGNOSTIC]

ALL: List[Scenario] = CONFIGS + EDGE
BY_NAME: Dict[str, Scenario] = {s.name: s for s in ALL}
assert len(BY_NAME) == len(ALL), "duplicate scenario names"

# ---------------------------------------------------------------------------
# interactive cases: (scenario, answers typed at the terminal one prompt at a time)
# ---------------------------------------------------------------------------
TTY_CASES = [
    (Scenario(name="tty_quickstart", note="README.md:112-121: every prompt with its default, '2' typed at the node count"),
     ["", "", "", "", "2", "", "", "", "yes"]),
    (Scenario(name="tty_reprompts", note="invalid answers: error lines, then the same prompt again"),
     ["my env", "", "1bad", "km1", "w", "wk", "0", "3", "9", "2,4", "x", "1", "99", "10", "maybe", "yes"]),
    (Scenario(name="tty_decline", note="'no' at the confirmation"), ["", "", "", "", "", "", "", "", "no"]),
    (Scenario(name="tty_clean", args=["-c"], pre_files={"terraform/masters.ip": "10.0.0.1\n", "terraform/hosts.ip": "10.0.1.1\n"},
              note="clean prompt wording when machines exist (setup.sh:487-493)"), ["what", "yes"]),
    (Scenario(name="tty_clean_fresh", args=["-c"], note="clean prompt wording on a fresh tree (setup.sh:495)"), ["no"]),
    (Scenario(name="tty_key_not_found", keygen={"id_rsa": "2048 MD5:ff:ff:ff:ff other (RSA)\n"},
              note="no key matches the profile: the three error lines, then the clean prompt"), ["maybe", "yes"]),
    (Scenario(name="tty_stale_config", pre_files={"config": "OLD=1\n"},
              note="an old ./config: clean is offered first, then the ordinary prompts follow"),
     ["yes", "", "", "", "", "", "", "", "", "no"]),
    (Scenario(name="tty_clock_is_not_a_number", date_seq=["08", "15 30"], curl_seq=["<html>starting</html>\n"],
              note="bash's arithmetic diagnostics land on the terminal between the progress dots"),
     ["", "", "", "", "", "", "", "", "yes"]),
]
