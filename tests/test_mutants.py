"""Do the goldens have teeth?  Each mutant below breaks one behaviour of the
path the way a careless rewrite would (numeric instead of lexical sort, quotes
not stripped, a wrong exit status ...).  Every mutant must be caught by at least
one recorded scenario, and the scenario that catches it is named, so the
catalogue's coverage claims are checked rather than asserted."""
import os

import pytest

from tritonk8ssupervisor_b200 import bashsem, supervisor, textops
from tritonk8ssupervisor_b200.conformance import replay as rp

GOLDENS = {os.path.basename(p)[:-5]: rp.load_golden(p) for p in rp.list_goldens()}


def failing(names=None):
    out = []
    for name, doc in GOLDENS.items():
        if names and name not in names:
            continue
        try:
            expected, got = rp.replay_in_process(doc)
        except Exception:          # a mutant may crash or trip the runaway-loop guard: that counts as caught
            out.append(name)
            continue
        if expected.diff(got):
            out.append(name)
    return out


def test_unmutated_passes_everything():
    assert failing() == []


def _numeric_sort(text):
    ls = textops.lines(text)
    return textops.unlines(sorted(ls, key=lambda l: (0, int(l)) if l.isdigit() else (1, l)))


MUTANTS = [
    # (description, object, attribute, replacement, scenarios that must catch it)
    ("index lists sorted numerically instead of lexically", textops, "sort_lines", _numeric_sort, {"net_lexical_sort"}),
    ("duplicates in index lists not removed", textops, "uniq", lambda t: t, {"config3_custom_names_multinet", "net_out_of_range"}),
    ("double quotes not stripped", textops, "strip_dquotes", lambda t: t, {"config1_defaults_1worker"}),
    ("read keeps backslashes (read -r)", bashsem, "read_builtin",
     lambda nb: (lambda raw: (raw.decode("utf-8", "surrogateescape").strip(" \t"), True))(
         bytes(iter(lambda: (lambda b: b if b not in (None, 10) else None)(nb()), None))), {"text_backslashes"}),
    ("echo does not take -n/-e from the data", bashsem, "echo_builtin", lambda args: " ".join(args) + "\n", {"text_echo_options"}),
    ("no pathname expansion of answers", bashsem, "glob_word", lambda w, cwd: [w], {"text_glob_chars"}),
    ("no field splitting (blanks kept)", bashsem, "split_fields", lambda s: [s] if s else [], {"text_quotes_and_blanks"}),
    ("trailing newlines kept by $(...)", bashsem, "cmdsubst", lambda o: o.replace("\x00", ""), {"config1_defaults_1worker"}),
    ("[[ == ]] compares literally", bashsem, "pattern_match", lambda s, p: s == p, {"net_default_is_a_pattern"}),
    ("hostname may be one character", supervisor, "HOSTNAME_RE", __import__("re").compile(r"[a-zA-Z][0-9a-zA-Z]*"), {"bad_hostnames"}),
    ("ten nodes allowed", supervisor, "NODE_COUNT_RE", __import__("re").compile(r"[1-9][0-9]?"),
     {"config5_10workers_rejected_then_9", "bad_node_counts"}),
    ("three-digit indices allowed", supervisor, "INDEX_LIST_RE", r"^[1-9][0-9]*(,[1-9][0-9]*)*$", {"net_three_digit_indices"}),
    ("awk prints the last field of one-field rows too", textops, "awk_last_field_if_many",
     lambda t: textops.unlines([l.split()[-1] for l in textops.lines(t) if l.split()]), {"net_header_lookalike"}),
    ("vars.yml takes the first master instead of the last", textops, "tail_1",
     lambda t: t.split("\n")[0] + "\n" if t else "", {"stale_ip_files_are_appended_to"}),
    ("word count instead of exact container count: wc -w counts lines", textops, "wc_w", lambda t: t.count("\n") + 5,
     {"wait_stuck_dashboard_killed", "wait_two_containers_but_other_error"}),
    ("arithmetic does not wrap at 2**64", supervisor, "_to_int64", lambda t: int(t) if t.isdigit() else 0, {"pkg_huge_index"}),
]


@pytest.mark.parametrize("desc,obj,attr,repl,must_catch", MUTANTS, ids=[m[0] for m in MUTANTS])
def test_mutant_is_caught(monkeypatch, desc, obj, attr, repl, must_catch):
    monkeypatch.setattr(obj, attr, repl)
    caught = failing(must_catch or None)
    if must_catch:
        assert set(caught) == must_catch, "%s: expected %s to fail, got %s" % (desc, sorted(must_catch), caught)
    else:
        assert caught, "%s: no recorded scenario notices" % desc


def _patched_method(name, fn):
    """(description helper) replace one Supervisor method for the in-process replay."""
    return (supervisor.Supervisor, name, fn)


def _menu_rows_sorted_raw(self, argv, kvm_only):
    text = self._triton_table(argv, 259 if kvm_only else 257)
    if kvm_only:
        text = textops.grep(text, r"-kvm-")
    text = textops.grep(text, r"^NAME.*ID$", invert=True)
    # sort BEFORE the blank->'=' rewrite (i.e. the resolver's order): looks harmless, is not
    return bashsem.cmdsubst(textops.tr(textops.tr_squeeze_spaces(textops.sort_lines(text)), " ", "="))


def _sed_literal_dot(orig):
    def sed_sub(text, pattern, repl, count=1):
        return orig(text, r"\.pub$" if pattern == r".pub$" else pattern, repl, count)
    return sed_sub


def _grep_case_sensitive(orig):
    def grep(text, pattern, invert=False, ignore_case=False):
        return orig(text, pattern, invert, False)
    return grep


def _grep_same_header_rule(orig):
    def grep(text, pattern, invert=False, ignore_case=False):
        return orig(text, r"^NAME.*ID$" if pattern == r"^NAME *ID$" else pattern, invert, ignore_case)
    return grep


def _export_without_second_read(self):
    for line in textops.lines(textops.grep(self.rt.read_text("config"), r"^$", invert=True)):
        self._export(line)


def _clean_also_ansible_tmp(orig):
    def _clean(self):
        orig(self)
        import shutil
        shutil.rmtree(self.rt.path("ansible/tmp"), ignore_errors=True)
    return _clean


def _clean_dotfiles_too(orig):
    def _clean(self):
        orig(self)
        import shutil
        shutil.rmtree(self.rt.path("tmp"), ignore_errors=True)
    return _clean


MUTANTS_2 = [
    ("menu rows sorted before the blank->'=' rewrite", supervisor.Supervisor, "_menu_rows", _menu_rows_sorted_raw,
     {"net_prefix_collision", "pkg_prefix_collision"}),
    (".pub stripped with a literal dot", textops, "sed_sub", _sed_literal_dot(textops.sed_sub), {"key_dot_matches_any_char"}),
    ("dashboard greps are case sensitive", textops, "grep", _grep_case_sensitive(textops.grep),
     {"wait_not_ready_then_up", "wait_stuck_dashboard_killed"}),
    ("resolver drops header look-alikes like the menu does", textops, "grep", _grep_same_header_rule(textops.grep),
     {"net_header_lookalike"}),
    ("exportVars does not run the lines through read again", supervisor.Supervisor, "export_vars", _export_without_second_read,
     {"text_backslashes"}),
    ("clean also empties ansible/tmp", supervisor.Supervisor, "_clean", _clean_also_ansible_tmp(supervisor.Supervisor._clean),
     {"clean_after_run", "clean_fresh_yes"}),
    ("clean removes dot files under tmp/ as well", supervisor.Supervisor, "_clean", _clean_dotfiles_too(supervisor.Supervisor._clean),
     {"clean_after_run"}),
]


@pytest.mark.parametrize("desc,obj,attr,repl,must_catch", MUTANTS_2, ids=[m[0] for m in MUTANTS_2])
def test_structural_mutant_is_caught(monkeypatch, desc, obj, attr, repl, must_catch):
    monkeypatch.setattr(obj, attr, repl)
    caught = set(failing())
    assert must_catch <= caught, "%s: expected %s among the failures, got %s" % (desc, sorted(must_catch), sorted(caught))


class _WrongExit(supervisor.Supervisor):
    def _main(self, argv):
        try:
            super()._main(argv)
        except supervisor.ScriptExit as e:
            # "fix" the statuses that look wrong: guard and `no` should fail, shouldn't they?
            raise supervisor.ScriptExit(1 if e.status == 0 and argv[:1] != ["-c"] else e.status)


def test_wrong_exit_statuses_are_caught(monkeypatch):
    from tritonk8ssupervisor_b200.conformance import modeltools
    monkeypatch.setattr(modeltools, "Supervisor", _WrongExit)
    caught = set(failing({"guard_previous_run", "verify_no", "guard_stale_config_clean_no", "key_not_found_decline_clean",
                          "config1_defaults_1worker"}))
    assert caught == {"guard_previous_run", "verify_no", "guard_stale_config_clean_no", "key_not_found_decline_clean"}
