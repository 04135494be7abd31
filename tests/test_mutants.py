"""Do the goldens have teeth?  Each mutant below breaks one behaviour of the
path the way a careless rewrite would (numeric instead of lexical sort, quotes
not stripped, a wrong exit status ...).  Every mutant must be caught by at least
one recorded scenario, and the scenario that catches it is named, so the
catalogue's coverage claims are checked rather than asserted."""
import os

import pytest

from tritonk8ssupervisor_b200 import basharith, bashsem, runtime, supervisor, textops
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
    ("arithmetic does not wrap at 2**64", basharith, "wrap", lambda v: v, {"pkg_huge_index"}),
    ("the result of ~ is pathname-expanded like the rest of the word", bashsem, "quote_for_glob", lambda t: t, {"home_glob_chars"}),
    ("HOME unset means the empty string", runtime.Runtime, "passwd_home", lambda self: "", {"home_unset"}),
    ("the previous-run guard sees dangling links ([ -L ] instead of [ -e ])", runtime.Runtime, "exists",
     lambda self, rel: os.path.lexists(self.path(rel)), {"ws_rancher_tf_dangling_symlink"}),
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


def _arith_digits_only(orig):
    def arith_eval(self, expr, lineno, command, last_status=0):
        if lineno == 64 and not expr[:-len(" % 15")].isdigit():
            return 1                     # "not a number: keep sleeping" -- what round 1 did
        return orig(self, expr, lineno, command, last_status)
    return arith_eval


def _arith_error_is_zero(orig):
    def arith_eval(self, expr, lineno, command, last_status=0):
        val = orig(self, expr, lineno, command, last_status)
        return 0 if val is None and command == "[[" else val
    return arith_eval


def _for_error_exits(orig):
    def arith_eval(self, expr, lineno, command, last_status=0):
        val = orig(self, expr, lineno, command, last_status)
        if val is None and lineno == 148:
            raise supervisor.ScriptExit(1)
        return val
    return arith_eval


def _mv_writes_through_links(src, dst):
    with open(src, "rb") as fh:
        data = fh.read()
    with open(dst, "wb") as fh:          # follows a symbolic link at dst; mv replaces it
        fh.write(data)
    os.remove(src)


_real_rm_rf = supervisor._rm_rf


def _rm_rf_following(path):
    _real_rm_rf(os.path.realpath(path))
    _real_rm_rf(path)


def _read_in_blocks(self, prompt):
    """stdin slurped into a private buffer on the first read: children that
    inherit stdin then find it empty."""
    if not hasattr(self, "_private"):
        self._private, self._ppos = self._stdin[self._pos:], 0
        self._pos = len(self._stdin)

    def nb():
        if self._ppos >= len(self._private):
            return None
        self._ppos += 1
        return self._private[self._ppos - 1]
    return bashsem.read_builtin(nb)


def _clean_shrugs_off_rm_failures(orig):
    def _clean(self):
        try:
            orig(self)
        except supervisor.ScriptExit:
            pass
    return _clean


MUTANTS_2 = [
    ("clean carries on when rm could not remove something", supervisor.Supervisor, "_clean",
     _clean_shrugs_off_rm_failures(supervisor.Supervisor._clean), {"ws_clean_retry_is_a_directory", "clean_keygen_R_fails"}),
    ("a directory is a file that cannot be read, says sed", supervisor.os.path, "isdir", lambda p: False,
     {"ws_cfg_is_a_directory", "ws_clean_cfg_is_a_directory"}),
    ("a clock reading that is not a number is skipped", supervisor.Supervisor, "arith_eval",
     _arith_digits_only(supervisor.Supervisor.arith_eval),
     {"wait_clock_word", "wait_clock_negative", "wait_clock_empty", "wait_clock_two_words",
      "wait_clock_bases_and_operators", "wait_clock_names_script_variables", "wait_clock_subscripts", "wait_clock_subscript_error"}),
    ("an operand [[ -gt ]] cannot evaluate counts as 0", supervisor.Supervisor, "arith_eval",
     _arith_error_is_zero(supervisor.Supervisor.arith_eval), {"pkg_arith_operands"}),
    ("an arithmetic error in the node loop ends the run", supervisor.Supervisor, "arith_eval",
     _for_error_exits(supervisor.Supervisor.arith_eval), {"count_invalid_via_name"}),
    ("mv writes through a symbolic link", supervisor.os, "replace", _mv_writes_through_links, {"ws_ansible_cfg_is_a_symlink"}),
    ("stdin read in blocks", __import__("tritonk8ssupervisor_b200.conformance.modeltools", fromlist=["x"]).ModelRuntime, "read",
     _read_in_blocks, {"stdin_taken_by_triton", "stdin_taken_by_terraform", "stdin_taken_by_ssh_in_wait_loop",
                       "stdin_taken_by_destroy_on_clean"}),
    ("rm -rf follows symbolic links", supervisor, "_rm_rf", lambda path: _rm_rf_following(path), {"ws_clean_with_symlinks"}),
    ("leading zeros are decimal", basharith.Evaluator, "_strlong", lambda self, num: int(num) if num.isdigit() else 0,
     {"wait_clock_leading_zero", "pkg_arith_octal_operand"}),
    ("diagnostics always labelled setup.sh", __import__("tritonk8ssupervisor_b200.conformance.scenario", fromlist=["x"]).Scenario,
     "program_label", property(lambda self: "setup.sh"), {"text_embedded_newline", "missing_terraform", "wait_clock_empty"}),

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
