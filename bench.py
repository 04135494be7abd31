#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric for cheapRoc/tritonK8ssupervisor.

    metric: offline validate/syntax-check pass rate per config   (unit: fraction of 5)

THERE IS NO GPU WORK IN THIS BENCHMARK, and none in this repository.
BASELINE.json flags the reference "TIER MISMATCH, no GPU graft target": it is
551 lines of bash generating < 3.5 KB of text per run.  The
line printed here therefore says ``"gpu_launches": 0`` and ``"roofline":
{"bound": "n/a", ...}`` -- a roofline fraction for this path would be a made-up
number -- and reports what *is* measurable:

  value         pass rate over the five configs with the path run in-process
                against in-memory tool models (inputs resident in memory: no
                fork/exec, no pipes), + structural checks
  e2e           the same pass rate through the public CLI
                (``python -m tritonk8ssupervisor_b200``) as a child process in the
                offline sandbox, talking to stub tool processes over pipes/files
  cpu_baseline  the unmodified reference ``setup.sh`` under bash in the same
                sandbox (kind "reference"), 1 core -- the script is sequential

Every leg measures all five pass criteria, determinism included (each config is
run twice and the two observations compared).  ``catalogue`` adds the whole
recorded parity surface: every fixture under tests/golden replayed in-process.

A "step" is one pass over the five configs.  Speed is secondary (SURVEY.md §8d:
"Speed is not the metric"); it is reported as ms per step and configs/s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import statistics
import subprocess
import sys
import time
from typing import Any, Dict, List, Optional

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
os.environ.setdefault("LC_ALL", "C")

from tritonk8ssupervisor_b200.conformance import harness, replay as rp  # noqa: E402
from tritonk8ssupervisor_b200.conformance.catalog import CONFIGS  # noqa: E402
from tritonk8ssupervisor_b200.conformance.modeltools import ModelSandbox  # noqa: E402
from tritonk8ssupervisor_b200.conformance.sandbox import Observation, Sandbox  # noqa: E402
from tritonk8ssupervisor_b200.conformance.scenario import Scenario  # noqa: E402

METRIC = "offline validate/syntax-check pass rate per config"
UNIT = "fraction of the 5 BASELINE configs passing"
WORKLOAD = "5 BASELINE.json configs (1, 2, 3, 5, 10->9 workers); one step = one pass over all five"
NOT_APPLICABLE = ("not applicable: the path is a control path (bash integer counters, <= 3.5 KB of text written per "
                  "config, 0 flops); BASELINE.json: 'TIER MISMATCH, no GPU graft target' (SURVEY.md §0, §8d)")


# ---------------------------------------------------------------------------
# the three ways of running one pass
# ---------------------------------------------------------------------------
def step_in_process(docs: List[dict]) -> List[harness.ConfigResult]:
    """Resident leg: path in this process against in-memory tool models."""
    out = []
    for doc in docs:
        sc = Scenario.from_json(doc["scenario"])
        expected = Observation.from_json(doc["expected"])
        res = harness.ConfigResult(name=sc.name)
        t0 = time.perf_counter()
        with ModelSandbox(harness.WORKSPACE_TEMPLATE, sc, input_overrides=doc["inputs"]) as sbx:
            obs = sbx.run()
            diffs = expected.diff(obs)
            res.parity = not diffs
            res.details += ["parity: %s" % d for d in diffs]
            harness.check_generated_tree(sbx.workspace, obs, res)
        with ModelSandbox(harness.WORKSPACE_TEMPLATE, sc, input_overrides=doc["inputs"]) as sbx:
            res.deterministic = not obs.diff(sbx.run())         # criterion (iv): a second run, same observation
        if not res.deterministic:
            res.details.append("determinism: two runs differ")
        res.wall_s = time.perf_counter() - t0
        out.append(res)
    return out


def step_cli(docs: List[dict], repeat: int) -> List[harness.ConfigResult]:
    """End-to-end leg: the public CLI as a child process in the sandbox."""
    return [harness.run_config(doc, repeat=repeat) for doc in docs]


def step_reference(docs: List[dict], ref_root: str) -> List[harness.ConfigResult]:
    """Baseline leg: the unmodified reference under bash in the same sandbox; its
    output is held to the goldens and to the same structural checks (run on the
    reference's own terraform/ansible trees)."""
    out = []
    for doc in docs:
        sc = Scenario.from_json(doc["scenario"])
        expected = Observation.from_json(doc["expected"])
        res = harness.ConfigResult(name=sc.name)
        t0 = time.perf_counter()
        with Sandbox(ref_root, sc) as sbx:
            obs = sbx.run(sc.reference_argv())
            diffs = expected.diff(obs)
            res.parity = not diffs
            res.details += ["parity: %s" % d for d in diffs]
            harness.check_generated_tree(sbx.workspace, obs, res)
        with Sandbox(ref_root, sc) as sbx:
            res.deterministic = not obs.diff(sbx.run(sc.reference_argv()))     # criterion (iv), measured here too
        if not res.deterministic:
            res.details.append("determinism: two runs differ")
        res.wall_s = time.perf_counter() - t0
        out.append(res)
    return out


def median_wall_ms(argv, env, workspace_src: str, docs: List[dict], runs: int, overrides: bool) -> Dict[str, float]:
    """wall clock around the subprocess alone (no checks), fresh
    sandbox per run, median of ``runs`` runs per config."""
    out = {}
    for doc in docs:
        sc = Scenario.from_json(doc["scenario"])
        walls = []
        for _ in range(runs):
            with Sandbox(workspace_src, sc, input_overrides=doc["inputs"] if overrides else None) as sbx:
                walls.append(sbx.run(argv, env).wall_s)
        out[sc.name] = round(1e3 * statistics.median(walls), 2)
    return out


def timed(fn, steps: int, warmup: int):
    """(results of the last step, [seconds per timed step])."""
    results = None
    for _ in range(warmup):
        results = fn()
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        results = fn()
        times.append(time.perf_counter() - t0)
    return results, times


def sample_clocks() -> Optional[Dict[str, Any]]:
    """GPU clocks are irrelevant to this benchmark; sampled once when
    nvidia-smi exists so the line has the key the driver looks for."""
    exe = shutil.which("nvidia-smi")
    if not exe:
        return None
    try:
        out = subprocess.run([exe, "--query-gpu=clocks.sm,clocks.max.sm,clocks_throttle_reasons.active",
                              "--format=csv,noheader,nounits", "-i", "0"], capture_output=True, text=True, timeout=20).stdout
        sm, mx, reasons = [x.strip() for x in out.strip().split(",")[:3]]
        return {"sm_mhz": float(sm), "sm_max_mhz": float(mx), "reasons": [reasons],
                "note": "GPU idle by design: this benchmark launches no kernels"}
    except Exception as exc:  # pragma: no cover - diagnostics only
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "nvidia-smi failed: %s" % exc}


# ---------------------------------------------------------------------------
# multi-process plumbing: the path does not use GPUs and one run of it does not
# shard.  What can be divided is the list of independent units:
# with N ranks the five configs and the recorded catalogue are dealt out round
# robin (rank r takes items r, r+N, ...), pass counts are summed over gloo and the
# time is the slowest rank's.  Total work is fixed, hence "scaling": "strong".
# ---------------------------------------------------------------------------
class Ranks:
    def __init__(self, distributed: bool = True) -> None:
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1")) if distributed else 1
        self.dist = None
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            import torch.distributed as dist  # gloo: there is nothing on the device to reduce
            dist.init_process_group(backend="gloo", rank=self.rank, world_size=self.world)
            self.dist = dist

    def barrier(self) -> None:
        if self.dist is not None:
            self.dist.barrier()

    def reduce(self, value: float, op: str) -> float:
        if self.dist is None:
            return value
        import torch
        t = torch.tensor([value], dtype=torch.float64)
        self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op))
        return float(t.item())

    def close(self) -> None:
        if self.dist is not None:
            self.dist.destroy_process_group()


def summarise(results: List[harness.ConfigResult]) -> Dict[str, Any]:
    return {r.name: {"passed": r.passed, "parity": r.parity, "terraform": r.terraform, "ansible": r.ansible,
                     "inventory": r.inventory, "deterministic": r.deterministic, "modules": r.n_modules,
                     "hosts": r.n_hosts, "details": r.details[:3]} for r in results}


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1, help="number of ranks (CPU replicas; no GPU is used)")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--triton-latency-ms", type=int, default=0,
                    help="opt-in latency MODEL: make every stubbed `triton` call take this long (a real call is a node "
                         "start-up plus a CloudAPI round trip) and time config 2 under the reference, the replacement, "
                         "and the replacement with TK8S_CACHE_TRITON=1")
    ap.add_argument("--median-runs", type=int, default=20,
                    help="runs per config for the per-config median wall times (SURVEY.md §8d asks for >= 20)")
    args = ap.parse_args(argv)
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    docs = [harness.golden_for(sc) for sc in CONFIGS]
    if args.impl == "reference" and int(os.environ.get("RANK", "0")) != 0:
        return 0          # under torchrun rank 0 alone runs the reference arm; no process group is formed
    ranks = Ranks(distributed=args.impl != "reference")

    from oracle.reference_runner import find_reference  # the baseline legs are the only users of oracle/
    ref_root = find_reference()

    base = {"metric": METRIC, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bytes (ASCII text)",
            "data": "synthetic: fixture triton tables from the reference's walkthrough (docs/detailed.md:156-195), "
                    "stub tools, fixed answers", "gpu_launches": 0,
            "config": {"workload": WORKLOAD, "l2_policy": "n/a (no device buffers)",
                       "parallelism": "%d CPU process(es), no GPU; configs and catalogue scenarios dealt out round robin, "
                                      "no data-path collective (SURVEY.md §8e)" % ranks.world,
                       "value_checks": "parity, terraform, ansible, inventory, determinism -- all five measured in every leg"},
            "roofline": {"bound": "n/a", "achieved": None, "peak": None, "unit": None, "frac": None, "traffic": None,
                         "note": NOT_APPLICABLE}}

    if args.impl == "reference":
        if ref_root is None:
            print(json.dumps({"impl": "reference", "unavailable": "no reference tree: neither /root/reference nor the staged oracle/_ref/reference"}))
            return 0
        results, times = timed(lambda: step_reference(docs, ref_root), steps, warmup)
        ms = 1e3 * statistics.mean(times)
        rate = harness.pass_rate(results)
        line = dict(base, impl="reference", value=rate, ms_per_step=ms, configs_per_s=len(docs) / (ms / 1e3),
                    clocks=sample_clocks(), per_config=summarise(results),
                    cpu_baseline={"value": rate, "unit": UNIT, "cores": 1, "kind": "reference",
                                  "sample": "all 5 configs x %d steps, unmodified setup.sh under bash in the stub sandbox (%s)" % (steps, ref_root),
                                  "ms_per_step": ms},
                    e2e={"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
        print(json.dumps(line))
        ranks.close()
        return 0

    # -- our arm -------------------------------------------------------------
    n_configs = len(docs)
    mine = docs[ranks.rank::ranks.world]          # this rank's share of the five configs
    results: List[harness.ConfigResult] = []
    for _ in range(warmup):
        step_in_process(mine)
    ranks.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        results = step_in_process(mine)
    local = time.perf_counter() - t0
    ranks.barrier()
    total = ranks.reduce(local, "MAX")            # max over ranks
    rate = ranks.reduce(float(sum(1 for r in results if r.passed)), "SUM") / n_configs
    ms = 1e3 * total / steps

    # end to end through the CLI: two runs per config (the second is criterion iv), both timed
    e2e_results, e2e_times = timed(lambda: step_cli(mine, repeat=2), max(1, min(steps, 3)), 1)
    e2e_rate = ranks.reduce(float(sum(1 for r in e2e_results if r.passed)), "SUM") / n_configs
    e2e_ms = 1e3 * ranks.reduce(statistics.mean(e2e_times), "MAX")
    stdin_bytes = sum(len(d["scenario"]["stdin"]) for d in docs)
    out_bytes = sum(len(c) for d in docs for c in d["expected"]["files"].values()) + sum(len(d["expected"]["stdout"]) for d in docs)

    # the whole recorded parity surface, in-process, outside the timed region
    t0 = time.perf_counter()
    cat = harness.replay_catalogue([rp.load_golden(p) for p in rp.list_goldens()[ranks.rank::ranks.world]])
    cat_s = ranks.reduce(time.perf_counter() - t0, "MAX")
    catalogue = {"replayed": int(ranks.reduce(float(cat["replayed"]), "SUM")), "passed": int(ranks.reduce(float(cat["passed"]), "SUM")),
                 "skipped_not_reproducible_here": int(ranks.reduce(float(len(cat["skipped"])), "SUM")),
                 "failed_on_rank0": cat["failed"], "seconds": round(cat_s, 2),
                 "note": "every fixture under tests/golden replayed in-process against the in-memory tool models"}

    cpu_baseline: Dict[str, Any]
    if ranks.rank == 0 and args.gpus == 1 and ref_root is not None:
        ref_results, ref_times = timed(lambda: step_reference(docs, ref_root), 3, 1)
        ref_ms = 1e3 * statistics.mean(ref_times)
        from oracle.reference_runner import REFERENCE_ARGV
        cpu_baseline = {"value": harness.pass_rate(ref_results), "unit": UNIT, "cores": 1, "kind": "reference",
                        "sample": "all 5 configs x 3 steps (with checks) + %d bare runs per config, unmodified setup.sh "
                                  "under bash in the stub sandbox" % args.median_runs,
                        "ms_per_step": ref_ms, "configs_per_s": len(docs) / (ref_ms / 1e3),
                        "median_run_ms_per_config": median_wall_ms(REFERENCE_ARGV, None, ref_root, docs, args.median_runs, False)}
    else:
        cpu_baseline = {"value": None, "unit": UNIT, "cores": 1, "kind": "reference",
                        "sample": "not run (rank != 0, N > 1, or no reference tree on this machine)"}

    ours_median = (median_wall_ms(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, harness.WORKSPACE_TEMPLATE, docs,
                                  args.median_runs, True) if ranks.rank == 0 and args.gpus == 1 else None)
    latency_model = None
    if args.triton_latency_ms > 0 and ranks.rank == 0:
        latency_model = triton_latency_model(docs[1], args.triton_latency_ms, ref_root)
    extension = harness.extension_count_scaling(10) if ranks.rank == 0 else None
    if ranks.rank == 0:
        line = dict(base, value=rate, ms_per_step=ms, configs_per_s=len(docs) / (ms / 1e3),
                    clocks=sample_clocks(), per_config=summarise(results), per_config_e2e=summarise(e2e_results),
                    catalogue=catalogue,
                    e2e={"value": e2e_rate, "unit": UNIT, "ms_per_step": e2e_ms,
                         "configs_per_s": len(docs) / (e2e_ms / 1e3),
                         "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                         "stdin_bytes_per_step": stdin_bytes, "output_bytes_per_step": out_bytes,
                         "median_run_ms_per_config": ours_median,
                         "note": "public CLI as a child process, stub tools as child processes; 2 timed runs per config (the second is the determinism check); no device"},
                    cpu_baseline=cpu_baseline,
                    extension_config5b={k: extension[k] for k in ("ok", "n_workers", "modules", "planned_machines",
                                                                   "inventory_hosts", "rancher_tf_bytes", "size_law")})
        if latency_model is not None:
            line["triton_latency_model"] = latency_model
        print(json.dumps(line))
    ranks.close()
    return 0


def triton_latency_model(doc: dict, ms: int, ref_root: Optional[str], runs: int = 3) -> Dict[str, Any]:
    """A MODEL, not a measurement of CloudAPI: each stubbed `triton` call sleeps ``ms``.
    Shows what the per-index re-fetching of setup.sh:536,541 costs and what the
    opt-in table cache saves; files are byte-identical in all three arms."""
    sc = Scenario.from_json(doc["scenario"])

    def med(argv, env, src, overrides):
        walls, calls = [], 0
        for _ in range(runs):
            with Sandbox(src, sc, input_overrides=doc["inputs"] if overrides else None, latency_ms={"triton": ms}) as sbx:
                obs = sbx.run(argv, env)
                walls.append(obs.wall_s)
                calls = len([c for c in obs.calls if c.startswith("triton|")])
        return {"median_run_ms": round(1e3 * statistics.median(walls), 1), "triton_calls": calls}
    out: Dict[str, Any] = {"model": "every `triton` call takes %d ms" % ms, "config": sc.name,
                           "replacement": med(rp.REPLACEMENT_ARGV, rp.REPLACEMENT_ENV, harness.WORKSPACE_TEMPLATE, True),
                           "replacement_cached": med(rp.REPLACEMENT_ARGV, dict(rp.REPLACEMENT_ENV, TK8S_CACHE_TRITON="1"),
                                                     harness.WORKSPACE_TEMPLATE, True)}
    if ref_root is not None:
        from oracle.reference_runner import REFERENCE_ARGV
        out["reference"] = med(REFERENCE_ARGV, None, ref_root, False)
    return out


if __name__ == "__main__":
    sys.exit(main())
