#!/usr/bin/env python
"""bench.py -- CV fits/sec of the fitness-evaluation hot path on BASELINE.json's shapes.

A step = one batch of individuals (members of the seeded population of the chosen config) evaluated by ONE call of
the C ABI (gt_engine_eval).  `value` times that call with the binned folds already resident in HBM; `e2e` times the
call a gentun user makes -- Population(XgboostIndividual, x, y, ...).get_fittest() on a float64 pandas DataFrame --
from host arrays to the fittest individual: ingestion (H2D + cast), folds, quantile sketch, binning, training and
the fitness read-back, i.e. everything xgb.DMatrix + xgb.cv do per fit in the reference.

    python bench.py --gpus N --steps K --warmup W                    # our arm, headline config C3
    python bench.py --workload c1|c2|c4|c5 ...                        # the other named shapes of BASELINE.json
    python bench.py --impl reference --gpus N --steps K --warmup W    # CPU restatement of the reference path
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "cv_fits_per_sec"
UNIT = "fits/s"
WINE = os.path.join(ROOT, "tests", "data", "winequality-white.csv")
# nfold / rounds / early stopping are what the config passes through additional_parameters; `ips` is the
# number of individuals one GPU evaluates per step
WORKLOADS = {
    "c1": dict(data="wine", population=100, grid=False, nfold=3, rounds=5000, esr=100, ips=100, cls=False,
               text="C1 of BASELINE.json: XgboostIndividual regressor on winequality-white.csv (4898x11), Population size=100, "
                    "nfold=3, up to 5000 rounds with early stopping 100, one generation per step"),
    "c2": dict(data="wine", population=200, grid=True, nfold=3, rounds=5000, esr=100, ips=200, cls=False,
               text="C2 of BASELINE.json: GridPopulation eta x max_depth x colsample_bytree (5x8x5 = 200 individuals) on "
                    "winequality-white.csv, nfold=3, up to 5000 rounds with early stopping 100, the whole grid per step"),
    "c3": dict(data="synthetic", rows=1_000_000, features=50, seed=1003, population=256, nfold=5, rounds=100, esr=100, ips=32,
               cls=False, text="C3 of BASELINE.json: XgboostIndividual regressor, synthetic 1000000x50, Population size=256, "
                               "nfold=5, 100 boosting rounds"),
    "c4": dict(data="synthetic", rows=10_000_000, features=100, seed=1004, population=128, nfold=3, rounds=100, esr=100, ips=16,
               cls=True, text="C4 of BASELINE.json: binary:logistic / logloss, synthetic 10000000x100, Population size=128, "
                              "nfold=3, 100 boosting rounds, max_depth up to 10"),
    "c5": dict(data="synthetic", rows=4_000_000, features=200, seed=1005, population=1024, nfold=5, rounds=100, esr=100, ips=32,
               cls=False, text="C5 of BASELINE.json: regressor, synthetic 4000000x200, Population size=1024, nfold=5, "
                               "100 boosting rounds, population sharded over the GPUs with a fitness all-gather"),
    # quick checks only, never a bench line
    "c3-small": dict(data="synthetic", rows=100_000, features=50, seed=1003, population=256, nfold=5, rounds=20, esr=100, ips=32,
                     cls=False, text="c3-small (quick check, not a bench line)"),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--inds-per-step", type=int, default=0, help="individuals per step per GPU (default: the workload's)")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--cpu-sample-rounds", type=int, default=10, help="boosting rounds of the bounded cpu_baseline sample in our arm")
    ap.add_argument("--cpu-budget-s", type=float, default=900.0,
                    help="reference arm: full fits are timed until this much wall clock is used; later steps are sampled")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--rounds", type=int, default=0, help="override boosting rounds (profiling only; not a bench line)")
    return ap.parse_args()


def load_workload(name, rounds_override=0):
    """(x32, y, x_api, y_api, genes, params, spec): x32/y feed the C ABI directly, x_api/y_api are what a gentun script
    holds (a float64 DataFrame and a Series), genes is the seeded population."""
    import pandas as pd
    from tests.synth import make_dataset, seeded_population_genes
    spec = dict(WORKLOADS[name])
    if rounds_override:
        spec["rounds"] = rounds_override
        spec["text"] += " [rounds overridden to %d: profiling run, not a bench line]" % rounds_override
    if spec["data"] == "wine":
        df = pd.read_csv(WINE, delimiter=";")
        y_api, x_api = df["quality"], df.drop(["quality"], axis=1)     # tests/test_wine-quality.py:17-19
        x32 = np.ascontiguousarray(x_api.values, dtype=np.float32)
        y = np.ascontiguousarray(y_api.values, dtype=np.float32)
        spec["rows"], spec["features"] = x32.shape
        if spec["grid"]:
            genes = grid_genes()
        else:
            genes = seeded_population_genes(spec["population"])
    else:
        x32, y = make_dataset(spec["rows"], spec["features"], spec["seed"], classification=spec["cls"])
        x_api = pd.DataFrame(np.asfortranarray(x32, dtype=np.float64))   # column-major float64, like read_csv gives
        y_api = pd.Series(y.astype(np.float64))
        genes = seeded_population_genes(spec["population"])
    params = dict(nfold=spec["nfold"], num_boost_round=spec["rounds"], early_stopping_rounds=spec["esr"])
    if spec["cls"]:
        params.update(objective="binary:logistic", eval_metric="logloss")
    return x32, y, x_api, y_api, genes, params, spec


def grid_genes():
    """The 200 gene sets of tests/test_grid_wine-quality.py:20-30, built by GridPopulation itself."""
    import contextlib
    import io
    import random
    from gentun_b200 import GridPopulation, XgboostIndividual
    grid = {'eta': [0.001, 0.005, 0.01, 0.015, 0.2], 'max_depth': range(3, 11),
            'colsample_bytree': [0.80, 0.85, 0.90, 0.95, 1.0]}
    state = random.getstate()
    random.seed(20240601)
    with contextlib.redirect_stdout(io.StringIO()):
        pop = GridPopulation(XgboostIndividual, None, None, genes_grid=grid, additional_parameters={'nfold': 3}, maximize=False)
    random.setstate(state)
    return [dict(pop[i].get_genes()) for i in range(pop.get_size())]


def fits_l2(spec):
    return spec["nfold"] * spec["rows"] * ((spec["features"] + 31) // 32 * 32) <= (126 << 20)


def describe(spec, ips, world):
    n, f, nfold = spec["rows"], spec["features"], spec["nfold"]
    shards = max(1, spec["population"] // ips)
    return {
        "workload": spec["text"], "rows": n, "features": f, "nfold": nfold, "num_boost_round": spec["rounds"],
        "early_stopping_rounds": spec["esr"], "population": spec["population"], "individuals_per_step_per_gpu": ips,
        "step": "the population seeded with random.seed(20240601) is dealt once into %d cost-balanced shard(s) of %d individuals "
                "(gentun_b200.sharding.shard_plan); in step s rank r runs ONE gt_engine_eval on shard (s*n_gpus + r) mod %d, so "
                "per-GPU work is the same at every n_gpus" % (shards, ips, shards),
        "l2": ("inputs exceed L2: binned folds %d MB + per-job state" % (nfold * n * ((f + 31) // 32 * 32) >> 20)) if not fits_l2(spec)
              else "inputs fit L2 (%d KB of bins per fold): 256 MB written between timed steps to flush it" % (n * ((f + 31) // 32 * 32) >> 10),
        "parallelism": "population sharded over %d GPU(s), fitness all-gather" % world if spec["data"] != "wine"
                       else "%d replica(s) of the same generation (a 36 KB bin matrix gives 8 GPUs nothing to share out)" % world,
    }


class ClockSampler(object):
    """SM clock and throttle reasons during the timed region (the fields of the nvidia-smi line in
    B200_PROFILING.md), read through NVML from a thread every 200 ms; falls back to nvidia-smi."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device = device
        self.sm, self.mx, self.reasons = [], [], set()
        self.proc = None
        self.stop_flag = threading.Event()
        self.thread = None
        self.source = None

    def _nvml_loop(self, nv, handle):
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
        while not self.stop_flag.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(handle, nv.NVML_CLOCK_SM)))
                self.mx.append(float(nv.nvmlDeviceGetMaxClockInfo(handle, nv.NVML_CLOCK_SM)))
                mask = nv.nvmlDeviceGetCurrentClocksEventReasons(handle)
                for name, bit in bits.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def _smi_loop(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                self.sm.append(float(parts[0])); self.mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    self.reasons.add(nm)

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            index = int(visible.split(",")[self.device]) if visible and visible.split(",")[self.device].isdigit() else self.device
            handle = nv.nvmlDeviceGetHandleByIndex(index)
            nv.nvmlDeviceGetClockInfo(handle, nv.NVML_CLOCK_SM)
            self.source = "nvml"
            self.thread = threading.Thread(target=self._nvml_loop, args=(nv, handle), daemon=True)
            self.thread.start()
            return
        except Exception:
            pass
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.source = "nvidia-smi"
            self.thread = threading.Thread(target=self._smi_loop, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def stop(self):
        self.stop_flag.set()
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        if self.thread is not None:
            self.thread.join(timeout=2)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock sampling unavailable"], "samples": 0}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": max(self.mx), "reasons": sorted(self.reasons),
                "samples": len(self.sm), "source": self.source}


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(workload):
    """DRAM bytes per histogram launch from the committed ncu capture of this workload, with the algorithmic bytes of
    the SAME launches beside it (profiles/hist_traffic.json), or (None, None)."""
    path = os.path.join(ROOT, "profiles", "hist_traffic.json")
    if os.path.exists(path):
        try:
            d = json.load(open(path))
            if d.get("workload", "c3") == workload.split("-")[0]:
                return d.get("dram_bytes_per_launch"), d.get("algorithmic_bytes_per_launch")
        except Exception:
            pass
    return None, None


def cpu_fit(x, y, genes, params, nthread, rounds_sample=None):
    """The reference's per-fit CPU path with the oracle port: the full per-fit set-up (fold slices, quantile sketch,
    binning: what xgb.DMatrix + the lazy Booster init redo for every individual) plus the boosting rounds -- all of them,
    or `rounds_sample` of them extrapolated linearly.  Returns (seconds per fit, detail)."""
    from oracle import orc
    rounds_total = params["num_boost_round"]
    run = rounds_total if not rounds_sample else min(rounds_sample, rounds_total)
    opts = dict(params, num_boost_round=run, nthread=nthread)
    if run < rounds_total:
        opts["early_stopping_rounds"] = 0
    t0 = time.perf_counter()
    ctx = orc.Context(x, y, params["nfold"], nthread=nthread)
    t_setup = time.perf_counter() - t0
    t0 = time.perf_counter()
    res = ctx.cv(genes, orc.make_opts(**opts))
    t_rounds = time.perf_counter() - t0
    ctx.close()
    full = run == rounds_total
    sec = t_setup + (t_rounds if full else t_rounds * (rounds_total / float(run)))
    return sec, {"setup_s": t_setup, "rounds_s": t_rounds, "rounds_run": res["rounds"], "full": full, "fitness": res["fitness"]}


def run_reference(args):
    """The reference's own CPU path for the same config: one individual's complete fit (set-up + every boosting round,
    early stopping as configured) per step, on all host cores.  The thread count is passed explicitly so that a
    launcher's OMP_NUM_THREADS=1 cannot shrink it, and `cores` is what OpenMP really used."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import orc
    x, y, _, _, genes, params, spec = load_workload(args.workload, args.rounds)
    nthread = orc.host_cores()
    cores = orc.threads_in_use(nthread)
    ips = args.inds_per_step or spec["ips"]
    times, fulls, t_start = [], 0, time.perf_counter()
    n_steps = 1 + args.steps if args.warmup else args.steps      # one warm-up fit is enough for a CPU loop
    for s in range(n_steps):
        over = (time.perf_counter() - t_start) > args.cpu_budget_s
        sample = max(2, params["num_boost_round"] // 20) if over else None
        sec, detail = cpu_fit(x, y, genes[s % len(genes)], params, nthread, rounds_sample=sample)
        if args.warmup and s == 0:
            continue
        times.append(sec)
        fulls += 1 if detail["full"] else 0
    sec_per_fit = float(np.mean(times))
    value = 1.0 / sec_per_fit
    sample = ("oracle port of the reference's per-fit path (no xgboost on this box): per individual, full set-up (fold slices + "
              "quantile sketch + binning of %dx%d x %d folds) + ALL boosting rounds (early stopping as configured), %d threads; "
              "%d individuals of the seeded population, %d timed in full, %d sampled and extrapolated after the %.0f s budget"
              % (spec["rows"], spec["features"], spec["nfold"], cores, len(times), fulls, len(times) - fulls, args.cpu_budget_s))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec_per_fit * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64 accumulators over f32 gradients", "data": "synthetic" if spec["data"] == "synthetic" else "winequality-white.csv",
        "config": describe(spec, ips, 1),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "extrapolated_fraction": 1.0 - fulls / float(len(times))},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)
    return 0


_REAL_STDOUT = None


def quiet_stdout():
    """stdout must carry the one JSON line and nothing else, but libraries write there too (NCCL prints its version
    banner to stdout when the box sets NCCL_DEBUG): point fd 1 at stderr for the run and keep the real stdout aside."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (json.dumps(line) + "\n").encode())


def main():
    args = parse_args()
    quiet_stdout()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    import torch.distributed as dist
    import gentun_b200 as gentun
    from gentun_b200.engine import Engine, genes_matrix
    from gentun_b200.sharding import fit_cost, shard_plan

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the engine has no CPU path)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    x, y, x_api, y_api, genes, params, spec = load_workload(args.workload, args.rounds)
    pop = len(genes)
    ips = args.inds_per_step or spec["ips"]
    if pop % ips:
        raise SystemExit("--inds-per-step must divide the population size %d" % pop)
    shards = shard_plan([fit_cost(g) for g in genes], pop // ips)
    replicas = spec["data"] == "wine"     # a 36 KB bin matrix: every GPU evaluates the same generation

    def step_genes(s, r=None):
        """Rank r's share of step s: one of the population's cost-balanced shards (the same dealing
        gentun_b200.sharding does for a generation on pop // ips GPUs), so every rank always has one shard's work."""
        r = rank if r is None else r
        return [genes[i] for i in shards[(s * (1 if replicas else world) + (0 if replicas else r)) % len(shards)]]

    small = fits_l2(spec)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda") if small else None

    def flush_l2():
        if flush_buf is not None:
            flush_buf.fill_(1)

    # ---- resident arm: binned folds in HBM before the timed region
    eng = Engine(x, y, device=local_rank, collect_timing=True)
    eng.prepare(spec["nfold"])
    prep_ms = eng.stats()["prepare_ms"]
    gather_buf = torch.empty(world * ips, dtype=torch.float64, device="cuda") if world > 1 else None

    def one_step(s):
        flush_l2()
        fit, _ = eng.evaluate(step_genes(s), **params)
        if world > 1:   # the path's one collective: fitness scalars back to every rank
            dist.all_gather_into_tensor(gather_buf, torch.from_numpy(fit).cuda())
        return fit

    for s in range(args.warmup):
        one_step(s)
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    t0 = time.perf_counter()
    dev_ms = hist_ms = hist_bytes = hist_rf = 0.0
    launches = hist_launches = 0
    for s in range(args.warmup, args.warmup + args.steps):
        one_step(s)
        st = eng.stats()
        dev_ms += st["eval_ms"]; hist_ms += st["hist_ms"]; hist_bytes += st["hist_alg_bytes"]; hist_rf += st["hist_row_features"]
        launches += st["kernel_launches"]; hist_launches += st["hist_launches"]
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop()
    t = torch.tensor([dev_ms, wall_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms_max, wall_ms_max = float(t[0]), float(t[1])
    rank_ms = torch.zeros(world, dtype=torch.float64, device="cuda")
    rank_ms[rank] = dev_ms / args.steps
    if world > 1:
        dist.all_reduce(rank_ms)
    value = world * ips * args.steps / (wall_ms_max / 1e3)
    eng.close()
    del eng

    # ---- end-to-end arm: the call a gentun user makes.  A float64 DataFrame and a Series go into a Population, and
    # get_fittest() returns the fittest individual: ingestion (pinned-tile H2D + cast on the device), folds, sketch,
    # binning, training and the fitness read-back are all inside the timed region.  Under torchrun every rank holds the
    # same population of world * ips individuals and gentun_b200.sharding deals it over the ranks (fitness all-gather).
    e2e_ms = 0.0
    for s in range(args.e2e_steps + 1 if args.e2e_steps else 0):
        gs = []
        for r in range(1 if replicas else world):
            gs += step_genes(args.warmup + args.steps + s, r)
        flush_l2()
        barrier()
        t1 = time.perf_counter()
        gentun.clear_engines()          # a new dataset object every time a user script starts: nothing is cached
        inds = [gentun.XgboostIndividual(x_api, y_api, genes=dict(g), **params) for g in gs]
        best = gentun.Population(gentun.XgboostIndividual, x_api, y_api, individual_list=inds, maximize=False).get_fittest()
        assert np.isfinite(best.get_fitness())
        barrier()
        if s > 0:   # first pass is a warm-up
            e2e_ms += (time.perf_counter() - t1) * 1e3
    gentun.clear_engines()
    te = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    n_e2e = ips * (1 if replicas else world)   # wine: the one generation, dealt over the ranks by the product itself
    e2e_value = n_e2e * args.e2e_steps / (float(te[0]) / 1e3) if args.e2e_steps else None
    x_bytes = int(np.asarray(x_api).nbytes + np.asarray(y_api).nbytes)
    h2d = x_bytes + genes_matrix(step_genes(0)).nbytes
    d2h = ips * 12 + spec["nfold"] * spec["features"] * 4

    if rank == 0:
        peak, peak_src = measured_peak()
        achieved = hist_bytes / (hist_ms * 1e-3) / 1e9 if hist_ms > 0 else None
        traffic, traffic_alg = ncu_traffic(args.workload)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": wall_ms_max / args.steps, "device_ms_per_step": dev_ms_max / args.steps,
            "device_ms_per_step_by_rank": [round(v, 1) for v in rank_ms.tolist()],
            "s_per_generation": spec["population"] / value if value else None,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int32/int64 fixed-point histograms over f32 gradients (f64 split gains)",
            "data": "synthetic" if spec["data"] == "synthetic" else "winequality-white.csv",
            "config": describe(spec, ips, world),
            "roofline": {
                "bound": "hbm", "kernel": "k_hist", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None, "traffic": traffic,
                "traffic_algorithmic_bytes_same_launches": traffic_alg, "peak_source": peak_src,
                "frac_of_8tbs_spec": (achieved / 8000.0) if achieved else None,
                "launches": hist_launches, "avg_launch_ms": hist_ms / max(hist_launches, 1),
                "algorithmic_bytes_per_launch": hist_bytes / max(hist_launches, 1),
                "row_features_per_s": hist_rf / (hist_ms * 1e-3) if hist_ms > 0 else None,
                "share_of_step": hist_ms / dev_ms if dev_ms > 0 else None,
            },
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "includes": "Population(XgboostIndividual, float64 DataFrame, Series, individual_list).get_fittest(): ingestion "
                                "(pinned-tile H2D + cast), folds, sketch, binning, gt_engine_eval, fitness read-back, engine teardown"},
            "gpu_launches": int(launches), "prepare_ms": prep_ms, "clocks": clocks,
        }
        if not args.no_cpu_baseline and world == 1:
            from oracle import orc
            nthread = orc.host_cores()
            wine = spec["data"] == "wine"
            sec, detail = cpu_fit(x, y, genes[0], params, nthread, rounds_sample=None if wine else args.cpu_sample_rounds)
            line["cpu_baseline"] = {
                "value": 1.0 / sec, "unit": UNIT, "cores": orc.threads_in_use(nthread), "kind": "port",
                "sample": "oracle port (no xgboost on this box): individual 0 of the seeded population, full per-fit set-up (folds + "
                          "sketch + binning, %.1f s) + %d of %d rounds x %d folds (%.1f s)%s; the full-length measurement is "
                          "`bench.py --impl reference`" % (detail["setup_s"], detail["rounds_run"], params["num_boost_round"], spec["nfold"],
                                                          detail["rounds_s"], "" if detail["full"] else ", extrapolated linearly")}
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
