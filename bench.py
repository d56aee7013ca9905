#!/usr/bin/env python
"""bench.py -- CV fits/sec of the fitness-evaluation hot path on BASELINE.json's headline shape.

A step = one batch of individuals (consecutive members of the seeded 256-individual population of
config C3: synthetic 1M x 50 regressor, nfold=5, 100 boosting rounds) evaluated by ONE call of the
C ABI (gt_engine_eval).  `value` times that call with the binned folds already resident in HBM;
`e2e` times gt_engine_create (host X, y -> device, sketch, binning) + gt_engine_eval + fitness
read-back, i.e. everything xgb.DMatrix + xgb.cv do per fit in the reference.

    python bench.py --gpus N --steps K --warmup W            # our arm
    python bench.py --impl reference --gpus N --steps K ...   # CPU restatement of the reference path
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
WORKLOADS = {
    # name: (rows, features, nfold, rounds, population, classification, seed)
    "c3": (1_000_000, 50, 5, 100, 256, False, 1003),
    "c3-small": (100_000, 50, 5, 20, 256, False, 1003),   # for quick checks only, never a bench line
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--inds-per-step", type=int, default=32,
                    help="individuals per step per GPU (32 = one GPU's share of the 256-individual generation on 8 GPUs)")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--cpu-rounds", type=int, default=2, help="boosting rounds in the CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--rounds", type=int, default=0, help="override boosting rounds (profiling only; not a bench line)")
    return ap.parse_args()


def describe(wl, args, world):
    n, f, nfold, rounds, pop, cls, seed = WORKLOADS[wl]
    return {
        "workload": "C3 of BASELINE.json: XgboostIndividual regressor, synthetic %dx%d, Population size=%d, nfold=%d, "
                    "%d boosting rounds" % (n, f, pop, nfold, rounds) if wl == "c3" else wl,
        "rows": n, "features": f, "nfold": nfold, "num_boost_round": rounds, "population": pop,
        "individuals_per_step_per_gpu": args.inds_per_step,
        "step": "the population seeded with random.seed(20240601) is dealt once into %d cost-balanced shards of %d "
                "individuals (gentun_b200.sharding.shard_plan); in step s rank r runs ONE gt_engine_eval on shard "
                "(s*n_gpus + r) mod %d, so per-GPU work is the same at every n_gpus and 8 GPUs evaluate one whole "
                "generation per step" % (pop // args.inds_per_step, args.inds_per_step, pop // args.inds_per_step),
        "l2": "inputs exceed L2: binned folds %d MB + per-job state > 126 MB" % (nfold * n * ((f + 31) // 32 * 32) >> 20),
        "parallelism": "population sharded over %d GPU(s), fitness all-gather" % world,
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


def ncu_traffic():
    """dram bytes per histogram launch from the committed ncu capture, if one exists."""
    path = os.path.join(ROOT, "profiles", "hist_traffic.json")
    if os.path.exists(path):
        try:
            return json.load(open(path)).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


def cpu_sample(x, y, genes, nfold, rounds_total, rounds_sample, nthread=0):
    """Bounded sample of the reference's per-fit CPU path with the oracle port: the full per-fit set-up
    (fold slices, quantile sketch, binning: what xgb.DMatrix + the lazy Booster init redo for every
    individual) plus `rounds_sample` boosting rounds over all folds, extrapolated linearly to
    rounds_total.  Returns (seconds per fit, detail)."""
    from oracle import orc
    t0 = time.perf_counter()
    ctx = orc.Context(x, y, nfold, nthread=nthread)
    t_setup = time.perf_counter() - t0
    t0 = time.perf_counter()
    ctx.cv(genes, orc.make_opts(nfold=nfold, num_boost_round=rounds_sample, early_stopping_rounds=100, nthread=nthread))
    t_rounds = time.perf_counter() - t0
    ctx.close()
    return t_setup + t_rounds * (rounds_total / float(rounds_sample)), {"setup_s": t_setup, "rounds_s": t_rounds}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import orc
    from tests.synth import make_dataset, seeded_population_genes
    n, f, nfold, rounds, pop, cls, seed = WORKLOADS[args.workload]
    x, y = make_dataset(n, f, seed, classification=cls)
    genes = seeded_population_genes(pop)
    cores = orc.max_threads()
    times = []
    for s in range(args.warmup + args.steps):
        if s >= 1 and s < args.warmup:
            continue   # one warm-up sample is enough for a CPU loop; keeps the run within minutes
        sec, _ = cpu_sample(x, y, genes[s % pop], nfold, rounds, args.cpu_rounds)
        if s >= args.warmup:
            times.append(sec)
    sec_per_fit = float(np.mean(times))
    value = 1.0 / sec_per_fit
    sample = ("oracle port of the reference's per-fit path (no xgboost on this box): per individual, full set-up "
              "(fold slices + quantile sketch + binning of %dx%d x %d folds) + %d of %d boosting rounds x %d folds, "
              "extrapolated linearly to %d rounds; %d individuals of the seeded population"
              % (n, f, nfold, args.cpu_rounds, rounds, nfold, rounds, len(times)))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec_per_fit * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64 accumulators over f32 gradients", "data": "synthetic",
        "config": describe(args.workload, args, 1),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
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
    if args.rounds:
        wl = list(WORKLOADS[args.workload]); wl[3] = args.rounds
        WORKLOADS[args.workload + "-profile"] = tuple(wl)
        args.workload = args.workload + "-profile"
    if args.impl == "reference":
        return run_reference(args)
    import torch
    import torch.distributed as dist
    from gentun_b200.engine import Engine, genes_matrix
    from tests.synth import make_dataset, seeded_population_genes

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

    n, f, nfold, rounds, pop, cls, seed = WORKLOADS[args.workload]
    x, y = make_dataset(n, f, seed, classification=cls)
    genes = seeded_population_genes(pop)
    params = dict(nfold=nfold, num_boost_round=rounds, early_stopping_rounds=100)
    ips = args.inds_per_step

    from gentun_b200.sharding import fit_cost, shard_plan
    if pop % ips:
        raise SystemExit("--inds-per-step must divide the population size %d" % pop)
    shards = shard_plan([fit_cost(g) for g in genes], pop // ips)

    def step_genes(s):
        """This rank's share of step s: one of the population's cost-balanced shards (the same dealing
        gentun_b200.sharding does for a generation on pop // ips GPUs), so every rank always has one shard's work."""
        return [genes[i] for i in shards[(s * world + rank) % len(shards)]]

    # ---- resident arm: binned folds in HBM before the timed region
    eng = Engine(x, y, device=local_rank, collect_timing=True)
    eng.prepare(nfold)
    prep_ms = eng.stats()["prepare_ms"]
    gather_buf = torch.empty(world * ips, dtype=torch.float64, device="cuda") if world > 1 else None

    def one_step(s):
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

    # ---- end-to-end arm: host buffers in, fitness out, through the C ABI; includes H2D of X and y,
    # fold construction, quantile sketch and binning (what the reference redoes per fit)
    eng.close()
    # the end-to-end arm reads its inputs from pinned host memory, as a caller that cares about H2D speed would
    xp = torch.from_numpy(x).pin_memory().numpy()
    yp = torch.from_numpy(y).pin_memory().numpy()
    e2e_ms = 0.0
    gm_bytes = 0
    for s in range(args.e2e_steps + 1 if args.e2e_steps else 0):
        barrier()
        t1 = time.perf_counter()
        e2 = Engine(xp, yp, device=local_rank)
        gs = step_genes(args.warmup + args.steps + s)
        fit, _ = e2.evaluate(gs, **params)
        if world > 1:
            dist.all_gather_into_tensor(gather_buf, torch.from_numpy(fit).cuda())
        e2.close()
        barrier()
        if s > 0:   # first pass is a warm-up
            e2e_ms += (time.perf_counter() - t1) * 1e3
        gm_bytes = genes_matrix(gs).nbytes
    te = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * ips * args.e2e_steps / (float(te[0]) / 1e3) if args.e2e_steps else None
    h2d = x.nbytes + y.nbytes + gm_bytes
    d2h = ips * 12 + nfold * f * 4 + y.nbytes * 0

    if rank == 0:
        peak, peak_src = measured_peak()
        achieved = hist_bytes / (hist_ms * 1e-3) / 1e9 if hist_ms > 0 else None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": wall_ms_max / args.steps, "device_ms_per_step": dev_ms_max / args.steps,
            "device_ms_per_step_by_rank": [round(v, 1) for v in rank_ms.tolist()],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int64 fixed-point histograms over f32 gradients (f64 split gains)", "data": "synthetic",
            "config": describe(args.workload, args, world),
            "roofline": {
                "bound": "hbm", "kernel": "k_hist", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None, "traffic": ncu_traffic(), "peak_source": peak_src,
                "frac_of_8tbs_spec": (achieved / 8000.0) if achieved else None,
                "launches": hist_launches, "avg_launch_ms": hist_ms / max(hist_launches, 1),
                "algorithmic_bytes_per_launch": hist_bytes / max(hist_launches, 1),
                "row_features_per_s": hist_rf / (hist_ms * 1e-3) if hist_ms > 0 else None,
                "share_of_step": hist_ms / dev_ms if dev_ms > 0 else None,
            },
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "includes": "gt_engine_create (H2D of X,y + folds + sketch + binning) + gt_engine_eval + destroy"},
            "gpu_launches": int(launches), "prepare_ms": prep_ms, "clocks": clocks,
        }
        if not args.no_cpu_baseline and world == 1:
            from oracle import orc
            sec, detail = cpu_sample(x, y, genes[0], nfold, rounds, args.cpu_rounds)
            line["cpu_baseline"] = {
                "value": 1.0 / sec, "unit": UNIT, "cores": orc.max_threads(), "kind": "port",
                "sample": "oracle port (no xgboost on this box): individual 0 of the seeded population, full per-fit "
                          "set-up (folds + sketch + binning, %.1f s) + %d of %d rounds x %d folds (%.1f s), extrapolated "
                          "linearly to %d rounds" % (detail["setup_s"], args.cpu_rounds, rounds, nfold, detail["rounds_s"], rounds)}
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
