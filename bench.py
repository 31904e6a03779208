#!/usr/bin/env python
"""Benchmark of the UASE embed-and-rank hot path (BASELINE.json metric) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c3|c2]

A step = one pass of the hot path over one synthetic multi-layer SBM: UASE solve at k = 24
(streaming tcgen05 products + CholeskyQR2 + Jacobi), distances for every swept d (4..24, euclidean +
cosine), per-column rank and Borda.  `value` = wall time per step with the preprocessed layers
resident in HBM (CUDA events, max over ranks); `e2e` = the same through the public N2V2R API with
the transformed graphs in pinned HOST memory (H2D of the graphs, plane build, solve, rank, D2H of
the tables inside the timed region).  Layers are sharded over the ranks (strong scaling: the
workload is fixed); the only data-path collective is the all-reduce of Y (n x l) per operator
application plus one all-gather of the embeddings.

`--impl reference` times the reference's own CPU path (oracle/reference_port.py: scipy svds/ARPACK on
CSC layers, scipy.spatial distances, pandas sort + joblib Borda -- SURVEY 8c) on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np

DIMS = list(range(4, 25, 2))
METRICS = ["euclidean", "cosine"]
WORKLOADS = {
    # BASELINE.json configs[2]: the configuration the metric is quoted on (fits one B200: 80 GB of planes)
    "c3": dict(n=50000, T=8, K=32, weighted=True, keep=10, binarize=False, strategy="one_vs_rest",
               name="synthetic 8-layer weighted SBM n=50k, top_percent_keep=10, one_vs_rest, d=4..24, euclidean+cosine, Borda"),
    # the same with 24-bit weights fp32(u') + 2^-24: three bf16 plane terms are needed for exactness
    "c3w24": dict(n=50000, T=8, K=32, weighted=2, keep=10, binarize=False, strategy="one_vs_rest",
                  name="synthetic 8-layer SBM with 24-bit weights n=50k, top_percent_keep=10, one_vs_rest, d=4..24, euclidean+cosine, Borda"),
    # BASELINE.json configs[1]
    "c2": dict(n=20000, T=2, K=32, weighted=False, keep=100, binarize=True, strategy="sequential",
               name="synthetic 2-layer dense SBM n=20k, binarize, sequential, d=4..24, euclidean+cosine, Borda"),
    # BASELINE.json configs[3]: bipartite layers -> W^T W on the tensor cores (K2) -> UASE; the projection is inside the step
    "c4": dict(n=30000, rows=2000, T=2, K=32, weighted=True, keep=100, binarize=False, strategy="sequential", bipartite=True,
               name="bipartite PANDA-shaped 2-layer graph (2k x 30k), project_unipartite_on=columns (W^T W, tcgen05), UASE d=4..24, "
                    "euclidean+cosine, Borda"),
    # BASELINE.json configs[4]: stability run, seeds sharded over the ranks, layers replicated
    "c5": dict(n=30000, T=4, K=32, weighted=False, keep=100, binarize=True, strategy="sequential", seeds=list(range(16)),
               name="stability run: 16 seeds x 4-layer SBM n=30k, binarize, sequential, d=4..24, euclidean+cosine, Borda across runs"),
}
CPU_SAMPLE_N = 3000
GRAPH_SEED = 1
SOLVER_SEED = 42


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--host-transfer", default="auto", choices=["auto", "dense", "packed"],
                    help="e2e leg: how the host float32 graphs cross PCIe (device.layers_from_host)")
    ap.add_argument("--n", type=int, default=0, help="override the node count (development only; named in config.workload)")
    ap.add_argument("--block-width", type=int, default=0,
                    help="solver block width l (default 0 = device.default_block_width(24, n, storage): 32 on row-list layers, 48 otherwise)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer end-to-end leg (development only)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (development only)")
    ap.add_argument("--layer-storage", default="auto", choices=["auto", "planes", "tiles", "e4m3", "rows"],
                    help="how preprocessed layers are held in HBM (device.make_layer): dense bf16 planes, or compressed tiles "
                         "expanded in shared memory for sparse layers; auto = the product default")
    ap.add_argument("--no-cross-check", action="store_true",
                    help="skip the untimed second solve (other seed, l = 80) whose subspace angle against the timed solve is reported")
    ap.add_argument("--pack-dense-every", type=int, default=-1,
                    help="e2e leg (development): every N-th row chunk of a pinned host graph crosses PCIe dense while the host packs the "
                         "others (device.PACK_DENSE_EVERY); 0 = pack every chunk; -1 = the product default")
    ap.add_argument("--seed-batch", type=int, default=6,
                    help="workload c5: solves that run concurrently and share every pass over a layer (1 = one after the other)")
    ap.add_argument("--cpu-ladder", default="2000,4000,8000",
                    help="--impl reference: node counts the reference path is timed at (all layers), ascending")
    ap.add_argument("--cpu-budget-s", type=float, default=330.0, help="--impl reference: wall-clock budget of the ladder")
    return ap.parse_args()


# ---- clocks ------------------------------------------------------------------------------------
class SmiClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.path = tempfile.mktemp(prefix="n2v2r_clocks_", suffix=".csv")
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-lms", os.environ.get("N2V2R_CLOCK_MS", "200")],      # the profiling recipe's interval
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def mark(self):
        """Start of the timed region: samples before this line are warm-up (nvidia-smi's own start-up, which enumerates
        every GPU of the box and can stall launches for ~0.1 s, is kept out of the timed steps by starting it early)."""
        try:
            self.skip = sum(1 for _ in open(self.path))
        except OSError:
            self.skip = 0

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in list(open(self.path))[getattr(self, "skip", 0):]:
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2]))
                except ValueError:
                    continue
                for nm, v in zip(names, f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            os.unlink(self.path)
        except OSError:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


class ClockSampler:
    """clocks.sm / clocks.max.sm / clocks_event_reasons of one GPU every 200 ms during the timed region (the profiling recipe's
    nvidia-smi line, read through NVML in-process: a polling nvidia-smi process re-enumerates the box and was seen to stall
    kernel launches for 0.1-0.3 s inside single timed steps).  Falls back to the nvidia-smi process when NVML is unavailable."""
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.samples, self.skip, self.thread, self.fallback = [], 0, None, None
        self.stop_flag = threading.Event()
        self.period = float(os.environ.get("N2V2R_CLOCK_MS", "200")) / 1e3

    def _handle(self, nv):
        try:
            import torch
            pr = torch.cuda.get_device_properties(self.gpu)
            bus = f"{pr.pci_domain_id:08X}:{pr.pci_bus_id:02X}:{pr.pci_device_id:02X}.0"
            return nv.nvmlDeviceGetHandleByPciBusId(bus.encode())
        except Exception:
            return nv.nvmlDeviceGetHandleByIndex(self.gpu)

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = self._handle(nv)
            mx = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
        except Exception:
            self.fallback = SmiClockSampler(self.gpu)
            self.fallback.start()
            return

        def loop():
            while not self.stop_flag.is_set():
                try:
                    self.samples.append((float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), mx,
                                         int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))))
                except Exception:
                    pass
                self.stop_flag.wait(self.period)

        self.thread = threading.Thread(target=loop, daemon=True)
        self.thread.start()

    def mark(self):
        if self.fallback is not None:
            return self.fallback.mark()
        self.skip = len(self.samples)

    def stop(self):
        if self.fallback is not None:
            return self.fallback.stop()
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.thread is None:
            return out
        self.stop_flag.set()
        self.thread.join(timeout=2)
        got = self.samples[self.skip:] or self.samples
        if got:
            bits = 0
            for _, _, r in got:
                bits |= r
            out.update(sm_mhz=float(np.median([g[0] for g in got])), sm_max_mhz=float(max(g[1] for g in got)),
                       reasons=sorted(nm for nm, bit in self.REASONS if bits & bit), samples=len(got), source="nvml")
        return out


# ---- workload helpers --------------------------------------------------------------------------
def workload_of(args):
    w = dict(WORKLOADS[args.workload])
    if args.n:
        w["n"] = args.n
        w["name"] += f" [development override n={args.n}]"
    return w


def block_probabilities(w):
    from node2vec2rank_b200.synthetic import block_matrices
    return block_matrices(w["K"], w["T"], seed=GRAPH_SEED)[0]


# ---- CPU legs: the reference's own algorithm through the oracle port -----------------------------
def cpu_sample_graphs(w, n):
    """Transformed fp32 layers of the sample, made like the reference does (preprocessing_utils.py:116-141)."""
    from node2vec2rank_b200.synthetic import sbm_layer_numpy
    from oracle import reference_port as port
    Bs = block_probabilities(w)
    out = []
    for t in range(w["T"]):
        raw = sbm_layer_numpy(n, w["K"], Bs[t], GRAPH_SEED, t, w["weighted"])
        out.append(port.network_transform(raw.astype(np.float64), top_percent_keep=w["keep"], binarize=w["binarize"]))
    return out


def cpu_step(graphs, w):
    """model.py:98-201 on transformed graphs: csc_matrix + UASE(svds/ARPACK) + distances + sort + joblib Borda."""
    from oracle import reference_port as port
    tm = {}
    np.random.seed(SOLVER_SEED)
    port.fit_rank_borda(graphs, DIMS, METRICS, w["strategy"], faithful=True, timings=tm)
    return tm


def cpu_cores_used():
    n = os.cpu_count() or 1
    return max(1, n - 1)        # joblib n_jobs=-2 in Borda; ARPACK / csc matvec / scipy distances are single-threaded


# How each phase of the reference path grows with the node count n (T, dims, metrics fixed), used to extrapolate the largest
# MEASURED size to the full one.  Measured on the ladder n = 2000 / 4000 / 8000 (8 cores, profiles/README.md): the fit (ARPACK
# on CSC layers) grows only ~n^0.9 there although a pass over the layers costs n^2 -- the planted spectrum separates from the
# noise bulk as n grows and ARPACK needs fewer restarts; the distances are n Python calls per column (linear; small sizes are
# overhead-bound, measured 0.6-0.7); the faithful Borda is list.index per node per column in joblib workers (quadratic;
# measured 1.8 between the two largest rungs, less below where process start-up dominates).  The extrapolation takes the
# LOCAL slope of the two largest measured sizes, clamped to [floor, cap]: never below linear (every phase touches every node
# at least once per column / per pass), never above the phase's asymptotic law.  PHASE_EXPONENT is used when only one size
# was measured (the in-run cpu_baseline of the B200 arm).
PHASE_EXPONENT = {"fit": 1.0, "distances": 1.0, "borda": 1.8}
PHASE_FLOOR = {"fit": 1.0, "distances": 1.0, "borda": 1.0}
PHASE_CAP = {"fit": 2.4, "distances": 1.0, "borda": 2.0}


def fit_exponents(points):
    """Per phase: slope of log(seconds) against log(n) between the two largest measured sizes, clamped to [floor, cap]."""
    out = dict(PHASE_EXPONENT)
    if len(points) >= 2:
        a, b = points[-2], points[-1]
        for ph in PHASE_EXPONENT:
            slope = float(np.log(max(b[ph], 1e-6) / max(a[ph], 1e-6)) / np.log(b["n"] / a["n"]))
            out[ph] = float(min(PHASE_CAP[ph], max(PHASE_FLOOR[ph], slope)))
    return out


def raw_slopes(points):
    """The unclamped slopes (reported next to the clamped exponents)."""
    if len(points) < 2:
        return {}
    a, b = points[-2], points[-1]
    return {ph: float(np.log(max(b[ph], 1e-6) / max(a[ph], 1e-6)) / np.log(b["n"] / a["n"])) for ph in PHASE_EXPONENT}


def extrapolate(point, n_full, exps):
    """Per-phase extrapolation of one measured point to n_full nodes."""
    f = n_full / point["n"]
    parts = {ph: point[ph] * f ** exps[ph] for ph in PHASE_EXPONENT}
    return sum(parts.values()), parts


def cpu_ladder_point(w, n):
    """One measured point: the reference path (csc + svds/ARPACK, scipy distances, pandas sort + joblib Borda) at n nodes,
    all layers of the workload, the workload's transform and comparison strategy."""
    t0 = time.perf_counter()
    graphs = cpu_sample_graphs(w, n)
    t_gen = time.perf_counter() - t0
    tm = cpu_step(graphs, w)
    return {"n": int(n), "fit": tm["fit"], "distances": tm["distances"], "borda": tm["borda"], "total": tm["total"],
            "generate_and_transform_s": t_gen}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    w = workload_of(args)
    if w.get("bipartite"):
        for _ in range(min(args.warmup, 1)):
            cpu_c4_leg(w)
        legs = [cpu_c4_leg(w) for _ in range(max(min(args.steps, 3), 1))]
        full, dt, sample = float(np.mean([x["value"] for x in legs])), float(np.mean([x["sample_seconds"] for x in legs])), legs[-1]["sample"]
        cores = legs[-1]["cores"]
        extra = {}
    else:
        # The budget goes into SIZE, not repeats: every rung of the ladder is the whole reference path once, at a larger n.
        # --steps / --warmup are echoed; the path has no warm-up effect beyond imports (done by the first rung).
        runs = len(w.get("seeds", [0]))
        sizes = sorted({min(int(x), w["n"]) for x in args.cpu_ladder.split(",") if x.strip()})
        t_begin = time.perf_counter()
        points = []
        for n_s in sizes:
            if points:                       # predicted cost of this rung from the last one (generation ~ n^2, path ~ fitted law)
                last = points[-1]
                est = extrapolate(last, n_s, fit_exponents(points))[0] + last["generate_and_transform_s"] * (n_s / last["n"]) ** 2
                if time.perf_counter() - t_begin + est > args.cpu_budget_s:
                    break
            points.append(cpu_ladder_point(w, n_s))
        exps = fit_exponents(points)
        top = points[-1]
        one_run, parts = extrapolate(top, w["n"], exps)
        full = one_run * runs
        dt = top["total"]
        cores = cpu_cores_used()
        sample = (f"MEASURED: the reference path once per size, all {w['T']} layers, sizes n = {[p['n'] for p in points]}: totals "
                  f"{[round(p['total'], 2) for p in points]} s (largest: fit {top['fit']:.2f}, distances {top['distances']:.2f}, borda "
                  f"{top['borda']:.2f}).  EXTRAPOLATED to n = {w['n']}: each phase of the largest measured size scaled by "
                  f"(n_full/n)^e, e = the slope between the two largest measured sizes clamped to [1, the phase's asymptotic law] "
                  f"(fit {exps['fit']:.2f}, distances {exps['distances']:.2f}, borda {exps['borda']:.2f}) -> fit {parts['fit']:.0f} s, distances {parts['distances']:.0f} s, borda {parts['borda']:.0f} s"
                  + (f", x {runs} runs" if runs > 1 else "") + ".  value is that extrapolation, not a measurement at full size "
                  f"(the full size needs > 150 GB of host RAM and hours).")
        extra = {"measured": points, "measured_n": top["n"], "measured_s": top["total"], "exponents": exps,
                 "measured_slopes_last_two": raw_slopes(points),
                 "extrapolated_s": full, "extrapolated_phases_s": parts,
                 "cores_by_phase": {"fit": 1, "distances": 1, "borda": cores},
                 "ladder_wall_s": time.perf_counter() - t_begin}
    line = {
        "impl": "reference", "metric": "UASE sweep+rank wall time", "value": full, "unit": "s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": full * 1e3, "higher_is_better": False, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": w["name"], "n": w["n"], "layers": w["T"], "communities": w["K"]},
        "cpu_baseline": dict({"value": full, "unit": "s", "cores": cores, "kind": "port", "sample": sample,
                              "sample_seconds": dt}, **extra),
        "e2e": {"value": full, "unit": "s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


DTYPE_BY_STORAGE = {
    "planes": "bf16 planes x3-split bf16 operand, fp32 accumulate (tcgen05), fp64 small dense",
    "tiles": "bf16 planes x3-split bf16 operand, fp32 accumulate (tcgen05), fp64 small dense",
    "e4m3": "e4m3 plane x 6-term scaled e4m3 operand, fp32 accumulate (tcgen05 kind::f8f6f4), fp64 small dense",
    "rows": "f32 (exact fp32 layer values x fp32 operand, fp32 FMA accumulate on the CUDA cores), fp64 small dense",
}


# ---- B200 arm ------------------------------------------------------------------------------------
class Env:
    """One process per GPU: device, rendezvous, barrier and max-over-ranks (torch.distributed over NCCL)."""

    def __init__(self):
        import torch
        import torch.distributed as tdist
        from node2vec2rank_b200 import _native as N
        self.torch, self.tdist = torch, tdist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local_rank)
        self.dev = N.require_gpu()
        if self.world > 1:
            if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":      # keep stdout to the one JSON line
                os.environ["NCCL_DEBUG"] = "WARN"
            tdist.init_process_group("nccl", device_id=self.dev)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except (OSError, ValueError):
            pass
        self.hbm_peak, self.hbm_src = (float(peaks["hbm_gbs"]), "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
        key = "bf16_tflops_sustained" if "bf16_tflops_sustained" in peaks else "bf16_tflops"
        self.tc_peak, self.tc_src = (float(peaks[key]), f"measured ({key})") if key in peaks else (1394.0, "fallback")
        self.sm_count = torch.cuda.get_device_properties(self.dev).multi_processor_count
        self.sm_max_mhz = 1965.0           # clocks.max.sm of this pool's B200s (B200_PROFILING.md); the shared-memory peak is per clock

    def barrier(self):
        if self.world > 1:
            self.tdist.barrier()
        self.torch.cuda.synchronize()

    def reduce_max(self, x):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.tdist.all_reduce(t, op=self.tdist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, step, steps, warmup):
        """W untimed steps, then exactly K steps between barrier + synchronize, CUDA events, max over ranks.
        Returns (ms per step, list of step results, clocks dict on rank 0)."""
        torch = self.torch
        sampler = ClockSampler(self.local_rank)
        if self.rank == 0:
            sampler.start()
        for _ in range(warmup):
            step()
        self.barrier()
        if self.rank == 0:
            sampler.mark()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        results = [step() for _ in range(steps)]
        e1.record()
        self.barrier()
        ms = self.reduce_max(e0.elapsed_time(e1) / max(steps, 1))
        return ms, results, (sampler.stop() if self.rank == 0 else None)

    def timed_host(self, step, steps, warmup):
        """The end-to-end leg: wall clock around K calls of the public API (host buffers in, host tables out)."""
        for _ in range(warmup):
            step()
        self.barrier()
        t0 = time.perf_counter()
        out = None
        for _ in range(steps):
            out = step()
        self.barrier()
        return self.reduce_max((time.perf_counter() - t0) / max(steps, 1)), out

    def finish(self):
        if self.world > 1:
            self.tdist.destroy_process_group()


def allocator_counters():
    """torch caching-allocator counters (device mallocs / frees / free-and-retry rounds so far): their change over a timed
    region shows whether a step went to the driver for memory -- cudaMalloc and cudaFree of multi-GB blocks stall the host."""
    import torch
    st = torch.cuda.memory_stats()
    return {k: int(st.get(k, 0)) for k in ("num_device_alloc", "num_device_free", "num_alloc_retries")}


def allocator_delta(before):
    after = allocator_counters()
    return {k.replace("num_", ""): after[k] - before[k] for k in after}


def transfer_note(x):
    """How the host graphs crossed PCIe in the e2e leg (device.layers_from_host.last_stats)."""
    if not x or not x.get("packed_chunks"):
        return "dense float32 copies"
    return (f"packed on the host ({x['packed_chunks']} row chunks as bitmap + non-zero values, {x['dense_chunks']} dense"
            + (" on a second copy stream while the host cores pack" if x.get("hybrid") else "") + "; "
            f"{x.get('threads', 0)} host threads, {x.get('pack_s', 0):.3f} s packing, producer thread done after {x.get('producer_s', 0):.3f} s, "
            f"{x.get('total_s', 0):.3f} s transfer + load pass), "
            f"expanded on the device")


def cpu_baseline_leg(w, n, T):
    """The bounded CPU sample inside the B200 run (rank 0, N = 1): one measured point at n = CPU_SAMPLE_N, each phase
    extrapolated by its own law (PHASE_EXPONENT; the reference arm's ladder measures the exponents)."""
    n_s = min(CPU_SAMPLE_N, n)
    pt = cpu_ladder_point(w, n_s)
    runs = len(w.get("seeds", [0]))
    one_run, parts = extrapolate(pt, n, PHASE_EXPONENT)
    return {"value": one_run * runs, "unit": "s", "cores": cpu_cores_used(), "kind": "port",
            "cores_by_phase": {"fit": 1, "distances": 1, "borda": cpu_cores_used()},
            "sample": f"MEASURED at n={n_s} of {n} nodes, all {T} layers, one run of {runs}: {pt['total']:.2f} s (fit {pt['fit']:.2f}, distances "
                      f"{pt['distances']:.2f}, borda {pt['borda']:.2f}).  EXTRAPOLATED per phase with the exponents the reference arm's ladder "
                      f"gives (`bench.py --impl reference`, n up to 8000): fit x (n/n_s)^{PHASE_EXPONENT['fit']:g}, distances x "
                      f"(n/n_s)^{PHASE_EXPONENT['distances']:g}, borda x (n/n_s)^{PHASE_EXPONENT['borda']:g} -> fit {parts['fit']:.0f} s, "
                      f"distances {parts['distances']:.0f} s, borda {parts['borda']:.0f} s" + (f", x {runs} runs" if runs > 1 else ""),
            "sample_seconds": pt["total"], "measured_n": n_s, "measured_s": pt["total"],
            "extrapolated_s": one_run * runs, "extrapolated_phases_s": parts}


def build_workload_layers(w, n, layer_ids, want_f32=False, storage="auto", graph_seed=None, block_width=None):
    """Set-up of one workload (untimed in `value`): K0 generate + K1 load pass per layer, exactly what the timed step
    works on.  Returns (layers, fp32 device matrices or None per layer, [(event, event)] around every load pass).
    tests/test_gpu_bench_shapes.py runs the benchmark shapes in miniature through this function and hot_path()."""
    import torch
    from node2vec2rank_b200 import device as D
    Bs = block_probabilities(w)
    seed = GRAPH_SEED if graph_seed is None else graph_seed
    layers, f32s, events = [], [], []
    for t in layer_ids:
        raw = D.sbm_generate(n, w["K"], Bs[t], seed=seed, layer=t, weighted=w["weighted"])
        pa, pb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        pa.record()
        layer, f32 = D.preprocess_to_layer(raw, top_percent_keep=w["keep"], binarize=w["binarize"], want_f32=want_f32, storage=storage,
                                           block_width=block_width or D.default_block_width(max(DIMS), n, "rows"))
        pb.record()
        events.append((pa, pb))
        del raw
        layers.append(layer)
        f32s.append(f32)
    return layers, f32s, events


def hot_path(layers, w, l, seed=SOLVER_SEED, T_global=None, layer_offset=0, gather=None, dims=None, metrics=None, shard=None,
             slab_group=None):
    """One step of the benchmark: UASE solve at k = max(dims) with the gap-aware stop at every swept d, distances for every
    (d, metric), rank + Borda of all comparisons.  Returns (info, emb (T, n, k), scores (C, cols, n), bordas (C, n)).
    shard = (rank, world, all_reduce): the (d, metric) columns of the sweep are split over the ranks (every rank holds all
    embeddings after the gather), each rank scores and ranks its columns, and the Borda points are summed with one
    all-reduce; `scores` then holds this rank's columns only."""
    from node2vec2rank_b200 import _native as N
    from node2vec2rank_b200 import device as D
    from node2vec2rank_b200 import dist as ND
    dims = DIMS if dims is None else dims
    names, cdims, cmets = D.column_specs(dims, METRICS if metrics is None else metrics)
    emb, sigma, _, info = D.uase(layers, max(dims), l=l, tol=1e-5, max_iters=400, seed=seed, check_dims=dims,
                                 T_global=T_global or len(layers), layer_offset=layer_offset, slab_group=slab_group)
    info["sigma"] = sigma
    if gather is not None:
        emb = gather(emb)
    if shard is not None and shard[1] > 1:
        mine = ND.layer_partition(len(cdims), shard[1], shard[0])
        cdims, cmets = cdims[mine.start:mine.stop], cmets[mine.start:mine.stop]
    scores = D.distances(emb, cdims, cmets, N.STRATEGIES[w["strategy"]])
    if scores.shape[1]:
        bordas = D.rank_borda(scores)[0]                   # all comparisons in one launch
    else:
        bordas = scores.new_zeros((scores.shape[0], scores.shape[2]), dtype=__import__("torch").int64)
    if shard is not None and shard[1] > 1:
        shard[2](bordas)                                   # Borda points are additive over the columns
    return info, emb, scores, bordas


def cross_solve_check(layers, w, l, emb, bordas, T_global, layer_offset, gather, shard=None, slab_group=None):
    """UNTIMED parity evidence at full size: a second, independent solve (other start block, wider block l = 80) and the
    principal angle between the two embeddings' subspaces at every swept d (device.subspace_sin_angles), plus the agreement
    of the two Borda tables.  Two solves that share nothing but the layers agree to 1e-4 only if both are converged."""
    from node2vec2rank_b200 import device as D
    l2 = 80 if l != 80 else 64
    info2, emb2, _, bordas2 = hot_path(layers, w, l2, seed=SOLVER_SEED + 1000003, T_global=T_global, layer_offset=layer_offset,
                                       gather=gather, shard=shard, slab_group=slab_group)
    ang = D.subspace_sin_angles(emb, emb2, DIMS)
    a, b = bordas.double(), bordas2.double()
    a = a - a.mean(dim=1, keepdim=True)
    b = b - b.mean(dim=1, keepdim=True)
    corr = ((a * b).sum(dim=1) / (a.norm(dim=1) * b.norm(dim=1))).min().item()
    return {"other_seed": SOLVER_SEED + 1000003, "other_block_width": l2, "other_operator_applications": info2["iterations"],
            "sin_angle_per_d": {str(d): float(f"{v:.3e}") for d, v in ang.items()}, "max_sin_angle": max(ang.values()),
            "borda_pearson_min_over_comparisons": corr}


def run_b200_arm(args):
    w = workload_of(args)
    if args.pack_dense_every >= 0:
        from node2vec2rank_b200 import device as _D
        _D.PACK_DENSE_EVERY = args.pack_dense_every
    if w.get("bipartite"):
        return run_b200_c4(args, w)
    if w.get("seeds"):
        return run_b200_c5(args, w)
    import torch
    import torch.distributed as tdist
    from node2vec2rank_b200 import _native as N
    from node2vec2rank_b200 import device as D
    from node2vec2rank_b200 import dist as ND
    from node2vec2rank_b200.model import N2V2R

    env = Env()
    world, rank, local_rank, dev = env.world, env.rank, env.local_rank, env.dev
    barrier, reduce_max = env.barrier, env.reduce_max
    hbm_peak, peak_src = env.hbm_peak, env.hbm_src
    n, T, K, l = w["n"], w["T"], w["K"], args.block_width
    slab_group, slab_rows = None, None
    if world > T:
        # fewer layers than GPUs: every layer is cut into world / T row slabs; rank r holds slab r % S of layer r // S
        if world % T:
            raise SystemExit(f"workload {args.workload} has {T} layers: the number of ranks must be a multiple of {T} (row slabs per layer)")
        S = world // T
        mine = range(rank // S, rank // S + 1)
        slab_rows = ND.layer_partition(n, S, rank % S)
        for t in range(T):                         # every rank creates every group (torch.distributed rule)
            g = tdist.new_group(list(range(t * S, (t + 1) * S)))
            if t == rank // S:
                slab_group = g
    else:
        mine = ND.layer_partition(T, world, rank)
    Bs = block_probabilities(w)

    # ---- setup (untimed): K0 generate + K1 load pass per local layer -------------------------------
    t_setup = time.perf_counter()
    want_host = not args.no_e2e and slab_rows is None
    layers, f32s, pre_events = build_workload_layers(w, n, mine, want_f32=want_host or slab_rows is not None,
                                                     storage="planes" if slab_rows is not None else args.layer_storage,
                                                     block_width=l or None)
    if slab_rows is not None:                      # keep only this rank's rows of its layer (and their transpose)
        layers = [D.make_slab_layer(f, slab_rows.start, len(slab_rows), n_planes=L.n_planes) for f, L in zip(f32s, layers)]
    host_graphs = []
    for f32 in f32s:
        if f32 is not None and want_host:  # what the reference's DataLoader hands to N2V2R: float32 (n, n) on the HOST (pinned)
            h = torch.empty((n, n), dtype=torch.float32, pin_memory=True)
            h.copy_(f32)
            host_graphs.append(h.numpy())
    del f32s
    torch.cuda.synchronize()
    n_planes = layers[0].n_planes if layers else 0
    storage = layers[0].storage if layers else "planes"
    if not l:                                      # default block width: by the layers' storage (the same on every rank)
        l = D.default_block_width(max(DIMS), n, storage)
    setup_s = time.perf_counter() - t_setup
    preprocess_ms = float(np.mean([a.elapsed_time(b) for a, b in pre_events][1:] or [a.elapsed_time(b) for a, b in pre_events]))
    layers_gb = sum(L.nbytes for L in layers) * (world if slab_rows is not None else T / max(len(layers), 1)) / 1e9
    offset = mine.start if len(mine) else 0
    if slab_rows is not None:                      # every rank of a layer group holds the layer's embedding: take one per group
        S_ = world // T

        def gather(e):
            out = torch.empty((world,) + tuple(e.shape[1:]), dtype=e.dtype, device=e.device)
            tdist.all_gather_into_tensor(out, e.contiguous())
            return out[::S_].contiguous()
    else:
        gather = (lambda e: ND.gather_layers(e, T)) if world > 1 else None
    shard = (rank, world, lambda t: tdist.all_reduce(t)) if world > 1 else None

    def step():
        info, emb, scores, bordas = hot_path(layers, w, l, T_global=T, layer_offset=offset, gather=gather, shard=shard,
                                             slab_group=slab_group)
        launches = info["kernel_launches"] + 2
        return info, scores, bordas, launches, emb

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()                  # before the warm-up: nvidia-smi's start-up stall stays out of the timed steps
    for _ in range(args.warmup):
        step()
    barrier()
    if rank == 0:
        sampler.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    alloc0 = allocator_counters()
    e0.record()
    infos, launches = [], 0
    for _ in range(args.steps):
        info, scores, bordas, nl, emb = step()
        infos.append(info)
        launches += nl
    e1.record()
    barrier()
    ms_step = reduce_max(e0.elapsed_time(e1) / max(args.steps, 1))
    alloc_step = allocator_delta(alloc0)
    clocks = sampler.stop() if rank == 0 else None
    sweep_ms = reduce_max(float(np.mean([max(0.0, ms_step - i["total_ms"]) for i in infos])))      # step minus solve
    # ---- untimed parity evidence: independent second solve, subspace angle per swept d ----------------
    cross = None if args.no_cross_check else cross_solve_check(layers, w, l, emb, bordas, T, offset, gather, shard, slab_group)
    del emb

    # ---- roofline of the dominant kernel (tcgen05 streaming product), measured live in the timed steps ----
    roofline = stream_roofline(env, infos, ms_step, args.steps, n, n_planes, l, args.workload, layers)

    # ---- e2e: the public API on host buffers ------------------------------------------------------------
    e2e = None
    if not args.no_e2e and slab_rows is None:
        del layers, scores, bordas
        torch.cuda.empty_cache()
        cfg = dict(embed_dimensions=DIMS, distance_metrics=METRICS, comp_strategy=w["strategy"], seed=SOLVER_SEED,
                   save_dir=None, verbose=-1, block_width=l, layer_sharded=world > 1, host_transfer=args.host_transfer,
                   layer_storage=args.layer_storage)
        node_names = np.arange(n)
        xfer = {}

        phases = []

        def e2e_step():
            ta = time.perf_counter()
            m = N2V2R(graphs=host_graphs, nodes=node_names, config=cfg)
            ranks = m.fit_transform_rank()              # H2D of the graphs + plane build + solve + distances + D2H of the tables
            tb = time.perf_counter()
            agg = m.aggregate_transform()               # rank + Borda + D2H
            phases.append({"fit_transform_rank_s": round(tb - ta, 4), "aggregate_s": round(time.perf_counter() - tb, 4),
                           "transfer_and_load_s": round((m.transfer_stats or {}).get("total_s", 0.0), 4),
                           "solver_s": round(m.solver_info["total_ms"] / 1e3, 4)})
            d2h = m.node_embeddings.shape[0] * n * max(DIMS) * 4 + sum(v.shape[0] * v.shape[1] * 8 for v in ranks.values()) \
                + sum(v.shape[0] * 8 for v in agg.values())
            xfer.update(m.transfer_stats or {})
            return d2h

        for _ in range(args.warmup):
            e2e_step()
        barrier()
        alloc0 = allocator_counters()
        t0 = time.perf_counter()
        d2h = 0
        for _ in range(args.steps):
            d2h = e2e_step()
        barrier()
        e2e_s = reduce_max((time.perf_counter() - t0) / max(args.steps, 1))
        alloc_e2e = allocator_delta(alloc0)
        # whole job: this rank's bytes scaled from its layers to all T (the layers have the same density)
        e2e = {"value": e2e_s, "unit": "s",
               "h2d_bytes_per_step": int(xfer.get("h2d_bytes", len(mine) * n * n * 4) * T / max(len(mine), 1)),
               "d2h_bytes_per_step": int(d2h), "host_graph_bytes_per_step": int(T) * n * n * 4,
               "host_transfer": transfer_note(xfer), "phases_per_step": phases[-args.steps:],
               "allocator": alloc_e2e,
               "api": "N2V2R(graphs=float32 host arrays).fit_transform_rank() + aggregate_transform()"}

    if rank != 0:
        env.finish()
        return 0
    # ---- cpu_baseline (rank 0, N = 1 only) -------------------------------------------------------------
    cpu = cpu_baseline_leg(w, n, T) if (world == 1 and not args.no_cpu) else None
    last = infos[-1]
    line = {
        "metric": "UASE sweep+rank wall time", "value": ms_step / 1e3, "unit": "s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": DTYPE_BY_STORAGE.get(storage, DTYPE_BY_STORAGE["planes"]),
        "data": "synthetic",
        "config": {"workload": w["name"], "n": n, "layers": T, "communities": K, "block_width": l, "planes_per_layer": n_planes,
                   "layers_per_gpu": len(mine), "layer_storage": storage,
                   "row_slabs_per_layer": (world // T if slab_rows is not None else 1),
                   "l2_policy": f"inputs larger than L2: the {T} layers ({layers_gb:.1f} GB in "
                                f"total, every pass streams one of them) against 0.126 GB of L2; no flush needed",
                   "operator_applications": last["iterations"], "layer_passes": last["layer_passes"],
                   "outer_iterations": last["outer_iterations"],
                   # the solver's own stopping record: status 1 = bound <= tol or all residuals at their fp32 target,
                   # 2 = residuals stagnated above the target; angle_bound is the Davis-Kahan UPPER bound it stopped at
                   "converged": last["converged"], "converged_status": last["converged_status"], "bound_met": last["bound_met"],
                   "angle_bound": last["angle_bound"],
                   "angle_bound_per_d": {str(d): float(f"{v:.3e}") for d, v in last["angle_bounds"].items()},
                   "cross_solve": cross, "allreduce_path": last["allreduce_path"],
                   "non_stream_solver_ms": round(float(np.mean([i["total_ms"] - i["stream_ms"] for i in infos])), 2),
                   "sweep_ms": round(sweep_ms, 2), "setup_s": setup_s,
                   "solver_ms": last["total_ms"], "solver_stream_ms": last["stream_ms"], "solver_mma_ms": last["mma_ms"],
                   "solver_ms_per_step": [round(i["total_ms"], 2) for i in infos], "allocator": alloc_step,
                   # wall-time segments: device-side load pass per layer (select + transform + planes; untimed set-up),
                   # the solve, and the sweep (embedding gather + distances for every d + rank + Borda)
                   "segments_ms": {"preprocess_per_layer": round(preprocess_ms, 2),
                                   "solve": round(float(np.mean([i["total_ms"] for i in infos])), 2),
                                   "sweep_rank_borda": round(sweep_ms, 2)}},
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    env.finish()
    return 0


def stream_roofline(env, infos, ms_step, steps, n, n_planes, l, workload, layers=None):
    """Roofline of the dominant kernel, from the per-launch CUDA-event times the solver records around every launch of it.

    Dense bf16 planes (stream_mma_kernel): bound by HBM; achieved = (2 P n^2 + 8 l n) bytes / mean launch time.
    Compressed tiles (sparse_mma_kernel): the layer is expanded in shared memory, so HBM carries only the entry lists and
    the kernel is bound by the tensor pipe: achieved = the bf16 MMA work one launch issues, 2 * rows_pad * n_pad * l * terms
    (terms = 3 operand splits for plane 0, 2 for plane 1, 1 for plane 2: what an fp32-exact product of these weights needs)
    / mean launch time, against the measured sustained bf16 peak; the HBM side of the same launch is reported under "hbm"."""
    mma_ms = sum(i["mma_ms"] for i in infos)
    mma_n = sum(i["mma_launches"] for i in infos)
    avg_s = mma_ms / max(mma_n, 1) * 1e-3
    stream_ms = sum(i["stream_ms"] for i in infos)
    common = {"launches_per_step": mma_n / max(steps, 1), "avg_launch_ms": mma_ms / max(mma_n, 1),
              "share_of_step": mma_ms / max(ms_step * steps, 1e-9),
              "streaming_products_share_of_step": stream_ms / max(ms_step * steps, 1e-9)}
    tiles = bool(layers) and layers[0].storage == "tiles"
    if bool(layers) and layers[0].storage == "rows":
        # row lists (rows_product_kernel): HBM carries the lists once (8 bytes per non-zero + the u16 row headers) and the skinny
        # operand / result; every CTA also reads the whole operand through L2 (panel by panel) and every entry gathers l x 4 bytes
        # of it from shared memory -- that gather, at 128 B/clk/SM, is what bounds the launch, so it is reported beside the HBM line
        L = layers[0]
        list_bytes = float(np.mean([x.n_units * 16 + x.tile_ptr.numel() * 4 for x in layers]))
        nnz = float(np.mean([x.n_entries for x in layers]))
        alg_bytes = list_bytes + 4.0 * l * 2.0 * n
        achieved = alg_bytes / avg_s / 1e9 if mma_n else 0.0
        n_blocks = -(-n // L.tile_rows)
        sms, clk_hz = env.sm_count, env.sm_max_mhz * 1e6
        gather = nnz * l * 4.0
        roof = {"bound": "hbm", "achieved": achieved, "peak": env.hbm_peak, "unit": "GB/s", "frac": achieved / env.hbm_peak,
                "traffic": None, "kernel": "rows_product_kernel", "peak_source": env.hbm_src,
                "algorithmic_bytes_per_launch": alg_bytes, "nonzeros_per_launch": nnz,
                "fp32_flop_per_launch": 2.0 * nnz * l,
                "shared_memory_gather": {"bytes_per_launch": gather, "achieved_bytes_per_clk_per_sm": (gather / avg_s / clk_hz / sms) if mma_n else 0.0,
                                         "peak_bytes_per_clk_per_sm": 128.0,
                                         "frac": (gather / avg_s / clk_hz / sms / 128.0) if mma_n else 0.0,
                                         "clock_mhz_assumed": env.sm_max_mhz},
                "l2_operand_bytes_per_launch": float(n_blocks) * n * l * 4.0,
                "dense_planes_bytes_per_launch": float(n) * n * 2.0 * 2 + 8.0 * l * n,
                "note": "a different algorithm from the dense-plane stream and from the compressed tiles (profiles/: both lines are kept "
                        "beside it): the layer crosses HBM as row lists of its non-zeros (exact fp32 values) and is multiplied entry by "
                        "entry on the FMA pipes; the launch is bound by the shared-memory gather of the operand rows (l x 4 bytes per "
                        "8-byte entry), not by HBM, so the HBM fraction is low by construction"}
        key = f"{workload}_rows_bytes_per_launch" if l == 32 else f"{workload}_rows_l{l}_bytes_per_launch"
    elif tiles:
        L = layers[0]
        terms = {1: 3, 2: 5, 3: 6}[n_planes]
        rows_pad = -(-n // L.tile_rows) * L.tile_rows
        n_pad = -(-n // 64) * 64
        flop = 2.0 * rows_pad * n_pad * l * terms
        entry_bytes = float(np.mean([x.n_entries * 8 + x.tile_ptr.numel() * 4 for x in layers]))
        alg_bytes = entry_bytes + 4.0 * l * 2.0 * n
        achieved = flop / avg_s / 1e12 if mma_n else 0.0
        roof = {"bound": "tensor", "achieved": achieved, "peak": env.tc_peak, "unit": "TFLOP/s", "frac": achieved / env.tc_peak,
                "traffic": None, "kernel": "sparse_mma_kernel", "peak_source": env.tc_src,
                "algorithmic_flop_per_launch": flop, "mma_terms": terms,
                "fp32_equivalent_flop_per_launch": 2.0 * n * n * l,
                "hbm": {"algorithmic_bytes_per_launch": alg_bytes, "achieved_gbs": alg_bytes / avg_s / 1e9 if mma_n else 0.0,
                        "peak_gbs": env.hbm_peak, "frac": (alg_bytes / avg_s / 1e9 / env.hbm_peak) if mma_n else 0.0,
                        "dense_planes_bytes_per_launch": float(n) * n * 2.0 * n_planes + 8.0 * l * n},
                "note": "a different algorithm from the dense-plane stream (profiles/: the dense line is kept beside it): the layer "
                        "crosses HBM as 8-byte entries of its non-zeros and is expanded into dense bf16 tiles in shared memory, so "
                        "the dense tile MMA work, not HBM, bounds the launch"}
        key = f"{workload}_tiles_bytes_per_launch"
    else:
        e4m3 = bool(layers) and layers[0].storage == "e4m3"
        per_entry = 1.0 if e4m3 else 2.0 * n_planes
        rows = layers[0].rows if (layers and layers[0].rows) else n              # a row slab streams rows x n entries per pass
        alg_bytes = float(rows) * n * per_entry + 4.0 * l * (n + rows)         # matrix once + skinny in + skinny out
        achieved = alg_bytes / avg_s / 1e9 if mma_n else 0.0
        roof = {"bound": "hbm", "achieved": achieved, "peak": env.hbm_peak, "unit": "GB/s", "frac": achieved / env.hbm_peak,
                "traffic": None, "kernel": "stream_mma8_kernel" if e4m3 else "stream_mma_kernel", "peak_source": env.hbm_src,
                "algorithmic_bytes_per_launch": alg_bytes}
        key = f"{workload}_e4m3_bytes_per_launch" if e4m3 else f"{workload}_bytes_per_launch"
    roof.update(common)
    # ncu dram__bytes_read + write per launch, captured once per (workload, n, planes, l, storage): only quoted when it matches
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        ent = tr.get(key)
        if isinstance(ent, dict) and (ent.get("n"), ent.get("planes"), ent.get("l")) == (n, n_planes, l):
            roof["traffic"] = ent.get("bytes_per_launch")
            roof["traffic_source"] = ent.get("source")
            wf = ent.get("shared_wavefronts_per_launch")
            if wf and "shared_memory_gather" in roof and mma_n:
                # what actually bounds rows_product_kernel: shared-memory wavefronts (ncu count of the same launch shape, one
                # per clock and SM at most) over the clocks of the live-timed launch
                g = roof["shared_memory_gather"]
                g["wavefronts_per_launch_ncu"] = wf
                g["wavefront_frac"] = wf / (avg_s * env.sm_max_mhz * 1e6 * env.sm_count)
    except (OSError, ValueError):
        pass
    return roof


def run_b200_c5(args, w):
    """BASELINE configs[4]: S seeds x one UASE run each on the same resident layers, Borda across runs.
    Seeds are split over the ranks (every rank holds all layers); score tables are all-gathered."""
    import torch
    from node2vec2rank_b200 import _native as N
    from node2vec2rank_b200 import device as D
    from node2vec2rank_b200 import dist as ND
    from node2vec2rank_b200.stability import StabilityRun

    env = Env()
    n, T, K, l, seeds = w["n"], w["T"], w["K"], args.block_width or 48, w["seeds"]
    mine = [seeds[i] for i in ND.layer_partition(len(seeds), env.world, env.rank)]
    Bs = block_probabilities(w)
    t_setup = time.perf_counter()
    layers, host_graphs = [], []
    for t in range(T):
        raw = D.sbm_generate(n, K, Bs[t], seed=GRAPH_SEED, layer=t, weighted=w["weighted"])
        # seed batching multiplies bf16 planes (three operands per pass); e4m3 layers are served one solve at a time
        c5_storage = "planes" if (args.layer_storage == "auto" and args.seed_batch > 1) else args.layer_storage
        layer, f32 = D.preprocess_to_layer(raw, top_percent_keep=w["keep"], binarize=w["binarize"], want_f32=not args.no_e2e,
                                           storage=c5_storage)
        del raw
        if f32 is not None:
            h = torch.empty((n, n), dtype=torch.float32, pin_memory=True)
            h.copy_(f32)
            host_graphs.append(h.numpy())
            del f32
        layers.append(layer)
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t_setup
    names, cdims, cmets = D.column_specs(DIMS, METRICS)
    strategy = N.STRATEGIES[w["strategy"]]

    group = max(1, min(16, args.seed_batch))
    batchable = group > 1 and layers[0].storage == "planes"
    batch_log = []

    def step():
        infos, local = [], []
        for g0 in range(0, len(mine), group if batchable else 1):
            chunk = mine[g0:g0 + (group if batchable else 1)]
            if batchable and len(chunk) > 1:         # concurrent solves sharing every pass over a layer (device.uase_batched)
                solved, bstats = D.uase_batched(layers, max(DIMS), chunk, l=l, tol=1e-5, max_iters=400, check_dims=DIMS)
                batch_log.append(bstats)
            else:
                solved = [(lambda r: (r[0], r[1], r[3]))(D.uase(layers, max(DIMS), l=l, tol=1e-5, max_iters=400, seed=seed, check_dims=DIMS))
                          for seed in chunk]
            for emb, _, info in solved:
                infos.append(info)
                local.append(D.distances(emb, cdims, cmets, strategy))
        sc = torch.stack(local) if local else torch.empty((0, T - 1, len(cdims), n), dtype=torch.float64, device=env.dev)
        if env.world > 1:
            sc = ND.gather_blocks(sc, len(seeds))
        S, C, cols, _ = sc.shape
        borda = D.rank_borda(sc.permute(1, 0, 2, 3).reshape(C, S * cols, n))[0]
        return infos, borda, sum(i["kernel_launches"] + 1 for i in infos) + 1

    ms_step, results, clocks = env.timed(step, args.steps, args.warmup)
    infos = [i for r in results for i in r[0]]
    launches = sum(r[2] for r in results)
    roofline = stream_roofline(env, infos, ms_step, args.steps, n, layers[0].n_planes, l, args.workload, layers)
    layer_storage = layers[0].storage
    e2e = None
    if not args.no_e2e:
        del layers
        torch.cuda.empty_cache()
        cfg = dict(embed_dimensions=DIMS, distance_metrics=METRICS, comp_strategy=w["strategy"], seed=None, save_dir=None,
                   verbose=-1, block_width=l, seed_sharded=env.world > 1, host_transfer=args.host_transfer)
        xfer = {}

        def e2e_step():
            run = StabilityRun(graphs=host_graphs, nodes=np.arange(n), config=cfg, seeds=seeds)
            agg = run.run()                       # H2D of the graphs, plane build, all solves, Borda across runs, D2H of the tables
            xfer.update(run._model.transfer_stats or {})
            return sum(v.shape[0] * v.shape[1] * 8 for v in run.pairwise_ranks.values()) + sum(v.shape[0] * 8 for v in agg.values()) \
                + sum(v.shape[0] * v.shape[1] * 8 for v in run.per_seed_aggregate_ranks.values())

        e2e_s, d2h = env.timed_host(e2e_step, args.steps, min(args.warmup, 1))
        e2e = {"value": e2e_s, "unit": "s", "h2d_bytes_per_step": int(xfer.get("h2d_bytes", int(T) * n * n * 4)),
               "d2h_bytes_per_step": int(d2h), "host_graph_bytes_per_step": int(T) * n * n * 4, "host_transfer": transfer_note(xfer),
               "api": "StabilityRun(graphs=float32 host arrays, seeds).run()"}
    if env.rank != 0:
        env.finish()
        return 0
    cpu = cpu_baseline_leg(w, n, T) if (env.world == 1 and not args.no_cpu) else None
    line = {
        "metric": "UASE sweep+rank wall time", "value": ms_step / 1e3, "unit": "s", "n_gpus": env.world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": DTYPE_BY_STORAGE.get(layer_storage, DTYPE_BY_STORAGE["planes"]), "data": "synthetic",
        "config": {"workload": w["name"], "n": n, "layers": T, "communities": K, "block_width": l, "seeds": len(seeds),
                   "seeds_per_gpu": len(mine), "planes_per_layer": 1, "layer_storage": layer_storage,
                   "l2_policy": "inputs larger than L2 (each layer pass streams 0.9-1.8 GB)",
                   "operator_applications_per_run": [i["iterations"] for i in results[-1][0]],
                   "all_converged": all(i["converged"] for i in infos), "setup_s": setup_s,
                   "seed_batch": group if batchable else 1,
                   "layer_streams_per_step": (sum(b["layer_streams"] for b in batch_log[-max(len(batch_log) // max(args.steps + args.warmup, 1), 1):])
                                              if batch_log else sum(i["layer_passes"] for i in results[-1][0])),
                   "layer_passes_requested_per_step": sum(i["layer_passes"] for i in results[-1][0])},
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    env.finish()
    return 0


def run_b200_c4(args, w):
    """BASELINE configs[3]: two bipartite layers (2k x 30k) -> W^T W on the tensor cores (K2) -> cut -> planes ->
    UASE -> rank.  The projection is inside the step; its roofline is the tensor pipe."""
    import torch
    from node2vec2rank_b200 import _native as N
    from node2vec2rank_b200 import device as D
    from node2vec2rank_b200.model import N2V2R
    from node2vec2rank_b200.preprocessing_utils import network_transform_device
    from node2vec2rank_b200.synthetic import bipartite_layer_numpy

    env = Env()
    if env.world > 1:
        raise SystemExit("workload c4 is a one-GPU configuration (BASELINE configs[3])")
    n, r, T, K, l = w["n"], w["rows"], w["T"], w["K"], args.block_width
    t_setup = time.perf_counter()
    hosts = []
    for t in range(T):
        h = torch.empty((r, n), dtype=torch.float32, pin_memory=True)
        h.copy_(torch.from_numpy(bipartite_layer_numpy(r, n, K, GRAPH_SEED, t)))
        hosts.append(h)
    Ws = [h.to(env.dev) for h in hosts]
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t_setup
    names, cdims, cmets = D.column_specs(DIMS, METRICS)
    strategy = N.STRATEGIES[w["strategy"]]
    proj_events = []

    def step():
        layers = []
        for W in Ws:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            P = D.project(W, "columns")
            b.record()
            proj_events.append((a, b))
            layers.append(D.preprocess_to_layer(P, top_percent_keep=w["keep"])[0])
            del P
        lw = l or D.default_block_width(max(DIMS), n, layers[0].storage)
        emb, _, _, info = D.uase(layers, max(DIMS), l=lw, tol=1e-5, max_iters=400, seed=SOLVER_SEED, check_dims=DIMS)
        scores = D.distances(emb, cdims, cmets, strategy)
        borda = D.rank_borda(scores)[0]
        info["n_planes"] = layers[0].n_planes
        info["block_width"], info["layer_storage"] = lw, layers[0].storage
        return info, borda, info["kernel_launches"] + 2 + T * 12

    ms_step, results, clocks = env.timed(step, args.steps, args.warmup)
    infos = [x[0] for x in results]
    launches = sum(x[2] for x in results)
    timed_events = proj_events[-args.steps * T:]
    proj_ms = float(np.mean([a.elapsed_time(b) for a, b in timed_events]))
    flop = 2.0 * r * n * n
    achieved = flop / (proj_ms * 1e-3) / 1e12
    roofline = {"bound": "tensor", "achieved": achieved, "peak": env.tc_peak, "unit": "TFLOP/s", "frac": achieved / env.tc_peak,
                "traffic": None, "kernel": "stream_mma_kernel (SYRK mode) + split_planes_kernel + mirror_reduce_kernel",
                "peak_source": env.tc_src, "algorithmic_flop_per_launch": flop, "launches_per_step": T, "avg_launch_ms": proj_ms,
                "share_of_step": proj_ms * T / ms_step,
                "note": "algorithmic 2*r*c*n_out FLOP of the full product; the kernel issues 6 bf16 MMA terms (3x3 split, 6 kept) on "
                        "the upper triangle only, i.e. 3x the algorithmic MMA work",
                "streaming_products": stream_roofline(env, infos, ms_step, args.steps, n, infos[-1]["n_planes"], infos[-1]["block_width"], args.workload)}
    e2e = None
    if not args.no_e2e:
        del Ws
        torch.cuda.empty_cache()
        cfg = dict(embed_dimensions=DIMS, distance_metrics=METRICS, comp_strategy=w["strategy"], seed=SOLVER_SEED, save_dir=None,
                   verbose=-1, block_width=l or None)

        def e2e_step():
            graphs = [network_transform_device(h.numpy(), top_percent_keep=w["keep"], project_unipartite_on="columns")[0] for h in hosts]
            m = N2V2R(graphs=graphs, nodes=np.arange(n), config=cfg)
            ranks = m.fit_transform_rank()
            agg = m.aggregate_transform()
            return T * n * max(DIMS) * 4 + sum(v.shape[0] * v.shape[1] * 8 for v in ranks.values()) + sum(v.shape[0] * 8 for v in agg.values())

        e2e_s, d2h = env.timed_host(e2e_step, args.steps, args.warmup)
        e2e = {"value": e2e_s, "unit": "s", "h2d_bytes_per_step": T * r * n * 4, "d2h_bytes_per_step": int(d2h),
               "api": "network_transform_device(host float32 bipartite, project_unipartite_on='columns') + N2V2R(...).fit_transform_rank() "
                      "+ aggregate_transform()"}
    cpu = None
    if not args.no_cpu:
        cpu = cpu_c4_leg(w)
    last = infos[-1]
    line = {
        "metric": "UASE sweep+rank wall time", "value": ms_step / 1e3, "unit": "s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": "projection: 3x3 split-bf16 (6 terms) fp32 accumulate (tcgen05); solve: bf16 planes x3-split operand; fp64 small dense",
        "data": "synthetic",
        "config": {"workload": w["name"], "n": n, "bipartite_rows": r, "layers": T, "communities": K, "block_width": infos[-1]["block_width"],
                   "layer_storage": infos[-1]["layer_storage"],
                   "planes_per_layer": last["n_planes"], "l2_policy": "inputs larger than L2 (each projected layer is 3.6 GB fp32)",
                   "operator_applications": last["iterations"], "layer_passes": last["layer_passes"], "converged": last["converged"],
                   "angle_bound": last["angle_bound"], "projection_ms_per_layer": proj_ms, "solver_ms": last["total_ms"], "setup_s": setup_s},
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    env.finish()
    return 0


def cpu_c4_leg(w, cols_sample=4000):
    """The reference's path on a bounded C4 sample: float64 W^T W (np.matmul, all BLAS threads) + transform + fit/rank/Borda
    at n = cols_sample; scaled by (n/n_s)^2 for everything quadratic in n (a lower bound for the fit)."""
    from node2vec2rank_b200.synthetic import bipartite_layer_numpy
    from oracle import reference_port as port
    r, n = w["rows"], w["n"]
    t0 = time.perf_counter()
    graphs = [port.network_transform(bipartite_layer_numpy(r, cols_sample, w["K"], GRAPH_SEED, t).astype(np.float64),
                                     top_percent_keep=w["keep"], project_unipartite_on="columns") for t in range(w["T"])]
    t1 = time.perf_counter()
    tm = cpu_step(graphs, w)
    total = (t1 - t0) + tm["total"]
    return {"value": total * (n / cols_sample) ** 2, "unit": "s", "cores": os.cpu_count() or 1, "kind": "port",
            "sample": f"{r} x {cols_sample} of {r} x {n} per layer; measured {total:.2f} s (projection+transform {t1 - t0:.2f}, fit {tm['fit']:.2f}, "
                      f"distances {tm['distances']:.2f}, borda {tm['borda']:.2f}); value = measured x (n_full/n_sample)^2 (lower bound)",
            "sample_seconds": total}


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_b200_arm(args)


if __name__ == "__main__":
    sys.exit(main())
