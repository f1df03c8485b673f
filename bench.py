#!/usr/bin/env python
"""bench.py — pod x node Filter+Score+NormalizeScore evaluations per second.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one pass of the hot path over one batch: 50 000 pending pods against 5 000
nodes x 8 GPUs per rank (BASELINE.json configs[1]); with N ranks the nodes are sharded
(5 000 per rank, weak scaling) and the per-pod extrema are all-reduced between the scoring
and the normalise kernels (the path's one exchange step).

Prints ONE JSON line (rank 0).  `value` = device-resident throughput (inputs in HBM when
the timed region starts); `e2e` = the same metric through the host-buffer C-ABI call
gs_score_batch, host<->device copies inside the timed region; `roofline` = the dominant
kernel against the measured HBM bandwidth; `cpu_baseline` = the CPU oracle (a port of the
reference's Go algorithm) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

N_PODS, N_NODES, GPUS_PER_NODE, N_APPS = 50_000, 5_000, 8, 1024
METRIC = "pod x node Filter+Score+NormalizeScore evals/sec (50k pods x 5k nodes x 8 GPUs/node per rank)"
UNIT = "evals/s"
# what the library's event bracket (gs_last_score_kernel_ms) spans
BRACKET_KERNELS = "score_pods_kernel (stage_fold_kernel is bracketed separately: stage_a_ms)"


def source_sha16():
    """Hash of the CUDA sources the library is built from: ties a committed ncu capture to a build."""
    import hashlib

    h = hashlib.sha256()
    csrc = os.path.join(ROOT, "k8s-gpu-scheduler_b200", "csrc")
    for name in sorted(os.listdir(csrc)):
        if name.endswith((".cu", ".cuh", ".h")):
            with open(os.path.join(csrc, name), "rb") as f:
                h.update(name.encode() + b"\0" + f.read())
    return h.hexdigest()[:16]


def measured_traffic(workload_key):
    """dram__bytes_read.sum + dram__bytes_write.sum per step from `ncu --set full` of THIS build
    (profiles/r2/traffic.json, written by profiles/ncu_traffic.py from the capture); None when the
    sources changed since that capture or the workload differs."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2", "traffic.json")) as f:
            t = json.load(f)
        if t.get("src_sha16") == source_sha16() and t.get("workload") == workload_key:
            return t
    except Exception:
        pass
    return None


def workload_name(n_ranks, pods=N_PODS, nodes=N_NODES):
    return (f"batched Filter+Score+NormalizeScore, {pods} pods x {nodes} nodes x {GPUS_PER_NODE} GPUs/node per rank"
            f" ({n_ranks} rank(s), nodes sharded)")


def algorithmic_bytes(n_pods, n_nodes, n_res, n_apps, n_cols, gpn):
    """2 B written per eval + every input table read once + 4 B/pod best node."""
    return (2 * n_pods * n_nodes + 8 * n_pods + n_nodes * 2 + 4 * (n_nodes * gpn + 1) + 8 * n_res
            + 4 * n_apps * n_cols + 8 * n_apps * n_apps + 4 * n_pods)


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region: NVML polled every few ms from a thread
    (nvidia-smi as a fallback, which is too slow for short regions)."""

    REASONS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, index):
        self.index, self.sm, self.mask, self.max_sm = index, [], 0, None
        self._stop, self._t, self._nvml, self._h = threading.Event(), None, None, None
        try:
            import pynvml

            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else index
            self._h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self._nvml = pynvml
        except Exception:
            self._nvml = None

    def _poll_nvml(self):
        n = self._nvml
        while not self._stop.is_set():
            try:
                self.sm.append(float(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)))
                self.mask |= int(n.nvmlDeviceGetCurrentClocksEventReasons(self._h))
            except Exception:
                try:
                    self.mask |= int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                except Exception:
                    pass
            self._stop.wait(0.003)

    def _poll_smi(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.active"
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                f = [x.strip() for x in out.splitlines()[0].split(",")]
                self.sm.append(float(f[0]))
                self.max_sm = float(f[1])
                self.mask |= int(f[2], 16)
            except Exception:
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        self._t = threading.Thread(target=self._poll_nvml if self._nvml else self._poll_smi, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=10)

    def summary(self):
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.max_sm, "reasons": ["unsampled"], "samples": 0}
        sm = sorted(self.sm)
        reasons = [k for k, bit in self.REASONS.items() if self.mask & bit]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.max_sm, "reasons": reasons, "samples": len(sm),
                "source": "nvml" if self._nvml else "nvidia-smi"}


def cpu_reference_leg(preds, cluster, pods, budget_evals_per_thread=3_000_000, steps=1, warmup=0):
    """The CPU restatement of the reference algorithm on a bounded sample (first pods of the batch)."""
    from oracle.score_oracle import Oracle, num_threads

    # all host cores this process may use (torchrun exports OMP_NUM_THREADS=1; the oracle sets its
    # own thread count explicitly)
    try:
        threads = len(os.sched_getaffinity(0))
    except AttributeError:
        threads = os.cpu_count() or num_threads()
    n_sample = max(8, min(pods.n_pods, int(budget_evals_per_thread * threads / max(1, cluster.n_nodes))))
    sub = type(pods)(app=pods.app[:n_sample], slo=pods.slo[:n_sample])
    ora = Oracle(preds, cluster)
    for _ in range(warmup):
        ora.score_batch(type(pods)(app=pods.app[:max(8, n_sample // 8)], slo=pods.slo[:max(8, n_sample // 8)]), threads=threads)
    times = []
    for _ in range(max(1, steps)):
        t0 = time.perf_counter()
        ora.score_batch(sub, normalize=True, threads=threads)
        times.append(time.perf_counter() - t0)
    evals = n_sample * cluster.n_nodes
    # the same restatement on one thread (report 1 thread and all cores)
    n_one = max(8, n_sample // max(1, threads))
    one = type(pods)(app=pods.app[:n_one], slo=pods.slo[:n_one])
    t0 = time.perf_counter()
    ora.score_batch(one, normalize=True, threads=1)
    one_thread = n_one * cluster.n_nodes / (time.perf_counter() - t0)
    return {"value": evals * len(times) / sum(times), "unit": UNIT, "cores": threads, "kind": "port",
            "single_thread_value": one_thread, "sampled_pods": n_sample,
            "sample": f"first {n_sample} of {pods.n_pods} pods x all {cluster.n_nodes} nodes ({evals} evals/step), "
                      f"oracle/score_oracle.c -O2, OpenMP over pods"}, sum(times) / len(times)


def bench_config(pkg, world, P, NL, n_apps, app_dist, pods, n_res_rank0):
    """The `config` object: identical in both arms (the driver compares them)."""
    return {"workload": workload_name(world, P, NL), "pods": P, "nodes_per_rank": NL, "gpus_per_node": GPUS_PER_NODE,
            "apps": n_apps, "app_distribution": app_dist, "residents_rank0": n_res_rank0,
            "distinct_app_slo_classes": pkg.synth.distinct_classes(pods), "seed": hex(pkg.synth.SEED),
            "l2": ("flushed between timed iterations (256 MB memset, untimed); outputs 500 MB > L2" if world == 1 else
                   "no flush: every step writes 500 MB of results per rank, four times the L2"),
            "timing": ("CUDA events per step on the launching stream, summed" if world == 1 else
                       "CUDA events around the K steps; step k's exchange (all-reduce + NormalizeScore, second stream) "
                       "overlaps step k+1's scoring; max over ranks")}


def run_reference(args, rank, world):
    if rank != 0:
        return
    import __graft_entry__ as entry

    pkg = entry.load_package()
    P, NL, n_apps = args.pods, args.nodes, args.apps
    preds = pkg.synth.make_predictions(n_apps)
    cluster_all = pkg.synth.make_cluster(NL * world, n_apps, GPUS_PER_NODE)
    cluster = cluster_all.shard(0, NL) if world > 1 else cluster_all
    pods = pkg.synth.make_pods(P, n_apps, app_dist=args.app_dist)
    cb, sec = cpu_reference_leg(preds, cluster, pods, steps=args.steps, warmup=min(args.warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32+f64", "data": "synthetic",
        "config": bench_config(pkg, world, P, NL, n_apps, args.app_dist, pods, int(cluster.res_app.size)),
        "reference_note": "CPU port of the reference's Go Logic/NormalizeScore (Go toolchain absent); each step is a "
                          "bounded sample of the workload (cpu_baseline.sample); the real reference also pays "
                          "2*G*(R+1) gRPC dials + Redis + API-server writes per Score, omitted here",
        "sampled_pods": cb["sampled_pods"],
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def completion_leg(pkg, sch, rank, world, n=20000, d=512, rank_k=64):
    """Recommender completion proxy: Imputer.fit on n x d (rows sharded over the ranks at N > 1, per-feature
    moments all-reduced through the library's callback), achieved vs the hidden ground truth, and the
    200 x 12 golden of the reference's Imputer refitted here."""
    rng = np.random.default_rng(5)
    X = rng.normal(size=(n, rank_k)) @ rng.normal(size=(rank_k, d)) + rng.normal(scale=0.01, size=(n, d))
    truth = X.copy()
    mask = rng.random((n, d)) < 0.30
    mask[rng.integers(0, n, size=d), np.arange(d)] = False
    X[mask] = np.nan
    lo, hi = rank * n // world, (rank + 1) * n // world
    e0 = pkg.eig_stats()
    t0 = time.perf_counter()
    if world > 1:
        m = pkg.complete_fit_sharded(sch, X[lo:hi], pkg.torch_allreduce())
    else:
        m = pkg.complete_fit(sch, X)
    fit_s = time.perf_counter() - t0
    e1 = pkg.eig_stats()
    tm = sch.complete_last_timing()
    out = m.transform(X[lo:hi])
    hid = mask[lo:hi]
    rmse = float(np.sqrt(np.mean((out[hid] - truth[lo:hi][hid]) ** 2))) if hid.any() else 0.0
    res = {"table": f"{n}x{d}, rank {rank_k} + N(0, 0.01^2) noise, 30 % missing", "rows_per_rank": hi - lo, "fit_s": fit_s,
           "rounds": m.n_rounds, "estimator_fits": m.n_steps, "ms_per_fit": 1e3 * fit_s / max(1, m.n_steps),
           "rmse_hidden_vs_truth": rmse,
           # device time per estimator fit, CUDA events on a sample of the fits (gs_complete_last_timing)
           "moments_ms_per_fit": tm["moments_ms"], "gram_ms_per_fit": tm["gram_ms"],
           "estimator_ms_per_fit": tm["estimator_ms"], "fits_timed": tm["fits_sampled"],
           # the per-fit Gram runs over the rows where the feature is missing (30 % of this rank's), on the FP64 tensor cores
           "gram_fp64_tflops": (2.0 * float(hid.sum()) / d * d * d / (tm["gram_ms"] * 1e-3) / 1e12) if tm["gram_ms"] > 0 else None,
           "solver": ("block Jacobi eigensolver (csrc/eig_jacobi.cu), %.1f sweeps per fit" % ((e1[1] - e0[1]) / (e1[0] - e0[0]))
                      if e1[0] > e0[0] else
                      "Householder reduction to tridiagonal form + evidence iteration in that basis (csrc/tridiag.cu)")}
    m.close()
    gpath = os.path.join(ROOT, "tests", "golden", "impute_synth_200x12.npz")
    if os.path.exists(gpath):
        g = np.load(gpath)
        m2 = pkg.complete_fit(sch, g["X"])
        got, want = m2.transform(g["X"]), g["completed"]
        floor = 1e-2 * float(np.mean(np.abs(want)))
        res["golden_max_rel_err"] = float(np.max(np.abs(got - want) / np.maximum(np.abs(want), floor)))
        res["golden"] = "tests/golden/impute_synth_200x12.npz (reference Imputer, scikit-learn " + str(g["sklearn_version"]) + ")"
        m2.close()
    return res


def parity_check(pkg, preds, cluster_all, pods, rank, NL, d_score, d_sel, d_best, n_check=64):
    """Untimed: n_check pods of this rank's node shard against the CPU oracle over the WHOLE cluster
    (global NormalizeScore extrema, global best node) — bit for bit."""
    from oracle.score_oracle import Oracle

    pick = np.random.default_rng(1234 + rank).choice(pods.n_pods, size=min(n_check, pods.n_pods), replace=False)
    pick.sort()
    sub = type(pods)(app=pods.app[pick], slo=pods.slo[pick])
    w_score, w_sel, w_best = Oracle(preds, cluster_all).score_batch(sub, normalize=True)
    cols = slice(rank * NL, (rank + 1) * NL)
    import torch

    idx = torch.from_numpy(pick).to(d_score.device)
    g_score = d_score.index_select(0, idx).cpu().numpy()
    g_sel = d_sel.index_select(0, idx).cpu().numpy()
    g_best = d_best.index_select(0, idx).cpu().numpy()
    return bool(np.array_equal(g_score, w_score[:, cols]) and np.array_equal(g_sel, w_sel[:, cols])
                and np.array_equal(g_best, w_best))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--pods", type=int, default=N_PODS)
    ap.add_argument("--nodes", type=int, default=N_NODES)
    ap.add_argument("--apps", type=int, default=N_APPS, help="sensitivity runs only; the metric is quoted at 1024")
    ap.add_argument("--app-dist", default="zipf", choices=["zipf", "uniform"], help="sensitivity runs only")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist

    import __graft_entry__ as entry

    pkg = entry.load_package()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    P, NL = args.pods, args.nodes                   # nodes per rank (weak scaling)
    n_apps = args.apps
    preds = pkg.synth.make_predictions(n_apps)
    cluster_all = pkg.synth.make_cluster(NL * world, n_apps, GPUS_PER_NODE)
    cluster = cluster_all.shard(rank * NL, (rank + 1) * NL) if world > 1 else cluster_all
    pods = pkg.synth.make_pods(P, n_apps, app_dist=args.app_dist)

    sch = pkg.Scheduler(local_rank)
    sch.upload_predictions(preds)
    sch.upload_cluster(cluster, node_base=rank * NL)
    sch.set_profiling(True)

    d_app = torch.from_numpy(pods.app.view(np.int32)).to(dev)
    d_slo = torch.from_numpy(pods.slo).to(dev)
    d_score = torch.empty((P, NL), dtype=torch.uint8, device=dev)
    d_sel = torch.empty((P, NL), dtype=torch.uint8, device=dev)
    d_ext = torch.zeros((2, P), dtype=torch.int64, device=dev)   # [hi_arg | neg_lo], one all-reduce(MAX)
    d_best = torch.empty(P, dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    # a non-default stream: a NULL stream argument would mean "the context's own stream"
    work_stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(work_stream)

    def step():
        st = torch.cuda.current_stream().cuda_stream
        assert st != 0
        sch.score_raw_dev(d_app, d_slo, P, d_score, d_sel, d_ext[0], d_ext[1], stream=st)
        if world > 1:
            dist.all_reduce(d_ext, op=dist.ReduceOp.MAX)
        sch.normalize_dev(P, d_score, d_ext[0], d_ext[1], d_best, stream=st)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()

    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    pipelined = world > 1
    if pipelined:
        # N > 1: the exchange of step k (all-reduce of the extrema + NormalizeScore) runs on a second stream
        # and overlaps the scoring of step k + 1 (two result-buffer sets) — a serving loop; no L2 flush inside
        # (each step writes 500 MB of results, four times the L2)
        sets = [(d_score, d_sel, d_ext, d_best),
                (torch.empty_like(d_score), torch.empty_like(d_sel), torch.zeros_like(d_ext), torch.empty_like(d_best))]
        xchg_stream = torch.cuda.Stream(device=dev)
        ev_scored = [torch.cuda.Event() for _ in range(2)]
        ev_done = [torch.cuda.Event() for _ in range(2)]

        def run_pipelined(k_steps):
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record(work_stream)
            for k in range(k_steps):
                sc, se, ex, be = sets[k % 2]
                if k >= 2:
                    work_stream.wait_event(ev_done[k % 2])     # the buffers of step k - 2 are free again
                sch.score_raw_dev(d_app, d_slo, P, sc, se, ex[0], ex[1], stream=work_stream.cuda_stream)
                ev_scored[k % 2].record(work_stream)
                with torch.cuda.stream(xchg_stream):
                    xchg_stream.wait_event(ev_scored[k % 2])
                    dist.all_reduce(ex, op=dist.ReduceOp.MAX)
                    sch.normalize_dev(P, sc, ex[0], ex[1], be, stream=xchg_stream.cuda_stream)
                    ev_done[k % 2].record(xchg_stream)
            xchg_stream.wait_stream(work_stream)
            t1.record(xchg_stream)
            return t0, t1

        run_pipelined(2)
        torch.cuda.synchronize()
    with ClockSampler(local_rank) as clocks:
        barrier()
        launches0 = pkg.launch_count()
        t_wall0 = time.perf_counter()
        if pipelined:
            t0, t1 = run_pipelined(args.steps)
        else:
            for a, b in ev:
                flush.zero_()          # evict L2 between timed iterations (not timed)
                a.record()
                step()
                b.record()
        barrier()
        t_wall = time.perf_counter() - t_wall0
    step_ms = [t0.elapsed_time(t1)] if pipelined else [a.elapsed_time(b) for a, b in ev]
    launches = pkg.launch_count() - launches0
    # dominant-kernel duration, live, from the library's own event bracket on the same stream
    kms, ams = [], []
    for _ in range(3):
        flush.zero_()
        step()
        torch.cuda.synchronize()
        kms.append(sch.last_score_kernel_ms())
        ams.append(sch.last_stage_a_ms())
    # untimed: this rank's results against the CPU oracle (after a fresh step, so the buffers hold its output)
    step()
    torch.cuda.synchronize()
    ok = parity_check(pkg, preds, cluster_all, pods, rank, NL, d_score, d_sel, d_best)
    ok_t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
    if world > 1:
        dist.all_reduce(ok_t, op=dist.ReduceOp.MIN)
    parity_ok = bool(int(ok_t.item()))
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    evals_per_step = P * NL * world
    value = evals_per_step * args.steps / (total_ms * 1e-3)

    # ---- extras (N = 1): what class dedupe buys, and the same step on a Deployment-like batch
    extras = {}
    if world == 1 and not args.no_extras:
        def timed(app_t, slo_t, flags, n=5):
            def one():
                st = torch.cuda.current_stream().cuda_stream
                sch.score_raw_dev(app_t, slo_t, P, d_score, d_sel, d_ext[0], d_ext[1], stream=st, flags=flags)
                sch.normalize_dev(P, d_score, d_ext[0], d_ext[1], d_best, stream=st)
            for _ in range(3):
                one()
            evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
            torch.cuda.synchronize()
            for a, b in evs:
                flush.zero_()
                a.record()
                one()
                b.record()
            torch.cuda.synchronize()
            ms = sum(a.elapsed_time(b) for a, b in evs) / n
            return {"ms_per_step": ms, "value": P * NL / (ms * 1e-3), "unit": UNIT,
                    "hbm_frac_of_whole_step": abytes_all / (ms * 1e-3) / 1e9 / measured_peak_gbs()[0]}

        abytes_all = algorithmic_bytes(P, NL, int(cluster.res_app.size), n_apps, preds.conf.shape[1], GPUS_PER_NODE)
        extras["all_classes_unique"] = dict(timed(d_app, d_slo, pkg.SCORE_NO_DEDUPE),
                                            note="same batch, every pod scored itself (GS_SCORE_NO_DEDUPE): the worst case")
        dp = pkg.synth.make_pods(P, n_apps, app_dist=args.app_dist, slo_levels=16)
        dd_app = torch.from_numpy(dp.app.view(np.int32)).to(dev)
        dd_slo = torch.from_numpy(dp.slo).to(dev)
        r = timed(dd_app, dd_slo, 0)
        sch.score_raw_dev(dd_app, dd_slo, P, d_score, d_sel, d_ext[0], d_ext[1],
                          stream=torch.cuda.current_stream().cuda_stream)
        sch.normalize_dev(P, d_score, d_ext[0], d_ext[1], d_best, stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        r.update(distinct_app_slo_classes=pkg.synth.distinct_classes(dp),
                 parity_ok=parity_check(pkg, preds, cluster_all, dp, rank, NL, d_score, d_sel, d_best),
                 note="Deployment-like batch: 16 SLO values per app, same apps and cluster; pods with equal "
                      "(app, SLO) are scored once and their rows copied")
        extras["discrete_slo"] = r
        r2 = timed(dd_app, dd_slo, pkg.SCORE_NO_DEDUPE)
        extras["discrete_slo"]["ms_per_step_without_dedupe"] = r2["ms_per_step"]

    # ---- completion (BASELINE configs[2] in miniature): fit of a rank-16 + noise table with 30 % missing
    # entries, and a small table against its committed golden (the reference's own Imputer) in the same run
    completion = None
    if not args.no_extras:
        completion = completion_leg(pkg, sch, rank, world)

    # ---- BASELINE configs[3] at N = 8: 1M pending pods x 20k nodes, 2 500 nodes per rank
    config4 = None
    if world == 8 and not args.no_extras:
        P4, NL4 = 1_000_000, 2_500
        cl4_all = pkg.synth.make_cluster(NL4 * world, n_apps, GPUS_PER_NODE, seed=pkg.synth.SEED + 4)
        pods4 = pkg.synth.make_pods(P4, n_apps, seed=pkg.synth.SEED + 4)
        s4 = pkg.Scheduler(local_rank)
        s4.upload_predictions(preds)
        s4.upload_cluster(cl4_all.shard(rank * NL4, (rank + 1) * NL4), node_base=rank * NL4)
        a4 = torch.from_numpy(pods4.app.view(np.int32)).to(dev)
        l4 = torch.from_numpy(pods4.slo).to(dev)
        sc4 = torch.empty((P4, NL4), dtype=torch.uint8, device=dev)
        se4 = torch.empty((P4, NL4), dtype=torch.uint8, device=dev)
        ex4 = torch.zeros((2, P4), dtype=torch.int64, device=dev)
        be4 = torch.empty(P4, dtype=torch.int32, device=dev)

        def step4():
            st = torch.cuda.current_stream().cuda_stream
            s4.score_raw_dev(a4, l4, P4, sc4, se4, ex4[0], ex4[1], stream=st)
            dist.all_reduce(ex4, op=dist.ReduceOp.MAX)
            s4.normalize_dev(P4, sc4, ex4[0], ex4[1], be4, stream=st)

        step4()
        barrier()
        t0e, t1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0e.record()
        for _ in range(3):
            step4()
        t1e.record()
        barrier()
        ms4 = torch.tensor([t0e.elapsed_time(t1e) / 3], dtype=torch.float64, device=dev)
        dist.all_reduce(ms4, op=dist.ReduceOp.MAX)
        ok4 = parity_check(pkg, preds, cl4_all, pods4, rank, NL4, sc4, se4, be4, n_check=8)
        ok4_t = torch.tensor([1 if ok4 else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(ok4_t, op=dist.ReduceOp.MIN)
        config4 = {"workload": f"{P4} pods x {NL4 * world} nodes x {GPUS_PER_NODE} GPUs/node, nodes sharded over {world} ranks",
                   "ms_per_step": float(ms4.item()), "value": P4 * NL4 * world / (float(ms4.item()) * 1e-3), "unit": UNIT,
                   "parity_ok": bool(int(ok4_t.item())), "distinct_app_slo_classes": pkg.synth.distinct_classes(pods4)}
        del sc4, se4
        s4.close()

    # ---- e2e: host buffers through gs_score_batch (pinned), copies inside the timed region
    e2e = None
    if not args.no_e2e:
        h_score = torch.empty((P, NL), dtype=torch.uint8).pin_memory()
        h_sel = torch.empty((P, NL), dtype=torch.uint8).pin_memory()
        h_best = torch.empty(P, dtype=torch.int32).pin_memory()
        h_app = torch.from_numpy(pods.app.view(np.int32)).pin_memory()
        h_slo = torch.from_numpy(pods.slo).pin_memory()
        hp = pkg.Pods(app=h_app.numpy().view(np.uint32), slo=h_slo.numpy())
        out = (h_score.numpy(), h_sel.numpy(), h_best.numpy())
        if world == 1:
            for _ in range(2):
                sch.score_batch(hp, normalize=True, out=out)
            n_e2e = max(3, min(args.steps, 10))
            barrier()
            t0 = time.perf_counter()
            for _ in range(n_e2e):
                sch.score_batch(hp, normalize=True, out=out)
            barrier()
            dt = (time.perf_counter() - t0) / n_e2e
            e2e = {"value": P * NL / dt, "unit": UNIT, "h2d_bytes_per_step": 8 * P,
                   "d2h_bytes_per_step": 2 * P * NL + 4 * P, "ms_per_step": dt * 1e3,
                   "api": "gs_score_batch (host buffers in, score+selection matrices and best node out)"}
        else:
            # N>1: device API + NCCL between the phases, host copies around it
            def e2e_step():
                st = torch.cuda.current_stream().cuda_stream
                d_app.copy_(h_app, non_blocking=True)
                d_slo.copy_(h_slo, non_blocking=True)
                sch.score_raw_dev(d_app, d_slo, P, d_score, d_sel, d_ext[0], d_ext[1], stream=st)
                dist.all_reduce(d_ext, op=dist.ReduceOp.MAX)
                sch.normalize_dev(P, d_score, d_ext[0], d_ext[1], d_best, stream=st)
                h_score.copy_(d_score, non_blocking=True)
                h_sel.copy_(d_sel, non_blocking=True)
                h_best.copy_(d_best, non_blocking=True)
                torch.cuda.synchronize()
            e2e_step()
            n_e2e = max(3, min(args.steps, 10))
            barrier()
            t0 = time.perf_counter()
            for _ in range(n_e2e):
                e2e_step()
            barrier()
            dt = torch.tensor([(time.perf_counter() - t0) / n_e2e], dtype=torch.float64, device=dev)
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            dt = float(dt.item())
            e2e = {"value": evals_per_step / dt, "unit": UNIT, "h2d_bytes_per_step": 8 * P,
                   "d2h_bytes_per_step": 2 * P * NL + 4 * P, "ms_per_step": dt * 1e3,
                   "api": "gs_score_raw_dev + all_reduce(MAX) + gs_normalize_dev with pinned host copies, per rank"}

    # ---- e2e, decisions only: host pods in, 8 bytes per pod out (node, selected GPU there, extrema, MPS choice)
    e2e_dec = None
    if not args.no_e2e:
        h_dec = torch.empty((P, 8), dtype=torch.uint8).pin_memory()
        if world == 1:
            dec_out = h_dec.numpy().view(sch.DECISION_DTYPE).reshape(P)
            for _ in range(2):
                sch.score_decisions(hp, out=dec_out)
            n_e2e = max(3, min(args.steps, 10))
            barrier()
            t0 = time.perf_counter()
            for _ in range(n_e2e):
                sch.score_decisions(hp, out=dec_out)
            barrier()
            dt = (time.perf_counter() - t0) / n_e2e
            api = "gs_score_decisions (host pods in, 8-byte decisions out)"
        else:
            d_dec = torch.empty((P, 8), dtype=torch.uint8, device=dev)

            def dec_step():
                st = torch.cuda.current_stream().cuda_stream
                d_app.copy_(h_app, non_blocking=True)
                d_slo.copy_(h_slo, non_blocking=True)
                sch.score_raw_dev(d_app, d_slo, P, d_score, d_sel, d_ext[0], d_ext[1], stream=st)
                dist.all_reduce(d_ext, op=dist.ReduceOp.MAX)
                sch.decisions_dev(d_app, d_slo, P, d_sel, d_ext[0], d_ext[1], d_dec, stream=st)
                h_dec.copy_(d_dec, non_blocking=True)
                torch.cuda.synchronize()
            dec_step()
            n_e2e = max(3, min(args.steps, 10))
            barrier()
            t0 = time.perf_counter()
            for _ in range(n_e2e):
                dec_step()
            barrier()
            dtt = torch.tensor([(time.perf_counter() - t0) / n_e2e], dtype=torch.float64, device=dev)
            dist.all_reduce(dtt, op=dist.ReduceOp.MAX)
            dt = float(dtt.item())
            api = "gs_score_raw_dev + all_reduce(MAX) + gs_decisions_dev with pinned host copies, per rank"
        e2e_dec = {"value": evals_per_step / dt, "unit": UNIT, "h2d_bytes_per_step": 8 * P, "d2h_bytes_per_step": 8 * P,
                   "ms_per_step": dt * 1e3, "api": api}

    if rank == 0:
        peak, peak_kind = measured_peak_gbs()
        kernel_ms = float(np.median(kms))
        abytes = algorithmic_bytes(P, NL, int(cluster.res_app.size), n_apps, preds.conf.shape[1], GPUS_PER_NODE)
        achieved = abytes / (kernel_ms * 1e-3) / 1e9
        traffic = measured_traffic(f"{P}x{NL}x{GPUS_PER_NODE}:{n_apps}:{args.app_dist}")
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32+f64", "data": "synthetic",
            "config": bench_config(pkg, world, P, NL, n_apps, args.app_dist, pods, int(cluster.res_app.size)),
            "parity_ok": parity_ok,
            "parity": "64 pods x this rank's node shard per rank vs the CPU oracle over the whole cluster, bit-exact "
                      "(normalised score, selection byte, best node); all ranks must agree",
            "gpu_launches": launches,  # counted by the library (gs_launch_count) over the timed region
            "wall_ms_per_step_incl_flush": t_wall * 1e3 / args.steps,
            "clocks": clocks.summary(),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         # dram__bytes_read.sum + dram__bytes_write.sum of the bracketed kernels, from an ncu
                         # capture of THIS build (profiles/r2/traffic.json), else null
                         "traffic": traffic["traffic_bytes"] if traffic else None,
                         "traffic_source": traffic["source"] if traffic else None,
                         "peak_kind": peak_kind, "kernel": BRACKET_KERNELS,
                         "kernel_ms": kernel_ms, "stage_a_ms": float(np.median(ams)), "algorithmic_bytes": abytes,
                         "frac_of_whole_step": abytes / (total_ms / args.steps * 1e-3) / 1e9 / peak,
                         "note": "achieved = the step's algorithmic bytes / the scoring kernels' duration; the kernel is bound by the FP32/ALU pipes (about 13 instructions per (pod,node,GPU) evaluation), not by HBM: DESIGN.md 4.2 states why 50 % of HBM is out of reach without (app, SLO) duplicates"},
        }
        if extras:
            line["extras"] = extras
        if completion:
            line["completion"] = completion
        if config4:
            line["config4"] = config4
        if e2e:
            line["e2e"] = e2e
        if e2e_dec:
            line["e2e_decisions"] = e2e_dec
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"], _ = cpu_reference_leg(preds, cluster, pods)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
