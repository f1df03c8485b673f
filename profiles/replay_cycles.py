"""BASELINE config 5 stand-in: replay of scheduling cycles through the plugin surface.

No Go toolchain / kube-scheduler framework exists here, so the C++ host mirror replays the
framework's call sequence for one pod per cycle: PreScore (one GPU call scoring the pod against
every node), Score for every node (row lookup), NormalizeScore on the list, pick the best node.
Reports p50/p99 per-cycle latency on the host clock.

    python profiles/replay_cycles.py [--nodes 5000] [--arrivals 10000]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=5000)
    ap.add_argument("--arrivals", type=int, default=10000)
    args = ap.parse_args()
    pkg = entry.load_package()
    from k8s_gpu_scheduler_b200 import host

    rng = np.random.default_rng(1)
    A = 256
    apps = [f"app_{i:03d}" for i in range(A)]
    cols = [f"{p}P_{m}" for m in ("A30", "V100") for p in (1, 2, 4, 8)]
    preds = pkg.synth.make_predictions(A)
    b = host.SnapshotBuilder()
    b.set_configurations(apps, cols, preds.conf)
    b.set_interference([f"{a}_{m}" for m in ("A30", "V100") for a in apps], apps,
                       np.concatenate([preds.interf[0], preds.interf[1]], axis=0))
    names = []
    for n in range(args.nodes):
        kind = rng.random()
        name = (f"k8s-a30-{n}" if kind < 0.5 else f"gpu-node-{n}" if kind < 0.95 else f"cpu-node-{n}")
        uu = [f"GPU-{n}-{g}" for g in range(8)]
        b.add_node(name, "[" + ",".join(f'"{u}"' for u in uu) + "]")
        names.append(name)
        for g in range(8):
            for _ in range(min(4, rng.poisson(1.5))):
                b.add_pod(f"{apps[rng.integers(0, A)].replace('_', '-')}-{rng.integers(1 << 30):x}", "Running", uu[g],
                          f"{rng.uniform(10, 600):.2f}")
    sch = pkg.Scheduler(0)
    plugin = host.Plugin(sch, b)
    scores = np.empty(args.nodes, dtype=np.int64)
    pods = [(f"{apps[rng.integers(0, A)].replace('_', '-')}-{i:05d}", f"{rng.uniform(10, 600):.2f}") for i in range(args.arrivals)]
    for name, slo in pods[:50]:   # warm-up
        plugin.pre_score(name, slo)
    lat = np.empty(args.arrivals)
    for i, (name, slo) in enumerate(pods):
        t0 = time.perf_counter()
        rc, _ = plugin.pre_score(name, slo)
        plugin.score_all(args.nodes, scores)
        norm = plugin.normalize_score(scores)
        best = int(np.argmax(norm))
        lat[i] = time.perf_counter() - t0
        assert rc == 0 and best >= 0
    lat_us = np.sort(lat) * 1e6
    print(json.dumps({"config": f"replay of {args.arrivals} pod arrivals x {args.nodes} nodes x 8 GPUs through the C++ plugin mirror "
                                "(PreScore on the GPU + Score for every node + NormalizeScore), one pod per cycle",
                      "p50_us": float(lat_us[len(lat_us) // 2]), "p99_us": float(lat_us[int(len(lat_us) * 0.99)]),
                      "mean_us": float(lat_us.mean()), "cycles_per_s": float(1e6 / lat_us.mean())}))
    plugin.close()
    b.close()
    sch.close()


if __name__ == "__main__":
    main()
