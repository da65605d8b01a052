"""Can the pass count of a pair be predicted from its first passes ("start pairs
longest-first from a cheap predictor measured in pass 1")?  Runs the bench batch with max_iters = 1..5 to
get the error after each of the first passes, and to convergence for the true pass counts; prints rank
correlations of candidate predictors with the passes still to go, and what a two-phase launch (k passes for
everybody, then the rest longest-predicted-first) would gain in the queue model of sim_queue_tail.py.
   python tests/devtools/probe_pass_predictor.py [n_pairs]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from lidar_slam_b200 import engine, synth
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from sim_queue_tail import simulate

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4955
b = synth.odometry_batch(n, seed=20, device="cuda", n_traj=1)
run = lambda mi: engine.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3, max_iters=mi,
                                      stopping_thresh=1e-4, planar_fallback=False)
full = run(2000)
passes = full.iters.cpu().numpy().astype(int)
K = 5
early = [run(k) for k in range(1, K + 1)]
err = np.stack([r.error.cpu().numpy() for r in early], axis=1)      # err[:, k-1] = error of the last pass of run k
seen = [int(r.iters.max().item()) for r in early]                   # passes a run with max_iters = k executes
print("passes executed with max_iters = 1..%d: %s" % (K, seen))
print(f"pairs {n}: passes mean {passes.mean():.2f} p90 {np.percentile(passes, 90):.0f} max {passes.max()}")


def rank_corr(x, y):
    rx, ry = np.argsort(np.argsort(x)), np.argsort(np.argsort(y))
    return float(np.corrcoef(rx, ry)[0, 1])


for k in range(2, K + 1):
    done_k = seen[k - 1]
    live = passes > done_k                  # pairs still running after the passes run k executed
    rem = passes[live] - done_k
    e = err[live]
    d_last = np.abs(e[:, k - 2] - e[:, k - 1])                 # |error change| of the last pass seen
    d_prev = np.abs(e[:, k - 3] - e[:, k - 2]) if k >= 3 else None
    feats = {"error of pass k": e[:, k - 1], "|d error| of pass k": d_last}
    if d_prev is not None:
        rate = np.clip(d_last / np.maximum(d_prev, 1e-300), 1e-6, 0.999999)
        feats["decay rate d_k / d_(k-1)"] = rate
        feats["geometric extrapolation to 1e-4"] = np.log(np.maximum(d_last, 1e-300) / 1e-4) / -np.log(rate)
    print(f"after {done_k} passes: {live.sum()} pairs still running, passes to go mean {rem.mean():.2f} max {rem.max()}")
    best_name, best_c, best_f = None, -2, None
    for name, f in feats.items():
        c = rank_corr(f, rem)
        print(f"    Spearman({name}, passes to go) = {c:+.3f}")
        if abs(c) > best_c:
            best_name, best_c, best_f = name, abs(c), f * np.sign(c)
    # queue model: phase 1 = k passes for everybody (no tail: equal work), phase 2 = the rest in some order
    ideal2 = rem.sum() / 592
    for label, order in (("index order", np.arange(len(rem))), ("true longest first", np.argsort(-rem)),
                         (f"by '{best_name}'", np.argsort(-best_f))):
        t, act = simulate(rem[order], "current")
        print(f"    phase 2 {label:42s}: makespan {t:6.1f} pass-times ({t / ideal2:.3f} x ideal), SMs active {act:.3f}")
t, act = simulate(passes, "current")
print(f"one launch, index order: makespan {t:.1f} pass-times ({t / (passes.sum() / 592):.3f} x ideal), SMs active {act:.3f}")
