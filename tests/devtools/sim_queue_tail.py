"""Why the planar kernel's SMs are active 0.88 of a 4955-pair launch, and what would change it.

A list-scheduling model of the persistent-CTA work queue: 148 SMs x 4 resident CTAs pull pairs in
index order; a pair costs its pass count (drawn from a log-normal fitted to the bench histogram:
mean 11.6, p50 11, p90 18, p99 27, max ~46, or read from an .npy of real pass counts); a CTA whose
SM neighbours have retired runs a little faster (latency-bound kernel: 1.0 / 1.15 / 1.35 / 1.6x with
4 / 3 / 2 / 1 resident).  Policies:
  current   pairs in index order, one CTA per pair until it converges
  helpers   once the queue is empty, the idle warps of an SM join the pairs still running there
  requeue   a pair is parked after Q passes and re-queued at the back (time slicing)
  lpt       longest first (needs an oracle for the pass counts: the bound, not an option)

    python tests/devtools/sim_queue_tail.py [passes.npy]
"""
import sys

import numpy as np

SPEED = {4: 1.0, 3: 1.15, 2: 1.35, 1: 1.6, 0: 0.0}


def draw(n, seed):
    rng = np.random.default_rng(seed)
    return np.clip(np.round(np.exp(rng.normal(np.log(10.6), 0.42, n))), 2, 60).astype(int)


def simulate(passes, policy, n_sm=148, slots=4, dt=0.05, quantum=16):
    queue = [(int(p), 0) for p in (sorted(passes, reverse=True) if policy == "lpt" else passes)]
    head = 0
    rem = np.zeros((n_sm, slots))
    left = np.zeros((n_sm, slots), dtype=int)          # passes left after this quantum (requeue)
    def pop(sm, sl):
        nonlocal head
        if head < len(queue):
            p, _ = queue[head]; head += 1
            if policy == "requeue" and p > quantum:
                rem[sm, sl], left[sm, sl] = quantum, p - quantum
            else:
                rem[sm, sl], left[sm, sl] = p, 0
        else:
            rem[sm, sl] = 0
    for k in range(n_sm * slots):
        pop(k % n_sm, k // n_sm)
    t = busy = 0.0
    while True:
        active = rem > 1e-9
        a = active.sum(1)
        if head >= len(queue) and a.sum() == 0:
            break
        for sm in np.nonzero(a)[0]:
            per = SPEED[a[sm]]
            if policy == "helpers" and head >= len(queue) and a[sm] < slots:
                per = min(max(3.4 / a[sm], per), 2.8)
            rem[sm, active[sm]] -= per * dt
        busy += (a > 0).sum() * dt
        t += dt
        for sm, sl in np.argwhere((rem <= 1e-9) & active):
            if left[sm, sl] > 0:
                queue.append((int(left[sm, sl]), 1)); left[sm, sl] = 0
            pop(sm, sl)
    return t, busy / (t * n_sm)


if __name__ == "__main__":
    sets = [np.load(sys.argv[1])] if len(sys.argv) > 1 else [draw(4955, s) for s in range(3)]
    for p in sets:
        ideal = p.sum() / (148 * 4)
        print(f"pairs {len(p)} mean {p.mean():.2f} p90 {np.percentile(p, 90):.0f} p99 {np.percentile(p, 99):.0f} "
              f"max {p.max()}  ideal makespan {ideal:.1f} pass-times")
        for policy in ("current", "helpers", "requeue", "lpt"):
            t, act = simulate(p, policy)
            print(f"  {policy:8s} makespan {t:6.1f}  ({t / ideal:.3f} x ideal)  SM active/elapsed {act:.3f}")
