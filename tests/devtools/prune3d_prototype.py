"""CPU prototype (numpy) of block pruning for the general 3-D kernel -- planning evidence for the next
round, no device code involved.  Object clouds have no scan order, so source and target are sorted
along a Morton curve first; rows of 32 sorted source points and blocks of 32 sorted target points then
have compact boxes and the bound of csrc/icp2d.cuh (lb(Q, B) with the candidate distance's own
monotone float32 arithmetic, here on three axes) prunes.  The script runs the float64 ICP of the
oracle on a yaw-sweep problem and reports, pass by pass, which share of the (row, block) tiles a
pruned sweep would still evaluate (seeds = two nearest boxes, then every block whose bound is within
the largest current nearest-neighbour distance of the row -- the float64 stand-in of the kernel's
threshold), and checks that the pruned search returns the exhaustive nearest neighbours.

    python tests/devtools/prune3d_prototype.py [n_source_points]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from lidar_slam_b200 import synth
from oracle import icp_oracle as orc


def morton_order(p, bits=10):
    lo, hi = p.min(0), p.max(0)
    g = np.minimum(((p - lo) / np.maximum(hi - lo, 1e-300) * (1 << bits)).astype(np.int64), (1 << bits) - 1)
    key = np.zeros(len(p), np.int64)
    for b in range(bits):
        for a in range(p.shape[1]):
            key |= ((g[:, a] >> b) & 1) << (p.shape[1] * b + a)
    return np.argsort(key, kind="stable")


def boxes(p, size=32):
    n = (len(p) + size - 1) // size
    lo = np.stack([p[k * size:(k + 1) * size].min(0) for k in range(n)])
    hi = np.stack([p[k * size:(k + 1) * size].max(0) for k in range(n)])
    return lo, hi


def pruned_nn(q, t, tlo, thi):
    """Nearest neighbours of the rows of q in t with block pruning; returns (indices, tiles swept)."""
    nblk = len(tlo)
    out = np.empty(len(q), np.int64)
    swept = 0
    for r0 in range(0, len(q), 32):
        row = q[r0:r0 + 32]
        qlo, qhi = row.min(0), row.max(0)
        gap = np.maximum(np.maximum(tlo - qhi, qlo - thi), 0.0)
        lb = (gap ** 2).sum(1)
        order = np.argsort(lb, kind="stable")
        best_d = np.full(len(row), np.inf)
        best_i = np.zeros(len(row), np.int64)
        done = np.zeros(nblk, bool)

        def sweep(k):
            nonlocal best_d, best_i
            blk = t[k * 32:(k + 1) * 32]
            d = ((row[:, None, :] - blk[None, :, :]) ** 2).sum(-1)
            j = d.argmin(1)
            dj = d[np.arange(len(row)), j]
            better = (dj < best_d) | ((dj == best_d) & (k * 32 + j < best_i))
            best_d = np.where(better, dj, best_d)
            best_i = np.where(better, k * 32 + j, best_i)
            done[k] = True
        for k in order[:2]:
            sweep(k)
        bound = best_d.max()
        for k in range(nblk):
            if not done[k] and lb[k] <= bound:
                sweep(k)
        swept += int(done.sum())
        out[r0:r0 + 32] = best_i
    return out, swept


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
    src = synth.drill_cloud(n, seed=1)
    tgt = synth.partial_view(src, seed=2)
    inits = synth.yaw_sweep_inits(src, tgt, 24)
    ps, pt = morton_order(src), morton_order(tgt)
    src_s, tgt_s = src[ps], tgt[pt]
    tlo, thi = boxes(tgt_s)
    rows, nblk = (len(src) + 31) // 32, len(tlo)
    print(f"source {len(src)} points ({rows} rows), target {len(tgt)} points ({nblk} blocks), Morton-sorted")
    for yaw in (0, 3, 12):
        T = inits[yaw].copy()
        shares = []
        for it in range(40):
            q = src_s @ T[:3, :3].T + T[:3, 3]
            nn, swept = pruned_nn(q, tgt_s, tlo, thi)
            d = ((q[:, None, :] - tgt_s[None, :, :]) ** 2).sum(-1) if len(src) <= 8000 else None
            if d is not None:
                assert np.array_equal(nn, d.argmin(1)), "pruned search differs from the exhaustive one"
            shares.append(swept / (rows * nblk))
            dT = orc.rigid_fit(q, tgt_s[nn])
            err = orc.alignment_error(q, tgt_s[nn], normalize=True)
            T = dT @ T
            if err < 1e-3 or (it and abs(prev - err) < 1e-4):       # the reference's stop tests (icp.py:172-178)
                break
            prev = err
        print(f"yaw init {yaw:2d}: {len(shares)} passes, evaluated share of the tiles per pass: first {shares[0]:.3f}, "
              f"median {np.median(shares):.3f}, last {shares[-1]:.3f}, mean {np.mean(shares):.3f}")


if __name__ == "__main__":
    main()
