"""Multi-GPU exchange on real devices (skipped unless >= 2 GPUs are visible): the fused peer-store
epilogue delivers exactly the rows the NCCL all-gather path delivers, in pair order, on every rank."""
import os
import subprocess
import sys
import textwrap

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, %r)
    import torch, torch.distributed as dist
    from lidar_slam_b200 import engine, synth
    rank = int(os.environ["RANK"]); torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    b = synth.odometry_batch(37, seed=11, device="cuda")        # 37 pairs: ragged shards + padding rows
    a, ra = engine.run_icp_sharded(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3, exchange="nccl")
    a = a.clone()
    p, rp = engine.run_icp_sharded(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3, exchange="peer")
    assert a.shape == p.shape == (37, 4), (a.shape, p.shape)
    assert torch.equal(a, p), (a - p).abs().max()
    full = engine.run_icp_batch(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3)
    assert torch.equal(p, engine.pack_poses(full.T, full.error))
    p2, _ = engine.run_icp_sharded(b["scans"], b["counts"], b["pairs"], b["init"], epsilon=1e-3, exchange="peer")
    assert torch.equal(p2, a)                                    # the other row buffer (parity 1)
    assert torch.equal(p, a)                                     # ... and the first one is still intact
    for step in range(5):                                        # buffers and arrival flags reused, no barrier between
        c = synth.odometry_batch(23 + step, seed=30 + step, device="cuda")
        want, _ = engine.run_icp_sharded(c["scans"], c["counts"], c["pairs"], c["init"], epsilon=1e-3, exchange="nccl")
        got, _ = engine.run_icp_sharded(c["scans"], c["counts"], c["pairs"], c["init"], epsilon=1e-3, exchange="peer")
        assert torch.equal(got, want), step
    for pg in engine._PEER_BUFFERS.values():
        pg.check()
    # the general (3-D) kernel has the same epilogue
    import numpy as np
    src = synth.drill_cloud(900, seed=1); tgt = synth.partial_view(src, seed=2)
    pool, counts = engine.pack_scans([src, tgt])
    inits = torch.tensor(np.stack(synth.yaw_sweep_inits(src, tgt, 24)[:7])).cuda()
    prs = torch.tensor([[0, 1]] * 7, dtype=torch.int32).cuda()
    want, _ = engine.run_icp_sharded(pool, counts, prs, inits, epsilon=1e-3, normalize_error=True, exchange="nccl")
    got, _ = engine.run_icp_sharded(pool, counts, prs, inits, epsilon=1e-3, normalize_error=True, exchange="peer")
    assert torch.equal(got, want)
    dist.barrier(); dist.destroy_process_group()
    print("rank", rank, "ok")
""") % ROOT


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_fused_peer_exchange_equals_nccl_all_gather(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert out.stdout.count("ok") >= 2
