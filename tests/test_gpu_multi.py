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
    assert torch.equal(p2, a)                                    # buffer reuse across calls
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
