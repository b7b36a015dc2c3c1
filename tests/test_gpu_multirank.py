"""Multi-rank correctness ON THE GPUs: the sharded paths over NCCL equal the one-GPU call.

tests/test_sharding_gloo.py drives basebandboard_b200/sharding.py over a CPU stand-in; this file launches
`python -m torch.distributed.run --nproc-per-node 2` on itself when at least two GPUs are visible (skipped
otherwise) and, in every rank, compares

  sharding.clt_hist_sharded              histogram | sum g | sum g^2 all-reduced over NCCL
  sharding.prbs_detect_sharded           detector state hand-off + re-run + counter all-reduce
  sharding.prbs_slicer_detect_sharded    the TX noise -> slicer -> detector loop
  sharding.rnghunt_sweep_sharded         interleaved candidates, all-gathered and merged best-k

with the same call on one GPU (rank's own device, whole range).  Reference behaviour being sharded:
gateware/bbb/rng.py:38-40,96-108, gateware/bbb/prbs.py:61-99, software/rnghunt/src/bin/rnghunt.rs:23-57."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker():
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from basebandboard_b200 import lib, sharding

    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    ctx = lib.Context(local)
    reset = lambda k: lib.DetState(0, 1, (1 << k) - 1, 0)     # noqa: E731  (prbs.py:62,80,90)

    # CLT-GRNG: ragged count so the shards differ in length
    ctx.lutopt_load_shipped(256)
    count = (1 << 30) + 12345
    h, s, ss = sharding.clt_hist_sharded(ctx, 1, 777, count, lib.MODEL_GATEWARE)
    h1, s1, ss1 = ctx.clt_hist(1, 777, count, lib.MODEL_GATEWARE)
    assert np.array_equal(h, h1) and (s, ss) == (s1, ss1), "clt_hist_sharded != one GPU"
    assert int(h.sum()) == count

    # PRBS detect: light errors (speculation right) and heavy errors (hand-off re-runs)
    for k, m, nbits in ((31, 10, (1 << 30) + 64 * 3), (15, 6, 1 << 26), (7, 2, (1 << 22) + 17)):
        ctr, final, runs = sharding.prbs_detect_sharded(ctx, k, nbits, m=m, seed=1)
        det = reset(k)
        one = ctx.prbs_detect(k, nbits, m=m, seed=1, det=det)
        assert ctr == one, ("prbs_detect_sharded", k, m, ctr, one)
        assert final == sharding._det_tuple(det), ("prbs_detect_sharded end state", k, m)

    # TX noise -> slicer -> detector
    nb = 1 << 24
    ctr, final, runs = sharding.prbs_slicer_detect_sharded(ctx, 15, nb, init=1, amplitude=200, noise_var=15, spb=8,
                                                           delay=5)
    det = reset(15)
    one = ctx.prbs_slicer_detect(15, nb, init=1, amplitude=200, noise_var=15, spb=8, delay=5, det=det)
    assert ctr == one and final == sharding._det_tuple(det), ("prbs_slicer_detect_sharded", ctr, one)
    assert ctr["injected"] > 0

    # rnghunt: merged ranking independent of the number of ranks
    r = sharding.rnghunt_sweep_sharded(ctx, 192, 1, 0, 1 << 18, k=16)
    r1 = ctx.rnghunt_sweep(192, 1, 0, 1 << 18, k=16)
    assert r["best"] == r1["best"] and len(r["best"]) == 16, ("rnghunt best", r["best"], r1["best"])
    assert np.array_equal(r["stage_counts"], r1["stage_counts"])

    dist.barrier()
    if rank == 0:
        print("MULTIRANK OK world=%d" % world, flush=True)
    ctx.close()
    dist.destroy_process_group()


@pytest.mark.gpu
def test_sharded_paths_over_nccl_equal_one_gpu():
    import torch
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs (one process per GPU over NCCL)")
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__), "--worker"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900, cwd=ROOT)
    assert res.returncode == 0 and "MULTIRANK OK world=2" in res.stdout, res.stdout[-4000:]


if __name__ == "__main__":
    if "--worker" in sys.argv:
        _worker()
