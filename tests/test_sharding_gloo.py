"""Multi-rank host logic (basebandboard_b200/sharding.py) on CPU: world_size 2 and 3 over gloo.

The CUDA engine is replaced by a stand-in that answers each shard's request with the CPU oracle,
so what is tested is the partitioning, the detector-state hand-off and the final reductions: the
sharded result must equal the one-shot oracle result exactly."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleEngine:
    """Same method names and argument meaning as basebandboard_b200.lib.Context."""

    def __init__(self):
        from oracle import oracle
        from basebandboard_b200 import recurrences
        self.O, self.rec = oracle, recurrences
        self.n = 256

    def clt_hist(self, init, first, count, model=0):
        r = self.O.clt_run(self.rec.packed(self.n), init, first, count, model=model)
        return r["hist"], r["sum"], r["sumsq"]

    def prbs_spec_state(self, k, first_bit, m=0, seed=1):
        from basebandboard_b200 import lib
        return lib.prbs_spec_state(k, first_bit, m, seed)    # host arithmetic of the product library

    def prbs_detect(self, k, nbits, m=0, seed=1, first_bit=0, det=None):
        ctr, _, _ = self.O.prbs_detect(k, nbits, m=m, seed=seed, first_bit=first_bit, det=det)
        return ctr

    def prbs_slicer_detect(self, k, nbits, init=1, amplitude=256, noise_var=8, spb=8, delay=0, first_bit=0, det=None):
        """tx.py:70-81 + rx.py:28-46 composed from the oracle's pieces, as tests/test_gpu_prbs.py does."""
        O = self.O
        g = O.clt_run(self.rec.packed(self.n), init, first_bit * spb, nbits * spb, model=0, want=("out8",))["out8"]
        tx, _ = O.prbs_bits(k, nbits, state=O.lib().orc_prbs_advance(k, 1, first_bit))
        rx = O.slicer_rx_bits(tx, g, nbits, amplitude, noise_var, spb, delay)
        ctr, _ = O.det_run_bits(k, rx, nbits, tx_words=tx, det=det)
        return ctr

    def rnghunt_sweep(self, n, seed, first, count, k=16, stride=1):
        counts = np.zeros(5, dtype=np.uint64)
        best = []
        for i in range(count):
            st, c = self.O.rnghunt_sweep(n, seed, first + i * stride, 1)
            counts += c
            if st[0] == 4:
                best.append(first + i * stride)
        return {"best": best[:k], "stage_counts": counts}


# quiet channel (the warm-up window gives the exact start state), and one in permanent reload storms
SLICER_CASES = {"slicer_quiet": (15, 64 * 300, 300, 8), "slicer_storm": (7, 64 * 200, 40, 15)}


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from basebandboard_b200 import sharding
    from oracle import oracle as O
    eng = OracleEngine()
    out = {}
    out["clt"] = sharding.clt_hist_sharded(eng, 1, 1000, 100003, 0)
    out["clt"] = (out["clt"][0].tolist(), out["clt"][1], out["clt"][2])
    # low error rate: the locked-state guess is right; heavy error rate: shards are re-run
    for name, (k, nbits, m) in {"det_low": (31, 50000, 10), "det_heavy": (7, 30001, 1),
                                "det_mid": (15, 40000, 3)}.items():
        ctr, final, runs = sharding.prbs_detect_sharded(eng, k, nbits, m=m, seed=1, make_det=O.DetState)
        out[name] = (ctr, final, runs)
    for name, (k, nbits, amp, nv) in SLICER_CASES.items():
        out[name] = sharding.prbs_slicer_detect_sharded(eng, k, nbits, amplitude=amp, noise_var=nv, spb=4, delay=1,
                                                        first_bit=128, make_det=O.DetState)
    r = sharding.rnghunt_sweep_sharded(eng, 16, 1, 100, 900, k=5)
    out["hunt"] = (r["best"], r["stage_counts"].tolist())
    out["stream"] = sharding.stream_shard(128, 1000003)
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_results_equal_one_shot_oracle(orc, world):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0

    from basebandboard_b200 import recurrences
    one = orc.clt_run(recurrences.packed(256), 1, 1000, 100003, model=0)
    want_clt = (one["hist"].tolist(), one["sum"], one["sumsq"])
    stages, counts = orc.rnghunt_sweep(16, 1, 100, 900, threads=4)
    want_best = [100 + int(i) for i in np.nonzero(stages == 4)[0]][:5]
    any_rerun = False
    eng = OracleEngine()
    for rank in range(world):
        out = res[rank]
        assert out["clt"] == want_clt
        for name, (k, nbits, m) in {"det_low": (31, 50000, 10), "det_heavy": (7, 30001, 1),
                                    "det_mid": (15, 40000, 3)}.items():
            wctr, _, wdet = orc.prbs_detect(k, nbits, m=m, seed=1)
            ctr, final, runs = out[name]
            assert ctr == wctr, (name, rank)
            assert final == (wdet.bit_in, wdet.prbs, wdet.err_sr, wdet.reload_ctr)
            if name == "det_low":
                assert runs == 1
            any_rerun |= runs > 1
        for name, (k, nbits, amp, nv) in SLICER_CASES.items():
            det = orc.DetState()
            orc.lib().orc_det_reset(k, __import__("ctypes").byref(det))
            wctr = eng.prbs_slicer_detect(k, nbits, amplitude=amp, noise_var=nv, spb=4, delay=1, first_bit=128, det=det)
            ctr, final, runs = out[name]
            assert ctr == wctr, (name, rank)
            assert final == (det.bit_in, det.prbs, det.err_sr, det.reload_ctr)
            if name == "slicer_storm":
                assert wctr["reload_events"] > 5
        assert out["hunt"] == (want_best, counts.tolist())
    assert any_rerun          # the hand-off path really ran
    # stream shards tile the range exactly
    spans = sorted(res[r]["stream"] for r in range(world))
    assert spans[0][0] == 128 and sum(c for _, c in spans) == 1000003
    for (a, ca), (b, _) in zip(spans, spans[1:]):
        assert a + ca == b and (b - 128) % 64 == 0


def test_split_range_covers_everything():
    from basebandboard_b200.sharding import split_range
    for count in (0, 1, 7, 64, 1000, 12345):
        for world in (1, 2, 3, 8):
            parts = [split_range(10, count, r, world) for r in range(world)]
            assert parts[0][0] == 10 and sum(c for _, c in parts) == count
            for (a, ca), (b, _) in zip(parts, parts[1:]):
                assert a + ca == b
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1
