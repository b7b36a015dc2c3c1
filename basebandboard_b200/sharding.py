"""Multi-GPU sharding of the three paths: one process per GPU, no data-path collective.

Every path is cut into independent units; the only exchange is one small
reduction / gather of results at the end (torch.distributed: NCCL on the GPU box, gloo in the CPU
tests):

  CLT-GRNG fused      rank r owns sample indices [first + r*per, ...) -> all-reduce(sum) of
                      u64[n+3] = histogram | sum g | sum g^2
  PRBS detect         rank r owns a contiguous bit range and starts from the locked detector state
                      (bbb_prbs_spec_state); (start, end, counters) are all-gathered and a rank whose
                      left neighbour ended in a different state re-runs its range from that state, until
                      every rank started where its neighbour ended; the gathered counters are summed
  rnghunt sweep       rank r owns candidates first + r, first + r + world, ... -> one all-gather of the
                      k smallest accepted indices and the 5 stage counts per rank, merged / summed
  materialised streams  disjoint index ranges, each rank writes its own HBM; nothing is exchanged

`engine` is a basebandboard_b200.lib.Context (the CUDA library).  The functions only use its
public methods, so the CPU tests drive the same logic with a stand-in engine over the oracle.
"""
import numpy as np


def split_range(first, count, rank, world):
    """Contiguous near-equal split of [first, first+count): returns (first_r, count_r)."""
    per, extra = divmod(count, world)
    lo = first + rank * per + min(rank, extra)
    return lo, per + (1 if rank < extra else 0)


def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


def _device():
    import torch
    dist = _dist()
    if dist is not None and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def _all_reduce_sum_u64(values):
    """Sum of uint64 vectors over ranks (two's complement wrap-around, carried as int64)."""
    import torch
    dist = _dist()
    a = np.ascontiguousarray(values, dtype=np.uint64)
    if dist is None or dist.get_world_size() == 1:
        return a.copy()
    t = torch.from_numpy(a.view(np.int64).copy()).to(_device())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.cpu().numpy().view(np.uint64)


def _all_gather_i64(values):
    """[world, len(values)] int64 on every rank."""
    import torch
    dist = _dist()
    a = np.ascontiguousarray(values, dtype=np.int64)
    if dist is None or dist.get_world_size() == 1:
        return a[None, :].copy()
    t = torch.from_numpy(a.copy()).to(_device())
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return np.stack([o.cpu().numpy() for o in out])


def rank_world():
    dist = _dist()
    return (dist.get_rank(), dist.get_world_size()) if dist is not None else (0, 1)


# ------------------------------------------------------------------------------------------------
def clt_hist_sharded(engine, init, first, count, model=0):
    """Histogram and moments of g[first, first+count) computed by all ranks together."""
    rank, world = rank_world()
    lo, cnt = split_range(first, count, rank, world)
    hist, s, ss = engine.clt_hist(init, lo, cnt, model)
    acc = np.concatenate([hist.astype(np.uint64), np.array([s], dtype=np.int64).view(np.uint64),
                          np.array([ss], dtype=np.uint64)])
    acc = _all_reduce_sum_u64(acc)
    return acc[:-2].copy(), int(acc[-2:-1].view(np.int64)[0]), int(acc[-1])


_KEYS = ("err", "err_valid", "reload_events", "reload_cycles", "injected")


def _det_tuple(det):
    return (int(det.bit_in), int(det.prbs), int(det.err_sr), int(det.reload_ctr))


def prbs_detect_sharded(engine, k, nbits, m=0, seed=1, first_bit=0, make_det=None):
    """Counters of the detector over bits [first_bit, first_bit+nbits) cut contiguously over the
    ranks.  `make_det(bit_in, prbs, err_sr, reload_ctr)` builds the engine's detector-state object
    (default: basebandboard_b200.lib.DetState).  Returns (counters, final detector state tuple,
    how many times this rank ran its range)."""
    rank, world = rank_world()
    if make_det is None:
        from .lib import DetState as make_det
    # cut on 64-bit word boundaries relative to first_bit so that every shard but the last is whole words
    words = (nbits + 63) // 64
    wlo, wcnt = split_range(0, words, rank, world)
    lo = first_bit + 64 * wlo
    cnt = max(0, min(64 * wcnt, first_bit + nbits - lo))
    if rank == 0:
        start = (0, 1, (1 << k) - 1, 0)               # reset values, prbs.py:62,80,90
    else:
        start = _det_tuple(engine.prbs_spec_state(k, lo, m, seed)) if lo >= 64 else (0, 1, (1 << k) - 1, 0)
    runs = 0
    ran_from = None
    while True:
        if ran_from != start:                             # only a rank whose start state changed runs again
            det = make_det(*start)
            ctr = engine.prbs_detect(k, cnt, m=m, seed=seed, first_bit=lo, det=det) if cnt else \
                {"err": 0, "err_valid": 0, "reload_events": 0, "reload_cycles": 0, "injected": 0}
            runs += 1
            end = _det_tuple(det) if cnt else start
            ran_from = start
        # one exchange per round: (start, end, counters) of every rank; when the chain is consistent the counters of
        # that same table are the result, so no separate reduction follows
        table = _all_gather_i64(list(start) + list(end) + [ctr[key] for key in _KEYS])     # [world, 13]
        # the state each rank should have started from: its left neighbour's end state
        wrong = [r for r in range(1, world) if tuple(table[r][:4]) != tuple(table[r - 1][4:8])]
        if not wrong:
            break
        if rank in wrong:
            start = tuple(int(v) for v in table[rank - 1][4:8])
    tot = table[:, 8:].astype(np.uint64).sum(axis=0, dtype=np.uint64)
    final = tuple(int(v) for v in table[world - 1][4:8])
    return dict(zip(_KEYS, (int(v) for v in tot))), final, runs


def _loop_sharded(run, spec_state, k, nbits, first_bit, warm_bits=1024):
    """Shared shape of the sharded detector loops: contiguous word-aligned shards, a speculative start state per
    rank, all-gather of (start, end) states, re-run of the ranks whose start differs from the left neighbour's
    end, until consistent (= the sequential result, by induction from rank 0).  `run(lo, cnt, start)` returns
    (counters dict, end state tuple).  Speculation for rank > 0: run the `warm_bits` bits before the shard from
    the error-free locked state -- exact unless the detector was reloading just there."""
    rank, world = rank_world()
    words = (nbits + 63) // 64
    wlo, wcnt = split_range(0, words, rank, world)
    lo = first_bit + 64 * wlo
    cnt = max(0, min(64 * wcnt, first_bit + nbits - lo))
    reset = (0, 1, (1 << k) - 1, 0)                   # reset values, prbs.py:62,80,90
    zero = {"err": 0, "err_valid": 0, "reload_events": 0, "reload_cycles": 0, "injected": 0}
    if rank == 0 or lo - first_bit < warm_bits + 64 or lo - warm_bits < 64:
        # short ranges: the chain from the true start is as cheap as a guess
        start = reset if lo == first_bit else run(first_bit, lo - first_bit, reset)[1]
    else:
        start = run(lo - warm_bits, warm_bits, _det_tuple(spec_state(k, lo - warm_bits)))[1]
    runs = 0
    ran_from = None
    while True:
        if ran_from != start:                             # only a rank whose start state changed runs again
            ctr, end = run(lo, cnt, start) if cnt else (zero, start)
            runs += 1
            ran_from = start
        table = _all_gather_i64(list(start) + list(end) + [ctr[key] for key in _KEYS])     # [world, 13]
        wrong = [r for r in range(1, world) if tuple(table[r][:4]) != tuple(table[r - 1][4:8])]
        if not wrong:
            break
        if rank in wrong:
            start = tuple(int(v) for v in table[rank - 1][4:8])
    tot = table[:, 8:].astype(np.uint64).sum(axis=0, dtype=np.uint64)
    final = tuple(int(v) for v in table[world - 1][4:8])
    return dict(zip(_KEYS, (int(v) for v in tot))), final, runs


def prbs_slicer_detect_sharded(engine, k, nbits, init=1, amplitude=256, noise_var=8, spb=8, delay=0, first_bit=0,
                               make_det=None):
    """The TX noise word -> slicer -> PRBSErrorDetector loop (tx.py:70-81, rx.py:28-46) over bits
    [first_bit, first_bit+nbits) cut contiguously over the ranks; PRBS and noise are indexed by absolute bit,
    so a shard is just the engine's call with its own first_bit.  Returns (counters, final state, runs)."""
    if make_det is None:
        from .lib import DetState as make_det

    def run(lo, cnt, start):
        det = make_det(*start)
        ctr = engine.prbs_slicer_detect(k, cnt, init=init, amplitude=amplitude, noise_var=noise_var, spb=spb,
                                        delay=delay, first_bit=lo, det=det)
        return ctr, _det_tuple(det)

    return _loop_sharded(run, lambda kk, at: engine.prbs_spec_state(kk, at, 0, 1), k, nbits, first_bit)


def prbs_shaped_slicer_detect_sharded(engine, k, nbits, coeff, init=1, noise_var=8, delay=4, first_bit=0, make_det=None):
    """Same for the pulse-shaped channel (bitshaper.py, bbb_prbs_shaped_slicer_detect)."""
    if make_det is None:
        from .lib import DetState as make_det

    def run(lo, cnt, start):
        det = make_det(*start)
        ctr = engine.prbs_shaped_slicer_detect(k, cnt, coeff, init=init, noise_var=noise_var, delay=delay,
                                               first_bit=lo, det=det)
        return ctr, _det_tuple(det)

    return _loop_sharded(run, lambda kk, at: engine.prbs_spec_state(kk, at, 0, 1), k, nbits, first_bit)


def rnghunt_sweep_sharded(engine, n, seed, first, count, k=16):
    """Sweep candidates first .. first+count-1, rank r taking every world-th one; returns the k
    smallest accepted candidate indices overall and the summed stage counts."""
    rank, world = rank_world()
    mine = len(range(rank, count, world))
    res = engine.rnghunt_sweep(n, seed, first + rank, mine, k=k, stride=world) if mine else \
        {"best": [], "stage_counts": np.zeros(5, dtype=np.uint64)}
    best = list(res["best"])[:k]
    pad = np.full(k + 1 + 5, -1, dtype=np.int64)       # one exchange: count | k indices | 5 stage counts
    pad[0] = len(best)
    pad[1:1 + len(best)] = best
    pad[k + 1:] = np.asarray(res["stage_counts"], dtype=np.uint64).view(np.int64)
    allb = _all_gather_i64(pad)
    merged = sorted(int(v) for row in allb for v in row[1:1 + int(row[0])])[:k]
    counts = allb[:, k + 1:].astype(np.uint64).sum(axis=0, dtype=np.uint64)
    return {"best": merged, "stage_counts": counts}


def stream_shard(first, count, align=64):
    """This rank's disjoint slice of a materialised stream, aligned to `align` units."""
    rank, world = rank_world()
    units = (count + align - 1) // align
    ulo, ucnt = split_range(0, units, rank, world)
    lo = first + ulo * align
    return lo, max(0, min(ucnt * align, first + count - lo))
