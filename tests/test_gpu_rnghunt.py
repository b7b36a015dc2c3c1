"""GPU parity: the rnghunt worker loop through the C ABI vs the CPU oracle (identical verdicts).

Reference behaviour: software/rnghunt/src/bin/rnghunt.rs:23-57 over binary_matrix.rs:53-101,
berlekamp_massey.rs:5-31 and binary_polynomial.rs:136-212.  The oracle (oracle/gf2.c) restates those
line by line and is pinned by the crate's 33 unit tests and the 7 shipped matrices
(tests/test_oracle_gf2.py); the candidate stream is this build's own definition, restated
independently in oracle/rnghunt_ref.c."""
import os

import numpy as np
import pytest

from basebandboard_b200 import recurrences

pytestmark = pytest.mark.gpu

STAGE_DEGREE, STAGE_FILTER, STAGE_ORDER, STAGE_FACTOR, STAGE_PRIMITIVE = range(5)


@pytest.fixture(scope="module")
def ctx():
    from basebandboard_b200 import lib
    c = lib.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("n", [8, 16, 24, 31, 32, 33, 64, 100, 128, 192, 255, 256, 320, 512])
def test_candidate_matrix_matches_oracle(ctx, orc, n):
    """BinaryMatrix::random with the build's counter-based draws: same matrix, BinaryMatrix layout."""
    for seed, c in ((1, 0), (1, 12345), (7, 1 << 40)):
        got = ctx.rnghunt_candidate(n, seed, c)
        want = orc.BinaryMatrix.candidate(n, seed, c).data
        assert np.array_equal(got, want), (n, seed, c)


def test_candidate_row_and_column_weights(ctx):
    """binary_matrix.rs:195-209 (test_random): row weights come from [3,4,...]; columns are balanced."""
    n = 192
    w = ctx.rnghunt_candidate(n, 3, 99).reshape(n, 3)
    rows = np.zeros(n, dtype=int)
    cols = np.zeros(n, dtype=int)
    for c in range(n):
        for r in range(n):
            if (int(w[c, r // 64]) >> (63 - r % 64)) & 1:
                rows[r] += 1
                cols[c] += 1
    assert set(rows.tolist()) <= {3, 4}
    assert cols.max() - cols.min() <= 1


@pytest.mark.parametrize("n,count", [(8, 2000), (16, 4000), (32, 4000), (33, 1500), (64, 3000), (100, 1000),
                                     (128, 1000), (192, 600), (256, 300), (320, 64), (512, 40)])
def test_sweep_stages_match_oracle(ctx, orc, n, count):
    got = ctx.rnghunt_sweep(n, 1, 0, count, k=8, want_stages=True)
    stages, counts = orc.rnghunt_sweep(n, 1, 0, count, threads=os.cpu_count() or 1)
    assert np.array_equal(got["stages"], stages)
    assert np.array_equal(got["stage_counts"], counts)
    hits = [int(i) for i in np.nonzero(stages == STAGE_PRIMITIVE)[0]]
    assert got["best"] == hits[:8]
    if n <= 64:
        assert len(hits) > 0 and counts[STAGE_ORDER] > 0 and counts[STAGE_FILTER] > 0


def test_sweep_small_n_many_hits(ctx, orc):
    """n = 8: almost 1 % of the candidates are accepted, so a batch returns more than a thousand hits (and their
    polynomials) through the pinned hit list."""
    n, seed, count = 8, 4, 150000
    stages, counts = orc.rnghunt_sweep(n, seed, 0, count, threads=os.cpu_count() or 1)
    hits = [int(i) for i in np.nonzero(stages == STAGE_PRIMITIVE)[0]]
    assert len(hits) > 1000
    got = ctx.rnghunt_sweep(n, seed, 0, count, k=count)
    assert got["best"] == hits and np.array_equal(got["stage_counts"], counts)
    gp = ctx.rnghunt_sweep(n, seed, 0, count, k=count, want_polys=True)
    assert gp["best"] == hits and all(w % 2 == 1 and 3 <= w <= 9 for w in gp["weights"])


def test_sweep_stride_offset_and_batches(ctx, orc):
    """Interleaved candidate indices (the multi-GPU partition) and small device batches."""
    n, seed = 32, 5
    stages, _ = orc.rnghunt_sweep(n, seed, 1000, 6000, threads=os.cpu_count() or 1)
    ctx.set_tuning("hunt_batch", 257)
    for rank in range(3):
        cnt = len(range(rank, 6000, 3))
        got = ctx.rnghunt_sweep(n, seed, 1000 + rank, cnt, k=4, stride=3, want_stages=True)
        assert np.array_equal(got["stages"], stages[rank::3])
        hits = [1000 + int(i) for i in np.nonzero(stages == STAGE_PRIMITIVE)[0] if i % 3 == rank]
        assert got["best"] == hits[:4]
    ctx.set_tuning("hunt_batch", 0)


@pytest.mark.parametrize("n", [16, 32, 64, 128, 192, 256, 512])
def test_eval_shipped_matrices_are_accepted(ctx, orc, n):
    """The archived rnghunt outputs (software/rnghunt/matrices/N == rng_recurrences.nN) pass the
    reference's acceptance test; the Berlekamp-Massey polynomial equals the oracle's."""
    m = orc.BinaryMatrix.from_packed(recurrences.packed(n))
    st, poly = ctx.rnghunt_eval(n, m.data, want_poly=True)
    ost, opoly = m.evaluate(want_poly=True)
    assert ost == STAGE_PRIMITIVE and int(st[0]) == STAGE_PRIMITIVE
    pw = (n + 1 + 63) // 64
    assert np.array_equal(poly[0], opoly[:pw])


@pytest.mark.parametrize("n", [16, 64, 192])
def test_eval_batch_matches_oracle_with_polynomials(ctx, orc, n):
    """bbb_rnghunt_eval on caller-supplied matrices: generated candidates, a singular matrix, the
    identity, and a row weight above 4 (the dense kernel does not assume rnghunt's weights)."""
    mats = [orc.BinaryMatrix.candidate(n, 2, c) for c in range(300 if n < 192 else 80)]
    wpc = (n + 63) // 64
    ident = np.zeros(n * wpc, dtype=np.uint64)
    heavy = np.zeros(n * wpc, dtype=np.uint64)
    for r in range(n):
        ident[r * wpc + r // 64] |= np.uint64(1) << np.uint64(63 - r % 64)
        for c in {r, (r + 1) % n, (3 * r + 2) % n, (5 * r + 1) % n, (7 * r + 3) % n, (r * r + 1) % n}:
            heavy[c * wpc + r // 64] |= np.uint64(1) << np.uint64(63 - r % 64)
    mats += [orc.BinaryMatrix.from_words(n, n, np.zeros(n * wpc, dtype=np.uint64)),
             orc.BinaryMatrix.from_words(n, n, ident), orc.BinaryMatrix.from_words(n, n, heavy)]
    data = np.concatenate([m.data for m in mats])
    st, poly = ctx.rnghunt_eval(n, data, want_poly=True)
    pw = (n + 1 + 63) // 64
    for i, m in enumerate(mats):
        ost, opoly = m.evaluate(want_poly=True)
        assert int(st[i]) == ost, i
        if ost != STAGE_DEGREE:
            assert np.array_equal(poly[i], opoly[:pw]), i
        else:
            assert not poly[i].any()


def test_sweep_agrees_with_eval_of_generated_matrices(ctx):
    """The sparse tap-list front end and the dense front end see the same candidates."""
    n = 64
    got = ctx.rnghunt_sweep(n, 9, 500, 400, k=4, want_stages=True)
    data = np.concatenate([ctx.rnghunt_candidate(n, 9, 500 + i) for i in range(400)])
    st = ctx.rnghunt_eval(n, data)
    assert np.array_equal(st, got["stages"])


def test_prbs_trinomials_are_primitive_through_eval(ctx):
    """prbs.py:14: the companion matrix of every x^k + x^tap + 1 is accepted (SURVEY 8a KAT)."""
    for k, tap in ((7, 6), (9, 5), (11, 9), (15, 14), (20, 3), (23, 18), (31, 28)):
        n = k if k >= 8 else None
        if n is None:
            continue
        wpc = (n + 63) // 64
        a = np.zeros(n * wpc, dtype=np.uint64)

        def setbit(r, c):
            a[c * wpc + r // 64] |= np.uint64(1) << np.uint64(63 - r % 64)
        # Fibonacci LFSR: x'[0] = x[k-1] ^ x[tap-1]; x'[i] = x[i-1]
        setbit(0, k - 1)
        setbit(0, tap - 1)
        for i in range(1, n):
            setbit(i, i - 1)
        assert int(ctx.rnghunt_eval(n, a)[0]) == STAGE_PRIMITIVE


def test_bad_arguments(ctx):
    from basebandboard_b200 import lib
    with pytest.raises(lib.BBBError):
        ctx.rnghunt_sweep(4, 1, 0, 10)
    with pytest.raises(lib.BBBError):
        ctx.rnghunt_sweep(1024, 1, 0, 10)
    with pytest.raises(lib.BBBError):
        ctx.rnghunt_sweep(32, 1, 0, 10, stride=0)


@pytest.mark.parametrize("n", [32, 64, 192])
def test_table_and_shift_add_primitivity_kernels_agree(ctx, n):
    """hunt_check32_kernel (table-folded squaring, n % 32 == 0) vs hunt_check_kernel (shift-and-add)."""
    count = 20000 if n < 192 else 6000
    fast = ctx.rnghunt_sweep(n, 11, 0, count, k=32, want_stages=True)
    ctx.set_tuning("hunt_slow_check", 1)
    try:
        slow = ctx.rnghunt_sweep(n, 11, 0, count, k=32, want_stages=True)
    finally:
        ctx.set_tuning("hunt_slow_check", 0)
    assert np.array_equal(fast["stages"], slow["stages"]) and fast["best"] == slow["best"]
    assert fast["stage_counts"][STAGE_ORDER] > 0


@pytest.mark.parametrize("n", [8, 16, 33, 100, 128, 192, 250, 256])
def test_power_warp_and_thread_front_kernels_agree(ctx, n):
    """Three independent implementations of binary_matrix.rs:53-76 + berlekamp_massey.rs:5-31:
    hunt_front_pow_kernel (one warp per candidate, A^K applied per pass with K = 8 and K = 4, K-1 extra rows for the
    output bits in between), hunt_front_warp_kernel (one clock per pass; both widths of its generation pass) and
    hunt_front_kernel (one thread per candidate, byte state + tap lists in shared memory)."""
    count = 5000 if n < 250 else 1500
    runs = {}
    for name, knobs in (("pow8", {}), ("pow4", {"hunt_pow_k": 4}), ("warp32", {"hunt_smem_taps": 2}),
                        ("warp16", {"hunt_smem_taps": 2, "hunt_cpw": 16}), ("thread", {"hunt_smem_taps": 1})):
        for k, v in knobs.items():
            ctx.set_tuning(k, v)
        try:
            runs[name] = ctx.rnghunt_sweep(n, 21, 7, count, k=8, want_stages=True)
        finally:
            for k in knobs:
                ctx.set_tuning(k, 0)
    ref = runs["thread"]
    for name, r in runs.items():
        assert np.array_equal(r["stages"], ref["stages"]) and r["best"] == ref["best"], name
        assert np.array_equal(r["stage_counts"], ref["stage_counts"]), name


def test_power_kernel_ragged_batches_and_strides(ctx, orc):
    """Group counts that are not multiples of 32, several batches, interleaved candidates: == oracle."""
    n, seed = 64, 9
    stages, _ = orc.rnghunt_sweep(n, seed, 0, 3000, threads=os.cpu_count() or 1)
    ctx.set_tuning("hunt_batch", 1000 + 37)
    try:
        for stride, first in ((1, 0), (3, 1), (7, 5)):
            cnt = len(range(first, 3000, stride))
            got = ctx.rnghunt_sweep(n, seed, first, cnt, k=4, stride=stride, want_stages=True)
            assert np.array_equal(got["stages"], stages[first::stride])
    finally:
        ctx.set_tuning("hunt_batch", 0)


@pytest.mark.parametrize("n", [16, 64, 192])
def test_sweep_polys_are_the_oracle_polynomials(ctx, orc, n):
    """bbb_rnghunt_sweep_polys: the characteristic polynomial of every returned hit, word for word the
    oracle's Berlekamp-Massey polynomial (berlekamp_massey.rs:5-31), and its weight (util/verify.py:19-28)."""
    count = 3000 if n < 192 else 20000
    got = ctx.rnghunt_sweep(n, 1, 0, count, k=64, want_polys=True)
    plain = ctx.rnghunt_sweep(n, 1, 0, count, k=64)
    assert got["best"] == plain["best"] and len(got["best"]) > 0
    assert np.array_equal(got["stage_counts"], plain["stage_counts"])
    pw = (n + 1 + 63) // 64
    for c, poly, wt in zip(got["best"], got["polys"], got["weights"]):
        st, opoly = orc.BinaryMatrix.candidate(n, 1, c).evaluate(want_poly=True)
        assert st == STAGE_PRIMITIVE
        assert np.array_equal(poly, opoly[:pw]), c
        assert wt == sum(bin(int(w)).count("1") for w in opoly[:pw]) and wt % 2 == 1
