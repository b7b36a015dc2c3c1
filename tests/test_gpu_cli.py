"""GPU: the drop-in command-line tools keep the reference's stdout lines and `out` file format.

rnghunt: software/rnghunt/src/bin/rnghunt.rs:15,45-53; clt-grng: software/clt-grng/clt-grng-evaluate.py:30-31."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_rnghunt_cli_stdout_and_out_file(tmp_path, orc):
    exe = os.path.join(ROOT, "cli", "rnghunt")
    assert os.path.exists(exe), "cli/rnghunt is not built (make)"
    n, seed = 32, 1
    res = subprocess.run([exe, "--n", str(n), "--seed", str(seed), "--stats"], cwd=tmp_path, capture_output=True,
                         text=True, timeout=300)
    assert res.returncode == 0, res.stderr
    assert res.stdout == "Starting search for n=%d\nFound suitable recurrence matrix for n=%d, writing out\n" % (n, n)
    lines = (tmp_path / "out").read_text().split("\n")
    assert lines[-1] == "" and len(lines) == n + 1
    assert all(len(ln) == n and set(ln) <= {"0", "1"} for ln in lines[:n])
    # the file is the first accepted candidate of the stream, and the oracle accepts it
    stages, _ = orc.rnghunt_sweep(n, seed, 0, 4096, threads=os.cpu_count() or 1)
    first_hit = int(np.nonzero(stages == 4)[0][0])
    want = orc.BinaryMatrix.candidate(n, seed, first_hit)
    for r in range(n):
        assert lines[r] == want.row(r).to_bitstring()
    packed = [[c for c, ch in enumerate(lines[r]) if ch == "1"] for r in range(n)]
    assert orc.BinaryMatrix.from_packed(packed).evaluate() == 4


def test_rnghunt_cli_default_is_n192(tmp_path):
    """No arguments: the reference's hard-coded n = 192 (rnghunt.rs:14); first line is printed at once."""
    exe = os.path.join(ROOT, "cli", "rnghunt")
    res = subprocess.run([exe, "--seed", "1"], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr
    assert res.stdout.split("\n")[:2] == ["Starting search for n=192",
                                          "Found suitable recurrence matrix for n=192, writing out"]
    assert len((tmp_path / "out").read_text()) == 192 * 193


def _run_hunt(tmp_path, *args):
    exe = os.path.join(ROOT, "cli", "rnghunt")
    res = subprocess.run([exe] + [str(a) for a in args], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr
    best = []
    for ln in res.stderr.split("\n"):
        if ln.startswith("best["):
            parts = ln.split()
            idx = int(parts[0].split("=")[1])
            wt = int(parts[1].split("=")[1]) if len(parts) > 1 else None
            best.append((idx, wt))
    return res, best, (tmp_path / "out").read_text()


def test_rnghunt_cli_same_answer_with_one_and_with_all_workers(tmp_path):
    """rnghunt.rs:16-18 uses every CPU; the drop-in uses every GPU (one host thread + ctx each).  The first hit
    and the best-k list must not depend on how many workers share the candidates (here: one device, the same
    device three times, and every visible device)."""
    n, seed = 64, 3
    runs = []
    for devs in (["--devices", "0"], ["--devices", "0,0,0"], []):
        _, best, out = _run_hunt(tmp_path, "--n", n, "--seed", seed, "--all", 20000, "--k", 8, *devs)
        runs.append(([b for b, _ in best], out))
    assert len(runs[0][0]) == 8
    assert runs[0] == runs[1] == runs[2]
    firsts = []
    for devs in (["--devices", "0"], ["--devices", "0,0"], []):
        res, best, out = _run_hunt(tmp_path, "--n", n, "--seed", seed, "--batch", 128, "--stats", *devs)
        firsts.append((best, out))
    assert firsts[0] == firsts[1] == firsts[2]
    assert firsts[0][0][0][0] == runs[0][0][0]            # first hit == smallest index of the best-k run
    assert firsts[0][1] == runs[0][1]


def test_rnghunt_cli_rank_by_weight(tmp_path, orc):
    """--all --rank weight: accepted candidates ordered by the number of terms of their characteristic
    polynomial (util/verify.py:19-28), weights equal to the oracle's Berlekamp-Massey polynomial."""
    n, seed, count = 32, 1, 4000
    _, best, out = _run_hunt(tmp_path, "--n", n, "--seed", seed, "--all", count, "--k", 12, "--rank", "weight",
                             "--devices", "0,0")
    stages, _ = orc.rnghunt_sweep(n, seed, 0, count, threads=os.cpu_count() or 1)
    hits = [int(i) for i in np.nonzero(stages == 4)[0]]
    pw = (n + 1 + 63) // 64
    weight = {}
    for c in hits:
        st, poly = orc.BinaryMatrix.candidate(n, seed, c).evaluate(want_poly=True)
        assert st == 4
        weight[c] = sum(bin(int(w)).count("1") for w in poly[:pw])
    want = sorted(hits, key=lambda c: (weight[c], c))[:12]
    assert [b for b, _ in best] == want
    assert [w for _, w in best] == [weight[c] for c in want]
    m = orc.BinaryMatrix.candidate(n, seed, want[0])
    assert out.split("\n")[:n] == [m.row(r).to_bitstring() for r in range(n)]


def test_clt_grng_evaluate_prints_the_reference_lines():
    res = subprocess.run([sys.executable, os.path.join(ROOT, "cli", "clt_grng_evaluate.py")], capture_output=True,
                         text=True, timeout=300)
    assert res.returncode == 0, res.stderr
    out = res.stdout.split("\n")
    assert out[0] == "Theoretical mean μ=0.0000e+00, variance σ²=6.4000e+01."
    assert out[1].startswith("Sample mean μ=") and ", variance σ²=" in out[1] and out[1].endswith(".")
    var = float(out[1].split("σ²=")[1].rstrip("."))
    assert 60.0 < var < 68.0
