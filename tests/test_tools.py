"""Format tools of SURVEY.md 8f-2 (cli/pack.py, cli/verify.py) against fixtures produced by running
the reference's own scripts (tests/golden/make_golden_tools.py): software/rnghunt/util/pack.py (whole
stdout) and util/verify.py:37-52 (the dieharder `outnums` writer)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, os.path.join(ROOT, "cli"))


def _matrix_file(tmp_path, n):
    from basebandboard_b200 import recurrences
    p = tmp_path / ("m%d" % n)
    p.write_text(recurrences.to_matrix_text(recurrences.packed(n)))   # == matrices/N of the reference
    return str(p)


@pytest.mark.parametrize("n", [16, 64])
def test_pack_stdout_is_the_reference_scripts(tmp_path, n):
    res = subprocess.run([sys.executable, os.path.join(ROOT, "cli", "pack.py"), _matrix_file(tmp_path, n)],
                         capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stderr
    assert res.stdout == open(os.path.join(GOLD, "pack_%d.txt" % n)).read()


def test_pack_module_block_round_trips(tmp_path):
    from basebandboard_b200 import recurrences
    res = subprocess.run([sys.executable, os.path.join(ROOT, "cli", "pack.py"), _matrix_file(tmp_path, 32), "--module"],
                         capture_output=True, text=True, timeout=120)
    block = res.stdout[res.stdout.index("n32 = ["):]
    ns = {}
    exec(block, ns)                                  # it is valid Python in the rng_recurrences.py layout
    assert ns["n32"] == recurrences.packed(32)
    assert block.split("\n")[1].startswith("    [5, 15, 19], [11, 25, 30, 31], ")   # rng_recurrences.py:19-20


def test_binary_matrix_layout_matches_oracle(orc):
    from basebandboard_b200 import recurrences
    for n in (16, 64, 192):
        want = orc.BinaryMatrix.from_packed(recurrences.packed(n)).data
        assert np.array_equal(recurrences.to_binary_matrix(recurrences.packed(n)), want)


def test_dieharder_words_bit_order(orc, tmp_path):
    """outnums of verify.py:37-52 (n = 32, start = LUTOPT reset value, 64 steps) from oracle URNG words."""
    import verify
    from basebandboard_b200 import recurrences
    words = orc.clt_run(recurrences.packed(32), 1, 0, 64, model=1, want=("words",))["words"].reshape(64, 1)
    out = tmp_path / "outnums"
    verify.write_outnums(str(out), verify.dieharder_words(words, 32))
    assert out.read_text() == open(os.path.join(GOLD, "outnums_32.txt")).read()


@pytest.mark.gpu
def test_verify_cli_writes_reference_outnums(tmp_path):
    res = subprocess.run([sys.executable, os.path.join(ROOT, "cli", "verify.py"), _matrix_file(tmp_path, 32),
                          "--nsamples", "64", "--outnums", str(tmp_path / "outnums")],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    lines = res.stdout.split("\n")
    assert lines[0] == "Detected n=32" and lines[1] == "Loaded matrix" and lines[2].startswith("Row weights: [")
    assert "Verdict: maximum period (degree 32, primitive)" in res.stdout
    assert (tmp_path / "outnums").read_text() == open(os.path.join(GOLD, "outnums_32.txt")).read()


@pytest.mark.gpu
def test_pack_check_verdicts(tmp_path):
    from basebandboard_b200 import recurrences
    good = _matrix_file(tmp_path, 64)
    res = subprocess.run([sys.executable, os.path.join(ROOT, "cli", "pack.py"), good, "--check"],
                         capture_output=True, text=True, timeout=300)
    assert res.stdout.rstrip().endswith("Verdict: maximum period (degree 64, primitive)")
    rows = recurrences.packed(64)
    rows[3] = rows[4]                                   # singular: two equal rows
    bad = tmp_path / "bad"
    bad.write_text(recurrences.to_matrix_text(rows))
    res = subprocess.run([sys.executable, os.path.join(ROOT, "cli", "pack.py"), str(bad), "--check"],
                         capture_output=True, text=True, timeout=300)
    assert "rejected at stage 'degree'" in res.stdout
