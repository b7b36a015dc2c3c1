"""CPU checks of things around the kernels: the committed straight-line kernel bodies are what the generator
emits (tools/gen_clt256_v2.py is deterministic for the committed load order), and the reference arm of
bench.py prints the contract's JSON line (metric, config, cpu_baseline, e2e, the rnghunt leg) without a GPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_generated_kernel_bodies_are_reproducible(tmp_path):
    out = tmp_path / "regen.inc"
    env = dict(os.environ, CLT256V2_OUT=str(out))
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_clt256_v2.py")], env=env,
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr
    committed = open(os.path.join(ROOT, "basebandboard_b200", "csrc", "clt256_v2_gen.inc")).read()
    assert out.read_text() == committed
    # 492 recurrence LOP3 + 247 full adders + 8 half adders per clock
    assert "recurrence LOP3: 492, carry-save adders: 247, half adders: 8" in committed


def test_generated_plain_clock_bodies_equal_the_recurrence():
    """P_BODY (104 resident planes) and P2_BODY (151, with its spill / fill sections) of the committed .inc, run as
    Python on random words, == five clocks of the shipped n256 recurrence (tools/check_clt256_bodies.py)."""
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "check_clt256_bodies.py")],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr
    assert res.stdout.startswith("ok")


def test_tuning_keys_in_the_header_are_the_ones_the_library_accepts():
    import re
    h = open(os.path.join(ROOT, "include", "bbb_b200.h")).read()
    a = open(os.path.join(ROOT, "basebandboard_b200", "csrc", "api.cu")).read()
    doc = set(re.findall(r'"([a-z0-9_]+)"', h[h.index("Tuning knobs"):h.index("int  bbb_set_tuning")]))
    impl = set(re.findall(r'strcmp\(key, "([a-z0-9_]+)"\)', a))
    assert doc == impl, (sorted(doc - impl), sorted(impl - doc))


def test_bench_reference_arm_prints_the_contract_line():
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "1", "--ref-step-seconds", "0.3", "--cpu-seconds", "0.5"],
                         capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert res.returncode == 0, res.stderr
    lines = [ln for ln in res.stdout.split("\n") if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "clt_grng_gsamples_per_s" and d["unit"] == "Gsamples/s"
    assert d["higher_is_better"] is True and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": "Gsamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    hunt = d["secondary"]["rnghunt"]
    assert hunt["metric"] == "rnghunt_candidates_per_s" and hunt["value"] > 0
    assert "configs[2]" in d["config"]["workload"]


def test_bench_algorithmic_work_constants():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.ALG_LOP3_PER_STEP32 == 492 + 494
    # n = 192, K = 8: 48 passes x 7 row groups x 7 + 7 products x 6 rows x 4 taps x 6 words, per lane, x 32 lanes
    assert bench.hunt_alg_lop3_per_candidate(192) == 32 * (48 * 49 + 7 * 6 * 4 * 6)
