#!/usr/bin/env python3
"""bench.py -- the headline measurement (BASELINE.json: bit-exact CLT-GRNG Gsamples/s/GPU).

Workload (BASELINE.json configs[2], the configuration the metric is quoted on): LUTOPT(n256,
init=1) -> CLTGRNG, 2^40 samples per GPU per step, fused generate + histogram + moments, no
stream stored.  Rank r owns samples [r*2^40, (r+1)*2^40) (jump-ahead seeding, no data-path
collective); the only exchange is one NCCL all-reduce of u64[259] (histogram | sum | sumsq) per step.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--log2-samples 40] [--impl reference]

`value`     device-timed (CUDA events on the library's stream, max over ranks), results stay on device.
`e2e`       the same step through the host-buffer C-ABI call bbb_clt_hist (host seed in, host
            histogram/moments out, all copies inside the timed region).
`roofline`  the fused kernel is bound by the integer ALU pipe (no HBM traffic, no tensor work):
            achieved = samples/s * 986/32 LOP3-equivalent thread-instructions, peak =
            LOP3 issue rate measured live by bbb_measure_lop3_peak.
`cpu_baseline`  the CPU oracle port (oracle/, C, all host cores) on a bounded sample.
The oracle is imported only for the cpu_baseline leg and for --impl reference.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_STATE = 256
ALG_LOP3_PER_STEP32 = 986          # 492 (recurrence) + 494 (popcount tree) per 32 bit-sliced samples
METRIC = "clt_grng_gsamples_per_s"
UNIT = "Gsamples/s"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index, period=0.2):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(self.period)

    def finish(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def physical_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_port_rate(count, threads, first=0):
    """Oracle port (oracle/gen.c orc_clt_run_mt) on `threads` host threads; returns samples/s."""
    from oracle import oracle
    from basebandboard_b200 import recurrences
    packed = recurrences.packed(N_STATE)
    oracle.clt_run(packed, 1, first, 1 << 16, model=0, threads=threads)       # warm
    t0 = time.perf_counter()
    oracle.clt_run(packed, 1, first, count, model=0, threads=threads)
    return count / (time.perf_counter() - t0)


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path.  The reference (Rust
    crate + numpy model lines) cannot be built or shipped (no Rust toolchain; /root/reference is
    absent on the GPU box), so this times the oracle port on all host cores; each step is a bounded
    sample of the same workload (same recurrence, seed, indexing, fused histogram + moments)."""
    if rank != 0:
        return
    cores = host_cores()
    cal = cpu_port_rate(1 << 21, cores)
    per_step = max(1 << 20, int(cal * args.ref_step_seconds))
    from oracle import oracle
    from basebandboard_b200 import recurrences
    packed = recurrences.packed(N_STATE)
    for _ in range(args.warmup):
        oracle.clt_run(packed, 1, 0, min(per_step, 1 << 20), model=0, threads=cores)
    t0 = time.perf_counter()
    for s in range(args.steps):
        oracle.clt_run(packed, 1, s * per_step, per_step, model=0, threads=cores)
    dt = time.perf_counter() - t0
    val = args.steps * per_step / dt / 1e9
    sample = "%d samples per step from sample index 0 (of the 2^%d per-GPU workload)" % (per_step, args.log2_samples)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32 (GF(2) bit ops)",
        "data": "synthetic", "config": workload_config(args, world),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if not args.no_secondary:
        # the metric's second half on the same arm: the oracle port of the rnghunt worker loop, all host cores
        rate, cnt, counts = cpu_rnghunt_rate(min(args.cpu_seconds, 8.0), cores)
        line["secondary"] = {"rnghunt": {
            "metric": "rnghunt_candidates_per_s", "value": rate, "unit": "candidates/s", "higher_is_better": True,
            "config": {"workload": "rnghunt sweep n=%d, seed %d (BASELINE.json configs[3]); bounded sample of the fixed "
                                   "candidate set [0, 2^%d)" % (HUNT_N, HUNT_SEED, args.hunt_log2)},
            "cpu_baseline": {"value": rate, "unit": "candidates/s", "cores": cores, "kind": "port",
                             "sample": "candidates [0, %d) (oracle/rnghunt_ref.c, %d threads), stage counts %s"
                                       % (cnt, cores, counts)},
            "e2e": {"value": rate, "unit": "candidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}}
    emit_json(line)


def workload_config(args, world):
    return {
        "workload": "clt-grng 2^%d samples/GPU fused generate+histogram+moments, no stream stored "
                    "(BASELINE.json configs[2]); LUTOPT n256 init=1, gateware widths" % args.log2_samples,
        "n": N_STATE, "log2_samples_per_gpu": args.log2_samples, "seed_init": 1,
        "partition": "rank r owns sample indices [r*2^%d, (r+1)*2^%d)" % (args.log2_samples, args.log2_samples),
        "collective": "all-reduce of u64[259] per step" if world > 1 else "none",
        "l2": "no HBM-resident input: each step regenerates its seed planes (a few MB) and keeps "
              "generator state on chip; nothing is re-read across steps, so no L2 flush applies",
    }


HUNT_N, HUNT_SEED, HUNT_LOG2, HUNT_K = 192, 1, 24, 16     # BASELINE.json configs[3] as SURVEY.md 8d restates it


def hunt_alg_lop3_per_candidate(n, k=8):
    """Algorithmic ALU work of the sweep's dominant kernel (hunt_front_pow_kernel) per candidate, as LOP3-equivalent
    thread-instructions: the 2n clocks of x' = A.x (binary_matrix.rs:68-76) are applied K at a time through A^K kept
    as dense row masks -- (2n/K) passes of (NW+1) row groups x (NW mask LOP3 + 1 parity predicate) per lane -- after
    K-1 sparse products A.P that build A^K (4 taps x NW word-XORs per row).  Candidate generation, Berlekamp-Massey,
    votes and addressing are NOT counted.  (The same clocks as one-at-a-time sparse gathers, the round-1
    formulation: 2n x 31n/8 = 286 k byte gathers at n = 192.)"""
    nw = (n + 31) // 32
    passes = -(-2 * n // k) * (nw + 1) * (nw + 1)
    powers = (k - 1) * nw * 4 * nw
    return 32 * (passes + powers)


def crc_u64(words):
    import zlib
    import numpy as np
    return zlib.crc32(np.ascontiguousarray(words, dtype=np.uint64).tobytes()) & 0xFFFFFFFF


def cpu_rnghunt_rate(seconds, threads):
    """Oracle port of the rnghunt worker loop (oracle/rnghunt_ref.c) on `threads` host threads: candidates/s
    on a bounded prefix of the same candidate stream."""
    from oracle import oracle
    t0 = time.perf_counter()
    oracle.rnghunt_sweep(HUNT_N, HUNT_SEED, 0, 64 * threads, threads=threads)
    cal = 64 * threads / (time.perf_counter() - t0)
    cnt = max(64 * threads, int(cal * seconds))
    t0 = time.perf_counter()
    _, counts = oracle.rnghunt_sweep(HUNT_N, HUNT_SEED, 0, cnt, threads=threads)
    return cnt / (time.perf_counter() - t0), cnt, [int(v) for v in counts]


def python_loop_rate(nsamp=4000):
    """The reference's literal implementation, clt-grng-evaluate.py:8-16, restated in oracle/oracle.py
    (py_clt_evaluate_loop): samples/s on one core."""
    from oracle import oracle
    t0 = time.perf_counter()
    oracle.py_clt_evaluate_loop(nsamp)
    return nsamp / (time.perf_counter() - t0)


def run_rnghunt_leg(args, ctx, rank, world, local_rank, lop3_peak, barrier):
    """The metric's second half (BASELINE.json: "rnghunt candidates/s at 1/2/4/8 B200", configs[3]): n = 192,
    seed 1, the FIXED candidate set [0, 2^24) interleaved c = rank (mod N) (strong scaling), per-rank best-16
    all-gathered and merged inside the timed region (rnghunt.rs:14,23-57).  Returns the dict that goes under
    "secondary": {"rnghunt": ...} of the JSON line (rank 0), None elsewhere."""
    import torch
    import torch.distributed as dist
    from basebandboard_b200 import sharding
    total = 1 << args.hunt_log2
    res = {}

    def step():
        res["r"] = sharding.rnghunt_sweep_sharded(ctx, HUNT_N, HUNT_SEED, 0, total, k=HUNT_K)
        return ctx.last_launches

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(physical_gpu_index(local_rank))
    sampler.start()
    launches = 0
    t0 = time.perf_counter()
    ctx.timer_start()
    for _ in range(args.steps):
        launches += step()
    dev_ms = ctx.timer_stop()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.finish()
    h2d, d2h = ctx.last_h2d_bytes, ctx.last_d2h_bytes
    sweep_ms = ctx.last_device_ms                    # this rank's share of the sweep, device-timed, one call
    barrier()
    # e2e: the same public call, host arguments in, merged host ranking out, wall clock
    t1 = time.perf_counter()
    step()
    e2e_ms = (time.perf_counter() - t1) * 1e3
    t = torch.tensor([dev_ms, wall_ms, e2e_ms, sweep_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms, e2e_ms, sweep_ms = [float(v) for v in t.cpu()]
    if rank != 0:
        return None
    value = total * args.steps / (dev_ms * 1e-3)
    alg = hunt_alg_lop3_per_candidate(HUNT_N)
    per_gpu = (total / world) / (sweep_ms * 1e-3)
    ach = per_gpu * alg / 1e12
    out = {
        "metric": "rnghunt_candidates_per_s", "value": value, "unit": "candidates/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "dtype": "u32 GF(2)",
        "config": {"workload": "rnghunt sweep n=%d, seed %d, fixed candidate set [0, 2^%d) interleaved c = rank (mod N), "
                               "merged best-%d (BASELINE.json configs[3])" % (HUNT_N, HUNT_SEED, args.hunt_log2, HUNT_K),
                   "collective": "one all-gather of k indices + 5 stage counts per rank, inside the timed region"
                                 if world > 1 else "none"},
        "wall_ms_per_step": wall_ms / args.steps, "clocks": clocks,
        "e2e": {"value": total / (e2e_ms * 1e-3), "unit": "candidates/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "api": "bbb_rnghunt_sweep per rank + merge (host scalars in, host ranking out)"},
        "gpu_launches": launches,
        "roofline": {"bound": "alu", "achieved": ach, "peak": lop3_peak / 1e12,
                     "unit": "T thread-instr/s (LOP3-equivalent)", "frac": ach / (lop3_peak / 1e12), "traffic": None,
                     "launch_ms_per_2p20": sweep_ms / max(1.0, total / world / (1 << 20)),
                     "kernel": "hunt_front_pow<6,8>",
                     "alu_pipe_busy_ncu": 0.83,
                     "survey_tap_xors_per_candidate": int(2 * HUNT_N * HUNT_N * 31 / 8),
                     "survey_tap_xor_rate_T_per_s": per_gpu * (2 * HUNT_N * HUNT_N * 31 / 8) / 1e12,
                     "note": "algorithmic work = (2n/8) passes of A^8 over 7 row groups x (6 mask LOP3 + 1 parity predicate) "
                             "+ 7 sparse products A.P (4 taps x 6 word-XORs per row) = %d thread-instr per candidate "
                             "(binary_matrix.rs:68-76); candidate generation (binary_matrix.rs:81-101), Berlekamp-Massey, "
                             "votes and the primitivity stage are NOT counted, which is why frac is far below the ALU-pipe "
                             "utilisation ncu reports for the kernel (83 %%, profiles/r2_rnghunt_pow8_ncu_full.txt); peak = "
                             "LOP3 issue rate measured live (bbb_measure_lop3_peak); HBM traffic ~0.  For continuity with "
                             "round 1 (one operation per tap of SURVEY 8d's 2n x 31n/8 tap-XORs): survey_tap_xor_rate_T_per_s, "
                             "which A^K lets exceed one-instruction-per-tap formulations" % alg},
        "result_check": {"best": [int(b) for b in res["r"]["best"]],
                         "stage_counts": [int(v) for v in res["r"]["stage_counts"]]},
    }
    if not args.no_cpu_baseline:
        cores = host_cores()
        rate, cnt, counts = cpu_rnghunt_rate(min(args.cpu_seconds, 8.0), cores)
        out["cpu_baseline"] = {"value": rate, "unit": "candidates/s", "cores": cores, "kind": "port",
                               "sample": "candidates [0, %d) of the same stream (oracle/rnghunt_ref.c, %d threads), "
                                         "stage counts %s" % (cnt, cores, counts)}
    return out


def run_secondary(args, ctx, rank, world, local_rank):
    """The other BASELINE.json configs through the same contract (device-timed value, e2e through the
    host-buffer API, max over ranks, clocks).  One JSON line from rank 0."""
    import torch
    import torch.distributed as dist
    from basebandboard_b200 import lib, sharding

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sm_count = torch.cuda.get_device_properties(local_rank).multi_processor_count
    hbm_peak = None
    try:
        hbm_peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    w = args.workload
    extra = {}
    if w == "prbs_stream":
        nbits = 1 << args.log2_bits
        first = rank * nbits
        buf = torch.empty(nbits // 64, dtype=torch.int64, device="cuda")
        def step():
            ctx.prbs_stream_dev(args.k, nbits, buf.data_ptr(), first_bit=first)
            return ctx.last_launches
        host = {}
        def step_e2e():          # host buffer out: the D2H copies of the packed stream are inside the timed region
            n = min(nbits, 1 << 35)
            if "b" not in host:  # caller-owned buffer, touched once (a fresh calloc page-faults on first write)
                import numpy as np
                host["b"] = np.ones(n // 64, dtype=np.uint64)
                ctx.prbs_stream(args.k, n, first_bit=first, out=host["b"])   # untimed: the ctx allocates its staging buffers
                return None
            ctx.prbs_stream(args.k, n, first_bit=first, out=host["b"])
            return n, ctx.last_h2d_bytes, ctx.last_d2h_bytes
        units, metric, unit, scale = nbits, "prbs_stream_gbytes_per_s", "GB/s", 1 / 8e9
        workload = "PRBS-%d packed stream materialised to HBM, 2^%d bits/GPU (BASELINE.json configs[4])" % (args.k, args.log2_bits)
        dtype = "u64 (GF(2) word recurrence)"
        l2 = "output (%d MiB) exceeds L2; nothing is read" % (nbits >> 23)
    elif w == "prbs_detect":
        # rank r owns bits [r*2^b, (r+1)*2^b); the detector state is handed from rank to rank (speculate the locked
        # state, all-gather end states, re-run on mismatch) and the 5 counters are all-reduced -- all inside step()
        nbits = 1 << args.log2_bits
        def step():
            ctr, final, runs = sharding.prbs_detect_sharded(ctx, args.k, world * nbits, m=args.err_m, seed=1)
            extra["ctr"], extra["rounds"], extra["runs"], extra["final"] = ctr, ctx.last_rounds, runs, final
            return ctx.last_launches * runs
        def step_e2e():
            step()
            return nbits, ctx.last_h2d_bytes + 16, ctx.last_d2h_bytes
        units, metric, unit, scale = nbits, "prbs_detect_tbit_per_s", "Tbit/s", 1e-12
        workload = ("PRBS-%d generate + Bernoulli(2^-%d) channel + self-synchronising error count, 2^%d bits/GPU, "
                    "contiguous bit ranges with detector-state hand-off (BASELINE.json configs[1])"
                    % (args.k, args.err_m, args.log2_bits))
        dtype = "u32/u64 integer"
        l2 = "no HBM-resident input"
    elif w == "prbs_slicer":
        # SURVEY 8f-1 composite; contiguous bit ranges per rank (PRBS and noise are indexed by absolute bit), start
        # state speculated from a 1024-bit warm-up window, hand-off loop and counter all-reduce inside step()
        nbits = 1 << args.log2_bits
        ctx.lutopt_load_shipped(N_STATE)
        def step():
            ctr, final, runs = sharding.prbs_slicer_detect_sharded(ctx, args.k, world * nbits, init=1, amplitude=args.amplitude,
                                                                   noise_var=args.noise_var, spb=args.spb, delay=0)
            extra["ctr"], extra["rounds"], extra["runs"], extra["final"] = ctr, ctx.last_rounds, runs, final
            return ctx.last_launches * runs
        def step_e2e():
            step()
            return nbits, ctx.last_h2d_bytes + 48, ctx.last_d2h_bytes
        units, metric, unit, scale = nbits, "prbs_slicer_gbit_per_s", "Gbit/s", 1e-9
        workload = ("PRBS-%d -> +-%d + CLT-GRNG(n256) x %d (12-bit TX word, %d samples/bit) -> slicer -> error detector, "
                    "2^%d bits per GPU (SURVEY 8f-1; contiguous bit ranges per rank)" % (args.k, args.amplitude, args.noise_var,
                                                                            args.spb, args.log2_bits))
        dtype = "u32 GF(2) + int8/int32"
        l2 = "intermediate buffer = the one kept noise sample per bit (%d MiB per pass) exceeds L2" % (min(nbits, 1 << 32) >> 20)
    elif w == "prbs_shaped_slicer":
        # SURVEY 8f-3 composite: shaped TX word + noise -> slicer at sample `delay` -> detector; ranks as for prbs_slicer
        from basebandboard_b200 import shaper
        nbits = 1 << args.log2_bits
        ctx.lutopt_load_shipped(N_STATE)
        coeff = shaper.tx_sets()[16]
        def step():
            ctr, final, runs = sharding.prbs_shaped_slicer_detect_sharded(ctx, args.k, world * nbits, coeff, init=1,
                                                                          noise_var=args.noise_var, delay=4)
            extra["ctr"], extra["rounds"], extra["runs"], extra["final"] = ctr, ctx.last_rounds, runs, final
            return ctx.last_launches * runs
        def step_e2e():
            step()
            return nbits, ctx.last_h2d_bytes + 48 + 128, ctx.last_d2h_bytes
        units, metric, unit, scale = nbits, "prbs_shaped_slicer_gbit_per_s", "Gbit/s", 1e-9
        workload = ("PRBS-%d through PRBSShaper set 16 + CLT-GRNG(n256) x %d (12-bit TX word, 8 samples/bit) -> slicer at "
                    "sample 4 -> error detector, 2^%d bits per GPU (SURVEY 8f-3; contiguous bit ranges per rank)"
                    % (args.k, args.noise_var, args.log2_bits))
        dtype = "u32 GF(2) + int8/int32"
        l2 = "intermediate buffer = the one kept noise sample per bit (%d MiB per pass) exceeds L2" % (min(nbits, 1 << 32) >> 20)
    elif w == "tx_stream":
        from basebandboard_b200 import shaper
        nbits = 1 << (args.log2_bits - 7)                 # 16 B per bit: 2^36 "bits" of budget = 8 GiB of TX words
        first = rank * nbits
        coeff = shaper.tx_sets()[16]
        buf = torch.empty(nbits * 8, dtype=torch.int16, device="cuda")
        def step():
            ctx.tx_stream(args.k, nbits, coeff, first_bit=first, dev_ptr=buf.data_ptr())
            return ctx.last_launches
        host = {}
        def step_e2e():
            n = min(nbits, 1 << 28)
            if "b" not in host:
                import numpy as np
                host["b"] = np.ones(n * 8, dtype=np.int16)
                ctx.tx_stream(args.k, n, coeff, first_bit=first, out=host["b"])
                return None
            ctx.tx_stream(args.k, n, coeff, first_bit=first, out=host["b"])
            return n * 16, ctx.last_h2d_bytes, ctx.last_d2h_bytes
        units, metric, unit, scale = nbits * 16, "tx_stream_gbytes_per_s", "GB/s", 1e-9
        workload = ("shaped PRBS-%d TX words (PRBSShaper set 16, noise off), 8 int16 per bit materialised to HBM, "
                    "2^%d bits/GPU (SURVEY 8f-3)" % (args.k, args.log2_bits - 7))
        dtype = "int16 sums of 9-bit coefficients"
        l2 = "output (%d MiB) exceeds L2; nothing is read back" % ((nbits * 16) >> 20)
    elif w == "lutopt_stream":
        # BASELINE.json configs[4](ii): the raw URNG word of every clock (rng.py:27,38-40), 32 B per step
        count = 1 << args.log2_steps
        first = rank * count
        ctx.lutopt_load_shipped(N_STATE)
        buf = torch.empty(count * 4, dtype=torch.int64, device="cuda")
        def step():
            ctx.lutopt_stream_dev(1, first, count, buf.data_ptr())
            return ctx.last_launches
        host = {}
        def step_e2e():
            n = min(count, 1 << 27)
            if "b" not in host:
                import numpy as np
                host["b"] = np.ones(n * 4, dtype=np.uint64)
                ctx.lutopt_stream(1, first, n, out=host["b"])
                return None
            ctx.lutopt_stream(1, first, n, out=host["b"])
            return n * 32, ctx.last_h2d_bytes, ctx.last_d2h_bytes
        units, metric, unit, scale = count * 32, "lutopt_stream_gbytes_per_s", "GB/s", 1e-9
        workload = ("raw LUTOPT-256 URNG words (init 1) materialised to HBM, 2^%d steps x 32 B per GPU "
                    "(BASELINE.json configs[4]; rng.py:27)" % args.log2_steps)
        dtype = "u32 (GF(2) bit ops)"
        l2 = "output (%d MiB) exceeds L2; nothing is read back" % ((count * 32) >> 20)
    elif w == "clt_stream":
        count = 1 << args.log2_samples
        first = rank * count
        ctx.lutopt_load_shipped(N_STATE)
        eb = 2 if args.noise_var_stream >= 0 else 1
        buf = torch.empty(count * eb, dtype=torch.uint8, device="cuda")
        def step():
            ctx.clt_stream_dev(1, first, count, 0, max(args.noise_var_stream, 1), buf.data_ptr(), eb)
            return ctx.last_launches
        host = {}
        def step_e2e():
            n = min(count, 1 << 32)
            if "b" not in host:
                import numpy as np
                host["b"] = np.ones(n, dtype=np.int8 if eb == 1 else np.int16)
                ctx.clt_stream(1, first, n, model=0, noise_var=None if eb == 1 else args.noise_var_stream, out=host["b"])
                return None
            ctx.clt_stream(1, first, n, model=0, noise_var=None if eb == 1 else args.noise_var_stream, out=host["b"])
            return n, ctx.last_h2d_bytes, ctx.last_d2h_bytes
        units, metric, unit, scale = count, "clt_stream_gsamples_per_s", "Gsamples/s", 1e-9
        workload = ("clt-grng n256 %s samples materialised to HBM, 2^%d samples/GPU (BASELINE.json configs[4])"
                    % ("int8" if eb == 1 else "int16 x noise_var %d" % args.noise_var_stream, args.log2_samples))
        dtype = "u32 (GF(2) bit ops)"
        l2 = "output (%d MiB) exceeds L2; nothing is read back" % ((count * eb) >> 20)
    else:
        cnt = 1 << args.log2_candidates
        def step():
            extra["res"] = ctx.rnghunt_sweep(args.hunt_n, 1, rank, cnt, k=16, stride=world)
            return ctx.last_launches
        def step_e2e():
            step()
            return cnt, ctx.last_h2d_bytes, ctx.last_d2h_bytes
        units, metric, unit, scale = cnt, "rnghunt_candidates_per_s", "candidates/s", 1.0
        workload = ("rnghunt sweep n=%d, seed 1, 2^%d candidates/GPU interleaved over ranks, best-16 "
                    "(BASELINE.json configs[3])" % (args.hunt_n, args.log2_candidates))
        dtype = "u32 GF(2)"
        l2 = "no HBM-resident input"

    lop3_peak = max(ctx.measure_lop3_peak() for _ in range(3)) / 1e12 if w in ("prbs_slicer", "prbs_shaped_slicer", "clt_stream", "prbs_detect") else None
    for _ in range(args.warmup):
        step()
        ctx.sync()
    barrier()
    sampler = ClockSampler(physical_gpu_index(local_rank))
    sampler.start()
    launches = 0
    kernel_ms = 0.0
    t0 = time.perf_counter()
    ctx.timer_start()                       # CUDA events on the library's stream around the K steps: host-side hand-off
    for _ in range(args.steps):             # and the small collectives of the sharded legs are inside
        launches += step()
        ctx.sync()
        kernel_ms += ctx.last_device_ms
    dev_ms = ctx.timer_stop()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.finish()
    barrier()
    if step_e2e() is None:           # first call of a stream workload only allocates the caller's host buffer
        barrier()
    t1 = time.perf_counter()
    e_units, h2d, d2h = step_e2e()
    e2e_ms = (time.perf_counter() - t1) * 1e3
    e2e_pinned_ms = 0.0
    if w in ("prbs_stream", "tx_stream", "clt_stream", "lutopt_stream"):
        # the same call into PAGE-LOCKED caller memory (a pinned torch tensor): the library copies each chunk straight
        # into it, no staging copy on the host
        pageable = host["b"]
        host["b"] = torch.empty(pageable.nbytes, dtype=torch.uint8, pin_memory=True).numpy().view(pageable.dtype)
        step_e2e()
        barrier()
        t2 = time.perf_counter()
        step_e2e()
        e2e_pinned_ms = (time.perf_counter() - t2) * 1e3
        host["b"] = pageable
    t = torch.tensor([dev_ms, wall_ms, e2e_ms, e2e_pinned_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms, e2e_ms, e2e_pinned_ms = [float(v) for v in t.cpu()]
    if rank == 0:
        value = world * units * args.steps / (dev_ms * 1e-3) * scale
        line = {"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": dtype, "data": "synthetic",
                "config": {"workload": workload, "l2": l2},
                "wall_ms_per_step": wall_ms / args.steps, "clocks": clocks,
                "e2e": {"value": world * e_units / (e2e_ms * 1e-3) * scale, "unit": unit,
                        "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                "gpu_launches": launches}
        line["e2e"]["host_buffer"] = "pageable numpy array owned by the caller"
        if e2e_pinned_ms:
            line["e2e_pinned"] = {"value": world * e_units / (e2e_pinned_ms * 1e-3) * scale, "unit": unit,
                                  "host_buffer": "page-locked (pinned torch tensor): chunks are copied straight into it"}
        if w == "tx_stream" and hbm_peak:
            ach = units / (dev_ms / args.steps * 1e-3) / 1e9
            line["roofline"] = {"bound": "hbm", "kernel": "shaper_stream", "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                                "frac": ach / hbm_peak,
                                "traffic": 8597517000 if (args.log2_bits == 36 and not args.tune and sm_count == 148) else None,
                                "traffic_note": "NOT re-measured in this run (null under --tune, another size or SM count): dram__bytes_read.sum + dram__bytes_write.sum of one shaper_stream_kernel launch at "
                                                "2^29 bits (ncu --set full, profiles/r2_tx_words_kernels_ncu_full.txt): 8.53 GB written "
                                                "for 8.59 GB of TX words, 67 MB read (the packed PRBS words)",
                                "note": "algorithmic bytes = 16 B written per PRBS bit (8 int16 TX words); "
                                "peak = measured copy bandwidth (MEASURED_PEAKS.json), per GPU"}
        if w == "prbs_stream" and hbm_peak:
            ach = units / 8 / (dev_ms / args.steps * 1e-3) / 1e9
            line["roofline"] = {"bound": "hbm", "kernel": "prbs_stream", "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                                "frac": ach / hbm_peak,
                                "traffic": 8529677576 if (args.log2_bits == 36 and args.k == 31 and not args.tune and sm_count == 148) else None,
                                "traffic_note": "NOT re-measured in this run (null under --tune, another size, k or SM count): dram__bytes_read.sum + dram__bytes_write.sum of the 2^36-bit prbs_stream_kernel "
                                                "launch (ncu --set full, profiles/r1_prbs_stream_ncu_full.txt): 8.53 GB written "
                                                "for 8.59 GB of packed bits (the last lines are still in L2 at kernel end), 0.1 MB read",
                                "note": "algorithmic bytes = 1/8 B per PRBS bit; peak = measured copy bandwidth "
                                        "(MEASURED_PEAKS.json), per GPU"}
        if w == "lutopt_stream" and hbm_peak:
            ach = units / (dev_ms / args.steps * 1e-3) / 1e9
            line["roofline"] = {"bound": "hbm", "kernel": ctx.last_kernel, "achieved": ach, "peak": hbm_peak, "unit": "GB/s",
                                "frac": ach / hbm_peak, "traffic": None,
                                "traffic_note": "ncu --set full of clt256_words7_kernel at 2^27 steps (profiles/r2_words7_ncu_full.txt): "
                                                "4.214 GB written + 0.017 GB read for 4.295 GB of output = 0.99x the algorithmic bytes "
                                                "(kernel alone 1.30 ms = 3.30 TB/s; the call adds the seeding kernel); not re-measured at this size",
                                "note": "algorithmic bytes = 32 B written per LUTOPT step (the 256-bit state, rng.py:27); "
                                        "peak = measured copy bandwidth (MEASURED_PEAKS.json), per GPU"}
        if w == "clt_stream":
            # still integer-pipe bound when stored (30.8 thread-instr per byte), SURVEY 8d
            ach = value / world * 1e9 * 30.8125 / 1e12
            line["roofline"] = {"bound": "alu", "kernel": "clt256_s8v2", "achieved": ach, "unit": "T thread-instr/s (LOP3-equivalent)",
                                "peak": lop3_peak, "frac": ach / lop3_peak if lop3_peak else None, "traffic": None,
                                "hbm_write_gbs": units * (2 if args.noise_var_stream >= 0 else 1) / (dev_ms / args.steps * 1e-3) / 1e9,
                                "note": "986 LOP3 per 32 samples as for the fused kernel; peak = LOP3 issue rate measured live "
                                        "(bbb_measure_lop3_peak); HBM write rate given for reference"}
        if w in ("prbs_slicer", "prbs_shaped_slicer"):
            # dominant kernel clt256_dec: spb recurrence clocks (492 LOP3 each) + one CLT tree (494) per bit and 32 lanes
            spb_ = args.spb if w == "prbs_slicer" else 8
            per_bit = (spb_ * 492 + 494) / 32.0
            ach = value / world * 1e9 * per_bit / 1e12
            line["roofline"] = {"bound": "alu", "kernel": "clt256_dec", "achieved": ach, "peak": lop3_peak,
                                "unit": "T thread-instr/s (LOP3-equivalent)", "frac": ach / lop3_peak if lop3_peak else None,
                                "traffic": None,
                                "note": "algorithmic work per bit = (spb x 492 recurrence + 494 popcount-tree LOP3) / 32 lanes; "
                                        "peak = LOP3 issue rate measured live (bbb_measure_lop3_peak); ncu: the kernel's "
                                        "shared-memory pipe is the busier one (82 % of its wavefront peak, ALU 75 %), "
                                        "profiles/r1_dec_kernels_ncu_full.txt"}
        if w == "prbs_detect":
            # ALU-pipe work per 64-bit word: m channel draws x (3 xorshifts of a 64-bit value = 6 SHF + 6 LOP3, key add
            # 2 IADD3, AND-accumulate 2 LOP3 = 16) + 8 for the PRBS word recurrence, the compare and the counters; the six
            # 32-bit multiply-adds of a draw's two 64-bit multiplies run on the FMA pipe and are not counted
            per_word = 16 * args.err_m + 8
            ach = value / world * 1e12 / 64.0 * per_word / 1e12
            line["roofline"] = {"bound": "alu", "kernel": "prbs_detect", "achieved": ach, "peak": lop3_peak,
                                "unit": "T thread-instr/s (ALU pipe)", "frac": ach / lop3_peak if lop3_peak else None,
                                "traffic": None, "alu_instr_per_64bit_word": per_word,
                                "note": "algorithmic ALU-pipe instructions per 64-bit word = 16 x err_m (splitmix64 channel draws) "
                                        "+ 8 (word recurrence, compare, counters); peak = LOP3 issue rate measured live "
                                        "(bbb_measure_lop3_peak); no HBM traffic"}
        if "ctr" in extra:
            line["result_check"] = dict(extra["ctr"], rounds=extra["rounds"], runs=extra["runs"],
                                        final_state=list(extra["final"]), counters="all-reduced over the ranks")
        if w == "rnghunt":
            lop3 = max(ctx.measure_lop3_peak() for _ in range(3)) / 1e12
            alg = hunt_alg_lop3_per_candidate(args.hunt_n)
            ach = value / world * alg / 1e12
            line["roofline"] = {"bound": "alu", "kernel": "hunt_front_pow", "achieved": ach, "peak": lop3,
                                "unit": "T thread-instr/s (LOP3-equivalent)", "frac": ach / lop3, "traffic": None,
                                "note": "algorithmic work = (2n/8) passes of A^8 as dense row masks + 7 sparse products A.P = %d "
                                        "thread-instr per candidate (see hunt_alg_lop3_per_candidate); generation, "
                                        "Berlekamp-Massey, votes and the primitivity stage not counted" % alg}
        if "res" in extra:
            line["result_check"] = {"best": extra["res"]["best"],
                                    "stage_counts": [int(v) for v in extra["res"]["stage_counts"]]}
        emit_json(line)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


_REAL_STDOUT = None


def quiet_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version
    banner to stdout at NCCL_DEBUG=VERSION/WARN), so file descriptor 1 is pointed at stderr for the
    whole run and the JSON line goes to a saved duplicate of the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit_json(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--log2-samples", type=int, default=None,
                    help="samples per GPU: default 2^40 for clt_fused, 2^33 (int8) / 2^32 (int16) = 8 GiB for clt_stream")
    ap.add_argument("--workload", default="clt_fused",
                    choices=["clt_fused", "prbs_stream", "prbs_detect", "rnghunt", "prbs_slicer", "prbs_shaped_slicer", "clt_stream",
                             "tx_stream", "lutopt_stream"],
                    help="clt_fused = the headline (BASELINE.json configs[2]); the others are the remaining "
                         "BASELINE configs, same JSON contract, their own metric")
    ap.add_argument("--k", type=int, default=31, help="PRBS order for the prbs_* workloads")
    ap.add_argument("--log2-bits", type=int, default=36, help="PRBS bits per GPU per step")
    ap.add_argument("--log2-steps", type=int, default=28, help="lutopt_stream: LUTOPT steps (32 B each) per GPU per step")
    ap.add_argument("--err-m", type=int, default=10, help="prbs_detect channel error probability 2^-m")
    ap.add_argument("--amplitude", type=int, default=256, help="prbs_slicer: TX bit amplitude (12-bit word)")
    ap.add_argument("--noise-var", type=int, default=8, help="prbs_slicer: noise_var 0..15 (tx.py:75-76; reset value 8)")
    ap.add_argument("--spb", type=int, default=8, help="prbs_slicer: samples per bit (tx.py:36-37: data clock = 1/8)")
    ap.add_argument("--noise-var-stream", type=int, default=-1, help="clt_stream: >= 0 stores int16 g*noise_var")
    ap.add_argument("--hunt-n", type=int, default=192)
    ap.add_argument("--log2-candidates", type=int, default=20, help="rnghunt candidates per GPU per step")
    ap.add_argument("--hunt-log2", type=int, default=HUNT_LOG2,
                    help="default line, secondary.rnghunt: log2 of the fixed candidate set shared by all ranks")
    ap.add_argument("--no-secondary", action="store_true", help="default line without the rnghunt leg")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU baseline sample length")
    ap.add_argument("--ref-step-seconds", type=float, default=4.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--tune", action="append", default=[], metavar="KEY=VALUE",
                    help="bbb_set_tuning knob (experiments only), e.g. clt_spec_variant=1")
    args = ap.parse_args()
    if args.log2_samples is None:
        args.log2_samples = 40 if args.workload != "clt_stream" else (32 if args.noise_var_stream >= 0 else 33)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return 0

    import torch
    import torch.distributed as dist
    from basebandboard_b200 import lib

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # keep stdout to the one JSON line: NCCL writes its version banner (level VERSION and above) to
        # stdout unless told otherwise
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx = lib.Context(local_rank)
    for kv in args.tune:
        key, _, val = kv.partition("=")
        ctx.set_tuning(key, int(val))
    if args.workload != "clt_fused":
        return run_secondary(args, ctx, rank, world, local_rank)
    ctx.lutopt_load_shipped(N_STATE)
    count = 1 << args.log2_samples
    first = rank * count
    acc = torch.zeros(N_STATE + 3, dtype=torch.int64, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        acc.zero_()
        torch.cuda.current_stream().synchronize()
        ctx.clt_hist_dev(1, first, count, lib.MODEL_GATEWARE, acc.data_ptr())
        launches = ctx.last_launches
        if world > 1:
            ctx.sync()
            dist.all_reduce(acc)
        return launches

    lop3_peak = max(ctx.measure_lop3_peak() for _ in range(3))
    sm_count = torch.cuda.get_device_properties(local_rank).multi_processor_count

    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(physical_gpu_index(local_rank))
    sampler.start()
    launches = 0
    kernel_ms = 0.0
    t0 = time.perf_counter()
    ctx.timer_start()
    for _ in range(args.steps):
        launches += step_device()
        kernel_ms += 0.0
    dev_ms = ctx.timer_stop()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.finish()
    barrier()
    # per-launch duration of the dominant kernel, from the library's own events around one call
    acc.zero_()
    torch.cuda.synchronize()
    ctx.clt_hist_dev(1, first, count, lib.MODEL_GATEWARE, acc.data_ptr())
    ctx.sync()
    one_call_ms = ctx.last_device_ms
    kernel_name = ctx.last_kernel
    hist_dev = acc.cpu().numpy().astype("uint64")
    # the reduced result of one step, read back: every rank's shard must be in it
    if world > 1:
        dist.all_reduce(acc)
    acc_all = acc.cpu().numpy().astype("uint64")
    assert int(acc_all[:N_STATE + 1].sum()) == world * count, "all-reduced histogram does not hold world * count samples"

    # e2e: host buffers through bbb_clt_hist, copies inside the timed region
    e2e_steps = max(1, min(args.steps, 2))
    ctx.clt_hist(1, first, min(count, 1 << 24), lib.MODEL_GATEWARE)
    barrier()
    t1 = time.perf_counter()
    h2d = d2h = 0
    for _ in range(e2e_steps):
        h, s1, s2 = ctx.clt_hist(1, first, count, lib.MODEL_GATEWARE)
        h2d, d2h = ctx.last_h2d_bytes + 8 * ((N_STATE + 63) // 64), ctx.last_d2h_bytes
    e2e_ms = (time.perf_counter() - t1) * 1e3 / e2e_steps
    assert int(h.sum()) == count
    if world == 1:
        assert (hist_dev[:N_STATE] == h).all() and int(hist_dev[N_STATE + 2]) == s2, "device and host-API results differ"

    t = torch.tensor([dev_ms, wall_ms, e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms, e2e_ms = [float(v) for v in t.cpu()]

    if rank == 0:
        total = world * count
        value = total * args.steps / (dev_ms * 1e-3) / 1e9
        per_gpu_rate = count / (one_call_ms * 1e-3)
        achieved = per_gpu_rate * ALG_LOP3_PER_STEP32 / 32.0
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32 (GF(2) bit ops)", "data": "synthetic",
            "config": workload_config(args, world),
            "wall_ms_per_step": wall_ms / args.steps,
            "clocks": clocks,
            "e2e": {"value": total / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "api": "bbb_clt_hist (host seed in, host hist/sum/sumsq out)"},
            "gpu_launches": launches,
            "roofline": {"bound": "alu", "kernel": kernel_name, "achieved": achieved / 1e12, "peak": lop3_peak / 1e12,
                         "unit": "T thread-instr/s (LOP3-equivalent)", "frac": achieved / lop3_peak,
                         "traffic": 38887680 if (kernel_name == "clt256_fused2" and not args.tune and sm_count == 148) else None,
                         "traffic_note": "NOT re-measured in this run: dram__bytes_read.sum + dram__bytes_write.sum of one "
                                         "clt256_hist2_kernel launch at this geometry (148 SMs x 2 CTAs x 128 threads, default "
                                         "tuning; ncu --set full, profiles/r1_clt256_fused2_ncu_full.txt) = the 256 seed planes "
                                         "x 37888 threads x 4 B, read once per launch whatever the sample count; 0 B written; "
                                         "null under --tune or another SM count",
                         "launch_ms": one_call_ms,
                         "note": "integer-pipe bound, not HBM/tensor: algorithmic work = 986 LOP3 per 32 "
                                 "samples (492 recurrence + 494 popcount tree); peak = LOP3 issue rate measured "
                                 "live on this GPU (bbb_measure_lop3_peak); HBM traffic ~0 B/sample"},
            "result_check": {"hist_total": int(h.sum()), "sum": int(s1), "sumsq": int(s2),
                             "reduced_hist_total": int(acc_all[:N_STATE + 1].sum()), "reduced_expected": world * count,
                             "reduced_sum": int(acc_all[N_STATE + 1:N_STATE + 2].view("int64")[0]),
                             "reduced_sumsq": int(acc_all[N_STATE + 2]),
                             "reduced_crc32_u64x259": crc_u64(acc_all)},
        }
        if not args.no_cpu_baseline:
            cores = host_cores()
            cal = cpu_port_rate(1 << 21, cores)
            n_cpu = max(1 << 21, int(cal * args.cpu_seconds))
            rate = cpu_port_rate(n_cpu, cores)
            line["cpu_baseline"] = {"value": rate / 1e9, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "first %d samples of the workload (oracle/gen.c, %d threads)" % (n_cpu, cores),
                                    "python_loop": {"value": python_loop_rate() / 1e9, "unit": UNIT, "cores": 1,
                                                    "sample": "4000 samples through the reference's literal loop, "
                                                              "clt-grng-evaluate.py:8-16 (numpy scalar subtractions)"}}
    if not args.no_secondary:
        hunt = run_rnghunt_leg(args, ctx, rank, world, local_rank, lop3_peak, barrier)
        if rank == 0:
            line["secondary"] = {"rnghunt": hunt}
            line["gpu_launches_secondary"] = hunt["gpu_launches"]
    if rank == 0:
        emit_json(line)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
