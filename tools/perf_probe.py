#!/usr/bin/env python3
"""Ad-hoc device timings of the secondary paths (PRBS stream/detect, materialised CLT/LUTOPT
streams, rnghunt sweep).  Prints one JSON line per measurement.  usage: tools/perf_probe.py [what ...]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from basebandboard_b200 import lib  # noqa: E402


def main():
    what = sys.argv[1:] or ["prbs_stream", "prbs_detect"]
    c = lib.Context(0)
    if "prbs_stream" in what:
        nbits = 1 << 36
        buf = torch.empty(nbits // 64, dtype=torch.int64, device="cuda")
        for k in (7, 15, 23, 31):
            for _ in range(3):
                c.prbs_stream_dev(k, nbits, buf.data_ptr())
            c.sync()
            ms = []
            for _ in range(5):
                c.prbs_stream_dev(k, nbits, buf.data_ptr())
                c.sync()
                ms.append(c.last_device_ms)
            best, med = min(ms), sorted(ms)[2]
            print(json.dumps({"what": "prbs_stream", "k": k, "bits": nbits, "ms_median": med, "ms_best": best,
                              "GBps_median": nbits / 8 / med / 1e6, "Tbit_s": nbits / med / 1e9}), flush=True)
        del buf
    if "prbs_detect" in what:
        for m in (10, 0, 4):
            for k in (7, 15, 23, 31):
                nbits = 1 << 36 if m else 1 << 38
                c.prbs_detect(k, 1 << 30, m=m, seed=1)
                r = c.prbs_detect(k, nbits, m=m, seed=1)
                ms = c.last_device_ms
                print(json.dumps({"what": "prbs_detect", "k": k, "m": m, "bits": nbits, "ms": ms,
                                  "rounds": c.last_rounds, "launches": c.last_launches,
                                  "Tbit_s": nbits / ms / 1e9, "counters": r}), flush=True)
    if "streams" in what:
        c.lutopt_load_shipped(256)
        for name, count, bytes_per in (("clt_stream_i8", 1 << 33, 1), ("clt_stream_i16", 1 << 32, 2),
                                       ("lutopt_stream", 1 << 28, 32), ("lutopt_stream_v5", 1 << 28, 32)):
            c.set_tuning("words_variant", 5 if name.endswith("v5") else 0)
            buf = torch.empty(count * bytes_per, dtype=torch.uint8, device="cuda")
            ms = []
            for _ in range(3):
                if name.startswith("lutopt_stream"):
                    c.lutopt_stream_dev(1, 0, count, buf.data_ptr())
                else:
                    c.clt_stream_dev(1, 0, count, 0, 8 if bytes_per == 2 else 1, buf.data_ptr(), bytes_per)
                c.sync()
                ms.append(c.last_device_ms)
            print(json.dumps({"what": name, "count": count, "bytes": count * bytes_per, "ms": ms,
                              "kernel": c.last_kernel, "launches": c.last_launches,
                              "GBps_best": count * bytes_per / min(ms) / 1e6,
                              "Gsamples_s": count / min(ms) / 1e6}), flush=True)
            del buf
    if "rnghunt" in what:
        for name, knobs in (("pow8", {}), ("pow4", {"hunt_pow_k": 4}), ("warp1", {"hunt_smem_taps": 2})):
            for k, v in knobs.items():
                c.set_tuning(k, v)
            for n in (64, 128, 192, 256):
                cnt = 1 << 22 if n <= 192 else 1 << 20
                c.rnghunt_sweep(n, 1, 0, 1 << 14, k=16)
                r = c.rnghunt_sweep(n, 1, 0, cnt, k=16)
                ms = c.last_device_ms
                print(json.dumps({"what": "rnghunt_sweep", "n": n, "kernel": name, "candidates": cnt, "ms": ms,
                                  "launches": c.last_launches, "cand_per_s": cnt / ms * 1e3,
                                  "stage_counts": [int(v) for v in r["stage_counts"]], "best": r["best"][:4]}), flush=True)
            for k in knobs:
                c.set_tuning(k, 0)
    c.close()


if __name__ == "__main__":
    main()
