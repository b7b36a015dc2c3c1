"""ctypes binding of libbbb_b200.so (the C ABI of include/bbb_b200.h).

There is no CPU fallback: if the library is missing, or no B200 is visible, the product path
raises.  `build()` (also called by __graft_entry__.build) compiles the library in-tree with nvcc
for sm_100a.
"""
import ctypes as C
import functools
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
# BBB_B200_LIB: kernel experiments only (another build of the same library, e.g. build/variants/*.so)
SO_PATH = os.environ.get("BBB_B200_LIB") or os.path.join(_HERE, "libbbb_b200.so")

MODEL_GATEWARE, MODEL_NUMPY = 0, 1
STAGE_NAMES = ("degree", "filter", "order", "factor", "primitive")
STAGE_PRIMITIVE = 4
TX_SRC_PULSE = 0          # bbb_tx_stream(k = BBB_TX_SRC_PULSE): the Pulser source of tx.py:20-31,59-65


class BBBError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("bbb_b200 error %d: %s" % (code, msg))
        self.code = code


class DetState(C.Structure):
    _fields_ = [("bit_in", C.c_uint32), ("prbs", C.c_uint32), ("err_sr", C.c_uint32),
                ("reload_ctr", C.c_uint32)]


class DetCounters(C.Structure):
    _fields_ = [("err", C.c_uint64), ("err_valid", C.c_uint64), ("reload_events", C.c_uint64),
                ("reload_cycles", C.c_uint64), ("injected", C.c_uint64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


_P = C.POINTER
_u64p, _u32p, _u16p, _u8p = _P(C.c_uint64), _P(C.c_uint32), _P(C.c_uint16), _P(C.c_uint8)
_ctx = C.c_void_p

# name -> (restype, argtypes); every function include/bbb_b200.h declares
SIGNATURES = {
    "bbb_version": (C.c_int, []),
    "bbb_device_count": (C.c_int, []),
    "bbb_create": (C.c_int, [C.c_int, _P(_ctx)]),
    "bbb_destroy": (None, [_ctx]),
    "bbb_last_error": (C.c_char_p, [_ctx]),
    "bbb_last_device_ms": (C.c_double, [_ctx]),
    "bbb_last_launches": (C.c_uint32, [_ctx]),
    "bbb_last_kernel": (C.c_char_p, [_ctx]),
    "bbb_sync": (C.c_int, [_ctx]),
    "bbb_last_h2d_bytes": (C.c_uint64, [_ctx]),
    "bbb_last_d2h_bytes": (C.c_uint64, [_ctx]),
    "bbb_set_tuning": (C.c_int, [_ctx, C.c_char_p, C.c_int64]),
    "bbb_timer_start": (C.c_int, [_ctx]),
    "bbb_timer_stop": (C.c_int, [_ctx, _P(C.c_double)]),
    "bbb_lutopt_load": (C.c_int, [_ctx, C.c_uint32, _u32p, _u16p]),
    "bbb_lutopt_load_shipped": (C.c_int, [_ctx, C.c_uint32]),
    "bbb_lutopt_get": (C.c_int, [_ctx, _u32p, _u32p, _u32p, _u16p]),
    "bbb_lutopt_jump": (C.c_int, [_ctx, _u64p, C.c_uint64, _u64p]),
    "bbb_lutopt_stream": (C.c_int, [_ctx, _u64p, C.c_uint64, C.c_uint64, _u64p]),
    "bbb_lutopt_stream_dev": (C.c_int, [_ctx, _u64p, C.c_uint64, C.c_uint64, C.c_void_p]),
    "bbb_clt_hist": (C.c_int, [_ctx, _u64p, C.c_uint64, C.c_uint64, C.c_uint32, _u64p,
                               _P(C.c_int64), _u64p]),
    "bbb_clt_hist_dev": (C.c_int, [_ctx, _u64p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p]),
    "bbb_clt_hist_dev_static": (C.c_int, [_ctx, _u64p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p]),
    "bbb_clt_stream": (C.c_int, [_ctx, _u64p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32,
                                 _P(C.c_int8), _P(C.c_int16)]),
    "bbb_clt_stream_dev": (C.c_int, [_ctx, _u64p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32,
                                     C.c_void_p, C.c_void_p]),
    "bbb_clt_stream_dec": (C.c_int, [_ctx, _u64p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, _P(C.c_int8)]),
    "bbb_clt_stream_dec_dev": (C.c_int, [_ctx, _u64p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p]),
    "bbb_clt_value": (C.c_int, [C.c_uint32, _u64p, C.c_uint32, _P(C.c_int32)]),
    "bbb_clt_stats": (C.c_int, [C.c_uint32, C.c_uint32, _u64p, _P(C.c_double)]),
    "bbb_prbs_tap": (C.c_uint32, [C.c_uint32]),
    "bbb_prbs_stream": (C.c_int, [_ctx, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint64, _u64p]),
    "bbb_prbs_stream_dev": (C.c_int, [_ctx, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint64,
                                      C.c_void_p]),
    "bbb_prbs_detect": (C.c_int, [_ctx, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint32,
                                  C.c_uint64, _P(DetState), _P(DetCounters)]),
    "bbb_prbs_detect_errs_dev": (C.c_int, [_ctx, C.c_uint32, C.c_uint64, C.c_uint64, C.c_void_p,
                                           _P(DetState), _P(DetCounters)]),
    "bbb_prbs_slicer_detect": (C.c_int, [_ctx, C.c_uint32, C.c_uint64, C.c_uint64, _u64p, C.c_uint32, C.c_uint32,
                                         C.c_uint32, C.c_uint32, _P(DetState), _P(DetCounters)]),
    "bbb_tx_stream": (C.c_int, [_ctx, C.c_uint32, C.c_uint64, C.c_uint64, _P(C.c_int16), C.c_uint32, _u64p, C.c_uint32,
                                _P(C.c_int16)]),
    "bbb_tx_stream_dev": (C.c_int, [_ctx, C.c_uint32, C.c_uint64, C.c_uint64, _P(C.c_int16), C.c_uint32, _u64p,
                                    C.c_uint32, C.c_void_p]),
    "bbb_prbs_shaped_slicer_detect": (C.c_int, [_ctx, C.c_uint32, C.c_uint64, C.c_uint64, _P(C.c_int16), _u64p,
                                                C.c_uint32, C.c_uint32, _P(DetState), _P(DetCounters)]),
    "bbb_prbs_spec_state": (C.c_int, [_ctx, C.c_uint32, C.c_uint64, C.c_uint32, C.c_uint64, _P(DetState)]),
    "bbb_last_rounds": (C.c_uint32, [_ctx]),
    "bbb_rnghunt_eval": (C.c_int, [_ctx, C.c_uint32, C.c_uint64, _u64p, _u8p, _u64p]),
    "bbb_rnghunt_candidate": (C.c_int, [_ctx, C.c_uint32, C.c_uint64, C.c_uint64, _u64p]),
    "bbb_rnghunt_sweep": (C.c_int, [_ctx, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint64,
                                    C.c_uint64, C.c_uint32, _u64p, _u32p, _u64p, _u8p]),
    "bbb_rnghunt_sweep_polys": (C.c_int, [_ctx, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint64,
                                          C.c_uint64, C.c_uint32, _u64p, _u32p, _u64p, _u64p, _u32p]),
    "bbb_measure_lop3_peak": (C.c_int, [_ctx, _P(C.c_double)]),
}

_lib = None


def build(verbose=False):
    """Compile libbbb_b200.so in-tree (nvcc, sm_100a).  Cross-compiles without a GPU."""
    res = subprocess.run(["make", "-C", ROOT, "-j8", "basebandboard_b200/libbbb_b200.so"],
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or res.returncode:
        print(res.stdout)
    if res.returncode:
        raise RuntimeError("building libbbb_b200.so failed")
    return SO_PATH


def load():
    """dlopen the library and attach prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError("%s is missing: run `make` (there is no CPU fallback)" % SO_PATH)
    L = C.CDLL(SO_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def state_words(value, n):
    """Python int (bit i = weight 2^i, rng.py:135) -> little-endian u64 words."""
    nw = (n + 63) // 64
    return np.array([(int(value) >> (64 * w)) & 0xFFFFFFFFFFFFFFFF for w in range(nw)],
                    dtype=np.uint64)


def state_int(words):
    return sum(int(w) << (64 * i) for i, w in enumerate(words))


def _ptr(a, typ):
    return a.ctypes.data_as(_P(typ))


def _check_out(out, dtype, size):
    if not (isinstance(out, np.ndarray) and out.dtype == dtype and out.size == size and out.flags.c_contiguous):
        raise ValueError("out must be a C-contiguous %s array of %d elements" % (np.dtype(dtype).name, size))
    return out


class Context:
    """One per GPU per host thread (the `bbb_ctx*` of the C ABI)."""

    def __init__(self, device=0):
        self._L = load()
        h = _ctx()
        rc = self._L.bbb_create(int(device), C.byref(h))
        if rc != 0:
            raise BBBError(rc, "bbb_create(device=%d) failed: no usable sm_100 GPU" % device)
        self._h = h
        self.device = int(device)
        self.n = 0

    def close(self):
        if getattr(self, "_h", None):
            self._L.bbb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc):
        if rc != 0:
            raise BBBError(rc, self._L.bbb_last_error(self._h).decode())

    # ---- measurement -----------------------------------------------------------------------
    def sync(self):
        self._check(self._L.bbb_sync(self._h))

    def set_tuning(self, key, value):
        self._check(self._L.bbb_set_tuning(self._h, key.encode(), int(value)))

    def timer_start(self):
        self._check(self._L.bbb_timer_start(self._h))

    def timer_stop(self):
        """Device milliseconds (CUDA events on the ctx's stream) since timer_start()."""
        ms = C.c_double(0)
        self._check(self._L.bbb_timer_stop(self._h, C.byref(ms)))
        return float(ms.value)

    @property
    def last_h2d_bytes(self):
        return int(self._L.bbb_last_h2d_bytes(self._h))

    @property
    def last_d2h_bytes(self):
        return int(self._L.bbb_last_d2h_bytes(self._h))

    @property
    def last_device_ms(self):
        return float(self._L.bbb_last_device_ms(self._h))

    @property
    def last_launches(self):
        return int(self._L.bbb_last_launches(self._h))

    @property
    def last_kernel(self):
        return self._L.bbb_last_kernel(self._h).decode()

    def measure_lop3_peak(self):
        v = C.c_double(0)
        self._check(self._L.bbb_measure_lop3_peak(self._h, C.byref(v)))
        return float(v.value)

    # ---- LUTOPT ----------------------------------------------------------------------------
    def lutopt_load(self, packed_rows):
        from .recurrences import to_csr
        rp, tp = to_csr(packed_rows)
        self._check(self._L.bbb_lutopt_load(self._h, len(packed_rows), _ptr(rp, C.c_uint32),
                                            _ptr(tp, C.c_uint16)))
        self.n = len(packed_rows)

    def lutopt_load_shipped(self, n):
        self._check(self._L.bbb_lutopt_load_shipped(self._h, n))
        self.n = n

    def lutopt_jump(self, x, e):
        xin = state_words(x, self.n)
        out = np.zeros_like(xin)
        self._check(self._L.bbb_lutopt_jump(self._h, _ptr(xin, C.c_uint64), e, _ptr(out, C.c_uint64)))
        return state_int(out)

    def lutopt_stream(self, init, first, count, out=None):
        """`out`: caller-owned uint64 array of count * ceil(n/64) words to fill (reused across calls)."""
        nw = (self.n + 63) // 64
        out = np.zeros(count * nw, dtype=np.uint64) if out is None else _check_out(out, np.uint64, count * nw)
        x0 = state_words(init, self.n)
        self._check(self._L.bbb_lutopt_stream(self._h, _ptr(x0, C.c_uint64), first, count,
                                              _ptr(out, C.c_uint64)))
        return out.reshape(count, nw)

    def lutopt_stream_dev(self, init, first, count, dev_ptr):
        x0 = state_words(init, self.n)
        self._check(self._L.bbb_lutopt_stream_dev(self._h, _ptr(x0, C.c_uint64), first, count,
                                                  C.c_void_p(dev_ptr)))

    # ---- CLT -------------------------------------------------------------------------------
    def clt_hist(self, init, first, count, model=MODEL_GATEWARE):
        n = self.n
        hist = np.zeros(n + (1 if model == MODEL_NUMPY else 0), dtype=np.uint64)
        s, ss = C.c_int64(0), C.c_uint64(0)
        x0 = state_words(init, n)
        self._check(self._L.bbb_clt_hist(self._h, _ptr(x0, C.c_uint64), first, count, model,
                                         _ptr(hist, C.c_uint64), C.byref(s), C.byref(ss)))
        return hist, int(s.value), int(ss.value)

    def clt_hist_dev(self, init, first, count, model, dev_ptr):
        """Adds hist[n+1] | sum | sumsq (u64[n+3]) into device memory; asynchronous."""
        x0 = state_words(init, self.n)
        self._check(self._L.bbb_clt_hist_dev(self._h, _ptr(x0, C.c_uint64), first, count, model,
                                             C.c_void_p(dev_ptr)))

    def clt_hist_dev_static(self, init, first, count, model, dev_ptr):
        """clt_hist_dev through the no-allocation entry point (static device scratch, count <= 2^32)."""
        x0 = state_words(init, self.n)
        self._check(self._L.bbb_clt_hist_dev_static(self._h, _ptr(x0, C.c_uint64), first, count, model,
                                                    C.c_void_p(dev_ptr)))

    def clt_stream(self, init, first, count, model=MODEL_GATEWARE, noise_var=None, out=None):
        x0 = state_words(init, self.n)
        if noise_var is None:
            out = np.zeros(count, dtype=np.int8) if out is None else _check_out(out, np.int8, count)
            self._check(self._L.bbb_clt_stream(self._h, _ptr(x0, C.c_uint64), first, count, model, 1,
                                               _ptr(out, C.c_int8), None))
        else:
            out = np.zeros(count, dtype=np.int16) if out is None else _check_out(out, np.int16, count)
            self._check(self._L.bbb_clt_stream(self._h, _ptr(x0, C.c_uint64), first, count, model,
                                               noise_var, None, _ptr(out, C.c_int16)))
        return out

    def clt_stream_dev(self, init, first, count, model, noise_var, dev_ptr, elem_bytes):
        x0 = state_words(init, self.n)
        p8 = C.c_void_p(dev_ptr) if elem_bytes == 1 else None
        p16 = C.c_void_p(dev_ptr) if elem_bytes == 2 else None
        self._check(self._L.bbb_clt_stream_dev(self._h, _ptr(x0, C.c_uint64), first, count, model,
                                               noise_var, p8, p16))

    def clt_stream_dec(self, init, first, count, spb, delay):
        """out[j] = g[first + j*spb + delay]: the sample of each bit period the receiver keeps (rx.py:18-19)."""
        x0 = state_words(init, self.n)
        out = np.zeros(count, dtype=np.int8)
        self._check(self._L.bbb_clt_stream_dec(self._h, _ptr(x0, C.c_uint64), first, count, spb, delay,
                                               _ptr(out, C.c_int8)))
        return out

    def clt_stream_dec_dev(self, init, first, count, spb, delay, dev_ptr):
        x0 = state_words(init, self.n)
        self._check(self._L.bbb_clt_stream_dec_dev(self._h, _ptr(x0, C.c_uint64), first, count, spb, delay,
                                                   C.c_void_p(dev_ptr)))

    # ---- PRBS ------------------------------------------------------------------------------
    def prbs_stream(self, k, nbits, state0=1, first_bit=0, out=None):
        out = np.zeros(nbits // 64, dtype=np.uint64) if out is None else _check_out(out, np.uint64, nbits // 64)
        self._check(self._L.bbb_prbs_stream(self._h, k, state0, first_bit, nbits,
                                            _ptr(out, C.c_uint64)))
        return out

    def prbs_stream_dev(self, k, nbits, dev_ptr, state0=1, first_bit=0):
        self._check(self._L.bbb_prbs_stream_dev(self._h, k, state0, first_bit, nbits,
                                                C.c_void_p(dev_ptr)))

    def prbs_detect(self, k, nbits, m=0, seed=1, first_bit=0, det=None):
        ctr = DetCounters()
        self._check(self._L.bbb_prbs_detect(self._h, k, first_bit, nbits, m, seed,
                                            C.byref(det) if det is not None else None,
                                            C.byref(ctr)))
        return ctr.as_dict()

    def prbs_detect_errs_dev(self, k, nbits, err_words_dev_ptr, first_bit=0, det=None):
        """PRBSErrorDetector (prbs.py:43-99) behind a caller-supplied error pattern (device u64 words)."""
        ctr = DetCounters()
        self._check(self._L.bbb_prbs_detect_errs_dev(self._h, k, first_bit, nbits, C.c_void_p(err_words_dev_ptr),
                                                     C.byref(det) if det is not None else None, C.byref(ctr)))
        return ctr.as_dict()

    def prbs_slicer_detect(self, k, nbits, init=1, amplitude=512, noise_var=8, spb=8, delay=0, first_bit=0,
                           det=None):
        """TX (PRBS +-amplitude + CLT-GRNG * noise_var, tx.py:70-81) -> slicer (rx.py:28) -> detector."""
        ctr = DetCounters()
        x0 = state_words(init, self.n)
        self._check(self._L.bbb_prbs_slicer_detect(self._h, k, first_bit, nbits, _ptr(x0, C.c_uint64), amplitude,
                                                   noise_var, spb, delay,
                                                   C.byref(det) if det is not None else None, C.byref(ctr)))
        return ctr.as_dict()

    def tx_stream(self, k, nbits, coeff, first_bit=0, bit_en=1, init=None, noise_var=0, dev_ptr=None, out=None):
        """The 12-bit TX words (tx.py:52-81: shaped PRBS + CLT-GRNG * noise_var), 8 per bit, as int16.
        init None = noise off.  dev_ptr: write to device memory instead of returning an array."""
        cf = np.ascontiguousarray(coeff, dtype=np.int16)
        assert cf.size == 64
        x0 = state_words(init, self.n) if init is not None else None
        px0 = _ptr(x0, C.c_uint64) if x0 is not None else None
        if dev_ptr is not None:
            self._check(self._L.bbb_tx_stream_dev(self._h, k, first_bit, nbits, _ptr(cf, C.c_int16), bit_en, px0,
                                                  noise_var, C.c_void_p(dev_ptr)))
            return None
        out = np.zeros(nbits * 8, dtype=np.int16) if out is None else _check_out(out, np.int16, nbits * 8)
        self._check(self._L.bbb_tx_stream(self._h, k, first_bit, nbits, _ptr(cf, C.c_int16), bit_en, px0, noise_var,
                                          _ptr(out, C.c_int16)))
        return out

    def prbs_shaped_slicer_detect(self, k, nbits, coeff, init=1, noise_var=8, delay=4, first_bit=0, det=None):
        """TX through the pulse shaper (bitshaper.py) + noise -> slicer at sample `delay` -> detector."""
        ctr = DetCounters()
        cf = np.ascontiguousarray(coeff, dtype=np.int16)
        x0 = state_words(init, self.n)
        self._check(self._L.bbb_prbs_shaped_slicer_detect(self._h, k, first_bit, nbits, _ptr(cf, C.c_int16),
                                                          _ptr(x0, C.c_uint64), noise_var, delay,
                                                          C.byref(det) if det is not None else None, C.byref(ctr)))
        return ctr.as_dict()

    def prbs_spec_state(self, k, first_bit, m=0, seed=1):
        return prbs_spec_state(k, first_bit, m, seed)

    @property
    def last_rounds(self):
        return int(self._L.bbb_last_rounds(self._h))

    # ---- rnghunt ---------------------------------------------------------------------------
    def rnghunt_eval(self, n, matrices, want_poly=False):
        """matrices: uint64 array [ncand, n*wpc] in BinaryMatrix layout."""
        m = np.ascontiguousarray(matrices, dtype=np.uint64)
        wpc = (n + 63) // 64
        ncand = m.size // (n * wpc)
        stage = np.zeros(ncand, dtype=np.uint8)
        pw = (n + 1 + 63) // 64
        poly = np.zeros(ncand * pw, dtype=np.uint64) if want_poly else None
        self._check(self._L.bbb_rnghunt_eval(self._h, n, ncand, _ptr(m, C.c_uint64),
                                             _ptr(stage, C.c_uint8),
                                             _ptr(poly, C.c_uint64) if want_poly else None))
        return (stage, poly.reshape(ncand, pw)) if want_poly else stage

    def rnghunt_candidate(self, n, seed, c):
        wpc = (n + 63) // 64
        out = np.zeros(n * wpc, dtype=np.uint64)
        self._check(self._L.bbb_rnghunt_candidate(self._h, n, seed, c, _ptr(out, C.c_uint64)))
        return out

    def rnghunt_sweep(self, n, seed, first, count, k=16, stride=1, want_stages=False, want_polys=False):
        best = np.zeros(max(k, 1), dtype=np.uint64)
        nfound = C.c_uint32(0)
        counts = np.zeros(5, dtype=np.uint64)
        if want_polys:
            pw = (n + 1 + 63) // 64
            polys = np.zeros(max(k, 1) * pw, dtype=np.uint64)
            weights = np.zeros(max(k, 1), dtype=np.uint32)
            self._check(self._L.bbb_rnghunt_sweep_polys(self._h, n, seed, first, stride, count, k,
                                                        _ptr(best, C.c_uint64), C.byref(nfound),
                                                        _ptr(counts, C.c_uint64), _ptr(polys, C.c_uint64),
                                                        _ptr(weights, C.c_uint32)))
            nf = nfound.value
            return {"best": [int(b) for b in best[:nf]], "stage_counts": counts,
                    "polys": polys.reshape(-1, pw)[:nf], "weights": [int(w) for w in weights[:nf]]}
        stages = np.zeros(count, dtype=np.uint8) if want_stages else None
        self._check(self._L.bbb_rnghunt_sweep(self._h, n, seed, first, stride, count, k,
                                              _ptr(best, C.c_uint64), C.byref(nfound),
                                              _ptr(counts, C.c_uint64),
                                              _ptr(stages, C.c_uint8) if want_stages else None))
        res = {"best": [int(b) for b in best[:nfound.value]], "stage_counts": counts}
        if want_stages:
            res["stages"] = stages
        return res


@functools.lru_cache(maxsize=1024)
def _prbs_spec_state_tuple(k, first_bit, m, seed):
    det = DetState()
    rc = load().bbb_prbs_spec_state(None, k, first_bit, m, seed, C.byref(det))
    if rc != 0:
        raise BBBError(rc, "bbb_prbs_spec_state(k=%d, first_bit=%d, m=%d)" % (k, first_bit, m))
    return (int(det.bit_in), int(det.prbs), int(det.err_sr), int(det.reload_ctr))


def prbs_spec_state(k, first_bit, m=0, seed=1):
    """State of a locked detector just before bit `first_bit` (bbb_prbs_spec_state; host arithmetic,
    needs no GPU).  Memoised: a sharded loop asks for the same boundary every step, and the call rebuilds the
    k x k jump tables."""
    return DetState(*_prbs_spec_state_tuple(int(k), int(first_bit), int(m), int(seed)))


def clt_value(n, x, model=MODEL_GATEWARE):
    """CLTGRNG of one URNG word (bbb_clt_value; host arithmetic, needs no GPU)."""
    w = state_words(x, n)
    g = C.c_int32(0)
    rc = load().bbb_clt_value(n, _ptr(w, C.c_uint64), model, C.byref(g))
    if rc != 0:
        raise BBBError(rc, "bbb_clt_value(n=%d)" % n)
    return int(g.value)


def hw_latency_split(n, first, count):
    """The hardware's output sequence (rng.py:70-72,189-190): hw[h] = 0 for h < log2(n) (reset registers),
    hw[log2 n] = clt(x0), hw[log2 n + 1 + i] = g[i].  For samples hw[first, first+count) returns
    (number of leading zeros, whether clt(x0) is included, first g index, number of g samples)."""
    logn = n.bit_length() - 1
    end = first + count
    zeros = max(0, min(end, logn) - first)
    has_x0 = first <= logn < end
    g_lo = max(first, logn + 1)
    g_cnt = max(0, end - g_lo)
    return zeros, has_x0, g_lo - (logn + 1), g_cnt


def clt_stats(n, hist, model=MODEL_GATEWARE):
    """mean, variance (ddof=0), excess kurtosis, chi2 vs N(0,n/4), chi2 vs Binomial, theory variance."""
    h = np.ascontiguousarray(hist, dtype=np.uint64)
    out = (C.c_double * 6)()
    rc = load().bbb_clt_stats(n, model, _ptr(h, C.c_uint64), out)
    if rc != 0:
        raise BBBError(rc, "bbb_clt_stats")
    return dict(zip(("mean", "variance", "kurtosis", "chi2_normal", "chi2_binomial", "variance_theory"),
                    [float(v) for v in out]))
