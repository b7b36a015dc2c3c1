"""ctypes front-end of the CPU ORACLE (test infrastructure, NOT the product path).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  It loads oracle/_build/liboracle.so (built from gf2.c, gen.c,
rnghunt_ref.c by oracle/Makefile; built on demand with gcc when missing) and exposes

  * BinaryVector / BinaryMatrix / BinaryPolynomial / berlekamp_massey with the reference
    crate's semantics (software/rnghunt/src/lib.rs:12-21),
  * the generator models (LUTOPT, CLT tree, PRBS, PRBSErrorDetector) of gateware/bbb,
  * numpy restatements of the reference's own model lines (rng.py:134-135, :173-181,
    prbs.py:112-113) used to cross-check the C code.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None

MODEL_GATEWARE = 0
MODEL_NUMPY = 1
STAGE_NAMES = ("degree", "filter", "order", "factor", "primitive")


class _BVec(C.Structure):
    _fields_ = [("n", C.c_size_t), ("len", C.c_size_t), ("data", C.POINTER(C.c_uint64))]


class _BMat(C.Structure):
    _fields_ = [("nrows", C.c_size_t), ("ncols", C.c_size_t), ("words_per_col", C.c_size_t),
                ("data", C.POINTER(C.c_uint64))]


class _BPoly(C.Structure):
    _fields_ = [("c", _BVec)]


class DetState(C.Structure):
    _fields_ = [("bit_in", C.c_uint32), ("prbs", C.c_uint32), ("err_sr", C.c_uint32),
                ("reload_ctr", C.c_uint32)]


class DetCounters(C.Structure):
    _fields_ = [("err", C.c_uint64), ("err_valid", C.c_uint64), ("reload_events", C.c_uint64),
                ("reload_cycles", C.c_uint64), ("injected", C.c_uint64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("gf2.c", "gen.c", "rnghunt_ref.c", "gf2.h", "gen.h",
                                             "rnghunt_ref.h", "mersenne_factors.inc", "Makefile")]
    stale = (not os.path.exists(_SO)) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if force or stale:
        subprocess.run(["make", "-C", _HERE], check=True, stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is not None:
        return _lib
    # On the GPU box the prebuilt .so travels with the snapshot; rebuild only if absent/stale
    # and a compiler is present.
    try:
        build()
    except (OSError, subprocess.CalledProcessError):
        if not os.path.exists(_SO):
            raise
    L = C.CDLL(_SO)
    P = C.POINTER
    u64p, u32p, u16p, u8p = P(C.c_uint64), P(C.c_uint32), P(C.c_uint16), P(C.c_uint8)
    bvp, bmp, bpp = P(_BVec), P(_BMat), P(_BPoly)
    sig = {
        "bv_alloc": (_BVec, [C.c_size_t, C.c_size_t]),
        "bv_free": (None, [bvp]),
        "bv_clone": (_BVec, [bvp]),
        "bv_get": (C.c_int, [bvp, C.c_size_t]),
        "bv_from_words": (_BVec, [C.c_size_t, u64p, C.c_size_t]),
        "bv_from_bits": (_BVec, [u8p, C.c_size_t]),
        "bv_slice": (_BVec, [bvp, C.c_size_t, C.c_size_t]),
        "bv_firstbit": (C.c_size_t, [bvp]),
        "bv_count_ones": (C.c_uint32, [bvp]),
        "bv_and": (_BVec, [bvp, bvp]), "bv_and_assign": (None, [bvp, bvp]),
        "bv_or": (_BVec, [bvp, bvp]), "bv_or_assign": (None, [bvp, bvp]),
        "bv_xor": (_BVec, [bvp, bvp]), "bv_xor_assign": (None, [bvp, bvp]),
        "bv_shl": (_BVec, [bvp, C.c_size_t]), "bv_shl_assign": (None, [bvp, C.c_size_t]),
        "bv_shr": (_BVec, [bvp, C.c_size_t]), "bv_shr_assign": (None, [bvp, C.c_size_t]),
        "bv_eq": (C.c_int, [bvp, bvp]),
        "bm_from_words": (_BMat, [C.c_size_t, C.c_size_t, u64p]),
        "bm_free": (None, [bmp]),
        "bm_col": (_BVec, [bmp, C.c_size_t]),
        "bm_row": (_BVec, [bmp, C.c_size_t]),
        "bm_dot": (_BVec, [bmp, bvp]),
        "bm_recur": (_BVec, [bmp, bvp, C.c_size_t]),
        "bp_free": (None, [bpp]),
        "bp_degree": (C.c_long, [bpp]),
        "bp_eval": (C.c_int, [bpp, bvp]),
        "bp_modmult": (_BPoly, [bpp, bpp, bpp]),
        "bp_modexp": (_BPoly, [bpp, bvp]),
        "bp_check_integer": (C.c_int, [bpp, bvp]),
        "bp_is_primitive": (C.c_int, [bpp]),
        "bp_display": (C.c_size_t, [bpp, C.c_char_p, C.c_size_t]),
        "gf2_berlekamp_massey": (_BPoly, [bvp]),
        "gf2_factor_count": (C.c_size_t, [C.c_size_t]),
        "gf2_factor_words": (C.c_size_t, [C.c_size_t, C.c_size_t, P(u64p)]),
        "orc_lutopt_step": (None, [C.c_uint32, u32p, u16p, u64p, u64p]),
        "orc_lutopt_jump": (None, [C.c_uint32, u32p, u16p, C.c_uint64, u64p, u64p]),
        "orc_clt": (C.c_int32, [C.c_uint32, u64p, C.c_int]),
        "orc_clt_run": (None, [C.c_uint32, u32p, u16p, u64p, C.c_uint64, C.c_int, C.c_uint32,
                               u64p, P(C.c_int64), u64p, P(C.c_int8), P(C.c_int16), u64p]),
        "orc_clt_run_mt": (None, [C.c_uint32, u32p, u16p, u64p, C.c_uint64, C.c_uint64, C.c_int,
                                  u64p, P(C.c_int64), u64p, C.c_int]),
        "orc_prbs_tap": (C.c_uint32, [C.c_uint32]),
        "orc_prbs_bits": (None, [C.c_uint32, u32p, C.c_uint64, u64p]),
        "orc_prbs_advance": (C.c_uint32, [C.c_uint32, C.c_uint32, C.c_uint64]),
        "orc_err_word": (C.c_uint64, [C.c_uint64, C.c_uint64, C.c_uint32]),
        "orc_det_reset": (None, [C.c_uint32, P(DetState)]),
        "orc_det_clock": (C.c_uint32, [C.c_uint32, P(DetState), C.c_uint32]),
        "orc_det_run_bits": (None, [C.c_uint32, P(DetState), P(C.c_uint64), P(C.c_uint64), C.c_uint64,
                                    P(DetCounters)]),
        "orc_prbs_detect": (None, [C.c_uint32, u32p, P(DetState), C.c_uint64, C.c_uint64,
                                   C.c_uint32, C.c_uint64, P(DetCounters)]),
        "orc_rnghunt_candidate": (_BMat, [C.c_uint32, C.c_uint64, C.c_uint64]),
        "orc_rnghunt_eval": (C.c_int, [bmp, u64p, C.c_uint32]),
        "orc_rnghunt_eval_list": (None, [C.c_uint32, C.c_uint64, u64p, C.c_uint64, u8p, u64p, C.c_int]),
        "orc_rnghunt_sweep": (None, [C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint64, u8p, u64p,
                                     C.c_int]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def _u64p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint64))


def _u32p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32))


def _u16p(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint16))


# --------------------------------------------------------------------------------------------
# The reference crate's types (lib.rs:12-21)
# --------------------------------------------------------------------------------------------
class BinaryVector:
    def __init__(self, raw):
        self._v = raw

    def __del__(self):
        try:
            lib().bv_free(C.byref(self._v))
        except Exception:
            pass

    # constructors -------------------------------------------------------------------------
    @classmethod
    def from_words(cls, n, words):
        w = np.asarray(words, dtype=np.uint64)
        return cls(lib().bv_from_words(n, _u64p(w), len(w)))

    @classmethod
    def raw(cls, n, words):
        """`BinaryVector { n, data: vec![...] }` literal."""
        return cls.from_words(n, words)

    @classmethod
    def from_bits(cls, bits):
        b = np.asarray(bits, dtype=np.uint8)
        return cls(lib().bv_from_bits(b.ctypes.data_as(C.POINTER(C.c_uint8)), len(b)))

    @classmethod
    def from_bitstring(cls, s):
        return cls.from_bits([int(c) for c in s])

    # accessors ----------------------------------------------------------------------------
    @property
    def n(self):
        return int(self._v.n)

    @property
    def data(self):
        return [int(self._v.data[i]) for i in range(self._v.len)]

    def __getitem__(self, i):
        return bool(lib().bv_get(C.byref(self._v), i))

    def to_bits(self):
        return [int(lib().bv_get(C.byref(self._v), i)) for i in range(self.n)]

    def to_bitstring(self):
        return "".join(str(b) for b in self.to_bits())

    def slice(self, start, end):
        return BinaryVector(lib().bv_slice(C.byref(self._v), start, end))

    def firstbit(self):
        return int(lib().bv_firstbit(C.byref(self._v)))

    def count_ones(self):
        return int(lib().bv_count_ones(C.byref(self._v)))

    def clone(self):
        return BinaryVector(lib().bv_clone(C.byref(self._v)))

    # operators ----------------------------------------------------------------------------
    def _bin(self, name, other):
        return BinaryVector(getattr(lib(), name)(C.byref(self._v), C.byref(other._v)))

    def _ibin(self, name, other):
        getattr(lib(), name)(C.byref(self._v), C.byref(other._v))
        return self

    def __and__(self, o): return self._bin("bv_and", o)
    def __or__(self, o): return self._bin("bv_or", o)
    def __xor__(self, o): return self._bin("bv_xor", o)
    def __iand__(self, o): return self._ibin("bv_and_assign", o)
    def __ior__(self, o): return self._ibin("bv_or_assign", o)
    def __ixor__(self, o): return self._ibin("bv_xor_assign", o)
    def __lshift__(self, i): return BinaryVector(lib().bv_shl(C.byref(self._v), i))
    def __rshift__(self, i): return BinaryVector(lib().bv_shr(C.byref(self._v), i))

    def __ilshift__(self, i):
        lib().bv_shl_assign(C.byref(self._v), i)
        return self

    def __irshift__(self, i):
        lib().bv_shr_assign(C.byref(self._v), i)
        return self

    def __eq__(self, o):
        return bool(lib().bv_eq(C.byref(self._v), C.byref(o._v)))

    def __ne__(self, o):
        return not self.__eq__(o)

    __hash__ = None


class BinaryMatrix:
    def __init__(self, raw):
        self._m = raw

    def __del__(self):
        try:
            lib().bm_free(C.byref(self._m))
        except Exception:
            pass

    @classmethod
    def from_words(cls, nrows, ncols, words):
        w = np.asarray(words, dtype=np.uint64)
        assert len(w) == ((nrows + 63) // 64) * ncols
        return cls(lib().bm_from_words(nrows, ncols, _u64p(w)))

    @classmethod
    def from_packed(cls, packed_rows):
        """packed tap lists (row -> columns) -> column-major BinaryMatrix (binary_matrix.rs:15-19)."""
        n = len(packed_rows)
        wpc = (n + 63) // 64
        w = np.zeros(n * wpc, dtype=np.uint64)
        for r, row in enumerate(packed_rows):
            for c in row:
                w[c * wpc + r // 64] |= np.uint64(1) << np.uint64(63 - r % 64)
        return cls.from_words(n, n, w)

    @classmethod
    def candidate(cls, n, seed, c):
        return cls(lib().orc_rnghunt_candidate(n, seed, c))

    @property
    def nrows(self): return int(self._m.nrows)
    @property
    def ncols(self): return int(self._m.ncols)
    @property
    def words_per_col(self): return int(self._m.words_per_col)

    @property
    def data(self):
        cnt = self.words_per_col * self.ncols
        return np.ctypeslib.as_array(self._m.data, shape=(cnt,)).copy()

    def col(self, c): return BinaryVector(lib().bm_col(C.byref(self._m), c))
    def row(self, r): return BinaryVector(lib().bm_row(C.byref(self._m), r))
    def dot(self, x): return BinaryVector(lib().bm_dot(C.byref(self._m), C.byref(x._v)))
    def recur(self, x, n): return BinaryVector(lib().bm_recur(C.byref(self._m), C.byref(x._v), n))

    def to_packed(self):
        rows = []
        for r in range(self.nrows):
            bits = self.row(r).to_bits()
            rows.append([c for c, b in enumerate(bits) if b])
        return rows

    def evaluate(self, want_poly=False):
        """rnghunt.rs:27-44 on this matrix -> stage (and the BM polynomial's data words)."""
        cap = (2 * self.nrows + 63) // 64 + 1
        pw = np.zeros(cap, dtype=np.uint64)
        st = lib().orc_rnghunt_eval(C.byref(self._m), _u64p(pw), cap)
        return (int(st), pw) if want_poly else int(st)


class BinaryPolynomial:
    def __init__(self, coefficients):
        """Takes ownership of a BinaryVector (coefficient 0 = highest power)."""
        self.coefficients = coefficients

    @classmethod
    def from_coefficients(cls, c):
        return cls(BinaryVector.from_bits(c))

    def _p(self):
        # _BPoly is layout-identical to its only member
        return C.cast(C.byref(self.coefficients._v), C.POINTER(_BPoly))

    def degree(self): return int(lib().bp_degree(self._p()))
    def eval(self, x): return int(lib().bp_eval(self._p(), C.byref(x._v)))

    def modmult(self, g, p):
        return BinaryPolynomial(BinaryVector(lib().bp_modmult(self._p(), g._p(), p._p()).c))

    def modexp(self, k):
        return BinaryPolynomial(BinaryVector(lib().bp_modexp(self._p(), C.byref(k._v)).c))

    def check_integer(self, k): return bool(lib().bp_check_integer(self._p(), C.byref(k._v)))
    def is_primitive(self): return bool(lib().bp_is_primitive(self._p()))

    def __str__(self):
        buf = C.create_string_buffer(16384)
        lib().bp_display(self._p(), buf, len(buf))
        return buf.value.decode()


def berlekamp_massey(s):
    return BinaryPolynomial(BinaryVector(lib().gf2_berlekamp_massey(C.byref(s._v)).c))


def get_factors(n):
    """factors.rs:7-16 as Python ints: [2^n-1, (2^n-1)/q1, ...]."""
    out = []
    for i in range(lib().gf2_factor_count(n)):
        p = C.POINTER(C.c_uint64)()
        nw = lib().gf2_factor_words(n, i, C.byref(p))
        v = 0
        for w in range(nw):
            v = (v << 64) | int(p[w])
        out.append((v, nw))
    return out


# --------------------------------------------------------------------------------------------
# Generator models (C)
# --------------------------------------------------------------------------------------------
def _csr(packed_rows):
    row_ptr = np.zeros(len(packed_rows) + 1, dtype=np.uint32)
    taps = []
    for r, row in enumerate(packed_rows):
        taps.extend(row)
        row_ptr[r + 1] = len(taps)
    return row_ptr, np.asarray(taps, dtype=np.uint16)


def state_words(value, n):
    """Python int (bit i = weight 2^i, rng.py:135) -> little-endian u64 words."""
    nw = (n + 63) // 64
    return np.array([(value >> (64 * w)) & 0xFFFFFFFFFFFFFFFF for w in range(nw)], dtype=np.uint64)


def state_int(words):
    return sum(int(w) << (64 * i) for i, w in enumerate(words))


def lutopt_step(packed_rows, x):
    n = len(packed_rows)
    rp, tp = _csr(packed_rows)
    xin = state_words(x, n)
    out = np.zeros_like(xin)
    lib().orc_lutopt_step(n, _u32p(rp), _u16p(tp), _u64p(xin), _u64p(out))
    return state_int(out)


def lutopt_jump(packed_rows, x, e):
    n = len(packed_rows)
    rp, tp = _csr(packed_rows)
    xin = state_words(x, n)
    out = np.zeros_like(xin)
    lib().orc_lutopt_jump(n, _u32p(rp), _u16p(tp), e, _u64p(xin), _u64p(out))
    return state_int(out)


def clt(n, x, model=MODEL_GATEWARE):
    w = state_words(x, n)
    return int(lib().orc_clt(n, _u64p(w), model))


def clt_run(packed_rows, init=1, first=0, count=0, model=MODEL_GATEWARE, noise_var=1,
            want=("hist",), threads=1):
    """Samples g[first, first+count) of LUTOPT(packed, init) -> CLTGRNG.  Returns a dict with the
    requested outputs: hist, sum, sumsq (always), out8, out16, words, final_state."""
    n = len(packed_rows)
    rp, tp = _csr(packed_rows)
    bins = n + (1 if model == MODEL_NUMPY else 0)
    hist = np.zeros(bins, dtype=np.uint64)
    s, ss = C.c_int64(0), C.c_uint64(0)
    res = {}
    if threads > 1:
        x0 = state_words(init, n)
        lib().orc_clt_run_mt(n, _u32p(rp), _u16p(tp), _u64p(x0), first, count, model, _u64p(hist),
                             C.byref(s), C.byref(ss), threads)
    else:
        st = state_words(lutopt_jump(packed_rows, init, first) if first else init, n)
        out8 = np.zeros(count, dtype=np.int8) if "out8" in want else None
        out16 = np.zeros(count, dtype=np.int16) if "out16" in want else None
        words = np.zeros(count * ((n + 63) // 64), dtype=np.uint64) if "words" in want else None
        lib().orc_clt_run(
            n, _u32p(rp), _u16p(tp), _u64p(st), count, model, noise_var, _u64p(hist),
            C.byref(s), C.byref(ss),
            out8.ctypes.data_as(C.POINTER(C.c_int8)) if out8 is not None else None,
            out16.ctypes.data_as(C.POINTER(C.c_int16)) if out16 is not None else None,
            _u64p(words) if words is not None else None)
        res.update(out8=out8, out16=out16, words=words, final_state=state_int(st))
    res.update(hist=hist, sum=int(s.value), sumsq=int(ss.value))
    return res


def prbs_bits(k, nbits, state=1):
    st = C.c_uint32(state)
    out = np.zeros((nbits + 63) // 64, dtype=np.uint64)
    lib().orc_prbs_bits(k, C.byref(st), nbits, _u64p(out))
    return out, int(st.value)


def prbs_detect(k, nbits, m=0, seed=1, first_bit=0, gen_state=None, det=None):
    """Generator -> channel -> PRBSErrorDetector over bits [first_bit, first_bit+nbits)."""
    L = lib()
    gs = C.c_uint32(L.orc_prbs_advance(k, 1, first_bit) if gen_state is None else gen_state)
    if det is None:
        det = DetState()
        L.orc_det_reset(k, C.byref(det))
    ctr = DetCounters()
    L.orc_prbs_detect(k, C.byref(gs), C.byref(det), first_bit, nbits, m, seed, C.byref(ctr))
    return ctr.as_dict(), int(gs.value), det


def det_run_bits(k, rx_words, nbits, tx_words=None, det=None):
    """PRBSErrorDetector (prbs.py:43-99) clocked with caller-supplied received bits."""
    L = lib()
    if det is None:
        det = DetState()
        L.orc_det_reset(k, C.byref(det))
    ctr = DetCounters()
    rx = np.ascontiguousarray(rx_words, dtype=np.uint64)
    tx = None if tx_words is None else np.ascontiguousarray(tx_words, dtype=np.uint64)
    L.orc_det_run_bits(k, C.byref(det), _u64p(rx), _u64p(tx) if tx is not None else None, nbits, C.byref(ctr))
    return ctr.as_dict(), det


def slicer_rx_bits(tx_words, g, nbits, amplitude, noise_var, spb, delay):
    """Received bits of the board's noise loop without the pulse shaper: TX sample of bit b =
    (+-amplitude) + g*noise_var as a 12-bit signed register (tx.py:75-81), RX slices at zero
    (rx.py:28: sliced = ~sample[-1]) and keeps sample `delay` of the spb samples of each bit.
    g: int8 samples covering bit i at g[i*spb + delay]."""
    tx = np.unpackbits(np.ascontiguousarray(tx_words, dtype="<u8").view(np.uint8), bitorder="little")[:nbits]
    noise = g[delay::spb][:nbits].astype(np.int64) * int(noise_var)
    x = np.where(tx == 1, int(amplitude), -int(amplitude)) + noise
    x12 = ((x + 2048) & 4095) - 2048                    # Signal((12, True)), tx.py:80-81
    rx = (x12 >= 0).astype(np.uint8)
    pad = (-nbits) % 64
    rxw = np.packbits(np.concatenate([rx, np.zeros(pad, np.uint8)]), bitorder="little").view("<u8")
    return rxw.astype(np.uint64)


def shaper_stream(tx_bits, coeff):
    """PRBSShaper as a sequence (bitshaper.py:12-91; the relation its test pins at :150-163):
    y[8m+4] = 2 b[m] - 1 else 0, S = lfilter(coeff, 1, y), with b[m] = 0 (-1 pulses) for the 8 bit
    periods before tx_bits[0] unless the caller passes history.  tx_bits: 0/1 array whose FIRST 8
    entries are the history b[first-8 .. first-1]; returns S[8*first .. ) for the remaining bits."""
    b = np.asarray(tx_bits, dtype=np.int64)
    y = np.zeros(8 * b.size, dtype=np.int64)
    y[4::8] = 2 * b - 1
    c = np.asarray(coeff, dtype=np.int64)
    f = np.convolve(y, c)[: y.size]
    return f[64:]


def pulser_bits(first, nbits):
    """Pulser (tx.py:20-31) in the shaper's bit-clock domain (bitshaper.py:35): an 8-bit counter from 0,
    x = (counter == 0) -> source bit m is 1 when m % 256 == 0.  Returns bits [first, first+nbits) as 0/1."""
    m = np.arange(first, first + nbits, dtype=np.int64)
    return (m % 256 == 0).astype(np.uint8)


def tx_words(shaped, g, noise_var, bit_en=1):
    """tx.py:75-81: x = bitmux + noisemux registered as Signal((12, True)) -> wraps to 12 bits."""
    x = (np.asarray(shaped, dtype=np.int64) if bit_en else 0) + np.asarray(g, dtype=np.int64) * int(noise_var)
    return (((x + 2048) & 4095) - 2048).astype(np.int16)


def rnghunt_sweep(n, seed, first, count, threads=1):
    stages = np.zeros(count, dtype=np.uint8)
    counts = np.zeros(5, dtype=np.uint64)
    lib().orc_rnghunt_sweep(n, seed, first, count, stages.ctypes.data_as(C.POINTER(C.c_uint8)),
                            _u64p(counts), threads)
    return stages, counts


def rnghunt_eval_list(n, seed, cands, threads=1):
    """Stage of every candidate index in `cands` (rnghunt.rs:27-44 on candidate c of `seed`)."""
    idx = np.ascontiguousarray(cands, dtype=np.uint64)
    stages = np.zeros(idx.size, dtype=np.uint8)
    counts = np.zeros(5, dtype=np.uint64)
    if idx.size:
        lib().orc_rnghunt_eval_list(n, seed, _u64p(idx), idx.size, stages.ctypes.data_as(C.POINTER(C.c_uint8)),
                                    _u64p(counts), threads)
    return stages, counts


# --------------------------------------------------------------------------------------------
# numpy restatements of the reference's own model lines (cross-checks for the C code)
# --------------------------------------------------------------------------------------------
def np_lutopt_states(packed_rows, nsteps, init=1):
    """rng.py:122-135: x = (A.x) mod 2, integer with bit 0 = LSB; returns x_1..x_nsteps."""
    k = len(packed_rows)
    a = np.zeros((k, k), dtype=np.uint8)          # rng.py:50-54
    for row in range(k):
        for idx in packed_rows[row]:
            a[row, idx] = 1
    x = np.array([[(init >> i) & 1] for i in range(k)], dtype=np.uint8)
    out = []
    for _ in range(nsteps):
        x = np.mod(np.dot(a, x), 2)
        out.append(int("".join(str(int(xi)) for xi in x[::-1].flatten()), 2))
    return out


def np_clt(n, x_int):
    """rng.py:173-181 on one URNG value (unwrapped int16 model)."""
    logn = int(np.log2(n))
    x = np.array([int(c) for c in bin(int(x_int))[2:].rjust(n, "0")[::-1]])
    for level in range(logn):
        level_n = 2 ** (logn - level)
        y = np.zeros(level_n // 2, dtype=np.int16)
        for pair in range(0, level_n, 2):
            y[pair // 2] = x[pair] - x[pair + 1]
        x = y
    return int(x[0])


def py_clt_evaluate_loop(nsamp, n=256, seed=None):
    """clt-grng-evaluate.py:4-16 restated literally (the reference's own implementation of the CLT tree:
    255 scalar numpy subtractions per sample on random bits).  Returns the int16 samples; bench.py times it
    once as cpu_baseline.python_loop."""
    logn = int(np.log2(n))
    rng = np.random.RandomState(seed)
    samples = np.empty(nsamp, dtype=np.int16)
    for sampidx in range(nsamp):
        x = rng.randint(2, size=n).astype(np.int16)
        for level in range(logn):
            level_n = 2 ** (logn - level)
            y = np.zeros(level_n // 2, dtype=np.int16)
            for pair in range(0, level_n, 2):
                y[pair // 2] = x[pair] - x[pair + 1]
            x = y
        samples[sampidx] = x[0]
    return samples


def py_prbs_bits(k, nbits, lfsr=1):
    """prbs.py:112-113."""
    taps = {7: 6, 9: 5, 11: 9, 15: 14, 20: 3, 23: 18, 31: 28}
    out = []
    for _ in range(nbits):
        bit = ((lfsr >> (k - 1)) ^ (lfsr >> taps[k] - 1)) & 1
        lfsr = ((lfsr << 1) | bit) & ((1 << k) - 1)
        out.append(bit)
    return out
