"""Recurrence matrices for the LUTOPT uniform RNG (reference gateware/bbb/rng_recurrences.py).

`packed(n)` returns the same list-of-lists the reference's `rng_recurrences.nN` holds
(row -> sorted column indices), ready for `LUTOPT.from_packed` (rng.py:42-55).
`csr(n)` returns the (row_ptr, taps) arrays the C ABI takes (include/bbb_b200.h).
"""
import numpy as np

from ._recurrence_rows import _ROWS

SIZES = tuple(sorted(_ROWS))


def packed(n):
    if n not in _ROWS:
        raise ValueError("no shipped recurrence for n={}".format(n))
    text = "".join(_ROWS[n])
    return [[int(c) for c in row.split(",")] for row in text.split(";")]


def to_csr(packed_rows):
    """packed list-of-lists -> (row_ptr uint32[n+1], taps uint16[nnz])."""
    row_ptr = np.zeros(len(packed_rows) + 1, dtype=np.uint32)
    taps = []
    for r, row in enumerate(packed_rows):
        taps.extend(row)
        row_ptr[r + 1] = len(taps)
    return row_ptr, np.asarray(taps, dtype=np.uint16)


def csr(n):
    return to_csr(packed(n))


def from_matrix_text(text):
    """Parse an rnghunt `out` file (rnghunt.rs:46-53): n lines of n '0'/'1' chars."""
    rows = [line.strip() for line in text.splitlines() if line.strip()]
    n = len(rows)
    if any(len(r) != n or set(r) - {"0", "1"} for r in rows):
        raise ValueError("not an n x n 0/1 matrix")
    return [[c for c, ch in enumerate(r) if ch == "1"] for r in rows]


def to_matrix_text(packed_rows):
    n = len(packed_rows)
    lines = []
    for row in packed_rows:
        chars = ["0"] * n
        for c in row:
            chars[c] = "1"
        lines.append("".join(chars))
    return "\n".join(lines) + "\n"


def to_binary_matrix(packed_rows):
    """Packed tap lists -> u64 words in the crate's BinaryMatrix layout (binary_matrix.rs:10-20):
    column-major, word data[col*wpc + r//64], row r at bit 63 - r%64.  What bbb_rnghunt_eval takes."""
    import numpy as np
    n = len(packed_rows)
    wpc = (n + 63) // 64
    data = np.zeros(n * wpc, dtype=np.uint64)
    for r, row in enumerate(packed_rows):
        for c in row:
            data[c * wpc + r // 64] |= np.uint64(1 << (63 - r % 64))
    return data


# names matching the reference module (rng_recurrences.py)
n16, n32, n64, n128, n192, n256 = (packed(n) for n in (16, 32, 64, 128, 192, 256))
n512 = packed(512)
