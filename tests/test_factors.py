"""The Mersenne cofactor table (factors.rs:7-16 / factors_data.rs) verified from first principles:
entry 0 is 2^n-1, every other entry is (2^n-1)/q with q prime, and the q's are ALL the primes of
2^n-1.  When the reference checkout is present the table is also compared with it word for word."""
import os
import re

import pytest

from oracle import oracle as O


def _is_probable_prime(q):
    if q < 2:
        return False
    small = (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37)
    for p in small:
        if q % p == 0:
            return q == p
    d, s = q - 1, 0
    while d % 2 == 0:
        d //= 2
        s += 1
    for a in small:
        x = pow(a, d, q)
        if x in (1, q - 1):
            continue
        for _ in range(s - 1):
            x = x * x % q
            if x == q - 1:
                break
        else:
            return False
    return True


@pytest.mark.parametrize("lo", [1, 129, 257, 385])
def test_table_is_the_complete_factorisation(lo):
    for n in range(lo, lo + 128):
        ents = O.get_factors(n)
        r = (1 << n) - 1
        assert ents[0][0] == r
        assert ents[0][1] == max(1, (n + 63) // 64)
        rest = r
        primes = []
        for v, _ in ents[1:]:
            assert r % v == 0
            q = r // v
            if q == 1:                      # n == 1: 2^1-1 = 1 has the single listed "factor" 1
                assert n == 1
                continue
            assert _is_probable_prime(q), (n, q)
            primes.append(q)
            while rest % q == 0:
                rest //= q
        assert rest == 1, n
        assert len(set(primes)) == len(primes)


@pytest.mark.skipif(not os.path.exists("/root/reference/software/rnghunt/src/factors_data.rs"),
                    reason="reference checkout not present")
def test_table_equals_reference_words():
    text = open("/root/reference/software/rnghunt/src/factors_data.rs").read()
    body = text[text.index("= [") + 3:]
    groups = re.findall(r"&\[\s*((?:&\[[^\]]*\],\s*)+)\]", body)
    assert len(groups) == 512
    for n, grp in enumerate(groups, start=1):
        ref = []
        for ent in re.findall(r"&\[([^\]]*)\]", grp):
            ws = [int(w, 16) for w in ent.replace(" ", "").split(",") if w]
            v = 0
            for w in ws:
                v = (v << 64) | w
            ref.append((v, len(ws)))
        assert ref == O.get_factors(n), n
