"""Pins the GF(2) oracle (oracle/gf2.c) against the reference crate's own unit tests.

Each test is the reference's inline `#[test]` of the same name, value for value:
  binary_vector.rs:326-624 (20 tests), binary_matrix.rs:104-210 (4),
  binary_polynomial.rs:215-372 (7), berlekamp_massey.rs:33-67 (2)
plus the shipped matrices (software/rnghunt/matrices/N) as positive known answers.
"""
import numpy as np
import pytest

from oracle.oracle import BinaryMatrix as BM, BinaryPolynomial as BP, BinaryVector as BV
from oracle.oracle import berlekamp_massey

X96 = [0x883C385E66BD8704, 0xE43A5DF300000000]
Y96 = [0xA8F3B1900CC10FFF, 0x38A599C200000000]
X256 = [0xA8F3B1900CC10FFF, 0x38A599C25A5F60A2, 0x5BC898AB15066BA4, 0x44F3A512D7C744EC]
BITS96 = [
    1, 0, 0, 0, 1, 0, 0, 0, 0, 0, 1, 1, 1, 1, 0, 0, 0, 0, 1, 1, 1, 0, 0, 0, 0, 1, 0, 1, 1,
    1, 1, 0, 0, 1, 1, 0, 0, 1, 1, 0, 1, 0, 1, 1, 1, 1, 0, 1, 1, 0, 0, 0, 0, 1, 1, 1, 0, 0,
    0, 0, 0, 1, 0, 0, 1, 1, 1, 0, 0, 1, 0, 0, 0, 0, 1, 1, 1, 0, 1, 0, 0, 1, 0, 1, 1, 1, 0,
    1, 1, 1, 1, 1, 0, 0, 1, 1]
BITS64 = [
    1, 0, 0, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 1, 1, 0, 1, 0, 0, 0, 1, 0, 0, 1, 0, 1, 0, 0,
    1, 0, 0, 1, 1, 1, 1, 0, 1, 0, 0, 0, 1, 0, 0, 1, 1, 0, 1, 0, 1, 1, 0, 1, 1, 0, 1, 0, 0,
    0, 1, 0, 0, 0, 0]


# ---- binary_vector.rs -------------------------------------------------------------------
def test_index():
    x = BV.raw(8, [0x8300_0000_0000_0000])
    assert [x[i] for i in range(8)] == [True, False, False, False, False, False, True, True]


def test_to_bits():
    assert BV.raw(8, [0x8300_0000_0000_0000]).to_bits() == [1, 0, 0, 0, 0, 0, 1, 1]
    assert BV.raw(64, [0x9F8344A4F44D6D10]).to_bits() == BITS64
    assert BV.raw(96, X96).to_bits() == BITS96


def test_from_words():
    assert BV.from_words(96, X96).to_bits() == BITS96


def test_from_bits():
    assert BV.from_bits(BITS64).data == [0x9F8344A4F44D6D10]


def test_from_bitstring():
    assert BV.from_bitstring("0000000100100011010001010110011110001001").data == [0x0123456789000000]


def test_to_bitstring():
    s = "0000000100100011010001010110011110001001"
    assert BV.from_bitstring(s).to_bitstring() == s


def test_slice():
    x = BV.from_words(256, X256)
    bits = x.to_bits()
    for a, b in [(0, 1), (1, 2), (0, 2), (4, 8), (60, 70), (60, 130), (1, 255), (64, 256),
                 (64, 128), (64, 129), (255, 256), (0, 256)]:
        assert x.slice(a, b).to_bits() == bits[a:b]


def _z(op):
    return BV.from_words(96, [op(X96[0], Y96[0]), op(X96[1], Y96[1])])


def test_bitand():
    assert (BV.from_words(96, X96) & BV.from_words(96, Y96)) == _z(lambda a, b: a & b)


def test_bitandassign():
    x = BV.from_words(96, X96)
    x &= BV.from_words(96, Y96)
    assert x == _z(lambda a, b: a & b)


def test_bitor():
    assert (BV.from_words(96, X96) | BV.from_words(96, Y96)) == _z(lambda a, b: a | b)


def test_bitorassign():
    x = BV.from_words(96, X96)
    x |= BV.from_words(96, Y96)
    assert x == _z(lambda a, b: a | b)


def test_bitxor():
    assert (BV.from_words(96, X96) ^ BV.from_words(96, Y96)) == _z(lambda a, b: a ^ b)


def test_bitxorassign():
    x = BV.from_words(96, X96)
    x ^= BV.from_words(96, Y96)
    assert x == _z(lambda a, b: a ^ b)


def test_shl():
    x = BV.from_bits([0, 1, 0, 1, 1, 0, 1, 1])
    want = [[1, 0, 1, 1, 0, 1, 1], [0, 1, 1, 0, 1, 1], [1, 1, 0, 1, 1], [1, 0, 1, 1], [0, 1, 1],
            [1, 1], [1], []]
    for i, w in enumerate(want, start=1):
        assert (x << i).to_bits() == w
    x = BV.from_words(96, X96)
    assert (x << 4) == BV.from_words(92, [0x83C385E66BD8704E, 0x43A5DF3000000000])
    assert (x << 64) == BV.from_words(32, [0xE43A5DF300000000])
    x = BV.from_words(256, X256)
    assert (x << 8) == BV.from_words(248, [0xF3B1900CC10FFF38, 0xA599C25A5F60A25B,
                                           0xC898AB15066BA444, 0xF3A512D7C744EC00])
    assert (x << 128) == BV.from_words(128, [0x5BC898AB15066BA4, 0x44F3A512D7C744EC])


def test_shl_assign():
    x = BV.from_bits([0, 1, 0, 1, 1, 0, 1, 1])
    x <<= 4
    assert x.to_bits() == [1, 0, 1, 1]
    x = BV.from_words(96, X96)
    x <<= 4
    assert x == BV.from_words(92, [0x83C385E66BD8704E, 0x43A5DF3000000000])
    x = BV.from_words(96, X96)
    x <<= 64
    assert x == BV.from_words(32, [0xE43A5DF300000000])
    x = BV.from_words(256, X256)
    x <<= 8
    assert x == BV.from_words(248, [0xF3B1900CC10FFF38, 0xA599C25A5F60A25B,
                                    0xC898AB15066BA444, 0xF3A512D7C744EC00])
    x = BV.from_words(256, X256)
    x <<= 128
    assert x == BV.from_words(128, [0x5BC898AB15066BA4, 0x44F3A512D7C744EC])


Y100 = [0x0883C385E66BD870, 0x4E43A5DF30000000]
Z160 = [0x0000000000000000, 0x883C385E66BD8704, 0xE43A5DF300000000]
Y264 = [0x00A8F3B1900CC10F, 0xFF38A599C25A5F60, 0xA25BC898AB15066B, 0xA444F3A512D7C744,
        0xEC00000000000000]
Z384 = [0, 0] + X256


def test_shr():
    x = BV.from_bits([0, 1, 0, 1, 1, 0, 1, 1])
    assert (x >> 1).to_bits() == [0, 0, 1, 0, 1, 1, 0, 1, 1]
    assert (x >> 2).to_bits() == [0, 0, 0, 1, 0, 1, 1, 0, 1, 1]
    assert (x >> 3).to_bits() == [0, 0, 0, 0, 1, 0, 1, 1, 0, 1, 1]
    x = BV.from_words(96, X96)
    assert (x >> 4) == BV.from_words(100, Y100)
    assert (x >> 64) == BV.from_words(160, Z160)
    x = BV.from_words(256, X256)
    assert (x >> 8) == BV.from_words(264, Y264)
    assert (x >> 128) == BV.from_words(384, Z384)


def test_shr_assign():
    x = BV.from_bits([0, 1, 0, 1, 1, 0, 1, 1])
    x >>= 3
    assert x.to_bits() == [0, 0, 0, 0, 1, 0, 1, 1, 0, 1, 1]
    x = BV.from_words(96, X96)
    x >>= 4
    assert x == BV.from_words(100, Y100)
    x = BV.from_words(96, X96)
    x >>= 64
    assert x == BV.from_words(160, Z160)
    x = BV.from_words(256, X256)
    x >>= 8
    assert x == BV.from_words(264, Y264)
    x = BV.from_words(256, X256)
    x >>= 128
    assert x == BV.from_words(384, Z384)


def test_eq():
    x, y = BV.from_words(96, X96), BV.from_words(96, X96)
    z, w = BV.from_words(96, Y96), BV.from_words(97, Y96)
    assert x == y
    assert y != z
    assert z != w


def test_firstbit():
    assert BV.from_bits([0, 0, 0, 0]).firstbit() == 4
    assert BV.from_bits([0, 1, 0, 0]).firstbit() == 1
    assert BV.from_words(96, X96).firstbit() == 0
    assert BV.from_words(96, [0x003C385E66BD8704, 0xE43A5DF300000000]).firstbit() == 10
    assert BV.from_words(96, [0x0, 0x743A5DF300000000]).firstbit() == 65


def test_count_ones():
    assert BV.from_bits([0, 1, 0, 1, 0]).count_ones() == 2
    assert BV.from_words(128, [0x1111, 0x2222]).count_ones() == 8


# ---- binary_matrix.rs -------------------------------------------------------------------
def test_matrix_from_words():
    words = [0x4D00000000000000, 0x6600000000000000, 0x2800000000000000, 0xC800000000000000,
             0x3E00000000000000, 0x9600000000000000, 0x2600000000000000, 0xB400000000000000,
             0xD700000000000000, 0x0A00000000000000, 0xD600000000000000, 0xA300000000000000]
    x = BM.from_words(8, 12, words)
    assert x.col(0).to_bits() == [0, 1, 0, 0, 1, 1, 0, 1]
    assert x.col(11).to_bits() == [1, 0, 1, 0, 0, 0, 1, 1]
    assert x.row(0).to_bits() == [0, 0, 0, 1, 0, 1, 0, 1, 1, 0, 1, 1]
    assert x.row(7).to_bits() == [1, 0, 0, 0, 0, 0, 0, 0, 1, 0, 0, 1]
    words = [0xD1571BCA685A723F, 0x8FACE88700000000, 0x03CA88D60FA542A5, 0x982C521100000000]
    x = BM.from_words(96, 2, words)
    assert x.row(0).to_bits() == [1, 0]
    assert x.col(0).to_bits() == [
        1, 1, 0, 1, 0, 0, 0, 1, 0, 1, 0, 1, 0, 1, 1, 1, 0, 0, 0, 1, 1, 0, 1, 1, 1, 1, 0, 0, 1,
        0, 1, 0, 0, 1, 1, 0, 1, 0, 0, 0, 0, 1, 0, 1, 1, 0, 1, 0, 0, 1, 1, 1, 0, 0, 1, 0, 0, 0,
        1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 1, 1, 1, 1, 1, 0, 1, 0, 1, 1, 0, 0, 1, 1, 1, 0, 1, 0, 0,
        0, 1, 0, 0, 0, 0, 1, 1, 1]


DOT_WORDS = [
    0x77C560511586E4DA, 0xB11A0E7E1E9CF914, 0x1A157938C089896A, 0x17D052B7EEF8E6D0,
    0x5DFD087B667BA4DF, 0x95328B51CD0E784A, 0x59911C3B44ECA753, 0x1730A89BCE0AB547,
    0xE29DDCB392C0E5B8, 0x80370BE533EEDCD8, 0x95F689E62570FE53, 0x4A1CFF8243D047CB,
    0x1E861C5EBF131D83, 0x813EC3A71242DA8F, 0xD7D26080432B589C, 0x7AFA0A8C881BB4F8,
    0x71C554B7F6E533E5, 0x3BFBE4E054B8E22F, 0x22C0B1C0E94BE294, 0x28C363CBD00494A3,
    0x573843D2DF76065E, 0x89910B59119E0632, 0x1F73D8036DE4EE5E, 0x22F710A3C3DB93E9,
    0xDF1653515BC13BF4, 0x63F7DD10128C77D3, 0xD41CCE1357E42511, 0x9E4EB4BC58DC4FEE,
    0x54A95557FFDE73B2, 0xBA2681F23B0CEDBD, 0xD02049299E685EB2, 0x9B2C183751C2B869,
    0x73CC438449CC6271, 0xFA43C8326C5C43E7, 0xB6AE643BE51D627A, 0x6759DDEC081A11E4,
    0x69FFA1DA9AFB9A16, 0x5C4AA766A50FF05E, 0x51669CC64A5E8803, 0xB2D66A2CFECA1000,
    0x26AD814E2AFA53B8, 0x0CDDA1532B0A52DA, 0xD8197312D918C4BF, 0x24B62C18F1C19346,
    0xE44320950703CFAE, 0xB034B92097062BAB, 0x196C0A40C6FC7CEE, 0x629BB5997E84BD48,
    0x75137E75630C2801, 0xBCCFC86141718941, 0xF6744675B715DE94, 0x0C519E42973468CA,
    0x5D892F760BE9501A, 0x5DB84E56D9DB94CB, 0x4D6F4B49588618C8, 0x14745934B5F54369,
    0x00BFF2C0CCA7DA5E, 0x698EC12ECC5F55A3, 0xE7607AFD0411A1EF, 0xC14B279EB4581717,
    0x2BEB989C25DE87F9, 0x65789946989E55D2, 0xA78CCAAD75F2AC67, 0x8F0FF89D1069A684,
    0x767EC8FC4A0DD689, 0x8011EFE294143EDA, 0x5579A9382B1D2F76, 0x171A317F6B45520B,
    0xD173916223164143, 0xBC3696451031D3A7, 0xA729E856902C22B7, 0xC85D913AAD78816D,
    0x75DE3C3712A5FC1E, 0x2C7334B3A58835D6, 0x5BBEB8A62019E907, 0xD93376BA76A65A07,
    0xD3ABFE3035C92107, 0x606E4AA6D487EC9A, 0xD8D3C0FDA50A2846, 0xDD12369F645ED991,
    0xF941D4447A2790F6, 0x07E6D238258D54B1, 0x60D649AE2F11152A, 0xF69989EF3C27C602,
    0xC7E838E302AD16C3, 0xF674DF92A8B704BB, 0x251BA8369ED7177E, 0xB8D3796A15844B22,
    0x75834CAF996681D9, 0x994211A62E69066D, 0xDE22B0148825CEEF, 0x5E18659537EC67EC,
    0x62EF086B0EBF786C, 0xDE5BF1513E53B856, 0x5556C74AC8BFD027, 0x1DF49C94CA385BA7,
    0xF9C1DFC77923294E, 0x8891C3FB31495C6F, 0x9784EAF4B7C670B3, 0x95DBACEF3679D924,
    0x9CF92F4F3A98606A, 0x52BC806C1806AC3C, 0xAE287240FB221ECC, 0xCCB5B93ED7BA3275,
    0xE5373152391CE2CF, 0x1974FB540F1AA6A4, 0x060F78470211F3B3, 0xD150DDD61DFF23A6,
    0x3CBCA58CB9685E3F, 0x6E98872422D2C506, 0x0B14E36DAF9CD11D, 0xC62E02E3482C1EBA,
    0xB3855792B2922279, 0x96155FD5DCB4F8D5, 0xE22C4340366D1826, 0x3A715D1DD0FAB8B8,
    0x6026404DD0852CD4, 0xCD5651D4A18F956C, 0xFCC6537EC81DDEE4, 0xC8C43B4084E7A22C,
    0x43F75AD11CEFA37B, 0x84FCDB8697552878, 0xD7ED4A855DAC4016, 0x81AB173B3634A266,
    0x3353019B01009DD4, 0xD3DEF8BC93D6DEBD, 0x43E1C4E8BCA565CB, 0x3B978613171F4F20]


def test_dot():
    a = BM.from_words(64, 128, DOT_WORDS)
    b = BV.from_bits([
        1, 0, 1, 0, 1, 1, 0, 1, 0, 1, 1, 0, 1, 1, 0, 1, 1, 0, 0, 0, 1, 1, 0, 1, 0, 0, 1, 1, 0,
        0, 0, 0, 0, 1, 0, 1, 0, 1, 1, 1, 0, 0, 0, 0, 0, 0, 1, 1, 1, 0, 1, 1, 1, 1, 0, 0, 1, 0,
        0, 1, 0, 0, 1, 0, 1, 0, 0, 1, 0, 0, 1, 1, 0, 1, 0, 1, 1, 1, 0, 1, 0, 1, 0, 1, 1, 0, 0,
        0, 0, 1, 1, 1, 1, 0, 0, 0, 0, 1, 0, 1, 1, 0, 1, 0, 1, 0, 1, 0, 0, 0, 1, 1, 0, 0, 1, 1,
        0, 0, 0, 0, 0, 1, 1, 1, 1, 0, 1, 0])
    assert a.dot(b).to_bits() == [
        1, 0, 0, 1, 0, 1, 1, 0, 0, 1, 0, 0, 0, 1, 1, 0, 0, 0, 0, 1, 1, 0, 0, 0, 1, 0, 0, 0, 0,
        1, 0, 1, 0, 1, 0, 1, 0, 1, 0, 1, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 0, 0, 1, 1, 1, 1, 0,
        0, 1, 0, 0, 1, 0]


def test_recur():
    words = [0x7400000000000000, 0x5800000000000000, 0xC500000000000000, 0xD000000000000000,
             0xD500000000000000, 0xE600000000000000, 0xF100000000000000, 0x4700000000000000]
    a = BM.from_words(8, 8, words)
    b = BV.from_bits([1, 0, 1, 0, 1, 0, 1, 0])
    assert a.recur(b, 24).to_bits() == [
        1, 0, 1, 1, 0, 0, 0, 1, 1, 1, 0, 0, 0, 1, 1, 1, 0, 0, 0, 1, 1, 1, 0, 0]


def test_random():
    """binary_matrix.rs:195-209 checks row weights only; the candidate stream always draws from
    [3,4,4,4,4,4,4,4] (rnghunt.rs:25), so check that and the greedy column balance."""
    for n in (8, 16, 64, 192):
        for c in range(4):
            a = BM.candidate(n, 1, c)
            rows = a.to_packed()
            assert all(len(r) in (3, 4) for r in rows)
            colw = np.zeros(n, dtype=int)
            for r in rows:
                assert len(set(r)) == len(r)
                for col in r:
                    colw[col] += 1
            assert colw.max() - colw.min() <= 1


# ---- binary_polynomial.rs ---------------------------------------------------------------
def test_degree():
    assert BP.from_coefficients([1, 1, 1, 1, 1, 1, 1]).degree() == 6
    assert BP.from_coefficients([0, 1, 0, 0, 0]).degree() == 3
    assert BP.from_coefficients([0, 0, 0, 0, 0, 0, 1]).degree() == 0
    assert BP.from_coefficients([0]).degree() == -1


def test_display():
    assert str(BP.from_coefficients([1, 0, 0, 1, 1, 1, 1])) == "x^6 + x^3 + x^2 + x + 1"
    assert str(BP.from_coefficients([0])) == "0"


def test_eval():
    p = BP.from_coefficients([1, 0, 1, 0, 1, 1])
    assert p.eval(BV.from_bits([1, 1, 1, 1, 1, 0])) == 1
    assert p.eval(BV.from_bits([1, 1, 1, 1, 1, 1])) == 0
    p = BP(BV.from_words(256, X256))
    x = BV.from_words(256, [0x72A750F0B61C8083, 0xD6B446B28D7709F5, 0x4A2A35BB6B4A9D57,
                            0x36B60EF547DD7363])
    assert p.eval(x) == 1
    y = BV.from_words(256, [0x6375252384DE2649, 0x15C4FC9E04004D91, 0x2FE1EF103A5A9D6D,
                            0xBA9BA69D6FE360E0])
    assert p.eval(y) == 0


def test_modmult():
    p = BP.from_coefficients([1, 0, 0, 0, 0, 0, 0])
    f = BP.from_coefficients([0, 0, 0, 0, 1, 0, 0])
    g = BP.from_coefficients([0, 0, 0, 0, 0, 1, 1])
    assert f.modmult(g, p).coefficients.to_bits() == [0, 0, 0, 1, 1, 0, 0]
    p = BP.from_coefficients([0, 1, 1, 0, 0, 1])
    f = BP.from_coefficients([0, 0, 0, 0, 1, 0])
    x = BP.from_coefficients([0, 0, 0, 0, 1, 0])
    f = f.modmult(f, p)
    assert f.coefficients.to_bits() == [0, 0, 0, 1, 0, 0]
    f = f.modmult(f, p)
    assert f.coefficients.to_bits() == [0, 0, 1, 0, 0, 1]
    f = f.modmult(x, p)
    assert f.coefficients.to_bits() == [0, 0, 1, 0, 1, 1]
    p = BP(BV.from_words(256, [0x0100000000000000, 0, 0, 0]))
    f = BP(BV.from_words(256, [0, 0x1000000010000000, 0x1000000010000000, 0]))
    g = BP(BV.from_words(256, [0, 0, 0, 0x0000000000003010]))
    fg = [0x0000000000000301, 0x0000030100000301, 0x0000030100000000, 0]
    assert f.modmult(g, p).coefficients.data == fg


def test_modexp():
    p = BP(BV.from_words(256, [0x0000000000000001, 0, 0, 0]))
    for k, s in [(0, "1"), (1, "x"), (2, "x^2"), (20, "x^20"), (100, "x^100")]:
        assert str(p.modexp(BV.from_words(64, [k]))) == s
    p = BP.from_coefficients([1, 1, 0, 0, 1])
    assert str(p.modexp(BV.from_words(64, [15]))) == "1"


P200 = [0x0000000000000100, 0, 0, 0x000000000000002d]


def test_check_integer():
    p = BP.from_coefficients([1, 1, 0, 0, 1])
    for k in range(1, 15):
        assert not p.check_integer(BV.from_words(64, [k]))
    assert p.check_integer(BV.from_words(64, [15]))
    p = BP(BV.from_words(256, P200))
    assert str(p) == "x^200 + x^5 + x^3 + x^2 + 1"
    for k in range(1, 100):
        assert not p.check_integer(BV.from_words(64, [k]))
    r = BV.from_words(256, [0xff, 0xffffffffffffffff, 0xffffffffffffffff, 0xffffffffffffffff])
    assert p.check_integer(r)


def test_is_primitive():
    assert BP.from_coefficients([1, 1, 0, 0, 1]).is_primitive()
    assert not BP.from_coefficients([1, 0, 1, 1, 1]).is_primitive()
    assert BP.from_coefficients([
        1, 1, 0, 0, 1, 0, 0, 0, 1, 1, 0, 0, 0, 0, 1, 1, 0,
        1, 0, 0, 0, 1, 0, 0, 1, 1, 0, 1, 1, 1, 1, 0, 1, 1]).is_primitive()
    assert BP(BV.from_words(256, P200)).is_primitive()


# ---- berlekamp_massey.rs ----------------------------------------------------------------
def test_berlekamp_massey():
    assert str(berlekamp_massey(BV.from_bitstring("0000100011000010011"))) == "x^9 + x^5 + 1"
    assert str(berlekamp_massey(BV.from_bitstring("00000000101000000100010"))) == "x^11 + x^9 + 1"
    s = "01000100111000101110110000100011"
    assert str(berlekamp_massey(BV.from_bitstring(s))) == "x^16 + x^14 + x^13 + x^11 + 1"


def test_berlekamp_massey_long(golden):
    s = "101010100110010000111101100101010011111000110110100010111010101011"
    assert str(berlekamp_massey(BV.from_bitstring(s))) == \
        "x^33 + x^31 + x^29 + x^26 + x^24 + x^22 + x^19 + x^14 + x^8 + x^7 + x^2 + 1"
    s = ("1011010010110100101101001011010010110100101101001011010010110101"
         "0111110101111101011111010111110101111101011111010111110101110010")
    assert str(berlekamp_massey(BV.from_bitstring(s))) == "x^64 + x^62 + x^61 + x + 1"
    # the strings above are what util/lfsr.py prints (regenerated by tests/golden/make_golden.py)
    assert golden["bm_strings_from_lfsr_py"][1] == s
    assert golden["bm_strings_from_lfsr_py"][0] == "01000100111000101110110000100011"


# ---- shipped matrices are positive known answers (rnghunt.rs:27-44) ----------------------
@pytest.mark.parametrize("n,weight", [(16, 9), (32, 23), (64, 31), (128, 65), (192, 103), (256, 125)])
def test_shipped_matrices_accepted(n, weight):
    from basebandboard_b200 import recurrences
    a = BM.from_packed(recurrences.packed(n))
    stage, poly = a.evaluate(want_poly=True)
    assert stage == 4
    # characteristic-polynomial weights recorded in SURVEY.md section 8(c)
    assert sum(bin(int(w)).count("1") for w in poly) == weight


def test_shipped_matrix_512_accepted():
    from basebandboard_b200 import recurrences
    assert BM.from_packed(recurrences.packed(512)).evaluate() == 4


@pytest.mark.parametrize("k,tap", [(7, 6), (9, 5), (11, 9), (15, 14), (20, 3), (23, 18), (31, 28)])
def test_prbs_trinomials_primitive(k, tap):
    """prbs.py:14: every x^k + x^tap + 1 is primitive (SURVEY 8a KAT)."""
    c = [0] * (k + 1)
    c[0] = c[k - tap] = c[k] = 1
    assert BP.from_coefficients(c).is_primitive()
