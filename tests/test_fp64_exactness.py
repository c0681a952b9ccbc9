"""The Poseidon MDS runs on the FP64 pipe (csrc/poseidon.cu: mds_dft / mds_blocks / mds_idft and the paired partial
rounds).  That is bit-exact only if every intermediate is an integer multiple of a fixed power of two whose magnitude
stays below 2^53.  This test re-states the data flow of those functions on (a) exact rationals — the result must equal
the plain matrix product — and (b) worst-case magnitude bounds (every input half at 2^32 - 1, signs chosen
adversarially at every addition), and checks the 2^53 margin for a single product and for a pair."""
from fractions import Fraction as Fr
import math
import random

ROW0 = [17, 20, 34, 18, 39, 13, 13, 28, 2, 16, 41, 15]


def mds_entry(j, i):
    return ROW0[(i - j) % 12] + (8 if i == 0 and j == 0 else 0)


def direct(x):
    return [sum(mds_entry(j, i) * x[j] for j in range(12)) for i in range(12)]


class Track:
    """Arithmetic on values (exact) or on magnitude bounds (|a op b| <= |a| + |b|).  `peak` is the largest mantissa demand
    seen: |value| * 2^frac, where frac is the number of fractional bits the quantities of the current section may carry
    (0 except for the rescaled spectrum between the two products of a pair)."""

    def __init__(self, bound_mode):
        self.b, self.peak, self.frac = bound_mode, 0, 0

    def see(self, v):
        self.peak = max(self.peak, abs(v) * 2**self.frac)
        return v

    def add(self, a, b):
        return self.see(abs(a) + abs(b) if self.b else a + b)

    def sub(self, a, b):
        return self.see(abs(a) + abs(b) if self.b else a - b)

    def fma(self, c, a, b):  # c*a + b, c a constant
        return self.see(abs(c) * abs(a) + abs(b) if self.b else c * a + b)


def dft(t, x):
    s = {"f0": [0] * 3, "f2": [0] * 3, "u": [0] * 3, "v": [0] * 3}
    for r in range(3):
        s02, s13 = t.add(x[r], x[6 + r]), t.add(x[3 + r], x[9 + r])
        s["u"][r], s["v"][r] = t.sub(x[r], x[6 + r]), t.sub(x[3 + r], x[9 + r])
        s["f0"][r], s["f2"][r] = t.add(s02, s13), t.sub(s02, s13)
    return s


def blocks(t, s, frac=(0, 0, 0)):
    """frac: fractional bits carried by the f0, f2 and u/v blocks (non-zero only between the two products of a pair)."""
    t.frac = frac[0]
    a0, a1, a2 = s["f0"]
    T = t.add(t.add(a0, a1), a2)
    s["f0"] = [t.add(T, a2), t.add(T, a0), t.add(T, a1)]
    t.frac = frac[1]
    b0, b1, b2 = s["f2"]
    s["f2"] = [t.fma(8, b2, t.fma(-2, b1, -b0)), t.fma(-8, b0, t.fma(-2, b2, -b1)), t.fma(2, b0, t.fma(-8, b1, -b2))]
    t.frac = frac[2]
    U0, U1, U2 = s["u"]
    V0, V1, V2 = s["v"]
    f, ad, sb = t.fma, t.add, t.sub
    s["u"] = [f(4, V2, ad(f(-16, V1, ad(f(2, U0, -V0), U1)), U2)),
              f(-16, V2, ad(f(2, U1, sb(f(-4, U0, V0), V1)), U2)),
              f(2, U2, sb(f(-4, U1, ad(f(16, U0, V0), V1)), V2))]
    s["v"] = [f(-4, U2, ad(f(16, U1, ad(f(2, V0, U0), V1)), V2)),
              f(16, U2, ad(f(2, V1, ad(f(-4, V0, -U0), U1)), V2)),
              f(2, V2, ad(f(-4, V1, sb(f(16, V0, -U0), U1)), U2))]
    t.frac = 0


def idft(t, s, x0, c0, c1):
    x = [0] * 12
    for r in range(3):
        A, B = t.fma(c0, s["f0"][r], s["f2"][r]), t.fma(c0, s["f0"][r], -s["f2"][r])
        x[r], x[3 + r] = t.fma(c1, s["u"][r], A), t.fma(c1, s["v"][r], B)
        x[6 + r], x[9 + r] = t.fma(-c1, s["u"][r], A), t.fma(-c1, s["v"][r], B)
    x[0] = t.fma(8, x0, x[0])
    return x


def single(t, x):
    s = dft(t, x)
    blocks(t, s)
    return idft(t, s, x[0], 16, 1)


def pair(t, x, new0):
    """Two products with word 0 replaced by new0 in between (poseidon_permute_state, ZK_LAZY2 == 2)."""
    s = dft(t, x)
    blocks(t, s)
    y = t.add(t.fma(16, s["f0"][0], s["f2"][0]), s["u"][0])
    w = t.fma(8, x[0], y)  # word 0 of the first product: goes to the integer side
    d = t.sub(new0, y)
    # between the two products the spectrum is kept in the scale mds_blocks left it in, so the f0 block carries d/64
    # (multiples of 2^-6) and the u block d/2; the f2 block and everything after the inverse DFT are integers again
    t.frac = 6
    s["f0"][0] = t.fma(Fr(1, 64), d, s["f0"][0])
    t.frac = 0
    s["f2"] = [t.fma(4, s["f2"][0], d), t.see(4 * s["f2"][1]), t.see(4 * s["f2"][2])]
    t.frac = 1
    s["u"][0] = t.fma(Fr(1, 2), d, s["u"][0])
    blocks(t, s, (6, 0, 1))
    return w, idft(t, s, new0, 1024, 2)


def test_dft_form_equals_matrix_product():
    rng = random.Random(3)
    t = Track(False)
    for _ in range(50):
        x = [rng.getrandbits(32) for _ in range(12)]
        assert single(t, x) == direct(x)
        new0 = rng.getrandbits(32)
        w, z = pair(t, x, new0)
        y = direct(x)
        assert w == y[0]
        y[0] = new0
        assert z == direct(y)


def test_every_intermediate_stays_exact_in_a_double():
    half = 2**32 - 1
    t = Track(True)
    out = single(t, [half] * 12)
    assert t.peak < 2**53 and max(out) < 2**53
    t = Track(True)
    w, z = pair(t, [half] * 12, half)
    # bounds with adversarial signs at every addition are loose (the true outputs of a pair are below 2^49) and still
    # inside the 53-bit mantissa
    assert t.peak < 2**53, "mantissa demand 2^%.2f" % math.log2(t.peak)


def test_exact_rationals_fit_the_mantissa_on_extreme_inputs():
    """Exact run on inputs at the edges (all halves 0 / 2^32-1 patterns): every intermediate, written as an odd integer
    times a power of two, has an integer part of at most 53 bits."""

    class Exact(Track):
        def see(self, v):
            v = Fr(v)
            n = abs(v.numerator)
            while n and n % 2 == 0:
                n //= 2
            assert n.bit_length() <= 53 and (v.denominator & (v.denominator - 1)) == 0, v
            return v

    rng = random.Random(9)
    half = 2**32 - 1
    for _ in range(200):
        x = [rng.choice((0, 1, half, half - 1, rng.getrandbits(32))) for _ in range(12)]
        t = Exact(False)
        single(t, x)
        pair(t, x, rng.choice((0, half, rng.getrandbits(32))))
