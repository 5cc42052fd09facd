"""The multicorrelator's short chip-phase form (csrc/multicorr.cu: prologue tables s_k / s_short, sample_short), restated in
numpy binary32 and checked on the CPU. The exact form evaluates, per tap, fl(fl(a + shift) - rem) — the resampler's index
expression of volk_gnsssdr_32f_xn_resampler_32f_xn.h:63-80 — the short form ONE addition fl(a + k), with

    u = 2^(e-23) the spacing of the binade [2^e, 2^(e+1)) that holds a,   RN_u(x) = fl(fl(x + 1.5 2^e) - 1.5 2^e),
    k = RN_u(shift) + RN_u(-rem),

claimed equal bit for bit for every binary32 a in [2^e + m, 2^(e+1) - m], m = max|shift| + |rem| + 4u, unless one of the two
roundings is a tie (then the kernel does not use the short form in that binade). This test draws shifts / remainders (random,
dyadic = tie candidates, tiny) and checks the claim on dense samples of a, including the ends of the range."""
import numpy as np
import pytest

f32 = np.float32


def rn_u(x, e):
    big = f32(1.5) * f32(2.0 ** e)
    return f32(f32(x + big) - big)


def short_form_tables(shifts, rem, e):
    """(k per tap, lo, hi) as the kernel's prologue computes them; lo > hi when the binade is ruled out"""
    lo2 = f32(2.0 ** e)
    u = f32(lo2 * f32(1.1920928955078125e-07))
    nrem = f32(-rem)
    r = rn_u(nrem, e)
    ok = not (f32(abs(f32(nrem - r))) * f32(2.0) == u)
    ks = []
    for sh in shifts:
        q = rn_u(sh, e)
        ok = ok and not (f32(abs(f32(sh - q))) * f32(2.0) == u)
        ks.append(f32(q + r))
    margin = f32(f32(max(abs(min(shifts)), abs(max(shifts)))) + f32(abs(nrem)) + f32(4.0) * u)
    ok = ok and bool(margin * f32(4.0) < lo2)
    return ks, (f32(lo2 + margin), f32(f32(2.0) * lo2 - margin)) if ok else (f32(3.4e38), f32(-3.4e38))


def exact_form(a, sh, rem):
    return f32(f32(a + sh) + f32(-rem))


CASES = {
    "random": lambda rng: (rng.uniform(-1.0, 1.0, 11).astype(f32), f32(rng.uniform(0, 1))),
    "veml_taps": lambda rng: (np.array([-1, -0.5, -0.25, -0.1, -0.05, 0, 0.05, 0.1, 0.25, 0.5, 1], f32), f32(rng.uniform(0, 1))),
    "dyadic": lambda rng: ((rng.integers(-64, 65, 11) / 64.0).astype(f32), f32(rng.integers(0, 1024) / 1024.0)),   # tie candidates
    "tiny": lambda rng: (rng.uniform(-1e-3, 1e-3, 5).astype(f32), f32(rng.uniform(0, 1e-4))),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_short_form_equals_exact_form_bit_for_bit(case):
    rng = np.random.default_rng(sorted(CASES).index(case))
    checked = excluded = 0
    for trial in range(40):
        shifts, rem = CASES[case](rng)
        shifts = [f32(s) for s in shifts]
        for e in range(2, 22):                       # chip phases of 4 ... 4 M chips (the kernel keeps 22 binades)
            ks, (lo, hi) = short_form_tables(shifts, rem, e)
            if lo > hi:
                excluded += 1
                continue
            u = 2.0 ** (e - 23)
            # binary32 values of the binade are the multiples of u: the ends of the proven range, and random points inside
            i_lo, i_hi = int(np.ceil(float(lo) / u)), int(np.floor(float(hi) / u))
            idx = np.unique(np.concatenate([np.arange(i_lo, i_lo + 8), np.arange(i_hi - 7, i_hi + 1), rng.integers(i_lo, i_hi + 1, 48)]))
            a = (idx.astype(np.float64) * u).astype(f32)
            assert np.all(a.astype(np.float64) == idx * u)
            for sh, k in zip(shifts, ks):
                want = (a + sh).astype(f32) + f32(-rem)            # two roundings
                got = a + k                                         # one
                assert want.dtype == f32 and got.dtype == f32
                assert np.array_equal(want.view(np.uint32), got.view(np.uint32)), (case, trial, e, float(sh), float(rem))
                checked += a.size
    assert checked > 10000
    if case == "dyadic":
        assert excluded > 0                          # ties do occur there, and are ruled out instead of trusted


def test_a_tie_really_breaks_the_short_form():
    """Why ties are excluded: with shift = u / 2 the exact form rounds to even, i.e. differently for even and odd a / u, and no
    single offset k can reproduce both."""
    e = 10
    u = f32(2.0 ** (e - 23))
    sh, rem = f32(u / 2), f32(0.0)
    a_even, a_odd = f32(2.0 ** e + 100 * float(u)), f32(2.0 ** e + 101 * float(u))
    d_even = float(exact_form(a_even, sh, rem)) - float(a_even)
    d_odd = float(exact_form(a_odd, sh, rem)) - float(a_odd)
    assert d_even != d_odd
    _, (lo, hi) = short_form_tables([sh], rem, e)
    assert lo > hi                                   # ... and the prologue's test sees it
