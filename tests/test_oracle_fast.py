"""The vectorised numpy/scipy CPU baseline (oracle/oracle_fast.py) against the C oracle — CPU only."""
import numpy as np
import pytest

import synth


def test_fast_acquisition_matches_c_oracle(oracle, oracle_mod):
    import oracle_fast
    geo, codes, iq, truth = synth.config_scene("C1", n_prn=4)
    fa = oracle_fast.FastAcq(geo["fft_size"], geo["consumed_samples"], geo["doppler_max"], geo["doppler_step"], geo["fs"],
                             geo["samples_per_chip"], use_cfar=True, workers=2)
    specs = [fa.code_spectrum(c) for c in codes]
    for mode in (fa.dwell_all, fa.dwell_all_like_reference):
        got = mode(iq, specs)
        for p in range(4):
            a = oracle.acq(fft_size=geo["fft_size"], consumed_samples=geo["consumed_samples"], doppler_max=geo["doppler_max"],
                           doppler_step=geo["doppler_step"], fs=geo["fs"], samples_per_chip=geo["samples_per_chip"],
                           samples_per_code=geo["samples_per_code"], use_cfar=True)
            a.set_local_code(codes[p])
            a.reset()
            w = a.dwell(iq)
            assert got[p]["test_statistic"] == pytest.approx(w["test_statistic"], rel=1e-4)
            if p in (0, 2):  # PRN 1 and 3 are in the scene
                assert got[p]["code_phase"] == w["code_phase"] and got[p]["doppler_index"] == w["doppler_index"]
    # wipe-off table: same float32 running sum as the C oracle
    assert np.abs(fa.wipe - a.wipeoffs()).max() < 3e-7  # numpy and glibc cosf/sinf differ by an ulp


def test_fast_multicorrelator_matches_c_oracle(oracle):
    import oracle_fast
    rng = np.random.default_rng(3)
    n = 8000
    sig = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    code = rng.choice([-1.0, 1.0], 1023).astype(np.float32)
    shifts = np.array([-0.5, 0.0, 0.5], np.float32)
    got = oracle_fast.mc_real_fast(sig, code, shifts, 0.4, 0.003, 0.3, 0.25575)
    want = oracle.mc_real(sig, code, shifts, 0.4, 0.003, 0.0, 0.3, 0.25575, 0.0)
    assert np.abs(got - want).max() <= 1e-4 * np.abs(want).max()
