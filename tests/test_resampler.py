"""The acquisition resampler (gnss_flowgraph.cc:1063-1118, acq_conf.cc:93-111): GNU Radio's low-pass tap
design and decimating FIR, restated in oracle/oracle_resampler.py, against the engine's C-ABI."""
import importlib.util
import os

import numpy as np
import pytest

import synth
from conftest import ROOT


@pytest.fixture(scope="module")
def orx():
    spec = importlib.util.spec_from_file_location("oracle_resampler", os.path.join(ROOT, "oracle", "oracle_resampler.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


CASES = [(25e6, 4e6), (30e6, 4e6), (8e6, 4e6), (12.5e6, 2e6), (4e6, 4e6)]      # (fs_in, opt_acq_fs) of the adapters


def test_decimation_rule(orx):
    # ConfigureAutomaticResampler: largest divisor of fs_in not above fs_in / opt_freq
    assert [orx.automatic_resampler(fs, opt)[0] for fs, opt in CASES] == [5, 6, 2, 5, 1]
    assert orx.automatic_resampler(25e6, 4e6)[1] == 5000000


def test_tap_design_equals_an_independent_windowed_sinc(orx, b200):
    """firdes::low_pass restated twice (oracle in numpy, engine in C) and pinned on scipy.signal.firwin, which implements the
    same design (Hamming-windowed sinc, unity DC gain); the C-ABI call needs no device."""
    from scipy.signal import firwin
    for fs, opt in CASES[:4]:
        D, acq_fs = orx.automatic_resampler(fs, opt)
        want = orx.firdes_low_pass(1.0, fs, acq_fs / 2.1, acq_fs / 2)
        assert want.size % 2 == 1 and want.size == (int(53.0 * fs / (22.0 * acq_fs / 2)) | 1)
        ref = firwin(want.size, acq_fs / 2.1, window="hamming", fs=fs)
        np.testing.assert_allclose(want, ref, rtol=0, atol=2e-7)
        assert abs(float(want.astype(np.float64).sum()) - 1.0) < 1e-6
        got = b200.firdes_low_pass(1.0, fs, acq_fs / 2.1, acq_fs / 2)
        assert got.size == want.size
        np.testing.assert_allclose(got, want, rtol=0, atol=1e-8)
    with pytest.raises(b200.B200Error):
        b200.firdes_low_pass(1.0, 4e6, 3e6, 1e6)       # cut-off above fs / 2


# known_taps of GNU Radio's own QA for this function (gr-filter/python/filter/qa_firdes.py, test_low_pass:
# firdes.low_pass(1, 1, 0.4, 0.2, WIN_HAMMING)), as published in the GNU Radio source tree. GNU Radio is not installed here and
# there is no network to re-fetch the file: the vector is quoted, and both restatements reproduce it to the last printed digit.
GR_QA_LOW_PASS_1_1_04_02 = (0.0024871660862118006, -4.403502608370943e-18, -0.014456653036177158, 0.0543283149600029, -0.116202212870121,
                            0.17504146695137024, 0.7976038455963135, 0.17504146695137024, -0.116202212870121, 0.0543283149600029,
                            -0.014456653036177158, -4.403502608370943e-18, 0.0024871660862118006)


def test_tap_design_reproduces_gnuradio_qa_vector(orx, b200):
    want = np.array(GR_QA_LOW_PASS_1_1_04_02, dtype=np.float32)
    for taps in (orx.firdes_low_pass(1.0, 1.0, 0.4, 0.2), b200.firdes_low_pass(1.0, 1.0, 0.4, 0.2)):
        assert taps.dtype == np.float32 and taps.size == 13
        # GNU Radio's QA compares with assertFloatTuplesAlmostEqual(..., 5); here: the same binary32 values but for the 1e-18 zero crossing
        assert np.array_equal(np.delete(taps, [1, 11]), np.delete(want, [1, 11]))
        assert abs(taps[1]) < 1e-9 and abs(taps[11]) < 1e-9


@pytest.mark.gpu
def test_fir_decimator_streams_like_fir_filter_ccf(orx, b200):
    rng = np.random.default_rng(3)
    fs, opt = 25e6, 4e6
    D, acq_fs = orx.automatic_resampler(fs, opt)
    taps = orx.firdes_low_pass(1.0, fs, acq_fs / 2.1, acq_fs / 2)
    x = (rng.standard_normal(200000) + 1j * rng.standard_normal(200000)).astype(np.complex64)
    want = orx.FirDecimator(D, taps).work(x)
    r = b200.AcqResamplerB200(D, taps, 100000)
    assert r.latency == (taps.size - 1) // 2
    # one stream in uneven work() calls (multiples of the decimation, one shorter than the history) == one call
    cuts = [0, 50000, 50000 + D * 3, 120000, 200000]
    got = np.concatenate([r.work(x[a:b])[0] for a, b in zip(cuts[:-1], cuts[1:])])
    assert got.size == want.size == x.size // D
    np.testing.assert_allclose(got, want, rtol=0, atol=2e-6 * np.abs(want).max())
    # the very first output only sees the first input item (zeros before the stream)
    assert got[0] == pytest.approx(np.complex64(taps[0]) * x[0], rel=1e-6)
    r.reset()
    again, _ = r.work(x[:50000])
    np.testing.assert_array_equal(again, got[:50000 // D])
    with pytest.raises(b200.B200Error):
        r.work(x[:D + 1])                             # not a multiple of the decimation
    r.close()


@pytest.mark.gpu
def test_resampled_block_feeds_the_search_on_the_device(orx, b200, oracle):
    """25 MS/s antenna stream -> decimate by 5 on the device -> PCPS search at 5 MS/s straight from the device buffer; the
    oracle searches the oracle-decimated block. Acq_delay_samples goes back to the input rate with the resampler ratio
    and the filter latency (pcps_acquisition.cc:699-705)."""
    fs, opt = 25e6, 4e6
    D, acq_fs = orx.automatic_resampler(fs, opt)
    taps = orx.firdes_low_pass(1.0, fs, acq_fs / 2.1, acq_fs / 2)
    n_in = 25000 + 5 * (taps.size - 1)                 # one code period plus the filter's fill
    truth = [dict(system="gps_l1_ca", prn=9, delay=7013, doppler=-1830.0, cn0=52.0)]
    x = synth.make_iq(n_in, fs, truth, seed=8)
    r = b200.AcqResamplerB200(D, taps, n_in)
    dec, dptr = r.work(x)
    want_dec = orx.FirDecimator(D, taps).work(x)
    skip = taps.size - 1                               # decimated items: drop the start-up transient
    N = 5000
    code = synth.adapter_code("gps_l1_ca", 9, acq_fs, 1)[:N]
    e = b200.PcpsAcquisitionB200(N, N, 5000, 250, acq_fs, 5, 1)
    e.set_local_codes([0], code[None, :])
    e.dwell_device_async(dptr + 8 * skip, np.array([0], dtype=np.uint32))
    got = e.sync(1)[0]
    a = oracle.acq(fft_size=N, consumed_samples=N, doppler_max=5000, doppler_step=250, fs=acq_fs, samples_per_chip=5, samples_per_code=float(N))
    a.set_local_code(code)
    want = a.dwell(want_dec[skip:skip + N])
    assert int(got["code_phase"]) == want["code_phase"] and int(got["doppler_hz"]) == want["doppler_hz"]
    assert got["test_statistic"] == pytest.approx(want["test_statistic"], rel=1e-4)
    # back at the input rate: delay * ratio - latency lands on the injected code phase (within a decimated sample)
    delay_in = (float(got["code_phase"]) + skip) * D - r.latency
    assert abs((delay_in - truth[0]["delay"] + 12500) % 25000 - 12500) <= D
    assert abs(got["doppler_hz"] - truth[0]["doppler"]) <= 250
    a.close()
    e.close()
    r.close()
