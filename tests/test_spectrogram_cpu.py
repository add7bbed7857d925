"""Pin the spectrogram oracle (oracle/spec_oracle.py) to scipy.signal.spectrogram itself, the
third-party routine datasets/dataloader.py:94 calls.  CPU only."""
import numpy as np
import pytest

from oracle import spec_oracle as so

signal = pytest.importorskip("scipy.signal")


@pytest.mark.parametrize("fs,seconds", [(16000, 10.0), (22050, 10.0), (8000, 3.7)])
def test_oracle_equals_scipy_spectrogram(fs, seconds):
    x = so.synthetic_waveforms(2, fs, 5, seconds)[1]
    re = so.prepare_samples(x, fs)
    assert re.shape[0] == fs * 10 and re.max() <= 1.0 and re.min() >= -1.0
    f, t, ref = signal.spectrogram(re, fs, nperseg=512, noverlap=1)         # dataloader.py:94
    got = so.psd_spectrogram(re, fs)
    assert got.shape == ref.shape == (257, (fs * 10 - 1) // 511)
    assert np.abs(got - ref).max() <= 1e-12 * ref.max()
    want = (np.log(ref + 1e-7) / 12.0).astype(np.float32)
    assert np.abs(so.log_spectrogram(x, fs)[0] - want).max() < 1e-6


def test_window_is_scipys_default():
    assert np.array_equal(so.tukey_periodic(), signal.get_window(("tukey", 0.25), 512))
    from audio_visual_tubes_b200.audio import tukey_window_periodic
    assert np.array_equal(tukey_window_periodic(), so.tukey_periodic())


def test_oracle_matches_committed_scipy_golden():
    """tests/golden/make_golden_spectrogram.py ran dataloader.py:88-96 with scipy and froze the result."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "spectrogram_golden.npz"))
    fs, wav = int(g["fs"]), so.synthetic_waveforms(2, int(g["fs"]), seed=int(g["seed"]), seconds=10.0)
    for i, n in enumerate(g["lengths"]):
        got = so.log_spectrogram(wav[i, :n], fs)[0][g["bins"]]
        assert np.abs(got - g["spec"][i]).max() <= 1e-6
