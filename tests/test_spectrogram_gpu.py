"""GPU parity of the spectrogram front-end (avt_log_spectrogram through audio.log_spectrogram)
against the CPU oracle, which tests/test_spectrogram_cpu.py pins to scipy.signal.spectrogram.
Tolerance: both sides compute in fp64 and round to fp32 once -> 2e-6 absolute on log(S)/12
(values span about [-1.4, 0.3]); the FFT orders differ, so not bit-exact."""
import numpy as np
import pytest
import torch

from oracle import spec_oracle as so

pytestmark = pytest.mark.gpu

TOL = 2e-6


def _gpu(x, fs, lengths=None):
    from audio_visual_tubes_b200.audio import log_spectrogram
    return log_spectrogram(torch.from_numpy(x).cuda(), fs, lengths).cpu().numpy()


@pytest.mark.parametrize("fs", [16000, 22050, 44100])
def test_ten_second_clips_match_oracle(fs):
    x = so.synthetic_waveforms(3, fs, seed=fs)
    got = _gpu(x, fs)
    t = (fs * 10 - 1) // 511
    assert got.shape == (3, 1, 257, t) and got.dtype == np.float32
    for i in range(3):
        want = so.log_spectrogram(x[i], fs)
        assert np.abs(got[i] - want).max() <= TOL


def test_committed_scipy_golden():
    """The frozen output of scipy.signal.spectrogram through dataloader.py:88-96
    (tests/golden/make_golden_spectrogram.py); needs neither scipy nor the oracle's transform."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "spectrogram_golden.npz"))
    fs = int(g["fs"])
    wav = so.synthetic_waveforms(2, fs, seed=int(g["seed"]), seconds=10.0)
    got = _gpu(wav, fs, [int(n) for n in g["lengths"]])
    assert np.abs(got[:, 0][:, g["bins"]] - g["spec"]).max() <= TOL


def test_short_clips_are_tiled_like_the_dataloader():
    fs = 16000
    lens = [fs * 10, 37123, 511, 1, 160001]
    x = so.synthetic_waveforms(len(lens), fs, seed=3, seconds=10.5)
    got = _gpu(x, fs, lens)
    for i, n in enumerate(lens):
        want = so.log_spectrogram(x[i, :n], fs)
        assert np.abs(got[i] - want).max() <= TOL, f"clip {i} (length {n})"


def test_pcm16_input_is_the_wav_file_read_by_soundfile():
    """int16 samples / 32768 is what sf.read returns for a PCM_16 file (dataloader.py:86)."""
    fs = 22050
    lens = [fs * 10, 50001, 700]
    x = so.synthetic_waveforms(3, fs, seed=8)
    pcm = np.clip(np.round(x * 32768.0), -32768, 32767).astype(np.int16)
    pcm[0, :100] = [-32768, 32767] * 50
    got = _gpu(pcm, fs, lens)
    for i, n in enumerate(lens):
        want = so.log_spectrogram(pcm[i, :n].astype(np.float64) / 32768.0, fs)
        assert np.abs(got[i] - want).max() <= TOL, f"clip {i}"
    assert np.array_equal(got, _gpu((pcm.astype(np.float32) / 32768.0), fs, lens))


def test_silence_and_full_scale_clip():
    fs = 16000
    x = np.zeros((3, fs * 10), np.float32)
    x[1] = 3.0                                     # clipped to 1, constant -> detrended to zero
    x[2, ::2] = 1.5; x[2, 1::2] = -1.5             # Nyquist square wave at full scale after the clip
    got = _gpu(x, fs)
    floor = np.float32(np.log(1e-7) / 12.0)
    assert np.all(got[0] == floor) and np.all(got[1] == floor)
    want = so.log_spectrogram(x[2], fs)
    assert np.abs(got[2] - want).max() <= TOL


def test_batch_rows_are_independent_at_training_batch_size():
    fs, b = 16000, 128                             # config 2's batch
    x = so.synthetic_waveforms(4, fs, seed=11)
    xb = np.tile(x, (b // 4, 1))
    rng = np.random.default_rng(0)
    perm = rng.permutation(b)
    got = _gpu(xb[perm], fs)
    single = _gpu(x, fs)
    assert np.array_equal(got, single[perm % 4])
    assert np.isfinite(got).all()
    for _ in range(10):                            # same bits every launch (no race in the exchanges)
        assert np.array_equal(_gpu(xb[perm], fs), got)


def test_row_stride_and_empty_batch():
    from audio_visual_tubes_b200.audio import log_spectrogram
    fs = 16000
    x = so.synthetic_waveforms(2, fs, seed=5, seconds=10.5)     # rows longer than n_keep
    got = _gpu(x, fs)
    assert np.abs(got[1] - so.log_spectrogram(x[1], fs)).max() <= TOL
    e = log_spectrogram(torch.empty(0, fs * 10, device="cuda"), fs)
    assert e.shape == (0, 1, 257, 313)


def test_bad_arguments_raise():
    from audio_visual_tubes_b200 import _lib
    from audio_visual_tubes_b200.audio import log_spectrogram
    with pytest.raises(ValueError):
        log_spectrogram(torch.zeros(2, 1000, device="cuda"), 16000)
    with pytest.raises(ValueError):
        log_spectrogram(torch.zeros(2, 160000, device="cuda", dtype=torch.float64), 16000)
    with pytest.raises(ValueError):
        log_spectrogram(torch.zeros(2, 160000, device="cuda"), 16000, lengths=[0, 5])
    lib = _lib.load()
    x = torch.zeros(1, 160000, device="cuda"); w = torch.ones(512, device="cuda", dtype=torch.float64)
    o = torch.empty(1, 257, 313, device="cuda")
    rc = lib.avt_log_spectrogram(_lib.ptr(x), None, 1, 160000, 160000, _lib.ptr(w), 1.0, 1e-7, 12.0,
                                 _lib.ptr(o), 312, _lib.stream_ptr())
    assert rc != 0
    with pytest.raises(RuntimeError):
        _lib.check(rc, "avt_log_spectrogram")


def test_feeds_the_model_like_the_dataloader_output():
    """waveform batch -> log_spectrogram -> AVENet.forward: the tensor takes the place of the
    DataLoader's `spec` (train_hardway_1frame.py:127-129) without reshaping."""
    import types
    import audio_visual_tubes_b200 as avt
    fs, b = 16000, 4
    spec = avt.log_spectrogram(torch.from_numpy(so.synthetic_waveforms(b, fs, seed=2)).cuda(), fs)
    args = types.SimpleNamespace(epsilon=0.65, epsilon2=0.4, tri_map=True, Neg=True)
    torch.manual_seed(0)
    net = avt.AVENet(args, False).cuda().eval()
    with torch.no_grad():
        A, logits, Pos, Neg = net(torch.randn(b, 3, 224, 224, device="cuda"), spec)
    assert A.shape == (b, 1, 14, 14) and logits.shape == (b, b + 2)
    assert torch.isfinite(logits).all() and torch.isfinite(A).all()
