"""Golden vectors for the spectrogram front-end, made with scipy.signal.spectrogram itself by the
statements of datasets/dataloader.py:88-96 (the reference's dataset module does not import here:
soundfile / torchvideotransforms are absent, so its lines are replayed on a synthetic waveform).

    python tests/golden/make_golden_spectrogram.py      # writes spectrogram_golden.npz (~30 KB)
"""
import os
import sys

import numpy as np
from scipy import signal

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from oracle import spec_oracle as so  # noqa: E402  (only its seeded waveform generator)

FS = 8000
KEEP_BINS = slice(0, 257, 4)          # 65 of the 257 bins, all 156 frames


def reference_getitem(samples, samplerate):
    """dataloader.py:88-96 verbatim in effect."""
    if samples.shape[0] < samplerate * 10:
        n = int(samplerate * 10 / samples.shape[0]) + 1
        samples = np.tile(samples, n)
    resamples = samples[:samplerate * 10]
    resamples[resamples > 1.] = 1.
    resamples[resamples < -1.] = -1.
    frequencies, times, spectrogram = signal.spectrogram(resamples, samplerate, nperseg=512, noverlap=1)
    spectrogram = np.log(spectrogram + 1e-7)
    return ((spectrogram - 0.0) / 12.0).astype(np.float32)        # ToTensor + Normalize(0, 12) + .float()


if __name__ == "__main__":
    wav = so.synthetic_waveforms(2, FS, seed=21, seconds=10.0)
    lengths = np.array([FS * 10, 30011], np.int32)
    out = np.stack([reference_getitem(wav[i, :lengths[i]].astype(np.float64).copy(), FS)[KEEP_BINS]
                    for i in range(2)])
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "spectrogram_golden.npz")
    np.savez_compressed(path, fs=FS, seed=21, lengths=lengths, spec=out, bins=np.arange(257)[KEEP_BINS])
    print(path, out.shape, os.path.getsize(path))
