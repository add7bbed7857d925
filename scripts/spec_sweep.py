"""Time avt_log_spectrogram (128 clips, 16 kHz, L2 flushed) -- the standalone timing loop the ncu captures in profiles/ run."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import spec_oracle as so
from audio_visual_tubes_b200.audio import log_spectrogram
fs, bs = int(os.environ.get("FS", 16000)), 128
dev = "cuda"
wav = torch.from_numpy(so.synthetic_waveforms(4, fs, 9)).to(dev).repeat(bs // 4, 1).contiguous()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for _ in range(20): log_spectrogram(wav, fs)
tot = 0.0; ts = []
for _ in range(50):
    flush.fill_(1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); o = log_spectrogram(wav, fs); b.record(); torch.cuda.synchronize()
    ts.append(a.elapsed_time(b))
ts.sort(); ms = ts[len(ts) // 2]
byt = bs * (fs * 10 * 4 + 257 * o.shape[-1] * 4)
ref = so.log_spectrogram(wav[1].cpu().numpy(), fs)
print(f"fs={fs} ms={ms:.4f} (min {ts[0]:.4f}) GB/s={byt/ms/1e6:.0f} err={abs(o[1].cpu().numpy()-ref).max():.2e}")
