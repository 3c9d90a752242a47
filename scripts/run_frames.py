"""frames_fft over one chunk (8 clips x 240 s at 24576 Hz, n_fft 2048, hop 1024): the onset mel launch, the full launch over
the audio (complex + magnitude + mel + flatness + rms) and the complex-only launch.  The command ncu profiles."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
import torch  # noqa: E402

from ssar_b200 import ops, plan  # noqa: E402
from ssar_b200.synth import synth_batch_device  # noqa: E402

sr, n = 24576, 24576 * 240
x = synth_batch_device(8, n, sr, torch.device("cuda"), first_index=0)
T = n // 1024
onset_bank = plan.mel_bank(float(sr), 2048, 128, 0.0, 11025.0, "cuda")
mel_bank = plan.mel_bank(float(sr), 2048, 128, 0.0, None, "cuda")
cases = {
    "onset mel": lambda: ops.frames_fft(x, 2048, 1024, T, torch.hann_window, banks=[(onset_bank, True, True)]),
    "full": lambda: ops.frames_fft(x, 2048, 1024, T + 1, torch.hann_window, want_complex=True, spec_power=1, want_flatness=True,
                                   want_rms=True, banks=[(mel_bank, True, True)], aux_frames=T),
    "complex": lambda: ops.frames_fft(x, 2048, 1024, T + 1, torch.hann_window, want_complex=True, want_rms=True, aux_frames=T),
}
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
for name, fn in cases.items():
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name}: {e0.elapsed_time(e1) / reps * 1e3:.1f} us")
