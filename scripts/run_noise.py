"""configs[3] legs on their own (bench.time_cfg4): noise sweeps at downscale 4 / 1 and the fixed decoder."""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

import bench  # noqa: E402
from ssar_b200.synth import synth_patch_inputs  # noqa: E402

dev = torch.device("cuda:0")
_, segs, palette = synth_patch_inputs(5760, 0)
print(json.dumps(bench.time_cfg4(dev, palette.to(dev), segs, 5760, "cuda"), indent=1))
