#!/usr/bin/env python
"""bench.py — audio-seconds/second of the hot path on N B200s of one node (contract in the task statement).

Workloads (a "step" is one pass over one batch of synthetic clips per GPU):
  full (default) : BASELINE.json's metric "full features + latents": audio2features [T, 59] (resample, HPSS x2, CQT
                   chroma, tonnetz, mfcc, contrast, flatness, 4 onset envelopes, PLP, 8 RMS / drop-strength columns,
                   percentile clamps, Gaussian smoothing) + the Patch latent trajectory [T, 18, 512] for every clip;
                   --clips 4-minute 44.1 kHz clips per GPU (configs[2] + configs[3] at the batch size of configs[1]).
  cfg2           : configs[1]: STFT power spectrogram + mel + onset strength + RMS, 64 clips, n_fft 2048, hop 512.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload full|cfg2] [--clips B]

value       : whole-job audio-s/s with the waveforms already resident in HBM (CUDA events, max over ranks)
e2e         : same metric through the public API from pinned HOST buffers, H2D + D2H copies inside the timed region
roofline    : the dominant kernel of the step, timed live with CUDA events around each of its launches
cpu_baseline: the CPU oracle port timed on this box's host cores on a bounded sample (rank 0, N=1 only)
--impl reference : times the CPU implementation of the same workload (oracle port: the reference is Python + torch and
                   cannot travel to the GPU box; the port is pinned to it at 2e-7) on all host threads.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
for p in (ROOT / "self-supervised-audio-reactive_b200", ROOT / "oracle"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))

import torch  # noqa: E402

SR, DUR_S, N_FFT, HOP2 = 44100, 240, 2048, 512
FPS = 24
DEFAULT_CLIPS = {"full": 32, "cfg2": 64, "cfg5": 1000}
CFG5_SECONDS = (120.0, 360.0)  # SURVEY 8d cfg 5: track durations ~ U[120, 360] s, seeded


def workload_name(kind, clips):
    if kind == "cfg2":
        return f"cfg2: STFT+mel+onset+RMS, {clips} clips x {DUR_S} s @44.1kHz per GPU, n_fft {N_FFT} hop {HOP2}"
    if kind == "cfg5":
        return (f"cfg5: audio2features + Patch latents for {clips} tracks of U[{CFG5_SECONDS[0]:.0f}, {CFG5_SECONDS[1]:.0f}] s "
                f"@44.1kHz IN TOTAL (strong scaling), sharded by track (longest-first bin packing), ragged chunks of 8, "
                f"per-feature (count, sum, sum^2) all-reduced over NCCL")
    return (f"full: audio2features [5760,59] + Patch latents [5760,18,512], {clips} clips x {DUR_S} s @44.1kHz per GPU "
            f"(resampled to {FPS * 1024} Hz, n_fft {N_FFT} hop 1024, {FPS} fps)")


def measured_peaks():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        d = json.loads(f.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, indices, enabled=True):
        """indices: GPU indices to watch (ONE nvidia-smi process for all of them: rank 0 samples the whole job, so N ranks
        do not run N pollers against the driver during the timed region)."""
        self.index, self.rows, self.proc, self.enabled = ",".join(str(i) for i in indices), [], None, enabled

    def _nvml_loop(self, handles, mx):
        import pynvml as nv
        bits = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        # ONE GPU per tick, round robin (two NVML calls per 50 ms): polling all 8 GPUs of a busy node in one burst
        # contends with the 8 launching processes for the driver's locks (it cost milliseconds per step and, on one box,
        # returned nothing inside a 280 ms timed region)
        i = 0
        while not self.stop.is_set():
            h, m = handles[i % len(handles)], mx[i % len(handles)]
            i += 1
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.rows.append([str(sm), str(m)] + ["Active" if rs & bits[n] else "Not Active" for n in
                                                     ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")])
            except Exception:
                pass
            self.stop.wait(0.05)

    def __enter__(self):
        if not self.enabled:
            return self
        # NVML in this process (the library nvidia-smi itself reads): two calls per GPU every 100 ms.  A looping
        # nvidia-smi process polling 8 GPUs cost 2.4 ms of a 48 ms step at N = 8 (driver locks); it stays as the fallback.
        try:
            import pynvml as nv
            nv.nvmlInit()
            handles = [nv.nvmlDeviceGetHandleByIndex(int(i)) for i in self.index.split(",")]
            mx = [nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM) for h in handles]  # (before the timed region)
            self.stop = threading.Event()
            self.source = "nvml"
            self.thread = threading.Thread(target=self._nvml_loop, args=(handles, mx), daemon=True)
            self.thread.start()
            return self
        except Exception:
            self.stop = None
        try:
            self.source = "nvidia-smi"
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if getattr(self, "stop", None) is not None:
            time.sleep(0.12)
            self.stop.set()
            self.thread.join(timeout=2)
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for n, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm), "source": getattr(self, "source", None)}


# ------------------------------------------------------------------------------------------------
# CPU arm (oracle port) — used by cpu_baseline and by --impl reference
# ------------------------------------------------------------------------------------------------
def cpu_pass(kind, x, sr):
    from ssar_oracle import dsp

    if kind == "cfg2":
        S = dsp.spectrogram(x, N_FFT, HOP2, power=2)
        return S, dsp.mel_basis(sr, N_FFT) @ S, dsp.onset_strength(x, sr, HOP2, N_FFT), dsp.rms(x, sr, N_FFT, HOP2)
    from ssar_b200.synth import synth_patch_inputs
    from ssar_oracle import features, traj

    # the reference's two call sites for one track: audio2features (supervised/data.py:60-107) for the dataset features
    # and retrieve_music_information's feature part (random/sample.py:34,61; mir.py:26,44) -> Patch -> latents
    f = features.audio2features(x, sr, FPS)
    _, segs, palette = synth_patch_inputs(f.shape[0], 0)
    pf = features.retrieve_music_features(dsp.resample(x, sr, FPS * 1024), FPS * 1024)
    lat, _ = traj.Patch(pf, segs, 120.0, fps=FPS, seed=42).forward(palette, downscale_factor=4)
    return f, lat


def cpu_sample_seconds(kind):
    return 60 if kind == "cfg2" else DUR_S  # full: one whole clip of the GPU arm's workload (240 s)


def use_all_host_threads():
    """torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the CPU arm is the reference's torch-CPU code with
    every host core it can use, so the thread count is set explicitly (and reported)."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    torch.set_num_threads(max(1, n))
    return torch.get_num_threads()


def time_cpu(kind, steps, warmup):
    from ssar_b200.synth import synth_clip

    s = cpu_sample_seconds(kind)
    x = synth_clip(SR * s, SR, 0)
    for _ in range(warmup):
        cpu_pass(kind, x, SR)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_pass(kind, x, SR)
    dt = (time.perf_counter() - t0) / steps
    return s / dt, dt, s


def run_reference(args, rank):
    if rank != 0:
        return
    cores = use_all_host_threads()
    v, dt, s = time_cpu(args.workload, args.steps, 1 if args.warmup > 0 else 0)
    sample = (f"1 clip x {s} s of the workload per step (oracle port of the reference's torch-CPU code, fp32, "
              f"{cores} threads)")
    print(json.dumps({
        "impl": "reference", "metric": "audio-sec/sec", "value": v, "unit": "audio-s/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args.workload, args.clips), "sample": sample},
        "cpu_baseline": {"value": v, "unit": "audio-s/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def time_cfg4(dev, palette, segs, T, rng_device):
    """configs[3]: (a) noise delivery of one Patch(seed 42): every `forward(i, 32)` batch of all 17 layers over the 5760
    frames at downscale 4 and 1, CUDA events around the whole sweep (after one warm-up sweep of the first 4 batches),
    GB/s against 4 * sum(H * W) bytes per frame written; (b) FixedLatentNoiseDecoder(latents [9, 18, 512], H 3, S 3,
    n 4) on [1, 5760, 17] envelopes -> latents [1, 5760, 18, 512] + noise 4..32."""
    from ssar_b200.models.latent_n_noise2 import FixedLatentNoiseDecoder
    from ssar_b200.random.patch import Patch
    from ssar_b200.synth import synth_features, synth_patch_inputs

    peak, _ = measured_peaks()
    feats, _, _ = synth_patch_inputs(T, 0)
    feats = {k: v.to(dev) for k, v in feats.items()}
    out = {}
    for ds in (4, 1):
        patch = Patch(feats, segs, 120.0, fps=FPS, seed=42, device=rng_device)
        _, noise = patch.forward(palette, downscale_factor=ds)
        for i in range(0, 4 * 32, 32):
            for m in noise:
                m.forward(i, 32)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        nbytes, n_calls = 0, 0
        e0.record()
        for i in range(0, T, 32):
            for m in noise:
                o = m.forward(i, min(32, T - i))
                nbytes += o.numel() * 4
                n_calls += 1
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        out[f"noise_downscale_{ds}"] = {"ms": round(ms, 2), "launch_calls": n_calls, "bytes_written": nbytes,
                                         "GBps": round(nbytes / (ms * 1e-3) / 1e9, 1),
                                         "frac_of_hbm_peak": round(nbytes / (ms * 1e-3) / 1e9 / peak, 4),
                                         "audio_s_per_s": round(T / FPS / (ms * 1e-3), 1)}
        del noise, patch
    x = synth_features(T, 3 * 3 + 2 * 4, 31).abs().add_(0.01)[None].to(dev)
    dec = FixedLatentNoiseDecoder(palette[:9], hidden_size=3, n_latent_split=3, n_noise=4)
    for _ in range(2):
        dec(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        lat, nz = dec(x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    nb = lat.numel() * 4 + sum(n.numel() * 4 for n in nz)
    out["fixed_decoder"] = {"ms": round(ms, 3), "bytes_written": nb, "GBps": round(nb / (ms * 1e-3) / 1e9, 1),
                            "frac_of_hbm_peak": round(nb / (ms * 1e-3) / 1e9 / peak, 4),
                            "audio_s_per_s": round(T / FPS / (ms * 1e-3), 1)}
    return out


class EventLog(dict):
    """cabi.event_log that records every entry point."""

    def get(self, k, d=None):
        return self.setdefault(k, [])


def cfg5_durations(n_tracks):
    g = torch.Generator("cpu").manual_seed(2025)
    return (CFG5_SECONDS[0] + (CFG5_SECONDS[1] - CFG5_SECONDS[0]) * torch.rand(n_tracks, generator=g)).tolist()


def run_cfg5(args, rank, world, local_rank):
    """configs[4]: 1000 tracks of different durations over the node's GPUs.  One step = one pass over ALL tracks (strong
    scaling): every rank processes its shard (longest-first greedy bin packing by duration) through the ragged pipeline,
    accumulates the per-feature (count, sum, sum^2) of its frames in fp64 and the ranks all-reduce them (NCCL) into the
    dataset mean / std."""
    import torch.distributed as dist

    from ssar_b200 import _cabi as cabi
    from ssar_b200 import distributed, ops, pipeline
    from ssar_b200.synth import synth_batch_device, synth_clip, synth_patch_inputs

    sys.stdout.flush()
    result_fd = os.dup(1)
    os.dup2(2, 1)
    result_out = os.fdopen(result_fd, "w")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    durations = cfg5_durations(args.clips)
    mine = distributed.shard_clips(durations, world)[rank]
    lengths = {i: int(round(durations[i] * SR)) for i in mine}
    clips = [synth_batch_device(1, lengths[i], SR, dev, first_index=i)[0] for i in mine]  # resident in HBM
    Tmax = int(CFG5_SECONDS[1] * FPS) + 2
    _, segs, palette = synth_patch_inputs(Tmax, 0)
    palette, segs = palette.to(dev), {k: v.to(dev) for k, v in segs.items()}
    stats = torch.zeros(3, 59, dtype=torch.float64, device=dev)
    total_audio_s = sum(durations)
    steps, warmup = max(1, args.steps), max(1, min(args.warmup, 2))

    def step(src, on_clip=None):
        stats.zero_()
        pipeline.features_and_latents_ragged(src, SR, FPS, palette, segs, rng_device=args.rng, chunk=args.chunk,
                                             on_clip=on_clip or (lambda i, f, lat: None), stats=stats)
        if world > 1:
            dist.all_reduce(stats)
        return stats

    for _ in range(warmup):
        step(clips)
    barrier()
    launches0 = cabi.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(range(world) if world > 1 else [local_rank], enabled=(rank == 0)) as clocks:
        barrier()
        ev0.record()
        for _ in range(steps):
            step(clips)
        ev1.record()
        barrier()
    ms = ev0.elapsed_time(ev1)
    launches = cabi.launch_count - launches0
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = total_audio_s * steps / (ms_max * 1e-3)
    mean, std = distributed.mean_std_from_stats(stats.clone())

    # e2e: the rank's first tracks from PINNED HOST memory (staged per chunk inside the call), features + a latent checksum
    # back to pinned host per track
    e2e = None
    if not args.no_e2e:
        sub = list(range(min(len(clips), 96)))
        host = [clips[j].cpu().pin_memory() for j in sub]
        host_f = [torch.empty(int(lengths[mine[j]] * 2048 // 3675 + 1) // 1024, 59, pin_memory=True) for j in sub]
        host_c = torch.zeros(len(sub), pin_memory=True)

        def sink(i, f, lat):
            host_f[i][: f.shape[0]].copy_(f, non_blocking=True)
            host_c[i : i + 1].copy_(lat.sum().reshape(1), non_blocking=True)

        step(host, sink)
        barrier()
        t0 = time.perf_counter()
        step(host, sink)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        td = torch.tensor([dt, sum(durations[mine[j]] for j in sub)], device=dev, dtype=torch.float64)
        if world > 1:
            tmax = td[:1].clone()
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            dist.all_reduce(td[1:], op=dist.ReduceOp.SUM)
            td[0] = tmax[0]
        e2e = {"value": float(td[1] / td[0]), "unit": "audio-s/s",
               "h2d_bytes_per_step": int(sum(h.numel() * 4 for h in host)), "d2h_bytes_per_step": int(sum(h.numel() * 4 for h in host_f) + 4 * len(sub)),
               "steps": 1, "note": f"per rank: its first {len(sub)} tracks from pinned host memory, features + latent checksum back per track"}

    # per-kernel table of one eager ragged chunk (8 tracks), CUDA events around each launch
    barrier()
    per_kernel = {}
    if rank == 0:
        cabi.event_log = EventLog()
        pipeline.features_and_latents_ragged(clips[:8], SR, FPS, palette, segs, rng_device=args.rng, chunk=8, use_graphs=False,
                                             on_clip=lambda i, f, lat: None)
        torch.cuda.synchronize()
        for name, evs in cabi.event_log.items():
            per_kernel[name] = (sum(a.elapsed_time(b) for a, b, _ in evs), sum(c for _, _, c in evs if c is not None), len(evs))
        cabi.event_log = None
    verify = None
    if rank == 0 and world == 1 and args.verify_tracks > 0:
        from ssar_oracle import features as ofeat
        import numpy as np
        idx = list(range(min(args.verify_tracks, len(clips))))
        sub_stats = torch.zeros(3, 59, dtype=torch.float64, device=dev)
        pipeline.features_and_latents_ragged([clips[j] for j in idx], SR, FPS, palette, segs, rng_device=args.rng,
                                             on_clip=lambda i, f, lat: None, stats=sub_stats)
        m_gpu, s_gpu = distributed.mean_std_from_stats(sub_stats)
        ref = np.concatenate([ofeat.audio2features(clips[j].cpu(), SR, FPS).numpy() for j in idx])
        keep = [j for j in range(59) if j != 50]
        verify = {"tracks": len(idx), "mean_rel_l2": float(np.linalg.norm(m_gpu.cpu().numpy()[keep] - ref.mean(0)[keep]) / np.linalg.norm(ref.mean(0)[keep])),
                  "std_rel_l2": float(np.linalg.norm(s_gpu.cpu().numpy()[keep] - ref.std(0)[keep]) / np.linalg.norm(ref.std(0)[keep]))}
    if rank == 0:
        peak, peak_src = measured_peaks()
        top = sorted(per_kernel.items(), key=lambda kv: -kv[1][0])
        name, (k_ms, k_bytes, k_n) = top[0]
        roof = {"bound": "hbm", "kernel": name, "achieved": k_bytes / (k_ms * 1e-3) / 1e9 if k_bytes else None, "peak": peak,
                "peak_source": peak_src, "unit": "GB/s", "traffic": None,
                "timing": "CUDA events around each launch of ONE eager ragged chunk (the 8 longest tracks of rank 0)",
                "top5_ms_per_chunk": {k: round(v[0], 3) for k, v in top[:5]}}
        roof["frac"] = roof["achieved"] / peak if roof["achieved"] else None
        cpu = None
        if world == 1 and not args.no_cpu:
            n_threads = use_all_host_threads()
            x = synth_clip(int(SR * CFG5_SECONDS[0]), SR, 0)
            t0 = time.perf_counter()
            cpu_pass("full", x, SR)
            dt = time.perf_counter() - t0
            cpu = {"value": CFG5_SECONDS[0] / dt, "unit": "audio-s/s", "cores": n_threads, "kind": "port",
                   "sample": f"1 track x {CFG5_SECONDS[0]:.0f} s (the shortest duration of the workload), oracle port, {dt:.2f} s"}
        result_out.write(json.dumps({
            "metric": "audio-sec/sec", "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms_max / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": workload_name("cfg5", args.clips), "tracks_total": args.clips, "tracks_rank0": len(mine),
                       "audio_seconds_total": total_audio_s, "chunk": args.chunk,
                       "l2": f"inputs_exceed_l2 ({sum(lengths.values()) * 4 / 1e9:.1f} GB of waveforms on rank 0)",
                       "parallelism": f"track-sharded x{world} (longest-first greedy bin packing); one NCCL all-reduce of 59x3 fp64 per step",
                       "dataset_mean_head": [round(float(v), 5) for v in mean[:4]], "dataset_std_head": [round(float(v), 5) for v in std[:4]],
                       "verify_vs_oracle": verify},
            "clocks": clocks.summary(), "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu,
        }) + "\n")
        result_out.flush()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="full", choices=["full", "cfg2", "cfg5"])
    ap.add_argument("--verify-tracks", type=int, default=0, help="cfg5: compare mean / std of this many tracks with the CPU oracle")
    ap.add_argument("--clips", type=int, default=None)
    ap.add_argument("--rng", default="cuda", choices=["cpu", "cuda"], help="device of the per-clip Patch generator")
    ap.add_argument("--workers", type=int, default=None, help="host threads building the per-clip patches")
    ap.add_argument("--chunk", type=int, default=8, help="clips per feature-graph replay / sampler batch (full workload)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-cfg4", action="store_true", help="skip the configs[3] noise / decoder legs")
    args = ap.parse_args()
    if args.clips is None:
        args.clips = DEFAULT_CLIPS[args.workload]
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if args.workload == "cfg5":
        return run_cfg5(args, rank, world, local_rank)
    args.warmup = max(args.warmup, 3)

    import torch.distributed as dist

    from ssar_b200 import _cabi as cabi
    from ssar_b200 import ops, pipeline
    from ssar_b200.synth import synth_batch_device, synth_patch_inputs

    # stdout carries exactly ONE JSON line: libraries that write to fd 1 (NCCL prints its version banner there) are
    # pointed at stderr, and the result line goes out through a private duplicate of the original stdout
    sys.stdout.flush()
    result_fd = os.dup(1)
    os.dup2(2, 1)
    result_out = os.fdopen(result_fd, "w")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    B, n = args.clips, SR * DUR_S
    kind = args.workload
    x = synth_batch_device(B, n, SR, dev, first_index=rank * B)
    audio_s_per_step = B * DUR_S  # per rank
    stats = torch.zeros(3, 59, dtype=torch.float64, device=dev)

    if kind == "cfg2":
        T = n // HOP2

        def step(xd, serial=False):
            return pipeline.spectral_frontend(xd, SR, N_FFT, HOP2)

        def results(o):
            return [o["mel"], o["onset"], o["rms"]]
        d2h_note = "mel+onset+rms out to pinned host; the power spectrogram stays in HBM for downstream kernels"
    else:
        T = 5760
        _, segs, palette = synth_patch_inputs(T, 0)
        palette = palette.to(dev)

        stats_ring = [stats, torch.zeros_like(stats)]
        pending = [None, None]
        step_no = [0]

        def step(xd, serial=False):
            feats, latents, _ = pipeline.features_and_latents(xd, SR, FPS, palette, segs, rng_device=args.rng,
                                                              workers=args.workers, serial=serial, chunk=args.chunk)
            # the one cross-clip reduction of the path: per-feature (count, sum, sum^2) for the dataset mean / std.
            # The all-reduce is asynchronous and double buffered: step k's reduction is only waited for when step k + 2
            # reuses its buffer, so the ranks are not put in lock-step by a 1.4 kB collective (the sum of per-step maxima
            # over 8 ranks is what a blocking all-reduce per step measures)
            slot = step_no[0] % 2
            step_no[0] += 1
            buf = stats_ring[slot]
            if pending[slot] is not None:
                pending[slot].wait()
            buf.zero_()
            ops.feature_stats(feats, buf)
            if world > 1:
                pending[slot] = dist.all_reduce(buf, async_op=True)  # NCCL: 59 x 3 fp64
            return {"feats": feats, "latents": latents}

        def results(o):
            # latents feed the generator on the device (maua StyleGAN2 wrapper); the host gets the features and a
            # per-clip fingerprint of the latents
            return [o["feats"], o["latents"][:, ::64].sum(dim=(1, 2, 3))]
        d2h_note = ("features [B,5760,59] + a per-clip latent fingerprint (sum over every 64th frame) out to pinned host; "
                    "latents stay in HBM for the generator")

    for _ in range(args.warmup):
        out = step(x)
    barrier()
    launches0 = cabi.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(range(world) if world > 1 else [local_rank], enabled=(rank == 0 and not os.environ.get("SSAR_BENCH_NO_CLOCKS"))) as clocks:
        barrier()
        ev0.record()
        for _ in range(args.steps):
            out = step(x)
        ev1.record()
        barrier()
    ms = ev0.elapsed_time(ev1)
    launches = cabi.launch_count - launches0
    # per-kernel times: ONE extra (untimed) step with CUDA events around every launch of the C ABI, on the launching
    # stream; kept out of the timed region because two event records per launch cost host time.  The step is queued
    # serially (one thread, one stream): with the sampler / copy side streams active an event interval would also
    # cover whatever else the GPU interleaves, and small kernels would look tens of times slower than they are.
    barrier()
    cabi.event_log = EventLog()
    out = step(x, serial=True)
    barrier()
    per_kernel = {}
    for name, evs in cabi.event_log.items():
        t_ms = sum(a.elapsed_time(b) for a, b, _ in evs) * args.steps  # scaled to the timed region's step count
        nb = sum(c for _, _, c in evs if c is not None) * args.steps
        per_kernel[name] = (t_ms, nb, len(evs) * args.steps)
    cabi.event_log = None
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = audio_s_per_step * world * args.steps / (ms_max * 1e-3)

    # ---- e2e: pinned host waveforms -> H2D -> public API -> D2H of the results, chunked over two streams
    e2e = None
    if not args.no_e2e:
        del out
        chunk = min(8, B)
        host_x = torch.empty(B, n, dtype=torch.float32, pin_memory=True)
        host_x.copy_(x)
        probe = results(step(x[:chunk]))
        host_out = [torch.empty((B,) + tuple(p.shape[1:]), dtype=p.dtype, pin_memory=True) for p in probe]
        del probe
        streams = [torch.cuda.Stream(dev), torch.cuda.Stream(dev)]

        staged = [pipeline.stage_batch(host_x, dev)] if kind == "full" else None

        def e2e_step(drain=True):
            if kind == "full":
                # a dataset job is software-pipelined over its batches (pipeline.stage_batch): every step starts the H2D
                # copy of the NEXT batch from pinned host memory, then runs the public call on the batch staged one step
                # earlier, so a step's interval holds one whole batch copy and one whole batch of compute
                cur = staged[0]
                staged[0] = pipeline.stage_batch(host_x, dev)
                for h, r in zip(host_out, results(step(cur))):
                    h.copy_(r, non_blocking=True)
                if drain:
                    torch.cuda.current_stream().synchronize()
                    staged[0].ready.synchronize()  # the copy this step started has landed before the clock stops
                return
            for ci, c0 in enumerate(range(0, B, chunk)):
                s = streams[ci % 2]
                with torch.cuda.stream(s):
                    xd = host_x[c0 : c0 + chunk].to(dev, non_blocking=True)
                    for h, r in zip(host_out, results(step(xd))):
                        h[c0 : c0 + chunk].copy_(r, non_blocking=True)
            for s in streams:
                s.synchronize()

        for _ in range(3):  # warm-up: the staged path's chunk shapes are captured on their second sight
            e2e_step()
        import gc
        gc.collect()
        gc.disable()  # a generation-2 collection in the middle of a step stalls the launching threads for ~30 ms
        barrier()
        t0 = time.perf_counter()
        e_steps = max(2, min(args.steps, 5))
        per_step = []
        for i in range(e_steps):
            # a dataset job does not drain the device between batches; K batch copies (H2D) and K result copies (D2H) are
            # issued and completed inside the timed region (the batch the first timed step computes on was staged by the
            # last warm-up step, the copy the last timed step starts is waited for before the clock stops)
            ts = time.perf_counter()
            e2e_step(drain=(i == e_steps - 1))
            per_step.append(round((time.perf_counter() - ts) * 1e3, 2))
        gc.enable()
        barrier()
        print(json.dumps({"e2e_ms_per_step": per_step}), file=sys.stderr)
        dt = time.perf_counter() - t0
        td = torch.tensor([dt], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(td, op=dist.ReduceOp.MAX)
        e2e = {
            "value": audio_s_per_step * world * e_steps / float(td.item()), "unit": "audio-s/s",
            "h2d_bytes_per_step": B * n * 4, "d2h_bytes_per_step": sum(h.numel() * h.element_size() for h in host_out),
            "steps": e_steps, "note": "per rank, software-pipelined (stage_batch: the H2D of batch k + 1 overlaps the kernels of batch k; one batch copy per step inside the timed region): pinned host waveforms in, " + d2h_note,
        }

    # the same end-to-end step from 16-bit PCM host buffers (the on-disk width of the audio the reference loads with
    # torchaudio.load; converted on the device as x / 32768): half the H2D bytes.  Informational — `e2e` above stays
    # the fp32-waveform interface of the reference
    e2e_pcm16 = None
    if not args.no_e2e and kind == "full":
        host_pcm = torch.empty(B, n, dtype=torch.int16, pin_memory=True)
        host_pcm.copy_((x * 32767).round_().clamp_(-32768, 32767).to(torch.int16))

        staged_pcm = [pipeline.stage_batch(host_pcm, dev)]

        def pcm_step(drain=True):
            cur = staged_pcm[0]
            staged_pcm[0] = pipeline.stage_batch(host_pcm, dev)
            for h, r in zip(host_out, results(step(cur))):
                h.copy_(r, non_blocking=True)
            if drain:
                torch.cuda.current_stream().synchronize()
                staged_pcm[0].ready.synchronize()

        for _ in range(3):
            pcm_step()
        barrier()
        t0 = time.perf_counter()
        for i in range(e_steps):
            pcm_step(drain=(i == e_steps - 1))
        barrier()
        td = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(td, op=dist.ReduceOp.MAX)
        e2e_pcm16 = {"value": audio_s_per_step * world * e_steps / float(td.item()), "unit": "audio-s/s",
                     "h2d_bytes_per_step": B * n * 2, "steps": e_steps,
                     "note": "same step from pinned int16 PCM host buffers (x / 32768 on the device)"}
        del host_pcm

    # latency of ONE 4-minute track through the public call (north-star target: < 50 ms), host clock, device drained
    single = None
    if kind == "full" and world == 1:  # (step() holds a collective when world > 1: all ranks or none)
        x1 = x[:1]
        for _ in range(2):
            step(x1)
        torch.cuda.synchronize()
        ts = []
        for _ in range(5):
            t0 = time.perf_counter()
            step(x1)
            torch.cuda.synchronize()
            ts.append((time.perf_counter() - t0) * 1e3)
        ts.sort()
        single = {"ms_median": ts[len(ts) // 2], "ms_min": ts[0], "clips": 1,
                  "note": "one 240 s clip resident in HBM -> features [5760,59] + latents [5760,18,512], wall clock"}

    # configs[3] (SURVEY 8d cfg 4) legs, N = 1 only: (a) every noise batch of one 4-minute Patch — all 180
    # `forward(i, 32)` calls of the 17 layers at downscale 4 and 1 — and (b) the fixed decoder on [1, 5760, 17] envelopes
    cfg4 = None
    if kind == "full" and world == 1 and not args.no_cfg4:
        cfg4 = time_cfg4(dev, palette, segs, T, args.rng)

    if rank == 0:
        peak, peak_src = measured_peaks()
        top = sorted(per_kernel.items(), key=lambda kv: -kv[1][0])
        name, (k_ms, k_bytes, k_n) = top[0]
        kernel_total = sum(v[0] for v in per_kernel.values())
        table = {k: {"ms_per_step": round(v[0] / args.steps, 4), "launches_per_step": v[2] // args.steps,
                     "GBps": round(v[1] / (v[0] * 1e-3) / 1e9, 1) if v[1] and v[0] > 0 else None,
                     "frac": round(v[1] / (v[0] * 1e-3) / 1e9 / peak, 4) if v[1] and v[0] > 0 else None}
                 for k, v in top}
        print(json.dumps({"per_kernel_serial_pass": table}), file=sys.stderr)
        roof = {
            "bound": "hbm", "kernel": name, "achieved": (k_bytes / (k_ms * 1e-3) / 1e9) if k_bytes else None,
            "peak": peak, "peak_source": peak_src, "unit": "GB/s", "traffic": None,
            "kernel_ms_per_step": k_ms / args.steps, "launches_per_step": k_n / args.steps,
            "algorithmic_bytes_per_step": k_bytes / args.steps, "share_of_step": k_ms / ms if ms > 0 else None,
            "all_kernels_share_of_step": kernel_total / ms if ms > 0 else None,
            "top5_ms_per_step": {k: round(v[0] / args.steps, 3) for k, v in top[:5]},
            "top5_frac": {k: table[k]["frac"] for k, _ in top[:5]},
            "timing": "CUDA events around each launch on the launching stream, one extra step queued serially",
        }
        roof["frac"] = roof["achieved"] / peak if roof["achieved"] else None
        # DRAM traffic of the dominant kernel from the committed ncu --set full captures (dram__bytes_read + write summed
        # over its launches of one 8-clip chunk, scaled to this run's clip count): profiles/r02_ncu_sel_v2.md (full: the 8
        # frames_fft launches of a chunk move 2.396 GB) and profiles/r01_frames_fft_v1.md (cfg2: 2.316 GB per 16 clips)
        ncu_bytes_per_clip = {"full": 2.396e9 / 8, "cfg2": 2.3160e9 / 16}
        if name == "ssar_frames_fft":
            roof["traffic_per_step"] = ncu_bytes_per_clip[kind] * B
            roof["traffic"] = roof["traffic_per_step"] / max(k_n / args.steps, 1)  # per launch, like `achieved`
            roof["algorithmic_bytes_per_launch"] = k_bytes / max(k_n, 1)
            roof["traffic_source"] = ("ncu --set full capture of the same kernels (profiles/r02_ncu_sel_v2.md), average per launch"
                                      if kind == "full" else "ncu capture profiles/r01_frames_fft_v1.md, average per launch")
        cpu = None
        if world == 1 and not args.no_cpu:
            n_threads = use_all_host_threads()
            v, dt, s = time_cpu(kind, 1 if kind == "full" else 3, 0 if kind == "full" else 1)
            cpu = {"value": v, "unit": "audio-s/s", "cores": n_threads, "kind": "port",
                   "sample": f"1 clip x {s} s of the same workload, oracle port (torch-CPU fp32), {dt:.2f} s per pass"}
        result_out.write(json.dumps({
            "metric": "audio-sec/sec", "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(kind, B), "clips_per_gpu": B, "frames_per_clip": T,
                       "single_track": single, "cfg4": cfg4,
                       "l2": f"inputs_exceed_l2 ({B * n * 4 / 1e9:.2f} GB of waveforms per step)",
                       "parallelism": f"clip-sharded x{world}; one asynchronous NCCL all-reduce of 59x3 fp64 feature stats per step"
                       if kind == "full" else f"clip-sharded x{world}, no data-path collective",
                       "patch_rng": args.rng if kind == "full" else None,
                       "host_threads": (args.workers or pipeline.default_patch_workers(args.rng)) if kind == "full" else 1},
            "clocks": clocks.summary(), "e2e": e2e, "e2e_pcm16": e2e_pcm16, "gpu_launches": launches, "roofline": roof,
            "cpu_baseline": cpu,
        }) + "\n")
        result_out.flush()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
