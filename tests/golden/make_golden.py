"""Generates the committed golden fixtures by running the UNMODIFIED reference (/root/reference) on CPU.

Run in the build container only:  python tests/golden/make_golden.py [group ...]
Each group writes tests/golden/<group>.npz holding the seeded synthetic inputs' parameters and the reference's
outputs.  Inputs are regenerated from `ssar_b200.synth.synth_clip` (deterministic), so only outputs are stored.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, str(ROOT / "self-supervised-audio-reactive_b200"))

from ref_loader import load_reference  # noqa: E402
from ssar_b200.synth import projection_matrix, synth_clip, synth_features, synth_patch_inputs  # noqa: E402

SR = 24576


def save(group, **arrays):
    out = {}
    for k, v in arrays.items():
        if torch.is_tensor(v):
            v = v.detach().cpu().numpy()
        out[k] = np.asarray(v)
    np.savez_compressed(HERE / f"{group}.npz", **out)
    print(group, {k: v.shape for k, v in out.items()})


def g_spectral(ref):
    """short ragged clip: frame placement + spectra (n_fft 2048 and the CQT-style rectangular 1024)."""
    n = 8192 + 100
    x = synth_clip(n, SR, 0)
    D = ref.spectral.stft(x)
    save(
        "spectral_short", n=n, sr=SR,
        stft_re=D.real, stft_im=D.imag,
        spec_p1=ref.spectral.spectrogram(x, power=1),
        spec_p2_hop512=ref.spectral.spectrogram(x, hop_length=512, power=2),
        stft_rect1024_re=ref.spectral.stft(x, n_fft=1024, hop_length=256, window=None).real,
        stft_rect1024_im=ref.spectral.stft(x, n_fft=1024, hop_length=256, window=None).imag,
        mel_basis_sum=ref.spectral.mel(SR, 2048).sum(1),
        mel_basis_11025_sum=ref.spectral.mel(SR, 2048, fmax=11025.0).sum(1),
    )


def g_frame_features(ref):
    """2 s (+300 samples) clip: mel / onset / rms / mfcc / flatness at both hops used by the configs."""
    n = SR * 2 + 300
    x = synth_clip(n, SR, 1)
    med = lambda *a, **k: torch.median(*a, **k).values
    save(
        "frame_features", n=n, sr=SR, clip_index=1,
        mel=ref.spectral.melspectrogram(x, SR),
        mel_hop512=ref.spectral.melspectrogram(x, SR, hop_length=512),
        onset=ref.beat.onset_strength(x, SR),
        onset_hop512=ref.beat.onset_strength(x, SR, hop_length=512),
        onset_median=ref.beat.onset_strength(x, SR, aggregate=med),
        rms=ref.audio.rms(x, SR),
        rms_hop512=ref.audio.rms(x, SR, hop_length=512),
        mfcc=ref.audio.mfcc(x, SR),
        flatness=ref.audio.spectral_flatness(x, SR),
    )
    # 44.1 kHz variant of BASELINE config 2 (n_fft 2048, hop 512)
    n2 = 44100 + 77
    x2 = synth_clip(n2, 44100, 2)
    save(
        "frame_features_44k", n=n2, sr=44100, clip_index=2,
        spec_p2=ref.spectral.spectrogram(x2, 2048, 512, power=2)[::8],
        mel=ref.spectral.melspectrogram(x2, 44100, 2048, 512),
        onset=ref.beat.onset_strength(x2, 44100, 512, 2048),
        rms=ref.audio.rms(x2, 44100, 2048, 512),
    )


def g_hpss_filters(ref):
    """HPSS waveforms, biquad band splits and the entry resampler on short clips."""
    n = SR * 2 + 300
    x = synth_clip(n, SR, 3)
    D = ref.spectral.stft(x)
    H, P = ref.spectral.hpss(D, margin=8.0)
    save(
        "hpss_filters", n=n, sr=SR, clip_index=3,
        hpss_h_abs=H.abs()[::4], hpss_p_abs=P.abs()[::4],
        harmonic=ref.audio.harmonic(x), percussive=ref.audio.percussive(x),
        low=ref.processing.low_pass(x, SR), mid=ref.processing.mid_pass(x, SR), high=ref.processing.high_pass(x, SR),
    )
    import torchaudio

    n2 = 44100 + 17
    x2 = synth_clip(n2, 44100, 4)
    y2 = torchaudio.functional.resample(x2, 44100, SR)
    y3 = torchaudio.functional.resample(y2, SR, SR / 2, resampling_method="kaiser_window")
    save("resample", n=n2, clip_index=4, to24576=y2, half_kaiser=y3)


def g_post(ref):
    """processing.py on a synthetic [T, C] feature matrix (+ short-sequence, 1-D, 3-D and 4-D gaussian cases)."""
    P = ref.processing
    x = synth_features(700, 7, 0)
    x3 = synth_features(64, 6, 1).reshape(64, 2, 3)
    x4 = synth_features(48, 8, 2).reshape(48, 2, 2, 2)
    short = synth_features(10, 3, 3)
    import contextlib, io

    with contextlib.redirect_stdout(io.StringIO()):
        g_short = P.gaussian_filter(short, 5.0)
    save(
        "post",
        gauss_s2=P.gaussian_filter(x, 2.4), gauss_s10_reflect=P.gaussian_filter(x, 10, mode="reflect"),
        gauss_1d=P.gaussian_filter(x[:, 0], 3), gauss_col1=P.gaussian_filter(x[:, :1], 3),
        gauss_3d=P.gaussian_filter(x3, 1), gauss_4d=P.gaussian_filter(x4, 2), gauss_short=g_short,
        normalize=P.normalize(x), standardize=P.standardize(x),
        q25=P.quantile(x, 0.25), q975=P.quantile(x[:, 3].contiguous(), 0.975),
        flux=P.spectral_flux(x), onset_env=P.onset_envelope(P.spectral_flux(x)),
        clamp_peaks=P.clamp_peaks_percentile(x, 97.5), clamp_peaks_1d=P.clamp_peaks_percentile(x[:, 2], 90),
        clamp_upper=P.clamp_upper_percentile(x, 95), clamp_lower=P.clamp_lower_percentile(x, 10),
        emphasize=P.emphasize(x, 2, 75), emphasize_10_50=P.emphasize(x[:, :1], 10, 50),
    )


def g_chroma(ref):
    """tuning estimate, CQT, chroma (cqt / cens), chromagram, tonnetz, spectral contrast on a 4 s ragged clip."""
    n = SR * 4 + 300
    x = synth_clip(n, SR, 5)
    tun = ref.pitch.estimate_tuning(x, SR, bins_per_octave=36)
    tun12 = ref.pitch.estimate_tuning(x, SR)
    C = ref.constantq.cqt(x, SR, n_bins=252, bins_per_octave=36, tuning=None)
    C0 = ref.constantq.cqt(x, SR, n_bins=252, bins_per_octave=36, tuning=0.0)
    cens = ref.spectral.chroma_cens(x, SR)
    save(
        "chroma", n=n, sr=SR, clip_index=5, tuning36=tun, tuning12=tun12,
        cqt_re=C.real, cqt_im=C.imag, cqt0_abs=C0.abs(),
        chroma_cqt=ref.spectral.chroma_cqt(x, SR), cens=cens,
        chromagram=ref.audio.chromagram(x, SR), tonnetz=ref.audio.tonnetz(x, SR),
        tonnetz_of_cens=ref.audio.tonnetz(x, SR, chroma_fn=lambda a, sr: cens),
        contrast=ref.audio.spectral_contrast(x, SR), contrast_linear=ref.audio.spectral_contrast(x, SR, linear=True),
    )


def g_a2f(ref):
    """plp / pulse / onsets / drop_strength on a 45 s clip (T = 1080 >= the 1024-frame tempogram window) and the
    full audio2features [1080, 59] of a 45 s 44.1 kHz clip (supervised/data.py:60-107)."""
    n = SR * 45 + 100
    x = synth_clip(n, SR, 6)
    x44 = synth_clip(44100 * 45 + 100, 44100, 8)
    save(
        "a2f", n=n, sr=SR, clip_index=6, n44=len(x44), clip_index44=8,
        plp=ref.beat.plp(x, SR), pulse=ref.audio.pulse(x, SR), onsets=ref.audio.onsets(x, SR),
        drop_strength=ref.audio.drop_strength(x, SR),
        features=ref.data.audio2features(x44, 44100, 24),
        features_raw=ref.data.audio2features(x44, 44100, 24, clamp=False, smooth=False),
    )


def g_a2f_short(ref):
    """audio2features on clips SHORTER than the 1024-frame tempogram window (rosa/beat.py:46-47: win_length = T): 10 s,
    16 s (the reference generators' own clip length, features/generate.py:51-54) and 42 s at 44.1 kHz -> T = 240, 384,
    1008; plus plp / pulse on a 24 576 Hz clip of T = 300 and an odd T = 333."""
    import contextlib, io

    out = {}
    for secs, ci in ((10, 11), (16, 12), (42, 13)):
        x44 = synth_clip(44100 * secs, 44100, ci)
        with contextlib.redirect_stdout(io.StringIO()):  # T = 240 prints the short-sequence Gaussian warning (sigma 10)
            out[f"features_{secs}s"] = ref.data.audio2features(x44, 44100, 24)
            out[f"features_raw_{secs}s"] = ref.data.audio2features(x44, 44100, 24, clamp=False, smooth=False)
    for T, ci in ((300, 14), (333, 15)):
        x = synth_clip(1024 * T + 17, SR, ci)
        out[f"plp_T{T}"] = ref.beat.plp(x, SR)
        med = lambda *a, **k: torch.median(*a, **k).values
        out[f"onset_median_T{T}"] = ref.beat.onset_strength(x, SR, aggregate=med)
    save("a2f_short", **out)


def g_mir(ref):
    """The UNMODIFIED retrieve_music_information (random/mir.py:25-46) on a 20 s clip: its feature dict (the part on the
    path: 8 feature functions -> gaussian sigma 2 -> salience_weighted -> normalize).  Tempo / beats / segmentations come
    from stand-ins (ref_loader) and are not stored."""
    n = SR * 20 + 100
    x = synth_clip(n, SR, 16)
    feats, _, _ = ref.mir.retrieve_music_information(x, SR, device="cpu")
    env = synth_features(500, 3, 21).abs()
    save("mir", n=n, sr=SR, clip_index=16, salience_2d=ref.mir.salience_weighted(env), salience_1d=ref.mir.salience_weighted(env[:, 0]),
         **{f"feat_{k}": v for k, v in feats.items()})


def g_traj(ref):
    """Patch(seed).forward latents / noise batches, spline loops, latent_patch, FixedLatentNoiseDecoder.
    Full [T, 18, 512] tensors are stored as a fixed random projection (512 -> 8) plus a few whole frames."""
    T = 600
    R = projection_matrix()
    out = {}
    for seed in (42, 7):
        feats, segs, palette = synth_patch_inputs(T, 0)
        patch = ref.patch.Patch(feats, segs, 120.0, fps=24, seed=seed, device="cpu")
        lat, noise = patch.forward(palette, downscale_factor=4)
        out[f"lat_proj_{seed}"] = lat @ R
        out[f"lat_frames_{seed}"] = lat[::120]
        for j in (0, 3, 8, 12, 16):
            nz = noise[j].forward(10, 7)
            out[f"noise_{seed}_{j}"] = nz if nz.shape[-1] <= 64 else nz[:, ::8, ::8]
        out[f"noise_tail_{seed}"] = noise[6].forward(T - 3, 32)
        out[f"n_sub_{seed}"] = np.array([len(patch.latent_patches), len(patch.noise_patches)])
    feats, segs, palette = synth_patch_inputs(T, 0)
    out["spline_proj"] = ref.latent.spline_loop_latents(palette[:8], T, 1) @ R
    out["spline_loops_proj"] = ref.latent.spline_loop_latents(palette[8:13], T, 2.37) @ R
    # one explicit latent_patch of every kind on a fresh base (in place)
    rng = torch.Generator("cpu").manual_seed(5)
    lat = ref.latent.spline_loop_latents(palette[:6], T)
    for kind, merge, depth, sf in (("segmentation", "average", "low", "mfcc"), ("feature", "modulate", "midhigh", "chromagram"),
                                   ("feature", "modulate", "all", "onsets"), ("loop", "overwrite", "high", "tonnetz")):
        lat = ref.latent.latent_patch(rng, lat, palette, segs, feats, 120.0, 24, kind, 6, 8, sf, 0.8, "rms", 0.9, merge, depth)
    out["latent_patch_proj"] = lat @ R
    # FixedLatentNoiseDecoder
    torch.manual_seed(0)
    dec = ref.lnn2.FixedLatentNoiseDecoder(palette[:9], hidden_size=3, n_latent_split=3, n_noise=4)
    x = synth_features(2 * 300, 17, 77).abs().reshape(2, 300, 17) + 0.1
    torch.manual_seed(0)
    dl, dn = dec.forward(x)
    out["dec_proj"] = dl @ R
    for i, n in enumerate(dn):
        out[f"dec_noise_{i}"] = n[:, ::10]
    save("traj", **out)


def g_maua(ref):
    """maua audioreactive/latent.py + signal.py run unmodified: slerp / spline loops, weighted sums, quantised selection,
    causal Gaussian.  [T, 18, 512] outputs stored as the fixed 512 -> 8 projection plus a few whole frames."""
    from ref_loader import load_maua_audioreactive

    ma = load_maua_audioreactive()
    T = 600
    R = projection_matrix()
    _, _, palette = synth_patch_inputs(T, 0)
    out = {}
    for name, (k0, k1, loops) in {"a": (0, 5, 1), "b": (5, 8, 3), "c": (8, 20, 2)}.items():
        sl = ma.latent.slerp_loops(palette[k0:k1], T, loops)
        out[f"slerp_{name}_proj"], out[f"slerp_{name}_frames"] = sl @ R, sl[::97]
        sp = ma.latent.spline_loops(palette[k0:k1], T, loops)
        out[f"spline_{name}_proj"] = sp @ R
    out["tempo_slerp_proj"] = ma.latent.tempo_loops(palette[:4], T, 24, 96.0, type="slerp") @ R
    env = synth_features(T, 1, 11).abs().reshape(T)
    env = env / env.max()
    out["single_proj"] = ma.latent.single_weighted(palette[0], palette[1], env) @ R
    envs = synth_features(T, 7, 12).abs() + 0.05
    out["multi_proj"] = ma.latent.multi_weighted(palette[:7], envs.clone()) @ R
    out["multi_wrap_proj"] = ma.latent.multi_weighted(palette[:3], envs.clone()) @ R
    sm = ma.latent.select_modulo(palette[:6], env.clone(), smooth=2)
    out["select_proj"], out["select_frames"] = sm @ R, sm[::97]
    x = synth_features(T, 5, 13)
    out["gauss_causal_none"] = ma.signal.gaussian_filter(x, 3.0)
    out["gauss_causal_0"] = ma.signal.gaussian_filter(x, 3.0, causal=0)
    out["gauss_causal_half"] = ma.signal.gaussian_filter(x, 3.0, causal=0.5)
    out["normalize"] = ma.signal.normalize(x)
    save("maua", **out)


def augmenter_ws():
    """The mapped-latent bank a LatentAugmenter holds (16 384 rows): a seeded [256, 18, 512] base repeated 64 times."""
    return torch.randn(256, 18, 512, generator=torch.Generator("cpu").manual_seed(4242)).repeat(64, 1, 1)


def g_widened(ref):
    """Functions whose MODULES cannot be imported (a GAN pickle / decord / maua / matplotlib at import time): their
    unmodified source text is compiled straight from the reference files (ref_loader.exec_reference_defs) and run with
    the reference's own feature functions plus stand-ins for what is off the path (random audio picker, GAN mapper,
    get_video).  generate.py:106-136 audio2features variant, :155-230 builders; optimize.py:103-142 decoder;
    supervised/latent_augmenter.py:14-105; maua StyleGAN2 wrapper make_noise_pyramid :193-210."""
    import random
    import types

    from ref_loader import TorchOnCpu, exec_reference_defs, exec_reference_methods

    R = projection_matrix()
    out = {}
    # ---- generate.py: audio2features variant + builders on a 16 s clip (DUR = 16, generate.py:52)
    n = SR * 16
    audio = synth_clip(n, SR, 18)
    _, _, palette = synth_patch_inputs(8, 1)
    tcpu = TorchOnCpu()
    ns = dict(torch=tcpu, np=np, FPS=24, DUR=16, N=16 * 24, SR=SR,
              harmonic=ref.audio.harmonic, percussive=ref.audio.percussive, mfcc=ref.audio.mfcc,
              chromagram=ref.audio.chromagram, tonnetz=ref.audio.tonnetz, spectral_contrast=ref.audio.spectral_contrast,
              spectral_flatness=ref.audio.spectral_flatness, rms=ref.audio.rms, drop_strength=ref.audio.drop_strength,
              onsets=ref.audio.onsets, onset_strength=ref.beat.onset_strength, plp=ref.beat.plp,
              low_pass=ref.processing.low_pass, mid_pass=ref.processing.mid_pass, high_pass=ref.processing.high_pass,
              gaussian_filter=ref.data.gaussian_filter, normalize=ref.processing.normalize,
              clamp_peaks_percentile=ref.processing.clamp_peaks_percentile,
              clamp_lower_percentile=ref.processing.clamp_lower_percentile,
              get_random_audio=lambda: (audio, SR), get_video=lambda latents, noise: (latents, noise),
              G=types.SimpleNamespace(mapper=lambda z: palette[: len(z)]))
    exec_reference_defs("ssar/features/generate.py", ["audio2features", "high_chroma_correlation", "high_onset_correlation",
                                                      "high_both_correlation"], ns)
    out["gen_n"], out["gen_clip_index"] = n, 18
    out["gen_a2f_raw"] = ns["audio2features"](audio, SR, clamp=False, smooth=False)
    out["gen_a2f"] = ns["audio2features"](audio, SR)
    out["gen_chroma"], out["gen_onsets"] = ref.audio.chromagram(audio, SR), ref.audio.onsets(audio, SR)
    for name, k, size in (("high_chroma_correlation", 12, 32), ("high_onset_correlation", 2, 64), ("high_both_correlation", 14, 64)):
        torch.manual_seed(100 + k)  # the test repeats: manual_seed, randn(k, 512) (the mapper's z), randn(T, 1, s, s)
        _, _, (lat, noise), _ = ns[name]()
        out[f"{name}_proj"], out[f"{name}_noise"] = lat @ R, noise[::16, :, ::4, ::4]
        out[f"{name}_noise_std"] = noise.std(dim=(1, 2, 3))
    # ---- optimize.py FixedLatentNoiseDecoder
    T = 720
    ns = dict(torch=torch, gaussian_filter=ref.processing.gaussian_filter)
    exec_reference_defs("ssar/optimize.py", ["FixedLatentNoiseDecoder"], ns)
    x = synth_features(T, 27 + 8, 31)
    dec = ns["FixedLatentNoiseDecoder"](palette_27(), 3, 3, 3)
    torch.manual_seed(0)  # the test repeats: manual_seed(0), randn(T, s, s) for s = 4, 8, 16, 32
    lat, noise = dec.forward(x)
    out["opt_proj"] = lat @ R
    for i, nz in enumerate(noise):
        out[f"opt_noise_{i}"] = nz[:, ::2, ::2] if nz.shape[-1] > 8 else nz
    # ---- maua StyleGAN2Synthesizer.make_noise_pyramid (+ the synthesizer's own resize of mismatched noise, :86-96)
    sizes = [4, 8, 8, 16, 16, 32, 32, 64, 64, 128, 128]  # noise_const sizes of the first 11 synthesis layers
    bs = []
    names = []
    for li, sz in enumerate(sizes):
        b, c = (li + 1) // 2, "conv1" if li % 2 == 0 else "conv0"
        while len(bs) <= b:
            bs.append(types.SimpleNamespace())
        setattr(bs[b], c, types.SimpleNamespace(noise_const=torch.empty(sz, sz)))
        names.append(f"bs.{b}.{c}")
    fake = types.SimpleNamespace(layer_names=names, G_synth=types.SimpleNamespace(bs=bs))
    ns = exec_reference_methods("maua/maua/GAN/wrappers/stylegan2.py", "StyleGAN2Synthesizer", ["make_noise_pyramid"],
                                dict(torch=torch))
    for tag, hw in (("16", (16, 16)), ("32", (32, 32)), ("64", (64, 64)), ("12x20", (12, 20))):
        nz = torch.randn(40, 1, *hw, generator=torch.Generator("cpu").manual_seed(77))
        pyr = ns["make_noise_pyramid"](fake, nz)
        assert list(pyr) == [f"noise{l}" for l in range(9)]
        for l in range(9):
            out[f"pyr{tag}_{l}"] = pyr[f"noise{l}"][::8] if pyr[f"noise{l}"].shape[-1] <= 32 else pyr[f"noise{l}"][::8, :, ::2, ::2]
    nz = torch.randn(6, 1, 16, 16, generator=torch.Generator("cpu").manual_seed(78))
    out["resize_16_to_64"] = torch.nn.functional.interpolate(nz, (64, 64), mode="bicubic", align_corners=False)  # stylegan2.py:95
    # ---- supervised/latent_augmenter.py: spline_loop_latents twin + LatentAugmenter.random_patch
    from ref_loader import _NaturalCubicSpline, _natural_cubic_spline_coeffs

    ws = augmenter_ws()

    class Mapper:  # StyleGAN2Mapper stand-in: the bank is an input here
        def __init__(self, *a, **k):
            pass

        def eval(self):
            return self

        def forward(self, latent_z):
            return ws

    ns = dict(torch=torch, np=np, random=random, NaturalCubicSpline=_NaturalCubicSpline,
              natural_cubic_spline_coeffs=_natural_cubic_spline_coeffs, normalize=ref.processing.normalize,
              StyleGAN2Mapper=Mapper)
    exec_reference_defs("ssar/supervised/latent_augmenter.py", ["spline_loop_latents", "LatentAugmenter"], ns)
    feats = synth_features(384, 59, 33)
    for seed in (0, 1, 2, 3):
        la = ns["LatentAugmenter"]("ckpt", n_patches=5 if seed < 3 else 9)
        np.random.seed(seed)
        random.seed(seed)
        res, off = la.random_patch(feats.clone())
        out[f"aug_res_proj_{seed}"], out[f"aug_off_{seed}"] = res @ R, off
        out[f"aug_res_frames_{seed}"] = res[::128]
    save("widened", **out)


def g_slices(ref):
    """supervised/data.py:255-258 overlapping_slices and the :317-319 window statistics (np.mean / np.std over all windows
    and frames) on synthetic feature matrices."""
    out = {}
    for T, L in ((1000, 192), (768, 192), (777, 96), (5760, 192)):
        x = synth_features(T, 59, 50 + T % 7)
        w = ref.data.overlapping_slices(x, L)
        out[f"first_row_{T}_{L}"] = torch.tensor([int((x[:, 0] == v).nonzero()[0]) for v in w[:, 0, 0]])
        out[f"n_{T}_{L}"] = w.shape[0]
    tracks = [synth_features(T, 59, 60 + i) for i, T in enumerate((1000, 2111, 777))]
    allw = np.concatenate([ref.data.overlapping_slices(f, 192).numpy() for f in tracks])
    out["win_mean"], out["win_std"] = np.mean(allw, axis=(0, 1)), np.std(allw, axis=(0, 1))
    save("slices", **out)


from video_synth import synth_video  # noqa: E402


def g_video(ref):
    """ssar/features/video.py:11-31, 58-73 (module imports cv2 + kornia: the functions' own text is compiled and run)."""
    from ref_loader import exec_reference_defs

    ns = exec_reference_defs("ssar/features/video.py", ["redogram", "greenogram", "blueogram", "rgb_hist", "visual_variance",
                                                      "absdiff"], dict(torch=torch))
    out = {}
    for tag, kw in (("a", dict()), ("b", dict(T=129, H=33, W=65, seed=6)), ("c", dict(T=2, H=16, W=16, seed=7))):
        v = synth_video(**kw)
        out[f"red_{tag}"], out[f"green_{tag}"], out[f"blue_{tag}"] = ns["redogram"](v), ns["greenogram"](v, 20), ns["blueogram"](v)
        out[f"rgb_{tag}"] = ns["rgb_hist"](v)
        out[f"var_{tag}"], out[f"absdiff_{tag}"] = ns["visual_variance"](v), ns["absdiff"](v)
    save("video", **out)


def palette_27():
    return synth_patch_inputs(8, 1)[2][:27]


GROUPS = {"video": g_video, "slices": g_slices, "widened": g_widened, "mir": g_mir, "a2f_short": g_a2f_short, "maua": g_maua, "traj": g_traj, "a2f": g_a2f, "chroma": g_chroma, "post": g_post, "spectral": g_spectral, "frame_features": g_frame_features, "hpss_filters": g_hpss_filters}

if __name__ == "__main__":
    torch.manual_seed(0)
    ref = load_reference()
    names = sys.argv[1:] or list(GROUPS)
    with torch.inference_mode():
        for name in names:
            GROUPS[name](ref)
