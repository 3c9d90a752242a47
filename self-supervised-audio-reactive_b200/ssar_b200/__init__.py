"""ssar_b200 — B200-native (sm_100a) drop-in for the audio-feature + latent-trajectory hot path of
JCBrouwer/self-supervised-audio-reactive.  Mirrors the reference's module layout:

    ssar_b200.features.audio / .processing / .rosa.{spectral,convert,beat,constantq,pitch}
    ssar_b200.supervised.data.audio2features
    ssar_b200.random.{latent,noise,patch}
    ssar_b200.models.latent_n_noise2.FixedLatentNoiseDecoder

All compute runs in hand-written CUDA kernels behind the C ABI in include/ssar_b200.h; there is no CPU path.
"""
__version__ = "0.1.0"
