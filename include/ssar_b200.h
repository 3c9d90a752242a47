/* ssar_b200 — C ABI of the B200-native (sm_100a) audio-feature / latent-trajectory hot path.
 *
 * The reference (JCBrouwer/self-supervised-audio-reactive) has no FFI layer for this path: its boundary is
 * the Python function surface of ssar/features/{audio,processing}.py, ssar/features/rosa/*.py,
 * ssar/supervised/data.py:60-107 (audio2features) and ssar/random/{latent,noise,patch}.py.  Every entry point
 * below replaces the ATen/torchaudio work behind one (or a fused group) of those functions; the reference
 * file:line each one stands in for is cited on the declaration.  The Python package `ssar_b200` keeps the
 * reference's names and argument meaning and binds these symbols with ctypes.
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless the name ends in `_host`;
 *   - the caller allocates every output and scratch buffer and passes the CUDA stream to launch on
 *     (a `cudaStream_t` cast to void*; NULL = legacy default stream);
 *   - spectra are FRAME-MAJOR: [clip, frame, bin] (the reference's [bin, frame] is a transposed view);
 *   - return value: 0 on success, negative on error (SSAR_ERR_*); `ssar_last_error()` returns a message;
 *   - all entry points are asynchronous with respect to the host unless stated otherwise.
 */
#ifndef SSAR_B200_H
#define SSAR_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSAR_OK 0
#define SSAR_ERR_ARG (-1)
#define SSAR_ERR_CUDA (-2)
#define SSAR_ERR_UNSUPPORTED (-3)

#define SSAR_ABI_VERSION 1

int ssar_abi_version(void);
const char* ssar_last_error(void);

/* ------------------------------------------------------------------------------------------------
 * Framed, windowed real FFT with fused epilogues.
 * Replaces: torch.stft behind rosa/spectral.py:10-21 (stft), :59-62 (spectrogram), :65-70
 * (melspectrogram = mel_basis @ |STFT|^2), features/audio.py:31-37 (rms), rosa/constantq.py:189-191
 * (__cqt_response = fft_basis.mm(stft(window=None))), rosa/pitch.py:40 (piptrack's spectrogram) and
 * rosa/beat.py:33-38 (fourier_tempogram).
 *
 * Frame t of a clip covers padded samples [t*hop, t*hop + n_fft) of the clip reflect-padded by n_fft/2 on
 * both sides (torch.stft center=True, pad_mode="reflect").  The caller chooses n_frames (1 + N/hop for stft,
 * N/hop for spectrogram-style callers that drop the last frame).
 * ---------------------------------------------------------------------------------------------- */

/* a banded real filterbank applied to the per-frame spectrum: out[band] = sum_k w[offset+k] * S[start+k] */
typedef struct {
    int n_bands;            /* 0 = unused */
    int on_power;           /* 1: S = |X|^2, 0: S = |X| */
    const int* start;       /* [n_bands] first bin */
    const int* count;       /* [n_bands] number of bins */
    const int* offset;      /* [n_bands] offset into weights */
    const float* weights;   /* packed band weights */
    float* out;             /* [n_clips, n_frames, ld] */
    long long ld;
    uint32_t* max_key;      /* optional [n_clips]: running max of `out` per clip as an order-preserving key
                               (ssar_fill_u32 it with 0 first); feeds power_to_db's global-max floor */
} ssar_filterbank;

/* a banded complex projection of the complex spectrum: out[row] = scale * sum_k w[offset+k] * X[start+k] */
typedef struct {
    int n_rows;             /* 0 = unused */
    const int* start;
    const int* count;
    const int* offset;
    const float* weights;   /* packed complex (re, im) pairs */
    float* out;             /* complex [n_clips, n_frames, ld] (re, im) */
    long long ld;           /* in complex elements */
    float scale;
} ssar_projection;

typedef struct {
    const float* x;         /* [n_clips, clip_stride] waveform */
    long long n_samples;    /* valid samples per clip */
    long long clip_stride;
    int n_clips;
    int n_fft;              /* 1024 or 2048 */
    int hop;
    int n_frames;           /* frames per clip to produce */
    const float* window;    /* [n_fft] or NULL (rectangular) */
    const float* twiddle;   /* [n_fft] complex (re, im): exp(-2*pi*i*k/n_fft) */
    float* out_complex;     /* optional complex [n_clips, n_frames, ldc] */
    long long ldc;
    float* out_spec;        /* optional [n_clips, n_frames, lds] = |X|^spec_power */
    long long lds;
    int spec_power;         /* 1 or 2 */
    float* out_rms;         /* optional [n_clips, n_frames]: sqrt(mean(x^2)) over the unwindowed frame */
    float* out_flatness;    /* optional [n_clips, n_frames]: exp(mean(log(max(1e-10,|X|^2))))/mean(max(1e-10,|X|^2)),
                               features/audio.py:126-133 */
    ssar_filterbank fb[2];
    ssar_projection proj;
} ssar_frames_args;

int ssar_frames_fft(const ssar_frames_args* args, void* stream);

/* fills n 32-bit words with `value` (used to reset max/min key accumulators) */
int ssar_fill_u32(uint32_t* dst, long long n, uint32_t value, void* stream);


/* ------------------------------------------------------------------------------------------------
 * Per-frame spectral features.  `max_key` arrays hold one order-preserving uint32 key per clip (written by
 * ssar_frames_fft's filterbank epilogue or by ssar_minmax_keys); they carry the clip-global maximum that
 * power_to_db (rosa/convert.py:7-12) floors against.
 * ---------------------------------------------------------------------------------------------- */

/* rosa/beat.py:10-23 onset_strength (after the mel projection): dB -> first difference along time -> relu ->
 * mean (aggregate 0) or torch.median (aggregate 1, rosa/beat.py:42-44) over bands -> left pad `pad_width`
 * zeros, truncate to n_frames.  mel: [n_clips, n_frames, n_bands], out: [n_clips, n_frames]. */
int ssar_onset_flux(const float* mel, const uint32_t* max_key, int n_clips, int n_frames, int n_bands,
                    int pad_width, int aggregate, float top_db, float* out, void* stream);

/* features/audio.py:60-65 mfcc: power_to_db(mel) then the orthonormal DCT-II of rosa/spectral.py:35-56 as an
 * explicit [n_mfcc, n_bands] matrix.  out: [n_clips, n_frames, n_mfcc]. */
int ssar_mfcc(const float* mel, const uint32_t* max_key, const float* dct, int n_clips, int n_frames,
              int n_bands, int n_mfcc, float top_db, float* out, void* stream);

/* rosa/convert.py:7-12 power_to_db with ref 1, amin 1e-10; has_top=0 disables the floor. */
int ssar_power_to_db(const float* x, int n_clips, long long n_per_clip, const uint32_t* max_key,
                     float top_db, int has_top, float* out, void* stream);

/* per-clip min / max as keys (atomically merged into pre-initialised accumulators; either may be NULL) */
int ssar_minmax_keys(const float* x, int n_clips, long long n_per_clip, uint32_t* min_key, uint32_t* max_key,
                     void* stream);

/* features/audio.py:31-37 rms: sqrt(mean(x^2)) over reflect-padded frames of `frame_length` (no window).
 * out: [n_clips, n_frames]; the caller passes n_frames = n_samples / hop (the reference drops the last frame). */
int ssar_rms_frames(const float* x, long long n_samples, long long clip_stride, int n_clips,
                    int frame_length, int hop, int n_frames, float* out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * HPSS, inverse STFT, IIR band splits, resampling
 * ---------------------------------------------------------------------------------------------- */

/* rosa/spectral.py:145-161 hpss on a complex STFT [n_clips, n_frames, ld] (complex, frame-major): 31-tap medians
 * along time / frequency (features/processing.py:74-84, reflect padding) + soft masks (rosa/spectral.py:120-142,
 * power 2, margin).  Either output may be NULL.  Outputs have the input's layout. */
int ssar_hpss(const float* stft, int n_clips, int n_frames, int n_bins, long long ld, float margin_harm,
              float margin_perc, float* out_harm, float* out_perc, void* stream);

/* torch.istft (rosa/spectral.py:24-32), step 1: per-frame inverse real FFT times the synthesis window.
 * spec: complex [n_clips, n_frames, ld]; frames: [n_clips, n_frames, n_fft]. n_fft 1024 or 2048. */
int ssar_irfft_frames(const float* spec, long long ld, int n_clips, int n_frames, int n_fft,
                      const float* window, const float* twiddle, float* frames, void* stream);

/* torch.istft step 2: overlap-add of `frames`, division by the summed squared window, drop n_fft/2 leading
 * samples, produce `length` samples per clip (zeros past the covered range). out: [n_clips, length]. */
int ssar_overlap_add(const float* frames, const float* window, int n_clips, int n_frames, int n_fft, int hop,
                     long long length, float* out, void* stream);

/* torchaudio lfilter for one biquad (features/processing.py:141-150): coefficients already divided by a0.
 * scratch: ssar_biquad_scratch_bytes(n_clips, n_samples) bytes of device memory.  y must not alias x. */
long long ssar_biquad_scratch_bytes(int n_clips, long long n_samples);
int ssar_biquad(const float* x, long long n_samples, long long stride_x, int n_clips, float b0, float b1,
                float b2, float a1, float a2, int clamp, float* y, long long stride_y, void* scratch,
                void* stream);

/* torchaudio.functional.resample (supervised/data.py:63; rosa/constantq.py:83) with banded polyphase taps:
 * taps [max_taps, newf] tap-major, first [newf] = index of the first kept tap of each phase in torchaudio's
 * dense kernel; y[j] = post_scale * sum_k taps[k][j % newf] * xpad[(j / newf) * orig + first + k]. */
int ssar_resample(const float* x, long long n_in, long long stride_in, int n_clips, const float* taps,
                  const int* first, int orig, int newf, int width, int max_taps, float post_scale, float* y,
                  long long n_out, long long stride_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SSAR_B200_H */
