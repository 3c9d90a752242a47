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
    /* optional (all four or none; on_power banks only): the same weights lane-interleaved for the kernel's warps.
       Bands go in passes of 32 (band 32 p + lane); band b's window starts at the 4-aligned bin lane_start4[b]
       ([n_pass*32]) and every band of pass p spans 4 * lane_count4[p] bins, zero weights padding the rest;
       lane_weights[((lane_offset4[p] + k4) * 32 + lane) * 4 + c] multiplies bin lane_start4[b] + 4 k4 + c.  Windows
       may run up to 64 bins past the spectrum (the kernel keeps that tail zero).  Results are bit-identical to the
       packed form: the extra terms are fmaf(0, finite, acc). */
    const int* lane_start4;
    const int* lane_count4;
    const int* lane_offset4;
    const float* lane_weights;
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
    const int* clip_slot;   /* optional [n_clips]: clip c uses table rows [clip_slot[c]*n_rows, +n_rows) (start/count/
                               offset hold n_slots*n_rows entries) - the per-clip tuned CQT bases */
    const float* row_scale; /* optional: out[row] *= row_scale[slot*row_scale_ld + row] (CQT 1/sqrt(filter length)) */
    long long row_scale_ld;
} ssar_projection;

/* rosa/pitch.py:27-83 piptrack as an epilogue of the frame transform (the magnitude spectrogram it looks at never leaves
 * shared memory): same arithmetic and outputs as ssar_piptrack on sqrt(|X|^2) */
typedef struct {
    int* count;             /* NULL = unused; [n_clips], zeroed by ssar_frames_fft */
    float* pitch;           /* [n_clips, capacity] */
    float* mag;             /* [n_clips, capacity] */
    long long capacity;     /* >= ssar_piptrack_capacity(n_frames, k_lo, k_hi) */
    int k_lo, k_hi;         /* bins fmin <= f < fmax */
    float threshold;
    float sr;
} ssar_piptrack_epilogue;

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
    int aux_frames;         /* 0 = n_frames.  Otherwise every output EXCEPT out_complex (spec, rms, flatness, filterbanks and
                               their max keys, projection) covers only frames [0, aux_frames) of each clip and is laid out
                               [n_clips, aux_frames, ...]: one launch then serves stft (1 + N/hop frames, complex) and the
                               spectrogram-style features that drop the last frame (rosa/spectral.py:60) of the same signal */
    ssar_piptrack_epilogue pip;
    /* RAGGED batches (clips of different lengths in one launch; both NULL = every clip has n_samples / n_frames):
       clip c holds clip_samples[c] <= n_samples valid samples (reflect padding happens at ITS end) and produces
       clip_frames[c] <= n_frames frames (its aux outputs cover clip_frames[c] - (n_frames - aux_frames) frames).  Rows of
       frames beyond a clip's count are written as ZEROS in every output, so whole-tensor maxima of the non-negative
       spectra and every later kernel that honours the same lengths see exactly what a per-clip call would.  Device arrays. */
    const long long* clip_samples;
    const int* clip_frames;
    /* TIME SEGMENTS of long tracks (clip_total = 0: unused).  x[clip] then holds the samples [sample_offset, sample_offset +
       n_samples) of a clip that is clip_total samples long, and local frame t is the clip's frame frame_offset + t: the
       frame is cut at (frame_offset + t) * hop - n_fft / 2 in CLIP coordinates, reflected at the clip's true ends only,
       and read from the segment — the caller ships each segment with an n_fft / 2 halo on both sides (clamped to the
       clip).  Segments of one track then reproduce the whole-track transform frame for frame.  Not with clip_samples. */
    long long sample_offset;
    long long clip_total;
    int frame_offset;
    /* sub_every = s > 1: out_complex and out_rms are written for every s-th frame only (frame t = s u -> row u; complex has
       ceil(n_frames / s) rows per clip, rms aux_frames / s): a transform at hop h also serves the consumer of the SAME signal's
       transform at hop s h (rosa/pitch.py:27-83 runs its STFT at hop 512, features/audio.py:13-24 at hop 1024: frame 2 u of the
       first is frame u of the second, bit for bit).  0 / 1 = every frame.  Not with clip_frames / time segments. */
    int sub_every;
} ssar_frames_args;

int ssar_frames_fft(const ssar_frames_args* args, void* stream);

/* The same transform for ARBITRARY n_fft in [2, 4096] (direct DFT, no radix restriction), restricted to the bins
 * [k_lo, k_hi) of the one-sided spectrum: torch.stft(center=True, pad_mode="reflect") behind rosa/spectral.py:10-21 as
 * rosa/beat.py:46-49 reaches it for a clip shorter than win_length frames (n_fft = min(len(onset_envelope), 1024), hop 1;
 * only the 60-180 BPM bins survive rosa/beat.py:54-57).  n_frames <= 1 + (n_samples + 2*(n_fft/2) - n_fft) / hop.
 * out_complex: [n_clips, n_frames, k_hi - k_lo] (re, im). */
int ssar_frames_dft(const float* x, long long n_samples, long long clip_stride, int n_clips, int n_fft, int hop,
                    int n_frames, const float* window, int k_lo, int k_hi, float* out_complex, void* stream);

/* torch.istft(center=True, length) (rosa/spectral.py:24-32, as rosa/beat.py:69 calls it) for arbitrary n_fft from a
 * spectrum whose only non-zero bins are [k_lo, k_hi): spec complex [n_clips, n_frames, k_hi - k_lo]; out [n_clips, length]. */
int ssar_istft_band(const float* spec, int n_clips, int n_frames, int n_fft, int hop, const float* window, int k_lo,
                    int k_hi, long long length, float* out, void* stream);

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
/* clip_frames: NULL, or int [n_clips] (device) for a RAGGED batch: clip c has clip_frames[c] <= n_frames frames (the time
 * medians reflect at ITS last frame); the rows beyond are written as zeros. */
int ssar_hpss(const float* stft, int n_clips, int n_frames, int n_bins, long long ld, float margin_harm,
              float margin_perc, float* out_harm, float* out_perc, const int* clip_frames, void* stream);

/* torch.istft (rosa/spectral.py:24-32), step 1: per-frame inverse real FFT times the synthesis window.
 * spec: complex [n_clips, n_frames, ld]; frames: [n_clips, n_frames, n_fft]. n_fft 1024 or 2048. */
int ssar_irfft_frames(const float* spec, long long ld, int n_clips, int n_frames, int n_fft,
                      const float* window, const float* twiddle, float* frames, void* stream);

/* torch.istft (rosa/spectral.py:24-32) for hop = n_fft / 2 and n_fft = 2048 in ONE kernel: inverse real FFT per frame,
 * synthesis window, overlap-add, division by the summed squared window, centre trim — the [frames, n_fft] intermediate
 * of ssar_irfft_frames + ssar_overlap_add is never written.  spec: frame-major complex [n_clips, n_frames, ld];
 * out [n_clips, length] (length even, 8-byte aligned). */
/* clip_frames / clip_len: NULL, or device arrays [n_clips] for a RAGGED batch (clip c: clip_frames[c] <= n_frames frames,
 * clip_len[c] <= length samples; the tail of its output row is zeros). */
int ssar_istft_half_overlap(const float* spec, long long ld, int n_clips, int n_frames, int n_fft, const float* window,
                            const float* twiddle, long long length, float* out, const int* clip_frames,
                            const long long* clip_len, void* stream);

/* torch.istft step 2: overlap-add of `frames`, division by the summed squared window, drop n_fft/2 leading
 * samples, produce `length` samples per clip (zeros past the covered range). out: [n_clips, length]. */
int ssar_overlap_add(const float* frames, const float* window, int n_clips, int n_frames, int n_fft, int hop,
                     long long length, float* out, const int* clip_frames, const long long* clip_len, void* stream);

/* torchaudio lfilter for one biquad (features/processing.py:141-150): coefficients already divided by a0.
 * scratch: ssar_biquad_scratch_bytes(n_clips, n_samples) bytes of device memory.  y must not alias x. */
long long ssar_biquad_scratch_bytes(int n_clips, long long n_samples);
int ssar_biquad(const float* x, long long n_samples, long long stride_x, int n_clips, float b0, float b1,
                float b2, float a1, float a2, int clamp, float* y, long long stride_y, void* scratch,
                void* stream);

/* torchaudio.functional.resample (supervised/data.py:63; rosa/constantq.py:83) with banded polyphase taps:
 * taps [max_taps, newf] tap-major, first [newf] = index of the first kept tap of each phase in torchaudio's
 * dense kernel; y[j] = post_scale * sum_k taps[k][j % newf] * xpad[(j / newf) * orig + first + k]. */
/* clip_n_in: NULL, or long long [n_clips] (device) for a RAGGED batch: clip c holds clip_n_in[c] <= n_in samples followed by
 * zeros; its outputs past ceil(newf * clip_n_in[c] / orig) are written as zeros. */
int ssar_resample(const float* x, long long n_in, long long stride_in, int n_clips, const float* taps,
                  const int* first, int orig, int newf, int width, int max_taps, float post_scale, float* y,
                  long long n_out, long long stride_out, const long long* clip_n_in, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Tuning estimate, chroma, CENS, tonnetz
 * ---------------------------------------------------------------------------------------------- */

/* rosa/pitch.py:27-83 piptrack on a magnitude spectrogram [n_clips, n_frames, ld] (n_bins valid bins): per frame,
 * local maxima above threshold * column max inside bins [k_lo, k_hi), parabolic interpolation.  The (pitch, mag)
 * pairs with pitch > 0 are appended to per-clip lists pitch / mag [n_clips, capacity] (unordered), count [n_clips]
 * receives their number.  capacity >= ssar_piptrack_capacity(n_frames, k_lo, k_hi). */
long long ssar_piptrack_capacity(int n_frames, int k_lo, int k_hi);
int ssar_piptrack(const float* mag_spec, long long ld, int n_clips, int n_frames, int n_bins, int k_lo, int k_hi,
                  float threshold, float sr, int n_fft, long long capacity, float* pitch, float* mag, int* count,
                  void* stream);

/* rosa/pitch.py:9-24 + :98-120 estimate_tuning tail: threshold = lower median of the tracked magnitudes; histogram
 * (n_hist bins with the n_hist + 1 `edges` of torch.linspace(-0.5, 0.5, n_hist + 1)) of the bins_per_octave residuals
 * of the pitches with mag >= threshold; tuning_idx[clip] = arg-max bin, tuning_val[clip] = edges[arg-max]. */
int ssar_tuning_from_peaks(const float* pitch, const float* mag, int n_clips, long long capacity, const int* count,
                           const float* edges, int n_hist, float bins_per_octave, int* tuning_idx,
                           float* tuning_val, void* stream);

/* rosa/spectral.py:286-325 chroma_cqt (norm=False) and, with quantize != 0, the first half of chroma_cens
 * (:264-271): |cqt| [n_clips, n_frames, ldc] complex -> fold bins onto chroma_of_bin[n_bins] classes -> (L1 norm ->
 * spline map with the natural-cubic-spline table (knots[n_knots], a, b, c, d [n_knots-1]; c, d used as written in
 * :192-202) -> sigmoid staircase step_h * (floor(w-.5) + ...)).  out: [n_clips, n_frames, n_chroma]. */
int ssar_chroma_from_cqt(const float* cqt, long long ldc, int n_clips, int n_frames, int n_bins,
                         const int* chroma_of_bin, int n_chroma, int quantize, const float* knots,
                         const float* sa, const float* sb, const float* sc, const float* sd, int n_knots,
                         float step_h, float step_alpha, float* out, void* stream);

/* rosa/spectral.py:273-280: correlation along time with window[win_len] (already normalised), zero "same" padding,
 * then (l2_normalize) division by the per-frame L2 norm.  q, out: [n_clips, n_frames, n_chroma]. */
/* clip_frames: NULL, or int [n_clips] (device) for a RAGGED batch (the zero padding starts at each clip's own last frame;
 * rows past it are written as zeros). */
int ssar_cens_smooth(const float* q, int n_clips, int n_frames, int n_chroma, const float* window, int win_len,
                     int l2_normalize, float* out, const int* clip_frames, void* stream);

/* features/audio.py:48-57 tonnetz: out[row] = phi[n_out, n_chroma] @ (chroma[row] / ||chroma[row]||_1) */
int ssar_tonnetz(const float* chroma, long long n_rows, int n_chroma, const float* phi, int n_out, float* out,
                 void* stream);

/* rosa/beat.py:54-66 plp, spectral step, IN PLACE on the Fourier tempogram (complex [n_clips, n_frames, ld], n_bins
 * valid): zero outside bins [k_lo, k_hi) (the 60-180 BPM range), keep only the per-frame maxima of log1p(1e6 |X|),
 * scale by 1 / (sqrt(tiny) + max |X|).  The caller follows with ssar_irfft_frames + ssar_overlap_add (hop 1). */
int ssar_plp_peaks(float* tempogram, long long ld, int n_clips, int n_frames, int n_bins, int k_lo, int k_hi,
                   void* stream);

/* features/audio.py:72-123 spectral_contrast, step 1: per (frame, band) mean of the band_q smallest (valley) and
 * largest (peak) magnitudes of bins [band_start, band_start + band_count) of mag [n_clips, n_frames, ld]; the
 * clip-global maxima of both are merged into keys (ssar_fill_u32 them with the key of -inf first).
 * peak, valley: [n_clips, n_frames, n_bands]. band_q <= 32 and <= band_count; max_count = max(band_count). */
int ssar_contrast_bands(const float* mag, long long ld, int n_clips, int n_frames, int n_bands,
                        const int* band_start, const int* band_count, const int* band_q, int max_count,
                        float* peak, float* valley, uint32_t* peak_key, uint32_t* valley_key, void* stream);

/* step 2: out = power_to_db(peak) - power_to_db(valley) (each floored at its own clip maximum - top_db), or
 * peak - valley when linear != 0 (features/audio.py:120-123) */
int ssar_contrast_finish(const float* peak, const float* valley, const uint32_t* peak_key,
                         const uint32_t* valley_key, int n_clips, long long n_per_clip, int linear, float top_db,
                         float* out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Post-processing of feature matrices [n_clips, n_frames, n_cols] (time-major, columns contiguous)
 * ---------------------------------------------------------------------------------------------- */

/* features/processing.py:10-48 gaussian_filter along time of [n_clips, n_frames, inner]: weights [2*radius+1] are
 * the normalised taps (built by the host exactly like :21-24); mode 0 circular (default), 1 reflect, 2 replicate;
 * radius > n_frames reproduces the short-sequence branch (:30-35: mode-pad by n_frames, then replicate). */
/* clip_frames: NULL, or int [n_clips] (device) for a RAGGED batch: clip c has clip_frames[c] <= n_frames rows (the wrap /
 * reflection happens at ITS length); the rows beyond are written as zeros. */
int ssar_gauss1d(const float* x, int n_clips, int n_frames, long long inner, const float* weights, int radius,
                 int mode, float* out, const int* clip_frames, void* stream);

/* per-(clip, column) minimum / maximum -> [n_clips, n_cols] (either output may be NULL) */
/* clip_frames: NULL, or int [n_clips] (device) for a RAGGED batch: only the first clip_frames[c] rows of clip c are real. */
int ssar_col_minmax(const float* x, int n_clips, int n_frames, int n_cols, float* min_out, float* max_out,
                    const int* clip_frames, void* stream);

/* per-(clip, column) quantile by radix select (exact order statistics).
 *   subset 0: all non-NaN values; 1: strict interior local maxima (features/processing.py:107-115)
 *   interp 0: torch.quantile "linear" (:117,125,129,137); 3: efficient_quantile midpoint
 *             (ssar/analysis/efficient_quantile/efficient_quantile.cpp:72-75, wrapper __init__.py:5-6)
 *   pre_sub/pre_div [n_clips, n_cols] or NULL: select on (x - pre_sub) / pre_div (emphasize, :133-137)
 *   count_out [n_clips, n_cols] or NULL: size of the subset (0 -> out = NaN; the host raises like torch) */
/* clip_frames: NULL, or int [n_clips] (device) for a RAGGED batch: only the first clip_frames[c] rows of clip c are real. */
int ssar_col_quantile(const float* x, int n_clips, int n_frames, int n_cols, int subset, double q, int interp,
                      const float* pre_sub, const float* pre_div, float* out, int* count_out,
                      const int* clip_frames, void* stream);

/* torch.clamp by per-(clip, column) bounds [n_clips, n_cols] (features/processing.py:117,124-129) */
int ssar_clamp_cols(const float* x, int n_clips, int n_frames, int n_cols, const float* lo, const float* hi,
                    float* out, void* stream);

/* features/processing.py:52-55 normalize over each whole clip: (x - min) / ((max - min) + eps); keys from
 * ssar_minmax_keys */
int ssar_normalize(const float* x, int n_clips, long long n_per_clip, const uint32_t* min_key,
                   const uint32_t* max_key, float eps, float* out, void* stream);

/* features/processing.py:132-138 emphasize; col_min/col_max from ssar_col_minmax; scratch: 2*n_clips*n_cols floats */
/* clip_frames: NULL, or int [n_clips] (device) for a RAGGED batch: only the first clip_frames[c] rows of clip c are real. */
int ssar_emphasize(const float* x, int n_clips, int n_frames, int n_cols, const float* col_min,
                   const float* col_max, float strength, double q, float* scratch, float* out,
                   const int* clip_frames, void* stream);

/* features/processing.py:88-89 spectral_flux: out[t] = x[t+1] - x[t] with a zero row appended */
int ssar_spectral_flux(const float* x, int n_clips, int n_frames, long long inner, float* out, void* stream);

/* features/processing.py:92: out[row] = sum_j 0.5 * (x + |x|) */
int ssar_hwr_rowsum(const float* x, long long n_rows, int inner, float* out, void* stream);

/* ssar/supervised/data.py:317-319 dataset statistics: accumulates (count, sum, sum of squares) per column in fp64
 * into stats[3, n_cols] (caller zeroes it; all-reduce it over ranks, then mean = s/n, std = sqrt(ss/n - mean^2)) */
/* clip_frames: NULL, or int [n_rows / frames_per_clip] (device) for a RAGGED batch: rows t >= clip_frames[c] of clip c are
 * skipped (x is [n_clips, frames_per_clip, n_cols]). */
int ssar_feature_stats(const float* x, long long n_rows, int n_cols, double* stats, int frames_per_clip,
                       const int* clip_frames, void* stream);

/* ssar/random/mir.py:14-22 (salience_weighted) after the three Gaussians (ssar_gauss1d: sigma 2 circular -> env, then
 * sigma 5 and sigma 80 in reflect mode of env): out = (smooth_short / smooth_long)^2 * env for a batch of feature matrices
 * [n_clips, n_frames, n_cols] whose columns are grouped into features (col_group[n_cols] in [0, n_groups)), plus the
 * running min / max of every (clip, group) as order-preserving keys [n_clips, n_groups] (ssar_fill_u32 them with
 * 0xFF800000 / 0x007FFFFF first). */
/* clip_frames: NULL, or int [n_clips] (device) for a RAGGED batch: only the first clip_frames[c] rows of clip c are real. */
int ssar_salience_weight(const float* env, const float* smooth_short, const float* smooth_long, int n_clips, int n_frames,
                         int n_cols, const int* col_group, int n_groups, float* out, uint32_t* min_key, uint32_t* max_key,
                         const int* clip_frames, void* stream);

/* ssar/random/mir.py:44's closing normalize (features/processing.py:52-55) per (clip, feature): (w - min) / ((max - min) +
 * eps) with the keys of ssar_salience_weight.  group_start[n_groups]: first column of each group (groups are contiguous
 * column ranges).  out is FEATURE-BLOCKED: clip b, group g = contiguous [n_frames, C_g] at
 * out + b * n_frames * n_cols + n_frames * group_start[g]. */
int ssar_group_normalize(const float* w, int n_clips, int n_frames, int n_cols, const int* col_group, const int* group_start,
                         int n_groups, const uint32_t* min_key, const uint32_t* max_key, float eps, float* out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Trajectory construction: latent sequences [T, num_ws, 512] and per-layer noise maps
 * ---------------------------------------------------------------------------------------------- */

/* out[t, n] = row_alpha[t] * out[t, n] + row_beta[t] * sum_k weights_{g(n)}[t, k] * basis[k, n] for n in
 * [col_begin, col_end), g(n) = n / group_cols, weights_g = weights + g * group_stride (row stride ldw).
 * row_alpha NULL = overwrite (0), row_beta NULL = 1.  Stands in for einsum("TN,NWL->TWL") and every merge of
 * ssar/random/latent.py:35-78, the spline evaluation of :7-13, features/generate.py:167,181,216-218 and
 * models/latent_n_noise2.py:150-161 (per-split weights = one group per split).
 * col_begin and group_cols multiples of 128; basis / out rows 16-byte aligned. */
int ssar_mix_gemm(const float* weights, long long ldw, long long group_stride, int group_cols,
                  const float* basis, long long ldb, float* out, long long ldo, int n_rows, int n_basis,
                  int col_begin, int col_end, const float* row_alpha, const float* row_beta, void* stream);

/* The same product as ssar_mix_gemm for whole rows (no row_alpha / row_beta, all n_cols columns, n_basis <= 192) on the
 * tensor cores: tcgen05.mma kind::f16 with a 3-term fp16 split of both operands (fp32-accurate, fp32 accumulation in
 * TMEM), TMA-fed, A-stationary (csrc/traj_mix_tc.cu).  n_cols / group_cols weight groups; group_cols a multiple of 128.
 * Both operands are scaled by a power of two taken from their amax before the split (any fp32 range is safe) and the
 * scale is undone in the epilogue.  basis_split: NULL, or a block written by ssar_mix_gemm_tc_split_basis for this
 * (basis, n_basis, n_cols) — one palette split for all clips of a batch; `basis` may then be NULL.
 * scratch: ssar_mix_gemm_tc_scratch_bytes(...) bytes; basis_split: ssar_mix_gemm_tc_basis_bytes(...) bytes. */
long long ssar_mix_gemm_tc_scratch_bytes(int n_rows, int n_groups, int n_basis, int n_cols);
long long ssar_mix_gemm_tc_basis_bytes(int n_basis, int n_cols);
int ssar_mix_gemm_tc_split_basis(const float* basis, long long ldb, int n_basis, int n_cols, void* basis_split, void* stream);
int ssar_mix_gemm_tc(const float* weights, long long ldw, long long group_stride, int group_cols, const float* basis,
                     long long ldb, float* out, long long ldo, int n_rows, int n_basis, int n_cols, const void* basis_split,
                     void* scratch, void* stream);

/* weights[t, g, k] = alpha[t] * weights[t, g, k] + beta[t] * seq[t, k] for g in [g_begin, g_end): the average /
 * modulate / overwrite merges of latent.py:68-78 applied to weight matrices [n_rows, n_groups, n_basis].
 * alpha / beta NULL -> the scalars. */
int ssar_weights_merge(float* weights, int n_rows, int n_groups, int n_basis, int g_begin, int g_end,
                       const float* seq, const float* alpha, const float* beta, float alpha_scalar,
                       float beta_scalar, void* stream);

/* One latent sub-patch (ssar/random/latent.py:35-78) applied to weight matrices [n_rows, n_groups, n_basis] in place:
 * raw[t, k] = src_scale * src[t * src_ld + k] (src_mode 0: einsum / spline / smoothed one-hot weights, :39-52) or
 * (f, 1 - f) with f = src_scale * src[t * src_ld] (src_mode 1: the 2-latent lerp of :44-45, n_select == 2);
 * S = circular Gaussian of raw along t with `taps` [2 * radius + 1] (:54; radius 0 = none);
 * beta[t] = mod_scale * mod[t * mod_ld], alpha = 1 - beta when mod != NULL (modulate, :72-75), else the scalars
 * (average 0.5 / 0.5, overwrite 0 / 1; alpha_scalar == 0 never reads the old weights);
 * weights[t, g, p] = alpha * weights[t, g, p] + beta * (p == selection[k] ? S[t, k] : 0) for g in [g_begin, g_end).
 * selection: n_select distinct int64 palette indices in device memory (the randperm prefix of :36). */
int ssar_subpatch_merge(float* weights, int n_rows, int n_groups, int n_basis, int g_begin, int g_end,
                        const long long* selection, int n_select, const float* src, long long src_ld, float src_scale,
                        int src_mode, const float* taps, int radius, const float* mod, long long mod_ld,
                        float mod_scale, float alpha_scalar, float beta_scalar, void* stream);

/* out[t, k] = sum_i taps[i] * (labels[(t + i - radius) mod n_rows] == k): latent.py:39-41's one-hot gather followed by
 * its circular Gaussian, as weights [n_rows, n_classes]. labels: int64 [n_rows] in device memory. */
int ssar_onehot_gauss(const long long* labels, int n_rows, int n_classes, const float* taps, int radius, float* out,
                      void* stream);

/* out[t, j] = mu[t] + sig[t] * x[t, j] (models/latent_n_noise2.py:163-176); mu / sig NULL -> 0 / 1 */
int ssar_rows_affine(const float* x, long long n_rows, long long inner, const float* mu, const float* sig,
                     float* out, void* stream);

/* One op of a compiled noise program (ssar/random/noise.py:4-86).  Leaves write `cur` (or `acc` when to_acc != 0). */
#define SSAR_NOISE_LOOP 0      /* noise [3,H,W], mod = idx [length], p0 = sigma / 50                    (:42-53) */
#define SSAR_NOISE_BLEND 1     /* noise [2,M,H,W], mod [length, M]                                      (:11-24) */
#define SSAR_NOISE_MULTIPLY 2  /* noise [M,H,W], mod [length, M]                                        (:27-39) */
#define SSAR_NOISE_BUFFER 3    /* noise = precomputed [n_frames, H*W] for this batch                             */
#define SSAR_NOISE_AVERAGE 4   /* acc = (acc + cur) / 2                                                 (:56-63) */
#define SSAR_NOISE_MODULATE 5  /* mod [length]: acc = acc * mod + cur * (1 - mod)                       (:66-75) */
#define SSAR_NOISE_SCALEBIAS 6 /* acc = p0 * acc + p1                                                   (:78-86) */
typedef struct {
    int kind;
    int M;
    int to_acc;
    int reserved;
    const float* noise;
    const float* mod;
    float p0, p1;
} ssar_noise_op;

/* Evaluates a program for frames [first_frame, first_frame + n_frames) -> out [n_frames, height * width].
 * ops_device / ops_host: the same n_ops ops in device and host memory (the host copy is only validated);
 * loop_scratch: (number of LOOP ops) * n_frames floats of device scratch (may be NULL without LOOP ops). */
int ssar_noise_eval(const ssar_noise_op* ops_device, const ssar_noise_op* ops_host, int n_ops, int first_frame,
                    int n_frames, int height, int width, float* loop_scratch, float* out, void* stream);

/* ---- generator draws, batched (csrc/philox_draws.cu) -------------------------------------------------------------
 * The reference draws with torch calls on a seeded torch.Generator; on a CUDA generator these entry points produce the
 * values the per-call sequence produces (same Philox4x32-10 seed / subsequence / offset per element).              */

/* Generator-offset increment of ONE torch.randn(numel) call on the current device (ATen nullary kernel geometry). */
long long ssar_philox_normal_increment(long long numel);

/* torch.randn(numel[s], generator) for s = 0..n_segments-1 issued back to back from generator offset `first_offset`
 * (ssar/random/patch.py:123-131 base noise, ssar/random/noise.py:15,31,46 module bases).  dst / numel: HOST arrays (dst
 * holds device pointers).  *end_offset (host, optional) = generator offset after the last call.                     */
int ssar_philox_normal_segments(float* const* dst, const long long* numel, int n_segments, long long seed,
                                long long first_offset, long long* end_offset, void* stream);

/* Draw ops of a sampler program (one torch call each on <= 256 elements). */
#define SSAR_DRAW_RANDN 0       /* torch.randn(())                        -> 1 float                                 */
#define SSAR_DRAW_EXPONENTIAL 1 /* empty(n).exponential_(1) (multinomial) -> n floats                                */
/* Patch.__init__'s config sampler (ssar/random/patch.py:41-66: randint / rand / random_choice / skewnorm draws) for
 * n_seeds generators seeded seeds[i], one warp per seed.  program (device): int[2 * n_ops] = kinds then sizes of the
 * head ops (after the randint + rand of :41-42), of one latent config (:77-91) and of one noise config (:93-106).
 * out [n_seeds, capacity] floats: num_base_latents, rand() for sigma, head draws, latent configs, noise configs;
 * counts [n_seeds, 3] = (n latent configs, n noise configs, floats produced); end_offset [n_seeds] generator offsets. */
int ssar_patch_sample(const long long* seeds, int n_seeds, const int* program, int n_head_ops, int n_latent_ops,
                      int n_noise_ops, int min_subpatches, int max_subpatches, int capacity, float* out, int* counts,
                      long long* end_offset, void* stream);

/* `count` consecutive torch.randperm(n, generator) calls (ssar/random/patch.py:113, ssar/random/latent.py:37: one per
 * latent sub-patch) from generator offset first_offset, call k at first_offset + k * offset_stride.  bits = ATen's key
 * width for n (ceil(log2(n - (6 n^2 + 1) / (12 ln 0.9)))); n <= 1024.  out: int64 [count, n] (device).            */
int ssar_philox_randperm(long long seed, long long first_offset, long long offset_stride, int n, int bits, int count,
                         long long* out, void* stream);

/* One latent sub-patch of a stack (the arguments of ssar_subpatch_merge, plus src_mode 2: segmentation labels).     */
typedef struct {
    const long long* selection; /* device int64 [n_select]: palette columns (a torch.randperm prefix)               */
    const void* src;            /* src_mode 0 / 1: float [T, src_ld];  src_mode 2: int64 labels [T]                  */
    const float* mod;           /* modulation feature [T, mod_ld] (column 0) or NULL                                 */
    const float* pre_taps;      /* src_mode 2: Gaussian taps [2 * pre_radius + 1] of the one-hot smoothing (sigma 5) */
    long long src_ld, mod_ld;
    float src_scale, mod_scale, alpha, beta;
    int g_begin, g_end, n_select, src_mode, radius, pre_radius;
} ssar_subpatch_desc;

/* A whole stack of latent sub-patches (ssar/random/patch.py:113-121 -> latent.py:35-78 per sub-patch) applied to the
 * weights [n_rows, n_groups, n_basis] in ONE launch: a CTA keeps its 8 frames of weights in shared memory and applies the
 * sub-patches in order (each exactly as ssar_subpatch_merge would; src_mode 2 = ssar_onehot_gauss fused in).  descs: HOST
 * array of n_descs <= 24 descriptors.  taps: Gaussian taps of the sigma = 1 smoothing for descriptors with radius > 0
 * (all share `radius`).  load_initial = 0 when descs[0] overwrites every group (the base loop latents).             */
int ssar_latent_stack(float* weights, int n_rows, int n_groups, int n_basis, const ssar_subpatch_desc* descs, int n_descs,
                      const float* taps, int load_initial, void* stream);

/* maua/maua/audiovisual/audioreactive/latent.py:55-81 (slerp, slerp_loops): spherical interpolation between the
 * consecutive unit-normalised knots cat([y] * n_loops + [y[0]]) on linspace(0, 1, n_t), rows in the reference's
 * (t, segment) order, linearly resampled along time to `size` (F.interpolate, align_corners=False), written straight
 * into out [size, n_layers, latent_dim].  y: [n_latents, n_layers, latent_dim]; n_t = round(size / (n_latents * n_loops
 * + 1)) (computed by the caller like the reference); scratch: ssar_slerp_loops_scratch_floats(...) floats.          */
long long ssar_slerp_loops_scratch_floats(int n_latents, int n_loops, int n_layers, int latent_dim);
int ssar_slerp_loops(const float* y, int n_latents, int n_layers, int latent_dim, int n_loops, int n_t, int size,
                     float* scratch, float* out, void* stream);

/* processing.py:52-55 normalize for per-clip arrays of <= 262 144 elements: min, max and the rescale in ONE launch (one
 * CTA per clip) — same result as ssar_minmax_keys + ssar_normalize.                                                  */
/* clip_rows: NULL, or int [n_clips] (device) for a RAGGED batch: the first clip_rows[c] * row_elems elements of clip c are
 * real (min / max over them only); the rest of the row is written as zeros. */
int ssar_normalize_fused(const float* x, int n_clips, long long n_per_clip, float eps, float* out,
                         const int* clip_rows, long long row_elems, void* stream);

/* ssar_biquad that also leaves the sums of squares of its (clamped) output per 1024-sample block in
 * block_ss [n_clips, ceil(n_samples / 1024) + 2] (the last two entries of a row are y[0] and y[1024]); y may be NULL when
 * only the framed RMS of the filtered signal is wanted (supervised/data.py:75-86: rms(low_pass(..)), ...).
 * ssar_rms_from_block_ss turns those sums into rms(y, 2048, 1024) [n_clips, n_frames] (features/audio.py:31-37):
 * frame t >= 1 = blocks t - 1 and t, frame 0 = the reflected edge.                                               */
int ssar_biquad_ss(const float* x, long long n_samples, long long stride_x, int n_clips, float b0, float b1, float b2,
                   float a1, float a2, int clamp, float* y, long long stride_y, float* block_ss, void* scratch,
                   void* stream);
int ssar_rms_from_block_ss(const float* block_ss, long long n_blocks, int n_clips, int n_frames, float* out, void* stream);

/* ssar_biquad WITHOUT caller scratch and without any allocation, for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the reference.
 * The scan scratch is a file-scope device array of SSAR_BIQUAD_STATIC_SCRATCH_BYTES inside the library
 * (every GPU that loads the library holds it until exit); sizes whose ssar_biquad_scratch_bytes(n_clips, n_samples) exceeds
 * it are rejected with SSAR_ERR_ARG (the scan keeps 96 bytes per 8192-sample span: 2 MiB = 21 845 spans = 30 clips x 240 s at
 * 24576 Hz).  Same kernel and results as ssar_biquad.  Every call
 * on a GPU shares the array: calls that overlap on the GPU (different streams or threads) race; serialise them. */
#define SSAR_BIQUAD_STATIC_SCRATCH_BYTES (2LL << 20)
int ssar_biquad_static(const float* x, long long n_samples, long long stride_x, int n_clips, float b0, float b1, float b2,
                       float a1, float a2, int clamp, float* y, long long stride_y, void* stream);

/* TWO biquads over the SAME input in one pass (features/processing.py:141-150: low_pass(x) and high_pass(x), which
 * also feeds mid_pass): the input is read once instead of once per filter.  coeffs: HOST array of 10
 * floats (b0, b1, b2, a1, a2 of filter 0, then of filter 1).  y0 / y1: output signals or NULL; block_ss0 / block_ss1: as in
 * ssar_biquad_ss or NULL; every filter needs at least one of the two.  Same arithmetic per filter as ssar_biquad. */
int ssar_biquad_pair(const float* x, long long n_samples, long long stride_x, int n_clips, const float* coeffs, int clamp,
                     float* y0, float* y1, long long stride_y, float* block_ss0, float* block_ss1, void* scratch,
                     void* stream);

/* low_pass(x), high_pass(x) AND mid_pass(x) = low_pass(high_pass(x, 4000), 200) (features/processing.py:141-150) from ONE
 * read of x: filters 0 and 1 run over x, filter 2 over filter 1's output (clamped when clamp != 0, as lfilter(clamp=True)
 * returns it) without that signal leaving shared memory.  coeffs: HOST array of 15 floats (b0, b1, b2, a1, a2 per filter).
 * y0..y2 / block_ss0..2: as in ssar_biquad_pair; filter 1 may have no output of its own.  Per filter the arithmetic inside a
 * 32-sample chunk is ssar_biquad's; filter 2 sees filter 1's true entering state as its input history at chunk starts. */
int ssar_biquad_chain(const float* x, long long n_samples, long long stride_x, int n_clips, const float* coeffs, int clamp,
                      float* y0, float* y1, float* y2, long long stride_y, float* block_ss0, float* block_ss1,
                      float* block_ss2, void* scratch, void* stream);

/* rosa/constantq.py:77-100, :189-191 (__cqt_response for every octave) as ONE tensor-core launch: the response of octave
 * o is the K = 1024 contraction of the rectangular 1024-sample frames of signals[o] (hop >> o, centre reflect padding)
 * with the time-domain form of the tuned basis, b[r, n] = sum_k fft_basis[r, k] exp(-2 pi i k n / 1024):
 *   out[clip, t, n_bins - 36 (o + 1) + r] = sqrt(2^o) * row_scale[slot, .] * sum_n y_o[t (hop >> o) - 512 + n] b[slot][r, n]
 * signals / n_samples: HOST arrays [n_octaves] of dense device signals [n_clips, n_samples[o]] (o > 0: the kaiser / 2
 * decimations scaled by sqrt(2), constantq.py:83-86).  basis_hi / basis_lo: fp16 [n_slots * 80, 1024] (rows 2r / 2r + 1 of
 * a slot = Re / Im of basis_scale * b[r, :], rows 72..79 zero), fp32 = hi + lo.  clip_slot: int32 [n_clips] (device; the
 * tuning index of each clip) or NULL.  out: [n_clips, n_frames, 36 * n_octaves, 2].  tcgen05.mma kind::f16 on 3-term
 * fp16 splits of both operands with fp32 accumulation; TMA tensor maps with overlapping rows read the frames straight
 * from the padded signals.  scratch: ssar_cqt_tc_scratch_bytes(...) bytes.  clip_samples / clip_frames: NULL, or device
 * arrays [n_clips] for a RAGGED batch (octave-0 samples of each clip — octave o holds ceil(.. / 2^o) — and its frame count;
 * reflection at the clip's own end, rows past its frame count written as zeros).                                     */
long long ssar_cqt_tc_scratch_bytes(int n_clips, const long long* n_samples, int n_octaves);
int ssar_cqt_tc(const float* const* signals, const long long* n_samples, int n_octaves, int n_clips, int hop, int n_frames,
                const void* basis_hi, const void* basis_lo, int n_slots, const int* clip_slot, const float* row_scale,
                long long row_scale_ld, float basis_scale, void* scratch, float* out, const long long* clip_samples,
                const int* clip_frames, void* stream);

/* ssar/random/latent.py:7-13 spline_loop_latents as weights: out[t, k] with loop(y)[t] = sum_k out[t, k] y[k] for the closed
 * natural cubic spline through (y_0 .. y_{K-1}, y_0) on linspace(0, 1, K + 1), evaluated at linspace(0, end, size) % 1.
 * coef: fp64 [K, 4, K + 1] cardinal-spline coefficients (a, b, two_c, three_d) per segment; knots: fp32 [K + 1];
 * n_knots_closed = K + 1 >= 3; out: fp32 [size, K] (the closing knot's weight folded onto knot 0). */
int ssar_spline_loop_weights(const double* coef, const float* knots, int n_knots_closed, int size, float end, float* out,
                             void* stream);

/* 16-bit PCM -> fp32: out[i] = pcm[i] / 32768 (exact; torchaudio.load's normalisation of the PCM files the reference reads,
 * supervised/data.py:127, :180).  A host can ship audio in its on-disk width: half the host-to-device bytes. */
int ssar_pcm16_to_float(const short* pcm, long long n, float* out, void* stream);

/* ssar/supervised/data.py:255-258 overlapping_slices (the dataset's training windows, also :119-132): window k = rows
 * [offsets[k], offsets[k] + window) of x [n_rows, inner] -> out [n_windows, window, inner].  offsets: device int64, every
 * offsets[k] + window <= n_rows (the caller lists the reference's windows: starts 0, L/4, L/2, 3L/4, splits of L, the last
 * split of every start dropped). */
int ssar_window_slices(const float* x, long long n_rows, long long inner, const long long* offsets, int n_windows, int window,
                       float* out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Video-side feature twins (ssar/features/video.py), video: [n_frames, n_channels, height, width] fp32
 * ---------------------------------------------------------------------------------------------- */

/* video.py:11-31 redogram / greenogram / blueogram: per frame torch.histc(video[t, channel], bins) over the frame's own
 * [min, max] (a constant frame: [v - 1, v + 1]), divided by the largest count.  Written to
 * out[t * out_ld + out_col0 + b] so rgb_hist's three histograms land side by side. */
int ssar_video_hist(const float* video, int n_frames, int n_channels, int height, int width, int channel, int bins,
                    float* out, long long out_ld, int out_col0, void* stream);
/* video.py:58-59 visual_variance: (std over (C, H, W), unbiased) ** 2 per frame -> out [n_frames] */
int ssar_video_variance(const float* video, int n_frames, long long n_per_frame, float* out, void* stream);
/* video.py:62-73 absdiff: out[t] = sum |video[t + 1] - video[t]| for t < n_frames - 1, out[n_frames - 1] = out[n_frames - 2] */
int ssar_video_absdiff(const float* video, int n_frames, long long n_per_frame, float* out, void* stream);

/* maua/maua/GAN/wrappers/stylegan2.py:193-210 make_noise_pyramid (normalize = 1) and the synthesizer's own resize of a
 * mismatched noise input, :86-96 (normalize = 0): noise [n_frames, in_h, in_w] -> outs[l] [n_frames, heights[l], widths[l]]
 * for l < n_layers <= 16, bicubic (torch.nn.functional.interpolate, mode="bicubic", align_corners=False); with
 * normalize each frame of each layer is divided by its own unbiased standard deviation (noises.std((1, 2, 3))).
 * heights / widths / outs: HOST arrays.  One launch for all layers.                                                 */
int ssar_noise_pyramid(const float* noise, int n_frames, int in_h, int in_w, int n_layers, const int* heights,
                       const int* widths, float* const* outs, int normalize, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SSAR_B200_H */
