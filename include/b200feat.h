/* libb200feat -- C ABI of the B200 (sm_100a) waveform -> log-mel / mel / fbank featurizers.
 *
 * This is the drop-in boundary of the hot path.  The reference
 * (RobMills28/Expressive-Speech-Translation) is pure Python and reaches the path through three
 * third-party Python calls; each entry point below is what a ctypes binding of that call would bind
 * (INTEGRATION.md shows the stubs):
 *
 *   b200feat_whisper_logmel  <- whisper.audio.log_mel_spectrogram(audio, n_mels, padding, device)
 *        called inside asr_model.transcribe(...) at
 *        /root/reference/Backend/services/cascaded_backend.py:158,
 *        /root/reference/Backend/speech_translation_pipeline.py:140,
 *        /root/reference/Backend/gtts_pipeline_fixed.py:64 and (n_mels = 128) by the CosyVoice frontend
 *        behind /root/reference/Docker/cosyvoice_api.py:128.
 *   b200feat_cosy_mel        <- matcha.utils.audio.mel_spectrogram(y, 1920, 80, 24000, 480, 1920, 0, 8000,
 *        center=False), bound at /root/reference/greek_sft.yaml:55-63 (feat_extractor) and used by
 *        frontend.zero_shot_frontend behind /root/reference/Docker/cosyvoice_api.py:128.
 *   b200feat_kaldi_fbank     <- torchaudio.compliance.kaldi.fbank(waveform, num_mel_bins=80, dither=0,
 *        sample_frequency=16000) [site-packages/torchaudio/compliance/kaldi.py:514-645] followed by the
 *        CosyVoice frontend's feat - feat.mean(0) (speaker embedding input, model named at
 *        /root/reference/Docker/debug_trigger_test_cosyvoice.py:102,258).
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ or torch types cross this boundary.
 *   - `wave`, `out`, `lengths`, `workspace` are DEVICE pointers owned by the caller (the *_host variants
 *     take HOST pointers, ideally pinned, and stage through a library-owned device arena).
 *   - `lengths` may be NULL (every clip has L samples); otherwise int64[B] on the device, values are
 *     clamped to [0, L].  Frames past a clip's own frame count are written as 0.
 *   - launches are asynchronous on `cuda_stream` (a cudaStream_t / CUstream; NULL = legacy default
 *     stream); the library never synchronises the device in the device-pointer entry points.
 *   - every function returns 0 on success and a negative B200FEAT_E* code otherwise;
 *     b200feat_last_error() returns a thread-local description of the last failure.
 *   - reentrant and thread-safe.  Device-pointer entry points keep no mutable state of their own (scratch = the
 *     caller's workspace; constant tables live in the module image) -- the one exception is
 *     b200feat_whisper_logmel_noalloc, whose static scratch array is shared by all calls on a GPU (see there).
 *     The *_host entry points lease a staging arena (device buffer + streams) per call from a per-device pool, so
 *     concurrent calls overlap; pageable host buffers are accepted (the driver stages them: measured penalty in
 *     DESIGN.md), pinned ones run at PCIe speed.
 */
#ifndef B200FEAT_H
#define B200FEAT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200FEAT_VERSION 101 /* major*10000 + minor*100 + patch */

enum {
    B200FEAT_OK = 0,
    B200FEAT_EINVAL = -1,     /* bad argument (message says which) */
    B200FEAT_ECUDA = -2,      /* a CUDA runtime call failed */
    B200FEAT_EWORKSPACE = -3, /* workspace too small */
    B200FEAT_EARCH = -4       /* device is not sm_100 */
};

enum { B200FEAT_KIND_WHISPER = 0, B200FEAT_KIND_COSY = 1, B200FEAT_KIND_KALDI = 2 };
enum { B200FEAT_NORM_PER_CLIP = 0, B200FEAT_NORM_PER_CALL = 1 };

int b200feat_version(void);
const char* b200feat_last_error(void);

/* frames produced for a clip of n_samples (padding only applies to Whisper) */
int64_t b200feat_num_frames(int32_t kind, int64_t n_samples, int32_t padding);

/* bytes of device workspace the call needs (>= 16, never 0) */
int64_t b200feat_workspace_bytes(int32_t kind, int32_t B, int64_t L);

/* Whisper log-mel.  out: (B, n_mels, T_out) fp32, out_batch_stride floats between clips,
 * T_out >= num_frames(L, padding) is the row stride.  n_mels in {80, 128}.
 * norm_mode: PER_CLIP = max over each clip (HF WhisperFeatureExtractor / one transcribe() call per
 * clip); PER_CALL = max over the whole call (whisper.audio semantics for a batched tensor). */
int b200feat_whisper_logmel(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                            int64_t wave_stride, int32_t n_mels, int32_t padding, int32_t norm_mode,
                            float* out, int64_t out_batch_stride, int64_t T_out,
                            void* workspace, int64_t workspace_bytes, void* cuda_stream);

/* Extended Whisper entry point (the bytes either side of the path).
 *   wave_dtype B200FEAT_S16: `wave` is int16 PCM as ffmpeg hands it to whisper.audio.load_audio; the loader scales by
 *     1/32768 exactly like load_audio's `.astype(np.float32) / 32768.0` (UPSTREAM openai-whisper audio.py, reached from
 *     /root/reference/Backend/services/cascaded_backend.py:158) -- half the input bytes.
 *   out_dtype B200FEAT_F16: `out` is (B, n_mels, T_out) IEEE binary16; every value is the float32 result rounded to
 *     nearest even, i.e. bit-identical to `mel.to(torch.float16)`, the cast whisper.transcribe applies to each segment
 *     with fp16=True (how cascaded_backend.py:158 calls it) -- half the output bytes.
 * Strides are in elements of the respective dtype.  B200FEAT_F32 / B200FEAT_F32 is b200feat_whisper_logmel. */
#define B200FEAT_F32 0
#define B200FEAT_F16 1
#define B200FEAT_S16 2
int b200feat_whisper_logmel_ex(const void* wave, int32_t wave_dtype, const int64_t* lengths, int32_t B, int64_t L,
                               int64_t wave_stride, int32_t n_mels, int32_t padding, int32_t norm_mode,
                               void* out, int32_t out_dtype, int64_t out_batch_stride, int64_t T_out,
                               void* workspace, int64_t workspace_bytes, void* cuda_stream);

/* Whisper log-mel as encoder-ready WINDOWS.  Replaces, for a batch of clips, what
 * whisper.transcribe does with the mel of a clip (UPSTREAM openai-whisper transcribe.py, reached from
 * /root/reference/Backend/services/cascaded_backend.py:158 `self.asr_model.transcribe(...)`):
 *     mel = log_mel_spectrogram(audio, n_mels, padding=N_SAMPLES)          # (n_mels, content_frames + 3000)
 *     mel_segment = mel[:, seek : seek + N_FRAMES]; pad_or_trim(mel_segment, N_FRAMES).to(dtype)    # seek = 0, 3000, ...
 * out: (B, n_windows, n_mels, win_frames), float32 or binary16 (out_dtype), out_batch_stride elements between clips:
 * window w of clip b = frames [w * win_frames, (w + 1) * win_frames) of the clip's padded log-mel, written by the
 * kernel's epilogue straight into that layout -- no (n_mels, T) intermediate, no slicing copy, no cast pass.
 *   - The clamp max - 8 uses the maximum over the clip's WHOLE padded mel, as transcribe's call does, including the
 *     frames behind the last window (they are transformed, not stored).
 *   - The windows must be complete slices: n_windows * win_frames <= (L + padding) / 160, and they must cover the
 *     content: n_windows * win_frames >= L / 160.  With transcribe's padding = 480000 and win_frames = 3000:
 *     n_windows = ceil((L / 160) / 3000).
 *   - Ragged clips (`lengths`): windows behind a clip's own content hold what transcribe's padding gives (the clip's
 *     clamp threshold) up to frame (len + padding) / 160 and 0 behind it.
 *   - wave: float32, 16-byte aligned rows (pointer and wave_stride * 4 multiples of 16); workspace:
 *     b200feat_workspace_bytes(B200FEAT_KIND_WHISPER, B, L + padding + 160 * 128 * (n_windows + 1)) bytes. */
int b200feat_whisper_logmel_windows(const float* wave, const int64_t* lengths, int32_t B, int64_t L, int64_t wave_stride,
                                    int32_t n_mels, int32_t padding, int32_t norm_mode, int32_t win_frames, int32_t n_windows,
                                    void* out, int32_t out_dtype, int64_t out_batch_stride,
                                    void* workspace, int64_t workspace_bytes, void* cuda_stream);

/* Whisper log-mel WITHOUT any allocation and without a caller workspace.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arithmetic, kernels and result as b200feat_whisper_logmel (float32 in / out);
 * the scratch the kernels need (per-clip maximum keys, per-unit minimum / maximum) lives in ONE library-owned static
 * device array of b200feat_noalloc_workspace_max() = 1 MiB per GPU (held by every GPU that loads the library until exit).
 *   limit    a call needs 32 + 4 * B + 8 * B * ceil(max(T_out, (L + padding) / 160 + 1) / 64) bytes (e.g. 1024 clips of
 *            30 s: 380 KiB; 2048: 760 KiB); larger calls are rejected with B200FEAT_EWORKSPACE and a message -- split
 *            the batch, or use b200feat_whisper_logmel with your own workspace.
 *   sharing  every call on a device uses the same array.  The part a call uses is cleared on cuda_stream before its
 *            kernels run, so calls that are ordered on the GPU (same stream, or ordered by events) are independent of
 *            each other; calls that OVERLAP on the GPU (different streams, no ordering) race on it and give wrong
 *            results -- serialise them or use the workspace entry point.
 *   The entry point calls no cudaMalloc / cudaFree / allocating library routine; its launches are asynchronous. */
int64_t b200feat_noalloc_workspace_max(void);
int b200feat_whisper_logmel_noalloc(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                                    int64_t wave_stride, int32_t n_mels, int32_t padding, int32_t norm_mode,
                                    float* out, int64_t out_batch_stride, int64_t T_out, void* cuda_stream);

/* CosyVoice 24 kHz mel (n_fft 1920, hop 480, 80 mels, ln, no normalisation).
 * out: (B, 80, T_out) fp32. */
int b200feat_cosy_mel(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                      int64_t wave_stride, float* out, int64_t out_batch_stride, int64_t T_out,
                      void* workspace, int64_t workspace_bytes, void* cuda_stream);

/* The same mel written frame-major, out: (B, T_out, 80): the layout the CosyVoice frontend builds with
 * `feat_extractor(speech).squeeze(0).transpose(0, 1)` before it hands the prompt features to the flow model (UPSTREAM
 * cosyvoice/cli/frontend.py `_extract_speech_feat`, reached from /root/reference/Docker/cosyvoice_api.py:128;
 * SURVEY.md section 8(f) item 2).  Rows past a ragged clip's own frame count are zero. */
int b200feat_cosy_mel_frames(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                             int64_t wave_stride, float* out, int64_t out_batch_stride, int64_t T_out,
                             void* workspace, int64_t workspace_bytes, void* cuda_stream);

/* kaldi fbank (80 bins, 25 ms / 10 ms, povey, dither 0, snip_edges), optional per-clip mean
 * subtraction over frames.  out: (B, T_out, 80) fp32 frame-major. */
int b200feat_kaldi_fbank(const float* wave, const int64_t* lengths, int32_t B, int64_t L,
                         int64_t wave_stride, int32_t subtract_mean, float* out,
                         int64_t out_batch_stride, int64_t T_out,
                         void* workspace, int64_t workspace_bytes, void* cuda_stream);

/* Waveform conditioning in front of the featurizers: interleaved PCM -> mono mix ->
 * band-limited polyphase resampling, fused in one kernel (the PCM is read once, nothing intermediate touches HBM).
 * Replaces
 *   torchaudio.transforms.Resample(orig_freq, 16000, lowpass_filter_width=128, rolloff=0.95,
 *       resampling_method='kaiser_window', beta=14.769656459379492)(audio) followed by the stereo -> mono mix at
 *       /root/reference/Backend/services/audio_processor.py:295-319 (mean of the channels, or mid + 0.3 * side =
 *       0.65 L + 0.35 R: both are `mix` weights; resampling and mixing are linear, so their order does not matter),
 *   torchaudio.functional.resample(wave, 44100, 16000) of SURVEY.md's cfg1 recipe (Backend/harvard.wav), and the
 *   CosyVoice frontend's Resample(16000 -> 24000) (UPSTREAM, reached from /root/reference/Docker/cosyvoice_api.py:128).
 * Arithmetic: site-packages/torchaudio/functional/functional.py:1305-1432 (`_get_sinc_resample_kernel`, taps computed
 * in float64 and stored as float32 like transforms.Resample does; `_apply_sinc_resample_kernel`), float32 accumulation,
 * only the taps inside the window's support are evaluated.
 *   pcm        device pointer, (B, L_in, channels) interleaved; pcm_dtype B200FEAT_F32 or B200FEAT_S16 (int16 is scaled by
 *              1/32768 like whisper.audio.load_audio); pcm_stride = elements between clips (>= L_in * channels)
 *   channels   1 or 2;  mix_host: HOST pointer to `channels` weights, NULL = mean of the channels
 *   lengths    device int64[B] input frames per clip or NULL; a clip yields ceil(new * len / orig) samples, the rest
 *              of its row up to L_out is written as 0
 *   method     B200FEAT_RESAMPLE_HANN ("sinc_interp_hann") or B200FEAT_RESAMPLE_KAISER ("sinc_interp_kaiser", beta)
 *   out        (B, out_stride) float32, L_out <= b200feat_resample_out_len(L_in, orig, new) samples per row
 * The tap table of a (device, ratio, filter) is built and uploaded on its first use (that one call copies
 * synchronously) and owned by the library; every other call is asynchronous on cuda_stream. */
enum { B200FEAT_RESAMPLE_HANN = 0, B200FEAT_RESAMPLE_KAISER = 1 };
int64_t b200feat_resample_out_len(int64_t n_frames, int32_t orig_freq, int32_t new_freq);
int b200feat_resample(const void* pcm, int32_t pcm_dtype, int32_t channels, const float* mix_host,
                      const int64_t* lengths, int32_t B, int64_t L_in, int64_t pcm_stride,
                      int32_t orig_freq, int32_t new_freq, int32_t lowpass_filter_width, double rolloff,
                      int32_t method, double beta, float* out, int64_t out_stride, int64_t L_out, void* cuda_stream);

/* Wav2Lip / Diff2Lip mel.  Replaces, for a batch of clips,
 *     /root/reference/Backend/diff2lip/audio/audio.py:45-51  melspectrogram(wav)
 * with the parameters of /root/reference/Backend/diff2lip/audio/hparams.py:22-65, as called at
 * /root/reference/Backend/diff2lip/generate.py:53 and generate_dist.py:49 (`audio.melspectrogram(wav).T`):
 *   pre-emphasis y[n] = x[n] - 0.97 x[n-1] (audio.py:19-22) | librosa.stft(n_fft 800, hop 200, win 800, periodic Hann,
 *   center=True) | magnitude | 80 Slaney filters 55..7600 Hz (librosa.filters.mel, audio.py:99-103) |
 *   20 log10(max(1e-5, .)) - 20 (audio.py:105-107, :47) | clip(8 (S + 100) / 100 - 4, -4, 4) (audio.py:112-118).
 * wave: (B, wave_stride) float32 at 16 kHz; out: (B, 80, T_out) float32 -- melspectrogram's layout -- or, frame_major != 0,
 * (B, T_out, 80), the `.T` the callers take; T_out >= b200feat_wav2lip_num_frames(L) = 1 + L / 200.  Ragged clips
 * (`lengths`): frames past 1 + len / 200 are 0.
 * pad_mode: the padding librosa.stft applies after the pre-emphasis -- B200FEAT_PAD_REFLECT is librosa 0.9.2's default, the
 * version the reference pins (/root/reference/Backend/requirements.txt:14; needs clips of more than 400 samples);
 * B200FEAT_PAD_ZEROS is the default of librosa >= 0.10.
 * The reference computes this mel in float64 (lfilter promotes); this is float32, within 1e-4 of it in the [-4, 4] scale. */
enum { B200FEAT_PAD_REFLECT = 0, B200FEAT_PAD_ZEROS = 1 };
int64_t b200feat_wav2lip_num_frames(int64_t n_samples);
int b200feat_wav2lip_mel(const float* wave, const int64_t* lengths, int32_t B, int64_t L, int64_t wave_stride,
                         int32_t pad_mode, int32_t frame_major, float* out, int64_t out_batch_stride, int64_t T_out,
                         void* cuda_stream);

/* The reference's stereo -> mono rule on the device.  Replaces, for a batch of stereo clips,
 * /root/reference/Backend/services/audio_processor.py:294-319: Resample both channels, then
 *     correlation = sum(r0 * r1) / sqrt(sum(r0 ** 2) * sum(r1 ** 2))
 *     mono = mean(r0, r1) if correlation > 0.5 else mid + 0.3 * side        (mid = (r0 + r1) / 2, side = (r0 - r1) / 2)
 * per clip, with the correlation taken over the RESAMPLED channels as the reference does.  Two passes of the resampler (one
 * per channel; `out` receives r0, `scratch` r1) and one CTA per clip for the sums (float64 accumulators, fixed order) and the
 * mix, which uses the reference's float32 operations in its order.
 *   pcm      device pointer, (B, L_in, 2) interleaved, B200FEAT_F32 or B200FEAT_S16; lengths as in b200feat_resample
 *   scratch  device, >= B * out_stride floats;  decisions (int32[B], 1 = mean, 0 = mid/side) and correlations (float[B])
 *            are optional outputs (NULL to skip).
 * The reference's filter is lowpass_filter_width 128, rolloff 0.95, B200FEAT_RESAMPLE_KAISER, beta 14.769656459379492. */
int b200feat_condition_stereo_auto(const void* pcm, int32_t pcm_dtype, const int64_t* lengths, int32_t B, int64_t L_in,
                                   int64_t pcm_stride, int32_t orig_freq, int32_t new_freq, int32_t lowpass_filter_width,
                                   double rolloff, int32_t method, double beta, float* out, int64_t out_stride, int64_t L_out,
                                   float* scratch, int64_t scratch_floats, int32_t* decisions, float* correlations,
                                   void* cuda_stream);

/* Spectral-gate denoiser.  Replaces, for a batch of mono clips,
 * /root/reference/Backend/services/audio_processor.py:322-358: torch.stft(1024, 256, Hann) | noise floor = 10 % quantile of
 * each bin's magnitudes over time | gain = clamp((mag - 1.5 floor) / (mag + 1e-8), 0, 1) | x 1.2 on the range
 * [int(T * 0.0625), int(T * 0.25)) of `mag.shape[1]` -- which is the FRAME axis in the reference's (513, T) tensor, and is
 * reproduced as such | torch.istft | audio / max|audio| * 0.95.
 *   wave (B, wave_stride) float32, L > 512 samples per row; lengths device int64[B] or NULL (a clip of <= 512 samples yields
 *   zeros); out (B, out_stride) float32, b200feat_spectral_gate_out_len(L) = 256 * (L / 256) samples per row, a ragged clip's
 *   row is zero behind its own 256 * (len / 256); workspace: b200feat_spectral_gate_workspace_bytes(B, L) (the complex
 *   spectrum, 513 * (1 + L / 256) * 8 bytes per clip, the noise floors and the peaks).  Float32 throughout, like the reference. */
int64_t b200feat_spectral_gate_out_len(int64_t n_samples);
int64_t b200feat_spectral_gate_workspace_bytes(int32_t B, int64_t L);
int b200feat_spectral_gate(const float* wave, const int64_t* lengths, int32_t B, int64_t L, int64_t wave_stride,
                           float* out, int64_t out_stride, void* workspace, int64_t workspace_bytes, void* cuda_stream);

/* Host-buffer twin of b200feat_resample: pcm_host (B, L_in, channels) densely packed, out_host (B, L_out) float32 with
 * L_out = b200feat_resample_out_len(L_in, orig_freq, new_freq); chunked H2D / kernel / D2H pipeline, blocking. */
int b200feat_resample_host(const void* pcm_host, int32_t pcm_dtype, int32_t channels, const float* mix_host,
                           const int64_t* lengths_host, int32_t B, int64_t L_in, int32_t orig_freq, int32_t new_freq,
                           int32_t lowpass_filter_width, double rolloff, int32_t method, double beta, float* out_host, int32_t device);

/* PCM in, Whisper log-mel out, ONE call ("int16 PCM goes in and features come out"):
 * b200feat_resample (mono mix + sinc resampler, arguments as there) followed by b200feat_whisper_logmel_ex on the same
 * stream; the float32 waveform between them lives in `workspace` (b200feat_pcm_whisper_workspace_bytes), never in a
 * caller tensor.  Replaces load_audio's int16 -> float cast + the resample + log_mel_spectrogram chain in front of
 * /root/reference/Backend/services/cascaded_backend.py:158 (SURVEY.md cfg1 recipe: Backend/harvard.wav, 44.1 kHz stereo).
 * lengths: input FRAMES per clip (device int64[B] or NULL).  out / out_dtype / T_out / norm_mode as in
 * b200feat_whisper_logmel_ex with L = b200feat_resample_out_len(L_in, orig_freq, new_freq). */
int64_t b200feat_pcm_whisper_workspace_bytes(int32_t B, int64_t L_in, int32_t orig_freq, int32_t new_freq, int32_t padding);
int b200feat_pcm_whisper_logmel(const void* pcm, int32_t pcm_dtype, int32_t channels, const float* mix_host,
                                const int64_t* lengths, int32_t B, int64_t L_in, int64_t pcm_stride,
                                int32_t orig_freq, int32_t new_freq, int32_t lowpass_filter_width, double rolloff,
                                int32_t method, double beta, int32_t n_mels, int32_t padding, int32_t norm_mode,
                                void* out, int32_t out_dtype, int64_t out_batch_stride, int64_t T_out,
                                void* workspace, int64_t workspace_bytes, void* cuda_stream);
/* the same through the chunked host pipeline (pinned or pageable host buffers, blocking): PCM crosses PCIe, not float32 */
int b200feat_pcm_whisper_logmel_host(const void* pcm_host, int32_t pcm_dtype, int32_t channels, const float* mix_host,
                                     const int64_t* lengths_host, int32_t B, int64_t L_in, int32_t orig_freq, int32_t new_freq,
                                     int32_t lowpass_filter_width, double rolloff, int32_t method, double beta,
                                     int32_t n_mels, int32_t padding, int32_t norm_mode, void* out_host, int32_t out_dtype,
                                     int32_t device);

/* Host-buffer variants (the end-to-end path): wave_host / out_host are HOST pointers (pinned memory
 * gives full PCIe bandwidth); clips are streamed through a library-owned device arena in chunks with
 * H2D copy, kernels and D2H copy overlapped on internal streams.  Blocks until out_host is complete.
 * lengths_host may be NULL.  Output layouts as above, densely packed (T_out = num_frames(L)). */
int b200feat_whisper_logmel_host(const float* wave_host, const int64_t* lengths_host, int32_t B, int64_t L,
                                 int32_t n_mels, int32_t padding, int32_t norm_mode, float* out_host,
                                 int32_t device);
int b200feat_whisper_logmel_ex_host(const void* wave_host, int32_t wave_dtype, const int64_t* lengths_host, int32_t B,
                                    int64_t L, int32_t n_mels, int32_t padding, int32_t norm_mode, void* out_host,
                                    int32_t out_dtype, int32_t device);
int b200feat_cosy_mel_host(const float* wave_host, const int64_t* lengths_host, int32_t B, int64_t L,
                           float* out_host, int32_t device);
int b200feat_kaldi_fbank_host(const float* wave_host, const int64_t* lengths_host, int32_t B, int64_t L,
                              int32_t subtract_mean, float* out_host, int32_t device);

/* frees the host-variant staging arenas of the calling process (optional; also done at unload) */
void b200feat_release(void);

/* number of kernel launches this library has issued in this process (bench.py's gpu_launches) */
int64_t b200feat_launch_count(void);

/* Optional per-launch timing for the roofline report: between profile_begin and profile_end every kernel
 * launch is bracketed by CUDA events on its launching stream.  profile_end synchronises those events and
 * returns the summed device time and launch count of the featurizer kernels and of the epilogue kernels
 * (Whisper normalisation / kaldi CMN).  Off by default; not meant for production calls. */
int b200feat_profile_begin(void);
int b200feat_profile_end(double* feature_ms, int64_t* feature_launches, double* epilogue_ms,
                         int64_t* epilogue_launches);

#ifdef __cplusplus
}
#endif
#endif /* B200FEAT_H */
