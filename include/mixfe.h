/* libmixfe -- B200-native (sm_100a) DSP front end for stem-aware mixing-style learning.
 *
 * C ABI of the drop-in boundary.  The reference (barry-mir/mixing-style-transfer) has no FFI
 * layer of its own; each entry point below names the reference Python interface it replaces
 * (paths relative to the reference checkout).  All pointers marked "device" are CUDA device
 * pointers on the plan's device; every call is stream-ordered and asynchronous on `cuda_stream`
 * (a cudaStream_t passed as void*), never synchronises the host in steady state and never
 * allocates once the plan's workspace has reached its high-water mark.  A plan serves ONE stream at a time: its
 * workspace, its parameter staging ring and the work counters of its persistent kernels are per plan, so calls on one
 * plan must not overlap on the GPU (use one plan per stream; plans are cheap).
 *
 * Return value: 0 on success, a negative MIXFE_E* code on failure; mixfe_last_error() returns a
 * thread-local human-readable message for the last failure on the calling thread.
 */
#ifndef MIXFE_H
#define MIXFE_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MIXFE_ABI_VERSION 2

#define MIXFE_OK 0
#define MIXFE_EINVAL (-1)       /* bad argument (shape, stride, NULL) */
#define MIXFE_EUNSUPPORTED (-2) /* configuration the kernels do not cover */
#define MIXFE_ECUDA (-3)        /* CUDA runtime error (message has the cudaError string) */
#define MIXFE_ESHORT (-4)       /* T <= n_fft/2: reflect padding impossible (torch.stft raises RuntimeError) */

#define MIXFE_NUM_STEMS 4   /* vocals, bass, drums, other -- src/model.py:54, src/mixing_utils.py:287 */

typedef struct mixfe_plan mixfe_plan;

/* Augmentation decisions for ONE clip, drawn on the host in the reference's order
 * (src/mixing_utils.py:372-419; RNG draw order in SURVEY.md section 3.5).  Index = stem in dict
 * order vocals, bass, drums, other. */
typedef struct mixfe_aug_params {
    float gain_lin[MIXFE_NUM_STEMS];     /* 10^(gain_db/20) as fp32, 1.0f when the gain gate did not fire (:389-392) */
    int tilt_on[MIXFE_NUM_STEMS];        /* spectral-tilt gate (:395) */
    double tilt_sos[MIXFE_NUM_STEMS][6]; /* one biquad b0 b1 b2 a0 a1 a2 = scipy butter(2, ...) (:426-429) */
    int comp_on[MIXFE_NUM_STEMS];        /* static compressor gate (:399) */
    float comp_threshold_db[MIXFE_NUM_STEMS]; /* apply_compression(threshold=-20) (:435); augment_stems uses -20 */
    float comp_ratio[MIXFE_NUM_STEMS];   /* apply_compression(ratio=4) (:435), > 0; augment_stems uses 4 */
    int bw_on[MIXFE_NUM_STEMS];          /* bandwidth-limit gate (:403) */
    double bw_sos[MIXFE_NUM_STEMS][12];  /* two biquads = scipy butter(4, cutoff, 'low') (:452) */
    int reverb_on;                       /* reverb gate (:409); impulse row must be valid when set */
    int _pad;
} mixfe_aug_params;

int mixfe_abi_version(void);
const char* mixfe_last_error(void);

/* Replaces MixingFeatureExtractor.__init__ (src/mixing_utils.py:19-51) and
 * MelSpectrogramPreprocessor.__init__ (src/model.py:25-39): builds the device-side constants
 * (FFT twiddles, 2-tap sparse mel table from an HTK filterbank with f_min=0, f_max=sr/2,
 * norm=None) for one (sample_rate, n_fft, hop, n_mels).  n_spectral_bins = 0 selects the 3-band
 * spectral features (D = 64), > 0 the detailed curve (D = 4*(n_bins+2+9)+8, :53-69). */
int mixfe_plan_create(mixfe_plan** out, int device, int sample_rate, int n_fft, int hop,
                      int n_mels, int n_spectral_bins);
int mixfe_plan_destroy(mixfe_plan* plan);

/* Replaces AudioAugmenter.__init__ (src/mixing_utils.py:367-370): a plan that serves only mixfe_augment,
 * mixfe_effects and mixfe_reverb (FFT twiddles of the reverb; no mel / STFT tables).  Every feature entry point
 * returns MIXFE_EINVAL on such a plan. */
int mixfe_plan_create_augment(mixfe_plan** out, int device, int sample_rate);

/* Optional: replace the plan's filterbank by a host matrix fb[n_fft/2+1][n_mels] (row-major),
 * e.g. the `mel_scale.fb` buffer of a reference checkpoint. */
int mixfe_plan_set_filterbank(mixfe_plan* plan, const float* fb_host);

int mixfe_feature_dim(const mixfe_plan* plan);          /* get_feature_dim(), src/mixing_utils.py:53-69 */
int mixfe_num_frames(const mixfe_plan* plan, int T);    /* 1 + T / hop (center=True framing of torch.stft) */
size_t mixfe_workspace_bytes(const mixfe_plan* plan);   /* current size of the plan-owned device workspace */

/* Replaces MixingFeatureExtractor.extract_all_features (src/mixing_utils.py:71-105, per clip) and
 * MelSpectrogramPreprocessor.forward (src/model.py:41-67, batched) in ONE pass over the audio.
 *   stems[s]        device, fp32; element (b, c, t) of stem s at stems[s][b*stride_b + c*stride_c + t]
 *   features        device [B][D] fp32 in the reference's sorted-key layout (:320-357), or NULL
 *   logmel          device, (b, ch, m, f) at logmel[b*logmel_stride_b + (ch*n_mels + m)*F + f],
 *                   ch = 2*stem + channel, F = mixfe_num_frames(T); natural log of (mel power + 1e-10); or NULL
 * Returns MIXFE_ESHORT when T <= n_fft/2. */
int mixfe_features_logmel(mixfe_plan* plan, const float* const stems[MIXFE_NUM_STEMS],
                          long long stride_b, long long stride_c, int B, int T,
                          float* features, float* logmel, long long logmel_stride_b,
                          void* cuda_stream);

/* Replaces AudioAugmenter.augment_stems (src/mixing_utils.py:372-419) for B clips with the
 * decisions of `params` (host array of B entries).  in/out use the same addressing as above;
 * out must not alias in.  impulse: device [B][L] fp32 reverb impulses (row b read only when
 * params[b].reverb_on), L = int(sample_rate*0.5) in the reference (:460). */
int mixfe_augment(mixfe_plan* plan, const float* const in[MIXFE_NUM_STEMS],
                  float* const out[MIXFE_NUM_STEMS], long long stride_b, long long stride_c,
                  int B, int T, const mixfe_aug_params* params, const float* impulse, int L,
                  void* cuda_stream);

/* Replaces the single-effect methods AudioAugmenter.apply_spectral_tilt / apply_compression(audio, threshold, ratio) /
 * apply_bandwidth_limit (src/mixing_utils.py:421-456) and the gain of :389-392 on ONE tensor: in/out device fp32,
 * element (b, c, t) at in[b*stride_b + c*stride_c + t], C channels per clip.  The decisions of params[b] with STEM
 * INDEX 0 (gain_lin[0], tilt_on[0], tilt_sos[0], comp_*[0], bw_*[0]) are applied to every channel of clip b;
 * reverb_on is ignored.  out must not alias in. */
int mixfe_effects(mixfe_plan* plan, const float* in, float* out, long long stride_b, long long stride_c,
                  int B, int C, int T, const mixfe_aug_params* params, void* cuda_stream);

/* Replaces AudioAugmenter.apply_reverb(audio, decay) (src/mixing_utils.py:458-479) on ONE tensor:
 * out = 0.7 in + 0.3 r,  r[c][n] = sum_k in[c][n + k - L/2] h_b[k]  (F.conv1d with padding L/2, cropped to T), without
 * the stem sum and the energy redistribution augment_stems adds (:409-417).  C = 1 or 2 channels per clip (wider
 * tensors: call per channel pair).  impulse: device fp32, clip b's L taps at impulse + b*impulse_stride_b;
 * impulse_stride_b = 0 shares one impulse (what the reference does for the channels of one call).
 * L = int(sample_rate * decay) in the reference (:460); L <= 2^19. */
int mixfe_reverb(mixfe_plan* plan, const float* in, float* out, long long stride_b, long long stride_c,
                 int B, int C, int T, const float* impulse, long long impulse_stride_b, int L, void* cuda_stream);

/* Replaces autograd through MelSpectrogramPreprocessor.forward (src/model.py:41-67), the path the style-transfer
 * loop differentiates (src/train_style_transfer.py:206-215): given d loss / d logmel, WRITES d loss / d stems into
 * grad_stems (every sample of the T-long rows is overwritten; no initialisation needed).  stems: the forward inputs,
 * addressed as in mixfe_features_logmel; logmel: the tensor the forward wrote for them (same addressing as there) or
 * NULL -- with it, the (1024, 256) and (2048, 512) configurations run the fast kernels K6F / K6E (dmel = g exp(-logmel),
 * no mel recomputation, every gradient sample written once); without it, or for any other configuration, the generic kernel K6 recomputes
 * the mel spectrum; grad_logmel: device, same addressing as logmel; grad_stems[s]: device, element (b, c, t) at
 * grad_stems[s][b*grad_stride_b + c*grad_stride_c + t].  Returns MIXFE_ESHORT when T <= n_fft/2. */
int mixfe_logmel_backward(mixfe_plan* plan, const float* const stems[MIXFE_NUM_STEMS],
                          long long stride_b, long long stride_c, int B, int T,
                          const float* logmel, long long logmel_stride_b,
                          const float* grad_logmel, long long grad_logmel_stride_b,
                          float* const grad_stems[MIXFE_NUM_STEMS], long long grad_stride_b,
                          long long grad_stride_c, void* cuda_stream);

/* EXTENSION (no counterpart in the reference, whose compute_loudness is un-weighted and un-gated,
 * src/mixing_utils.py:311-318): integrated loudness after ITU-R BS.1770-4 --
 * K-weighting, 400 ms blocks with 75 % overlap, absolute gate -70 LUFS, relative gate -10 LU -- of each stem and
 * of the mixture.  lufs: device [B][5] fp32 = vocals, bass, drums, other, mixture (-inf when no block passes the
 * gates).  Needs sample_rate % 10 == 0 and T >= 0.4 s (MIXFE_EINVAL otherwise).  Not part of the feature vector. */
int mixfe_gated_loudness(mixfe_plan* plan, const float* const stems[MIXFE_NUM_STEMS],
                         long long stride_b, long long stride_c, int B, int T, float* lufs, void* cuda_stream);

/* EXTENSION (same status as mixfe_gated_loudness): the proposal's dynamic-range descriptor
 * DR_IQR = IQR(L_short(t)) (Mixing_Style_Representation_Learning_Proposal.pdf p.2 section 2.1; no code in the
 * reference).  Short-term loudness = K-weighted loudness of 3 s windows every 100 ms (EBU Tech 3341); windows under
 * -70 LUFS are dropped; iqr = q75 - q25 of the rest, linear interpolation between order statistics.  iqr: device
 * [B][5] fp32 in LU = vocals, bass, drums, other, mixture (0 when fewer than two windows remain).  Needs
 * sample_rate % 10 == 0 and 3 s <= T <= 6 min 52 s (MIXFE_EINVAL / MIXFE_EUNSUPPORTED otherwise). */
int mixfe_short_term_loudness_iqr(mixfe_plan* plan, const float* const stems[MIXFE_NUM_STEMS],
                                  long long stride_b, long long stride_c, int B, int T, float* iqr, void* cuda_stream);

/* EXTENSION (same status): the proposal's activity- and band-weighted inter-stem masking (PDF p.2 section 2.4; the
 * reference implements only the un-weighted mean, src/mixing_utils.py:296-306):
 *   M_i = sum_{k,t} a_i(t) A(k) sigmoid((beta - Delta_i(k,t)) / tau) / sum_{k,t} a_i(t) A(k),
 *   Delta_i = S_i - max_{j != i} S_j on the channel-mean mel power S.  Choices the proposal leaves open: a_i(t) = E_i / (E_i + E_0)
 * with E_i(t) the frame's mel energy and E_0 that of a -60 dBFS white-noise frame (0 under 1e-6 E_0); A(k) = band_weights[k] (device,
 * n_mels floats; the Python layer passes the IEC 61672 A-weighting at the band centres), band_weight_sum their sum.
 * logmel: device, the tensor mixfe_features_logmel wrote for the same clips (same addressing); out: device [B][4] fp32
 * = vocals, bass, drums, other (0 for a stem that is never active).  beta = 0, tau = 1 are the reference's constants. */
int mixfe_weighted_masking(mixfe_plan* plan, const float* logmel, long long logmel_stride_b, int B, int T,
                           const float* band_weights, double band_weight_sum, float beta, float tau, float* out,
                           void* cuda_stream);

/* NO COUNTERPART IN THE REFERENCE.  mixfe_weighted_masking for callers that
 * must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): same
 * arguments, same two launches, same result, but the partial sums the two launches exchange live in a static 256 KB
 * device array inside libmixfe (held by every GPU that loads the library until exit) instead of the plan's workspace.
 * The call allocates nothing (no cudaMalloc / cudaFree / events; the plan is only read).
 * Limit: B * ceil((1 + T / hop) / 128) <= MIXFE_NOALLOC_MASKING_MAX_BLOCKS, else MIXFE_EUNSUPPORTED (e.g. 10 s clips at
 * hop 256: 14 blocks per clip, 585 clips per call).  ALL calls on one GPU share the array: calls that overlap on the GPU
 * (different streams, or graphs replayed concurrently) race; serialise them.  Nothing in it needs clearing between calls. */
#define MIXFE_NOALLOC_MASKING_MAX_BLOCKS 8192
int mixfe_weighted_masking_noalloc(mixfe_plan* plan, const float* logmel, long long logmel_stride_b, int B, int T,
                                   const float* band_weights, double band_weight_sum, float beta, float tau, float* out,
                                   void* cuda_stream);

/* EXTENSION (same status): a compressor with an envelope follower, which the brief names and the reference does not have
 * (its apply_compression is memoryless, src/mixing_utils.py:435-447).  Decoupled smooth peak detector (Giannoulis,
 * Massberg, Reiss, JAES 2012): p[n] = max(|x[n]|, release_coeff p[n-1]); s[n] = attack_coeff s[n-1] + (1 - attack_coeff) p[n];
 * the reference's static curve (threshold_db, ratio) is applied to the level of s and y[n] = x[n] gain(s[n]).  Coefficients
 * are per sample, exp(-1 / (time_constant_seconds * sample_rate)), in [0, 1); both 0 gives the memoryless compressor.
 * in/out: one tensor, element (b, c, t) at in[b*stride_b + c*stride_c + t]; every row is an independent detector. */
int mixfe_envelope_compress(mixfe_plan* plan, const float* in, float* out, long long stride_b, long long stride_c,
                            int B, int C, int T, float threshold_db, float ratio, float attack_coeff, float release_coeff,
                            void* cuda_stream);

/* Replaces ONE resolution of MultiResolutionSTFTLoss.forward (src/loss.py:332-448; the cycle loss of
 * src/train_style_transfer.py:239) and its autograd.  The plan's (n_fft, hop) is the resolution (win = n_fft, periodic
 * Hann, center=True, reflect).  x (prediction), y (target): device fp32, element (b, c, t) at
 * x[b*x_stride_b + c*x_stride_c + t].
 *   forward : state (device, 8 doubles) <- { sum (|Y|-|X|)^2, sum |Y|^2, sum |log(|X|+1e-5) - log(|Y|+1e-5)|,
 *             spectral convergence (:395), log-magnitude L1 (:408), and the two gradient factors, element count }.
 *             The resolution's loss is state[3] + state[4]; the module averages over resolutions (:444).
 *   backward: ADDS grad_scale[0] * d(state[3] + state[4]) / dx into grad_x (zero-initialised by the caller);
 *             grad_scale is a device scalar = upstream gradient / number of resolutions; `state` is the forward's. */
int mixfe_stft_loss_forward(mixfe_plan* plan, const float* x, const float* y, int B, int C, int T,
                            long long x_stride_b, long long x_stride_c, long long y_stride_b, long long y_stride_c,
                            double* state, void* cuda_stream);
int mixfe_stft_loss_backward(mixfe_plan* plan, const float* x, const float* y, int B, int C, int T,
                             long long x_stride_b, long long x_stride_c, long long y_stride_b, long long y_stride_c,
                             const double* state, const float* grad_scale, float* grad_x,
                             long long g_stride_b, long long g_stride_c, void* cuda_stream);

/* Test hook: mel POWER (before the log) for every channel, same addressing as logmel. */
int mixfe_debug_mel_power(mixfe_plan* plan, const float* const stems[MIXFE_NUM_STEMS],
                          long long stride_b, long long stride_c, int B, int T,
                          float* melpow, long long melpow_stride_b, void* cuda_stream);

/* Test hook: launch shape of the fused feature kernel K1 -- resident CTAs per SM (2 on a B200 for the
 * reference configuration), its dynamic shared memory, and the steps of the warp-local mel walk. */
int mixfe_debug_feature_kernel_config(const mixfe_plan* plan, int* ctas_per_sm, int* smem_bytes, int* mel_steps);

/* Device-to-device copy by a copy engine, stream-ordered on `cuda_stream`.  dst may be the NVSwitch multicast
 * mapping of a symmetric allocation: one copy then lands in every rank's buffer (distributed.GlobalBatch,
 * the all-gather of BASELINE config 5 without kernels). */
int mixfe_copy_async(void* dst, const void* src, size_t bytes, void* cuda_stream);

/* Strided copy by a copy engine, stream-ordered: `blocks` blocks of `rows` runs of width_bytes; run r of block k goes from
 * src + (k*src_block_rows + r)*src_pitch to dst + (k*dst_block_rows + r)*dst_pitch (blocks = 1: the block strides are
 * ignored).  Either side may be pinned host memory (direction inferred).  The host pipelines (pipeline.py) use it to move
 * the anchor and the positive rows of a triplet chunk -- two ranges a constant stride apart in the caller's (2G, 2, T)
 * stem tensors (the reference's collate layout, src/data.py:231-265) -- in ONE transfer instead of two, and all four
 * stems in one when the caller's stem tensors are equally spaced views of one allocation. */
int mixfe_copy_strided_async(void* dst, size_t dst_pitch, size_t dst_block_rows, const void* src, size_t src_pitch,
                             size_t src_block_rows, size_t width_bytes, size_t rows, size_t blocks, void* cuda_stream);

/* The all-gather of BASELINE config 5 driven by SMs: copies `bytes` (a multiple of 16) from src to the same position of
 * n_peers peer buffers (dst: HOST array of device pointers that are NVLink peer mappings of this process, e.g. torch
 * symmetric-memory buffers; <= 15), 128-bit loads and stores, n_ctas CTAs of 512 threads, stream-ordered.  Meant to run beside
 * K1 with mixfe_plan_reserve_sms(n_ctas) so that the gather of one chunk overlaps the kernels of the next. */
int mixfe_push_peers(void* const* dst, int n_peers, const void* src, size_t bytes, int n_ctas, void* cuda_stream);

/* K1 runs as a persistent grid that fills every SM; a collective launched beside it (the all-gather of the
 * previous chunk, distributed.extract_global_batch) cannot start until CTAs retire.  Leaving n_sms SMs out
 * of K1's grid lets it run concurrently.  0 restores the full grid. */
int mixfe_plan_reserve_sms(mixfe_plan* plan, int n_sms);

/* Number of kernel launches issued through this library since load (bench.py's gpu_launches). */
long long mixfe_launch_count(void);

/* Measurement hook: when enabled, every launch of the plan's kernels is bracketed by CUDA events on the
 * launching stream.  mixfe_plan_kernel_time() synchronises those events, returns the accumulated device
 * time (ms) and launch count of kernel `which` since the last query, and resets the accumulator. */
#define MIXFE_KERNEL_FEATURES 0   /* K1: fused STFT/mel/log/statistics */
#define MIXFE_KERNEL_FINALIZE 1   /* K2: per-clip feature finalize */
#define MIXFE_KERNEL_AUGMENT 2    /* K3: gain/biquad/compressor scan */
#define MIXFE_KERNEL_REVERB 3     /* K4/K5: FFT-domain reverb + redistribution (all its launches) */
#define MIXFE_KERNEL_LOGMEL_BWD 4 /* K6: backward of the log-mel */
#define MIXFE_KERNEL_LOUDNESS 5   /* K7: K-weighting scan + sub-block energies + gating */
#define MIXFE_KERNEL_STFT_LOSS 6  /* K8: STFT-loss partials / finish / backward */
#define MIXFE_KERNEL_EXTENSION 7  /* K9 / K10: weighted masking, envelope-follower compressor */
#define MIXFE_NUM_KERNELS 8
int mixfe_plan_enable_timing(mixfe_plan* plan, int enable);
int mixfe_plan_kernel_time(mixfe_plan* plan, int which, double* total_ms, long long* launches);

#ifdef __cplusplus
}
#endif
#endif /* MIXFE_H */
