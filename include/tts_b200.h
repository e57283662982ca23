/*
 * tts_b200 — C ABI of the B200-native MelGAN / MB-MelGAN / FB-MelGAN vocoder inference path.
 *
 * The reference (wjqkkky/TTS, Mozilla TTS 0.0.5) has no native boundary for this path: its
 * "plugin API" is the Python surface of
 *     TTS/vocoder/models/melgan_generator.py:8-97            (MelganGenerator)
 *     TTS/vocoder/models/multiband_melgan_generator.py:7-39  (MultibandMelganGenerator)
 *     TTS/vocoder/models/fullband_melgan_generator.py:6-30   (FullbandMelganGenerator)
 *     TTS/vocoder/layers/melgan.py:5-45                      (ResidualStack)
 *     TTS/vocoder/layers/pqmf.py:10-56                       (PQMF)
 * The functions below are what a maintainer's binding for those classes calls (ctypes stub in
 * INTEGRATION.md; the shipped binding is tts_b200/_lib.py).  Plain C: pointers, sizes, a stream
 * handle.  No torch types.  All device pointers are CUDA global-memory pointers on the plan's
 * device; `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *
 * Every function returns 0 on success or a negative ttsb200_status; the message of the last
 * failure on the calling thread is ttsb200_last_error().  Nothing here falls back to the CPU.
 */
#ifndef TTS_B200_H_
#define TTS_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define TTSB200_API __attribute__((visibility("default")))
#else
#define TTSB200_API
#endif

#define TTSB200_ABI_VERSION 2
#define TTSB200_MAX_STAGES 8

typedef enum ttsb200_status {
  TTSB200_OK = 0,
  TTSB200_ERR_INVALID = -1,   /* bad argument / unsupported configuration                        */
  TTSB200_ERR_SHAPE = -2,     /* input too short for the reflect padding (reference raises too)  */
  TTSB200_ERR_CUDA = -3,      /* a CUDA runtime call failed                                      */
  TTSB200_ERR_WORKSPACE = -4, /* workspace missing or smaller than ttsb200_workspace_bytes()     */
  TTSB200_ERR_ARCH = -5       /* device is not sm_100 (the library carries sm_100a code only)    */
} ttsb200_status;

typedef enum ttsb200_precision {
  TTSB200_F32 = 0, /* fp32 FFMA path: max-abs waveform error <= 1e-5 vs the reference              */
  TTSB200_BF16 = 1 /* BF16 operands, fp32 accumulate on tcgen05/TMEM: SNR >= 40 dB vs fp32         */
} ttsb200_precision;

/* Constructor arguments of the generator classes (melgan_generator.py:9-16,
 * multiband_melgan_generator.py:8-16, fullband_melgan_generator.py:7-14) plus the PQMF the
 * multi-band class owns (multiband_melgan_generator.py:24; pqmf.py:11). */
typedef struct ttsb200_generator_config {
  int32_t in_channels;      /* mel bins (80)                                                     */
  int32_t out_channels;     /* 1, or 4 sub-bands for MB-MelGAN                                   */
  int32_t proj_kernel;      /* odd (7)                                                           */
  int32_t base_channels;    /* 512 / 384                                                         */
  int32_t n_stages;         /* len(upsample_factors) <= TTSB200_MAX_STAGES                       */
  int32_t upsample_factors[TTSB200_MAX_STAGES];
  int32_t res_kernel;       /* odd (3); block b has dilation res_kernel**b (melgan.py:14)        */
  int32_t num_res_blocks;
  int32_t pqmf_bands;       /* 0 = no PQMF; else N (4) and out_channels must equal it            */
  int32_t pqmf_taps;        /* `taps` of PQMF (62): the synthesis filter has taps+1 coefficients */
  float act_slope;          /* LeakyReLU slope (0.2)                                             */
} ttsb200_generator_config;

/* Element type of the waveform a forward call writes.  Every reference caller ends in a 16-bit WAV
 * (TTS/utils/audio.py:338-340 save_wav -> int16 after a max-normalise); fp16 keeps |wav| <= 1 to a
 * relative 2^-11 and halves the device->host bytes of the result. */
typedef enum ttsb200_out_dtype {
  TTSB200_OUT_F32 = 0,
  TTSB200_OUT_F16 = 1
} ttsb200_out_dtype;

typedef struct ttsb200_plan ttsb200_plan;       /* immutable after creation; shareable across threads */
typedef struct ttsb200_profile ttsb200_profile; /* per-launch CUDA-event timings; one user at a time  */

typedef struct ttsb200_forward_args {
  const float* mel;         /* device, [batch, in_channels, frames] fp32, contiguous             */
  int32_t batch;
  int32_t frames;
  int32_t pad;              /* inference_padding: replicate-pad frames per side                  */
                            /* (melgan_generator.py:85-88); 0 for forward()                      */
  int32_t precision;        /* ttsb200_precision                                                 */
  void* out;                /* device, [batch, 1 or out_channels, ttsb200_output_length()],      */
                            /*   fp32, or fp16 when out_dtype == TTSB200_OUT_F16                 */
  void* workspace;          /* device scratch, >= ttsb200_workspace_bytes(), 256-byte aligned    */
  size_t workspace_bytes;
  void* stream;             /* cudaStream_t                                                      */
  int32_t skip_pqmf;        /* 1: stop after tanh and write the sub-bands (== forward())         */
  int32_t n_taps;           /* optional: copy outputs of reference `layers[i]` (per-layer parity,*/
  const int32_t* tap_layers;/*   the indexing of bin/convert_melgan_torch_to_tf.py:89-101)       */
  float* const* tap_out;    /*   each device [batch, C_i, L_i] fp32                              */
  ttsb200_profile* profile; /* optional                                                          */
  int32_t out_dtype;        /* ttsb200_out_dtype of `out` (taps are always fp32)                 */
  /* Optional output window (all zero = the whole result, densely packed): only samples             */
  /* [out_keep_lo, out_keep_lo + out_keep_len) of every row are stored, row b starting at           */
  /* out + b * out_batch_stride elements.  A long-form rank or a streaming window computes halo     */
  /* samples it must not emit; with a window its last kernel stores the kept samples straight into  */
  /* their place in the final buffer — which may be a peer GPU's memory mapped over NVLink — and     */
  /* nothing is gathered or concatenated afterwards (tts_b200/longform.py).                         */
  int64_t out_keep_lo;
  int64_t out_keep_len;
  int64_t out_batch_stride;
} ttsb200_forward_args;

TTSB200_API int32_t ttsb200_abi_version(void);
TTSB200_API const char* ttsb200_last_error(void);

/* Device check: returns 0 when `device` is an sm_100 part this library can run on. */
TTSB200_API int32_t ttsb200_device_ok(int32_t device);

/* Build a plan from folded (weight-norm removed: w = g*v/||v||, melgan_generator.py:91-97) fp32
 * HOST tensors, passed in the reference's state_dict() order of parametrised layers:
 *   layers.1 {weight [C0,in,k], bias}, then for each stage s:
 *     layers.{3+3s} ConvTranspose1d {weight [Cin,Cout,2f], bias},
 *     for b in blocks: layers.{4+3s}.blocks.b.2 {weight [C,C,res_kernel], bias}, .blocks.b.4 {weight [C,C,1], bias}
 *     for b in blocks: layers.{4+3s}.shortcuts.b {weight [C,C,1], bias}
 *   layers.{4+3S} head {weight [out,Clast,k], bias}
 *   and, when pqmf_bands > 0, the PQMF synthesis buffer G [1,N,taps+1] (pqmf.py:32,35).
 * n_tensors must equal ttsb200_expected_tensors(cfg).  Weights are re-laid-out for the kernels
 * and uploaded to `device`; the host tensors are not referenced after the call returns. */
TTSB200_API int32_t ttsb200_expected_tensors(const ttsb200_generator_config* cfg);
TTSB200_API int32_t ttsb200_plan_create(const ttsb200_generator_config* cfg, const float* const* tensors,
                            int32_t n_tensors, int32_t device, ttsb200_plan** out_plan);
TTSB200_API void ttsb200_plan_destroy(ttsb200_plan* plan);

/* Samples per item: (frames + 2*pad) * prod(upsample_factors) * (pqmf_bands or 1). */
TTSB200_API int64_t ttsb200_output_length(const ttsb200_plan* plan, int32_t frames, int32_t pad, int32_t skip_pqmf);
TTSB200_API int64_t ttsb200_workspace_bytes(const ttsb200_plan* plan, int32_t batch, int32_t frames, int32_t pad,
                                int32_t precision);
/* Shape [batch, C_i, L_i] of reference layers[layer] output, for sizing tap buffers. */
TTSB200_API int32_t ttsb200_layer_shape(const ttsb200_plan* plan, int32_t layer, int32_t frames, int32_t pad,
                            int32_t* channels, int64_t* length);

/* generator.inference(mel) / generator.forward(mel): asynchronous on args->stream, device buffers. */
TTSB200_API int32_t ttsb200_generator_forward(const ttsb200_plan* plan, const ttsb200_forward_args* args);

/* The same forward with NO allocation anywhere on the call: args->workspace / workspace_bytes are ignored and the
 * scratch is a file-scope `static __device__` array of TTSB200_STATIC_WORKSPACE_BYTES inside the library (every GPU
 * that loads the library holds those bytes until exit).  For callers that cannot call cudaMalloc: code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it.  No counterpart in the reference.
 * Shapes whose ttsb200_workspace_bytes() exceeds the array are refused with
 * TTSB200_ERR_WORKSPACE (the message names both sizes); args->profile must be NULL (a profile creates events).
 * Every call on a device shares the one array, which keeps the previous call's contents (nothing in it has to start
 * at zero: each region is written before it is read): calls that overlap on the GPU — two streams, two threads —
 * race; order them on one stream or with events.  The first call per device resolves the array's address
 * (cudaGetSymbolAddress); make it outside a graph capture. */
#define TTSB200_STATIC_WORKSPACE_BYTES (32u << 20)
TTSB200_API int32_t ttsb200_generator_forward_static(const ttsb200_plan* plan, const ttsb200_forward_args* args);

/* Same with HOST buffers (pageable or pinned): copies mel in, runs, copies the waveform out and
 * synchronises — the call TTS/server/synthesizer.py:153-159 amounts to.  Uses plan-owned device
 * and pinned staging buffers (serialised by a mutex). */
TTSB200_API int32_t ttsb200_generator_forward_host(ttsb200_plan* plan, const float* mel_host, int32_t batch,
                                       int32_t frames, int32_t pad, int32_t precision,
                                       float* out_host);

/* PQMF.synthesis(x) (pqmf.py:51-56): x [batch, N, Ts] -> y [batch, 1, N*Ts + 2*(taps/2) - taps].
 * G is the module's device buffer [1, N, taps+1]. */
TTSB200_API int32_t ttsb200_pqmf_synthesis(const float* x, const float* G, int32_t batch, int32_t bands,
                               int32_t taps, int64_t Ts, float* y, void* stream);
/* PQMF.analysis(x) (pqmf.py:48-49): x [batch, 1, T] -> y [batch, N, (T + 2*(taps/2) - taps - 1)/N + 1].
 * H is the module's device buffer [N, 1, taps+1]. */
TTSB200_API int32_t ttsb200_pqmf_analysis(const float* x, const float* H, int32_t batch, int32_t bands,
                              int32_t taps, int64_t T, float* y, void* stream);

/* Host-only helper: factorises the 4-band / 63-tap synthesis bank G [4][63] (host pointer) into the table the
 * generator's fast synthesis kernel uses — G[k][j] = q[j] * C[k][j] with the cosine modulation of pqmf.py:22-32,
 * table = C[k][p] as [8][4] then the signed prototype taps as [8][8] (96 floats).  Returns 1 when G has that
 * structure (ttsb200_plan_create then uses the 24-MAC kernel), 0 when it does not (direct 63-tap kernel), negative
 * on error.  Exposed so that the factorisation can be checked against the oracle without a GPU. */
TTSB200_API int32_t ttsb200_pqmf_factor_4x63(const float* G_host, float* table96_host);

/* Per-launch timing.  Attach via args->profile; after the stream has been synchronised,
 * ttsb200_profile_read() returns one record per kernel launched since the last reset. */
typedef struct ttsb200_launch_record {
  const char* name;   /* static string: kernel family + role, e.g. "f32.resblock.conv2+shortcut" */
  int32_t layer;      /* reference layers[i] index the launch belongs to                         */
  float ms;           /* CUDA-event duration on the launch stream                                */
  double flops;       /* algorithmic FLOPs of the launch (2 x MACs)                              */
  double bytes;       /* algorithmic HBM bytes (inputs read once + outputs written once)         */
} ttsb200_launch_record;

TTSB200_API int32_t ttsb200_profile_create(ttsb200_profile** out_profile);
TTSB200_API void ttsb200_profile_destroy(ttsb200_profile* profile);
TTSB200_API void ttsb200_profile_reset(ttsb200_profile* profile);
TTSB200_API int32_t ttsb200_profile_read(ttsb200_profile* profile, ttsb200_launch_record* records, int32_t max_records);

/* Number of kernels one ttsb200_generator_forward() launches for this shape (bench "gpu_launches"). */
TTSB200_API int32_t ttsb200_launch_count(const ttsb200_plan* plan, int32_t precision, int32_t skip_pqmf);

/* ---- Mel front-end (the step in front of the vocoder) --------------------------------------
 * AudioProcessor.melspectrogram of TTS/utils/audio.py:216-222 in one kernel: pre-emphasis (:198-201), centred
 * reflect-padded STFT with a periodic Hann window (:258-266, librosa.stft), magnitude, mel filterbank (:203-204),
 * spec_gain * log10(max(1e-5, x)) (:190-191) and range or mean-var normalisation (:106-134).  The constructor
 * arguments it reads (:15-71): */
typedef struct ttsb200_mel_config {
  int32_t sample_rate;
  int32_t n_fft;          /* fft_size: a power of two in [256, 2048]                                   */
  int32_t hop_length;
  int32_t win_length;     /* <= n_fft (audio.py:72)                                                     */
  int32_t num_mels;       /* <= 128                                                                     */
  float preemphasis;      /* 0 = none                                                                   */
  float spec_gain;
  float ref_level_db;
  float min_level_db;     /* != 0 when range normalisation is used (audio.py:71)                        */
  float max_norm;
  int32_t signal_norm;
  int32_t symmetric_norm;
  int32_t clip_norm;
} ttsb200_mel_config;
typedef struct ttsb200_mel_plan ttsb200_mel_plan;

/* mel_basis: HOST [num_mels][n_fft/2 + 1] fp32, the matrix AudioProcessor._build_mel_basis returns (:85-93);
 * window: HOST [win_length] fp32 (scipy.signal.get_window('hann', win_length, fftbins=True));
 * mean / scale: HOST [num_mels] StandardScaler statistics of a stats_path config (:185-189), or both NULL. */
TTSB200_API int32_t ttsb200_mel_plan_create(const ttsb200_mel_config* cfg, const float* mel_basis, const float* window,
                                            const float* mean, const float* scale, int32_t device,
                                            ttsb200_mel_plan** out_plan);
TTSB200_API void ttsb200_mel_plan_destroy(ttsb200_mel_plan* plan);
/* 1 + n_samples / hop_length frames (librosa.stft, center=True). */
TTSB200_API int64_t ttsb200_mel_frames(const ttsb200_mel_plan* plan, int64_t n_samples);
/* wav: device [batch][n_samples] fp32 -> out: device [batch][num_mels][frames] fp32 (the vocoder's input layout). */
TTSB200_API int32_t ttsb200_mel_forward(const ttsb200_mel_plan* plan, const float* wav, int32_t batch, int64_t n_samples,
                                        float* out, void* stream);
/* Frame statistics for find_endpoint / trim_silence (audio.py:302-316): for frame f = samples [f*hop, f*hop +
 * frame_length) clipped to the signal, out[b][f] = {max x, max |x|, mean x^2}; ceil(n_samples / hop) frames. */
TTSB200_API int32_t ttsb200_frame_stats(const float* wav, int32_t batch, int64_t n_samples, int32_t frame_length,
                                        int32_t hop, float* out, void* stream);

/* Diagnostics for the roofline denominators MEASURED_PEAKS.json does not carry (bench.py).
 * fp32 FFMA rate of `device` in TFLOP/s (register-resident FMA loop on every SM, CUDA events) — the ceiling of the
 * TTSB200_F32 path; pinned host<->device copy rate in GB/s over `streams` concurrent copies of `bytes` each
 * (direction 0 = host to device, 1 = device to host) — the ceiling of the end-to-end figure.  Negative on error. */
TTSB200_API double ttsb200_diag_ffma_tflops(int32_t device);
TTSB200_API double ttsb200_diag_hostlink_gbs(int32_t device, int32_t direction, int64_t bytes, int32_t streams);

#ifdef __cplusplus
}
#endif
#endif /* TTS_B200_H_ */
