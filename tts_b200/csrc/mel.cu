// Mel front-end on the GPU: AudioProcessor.melspectrogram (TTS/utils/audio.py:216-222) =
// pre-emphasis (:198-201) -> librosa.stft, centred, reflect-padded, periodic Hann (:258-266) -> |.| -> mel filterbank
// (:203-204) -> spec_gain * log10(max(1e-5, .)) (:190-191) -> range or mean-var normalisation (:106-134),
// in ONE kernel; plus the frame statistics find_endpoint / trim_silence need (:302-316).
//
// HBM-bound by construction: a waveform sample is read once per overlapping frame (n_fft / hop times, served by
// L1 / L2), a mel value written once — 4 B in per sample + 4 * num_mels / hop B out per sample; the arithmetic (a
// 1,024-point FFT per frame) is ~50 FLOPs per input byte and stays in shared memory.
//
// One CTA = FR consecutive frames of one utterance.  Two real frames ride one complex FFT (frame A in the real part,
// frame B in the imaginary part; X_A[k] = (Z[k] + conj Z[N-k]) / 2, X_B[k] = (Z[k] - conj Z[N-k]) / 2i), computed
// with a radix-2 Stockham autosort FFT ping-ponging between two shared-memory buffers: no bit reversal, unit-stride
// conflict-free reads, twiddles from a table built in double on the host.  The filterbank is applied from a CSR
// copy of the reference's dense mel_basis (each triangle covers a contiguous run of bins), and the finished
// [num_mels x FR] tile is written with FR-frame runs per mel row.
#include <cmath>
#include <vector>

#include "common.cuh"

struct ttsb200_mel_plan {
  ttsb200_mel_config cfg{};
  int device = 0;
  int log2n = 0;
  float* window = nullptr;      // [n_fft] window, zero-padded about its centre
  float2* twiddle = nullptr;    // [n_fft / 2] exp(-2 pi i k / n_fft)
  int* row_ptr = nullptr;       // [num_mels + 1] CSR of mel_basis
  int* row_first = nullptr;     // [num_mels] first FFT bin of each row's run
  float* weights = nullptr;     // [nnz]
  float* mean = nullptr;        // [num_mels] or null
  float* inv_scale = nullptr;   // [num_mels] or null
};

namespace ttsb {
namespace {

constexpr int MEL_THREADS = 256;
constexpr int MEL_FR = 8;          // frames per CTA (4 complex FFTs)
constexpr int MEL_MAX_MELS = 128;

__device__ __forceinline__ int reflect_index(int i, int n) {   // numpy 'reflect' about the end samples
  if (n == 1) return 0;
  const int period = 2 * (n - 1);
  i = i < 0 ? -i : i;
  i %= period;
  return i < n ? i : period - i;
}

struct MelParams {
  const float* wav;       // [B][n_samples]
  int n_samples, n_frames, hop, num_mels, n_fft;
  const float* window;
  const float2* twiddle;
  const int* row_ptr;
  const int* row_first;
  const float* weights;
  const float* mean;
  const float* inv_scale;
  float preemph, spec_gain, ref_db, min_db, max_norm;
  int signal_norm, symmetric, clip;
  float* out;             // [B][num_mels][n_frames]
};

template <int LOG2N>
__global__ void __launch_bounds__(MEL_THREADS) mel_kernel(const MelParams p) {
  constexpr int N = 1 << LOG2N, H = N / 2;
  __shared__ float2 buf[2][N];
  constexpr bool TW_SMEM = LOG2N <= 10;          // 2,048 points: the table stays in global memory (48 KB static limit)
  __shared__ float2 tw_s[TW_SMEM ? H : 1];
  __shared__ float mag[2][H + 1];
  __shared__ float tile[MEL_MAX_MELS][MEL_FR + 1];

  const int tid = threadIdx.x;
  const int b = blockIdx.y;
  const int f0 = blockIdx.x * MEL_FR;
  const float* y = p.wav + (int64_t)b * p.n_samples;
  if (TW_SMEM)
    for (int k = tid; k < H; k += MEL_THREADS) tw_s[k] = p.twiddle[k];
  const float2* tw = TW_SMEM ? tw_s : p.twiddle;

  auto sample = [&](int frame, int j) -> float {
    if (frame >= p.n_frames) return 0.f;
    const int i = reflect_index(frame * p.hop - H + j, p.n_samples);
    float v = y[i];
    if (p.preemph != 0.f && i > 0) v -= p.preemph * y[i - 1];   // lfilter([1, -a], [1], y), zero initial state
    return v;
  };

  for (int pair = 0; pair < MEL_FR / 2; ++pair) {
    const int fa = f0 + 2 * pair, fb = fa + 1;
    if (fa >= p.n_frames) break;                  // uniform over the CTA
    __syncthreads();                              // tw ready; previous pair's mag / buf consumed
    for (int j = tid; j < N; j += MEL_THREADS) {
      const float w = p.window[j];
      buf[0][j] = make_float2(w * sample(fa, j), w * sample(fb, j));
    }
    __syncthreads();
    // Stockham radix-2, decimation in time: stage s combines sub-transforms of length 2^s
    int src = 0;
#pragma unroll
    for (int s = 0; s < LOG2N; ++s) {
      const int m = 1 << s;                       // half-length of the transforms being built
      for (int i = tid; i < H; i += MEL_THREADS) {
        const int k = i & (m - 1);                // index within the sub-transform
        const int j = i >> s;                     // which sub-transform
        const float2 a = buf[src][j * m + k];
        const float2 c = buf[src][j * m + k + H];
        const float2 t = tw[k << (LOG2N - 1 - s)];
        const float2 ct = make_float2(c.x * t.x - c.y * t.y, c.x * t.y + c.y * t.x);
        buf[src ^ 1][j * 2 * m + k] = make_float2(a.x + ct.x, a.y + ct.y);
        buf[src ^ 1][j * 2 * m + k + m] = make_float2(a.x - ct.x, a.y - ct.y);
      }
      src ^= 1;
      __syncthreads();
    }
    // split the two real spectra and take magnitudes
    for (int k = tid; k <= H; k += MEL_THREADS) {
      const float2 z = buf[src][k];
      const float2 zc = buf[src][(N - k) & (N - 1)];
      const float ar = 0.5f * (z.x + zc.x), ai = 0.5f * (z.y - zc.y);
      const float br = 0.5f * (z.y + zc.y), bi = -0.5f * (z.x - zc.x);
      mag[0][k] = sqrtf(ar * ar + ai * ai);
      mag[1][k] = sqrtf(br * br + bi * bi);
    }
    __syncthreads();
    // filterbank + dB + normalisation: thread (which frame of the pair, mel row)
    for (int e = tid; e < 2 * p.num_mels; e += MEL_THREADS) {
      const int which = e / p.num_mels, m = e - which * p.num_mels;
      const int r0 = p.row_ptr[m], r1 = p.row_ptr[m + 1], k0 = p.row_first[m];
      float acc = 0.f;
      for (int r = r0; r < r1; ++r) acc = fmaf(p.weights[r], mag[which][k0 + r - r0], acc);
      float S = p.spec_gain * log10f(fmaxf(1e-5f, acc));
      if (p.signal_norm) {
        if (p.mean) {
          S = (S - p.mean[m]) * p.inv_scale[m];
        } else {
          S -= p.ref_db;
          float sn = (S - p.min_db) / (-p.min_db);
          if (p.symmetric) {
            sn = 2.f * p.max_norm * sn - p.max_norm;
            if (p.clip) sn = fminf(fmaxf(sn, -p.max_norm), p.max_norm);
          } else {
            sn = p.max_norm * sn;
            if (p.clip) sn = fminf(fmaxf(sn, 0.f), p.max_norm);
          }
          S = sn;
        }
      }
      tile[m][2 * pair + which] = S;
    }
  }
  __syncthreads();
  const int nf = min(MEL_FR, p.n_frames - f0);
  for (int e = tid; e < p.num_mels * MEL_FR; e += MEL_THREADS) {
    const int m = e / MEL_FR, i = e - m * MEL_FR;
    if (i < nf) p.out[((int64_t)b * p.num_mels + m) * p.n_frames + f0 + i] = tile[m][i];
  }
}

// Per-frame statistics of a waveform: max of the signed samples (what find_endpoint thresholds, audio.py:307), max |x|
// and mean square (librosa.effects.trim's RMS) over [f * hop, f * hop + frame_length) clipped to the signal.
__global__ void frame_stats_kernel(const float* __restrict__ wav, int n_samples, int frame_length, int hop, int n_frames,
                                   float* __restrict__ out) {
  const int f = blockIdx.x, b = blockIdx.y;
  const float* y = wav + (int64_t)b * n_samples;
  const int lo = f * hop, hi = min(n_samples, lo + frame_length);
  float mx = -INFINITY, amx = 0.f, sq = 0.f;
  for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const float v = y[i];
    mx = fmaxf(mx, v);
    amx = fmaxf(amx, fabsf(v));
    sq = fmaf(v, v, sq);
  }
  __shared__ float s[3][32];
  for (int o = 16; o > 0; o >>= 1) {
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    amx = fmaxf(amx, __shfl_xor_sync(0xffffffffu, amx, o));
    sq += __shfl_xor_sync(0xffffffffu, sq, o);
  }
  if ((threadIdx.x & 31) == 0) {
    s[0][threadIdx.x >> 5] = mx; s[1][threadIdx.x >> 5] = amx; s[2][threadIdx.x >> 5] = sq;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int nw = (blockDim.x + 31) / 32;
    for (int w = 1; w < nw; ++w) {
      mx = fmaxf(mx, s[0][w]); amx = fmaxf(amx, s[1][w]); sq += s[2][w];
    }
    float* o = out + ((int64_t)b * n_frames + f) * 3;
    o[0] = mx; o[1] = amx; o[2] = hi > lo ? sq / (float)(hi - lo) : 0.f;
  }
}

}  // namespace
}  // namespace ttsb

using namespace ttsb;

extern "C" {

int32_t ttsb200_mel_plan_create(const ttsb200_mel_config* cfg, const float* mel_basis, const float* window,
                                const float* mean, const float* scale, int32_t device, ttsb200_mel_plan** out_plan) {
  if (!cfg || !mel_basis || !window || !out_plan) return fail(TTSB200_ERR_INVALID, "NULL argument");
  int log2n = 0;
  while ((1 << log2n) < cfg->n_fft) ++log2n;
  if ((1 << log2n) != cfg->n_fft || log2n < 8 || log2n > 11)
    return fail(TTSB200_ERR_INVALID, "n_fft must be a power of two in [256, 2048]");
  if (cfg->win_length <= 0 || cfg->win_length > cfg->n_fft)
    return fail(TTSB200_ERR_INVALID, " [!] win_length cannot be larger than fft_size");   // audio.py:72
  if (cfg->hop_length <= 0 || cfg->num_mels <= 0 || cfg->num_mels > MEL_MAX_MELS)
    return fail(TTSB200_ERR_INVALID, "hop_length must be positive and num_mels in [1, 128]");
  if (cfg->signal_norm && !mean && cfg->min_level_db == 0.f)
    return fail(TTSB200_ERR_INVALID, " [!] min_level_db is 0");                            // audio.py:71
  if ((mean == nullptr) != (scale == nullptr)) return fail(TTSB200_ERR_INVALID, "mean and scale go together");
  if (int rc = ttsb200_device_ok(device)) return rc;

  const int N = cfg->n_fft, H = N / 2, nbins = H + 1, M = cfg->num_mels;
  // window zero-padded about its centre (librosa.util.pad_center)
  std::vector<float> win(N, 0.f);
  const int lpad = (N - cfg->win_length) / 2;
  for (int i = 0; i < cfg->win_length; ++i) win[lpad + i] = window[i];
  std::vector<float2> tw(H);
  const double PI = 3.14159265358979323846;
  for (int k = 0; k < H; ++k) tw[k] = make_float2((float)std::cos(-2.0 * PI * k / N), (float)std::sin(-2.0 * PI * k / N));
  // CSR of the dense filterbank: each row's non-zeros must be one contiguous run of bins (triangular filters are)
  std::vector<int> row_ptr(M + 1, 0), row_first(M, 0);
  std::vector<float> weights;
  for (int m = 0; m < M; ++m) {
    int first = -1, last = -1;
    for (int k = 0; k < nbins; ++k)
      if (mel_basis[(size_t)m * nbins + k] != 0.f) {
        if (first < 0) first = k;
        last = k;
      }
    row_first[m] = first < 0 ? 0 : first;
    if (first >= 0)
      for (int k = first; k <= last; ++k) weights.push_back(mel_basis[(size_t)m * nbins + k]);   // zeros inside a run kept
    row_ptr[m + 1] = (int)weights.size();
  }
  if (weights.empty()) weights.push_back(0.f);
  std::vector<float> inv_scale;
  if (scale) {
    inv_scale.resize(M);
    for (int m = 0; m < M; ++m) {
      if (scale[m] == 0.f) return fail(TTSB200_ERR_INVALID, "scale has a zero entry");
      inv_scale[m] = 1.f / scale[m];
    }
  }

  ttsb200_mel_plan* plan = new ttsb200_mel_plan();
  plan->cfg = *cfg;
  plan->device = device;
  plan->log2n = log2n;
  int prev = 0;
  cudaGetDevice(&prev);
  cudaError_t e = cudaSetDevice(device);
  auto up = [&](void** dst, const void* src, size_t bytes) {
    if (e == cudaSuccess) e = cudaMalloc(dst, bytes);
    if (e == cudaSuccess) e = cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice);
  };
  up(reinterpret_cast<void**>(&plan->window), win.data(), win.size() * sizeof(float));
  up(reinterpret_cast<void**>(&plan->twiddle), tw.data(), tw.size() * sizeof(float2));
  up(reinterpret_cast<void**>(&plan->row_ptr), row_ptr.data(), row_ptr.size() * sizeof(int));
  up(reinterpret_cast<void**>(&plan->row_first), row_first.data(), row_first.size() * sizeof(int));
  up(reinterpret_cast<void**>(&plan->weights), weights.data(), weights.size() * sizeof(float));
  if (mean) {
    up(reinterpret_cast<void**>(&plan->mean), mean, M * sizeof(float));
    up(reinterpret_cast<void**>(&plan->inv_scale), inv_scale.data(), M * sizeof(float));
  }
  cudaSetDevice(prev);
  if (e != cudaSuccess) {
    ttsb200_mel_plan_destroy(plan);
    return fail(TTSB200_ERR_CUDA, std::string("mel plan upload: ") + cudaGetErrorString(e));
  }
  *out_plan = plan;
  return 0;
}

void ttsb200_mel_plan_destroy(ttsb200_mel_plan* plan) {
  if (!plan) return;
  for (void* p : {(void*)plan->window, (void*)plan->twiddle, (void*)plan->row_ptr, (void*)plan->row_first,
                  (void*)plan->weights, (void*)plan->mean, (void*)plan->inv_scale})
    if (p) cudaFree(p);
  delete plan;
}

int64_t ttsb200_mel_frames(const ttsb200_mel_plan* plan, int64_t n_samples) {
  if (!plan) return fail(TTSB200_ERR_INVALID, "plan is NULL");
  return 1 + n_samples / plan->cfg.hop_length;   // librosa.stft, center=True
}

int32_t ttsb200_mel_forward(const ttsb200_mel_plan* plan, const float* wav, int32_t batch, int64_t n_samples, float* out,
                            void* stream) {
  if (!plan || !wav || !out) return fail(TTSB200_ERR_INVALID, "NULL argument");
  if (batch <= 0 || batch > 65535 || n_samples >= (1ll << 30))
    return fail(TTSB200_ERR_INVALID, "batch / n_samples out of range");
  if (n_samples <= plan->cfg.n_fft / 2)   // numpy's reflect padding needs pad < length; the reference raises too
    return fail(TTSB200_ERR_SHAPE, "waveform shorter than fft_size // 2 + 1 samples: reflect padding is undefined");
  int prev = 0;
  TTSB_CUDA(cudaGetDevice(&prev));
  if (prev != plan->device) TTSB_CUDA(cudaSetDevice(plan->device));
  struct Restore {
    int prev, cur;
    ~Restore() { if (prev != cur) cudaSetDevice(prev); }
  } restore{prev, plan->device};
  const ttsb200_mel_config& c = plan->cfg;
  MelParams p{};
  p.wav = wav; p.n_samples = (int)n_samples; p.n_frames = (int)(1 + n_samples / c.hop_length); p.hop = c.hop_length;
  p.num_mels = c.num_mels; p.n_fft = c.n_fft; p.window = plan->window; p.twiddle = plan->twiddle;
  p.row_ptr = plan->row_ptr; p.row_first = plan->row_first; p.weights = plan->weights;
  p.mean = plan->mean; p.inv_scale = plan->inv_scale;
  p.preemph = c.preemphasis; p.spec_gain = c.spec_gain; p.ref_db = c.ref_level_db; p.min_db = c.min_level_db;
  p.max_norm = c.max_norm; p.signal_norm = c.signal_norm; p.symmetric = c.symmetric_norm; p.clip = c.clip_norm;
  p.out = out;
  const dim3 grid(ceil_div(p.n_frames, MEL_FR), batch);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (plan->log2n) {
    case 8: mel_kernel<8><<<grid, MEL_THREADS, 0, st>>>(p); break;
    case 9: mel_kernel<9><<<grid, MEL_THREADS, 0, st>>>(p); break;
    case 10: mel_kernel<10><<<grid, MEL_THREADS, 0, st>>>(p); break;
    case 11: mel_kernel<11><<<grid, MEL_THREADS, 0, st>>>(p); break;
    default: return fail(TTSB200_ERR_INVALID, "unsupported n_fft");
  }
  TTSB_CUDA(cudaGetLastError());
  return 0;
}

int32_t ttsb200_frame_stats(const float* wav, int32_t batch, int64_t n_samples, int32_t frame_length, int32_t hop,
                            float* out, void* stream) {
  if (!wav || !out) return fail(TTSB200_ERR_INVALID, "NULL argument");
  if (batch <= 0 || batch > 65535 || n_samples <= 0 || n_samples >= (1ll << 30) || frame_length <= 0 || hop <= 0)
    return fail(TTSB200_ERR_INVALID, "batch / n_samples / frame_length / hop out of range");
  const int n_frames = (int)((n_samples + hop - 1) / hop);
  frame_stats_kernel<<<dim3(n_frames, batch), 128, 0, static_cast<cudaStream_t>(stream)>>>(wav, (int)n_samples, frame_length,
                                                                                            hop, n_frames, out);
  TTSB_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
