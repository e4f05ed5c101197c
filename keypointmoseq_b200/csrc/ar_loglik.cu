// a1: AR emission log-likelihood for every (session, frame, state).
// Reference: gibbs.py:100-104,110 (_ar_log_likelihood, lax.map over states);
// util.py:192-211 (get_lags), util.py:169-176 (pad_affine);
// tfd.MultivariateNormalFullCovariance(mu, Q).log_prob.
//
// Whitened form: with Q_k = Lq Lq^T,  W_k = Lq^-1 [ -A_k | I | -b_k ]  (D x P, P = D*L+D+1)
//   r = W_k f,  f = [x[t-L..t-1] (oldest first), x[t], 1]
//   ll = -1/2 |r|^2 - sum log diag Lq - D/2 log 2pi
// so the whole op is one (frames x P) . (P x K*D) contraction plus a square-and-group-sum
// epilogue; the lag embedding is an overlapping window of x (never materialised).
#include "common.cuh"
#include "dispatch.cuh"

#define AR_DMAX 16

static inline int ar_pp(int D, int L) { return (D * L + D + 1 + 3) / 4 * 4; }

size_t kpms_ar_tc_tile_bytes(int K, int D, int L);  // ar_loglik_tc.cu

// workspace: [Wf: K*D*PP f32][cst: K f32][TF32 hi/lo weight tiles of the tensor-core kernel]
extern "C" size_t kpms_ar_loglik_workspace_bytes(int K, int D, int L) {
  if (K <= 0 || D <= 0 || L <= 0) return 0;
  size_t w = (size_t)K * D * ar_pp(D, L) * sizeof(float);
  return kpms_round_up(w, 256) + kpms_round_up((size_t)K * sizeof(float), 256) +
         kpms_ar_tc_tile_bytes(K, D, L);
}

// One warp per state: lane 0 factors Q_k, then the P columns of W_k are solved in parallel.
__global__ void ar_prepare_kernel(const double* __restrict__ Ab, const double* __restrict__ Q,
                                  int D, int L, int PP, float* __restrict__ Wf,
                                  float* __restrict__ cst) {
  const int k = blockIdx.x, lane = threadIdx.x;
  const int DL = D * L, P = DL + D + 1;
  __shared__ double Lq[AR_DMAX * AR_DMAX];
  __shared__ double ldsum;
  if (lane == 0) {
    const double* Qk = Q + (size_t)k * D * D;
    double ls = 0.0;
    for (int j = 0; j < D; j++) {
      double s = Qk[j * D + j];
      for (int m = 0; m < j; m++) s -= Lq[j * AR_DMAX + m] * Lq[j * AR_DMAX + m];
      double d = sqrt(s);
      Lq[j * AR_DMAX + j] = d;
      ls += log(d);
      for (int i = j + 1; i < D; i++) {
        double t = Qk[i * D + j];
        for (int m = 0; m < j; m++) t -= Lq[i * AR_DMAX + m] * Lq[j * AR_DMAX + m];
        Lq[i * AR_DMAX + j] = t / d;
      }
    }
    ldsum = ls;
  }
  __syncwarp();
  const double* Abk = Ab + (size_t)k * D * (DL + 1);
  for (int c = lane; c < PP; c += 32) {
    double col[AR_DMAX];
    for (int i = 0; i < D; i++) {
      double rhs;
      if (c < DL) rhs = -Abk[i * (DL + 1) + c];
      else if (c < DL + D) rhs = (c - DL == i) ? 1.0 : 0.0;
      else if (c == P - 1) rhs = -Abk[i * (DL + 1) + DL];
      else rhs = 0.0;
      for (int m = 0; m < i; m++) rhs -= Lq[i * AR_DMAX + m] * col[m];
      col[i] = rhs / Lq[i * AR_DMAX + i];
    }
    for (int i = 0; i < D; i++) Wf[((size_t)k * D + i) * PP + c] = (float)col[i];
  }
  if (lane == 0) cst[k] = (float)(-ldsum - 0.5 * D * KPMS_LOG_2PI);
}

// CUDA-core fp32 kernel: one thread per frame, the P inputs in registers, a tile of KT
// states' weights in shared memory (broadcast reads).
template <int DD, int LL, int KT>
__global__ void __launch_bounds__(256)
    ar_loglik_simt_kernel(const float* __restrict__ x, const float* __restrict__ Wf,
                          const float* __restrict__ cst, int N, int T, int K,
                          float* __restrict__ ll) {
  constexpr int DL = DD * LL, P = DL + DD + 1, PP = (P + 3) / 4 * 4;
  extern __shared__ __align__(16) float sm[];
  float* Ws = sm;                 // KT*DD*PP
  float* cs = sm + KT * DD * PP;  // KT
  const int Tm = T - LL;
  const long long total = (long long)N * Tm;
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = g < total;
  float f[PP];
#pragma unroll
  for (int p = 0; p < PP; p++) f[p] = 0.0f;
  long long obase = 0;
  if (live) {
    const int n = (int)(g / Tm), tp = (int)(g % Tm);
    const float* xp = x + ((size_t)n * T + tp) * DD;  // x[n][tp .. tp+L] = lags then current
#pragma unroll
    for (int p = 0; p < DL + DD; p++) f[p] = xp[p];
    f[P - 1] = 1.0f;
    obase = g * K;
  }
  for (int k0 = 0; k0 < K; k0 += KT) {
    const int kt = min(KT, K - k0);
    __syncthreads();
    for (int i = threadIdx.x; i < kt * DD * PP / 4; i += blockDim.x)
      reinterpret_cast<float4*>(Ws)[i] =
          reinterpret_cast<const float4*>(Wf + (size_t)k0 * DD * PP)[i];
    for (int i = threadIdx.x; i < kt; i += blockDim.x) cs[i] = cst[k0 + i];
    __syncthreads();
    if (!live) continue;
    for (int kk = 0; kk < kt; kk++) {
      float q = 0.0f;
      const float4* wrow = reinterpret_cast<const float4*>(Ws + kk * DD * PP);
#pragma unroll
      for (int d = 0; d < DD; d++) {
        float a0 = 0.0f, a1 = 0.0f;
#pragma unroll
        for (int p4 = 0; p4 < PP / 4; p4++) {
          float4 w = wrow[d * (PP / 4) + p4];
          a0 = fmaf(w.x, f[4 * p4 + 0], a0);
          a1 = fmaf(w.y, f[4 * p4 + 1], a1);
          a0 = fmaf(w.z, f[4 * p4 + 2], a0);
          a1 = fmaf(w.w, f[4 * p4 + 3], a1);
        }
        float r = a0 + a1;
        q = fmaf(r, r, q);
      }
      ll[obase + k0 + kk] = cs[kk] - 0.5f * q;
    }
  }
}

template <int DD, int LL>
static int launch_simt(const float* x, const float* Wf, const float* cst, int N, int T, int K,
                       float* ll, cudaStream_t st) {
  constexpr int KT = 32;
  constexpr int P = DD * LL + DD + 1, PP = (P + 3) / 4 * 4;
  const size_t smem = (size_t)(KT * DD * PP + KT) * sizeof(float);
  auto kern = ar_loglik_simt_kernel<DD, LL, KT>;
  KPMS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long total = (long long)N * (T - LL);
  const unsigned grid = (unsigned)((total + 255) / 256);
  kern<<<grid, 256, smem, st>>>(x, Wf, cst, N, T, K, ll);
  return kpms_last_launch();
}

int kpms_ar_loglik_tc(const float* x, const float* Wf, const float* cst, int N, int T, int D, int L,
                      int K, float* ll, void* tiles, size_t tile_bytes, int tile_cl, int tile_nch,
                      cudaStream_t st);  // ar_loglik_tc.cu

// tile_cl > 0 asks for ll in the HMM's interleaved layout (tensor-core kernel only); *wrote_tiled
// reports whether that happened — otherwise ll is (N, T-L, K) row-major.
int kpms_ar_loglik_impl(const float* x, const double* Ab, const double* Q, int N, int T, int D, int L,
                        int K, float* ll, void* workspace, size_t workspace_bytes, int impl,
                        int tile_cl, int tile_nch, int* wrote_tiled, void* stream) {
  if (wrote_tiled) *wrote_tiled = 0;
  KPMS_REQUIRE(x && Ab && Q && ll && workspace, KPMS_ERR_NULL);
  KPMS_REQUIRE(N > 0 && T > L && D > 0 && L > 0 && K > 0, KPMS_ERR_SHAPE);
  KPMS_REQUIRE(D <= AR_DMAX && kpms_dl_supported(D, L), KPMS_ERR_UNSUPPORTED);
  KPMS_REQUIRE(kpms_aligned(workspace, 256) &&
                   workspace_bytes >= kpms_ar_loglik_workspace_bytes(K, D, L),
               KPMS_ERR_WORKSPACE);
  cudaStream_t st = (cudaStream_t)stream;
  const int PP = ar_pp(D, L);
  float* Wf = (float*)workspace;
  float* cst = (float*)((char*)workspace + kpms_round_up((size_t)K * D * PP * sizeof(float), 256));
  void* tiles = (char*)cst + kpms_round_up((size_t)K * sizeof(float), 256);
  ar_prepare_kernel<<<K, 32, 0, st>>>(Ab, Q, D, L, PP, Wf, cst);
  int rc = kpms_last_launch();
  if (rc) return rc;
  if (impl == 0) {
    if (tile_cl > 0) {
      rc = kpms_ar_loglik_tc(x, Wf, cst, N, T, D, L, K, ll, tiles, kpms_ar_tc_tile_bytes(K, D, L), tile_cl,
                             tile_nch, st);
      if (rc == KPMS_OK && wrote_tiled) *wrote_tiled = 1;
      if (rc != KPMS_ERR_UNSUPPORTED) return rc;
    }
    rc = kpms_ar_loglik_tc(x, Wf, cst, N, T, D, L, K, ll, tiles, kpms_ar_tc_tile_bytes(K, D, L), 0, 0, st);
    if (rc != KPMS_ERR_UNSUPPORTED) return rc;
  }
#define KPMS_X(d, l) \
  if (D == d && L == l) return launch_simt<d, l>(x, Wf, cst, N, T, K, ll, st);
  KPMS_FOR_EACH_DL(KPMS_X)
#undef KPMS_X
  return KPMS_ERR_UNSUPPORTED;
}

extern "C" int kpms_ar_loglik(const float* x, const double* Ab, const double* Q, int N, int T,
                              int D, int L, int K, float* ll, void* workspace,
                              size_t workspace_bytes, int impl, void* stream) {
  return kpms_ar_loglik_impl(x, Ab, Q, N, T, D, L, K, ll, workspace, workspace_bytes, impl, 0, 0, nullptr,
                             stream);
}

// ======================================================================================
// Allocation-free variant of kpms_ar_loglik.
//
// Who needs it: a caller that is handed its buffers by someone else and is not allowed to
// create device memory of its own.  The obvious case for this library is the reference itself:
// it is a JAX program, and the natural way to plug these kernels into it is an XLA custom call.
// Such a handler receives operand and result buffers from XLA and nothing more; XLA has usually
// claimed most of the GPU's memory for its own pool already, so a cudaMalloc from inside the
// handler can fail outright, and when it succeeds it stalls the whole device while the driver
// maps the memory.  The same holds for any host loop that must not block on the driver.
//
// kpms_ar_loglik needs a few hundred kilobytes of scratch for the whitened weights; here that
// scratch is a fixed array that lives in the library image on every GPU that loads it (and stays
// there until the process exits), so the caller passes only inputs and the output.
//
// Rules of use
//   * Limit: the weights of all states must fit KPMS_AR_STATIC_FLOATS floats
//     (round_up(K*D*PP, 64) + K <= 65536, PP = D*L+D+1 rounded up to 4; K = 128 states at
//     D = 10, L = 3 fit).  Larger problems are refused with KPMS_ERR_TOO_LARGE before anything
//     is launched; use kpms_ar_loglik with a caller workspace instead.
//   * The first call on each device looks up the device address of the array (with that
//     device current).  Make that first call outside of any stream capture.
//   * There is ONE array per device and every call writes it.  Two calls whose kernels can be
//     in flight together on the same GPU (different streams, different host threads) race on
//     it; order them on one stream or with events.  The array keeps whatever the previous call
//     left behind; this entry point never depends on that, because the prepare kernel rewrites
//     every weight and constant it later reads, so nothing has to be cleared.
#define KPMS_AR_STATIC_FLOATS (64 * 1024)
#define KPMS_AR_STATIC_MAX_DEVICES 64

static __device__ __align__(256) float g_ar_static_scratch[KPMS_AR_STATIC_FLOATS];

#include <mutex>
static std::mutex g_ar_static_mutex;
static float* g_ar_static_dev_ptr[KPMS_AR_STATIC_MAX_DEVICES] = {nullptr};

// Device address of the scratch array on the current device.  In host code the name
// `g_ar_static_scratch` is only a handle for the runtime, never an address a kernel may use;
// the real address comes from cudaGetSymbolAddress and is cached per device.
static int ar_static_scratch_ptr(float** out) {
  int dev = -1;
  KPMS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= KPMS_AR_STATIC_MAX_DEVICES) return KPMS_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_ar_static_mutex);
  if (g_ar_static_dev_ptr[dev] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_ar_static_scratch);
    if (e != cudaSuccess) return (int)e;
    if (p == nullptr) return KPMS_ERR_NULL;
    g_ar_static_dev_ptr[dev] = static_cast<float*>(p);
  }
  *out = g_ar_static_dev_ptr[dev];
  return KPMS_OK;
}

extern "C" int kpms_ar_loglik_static(const float* x, const double* Ab, const double* Q, int N,
                                     int T, int D, int L, int K, float* ll, int impl,
                                     void* stream) {
  KPMS_REQUIRE(x && Ab && Q && ll, KPMS_ERR_NULL);
  KPMS_REQUIRE(N > 0 && T > L && D > 0 && L > 0 && K > 0, KPMS_ERR_SHAPE);
  KPMS_REQUIRE(D <= AR_DMAX && kpms_dl_supported(D, L), KPMS_ERR_UNSUPPORTED);
  const int PP = ar_pp(D, L);
  const size_t wfloats = kpms_round_up((size_t)K * D * PP, 64);  // keeps cst 256-byte aligned
  if (wfloats + (size_t)K > (size_t)KPMS_AR_STATIC_FLOATS) return KPMS_ERR_TOO_LARGE;
  float* base = nullptr;
  int rc = ar_static_scratch_ptr(&base);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  float* Wf = base;
  float* cst = base + wfloats;
  ar_prepare_kernel<<<K, 32, 0, st>>>(Ab, Q, D, L, PP, Wf, cst);
  rc = kpms_last_launch();
  if (rc) return rc;
  if (impl == 0) {
    // the tensor-core kernel wants room for its packed TF32 tiles; what is left of the fixed
    // array after the fp32 weights is offered, and when that is too small (more than a few
    // dozen states) the CUDA-core kernel computes the same result
    float* tiles = cst + kpms_round_up((size_t)K, 64);
    const size_t used = (size_t)(tiles - base);
    const size_t left = used < (size_t)KPMS_AR_STATIC_FLOATS ? ((size_t)KPMS_AR_STATIC_FLOATS - used) * sizeof(float) : 0;
    rc = kpms_ar_loglik_tc(x, Wf, cst, N, T, D, L, K, ll, tiles, left, 0, 0, st);
    if (rc != KPMS_ERR_UNSUPPORTED) return rc;
  }
#define KPMS_X(d, l) \
  if (D == d && L == l) return launch_simt<d, l>(x, Wf, cst, N, T, K, ll, st);
  KPMS_FOR_EACH_DL(KPMS_X)
#undef KPMS_X
  return KPMS_ERR_UNSUPPORTED;
}
