// a4: resample_stateseqs = emission log-likelihoods (a1) then forward-filter /
// backward-sample (a2+a3).  Reference: gibbs.py:107-115.
#include "common.cuh"
#include <stdlib.h>

int kpms_ar_loglik_impl(const float* x, const double* Ab, const double* Q, int N, int T, int D, int L,
                        int K, float* ll, void* workspace, size_t workspace_bytes, int impl,
                        int tile_cl, int tile_nch, int* wrote_tiled, void* stream);  // ar_loglik.cu
int kpms_hmm_ffbs_impl(const float* ll, int ll_tiled, const float* mask, int64_t mask_stride,
                       const double* pi, const float* u, int N, int Tm, int K, int chunk_len,
                       int burn_in, int32_t* z, double* loglik, float* boundary_err,
                       void* workspace, size_t workspace_bytes, void* stream);  // hmm_ffbs.cu
bool kpms_hmm_tc_native(int Tm, int K, int chunk_len);                          // hmm_tc.cu

// ll is (N, T-L, K) row-major, or — when the tensor-core filter will consume it — in the filter's
// chain-interleaved layout, whose chunks are padded to full length and whose chains to 128.
static size_t ss_ll_bytes(int N, int T, int L, int K, int chunk_len) {
  const int Tm = T - L;
  size_t elems = (size_t)N * Tm * K;
  if (chunk_len > 0 && chunk_len < Tm) {
    const size_t nch = (size_t)(Tm + chunk_len - 1) / chunk_len;
    const size_t tiled = ((size_t)N * nch + 127) / 128 * 128 * chunk_len * ((K + 3) & ~3);
    if (tiled > elems) elems = tiled;
  }
  return kpms_round_up(elems * sizeof(float), 256);
}

extern "C" size_t kpms_resample_stateseqs_workspace_bytes(int N, int T, int D, int L, int K,
                                                          int chunk_len) {
  if (N <= 0 || T <= L || D <= 0 || L <= 0 || K <= 0) return 0;
  return kpms_round_up(kpms_ar_loglik_workspace_bytes(K, D, L), 256) + ss_ll_bytes(N, T, L, K, chunk_len) +
         kpms_hmm_ffbs_workspace_bytes(N, T - L, K, chunk_len);
}

extern "C" int kpms_resample_stateseqs(const float* x, const float* mask, const double* Ab,
                                       const double* Q, const double* pi, const float* u, int N,
                                       int T, int D, int L, int K, int chunk_len, int burn_in,
                                       int32_t* z, double* loglik, float* boundary_err,
                                       void* workspace, size_t workspace_bytes, int impl,
                                       void* stream) {
  KPMS_REQUIRE(x && mask && Ab && Q && pi && u && z && workspace, KPMS_ERR_NULL);
  KPMS_REQUIRE(N > 0 && T > L && D > 0 && L > 0 && K > 0, KPMS_ERR_SHAPE);
  KPMS_REQUIRE(kpms_aligned(workspace, 256) &&
                   workspace_bytes >= kpms_resample_stateseqs_workspace_bytes(N, T, D, L, K, chunk_len),
               KPMS_ERR_WORKSPACE);
  char* base = (char*)workspace;
  const size_t w1 = kpms_round_up(kpms_ar_loglik_workspace_bytes(K, D, L), 256);
  const size_t w2 = ss_ll_bytes(N, T, L, K, chunk_len);
  float* ll = (float*)(base + w1);
  const int Tm = T - L;
  int tile_cl = 0, tile_nch = 0, tiled = 0;
  // Writing ll straight into the filter's interleaved layout makes the emission kernel gather its
  // lag windows from scattered frames, which costs more (6.5 vs 3.3 ms on the bench shard) than the
  // filter's own transposition pass (0.65 ms); opt in with kpms_set_option("ll_tiled", 1).
  if (kpms_opts.ll_tiled == 1 && kpms_hmm_tc_native(Tm, K, chunk_len)) {
    tile_cl = chunk_len;
    tile_nch = (Tm + chunk_len - 1) / chunk_len;
  }
  int rc = kpms_ar_loglik_impl(x, Ab, Q, N, T, D, L, K, ll, base, w1, impl, tile_cl, tile_nch, &tiled, stream);
  if (rc) return rc;
  // the reference hands the filter mask[:, nlags:] (gibbs.py:114)
  return kpms_hmm_ffbs_impl(ll, tiled, mask + L, (int64_t)T, pi, u, N, Tm, K, chunk_len, burn_in, z, loglik,
                            boundary_err, base + w1 + w2, workspace_bytes - w1 - w2, stream);
}
