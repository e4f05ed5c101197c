// a4: resample_stateseqs = emission log-likelihoods (a1) then forward-filter /
// backward-sample (a2+a3).  Reference: gibbs.py:107-115.
#include "common.cuh"

static size_t ss_ll_bytes(int N, int T, int L, int K) {
  return kpms_round_up((size_t)N * (T - L) * K * sizeof(float), 256);
}

extern "C" size_t kpms_resample_stateseqs_workspace_bytes(int N, int T, int D, int L, int K,
                                                          int chunk_len) {
  if (N <= 0 || T <= L || D <= 0 || L <= 0 || K <= 0) return 0;
  return kpms_round_up(kpms_ar_loglik_workspace_bytes(K, D, L), 256) + ss_ll_bytes(N, T, L, K) +
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
  const size_t w2 = ss_ll_bytes(N, T, L, K);
  float* ll = (float*)(base + w1);
  int rc = kpms_ar_loglik(x, Ab, Q, N, T, D, L, K, ll, base, w1, impl, stream);
  if (rc) return rc;
  // the reference hands the filter mask[:, nlags:] (gibbs.py:114)
  return kpms_hmm_ffbs(ll, mask + L, (int64_t)T, pi, u, N, T - L, K, chunk_len, burn_in, z, loglik,
                       boundary_err, base + w1 + w2, workspace_bytes - w1 - w2, stream);
}
