// Sufficient-statistic kernels: AR second moments (a8), transition counts (a10), CRP
// table counts with sticky thinning (a11).
#include "common.cuh"

// ======================================================================== a8: AR statistics
// Reference: gibbs.py:129-131 (S_out_out, S_out_in, S_in_in), :141-143 (dense one-hot masks).
// (Fallback path; the production kernel is ar_stats.cu.)  Instead of the reference's dense (K, frames) one-hot, CTA (part, k) scans its slice of z,
// compacts the frames whose state is k (deterministic block prefix-sum order), stages their
// regressors f = [x[t-L..t-1], 1, x[t]] in shared memory as fp64 and accumulates the lower
// triangle of sum f f^T with one thread per matrix entry.  Per-part partials are combined in
// fixed order by a second kernel, so the result is bit-reproducible.
#define ARS_THREADS 896
#define ARS_BATCH 64
#define ARS_ACC 3

static inline int ars_parts(long long frames, int K) {
  long long want = (4LL * 148 * 2 + K - 1) / K;
  long long maxp = (frames + ARS_THREADS - 1) / ARS_THREADS;
  long long p = want < 1 ? 1 : want;
  if (p > 64) p = 64;
  if (p > maxp) p = maxp;
  if (p < 1) p = 1;
  return (int)p;
}

size_t kpms_ar_stats_scan_workspace_bytes(int N, int T, int D, int L, int K) {
  if (N <= 0 || T <= L || D <= 0 || L <= 0 || K <= 0) return 0;
  const int P = D * L + 1 + D;
  const int parts = ars_parts((long long)N * (T - L), K);
  return kpms_round_up((size_t)parts * K * (P * (P + 1) / 2) * sizeof(double), 256);
}

__global__ void __launch_bounds__(ARS_THREADS)
    ar_stats_kernel(const float* __restrict__ x, const int* __restrict__ z,
                    const float* __restrict__ mask, int N, int T, int D, int L,
                    long long per_part, double* __restrict__ partial) {
  const int P = D * L + 1 + D, DL = D * L, Tm = T - L, ntri = P * (P + 1) / 2;
  const int part = blockIdx.x, k = blockIdx.y, tid = threadIdx.x, lane = tid & 31,
            warp = tid >> 5;
  constexpr int NWARP = ARS_THREADS / 32;
  extern __shared__ __align__(16) unsigned char sm_raw[];
  double* Fs = reinterpret_cast<double*>(sm_raw);                       // ARS_BATCH * P
  long long* sel = reinterpret_cast<long long*>(Fs + ARS_BATCH * P);    // ARS_THREADS
  int* wc = reinterpret_cast<int*>(sel + ARS_THREADS);                  // NWARP
  int ei[ARS_ACC], ej[ARS_ACC];
  double acc[ARS_ACC];
#pragma unroll
  for (int a = 0; a < ARS_ACC; a++) {
    acc[a] = 0.0;
    int e = tid + a * ARS_THREADS;
    ei[a] = -1;
    ej[a] = 0;
    if (e < ntri) {  // e = i(i+1)/2 + j, j <= i
      int i = (int)((sqrt(8.0 * (double)e + 1.0) - 1.0) * 0.5);
      while (i * (i + 1) / 2 > e) i--;
      while ((i + 1) * (i + 2) / 2 <= e) i++;
      ei[a] = i;
      ej[a] = e - i * (i + 1) / 2;
    }
  }
  const long long total = (long long)N * Tm;
  const long long g0 = (long long)part * per_part;
  const long long g1 = (g0 + per_part < total) ? g0 + per_part : total;
  for (long long base = g0; base < g1; base += ARS_THREADS) {
    const long long g = base + tid;
    bool pred = false;
    if (g < g1) {
      const int n = (int)(g / Tm), tp = (int)(g % Tm);
      pred = (z[g] == k) && (mask[(size_t)n * T + tp + L] > 0.0f);
    }
    const unsigned bal = __ballot_sync(0xffffffffu, pred);
    if (lane == 0) wc[warp] = __popc(bal);
    __syncthreads();
    int off = 0, cnt = 0;
    for (int w = 0; w < NWARP; w++) {
      const int c = wc[w];
      if (w < warp) off += c;
      cnt += c;
    }
    if (pred) sel[off + __popc(bal & ((1u << lane) - 1u))] = g;
    __syncthreads();
    for (int b0 = 0; b0 < cnt; b0 += ARS_BATCH) {
      const int nb = min(ARS_BATCH, cnt - b0);
      for (int idx = tid; idx < nb * P; idx += ARS_THREADS) {
        const int b = idx / P, p = idx % P;
        const long long gg = sel[b0 + b];
        const int n = (int)(gg / Tm), tp = (int)(gg % Tm);
        const float* win = x + ((size_t)n * T + tp) * D;  // x[tp .. tp+L], contiguous
        double v;
        if (p < DL) v = (double)win[p];
        else if (p == DL) v = 1.0;
        else v = (double)win[p - 1];
        Fs[b * P + p] = v;
      }
      __syncthreads();
#pragma unroll
      for (int a = 0; a < ARS_ACC; a++) {
        if (ei[a] >= 0) {
          double s = acc[a];
          const int i = ei[a], j = ej[a];
          for (int b = 0; b < nb; b++) s = fma(Fs[b * P + i], Fs[b * P + j], s);
          acc[a] = s;
        }
      }
      __syncthreads();
    }
  }
  double* out = partial + ((size_t)part * gridDim.y + k) * ntri;
#pragma unroll
  for (int a = 0; a < ARS_ACC; a++)
    if (ei[a] >= 0) out[tid + a * ARS_THREADS] = acc[a];
}

__global__ void ar_stats_reduce_kernel(const double* __restrict__ partial, int parts, int K, int P,
                                       double* __restrict__ F) {
  const int ntri = P * (P + 1) / 2;
  const int k = blockIdx.x;
  for (int e = threadIdx.x; e < ntri; e += blockDim.x) {
    double s = 0.0;
    for (int p = 0; p < parts; p++) s += partial[((size_t)p * K + k) * ntri + e];
    int i = (int)((sqrt(8.0 * (double)e + 1.0) - 1.0) * 0.5);
    while (i * (i + 1) / 2 > e) i--;
    while ((i + 1) * (i + 2) / 2 <= e) i++;
    const int j = e - i * (i + 1) / 2;
    F[((size_t)k * P + i) * P + j] = s;
    F[((size_t)k * P + j) * P + i] = s;
  }
}

// Generic fallback of kpms_ar_stats (ar_stats.cu) for regressor lengths beyond its DMMA tiling.
int kpms_ar_stats_scan(const float* x, const int32_t* z, const float* mask, int N, int T,
                       int D, int L, int K, double* F, void* workspace,
                       size_t workspace_bytes, void* stream) {
  KPMS_REQUIRE(x && z && mask && F && workspace, KPMS_ERR_NULL);
  KPMS_REQUIRE(N > 0 && T > L && D > 0 && L > 0 && K > 0, KPMS_ERR_SHAPE);
  const int P = D * L + 1 + D;
  KPMS_REQUIRE(P * (P + 1) / 2 <= ARS_ACC * ARS_THREADS && K <= 65535, KPMS_ERR_UNSUPPORTED);
  KPMS_REQUIRE(kpms_aligned(workspace, 256) &&
                   workspace_bytes >= kpms_ar_stats_scan_workspace_bytes(N, T, D, L, K),
               KPMS_ERR_WORKSPACE);
  cudaStream_t st = (cudaStream_t)stream;
  const long long frames = (long long)N * (T - L);
  const int parts = ars_parts(frames, K);
  long long per_part = (frames + parts - 1) / parts;
  per_part = (per_part + ARS_THREADS - 1) / ARS_THREADS * ARS_THREADS;
  const size_t smem = (size_t)ARS_BATCH * P * sizeof(double) + ARS_THREADS * sizeof(long long) +
                      (ARS_THREADS / 32) * sizeof(int);
  KPMS_CUDA(cudaFuncSetAttribute(ar_stats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
  ar_stats_kernel<<<dim3(parts, K), ARS_THREADS, smem, st>>>(x, z, mask, N, T, D, L, per_part,
                                                            (double*)workspace);
  int rc = kpms_last_launch();
  if (rc) return rc;
  ar_stats_reduce_kernel<<<K, 256, 0, st>>>((const double*)workspace, parts, K, P, F);
  return kpms_last_launch();
}

// ======================================================================== a10: transition counts
// Reference: util.py:178-189.  Integer histogram: shared-memory privatised (int32) when
// K*K*4 fits, flushed with 64-bit global atomics; order-independent, hence bit-exact.
__global__ void count_transitions_kernel(const int* __restrict__ z, const float* __restrict__ mask,
                                         long long mask_stride, int N, int Tm, int K, int use_smem,
                                         unsigned long long* __restrict__ counts) {
  extern __shared__ int hist[];
  const int KK = K * K;
  if (use_smem) {
    for (int i = threadIdx.x; i < KK; i += blockDim.x) hist[i] = 0;
    __syncthreads();
  }
  const long long pairs = (long long)N * (Tm - 1);
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < pairs;
       g += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(g / (Tm - 1)), t = (int)(g % (Tm - 1));
    const float* mk = mask + (size_t)n * mask_stride + t;
    if (mk[0] > 0.0f && mk[1] > 0.0f) {
      const int a = z[(size_t)n * Tm + t], b = z[(size_t)n * Tm + t + 1];
      if ((unsigned)a < (unsigned)K && (unsigned)b < (unsigned)K) {
        if (use_smem) atomicAdd(&hist[a * K + b], 1);
        else atomicAdd(&counts[a * K + b], 1ULL);
      }
    }
  }
  if (use_smem) {
    __syncthreads();
    for (int i = threadIdx.x; i < KK; i += blockDim.x) {
      const int c = hist[i];
      if (c) atomicAdd(&counts[i], (unsigned long long)c);
    }
  }
}

extern "C" int kpms_count_transitions(const int32_t* z, const float* mask, int64_t mask_stride,
                                      int N, int Tm, int K, int64_t* counts, void* stream) {
  KPMS_REQUIRE(z && mask && counts, KPMS_ERR_NULL);
  KPMS_REQUIRE(N > 0 && Tm > 0 && K > 0 && mask_stride >= Tm, KPMS_ERR_SHAPE);
  cudaStream_t st = (cudaStream_t)stream;
  KPMS_CUDA(cudaMemsetAsync(counts, 0, (size_t)K * K * sizeof(int64_t), st));
  if (Tm < 2) return KPMS_OK;
  const size_t smem = (size_t)K * K * sizeof(int);
  const int use_smem = smem <= 160 * 1024;
  if (use_smem)
    KPMS_CUDA(cudaFuncSetAttribute(count_transitions_kernel,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long pairs = (long long)N * (Tm - 1);
  long long blocks = (pairs + 255) / 256;
  if (blocks > 148 * 4) blocks = 148 * 4;
  count_transitions_kernel<<<(unsigned)blocks, 256, use_smem ? smem : 0, st>>>(
      z, mask, mask_stride, N, Tm, K, use_smem, (unsigned long long*)counts);
  return kpms_last_launch();
}

// ======================================================================== a11: CRP table counts
// Reference: transitions.py:7-21 (sample_crp_tablecounts) and :23-33 (sample_ms).
// One CTA per (i,j) cell; customer c of the cell is seated at a new table when
//   randseq[starts[i,j]+c] < alpha*beta_j / (c + alpha*beta_j)          (transitions.py:19-20)
// Diagonal cells are then thinned: ms_ii <- sum_{c<ms_ii} [u_thin[i,c] < rho_i],
//   rho_i = beta_i*alpha / (beta_i*alpha + kappa)                        (transitions.py:28-31)
// which is the Binomial(ms_ii, rho_i) of the reference written as its Bernoulli sum.
__global__ void crp_tablecounts_kernel(const long long* __restrict__ counts,
                                       const long long* __restrict__ starts,
                                       const double* __restrict__ betas, double alpha, double kappa,
                                       const float* __restrict__ randseq, long long randseq_len,
                                       const float* __restrict__ u_thin, int thin_cap, int K,
                                       long long* __restrict__ ms, long long* __restrict__ ms_raw,
                                       int* __restrict__ overflow) {
  const int cell = blockIdx.x, i = cell / K, j = cell % K;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ long long red[8];
  __shared__ long long m_sh;
  const long long c = counts[cell];
  long long m = 0;
  if (c > 0 && starts[cell] + c > randseq_len) {  // the caller's uniforms do not cover this cell
    if (tid == 0 && overflow) atomicOr(overflow, 2);
  } else if (c > 0) {
    const double cw = alpha * betas[j];
    const float* rs = randseq + starts[cell];
    for (long long q = tid; q < c; q += blockDim.x)
      m += ((double)rs[q] < cw / ((double)q + cw)) ? 1 : 0;
  }
  m = warp_sum_ll(m);
  if (lane == 0) red[warp] = m;
  __syncthreads();
  if (tid == 0) {
    long long s = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) s += red[w];
    m_sh = s;
    if (ms_raw) ms_raw[cell] = s;
    if (i != j) ms[cell] = s;
  }
  if (i != j) return;
  __syncthreads();
  const long long mraw = m_sh;
  const double rho = betas[i] * alpha / (betas[i] * alpha + kappa);
  long long lim = mraw;
  if (lim > thin_cap) {
    lim = thin_cap;
    if (tid == 0 && overflow) atomicOr(overflow, 1);
  }
  long long kept = 0;
  for (long long q = tid; q < lim; q += blockDim.x)
    kept += ((double)u_thin[(size_t)i * thin_cap + q] < rho) ? 1 : 0;
  kept = warp_sum_ll(kept);
  __syncthreads();
  if (lane == 0) red[warp] = kept;
  __syncthreads();
  if (tid == 0) {
    long long s = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) s += red[w];
    ms[cell] = s;
  }
}

extern "C" int kpms_crp_tablecounts(const int64_t* counts, const int64_t* starts,
                                    const double* betas, double alpha, double kappa,
                                    const float* randseq, int64_t randseq_len, const float* u_thin,
                                    int thin_cap, int K, int64_t* ms, int64_t* ms_raw, int32_t* overflow,
                                    void* stream) {
  KPMS_REQUIRE(counts && starts && betas && randseq && u_thin && ms, KPMS_ERR_NULL);
  KPMS_REQUIRE(K > 0 && thin_cap > 0 && randseq_len >= 0, KPMS_ERR_SHAPE);
  cudaStream_t st = (cudaStream_t)stream;
  if (overflow) KPMS_CUDA(cudaMemsetAsync(overflow, 0, sizeof(int32_t), st));
  crp_tablecounts_kernel<<<K * K, 256, 0, st>>>((const long long*)counts, (const long long*)starts,
                                                betas, alpha, kappa, randseq, (long long)randseq_len, u_thin,
                                                thin_cap, K,
                                                (long long*)ms, (long long*)ms_raw, overflow);
  return kpms_last_launch();
}

// Same table counts with the uniforms generated in the kernel: customer c of cell (i,j) uses
// uniform number starts[i,j]+c of the Philox stream (seed, stream_crp) — exactly the value
// kpms_fill_uniform would have written to randseq[starts[i,j]+c] — and thinning draw q of state i
// uses uniform number (i << 32) + q of stream (seed, stream_thin).  Nothing of the size of the data
// is materialised, the thinning has no capacity limit, and every rank of a sharded job computes
// identical counts from the all-reduced `counts` without exchanging anything.
__global__ void crp_tablecounts_philox_kernel(const long long* __restrict__ counts,
                                              const long long* __restrict__ starts,
                                              const double* __restrict__ betas, double alpha, double kappa,
                                              uint2 key, uint32_t sid_crp, uint32_t sid_thin, int K,
                                              long long* __restrict__ ms, long long* __restrict__ ms_raw) {
  const int cell = blockIdx.x, i = cell / K, j = cell % K;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ long long red[8];
  __shared__ long long m_sh;
  const long long c = counts[cell];
  long long m = 0;
  if (c > 0) {
    const double cw = alpha * betas[j];
    const unsigned long long e0 = (unsigned long long)starts[cell], e1 = e0 + (unsigned long long)c;
    for (unsigned long long G = (e0 >> 2) + tid; G <= ((e1 - 1) >> 2); G += blockDim.x) {
      const uint4 r = philox4x32_10(make_uint4((uint32_t)G, (uint32_t)(G >> 32), 0u, sid_crp), key);
      const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
      for (int q4 = 0; q4 < 4; q4++) {
        const unsigned long long e = G * 4ULL + q4;
        if (e >= e0 && e < e1) m += ((double)u01(w[q4]) < cw / ((double)(e - e0) + cw)) ? 1 : 0;
      }
    }
  }
  m = warp_sum_ll(m);
  if (lane == 0) red[warp] = m;
  __syncthreads();
  if (tid == 0) {
    long long s = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) s += red[w];
    m_sh = s;
    if (ms_raw) ms_raw[cell] = s;
    if (i != j) ms[cell] = s;
  }
  if (i != j) return;
  __syncthreads();
  const long long mraw = m_sh;
  const double rho = betas[i] * alpha / (betas[i] * alpha + kappa);
  long long kept = 0;
  const unsigned long long base = (unsigned long long)i << 32;
  for (unsigned long long G = tid; G * 4ULL < (unsigned long long)mraw; G += blockDim.x) {
    const unsigned long long GG = (base >> 2) + G;
    const uint4 r = philox4x32_10(make_uint4((uint32_t)GG, (uint32_t)(GG >> 32), 0u, sid_thin), key);
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int q4 = 0; q4 < 4; q4++)
      if ((long long)(G * 4ULL + q4) < mraw) kept += ((double)u01(w[q4]) < rho) ? 1 : 0;
  }
  kept = warp_sum_ll(kept);
  __syncthreads();
  if (lane == 0) red[warp] = kept;
  __syncthreads();
  if (tid == 0) {
    long long s = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) s += red[w];
    ms[cell] = s;
  }
}

extern "C" int kpms_crp_tablecounts_philox(const int64_t* counts, const int64_t* starts,
                                           const double* betas, double alpha, double kappa, uint64_t seed,
                                           uint64_t stream_crp, uint64_t stream_thin, int K, int64_t* ms,
                                           int64_t* ms_raw, void* stream) {
  KPMS_REQUIRE(counts && starts && betas && ms, KPMS_ERR_NULL);
  KPMS_REQUIRE(K > 0, KPMS_ERR_SHAPE);
  crp_tablecounts_philox_kernel<<<K * K, 256, 0, (cudaStream_t)stream>>>(
      (const long long*)counts, (const long long*)starts, betas, alpha, kappa, split_seed(seed),
      (uint32_t)stream_crp, (uint32_t)stream_thin, K, (long long*)ms, (long long*)ms_raw);
  return kpms_last_launch();
}
