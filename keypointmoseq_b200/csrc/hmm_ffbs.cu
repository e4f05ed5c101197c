// a2+a3: HMM forward filter / backward sampler, one CTA per session.
// Reference: distributions.py:46-79 (sample_hmm_stateseq), vmapped at gibbs.py:111-114.
//
// Forward (distributions.py:51-60): scaled linear-space filter, uniform initial
// distribution, mask-blended carry, +1e-16 guard, marginal log-likelihood.
//   K <= 128: thread j owns state j and keeps column j of pi in REGISTERS; one
//   __syncthreads per frame (the unnormalised message is double-buffered in shared memory,
//   every thread re-derives the normaliser from the same shared values, and max_j ll[t+1]
//   is reduced one frame ahead, off the critical path).  ll is prefetched 8 frames ahead.
//   K  > 128: generic kernel, pi streamed from L2.
// Backward (distributions.py:62-72): z_t ~ alphan_t * pi[:, z_{t+1}], inverse CDF on one
// uniform per frame; done by a single warp (K<=128) with shuffle scans, no block barrier.
#include "common.cuh"

#define HMM_RING 8

extern "C" size_t kpms_hmm_ffbs_workspace_bytes(int N, int Tm, int K) {
  if (N <= 0 || Tm <= 0 || K <= 0) return 0;
  size_t al = kpms_round_up((size_t)N * Tm * K * sizeof(float), 256);   // alphan
  size_t pf = kpms_round_up((size_t)K * K * sizeof(float), 256) * 2;    // pi, pi^T in f32 (K>128)
  return al + pf;
}

template <int KP>
__global__ void __launch_bounds__(KP)
    hmm_ffbs_reg_kernel(const float* __restrict__ ll, const float* __restrict__ mask,
                        long long mask_stride, const double* __restrict__ pi,
                        const float* __restrict__ u, int Tm, int K, int* __restrict__ z,
                        double* __restrict__ loglik, float* __restrict__ alphan) {
  constexpr int NW = KP / 32;
  const int n = blockIdx.x, j = threadIdx.x, lane = j & 31, warp = j >> 5;
  extern __shared__ __align__(16) float piT[];  // [KP][KP]: piT[s*KP+i] = pi[i][s]
  __shared__ __align__(16) float a_s[2][KP];
  __shared__ float wmax_s[2][NW];

  float pcol[KP];
#pragma unroll
  for (int i = 0; i < KP; i++) pcol[i] = (i < K && j < K) ? (float)pi[(size_t)i * K + j] : 0.0f;
  for (int idx = j; idx < KP * KP; idx += KP) {
    int s = idx / KP, i = idx % KP;
    piT[idx] = (s < K && i < K) ? (float)pi[(size_t)i * K + s] : 0.0f;
  }

  const float* ll_n = ll + (size_t)n * Tm * K;
  float* al_n = alphan + (size_t)n * Tm * K;
  const float* mk_n = mask + (size_t)n * mask_stride;
  const float* u_n = u + (size_t)n * Tm;
  int* z_n = z + (size_t)n * Tm;

  // ------------------------------------------------------------------ forward
  float ring[HMM_RING];
#pragma unroll
  for (int i = 0; i < HMM_RING; i++)
    ring[i] = (j < K && i < Tm) ? ll_n[(size_t)i * K + j] : -INFINITY;
  {
    float wm = warp_max_f(ring[0]);
    if (lane == 0) wmax_s[0][warp] = wm;
  }
  __syncthreads();
  float in_pot = (j < K) ? 1.0f / (float)K : 0.0f;
  double logtot = 0.0;
  for (int tb = 0; tb < Tm; tb += HMM_RING) {
#pragma unroll
    for (int i = 0; i < HMM_RING; i++) {
      const int t = tb + i;
      if (t < Tm) {  // uniform
        const int buf = t & 1;
        float cmax = wmax_s[buf][0];
#pragma unroll
        for (int w = 1; w < NW; w++) cmax = fmaxf(cmax, wmax_s[buf][w]);
        const float llv = ring[i];
        const float a = (j < K) ? in_pot * expf(llv - cmax) : 0.0f;
        a_s[buf][j] = a;
        const float m = mk_n[t];
        // next frame's max, one frame ahead (ring[(i+1)%RING] already holds ll[t+1])
        {
          const float nxt = ring[(i + 1) % HMM_RING];
          float wm = warp_max_f(nxt);
          if (lane == 0) wmax_s[buf ^ 1][warp] = wm;
        }
        // refill this ring slot with ll[t+RING]
        ring[i] = (j < K && t + HMM_RING < Tm) ? ll_n[(size_t)(t + HMM_RING) * K + j] : -INFINITY;
        __syncthreads();
        // normaliser: every warp reduces the KP shared values itself (identical order in
        // every warp, so all threads agree bit-for-bit); the shuffles overlap the FMA stream
        float part = 0.0f;
#pragma unroll
        for (int q = 0; q < NW; q++) part += a_s[buf][lane + 32 * q];
        float acc[8];
#pragma unroll
        for (int q = 0; q < 8; q++) acc[q] = 0.0f;
        const float4* av = reinterpret_cast<const float4*>(a_s[buf]);
#pragma unroll
        for (int q = 0; q < KP / 4; q++) {
          const float4 v = av[q];
          acc[(4 * q + 0) & 7] = fmaf(v.x, pcol[4 * q + 0], acc[(4 * q + 0) & 7]);
          acc[(4 * q + 1) & 7] = fmaf(v.y, pcol[4 * q + 1], acc[(4 * q + 1) & 7]);
          acc[(4 * q + 2) & 7] = fmaf(v.z, pcol[4 * q + 2], acc[(4 * q + 2) & 7]);
          acc[(4 * q + 3) & 7] = fmaf(v.w, pcol[4 * q + 3], acc[(4 * q + 3) & 7]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
        const float acc0 = (acc[0] + acc[1]) + (acc[2] + acc[3]);
        const float acc1 = (acc[4] + acc[5]) + (acc[6] + acc[7]);
        const float s0 = part, s1 = 0.0f;
        const float norm = (s0 + s1) + 1e-16f;
        const float inv = 1.0f / norm;
        if (j < K) al_n[(size_t)t * K + j] = a * inv;
        if (m > 0.0f) in_pot = (acc0 + acc1) * inv;
        if (j == 0 && m > 0.0f) logtot += (double)logf(norm) + (double)cmax;
      }
    }
  }
  if (j == 0 && loglik != nullptr) loglik[n] = logtot;
  __syncthreads();  // alphan (global) written by this CTA is visible to warp 0 below
  if (warp != 0) return;

  // ------------------------------------------------------------------ backward (warp 0)
  constexpr int PER = KP / 32;  // states per lane: lane owns states lane*PER .. lane*PER+PER-1
  int s_next = -1;              // -1: next_potential = ones (distributions.py:76)
  float aring[HMM_RING][PER];
  float uring[HMM_RING], mring[HMM_RING];
#pragma unroll
  for (int i = 0; i < HMM_RING; i++) {
    const int t = Tm - 1 - i;
#pragma unroll
    for (int e = 0; e < PER; e++) {
      const int st = lane * PER + e;
      aring[i][e] = (t >= 0 && st < K) ? al_n[(size_t)t * K + st] : 0.0f;
    }
    uring[i] = (t >= 0) ? u_n[t] : 0.0f;
    mring[i] = (t >= 0) ? mk_n[t] : 0.0f;
  }
  for (int tb = Tm - 1; tb >= 0; tb -= HMM_RING) {
#pragma unroll
    for (int i = 0; i < HMM_RING; i++) {
      const int t = tb - i;
      if (t >= 0) {
        int zt = 0;
        if (mring[i] > 0.0f) {
          float w[PER];
          float run = 0.0f;
#pragma unroll
          for (int e = 0; e < PER; e++) {
            const float np = (s_next < 0) ? 1.0f : piT[s_next * KP + lane * PER + e];
            run += np * aring[i][e];
            w[e] = run;  // inclusive prefix within the lane
          }
          float incl = run;  // inclusive scan of lane totals
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            float v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
          }
          const float total = __shfl_sync(0xffffffffu, incl, 31);
          const float thr = uring[i] * total;
          const float excl = incl - run;
          const unsigned bal = __ballot_sync(0xffffffffu, incl > thr);
          int first = K - 1;
          if (bal != 0u) {
            const int fl = __ffs(bal) - 1;
            int cand = K - 1;
            if (lane == fl) {
              cand = lane * PER + PER - 1;
#pragma unroll
              for (int e = PER - 1; e >= 0; e--)
                if (excl + w[e] > thr) cand = lane * PER + e;
            }
            first = __shfl_sync(0xffffffffu, cand, fl);
          }
          zt = min(first, K - 1);
          s_next = zt;
        }
        if (lane == 0) z_n[t] = zt;
        // refill
        const int tn = t - HMM_RING;
#pragma unroll
        for (int e = 0; e < PER; e++) {
          const int st = lane * PER + e;
          aring[i][e] = (tn >= 0 && st < K) ? al_n[(size_t)tn * K + st] : 0.0f;
        }
        uring[i] = (tn >= 0) ? u_n[tn] : 0.0f;
        mring[i] = (tn >= 0) ? mk_n[tn] : 0.0f;
      }
    }
  }
}

// Generic kernel for K > 128 (K <= 1024): thread j owns state j, pi / pi^T in fp32 in global
// memory (L2-resident), two block barriers per frame.
__global__ void hmm_prep_pi_kernel(const double* __restrict__ pi, int K, float* __restrict__ pf,
                                   float* __restrict__ pT) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= K * K) return;
  const int i = idx / K, s = idx % K;
  const float v = (float)pi[idx];
  pf[idx] = v;
  pT[(size_t)s * K + i] = v;
}

__global__ void hmm_ffbs_generic_kernel(const float* __restrict__ ll, const float* __restrict__ mask,
                                        long long mask_stride, const float* __restrict__ pf,
                                        const float* __restrict__ pT, const float* __restrict__ u,
                                        int Tm, int K, int* __restrict__ z,
                                        double* __restrict__ loglik, float* __restrict__ alphan) {
  const int n = blockIdx.x, j = threadIdx.x, lane = j & 31, warp = j >> 5;
  const int nw = blockDim.x >> 5;
  extern __shared__ __align__(16) float smg[];
  float* a_s = smg;            // K
  float* red = smg + K;        // 32
  float* red2 = red + 32;      // 32
  __shared__ int s_pick;
  const float* ll_n = ll + (size_t)n * Tm * K;
  float* al_n = alphan + (size_t)n * Tm * K;
  const float* mk_n = mask + (size_t)n * mask_stride;
  const float* u_n = u + (size_t)n * Tm;
  int* z_n = z + (size_t)n * Tm;
  float in_pot = (j < K) ? 1.0f / (float)K : 0.0f;
  double logtot = 0.0;
  for (int t = 0; t < Tm; t++) {
    const float llv = (j < K) ? ll_n[(size_t)t * K + j] : -INFINITY;
    float wm = warp_max_f(llv);
    if (lane == 0) red[warp] = wm;
    __syncthreads();
    float cmax = red[0];
    for (int w = 1; w < nw; w++) cmax = fmaxf(cmax, red[w]);
    const float a = (j < K) ? in_pot * expf(llv - cmax) : 0.0f;
    if (j < K) a_s[j] = a;
    __syncthreads();
    float acc = 0.0f, sum = 0.0f;
    if (j < K) {
      for (int i = 0; i < K; i++) {
        const float v = a_s[i];
        acc = fmaf(v, pf[(size_t)i * K + j], acc);
        sum += v;
      }
    }
    const float m = mk_n[t];
    // every thread j < K holds the same `sum` (same values, same order)
    const float norm = sum + 1e-16f;
    const float inv = 1.0f / norm;
    if (j < K) {
      al_n[(size_t)t * K + j] = a * inv;
      if (m > 0.0f) in_pot = acc * inv;
    }
    if (j == 0 && m > 0.0f) logtot += (double)logf(norm) + (double)cmax;
    __syncthreads();
  }
  if (j == 0 && loglik != nullptr) loglik[n] = logtot;
  // backward
  int s_next = -1;
  for (int t = Tm - 1; t >= 0; t--) {
    int zt = 0;
    if (mk_n[t] > 0.0f) {
      float w = 0.0f;
      if (j < K) w = ((s_next < 0) ? 1.0f : pT[(size_t)s_next * K + j]) * al_n[(size_t)t * K + j];
      float incl = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        float v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
      }
      if (lane == 31) red[warp] = incl;
      if (j == 0) s_pick = K - 1;
      __syncthreads();
      float off = 0.0f, total = 0.0f;
      for (int w2 = 0; w2 < nw; w2++) {
        const float v = red[w2];
        if (w2 < warp) off += v;
        total += v;
      }
      const float thr = u_n[t] * total;
      if (j < K && off + incl > thr) atomicMin(&s_pick, j);
      __syncthreads();
      zt = s_pick;
      s_next = zt;
      __syncthreads();
    }
    if (j == 0) z_n[t] = zt;
  }
  (void)red2;
}

template <int KP>
static int launch_reg(const float* ll, const float* mask, long long ms, const double* pi,
                      const float* u, int N, int Tm, int K, int* z, double* loglik, float* alphan,
                      cudaStream_t st) {
  auto kern = hmm_ffbs_reg_kernel<KP>;
  const size_t smem = (size_t)KP * KP * sizeof(float);
  KPMS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<N, KP, smem, st>>>(ll, mask, ms, pi, u, Tm, K, z, loglik, alphan);
  return kpms_last_launch();
}

extern "C" int kpms_hmm_ffbs(const float* ll, const float* mask, int64_t mask_stride,
                             const double* pi, const float* u, int N, int Tm, int K, int32_t* z,
                             double* loglik, void* workspace, size_t workspace_bytes,
                             void* stream) {
  KPMS_REQUIRE(ll && mask && pi && u && z && workspace, KPMS_ERR_NULL);
  KPMS_REQUIRE(N > 0 && Tm > 0 && K > 0 && mask_stride >= Tm, KPMS_ERR_SHAPE);
  KPMS_REQUIRE(K <= 1024, KPMS_ERR_UNSUPPORTED);
  KPMS_REQUIRE(kpms_aligned(workspace, 256) &&
                   workspace_bytes >= kpms_hmm_ffbs_workspace_bytes(N, Tm, K),
               KPMS_ERR_WORKSPACE);
  cudaStream_t st = (cudaStream_t)stream;
  float* alphan = (float*)workspace;
  if (K <= 32) return launch_reg<32>(ll, mask, mask_stride, pi, u, N, Tm, K, z, loglik, alphan, st);
  if (K <= 64) return launch_reg<64>(ll, mask, mask_stride, pi, u, N, Tm, K, z, loglik, alphan, st);
  if (K <= 96) return launch_reg<96>(ll, mask, mask_stride, pi, u, N, Tm, K, z, loglik, alphan, st);
  if (K <= 128) return launch_reg<128>(ll, mask, mask_stride, pi, u, N, Tm, K, z, loglik, alphan, st);
  char* base = (char*)workspace + kpms_round_up((size_t)N * Tm * K * sizeof(float), 256);
  float* pf = (float*)base;
  float* pT = (float*)(base + kpms_round_up((size_t)K * K * sizeof(float), 256));
  hmm_prep_pi_kernel<<<(K * K + 255) / 256, 256, 0, st>>>(pi, K, pf, pT);
  int rc = kpms_last_launch();
  if (rc) return rc;
  const int threads = (K + 31) / 32 * 32;
  const size_t smem = (size_t)(K + 64) * sizeof(float);
  hmm_ffbs_generic_kernel<<<N, threads, smem, st>>>(ll, mask, mask_stride, pf, pT, u, Tm, K, z,
                                                    loglik, alphan);
  return kpms_last_launch();
}
