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
#define HMM_STAGE 16   // frames staged ahead by the backward sampler

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
  // ll rows (and the mask) are staged HMM_STAGE frames ahead into shared memory with cp.async:
  // thread j copies and later reads only its own elements, so the staging needs no barrier, and
  // unlike a register ring it does not serialise on the six scoreboard slots.
  float* stage = piT + KP * KP;              // [HMM_STAGE][KP]  ll rows (alphan rows in the backward pass)
  float* mstage = stage + HMM_STAGE * KP;    // [HMM_STAGE][KP]  per-thread copy of the mask
  float* um = mstage + HMM_STAGE * KP;       // [HMM_STAGE][2]   backward: u, mask
  float* lt = um + HMM_STAGE * 2;            // [32]             backward: lane totals
  const unsigned stage_sa = (unsigned)__cvta_generic_to_shared(stage);
  const unsigned mstage_sa = (unsigned)__cvta_generic_to_shared(mstage);
  const unsigned um_sa = (unsigned)__cvta_generic_to_shared(um);
  auto issue_f = [&](int t) {
    if (t < Tm) {
      const int slot = t % HMM_STAGE;
      if (j < K) {
        const float* src = ll_n + (size_t)t * K + j;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(stage_sa + (unsigned)((slot * KP + j) * 4)), "l"(src));
      }
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(mstage_sa + (unsigned)((slot * KP + j) * 4)), "l"(mk_n + t));
    }
    asm volatile("cp.async.commit_group;\n" ::);
  };
#pragma unroll 1
  for (int i = 0; i < HMM_STAGE; i++) issue_f(i);
  asm volatile("cp.async.wait_group %0;\n" ::"n"(HMM_STAGE - 2));
  {  // maxima of frames 0 and 1; afterwards frame t+2 is reduced during frame t
    const float l0 = (j < K) ? stage[j] : -INFINITY;
    const float l1 = (j < K && Tm > 1) ? stage[KP + j] : -INFINITY;
    const float wm0 = warp_max_f(l0), wm1 = warp_max_f(l1);
    if (lane == 0) {
      wmax_s[0][warp] = wm0;
      wmax_s[1][warp] = wm1;
    }
  }
  __syncthreads();
  float in_pot = (j < K) ? 1.0f / (float)K : 0.0f;
  double logtot = 0.0;
#pragma unroll 1
  for (int t = 0; t < Tm; t++) {
    const int buf = t & 1, slot = t % HMM_STAGE;
    float cmax = wmax_s[buf][0];
#pragma unroll
    for (int w = 1; w < NW; w++) cmax = fmaxf(cmax, wmax_s[buf][w]);
    asm volatile("cp.async.wait_group %0;\n" ::"n"(HMM_STAGE - 3));  // frames <= t+2 have landed
    const float llv = (j < K) ? stage[slot * KP + j] : -INFINITY;
    const float m = mstage[slot * KP + j];
    const float a = (j < K) ? in_pot * expf(llv - cmax) : 0.0f;
    a_s[buf][j] = a;
    const float nxt2 = (j < K && t + 2 < Tm) ? stage[((t + 2) % HMM_STAGE) * KP + j] : -INFINITY;
    issue_f(t + HMM_STAGE);  // refills the slot this thread has just read
    __syncthreads();
    // max of frame t+2, written after this frame's barrier into the slot this frame has just
    // finished reading and read two barriers later: the shuffle chain overlaps the FMA stream
    {
      const float wm = warp_max_f(nxt2);
      if (lane == 0) wmax_s[buf][warp] = wm;
    }
    // mat-vec and normaliser from the same shared values; 8 + 8 independent chains
    float acc[8], sm8[8];
#pragma unroll
    for (int q = 0; q < 8; q++) acc[q] = 0.0f, sm8[q] = 0.0f;
    const float4* av = reinterpret_cast<const float4*>(a_s[buf]);
#pragma unroll
    for (int q = 0; q < KP / 4; q++) {
      const float4 v = av[q];
      acc[(4 * q + 0) & 7] = fmaf(v.x, pcol[4 * q + 0], acc[(4 * q + 0) & 7]);
      acc[(4 * q + 1) & 7] = fmaf(v.y, pcol[4 * q + 1], acc[(4 * q + 1) & 7]);
      acc[(4 * q + 2) & 7] = fmaf(v.z, pcol[4 * q + 2], acc[(4 * q + 2) & 7]);
      acc[(4 * q + 3) & 7] = fmaf(v.w, pcol[4 * q + 3], acc[(4 * q + 3) & 7]);
      sm8[(4 * q + 0) & 7] += v.x;
      sm8[(4 * q + 1) & 7] += v.y;
      sm8[(4 * q + 2) & 7] += v.z;
      sm8[(4 * q + 3) & 7] += v.w;
    }
    const float acc0 = (acc[0] + acc[1]) + (acc[2] + acc[3]);
    const float acc1 = (acc[4] + acc[5]) + (acc[6] + acc[7]);
    const float s0 = (sm8[0] + sm8[1]) + (sm8[2] + sm8[3]);
    const float s1 = (sm8[4] + sm8[5]) + (sm8[6] + sm8[7]);
    const float norm = (s0 + s1) + 1e-16f;
    const float inv = 1.0f / norm;
    if (j < K) al_n[(size_t)t * K + j] = a * inv;
    if (m > 0.0f) in_pot = (acc0 + acc1) * inv;
    if (j == 0 && m > 0.0f) logtot += (double)__logf(norm) + (double)cmax;
  }
  if (j == 0 && loglik != nullptr) loglik[n] = logtot;
  asm volatile("cp.async.wait_group 0;\n" ::);
  __syncthreads();  // alphan (global) written by this CTA is visible to warp 0 below
  if (warp != 0) return;

  // ------------------------------------------------------------------ backward (warp 0)
  // alphan rows, u and mask are staged HMM_STAGE frames ahead into shared memory with
  // cp.async (a register ring would serialise on the 6 scoreboard slots); the prefix sum over
  // the 32 lane totals goes through shared memory (one store, 8 broadcast loads) because a
  // 5-step shuffle scan costs ~400 cycles of dependent latency per frame; the sampled state
  // is the number of states whose cumulative weight is <= u * total (4 ballots + popc).
  constexpr int PER = KP / 32;  // states per lane: lane owns states lane*PER .. lane*PER+PER-1
  static_assert(PER <= 4, "one 16-byte slot per lane");
  auto issue = [&](int t) {  // copy frame t into its slot (t >= 0), then close the group
    if (t >= 0) {
      const int slot = t % HMM_STAGE;
#pragma unroll
      for (int e = 0; e < PER; e++) {
        const int st = lane * PER + e;
        if (st < K) {
          const float* src = al_n + (size_t)t * K + st;
          const unsigned dst = stage_sa + (unsigned)((slot * KP + st) * 4);
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(dst), "l"(src));
        }
      }
      if (lane < 2) {
        const float* src = (lane == 0) ? (u_n + t) : (mk_n + t);
        const unsigned dst = um_sa + (unsigned)((slot * 2 + lane) * 4);
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(dst), "l"(src));
      }
    }
    asm volatile("cp.async.commit_group;\n" ::);
  };
  for (int e = lane; e < HMM_STAGE * KP; e += 32) stage[e] = 0.0f;  // padded states stay 0
  __syncwarp();
#pragma unroll 1
  for (int i = 0; i < HMM_STAGE; i++) issue(Tm - 1 - i);
  int s_next = -1;  // -1: next_potential = ones (distributions.py:76)
#pragma unroll 1
  for (int t = Tm - 1; t >= 0; t--) {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(HMM_STAGE - 1));
    __syncwarp();
    const int slot = t % HMM_STAGE;
    const float uu = um[slot * 2 + 0], mm = um[slot * 2 + 1];
    int zt = 0;
    if (mm > 0.0f) {
      float w[PER];
      float run = 0.0f;
#pragma unroll
      for (int e = 0; e < PER; e++) {
        const float np = (s_next < 0) ? 1.0f : piT[s_next * KP + lane * PER + e];
        run = fmaf(np, stage[slot * KP + lane * PER + e], run);
        w[e] = run;  // inclusive prefix within the lane
      }
      lt[lane] = run;
      __syncwarp();
      // exclusive prefix over lane totals: 8 group sums of 4 lanes, then the part of the own group
      const int grp = lane >> 2, sub = lane & 3;
      float excl = 0.0f, total = 0.0f;
#pragma unroll
      for (int q = 0; q < 8; q++) {
        const float4 v = reinterpret_cast<const float4*>(lt)[q];
        const float g4 = (v.x + v.y) + (v.z + v.w);
        const float own = ((sub > 0) ? v.x : 0.0f) + ((sub > 1) ? v.y : 0.0f) + ((sub > 2) ? v.z : 0.0f);
        excl += (q < grp) ? g4 : ((q == grp) ? own : 0.0f);
        total += g4;
      }
      const float thr = uu * total;
      int cnt = 0;
#pragma unroll
      for (int e = 0; e < PER; e++)
        cnt += __popc(__ballot_sync(0xffffffffu, !(excl + w[e] > thr)));
      zt = min(cnt, K - 1);
      s_next = zt;
    }
    if (lane == 0) z_n[t] = zt;
    __syncwarp();  // every lane is done with this slot and with lt
    issue(t - HMM_STAGE);
  }
  asm volatile("cp.async.wait_group 0;\n" ::);
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
  const size_t smem = (size_t)(KP * KP + 2 * HMM_STAGE * KP + HMM_STAGE * 2 + 32) * sizeof(float);
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
