// a2+a3: HMM forward filter / backward sampler.
// Reference: distributions.py:46-79 (sample_hmm_stateseq), vmapped at gibbs.py:111-114.
//
// Forward (distributions.py:51-60): scaled linear-space filter, uniform initial distribution,
// mask-blended carry, +1e-16 guard, marginal log-likelihood.
//   K <= 128, kernel F: one CTA per session (or per session chunk); thread j owns state j and keeps
//   column j of pi in REGISTERS; one __syncthreads per frame (the unnormalised message is
//   double-buffered in shared memory, every thread re-derives the normaliser from the same shared
//   values, max_j ll[t+2] is reduced two frames ahead); ll / mask / u are staged 16 frames ahead
//   with cp.async.  While running the dot product the thread also learns
//     T_j = sum_i a_i pi[i][j],  P_j = sum_{i<j} a_i pi[i][j]  and  a_j pi[j][j]:
//   IF the next state is j, the backward sampler keeps state j at this frame exactly when
//     P_j <= u_t T_j < P_j + a_j pi[j][j]          (inverse CDF, distributions.py:65).
//   One such "stay" bit per (frame, state) is written out.
//   Chunks: with chunk_len > 0 a session is cut into chunks run by independent CTAs (3 per SM),
//   each warmed up over `burn_in` earlier frames from the uniform distribution; the mismatch
//   between a warmed-up start and its predecessor's end is reported for verification.
// Backward (distributions.py:62-72), kernel B: one warp per session settles whole runs of frames
//   (masked, or stay bit set for the current next state) with one ballot and evaluates the full
//   inverse CDF z = min{ j : cumsum_j > u * total } only where the state changes.
//   K > 128: generic kernel, pi streamed from L2.
#include "common.cuh"
#include <stdlib.h>

bool kpms_hmm_big_supported(int Tm, int K, int chunk_len);  // hmm_big.cu
int kpms_hmm_big_kp(int K);
int kpms_hmm_big(const float* ll, const float* mask, long long mask_stride, const double* pi, const float* u,
                 const int* tv, int N, int Tm, int K, int Kl, int chunk_len, int burn_in, int bwd_burn, int nch,
                 float* pf, float* pT, float* alphan, double* llpart, float* bnd, int* carry_warm, int* carry_out,
                 int* z, int forward_only, int* repair_scratch, cudaStream_t st);

#define HMM_STAGE 16    // frames of ll / mask / u staged ahead by the forward pass
#define HMM_BRING 128   // frames of stay bits / u / mask staged ahead by the backward sampler

__host__ __device__ static inline int hmm_nchunks(int Tm, int chunk_len) {
  if (chunk_len <= 0 || chunk_len >= Tm) return 1;
  return (Tm + chunk_len - 1) / chunk_len;
}

struct HmmWs {
  float* alphan;
  float* pf;
  float* pT;
  unsigned* stay;
  double* llpart;
  float* bnd;   // [N][nch][2][K]
  int* tv;
  int* carry_warm;  // [N][nch] carry a backward chunk starts from (after its warm-up)
  int* carry_out;   // [N][nch] carry a backward chunk leaves behind
  float* Tbuf;    // a @ pi of the tensor-core forward pass (chunked mode only), tiled like alphan
  float* lltile;  // ll in the interleaved layout when the caller's is row-major
  // Message layout.  tiled == 0: alphan[n][t][j].  tiled == 1 (tensor-core forward pass): the
  // chunk chains u = n*nch + t/chunk_len are interleaved 128 at a time so that a CTA stepping 128
  // chains moves whole 512-byte lines:  float4 index (((u/128)*CL + t%CL) * K/4 + j/4) * 128 + u%128.
  int tiled, tile_cl, tile_nch;
  int Kl;  // layout width of the messages and of the boundary records: K rounded up to a multiple of 4
};

__device__ __forceinline__ size_t hmm_msg_base4(const HmmWs& ws, int n, int t, int Tm, int K) {
  // float4 index of chunk 0 of message (n, t); chunk kc lives kc * (tiled ? 128 : 1) further
  if (!ws.tiled) return ((size_t)n * Tm + t) * (K >> 2);
  const int cc = t / ws.tile_cl, r = t - cc * ws.tile_cl;
  const int u = n * ws.tile_nch + cc;
  return (((size_t)(u >> 7) * ws.tile_cl + r) * (ws.Kl >> 2)) * 128 + (u & 127);
}

static size_t hmm_layout(int N, int Tm, int K, int chunk_len, void* base, HmmWs* ws) {
  const int nch = hmm_nchunks(Tm, chunk_len);
  const int Kl = (K + 3) & ~3;
  size_t off = 0;
  char* b = (char*)base;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += kpms_round_up(bytes, 256);
    return b ? (void*)(b + o) : nullptr;
  };
  void* p;
  // chunked mode may use the tiled layout: units padded to 128, chunks to their full length
  const size_t msg = (nch > 1) ? (size_t)((N * (size_t)nch + 127) / 128 * 128) * chunk_len * Kl : (size_t)N * Tm * K;
  p = take(msg * sizeof(float)); if (ws) ws->alphan = (float*)p;
  const size_t KPP = (K > 128 && K <= 512) ? (size_t)kpms_hmm_big_kp(K) : (size_t)K;  // padded fp32 tables of hmm_big.cu
  p = take(KPP * KPP * sizeof(float)); if (ws) ws->pf = (float*)p;
  p = take(KPP * KPP * sizeof(float)); if (ws) ws->pT = (float*)p;
  p = take((size_t)N * Tm * 4 * sizeof(unsigned)); if (ws) ws->stay = (unsigned*)p;
  p = take((size_t)N * nch * sizeof(double)); if (ws) ws->llpart = (double*)p;
  p = take((size_t)N * nch * 2 * Kl * sizeof(float)); if (ws) ws->bnd = (float*)p;
  p = take((size_t)N * sizeof(int)); if (ws) ws->tv = (int*)p;
  p = take((size_t)N * nch * sizeof(int)); if (ws) ws->carry_warm = (int*)p;
  p = take((size_t)N * nch * sizeof(int)); if (ws) ws->carry_out = (int*)p;
  if (nch > 1 && K <= 104) {
    p = take(msg * sizeof(float)); if (ws) ws->Tbuf = (float*)p;
    p = take(msg * sizeof(float)); if (ws) ws->lltile = (float*)p;
  } else if (ws) {
    ws->Tbuf = nullptr;
    ws->lltile = nullptr;
  }
  if (ws) {
    ws->tiled = 0;
    ws->tile_cl = chunk_len;
    ws->tile_nch = nch;
    ws->Kl = Kl;
  }
  return off;
}

extern "C" size_t kpms_hmm_ffbs_workspace_bytes(int N, int Tm, int K, int chunk_len) {
  if (N <= 0 || Tm <= 0 || K <= 0) return 0;
  return hmm_layout(N, Tm, K, chunk_len, nullptr, nullptr);
}

// number of unmasked frames per session (trailing padding assumed for the warm-up start)
__global__ void hmm_tv_kernel(const float* __restrict__ mask, long long mask_stride, int Tm,
                              int* __restrict__ tv) {
  const int n = blockIdx.x;
  __shared__ int cnt;
  if (threadIdx.x == 0) cnt = 0;
  __syncthreads();
  int c = 0;
  for (int t = threadIdx.x; t < Tm; t += blockDim.x) c += mask[(size_t)n * mask_stride + t] > 0.0f;
  atomicAdd(&cnt, c);
  __syncthreads();
  if (threadIdx.x == 0) tv[n] = cnt;
}

bool kpms_hmm_tc_supported(int Tm, int K, int chunk_len);  // hmm_tc.cu
int kpms_hmm_mask_check(const float* mask, long long mask_stride, int Tm, const int* tv, int N,
                        unsigned* err_bits, cudaStream_t st);  // hmm_tc.cu
int kpms_hmm_forward_tc(const float* ll, int ll_tiled, float* ll_scratch, const float* mask,
                        long long mask_stride, const double* pi, const float* u, const int* tv, int N,
                        int Tm, int K, int chunk_len, int burn_in, int nch, float* a_out, float* T_out,
                        unsigned* stay, double* llpart, float* bnd, unsigned* err_bits, cudaStream_t st);

// ------------------------------------------------------------------ kernel F: forward filter
template <int KP>
__global__ void __launch_bounds__(KP, 3)
    hmm_forward_kernel(const float* __restrict__ ll, const float* __restrict__ mask,
                       long long mask_stride, const double* __restrict__ pi,
                       const float* __restrict__ u, int Tm, int K, int chunk_len, int burn_in,
                       double* __restrict__ loglik, HmmWs ws) {
  constexpr int NW = KP / 32;
  const int nch = hmm_nchunks(Tm, chunk_len);
  const int n = blockIdx.x / nch, cc = blockIdx.x % nch;
  const int j = threadIdx.x, lane = j & 31, warp = j >> 5;
  const int t0 = (nch == 1) ? 0 : cc * chunk_len;
  const int t1 = (nch == 1) ? Tm : min(Tm, t0 + chunk_len);
  const int tb = (cc > 0) ? max(0, min(t0, ws.tv[n]) - burn_in) : 0;
  __shared__ __align__(16) float a_s[2][KP];
  __shared__ float wmax_s[2][NW];
  __shared__ __align__(16) float stage[HMM_STAGE][KP];    // ll rows
  __shared__ float mstage[HMM_STAGE][KP];                 // per-thread copies of the mask
  __shared__ float ustage[HMM_STAGE][KP];                 // per-thread copies of u

  float pcol[KP];
#pragma unroll
  for (int i = 0; i < KP; i++) pcol[i] = (i < K && j < K) ? (float)pi[(size_t)i * K + j] : 0.0f;
  const float pjj = (j < K) ? (float)pi[(size_t)j * K + j] : 0.0f;

  const float* ll_n = ll + (size_t)n * Tm * K;
  float* al_n = ws.alphan + (size_t)n * Tm * K;
  const float* mk_n = mask + (size_t)n * mask_stride;
  const float* u_n = u + (size_t)n * Tm;
  unsigned* stay_n = ws.stay + (size_t)n * Tm * 4;

  const unsigned stage_sa = (unsigned)__cvta_generic_to_shared(&stage[0][0]);
  const unsigned mstage_sa = (unsigned)__cvta_generic_to_shared(&mstage[0][0]);
  const unsigned ustage_sa = (unsigned)__cvta_generic_to_shared(&ustage[0][0]);
  auto issue_f = [&](int t) {
    if (t < t1) {
      const int slot = t % HMM_STAGE;
      if (j < K) {
        const float* src = ll_n + (size_t)t * K + j;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(stage_sa + (unsigned)((slot * KP + j) * 4)), "l"(src));
      }
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(mstage_sa + (unsigned)((slot * KP + j) * 4)), "l"(mk_n + t));
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(ustage_sa + (unsigned)((slot * KP + j) * 4)), "l"(u_n + t));
    }
    asm volatile("cp.async.commit_group;\n" ::);
  };
#pragma unroll 1
  for (int i = 0; i < HMM_STAGE; i++) issue_f(tb + i);
  asm volatile("cp.async.wait_group %0;\n" ::"n"(HMM_STAGE - 2));
  {  // maxima of frames tb and tb+1; afterwards frame t+2 is reduced during frame t
    const float l0 = (j < K) ? stage[tb % HMM_STAGE][j] : -INFINITY;
    const float l1 = (j < K && tb + 1 < t1) ? stage[(tb + 1) % HMM_STAGE][j] : -INFINITY;
    const float wm0 = warp_max_f(l0), wm1 = warp_max_f(l1);
    if (lane == 0) {
      wmax_s[tb & 1][warp] = wm0;
      wmax_s[(tb + 1) & 1][warp] = wm1;
    }
  }
  __syncthreads();
  float in_pot = (j < K) ? 1.0f / (float)K : 0.0f;
  double logtot = 0.0;
#pragma unroll 1
  for (int t = tb; t < t1; t++) {
    if (t == t0 && cc > 0 && j < K) ws.bnd[((size_t)(n * nch + cc) * 2 + 0) * ws.Kl + j] = in_pot;
    const int buf = t & 1, slot = t % HMM_STAGE;
    float cmax = wmax_s[buf][0];
#pragma unroll
    for (int w = 1; w < NW; w++) cmax = fmaxf(cmax, wmax_s[buf][w]);
    asm volatile("cp.async.wait_group %0;\n" ::"n"(HMM_STAGE - 3));  // frames <= t+2 have landed
    const float llv = (j < K) ? stage[slot][j] : -INFINITY;
    const float m = mstage[slot][j];
    const float uu = ustage[slot][j];
    const float a = (j < K) ? in_pot * __expf(llv - cmax) : 0.0f;
    a_s[buf][j] = a;
    const float nxt2 = (j < K && t + 2 < t1) ? stage[(t + 2) % HMM_STAGE][j] : -INFINITY;
    issue_f(t + HMM_STAGE);  // refills the slot this thread has just read
    __syncthreads();
    // max of frame t+2, written after this frame's barrier into the slot this frame has just
    // finished reading and read two barriers later: the shuffle chain overlaps the FMA stream
    {
      const float wm = warp_max_f(nxt2);
      if (lane == 0) wmax_s[buf][warp] = wm;
    }
    // One pass over the shared message: total T_j, strictly-lower-index part P_j, normaliser.
    // Rows are split into 32-row segments; for warp w, segments below w lie entirely under the
    // diagonal (i < j), segments above entirely over it, and only segment w needs the i < j mask.
    float lo[4] = {0.f, 0.f, 0.f, 0.f}, hi[4] = {0.f, 0.f, 0.f, 0.f}, md[4] = {0.f, 0.f, 0.f, 0.f};
    float sm8[4] = {0.f, 0.f, 0.f, 0.f};
    const float4* av = reinterpret_cast<const float4*>(a_s[buf]);
#pragma unroll
    for (int sg = 0; sg < NW; sg++) {
      if (sg < warp) {
#pragma unroll
        for (int q = 8 * sg; q < 8 * sg + 8; q++) {
          const float4 v = av[q];
          lo[0] = fmaf(v.x, pcol[4 * q + 0], lo[0]);
          lo[1] = fmaf(v.y, pcol[4 * q + 1], lo[1]);
          lo[2] = fmaf(v.z, pcol[4 * q + 2], lo[2]);
          lo[3] = fmaf(v.w, pcol[4 * q + 3], lo[3]);
          sm8[0] += v.x; sm8[1] += v.y; sm8[2] += v.z; sm8[3] += v.w;
        }
      } else if (sg > warp) {
#pragma unroll
        for (int q = 8 * sg; q < 8 * sg + 8; q++) {
          const float4 v = av[q];
          hi[0] = fmaf(v.x, pcol[4 * q + 0], hi[0]);
          hi[1] = fmaf(v.y, pcol[4 * q + 1], hi[1]);
          hi[2] = fmaf(v.z, pcol[4 * q + 2], hi[2]);
          hi[3] = fmaf(v.w, pcol[4 * q + 3], hi[3]);
          sm8[0] += v.x; sm8[1] += v.y; sm8[2] += v.z; sm8[3] += v.w;
        }
      } else {
#pragma unroll
        for (int q = 8 * sg; q < 8 * sg + 8; q++) {
          const float4 v = av[q];
          const float p0 = v.x * pcol[4 * q + 0], p1 = v.y * pcol[4 * q + 1];
          const float p2 = v.z * pcol[4 * q + 2], p3 = v.w * pcol[4 * q + 3];
          hi[0] += p0; hi[1] += p1; hi[2] += p2; hi[3] += p3;     // full segment into the total
          md[0] += (4 * q + 0 < j) ? p0 : 0.0f;                   // masked part into P
          md[1] += (4 * q + 1 < j) ? p1 : 0.0f;
          md[2] += (4 * q + 2 < j) ? p2 : 0.0f;
          md[3] += (4 * q + 3 < j) ? p3 : 0.0f;
          sm8[0] += v.x; sm8[1] += v.y; sm8[2] += v.z; sm8[3] += v.w;
        }
      }
    }
    const float plo = (lo[0] + lo[1]) + (lo[2] + lo[3]);
    const float P = plo + ((md[0] + md[1]) + (md[2] + md[3]));
    const float tot = plo + ((hi[0] + hi[1]) + (hi[2] + hi[3]));
    const float norm = ((sm8[0] + sm8[1]) + (sm8[2] + sm8[3])) + 1e-16f;
    const float inv = __frcp_rn(norm);
    const bool store = t >= t0;
    {  // stay bit (unnormalised units throughout: only ratios matter)
      const float thr = uu * tot;
      const bool st = (j < K) && (P <= thr) && (P + a * pjj > thr);
      const unsigned word = __ballot_sync(0xffffffffu, st);
      if (store && lane == 0) stay_n[(size_t)t * 4 + warp] = word;
    }
    if (store && j < K) al_n[(size_t)t * K + j] = a * inv;
    if (m > 0.0f) in_pot = tot * inv;
    if (store && j == 0 && m > 0.0f) logtot += (double)__logf(norm) + (double)cmax;
  }
  if (nch > 1 && j < K) ws.bnd[((size_t)(n * nch + cc) * 2 + 1) * ws.Kl + j] = in_pot;
  if (j == 0) ws.llpart[(size_t)n * nch + cc] = logtot;
  asm volatile("cp.async.wait_group 0;\n" ::);
  (void)loglik;
}

// loglik[n] = sum of chunk partials (fixed order); largest mismatch between a warmed-up chunk
// start and its predecessor's end (max abs difference of the normalised potentials)
__global__ void hmm_finish_kernel(HmmWs ws, int nch, int K, int chunk_len, double* __restrict__ loglik,
                                  unsigned* __restrict__ err_bits) {
  const int n = blockIdx.x;
  if (threadIdx.x == 0 && loglik) {
    double s = 0.0;
    for (int c = 0; c < nch; c++) s += ws.llpart[(size_t)n * nch + c];
    loglik[n] = s;
  }
  if (!err_bits || nch < 2) return;
  float worst = 0.0f;
  const int tvn = ws.tv[n];
  for (int c = 0; c + 1 < nch; c++) {
    if ((c + 1) * chunk_len >= tvn) break;  // the next chunk starts inside the trailing padding
    const float* a = ws.bnd + ((size_t)(n * nch + c) * 2 + 1) * ws.Kl;
    const float* b = ws.bnd + ((size_t)(n * nch + c + 1) * 2 + 0) * ws.Kl;
    for (int e = threadIdx.x; e < K; e += blockDim.x) {
      const float d = fabsf(a[e] - b[e]);
      worst = (d != d) ? INFINITY : fmaxf(worst, d);  // a NaN boundary must fail the check, not vanish in fmax
    }
  }
  for (int o = 16; o > 0; o >>= 1) worst = fmaxf(worst, __shfl_xor_sync(0xffffffffu, worst, o));
  if ((threadIdx.x & 31) == 0) atomicMax(err_bits, __float_as_uint(worst));
}

// ------------------------------------------------------------------ kernel B: backward sampler
// One warp per session chunk; the warps of a CTA share pi^T.  z_t depends on z_{t+1}, but two
// backward chains driven by the SAME uniforms coalesce as soon as they pick the same state once,
// which sticky transitions make a matter of a few frames.  A chunk [t0, t1) therefore starts
// `burn_in` frames above t1 with an empty carry, discards those frames, and records the carry it
// arrives with at t1.  hmm_brepair_kernel then walks every session's chunks from the top and
// redoes, from the true carry, the few chunks whose warm-up had not coalesced: the result is
// exactly the sequential sampler's for the given messages.
#define HMM_BWARPS 8

// Walk frames ts, ts-1, ..., t0 of session n.  Frames >= t1 are a warm-up (nothing written);
// `s_next` is the carry to start from (-1: next_potential = ones, distributions.py:76).
// Returns the carry after frame t0; *carry_at_t1 receives the carry in use below t1.
template <int KP>
__device__ __forceinline__ int hmm_backward_walk(const float* __restrict__ mk_n, const float* __restrict__ u_n,
                                                 const unsigned* __restrict__ stay_n, int* __restrict__ z_n,
                                                 const HmmWs& ws, const float* piT, unsigned* bring, float* lt,
                                                 int n, int Tm, int K, int t0, int t1, int ts, int s_next,
                                                 int* carry_at_t1) {
  const int lane = threadIdx.x & 31;
  const float* al_n = ws.alphan + (size_t)n * Tm * K;
  // z_t ~ alphan_t * pi[:, z_{t+1}] by inverse CDF on u_t (distributions.py:62-72).  Lane l looks
  // at frame t-l: if that frame is masked (z = 0, carry unchanged) or its stay bit for the current
  // next state is set, the frame is settled; one ballot settles the longest such prefix, and only
  // the first unsettled frame pays for the full prefix sum over the K weights.
  constexpr int PER = KP / 32;  // states per lane in the slow path: lane*PER .. lane*PER+PER-1
  const unsigned bring_sa = (unsigned)__cvta_generic_to_shared(bring);
  auto issue_b = [&](int f0, int cnt) {  // frames f0, f0-1, ..., f0-cnt+1 (those >= t0)
    const int f = f0 - lane;
    if (lane < cnt && f >= t0) {
      const unsigned dst = bring_sa + (unsigned)((f % HMM_BRING) * 32);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(stay_n + (size_t)f * 4));
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(dst + 16), "l"(u_n + f));
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(dst + 20), "l"(mk_n + f));
    }
    asm volatile("cp.async.commit_group;\n" ::);
  };
#pragma unroll 1
  for (int b0 = 0; b0 < HMM_BRING / 32; b0++) issue_b(ts - 32 * b0, 32);
  int carry1 = s_next;  // the carry in use when there is no warm-up
  int t = ts;
  int lb = (ts >= t1) ? t1 : t0;  // warm-up phase first (frames >= t1, nothing written), then the chunk
#pragma unroll 1
  while (t >= t0) {
    if (t < lb) {  // warm-up finished: remember the carry this chunk starts from
      carry1 = s_next;
      lb = t0;
    }
    asm volatile("cp.async.wait_group 1;\n" ::);
    __syncwarp();
    const int f = t - lane;
    const bool valid = f >= lb;
    const bool keep = lb == t0;  // frames of the chunk proper are written
    const unsigned* e = bring + (valid ? (f % HMM_BRING) : 0) * 8;
    const bool masked = valid && !(__uint_as_float(e[5]) > 0.0f);
    const bool stayb = valid && s_next >= 0 && ((e[s_next >> 5] >> (s_next & 31)) & 1u);
    const unsigned bal = __ballot_sync(0xffffffffu, masked || stayb);
    const int r = (bal == 0xffffffffu) ? 32 : (__ffs(~bal) - 1);  // settled frames t, t-1, ..., t-r+1
    if (keep && lane < r) z_n[f] = masked ? 0 : s_next;
    int consumed = r;
    const int fb = t - r;  // first unsettled frame (unmasked by construction) if it exists
    if (r < 32 && fb >= lb) {
      const float ub = __uint_as_float(bring[(fb % HMM_BRING) * 8 + 4]);
      const size_t mb4 = ws.tiled ? hmm_msg_base4(ws, n, fb, Tm, K) : 0;
      float w[PER];
      float run = 0.0f;
      float al4[4] = {0.f, 0.f, 0.f, 0.f};
      if (PER == 4 && ws.tiled) {  // the lane's four states are one 16-byte piece of the tile
        if (lane * 4 < ws.Kl) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(ws.alphan) + mb4 + (size_t)lane * 128);
          al4[0] = v.x; al4[1] = v.y; al4[2] = v.z; al4[3] = v.w;
        }
      }
#pragma unroll
      for (int q = 0; q < PER; q++) {
        const int st = lane * PER + q;
        float al = 0.0f;
        if (PER == 4 && ws.tiled) al = al4[q & 3];
        else if (st < K) {
          if (ws.tiled) al = ws.alphan[(mb4 + (size_t)(st >> 2) * 128) * 4 + (st & 3)];
          else al = al_n[(size_t)fb * K + st];
        }
        const float np = (s_next < 0) ? 1.0f : piT[s_next * KP + st];
        run = fmaf(np, al, run);
        w[q] = run;  // inclusive prefix within the lane
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
      const float thr = ub * total;
      int cnt = 0;
#pragma unroll
      for (int q = 0; q < PER; q++) cnt += __popc(__ballot_sync(0xffffffffu, !(excl + w[q] > thr)));
      const int zt = min(cnt, K - 1);
      if (keep && lane == 0) z_n[fb] = zt;
      s_next = zt;
      consumed = r + 1;
    }
    __syncwarp();  // every lane is done with the consumed slots and with lt
    issue_b(t - HMM_BRING, consumed);
    t -= consumed;
  }
  asm volatile("cp.async.wait_group 0;\n" ::);
  __syncwarp();
  *carry_at_t1 = carry1;
  return s_next;
}

template <int KP>
__device__ __forceinline__ void hmm_load_piT(const double* __restrict__ pi, int K, float* piT) {
  for (int idx = threadIdx.x; idx < KP * KP; idx += blockDim.x) {
    int s = idx / KP, i = idx % KP;
    piT[idx] = (s < K && i < K) ? (float)pi[(size_t)i * K + s] : 0.0f;  // piT[s*KP+i] = pi[i][s]
  }
  __syncthreads();
}

template <int KP>
__global__ void __launch_bounds__(HMM_BWARPS * 32)
    hmm_backward_kernel(const float* __restrict__ mask, long long mask_stride,
                        const double* __restrict__ pi, const float* __restrict__ u, int N, int Tm, int K,
                        int chunk_len, int burn_in, int* __restrict__ z, HmmWs ws) {
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  extern __shared__ __align__(16) float piT[];  // [KP][KP]
  unsigned* bring = reinterpret_cast<unsigned*>(piT + KP * KP) + (size_t)wib * (HMM_BRING * 8 + 32);
  float* lt = reinterpret_cast<float*>(bring + HMM_BRING * 8);    // [32] lane totals (slow path)
  hmm_load_piT<KP>(pi, K, piT);
  const int nch = hmm_nchunks(Tm, chunk_len);
  const long long unit = (long long)blockIdx.x * HMM_BWARPS + wib;
  if (unit >= (long long)N * nch) return;
  const int n = (int)(unit / nch), cc = (int)(unit % nch);
  const int t0 = (nch == 1) ? 0 : cc * chunk_len;
  const int t1 = (nch == 1) ? Tm : min(Tm, t0 + chunk_len);
  const int ts = (cc == nch - 1) ? t1 - 1 : min(Tm, t1 + burn_in) - 1;  // first frame visited
  int carry1;
  const int out = hmm_backward_walk<KP>(mask + (size_t)n * mask_stride, u + (size_t)n * Tm,
                                        ws.stay + (size_t)n * Tm * 4, z + (size_t)n * Tm, ws, piT, bring, lt,
                                        n, Tm, K, t0, t1, ts, -1, &carry1);
  if (lane == 0) {
    ws.carry_warm[unit] = carry1;
    ws.carry_out[unit] = out;
  }
}

// One warp per session walks its chunks from the top: a chunk whose warm-up did not arrive at the
// carry the chunk above really left behind is sampled again from that carry (sequentially, its
// own frames only).  Usually nothing, or a handful of chunks, needs redoing.
template <int KP>
__global__ void __launch_bounds__(HMM_BWARPS * 32)
    hmm_brepair_kernel(const float* __restrict__ mask, long long mask_stride,
                       const double* __restrict__ pi, const float* __restrict__ u, int N, int Tm, int K,
                       int chunk_len, int* __restrict__ z, HmmWs ws) {
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  extern __shared__ __align__(16) float piT[];
  unsigned* bring = reinterpret_cast<unsigned*>(piT + KP * KP) + (size_t)wib * (HMM_BRING * 8 + 32);
  float* lt = reinterpret_cast<float*>(bring + HMM_BRING * 8);
  const int nch = hmm_nchunks(Tm, chunk_len);
  const int n = blockIdx.x * HMM_BWARPS + wib;
  // skip the table load when no session of this CTA has a mismatch
  int need = 0;
  if (n < N)
    for (int cc = lane; cc + 1 < nch; cc += 32)
      need |= __ldcg(&ws.carry_warm[(size_t)n * nch + cc]) != __ldcg(&ws.carry_out[(size_t)n * nch + cc + 1]);
  if (!__syncthreads_or(need)) return;
  hmm_load_piT<KP>(pi, K, piT);
  if (n >= N) return;
  for (int cc = nch - 2; cc >= 0; cc--) {
    const size_t unit = (size_t)n * nch + cc;
    const int want = __ldcg(&ws.carry_out[unit + 1]);  // L2: lane 0 may have just rewritten it
    if (__ldcg(&ws.carry_warm[unit]) == want) continue;  // warp-uniform
    const int t0 = cc * chunk_len, t1 = min(Tm, t0 + chunk_len);
    int carry1;
    const int out = hmm_backward_walk<KP>(mask + (size_t)n * mask_stride, u + (size_t)n * Tm,
                                          ws.stay + (size_t)n * Tm * 4, z + (size_t)n * Tm, ws, piT, bring,
                                          lt, n, Tm, K, t0, t1, t1 - 1, want, &carry1);
    __syncwarp();
    if (lane == 0) {
      ws.carry_warm[unit] = want;
      ws.carry_out[unit] = out;
    }
    __syncwarp();
    __threadfence_block();
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
static int launch_reg(const float* ll, int ll_tiled, const float* mask, long long ms, const double* pi,
                      const float* u, int N, int Tm, int K, int chunk_len, int burn_in, int* z,
                      double* loglik, float* boundary_err, HmmWs ws, cudaStream_t st) {
  const int nch = hmm_nchunks(Tm, chunk_len);
  if (nch > 1) {
    hmm_tv_kernel<<<N, 256, 0, st>>>(mask, ms, Tm, ws.tv);
    int rc = kpms_last_launch();
    if (rc) return rc;
  }
  if (boundary_err) KPMS_CUDA(cudaMemsetAsync(boundary_err, 0, sizeof(float), st));
  int rc = KPMS_ERR_UNSUPPORTED;
  if (nch > 1) {  // both chunked filters assume trailing padding; anything else is flagged as 1e30
    rc = kpms_hmm_mask_check(mask, ms, Tm, ws.tv, N, (unsigned*)boundary_err, st);
    if (rc) return rc;
    rc = KPMS_ERR_UNSUPPORTED;
  }
  if (nch > 1)
    rc = kpms_hmm_forward_tc(ll, ll_tiled, ws.lltile, mask, ms, pi, u, ws.tv, N, Tm, K, chunk_len, burn_in, nch,
                             ws.alphan, ws.Tbuf, ws.stay, ws.llpart, ws.bnd, (unsigned*)boundary_err, st);
  if (rc == KPMS_ERR_UNSUPPORTED && ll_tiled) return KPMS_ERR_UNSUPPORTED;  // only the tensor-core pass reads tiles
  if (rc == KPMS_ERR_UNSUPPORTED) {
    hmm_forward_kernel<KP><<<N * nch, KP, 0, st>>>(ll, mask, ms, pi, u, Tm, K, chunk_len, burn_in, loglik, ws);
    rc = kpms_last_launch();
  } else {
    ws.tiled = 1;  // the tensor-core pass wrote interleaved messages
  }
  if (rc) return rc;
  hmm_finish_kernel<<<N, 128, 0, st>>>(ws, nch, K, chunk_len, loglik, (unsigned*)boundary_err);
  rc = kpms_last_launch();
  if (rc) return rc;
  if (kpms_opts.hmm_flags & 1) return KPMS_OK;  // debugging: forward pass only
  auto kern = hmm_backward_kernel<KP>;
  const size_t smem = (size_t)(KP * KP + HMM_BWARPS * (HMM_BRING * 8 + 32)) * sizeof(float);
  KPMS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long units = (long long)N * nch;
  const int bwd_burn = kpms_opts.hmm_bwd_burn;  // testing: warm-up of the sampler only (0 = every chunk repaired)
  kern<<<(unsigned)((units + HMM_BWARPS - 1) / HMM_BWARPS), HMM_BWARPS * 32, smem, st>>>(
      mask, ms, pi, u, N, Tm, K, chunk_len, bwd_burn >= 0 ? bwd_burn : burn_in, z, ws);
  rc = kpms_last_launch();
  if (rc || nch == 1) return rc;
  auto rep = hmm_brepair_kernel<KP>;
  KPMS_CUDA(cudaFuncSetAttribute(rep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  rep<<<(unsigned)((N + HMM_BWARPS - 1) / HMM_BWARPS), HMM_BWARPS * 32, smem, st>>>(mask, ms, pi, u, N, Tm, K,
                                                                                   chunk_len, z, ws);
  return kpms_last_launch();
}

// ll_tiled != 0: ll is in the interleaved layout of HmmWs for this (chunk_len); only valid when
// kpms_hmm_tc_supported(Tm, K, chunk_len).
int kpms_hmm_ffbs_impl(const float* ll, int ll_tiled, const float* mask, int64_t mask_stride,
                       const double* pi, const float* u, int N, int Tm, int K, int chunk_len,
                       int burn_in, int32_t* z, double* loglik, float* boundary_err,
                       void* workspace, size_t workspace_bytes, void* stream) {
  KPMS_REQUIRE(ll && mask && pi && u && z && workspace, KPMS_ERR_NULL);
  KPMS_REQUIRE(N > 0 && Tm > 0 && K > 0 && mask_stride >= Tm, KPMS_ERR_SHAPE);
  KPMS_REQUIRE(K <= 1024, KPMS_ERR_UNSUPPORTED);
  KPMS_REQUIRE(chunk_len <= 0 || burn_in >= 0, KPMS_ERR_SHAPE);
  if (K > 128 && !kpms_hmm_big_supported(Tm, K, chunk_len)) chunk_len = 0;  // the generic kernel walks whole sessions
  // chunked mode reports what it could not verify through boundary_err: it is not optional there
  KPMS_REQUIRE(hmm_nchunks(Tm, chunk_len) == 1 || boundary_err != nullptr, KPMS_ERR_NULL);
  HmmWs ws;
  const size_t need = hmm_layout(N, Tm, K, chunk_len, workspace, &ws);
  KPMS_REQUIRE(kpms_aligned(workspace, 256) && workspace_bytes >= need, KPMS_ERR_WORKSPACE);
  cudaStream_t st = (cudaStream_t)stream;
#define KPMS_HMM_GO(KPV) \
  return launch_reg<KPV>(ll, ll_tiled, mask, mask_stride, pi, u, N, Tm, K, chunk_len, burn_in, z, loglik, boundary_err, ws, st)
  if (K <= 32) KPMS_HMM_GO(32);
  if (K <= 64) KPMS_HMM_GO(64);
  if (K <= 96) KPMS_HMM_GO(96);
  if (K <= 128) KPMS_HMM_GO(128);
#undef KPMS_HMM_GO
  if (boundary_err) KPMS_CUDA(cudaMemsetAsync(boundary_err, 0, sizeof(float), st));
  if (kpms_hmm_big_supported(Tm, K, chunk_len)) {
    // chunk-parallel, 32 chains per CTA (hmm_big.cu); same verification protocol as K <= 128
    const int nch = hmm_nchunks(Tm, chunk_len);
    hmm_tv_kernel<<<N, 256, 0, st>>>(mask, mask_stride, Tm, ws.tv);
    int rc = kpms_last_launch();
    if (rc) return rc;
    rc = kpms_hmm_mask_check(mask, mask_stride, Tm, ws.tv, N, (unsigned*)boundary_err, st);
    if (rc) return rc;
    rc = kpms_hmm_big(ll, mask, mask_stride, pi, u, ws.tv, N, Tm, K, ws.Kl, chunk_len, burn_in, kpms_opts.hmm_bwd_burn,
                      nch, ws.pf, ws.pT, ws.alphan, ws.llpart, ws.bnd, ws.carry_warm, ws.carry_out, z,
                      kpms_opts.hmm_flags & 1, (int*)ws.stay /* no stay words on this path: repair list */, st);
    if (rc) return rc;
    hmm_finish_kernel<<<N, 128, 0, st>>>(ws, nch, K, chunk_len, loglik, (unsigned*)boundary_err);
    return kpms_last_launch();
  }
  hmm_prep_pi_kernel<<<(K * K + 255) / 256, 256, 0, st>>>(pi, K, ws.pf, ws.pT);
  int rc = kpms_last_launch();
  if (rc) return rc;
  const int threads = (K + 31) / 32 * 32;
  const size_t smem = (size_t)(K + 64) * sizeof(float);
  hmm_ffbs_generic_kernel<<<N, threads, smem, st>>>(ll, mask, mask_stride, ws.pf, ws.pT, u, Tm, K, z,
                                                    loglik, ws.alphan);
  return kpms_last_launch();
}

extern "C" int kpms_hmm_ffbs(const float* ll, const float* mask, int64_t mask_stride,
                             const double* pi, const float* u, int N, int Tm, int K, int chunk_len,
                             int burn_in, int32_t* z, double* loglik, float* boundary_err,
                             void* workspace, size_t workspace_bytes, void* stream) {
  return kpms_hmm_ffbs_impl(ll, 0, mask, mask_stride, pi, u, N, Tm, K, chunk_len, burn_in, z, loglik,
                            boundary_err, workspace, workspace_bytes, stream);
}
