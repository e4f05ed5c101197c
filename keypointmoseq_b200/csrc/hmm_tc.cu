// HMM forward filter on the 5th-generation tensor cores (tcgen05, 3xTF32, TMEM accumulators).
// Reference: distributions.py:51-60 (_forward_message) — per frame  in_pot' = (a / sum a) @ pi,
// a = in_pot * exp(ll - max ll).
//
// A single session is a chain of K x K mat-vecs, which no tensor core can use.  Cutting every
// session into chunks with a warm-up (hmm_ffbs.cu) yields thousands of independent chains, and 128
// chains stepping in lock-step make the per-frame work a real GEMM:
//     T (128 chains x K) = A (128 x K: the chains' current messages a) @ pi (K x K)
// Kernel hmm_forward_tc: one CTA = 128 chains, 512 threads: thread = (chain, quarter of the states).
// pi (hi / lo TF32 parts) stays resident in shared memory in the UMMA K-major layout; per frame a
// thread reads its part of T from TMEM, multiplies by exp(ll - row max) (row max and normaliser are
// exchanged between the four parts through shared memory), accumulates the marginal log-likelihood,
// writes the message (global, interleaved layout, for the sampler) and its hi / lo split (shared,
// next A operand), then one elected thread issues 3 x KD/8 tcgen05.mma.  ll is prefetched two
// frames ahead, while the tensor core works.
// Kernel hmm_stay_tc (frame-parallel): P = a @ pi_su (pi_su[i][j] = pi[i][j] for i < j) and the
// "stay" bits  P_j <= u T_j < P_j + a_j pi[j][j]  that let the backward sampler skip runs.
// Messages are stored unnormalised (a, and T = a @ pi in the same units): every consumer only
// needs ratios.  Masks must be trailing padding (merge_data, util.py:47-60); anything else is
// flagged through boundary_err so that the caller falls back to the sequential kernel.
#include "tc_common.cuh"
#include <stdlib.h>

struct HmmTcArgs {
  const float* ll;
  const double* pi;
  const float* u;
  const int* tv;
  float* a_out;    // [N][Tm][K] unnormalised messages
  float* T_out;    // [N][Tm][K] a @ pi
  unsigned* stay;  // [N][Tm][4]
  double* llpart;  // [N][nch]
  float* bnd;      // [N][nch][2][K]
  int N, Tm, K, chunk_len, burn_in, nch;
  int Kt;  // true number of states; K is the layout width (Kt rounded up to a multiple of 4, padded
           // states carry ll = -inf and zero transition probability)
  int ll_rowmajor;  // 1: `ll` is the caller's (N, Tm, K) row-major array, read in place (K % 4 == 0, 16-byte aligned)
};

template <int KD, int NP>
struct HTC {
  static constexpr int SBO = (KD / 4) * 128;
  static constexpr int A_BYTES = 16 * SBO;
  static constexpr int B_BYTES = (NP / 8) * SBO;
  static constexpr int KSTEPS = KD / 8;
  static constexpr int TMEM_COLS = NP <= 32 ? 32 : (NP <= 64 ? 64 : (NP <= 128 ? 128 : 256));
  static constexpr uint32_t IDESC = umma_idesc_tf32_m128(NP);
  static constexpr size_t SMEM = 1024 + 2 * A_BYTES + 2 * B_BYTES + 256;
};

// flag (as +1e30 in err[0]) sessions whose mask is not of the form  t < tv[n]
__global__ void hmm_mask_check_kernel(const float* __restrict__ mask, long long mask_stride, int Tm,
                                      const int* __restrict__ tv, unsigned* __restrict__ err_bits) {
  const int n = blockIdx.x;
  const int tvn = tv[n];
  bool bad = false;
  for (int t = threadIdx.x; t < Tm; t += blockDim.x)
    bad |= ((mask[(size_t)n * mask_stride + t] > 0.0f) != (t < tvn));
  if (__syncthreads_or(bad) && threadIdx.x == 0) atomicMax(err_bits, __float_as_uint(1e30f));
}

template <int KD, int NP, bool SU>
__device__ __forceinline__ void load_pi_tile(const double* __restrict__ pi, int K, unsigned char* sB_hi,
                                             unsigned char* sB_lo) {
  using C = HTC<KD, NP>;
  for (int e = threadIdx.x; e < NP * KD; e += blockDim.x) {
    const int jn = e / KD, i = e % KD;  // B[n = output state j][k = input state i] = pi[i][j]
    float v = (i < K && jn < K) ? (float)pi[(size_t)i * K + jn] : 0.0f;  // K here = true state count
    if (SU && !(i < jn)) v = 0.0f;
    const uint32_t hi = tf32_rna(v);
    const uint32_t lo = tf32_rna(v - __uint_as_float(hi));
    const size_t off = umma_off(jn, i, C::SBO);
    *reinterpret_cast<uint32_t*>(sB_hi + off) = hi;
    *reinterpret_cast<uint32_t*>(sB_lo + off) = lo;
  }
}

template <int KD, int NP>
__device__ __forceinline__ void issue_3xtf32(uint32_t tmem, uint32_t a_hi, uint32_t a_lo, uint32_t b_hi,
                                             uint32_t b_lo, uint32_t mbar_sa) {
  using C = HTC<KD, NP>;
  uint32_t acc = 0;
#pragma unroll
  for (int ks = 0; ks < C::KSTEPS; ks++) {
    const uint32_t koff = ks * 2 * TC_LBO;
    const uint64_t dah = umma_desc(a_hi + koff, TC_LBO, C::SBO), dal = umma_desc(a_lo + koff, TC_LBO, C::SBO);
    const uint64_t dbh = umma_desc(b_hi + koff, TC_LBO, C::SBO), dbl = umma_desc(b_lo + koff, TC_LBO, C::SBO);
    umma_tf32(tmem, dah, dbh, C::IDESC, acc);
    umma_tf32(tmem, dah, dbl, C::IDESC, 1u);
    umma_tf32(tmem, dal, dbh, C::IDESC, 1u);
    acc = 1u;
  }
  umma_commit(mbar_sa);
}

__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Work split inside a CTA of 512 threads: thread = (row = tid % 128, part = tid / 128).  The warp
// index modulo 4 equals row / 32, which is the TMEM lane quarter that warp may read.  A row's
// KD columns are cut into groups of 8; part p owns groups p, p + 4, p + 8, ...  Four warps per
// scheduler hide the fixed-latency dependencies that a single warp cannot.
#define HTC_THREADS 512
template <int KD>
struct HGroups {
  static constexpr int G = KD / 8;
  static constexpr int MAXG = (G + 3) / 4;
};

__device__ __forceinline__ void ldg8(const float* __restrict__ src, bool ok, float fill, float4& v0, float4& v1) {
  if (ok) {
    v0 = __ldg(reinterpret_cast<const float4*>(src));
    v1 = __ldg(reinterpret_cast<const float4*>(src) + 1);
  } else {
    v0 = make_float4(fill, fill, fill, fill);
    v1 = v0;
  }
}

// ------------------------------------------------------------------ forward filter, 128 chains per CTA
template <int KD, int NP>
__global__ void __launch_bounds__(HTC_THREADS, 1) hmm_forward_tc_kernel(HmmTcArgs p) {
  using C = HTC<KD, NP>;
  using HG = HGroups<KD>;
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sA_hi = smem;
  unsigned char* sA_lo = sA_hi + C::A_BYTES;
  unsigned char* sB_hi = sA_lo + C::A_BYTES;
  unsigned char* sB_lo = sB_hi + C::B_BYTES;
  uint64_t* mbar = reinterpret_cast<uint64_t*>(sB_lo + C::B_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbar + 1);
  int* smax = reinterpret_cast<int*>(tmem_slot + 1);
  __shared__ float psum[2][4][128];  // per-part row sums of frame s (double-buffered by s & 1)
  __shared__ float pmax[2][4][128];  // per-part row maxima of ll for frame s
  const int tid = threadIdx.x, warp = tid >> 5;
  const int row = tid & 127, part = tid >> 7;
  const int K = p.K, Tm = p.Tm, nch = p.nch;
  const long long unit = (long long)blockIdx.x * 128 + row;
  const bool has = unit < (long long)p.N * nch;
  int n = 0, cc = 0, t0 = 0, t1 = 0, tb = 0, tend = 0;
  if (has) {
    n = (int)(unit / nch);
    cc = (int)(unit % nch);
    t0 = (nch == 1) ? 0 : cc * p.chunk_len;
    t1 = (nch == 1) ? Tm : min(Tm, t0 + p.chunk_len);
    const int tvn = p.tv[n];
    tb = (cc > 0) ? max(0, min(t0, tvn) - p.burn_in) : 0;
    tend = min(t1, tvn);  // trailing padding: the chain stops at the first masked frame
  }
  load_pi_tile<KD, NP, false>(p.pi, p.Kt, sB_hi, sB_lo);
  if (tid == 0) *smax = 0;
  const uint32_t mbar_sa = (uint32_t)__cvta_generic_to_shared(mbar);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(mbar_sa));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::);
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(tmem_slot)),
                 "n"(C::TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::);
  }
  __syncthreads();
  if (part == 0) atomicMax(smax, max(tend - tb, 0));
  const size_t rowbase = (size_t)n * Tm;
  const float* __restrict__ llp = p.ll;
  float* __restrict__ aout = p.a_out;
  float* __restrict__ tout = p.T_out;

  // ll of the current and of the next frame live in registers (this thread's column groups); the
  // frame after that is fetched while the tensor core works on the current one.
  float4 llc[HG::MAXG][2], lln[HG::MAXG][2];
  // ll arrives in the interleaved layout (see HmmWs): float4 index ((b*CL + r)*K/4 + kc)*128 + row
  // of the chain that OWNS frame t — during the warm-up that is the previous chunk of the session.
  const float4* __restrict__ ll4 = reinterpret_cast<const float4*>(llp);
  const int KC4f = K >> 2;
  auto fetch_ll = [&](int t, float4 (&dst)[HG::MAXG][2]) {
    const bool act = has && (t < tend);
    const int tt = act ? t : 0;
    const int co = tt / p.chunk_len, ro = tt - co * p.chunk_len;
    const int uo = n * nch + co;
    // interleaved tiles: float4 piece kc of a frame at stride 128; row-major: consecutive
    const float4* src = p.ll_rowmajor ? ll4 + ((size_t)n * Tm + tt) * KC4f
                                      : ll4 + ((size_t)(uo >> 7) * p.chunk_len + ro) * KC4f * 128 + (uo & 127);
    const size_t kstride = p.ll_rowmajor ? 1 : 128;
#pragma unroll
    for (int i = 0; i < HG::MAXG; i++) {
      const int gi = part + 4 * i;
      const int c = 8 * gi;
      const float f = act ? -INFINITY : 0.0f;  // columns beyond K contribute exp(-inf) = 0
      dst[i][0] = (act && c < K) ? __ldg(src + (size_t)(2 * gi) * kstride) : make_float4(f, f, f, f);
      dst[i][1] = (act && c + 4 < K) ? __ldg(src + (size_t)(2 * gi + 1) * kstride) : make_float4(f, f, f, f);
    }
  };
  auto row_max = [&](const float4 (&v)[HG::MAXG][2]) {
    float m = -INFINITY;
#pragma unroll
    for (int i = 0; i < HG::MAXG; i++) {
      m = fmaxf(m, fmaxf(fmaxf(v[i][0].x, v[i][0].y), fmaxf(v[i][0].z, v[i][0].w)));
      m = fmaxf(m, fmaxf(fmaxf(v[i][1].x, v[i][1].y), fmaxf(v[i][1].z, v[i][1].w)));
    }
    return m;
  };
  fetch_ll(tb, llc);
  fetch_ll(tb + 1, lln);
  pmax[0][part][row] = row_max(llc);
  TC_FENCE_PROXY();
  TC_FENCE_BEFORE();
  __syncthreads();
  TC_FENCE_AFTER();
  const int S = *smax;
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
  const uint32_t a_hi_sa = (uint32_t)__cvta_generic_to_shared(sA_hi), a_lo_sa = (uint32_t)__cvta_generic_to_shared(sA_lo);
  const uint32_t b_hi_sa = (uint32_t)__cvta_generic_to_shared(sB_hi), b_lo_sa = (uint32_t)__cvta_generic_to_shared(sB_lo);
  const size_t rowoff = (size_t)(row / 8) * C::SBO + (row % 8) * 16;
  const float LOG2E = 1.4426950408889634f;
  const int KC4 = K >> 2;

  double logtot = 0.0, prod = 1.0;  // marginal log-likelihood = log prod(norm) + sum(cmax)
  float cmax_prev = 0.0f;
  uint32_t parity = 0;
  for (int s = 0; s <= S; s++) {
    const int t = tb + s;
    const int buf = s & 1;
    const bool active = has && (s < S) && (t < tend);
    const bool prev_active = has && (s > 0) && (t - 1 < tend);
    const bool prev_store = prev_active && (t - 1 >= t0);
    const bool store = active && (t >= t0);
    const bool bstart = active && (t == t0) && (cc > 0);
    const bool bend = prev_active && (t == tend) && (nch > 1);
    // normaliser of the previous frame and maximum of this one, summed / maxed over the four parts
    float inv = 1.0f / (float)p.Kt;
    if (s > 0) {
      const float norm = ((psum[buf ^ 1][0][row] + psum[buf ^ 1][1][row]) +
                          (psum[buf ^ 1][2][row] + psum[buf ^ 1][3][row])) + 1e-16f;
      inv = __frcp_rn(norm);
      if (part == 0 && prev_store) {
        prod *= (double)norm;
        logtot += (double)cmax_prev;
        if ((s & 15) == 0) {
          logtot += log(prod);
          prod = 1.0;
        }
      }
    }
    const float cmax = active ? fmaxf(fmaxf(pmax[buf][0][row], pmax[buf][1][row]),
                                      fmaxf(pmax[buf][2][row], pmax[buf][3][row]))
                              : 0.0f;
    cmax_prev = cmax;
    const float inva = active ? inv : 0.0f;
    // interleaved message tiles (hmm_ffbs.cu: HmmWs): float4 index ((b*CL + r)*K/4 + kc)*128 + row
    float4* ar4 = reinterpret_cast<float4*>(aout) + ((size_t)blockIdx.x * p.chunk_len + (t - t0)) * KC4 * 128 + row;
    float4* Tr4 = reinterpret_cast<float4*>(tout) + ((size_t)blockIdx.x * p.chunk_len + (t - 1 - t0)) * KC4 * 128 + row;
    float sum = 0.0f;
#pragma unroll
    for (int i = 0; i < HG::MAXG; i++) {
      const int g = part + 4 * i;
      if (g < HG::G) {  // warp-uniform
        const int c = 8 * g;
        float Tv[8];
        if (s > 0) {
          uint32_t tr[16];
          tmem_ld8(taddr + (uint32_t)c, tr);
#pragma unroll
          for (int q = 0; q < 8; q++) Tv[q] = __uint_as_float(tr[q]);
        } else {
#pragma unroll
          for (int q = 0; q < 8; q++) Tv[q] = 1.0f;
        }
        if (prev_store) {
          if (c < K) Tr4[(size_t)(2 * g) * 128] = make_float4(Tv[0], Tv[1], Tv[2], Tv[3]);
          if (c + 4 < K) Tr4[(size_t)(2 * g + 1) * 128] = make_float4(Tv[4], Tv[5], Tv[6], Tv[7]);
        }
        if (bstart || bend) {
          float* bd = p.bnd + ((size_t)(n * nch + cc) * 2 + (bstart ? 0 : 1)) * K;
          for (int q = 0; q < 8 && c + q < K; q++) bd[c + q] = Tv[q] * inv;
        }
        const float l8[8] = {llc[i][0].x, llc[i][0].y, llc[i][0].z, llc[i][0].w,
                             llc[i][1].x, llc[i][1].y, llc[i][1].z, llc[i][1].w};
        float a8[8];
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int q = 0; q < 8; q++) {
          a8[q] = (Tv[q] * inva) * ex2_ftz((l8[q] - cmax) * LOG2E);
          hi[q] = tf32_rna(a8[q]);
          lo[q] = tf32_rna(a8[q] - __uint_as_float(hi[q]));
        }
        sum += ((a8[0] + a8[1]) + (a8[2] + a8[3])) + ((a8[4] + a8[5]) + (a8[6] + a8[7]));
        if (store) {
          if (c < K) ar4[(size_t)(2 * g) * 128] = make_float4(a8[0], a8[1], a8[2], a8[3]);
          if (c + 4 < K) ar4[(size_t)(2 * g + 1) * 128] = make_float4(a8[4], a8[5], a8[6], a8[7]);
        }
        const size_t off = rowoff + (size_t)(2 * g) * TC_LBO;
        *reinterpret_cast<uint4*>(sA_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(sA_hi + off + TC_LBO) = make_uint4(hi[4], hi[5], hi[6], hi[7]);
        *reinterpret_cast<uint4*>(sA_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        *reinterpret_cast<uint4*>(sA_lo + off + TC_LBO) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
      }
    }
    if (s == S) break;
    psum[buf][part][row] = sum;
    pmax[buf ^ 1][part][row] = row_max(lln);
    TC_FENCE_PROXY();
    TC_FENCE_BEFORE();
    __syncthreads();
    if (tid == 0) {
      TC_FENCE_AFTER();
      issue_3xtf32<KD, NP>(tmem_base, a_hi_sa, a_lo_sa, b_hi_sa, b_lo_sa, mbar_sa);
    }
#pragma unroll
    for (int i = 0; i < HG::MAXG; i++) {
      llc[i][0] = lln[i][0];
      llc[i][1] = lln[i][1];
    }
    fetch_ll(t + 2, lln);  // in flight while the tensor core works
    mbar_wait(mbar_sa, parity);
    parity ^= 1u;
    TC_FENCE_AFTER();
  }
  if (has && part == 0) p.llpart[(size_t)n * nch + cc] = logtot + log(prod);
  TC_FENCE_BEFORE();
  __syncthreads();
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "n"(C::TMEM_COLS));
}

// ------------------------------------------------------------------ stay bits, frame-parallel
template <int KD, int NP>
__global__ void __launch_bounds__(HTC_THREADS, 1) hmm_stay_tc_kernel(HmmTcArgs p) {
  using C = HTC<KD, NP>;
  using HG = HGroups<KD>;
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sA_hi = smem;
  unsigned char* sA_lo = sA_hi + C::A_BYTES;
  unsigned char* sB_hi = sA_lo + C::A_BYTES;
  unsigned char* sB_lo = sB_hi + C::B_BYTES;
  uint64_t* mbar = reinterpret_cast<uint64_t*>(sB_lo + C::B_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbar + 1);
  __shared__ float pjj_s[128];
  __shared__ __align__(16) unsigned char sbits[128][16];  // one byte per 8-column group
  const int tid = threadIdx.x, warp = tid >> 5;
  const int row = tid & 127, part = tid >> 7;
  const int K = p.K, Tm = p.Tm;
  load_pi_tile<KD, NP, true>(p.pi, p.Kt, sB_hi, sB_lo);
  if (tid < 128) pjj_s[tid] = (tid < p.Kt) ? (float)p.pi[(size_t)tid * p.Kt + tid] : 0.0f;
  for (int e = tid; e < 128 * 4; e += HTC_THREADS) reinterpret_cast<unsigned*>(&sbits[0][0])[e] = 0u;
  const uint32_t mbar_sa = (uint32_t)__cvta_generic_to_shared(mbar);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(mbar_sa));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::);
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(tmem_slot)),
                 "n"(C::TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::);
  }
  TC_FENCE_PROXY();
  TC_FENCE_BEFORE();
  __syncthreads();
  TC_FENCE_AFTER();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
  const uint32_t a_hi_sa = (uint32_t)__cvta_generic_to_shared(sA_hi), a_lo_sa = (uint32_t)__cvta_generic_to_shared(sA_lo);
  const uint32_t b_hi_sa = (uint32_t)__cvta_generic_to_shared(sB_hi), b_lo_sa = (uint32_t)__cvta_generic_to_shared(sB_lo);
  const size_t rowoff = (size_t)(row / 8) * C::SBO + (row % 8) * 16;
  // one tile = the 128 chains of a forward CTA at one in-chunk offset r (interleaved layout)
  const long long units = (long long)p.N * p.nch;
  const long long nblk = ((units + 127) / 128) * p.chunk_len;
  const int KC4 = K >> 2;
  const float4* __restrict__ ap = reinterpret_cast<const float4*>(p.a_out);
  const float4* __restrict__ tp = reinterpret_cast<const float4*>(p.T_out);
  uint32_t parity = 0;

  // frame of this thread's row in tile blk: -1 when the row carries no message there
  auto frame_of = [&](long long blk, bool& inrange) -> long long {
    inrange = false;
    if (blk >= nblk) return -1;
    const long long b = blk / p.chunk_len;
    const int r = (int)(blk - b * p.chunk_len);
    const long long u = b * 128 + row;
    if (u >= units) return -1;
    const int n = (int)(u / p.nch), cc = (int)(u % p.nch);
    const int t = cc * p.chunk_len + r;
    if (t >= Tm) return -1;
    inrange = true;
    const long long g = (long long)n * Tm + t;
    return (t < p.tv[n]) ? g : -(g + 2);  // masked frames carry no message: encoded as negative
  };
  auto fetch_row = [&](const float4* __restrict__ base, long long blk, bool live, float4 (&dst)[HG::MAXG][2]) {
    const float4* r = base + (size_t)blk * KC4 * 128 + row;
#pragma unroll
    for (int i = 0; i < HG::MAXG; i++) {
      const int gi = part + 4 * i;
      const int c = 8 * gi;
      dst[i][0] = (live && c < K) ? __ldg(r + (size_t)(2 * gi) * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
      dst[i][1] = (live && c + 4 < K) ? __ldg(r + (size_t)(2 * gi + 1) * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };

  float4 arow[HG::MAXG][2], anext[HG::MAXG][2], trow[HG::MAXG][2];
  long long blk = blockIdx.x;
  bool inr = false;
  long long g = frame_of(blk, inr);
  bool live = g >= 0;
  if (blk < nblk) fetch_row(ap, blk, live, arow);
  for (; blk < nblk; blk += gridDim.x) {
#pragma unroll
    for (int i = 0; i < HG::MAXG; i++) {
      const int gi = part + 4 * i;
      if (gi < HG::G) {
        const float a8[8] = {arow[i][0].x, arow[i][0].y, arow[i][0].z, arow[i][0].w,
                             arow[i][1].x, arow[i][1].y, arow[i][1].z, arow[i][1].w};
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int q = 0; q < 8; q++) {
          hi[q] = tf32_rna(a8[q]);
          lo[q] = tf32_rna(a8[q] - __uint_as_float(hi[q]));
        }
        const size_t off = rowoff + (size_t)(2 * gi) * TC_LBO;
        *reinterpret_cast<uint4*>(sA_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(sA_hi + off + TC_LBO) = make_uint4(hi[4], hi[5], hi[6], hi[7]);
        *reinterpret_cast<uint4*>(sA_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        *reinterpret_cast<uint4*>(sA_lo + off + TC_LBO) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
      }
    }
    TC_FENCE_PROXY();
    TC_FENCE_BEFORE();
    __syncthreads();
    if (tid == 0) {
      TC_FENCE_AFTER();
      issue_3xtf32<KD, NP>(tmem_base, a_hi_sa, a_lo_sa, b_hi_sa, b_lo_sa, mbar_sa);
    }
    // while the tensor core works: T row and u of this tile, message rows of the next tile
    fetch_row(tp, blk, live, trow);
    const float uu = live ? __ldg(p.u + g) : 0.0f;
    bool inr_next = false;
    const long long g_next = frame_of(blk + gridDim.x, inr_next);
    const bool live_next = g_next >= 0;
    fetch_row(ap, blk + gridDim.x, live_next, anext);
    mbar_wait(mbar_sa, parity);
    parity ^= 1u;
    TC_FENCE_AFTER();
#pragma unroll
    for (int i = 0; i < HG::MAXG; i++) {
      const int gi = part + 4 * i;
      if (gi < HG::G) {
        const int c = 8 * gi;
        uint32_t tr[16];
        tmem_ld8(taddr + (uint32_t)c, tr);
        const float a8[8] = {arow[i][0].x, arow[i][0].y, arow[i][0].z, arow[i][0].w,
                             arow[i][1].x, arow[i][1].y, arow[i][1].z, arow[i][1].w};
        const float t8[8] = {trow[i][0].x, trow[i][0].y, trow[i][0].z, trow[i][0].w,
                             trow[i][1].x, trow[i][1].y, trow[i][1].z, trow[i][1].w};
        unsigned bits = 0u;
#pragma unroll
        for (int q = 0; q < 8; q++) {
          const float P = __uint_as_float(tr[q]);
          const float thr = uu * t8[q];
          const bool st = live && (c + q < p.Kt) && (P <= thr) && (P + a8[q] * pjj_s[(c + q) & 127] > thr);
          bits |= (st ? 1u : 0u) << q;
        }
        sbits[row][gi] = (unsigned char)bits;
      }
    }
    TC_FENCE_BEFORE();
    __syncthreads();
    if (part == 0 && inr) {
      const long long gg = (g >= 0) ? g : -(g + 2);
      *reinterpret_cast<uint4*>(p.stay + gg * 4) = *reinterpret_cast<const uint4*>(&sbits[row][0]);
    }
    live = live_next;
    g = g_next;
    inr = inr_next;
#pragma unroll
    for (int i = 0; i < HG::MAXG; i++) {
      arow[i][0] = anext[i][0];
      arow[i][1] = anext[i][1];
    }
  }
  if (warp == 0)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "n"(C::TMEM_COLS));
}

template <int KD, int NP>
static int launch_hmm_tc(const HmmTcArgs& a, cudaStream_t st) {
  using C = HTC<KD, NP>;
  auto kf = hmm_forward_tc_kernel<KD, NP>;
  auto ks = hmm_stay_tc_kernel<KD, NP>;
  KPMS_CUDA(cudaFuncSetAttribute(kf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
  KPMS_CUDA(cudaFuncSetAttribute(ks, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
  const long long units = (long long)a.N * a.nch;
  kf<<<(unsigned)((units + 127) / 128), HTC_THREADS, C::SMEM, st>>>(a);
  int rc = kpms_last_launch();
  if (rc) return rc;
  const long long nblk = ((units + 127) / 128) * a.chunk_len;
  ks<<<(unsigned)(nblk < 148 ? nblk : 148), HTC_THREADS, C::SMEM, st>>>(a);
  return kpms_last_launch();
}

// Row-major ll (N, Tm, Kt) -> interleaved tiles of layout width K = round_up(Kt, 4) (padded states
// get -inf).  One CTA per tile (b, r): 128 chains x K values, read as whole rows (coalesced),
// transposed through shared memory, written as whole lines.
__global__ void __launch_bounds__(256) hmm_ll_tile_kernel(const float* __restrict__ ll, int N, int Tm, int Kt, int K,
                                                          int chunk_len, int nch, float* __restrict__ out) {
  extern __shared__ float4 tl[];  // [K/4][129]
  __shared__ long long src_row[128];
  const int KC4 = K >> 2;
  const long long blk = blockIdx.x;
  const long long b = blk / chunk_len;
  const int r = (int)(blk - b * chunk_len);
  const long long units = (long long)N * nch;
  if (threadIdx.x < 128) {
    const long long u = b * 128 + threadIdx.x;
    long long g = -1;
    if (u < units) {
      const int n = (int)(u / nch), cc = (int)(u % nch);
      const int t = cc * chunk_len + r;
      if (t < Tm) g = (long long)n * Tm + t;
    }
    src_row[threadIdx.x] = g;
  }
  __syncthreads();
  if (Kt == K) {
    const float4* ll4 = reinterpret_cast<const float4*>(ll);
    for (int i = threadIdx.x; i < 128 * KC4; i += 256) {
      const int row = i / KC4, kc = i - row * KC4;
      const long long g = src_row[row];
      tl[kc * 129 + row] = (g >= 0) ? __ldg(ll4 + (size_t)g * KC4 + kc) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  } else {  // rows are not 16-byte aligned: element loads, padding filled with -inf
    float* tlf = reinterpret_cast<float*>(tl);
    for (int i = threadIdx.x; i < 128 * K; i += 256) {
      const int row = i / K, j = i - row * K;
      const long long g = src_row[row];
      const float v = (g >= 0) ? ((j < Kt) ? __ldg(ll + (size_t)g * Kt + j) : -INFINITY) : 0.0f;
      tlf[((size_t)(j >> 2) * 129 + row) * 4 + (j & 3)] = v;
    }
  }
  __syncthreads();
  float4* dst = reinterpret_cast<float4*>(out) + (size_t)blk * KC4 * 128;
  for (int i = threadIdx.x; i < 128 * KC4; i += 256) {
    const int kc = i >> 7, row = i & 127;
    dst[i] = tl[kc * 129 + row];
  }
}

// Chunked passes assume trailing padding: flags anything else as 1e30 in *err_bits.
int kpms_hmm_mask_check(const float* mask, long long mask_stride, int Tm, const int* tv, int N,
                        unsigned* err_bits, cudaStream_t st) {
  hmm_mask_check_kernel<<<N, 256, 0, st>>>(mask, mask_stride, Tm, tv, err_bits);
  return kpms_last_launch();
}

// any K <= 104: the layout width is K rounded up to a multiple of 4
bool kpms_hmm_tc_supported(int Tm, int K, int chunk_len) {
  if (kpms_opts.hmm_tc == 0) return false;  // option: CUDA-core forward kernel in chunked mode
  return chunk_len > 0 && chunk_len < Tm && K <= 104;
}
// the emission kernel can write the interleaved layout itself only when no padding is involved
bool kpms_hmm_tc_native(int Tm, int K, int chunk_len) {
  return kpms_hmm_tc_supported(Tm, K, chunk_len) && (K & 3) == 0;
}

// Tensor-core forward filter + stay bits.  `ll_tiled` != 0: ll is already in the interleaved
// layout; otherwise it is (N, Tm, K) row-major and is transposed into `ll_scratch` first.
// Returns KPMS_ERR_UNSUPPORTED when K has no instantiation.
int kpms_hmm_forward_tc(const float* ll, int ll_tiled, float* ll_scratch, const float* mask,
                        long long mask_stride, const double* pi, const float* u, const int* tv, int N,
                        int Tm, int K, int chunk_len, int burn_in, int nch, float* a_out, float* T_out,
                        unsigned* stay, double* llpart, float* bnd, unsigned* err_bits, cudaStream_t st) {
  if (!kpms_hmm_tc_supported(Tm, K, chunk_len)) return KPMS_ERR_UNSUPPORTED;
  int rc = 0;
  (void)mask; (void)mask_stride; (void)err_bits;  // the caller has run kpms_hmm_mask_check
  const int Kl = (K + 3) & ~3;
  if (ll_tiled && Kl != K) return KPMS_ERR_UNSUPPORTED;
  // row-major ll can be read in place when its rows are whole float4s (no re-tiling pass, no second copy of ll)
  const int rowmajor = !ll_tiled && Kl == K && kpms_aligned(ll, 16) && kpms_opts.hmm_ll_rowmajor;
  if (!ll_tiled && !rowmajor) {
    const long long tiles = (((long long)N * nch + 127) / 128) * chunk_len;
    const size_t smem = (size_t)(Kl >> 2) * 129 * sizeof(float4);
    KPMS_CUDA(cudaFuncSetAttribute(hmm_ll_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    hmm_ll_tile_kernel<<<(unsigned)tiles, 256, smem, st>>>(ll, N, Tm, K, Kl, chunk_len, nch, ll_scratch);
    rc = kpms_last_launch();
    if (rc) return rc;
    ll = ll_scratch;
  }
  HmmTcArgs a{ll, pi, u, tv, a_out, T_out, stay, llpart, bnd, N, Tm, Kl, chunk_len, burn_in, nch, K, rowmajor};
  if (Kl <= 32) return launch_hmm_tc<32, 32>(a, st);
  if (Kl <= 64) return launch_hmm_tc<64, 64>(a, st);
  return launch_hmm_tc<104, 112>(a, st);
}
