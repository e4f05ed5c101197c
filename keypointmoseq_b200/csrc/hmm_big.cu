// a2+a3 for large state counts (128 < K <= 512), chunk-parallel.
// Reference: distributions.py:46-79 (sample_hmm_stateseq) — the same recursion at every K.
//
// With K = 500 the transition matrix (1 MB in fp32) fits neither registers nor shared memory, and
// one CTA per session (the K-generic kernel of hmm_ffbs.cu) leaves a 31 250-step serial chain on
// 32 SMs.  Here every session is cut into chunks with a warm-up (as for K <= 128), and a CTA steps
// HB_CH = 32 chunk chains in lock-step, so that one frame of work is a real
//   (32 chains x K) @ (K x K)
// product: pi is streamed from L2 through a double-buffered cp.async ring of 16-row tiles, the
// chains' messages sit in shared memory (state-major, so that the 8 chains a thread owns are one
// 32-byte broadcast load), each thread accumulates an 8-chain x KP/64-state register tile.
// The epilogue of a frame (normaliser, next frame's row maximum, marginal log-likelihood, the next
// message) is thread-local plus two 64-thread reductions per chain.  Messages are stored
// unnormalised (the sampler only needs ratios within a frame).
// The backward sampler runs one warp per chunk without stay bits (every unmasked frame evaluates the
// K-term inverse CDF: two coalesced 2 KB rows per frame), with the same coalescence warm-up and
// device-side repair as the K <= 128 sampler.
#include "common.cuh"

#define HB_CH 32
#define HB_TK 16
#define HB_NT 256
#define HB_RING 4  // transition-table tiles in shared memory: three in flight while one is read (a repair round with one
                   // live chain is a pure stream of the table: its rate is the depth of this ring over the L2 latency)
#define HB_AS 36  // row stride (floats) of the state-major message buffer: 144 B keeps 16-byte stores conflict-free

struct HmmBigArgs {
  const float* ll;       // (N, Tm, K) row-major
  const float* mask;
  long long mask_stride;
  const float* pf;       // (KP, KP) fp32, pf[i*KP + j] = pi[i][j], zero padded
  const float* pT;       // (KP, KP) fp32, pT[s*KP + i] = pi[i][s], zero padded
  const float* u;
  const int* tv;
  float* alphan;         // (N, Tm, K) unnormalised messages
  double* llpart;        // [N][nch]
  float* bnd;            // [N][nch][2][Kl]
  int* carry_warm;
  int* carry_out;
  int N, Tm, K, Kl, chunk_len, burn_in, nch;
  int bwd_burn;          // warm-up of the sampler's chunks (normally = burn_in)
  // forward repair (see hmm_big_mark_kernel): rlist[0] = number of chunks to re-filter, rlist[1..] their
  // unit ids n*nch + c; rmode = 1: the forward kernel processes exactly those, each from the carry its
  // predecessor left behind (no warm-up)
  int* rlist;
  int rmode;
  float rtol;
};
#define HB_RMAX 128      // chunks one repair round can take (4 CTAs)

__global__ void hmm_big_prep_pi_kernel(const double* __restrict__ pi, int K, int KP, float* __restrict__ pf,
                                       float* __restrict__ pT) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= KP * KP) return;
  const int i = idx / KP, j = idx % KP;
  const float v = (i < K && j < K) ? (float)pi[(size_t)i * K + j] : 0.0f;
  pf[idx] = v;
  pT[(size_t)j * KP + i] = v;
}

// Reduce `v` over the 64 threads (two warps) that share a chain group; `red` is [HB_CH][2] floats.
// Every thread of the CTA calls this for each of its 8 chains (call sites are uniform).
template <bool MAX>
__device__ __forceinline__ void hb_reduce8(float (&v)[8], float* red, int cg, int lane, int wpair) {
#pragma unroll
  for (int c = 0; c < 8; c++) {
    float x = v[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float y = __shfl_xor_sync(0xffffffffu, x, o);
      x = MAX ? fmaxf(x, y) : x + y;
    }
    if (lane == 0) red[(cg * 8 + c) * 2 + wpair] = x;
  }
  __syncthreads();
#pragma unroll
  for (int c = 0; c < 8; c++) {
    const float a = red[(cg * 8 + c) * 2], b = red[(cg * 8 + c) * 2 + 1];
    v[c] = MAX ? fmaxf(a, b) : a + b;
  }
  __syncthreads();
}

template <int KP>
__global__ void __launch_bounds__(HB_NT, 1) hmm_big_forward_kernel(HmmBigArgs p) {
  constexpr int SN = KP / 64;        // states per thread: j = q*64 + sg
  constexpr int NTILE = KP / HB_TK;  // pi tiles per frame
  extern __shared__ __align__(16) float smf[];
  float* a_s = smf;                          // [KP][HB_AS]: a_s[i*HB_AS + chain]
  float* pi_s = a_s + (size_t)KP * HB_AS;    // HB_RING x [HB_TK][KP]
  float* red = pi_s + HB_RING * HB_TK * KP;  // [HB_CH][2]
  __shared__ int s_info[HB_CH][6];           // n, cc, t0, tb, tend, -
  __shared__ int s_S;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int sg = tid & 63, cg = tid >> 6, wpair = warp & 1;
  const int K = p.K, Tm = p.Tm, nch = p.nch;
  const int rcount = p.rmode ? min(p.rlist[0], HB_RMAX) : 0;
  if (p.rmode && (int)blockIdx.x * HB_CH >= rcount) return;   // nothing (left) to repair: the whole CTA leaves
  if (tid == 0) s_S = 0;
  __syncthreads();
  if (tid < HB_CH) {
    long long unit = (long long)blockIdx.x * HB_CH + tid;
    bool valid = unit < (long long)p.N * nch;
    if (p.rmode) {
      valid = unit < rcount;
      unit = valid ? p.rlist[1 + unit] : 0;
    }
    int n = 0, cc = 0, t0 = 0, tb = 0, tend = 0;
    if (valid) {
      n = (int)(unit / nch);
      cc = (int)(unit % nch);
      t0 = cc * p.chunk_len;
      const int t1 = min(Tm, t0 + p.chunk_len);
      const int tvn = p.tv[n];
      tb = (cc > 0) ? max(0, min(t0, tvn) - (p.rmode ? 0 : p.burn_in)) : 0;
      tend = min(t1, tvn);  // trailing padding: the chain stops at the first masked frame
    }
    s_info[tid][0] = n; s_info[tid][1] = cc; s_info[tid][2] = t0; s_info[tid][3] = tb; s_info[tid][4] = tend;
    s_info[tid][5] = valid ? 1 : 0;
    atomicMax(&s_S, max(tend - tb, 0));
  }
  __syncthreads();
  const int S = s_S;
  int cn[8], ccc[8], ct0[8], ctb[8], cte[8];
  bool cva[8];
#pragma unroll
  for (int c = 0; c < 8; c++) {
    const int ch = cg * 8 + c;
    cn[c] = s_info[ch][0]; ccc[c] = s_info[ch][1]; ct0[c] = s_info[ch][2]; ctb[c] = s_info[ch][3]; cte[c] = s_info[ch][4];
    cva[c] = s_info[ch][5] != 0;
  }
  // chains of this thread's group that carry work: all eight in a normal pass; in a repair round the listed
  // chunks fill the slots in order, so the first `nact` of the group are the live ones
  const int nact = p.rmode ? max(0, min(8, rcount - ((int)blockIdx.x * HB_CH + cg * 8))) : 8;
  const unsigned pis_sa = (unsigned)__cvta_generic_to_shared(pi_s);
  auto issue_tile = [&](int g) {  // global tile counter g: tile (g % NTILE) into buffer (g % HB_RING)
    const float* src = p.pf + (size_t)(g % NTILE) * HB_TK * KP;
    const unsigned dst = pis_sa + (unsigned)((g % HB_RING) * HB_TK * KP * 4);
#pragma unroll
    for (int e = 0; e < HB_TK * KP / 4 / HB_NT; e++) {
      const int piece = tid + e * HB_NT;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst + (unsigned)piece * 16), "l"(src + (size_t)piece * 4));
    }
    asm volatile("cp.async.commit_group;\n" ::);
  };
  // ll of frame tb + s for this thread's chains and states (0 for inactive chains / padded states)
  auto load_ll = [&](int s, float (&dst)[8][SN]) {
#pragma unroll
    for (int c = 0; c < 8; c++) {
      const int t = ctb[c] + s;
      const bool act = t < cte[c];
      const float* row = p.ll + ((size_t)cn[c] * Tm + (act ? t : 0)) * K;
#pragma unroll
      for (int q = 0; q < SN; q++) {
        const int j = q * 64 + sg;
        dst[c][q] = (act && j < K) ? __ldg(row + j) : -INFINITY;
      }
    }
  };

  float T[8][SN];     // in_pot (unnormalised): T = a @ pi of the previous frame, 1 before the first
  float inv[8];       // 1 / normaliser of the previous frame (1/K before the first)
  float lln[8][SN];
  double logtot[8];
#pragma unroll
  for (int c = 0; c < 8; c++) {
    inv[c] = 1.0f / (float)K;
    logtot[c] = 0.0;
#pragma unroll
    for (int q = 0; q < SN; q++) T[c][q] = (q * 64 + sg < K) ? 1.0f : 0.0f;
    if (p.rmode && cva[c] && ccc[c] > 0) {
      // repair: start from the (normalised) carry the predecessor chunk left behind
      const float* pe = p.bnd + ((size_t)(cn[c] * nch + ccc[c] - 1) * 2 + 1) * p.Kl;
      inv[c] = 1.0f;
#pragma unroll
      for (int q = 0; q < SN; q++) T[c][q] = (q * 64 + sg < K) ? pe[q * 64 + sg] : 0.0f;
    }
  }
  load_ll(0, lln);
#pragma unroll
  for (int g0 = 0; g0 < HB_RING - 1; g0++) issue_tile(g0);   // HB_RING - 1 tiles in flight
  int g = 0;  // global tile counter
  for (int s = 0; s < S; s++) {
    // ---- message of frame tb + s:  a = in_pot * exp(ll - max ll),  norm = sum a + 1e-16
    float cmax[8];
#pragma unroll
    for (int c = 0; c < 8; c++) {
      float m = -INFINITY;
#pragma unroll
      for (int q = 0; q < SN; q++) m = fmaxf(m, lln[c][q]);
      cmax[c] = m;
    }
    hb_reduce8<true>(cmax, red, cg, lane, wpair);
    float asum[8];
#pragma unroll
    for (int c = 0; c < 8; c++) {
      const int t = ctb[c] + s;
      const bool act = t < cte[c];
      // boundary record: the warmed-up (normalised) carry a chunk starts from
      if (act && t == ct0[c] && ccc[c] > 0) {
        float* bd = p.bnd + ((size_t)(cn[c] * nch + ccc[c]) * 2 + 0) * p.Kl;
#pragma unroll
        for (int q = 0; q < SN; q++)
          if (q * 64 + sg < K) bd[q * 64 + sg] = T[c][q] * inv[c];
      }
      float sum = 0.0f;
      float* arow = p.alphan + ((size_t)cn[c] * Tm + (act ? t : 0)) * K;
      const bool store = act && t >= ct0[c];
#pragma unroll
      for (int q = 0; q < SN; q++) {
        const float a = act ? (T[c][q] * inv[c]) * __expf(lln[c][q] - cmax[c]) : 0.0f;
        T[c][q] = a;  // reused as the message until the product below overwrites it
        sum += a;
        if (store && q * 64 + sg < K) arow[q * 64 + sg] = a;
      }
      asum[c] = sum;
    }
    // state-major copy for the product (every thread: its SN states x 8 chains)
#pragma unroll
    for (int q = 0; q < SN; q++) {
      float* dst = a_s + (size_t)(q * 64 + sg) * HB_AS + cg * 8;
      *reinterpret_cast<float4*>(dst) = make_float4(T[0][q], T[1][q], T[2][q], T[3][q]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(T[4][q], T[5][q], T[6][q], T[7][q]);
    }
    hb_reduce8<false>(asum, red, cg, lane, wpair);   // its barriers also publish a_s
#pragma unroll
    for (int c = 0; c < 8; c++) {
      const int t = ctb[c] + s;
      const float norm = asum[c] + 1e-16f;
      if (t < cte[c]) {
        inv[c] = 1.0f / norm;
        if (t >= ct0[c] && sg == 0) logtot[c] += (double)__logf(norm) + (double)cmax[c];
      }
    }
    load_ll(s + 1, lln);   // in flight during the product
    // ---- T = a @ pi, pi streamed in HB_TK-row tiles
#pragma unroll
    for (int c = 0; c < 8; c++)
#pragma unroll
      for (int q = 0; q < SN; q++) T[c][q] = 0.0f;
#pragma unroll 1
    for (int tile = 0; tile < NTILE; tile++, g++) {
      asm volatile("cp.async.wait_group %0;\n" ::"n"(HB_RING - 2));   // tile g has landed (the newer ones may be on their way)
      __syncthreads();
      issue_tile(g + HB_RING - 1);   // into the buffer tile g-1 has just been read from
      const float* pt = pi_s + (size_t)(g % HB_RING) * HB_TK * KP;
      if (nact > 2) {
#pragma unroll
        for (int ii = 0; ii < HB_TK; ii++) {
          const float* ar = a_s + (size_t)(tile * HB_TK + ii) * HB_AS + cg * 8;
          const float4 a0 = *reinterpret_cast<const float4*>(ar);
          const float4 a1 = *reinterpret_cast<const float4*>(ar + 4);
          const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
          for (int q = 0; q < SN; q++) {
            const float pv = pt[ii * KP + q * 64 + sg];
#pragma unroll
            for (int c = 0; c < 8; c++) T[c][q] = fmaf(av[c], pv, T[c][q]);
          }
        }
      } else if (nact > 0) {   // a repair round with one or two chunks in this group: an eighth / a quarter of the products
#pragma unroll
        for (int ii = 0; ii < HB_TK; ii++) {
          const float2 a0 = *reinterpret_cast<const float2*>(a_s + (size_t)(tile * HB_TK + ii) * HB_AS + cg * 8);
#pragma unroll
          for (int q = 0; q < SN; q++) {
            const float pv = pt[ii * KP + q * 64 + sg];
            T[0][q] = fmaf(a0.x, pv, T[0][q]);
            T[1][q] = fmaf(a0.y, pv, T[1][q]);
          }
        }
      }
    }
    __syncthreads();   // everyone is done with a_s before the next frame overwrites it
    // a chain that has just filtered its last frame leaves its (normalised) carry behind, for
    // verification against the next chunk's warmed-up start (later steps of this CTA zero its T)
    if (nch > 1) {
#pragma unroll
      for (int c = 0; c < 8; c++) {
        if (ctb[c] + s == cte[c] - 1) {
          float* bd = p.bnd + ((size_t)(cn[c] * nch + ccc[c]) * 2 + 1) * p.Kl;
#pragma unroll
          for (int q = 0; q < SN; q++)
            if (q * 64 + sg < K) bd[q * 64 + sg] = T[c][q] * inv[c];
        }
      }
    }
  }
  asm volatile("cp.async.wait_group 0;\n" ::);
#pragma unroll
  for (int c = 0; c < 8; c++) {
    if (cva[c] && sg == 0) p.llpart[(size_t)cn[c] * nch + ccc[c]] = logtot[c];
  }
}

// ------------------------------------------------------------------ backward sampler, K <= 512
#define HBB_WARPS 8

// Walk frames ts, ts-1, ..., t0 of session n (frames >= t1 are a warm-up: nothing written).
// s_next = -1: next_potential = ones (distributions.py:76).  Returns the carry after frame t0.
template <int KP>
__device__ __forceinline__ int hb_backward_walk(const HmmBigArgs& p, const float* __restrict__ mk_n,
                                                const float* __restrict__ u_n, int* __restrict__ z_n, int n, int t0,
                                                int t1, int ts, int s_next, int* carry_at_t1) {
  constexpr int PER = KP / 32;
  const int lane = threadIdx.x & 31;
  const int K = p.K, Tm = p.Tm;
  const float* al_n = p.alphan + (size_t)n * Tm * K;
  int carry1 = s_next;
  bool warm = ts >= t1;
  for (int t = ts; t >= t0; t--) {
    if (warm && t < t1) {
      carry1 = s_next;
      warm = false;
    }
    const bool live = mk_n[t] > 0.0f;
    int zt = 0;
    if (live) {
      const float* ar = al_n + (size_t)t * K;
      const float* pr = p.pT + (size_t)max(s_next, 0) * KP;
      float w[PER];
      float tot = 0.0f;
#pragma unroll
      for (int q = 0; q < PER; q++) {
        const int st = 32 * q + lane;
        const float al = (st < K) ? __ldg(ar + st) : 0.0f;
        const float np = (s_next < 0) ? 1.0f : __ldg(pr + st);
        w[q] = np * al;
        tot += w[q];
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
      const float thr = u_n[t] * tot;
      // z = number of states whose inclusive prefix sum is <= thr (inverse CDF in state order)
      float base = 0.0f;
      int cnt = 0;
#pragma unroll
      for (int q = 0; q < PER; q++) {
        float incl = w[q];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const float y = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += y;
        }
        cnt += __popc(__ballot_sync(0xffffffffu, !(base + incl > thr)));
        base += __shfl_sync(0xffffffffu, incl, 31);
      }
      zt = min(cnt, K - 1);
      s_next = zt;
    }
    if (!warm && lane == 0) z_n[t] = zt;
  }
  if (warm) carry1 = s_next;  // the chunk itself was empty (t1 <= t0 cannot happen; kept for safety)
  *carry_at_t1 = carry1;
  return s_next;
}

template <int KP>
__global__ void __launch_bounds__(HBB_WARPS * 32) hmm_big_backward_kernel(HmmBigArgs p, int* __restrict__ z) {
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const long long unit = (long long)blockIdx.x * HBB_WARPS + wib;
  if (unit >= (long long)p.N * p.nch) return;
  const int n = (int)(unit / p.nch), cc = (int)(unit % p.nch);
  const int t0 = cc * p.chunk_len, t1 = min(p.Tm, t0 + p.chunk_len);
  const int ts = (cc == p.nch - 1) ? t1 - 1 : min(p.Tm, t1 + p.bwd_burn) - 1;
  int carry1;
  const int out = hb_backward_walk<KP>(p, p.mask + (size_t)n * p.mask_stride, p.u + (size_t)n * p.Tm,
                                       z + (size_t)n * p.Tm, n, t0, t1, ts, -1, &carry1);
  if (lane == 0) {
    p.carry_warm[unit] = carry1;
    p.carry_out[unit] = out;
  }
}

// One warp per session walks its chunks from the top: a chunk whose warm-up did not arrive at the
// carry the chunk above really left behind is sampled again from that carry.
template <int KP>
__global__ void __launch_bounds__(HBB_WARPS * 32) hmm_big_brepair_kernel(HmmBigArgs p, int* __restrict__ z) {
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int n = blockIdx.x * HBB_WARPS + wib;
  if (n >= p.N) return;
  const int nch = p.nch;
  for (int cc = nch - 2; cc >= 0; cc--) {
    const size_t unit = (size_t)n * nch + cc;
    const int want = __ldcg(&p.carry_out[unit + 1]);
    if (__ldcg(&p.carry_warm[unit]) == want) continue;  // warp-uniform
    const int t0 = cc * p.chunk_len, t1 = min(p.Tm, t0 + p.chunk_len);
    int carry1;
    const int out = hb_backward_walk<KP>(p, p.mask + (size_t)n * p.mask_stride, p.u + (size_t)n * p.Tm,
                                         z + (size_t)n * p.Tm, n, t0, t1, t1 - 1, want, &carry1);
    __syncwarp();
    if (lane == 0) {
      p.carry_warm[unit] = want;
      p.carry_out[unit] = out;
    }
    __syncwarp();
    __threadfence_block();
  }
}

// Forward repair, step 1: list the chunks whose warmed-up start differs from the carry their predecessor
// arrived with by more than rtol (a NaN counts).  One CTA per session.
__global__ void hmm_big_mark_kernel(HmmBigArgs p) {
  const int n = blockIdx.x, nch = p.nch;
  const int tvn = p.tv[n];
  __shared__ float red[4];
  for (int c = 0; c + 1 < nch; c++) {
    if ((c + 1) * p.chunk_len >= tvn) break;   // the next chunk starts inside the trailing padding
    const float* a = p.bnd + ((size_t)(n * nch + c) * 2 + 1) * p.Kl;
    const float* b = p.bnd + ((size_t)(n * nch + c + 1) * 2 + 0) * p.Kl;
    float worst = 0.0f;
    for (int e = threadIdx.x; e < p.K; e += blockDim.x) {
      const float d = fabsf(a[e] - b[e]);
      worst = (d != d) ? INFINITY : fmaxf(worst, d);
    }
    for (int o = 16; o > 0; o >>= 1) worst = fmaxf(worst, __shfl_xor_sync(0xffffffffu, worst, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = worst;
    __syncthreads();
    if (threadIdx.x == 0) {
      const float w = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
      if (w > p.rtol) {
        const int idx = atomicAdd(p.rlist, 1);
        if (idx < HB_RMAX) p.rlist[1 + idx] = n * nch + c + 1;
      }
    }
    __syncthreads();
  }
}

template <int KP>
static int launch_hmm_big(const HmmBigArgs& a, int* z, int forward_only, cudaStream_t st) {
  const size_t smem = ((size_t)KP * HB_AS + HB_RING * HB_TK * KP + HB_CH * 2) * sizeof(float);
  auto kf = hmm_big_forward_kernel<KP>;
  KPMS_CUDA(cudaFuncSetAttribute(kf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long units = (long long)a.N * a.nch;
  kf<<<(unsigned)((units + HB_CH - 1) / HB_CH), HB_NT, smem, st>>>(a);
  int rc = kpms_last_launch();
  if (rc) return rc;
  // Forward repair: a chunk whose warm-up did not reach the true carry (a numerically reducible chain: which of
  // two similar states it sits in can depend on history further back than any warm-up) is filtered again from
  // its predecessor's end.  Each round fixes the listed chunks; a chunk whose new end no longer matches its
  // successor's start is listed in the next round.  No host round trip: a round with an empty list costs two
  // empty launches.  What is still mismatched afterwards is reported through boundary_err as before.
  if (a.nch > 1 && a.rlist != nullptr) {
    HmmBigArgs r = a;
    r.rmode = 1;
    for (int round = 0; round < kpms_opts.hmm_repair_rounds; round++) {
      KPMS_CUDA(cudaMemsetAsync(a.rlist, 0, sizeof(int), st));
      hmm_big_mark_kernel<<<a.N, 128, 0, st>>>(a);
      rc = kpms_last_launch();
      if (rc) return rc;
      kf<<<HB_RMAX / HB_CH, HB_NT, smem, st>>>(r);
      rc = kpms_last_launch();
      if (rc) return rc;
    }
  }
  if (forward_only) return rc;
  hmm_big_backward_kernel<KP><<<(unsigned)((units + HBB_WARPS - 1) / HBB_WARPS), HBB_WARPS * 32, 0, st>>>(a, z);
  rc = kpms_last_launch();
  if (rc || a.nch == 1) return rc;
  hmm_big_brepair_kernel<KP><<<(unsigned)((a.N + HBB_WARPS - 1) / HBB_WARPS), HBB_WARPS * 32, 0, st>>>(a, z);
  return kpms_last_launch();
}

bool kpms_hmm_big_supported(int Tm, int K, int chunk_len) {
  return K > 128 && K <= 512 && chunk_len > 0 && chunk_len < Tm;
}

// padded width of the fp32 transition tables of this path
int kpms_hmm_big_kp(int K) { return K <= 256 ? 256 : 512; }

// Forward filter + backward sampler; tv (unmasked frames per session) and the mask check have been
// run by the caller, which also runs the finish kernel (log-likelihood sums, boundary mismatch).
int kpms_hmm_big(const float* ll, const float* mask, long long mask_stride, const double* pi, const float* u,
                 const int* tv, int N, int Tm, int K, int Kl, int chunk_len, int burn_in, int bwd_burn, int nch,
                 float* pf, float* pT, float* alphan, double* llpart, float* bnd, int* carry_warm, int* carry_out,
                 int* z, int forward_only, int* repair_scratch, cudaStream_t st) {
  const int KP = kpms_hmm_big_kp(K);
  hmm_big_prep_pi_kernel<<<(KP * KP + 255) / 256, 256, 0, st>>>(pi, K, KP, pf, pT);
  int rc = kpms_last_launch();
  if (rc) return rc;
  HmmBigArgs a{ll, mask, mask_stride, pf, pT, u, tv, alphan, llpart, bnd, carry_warm, carry_out,
               N, Tm, K, Kl, chunk_len, burn_in, nch, bwd_burn >= 0 ? bwd_burn : burn_in,
               repair_scratch, 0, 1e-6f * (float)kpms_opts.hmm_repair_tol_ppm};
  return (KP == 256) ? launch_hmm_big<256>(a, z, forward_only, st) : launch_hmm_big<512>(a, z, forward_only, st);
}
