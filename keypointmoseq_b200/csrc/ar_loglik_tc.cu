// a1 on the 5th-generation tensor cores: tcgen05.mma kind::tf32, 3xTF32 split, FP32 accumulators
// in TMEM.  Reference: gibbs.py:100-104,110 (the (frames x (D*L+1)) . ((D*L+1) x D) contraction per
// state) + tfd.MultivariateNormalFullCovariance.log_prob.
//
//   D[frame, k*DD+d] = sum_p  f[frame, p] * W[k, d, p]        (M = 128 frames, N = 16 states * DD,
//   ll[frame, k]     = cst[k] - 1/2 sum_d D[frame, k*DD+d]^2    contraction KD = round_up(P, 8))
//
// Work split: CTA (tile, stripe) keeps ONE weight tile (16 states; hi and lo TF32 parts, packed by
// ar_tc_pack_kernel in the UMMA K-major no-swizzle layout) resident in shared memory for the whole
// launch and walks its stripe of 128-frame blocks: build the A tile (lag window of x read in place,
// hi/lo split) -> 3 x KD/8 tcgen05.mma into TMEM -> tcgen05.ld epilogue: each thread owns one
// frame row, squares and group-sums DD consecutive columns per state, adds the constant, stores.
// The weights are never re-read from L2/HBM and x is read once per weight tile.
//
// UMMA K-major no-swizzle layout (16-byte units): element (row r, k) lives at
//   (r / 8) * SBO + (k / 4) * LBO + (r % 8) * 16 + (k % 4) * 4   bytes,  LBO = 128, SBO = (KD/4)*128.
#include "tc_common.cuh"
#include "dispatch.cuh"

#define TC_STATES 16   // states per weight tile
#define TC_M 128       // frames per block (UMMA M)
#define TC_THREADS 256 // thread = (frame row, half of the K chunks / half of the states)
// pipelined kernel: threads per frame row (each builds 1/PARTS of the row's chunks and reads 1/PARTS of the states).
// Two: with four (512 threads; possible when 4 states x D is a whole number of 8-column TMEM loads) the kernel was
// measured slower, 2.86 vs 2.68 ms on the bench shard.
#define TCP_PARTS_OF(DD) 2

template <int DD, int LL>
struct TC {
  static constexpr int P = DD * LL + DD + 1;
  static constexpr int PP = (P + 3) / 4 * 4;       // row length of the fp32 weights (ar_prepare_kernel)
  static constexpr int KD = (P + 7) / 8 * 8;       // contraction length, multiple of the UMMA K (8 for tf32)
  static constexpr int KC = KD / 4;                // 16-byte chunks along K
  static constexpr int LBO = 128;
  static constexpr int SBO = KC * 128;
  static constexpr int N = TC_STATES * DD;         // UMMA N (multiple of 16, <= 256 for DD <= 16)
  static constexpr int A_BYTES = (TC_M / 8) * SBO;
  static constexpr int B_BYTES = (N / 8) * SBO;
  static constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) |
                                    ((uint32_t)(TC_M >> 4) << 24);  // F32 accum, TF32 x TF32, K-major both
  static_assert(N % 16 == 0 && N <= 256, "invalid UMMA N");
};

// Pack the fp32 whitened weights into per-tile hi / lo TF32 operands in the UMMA layout.
template <int DD, int LL>
__global__ void ar_tc_pack_kernel(const float* __restrict__ Wf, int K, unsigned char* __restrict__ tiles) {
  using C = TC<DD, LL>;
  const int tile = blockIdx.x;
  unsigned char* base = tiles + (size_t)tile * 2 * C::B_BYTES;
  for (int e = threadIdx.x; e < C::N * C::KD; e += blockDim.x) {
    const int n = e / C::KD, p = e % C::KD;
    const int k = tile * TC_STATES + n / DD, d = n % DD;
    const float v = (k < K && p < C::PP) ? Wf[((size_t)k * DD + d) * C::PP + p] : 0.0f;
    const uint32_t hi = tf32_rna(v);
    const uint32_t lo = tf32_rna(v - __uint_as_float(hi));
    const size_t off = (size_t)(n / 8) * C::SBO + (size_t)(p / 4) * C::LBO + (n % 8) * 16 + (p % 4) * 4;
    *reinterpret_cast<uint32_t*>(base + off) = hi;
    *reinterpret_cast<uint32_t*>(base + C::B_BYTES + off) = lo;
  }
}

template <int DD, int LL>
__global__ void __launch_bounds__(TC_THREADS)
    ar_loglik_tc_kernel(const float* __restrict__ x, const unsigned char* __restrict__ tiles,
                        const float* __restrict__ cst, int N, int T, int K, int nstripes,
                        int tile_cl, int tile_nch, float* __restrict__ ll) {
  using C = TC<DD, LL>;
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sA_hi = smem;
  unsigned char* sA_lo = sA_hi + C::A_BYTES;
  unsigned char* sB_hi = sA_lo + C::A_BYTES;
  unsigned char* sB_lo = sB_hi + C::B_BYTES;
  float* s_cst = reinterpret_cast<float*>(sB_lo + C::B_BYTES);                    // TC_STATES
  uint64_t* mbar = reinterpret_cast<uint64_t*>(s_cst + TC_STATES);                // 8-byte aligned
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbar + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row = tid & (TC_M - 1), half = tid >> 7;
  const int tile = blockIdx.x, stripe = blockIdx.y;
  const int k0 = tile * TC_STATES;
  const int Tm = T - LL;
  const long long frames = (long long)N * Tm;
  // tile_cl > 0: block = one tile (b, r) of the HMM's interleaved layout (hmm_ffbs.cu: HmmWs) — row i is
  // frame t = cc*tile_cl + r of chain u = 128 b + i = n*tile_nch + cc, and ll goes out as whole lines.
  const long long units = (long long)N * tile_nch;
  const long long nblk = (tile_cl > 0) ? ((units + 127) / 128) * tile_cl : (frames + TC_M - 1) / TC_M;

  // ---- one-time setup: weight tile, constants, barrier, tensor memory
  {
    const uint4* src = reinterpret_cast<const uint4*>(tiles + (size_t)tile * 2 * C::B_BYTES);
    uint4* dst = reinterpret_cast<uint4*>(sB_hi);
    for (int i = tid; i < 2 * C::B_BYTES / 16; i += TC_THREADS) dst[i] = src[i];
    if (tid < TC_STATES) s_cst[tid] = (k0 + tid < K) ? cst[k0 + tid] : 0.0f;
  }
  const uint32_t mbar_sa = (uint32_t)__cvta_generic_to_shared(mbar);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(mbar_sa));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::);
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;\n" ::"r"(
        (uint32_t)__cvta_generic_to_shared(tmem_slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::);
  }
  asm volatile("fence.proxy.async.shared::cta;\n" ::);  // weight tile written with generic stores
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::);
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::);
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t a_hi_sa = (uint32_t)__cvta_generic_to_shared(sA_hi);
  const uint32_t a_lo_sa = (uint32_t)__cvta_generic_to_shared(sA_lo);
  const uint32_t b_hi_sa = (uint32_t)__cvta_generic_to_shared(sB_hi);
  const uint32_t b_lo_sa = (uint32_t)__cvta_generic_to_shared(sB_lo);
  uint32_t parity = 0;

  for (long long blk = stripe; blk < nblk; blk += nstripes) {
    // ---- A tile: thread r builds row r = frame blk*128 + r (hi and lo parts)
    long long g = blk * TC_M + row;
    bool live = g < frames;
    {
      const float* xp = x;
      if (tile_cl > 0) {
        const long long b = blk / tile_cl;
        const int r = (int)(blk - b * tile_cl);
        const long long uu = b * 128 + row;
        live = false;
        if (uu < units) {
          const int n = (int)(uu / tile_nch), cc = (int)(uu % tile_nch);
          const int tp = cc * tile_cl + r;
          if (tp < Tm) {
            live = true;
            xp = x + ((size_t)n * T + tp) * DD;
          }
        }
      } else if (live) {
        const int n = (int)(g / Tm), tp = (int)(g % Tm);
        xp = x + ((size_t)n * T + tp) * DD;  // lag window x[tp .. tp+L]: DL+DD contiguous floats
      }
      const size_t rowoff = (size_t)(row / 8) * C::SBO + (row % 8) * 16;
#pragma unroll
      for (int kc = 0; kc < C::KC; kc++) {
        if ((kc & 1) != half) continue;   // the two threads of a row share the chunks
        uint32_t hi[4], lo[4];
        float v4[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        if (live) {
          if ((DD & 1) == 0 && kc * 4 + 3 < DD * LL + DD) {   // windows start on 8-byte boundaries when D is even
            const float2 a = __ldg(reinterpret_cast<const float2*>(xp + kc * 4));
            const float2 b = __ldg(reinterpret_cast<const float2*>(xp + kc * 4 + 2));
            v4[0] = a.x; v4[1] = a.y; v4[2] = b.x; v4[3] = b.y;
          } else {
#pragma unroll
            for (int q = 0; q < 4; q++) {
              const int p = kc * 4 + q;
              if (p < DD * LL + DD) v4[q] = xp[p];
              else if (p == C::P - 1) v4[q] = 1.0f;
            }
          }
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const float v = v4[q];
          hi[q] = tf32_rna(v);
          lo[q] = tf32_rna(v - __uint_as_float(hi[q]));
        }
        *reinterpret_cast<uint4*>(sA_hi + rowoff + kc * C::LBO) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<uint4*>(sA_lo + rowoff + kc * C::LBO) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
      }
    }
    asm volatile("fence.proxy.async.shared::cta;\n" ::);
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::);
    __syncthreads();
    // ---- 3xTF32: hi*hi + hi*lo + lo*hi, KD/8 k-steps each, one elected thread issues
    if (tid == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;\n" ::);
      uint32_t acc = 0;
#pragma unroll
      for (int ks = 0; ks < C::KD / 8; ks++) {
        const uint32_t koff = ks * 2 * C::LBO;
        const uint64_t dah = umma_desc(a_hi_sa + koff, C::LBO, C::SBO), dal = umma_desc(a_lo_sa + koff, C::LBO, C::SBO);
        const uint64_t dbh = umma_desc(b_hi_sa + koff, C::LBO, C::SBO), dbl = umma_desc(b_lo_sa + koff, C::LBO, C::SBO);
#pragma unroll
        for (int term = 0; term < 3; term++) {
          const uint64_t da = (term == 2) ? dal : dah;
          const uint64_t db = (term == 1) ? dbl : dbh;
          asm volatile(
              "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
              "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_base),
              "l"(da), "l"(db), "r"(C::IDESC), "r"(acc));
          acc = 1;
        }
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(mbar_sa));
    }
    // ---- wait for the accumulator
    {
      uint32_t done = 0;
      while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}\n"
            : "=r"(done)
            : "r"(mbar_sa), "r"(parity));
      }
      parity ^= 1u;
    }
    asm volatile("tcgen05.fence::after_thread_sync;\n" ::);
    // ---- epilogue: thread = frame row; warp w reads TMEM lanes 32w..32w+31
    {
      // epilogue: thread = (frame row, half): states 8*half .. 8*half+7 of the tile, i.e. TMEM
      // columns [8*half*DD, 8*(half+1)*DD); warp w reads TMEM lanes 32*(w%4) .. +31
      constexpr int HS = TC_STATES / 2, HC = HS * DD, W = (HC % 16 == 0) ? 16 : 8;
      float q[HS];
#pragma unroll
      for (int sidx = 0; sidx < HS; sidx++) q[sidx] = 0.0f;
      const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(half * HC);
#pragma unroll
      for (int c0 = 0; c0 < HC; c0 += W) {
        uint32_t r[16];
        if (W == 16) tmem_ld16(taddr + (uint32_t)c0, r);
        else tmem_ld8(taddr + (uint32_t)c0, r);
#pragma unroll
        for (int i = 0; i < W; i++) {
          const float v = __uint_as_float(r[i]);
          q[(c0 + i) / DD] = fmaf(v, v, q[(c0 + i) / DD]);
        }
      }
      const int ks = k0 + half * HS;
      const float* cs = s_cst + half * HS;
      if (live && tile_cl > 0) {
        // interleaved layout: float4 index (blk * K/4 + kc) * 128 + row  (K % 4 == 0 guaranteed by the caller)
        float4* out4 = reinterpret_cast<float4*>(ll) + ((size_t)blk * (K >> 2) + (ks >> 2)) * 128 + row;
#pragma unroll
        for (int s4 = 0; s4 < HS; s4 += 4)
          if (ks + s4 < K)
            out4[(size_t)(s4 >> 2) * 128] =
                make_float4(cs[s4] - 0.5f * q[s4], cs[s4 + 1] - 0.5f * q[s4 + 1],
                            cs[s4 + 2] - 0.5f * q[s4 + 2], cs[s4 + 3] - 0.5f * q[s4 + 3]);
      } else if (live) {
        float* out = ll + g * K + ks;
        if ((K & 3) == 0 && ks + HS <= K) {
#pragma unroll
          for (int s4 = 0; s4 < HS; s4 += 4)
            *reinterpret_cast<float4*>(out + s4) =
                make_float4(cs[s4] - 0.5f * q[s4], cs[s4 + 1] - 0.5f * q[s4 + 1],
                            cs[s4 + 2] - 0.5f * q[s4 + 2], cs[s4 + 3] - 0.5f * q[s4 + 3]);
        } else {
#pragma unroll
          for (int sidx = 0; sidx < HS; sidx++)
            if (ks + sidx < K) out[sidx] = cs[sidx] - 0.5f * q[sidx];
        }
      }
    }
    // the next block overwrites the A tile and the accumulator
    asm volatile("tcgen05.fence::before_thread_sync;\n" ::);
    __syncthreads();
  }
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;\n" ::"r"(tmem_base));
  }
}

// The same contraction, software-pipelined inside one CTA per SM: two A stages in shared memory and two
// accumulators in TMEM (columns 0.. and 256..), so that while the tensor core works on block i+1 the
// threads read block i's accumulator (epilogue) and build the A tile of block i+2.  x of block i+2 is
// loaded into registers before the epilogue, so its latency is covered as well.  Per block the CTA costs
// max(MMA, epilogue + A build) instead of their sum.
template <int DD, int LL>
__global__ void __launch_bounds__(128 * TCP_PARTS_OF(DD), 1)
    ar_loglik_tc_pipe_kernel(const float* __restrict__ x, const unsigned char* __restrict__ tiles,
                             const float* __restrict__ cst, int N, int T, int K, int nstripes,
                             int tile_cl, int tile_nch, float* __restrict__ ll) {
  using C = TC<DD, LL>;
  constexpr int TCP_PARTS = TCP_PARTS_OF(DD), TCP_THREADS = 128 * TCP_PARTS;
  constexpr int NCH = (C::KC + TCP_PARTS - 1) / TCP_PARTS;   // 16-byte chunks of a row built by one thread
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sA = smem;                                   // [stage][hi | lo]
  unsigned char* sB_hi = sA + 4 * C::A_BYTES;
  unsigned char* sB_lo = sB_hi + C::B_BYTES;
  float* s_cst = reinterpret_cast<float*>(sB_lo + C::B_BYTES);
  uint64_t* mbar = reinterpret_cast<uint64_t*>(s_cst + TC_STATES);   // two barriers, one per stage
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbar + 2);
  const int tid = threadIdx.x, warp = tid >> 5;
  const int row = tid & (TC_M - 1), half = tid >> 7;   // `half`: which of the TCP_PARTS threads of a row
  const int tile = blockIdx.x, stripe = blockIdx.y;
  const int k0 = tile * TC_STATES;
  const int Tm = T - LL;
  const long long frames = (long long)N * Tm;
  const long long units = (long long)N * tile_nch;
  const long long nblk = (tile_cl > 0) ? ((units + 127) / 128) * tile_cl : (frames + TC_M - 1) / TC_M;
  {
    const uint4* src = reinterpret_cast<const uint4*>(tiles + (size_t)tile * 2 * C::B_BYTES);
    uint4* dst = reinterpret_cast<uint4*>(sB_hi);
    for (int i = tid; i < 2 * C::B_BYTES / 16; i += TCP_THREADS) dst[i] = src[i];
    if (tid < TC_STATES) s_cst[tid] = (k0 + tid < K) ? cst[k0 + tid] : 0.0f;
  }
  const uint32_t mbar_sa = (uint32_t)__cvta_generic_to_shared(mbar);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(mbar_sa));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(mbar_sa + 8));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::);
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"r"(
        (uint32_t)__cvta_generic_to_shared(tmem_slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::);
  }
  TC_FENCE_PROXY();
  TC_FENCE_BEFORE();
  __syncthreads();
  TC_FENCE_AFTER();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t a_sa = (uint32_t)__cvta_generic_to_shared(sA);
  const uint32_t b_hi_sa = (uint32_t)__cvta_generic_to_shared(sB_hi);
  const uint32_t b_lo_sa = (uint32_t)__cvta_generic_to_shared(sB_lo);

  // where row `row` of block blk lives: frame index g (row-major ll), liveness, start of its lag window
  auto locate = [&](long long blk, long long& g, const float*& xp) -> bool {
    g = blk * TC_M + row;
    xp = x;
    if (tile_cl > 0) {
      const long long b = blk / tile_cl;
      const int r = (int)(blk - b * tile_cl);
      const long long uu = b * 128 + row;
      if (uu >= units) return false;
      const int n = (int)(uu / tile_nch), cc = (int)(uu % tile_nch);
      const int tp = cc * tile_cl + r;
      if (tp >= Tm) return false;
      xp = x + ((size_t)n * T + tp) * DD;
      return true;
    }
    if (g >= frames) return false;
    const int n = (int)(g / Tm), tp = (int)(g % Tm);
    xp = x + ((size_t)n * T + tp) * DD;
    return true;
  };
  // this thread's chunks of its row: global -> registers
  auto load_row = [&](long long blk, float4 (&pre)[NCH]) {
    long long g;
    const float* xp;
    const bool live = locate(blk, g, xp);
#pragma unroll
    for (int j = 0; j < NCH; j++) {
      const int kc = TCP_PARTS * j + half;
      float4 v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
      if (live && kc < C::KC) {
        if ((DD & 1) == 0 && kc * 4 + 3 < DD * LL + DD) {
          const float2 a = __ldg(reinterpret_cast<const float2*>(xp + kc * 4));
          const float2 b = __ldg(reinterpret_cast<const float2*>(xp + kc * 4 + 2));
          v = make_float4(a.x, a.y, b.x, b.y);
        } else {
          float q4[4];
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const int p = kc * 4 + q;
            q4[q] = (p < DD * LL + DD) ? xp[p] : (p == C::P - 1 ? 1.0f : 0.0f);
          }
          v = make_float4(q4[0], q4[1], q4[2], q4[3]);
        }
      }
      pre[j] = v;
    }
  };
  // registers -> hi / lo TF32 parts in stage `st`
  auto store_row = [&](int st, const float4 (&pre)[NCH]) {
    unsigned char* ah = sA + (size_t)st * 2 * C::A_BYTES + (size_t)(row / 8) * C::SBO + (row % 8) * 16;
#pragma unroll
    for (int j = 0; j < NCH; j++) {
      const int kc = TCP_PARTS * j + half;
      if (kc >= C::KC) continue;
      const float v4[4] = {pre[j].x, pre[j].y, pre[j].z, pre[j].w};
      uint32_t hi[4], lo[4];
#pragma unroll
      for (int q = 0; q < 4; q++) {
        hi[q] = tf32_rna(v4[q]);
        lo[q] = tf32_rna(v4[q] - __uint_as_float(hi[q]));
      }
      *reinterpret_cast<uint4*>(ah + kc * C::LBO) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
      *reinterpret_cast<uint4*>(ah + C::A_BYTES + kc * C::LBO) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
  };
  // 3xTF32 (hi*hi + hi*lo + lo*hi) of stage `st` into accumulator `st`; one thread issues
  auto issue = [&](int st) {
    TC_FENCE_AFTER();
    const uint32_t ah = a_sa + (uint32_t)(st * 2 * C::A_BYTES), al = ah + (uint32_t)C::A_BYTES;
    uint32_t acc = 0;
#pragma unroll
    for (int ks = 0; ks < C::KD / 8; ks++) {
      const uint32_t koff = ks * 2 * C::LBO;
      const uint64_t dah = umma_desc(ah + koff, C::LBO, C::SBO), dal = umma_desc(al + koff, C::LBO, C::SBO);
      const uint64_t dbh = umma_desc(b_hi_sa + koff, C::LBO, C::SBO), dbl = umma_desc(b_lo_sa + koff, C::LBO, C::SBO);
#pragma unroll
      for (int term = 0; term < 3; term++) {
        umma_tf32(tmem_base + (uint32_t)(st * 256), (term == 2) ? dal : dah, (term == 1) ? dbl : dbh, C::IDESC, acc);
        acc = 1;
      }
    }
    umma_commit(mbar_sa + 8 * st);
  };

  const long long nmine = (nblk > stripe) ? (nblk - stripe + nstripes - 1) / nstripes : 0;
  float4 pre[NCH];
  if (nmine > 0) {
    load_row(stripe, pre);
    store_row(0, pre);
  }
  if (nmine > 1) {
    load_row(stripe + nstripes, pre);
    store_row(1, pre);
  }
  TC_FENCE_PROXY();
  TC_FENCE_BEFORE();
  __syncthreads();
  if (nmine > 0 && tid == 0) issue(0);
  uint32_t parity[2] = {0u, 0u};
  for (long long i = 0; i < nmine; i++) {
    const int st = (int)(i & 1);
    const long long blk = stripe + i * nstripes;
    if (i + 1 < nmine && tid == 0) issue(st ^ 1);          // block i+1: its A stage was completed before the last barrier
    if (i + 2 < nmine) load_row(blk + 2 * nstripes, pre);  // x of block i+2 on its way while we wait and read TMEM
    mbar_wait(mbar_sa + 8 * st, st ? parity[1] : parity[0]);
    if (st) parity[1] ^= 1u; else parity[0] ^= 1u;
    TC_FENCE_AFTER();
    {
      constexpr int HS = TC_STATES / TCP_PARTS, HC = HS * DD, W = (HC % 16 == 0) ? 16 : 8;
      static_assert(HC % 8 == 0 && HS % 4 == 0, "a thread reads whole 8-column groups and writes whole float4s");
      long long g;
      const float* xp;
      const bool live = locate(blk, g, xp);
      float q[HS];
#pragma unroll
      for (int sidx = 0; sidx < HS; sidx++) q[sidx] = 0.0f;
      const uint32_t taddr = tmem_base + (uint32_t)(st * 256) + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(half * HC);
      uint32_t r[HC];   // all of this thread's columns in flight at once, one wait
#pragma unroll
      for (int c0 = 0; c0 < HC; c0 += W) {
        if (W == 16) tmem_ld16_nowait(taddr + (uint32_t)c0, r + c0);
        else tmem_ld8_nowait(taddr + (uint32_t)c0, r + c0);
      }
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < HC; e++) asm volatile("" : "+r"(r[e]));   // no use of r may be scheduled above the wait
#pragma unroll
      for (int e = 0; e < HC; e++) {
        const float v = __uint_as_float(r[e]);
        q[e / DD] = fmaf(v, v, q[e / DD]);
      }
      const int ks = k0 + half * HS;
      const float* cs = s_cst + half * HS;
      if (live && tile_cl > 0) {
        float4* out4 = reinterpret_cast<float4*>(ll) + ((size_t)blk * (K >> 2) + (ks >> 2)) * 128 + row;
#pragma unroll
        for (int s4 = 0; s4 < HS; s4 += 4)
          if (ks + s4 < K)
            out4[(size_t)(s4 >> 2) * 128] =
                make_float4(cs[s4] - 0.5f * q[s4], cs[s4 + 1] - 0.5f * q[s4 + 1],
                            cs[s4 + 2] - 0.5f * q[s4 + 2], cs[s4 + 3] - 0.5f * q[s4 + 3]);
      } else if (live) {
        float* out = ll + g * K + ks;
        if ((K & 3) == 0 && ks + HS <= K) {
#pragma unroll
          for (int s4 = 0; s4 < HS; s4 += 4)
            *reinterpret_cast<float4*>(out + s4) =
                make_float4(cs[s4] - 0.5f * q[s4], cs[s4 + 1] - 0.5f * q[s4 + 1],
                            cs[s4 + 2] - 0.5f * q[s4 + 2], cs[s4 + 3] - 0.5f * q[s4 + 3]);
        } else {
#pragma unroll
          for (int sidx = 0; sidx < HS; sidx++)
            if (ks + sidx < K) out[sidx] = cs[sidx] - 0.5f * q[sidx];
        }
      }
    }
    // stage st is free (its MMAs completed): A tile of block i+2
    if (i + 2 < nmine) store_row(st, pre);
    TC_FENCE_PROXY();
    TC_FENCE_BEFORE();
    __syncthreads();
  }
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" ::"r"(tmem_base));
  }
}

size_t kpms_ar_tc_tile_bytes(int K, int D, int L) {
  const int P = D * L + D + 1, KD = (P + 7) / 8 * 8, SBO = (KD / 4) * 128;
  const int ntiles = (K + TC_STATES - 1) / TC_STATES;
  return kpms_round_up((size_t)ntiles * 2 * ((TC_STATES * D) / 8) * SBO, 256);
}

template <int DD, int LL>
static int launch_tc(const float* x, const float* Wf, const float* cst, int N, int T, int K, float* ll,
                     unsigned char* tiles, int tile_cl, int tile_nch, cudaStream_t st) {
  using C = TC<DD, LL>;
  const int ntiles = (K + TC_STATES - 1) / TC_STATES;
  ar_tc_pack_kernel<DD, LL><<<ntiles, 256, 0, st>>>(Wf, K, tiles);
  int rc = kpms_last_launch();
  if (rc) return rc;
  const long long nblk0 = (tile_cl > 0) ? (((long long)N * tile_nch + 127) / 128) * tile_cl
                                        : ((long long)N * (T - LL) + TC_M - 1) / TC_M;
  if (kpms_opts.emission_pipe && ntiles <= 148) {
    // pipelined kernel: one CTA per SM (it owns all 512 TMEM columns; the shared-memory request is kept
    // above half of an SM's so that two CTAs can never meet on one SM and wait for each other's columns)
    size_t smp = 1024 + 4 * C::A_BYTES + 2 * C::B_BYTES + TC_STATES * sizeof(float) + 32;
    if (smp < 120 * 1024) smp = 120 * 1024;
    if (smp <= 227 * 1024) {
      auto kp = ar_loglik_tc_pipe_kernel<DD, LL>;
      KPMS_CUDA(cudaFuncSetAttribute(kp, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smp));
      long long stp = 148 / ntiles;
      if (stp > nblk0) stp = nblk0;
      if (stp < 1) stp = 1;
      kp<<<dim3(ntiles, (unsigned)stp), 128 * TCP_PARTS_OF(DD), smp, st>>>(x, tiles, cst, N, T, K, (int)stp, tile_cl, tile_nch, ll);
      return kpms_last_launch();
    }
  }
  const size_t smem = 1024 + 2 * C::A_BYTES + 2 * C::B_BYTES + TC_STATES * sizeof(float) + 16;
  auto kern = ar_loglik_tc_kernel<DD, LL>;
  KPMS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long nblk = (tile_cl > 0) ? (((long long)N * tile_nch + 127) / 128) * tile_cl
                                       : ((long long)N * (T - LL) + TC_M - 1) / TC_M;
  // two CTAs per SM when the operand tiles allow it: one CTA's epilogue overlaps the other's MMAs
  long long stripes = (smem <= 112 * 1024 ? 2 : 1) * 148 / ntiles;
  if (stripes < 1) stripes = 1;
  if (stripes > nblk) stripes = nblk;
  if (ntiles > 65535) return KPMS_ERR_UNSUPPORTED;
  kern<<<dim3(ntiles, (unsigned)stripes), TC_THREADS, smem, st>>>(x, tiles, cst, N, T, K, (int)stripes, tile_cl, tile_nch, ll);
  return kpms_last_launch();
}

// tiles == nullptr or too small -> KPMS_ERR_UNSUPPORTED (caller falls back to the CUDA-core kernel)
// tile_cl > 0: write ll in the HMM's interleaved layout for that chunk length (needs K % 4 == 0)
int kpms_ar_loglik_tc(const float* x, const float* Wf, const float* cst, int N, int T, int D, int L,
                      int K, float* ll, void* tiles, size_t tile_bytes, int tile_cl, int tile_nch,
                      cudaStream_t st) {
  if (tile_cl > 0 && (K & 3)) return KPMS_ERR_UNSUPPORTED;
  if (tiles == nullptr || tile_bytes < kpms_ar_tc_tile_bytes(K, D, L) || !kpms_aligned(tiles, 256))
    return KPMS_ERR_UNSUPPORTED;
#define KPMS_X(d, l) \
  if (D == d && L == l) return launch_tc<d, l>(x, Wf, cst, N, T, K, ll, (unsigned char*)tiles, tile_cl, tile_nch, st);
  KPMS_FOR_EACH_DL(KPMS_X)
#undef KPMS_X
  return KPMS_ERR_UNSUPPORTED;
}
