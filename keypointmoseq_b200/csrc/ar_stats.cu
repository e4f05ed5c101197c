// a8: AR sufficient statistics  F[k] = sum_{t: z_t = k, unmasked} f_t f_t^T,
// f_t = [x[t-L..t-1] (oldest first), 1, x[t]]  (P = D*L + 1 + D entries).
// Reference: gibbs.py:129-131 (S_out_out, S_out_in, S_in_in), :141-143 (dense one-hot masks).
//
// The reference multiplies by a dense (K, frames) one-hot matrix.  Here the frames are first
// grouped by state with a deterministic counting sort, then every state's frame list is reduced
// with the FP64 tensor cores (mma.sync m8n8k4.f64):
//   1. hist     one warp per chunk of 512 frames counts its states (__match_any_sync, no atomics)
//   2. scan     per state: exclusive prefix over the chunks; plan: state starts, slices per state
//   3. scatter  the same warps recompute their ranks and write perm[] (window offset of each
//               frame), grouped by state, in frame order inside a state
//   4. syrk     CTA = (state, slice of its list): every warp streams groups of 4 frames, gathers
//               the regressor K-fragments straight from x (lane (g, qd) holds f_{4i+qd}[8b+g]),
//               and accumulates the lower block triangle of F in DMMA accumulators; warps are
//               combined in fixed order through shared memory
//   5. reduce   slices of a state are added in fixed order and mirrored into F (K, P, P)
// Every step is order-deterministic, so F is bit-reproducible run to run.
#include "common.cuh"

#define ARS_CH 512        // frames per warp chunk of the counting sort
#define ARS_SORT_WARPS 8
#define ARS_SYRK_WARPS 8
#define ARS_MAXPB 6       // regressor length up to 48

size_t kpms_ar_stats_scan_workspace_bytes(int N, int T, int D, int L, int K);  // stats.cu
int kpms_ar_stats_scan(const float* x, const int32_t* z, const float* mask, int N, int T, int D, int L,
                       int K, double* F, void* workspace, size_t workspace_bytes, void* stream);

struct ArsPlan {
  int* hist;        // [K][nchunks] counts, then exclusive offsets inside the state
  int* total;       // [K]
  int* start;       // [K + 1] first position of a state in perm
  int* slice_start; // [K + 1] first CTA of a state
  int* perm;        // [frames] window offsets n*T + t', grouped by state
  double* partial;  // [max_ctas][nblk * 64]
  int nchunks, slice, max_ctas;
};

static inline int ars_slice(long long frames) {
  long long s = (frames / (4 * 148) + 31) / 32 * 32;
  if (s < 1024) s = 1024;
  return (int)s;
}

static size_t ars_layout(int N, int T, int D, int L, int K, void* base, ArsPlan* pl) {
  const long long frames = (long long)N * (T - L);
  const int P = D * L + 1 + D, PB = (P + 7) / 8, nblk = PB * (PB + 1) / 2;
  const int nchunks = (int)((frames + ARS_CH - 1) / ARS_CH);
  const int slice = ars_slice(frames);
  const int max_ctas = (int)(frames / slice) + K + 1;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += kpms_round_up(bytes, 256);
    return o;
  };
  const size_t o_hist = take((size_t)K * nchunks * sizeof(int));
  const size_t o_total = take((size_t)K * sizeof(int));
  const size_t o_start = take((size_t)(K + 1) * sizeof(int));
  const size_t o_ss = take((size_t)(K + 1) * sizeof(int));
  const size_t o_perm = take((size_t)frames * sizeof(int));
  const size_t o_part = take((size_t)max_ctas * nblk * 64 * sizeof(double));
  if (pl) {
    char* b = (char*)base;
    pl->hist = (int*)(b + o_hist);
    pl->total = (int*)(b + o_total);
    pl->start = (int*)(b + o_start);
    pl->slice_start = (int*)(b + o_ss);
    pl->perm = (int*)(b + o_perm);
    pl->partial = (double*)(b + o_part);
    pl->nchunks = nchunks;
    pl->slice = slice;
    pl->max_ctas = max_ctas;
  }
  return off;
}

static inline bool ars_fast_path(int N, int T, int D, int L, int K) {
  const int P = D * L + 1 + D;
  return (P + 7) / 8 <= ARS_MAXPB && K <= 1024 && (long long)N * T < (1LL << 31);
}

extern "C" size_t kpms_ar_stats_workspace_bytes(int N, int T, int D, int L, int K) {
  if (N <= 0 || T <= L || D <= 0 || L <= 0 || K <= 0) return 0;
  if (!ars_fast_path(N, T, D, L, K)) return kpms_ar_stats_scan_workspace_bytes(N, T, D, L, K);
  return ars_layout(N, T, D, L, K, nullptr, nullptr);
}

// ------------------------------------------------------------------ 1 + 3: counting sort passes
// SCATTER = false: hist[k][chunk] = number of frames of state k in the chunk.
// SCATTER = true : perm[start[k] + hist[k][chunk] + rank] = n*T + t'   (hist now holds offsets).
template <bool SCATTER>
__global__ void __launch_bounds__(ARS_SORT_WARPS * 32)
    ars_sort_kernel(const int* __restrict__ z, const float* __restrict__ mask, int T, int L, int K,
                    long long frames, ArsPlan pl) {
  extern __shared__ int cnt_s[];  // [ARS_SORT_WARPS][K]
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int chunk = blockIdx.x * ARS_SORT_WARPS + w;
  int* cnt = cnt_s + (size_t)w * K;
  for (int k = lane; k < K; k += 32) cnt[k] = 0;
  __syncwarp();
  if (chunk >= pl.nchunks) return;
  const int Tm = T - L;
  const unsigned lt = (1u << lane) - 1u;
  for (int r = 0; r < ARS_CH / 32; r++) {
    const long long g = (long long)chunk * ARS_CH + r * 32 + lane;
    int key = -1;
    int woff = 0;
    if (g < frames) {
      const int n = (int)(g / Tm), tp = (int)(g % Tm);
      const int zz = z[g];
      if (mask[(size_t)n * T + tp + L] > 0.0f && zz >= 0 && zz < K) key = zz;
      woff = n * T + tp;
    }
    const unsigned m = __match_any_sync(0xffffffffu, key);
    int rank = 0;
    if (key >= 0) rank = cnt[key] + __popc(m & lt);
    __syncwarp();
    if (key >= 0 && (m & lt) == 0u) cnt[key] += __popc(m);  // first lane of each group
    __syncwarp();
    if (SCATTER && key >= 0)
      pl.perm[pl.start[key] + pl.hist[(size_t)key * pl.nchunks + chunk] + rank] = woff;
  }
  if (!SCATTER)
    for (int k = lane; k < K; k += 32) pl.hist[(size_t)k * pl.nchunks + chunk] = cnt[k];
}

// ------------------------------------------------------------------ 2: scan + plan
__global__ void __launch_bounds__(256) ars_scan_kernel(ArsPlan pl) {
  const int k = blockIdx.x, tid = threadIdx.x;
  int* h = pl.hist + (size_t)k * pl.nchunks;
  const int per = (pl.nchunks + 255) / 256;
  const int b = min(tid * per, pl.nchunks), e = min(b + per, pl.nchunks);
  int s = 0;
  for (int i = b; i < e; i++) s += h[i];
  __shared__ int part[256];
  part[tid] = s;
  __syncthreads();
  if (tid == 0) {
    int run = 0;
    for (int i = 0; i < 256; i++) {
      const int v = part[i];
      part[i] = run;
      run += v;
    }
    pl.total[k] = run;
  }
  __syncthreads();
  int run = part[tid];
  for (int i = b; i < e; i++) {
    const int v = h[i];
    h[i] = run;
    run += v;
  }
}

__global__ void ars_plan_kernel(ArsPlan pl, int K) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  int pos = 0, cta = 0;
  for (int k = 0; k < K; k++) {
    pl.start[k] = pos;
    pl.slice_start[k] = cta;
    pos += pl.total[k];
    cta += (pl.total[k] + pl.slice - 1) / pl.slice;
  }
  pl.start[K] = pos;
  pl.slice_start[K] = cta;
}

// ------------------------------------------------------------------ 4: per-state SYRK on DMMA
template <int PB>
__global__ void __launch_bounds__(ARS_SYRK_WARPS * 32, 1)
    ars_syrk_kernel(const float* __restrict__ x, int D, int L, int K, ArsPlan pl) {
  constexpr int NBLK = PB * (PB + 1) / 2;
  extern __shared__ __align__(16) double red[];  // [ARS_SYRK_WARPS][NBLK][64]
  __shared__ int seg[3];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int g = lane >> 2, qd = lane & 3;
  const int DL = D * L, P = DL + 1 + D;
  if (threadIdx.x == 0) {
    const int b = blockIdx.x;
    int lo = 0, hi = K;  // last k with slice_start[k] <= b
    if (b >= pl.slice_start[K]) {
      seg[0] = -1;
    } else {
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (pl.slice_start[mid] <= b) lo = mid;
        else hi = mid;
      }
      // states without frames own no CTA: skip forward to the owner of b
      while (pl.slice_start[lo + 1] <= b) lo++;
      const int j = b - pl.slice_start[lo];
      seg[0] = lo;
      seg[1] = pl.start[lo] + j * pl.slice;
      seg[2] = min(pl.start[lo] + pl.total[lo], seg[1] + pl.slice);
    }
  }
  __syncthreads();
  if (seg[0] < 0) return;
  const int beg = seg[1], end = seg[2];

  double acc[NBLK][2];
#pragma unroll
  for (int e = 0; e < NBLK; e++) acc[e][0] = acc[e][1] = 0.0;

  // element p = 8b + g of the regressor: x window element (p < DL: p, p > DL: p - 1), the
  // constant 1 at p == DL, zero padding beyond P
  int src[PB];
  float cst[PB];
#pragma unroll
  for (int b = 0; b < PB; b++) {
    const int p = 8 * b + g;
    src[b] = (p < DL) ? p : ((p > DL && p < P) ? p - 1 : -1);
    cst[b] = (p == DL) ? 1.0f : 0.0f;
  }
  constexpr int GB = 4;                                  // groups of 4 frames per batch
  const int nbatch = (end - beg + 4 * GB - 1) / (4 * GB);
  auto load_perm = [&](int bt) {                         // lanes 0..15: the 16 frames of batch bt
    const int idx = beg + bt * 4 * GB + lane;
    return (bt < nbatch && lane < 4 * GB && idx < end) ? __ldg(pl.perm + idx) : -1;
  };
  float fr[GB][PB];
  auto load_x = [&](int pv) {
#pragma unroll
    for (int gi = 0; gi < GB; gi++) {
      const int off = __shfl_sync(0xffffffffu, pv, 4 * gi + qd);
      const float* win = x + (size_t)max(off, 0) * D;
#pragma unroll
      for (int b = 0; b < PB; b++)
        fr[gi][b] = (off >= 0) ? ((src[b] >= 0) ? __ldg(win + src[b]) : cst[b]) : 0.0f;
    }
  };
  int pv_next = load_perm(w);
  int bt = w;
  if (bt < nbatch) {
    load_x(pv_next);
    pv_next = load_perm(bt + ARS_SYRK_WARPS);
  }
  for (; bt < nbatch; bt += ARS_SYRK_WARPS) {
    double fd[GB][PB];
#pragma unroll
    for (int gi = 0; gi < GB; gi++)
#pragma unroll
      for (int b = 0; b < PB; b++) fd[gi][b] = (double)fr[gi][b];
    // fetch the next batch of this warp while the tensor cores work on the current one
    const int pv = pv_next;
    pv_next = load_perm(bt + 2 * ARS_SYRK_WARPS);
    if (bt + ARS_SYRK_WARPS < nbatch) load_x(pv);
#pragma unroll
    for (int gi = 0; gi < GB; gi++)
#pragma unroll
      for (int i = 0; i < PB; i++)
#pragma unroll
        for (int j = 0; j <= i; j++)
          asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                       : "+d"(acc[i * (i + 1) / 2 + j][0]), "+d"(acc[i * (i + 1) / 2 + j][1])
                       : "d"(fd[gi][i]), "d"(fd[gi][j]));
  }
  // combine the warps in fixed order; accumulator layout: row g, columns 2qd, 2qd+1
#pragma unroll
  for (int e = 0; e < NBLK; e++)
    *reinterpret_cast<double2*>(red + ((size_t)w * NBLK + e) * 64 + g * 8 + 2 * qd) = make_double2(acc[e][0], acc[e][1]);
  __syncthreads();
  double* out = pl.partial + (size_t)blockIdx.x * NBLK * 64;
  for (int e = threadIdx.x; e < NBLK * 64; e += blockDim.x) {
    double s = 0.0;
#pragma unroll
    for (int ww = 0; ww < ARS_SYRK_WARPS; ww++) s += red[(size_t)ww * NBLK * 64 + e];
    out[e] = s;
  }
}

// ------------------------------------------------------------------ 5: slices -> F (K, P, P)
__global__ void __launch_bounds__(256) ars_reduce_kernel(ArsPlan pl, int P, int PB, double* __restrict__ F) {
  const int k = blockIdx.x;
  const int nblk = PB * (PB + 1) / 2;
  const int c0 = pl.slice_start[k], c1 = pl.slice_start[k + 1];
  for (int e = threadIdx.x; e < nblk * 64; e += blockDim.x) {
    const int blk = e >> 6, r = (e >> 3) & 7, c = e & 7;
    int bi = 0;
    while ((bi + 1) * (bi + 2) / 2 <= blk) bi++;
    const int bj = blk - bi * (bi + 1) / 2;
    const int i = 8 * bi + r, j = 8 * bj + c;
    if (i >= P || j >= P || j > i) continue;
    double s = 0.0;
    for (int cta = c0; cta < c1; cta++) s += pl.partial[(size_t)cta * nblk * 64 + e];
    F[((size_t)k * P + i) * P + j] = s;
    F[((size_t)k * P + j) * P + i] = s;
  }
}

template <int PB>
static int launch_syrk(const float* x, int D, int L, int K, const ArsPlan& pl, cudaStream_t st) {
  constexpr int NBLK = PB * (PB + 1) / 2;
  const size_t smem = (size_t)ARS_SYRK_WARPS * NBLK * 64 * sizeof(double);
  auto kern = ars_syrk_kernel<PB>;
  KPMS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<pl.max_ctas, ARS_SYRK_WARPS * 32, smem, st>>>(x, D, L, K, pl);
  return kpms_last_launch();
}

extern "C" int kpms_ar_stats(const float* x, const int32_t* z, const float* mask, int N, int T,
                             int D, int L, int K, double* F, void* workspace,
                             size_t workspace_bytes, void* stream) {
  KPMS_REQUIRE(x && z && mask && F && workspace, KPMS_ERR_NULL);
  KPMS_REQUIRE(N > 0 && T > L && D > 0 && L > 0 && K > 0, KPMS_ERR_SHAPE);
  if (!ars_fast_path(N, T, D, L, K))
    return kpms_ar_stats_scan(x, z, mask, N, T, D, L, K, F, workspace, workspace_bytes, stream);
  KPMS_REQUIRE(kpms_aligned(workspace, 256) &&
                   workspace_bytes >= kpms_ar_stats_workspace_bytes(N, T, D, L, K),
               KPMS_ERR_WORKSPACE);
  cudaStream_t st = (cudaStream_t)stream;
  ArsPlan pl;
  ars_layout(N, T, D, L, K, workspace, &pl);
  const long long frames = (long long)N * (T - L);
  const int P = D * L + 1 + D, PB = (P + 7) / 8;
  const int sort_ctas = (pl.nchunks + ARS_SORT_WARPS - 1) / ARS_SORT_WARPS;
  const size_t sort_smem = (size_t)ARS_SORT_WARPS * K * sizeof(int);
  ars_sort_kernel<false><<<sort_ctas, ARS_SORT_WARPS * 32, sort_smem, st>>>(z, mask, T, L, K, frames, pl);
  int rc = kpms_last_launch();
  if (rc) return rc;
  ars_scan_kernel<<<K, 256, 0, st>>>(pl);
  rc = kpms_last_launch();
  if (rc) return rc;
  ars_plan_kernel<<<1, 32, 0, st>>>(pl, K);
  rc = kpms_last_launch();
  if (rc) return rc;
  ars_sort_kernel<true><<<sort_ctas, ARS_SORT_WARPS * 32, sort_smem, st>>>(z, mask, T, L, K, frames, pl);
  rc = kpms_last_launch();
  if (rc) return rc;
  switch (PB) {
    case 1: rc = launch_syrk<1>(x, D, L, K, pl, st); break;
    case 2: rc = launch_syrk<2>(x, D, L, K, pl, st); break;
    case 3: rc = launch_syrk<3>(x, D, L, K, pl, st); break;
    case 4: rc = launch_syrk<4>(x, D, L, K, pl, st); break;
    case 5: rc = launch_syrk<5>(x, D, L, K, pl, st); break;
    default: rc = launch_syrk<6>(x, D, L, K, pl, st); break;
  }
  if (rc) return rc;
  ars_reduce_kernel<<<K, 256, 0, st>>>(pl, P, PB, F);
  return kpms_last_launch();
}
