// Frame-parallel observation passes and the location sampler.
//   a13 obs-variance statistic  (gibbs.py:84-94)
//   a14 scales                  (gibbs.py:117-125, distributions.py:8-15)
//   a15 heading                 (gibbs.py:25-35, util.py:66-82, distributions.py:5-6)
//   a16 location                (gibbs.py:38-54 + kalman.py:80-156 specialised to the
//                                isotropic A=I, C=I, Q=sigmasq_loc*I, R=gammasq*I case)
// All share the prologue  Ybar = Cdfull [x;1]  (Cdfull = kron(Gamma,I) Cd, gibbs.py:14,29)
// and the rigid transform of util.py:112-166.  Arithmetic is fp64; storage fp32.
#include "common.cuh"

#define OBS_XMAX 16   // latent dim D <= 16
#define OBS_KPMAX 32  // keypoints k <= 32
#define OBS_DKMAX 3   // keypoint dim dk in {2,3}

struct ObsDims {
  int N, T, k, dk, D;
};

static inline int obs_check(int N, int T, int k, int dk, int D) {
  if (N <= 0 || T <= 0 || k <= 0 || D <= 0) return KPMS_ERR_SHAPE;
  if (dk < 2 || dk > OBS_DKMAX || k > OBS_KPMAX || D > OBS_XMAX) return KPMS_ERR_UNSUPPORTED;
  return KPMS_OK;
}

// shared-memory copy of Cdfull (k*dk x (D+1)) and per-keypoint constants
__device__ __forceinline__ void load_cd(const double* __restrict__ Cdfull, int rows, int D,
                                        double* Cs) {
  for (int i = threadIdx.x; i < rows * (D + 1); i += blockDim.x) Cs[i] = Cdfull[i];
}

__device__ __forceinline__ void load_x(const float* __restrict__ x, size_t frame, int D,
                                       double (&xr)[OBS_XMAX]) {
#pragma unroll
  for (int d = 0; d < OBS_XMAX; d++) xr[d] = (d < D) ? (double)x[frame * D + d] : 0.0;
}

// Ybar for one keypoint: yb[dim] = Cs[(kp*dk+dim)] . [x;1]
__device__ __forceinline__ void ybar_kp(const double* Cs, int kp, int dk, int D,
                                        const double (&xr)[OBS_XMAX], double (&yb)[OBS_DKMAX]) {
#pragma unroll
  for (int dim = 0; dim < OBS_DKMAX; dim++) {
    yb[dim] = 0.0;
    if (dim < dk) {
      const double* row = Cs + (size_t)(kp * dk + dim) * (D + 1);
      double a = row[D];
#pragma unroll
      for (int d = 0; d < OBS_XMAX; d++)
        if (d < D) a = fma(row[d], xr[d], a);
      yb[dim] = a;
    }
  }
}

// Ytild = R(-h)(Y - v) for one keypoint
__device__ __forceinline__ void ytild_kp(const float* __restrict__ Yf, const double (&vv)[OBS_DKMAX],
                                         double c, double s, int dk, double (&yt)[OBS_DKMAX]) {
  double d0 = (double)Yf[0] - vv[0], d1 = (double)Yf[1] - vv[1];
  rot_inv(c, s, d0, d1, yt[0], yt[1]);
  yt[2] = (dk > 2) ? (double)Yf[2] - vv[2] : 0.0;
}

__device__ __forceinline__ void ytild3(float y0, float y1, float y2, const double (&vv)[OBS_DKMAX], double c, double s,
                                       int dk, double (&yt)[OBS_DKMAX]) {
  rot_inv(c, s, (double)y0 - vv[0], (double)y1 - vv[1], yt[0], yt[1]);
  yt[2] = (dk > 2) ? (double)y2 - vv[2] : 0.0;
}

// x row as float2 pairs when D is even (rows are then 8-byte aligned)
__device__ __forceinline__ void load_x2(const float* __restrict__ x, size_t frame, int D, double (&xr)[OBS_XMAX]) {
  if ((D & 1) == 0) {
    const float2* p = reinterpret_cast<const float2*>(x + frame * D);
#pragma unroll
    for (int d = 0; d < OBS_XMAX / 2; d++) {
      float2 v = make_float2(0.f, 0.f);
      if (2 * d < D) v = __ldg(p + d);
      xr[2 * d] = (double)v.x;
      xr[2 * d + 1] = (double)v.y;
    }
    if (OBS_XMAX & 1) xr[OBS_XMAX - 1] = 0.0;
  } else {
    load_x(x, frame, D, xr);
  }
}

// ======================================================================== a13
#define OV_BLOCKS (148 * 4)
#define OV_THREADS 256

extern "C" size_t kpms_obs_variance_stats_workspace_bytes(int N, int T, int k) {
  if (N <= 0 || T <= 0 || k <= 0) return 0;
  return kpms_round_up((size_t)OV_BLOCKS * (OBS_KPMAX + 1) * sizeof(double), 256);
}

__global__ void __launch_bounds__(OV_THREADS)
    obs_variance_kernel(const float* __restrict__ Y, const float* __restrict__ mask,
                        const float* __restrict__ v, const float* __restrict__ h,
                        const float* __restrict__ x, const float* __restrict__ s,
                        const double* __restrict__ Cdfull, ObsDims dm,
                        double* __restrict__ partial) {
  extern __shared__ __align__(16) double Cs[];
  __shared__ double red[OV_THREADS / 32];
  const int k = dm.k, dk = dm.dk, D = dm.D;
  load_cd(Cdfull, k * dk, D, Cs);
  __syncthreads();
  // per-thread accumulators live in shared memory ([entry][thread], conflict-free): keeping 33
  // doubles in registers across the unrolled keypoint loop costs 255 registers and occupancy
  double* accs = Cs + (size_t)k * dk * (D + 1);
  for (int i = 0; i <= k; i++) accs[i * OV_THREADS + threadIdx.x] = 0.0;
  const long long frames = (long long)dm.N * dm.T;
  const bool vec = kp4_vec_ok(Y, s, k, dk);
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < frames;
       g += (long long)gridDim.x * blockDim.x) {
    const double m = (double)mask[g];
    if (m > 0.0) {
      double xr[OBS_XMAX];
      load_x2(x, (size_t)g, D, xr);
      double sn, cs;
      sincos((double)h[g], &sn, &cs);
      double vv[OBS_DKMAX];
#pragma unroll
      for (int dim = 0; dim < OBS_DKMAX; dim++) vv[dim] = (dim < dk) ? (double)v[g * dk + dim] : 0.0;
      for (int kp0 = 0; kp0 < k; kp0 += 4) {
        Kp4 q;
        load_kp4(Y + (size_t)g * k * dk, s + (size_t)g * k, kp0, k, dk, vec, q);
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const int kp = kp0 + j;
          if (kp < k) {
            double yb[OBS_DKMAX], yt[OBS_DKMAX];
            ybar_kp(Cs, kp, dk, D, xr, yb);
            ytild3(KP4_Y(q, j, dk, 0), KP4_Y(q, j, dk, 1), KP4_Y(q, j, dk, 2), vv, cs, sn, dk, yt);
            double e = 0.0;
#pragma unroll
            for (int dim = 0; dim < OBS_DKMAX; dim++) {
              const double r = yt[dim] - yb[dim];
              e = fma(r, r, e);
            }
            accs[kp * OV_THREADS + threadIdx.x] += m * e * rcp_pos_d((double)q.s[j]);   // same arithmetic as scales_stats_kernel
          }
        }
      }
      accs[k * OV_THREADS + threadIdx.x] += m;
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = 0; i <= k; i++) {
    double w = warp_sum_d(accs[i * OV_THREADS + threadIdx.x]);
    __syncthreads();
    if (lane == 0) red[warp] = w;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int q = 0; q < OV_THREADS / 32; q++) t += red[q];
      partial[(size_t)blockIdx.x * (OBS_KPMAX + 1) + (i < k ? i : OBS_KPMAX)] = t;
    }
  }
}

__global__ void obs_variance_reduce_kernel(const double* __restrict__ partial, int blocks, int k,
                                           double* __restrict__ S_y, double* __restrict__ msum) {
  const int i = threadIdx.x;
  if (i < k) {
    double t = 0.0;
    for (int b = 0; b < blocks; b++) t += partial[(size_t)b * (OBS_KPMAX + 1) + i];
    S_y[i] = t;
  } else if (i == OBS_KPMAX) {
    double t = 0.0;
    for (int b = 0; b < blocks; b++) t += partial[(size_t)b * (OBS_KPMAX + 1) + OBS_KPMAX];
    msum[0] = t;
  }
}

extern "C" int kpms_obs_variance_stats(const float* Y, const float* mask, const float* v,
                                       const float* h, const float* x, const float* s,
                                       const double* Cdfull, int N, int T, int k, int dk, int D,
                                       double* S_y, double* mask_sum, void* workspace,
                                       size_t workspace_bytes, void* stream) {
  KPMS_REQUIRE(Y && mask && v && h && x && s && Cdfull && S_y && mask_sum && workspace,
               KPMS_ERR_NULL);
  int rc = obs_check(N, T, k, dk, D);
  if (rc) return rc;
  KPMS_REQUIRE(kpms_aligned(workspace, 256) &&
                   workspace_bytes >= kpms_obs_variance_stats_workspace_bytes(N, T, k),
               KPMS_ERR_WORKSPACE);
  cudaStream_t st = (cudaStream_t)stream;
  ObsDims dm{N, T, k, dk, D};
  const size_t smem = ((size_t)k * dk * (D + 1) + (size_t)(k + 1) * OV_THREADS) * sizeof(double);
  KPMS_CUDA(cudaFuncSetAttribute(obs_variance_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  obs_variance_kernel<<<OV_BLOCKS, OV_THREADS, smem, st>>>(Y, mask, v, h, x, s, Cdfull, dm,
                                                          (double*)workspace);
  rc = kpms_last_launch();
  if (rc) return rc;
  obs_variance_reduce_kernel<<<1, OBS_KPMAX + 1, 0, st>>>((const double*)workspace, OV_BLOCKS, k,
                                                         S_y, mask_sum);
  return kpms_last_launch();
}

// ======================================================================== (f)1: moments for Cd
// Reference: gibbs.py:58-82 (resample_obs_params).  The reference materialises per-frame
// (k-1)x(k-1) and k*d x (D+1)(k-1)d temporaries; all it needs from the data are, per keypoint j,
//   W[j] = sum_f w_fj xt_f xt_f^T   (D+1 x D+1),   V[j] = sum_f w_fj xt_f Ytild_fj^T   (D+1 x dk),
//   w_fj = mask_f / (s_fj sigmasq_j),  xt = [x; 1],  Ytild = R(-h)(Y - v)        (util.py:140-166)
// (S_xx = sum_j Gamma_j Gamma_j^T (x) W[j], S_yx = sum_j Gamma_j (x) V[j]; the 330-dim Gaussian draw
// stays with the host-side dense algebra).  CTA = tile of OP_TILE frames staged in shared memory in
// fp64; thread <-> up to OP_PAIRS (keypoint, row a) pairs, whose D+1+dk accumulators stay in
// registers for the whole launch; per-CTA partials are combined in fixed order (reproducible).
#define OP_THREADS 256
#define OP_TILE 64
#define OP_PAIRS 3
#define OP_BLOCKS (148 * 2)
#define OP_ACC (OBS_XMAX + 1 + OBS_DKMAX)

extern "C" size_t kpms_obs_params_stats_workspace_bytes(int N, int T, int k, int dk, int D) {
  if (N <= 0 || T <= 0 || k <= 0 || D <= 0) return 0;
  (void)dk;
  return kpms_round_up((size_t)OP_BLOCKS * k * (D + 1) * OP_ACC * sizeof(double), 256);
}

__global__ void __launch_bounds__(OP_THREADS)
    obs_params_stats_kernel(const float* __restrict__ Y, const float* __restrict__ mask,
                            const float* __restrict__ v, const float* __restrict__ h,
                            const float* __restrict__ x, const float* __restrict__ s,
                            const double* __restrict__ sigmasq, ObsDims dm, double* __restrict__ partial) {
  extern __shared__ __align__(16) double sm[];
  const int k = dm.k, dk = dm.dk, D = dm.D, D1 = D + 1;
  double* xt = sm;                         // [OP_TILE][D1]
  double* yt = xt + OP_TILE * D1;          // [OP_TILE][k*dk]
  double* wt = yt + OP_TILE * k * dk;      // [OP_TILE][k]
  double* rot = wt + OP_TILE * k;          // [OP_TILE][2 + dk + 1]: cos, sin, v, mask
  double* isg = rot + OP_TILE * (3 + dk);  // [k] 1 / sigmasq
  for (int i = threadIdx.x; i < k; i += blockDim.x) isg[i] = 1.0 / sigmasq[i];
  const int npairs = k * D1;
  int pj[OP_PAIRS], pa[OP_PAIRS];
  double acc[OP_PAIRS][OP_ACC];
#pragma unroll
  for (int q = 0; q < OP_PAIRS; q++) {
    const int p = threadIdx.x + q * OP_THREADS;
    pj[q] = (p < npairs) ? p / D1 : -1;
    pa[q] = (p < npairs) ? p % D1 : 0;
#pragma unroll
    for (int e = 0; e < OP_ACC; e++) acc[q][e] = 0.0;
  }
  const long long frames = (long long)dm.N * dm.T;
  for (long long f0 = (long long)blockIdx.x * OP_TILE; f0 < frames; f0 += (long long)gridDim.x * OP_TILE) {
    const int rows = (int)min((long long)OP_TILE, frames - f0);
    __syncthreads();  // previous tile fully consumed
    for (int r = threadIdx.x; r < OP_TILE; r += blockDim.x) {
      const long long f = f0 + r;
      double* ro = rot + r * (3 + dk);
      if (r < rows) {
        double sn, cs;
        sincos((double)h[f], &sn, &cs);
        ro[0] = cs;
        ro[1] = sn;
        for (int c = 0; c < dk; c++) ro[2 + c] = (double)v[f * dk + c];
        ro[2 + dk] = (double)mask[f];
        for (int a = 0; a < D; a++) xt[r * D1 + a] = (double)x[f * D + a];
        xt[r * D1 + D] = 1.0;
      } else {
        ro[0] = 1.0;
        ro[1] = 0.0;
        for (int c = 0; c < dk; c++) ro[2 + c] = 0.0;
        ro[2 + dk] = 0.0;
        for (int a = 0; a < D1; a++) xt[r * D1 + a] = 0.0;
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < OP_TILE * k; i += blockDim.x) {   // coalesced over (frame, keypoint)
      const int r = i / k, j = i - r * k;
      const double* ro = rot + r * (3 + dk);
      double w = 0.0, y0 = 0.0, y1 = 0.0, y2 = 0.0;
      if (r < rows) {
        const long long f = f0 + r;
        const float* Yf = Y + ((size_t)f * k + j) * dk;
        rot_inv(ro[0], ro[1], (double)Yf[0] - ro[2], (double)Yf[1] - ro[3], y0, y1);
        if (dk > 2) y2 = (double)Yf[2] - ro[4];
        w = ro[2 + dk] * isg[j] / (double)s[(size_t)f * k + j];
      }
      wt[r * k + j] = w;
      yt[(r * k + j) * dk + 0] = y0;
      yt[(r * k + j) * dk + 1] = y1;
      if (dk > 2) yt[(r * k + j) * dk + 2] = y2;
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < OP_PAIRS; q++) {
      if (pj[q] < 0) continue;
      const int j = pj[q], a = pa[q];
      for (int r = 0; r < OP_TILE; r++) {
        const double wa = wt[r * k + j] * xt[r * D1 + a];
#pragma unroll
        for (int b = 0; b <= OBS_XMAX; b++)
          if (b < D1) acc[q][b] = fma(wa, xt[r * D1 + b], acc[q][b]);
#pragma unroll
        for (int c = 0; c < OBS_DKMAX; c++)
          if (c < dk) acc[q][OBS_XMAX + 1 + c] = fma(wa, yt[(r * k + j) * dk + c], acc[q][OBS_XMAX + 1 + c]);
      }
    }
  }
#pragma unroll
  for (int q = 0; q < OP_PAIRS; q++) {
    const int p = threadIdx.x + q * OP_THREADS;
    if (p < npairs) {
      double* o = partial + ((size_t)blockIdx.x * npairs + p) * OP_ACC;
#pragma unroll
      for (int e = 0; e < OP_ACC; e++) o[e] = acc[q][e];
    }
  }
}

__global__ void obs_params_reduce_kernel(const double* __restrict__ partial, int blocks, int k, int D, int dk,
                                         double* __restrict__ W, double* __restrict__ V) {
  const int D1 = D + 1, npairs = k * D1;
  const int p = blockIdx.x, e = threadIdx.x;  // pair, accumulator entry
  if (p >= npairs || e >= OP_ACC) return;
  double t = 0.0;
  for (int b = 0; b < blocks; b++) t += partial[((size_t)b * npairs + p) * OP_ACC + e];
  const int j = p / D1, a = p % D1;
  if (e < D1) W[((size_t)j * D1 + a) * D1 + e] = t;
  else if (e >= OBS_XMAX + 1 && e - (OBS_XMAX + 1) < dk) V[((size_t)j * D1 + a) * dk + (e - (OBS_XMAX + 1))] = t;
}

extern "C" int kpms_obs_params_stats(const float* Y, const float* mask, const float* v, const float* h,
                                     const float* x, const float* s, const double* sigmasq, int N, int T,
                                     int k, int dk, int D, double* W, double* V, void* workspace,
                                     size_t workspace_bytes, void* stream) {
  KPMS_REQUIRE(Y && mask && v && h && x && s && sigmasq && W && V && workspace, KPMS_ERR_NULL);
  int rc = obs_check(N, T, k, dk, D);
  if (rc) return rc;
  KPMS_REQUIRE(k * (D + 1) <= OP_PAIRS * OP_THREADS, KPMS_ERR_UNSUPPORTED);
  KPMS_REQUIRE(kpms_aligned(workspace, 256) &&
                   workspace_bytes >= kpms_obs_params_stats_workspace_bytes(N, T, k, dk, D),
               KPMS_ERR_WORKSPACE);
  cudaStream_t st = (cudaStream_t)stream;
  ObsDims dm{N, T, k, dk, D};
  const size_t smem = ((size_t)OP_TILE * ((D + 1) + k * dk + k + 3 + dk) + k) * sizeof(double);
  KPMS_CUDA(cudaFuncSetAttribute(obs_params_stats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  obs_params_stats_kernel<<<OP_BLOCKS, OP_THREADS, smem, st>>>(Y, mask, v, h, x, s, sigmasq, dm,
                                                              (double*)workspace);
  rc = kpms_last_launch();
  if (rc) return rc;
  obs_params_reduce_kernel<<<k * (D + 1), 32, 0, st>>>((const double*)workspace, OP_BLOCKS, k, D, dk, W, V);
  return kpms_last_launch();
}

// ======================================================================== (f)2: log joint
// Reference: util.py:232-275 (log_joint_likelihood): five masked sums over frames.  One thread
// per frame evaluates its share of every term; per-CTA partials, fixed-order reduction.
//   'Y': -1/2 |Ytild - Ybar|^2 / (s sigmasq) - dk/2 log(2 pi s sigmasq)       weight mask[t]
//   'x': log N(x_t; Ab[z] [x_{t-L..t-1}; 1], Q[z])                            weight mask[t], t >= L
//   'z': log pi[z_{t-L}, z_{t-L+1}]                                           weight mask[t+1], t >= L
//   'v': -|v_t - v_{t+1}|^2 / (2 sigmasq_loc) - dk/2 log(2 pi sigmasq_loc)    weight mask[t+1]
//   's': -nu_s s_0 / (2 s) - (1 + nu_s/2) log s                               weight mask[t]
#define LJ_THREADS 256
#define LJ_BLOCKS (148 * 4)

extern "C" size_t kpms_log_joint_workspace_bytes(void) { return kpms_round_up((size_t)LJ_BLOCKS * 5 * sizeof(double), 256); }

__global__ void __launch_bounds__(LJ_THREADS)
    log_joint_kernel(const float* __restrict__ Y, const float* __restrict__ mask, const float* __restrict__ x,
                     const float* __restrict__ s, const float* __restrict__ v, const float* __restrict__ h,
                     const int* __restrict__ z, const double* __restrict__ pi, const double* __restrict__ Ab,
                     const double* __restrict__ Qinv, const double* __restrict__ logdetQ,
                     const double* __restrict__ Cdfull, const double* __restrict__ sigmasq, double sigmasq_loc,
                     double nu_s, double s_0, ObsDims dm, int L, int K, double* __restrict__ partial) {
  extern __shared__ __align__(16) double Cs[];
  __shared__ double red[LJ_THREADS / 32][5];
  const int k = dm.k, dk = dm.dk, D = dm.D, T = dm.T, Tm = T - L, P1 = D * L + 1;
  double* sig = Cs + (size_t)k * dk * (D + 1);
  load_cd(Cdfull, k * dk, D, Cs);
  for (int i = threadIdx.x; i < k; i += blockDim.x) sig[i] = sigmasq[i];
  __syncthreads();
  double acc[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
  const double lv = -0.5 * (double)dk * log(2.0 * KPMS_PI * sigmasq_loc);
  const long long frames = (long long)dm.N * T;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < frames;
       g += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(g / T), t = (int)(g % T);
    const double m = (double)mask[g];
    const double m1 = (t + 1 < T) ? (double)mask[g + 1] : 0.0;
    if (m != 0.0) {
      double xr[OBS_XMAX];
      load_x(x, (size_t)g, D, xr);
      double sn, cs;
      sincos((double)h[g], &sn, &cs);
      double vv[OBS_DKMAX];
#pragma unroll
      for (int dim = 0; dim < OBS_DKMAX; dim++) vv[dim] = (dim < dk) ? (double)v[g * dk + dim] : 0.0;
      double aY = 0.0, aS = 0.0;
      for (int kp = 0; kp < k; kp++) {
        double yb[OBS_DKMAX], yt[OBS_DKMAX];
        ybar_kp(Cs, kp, dk, D, xr, yb);
        ytild_kp(Y + ((size_t)g * k + kp) * dk, vv, cs, sn, dk, yt);
        double e = 0.0;
#pragma unroll
        for (int dim = 0; dim < OBS_DKMAX; dim++) {
          const double r = yt[dim] - yb[dim];
          e = fma(r, r, e);
        }
        const double sv = (double)s[(size_t)g * k + kp], R = sv * sig[kp];
        aY += -0.5 * e / R - 0.5 * (double)dk * log(2.0 * KPMS_PI * R);
        aS += -nu_s * s_0 / sv * 0.5 - (1.0 + 0.5 * nu_s) * log(sv);
      }
      acc[0] = fma(m, aY, acc[0]);
      acc[4] = fma(m, aS, acc[4]);
      if (t >= L) {
        const int zt = z[(size_t)n * Tm + (t - L)];
        const double* A = Ab + (size_t)zt * D * P1;
        const float* win = x + ((size_t)n * T + (t - L)) * D;   // x[t-L .. t-1], oldest first
        double d[OBS_XMAX];
#pragma unroll
        for (int a = 0; a < OBS_XMAX; a++) {
          d[a] = 0.0;
          if (a < D) {
            double p = A[a * P1 + P1 - 1];
            for (int q = 0; q < P1 - 1; q++) p = fma(A[a * P1 + q], (double)win[q], p);
            d[a] = p - xr[a];
          }
        }
        const double* Qi = Qinv + (size_t)zt * D * D;
        double qf = 0.0;
#pragma unroll
        for (int a = 0; a < OBS_XMAX; a++)
          if (a < D) {
            double r = 0.0;
#pragma unroll
            for (int b = 0; b < OBS_XMAX; b++)
              if (b < D) r = fma(Qi[a * D + b], d[b], r);
            qf = fma(d[a], r, qf);
          }
        acc[1] = fma(m, -0.5 * qf - 0.5 * logdetQ[zt] - 0.5 * (double)D * KPMS_LOG_2PI, acc[1]);
      }
    }
    if (m1 != 0.0) {
      double d2 = 0.0;
      for (int dim = 0; dim < dk; dim++) {
        const double dd = (double)v[g * dk + dim] - (double)v[(g + 1) * dk + dim];
        d2 = fma(dd, dd, d2);
      }
      acc[3] = fma(m1, -0.5 * d2 / sigmasq_loc + lv, acc[3]);
      if (t >= L) {
        const int za = z[(size_t)n * Tm + (t - L)], zb = z[(size_t)n * Tm + (t - L) + 1];
        acc[2] = fma(m1, log(pi[(size_t)za * K + zb]), acc[2]);
      }
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < 5; i++) {
    const double w = warp_sum_d(acc[i]);
    if (lane == 0) red[warp][i] = w;
  }
  __syncthreads();
  if (threadIdx.x < 5) {
    double t = 0.0;
    for (int q = 0; q < LJ_THREADS / 32; q++) t += red[q][threadIdx.x];
    partial[(size_t)blockIdx.x * 5 + threadIdx.x] = t;
  }
}

__global__ void log_joint_reduce_kernel(const double* __restrict__ partial, int blocks, double* __restrict__ out) {
  const int i = threadIdx.x;
  if (i >= 5) return;
  double t = 0.0;
  for (int b = 0; b < blocks; b++) t += partial[(size_t)b * 5 + i];
  out[i] = t;
}

extern "C" int kpms_log_joint(const float* Y, const float* mask, const float* x, const float* s, const float* v,
                              const float* h, const int32_t* z, const double* pi, const double* Ab,
                              const double* Qinv, const double* logdetQ, const double* Cdfull,
                              const double* sigmasq, double sigmasq_loc, double nu_s, double s_0, int N, int T,
                              int k, int dk, int D, int L, int K, double* out5, void* workspace,
                              size_t workspace_bytes, void* stream) {
  KPMS_REQUIRE(Y && mask && x && s && v && h && z && pi && Ab && Qinv && logdetQ && Cdfull && sigmasq && out5 &&
                   workspace,
               KPMS_ERR_NULL);
  int rc = obs_check(N, T, k, dk, D);
  if (rc) return rc;
  KPMS_REQUIRE(L > 0 && T > L && K > 0, KPMS_ERR_SHAPE);
  KPMS_REQUIRE(kpms_aligned(workspace, 256) && workspace_bytes >= kpms_log_joint_workspace_bytes(),
               KPMS_ERR_WORKSPACE);
  cudaStream_t st = (cudaStream_t)stream;
  ObsDims dm{N, T, k, dk, D};
  const size_t smem = ((size_t)k * dk * (D + 1) + k) * sizeof(double);
  log_joint_kernel<<<LJ_BLOCKS, LJ_THREADS, smem, st>>>(Y, mask, x, s, v, h, z, pi, Ab, Qinv, logdetQ, Cdfull,
                                                       sigmasq, sigmasq_loc, nu_s, s_0, dm, L, K,
                                                       (double*)workspace);
  rc = kpms_last_launch();
  if (rc) return rc;
  log_joint_reduce_kernel<<<1, 32, 0, st>>>((const double*)workspace, LJ_BLOCKS, out5);
  return kpms_last_launch();
}

// ======================================================================== a14
__global__ void __launch_bounds__(256)
    scales_kernel(const float* __restrict__ Y, const float* __restrict__ v,
                  const float* __restrict__ h, const float* __restrict__ x,
                  const double* __restrict__ Cdfull, const double* __restrict__ sigmasq,
                  double nu_s, double s_0, const float* __restrict__ g, ObsDims dm,
                  float* __restrict__ s_out) {
  extern __shared__ __align__(16) double Cs[];
  const int k = dm.k, dk = dm.dk, D = dm.D;
  double* isig = Cs + (size_t)k * dk * (D + 1);
  load_cd(Cdfull, k * dk, D, Cs);
  for (int i = threadIdx.x; i < k; i += blockDim.x) isig[i] = 1.0 / sigmasq[i];
  __syncthreads();
  const long long frames = (long long)dm.N * dm.T;
  const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= frames) return;
  double xr[OBS_XMAX];
  load_x(x, (size_t)f, D, xr);
  double sn, cs;
  sincos((double)h[f], &sn, &cs);
  double vv[OBS_DKMAX];
#pragma unroll
  for (int dim = 0; dim < OBS_DKMAX; dim++) vv[dim] = (dim < dk) ? (double)v[f * dk + dim] : 0.0;
  const double base = s_0 * nu_s;
  for (int kp = 0; kp < k; kp++) {
    double yb[OBS_DKMAX], yt[OBS_DKMAX];
    ybar_kp(Cs, kp, dk, D, xr, yb);
    ytild_kp(Y + ((size_t)f * k + kp) * dk, vv, cs, sn, dk, yt);
    double e = 0.0;
#pragma unroll
    for (int dim = 0; dim < OBS_DKMAX; dim++) {
      const double r = yt[dim] - yb[dim];
      e = fma(r, r, e);
    }
    const double gg = (double)g[(size_t)f * k + kp];
    s_out[(size_t)f * k + kp] = (float)((e * isig[kp] + base) * 0.5 / gg);
  }
}

extern "C" int kpms_resample_scales(const float* Y, const float* v, const float* h, const float* x,
                                    const double* Cdfull, const double* sigmasq, double nu_s,
                                    double s_0, const float* g, int N, int T, int k, int dk, int D,
                                    float* s_out, void* stream) {
  KPMS_REQUIRE(Y && v && h && x && Cdfull && sigmasq && g && s_out, KPMS_ERR_NULL);
  int rc = obs_check(N, T, k, dk, D);
  if (rc) return rc;
  ObsDims dm{N, T, k, dk, D};
  const size_t smem = ((size_t)k * dk * (D + 1) + k) * sizeof(double);
  const long long frames = (long long)N * T;
  scales_kernel<<<(unsigned)((frames + 255) / 256), 256, smem, (cudaStream_t)stream>>>(
      Y, v, h, x, Cdfull, sigmasq, nu_s, s_0, g, dm, s_out);
  return kpms_last_launch();
}

// Fused variant: the Gamma((nu_s+3)/2, 1) variates are drawn inside the kernel when g == NULL
// (Philox4x32-10 keyed by the GLOBAL element index elem_offset + f*k + kp: the value
// kpms_fill_gamma_at would have written there), and, when S_y is requested, the kernel also
// accumulates  S_y(k) = sum_f mask ||Ytild - Ybar||^2 / s_new  and  sum_f mask  for the state it has
// just produced — the statistics resample_obs_variance needs at the start of the NEXT sweep
// (gibbs.py:86-94), so that sweep does not have to pass over Y once more.
__global__ void __launch_bounds__(OV_THREADS)
    scales_stats_kernel(const float* __restrict__ Y, const float* __restrict__ mask, const float* __restrict__ v,
                        const float* __restrict__ h, const float* __restrict__ x,
                        const double* __restrict__ Cdfull, const double* __restrict__ sigmasq, double nu_s,
                        double s_0, const float* __restrict__ g, uint2 key, uint32_t sid,
                        unsigned long long elem_offset, ObsDims dm, float* __restrict__ s_out,
                        double* __restrict__ partial) {
  extern __shared__ __align__(16) double Cs[];
  __shared__ double red[OV_THREADS / 32];
  const int k = dm.k, dk = dm.dk, D = dm.D;
  double* isig = Cs + (size_t)k * dk * (D + 1);
  double* accs = isig + k;   // [k+1][OV_THREADS], only when partial != NULL
  load_cd(Cdfull, k * dk, D, Cs);
  for (int i = threadIdx.x; i < k; i += blockDim.x) isig[i] = 1.0 / sigmasq[i];
  if (partial)
    for (int i = 0; i <= k; i++) accs[i * OV_THREADS + threadIdx.x] = 0.0;
  __syncthreads();
  const long long frames = (long long)dm.N * dm.T;
  const double base = s_0 * nu_s;
  const float shape = (float)((nu_s + 3.0) * 0.5);
  const bool vec = kp4_vec_ok(Y, nullptr, k, dk) && ((uintptr_t)s_out % 16 == 0) && (g == nullptr || (uintptr_t)g % 16 == 0);
  for (long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x; f < frames;
       f += (long long)gridDim.x * blockDim.x) {
    double xr[OBS_XMAX];
    load_x2(x, (size_t)f, D, xr);
    double sn, cs;
    sincos((double)h[f], &sn, &cs);
    double vv[OBS_DKMAX];
#pragma unroll
    for (int dim = 0; dim < OBS_DKMAX; dim++) vv[dim] = (dim < dk) ? (double)v[f * dk + dim] : 0.0;
    const double m = partial ? (double)mask[f] : 0.0;
    for (int kp0 = 0; kp0 < k; kp0 += 4) {
      Kp4 q;
      load_kp4(Y + (size_t)f * k * dk, g ? g + (size_t)f * k : nullptr, kp0, k, dk, vec, q);   // q.s = injected gammas
      float sv4[4];
      if (!g) {   // in-kernel variates: one Philox block per pair of elements (pairs never straddle a frame when k is even)
        const unsigned long long el0 = elem_offset + (unsigned long long)f * k + kp0;
        if ((el0 & 1ull) == 0 && kp0 + 3 < k) {
          philox_gamma_pair(el0, shape, key, sid, q.s[0], q.s[1]);
          philox_gamma_pair(el0 + 2, shape, key, sid, q.s[2], q.s[3]);
        } else {
#pragma unroll
          for (int j = 0; j < 4; j++)
            if (kp0 + j < k) q.s[j] = philox_gamma(el0 + j, shape, key, sid);
        }
      }
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int kp = kp0 + j;
        sv4[j] = 0.0f;
        if (kp < k) {
          double yb[OBS_DKMAX], yt[OBS_DKMAX];
          ybar_kp(Cs, kp, dk, D, xr, yb);
          ytild3(KP4_Y(q, j, dk, 0), KP4_Y(q, j, dk, 1), KP4_Y(q, j, dk, 2), vv, cs, sn, dk, yt);
          double e = 0.0;
#pragma unroll
          for (int dim = 0; dim < OBS_DKMAX; dim++) {
            const double r = yt[dim] - yb[dim];
            e = fma(r, r, e);
          }
          const size_t el = (size_t)f * k + kp;
          // gamma variates and the new scale are positive and normal: reciprocals by seed + one refinement
          sv4[j] = (float)((e * isig[kp] + base) * 0.5 * rcp_pos_d((double)q.s[j]));
          if (!vec) s_out[el] = sv4[j];
          if (partial && m > 0.0) accs[kp * OV_THREADS + threadIdx.x] += m * e * rcp_pos_d((double)sv4[j]);
        }
      }
      if (vec) *reinterpret_cast<float4*>(s_out + (size_t)f * k + kp0) = make_float4(sv4[0], sv4[1], sv4[2], sv4[3]);
    }
    if (partial && m > 0.0) accs[k * OV_THREADS + threadIdx.x] += m;
  }
  if (!partial) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = 0; i <= k; i++) {
    double w = warp_sum_d(accs[i * OV_THREADS + threadIdx.x]);
    __syncthreads();
    if (lane == 0) red[warp] = w;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int q = 0; q < OV_THREADS / 32; q++) t += red[q];
      partial[(size_t)blockIdx.x * (OBS_KPMAX + 1) + (i < k ? i : OBS_KPMAX)] = t;
    }
  }
}

extern "C" int kpms_resample_scales_stats(const float* Y, const float* mask, const float* v, const float* h,
                                          const float* x, const double* Cdfull, const double* sigmasq,
                                          double nu_s, double s_0, const float* g, uint64_t seed,
                                          uint64_t stream_id, uint64_t elem_offset, int N, int T, int k, int dk,
                                          int D, float* s_out, double* S_y, double* mask_sum, void* workspace,
                                          size_t workspace_bytes, void* stream) {
  KPMS_REQUIRE(Y && v && h && x && Cdfull && sigmasq && s_out, KPMS_ERR_NULL);
  int rc = obs_check(N, T, k, dk, D);
  if (rc) return rc;
  KPMS_REQUIRE(g != nullptr || nu_s + 3.0 >= 2.0, KPMS_ERR_UNSUPPORTED);   // in-kernel sampler needs shape >= 1
  const bool stats = S_y != nullptr;
  if (stats) {
    KPMS_REQUIRE(mask && mask_sum && workspace, KPMS_ERR_NULL);
    KPMS_REQUIRE(kpms_aligned(workspace, 256) && workspace_bytes >= kpms_obs_variance_stats_workspace_bytes(N, T, k),
                 KPMS_ERR_WORKSPACE);
  }
  cudaStream_t st = (cudaStream_t)stream;
  ObsDims dm{N, T, k, dk, D};
  const size_t smem = ((size_t)k * dk * (D + 1) + k + (stats ? (size_t)(k + 1) * OV_THREADS : 0)) * sizeof(double);
  KPMS_CUDA(cudaFuncSetAttribute(scales_stats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long frames = (long long)N * T;
  (void)frames;
  const long long blocks = OV_BLOCKS;   // the grid of obs_variance_kernel: the statistics come out bit-identical to that pass
  scales_stats_kernel<<<(unsigned)blocks, OV_THREADS, smem, st>>>(Y, mask, v, h, x, Cdfull, sigmasq, nu_s, s_0, g,
                                                                  split_seed(seed), (uint32_t)stream_id,
                                                                  (unsigned long long)elem_offset, dm, s_out,
                                                                  stats ? (double*)workspace : nullptr);
  rc = kpms_last_launch();
  if (rc || !stats) return rc;
  obs_variance_reduce_kernel<<<1, OBS_KPMAX + 1, 0, st>>>((const double*)workspace, (int)blocks, k, S_y, mask_sum);
  return kpms_last_launch();
}

// ======================================================================== a15
// Best & Fisher (1979) von Mises rejection sampler on injected uniforms; contract shared
// with oracle/kpms_oracle.py:sample_vonmises.
__device__ __forceinline__ double vonmises_dev(double kappa, const float* __restrict__ u,
                                               int rounds) {
  if (kappa < 1e-8) return KPMS_PI * (2.0 * (double)u[2] - 1.0);
  const double tau = 1.0 + sqrt(1.0 + 4.0 * kappa * kappa);
  const double rho = (tau - sqrt(2.0 * tau)) / (2.0 * kappa);
  const double r = (1.0 + rho * rho) / (2.0 * rho);
  for (int rd = 0; rd < rounds; rd++) {
    const double u1 = (double)u[rd * 3 + 0], u2 = (double)u[rd * 3 + 1], u3 = (double)u[rd * 3 + 2];
    const double zc = cospi(u1);
    double f = (1.0 + r * zc) / (r + zc);
    const double c = kappa * (r - f);
    if (c * (2.0 - c) - u2 > 0.0 || log(c / u2) + 1.0 - c >= 0.0) {
      f = fmin(1.0, fmax(-1.0, f));
      return (u3 > 0.5) ? acos(f) : -acos(f);
    }
  }
  return 0.0;
}

// The same sampler with its uniforms drawn in place: round rd of GLOBAL frame fg uses the Philox
// block (counter = (fg, rd), key = seed, stream id), so no uniform buffer exists and the proposal
// loop simply runs until it accepts (Best-Fisher accepts with probability >= 0.66 per round; the
// cap of 64 rounds leaves < 1e-30 and is counted in `exhausted` should it ever be hit).
#define VM_MAX_ROUNDS 64
__device__ __forceinline__ double vonmises_philox(double kappa, unsigned long long fg, uint2 key, uint32_t sid,
                                                  int* __restrict__ exhausted) {
  if (kappa < 1e-8) {
    const uint4 r = philox4x32_10(make_uint4((uint32_t)fg, (uint32_t)(fg >> 32), 0u, sid), key);
    return KPMS_PI * (2.0 * (double)u01(r.z) - 1.0);
  }
  const double tau = 1.0 + sqrt(1.0 + 4.0 * kappa * kappa);
  const double rho = (tau - sqrt(2.0 * tau)) / (2.0 * kappa);
  const double rr = (1.0 + rho * rho) / (2.0 * rho);
  for (uint32_t rd = 0; rd < VM_MAX_ROUNDS; rd++) {
    const uint4 r = philox4x32_10(make_uint4((uint32_t)fg, (uint32_t)(fg >> 32), rd, sid), key);
    const double u1 = (double)u01(r.x), u2 = (double)u01_open0(r.y), u3 = (double)u01(r.z);
    const double zc = cospi(u1);
    double f = (1.0 + rr * zc) / (rr + zc);
    const double c = kappa * (rr - f);
    if (c * (2.0 - c) - u2 > 0.0 || log(c / u2) + 1.0 - c >= 0.0) {
      f = fmin(1.0, fmax(-1.0, f));
      return (u3 > 0.5) ? acos(f) : -acos(f);
    }
  }
  if (exhausted) atomicAdd(exhausted, 1);
  return 0.0;
}

template <bool PHILOX>
__global__ void __launch_bounds__(256)
    heading_kernel(const float* __restrict__ Y, const float* __restrict__ v,
                   const float* __restrict__ x, const float* __restrict__ s,
                   const double* __restrict__ Cdfull, const double* __restrict__ sigmasq,
                   const float* __restrict__ u, int rounds, uint2 key, uint32_t sid,
                   unsigned long long frame_offset, int* __restrict__ exhausted, ObsDims dm,
                   float* __restrict__ h_out, double* __restrict__ theta_out, double* __restrict__ kappa_out) {
  extern __shared__ __align__(16) double Cs[];
  const int k = dm.k, dk = dm.dk, D = dm.D;
  double* isig = Cs + (size_t)k * dk * (D + 1);
  load_cd(Cdfull, k * dk, D, Cs);
  for (int i = threadIdx.x; i < k; i += blockDim.x) isig[i] = sigmasq[i];
  __syncthreads();
  const long long frames = (long long)dm.N * dm.T;
  const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= frames) return;
  double xr[OBS_XMAX];
  load_x2(x, (size_t)f, D, xr);
  const double v0 = (double)v[f * dk + 0], v1 = (double)v[f * dk + 1];
  double S00 = 0.0, S01 = 0.0, S10 = 0.0, S11 = 0.0;
  const bool vec = kp4_vec_ok(Y, s, k, dk);
  for (int kp0 = 0; kp0 < k; kp0 += 4) {
    Kp4 q;
    load_kp4(Y + (size_t)f * k * dk, s + (size_t)f * k, kp0, k, dk, vec, q);
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int kp = kp0 + j;
      if (kp < k) {
        double yb[OBS_DKMAX];
        ybar_kp(Cs, kp, dk, D, xr, yb);
        const double w = 1.0 / ((double)q.s[j] * isig[kp]);
        const double y0 = ((double)KP4_Y(q, j, dk, 0) - v0) * w, y1 = ((double)KP4_Y(q, j, dk, 1) - v1) * w;
        S00 = fma(yb[0], y0, S00);
        S01 = fma(yb[0], y1, S01);
        S10 = fma(yb[1], y0, S10);
        S11 = fma(yb[1], y1, S11);
      }
    }
  }
  const double kc = S00 + S11, ks = S01 - S10;
  // util.py:80-82 vector_to_angle
  const double xx = kc + 1e-10;
  double ang = atan(ks / xx) + ((xx > 0.0) ? KPMS_PI : 0.0);
  ang = ang - 2.0 * KPMS_PI * floor(ang / (2.0 * KPMS_PI)) - KPMS_PI;
  const double kappa = sqrt(kc * kc + ks * ks);
  if (theta_out) theta_out[f] = ang;
  if (kappa_out) kappa_out[f] = kappa;
  double hh = ang + (PHILOX ? vonmises_philox(kappa, frame_offset + (unsigned long long)f, key, sid, exhausted)
                            : vonmises_dev(kappa, u + (size_t)f * rounds * 3, rounds));
  hh = hh - 2.0 * KPMS_PI * rint(hh / (2.0 * KPMS_PI));
  h_out[f] = (float)hh;
}

extern "C" int kpms_resample_heading(const float* Y, const float* v, const float* x, const float* s,
                                     const double* Cdfull, const double* sigmasq, const float* u,
                                     int rounds, int N, int T, int k, int dk, int D, float* h_out,
                                     double* theta_out, double* kappa_out, void* stream) {
  KPMS_REQUIRE(Y && v && x && s && Cdfull && sigmasq && u && h_out, KPMS_ERR_NULL);
  int rc = obs_check(N, T, k, dk, D);
  if (rc) return rc;
  KPMS_REQUIRE(rounds > 0, KPMS_ERR_SHAPE);
  ObsDims dm{N, T, k, dk, D};
  const size_t smem = ((size_t)k * dk * (D + 1) + k) * sizeof(double);
  const long long frames = (long long)N * T;
  heading_kernel<false><<<(unsigned)((frames + 255) / 256), 256, smem, (cudaStream_t)stream>>>(
      Y, v, x, s, Cdfull, sigmasq, u, rounds, make_uint2(0u, 0u), 0u, 0ULL, nullptr, dm, h_out, theta_out, kappa_out);
  return kpms_last_launch();
}

extern "C" int kpms_resample_heading_philox(const float* Y, const float* v, const float* x, const float* s,
                                            const double* Cdfull, const double* sigmasq, uint64_t seed,
                                            uint64_t stream_id, uint64_t frame_offset, int N, int T, int k, int dk,
                                            int D, float* h_out, double* theta_out, double* kappa_out,
                                            int32_t* exhausted, void* stream) {
  KPMS_REQUIRE(Y && v && x && s && Cdfull && sigmasq && h_out, KPMS_ERR_NULL);
  int rc = obs_check(N, T, k, dk, D);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (exhausted) KPMS_CUDA(cudaMemsetAsync(exhausted, 0, sizeof(int32_t), st));
  ObsDims dm{N, T, k, dk, D};
  const size_t smem = ((size_t)k * dk * (D + 1) + k) * sizeof(double);
  const long long frames = (long long)N * T;
  heading_kernel<true><<<(unsigned)((frames + 255) / 256), 256, smem, st>>>(
      Y, v, x, s, Cdfull, sigmasq, nullptr, 0, split_seed(seed), (uint32_t)stream_id,
      (unsigned long long)frame_offset, exhausted, dm, h_out, theta_out, kappa_out);
  return kpms_last_launch();
}

// ======================================================================== a16
// Stage 1 (frame-parallel, gibbs.py:40-46): gammasq = 1/sum_k 1/(s sigmasq),
//   mu = sum_k (Y_k - R(h) Ybar_k) gammasq/(s sigmasq); stored in information form
//   rg = 1/gammasq, em = mu/gammasq, ct = sum_dim -1/2 log(2 pi gammasq) - 1/2 mu^2/gammasq.
// Stage 2 (one thread per session): the isotropic model keeps every covariance a multiple of
//   the identity, so kalman_sample reduces to scalar recursions (information form, same
//   algebra as kalman.cu): lam_c = lam_p + rg, eta_c = eta_p + em,
//   S_cond = 1/(lam_c + 1/q) (q = sigmasq_loc), lam_p' = lam_c/(q lam_c + 1), eta_p' likewise;
//   backward x = S_cond (eta_c + x'/q) + sqrt(S_cond) w.
#define LOC_UNROLL 8
#define LOC_CHUNK 256   // steps per independent chunk
#define LOC_BURN 128    // warm-up steps: the scalar filter forgets by >= 1/(q*rg+1) per step

__host__ __device__ static inline int loc_nchunks(int T) {
  const int Tt = T - 1;
  return Tt <= LOC_CHUNK ? 1 : (Tt + LOC_CHUNK - 1) / LOC_CHUNK;
}

extern "C" size_t kpms_resample_location_workspace_bytes(int N, int T, int dk) {
  if (N <= 0 || T <= 0 || dk <= 0) return 0;
  // stats: (2+dk) doubles per frame; filter: (1+dk) doubles per frame; ll partials; tv
  const size_t units = (size_t)N * ((T - 1 + 127) / 128 + 1);   // chunks of the scan variant (LSC_CHUNK = 128) >= those of the warm-up variant
  return kpms_round_up((size_t)N * T * (2 + dk) * sizeof(double), 256) +
         kpms_round_up((size_t)N * T * (1 + dk) * sizeof(double), 256) +
         kpms_round_up(units * sizeof(double), 256) +
         kpms_round_up((size_t)N * sizeof(int), 256) +
         kpms_round_up(units * (6 + 2 * dk) * sizeof(double), 256);
}

__global__ void __launch_bounds__(256)
    location_stats_kernel(const float* __restrict__ Y, const float* __restrict__ h,
                          const float* __restrict__ x, const float* __restrict__ s,
                          const double* __restrict__ Cdfull, const double* __restrict__ sigmasq,
                          ObsDims dm, double* __restrict__ st) {
  extern __shared__ __align__(16) double Cs[];
  const int k = dm.k, dk = dm.dk, D = dm.D;
  double* sig = Cs + (size_t)k * dk * (D + 1);
  load_cd(Cdfull, k * dk, D, Cs);
  for (int i = threadIdx.x; i < k; i += blockDim.x) sig[i] = sigmasq[i];
  __syncthreads();
  const long long frames = (long long)dm.N * dm.T;
  const long long f = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= frames) return;
  double xr[OBS_XMAX];
  load_x2(x, (size_t)f, D, xr);
  double sn, cs;
  sincos((double)h[f], &sn, &cs);
  double wsum = 0.0, acc[OBS_DKMAX] = {0.0, 0.0, 0.0};
  const bool vec = kp4_vec_ok(Y, s, k, dk);
  for (int kp0 = 0; kp0 < k; kp0 += 4) {
    Kp4 q;
    load_kp4(Y + (size_t)f * k * dk, s + (size_t)f * k, kp0, k, dk, vec, q);
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int kp = kp0 + j;
      if (kp < k) {
        double yb[OBS_DKMAX];
        ybar_kp(Cs, kp, dk, D, xr, yb);
        const double w = 1.0 / ((double)q.s[j] * sig[kp]);
        // R(h) Ybar  (util.py:85-109): [c -s; s c] on the first two coordinates
        const double r0 = cs * yb[0] - sn * yb[1], r1 = sn * yb[0] + cs * yb[1];
        acc[0] = fma((double)KP4_Y(q, j, dk, 0) - r0, w, acc[0]);
        acc[1] = fma((double)KP4_Y(q, j, dk, 1) - r1, w, acc[1]);
        if (dk > 2) acc[2] = fma((double)KP4_Y(q, j, dk, 2) - yb[2], w, acc[2]);
        wsum += w;
      }
    }
  }
  // mu = acc * gammasq with gammasq = 1/wsum  ->  em = mu/gammasq, rg = wsum
  const double g2 = 1.0 / wsum;
  double ct = 0.0;
  double* o = st + (size_t)f * (2 + dk);
  o[0] = wsum;
  for (int dim = 0; dim < dk; dim++) {
    const double mu = acc[dim] * g2;
    o[1 + dim] = mu * wsum;
    ct += -0.5 * (KPMS_LOG_2PI + log(g2)) - 0.5 * mu * mu * wsum;
  }
  o[1 + dk] = ct;
}

// unmasked steps per session (mask[:, :-1]); trailing padding assumed for the warm-up start
__global__ void location_tv_kernel(const float* __restrict__ mask, int T, int* __restrict__ tv) {
  const int n = blockIdx.x;
  __shared__ int cnt;
  if (threadIdx.x == 0) cnt = 0;
  __syncthreads();
  int c = 0;
  for (int t = threadIdx.x; t < T - 1; t += blockDim.x) c += mask[(size_t)n * T + t] > 0.0f;
  atomicAdd(&cnt, c);
  __syncthreads();
  if (threadIdx.x == 0) tv[n] = cnt;
}

// Forward filter: one thread per (session, chunk).  A chunk other than the first is warmed up
// from the prior over LOC_BURN earlier steps; with a contraction of at least 1/(q*rg+1) per
// step (~0.1 here) the carry it reaches is the sequential one to the last bit.
template <int DK>
__global__ void __launch_bounds__(64)
    location_forward_kernel(const double* __restrict__ st, const float* __restrict__ mask,
                            const float* __restrict__ w_last, const int* __restrict__ tv, double q,
                            int N, int T, int want_ll, double* __restrict__ fw,
                            double* __restrict__ llpart, float* __restrict__ v_out) {
  const int nch = loc_nchunks(T);
  const int unit = blockIdx.x * blockDim.x + threadIdx.x;
  if (unit >= N * nch) return;
  const int n = unit / nch, cc = unit % nch;
  const int Tt = T - 1;
  const int t0 = (nch == 1) ? 0 : cc * LOC_CHUNK;
  const int t1 = (nch == 1) ? Tt : min(Tt, t0 + LOC_CHUNK);
  const int tb = (cc > 0) ? max(0, min(t0, tv[n]) - LOC_BURN) : 0;
  const double* st_n = st + (size_t)n * T * (2 + DK);
  double* fw_n = fw + (size_t)n * T * (1 + DK);
  const float* mk = mask + (size_t)n * T;
  const double lq = log(q);
  double lam = 1e-6, eta[DK];   // gibbs.py:48: m0 = 0, S0 = 1e6 I
#pragma unroll
  for (int d = 0; d < DK; d++) eta[d] = 0.0;
  double ll = (cc == 0) ? 0.5 * DK * log(1e-6) : 0.0;
  for (int tbk = tb; tbk < t1; tbk += LOC_UNROLL) {
    double rg[LOC_UNROLL], em[LOC_UNROLL][DK], ct[LOC_UNROLL];
    float m[LOC_UNROLL];
#pragma unroll
    for (int i = 0; i < LOC_UNROLL; i++) {
      const int t = tbk + i;
      m[i] = 0.0f;
      if (t < t1) {
        m[i] = mk[t];
        const double* o = st_n + (size_t)t * (2 + DK);
        rg[i] = o[0];
#pragma unroll
        for (int d = 0; d < DK; d++) em[i][d] = o[1 + d];
        ct[i] = o[1 + DK];
      }
    }
#pragma unroll
    for (int i = 0; i < LOC_UNROLL; i++) {
      const int t = tbk + i;
      if (t < t1 && m[i] > 0.0f) {
        const double lc = lam + rg[i];
        const double r = 1.0 / fma(q, lc, 1.0);
        const double Sc = q * r;
        const double Lc = sqrt(Sc);
        const bool store = t >= t0;
        double* o = fw_n + (size_t)t * (1 + DK);
        if (store) o[0] = Lc;
        double uu = 0.0;
#pragma unroll
        for (int d = 0; d < DK; d++) {
          const double ec = eta[d] + em[i][d];
          if (store) o[1 + d] = ec;
          eta[d] = ec * r;
          uu = fma(Lc * ec, Lc * ec, uu);
        }
        lam = lc * r;
        if (want_ll && store) ll += ct[i] - 0.5 * DK * lq + DK * log(Lc) + 0.5 * uu;
      }
    }
  }
  if (cc == nch - 1) {
    // unconditional final update (kalman.py:117) and draw of the last state (kalman.py:150)
    const double* o = st_n + (size_t)(T - 1) * (2 + DK);
    const double lc = lam + o[0];
    const double Ll = 1.0 / sqrt(lc);
    double tt = 0.0;
    double* xl = fw_n + (size_t)(T - 1) * (1 + DK);
#pragma unroll
    for (int d = 0; d < DK; d++) {
      const double ec = eta[d] + o[1 + d];
      const double td = Ll * ec;
      tt = fma(td, td, tt);
      const double xv = Ll * (td + (double)w_last[(size_t)n * DK + d]);
      xl[1 + d] = xv;
      v_out[((size_t)n * T + (T - 1)) * DK + d] = (float)xv;
    }
    ll += o[1 + DK] + DK * log(Ll) + 0.5 * tt;
  }
  llpart[(size_t)n * nch + cc] = ll;
}

// Backward sampler, one thread per (session, chunk): x_t = S_cond (eta_c + x_{t+1}/q) + sqrt(S_cond) w_t.
// A chunk other than the last starts LOC_BURN steps later from 0 (or from x_last inside the
// trailing padding) and couples to the sequential path at the same geometric rate.
template <int DK>
__global__ void __launch_bounds__(64)
    location_backward_kernel(const double* __restrict__ fw, const float* __restrict__ mask,
                             const float* __restrict__ w, double q, int N, int T,
                             const double* __restrict__ llpart, float* __restrict__ v_out,
                             double* __restrict__ loglik) {
  const int nch = loc_nchunks(T);
  const int unit = blockIdx.x * blockDim.x + threadIdx.x;
  if (unit >= N * nch) return;
  const int n = unit / nch, cc = unit % nch;
  const int Tt = T - 1;
  const int t0 = (nch == 1) ? 0 : cc * LOC_CHUNK;
  const int t1 = (nch == 1) ? Tt : min(Tt, t0 + LOC_CHUNK);
  const double* fw_n = fw + (size_t)n * T * (1 + DK);
  const float* mk = mask + (size_t)n * T;
  const float* w_n = w + (size_t)n * Tt * DK;
  const double iq = 1.0 / q;
  if (cc == 0 && loglik != nullptr) {
    double s = 0.0;
    for (int c = 0; c < nch; c++) s += llpart[(size_t)n * nch + c];
    loglik[n] = s;
  }
  int ts = t1;
  bool exact = true;
  if (cc < nch - 1) {
    ts = min(Tt, t1 + LOC_BURN);
    exact = (ts == Tt) || !(mk[ts] > 0.0f);
  }
  double xv[DK];
#pragma unroll
  for (int d = 0; d < DK; d++) xv[d] = exact ? fw_n[(size_t)(T - 1) * (1 + DK) + 1 + d] : 0.0;
  for (int tbk = ts - 1; tbk >= t0; tbk -= LOC_UNROLL) {
    double Lc[LOC_UNROLL], ec[LOC_UNROLL][DK];
    float wv[LOC_UNROLL][DK], m[LOC_UNROLL];
#pragma unroll
    for (int i = 0; i < LOC_UNROLL; i++) {
      const int t = tbk - i;
      m[i] = 0.0f;
      if (t >= t0) {
        m[i] = mk[t];
        if (m[i] > 0.0f) {
          const double* o = fw_n + (size_t)t * (1 + DK);
          Lc[i] = o[0];
#pragma unroll
          for (int d = 0; d < DK; d++) {
            ec[i][d] = o[1 + d];
            wv[i][d] = w_n[(size_t)t * DK + d];
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < LOC_UNROLL; i++) {
      const int t = tbk - i;
      if (t >= t0) {
        if (m[i] > 0.0f) {
#pragma unroll
          for (int d = 0; d < DK; d++) {
            const double hh = fma(xv[d], iq, ec[i][d]);
            xv[d] = Lc[i] * fma(Lc[i], hh, (double)wv[i][d]);
            if (t < t1) v_out[((size_t)n * T + t) * DK + d] = (float)xv[d];
          }
        } else if (t < t1) {
#pragma unroll
          for (int d = 0; d < DK; d++) v_out[((size_t)n * T + t) * DK + d] = 0.0f;  // kalman.py:142-143
        }
      }
    }
  }
}

// ------------------------------------------------------------------ exact time-parallel variant
// In this scalar model every pass over time is a composition of maps that compose in closed form:
//   filter precision   lam' = (lam + rg) / (q lam + q rg + 1)          a Moebius map (2 x 2 matrix)
//   filter mean        eta' = r_t (eta + em_t),  r_t = 1/(q (lam_t + rg_t) + 1)   affine given the lam path
//   backward sample    x_t  = a_t x_{t+1} + b_t,  a_t = S_t / q,  b_t = S_t ec_t + sqrt(S_t) w_t   affine
// so the session is cut into chunks, every chunk composes its map, a chunk's exact starting value
// is the composition of the maps of the chunks before (after) it applied to the prior (to x_last), and
// the chunk is then run from that value: an associative scan over time, two levels deep, with no
// warm-up, no tolerance and no fallback (kalman.py:86-95,129-140 specialised as in gibbs.py:38-54).
#define LSC_CHUNK 128
__host__ __device__ static inline int lsc_nchunks(int T) { return (T - 1 + LSC_CHUNK - 1) / LSC_CHUNK; }

struct LocScan {
  double* mob;   // [units][4]  Moebius map of the chunk's filter-precision recursion
  double* fA;    // [units]     affine map of the chunk's filter-mean recursion: eta_out = fA eta_in + fB
  double* fB;    // [units][DK]
  double* bA;    // [units]     affine map of the chunk's backward recursion: x_out = bA x_in + bB
  double* bB;    // [units][DK]
};

__device__ __forceinline__ double lsc_lam_start(const double* __restrict__ mob, int n, int nch, int cc) {
  double lam = 1e-6;   // gibbs.py:48: S0 = 1e6 I
  const double* m = mob + (size_t)n * nch * 4;
  for (int c = 0; c < cc; c++) lam = fma(m[4 * c], lam, m[4 * c + 1]) / fma(m[4 * c + 2], lam, m[4 * c + 3]);
  return lam;
}

// pass 1: Moebius composite of every chunk
__global__ void __launch_bounds__(64) lsc_mobius_kernel(const double* __restrict__ st, const float* __restrict__ mask,
                                                         double q, int N, int T, int DK, LocScan ls) {
  const int nch = lsc_nchunks(T);
  const int unit = blockIdx.x * blockDim.x + threadIdx.x;
  if (unit >= N * nch) return;
  const int n = unit / nch, cc = unit % nch;
  const int t0 = cc * LSC_CHUNK, t1 = min(T - 1, t0 + LSC_CHUNK);
  const double* st_n = st + (size_t)n * T * (2 + DK);
  const float* mk = mask + (size_t)n * T;
  double a = 1.0, b = 0.0, c = 0.0, d = 1.0;
  int since = 0;
  for (int t = t0; t < t1; t++) {
    if (!(mk[t] > 0.0f)) continue;
    const double rg = st_n[(size_t)t * (2 + DK)];
    // [1 rg; q q rg + 1] . [a b; c d]
    const double na = fma(rg, c, a), nb = fma(rg, d, b);
    const double nc = fma(q, na, c), nd = fma(q, nb, d);
    a = na; b = nb; c = nc; d = nd;
    if (++since == 8) {   // the map is projective: rescale before the entries grow out of range
      const double sc = 1.0 / d;
      a *= sc; b *= sc; c *= sc; d = 1.0;
      since = 0;
    }
  }
  double* m = ls.mob + (size_t)unit * 4;
  m[0] = a; m[1] = b; m[2] = c; m[3] = d;
}

// pass 2: with the exact starting precision, the affine map of the chunk's mean recursion
template <int DK>
__global__ void __launch_bounds__(64) lsc_mean_map_kernel(const double* __restrict__ st, const float* __restrict__ mask,
                                                           double q, int N, int T, LocScan ls) {
  const int nch = lsc_nchunks(T);
  const int unit = blockIdx.x * blockDim.x + threadIdx.x;
  if (unit >= N * nch) return;
  const int n = unit / nch, cc = unit % nch;
  const int t0 = cc * LSC_CHUNK, t1 = min(T - 1, t0 + LSC_CHUNK);
  const double* st_n = st + (size_t)n * T * (2 + DK);
  const float* mk = mask + (size_t)n * T;
  double lam = lsc_lam_start(ls.mob, n, nch, cc);
  double A = 1.0, Bv[DK];
#pragma unroll
  for (int dd = 0; dd < DK; dd++) Bv[dd] = 0.0;
  for (int t = t0; t < t1; t++) {
    if (!(mk[t] > 0.0f)) continue;
    const double* o = st_n + (size_t)t * (2 + DK);
    const double lc = lam + o[0];
    const double r = 1.0 / fma(q, lc, 1.0);
    A *= r;
#pragma unroll
    for (int dd = 0; dd < DK; dd++) Bv[dd] = (Bv[dd] + o[1 + dd]) * r;
    lam = lc * r;
  }
  ls.fA[unit] = A;
#pragma unroll
  for (int dd = 0; dd < DK; dd++) ls.fB[(size_t)unit * DK + dd] = Bv[dd];
}

// pass 3: the filter itself from the exact chunk start (records for the sampler, log-likelihood, last state),
// and the affine map of the chunk's backward recursion
template <int DK>
__global__ void __launch_bounds__(64) lsc_filter_kernel(const double* __restrict__ st, const float* __restrict__ mask,
                                                         const float* __restrict__ w, const float* __restrict__ w_last,
                                                         double q, int N, int T, int want_ll, double* __restrict__ fw,
                                                         double* __restrict__ llpart, float* __restrict__ v_out,
                                                         LocScan ls) {
  const int nch = lsc_nchunks(T);
  const int unit = blockIdx.x * blockDim.x + threadIdx.x;
  if (unit >= N * nch) return;
  const int n = unit / nch, cc = unit % nch;
  const int Tt = T - 1;
  const int t0 = cc * LSC_CHUNK, t1 = min(Tt, t0 + LSC_CHUNK);
  const double* st_n = st + (size_t)n * T * (2 + DK);
  double* fw_n = fw + (size_t)n * T * (1 + DK);
  const float* mk = mask + (size_t)n * T;
  const float* w_n = w + (size_t)n * Tt * DK;
  const double lq = log(q), iq = 1.0 / q;
  double lam = lsc_lam_start(ls.mob, n, nch, cc);
  double eta[DK];
#pragma unroll
  for (int dd = 0; dd < DK; dd++) eta[dd] = 0.0;   // gibbs.py:48: m0 = 0
  for (int c = 0; c < cc; c++) {
    const double A = ls.fA[(size_t)n * nch + c];
#pragma unroll
    for (int dd = 0; dd < DK; dd++) eta[dd] = fma(A, eta[dd], ls.fB[((size_t)n * nch + c) * DK + dd]);
  }
  double ll = (cc == 0) ? 0.5 * DK * log(1e-6) : 0.0;
  // records first (forward in time), then the backward map is composed from them (backward in time)
  for (int t = t0; t < t1; t++) {
    if (!(mk[t] > 0.0f)) continue;
    const double* o = st_n + (size_t)t * (2 + DK);
    const double lc = lam + o[0];
    const double r = 1.0 / fma(q, lc, 1.0);
    const double Lc = sqrt(q * r);
    double* fo = fw_n + (size_t)t * (1 + DK);
    fo[0] = Lc;
    double uu = 0.0;
#pragma unroll
    for (int dd = 0; dd < DK; dd++) {
      const double ec = eta[dd] + o[1 + dd];
      fo[1 + dd] = ec;
      eta[dd] = ec * r;
      uu = fma(Lc * ec, Lc * ec, uu);
    }
    lam = lc * r;
    if (want_ll) ll += o[1 + DK] - 0.5 * DK * lq + DK * log(Lc) + 0.5 * uu;
  }
  if (cc == nch - 1) {
    // unconditional final update (kalman.py:117) and draw of the last state (kalman.py:150)
    const double* o = st_n + (size_t)(T - 1) * (2 + DK);
    const double lc = lam + o[0];
    const double Ll = 1.0 / sqrt(lc);
    double tt = 0.0;
    double* xl = fw_n + (size_t)(T - 1) * (1 + DK);
#pragma unroll
    for (int dd = 0; dd < DK; dd++) {
      const double ec = eta[dd] + o[1 + dd];
      const double td = Ll * ec;
      tt = fma(td, td, tt);
      const double xv = Ll * (td + (double)w_last[(size_t)n * DK + dd]);
      xl[1 + dd] = xv;
      v_out[((size_t)n * T + (T - 1)) * DK + dd] = (float)xv;
    }
    ll += o[1 + DK] + DK * log(Ll) + 0.5 * tt;
  }
  llpart[(size_t)n * nch + cc] = ll;
  double A = 1.0, Bv[DK];
#pragma unroll
  for (int dd = 0; dd < DK; dd++) Bv[dd] = 0.0;
  for (int t = t1 - 1; t >= t0; t--) {
    if (!(mk[t] > 0.0f)) continue;   // kalman.py:142-143: the carried sample is unchanged
    const double* fo = fw_n + (size_t)t * (1 + DK);
    const double Lc = fo[0], at = Lc * Lc * iq;
    A *= at;
#pragma unroll
    for (int dd = 0; dd < DK; dd++)
      Bv[dd] = fma(at, Bv[dd], Lc * fma(Lc, fo[1 + dd], (double)w_n[(size_t)t * DK + dd]));
  }
  ls.bA[unit] = A;
#pragma unroll
  for (int dd = 0; dd < DK; dd++) ls.bB[(size_t)unit * DK + dd] = Bv[dd];
}

// pass 4: backward sampler from the exact chunk start
template <int DK>
__global__ void __launch_bounds__(64) lsc_sample_kernel(const double* __restrict__ fw, const float* __restrict__ mask,
                                                         const float* __restrict__ w, double q, int N, int T,
                                                         const double* __restrict__ llpart, float* __restrict__ v_out,
                                                         double* __restrict__ loglik, LocScan ls) {
  const int nch = lsc_nchunks(T);
  const int unit = blockIdx.x * blockDim.x + threadIdx.x;
  if (unit >= N * nch) return;
  const int n = unit / nch, cc = unit % nch;
  const int Tt = T - 1;
  const int t0 = cc * LSC_CHUNK, t1 = min(Tt, t0 + LSC_CHUNK);
  const double* fw_n = fw + (size_t)n * T * (1 + DK);
  const float* mk = mask + (size_t)n * T;
  const float* w_n = w + (size_t)n * Tt * DK;
  const double iq = 1.0 / q;
  if (cc == 0 && loglik != nullptr) {
    double s = 0.0;
    for (int c = 0; c < nch; c++) s += llpart[(size_t)n * nch + c];
    loglik[n] = s;
  }
  double xv[DK];
#pragma unroll
  for (int dd = 0; dd < DK; dd++) xv[dd] = fw_n[(size_t)(T - 1) * (1 + DK) + 1 + dd];
  for (int c = nch - 1; c > cc; c--) {
    const double A = ls.bA[(size_t)n * nch + c];
#pragma unroll
    for (int dd = 0; dd < DK; dd++) xv[dd] = fma(A, xv[dd], ls.bB[((size_t)n * nch + c) * DK + dd]);
  }
  for (int t = t1 - 1; t >= t0; t--) {
    if (mk[t] > 0.0f) {
      const double* fo = fw_n + (size_t)t * (1 + DK);
      const double Lc = fo[0];
#pragma unroll
      for (int dd = 0; dd < DK; dd++) {
        const double hh = fma(xv[dd], iq, fo[1 + dd]);
        xv[dd] = Lc * fma(Lc, hh, (double)w_n[(size_t)t * DK + dd]);
        v_out[((size_t)n * T + t) * DK + dd] = (float)xv[dd];
      }
    } else {
#pragma unroll
      for (int dd = 0; dd < DK; dd++) v_out[((size_t)n * T + t) * DK + dd] = 0.0f;  // kalman.py:142-143
    }
  }
}

template <int DK>
static int launch_location_scan(const double* stats, const float* mask, const float* w, const float* w_last, double q,
                                int N, int T, double* fw, double* llpart, float* v_out, double* loglik, LocScan ls,
                                cudaStream_t st) {
  const int nch = lsc_nchunks(T);
  const unsigned grid = (unsigned)((N * nch + 63) / 64);
  lsc_mobius_kernel<<<grid, 64, 0, st>>>(stats, mask, q, N, T, DK, ls);
  int rc = kpms_last_launch();
  if (rc) return rc;
  lsc_mean_map_kernel<DK><<<grid, 64, 0, st>>>(stats, mask, q, N, T, ls);
  rc = kpms_last_launch();
  if (rc) return rc;
  lsc_filter_kernel<DK><<<grid, 64, 0, st>>>(stats, mask, w, w_last, q, N, T, loglik != nullptr, fw, llpart, v_out, ls);
  rc = kpms_last_launch();
  if (rc) return rc;
  lsc_sample_kernel<DK><<<grid, 64, 0, st>>>(fw, mask, w, q, N, T, llpart, v_out, loglik, ls);
  return kpms_last_launch();
}

extern "C" int kpms_resample_location(const float* Y, const float* mask, const float* h,
                                      const float* x, const float* s, const double* Cdfull,
                                      const double* sigmasq, double sigmasq_loc, const float* w,
                                      const float* w_last, int N, int T, int k, int dk, int D,
                                      float* v_out, double* loglik, void* workspace,
                                      size_t workspace_bytes, void* stream) {
  KPMS_REQUIRE(Y && mask && h && x && s && Cdfull && sigmasq && w && w_last && v_out && workspace,
               KPMS_ERR_NULL);
  int rc = obs_check(N, T, k, dk, D);
  if (rc) return rc;
  KPMS_REQUIRE(T >= 2 && sigmasq_loc > 0.0, KPMS_ERR_SHAPE);
  KPMS_REQUIRE(kpms_aligned(workspace, 256) &&
                   workspace_bytes >= kpms_resample_location_workspace_bytes(N, T, dk),
               KPMS_ERR_WORKSPACE);
  cudaStream_t st = (cudaStream_t)stream;
  ObsDims dm{N, T, k, dk, D};
  double* stats = (double*)workspace;
  double* fw = (double*)((char*)workspace +
                         kpms_round_up((size_t)N * T * (2 + dk) * sizeof(double), 256));
  const size_t smem = ((size_t)k * dk * (D + 1) + k) * sizeof(double);
  const long long frames = (long long)N * T;
  location_stats_kernel<<<(unsigned)((frames + 255) / 256), 256, smem, st>>>(Y, h, x, s, Cdfull,
                                                                           sigmasq, dm, stats);
  rc = kpms_last_launch();
  if (rc) return rc;
  char* p2 = (char*)fw + kpms_round_up((size_t)N * T * (1 + dk) * sizeof(double), 256);
  double* llpart = (double*)p2;
  const int nch = loc_nchunks(T);
  const size_t units = (size_t)N * ((T - 1 + 127) / 128 + 1);
  int* tv = (int*)(p2 + kpms_round_up(units * sizeof(double), 256));
  if (kpms_opts.location_impl != 1) {   // default: exact two-level scan over time
    double* sc = (double*)((char*)tv + kpms_round_up((size_t)N * sizeof(int), 256));
    LocScan ls{sc, sc + units * 4, sc + units * 5, sc + units * (5 + dk), sc + units * (6 + dk)};
    return (dk == 2) ? launch_location_scan<2>(stats, mask, w, w_last, sigmasq_loc, N, T, fw, llpart, v_out, loglik, ls, st)
                     : launch_location_scan<3>(stats, mask, w, w_last, sigmasq_loc, N, T, fw, llpart, v_out, loglik, ls, st);
  }
  location_tv_kernel<<<N, 128, 0, st>>>(mask, T, tv);
  rc = kpms_last_launch();
  if (rc) return rc;
  const unsigned grid = (unsigned)((N * nch + 63) / 64);
  if (dk == 2) {
    location_forward_kernel<2><<<grid, 64, 0, st>>>(stats, mask, w_last, tv, sigmasq_loc, N, T,
                                                    loglik != nullptr, fw, llpart, v_out);
    rc = kpms_last_launch();
    if (rc) return rc;
    location_backward_kernel<2><<<grid, 64, 0, st>>>(fw, mask, w, sigmasq_loc, N, T, llpart, v_out,
                                                     loglik);
  } else {
    location_forward_kernel<3><<<grid, 64, 0, st>>>(stats, mask, w_last, tv, sigmasq_loc, N, T,
                                                    loglik != nullptr, fw, llpart, v_out);
    rc = kpms_last_launch();
    if (rc) return rc;
    location_backward_kernel<3><<<grid, 64, 0, st>>>(fw, mask, w, sigmasq_loc, N, T, llpart, v_out,
                                                     loglik);
  }
  return kpms_last_launch();
}
