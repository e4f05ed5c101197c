// a5-a7: continuous latents — Kalman filter + backward sampler for the switching AR model.
// Reference: gibbs.py:10-23 (resample_latents), kalman.py:80-156 (kalman_filter,
// kalman_sample), util.py:214-230 (ar_to_lds).
//
// The reference conditions in covariance form with four dense n x n inversions and one
// Cholesky per step (n = D*L).  This file computes the SAME Gaussian conditionals in
// information form (prototype + proof of equivalence: tools/proto_kalman_info.py):
//
//   carry (Lam_p, eta_p) = (S_pred^-1, S_pred^-1 m_pred), start Lam = I/10, eta = 0 (gibbs.py:18)
//   update   Lam_c = Lam_p + C^T R^-1 C         (touches only the newest D x D block, because
//            eta_c = eta_p + C^T R^-1 (y - d)    C is zero outside the last D columns, util.py:228)
//   smooth   Lam_s = Lam_c + A_z^T Qinv_z A_z  = inverse of kalman.py:131's S_cond
//            Lam_s = M^T M (M lower, factored bottom-up)  =>  L = M^-1 is exactly
//            cholesky(S_cond) of kalman.py:138 (lower, positive diagonal, unique)
//   predict  V = L^T [A_z^T Qinv_z | eta_c - A_z^T Qinv_z B_z]
//            Lam_p' = Qinv_z - V^T V,  eta_p' = Qinv_z B_z + V^T u            (Woodbury)
//   backward x_t = L (L^T (eta_c + A_z^T Qinv_z (x_{t+1} - B_z)) + w_t)       (kalman.py:129-140)
//
// The companion structure (util.py:214-230) makes A^T Qinv = [100 E | F_D] with E a shifted
// identity, so its first n-D columns of V are 100x columns of L^T that the triangular inverse
// produces anyway.  Per step: one bottom-up Cholesky, one 41-column triangular solve, one
// symmetric rank-n update — ~25k fp64 FMAs instead of the reference's ~250k flops.
//
// Kernels: per-state tables -> per-frame observation information (frame-parallel) ->
// forward filter (CTA per session or per session chunk) -> backward sampler (warp per
// session/chunk).  Masked steps leave the carry untouched and emit zeros; the final update
// is unconditional (kalman.py:109-117,142-143).
#include "common.cuh"
#include "dispatch.cuh"
#include <stdlib.h>

template <int NN, int DD>
struct KL {
  static constexpr int NTRI = NN * (NN + 1) / 2;
  static constexpr int NS = NN | 1;        // odd row stride (doubles) for W / Lp / Lfull
  static constexpr int NV = (NN + 1) | 1;  // odd row stride for V
  static constexpr int NC = NN + DD + 1;   // columns solved per step
  static constexpr int NJ = DD * (DD + 1) / 2;
  // per-state table
  static constexpr int TAB_G = 0;
  static constexpr int TAB_FD = TAB_G + NN * NN;
  static constexpr int TAB_QI = TAB_FD + NN * DD;
  static constexpr int TAB_F = TAB_QI + DD * DD;
  static constexpr int TAB_QB = TAB_F + NN;
  static constexpr int TAB_B = TAB_QB + DD;
  static constexpr int TAB_SC = TAB_B + DD;  // c_z, logdet Q_
  static constexpr int TAB_SIZE = (TAB_SC + 2 + 1) / 2 * 2;
  // per-frame observation information
  static constexpr int OBS_J = 0;
  static constexpr int OBS_E = NJ;
  static constexpr int OBS_C = NJ + DD;
  static constexpr int OBS_SIZE = (OBS_C + 1 + 1) / 2 * 2;
  // per-frame filter output
  static constexpr int FW_L = 0;
  static constexpr int FW_ETA = NTRI;
  static constexpr int FW_SIZE = (NTRI + NN + 1) / 2 * 2;
  static constexpr int BND = NTRI + NN;
};

struct KalWs {
  double* tab;
  double* tabw;  // packed per-state tables for the warp kernel (KW layout)
  double* obs;
  double* fw;
  double* llpart;
  double* bndf;
  double* bndx;
  unsigned long long* errs;
  int* tv;
  double* tabg;  // per-state G in MMA accumulator layout (blocked kernel)
  double* tabf;  // per-state [unit columns | F] in MMA B-operand layout (blocked kernel)
  double* fin;   // [N][BND] carry after the last transition step (input of kalman_final_kernel)
};

__host__ __device__ static inline int kal_nchunks(int Tt, int chunk_len) {
  if (chunk_len <= 0 || chunk_len >= Tt) return 1;
  return (Tt + chunk_len - 1) / chunk_len;
}

static size_t kal_layout(int N, int T, int D, int L, int K, int chunk_len, void* base, KalWs* ws) {
  const int NN = D * L, NTRI = NN * (NN + 1) / 2, NJ = D * (D + 1) / 2;
  const int TAB = (NN * NN + NN * D + D * D + NN + D + D + 2 + 1) / 2 * 2;
  const int OBS = (NJ + D + 1 + 1) / 2 * 2;
  const int FW = (NTRI + NN + 1) / 2 * 2;
  const int Tp = T - L + 1, Tt = Tp - 1;
  const int nch = kal_nchunks(Tt, chunk_len);
  size_t off = 0;
  char* b = (char*)base;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off += kpms_round_up(bytes, 256);
    return b ? (void*)(b + o) : nullptr;
  };
  void* p;
  p = take((size_t)K * TAB * 8); if (ws) ws->tab = (double*)p;
  const int TABW = (NTRI + NN * D + D * D + NN + D + 2 + 1) / 2 * 2;
  p = take((size_t)K * TABW * 8); if (ws) ws->tabw = (double*)p;
  p = take((size_t)N * Tp * OBS * 8); if (ws) ws->obs = (double*)p;
  p = take((size_t)N * Tp * FW * 8); if (ws) ws->fw = (double*)p;
  p = take((size_t)N * nch * 8); if (ws) ws->llpart = (double*)p;
  p = take((size_t)N * nch * 2 * (NTRI + NN) * 8); if (ws) ws->bndf = (double*)p;
  p = take((size_t)N * nch * 2 * NN * 8); if (ws) ws->bndx = (double*)p;
  p = take(2 * 8); if (ws) ws->errs = (unsigned long long*)p;
  p = take((size_t)N * 4); if (ws) ws->tv = (int*)p;
  {  // blocked kernel (kalman_blk.cuh: KB<NN, DD>)
    const int NBK = (NN + 1 + 7) / 8, NBLK = NBK * (NBK + 1) / 2, CB0 = (NN - D) / 8;
    p = take((size_t)K * NBLK * 64 * 8); if (ws) ws->tabg = (double*)p;
    p = take((size_t)K * (2 * NBK) * (NBK - CB0) * 32 * 8); if (ws) ws->tabf = (double*)p;
    p = take((size_t)N * (NTRI + NN) * 8); if (ws) ws->fin = (double*)p;
  }
  return off;
}

extern "C" size_t kpms_resample_latents_workspace_bytes(int N, int T, int k, int dk, int D, int L,
                                                        int K, int chunk_len) {
  if (N <= 0 || T <= L || D <= 0 || L <= 0 || K <= 0) return 0;
  (void)k; (void)dk;
  return kal_layout(N, T, D, L, K, chunk_len, nullptr, nullptr);
}

// ------------------------------------------------------------------ per-state tables
template <int NN, int DD>
struct KW {  // packed per-state table + per-warp shared-memory layout of the warp kernel
  static constexpr int NTRI = NN * (NN + 1) / 2;
  static constexpr int G = 0;                  // packed lower: [i(i+1)/2 + c]
  static constexpr int FD = G + NTRI;          // [q*DD + a]
  static constexpr int QI = FD + NN * DD;      // [a*DD + b]
  static constexpr int F = QI + DD * DD;       // NN
  static constexpr int QB = F + NN;            // DD
  static constexpr int SC = QB + DD;           // c_z, logdet Q_
  static constexpr int SIZE = (SC + 2 + 1) / 2 * 2;
  static constexpr int RS = 36;                // row stride (doubles) of the 32-row L / Z buffer
  static constexpr int OFF_ = NN - DD;
  static constexpr int NB = (NN + 1 + 7) / 8;  // 8-column blocks of T (NN columns of V and u)
  static constexpr int CB0 = OFF_ / 8;         // first column block holding a dense column of F
  static constexpr int TS = 8 * (NB - CB0) + 4;  // row stride of the dense blocks of T
  // rows of M packed back to back, each padded to an even length (16-byte aligned row starts)
  __host__ __device__ static constexpr int moff(int k) {
    int o = 0;
    for (int r = 0; r < k; r++) o += (r + 2) & ~1;
    return o;
  }
  static constexpr int MSZ = ((moff(NN) > 32 * TS ? moff(NN) : 32 * TS) + 1) & ~1;
  static constexpr int TABS = (SIZE - FD + 1) & ~1;  // shared-memory part of the table (all but G)
  // Scratch timeline of one step: [M rows] -> [dense T | packed L] -> [Z over both].
  static constexpr int SM_MS = 0;              // M rows (S2/S3), then the dense blocks of T
  static constexpr int SM_LS = SM_MS + MSZ;    // L packed by rows: L[q][r] at q(q+1)/2 + r
  static constexpr int LPK = (NTRI + 1) & ~1;
  static constexpr int ZSZ = (NN + 1) * RS;    // Z: rows 0..NN, row stride RS, starts at SM_MS
  static constexpr int SCR = (MSZ + LPK > ZSZ ? MSZ + LPK : ZSZ);
  static constexpr int SM_TAB = SM_MS + ((SCR + 1) & ~1);
  static constexpr int SM_DV = SM_TAB + TABS;
  static constexpr int SM_HV = SM_DV + 32;
  static constexpr int SM_TOTAL = SM_HV + 32;
};

template <int NN, int DD>
__global__ void kalman_tables_kernel(const double* __restrict__ Ab, const double* __restrict__ Q,
                                     double* __restrict__ tab, double* __restrict__ tabw) {
  using Y = KL<NN, DD>;
  using Wt = KW<NN, DD>;
  double* tw = tabw + (size_t)blockIdx.x * Wt::SIZE;
  const int k = blockIdx.x, lane = threadIdx.x;
  __shared__ double Lq[DD * DD], Li[DD * DD], Qi[DD * DD], QA[DD * NN], bb[DD], qb[DD];
  const double* Abk = Ab + (size_t)k * DD * (NN + 1);
  const double* Qk = Q + (size_t)k * DD * DD;
  double* t = tab + (size_t)k * Y::TAB_SIZE;
  if (lane == 0) {
    double ls = 0.0;
    for (int j = 0; j < DD; j++) {
      double s = Qk[j * DD + j];
      for (int m = 0; m < j; m++) s -= Lq[j * DD + m] * Lq[j * DD + m];
      const double d = sqrt(s);
      Lq[j * DD + j] = d;
      ls += log(d);
      for (int i = j + 1; i < DD; i++) {
        double v = Qk[i * DD + j];
        for (int m = 0; m < j; m++) v -= Lq[i * DD + m] * Lq[j * DD + m];
        Lq[i * DD + j] = v / d;
      }
    }
    // Li = Lq^-1 (lower)
    for (int c = 0; c < DD; c++) {
      for (int i = 0; i < DD; i++) {
        double v = (i == c) ? 1.0 : 0.0;
        for (int m = c; m < i; m++) v -= Lq[i * DD + m] * Li[m * DD + c];
        Li[i * DD + c] = (i < c) ? 0.0 : v / Lq[i * DD + i];
      }
    }
    t[Y::TAB_SC + 1] = 2.0 * ls + (double)(NN - DD) * log(1e-2);
    tw[Wt::SC + 1] = t[Y::TAB_SC + 1];
  }
  __syncwarp();
  for (int e = lane; e < DD * DD; e += 32) {
    const int a = e / DD, b = e % DD;
    double s = 0.0;
    for (int m = (a > b ? a : b); m < DD; m++) s += Li[m * DD + a] * Li[m * DD + b];
    Qi[e] = s;
    t[Y::TAB_QI + e] = s;
    tw[Wt::QI + e] = s;
  }
  for (int a = lane; a < DD; a += 32) bb[a] = Abk[a * (NN + 1) + NN];
  __syncwarp();
  for (int e = lane; e < DD * NN; e += 32) {
    const int a = e / NN, c = e % NN;
    double s = 0.0;
    for (int b = 0; b < DD; b++) s += Qi[a * DD + b] * Abk[b * (NN + 1) + c];
    QA[e] = s;
  }
  for (int a = lane; a < DD; a += 32) {
    double s = 0.0;
    for (int b = 0; b < DD; b++) s += Qi[a * DD + b] * bb[b];
    qb[a] = s;
    t[Y::TAB_QB + a] = s;
    tw[Wt::QB + a] = s;
    t[Y::TAB_B + a] = bb[a];
  }
  __syncwarp();
  for (int e = lane; e < NN * NN; e += 32) {
    const int r = e / NN, c = e % NN;
    double s = (r == c && r >= DD) ? 100.0 : 0.0;
    for (int a = 0; a < DD; a++) s += Abk[a * (NN + 1) + r] * QA[a * NN + c];
    t[Y::TAB_G + e] = s;
    if (c <= r) tw[Wt::G + r * (r + 1) / 2 + c] = s;
  }
  for (int e = lane; e < NN * DD; e += 32) {
    const int r = e / DD, a = e % DD;
    t[Y::TAB_FD + e] = QA[a * NN + r];
    tw[Wt::FD + e] = QA[a * NN + r];
  }
  for (int r = lane; r < NN; r += 32) {
    double s = 0.0;
    for (int a = 0; a < DD; a++) s += QA[a * NN + r] * bb[a];
    t[Y::TAB_F + r] = s;
    tw[Wt::F + r] = s;
  }
  if (lane == 0) {
    double s = 0.0;
    for (int a = 0; a < DD; a++) s += bb[a] * qb[a];
    t[Y::TAB_SC + 0] = s;
    tw[Wt::SC + 0] = s;
  }
}

// ------------------------------------------------------------------ observation information
// One thread per LDS step tau (frame tau+L-1):  J = C_D^T R^-1 C_D (packed lower),
// e = C_D^T R^-1 (ytild - d),  c = -1/2 sum_p [log(2 pi R_p) + (ytild_p - d_p)^2 / R_p],
// ytild = R(-h)(Y - v) (util.py:140-166), R_p = s_kp sigmasq_kp repeated over dk (gibbs.py:17).
template <int NN, int DD>
__global__ void __launch_bounds__(128)
    kalman_obs_kernel(const float* __restrict__ Y, const float* __restrict__ v,
                      const float* __restrict__ h, const float* __restrict__ s,
                      const double* __restrict__ Cdfull, const double* __restrict__ sigmasq,
                      int N, int T, int k, int dk, int Lg, double* __restrict__ obs) {
  using Y_ = KL<NN, DD>;
  extern __shared__ __align__(16) double sm[];
  double* Cs = sm;                          // (k*dk) x (DD+1)
  double* CC = Cs + (size_t)k * dk * (DD + 1);  // k x NJ
  double* sg = CC + (size_t)k * Y_::NJ;     // k
  for (int i = threadIdx.x; i < k * dk * (DD + 1); i += blockDim.x) Cs[i] = Cdfull[i];
  for (int i = threadIdx.x; i < k; i += blockDim.x) sg[i] = sigmasq[i];
  __syncthreads();
  for (int i = threadIdx.x; i < k * Y_::NJ; i += blockDim.x) {
    const int kp = i / Y_::NJ, ab = i % Y_::NJ;
    int a = 0;
    while ((a + 1) * (a + 2) / 2 <= ab) a++;
    const int b = ab - a * (a + 1) / 2;
    double sacc = 0.0;
    for (int dim = 0; dim < dk; dim++)
      sacc += Cs[(size_t)(kp * dk + dim) * (DD + 1) + a] * Cs[(size_t)(kp * dk + dim) * (DD + 1) + b];
    CC[i] = sacc;
  }
  __syncthreads();
  const int Tp = T - Lg + 1;
  const long long total = (long long)N * Tp;
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  // per-warp staging of the 32 records of a warp: threads fill their own record (row stride odd:
  // conflict-free), then the warp writes the 32 x OBS_SIZE doubles, which are contiguous in `obs`,
  // as whole 256-byte lines (a thread writing its own 528-byte record touches 32 sectors per store)
  constexpr int SST = Y_::OBS_SIZE | 1;
  double* stage = sg + k + (size_t)wib * 32 * SST;
  const bool act = g < total;
  if (act) {
    const int n = (int)(g / Tp), tau = (int)(g % Tp);
    const size_t f = (size_t)n * T + tau + Lg - 1;
    double J[Y_::NJ], ev[DD];
#pragma unroll
    for (int i = 0; i < Y_::NJ; i++) J[i] = 0.0;
#pragma unroll
    for (int i = 0; i < DD; i++) ev[i] = 0.0;
    double ct = 0.0;
    double sn, cs;
    sincos((double)h[f], &sn, &cs);
    const double v0 = (double)v[f * dk], v1 = (double)v[f * dk + 1];
    const double v2 = (dk > 2) ? (double)v[f * dk + 2] : 0.0;
    const bool vec = kp4_vec_ok(Y, s, k, dk);
    for (int kp0 = 0; kp0 < k; kp0 += 4) {
      Kp4 q;
      load_kp4(Y + f * k * dk, s + f * k, kp0, k, dk, vec, q);
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int kp = kp0 + j;
        if (kp >= k) continue;
        const double R = (double)q.s[j] * sg[kp];
        const double ri = 1.0 / R;
        double yt[3];
        rot_inv(cs, sn, (double)KP4_Y(q, j, dk, 0) - v0, (double)KP4_Y(q, j, dk, 1) - v1, yt[0], yt[1]);
        yt[2] = (dk > 2) ? (double)KP4_Y(q, j, dk, 2) - v2 : 0.0;
        double qq = 0.0;
#pragma unroll
        for (int dim = 0; dim < 3; dim++) {
          if (dim < dk) {
            const double* row = Cs + (size_t)(kp * dk + dim) * (DD + 1);
            const double yy = yt[dim] - row[DD];
            qq = fma(yy, yy, qq);
            const double wgt = yy * ri;
#pragma unroll
            for (int a = 0; a < DD; a++) ev[a] = fma(row[a], wgt, ev[a]);
          }
        }
        ct += -0.5 * (double)dk * (KPMS_LOG_2PI + log(R)) - 0.5 * qq * ri;
        const double* cc = CC + (size_t)kp * Y_::NJ;
#pragma unroll
        for (int i = 0; i < Y_::NJ; i++) J[i] = fma(cc[i], ri, J[i]);
      }
    }
    double* o = stage + (size_t)lane * SST;
#pragma unroll
    for (int i = 0; i < Y_::NJ; i++) o[Y_::OBS_J + i] = J[i];
#pragma unroll
    for (int i = 0; i < DD; i++) o[Y_::OBS_E + i] = ev[i];
    o[Y_::OBS_C] = ct;
    if (Y_::OBS_SIZE > Y_::OBS_C + 1) o[Y_::OBS_C + 1] = 0.0;
  }
  __syncwarp();
  const long long g0 = g - lane;                      // first record of this warp
  const long long left = total - g0;
  const int nrec = left >= 32 ? 32 : (left > 0 ? (int)left : 0);
  double* dst = obs + (size_t)g0 * Y_::OBS_SIZE;
  for (int i = lane; i < nrec * Y_::OBS_SIZE; i += 32) dst[i] = stage[(size_t)(i / Y_::OBS_SIZE) * SST + (i % Y_::OBS_SIZE)];
}

// The same records by FP64 tensor cores, for dk == 2 and k % 4 == 0 (every BASELINE config).  Both
// J = sum_kp (1/R_kp) C_kp^T C_kp and e = sum_(kp,dim) C_(kp,dim) (ytild - d)/R are products of a
// per-frame weight row with a table that depends on the keypoint only, so eight frames make the M
// dimension of mma.m8n8k4: lane (g, qd) prepares frame g's keypoints qd, qd+4, ... (its A operands),
// the tables come from shared memory as B operands, and the accumulator rows are the records.
// A warp takes 32 frames at a time: sincos(h) once per lane, then four batches of eight frames.
// The thread-per-frame kernel above needs ~2400 instructions per frame at 242 registers; this one
// ~1000 per 32 frames, so the pass is bound by its record writes instead of instruction issue.
#define KO_DMMA(c, a, b)                                                                           \
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" \
               : "+d"((c)[0]), "+d"((c)[1])                                                        \
               : "d"(a), "d"(b))
template <int NN, int DD>
__global__ void __launch_bounds__(256)
    kalman_obs_mma_kernel(const float* __restrict__ Y, const float* __restrict__ v,
                          const float* __restrict__ h, const float* __restrict__ s,
                          const double* __restrict__ Cdfull, const double* __restrict__ sigmasq,
                          int N, int T, int k, int Lg, int want_c, double* __restrict__ obs) {
  using Y_ = KL<NN, DD>;
  constexpr int NJ = Y_::NJ, NJB = (NJ + 7) / 8, NEB = (DD + 7) / 8;
  constexpr int NJP = 8 * NJB + 4;   // row strides (doubles) of the two tables: the four k rows of a B
  constexpr int CSP = 8 * NEB + 2;   // fragment land on distinct bank pairs (60 -> 12, 2*18 -> 4 mod 16)
  constexpr unsigned FULL = 0xffffffffu;
  extern __shared__ __align__(16) double sm[];
  double* CC = sm;                          // k x NJP : C_kp^T C_kp packed lower, zero padded
  double* Cs = CC + (size_t)k * NJP;        // (2k) x CSP : rows of C_D, zero padded
  double* dd = Cs + (size_t)2 * k * CSP;    // 2k : d
  double* sg = dd + 2 * k;                  // k : sigmasq
  for (int i = threadIdx.x; i < 2 * k * CSP; i += blockDim.x) {
    const int r = i / CSP, c = i % CSP;
    Cs[i] = c < DD ? Cdfull[(size_t)r * (DD + 1) + c] : 0.0;
  }
  for (int i = threadIdx.x; i < 2 * k; i += blockDim.x) dd[i] = Cdfull[(size_t)i * (DD + 1) + DD];
  for (int i = threadIdx.x; i < k; i += blockDim.x) sg[i] = sigmasq[i];
  __syncthreads();
  for (int i = threadIdx.x; i < k * NJP; i += blockDim.x) {
    const int kp = i / NJP, ab = i % NJP;
    double sacc = 0.0;
    if (ab < NJ) {
      int a = 0;
      while ((a + 1) * (a + 2) / 2 <= ab) a++;
      const int b = ab - a * (a + 1) / 2;
      sacc = Cs[(2 * kp) * CSP + a] * Cs[(2 * kp) * CSP + b] + Cs[(2 * kp + 1) * CSP + a] * Cs[(2 * kp + 1) * CSP + b];
    }
    CC[i] = sacc;
  }
  __syncthreads();
  const int Tp = T - Lg + 1;
  const long long total = (long long)N * Tp;
  const int lane = threadIdx.x & 31, g = lane >> 2, qd = lane & 3;
  const long long nwarp = (long long)gridDim.x * (blockDim.x >> 5);
  const int KS = k >> 2;
  for (long long grp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); grp * 32 < total; grp += nwarp) {
    const long long g0 = grp * 32;
    // lane i: frame g0 + i
    double cs_i = 1.0, sn_i = 0.0;
    float v0_i = 0.0f, v1_i = 0.0f;
    long long f_i = 0;
    if (g0 + lane < total) {
      const int n = (int)((g0 + lane) / Tp), tau = (int)((g0 + lane) % Tp);
      f_i = (long long)n * T + tau + Lg - 1;
      sincos((double)h[f_i], &sn_i, &cs_i);
      const float2 vv = __ldg(reinterpret_cast<const float2*>(v) + f_i);
      v0_i = vv.x;
      v1_i = vv.y;
    }
#pragma unroll 1
    for (int b = 0; b < 4; b++) {
      const int src = 8 * b + g;
      const double cs = __shfl_sync(FULL, cs_i, src), sn = __shfl_sync(FULL, sn_i, src);
      const double v0 = (double)__shfl_sync(FULL, v0_i, src), v1 = (double)__shfl_sync(FULL, v1_i, src);
      const long long f = __shfl_sync(FULL, f_i, src);
      const bool act = g0 + src < total;
      const float2* Yf = reinterpret_cast<const float2*>(Y) + (size_t)f * k;
      const float* sf = s + (size_t)f * k;
      double Ja[NJB][2], Ea[NEB][2];
#pragma unroll
      for (int i = 0; i < NJB; i++) Ja[i][0] = Ja[i][1] = 0.0;
#pragma unroll
      for (int i = 0; i < NEB; i++) Ea[i][0] = Ea[i][1] = 0.0;
      double qq = 0.0, lr = 0.0, pr = 1.0;
      for (int ks = 0; ks < KS; ks++) {
        const int kp = 4 * ks + qd;
        double ri = 0.0, w0 = 0.0, w1 = 0.0;
        if (act) {
          const float2 yv = __ldg(Yf + kp);
          const double R = (double)__ldg(sf + kp) * sg[kp];
          ri = rcp_pos_d(R);
          pr *= R;
          double y0, y1;
          rot_inv(cs, sn, (double)yv.x - v0, (double)yv.y - v1, y0, y1);
          y0 -= dd[2 * kp];
          y1 -= dd[2 * kp + 1];
          w0 = y0 * ri;
          w1 = y1 * ri;
          qq = fma(y0, w0, qq);
          qq = fma(y1, w1, qq);
        }
        if (want_c && (ks & 3) == 3) {   // fold the running product of R into a log before it can leave the double range
          lr += log(pr);
          pr = 1.0;
        }
        const double* ccr = CC + (size_t)kp * NJP + g;
#pragma unroll
        for (int nb = 0; nb < NJB; nb++) KO_DMMA(Ja[nb], ri, ccr[8 * nb]);
        const double* c0 = Cs + (size_t)(2 * kp) * CSP + g;
#pragma unroll
        for (int nb = 0; nb < NEB; nb++) {
          KO_DMMA(Ea[nb], w0, c0[8 * nb]);
          KO_DMMA(Ea[nb], w1, c0[CSP + 8 * nb]);
        }
      }
      // c = -1/2 sum_p [log(2 pi R_p) + (ytild_p - d_p)^2 / R_p], two coordinates per keypoint; only the marginal
      // log-likelihood reads it
      double ct = 0.0;
      if (want_c) {
        if (KS & 3) lr += log(pr);
        ct = -lr - 0.5 * qq;
        ct += __shfl_xor_sync(FULL, ct, 1);
        ct += __shfl_xor_sync(FULL, ct, 2);
        ct -= (double)k * KPMS_LOG_2PI;
      }
      if (act) {
        double* o = obs + (size_t)(g0 + src) * Y_::OBS_SIZE;
#pragma unroll
        for (int nb = 0; nb < NJB; nb++) {
          const int i0 = 8 * nb + 2 * qd;
          if (i0 + 1 < NJ) *reinterpret_cast<double2*>(o + Y_::OBS_J + i0) = make_double2(Ja[nb][0], Ja[nb][1]);
          else if (i0 < NJ) o[Y_::OBS_J + i0] = Ja[nb][0];
        }
#pragma unroll
        for (int nb = 0; nb < NEB; nb++)
#pragma unroll
          for (int e = 0; e < 2; e++) {
            const int a = 8 * nb + 2 * qd + e;
            if (a < DD) o[Y_::OBS_E + a] = Ea[nb][e];
          }
        if (qd == 0) {
          o[Y_::OBS_C] = ct;
          if (Y_::OBS_SIZE > Y_::OBS_C + 1) o[Y_::OBS_C + 1] = 0.0;
        }
      }
    }
  }
}

// number of unmasked transition steps per session (trailing padding assumed for chunking)
__global__ void kalman_tv_kernel(const float* __restrict__ mask, int T, int Lg, int Tt,
                                 int* __restrict__ tv) {
  const int n = blockIdx.x;
  __shared__ int cnt;
  if (threadIdx.x == 0) cnt = 0;
  __syncthreads();
  int c = 0;
  for (int t = threadIdx.x; t < Tt; t += blockDim.x) c += mask[(size_t)n * T + t + Lg - 1] > 0.0f;
  atomicAdd(&cnt, c);
  __syncthreads();
  if (threadIdx.x == 0) tv[n] = cnt;
}

// Bottom-up Cholesky  W = M^T M  (M lower, stored in the lower triangle of W, row stride NS),
// dinv[i] = 1 / M[i][i].  All threads of the CTA call this.
template <int NN, int NS, int NT>
__device__ __forceinline__ void rev_cholesky(double* W, double* dinv, const unsigned char* ti,
                                             const unsigned char* tj) {
  const int tid = threadIdx.x;
#pragma unroll 1
  for (int kk = NN - 1; kk >= 1; kk--) {
    __syncthreads();
    const double ip = 1.0 / W[kk * NS + kk];
    const int cnt = kk * (kk + 1) / 2;
    for (int e = tid; e < cnt; e += NT) {
      const int i = ti[e], j = tj[e];
      W[i * NS + j] = fma(-W[kk * NS + i] * ip, W[kk * NS + j], W[i * NS + j]);
    }
  }
  __syncthreads();
  if (tid < NN) dinv[tid] = 1.0 / sqrt(W[tid * NS + tid]);
  __syncthreads();
  for (int e = tid; e < NN * (NN + 1) / 2; e += NT) {
    const int i = ti[e], j = tj[e];
    W[i * NS + j] *= dinv[i];  // diagonal: W_ii / sqrt(W_ii) = sqrt(W_ii)
  }
  __syncthreads();
}

#define KAL_NT 128

template <int NN, int DD>
__global__ void __launch_bounds__(KAL_NT, 1)
    kalman_forward_kernel(const float* __restrict__ mask, const int* __restrict__ z,
                          const float* __restrict__ w_last, KalWs ws, int N, int T, int Lg,
                          int chunk_len, int burn_in, int want_ll) {
  using Y = KL<NN, DD>;
  constexpr int NS = Y::NS, NV = Y::NV, NTRI = Y::NTRI, NC = Y::NC, NT = KAL_NT, OFF = NN - DD;
  static_assert(NC <= NT && Y::OBS_SIZE <= NT, "thread count too small");
  extern __shared__ __align__(16) double sm[];
  double* W = sm;                   // NN*NS
  double* Vm = W + NN * NS;         // NN*NV
  double* Lp = Vm + NN * NV;        // NN*NS
  double* eta_p = Lp + NN * NS;     // NN
  double* eta_c = eta_p + NN;       // NN
  double* dinv = eta_c + NN;        // NN
  double* Gs = dinv + NN;           // NN*NN
  double* FDs = Gs + NN * NN;       // NN*DD
  double* Qis = FDs + NN * DD;      // DD*DD
  double* fs = Qis + DD * DD;       // NN
  double* qbs = fs + NN;            // DD
  double* scs = qbs + DD;           // 2
  double* obs_s = scs + 2;          // OBS_SIZE
  unsigned char* ti = reinterpret_cast<unsigned char*>(obs_s + Y::OBS_SIZE);
  unsigned char* tj = ti + NTRI;
  const int tid = threadIdx.x;
  const int Tp = T - Lg + 1, Tt = Tp - 1;
  const int nch = kal_nchunks(Tt, chunk_len);
  const int n = blockIdx.x / nch, c = blockIdx.x % nch;
  const int t0 = (nch == 1) ? 0 : c * chunk_len;
  const int t1 = (nch == 1) ? Tt : min(Tt, t0 + chunk_len);
  int tb = 0;
  if (c > 0) {
    const int tvn = ws.tv[n];
    tb = max(0, min(t0, tvn) - burn_in);
  }
  for (int e = tid; e < NTRI; e += NT) {
    int i = 0;
    while ((i + 1) * (i + 2) / 2 <= e) i++;
    ti[e] = (unsigned char)i;
    tj[e] = (unsigned char)(e - i * (i + 1) / 2);
  }
  // prior carry (gibbs.py:18: mu0 = 0, S0 = 10 I)
  for (int e = tid; e < NN * NS; e += NT) Lp[e] = 0.0;
  __syncthreads();
  if (tid < NN) {
    Lp[tid * NS + tid] = 0.1;
    eta_p[tid] = 0.0;
  }
  double llacc = (c == 0) ? 0.5 * (double)NN * log(0.1) : 0.0;
  const float* mk = mask + (size_t)n * T + (Lg - 1);
  const int* zn = z + (size_t)n * Tt;
  const double* obs_n = ws.obs + (size_t)n * Tp * Y::OBS_SIZE;
  double* fw_n = ws.fw + (size_t)n * Tp * Y::FW_SIZE;
  int cur_z = -1;
  double obs_reg = 0.0;
  if (tid < Y::OBS_SIZE && tb < Tp) obs_reg = obs_n[(size_t)tb * Y::OBS_SIZE + tid];
  __syncthreads();

  for (int tau = tb; tau < t1; tau++) {
    if (tau == t0 && c > 0) {  // warmed-up carry at the chunk start, for verification
      double* bd = ws.bndf + ((size_t)(n * nch + c) * 2 + 0) * Y::BND;
      for (int e = tid; e < NTRI; e += NT) bd[e] = Lp[ti[e] * NS + tj[e]];
      if (tid < NN) bd[NTRI + tid] = eta_p[tid];
    }
    const bool live = mk[tau] > 0.0f;
    const double obs_cur = obs_reg;
    if (tid < Y::OBS_SIZE) obs_reg = obs_n[(size_t)(tau + 1) * Y::OBS_SIZE + tid];  // tau+1 <= Tp-1
    if (!live) continue;  // kalman.py:109-111: carry unchanged (uniform branch)
    const bool store = tau >= t0;
    const int zt = zn[tau];
    if (zt != cur_z) {
      const double* tb_ = ws.tab + (size_t)zt * Y::TAB_SIZE;
      for (int e = tid; e < NN * NN; e += NT) Gs[e] = tb_[Y::TAB_G + e];
      for (int e = tid; e < NN * DD; e += NT) FDs[e] = tb_[Y::TAB_FD + e];
      for (int e = tid; e < DD * DD; e += NT) Qis[e] = tb_[Y::TAB_QI + e];
      if (tid < NN) fs[tid] = tb_[Y::TAB_F + tid];
      if (tid < DD) qbs[tid] = tb_[Y::TAB_QB + tid];
      if (tid < 2) scs[tid] = tb_[Y::TAB_SC + tid];
      cur_z = zt;
    }
    if (tid < Y::OBS_SIZE) obs_s[tid] = obs_cur;
    __syncthreads();
    // S1: Lam_s = Lam_p + C^T R^-1 C + A^T Qinv A (lower triangle), eta_c
    for (int e = tid; e < NTRI; e += NT) {
      const int i = ti[e], j = tj[e];
      double v = Lp[i * NS + j] + Gs[i * NN + j];
      if (j >= OFF) v += obs_s[Y::OBS_J + (i - OFF) * (i - OFF + 1) / 2 + (j - OFF)];
      W[i * NS + j] = v;
    }
    if (tid < NN) {
      const double ec = eta_p[tid] + ((tid >= OFF) ? obs_s[Y::OBS_E + tid - OFF] : 0.0);
      eta_c[tid] = ec;
      if (store) fw_n[(size_t)tau * Y::FW_SIZE + Y::FW_ETA + tid] = ec;
    }
    // S2
    rev_cholesky<NN, NS, NT>(W, dinv, ti, tj);
    // S3: solve M^T X = [I | F_D | eta_c - f]; thread c owns one column in registers
    if (tid < NC) {
      double col[NN];
      int rstart = NN - 1;
      if (tid < NN) {
#pragma unroll
        for (int r = 0; r < NN; r++) col[r] = (r == tid) ? 1.0 : 0.0;
        rstart = tid;
      } else if (tid < NN + DD) {
#pragma unroll
        for (int r = 0; r < NN; r++) col[r] = FDs[r * DD + (tid - NN)];
      } else {
#pragma unroll
        for (int r = 0; r < NN; r++) col[r] = eta_c[r] - fs[r];
      }
#pragma unroll
      for (int r = NN - 1; r >= 0; r--) {
        if (r <= rstart) {
          const double vv = col[r] * dinv[r];
          col[r] = vv;
#pragma unroll
          for (int q = 0; q < r; q++) col[q] = fma(-W[r * NS + q], vv, col[q]);
        }
      }
      if (tid < NN) {
        if (store) {  // row tid of L = cholesky(S_cond): L[tid][r] = (M^-T)[r][tid]
          double* lrow = fw_n + (size_t)tau * Y::FW_SIZE + Y::FW_L + tid * (tid + 1) / 2;
#pragma unroll
          for (int r = 0; r < NN; r++)
            if (r <= tid) lrow[r] = col[r];
        }
        if (tid >= DD) {
#pragma unroll
          for (int r = 0; r < NN; r++) Vm[r * NV + (tid - DD)] = 100.0 * col[r];
        }
      } else if (tid < NN + DD) {
#pragma unroll
        for (int r = 0; r < NN; r++) Vm[r * NV + OFF + (tid - NN)] = col[r];
      } else {
#pragma unroll
        for (int r = 0; r < NN; r++) Vm[r * NV + NN] = col[r];
      }
    }
    __syncthreads();
    // S4: Lam_p' = Qinv_ - V^T V,  eta_p' = Qinv_ B_ + V^T u
    for (int e = tid; e < NTRI + NN; e += NT) {
      if (e < NTRI) {
        const int a = ti[e], b = tj[e];          // a >= b
        const int rmax = (b < OFF) ? DD + b : NN - 1;  // column b of V is zero below this row
        double acc = 0.0;
        for (int r = 0; r <= rmax; r++) acc = fma(Vm[r * NV + a], Vm[r * NV + b], acc);
        double qv;
        if (a < OFF) qv = (a == b) ? 100.0 : 0.0;
        else qv = (b >= OFF) ? Qis[(a - OFF) * DD + (b - OFF)] : 0.0;
        Lp[a * NS + b] = qv - acc;
      } else {
        const int a = e - NTRI;
        const int rmax = (a < OFF) ? DD + a : NN - 1;
        double acc = 0.0;
        for (int r = 0; r <= rmax; r++) acc = fma(Vm[r * NV + a], Vm[r * NV + NN], acc);
        eta_p[a] = ((a >= OFF) ? qbs[a - OFF] : 0.0) + acc;
      }
    }
    if (want_ll && store && tid == NT - 1) {
      double uu = 0.0, pd = 1.0;
      for (int r = 0; r < NN; r++) {
        uu = fma(Vm[r * NV + NN], Vm[r * NV + NN], uu);
        pd *= dinv[r];
      }
      llacc += obs_s[Y::OBS_C] - 0.5 * scs[1] + log(pd) + 0.5 * (uu - scs[0]);
    }
    __syncthreads();
  }
  // carry at the chunk end, for verification against the next chunk's warmed-up start
  if (nch > 1) {
    double* bd = ws.bndf + ((size_t)(n * nch + c) * 2 + 1) * Y::BND;
    for (int e = tid; e < NTRI; e += NT) bd[e] = Lp[ti[e] * NS + tj[e]];
    if (tid < NN) bd[NTRI + tid] = eta_p[tid];
  }
  if (c == nch - 1) {
    // unconditional final update (kalman.py:117) and the draw of the last state (kalman.py:150)
    if (tid < Y::OBS_SIZE) obs_s[tid] = obs_n[(size_t)(Tp - 1) * Y::OBS_SIZE + tid];
    __syncthreads();
    for (int e = tid; e < NTRI; e += NT) {
      const int i = ti[e], j = tj[e];
      double v = Lp[i * NS + j];
      if (j >= OFF) v += obs_s[Y::OBS_J + (i - OFF) * (i - OFF + 1) / 2 + (j - OFF)];
      W[i * NS + j] = v;
    }
    if (tid < NN) eta_c[tid] = eta_p[tid] + ((tid >= OFF) ? obs_s[Y::OBS_E + tid - OFF] : 0.0);
    rev_cholesky<NN, NS, NT>(W, dinv, ti, tj);
    if (tid == 0) {
      // runs once per session: serial, shared-memory scratch (column NN of Vm), no unrolling
      double* col = Vm + NN;  // element r at col[r * NV]
      double tt = 0.0, pd = 1.0;
#pragma unroll 1
      for (int r = 0; r < NN; r++) col[r * NV] = eta_c[r];
#pragma unroll 1
      for (int r = NN - 1; r >= 0; r--) {  // t = M^-T eta_c
        const double vv = col[r * NV] * dinv[r];
        col[r * NV] = vv;
        tt = fma(vv, vv, tt);
        pd *= dinv[r];
#pragma unroll 1
        for (int q = 0; q < r; q++) col[q * NV] = fma(-W[r * NS + q], vv, col[q * NV]);
      }
#pragma unroll 1
      for (int r = 0; r < NN; r++) col[r * NV] += (double)w_last[(size_t)n * NN + r];
      double* xl = fw_n + (size_t)(Tp - 1) * Y::FW_SIZE;
#pragma unroll 1
      for (int r = 0; r < NN; r++) {  // x = M^-1 (t + w): forward substitution, column sweep
        const double xv = col[r * NV] * dinv[r];
        xl[r] = xv;
#pragma unroll 1
        for (int q = r + 1; q < NN; q++) col[q * NV] = fma(-W[q * NS + r], xv, col[q * NV]);
      }
      ws.llpart[(size_t)n * nch + c] = obs_s[Y::OBS_C] + log(pd) + 0.5 * tt;
    }
    // the running sum lives in thread NT-1
    __syncthreads();
    if (tid == NT - 1) ws.llpart[(size_t)n * nch + c] += llacc;
  } else if (tid == NT - 1) {
    ws.llpart[(size_t)n * nch + c] = llacc;
  }
}

// ------------------------------------------------------------------ forward filter, warp per chunk
// One WARP walks one session (or session chunk); no block barriers.  Lane c owns COLUMN c of
// every n x n matrix in registers (static indexing, loops fully unrolled); a row is handed to
// the other lanes by publishing it to a per-warp shared-memory row buffer (one store, every lane
// writes its own element) and is consumed with broadcast loads.  Per step:
//   W = Lam_p + G_z + J                      (column c, rows >= c)
//   bottom-up Cholesky  W = M^T M            (30 pivots: shuffle pivot, rsqrt, publish row)
//   L = M^-1                                 (forward substitution against the published rows)
//   T = L^T [F | eta_c - f]                  (lane c: column c of F; lane NN: the mean column)
//   Lam_p' = Qinv_ - T^T T, eta_p' = qb_ + T^T u      (rows of T published, broadcast back)
// ~4.2 k warp instructions per step, of which ~2.1 k are DFMA; eight such warps per SM.
#define KW_WARPS 12
#define KW_LOCKSTEP 1

// 1/sqrt(x) for the pivots of an SPD matrix (x > 0, normal): hardware seed (2^-22) and one
// third-order correction, without rsqrt()'s special-case handling.
__device__ __forceinline__ double rsqrt_pos(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-x, y * y, 1.0);
  const double p = fma(0.375, e, 0.5);
  return fma(y * e, p, y);
}

// 1/x for the pivots of an SPD matrix (x > 0, normal): hardware seed and one third-order correction.
__device__ __forceinline__ double rcp_pos(double x) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-x, y, 1.0);
  return fma(y, fma(e, e, e), y);
}

template <int NN, int DD>
__global__ void __launch_bounds__(KW_WARPS * 32, 1)
    kalman_forward_warp_kernel(const float* __restrict__ mask, const int* __restrict__ z,
                               const float* __restrict__ w_last, KalWs ws, int N, int T, int Lg,
                               int chunk_len, int burn_in, int want_ll) {
  using Y = KL<NN, DD>;
  using Wt = KW<NN, DD>;
  static_assert(NN <= 30, "lane NN carries the mean column");
  constexpr int OFF = NN - DD, RS = Wt::RS, NTRI = Y::NTRI;
  constexpr unsigned FULL = 0xffffffffu;
  extern __shared__ __align__(16) double sm[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  double* S = sm + (size_t)wib * Wt::SM_TOTAL;
  double* Ms = S + Wt::SM_MS;   // rows of M, later rows of T
  double* Lk = S + Wt::SM_LS;   // L packed by rows
  double* Zs = S + Wt::SM_MS;   // Z (after T and L are dead); scratch of the final section
  double* Tb = S + Wt::SM_TAB - Wt::FD;  // tables of the current state except G (indexed with the KW offsets)
  const double* Gg = ws.tabw;   // G of the current state is read from global memory (L1/L2 resident)
  double* dv = S + Wt::SM_DV;   // 1 / M[i][i]
  double* hv = S + Wt::SM_HV;   // eta_c - f  (then scratch)
  const int Tp = T - Lg + 1, Tt = Tp - 1;
  const int nch = kal_nchunks(Tt, chunk_len);
  const int unit0 = blockIdx.x * KW_WARPS + wib;
  const bool has_unit = unit0 < N * nch;
  const int unit = has_unit ? unit0 : 0;
  const int n = unit / nch, cc = unit % nch;
  const int t0 = (nch == 1) ? 0 : cc * chunk_len;
  int t1 = (nch == 1) ? Tt : min(Tt, t0 + chunk_len);
  int tb = 0;
  if (cc > 0) tb = max(0, min(t0, ws.tv[n]) - burn_in);
  if (!has_unit) t1 = tb;
  const int c = lane;
  const bool col = c < NN;
  const int ca = min(max(c - OFF, 0), DD - 1);  // clamped index into the D-blocks
  for (int e = lane; e < Wt::SM_TOTAL; e += 32) S[e] = 0.0;  // zero padding of the 32 x 32 operands
  __syncwarp();

  double Lp[NN];
#pragma unroll
  for (int i = 0; i < NN; i++) Lp[i] = (i == c) ? 0.1 : 0.0;  // gibbs.py:18: S0 = 10 I
  double eta_p = 0.0;
  double llacc = (cc == 0) ? 0.5 * (double)NN * log(0.1) : 0.0;
  const float* mk = mask + (size_t)n * T + (Lg - 1);
  const int* zn = z + (size_t)n * Tt;
  const double* obs_n = ws.obs + (size_t)n * Tp * Y::OBS_SIZE;
  double* fw_n = ws.fw + (size_t)n * Tp * Y::FW_SIZE;
  int cur_z = -1;

  // observation information of the next step, prefetched one step ahead
  double Jn[DD], en, ctn;
  auto load_obs = [&](int tau) {
    const double* o = obs_n + (size_t)tau * Y::OBS_SIZE;
#pragma unroll
    for (int a = 0; a < DD; a++) Jn[a] = (c >= OFF && col && a >= c - OFF) ? o[Y::OBS_J + a * (a + 1) / 2 + (c - OFF)] : 0.0;
    en = (c >= OFF && col) ? o[Y::OBS_E + c - OFF] : 0.0;
    ctn = o[Y::OBS_C];
  };
  load_obs(tb);
  float mk_next = mk[tb];
  int z_next = (tb < Tt) ? zn[tb] : 0;

  // every warp of the CTA takes the same number of trips (the longest chain's)
  __shared__ int maxit_s;
  if (threadIdx.x == 0) maxit_s = 0;
  __syncthreads();
  if (lane == 0) atomicMax(&maxit_s, t1 - tb);
  __syncthreads();
  const int maxit = maxit_s;
  for (int it = 0; it < maxit; it++) {
    __syncthreads();   // keeps the warps that share a scheduler on the same code (instruction cache)
    const int tau = tb + it;
    if (tau >= t1) continue;
    if (tau == t0 && cc > 0) {
      double* bd = ws.bndf + ((size_t)(n * nch + cc) * 2 + 0) * Y::BND;
#pragma unroll
      for (int i = 0; i < NN; i++)
        if (c <= i) bd[i * (i + 1) / 2 + c] = Lp[i];
      if (col) bd[NTRI + c] = eta_p;
    }
    const bool live = mk_next > 0.0f;
    const int zt = z_next;
    mk_next = mk[tau + 1];                 // tau + 1 <= Tt: in range (mask row has T - L + 1 usable entries)
    z_next = (tau + 1 < Tt) ? zn[tau + 1] : 0;
    double Jc[DD];
#pragma unroll
    for (int a = 0; a < DD; a++) Jc[a] = Jn[a];
    const double ec = en, ct = ctn;
    load_obs(tau + 1);  // tau + 1 <= Tp - 1
    if (!live) continue;  // kalman.py:109-111
    const bool store = tau >= t0;
    if (zt != cur_z) {
      __syncwarp();
      const double* src = ws.tabw + (size_t)zt * Wt::SIZE;
      for (int e = Wt::FD + lane; e < Wt::SIZE; e += 32) Tb[e] = src[e];
      Gg = src;
      cur_z = zt;
      __syncwarp();
    }
    // ---- S1: column c of Lam_s = Lam_p + A^T Qinv A + C^T R^-1 C
    double Wc[NN];
#pragma unroll
    for (int i = 0; i < NN; i++) Wc[i] = Lp[i] + ((i >= c) ? __ldg(Gg + Wt::G + i * (i + 1) / 2 + c) : 0.0);
#pragma unroll
    for (int a = 0; a < DD; a++) Wc[OFF + a] += Jc[a];
    const double eta_c = eta_p + ec;
    // ---- S2: bottom-up Cholesky, rows of M published to Ms
    double pd = 1.0;
    double piv = __shfl_sync(FULL, Wc[NN - 1], NN - 1);
#pragma unroll
    for (int k = NN - 1; k >= 0; k--) {
      const double rs = rsqrt_pos(piv);
      const double m = Wc[k] * rs;
      if (k > 0) {
        // next pivot W[k-1][k-1] - M[k][k-1]^2 is known to lane k-1 from its own registers:
        // broadcast it now, so the shared-memory round trip below is off the pivot chain
        piv = __shfl_sync(FULL, fma(-m, m, Wc[k - 1]), k - 1);
      }
      if (c <= k) Ms[Wt::moff(k) + c] = m;
      if (c == 0) dv[k] = rs;
      pd *= rs;
      __syncwarp();
#pragma unroll
      for (int i = 0; i + 1 < k; i += 2) {
        const double2 mm = *reinterpret_cast<const double2*>(Ms + Wt::moff(k) + i);
        Wc[i] = fma(-mm.x, m, Wc[i]);
        Wc[i + 1] = fma(-mm.y, m, Wc[i + 1]);
      }
      if (k & 1) Wc[k - 1] = fma(-Ms[Wt::moff(k) + k - 1], m, Wc[k - 1]);
    }
    // ---- S3: L = M^-1, column c (zero above the diagonal by construction)
    double Lc[NN];
#pragma unroll
    for (int i = 0; i < NN; i++) {
      double a0 = (i == c) ? 1.0 : 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
      for (int q = 0; q + 1 < i; q += 2) {
        const double2 mm = *reinterpret_cast<const double2*>(Ms + Wt::moff(i) + q);
        if (q & 2) {
          a2 = fma(-mm.x, Lc[q], a2);
          a3 = fma(-mm.y, Lc[q + 1], a3);
        } else {
          a0 = fma(-mm.x, Lc[q], a0);
          a1 = fma(-mm.y, Lc[q + 1], a1);
        }
      }
      if (i & 1) a0 = fma(-Ms[Wt::moff(i) + i - 1], Lc[i - 1], a0);
      Lc[i] = ((a0 + a1) + (a2 + a3)) * dv[i];
    }
    // rows of L to shared memory, eta_c - f for the mean column, L and eta_c to the workspace
#pragma unroll
    for (int i = 0; i < NN; i++)
      if (c <= i) Lk[i * (i + 1) / 2 + c] = Lc[i];
    if (col) hv[c] = eta_c - Tb[Wt::F + c];
    if (store) {
      double* fwp = fw_n + (size_t)tau * Y::FW_SIZE;
#pragma unroll
      for (int i = 0; i < NN; i++)
        if (c <= i) fwp[Y::FW_L + i * (i + 1) / 2 + c] = Lc[i];
      if (col) fwp[Y::FW_ETA + c] = eta_c;
    }
    __syncwarp();
    // ---- T = L^T [F | eta_c - f] and Lam_p' = Qinv_ - T^T T on the FP64 tensor cores
    // (mma.sync m8n8k4.f64): operands are register fragments fetched once from the row buffers,
    // so the n^3-type work no longer pays one broadcast load per FMA.
    //   "K-fragment" of a row-major X: lane (g, qd) holds X[4*kq + qd][8*b + g]; it is both the
    //   A operand (X^T tile) and the B operand (X tile) of a product X^T Y.
    {
      constexpr int NB = (NN + 1 + 7) / 8;   // 8-column blocks of T (NN columns of V and u)
      constexpr int KQ = (NN + 3) / 4;       // 4-row chunks of the contraction
      constexpr int RB = (NN + 7) / 8;       // 8-row blocks of T
      constexpr int CB0 = OFF / 8;           // first column block that holds a dense column
      const int g = lane >> 2, qd = lane & 3;
      // Column c < OFF of T is 100 x row DD + c of L, so block b < CB0 of T is zero below row
      // DD + 8 b + 7: its last contributing 4-row chunk.
      auto tkmax = [](int b) constexpr { return (b < CB0) ? ((DD + 8 * b + 7) / 4 < KQ - 1 ? (DD + 8 * b + 7) / 4 : KQ - 1) : KQ - 1; };
      // dense column blocks of T by DMMA: T[:, 8cb..8cb+7] = L^T F[:, 8cb..]
      double Tacc[RB][NB - CB0][2];
#pragma unroll
      for (int rb = 0; rb < RB; rb++)
#pragma unroll
        for (int cb = 0; cb < NB - CB0; cb++) Tacc[rb][cb][0] = Tacc[rb][cb][1] = 0.0;
#pragma unroll
      for (int kq = 0; kq < KQ; kq++) {
        const int q = 4 * kq + qd;
        double bF[NB - CB0];
#pragma unroll
        for (int cb = 0; cb < NB - CB0; cb++) {
          const int cf = 8 * (cb + CB0) + g;   // column of F held by this lane
          const int qq = min(q, NN - 1);
          const double fd = Tb[Wt::FD + qq * DD + min(max(cf - OFF, 0), DD - 1)];
          const double hq = hv[qq];
          double v = (cf < OFF) ? ((q == DD + cf) ? 100.0 : 0.0) : ((cf < NN) ? fd : ((cf == NN) ? hq : 0.0));
          v = (q < NN) ? v : 0.0;
          bF[cb] = v;
        }
#pragma unroll
        for (int rb = 0; rb < RB; rb++) {
          if (kq < 2 * rb) continue;  // L is lower triangular: L[q][r] = 0 for q < r (8 rb <= r)
          const int rr = 8 * rb + g;
          const double av = (q < NN && rr <= q) ? Lk[q * (q + 1) / 2 + rr] : 0.0;   // L[q][r], lower triangular
#pragma unroll
          for (int cb = 0; cb < NB - CB0; cb++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                         : "+d"(Tacc[rb][cb][0]), "+d"(Tacc[rb][cb][1])
                         : "d"(av), "d"(bF[cb]));
        }
      }
      // dense blocks of T to the row buffer (accumulator layout: row 8rb+g, columns 8cb+2qd, +1)
#pragma unroll
      for (int rb = 0; rb < RB; rb++)
#pragma unroll
        for (int cb = 0; cb < NB - CB0; cb++)
          *reinterpret_cast<double2*>(Ms + (8 * rb + g) * Wt::TS + 8 * cb + 2 * qd) =
              make_double2(Tacc[rb][cb][0], Tacc[rb][cb][1]);
      __syncwarp();
      // K-fragments of T: unit-column blocks straight from L (T[r][c] = 100 L[D+c][r]), the rest
      // from the buffer just written
      double Tk[KQ][NB];
#pragma unroll
      for (int kq = 0; kq < KQ; kq++) {
        const int r = 4 * kq + qd;
#pragma unroll
        for (int b = 0; b < NB; b++) {
          const int cf = 8 * b + g;
          if (b < CB0) Tk[kq][b] = (kq <= tkmax(b) && r <= DD + cf) ? 100.0 * Lk[(DD + cf) * (DD + cf + 1) / 2 + r] : 0.0;
          else Tk[kq][b] = Ms[r * Wt::TS + (cf - 8 * CB0)];
        }
      }
      __syncwarp();
      // Z = T^T T, lower block triangle; row NN of Z is T^T u, Z[NN][NN] = u^T u
      double Zacc[NB * (NB + 1) / 2][2];
#pragma unroll
      for (int e = 0; e < NB * (NB + 1) / 2; e++) Zacc[e][0] = Zacc[e][1] = 0.0;
#pragma unroll
      for (int kq = 0; kq < KQ; kq++) {
#pragma unroll
        for (int ib = 0; ib < NB; ib++)
#pragma unroll
          for (int jb = 0; jb <= ib; jb++) {
            if (kq > tkmax(jb) || kq > tkmax(ib)) continue;  // structural zeros of the unit-column blocks
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                         : "+d"(Zacc[ib * (ib + 1) / 2 + jb][0]), "+d"(Zacc[ib * (ib + 1) / 2 + jb][1])
                         : "d"(Tk[kq][ib]), "d"(Tk[kq][jb]));
          }
      }
#pragma unroll
      for (int ib = 0; ib < NB; ib++)
#pragma unroll
        for (int jb = 0; jb <= ib; jb++)
          if (8 * ib + g <= NN)
            *reinterpret_cast<double2*>(Zs + (8 * ib + g) * RS + 8 * jb + 2 * qd) =
              make_double2(Zacc[ib * (ib + 1) / 2 + jb][0], Zacc[ib * (ib + 1) / 2 + jb][1]);
      __syncwarp();
    }
    // back to one column per lane: Lam_p'[i][c] = Qinv_[i][c] - Z[i][c] (i >= c), eta_p' = qb_ + Z[NN][c]
#pragma unroll
    for (int i = 0; i < NN; i++) {
      double qv;
      if (c < OFF) qv = (i == c) ? 100.0 : 0.0;
      else qv = (i >= OFF) ? Tb[Wt::QI + (i - OFF) * DD + ca] : 0.0;
      Lp[i] = qv - ((i >= c && col) ? Zs[i * RS + c] : 0.0);
    }
    eta_p = ((c >= OFF) ? Tb[Wt::QB + ca] : 0.0) + (col ? Zs[NN * RS + c] : 0.0);
    const double uu_z = Zs[NN * RS + NN];
    __syncwarp();
    if (want_ll && store) llacc += ct - 0.5 * Tb[Wt::SC + 1] + log(pd) + 0.5 * (uu_z - Tb[Wt::SC + 0]);
    __syncwarp();
  }
  if (!has_unit) return;
  if (nch > 1) {
    double* bd = ws.bndf + ((size_t)(n * nch + cc) * 2 + 1) * Y::BND;
#pragma unroll
    for (int i = 0; i < NN; i++)
      if (c <= i) bd[i * (i + 1) / 2 + c] = Lp[i];
    if (col) bd[NTRI + c] = eta_p;
  }
  if (cc == nch - 1) {
    // unconditional final update (kalman.py:117) and the draw of the last state (kalman.py:150)
    double Wc[NN];
#pragma unroll
    for (int i = 0; i < NN; i++) Wc[i] = Lp[i];
#pragma unroll
    for (int a = 0; a < DD; a++) Wc[OFF + a] += Jn[a];
    const double eta_c = eta_p + en;
    double pd = 1.0;
    __syncwarp();
#pragma unroll
    for (int k = NN - 1; k >= 0; k--) {
      const double piv = __shfl_sync(FULL, Wc[k], k);
      const double rs = rsqrt_pos(piv);
      const double m = Wc[k] * rs;
      if (c <= k) Zs[k * RS + c] = m;
      if (c == 0) dv[k] = rs;
      pd *= rs;
      __syncwarp();
#pragma unroll
      for (int i = 0; i < k; i++) Wc[i] = fma(-Zs[k * RS + i], m, Wc[i]);
    }
    double Lc[NN];
#pragma unroll
    for (int i = 0; i < NN; i++) {
      double a0 = (i == c) ? 1.0 : 0.0;
#pragma unroll
      for (int q = 0; q < i; q++) a0 = fma(-Zs[i * RS + q], Lc[q], a0);
      Lc[i] = a0 * dv[i];
    }
    if (col) hv[c] = eta_c;
    __syncwarp();
    // t_c = sum_q L[q][c] eta_c[q];  y = t + w_last;  x_i = sum_c L[i][c] y_c
    double tc = 0.0;
#pragma unroll
    for (int q = 0; q < NN; q++) tc = fma(Lc[q], hv[q], tc);
    if (!col) tc = 0.0;
    double tt = tc * tc;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tt += __shfl_xor_sync(FULL, tt, o);
    const double yc = col ? tc + (double)w_last[(size_t)n * NN + c] : 0.0;
    double* xl = fw_n + (size_t)(Tp - 1) * Y::FW_SIZE;
#pragma unroll
    for (int i = 0; i < NN; i++) {
      double v = Lc[i] * yc;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
      if (lane == 0) xl[i] = v;
    }
    llacc += ctn + log(pd) + 0.5 * tt;
  }
  if (lane == NN) ws.llpart[(size_t)n * nch + cc] = llacc;
}

#include "kalman_blk.cuh"

// ------------------------------------------------------------------ backward sampler
// One warp per session (or session chunk).  x_t = L (L^T h + w_t),
// h = eta_c + A^T Qinv (x_{t+1} - B) with A^T Qinv = [100 E | F_D] (see file header).
// The packed filter record of a step (lower-triangular L row by row, then eta_c) is copied
// global -> shared with cp.async, one step ahead, and read IN PLACE: L[q][c] sits at
// q(q+1)/2 + c, so "column c, all rows" is contiguous across lanes, and "row c" starts at the
// triangular number c(c+1)/2, which hits sixteen distinct 8-byte banks for sixteen consecutive c.
template <int NN, int DD>
__global__ void __launch_bounds__(32)
    kalman_backward_kernel(const float* __restrict__ mask, const int* __restrict__ z,
                           const float* __restrict__ w, KalWs ws, int N, int T, int Lg,
                           int chunk_len, int burn_in, float* __restrict__ x_out) {
  using Y = KL<NN, DD>;
  constexpr int NTRI = Y::NTRI, OFF = NN - DD, FW = Y::FW_SIZE;
  constexpr int NCP = (FW * 8 + 16 * 32 - 1) / (16 * 32);   // 16-byte pieces per lane and record
  extern __shared__ __align__(16) double sm[];
  double* rec = sm;            // 2 x FW: packed records, double-buffered
  constexpr int NNP = (NN + 1) / 2 * 2;   // the vectors are read as broadcast double2: keep them 16-byte aligned
  double* hs = rec + 2 * FW;   // NNP
  double* ys = hs + NNP;       // NNP
  double* xs = ys + NNP;       // NNP
  double* FDs = xs + NNP;      // NN*DD
  double* bs = FDs + NN * DD;  // DD
  const int lane = threadIdx.x;
  const int Tp = T - Lg + 1, Tt = Tp - 1;
  const int nch = kal_nchunks(Tt, chunk_len);
  const int n = blockIdx.x / nch, c = blockIdx.x % nch;
  const int t0 = (nch == 1) ? 0 : c * chunk_len;
  const int t1 = (nch == 1) ? Tt : min(Tt, t0 + chunk_len);
  const float* mk = mask + (size_t)n * T + (Lg - 1);
  const int* zn = z + (size_t)n * Tt;
  const double* fw_n = ws.fw + (size_t)n * Tp * FW;
  const float* w_n = w + (size_t)n * Tt * NN;
  float* xo = x_out + (size_t)n * T * DD;
  const double* xlast = fw_n + (size_t)(Tp - 1) * FW;
  int ts = t1;
  bool exact_start = true;
  if (c < nch - 1) {
    ts = min(Tt, t1 + burn_in);
    exact_start = (ts == Tt) || !(mk[ts] > 0.0f);  // trailing padding: x is still x_last there
  }
  if (lane < NN) xs[lane] = exact_start ? xlast[lane] : 0.0;
  if (c == nch - 1 && lane < DD) xo[(size_t)(T - 1) * DD + lane] = (float)xlast[OFF + lane];
  const unsigned rec_sa = (unsigned)__cvta_generic_to_shared(rec);
  auto fetch = [&](int tau, int buf) {   // record of step tau -> buffer buf (asynchronous)
    if (tau >= t0) {
      const char* src = reinterpret_cast<const char*>(fw_n + (size_t)tau * FW);
#pragma unroll
      for (int i = 0; i < NCP; i++) {
        const int piece = lane + 32 * i;
        if (piece * 16 < FW * 8)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(rec_sa + (unsigned)(buf * FW * 8 + piece * 16)),
                       "l"(src + piece * 16));
      }
    }
    asm volatile("cp.async.commit_group;\n" ::);
  };
  fetch(ts - 1, 0);
  __syncwarp();
  int cur_z = -1;
  float wpre = (ts - 1 >= t0 && lane < NN) ? w_n[(size_t)(ts - 1) * NN + lane] : 0.0f;
  // mask and state of a step are loaded one step ahead too: a global load at the top of a step would sit
  // on the serial chain of this one-warp recursion
  float mpre = (ts - 1 >= t0) ? mk[ts - 1] : 0.0f;
  int zpre = (ts - 1 >= t0) ? zn[ts - 1] : 0;
  const int rowbase = lane * (lane + 1) / 2;
  int buf = 0;
  for (int tau = ts - 1; tau >= t0; tau--, buf ^= 1) {
    if (tau == t1 - 1 && c < nch - 1 && lane < NN)
      ws.bndx[((size_t)(n * nch + c) * 2 + 0) * NN + lane] = xs[lane];
    const bool live = mpre > 0.0f;
    const int zt = zpre;
    fetch(tau - 1, buf ^ 1);          // next step's record, one step ahead
    const float wcur = wpre;
    if (tau - 1 >= t0) {
      if (lane < NN) wpre = w_n[(size_t)(tau - 1) * NN + lane];
      mpre = mk[tau - 1];
      zpre = zn[tau - 1];
    }
    asm volatile("cp.async.wait_group 1;\n" ::);   // this step's record has landed
    __syncwarp();
    if (live) {
      const double* Pk = rec + buf * FW;
      if (zt != cur_z) {
        const double* tb_ = ws.tab + (size_t)zt * Y::TAB_SIZE;
        for (int e = lane; e < NN * DD; e += 32) FDs[e] = tb_[Y::TAB_FD + e];
        if (lane < DD) bs[lane] = tb_[Y::TAB_B + lane];
        cur_z = zt;
      }
      __syncwarp();
      if (lane < NN) {
        double hh = Pk[NTRI + lane] + ((lane >= DD) ? 100.0 * xs[lane - DD] : 0.0), h2 = 0.0;
#pragma unroll
        for (int a = 0; a + 1 < DD; a += 2) {
          hh = fma(FDs[lane * DD + a], xs[OFF + a] - bs[a], hh);
          h2 = fma(FDs[lane * DD + a + 1], xs[OFF + a + 1] - bs[a + 1], h2);
        }
        if (DD & 1) hh = fma(FDs[lane * DD + DD - 1], xs[OFF + DD - 1] - bs[DD - 1], hh);
        hs[lane] = hh + h2;
      }
      __syncwarp();
      if (lane < NN) {   // four partial sums: the dependent-FMA chain is the latency of this step
        double ta[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int q = 0; q < NN; q += 2) {
          const double2 h2 = *reinterpret_cast<const double2*>(hs + q);   // one broadcast load for two rows
          if (q >= lane) ta[q & 3] = fma(Pk[q * (q + 1) / 2 + lane], h2.x, ta[q & 3]);
          if (q + 1 < NN && q + 1 >= lane) ta[(q + 1) & 3] = fma(Pk[(q + 1) * (q + 2) / 2 + lane], h2.y, ta[(q + 1) & 3]);
        }
        ys[lane] = ((ta[0] + ta[1]) + (ta[2] + ta[3])) + (double)wcur;
      }
      __syncwarp();
      double xn = 0.0;
      if (lane < NN) {
        double xa[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int q = 0; q < NN; q += 2) {
          const double2 y2 = *reinterpret_cast<const double2*>(ys + q);
          if (q <= lane) xa[q & 3] = fma(Pk[rowbase + q], y2.x, xa[q & 3]);
          if (q + 1 < NN && q + 1 <= lane) xa[(q + 1) & 3] = fma(Pk[rowbase + q + 1], y2.y, xa[(q + 1) & 3]);
        }
        xn = (xa[0] + xa[1]) + (xa[2] + xa[3]);
      }
      __syncwarp();
      if (lane < NN) xs[lane] = xn;
      __syncwarp();
    }
    if (tau < t1) {
      // gibbs.py:22: x[t] = newest block of LDS state t-L+1; x[0..L-2] = older blocks of state 0
      if (lane < DD) xo[(size_t)(tau + Lg - 1) * DD + lane] = live ? (float)xs[OFF + lane] : 0.0f;
      if (tau == 0 && lane < OFF) xo[lane] = live ? (float)xs[lane] : 0.0f;
    }
  }
  asm volatile("cp.async.wait_group 0;\n" ::);
  if (nch > 1 && c > 0 && lane < NN) ws.bndx[((size_t)(n * nch + c) * 2 + 1) * NN + lane] = xs[lane];
}

// loglik[n] = sum of chunk partials (fixed order); boundary mismatch of chunked mode.
__global__ void kalman_finish_kernel(KalWs ws, int nch, int BND, int NN, double* __restrict__ loglik,
                                     unsigned long long* __restrict__ err_bits) {
  const int n = blockIdx.x, tid = threadIdx.x;
  if (tid == 0 && loglik) {
    double s = 0.0;
    for (int c = 0; c < nch; c++) s += ws.llpart[(size_t)n * nch + c];
    loglik[n] = s;
  }
  if (!err_bits || nch < 2) return;
  __shared__ double red[2][32];
  for (int which = 0; which < 2; which++) {
    double worst = 0.0;
    for (int c = 0; c + 1 < nch; c++) {
      const double *a, *b;
      int len;
      if (which == 0) {  // filter: end carry of chunk c vs warmed-up start of chunk c+1
        a = ws.bndf + ((size_t)(n * nch + c) * 2 + 1) * BND;
        b = ws.bndf + ((size_t)(n * nch + c + 1) * 2 + 0) * BND;
        len = BND;
      } else {           // sampler: warmed-up x of chunk c vs x at the start of chunk c+1
        a = ws.bndx + ((size_t)(n * nch + c + 1) * 2 + 1) * NN;
        b = ws.bndx + ((size_t)(n * nch + c) * 2 + 0) * NN;
        len = NN;
      }
      double md = 0.0, ma = 0.0;
      for (int e = tid; e < len; e += blockDim.x) {
        const double d = fabs(a[e] - b[e]);
        md = (d != d) ? INFINITY : fmax(md, d);  // a NaN boundary must fail the check, not vanish in fmax
        ma = fmax(ma, fabs(a[e]));
      }
      for (int o = 16; o > 0; o >>= 1) {
        md = fmax(md, __shfl_xor_sync(0xffffffffu, md, o));
        ma = fmax(ma, __shfl_xor_sync(0xffffffffu, ma, o));
      }
      __syncthreads();
      if ((tid & 31) == 0) {
        red[0][tid >> 5] = md;
        red[1][tid >> 5] = ma;
      }
      __syncthreads();
      if (tid == 0) {
        double d = 0.0, m = 0.0;
        for (int q = 0; q < (int)(blockDim.x >> 5); q++) {
          d = fmax(d, red[0][q]);
          m = fmax(m, red[1][q]);
        }
        worst = fmax(worst, d / (m + 1e-300));
      }
    }
    if (tid == 0) atomicMax(&err_bits[which], (unsigned long long)__double_as_longlong(worst));
  }
}

template <int NN, int DD>
static int launch_latents(const float* Y, const float* mask, const float* v, const float* h,
                          const int* z, const float* s, const double* Cdfull, const double* sigmasq,
                          const double* Ab, const double* Q, const float* w, const float* w_last,
                          int N, int T, int k, int dk, int Lg, int K, int chunk_len, int burn_in,
                          float* x, double* loglik, double* boundary_err, KalWs ws,
                          cudaStream_t st) {
  using YL = KL<NN, DD>;
  const int Tp = T - Lg + 1, Tt = Tp - 1;
  const int nch = kal_nchunks(Tt, chunk_len);
  kalman_tables_kernel<NN, DD><<<K, 32, 0, st>>>(Ab, Q, ws.tab, ws.tabw);
  int rc = kpms_last_launch();
  if (rc) return rc;
  {
    const long long tot = (long long)N * Tp;
    if (dk == 2 && k % 4 == 0 && k <= 64 && kpms_opts.kalman_obs_mma && kpms_aligned(Y, 8) && kpms_aligned(v, 8)) {
      // tensor-core records: a warp per 32 frames, a few waves of CTAs striding over the frames
      constexpr int NJP = 8 * ((YL::NJ + 7) / 8) + 4, CSP = 8 * ((DD + 7) / 8) + 2;
      const size_t smem = ((size_t)k * NJP + (size_t)2 * k * CSP + 3 * (size_t)k) * sizeof(double);
      auto kern = kalman_obs_mma_kernel<NN, DD>;
      KPMS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      const long long groups = (tot + 31) / 32;
      const long long want = (groups + 7) / 8;
      const unsigned grid = (unsigned)(want < 148LL * 8 ? (want > 0 ? want : 1) : 148LL * 8);
      kern<<<grid, 256, smem, st>>>(Y, v, h, s, Cdfull, sigmasq, N, T, k, Lg, loglik != nullptr, ws.obs);
    } else {
      const size_t smem = ((size_t)k * dk * (DD + 1) + (size_t)k * YL::NJ + k + 4 * 32 * (size_t)(YL::OBS_SIZE | 1)) * sizeof(double);
      auto kern = kalman_obs_kernel<NN, DD>;
      KPMS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      kern<<<(unsigned)((tot + 127) / 128), 128, smem, st>>>(Y, v, h, s, Cdfull, sigmasq, N, T, k, dk,
                                                            Lg, ws.obs);
    }
    rc = kpms_last_launch();
    if (rc) return rc;
  }
  if (nch > 1) {
    kalman_tv_kernel<<<N, 256, 0, st>>>(mask, T, Lg, Tt, ws.tv);
    rc = kpms_last_launch();
    if (rc) return rc;
  }
  {
    const size_t smem = (size_t)(2 * NN * YL::NS + NN * YL::NV + 3 * NN + NN * NN + NN * DD + DD * DD +
                                 NN + DD + 2 + YL::OBS_SIZE) * sizeof(double) + 2 * YL::NTRI;
    if (kpms_opts.kalman_impl >= 2 && NN <= 31) {
      // blocked DMMA kernel (default) + the final update / last-state draw as a kernel of its own
      using KBt = KB<NN, DD>;
      kalman_tables_blk_kernel<NN, DD><<<K, 32, 0, st>>>(ws.tab, ws.tabg, ws.tabf);
      rc = kpms_last_launch();
      if (rc) return rc;
      const int wpb = kpms_opts.kalman_warps;
      const int lockstep = kpms_opts.kalman_lockstep;
#define KPMS_BLK_GO(W)                                                                                              \
  {                                                                                                                 \
    const size_t smb = (size_t)(W) * KBt::SM_TOTAL * sizeof(double);                                                \
    auto kern = kalman_forward_blk_kernel<NN, DD, W>;                                                               \
    KPMS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smb));                   \
    kern<<<(N * nch + (W) - 1) / (W), (W) * 32, smb, st>>>(mask, z, ws, N, T, Lg, chunk_len, burn_in,               \
                                                          loglik != nullptr, lockstep);                             \
  }
      if (wpb == 8) KPMS_BLK_GO(8) else if (wpb == 16) KPMS_BLK_GO(16) else KPMS_BLK_GO(12)
#undef KPMS_BLK_GO
      rc = kpms_last_launch();
      if (rc) return rc;
      kalman_final_kernel<NN, DD><<<N, 32, 0, st>>>(w_last, ws, N, T, Lg, nch);
    } else if (kpms_opts.kalman_impl == 0 || NN > 30) {  // CTA-per-chunk shared-memory kernel (cross-checks; n > 30)
      auto kern = kalman_forward_kernel<NN, DD>;
      KPMS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      kern<<<N * nch, KAL_NT, smem, st>>>(mask, z, w_last, ws, N, T, Lg, chunk_len, burn_in,
                                          loglik != nullptr);
    } else {
      if constexpr (NN <= 30) {
        const size_t smw = (size_t)KW_WARPS * KW<NN, DD>::SM_TOTAL * sizeof(double);
        auto kern = kalman_forward_warp_kernel<NN, DD>;
        KPMS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smw));
        kern<<<(N * nch + KW_WARPS - 1) / KW_WARPS, KW_WARPS * 32, smw, st>>>(
            mask, z, w_last, ws, N, T, Lg, chunk_len, burn_in, loglik != nullptr);
      }
    }
    rc = kpms_last_launch();
    if (rc) return rc;
  }
  {
    const size_t smem = (size_t)(2 * YL::FW_SIZE + 3 * ((NN + 1) / 2 * 2) + NN * DD + DD) * sizeof(double);
    auto kern = kalman_backward_kernel<NN, DD>;
    KPMS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<N * nch, 32, smem, st>>>(mask, z, w, ws, N, T, Lg, chunk_len, burn_in, x);
    rc = kpms_last_launch();
    if (rc) return rc;
  }
  if (boundary_err) KPMS_CUDA(cudaMemsetAsync(boundary_err, 0, 2 * sizeof(double), st));
  if (loglik || (boundary_err && nch > 1)) {
    kalman_finish_kernel<<<N, 128, 0, st>>>(ws, nch, YL::BND, NN, loglik,
                                            (unsigned long long*)boundary_err);
    rc = kpms_last_launch();
  }
  return rc;
}

extern "C" int kpms_resample_latents(const float* Y, const float* mask, const float* v,
                                     const float* h, const int32_t* z, const float* s,
                                     const double* Cdfull, const double* sigmasq, const double* Ab,
                                     const double* Q, const float* w, const float* w_last, int N,
                                     int T, int k, int dk, int D, int L, int K, int chunk_len,
                                     int burn_in, float* x, double* loglik, double* boundary_err,
                                     void* workspace, size_t workspace_bytes, void* stream) {
  KPMS_REQUIRE(Y && mask && v && h && z && s && Cdfull && sigmasq && Ab && Q && w && w_last && x &&
                   workspace,
               KPMS_ERR_NULL);
  KPMS_REQUIRE(N > 0 && T > L && k > 0 && D > 0 && L > 0 && K > 0, KPMS_ERR_SHAPE);
  KPMS_REQUIRE(dk >= 2 && dk <= 3 && kpms_dl_supported(D, L) && k <= 64, KPMS_ERR_UNSUPPORTED);
  KPMS_REQUIRE(chunk_len <= 0 || burn_in >= 0, KPMS_ERR_SHAPE);
  KalWs ws;
  const size_t need = kal_layout(N, T, D, L, K, chunk_len, workspace, &ws);
  KPMS_REQUIRE(kpms_aligned(workspace, 256) && workspace_bytes >= need, KPMS_ERR_WORKSPACE);
  cudaStream_t st = (cudaStream_t)stream;
#define KPMS_X(d, l)                                                                              \
  if (D == d && L == l)                                                                           \
    return launch_latents<d * l, d>(Y, mask, v, h, z, s, Cdfull, sigmasq, Ab, Q, w, w_last, N, T, \
                                    k, dk, L, K, chunk_len, burn_in, x, loglik, boundary_err, ws, \
                                    st);
  KPMS_FOR_EACH_DL(KPMS_X)
#undef KPMS_X
  return KPMS_ERR_UNSUPPORTED;
}
