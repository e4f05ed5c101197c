// Blocked (8 x 8) Kalman forward filter: one warp per session chunk, all O(n^3) work as FP64
// tensor-core products (mma.sync.m8n8k4.f64) on register fragments.  Included by kalman.cu.
//
// Same information-form recursion as kalman_forward_warp_kernel (see the header of kalman.cu), other
// data distribution.  The n x n carry Lam_p, augmented by eta_p as row n, lives in the accumulator
// layout of the MMA (lane = 4 g + qd holds rows 8 i + g, columns 8 j + 2 qd, + 1 of every lower
// block (i, j)), so it is produced by the last product of one step and consumed by the first sum of
// the next without leaving the registers.  Per live step:
//   S1  W = Lam_p + G_z + J  in place (G_z comes in the same layout from a per-state table)
//   A   bottom-up Cholesky W = M^T M by block rows k = NB-1 .. 0, left-looking: the block row is
//       accumulated in registers (acc_kj -= M_k'k^T M_k'j, k' > k), then factored as an 8-row strip
//       with scalar pivots, lane c owning column c (diagonal block and panel together; the next
//       pivot is broadcast by its owner lane before the multipliers pass through shared memory, so a
//       pivot costs one reciprocal, two FMAs and a shuffle on the critical path).  M goes to a
//       per-warp P x (P+4) buffer whose row stride makes all fragment access patterns conflict free.
//   B   the diagonal blocks of M are inverted all at once (lane = 8 b + c: column c of block b), then
//       L = M^-1 in place, L_ij = -(sum_{j<k<=i} L_ik M_kj) Li_jj, diagonal by diagonal; finished
//       blocks go straight from the accumulators to the filter record in global memory.
//   C   T = L^T [F | eta_c - f], Z = T^T T  (as before; Z's accumulators ARE the next carry:
//       Lam_p' = Qinv_ - Z, eta_p' = qb_ + Z[n][:]).
// 64 + ~105 DMMA per step; shared-memory traffic drops from ~1400 to ~600 wavefronts per step and
// the instruction count to less than half (tools/proto_kalman_frag.py emulates this kernel lane by
// lane in NumPy and checks it against a dense step for every (D, L) pair of dispatch.cuh).
// The unconditional final update and the draw of the last state run in kalman_final_kernel.
#pragma once

#define KB_WARPS_DEFAULT 12

template <int NN, int DD>
struct KB {
  static constexpr int NBK = (NN + 1 + 7) / 8;       // 8-blocks covering the augmented matrix (rows 0..NN)
  static constexpr int P = 8 * NBK;
  static constexpr int SA = P + 4;                     // row stride of the M / L buffer (4 mod 8)
  static constexpr int NBLK = NBK * (NBK + 1) / 2;
  static constexpr int OFF = NN - DD;
  static constexpr int CB0 = OFF / 8;                  // first column block of T with a dense column
  static constexpr int NDB = NBK - CB0;                // dense column blocks of T
  static constexpr int TS = 8 * NDB + 4;               // row stride of the T buffer (4 mod 8)
  static constexpr int KQ = 2 * NBK;                   // 4-row chunks of a contraction over P rows
  static constexpr int IB = NN / 8, GE = NN % 8;       // block row / row within the block of the eta row
  static constexpr int SM_BUF = 0;
  static constexpr int SM_TB = SM_BUF + P * SA;        // T dense blocks; scratch of the diagonal factorisation
  static constexpr int SM_TOTAL = (SM_TB + (P * TS > 64 ? P * TS : 64) + 1) & ~1;
  // eta_c - f and 1 / M_dd live in the four padding columns of the M / L buffer (element c at row c)
  static constexpr int HV_COL = P, DV_COL = P + 1;
  static constexpr int TABG = NBLK * 64;               // per-state G in accumulator layout
  static constexpr int TABF = KQ * NDB * 32;           // per-state [unit columns | F] in B-operand layout
  __host__ __device__ static constexpr int blk(int i, int j) { return i * (i + 1) / 2 + j; }
};

#define KB_DMMA(c, a, b)                                                                           \
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" \
               : "+d"((c)[0]), "+d"((c)[1])                                                        \
               : "d"(a), "d"(b))

// per-state tables in fragment order (from the dense tables of kalman_tables_kernel)
template <int NN, int DD>
__global__ void kalman_tables_blk_kernel(const double* __restrict__ tab, double* __restrict__ tabg,
                                         double* __restrict__ tabf) {
  using Y = KL<NN, DD>;
  using B = KB<NN, DD>;
  const int k = blockIdx.x, lane = threadIdx.x;
  const int g = lane >> 2, qd = lane & 3;
  const double* t = tab + (size_t)k * Y::TAB_SIZE;
  double* tg = tabg + (size_t)k * B::TABG;
  double* tf = tabf + (size_t)k * B::TABF;
  for (int i = 0; i < B::NBK; i++)
    for (int j = 0; j <= i; j++)
      for (int e = 0; e < 2; e++) {
        const int row = 8 * i + g, col = 8 * j + 2 * qd + e;
        double v = 0.0;
        if (row < NN && col < NN) v = t[Y::TAB_G + row * NN + col];
        else if (row == col) v = 1.0;  // identity padding keeps W positive definite
        tg[(B::blk(i, j) * 32 + lane) * 2 + e] = v;
      }
  for (int kq = 0; kq < B::KQ; kq++)
    for (int cb = 0; cb < B::NDB; cb++) {
      const int q = 4 * kq + qd, cf = 8 * (cb + B::CB0) + g;
      double v = 0.0;
      if (q < NN) {
        if (cf < B::OFF) v = (q == DD + cf) ? 100.0 : 0.0;
        else if (cf < NN) v = t[Y::TAB_FD + q * DD + (cf - B::OFF)];
      }
      tf[(kq * B::NDB + cb) * 32 + lane] = v;
    }
}

__device__ __forceinline__ void kb_prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

template <int NN, int DD, int KB_WARPS>
__global__ void __launch_bounds__(KB_WARPS * 32, 1)
    kalman_forward_blk_kernel(const float* __restrict__ mask, const int* __restrict__ z, KalWs ws, int N, int T,
                              int Lg, int chunk_len, int burn_in, int want_ll, int lockstep) {
  using Y = KL<NN, DD>;
  using B = KB<NN, DD>;
  constexpr int NBK = B::NBK, SA = B::SA, OFF = B::OFF, NTRI = Y::NTRI, TS = B::TS, NDB = B::NDB, CB0 = B::CB0;
  constexpr unsigned FULL = 0xffffffffu;
  extern __shared__ __align__(16) double sm[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int g = lane >> 2, qd = lane & 3;
  double* S = sm + (size_t)wib * B::SM_TOTAL;
  double* buf = S + B::SM_BUF;   // M off-diagonal blocks / Li_kk, then L
  double* Tb = S + B::SM_TB;     // dense blocks of T; scratch of the diagonal factorisation
  double* hv = buf + B::HV_COL;  // eta_c - f: element c at hv[c * SA]
  double* dvs = buf + B::DV_COL; // 1 / M_dd: element c at dvs[c * SA]
  const int Tp = T - Lg + 1, Tt = Tp - 1;
  const int nch = kal_nchunks(Tt, chunk_len);
  const int unit0 = blockIdx.x * KB_WARPS + wib;
  const bool has_unit = unit0 < N * nch;
  const int unit = has_unit ? unit0 : 0;
  const int n = unit / nch, cc = unit % nch;
  const int t0 = (nch == 1) ? 0 : cc * chunk_len;
  int t1 = (nch == 1) ? Tt : min(Tt, t0 + chunk_len);
  int tb = 0;
  if (cc > 0) tb = max(0, min(t0, ws.tv[n]) - burn_in);
  if (!has_unit) t1 = tb;
  for (int e = lane; e < B::SM_TOTAL; e += 32) S[e] = 0.0;
  __syncwarp();

  // carry in accumulator layout: C[blk(i,j)][e] = element (8i+g, 8j+2qd+e); row NN holds eta_p
  double C[B::NBLK][2];
#pragma unroll
  for (int i = 0; i < NBK; i++)
#pragma unroll
    for (int j = 0; j <= i; j++)
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const int row = 8 * i + g, col = 8 * j + 2 * qd + e;
        C[B::blk(i, j)][e] = (row == col && row < NN) ? 0.1 : 0.0;  // gibbs.py:18: S0 = 10 I, mu0 = 0
      }
  double llacc = (cc == 0) ? 0.5 * (double)NN * log(0.1) : 0.0;
  const float* mk = mask + (size_t)n * T + (Lg - 1);
  const int* zn = z + (size_t)n * Tt;
  const double* obs_n = ws.obs + (size_t)n * Tp * Y::OBS_SIZE;
  double* fw_n = ws.fw + (size_t)n * Tp * Y::FW_SIZE;

  // observation information of the next step, prefetched one step ahead, already in the slots of the
  // accumulator elements it is added to
  double Jn[B::NBLK][2], En[NBK][2], ctn;
  auto load_obs = [&](int tau) {
    const double* o = obs_n + (size_t)tau * Y::OBS_SIZE;
#pragma unroll
    for (int i = 0; i < NBK; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) {
        if (8 * i + 7 < OFF || 8 * j + 7 < OFF) continue;   // compile-time: block outside the observed corner
#pragma unroll
        for (int e = 0; e < 2; e++) {
          const int row = 8 * i + g, col = 8 * j + 2 * qd + e;
          const bool in = row < NN && col <= row && col >= OFF && row >= OFF;
          const int a = in ? row - OFF : 0, b = in ? col - OFF : 0;
          Jn[B::blk(i, j)][e] = in ? o[Y::OBS_J + a * (a + 1) / 2 + b] : 0.0;
        }
      }
#pragma unroll
    for (int j = 0; j <= B::IB; j++) {
      if (8 * j + 7 < OFF) continue;
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const int col = 8 * j + 2 * qd + e;
        const bool in = g == B::GE && col >= OFF && col < NN;
        En[j][e] = in ? o[Y::OBS_E + (in ? col - OFF : 0)] : 0.0;
      }
    }
    ctn = o[Y::OBS_C];
  };
  load_obs(tb);
  float mk_next = mk[tb];
  int z_next = (tb < Tt) ? zn[tb] : 0;

  __shared__ int maxit_s;
  if (threadIdx.x == 0) maxit_s = 0;
  __syncthreads();
  if (lane == 0) atomicMax(&maxit_s, t1 - tb);
  __syncthreads();
  const int maxit = maxit_s;
  auto record_carry = [&](double* bd) {   // packed lower Lam_p, then eta_p (the layout of KL::BND)
#pragma unroll
    for (int i = 0; i < NBK; i++)
#pragma unroll
      for (int j = 0; j <= i; j++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
          const int row = 8 * i + g, col = 8 * j + 2 * qd + e;
          if (row < NN && col <= row) bd[row * (row + 1) / 2 + col] = C[B::blk(i, j)][e];
          if (row == NN && col < NN) bd[NTRI + col] = C[B::blk(i, j)][e];
        }
  };
  for (int it = 0; it < maxit; it++) {
    if (lockstep & 1) __syncthreads();   // keeps the warps that share a scheduler on the same code (instruction cache)
    const int tau = tb + it;
    if (tau >= t1) continue;
    if (tau == t0 && cc > 0) record_carry(ws.bndf + ((size_t)(n * nch + cc) * 2 + 0) * Y::BND);
    const bool live = mk_next > 0.0f;
    const int zt = z_next;
    mk_next = mk[tau + 1];                 // tau + 1 <= Tt: in range
    z_next = (tau + 1 < Tt) ? zn[tau + 1] : 0;
    double Jc[B::NBLK][2], Ec[NBK][2];
#pragma unroll
    for (int b = 0; b < B::NBLK; b++) { Jc[b][0] = Jn[b][0]; Jc[b][1] = Jn[b][1]; }
#pragma unroll
    for (int j = 0; j < NBK; j++) { Ec[j][0] = En[j][0]; Ec[j][1] = En[j][1]; }
    const double ct = ctn;
    load_obs(tau + 1);  // tau + 1 <= Tp - 1
    if (!live) continue;  // kalman.py:109-111
    const bool store = tau >= t0;
    const double* tabz = ws.tab + (size_t)zt * Y::TAB_SIZE;
    const double* tgz = ws.tabg + (size_t)zt * B::TABG;
    const double* tfz = ws.tabf + (size_t)zt * B::TABF;
    double* fwp = fw_n + (size_t)tau * Y::FW_SIZE;
    if (lockstep & 2) {   // bit 1: pull the tables phase C and the carry update read (F fragments, Qinv, qb) into L1 now
      if (lane * 16 < B::TABF) kb_prefetch_l1(tfz + lane * 16);
      if (lane * 16 < Y::TAB_SIZE - Y::TAB_QI + 16) kb_prefetch_l1(tabz + Y::TAB_QI + min(lane * 16, Y::TAB_SIZE - Y::TAB_QI - 1));
    }
    // ---- eta_c = eta_p + e (row NN of the carry), h = eta_c - f to shared memory, eta_c to the record
#pragma unroll
    for (int j = 0; j <= B::IB; j++)
#pragma unroll
      for (int e = 0; e < 2; e++) {
        const int col = 8 * j + 2 * qd + e;
        if (g == B::GE && col < NN) {
          double ec = C[B::blk(B::IB, j)][e];
          if (8 * j + 7 >= OFF) ec += Ec[j][e];
          hv[col * SA] = ec - __ldg(tabz + Y::TAB_F + col);
          if (store) fwp[Y::FW_ETA + col] = ec;
        }
      }
    // ---- S1 (in place): W = Lam_p + G_z + J on rows / columns < NN, the table's identity elsewhere
#pragma unroll
    for (int i = 0; i < NBK; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) {
        const double2 gz = __ldg(reinterpret_cast<const double2*>(tgz) + B::blk(i, j) * 32 + lane);
#pragma unroll
        for (int e = 0; e < 2; e++) {
          const int row = 8 * i + g, col = 8 * j + 2 * qd + e;
          double w = (row < NN && col < NN) ? C[B::blk(i, j)][e] : 0.0;
          w += e ? gz.y : gz.x;
          if (8 * i + 7 >= OFF && 8 * j + 7 >= OFF) w += Jc[B::blk(i, j)][e];
          C[B::blk(i, j)][e] = w;
        }
      }
    // ---- phase A: blocked bottom-up Cholesky, left-looking
    double pd = 1.0;
#pragma unroll
    for (int k = NBK - 1; k >= 0; k--) {
#pragma unroll
      for (int kp = k + 1; kp < NBK; kp++)
#pragma unroll
        for (int half = 0; half < 2; half++) {
          const double* rowp = buf + (8 * kp + 4 * half + qd) * SA + g;
          const double a = -rowp[8 * k];
#pragma unroll
          for (int j = 0; j <= k; j++) KB_DMMA(C[B::blk(k, j)], a, rowp[8 * j]);
        }
      // block row k by scalar pivots: lane c owns column c of the strip (rows 8k..8k+7); the diagonal
      // block and the panel are factored together, so no inverse of the diagonal block is needed here
#pragma unroll
      for (int j = 0; j <= k; j++)
        *reinterpret_cast<double2*>(buf + (8 * k + g) * SA + 8 * j + 2 * qd) = make_double2(C[B::blk(k, j)][0], C[B::blk(k, j)][1]);
      __syncwarp();
      {
        double R[8], Mr[8];
#pragma unroll
        for (int i = 0; i < 8; i++) R[i] = buf[(8 * k + i) * SA + lane];   // columns right of the strip read zeros
        double* scr = Tb;   // scr[d*8 + i] = W[d][8k+i] / pivot_d, the multipliers of pivot d
        double piv = __shfl_sync(FULL, R[7], 8 * k + 7);
#pragma unroll
        for (int d = 7; d >= 0; d--) {
          const double pcur = piv;
          const double t = R[d] * rcp_pos(pcur);
          // the next pivot W[d-1][d-1] - W[d][d-1]^2 / pivot_d is known to its owner lane now: broadcast it
          // before the multipliers make their round trip through shared memory
          if (d > 0) piv = __shfl_sync(FULL, fma(-t, R[d], R[d - 1]), 8 * k + d - 1);
          if (lane >= 8 * k && lane < 8 * k + d) scr[d * 8 + (lane - 8 * k)] = t;
          const double rs = rsqrt_pos(pcur);
          Mr[d] = R[d] * rs;
          if (lane == 8 * k + d) dvs[(8 * k + d) * SA] = rs;
          if (want_ll) pd *= rs;
          __syncwarp();
#pragma unroll
          for (int i = 0; i + 1 < d; i += 2) {
            const double2 tt = *reinterpret_cast<const double2*>(scr + d * 8 + i);
            R[i] = fma(-tt.x, R[d], R[i]);
            R[i + 1] = fma(-tt.y, R[d], R[i + 1]);
          }
          if (d & 1) R[d - 1] = fma(-scr[d * 8 + d - 1], R[d], R[d - 1]);
        }
        if (lane < 8 * (k + 1)) {
#pragma unroll
          for (int d = 0; d < 8; d++) buf[(8 * k + d) * SA + lane] = (lane <= 8 * k + d) ? Mr[d] : 0.0;
        }
        __syncwarp();
      }
    }
    // ---- inverses of all diagonal blocks in one pass: lane = 8 b + c computes column c of M_bb^-1
    {
      const int b = lane >> 3, c = lane & 7;
      const bool valid = b < NBK;
      const double* Mb = buf + (8 * (valid ? b : 0)) * SA + 8 * (valid ? b : 0);
      double Lc[8];
#pragma unroll
      for (int i = 0; i < 8; i++) {
        double a0 = (i == c) ? 1.0 : 0.0, a1 = 0.0;
#pragma unroll
        for (int q = 0; q + 1 < i; q += 2) {
          const double2 mm = *reinterpret_cast<const double2*>(Mb + i * SA + q);
          a0 = fma(-mm.x, Lc[q], a0);
          a1 = fma(-mm.y, Lc[q + 1], a1);
        }
        if (i & 1) a0 = fma(-Mb[i * SA + i - 1], Lc[i - 1], a0);
        Lc[i] = (a0 + a1) * dvs[(8 * (valid ? b : 0) + i) * SA];
      }
      __syncwarp();
      if (valid) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
          buf[(8 * b + i) * SA + 8 * b + c] = Lc[i];   // zero above the diagonal by construction
          const int row = 8 * b + i, col = 8 * b + c;
          if (store && i >= c && row < NN) fwp[Y::FW_L + row * (row + 1) / 2 + col] = Lc[i];
        }
      }
      __syncwarp();
    }
    // ---- phase B: L_ij = -(sum_{j<k<=i} L_ik M_kj) Li_jj.  First every off-diagonal block of M is scaled
    //      on the right in place, N_kj = M_kj Li_jj (independent products); then L_ij = -sum_k L_ik N_kj
    //      has no layout change inside its chain, diagonal by diagonal (blocks at the same distance from
    //      the diagonal are independent).  N stays in the lower blocks until the end: an L block is
    //      parked, untransposed, in the free upper block (j, i), where the next diagonals read it as an
    //      A operand; the finished blocks stay in registers and replace N at the end.
    {
      double Lr[B::NBLK][2];
#pragma unroll
      for (int k = 1; k < NBK; k++)
#pragma unroll
        for (int j = 0; j < k; j++) {
          double* acc = Lr[B::blk(k, j)];
          acc[0] = acc[1] = 0.0;
#pragma unroll
          for (int half = 0; half < 2; half++)
            KB_DMMA(acc, buf[(8 * k + g) * SA + 8 * j + 4 * half + qd], buf[(8 * j + 4 * half + qd) * SA + 8 * j + g]);
        }
      __syncwarp();
#pragma unroll
      for (int k = 1; k < NBK; k++)
#pragma unroll
        for (int j = 0; j < k; j++)
          *reinterpret_cast<double2*>(buf + (8 * k + g) * SA + 8 * j + 2 * qd) =
              make_double2(Lr[B::blk(k, j)][0], Lr[B::blk(k, j)][1]);
      __syncwarp();
#pragma unroll
      for (int sd = 1; sd < NBK; sd++) {
#pragma unroll
        for (int i = sd; i < NBK; i++) {
          const int j = i - sd;
          double* acc = Lr[B::blk(i, j)];
          acc[0] = acc[1] = 0.0;
#pragma unroll
          for (int k = j + 1; k <= i; k++)
#pragma unroll
            for (int half = 0; half < 2; half++) {
              // A operand L_ik: the diagonal block in place, an off-diagonal one from its parking block (k, i)
              const double av = buf[(8 * k + g) * SA + 8 * i + 4 * half + qd];
              KB_DMMA(acc, -av, buf[(8 * k + 4 * half + qd) * SA + 8 * j + g]);
            }
        }
        if (sd + 1 < NBK) {
#pragma unroll
          for (int i = sd; i < NBK; i++)
            *reinterpret_cast<double2*>(buf + (8 * (i - sd) + g) * SA + 8 * i + 2 * qd) =
                make_double2(Lr[B::blk(i, i - sd)][0], Lr[B::blk(i, i - sd)][1]);
        }
        __syncwarp();
      }
#pragma unroll
      for (int i = 1; i < NBK; i++)
#pragma unroll
        for (int j = 0; j < i; j++) {
          const double* l2 = Lr[B::blk(i, j)];
          *reinterpret_cast<double2*>(buf + (8 * i + g) * SA + 8 * j + 2 * qd) = make_double2(l2[0], l2[1]);
          if (store) {
            const int row = 8 * i + g, col = 8 * j + 2 * qd;
            if (row < NN) {
              fwp[Y::FW_L + row * (row + 1) / 2 + col] = l2[0];
              fwp[Y::FW_L + row * (row + 1) / 2 + col + 1] = l2[1];
            }
          }
        }
      __syncwarp();
    }
    // ---- phase C: T = L^T [unit columns | F | h], dense column blocks by DMMA
    if ((lockstep & 2) && tau + 1 < Tt) {   // the G fragments of the next step's state
      const double* tgn = ws.tabg + (size_t)z_next * B::TABG;
#pragma unroll
      for (int o = 0; o < B::TABG; o += 512)
        if (o + lane * 16 < B::TABG) kb_prefetch_l1(tgn + o + lane * 16);
    }
    {
      double Tacc[NBK][NDB][2];
#pragma unroll
      for (int rb = 0; rb < NBK; rb++)
#pragma unroll
        for (int cb = 0; cb < NDB; cb++) Tacc[rb][cb][0] = Tacc[rb][cb][1] = 0.0;
#pragma unroll
      for (int kq = 0; kq < B::KQ; kq++) {
        const int q = 4 * kq + qd;
        double bF[NDB];
#pragma unroll
        for (int cb = 0; cb < NDB; cb++) {
          double v = __ldg(tfz + (kq * NDB + cb) * 32 + lane);
          if (8 * (cb + CB0) + g == NN && q < NN) v = hv[q * SA];   // the mean column
          bF[cb] = v;
        }
#pragma unroll
        for (int rb = 0; rb < NBK; rb++) {
          if (kq < 2 * rb) continue;   // L is lower triangular
          const double av = buf[q * SA + 8 * rb + g];   // (L^T)[8rb+g][q]; zero above the diagonal, identity padding meets zero rows of B
#pragma unroll
          for (int cb = 0; cb < NDB; cb++) KB_DMMA(Tacc[rb][cb], av, bF[cb]);
        }
      }
#pragma unroll
      for (int rb = 0; rb < NBK; rb++)
#pragma unroll
        for (int cb = 0; cb < NDB; cb++)
          *reinterpret_cast<double2*>(Tb + (8 * rb + g) * TS + 8 * cb + 2 * qd) = make_double2(Tacc[rb][cb][0], Tacc[rb][cb][1]);
      __syncwarp();
    }
    // ---- Z = T^T T into the carry registers; block b < CB0 of T is 100 x rows DD + 8b.. of L
    {
      auto tkmax = [](int b) constexpr {
        return (b < CB0) ? ((DD + 8 * b + 7) / 4 < B::KQ - 1 ? (DD + 8 * b + 7) / 4 : B::KQ - 1) : B::KQ - 1;
      };
#pragma unroll
      for (int b = 0; b < B::NBLK; b++) C[b][0] = C[b][1] = 0.0;
#pragma unroll
      for (int kq = 0; kq < B::KQ; kq++) {
        const int r = 4 * kq + qd;
        double Tk[NBK];
#pragma unroll
        for (int b = 0; b < NBK; b++) {
          const int cf = 8 * b + g;
          if (b < CB0) Tk[b] = (kq <= tkmax(b) && r <= DD + cf) ? 100.0 * buf[(DD + cf) * SA + r] : 0.0;
          else Tk[b] = Tb[r * TS + (cf - 8 * CB0)];
        }
#pragma unroll
        for (int ib = 0; ib < NBK; ib++)
#pragma unroll
          for (int jb = 0; jb <= ib; jb++) {
            if (kq > tkmax(jb) || kq > tkmax(ib)) continue;   // structural zeros of the unit-column blocks
            KB_DMMA(C[B::blk(ib, jb)], Tk[ib], Tk[jb]);
          }
      }
      __syncwarp();
    }
    // ---- new carry: Lam_p' = Qinv_ - Z, eta_p' = qb_ + Z[NN][:]; u^T u = Z[NN][NN]
    const double uu_z = __shfl_sync(FULL, C[B::blk(B::IB, B::IB)][B::GE & 1], 4 * B::GE + (B::GE >> 1));
#pragma unroll
    for (int i = 0; i < NBK; i++)
#pragma unroll
      for (int j = 0; j <= i; j++)
#pragma unroll
        for (int e = 0; e < 2; e++) {
          const int row = 8 * i + g, col = 8 * j + 2 * qd + e;
          double qv = (row < OFF && row == col) ? 100.0 : 0.0;
          if (8 * i + 7 >= OFF && 8 * j + 7 >= OFF) {   // blocks that meet Qinv_ / qb_: one unconditional load, then a select
            const bool in = row >= OFF && row <= NN && col >= OFF && col < NN;
            const int idx = (row < NN) ? Y::TAB_QI + (row - OFF) * DD + (col - OFF) : Y::TAB_QB + (col - OFF);
            const double tv = __ldg(tabz + (in ? idx : Y::TAB_QI));
            if (in) qv = tv;
          }
          const double zv = C[B::blk(i, j)][e];
          C[B::blk(i, j)][e] = (row == NN) ? qv + zv : qv - zv;
        }
    if (want_ll && store) llacc += ct - 0.5 * __ldg(tabz + Y::TAB_SC + 1) + log(pd) + 0.5 * (uu_z - __ldg(tabz + Y::TAB_SC + 0));
  }
  if (!has_unit) return;
  if (nch > 1) record_carry(ws.bndf + ((size_t)(n * nch + cc) * 2 + 1) * Y::BND);
  if (cc == nch - 1) record_carry(ws.fin + (size_t)n * Y::BND);   // for kalman_final_kernel
  if (lane == 0) ws.llpart[(size_t)n * nch + cc] = llacc;
}

// Unconditional final update (kalman.py:117) and the draw of the last state (kalman.py:150) from the
// carry the last chunk left behind; one warp per session, lane c owns column c.
template <int NN, int DD>
__global__ void __launch_bounds__(32) kalman_final_kernel(const float* __restrict__ w_last, KalWs ws, int N, int T,
                                                          int Lg, int nch) {
  using Y = KL<NN, DD>;
  constexpr int OFF = NN - DD, NTRI = Y::NTRI, RS = NN | 1;
  constexpr unsigned FULL = 0xffffffffu;
  static_assert(NN <= 32, "one lane per column");
  __shared__ double Zs[NN * RS];
  __shared__ double dv[NN], hv[NN];
  const int n = blockIdx.x, c = threadIdx.x;
  const bool col = c < NN;
  const int Tp = T - Lg + 1;
  const double* fin = ws.fin + (size_t)n * Y::BND;
  const double* o = ws.obs + ((size_t)n * Tp + (Tp - 1)) * Y::OBS_SIZE;
  double Wc[NN];
#pragma unroll
  for (int i = 0; i < NN; i++) {
    double v = (col && i >= c) ? fin[i * (i + 1) / 2 + c] : 0.0;
    if (col && i >= c && c >= OFF) v += o[Y::OBS_J + (i - OFF) * (i - OFF + 1) / 2 + (c - OFF)];
    Wc[i] = v;
  }
  const double eta_c = col ? fin[NTRI + c] + ((c >= OFF) ? o[Y::OBS_E + c - OFF] : 0.0) : 0.0;
  double pd = 1.0;
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
  for (int s = 16; s > 0; s >>= 1) tt += __shfl_xor_sync(FULL, tt, s);
  const double yc = col ? tc + (double)w_last[(size_t)n * NN + c] : 0.0;
  double* xl = ws.fw + ((size_t)n * Tp + (Tp - 1)) * Y::FW_SIZE;
#pragma unroll
  for (int i = 0; i < NN; i++) {
    double v = Lc[i] * yc;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(FULL, v, s);
    if (c == 0) xl[i] = v;
  }
  if (c == 0) ws.llpart[(size_t)n * nch + nch - 1] += o[Y::OBS_C] + log(pd) + 0.5 * tt;
}
