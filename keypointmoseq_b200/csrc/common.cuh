// Shared device/host helpers for libkpms_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#include "../../include/kpms_b200.h"

#define KPMS_REQUIRE(cond, code) \
  do {                           \
    if (!(cond)) return (code);  \
  } while (0)

#define KPMS_CUDA(call)                      \
  do {                                       \
    cudaError_t e__ = (call);                \
    if (e__ != cudaSuccess) return (int)e__; \
  } while (0)

// debugging / cross-check switches (api.cu): read from the environment once at load time,
// changed afterwards only through kpms_set_option
struct KpmsOptions {
  int kalman_impl;   // 2 (default): blocked DMMA warp kernel; 1: lane-per-column warp kernel; 0: CTA-per-chunk kernel
  int hmm_tc;        // 0: CUDA-core chunked HMM filter instead of the tcgen05 one
  int hmm_bwd_burn;  // >= 0: warm-up of the HMM sampler only (0 forces every chunk through the repair pass)
  int ll_tiled;      // 1: the emission kernel writes the interleaved HMM layout directly
  int hmm_flags;     // bit 0: forward pass only (timing)
  int location_impl; // 0 (default): exact associative scan over time; 1: chunks with a fixed warm-up (no verification)
  int kalman_warps;  // warps per CTA of the blocked Kalman filter: 8, 12 (default) or 16
  int kalman_lockstep;  // blocked filter, bit mask: 1 = the warps of a CTA step together (no gain measured), 2 = L1 prefetch of the per-state tables; default 0
  int emission_pipe;    // 1 (default): software-pipelined emission kernel (two A stages, two TMEM accumulators); 0: one block at a time, two CTAs per SM
  int hmm_repair_rounds;   // K > 128 forward filter: rounds of device-side re-filtering of mismatched chunks (default 4; 0 = off)
  int hmm_repair_tol_ppm;  // ... and the mismatch (in 1e-6 of probability mass) above which a chunk is re-filtered (default 1000)
  int hmm_ll_rowmajor;     // K <= 104 filter: 1 = read a row-major ll in place when K % 4 == 0 instead of re-tiling it first
  int kalman_obs_mma;   // 1 (default): observation records by FP64 tensor cores when dk == 2 and k % 4 == 0; 0: thread-per-frame kernel
};
extern KpmsOptions kpms_opts;

// host-side count of kernel launches issued by this library (one bump per launch site)
extern unsigned long long kpms_launch_counter;
static inline int kpms_last_launch() {
  ++kpms_launch_counter;
  return (int)cudaGetLastError();
}
static inline bool kpms_aligned(const void* p, size_t a) { return ((uintptr_t)p % a) == 0; }
static inline size_t kpms_round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

#define KPMS_PI 3.14159265358979323846
#define KPMS_LOG_2PI 1.8378770664093454835606594728112

// ---------------------------------------------------------------- warp helpers
__device__ __forceinline__ float warp_max_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// 1/x for x > 0, normal: hardware seed (2^-22) and one third-order correction; relative error ~2^-66
// before the final rounding, without the special-case handling of the IEEE division sequence.
__device__ __forceinline__ double rcp_pos_d(double x) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-x, y, 1.0);
  return fma(y, fma(e, e, e), y);
}

// ---------------------------------------------------------------- Philox4x32-10
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < 10; i++) {
    uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}
static inline uint2 split_seed(uint64_t seed) { return make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)); }
// [0,1) with 24 random bits (exactly representable in fp32)
__device__ __forceinline__ float u01(uint32_t r) { return (float)(r >> 8) * (1.0f / 16777216.0f); }
// (0,1]
__device__ __forceinline__ float u01_open0(uint32_t r) {
  return ((float)(r >> 8) + 1.0f) * (1.0f / 16777216.0f);
}
__device__ __forceinline__ float2 box_muller(uint32_t a, uint32_t b) {
  float r = sqrtf(-2.0f * logf(u01_open0(a)));
  float s, c;
  sincospif(2.0f * u01(b), &s, &c);
  return make_float2(r * c, r * s);
}

// Gamma(shape, 1), shape >= 1 (Marsaglia & Tsang 2000) for GLOBAL element index e; the value depends
// on (seed, stream id, e) only.  Elements 2m and 2m+1 share the Philox block of their first round
// (counter (m, 1, sid): its two Box-Muller normals and two uniforms serve one element each), which
// halves the cost for kernels that draw for consecutive elements; later rounds (about one element
// in twenty needs one) are keyed by the element itself, third counter word 2 + round.
__device__ __forceinline__ bool gamma_mt_try(float xx, float u, float d, float c, float& res) {
  float v = 1.0f + c * xx;
  if (v <= 0.0f) return false;
  v = v * v * v;
  if (logf(u) < 0.5f * xx * xx + d - d * v + d * logf(v)) {
    res = d * v;
    return true;
  }
  return false;
}
static __device__ __noinline__ float philox_gamma_retry(unsigned long long e, float d, float c, uint2 key, uint32_t sid) {
  float res = d;
  for (uint32_t round = 0; round < 64u; round++) {
    const uint4 r = philox4x32_10(make_uint4((uint32_t)e, (uint32_t)(e >> 32), 2u + round, sid), key);
    const float2 nn = box_muller(r.x, r.y);
    if (gamma_mt_try(nn.x, u01_open0(r.z), d, c, res)) break;
  }
  return res;
}
// the variates of elements e0 and e0 + 1, e0 even
__device__ __forceinline__ void philox_gamma_pair(unsigned long long e0, float shape, uint2 key, uint32_t sid, float& g0,
                                                  float& g1) {
  const float d = shape - 1.0f / 3.0f;
  const float c = rsqrtf(9.0f * d);
  const unsigned long long m = e0 >> 1;
  const uint4 r = philox4x32_10(make_uint4((uint32_t)m, (uint32_t)(m >> 32), 1u, sid), key);
  const float2 nn = box_muller(r.x, r.y);
  if (!gamma_mt_try(nn.x, u01_open0(r.z), d, c, g0)) g0 = philox_gamma_retry(e0, d, c, key, sid);
  if (!gamma_mt_try(nn.y, u01_open0(r.w), d, c, g1)) g1 = philox_gamma_retry(e0 + 1, d, c, key, sid);
}
__device__ __forceinline__ float philox_gamma(unsigned long long e, float shape, uint2 key, uint32_t sid) {
  float g0, g1;
  philox_gamma_pair(e & ~1ull, shape, key, sid, g0, g1);
  return (e & 1ull) ? g1 : g0;
}
// uniform [0,1) number `e` of the stream a kpms_fill_uniform(seed, sid) call would have produced
__device__ __forceinline__ float philox_uniform_at(unsigned long long e, uint2 key, uint32_t sid) {
  const unsigned long long G = e >> 2;
  const uint4 r = philox4x32_10(make_uint4((uint32_t)G, (uint32_t)(G >> 32), 0u, sid), key);
  const uint32_t w = (e & 3) == 0 ? r.x : ((e & 3) == 1 ? r.y : ((e & 3) == 2 ? r.z : r.w));
  return u01(w);
}

// Four keypoints of one frame: Y (dk coordinates each) and, optionally, s.  Thread-per-frame kernels
// walk their frame's keypoints in groups of four: with dk == 2 and k % 4 == 0 (every BASELINE config)
// a group is two 16-byte loads of Y and one of s instead of twelve 4-byte ones — the scalar loads of a
// warp hit 32 different sectors each, which made these kernels LSU-bound at a fifth of the HBM rate.
struct Kp4 {
  float y[12];   // [j*dk + dim]
  float s[4];
};
__device__ __forceinline__ void load_kp4(const float* __restrict__ Yf, const float* __restrict__ sf, int kp0, int k,
                                         int dk, bool vec, Kp4& o) {
  if (vec) {   // dk == 2, k % 4 == 0, 16-byte aligned rows
    const float4 a = __ldg(reinterpret_cast<const float4*>(Yf + kp0 * 2));
    const float4 b = __ldg(reinterpret_cast<const float4*>(Yf + kp0 * 2) + 1);
    o.y[0] = a.x; o.y[1] = a.y; o.y[2] = a.z; o.y[3] = a.w;
    o.y[4] = b.x; o.y[5] = b.y; o.y[6] = b.z; o.y[7] = b.w;
    if (sf) {
      const float4 c = __ldg(reinterpret_cast<const float4*>(sf + kp0));
      o.s[0] = c.x; o.s[1] = c.y; o.s[2] = c.z; o.s[3] = c.w;
    }
  } else {
#pragma unroll
    for (int e = 0; e < 12; e++) o.y[e] = 0.0f;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      o.s[j] = 1.0f;
      if (kp0 + j < k) {
        const float* yp = Yf + (kp0 + j) * dk;
        if (dk == 2) {
          o.y[2 * j] = yp[0];
          o.y[2 * j + 1] = yp[1];
        } else {
          o.y[3 * j] = yp[0];
          o.y[3 * j + 1] = yp[1];
          o.y[3 * j + 2] = yp[2];
        }
        if (sf) o.s[j] = sf[kp0 + j];
      }
    }
  }
}
// coordinates of keypoint j (compile-time) of the group, without dynamic register indexing
#define KP4_Y(q, j, dk, dim) ((dk) == 2 ? ((dim) < 2 ? (q).y[2 * (j) + ((dim) < 2 ? (dim) : 0)] : 0.0f) : (q).y[3 * (j) + (dim)])
__device__ __forceinline__ bool kp4_vec_ok(const float* Y, const float* s, int k, int dk) {
  return dk == 2 && (k & 3) == 0 && ((uintptr_t)Y % 16 == 0) && (s == nullptr || (uintptr_t)s % 16 == 0);
}

// R(-h)(y - v) in the first two coordinates (reference util.py:140-166)
__device__ __forceinline__ void rot_inv(double c, double s, double y0, double y1, double& o0,
                                        double& o1) {
  o0 = c * y0 + s * y1;
  o1 = -s * y0 + c * y1;
}
