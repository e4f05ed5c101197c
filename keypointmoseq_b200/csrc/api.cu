// Version / status strings and the Philox fill kernels of libkpms_b200.
#include "common.cuh"

#include <stdlib.h>
#include <string.h>

unsigned long long kpms_launch_counter = 0;

// Debugging / cross-check switches.  The environment is read ONCE, when the library is loaded;
// afterwards only kpms_set_option changes them (no getenv on any launch path).
static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return (e && e[0]) ? atoi(e) : dflt;
}
KpmsOptions kpms_opts = {env_int("KPMS_KALMAN_IMPL", 2), env_int("KPMS_HMM_TC", 1), env_int("KPMS_HMM_BWD_BURN", -1),
                         env_int("KPMS_LL_TILED", 0), env_int("KPMS_HMM_FLAGS", 0), env_int("KPMS_LOCATION_IMPL", 0), env_int("KPMS_KALMAN_WARPS", 12),
                         env_int("KPMS_KALMAN_LOCKSTEP", 0), env_int("KPMS_EMISSION_PIPE", 1), env_int("KPMS_HMM_REPAIR_ROUNDS", 4), env_int("KPMS_HMM_REPAIR_TOL_PPM", 1000), env_int("KPMS_HMM_LL_ROWMAJOR", 0),
                         env_int("KPMS_KALMAN_OBS_MMA", 1)};

static int* option_slot(const char* name) {
  if (!name) return nullptr;
  if (!strcmp(name, "kalman_impl")) return &kpms_opts.kalman_impl;
  if (!strcmp(name, "hmm_tc")) return &kpms_opts.hmm_tc;
  if (!strcmp(name, "hmm_bwd_burn")) return &kpms_opts.hmm_bwd_burn;
  if (!strcmp(name, "ll_tiled")) return &kpms_opts.ll_tiled;
  if (!strcmp(name, "hmm_flags")) return &kpms_opts.hmm_flags;
  if (!strcmp(name, "location_impl")) return &kpms_opts.location_impl;
  if (!strcmp(name, "kalman_warps")) return &kpms_opts.kalman_warps;
  if (!strcmp(name, "kalman_lockstep")) return &kpms_opts.kalman_lockstep;
  if (!strcmp(name, "kalman_obs_mma")) return &kpms_opts.kalman_obs_mma;
  if (!strcmp(name, "emission_pipe")) return &kpms_opts.emission_pipe;
  if (!strcmp(name, "hmm_repair_rounds")) return &kpms_opts.hmm_repair_rounds;
  if (!strcmp(name, "hmm_repair_tol_ppm")) return &kpms_opts.hmm_repair_tol_ppm;
  if (!strcmp(name, "hmm_ll_rowmajor")) return &kpms_opts.hmm_ll_rowmajor;
  return nullptr;
}
extern "C" int kpms_set_option(const char* name, int value) {
  int* slot = option_slot(name);
  KPMS_REQUIRE(slot != nullptr, KPMS_ERR_UNSUPPORTED);
  *slot = value;
  return KPMS_OK;
}
extern "C" int kpms_get_option(const char* name, int* value) {
  int* slot = option_slot(name);
  KPMS_REQUIRE(slot != nullptr && value != nullptr, KPMS_ERR_UNSUPPORTED);
  *value = *slot;
  return KPMS_OK;
}

extern "C" int kpms_version(void) { return 200; }
extern "C" uint64_t kpms_launch_count(void) { return kpms_launch_counter; }

extern "C" const char* kpms_status_string(int status) {
  switch (status) {
    case KPMS_OK: return "ok";
    case KPMS_ERR_NULL: return "kpms: a required pointer is NULL";
    case KPMS_ERR_SHAPE: return "kpms: a dimension is <= 0 or inconsistent";
    case KPMS_ERR_UNSUPPORTED: return "kpms: size not supported by the compiled kernels";
    case KPMS_ERR_WORKSPACE: return "kpms: workspace too small or misaligned";
    case KPMS_ERR_ALIGN: return "kpms: pointer not aligned as required";
    case KPMS_ERR_TOO_LARGE:
      return "kpms: problem too large for the built-in fixed-size scratch; use the entry point that takes a workspace";
    default: break;
  }
  if (status > 0) return cudaGetErrorString((cudaError_t)status);
  return "kpms: unknown status";
}

// counter = (element-group index lo, hi, round, stream_id); key = seed.  One Philox call
// yields 4 uniforms / 4 normals: group G holds the elements 4G..4G+3 of the GLOBAL index space
// (global index = offset + local index), so a shard that fills only its own sessions gets exactly
// the values a single process filling the whole array would have put there.
template <int NORMAL>
__global__ void fill4_kernel(float* __restrict__ out, size_t n, unsigned long long offset, uint2 key,
                             uint32_t sid) {
  const unsigned long long G = (offset >> 2) + (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long i = (long long)(G * 4ULL - offset);   // local index of the group's first element (may be < 0)
  if (i >= (long long)n) return;
  uint4 r = philox4x32_10(make_uint4((uint32_t)G, (uint32_t)(G >> 32), NORMAL ? 1u : 0u, sid), key);
  float v[4];
  if (NORMAL) {
    float2 a = box_muller(r.x, r.y), b = box_muller(r.z, r.w);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
  } else {
    v[0] = u01(r.x); v[1] = u01(r.y); v[2] = u01(r.z); v[3] = u01(r.w);
  }
  if (i >= 0 && i + 3 < (long long)n && ((uintptr_t)(out + i) % 16 == 0)) {
    *reinterpret_cast<float4*>(out + i) = make_float4(v[0], v[1], v[2], v[3]);
  } else {
    for (int j = 0; j < 4; j++)
      if (i + j >= 0 && i + j < (long long)n) out[i + j] = v[j];
  }
}

// Marsaglia & Tsang (2000) for shape >= 1; one element per thread, rejection rounds
// advance the third counter word.
__global__ void fill_gamma_kernel(float* __restrict__ out, size_t n, unsigned long long offset, float shape,
                                  uint2 key, uint32_t sid) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = philox_gamma(offset + i, shape, key, sid);
}


static int fill4(int normal, float* out, size_t n, uint64_t offset, uint64_t seed, uint64_t stream_id, void* stream) {
  KPMS_REQUIRE(out != nullptr || n == 0, KPMS_ERR_NULL);
  if (n == 0) return KPMS_OK;
  const size_t groups = (size_t)(((offset + n + 3) >> 2) - (offset >> 2));
  const unsigned blocks = (unsigned)((groups + 255) / 256);
  if (normal)
    fill4_kernel<1><<<blocks, 256, 0, (cudaStream_t)stream>>>(out, n, offset, split_seed(seed), (uint32_t)stream_id);
  else
    fill4_kernel<0><<<blocks, 256, 0, (cudaStream_t)stream>>>(out, n, offset, split_seed(seed), (uint32_t)stream_id);
  return kpms_last_launch();
}

extern "C" int kpms_fill_uniform_at(float* out, size_t n, uint64_t offset, uint64_t seed, uint64_t stream_id,
                                    void* stream) {
  return fill4(0, out, n, offset, seed, stream_id, stream);
}
extern "C" int kpms_fill_normal_at(float* out, size_t n, uint64_t offset, uint64_t seed, uint64_t stream_id,
                                   void* stream) {
  return fill4(1, out, n, offset, seed, stream_id, stream);
}
extern "C" int kpms_fill_gamma_at(float* out, size_t n, uint64_t offset, double shape, uint64_t seed,
                                  uint64_t stream_id, void* stream) {
  KPMS_REQUIRE(out != nullptr || n == 0, KPMS_ERR_NULL);
  KPMS_REQUIRE(shape >= 1.0, KPMS_ERR_UNSUPPORTED);
  if (n == 0) return KPMS_OK;
  fill_gamma_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      out, n, offset, (float)shape, split_seed(seed), (uint32_t)stream_id);
  return kpms_last_launch();
}
extern "C" int kpms_fill_uniform(float* out, size_t n, uint64_t seed, uint64_t stream_id, void* stream) {
  return fill4(0, out, n, 0, seed, stream_id, stream);
}
extern "C" int kpms_fill_normal(float* out, size_t n, uint64_t seed, uint64_t stream_id, void* stream) {
  return fill4(1, out, n, 0, seed, stream_id, stream);
}
extern "C" int kpms_fill_gamma(float* out, size_t n, double shape, uint64_t seed, uint64_t stream_id,
                               void* stream) {
  return kpms_fill_gamma_at(out, n, 0, shape, seed, stream_id, stream);
}
