// Version / status strings and the Philox fill kernels of libkpms_b200.
#include "common.cuh"

unsigned long long kpms_launch_counter = 0;

extern "C" int kpms_version(void) { return 100; }
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
// yields 4 uniforms / 4 normals, so thread g fills elements 4g..4g+3.
__global__ void fill_uniform_kernel(float* __restrict__ out, size_t n, uint2 key, uint32_t sid) {
  size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t i = g * 4;
  if (i >= n) return;
  uint4 r = philox4x32_10(make_uint4((uint32_t)g, (uint32_t)(g >> 32), 0u, sid), key);
  float v[4] = {u01(r.x), u01(r.y), u01(r.z), u01(r.w)};
  if (i + 3 < n && ((uintptr_t)(out + i) % 16 == 0)) {
    *reinterpret_cast<float4*>(out + i) = make_float4(v[0], v[1], v[2], v[3]);
  } else {
    for (int j = 0; j < 4 && i + j < n; j++) out[i + j] = v[j];
  }
}

__global__ void fill_normal_kernel(float* __restrict__ out, size_t n, uint2 key, uint32_t sid) {
  size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t i = g * 4;
  if (i >= n) return;
  uint4 r = philox4x32_10(make_uint4((uint32_t)g, (uint32_t)(g >> 32), 1u, sid), key);
  float2 a = box_muller(r.x, r.y), b = box_muller(r.z, r.w);
  float v[4] = {a.x, a.y, b.x, b.y};
  if (i + 3 < n && ((uintptr_t)(out + i) % 16 == 0)) {
    *reinterpret_cast<float4*>(out + i) = make_float4(v[0], v[1], v[2], v[3]);
  } else {
    for (int j = 0; j < 4 && i + j < n; j++) out[i + j] = v[j];
  }
}

// Marsaglia & Tsang (2000) for shape >= 1; one element per thread, rejection rounds
// advance the third counter word.
__global__ void fill_gamma_kernel(float* __restrict__ out, size_t n, float shape, uint2 key,
                                  uint32_t sid) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float d = shape - 1.0f / 3.0f;
  const float c = rsqrtf(9.0f * d);
  float res = d;
  for (uint32_t round = 0; round < 64u; round++) {
    uint4 r = philox4x32_10(make_uint4((uint32_t)i, (uint32_t)(i >> 32), 2u + round, sid), key);
    float2 nn = box_muller(r.x, r.y);
    float xx = nn.x;
    float v = 1.0f + c * xx;
    if (v <= 0.0f) continue;
    v = v * v * v;
    float u = u01_open0(r.z);
    if (logf(u) < 0.5f * xx * xx + d - d * v + d * logf(v)) {
      res = d * v;
      break;
    }
  }
  out[i] = res;
}

static inline uint2 split_seed(uint64_t seed) {
  return make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
}

extern "C" int kpms_fill_uniform(float* out, size_t n, uint64_t seed, uint64_t stream_id,
                                 void* stream) {
  KPMS_REQUIRE(out != nullptr || n == 0, KPMS_ERR_NULL);
  if (n == 0) return KPMS_OK;
  size_t groups = (n + 3) / 4;
  fill_uniform_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      out, n, split_seed(seed), (uint32_t)stream_id);
  return kpms_last_launch();
}

extern "C" int kpms_fill_normal(float* out, size_t n, uint64_t seed, uint64_t stream_id,
                                void* stream) {
  KPMS_REQUIRE(out != nullptr || n == 0, KPMS_ERR_NULL);
  if (n == 0) return KPMS_OK;
  size_t groups = (n + 3) / 4;
  fill_normal_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      out, n, split_seed(seed), (uint32_t)stream_id);
  return kpms_last_launch();
}

extern "C" int kpms_fill_gamma(float* out, size_t n, double shape, uint64_t seed,
                               uint64_t stream_id, void* stream) {
  KPMS_REQUIRE(out != nullptr || n == 0, KPMS_ERR_NULL);
  KPMS_REQUIRE(shape >= 1.0, KPMS_ERR_UNSUPPORTED);
  if (n == 0) return KPMS_OK;
  fill_gamma_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      out, n, (float)shape, split_seed(seed), (uint32_t)stream_id);
  return kpms_last_launch();
}
