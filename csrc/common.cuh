// Shared host/device helpers for libb200sys (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>

#include "b200sys.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ != 1000)
#error "libb200sys is written for sm_100a only (compile with -gencode arch=compute_100a,code=sm_100a)"
#endif

namespace b200 {

// ---- error reporting through the C ABI -------------------------------------------------------
void set_error(const char* fmt, ...);
const char* get_error();

#define B200_CHECK_CUDA(expr)                                                              \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      ::b200::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return B200_ERR_CUDA;                                                                \
    }                                                                                      \
  } while (0)

#define B200_REQUIRE(cond, ...)                          \
  do {                                                   \
    if (!(cond)) {                                       \
      ::b200::set_error(__VA_ARGS__);                    \
      return B200_ERR_ARG;                               \
    }                                                    \
  } while (0)

// Opt a kernel into more than 48 KB of dynamic shared memory, once per (launch site, device): the attribute is a
// per-device property, so a process that drives several GPUs must set it on each of them.
#define B200_ENSURE_DYN_SMEM(kern, bytes)                                                              \
  do {                                                                                                 \
    static unsigned long long _done_mask = 0ull; /* one bit per device ordinal (mod 64) */             \
    int _dev = 0;                                                                                      \
    B200_CHECK_CUDA(cudaGetDevice(&_dev));                                                             \
    const unsigned long long _bit = 1ull << (_dev & 63);                                               \
    if (!(_done_mask & _bit)) {                                                                        \
      B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
      _done_mask |= _bit;                                                                              \
    }                                                                                                  \
  } while (0)

// Every kernel launch of the library is counted (bench.py reports the number as "gpu_launches"):
// B200_LAUNCH_CHECK() follows each <<<>>> launch, launch_pdl() counts its own.
void count_launch(int n = 1);
long long launch_count();

#define B200_LAUNCH_CHECK()                  \
  do {                                       \
    ::b200::count_launch();                  \
    B200_CHECK_CUDA(cudaGetLastError());     \
  } while (0)

// Launch with the programmatic-stream-serialization attribute (PDL): the kernel may begin while its predecessor
// drains; kernels order themselves with griddepcontrol.wait.  B200SYS_PDL=0 falls back to plain launches.
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  count_launch();
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// Same, with a thread-block cluster of `cluster_x` CTAs along x chosen at launch (1 = no cluster): lets one kernel
// template serve both the clustered and the plain launch.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                      int cluster_x, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (cluster_x > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster_x;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  count_launch();
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

inline int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---- device helpers ----------------------------------------------------------------------------
#ifdef __CUDACC__

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum for up to 1024 threads; every thread gets the result. `red` is >= 32 floats of smem.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` from a previous use
  if (lane == 0) red[w] = v;
  __syncthreads();
  float t = (lane < nw) ? red[lane] : 0.f;
  return warp_sum(t);
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float t = (lane < nw) ? red[lane] : -INFINITY;
  return warp_max(t);
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}

// 16-byte vector views
struct __align__(16) f32x4 { float x, y, z, w; };
struct __align__(16) u32x4 { uint32_t x, y, z, w; };

#endif  // __CUDACC__

}  // namespace b200
