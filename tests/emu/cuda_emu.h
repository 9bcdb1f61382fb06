// TEST INFRASTRUCTURE: a minimal single-process CUDA emulator so the hand-written kernels of
// ska_sdp_screen_fitting_b200/csrc/*.cu (everything except the tcgen05 kernel) can be
// compiled with g++ and checked against the oracle in the GPU-less build container.
//
// Never part of the product: the product library is built by nvcc for sm_100a only and has
// no CPU path.  This header is force-included (-include) when building tests/emu/_build/.
//
// Model: blocks run one after another; the threads of a block are ucontext coroutines on one
// OS thread, so `__shared__` can be a function-local static, `__syncthreads()` and the warp
// shuffles are cooperative yields, and execution is deterministic.
#pragma once
#define SSF_CPU_EMU 1

#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <functional>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __shared__ static
#define __constant__ static
#define __align__(n) __attribute__((aligned(n)))

struct uint3 { unsigned x, y, z; };
struct dim3 {
  unsigned x, y, z;
  dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct float2 { float x, y; };
struct uchar2 { unsigned char x, y; };
static inline uchar2 make_uchar2(unsigned char x, unsigned char y) { return uchar2{x, y}; }
struct alignas(16) float4 { float x, y, z, w; };
struct alignas(16) int4 { int x, y, z, w; };
struct alignas(16) uint4 { unsigned x, y, z, w; };
static inline float2 make_float2(float x, float y) { return float2{x, y}; }
static inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }

namespace emu {
extern uint3 g_threadIdx, g_blockIdx;
extern dim3 g_blockDim, g_gridDim;
extern unsigned char *g_dyn_smem;
void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()> &body);
void sync_block();
uint32_t shfl_u32(uint32_t v, int src_lane);
}  // namespace emu

#define threadIdx emu::g_threadIdx
#define blockIdx emu::g_blockIdx
#define blockDim emu::g_blockDim
#define gridDim emu::g_gridDim

static inline void __syncthreads() { emu::sync_block(); }
static inline void __syncwarp(unsigned = 0xffffffffu) {}

template <typename T>
static inline T __shfl_sync(unsigned, T v, int src) {
  static_assert(sizeof(T) == 4 || sizeof(T) == 8, "shfl size");
  if (sizeof(T) == 4) {
    uint32_t u;
    memcpy(&u, &v, 4);
    u = emu::shfl_u32(u, src);
    memcpy(&v, &u, 4);
    return v;
  }
  uint32_t u[2];
  memcpy(u, &v, 8);
  u[0] = emu::shfl_u32(u[0], src);
  u[1] = emu::shfl_u32(u[1], src);
  memcpy(&v, u, 8);
  return v;
}
template <typename T>
static inline T __shfl_xor_sync(unsigned m, T v, int lanemask) {
  int lane = (threadIdx.x + threadIdx.y * blockDim.x + threadIdx.z * blockDim.x * blockDim.y) & 31;
  return __shfl_sync(m, v, lane ^ lanemask);
}

template <typename T>
static inline T __shfl_up_sync(unsigned m, T v, int delta) {
  int lane = (threadIdx.x + threadIdx.y * blockDim.x + threadIdx.z * blockDim.x * blockDim.y) & 31;
  return __shfl_sync(m, v, lane >= delta ? lane - delta : lane);
}

// warp vote: OR-butterfly over the lanes' bits (every lane of the warp calls, as on the device)
static inline unsigned __ballot_sync(unsigned m, int pred) {
  int lane = (threadIdx.x + threadIdx.y * blockDim.x + threadIdx.z * blockDim.x * blockDim.y) & 31;
  uint32_t v = pred ? (1u << lane) : 0u;
  for (int o = 16; o; o >>= 1) v |= __shfl_xor_sync(m, v, o);
  return v;
}
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
// barrier with an OR reduction over the block (one block runs at a time: a static is the block's)
static inline int __syncthreads_or(int pred) {
  static int acc;
  emu::sync_block();
  if (threadIdx.x == 0 && threadIdx.y == 0 && threadIdx.z == 0) acc = 0;
  emu::sync_block();
  if (pred) acc = 1;                       // a boolean OR, as on the device (not a bitwise one)
  emu::sync_block();
  const int r = acc;
  emu::sync_block();
  return r;
}

template <typename T> static inline T __ldg(const T *p) { return *p; }
template <typename T> static inline void __stcs(T *p, T v) { *p = v; }
static inline double __dadd_rn(double a, double b) { volatile double r = a + b; return r; }
static inline double __dmul_rn(double a, double b) { volatile double r = a * b; return r; }
static inline float __fmul_rn(float a, float b) { volatile float r = a * b; return r; }
static inline float __fadd_rn(float a, float b) { volatile float r = a + b; return r; }
static inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
static inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
static inline float __uint_as_float(uint32_t i) { float f; memcpy(&f, &i, 4); return f; }
static inline uint32_t __float_as_uint(float f) { uint32_t i; memcpy(&i, &f, 4); return i; }
template <typename T> static inline T atomicExch(T *p, T v) { T o = *p; *p = v; return o; }
template <typename T> static inline T atomicAdd(T *p, T v) { T o = *p; *p = o + v; return o; }
template <typename T> static inline T atomicSub(T *p, T v) { T o = *p; *p = o - v; return o; }
template <typename T> static inline T atomicCAS(T *p, T cmp, T v) { T o = *p; if (o == cmp) *p = v; return o; }
static inline void __threadfence() {}
static inline uint32_t __byte_perm(uint32_t a, uint32_t b, uint32_t sel) {
  uint64_t v = ((uint64_t)b << 32) | a;
  uint32_t r = 0;
  for (int i = 0; i < 4; ++i) r |= (uint32_t)((v >> (8 * ((sel >> (4 * i)) & 7))) & 0xff) << (8 * i);
  return r;
}
// glibc declares __sinf/__cosf itself: map the CUDA intrinsics with macros (after <math.h>)
#define __sinf(x) sinf(x)
#define __cosf(x) cosf(x)
static inline int __popcll(unsigned long long v) { return __builtin_popcountll(v); }

// ---- the slice of the runtime API the library uses ---------------------------------------
typedef int cudaError_t;
typedef void *cudaStream_t;
enum { cudaSuccess = 0 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize };
static inline const char *cudaGetErrorString(cudaError_t) { return "emulated"; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaMalloc(void **p, size_t n) { *p = malloc(n); return *p ? 0 : 2; }
static inline cudaError_t cudaFree(void *p) { free(p); return 0; }
static inline cudaError_t cudaMallocAsync(void **p, size_t n, cudaStream_t) { *p = malloc(n); return *p ? 0 : 2; }
static inline cudaError_t cudaFreeAsync(void *p, cudaStream_t) { free(p); return 0; }
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t) { memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemcpy2DAsync(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h, cudaMemcpyKind,
                                            cudaStream_t) {
  for (size_t r = 0; r < h; ++r) memcpy((char *)d + r * dp, (const char *)s + r * sp, w);
  return 0;
}
static inline cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t) { memset(d, v, n); return 0; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
static inline cudaError_t cudaDeviceSynchronize() { return 0; }
template <typename F> static inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return 0; }
static inline cudaError_t cudaGetDevice(int *d) { *d = 0; return 0; }
// a __device__ array is a plain host array here, so its "device address" is its own address
template <typename T> static inline cudaError_t cudaGetSymbolAddress(void **p, T &sym) { *p = (void *)&sym; return 0; }
template <typename T>
static inline cudaError_t cudaMemcpyToSymbolAsync(T &sym, const void *s, size_t n, size_t off, cudaMemcpyKind, cudaStream_t) {
  memcpy((char *)&sym + off, s, n);
  return 0;
}

#define SSF_LAUNCH(kernel, grid, block, smem, stream, ...) \
  emu::launch(dim3(grid), dim3(block), (smem), [&]() { kernel(__VA_ARGS__); })
#define SSF_DYN_SMEM(type, name) type *name = reinterpret_cast<type *>(emu::g_dyn_smem)
