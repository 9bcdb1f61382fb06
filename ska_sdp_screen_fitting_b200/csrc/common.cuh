// Shared helpers of libssf_b200: error reporting, launch accounting, small device utilities.
#pragma once
#ifndef SSF_CPU_EMU
#include <cuda_runtime.h>
// Launch / dynamic-shared-memory spellings go through two macros so that the GPU-less test
// build (tests/emu, g++) can compile the same kernel sources; the product build is nvcc only.
#define SSF_LAUNCH(kernel, grid, block, smem, stream, ...) \
  kernel<<<grid, block, smem, stream>>>(__VA_ARGS__)
#define SSF_DYN_SMEM(type, name)                                   \
  extern __shared__ __align__(16) unsigned char name##_raw_[];     \
  type *name = reinterpret_cast<type *>(name##_raw_)
#endif
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ssf.h"

namespace ssf {

void set_error(const char *fmt, ...);
void count_launch(int n = 1);
// Streaming multiprocessors of the current device (queried once per device; grids of the
// persistent / capped kernels are sized from it).
int sm_count();

inline int div_up(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

}  // namespace ssf

#define SSF_REQUIRE(cond, ...)            \
  do {                                    \
    if (!(cond)) {                        \
      ssf::set_error(__VA_ARGS__);        \
      return SSF_EINVAL;                  \
    }                                     \
  } while (0)

#define SSF_CUDA(call)                                                              \
  do {                                                                              \
    cudaError_t e_ = (call);                                                        \
    if (e_ != cudaSuccess) {                                                        \
      ssf::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, \
                     __LINE__);                                                     \
      return SSF_ECUDA;                                                             \
    }                                                                               \
  } while (0)

// After a <<<>>> launch: count it and surface launch-configuration errors.
#define SSF_LAUNCHED()                    \
  do {                                    \
    ssf::count_launch();                  \
    SSF_CUDA(cudaGetLastError());         \
  } while (0)

namespace ssf {

// Structure-function kernel value for a squared distance (in screen pixels^2):
//   -((d2 / r0^2)^(beta/2)) / 2        stationscreen.py:424-426, kl_screen.py:447-448
__device__ __forceinline__ double structure_fn(double d2, double inv_r0sq, double half_beta) {
  double q = d2 * inv_r0sq;
  return q > 0.0 ? -0.5 * pow(q, half_beta) : 0.0;
}

// NaN fix of screen.py:364-378: real planes -> 1, imaginary planes -> 0.
__device__ __forceinline__ float fix_re(float v) { return v != v ? 1.0f : v; }
__device__ __forceinline__ float fix_im(float v) { return v != v ? 0.0f : v; }

// Output format of the cube kernels (SSF_OUT_* flags of ssf.h), resolved once on the host:
//   prmt   = byte-permute selector applied to every stored word: 0x3210 keeps the native
//            little-endian float32, 0x0123 stores it byte-swapped = the big-endian order of a
//            BITPIX -32 FITS cube (the swap of hdu.writeto, screen.py:381-382, fused into the
//            store: one PRMT per value, no second pass over the slab)
//   planes = 4: slab [ncol, 4, npix]; 2: only planes 0 and 1 are stored, slab [ncol, 2, npix]
//            (phase-only solutions: planes 2, 3 repeat 0, 1 -- kl_screen.py:368-373,
//            voronoi_screen.py:194-198 -- and the FITS writer places the pair twice);
//            1: one plane of raw screen values (tec cube, processing_utils.py:187-190)
struct OutFmt {
  uint32_t prmt;
  int planes;
};
inline bool out_fmt(int flags, OutFmt *f) {
  if (flags & ~(SSF_OUT_BIG_ENDIAN | SSF_OUT_PLANES2)) return false;
  f->prmt = (flags & SSF_OUT_BIG_ENDIAN) ? 0x0123u : 0x3210u;
  f->planes = (flags & SSF_OUT_PLANES2) ? 2 : 4;
  return true;
}
__device__ __forceinline__ float out_word(float v, uint32_t prmt) {
  return __uint_as_float(__byte_perm(__float_as_uint(v), 0u, prmt));
}

}  // namespace ssf
