// Voronoi (tessellated) screen kernels: nearest-patch template, cos/sin LUT, cube fill,
// indicator images, NaN fix.  All are HBM-bound streaming kernels: 16 B written per voxel.
#include <stdlib.h>

#include "common.cuh"

namespace {

// ------------------------------------------------------------------------------------
// Template: one thread per pixel, patch sites staged in shared memory, float64 distance
// compare, strict '<' so the lowest patch index wins ties (bit-exact against the float64
// CPU definition oracle.port.screens.nearest_patch_template).
__global__ void assign_kernel(const double *__restrict__ px, const double *__restrict__ py,
                              int ndir, int nx, int ny, int32_t *__restrict__ tmpl) {
  __shared__ double sx[SSF_MAX_DIR * 8], sy[SSF_MAX_DIR * 8];
  for (int d = threadIdx.x; d < ndir; d += blockDim.x) {
    sx[d] = px[d];
    sy[d] = py[d];
  }
  __syncthreads();
  int64_t npix = (int64_t)nx * ny;
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < npix;
       p += (int64_t)gridDim.x * blockDim.x) {
    double x = (double)(p % nx), y = (double)(p / nx);
    double best = INFINITY;
    int arg = 0;
    for (int d = 0; d < ndir; ++d) {
      // same association as numpy: (x-px)^2 + (y-py)^2, no FMA contraction
      double dx = x - sx[d], dy = y - sy[d];
      double d2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
      if (d2 < best) {
        best = d2;
        arg = d;
      }
    }
    tmpl[p] = arg + 1;
  }
}

// LUT: lut[col, d, 0] = (float)(amp*cos(phi)), lut[col, d, 1] = (float)(amp*sin(phi)),
// evaluated in FP64 so the single rounding to float32 matches numpy's.
__global__ void lut_kernel(const double *__restrict__ phase, const double *__restrict__ amp,
                           int64_t n, float2 *__restrict__ lut) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    double s, c;
    sincos(phase[i], &s, &c);
    double a = amp ? amp[i] : 1.0;
    lut[i] = make_float2((float)(a * c), (float)(a * s));
  }
}

// Fill: work item = (column, run of 4*VEC consecutive pixels).  The template is tiny and
// L2/L1 resident; the LUT row of a column is 8*ndir bytes and L1 resident; the only HBM
// stream is the 4-plane store.  float4 stores when the plane size allows 16 B alignment.
template <bool VEC4>
__global__ void fill_kernel(const int32_t *__restrict__ tmpl, const float2 *__restrict__ lut,
                            int ndir, int64_t npix, int64_t ncol, float *__restrict__ slab,
                            int fix_nan, ssf::OutFmt fmt) {
  const int64_t per_col = VEC4 ? npix / 4 : npix;
  const int64_t total = per_col * ncol;
  const bool four = fmt.planes == 4;
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < total;
       w += (int64_t)gridDim.x * blockDim.x) {
    int64_t col = w / per_col;
    int64_t q = w - col * per_col;
    const float2 *row = lut + col * ndir;
    float *out = slab + col * fmt.planes * npix;
    if (VEC4) {
      int4 t = __ldg(reinterpret_cast<const int4 *>(tmpl) + q);
      float2 a = __ldg(row + (t.x - 1)), b = __ldg(row + (t.y - 1));
      float2 c = __ldg(row + (t.z - 1)), d = __ldg(row + (t.w - 1));
      float4 re = make_float4(a.x, b.x, c.x, d.x), im = make_float4(a.y, b.y, c.y, d.y);
      if (fix_nan) {
        re = make_float4(ssf::fix_re(re.x), ssf::fix_re(re.y), ssf::fix_re(re.z), ssf::fix_re(re.w));
        im = make_float4(ssf::fix_im(im.x), ssf::fix_im(im.y), ssf::fix_im(im.z), ssf::fix_im(im.w));
      }
      re = make_float4(ssf::out_word(re.x, fmt.prmt), ssf::out_word(re.y, fmt.prmt), ssf::out_word(re.z, fmt.prmt),
                       ssf::out_word(re.w, fmt.prmt));
      im = make_float4(ssf::out_word(im.x, fmt.prmt), ssf::out_word(im.y, fmt.prmt), ssf::out_word(im.z, fmt.prmt),
                       ssf::out_word(im.w, fmt.prmt));
      float4 *o = reinterpret_cast<float4 *>(out) + q;
      const int64_t ps = npix / 4;
      __stcs(o, re);
      __stcs(o + ps, im);
      if (four) {
        __stcs(o + 2 * ps, re);
        __stcs(o + 3 * ps, im);
      }
    } else {
      float2 a = __ldg(row + (__ldg(tmpl + q) - 1));
      float re = fix_nan ? ssf::fix_re(a.x) : a.x, im = fix_nan ? ssf::fix_im(a.y) : a.y;
      re = ssf::out_word(re, fmt.prmt);
      im = ssf::out_word(im, fmt.prmt);
      out[q] = re;
      out[npix + q] = im;
      if (four) {
        out[2 * npix + q] = re;
        out[3 * npix + q] = im;
      }
    }
  }
}

// Gain (amplitude + phase) solutions: planes 0/1 carry the XX amplitude, planes 2/3 the YY one
// (voronoi_screen.py:177-214).  lut4[col, d] = (axx cos, axx sin, ayy cos, ayy sin), the FP64
// products rounded once to float32.
__global__ void lut_gain_kernel(const double *__restrict__ phase, const double *__restrict__ amp,
                                int64_t n, float4 *__restrict__ lut) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    double s, c;
    sincos(phase[i], &s, &c);
    const double ax = amp[2 * i], ay = amp[2 * i + 1];
    lut[i] = make_float4((float)(ax * c), (float)(ax * s), (float)(ay * c), (float)(ay * s));
  }
}

template <bool VEC4>
__global__ void fill_gain_kernel(const int32_t *__restrict__ tmpl, const float4 *__restrict__ lut,
                                 int ndir, int64_t npix, int64_t ncol, float *__restrict__ slab,
                                 int fix_nan, uint32_t prmt) {
  const int64_t per_col = VEC4 ? npix / 4 : npix;
  const int64_t total = per_col * ncol;
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < total;
       w += (int64_t)gridDim.x * blockDim.x) {
    int64_t col = w / per_col;
    int64_t q = w - col * per_col;
    const float4 *row = lut + col * ndir;
    float *out = slab + col * 4 * npix;
    if (VEC4) {
      int4 t = __ldg(reinterpret_cast<const int4 *>(tmpl) + q);
      float4 a = __ldg(row + (t.x - 1)), b = __ldg(row + (t.y - 1));
      float4 c = __ldg(row + (t.z - 1)), d = __ldg(row + (t.w - 1));
      float4 p0 = make_float4(a.x, b.x, c.x, d.x), p1 = make_float4(a.y, b.y, c.y, d.y);
      float4 p2 = make_float4(a.z, b.z, c.z, d.z), p3 = make_float4(a.w, b.w, c.w, d.w);
      if (fix_nan) {
        p0 = make_float4(ssf::fix_re(p0.x), ssf::fix_re(p0.y), ssf::fix_re(p0.z), ssf::fix_re(p0.w));
        p1 = make_float4(ssf::fix_im(p1.x), ssf::fix_im(p1.y), ssf::fix_im(p1.z), ssf::fix_im(p1.w));
        p2 = make_float4(ssf::fix_re(p2.x), ssf::fix_re(p2.y), ssf::fix_re(p2.z), ssf::fix_re(p2.w));
        p3 = make_float4(ssf::fix_im(p3.x), ssf::fix_im(p3.y), ssf::fix_im(p3.z), ssf::fix_im(p3.w));
      }
      float4 *o = reinterpret_cast<float4 *>(out) + q;
      const int64_t ps = npix / 4;
      float4 pl[4] = {p0, p1, p2, p3};
#pragma unroll
      for (int k = 0; k < 4; ++k)
        __stcs(o + k * ps, make_float4(ssf::out_word(pl[k].x, prmt), ssf::out_word(pl[k].y, prmt),
                                       ssf::out_word(pl[k].z, prmt), ssf::out_word(pl[k].w, prmt)));
    } else {
      float4 a = __ldg(row + (__ldg(tmpl + q) - 1));
      out[q] = ssf::out_word(fix_nan ? ssf::fix_re(a.x) : a.x, prmt);
      out[npix + q] = ssf::out_word(fix_nan ? ssf::fix_im(a.y) : a.y, prmt);
      out[2 * npix + q] = ssf::out_word(fix_nan ? ssf::fix_re(a.z) : a.z, prmt);
      out[3 * npix + q] = ssf::out_word(fix_nan ? ssf::fix_im(a.w) : a.w, prmt);
    }
  }
}

__global__ void indicator_kernel(const int32_t *__restrict__ tmpl, int ndir, int64_t npix,
                                 float *__restrict__ ind) {
  // ind[d, pix] = (tmpl[pix] == d + 1)
  int64_t total = npix * ndir;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    int64_t d = i / npix, p = i - d * npix;
    ind[i] = (tmpl[p] == (int32_t)d + 1) ? 1.0f : 0.0f;
  }
}

__global__ void nan_fix_kernel(float *__restrict__ slab, int64_t ncol, int64_t npix) {
  int64_t total = ncol * 4 * npix;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    float v = slab[i];
    if (v != v) slab[i] = ((i / npix) & 1) ? 0.0f : 1.0f;
  }
}

// float32 little-endian -> big-endian in place (FITS BITPIX -32 storage order), 16 B per thread.
__global__ void bswap32_kernel(uint32_t *__restrict__ data, int64_t n) {
  const int64_t n4 = n / 4;
  uint4 *v = reinterpret_cast<uint4 *>(data);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (int64_t)gridDim.x * blockDim.x) {
    uint4 a = v[i];
    a.x = __byte_perm(a.x, 0, 0x0123);
    a.y = __byte_perm(a.y, 0, 0x0123);
    a.z = __byte_perm(a.z, 0, 0x0123);
    a.w = __byte_perm(a.w, 0, 0x0123);
    v[i] = a;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    int64_t i = n4 * 4 + threadIdx.x;
    data[i] = __byte_perm(data[i], 0, 0x0123);
  }
}

int grid_for(int64_t work, int threads, int ctas_per_sm) {
  int64_t blocks = (work + threads - 1) / threads;
  int64_t cap = (int64_t)ssf::sm_count() * ctas_per_sm;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

// Grid of the fill kernels: 512 CTAs per SM (8 are resident), each thread a few grid-strided
// passes.  A grid of exactly the resident CTAs looping to the end (round 1: 8 per SM) is a static
// partition, and a write-only stream then ends with its slowest SM: 0.90 of the copy bandwidth
// at 512^2, 0.80 at 2048^2; with 512 CTAs per SM the hardware hands out the work and the same
// kernel reaches 1.09 / 1.00 (a write-only stream is not bound by the copy figure; the probe
// scripts/stream_write_probe.cu measures 7.5 TB/s for one-pass CTAs).  One pass per thread is
// best for the four-plane phase fill (1.15 / 1.06: 64 B of stores per thread and one LUT row per
// CTA) but starts too many CTAs per byte for the plane-pair fill (0.73) and re-reads the wider
// LUT row too often for the gain fill (0.99): `one_pass` is set for the former only.
// Experiment hook SSF_FILL_CTAS: CTAs per SM for every fill, 0 = one pass per thread
// (profiles/r02_fill_grid.txt).
int fill_grid(int64_t work, bool one_pass) {
  static const int forced = [] {
    const char *e = getenv("SSF_FILL_CTAS");
    return e ? atoi(e) : -1;
  }();
  const int ctas = forced >= 0 ? forced : (one_pass ? 0 : 512);
  const int64_t blocks = (work + 255) / 256;
  if (ctas <= 0) return (int)(blocks < 0x7fffffff ? (blocks > 0 ? blocks : 1) : 0x7fffffff);
  return grid_for(work, 256, ctas);
}

}  // namespace

extern "C" int ssf_voronoi_assign(const double *px, const double *py, int ndir, int nx, int ny,
                                  int32_t *tmpl, void *stream) {
  SSF_REQUIRE(px && py && tmpl, "ssf_voronoi_assign: null pointer");
  SSF_REQUIRE(ndir >= 1 && ndir <= SSF_MAX_DIR * 8, "ssf_voronoi_assign: ndir=%d out of range", ndir);
  SSF_REQUIRE(nx >= 1 && ny >= 1, "ssf_voronoi_assign: empty image");
  int64_t npix = (int64_t)nx * ny;
  SSF_LAUNCH(assign_kernel, grid_for(npix, 256, 8), 256, 0, (cudaStream_t)stream, px, py, ndir, nx, ny, tmpl);
  SSF_LAUNCHED();
  return SSF_OK;
}

extern "C" int ssf_voronoi_lut(const double *phase, const double *amp, int64_t ncol, int ndir,
                               float *lut, void *stream) {
  SSF_REQUIRE(phase && lut, "ssf_voronoi_lut: null pointer");
  SSF_REQUIRE(ncol >= 0 && ndir >= 1, "ssf_voronoi_lut: bad sizes");
  if (ncol == 0) return SSF_OK;
  int64_t n = ncol * ndir;
  SSF_LAUNCH(lut_kernel, grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream, phase, amp, n, (float2 *)lut);
  SSF_LAUNCHED();
  return SSF_OK;
}

extern "C" int ssf_voronoi_fill(const int32_t *tmpl, const float *lut, int ndir, int64_t npix,
                                int64_t ncol, float *slab, int fix_nan, int out_flags, void *stream) {
  SSF_REQUIRE(tmpl && lut && slab, "ssf_voronoi_fill: null pointer");
  ssf::OutFmt fmt;
  SSF_REQUIRE(ssf::out_fmt(out_flags, &fmt), "ssf_voronoi_fill: unknown out_flags %d", out_flags);
  SSF_REQUIRE(ndir >= 1 && npix >= 1 && ncol >= 0, "ssf_voronoi_fill: bad sizes");
  if (ncol == 0) return SSF_OK;
  bool vec = (npix % 4 == 0) && ((uintptr_t)slab % 16 == 0) && ((uintptr_t)tmpl % 16 == 0);
  if (vec) {
    SSF_LAUNCH(fill_kernel<true>, fill_grid(npix / 4 * ncol, fmt.planes == 4), 256, 0, (cudaStream_t)stream,
               tmpl, (const float2 *)lut, ndir, npix, ncol, slab, fix_nan, fmt);
  } else {
    SSF_LAUNCH(fill_kernel<false>, fill_grid(npix * ncol, false), 256, 0, (cudaStream_t)stream,
               tmpl, (const float2 *)lut, ndir, npix, ncol, slab, fix_nan, fmt);
  }
  SSF_LAUNCHED();
  return SSF_OK;
}

extern "C" int ssf_voronoi_lut_gain(const double *phase, const double *amp, int64_t ncol, int ndir,
                                    float *lut4, void *stream) {
  SSF_REQUIRE(phase && amp && lut4, "ssf_voronoi_lut_gain: null pointer");
  SSF_REQUIRE(ncol >= 0 && ndir >= 1, "ssf_voronoi_lut_gain: bad sizes");
  SSF_REQUIRE((uintptr_t)lut4 % 16 == 0, "ssf_voronoi_lut_gain: lut4 must be 16-byte aligned");
  if (ncol == 0) return SSF_OK;
  int64_t n = ncol * ndir;
  SSF_LAUNCH(lut_gain_kernel, grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream, phase, amp, n, (float4 *)lut4);
  SSF_LAUNCHED();
  return SSF_OK;
}

extern "C" int ssf_voronoi_fill_gain(const int32_t *tmpl, const float *lut4, int ndir, int64_t npix,
                                     int64_t ncol, float *slab, int fix_nan, int out_flags, void *stream) {
  SSF_REQUIRE(tmpl && lut4 && slab, "ssf_voronoi_fill_gain: null pointer");
  ssf::OutFmt fmt;
  SSF_REQUIRE(ssf::out_fmt(out_flags, &fmt) && fmt.planes == 4,
              "ssf_voronoi_fill_gain: out_flags may only hold SSF_OUT_BIG_ENDIAN (the four planes differ)");
  SSF_REQUIRE(ndir >= 1 && npix >= 1 && ncol >= 0, "ssf_voronoi_fill_gain: bad sizes");
  SSF_REQUIRE((uintptr_t)lut4 % 16 == 0, "ssf_voronoi_fill_gain: lut4 must be 16-byte aligned");
  if (ncol == 0) return SSF_OK;
  bool vec = (npix % 4 == 0) && ((uintptr_t)slab % 16 == 0) && ((uintptr_t)tmpl % 16 == 0);
  if (vec) {
    SSF_LAUNCH(fill_gain_kernel<true>, fill_grid(npix / 4 * ncol, false), 256, 0, (cudaStream_t)stream,
               tmpl, (const float4 *)lut4, ndir, npix, ncol, slab, fix_nan, fmt.prmt);
  } else {
    SSF_LAUNCH(fill_gain_kernel<false>, fill_grid(npix * ncol, false), 256, 0, (cudaStream_t)stream,
               tmpl, (const float4 *)lut4, ndir, npix, ncol, slab, fix_nan, fmt.prmt);
  }
  SSF_LAUNCHED();
  return SSF_OK;
}

extern "C" int ssf_voronoi_indicators(const int32_t *tmpl, int ndir, int64_t npix, float *ind,
                                      void *stream) {
  SSF_REQUIRE(tmpl && ind, "ssf_voronoi_indicators: null pointer");
  SSF_LAUNCH(indicator_kernel, grid_for(npix * ndir, 256, 512), 256, 0, (cudaStream_t)stream, tmpl, ndir, npix, ind);
  SSF_LAUNCHED();
  return SSF_OK;
}

extern "C" int ssf_nan_fix(float *slab, int64_t ncol, int64_t npix, void *stream) {
  SSF_REQUIRE(slab, "ssf_nan_fix: null pointer");
  if (ncol == 0) return SSF_OK;
  SSF_LAUNCH(nan_fix_kernel, grid_for(ncol * 4 * npix, 256, 512), 256, 0, (cudaStream_t)stream, slab, ncol, npix);
  SSF_LAUNCHED();
  return SSF_OK;
}

extern "C" int ssf_bswap32(void *data, int64_t count, void *stream) {
  SSF_REQUIRE(data, "ssf_bswap32: null pointer");
  SSF_REQUIRE(((uintptr_t)data % 16) == 0, "ssf_bswap32: buffer must be 16-byte aligned");
  if (count <= 0) return SSF_OK;
  SSF_LAUNCH(bswap32_kernel, grid_for(count / 4 + 1, 256, 512), 256, 0, (cudaStream_t)stream,
             (uint32_t *)data, count);
  SSF_LAUNCHED();
  return SSF_OK;
}
