// libssf_b200 core: error string, launch accounting, and the host-buffer convenience entry
// points (allocate, H2D, kernels, D2H, free) a ctypes/cffi binding of the reference calls.
#include <stdarg.h>
#include <string.h>

#include <atomic>
#include <thread>
#include <vector>

#include "common.cuh"

namespace {
thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};
}  // namespace

namespace ssf {
void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }
int sm_count() {
#ifdef SSF_CPU_EMU
  return 148;
#else
  static std::atomic<int> cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  int n = cached[dev].load(std::memory_order_relaxed);
  if (n > 0) return n;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1) n = 148;
  cached[dev].store(n, std::memory_order_relaxed);
  return n;
#endif
}
}  // namespace ssf

extern "C" int ssf_version(void) { return 100; }
extern "C" const char *ssf_last_error(void) { return g_err; }
extern "C" uint64_t ssf_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

namespace {

// Frees every registered device allocation when it goes out of scope.
struct DeviceArena {
  void *ptrs[32];
  int n = 0;
  ~DeviceArena() {
    for (int i = 0; i < n; ++i) cudaFree(ptrs[i]);
  }
  template <typename T>
  int alloc(T **p, size_t count) {
    SSF_CUDA(cudaMalloc((void **)p, sizeof(T) * (count ? count : 1)));
    ptrs[n++] = *p;
    return SSF_OK;
  }
};

#define SSF_TRY(call)           \
  do {                          \
    int rc_ = (call);           \
    if (rc_ != SSF_OK) return rc_; \
  } while (0)

}  // namespace

extern "C" int ssf_copy_rows_d2h(void *dst_host, int64_t dst_pitch, const void *src_dev, int64_t src_pitch,
                                 int64_t width_bytes, int64_t rows, void *stream) {
  SSF_REQUIRE(dst_host && src_dev, "ssf_copy_rows_d2h: null pointer");
  SSF_REQUIRE(width_bytes >= 0 && rows >= 0 && dst_pitch >= width_bytes && src_pitch >= width_bytes,
              "ssf_copy_rows_d2h: pitches must be >= width");
  if (rows == 0 || width_bytes == 0) return SSF_OK;
  if (dst_pitch == width_bytes && src_pitch == width_bytes) {
    SSF_CUDA(cudaMemcpyAsync(dst_host, src_dev, (size_t)(width_bytes * rows), cudaMemcpyDeviceToHost,
                             (cudaStream_t)stream));
  } else {
    SSF_CUDA(cudaMemcpy2DAsync(dst_host, (size_t)dst_pitch, src_dev, (size_t)src_pitch, (size_t)width_bytes,
                               (size_t)rows, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  }
  return SSF_OK;
}

// Host-side placement of a staged slab into the (memory-mapped) FITS cube: column c of `src`
// (src_bytes) goes to dst + c * copies * src_bytes, `copies` times back to back.  Threads take
// contiguous runs of columns; memcpy switches to streaming stores for copies of this size, and
// the first touch of the pages of a fresh tmpfs / page-cache mapping (the dominant cost) runs in
// parallel -- unlike pwrite calls on one file, which serialise on the inode lock.
extern "C" int ssf_host_place(void *dst, const void *src, int64_t ncol, int64_t src_bytes, int copies,
                              int threads) {
  SSF_REQUIRE(dst && src, "ssf_host_place: null pointer");
  SSF_REQUIRE(ncol >= 0 && src_bytes >= 0 && copies >= 1 && copies <= 4, "ssf_host_place: bad sizes");
  if (ncol == 0 || src_bytes == 0) return SSF_OK;
  const int64_t total = ncol * src_bytes;
  int nt = threads < 1 ? 1 : threads > 64 ? 64 : threads;
  // at least ~4 MiB per thread
  if ((int64_t)nt > (total >> 22) + 1) nt = (int)((total >> 22) + 1);
  char *d = static_cast<char *>(dst);
  const char *s = static_cast<const char *>(src);
  // a thread's share is a byte range of the SOURCE (so that few very large columns, e.g. 2048^2
  // SKA-Low planes, still spread over all threads), mapped piecewise onto the destination
  auto work = [=](int64_t lo, int64_t hi) {
    const int64_t piece = 1 << 21;
    for (int64_t p = lo; p < hi;) {
      const int64_t c = p / src_bytes, within = p - c * src_bytes;
      int64_t n = src_bytes - within;
      if (n > hi - p) n = hi - p;
      if (n > piece) n = piece;
      for (int k = 0; k < copies; ++k) memcpy(d + (c * copies + k) * src_bytes + within, s + p, (size_t)n);
      p += n;
    }
  };
  if (nt == 1) {
    work(0, total);
    return SSF_OK;
  }
  std::vector<std::thread> pool;
  pool.reserve(nt);
  try {
    for (int i = 0; i < nt; ++i) {
      const int64_t lo = total * i / nt / 64 * 64, hi = i + 1 == nt ? total : total * (i + 1) / nt / 64 * 64;
      pool.emplace_back(work, lo, hi);
    }
  } catch (...) {
    for (auto &t : pool) t.join();
    ssf::set_error("ssf_host_place: could not start a worker thread");
    return SSF_EINVAL;
  }
  for (auto &t : pool) t.join();
  return SSF_OK;
}

extern "C" int ssf_voronoi_cube_host(const int32_t *tmpl_host, const double *phase_host, int ndir,
                                     int ny, int nx, int64_t ncol, double smooth_pix,
                                     float *slab_host) {
  SSF_REQUIRE(tmpl_host && phase_host && slab_host, "ssf_voronoi_cube_host: null pointer");
  SSF_REQUIRE(ndir >= 1 && ndir <= SSF_MAX_DIR && nx >= 1 && ny >= 1 && ncol >= 0,
              "ssf_voronoi_cube_host: bad sizes");
  if (ncol == 0) return SSF_OK;
  const int64_t npix = (int64_t)nx * ny;
  DeviceArena arena;
  int32_t *d_tmpl;
  double *d_phase;
  float *d_lut, *d_slab;
  SSF_TRY(arena.alloc(&d_tmpl, npix));
  SSF_TRY(arena.alloc(&d_phase, ncol * ndir));
  SSF_TRY(arena.alloc(&d_lut, ncol * ndir * 2));
  SSF_TRY(arena.alloc(&d_slab, ncol * 4 * npix));
  SSF_CUDA(cudaMemcpy(d_tmpl, tmpl_host, sizeof(int32_t) * npix, cudaMemcpyHostToDevice));
  SSF_CUDA(cudaMemcpy(d_phase, phase_host, sizeof(double) * ncol * ndir, cudaMemcpyHostToDevice));
  SSF_TRY(ssf_voronoi_lut(d_phase, nullptr, ncol, ndir, d_lut, nullptr));
  if (smooth_pix > 0.0) {
    // NaN phases cannot go through the linear (contraction) form: 0 * NaN would spread them
    // beyond the filter radius.  The host entry point always takes the direct form.
    float *d_scratch;
    SSF_TRY(arena.alloc(&d_scratch, ncol * 4 * npix));
    SSF_TRY(ssf_voronoi_fill(d_tmpl, d_lut, ndir, npix, ncol, d_slab, 0, 0, nullptr));
    SSF_TRY(ssf_gauss_smooth(d_slab, ncol * 4, ny, nx, smooth_pix, d_scratch, nullptr));
    SSF_TRY(ssf_nan_fix(d_slab, ncol, npix, nullptr));
  } else {
    SSF_TRY(ssf_voronoi_fill(d_tmpl, d_lut, ndir, npix, ncol, d_slab, 1, 0, nullptr));
  }
  SSF_CUDA(cudaMemcpy(slab_host, d_slab, sizeof(float) * ncol * 4 * npix, cudaMemcpyDeviceToHost));
  return SSF_OK;
}

extern "C" int ssf_kl_cube_host(const double *x_host, const double *y_host, int nx, int ny,
                                const double *pp_host, int ndir, const double *white_host,
                                int64_t ncol, double r0, double beta, double pinv_cut,
                                float *slab_host) {
  SSF_REQUIRE(x_host && y_host && pp_host && white_host && slab_host, "ssf_kl_cube_host: null pointer");
  SSF_REQUIRE(ndir >= 2 && ndir <= SSF_MAX_DIR && nx >= 1 && ny >= 1 && ncol >= 0,
              "ssf_kl_cube_host: bad sizes");
  if (ncol == 0) return SSF_OK;
  const int64_t npix = (int64_t)nx * ny;
  const int ldg = (ndir + 3) / 4 * 4;
  DeviceArena arena;
  double *d_x, *d_y, *d_pp, *d_white, *d_C, *d_pinv, *d_U, *d_sv;
  uint64_t *d_mask;
  float *d_Bt, *d_G, *d_slab;
  SSF_TRY(arena.alloc(&d_x, nx));
  SSF_TRY(arena.alloc(&d_y, ny));
  SSF_TRY(arena.alloc(&d_pp, 2 * ndir));
  SSF_TRY(arena.alloc(&d_white, ncol * ndir));
  SSF_TRY(arena.alloc(&d_C, (size_t)ndir * ndir));
  SSF_TRY(arena.alloc(&d_pinv, (size_t)ndir * ndir));
  SSF_TRY(arena.alloc(&d_U, (size_t)ndir * ndir));
  SSF_TRY(arena.alloc(&d_sv, ndir));
  SSF_TRY(arena.alloc(&d_mask, 2));
  SSF_TRY(arena.alloc(&d_Bt, (size_t)ndir * npix));
  SSF_TRY(arena.alloc(&d_G, ncol * ldg));
  SSF_TRY(arena.alloc(&d_slab, ncol * 4 * npix));
  uint64_t mask[2] = {0, 0};
  for (int d = 0; d < ndir; ++d) mask[d / 64] |= 1ull << (d % 64);
  SSF_CUDA(cudaMemcpy(d_x, x_host, sizeof(double) * nx, cudaMemcpyHostToDevice));
  SSF_CUDA(cudaMemcpy(d_y, y_host, sizeof(double) * ny, cudaMemcpyHostToDevice));
  SSF_CUDA(cudaMemcpy(d_pp, pp_host, sizeof(double) * 2 * ndir, cudaMemcpyHostToDevice));
  SSF_CUDA(cudaMemcpy(d_white, white_host, sizeof(double) * ncol * ndir, cudaMemcpyHostToDevice));
  SSF_CUDA(cudaMemcpy(d_mask, mask, sizeof(mask), cudaMemcpyHostToDevice));
  SSF_TRY(ssf_kl_decompose(d_pp, ndir, d_mask, 1, r0, beta, pinv_cut, d_C, d_pinv, d_U, d_sv, nullptr));
  SSF_TRY(ssf_kl_basis(d_x, d_y, nx, ny, d_pp, ndir, d_pinv, r0, beta, d_Bt, nullptr));
  SSF_TRY(ssf_kl_coeffs(d_white, ncol, ndir, d_C, d_G, ldg, nullptr));
  SSF_TRY(ssf_kl_eval(d_Bt, d_G, ndir, ldg, npix, ncol, d_slab, -1, 0, nullptr));
  SSF_CUDA(cudaMemcpy(slab_host, d_slab, sizeof(float) * ncol * 4 * npix, cudaMemcpyDeviceToHost));
  return SSF_OK;
}
