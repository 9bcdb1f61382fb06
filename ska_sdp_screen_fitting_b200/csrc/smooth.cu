// Separable Gaussian smoothing (ndimage.gaussian_filter semantics, screen.py:353-362):
// y pass then x pass, 'reflect' (half-sample symmetric) boundary, radius int(4 sigma + .5),
// float32 storage between passes.  Two x-pass variants: warp-shuffle halo exchange for
// small radii (each warp loads 32 consecutive pixels once, neighbours come from shuffles),
// shared-memory row tiles otherwise.
#include "common.cuh"

#include <mutex>

namespace {

constexpr int MAX_RADIUS = 1024;
// The taps travel as a per-call device buffer (stream-ordered allocation), not a __constant__
// symbol: two calls with different sigma on different streams or threads must not share state.
// Every block copies them to shared memory first; all lanes then read the same word (broadcast).

// scipy 'reflect': (d c b a | a b c d | d c b a), any distance outside the line.
__device__ __forceinline__ int reflect(int i, int n) {
  int period = 2 * n;
  i %= period;
  if (i < 0) i += period;
  return i >= n ? period - 1 - i : i;
}

// y pass: thread = (x, y) output; column tile staged in shared memory with its halo.
// Block 32 (x) by TY (y) threads; smem [(TY + 2r)][32].
template <int TY>
__global__ void ypass_kernel(const float *__restrict__ in, float *__restrict__ out, int ny, int nx,
                             int r, const float *__restrict__ taps) {
  SSF_DYN_SMEM(float, c_w);                          // [2r + 1] taps, then the tile
  float *tile = c_w + 2 * r + 1;
  for (int k = threadIdx.y * 32 + threadIdx.x; k <= 2 * r; k += 32 * TY) c_w[k] = taps[k];
  const int64_t img = blockIdx.z;
  const float *src = in + img * (int64_t)ny * nx;
  float *dst = out + img * (int64_t)ny * nx;
  const int x = blockIdx.x * 32 + threadIdx.x;
  const int y0 = blockIdx.y * TY;
  const int rows = TY + 2 * r;
  for (int j = threadIdx.y; j < rows; j += TY) {
    int yy = reflect(y0 + j - r, ny);
    tile[j * 32 + threadIdx.x] = x < nx ? src[(int64_t)yy * nx + x] : 0.0f;
  }
  __syncthreads();
  const int y = y0 + threadIdx.y;
  if (x < nx && y < ny) {
    float acc = 0.0f;
    for (int k = 0; k <= 2 * r; ++k) acc = fmaf(c_w[k], tile[(threadIdx.y + k) * 32 + threadIdx.x], acc);
    dst[(int64_t)y * nx + x] = acc;
  }
}

// x pass, shared-memory variant: block = one row segment of TX outputs.
template <int TX>
__global__ void xpass_smem_kernel(const float *__restrict__ in, float *__restrict__ out, int ny,
                                  int nx, int r, const float *__restrict__ taps) {
  SSF_DYN_SMEM(float, c_w);                          // [2r + 1] taps, then the row segment
  float *row = c_w + 2 * r + 1;
  for (int k = threadIdx.x; k <= 2 * r; k += TX) c_w[k] = taps[k];
  const int64_t img = blockIdx.z;
  const int y = blockIdx.y;
  const float *src = in + (img * ny + y) * (int64_t)nx;
  float *dst = out + (img * ny + y) * (int64_t)nx;
  const int x0 = blockIdx.x * TX;
  for (int j = threadIdx.x; j < TX + 2 * r; j += TX) row[j] = src[reflect(x0 + j - r, nx)];
  __syncthreads();
  const int x = x0 + threadIdx.x;
  if (x < nx) {
    float acc = 0.0f;
    for (int k = 0; k <= 2 * r; ++k) acc = fmaf(c_w[k], row[threadIdx.x + k], acc);
    dst[x] = acc;
  }
}

// x pass, warp-shuffle halo exchange (r <= 8): a warp owns 32 - 2r outputs of one row; it
// loads 32 consecutive (reflected) pixels, one per lane, and every tap is a shuffle.
__global__ void xpass_shfl_kernel(const float *__restrict__ in, float *__restrict__ out, int ny,
                                  int nx, int r, const float *__restrict__ taps) {
  __shared__ float c_w[17];                          // r <= 8
  if (threadIdx.x <= 2 * r) c_w[threadIdx.x] = taps[threadIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int span = 32 - 2 * r;
  const int64_t img = blockIdx.z;
  const int y = blockIdx.y;
  const int seg = blockIdx.x * (blockDim.x >> 5) + warp;
  const int x0 = seg * span;
  if (x0 >= nx) return;  // whole warp exits together
  const float *src = in + (img * ny + y) * (int64_t)nx;
  float *dst = out + (img * ny + y) * (int64_t)nx;
  float v = src[reflect(x0 + lane - r, nx)];
  float acc = 0.0f;
  for (int k = 0; k <= 2 * r; ++k) {
    float t = __shfl_sync(0xffffffffu, v, (lane + k) & 31);
    acc = fmaf(c_w[k], t, acc);
  }
  const int x = x0 + lane;
  if (lane < span && x < nx) dst[x] = acc;
}

// ---- register-blocked passes (the default while the tile fits shared memory) ----------------
// Every thread produces BLK consecutive outputs ALONG the filter direction from a sliding window
// of BLK registers: per tap one broadcast load of the weight, one load of the sample entering the
// window and BLK FMAs -- the tap-per-output kernels above pay one shared-memory load per FMA
// (0.11 of the HBM bound at sigma = 5 px).  The accumulation order (taps ascending, one FMA each)
// is the same as theirs, so the results are identical bit for bit.  Taps beyond 2r are never
// executed (a zero weight times a NaN sample would be a NaN).
constexpr int BLK = 8;

// window step for tap k (compile-time slot u = k mod 8): accumulate, then replace the oldest sample
#define SSF_SMOOTH_TAPS(LOAD)                                                          \
  float acc[BLK], win[BLK];                                                            \
  _Pragma("unroll") for (int j = 0; j < BLK; ++j) {                                    \
    acc[j] = 0.0f;                                                                     \
    win[j] = LOAD(j);                                                                  \
  }                                                                                    \
  const int ntap = 2 * r + 1;                                                          \
  int k0 = 0;                                                                          \
  for (; k0 + BLK <= ntap; k0 += BLK) {                                                \
    _Pragma("unroll") for (int u = 0; u < BLK; ++u) {                                  \
      const float w = c_w[k0 + u];                                                     \
      _Pragma("unroll") for (int j = 0; j < BLK; ++j) acc[j] = fmaf(w, win[(j + u) & (BLK - 1)], acc[j]); \
      win[u] = LOAD(k0 + u + BLK);                                                     \
    }                                                                                  \
  }                                                                                    \
  _Pragma("unroll") for (int u = 0; u < BLK; ++u) {                                    \
    if (k0 + u < ntap) {                                                               \
      const float w = c_w[k0 + u];                                                     \
      _Pragma("unroll") for (int j = 0; j < BLK; ++j) acc[j] = fmaf(w, win[(j + u) & (BLK - 1)], acc[j]); \
      win[u] = LOAD(k0 + u + BLK);                                                     \
    }                                                                                  \
  }

// y pass: block 32 (x) by YT (y groups); thread (tx, ty) owns rows y0 + 8 ty .. + 7 of column x.
// smem: taps[2r + 1], tile[(8 YT + 2r + 1)][32] (the last row is read by the final window
// refill and never used).
template <int YT>
__global__ void ypass_blocked_kernel(const float *__restrict__ in, float *__restrict__ out, int ny, int nx, int r,
                                     const float *__restrict__ taps) {
  SSF_DYN_SMEM(float, c_w);
  float *tile = c_w + 2 * r + 1;
  const int tx = threadIdx.x, ty = threadIdx.y;
  for (int k = ty * 32 + tx; k <= 2 * r; k += 32 * YT) c_w[k] = taps[k];
  const int64_t img = blockIdx.z;
  const float *src = in + img * (int64_t)ny * nx;
  float *dst = out + img * (int64_t)ny * nx;
  const int x = blockIdx.x * 32 + tx;
  const int y0 = blockIdx.y * (YT * BLK);
  const int rows = YT * BLK + 2 * r + 1;
  for (int j = ty; j < rows; j += YT) {
    const int yy = reflect(y0 + j - r, ny);
    tile[j * 32 + tx] = (x < nx && j < rows - 1) ? src[(int64_t)yy * nx + x] : 0.0f;
  }
  __syncthreads();
  const int base = ty * BLK;
#define SSF_LOAD_Y(i) tile[(base + (i)) * 32 + tx]
  SSF_SMOOTH_TAPS(SSF_LOAD_Y)
#undef SSF_LOAD_Y
  if (x < nx) {
#pragma unroll
    for (int j = 0; j < BLK; ++j) {
      const int y = y0 + base + j;
      if (y < ny) dst[(int64_t)y * nx + x] = acc[j];
    }
  }
}

// x pass: block 32 (rows, lane = row so that the lanes of a warp walk different rows of an
// odd-pitched tile: no bank conflicts) by XT (x groups); thread (lane, tg) owns columns
// x0 + 8 tg .. + 7 of its row.  Results go back through the tile so that the global stores are
// coalesced along x; `fix` (0 none, 1 + plane parity) fuses the NaN fix of screen.py:364-378.
template <int XT>
__global__ void xpass_blocked_kernel(const float *__restrict__ in, float *__restrict__ out, int ny, int nx, int r,
                                     const float *__restrict__ taps, int fix_nan) {
  SSF_DYN_SMEM(float, c_w);
  float *tile = c_w + 2 * r + 1;
  const int lane = threadIdx.x, tg = threadIdx.y, tid = tg * 32 + lane;
  for (int k = tid; k <= 2 * r; k += 32 * XT) c_w[k] = taps[k];
  const int64_t img = blockIdx.z;
  const float *src = in + img * (int64_t)ny * nx;
  float *dst = out + img * (int64_t)ny * nx;
  const int x0 = blockIdx.x * (XT * BLK), y0 = blockIdx.y * 32;
  const int width = XT * BLK + 2 * r + 1;            // last column: read by the final refill, never used
  const int pitch = width | 1;
  for (int i = tid; i < 32 * width; i += 32 * XT) {
    const int rr = i / width, cc = i - rr * width;
    const int y = y0 + rr;
    tile[rr * pitch + cc] = (y < ny && cc < width - 1) ? src[(int64_t)y * nx + reflect(x0 + cc - r, nx)] : 0.0f;
  }
  __syncthreads();
  const int base = tg * BLK;
  const float *row = tile + lane * pitch;
#define SSF_LOAD_X(i) row[base + (i)]
  SSF_SMOOTH_TAPS(SSF_LOAD_X)
#undef SSF_LOAD_X
  __syncthreads();                                   // everybody has read its inputs: reuse the tile for the outputs
#pragma unroll
  for (int j = 0; j < BLK; ++j) tile[lane * pitch + base + j] = acc[j];
  __syncthreads();
  const float nan_to = (img & 1) ? 0.0f : 1.0f;      // planes 0, 2 (real) -> 1; planes 1, 3 (imaginary) -> 0
  for (int i = tid; i < 32 * XT * BLK; i += 32 * XT) {
    const int rr = i / (XT * BLK), cc = i - rr * (XT * BLK);
    const int y = y0 + rr, x = x0 + cc;
    if (y < ny && x < nx) {
      float v = tile[rr * pitch + cc];
      if (fix_nan && v != v) v = nan_to;
      dst[(int64_t)y * nx + x] = v;
    }
  }
}
#undef SSF_SMOOTH_TAPS

}  // namespace

namespace {
int gauss_smooth_impl(const char *who, float *images, int64_t nimg, int ny, int nx, double sigma, float *scratch,
                      int fix_nan, cudaStream_t stream) {
  SSF_REQUIRE(images && scratch, "%s: null pointer", who);
  SSF_REQUIRE(ny >= 1 && nx >= 1 && nimg >= 0, "%s: bad sizes", who);
  SSF_REQUIRE(sigma > 0.0, "%s: sigma must be > 0", who);
  SSF_REQUIRE(!fix_nan || nimg % 4 == 0, "%s: the NaN fix needs whole columns of 4 planes", who);
  if (nimg == 0) return SSF_OK;
  const int r = (int)(4.0 * sigma + 0.5);
  SSF_REQUIRE(r <= MAX_RADIUS, "%s: radius %d exceeds %d", who, r, MAX_RADIUS);
  // weights in float64 exactly as scipy's _gaussian_kernel1d, rounded once to float32
  static thread_local float hw[2 * MAX_RADIUS + 1];
  double sum = 0.0;
  static thread_local double dw[2 * MAX_RADIUS + 1];
  for (int k = -r; k <= r; ++k) {
    dw[k + r] = exp(-0.5 / (sigma * sigma) * (double)(k * k));
    sum += dw[k + r];
  }
  for (int k = 0; k <= 2 * r; ++k) hw[k] = (float)(dw[k] / sum);
  float *taps = nullptr;
  SSF_CUDA(cudaMallocAsync(reinterpret_cast<void **>(&taps), sizeof(float) * (2 * r + 1), stream));
  struct Free {
    float *p;
    cudaStream_t s;
    ~Free() { cudaFreeAsync(p, s); }
  } free_taps{taps, stream};
  // pageable source: the runtime stages it before the call returns, so hw may be reused at once
  SSF_CUDA(cudaMemcpyAsync(taps, hw, sizeof(float) * (2 * r + 1), cudaMemcpyHostToDevice, stream));
  // register-blocked passes while their tiles fit (radius <= ~190 pixels), else the tap-per-output ones
  constexpr int YT = 16, XT = 16;
  const size_t smem_y = sizeof(float) * ((size_t)(YT * BLK + 2 * r + 1) * 32 + 2 * r + 1);
  const size_t smem_x = sizeof(float) * ((size_t)32 * ((XT * BLK + 2 * r + 1) | 1) + 2 * r + 1);
  const bool blocked = smem_y <= 100 * 1024 && smem_x <= 100 * 1024;
  bool fixed = false;
  // gridDim.z is limited to 65535 images per launch (kept a multiple of 4: plane parity = image parity)
  for (int64_t i0 = 0; i0 < nimg; i0 += 65532) {
    int nz = (int)((nimg - i0) < 65532 ? (nimg - i0) : 65532);
    float *img = images + i0 * (int64_t)ny * nx;
    float *tmp = scratch + i0 * (int64_t)ny * nx;
    if (blocked) {
      {
        dim3 grid(ssf::div_up(nx, 32), ssf::div_up(ny, YT * BLK), nz), block(32, YT);
        if (smem_y > 48 * 1024)
          SSF_CUDA(cudaFuncSetAttribute(ypass_blocked_kernel<YT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_y));
        SSF_LAUNCH(ypass_blocked_kernel<YT>, grid, block, smem_y, stream, img, tmp, ny, nx, r, taps);
        SSF_LAUNCHED();
      }
      {
        dim3 grid(ssf::div_up(nx, XT * BLK), ssf::div_up(ny, 32), nz), block(32, XT);
        if (smem_x > 48 * 1024)
          SSF_CUDA(cudaFuncSetAttribute(xpass_blocked_kernel<XT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_x));
        SSF_LAUNCH(xpass_blocked_kernel<XT>, grid, block, smem_x, stream, tmp, img, ny, nx, r, taps, fix_nan);
        SSF_LAUNCHED();
        fixed = true;
      }
      continue;
    }
    {
      constexpr int TY = 16;
      dim3 grid(ssf::div_up(nx, 32), ssf::div_up(ny, TY), nz), block(32, TY);
      size_t smem = sizeof(float) * (32 * (TY + 2 * r) + 2 * r + 1);
      SSF_REQUIRE(smem <= 200 * 1024, "%s: radius too large for the y tile", who);
      if (smem > 48 * 1024)
        SSF_CUDA(cudaFuncSetAttribute(ypass_kernel<TY>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem));
      SSF_LAUNCH(ypass_kernel<TY>, grid, block, smem, stream, img, tmp, ny, nx, r, taps);
      SSF_LAUNCHED();
    }
    if (r <= 8) {
      const int span = 32 - 2 * r, warps = 8;
      dim3 grid(ssf::div_up(ssf::div_up(nx, span), warps), ny, nz);
      SSF_LAUNCH(xpass_shfl_kernel, grid, warps * 32, 0, stream, tmp, img, ny, nx, r, taps);
      SSF_LAUNCHED();
    } else {
      constexpr int TX = 256;
      dim3 grid(ssf::div_up(nx, TX), ny, nz);
      size_t smem = sizeof(float) * (TX + 2 * r + 2 * r + 1);
      SSF_LAUNCH(xpass_smem_kernel<TX>, grid, TX, smem, stream, tmp, img, ny, nx, r, taps);
      SSF_LAUNCHED();
    }
  }
  if (fix_nan && !fixed) return ssf_nan_fix(images, nimg / 4, (int64_t)ny * nx, (void *)stream);
  return SSF_OK;
}
}  // namespace

extern "C" int ssf_gauss_smooth(float *images, int64_t nimg, int ny, int nx, double sigma,
                                float *scratch, void *stream) {
  return gauss_smooth_impl("ssf_gauss_smooth", images, nimg, ny, nx, sigma, scratch, 0, (cudaStream_t)stream);
}

extern "C" int ssf_gauss_smooth_slab(float *slab, int64_t ncol, int ny, int nx, double sigma, float *scratch,
                                     int fix_nan, void *stream) {
  SSF_REQUIRE(ncol >= 0, "ssf_gauss_smooth_slab: bad sizes");
  return gauss_smooth_impl("ssf_gauss_smooth_slab", slab, ncol * 4, ny, nx, sigma, scratch, fix_nan,
                           (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// ssf_gauss_smooth_static: the same filter for callers that may not allocate (CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The intermediate image of the two passes and the taps
// live in ONE file-scope device array of fixed size, held by every GPU that loads the library
// until the process exits: SSF_SMOOTH_STATIC_MAX_FLOATS floats (16 MiB) of intermediate, then the
// taps.  All calls share it: it keeps the last call's contents, and two calls that overlap on the
// GPU (different streams, or different host threads) race on it -- callers serialise them.
// Nothing in it has to start at zero: the first pass writes every intermediate value the second
// pass reads, and the taps [0, 2r] are rewritten by every call before the first pass runs.
// The existing entry points above are untouched; this one only launches the same kernels.
namespace {
constexpr int64_t STATIC_MAX_FLOATS = SSF_SMOOTH_STATIC_MAX_FLOATS;
constexpr int STATIC_MAX_DEVICES = 64;
static __device__ __align__(256) float g_static_scratch[STATIC_MAX_FLOATS + 2 * MAX_RADIUS + 64];

// In host code `g_static_scratch` is a stand-in for the symbol, not a device address: the address
// comes from cudaGetSymbolAddress, once per device, with that device current.
int static_scratch_address(float **out) {
  static std::mutex lock;
  static float *address[STATIC_MAX_DEVICES] = {};
  int device = -1;
  SSF_CUDA(cudaGetDevice(&device));
  SSF_REQUIRE(device >= 0 && device < STATIC_MAX_DEVICES, "ssf_gauss_smooth_static: device %d out of range", device);
  std::lock_guard<std::mutex> hold(lock);
  if (!address[device]) {
    float *ptr = nullptr;
    SSF_CUDA(cudaGetSymbolAddress(reinterpret_cast<void **>(&ptr), g_static_scratch));
    SSF_REQUIRE(ptr != nullptr, "ssf_gauss_smooth_static: no address for the scratch symbol on device %d", device);
    address[device] = ptr;
  }
  *out = address[device];
  return SSF_OK;
}
}  // namespace

extern "C" int64_t ssf_gauss_smooth_static_max_floats(void) { return STATIC_MAX_FLOATS; }

extern "C" int ssf_gauss_smooth_static(float *images, int64_t nimg, int ny, int nx, double sigma, void *stream_) {
  const char *who = "ssf_gauss_smooth_static";
  cudaStream_t stream = (cudaStream_t)stream_;
  SSF_REQUIRE(images, "%s: null pointer", who);
  SSF_REQUIRE(ny >= 1 && nx >= 1 && nimg >= 0, "%s: bad sizes", who);
  SSF_REQUIRE(sigma > 0.0, "%s: sigma must be > 0", who);
  SSF_REQUIRE(nimg <= STATIC_MAX_FLOATS / ((int64_t)ny * nx),
              "%s: %lld images of %d x %d need %lld floats of scratch, the static array holds %lld "
              "(use ssf_gauss_smooth with caller-owned scratch)",
              who, (long long)nimg, ny, nx, (long long)nimg * ny * nx, (long long)STATIC_MAX_FLOATS);
  if (nimg == 0) return SSF_OK;
  const int r = (int)(4.0 * sigma + 0.5);
  SSF_REQUIRE(r <= MAX_RADIUS, "%s: radius %d exceeds %d", who, r, MAX_RADIUS);
  float *scratch = nullptr;
  if (int rc = static_scratch_address(&scratch)) return rc;
  float *taps = scratch + STATIC_MAX_FLOATS;          // 256-byte aligned: STATIC_MAX_FLOATS is a multiple of 64
  // weights as in gauss_smooth_impl: float64 like scipy's _gaussian_kernel1d, rounded once to float32
  static thread_local float hw[2 * MAX_RADIUS + 1];
  static thread_local double dw[2 * MAX_RADIUS + 1];
  double sum = 0.0;
  for (int k = -r; k <= r; ++k) {
    dw[k + r] = exp(-0.5 / (sigma * sigma) * (double)(k * k));
    sum += dw[k + r];
  }
  for (int k = 0; k <= 2 * r; ++k) hw[k] = (float)(dw[k] / sum);
  // pageable source: staged by the runtime before the call returns (no allocation of ours)
  SSF_CUDA(cudaMemcpyAsync(taps, hw, sizeof(float) * (2 * r + 1), cudaMemcpyHostToDevice, stream));
  constexpr int YT = 16, XT = 16;
  const size_t smem_y = sizeof(float) * ((size_t)(YT * BLK + 2 * r + 1) * 32 + 2 * r + 1);
  const size_t smem_x = sizeof(float) * ((size_t)32 * ((XT * BLK + 2 * r + 1) | 1) + 2 * r + 1);
  const bool blocked = smem_y <= 100 * 1024 && smem_x <= 100 * 1024;
  for (int64_t i0 = 0; i0 < nimg; i0 += 65532) {
    const int nz = (int)((nimg - i0) < 65532 ? (nimg - i0) : 65532);
    float *img = images + i0 * (int64_t)ny * nx;
    float *tmp = scratch + i0 * (int64_t)ny * nx;
    if (blocked) {
      dim3 gy(ssf::div_up(nx, 32), ssf::div_up(ny, YT * BLK), nz), by(32, YT);
      if (smem_y > 48 * 1024)
        SSF_CUDA(cudaFuncSetAttribute(ypass_blocked_kernel<YT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_y));
      SSF_LAUNCH(ypass_blocked_kernel<YT>, gy, by, smem_y, stream, img, tmp, ny, nx, r, taps);
      SSF_LAUNCHED();
      dim3 gx(ssf::div_up(nx, XT * BLK), ssf::div_up(ny, 32), nz), bx(32, XT);
      if (smem_x > 48 * 1024)
        SSF_CUDA(cudaFuncSetAttribute(xpass_blocked_kernel<XT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_x));
      SSF_LAUNCH(xpass_blocked_kernel<XT>, gx, bx, smem_x, stream, tmp, img, ny, nx, r, taps, 0);
      SSF_LAUNCHED();
      continue;
    }
    {
      constexpr int TY = 16;
      dim3 grid(ssf::div_up(nx, 32), ssf::div_up(ny, TY), nz), block(32, TY);
      const size_t smem = sizeof(float) * (32 * (TY + 2 * r) + 2 * r + 1);
      SSF_REQUIRE(smem <= 200 * 1024, "%s: radius too large for the y tile", who);
      if (smem > 48 * 1024)
        SSF_CUDA(cudaFuncSetAttribute(ypass_kernel<TY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      SSF_LAUNCH(ypass_kernel<TY>, grid, block, smem, stream, img, tmp, ny, nx, r, taps);
      SSF_LAUNCHED();
    }
    {
      constexpr int TX = 256;
      dim3 grid(ssf::div_up(nx, TX), ny, nz);
      const size_t smem = sizeof(float) * (TX + 2 * r + 2 * r + 1);
      SSF_LAUNCH(xpass_smem_kernel<TX>, grid, TX, smem, stream, tmp, img, ny, nx, r, taps);
      SSF_LAUNCHED();
    }
  }
  return SSF_OK;
}
