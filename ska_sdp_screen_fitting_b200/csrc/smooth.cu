// Separable Gaussian smoothing (ndimage.gaussian_filter semantics, screen.py:353-362):
// y pass then x pass, 'reflect' (half-sample symmetric) boundary, radius int(4 sigma + .5),
// float32 storage between passes.  Two x-pass variants: warp-shuffle halo exchange for
// small radii (each warp loads 32 consecutive pixels once, neighbours come from shuffles),
// shared-memory row tiles otherwise.
#include "common.cuh"

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

}  // namespace

extern "C" int ssf_gauss_smooth(float *images, int64_t nimg, int ny, int nx, double sigma,
                                float *scratch, void *stream_) {
  SSF_REQUIRE(images && scratch, "ssf_gauss_smooth: null pointer");
  SSF_REQUIRE(ny >= 1 && nx >= 1 && nimg >= 0, "ssf_gauss_smooth: bad sizes");
  SSF_REQUIRE(sigma > 0.0, "ssf_gauss_smooth: sigma must be > 0");
  if (nimg == 0) return SSF_OK;
  cudaStream_t stream = (cudaStream_t)stream_;
  const int r = (int)(4.0 * sigma + 0.5);
  SSF_REQUIRE(r <= MAX_RADIUS, "ssf_gauss_smooth: radius %d exceeds %d", r, MAX_RADIUS);
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
  // gridDim.z is limited to 65535 images per launch
  for (int64_t i0 = 0; i0 < nimg; i0 += 65535) {
    int nz = (int)((nimg - i0) < 65535 ? (nimg - i0) : 65535);
    float *img = images + i0 * (int64_t)ny * nx;
    float *tmp = scratch + i0 * (int64_t)ny * nx;
    {
      constexpr int TY = 16;
      dim3 grid(ssf::div_up(nx, 32), ssf::div_up(ny, TY), nz), block(32, TY);
      size_t smem = sizeof(float) * (32 * (TY + 2 * r) + 2 * r + 1);
      SSF_REQUIRE(smem <= 200 * 1024, "ssf_gauss_smooth: radius too large for the y tile");
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
  return SSF_OK;
}
