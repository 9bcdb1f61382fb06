// Write-only bandwidth of the slab layout [col][plane][plane_bytes] as a function of the plane
// size, the number of planes a thread writes at once and the order of the stores -- no reads, no
// arithmetic: what the memory system itself gives the cube kernels' store pattern.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scripts/_stream_write_probe scripts/stream_write_probe.cu
//   scripts/_stream_write_probe [total GiB, default 23]
//
// Modes: 0  grid-stride, thread stores one float4 to each of the S planes (the fill kernel's order)
//        1  as 0, but a CTA covers CH consecutive 4 KB pieces of ONE plane before it moves to the next
//           plane (16 KB contiguous per CTA and plane)
//        2  plane-major: the whole grid sweeps plane 0 of a group of columns, then plane 1, ...
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x)                                                                              \
  do {                                                                                     \
    cudaError_t e = (x);                                                                   \
    if (e != cudaSuccess) {                                                                \
      fprintf(stderr, "%s: %s (line %d)\n", #x, cudaGetErrorString(e), __LINE__);          \
      exit(1);                                                                             \
    }                                                                                      \
  } while (0)

template <int MODE, int CH>
__global__ void __launch_bounds__(256) probe(float4 *__restrict__ out, int64_t plane_vec, int planes, int64_t ncol) {
  const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
  if (MODE == 0) {
    const int64_t total = plane_vec * ncol;
    for (int64_t w = (int64_t)blockIdx.x * 256 + threadIdx.x; w < total; w += (int64_t)gridDim.x * 256) {
      const int64_t col = w / plane_vec, q = w - col * plane_vec;
      float4 *o = out + col * planes * plane_vec + q;
      for (int p = 0; p < planes; ++p) __stcs(o + p * plane_vec, v);
    }
  } else if (MODE == 1) {
    const int64_t chunks = plane_vec / (256 * CH) * ncol;          // plane_vec is a multiple of 256 * CH
    for (int64_t c = blockIdx.x; c < chunks; c += gridDim.x) {
      const int64_t per = plane_vec / (256 * CH);
      const int64_t col = c / per, q0 = (c - col * per) * (256 * CH);
      float4 *o = out + col * planes * plane_vec + q0 + threadIdx.x;
      for (int p = 0; p < planes; ++p)
#pragma unroll
        for (int i = 0; i < CH; ++i) __stcs(o + p * plane_vec + i * 256, v);
    }
  } else if (MODE == 3) {
    // fill order with the fill kernel's read: one int4 of a plane-sized template per thread and column
    const int4 *tmpl = reinterpret_cast<const int4 *>(out + plane_vec * planes * ncol);
    const int64_t total = plane_vec * ncol;
    for (int64_t w = (int64_t)blockIdx.x * 256 + threadIdx.x; w < total; w += (int64_t)gridDim.x * 256) {
      const int64_t col = w / plane_vec, q = w - col * plane_vec;
      const int4 t = __ldg(tmpl + q);
      const float4 x = make_float4(__int_as_float(t.x), __int_as_float(t.y), __int_as_float(t.z), __int_as_float(t.w));
      float4 *o = out + col * planes * plane_vec + q;
      for (int p = 0; p < planes; ++p) __stcs(o + p * plane_vec, x);
    }
  } else if (MODE == 4) {
    // persistent grid, R = CH regions of the slab, each swept cyclically by gridDim.x / R CTAs
    // (R = gridDim.x: blocked partition, every CTA streams through its own contiguous share)
    const int64_t chunks = plane_vec / 256 * ncol;                 // 4 KB per plane each
    const int R = CH > 0 ? CH : gridDim.x;
    const int64_t per_region = (chunks + R - 1) / R;
    const int region = blockIdx.x % R, lane = blockIdx.x / R, width = (gridDim.x + R - 1 - region) / R;
    for (int64_t k = lane; k < per_region; k += width) {
      const int64_t c = (int64_t)region * per_region + k;
      if (c >= chunks) break;
      const int64_t per = plane_vec / 256;
      const int64_t col = c / per, q = (c - col * per) * 256 + threadIdx.x;
      float4 *o = out + col * planes * plane_vec + q;
      for (int p = 0; p < planes; ++p) __stcs(o + p * plane_vec, v);
    }
  } else if (MODE == 5) {
    // one 4 KB x planes chunk per CTA, grid = number of chunks (hardware dispatch order)
    const int64_t c = blockIdx.x;
    const int64_t per = plane_vec / 256;
    const int64_t col = c / per, q = (c - col * per) * 256 + threadIdx.x;
    if (col < ncol) {
      float4 *o = out + col * planes * plane_vec + q;
      for (int p = 0; p < planes; ++p) __stcs(o + p * plane_vec, v);
    }
  } else if (MODE == 6) {
    // one chunk per CTA, chunk order scrambled (multiplication by an odd constant modulo the count)
    const int64_t per = plane_vec / 256, chunks = per * ncol;
    const int64_t c = (int64_t)(((unsigned __int128)blockIdx.x * 2654435761ull) % (uint64_t)chunks);
    const int64_t col = c / per, q = (c - col * per) * 256 + threadIdx.x;
    float4 *o = out + col * planes * plane_vec + q;
    for (int p = 0; p < planes; ++p) __stcs(o + p * plane_vec, v);
  } else if (MODE == 7) {
    // persistent, blocked: a CTA writes CH consecutive 4 KB pieces of one plane, then the same run
    // of the next plane, then takes its next run (cyclic over runs)
    const int64_t runs_per_plane = plane_vec / (256 * (int64_t)CH), runs = runs_per_plane * ncol;
    for (int64_t r = blockIdx.x; r < runs; r += gridDim.x) {
      const int64_t col = r / runs_per_plane, q0 = (r - col * runs_per_plane) * (256 * (int64_t)CH);
      float4 *o = out + col * planes * plane_vec + q0 + threadIdx.x;
      for (int p = 0; p < planes; ++p)
#pragma unroll 4
        for (int i = 0; i < CH; ++i) __stcs(o + p * plane_vec + (int64_t)i * 256, v);
    }
  } else {
    // plane-major over groups of G columns: w enumerates (group, plane, col in group, q)
    const int G = CH;
    const int64_t total = plane_vec * ncol * planes;
    for (int64_t w = (int64_t)blockIdx.x * 256 + threadIdx.x; w < total; w += (int64_t)gridDim.x * 256) {
      const int64_t per_group = plane_vec * G * planes;
      const int64_t g = w / per_group;
      int64_t r = w - g * per_group;
      const int64_t p = r / (plane_vec * G);
      r -= p * plane_vec * G;
      const int64_t cg = r / plane_vec, q = r - cg * plane_vec;
      const int64_t col = g * G + cg;
      if (col < ncol) __stcs(out + (col * planes + p) * plane_vec + q, v);
    }
  }
}

// The cube kernels' store pattern: a CTA owns a 128-pixel tile and walks `cols` columns of 4 planes
// (grid = (column chunks, pixel tiles), dynamic shared memory pads the residency to 2 CTAs per SM).
//   KIND 0  thread = pixel, one 32-bit store per (column, plane): 128 B per warp instruction
//   KIND 1  a warp covers the 128 pixels of one (column, plane) row with one 128-bit store per lane
//   KIND 2  rows leave from shared memory as 512-byte bulk copies (cp.async.bulk), one lane per row
template <int KIND>
__global__ void __launch_bounds__(128) tile_probe(float *__restrict__ out, int64_t npix, int64_t ncol, int cols) {
  extern __shared__ __align__(128) float stage[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t pix0 = (int64_t)blockIdx.y * 128;
  const int64_t c0 = (int64_t)blockIdx.x * cols;
  const int64_t c1 = c0 + cols < ncol ? c0 + cols : ncol;
  if (KIND == 0) {
    for (int64_t c = c0; c < c1; ++c) {
      float *o = out + c * 4 * npix + pix0 + tid;
#pragma unroll
      for (int p = 0; p < 4; ++p) __stcs(o + p * npix, 1.0f);
    }
  } else if (KIND == 1) {
    const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
    for (int64_t c = c0; c < c1; ++c)       // warp w stores plane w of the column
      __stcs(reinterpret_cast<float4 *>(out + (c * 4 + warp) * npix + pix0) + lane, v);
  } else {
    for (int i = tid; i < 128 * 8; i += 128) stage[i] = 1.0f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const unsigned src = (unsigned)__cvta_generic_to_shared(stage);
    // 4 warps x 32 lanes: lane l of warp w issues row (c, p) with (c - c0) * 4 + p = k * 128 + tid
    const int64_t rows = (c1 - c0) * 4;
    for (int64_t r = tid; r < rows; r += 128) {
      float *dst = out + (c0 * 4 + r) * npix + pix0;
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], 512;" ::"l"(dst), "r"(src + (unsigned)(r & 7) * 512)
                   : "memory");
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

// The 32-bit store pattern above plus an independent stream of L2-resident reads (ld.global.cg from
// a 32 MB window): 16 bytes per thread (cp.async, not waited for) every `every` columns = 100 % / every of the bytes written.
__global__ void __launch_bounds__(128) tile_read_probe(float *__restrict__ out, const float *__restrict__ window, int64_t npix,
                                                        int64_t ncol, int cols, int every, uint32_t window_bytes) {
  const int tid = threadIdx.x;
  const int64_t pix0 = (int64_t)blockIdx.y * 128;
  const int64_t c0 = (int64_t)blockIdx.x * cols;
  const int64_t c1 = c0 + cols < ncol ? c0 + cols : ncol;
  extern __shared__ __align__(128) float stage[];
  const uint32_t mask = window_bytes / 16 - 1;
  uint32_t pos = ((uint32_t)blockIdx.y * 977u + (uint32_t)blockIdx.x * 131071u) * 128u;
  const unsigned dst = (unsigned)__cvta_generic_to_shared(stage) + tid * 16;
  const float4 *w4 = reinterpret_cast<const float4 *>(window);
  int slot = 0;
  const bool keep = every < 0;              // negative: the reads carry the L2 evict_last policy
  if (keep) every = -every;
  uint64_t policy = 0;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(policy));
  for (int64_t c = c0; c < c1; ++c) {
    if (every > 0 && ((int)(c - c0) & (every - 1)) == 0) {
      // asynchronous 16-byte copy global -> shared (never waited for inside the loop): 2 KB per CTA
      if (keep)
        asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(dst + (unsigned)slot * 2048),
                     "l"(w4 + ((pos + tid) & mask)), "l"(policy)
                     : "memory");
      else
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (unsigned)slot * 2048), "l"(w4 + ((pos + tid) & mask)) : "memory");
      asm volatile("cp.async.commit_group;" ::: "memory");
      slot = (slot + 1) & 15;
      pos += 128u * 4099u;
    }
    float *o = out + c * 4 * npix + pix0 + tid;
#pragma unroll
    for (int p = 0; p < 4; ++p) __stcs(o + p * npix, 1.0f);
  }
  asm volatile("cp.async.wait_all;" ::: "memory");
}

static cudaStream_t g_stream = nullptr;     // launches of run_tile_read go here (access policy window experiments)
static double run_tile_read(float *buf, int64_t npix, int64_t total_bytes, int cols, int smem_kb, int every, uint32_t window_bytes = 32u << 20) {
  const int64_t ncol = total_bytes / (npix * 16);
  const float *window = buf + ncol * npix * 4;     // the 64 MB of slack behind the slab
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  CK(cudaFuncSetAttribute(tile_read_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_kb * 1024));
  dim3 grid((unsigned)((ncol + cols - 1) / cols), (unsigned)(npix / 128));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaEventRecord(a, g_stream));
    tile_read_probe<<<grid, 128, smem_kb * 1024, g_stream>>>(buf, window_bytes > (64u << 20) ? buf : window, npix, ncol, cols, every, window_bytes);
    CK(cudaEventRecord(b, g_stream));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    if (rep > 0 && ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return (double)ncol * npix * 16 / best / 1e6;
}

// The cube kernels' operand read: every CTA first copies the 24 KB basis tile of its pixel tile
// (DRAM: each tile is read by the `gx` CTAs of its pixel tile only) and then stores its columns.
// group = 0: the reads trickle in at use time.  group = G > 0: the first CTA of every G-th pixel
// tile also prefetches the tiles of the NEXT group into L2 in one burst (cp.async.bulk.prefetch.L2),
// so that DRAM sees the reads in bursts instead of one here and there between the writes.
__global__ void __launch_bounds__(128) tile_operand_probe(float *__restrict__ out, const float *__restrict__ basis, int64_t npix,
                                                           int64_t ncol, int cols, int group, int tile_bytes) {
  extern __shared__ __align__(128) float stage[];
  const int tid = threadIdx.x;
  const int64_t pix0 = (int64_t)blockIdx.y * 128;
  const int64_t c0 = (int64_t)blockIdx.x * cols;
  const int64_t c1 = c0 + cols < ncol ? c0 + cols : ncol;
  const char *mine = reinterpret_cast<const char *>(basis) + (int64_t)blockIdx.y * tile_bytes;
  const bool hints = group < 0;            // negative: prefetch with evict_last, use with evict_first
  if (hints) group = -group;
  uint64_t keep_policy = 0, drop_policy = 0;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(keep_policy));
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(drop_policy));
  if (group > 0 && blockIdx.x == 0 && blockIdx.y % group == 0 && tid == 0) {
    const int64_t first = (int64_t)blockIdx.y + group, last = first + group < gridDim.y ? first + group : gridDim.y;
    for (int64_t t = first; t < last; ++t)
      if (hints)
        asm volatile("cp.async.bulk.prefetch.L2.global.L2::cache_hint [%0], %1, %2;" ::"l"(reinterpret_cast<const char *>(basis) + t * tile_bytes),
                     "r"(tile_bytes), "l"(keep_policy)
                     : "memory");
      else
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(reinterpret_cast<const char *>(basis) + t * tile_bytes), "r"(tile_bytes)
                   : "memory");
  }
  const unsigned dst = (unsigned)__cvta_generic_to_shared(stage);
  for (int off = tid * 16; off < tile_bytes; off += 128 * 16)
    if (hints && blockIdx.x == gridDim.x - 1)
      asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(dst + off), "l"(mine + off), "l"(drop_policy) : "memory");
    else
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + off), "l"(mine + off) : "memory");
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncthreads();
  const float v = stage[tid];
  for (int64_t c = c0; c < c1; ++c) {
    float *o = out + c * 4 * npix + pix0 + tid;
#pragma unroll
    for (int p = 0; p < 4; ++p) __stcs(o + p * npix, v);
  }
}

static double run_tile_operand(float *buf, int64_t npix, int64_t total_bytes, int cols, int group, int tile_bytes) {
  const int64_t ncol = total_bytes / (npix * 16);
  const float *basis = buf + ncol * npix * 4;      // behind the slab (1 GiB of slack)
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  CK(cudaFuncSetAttribute(tile_operand_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  dim3 grid((unsigned)((ncol + cols - 1) / cols), (unsigned)(npix / 128));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaEventRecord(a));
    tile_operand_probe<<<grid, 128, 100 * 1024>>>(buf, basis, npix, ncol, cols, group, tile_bytes);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    if (rep > 0 && ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return (double)ncol * npix * 16 / best / 1e6;
}

template <int KIND>
static double run_tile(float *buf, int64_t npix, int64_t total_bytes, int cols, int smem_kb) {
  const int64_t ncol = total_bytes / (npix * 16);
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  CK(cudaFuncSetAttribute(tile_probe<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_kb * 1024));
  dim3 grid((unsigned)((ncol + cols - 1) / cols), (unsigned)(npix / 128));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaEventRecord(a));
    tile_probe<KIND><<<grid, 128, smem_kb * 1024>>>(buf, npix, ncol, cols);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    if (rep > 0 && ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return (double)ncol * npix * 16 / best / 1e6;
}

template <int MODE, int CH>
static double run(float4 *buf, int64_t plane_bytes, int planes, int64_t total_bytes, int sms, int ctas) {
  const int64_t plane_vec = plane_bytes / 16;
  const int64_t ncol = total_bytes / (plane_bytes * planes);
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    CK(cudaEventRecord(a));
    const int64_t grid = ctas > 0 ? (int64_t)sms * ctas : plane_vec / 256 * ncol;
    probe<MODE, CH><<<(unsigned)grid, 256>>>(buf, plane_vec, planes, ncol);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    if (rep > 0 && ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return (double)ncol * planes * plane_bytes / best / 1e6;   // GB/s
}

int main(int argc, char **argv) {
  const double gib = argc > 1 ? atof(argv[1]) : 23.0;
  const int64_t total = (int64_t)(gib * (1 << 30));
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  float4 *buf = nullptr;
  CK(cudaMalloc(&buf, total + (1 << 30)));
  printf("# write-only probe: %.1f GiB per launch, %d SMs, grid = SMs x 8 CTAs x 256 threads, best of 3 (GB/s)\n", gib, sms);
  printf("# plane_bytes planes : fill order | 16 KB per CTA and plane | plane-major (groups of 4 columns)\n");
  const int64_t sizes[] = {256 << 10, 1 << 20, 4 << 20, 16 << 20, 64 << 20, (16 << 20) + (68 << 10), 17 << 20};
  for (int64_t d : sizes)
    for (int planes : {1, 2, 4, 8}) {
      if (planes != 4 && d != (16 << 20) && d != (1 << 20)) continue;
      const double a = run<0, 1>(buf, d, planes, total, sms, 8);
      const double b = run<1, 4>(buf, d, planes, total, sms, 8);
      const double c = 0.0;  // plane-major (mode 2) is bound by its index arithmetic: 3600 GB/s everywhere, dropped
      printf("%10lld %d : %7.0f | %7.0f | %7.0f\n", (long long)d, planes, a, b, c);
      fflush(stdout);
    }
  // grid width, 4 planes, fill order; then the same with the template read of the fill kernel
  printf("# CTAs per SM in the grid (8 are resident) : write only, 16 MB | 1 MB planes || with template read, 16 MB | 4 MB | 1 MB planes\n");
  for (int ctas : {1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 16, 20, 24, 32, 48, 64, 128, 256}) {
    const double a = run<0, 1>(buf, 16 << 20, 4, total, sms, ctas);
    const double a1 = run<0, 1>(buf, 1 << 20, 4, total, sms, ctas);
    const double r16 = run<3, 1>(buf, 16 << 20, 4, total, sms, ctas);
    const double r4 = run<3, 1>(buf, 4 << 20, 4, total, sms, ctas);
    const double r1 = run<3, 1>(buf, 1 << 20, 4, total, sms, ctas);
    printf("%3d : %7.0f | %7.0f || %7.0f | %7.0f | %7.0f\n", ctas, a, a1, r16, r4, r1);
    fflush(stdout);
  }
  printf("# persistent grid of 8 CTAs per SM, 4 planes, write only; 16 MB | 1 MB planes\n");
  printf("one chunk per CTA (grid = chunks)  : %7.0f | %7.0f\n", run<5, 1>(buf, 16 << 20, 4, total, sms, 0), run<5, 1>(buf, 1 << 20, 4, total, sms, 0));
  printf("cyclic (fill order)                : %7.0f | %7.0f\n", run<0, 1>(buf, 16 << 20, 4, total, sms, 8), run<0, 1>(buf, 1 << 20, 4, total, sms, 8));
  printf("2 regions                          : %7.0f | %7.0f\n", run<4, 2>(buf, 16 << 20, 4, total, sms, 8), run<4, 2>(buf, 1 << 20, 4, total, sms, 8));
  printf("4 regions                          : %7.0f | %7.0f\n", run<4, 4>(buf, 16 << 20, 4, total, sms, 8), run<4, 4>(buf, 1 << 20, 4, total, sms, 8));
  printf("8 regions                          : %7.0f | %7.0f\n", run<4, 8>(buf, 16 << 20, 4, total, sms, 8), run<4, 8>(buf, 1 << 20, 4, total, sms, 8));
  printf("16 regions                         : %7.0f | %7.0f\n", run<4, 16>(buf, 16 << 20, 4, total, sms, 8), run<4, 16>(buf, 1 << 20, 4, total, sms, 8));
  printf("37 regions                         : %7.0f | %7.0f\n", run<4, 37>(buf, 16 << 20, 4, total, sms, 8), run<4, 37>(buf, 1 << 20, 4, total, sms, 8));
  printf("148 regions                        : %7.0f | %7.0f\n", run<4, 148>(buf, 16 << 20, 4, total, sms, 8), run<4, 148>(buf, 1 << 20, 4, total, sms, 8));
  printf("blocked (one region per CTA)       : %7.0f | %7.0f\n", run<4, 0>(buf, 16 << 20, 4, total, sms, 8), run<4, 0>(buf, 1 << 20, 4, total, sms, 8));
  printf("# the same with 2 CTAs per SM (the cube kernels' residency): cyclic, 16 regions, blocked\n");
  printf("cyclic %7.0f | 16 regions %7.0f | 148 regions %7.0f | blocked %7.0f   (16 MB planes)\n", run<0, 1>(buf, 16 << 20, 4, total, sms, 2),
         run<4, 16>(buf, 16 << 20, 4, total, sms, 2), run<4, 148>(buf, 16 << 20, 4, total, sms, 2), run<4, 0>(buf, 16 << 20, 4, total, sms, 2));
  printf("# one chunk per CTA, scrambled chunk order: %7.0f | %7.0f\n", run<6, 1>(buf, 16 << 20, 4, total, sms, 0), run<6, 1>(buf, 1 << 20, 4, total, sms, 0));
  printf("# persistent (8 CTAs per SM), contiguous run per CTA and plane: 4 KB %7.0f, 16 KB %7.0f, 64 KB %7.0f, 256 KB %7.0f, 1 MB %7.0f, 4 MB %7.0f  (16 MB planes)\n",
         run<7, 1>(buf, 16 << 20, 4, total, sms, 8), run<7, 4>(buf, 16 << 20, 4, total, sms, 8), run<7, 16>(buf, 16 << 20, 4, total, sms, 8),
         run<7, 64>(buf, 16 << 20, 4, total, sms, 8), run<7, 256>(buf, 16 << 20, 4, total, sms, 8), run<7, 1024>(buf, 16 << 20, 4, total, sms, 8));
  printf("# the same with 2 CTAs per SM: 4 KB %7.0f, 64 KB %7.0f, 1 MB %7.0f\n", run<7, 1>(buf, 16 << 20, 4, total, sms, 2),
         run<7, 16>(buf, 16 << 20, 4, total, sms, 2), run<7, 256>(buf, 16 << 20, 4, total, sms, 2));
  printf("# cube-kernel store pattern (128-pixel tile x columns per CTA), 512^2 | 2048^2 images: 32-bit stores, 128-bit stores, 512 B bulk copies\n");
  for (int smem_kb : {100, 48, 24}) for (int cols : {256, 64, 16}) {
    float *f = reinterpret_cast<float *>(buf);
    printf("%3d KB smem (%d CTAs/SM) %3d columns per CTA : %6.0f %6.0f %6.0f | %6.0f %6.0f %6.0f\n", smem_kb, 227 / (smem_kb + 1), cols,
           run_tile<0>(f, 512 * 512, total, cols, smem_kb), run_tile<1>(f, 512 * 512, total, cols, smem_kb), run_tile<2>(f, 512 * 512, total, cols, smem_kb),
           run_tile<0>(f, 2048 * 2048, total, cols, smem_kb), run_tile<1>(f, 2048 * 2048, total, cols, smem_kb), run_tile<2>(f, 2048 * 2048, total, cols, smem_kb));
    fflush(stdout);
  }
  printf("# 32-bit store pattern, 2 CTAs per SM, plus asynchronous L2-resident reads of 0 / 3 / 6 / 12.5 / 25 / 50 %% of the bytes written (GB/s written)\n");
  for (int cols : {256, 64}) for (int64_t n : {512, 2048}) {
    float *f = reinterpret_cast<float *>(buf);
    printf("%4lld^2 %3d columns per CTA : %6.0f %6.0f %6.0f %6.0f %6.0f", (long long)n, cols, run_tile_read(f, n * n, total, cols, 100, 0),
           run_tile_read(f, n * n, total, cols, 100, 32), run_tile_read(f, n * n, total, cols, 100, 16), run_tile_read(f, n * n, total, cols, 100, 8),
           run_tile_read(f, n * n, total, cols, 100, 4));
    printf(" %6.0f", run_tile_read(f, n * n, total, cols, 100, 2));
    printf("\n");
    fflush(stdout);
  }
  printf("# the same at 512^2, 64 columns per CTA, by size of the window the reads come from (2 GB: the slab itself, DRAM): reads of 3 %% | 25 %%\n");
  for (uint32_t wb : {256u << 10, 1u << 20, 8u << 20, 32u << 20, 64u << 20, 2048u << 20}) {
    float *f = reinterpret_cast<float *>(buf);
    printf("%7u KB window : %6.0f | %6.0f\n", wb >> 10, run_tile_read(f, 512 * 512, total, 64, 100, 32, wb), run_tile_read(f, 512 * 512, total, 64, 100, 4, wb));
    fflush(stdout);
  }
  {
    float *f = reinterpret_cast<float *>(buf);
    printf("# 512^2, 256 columns per CTA, rare DRAM reads (2 GB window): 0 / 0.4 / 0.8 / 1.6 / 3 %% of the bytes written\n");
    printf("  %6.0f %6.0f %6.0f %6.0f %6.0f\n", run_tile_read(f, 512 * 512, total, 256, 100, 0, 2048u << 20), run_tile_read(f, 512 * 512, total, 256, 100, 256, 2048u << 20),
           run_tile_read(f, 512 * 512, total, 256, 100, 128, 2048u << 20), run_tile_read(f, 512 * 512, total, 256, 100, 64, 2048u << 20),
           run_tile_read(f, 512 * 512, total, 256, 100, 32, 2048u << 20));
    printf("# 512^2, 64 columns per CTA, reads with the L2 evict_last policy, by window: 3 %% | 25 %% of the bytes written\n");
    for (uint32_t wb : {8u << 20, 32u << 20, 64u << 20}) {
      printf("%7u KB window, evict_last : %6.0f | %6.0f\n", wb >> 10, run_tile_read(f, 512 * 512, total, 64, 100, -32, wb), run_tile_read(f, 512 * 512, total, 64, 100, -4, wb));
      fflush(stdout);
    }
    printf("# write only again, after the evict_last runs : %6.0f\n", run_tile_read(f, 512 * 512, total, 64, 100, 0));
  }
  printf("# operand tile per CTA (DRAM) + 32-bit stores, 2 CTAs per SM, 2048^2, by prefetch group (0 = reads at use time): GB/s written\n");
  for (int cols : {256, 128}) for (int tile_bytes : {24576}) {
    float *f = reinterpret_cast<float *>(buf);
    printf("%3d columns per CTA, %5d B tiles :", cols, tile_bytes);
    for (int group : {0, 64, 256, 1024, -32, -64, -128, -256, -512}) printf(" G=%d %6.0f", group, run_tile_operand(f, 2048 * 2048, total, cols, group, tile_bytes));
    printf("\n");
    fflush(stdout);
  }
  if (argc > 2) {
    // (second argument given) the same reads with the window marked persisting in L2 (set-aside of the
    // maximum size): the set-aside takes L2 from the write stream -- 2900 GB/s, and it stays that way
    // for the rest of the process, so this block runs last
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    printf("# L2 %d MB, persisting L2 maximum %d MB, access policy window maximum %d MB\n", prop.l2CacheSize >> 20, prop.persistingL2CacheMaxSize >> 20,
           prop.accessPolicyMaxWindowSize >> 20);
    CK(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, prop.persistingL2CacheMaxSize));
    CK(cudaStreamCreate(&g_stream));
    float *f = reinterpret_cast<float *>(buf);
    const int64_t ncol = total / (512 * 512 * 16);
    for (uint32_t wb : {8u << 20, 32u << 20, 64u << 20}) {
      cudaStreamAttrValue attr = {};
      attr.accessPolicyWindow.base_ptr = f + ncol * 512 * 512 * 4;
      attr.accessPolicyWindow.num_bytes = wb;
      attr.accessPolicyWindow.hitRatio = 1.0f;
      attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
      attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
      CK(cudaStreamSetAttribute(g_stream, cudaStreamAttributeAccessPolicyWindow, &attr));
      printf("%7u KB window, persisting : %6.0f | %6.0f\n", wb >> 10, run_tile_read(f, 512 * 512, total, 64, 100, 32, wb), run_tile_read(f, 512 * 512, total, 64, 100, 4, wb));
      fflush(stdout);
    }
    cudaStreamAttrValue off = {};
    off.accessPolicyWindow.num_bytes = 0;
    CK(cudaStreamSetAttribute(g_stream, cudaStreamAttributeAccessPolicyWindow, &off));
    CK(cudaCtxResetPersistingL2Cache());
    g_stream = nullptr;
  }
  return 0;
}
