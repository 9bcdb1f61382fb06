// Basis contraction, FP32 SIMT variant (variant 0): the parity reference for the tensor-core
// variant and the path for shapes the tcgen05 kernel does not take (tiny / ragged images).
//
//   KL:       Phi[pix, col] = sum_d Bt[d, pix] * G[col, d];  slab = (cos, sin, cos, sin)(Phi)
//   Voronoi:  Re[pix, col]  = sum_d St[d, pix] * lut[col, d, 0]
//             Im[pix, col]  = sum_d St[d, pix] * lut[col, d, 1]; slab = (Re, Im, Re, Im)
//
// Bt / St are stored direction-major ([ld][npix], pixel fastest) so a thread's 4 consecutive
// pixels are one 16 B load and a warp's loads and stores are fully coalesced.  A CTA owns a
// 512-pixel tile and walks column tiles; its basis tile stays L1-resident across them.
#include "common.cuh"

namespace {

constexpr int THREADS = 128;
constexpr int PIX_PER_THREAD = 4;
constexpr int TILE_PIX = THREADS * PIX_PER_THREAD;
constexpr int TC = 8;  // columns per inner tile

template <bool KL, bool VEC>
__global__ void __launch_bounds__(THREADS)
contract_simt_kernel(const float *__restrict__ Bt, const float *__restrict__ G, int ndir, int ldg,
                     int64_t npix, int64_t ncol, float *__restrict__ slab, int64_t col_tiles_per_cta,
                     ssf::OutFmt fmt) {
  // KL: coefficient tile gs[d][TC]; Voronoi: gs[d][2*TC] (re block then im block)
  SSF_DYN_SMEM(float, gs);
  constexpr int W = KL ? TC : 2 * TC;
  const int64_t p0 = (int64_t)blockIdx.x * TILE_PIX + (int64_t)threadIdx.x * PIX_PER_THREAD;
  const int64_t ct_begin = (int64_t)blockIdx.y * col_tiles_per_cta;
  const int64_t n_ct = (ncol + TC - 1) / TC;
  int64_t ct_end = ct_begin + col_tiles_per_cta;
  if (ct_end > n_ct) ct_end = n_ct;

  for (int64_t ct = ct_begin; ct < ct_end; ++ct) {
    const int64_t c0 = ct * TC;
    __syncthreads();
    for (int i = threadIdx.x; i < ndir * TC; i += THREADS) {
      int c = i / ndir, d = i - c * ndir;
      int64_t col = c0 + c;
      if (KL) {
        gs[d * W + c] = col < ncol ? G[col * ldg + d] : 0.0f;
      } else {
        float2 v = col < ncol ? reinterpret_cast<const float2 *>(G)[col * ldg + d] : make_float2(0.f, 0.f);
        gs[d * W + c] = v.x;
        gs[d * W + TC + c] = v.y;
      }
    }
    __syncthreads();
    if (p0 >= npix) continue;

    float acc[W][PIX_PER_THREAD];
#pragma unroll
    for (int c = 0; c < W; ++c)
#pragma unroll
      for (int j = 0; j < PIX_PER_THREAD; ++j) acc[c][j] = 0.0f;

    for (int d = 0; d < ndir; ++d) {
      float b[PIX_PER_THREAD];
      if (VEC) {
        float4 v = __ldg(reinterpret_cast<const float4 *>(Bt + (int64_t)d * npix + p0));
        b[0] = v.x; b[1] = v.y; b[2] = v.z; b[3] = v.w;
      } else {
#pragma unroll
        for (int j = 0; j < PIX_PER_THREAD; ++j)
          b[j] = (p0 + j < npix) ? __ldg(Bt + (int64_t)d * npix + p0 + j) : 0.0f;
      }
      const float4 *g4 = reinterpret_cast<const float4 *>(gs + d * W);
#pragma unroll
      for (int q = 0; q < W / 4; ++q) {
        float4 g = g4[q];
#pragma unroll
        for (int j = 0; j < PIX_PER_THREAD; ++j) {
          acc[4 * q + 0][j] = fmaf(b[j], g.x, acc[4 * q + 0][j]);
          acc[4 * q + 1][j] = fmaf(b[j], g.y, acc[4 * q + 1][j]);
          acc[4 * q + 2][j] = fmaf(b[j], g.z, acc[4 * q + 2][j]);
          acc[4 * q + 3][j] = fmaf(b[j], g.w, acc[4 * q + 3][j]);
        }
      }
    }

#pragma unroll
    for (int c = 0; c < TC; ++c) {
      const int64_t col = c0 + c;
      if (col >= ncol) break;
      float re[PIX_PER_THREAD], im[PIX_PER_THREAD];
#pragma unroll
      for (int j = 0; j < PIX_PER_THREAD; ++j) {
        if (KL && fmt.planes == 1) {             // raw screen values (tec cube): NaN kept
          re[j] = ssf::out_word(acc[c][j], fmt.prmt);
          im[j] = 0.0f;
          continue;
        }
        if (KL) {
          sincosf(acc[c][j], &im[j], &re[j]);
        } else {
          re[j] = acc[c][j];
          im[j] = acc[TC + c][j];
        }
        re[j] = ssf::out_word(ssf::fix_re(re[j]), fmt.prmt);
        im[j] = ssf::out_word(ssf::fix_im(im[j]), fmt.prmt);
      }
      float *out = slab + col * fmt.planes * npix + p0;
      if (VEC) {
        float4 r4 = make_float4(re[0], re[1], re[2], re[3]), i4 = make_float4(im[0], im[1], im[2], im[3]);
        __stcs(reinterpret_cast<float4 *>(out), r4);
        if (fmt.planes >= 2) __stcs(reinterpret_cast<float4 *>(out + npix), i4);
        if (fmt.planes == 4) {
          __stcs(reinterpret_cast<float4 *>(out + 2 * npix), r4);
          __stcs(reinterpret_cast<float4 *>(out + 3 * npix), i4);
        }
      } else {
#pragma unroll
        for (int j = 0; j < PIX_PER_THREAD; ++j) {
          if (p0 + j < npix) {
            out[j] = re[j];
            if (fmt.planes >= 2) out[npix + j] = im[j];
            if (fmt.planes == 4) {
              out[2 * npix + j] = re[j];
              out[3 * npix + j] = im[j];
            }
          }
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------
// Tiled variant (the fast path): a CTA stages its basis tile Bt[:, tile] in shared memory ONCE
// and walks a chunk of the columns in tiles of TC, so the inner loop touches shared memory
// only (1 full-width LDS for PPT pixels + TC/4 broadcast LDS.128 per PPT*TC FMAs).  The next
// coefficient tile is prefetched into registers while the current one is being contracted.
// Grid = (column chunks, pixel tiles) with the chunk index fastest, so CTAs that share a basis
// tile run together and the tile is served by L2; the chunk count is chosen so that the grid
// is a whole number of waves of the resident CTA capacity (2 per SM).
//
// sin/cos: explicit two-constant range reduction to [-pi, pi] followed by the MUFU
// approximations (abs. error < 5e-7 there), well inside the 1e-5 tolerance on the screens.
__device__ __forceinline__ void sincos_reduced(float x, float *s, float *c) {
  const float k = rintf(x * 0.15915494309189535f);
  float r = fmaf(-k, 6.2831854820251465f, x);   // 2*pi rounded to float
  r = fmaf(-k, -1.7484555e-7f, r);               // 2*pi - float(2*pi)
  *s = __sinf(r);
  *c = __cosf(r);
}

constexpr int MAX_PRE = (SSF_MAX_DIR * 16 + THREADS - 1) / THREADS;  // coefficient prefetch registers

template <bool KL, int PPT, int TCOL>
__global__ void __launch_bounds__(THREADS, 2)
contract_tile_kernel(const float *__restrict__ Bt, const float *__restrict__ G, int ndir, int ldg,
                     int64_t npix, int64_t ncol, float *__restrict__ slab, int tiles_per_chunk,
                     ssf::OutFmt fmt) {
  constexpr int TPIX = THREADS * PPT;
  constexpr int W = KL ? TCOL : 2 * TCOL;       // accumulator columns (Voronoi: re block, im block)
  SSF_DYN_SMEM(float, sm);
  float *bs = sm;                                // [ndir][TPIX]
  float *gs = sm + (size_t)ndir * TPIX;          // [ndir][W]
  const int tid = threadIdx.x;
  const int64_t pix0 = (int64_t)blockIdx.y * TPIX;
  const int64_t p0 = pix0 + (int64_t)tid * PPT;
  const bool active = p0 < npix;                 // npix % PPT == 0: a thread is all in or all out
  const int64_t n_ct = (ncol + TCOL - 1) / TCOL;
  const int64_t ct_begin = (int64_t)blockIdx.x * tiles_per_chunk;
  int64_t ct_end = ct_begin + tiles_per_chunk;
  if (ct_end > n_ct) ct_end = n_ct;
  if (ct_begin >= ct_end) return;

  // stage the basis tile (coalesced vector loads, conflict-free vector stores)
  for (int d = 0; d < ndir; ++d) {
    const float *src = Bt + (int64_t)d * npix + p0;
    float *dst = bs + d * TPIX + tid * PPT;
    if (PPT == 4) {
      float4 v = active ? __ldg(reinterpret_cast<const float4 *>(src)) : make_float4(0.f, 0.f, 0.f, 0.f);
      *reinterpret_cast<float4 *>(dst) = v;
    } else {
      float2 v = active ? __ldg(reinterpret_cast<const float2 *>(src)) : make_float2(0.f, 0.f);
      *reinterpret_cast<float2 *>(dst) = v;
    }
  }

  // coefficient prefetch: element e = c * ndir + d of the tile, e = tid + k * THREADS
  const int n_el = ndir * TCOL;
  float pre_a[MAX_PRE], pre_b[MAX_PRE];
  auto prefetch = [&](int64_t ct) {
    const int64_t c0 = ct * TCOL;
#pragma unroll
    for (int k = 0; k < MAX_PRE; ++k) {
      const int e = tid + k * THREADS;
      pre_a[k] = 0.f;
      pre_b[k] = 0.f;
      if (e < n_el) {
        const int c = e / ndir, d = e - c * ndir;
        const int64_t col = c0 + c;
        if (col < ncol) {
          if (KL) {
            pre_a[k] = __ldg(G + col * ldg + d);
          } else {
            float2 v = __ldg(reinterpret_cast<const float2 *>(G) + col * ldg + d);
            pre_a[k] = v.x;
            pre_b[k] = v.y;
          }
        }
      }
    }
  };
  prefetch(ct_begin);

  for (int64_t ct = ct_begin; ct < ct_end; ++ct) {
    __syncthreads();                             // previous tile fully consumed (and bs staged)
#pragma unroll
    for (int k = 0; k < MAX_PRE; ++k) {
      const int e = tid + k * THREADS;
      if (e < n_el) {
        const int c = e / ndir, d = e - c * ndir;
        gs[d * W + c] = pre_a[k];
        if (!KL) gs[d * W + TCOL + c] = pre_b[k];
      }
    }
    __syncthreads();
    if (ct + 1 < ct_end) prefetch(ct + 1);

    float acc[W][PPT];
#pragma unroll
    for (int c = 0; c < W; ++c)
#pragma unroll
      for (int j = 0; j < PPT; ++j) acc[c][j] = 0.0f;

#pragma unroll 3
    for (int d = 0; d < ndir; ++d) {
      float b[PPT];
      if (PPT == 4) {
        float4 v = *reinterpret_cast<const float4 *>(bs + d * TPIX + tid * 4);
        b[0] = v.x; b[1] = v.y; b[2] = v.z; b[3] = v.w;
      } else {
        float2 v = *reinterpret_cast<const float2 *>(bs + d * TPIX + tid * 2);
        b[0] = v.x; b[PPT - 1] = v.y;
      }
      const float4 *g4 = reinterpret_cast<const float4 *>(gs + d * W);
#pragma unroll
      for (int q = 0; q < W / 4; ++q) {
        const float4 g = g4[q];
#pragma unroll
        for (int j = 0; j < PPT; ++j) {
          acc[4 * q + 0][j] = fmaf(b[j], g.x, acc[4 * q + 0][j]);
          acc[4 * q + 1][j] = fmaf(b[j], g.y, acc[4 * q + 1][j]);
          acc[4 * q + 2][j] = fmaf(b[j], g.z, acc[4 * q + 2][j]);
          acc[4 * q + 3][j] = fmaf(b[j], g.w, acc[4 * q + 3][j]);
        }
      }
    }

    if (!active) continue;
    const int64_t c0 = ct * TCOL;
#pragma unroll
    for (int c = 0; c < TCOL; ++c) {
      const int64_t col = c0 + c;
      if (col >= ncol) break;
      float re[PPT], im[PPT];
#pragma unroll
      for (int j = 0; j < PPT; ++j) {
        if (KL && fmt.planes == 1) {             // raw screen values (tec cube): NaN kept
          re[j] = ssf::out_word(acc[c][j], fmt.prmt);
          im[j] = 0.0f;
          continue;
        }
        if (KL) {
          sincos_reduced(acc[c][j], &im[j], &re[j]);
        } else {
          re[j] = acc[c][j];
          im[j] = acc[TCOL + c][j];
        }
        re[j] = ssf::out_word(ssf::fix_re(re[j]), fmt.prmt);
        im[j] = ssf::out_word(ssf::fix_im(im[j]), fmt.prmt);
      }
      float *out = slab + col * fmt.planes * npix + p0;
      if (PPT == 4) {
        const float4 r4 = make_float4(re[0], re[1], re[2], re[3]);
        const float4 i4 = make_float4(im[0], im[1], im[2], im[3]);
        __stcs(reinterpret_cast<float4 *>(out), r4);
        if (fmt.planes >= 2) __stcs(reinterpret_cast<float4 *>(out + npix), i4);
        if (fmt.planes == 4) {
          __stcs(reinterpret_cast<float4 *>(out + 2 * npix), r4);
          __stcs(reinterpret_cast<float4 *>(out + 3 * npix), i4);
        }
      } else {
        const float2 r2 = make_float2(re[0], re[PPT - 1]);
        const float2 i2 = make_float2(im[0], im[PPT - 1]);
        __stcs(reinterpret_cast<float2 *>(out), r2);
        if (fmt.planes >= 2) __stcs(reinterpret_cast<float2 *>(out + npix), i2);
        if (fmt.planes == 4) {
          __stcs(reinterpret_cast<float2 *>(out + 2 * npix), r2);
          __stcs(reinterpret_cast<float2 *>(out + 3 * npix), i2);
        }
      }
    }
  }
}

// Column chunks per pixel tile.  A CTA stages its basis tile once (worth about 3 column tiles
// of time) and then walks `per` column tiles, so the launch takes about
// waves * (3 + per) tile times; take the chunk count that minimises that.  Few, long CTAs
// win unless the grid would leave SMs idle or the last wave nearly empty.
int pick_chunks(int pix_tiles, int64_t n_ct, int capacity) {
  int best = 1;
  double best_cost = 1e300;
  const int64_t max_ch = n_ct < 256 ? n_ct : 256;
  for (int64_t ch = 1; ch <= max_ch; ++ch) {
    const int64_t per = (n_ct + ch - 1) / ch;
    const int64_t real = (n_ct + per - 1) / per;      // chunks that actually have work
    const int64_t ctas = real * pix_tiles;
    const int64_t waves = (ctas + capacity - 1) / capacity;
    const double cost = (double)waves * (3.0 + (double)per);
    if (cost < best_cost - 1e-9) {
      best_cost = cost;
      best = (int)ch;
    }
  }
  return best;
}

template <bool KL, int PPT, int TCOL>
int launch_tile(const float *Bt, const float *G, int ndir, int ldg, int64_t npix, int64_t ncol,
                float *slab, ssf::OutFmt fmt, cudaStream_t stream) {
  constexpr int TPIX = THREADS * PPT;
  constexpr int W = KL ? TCOL : 2 * TCOL;
  const size_t smem = sizeof(float) * ((size_t)ndir * TPIX + (size_t)ndir * W);
  auto kfn = contract_tile_kernel<KL, PPT, TCOL>;
  SSF_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int pix_tiles = ssf::div_up(npix, TPIX);
  const int64_t n_ct = (ncol + TCOL - 1) / TCOL;
  const int per_sm = smem <= 113 * 1024 ? 2 : 1;
  const int chunks = pick_chunks(pix_tiles, n_ct, ssf::sm_count() * per_sm);
  const int per = (int)((n_ct + chunks - 1) / chunks);
  const int gx = (int)((n_ct + per - 1) / per);
  SSF_REQUIRE(pix_tiles <= 65535, "image too large for the tiled contraction grid");
  dim3 grid(gx, pix_tiles);
  SSF_LAUNCH(kfn, grid, THREADS, smem, stream, Bt, G, ndir, ldg, npix, ncol, slab, per, fmt);
  SSF_LAUNCHED();
  return SSF_OK;
}

template <bool KL>
int launch(const float *Bt, const float *G, int ndir, int ldg, int64_t npix, int64_t ncol,
           float *slab, ssf::OutFmt fmt, cudaStream_t stream) {
  const bool vec = (npix % 4 == 0) && ((uintptr_t)slab % 16 == 0) && ((uintptr_t)Bt % 16 == 0);
  const bool g_ok = KL || ((uintptr_t)G % 8 == 0);
  if (vec && g_ok && npix >= 1024 && ncol >= 4) {
    // tiled fast path; 4 pixels per thread while the basis tile leaves room for 2 CTAs per SM
    if (KL) {
      if (ndir <= 52) return launch_tile<true, 4, 16>(Bt, G, ndir, ldg, npix, ncol, slab, fmt, stream);
      return launch_tile<true, 2, 16>(Bt, G, ndir, ldg, npix, ncol, slab, fmt, stream);
    }
    if (ndir <= 52) return launch_tile<false, 4, 8>(Bt, G, ndir, ldg, npix, ncol, slab, fmt, stream);
    return launch_tile<false, 2, 8>(Bt, G, ndir, ldg, npix, ncol, slab, fmt, stream);
  }
  const int pix_tiles = ssf::div_up(npix, TILE_PIX);
  const int64_t n_ct = (ncol + TC - 1) / TC;
  // enough CTAs for ~8 per SM, but keep each CTA on one pixel tile for many column tiles
  int64_t want_y = (ssf::sm_count() * 8 + pix_tiles - 1) / pix_tiles;
  if (want_y > n_ct) want_y = n_ct;
  if (want_y > 65535) want_y = 65535;
  if (want_y < 1) want_y = 1;
  const int64_t per = (n_ct + want_y - 1) / want_y;
  const int gy = (int)((n_ct + per - 1) / per);
  dim3 grid(pix_tiles, gy);
  size_t smem = sizeof(float) * ndir * (KL ? TC : 2 * TC);
  auto kfn = vec ? contract_simt_kernel<KL, true> : contract_simt_kernel<KL, false>;
  SSF_LAUNCH(kfn, grid, THREADS, smem, stream, Bt, G, ndir, ldg, npix, ncol, slab, per, fmt);
  SSF_LAUNCHED();
  return SSF_OK;
}

}  // namespace

namespace ssf {
int kl_eval_simt(const float *Bt, const float *G, int ndir, int ldg, int64_t npix, int64_t ncol,
                 float *slab, OutFmt fmt, cudaStream_t stream) {
  return launch<true>(Bt, G, ndir, ldg, npix, ncol, slab, fmt, stream);
}
int voronoi_contract_simt(const float *St, const float *lut, int ndir, int64_t npix, int64_t ncol,
                          float *slab, OutFmt fmt, cudaStream_t stream) {
  return launch<false>(St, lut, ndir, ndir, npix, ncol, slab, fmt, stream);
}
}  // namespace ssf
