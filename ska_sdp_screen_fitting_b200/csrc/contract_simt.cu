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
                     int64_t npix, int64_t ncol, float *__restrict__ slab, int64_t col_tiles_per_cta) {
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
        if (KL) {
          sincosf(acc[c][j], &im[j], &re[j]);
        } else {
          re[j] = acc[c][j];
          im[j] = acc[TC + c][j];
        }
        re[j] = ssf::fix_re(re[j]);
        im[j] = ssf::fix_im(im[j]);
      }
      float *out = slab + col * 4 * npix + p0;
      if (VEC) {
        float4 r4 = make_float4(re[0], re[1], re[2], re[3]), i4 = make_float4(im[0], im[1], im[2], im[3]);
        __stcs(reinterpret_cast<float4 *>(out), r4);
        __stcs(reinterpret_cast<float4 *>(out + npix), i4);
        __stcs(reinterpret_cast<float4 *>(out + 2 * npix), r4);
        __stcs(reinterpret_cast<float4 *>(out + 3 * npix), i4);
      } else {
#pragma unroll
        for (int j = 0; j < PIX_PER_THREAD; ++j) {
          if (p0 + j < npix) {
            out[j] = re[j];
            out[npix + j] = im[j];
            out[2 * npix + j] = re[j];
            out[3 * npix + j] = im[j];
          }
        }
      }
    }
  }
}

template <bool KL>
int launch(const float *Bt, const float *G, int ndir, int ldg, int64_t npix, int64_t ncol,
           float *slab, cudaStream_t stream) {
  const bool vec = (npix % 4 == 0) && ((uintptr_t)slab % 16 == 0) && ((uintptr_t)Bt % 16 == 0);
  const int pix_tiles = ssf::div_up(npix, TILE_PIX);
  const int64_t n_ct = (ncol + TC - 1) / TC;
  // enough CTAs for ~8 per SM, but keep each CTA on one pixel tile for many column tiles
  int64_t want_y = (148 * 8 + pix_tiles - 1) / pix_tiles;
  if (want_y > n_ct) want_y = n_ct;
  if (want_y > 65535) want_y = 65535;
  if (want_y < 1) want_y = 1;
  const int64_t per = (n_ct + want_y - 1) / want_y;
  const int gy = (int)((n_ct + per - 1) / per);
  dim3 grid(pix_tiles, gy);
  size_t smem = sizeof(float) * ndir * (KL ? TC : 2 * TC);
  auto kfn = vec ? contract_simt_kernel<KL, true> : contract_simt_kernel<KL, false>;
  SSF_LAUNCH(kfn, grid, THREADS, smem, stream, Bt, G, ndir, ldg, npix, ncol, slab, per);
  SSF_LAUNCHED();
  return SSF_OK;
}

}  // namespace

namespace ssf {
int kl_eval_simt(const float *Bt, const float *G, int ndir, int ldg, int64_t npix, int64_t ncol,
                 float *slab, cudaStream_t stream) {
  return launch<true>(Bt, G, ndir, ldg, npix, ncol, slab, stream);
}
int voronoi_contract_simt(const float *St, const float *lut, int ndir, int64_t npix, int64_t ncol,
                          float *slab, cudaStream_t stream) {
  return launch<false>(St, lut, ndir, ndir, npix, ncol, slab, stream);
}
}  // namespace ssf
