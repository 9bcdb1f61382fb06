// KL evaluation support: the per-pixel interpolation basis (FP64, once per run), the
// per-column coefficients (FP64 GEMV per column), and the ssf_kl_eval / ssf_voronoi_contract
// dispatch between the SIMT and tcgen05 contraction kernels.
#include "common.cuh"

namespace ssf {
int kl_eval_simt(const float *, const float *, int, int, int64_t, int64_t, float *, OutFmt, cudaStream_t);
int voronoi_contract_simt(const float *, const float *, int, int64_t, int64_t, float *, OutFmt, cudaStream_t);
int contract_tc(bool kl, const float *, const float *, int, int, int64_t, int64_t, float *, OutFmt, cudaStream_t);
bool contract_tc_supported(bool kl, int ndir, int64_t npix, int64_t ncol, const void *bt, const void *slab);
int contract_tc_gain(const float *, const float *, int, int, int64_t, int64_t, float *, OutFmt, cudaStream_t);
}  // namespace ssf

namespace {

// One thread per pixel.  K(pix, e) in FP64 for all e, then the ndir x ndir product with
// pinvC (shared memory).  Bt is written pixel-fastest -> coalesced.
__global__ void __launch_bounds__(128)
basis_kernel(const double *__restrict__ xc, const double *__restrict__ yc, int nx, int ny,
             const double *__restrict__ pp, int ndir, const double *__restrict__ pinvC, double r0sq,
             double half_beta, float *__restrict__ Bt) {
  SSF_DYN_SMEM(double, sm);
  double *s_pinv = sm;                 // [ndir][ndir]
  double *s_pp = sm + ndir * ndir;     // [ndir][2]
  for (int i = threadIdx.x; i < ndir * ndir; i += blockDim.x) s_pinv[i] = pinvC[i];
  for (int i = threadIdx.x; i < ndir * 2; i += blockDim.x) s_pp[i] = pp[i];
  __syncthreads();
  const int64_t npix = (int64_t)nx * ny;
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npix) return;
  const double x = xc[p % nx], y = yc[p / nx];
  double k[SSF_MAX_DIR];
  for (int e = 0; e < ndir; ++e) {
    // kl_screen.py:446-448: sum of squares over (x, y, 0), then -((d2/r0^2)^(beta/2))/2
    double dx = s_pp[2 * e] - x, dy = s_pp[2 * e + 1] - y;
    double d2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
    double q = d2 / r0sq;
    k[e] = q > 0.0 ? -0.5 * pow(q, half_beta) : 0.0;
  }
  for (int d = 0; d < ndir; ++d) {
    double acc = 0.0;
    for (int e = 0; e < ndir; ++e) acc = fma(k[e], s_pinv[e * ndir + d], acc);
    Bt[(int64_t)d * npix + p] = (float)acc;
  }
}

// G[col, d] = sum_e C[d, e] * white[col, e]; block = 128 threads, each block handles a run
// of columns; thread t handles (col_local, d) pairs.  C (symmetric) lives in shared memory.
__global__ void __launch_bounds__(128)
coeffs_kernel(const double *__restrict__ white, int64_t ncol, int ndir, const double *__restrict__ C,
              float *__restrict__ G, int ldg, int cols_per_block) {
  SSF_DYN_SMEM(double, sm);
  double *s_c = sm;                    // [ndir][ndir]
  double *s_w = sm + ndir * ndir;      // [cols_per_block][ndir]
  for (int i = threadIdx.x; i < ndir * ndir; i += blockDim.x) s_c[i] = C[i];
  for (int64_t c0 = (int64_t)blockIdx.x * cols_per_block; c0 < ncol;
       c0 += (int64_t)gridDim.x * cols_per_block) {
    __syncthreads();
    for (int i = threadIdx.x; i < cols_per_block * ndir; i += blockDim.x) {
      int64_t col = c0 + i / ndir;
      s_w[i] = col < ncol ? white[col * ndir + (i % ndir)] : 0.0;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < cols_per_block * ldg; i += blockDim.x) {
      int cl = i / ldg, d = i - cl * ldg;
      int64_t col = c0 + cl;
      if (col >= ncol) continue;
      double acc = 0.0;
      if (d < ndir) {
        const double *w = s_w + cl * ndir;
        // C is symmetric: read column d as row d (stride-1 in e would bank-conflict across d)
        for (int e = 0; e < ndir; ++e) acc = fma(s_c[e * ndir + d], w[e], acc);
      }
      G[col * ldg + d] = (float)acc;
    }
  }
}


// Gain solutions (kl_screen.py:319-378): three contractions per voxel against the same basis --
// the phase screen and the log10 XX / YY amplitude screens -- and
//   slab = (10^Axx cos Phi, 10^Axx sin Phi, 10^Ayy cos Phi, 10^Ayy sin Phi).
// G3 [ncol][3][ldg] = coefficients (phase, Axx, Ayy).  FP32 SIMT: a CTA owns a 512-pixel basis
// tile (L1-resident across its columns), a thread 4 consecutive pixels.
constexpr int GAIN_THREADS = 128;

template <bool VEC>
__global__ void __launch_bounds__(GAIN_THREADS)
kl_gain_kernel(const float *__restrict__ Bt, const float *__restrict__ G3, int ndir, int ldg, int64_t npix,
               int64_t ncol, float *__restrict__ slab, int64_t cols_per_cta, uint32_t prmt) {
  SSF_DYN_SMEM(float, gs);                       // [3][ndir] of the current column
  const int64_t p0 = ((int64_t)blockIdx.x * GAIN_THREADS + threadIdx.x) * 4;
  const int64_t c_begin = (int64_t)blockIdx.y * cols_per_cta;
  int64_t c_end = c_begin + cols_per_cta;
  if (c_end > ncol) c_end = ncol;
  for (int64_t col = c_begin; col < c_end; ++col) {
    __syncthreads();
    for (int i = threadIdx.x; i < 3 * ndir; i += GAIN_THREADS) {
      const int k = i / ndir, d = i - k * ndir;
      gs[i] = G3[(col * 3 + k) * ldg + d];
    }
    __syncthreads();
    if (p0 >= npix) continue;
    float acc[3][4];
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[k][j] = 0.0f;
    for (int d = 0; d < ndir; ++d) {
      float b[4];
      if (VEC) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(Bt + (int64_t)d * npix + p0));
        b[0] = v.x; b[1] = v.y; b[2] = v.z; b[3] = v.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) b[j] = (p0 + j < npix) ? __ldg(Bt + (int64_t)d * npix + p0 + j) : 0.0f;
      }
      const float g0 = gs[d], g1 = gs[ndir + d], g2 = gs[2 * ndir + d];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[0][j] = fmaf(b[j], g0, acc[0][j]);
        acc[1][j] = fmaf(b[j], g1, acc[1][j]);
        acc[2][j] = fmaf(b[j], g2, acc[2][j]);
      }
    }
    float pl[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float sn, cs;
      sincosf(acc[0][j], &sn, &cs);
      const float ax = exp10f(acc[1][j]), ay = exp10f(acc[2][j]);
      pl[0][j] = ssf::out_word(ssf::fix_re(ax * cs), prmt);
      pl[1][j] = ssf::out_word(ssf::fix_im(ax * sn), prmt);
      pl[2][j] = ssf::out_word(ssf::fix_re(ay * cs), prmt);
      pl[3][j] = ssf::out_word(ssf::fix_im(ay * sn), prmt);
    }
    float *out = slab + col * 4 * npix + p0;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      if (VEC) {
        __stcs(reinterpret_cast<float4 *>(out + p * npix), make_float4(pl[p][0], pl[p][1], pl[p][2], pl[p][3]));
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (p0 + j < npix) out[p * npix + j] = pl[p][j];
      }
    }
  }
}

}  // namespace

extern "C" int ssf_kl_basis(const double *x_coord, const double *y_coord, int nx, int ny,
                            const double *pp, int ndir, const double *pinvC, double r0, double beta,
                            float *Bt, void *stream) {
  SSF_REQUIRE(x_coord && y_coord && pp && pinvC && Bt, "ssf_kl_basis: null pointer");
  SSF_REQUIRE(ndir >= 1 && ndir <= SSF_MAX_DIR, "ssf_kl_basis: ndir=%d out of range [1,%d]", ndir, SSF_MAX_DIR);
  SSF_REQUIRE(nx >= 1 && ny >= 1, "ssf_kl_basis: empty image");
  size_t smem = sizeof(double) * ((size_t)ndir * ndir + 2 * ndir);
  if (smem > 48 * 1024)
    SSF_CUDA(cudaFuncSetAttribute(basis_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t npix = (int64_t)nx * ny;
  SSF_LAUNCH(basis_kernel, ssf::div_up(npix, 128), 128, smem, (cudaStream_t)stream,
             x_coord, y_coord, nx, ny, pp, ndir, pinvC, r0 * r0, beta / 2.0, Bt);
  SSF_LAUNCHED();
  return SSF_OK;
}

extern "C" int ssf_kl_coeffs(const double *white, int64_t ncol, int ndir, const double *C, float *G,
                             int ldg, void *stream) {
  SSF_REQUIRE(white && C && G, "ssf_kl_coeffs: null pointer");
  SSF_REQUIRE(ndir >= 1 && ndir <= SSF_MAX_DIR && ldg >= ndir, "ssf_kl_coeffs: bad ndir/ldg");
  if (ncol <= 0) return SSF_OK;
  const int cpb = 16;
  size_t smem = sizeof(double) * ((size_t)ndir * ndir + (size_t)cpb * ndir);
  if (smem > 48 * 1024)
    SSF_CUDA(cudaFuncSetAttribute(coeffs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t blocks = (ncol + cpb - 1) / cpb;
  if (blocks > ssf::sm_count() * 8) blocks = ssf::sm_count() * 8;
  SSF_LAUNCH(coeffs_kernel, (int)blocks, 128, smem, (cudaStream_t)stream, white, ncol, ndir, C, G, ldg, cpb);
  SSF_LAUNCHED();
  return SSF_OK;
}

namespace {
int kl_eval_any(const char *who, const float *Bt, const float *G, int ndir, int ldg, int64_t npix, int64_t ncol,
                float *slab, int variant, ssf::OutFmt fmt, cudaStream_t stream) {
  SSF_REQUIRE(Bt && G && slab, "%s: null pointer", who);
  SSF_REQUIRE(ndir >= 1 && ndir <= SSF_MAX_DIR && ldg >= ndir && npix >= 1, "%s: bad sizes", who);
  if (ncol <= 0) return SSF_OK;
  bool tc_ok = ssf::contract_tc_supported(true, ndir, npix, ncol, Bt, slab) && (ldg % 4 == 0);
  if (variant == 1 && !tc_ok) {
    ssf::set_error("%s: tcgen05 variant needs npix %% 128 == 0, 16 B aligned buffers, ldg %% 4 == 0", who);
    return SSF_ENOTSUP;
  }
  if (variant == 1 || (variant < 0 && tc_ok)) return ssf::contract_tc(true, Bt, G, ndir, ldg, npix, ncol, slab, fmt, stream);
  return ssf::kl_eval_simt(Bt, G, ndir, ldg, npix, ncol, slab, fmt, stream);
}
}  // namespace

extern "C" int ssf_kl_eval(const float *Bt, const float *G, int ndir, int ldg, int64_t npix,
                           int64_t ncol, float *slab, int variant, int out_flags, void *stream) {
  ssf::OutFmt fmt;
  SSF_REQUIRE(ssf::out_fmt(out_flags, &fmt), "ssf_kl_eval: unknown out_flags %d", out_flags);
  return kl_eval_any("ssf_kl_eval", Bt, G, ndir, ldg, npix, ncol, slab, variant, fmt, (cudaStream_t)stream);
}

extern "C" int ssf_kl_eval_tec(const float *Bt, const float *G, int ndir, int ldg, int64_t npix,
                               int64_t ncol, float *slab, int variant, int out_flags, void *stream) {
  ssf::OutFmt fmt;
  SSF_REQUIRE(ssf::out_fmt(out_flags, &fmt) && fmt.planes == 4, "ssf_kl_eval_tec: out_flags may only hold SSF_OUT_BIG_ENDIAN");
  fmt.planes = 1;
  return kl_eval_any("ssf_kl_eval_tec", Bt, G, ndir, ldg, npix, ncol, slab, variant, fmt, (cudaStream_t)stream);
}

extern "C" int ssf_voronoi_contract(const float *S, const float *lut, int ndir, int64_t npix,
                                    int64_t ncol, float *slab, int variant, int out_flags, void *stream) {
  SSF_REQUIRE(S && lut && slab, "ssf_voronoi_contract: null pointer");
  ssf::OutFmt fmt;
  SSF_REQUIRE(ssf::out_fmt(out_flags, &fmt), "ssf_voronoi_contract: unknown out_flags %d", out_flags);
  SSF_REQUIRE(ndir >= 1 && ndir <= SSF_MAX_DIR && npix >= 1, "ssf_voronoi_contract: bad sizes");
  if (ncol <= 0) return SSF_OK;
  bool tc_ok = ssf::contract_tc_supported(false, ndir, npix, ncol, S, slab);
  if (variant == 1 && !tc_ok) {
    ssf::set_error("ssf_voronoi_contract: tcgen05 variant needs npix %% 128 == 0 and 16 B aligned buffers");
    return SSF_ENOTSUP;
  }
  if (variant == 1 || (variant < 0 && tc_ok))
    return ssf::contract_tc(false, S, lut, ndir, ndir, npix, ncol, slab, fmt, (cudaStream_t)stream);
  return ssf::voronoi_contract_simt(S, lut, ndir, npix, ncol, slab, fmt, (cudaStream_t)stream);
}


extern "C" int ssf_kl_eval_gain(const float *Bt, const float *G3, int ndir, int ldg, int64_t npix,
                                int64_t ncol, float *slab, int variant, int out_flags, void *stream) {
  SSF_REQUIRE(Bt && G3 && slab, "ssf_kl_eval_gain: null pointer");
  ssf::OutFmt fmt;
  SSF_REQUIRE(ssf::out_fmt(out_flags, &fmt) && fmt.planes == 4,
              "ssf_kl_eval_gain: out_flags may only hold SSF_OUT_BIG_ENDIAN (the four planes differ)");
  SSF_REQUIRE(ndir >= 1 && ndir <= SSF_MAX_DIR && ldg >= ndir && npix >= 1, "ssf_kl_eval_gain: bad sizes");
  if (ncol <= 0) return SSF_OK;
  const bool tc_ok = ssf::contract_tc_supported(true, ndir, npix, ncol, Bt, slab);
  if (variant == 1 && !tc_ok) {
    ssf::set_error("ssf_kl_eval_gain: tcgen05 variant needs npix %% 128 == 0 and 16 B aligned buffers");
    return SSF_ENOTSUP;
  }
  if (variant == 1 || (variant < 0 && tc_ok))
    return ssf::contract_tc_gain(Bt, G3, ndir, ldg, npix, ncol, slab, fmt, (cudaStream_t)stream);
  const int64_t pix_tiles = (npix + 4 * GAIN_THREADS - 1) / (4 * GAIN_THREADS);
  // enough column chunks for ~8 CTAs per SM; a chunk keeps its basis tile in L1 across columns
  int64_t chunks = (ssf::sm_count() * 8 + pix_tiles - 1) / pix_tiles;
  if (chunks > ncol) chunks = ncol;
  if (chunks > 65535) chunks = 65535;
  const int64_t per = (ncol + chunks - 1) / chunks;
  dim3 grid((unsigned)pix_tiles, (unsigned)((ncol + per - 1) / per));
  const size_t smem = sizeof(float) * 3 * (size_t)ndir;
  const bool vec = npix % 4 == 0 && (uintptr_t)Bt % 16 == 0 && (uintptr_t)slab % 16 == 0;
  if (vec)
    SSF_LAUNCH(kl_gain_kernel<true>, grid, GAIN_THREADS, smem, (cudaStream_t)stream, Bt, G3, ndir, ldg, npix, ncol,
               slab, per, fmt.prmt);
  else
    SSF_LAUNCH(kl_gain_kernel<false>, grid, GAIN_THREADS, smem, (cudaStream_t)stream, Bt, G3, ndir, ldg, npix, ncol,
               slab, per, fmt.prmt);
  SSF_LAUNCHED();
  return SSF_OK;
}
