// Batched Karhunen-Loeve phase-screen fit.  One CTA per system (one (time, freq, station)
// cell = one ndir-vector of phases + weights), FP64 throughout, everything in shared memory
// except the pattern matrices (C_S, pinv C_S, U_S), which are shared by all systems with the
// same flag pattern and are read through L1/L2.
//
// Replaces _fit_screen (stationscreen.py:433-594), the per-time body of _process_station
// (:681-779), _circ_chi2 (:353-387) and _flag_outliers (:303-350).
#include <stdlib.h>

#include <atomic>

#include "common.cuh"

namespace {

std::atomic<int> g_fit_ctas_per_sm{16};

// Threads per CTA of the fit kernel: a template parameter FT >= ndir (64 up to 64 directions,
// 128 above): half the idle lanes, registers and barrier width at the LOFAR direction count.
constexpr double TWO_PI = 6.283185307179586476925286766559;
constexpr double PI = 3.141592653589793238462643383279;

struct Pattern {
  const double *C, *P, *U;  // [n][n] blocks with row stride ld
  const double *Pfull;      // pinv of the full pattern (pattern 0)
  const double *Cfull;      // structure matrix of the full pattern: C(e, d) for ALL direction pairs
  int ld;
};

struct Smem {
  double *ph, *w;                   // [nall] phases, weights
  double *rc, *rs, *wc;             // [n] compact cos, sin, weight
  double *zc, *zs;                  // [k]
  double *yc, *ys, *gc, *gs_;       // [n]
  double *fit, *wh, *fall;          // [n], [n], [nall]
  double *A;                        // [k][k] (non-uniform weights only)
  int *idx;                         // [n] compact -> direction
};

__device__ Smem carve(double *base, int nall) {
  Smem s;
  double *p = base;
  s.ph = p; p += nall;
  s.w = p; p += nall;
  s.rc = p; p += nall;
  s.rs = p; p += nall;
  s.wc = p; p += nall;
  s.zc = p; p += nall;
  s.zs = p; p += nall;
  s.yc = p; p += nall;
  s.ys = p; p += nall;
  s.gc = p; p += nall;
  s.gs_ = p; p += nall;
  s.fit = p; p += nall;
  s.wh = p; p += nall;
  s.fall = p; p += nall;
  s.idx = reinterpret_cast<int *>(p); p += (nall + 1) / 2;
  s.A = p;
  return s;
}

__host__ __device__ inline size_t smem_doubles(int nall) {
  return (size_t)14 * nall + (nall + 1) / 2 + (size_t)nall * nall;
}

// One screen fit at `order` (stationscreen.py:501-590, phase type).  Whole CTA participates.
// On return s.wh/s.fall hold the results: white_out/resid_out are written to global.
// Weights all equal: closed form; all > cut: Cholesky solve; some weight in (0, cut]: the general
// pinv by one-sided Jacobi (below).  Always returns true (kept for the callers' status plumbing).
template <int FT>
__device__ bool fit_once(const Smem &s, const Pattern &pt, const double *__restrict__ pp, int nall,
                         int n, int k, double r0sq, double half_beta, double cut,
                         double *__restrict__ white_out, double *__restrict__ resid_out, int kind) {
  const bool amp = kind == 1;           // kind: 0 phase, 1 amplitude (log10), 2 tec (linear, :549-562)
  const int tid = threadIdx.x;
  const int ld = pt.ld;
  if (tid < n) {
    double phi = s.ph[s.idx[tid]];
    double sn = 0.0, cs;
    if (amp) cs = log10(phi);          // amplitude type fits log10(amp) (:535-541); one right-hand side
    else if (kind == 2) cs = phi;      // tec type fits the values themselves (:549-556)
    else sincos(phi, &sn, &cs);
    s.rc[tid] = cs;
    s.rs[tid] = sn;
    s.wc[tid] = s.w[s.idx[tid]];
  }
  __syncthreads();
  // bit 0: some weight differs from the first one; bit 1: some weight is not above the cutoff
  const int my_bad = tid < n ? ((s.wc[tid] == s.wc[0] ? 0 : 1) | (s.wc[tid] > cut ? 0 : 2)) : 0;
  // rr1 = U_k^T W r   (:513-516, :522-525)
  if (tid < k) {
    double ac = 0.0, as = 0.0;
    for (int i = 0; i < n; ++i) {
      double u = pt.U[i * ld + tid] * s.wc[i];
      ac = fma(u, s.rc[i], ac);
      as = fma(u, s.rs[i], as);
    }
    s.zc[tid] = ac;
    s.zs[tid] = as;
  }
  // the barrier the stage needs anyway, with the two votes (the reduction is a boolean OR)
  const bool uniform = !__syncthreads_or(my_bad & 1);               // both uniform over the CTA
  const bool general = __syncthreads_or(my_bad & 2) && !uniform;
  if (general) {
    // Some weight lies in (0, cut]: A = U_k^T W U_k may have eigenvalues <= cut, which
    // pinv(A, rcond=1e-3) drops (:501-504; absolute cutoff).  One-sided (Hestenes) Jacobi on
    // B = W^(1/2) U_k  (A = B^T B = V S^2 V^T): rotating column pairs until they are orthogonal
    // gives B' = B V with |b'_j|^2 = lambda_j, and with g = W^(1/2) r
    //     U_k pinv(A) U_k^T W r = W^(-1/2) B' diag(lambda_j > cut ? 1 / lambda_j : 0) B'^T g,
    // so V never has to be formed: B lives in s.A (n k <= nall^2 doubles), y is produced here.
    for (int e = tid; e < n * k; e += FT) {
      const int i = e / k, j = e - i * k;
      s.A[e] = sqrt(s.wc[i]) * pt.U[i * ld + j];
    }
    __shared__ int sh_rotated;
    const int warp = tid >> 5, lane = tid & 31, m = k + (k & 1);
    for (int sweep = 0; sweep < 40; ++sweep) {
      __syncthreads();
      if (tid == 0) sh_rotated = 0;
      __syncthreads();
      for (int step = 0; step < m - 1; ++step) {
        // round-robin tournament: m / 2 disjoint column pairs per step, one warp per pair
        for (int pr = warp; pr < m / 2; pr += FT / 32) {
          int p = pr == 0 ? m - 1 : (step + pr) % (m - 1);
          int q = pr == 0 ? step : (step - pr + m - 1) % (m - 1);
          if (p >= k || q >= k) continue;
          if (p > q) {
            const int t = p;
            p = q;
            q = t;
          }
          double al = 0.0, be = 0.0, ga = 0.0;
          for (int i = lane; i < n; i += 32) {
            const double bp = s.A[i * k + p], bq = s.A[i * k + q];
            al = fma(bp, bp, al);
            be = fma(bq, bq, be);
            ga = fma(bp, bq, ga);
          }
          for (int o = 16; o; o >>= 1) {
            al += __shfl_xor_sync(0xffffffffu, al, o);
            be += __shfl_xor_sync(0xffffffffu, be, o);
            ga += __shfl_xor_sync(0xffffffffu, ga, o);
          }
          if (fabs(ga) <= 1e-13 * sqrt(al * be)) continue;      // uniform over the warp (lambda errs by ~ga^2)
          if (lane == 0) sh_rotated = 1;
          const double zeta = (be - al) / (2.0 * ga);
          const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
          const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
          for (int i = lane; i < n; i += 32) {
            const double bp = s.A[i * k + p], bq = s.A[i * k + q];
            s.A[i * k + p] = c * bp - sn * bq;
            s.A[i * k + q] = sn * bp + c * bq;
          }
        }
        __syncthreads();
      }
      if (!sh_rotated) break;
    }
    __syncthreads();
    // coefficients a_j = (b'_j . g) / lambda_j for lambda_j > cut, else 0 (both right-hand sides)
    if (tid < k) {
      double lam = 0.0, yc = 0.0, ys = 0.0;
      for (int i = 0; i < n; ++i) {
        const double b = s.A[i * k + tid], sw = sqrt(s.wc[i]);
        lam = fma(b, b, lam);
        yc = fma(b, sw * s.rc[i], yc);
        ys = fma(b, sw * s.rs[i], ys);
      }
      const double inv = lam > cut ? 1.0 / lam : 0.0;
      s.zc[tid] = yc * inv;
      s.zs[tid] = ys * inv;
    }
    __syncthreads();
    // y = U_k pinv(A) z = W^(-1/2) B' a
    if (tid < n) {
      double ac = 0.0, as = 0.0;
      for (int j = 0; j < k; ++j) {
        const double b = s.A[tid * k + j];
        ac = fma(b, s.zc[j], ac);
        as = fma(b, s.zs[j], as);
      }
      const double isw = 1.0 / sqrt(s.wc[tid]);
      s.yc[tid] = ac * isw;
      s.ys[tid] = as * isw;
    }
  } else if (uniform) {
    // A = w0 U_k^T U_k = w0 I  ->  pinv(A) = I / w0 when w0 > cut, else 0   (:501-504)
    if (tid < k) {
      double w0 = s.wc[0];
      double inv = w0 > cut ? 1.0 / w0 : 0.0;
      s.zc[tid] *= inv;
      s.zs[tid] *= inv;
    }
  } else {
    // A = U_k^T W U_k is SPD with eigenvalues in [min w, max w] (Rayleigh quotient, U_k
    // orthonormal); min w > cut so pinv(A, cut) = inv(A): Cholesky solve.
    for (int e = tid; e < k * k; e += FT) {
      int a = e / k, b = e - a * k;
      if (b > a) continue;
      double acc = 0.0;
      for (int i = 0; i < n; ++i) acc = fma(pt.U[i * ld + a] * s.wc[i], pt.U[i * ld + b], acc);
      s.A[a * k + b] = acc;
      s.A[b * k + a] = acc;
    }
    __syncthreads();
    for (int j = 0; j < k; ++j) {  // right-looking Cholesky, lower triangle
      if (tid == 0) s.A[j * k + j] = sqrt(s.A[j * k + j]);
      __syncthreads();
      double d = s.A[j * k + j];
      for (int i = j + 1 + tid; i < k; i += FT) s.A[i * k + j] /= d;
      __syncthreads();
      for (int e = tid; e < (k - j - 1) * (k - j - 1); e += FT) {
        int a = j + 1 + e / (k - j - 1), b = j + 1 + e % (k - j - 1);
        if (b <= a) s.A[a * k + b] -= s.A[a * k + j] * s.A[b * k + j];
      }
      __syncthreads();
    }
    if (tid < 2) {  // two right-hand sides, forward then backward substitution
      double *z = tid == 0 ? s.zc : s.zs;
      for (int i = 0; i < k; ++i) {
        double v = z[i];
        for (int j = 0; j < i; ++j) v -= s.A[i * k + j] * z[j];
        z[i] = v / s.A[i * k + i];
      }
      for (int i = k - 1; i >= 0; --i) {
        double v = z[i];
        for (int j = i + 1; j < k; ++j) v -= s.A[j * k + i] * z[j];
        z[i] = v / s.A[i * k + i];
      }
    }
  }
  __syncthreads();
  // y = U_k z
  if (!general && tid < n) {
    double ac = 0.0, as = 0.0;
    for (int j = 0; j < k; ++j) {
      double u = pt.U[tid * ld + j];
      ac = fma(u, s.zc[j], ac);
      as = fma(u, s.zs[j], as);
    }
    s.yc[tid] = ac;
    s.ys[tid] = as;
  }
  __syncthreads();
  // g = pinvC y   (:517-519, :526-528); pinvC symmetric -> column reads are coalesced
  if (tid < n) {
    double ac = 0.0, as = 0.0;
    for (int e = 0; e < n; ++e) {
      double p = pt.P[e * ld + tid];
      ac = fma(p, s.yc[e], ac);
      as = fma(p, s.ys[e], as);
    }
    s.gc[tid] = ac;
    s.gs_[tid] = as;
  }
  __syncthreads();
  // screen_fit = atan2(C g_imag, C g_real)   (:531-533)
  if (tid < n) {
    double ac = 0.0, as = 0.0;
    for (int e = 0; e < n; ++e) {
      double c = pt.C[e * ld + tid];
      ac = fma(c, s.gc[e], ac);
      as = fma(c, s.gs_[e], as);
    }
    s.fit[tid] = kind != 0 ? ac : atan2(as, ac);  // amplitude / tec: screen_fit = C fit (:547, :561)
  }
  __syncthreads();
  // white = pinvC fit   (:534)
  if (tid < n) {
    double a = 0.0;
    for (int e = 0; e < n; ++e) a = fma(pt.P[e * ld + tid], s.fit[e], a);
    s.wh[tid] = a;
  }
  __syncthreads();
  if (n == nall) {
    // resid = phase - C white   (:590)
    if (tid < n) {
      double a = 0.0;
      for (int e = 0; e < n; ++e) a = fma(pt.C[e * ld + tid], s.wh[e], a);
      white_out[tid] = s.wh[tid];
      resid_out[tid] = s.ph[tid] - (amp ? pow(10.0, a) : a);     // :584-590
    }
  } else {
    // flagged directions are predicted from the unflagged screen   (:565-574)
    if (tid < nall) s.fall[tid] = 0.0;
    __syncthreads();
    if (tid < n) s.fall[s.idx[tid]] = s.fit[tid];
    __syncthreads();
    if (tid < nall && !(s.w[tid] > 0.0)) {
      // c_val = -((|pp_e - pp_tid|^2 / r0^2)^(beta/2)) / 2 is entry (e, tid) of the FULL pattern's
      // structure matrix, which decompose_one built from the same expression (kldecomp.cu): read
      // it (coalesced over tid) instead of n pow() calls in the few threads of flagged directions
      // while the rest of the CTA waits -- 19 % of the kernel's stall samples in round 2
      double a = 0.0;
      for (int i = 0; i < n; ++i) a = fma(pt.Cfull[s.idx[i] * ld + tid], s.wh[i], a);
      s.fall[tid] = a;
    }
    __syncthreads();
    if (tid < nall) {
      double a = 0.0;
      for (int e = 0; e < nall; ++e) a = fma(pt.Pfull[e * ld + tid], s.fall[e], a);
      white_out[tid] = a;
      resid_out[tid] = s.ph[tid] - (amp ? pow(10.0, s.fall[tid]) : s.fall[tid]);   // :576-583
    }
  }
  __syncthreads();
  return true;
}

template <int FT>
__global__ void __launch_bounds__(FT)
fit_pass_kernel(const double *__restrict__ phase, const float *__restrict__ weights,
                const int32_t *__restrict__ pat, const double *__restrict__ pp, int nall, int64_t nsys,
                const double *__restrict__ C, const double *__restrict__ P, const double *__restrict__ U,
                double r0sq, double half_beta, double cut, int mode, double *__restrict__ order,
                const double *__restrict__ station_order, double *__restrict__ white,
                double *__restrict__ resid, int32_t *__restrict__ status) {
  SSF_DYN_SMEM(double, sm);
  __shared__ int sh_n, sh_k, sh_go;
  __shared__ unsigned sh_vote[FT / 32];
  __shared__ double sh_t1[FT], sh_t2[FT];
  const int tid = threadIdx.x;
  Smem s = carve(sm, nall);

  for (int64_t sys = blockIdx.x; sys < nsys; sys += gridDim.x) {
    __syncthreads();
    if (tid < nall) {
      s.ph[tid] = phase[sys * nall + tid];
      s.w[tid] = (double)weights[sys * nall + tid];
    }
    // compact list of the unflagged directions, in direction order: warp votes + popcounts
    const bool keep = tid < nall && s.w[tid] > 0.0;
    const unsigned vote = __ballot_sync(0xffffffffu, keep);
    if ((tid & 31) == 0) sh_vote[tid >> 5] = vote;
    __syncthreads();
    {
      int before = 0, all = 0;
      for (int w = 0; w < FT / 32; ++w) {
        const int c = __popc(sh_vote[w]);
        if (w < (tid >> 5)) before += c;
        all += c;
      }
      if (keep) s.idx[before + __popc(vote & ((1u << (tid & 31)) - 1u))] = tid;
      if (tid == 0) sh_n = all;
    }
    __syncthreads();
    const int n = sh_n;
    if (n == 0) continue;  // :683-684
    // no slot: the pattern cache overflowed in this pass (ssf_kl_patterns_update sets its
    // overflow counter, which the caller checks); never fit against another pattern's matrices
    if (pat[sys] < 0) continue;
    Pattern pt;
    const size_t off = (size_t)pat[sys] * nall * nall;
    pt.C = C + off;
    pt.P = P + off;
    pt.U = U + off;
    pt.Pfull = P;
    pt.Cfull = C;
    pt.ld = nall;
    double *w_out = white + sys * nall;
    double *r_out = resid + sys * nall;

    double ord = order[sys];
    if (ord > (double)(n - 1)) ord = (double)(n - 1);  // :685-686
    const double st_order = station_order[sys];

    const int kind = (mode & 2) ? 1 : ((mode & 4) ? 2 : 0);
    const bool amp = kind == 1;
    if ((mode & 1) == 0) {
      bool ok = fit_once<FT>(s, pt, pp, nall, n, (int)ord, r0sq, half_beta, cut, w_out, r_out, kind);
      if (!ok && tid == 0) atomicExch(status, 1);
      if (tid == 0) order[sys] = ord;
      continue;
    }

    // mode 1: order-adjust loop (:692-779).  The reference's flagged weights alias its
    // "previous" weights, so the first order iteration never refits (:694-705).
    bool hit_upper = false, hit_lower = false, hit_upper2 = false, hit_lower2 = false;
    double sign = 1.0, prev_redchi2 = 0.0;
    for (int oi = 0; oi < 4; ++oi) {
      if (oi > 0) {
        bool ok = fit_once<FT>(s, pt, pp, nall, n, (int)ord, r0sq, half_beta, cut, w_out, r_out, kind);
        if (!ok && tid == 0) atomicExch(status, 1);
      }
      if (hit_lower2 || hit_upper2) break;
      // redchi2 = circ_chi2(resid, w) / (n_unfl - order)   (:729-731, :353-387)
      __syncthreads();
      if (tid < nall) {
        double r = r_out[tid];
        if (amp) {
          // log residual (:732-741): over ALL directions, NaN/inf times a zero weight included
          const double v = s.ph[tid];
          const double diff = log10(v) - log10(fabs(v - r));
          sh_t1[tid] = diff * diff * s.w[tid];
        } else if (kind == 2) {
          sh_t1[tid] = r * r * s.w[tid];                          // tec: plain chi2 (:742-746)
        } else {
          double sn, cs;
          sincos(r, &sn, &cs);
          sh_t1[tid] = s.w[tid] * sn * sn;
          sh_t2[tid] = s.w[tid] * cs * cs;
        }
      }
      __syncthreads();
      if (tid == 0) {
        double chi2;
        if (kind != 0) {
          chi2 = 0.0;
          for (int d = 0; d < nall; ++d) chi2 += sh_t1[d];
        } else {
          double a1 = 0.0, a2 = 0.0, sw = 0.0;
          for (int i = 0; i < n; ++i) {
            int d = s.idx[i];
            a1 += sh_t1[d];
            a2 += sh_t2[d];
            sw += s.w[d];
          }
          chi2 = (1.0 - hypot(a1 / sw, a2 / sw)) * sw;
        }
        double redchi2 = chi2 / ((double)n - ord);
        if (oi > 0) {
          if (redchi2 > 1.0 && prev_redchi2 < redchi2) sign = -sign;
          if (redchi2 < 1.0 && prev_redchi2 > redchi2) sign = -sign;
        }
        prev_redchi2 = redchi2;
        double factor = pow((double)n - ord, 0.2);
        double target = ord - sign * factor * (1.0 - redchi2);   // :753-756
        target = fmax(st_order, target);                          // :757
        double t_int = fmin(rint(target), (double)(n - 1));       // :758-760 (round half even)
        if (t_int <= 0.0) t_int = fmin(st_order, (double)(n - 1));
        int go = 1;
        if (t_int == ord) {
          go = 0;                                                 // :763-766
        } else {
          if (t_int == (double)(n - 1)) {
            if (hit_upper) hit_upper2 = true;
            hit_upper = true;
          }
          if (t_int == st_order) {
            if (hit_lower) hit_lower2 = true;
            hit_lower = true;
          }
          ord = t_int;
        }
        sh_go = go;
        sh_k = (int)ord;
        sh_t1[0] = ord;
        sh_t2[0] = (hit_upper2 || hit_lower2) ? 1.0 : 0.0;
      }
      __syncthreads();
      const int go = sh_go;
      ord = sh_t1[0];
      const bool hit2 = sh_t2[0] != 0.0;
      __syncthreads();
      hit_upper2 = hit2;  // only their OR is used by the other threads
      hit_lower2 = false;
      if (tid == 0) {
        // keep thread 0's private copies coherent with what it computed
        hit_upper2 = hit2;
      }
      if (!go) break;
    }
    if (tid == 0) order[sys] = ord;
  }
}

// _flag_outliers (stationscreen.py:303-350) per system; one warp per system.
__global__ void __launch_bounds__(128)
flag_kernel(const double *__restrict__ resid, float *__restrict__ weights, int nall, int64_t nsys,
            double nsigma, uint64_t *__restrict__ masks) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t sys = warp; sys < nsys; sys += nwarps) {
    const double *r = resid + sys * nall;
    float *w = weights + sys * nall;
    double ssum = 0.0, csum = 0.0;
    int cnt = 0;
    double rn[(SSF_MAX_DIR + 31) / 32];
#pragma unroll
    for (int j = 0; j < (SSF_MAX_DIR + 31) / 32; ++j) {
      int d = lane + 32 * j;
      rn[j] = 0.0;
      if (d < nall) {
        // normalize_phase (processing_utils.py:73-98)
        double v = fmod(r[d], TWO_PI);
        if (v < -PI) v += TWO_PI;
        if (v > PI) v -= TWO_PI;
        rn[j] = v;
        if (w[d] != 0.0f && v == v) {  // nanmean over the non-flagged, non-NaN entries
          double sn, cs;
          sincos(v, &sn, &cs);
          ssum += sn;
          csum += cs;
          ++cnt;
        }
      }
    }
    for (int o = 16; o > 0; o >>= 1) {
      ssum += __shfl_xor_sync(0xffffffffu, ssum, o);
      csum += __shfl_xor_sync(0xffffffffu, csum, o);
      cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    // nancircstd (processing_utils.py:124-132); cnt == 0 -> NaN -> nothing is flagged
    double rv = hypot(ssum / (double)cnt, csum / (double)cnt);
    double sd = sqrt(-2.0 * log(rv));
    uint64_t m0 = 0, m1 = 0;
#pragma unroll
    for (int j = 0; j < (SSF_MAX_DIR + 31) / 32; ++j) {
      int d = lane + 32 * j;
      if (d < nall) {
        float wd = w[d];
        if (fabs(rn[j]) > nsigma * sd) {
          wd = 0.0f;
          w[d] = 0.0f;
        }
        if (wd > 0.0f) {
          if (d < 64) m0 |= 1ull << d;
          else m1 |= 1ull << (d - 64);
        }
      }
    }
    for (int o = 16; o > 0; o >>= 1) {
      m0 |= __shfl_xor_sync(0xffffffffu, m0, o);
      m1 |= __shfl_xor_sync(0xffffffffu, m1, o);
    }
    if (lane == 0) {
      masks[2 * sys] = m0;
      masks[2 * sys + 1] = m1;
    }
  }
}


// Amplitude type (_flag_outliers, stationscreen.py:341-348, with the log residual of :664-668):
// ONE weighted rms per (freq, station, pol) group over all directions and times.  Systems are
// time-major: sys = t * ngroups + g.  One warp per group, fixed summation order.
// val == nullptr: tec type, the residual itself is clipped (:661-663).
__device__ __forceinline__ double amp_log_resid(const double *val, int64_t i, double r) {
  if (val == nullptr) return r;
  const double v = val[i];
  return log10(v) - log10(fabs(v - r));
}

__global__ void __launch_bounds__(128)
amp_std_kernel(const double *__restrict__ val, const double *__restrict__ resid,
               const float *__restrict__ weights, int nall, int64_t nsys, int64_t ngroups,
               double *__restrict__ group_std) {
  const int lane = threadIdx.x & 31;
  const int64_t g = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (g >= ngroups) return;
  const int64_t nt = nsys / ngroups;
  double acc = 0.0, sw = 0.0;
  int cnt = 0;
  for (int64_t t = 0; t < nt; ++t) {
    const int64_t base = (t * ngroups + g) * nall;
    for (int d = lane; d < nall; d += 32) {
      const double w = (double)weights[base + d];
      if (w > 0.0) {
        const double diff = amp_log_resid(val, base + d, resid[base + d]);
        acc = fma(diff * diff, w, acc);
        sw += w;
        ++cnt;
      }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    acc += __shfl_xor_sync(0xffffffffu, acc, o);
    sw += __shfl_xor_sync(0xffffffffu, sw, o);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  // no unflagged solution: _flag_outliers returns early (:329-330) -> nothing can exceed +inf
  if (lane == 0) group_std[g] = cnt > 0 ? sqrt(acc / sw) : INFINITY;
}

__global__ void __launch_bounds__(128)
amp_flag_kernel(const double *__restrict__ val, const double *__restrict__ resid, float *__restrict__ weights,
                int nall, int64_t nsys, int64_t ngroups, double nsigma, const double *__restrict__ group_std,
                uint64_t *__restrict__ masks) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t sys = warp; sys < nsys; sys += nwarps) {
    float *w = weights + sys * nall;
    const double lim = group_std ? nsigma * group_std[sys % ngroups] : INFINITY;
    uint64_t m0 = 0, m1 = 0;
    for (int d = lane; d < nall; d += 32) {
      float wd = w[d];
      if (group_std) {
        const double diff = amp_log_resid(val, sys * nall + d, resid[sys * nall + d]);
        if (fabs(diff) > lim) {
          wd = 0.0f;
          w[d] = 0.0f;
        }
      }
      if (wd > 0.0f) {
        if (d < 64) m0 |= 1ull << d;
        else m1 |= 1ull << (d - 64);
      }
    }
    for (int o = 16; o > 0; o >>= 1) {
      m0 |= __shfl_xor_sync(0xffffffffu, m0, o);
      m1 |= __shfl_xor_sync(0xffffffffu, m1, o);
    }
    if (lane == 0) {
      masks[2 * sys] = m0;
      masks[2 * sys + 1] = m1;
    }
  }
}

}  // namespace

// Grid width of ssf_kl_fit_pass in CTAs per SM (process-wide; 0 restores the default of 16).  A
// fit that runs beside the cube kernels on a side stream is better off narrow: with 1-2 CTAs per
// SM (64 threads, 21 KB of shared memory at 45 directions) it fits NEXT to the two resident
// evaluation CTAs, with 16 it takes their place as they retire (measured on the LOFAR step:
// evaluation kernel 0.871 -> 0.885-0.893 of the HBM peak inside the step; the fit call itself
// 1.2 -> 2.8 ms per 23 808 systems, still far inside the 17.6 ms it has).
extern "C" int ssf_kl_fit_ctas_per_sm(int ctas) {
  SSF_REQUIRE(ctas >= 0 && ctas <= 32, "ssf_kl_fit_ctas_per_sm: 0 (default) .. 32");
  g_fit_ctas_per_sm.store(ctas == 0 ? 16 : ctas, std::memory_order_relaxed);
  return SSF_OK;
}

extern "C" int ssf_kl_fit_pass(const double *phase, const float *weights, const int32_t *pat,
                               const double *pp, int ndir, int64_t nsys, const double *C,
                               const double *pinvC, const double *U, double r0, double beta,
                               double pinv_cut, int mode, double *order, const double *station_order,
                               double *white, double *resid, int32_t *status, void *stream) {
  SSF_REQUIRE(phase && weights && pat && pp && C && pinvC && U && order && station_order && white &&
                  resid && status,
              "ssf_kl_fit_pass: null pointer");
  SSF_REQUIRE(ndir >= 2 && ndir <= SSF_MAX_DIR, "ssf_kl_fit_pass: ndir=%d out of range [2,%d]", ndir, SSF_MAX_DIR);
  SSF_REQUIRE(mode >= 0 && mode <= 5, "ssf_kl_fit_pass: mode must be in 0..5 (bit 1 amplitude and bit 2 tec exclude each other)");
  if (nsys <= 0) return SSF_OK;
  size_t smem = sizeof(double) * smem_doubles(ndir);
  // CTAs per SM of the grid-stride fit: 16 unless ssf_kl_fit_ctas_per_sm() set another value
  const int per_sm = g_fit_ctas_per_sm.load(std::memory_order_relaxed);
  int64_t blocks = nsys < (int64_t)ssf::sm_count() * per_sm ? nsys : (int64_t)ssf::sm_count() * per_sm;
  if (ndir <= 64) {
    if (smem > 48 * 1024)
      SSF_CUDA(cudaFuncSetAttribute(fit_pass_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    SSF_LAUNCH(fit_pass_kernel<64>, (int)blocks, 64, smem, (cudaStream_t)stream, phase, weights, pat, pp, ndir,
               nsys, C, pinvC, U, r0 * r0, beta / 2.0, pinv_cut, mode, order, station_order, white, resid,
               status);
  } else {
    if (smem > 48 * 1024)
      SSF_CUDA(cudaFuncSetAttribute(fit_pass_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    SSF_LAUNCH(fit_pass_kernel<128>, (int)blocks, 128, smem, (cudaStream_t)stream, phase, weights, pat, pp, ndir,
               nsys, C, pinvC, U, r0 * r0, beta / 2.0, pinv_cut, mode, order, station_order, white, resid,
               status);
  }
  SSF_LAUNCHED();
  return SSF_OK;
}

extern "C" int ssf_kl_flag_outliers(const double *resid, float *weights, int ndir, int64_t nsys,
                                    double nsigma, uint64_t *masks_out, void *stream) {
  SSF_REQUIRE(resid && weights && masks_out, "ssf_kl_flag_outliers: null pointer");
  SSF_REQUIRE(ndir >= 1 && ndir <= SSF_MAX_DIR, "ssf_kl_flag_outliers: ndir out of range");
  if (nsys <= 0) return SSF_OK;
  int64_t blocks = (nsys + 3) / 4;
  if (blocks > ssf::sm_count() * 16) blocks = ssf::sm_count() * 16;
  SSF_LAUNCH(flag_kernel, (int)blocks, 128, 0, (cudaStream_t)stream, resid, weights, ndir, nsys, nsigma,
             masks_out);
  SSF_LAUNCHED();
  return SSF_OK;
}


extern "C" int ssf_kl_flag_outliers_amp(const double *val, const double *resid, float *weights, int ndir,
                                        int64_t nsys, int64_t ngroups, double nsigma, double *group_std,
                                        uint64_t *masks_out, void *stream) {
  SSF_REQUIRE(resid && weights && masks_out, "ssf_kl_flag_outliers_amp: null pointer");   // val may be null (tec)
  SSF_REQUIRE(ndir >= 1 && ndir <= SSF_MAX_DIR, "ssf_kl_flag_outliers_amp: ndir out of range");
  if (nsys <= 0) return SSF_OK;
  SSF_REQUIRE(ngroups >= 1 && nsys % ngroups == 0, "ssf_kl_flag_outliers_amp: nsys must be a multiple of ngroups");
  const bool flag = nsigma < INFINITY;
  SSF_REQUIRE(!flag || group_std, "ssf_kl_flag_outliers_amp: group_std scratch needed when flagging");
  if (flag) {
    SSF_LAUNCH(amp_std_kernel, ssf::div_up(ngroups * 32, 128), 128, 0, (cudaStream_t)stream, val, resid, weights,
               ndir, nsys, ngroups, group_std);
    SSF_LAUNCHED();
  }
  int64_t blocks = (nsys + 3) / 4;
  if (blocks > ssf::sm_count() * 16) blocks = ssf::sm_count() * 16;
  SSF_LAUNCH(amp_flag_kernel, (int)blocks, 128, 0, (cudaStream_t)stream, val, resid, weights, ndir, nsys, ngroups,
             nsigma, flag ? (const double *)group_std : (const double *)nullptr, masks_out);
  SSF_LAUNCHED();
  return SSF_OK;
}
