// Batched KL decomposition: for every flag pattern (subset S of the pierce points) build the
// structure matrix C_S, its singular vectors U (by |eigenvalue| descending) and pinv(C_S)
// with an absolute cutoff.  One CTA per pattern, parallel two-sided cyclic Jacobi in shared
// memory (round-robin pairing: m/2 disjoint rotations per step), FP64 throughout.
//
// Replaces _calculate_svd (stationscreen.py:390-430).  C_S is symmetric, zero-diagonal and
// indefinite: its SVD is its eigen-decomposition with sigma = |lambda| and U = V up to
// column signs, which cancel everywhere U is used (U_k (.) U_k^T).
#include "common.cuh"

namespace {

// threads per CTA: 256 for up to 64 directions; 512 above, where a pattern's two 100 x 100
// FP64 matrices leave room for only one CTA per SM and more warps hide the latency
constexpr int MAX_SWEEPS = 40;
constexpr int WARM_MAX_FLAGGED = 16;   // warm start (see decompose_kernel) up to this many flagged directions
static_assert(SSF_MAX_DIR <= 128, "warm start: 4 columns per lane of one warp");
// off-diagonal / total ENERGY below which one more sweep is the last (see the sweep loop)
#ifndef SSF_JACOBI_LAST
#define SSF_JACOBI_LAST 1e-24
#endif
#ifndef SSF_DECOMP_BIG
#define SSF_DECOMP_BIG 512
#endif
constexpr int BIG = SSF_DECOMP_BIG;   // threads per CTA above 64 directions

#ifdef SSF_CPU_EMU
inline double rsqrt(double x) { return 1.0 / sqrt(x); }
inline double __drcp_rn(double x) { return 1.0 / x; }
#endif

// One flag pattern (slot `pat`) by the whole CTA; `sm` = the CTA's dynamic shared memory.
template <int THREADS>
__device__ void decompose_one(double *sm, const int pat, const double *__restrict__ pp, int ndir,
                              const uint64_t *__restrict__ masks, double r0sq, double half_beta, double cut,
                              double *__restrict__ Cout, double *__restrict__ Pout, double *__restrict__ Uout,
                              double *__restrict__ svout, int32_t *__restrict__ status, const double *U0) {
  __shared__ int s_idx[SSF_MAX_DIR];
  __shared__ double s_c[2][SSF_MAX_DIR / 2 + 1], s_s[2][SSF_MAX_DIR / 2 + 1];   // double-buffered rotations
  __shared__ int s_p[2][SSF_MAX_DIR / 2 + 1], s_q[2][SSF_MAX_DIR / 2 + 1];
  __shared__ int s_rank[SSF_MAX_DIR];
  __shared__ int s_n;
  __shared__ double s_red[THREADS / 32], s_red2[THREADS / 32];
  __shared__ uchar2 s_blk[(SSF_MAX_DIR / 2) * (SSF_MAX_DIR / 2 + 1) / 2];
  __shared__ double s_off, s_tot;
  __shared__ double s_inv[SSF_MAX_DIR + 16], s_gsu[SSF_MAX_DIR + 16], s_gsa[SSF_MAX_DIR + 16];

  const int tid = threadIdx.x;
  const uint64_t m0 = masks[2 * pat], m1 = masks[2 * pat + 1];
  if (tid == 0) {
    int n = 0;
    for (int d = 0; d < ndir; ++d)
      if (((d < 64 ? m0 >> d : m1 >> (d - 64)) & 1ull) != 0) s_idx[n++] = d;
    s_n = n;
  }
  __syncthreads();
  const int n = s_n;
  const int m = (n + 1) & ~1;   // padded to an even size: the extra row / column is zero, so every
                                // rotation that touches it is the identity (apq == 0)
  const int npairs = m / 2;
  const int ld = m + 1;         // odd leading dimension: column walks (the mirror stores of the
                                // A update) hit every bank instead of 4 of them
  double *A = sm;               // [m][ld]
  double *V = sm + m * ld;      // [m][ld]
  const int warp = tid >> 5, lane = tid & 31;
  double *C = Cout + (size_t)pat * ndir * ndir;
  double *P = Pout + (size_t)pat * ndir * ndir;
  double *U = Uout + (size_t)pat * ndir * ndir;
  double *sv = svout + (size_t)pat * ndir;

  // zero the whole output block so the unused part is defined
  for (int i = tid; i < ndir * ndir; i += THREADS) {
    C[i] = 0.0;
    P[i] = 0.0;
    U[i] = 0.0;
  }
  for (int i = tid; i < ndir; i += THREADS) sv[i] = 0.0;
  // upper-triangular list of pair blocks (k1 <= k2), built once
  for (int e = tid; e < npairs * npairs; e += THREADS) {
    const int k1 = e / npairs, k2 = e - k1 * npairs;
    if (k1 <= k2) s_blk[k1 * npairs - k1 * (k1 - 1) / 2 + (k2 - k1)] = make_uchar2((unsigned char)k1, (unsigned char)k2);
  }
  __syncthreads();
  if (n == 0) return;
  const int nblocks = npairs * (npairs + 1) / 2;

  // C_S (stationscreen.py:417-426); same association as numpy: dx^2 + dy^2 (+ 0)
  for (int e = tid; e < m * m; e += THREADS) {
    int i = e / m, j = e - i * m;
    double v = 0.0;
    if (i < n && j < n) {
      double dx = pp[2 * s_idx[i]] - pp[2 * s_idx[j]];
      double dy = pp[2 * s_idx[i] + 1] - pp[2 * s_idx[j] + 1];
      double d2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
      double q = d2 / r0sq;
      v = q > 0.0 ? -0.5 * pow(q, half_beta) : 0.0;
      C[i * ndir + j] = v;
    }
    A[i * ld + j] = v;
    V[i * ld + j] = (i == j) ? 1.0 : 0.0;
  }
  __syncthreads();

  // ---- warm start (device pattern cache only) -------------------------------------------------
  // Every flag pattern is a principal submatrix of the SAME full matrix, whose eigenvectors U0
  // (columns by |lambda| descending) are in slot 0.  Deleting row j from an orthogonal Q and one
  // column leaves columns w_c with Gram matrix I - u u^T (u = row j); Gram-Schmidt over them IN
  // |lambda| ORDER has the closed form (s_c = 1 - sum_{c' < c} u_c'^2)
  //     q_c = (w_c + (u_c / s_c) sum_{c' < c} u_c' w_c') sqrt(s_c / s_{c+1}),
  // a prefix sum along each row.  The leading eigenvectors stay almost untouched and the
  // error lands on the small eigenvalues, so C_S is already nearly diagonal in this basis
  // (off-diagonal energy ~1e-4 of the total) and Jacobi needs half the sweeps of a cold start.
  // Only for up to WARM_MAX_FLAGGED flagged directions: every deflation costs a little
  // orthogonality (1e-15 per step up to ~16, 1e-13 after 60), which ends up in the eigenvalues;
  // heavily flagged patterns are small matrices and start cold.
  if (U0 != nullptr && n < ndir && n > 1 && ndir - n <= WARM_MAX_FLAGGED) {
    const int mfull = (ndir + 1) & ~1;
    const int ldf = mfull + 1;
    double *F = sm + (size_t)mfull * ldf;         // second half of the allocation: [ndir][ldf]
    int *s_col = s_rank;                          // reused before the ranking needs it
    for (int e = tid; e < ndir * ndir; e += THREADS) {
      const int i = e / ndir, c = e - i * ndir;
      F[i * ldf + c] = U0[e];
    }
    for (int c = tid; c < ndir; c += THREADS) s_col[c] = 1;   // column still active
    __syncthreads();
    for (int j = 0; j < ndir; ++j) {
      if (((j < 64 ? m0 >> j : m1 >> (j - 64)) & 1ull) != 0) continue;   // uniform: kept direction
      if (warp == 0) {
        // the dropped column: largest |Q[j][c]| (keeps s_end = u_c^2 >= 1 / n: well conditioned)
        double best = -1.0;
        int bc = 0;
        for (int c = lane; c < ndir; c += 32) {
          const double v = s_col[c] ? fabs(F[j * ldf + c]) : -1.0;
          if (v > best) {
            best = v;
            bc = c;
          }
        }
        for (int o = 16; o > 0; o >>= 1) {
          const double ov = __shfl_xor_sync(0xffffffffu, best, o);
          const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
          if (ov > best || (ov == best && oc < bc)) {
            best = ov;
            bc = oc;
          }
        }
        if (lane == 0) s_col[bc] = 0;
        __syncwarp();
        // s_c = 1 - sum_{c' < c} u_c'^2 over the surviving columns: 4 columns per lane + warp scan
        double x[4], loc = 0.0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int c = 4 * lane + q;
          const double uc = (c < ndir && s_col[c]) ? F[j * ldf + c] : 0.0;
          x[q] = uc;
          loc = fma(uc, uc, loc);
        }
        double incl = loc;
        for (int o = 1; o < 32; o <<= 1) {
          const double t = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += t;
        }
        double sc = 1.0 - (incl - loc);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int c = 4 * lane + q;
          if (c < ndir) {
            const double sn = sc - x[q] * x[q];
            s_gsu[c] = x[q];
            s_gsa[c] = x[q] / sc;
            s_inv[c] = s_col[c] ? sqrt(sc / sn) : 0.0;     // 0 marks a dead column
            sc = sn;
          }
        }
      }
      __syncthreads();
      // Gram-Schmidt in closed form, one row per thread, columns in |lambda| order
      for (int i = tid; i < ndir; i += THREADS) {
        if (i == j) continue;
        double *fr = F + i * ldf;
        double P = 0.0;
        for (int c = 0; c < ndir; ++c) {
          const double g = s_inv[c];
          if (g == 0.0) continue;
          const double wc = fr[c];
          fr[c] = fma(s_gsa[c], P, wc) * g;
          P = fma(s_gsu[c], wc, P);
        }
      }
      __syncthreads();
    }
    if (tid == 0) {                               // compact list of the surviving columns
      int k = 0;
      for (int c = 0; c < ndir; ++c)
        if (s_col[c]) s_col[k++] = c;
      s_n = n;
    }
    __syncthreads();
    // 4 x 4 register tiles with strided rows / columns (element (a, b) of tile (ti, tj) is
    // (ti + a nt, tj + b nt)): lanes walk consecutive columns, no bank conflicts
    constexpr int MAXM = THREADS >= 512 ? ((SSF_MAX_DIR + 1) & ~1) : 64;   // the 256-thread variant serves <= 64
    constexpr int NT = ((MAXM + 3) / 4 * ((MAXM + 3) / 4) + THREADS - 1) / THREADS;
    const int nt = (m + 3) / 4;
    double acc[NT][16];
    // T = C_S Q
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const int e = tid + t * THREADS;
      const int ti = e / nt, tj = e - ti * nt;
#pragma unroll
      for (int x = 0; x < 16; ++x) acc[t][x] = 0.0;
      if (ti >= nt) continue;
      int fc[4];
#pragma unroll
      for (int b = 0; b < 4; ++b) fc[b] = (tj + b * nt < n) ? s_col[tj + b * nt] : -1;
      for (int l = 0; l < n; ++l) {
        const double *fr = F + s_idx[l] * ldf;
        double av[4], fv[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) av[a] = (ti + a * nt < n) ? A[(ti + a * nt) * ld + l] : 0.0;
#pragma unroll
        for (int b = 0; b < 4; ++b) fv[b] = fc[b] >= 0 ? fr[fc[b]] : 0.0;
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[t][a * 4 + b] = fma(av[a], fv[b], acc[t][a * 4 + b]);
      }
    }
    __syncthreads();
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const int e = tid + t * THREADS;
      const int ti = e / nt, tj = e - ti * nt;
      if (ti >= nt) continue;
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b)
          if (ti + a * nt < m && tj + b * nt < m) A[(ti + a * nt) * ld + tj + b * nt] = acc[t][a * 4 + b];
    }
    __syncthreads();
    // A0 = Q^T T
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const int e = tid + t * THREADS;
      const int ti = e / nt, tj = e - ti * nt;
#pragma unroll
      for (int x = 0; x < 16; ++x) acc[t][x] = 0.0;
      if (ti >= nt) continue;
      int fc[4];
#pragma unroll
      for (int a = 0; a < 4; ++a) fc[a] = (ti + a * nt < n) ? s_col[ti + a * nt] : -1;
      for (int l = 0; l < n; ++l) {
        const double *fr = F + s_idx[l] * ldf;
        double qv[4], tv[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) qv[a] = fc[a] >= 0 ? fr[fc[a]] : 0.0;
#pragma unroll
        for (int b = 0; b < 4; ++b) tv[b] = (tj + b * nt < n) ? A[l * ld + tj + b * nt] : 0.0;
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[t][a * 4 + b] = fma(qv[a], tv[b], acc[t][a * 4 + b]);
      }
    }
    __syncthreads();
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const int e = tid + t * THREADS;
      const int ti = e / nt, tj = e - ti * nt;
      if (ti >= nt) continue;
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int i = ti + a * nt, c = tj + b * nt;
          if (i < m && c < m) {
            A[i * ld + c] = acc[t][a * 4 + b];
            // gather the compact Q (V overlaps F: through registers, stored after the barrier)
            acc[t][a * 4 + b] = (i < n && c < n) ? F[s_idx[i] * ldf + s_col[c]] : (i == c ? 1.0 : 0.0);
          }
        }
    }
    __syncthreads();
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const int e = tid + t * THREADS;
      const int ti = e / nt, tj = e - ti * nt;
      if (ti >= nt) continue;
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int i = ti + a * nt, c = tj + b * nt;
          if (i < m && c < m) V[i * ld + c] = acc[t][a * 4 + b];
        }
    }
    __syncthreads();
  }

  // Rotation of pair k of round-robin step r (m / 2 disjoint pairs), published in buffer `buf`.
  auto publish = [&](int r, int buf) {
    const int k = lane + 32 * warp;       // called by warps 0 and 1
    if (k >= npairs) return;
    int p, q;
    if (k == 0) {
      p = m - 1;
      q = r;
    } else {
      p = r + k;
      if (p >= m - 1) p -= m - 1;
      q = r - k;
      if (q < 0) q += m - 1;
    }
    if (p > q) {
      int t = p;
      p = q;
      q = t;
    }
    double c = 1.0, sn = 0.0;
    const double apq = A[p * ld + q];
    if (apq != 0.0) {
      // t = sgn(tau) / (|tau| + sqrt(1 + tau^2)), tau = (aqq - app) / (2 apq), written
      // without the slow FP64 divide / sqrt: t = apq / (d + sgn(d) hypot(d, apq))
      const double d = 0.5 * (A[q * ld + q] - A[p * ld + p]);
      const double r2 = fma(d, d, apq * apq);
      if (r2 > 0.0) {
        const double den = d + copysign(r2 * rsqrt(r2), d);
        const double t = apq * __drcp_rn(den);
        c = rsqrt(fma(t, t, 1.0));
        sn = t * c;
      }
    }
    s_p[buf][k] = p;
    s_q[buf][k] = q;
    s_c[buf][k] = c;
    s_s[buf][k] = sn;
  };

  // Quadratic convergence: once the off-diagonal energy is below 1e-24 of the total, one
  // more sweep takes it to rounding level.
  bool converged = (n == 1);
  bool last = false;
  if (!converged && warp < 2) publish(0, 0);
  __syncthreads();
  int buf = 0;
  for (int sweep = 0; sweep < MAX_SWEEPS && !converged; ++sweep) {
    for (int r = 0; r < m - 1; ++r, buf ^= 1) {
      // A <- J^T A J by 2 x 2 pair blocks: the thread of block (k1, k2), k1 <= k2, reads its 4
      // elements, applies the column rotation of pair k2 and the row rotation of pair k1,
      // and writes them and their mirror images.  No element is touched by two threads.
      for (int e = tid; e < nblocks; e += THREADS) {
        const uchar2 kk = s_blk[e];
        const int p1 = s_p[buf][kk.x], q1 = s_q[buf][kk.x], p2 = s_p[buf][kk.y], q2 = s_q[buf][kk.y];
        const double c1 = s_c[buf][kk.x], s1 = s_s[buf][kk.x], c2 = s_c[buf][kk.y], s2 = s_s[buf][kk.y];
        const double a00 = A[p1 * ld + p2], a01 = A[p1 * ld + q2], a10 = A[q1 * ld + p2], a11 = A[q1 * ld + q2];
        const double b00 = c2 * a00 - s2 * a01, b01 = s2 * a00 + c2 * a01;
        const double b10 = c2 * a10 - s2 * a11, b11 = s2 * a10 + c2 * a11;
        const double r00 = c1 * b00 - s1 * b10, r10 = s1 * b00 + c1 * b10;
        const double r01 = c1 * b01 - s1 * b11, r11 = s1 * b01 + c1 * b11;
        A[p1 * ld + p2] = r00;
        A[q1 * ld + q2] = r11;
        if (kk.x == kk.y) {
          A[p1 * ld + q1] = r01;                     // the annihilated element (rounding level)
          A[q1 * ld + p1] = r01;
        } else {
          A[p1 * ld + q2] = r01;
          A[q1 * ld + p2] = r10;
          A[p2 * ld + p1] = r00;
          A[q2 * ld + p1] = r01;
          A[p2 * ld + q1] = r10;
          A[q2 * ld + q1] = r11;
        }
      }
      __syncthreads();
      // A is final for this step: warps 0 and 1 work out the NEXT step's rotations (a long
      // dependent FP64 chain) while the others apply this step's to V.
      if (warp < 2) publish(r + 1 < m - 1 ? r + 1 : 0, buf ^ 1);
      // V <- V J (columns).  Rows are owned by warps (row = warp, warp + THREADS / 32, ..): no
      // other warp ever touches them.  Lane = pair, 4 rows in flight per lane.
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        const int k = lane + 32 * g;
        if (k >= npairs) break;
        double *vp = V + s_p[buf][k], *vq = V + s_q[buf][k];
        const double c = s_c[buf][k], sn = s_s[buf][k];
        for (int i0 = warp; i0 < m; i0 += 4 * (THREADS / 32)) {
          double x[4], y[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int i = i0 + u * (THREADS / 32);
            if (i < m) {
              x[u] = vp[i * ld];
              y[u] = vq[i * ld];
            }
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int i = i0 + u * (THREADS / 32);
            if (i < m) {
              vp[i * ld] = c * x[u] - sn * y[u];
              vq[i * ld] = sn * x[u] + c * y[u];
            }
          }
        }
      }
      __syncthreads();
    }
    // convergence: off-diagonal energy against total energy
    double off = 0.0, tot = 0.0;
    for (int e = tid; e < m * m; e += THREADS) {
      int i = e / m, j = e - i * m;
      double v = A[i * ld + j] * A[i * ld + j];
      tot += v;
      if (i != j) off += v;
    }
    for (int o = 16; o > 0; o >>= 1) {
      off += __shfl_xor_sync(0xffffffffu, off, o);
      tot += __shfl_xor_sync(0xffffffffu, tot, o);
    }
    if (lane == 0) {
      s_red[warp] = off;
      s_red2[warp] = tot;
    }
    __syncthreads();
    if (tid == 0) {
      double a = 0.0, b = 0.0;
      for (int w = 0; w < THREADS / 32; ++w) {
        a += s_red[w];
        b += s_red2[w];
      }
      s_off = a;
      s_tot = b;
    }
    __syncthreads();
    if (last || s_off <= 1e-30 * s_tot) converged = true;   // 1e-30: what the extra sweep reaches
    if (s_off <= SSF_JACOBI_LAST * s_tot) last = true;
  }

  // order by |lambda| descending (ties: lower index first)
  for (int i = tid; i < n; i += THREADS) {
    double li = fabs(A[i * ld + i]);
    int rank = 0;
    for (int j = 0; j < n; ++j) {
      double lj = fabs(A[j * ld + j]);
      rank += (lj > li) || (lj == li && j < i);
    }
    s_rank[i] = rank;
    sv[rank] = converged ? li : nan("");
    if (!converged && status) *status = 1;
    s_inv[i] = li > cut ? 1.0 / A[i * ld + i] : 0.0;
  }
  __syncthreads();
  for (int e = tid; e < n * n; e += THREADS) {
    int a = e / n, i = e - a * n;
    U[a * ndir + s_rank[i]] = V[a * ld + i];
  }
  // pinv(C_S) = sum_{|lambda_i| > cut} v_i v_i^T / lambda_i
  for (int e = tid; e < n * n; e += THREADS) {
    int a = e / n, b = e - a * n;
    if (b < a) continue;
    double acc = 0.0;
    for (int i = 0; i < n; ++i) acc = fma(V[a * ld + i] * V[b * ld + i], s_inv[i], acc);
    P[a * ndir + b] = acc;
    P[b * ndir + a] = acc;
  }
}

// range == null: patterns 0 .. npat-1 (ssf_kl_decompose).  range != null: device-resident pattern
// cache -- the slots [range[1], range[0]) that ssf_kl_patterns_update has just handed out.  The
// grid strides over the patterns, so the launch needs no grid as large as the (host-unknown) upper
// bound of new patterns: a few CTAs per SM instead of one -- mostly empty -- CTA per SYSTEM of the
// pass (23 808 at the LOFAR shape), each of which had to be given a 35 KB shared-memory slot that
// an evaluation CTA could have had.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
decompose_kernel(const double *__restrict__ pp, int ndir, const uint64_t *__restrict__ masks, int npat,
                 double r0sq, double half_beta, double cut, double *__restrict__ Cout,
                 double *__restrict__ Pout, double *__restrict__ Uout, double *__restrict__ svout,
                 const int32_t *__restrict__ range, int32_t *__restrict__ status,
                 const double *U0) {
  SSF_DYN_SMEM(double, sm);
  const int first = range ? range[1] : 0, end = range ? range[0] : npat;
  for (int pat = first + (int)blockIdx.x; pat < end; pat += (int)gridDim.x) {
    decompose_one<THREADS>(sm, pat, pp, ndir, masks, r0sq, half_beta, cut, Cout, Pout, Uout, svout, status, U0);
    __syncthreads();
  }
}


// ------------------------------------------------------------------------------------ pattern cache
// Device-resident map  flag pattern (128-bit mask) -> slot of its decomposition, so that a fit
// pass needs no host round trip: open-addressing hash table, slots handed out by an atomic
// counter, new slots decomposed by the same launch sequence.
//   counters[0] = slots in use (slot 0 = the full pattern, decomposed by the caller)
//   counters[1] = slots already decomposed
//   counters[2] = 1 after an unrecoverable overflow (one pass has more patterns than slots)
//   counters[3] = systems of the current pass whose pattern is not cached yet (scratch)
constexpr int PC_EMPTY = -1, PC_BUSY = -2;

__device__ __forceinline__ bool pc_is_full(uint64_t lo, uint64_t hi, int ndir) {
  const uint64_t f0 = ndir >= 64 ? ~0ull : ((1ull << ndir) - 1);
  const uint64_t f1 = ndir <= 64 ? 0ull : (ndir >= 128 ? ~0ull : ((1ull << (ndir - 64)) - 1));
  return lo == f0 && hi == f1;
}

__device__ __forceinline__ int pc_hash(uint64_t lo, uint64_t hi, int hash_cap) {
  uint64_t h = (lo ^ (hi * 0x9E3779B97F4A7C15ull)) * 0xD6E8FEB86659FD93ull;
  h ^= h >> 32;
  return (int)(h & (uint64_t)(hash_cap - 1));
}

// Read-only probe: counts the flagged systems of this pass whose pattern is NOT in the table yet
// (an upper bound on the slots the pass can take: systems that share a new pattern count once
// each).  Nobody writes the table during this kernel.
__global__ void pc_count_kernel(const uint64_t *__restrict__ masks, int64_t nsys, int ndir,
                                const uint64_t *__restrict__ table_keys, const int32_t *__restrict__ table_slots,
                                int hash_cap, int32_t *__restrict__ counters) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int miss = 0;
  if (i < nsys) {
    const uint64_t lo = masks[2 * i], hi = masks[2 * i + 1];
    if (!pc_is_full(lo, hi, ndir)) {
      miss = 1;
      int pos = pc_hash(lo, hi, hash_cap);
      for (int probes = 0; probes < hash_cap; ++probes) {
        const int32_t s = table_slots[pos];
        if (s < 0) break;                          // empty: not cached
        if (table_keys[2 * pos] == lo && table_keys[2 * pos + 1] == hi) {
          miss = 0;
          break;
        }
        pos = (pos + 1) & (hash_cap - 1);
      }
    }
  }
#ifndef SSF_CPU_EMU
  miss = __reduce_add_sync(0xffffffffu, miss);
  if ((threadIdx.x & 31) == 0 && miss) atomicAdd(&counters[3], miss);
#else
  if (miss) atomicAdd(&counters[3], miss);
#endif
}

// If this pass could run out of slots, forget every cached pattern but the full one.  Every
// thread evaluates the same condition (nobody writes the counters during this kernel).
__global__ void pc_reset_kernel(int32_t *__restrict__ table_slots, int hash_cap, int max_slots,
                                const int32_t *__restrict__ counters) {
  if (counters[0] + counters[3] <= max_slots) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < hash_cap) table_slots[i] = PC_EMPTY;
}
__global__ void pc_reset_commit_kernel(int max_slots, int32_t *__restrict__ counters) {
  if (counters[0] + counters[3] > max_slots) {
    counters[0] = 1;
    counters[1] = 1;
  }
  counters[3] = 0;
}

__global__ void pc_lookup_kernel(const uint64_t *__restrict__ masks, int64_t nsys, int ndir,
                                 uint64_t *table_keys, int32_t *table_slots, int hash_cap,
                                 uint64_t *__restrict__ slot_masks, int max_slots, int32_t *counters,
                                 int32_t *__restrict__ pat) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nsys) return;
  const uint64_t lo = masks[2 * i], hi = masks[2 * i + 1];
  if (pc_is_full(lo, hi, ndir)) {
    pat[i] = 0;
    return;
  }
  int pos = pc_hash(lo, hi, hash_cap);
  volatile int32_t *vslots = table_slots;
  volatile uint64_t *vkeys = table_keys;
  // -1 = no slot: the pass holds more distinct patterns than the cache has slots (counters[2] is
  // set; ssf_kl_fit_pass skips such systems and raises its status) -- callers keep a pass to at
  // most max_slots - 1 systems, which cannot overflow
  int result = -1;
  for (int probes = 0; probes < hash_cap;) {
    const int32_t s = vslots[pos];
    if (s == PC_EMPTY) {
      if (atomicCAS(&table_slots[pos], PC_EMPTY, PC_BUSY) == PC_EMPTY) {
        const int slot = atomicAdd(&counters[0], 1);
        if (slot >= max_slots) {                 // a genuine overflow: more distinct patterns than slots
          counters[2] = 1;
          atomicSub(&counters[0], 1);
          atomicExch(&table_slots[pos], PC_EMPTY);
          break;
        }
        vkeys[2 * pos] = lo;
        vkeys[2 * pos + 1] = hi;
        slot_masks[2 * slot] = lo;
        slot_masks[2 * slot + 1] = hi;
        __threadfence();
        atomicExch(&table_slots[pos], slot);     // publish
        result = slot;
        break;
      }
      continue;                                  // somebody else claimed it: look again
    }
    if (s == PC_BUSY) continue;                  // being filled in: the key is published with the slot
    if (vkeys[2 * pos] == lo && vkeys[2 * pos + 1] == hi) {
      result = s;
      break;
    }
    pos = (pos + 1) & (hash_cap - 1);
    ++probes;
  }
  pat[i] = result;
}

__global__ void pc_commit_kernel(int32_t *__restrict__ counters) { counters[1] = counters[0]; }

}  // namespace

extern "C" int ssf_kl_decompose(const double *pp, int ndir, const uint64_t *masks, int npat,
                                double r0, double beta, double pinv_cut, double *C, double *pinvC,
                                double *U, double *sv, void *stream) {
  SSF_REQUIRE(pp && masks && C && pinvC && U && sv, "ssf_kl_decompose: null pointer");
  SSF_REQUIRE(ndir >= 1 && ndir <= SSF_MAX_DIR, "ssf_kl_decompose: ndir=%d out of range [1,%d]", ndir, SSF_MAX_DIR);
  if (npat <= 0) return SSF_OK;
  const int mmax = (ndir + 1) & ~1;
  size_t smem = sizeof(double) * 2 * (size_t)mmax * (mmax + 1);
  SSF_CUDA(cudaFuncSetAttribute(decompose_kernel<BIG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem > 48 * 1024 ? (int)smem : 48 * 1024));
  SSF_CUDA(cudaFuncSetAttribute(decompose_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem > 48 * 1024 ? (int)smem : 48 * 1024));
  if (ndir > 64) {
    SSF_LAUNCH(decompose_kernel<BIG>, npat, BIG, smem, (cudaStream_t)stream, pp, ndir, masks, npat, r0 * r0,
               beta / 2.0, pinv_cut, C, pinvC, U, sv, (const int32_t *)nullptr, (int32_t *)nullptr, (const double *)nullptr);
  } else {
    SSF_LAUNCH(decompose_kernel<256>, npat, 256, smem, (cudaStream_t)stream, pp, ndir, masks, npat, r0 * r0,
               beta / 2.0, pinv_cut, C, pinvC, U, sv, (const int32_t *)nullptr, (int32_t *)nullptr, (const double *)nullptr);
  }
  SSF_LAUNCHED();
  return SSF_OK;
}


extern "C" int ssf_kl_patterns_update(const uint64_t *masks, int64_t nsys, int ndir, const double *pp,
                                      double r0, double beta, double pinv_cut, uint64_t *table_keys,
                                      int32_t *table_slots, int hash_cap, uint64_t *slot_masks,
                                      int max_slots, double *C, double *pinvC, double *U, double *sv,
                                      int32_t *counters, int32_t *status, int32_t *pat, void *stream) {
  SSF_REQUIRE(masks && pp && table_keys && table_slots && slot_masks && C && pinvC && U && sv && counters && pat,
              "ssf_kl_patterns_update: null pointer");
  SSF_REQUIRE(ndir >= 1 && ndir <= SSF_MAX_DIR, "ssf_kl_patterns_update: ndir=%d out of range [1,%d]", ndir,
              SSF_MAX_DIR);
  SSF_REQUIRE(hash_cap >= 2 && (hash_cap & (hash_cap - 1)) == 0 && hash_cap >= 2 * max_slots && max_slots >= 2,
              "ssf_kl_patterns_update: hash_cap must be a power of two >= 2 * max_slots");
  if (nsys <= 0) return SSF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int nb = ssf::div_up(nsys, 256);
  SSF_LAUNCH(pc_count_kernel, nb, 256, 0, st, masks, nsys, ndir, (const uint64_t *)table_keys,
             (const int32_t *)table_slots, hash_cap, counters);
  SSF_LAUNCHED();
  SSF_LAUNCH(pc_reset_kernel, ssf::div_up(hash_cap, 256), 256, 0, st, table_slots, hash_cap, max_slots, counters);
  SSF_LAUNCHED();
  SSF_LAUNCH(pc_reset_commit_kernel, 1, 1, 0, st, max_slots, counters);
  SSF_LAUNCHED();
  SSF_LAUNCH(pc_lookup_kernel, nb, 256, 0, st, masks, nsys, ndir, table_keys, table_slots, hash_cap, slot_masks,
             max_slots, counters, pat);
  SSF_LAUNCHED();
  // decompose the slots handed out by this pass: the grid is an upper bound, CTAs beyond
  // counters[0] return at once
  const int mmax = (ndir + 1) & ~1;
  size_t smem = sizeof(double) * 2 * (size_t)mmax * (mmax + 1);
  static const bool warm = [] {                  // SSF_DECOMP_WARM=0: cold Jacobi start (experiments)
    const char *e = getenv("SSF_DECOMP_WARM");
    return !(e && e[0] == '0');
  }();
  int64_t bound = nsys < max_slots ? nsys : max_slots;
  const int64_t cap = (int64_t)ssf::sm_count() * 4;         // the kernel strides over the new slots
  if (bound > cap) bound = cap;
  if (ndir > 64) {
    SSF_CUDA(cudaFuncSetAttribute(decompose_kernel<BIG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem > 48 * 1024 ? (int)smem : 48 * 1024));
    SSF_LAUNCH(decompose_kernel<BIG>, (int)bound, BIG, smem, st, pp, ndir, slot_masks, 0, r0 * r0, beta / 2.0, pinv_cut,
               C, pinvC, U, sv, (const int32_t *)counters, status, warm ? (const double *)U : (const double *)nullptr);
  } else {
    SSF_CUDA(cudaFuncSetAttribute(decompose_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem > 48 * 1024 ? (int)smem : 48 * 1024));
    SSF_LAUNCH(decompose_kernel<256>, (int)bound, 256, smem, st, pp, ndir, slot_masks, 0, r0 * r0, beta / 2.0, pinv_cut,
               C, pinvC, U, sv, (const int32_t *)counters, status, warm ? (const double *)U : (const double *)nullptr);
  }
  SSF_LAUNCHED();
  SSF_LAUNCH(pc_commit_kernel, 1, 1, 0, st, counters);
  SSF_LAUNCHED();
  return SSF_OK;
}
