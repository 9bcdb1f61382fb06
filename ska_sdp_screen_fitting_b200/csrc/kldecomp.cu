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

constexpr int THREADS = 256;
constexpr int MAX_SWEEPS = 40;

__global__ void __launch_bounds__(THREADS)
decompose_kernel(const double *__restrict__ pp, int ndir, const uint64_t *__restrict__ masks,
                 double r0sq, double half_beta, double cut, double *__restrict__ Cout,
                 double *__restrict__ Pout, double *__restrict__ Uout, double *__restrict__ svout) {
  SSF_DYN_SMEM(double, sm);
  __shared__ int s_idx[SSF_MAX_DIR];
  __shared__ double s_c[SSF_MAX_DIR / 2 + 1], s_s[SSF_MAX_DIR / 2 + 1];
  __shared__ int s_p[SSF_MAX_DIR / 2 + 1], s_q[SSF_MAX_DIR / 2 + 1];
  __shared__ int s_rank[SSF_MAX_DIR];
  __shared__ int s_n;
  __shared__ double s_red[THREADS / 32];
  __shared__ double s_off, s_tot;

  const int pat = blockIdx.x;
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
  double *A = sm;           // [n][n]
  double *V = sm + n * n;   // [n][n]
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
  __syncthreads();
  if (n == 0) return;

  // C_S (stationscreen.py:417-426); same association as numpy: dx^2 + dy^2 (+ 0)
  for (int e = tid; e < n * n; e += THREADS) {
    int i = e / n, j = e - i * n;
    double dx = pp[2 * s_idx[i]] - pp[2 * s_idx[j]];
    double dy = pp[2 * s_idx[i] + 1] - pp[2 * s_idx[j] + 1];
    double d2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
    double q = d2 / r0sq;
    double v = q > 0.0 ? -0.5 * pow(q, half_beta) : 0.0;
    A[e] = v;
    V[e] = (i == j) ? 1.0 : 0.0;
    C[i * ndir + j] = v;
  }
  __syncthreads();

  const int m = (n + 1) & ~1;  // players (last one is a dummy when n is odd)
  const int npairs = m / 2;
  // Quadratic convergence: once the off-diagonal energy is below 1e-24 of the total, one
  // more sweep takes it to rounding level.
  bool converged = (n == 1);
  bool last = false;
  for (int sweep = 0; sweep < MAX_SWEEPS && !converged; ++sweep) {
    for (int r = 0; r < m - 1; ++r) {
      if (tid < npairs) {
        int p, q;
        if (tid == 0) {
          p = m - 1;
          q = r;
        } else {
          p = (r + tid) % (m - 1);
          q = (r - tid + (m - 1)) % (m - 1);
        }
        if (p > q) {
          int t = p;
          p = q;
          q = t;
        }
        double c = 1.0, s = 0.0;
        if (q < n) {
          double apq = A[p * n + q];
          if (apq != 0.0) {
            double app = A[p * n + p], aqq = A[q * n + q];
            double tau = (aqq - app) / (2.0 * apq);
            double t = (tau >= 0.0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
            c = 1.0 / sqrt(1.0 + t * t);
            s = t * c;
          }
        } else {
          q = -1;
        }
        s_p[tid] = p;
        s_q[tid] = q;
        s_c[tid] = c;
        s_s[tid] = s;
      }
      __syncthreads();
      // column update of A and V:  X <- X J
      for (int e = tid; e < n * npairs; e += THREADS) {
        int i = e / npairs, k = e - i * npairs;
        int q = s_q[k];
        if (q < 0) continue;
        int p = s_p[k];
        double c = s_c[k], s = s_s[k];
        double aip = A[i * n + p], aiq = A[i * n + q];
        A[i * n + p] = c * aip - s * aiq;
        A[i * n + q] = s * aip + c * aiq;
        double vip = V[i * n + p], viq = V[i * n + q];
        V[i * n + p] = c * vip - s * viq;
        V[i * n + q] = s * vip + c * viq;
      }
      __syncthreads();
      // row update of A:  A <- J^T A
      for (int e = tid; e < n * npairs; e += THREADS) {
        int k = e / n, j = e - k * n;
        int q = s_q[k];
        if (q < 0) continue;
        int p = s_p[k];
        double c = s_c[k], s = s_s[k];
        double apj = A[p * n + j], aqj = A[q * n + j];
        A[p * n + j] = c * apj - s * aqj;
        A[q * n + j] = s * apj + c * aqj;
      }
      __syncthreads();
    }
    // convergence: off-diagonal energy against total energy
    double off = 0.0, tot = 0.0;
    for (int e = tid; e < n * n; e += THREADS) {
      int i = e / n, j = e - i * n;
      double v = A[e] * A[e];
      tot += v;
      if (i != j) off += v;
    }
    for (int o = 16; o > 0; o >>= 1) {
      off += __shfl_xor_sync(0xffffffffu, off, o);
      tot += __shfl_xor_sync(0xffffffffu, tot, o);
    }
    if ((tid & 31) == 0) s_red[tid >> 5] = off;
    __syncthreads();
    if (tid == 0) {
      double a = 0.0;
      for (int w = 0; w < THREADS / 32; ++w) a += s_red[w];
      s_off = a;
    }
    __syncthreads();
    if ((tid & 31) == 0) s_red[tid >> 5] = tot;
    __syncthreads();
    if (tid == 0) {
      double a = 0.0;
      for (int w = 0; w < THREADS / 32; ++w) a += s_red[w];
      s_tot = a;
    }
    __syncthreads();
    if (last) converged = true;
    if (s_off <= 1e-24 * s_tot) last = true;
  }

  // order by |lambda| descending (ties: lower index first)
  for (int i = tid; i < n; i += THREADS) {
    double li = fabs(A[i * n + i]);
    int rank = 0;
    for (int j = 0; j < n; ++j) {
      double lj = fabs(A[j * n + j]);
      rank += (lj > li) || (lj == li && j < i);
    }
    s_rank[i] = rank;
    sv[rank] = converged ? li : nan("");
  }
  __syncthreads();
  for (int e = tid; e < n * n; e += THREADS) {
    int a = e / n, i = e - a * n;
    U[a * ndir + s_rank[i]] = V[e];
  }
  // pinv(C_S) = sum_{|lambda_i| > cut} v_i v_i^T / lambda_i
  for (int e = tid; e < n * n; e += THREADS) {
    int a = e / n, b = e - a * n;
    if (b < a) continue;
    double acc = 0.0;
    for (int i = 0; i < n; ++i) {
      double lam = A[i * n + i];
      if (fabs(lam) > cut) acc += V[a * n + i] * V[b * n + i] / lam;
    }
    P[a * ndir + b] = acc;
    P[b * ndir + a] = acc;
  }
}

}  // namespace

extern "C" int ssf_kl_decompose(const double *pp, int ndir, const uint64_t *masks, int npat,
                                double r0, double beta, double pinv_cut, double *C, double *pinvC,
                                double *U, double *sv, void *stream) {
  SSF_REQUIRE(pp && masks && C && pinvC && U && sv, "ssf_kl_decompose: null pointer");
  SSF_REQUIRE(ndir >= 1 && ndir <= SSF_MAX_DIR, "ssf_kl_decompose: ndir=%d out of range [1,%d]", ndir, SSF_MAX_DIR);
  if (npat <= 0) return SSF_OK;
  size_t smem = sizeof(double) * 2 * (size_t)ndir * ndir;
  if (smem > 48 * 1024)
    SSF_CUDA(cudaFuncSetAttribute(decompose_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  SSF_LAUNCH(decompose_kernel, npat, THREADS, smem, (cudaStream_t)stream, pp, ndir, masks, r0 * r0,
             beta / 2.0, pinv_cut, C, pinvC, U, sv);
  SSF_LAUNCHED();
  return SSF_OK;
}
