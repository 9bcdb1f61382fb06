// Basis contraction, tcgen05 variant (variant 1): the same contraction as contract_simt.cu
//
//   KL:       Phi[pix, col] = sum_d Bt[d, pix] * G[col, d];  slab = (cos, sin, cos, sin)(Phi)
//   Voronoi:  Re/Im[pix, col] = sum_d St[d, pix] * lut[col, d, 0/1]; slab = (Re, Im, Re, Im)
//
// with the multiply-accumulate moved to the 5th-generation tensor cores so that the SIMT
// pipes only run the epilogue (range reduction, MUFU sin/cos, NaN fix, 4 plane stores) and
// the kernel is bound by the 16 B/voxel store stream.
//
//   D[128 pixels, 64] (FP32, TMEM) = A[128, K] * B[64, K]^T        kind::tf32, K = ndir padded to 8
//
// FP32-grade accuracy from TF32 products (split TF32): every operand x is split with
// round-to-nearest into x = hi + lo, hi = tf32(x), lo = tf32(x - hi) (|x - hi - lo| <= 2^-24 |x|),
// and D accumulates A_lo B_lo + A_lo B_hi + A_hi B_lo + A_hi B_hi.  Both halves are exactly
// representable in TF32, so the tensor core's own operand conversion is the identity and
// every product is exact; the only rounding left is the accumulator's (see the issue loop).
//
// CTA = 8 epilogue warps + 2 producer/MMA warps (one per ring stage), one 128-pixel tile, a
// run of column tiles:
//   * the basis tile is split once and kept in shared memory in the canonical K-major
//     no-swizzle UMMA layout (8 x 16 B core matrices);
//   * producer warp s splits every second coefficient tile into stage s of a 2-stage
//     shared-memory ring, fences it to the async proxy and its lane 0 issues the 4 * K/8 tcgen05.mma into
//     a 2-stage TMEM accumulator ring, tcgen05.commit signalling an mbarrier;
//   * epilogue warp w reads TMEM lanes 32 (w % 4).. (its pixel = lane) and accumulator columns
//     32 (w / 4).. with tcgen05.ld, releases the TMEM stage, and streams the planes out: for a
//     fixed column the 32 lanes of a warp store 128 contiguous bytes per plane.
#include <stdlib.h>

#include "common.cuh"

#ifdef SSF_CPU_EMU
// The tensor-core variant has no CPU emulation: the emulator library reports "not supported".
namespace ssf {
bool contract_tc_supported(bool, int, int64_t, int64_t, const void *, const void *) { return false; }
int contract_tc(bool, const float *, const float *, int, int, int64_t, int64_t, float *, cudaStream_t) {
  set_error("tcgen05 contraction variant is not available in the CPU emulator");
  return SSF_ENOTSUP;
}
}  // namespace ssf
#else

namespace {

constexpr int TC_M = 128;                       // pixels per CTA tile = TMEM lanes
constexpr int TC_N = 64;                        // accumulator columns per tile
constexpr int EPI_WARPS = 8;
constexpr int PROD_WARPS = 2;                     // one producer / MMA-issuing warp per ring stage
constexpr int TC_THREADS = (EPI_WARPS + PROD_WARPS) * 32;
constexpr int TMEM_COLS = 2 * TC_N;             // 2 accumulator stages (power of two >= 32)

// ------------------------------------------------------------------------------------ PTX
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("{ .reg .b64 st; mbarrier.arrive.shared::cta.b64 st, [%0]; }" ::"r"(bar) : "memory");
}
// Bounded wait: a protocol error traps (the launch fails with an error) instead of hanging.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (int spin = 0; !done; ++spin) {
    asm volatile(
        "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (!done && spin > (1 << 22)) __trap();
  }
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
               : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], kind::tf32, issued by ONE thread.
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                            uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{ .reg .pred p; setp.ne.b32 p, %4, 0;\n"
      "  tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p; }"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ float to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
// 32 lanes x 32 consecutive accumulator columns -> 32 registers per thread.
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float *v) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]),
        "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]),
        "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]),
        "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor, K-major, no swizzle (cute::UMMA::SmemDescriptor):
//   [0,14) start address >> 4, [16,30) leading byte offset >> 4 (core matrix to core matrix
//   along K), [32,46) stride byte offset >> 4 (8-row group to 8-row group), [46,48) version = 1,
//   [61,64) layout type = 0.
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) |
         ((uint64_t)(sbo_bytes >> 4) << 32) | (1ull << 46);
}

// Instruction descriptor (cute::UMMA::InstrDescriptor): D = F32 (bit 4), A = B = TF32 (2 at
// bits 7 and 10), both K-major, N >> 3 at bit 17, M >> 4 at bit 24.
constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TC_N >> 3) << 17) |
                           ((uint32_t)(TC_M >> 4) << 24);

// Float offset of element (row r, k) in the canonical layout with kc4 = Kp / 4 core matrices
// per 8-row group: core matrix = 8 rows x 4 floats (128 B), K-adjacent matrices contiguous.
__device__ __forceinline__ int canon(int r, int k, int kc4) {
  return (((r >> 3) * kc4 + (k >> 2)) << 5) + ((r & 7) << 2) + (k & 3);
}

// sin/cos: explicit two-constant range reduction to [-pi, pi], then the MUFU approximations
// (abs. error < 5e-7 there) -- the same epilogue as the SIMT kernel.
__device__ __forceinline__ void sincos_reduced(float x, float *s, float *c) {
  const float k = rintf(x * 0.15915494309189535f);
  float r = fmaf(-k, 6.2831854820251465f, x);
  r = fmaf(-k, -1.7484555e-7f, r);
  *s = __sinf(r);
  *c = __cosf(r);
}

// Epilogue stores of one thread's 32 accumulator columns (its pixel, 32 slab columns): the 32
// lanes of the warp write 128 contiguous bytes per plane.  CHECK = ragged last tile.
template <bool CHECK>
__device__ __forceinline__ void store_kl(const float (&v)[32], float *out, int64_t npix, int nvalid) {
  const int64_t p2 = 2 * npix, p3 = 3 * npix, p4 = 4 * npix;
#pragma unroll
  for (int j = 0; j < 32; ++j) {                             // fully unrolled: v[j] stays in registers
    if (CHECK && j >= nvalid) break;
    float sn, cs;
    sincos_reduced(v[j], &sn, &cs);
    cs = ssf::fix_re(cs);
    sn = ssf::fix_im(sn);
    __stcs(out, cs);
    __stcs(out + npix, sn);
    __stcs(out + p2, cs);
    __stcs(out + p3, sn);
    out += p4;
  }
}
template <bool CHECK>
__device__ __forceinline__ void store_vo(const float (&v)[32], float *out, int64_t npix, int nvalid, int h) {
  const int64_t p2 = 2 * npix, p4 = 4 * npix;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    if (CHECK && j >= nvalid) break;
    const float x = h == 0 ? ssf::fix_re(v[j]) : ssf::fix_im(v[j]);
    __stcs(out, x);
    __stcs(out + p2, x);
    out += p4;
  }
}

// ------------------------------------------------------------------------------------ kernel
// Shared memory (floats): A_hi[128 Kp] A_lo[128 Kp] B[2 stages][hi, lo][64 Kp], then 4
// mbarriers (tmem_full[2], tmem_empty[2]) and the TMEM base address.
template <bool KL>
__global__ void __launch_bounds__(TC_THREADS, 2)
contract_tc_kernel(const float *__restrict__ Bt, const float *__restrict__ G, int ndir, int ldg,
                   int64_t npix, int64_t ncol, float *__restrict__ slab, int tiles_per_chunk, int Kp) {
  constexpr int CT = KL ? TC_N : TC_N / 2;       // slab columns per tile (Voronoi: re + im rows)
  SSF_DYN_SMEM(float, sm);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kc4 = Kp >> 2;
  float *a_hi = sm, *a_lo = sm + TC_M * Kp;
  float *b_base = sm + 2 * TC_M * Kp;            // stage s: hi at s*2*64*Kp, lo right after
  uint64_t *bars = reinterpret_cast<uint64_t *>(sm + (2 * TC_M + 4 * TC_N) * Kp);
  uint32_t *tmem_holder = reinterpret_cast<uint32_t *>(bars + 4);
  const uint32_t bar_full = smem_u32(bars), bar_empty = smem_u32(bars + 2);

  const int64_t pix0 = (int64_t)blockIdx.y * TC_M;
  const int64_t n_ct = (ncol + CT - 1) / CT;
  const int64_t ct_begin = (int64_t)blockIdx.x * tiles_per_chunk;
  int64_t ct_end = ct_begin + tiles_per_chunk;
  if (ct_end > n_ct) ct_end = n_ct;
  if (ct_begin >= ct_end) return;                // whole CTA, before any barrier / allocation
  const int n_tiles = (int)(ct_end - ct_begin);

  if (tid == 0) {
    mbar_init(bar_full, 1);
    mbar_init(bar_full + 8, 1);
    mbar_init(bar_empty, EPI_WARPS);
    mbar_init(bar_empty + 8, EPI_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == EPI_WARPS) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_holder)),
                 "n"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // basis tile: coalesced reads (pixel fastest), split, canonical layout
  // (8 loads in flight per thread: under a saturated store stream a dependent load costs us)
  for (int i0 = tid; i0 < TC_M * Kp; i0 += TC_THREADS * 8) {
    float x[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * TC_THREADS;
      const int r = i & (TC_M - 1), k = i >> 7;
      x[u] = k < ndir ? __ldg(Bt + (int64_t)k * npix + pix0 + r) : 0.0f;   // k >= ndir also covers i >= 128 Kp
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * TC_THREADS;
      if (i >= TC_M * Kp) break;
      const float hi = to_tf32(x[u]);
      const int o = canon(i & (TC_M - 1), i >> 7, kc4);
      a_hi[o] = hi;
      a_lo[o] = to_tf32(x[u] - hi);
    }
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp >= EPI_WARPS) {
    // ===== producers + MMA issuers: warp EPI_WARPS + s owns ring stage s (tiles it = s, s + 2, ..) =====
    const uint32_t lbo = 128, sbo = (uint32_t)kc4 * 128;
    const uint64_t da_hi = make_desc(smem_u32(a_hi), lbo, sbo), da_lo = make_desc(smem_u32(a_lo), lbo, sbo);
    const int s = warp - EPI_WARPS;
    for (int it = s; it < n_tiles; it += 2) {
      const int use = it >> 1;
      const int64_t c0 = (ct_begin + it) * CT;
      float *bh = b_base + (size_t)s * 2 * TC_N * Kp, *bl = bh + TC_N * Kp;
      // the MMAs of the tile that last used this stage have finished reading it
      if (use > 0) mbar_wait(bar_full + 8 * s, (use - 1) & 1);
      if (KL) {
        // lane -> coefficient rows lane and lane + 32 (one slab column each, K contiguous in
        // G); 4 core-matrix rows (float4) of both rows in flight per batch
        const float *src[2];
        bool ok[2];
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
          const int64_t col = c0 + lane + 32 * rr;
          ok[rr] = col < ncol;
          src[rr] = G + (ok[rr] ? col : 0) * ldg;
        }
        for (int kc0 = 0; kc0 < kc4; kc0 += 6) {
          float4 v[2][6];
#pragma unroll
          for (int rr = 0; rr < 2; ++rr)
#pragma unroll
            for (int u = 0; u < 6; ++u) {
              const int kc = kc0 + u;
              v[rr][u] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (ok[rr] && kc < kc4 && 4 * kc < ldg)
                v[rr][u] = __ldg(reinterpret_cast<const float4 *>(src[rr]) + kc);
            }
#pragma unroll
          for (int u = 0; u < 6; ++u) {
            const int kc = kc0 + u;
            if (kc >= kc4) break;                              // warp-uniform
            const bool tail = 4 * kc + 3 >= ndir;              // warp-uniform: zero the K padding
#pragma unroll
            for (int rr = 0; rr < 2; ++rr) {
              float4 x = v[rr][u];
              if (tail) {
                if (4 * kc + 0 >= ndir) x.x = 0.f;
                if (4 * kc + 1 >= ndir) x.y = 0.f;
                if (4 * kc + 2 >= ndir) x.z = 0.f;
                x.w = 0.f;
              }
              const float4 hv = make_float4(to_tf32(x.x), to_tf32(x.y), to_tf32(x.z), to_tf32(x.w));
              const float4 lv = make_float4(to_tf32(x.x - hv.x), to_tf32(x.y - hv.y), to_tf32(x.z - hv.z),
                                            to_tf32(x.w - hv.w));
              const int o = canon(lane + 32 * rr, 4 * kc, kc4);
              *reinterpret_cast<float4 *>(bh + o) = hv;
              *reinterpret_cast<float4 *>(bl + o) = lv;
            }
          }
        }
      } else {
        // lut [col][d][2]: lane -> slab column c0 + lane; row lane = Re, row 32 + lane = Im
        const int64_t col = c0 + lane;
        const bool ok = col < ncol;
        const float2 *src = reinterpret_cast<const float2 *>(G) + (ok ? col : 0) * ndir;
        for (int kc0 = 0; kc0 < kc4; kc0 += 2) {
          float2 v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int k = 4 * kc0 + u;
            v[u] = (ok && k < ndir) ? __ldg(src + k) : make_float2(0.f, 0.f);
          }
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int kc = kc0 + u;                            // kc4 is even: always < kc4
            const float4 re = make_float4(v[4 * u].x, v[4 * u + 1].x, v[4 * u + 2].x, v[4 * u + 3].x);
            const float4 im = make_float4(v[4 * u].y, v[4 * u + 1].y, v[4 * u + 2].y, v[4 * u + 3].y);
            const float4 rh = make_float4(to_tf32(re.x), to_tf32(re.y), to_tf32(re.z), to_tf32(re.w));
            const float4 ih = make_float4(to_tf32(im.x), to_tf32(im.y), to_tf32(im.z), to_tf32(im.w));
            const float4 rl = make_float4(to_tf32(re.x - rh.x), to_tf32(re.y - rh.y), to_tf32(re.z - rh.z),
                                          to_tf32(re.w - rh.w));
            const float4 il = make_float4(to_tf32(im.x - ih.x), to_tf32(im.y - ih.y), to_tf32(im.z - ih.z),
                                          to_tf32(im.w - ih.w));
            const int o_re = canon(lane, 4 * kc, kc4), o_im = canon(CT + lane, 4 * kc, kc4);
            *reinterpret_cast<float4 *>(bh + o_re) = rh;
            *reinterpret_cast<float4 *>(bl + o_re) = rl;
            *reinterpret_cast<float4 *>(bh + o_im) = ih;
            *reinterpret_cast<float4 *>(bl + o_im) = il;
          }
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        // the epilogue has drained the accumulator stage this tile writes
        if (use > 0) mbar_wait(bar_empty + 8 * s, (use - 1) & 1);
        tc_fence_after();
        const uint64_t db_hi = make_desc(smem_u32(bh), lbo, sbo), db_lo = make_desc(smem_u32(bl), lbo, sbo);
        const uint32_t d_tmem = tmem_base + (uint32_t)(s * TC_N);
        const int ksteps = Kp >> 3;
        // The tensor core truncates (round toward zero) once per accumulating instruction
        // (measured: -0.4 ulp of |D| per tcgen05.mma).  So the small correction products go
        // first, while |D| is ~2^-11 of its final size and the truncation is harmless, and
        // only the K/8 main products A_hi B_hi round at full magnitude.
        // One K step = 8 TF32 = 2 core matrices = 256 B -> start-address field + 16.
        for (int ks = 0; ks < ksteps; ++ks) {
          const uint64_t adv = (uint64_t)(ks * 16);
          tc_mma_tf32(d_tmem, da_lo + adv, db_lo + adv, IDESC, ks > 0);
          tc_mma_tf32(d_tmem, da_lo + adv, db_hi + adv, IDESC, 1);
          tc_mma_tf32(d_tmem, da_hi + adv, db_lo + adv, IDESC, 1);
        }
        for (int ks = 0; ks < ksteps; ++ks) {
          const uint64_t adv = (uint64_t)(ks * 16);
          tc_mma_tf32(d_tmem, da_hi + adv, db_hi + adv, IDESC, 1);
        }
        tc_commit(bar_full + 8 * s);
      }
      __syncwarp();
    }
  } else {
    // ===== epilogue warps =====
    const int q = warp & 3, h = warp >> 2;
    const int64_t pix = pix0 + 32 * q + lane;
    for (int it = 0; it < n_tiles; ++it) {
      const int s = it & 1, use = it >> 1;
      const int64_t c0 = (ct_begin + it) * CT;
      mbar_wait(bar_full + 8 * s, use & 1);
      tc_fence_after();
      float v[32];
      tmem_ld32(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(s * TC_N + 32 * h), v);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_empty + 8 * s);
      if (KL) {
        float *out = slab + (c0 + 32 * h) * 4 * npix + pix;
        const int64_t left = ncol - (c0 + 32 * h);
        if (left >= 32) store_kl<false>(v, out, npix, 32);    // branch-free common case
        else store_kl<true>(v, out, npix, (int)left);
      } else {
        // h = 0: Re rows -> planes 0, 2; h = 1: Im rows -> planes 1, 3
        float *out = slab + c0 * 4 * npix + (int64_t)h * npix + pix;
        const int64_t left = ncol - c0;
        if (left >= 32) store_vo<false>(v, out, npix, 32, h);
        else store_vo<true>(v, out, npix, (int)left, h);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == EPI_WARPS) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS)
                 : "memory");
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

template <bool KL>
int launch_tc(const float *Bt, const float *G, int ndir, int ldg, int64_t npix, int64_t ncol, float *slab,
              cudaStream_t stream) {
  constexpr int CT = KL ? TC_N : TC_N / 2;
  const int Kp = (ndir + 7) / 8 * 8;
  const size_t smem = sizeof(float) * (size_t)(2 * TC_M + 4 * TC_N) * Kp + 64;
  auto kfn = contract_tc_kernel<KL>;
  SSF_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int pix_tiles = (int)(npix / TC_M);
  const int64_t n_ct = (ncol + CT - 1) / CT;
  const int per_sm = 2 * (smem + 1024) <= 227 * 1024 ? 2 : 1;
  int chunks = pick_chunks(pix_tiles, n_ct, 148 * per_sm);
  if (const char *e = getenv("SSF_TC_CHUNKS")) {   // experiment hook
    chunks = atoi(e);
    if (chunks < 1) chunks = 1;
    if (chunks > n_ct) chunks = (int)n_ct;
  }
  const int per = (int)((n_ct + chunks - 1) / chunks);
  const int gx = (int)((n_ct + per - 1) / per);
  dim3 grid(gx, pix_tiles);
  SSF_LAUNCH(kfn, grid, TC_THREADS, smem, stream, Bt, G, ndir, ldg, npix, ncol, slab, per, Kp);
  SSF_LAUNCHED();
  return SSF_OK;
}

}  // namespace

namespace ssf {
bool contract_tc_supported(bool kl, int ndir, int64_t npix, int64_t ncol, const void *bt, const void *slab) {
  (void)kl;
  (void)ncol;
  return ndir >= 1 && ndir <= SSF_MAX_DIR && npix % TC_M == 0 && npix / TC_M <= 65535 &&
         ((uintptr_t)bt % 16 == 0) && ((uintptr_t)slab % 16 == 0);
}
int contract_tc(bool kl, const float *Bt, const float *G, int ndir, int ldg, int64_t npix, int64_t ncol,
                float *slab, cudaStream_t stream) {
  if (kl) {
    SSF_REQUIRE(ldg % 4 == 0 && (uintptr_t)G % 16 == 0, "tcgen05 KL contraction: G must be 16 B aligned, ldg %% 4 == 0");
    return launch_tc<true>(Bt, G, ndir, ldg, npix, ncol, slab, stream);
  }
  SSF_REQUIRE((uintptr_t)G % 8 == 0, "tcgen05 Voronoi contraction: lut must be 8 B aligned");
  return launch_tc<false>(Bt, G, ndir, ldg, npix, ncol, slab, stream);
}
}  // namespace ssf
#endif  // SSF_CPU_EMU
