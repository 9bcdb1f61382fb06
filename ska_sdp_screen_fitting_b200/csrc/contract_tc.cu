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
// Coefficient operand: a small pre-pass (split_b_kernel) splits G / the LUT once per launch into
// hi/lo tiles stored in global memory in exactly the shared-memory image the MMA reads (canonical
// K-major no-swizzle UMMA layout, 8 x 16 B core matrices), so the hot kernel stages a tile with
// ONE bulk-async copy (cp.async.bulk, the TMA engine, mbarrier complete_tx) and no thread touches it.
//
// CTA = NEPI epilogue warps + 1 TMA warp + 1 MMA warp, one 128-pixel tile, a run of column tiles:
//   * the basis tile is split once and kept in shared memory in the canonical layout;
//   * TMA warp (one lane): bulk copies of the coefficient tiles into a 2-stage shared-memory ring,
//     each stage re-armed as soon as the MMAs that read it have retired (tcgen05.commit);
//   * MMA warp (one lane): issues the 4 * K/8 tcgen05.mma of a tile into a 4-stage TMEM accumulator
//     ring (256 columns), tcgen05.commit signalling the epilogue's mbarrier;
//   * epilogue warp w reads TMEM lanes 32 (w % 4).. (its pixel = lane) and accumulator columns
//     32 ((w / 4) % 2).. with tcgen05.ld, releases the TMEM stage, and streams the planes out: for a
//     fixed column the 32 lanes of a warp store 128 contiguous bytes per plane.  NEPI = 8 with two
//     CTAs per SM, or 16 (two groups taking alternate tiles) when only one CTA fits (ndir > 56).
#include <stdlib.h>

#include <mutex>

#include "common.cuh"

#ifdef SSF_CPU_EMU
// The tensor-core variant has no CPU emulation: the emulator library reports "not supported".
namespace ssf {
bool contract_tc_supported(bool, int, int64_t, int64_t, const void *, const void *) { return false; }
int contract_tc(bool, const float *, const float *, int, int, int64_t, int64_t, float *, OutFmt, cudaStream_t) {
  set_error("tcgen05 contraction variant is not available in the CPU emulator");
  return SSF_ENOTSUP;
}
int contract_tc_gain(const float *, const float *, int, int, int64_t, int64_t, float *, OutFmt, cudaStream_t) {
  set_error("tcgen05 contraction variant is not available in the CPU emulator");
  return SSF_ENOTSUP;
}
}  // namespace ssf
#else
#include <cuda.h>

namespace {

constexpr int TC_M = 128;                       // pixels per CTA tile = TMEM lanes
constexpr int TC_N = 64;                        // accumulator columns per tile
// Register cap of the two-CTAs-per-SM variant.  The kernel needs no more than 64 (no spills);
// what is left of the 64 K register file decides whether the FP64 fit kernels of the next slabs
// (fit_pass: 128 x 80 registers, decompose: 256 x 70) can be resident NEXT TO two evaluation
// CTAs and run in their idle issue slots, or have to wait for an evaluation CTA to retire and
// take its place.
#ifndef SSF_TC_MAXREG
#define SSF_TC_MAXREG 72
#endif
// Products of the split-TF32 scheme.  4: A_lo B_lo + A_lo B_hi + A_hi B_lo + A_hi B_hi.  3 drops
// A_lo B_lo, whose terms (<= 2^-22 |a b|, rms 2^-23.6) are of the size of the rounding residual each
// split already leaves behind (lo is itself rounded to TF32), so the error grows by < 20 % while a
// quarter of the tensor-pipe time goes away -- at 100 directions the gain contraction is bound
// by the tensor pipe (3 x 104 x 4 TF32 MACs per 16 B voxel), not by HBM.
#ifndef SSF_TC_PRODUCTS
#define SSF_TC_PRODUCTS 3
#endif
// Store policy of the cube epilogues: 0 = st.global.cs (streaming, evict first), 1 = plain, 2 = write-through.
#ifndef SSF_TC_STORE
#define SSF_TC_STORE 0
#endif
constexpr int B_MAX = 4;                        // most shared-memory stages of coefficient tiles (ring depth nb <= B_MAX)
constexpr int T_STAGES = 4;                     // TMEM accumulator ring
constexpr int TMEM_COLS = T_STAGES * TC_N;      // 256 columns (two resident CTAs use all 512)
constexpr int N_BARS = 2 * B_MAX + 2 * T_STAGES;
// Slab columns per coefficient tile.  MODE 0 = Voronoi (Re + Im rows per column),
// 1 = KL phase screens (one row per column), 2 = KL gain (phase + log10 Axx + log10 Ayy rows).
__host__ __device__ constexpr int tile_columns(int mode) { return mode == 1 ? TC_N : mode == 0 ? TC_N / 2 : 16; }
// Rows (the MMA's N) of a coefficient tile: gain tiles are 3 x 16 rows -- a 64-row tile would
// carry 16 zero rows, a quarter of the tensor-pipe time and of the operand traffic for nothing.
__host__ __device__ constexpr int tile_rows(int mode) { return mode == 2 ? 48 : TC_N; }

// ------------------------------------------------------------------------------------ PTX
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
// One lane of a converged warp (elect.sync).  Unlike `lane == 0`, the compiler knows that exactly
// one thread runs the guarded code, so the operands of tcgen05.mma / the bulk copies stay in
// uniform registers; behind `lane == 0` every such instruction became a "waterfall" loop (ELECT,
// 4 x R2UR, the instruction, BRA.U.ANY: ~25 dependent instructions, ~90 cycles per MMA -- the
// single-thread issue rate, not the tensor pipe, bounded the 100-direction and gain launches).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{ .reg .pred p; elect.sync _|p, 0xffffffff; selp.u32 %0, 1, 0, p; }" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("{ .reg .b64 st; mbarrier.arrive.shared::cta.b64 st, [%0]; }" ::"r"(bar) : "memory");
}
// Bounded wait: a protocol error traps (the launch fails with an error) instead of hanging.
// The suspend-time hint lets the hardware park the waiting thread instead of having it spin:
// without it the TMA / MMA lanes re-issued try_wait every ~50 cycles and took ~14 % of the issue
// slots of their SM sub-partition from the epilogue warps that share it (ncu capture r02e: 63 M
// try_wait executions in the TMA lane alone), so the warps of that sub-partition fell behind
// and the other six waited for them at the accumulator hand-over.
#ifndef SSF_MBAR_HINT_NS
#define SSF_MBAR_HINT_NS 20000
#endif
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (int spin = 0; !done; ++spin) {
    asm volatile(
        "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3; selp.u32 %0, 1, 0, p; }"
        : "=r"(done)
        : "r"(bar), "r"(parity), "r"((uint32_t)SSF_MBAR_HINT_NS)
        : "memory");
    if (!done && spin > (1 << 22)) __trap();
  }
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Bulk-async (TMA engine) copy global -> shared, completion counted in bytes on an mbarrier.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// TMA tiled 2-D load (tensor map): box at element coordinates (x, y) -> dense rows in shared memory.
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, int x, int y, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
               "l"(map), "r"(x), "r"(y), "r"(bar)
               : "memory");
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

// Accumulator chains.  Consecutive tcgen05.mma instructions that accumulate into the SAME tensor-
// memory columns execute one after the other, each waiting out the pipeline latency of the one
// before: ~90 cycles per instruction measured, whatever N is, where a 128 x 48 x 8 / 128 x 64 x 8
// TF32 product is only 24 / 32 cycles of tensor-pipe work (a GEMM never sees this: its N = 256
// instructions are longer than the latency).  At 100 directions the 52 instructions of a tile
// then take ~5000 cycles -- 3.5 x the time its 32 KB gain tile needs to leave the SM.  So a tile's
// K steps are dealt round-robin onto C accumulators (chain c = K steps with ks % C == c, still
// corrections first within each chain), consecutive instructions alternate between them, and
// the epilogue adds the C partial sums.
__host__ __device__ constexpr int chains_for(int mode, bool bigk) {
#ifdef SSF_TC_CHAINS
  return SSF_TC_CHAINS;
#else
  (void)mode;
  (void)bigk;
  return 1;
#endif
}

// 32 lanes x 8 consecutive accumulator columns -> 8 registers per thread.
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float *v) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}

// v[0..8) += 32 lanes x 8 consecutive accumulator columns (the partial sum of another chain).
__device__ __forceinline__ void tmem_add8(uint32_t taddr, float *v) {
  float w[8];
  tmem_ld8(taddr, w);
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] += w[i];
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
__host__ __device__ constexpr uint32_t idesc_for(int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(TC_M >> 4) << 24);
}

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

__device__ __forceinline__ void st_cube(float *p, float v) {
#if SSF_TC_STORE == 1
  *p = v;
#elif SSF_TC_STORE == 2
  __stwt(p, v);
#else
  __stcs(p, v);
#endif
}

// Epilogue stores of one thread's 32 accumulator columns (its pixel, 32 slab columns): the 32
// lanes of the warp write 128 contiguous bytes per plane.  CHECK = ragged last tile.
template <bool CHECK>
__device__ __forceinline__ void store_kl(const float (&v)[32], float *out, int64_t npix, int nvalid, ssf::OutFmt fmt) {
  const int64_t p2 = 2 * npix, p3 = 3 * npix, pc = fmt.planes * npix;
  const bool four = fmt.planes == 4;                         // uniform: predicated stores, no divergence
#pragma unroll
  for (int j = 0; j < 32; ++j) {                             // fully unrolled: v[j] stays in registers
    if (CHECK && j >= nvalid) break;
    float sn, cs;
    sincos_reduced(v[j], &sn, &cs);
    cs = ssf::out_word(ssf::fix_re(cs), fmt.prmt);
    sn = ssf::out_word(ssf::fix_im(sn), fmt.prmt);
    st_cube(out, cs);
    st_cube(out + npix, sn);
    if (four) {
      st_cube(out + p2, cs);
      st_cube(out + p3, sn);
    }
    out += pc;
  }
}
// dTEC cube (ssf_kl_eval_tec): the raw screen value, one plane per column, NaN kept.
__device__ __forceinline__ void store_raw(const float (&v)[32], float *out, int64_t npix, int nvalid, ssf::OutFmt fmt) {
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    if (j >= nvalid) break;
    st_cube(out, ssf::out_word(v[j], fmt.prmt));
    out += npix;
  }
}
template <bool CHECK>
__device__ __forceinline__ void store_vo(const float (&v)[32], float *out, int64_t npix, int nvalid, int h,
                                         ssf::OutFmt fmt) {
  const int64_t p2 = 2 * npix, pc = fmt.planes * npix;
  const bool four = fmt.planes == 4;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    if (CHECK && j >= nvalid) break;
    const float x = ssf::out_word(h == 0 ? ssf::fix_re(v[j]) : ssf::fix_im(v[j]), fmt.prmt);
    st_cube(out, x);
    if (four) st_cube(out + p2, x);
    out += pc;
  }
}

// Gain tiles: 8 slab columns per epilogue warp; phase, log10 Axx and log10 Ayy accumulators.
// 10^x as ex2.approx(x log2 10): relative error < 5e-7 for |x| < 2 (tolerance 1e-5 of the
// amplitude) in 2 instructions; exp10f's compensated product costs 8.  CHECK = ragged last tile;
// the common case is branch-free so that the loop-invariant address pieces stay hoisted.
__device__ __forceinline__ float fast_exp10(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x * 3.3219280948873623f));
  return r;
}
template <bool CHECK>
__device__ __forceinline__ void store_gain(const float (&ph)[8], const float (&ax)[8], const float (&ay)[8],
                                           float *out, int64_t npix, int nvalid, uint32_t prmt) {
  float *o1 = out + npix, *o2 = out + 2 * npix, *o3 = out + 3 * npix;
  const int64_t p4 = 4 * npix;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (CHECK && j >= nvalid) break;
    float sn, cs;
    sincos_reduced(ph[j], &sn, &cs);
    const float axx = fast_exp10(ax[j]), ayy = fast_exp10(ay[j]);
    st_cube(out, ssf::out_word(ssf::fix_re(axx * cs), prmt));
    st_cube(o1, ssf::out_word(ssf::fix_im(axx * sn), prmt));
    st_cube(o2, ssf::out_word(ssf::fix_re(ayy * cs), prmt));
    st_cube(o3, ssf::out_word(ssf::fix_im(ayy * sn), prmt));
    out += p4;
    o1 += p4;
    o2 += p4;
    o3 += p4;
  }
}

// ------------------------------------------------------------------------------------ pre-pass
// Coefficient tiles, split and laid out as the MMA wants them:
//   Bs[tile][hi | lo][canon(row, k)]   (tile_rows(MODE) rows x Kp, zero padded)
// KL:      row r of tile t = slab column 64 t + r               (G[col, d], leading dimension ldg)
// Voronoi: rows r < 32 = Re, r >= 32 = Im of column 32 t + (r & 31)   (lut[col, d, 0/1])
// Gain:    48 rows: 16 g + r, g = 0 (phase), 1 (log10 Axx), 2 (log10 Ayy) of column 16 t + r
//                                                               (G3[col, g, d], leading dimension ldg)
template <int MODE>
__global__ void split_b_kernel(const float *__restrict__ G, int ndir, int ldg, int64_t ncol, int Kp,
                               int64_t n_ct, float *__restrict__ Bs) {
  constexpr int NR = tile_rows(MODE);
  const int kc4 = Kp >> 2;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_ct * NR * kc4) return;
  const int kc = (int)(idx % kc4);
  const int r = (int)((idx / kc4) % NR);
  const int64_t tile = idx / ((int64_t)kc4 * NR);
  float x[4];
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int k = 4 * kc + u;
    float v = 0.0f;
    if (MODE == 2) {
      const int64_t col = tile * 16 + (r & 15);
      if (col < ncol && k < ndir) v = G[(col * 3 + (r >> 4)) * ldg + k];
    } else if (MODE == 1) {
      const int64_t col = tile * TC_N + r;
      if (col < ncol && k < ndir) v = G[col * ldg + k];
    } else {
      const int64_t col = tile * (TC_N / 2) + (r & 31);
      if (col < ncol && k < ndir) v = G[(col * ndir + k) * 2 + (r >> 5)];
    }
    x[u] = v;
  }
  const float4 hv = make_float4(to_tf32(x[0]), to_tf32(x[1]), to_tf32(x[2]), to_tf32(x[3]));
  const float4 lv = make_float4(to_tf32(x[0] - hv.x), to_tf32(x[1] - hv.y), to_tf32(x[2] - hv.z),
                                to_tf32(x[3] - hv.w));
  float *dst = Bs + tile * (2 * NR * Kp) + canon(r, 4 * kc, kc4);
  *reinterpret_cast<float4 *>(dst) = hv;
  *reinterpret_cast<float4 *>(dst + NR * Kp) = lv;
}

// ------------------------------------------------------------------------------------ kernel
// Shared memory (floats): A_hi[128 Kp] A_lo[128 Kp] B[nb stages][hi, lo][NR Kp], then the
// mbarriers b_full[4] b_empty[4] t_full[4] t_empty[4] and the TMEM base address.
template <int MODE, int NEPI>
__global__ void __maxnreg__(NEPI == 8 ? SSF_TC_MAXREG : 96)
contract_tc_kernel(const float *__restrict__ Bt, const float *__restrict__ Bs, int ndir, int64_t npix,
                   int64_t ncol, float *__restrict__ slab, int tiles_per_chunk, int Kp, int nb, ssf::OutFmt fmt) {
  constexpr bool KL = MODE == 1;
  constexpr int NR = tile_rows(MODE);            // rows of a coefficient tile = N of the MMA
  constexpr uint32_t IDESC = idesc_for(NR);
  constexpr int CT = tile_columns(MODE);         // slab columns per tile
  constexpr int C = chains_for(MODE, false);     // accumulator chains per tile
  constexpr int TSTRIDE = C * NR;                // TMEM columns of one tile stage
  constexpr int TS = TMEM_COLS / TSTRIDE < T_STAGES ? TMEM_COLS / TSTRIDE : T_STAGES;   // tile stages
  constexpr int THREADS = (NEPI + 2) * 32;
  constexpr int GROUPS = NEPI / 8;               // epilogue groups taking alternate tiles
  SSF_DYN_SMEM(float, sm);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kc4 = Kp >> 2;
  float *a_hi = sm, *a_lo = sm + TC_M * Kp;
  float *b_base = sm + 2 * TC_M * Kp;            // stage s: hi at s*2*64*Kp, lo right after
  uint64_t *bars = reinterpret_cast<uint64_t *>(sm + (2 * TC_M + 2 * nb * NR) * Kp);
  uint32_t *tmem_holder = reinterpret_cast<uint32_t *>(bars + N_BARS);
  const uint32_t b_full = smem_u32(bars), b_empty = smem_u32(bars + B_MAX);
  const uint32_t t_full = smem_u32(bars + 2 * B_MAX), t_empty = smem_u32(bars + 2 * B_MAX + T_STAGES);
  const uint32_t stage_bytes = (uint32_t)(2 * NR * Kp * sizeof(float));

  const int64_t pix0 = (int64_t)blockIdx.y * TC_M;
  const int64_t n_ct = (ncol + CT - 1) / CT;
  const int64_t ct_begin = (int64_t)blockIdx.x * tiles_per_chunk;
  int64_t ct_end = ct_begin + tiles_per_chunk;
  if (ct_end > n_ct) ct_end = n_ct;
  if (ct_begin >= ct_end) return;                // whole CTA, before any barrier / allocation
  const int n_tiles = (int)(ct_end - ct_begin);
  const float *b_src = Bs + ct_begin * (2 * NR * Kp);

  if (tid == NEPI * 32) {
    // the TMA lane initialises the barriers and starts the first coefficient tiles at once, so
    // their latency runs under the staging of the basis tile below
    for (int i = 0; i < B_MAX; ++i) {
      mbar_init(b_full + 8 * i, 1);
      mbar_init(b_empty + 8 * i, 1);
    }
    for (int i = 0; i < T_STAGES; ++i) {
      mbar_init(t_full + 8 * i, 1);
      mbar_init(t_empty + 8 * i, 8);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    for (int it = 0; it < nb && it < n_tiles; ++it) {
      mbar_expect_tx(b_full + 8 * it, stage_bytes);
      bulk_g2s(smem_u32(b_base + (size_t)it * 2 * NR * Kp), b_src + (size_t)it * 2 * NR * Kp, stage_bytes,
               b_full + 8 * it);
    }
  }
  if (warp == NEPI + 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_holder)),
                 "n"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // basis tile: coalesced reads (pixel fastest), split, canonical layout
  // (8 loads in flight per thread: under a saturated store stream a dependent load costs us)
  for (int i0 = tid; i0 < TC_M * Kp; i0 += THREADS * 8) {
    float x[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * THREADS;
      const int r = i & (TC_M - 1), k = i >> 7;
#ifdef SSF_TC_WRAP
      // experiment (wrong results): every CTA takes its basis tile from the first SSF_TC_WRAP pixel
      // tiles, an L2-resident window -- what the DRAM reads of the basis cost a large image
      x[u] = k < ndir ? __ldg(Bt + (int64_t)k * npix + (int64_t)(blockIdx.y % SSF_TC_WRAP) * TC_M + r) : 0.0f;
#else
      x[u] = k < ndir ? __ldg(Bt + (int64_t)k * npix + pix0 + r) : 0.0f;   // k >= ndir also covers i >= 128 Kp
#endif
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = i0 + u * THREADS;
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

  if (warp == NEPI) {
    // ===== TMA lane: refill stage s once the MMAs of the tile that used it have retired =====
    if (elect_one()) {
      for (int it = nb; it < n_tiles; ++it) {
        const int s = it % nb, use = it / nb;
        mbar_wait(b_empty + 8 * s, (use - 1) & 1);
        mbar_expect_tx(b_full + 8 * s, stage_bytes);
        bulk_g2s(smem_u32(b_base + (size_t)s * 2 * NR * Kp), b_src + (size_t)it * 2 * NR * Kp, stage_bytes,
                 b_full + 8 * s);
      }
    }
  } else if (warp == NEPI + 1) {
    // ===== MMA lane =====
    if (elect_one()) {
      const uint32_t lbo = 128, sbo = (uint32_t)kc4 * 128;
      const uint64_t da_hi = make_desc(smem_u32(a_hi), lbo, sbo), da_lo = make_desc(smem_u32(a_lo), lbo, sbo);
      const int ksteps = Kp >> 3;
      for (int it = 0; it < n_tiles; ++it) {
        const int s = it % nb, t = it % TS;
        mbar_wait(b_full + 8 * s, (it / nb) & 1);                  // coefficient tile landed
        if (it >= TS) mbar_wait(t_empty + 8 * t, (it / TS - 1) & 1);   // accumulator drained
        tc_fence_after();
        const float *bh = b_base + (size_t)s * 2 * NR * Kp;
        const uint64_t db_hi = make_desc(smem_u32(bh), lbo, sbo), db_lo = make_desc(smem_u32(bh + NR * Kp), lbo, sbo);
        const uint32_t d_tmem = tmem_base + (uint32_t)(t * TSTRIDE);
        // The tensor core truncates (round toward zero) once per accumulating instruction
        // (measured: -0.4 ulp of |D| per tcgen05.mma).  So the small correction products go
        // first, while |D| is ~2^-11 of its final size and the truncation is harmless, and
        // only the K/8 main products A_hi B_hi round at full magnitude.
        // One K step = 8 TF32 = 2 core matrices = 256 B -> start-address field + 16.
        // Chain c = K steps with ks % C == c accumulates into columns d_tmem + c NR (see
        // chains_for): consecutive instructions alternate between the chains.
        if (SSF_TC_PRODUCTS == 4)
          for (int ks = 0; ks < ksteps; ++ks) {
            const uint64_t adv = (uint64_t)(ks * 16);
            tc_mma_tf32(d_tmem + (uint32_t)((ks % C) * NR), da_lo + adv, db_lo + adv, IDESC, ks >= C);
          }
        for (int ks = 0; ks < ksteps; ++ks) {
          const uint64_t adv = (uint64_t)(ks * 16);
          tc_mma_tf32(d_tmem + (uint32_t)((ks % C) * NR), da_lo + adv, db_hi + adv, IDESC,
                      SSF_TC_PRODUCTS == 4 || ks >= C);
        }
        for (int ks = 0; ks < ksteps; ++ks) {
          const uint64_t adv = (uint64_t)(ks * 16);
          tc_mma_tf32(d_tmem + (uint32_t)((ks % C) * NR), da_hi + adv, db_lo + adv, IDESC, 1);
        }
        for (int ks = 0; ks < ksteps; ++ks) {
          const uint64_t adv = (uint64_t)(ks * 16);
          tc_mma_tf32(d_tmem + (uint32_t)((ks % C) * NR), da_hi + adv, db_hi + adv, IDESC, 1);
        }
        tc_commit(b_empty + 8 * s);                  // shared-memory stage free for the next copy
        tc_commit(t_full + 8 * t);                   // accumulator ready for the epilogue
      }
    }
  } else {
    // ===== epilogue warps =====
    const int g = warp >> 3, q = warp & 3, h = (warp >> 2) & 1;
    const int64_t pix = pix0 + 32 * q + lane;
    for (int it = g; it < n_tiles; it += GROUPS) {
      const int t = it % TS;
      const int64_t c0 = (ct_begin + it) * CT;
      mbar_wait(t_full + 8 * t, (it / TS) & 1);
      tc_fence_after();
      if (MODE == 2) {
        // this warp: its 32 pixels x slab columns c0 + 8 h .. + 7, three accumulator groups
        float ph[8], ax[8], ay[8];
        const uint32_t ta = tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(t * TSTRIDE + 8 * h);
        tmem_ld8(ta, ph);
        tmem_ld8(ta + 16, ax);
        tmem_ld8(ta + 32, ay);
        const int chains = (Kp >> 3) < C ? (Kp >> 3) : C;        // chains that hold a partial sum
        for (int c = 1; c < chains; ++c) {
          tmem_add8(ta + (uint32_t)(c * NR), ph);
          tmem_add8(ta + (uint32_t)(c * NR) + 16, ax);
          tmem_add8(ta + (uint32_t)(c * NR) + 32, ay);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(t_empty + 8 * t);
        const int64_t left = ncol - (c0 + 8 * h);
        float *out = slab + (c0 + 8 * h) * 4 * npix + pix;
        if (left >= 8) store_gain<false>(ph, ax, ay, out, npix, 8, fmt.prmt);
        else if (left > 0) store_gain<true>(ph, ax, ay, out, npix, (int)left, fmt.prmt);
        continue;
      }
      float v[32];
      tmem_ld32(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(t * TSTRIDE + 32 * h), v);
      if (C > 1) {
        const int chains = (Kp >> 3) < C ? (Kp >> 3) : C;
        for (int c = 1; c < chains; ++c)
#pragma unroll
          for (int u = 0; u < 4; ++u)
            tmem_add8(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(t * TSTRIDE + c * NR + 32 * h + 8 * u), &v[8 * u]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(t_empty + 8 * t);
      if (KL) {
        float *out = slab + (c0 + 32 * h) * fmt.planes * npix + pix;
        const int64_t left = ncol - (c0 + 32 * h);
        if (fmt.planes == 1) store_raw(v, out, npix, left >= 32 ? 32 : (int)left, fmt);
        else if (left >= 32) store_kl<false>(v, out, npix, 32, fmt);    // branch-free common case
        else store_kl<true>(v, out, npix, (int)left, fmt);
      } else {
        // h = 0: Re rows -> planes 0, 2; h = 1: Im rows -> planes 1, 3
        float *out = slab + c0 * fmt.planes * npix + (int64_t)h * npix + pix;
        const int64_t left = ncol - c0;
        if (left >= 32) store_vo<false>(v, out, npix, 32, h, fmt);
        else store_vo<true>(v, out, npix, (int)left, h, fmt);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == NEPI + 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS)
                 : "memory");
  }
}

// ------------------------------------------------------------------------------------ big-K kernel
// More than 56 directions: the split basis tile alone is > 57 KB x 2, so only one CTA fits an SM
// and the kernel above cannot hide the staging of its basis tile behind a second CTA (0.68-0.72 of
// the measured HBM peak at 100 directions).  This variant is PERSISTENT (one CTA per SM works
// through a list of (pixel tile, column chunk) items) and keeps the split basis operand in TENSOR
// MEMORY (tcgen05.mma with A from TMEM: lane = pixel row, one TF32 per 32-bit column, 2 Kp columns
// next to the 4 x 64 accumulator columns).  Shared memory then only holds the coefficient ring
// and the RAW basis tile of the NEXT item, which the TMA lane prefetches with one tensor-map copy
// while the current item is being contracted; four converter warps split it into TMEM
// (tcgen05.st) as soon as the last MMA of the current item has retired, while the 16 epilogue warps
// still have accumulator stages to drain.
//
// Items are handed out DYNAMICALLY: a CTA's first item is its block index, every further one
// comes from an atomic counter.  SMs differ in their distance to the L2 slices / HBM stacks they
// write to (two dies; the map differs from GPU to GPU), so with a static round-robin list the
// slowest SM set the launch time: 0.67 of the measured HBM peak on one box, 0.82 on another.
// The TMA lane is the scheduler: it fetches the next item one item ahead and publishes it to the
// other roles through a small shared-memory ring (sched_full / sched_empty mbarriers); item -1
// ends the walk.
//   shared memory (floats): raw[Kp][128]  B[nb stages][hi, lo][NR Kp]
//   tensor memory (columns): D[TS tile stages][C chains][NR] ... | A_hi[Kp] | A_lo[Kp] at the top
#ifndef SSF_BK_EPI
#define SSF_BK_EPI 16
#endif
constexpr int BK_EPI = SSF_BK_EPI, BK_CONV = 4;   // epilogue warps (8 or 16), converter warps
constexpr int BK_GROUPS = BK_EPI / 8;            // epilogue groups taking alternate tiles
constexpr int BK_THREADS = (BK_EPI + 2 + BK_CONV) * 32;
constexpr int BK_MAXB = 6;                       // most coefficient stages
constexpr int BK_SR = 8;                         // scheduler ring slots
constexpr int BK_NBARS = 2 * BK_MAXB + 2 * T_STAGES + 4 + 2 * BK_SR;
constexpr int BK_CONSUMERS = 1 + BK_CONV + BK_EPI;   // warps that read every scheduler slot

// D[tmem] (+)= A[tmem] * B[smem desc], kind::tf32, issued by ONE thread.
__device__ __forceinline__ void tc_mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{ .reg .pred p; setp.ne.b32 p, %4, 0;\n"
      "  tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p; }"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 32 lanes x 8 consecutive columns <- 8 registers per thread.
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const float *v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr),
               "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
               "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
               "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7]))
               : "memory");
}
// Non-blocking probe of an mbarrier phase.
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{ .reg .pred p; mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
      : "=r"(done)
      : "r"(bar), "r"(parity)
      : "memory");
  return done != 0;
}

template <int MODE>
__global__ void __launch_bounds__(BK_THREADS, 1)
contract_tc_bigk_kernel(const __grid_constant__ CUtensorMap bt_map, const float *__restrict__ Bs, int ndir,
                        int64_t npix, int64_t ncol, float *__restrict__ slab, int tiles_per_chunk, int chunks,
                        int Kp, int nb, int n_items, int *__restrict__ next_item, ssf::OutFmt fmt) {
  constexpr int NR = tile_rows(MODE);            // rows of a coefficient tile = N of the MMA
  constexpr uint32_t IDESC = idesc_for(NR);
  constexpr int C = chains_for(MODE, true);      // accumulator chains per tile
  constexpr int TSTRIDE = C * NR;                // TMEM columns of one tile stage
  const int acol = 512 - 2 * Kp;                 // first TMEM column of A_hi; tile stages below it
  const int TS = acol / TSTRIDE < T_STAGES ? acol / TSTRIDE : T_STAGES;
  const int chains = (Kp >> 3) < C ? (Kp >> 3) : C;      // chains that hold a partial sum
  SSF_DYN_SMEM(float, sm);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kc4 = Kp >> 2;
  float *raw = sm;
#ifdef SSF_BK_SS
  // experiment: split basis operand in SHARED memory (canonical layout, SS-mode MMAs) instead of TMEM
  float *a_hi_s = sm + TC_M * Kp, *a_lo_s = a_hi_s + TC_M * Kp;
  float *b_base = sm + 3 * TC_M * Kp;
  uint64_t *bars = reinterpret_cast<uint64_t *>(sm + (3 * TC_M + 2 * nb * NR) * Kp);
#else
  float *b_base = sm + TC_M * Kp;                // stage s: hi at s*2*NR*Kp, lo right after
  uint64_t *bars = reinterpret_cast<uint64_t *>(sm + (TC_M + 2 * nb * NR) * Kp);
#endif
  uint32_t *tmem_holder = reinterpret_cast<uint32_t *>(bars + BK_NBARS);
  volatile int *sched = reinterpret_cast<volatile int *>(tmem_holder + 1);     // [BK_SR] published items
  const uint32_t b_full = smem_u32(bars), b_empty = smem_u32(bars + BK_MAXB);
  const uint32_t t_full = smem_u32(bars + 2 * BK_MAXB), t_empty = smem_u32(bars + 2 * BK_MAXB + T_STAGES);
  const uint32_t raw_full = smem_u32(bars + 2 * BK_MAXB + 2 * T_STAGES), raw_free = raw_full + 8;
  const uint32_t a_ready = raw_full + 16, a_free = raw_full + 24;
  const uint32_t sched_full = raw_full + 32, sched_empty = sched_full + 8 * BK_SR;
  const uint32_t stage_bytes = (uint32_t)(2 * NR * Kp * sizeof(float));
  constexpr int CT = tile_columns(MODE);         // slab columns per coefficient tile
  const int64_t n_ct = (ncol + CT - 1) / CT;

  if (tid == BK_EPI * 32) {
    for (int i = 0; i < BK_MAXB; ++i) {
      mbar_init(b_full + 8 * i, 1);
      mbar_init(b_empty + 8 * i, 1);
    }
    for (int i = 0; i < T_STAGES; ++i) {
      mbar_init(t_full + 8 * i, 1);
      mbar_init(t_empty + 8 * i, 8);
    }
    mbar_init(raw_full, 1);
    mbar_init(raw_free, BK_CONV * 32);
    mbar_init(a_ready, BK_CONV * 32);
    mbar_init(a_free, 1);
    for (int i = 0; i < BK_SR; ++i) {
      mbar_init(sched_full + 8 * i, 1);
      mbar_init(sched_empty + 8 * i, BK_CONSUMERS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == BK_EPI + 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_holder)),
                 "n"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  // item -> (pixel tile, first column tile, tile count)
#define BK_ITEM(item, pix0, ct0, nt)                                                   \
  const int64_t pix0 = (int64_t)((item) / chunks) * TC_M;                              \
  const int64_t ct0 = (int64_t)((item) % chunks) * tiles_per_chunk;                    \
  const int nt = (int)((ct0 + tiles_per_chunk <= n_ct ? ct0 + tiles_per_chunk : n_ct) - ct0)
  // consumers: the j-th item of this CTA (every lane of the warp reads it, lane 0 frees the slot)
#define BK_NEXT_ITEM(j, item)                                                          \
  mbar_wait(sched_full + 8 * ((j) % BK_SR), ((j) / BK_SR) & 1);                        \
  const int item = sched[(j) % BK_SR];                                                 \
  __syncwarp();                                                                        \
  if (lane == 0) mbar_arrive(sched_empty + 8 * ((j) % BK_SR))

  if (warp == BK_EPI) {
    // ===== TMA lane = scheduler: next item, its raw basis tile, the coefficient tiles of this one =====
    if (elect_one()) {
      auto load_raw = [&](int item) {
        // one tensor-map copy: box = Kp direction rows x 128 pixels of Bt[ndir][npix]; rows beyond
        // ndir are out of bounds and arrive as zeros
        mbar_expect_tx(raw_full, (uint32_t)Kp * TC_M * (uint32_t)sizeof(float));
        tma_load_2d(smem_u32(raw), &bt_map, (item / chunks) * TC_M, 0, raw_full);
      };
      auto publish = [&](int j, int item) {
        const int slot = j % BK_SR;
        if (j >= BK_SR) mbar_wait(sched_empty + 8 * slot, (j / BK_SR - 1) & 1);
        sched[slot] = item;
        mbar_arrive(sched_full + 8 * slot);        // release: the store above is visible to the waiters
      };
      // the item after the next is fetched while this item's tiles are issued: the atomic's
      // round trip (a microsecond under a saturated store stream) is never waited for
      int item = blockIdx.x;
#ifdef SSF_BK_STATIC
      int next = item + (int)gridDim.x;                // experiment: static round-robin item list
#else
      int next = (int)gridDim.x + atomicAdd(next_item, 1);
#endif
      if (next >= n_items) next = -1;
      publish(0, item);
      load_raw(item);
      uint32_t T = 0;
      for (int j = 0; item >= 0; ++j) {
        int next2 = -1;
#ifdef SSF_BK_STATIC
        if (next >= 0) next2 = next + (int)gridDim.x;
#else
        if (next >= 0) next2 = (int)gridDim.x + atomicAdd(next_item, 1);
#endif
        publish(j + 1, next);
        bool raw_sent = next < 0;
        BK_ITEM(item, pix0, ct0, nt);
        (void)pix0;
        for (int it = 0; it < nt; ++it, ++T) {
          // the raw tile of the next item goes out as soon as the converters have read this one's
#ifndef SSF_BK_LATE_RAW
          if (!raw_sent && mbar_test(raw_free, j & 1)) {
            load_raw(next);
            raw_sent = true;
          }
#endif
          const int s = T % nb;
          if (T >= (uint32_t)nb) mbar_wait(b_empty + 8 * s, (T / nb - 1) & 1);
          mbar_expect_tx(b_full + 8 * s, stage_bytes);
          bulk_g2s(smem_u32(b_base + (size_t)s * 2 * NR * Kp), Bs + (ct0 + it) * (2 * NR * Kp), stage_bytes,
                   b_full + 8 * s);
        }
        if (!raw_sent) {
          mbar_wait(raw_free, j & 1);
#ifdef SSF_BK_LATE_RAW
          mbar_wait(a_free, j & 1);                    // experiment: no prefetch under the contraction
#endif
          load_raw(next);
        }
        item = next;
        next = next2 >= n_items ? -1 : next2;
      }
    }
  }
  else if (warp == BK_EPI + 1) {
    // ===== MMA lane: A from tensor memory =====
    const uint32_t lbo = 128, sbo = (uint32_t)kc4 * 128;
    const uint32_t ta_hi = tmem_base + (uint32_t)acol, ta_lo = ta_hi + (uint32_t)Kp;
    const int ksteps = Kp >> 3;
    uint32_t T = 0;
    for (int j = 0;; ++j) {
      BK_NEXT_ITEM(j, item);
      if (item < 0) break;
      BK_ITEM(item, pix0, ct0, nt);
      (void)pix0;
      (void)ct0;
      mbar_wait(a_ready, j & 1);                    // split basis tile of this item is in TMEM
      tc_fence_after();
      // the whole warp walks the items and waits; one elected lane issues (see elect_one)
      if (elect_one()) {
        for (int it = 0; it < nt; ++it) {
          const uint32_t Tl = T + (uint32_t)it;
          const int s = Tl % nb, t = Tl % TS;
          mbar_wait(b_full + 8 * s, (Tl / nb) & 1);
          if (Tl >= (uint32_t)TS) mbar_wait(t_empty + 8 * t, (Tl / TS - 1) & 1);
          tc_fence_after();
          const float *bh = b_base + (size_t)s * 2 * NR * Kp;
          const uint64_t db_hi = make_desc(smem_u32(bh), lbo, sbo), db_lo = make_desc(smem_u32(bh + NR * Kp), lbo, sbo);
          const uint32_t d_tmem = tmem_base + (uint32_t)(t * TSTRIDE);
          // corrections first, K steps dealt onto the C accumulator chains (see chains_for);
          // one K step = 8 TF32 = 8 TMEM columns of A = 16 in the descriptor's start-address field
#ifdef SSF_BK_SS
          {
            const uint64_t da_hi = make_desc(smem_u32(a_hi_s), lbo, sbo), da_lo = make_desc(smem_u32(a_lo_s), lbo, sbo);
            for (int ks = 0; ks < ksteps; ++ks)
              tc_mma_tf32(d_tmem, da_lo + (uint64_t)(ks * 16), db_hi + (uint64_t)(ks * 16), IDESC, ks > 0);
            for (int ks = 0; ks < ksteps; ++ks)
              tc_mma_tf32(d_tmem, da_hi + (uint64_t)(ks * 16), db_lo + (uint64_t)(ks * 16), IDESC, 1);
            for (int ks = 0; ks < ksteps; ++ks)
              tc_mma_tf32(d_tmem, da_hi + (uint64_t)(ks * 16), db_hi + (uint64_t)(ks * 16), IDESC, 1);
          }
#else
          if (SSF_TC_PRODUCTS == 4)
            for (int ks = 0; ks < ksteps; ++ks)
              tc_mma_tf32_ts(d_tmem + (uint32_t)((ks % C) * NR), ta_lo + (uint32_t)(ks * 8),
                             db_lo + (uint64_t)(ks * 16), IDESC, ks >= C);
          for (int ks = 0; ks < ksteps; ++ks)
            tc_mma_tf32_ts(d_tmem + (uint32_t)((ks % C) * NR), ta_lo + (uint32_t)(ks * 8), db_hi + (uint64_t)(ks * 16),
                           IDESC, SSF_TC_PRODUCTS == 4 || ks >= C);
          for (int ks = 0; ks < ksteps; ++ks)
            tc_mma_tf32_ts(d_tmem + (uint32_t)((ks % C) * NR), ta_hi + (uint32_t)(ks * 8), db_lo + (uint64_t)(ks * 16),
                           IDESC, 1);
          for (int ks = 0; ks < ksteps; ++ks)
            tc_mma_tf32_ts(d_tmem + (uint32_t)((ks % C) * NR), ta_hi + (uint32_t)(ks * 8), db_hi + (uint64_t)(ks * 16),
                           IDESC, 1);
#endif
          tc_commit(b_empty + 8 * s);
          tc_commit(t_full + 8 * t);
        }
        tc_commit(a_free);                          // the basis operand may be overwritten
      }
      T += (uint32_t)nt;
      __syncwarp();
    }
  } else if (warp >= BK_EPI + 2) {
    // ===== converter warps: raw[k][row] -> split operand in TMEM, thread = pixel row = TMEM lane =====
    const int cw = warp & 3;                          // the TMEM lane quadrant this warp may touch
    const int r = 32 * cw + lane;
    const uint32_t ta_hi = tmem_base + ((uint32_t)(32 * cw) << 16) + (uint32_t)acol, ta_lo = ta_hi + (uint32_t)Kp;
    for (int j = 0;; ++j) {
      BK_NEXT_ITEM(j, item);
      if (item < 0) break;
      mbar_wait(raw_full, j & 1);
      if (j > 0) mbar_wait(a_free, (j - 1) & 1);
      tc_fence_after();
      for (int k0 = 0; k0 < Kp; k0 += 8) {
        float hi[8], lo[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const float x = (k0 + u < ndir) ? raw[(k0 + u) * TC_M + r] : 0.0f;
          hi[u] = to_tf32(x);
          lo[u] = to_tf32(x - hi[u]);
        }
#ifdef SSF_BK_SS
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          a_hi_s[canon(r, k0 + u, kc4)] = hi[u];
          a_lo_s[canon(r, k0 + u, kc4)] = lo[u];
        }
      }
      fence_proxy_async();
      tc_fence_before();
#else
        tmem_st8(ta_hi + (uint32_t)k0, hi);
        tmem_st8(ta_lo + (uint32_t)k0, lo);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
#endif
      mbar_arrive(raw_free);
      mbar_arrive(a_ready);
    }
  } else {
    // ===== epilogue warps: group g = warp / 8 takes the tiles T with T % 2 == g =====
    const int g = warp >> 3, q = warp & 3, h = (warp >> 2) & 1;
    uint32_t T = 0;
    for (int j = 0;; ++j) {
      BK_NEXT_ITEM(j, item);
      if (item < 0) break;
      BK_ITEM(item, pix0, ct0, nt);
      const int64_t pix = pix0 + 32 * q + lane;
      for (int it = 0; it < nt; ++it, ++T) {
        if ((int)(T % BK_GROUPS) != g) continue;
        const int t = T % TS;
        mbar_wait(t_full + 8 * t, (T / TS) & 1);
        tc_fence_after();
        const int64_t c0 = (ct0 + it) * CT;
        if (MODE == 2) {
          // gain: this warp's 32 pixels x slab columns c0 + 8 h .. + 7, three accumulator groups
          float ph[8], ax[8], ay[8];
          const uint32_t ta = tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(t * TSTRIDE + 8 * h);
          tmem_ld8(ta, ph);
          tmem_ld8(ta + 16, ax);
          tmem_ld8(ta + 32, ay);
          for (int c = 1; c < chains; ++c) {
            tmem_add8(ta + (uint32_t)(c * NR), ph);
            tmem_add8(ta + (uint32_t)(c * NR) + 16, ax);
            tmem_add8(ta + (uint32_t)(c * NR) + 32, ay);
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(t_empty + 8 * t);
          const int64_t left = ncol - (c0 + 8 * h);
          float *out = slab + (c0 + 8 * h) * 4 * npix + pix;
          if (left >= 8) store_gain<false>(ph, ax, ay, out, npix, 8, fmt.prmt);
          else if (left > 0) store_gain<true>(ph, ax, ay, out, npix, (int)left, fmt.prmt);
          continue;
        }
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(t * TSTRIDE + 32 * h), v);
        for (int c = 1; c < chains; ++c)
#pragma unroll
          for (int u = 0; u < 4; ++u)
            tmem_add8(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(t * TSTRIDE + c * NR + 32 * h + 8 * u), &v[8 * u]);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(t_empty + 8 * t);
        if (MODE == 1) {
          float *out = slab + (c0 + 32 * h) * fmt.planes * npix + pix;
          const int64_t left = ncol - (c0 + 32 * h);
          if (fmt.planes == 1) store_raw(v, out, npix, left >= 32 ? 32 : (int)left, fmt);
          else if (left >= 32) store_kl<false>(v, out, npix, 32, fmt);
          else if (left > 0) store_kl<true>(v, out, npix, (int)left, fmt);
        } else {
          // Voronoi: h = 0: Re rows -> planes 0, 2; h = 1: Im rows -> planes 1, 3
          float *out = slab + c0 * fmt.planes * npix + (int64_t)h * npix + pix;
          const int64_t left = ncol - c0;
          if (left >= 32) store_vo<false>(v, out, npix, 32, h, fmt);
          else store_vo<true>(v, out, npix, (int)left, h, fmt);
        }
      }
    }
  }
#undef BK_ITEM
#undef BK_NEXT_ITEM

  tc_fence_before();
  __syncthreads();
  if (warp == BK_EPI + 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
  }
}

// ------------------------------------------------------------------------------------ narrow-tile kernel
// More than 56 directions with TWO CTAs per SM.  The persistent kernel above is bound by its
// one-CTA-per-SM structure, not by its pipeline (every one-CTA variant of either kernel
// loses 10-15 points against two CTAs at the same shape).  Two CTAs fit an SM at 100 directions
// when half of the split basis operand lives in TENSOR memory and the tiles are narrow:
//   tensor memory, 256 columns per CTA:  D[TS][NT] at column 0 | A_hi[Kp] at the top
//                                        (Kp = 104: four stages of NT = 32 accumulator columns)
//   shared memory (floats):  A_lo[128 Kp] in the canonical layout, written IN PLACE over the raw
//                            basis tile raw[Kp][128] (one TMA tensor copy)  B[nb stages][hi, lo][NT Kp]
//                                        (Kp = 104: 53 KB + 2 x 27 KB per CTA)
// A CTA = one 128-pixel tile and a run of NT-row coefficient tiles (KL: NT slab columns, Voronoi:
// NT / 2 columns x Re / Im); NT = 32 while two CTAs fit, else 16.  The eight epilogue warps first
// split the raw basis tile: each thread reads its whole share (pixel row = lane quadrant w % 4, K
// half w / 4) into registers, all meet at a named barrier, then A_hi goes to tensor memory
// (tcgen05.st) and A_lo over the raw bytes.  The products are  A_lo B_hi (both operands from shared
// memory),  A_hi B_lo and A_hi B_hi (A from tensor memory).  Then the warps work as in the kernel
// of 4.1: thread = pixel, NT / 2 accumulator columns per warp and tile.  The second resident
// CTA's store stream covers this CTA's staging.  MODE 0 = Voronoi, 1 = KL.
constexpr int NW_MAXB = 6;
constexpr int NW_NBARS = 2 * NW_MAXB + 2 * T_STAGES + 2;
template <int MODE, int NT>
__global__ void split_b_narrow_kernel(const float *__restrict__ G, int ndir, int ldg, int64_t ncol, int Kp,
                                   int64_t n_ct, float *__restrict__ Bs) {
  const int kc4 = Kp >> 2;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_ct * NT * kc4) return;
  const int kc = (int)(idx % kc4);
  const int r = (int)((idx / kc4) % NT);
  const int64_t tile = idx / ((int64_t)kc4 * NT);
  float x[4];
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int k = 4 * kc + u;
    float v = 0.0f;
    if (MODE == 1) {
      const int64_t col = tile * NT + r;
      if (col < ncol && k < ndir) v = G[col * ldg + k];
    } else {
      const int64_t col = tile * (NT / 2) + (r % (NT / 2));
      if (col < ncol && k < ndir) v = G[(col * ndir + k) * 2 + (r / (NT / 2))];
    }
    x[u] = v;
  }
  const float4 hv = make_float4(to_tf32(x[0]), to_tf32(x[1]), to_tf32(x[2]), to_tf32(x[3]));
  const float4 lv = make_float4(to_tf32(x[0] - hv.x), to_tf32(x[1] - hv.y), to_tf32(x[2] - hv.z),
                                to_tf32(x[3] - hv.w));
  float *dst = Bs + tile * (2 * NT * Kp) + canon(r, 4 * kc, kc4);
  *reinterpret_cast<float4 *>(dst) = hv;
  *reinterpret_cast<float4 *>(dst + NT * Kp) = lv;
}

// 32 lanes x 16 consecutive accumulator columns -> 16 registers per thread.
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float *v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// Epilogue stores of one thread's NC accumulator columns (its pixel, NC slab columns).
template <int NC, bool CHECK>
__device__ __forceinline__ void store_kl_n(const float (&v)[NC], float *out, int64_t npix, int nvalid, ssf::OutFmt fmt) {
  const int64_t p2 = 2 * npix, p3 = 3 * npix, pc = fmt.planes * npix;
  const bool four = fmt.planes == 4;
#pragma unroll
  for (int j = 0; j < NC; ++j) {
    if (CHECK && j >= nvalid) break;
    if (fmt.planes == 1) {                                   // dTEC cube: the raw screen value
      st_cube(out, ssf::out_word(v[j], fmt.prmt));
      out += npix;
      continue;
    }
    float sn, cs;
    sincos_reduced(v[j], &sn, &cs);
    cs = ssf::out_word(ssf::fix_re(cs), fmt.prmt);
    sn = ssf::out_word(ssf::fix_im(sn), fmt.prmt);
    st_cube(out, cs);
    st_cube(out + npix, sn);
    if (four) {
      st_cube(out + p2, cs);
      st_cube(out + p3, sn);
    }
    out += pc;
  }
}
template <int NC, bool CHECK>
__device__ __forceinline__ void store_vo_n(const float (&v)[NC], float *out, int64_t npix, int nvalid, int h,
                                           ssf::OutFmt fmt) {
  const int64_t p2 = 2 * npix, pc = fmt.planes * npix;
  const bool four = fmt.planes == 4;
#pragma unroll
  for (int j = 0; j < NC; ++j) {
    if (CHECK && j >= nvalid) break;
    const float x = ssf::out_word(h == 0 ? ssf::fix_re(v[j]) : ssf::fix_im(v[j]), fmt.prmt);
    st_cube(out, x);
    if (four) st_cube(out + p2, x);
    out += pc;
  }
}

template <int MODE, int NT>
__global__ void __launch_bounds__(320, 2)
contract_tc_narrow_kernel(const __grid_constant__ CUtensorMap bt_map, const float *__restrict__ Bs, int ndir,
                       int64_t npix, int64_t ncol, float *__restrict__ slab, int tiles_per_chunk, int Kp, int nb,
                       ssf::OutFmt fmt) {
  constexpr int NEPI = 8;
  constexpr int NC = NT / 2;                     // accumulator columns per epilogue warp
  constexpr int CT = MODE == 1 ? NT : NT / 2;    // slab columns per coefficient tile
  constexpr uint32_t IDESC = idesc_for(NT);
  constexpr int MAXKS = (SSF_MAX_DIR + 7) / 8 / 2;   // K steps per converter thread (half of <= 14)
  SSF_DYN_SMEM(float, sm);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kc4 = Kp >> 2;
  float *raw = sm;                               // [Kp][128] until the split, then A_lo in the canonical layout
  float *b_base = sm + TC_M * Kp;                // stage s: hi at s*2*NT*Kp, lo right after
  uint64_t *bars = reinterpret_cast<uint64_t *>(sm + (TC_M + 2 * nb * NT) * Kp);
  uint32_t *tmem_holder = reinterpret_cast<uint32_t *>(bars + NW_NBARS);
  const uint32_t b_full = smem_u32(bars), b_empty = smem_u32(bars + NW_MAXB);
  const uint32_t t_full = smem_u32(bars + 2 * NW_MAXB), t_empty = smem_u32(bars + 2 * NW_MAXB + T_STAGES);
  const uint32_t raw_full = smem_u32(bars + 2 * NW_MAXB + 2 * T_STAGES), a_ready = raw_full + 8;
  const uint32_t stage_bytes = (uint32_t)(2 * NT * Kp * sizeof(float));
  const int acol = 256 - Kp;                     // first TMEM column of A_hi; accumulator stages below it
  const int TS = acol / NT < T_STAGES ? acol / NT : T_STAGES;

  const int64_t pix0 = (int64_t)blockIdx.y * TC_M;
  const int64_t n_ct = (ncol + CT - 1) / CT;
  const int64_t ct_begin = (int64_t)blockIdx.x * tiles_per_chunk;
  int64_t ct_end = ct_begin + tiles_per_chunk;
  if (ct_end > n_ct) ct_end = n_ct;
  if (ct_begin >= ct_end) return;                // whole CTA, before any barrier / allocation
  const int n_tiles = (int)(ct_end - ct_begin);
  const float *b_src = Bs + ct_begin * (2 * NT * Kp);

  if (warp == NEPI) {
    if (elect_one()) {
      for (int i = 0; i < NW_MAXB; ++i) {
        mbar_init(b_full + 8 * i, 1);
        mbar_init(b_empty + 8 * i, 1);
      }
      for (int i = 0; i < T_STAGES; ++i) {
        mbar_init(t_full + 8 * i, 1);
        mbar_init(t_empty + 8 * i, NEPI);
      }
      mbar_init(raw_full, 1);
      mbar_init(a_ready, NEPI);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      mbar_expect_tx(raw_full, (uint32_t)Kp * TC_M * (uint32_t)sizeof(float));
#ifdef SSF_TC_WRAP
      tma_load_2d(smem_u32(raw), &bt_map, (int)(blockIdx.y % SSF_TC_WRAP) * TC_M, 0, raw_full);   // experiment, see contract_tc_kernel
#else
      tma_load_2d(smem_u32(raw), &bt_map, (int)pix0, 0, raw_full);
#endif
      for (int it = 0; it < nb && it < n_tiles; ++it) {
        mbar_expect_tx(b_full + 8 * it, stage_bytes);
        bulk_g2s(smem_u32(b_base + (size_t)it * 2 * NT * Kp), b_src + (size_t)it * 2 * NT * Kp, stage_bytes,
                 b_full + 8 * it);
      }
    }
  }
  if (warp == NEPI + 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(tmem_holder)),
                 "n"(256)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == NEPI) {
    // ===== TMA lane =====
    if (elect_one()) {
      for (int it = nb; it < n_tiles; ++it) {
        const int s = it % nb, use = it / nb;
        mbar_wait(b_empty + 8 * s, (use - 1) & 1);
        mbar_expect_tx(b_full + 8 * s, stage_bytes);
        bulk_g2s(smem_u32(b_base + (size_t)s * 2 * NT * Kp), b_src + (size_t)it * 2 * NT * Kp, stage_bytes,
                 b_full + 8 * s);
      }
    }
  } else if (warp == NEPI + 1) {
    // ===== MMA lane: A_lo from shared memory, A_hi from tensor memory =====
    mbar_wait(a_ready, 0);
    tc_fence_after();
    if (elect_one()) {
      const uint32_t lbo = 128, sbo = (uint32_t)kc4 * 128;
      const uint64_t da_lo = make_desc(smem_u32(raw), lbo, sbo);
      const uint32_t ta_hi = tmem_base + (uint32_t)acol;
      const int ksteps = Kp >> 3;
      for (int it = 0; it < n_tiles; ++it) {
        const int s = it % nb, t = it % TS;
        mbar_wait(b_full + 8 * s, (it / nb) & 1);
        if (it >= TS) mbar_wait(t_empty + 8 * t, (it / TS - 1) & 1);
        tc_fence_after();
        const float *bh = b_base + (size_t)s * 2 * NT * Kp;
        const uint64_t db_hi = make_desc(smem_u32(bh), lbo, sbo), db_lo = make_desc(smem_u32(bh + NT * Kp), lbo, sbo);
        const uint32_t d_tmem = tmem_base + (uint32_t)(t * NT);
        // corrections first; one K step = 8 TMEM columns of A_hi = 16 in a descriptor's start field
        if (SSF_TC_PRODUCTS == 4)
          for (int ks = 0; ks < ksteps; ++ks)
            tc_mma_tf32(d_tmem, da_lo + (uint64_t)(ks * 16), db_lo + (uint64_t)(ks * 16), IDESC, ks > 0);
        for (int ks = 0; ks < ksteps; ++ks)
          tc_mma_tf32(d_tmem, da_lo + (uint64_t)(ks * 16), db_hi + (uint64_t)(ks * 16), IDESC,
                      SSF_TC_PRODUCTS == 4 || ks > 0);
        for (int ks = 0; ks < ksteps; ++ks)
          tc_mma_tf32_ts(d_tmem, ta_hi + (uint32_t)(ks * 8), db_lo + (uint64_t)(ks * 16), IDESC, 1);
        for (int ks = 0; ks < ksteps; ++ks)
          tc_mma_tf32_ts(d_tmem, ta_hi + (uint32_t)(ks * 8), db_hi + (uint64_t)(ks * 16), IDESC, 1);
        tc_commit(b_empty + 8 * s);
        tc_commit(t_full + 8 * t);
      }
    }
    __syncwarp();
  } else {
    // ===== epilogue warps: first the split of the basis tile (A_hi -> TMEM, A_lo in place) ... =====
    const int q = warp & 3, h = (warp >> 2) & 1;
    {
      const int r = 32 * q + lane;               // pixel row = TMEM lane
      const uint32_t ta_hi = tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)acol;
      const int ksteps = Kp >> 3, half = (ksteps + 1) >> 1;
      const int ks0 = h * half, ks1 = h == 0 ? half : ksteps;
      float x[MAXKS][8];
      mbar_wait(raw_full, 0);
#pragma unroll
      for (int i = 0; i < MAXKS; ++i)
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int k = (ks0 + i) * 8 + u;
          x[i][u] = (ks0 + i < ks1 && k < ndir) ? raw[k * TC_M + r] : 0.0f;
        }
      // every epilogue thread holds its share of the raw tile: the bytes may be overwritten
      asm volatile("bar.sync 1, 256;" ::: "memory");
#pragma unroll
      for (int i = 0; i < MAXKS; ++i) {
        if (ks0 + i < ks1) {
          const int k0 = (ks0 + i) * 8;
          float hi[8];
          float4 l0, l1;
#pragma unroll
          for (int u = 0; u < 8; ++u) hi[u] = to_tf32(x[i][u]);
          l0 = make_float4(to_tf32(x[i][0] - hi[0]), to_tf32(x[i][1] - hi[1]), to_tf32(x[i][2] - hi[2]),
                           to_tf32(x[i][3] - hi[3]));
          l1 = make_float4(to_tf32(x[i][4] - hi[4]), to_tf32(x[i][5] - hi[5]), to_tf32(x[i][6] - hi[6]),
                           to_tf32(x[i][7] - hi[7]));
          tmem_st8(ta_hi + (uint32_t)k0, hi);
          *reinterpret_cast<float4 *>(raw + canon(r, k0, kc4)) = l0;
          *reinterpret_cast<float4 *>(raw + canon(r, k0 + 4, kc4)) = l1;
        }
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      fence_proxy_async();                       // A_lo is read by the tensor core (async proxy)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(a_ready);
    }
    // ===== ... then the tiles: thread = pixel, NC accumulator columns per warp =====
    const int64_t pix = pix0 + 32 * q + lane;
    for (int it = 0; it < n_tiles; ++it) {
      const int t = it % TS;
      const int64_t c0 = (ct_begin + it) * CT;
      mbar_wait(t_full + 8 * t, (it / TS) & 1);
      tc_fence_after();
      float v[NC];
      const uint32_t ta = tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(t * NT + NC * h);
      if (NC == 16) tmem_ld16(ta, v);
      else tmem_ld8(ta, v);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(t_empty + 8 * t);
      if (MODE == 1) {
        float *out = slab + (c0 + NC * h) * fmt.planes * npix + pix;
        const int64_t left = ncol - (c0 + NC * h);
        if (left >= NC) store_kl_n<NC, false>(v, out, npix, NC, fmt);
        else if (left > 0) store_kl_n<NC, true>(v, out, npix, (int)left, fmt);
      } else {
        // rows 0..NC-1 = Re (h = 0) -> planes 0, 2; rows NC.. = Im (h = 1) -> planes 1, 3
        float *out = slab + c0 * fmt.planes * npix + (int64_t)h * npix + pix;
        const int64_t left = ncol - c0;
        if (left >= NC) store_vo_n<NC, false>(v, out, npix, NC, h, fmt);
        else if (left > 0) store_vo_n<NC, true>(v, out, npix, (int)left, h, fmt);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == NEPI + 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(256) : "memory");
  }
}

// Returns pool scratch to the stream-ordered pool on every exit path of a launcher.
struct ScratchGuard {
  void *ptr;
  cudaStream_t stream;
  ~ScratchGuard() {
    if (ptr) cudaFreeAsync(ptr, stream);
  }
  int release() {
    void *p = ptr;
    ptr = nullptr;
    SSF_CUDA(cudaFreeAsync(p, stream));
    return SSF_OK;
  }
};

// Column tiles per CTA.  Measured on B200 (profiles/r01c chunk sweeps): with the coefficient
// tiles arriving by TMA the launch is fastest when a CTA covers about 256 slab columns (4 KL
// tiles / 8 Voronoi tiles, 0.5 MB of cube): 0.93 of the measured HBM peak against 0.78 for CTAs
// that walk all 93 column tiles, and 0.67 for one-tile CTAs.  Short CTAs let the dynamic CTA
// scheduler even out SMs of unequal speed; the floor is the cost of staging the basis tile
// (128 x Kp x 4 B per CTA, an L2 hit for all but the first chunk of a pixel tile), which
// grows with the direction count -- hence the Kp / 12 term.
// Gain tiles (16 columns, a full MMA set per 32 KB of cube) pay more per CTA start: measured at
// 45 directions, 16 / 48 / 96 / 186 tiles per CTA give 0.69 / 0.73 / 0.71 / 0.72 -> 768 columns.
int pick_tiles_per_cta(int ct, int Kp, int64_t n_ct) {
  int64_t per = (ct == 16 ? 768 : 256) / ct;
  if (per < Kp / 12) per = Kp / 12;
  if (per > n_ct) per = n_ct;
  return (int)(per < 1 ? 1 : per);
}

typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                             const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                             CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
// cuTensorMapEncodeTiled through the runtime's driver entry point (no -lcuda), resolved once.
EncodeFn tensor_map_encoder() {
  static std::once_flag once;
  static EncodeFn encode = nullptr;
  std::call_once(once, [] {
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      encode = reinterpret_cast<EncodeFn>(fn);
  });
  return encode;
}

template <int MODE>
int launch_bigk(const float *Bt, const float *G, int ndir, int ldg, int64_t npix, int64_t ncol, float *slab,
                ssf::OutFmt fmt, cudaMemPool_t pool, cudaStream_t stream) {
  constexpr int CT = tile_columns(MODE);
  constexpr int NR = tile_rows(MODE);
  const int Kp = (ndir + 7) / 8 * 8;
  const int64_t n_ct = (ncol + CT - 1) / CT;
  // as many coefficient stages as fit next to the raw basis tile (KL at Kp = 104: 3, gain: 4)
#ifdef SSF_BK_SS
  auto smem_for = [&](int stages) { return sizeof(float) * (size_t)(3 * TC_M + 2 * stages * NR) * Kp + 512; };
#else
  auto smem_for = [&](int stages) { return sizeof(float) * (size_t)(TC_M + 2 * stages * NR) * Kp + 512; };
#endif
  int nb = 2;
  while (nb < BK_MAXB && smem_for(nb + 1) <= 227 * 1024) ++nb;
  if (const char *e = getenv("SSF_TC_NB")) {         // experiment hook: ring depth
    const int want = atoi(e);
    if (want >= 2 && want <= BK_MAXB && smem_for(want) <= 227 * 1024) nb = want;
  }
  const size_t smem = smem_for(nb);
  SSF_REQUIRE(smem <= 227 * 1024, "tcgen05 contraction: %d directions do not fit the persistent kernel", ndir);
  // scratch: the split coefficient tiles, then the item counter of the dynamic schedule
  const size_t bs_bytes = (sizeof(float) * (size_t)n_ct * 2 * NR * Kp + 255) / 256 * 256;
  float *Bs = nullptr;
  SSF_CUDA(cudaMallocFromPoolAsync(reinterpret_cast<void **>(&Bs), bs_bytes + 256, pool, stream));
  ScratchGuard guard{Bs, stream};
  int *next_item = reinterpret_cast<int *>(reinterpret_cast<char *>(Bs) + bs_bytes);
  SSF_CUDA(cudaMemsetAsync(next_item, 0, sizeof(int), stream));
  SSF_LAUNCH(split_b_kernel<MODE>, ssf::div_up(n_ct * NR * (Kp >> 2), 256), 256, 0, stream, G, ndir, ldg, ncol, Kp,
             n_ct, Bs);
  SSF_LAUNCHED();
  const int sms = ssf::sm_count();
  // items: one per pixel tile, or column chunks of >= 4 tiles when there are few pixel tiles
  // (about 8 items per SM at least, so that the dynamic schedule has something to balance)
  const int64_t pix_tiles = npix / TC_M;
  int64_t chunks = 1;
  if (pix_tiles < 8 * sms) {
    chunks = (8 * sms + pix_tiles - 1) / pix_tiles;
    if (chunks > (n_ct + 3) / 4) chunks = (n_ct + 3) / 4;
    if (chunks < 1) chunks = 1;
  }
  if (const char *e = getenv("SSF_TC_PER")) {        // experiment hook: column tiles per item
    const int64_t want = atoi(e);
    if (want >= 1) chunks = (n_ct + want - 1) / want;
  }
  const int per = (int)((n_ct + chunks - 1) / chunks);
  chunks = (n_ct + per - 1) / per;
  const int64_t n_items = pix_tiles * chunks;
  SSF_REQUIRE(n_items < (int64_t)1 << 30, "tcgen05 contraction: too many work items");
  // tensor map of the basis: [ndir rows][npix] float32, box = Kp rows x 128 pixels, no swizzle
  const EncodeFn encode = tensor_map_encoder();
  SSF_REQUIRE(encode != nullptr, "cuTensorMapEncodeTiled is not available");
  CUtensorMap bt_map;
  const cuuint64_t gdim[2] = {(cuuint64_t)npix, (cuuint64_t)ndir};
  const cuuint64_t gstride[1] = {(cuuint64_t)npix * sizeof(float)};
  const cuuint32_t box[2] = {(cuuint32_t)TC_M, (cuuint32_t)Kp};
  const cuuint32_t estride[2] = {1, 1};
  const CUresult cr = encode(&bt_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(Bt), gdim, gstride, box,
                             estride, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SSF_REQUIRE(cr == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d)", (int)cr);
  SSF_CUDA(cudaFuncSetAttribute(contract_tc_bigk_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = (int)(n_items < sms ? n_items : sms);
  SSF_LAUNCH(contract_tc_bigk_kernel<MODE>, grid, BK_THREADS, smem, stream, bt_map, Bs, ndir, npix, ncol, slab, per,
             (int)chunks, Kp, nb, (int)n_items, next_item, fmt);
  SSF_LAUNCHED();
  return guard.release();
}

template <int MODE, int NT>
int launch_narrow_nt(const float *Bt, const float *G, int ndir, int ldg, int64_t npix, int64_t ncol, float *slab,
                  ssf::OutFmt fmt, cudaMemPool_t pool, cudaStream_t stream, int nb, size_t smem) {
  constexpr int CT = MODE == 1 ? NT : NT / 2;
  const int Kp = (ndir + 7) / 8 * 8;
  const int64_t n_ct = (ncol + CT - 1) / CT;
  const size_t bs_bytes = sizeof(float) * (size_t)n_ct * 2 * NT * Kp;
  float *Bs = nullptr;
  SSF_CUDA(cudaMallocFromPoolAsync(reinterpret_cast<void **>(&Bs), bs_bytes, pool, stream));
  ScratchGuard guard{Bs, stream};
  SSF_LAUNCH((split_b_narrow_kernel<MODE, NT>), ssf::div_up(n_ct * NT * (Kp >> 2), 256), 256, 0, stream, G, ndir, ldg,
             ncol, Kp, n_ct, Bs);
  SSF_LAUNCHED();
  const EncodeFn encode = tensor_map_encoder();
  SSF_REQUIRE(encode != nullptr, "cuTensorMapEncodeTiled is not available");
  CUtensorMap bt_map;
  const cuuint64_t gdim[2] = {(cuuint64_t)npix, (cuuint64_t)ndir};
  const cuuint64_t gstride[1] = {(cuuint64_t)npix * sizeof(float)};
  const cuuint32_t box[2] = {(cuuint32_t)TC_M, (cuuint32_t)Kp};
  const cuuint32_t estride[2] = {1, 1};
  const CUresult cr = encode(&bt_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(Bt), gdim, gstride, box,
                             estride, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SSF_REQUIRE(cr == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d)", (int)cr);
  auto kfn = contract_tc_narrow_kernel<MODE, NT>;
  SSF_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t per = 256 / CT;                        // about 256 slab columns per CTA, as for the kernel of 4.1
  if (const char *e = getenv("SSF_TC_PER")) {
    per = atoi(e);
    if (per < 1) per = 1;
  }
  if (per > n_ct) per = n_ct;
  const int gx = (int)((n_ct + per - 1) / per);
  dim3 grid(gx, (unsigned)(npix / TC_M));
  SSF_LAUNCH(kfn, grid, 320, smem, stream, bt_map, Bs, ndir, npix, ncol, slab, (int)per, Kp, nb, fmt);
  SSF_LAUNCHED();
  return guard.release();
}

// 32-row tiles while two CTAs (basis tile + two coefficient stages each) fit an SM, else 16-row ones.
template <int MODE>
int launch_narrow(const float *Bt, const float *G, int ndir, int ldg, int64_t npix, int64_t ncol, float *slab,
               ssf::OutFmt fmt, cudaMemPool_t pool, cudaStream_t stream) {
  const int Kp = (ndir + 7) / 8 * 8;
  auto smem_for = [&](int nt, int stages) {
    return sizeof(float) * (size_t)(TC_M + 2 * stages * nt) * Kp + 8 * NW_NBARS + 64;
  };
  auto fits2 = [&](int nt, int stages) { return 2 * (smem_for(nt, stages) + 1024) <= 227 * 1024; };
  int nt = fits2(32, 2) ? 32 : 16;
  if (const char *e = getenv("SSF_TC_NT")) {        // experiment hook: tile rows
    const int want = atoi(e);
    if ((want == 16 || want == 32) && fits2(want, 2)) nt = want;
  }
  int nb = 2;
  while (nb < NW_MAXB && fits2(nt, nb + 1)) ++nb;
  if (const char *e = getenv("SSF_TC_NB")) {
    const int want = atoi(e);
    if (want >= 2 && want <= NW_MAXB && fits2(nt, want)) nb = want;
  }
  const size_t smem = smem_for(nt, nb);
  if (nt == 32) return launch_narrow_nt<MODE, 32>(Bt, G, ndir, ldg, npix, ncol, slab, fmt, pool, stream, nb, smem);
  return launch_narrow_nt<MODE, 16>(Bt, G, ndir, ldg, npix, ncol, slab, fmt, pool, stream, nb, smem);
}

// Stream-ordered scratch for the split coefficient tiles, from a library-private memory pool
// (kept warm: release threshold = max), one per device.
cudaMemPool_t scratch_pool(int device) {
  static std::mutex mu;
  static cudaMemPool_t pools[64] = {};
  std::lock_guard<std::mutex> lock(mu);
  if (device < 0 || device >= 64) return nullptr;
  if (!pools[device]) {
    cudaMemPoolProps props = {};
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = device;
    cudaMemPool_t pool = nullptr;
    if (cudaMemPoolCreate(&pool, &props) != cudaSuccess) return nullptr;
    uint64_t keep = UINT64_MAX;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    pools[device] = pool;
  }
  return pools[device];
}

template <int MODE, int NEPI>
int launch_kernel(const float *Bt, const float *Bs, int ndir, int64_t npix, int64_t ncol, float *slab, int Kp,
                  size_t smem, int nb, ssf::OutFmt fmt, cudaStream_t stream) {
  constexpr int CT = tile_columns(MODE);
  auto kfn = contract_tc_kernel<MODE, NEPI>;
  SSF_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int pix_tiles = (int)(npix / TC_M);
  const int64_t n_ct = (ncol + CT - 1) / CT;
  int per = pick_tiles_per_cta(CT, Kp, n_ct);
  if (const char *e = getenv("SSF_TC_PER")) {      // experiment hook: column tiles per CTA
    per = atoi(e);
    if (per < 1) per = 1;
    if (per > n_ct) per = (int)n_ct;
  }
  const int gx = (int)((n_ct + per - 1) / per);
  dim3 grid(gx, pix_tiles);
  SSF_LAUNCH(kfn, grid, (NEPI + 2) * 32, smem, stream, Bt, Bs, ndir, npix, ncol, slab, per, Kp, nb, fmt);
  SSF_LAUNCHED();
  return SSF_OK;
}

template <int MODE>
int launch_tc(const float *Bt, const float *G, int ndir, int ldg, int64_t npix, int64_t ncol, float *slab,
              ssf::OutFmt fmt, cudaStream_t stream) {
  constexpr int CT = tile_columns(MODE);
  constexpr int NR = tile_rows(MODE);
  const int Kp = (ndir + 7) / 8 * 8;
  const int64_t n_ct = (ncol + CT - 1) / CT;
  if (n_ct == 0) return SSF_OK;
  // + the mbarriers (N_BARS x 8 B), the TMEM base address and slack
  auto smem_for = [&](int stages) { return sizeof(float) * (size_t)(2 * TC_M + 2 * stages * NR) * Kp + 8 * N_BARS + 64; };
  const int per_sm = 2 * (smem_for(2) + 1024) <= 227 * 1024 ? 2 : 1;
  // Ring depth of the coefficient tiles: as deep as two resident CTAs allow.  A stage is only
  // re-armed when the MMAs that read it have retired, and its refill then has to come from L2
  // through a memory system saturated with stores (microseconds); the small gain tiles (32 KB of
  // cube per tile) need 3 stages to ride that out, the 128 KB KL tiles are fine with 2.
  int nb = 2;
  if (per_sm == 2)
    while (nb < B_MAX && 2 * (smem_for(nb + 1) + 1024) <= 227 * 1024) ++nb;
  if (const char *e = getenv("SSF_TC_NB")) {         // experiment hook: ring depth
    const int want = atoi(e);
    if (want >= 2 && want <= B_MAX && smem_for(want) + 1024 <= 227 * 1024) nb = want;
  }
  size_t smem = smem_for(nb);
  bool force_one = false;
  if (const char *e = getenv("SSF_TC_ONE_CTA")) {    // experiment hook: pad shared memory so that one CTA fits an SM
    if (e[0] == '1' && smem < 120 * 1024) {
      smem = 120 * 1024;
      force_one = true;
    }
  }

  int device = 0;
  SSF_CUDA(cudaGetDevice(&device));
  cudaMemPool_t pool = scratch_pool(device);
  SSF_REQUIRE(pool != nullptr, "tcgen05 contraction: could not create the scratch memory pool");
  {
    // more than 56 directions: persistent kernel with basis-tile prefetch
    // (experiment hooks: SSF_TC_BIGK=0 never, SSF_TC_BIGK=2 always)
    const char *e = getenv("SSF_TC_BIGK");
    const bool never = e && e[0] == '0', always = e && e[0] == '2';
    // KL screens above 56 directions: the narrow-tile kernel keeps two CTAs per SM
    // (SSF_TC_NARROW=0: the persistent kernel instead; =2: always, any mode but gain, for experiments).
    // Measured (profiles/r02_kernels_w_narrow.txt): KL 0.90 at 60 directions / 1024^2 and 0.80 at 64 / 2048^2
    // against 0.76-0.77 for the persistent kernel, 0.72-0.76 against 0.69-0.72 at 100 / 2048^2; the
    // Voronoi tiles of this layout (16 columns) stay behind the persistent kernel (0.66 against 0.70).
    const char *en = getenv("SSF_TC_NARROW");
    const bool narrow_never = en && en[0] == '0', narrow_always = en && en[0] == '2';
    if (MODE != 2 && !always && ((MODE == 1 && per_sm == 1 && !never && !narrow_never) || narrow_always))
      return launch_narrow<(MODE == 2 ? 1 : MODE)>(Bt, G, ndir, ldg, npix, ncol, slab, fmt, pool, stream);
    if ((per_sm == 1 && !never) || always) return launch_bigk<MODE>(Bt, G, ndir, ldg, npix, ncol, slab, fmt, pool, stream);
  }
  float *Bs = nullptr;
  const size_t bs_bytes = sizeof(float) * (size_t)n_ct * 2 * NR * Kp;
  SSF_CUDA(cudaMallocFromPoolAsync(reinterpret_cast<void **>(&Bs), bs_bytes, pool, stream));
  ScratchGuard guard{Bs, stream};
  const int64_t n_thr = n_ct * NR * (Kp >> 2);
  SSF_LAUNCH(split_b_kernel<MODE>, ssf::div_up(n_thr, 256), 256, 0, stream, G, ndir, ldg, ncol, Kp, n_ct, Bs);
  SSF_LAUNCHED();
  const int rc = (2 * (smem + 1024) <= 227 * 1024 || force_one)
                     ? launch_kernel<MODE, 8>(Bt, Bs, ndir, npix, ncol, slab, Kp, smem, nb, fmt, stream)
                     : launch_kernel<MODE, 16>(Bt, Bs, ndir, npix, ncol, slab, Kp, smem, nb, fmt, stream);
  if (rc != SSF_OK) return rc;
  return guard.release();
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
                float *slab, OutFmt fmt, cudaStream_t stream) {
  if (kl) {
    SSF_REQUIRE(ldg % 4 == 0 && (uintptr_t)G % 16 == 0, "tcgen05 KL contraction: G must be 16 B aligned, ldg %% 4 == 0");
    return launch_tc<1>(Bt, G, ndir, ldg, npix, ncol, slab, fmt, stream);
  }
  SSF_REQUIRE((uintptr_t)G % 8 == 0, "tcgen05 Voronoi contraction: lut must be 8 B aligned");
  return launch_tc<0>(Bt, G, ndir, ldg, npix, ncol, slab, fmt, stream);
}
int contract_tc_gain(const float *Bt, const float *G3, int ndir, int ldg, int64_t npix, int64_t ncol, float *slab,
                     OutFmt fmt, cudaStream_t stream) {
  return launch_tc<2>(Bt, G3, ndir, ldg, npix, ncol, slab, fmt, stream);
}
}  // namespace ssf
#endif  // SSF_CPU_EMU
