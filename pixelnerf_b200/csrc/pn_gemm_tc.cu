// Tensor-core GEMMs of the training path (dgrad + wgrad of the 512-wide linears), tcgen05 +
// TMEM + TMA, one CTA pair (cta_group::2, M = 256) per tile.
//
//   rows GEMM   D[rows, N]  = A[rows, K] * B[K, N]            (dgrad: A = dL/d(out) bf16, B = W[out, in] as stored)
//   wgrad GEMM  D[M, N]    += A[rows, M]^T * B[rows, N]       (A = dL/d(out) bf16, B = layer input fp16, split over rows)
//
// Both read their "transposed" operand straight from its row-major tensor: the shared-memory tile is MN-major
// (TMA box = 64 contiguous elements of the M/N dimension x 64 K rows, 128-byte swizzle), which the UMMA
// descriptor supports for 16-bit types (major bits 15/16 of the instruction descriptor), so no transposed copy
// of any activation, gradient or weight is ever written.  fp32 accumulation in TMEM; A and B are both fp16 or
// both bf16 (one of each is an illegal instruction on sm_100a, measured).
//
// Warp roles per CTA (320 threads): warp 0 = TMA producer, warp 1 = MMA issuer (leader CTA only), warps 2-9 =
// epilogue (lane quarter = warp & 3, column half = (warp - 2) >> 2).
#include <cuda_bf16.h>

#include "pn_common.cuh"
#include "pn_tc_ptx.cuh"
#include "pn_tc_host.cuh"

namespace pn {
namespace tc {

constexpr int kGThreads = 320;
constexpr int kGBox = 64 * 64 * 2;                 // one [64 x 64] 16-bit box
constexpr int kGStageRows = 32 * 1024;                       // rows GEMM: A [128 x 64] + B 2 boxes; 4-6 stages (see kernel)
constexpr int kGStageW = 48 * 1024, kGStagesW = 4;           // wgrad: A 2 boxes + B 2 halves x 2 boxes
constexpr int kGRing = 192 * 1024;
constexpr int kGOffBar = kGRing;
enum { GB_FULL = 0, GB_EMPTY = 8, GB_ACCF = 16, GB_ACCE = 18, GB_IN = 20, GB_COUNT = 36 };   // GB_IN: 2 per epilogue warp
constexpr int kGOffTmem = kGOffBar + GB_COUNT * 8;
constexpr int kGOffColsum = kGOffTmem + 16;            // 512 floats: per-CTA column sums (rows GEMM)
constexpr int kGOffOut = (kGOffColsum + 512 * 4 + 511) / 512 * 512;   // rows GEMM: per epilogue warp, one [32 x 32] bf16
constexpr int kGOutBytes = 32 * 64;                                   //   box (64-byte rows, 64-byte swizzle) = TMA-store source
constexpr int kGSmem = kGOffOut + 8 * kGOutBytes + 1024;
static_assert(kGSmem <= 227 * 1024, "shared memory budget");

// MN-major operand tile, 128-byte swizzle: 64 contiguous M/N elements per 128-byte row, 8 K rows per 1024-byte
// group (SBO), 64-element M/N blocks `lbo` bytes apart.
__device__ __forceinline__ uint64_t smem_desc_mn128(uint32_t smem_addr, uint32_t lbo) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// kind::f16 instruction descriptor with operand formats (0 = fp16, 1 = bf16) and majors (1 = MN-major).
__host__ __device__ constexpr uint32_t idesc_16(int m, int n, int a_fmt, int b_fmt, int a_mn, int b_mn) {
  return (1u << 4) | ((uint32_t)a_fmt << 7) | ((uint32_t)b_fmt << 10) | ((uint32_t)a_mn << 15) |
         ((uint32_t)b_mn << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}
__device__ __forceinline__ uint32_t pack_bf2(float a, float b) {   // a in the low half
  uint32_t d;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(b), "f"(a));
  return d;
}

struct RowsParams {
  int64_t n_rows;
  int n_tiles;           // 256-row tiles
  int k_chunks;          // K / 64
  int n_halves;          // N / 256
  int ldn;               // N: row pitch (elements) of mask / acc32 / out16
  uint32_t idesc;
  const __half* mask;    // optional [rows, N] fp16 or bf16 relu OUTPUT: D is zeroed where it is +-0 (relu backward)
  float* acc32;          // optional [rows, N] fp32: written with D (+ its previous content when `accumulate`)
  int accumulate;
  const __nv_bfloat16* resid16;  // optional [rows, N] bf16 added to D (may alias out16: same thread reads then writes)
  __nv_bfloat16* out16;  // optional [rows, N] bf16 copy of the result
  float* colsum0;        // optional [N] fp32: += column sums of the result (bias gradients), up to two destinations
  float* colsum1;
  float scale;           // D is multiplied by this first
  int a_slab_rows;       // 0: A is [rows, K].  > 0: A is K/512 slabs of [a_slab_rows, 512] stacked along the rows
                         // (K chunk kc lives in slab kc / 8): sum over several 512-wide operands in one launch
  unsigned int* err;
};

struct WgradParams {
  int64_t n_rows;
  int k_split;           // row ranges; grid = 2 (M halves) x k_split pairs
  int n_halves;          // accumulators of kNB x 128 columns
  int n_store;           // columns actually stored (<= n_halves * kNB * 128)
  int ldo;               // pitch of out (floats)
  uint32_t idesc;
  float* out;            // [512, ldo] fp32, accumulated with red.global.add
  unsigned int* err;
};

// Common prologue: barriers, TMEM.  Returns the TMEM base.
__device__ __forceinline__ uint32_t gemm_setup(uint32_t smem_base, int warp, int lane, int stages) {
  auto bar = [&](int i) { return smem_base + kGOffBar + i * 8; };
  cluster_sync_all();
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(bar(GB_FULL + s), 1);    // the leader's expect_tx arrive (both CTAs' bytes)
      mbar_init(bar(GB_EMPTY + s), 1);   // one multicast commit
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(bar(GB_ACCF + a), 1);
      mbar_init(bar(GB_ACCE + a), 16);   // 8 epilogue warps of each CTA of the pair
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_pair(smem_base + kGOffTmem, 512);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(smem_base + kGOffTmem));
  return tmem_base;
}

__device__ __forceinline__ void gemm_teardown(uint32_t tmem_base, int warp) {
  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

// ---- rows GEMM ----------------------------------------------------------------------------------------
// Work item = (256-row tile, 256-column half); consecutive items alternate between the two TMEM accumulators so
// the epilogue of one item overlaps the MMAs of the next.  Stage = A chunk [128 rows x 64 K] (K-major) + this
// CTA's share of the B chunk [64 K x 128 N] (MN-major, two boxes).
template <bool kAcc>   // kAcc: the fp32 accumulate / output path (acc32) is compiled in
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGThreads, 1)
gemm_rows_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_o, const __grid_constant__ CUtensorMap map_m,
                 const __grid_constant__ CUtensorMap map_r, const RowsParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int n_pairs = gridDim.x >> 1, pair_id = blockIdx.x >> 1;
  auto bar = [&](int i) { return smem_base + kGOffBar + i * 8; };
  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&map_a);
    prefetch_tensormap(&map_b);
    if (p.out16) prefetch_tensormap(&map_o);
    if (!kAcc && p.mask) prefetch_tensormap(&map_m);
    if (!kAcc && p.resid16) prefetch_tensormap(&map_r);
  }
  if (warp == 1 && lane == 0)   // published by the fence + cluster barrier inside gemm_setup
    for (int i = 0; i < 16; ++i) mbar_init(bar(GB_IN + i), 1);
  float* cs_smem = reinterpret_cast<float*>(smem_raw + (smem_base - smem_u32(smem_raw)) + kGOffColsum);
  for (int i = threadIdx.x; i < 512; i += kGThreads) cs_smem[i] = 0.0f;
  // Operand ring depth: 6 stages when the epilogue takes its inputs with plain loads (fp32 accumulate path); the
  // dgrad forms give 64 KB (mask + residual boxes) or 32 KB (mask only) of the ring to the epilogue's TMA inputs.
  const int n_stages = kAcc ? 6 : (p.resid16 ? 4 : 5);
  const uint32_t tmem_base = gemm_setup(smem_base, warp, lane, n_stages);   // cluster barrier inside

  if (warp == 0) {
    uint32_t s = 0, ph = 0;
#pragma unroll 1
    for (int tile = pair_id; tile < p.n_tiles; tile += n_pairs) {
#pragma unroll 1
      for (int half = 0; half < p.n_halves; ++half) {
        const int n0 = half * 256 + (int)rank * 128;
#pragma unroll 1
        for (int kc = 0; kc < p.k_chunks; ++kc) {
          wait_bar<false>(bar(GB_EMPTY + s), ph ^ 1, p.err, 0x300 + s);
          if (elect_one()) {
            const uint32_t full = bar(GB_FULL + s);
            const uint32_t dst = smem_base + s * kGStageRows;
            if (rank == 0) mbar_arrive_expect_tx(full, 2 * kGStageRows);
            const int a_col = p.a_slab_rows ? (kc & 7) * 64 : kc * 64;
            const int a_row = tile * 256 + (int)rank * 128 + (p.a_slab_rows ? (kc >> 3) * p.a_slab_rows : 0);
            tma_load_2d_pair(dst, &map_a, full, a_col, a_row);
            tma_load_2d_pair(dst + 2 * kGBox, &map_b, full, n0, kc * 64);
            tma_load_2d_pair(dst + 3 * kGBox, &map_b, full, n0 + 64, kc * 64);
          }
          __syncwarp();
          if (++s == (uint32_t)n_stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1 && rank == 0) {
    uint32_t s = 0, ph = 0, it = 0;
    const uint64_t a_hi = smem_desc_k128(0), b_hi = smem_desc_mn128(0, kGBox);
#pragma unroll 1
    for (int tile = pair_id; tile < p.n_tiles; tile += n_pairs) {
#pragma unroll 1
      for (int half = 0; half < p.n_halves; ++half, ++it) {
        const uint32_t acc = it & 1, accph = (it >> 1) & 1;
        wait_bar<true>(bar(GB_ACCE + acc), accph ^ 1, p.err, 0x310 + acc);
        tc_fence_after();
#pragma unroll 1
        for (int kc = 0; kc < p.k_chunks; ++kc) {
          wait_bar<false>(bar(GB_FULL + s), ph, p.err, 0x320 + s);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t a_lo = (smem_base + s * kGStageRows) >> 4;
            const uint32_t b_lo = (smem_base + s * kGStageRows + 2 * kGBox) >> 4;
#pragma unroll
            for (int j = 0; j < 4; ++j)
              umma_f16_pair(tmem_base + acc * 256, a_hi | (uint64_t)(a_lo + 2 * j), b_hi | (uint64_t)(b_lo + 128 * j),
                            p.idesc, (kc | j) ? 1u : 0u);
            umma_commit_pair(bar(GB_EMPTY + s), 0x3);
          }
          __syncwarp();
          if (++s == (uint32_t)n_stages) { s = 0; ph ^= 1; }
        }
        if (elect_one()) umma_commit_pair(bar(GB_ACCF + acc), 0x3);
        __syncwarp();
      }
    }
  } else if (warp >= 2) {
    const int q = warp & 3, ch = (warp - 2) >> 2;
    const uint32_t ostage = smem_base + kGOffOut + (uint32_t)(warp - 2) * kGOutBytes;
    // !kAcc (the dgrad forms): the [32 rows x 32 columns] mask / residual boxes of a chunk arrive by TMA in one of
    // the warp's two buffers (64-byte rows, 64-byte swizzle, their own mbarrier), requested two chunks ahead --
    // across item boundaries too -- so neither the lookahead nor the 32-lines-per-instruction global loads of the
    // earlier version cost registers or L1 wavefronts.  (in_t, in_h, in_c) is the next chunk to request; a
    // chunk's buffer is c & 1.  (Four mask-only buffers, four chunks ahead, measured no faster.)
    const uint32_t in_stride = p.resid16 ? 4096u : 2048u;
    const uint32_t in_base = smem_base + (uint32_t)n_stages * kGStageRows + (uint32_t)(warp - 2) * 2 * in_stride;
    const uint32_t in_bar0 = bar(GB_IN + 2 * (warp - 2));
    const uint32_t in_mask = 1u;
    uint32_t in_ph = 0;
    int in_t = pair_id, in_h = 0, in_c = 0;
    auto in_issue = [&]() {   // warp uniform; buffer = parity of the chunk (4 chunks per item)
      if (in_t < p.n_tiles) {
        if (lane == 0) {
          const uint32_t buf = (uint32_t)in_c & in_mask, ib = in_bar0 + 8 * buf;
          const int col = in_h * 256 + ch * 128 + in_c * 32;
          const int row0 = in_t * 256 + (int)rank * 128 + q * 32;
          mbar_arrive_expect_tx(ib, (p.mask ? 2048u : 0u) + (p.resid16 ? 2048u : 0u));
          if (p.mask) tma_load_2d_cta(in_base + buf * in_stride, &map_m, ib, col, row0);
          if (p.resid16) tma_load_2d_cta(in_base + buf * in_stride + 2048, &map_r, ib, col, row0);
        }
        if (++in_c == 4) {
          in_c = 0;
          if (++in_h == p.n_halves) { in_h = 0; in_t += n_pairs; }
        }
      }
    };
    auto in_take = [&](int c, uint32_t (&mk)[16], uint32_t (&rs)[16]) {
      const uint32_t buf = (uint32_t)c & in_mask;
      wait_bar<false>(in_bar0 + 8 * buf, (in_ph >> buf) & 1u, p.err, 0x340 + buf);
      in_ph ^= 1u << buf;
      const uint32_t src = in_base + buf * in_stride + lane * 64;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t o = (uint32_t)(j ^ ((lane >> 1) & 3)) << 4;
        if (p.mask)
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(mk[4 * j]), "=r"(mk[4 * j + 1]), "=r"(mk[4 * j + 2]), "=r"(mk[4 * j + 3]) : "r"(src + o) : "memory");
        if (p.resid16)
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(rs[4 * j]), "=r"(rs[4 * j + 1]), "=r"(rs[4 * j + 2]), "=r"(rs[4 * j + 3]) : "r"(src + 2048 + o) : "memory");
      }
      // The refill is an async-proxy write: order this lane's generic-proxy reads in front of it.  (Without this
      // fence a 25-run determinism test saw, about once in ten launches, a few 16-byte pieces of a mask box
      // replaced by the next chunk's -- tests/test_gpu_guards.py.)
      fence_proxy_async_smem();
      __syncwarp();   // every lane holds its copy: the buffer may be refilled
    };
    if (!kAcc) {
      in_issue();
      in_issue();
    }
    uint32_t it = 0;
#pragma unroll 1
    for (int tile = pair_id; tile < p.n_tiles; tile += n_pairs) {
      const int64_t row = (int64_t)tile * 256 + rank * 128 + q * 32 + lane;
      const bool live = row < p.n_rows;
#pragma unroll 1
      for (int half = 0; half < p.n_halves; ++half, ++it) {
        const uint32_t acc = it & 1, accph = (it >> 1) & 1;
        {
          // The epilogue's reads (mask, residual, fp32 accumulator) are per-thread row segments needed right
          // after the accumulator: pull the NEXT item's segments into L2 now, one item of lead time, so they cost
          // an L2 hit instead of a DRAM round trip when their turn comes (no registers, no shared memory).
          const int half2 = half + 1 < p.n_halves ? half + 1 : 0;
          const int tile2 = half + 1 < p.n_halves ? tile : tile + n_pairs;
          const int64_t row2 = (int64_t)tile2 * 256 + rank * 128 + q * 32 + lane;
          if (kAcc && tile2 < p.n_tiles && row2 < p.n_rows) {
            const size_t off2 = (size_t)row2 * p.ldn + half2 * 256 + ch * 128;
            if (p.mask) {
              prefetch_l2(p.mask + off2);
              prefetch_l2(p.mask + off2 + 64);
            }
            if (p.resid16) {
              prefetch_l2(p.resid16 + off2);
              prefetch_l2(p.resid16 + off2 + 64);
            }
            if (kAcc && p.accumulate) {
#pragma unroll
              for (int i = 0; i < 4; ++i) prefetch_l2(p.acc32 + off2 + 32 * i);
            }
          }
        }
        // Four 32-column chunks per warp.  The mask / residual segments of chunk c+1 are requested before chunk c
        // is processed (two register buffers, loop unrolled by hand), and those of chunk 0 before the accumulator
        // wait: a warp is otherwise one long dependent chain load -> tcgen05.ld -> math -> store per chunk, and
        // with two warps per scheduler nothing else hides it (ncu r01v: 7-9 us per item against 2.3 us of MMAs).
        auto issue = [&](int c, uint32_t (&mk)[16], uint32_t (&rs)[16]) {
          const size_t off = (size_t)row * p.ldn + half * 256 + ch * 128 + c * 32;
          if (live && p.resid16) {
            ldg_v8(p.resid16 + off, reinterpret_cast<uint32_t(&)[8]>(rs[0]));
            ldg_v8(p.resid16 + off + 16, reinterpret_cast<uint32_t(&)[8]>(rs[8]));
          }
          if (live && p.mask) {
            ldg_v8(p.mask + off, reinterpret_cast<uint32_t(&)[8]>(mk[0]));
            ldg_v8(p.mask + off + 16, reinterpret_cast<uint32_t(&)[8]>(mk[8]));
          }
        };
        auto process = [&](int c, const uint32_t (&mk)[16], const uint32_t (&rs)[16]) {
          const int lcol = ch * 128 + c * 32;
          const size_t off = (size_t)row * p.ldn + half * 256 + lcol;
          uint32_t prev[kAcc ? 32 : 1];
          if (kAcc && live && p.accumulate) {
#pragma unroll
            for (int i = 0; i < 4; ++i) ldg_v8(p.acc32 + off + 8 * i, reinterpret_cast<uint32_t(&)[8]>(prev[kAcc ? 8 * i : 0]));
          }
          uint32_t v[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * 256 + lcol, v);
          tmem_wait_ld();
          float f[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = 0.0f;
          if (live) {
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]) * p.scale;
            if (p.mask) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {   // relu outputs are >= 0: "not positive" = all value bits clear
                if ((mk[i] & 0x7FFFu) == 0) f[2 * i] = 0.0f;
                if ((mk[i] & 0x7FFF0000u) == 0) f[2 * i + 1] = 0.0f;
              }
            }
            if (kAcc && p.accumulate) {
#pragma unroll
              for (int i = 0; i < 32; ++i) f[i] += __uint_as_float(prev[kAcc ? i : 0]);
            }
            if (p.resid16) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                f[2 * i] += __uint_as_float(rs[i] << 16);
                f[2 * i + 1] += __uint_as_float(rs[i] & 0xFFFF0000u);
              }
            }
            if (kAcc && p.acc32) {
#pragma unroll
              for (int i = 0; i < 4; ++i)
                stg_v8(p.acc32 + off + 8 * i, __float_as_uint(f[8 * i]), __float_as_uint(f[8 * i + 1]),
                       __float_as_uint(f[8 * i + 2]), __float_as_uint(f[8 * i + 3]), __float_as_uint(f[8 * i + 4]),
                       __float_as_uint(f[8 * i + 5]), __float_as_uint(f[8 * i + 6]), __float_as_uint(f[8 * i + 7]));
            }
          }
          if (p.out16) {
            // The bf16 result leaves as one [32 rows x 32 columns] box per warp and chunk: staged in shared memory
            // (64-byte rows, 64-byte swizzle) and written by one TMA store, asynchronously -- a lane storing its own
            // row's 64 bytes touches 32 lines per instruction, and L1 wavefronts are what bounds this epilogue
            // (ncu r01v: l1tex 62-69 %).  Rows past the end are clipped by the tensor map; in-place use (out16 ==
            // resid16) is safe: the box is exactly what this warp has already read.
            if (lane == 0) bulk_wait_read0();   // the previous box has left the staging buffer
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j)
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};"
                           ::"r"(ostage + lane * 64 + ((uint32_t)(j ^ ((lane >> 1) & 3)) << 4)),
                             "r"(pack_bf2(f[8 * j], f[8 * j + 1])), "r"(pack_bf2(f[8 * j + 2], f[8 * j + 3])),
                             "r"(pack_bf2(f[8 * j + 4], f[8 * j + 5])), "r"(pack_bf2(f[8 * j + 6], f[8 * j + 7]))
                           : "memory");
            fence_proxy_async_smem();
            __syncwarp();
            const int64_t row0 = row - lane;
            if (lane == 0 && row0 < p.n_rows) {
              tma_store_2d(&map_o, ostage, half * 256 + lcol, (int)row0);
              bulk_commit_group();
            }
          }
          if (p.colsum0) {
            // column sums over the warp's 32 rows by a transposing butterfly (31 shuffles): lane l ends up
            // with column lcol + l; dead rows hold zeros
#pragma unroll
            for (int sft = 16; sft >= 1; sft >>= 1) {
              const bool up = (lane & sft) != 0;
#pragma unroll
              for (int i = 0; i < sft; ++i) {
                const float send = up ? f[i] : f[i + sft];
                const float recv = __shfl_xor_sync(0xffffffffu, send, sft);
                f[i] = (up ? f[i + sft] : f[i]) + recv;
              }
            }
            atomicAdd(cs_smem + half * 256 + lcol + lane, f[0]);
          }
        };
        if constexpr (kAcc) {   // the fp32 path needs its registers for the accumulator segment: no lookahead
          uint32_t mk[16], rs[16];
          wait_bar<false>(bar(GB_ACCF + acc), accph, p.err, 0x330 + acc);
          tc_fence_after();
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            issue(c, mk, rs);
            process(c, mk, rs);
          }
        } else {
          wait_bar<false>(bar(GB_ACCF + acc), accph, p.err, 0x330 + acc);
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t mk[16], rs[16];
            in_take(c, mk, rs);
            in_issue();          // chunk c + 2 (this item's or the next one's) into the buffer just read
            process(c, mk, rs);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(bar(GB_ACCE + acc), 0);
      }
    }
  }
  __syncwarp();
  bulk_wait0();                     // this thread's TMA stores are complete
  gemm_teardown(tmem_base, warp);   // cluster barrier: every epilogue warp's shared-memory atomics are done
  if (p.colsum0) {
    __syncthreads();
    for (int i = threadIdx.x; i < p.n_halves * 256; i += kGThreads) {
      const float v = cs_smem[i];
      atomicAdd(p.colsum0 + i, v);
      if (p.colsum1) atomicAdd(p.colsum1 + i, v);
    }
  }
}

// ---- wgrad GEMM -----------------------------------------------------------------------------------------
// Pair (mr, ks): output rows [mr*256, +256) x all N columns, rows [ks-th range) of the K dimension; both operands
// MN-major.  kNB = 64-column boxes of B per CTA per accumulator (2: N = 256 per accumulator, 1: N = 128).
// Stage = A 2 boxes + B n_halves * kNB boxes, all [64 x 64].  The partial sums leave through red.global.add.
template <int kNB>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGThreads, 1)
gemm_wgrad_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                  const WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int pair_id = blockIdx.x >> 1;
  const int mr = pair_id & 1, ks = pair_id >> 1;
  auto bar = [&](int i) { return smem_base + kGOffBar + i * 8; };
  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&map_a);
    prefetch_tensormap(&map_b);
  }
  const uint32_t tmem_base = gemm_setup(smem_base, warp, lane, kGStagesW);
  const int total_chunks = (int)((p.n_rows + 63) >> 6);
  const int per = (total_chunks + p.k_split - 1) / p.k_split;
  const int c_begin = ks * per;
  const int c_end = (c_begin + per < total_chunks) ? c_begin + per : total_chunks;
  const int n_acc_cols = kNB * 128;
  const uint32_t stage_bytes = (uint32_t)(2 + p.n_halves * kNB) * kGBox;

  if (c_begin < c_end) {
    if (warp == 0) {
      uint32_t s = 0, ph = 0;
      const int m0 = mr * 256 + (int)rank * 128;
#pragma unroll 1
      for (int c = c_begin; c < c_end; ++c) {
        wait_bar<false>(bar(GB_EMPTY + s), ph ^ 1, p.err, 0x400 + s);
        if (elect_one()) {
          const uint32_t full = bar(GB_FULL + s);
          const uint32_t dst = smem_base + s * kGStageW;
          if (rank == 0) mbar_arrive_expect_tx(full, 2 * stage_bytes);
          tma_load_2d_pair(dst, &map_a, full, m0, c * 64);
          tma_load_2d_pair(dst + kGBox, &map_a, full, m0 + 64, c * 64);
          for (int h = 0; h < p.n_halves; ++h)
#pragma unroll
            for (int i = 0; i < kNB; ++i)
              tma_load_2d_pair(dst + (2 + h * kNB + i) * kGBox, &map_b, full,
                               h * n_acc_cols + (int)rank * (kNB * 64) + i * 64, c * 64);
        }
        __syncwarp();
        if (++s == kGStagesW) { s = 0; ph ^= 1; }
      }
    } else if (warp == 1 && rank == 0) {
      uint32_t s = 0, ph = 0;
      const uint64_t a_hi = smem_desc_mn128(0, kGBox), b_hi = smem_desc_mn128(0, kGBox);
#pragma unroll 1
      for (int c = c_begin; c < c_end; ++c) {
        wait_bar<false>(bar(GB_FULL + s), ph, p.err, 0x420 + s);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_lo = (smem_base + s * kGStageW) >> 4;
          for (int h = 0; h < p.n_halves; ++h) {
            const uint32_t b_lo = (smem_base + s * kGStageW + (2 + h * kNB) * kGBox) >> 4;
#pragma unroll
            for (int j = 0; j < 4; ++j)
              umma_f16_pair(tmem_base + h * 256, a_hi | (uint64_t)(a_lo + 128 * j), b_hi | (uint64_t)(b_lo + 128 * j),
                            p.idesc, (c != c_begin || j) ? 1u : 0u);
          }
          umma_commit_pair(bar(GB_EMPTY + s), 0x3);
        }
        __syncwarp();
        if (++s == kGStagesW) { s = 0; ph ^= 1; }
      }
      if (elect_one()) umma_commit_pair(bar(GB_ACCF), 0x3);
      __syncwarp();
    } else if (warp >= 2) {
      const int q = warp & 3, ch = (warp - 2) >> 2;
      wait_bar<false>(bar(GB_ACCF), 0, p.err, 0x430);
      tc_fence_after();
      const int m = mr * 256 + (int)rank * 128 + q * 32 + lane;
      const bool vec = (p.ldo & 3) == 0;
      for (int h = 0; h < p.n_halves; ++h) {
#pragma unroll 1
        for (int c = ch; c < n_acc_cols / 32; c += 2) {
          const int n = h * n_acc_cols + c * 32;
          if (n >= p.n_store) continue;     // warp-uniform
          uint32_t v[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + h * 256 + c * 32, v);
          tmem_wait_ld();
          float* o = p.out + (size_t)m * p.ldo + n;
          if (vec && n + 32 <= p.n_store) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              red_add_v4(o + 4 * i, __uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                         __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (n + i < p.n_store) atomicAdd(o + i, __uint_as_float(v[i]));
          }
        }
      }
      tc_fence_before();
    }
  }
  __syncwarp();
  gemm_teardown(tmem_base, warp);
}

// ---- host side ----------------------------------------------------------------------------------------------
static int gemm_grid(int* sms_out) {
  int dev = 0, sms = 0;
  PN_CUDA_OK(cudaGetDevice(&dev));
  PN_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  *sms_out = sms;
  return PN_OK;
}

template <typename K>
static int gemm_smem_attr(K kernel) {
  PN_CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGSmem));
  return PN_OK;
}

}  // namespace tc
}  // namespace pn

static int gemm_tc_rows_impl(const void* a, int a_fmt, const void* b, int b_fmt, int64_t n_rows, int k, int n,
                             const void* mask_f16, float* acc32, int accumulate, const void* resid_bf16,
                             void* out_bf16, float scale, float* colsum0, float* colsum1,
                             unsigned int* err_word, pn_stream_t stream, int64_t a_slab_rows);

extern "C" int pn_gemm_tc_rows(const void* a, int a_fmt, const void* b, int b_fmt, int64_t n_rows, int k, int n,
                               const void* mask_f16, float* acc32, int accumulate, const void* resid_bf16,
                               void* out_bf16, float scale, float* colsum0, float* colsum1,
                               unsigned int* err_word, pn_stream_t stream) {
  return gemm_tc_rows_impl(a, a_fmt, b, b_fmt, n_rows, k, n, mask_f16, acc32, accumulate, resid_bf16, out_bf16, scale,
                           colsum0, colsum1, err_word, stream, 0);
}

extern "C" int pn_gemm_tc_rows_slabs(const void* a_slabs, int64_t slab_rows, int fmt, const void* b, int64_t n_rows,
                                     int k, int n, float* acc32, int accumulate, float scale, unsigned int* err_word,
                                     pn_stream_t stream) {
  PN_CHECK_ARG(k > 0 && k % 512 == 0, "pn_gemm_tc_rows_slabs: K must be a multiple of 512 (one slab per 512)");
  PN_CHECK_ARG(slab_rows >= n_rows && slab_rows % 256 == 0 && slab_rows * (k / 512) < (1ll << 31),
               "pn_gemm_tc_rows_slabs: slab pitch must be a multiple of 256 rows, >= n_rows");
  PN_CHECK_ARG(acc32, "pn_gemm_tc_rows_slabs: fp32 output required");
  return gemm_tc_rows_impl(a_slabs, fmt, b, fmt, n_rows, k, n, nullptr, acc32, accumulate, nullptr, nullptr, scale,
                           nullptr, nullptr, err_word, stream, slab_rows);
}

static int gemm_tc_rows_impl(const void* a, int a_fmt, const void* b, int b_fmt, int64_t n_rows, int k, int n,
                             const void* mask_f16, float* acc32, int accumulate, const void* resid_bf16,
                             void* out_bf16, float scale, float* colsum0, float* colsum1,
                             unsigned int* err_word, pn_stream_t stream, int64_t a_slab_rows) {
  using namespace pn::tc;
  PN_CHECK_ARG(a && b, "pn_gemm_tc_rows: null pointer");
  PN_CHECK_ARG((a_fmt | 1) == 1 && a_fmt == b_fmt,
               "pn_gemm_tc_rows: operands must have the same format, 0 (fp16) or 1 (bf16)");
  PN_CHECK_ARG(n_rows >= 0 && n_rows < (1ll << 31) - 256, "pn_gemm_tc_rows: bad row count");
  PN_CHECK_ARG(k > 0 && k % 64 == 0 && n > 0 && n % 256 == 0, "pn_gemm_tc_rows: K must be a multiple of 64, N of 256");
  PN_CHECK_ARG(acc32 || out_bf16, "pn_gemm_tc_rows: no output");
  PN_CHECK_ARG(!accumulate || acc32, "pn_gemm_tc_rows: accumulate needs acc32");
  PN_CHECK_ARG(colsum0 || !colsum1, "pn_gemm_tc_rows: colsum1 without colsum0");
  PN_CHECK_ARG(n <= 512 || !colsum0, "pn_gemm_tc_rows: column sums need n <= 512");
  PN_CHECK_ARG((((uintptr_t)a | (uintptr_t)b) & 15) == 0 &&
                   (((uintptr_t)mask_f16 | (uintptr_t)acc32 | (uintptr_t)out_bf16 | (uintptr_t)resid_bf16) & 31) == 0,
               "pn_gemm_tc_rows: operands must be 16-byte, mask / outputs 32-byte aligned");
  if (n_rows == 0) return PN_OK;
  static pn::DeviceOnce attr_set;   // per device (a second GPU of the process needs its own opt-in)
  int rc, attr_dev;
  if (attr_set.needed(&attr_dev)) {
    if ((rc = gemm_smem_attr(gemm_rows_kernel<false>))) return rc;
    if ((rc = gemm_smem_attr(gemm_rows_kernel<true>))) return rc;
    attr_set.mark(attr_dev);
  }
  CUtensorMap ma, mb;
  if (a_slab_rows > 0) {
    if ((rc = make_map(&ma, a, (uint64_t)a_slab_rows * (k / 512), 512, 1024, 128))) return rc;
  } else if ((rc = make_map(&ma, a, (uint64_t)n_rows, (uint64_t)k, (uint64_t)k * 2, 128))) return rc;
  if ((rc = make_map(&mb, b, (uint64_t)k, (uint64_t)n, (uint64_t)n * 2, 64))) return rc;
  RowsParams p = {};
  p.n_rows = n_rows;
  p.n_tiles = (int)((n_rows + 255) / 256);
  p.k_chunks = k / 64;
  p.n_halves = n / 256;
  p.ldn = n;
  p.idesc = idesc_16(256, 256, a_fmt, b_fmt, 0, 1);
  p.mask = (const __half*)mask_f16;
  p.acc32 = acc32;
  p.accumulate = accumulate;
  p.out16 = (__nv_bfloat16*)out_bf16;
  p.resid16 = (const __nv_bfloat16*)resid_bf16;
  p.colsum0 = colsum0;
  p.colsum1 = colsum1;
  p.scale = scale;
  p.a_slab_rows = (int)a_slab_rows;
  p.err = err_word;
  int sms = 0;
  if ((rc = gemm_grid(&sms))) return rc;
  const int pairs = p.n_tiles < sms / 2 ? p.n_tiles : sms / 2;
  CUtensorMap mo = ma;   // placeholder when there is no bf16 output (never read then)
  if (out_bf16 && (rc = make_store_map(&mo, out_bf16, (uint64_t)n_rows, (uint64_t)n))) return rc;
  CUtensorMap mm = ma, mr = ma;   // placeholders (never read without the matching pointer)
  if (!acc32 && mask_f16 && (rc = make_store_map(&mm, mask_f16, (uint64_t)n_rows, (uint64_t)n))) return rc;
  if (!acc32 && resid_bf16 && (rc = make_store_map(&mr, resid_bf16, (uint64_t)n_rows, (uint64_t)n))) return rc;
  if (acc32) gemm_rows_kernel<true><<<2 * pairs, kGThreads, kGSmem, (cudaStream_t)stream>>>(ma, mb, mo, mm, mr, p);
  else gemm_rows_kernel<false><<<2 * pairs, kGThreads, kGSmem, (cudaStream_t)stream>>>(ma, mb, mo, mm, mr, p);
  PN_LAUNCH_OK("gemm_rows_kernel");
  pn::count_launch();
  return PN_OK;
}

extern "C" int pn_gemm_tc_wgrad(const void* g, int g_fmt, int ldg, const void* x, int x_fmt, int ldx, int64_t n_rows,
                                int n_store, float* out, int ldo, unsigned int* err_word, pn_stream_t stream) {
  using namespace pn::tc;
  PN_CHECK_ARG(g && x && out, "pn_gemm_tc_wgrad: null pointer");
  PN_CHECK_ARG((g_fmt | 1) == 1 && g_fmt == x_fmt,
               "pn_gemm_tc_wgrad: operands must have the same format, 0 (fp16) or 1 (bf16)");
  PN_CHECK_ARG(n_rows >= 0 && n_rows < (1ll << 31) - 256, "pn_gemm_tc_wgrad: bad row count");
  PN_CHECK_ARG(ldg == 512, "pn_gemm_tc_wgrad: the gradient operand must be 512 wide");
  PN_CHECK_ARG(ldx == 512 || ldx == 64, "pn_gemm_tc_wgrad: the input operand must be 512 or 64 wide");
  PN_CHECK_ARG(n_store > 0 && n_store <= ldx && ldo >= n_store, "pn_gemm_tc_wgrad: bad output shape");
  PN_CHECK_ARG((((uintptr_t)g | (uintptr_t)x) & 15) == 0 && ((uintptr_t)out & 3) == 0 &&
                   ((ldo & 3) != 0 || ((uintptr_t)out & 15) == 0),
               "pn_gemm_tc_wgrad: misaligned buffer");
  if (n_rows == 0) return PN_OK;
  static pn::DeviceOnce attr_set;
  int rc, attr_dev;
  if (attr_set.needed(&attr_dev)) {
    if ((rc = gemm_smem_attr(gemm_wgrad_kernel<2>))) return rc;
    if ((rc = gemm_smem_attr(gemm_wgrad_kernel<1>))) return rc;
    attr_set.mark(attr_dev);
  }
  CUtensorMap ma, mb;
  if ((rc = make_map(&ma, g, (uint64_t)n_rows, (uint64_t)ldg, (uint64_t)ldg * 2, 64))) return rc;
  if ((rc = make_map(&mb, x, (uint64_t)n_rows, (uint64_t)ldx, (uint64_t)ldx * 2, 64))) return rc;
  WgradParams p = {};
  p.n_rows = n_rows;
  int sms = 0;
  if ((rc = gemm_grid(&sms))) return rc;
  const int chunks = (int)((n_rows + 63) / 64);
  const int max_split = sms / 4;
  p.k_split = chunks < max_split ? chunks : max_split;
  p.n_store = n_store;
  p.ldo = ldo;
  p.out = out;
  p.err = err_word;
  if (ldx == 512) {
    p.n_halves = 2;
    p.idesc = idesc_16(256, 256, g_fmt, x_fmt, 1, 1);
    gemm_wgrad_kernel<2><<<4 * p.k_split, kGThreads, kGSmem, (cudaStream_t)stream>>>(ma, mb, p);
  } else {
    p.n_halves = 1;   // N = 128 over the pair: rank 1's box lies past column 64 and is zero filled by TMA
    p.idesc = idesc_16(256, 128, g_fmt, x_fmt, 1, 1);
    gemm_wgrad_kernel<1><<<4 * p.k_split, kGThreads, kGSmem, (cudaStream_t)stream>>>(ma, mb, p);
  }
  PN_LAUNCH_OK("gemm_wgrad_kernel");
  pn::count_launch();
  return PN_OK;
}
