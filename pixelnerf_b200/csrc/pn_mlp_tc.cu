// ResnetFC chain on the 5th-generation tensor cores (reference src/model/resnetfc.py:132-184,
// block :53-62, head src/model/models.py:257-266): one persistent kernel runs
//   lin_in, 3 x lin_z, 5 x (fc_0, fc_1), lin_out
// for a tile of rows with every activation resident on chip.  lin_z is not a GEMM here: it is linear
// and a point's latent row is a convex blend of four texels, so lin_z[b](blend) = blend of per-texel
// tables lin_z[b](texel) (pn_prepare_latent_tables, once per encode): the epilogue gathers the four
// table rows of each point and adds their blend to the residual stream (22 % fewer tensor FLOPs,
// no 1 KB latent row per point in HBM).
//
// Work decomposition.  A CTA pair (2-CTA cluster, tcgen05 cta_group::2) owns a tile of 128
// feature rows, 64 per CTA.  Every GEMM is issued as M=128 x N=256 x K=16 tcgen05.mma
// instructions by one thread of the leader CTA; each CTA supplies its 64 rows of A and its
// 128-row half of the 256-wide weight block B from its own shared memory.
//
// Tensor memory (512 columns per CTA, all used).  In the M=128 pair layout a 64 x 256 fp32
// accumulator takes 128 columns: lanes 0-63 hold output features [0,128) of the block, lanes
// 64-127 features [128,256).  Columns [0,256) hold the fp32 residual stream h (two 256-wide
// blocks), columns [256,512) hold t = fc_0(relu(h)).  lin_out (4 outputs) and the head are evaluated
// in fp32 by the epilogue warps straight from h (head_ffma).  h is never
// rounded: fc_1 accumulates into it in place (tcgen05.mma accumulate); before blocks 0-2 the
// epilogue adds the lin_z table blend (all biases up to there folded into the tables) with a
// tcgen05.ld / tcgen05.st read-modify-write while h is quiescent, later biases are added as a
// cumulative per-column vector whenever h is read.
//
// Shared memory per CTA (209 KB): ACT (relu'd fp16 operand of the next GEMM, 8 K-chunks of
// [64 rows x 64 fp16], 128-byte swizzled, 64 KB), two 8 KB slots for the tile's positional-code
// chunk x (TMA) and a 4-stage ring of weight
// tiles [128 features x 128 K] (32 KB each, two TMA boxes).  128 KB of weights in flight per
// CTA is what covers the L2 round trip; one stage = 8 tcgen05.mma = 512 tensor-pipe clocks.
//
// Overlap.  A GEMM is issued in four quarters (block0,K-lo) (block1,K-lo) (block0,K-hi)
// (block1,K-hi): block 0's accumulator is complete after the third quarter, so the epilogue
// warps convert it to the next operand's K-chunks 0-3 while the fourth quarter runs, and
// the next GEMM's first two quarters only need chunks 0-3.  The table gather of an h epilogue
// is issued before that epilogue waits for its accumulator, so it is off the critical path.
#include <cstdio>
#include <cstdlib>
#include <type_traits>

#include "pn_common.cuh"
#include "pn_tc_ptx.cuh"
#include "pn_tc_host.cuh"

namespace pn {
namespace tc {

constexpr int kThreads = 384;          // warp 0 TMA, warp 1 MMA issue, warp 2 TMEM alloc, warps 4-11 epilogue
constexpr int kEpiWarp0 = 4;
constexpr int kRows = 64;              // rows per CTA
constexpr int kTileRows = 128;         // rows per CTA pair
constexpr int kChunkBytes = kRows * 128;   // [64 rows x 64 fp16]
constexpr int kWBoxBytes = 128 * 128;      // TMA box [128 features x 64 fp16]
constexpr int kWStageBytes = 2 * kWBoxBytes;   // one ring stage = K 128 = two boxes
constexpr int kStages = 4;
constexpr int kASlots = 2;                 // x chunk of the current and the next tile
constexpr int kOffAct = 0;
constexpr int kOffA = 8 * kChunkBytes;
constexpr int kOffW = kOffA + kASlots * kChunkBytes;
constexpr int kOffBar = kOffW + kStages * kWStageBytes;
enum {
  BAR_W_FULL = 0,
  BAR_W_EMPTY = kStages,
  BAR_A_FULL = 2 * kStages,
  BAR_A_EMPTY = 2 * kStages + kASlots,
  BAR_ACT = 2 * kStages + 2 * kASlots,   // 8: operand chunk c of the current generation is in shared memory
  BAR_ACC = BAR_ACT + 8,                 // 4: accumulator (region, block) complete
  BAR_XLO_FULL = BAR_ACC + 4,            // precise chains: the x_lo chunk of the coming lin_in has landed (in ACT chunk 7)
  BAR_XLO_FREE = BAR_XLO_FULL + 1,       //   the last GEMM before that lin_in is complete: ACT may be overwritten
  BAR_COUNT = BAR_XLO_FREE + 1
};
constexpr int kOffXlo = kOffAct + 7 * kChunkBytes;
constexpr int kOffTmemPtr = kOffBar + BAR_COUNT * 8;
constexpr int kOffStage = (kOffTmemPtr + 16 + 511) / 512 * 512;   // per epilogue warp: [32 rows x 32 x 16 bit], 64-byte swizzle: table-blend staging and TMA-store source
constexpr int kStageBytes = 32 * 64;
constexpr int kSmemBytes = kOffStage + 8 * kStageBytes;
constexpr int kSmemAlloc = kSmemBytes + 1024;  // slack for the manual 1024-byte alignment
static_assert(kSmemAlloc <= 227 * 1024, "shared memory budget");

constexpr uint32_t kColH = 0, kColT = 256;  // TMEM column bases
constexpr int kParkCtas = 192;              // fused view mean: CTAs the park buffer (64 KB each) is sized for
constexpr size_t kParkBytesPerCta = 128 * 1024;   // 256 epilogue threads x 128 fp32
constexpr uint32_t kIdesc = idesc_f16(128, 256);

// Packed weights: 15 layers of [512 features][512 K] fp16 (lin_in and lin_out zero padded).
constexpr int kLayers = 15;
constexpr int L_IN = 0, L_Z0 = 1, L_FC = 4, L_OUT = 14;
constexpr size_t kPackedWeightBytes = (size_t)kLayers * 512 * 512 * 2;
// Bias table: rows 0-5 cumulative residual bias cb_0..cb_5, rows 6-10 fc_0 biases, row 11 b_out[4], rows 12-15 the
// four fp32 rows of lin_out.weight (the head runs on the CUDA cores in fp32, see head_ffma).
constexpr int kBiasFloats = 16 * 512;
constexpr size_t kBiasBytes = ((size_t)kBiasFloats * 4 + 255) / 256 * 256;

struct ChainParams {
  const float* bias;     // table of the segment being run
  const __half* tables;  // [3][n_texels][512] per-texel lin_z tables (pn_prepare_latent_tables)
  int64_t table_stride;  // n_texels * 512
  float4* out;           // (rows, 4) [sigmoid rgb, relu sigma], segments ending in lin_out
  float4* pre;           // optional raw lin_out outputs
  __half* h_io;          // segment [0,3): h written here as fp16; segment [3,5]: view-mean of it read
  uint16_t* save;        // training: 11 slots of save_rows x 512 bf16, the relu'd operand of every 512-wide GEMM
  int64_t save_rows;     //   slot 2b = relu(h_b) (fc_0 input), 2b+1 = relu(fc_0 out) (fc_1 input), 10 = lin_out input;
                         //   rows per slot, padded to whole tiles (stores of a ragged last tile land in the padding)
  float* debug_h;        // optional (rows, 512) dump of h at stage debug_stage
  unsigned int* err;     // device word set before a watchdog trap
  int64_t n_rows;        // rows this launch processes
  int64_t ppo;           // points per object (view-mean gather)
  int n_tiles;
  int first_block, last_block;   // [0,5] full chain, [0,3) and [3,5] when NS > 1
  int n_views;
  int debug_stage;
  // fused view mean (kFuse, n_views > 1, pn_mlp_forward_tc_grouped): x rows arrive in grouped order (pn_point_taps_grouped:
  // rows (S * n_views + v) * 128 + m = view v of point S * 128 + m); a "tile" of the persistent loop is a super-tile =
  // n_views sub-tiles of blocks 0-2 (one per view, same 128 points) followed by ONE merged tile of blocks 3-4 + head
  int group;             // != 0: fused mode
  int64_t n_points;      // SB * ppo
  const float* bias2;    // bias table of the merged phase (cumulative from block 3)
  float* park;           // 128 KB per CTA: the running fp32 sum of h over the views done so far (L2 resident)
  // precise chain (kX3, mlp_precision = fp16x3): every 512-wide GEMM runs as three operand passes a_hi w_hi +
  // a_hi w_lo + a_lo w_hi (fp16 pairs, fp32 accumulate); everything else stays in fp32
  const float* tables32; // [3][n_texels][512] fp32 lin_z tables
  // two-pass chain (mlp_precision = fp16x2, pn_mlp_forward_tc_x2): the precise chain without its third pass --
  // a_hi W_hi + a_hi W_lo: weights as fp16 pairs, activations single fp16, tables / lin_in / head in fp32
  int drop_a_lo;         // 0 (three passes) or 1
};

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// relu + round-to-nearest fp16 pair in one instruction (a in the low half)
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t pack_relu_bf2(float a, float b) {
  uint32_t d;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(b), "f"(a));
  return d;
}
// Training: 32 relu'd values of each of the warp's 32 rows -> one [32 x 32] bf16 box in global memory.  A lane
// storing its own row's 64 bytes straight to global touches 32 lines per instruction on the L1/shared-memory data
// path that bounds this kernel (the forward with such stores ran 48 % slower per row); instead the box is written to
// the warp's staging buffer (64-byte rows, 64-byte swizzle, the same 4 st.shared.v4 an ACT chunk costs) and leaves
// through one TMA store issued by lane 0, asynchronously.  `stage` must be free (bulk_wait_read0 + __syncwarp).
__device__ __forceinline__ void save_relu32(const CUtensorMap* map, uint32_t stage, int lane, int col, int64_t row0,
                                            bool store, const float (&x)[32]) {
#ifndef PN_EXPERIMENT_SAVE_NOWAIT   // timing experiment only (races on the staging buffer)
  if (lane == 0) bulk_wait_read0();   // the previous box has left the staging buffer
#endif
  __syncwarp();
#pragma unroll
  for (int j = 0; j < 4; ++j)
    st_shared_v4(stage + lane * 64 + ((uint32_t)(j ^ ((lane >> 1) & 3)) << 4),
                 pack_relu_bf2(x[8 * j + 0], x[8 * j + 1]), pack_relu_bf2(x[8 * j + 2], x[8 * j + 3]),
                 pack_relu_bf2(x[8 * j + 4], x[8 * j + 5]), pack_relu_bf2(x[8 * j + 6], x[8 * j + 7]));
  fence_proxy_async_smem();
  __syncwarp();
  if (lane == 0 && store) {
    tma_store_2d(map, stage, col, (int)row0);
    bulk_commit_group();
  }
}
__device__ __forceinline__ uint32_t pack_relu_h2(float a, float b) {
  uint32_t d;
  asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(b), "f"(a));
  return d;
}

template <bool kSave, bool kFuse = false, bool kX3 = false>
__global__ void __launch_bounds__(kThreads, 1)
mlp_chain_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x,
                 const __grid_constant__ CUtensorMap map_s, const ChainParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // A cluster is two CTA pairs where the GPU can place four CTAs together (preferred cluster dimension 4),
  // one pair elsewhere (148 SMs hold 33 clusters of four and 8 of two on B200).  Each pair runs its own row
  // tiles; the pairs of a 4-cluster walk the same weight schedule in lockstep and every weight tile is
  // fetched from L2 once per cluster (TMA multicast), which relieves the L2 -> SM path.
  const bool quad = cluster_nctarank() == 4;
  const uint32_t rank4 = cluster_ctarank();
  const uint32_t rank = rank4 & 1;            // rank inside the pair (0 = MMA leader)
  const uint32_t pair = rank4 >> 1;           // which pair of the cluster
  const uint16_t pair_mask = (uint16_t)(3u << (2 * pair));
  const int n_clusters = gridDim.x >> 1;      // historical names: "cluster" below = one CTA pair
  const int cluster_id = blockIdx.x >> 1;
  // every pair runs the same number of tiles (lockstep ring); tiles past the end are dummies (TMA zero fill,
  // stores masked by row < n_rows)
  const int tile_end = (p.n_tiles + n_clusters - 1) / n_clusters * n_clusters;
  auto bar = [&](int i) { return smem_base + kOffBar + i * 8; };
  const bool full_front = p.first_block == 0;      // lin_in / lin_z present (needs X, Z)
  const bool has_head = p.last_block == 5;         // lin_out present

  cluster_sync_all();
  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&map_w);
    prefetch_tensormap(&map_x);
    if (kSave || kX3) prefetch_tensormap(&map_s);   // kSave: store map of the saved operands; kX3: the x_lo rows
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar(BAR_W_FULL + s), 1);   // the leader's expect_tx arrive (both CTAs' bytes)
      mbar_init(bar(BAR_W_EMPTY + s), quad ? 2 : 1);  // one multicast commit from each pair of the cluster
    }
    for (int a = 0; a < kASlots; ++a) {
      mbar_init(bar(BAR_A_FULL + a), 1);
      mbar_init(bar(BAR_A_EMPTY + a), 1);
    }
    for (int c = 0; c < 8; ++c) mbar_init(bar(BAR_ACT + c), 8);  // 4 warps per chunk per CTA
    for (int a = 0; a < 4; ++a) mbar_init(bar(BAR_ACC + a), 1);
    mbar_init(bar(BAR_XLO_FULL), 1);
    mbar_init(bar(BAR_XLO_FREE), 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_pair(smem_base + kOffTmemPtr, 512);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(smem_base + kOffTmemPtr));

  if (warp == 0) {
    // =============================== TMA producer (both CTAs) ===============================
    // The whole warp walks the schedule (warp-uniform control flow keeps addresses and
    // parities in uniform registers); one elected lane issues the copies.  Loads are issued
    // in exactly the order the MMA issuer consumes them, so the two rings cannot deadlock.
    // Weight stages are assigned statically (lin_in: 0,1; item i of a 512x512 GEMM: i % 4; lin_out: 0-3) so
    // that every shared-memory address and barrier is a constant plus one uniform base register; the phase of
    // each stage is one bit of `wpar`.  The issuer walks the same schedule.
    uint32_t wpar = 0;          // bit s: phase parity of weight stage s (number of times it was filled, mod 2)
    uint32_t as = 0, aph = 0;   // operand-chunk ring slot / phase
    // Weight tile for K-chunks kc0 and kc1 of block nb of `layer`: two boxes of [128 out x 64 in].
    auto load_w = [&](auto stage_c, int layer, int nb, int kc0, int kc1) {
      // precise chain: a two-stage ring (stages 2-3 of the ring hold the lo half of the ACT operand); its two-pass
      // form has no lo half and keeps all four stages
      const uint32_t ws = (kX3 && !p.drop_a_lo) ? (decltype(stage_c)::value & 1u) : decltype(stage_c)::value;
      wait_bar<false>(bar(BAR_W_EMPTY + ws), ((wpar >> ws) & 1u) ^ 1u, p.err, 0x100 + ws);
      wpar ^= 1u << ws;
      if (elect_one()) {
        const uint32_t full = bar(BAR_W_FULL + ws);
        // the leader expects both CTAs' bytes; the peer's copies credit the leader's barrier
        if (rank == 0) mbar_arrive_expect_tx(full, 4 * kWBoxBytes);
        const int row = layer * 512 + nb * 256 + (int)rank * 128;
        // this CTA and its twin in the other pair (same rank) need the same two boxes: each fetches one and
        // multicasts it to both
        if (quad) {
          const uint16_t twins = (uint16_t)((1u << rank) | (1u << (rank + 2)));
          tma_load_2d_pair_mc(smem_base + kOffW + ws * kWStageBytes + pair * kWBoxBytes, &map_w, full,
                              (pair == 0 ? kc0 : kc1) * 64, row, twins);
        } else {
          tma_load_2d_pair(smem_base + kOffW + ws * kWStageBytes, &map_w, full, kc0 * 64, row);
          tma_load_2d_pair(smem_base + kOffW + ws * kWStageBytes + kWBoxBytes, &map_w, full, kc1 * 64, row);
        }
      }
      __syncwarp();
    };
    // One [64 rows x 64] operand chunk (x) of this CTA's rows.
    auto load_a = [&](const CUtensorMap* map, int col, int row0) {
      wait_bar<false>(bar(BAR_A_EMPTY + as), aph ^ 1, p.err, 0x110 + as);
      if (elect_one()) {
        const uint32_t full = bar(BAR_A_FULL + as);
        if (rank == 0) mbar_arrive_expect_tx(full, 2 * kChunkBytes);
        tma_load_2d_pair(smem_base + kOffA + as * kChunkBytes, map, full, col, row0);
      }
      __syncwarp();
      if (++as == kASlots) { as = 0; aph ^= 1; }
    };
    // Precise chains: the fp16 REMAINDER of the positional code, x_lo = fp16(x - fp16(x)) (pn_point_taps_f32 writes it
    // as a second [rows x 64] fp16 tensor), so that lin_in runs on the tensor cores as x_hi W_hi + x_hi W_lo + x_lo W_hi,
    // exact to 2^-22.  Shared memory has no room for a third operand slot; the chunk borrows ACT chunk 7, which is idle
    // between the last GEMM before a lin_in (its MMAs complete: BAR_XLO_FREE, committed by the issuer) and that
    // lin_in's epilogue (which first waits for the lin_in accumulators, i.e. for the MMAs that read the chunk).  The
    // load overlaps the head / merge epilogue of the previous tile.
    uint32_t xlo_ph = 0;
    auto load_xlo = [&](int row0) {
      wait_bar<false>(bar(BAR_XLO_FREE), xlo_ph ^ 1, p.err, 0x120);
      xlo_ph ^= 1;
      if (elect_one()) {
        const uint32_t full = bar(BAR_XLO_FULL);
        if (rank == 0) mbar_arrive_expect_tx(full, 2 * kChunkBytes);
        tma_load_2d_pair(smem_base + kOffXlo, &map_s, full, 0, row0);
      }
      __syncwarp();
    };
    using S0 = std::integral_constant<uint32_t, 0>;
    using S1 = std::integral_constant<uint32_t, 1>;
    using S2 = std::integral_constant<uint32_t, 2>;
    using S3 = std::integral_constant<uint32_t, 3>;
    auto load_fc = [&](int layer0) {   // quarter order, two K=128 items per quarter: K-chunks (h, h+2) of the group
      const int n_ps = kX3 ? 3 - p.drop_a_lo : 1;
#pragma unroll 1
      for (int ps = 0; ps < n_ps; ++ps) {   // precise chain: W_hi, W_lo, W_hi (against a_hi, a_hi, a_lo); two-pass form: W_hi, W_lo
        const int layer = ps == 1 ? layer0 + kLayers : layer0;
#pragma unroll 1
        for (int kg = 0; kg < 2; ++kg) {
          load_w(S0{}, layer, 0, kg * 4 + 0, kg * 4 + 2);
          load_w(S1{}, layer, 0, kg * 4 + 1, kg * 4 + 3);
          load_w(S2{}, layer, 1, kg * 4 + 0, kg * 4 + 2);
          load_w(S3{}, layer, 1, kg * 4 + 1, kg * 4 + 3);
        }
      }
    };
#pragma unroll 1
    for (int tile = cluster_id; tile < tile_end; tile += n_clusters) {
      const int row0 = tile * kTileRows + (int)rank * kRows;
      if (!full_front && elect_one()) {
        // merged segment: pull the NEXT tile's h rows (one contiguous 64 KB run per view) into L2 while this
        // tile computes; the epilogue warps' view-mean load then streams from L2 instead of HBM
        const int64_t nrow = (int64_t)(tile + n_clusters) * kTileRows + (int64_t)rank * kRows;
        if (tile + n_clusters < tile_end && nrow + kRows <= p.n_rows) {
          const int64_t obj = nrow / p.ppo, pt = nrow - obj * p.ppo;
          if (pt + kRows <= p.ppo)
            for (int v = 0; v < p.n_views; ++v)
              bulk_prefetch_l2(p.h_io + ((obj * p.n_views + v) * p.ppo + pt) * 512, kRows * 512 * 2);
        }
      }
      __syncwarp();
      if constexpr (kFuse) {
        // super-tile: n_views sub-tiles of (x, lin_in, blocks 0-2), then the merged tile's blocks 3-4
#pragma unroll 1
        for (int k = 0; k < p.n_views; ++k) {
          load_a(&map_x, 0, ((tile * p.n_views + k) * 2 + (int)rank) * kRows);
          if constexpr (kX3) load_xlo(((tile * p.n_views + k) * 2 + (int)rank) * kRows);
          load_w(S0{}, L_IN, 0, 0, 1);
          load_w(S1{}, L_IN, 1, 0, 1);
#pragma unroll 1
          for (int b = 0; b < 3; ++b) {
            load_fc(L_FC + 2 * b);
            load_fc(L_FC + 2 * b + 1);
          }
        }
#pragma unroll 1
        for (int b = 3; b < 5; ++b) {
          load_fc(L_FC + 2 * b);
          load_fc(L_FC + 2 * b + 1);
        }
        continue;
      }
      if (full_front) {
        load_a(&map_x, 0, row0);
        if constexpr (kX3) load_xlo(row0);
        load_w(S0{}, L_IN, 0, 0, 1);           // hi and lo halves of lin_in
        load_w(S1{}, L_IN, 1, 0, 1);
      }
#pragma unroll 1
      for (int b = p.first_block; b < p.last_block; ++b) {
        load_fc(L_FC + 2 * b);
        load_fc(L_FC + 2 * b + 1);
      }
      // lin_out needs no weight tiles: the head is evaluated in fp32 by the epilogue warps (head_ffma)
    }
  } else if (warp == 1 && rank == 0) {
    // =============================== MMA issuer (leader CTA) ================================
    // Warp-uniform control flow; one elected lane issues tcgen05.mma / tcgen05.commit.
    using S0 = std::integral_constant<uint32_t, 0>;
    using S1 = std::integral_constant<uint32_t, 1>;
    using S2 = std::integral_constant<uint32_t, 2>;
    using S3 = std::integral_constant<uint32_t, 3>;
    uint32_t wpar = 0, as = 0, aph = 0, gen = 0;   // wpar bit s: parity to wait for on weight stage s
    const uint32_t w0 = smem_base + kOffW;
    // descriptor = constant high word | (address >> 4): only the low word changes per MMA
    const uint64_t desc_hi = smem_desc_k128(0);
    // One weight stage: nk K=16 steps; A steps [0,4) come from a0, [4,8) from a1 (64-wide chunks).
    auto item = [&](auto stage_c, uint32_t a0, uint32_t a1, uint32_t d_col, bool clear, uint32_t kmask, uint32_t a2 = 0) {
      constexpr uint32_t idesc = kIdesc;
      const uint32_t ws = (kX3 && !p.drop_a_lo) ? (decltype(stage_c)::value & 1u) : decltype(stage_c)::value;
      wait_bar<false>(bar(BAR_W_FULL + ws), (wpar >> ws) & 1u, p.err, 0x200 + ws);
      wpar ^= 1u << ws;
      tc_fence_after();
      if (elect_one()) {
        const uint32_t b_lo = (w0 + ws * kWStageBytes) >> 4;
        const uint32_t a0_lo = a0 >> 4, a1_lo = a1 >> 4;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if ((kmask >> j) & 1u) {
            const uint32_t a_lo = (j < 4 ? a0_lo : a1_lo) + 2 * (j & 3);
            const uint32_t bj = b_lo + (j < 4 ? 0 : (kWBoxBytes >> 4)) + 2 * (j & 3);
            umma_f16_pair(tmem_base + d_col, desc_hi | (uint64_t)a_lo, desc_hi | (uint64_t)bj, idesc,
                          (clear && j == 0) ? 0u : 1u);
          }
        if (a2) {   // lin_in of the precise chains: + x_lo W_hi (K steps 0-2 of the first box)
#pragma unroll
          for (int j = 0; j < 3; ++j)
            umma_f16_pair(tmem_base + d_col, desc_hi | (uint64_t)((a2 >> 4) + 2 * j), desc_hi | (uint64_t)(b_lo + 2 * j),
                          idesc, 1u);
        }
        umma_commit_pair(bar(BAR_W_EMPTY + ws), quad ? 0xF : 0x3);   // frees the stage in every CTA of the cluster
      }
      __syncwarp();
    };
    auto commit = [&](int b) {
      if (elect_one()) umma_commit_pair(bar(b), pair_mask);
      __syncwarp();
    };
    uint32_t xlo_fph = 0;
    auto take_xlo = [&]() -> uint32_t {     // precise chains: the x_lo chunk of this lin_in (0 otherwise)
      if constexpr (!kX3) return 0u;
      wait_bar<false>(bar(BAR_XLO_FULL), xlo_fph, p.err, 0x230);
      xlo_fph ^= 1;
      return smem_base + kOffXlo;
    };
    auto wait_chunk = [&](int c) { wait_bar<false>(bar(BAR_ACT + c), gen & 1, p.err, 0x210 + c); };
    auto wait_act = [&](int c0) {
#pragma unroll 1
      for (int c = c0; c < c0 + 4; ++c) wait_chunk(c);
      tc_fence_after();
    };
    // operand-chunk ring: wait for the next slot, return its address
    auto take_a = [&]() {
      wait_bar<false>(bar(BAR_A_FULL + as), aph, p.err, 0x220 + as);
      const uint32_t addr = smem_base + kOffA + as * kChunkBytes;
      if (++as == kASlots) { as = 0; aph ^= 1; }
      return addr;
    };
    auto release_a = [&](uint32_t addr) { commit(BAR_A_EMPTY + (int)((addr - (smem_base + kOffA)) / kChunkBytes)); };
    const uint32_t act = smem_base + kOffAct;
    const uint32_t act_lo = smem_base + kOffW + 2 * kWStageBytes;   // precise chain only: lo half of the operand
    // 512x512 GEMM from the ACT operand, quarter order (see header), D = 2 blocks at d_col.
    auto gemm_act = [&](uint32_t d_col, bool clear, int commit_region) {
      const int n_ps = kX3 ? 3 - p.drop_a_lo : 1;
#pragma unroll 1
      for (int ps = 0; ps < n_ps; ++ps) {
        // precise chain: three operand passes into the same accumulator -- a_hi W_hi, a_hi W_lo, a_lo W_hi (the
        // producer streams W_hi, W_lo, W_hi); the two-pass form stops after the second.  The accumulators are
        // complete, and committed, in the last pass only
        const uint32_t abase = (kX3 && ps == 2) ? act_lo : act;
        const bool fin = ps == n_ps - 1;
#pragma unroll 1
        for (int kg = 0; kg < 2; ++kg) {
          // K-chunks (h, h+2) of the group: the epilogue finishes chunks 0 and 2 of a block first
          const uint32_t a = abase + kg * 4 * kChunkBytes;
          const bool clr = clear && kg == 0 && ps == 0;
          wait_chunk(kg * 4 + 0);
          wait_chunk(kg * 4 + 2);
          tc_fence_after();
          item(S0{}, a, a + 2 * kChunkBytes, d_col, clr, 0xFFu);
          wait_chunk(kg * 4 + 1);
          wait_chunk(kg * 4 + 3);
          tc_fence_after();
          item(S1{}, a + kChunkBytes, a + 3 * kChunkBytes, d_col, false, 0xFFu);
          if (kg == 1 && fin) commit(BAR_ACC + commit_region * 2);
          item(S2{}, a, a + 2 * kChunkBytes, d_col + 128, clr, 0xFFu);
          item(S3{}, a + kChunkBytes, a + 3 * kChunkBytes, d_col + 128, false, 0xFFu);
          if (kg == 1 && fin) commit(BAR_ACC + commit_region * 2 + 1);
        }
      }
      ++gen;
    };
#pragma unroll 1
    for (int tile = cluster_id; tile < tile_end; tile += n_clusters) {
      if constexpr (kFuse) {
#pragma unroll 1
        for (int k = 0; k < p.n_views; ++k) {
          const uint32_t xa = take_a();
          const uint32_t xl = take_xlo();
          tc_fence_after();
          item(S0{}, xa, xa, kColH, true, 0x77u, xl);
          item(S1{}, xa, xa, kColH + 128, true, 0x77u, xl);
          release_a(xa);
          commit(BAR_ACC + 0);
          commit(BAR_ACC + 1);
#pragma unroll 1
          for (int b = 0; b < 3; ++b) {
            gemm_act(kColT, true, 1);
            gemm_act(kColH, false, 0);
          }
          if (k + 1 < p.n_views) {
            if constexpr (kX3) commit(BAR_XLO_FREE);   // the next sub-tile's x_lo may land in ACT chunk 7
            // the merge epilogue reads this sub-tile's h (and parks its view means): wait for its drain signal
            // before the next sub-tile's lin_in clears h.  After the LAST sub-tile the merge epilogue's signals are
            // the operand-ready signals of block 3's fc_0 instead (consumed by gemm_act below).
            wait_act(0);
            wait_act(4);
            ++gen;
          }
        }
#pragma unroll 1
        for (int b = 3; b < 5; ++b) {
          gemm_act(kColT, true, 1);
          gemm_act(kColH, false, 0);
        }
        if constexpr (kX3) commit(BAR_XLO_FREE);
        wait_act(0);     // the fp32 head has read h
        wait_act(4);
        ++gen;
        continue;
      }
      if (full_front) {
        const uint32_t xa = take_a();
        const uint32_t xl = take_xlo();
        tc_fence_after();
        item(S0{}, xa, xa, kColH, true, 0x77u, xl);  // lin_in: x (K = 48 of the zero-padded 64) times W_hi, then W_lo
        item(S1{}, xa, xa, kColH + 128, true, 0x77u, xl);
        release_a(xa);
        commit(BAR_ACC + 0);                   // the h_0 epilogue adds the lin_z[0] table blend
        commit(BAR_ACC + 1);
      }
#pragma unroll 1
      for (int b = p.first_block; b < p.last_block; ++b) {
        gemm_act(kColT, true, 1);              // t = fc_0(relu(h))
        gemm_act(kColH, false, 0);             // h += fc_1(relu(t))
      }
      if constexpr (kX3) commit(BAR_XLO_FREE);   // the next tile's x_lo may land in ACT chunk 7
      // The epilogue is still reading h out of tensor memory -- for the fp32 head (lin_out on the CUDA cores) or
      // for the dump that ends a per-view segment: wait for its drain signal before the next tile's lin_in clears h.
      wait_act(0);
      wait_act(4);
      ++gen;
    }
  } else if (warp >= kEpiWarp0) {
    // =============================== epilogue warps (both CTAs) =============================
    const int lq = warp & 3;             // TMEM lane quarter this warp may touch
    const int ch = (warp - kEpiWarp0) >> 2;  // which 64 of the 128 accumulator columns
    const int hf = lq >> 1;              // lanes 64-127 hold features [128,256) of a block
    const int r = (lq & 1) * 32 + lane;  // row within the CTA
    const uint32_t lane_addr = ((uint32_t)(lq * 32)) << 16;
    const uint32_t sw = (uint32_t)(r & 7);
    const uint32_t row_off = (uint32_t)((r >> 3) * 1024 + (r & 7) * 128);
    uint32_t n_acc[4] = {0, 0, 0, 0};    // completed waits per (region, block) barrier

    // This warp's share of one accumulator block (64 rows x 256 features per CTA): rows r, and of each
    // of the two 64-feature K-chunks its lane half holds (chunk = nb*4 + 2*hf + half) the 32 features
    // [32*ch, 32*ch+32).  Both warps of a lane quarter work on the same chunk, so chunks 2*hf of a block
    // are complete after half of the epilogue: the issuer consumes K in the order (0,2),(1,3).
    auto ld_bias32 = [&](const float* src, float (&bv)[32]) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(bv[4 * i]), "=f"(bv[4 * i + 1]), "=f"(bv[4 * i + 2]), "=f"(bv[4 * i + 3]) : "l"(src + 4 * i));
    };
    // g[0..32) = this lane's row of the lin_z table blend, 32 features starting at `tab`: sum over the row's
    // four bilinear taps of weight * table row.  Loaded cooperatively -- one 16-byte load instruction covers 8
    // rows x 64 contiguous bytes (lane = (row in group, 16-byte piece)), each lane blends the four taps of its
    // piece in fp32 -- then transposed to the tensor-memory layout (lane = row) through a 2 KB per-warp staging
    // buffer in fp16.  (One row per lane straight from global costs 32 L1 line lookups per instruction for
    // 16 useful bytes each: 8 % of the L1 data path, measured as 76 % busy epilogue warps, profiles/ r01j.)
    const uint32_t stage = smem_base + kOffStage + (uint32_t)(warp - kEpiWarp0) * kStageBytes;
    // Both 32-feature halves of one block for this warp's rows (tab, tab + 64).  All global loads of a half
    // are issued before any of them is consumed (one L2 round trip per half instead of one per 8 rows); the
    // staging accesses carry no compiler memory barrier so the second half's loads can move above the first
    // half's arithmetic -- __syncwarp() orders the staging writes against the reads.
    // The taps record of a row (4 texel indices, 4 weights) sits in bytes [96,128) of its row of the x chunk
    // (pn_point_taps), i.e. in 16-byte pieces 6 and 7 of the 128-byte swizzled row in shared memory; rows past
    // the end of the input are zero-filled by TMA (weight 0, texel 0).
    auto gather_block = [&](const __half* tab, uint32_t xslot, float (&g0)[32], float (&g1)[32]) {
      uint4 ti[4];
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const uint32_t rc = (uint32_t)((lq & 1) * 32 + it * 8 + (lane >> 2));   // row within the CTA
        ti[it] = ld_shared_v4(xslot + rc * 128 + ((6u ^ (rc & 7u)) << 4));
      }
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const uint4* src = reinterpret_cast<const uint4*>(tab + half * 64) + (lane & 3);
        uint4 t[4][4];   // every global load of the half is in flight before the first is consumed
        if (p.debug_stage == -96) {   // timing experiment only (wrong results): blend + staging without the loads
#pragma unroll
          for (int it = 0; it < 4; ++it)
#pragma unroll
            for (int k = 0; k < 4; ++k) t[it][k] = make_uint4(ti[it].x, ti[it].y, ti[it].z, ti[it].w);
        } else {
#pragma unroll
          for (int it = 0; it < 4; ++it) {
            t[it][0] = __ldg(src + (size_t)ti[it].x * 64);
            t[it][1] = __ldg(src + (size_t)ti[it].y * 64);
            t[it][2] = __ldg(src + (size_t)ti[it].z * 64);
            t[it][3] = __ldg(src + (size_t)ti[it].w * 64);
          }
        }
        if (p.debug_stage == -97) {   // timing experiment only (wrong results): the loads without blend and staging
          uint32_t keep = 0;
#pragma unroll
          for (int it = 0; it < 4; ++it)
#pragma unroll
            for (int k = 0; k < 4; ++k) keep ^= t[it][k].x ^ t[it][k].y ^ t[it][k].z ^ t[it][k].w;
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            if (half == 0) g0[i] = __uint_as_float(keep & 0x3f800000u);
            else g1[i] = __uint_as_float(keep & 0x3f800000u);
          }
          continue;
        }
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          const int rr = it * 8 + (lane >> 2);
          const uint32_t rc = (uint32_t)((lq & 1) * 32 + rr);
          const uint4 tw = ld_shared_v4(xslot + rc * 128 + ((7u ^ (rc & 7u)) << 4));
          const float wt[4] = {__uint_as_float(tw.x), __uint_as_float(tw.y), __uint_as_float(tw.z), __uint_as_float(tw.w)};
          float acc[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] = 0.0f;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint32_t tv[4] = {t[it][k].x, t[it][k].y, t[it][k].z, t[it][k].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&tv[j]));
              acc[2 * j] = fmaf(wt[k], f.x, acc[2 * j]);
              acc[2 * j + 1] = fmaf(wt[k], f.y, acc[2 * j + 1]);
            }
          }
          uint32_t pk[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const __half2 h2 = __floats2half2_rn(acc[2 * j], acc[2 * j + 1]);
            pk[j] = *reinterpret_cast<const uint32_t*>(&h2);
          }
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};"
                       ::"r"(stage + rr * 64 + ((uint32_t)((lane & 3) ^ ((rr >> 1) & 3)) << 4)),
                         "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]));
        }
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t tv[4];
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(tv[0]), "=r"(tv[1]), "=r"(tv[2]), "=r"(tv[3])
                       : "r"(stage + lane * 64 + ((uint32_t)(j ^ ((lane >> 1) & 3)) << 4)));
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&tv[i]));
            if (half == 0) { g0[8 * j + 2 * i] = f.x; g0[8 * j + 2 * i + 1] = f.y; }
            else { g1[8 * j + 2 * i] = f.x; g1[8 * j + 2 * i + 1] = f.y; }
          }
        }
        __syncwarp();
      }
    };
    // ---- precise chain (kX3) helpers ----
    // 32 consecutive features of this thread's row -> the 64 bytes of its ACT chunk row: relu + fp16, and for the
    // precise chain also the lo half fp16(relu(x) - hi) into the second operand buffer.
    constexpr uint32_t kActLoOff = (uint32_t)(kOffW + 2 * kWStageBytes - kOffAct);
    auto store_act = [&](uint32_t dst, const float (&b)[32]) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t at = dst + (((uint32_t)(ch * 4 + i) ^ sw) << 4);
        if constexpr (!kX3) {
          st_shared_v4(at, pack_relu_h2(b[8 * i + 0], b[8 * i + 1]), pack_relu_h2(b[8 * i + 2], b[8 * i + 3]),
                       pack_relu_h2(b[8 * i + 4], b[8 * i + 5]), pack_relu_h2(b[8 * i + 6], b[8 * i + 7]));
        } else if (p.drop_a_lo) {   // two-pass form: no a_lo operand
          st_shared_v4(at, pack_relu_h2(b[8 * i + 0], b[8 * i + 1]), pack_relu_h2(b[8 * i + 2], b[8 * i + 3]),
                       pack_relu_h2(b[8 * i + 4], b[8 * i + 5]), pack_relu_h2(b[8 * i + 6], b[8 * i + 7]));
        } else {
          uint32_t hi[4], lo[4];
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const float ra = fmaxf(b[8 * i + 2 * jj], 0.0f), rb = fmaxf(b[8 * i + 2 * jj + 1], 0.0f);
            const __half2 h2 = __floats2half2_rn(ra, rb);
            const float2 back = __half22float2(h2);
            const __half2 l2 = __floats2half2_rn(ra - back.x, rb - back.y);
            hi[jj] = *reinterpret_cast<const uint32_t*>(&h2);
            lo[jj] = *reinterpret_cast<const uint32_t*>(&l2);
          }
          st_shared_v4(at, hi[0], hi[1], hi[2], hi[3]);
          st_shared_v4(at + kActLoOff, lo[0], lo[1], lo[2], lo[3]);
        }
      }
    };
    // fp32 table blend of this warp's 32 rows, both 32-feature halves (tab32, tab32 + 64), for the precise / two-pass
    // chains.  Same cooperative form as gather_block, in fp32 throughout: one 16-byte load instruction covers 8 rows x
    // 64 contiguous bytes (lane = (row in group, 16-byte piece = 4 floats)), each lane blends the four taps of its
    // piece, and the 2 KB staging buffer transposes 32 rows x 16 features at a time to the tensor-memory layout (lane =
    // row), so a 32-feature half takes two passes.  (The first version had every lane read its own four table rows:
    // 32 L1 lines per load instruction for 16 useful bytes each, 4x the wavefronts of this form.)
    auto gather_block_x3 = [&](const float* tab32, uint32_t xslot, float (&g0)[32], float (&g1)[32]) {
      uint4 ti[4];
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const uint32_t rc = (uint32_t)((lq & 1) * 32 + it * 8 + (lane >> 2));   // row within the CTA
        ti[it] = ld_shared_v4(xslot + rc * 128 + ((6u ^ (rc & 7u)) << 4));
      }
#pragma unroll
      for (int pass = 0; pass < 4; ++pass) {      // (half, 16-feature quarter)
        const int half = pass >> 1, q = pass & 1;
        const float4* src = reinterpret_cast<const float4*>(tab32 + half * 64 + q * 16) + (lane & 3);
        float4 t[4][4];   // every global load of the pass is in flight before the first is consumed
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          t[it][0] = __ldg(src + (size_t)ti[it].x * 128);
          t[it][1] = __ldg(src + (size_t)ti[it].y * 128);
          t[it][2] = __ldg(src + (size_t)ti[it].z * 128);
          t[it][3] = __ldg(src + (size_t)ti[it].w * 128);
        }
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          const int rr = it * 8 + (lane >> 2);
          const uint32_t rc = (uint32_t)((lq & 1) * 32 + rr);
          const uint4 tw = ld_shared_v4(xslot + rc * 128 + ((7u ^ (rc & 7u)) << 4));
          const float wt[4] = {__uint_as_float(tw.x), __uint_as_float(tw.y), __uint_as_float(tw.z), __uint_as_float(tw.w)};
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            acc.x = fmaf(wt[k], t[it][k].x, acc.x);
            acc.y = fmaf(wt[k], t[it][k].y, acc.y);
            acc.z = fmaf(wt[k], t[it][k].z, acc.z);
            acc.w = fmaf(wt[k], t[it][k].w, acc.w);
          }
          asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};"
                       ::"r"(stage + rr * 64 + ((uint32_t)((lane & 3) ^ ((rr >> 1) & 3)) << 4)),
                         "f"(acc.x), "f"(acc.y), "f"(acc.z), "f"(acc.w));
        }
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float4 v;
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                       : "r"(stage + lane * 64 + ((uint32_t)(j ^ ((lane >> 1) & 3)) << 4)));
          if (half == 0) { g0[q * 16 + 4 * j] = v.x; g0[q * 16 + 4 * j + 1] = v.y; g0[q * 16 + 4 * j + 2] = v.z; g0[q * 16 + 4 * j + 3] = v.w; }
          else { g1[q * 16 + 4 * j] = v.x; g1[q * 16 + 4 * j + 1] = v.y; g1[q * 16 + 4 * j + 2] = v.z; g1[q * 16 + 4 * j + 3] = v.w; }
        }
        __syncwarp();
      }
    };
    // fp32 dump of (accumulator + bias) of one block: segment end (h_io) and the debug tap
    auto dump_block = [&](uint32_t tcol, const float* bias_vec, int nb, float* dst, int64_t grow) {
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
        const int n0 = 256 * nb + 128 * hf + 64 * half + 32 * ch;
        uint32_t v[32];
        tmem_ld_32x32(tcol + half * 64, v);
        tmem_wait_ld();
        if (grow < p.n_rows) {
          const float4* bp = reinterpret_cast<const float4*>(bias_vec + n0);
          float4* d = reinterpret_cast<float4*>(dst + grow * 512 + n0);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 bv = __ldg(bp + i);
            d[i] = make_float4(__uint_as_float(v[4 * i]) + bv.x, __uint_as_float(v[4 * i + 1]) + bv.y,
                               __uint_as_float(v[4 * i + 2]) + bv.z, __uint_as_float(v[4 * i + 3]) + bv.w);
          }
        }
      }
    };
    // fp16 dump of (accumulator + bias): the hand-off of h from the per-view segment to the merged one.  One
    // rounding of the residual stream to fp16 (the same rounding every GEMM operand of the chain gets) halves the
    // hand-off traffic: 1 KB written per (view, point) row and NS KB read per point.
    auto dump_block16 = [&](uint32_t tcol, const float* bias_vec, int nb, __half* dst, int64_t grow) {
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
        const int n0 = 256 * nb + 128 * hf + 64 * half + 32 * ch;
        uint32_t v[32];
        tmem_ld_32x32(tcol + half * 64, v);
        tmem_wait_ld();
        if (grow < p.n_rows) {
          const float4* bp = reinterpret_cast<const float4*>(bias_vec + n0);
          uint4* d = reinterpret_cast<uint4*>(dst + grow * 512 + n0);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 b0 = __ldg(bp + 2 * i), b1 = __ldg(bp + 2 * i + 1);
            const __half2 h0 = __floats2half2_rn(__uint_as_float(v[8 * i]) + b0.x, __uint_as_float(v[8 * i + 1]) + b0.y);
            const __half2 h1 = __floats2half2_rn(__uint_as_float(v[8 * i + 2]) + b0.z, __uint_as_float(v[8 * i + 3]) + b0.w);
            const __half2 h2 = __floats2half2_rn(__uint_as_float(v[8 * i + 4]) + b1.x, __uint_as_float(v[8 * i + 5]) + b1.y);
            const __half2 h3 = __floats2half2_rn(__uint_as_float(v[8 * i + 6]) + b1.z, __uint_as_float(v[8 * i + 7]) + b1.w);
            d[i] = make_uint4(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1),
                              *reinterpret_cast<const uint32_t*>(&h2), *reinterpret_cast<const uint32_t*>(&h3));
          }
        }
      }
    };
    auto signal_chunk = [&](int c) {
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(bar(BAR_ACT + c), rank4 & 2u);
    };
    // mode 0: relu -> ACT operand; mode 1: dump fp32 to h_io (segment end)
    // gather_b >= 0 (h before blocks 0-2): add the lin_z[gather_b] table blend instead of a bias vector and
    // write the sum back to tensor memory, so the residual stream carries it from here on.
    auto epi = [&](int region, const float* bias_vec, int gather_b, uint32_t xslot, int mode, int64_t grow, bool dbg,
                   int slot, int64_t xrow = 0) {
#pragma unroll 1
      for (int nb = 0; nb < 2; ++nb) {
        const int a = region * 2 + nb;
        const int c0 = nb * 4 + 2 * hf;
        const uint32_t tcol = tmem_base + lane_addr + (region ? kColT : kColH) + nb * 128 + ch * 32;
        // the addend (bias, or the gathered table blend) is fetched while the accumulator is still being produced
        if (p.debug_stage == -99 && mode == 0) {
          // timing experiment only (wrong results): an epilogue of zero length, to measure what the issuer,
          // the weight ring and the tensor pipe reach on their own
          wait_bar<false>(bar(BAR_ACC + a), n_acc[a] & 1, p.err, 0x300 + a);
          ++n_acc[a];
          tc_fence_after();
          tc_fence_before();
          signal_chunk(c0);
          signal_chunk(c0 + 1);
          continue;
        }
        float b0[32], b1[32];
        // block 0: the taps arrive with x, which is only known to be in shared memory (of both CTAs) once
        // lin_in has consumed it, so there the accumulator wait comes first
        if (gather_b == 0) wait_bar<false>(bar(BAR_ACC + a), n_acc[a] & 1, p.err, 0x300 + a);
        if (gather_b >= 0 && p.debug_stage == -98) {   // timing experiment only (wrong results): no table gather
#pragma unroll
          for (int i = 0; i < 32; ++i) b0[i] = b1[i] = 0.0f;
        } else if (gather_b >= 0) {
          if (kSave) {   // the staging buffer may still be the source of the previous TMA store
            if (lane == 0) bulk_wait_read0();
            __syncwarp();
          }
          if constexpr (kX3)
            gather_block_x3(p.tables32 + (size_t)gather_b * p.table_stride + c0 * 64 + 32 * ch, xslot, b0, b1);
          else
            gather_block(p.tables + (size_t)gather_b * p.table_stride + c0 * 64 + 32 * ch, xslot, b0, b1);
        } else {
          ld_bias32(bias_vec + c0 * 64 + 32 * ch, b0);
          ld_bias32(bias_vec + c0 * 64 + 64 + 32 * ch, b1);
        }
        if (gather_b != 0) wait_bar<false>(bar(BAR_ACC + a), n_acc[a] & 1, p.err, 0x300 + a);
        ++n_acc[a];
        tc_fence_after();
        if (dbg && gather_b < 0) dump_block(tcol, bias_vec, nb, p.debug_h, grow);
        if (mode == 1) {
          dump_block16(tcol, bias_vec, nb, p.h_io, grow);
          tc_fence_before();
          signal_chunk(c0);
          signal_chunk(c0 + 1);
          continue;
        }
        uint32_t v0[32], v1[32];
        const uint32_t dst = smem_base + kOffAct + c0 * kChunkBytes + row_off;
        {
          tmem_ld_32x32(tcol, v0);
          tmem_ld_32x32(tcol + 64, v1);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            b0[i] += __uint_as_float(v0[i]);
            b1[i] += __uint_as_float(v1[i]);
          }
        }
        if (gather_b >= 0) {   // h <- h + blend: read-modify-write while no MMA touches h
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            v0[i] = __float_as_uint(b0[i]);
            v1[i] = __float_as_uint(b1[i]);
          }
          tmem_st_32x32(tcol, v0);
          tmem_st_32x32(tcol + 64, v1);
          if (dbg && grow < p.n_rows) {
            float4* d0 = reinterpret_cast<float4*>(p.debug_h + grow * 512 + c0 * 64 + 32 * ch);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              d0[i] = make_float4(b0[4 * i], b0[4 * i + 1], b0[4 * i + 2], b0[4 * i + 3]);
              d0[16 + i] = make_float4(b1[4 * i], b1[4 * i + 1], b1[4 * i + 2], b1[4 * i + 3]);
            }
          }
        }
        store_act(dst, b0);
        fence_proxy_async_smem();
        if (gather_b >= 0) tmem_wait_st();
        tc_fence_before();
        signal_chunk(c0);
        if (kSave)
          save_relu32(&map_s, stage, lane, c0 * 64 + 32 * ch, (int64_t)slot * p.save_rows + (grow - lane),
                      grow - lane < p.n_rows, b0);
        store_act(dst + kChunkBytes, b1);
        fence_proxy_async_smem();
        signal_chunk(c0 + 1);
        if (kSave)
          save_relu32(&map_s, stage, lane, c0 * 64 + 64 + 32 * ch, (int64_t)slot * p.save_rows + (grow - lane),
                      grow - lane < p.n_rows, b1);
      }
    };
    // Segment [3,5] entry: h = mean over views of the fp32 h the first segment wrote.
    auto epi_load = [&](int64_t grow) {
      const bool valid = grow < p.n_rows;
      const int64_t obj = valid ? grow / p.ppo : 0, pt = valid ? grow - obj * p.ppo : 0;
      const __half* src = p.h_io + ((obj * p.n_views) * p.ppo + pt) * 512;
      const float inv = (float)p.n_views;
      for (int nb = 0; nb < 2; ++nb) {
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
          const int c = nb * 4 + 2 * hf + half;
          const int n0 = c * 64 + 32 * ch;
          const uint32_t chunk = smem_base + kOffAct + c * kChunkBytes + row_off;
          float x[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) x[i] = 0.0f;
          if (valid) {
            for (int v = 0; v < p.n_views; ++v) {
              const uint4* s4 = reinterpret_cast<const uint4*>(src + (int64_t)v * p.ppo * 512 + n0);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 t = s4[i];
                const uint32_t tw[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&tw[j]));
                  x[8 * i + 2 * j] += f.x; x[8 * i + 2 * j + 1] += f.y;
                }
              }
            }
#pragma unroll
            for (int i = 0; i < 32; ++i) x[i] = x[i] / inv;   // torch.mean: sum in view order, then divide
          }
          uint32_t v[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(x[i]);
          tmem_st_32x32(tmem_base + lane_addr + kColH + nb * 128 + half * 64 + ch * 32, v);
#pragma unroll
          for (int i = 0; i < 4; ++i)
            st_shared_v4(chunk + (((uint32_t)(ch * 4 + i) ^ sw) << 4), pack_relu_h2(x[8 * i + 0], x[8 * i + 1]),
                         pack_relu_h2(x[8 * i + 2], x[8 * i + 3]), pack_relu_h2(x[8 * i + 4], x[8 * i + 5]),
                         pack_relu_h2(x[8 * i + 6], x[8 * i + 7]));
          tmem_wait_st();
          fence_proxy_async_smem();
          tc_fence_before();
          signal_chunk(c);
          if (kSave)
            save_relu32(&map_s, stage, lane, n0, (int64_t)(2 * p.first_block) * p.save_rows + (grow - lane),
                        grow - lane < p.n_rows, x);
        }
      }
    };

    // lin_out + head (resnetfc.py:183, models.py:257-266) in fp32 on the CUDA cores.  The four outputs are the one
    // place where rounding the operand to fp16 shows: sigma = w . relu(h_5) sums 512 rounded activations against
    // the largest weights of the network, an error that goes straight into alpha (a float64 ablation on the
    // real-size DTU slice: 1.8e-3 of the 2.4e-3 depth error caused by activation rounding came from this one layer).
    // Every epilogue thread already holds its 128 of the row's 512 values of h_5 in fp32, so it forms its part of
    // the four dot products (512 FFMA per thread and tile against warp-uniform weight loads); the four threads of
    // a row meet through 4 KB of the tile's x slot (free since block 2's table gather) and one of them applies
    // sigmoid / relu and stores.  No tensor-core work, no ACT operand, no weight tiles for this layer.
    auto head_ffma = [&](int64_t grow, uint32_t xslot, const float* tbl) {
      const float* bias_vec = tbl + 5 * 512;
      const float* w_out = tbl + 12 * 512;
      float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
      const bool dbg = p.debug_h != nullptr && p.debug_stage == 5;
#pragma unroll 1
      for (int nb = 0; nb < 2; ++nb) {
        const int c0 = nb * 4 + 2 * hf;
        const uint32_t tcol = tmem_base + lane_addr + kColH + nb * 128 + ch * 32;
        wait_bar<false>(bar(BAR_ACC + nb), n_acc[nb] & 1, p.err, 0x320 + nb);
        ++n_acc[nb];
        tc_fence_after();
        if (dbg) dump_block(tcol, bias_vec, nb, p.debug_h, grow);
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
          const int n0 = c0 * 64 + 64 * half + 32 * ch;
          float x[32];
          ld_bias32(bias_vec + n0, x);
          uint32_t v[32];
          tmem_ld_32x32(tcol + half * 64, v);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) x[i] += __uint_as_float(v[i]);
          if (kSave)   // slot 10: relu(h_5), the operand of lin_out's weight gradient
            save_relu32(&map_s, stage, lane, n0, (int64_t)10 * p.save_rows + (grow - lane), grow - lane < p.n_rows, x);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4* wj = reinterpret_cast<const float4*>(w_out + j * 512 + n0);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 w4 = __ldg(wj + i);     // warp-uniform address: one broadcast transaction
              acc[j] = fmaf(fmaxf(x[4 * i], 0.0f), w4.x, acc[j]);
              acc[j] = fmaf(fmaxf(x[4 * i + 1], 0.0f), w4.y, acc[j]);
              acc[j] = fmaf(fmaxf(x[4 * i + 2], 0.0f), w4.z, acc[j]);
              acc[j] = fmaf(fmaxf(x[4 * i + 3], 0.0f), w4.w, acc[j]);
            }
          }
        }
        // h_5 of this block is in registers: the next tile's lin_in may overwrite it
        tc_fence_before();
        signal_chunk(c0);
        signal_chunk(c0 + 1);
      }
      const uint32_t slot4 = xslot + (uint32_t)(((hf * 2 + ch) * kRows + r) * 16);
      st_shared_v4(slot4, __float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
      asm volatile("bar.sync 1, 256;" ::: "memory");     // the eight epilogue warps of this CTA
      if (hf == 0 && ch == 0) {
        float s[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
        for (int g = 0; g < 4; ++g) {    // fixed order: deterministic from run to run
          const uint4 t = ld_shared_v4(xslot + (uint32_t)((g * kRows + r) * 16));
          s[0] += __uint_as_float(t.x); s[1] += __uint_as_float(t.y);
          s[2] += __uint_as_float(t.z); s[3] += __uint_as_float(t.w);
        }
        if (grow < p.n_rows) {
          const float4 bo = __ldg(reinterpret_cast<const float4*>(tbl + 11 * 512));
          const float s0 = s[0] + bo.x, s1 = s[1] + bo.y, s2 = s[2] + bo.z, s3 = s[3] + bo.w;
          if (p.pre) p.pre[grow] = make_float4(s0, s1, s2, s3);
          p.out[grow] = make_float4(1.0f / (1.0f + expf(-s0)), 1.0f / (1.0f + expf(-s1)),
                                    1.0f / (1.0f + expf(-s2)), fmaxf(s3, 0.0f));
        }
      }
      // the slot is refilled by TMA two tiles from now: order these generic-proxy accesses in front of it
      fence_proxy_async_smem();
    };

    // Fused view mean (kFuse; resnetfc.py:152-172, util.py:461-466), at the end of block 2 of sub-tile k_cur.
    // Sub-tile v of a super-tile is VIEW v of the super-tile's 128 points (row m of every sub-tile = point m), so the
    // sum over views is a per-thread running sum across sub-tiles: after sub-tile 0 the thread parks its h (fp32, 128
    // values, written and later read by the SAME thread, piece-major so that a warp's 16-byte accesses are contiguous;
    // L2 resident), after sub-tile v it adds its h to the parked sum -- in view order, like torch.mean's sum -- and
    // after the last one divides by n_views, rounds the mean to fp16 once, writes it to tensor memory as the merged
    // tile's residual stream and relu(mean) to the ACT operand: block 3 starts on the same 128 rows, now one per point.
    auto merge = [&](int k_cur) {
      const bool first = k_cur == 0, last = k_cur == p.n_views - 1;
      const float* bias_vec = p.bias + 3 * 512;   // fc_1 bias of block 2 (everything earlier sits in the tables)
      const float inv = (float)p.n_views;
      float4* park = reinterpret_cast<float4*>(p.park) + ((size_t)blockIdx.x * 8 + (size_t)(warp - kEpiWarp0)) * (32 * 32) + lane;
#pragma unroll 1
      for (int nb = 0; nb < 2; ++nb) {
        const int c0 = nb * 4 + 2 * hf;
        const uint32_t tcol = tmem_base + lane_addr + kColH + nb * 128 + ch * 32;
        wait_bar<false>(bar(BAR_ACC + nb), n_acc[nb] & 1, p.err, 0x330 + nb);
        ++n_acc[nb];
        tc_fence_after();
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
          const int piece = nb * 2 + half;
          const int n0 = c0 * 64 + 64 * half + 32 * ch;
          float4* pp = park + piece * 8 * 32;
          float x[32];
          if (first) {
            ld_bias32(bias_vec + n0, x);
          } else {   // the running sum of views 0 .. k_cur-1 (their biases included), requested before the accumulator
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 t4 = __ldcg(pp + i * 32);
              x[4 * i] = t4.x; x[4 * i + 1] = t4.y; x[4 * i + 2] = t4.z; x[4 * i + 3] = t4.w;
            }
          }
          uint32_t v[32];
          tmem_ld_32x32(tcol + half * 64, v);
          tmem_wait_ld();
          if (first) {
#pragma unroll
            for (int i = 0; i < 32; ++i) x[i] += __uint_as_float(v[i]);
          } else {
            float bv[32];
            ld_bias32(bias_vec + n0, bv);
#pragma unroll
            for (int i = 0; i < 32; ++i) x[i] += __uint_as_float(v[i]) + bv[i];     // sum_so_far + (h_v + bias)
          }
          if (!last) {
#pragma unroll
            for (int i = 0; i < 8; ++i) __stcg(pp + i * 32, make_float4(x[4 * i], x[4 * i + 1], x[4 * i + 2], x[4 * i + 3]));
            continue;
          }
          float f[32];
#pragma unroll
          for (int i = 0; i < 16; ++i) {   // mean; one fp16 rounding in the fast chain, none in the precise one
            if constexpr (kX3) {
              f[2 * i] = x[2 * i] / inv; f[2 * i + 1] = x[2 * i + 1] / inv;
            } else {
              const float2 t2 = __half22float2(__floats2half2_rn(x[2 * i] / inv, x[2 * i + 1] / inv));
              f[2 * i] = t2.x; f[2 * i + 1] = t2.y;
            }
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(f[i]);
          tmem_st_32x32(tcol + half * 64, v);
          store_act(smem_base + kOffAct + (c0 + half) * kChunkBytes + row_off, f);
          tmem_wait_st();
          fence_proxy_async_smem();
          tc_fence_before();
          signal_chunk(c0 + half);
        }
        if (!last) {   // h of this block is consumed: the next sub-tile's lin_in may clear it
          tc_fence_before();
          signal_chunk(c0);
          signal_chunk(c0 + 1);
        }
      }
    };

    uint32_t xs = 0;   // x slot of the current tile (the producer fills the slots round robin)
    for (int tile = cluster_id; tile < tile_end; tile += n_clusters) {
      if constexpr (kFuse) {
        const int64_t none = p.n_rows;      // no per-row side outputs (debug dump, saved operands) in this mode
        uint32_t xslot = 0;
#pragma unroll 1
        for (int k = 0; k < p.n_views; ++k) {
          xslot = smem_base + kOffA + xs * kChunkBytes;
          if (++xs == kASlots) xs = 0;
#pragma unroll 1
          for (int b = 0; b < 3; ++b) {
            epi(0, p.bias + b * 512, b, xslot, 0, none, false, 2 * b,
                ((int64_t)tile * p.n_views + k) * kTileRows + rank * kRows + r);
            epi(1, p.bias + (6 + b) * 512, -1, 0u, 0, none, false, 2 * b + 1);
          }
          merge(k);
        }
        epi(1, p.bias2 + (6 + 3) * 512, -1, 0u, 0, none, false, 7);
        epi(0, p.bias2 + 4 * 512, -1, 0u, 0, none, false, 8);
        epi(1, p.bias2 + (6 + 4) * 512, -1, 0u, 0, none, false, 9);
        // row m of the merged tile = point m of the super-tile
        head_ffma((int64_t)tile * kTileRows + rank * kRows + r, xslot, p.bias2);
        continue;
      }
      const int64_t grow = (int64_t)tile * kTileRows + rank * kRows + r;
      const uint32_t xslot = smem_base + kOffA + xs * kChunkBytes;
      if (++xs == kASlots) xs = 0;
      if (!full_front) epi_load(grow);
      for (int b = p.first_block; b < p.last_block; ++b) {
        if (b > p.first_block || full_front)
          epi(0, p.bias + b * 512, b < 3 ? b : -1, xslot, 0, grow, p.debug_h != nullptr && p.debug_stage == b, 2 * b, grow);
        epi(1, p.bias + (6 + b) * 512, -1, 0u, 0, grow, false, 2 * b + 1);
      }
      if (has_head) {
        head_ffma(grow, xslot, p.bias);
      } else {
        epi(0, p.bias + p.last_block * 512, -1, 0u, 1, grow, false, -1);
      }
    }
  }
  __syncwarp();
  if (kSave) bulk_wait0();   // this thread's TMA stores (lane 0 of the epilogue warps) are complete
  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) tmem_dealloc_pair(tmem_base, 512);
}

// ---- weight packing -----------------------------------------------------------------------
__global__ void __launch_bounds__(256)
pack_weights_kernel(pn_mlp_weights_t w, __half* __restrict__ dst) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)kLayers * 512 * 512) return;
  const int layer = (int)(i / (512 * 512));
  const int n = (int)((i / 512) % 512), k = (int)(i % 512);
  // fp16 rounding of the weights is an error that is coherent over all samples of a ray (the same perturbed
  // network everywhere); lin_in is cheap and carries a large share of it, so it is stored as
  // hi + lo: the lo part in K columns [64, 64+42) (the x chunk is multiplied by both).  lin_out no longer runs on
  // the tensor cores at all (fp32 head on the CUDA cores, head_ffma); its slot of the packed buffer is unused.
  float v = 0.0f;
  if (layer == L_IN) {
    if (k < PN_D_CODE) {
      v = w.lin_in_w[n * PN_D_CODE + k];
    } else if (k >= 64 && k < 64 + PN_D_CODE) {
      const float f = w.lin_in_w[n * PN_D_CODE + k - 64];
      v = f - __half2float(__float2half_rn(f));
    }
  } else if (layer < L_FC) {
    v = w.lin_z_w[layer - L_Z0][n * 512 + k];
  } else if (layer < L_OUT) {
    const int b = (layer - L_FC) >> 1;
    v = ((layer - L_FC) & 1) ? w.fc1_w[b][n * 512 + k] : w.fc0_w[b][n * 512 + k];
  } else if (n < 4) {
    v = w.lin_out_w[n * 512 + k];
  } else if (n < 8) {
    const float f = w.lin_out_w[(n - 4) * 512 + k];
    v = (f - __half2float(__float2half_rn(f))) * 1024.0f;
  }
  dst[i] = __float2half_rn(v);
}

// Precise chain (fp16x3): the lo halves of the ten 512x512 matrices, W_lo = fp16(W - fp16(W)), as layers
// [kLayers + L_FC, kLayers + L_OUT) of the packed buffer (the other slots of the second half stay unused).
__global__ void __launch_bounds__(256)
pack_weights_lo_kernel(pn_mlp_weights_t w, __half* __restrict__ dst) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)(L_OUT - L_FC) * 512 * 512) return;
  const int layer = L_FC + (int)(i / (512 * 512));
  const int n = (int)((i / 512) % 512), k = (int)(i % 512);
  const int b = (layer - L_FC) >> 1;
  const float v = ((layer - L_FC) & 1) ? w.fc1_w[b][n * 512 + k] : w.fc0_w[b][n * 512 + k];
  dst[((size_t)(kLayers + layer) * 512 + n) * 512 + k] = __float2half_rn(v - __half2float(__float2half_rn(v)));
}


// table A: cumulative from lin_in; table B: cumulative from block 3 (h handed over with biases in).
__global__ void __launch_bounds__(512)
pack_bias_kernel(pn_mlp_weights_t w, float* __restrict__ ta, float* __restrict__ tb) {
  const int n = threadIdx.x;
  // rows 0-2 unused: lin_in.bias, lin_z biases and fc_1 biases of blocks 0-1 are folded into the lin_z tables,
  // whose blend the epilogue writes into h; from block 3 on h is read as h + (fc_1 biases of blocks 2..b-1)
  ta[0 * 512 + n] = ta[1 * 512 + n] = ta[2 * 512 + n] = 0.0f;
  float cb = 0.0f;
  for (int b = 2; b < 5; ++b) {
    cb += w.fc1_b[b][n];
    ta[(b + 1) * 512 + n] = cb;
  }
  tb[0 * 512 + n] = tb[1 * 512 + n] = tb[2 * 512 + n] = tb[3 * 512 + n] = 0.0f;
  tb[4 * 512 + n] = w.fc1_b[3][n];
  tb[5 * 512 + n] = w.fc1_b[3][n] + w.fc1_b[4][n];
  for (int b = 0; b < 5; ++b) ta[(6 + b) * 512 + n] = tb[(6 + b) * 512 + n] = w.fc0_b[b][n];
  ta[11 * 512 + n] = tb[11 * 512 + n] = n < 4 ? w.lin_out_b[n] : 0.0f;
  for (int j = 0; j < 4; ++j) ta[(12 + j) * 512 + n] = tb[(12 + j) * 512 + n] = w.lin_out_w[j * 512 + n];
}

// ---- host side ----------------------------------------------------------------------------
static int launch_chain(const CUtensorMap& mw, const CUtensorMap& mx, const CUtensorMap& ms, ChainParams p,
                        cudaStream_t st) {
  static pn::DeviceOnce attr_set;  // per device; idempotent, a benign race at worst repeats the call
  int attr_dev;
  if (attr_set.needed(&attr_dev)) {
    PN_CUDA_OK(cudaFuncSetAttribute(mlp_chain_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemAlloc));
    PN_CUDA_OK(cudaFuncSetAttribute(mlp_chain_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemAlloc));
    PN_CUDA_OK(cudaFuncSetAttribute(mlp_chain_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemAlloc));
    PN_CUDA_OK(cudaFuncSetAttribute(mlp_chain_kernel<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemAlloc));
    PN_CUDA_OK(cudaFuncSetAttribute(mlp_chain_kernel<false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemAlloc));
    attr_set.mark(attr_dev);
  }
  int dev = 0, sms = 0;
  PN_CUDA_OK(cudaGetDevice(&dev));
  PN_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  p.n_tiles = (int)((p.n_rows + kTileRows - 1) / kTileRows);
  if (p.group > 0)   // fused view mean: the persistent loop runs over super-tiles of 128 points
    p.n_tiles = (int)((p.n_points + kTileRows - 1) / kTileRows);
  // One CTA per SM, in groups of four where the GPU can co-schedule four (preferred cluster dimension) and
  // pairs elsewhere; no more groups than the tiles need.
  int max_groups = sms / 4;
  if (p.group > 0 && max_groups > kParkCtas / 4) max_groups = kParkCtas / 4;   // the park buffer is sized for kParkCtas CTAs
  const int want = (p.n_tiles + 1) / 2;
  const int groups = want < max_groups ? want : max_groups;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(4 * groups);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kSmemAlloc;
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = 2; attrs[0].val.clusterDim.y = 1; attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributePreferredClusterDimension;
  attrs[1].val.preferredClusterDim.x = 4; attrs[1].val.preferredClusterDim.y = 1; attrs[1].val.preferredClusterDim.z = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = 2;
  if (p.tables32 && p.group > 0) PN_CUDA_OK(cudaLaunchKernelEx(&cfg, mlp_chain_kernel<false, true, true>, mw, mx, ms, p));
  else if (p.tables32) PN_CUDA_OK(cudaLaunchKernelEx(&cfg, mlp_chain_kernel<false, false, true>, mw, mx, ms, p));
  else if (p.group > 0) PN_CUDA_OK(cudaLaunchKernelEx(&cfg, mlp_chain_kernel<false, true>, mw, mx, ms, p));
  else if (p.save) PN_CUDA_OK(cudaLaunchKernelEx(&cfg, mlp_chain_kernel<true>, mw, mx, ms, p));
  else PN_CUDA_OK(cudaLaunchKernelEx(&cfg, mlp_chain_kernel<false>, mw, mx, ms, p));
  PN_LAUNCH_OK("mlp_chain_kernel");
  count_launch();
  return PN_OK;
}

}  // namespace tc
}  // namespace pn

extern "C" size_t pn_mlp_tc_packed_bytes(void) {
  return pn::tc::kPackedWeightBytes + 2 * pn::tc::kBiasBytes;
}

extern "C" int pn_prepare_weights(const pn_mlp_weights_t* w, void* packed, size_t packed_bytes, pn_stream_t stream) {
  using namespace pn::tc;
  PN_CHECK_ARG(w && packed, "pn_prepare_weights: null pointer");
  PN_CHECK_ARG(packed_bytes >= pn_mlp_tc_packed_bytes(), "pn_prepare_weights: buffer too small");
  PN_CHECK_ARG(((uintptr_t)packed & 255) == 0, "pn_prepare_weights: buffer must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t total = (int64_t)kLayers * 512 * 512;
  pack_weights_kernel<<<pn::ceil_div(total, 256), 256, 0, st>>>(*w, (__half*)packed);
  PN_LAUNCH_OK("pack_weights_kernel");
  float* ta = (float*)((char*)packed + kPackedWeightBytes);
  float* tb = (float*)((char*)packed + kPackedWeightBytes + kBiasBytes);
  pack_bias_kernel<<<1, 512, 0, st>>>(*w, ta, tb);
  PN_LAUNCH_OK("pack_bias_kernel");
  pn::count_launch(2);
  return PN_OK;
}

// packed buffer of the precise chain: [hi layers][lo layers][bias table A][bias table B]
constexpr size_t kX3BiasOff = 2 * pn::tc::kPackedWeightBytes;
extern "C" size_t pn_mlp_tc_packed_x3_bytes(void) { return kX3BiasOff + 2 * pn::tc::kBiasBytes; }

extern "C" int pn_prepare_weights_x3(const pn_mlp_weights_t* w, void* packed, size_t packed_bytes, pn_stream_t stream) {
  using namespace pn::tc;
  PN_CHECK_ARG(w && packed, "pn_prepare_weights_x3: null pointer");
  PN_CHECK_ARG(packed_bytes >= pn_mlp_tc_packed_x3_bytes(), "pn_prepare_weights_x3: buffer too small");
  PN_CHECK_ARG(((uintptr_t)packed & 255) == 0, "pn_prepare_weights_x3: buffer must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t total = (int64_t)kLayers * 512 * 512;
  pack_weights_kernel<<<pn::ceil_div(total, 256), 256, 0, st>>>(*w, (__half*)packed);
  PN_LAUNCH_OK("pack_weights_kernel");
  PN_CUDA_OK(cudaMemsetAsync((char*)packed + kPackedWeightBytes, 0, kPackedWeightBytes, st));
  pack_weights_lo_kernel<<<pn::ceil_div((int64_t)(L_OUT - L_FC) * 512 * 512, 256), 256, 0, st>>>(*w, (__half*)packed);
  PN_LAUNCH_OK("pack_weights_lo_kernel");
  pack_bias_kernel<<<1, 512, 0, st>>>(*w, (float*)((char*)packed + kX3BiasOff), (float*)((char*)packed + kX3BiasOff + kBiasBytes));
  PN_LAUNCH_OK("pack_bias_kernel");
  pn::count_launch(3);
  return PN_OK;
}

extern "C" size_t pn_mlp_tc_workspace_bytes(int64_t n_rows, int n_views) {
  // fp16 h of every (view, point) row handed from the per-view segment to the merged one, + error word
  return 256 + (n_views > 1 && n_rows > 0 ? (size_t)n_rows * 512 * sizeof(__half) : 0);
}

// rows per slot of the saved-operand buffer: whole 128-row tiles (the TMA stores of a ragged last tile end in the
// padding instead of the next slot)
static inline int64_t acts_rows(int64_t n_rows) { return (n_rows > 0 ? n_rows + 127 : 0) / 128 * 128; }

extern "C" size_t pn_mlp_tc_acts_bytes(int64_t n_rows) {
  return (size_t)acts_rows(n_rows) * 512 * 2 * 11 + 256;
}

static int mlp_forward_tc_impl(const void* packed, const void* tables, int64_t n_texels, const void* feat_x,
                               int n_objects, int n_views, int64_t points_per_object, float* out, float* pre_act,
                               void* workspace, size_t workspace_bytes, const pn_mlp_tc_debug_t* dbg, void* acts,
                               pn_stream_t stream);

extern "C" int pn_mlp_forward_tc(const void* packed, const void* tables, int64_t n_texels,
                                 const void* feat_x, int n_objects,
                                 int n_views, int64_t points_per_object, float* out, float* pre_act,
                                 void* workspace, size_t workspace_bytes, const pn_mlp_tc_debug_t* dbg,
                                 pn_stream_t stream) {
  return mlp_forward_tc_impl(packed, tables, n_texels, feat_x, n_objects, n_views, points_per_object, out, pre_act,
                             workspace, workspace_bytes, dbg, nullptr, stream);
}

extern "C" int pn_mlp_forward_tc_train(const void* packed, const void* tables, int64_t n_texels,
                                       const void* feat_x, int n_objects, int n_views, int64_t points_per_object,
                                       float* out, float* pre_act, void* workspace, size_t workspace_bytes,
                                       void* acts, size_t acts_bytes, pn_stream_t stream) {
  PN_CHECK_ARG(acts && ((uintptr_t)acts & 31) == 0, "pn_mlp_forward_tc_train: activation buffer missing or misaligned");
  PN_CHECK_ARG(acts_bytes >= pn_mlp_tc_acts_bytes((int64_t)n_objects * n_views * points_per_object),
               "pn_mlp_forward_tc_train: activation buffer too small");
  return mlp_forward_tc_impl(packed, tables, n_texels, feat_x, n_objects, n_views, points_per_object, out, pre_act,
                             workspace, workspace_bytes, nullptr, acts, stream);
}

static int mlp_forward_tc_impl(const void* packed, const void* tables, int64_t n_texels, const void* feat_x,
                               int n_objects, int n_views, int64_t points_per_object, float* out, float* pre_act,
                               void* workspace, size_t workspace_bytes, const pn_mlp_tc_debug_t* dbg, void* acts,
                               pn_stream_t stream) {
  using namespace pn::tc;
  PN_CHECK_ARG(packed && tables && feat_x && out && workspace, "pn_mlp_forward_tc: null pointer");
  PN_CHECK_ARG(n_texels > 0, "pn_mlp_forward_tc: no texels");
  PN_CHECK_ARG(n_objects > 0 && n_views > 0 && points_per_object >= 0, "pn_mlp_forward_tc: bad sizes");
  const int64_t n_rows = (int64_t)n_objects * n_views * points_per_object;
  const int64_t n_pts = (int64_t)n_objects * points_per_object;
  if (n_rows == 0) return PN_OK;
  PN_CHECK_ARG(n_rows < (1ll << 31) - 256, "pn_mlp_forward_tc: too many rows for one launch");
  PN_CHECK_ARG(workspace_bytes >= pn_mlp_tc_workspace_bytes(n_rows, n_views), "pn_mlp_forward_tc: workspace too small");
  PN_CHECK_ARG(((uintptr_t)workspace & 255) == 0 && ((uintptr_t)tables & 15) == 0 &&
                   ((uintptr_t)feat_x & 15) == 0,
               "pn_mlp_forward_tc: misaligned buffer");
  cudaStream_t st = (cudaStream_t)stream;
  CUtensorMap mw, mx;
  int rc;
  if ((rc = make_map(&mw, packed, (uint64_t)kLayers * 512, 512, 1024, 128))) return rc;
  if ((rc = make_map(&mx, feat_x, (uint64_t)n_rows, 64, 128, kRows))) return rc;
  unsigned int* err = (unsigned int*)workspace;
  PN_CUDA_OK(cudaMemsetAsync(err, 0, 4, st));
  ChainParams p = {};
  p.out = (float4*)out;
  p.pre = (float4*)pre_act;
  p.err = err;
  p.tables = (const __half*)tables;
  p.table_stride = n_texels * 512;
  p.ppo = points_per_object;
  p.n_views = n_views;
  p.debug_h = dbg ? dbg->h : nullptr;
  p.debug_stage = dbg ? dbg->stage : -1;
  p.save = (uint16_t*)acts;
  p.save_rows = acts_rows(n_rows);
  CUtensorMap ms = mw;   // placeholder when nothing is saved (the kernel variant without stores never reads it)
  if (acts && (rc = make_store_map(&ms, acts, (uint64_t)(11 * p.save_rows)))) return rc;
  const float* ta = (const float*)((const char*)packed + kPackedWeightBytes);
  const float* tb = (const float*)((const char*)packed + kPackedWeightBytes + kBiasBytes);
  if (n_views == 1) {
    p.bias = ta; p.n_rows = n_rows; p.first_block = 0; p.last_block = 5;
    return launch_chain(mw, mx, ms, p, st);
  }
  // NS > 1: blocks 0-2 per (view, point) row, then the view mean (resnetfc.py:152-172) feeds blocks 3-4
  p.h_io = (__half*)((char*)workspace + 256);
  p.bias = ta; p.n_rows = n_rows; p.first_block = 0; p.last_block = 3;
  if ((rc = launch_chain(mw, mx, ms, p, st))) return rc;
  p.bias = tb; p.n_rows = n_pts; p.first_block = 3; p.last_block = 5;
  p.debug_h = (dbg && dbg->stage >= 4) ? dbg->h : nullptr;
  return launch_chain(mw, mx, ms, p, st);
}

// ---- fused view mean: one launch for n_views > 1 -----------------------------------------------------------------
extern "C" size_t pn_mlp_tc_grouped_workspace_bytes(void) {
  // watchdog word + the park buffer (each CTA's running sum of h over the views of its super-tile)
  return 256 + (size_t)pn::tc::kParkCtas * pn::tc::kParkBytesPerCta;
}

extern "C" int pn_mlp_forward_tc_grouped(const void* packed, const void* tables, int64_t n_texels,
                                         const void* feat_x_grouped, int n_objects, int n_views,
                                         int64_t points_per_object, float* out, float* pre_act, void* workspace,
                                         size_t workspace_bytes, pn_stream_t stream) {
  using namespace pn::tc;
  PN_CHECK_ARG(packed && tables && feat_x_grouped && out && workspace, "pn_mlp_forward_tc_grouped: null pointer");
  PN_CHECK_ARG(n_texels > 0, "pn_mlp_forward_tc_grouped: no texels");
  PN_CHECK_ARG(n_objects > 0 && n_views > 1 && n_views <= 64 && points_per_object >= 0,
               "pn_mlp_forward_tc_grouped: bad sizes (2 to 64 views; one view: pn_mlp_forward_tc)");
  const int64_t n_pts = (int64_t)n_objects * points_per_object;
  if (n_pts == 0) return PN_OK;
  const int64_t n_rows = pn_point_taps_grouped_rows(n_pts, n_views);
  PN_CHECK_ARG(n_rows < (1ll << 31) - 4096, "pn_mlp_forward_tc_grouped: too many rows for one launch");
  PN_CHECK_ARG(workspace_bytes >= pn_mlp_tc_grouped_workspace_bytes(), "pn_mlp_forward_tc_grouped: workspace too small");
  PN_CHECK_ARG(((uintptr_t)workspace & 255) == 0 && ((uintptr_t)tables & 15) == 0 && ((uintptr_t)feat_x_grouped & 15) == 0,
               "pn_mlp_forward_tc_grouped: misaligned buffer");
  cudaStream_t st = (cudaStream_t)stream;
  CUtensorMap mw, mx;
  int rc;
  if ((rc = make_map(&mw, packed, (uint64_t)kLayers * 512, 512, 1024, 128))) return rc;
  if ((rc = make_map(&mx, feat_x_grouped, (uint64_t)n_rows, 64, 128, kRows))) return rc;
  unsigned int* err = (unsigned int*)workspace;
  PN_CUDA_OK(cudaMemsetAsync(err, 0, 4, st));
  ChainParams p = {};
  p.out = (float4*)out;
  p.pre = (float4*)pre_act;
  p.err = err;
  p.tables = (const __half*)tables;
  p.table_stride = n_texels * 512;
  p.ppo = points_per_object;
  p.n_views = n_views;
  p.debug_stage = -1;
  p.bias = (const float*)((const char*)packed + kPackedWeightBytes);
  p.bias2 = (const float*)((const char*)packed + kPackedWeightBytes + kBiasBytes);
  p.group = 1;
  p.n_points = n_pts;
  p.n_rows = n_pts;          // what the head's row check compares against
  p.first_block = 0;
  p.last_block = 5;
  p.park = (float*)((char*)workspace + 256);
  return launch_chain(mw, mx, mw, p, st);
}

// ---- precise chain: mlp_precision = fp16x3 ------------------------------------------------------------------------
static int mlp_forward_tc_pairs(const void* packed_x3, const float* tables32, int64_t n_texels, const void* feat_x,
                                const void* feat_x_lo, int n_objects, int n_views, int64_t points_per_object,
                                float* out, float* pre_act, void* workspace, size_t workspace_bytes, int drop_a_lo,
                                pn_stream_t stream) {
  using namespace pn::tc;
  PN_CHECK_ARG(packed_x3 && tables32 && feat_x && feat_x_lo && out && workspace, "pn_mlp_forward_tc_x3: null pointer");
  PN_CHECK_ARG(n_texels > 0, "pn_mlp_forward_tc_x3: no texels");
  PN_CHECK_ARG(n_objects > 0 && n_views > 0 && n_views <= 64 && points_per_object >= 0, "pn_mlp_forward_tc_x3: bad sizes");
  const int64_t n_pts = (int64_t)n_objects * points_per_object;
  if (n_pts == 0) return PN_OK;
  const bool fused = n_views > 1;
  const int64_t x_rows = fused ? pn_point_taps_grouped_rows(n_pts, n_views) : n_pts;
  PN_CHECK_ARG(x_rows < (1ll << 31) - 4096, "pn_mlp_forward_tc_x3: too many rows for one launch");
  PN_CHECK_ARG(workspace_bytes >= pn_mlp_tc_grouped_workspace_bytes(), "pn_mlp_forward_tc_x3: workspace too small");
  PN_CHECK_ARG(((uintptr_t)workspace & 255) == 0 && ((uintptr_t)tables32 & 15) == 0 && ((uintptr_t)feat_x & 15) == 0 &&
                   ((uintptr_t)feat_x_lo & 15) == 0 && ((uintptr_t)packed_x3 & 255) == 0,
               "pn_mlp_forward_tc_x3: misaligned buffer");
  cudaStream_t st = (cudaStream_t)stream;
  CUtensorMap mw, mx;
  int rc;
  if ((rc = make_map(&mw, packed_x3, (uint64_t)2 * kLayers * 512, 512, 1024, 128))) return rc;
  if ((rc = make_map(&mx, feat_x, (uint64_t)x_rows, 64, 128, kRows))) return rc;
  CUtensorMap ml;
  if ((rc = make_map(&ml, feat_x_lo, (uint64_t)x_rows, 64, 128, kRows))) return rc;
  unsigned int* err = (unsigned int*)workspace;
  PN_CUDA_OK(cudaMemsetAsync(err, 0, 4, st));
  ChainParams p = {};
  p.out = (float4*)out;
  p.pre = (float4*)pre_act;
  p.err = err;
  p.table_stride = n_texels * 512;
  p.tables32 = tables32;
  p.ppo = points_per_object;
  p.n_views = n_views;
  p.debug_stage = -1;
  p.bias = (const float*)((const char*)packed_x3 + kX3BiasOff);
  p.bias2 = (const float*)((const char*)packed_x3 + kX3BiasOff + kBiasBytes);
  p.n_points = n_pts;
  p.n_rows = n_pts;
  p.group = fused ? 1 : 0;
  p.first_block = 0;
  p.last_block = 5;
  p.park = (float*)((char*)workspace + 256);
  p.drop_a_lo = drop_a_lo;
  return launch_chain(mw, mx, ml, p, st);
}

extern "C" int pn_mlp_forward_tc_x3(const void* packed_x3, const float* tables32, int64_t n_texels, const void* feat_x,
                                    const void* feat_x_lo, int n_objects, int n_views, int64_t points_per_object,
                                    float* out, float* pre_act, void* workspace, size_t workspace_bytes,
                                    pn_stream_t stream) {
  return mlp_forward_tc_pairs(packed_x3, tables32, n_texels, feat_x, feat_x_lo, n_objects, n_views, points_per_object, out,
                              pre_act, workspace, workspace_bytes, 0, stream);
}

// mlp_precision = fp16x2: the precise chain without its third operand pass (a_lo W_hi).  Weight rounding and the fp16
// tables are the fast chain's errors that do not average out along a ray; activation rounding does.
extern "C" int pn_mlp_forward_tc_x2(const void* packed_x3, const float* tables32, int64_t n_texels, const void* feat_x,
                                    const void* feat_x_lo, int n_objects, int n_views, int64_t points_per_object,
                                    float* out, float* pre_act, void* workspace, size_t workspace_bytes,
                                    pn_stream_t stream) {
  return mlp_forward_tc_pairs(packed_x3, tables32, n_texels, feat_x, feat_x_lo, n_objects, n_views, points_per_object, out,
                              pre_act, workspace, workspace_bytes, 1, stream);
}
