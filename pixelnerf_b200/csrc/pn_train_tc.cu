// Tensor-core backward of the ResnetFC MLP: what torch autograd does for the reference's
// loss.backward() (train/train.py:271-281) over resnetfc.py:132-184, restated as 13 dgrad and 14 wgrad GEMMs on
// tcgen05 (csrc/pn_gemm_tc.cu) around the activations the forward chain kernel saved (pn_mlp_forward_tc_train).
//
// Precision: gradients travel between layers as bf16 (fp32 range, 8-bit mantissa), including the residual-stream
// gradient dL/dh (each block's contribution is added in fp32 registers and the sum rounded once); every GEMM and
// every bias/weight-gradient sum accumulates in fp32.  The saved activations are the bf16 relu outputs that fed
// each forward GEMM; the relu derivative is read off them (zero / non-zero).  An fp32 master copy of dL/dh was
// measured: it costs 4 KB of HBM traffic per row and block (the dgrad GEMMs are HBM-bound) for a gradient that
// is rounded to bf16 for the next GEMM anyway.
#include <cuda_bf16.h>

#include "pn_common.cuh"

extern "C" int pn_gemm_tc_rows(const void* a, int a_fmt, const void* b, int b_fmt, int64_t n_rows, int k, int n,
                               const void* mask_f16, float* acc32, int accumulate, const void* resid_bf16,
                               void* out_bf16, float scale, float* colsum0, float* colsum1,
                               unsigned int* err_word, pn_stream_t stream);
extern "C" int pn_gemm_tc_rows_slabs(const void* a_slabs, int64_t slab_rows, int fmt, const void* b, int64_t n_rows,
                                     int k, int n, float* acc32, int accumulate, float scale, unsigned int* err_word,
                                     pn_stream_t stream);
extern "C" int pn_gemm_tc_wgrad(const void* g, int g_fmt, int ldg, const void* x, int x_fmt, int ldx, int64_t n_rows,
                                int n_store, float* out, int ldo, unsigned int* err_word, pn_stream_t stream);

namespace pn {
namespace tc {

constexpr int kW16Mats = 13;                       // fc_0[0..4], fc_1[0..4], lin_z[0..2]
constexpr size_t kW16Elems = (size_t)512 * 512;

__device__ __forceinline__ uint32_t bf2(float a, float b) {   // a in the low half
  uint32_t d;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(b), "f"(a));
  return d;
}
__device__ __forceinline__ float bf_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf_hi(uint32_t v) { return __uint_as_float(v & 0xFFFF0000u); }

// bf16 copies of the 512x512 matrices exactly as stored ([out, in] row-major): the dgrad GEMM reads them as its
// MN-major B operand, so no transpose is needed.
__global__ void __launch_bounds__(256)
pack_w16_kernel(pn_mlp_weights_t w, uint32_t* __restrict__ dst) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 2;
  if (i >= (int64_t)kW16Mats * kW16Elems) return;
  const int mat = (int)(i / kW16Elems);
  const int64_t e = i - (int64_t)mat * kW16Elems;
  const float* src = mat < 5 ? w.fc0_w[mat] : mat < 10 ? w.fc1_w[mat - 5] : w.lin_z_w[mat - 10];
  const float2 v = *reinterpret_cast<const float2*>(src + e);
  dst[i >> 1] = bf2(v.x, v.y);
}

// Per row: z = bilinear blend of the four latent texels named by the taps record the forward left in bytes
// [96,128) of the row's code chunk (encoder.py:102-108), as bf16 (wgrad operand of lin_z), and the 42 code
// entries as bf16 padded to 64 (wgrad operand of lin_in).  One warp per row, 16-byte loads.
__global__ void __launch_bounds__(256)
blend_taps_kernel(const uint4* __restrict__ feat_x, const float* __restrict__ latent, uint2* __restrict__ z16,
                  uint4* __restrict__ x16, int64_t n_rows) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5), n_warps = (int64_t)gridDim.x * 8;
  for (int64_t row = warp0; row < n_rows; row += n_warps) {
    const uint4 idx = __ldg(feat_x + row * 8 + 6), wt = __ldg(feat_x + row * 8 + 7);
    const float w0 = __uint_as_float(wt.x), w1 = __uint_as_float(wt.y), w2 = __uint_as_float(wt.z), w3 = __uint_as_float(wt.w);
    const float4* t0 = reinterpret_cast<const float4*>(latent + (size_t)idx.x * 512);
    const float4* t1 = reinterpret_cast<const float4*>(latent + (size_t)idx.y * 512);
    const float4* t2 = reinterpret_cast<const float4*>(latent + (size_t)idx.z * 512);
    const float4* t3 = reinterpret_cast<const float4*>(latent + (size_t)idx.w * 512);
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int c = it * 32 + lane;
      const float4 a = __ldg(t0 + c), b = __ldg(t1 + c), d = __ldg(t2 + c), e = __ldg(t3 + c);
      z16[row * 128 + c] = make_uint2(bf2(a.x * w0 + b.x * w1 + d.x * w2 + e.x * w3, a.y * w0 + b.y * w1 + d.y * w2 + e.y * w3),
                                      bf2(a.z * w0 + b.z * w1 + d.z * w2 + e.z * w3, a.w * w0 + b.w * w1 + d.w * w2 + e.w * w3));
    }
    if (lane < 8) {
      uint4 o = make_uint4(0u, 0u, 0u, 0u);
      if (lane < 6) {
        const uint4 h = __ldg(feat_x + row * 8 + lane);
        const uint32_t hw[4] = {h.x, h.y, h.z, h.w};
        uint32_t r[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&hw[i]));
          r[i] = bf2(f.x, f.y);
        }
        o = make_uint4(r[0], r[1], r[2], r[3]);
      }
      x16[row * 8 + lane] = o;
    }
  }
}

// Head + lin_out backward (models.py:257-266, resnetfc.py:182): g_pre = [c(1-c) g_c, [sigma > 0] g_sigma];
// dL/dh5 = [x10 > 0] (g_pre . W_out) as bf16; dW_out += g_pre^T x10, db_out += sum g_pre, where
// x10 = relu(h5) is the saved bf16 operand.  One warp per row, block-level reduction of the weight gradient.
__global__ void __launch_bounds__(256)
head_bwd16_kernel(const uint2* __restrict__ x10, const float4* __restrict__ out, const float4* __restrict__ g_out,
                  const float* __restrict__ w_out, uint2* __restrict__ g_h16,
                  float* __restrict__ dw_out, float* __restrict__ db_out, int64_t rows, int rows_per_block) {
  __shared__ float sdw[4][512];
  __shared__ float sdb[4];
  for (int i = threadIdx.x; i < 4 * 512; i += 256) (&sdw[0][0])[i] = 0.0f;
  if (threadIdx.x < 4) sdb[threadIdx.x] = 0.0f;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  float4 wv[4][4];  // this lane's 16 columns (4 chunks of 4) of the 4 weight rows
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int k = 0; k < 4; ++k) wv[k][i] = __ldg(reinterpret_cast<const float4*>(w_out + k * 512 + i * 128 + lane * 4));
  float dwl[4][16];
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int j = 0; j < 16; ++j) dwl[k][j] = 0.0f;
  float dbl[4] = {0.f, 0.f, 0.f, 0.f};
  // kU rows per warp iteration with every load issued before the first use: one row at a time left 1 KB per warp
  // in flight (8 KB per SM at this register count) and the kernel at a third of the HBM rate
  constexpr int kU = 4;
  const int64_t r_end = r0 + rows_per_block < rows ? r0 + rows_per_block : rows;
  for (int64_t rb = r0 + warp; rb < r_end; rb += 8 * kU) {
    float4 o[kU], go[kU];
    uint2 xv[kU][4];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t r = rb + 8 * u;
      if (r < r_end) {
        o[u] = out[r];
        go[u] = g_out[r];
#pragma unroll
        for (int i = 0; i < 4; ++i) xv[u][i] = __ldg(x10 + (r * 512 + i * 128 + lane * 4) / 4);
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t r = rb + 8 * u;
      if (r >= r_end) break;
      const float gp[4] = {o[u].x * (1.0f - o[u].x) * go[u].x, o[u].y * (1.0f - o[u].y) * go[u].y,
                           o[u].z * (1.0f - o[u].z) * go[u].z, o[u].w > 0.0f ? go[u].w : 0.0f};
#pragma unroll
      for (int k = 0; k < 4; ++k) dbl[k] += gp[k];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int c = i * 128 + lane * 4;
        const float hh[4] = {bf_lo(xv[u][i].x), bf_hi(xv[u][i].x), bf_lo(xv[u][i].y), bf_hi(xv[u][i].y)};
        float gh[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float w0 = (&wv[0][i].x)[e], w1 = (&wv[1][i].x)[e], w2 = (&wv[2][i].x)[e], w3 = (&wv[3][i].x)[e];
          const float g = gp[0] * w0 + gp[1] * w1 + gp[2] * w2 + gp[3] * w3;
          gh[e] = hh[e] > 0.0f ? g : 0.0f;
#pragma unroll
          for (int k = 0; k < 4; ++k) dwl[k][i * 4 + e] += gp[k] * hh[e];
        }
        g_h16[(r * 512 + c) / 4] = make_uint2(bf2(gh[0], gh[1]), bf2(gh[2], gh[3]));
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int j = 0; j < 16; ++j) atomicAdd(&sdw[k][(j >> 2) * 128 + lane * 4 + (j & 3)], dwl[k][j]);
  if (lane == 0)
#pragma unroll
    for (int k = 0; k < 4; ++k) atomicAdd(&sdb[k], dbl[k]);
  __syncthreads();
  for (int i = threadIdx.x; i < 4 * 512; i += 256) atomicAdd(dw_out + i, (&sdw[0][0])[i]);
  if (threadIdx.x < 4) atomicAdd(db_out + threadIdx.x, sdb[threadIdx.x]);
}

// backward of combine_interleaved(average) (util.py:461-466): g (SB,P,512) -> (SB,NS,P,512), each view gets g / NS
__global__ void __launch_bounds__(256)
view_mean_bwd16_kernel(const uint2* __restrict__ g16, uint2* __restrict__ out16, int ns, int64_t ppo4, int64_t total4) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total4) return;
  const int64_t ov = i / ppo4, rem = i - ov * ppo4;
  const uint2 v = __ldg(g16 + (ov / ns) * ppo4 + rem);
  const float d = (float)ns;
  out16[i] = make_uint2(bf2(bf_lo(v.x) / d, bf_hi(v.x) / d), bf2(bf_lo(v.y) / d, bf_hi(v.y) / d));
}

// db[512] += column sums of the bf16 matrix G[M, 512]; thread t owns columns 2t, 2t+1
__global__ void __launch_bounds__(256)
colsum16_kernel(const uint32_t* __restrict__ G, float* __restrict__ db, int64_t M, int64_t rows_per_block) {
  const int64_t m0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t m1 = m0 + rows_per_block < M ? m0 + rows_per_block : M;
  float s0 = 0.f, s1 = 0.f;
  for (int64_t m = m0; m < m1; ++m) {
    const uint32_t v = __ldg(G + m * 256 + threadIdx.x);
    s0 += bf_lo(v);
    s1 += bf_hi(v);
  }
  atomicAdd(db + 2 * threadIdx.x, s0);
  atomicAdd(db + 2 * threadIdx.x + 1, s1);
}

}  // namespace tc
}  // namespace pn

extern "C" size_t pn_mlp_tc_w16_bytes(void) { return pn::tc::kW16Mats * pn::tc::kW16Elems * 2; }

extern "C" int pn_prepare_weights_bf16(const pn_mlp_weights_t* w, void* w16, size_t w16_bytes, pn_stream_t stream) {
  using namespace pn::tc;
  PN_CHECK_ARG(w && w16, "pn_prepare_weights_bf16: null pointer");
  PN_CHECK_ARG(w16_bytes >= pn_mlp_tc_w16_bytes() && ((uintptr_t)w16 & 15) == 0,
               "pn_prepare_weights_bf16: buffer too small or misaligned");
  const int64_t pairs = (int64_t)kW16Mats * kW16Elems / 2;
  pack_w16_kernel<<<pn::ceil_div(pairs, 256), 256, 0, (cudaStream_t)stream>>>(*w, (uint32_t*)w16);
  PN_LAUNCH_OK("pack_w16_kernel");
  pn::count_launch();
  return PN_OK;
}

// Workspace: [256 B error word] slab 0 | slab 1 | slab 2 (dL/dh_b, bf16, pitch rows rounded up to 256) | G1 | Z16
// (bf16 rows x 512) | x16 (rows x 64).  The three residual-stream gradients that feed lin_z stay side by side so
// that dL/dz = sum_b dL/dh_b W_z[b] is ONE K = 1536 GEMM over the slabs instead of three accumulating launches
// (each of which re-read and re-wrote the rows x 512 fp32 result).
static inline size_t slab_rows_of(int64_t n_rows) { return (size_t)((n_rows > 0 ? n_rows : 0) + 255) / 256 * 256; }

extern "C" size_t pn_mlp_backward_tc_workspace_bytes(int64_t n_rows, int n_views) {
  const size_t r = (size_t)(n_rows > 0 ? n_rows : 0);
  (void)n_views;
  return 256 + 3 * slab_rows_of(n_rows) * 512 * 2 + 2 * r * 512 * 2 + r * 64 * 2;
}

// dL/dh_0 (rows x 512 bf16, the gradient at lin_in's output) is left in the workspace at this byte offset (slab 0)
extern "C" size_t pn_mlp_backward_tc_gh0_offset(int64_t n_rows, int n_views) {
  (void)n_rows; (void)n_views;
  return 256;
}

extern "C" int pn_mlp_backward_tc(const pn_mlp_weights_t* w, const void* w16, const void* feat_x,
                                  const float* latent_nhwc, int n_objects, int n_views, int64_t points_per_object,
                                  const void* acts, const float* out, const float* g_out,
                                  const pn_mlp_grads_t* grads, float* g_feat_z, void* workspace,
                                  size_t workspace_bytes, pn_stream_t stream) {
  using namespace pn;
  using namespace pn::tc;
  PN_CHECK_ARG(w && w16 && feat_x && latent_nhwc && acts && out && g_out && grads && workspace,
               "pn_mlp_backward_tc: null pointer");
  PN_CHECK_ARG(n_objects > 0 && n_views > 0 && points_per_object >= 0, "pn_mlp_backward_tc: bad sizes");
  const int64_t n_rows = (int64_t)n_objects * n_views * points_per_object;
  const int64_t n_pts = (int64_t)n_objects * points_per_object;
  if (n_rows == 0) return PN_OK;
  PN_CHECK_ARG(workspace_bytes >= pn_mlp_backward_tc_workspace_bytes(n_rows, n_views), "pn_mlp_backward_tc: workspace too small");
  PN_CHECK_ARG((((uintptr_t)workspace | (uintptr_t)acts | (uintptr_t)feat_x | (uintptr_t)w16 | (uintptr_t)g_feat_z) & 31) == 0,
               "pn_mlp_backward_tc: misaligned buffer");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t R = (size_t)n_rows;
  char* ws = (char*)workspace;
  unsigned int* err = (unsigned int*)ws;
  const size_t Rp = slab_rows_of(n_rows);
  uint16_t* slab[3] = {(uint16_t*)(ws + 256), (uint16_t*)(ws + 256) + Rp * 512, (uint16_t*)(ws + 256) + 2 * Rp * 512};
  uint16_t* G1 = slab[2] + Rp * 512;
  uint16_t* Z16 = G1 + R * 512;
  uint16_t* X16 = Z16 + R * 512;
  // dL/dh_5 .. dL/dh_3 (one row per point) are updated in place: in slab 2 when NS = 1 (block 2 then continues in
  // place and leaves dL/dh_2 there), in slab 1 (free until block 1) when the view-mean backward expands them into
  // slab 2 first
  uint16_t* GH16 = n_views > 1 ? slab[1] : slab[2];
  const uint16_t* W16 = (const uint16_t*)w16;
  const size_t R_slot = (R + 127) / 128 * 128;   // slot pitch of the forward's saved operands (pn_mlp_tc_acts_bytes)
  auto X = [&](int slot) { return (const uint16_t*)acts + (size_t)slot * R_slot * 512; };
  auto Wm = [&](int mat) { return W16 + (size_t)mat * kW16Elems; };
  int rc;
  PN_CUDA_OK(cudaMemsetAsync(err, 0, 4, st));
  auto colsum = [&](const uint16_t* G, float* db, int64_t m) {
    const int64_t rpb = 256;
    colsum16_kernel<<<ceil_div(m, rpb), 256, 0, st>>>((const uint32_t*)G, db, m, rpb);
    count_launch();
  };
  auto wgrad = [&](const uint16_t* G, const uint16_t* Xin, int ldx, int n_store, float* dW, int64_t m) {
    return pn_gemm_tc_wgrad(G, 1, 512, Xin, 1, ldx, m, n_store, dW, n_store, err, stream);
  };
  {
    blend_taps_kernel<<<2048, 256, 0, st>>>((const uint4*)feat_x, latent_nhwc, (uint2*)Z16, (uint4*)X16, n_rows);
    PN_LAUNCH_OK("blend_taps_kernel");
    const int rpb = 1024;
    head_bwd16_kernel<<<ceil_div(n_pts, rpb), 256, 0, st>>>((const uint2*)X(10), (const float4*)out, (const float4*)g_out,
                                                            w->lin_out_w, (uint2*)GH16, grads->lin_out_w,
                                                            grads->lin_out_b, n_pts, rpb);
    PN_LAUNCH_OK("head_bwd16_kernel");
    count_launch(2);
  }
  int64_t m = n_pts;
  // Bias gradients are column sums of the gradient tensors: of dL/dh_5 by its own kernel, of everything a dgrad
  // GEMM produces by that GEMM's epilogue (dL/dh_b feeds fc_1.bias of block b-1, lin_z[b].bias and, for b = 0,
  // lin_in.bias; summed over the per-point rows before the view-mean backward, which is the same number).
  colsum(GH16, grads->fc1_b[PN_N_BLOCKS - 1], m);
  for (int b = PN_N_BLOCKS - 1; b >= 0; --b) {
    if (b == PN_COMBINE_LAYER - 1 && n_views > 1) {   // backward of the view mean (resnetfc.py:152-172)
      const int64_t ppo4 = points_per_object * 128, total4 = n_rows * 128;
      view_mean_bwd16_kernel<<<ceil_div(total4, 256), 256, 0, st>>>((const uint2*)GH16, (uint2*)slab[2], n_views, ppo4, total4);
      PN_LAUNCH_OK("view_mean_bwd16_kernel");
      count_launch();
      GH16 = slab[2];
      m = n_rows;
    }
    // h_{b+1} = h_b + fc_1(x_{2b+1}) + b1,  x_{2b+1} = relu(fc_0(x_{2b}) + b0),  x_{2b} = relu(h_b)
    if ((rc = wgrad(GH16, X(2 * b + 1), 512, 512, grads->fc1_w[b], m))) return rc;
    if ((rc = pn_gemm_tc_rows(GH16, 1, Wm(5 + b), 1, m, 512, 512, X(2 * b + 1), nullptr, 0, nullptr, G1, 1.0f,
                              grads->fc0_b[b], nullptr, err, stream)))
      return rc;
    if ((rc = wgrad(G1, X(2 * b), 512, 512, grads->fc0_w[b], m))) return rc;
    float* cs0 = b > 0 ? grads->fc1_b[b - 1] : grads->lin_in_b;
    float* cs1 = b < PN_COMBINE_LAYER ? grads->lin_z_b[b] : nullptr;
    // dL/dh_b = dL/dh_{b+1} + ...: in place, except that blocks 1 and 0 leave their result in their own slab
    uint16_t* gh_out = b < PN_COMBINE_LAYER - 1 ? slab[b] : GH16;
    if ((rc = pn_gemm_tc_rows(G1, 1, Wm(b), 1, m, 512, 512, X(2 * b), nullptr, 0, GH16, gh_out, 1.0f, cs0, cs1, err, stream)))
      return rc;
    GH16 = gh_out;
    if (b < PN_COMBINE_LAYER) {   // h_b = ... + lin_z[b](z)  (resnetfc.py:174-180)
      if ((rc = wgrad(GH16, Z16, 512, 512, grads->lin_z_w[b], m))) return rc;
    }
  }
  // dL/dz = [dL/dh_0 | dL/dh_1 | dL/dh_2] x [W_z0; W_z1; W_z2] (the three lin_z matrices are consecutive in w16)
  if (g_feat_z && (rc = pn_gemm_tc_rows_slabs(slab[0], (int64_t)Rp, 1, Wm(10), m, 3 * 512, 512, g_feat_z, 0, 1.0f, err,
                                             stream)))
    return rc;
  if ((rc = wgrad(GH16, X16, 64, PN_D_CODE, grads->lin_in_w, m))) return rc;
  PN_LAUNCH_OK("pn_mlp_backward_tc");
  return PN_OK;
}
