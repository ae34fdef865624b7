// ResnetFC forward in fp32 on CUDA cores (reference src/model/resnetfc.py:132-184,
// block :53-62, view mean src/util/util.py:461-466, head src/model/models.py:257-266).
//
// This is the fp32 parity path (1e-4 against the reference); the tensor-core chain
// is in pn_mlp_tc.cu.  One generic kernel computes
//     C[M,512] (+)= relu?(A[M,K]) * W[512,K]^T + bias
// with a 128x128x16 block tile, 8x8 register tile per thread and register-staged
// double buffering.  A and W are both K-major (torch Linear layout), so both tiles
// are transposed into shared memory as [k][m] for conflict-free float4 reads.
#include "pn_common.cuh"

namespace pn {

constexpr int BM = 128, BN = 128, BK = 16, TM = 8, TN = 8;
constexpr int kGemmThreads = (BM / TM) * (BN / TN);  // 256

// Up to four consecutive floats; vector load when all four exist and the address is
// 16-byte aligned (lin_in's 42-wide weight rows are not), zero fill past `remain`.
__device__ __forceinline__ float4 load4(const float* p, int remain) {
  if (remain >= 4 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) return *reinterpret_cast<const float4*>(p);
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (remain > 0) v.x = p[0];
  if (remain > 1) v.y = p[1];
  if (remain > 2) v.z = p[2];
  if (remain > 3) v.w = p[3];
  return v;
}

template <bool kReluA, bool kAccum>
__global__ void __launch_bounds__(kGemmThreads)
linear512_kernel(const float* __restrict__ A, int lda, const float* __restrict__ W, int ldw,
                 const float* __restrict__ bias, float* __restrict__ C, int64_t M, int K) {
  __shared__ __align__(16) float As[2][BK][BM + 4];
  __shared__ __align__(16) float Ws[2][BK][BN + 4];
  const int tid = threadIdx.x;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  // global->smem mapping: 128 rows x 16 k = 512 float4; each thread loads 2 per operand
  const int lrow = tid >> 2;        // 0..63 (+64)
  const int lk = (tid & 3) * 4;     // 0,4,8,12
  const int tx = tid & 15, ty = tid >> 4;

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.0f;

  float4 ra[2], rw[2];
  auto load_tiles = [&](int k0) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int row = lrow + h * 64;
      const int64_t gm = m0 + row;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (gm < M) {
        v = load4(A + gm * lda + k0 + lk, K - (k0 + lk));
      }
      if (kReluA) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
      ra[h] = v;
      rw[h] = load4(W + (size_t)(n0 + row) * ldw + k0 + lk, K - (k0 + lk));
    }
  };
  auto store_tiles = [&](int buf) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int row = lrow + h * 64;
      As[buf][lk + 0][row] = ra[h].x; As[buf][lk + 1][row] = ra[h].y;
      As[buf][lk + 2][row] = ra[h].z; As[buf][lk + 3][row] = ra[h].w;
      Ws[buf][lk + 0][row] = rw[h].x; Ws[buf][lk + 1][row] = rw[h].y;
      Ws[buf][lk + 2][row] = rw[h].z; Ws[buf][lk + 3][row] = rw[h].w;
    }
  };

  const int nk = (K + BK - 1) / BK;
  load_tiles(0);
  store_tiles(0);
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) load_tiles((kt + 1) * BK);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[TM], b[TN];
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Ws[buf][kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Ws[buf][kk][64 + tx * 4]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (kt + 1 < nk) store_tiles(buf ^ 1);
    __syncthreads();
  }
  // epilogue: rows ty*4..+3 and 64+ty*4..+3; cols tx*4..+3 and 64+tx*4..+3
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int64_t gm = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (gm >= M) continue;
#pragma unroll
    for (int jh = 0; jh < 2; ++jh) {
      const int gn = n0 + jh * 64 + tx * 4;
      float4 o = make_float4(acc[i][jh * 4 + 0], acc[i][jh * 4 + 1], acc[i][jh * 4 + 2], acc[i][jh * 4 + 3]);
      const float4 bv = *reinterpret_cast<const float4*>(bias + gn);
      o.x += bv.x; o.y += bv.y; o.z += bv.z; o.w += bv.w;
      float* dst = C + gm * PN_D_HIDDEN + gn;
      if (kAccum) {
        const float4 c = *reinterpret_cast<const float4*>(dst);
        o.x += c.x; o.y += c.y; o.z += c.z; o.w += c.w;
      }
      *reinterpret_cast<float4*>(dst) = o;
    }
  }
}

// combine_interleaved(average): h (SB, NS, P, 512) -> (SB, P, 512), torch.mean order
// (sum over views in order, then divide by NS).
__global__ void __launch_bounds__(256)
view_mean_kernel(const float4* __restrict__ h, float4* __restrict__ out, int ns, int64_t ppo4,
                 int64_t total4) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total4) return;
  const int64_t obj = i / ppo4, rem = i - obj * ppo4;
  const float4* src = h + obj * ns * ppo4 + rem;
  float4 s = src[0];
  for (int v = 1; v < ns; ++v) {
    const float4 t = src[(int64_t)v * ppo4];
    s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
  }
  const float d = (float)ns;
  out[i] = make_float4(s.x / d, s.y / d, s.z / d, s.w / d);
}

// lin_out(relu(h)) + head: one warp per row, 4 dot products of length 512.
__global__ void __launch_bounds__(256)
head_kernel(const float* __restrict__ h, const float* __restrict__ w, const float* __restrict__ b,
            float4* __restrict__ out, float4* __restrict__ pre, int64_t rows) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
  for (int i = 0; i < PN_D_HIDDEN / 128; ++i) {
    const int c = i * 128 + lane * 4;
    float4 v = *reinterpret_cast<const float4*>(h + row * PN_D_HIDDEN + c);
    v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f);
    const float4 w0 = __ldg(reinterpret_cast<const float4*>(w + 0 * PN_D_HIDDEN + c));
    const float4 w1 = __ldg(reinterpret_cast<const float4*>(w + 1 * PN_D_HIDDEN + c));
    const float4 w2 = __ldg(reinterpret_cast<const float4*>(w + 2 * PN_D_HIDDEN + c));
    const float4 w3 = __ldg(reinterpret_cast<const float4*>(w + 3 * PN_D_HIDDEN + c));
    s0 += v.x * w0.x + v.y * w0.y + v.z * w0.z + v.w * w0.w;
    s1 += v.x * w1.x + v.y * w1.y + v.z * w1.z + v.w * w1.w;
    s2 += v.x * w2.x + v.y * w2.y + v.z * w2.z + v.w * w2.w;
    s3 += v.x * w3.x + v.y * w3.y + v.z * w3.z + v.w * w3.w;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    s0 += __shfl_xor_sync(0xffffffffu, s0, d); s1 += __shfl_xor_sync(0xffffffffu, s1, d);
    s2 += __shfl_xor_sync(0xffffffffu, s2, d); s3 += __shfl_xor_sync(0xffffffffu, s3, d);
  }
  if (lane == 0) {
    s0 += b[0]; s1 += b[1]; s2 += b[2]; s3 += b[3];
    if (pre) pre[row] = make_float4(s0, s1, s2, s3);
    out[row] = make_float4(1.0f / (1.0f + expf(-s0)), 1.0f / (1.0f + expf(-s1)), 1.0f / (1.0f + expf(-s2)),
                           fmaxf(s3, 0.0f));
  }
}

template <bool kReluA, bool kAccum>
static int launch_linear(const float* A, int lda, const float* W, int ldw, const float* bias, float* C,
                         int64_t M, int K, cudaStream_t st) {
  dim3 grid(ceil_div(M, BM), PN_D_HIDDEN / BN);
  linear512_kernel<kReluA, kAccum><<<grid, kGemmThreads, 0, st>>>(A, lda, W, ldw, bias, C, M, K);
  PN_LAUNCH_OK("linear512_kernel");
  count_launch();
  return PN_OK;
}

}  // namespace pn

extern "C" size_t pn_mlp_fp32_workspace_bytes(int64_t n_rows) {
  return (size_t)(n_rows > 0 ? n_rows : 0) * PN_D_HIDDEN * sizeof(float) * 2 + 256;
}

extern "C" int pn_mlp_forward_fp32(const pn_mlp_weights_t* w, const float* feat_z, const float* feat_x,
                                   int ldx, int n_objects, int n_views, int64_t points_per_object,
                                   float* out, float* pre_act, void* workspace, size_t workspace_bytes,
                                   pn_stream_t stream) {
  using namespace pn;
  PN_CHECK_ARG(w && feat_z && feat_x && out && workspace, "pn_mlp_forward_fp32: null pointer");
  PN_CHECK_ARG(n_objects > 0 && n_views > 0 && points_per_object >= 0, "pn_mlp_forward_fp32: bad sizes");
  PN_CHECK_ARG(ldx >= PN_D_CODE && ldx % 4 == 0, "pn_mlp_forward_fp32: ldx=%d must be >= 42 and a multiple of 4", ldx);
  const int64_t n_rows = (int64_t)n_objects * n_views * points_per_object;
  const int64_t n_pts = (int64_t)n_objects * points_per_object;
  if (n_rows == 0) return PN_OK;
  PN_CHECK_ARG(workspace_bytes >= pn_mlp_fp32_workspace_bytes(n_rows), "pn_mlp_forward_fp32: workspace too small");
  PN_CHECK_ARG(((uintptr_t)workspace & 15) == 0, "pn_mlp_forward_fp32: workspace must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  float* h = (float*)workspace;
  float* t = h + (size_t)n_rows * PN_D_HIDDEN;
  int rc;
  // h = lin_in(x)                                                   resnetfc.py:146-147
  if ((rc = launch_linear<false, false>(feat_x, ldx, w->lin_in_w, PN_D_CODE, w->lin_in_b, h, n_rows, PN_D_CODE, st))) return rc;
  int64_t m = n_rows;
  for (int b = 0; b < PN_N_BLOCKS; ++b) {
    if (b == PN_COMBINE_LAYER && n_views > 1) {                    // resnetfc.py:152-172
      const int64_t ppo4 = points_per_object * (PN_D_HIDDEN / 4), total4 = n_pts * (PN_D_HIDDEN / 4);
      view_mean_kernel<<<ceil_div(total4, 256), 256, 0, st>>>((const float4*)h, (float4*)t, n_views, ppo4, total4);
      PN_LAUNCH_OK("view_mean_kernel");
      count_launch();
      float* tmp = h; h = t; t = tmp;
      m = n_pts;
    }
    if (b < PN_COMBINE_LAYER)                                        // h += lin_z[b](z)   :174-180
      if ((rc = launch_linear<false, true>(feat_z, PN_D_LATENT, w->lin_z_w[b], PN_D_LATENT, w->lin_z_b[b], h, m, PN_D_LATENT, st))) return rc;
    // t = fc_0(relu(h)); h += fc_1(relu(t))                         resnetfc.py:53-62
    if ((rc = launch_linear<true, false>(h, PN_D_HIDDEN, w->fc0_w[b], PN_D_HIDDEN, w->fc0_b[b], t, m, PN_D_HIDDEN, st))) return rc;
    if ((rc = launch_linear<true, true>(t, PN_D_HIDDEN, w->fc1_w[b], PN_D_HIDDEN, w->fc1_b[b], h, m, PN_D_HIDDEN, st))) return rc;
  }
  head_kernel<<<ceil_div(n_pts, 8), 256, 0, st>>>(h, w->lin_out_w, w->lin_out_b, (float4*)out, (float4*)pre_act, n_pts);
  PN_LAUNCH_OK("head_kernel");
  count_launch();
  return PN_OK;
}
