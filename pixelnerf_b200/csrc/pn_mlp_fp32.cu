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

// C[M,512] = [resid +] [mask > 0 ?] (act(A)[M,K] * W[512,K]^T [+ bias]).  resid may alias C (in-place
// accumulate); mask zeroes the product where mask[m,n] <= 0 (relu backward).
template <bool kReluA, bool kAccum>
__global__ void __launch_bounds__(kGemmThreads)
linear512_kernel(const float* __restrict__ A, int lda, const float* __restrict__ W, int ldw,
                 const float* __restrict__ bias, float* C, int64_t M, int K,
                 const float* resid = nullptr, const float* __restrict__ mask = nullptr) {
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
      if (bias) {
        const float4 bv = *reinterpret_cast<const float4*>(bias + gn);
        o.x += bv.x; o.y += bv.y; o.z += bv.z; o.w += bv.w;
      }
      if (mask) {
        const float4 mv = *reinterpret_cast<const float4*>(mask + gm * PN_D_HIDDEN + gn);
        o.x = mv.x > 0.f ? o.x : 0.f; o.y = mv.y > 0.f ? o.y : 0.f;
        o.z = mv.z > 0.f ? o.z : 0.f; o.w = mv.w > 0.f ? o.w : 0.f;
      }
      float* dst = C + gm * PN_D_HIDDEN + gn;
      if (kAccum) {
        const float4 c = *reinterpret_cast<const float4*>((resid ? resid : C) + gm * PN_D_HIDDEN + gn);
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
int launch_linear(const float* A, int lda, const float* W, int ldw, const float* bias, float* C,
                  int64_t M, int K, cudaStream_t st, const float* resid = nullptr, const float* mask = nullptr) {
  dim3 grid(ceil_div(M, BM), PN_D_HIDDEN / BN);
  linear512_kernel<kReluA, kAccum><<<grid, kGemmThreads, 0, st>>>(A, lda, W, ldw, bias, C, M, K, resid, mask);
  PN_LAUNCH_OK("linear512_kernel");
  count_launch();
  return PN_OK;
}


// ---- backward kernels (training; reference: autograd of resnetfc.py:132-184) -----------------

// dW[512, Ki] += G[M,512]^T * act(A)[M, Ki]   (split over row ranges, fp32 atomics)
// Both operands are row-major with the reduction index m slowest, so tiles load coalesced with
// no transpose: Gs[m][o], As[m][i].
constexpr int WK = 16;
template <bool kReluA>
__global__ void __launch_bounds__(kGemmThreads)
wgrad512_kernel(const float* __restrict__ G, const float* __restrict__ A, int lda, int ki,
                float* __restrict__ dW, int ldw, int64_t M, int64_t rows_per_split) {
  __shared__ __align__(16) float Gs[2][WK][BM + 4];
  __shared__ __align__(16) float As[2][WK][BN + 4];
  const int tid = threadIdx.x;
  const int o0 = blockIdx.x * BM, i0 = blockIdx.y * BN;
  const int64_t m_begin = (int64_t)blockIdx.z * rows_per_split;
  const int64_t m_end = m_begin + rows_per_split < M ? m_begin + rows_per_split : M;
  if (m_begin >= m_end) return;
  const int lm = tid >> 5;          // 0..7 (+8): row of the 16-row slab
  const int lc = (tid & 31) * 4;    // column within the 128-wide tile
  const int tx = tid & 15, ty = tid >> 4;
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.0f;
  float4 rg[2], ra[2];
  auto load_tiles = [&](int64_t m0) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int64_t m = m0 + lm + h * 8;
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f), a = g;
      if (m < m_end) {
        g = *reinterpret_cast<const float4*>(G + m * PN_D_HIDDEN + o0 + lc);
        a = load4(A + m * lda + i0 + lc, ki - (i0 + lc));
        if (kReluA) { a.x = fmaxf(a.x, 0.f); a.y = fmaxf(a.y, 0.f); a.z = fmaxf(a.z, 0.f); a.w = fmaxf(a.w, 0.f); }
      }
      rg[h] = g; ra[h] = a;
    }
  };
  auto store_tiles = [&](int buf) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      *reinterpret_cast<float4*>(&Gs[buf][lm + h * 8][lc]) = rg[h];
      *reinterpret_cast<float4*>(&As[buf][lm + h * 8][lc]) = ra[h];
    }
  };
  const int nk = (int)((m_end - m_begin + WK - 1) / WK);
  load_tiles(m_begin);
  store_tiles(0);
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) load_tiles(m_begin + (int64_t)(kt + 1) * WK);
#pragma unroll
    for (int kk = 0; kk < WK; ++kk) {
      float a[TM], b[TN];
      const float4 a0 = *reinterpret_cast<const float4*>(&Gs[buf][kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&Gs[buf][kk][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&As[buf][kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + tx * 4]);
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
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int o = o0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int c = i0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
      if (c < ki) atomicAdd(dW + (size_t)o * ldw + c, acc[i][j]);
    }
  }
}

// db[512] += column sums of G[M,512]
__global__ void __launch_bounds__(256)
colsum512_kernel(const float* __restrict__ G, float* __restrict__ db, int64_t M, int64_t rows_per_block) {
  const int64_t m0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t m1 = m0 + rows_per_block < M ? m0 + rows_per_block : M;
  float s0 = 0.f, s1 = 0.f;
  for (int64_t m = m0; m < m1; ++m) {
    s0 += G[m * PN_D_HIDDEN + threadIdx.x];
    s1 += G[m * PN_D_HIDDEN + 256 + threadIdx.x];
  }
  atomicAdd(db + threadIdx.x, s0);
  atomicAdd(db + 256 + threadIdx.x, s1);
}

// Head + lin_out backward.  out = [sigmoid(pre_rgb), relu(pre_sigma)]; g_out = dL/d out (P,4).
//   g_pre = [c(1-c) g_c, [pre_sigma > 0] g_sigma]
//   g_h[p,:] = [h5 > 0] * (g_pre[p] . W_out)          dW_out += g_pre^T relu(h5), db_out += sum g_pre
// One warp per row; weight-gradient partials are reduced per block in shared memory first.
__global__ void __launch_bounds__(256)
head_bwd_kernel(const float* __restrict__ h5, const float4* __restrict__ out, const float4* __restrict__ g_out,
                const float* __restrict__ w_out, float* __restrict__ g_h, float* __restrict__ dw_out,
                float* __restrict__ db_out, int64_t rows, int rows_per_block) {
  __shared__ float sdw[4][PN_D_HIDDEN];
  __shared__ float sdb[4];
  for (int i = threadIdx.x; i < 4 * PN_D_HIDDEN; i += 256) (&sdw[0][0])[i] = 0.0f;
  if (threadIdx.x < 4) sdb[threadIdx.x] = 0.0f;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  float4 wv[4][4];  // this lane's 16 columns (4 chunks of 4) of the 4 weight rows
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int k = 0; k < 4; ++k) wv[k][i] = __ldg(reinterpret_cast<const float4*>(w_out + k * PN_D_HIDDEN + i * 128 + lane * 4));
  float dwl[4][16];
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int j = 0; j < 16; ++j) dwl[k][j] = 0.0f;
  float dbl[4] = {0.f, 0.f, 0.f, 0.f};
  for (int64_t r = r0 + warp; r < r0 + rows_per_block && r < rows; r += 8) {
    const float4 o = out[r], go = g_out[r];
    const float gp[4] = {o.x * (1.0f - o.x) * go.x, o.y * (1.0f - o.y) * go.y, o.z * (1.0f - o.z) * go.z,
                         o.w > 0.0f ? go.w : 0.0f};
#pragma unroll
    for (int k = 0; k < 4; ++k) dbl[k] += gp[k];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = i * 128 + lane * 4;
      const float4 hv = *reinterpret_cast<const float4*>(h5 + r * PN_D_HIDDEN + c);
      const float hh[4] = {hv.x, hv.y, hv.z, hv.w};
      float gh[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float w0 = (&wv[0][i].x)[e], w1 = (&wv[1][i].x)[e], w2 = (&wv[2][i].x)[e], w3 = (&wv[3][i].x)[e];
        const float g = gp[0] * w0 + gp[1] * w1 + gp[2] * w2 + gp[3] * w3;
        const bool on = hh[e] > 0.0f;
        gh[e] = on ? g : 0.0f;
        const float a = on ? hh[e] : 0.0f;
#pragma unroll
        for (int k = 0; k < 4; ++k) dwl[k][i * 4 + e] += gp[k] * a;
      }
      *reinterpret_cast<float4*>(g_h + r * PN_D_HIDDEN + c) = make_float4(gh[0], gh[1], gh[2], gh[3]);
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
  for (int i = threadIdx.x; i < 4 * PN_D_HIDDEN; i += 256) atomicAdd(dw_out + i, (&sdw[0][0])[i]);
  if (threadIdx.x < 4) atomicAdd(db_out + threadIdx.x, sdb[threadIdx.x]);
}

// backward of combine_interleaved(average): g (SB,P,512) -> (SB,NS,P,512), each view gets g / NS
__global__ void __launch_bounds__(256)
view_mean_bwd_kernel(const float4* __restrict__ g, float4* __restrict__ out, int ns, int64_t ppo4, int64_t total4) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total4) return;
  const int64_t ov = i / ppo4, rem = i - ov * ppo4;
  const float4 v = g[(ov / ns) * ppo4 + rem];
  const float d = (float)ns;
  out[i] = make_float4(v.x / d, v.y / d, v.z / d, v.w / d);
}

template <bool kReluA>
int launch_wgrad(const float* G, const float* A, int lda, int ki, float* dW, int ldw, int64_t M, cudaStream_t st) {
  int splits = (int)((M + 4095) / 4096);
  if (splits > 64) splits = 64;
  if (splits < 1) splits = 1;
  int64_t rps = (M + splits - 1) / splits;
  rps = (rps + WK - 1) / WK * WK;
  dim3 grid(PN_D_HIDDEN / BM, ceil_div(ki, BN), splits);
  wgrad512_kernel<kReluA><<<grid, kGemmThreads, 0, st>>>(G, A, lda, ki, dW, ldw, M, rps);
  PN_LAUNCH_OK("wgrad512_kernel");
  count_launch();
  return PN_OK;
}
static int launch_colsum(const float* G, float* db, int64_t M, cudaStream_t st) {
  const int64_t rpb = 256;
  colsum512_kernel<<<ceil_div(M, rpb), 256, 0, st>>>(G, db, M, rpb);
  PN_LAUNCH_OK("colsum512_kernel");
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

// ---- per-texel lin_z tables for the tensor-core chain (include/pixelnerf_b200.h) --------------
namespace pn {
__global__ void __launch_bounds__(256)
table_finish_kernel(const float4* __restrict__ src, const float* __restrict__ c0, const float* __restrict__ c1,
                    uint2* __restrict__ dst, int64_t total4) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total4) return;
  const int n = (int)(i % (PN_D_HIDDEN / 4)) * 4;
  float4 v = src[i];
  v.x += c0[n] + c1[n]; v.y += c0[n + 1] + c1[n + 1]; v.z += c0[n + 2] + c1[n + 2]; v.w += c0[n + 3] + c1[n + 3];
  const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
  dst[i] = make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
}
}  // namespace pn

extern "C" size_t pn_latent_tables_bytes(int64_t n_texels) {
  return (size_t)(n_texels > 0 ? n_texels : 0) * PN_COMBINE_LAYER * PN_D_HIDDEN * sizeof(__half);
}

extern "C" int pn_prepare_latent_tables(const pn_mlp_weights_t* w, const float* latent_nhwc, int64_t n_texels,
                                        void* tables, void* workspace, size_t workspace_bytes, pn_stream_t stream) {
  using namespace pn;
  PN_CHECK_ARG(w && latent_nhwc && tables && workspace, "pn_prepare_latent_tables: null pointer");
  PN_CHECK_ARG(n_texels > 0, "pn_prepare_latent_tables: no texels");
  PN_CHECK_ARG(workspace_bytes >= (size_t)n_texels * PN_D_HIDDEN * sizeof(float), "pn_prepare_latent_tables: workspace too small");
  PN_CHECK_ARG(((uintptr_t)tables & 15) == 0 && ((uintptr_t)workspace & 15) == 0, "pn_prepare_latent_tables: misaligned buffer");
  cudaStream_t st = (cudaStream_t)stream;
  float* tmp = (float*)workspace;
  const int64_t total4 = n_texels * (PN_D_HIDDEN / 4);
  int rc;
  for (int b = 0; b < PN_COMBINE_LAYER; ++b) {
    // fp32 lin_z[b] of every texel (no bias yet), then + lin_z bias + the constant folded into this stage
    if ((rc = launch_linear<false, false>(latent_nhwc, PN_D_LATENT, w->lin_z_w[b], PN_D_LATENT, nullptr, tmp, n_texels,
                                          PN_D_LATENT, st))) return rc;
    const float* folded = b == 0 ? w->lin_in_b : w->fc1_b[b - 1];
    table_finish_kernel<<<ceil_div(total4, 256), 256, 0, st>>>(
        (const float4*)tmp, w->lin_z_b[b], folded, (uint2*)tables + (size_t)b * total4, total4);
    PN_LAUNCH_OK("table_finish_kernel");
    count_launch();
  }
  return PN_OK;
}

// The same tables kept in fp32 (2 KB per texel and layer) for the precise tensor-core chain (mlp_precision = fp16x3),
// where nothing but the 512-wide GEMM operands may be rounded.
namespace pn {
__global__ void __launch_bounds__(256)
table_bias_kernel(float4* __restrict__ t, const float* __restrict__ c0, const float* __restrict__ c1, int64_t total4) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total4) return;
  const int n = (int)(i % (PN_D_HIDDEN / 4)) * 4;
  float4 v = t[i];
  v.x += c0[n] + c1[n]; v.y += c0[n + 1] + c1[n + 1]; v.z += c0[n + 2] + c1[n + 2]; v.w += c0[n + 3] + c1[n + 3];
  t[i] = v;
}
}  // namespace pn

extern "C" int pn_prepare_latent_tables_f32(const pn_mlp_weights_t* w, const float* latent_nhwc, int64_t n_texels,
                                            float* tables32, pn_stream_t stream) {
  using namespace pn;
  PN_CHECK_ARG(w && latent_nhwc && tables32, "pn_prepare_latent_tables_f32: null pointer");
  PN_CHECK_ARG(n_texels > 0, "pn_prepare_latent_tables_f32: no texels");
  PN_CHECK_ARG(((uintptr_t)tables32 & 15) == 0, "pn_prepare_latent_tables_f32: misaligned buffer");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t total4 = n_texels * (PN_D_HIDDEN / 4);
  int rc;
  for (int b = 0; b < PN_COMBINE_LAYER; ++b) {
    float* dst = tables32 + (size_t)b * n_texels * PN_D_HIDDEN;
    if ((rc = launch_linear<false, false>(latent_nhwc, PN_D_LATENT, w->lin_z_w[b], PN_D_LATENT, nullptr, dst, n_texels,
                                          PN_D_LATENT, st))) return rc;
    // the same constants are folded in as in the fp16 tables (lin_in.bias into table 0, fc_1 biases of blocks 0-1 into
    // tables 1-2), so the chain's bias tables serve both precisions
    const float* folded = b == 0 ? w->lin_in_b : w->fc1_b[b - 1];
    table_bias_kernel<<<ceil_div(total4, 256), 256, 0, st>>>((float4*)dst, w->lin_z_b[b], folded, total4);
    PN_LAUNCH_OK("table_bias_kernel");
    count_launch();
  }
  return PN_OK;
}

// Training mode rebuilds the tables every step (the lin_z weights change): the same product on the tensor cores.
// latent16 = the fp16 channels-last latent; lin_z[b].weight enters as an fp16 pair hi + lo * 2^-10 (two GEMMs into
// the fp32 workspace), so the one coherent error source of DESIGN 3.1 -- weight rounding -- stays out of the tables;
// the latent's own fp16 rounding is independent per texel.
extern "C" int pn_gemm_tc_rows(const void* a, int a_fmt, const void* b, int b_fmt, int64_t n_rows, int k, int n,
                               const void* mask_f16, float* acc32, int accumulate, const void* resid_bf16,
                               void* out_bf16, float scale, float* colsum0, float* colsum1,
                               unsigned int* err_word, pn_stream_t stream);
namespace pn {
// dst[(2 b + part)][in][out] = fp16 hi (part 0) / (W - hi) * 2^10 (part 1) of lin_z[b].weight[out][in]
__global__ void __launch_bounds__(256)
wz_transpose16_kernel(pn_mlp_weights_t w, __half* __restrict__ dst) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const float* src = w.lin_z_w[b];
  const int o0 = blockIdx.y * 32, i0 = blockIdx.x * 32;
  for (int r = threadIdx.y; r < 32; r += 8) tile[r][threadIdx.x] = src[(size_t)(o0 + r) * 512 + i0 + threadIdx.x];
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += 8) {
    const float v = tile[threadIdx.x][r];           // W[o0 + tx][i0 + r]
    const __half hi = __float2half_rn(v);
    const size_t at = (size_t)(i0 + r) * 512 + o0 + threadIdx.x;
    dst[(size_t)(2 * b) * 512 * 512 + at] = hi;
    dst[(size_t)(2 * b + 1) * 512 * 512 + at] = __float2half_rn((v - __half2float(hi)) * 1024.0f);
  }
}
}  // namespace pn

extern "C" size_t pn_latent_tables_tc_workspace_bytes(int64_t n_texels) {
  return (size_t)(n_texels > 0 ? n_texels : 0) * PN_D_HIDDEN * sizeof(float) + (size_t)6 * 512 * 512 * sizeof(__half) + 256;
}

extern "C" int pn_prepare_latent_tables_tc(const pn_mlp_weights_t* w, const void* latent_nhwc_f16, int64_t n_texels,
                                           void* tables, void* workspace, size_t workspace_bytes,
                                           pn_stream_t stream) {
  using namespace pn;
  PN_CHECK_ARG(w && latent_nhwc_f16 && tables && workspace, "pn_prepare_latent_tables_tc: null pointer");
  PN_CHECK_ARG(n_texels > 0, "pn_prepare_latent_tables_tc: no texels");
  PN_CHECK_ARG(workspace_bytes >= pn_latent_tables_tc_workspace_bytes(n_texels), "pn_prepare_latent_tables_tc: workspace too small");
  PN_CHECK_ARG(((uintptr_t)tables & 15) == 0 && ((uintptr_t)workspace & 255) == 0 && ((uintptr_t)latent_nhwc_f16 & 15) == 0,
               "pn_prepare_latent_tables_tc: misaligned buffer");
  cudaStream_t st = (cudaStream_t)stream;
  __half* wt = (__half*)workspace;
  float* tmp = (float*)((char*)workspace + (size_t)6 * 512 * 512 * sizeof(__half));
  wz_transpose16_kernel<<<dim3(16, 16, PN_COMBINE_LAYER), dim3(32, 8), 0, st>>>(*w, wt);
  PN_LAUNCH_OK("wz_transpose16_kernel");
  count_launch();
  const int64_t total4 = n_texels * (PN_D_HIDDEN / 4);
  int rc;
  for (int b = 0; b < PN_COMBINE_LAYER; ++b) {
    const __half* hi = wt + (size_t)(2 * b) * 512 * 512;
    if ((rc = pn_gemm_tc_rows(latent_nhwc_f16, 0, hi, 0, n_texels, 512, 512, nullptr, tmp, 0, nullptr, nullptr, 1.0f,
                              nullptr, nullptr, nullptr, stream))) return rc;
    if ((rc = pn_gemm_tc_rows(latent_nhwc_f16, 0, hi + 512 * 512, 0, n_texels, 512, 512, nullptr, tmp, 1, nullptr, nullptr,
                              1.0f / 1024.0f, nullptr, nullptr, nullptr, stream))) return rc;
    const float* folded = b == 0 ? w->lin_in_b : w->fc1_b[b - 1];
    table_finish_kernel<<<ceil_div(total4, 256), 256, 0, st>>>(
        (const float4*)tmp, w->lin_z_b[b], folded, (uint2*)tables + (size_t)b * total4, total4);
    PN_LAUNCH_OK("table_finish_kernel");
    count_launch();
  }
  return PN_OK;
}

// ---- training: forward that keeps every activation, and the hand-written backward ------------
// Activation slots (each n_rows x 512 fp32): slot 2b = H_b (input of block b: residual stream after
// lin_z[b] / the view mean), slot 2b+1 = T_b (fc_0 output, pre-relu), slot 10 = H_5 (input of lin_out).
static inline float* act_slot(void* acts, int64_t n_rows, int i) {
  return (float*)acts + (size_t)i * n_rows * PN_D_HIDDEN;
}

extern "C" size_t pn_mlp_train_acts_bytes(int64_t n_rows) {
  return (size_t)(n_rows > 0 ? n_rows : 0) * PN_D_HIDDEN * sizeof(float) * 11 + 256;
}

extern "C" int pn_mlp_forward_fp32_train(const pn_mlp_weights_t* w, const float* feat_z, const float* feat_x,
                                         int ldx, int n_objects, int n_views, int64_t points_per_object,
                                         float* out, float* pre_act, void* acts, size_t acts_bytes,
                                         pn_stream_t stream) {
  using namespace pn;
  PN_CHECK_ARG(w && feat_z && feat_x && out && acts, "pn_mlp_forward_fp32_train: null pointer");
  PN_CHECK_ARG(n_objects > 0 && n_views > 0 && points_per_object >= 0, "pn_mlp_forward_fp32_train: bad sizes");
  PN_CHECK_ARG(ldx >= PN_D_CODE && ldx % 4 == 0, "pn_mlp_forward_fp32_train: ldx=%d", ldx);
  const int64_t n_rows = (int64_t)n_objects * n_views * points_per_object;
  const int64_t n_pts = (int64_t)n_objects * points_per_object;
  if (n_rows == 0) return PN_OK;
  PN_CHECK_ARG(acts_bytes >= pn_mlp_train_acts_bytes(n_rows), "pn_mlp_forward_fp32_train: activation buffer too small");
  PN_CHECK_ARG(((uintptr_t)acts & 15) == 0, "pn_mlp_forward_fp32_train: activation buffer must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  auto S = [&](int i) { return act_slot(acts, n_rows, i); };
  if ((rc = launch_linear<false, false>(feat_x, ldx, w->lin_in_w, PN_D_CODE, w->lin_in_b, S(0), n_rows, PN_D_CODE, st))) return rc;
  if ((rc = launch_linear<false, true>(feat_z, PN_D_LATENT, w->lin_z_w[0], PN_D_LATENT, w->lin_z_b[0], S(0), n_rows, PN_D_LATENT, st))) return rc;
  int64_t m = n_rows;
  for (int b = 0; b < PN_N_BLOCKS; ++b) {
    if ((rc = launch_linear<true, false>(S(2 * b), PN_D_HIDDEN, w->fc0_w[b], PN_D_HIDDEN, w->fc0_b[b], S(2 * b + 1), m, PN_D_HIDDEN, st))) return rc;
    const bool mean_next = (b + 1 == PN_COMBINE_LAYER) && n_views > 1;
    float* target = (b + 1 < PN_N_BLOCKS && !mean_next) ? S(2 * b + 2) : S(10);
    if ((rc = launch_linear<true, true>(S(2 * b + 1), PN_D_HIDDEN, w->fc1_w[b], PN_D_HIDDEN, w->fc1_b[b], target, m, PN_D_HIDDEN, st, S(2 * b)))) return rc;
    if (mean_next) {  // resnetfc.py:152-172
      const int64_t ppo4 = points_per_object * (PN_D_HIDDEN / 4), total4 = n_pts * (PN_D_HIDDEN / 4);
      view_mean_kernel<<<ceil_div(total4, 256), 256, 0, st>>>((const float4*)S(10), (float4*)S(2 * b + 2), n_views, ppo4, total4);
      PN_LAUNCH_OK("view_mean_kernel");
      count_launch();
      m = n_pts;
    } else if (b + 1 < PN_COMBINE_LAYER) {
      if ((rc = launch_linear<false, true>(feat_z, PN_D_LATENT, w->lin_z_w[b + 1], PN_D_LATENT, w->lin_z_b[b + 1], target, m, PN_D_LATENT, st))) return rc;
    }
  }
  head_kernel<<<ceil_div(n_pts, 8), 256, 0, st>>>(S(10), w->lin_out_w, w->lin_out_b, (float4*)out, (float4*)pre_act, n_pts);
  PN_LAUNCH_OK("head_kernel");
  count_launch();
  return PN_OK;
}

extern "C" size_t pn_mlp_backward_workspace_bytes(int64_t n_rows) {
  return (size_t)(n_rows > 0 ? n_rows : 0) * PN_D_HIDDEN * sizeof(float) * 2 + 256;
}

// dL/dh_0 (rows x 512 fp32, the gradient at lin_in's output) is left in the workspace at this byte offset
extern "C" size_t pn_mlp_backward_gh0_offset(int64_t n_rows, int n_views) {
  return n_views > 1 ? (size_t)(n_rows > 0 ? n_rows : 0) * PN_D_HIDDEN * sizeof(float) : 0;
}

extern "C" int pn_mlp_backward_fp32(const pn_mlp_weights_t* w, const pn_mlp_weights_t* wt, const float* feat_z,
                                    const float* feat_x, int ldx, int n_objects, int n_views,
                                    int64_t points_per_object, const void* acts, const float* out,
                                    const float* g_out, const pn_mlp_grads_t* grads, float* g_feat_z,
                                    void* workspace, size_t workspace_bytes, pn_stream_t stream) {
  using namespace pn;
  PN_CHECK_ARG(w && wt && feat_z && feat_x && acts && out && g_out && grads && workspace, "pn_mlp_backward_fp32: null pointer");
  PN_CHECK_ARG(n_objects > 0 && n_views > 0 && points_per_object >= 0, "pn_mlp_backward_fp32: bad sizes");
  const int64_t n_rows = (int64_t)n_objects * n_views * points_per_object;
  const int64_t n_pts = (int64_t)n_objects * points_per_object;
  if (n_rows == 0) return PN_OK;
  PN_CHECK_ARG(workspace_bytes >= pn_mlp_backward_workspace_bytes(n_rows), "pn_mlp_backward_fp32: workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  auto S = [&](int i) { return (const float*)act_slot(const_cast<void*>(acts), n_rows, i); };
  float* GA = (float*)workspace;
  float* GB = GA + (size_t)n_rows * PN_D_HIDDEN;
  {
    const int rpb = 64;
    head_bwd_kernel<<<ceil_div(n_pts, rpb), 256, 0, st>>>(S(10), (const float4*)out, (const float4*)g_out, w->lin_out_w,
                                                          GA, grads->lin_out_w, grads->lin_out_b, n_pts, rpb);
    PN_LAUNCH_OK("head_bwd_kernel");
    count_launch();
  }
  int64_t m = n_pts;
  for (int b = PN_N_BLOCKS - 1; b >= 0; --b) {
    if (b == PN_COMBINE_LAYER - 1 && n_views > 1) {  // backward of the view mean: every view gets g / NS
      const int64_t ppo4 = points_per_object * (PN_D_HIDDEN / 4), total4 = n_rows * (PN_D_HIDDEN / 4);
      view_mean_bwd_kernel<<<ceil_div(total4, 256), 256, 0, st>>>((const float4*)GA, (float4*)GB, n_views, ppo4, total4);
      PN_LAUNCH_OK("view_mean_bwd_kernel");
      count_launch();
      float* tmp = GA; GA = GB; GB = tmp;
      m = n_rows;
    }
    // h_out = H_b + fc_1(relu(T_b)),  T_b = fc_0(relu(H_b))
    if ((rc = launch_wgrad<true>(GA, S(2 * b + 1), PN_D_HIDDEN, PN_D_HIDDEN, grads->fc1_w[b], PN_D_HIDDEN, m, st))) return rc;
    if ((rc = launch_colsum(GA, grads->fc1_b[b], m, st))) return rc;
    if ((rc = launch_linear<false, false>(GA, PN_D_HIDDEN, wt->fc1_w[b], PN_D_HIDDEN, nullptr, GB, m, PN_D_HIDDEN, st, nullptr, S(2 * b + 1)))) return rc;
    if ((rc = launch_wgrad<true>(GB, S(2 * b), PN_D_HIDDEN, PN_D_HIDDEN, grads->fc0_w[b], PN_D_HIDDEN, m, st))) return rc;
    if ((rc = launch_colsum(GB, grads->fc0_b[b], m, st))) return rc;
    if ((rc = launch_linear<false, true>(GB, PN_D_HIDDEN, wt->fc0_w[b], PN_D_HIDDEN, nullptr, GA, m, PN_D_HIDDEN, st, nullptr, S(2 * b)))) return rc;
    if (b < PN_COMBINE_LAYER) {  // H_b = ... + lin_z[b](z)
      if ((rc = launch_wgrad<false>(GA, feat_z, PN_D_LATENT, PN_D_LATENT, grads->lin_z_w[b], PN_D_LATENT, m, st))) return rc;
      if ((rc = launch_colsum(GA, grads->lin_z_b[b], m, st))) return rc;
      if (g_feat_z) {
        if (b == PN_COMBINE_LAYER - 1) rc = launch_linear<false, false>(GA, PN_D_HIDDEN, wt->lin_z_w[b], PN_D_HIDDEN, nullptr, g_feat_z, m, PN_D_HIDDEN, st);
        else rc = launch_linear<false, true>(GA, PN_D_HIDDEN, wt->lin_z_w[b], PN_D_HIDDEN, nullptr, g_feat_z, m, PN_D_HIDDEN, st);
        if (rc) return rc;
      }
    }
  }
  if ((rc = launch_wgrad<false>(GA, feat_x, ldx, PN_D_CODE, grads->lin_in_w, PN_D_CODE, m, st))) return rc;
  if ((rc = launch_colsum(GA, grads->lin_in_b, m, st))) return rc;
  return PN_OK;
}
