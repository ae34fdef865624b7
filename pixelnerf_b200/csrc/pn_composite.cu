// Alpha compositing, forward and hand-written backward
// (reference src/render/nerf.py:178-182, 228-249; backward = autograd of the same).
//
// G lanes per ray (8 for K <= 64, 16 for K <= 128, 32 beyond), samples dealt round-robin: lane l of a
// group owns samples l, l+G, l+2G, ...  Every global access of a warp is then fully coalesced -- one
// rgb-sigma load instruction covers G consecutive float4 of each of the warp's 32/G rays (whole 128-byte
// lines), z / weights / gradients move as runs of G consecutive floats (whole 32-byte sectors).  (The
// earlier layout, lane l owning a contiguous segment of samples, had every load instruction touch 32
// different lines for 16 bytes each: 3 x the L1 wavefronts per byte, which is what bounded it.)
// The exclusive transmittance product is one G-wide multiplicative scan per round of G samples
// chained through a running product; the backward's suffix sums are the mirrored additive scans.
// HBM-bound: per ray the forward reads K*(16+4)+8 bytes and writes K*4+16.
#include "pn_common.cuh"

namespace pn {

// Block size / occupancy bound, from a sweep on B200 (tests/comp_sweep.sh, 2^20 rays; fraction of the measured
// copy bandwidth for K=64 / K=96): forward 4 warps x 6 blocks (72 registers, no spills) 1.00 / 0.93; backward
// 8 warps x 4 blocks (64 registers, 40 B of spills) 0.91 / 0.86 -- occupancy beats spill-free code there:
// (8 warps x 2 blocks, 96 registers) gave 0.71 / 0.63.
#ifndef PN_COMP_FWD_WARPS
#define PN_COMP_FWD_WARPS 4
#endif
#ifndef PN_COMP_BWD_WARPS
#define PN_COMP_BWD_WARPS 8
#endif
#ifndef PN_COMP_FWD_MINB
#define PN_COMP_FWD_MINB 6
#endif
#ifndef PN_COMP_BWD_MINB
#define PN_COMP_BWD_MINB 4
#endif
constexpr int kCompWarpsFwd = PN_COMP_FWD_WARPS;
constexpr int kCompWarpsBwd = PN_COMP_BWD_WARPS;

// inclusive product scan over the G lanes of a group; returns (exclusive product, group total)
template <int G>
__device__ __forceinline__ void group_prod_scan(float v, int gl, float& excl, float& total) {
  float x = v;
#pragma unroll
  for (int d = 1; d < G; d <<= 1) {
    const float y = __shfl_up_sync(0xffffffffu, x, d, G);
    if (gl >= d) x *= y;
  }
  const float e = __shfl_up_sync(0xffffffffu, x, 1, G);
  excl = gl == 0 ? 1.0f : e;
  total = __shfl_sync(0xffffffffu, x, G - 1, G);
}
// sum over the lanes after this one (exclusive suffix) and the group total
template <int G>
__device__ __forceinline__ void group_suffix_sum(float v, int gl, float& excl, float& total) {
  float x = v;
#pragma unroll
  for (int d = 1; d < G; d <<= 1) {
    const float y = __shfl_down_sync(0xffffffffu, x, d, G);
    if (gl + d < G) x += y;
  }
  excl = x - v;
  total = __shfl_sync(0xffffffffu, x, 0, G);
}
template <int G>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int d = G / 2; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d, G);
  return v;
}

// rgb-sigma and z of this lane's samples (sample i*G + gl in round i; z = far and sigma = 0 past the row
// end, which makes such samples transparent with alpha = 0 exactly) and z of each sample's successor:
// the next lane's value of the same round, or lane 0's of the next round; z_K := far (nerf.py:178-182).
template <int G, int CMAX>
__device__ __forceinline__ void load_samples(const float4* __restrict__ crow, const float* __restrict__ zrow, int k,
                                             int rounds, int gl, float far, float4 (&c)[CMAX], float (&zz)[CMAX],
                                             float (&zn)[CMAX]) {
#pragma unroll
  for (int i = 0; i < CMAX; ++i) {
    const int s = i * G + gl;
    if (i < rounds && s < k) {
      c[i] = crow[s];
      zz[i] = zrow[s];
    } else {
      c[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      zz[i] = far;
    }
  }
#pragma unroll
  for (int i = 0; i < CMAX; ++i) {
    const float wrap = i + 1 < CMAX ? zz[i + 1 < CMAX ? i + 1 : i] : far;
    const float give = gl == 0 ? wrap : zz[i];
    zn[i] = __shfl_sync(0xffffffffu, give, (gl + 1) & (G - 1), G);
  }
}

template <int G, int CMAX>
__global__ void __launch_bounds__(kCompWarpsFwd * 32, CMAX <= 8 ? PN_COMP_FWD_MINB : 2)
composite_fwd_kernel(const float4* __restrict__ rgbs, const float* __restrict__ z,
                     const RaySrc rays, float* __restrict__ weights,
                     float* __restrict__ rgb, float* __restrict__ depth, int64_t n_rays, int k,
                     int rounds, bool white) {
  const int lane = threadIdx.x & 31, gl = lane % G;
  const int64_t r_raw = ((int64_t)blockIdx.x * kCompWarpsFwd + (threadIdx.x >> 5)) * (32 / G) + lane / G;
  const bool live = r_raw < n_rays;
  const int64_t r = live ? r_raw : n_rays - 1;   // dead groups recompute the last ray and store nothing
  const float far = ray_far(rays, r);
  float4 c[CMAX];
  float zz[CMAX], zn[CMAX];
  load_samples<G, CMAX>(rgbs + r * k, z + r * k, k, rounds, gl, far, c, zz, zn);
  float run = 1.0f;   // transmittance in front of the current round
  float sr = 0.f, sg = 0.f, sb = 0.f, sd = 0.f, sw = 0.f;
#pragma unroll
  for (int i = 0; i < CMAX; ++i) {
    if (i < rounds) {
      const float al = 1.0f - expf(-(zn[i] - zz[i]) * fmaxf(c[i].w, 0.0f));
      float excl, total;
      group_prod_scan<G>(1.0f - al + 1e-10f, gl, excl, total);
      const float w = al * (run * excl);
      run *= total;
      const int s = i * G + gl;
      if (live && s < k) weights[r * k + s] = w;
      sr += w * c[i].x; sg += w * c[i].y; sb += w * c[i].z;
      sd += w * zz[i];
      sw += w;
    }
  }
  sr = group_sum<G>(sr); sg = group_sum<G>(sg); sb = group_sum<G>(sb); sd = group_sum<G>(sd); sw = group_sum<G>(sw);
  if (live && gl == 0) {
    if (white) { const float bg = 1.0f - sw; sr = sr + bg; sg = sg + bg; sb = sb + bg; }
    rgb[r * 3 + 0] = sr; rgb[r * 3 + 1] = sg; rgb[r * 3 + 2] = sb;
    depth[r] = sd;
  }
}

// Backward.  With s = relu(sigma), E = exp(-delta s), alpha = 1-E, q = 1-alpha+1e-10,
// T_k = prod_{j<k} q_j, w = alpha T:
//   g_w[k]     = gC.c_k + gD z_k + gW_k - [white] sum(gC)
//   g_c[k]     = w_k gC
//   g_alpha[k] = T_k g_w[k] - (1/q_k) sum_{m>k} w_m g_w[m]
//   g_sigma[k] = g_alpha[k] delta_k E_k [sigma_k > 0]
//   g_delta[k] = g_alpha[k] s_k E_k
//   g_z[k]     = g_delta[k-1] - g_delta[k] + w_k gD        (z_K = far carries no grad)
template <int G, int CMAX>
__global__ void __launch_bounds__(kCompWarpsBwd * 32, CMAX <= 8 ? PN_COMP_BWD_MINB : 2)
composite_bwd_kernel(const float4* __restrict__ rgbs, const float* __restrict__ z,
                     const float* __restrict__ rays, const float* __restrict__ g_rgb,
                     const float* __restrict__ g_depth, const float* __restrict__ g_weights,
                     float4* __restrict__ g_rgbs, float* __restrict__ g_z, int64_t n_rays, int k,
                     int rounds, bool white) {
  const int lane = threadIdx.x & 31, gl = lane % G;
  const int64_t r_raw = ((int64_t)blockIdx.x * kCompWarpsBwd + (threadIdx.x >> 5)) * (32 / G) + lane / G;
  const bool live = r_raw < n_rays;
  const int64_t r = live ? r_raw : n_rays - 1;
  const float far = rays[r * 8 + 7];
  const float gr = g_rgb[r * 3 + 0], gg = g_rgb[r * 3 + 1], gb = g_rgb[r * 3 + 2];
  const float gd = g_depth ? g_depth[r] : 0.0f;
  const float gbg = white ? (gr + gg + gb) : 0.0f;
  float4 c[CMAX];
  float zz[CMAX], zn[CMAX], gw[CMAX], E[CMAX], Tk[CMAX];
  load_samples<G, CMAX>(rgbs + r * k, z + r * k, k, rounds, gl, far, c, zz, zn);
#pragma unroll
  for (int i = 0; i < CMAX; ++i) {
    const int s = i * G + gl;
    gw[i] = (g_weights && i < rounds && s < k) ? g_weights[r * k + s] : 0.0f;
  }
  // forward quantities: E, T_k and the full g_w; v = w g_w feeds the suffix sums
  float run = 1.0f;
#pragma unroll
  for (int i = 0; i < CMAX; ++i) {
    if (i < rounds) {
      E[i] = expf(-(zn[i] - zz[i]) * fmaxf(c[i].w, 0.0f));
      float excl, total;
      group_prod_scan<G>(1.0f - (1.0f - E[i]) + 1e-10f, gl, excl, total);
      Tk[i] = run * excl;
      run *= total;
      gw[i] = gr * c[i].x + gg * c[i].y + gb * c[i].z + gd * zz[i] - gbg + gw[i];
    }
  }
  // rounds from last to first: suffix sum of w g_w over the later samples
  float later = 0.0f;
  float gdel[CMAX];
#pragma unroll
  for (int i = CMAX - 1; i >= 0; --i) {
    gdel[i] = 0.0f;
    if (i < rounds) {
      const int s = i * G + gl;
      const float al = 1.0f - E[i];
      const float w = al * Tk[i];
      float excl, total;
      group_suffix_sum<G>(s < k ? w * gw[i] : 0.0f, gl, excl, total);
      const float q = 1.0f - al + 1e-10f;
      const float ga = Tk[i] * gw[i] - (later + excl) / q;
      later += total;
      const float sg = fmaxf(c[i].w, 0.0f);
      gdel[i] = s < k ? ga * sg * E[i] : 0.0f;
      const float gs = c[i].w > 0.0f ? ga * (zn[i] - zz[i]) * E[i] : 0.0f;
      if (live && s < k) g_rgbs[r * k + s] = make_float4(w * gr, w * gg, w * gb, gs);
    }
  }
  if (g_z) {
    // g_delta of each sample's predecessor: the previous lane's value of the same round, or the last
    // lane's of the previous round
#pragma unroll
    for (int i = 0; i < CMAX; ++i) {
      if (i < rounds) {
        const float wrap = i > 0 ? gdel[i > 0 ? i - 1 : 0] : 0.0f;
        const float give = gl == G - 1 ? wrap : gdel[i];
        const float prev = __shfl_sync(0xffffffffu, give, (gl + G - 1) & (G - 1), G);
        const int s = i * G + gl;
        if (live && s < k) g_z[r * k + s] = prev - gdel[i] + (1.0f - E[i]) * Tk[i] * gd;
      }
    }
  }
}

}  // namespace pn

// lanes per ray and rounds capacity for K samples: (8,8) K<=64, (16,8) K<=128, (32,8) K<=256, (32,16) K<=512
#define PN_COMP_DISPATCH(KERNEL, K, ...)                                                        \
  do {                                                                                          \
    if ((K) <= 64) KERNEL<8, 8> __VA_ARGS__;                                                    \
    else if ((K) <= 128) KERNEL<16, 8> __VA_ARGS__;                                             \
    else if ((K) <= 256) KERNEL<32, 8> __VA_ARGS__;                                             \
    else KERNEL<32, 16> __VA_ARGS__;                                                            \
  } while (0)

static int comp_lanes(int k) { return k <= 64 ? 8 : k <= 128 ? 16 : 32; }
static int composite_args_ok(int64_t n_rays, int k) { return n_rays >= 0 && k > 0 && k <= 512; }

int pn::composite_fwd_src(const float* rgb_sigma, const float* z, const pn::RaySrc& rays,
                                float* weights, float* rgb, float* depth, int64_t n_rays,
                                int n_samples, int white_bkgd, pn_stream_t stream) {
  if (n_rays == 0) return PN_OK;  // empty batches carry null data pointers
  PN_CHECK_ARG(rgb_sigma && z && (rays.rays || rays.poses) && weights && rgb && depth, "pn_composite_fwd: null pointer");
  PN_CHECK_ARG(composite_args_ok(n_rays, n_samples), "pn_composite_fwd: need 1 <= K <= 512, got K=%d", n_samples);
  const int g = comp_lanes(n_samples);
  const int per_lane = (n_samples + g - 1) / g;
  const int blocks = pn::ceil_div(n_rays, pn::kCompWarpsFwd * (32 / g));
  PN_COMP_DISPATCH(pn::composite_fwd_kernel, n_samples,
                   <<<blocks, pn::kCompWarpsFwd * 32, 0, (cudaStream_t)stream>>>(
                       reinterpret_cast<const float4*>(rgb_sigma), z, rays, weights, rgb, depth, n_rays, n_samples,
                       per_lane, white_bkgd != 0));
  PN_LAUNCH_OK("composite_fwd_kernel");
  pn::count_launch();
  return PN_OK;
}

extern "C" int pn_composite_fwd(const float* rgb_sigma, const float* z, const float* rays,
                                float* weights, float* rgb, float* depth, int64_t n_rays,
                                int n_samples, int white_bkgd, pn_stream_t stream) {
  if (n_rays == 0) return PN_OK;  // empty batches carry null data pointers
  PN_CHECK_ARG(rays, "pn_composite_fwd: null pointer");
  return pn::composite_fwd_src(rgb_sigma, z, pn::ray_src(rays), weights, rgb, depth, n_rays, n_samples, white_bkgd, stream);
}

extern "C" int pn_composite_bwd(const float* rgb_sigma, const float* z, const float* rays,
                                const float* g_rgb, const float* g_depth, const float* g_weights,
                                float* g_rgb_sigma, float* g_z, int64_t n_rays, int n_samples,
                                int white_bkgd, pn_stream_t stream) {
  if (n_rays == 0) return PN_OK;  // empty batches carry null data pointers
  PN_CHECK_ARG(rgb_sigma && z && rays && g_rgb && g_rgb_sigma, "pn_composite_bwd: null pointer");
  PN_CHECK_ARG(composite_args_ok(n_rays, n_samples), "pn_composite_bwd: need 1 <= K <= 512, got K=%d", n_samples);
  const int g = comp_lanes(n_samples);
  const int per_lane = (n_samples + g - 1) / g;
  const int blocks = pn::ceil_div(n_rays, pn::kCompWarpsBwd * (32 / g));
  PN_COMP_DISPATCH(pn::composite_bwd_kernel, n_samples,
                   <<<blocks, pn::kCompWarpsBwd * 32, 0, (cudaStream_t)stream>>>(
                       reinterpret_cast<const float4*>(rgb_sigma), z, rays, g_rgb, g_depth, g_weights,
                       reinterpret_cast<float4*>(g_rgb_sigma), g_z, n_rays, n_samples, per_lane, white_bkgd != 0));
  PN_LAUNCH_OK("composite_bwd_kernel");
  pn::count_launch();
  return PN_OK;
}
