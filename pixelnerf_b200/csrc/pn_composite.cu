// Alpha compositing, forward and hand-written backward
// (reference src/render/nerf.py:178-182, 228-249; backward = autograd of the same).
//
// One warp per ray.  Lane l owns the contiguous samples [l*C, (l+1)*C) with
// C = ceil(K/32), so the exclusive transmittance product is a per-lane serial product
// followed by one warp scan, and every global access is a contiguous run per lane.
// HBM-bound: per ray the forward reads K*(16+4)+8 bytes and writes K*4+16.
#include "pn_common.cuh"

namespace pn {

constexpr int kCompWarps = 8;
constexpr int kMaxPerLane = 8;  // K <= 256

__device__ __forceinline__ float warp_excl_prod(float v, int lane) {
  // inclusive scan then shift
  float x = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    float y = __shfl_up_sync(0xffffffffu, x, d);
    if (lane >= d) x *= y;
  }
  float e = __shfl_up_sync(0xffffffffu, x, 1);
  return lane == 0 ? 1.0f : e;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

__global__ void __launch_bounds__(kCompWarps * 32)
composite_fwd_kernel(const float4* __restrict__ rgbs, const float* __restrict__ z,
                     const float* __restrict__ rays, float* __restrict__ weights,
                     float* __restrict__ rgb, float* __restrict__ depth, int64_t n_rays, int k,
                     int per_lane, bool white) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * kCompWarps + (threadIdx.x >> 5);
  if (r >= n_rays) return;
  const float far = rays[r * 8 + 7];
  const int k0 = lane * per_lane;
  float zz[kMaxPerLane + 1];
  float al[kMaxPerLane];
  float4 c[kMaxPerLane];
#pragma unroll
  for (int i = 0; i <= kMaxPerLane; ++i) {
    if (i <= per_lane) {
      const int kk = k0 + i;
      zz[i] = kk < k ? z[r * k + kk] : far;  // z_K := far gives delta_{K-1} = far - z_{K-1}
    }
  }
  float prod = 1.0f;
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i) {
    if (i < per_lane) {
      const int kk = k0 + i;
      if (kk < k) {
        c[i] = rgbs[r * k + kk];
        const float delta = zz[i + 1] - zz[i];
        al[i] = 1.0f - expf(-delta * fmaxf(c[i].w, 0.0f));
        prod *= (1.0f - al[i] + 1e-10f);
      } else {
        al[i] = 0.0f;
        c[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  float T = warp_excl_prod(prod, lane);
  float sr = 0.f, sg = 0.f, sb = 0.f, sd = 0.f, sw = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i) {
    if (i < per_lane) {
      const int kk = k0 + i;
      if (kk < k) {
        const float w = al[i] * T;
        T *= (1.0f - al[i] + 1e-10f);
        weights[r * k + kk] = w;
        sr += w * c[i].x; sg += w * c[i].y; sb += w * c[i].z;
        sd += w * zz[i];
        sw += w;
      }
    }
  }
  sr = warp_sum(sr); sg = warp_sum(sg); sb = warp_sum(sb); sd = warp_sum(sd); sw = warp_sum(sw);
  if (lane == 0) {
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
__global__ void __launch_bounds__(kCompWarps * 32)
composite_bwd_kernel(const float4* __restrict__ rgbs, const float* __restrict__ z,
                     const float* __restrict__ rays, const float* __restrict__ g_rgb,
                     const float* __restrict__ g_depth, const float* __restrict__ g_weights,
                     float4* __restrict__ g_rgbs, float* __restrict__ g_z, int64_t n_rays, int k,
                     int per_lane, bool white) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * kCompWarps + (threadIdx.x >> 5);
  if (r >= n_rays) return;
  const float far = rays[r * 8 + 7];
  const int k0 = lane * per_lane;
  const float gr = g_rgb[r * 3 + 0], gg = g_rgb[r * 3 + 1], gb = g_rgb[r * 3 + 2];
  const float gd = g_depth ? g_depth[r] : 0.0f;
  const float gbg = white ? (gr + gg + gb) : 0.0f;
  float zz[kMaxPerLane + 1], al[kMaxPerLane], Tk[kMaxPerLane], gw[kMaxPerLane], E[kMaxPerLane];
  float4 c[kMaxPerLane];
#pragma unroll
  for (int i = 0; i <= kMaxPerLane; ++i)
    if (i <= per_lane) { const int kk = k0 + i; zz[i] = kk < k ? z[r * k + kk] : far; }
  float prod = 1.0f;
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i) {
    if (i < per_lane) {
      const int kk = k0 + i;
      if (kk < k) {
        c[i] = rgbs[r * k + kk];
        E[i] = expf(-(zz[i + 1] - zz[i]) * fmaxf(c[i].w, 0.0f));
        al[i] = 1.0f - E[i];
        prod *= (1.0f - al[i] + 1e-10f);
      } else { al[i] = 0.f; E[i] = 1.f; c[i] = make_float4(0.f, 0.f, 0.f, 0.f); }
    }
  }
  float T = warp_excl_prod(prod, lane);
  // per-lane sum of w*g_w, then suffix sums across lanes
  float lane_tot = 0.0f;
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i) {
    if (i < per_lane) {
      const int kk = k0 + i;
      Tk[i] = T;
      gw[i] = 0.0f;
      if (kk < k) {
        gw[i] = gr * c[i].x + gg * c[i].y + gb * c[i].z + gd * zz[i] - gbg +
                (g_weights ? g_weights[r * k + kk] : 0.0f);
        lane_tot += al[i] * T * gw[i];
        T *= (1.0f - al[i] + 1e-10f);
      }
    }
  }
  // exclusive suffix sum over lanes: sum of lane_tot for lanes > lane
  float incl = lane_tot;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    float y = __shfl_down_sync(0xffffffffu, incl, d);
    if (lane + d < 32) incl += y;
  }
  float suffix = incl - lane_tot;  // samples in later lanes
  float gdel[kMaxPerLane];
  // walk this lane's samples from last to first, extending the suffix
#pragma unroll
  for (int i = kMaxPerLane - 1; i >= 0; --i) {
    if (i < per_lane) {
      const int kk = k0 + i;
      gdel[i] = 0.0f;
      if (kk < k) {
        const float w = al[i] * Tk[i];
        const float q = 1.0f - al[i] + 1e-10f;
        const float ga = Tk[i] * gw[i] - suffix / q;
        suffix += w * gw[i];
        const float delta = zz[i + 1] - zz[i];
        const float s = fmaxf(c[i].w, 0.0f);
        gdel[i] = ga * s * E[i];
        const float gs = c[i].w > 0.0f ? ga * delta * E[i] : 0.0f;
        g_rgbs[r * k + kk] = make_float4(w * gr, w * gg, w * gb, gs);
      }
    }
  }
  if (g_z) {
    // g_delta of the sample just before this lane's first one
    float last = 0.0f;
#pragma unroll
    for (int i = 0; i < kMaxPerLane; ++i)
      if (i < per_lane && i == per_lane - 1) last = gdel[i];
    float prev = __shfl_up_sync(0xffffffffu, last, 1);
    if (lane == 0) prev = 0.0f;
#pragma unroll
    for (int i = 0; i < kMaxPerLane; ++i) {
      if (i < per_lane) {
        const int kk = k0 + i;
        if (kk < k) g_z[r * k + kk] = prev - gdel[i] + al[i] * Tk[i] * gd;
        prev = gdel[i];
      }
    }
  }
}

}  // namespace pn

static int composite_args_ok(int64_t n_rays, int k) {
  return n_rays >= 0 && k > 0 && k <= 32 * pn::kMaxPerLane;
}

extern "C" int pn_composite_fwd(const float* rgb_sigma, const float* z, const float* rays,
                                float* weights, float* rgb, float* depth, int64_t n_rays,
                                int n_samples, int white_bkgd, pn_stream_t stream) {
  if (n_rays == 0) return PN_OK;  // empty batches carry null data pointers
  PN_CHECK_ARG(rgb_sigma && z && rays && weights && rgb && depth, "pn_composite_fwd: null pointer");
  PN_CHECK_ARG(composite_args_ok(n_rays, n_samples), "pn_composite_fwd: need 1 <= K <= %d, got K=%d",
               32 * pn::kMaxPerLane, n_samples);
  if (n_rays == 0) return PN_OK;
  const int per_lane = (n_samples + 31) / 32;
  pn::composite_fwd_kernel<<<pn::ceil_div(n_rays, pn::kCompWarps), pn::kCompWarps * 32, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4*>(rgb_sigma), z, rays, weights, rgb, depth, n_rays, n_samples,
      per_lane, white_bkgd != 0);
  PN_LAUNCH_OK("composite_fwd_kernel");
  pn::count_launch();
  return PN_OK;
}

extern "C" int pn_composite_bwd(const float* rgb_sigma, const float* z, const float* rays,
                                const float* g_rgb, const float* g_depth, const float* g_weights,
                                float* g_rgb_sigma, float* g_z, int64_t n_rays, int n_samples,
                                int white_bkgd, pn_stream_t stream) {
  if (n_rays == 0) return PN_OK;  // empty batches carry null data pointers
  PN_CHECK_ARG(rgb_sigma && z && rays && g_rgb && g_rgb_sigma, "pn_composite_bwd: null pointer");
  PN_CHECK_ARG(composite_args_ok(n_rays, n_samples), "pn_composite_bwd: need 1 <= K <= %d, got K=%d",
               32 * pn::kMaxPerLane, n_samples);
  if (n_rays == 0) return PN_OK;
  const int per_lane = (n_samples + 31) / 32;
  pn::composite_bwd_kernel<<<pn::ceil_div(n_rays, pn::kCompWarps), pn::kCompWarps * 32, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4*>(rgb_sigma), z, rays, g_rgb, g_depth, g_weights,
      reinterpret_cast<float4*>(g_rgb_sigma), g_z, n_rays, n_samples, per_lane, white_bkgd != 0);
  PN_LAUNCH_OK("composite_bwd_kernel");
  pn::count_launch();
  return PN_OK;
}
