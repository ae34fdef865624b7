// Alpha compositing, forward and hand-written backward
// (reference src/render/nerf.py:178-182, 228-249; backward = autograd of the same).
//
// G lanes per ray (G = 8 for the usual K <= 128, so one warp composites 4 rays).  Lane l of a
// group owns the contiguous samples [l*C, (l+1)*C), C = ceil(K/G): its rgb-sigma values are C
// consecutive float4 (whole 128-byte lines for C = 8), z and the weights move as float4 when the
// row fits exactly.  The exclusive transmittance product is a per-lane serial product plus one
// segmented scan over the G lanes (log2 G shuffles), the five ray sums one segmented butterfly.
// HBM-bound: per ray the forward reads K*(16+4)+8 bytes and writes K*4+16.
#include "pn_common.cuh"

namespace pn {

constexpr int kCompWarps = 8;

template <int G>
__device__ __forceinline__ float group_excl_prod(float v, int gl) {
  float x = v;
#pragma unroll
  for (int d = 1; d < G; d <<= 1) {
    const float y = __shfl_up_sync(0xffffffffu, x, d, G);
    if (gl >= d) x *= y;
  }
  const float e = __shfl_up_sync(0xffffffffu, x, 1, G);
  return gl == 0 ? 1.0f : e;
}
template <int G>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int d = G / 2; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d, G);
  return v;
}

// z of this lane's samples (far beyond the row end) and the first z of the next lane's segment
template <int G, int CMAX>
__device__ __forceinline__ void load_z(const float* __restrict__ zrow, int k, int per_lane, int k0, int gl,
                                       float far, bool exact4, float (&zz)[CMAX + 1]) {
  if (exact4) {
#pragma unroll
    for (int i = 0; i < CMAX; i += 4)
      if (i < per_lane) {
        const float4 t = *reinterpret_cast<const float4*>(zrow + k0 + i);
        zz[i] = t.x; zz[i + 1] = t.y; zz[i + 2] = t.z; zz[i + 3] = t.w;
      }
  } else {
#pragma unroll
    for (int i = 0; i < CMAX; ++i)
      if (i < per_lane) zz[i] = (k0 + i) < k ? zrow[k0 + i] : far;
  }
  const float nxt = __shfl_down_sync(0xffffffffu, zz[0], 1, G);
  const float zn = gl == G - 1 ? far : nxt;   // z_K := far gives delta_{K-1} = far - z_{K-1}
#pragma unroll
  for (int i = 0; i < CMAX; ++i)
    if (i == per_lane - 1) zz[i + 1] = zn;
}

template <int G, int CMAX>
__global__ void __launch_bounds__(kCompWarps * 32, CMAX <= 8 ? 4 : CMAX <= 12 ? 3 : 2)
composite_fwd_kernel(const float4* __restrict__ rgbs, const float* __restrict__ z,
                     const float* __restrict__ rays, float* __restrict__ weights,
                     float* __restrict__ rgb, float* __restrict__ depth, int64_t n_rays, int k,
                     int per_lane, bool white) {
  const int lane = threadIdx.x & 31, gl = lane % G;
  const int64_t r_raw = ((int64_t)blockIdx.x * kCompWarps + (threadIdx.x >> 5)) * (32 / G) + lane / G;
  const bool live = r_raw < n_rays;
  const int64_t r = live ? r_raw : n_rays - 1;   // dead groups recompute the last ray and store nothing
  const float far = rays[r * 8 + 7];
  const int k0 = gl * per_lane;
  const bool exact4 = (per_lane * G == k) && (per_lane % 4 == 0);
  float zz[CMAX + 1];
  float al[CMAX];
  float4 c[CMAX];
  load_z<G, CMAX>(z + r * k, k, per_lane, k0, gl, far, exact4, zz);
  float prod = 1.0f;
#pragma unroll
  for (int i = 0; i < CMAX; ++i) {
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
  float T = group_excl_prod<G>(prod, gl);
  float sr = 0.f, sg = 0.f, sb = 0.f, sd = 0.f, sw = 0.f;
  float wv[CMAX];
#pragma unroll
  for (int i = 0; i < CMAX; ++i) {
    if (i < per_lane) {
      const float w = al[i] * T;   // zero for samples past the row end (al = 0)
      T *= (1.0f - al[i] + 1e-10f);
      wv[i] = w;
      sr += w * c[i].x; sg += w * c[i].y; sb += w * c[i].z;
      sd += w * zz[i];
      sw += w;
    }
  }
  if (live) {
    if (exact4) {
#pragma unroll
      for (int i = 0; i < CMAX; i += 4)
        if (i < per_lane)
          *reinterpret_cast<float4*>(weights + r * k + k0 + i) = make_float4(wv[i], wv[i + 1], wv[i + 2], wv[i + 3]);
    } else {
#pragma unroll
      for (int i = 0; i < CMAX; ++i)
        if (i < per_lane && k0 + i < k) weights[r * k + k0 + i] = wv[i];
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
__global__ void __launch_bounds__(kCompWarps * 32, 2)
composite_bwd_kernel(const float4* __restrict__ rgbs, const float* __restrict__ z,
                     const float* __restrict__ rays, const float* __restrict__ g_rgb,
                     const float* __restrict__ g_depth, const float* __restrict__ g_weights,
                     float4* __restrict__ g_rgbs, float* __restrict__ g_z, int64_t n_rays, int k,
                     int per_lane, bool white) {
  const int lane = threadIdx.x & 31, gl = lane % G;
  const int64_t r_raw = ((int64_t)blockIdx.x * kCompWarps + (threadIdx.x >> 5)) * (32 / G) + lane / G;
  const bool live = r_raw < n_rays;
  const int64_t r = live ? r_raw : n_rays - 1;
  const float far = rays[r * 8 + 7];
  const int k0 = gl * per_lane;
  const bool exact4 = (per_lane * G == k) && (per_lane % 4 == 0);
  const float gr = g_rgb[r * 3 + 0], gg = g_rgb[r * 3 + 1], gb = g_rgb[r * 3 + 2];
  const float gd = g_depth ? g_depth[r] : 0.0f;
  const float gbg = white ? (gr + gg + gb) : 0.0f;
  // per sample: E = exp(-delta s) (alpha = 1 - E is recomputed where needed), T_k, g_w
  float zz[CMAX + 1], Tk[CMAX], gw[CMAX], E[CMAX];
  float4 c[CMAX];
  load_z<G, CMAX>(z + r * k, k, per_lane, k0, gl, far, exact4, zz);
  // upstream weight gradients of this lane's samples, as float4 when the row splits evenly
#pragma unroll
  for (int i = 0; i < CMAX; ++i) gw[i] = 0.0f;
  if (g_weights) {
    if (exact4) {
#pragma unroll
      for (int i = 0; i < CMAX; i += 4)
        if (i < per_lane) {
          const float4 t = *reinterpret_cast<const float4*>(g_weights + r * k + k0 + i);
          gw[i] = t.x; gw[i + 1] = t.y; gw[i + 2] = t.z; gw[i + 3] = t.w;
        }
    } else {
#pragma unroll
      for (int i = 0; i < CMAX; ++i)
        if (i < per_lane && k0 + i < k) gw[i] = g_weights[r * k + k0 + i];
    }
  }
  float prod = 1.0f;
#pragma unroll
  for (int i = 0; i < CMAX; ++i) {
    if (i < per_lane) {
      const int kk = k0 + i;
      if (kk < k) {
        c[i] = rgbs[r * k + kk];
        E[i] = expf(-(zz[i + 1] - zz[i]) * fmaxf(c[i].w, 0.0f));
        prod *= (1.0f - (1.0f - E[i]) + 1e-10f);
      } else { E[i] = 1.f; c[i] = make_float4(0.f, 0.f, 0.f, 0.f); }
    }
  }
  float T = group_excl_prod<G>(prod, gl);
  // per-lane sum of w*g_w, then suffix sums across the group's lanes
  float lane_tot = 0.0f;
#pragma unroll
  for (int i = 0; i < CMAX; ++i) {
    if (i < per_lane) {
      const int kk = k0 + i;
      Tk[i] = T;
      if (kk < k) {
        const float al = 1.0f - E[i];
        gw[i] = gr * c[i].x + gg * c[i].y + gb * c[i].z + gd * zz[i] - gbg + gw[i];
        lane_tot += al * T * gw[i];
        T *= (1.0f - al + 1e-10f);
      } else {
        gw[i] = 0.0f;
      }
    }
  }
  float incl = lane_tot;   // inclusive suffix sum over the lanes of the group
#pragma unroll
  for (int d = 1; d < G; d <<= 1) {
    const float y = __shfl_down_sync(0xffffffffu, incl, d, G);
    if (gl + d < G) incl += y;
  }
  float suffix = incl - lane_tot;  // samples owned by later lanes
  float gdel[CMAX];
  // walk this lane's samples from last to first, extending the suffix
#pragma unroll
  for (int i = CMAX - 1; i >= 0; --i) {
    if (i < per_lane) {
      const int kk = k0 + i;
      gdel[i] = 0.0f;
      if (kk < k) {
        const float al = 1.0f - E[i];
        const float w = al * Tk[i];
        const float q = 1.0f - al + 1e-10f;
        const float ga = Tk[i] * gw[i] - suffix / q;
        suffix += w * gw[i];
        const float delta = zz[i + 1] - zz[i];
        const float s = fmaxf(c[i].w, 0.0f);
        gdel[i] = ga * s * E[i];
        const float gs = c[i].w > 0.0f ? ga * delta * E[i] : 0.0f;
        if (live) g_rgbs[r * k + kk] = make_float4(w * gr, w * gg, w * gb, gs);
      }
    }
  }
  if (g_z) {
    // g_delta of the sample just before this lane's first one
    float last = 0.0f;
#pragma unroll
    for (int i = 0; i < CMAX; ++i)
      if (i == per_lane - 1) last = gdel[i];
    float prev = __shfl_up_sync(0xffffffffu, last, 1, G);
    if (gl == 0) prev = 0.0f;
#pragma unroll
    for (int i = 0; i < CMAX; ++i) {
      if (i < per_lane) {
        const int kk = k0 + i;
        if (live && kk < k) g_z[r * k + kk] = prev - gdel[i] + (1.0f - E[i]) * Tk[i] * gd;
        prev = gdel[i];
      }
    }
  }
}

}  // namespace pn

// lanes per ray and per-lane capacity for K samples: (8,8) K<=64, (8,16) K<=128, (16,16) K<=256, (32,16) K<=512
#define PN_COMP_DISPATCH(KERNEL, K, ...)                                                        \
  do {                                                                                          \
    if ((K) <= 64) KERNEL<8, 8> __VA_ARGS__;                                                    \
    else if ((K) <= 96) KERNEL<8, 12> __VA_ARGS__;                                              \
    else if ((K) <= 128) KERNEL<8, 16> __VA_ARGS__;                                             \
    else if ((K) <= 256) KERNEL<16, 16> __VA_ARGS__;                                            \
    else KERNEL<32, 16> __VA_ARGS__;                                                            \
  } while (0)

static int comp_lanes(int k) { return k <= 128 ? 8 : k <= 256 ? 16 : 32; }
static int composite_args_ok(int64_t n_rays, int k) { return n_rays >= 0 && k > 0 && k <= 512; }

extern "C" int pn_composite_fwd(const float* rgb_sigma, const float* z, const float* rays,
                                float* weights, float* rgb, float* depth, int64_t n_rays,
                                int n_samples, int white_bkgd, pn_stream_t stream) {
  if (n_rays == 0) return PN_OK;  // empty batches carry null data pointers
  PN_CHECK_ARG(rgb_sigma && z && rays && weights && rgb && depth, "pn_composite_fwd: null pointer");
  PN_CHECK_ARG(composite_args_ok(n_rays, n_samples), "pn_composite_fwd: need 1 <= K <= 512, got K=%d", n_samples);
  const int g = comp_lanes(n_samples);
  const int per_lane = (n_samples + g - 1) / g;
  const int blocks = pn::ceil_div(n_rays, pn::kCompWarps * (32 / g));
  PN_COMP_DISPATCH(pn::composite_fwd_kernel, n_samples,
                   <<<blocks, pn::kCompWarps * 32, 0, (cudaStream_t)stream>>>(
                       reinterpret_cast<const float4*>(rgb_sigma), z, rays, weights, rgb, depth, n_rays, n_samples,
                       per_lane, white_bkgd != 0));
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
  PN_CHECK_ARG(composite_args_ok(n_rays, n_samples), "pn_composite_bwd: need 1 <= K <= 512, got K=%d", n_samples);
  const int g = comp_lanes(n_samples);
  const int per_lane = (n_samples + g - 1) / g;
  const int blocks = pn::ceil_div(n_rays, pn::kCompWarps * (32 / g));
  PN_COMP_DISPATCH(pn::composite_bwd_kernel, n_samples,
                   <<<blocks, pn::kCompWarps * 32, 0, (cudaStream_t)stream>>>(
                       reinterpret_cast<const float4*>(rgb_sigma), z, rays, g_rgb, g_depth, g_weights,
                       reinterpret_cast<float4*>(g_rgb_sigma), g_z, n_rays, n_samples, per_lane, white_bkgd != 0));
  PN_LAUNCH_OK("composite_bwd_kernel");
  pn::count_launch();
  return PN_OK;
}
