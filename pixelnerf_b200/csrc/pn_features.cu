// Point features: world->view transform, projection, bilinear latent gather and
// positional encoding (reference src/model/models.py:157-227,
// src/model/encoder.py:80-109, src/model/code.py:30-42, src/util/util.py:58-65).
//
// Layout: the latent is channels-last (V,Hl,Wl,C) so each of the four bilinear taps is
// one contiguous C-vector; one warp produces one feature row, lanes stride the
// channel axis in 16-byte vectors (512 B per warp load instruction, fully coalesced).
// Rows are ordered (object, view, point) like the reference's repeat_interleave.
#include <algorithm>

#include "pn_common.cuh"

namespace pn {

// NCHW fp32 -> NHWC (fp32 or fp16) through a padded 32x32 shared tile.
template <typename OutT>
__global__ void __launch_bounds__(256)
nchw_to_nhwc_kernel(const float* __restrict__ src, OutT* __restrict__ dst, int channels, int hw) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z;
  const int c0 = blockIdx.y * 32, p0 = blockIdx.x * 32;
  const float* s = src + (size_t)n * channels * hw;
  OutT* d = dst + (size_t)n * channels * hw;
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int c = c0 + i, p = p0 + threadIdx.x;
    tile[i][threadIdx.x] = (c < channels && p < hw) ? s[(size_t)c * hw + p] : 0.0f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int p = p0 + i, c = c0 + threadIdx.x;
    if (p < hw && c < channels) d[(size_t)p * channels + c] = (OutT)tile[threadIdx.x][i];
  }
}

struct CamView {
  float r[9], t[3], fx, fy, cx, cy;
};

__device__ __forceinline__ void store_vec4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void store_vec4(__half* p, float4 v) {
  __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = u;
}
__device__ __forceinline__ float4 load_vec4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 load_vec4(const __half* p) {
  uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
  float2 a = __half22float2(*reinterpret_cast<__half2*>(&u.x));
  float2 b = __half22float2(*reinterpret_cast<__half2*>(&u.y));
  return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void store_one(float* p, float v) { *p = v; }
__device__ __forceinline__ void store_one(__half* p, float v) { *p = __float2half_rn(v); }

#ifndef PN_FEAT_WARPS
#define PN_FEAT_WARPS 8
#endif
constexpr int kFeatWarps = PN_FEAT_WARPS;

// 16-byte vector of latent / feature elements as floats
template <typename T> struct Vec16;
template <> struct Vec16<float> {
  static constexpr int N = 4;
  typedef float4 Raw;
  static __device__ __forceinline__ Raw load_raw(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
  static __device__ __forceinline__ void unpack(const Raw& t, float (&v)[4]) { v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
  static __device__ __forceinline__ void load(const float* p, float (&v)[4]) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(p));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <> struct Vec16<__half> {
  static constexpr int N = 8;
  typedef uint4 Raw;
  static __device__ __forceinline__ Raw load_raw(const __half* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
  static __device__ __forceinline__ void unpack(const Raw& t, float (&v)[8]) {
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
  }
  static __device__ __forceinline__ void load(const __half* p, float (&v)[8]) {
    const uint4 t = __ldg(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
  }
  static __device__ __forceinline__ void store(__half* p, const float (&v)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __half2 h = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  }
};

// Each warp produces 32 consecutive feature rows.  Phase 1: lane l projects row l (camera
// transform, uv, the four tap offsets and bilinear weights) and writes that row's positional code
// (its own 36 sinf) -- per-row scalar work is done once, not 32 times.  Phase 2: for each of the
// 32 rows the tap offsets / weights are broadcast by shuffle and the whole warp blends the four
// contiguous channels-last taps with 16-byte loads and stores (512 B per warp request).
// kTaps (tensor-core path, fp16): no blend; each row's 128 bytes = [code 88 B, zero 8 B, taps record 32 B] are
// staged in shared memory (row pitch 144 B: 8-byte stores of a half-warp land 2-way, 16-byte reads of a
// quarter-warp conflict free) and leave as fully coalesced 16-byte stores, 512 B per warp instruction --
// one lane storing its own row directly touches 32 lines for 8 bytes each per instruction.
constexpr int kTapsPitch = 144;

template <typename LatT, typename OutT, bool kTaps, bool kLo = false>   // kLo: also the x_lo rows (precise chains)
__global__ void __launch_bounds__(kFeatWarps * 32)
point_features_kernel(pn_camera_t cam, const LatT* __restrict__ latent,
                      const RaySrc rays, const float* __restrict__ zs, int n_samples,
                      const float* __restrict__ xyz_in, const float* __restrict__ dirs_in,
                      int64_t ppo, OutT* __restrict__ feat_z, OutT* __restrict__ feat_x, int ldx,
                      int64_t n_rows, uint4* __restrict__ taps, int group = 0, __half* __restrict__ xlo = nullptr) {
  __shared__ __align__(16) unsigned char stage_all[kTaps ? kFeatWarps * 32 * kTapsPitch : 16];
  const int lane = threadIdx.x & 31;
  const int64_t row0 = ((int64_t)blockIdx.x * kFeatWarps + (threadIdx.x >> 5)) * 32;
  if (row0 >= n_rows) return;
  const int64_t row = row0 + lane;
  bool valid = row < n_rows;
  unsigned char* stage = stage_all + (kTaps ? (threadIdx.x >> 5) * 32 * kTapsPitch : 0);
  uint32_t o00 = 0, o01 = 0, o10 = 0, o11 = 0;
  float w00 = 0.f, w01 = 0.f, w10 = 0.f, w11 = 0.f;
  uint2 lo_keep[kLo ? 11 : 1];     // precise chains: this lane's x_lo row (44 fp16), zero for rows that are not written
#pragma unroll
  for (int i = 0; i < (kLo ? 11 : 1); ++i) lo_keep[i] = make_uint2(0u, 0u);
  // row = (obj * NS + view) * P + p ...
  // (index arithmetic in 32 bits whenever the launch fits -- warp uniform; a 64-bit division is ~50 instructions and
  // this kernel is instruction bound: eight of them were a quarter of the taps kernel)
  const bool idx32 = n_rows <= 0x7fffffffll && (int64_t)cam.n_objects * ppo <= 0x7fffffffll;
  int64_t ov, p;
  if (idx32) {
    const uint32_t o32 = (uint32_t)row / (uint32_t)ppo;
    ov = o32;
    p = (uint32_t)row - o32 * (uint32_t)ppo;
  } else {
    ov = row / ppo;
    p = row - ov * ppo;
  }
  if (kTaps && group > 0 && idx32) {
    const uint32_t blk = (uint32_t)row >> 7, nv = (uint32_t)cam.n_views;
    const uint32_t s = blk / nv, v = blk - s * nv;
    const uint32_t q = s * 128u + ((uint32_t)row & 127u);
    valid = q < (uint32_t)cam.n_objects * (uint32_t)ppo;
    const uint32_t obj_q = q / (uint32_t)ppo;
    p = q - obj_q * (uint32_t)ppo;
    ov = obj_q * nv + v;
  } else if (kTaps && group > 0) {
    // ... or, grouped order (pn_point_taps_grouped, for the chain's on-chip view mean): super-tile S = 128 points
    // [128 S, 128 S + 128) of the flattened (object, point) index, one 128-row block per view:
    // row (S * n_views + v) * 128 + m = view v of point 128 S + m; points past the end are zero rows (zero taps weights)
    const int64_t blk = row >> 7;
    const int v = (int)(blk % cam.n_views);
    const int64_t q = (blk / cam.n_views) * 128 + (row & 127);
    valid = q < (int64_t)cam.n_objects * ppo;
    const int64_t obj_q = q / ppo;
    p = q - obj_q * ppo;
    ov = obj_q * cam.n_views + v;
  }
  if (kTaps && group > 0) {
    if (!valid) {
#pragma unroll
      for (int i = 0; i < 8; ++i) *reinterpret_cast<uint4*>(stage + lane * kTapsPitch + 16 * i) = make_uint4(0u, 0u, 0u, 0u);

    }
  }
  if (valid) {
    const int view = (int)ov;
    const int obj = view / cam.n_views;
    float x, y, z, dx, dy, dz;
    if (rays.rays || rays.poses) {  // x = o + z d  (src/render/nerf.py:185), separate mul and add like torch
      int64_t ray;
      int ks;
      if (idx32) {
        const uint32_t ns32 = (uint32_t)n_samples, pr = (uint32_t)p / ns32;
        ks = (int)((uint32_t)p - pr * ns32);
        ray = (int64_t)((uint32_t)obj * ((uint32_t)ppo / ns32) + pr);
      } else {
        ray = obj * (ppo / n_samples) + p / n_samples;
        ks = (int)(p % n_samples);
      }
      float4 r0, r1;
      ray_load(rays, ray, r0, r1);
      const float zz = zs[ray * n_samples + ks];
      dx = r0.w; dy = r1.x; dz = r1.y;
      x = __fadd_rn(r0.x, __fmul_rn(zz, dx));
      y = __fadd_rn(r0.y, __fmul_rn(zz, dy));
      z = __fadd_rn(r0.z, __fmul_rn(zz, dz));
    } else {
      const float* pp = xyz_in + ((int64_t)obj * ppo + p) * 3;
      x = pp[0]; y = pp[1]; z = pp[2];
      if (dirs_in) {
        const float* dd = dirs_in + ((int64_t)obj * ppo + p) * 3;
        dx = dd[0]; dy = dd[1]; dz = dd[2];
      } else { dx = dy = dz = 0.0f; }
    }
    const float* P = cam.poses + (size_t)view * 12;  // (3,4) row-major [R | t]
    const float r0 = P[0], r1 = P[1], r2 = P[2], t0 = P[3];
    const float r3 = P[4], r4 = P[5], r5 = P[6], t1 = P[7];
    const float r6 = P[8], r7 = P[9], r8 = P[10], t2 = P[11];
    // xyz_rot = R x (models.py:162-164); xyz = xyz_rot + t (:165)
    const float xr = r0 * x + r1 * y + r2 * z;
    const float yr = r3 * x + r4 * y + r5 * z;
    const float zr = r6 * x + r7 * y + r8 * z;
    const float xc = xr + t0, yc = yr + t1, zc = zr + t2;
    // uv = -xy/z * focal + c (models.py:206-212); focal = [fx, -fy]
    const float u = (-xc / zc) * cam.focal[view * 2 + 0] + cam.c[view * 2 + 0];
    const float v = (-yc / zc) * cam.focal[view * 2 + 1] + cam.c[view * 2 + 1];
    // encoder.index (encoder.py:98-108): uv*scale - 1, then grid_sample(align_corners, border)
    const float gx = u * cam.scale_x - 1.0f, gy = v * cam.scale_y - 1.0f;
    float ix = ((gx + 1.0f) * 0.5f) * (float)(cam.wl - 1);
    float iy = ((gy + 1.0f) * 0.5f) * (float)(cam.hl - 1);
    ix = fminf(fmaxf(ix, 0.0f), (float)(cam.wl - 1));
    iy = fminf(fmaxf(iy, 0.0f), (float)(cam.hl - 1));
    // NaN (point on the camera plane) propagates through fminf/fmaxf as the clamp bound: texel 0
    const float fx0 = floorf(ix), fy0 = floorf(iy);
    const int x0 = (int)fx0, y0 = (int)fy0;
    const int x1 = min(x0 + 1, cam.wl - 1), y1 = min(y0 + 1, cam.hl - 1);
    const float ax = ix - fx0, ay = iy - fy0;  // weight of the +1 tap (0 when clamped at the border)
    w00 = (1.0f - ax) * (1.0f - ay); w01 = ax * (1.0f - ay);
    w10 = (1.0f - ax) * ay; w11 = ax * ay;
    const uint32_t vbase = (uint32_t)view * (uint32_t)(cam.hl * cam.wl);
    o00 = vbase + (uint32_t)(y0 * cam.wl + x0);
    o01 = vbase + (uint32_t)(y0 * cam.wl + x1);
    o10 = vbase + (uint32_t)(y1 * cam.wl + x0);
    o11 = vbase + (uint32_t)(y1 * cam.wl + x1);
    if (kTaps) {  // tensor-core path: the chain kernel blends per-texel lin_z tables with these.  The record
      // rides in bytes [96,128) of the row's 128-byte positional-code chunk (fp16 columns 48-63, which the
      // K=48 lin_in MMA never reads), so it reaches shared memory with the TMA load of x.
      uint4* rec = reinterpret_cast<uint4*>(stage + lane * kTapsPitch + 96);
      rec[0] = make_uint4(o00, o01, o10, o11);
      rec[1] = make_uint4(__float_as_uint(w00), __float_as_uint(w01), __float_as_uint(w10), __float_as_uint(w11));
    }
    o00 *= PN_D_LATENT; o01 *= PN_D_LATENT; o10 *= PN_D_LATENT; o11 *= PN_D_LATENT;
    // code: [x(3), sin(f_j x + phase_j)(36), R d (3)], zero pad to ldx; 4 elements at a time
    OutT* ox = kTaps ? reinterpret_cast<OutT*>(stage + lane * kTapsPitch) : feat_x + (size_t)row * ldx;
    const float comp[3] = {xr, yr, zr};
    const float rd[3] = {r0 * dx + r1 * dy + r2 * dz, r3 * dx + r4 * dy + r5 * dz, r6 * dx + r7 * dy + r8 * dz};
#pragma unroll
    for (int e0 = 0; e0 < 44; e0 += 4) {   // 42 real entries; fully unrolled so every index is constant
      float q[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int e = e0 + i;
        float val = 0.0f;
        if (e < 3) {
          val = comp[e];
        } else if (e < 39) {
          const int j = (e - 3) / 3, d = (e - 3) - j * 3;
          const float freq = 1.5f * (float)(1 << (j >> 1));      // freq_factor * 2^k (conf/default.conf:19-23)
          const float phase = (j & 1) ? 1.57079637050628662109375f : 0.0f;  // float32(pi/2) (code.py:24-28)
          const float arg = __fadd_rn(phase, __fmul_rn(comp[d], freq));   // addcmul then sin (code.py:38)
          // The fp16 code (tensor-core paths: taps and the materialised training features, which must agree bit for
          // bit -- tests/test_gpu_guards.py; rounded to 2^-11 relative right below) takes the SFU sine:
          // |error| <= |arg| * 2^-24 + 2^-21.4 (arguments reach ~100: 6e-6), a fortieth of the fp16 quantum, against
          // ~25 instructions for sinf -- 36 of them per row were two thirds of this kernel.  The fp32 code (parity
          // path, and the fp16 hi + lo pair of the precise chains) keeps sinf.
          val = (sizeof(OutT) == 2 && !kLo) ? __sinf(arg) : sinf(arg);
        } else if (e < 42) {
          val = rd[e - 39];
        }
        q[i] = val;
      }
      if constexpr (kLo) {  // precise chains: the fp16 remainder of the code, x_lo = fp16(x - fp16(x)), as a second row
        float l[4];           // (kept in registers: it leaves through the staging buffer after the first row has)
#pragma unroll
        for (int i = 0; i < 4; ++i) l[i] = q[i] - __half2float(__float2half_rn(q[i]));
        const __half2 la = __floats2half2_rn(l[0], l[1]), lb = __floats2half2_rn(l[2], l[3]);
        lo_keep[e0 >> 2] = make_uint2(*reinterpret_cast<const uint32_t*>(&la), *reinterpret_cast<const uint32_t*>(&lb));
      }
      if (sizeof(OutT) == 4) {
        *reinterpret_cast<float4*>(reinterpret_cast<float*>(ox) + e0) = make_float4(q[0], q[1], q[2], q[3]);
      } else {
        const __half2 a = __floats2half2_rn(q[0], q[1]), b = __floats2half2_rn(q[2], q[3]);
        *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(ox) + e0) =
            make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
      }
    }
    for (int e0 = 44; e0 < (kTaps ? 48 : ldx); e0 += 4) {  // zero padding up to ldx (up to the taps record)
      if (sizeof(OutT) == 4) *reinterpret_cast<float4*>(reinterpret_cast<float*>(ox) + e0) = make_float4(0.f, 0.f, 0.f, 0.f);
      else *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(ox) + e0) = make_uint2(0u, 0u);
    }
  }
  if (kTaps) {
    __syncwarp();
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int j = it * 32 + lane, rr = j >> 3, piece = j & 7;
      if (row0 + rr < n_rows)
        taps[(row0 + rr) * 8 + piece] = *reinterpret_cast<const uint4*>(stage + rr * kTapsPitch + piece * 16);
    }
    if constexpr (kLo) {   // the x_lo rows take the same way out: staged, then coalesced 16-byte stores (zero rows stay zero)
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 11; ++i) *reinterpret_cast<uint2*>(stage + lane * kTapsPitch + 8 * i) = lo_keep[i];
#pragma unroll
      for (int i = 11; i < 16; ++i) *reinterpret_cast<uint2*>(stage + lane * kTapsPitch + 8 * i) = make_uint2(0u, 0u);
      __syncwarp();
      uint4* dst = reinterpret_cast<uint4*>(xlo);
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int j = it * 32 + lane, rr = j >> 3, piece = j & 7;
        if (row0 + rr < n_rows)
          dst[(row0 + rr) * 8 + piece] = *reinterpret_cast<const uint4*>(stage + rr * kTapsPitch + piece * 16);
      }
    }
    return;
  }
  // phase 2: the warp blends the taps of each of its rows
  constexpr int N = Vec16<LatT>::N;
  constexpr int kIters = PN_D_LATENT / (32 * N);
  const int n_here = (int)((n_rows - row0) < 32 ? (n_rows - row0) : 32);
  // Consecutive rows of a warp are consecutive samples of one ray, and several of them fall into the same texel
  // cell: the four tap vectors stay in registers and are only re-read when a row names different texels (the
  // condition is warp uniform).  The kernel is bound by the L1 data path (4 KB of taps read per 1 KB written), so
  // every skipped reload is time: 58 % -> 69 % of the HBM peak on the bench frame (taps held packed).  (A second buffer that requests
  // the next row's taps before the current row is blended was measured slower, 58 %: 125 registers and the
  // buffer moves cost more than the overlap gains.)
  uint32_t p00 = 0xffffffffu, p01 = 0xffffffffu, p10 = 0xffffffffu, p11 = 0xffffffffu;
  typename Vec16<LatT>::Raw ta[kIters], tb[kIters], tc[kIters], td[kIters];   // kept packed: half the registers for fp16
#pragma unroll 1
  for (int r = 0; r < n_here; ++r) {
    const uint32_t a00 = __shfl_sync(0xffffffffu, o00, r), a01 = __shfl_sync(0xffffffffu, o01, r);
    const uint32_t a10 = __shfl_sync(0xffffffffu, o10, r), a11 = __shfl_sync(0xffffffffu, o11, r);
    const float b00 = __shfl_sync(0xffffffffu, w00, r), b01 = __shfl_sync(0xffffffffu, w01, r);
    const float b10 = __shfl_sync(0xffffffffu, w10, r), b11 = __shfl_sync(0xffffffffu, w11, r);
    OutT* oz = feat_z + (size_t)(row0 + r) * PN_D_LATENT;
    if (a00 != p00 || a01 != p01 || a10 != p10 || a11 != p11) {
#pragma unroll
      for (int it = 0; it < kIters; ++it) {
        const int ch = (it * 32 + lane) * N;
        ta[it] = Vec16<LatT>::load_raw(latent + a00 + ch);
        tb[it] = Vec16<LatT>::load_raw(latent + a01 + ch);
        tc[it] = Vec16<LatT>::load_raw(latent + a10 + ch);
        td[it] = Vec16<LatT>::load_raw(latent + a11 + ch);
      }
      p00 = a00; p01 = a01; p10 = a10; p11 = a11;
    }
#pragma unroll
    for (int it = 0; it < kIters; ++it) {
      const int ch = (it * 32 + lane) * N;
      float va[N], vb[N], vc[N], vd[N], o[N];
      Vec16<LatT>::unpack(ta[it], va);
      Vec16<LatT>::unpack(tb[it], vb);
      Vec16<LatT>::unpack(tc[it], vc);
      Vec16<LatT>::unpack(td[it], vd);
#pragma unroll
      for (int i = 0; i < N; ++i) o[i] = va[i] * b00 + vb[i] * b01 + vc[i] * b10 + vd[i] * b11;
      Vec16<OutT>::store(oz + ch, o);
    }
  }
}

// Backward of the bilinear gather w.r.t. the latent (encoder.py:102-108 = grid_sample backward for
// its input): g_latent[view, tap] += w_tap * g_feat_z[row].  One warp per row, 16-byte vector
// atomics (red.global.add.v4.f32) on the channels-last fp32 gradient map.  Sample positions carry
// no gradient here ("detached sample positions").
__global__ void __launch_bounds__(kFeatWarps * 32)
point_features_bwd_kernel(pn_camera_t cam, const float* __restrict__ rays, const float* __restrict__ zs,
                          int n_samples, const float* __restrict__ xyz_in, int64_t ppo,
                          const float* __restrict__ g_feat_z, float* __restrict__ g_latent, int64_t n_rows) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * kFeatWarps + (threadIdx.x >> 5);
  if (row >= n_rows) return;
  const int64_t ov = row / ppo;
  const int64_t p = row - ov * ppo;
  const int view = (int)ov;
  const int obj = view / cam.n_views;
  float x, y, z;
  if (rays) {
    const int64_t ray = obj * (ppo / n_samples) + p / n_samples;
    const float* rr = rays + ray * 8;
    const float zz = zs[ray * n_samples + (p % n_samples)];
    x = __fadd_rn(rr[0], __fmul_rn(zz, rr[3]));
    y = __fadd_rn(rr[1], __fmul_rn(zz, rr[4]));
    z = __fadd_rn(rr[2], __fmul_rn(zz, rr[5]));
  } else {
    const float* pp = xyz_in + ((int64_t)obj * ppo + p) * 3;
    x = pp[0]; y = pp[1]; z = pp[2];
  }
  const float* P = cam.poses + (size_t)view * 12;
  const float xc = P[0] * x + P[1] * y + P[2] * z + P[3];
  const float yc = P[4] * x + P[5] * y + P[6] * z + P[7];
  const float zc = P[8] * x + P[9] * y + P[10] * z + P[11];
  const float u = (-xc / zc) * cam.focal[view * 2 + 0] + cam.c[view * 2 + 0];
  const float v = (-yc / zc) * cam.focal[view * 2 + 1] + cam.c[view * 2 + 1];
  const float gx = u * cam.scale_x - 1.0f, gy = v * cam.scale_y - 1.0f;
  float ix = ((gx + 1.0f) * 0.5f) * (float)(cam.wl - 1);
  float iy = ((gy + 1.0f) * 0.5f) * (float)(cam.hl - 1);
  ix = fminf(fmaxf(ix, 0.0f), (float)(cam.wl - 1));
  iy = fminf(fmaxf(iy, 0.0f), (float)(cam.hl - 1));
  const float fx0 = floorf(ix), fy0 = floorf(iy);
  const int x0 = (int)fx0, y0 = (int)fy0;
  const int x1 = min(x0 + 1, cam.wl - 1), y1 = min(y0 + 1, cam.hl - 1);
  const float ax = ix - fx0, ay = iy - fy0;
  const float w00 = (1.0f - ax) * (1.0f - ay), w01 = ax * (1.0f - ay);
  const float w10 = (1.0f - ax) * ay, w11 = ax * ay;
  float* base = g_latent + (size_t)view * cam.hl * cam.wl * PN_D_LATENT;
  float* p00 = base + ((size_t)y0 * cam.wl + x0) * PN_D_LATENT;
  float* p01 = base + ((size_t)y0 * cam.wl + x1) * PN_D_LATENT;
  float* p10 = base + ((size_t)y1 * cam.wl + x0) * PN_D_LATENT;
  float* p11 = base + ((size_t)y1 * cam.wl + x1) * PN_D_LATENT;
  const float* g = g_feat_z + (size_t)row * PN_D_LATENT;
#pragma unroll
  for (int i = 0; i < PN_D_LATENT / 128; ++i) {
    const int ch = i * 128 + lane * 4;
    const float4 gv = *reinterpret_cast<const float4*>(g + ch);
    atomicAdd(reinterpret_cast<float4*>(p00 + ch), make_float4(gv.x * w00, gv.y * w00, gv.z * w00, gv.w * w00));
    if (w01 != 0.0f) atomicAdd(reinterpret_cast<float4*>(p01 + ch), make_float4(gv.x * w01, gv.y * w01, gv.z * w01, gv.w * w01));
    if (w10 != 0.0f) atomicAdd(reinterpret_cast<float4*>(p10 + ch), make_float4(gv.x * w10, gv.y * w10, gv.z * w10, gv.w * w10));
    if (w11 != 0.0f) atomicAdd(reinterpret_cast<float4*>(p11 + ch), make_float4(gv.x * w11, gv.y * w11, gv.z * w11, gv.w * w11));
  }
}

// Gradient of the fine pass w.r.t. the coarse depth through the depth-guided sample positions
// (src/render/nerf.py:150-161 and :290-295): z_d = clamp(depth + noise * std, near, far) is the only sample set whose
// positions are not detached, so dL/d depth[ray] = sum over the ray's depth samples strictly inside (near, far) of
// dL/dz at that sample's place in the sorted row.  dL/dz has three parts:
//   compositing   g_z[k] from pn_composite_bwd (deltas and the depth output)
//   code rows     x = [R p, sin(f R p + phase), R d] (models.py:162-171, code.py:30-42):
//                 dL/dh0 . W_in . dx/dz with d(R p)/dz = R d and d sin(..)/dz = f cos(..) (R d)
//   latent rows   grid_sample's grid gradient (encoder.py:102-108; zero where the border clamp is active) through
//                 uv = -xy/z * focal + c (models.py:206-212)
// One warp per (ray, depth sample); lanes split the 512 hidden / latent channels; lin_in.weight sits transposed
// in shared memory ([code entry][hidden], conflict-free 16-byte reads).
constexpr int kDsWarps = 8;
constexpr int kDsCode = 39;   // code entries that depend on the position (the 3 view-direction entries do not)

template <bool kBf16>
__global__ void __launch_bounds__(kDsWarps * 32)
depth_sample_bwd_kernel(pn_camera_t cam, const float* __restrict__ latent, const float* __restrict__ rays,
                        const float* __restrict__ zs, int k, const float* __restrict__ z_dep, int kd,
                        const float* __restrict__ g_zd, const void* __restrict__ g_h0v,
                        const float* __restrict__ g_fz, const float* __restrict__ w_in, int64_t rays_per_object,
                        int64_t n_rays, float* __restrict__ g_depth) {
  extern __shared__ __align__(16) float ds_smem[];
  float* Ws = ds_smem;                                              // [39][512]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* dxs = ds_smem + kDsCode * 512 + warp * 40;                 // this warp's dx/dz vector
  for (int idx = threadIdx.x; idx < 512 * PN_D_CODE; idx += kDsWarps * 32) {
    const int o = idx / PN_D_CODE, i = idx - o * PN_D_CODE;
    if (i < kDsCode) Ws[i * 512 + o] = __ldg(w_in + idx);
  }
  __syncthreads();
  const int64_t n_items = n_rays * kd;
  for (int64_t item = (int64_t)blockIdx.x * kDsWarps + warp; item < n_items; item += (int64_t)gridDim.x * kDsWarps) {
    const int64_t ray = item / kd;
    const float* rr = rays + ray * 8;
    const float zd = __ldg(z_dep + item);
    if (!(zd > rr[6] && zd < rr[7])) continue;      // max(min(z, far), near): no gradient once a bound is active
    const float* zrow = zs + ray * k;
    int lo = 0, hi = k;
    while (lo < hi) {                               // the sample's place after the sort (nerf.py:295)
      const int mid = (lo + hi) >> 1;
      if (__ldg(zrow + mid) < zd) lo = mid + 1; else hi = mid;
    }
    if (lo >= k || __ldg(zrow + lo) != zd) continue;
    const int64_t obj = ray / rays_per_object, rio = ray - obj * rays_per_object;
    const float dwx = rr[3], dwy = rr[4], dwz = rr[5];
    const float x = __fadd_rn(rr[0], __fmul_rn(zd, dwx));
    const float y = __fadd_rn(rr[1], __fmul_rn(zd, dwy));
    const float z = __fadd_rn(rr[2], __fmul_rn(zd, dwz));
    float total = g_zd ? __ldg(g_zd + ray * k + lo) : 0.0f;
    for (int v = 0; v < cam.n_views; ++v) {
      const int view = (int)obj * cam.n_views + v;
      const int64_t row = ((int64_t)view * rays_per_object + rio) * k + lo;
      const float* P = cam.poses + (size_t)view * 12;
      const float comp[3] = {P[0] * x + P[1] * y + P[2] * z, P[4] * x + P[5] * y + P[6] * z, P[8] * x + P[9] * y + P[10] * z};
      const float rd[3] = {P[0] * dwx + P[1] * dwy + P[2] * dwz, P[4] * dwx + P[5] * dwy + P[6] * dwz,
                           P[8] * dwx + P[9] * dwy + P[10] * dwz};
      const float xc = comp[0] + P[3], yc = comp[1] + P[7], zc = comp[2] + P[11];
      // d(code entry)/dz
      for (int e = lane; e < kDsCode; e += 32) {
        float val;
        if (e < 3) {
          val = rd[e];
        } else {
          const int j = (e - 3) / 3, d = (e - 3) - j * 3;
          const float freq = 1.5f * (float)(1 << (j >> 1));
          const float phase = (j & 1) ? 1.57079637050628662109375f : 0.0f;
          val = freq * cosf(__fadd_rn(phase, __fmul_rn(comp[d], freq))) * rd[d];
        }
        dxs[e] = val;
      }
      __syncwarp();
      float s = 0.0f;
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int o = it * 128 + lane * 4;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 13
        for (int i = 0; i < kDsCode; ++i) {
          const float4 w = *reinterpret_cast<const float4*>(Ws + i * 512 + o);
          const float d = dxs[i];
          acc.x += w.x * d; acc.y += w.y * d; acc.z += w.z * d; acc.w += w.w * d;
        }
        float4 gh;
        if (kBf16) {
          const uint2 u = __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const uint16_t*>(g_h0v) + row * 512 + o));
          gh = make_float4(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xFFFF0000u),
                           __uint_as_float(u.y << 16), __uint_as_float(u.y & 0xFFFF0000u));
        } else {
          gh = __ldg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(g_h0v) + row * 512 + o));
        }
        s += acc.x * gh.x + acc.y * gh.y + acc.z * gh.z + acc.w * gh.w;
      }
      // grid gradient of the bilinear gather
      const float f0 = cam.focal[view * 2 + 0], f1 = cam.focal[view * 2 + 1];
      const float u = (-xc / zc) * f0 + cam.c[view * 2 + 0];
      const float vv = (-yc / zc) * f1 + cam.c[view * 2 + 1];
      const float gx = u * cam.scale_x - 1.0f, gy = vv * cam.scale_y - 1.0f;
      const float ixu = ((gx + 1.0f) * 0.5f) * (float)(cam.wl - 1);
      const float iyu = ((gy + 1.0f) * 0.5f) * (float)(cam.hl - 1);
      const bool in_x = ixu > 0.0f && ixu < (float)(cam.wl - 1), in_y = iyu > 0.0f && iyu < (float)(cam.hl - 1);
      const float ix = fminf(fmaxf(ixu, 0.0f), (float)(cam.wl - 1)), iy = fminf(fmaxf(iyu, 0.0f), (float)(cam.hl - 1));
      const float fx0 = floorf(ix), fy0 = floorf(iy);
      const int x0 = (int)fx0, y0 = (int)fy0;
      const int x1 = min(x0 + 1, cam.wl - 1), y1 = min(y0 + 1, cam.hl - 1);
      const float ax = ix - fx0, ay = iy - fy0;
      float gix = 0.0f, giy = 0.0f;
      if (in_x || in_y) {
        const float* base = latent + (size_t)view * cam.hl * cam.wl * PN_D_LATENT;
        const float* p00 = base + ((size_t)y0 * cam.wl + x0) * PN_D_LATENT;
        const float* p01 = base + ((size_t)y0 * cam.wl + x1) * PN_D_LATENT;
        const float* p10 = base + ((size_t)y1 * cam.wl + x0) * PN_D_LATENT;
        const float* p11 = base + ((size_t)y1 * cam.wl + x1) * PN_D_LATENT;
        const float* g = g_fz + (size_t)row * PN_D_LATENT;
#pragma unroll
        for (int it = 0; it < PN_D_LATENT / 128; ++it) {
          const int ch = it * 128 + lane * 4;
          const float4 gv = __ldg(reinterpret_cast<const float4*>(g + ch));
          const float4 a = load_vec4(p00 + ch), b = load_vec4(p01 + ch), c = load_vec4(p10 + ch), d = load_vec4(p11 + ch);
          gix += gv.x * ((1.0f - ay) * (b.x - a.x) + ay * (d.x - c.x)) + gv.y * ((1.0f - ay) * (b.y - a.y) + ay * (d.y - c.y)) +
                 gv.z * ((1.0f - ay) * (b.z - a.z) + ay * (d.z - c.z)) + gv.w * ((1.0f - ay) * (b.w - a.w) + ay * (d.w - c.w));
          giy += gv.x * ((1.0f - ax) * (c.x - a.x) + ax * (d.x - b.x)) + gv.y * ((1.0f - ax) * (c.y - a.y) + ax * (d.y - b.y)) +
                 gv.z * ((1.0f - ax) * (c.z - a.z) + ax * (d.z - b.z)) + gv.w * ((1.0f - ax) * (c.w - a.w) + ax * (d.w - b.w));
        }
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, off);
        gix += __shfl_xor_sync(0xffffffffu, gix, off);
        giy += __shfl_xor_sync(0xffffffffu, giy, off);
      }
      const float gu = in_x ? gix * (0.5f * (float)(cam.wl - 1)) * cam.scale_x : 0.0f;
      const float gv2 = in_y ? giy * (0.5f * (float)(cam.hl - 1)) * cam.scale_y : 0.0f;
      // u = -xc/zc f0 + c0, v = -yc/zc f1 + c1; d(xc, yc, zc)/dz = R d
      const float gxc = -gu * f0 / zc, gyc = -gv2 * f1 / zc, gzc = (gu * f0 * xc + gv2 * f1 * yc) / (zc * zc);
      total += s + gxc * rd[0] + gyc * rd[1] + gzc * rd[2];
      __syncwarp();
    }
    if (lane == 0) atomicAdd(g_depth + ray, total);
  }
}

}  // namespace pn

extern "C" int pn_point_features_bwd(const pn_camera_t* cam, const float* rays, const float* z, int n_samples,
                                     const float* xyz, int64_t points_per_object, const float* g_feat_z,
                                     float* g_latent_nhwc, pn_stream_t stream) {
  PN_CHECK_ARG(cam && g_feat_z && g_latent_nhwc, "pn_point_features_bwd: null pointer");
  PN_CHECK_ARG(cam->poses && cam->focal && cam->c, "pn_point_features_bwd: camera buffers missing");
  PN_CHECK_ARG((rays && z && n_samples > 0) || xyz, "pn_point_features_bwd: need rays+z or xyz");
  PN_CHECK_ARG(!rays || points_per_object % n_samples == 0, "pn_point_features_bwd: P must be a multiple of K");
  const int64_t n_rows = (int64_t)cam->n_objects * cam->n_views * points_per_object;
  if (n_rows == 0) return PN_OK;
  pn::point_features_bwd_kernel<<<pn::ceil_div(n_rows, pn::kFeatWarps), pn::kFeatWarps * 32, 0, (cudaStream_t)stream>>>(
      *cam, rays, z, n_samples, xyz, points_per_object, g_feat_z, g_latent_nhwc, n_rows);
  PN_LAUNCH_OK("point_features_bwd_kernel");
  pn::count_launch();
  return PN_OK;
}

extern "C" int pn_depth_sample_bwd(const pn_camera_t* cam, const float* latent_nhwc, const float* rays,
                                   const float* z_sorted, int n_samples, const float* z_depth, int n_depth,
                                   const float* g_z, const void* g_h0, int g_h0_dtype, const float* g_feat_z,
                                   const float* lin_in_w, int64_t rays_per_object, float* g_depth,
                                   pn_stream_t stream) {
  PN_CHECK_ARG(cam && latent_nhwc && rays && z_sorted && z_depth && g_h0 && g_feat_z && lin_in_w && g_depth,
               "pn_depth_sample_bwd: null pointer");
  PN_CHECK_ARG(cam->poses && cam->focal && cam->c, "pn_depth_sample_bwd: camera buffers missing");
  PN_CHECK_ARG(n_samples > 0 && n_depth >= 0 && rays_per_object >= 0, "pn_depth_sample_bwd: bad sizes");
  PN_CHECK_ARG(g_h0_dtype == PN_DTYPE_F32 || g_h0_dtype == PN_DTYPE_BF16, "pn_depth_sample_bwd: g_h0 must be f32 or bf16");
  PN_CHECK_ARG((((uintptr_t)g_h0 | (uintptr_t)g_feat_z | (uintptr_t)latent_nhwc) & 15) == 0, "pn_depth_sample_bwd: misaligned buffer");
  const int64_t n_rays = (int64_t)cam->n_objects * rays_per_object;
  const int64_t n_items = n_rays * n_depth;
  if (n_items == 0) return PN_OK;
  const size_t smem = (size_t)(pn::kDsCode * 512 + pn::kDsWarps * 40) * sizeof(float);
  const int blocks = (int)std::min<int64_t>(pn::ceil_div(n_items, pn::kDsWarps), 148 * 2);
  auto launch = [&](auto kern) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<blocks, pn::kDsWarps * 32, smem, (cudaStream_t)stream>>>(*cam, latent_nhwc, rays, z_sorted, n_samples, z_depth,
                                                                    n_depth, g_z, g_h0, g_feat_z, lin_in_w,
                                                                    rays_per_object, n_rays, g_depth);
  };
  if (g_h0_dtype == PN_DTYPE_BF16) launch(pn::depth_sample_bwd_kernel<true>);
  else launch(pn::depth_sample_bwd_kernel<false>);
  PN_LAUNCH_OK("depth_sample_bwd_kernel");
  pn::count_launch();
  return PN_OK;
}


extern "C" int pn_prepare_latent(const float* latent_nchw, void* latent_nhwc, int n_maps, int channels,
                                 int hl, int wl, int dtype, pn_stream_t stream) {
  PN_CHECK_ARG(latent_nchw && latent_nhwc, "pn_prepare_latent: null pointer");
  PN_CHECK_ARG(n_maps > 0 && channels > 0 && hl > 0 && wl > 0, "pn_prepare_latent: bad sizes");
  PN_CHECK_ARG(dtype == PN_DTYPE_F32 || dtype == PN_DTYPE_F16, "pn_prepare_latent: dtype must be f32 or f16");
  const int hw = hl * wl;
  dim3 grid(pn::ceil_div(hw, 32), pn::ceil_div(channels, 32), n_maps), block(32, 8);
  if (dtype == PN_DTYPE_F32)
    pn::nchw_to_nhwc_kernel<float><<<grid, block, 0, (cudaStream_t)stream>>>(latent_nchw, (float*)latent_nhwc, channels, hw);
  else
    pn::nchw_to_nhwc_kernel<__half><<<grid, block, 0, (cudaStream_t)stream>>>(latent_nchw, (__half*)latent_nhwc, channels, hw);
  PN_LAUNCH_OK("nchw_to_nhwc_kernel");
  pn::count_launch();
  return PN_OK;
}

extern "C" int pn_point_features(const pn_camera_t* cam, const void* latent_nhwc, int latent_dtype,
                                 const float* rays, const float* z, int n_samples, const float* xyz,
                                 const float* viewdirs, int64_t points_per_object, void* feat_z,
                                 void* feat_x, int ldx, int out_dtype, pn_stream_t stream) {
  PN_CHECK_ARG(cam && latent_nhwc && feat_z && feat_x, "pn_point_features: null pointer");
  PN_CHECK_ARG(cam->poses && cam->focal && cam->c, "pn_point_features: camera buffers missing (call encode first)");
  PN_CHECK_ARG((rays && z && n_samples > 0) || xyz, "pn_point_features: need rays+z or xyz");
  PN_CHECK_ARG(!rays || points_per_object % n_samples == 0, "pn_point_features: P must be a multiple of K");
  PN_CHECK_ARG(ldx >= 44 && ldx % 4 == 0, "pn_point_features: ldx=%d must be a multiple of 4 and >= 44", ldx);
  PN_CHECK_ARG(cam->hl > 1 && cam->wl > 1 && cam->n_objects > 0 && cam->n_views > 0, "pn_point_features: bad camera sizes");
  PN_CHECK_ARG(latent_dtype == out_dtype && (out_dtype == PN_DTYPE_F32 || out_dtype == PN_DTYPE_F16),
               "pn_point_features: latent/out dtype must both be f32 or both f16");
  const int64_t n_rows = (int64_t)cam->n_objects * cam->n_views * points_per_object;
  if (n_rows == 0) return PN_OK;
  const int blocks = pn::ceil_div(n_rows, pn::kFeatWarps * 32);
  if (out_dtype == PN_DTYPE_F32)
    pn::point_features_kernel<float, float, false><<<blocks, pn::kFeatWarps * 32, 0, (cudaStream_t)stream>>>(
        *cam, (const float*)latent_nhwc, pn::ray_src(rays), z, n_samples, xyz, viewdirs, points_per_object,
        (float*)feat_z, (float*)feat_x, ldx, n_rows, nullptr);
  else
    pn::point_features_kernel<__half, __half, false><<<blocks, pn::kFeatWarps * 32, 0, (cudaStream_t)stream>>>(
        *cam, (const __half*)latent_nhwc, pn::ray_src(rays), z, n_samples, xyz, viewdirs, points_per_object,
        (__half*)feat_z, (__half*)feat_x, ldx, n_rows, nullptr);
  PN_LAUNCH_OK("point_features_kernel");
  pn::count_launch();
  return PN_OK;
}

extern "C" int64_t pn_point_taps_grouped_rows(int64_t n_points, int n_views) {
  if (n_points <= 0 || n_views < 1) return 0;
  return (n_points + 127) / 128 * 128 * n_views;
}

// pn_point_taps / pn_point_taps_grouped with the rays coming from any RaySrc (an explicit tensor, or frames)
int pn::point_taps_src(const pn_camera_t* cam, const pn::RaySrc& rays, const float* z, int n_samples, const float* xyz,
                       const float* viewdirs, int64_t points_per_object, void* feat_x, bool grouped, pn_stream_t stream,
                       void* feat_x_lo) {
  const bool have_rays = rays.rays || rays.poses;
  PN_CHECK_ARG(cam && feat_x, "pn_point_taps: null pointer");
  PN_CHECK_ARG(cam->poses && cam->focal && cam->c, "pn_point_taps: camera buffers missing (call encode first)");
  PN_CHECK_ARG((have_rays && z && n_samples > 0) || xyz, "pn_point_taps: need rays+z or xyz");
  PN_CHECK_ARG(!have_rays || points_per_object % n_samples == 0, "pn_point_taps: P must be a multiple of K");
  PN_CHECK_ARG(cam->hl > 1 && cam->wl > 1 && cam->n_objects > 0 && cam->n_views > 0, "pn_point_taps: bad camera sizes");
  PN_CHECK_ARG((int64_t)cam->n_objects * cam->n_views * cam->hl * cam->wl < (1ll << 31), "pn_point_taps: too many texels");
  PN_CHECK_ARG(((uintptr_t)feat_x & 127) == 0, "pn_point_taps: feat_x must be 128-byte aligned");
  const int64_t n_rows = grouped ? pn_point_taps_grouped_rows((int64_t)cam->n_objects * points_per_object, cam->n_views)
                                 : (int64_t)cam->n_objects * cam->n_views * points_per_object;
  if (n_rows == 0) return PN_OK;
  const int blocks = pn::ceil_div(n_rows, pn::kFeatWarps * 32);
  if (feat_x_lo)
    pn::point_features_kernel<__half, __half, true, true><<<blocks, pn::kFeatWarps * 32, 0, (cudaStream_t)stream>>>(
        *cam, nullptr, rays, z, n_samples, xyz, viewdirs, points_per_object, nullptr, (__half*)feat_x, 64, n_rows,
        (uint4*)feat_x, grouped ? 1 : 0, (__half*)feat_x_lo);
  else
    pn::point_features_kernel<__half, __half, true, false><<<blocks, pn::kFeatWarps * 32, 0, (cudaStream_t)stream>>>(
        *cam, nullptr, rays, z, n_samples, xyz, viewdirs, points_per_object, nullptr, (__half*)feat_x, 64, n_rows,
        (uint4*)feat_x, grouped ? 1 : 0, nullptr);
  PN_LAUNCH_OK("point_features_kernel(taps)");
  pn::count_launch();
  return PN_OK;
}

extern "C" int pn_point_taps_grouped(const pn_camera_t* cam, const float* rays, const float* z, int n_samples,
                                     const float* xyz, const float* viewdirs, int64_t points_per_object,
                                     void* feat_x, pn_stream_t stream) {
  return pn::point_taps_src(cam, pn::ray_src(rays), z, n_samples, xyz, viewdirs, points_per_object, feat_x, true, stream);
}

extern "C" int pn_point_taps(const pn_camera_t* cam, const float* rays, const float* z, int n_samples,
                             const float* xyz, const float* viewdirs, int64_t points_per_object,
                             void* feat_x, pn_stream_t stream) {
  return pn::point_taps_src(cam, pn::ray_src(rays), z, n_samples, xyz, viewdirs, points_per_object, feat_x, false, stream);
}

// pn_point_taps / pn_point_taps_grouped that also write the fp16 REMAINDER of the positional code: feat_x_lo (rows, 64)
// fp16 in the same row order as feat_x, x_lo = fp16(x - fp16(x)) in columns [0,42), zeros elsewhere -- with feat_x the
// fp16 pair (22 significand bits) that the precise chains' tensor-core lin_in multiplies.
extern "C" int pn_point_taps_f32(const pn_camera_t* cam, const float* rays, const float* z, int n_samples,
                                 const float* xyz, const float* viewdirs, int64_t points_per_object,
                                 void* feat_x, void* feat_x_lo, int grouped, pn_stream_t stream) {
  PN_CHECK_ARG(feat_x_lo && ((uintptr_t)feat_x_lo & 127) == 0, "pn_point_taps_f32: feat_x_lo missing or misaligned");
  return pn::point_taps_src(cam, pn::ray_src(rays), z, n_samples, xyz, viewdirs, points_per_object, feat_x, grouped != 0,
                            stream, feat_x_lo);
}
