// Ray sampling kernels: stratified coarse samples, inverse-CDF importance samples,
// depth samples and the merge sort (reference src/render/nerf.py:98-161, 285-295).
//
// Bit-exactness contract (oracle/sampling_np.py is the executable statement of it):
// every fp32 operation is a separately rounded IEEE op (__f*_rn intrinsics, which
// the compiler never contracts into FMAs), the row sum follows ATen's CPU order
// (8 lanes x 4 interleaved accumulators), the CDF is an fp64 running sum rounded to
// fp32 per element, and searchsorted(right=True) is an upper-bound search.
#include "pn_common.cuh"

namespace pn {

// pn_set_device_seed: a device-resident seed word that the sampling launches of THIS host thread add to their host
// seed (NULL = none).  A captured CUDA graph bakes kernel arguments in; with the seed in device memory a replayed
// training step still draws new jitter every time (the owner advances the word with a captured device-side add).
static thread_local const uint64_t* g_seed_dev = nullptr;
__device__ __forceinline__ uint64_t eff_seed(uint64_t seed, const uint64_t* seed_dev) {
  return seed_dev ? seed + __ldg(seed_dev) : seed;
}

__device__ __forceinline__ float lerp_depth(float near, float far, float zs, bool lindisp) {
  float a = __fsub_rn(1.0f, zs);
  if (!lindisp) return __fadd_rn(__fmul_rn(near, a), __fmul_rn(far, zs));
  float inv_n = __fdiv_rn(1.0f, near), inv_f = __fdiv_rn(1.0f, far);
  return __fdiv_rn(1.0f, __fadd_rn(__fmul_rn(inv_n, a), __fmul_rn(inv_f, zs)));
}

// One thread per 4 consecutive samples (one Philox call yields their 4 draws, one 16-byte store);
// consecutive threads write consecutive z: fully coalesced.  Kc % 4 == 0 takes this path.
// The kernel is instruction bound (Philox4x32-10 is ~70 of its ~110 instructions per thread), so `lindisp` is a
// template parameter (its three IEEE divisions per sample stay out of the linear instantiation) and the index split is
// 32-bit when the launch fits.
template <bool kLindisp>
__global__ void __launch_bounds__(256)
sample_coarse4_kernel(const RaySrc rays, const float* __restrict__ u,
                      const float* __restrict__ lin, float* __restrict__ z, int64_t total4, int kc,
                      float step, uint64_t seed, const uint64_t* __restrict__ seed_dev) {
  constexpr bool lindisp = kLindisp;
  seed = eff_seed(seed, seed_dev);
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= total4) return;
  const int64_t i = q * 4;
  int64_t r;
  int k;
  if (total4 <= 0x3fffffffll) {   // uniform
    const uint32_t kc4 = (uint32_t)kc >> 2, r32 = (uint32_t)q / kc4;
    r = r32;
    k = (int)(((uint32_t)q - r32 * kc4) << 2);
  } else {
    r = i / kc;
    k = (int)(i - r * kc);
  }
  float4 uu;
  if (u) {
    uu = *reinterpret_cast<const float4*>(u + i);
  } else {
    const uint4 rnd = Philox(seed)((uint64_t)q, (uint64_t)PN_STREAM_COARSE);  // same draws as rand_uniform(i..i+3)
    uu = make_float4(u01(rnd.x), u01(rnd.y), u01(rnd.z), u01(rnd.w));
  }
  const float4 ln = *reinterpret_cast<const float4*>(lin + k);
  const float2 nf = ray_near_far(rays, r);
  float4 o;
  o.x = lerp_depth(nf.x, nf.y, __fadd_rn(ln.x, __fmul_rn(uu.x, step)), lindisp);
  o.y = lerp_depth(nf.x, nf.y, __fadd_rn(ln.y, __fmul_rn(uu.y, step)), lindisp);
  o.z = lerp_depth(nf.x, nf.y, __fadd_rn(ln.z, __fmul_rn(uu.z, step)), lindisp);
  o.w = lerp_depth(nf.x, nf.y, __fadd_rn(ln.w, __fmul_rn(uu.w, step)), lindisp);
  *reinterpret_cast<float4*>(z + i) = o;
}

// General Kc: one thread per sample.
__global__ void __launch_bounds__(256)
sample_coarse_kernel(const RaySrc rays, const float* __restrict__ u,
                     const float* __restrict__ lin, float* __restrict__ z, int64_t total, int kc,
                     float step, bool lindisp, uint64_t seed, const uint64_t* __restrict__ seed_dev) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  seed = eff_seed(seed, seed_dev);
  int64_t r = i / kc;
  int k = (int)(i - r * kc);
  float uu = u ? u[i] : rand_uniform(seed, PN_STREAM_COARSE, (uint64_t)i);
  float zs = __fadd_rn(lin[k], __fmul_rn(uu, step));
  float2 nf = ray_near_far(rays, r);
  z[i] = lerp_depth(nf.x, nf.y, zs, lindisp);
}

// torch.sum(row) in ATen's CPU order, computed redundantly-free by one warp.
// Thread (j = lane/8, l = lane%8) owns accumulator j of vector lane l.
__device__ float aten_row_sum(const float* __restrict__ w, int k, int lane) {
  const int nv = k >> 3, full = nv >> 2;
  const int j = lane >> 3, l = lane & 7;
  float acc = 0.0f;
  for (int i = 0; i < full; ++i) acc = __fadd_rn(acc, w[((i << 2) + j) * 8 + l]);
  if (j == 0)
    for (int i = full << 2; i < nv; ++i) acc = __fadd_rn(acc, w[i * 8 + l]);
  // A[l] = ((acc0 + acc1) + acc2) + acc3, held by threads j == 0
  float a1 = __shfl_sync(0xffffffffu, acc, 8 + l);
  float a2 = __shfl_sync(0xffffffffu, acc, 16 + l);
  float a3 = __shfl_sync(0xffffffffu, acc, 24 + l);
  float a = __fadd_rn(__fadd_rn(__fadd_rn(acc, a1), a2), a3);
  // scalar tail (k % 8), same 4-way structure, summed first
  float s = 0.0f;
  const int nt = k & 7;
  if (nt) {
    const float* t = w + (nv << 3);
    float t0 = 0.f, t1 = 0.f, t2 = 0.f, t3 = 0.f;
    if (nt >= 4) { t0 = t[0]; t1 = t[1]; t2 = t[2]; t3 = t[3]; }
    for (int i = (nt >> 2) << 2; i < nt; ++i) t0 = __fadd_rn(t0, t[i]);
    s = __fadd_rn(__fadd_rn(__fadd_rn(t0, t1), t2), t3);
  }
#pragma unroll
  for (int q = 0; q < 8; ++q) s = __fadd_rn(s, __shfl_sync(0xffffffffu, a, q));
  return s;  // identical in every lane
}

template <int EPL>
__device__ __forceinline__ void sort_store(const float* __restrict__ srt, float* __restrict__ dst, int k2, int lane) {
  constexpr int N = 32 * EPL;
  float v[EPL];
#pragma unroll
  for (int j = 0; j < EPL; ++j) v[j] = srt[lane * EPL + j];
#pragma unroll
  for (int size = 2; size <= N; size <<= 1) {
#pragma unroll
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      if (stride >= EPL) {
        const int e = lane * EPL;   // bits >= log2(EPL) of the element index are the lane's
        const bool keep_min = ((e & stride) == 0) == ((e & size) == 0);
#pragma unroll
        for (int j = 0; j < EPL; ++j) {
          const float p = __shfl_xor_sync(0xffffffffu, v[j], stride / EPL);
          v[j] = keep_min ? fminf(v[j], p) : fmaxf(v[j], p);
        }
      } else {
#pragma unroll
        for (int j = 0; j < EPL; ++j) {
          if ((j & stride) == 0) {
            const int e = lane * EPL + j;
            const bool up = (e & size) == 0;
            const float a = v[j], b = v[j | stride];
            v[j] = up ? fminf(a, b) : fmaxf(a, b);
            v[j | stride] = up ? fmaxf(a, b) : fminf(a, b);
          }
        }
      }
    }
  }
  const int e0 = lane * EPL;
  if (EPL >= 4 && (k2 % 4) == 0 && ((uintptr_t)dst & 15) == 0) {
#pragma unroll
    for (int j = 0; j < EPL; j += 4)
      if (e0 + j < k2) *reinterpret_cast<float4*>(dst + e0 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
  } else {
#pragma unroll
    for (int j = 0; j < EPL; ++j)
      if (e0 + j < k2) dst[e0 + j] = v[j];
  }
}

constexpr int kFineWarps = 4;

// One warp per ray.  Shared memory per warp: w/pdf (Kc) + cdf (Kc+1) + sort buffer (pow2 >= K2).
#ifndef PN_FINE_MINB
#define PN_FINE_MINB 10   // 48 registers: 1.62 ms per 2^20 rays against 1.78 ms at 72 (tests/fine_sweep.sh); instruction bound either way
#endif
__global__ void __launch_bounds__(kFineWarps * 32, PN_FINE_MINB)
sample_fine_kernel(const RaySrc rays, const float* __restrict__ z_coarse,
                   const float* __restrict__ weights, const float* __restrict__ depth,
                   const float* __restrict__ u_bin, const float* __restrict__ u_in,
                   const float* __restrict__ noise, float* __restrict__ z_sorted,
                   float* __restrict__ z_imp, float* __restrict__ z_depth, int64_t n_rays, int kc,
                   int kf, int kd, int sort_n, float depth_std, bool lindisp, uint64_t seed,
                   const uint64_t* __restrict__ seed_dev) {
  extern __shared__ float smem[];
  seed = eff_seed(seed, seed_dev);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per_warp = kc + (kc + 1) + sort_n;
  float* w = smem + warp * per_warp;
  float* cdf = w + kc;
  float* srt = cdf + kc + 1;
  const int k2 = kc + kf + kd;

  for (int64_t r = (int64_t)blockIdx.x * kFineWarps + warp; r < n_rays; r += (int64_t)gridDim.x * kFineWarps) {
    const float2 nf = ray_near_far(rays, r);
    for (int k = lane; k < kc; k += 32) srt[k] = z_coarse[r * kc + k];
    if (kf > 0) {
      for (int k = lane; k < kc; k += 32) w[k] = __fadd_rn(weights[r * kc + k], 1e-5f);
      __syncwarp();
      const float s = aten_row_sum(w, kc, lane);
      for (int k = lane; k < kc; k += 32) w[k] = __fdiv_rn(w[k], s);  // pdf
      __syncwarp();
      // torch.cumsum on CPU: fp64 running sum, rounded to fp32 per element.  Done as a warp scan:
      // every pdf is a 24-bit value >= 2^-23 (w >= 1e-5, S < 2^7) and the sums stay < 2, so all
      // partial sums fit in fp64's 53 bits exactly and the order of additions cannot matter.
      {
        const int per = (kc + 31) >> 5, e0 = lane * per;
        double run = 0.0;
        for (int i = 0; i < per; ++i)
          if (e0 + i < kc) run += (double)w[e0 + i];
        double incl = run;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const double y = __shfl_up_sync(0xffffffffu, incl, d);
          if (lane >= d) incl += y;
        }
        double acc = incl - run;   // sum of the elements owned by lower lanes
        if (lane == 0) cdf[0] = 0.0f;
        for (int i = 0; i < per; ++i)
          if (e0 + i < kc) {
            acc += (double)w[e0 + i];
            cdf[e0 + i + 1] = (float)acc;
          }
      }
      __syncwarp();
      for (int q = lane; q < kf; q += 32) {
        const int64_t di = r * kf + q;
        float ub, ui;
        if (u_bin && u_in) {
          ub = u_bin[di];
          ui = u_in[di];
        } else {   // both draws of a sample from one Philox block (x: bin, y: position inside the bin)
          const uint4 rnd = Philox(seed)((uint64_t)di, (uint64_t)PN_STREAM_BIN + 16);
          ub = u_bin ? u_bin[di] : u01(rnd.x);
          ui = u_in ? u_in[di] : u01(rnd.y);
        }
        int lo = 0, hi = kc + 1;  // upper bound: first index with cdf > ub
        while (lo < hi) {
          int mid = (lo + hi) >> 1;
          if (cdf[mid] <= ub) lo = mid + 1; else hi = mid;
        }
        const float ind = fmaxf(__fsub_rn((float)lo, 1.0f), 0.0f);
        const float zs = __fdiv_rn(__fadd_rn(ind, ui), (float)kc);
        const float zz = lerp_depth(nf.x, nf.y, zs, lindisp);
        srt[kc + q] = zz;
        if (z_imp) z_imp[di] = zz;
      }
    }
    if (kd > 0) {
      const float d = depth[r];
      for (int q = lane; q < kd; q += 32) {
        const int64_t di = r * kd + q;
        const float nz = noise ? noise[di] : rand_normal(seed, PN_STREAM_NOISE, (uint64_t)di);
        float zz = __fadd_rn(d, __fmul_rn(nz, depth_std));
        zz = fmaxf(fminf(zz, nf.y), nf.x);
        srt[kc + kf + q] = zz;
        if (z_depth) z_depth[di] = zz;
      }
    }
    __syncwarp();
    // The stratified samples arrive sorted (nerf.py:98-118 is monotone in k) and at most 32 samples are new: sort
    // the new ones across the lanes (15 compare-exchange steps) and merge by rank -- each element's place is its
    // own index plus the number of elements of the other list in front of it (ties: stratified first) -- instead
    // of a full 128-wide bitonic sort (28 steps over 4 elements per lane).  Same multiset, so the same sorted row
    // bit for bit.  Anything else (unsorted z_coarse, more than 32 new samples, Kc outside [32,128]) takes the
    // general bitonic path below.
    const int n_new = kf + kd;
    bool merge_ok = n_new <= 32 && kc >= 32 && kc <= 128;
    if (merge_ok) {
      bool sorted = true;
      for (int k = lane; k + 1 < kc; k += 32) sorted = sorted && (srt[k] <= srt[k + 1]);
      merge_ok = __all_sync(0xffffffffu, sorted);
    }
    if (merge_ok) {
      float v = lane < n_new ? srt[kc + lane] : __int_as_float(0x7f800000);
#pragma unroll
      for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
          const float p = __shfl_xor_sync(0xffffffffu, v, stride);
          const bool keep_min = ((lane & stride) == 0) == ((lane & size) == 0 || size == 32);
          v = keep_min ? fminf(v, p) : fmaxf(v, p);
        }
      }
      w[lane] = v;            // the pdf buffer is free by now (Kc >= 32 floats)
      __syncwarp();
      int rank_new = 0;
      if (lane < n_new) {     // stratified samples <= v
        int lo = 0, hi = kc;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (srt[mid] <= v) lo = mid + 1; else hi = mid;
        }
        rank_new = lane + lo;
      }
      float cval[4];
      int crank[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = lane + 32 * j;
        cval[j] = 0.0f;
        crank[j] = -1;
        if (k < kc) {
          const float c = srt[k];
          int lo = 0, hi = n_new;   // new samples < c
          while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (w[mid] < c) lo = mid + 1; else hi = mid;
          }
          cval[j] = c;
          crank[j] = k + lo;
        }
      }
      __syncwarp();
      if (lane < n_new) srt[rank_new] = v;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (crank[j] >= 0) srt[crank[j]] = cval[j];
      __syncwarp();
      float* dst = z_sorted + r * k2;
      if ((k2 & 3) == 0 && ((uintptr_t)dst & 15) == 0) {
        for (int e = lane * 4; e < k2; e += 128)
          *reinterpret_cast<float4*>(dst + e) = make_float4(srt[e], srt[e + 1], srt[e + 2], srt[e + 3]);
      } else {
        for (int e = lane; e < k2; e += 32) dst[e] = srt[e];
      }
      __syncwarp();
      continue;
    }
    for (int k = k2 + lane; k < sort_n; k += 32) srt[k] = __int_as_float(0x7f800000);  // +inf padding
    __syncwarp();
    // bitonic sort of sort_n = 32*EPL floats: lane l holds elements [l*EPL, (l+1)*EPL) in registers;
    // strides >= EPL exchange through shuffles, smaller strides compare within the lane
    if (sort_n == 128) sort_store<4>(srt, z_sorted + r * k2, k2, lane);
    else if (sort_n == 64) sort_store<2>(srt, z_sorted + r * k2, k2, lane);
    else if (sort_n == 256) sort_store<8>(srt, z_sorted + r * k2, k2, lane);
    else if (sort_n == 512) sort_store<16>(srt, z_sorted + r * k2, k2, lane);
    else sort_store<1>(srt, z_sorted + r * k2, k2, lane);
    __syncwarp();
  }
}

// The entry points below with the rays coming from any RaySrc (an explicit tensor, or frames: pn_render_frames).
int sample_coarse_src(const RaySrc& rays, const float* u, const float* lin, float* z, int64_t n_rays, int n_coarse,
                      int lindisp, uint64_t seed, pn_stream_t stream) {
  if (n_rays == 0) return PN_OK;  // empty batches carry null data pointers
  PN_CHECK_ARG((rays.rays || rays.poses) && lin && z, "pn_sample_coarse: null pointer");
  PN_CHECK_ARG(n_rays >= 0 && n_coarse > 0, "pn_sample_coarse: bad sizes R=%lld Kc=%d", (long long)n_rays, n_coarse);
  if (n_rays == 0) return PN_OK;
  const int64_t total = n_rays * n_coarse;
  const float step = (float)(1.0 / (double)n_coarse);
  const bool vec = n_coarse % 4 == 0 && ((uintptr_t)z & 15) == 0 && ((uintptr_t)lin & 15) == 0 &&
                   (!u || ((uintptr_t)u & 15) == 0);
  if (vec && lindisp)
    pn::sample_coarse4_kernel<true><<<pn::ceil_div(total / 4, 256), 256, 0, (cudaStream_t)stream>>>(
        rays, u, lin, z, total / 4, n_coarse, step, seed, pn::g_seed_dev);
  else if (vec)
    pn::sample_coarse4_kernel<false><<<pn::ceil_div(total / 4, 256), 256, 0, (cudaStream_t)stream>>>(
        rays, u, lin, z, total / 4, n_coarse, step, seed, pn::g_seed_dev);
  else
    pn::sample_coarse_kernel<<<pn::ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(
        rays, u, lin, z, total, n_coarse, step, lindisp != 0, seed, pn::g_seed_dev);
  PN_LAUNCH_OK("sample_coarse_kernel");
  pn::count_launch();
  return PN_OK;
}

int sample_fine_src(const RaySrc& rays, const float* z_coarse, const float* weights, const float* depth,
                    const float* u_bin, const float* u_in, const float* noise, float* z_sorted, float* z_imp,
                    float* z_depth, int64_t n_rays, int n_coarse, int n_imp, int n_depth, float depth_std, int lindisp,
                    uint64_t seed, pn_stream_t stream) {
  if (n_rays == 0) return PN_OK;  // empty batches carry null data pointers
  PN_CHECK_ARG((rays.rays || rays.poses) && z_coarse && z_sorted, "pn_sample_fine: null pointer");
  PN_CHECK_ARG(n_imp == 0 || weights, "pn_sample_fine: importance samples need coarse weights");
  PN_CHECK_ARG(n_depth == 0 || depth, "pn_sample_fine: depth samples need coarse depth");
  PN_CHECK_ARG(n_rays >= 0 && n_coarse > 0 && n_imp >= 0 && n_depth >= 0, "pn_sample_fine: bad sizes");
  PN_CHECK_ARG(n_coarse < 512, "pn_sample_fine: Kc=%d >= 512 is outside the emulated ATen row-sum regime", n_coarse);
  if (n_rays == 0) return PN_OK;
  const int k2 = n_coarse + n_imp + n_depth;
  int sort_n = 32;
  while (sort_n < k2) sort_n <<= 1;
  PN_CHECK_ARG(sort_n <= 512, "pn_sample_fine: more than 512 samples per ray");
  const size_t smem = (size_t)pn::kFineWarps * (n_coarse + n_coarse + 1 + sort_n) * sizeof(float);
  PN_CHECK_ARG(smem <= 48 * 1024, "pn_sample_fine: K too large for shared memory");
  const int blocks = (int)((n_rays + pn::kFineWarps - 1) / pn::kFineWarps);
  pn::sample_fine_kernel<<<blocks < 148 * 16 ? blocks : 148 * 16, pn::kFineWarps * 32, smem, (cudaStream_t)stream>>>(
      rays, z_coarse, weights, depth, u_bin, u_in, noise, z_sorted, z_imp, z_depth, n_rays, n_coarse,
      n_imp, n_depth, sort_n, depth_std, lindisp != 0, seed, pn::g_seed_dev);
  PN_LAUNCH_OK("sample_fine_kernel");
  pn::count_launch();
  return PN_OK;
}

}  // namespace pn

extern "C" void pn_set_device_seed(const uint64_t* seed_dev) { pn::g_seed_dev = seed_dev; }

extern "C" int pn_sample_coarse(const float* rays, const float* u, const float* lin, float* z,
                                int64_t n_rays, int n_coarse, int lindisp, uint64_t seed,
                                pn_stream_t stream) {
  return pn::sample_coarse_src(pn::ray_src(rays), u, lin, z, n_rays, n_coarse, lindisp, seed, stream);
}

extern "C" int pn_sample_fine(const float* rays, const float* z_coarse, const float* weights,
                              const float* depth, const float* u_bin, const float* u_in,
                              const float* noise, float* z_sorted, float* z_imp, float* z_depth,
                              int64_t n_rays, int n_coarse, int n_imp, int n_depth, float depth_std,
                              int lindisp, uint64_t seed, pn_stream_t stream) {
  return pn::sample_fine_src(pn::ray_src(rays), z_coarse, weights, depth, u_bin, u_in, noise, z_sorted, z_imp, z_depth,
                             n_rays, n_coarse, n_imp, n_depth, depth_std, lindisp, seed, stream);
}
