// BatchNorm2d (training and evaluation) fused with ReLU, and bilinear x2 up-sampling (align_corners=True), on NHWC
// bf16 activations (sm_100a, 256-bit accesses, fp32 math, deterministic two-stage reductions).
//
// These are the kernels of the north-star model family (the upstream DoubleConv / Up blocks that the reference only
// documents in model/modelsummary.txt:153-247: Conv3x3(bias=False) -> BatchNorm2d -> ReLU, and
// nn.Upsample(scale_factor=2, mode='bilinear', align_corners=True)).  In the reference's environment they would be
// cuDNN batch-norm and ATen upsample_bilinear2d kernels.
//
//   forward   y_raw = conv(x)            (tcgen05 kernel, no bias)
//             (sum, sumsq) per channel   dpt_bn_stats          (two-stage, fixed order)
//             mean/var/invstd, running   dpt_bn_finalize       (also folds gamma/beta into scale/shift)
//             y = relu(y_raw*scale+shift) dpt_bn_apply_relu
//   backward  g = dy * (y > 0); sum g, sum g*xhat   dpt_bn_bwd_reduce
//             dz = scale * (g - (sum g + xhat * sum g*xhat) / N)   dpt_bn_bwd_apply  (+ dgamma, dbeta)
#include <atomic>

#include "../../include/dpt_b200.h"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dpt {

extern std::atomic<long long> g_launch_count;

constexpr int kBnBlocks = 296;

static inline int grid_for(long long work, int threads, int per_sm = 8) {
  long long b = (work + threads - 1) / threads;
  const long long cap = static_cast<long long>(sm_count()) * per_sm;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

// Per-thread BatchNorm parameters of one 16-channel group, loaded once: every thread of these kernels keeps ONE channel
// group for its whole life (the thread count of a grid is a multiple of the group count), so the per-element work is
// pure FMA on registers.  (The first version re-read four parameters per element through __ldg: 64 loads for every two
// data loads, which capped the apply passes at ~40 % of the HBM rate.)
__device__ __forceinline__ void load16(const float* __restrict__ src, float (&dst)[16]) {
  const float4* s4 = reinterpret_cast<const float4*>(src);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float4 v = __ldg(s4 + i);
    dst[4 * i + 0] = v.x;
    dst[4 * i + 1] = v.y;
    dst[4 * i + 2] = v.z;
    dst[4 * i + 3] = v.w;
  }
}

// ---- two-value per-channel reduction: thread owns 16 channels, blocks own pixel ranges, 4 pixels in flight per thread ----
// MODE 0: (x, x*x)                                  -> BatchNorm forward statistics
// MODE 1: (g, g*(x - mean)), g = dy * (x*scale+shift > 0)   -> BatchNorm backward (x = y_raw; the finalize multiplies
//         the second sum by invstd)
template <int MODE>
__global__ void __launch_bounds__(256, 2)
bn_reduce_stage1(const bf16* __restrict__ a, const bf16* __restrict__ b, const float* __restrict__ bn, float* __restrict__ partial,
                 long long P, int C) {
  grid_dep_sync();
  // bn (MODE 1): [4][C] = mean, invstd, scale, shift
  __shared__ float sh[256][33];
  const int lanes = C / 16;
  const int ppi = 256 / lanes;  // C <= 4096 is checked by the host; for C > 4096/16*... lanes <= 256
  const int lane = threadIdx.x % lanes;
  const int psub = threadIdx.x / lanes;
  const long long per_block = (P + gridDim.x - 1) / gridDim.x;
  const long long p0 = blockIdx.x * per_block;
  long long p1 = p0 + per_block;
  if (p1 > P) p1 = P;
  float s0[16], s1[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) s0[j] = s1[j] = 0.f;
  float mean[16], scale[16], shift[16];
  if (MODE == 1) {
    load16(bn + lane * 16, mean);
    load16(bn + 2 * C + lane * 16, scale);
    load16(bn + 3 * C + lane * 16, shift);
  }
  auto accumulate = [&](const u32x8& va, const u32x8& vb) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float x0 = bf16lo(va.v[k]), x1 = bf16hi(va.v[k]);
      if (MODE == 0) {
        s0[2 * k] += x0;
        s0[2 * k + 1] += x1;
        s1[2 * k] = fmaf(x0, x0, s1[2 * k]);
        s1[2 * k + 1] = fmaf(x1, x1, s1[2 * k + 1]);
      } else {
        float g0 = bf16lo(vb.v[k]), g1 = bf16hi(vb.v[k]);
        if (!(fmaf(x0, scale[2 * k], shift[2 * k]) > 0.f)) g0 = 0.f;
        if (!(fmaf(x1, scale[2 * k + 1], shift[2 * k + 1]) > 0.f)) g1 = 0.f;
        s0[2 * k] += g0;
        s0[2 * k + 1] += g1;
        s1[2 * k] = fmaf(g0, x0 - mean[2 * k], s1[2 * k]);
        s1[2 * k + 1] = fmaf(g1, x1 - mean[2 * k + 1], s1[2 * k + 1]);
      }
    }
  };
  if (psub < ppi) {
    long long p = p0 + psub;
    const long long step = ppi;
    for (; p + 3 * step < p1; p += 4 * step) {
      u32x8 va[4], vb[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        va[u] = ldg256(a + (p + u * step) * C + lane * 16);
        if (MODE == 1) vb[u] = ldg256(b + (p + u * step) * C + lane * 16);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) accumulate(va[u], vb[u]);
    }
    for (; p < p1; p += step) {
      const u32x8 va = ldg256(a + p * C + lane * 16);
      u32x8 vb = va;
      if (MODE == 1) vb = ldg256(b + p * C + lane * 16);
      accumulate(va, vb);
    }
  }
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    sh[threadIdx.x][j] = s0[j];
    sh[threadIdx.x][16 + j] = s1[j];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += 256) {
    const int ln = c >> 4, j = c & 15;
    float r0 = 0.f, r1 = 0.f;
    for (int k = 0; k < ppi; ++k) {
      r0 += sh[k * lanes + ln][j];
      r1 += sh[k * lanes + ln][16 + j];
    }
    partial[(static_cast<size_t>(blockIdx.x) * 2) * C + c] = r0;
    partial[(static_cast<size_t>(blockIdx.x) * 2 + 1) * C + c] = r1;
  }
}

// Fixed-order sum of the partial pairs of channel c.  The rows of a channel are first + k * stride, k < nrows, with
// first = c / nt and stride = C / nt: a plain [rows][2][C] array is nt = C; the statistics epilogue of the convolution
// kernel writes one row per CTA and a CTA only covers the nt channels of its N tile (igemm_conv.cu).  8 threads per
// channel walk every 8th row, then an 8-way sum in shared memory (one thread walking ~300 partials serialised on L2
// latency).  Launch with 256 threads = 32 channels x 8 lanes; returns the sums to the lane-0 thread of each channel.
__device__ __forceinline__ bool bn_sum_partials(const float* __restrict__ partial, int nrows, int nt, int C, int& c, float& s0,
                                                float& s1) {
  __shared__ float sh[2][8][33];
  const int o = threadIdx.x & 31, sl = threadIdx.x >> 5;
  c = blockIdx.x * 32 + o;
  float a0 = 0.f, a1 = 0.f;
  if (c < C) {
    const int first = c / nt, stride = C / nt;
    for (int k = sl; k < nrows; k += 8) {
      const size_t row = static_cast<size_t>(first + k * stride);
      a0 += partial[(row * 2) * C + c];
      a1 += partial[(row * 2 + 1) * C + c];
    }
  }
  sh[0][sl][o] = a0;
  sh[1][sl][o] = a1;
  __syncthreads();
  s0 = s1 = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    s0 += sh[0][k][o];
    s1 += sh[1][k][o];
  }
  return sl == 0 && c < C;
}

// Forward finalize: mean / biased var -> invstd; running stats (momentum, unbiased var); scale / shift.
// bn: [4][C] = mean, invstd, scale, shift.
__global__ void bn_finalize_kernel(const float* __restrict__ partial, int nrows, int nt, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float* __restrict__ running_mean,
                                   float* __restrict__ running_var, float* __restrict__ bn, float* __restrict__ sums_out,
                                   int C, float count, float eps, float momentum) {
  grid_dep_sync();
  int c;
  float s0, s1;
  if (!bn_sum_partials(partial, nrows, nt, C, c, s0, s1)) return;
  if (sums_out != nullptr) {
    sums_out[c] = s0;
    sums_out[C + c] = s1;
  }
  const float mean = s0 / count;
  const float var = fmaxf(s1 / count - mean * mean, 0.f);
  const float invstd = rsqrtf(var + eps);
  const float sc = gamma[c] * invstd;
  bn[c] = mean;
  bn[C + c] = invstd;
  bn[2 * C + c] = sc;
  bn[3 * C + c] = beta[c] - mean * sc;
  if (running_mean != nullptr) {
    const float unbiased = count > 1.f ? var * count / (count - 1.f) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
  }
}

// Evaluation mode: scale / shift from the running statistics.
__global__ void bn_eval_prepare_kernel(const float* __restrict__ gamma, const float* __restrict__ beta,
                                       const float* __restrict__ running_mean, const float* __restrict__ running_var,
                                       float* __restrict__ bn, int C, float eps) {
  grid_dep_sync();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float invstd = rsqrtf(running_var[c] + eps);
  const float sc = gamma[c] * invstd;
  bn[c] = running_mean[c];
  bn[C + c] = invstd;
  bn[2 * C + c] = sc;
  bn[3 * C + c] = beta[c] - running_mean[c] * sc;
}

// y = relu(x*scale + shift) over 2x2 pixel quads (four 256-bit loads in flight per thread); POOL also writes
// nn.MaxPool2d(2,2) of y, so the Down block's pooling costs a quarter-size store instead of a read + write pass.
template <bool POOL>
__global__ void __launch_bounds__(256)
bn_apply_quad_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, bf16* __restrict__ pooled, const float* __restrict__ bn,
                     int B, int H, int W, int C, int relu) {
  grid_dep_sync();
  const int G = C / 16, Hq = H >> 1, Wq = W >> 1;
  const long long total = static_cast<long long>(B) * Hq * Wq * G;
  const long long tid0 = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  const long long T = static_cast<long long>(gridDim.x) * blockDim.x;  // multiple of 256, hence of G
  const int g = static_cast<int>(tid0 % G);
  float sc[16], sf[16];
  load16(bn + 2 * C + g * 16, sc);
  load16(bn + 3 * C + g * 16, sf);
  const long long rowstride = static_cast<long long>(W) * C;
  for (long long i = tid0; i < total; i += T) {
    long long r = i / G;
    const int qx = static_cast<int>(r % Wq);
    r /= Wq;
    const int qy = static_cast<int>(r % Hq);
    const long long b = r / Hq;
    const long long base = ((b * H + 2 * qy) * W + 2 * qx) * C + g * 16;
    u32x8 v[4];
    v[0] = ldg256(x + base);
    v[1] = ldg256(x + base + C);
    v[2] = ldg256(x + base + rowstride);
    v[3] = ldg256(x + base + rowstride + C);
    u32x8 o[4], om;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float mlo = -INFINITY, mhi = -INFINITY;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float lo = fmaf(bf16lo(v[u].v[k]), sc[2 * k], sf[2 * k]);
        float hi = fmaf(bf16hi(v[u].v[k]), sc[2 * k + 1], sf[2 * k + 1]);
        if (relu) {
          lo = fmaxf(lo, 0.f);
          hi = fmaxf(hi, 0.f);
        }
        o[u].v[k] = pack_bf16x2(lo, hi);
        mlo = fmaxf(mlo, lo);
        mhi = fmaxf(mhi, hi);
      }
      om.v[k] = pack_bf16x2(mlo, mhi);
    }
    stg256(y + base, o[0]);
    stg256(y + base + C, o[1]);
    stg256(y + base + rowstride, o[2]);
    stg256(y + base + rowstride + C, o[3]);
    if (POOL) stg256(pooled + ((b * Hq + qy) * Wq + qx) * C + g * 16, om);
  }
}

// The same without the pooled output on a flat pixel range (any pixel count): four pixels in flight per thread.
__global__ void __launch_bounds__(256)
bn_apply_flat_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, const float* __restrict__ bn, long long P, int C, int relu) {
  grid_dep_sync();
  const int G = C / 16;
  const long long total = P * G;
  const long long tid0 = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  const long long T = static_cast<long long>(gridDim.x) * blockDim.x;
  const int g = static_cast<int>(tid0 % G);
  float sc[16], sf[16];
  load16(bn + 2 * C + g * 16, sc);
  load16(bn + 3 * C + g * 16, sf);
  auto one = [&](const u32x8& v, long long i) {
    u32x8 o;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float lo = fmaf(bf16lo(v.v[k]), sc[2 * k], sf[2 * k]);
      float hi = fmaf(bf16hi(v.v[k]), sc[2 * k + 1], sf[2 * k + 1]);
      if (relu) {
        lo = fmaxf(lo, 0.f);
        hi = fmaxf(hi, 0.f);
      }
      o.v[k] = pack_bf16x2(lo, hi);
    }
    stg256(y + i * 16, o);
  };
  long long i = tid0;
  for (; i + 3 * T < total; i += 4 * T) {
    u32x8 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = ldg256(x + (i + u * T) * 16);
#pragma unroll
    for (int u = 0; u < 4; ++u) one(v[u], i + u * T);
  }
  for (; i < total; i += T) one(ldg256(x + i * 16), i);
}

// Backward finalize: dbeta = sum g, dgamma = sum g*xhat = invstd * sum g*(x - mean); and the per-channel coefficients of
// dz = scale*g + cb*x + ca  (= scale * (g - (dbeta + xhat*dgamma)/N)),  cb = -scale*dgamma*invstd/N,
// ca = -scale*dbeta/N - cb*mean.  red: [2][C] = cb, ca.
__global__ void bn_bwd_finalize_kernel(const float* __restrict__ partial, int nrows, const float* __restrict__ bn,
                                       float* __restrict__ red, float* __restrict__ dgamma, float* __restrict__ dbeta, int C,
                                       float inv_count, int accumulate) {
  grid_dep_sync();
  int c;
  float s0, s1;
  if (!bn_sum_partials(partial, nrows, C, C, c, s0, s1)) return;
  const float mean = bn[c], invstd = bn[C + c], sc = bn[2 * C + c];
  const float dg = s1 * invstd;
  const float cb = -sc * dg * invstd * inv_count;
  red[c] = cb;
  red[C + c] = -sc * s0 * inv_count - cb * mean;
  if (dbeta != nullptr) dbeta[c] = accumulate ? dbeta[c] + s0 : s0;
  if (dgamma != nullptr) dgamma[c] = accumulate ? dgamma[c] + dg : dg;
}

// dz = scale * g + cb * x + ca with g = dy * (x*scale + shift > 0); two pixels (four loads) in flight per thread.
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(const bf16* __restrict__ y_raw, const bf16* __restrict__ dy, bf16* __restrict__ dz,
                    const float* __restrict__ bn, const float* __restrict__ red, long long P, int C) {
  grid_dep_sync();
  const int G = C / 16;
  const long long total = P * G;
  const long long tid0 = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  const long long T = static_cast<long long>(gridDim.x) * blockDim.x;
  const int g = static_cast<int>(tid0 % G);
  float sc[16], sf[16], cb[16], ca[16];
  load16(bn + 2 * C + g * 16, sc);
  load16(bn + 3 * C + g * 16, sf);
  load16(red + g * 16, cb);
  load16(red + C + g * 16, ca);
  auto one = [&](const u32x8& vx, const u32x8& vg, long long i) {
    u32x8 o;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float x0 = bf16lo(vx.v[k]), x1 = bf16hi(vx.v[k]);
      float g0 = bf16lo(vg.v[k]), g1 = bf16hi(vg.v[k]);
      if (!(fmaf(x0, sc[2 * k], sf[2 * k]) > 0.f)) g0 = 0.f;
      if (!(fmaf(x1, sc[2 * k + 1], sf[2 * k + 1]) > 0.f)) g1 = 0.f;
      o.v[k] = pack_bf16x2(fmaf(sc[2 * k], g0, fmaf(cb[2 * k], x0, ca[2 * k])),
                           fmaf(sc[2 * k + 1], g1, fmaf(cb[2 * k + 1], x1, ca[2 * k + 1])));
    }
    stg256(dz + i * 16, o);
  };
  long long i = tid0;
  for (; i + T < total; i += 2 * T) {
    const u32x8 x0 = ldg256(y_raw + i * 16), x1 = ldg256(y_raw + (i + T) * 16);
    const u32x8 g0 = ldg256(dy + i * 16), g1 = ldg256(dy + (i + T) * 16);
    one(x0, g0, i);
    one(x1, g1, i + T);
  }
  if (i < total) one(ldg256(y_raw + i * 16), ldg256(dy + i * 16), i);
}

// ---- bilinear x2, align_corners=True -------------------------------------------------------------------
// forward: out[b, y, x, :] = lerp of in at (y*(H-1)/(2H-1), x*(W-1)/(2W-1))
__global__ void bilinear_up2x_fwd_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, int B, int H, int W, int C) {
  grid_dep_sync();
  const int Ho = 2 * H, Wo = 2 * W, G = C / 16;
  const float ry = Ho > 1 ? static_cast<float>(H - 1) / (Ho - 1) : 0.f;
  const float rx = Wo > 1 ? static_cast<float>(W - 1) / (Wo - 1) : 0.f;
  const long long total = static_cast<long long>(B) * Ho * Wo * G;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % G);
    long long r = i / G;
    const int xo = static_cast<int>(r % Wo);
    r /= Wo;
    const int yo = static_cast<int>(r % Ho);
    const int b = static_cast<int>(r / Ho);
    const float sy = yo * ry, sx = xo * rx;
    const int y0 = static_cast<int>(sy), x0 = static_cast<int>(sx);
    const int y1 = y0 + 1 < H ? y0 + 1 : H - 1, x1 = x0 + 1 < W ? x0 + 1 : W - 1;
    const float ly = sy - y0, lx = sx - x0;
    const bf16* base = in + static_cast<long long>(b) * H * W * C + g * 16;
    const u32x8 v00 = ldg256(base + (static_cast<long long>(y0) * W + x0) * C);
    const u32x8 v01 = ldg256(base + (static_cast<long long>(y0) * W + x1) * C);
    const u32x8 v10 = ldg256(base + (static_cast<long long>(y1) * W + x0) * C);
    const u32x8 v11 = ldg256(base + (static_cast<long long>(y1) * W + x1) * C);
    const float w00 = (1.f - ly) * (1.f - lx), w01 = (1.f - ly) * lx, w10 = ly * (1.f - lx), w11 = ly * lx;
    u32x8 o;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float a = w00 * bf16lo(v00.v[k]) + w01 * bf16lo(v01.v[k]) + w10 * bf16lo(v10.v[k]) + w11 * bf16lo(v11.v[k]);
      const float c = w00 * bf16hi(v00.v[k]) + w01 * bf16hi(v01.v[k]) + w10 * bf16hi(v10.v[k]) + w11 * bf16hi(v11.v[k]);
      o.v[k] = pack_bf16x2(a, c);
    }
    stg256(out + i * 16, o);
  }
}

// backward as a gather (deterministic): din[b, yi, xi] = sum over output pixels whose support touches (yi, xi)
__global__ void bilinear_up2x_bwd_kernel(const bf16* __restrict__ dout, bf16* __restrict__ din, int B, int H, int W, int C,
                                         int relu_mask_unused) {
  grid_dep_sync();
  const int Ho = 2 * H, Wo = 2 * W, G = C / 16;
  const float ry = Ho > 1 ? static_cast<float>(H - 1) / (Ho - 1) : 0.f;
  const float rx = Wo > 1 ? static_cast<float>(W - 1) / (Wo - 1) : 0.f;
  const long long total = static_cast<long long>(B) * H * W * G;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % G);
    long long r = i / G;
    const int xi = static_cast<int>(r % W);
    r /= W;
    const int yi = static_cast<int>(r % H);
    const int b = static_cast<int>(r / H);
    // output rows yo with |yo*ry - yi| < 1  (ry ~ 0.5 -> at most 4 candidates); same for columns
    int ylo = ry > 0.f ? static_cast<int>(ceilf((yi - 1) / ry)) : 0, yhi = ry > 0.f ? static_cast<int>(floorf((yi + 1) / ry)) : Ho - 1;
    int xlo = rx > 0.f ? static_cast<int>(ceilf((xi - 1) / rx)) : 0, xhi = rx > 0.f ? static_cast<int>(floorf((xi + 1) / rx)) : Wo - 1;
    if (ylo < 0) ylo = 0;
    if (xlo < 0) xlo = 0;
    if (yhi > Ho - 1) yhi = Ho - 1;
    if (xhi > Wo - 1) xhi = Wo - 1;
    float acc[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) acc[j] = 0.f;
    const bf16* base = dout + static_cast<long long>(b) * Ho * Wo * C + g * 16;
    for (int yo = ylo; yo <= yhi; ++yo) {
      // weight of input row yi in output row yo, replicating the forward's index clamping exactly
      const float sy = yo * ry;
      const int y0 = static_cast<int>(sy);
      const int y1 = y0 + 1 < H ? y0 + 1 : H - 1;
      const float ly = sy - y0;
      const float wy = (y0 == yi ? 1.f - ly : 0.f) + (y1 == yi ? ly : 0.f);
      if (wy == 0.f) continue;
      for (int xo = xlo; xo <= xhi; ++xo) {
        const float sx = xo * rx;
        const int x0 = static_cast<int>(sx);
        const int x1 = x0 + 1 < W ? x0 + 1 : W - 1;
        const float lx = sx - x0;
        const float wx = (x0 == xi ? 1.f - lx : 0.f) + (x1 == xi ? lx : 0.f);
        if (wx == 0.f) continue;
        const float w = wy * wx;
        const u32x8 v = ldg256(base + (static_cast<long long>(yo) * Wo + xo) * C);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          acc[2 * k] += w * bf16lo(v.v[k]);
          acc[2 * k + 1] += w * bf16hi(v.v[k]);
        }
      }
    }
    u32x8 o;
#pragma unroll
    for (int k = 0; k < 8; ++k) o.v[k] = pack_bf16x2(acc[2 * k], acc[2 * k + 1]);
    stg256(din + i * 16, o);
  }
}

}  // namespace dpt

using namespace dpt;

extern "C" size_t dpt_bn_scratch_bytes(int c) { return static_cast<size_t>(kBnBlocks) * 2 * c * sizeof(float); }

static int bn_check(int c, size_t scratch_bytes) {
  if (c % 16 || c <= 0 || c > 4096 || 256 % (c / 16 > 256 ? 256 : c / 16))
    return fail("batchnorm: c=%d must be 16 times a power of two (<= 4096)", c);
  if (scratch_bytes < dpt_bn_scratch_bytes(c)) return fail("batchnorm: scratch too small");
  return 0;
}

static int bn_apply_launch(const void* y_raw, void* y, void* pooled, const float* bn, int batch, int h, int w, int c, int relu,
                           cudaStream_t st) {
  if (c % 16 || c <= 0 || 256 % (c / 16)) return fail("batchnorm apply: c=%d must be 16 times a power of two (<= 4096)", c);
  if (pooled != nullptr || ((h | w) & 1) == 0) {
    if ((h | w) & 1) return fail("batchnorm apply with a pooled output: h=%d, w=%d must be even", h, w);
    const long long work = static_cast<long long>(batch) * (h / 2) * (w / 2) * (c / 16);
    const int grid = grid_for(work, 256);
    if (pooled != nullptr) {
      DPT_CUDA_OK(launch_chained(bn_apply_quad_kernel<true>, dim3(grid), dim3(256), 0, st, static_cast<const bf16*>(y_raw), static_cast<bf16*>(y),
                                                      static_cast<bf16*>(pooled), bn, batch, h, w, c, relu));
    } else {
      DPT_CUDA_OK(launch_chained(bn_apply_quad_kernel<false>, dim3(grid), dim3(256), 0, st, static_cast<const bf16*>(y_raw), static_cast<bf16*>(y), nullptr, bn,
                                                       batch, h, w, c, relu));
    }
  } else {  // odd extents (only possible without a pooled output): flat walk over the pixels
    const long long pixels = static_cast<long long>(batch) * h * w;
    DPT_CUDA_OK(launch_chained(bn_apply_flat_kernel, dim3(grid_for((pixels * (c / 16) + 3) / 4, 256)), dim3(256), 0, st, static_cast<const bf16*>(y_raw),
                                                                                   static_cast<bf16*>(y), bn, pixels, c, relu));
  }
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_bn_finalize(const float* partial, int rows_per_tile, int nt, const float* gamma, const float* beta,
                               float* running_mean, float* running_var, float* bn, long long pixels, int c, float eps,
                               float momentum, void* stream) {
  if (c <= 0 || nt <= 0 || c % nt || rows_per_tile <= 0) return fail("bn_finalize: bad geometry c=%d nt=%d rows=%d", c, nt, rows_per_tile);
  DPT_CUDA_OK(launch_chained(bn_finalize_kernel, dim3((c + 31) / 32), dim3(256), 0, static_cast<cudaStream_t>(stream), partial, rows_per_tile, nt, gamma, beta, running_mean, running_var, bn, nullptr, c, static_cast<float>(pixels), eps, momentum));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_bn_apply(const void* y_raw, void* y, void* pooled, const float* bn, int batch, int h, int w, int c, int relu,
                            void* stream) {
  return bn_apply_launch(y_raw, y, pooled, bn, batch, h, w, c, relu, static_cast<cudaStream_t>(stream));
}

extern "C" int dpt_bn_forward_train(const void* y_raw, void* y, const float* gamma, const float* beta, float* running_mean,
                                    float* running_var, float* bn, float* sums_out, long long pixels, int c, float eps,
                                    float momentum, int relu, void* scratch, size_t scratch_bytes, void* stream) {
  if (bn_check(c, scratch_bytes)) return -1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* partial = static_cast<float*>(scratch);
  DPT_CUDA_OK(launch_chained(bn_reduce_stage1<0>, dim3(kBnBlocks), dim3(256), 0, st, static_cast<const bf16*>(y_raw), nullptr, nullptr, partial, pixels, c));
  DPT_CUDA_OK(launch_chained(bn_finalize_kernel, dim3((c + 31) / 32), dim3(256), 0, st, partial, kBnBlocks, c, gamma, beta, running_mean, running_var, bn, sums_out, c,
                                                      static_cast<float>(pixels), eps, momentum));
  g_launch_count.fetch_add(2, std::memory_order_relaxed);
  // no pooled output: any factorisation of `pixels` is the same math (odd extents take the flat kernel)
  return bn_apply_launch(y_raw, y, nullptr, bn, 1, pixels % 4 ? 1 : 2, static_cast<int>(pixels % 4 ? pixels : pixels / 2), c, relu, st);
}

extern "C" int dpt_bn_forward_eval(const void* y_raw, void* y, const float* gamma, const float* beta, const float* running_mean,
                                   const float* running_var, float* bn, long long pixels, int c, float eps, int relu,
                                   void* stream) {
  if (c % 16 || c <= 0) return fail("batchnorm: c=%d must be a multiple of 16", c);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DPT_CUDA_OK(launch_chained(bn_eval_prepare_kernel, dim3((c + 127) / 128), dim3(128), 0, st, gamma, beta, running_mean, running_var, bn, c, eps));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  return bn_apply_launch(y_raw, y, nullptr, bn, 1, pixels % 4 ? 1 : 2, static_cast<int>(pixels % 4 ? pixels : pixels / 2), c, relu, st);
}

extern "C" int dpt_bn_eval_prepare(const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                                   float* bn, int c, float eps, void* stream) {
  DPT_CUDA_OK(launch_chained(bn_eval_prepare_kernel, dim3((c + 127) / 128), dim3(128), 0, static_cast<cudaStream_t>(stream), gamma, beta, running_mean, running_var,
                                                                                         bn, c, eps));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_bn_backward(const void* y_raw, const void* dy, void* dz, const float* bn, float* dgamma, float* dbeta,
                               long long pixels, int c, int accumulate, void* scratch, size_t scratch_bytes, void* stream) {
  if (bn_check(c, scratch_bytes)) return -1;
  if (scratch_bytes < dpt_bn_scratch_bytes(c) + 2 * c * sizeof(float)) return fail("batchnorm backward: scratch too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* partial = static_cast<float*>(scratch);
  float* red = partial + static_cast<size_t>(kBnBlocks) * 2 * c;
  DPT_CUDA_OK(launch_chained(bn_reduce_stage1<1>, dim3(kBnBlocks), dim3(256), 0, st, static_cast<const bf16*>(y_raw), static_cast<const bf16*>(dy), bn, partial, pixels, c));
  DPT_CUDA_OK(launch_chained(bn_bwd_finalize_kernel, dim3((c + 31) / 32), dim3(256), 0, st, partial, kBnBlocks, bn, red, dgamma, dbeta, c,
                                                          1.f / static_cast<float>(pixels), accumulate));
  DPT_CUDA_OK(launch_chained(bn_bwd_apply_kernel, dim3(grid_for(pixels * (c / 16), 256)), dim3(256), 0, st, static_cast<const bf16*>(y_raw), static_cast<const bf16*>(dy),
                                                                        static_cast<bf16*>(dz), bn, red, pixels, c));
  g_launch_count.fetch_add(3, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_bilinear_up2x_fwd(const void* in, void* out, int batch, int h, int w, int c, void* stream) {
  if (c % 16) return fail("bilinear: c=%d must be a multiple of 16", c);
  const long long work = static_cast<long long>(batch) * 4 * h * w * (c / 16);
  DPT_CUDA_OK(launch_chained(bilinear_up2x_fwd_kernel, dim3(grid_for(work, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), static_cast<const bf16*>(in), static_cast<bf16*>(out), batch, h, w, c));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_bilinear_up2x_bwd(const void* dout, void* din, int batch, int h, int w, int c, void* stream) {
  if (c % 16) return fail("bilinear: c=%d must be a multiple of 16", c);
  const long long work = static_cast<long long>(batch) * h * w * (c / 16);
  DPT_CUDA_OK(launch_chained(bilinear_up2x_bwd_kernel, dim3(grid_for(work, 128)), dim3(128), 0, static_cast<cudaStream_t>(stream), static_cast<const bf16*>(dout), static_cast<bf16*>(din), batch, h, w, c, 0));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}
