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

// ---- two-value per-channel reduction skeleton: thread owns 16 channels of a pixel, blocks own pixel ranges ----
// MODE 0: (x, x*x)                                  -> BatchNorm forward statistics
// MODE 1: (g, g*xhat), g = dy * (y_raw*scale+shift > 0), xhat = (y_raw - mean) * invstd   -> BatchNorm backward
template <int MODE>
__global__ void __launch_bounds__(256)
bn_reduce_stage1(const bf16* __restrict__ a, const bf16* __restrict__ b, const float* __restrict__ bn, float* __restrict__ partial,
                 long long P, int C) {
  // bn (MODE 1): [4][C] = mean, invstd, scale, shift
  __shared__ float sh[256][33];
  const int lanes = C / 16;
  const int ppi = 256 / lanes;
  const int lane = threadIdx.x % lanes;
  const int psub = threadIdx.x / lanes;
  const long long per_block = (P + gridDim.x - 1) / gridDim.x;
  const long long p0 = blockIdx.x * per_block;
  long long p1 = p0 + per_block;
  if (p1 > P) p1 = P;
  float s0[16], s1[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) s0[j] = s1[j] = 0.f;
  float mean[16], invstd[16], scale[16], shift[16];
  if (MODE == 1) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int c = lane * 16 + j;
      mean[j] = bn[c];
      invstd[j] = bn[C + c];
      scale[j] = bn[2 * C + c];
      shift[j] = bn[3 * C + c];
    }
  }
  if (psub < ppi) {
    for (long long p = p0 + psub; p < p1; p += ppi) {
      const u32x8 va = ldg256(a + p * C + lane * 16);
      u32x8 vb;
      if (MODE == 1) vb = ldg256(b + p * C + lane * 16);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float x0 = bf16lo(va.v[k]), x1 = bf16hi(va.v[k]);
        if (MODE == 0) {
          s0[2 * k] += x0;
          s0[2 * k + 1] += x1;
          s1[2 * k] += x0 * x0;
          s1[2 * k + 1] += x1 * x1;
        } else {
          // a = y_raw, b = dy
          float g0 = bf16lo(vb.v[k]), g1 = bf16hi(vb.v[k]);
          if (!(x0 * scale[2 * k] + shift[2 * k] > 0.f)) g0 = 0.f;
          if (!(x1 * scale[2 * k + 1] + shift[2 * k + 1] > 0.f)) g1 = 0.f;
          s0[2 * k] += g0;
          s0[2 * k + 1] += g1;
          s1[2 * k] += g0 * (x0 - mean[2 * k]) * invstd[2 * k];
          s1[2 * k + 1] += g1 * (x1 - mean[2 * k + 1]) * invstd[2 * k + 1];
        }
      }
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

// Fixed-order sum of the per-block partial pairs of channel c: 8 threads per channel walk every 8th block, then an 8-way
// sum in shared memory (one thread walking all ~300 partials serialised on L2 latency: ~20 us per BatchNorm layer and
// direction).  Launch with 256 threads = 32 channels x 8 lanes; returns the sums to the lane-0 thread of each channel.
__device__ __forceinline__ bool bn_sum_partials(const float* __restrict__ partial, int nblocks, int C, int& c, float& s0,
                                                float& s1) {
  __shared__ float sh[2][8][33];
  const int o = threadIdx.x & 31, sl = threadIdx.x >> 5;
  c = blockIdx.x * 32 + o;
  float a0 = 0.f, a1 = 0.f;
  if (c < C) {
    for (int b = sl; b < nblocks; b += 8) {
      a0 += partial[(static_cast<size_t>(b) * 2) * C + c];
      a1 += partial[(static_cast<size_t>(b) * 2 + 1) * C + c];
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
// bn: [4][C] = mean, invstd, scale, shift.  sums: [2][C] (already reduced over blocks, possibly over ranks).
__global__ void bn_finalize_kernel(const float* __restrict__ partial, int nblocks, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float* __restrict__ running_mean,
                                   float* __restrict__ running_var, float* __restrict__ bn, float* __restrict__ sums_out,
                                   int C, float count, float eps, float momentum) {
  int c;
  float s0, s1;
  if (!bn_sum_partials(partial, nblocks, C, c, s0, s1)) return;
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
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float invstd = rsqrtf(running_var[c] + eps);
  const float sc = gamma[c] * invstd;
  bn[c] = running_mean[c];
  bn[C + c] = invstd;
  bn[2 * C + c] = sc;
  bn[3 * C + c] = beta[c] - running_mean[c] * sc;
}

__global__ void bn_apply_relu_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, const float* __restrict__ bn,
                                     long long P, int C, int relu) {
  const int G = C / 16;
  const long long total = P * G;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % G);
    const u32x8 v = ldg256(x + i * 16);
    u32x8 o;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int c = g * 16 + 2 * k;
      float a = bf16lo(v.v[k]) * __ldg(bn + 2 * C + c) + __ldg(bn + 3 * C + c);
      float b = bf16hi(v.v[k]) * __ldg(bn + 2 * C + c + 1) + __ldg(bn + 3 * C + c + 1);
      if (relu) {
        a = fmaxf(a, 0.f);
        b = fmaxf(b, 0.f);
      }
      o.v[k] = pack_bf16x2(a, b);
    }
    stg256(y + i * 16, o);
  }
}

// Backward finalize + apply: dgamma = sum g*xhat, dbeta = sum g; dz = scale * (g - (dbeta + xhat*dgamma)/N)
__global__ void bn_bwd_finalize_kernel(const float* __restrict__ partial, int nblocks, float* __restrict__ red,
                                       float* __restrict__ dgamma, float* __restrict__ dbeta, int C, int accumulate) {
  int c;
  float s0, s1;
  if (!bn_sum_partials(partial, nblocks, C, c, s0, s1)) return;
  red[c] = s0;
  red[C + c] = s1;
  if (dbeta != nullptr) dbeta[c] = accumulate ? dbeta[c] + s0 : s0;
  if (dgamma != nullptr) dgamma[c] = accumulate ? dgamma[c] + s1 : s1;
}

__global__ void bn_bwd_apply_kernel(const bf16* __restrict__ y_raw, const bf16* __restrict__ dy, bf16* __restrict__ dz,
                                    const float* __restrict__ bn, const float* __restrict__ red, long long P, int C,
                                    float inv_count) {
  const int G = C / 16;
  const long long total = P * G;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % G);
    const u32x8 vx = ldg256(y_raw + i * 16);
    const u32x8 vg = ldg256(dy + i * 16);
    u32x8 o;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float r[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int c = g * 16 + 2 * k + h;
        const float x = h ? bf16hi(vx.v[k]) : bf16lo(vx.v[k]);
        float gr = h ? bf16hi(vg.v[k]) : bf16lo(vg.v[k]);
        const float mean = __ldg(bn + c), invstd = __ldg(bn + C + c), sc = __ldg(bn + 2 * C + c), sh = __ldg(bn + 3 * C + c);
        if (!(x * sc + sh > 0.f)) gr = 0.f;
        const float xhat = (x - mean) * invstd;
        r[h] = sc * (gr - (__ldg(red + c) + xhat * __ldg(red + C + c)) * inv_count);
      }
      o.v[k] = pack_bf16x2(r[0], r[1]);
    }
    stg256(dz + i * 16, o);
  }
}

// ---- bilinear x2, align_corners=True -------------------------------------------------------------------
// forward: out[b, y, x, :] = lerp of in at (y*(H-1)/(2H-1), x*(W-1)/(2W-1))
__global__ void bilinear_up2x_fwd_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, int B, int H, int W, int C) {
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
  if (c % 16 || c <= 0 || c > 4096) return fail("batchnorm: c=%d must be a multiple of 16 (<= 4096)", c);
  if (scratch_bytes < dpt_bn_scratch_bytes(c)) return fail("batchnorm: scratch too small");
  return 0;
}

extern "C" int dpt_bn_forward_train(const void* y_raw, void* y, const float* gamma, const float* beta, float* running_mean,
                                    float* running_var, float* bn, float* sums_out, long long pixels, int c, float eps,
                                    float momentum, int relu, void* scratch, size_t scratch_bytes, void* stream) {
  if (bn_check(c, scratch_bytes)) return -1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* partial = static_cast<float*>(scratch);
  bn_reduce_stage1<0><<<kBnBlocks, 256, 0, st>>>(static_cast<const bf16*>(y_raw), nullptr, nullptr, partial, pixels, c);
  bn_finalize_kernel<<<(c + 31) / 32, 256, 0, st>>>(partial, kBnBlocks, gamma, beta, running_mean, running_var, bn, sums_out, c,
                                                      static_cast<float>(pixels), eps, momentum);
  bn_apply_relu_kernel<<<grid_for(pixels * (c / 16), 256), 256, 0, st>>>(static_cast<const bf16*>(y_raw), static_cast<bf16*>(y), bn,
                                                                         pixels, c, relu);
  g_launch_count.fetch_add(3, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_bn_forward_eval(const void* y_raw, void* y, const float* gamma, const float* beta, const float* running_mean,
                                   const float* running_var, float* bn, long long pixels, int c, float eps, int relu,
                                   void* stream) {
  if (c % 16 || c <= 0) return fail("batchnorm: c=%d must be a multiple of 16", c);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  bn_eval_prepare_kernel<<<(c + 127) / 128, 128, 0, st>>>(gamma, beta, running_mean, running_var, bn, c, eps);
  bn_apply_relu_kernel<<<grid_for(pixels * (c / 16), 256), 256, 0, st>>>(static_cast<const bf16*>(y_raw), static_cast<bf16*>(y), bn,
                                                                         pixels, c, relu);
  g_launch_count.fetch_add(2, std::memory_order_relaxed);
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
  bn_reduce_stage1<1><<<kBnBlocks, 256, 0, st>>>(static_cast<const bf16*>(y_raw), static_cast<const bf16*>(dy), bn, partial, pixels, c);
  bn_bwd_finalize_kernel<<<(c + 31) / 32, 256, 0, st>>>(partial, kBnBlocks, red, dgamma, dbeta, c, accumulate);
  bn_bwd_apply_kernel<<<grid_for(pixels * (c / 16), 256), 256, 0, st>>>(static_cast<const bf16*>(y_raw), static_cast<const bf16*>(dy),
                                                                        static_cast<bf16*>(dz), bn, red, pixels, c,
                                                                        1.f / static_cast<float>(pixels));
  g_launch_count.fetch_add(3, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_bilinear_up2x_fwd(const void* in, void* out, int batch, int h, int w, int c, void* stream) {
  if (c % 16) return fail("bilinear: c=%d must be a multiple of 16", c);
  const long long work = static_cast<long long>(batch) * 4 * h * w * (c / 16);
  bilinear_up2x_fwd_kernel<<<grid_for(work, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const bf16*>(in), static_cast<bf16*>(out), batch, h, w, c);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_bilinear_up2x_bwd(const void* dout, void* din, int batch, int h, int w, int c, void* stream) {
  if (c % 16) return fail("bilinear: c=%d must be a multiple of 16", c);
  const long long work = static_cast<long long>(batch) * h * w * (c / 16);
  bilinear_up2x_bwd_kernel<<<grid_for(work, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const bf16*>(dout), static_cast<bf16*>(din), batch, h, w, c, 0);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}
