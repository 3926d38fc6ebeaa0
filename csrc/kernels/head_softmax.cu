// Multi-class output head for the north-star model family: 1x1 conv (OutConv) -> softmax -> cross-entropy + Dice
// loss, forward and backward, fp32 math on NHWC bf16 activations (sm_100a).
//
// Loss (upstream milesial recipe, named by the north star; the reference itself only has the 1-class BCE/Dice head):
//   loss = CE(logits, target) + 1 - dice,   dice = (2 * sum_p p_target(p) + eps) / (2 * P + eps),  eps = 1e-6
// (multiclass_dice_coeff with reduce_batch_first=True sums over batch, classes and pixels; softmax rows and one-hot
// rows both sum to 1, so the denominator is the constant 2P).
// state (8 floats): ce_sum, sum p_target, -, -, loss, pixel count, ce mean, dice loss.
#include <atomic>

#include "../../include/dpt_b200.h"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dpt {

extern std::atomic<long long> g_launch_count;

constexpr int kSmThreads = 256;
constexpr int kSmMaxC = 64;
constexpr int kSmMaxK = 4;
constexpr float kDiceEps = 1e-6f;

__device__ __forceinline__ float sm_block_sum(float v, float* sh) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float r = 0.f;
  for (int w = 0; w < (blockDim.x >> 5); ++w) r += sh[w];
  return r;
}

__device__ __forceinline__ int load_class(const void* t, int dtype, long long i) {
  return dtype == 1 ? static_cast<int>(static_cast<const uint8_t*>(t)[i])
                    : static_cast<int>(static_cast<const float*>(t)[i]);
}

__device__ __forceinline__ void logits_of(const bf16* x, long long i, int C, int K, const float* ws, const float* bs,
                                          float* z) {
#pragma unroll
  for (int k = 0; k < kSmMaxK; ++k) z[k] = k < K ? bs[k] : -INFINITY;
  for (int g = 0; g < C / 16; ++g) {
    const u32x8 v = ldg256(x + i * C + g * 16);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float x0 = bf16lo(v.v[j]), x1 = bf16hi(v.v[j]);
#pragma unroll
      for (int k = 0; k < kSmMaxK; ++k)
        if (k < K) z[k] += x0 * ws[k * kSmMaxC + g * 16 + 2 * j] + x1 * ws[k * kSmMaxC + g * 16 + 2 * j + 1];
    }
  }
}

// logits_out: fp32 planar [B][K][HW] (what UNet.forward returns for the milesial variant), may be NULL.
__global__ void __launch_bounds__(kSmThreads)
head_softmax_fwd_kernel(const bf16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                        const void* __restrict__ target, int target_dtype, float* __restrict__ logits_out,
                        float* __restrict__ partial, long long P, long long HW, int C, int K) {
  grid_dep_sync();
  __shared__ float sh[8];
  __shared__ float ws[kSmMaxK * kSmMaxC];
  __shared__ float bs[kSmMaxK];
  for (int i = threadIdx.x; i < K * C; i += blockDim.x) ws[(i / C) * kSmMaxC + (i % C)] = w[i];
  if (threadIdx.x < K) bs[threadIdx.x] = b[threadIdx.x];
  __syncthreads();
  float s_ce = 0.f, s_pt = 0.f;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < P;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float z[kSmMaxK];
    logits_of(x, i, C, K, ws, bs, z);
    if (logits_out != nullptr) {
      const long long img = i / HW, pix = i - img * HW;
#pragma unroll
      for (int k = 0; k < kSmMaxK; ++k)
        if (k < K) logits_out[(img * K + k) * HW + pix] = z[k];
    }
    if (target != nullptr) {
      float m = z[0];
#pragma unroll
      for (int k = 1; k < kSmMaxK; ++k) m = fmaxf(m, z[k]);
      float den = 0.f;
#pragma unroll
      for (int k = 0; k < kSmMaxK; ++k)
        if (k < K) den += __expf(z[k] - m);
      const int t = load_class(target, target_dtype, i);
      float zt = z[0];
#pragma unroll
      for (int k = 1; k < kSmMaxK; ++k)
        if (k == t) zt = z[k];
      const float logp = zt - m - __logf(den);
      s_ce -= logp;
      s_pt += __expf(logp);
    }
  }
  const float r0 = sm_block_sum(s_ce, sh), r1 = sm_block_sum(s_pt, sh);
  if (threadIdx.x == 0) {
    float4* o = reinterpret_cast<float4*>(partial + 4 * static_cast<size_t>(blockIdx.x));
    *o = make_float4(r0, r1, 0.f, 0.f);
  }
}

struct SmParts {
  const float* p[8];
  int n;
};
// Sums the per-block partials of up to 8 arrays (micro-batches of a pipeline stage, or one all-reduced array in
// DataParallel) in a fixed order; `pixels` is the pixel count of the WHOLE batch the loss is taken over.
__global__ void head_softmax_finalize_kernel(const SmParts parts, int n_blocks, float pixels, float* __restrict__ state) {
  grid_dep_sync();
  __shared__ float sh[8];
  float a0 = 0.f, a1 = 0.f;
  for (int k = 0; k < parts.n; ++k) {
    const float* partial = parts.p[k];
    for (int i = threadIdx.x; i < n_blocks; i += blockDim.x) {
      a0 += partial[4 * static_cast<size_t>(i)];
      a1 += partial[4 * static_cast<size_t>(i) + 1];
    }
  }
  const float ce = sm_block_sum(a0, sh), spt = sm_block_sum(a1, sh);
  if (threadIdx.x == 0) {
    const float dice = (2.f * spt + kDiceEps) / (2.f * pixels + kDiceEps);
    state[0] = ce;
    state[1] = spt;
    state[2] = 0.f;
    state[3] = 0.f;
    state[4] = ce / pixels + 1.f - dice;
    state[5] = pixels;
    state[6] = ce / pixels;
    state[7] = 1.f - dice;
  }
}

// Backward: dz_k = gscale * [ (p_k - [k==t]) / P - (2 / (2P + eps)) * p_t * ([k==t] - p_k) ]
// Writes dx = sum_k dz_k * W[k] (bf16, no mask: the producer's BatchNorm backward owns the ReLU mask) and dz (fp32 [P][K]).
__global__ void __launch_bounds__(kSmThreads)
head_softmax_bwd_kernel(const bf16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                        const void* __restrict__ target, int target_dtype, const float* __restrict__ state,
                        bf16* __restrict__ dx, float* __restrict__ dz_out, long long P, int C, int K, float gscale) {
  grid_dep_sync();
  __shared__ float ws[kSmMaxK * kSmMaxC];
  __shared__ float bs[kSmMaxK];
  for (int i = threadIdx.x; i < K * C; i += blockDim.x) ws[(i / C) * kSmMaxC + (i % C)] = w[i];
  if (threadIdx.x < K) bs[threadIdx.x] = b[threadIdx.x];
  __syncthreads();
  const float pixels = state[5];
  const float inv_p = 1.f / pixels, dcoef = 2.f / (2.f * pixels + kDiceEps);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < P;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float z[kSmMaxK];
    logits_of(x, i, C, K, ws, bs, z);
    float m = z[0];
#pragma unroll
    for (int k = 1; k < kSmMaxK; ++k) m = fmaxf(m, z[k]);
    float pk[kSmMaxK], den = 0.f;
#pragma unroll
    for (int k = 0; k < kSmMaxK; ++k) {
      pk[k] = k < K ? __expf(z[k] - m) : 0.f;
      den += pk[k];
    }
    const int t = load_class(target, target_dtype, i);
    float pt = 0.f;
#pragma unroll
    for (int k = 0; k < kSmMaxK; ++k) {
      pk[k] /= den;
      if (k == t) pt = pk[k];
    }
    float dz[kSmMaxK];
#pragma unroll
    for (int k = 0; k < kSmMaxK; ++k) {
      const float ind = k == t ? 1.f : 0.f;
      dz[k] = k < K ? gscale * ((pk[k] - ind) * inv_p - dcoef * pt * (ind - pk[k])) : 0.f;
      if (k < K) dz_out[i * K + k] = dz[k];
    }
    for (int g = 0; g < C / 16; ++g) {
      u32x8 o;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float d0 = 0.f, d1 = 0.f;
#pragma unroll
        for (int k = 0; k < kSmMaxK; ++k)
          if (k < K) {
            d0 += dz[k] * ws[k * kSmMaxC + g * 16 + 2 * j];
            d1 += dz[k] * ws[k * kSmMaxC + g * 16 + 2 * j + 1];
          }
        o.v[j] = pack_bf16x2(d0, d1);
      }
      stg256(dx + i * C + g * 16, o);
    }
  }
}

// dW[k][c] = sum_p dz[p][k] * x[p][c], db[k] = sum_p dz[p][k]: stage 1 per block, stage 2 fixed-order over blocks.
__global__ void __launch_bounds__(256)
head_softmax_wgrad_stage1(const bf16* __restrict__ x, const float* __restrict__ dz, float* __restrict__ partial, long long P,
                          int C, int K) {
  grid_dep_sync();
  // partial: [blocks][K][C+1]
  __shared__ float sh[256][17];
  const int lanes = C / 16;
  const int ppi = 256 / lanes;
  const int lane = threadIdx.x % lanes, psub = threadIdx.x / lanes;
  const long long per_block = (P + gridDim.x - 1) / gridDim.x;
  const long long p0 = blockIdx.x * per_block;
  long long p1 = p0 + per_block;
  if (p1 > P) p1 = P;
  for (int k = 0; k < K; ++k) {
    float acc[16], accb = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) acc[j] = 0.f;
    if (psub < ppi) {
      for (long long p = p0 + psub; p < p1; p += ppi) {
        const float d = dz[p * K + k];
        const u32x8 v = ldg256(x + p * C + lane * 16);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          acc[2 * j] += d * bf16lo(v.v[j]);
          acc[2 * j + 1] += d * bf16hi(v.v[j]);
        }
        if (lane == 0) accb += d;
      }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 16; ++j) sh[threadIdx.x][j] = acc[j];
    sh[threadIdx.x][16] = accb;
    __syncthreads();
    float* out = partial + (static_cast<size_t>(blockIdx.x) * K + k) * (C + 1);
    for (int c = threadIdx.x; c <= C; c += 256) {
      float a = 0.f;
      if (c < C) {
        const int ln = c >> 4, j = c & 15;
        for (int q = 0; q < ppi; ++q) a += sh[q * lanes + ln][j];
      } else {
        for (int q = 0; q < ppi; ++q) a += sh[q * lanes][16];
      }
      out[c] = a;
    }
  }
}
__global__ void head_softmax_wgrad_stage2(const float* __restrict__ partial, int nblocks, int C, int K, float* __restrict__ dw,
                                          float* __restrict__ db, int accumulate) {
  grid_dep_sync();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= K * (C + 1)) return;
  const int k = i / (C + 1), c = i % (C + 1);
  float acc = 0.f;
  for (int b = 0; b < nblocks; ++b) acc += partial[(static_cast<size_t>(b) * K + k) * (C + 1) + c];
  float* dst = c < C ? dw + k * C + c : db + k;
  *dst = accumulate ? *dst + acc : acc;
}

}  // namespace dpt

using namespace dpt;

static int sm_check(int c, int k) {
  if (c % 16 || c > kSmMaxC || c <= 0) return fail("softmax head: c=%d must be a multiple of 16, <= %d", c, kSmMaxC);
  if (k < 2 || k > kSmMaxK) return fail("softmax head: classes=%d must be in 2..%d", k, kSmMaxK);
  return 0;
}

extern "C" int dpt_head_softmax_fwd(const void* x, const float* w, const float* b, const void* target, int target_dtype,
                                    float* logits_out, float* partial, float* state, long long pixels, long long hw, int c,
                                    int classes, void* stream) {
  if (sm_check(c, classes)) return -1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int blocks = dpt_head_blocks();
  DPT_CUDA_OK(launch_chained(head_softmax_fwd_kernel, dim3(blocks), dim3(kSmThreads), 0, st, static_cast<const bf16*>(x), w, b, target, target_dtype, logits_out,
                                                         partial, pixels, hw, c, classes));
  int n = 1;
  if (target != nullptr && state != nullptr) {
    SmParts parts{};
    parts.p[0] = partial;
    parts.n = 1;
    DPT_CUDA_OK(launch_chained(head_softmax_finalize_kernel, dim3(1), dim3(256), 0, st, parts, blocks, static_cast<float>(pixels), state));
    n = 2;
  }
  g_launch_count.fetch_add(n, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_head_softmax_finalize(const float* const* partials, int n_parts, double total_pixels, float* state,
                                         void* stream) {
  if (n_parts < 1 || n_parts > 8 || !partials || !state) return fail("softmax head finalize: 1..8 partial arrays and a state");
  SmParts parts{};
  for (int k = 0; k < n_parts; ++k) parts.p[k] = partials[k];
  parts.n = n_parts;
  DPT_CUDA_OK(launch_chained(head_softmax_finalize_kernel, dim3(1), dim3(256), 0, static_cast<cudaStream_t>(stream), parts, dpt_head_blocks(),
                                                                                 static_cast<float>(total_pixels), state));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" size_t dpt_head_softmax_scratch_bytes(long long pixels, int c, int classes) {
  return static_cast<size_t>(pixels) * classes * sizeof(float) +
         static_cast<size_t>(296) * classes * (c + 1) * sizeof(float);
}

extern "C" int dpt_head_softmax_bwd(const void* x, const float* w, const float* b, const void* target, int target_dtype,
                                    const float* state, void* dx, float* dw, float* db, long long pixels, int c, int classes,
                                    float gscale, int accumulate, void* scratch, size_t scratch_bytes, void* stream) {
  if (sm_check(c, classes)) return -1;
  if (scratch_bytes < dpt_head_softmax_scratch_bytes(pixels, c, classes)) return fail("softmax head: scratch too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* dz = static_cast<float*>(scratch);
  float* partial = dz + static_cast<size_t>(pixels) * classes;
  DPT_CUDA_OK(launch_chained(head_softmax_bwd_kernel, dim3(dpt_head_blocks()), dim3(kSmThreads), 0, st, static_cast<const bf16*>(x), w, b, target, target_dtype, state,
                                                                    static_cast<bf16*>(dx), dz, pixels, c, classes, gscale));
  DPT_CUDA_OK(launch_chained(head_softmax_wgrad_stage1, dim3(296), dim3(256), 0, st, static_cast<const bf16*>(x), dz, partial, pixels, c, classes));
  const int n = classes * (c + 1);
  DPT_CUDA_OK(launch_chained(head_softmax_wgrad_stage2, dim3((n + 127) / 128), dim3(128), 0, st, partial, 296, c, classes, dw, db, accumulate));
  g_launch_count.fetch_add(3, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}
