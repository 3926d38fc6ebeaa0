// Output head + loss (forward and backward) and the flat-arena optimizers (sm_100a, fp32 math).
//
// Head, reference variant (model/unet_model.py:10-11,59-60 + utils/utils.py:9-25 of the reference):
//   z = w . x + b ; p = sigmoid(z) ; loss = mean BCE(p, t) - dice_weight * log(2 I / (sum p + sum[t==1] + 1e-15)),
//   I = sum p [t==1], all sums over the WHOLE batch.  One pass over the full-resolution activation produces
//   the probabilities and the four partial sums; the loss scalar stays on the device (no .item() sync,
//   utils/train_utils.py:67).  The backward pass fuses d(loss)/dz, the 1x1-conv input gradient, the ReLU
//   mask of the producing conv, and the weight/bias gradient partial sums.
// Optimizers: Adam with L2 weight decay exactly as torch.optim.Adam(weight_decay=...) (utils/train_utils.py:45)
//   and RMSprop with momentum (north-star recipe), over one flat fp32 arena.
#include <atomic>

#include "../../include/dpt_b200.h"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dpt {

extern std::atomic<long long> g_launch_count;

constexpr int kHeadThreads = 256;
constexpr int kHeadMaxC = 64;

__device__ __forceinline__ float block_sum(float v, float* sh) {
  // deterministic: warp shuffle tree then fixed-order sum over warps
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float r = 0.f;
  for (int w = 0; w < (blockDim.x >> 5); ++w) r += sh[w];
  return r;
}

__device__ __forceinline__ float load_target(const void* t, int dtype, long long i) {
  return dtype == 1 ? static_cast<float>(static_cast<const uint8_t*>(t)[i]) : static_cast<const float*>(t)[i];
}

// x: [P][C] bf16 (C multiple of 16, <= 64).  partial: [gridDim.x][4] = {bce_sum, inter, psum, tsum}
__global__ void __launch_bounds__(kHeadThreads)
head_sigmoid_fwd_kernel(const bf16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                        const void* __restrict__ target, int target_dtype, float* __restrict__ prob,
                        float* __restrict__ partial, long long P, int C) {
  grid_dep_sync();
  __shared__ float sh[8];
  __shared__ float ws[kHeadMaxC];
  if (threadIdx.x < C) ws[threadIdx.x] = w[threadIdx.x];
  __syncthreads();
  const float bias = b[0];
  float s_bce = 0.f, s_i = 0.f, s_p = 0.f, s_t = 0.f;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < P;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float z = bias;
    for (int g = 0; g < C / 16; ++g) {
      const u32x8 v = ldg256(x + i * C + g * 16);
#pragma unroll
      for (int k = 0; k < 8; ++k) z += bf16lo(v.v[k]) * ws[g * 16 + 2 * k] + bf16hi(v.v[k]) * ws[g * 16 + 2 * k + 1];
    }
    const float p = 1.f / (1.f + __expf(-z));
    if (prob != nullptr) prob[i] = p;
    if (target != nullptr) {
      const float t = load_target(target, target_dtype, i);
      // log p = -softplus(-z), log(1-p) = -softplus(z); BCELoss clamps both at -100
      const float sp_neg = z > 0.f ? log1pf(__expf(-z)) : -z + log1pf(__expf(z));  // softplus(-z)
      const float sp_pos = sp_neg + z;                                              // softplus(z)
      const float logp = fmaxf(-sp_neg, -100.f), log1mp = fmaxf(-sp_pos, -100.f);
      s_bce -= t * logp + (1.f - t) * log1mp;
      const float t1 = t == 1.f ? 1.f : 0.f;
      s_i += p * t1;
      s_p += p;
      s_t += t1;
    }
  }
  const float r0 = block_sum(s_bce, sh), r1 = block_sum(s_i, sh), r2 = block_sum(s_p, sh), r3 = block_sum(s_t, sh);
  if (threadIdx.x == 0) {
    float4* o = reinterpret_cast<float4*>(partial + 4 * static_cast<size_t>(blockIdx.x));
    *o = make_float4(r0, r1, r2, r3);
  }
}

// state: [0]=bce_sum [1]=I [2]=sum p [3]=sum t1 [4]=loss [5]=total pixel count [6]=bce mean [7]=dice term
struct HeadFinalizeParams {
  const float* partial[8];  // one partial array per replica (peer-readable); n_parts of them
  int n_blocks[8];
  int n_parts;
  float pixels;  // total pixels over all parts
  float dice_weight;
  float* state;
  float* loss_log;  // optional ring of past losses
  int log_slot;
};
__global__ void head_finalize_kernel(const HeadFinalizeParams p) {
  grid_dep_sync();
  __shared__ float sh[8];
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int part = 0; part < p.n_parts; ++part)
    for (int i = threadIdx.x; i < p.n_blocks[part]; i += blockDim.x) {
      const float4 v = *reinterpret_cast<const float4*>(p.partial[part] + 4 * static_cast<size_t>(i));
      acc[0] += v.x;
      acc[1] += v.y;
      acc[2] += v.z;
      acc[3] += v.w;
    }
  float r[4];
  for (int k = 0; k < 4; ++k) r[k] = block_sum(acc[k], sh);
  if (threadIdx.x == 0) {
    const float bce = r[0] / p.pixels;
    const float uni = r[2] + r[3] + 1e-15f;
    const float dice = p.dice_weight != 0.f ? -p.dice_weight * logf(2.f * r[1] / uni) : 0.f;
    p.state[0] = r[0];
    p.state[1] = r[1];
    p.state[2] = r[2];
    p.state[3] = r[3];
    p.state[4] = bce + dice;
    p.state[5] = p.pixels;
    p.state[6] = bce;
    p.state[7] = dice;
    if (p.loss_log != nullptr) p.loss_log[p.log_slot] = bce + dice;
  }
}

// Backward.  dz = gscale * [ (p - t)/Ptot + dice_weight * (-t1/I + 1/U) * p (1-p) ]
// dx[c] = dz * w[c] * (x[c] > 0 if relu_mask);  partial_w: [gridDim.x][C+1] = {sum dz*x[c]..., sum dz}
// CG = C / 16 is a template parameter so the per-thread weight-gradient accumulators and pixel registers are sized for
// the real channel count: with the 64-channel maximum compiled in, 166 registers left one 256-thread block per SM and the
// stream ran at 1.9 TB/s.
template <int CG>
__global__ void __launch_bounds__(kHeadThreads, CG <= 2 ? 2 : 1)
head_sigmoid_bwd_kernel(const bf16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                        const void* __restrict__ target, int target_dtype, const float* __restrict__ state,
                        bf16* __restrict__ dx, float* __restrict__ partial_w, long long P, int C, float gscale,
                        float dice_weight, int relu_mask) {
  grid_dep_sync();
  __shared__ float sh[8];
  __shared__ float ws[kHeadMaxC];
  if (threadIdx.x < C) ws[threadIdx.x] = w[threadIdx.x];
  __syncthreads();
  const float bias = b[0];
  const float inter = state[1], uni = state[2] + state[3] + 1e-15f, inv_p = 1.f / state[5];
  float gw[CG * 16];
#pragma unroll
  for (int c = 0; c < CG * 16; ++c) gw[c] = 0.f;
  float gb = 0.f;
  // two pixels per thread per iteration: twice the bytes in flight (the loop is a pure stream: 64 B in, 64 B out)
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i0 = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i0 < P; i0 += 2 * stride) {
    u32x8 v[2][CG];
    float tt[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long i = i0 + u * stride;
      if (i < P) {
#pragma unroll
        for (int g = 0; g < CG; ++g)
          if (g * 16 < C) v[u][g] = ldg256(x + i * C + g * 16);
        tt[u] = load_target(target, target_dtype, i);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long i = i0 + u * stride;
      if (i >= P) break;
      float z = bias;
#pragma unroll
      for (int g = 0; g < CG; ++g) {
        if (g * 16 < C) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            z += bf16lo(v[u][g].v[k]) * ws[g * 16 + 2 * k] + bf16hi(v[u][g].v[k]) * ws[g * 16 + 2 * k + 1];
        }
      }
      const float p = 1.f / (1.f + __expf(-z));
      const float t = tt[u];
      const float t1 = t == 1.f ? 1.f : 0.f;
      float dz = (p - t) * inv_p;
      if (dice_weight != 0.f) dz += dice_weight * (-t1 / inter + 1.f / uni) * p * (1.f - p);
      dz *= gscale;
      gb += dz;
#pragma unroll
      for (int g = 0; g < CG; ++g) {
        if (g * 16 < C) {
          u32x8 o;
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float x0 = bf16lo(v[u][g].v[k]), x1 = bf16hi(v[u][g].v[k]);
            gw[g * 16 + 2 * k] += dz * x0;
            gw[g * 16 + 2 * k + 1] += dz * x1;
            float d0 = dz * ws[g * 16 + 2 * k], d1 = dz * ws[g * 16 + 2 * k + 1];
            if (relu_mask) {
              if (!(x0 > 0.f)) d0 = 0.f;
              if (!(x1 > 0.f)) d1 = 0.f;
            }
            o.v[k] = pack_bf16x2(d0, d1);
          }
          stg256(dx + i * C + g * 16, o);
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < CG * 16; ++c) {
    if (c < C) {
      const float r = block_sum(gw[c], sh);
      if (threadIdx.x == 0) partial_w[static_cast<size_t>(blockIdx.x) * (C + 1) + c] = r;
    }
  }
  const float rb = block_sum(gb, sh);
  if (threadIdx.x == 0) partial_w[static_cast<size_t>(blockIdx.x) * (C + 1) + C] = rb;
}

// 64 channel slots x 8 lanes: each lane walks every 8th block partial, then a fixed-order 8-way sum (a single thread
// per channel walking ~300 partials serialised on L2 latency).
__global__ void __launch_bounds__(512) head_wgrad_finalize_kernel(const float* __restrict__ partial_w, int n_blocks, int C,
                                                                  float* __restrict__ dw, float* __restrict__ db,
                                                                  int accumulate) {
  grid_dep_sync();
  __shared__ float sh[8][65];
  const int c = threadIdx.x & 63, sl = threadIdx.x >> 6;
  float acc = 0.f;
  if (c <= C)
    for (int i = sl; i < n_blocks; i += 8) acc += partial_w[static_cast<size_t>(i) * (C + 1) + c];
  sh[sl][c] = acc;
  __syncthreads();
  if (sl != 0 || c > C) return;
  float v = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) v += sh[k][c];
  if (c < C)
    dw[c] = accumulate ? dw[c] + v : v;
  else
    db[0] = accumulate ? db[0] + v : v;
}

__global__ void opt_tick_kernel(float* st) {
  grid_dep_sync();
  const float step = st[1] + 1.f;
  st[1] = step;
  st[2] = 1.f - powf(st[4], step);
  st[3] = 1.f - powf(st[5], step);
}
__global__ void opt_init_kernel(float* st, float lr, float step, float beta1, float beta2) {
  grid_dep_sync();
  st[0] = lr;
  st[1] = step;
  st[4] = beta1;
  st[5] = beta2;
  st[2] = 1.f - powf(beta1, step);
  st[3] = 1.f - powf(beta2, step);
}
__global__ void opt_set_lr_kernel(float* st, float lr) {
  grid_dep_sync();
  st[0] = lr;
}

__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, float lr, float b1, float b2,
                                            float bc1, float bc2, float eps, float wd) {
  g += wd * p;
  m = b1 * m + (1.f - b1) * g;
  v = b2 * v + (1.f - b2) * g * g;
  const float denom = sqrtf(v) / sqrtf(bc2) + eps;
  p -= (lr / bc1) * (m / denom);
}

__global__ void adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                            long long n, const float* __restrict__ st, float eps, float wd, float gscale,
                            int zero_grad) {
  grid_dep_sync();
  const float lr = st[0], bc1 = st[2], bc2 = st[3], b1 = st[4], b2 = st[5];
  const long long n4 = n >> 2;
  float4* p4 = reinterpret_cast<float4*>(p);
  float4* g4 = reinterpret_cast<float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 pp = p4[i], gg = g4[i], mm = m4[i], vv = v4[i];
    adam_update(pp.x, gg.x * gscale, mm.x, vv.x, lr, b1, b2, bc1, bc2, eps, wd);
    adam_update(pp.y, gg.y * gscale, mm.y, vv.y, lr, b1, b2, bc1, bc2, eps, wd);
    adam_update(pp.z, gg.z * gscale, mm.z, vv.z, lr, b1, b2, bc1, bc2, eps, wd);
    adam_update(pp.w, gg.w * gscale, mm.w, vv.w, lr, b1, b2, bc1, bc2, eps, wd);
    p4[i] = pp;
    m4[i] = mm;
    v4[i] = vv;
    if (zero_grad) g4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const long long i = (n4 << 2) + threadIdx.x;
    float pp = p[i], mm = m[i], vv = v[i];
    adam_update(pp, g[i] * gscale, mm, vv, lr, b1, b2, bc1, bc2, eps, wd);
    p[i] = pp;
    m[i] = mm;
    v[i] = vv;
    if (zero_grad) g[i] = 0.f;
  }
}

// torch.optim.RMSprop(alpha, eps, weight_decay, momentum), not centered.  `sq` = square_avg, `buf` = momentum
// buffer (may be NULL when momentum == 0).
__global__ void rmsprop_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ sq,
                               float* __restrict__ buf, long long n, const float* __restrict__ st, float alpha, float eps,
                               float wd, float momentum, float gscale, int zero_grad) {
  grid_dep_sync();
  const float lr = st[0];
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float pp = p[i];
    float gg = g[i] * gscale + wd * pp;
    const float s = alpha * sq[i] + (1.f - alpha) * gg * gg;
    sq[i] = s;
    const float avg = sqrtf(s) + eps;
    if (momentum > 0.f) {
      const float bb = momentum * buf[i] + gg / avg;
      buf[i] = bb;
      pp -= lr * bb;
    } else {
      pp -= lr * gg / avg;
    }
    p[i] = pp;
    if (zero_grad) g[i] = 0.f;
  }
}

static inline int grid_for(long long work, int threads, int per_sm = 8) {
  long long b = (work + threads - 1) / threads;
  const long long cap = static_cast<long long>(sm_count()) * per_sm;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

}  // namespace dpt

using namespace dpt;

extern "C" int dpt_head_blocks(void) { return 4 * sm_count_real(); }
extern "C" int dpt_set_sm_reserve(int n) {
  if (n < 0 || n > sm_count_real() - 8) return fail("set_sm_reserve: %d is outside 0..%d", n, sm_count_real() - 8);
  sm_reserve_ref() = n;
  return 0;
}
extern "C" int dpt_get_sm_count(void) { return sm_count(); }

extern "C" int dpt_head_sigmoid_fwd(const void* x, const float* w, const float* b, const void* target, int target_dtype,
                                    float* prob, float* partial, long long pixels, int c, void* stream) {
  if (c % 16 || c > kHeadMaxC) return fail("head: c=%d must be a multiple of 16, <= %d", c, kHeadMaxC);
  DPT_CUDA_OK(launch_chained(head_sigmoid_fwd_kernel, dim3(dpt_head_blocks()), dim3(kHeadThreads), 0, static_cast<cudaStream_t>(stream), static_cast<const bf16*>(x), w, b, target, target_dtype, prob, partial, pixels, c));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_head_finalize(const float* const* partials, const int* n_blocks, int n_parts, double total_pixels,
                                 float dice_weight, float* state, float* loss_log, int log_slot, void* stream) {
  if (n_parts < 1 || n_parts > 8) return fail("head_finalize: n_parts=%d must be in 1..8", n_parts);
  HeadFinalizeParams p;
  memset(&p, 0, sizeof(p));
  for (int i = 0; i < n_parts; ++i) {
    p.partial[i] = partials[i];
    p.n_blocks[i] = n_blocks[i];
  }
  p.n_parts = n_parts;
  p.pixels = static_cast<float>(total_pixels);
  p.dice_weight = dice_weight;
  p.state = state;
  p.loss_log = loss_log;
  p.log_slot = log_slot;
  DPT_CUDA_OK(launch_chained(head_finalize_kernel, dim3(1), dim3(256), 0, static_cast<cudaStream_t>(stream), p));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_head_sigmoid_bwd_ex(const void* x, const float* w, const float* b, const void* target,
                                       int target_dtype, const float* state, void* dx, float* partial_w, float* dw,
                                       float* db, long long pixels, int c, float gscale, float dice_weight,
                                       int relu_mask, int accumulate, void* stream);
extern "C" int dpt_head_sigmoid_bwd(const void* x, const float* w, const float* b, const void* target, int target_dtype,
                                    const float* state, void* dx, float* partial_w, float* dw, float* db,
                                    long long pixels, int c, float gscale, float dice_weight, int relu_mask,
                                    void* stream) {
  return dpt_head_sigmoid_bwd_ex(x, w, b, target, target_dtype, state, dx, partial_w, dw, db, pixels, c, gscale,
                                 dice_weight, relu_mask, 0, stream);
}
extern "C" int dpt_head_sigmoid_bwd_ex(const void* x, const float* w, const float* b, const void* target,
                                       int target_dtype, const float* state, void* dx, float* partial_w, float* dw,
                                       float* db, long long pixels, int c, float gscale, float dice_weight,
                                       int relu_mask, int accumulate, void* stream) {
  if (c % 16 || c > kHeadMaxC) return fail("head: c=%d must be a multiple of 16, <= %d", c, kHeadMaxC);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int blocks = dpt_head_blocks();
  auto launch = [&](auto kernel) {
    DPT_CUDA_OK(launch_chained(kernel, dim3(blocks), dim3(kHeadThreads), 0, st, static_cast<const bf16*>(x), w, b, target, target_dtype, state,
                                            static_cast<bf16*>(dx), partial_w, pixels, c, gscale, dice_weight, relu_mask));
  };
  switch (c / 16) {
    case 1: launch(head_sigmoid_bwd_kernel<1>); break;
    case 2: launch(head_sigmoid_bwd_kernel<2>); break;
    case 3: launch(head_sigmoid_bwd_kernel<3>); break;
    default: launch(head_sigmoid_bwd_kernel<4>); break;
  }
  DPT_CUDA_OK(cudaGetLastError());
  DPT_CUDA_OK(launch_chained(head_wgrad_finalize_kernel, dim3(1), dim3(512), 0, st, partial_w, blocks, c, dw, db, accumulate));
  g_launch_count.fetch_add(2, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_opt_init(float* state, float lr, int step, float beta1, float beta2, void* stream) {
  DPT_CUDA_OK(launch_chained(opt_init_kernel, dim3(1), dim3(1), 0, static_cast<cudaStream_t>(stream), state, lr, static_cast<float>(step), beta1, beta2));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}
extern "C" int dpt_opt_set_lr(float* state, float lr, void* stream) {
  DPT_CUDA_OK(launch_chained(opt_set_lr_kernel, dim3(1), dim3(1), 0, static_cast<cudaStream_t>(stream), state, lr));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}
extern "C" int dpt_opt_tick(float* state, void* stream) {
  DPT_CUDA_OK(launch_chained(opt_tick_kernel, dim3(1), dim3(1), 0, static_cast<cudaStream_t>(stream), state));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_adam_step(float* p, float* g, float* m, float* v, long long n, const float* state, float eps,
                             float weight_decay, float grad_scale, int zero_grad, void* stream) {
  if ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
       reinterpret_cast<uintptr_t>(v)) & 15)
    return fail("adam: arenas must be 16-byte aligned");
  DPT_CUDA_OK(launch_chained(adam_kernel, dim3(grid_for(n / 4 + 1, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), p, g, m, v, n, state, eps,
                                                                                      weight_decay, grad_scale, zero_grad));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_rmsprop_step(float* p, float* g, float* sq, float* buf, long long n, const float* state, float alpha,
                                float eps, float weight_decay, float momentum, float grad_scale, int zero_grad,
                                void* stream) {
  if (momentum > 0.f && buf == nullptr) return fail("rmsprop: momentum needs a buffer");
  DPT_CUDA_OK(launch_chained(rmsprop_kernel, dim3(grid_for(n, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), p, g, sq, buf, n, state, alpha, eps, weight_decay, momentum, grad_scale, zero_grad));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}
