// Bandwidth-bound kernels of the UNet step (sm_100a, 256-bit global accesses, NHWC bf16 activations):
// input packing, 2x2 max-pool forward, fused pool-backward + skip-add + ReLU-backward, per-channel sums
// (bias gradients), kernel-side weight re-layouts.  They replace the ATen kernels listed in SURVEY.md
// K24-K27/K35 (ReLU, max_pool2d_with_indices, cat, threshold_backward, max_pool2d backward).
#include <atomic>

#include "../../include/dpt_b200.h"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dpt {

std::atomic<long long> g_launch_count{0};

static inline int grid_for(long long work, int threads, int per_sm = 8) {
  long long b = (work + threads - 1) / threads;
  const long long cap = static_cast<long long>(sm_count()) * per_sm;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

// ------------------------------------------------------------------------------------------------
// fp32 NCHW image -> bf16 NHWC padded to 16 channels (zero fill): the layout the first conv's TMA reads.
// Replaces the .to(torch.float32) + implicit cuDNN layout handling of utils/train_utils.py:60.
// ------------------------------------------------------------------------------------------------
// T = float: the reference's host pipeline (already divided by 255, utils/dataloading.py:40).  T = uint8_t: raw pixel values
// straight from the decoder, normalised here (`scale` = 1/255): a quarter of the host->device bytes.
template <typename T>
__global__ void pack_input_kernel(const T* __restrict__ src, bf16* __restrict__ dst, int B, int C, long long HW, float scale) {
  grid_dep_sync();
  const long long total = static_cast<long long>(B) * HW;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long b = i / HW, pix = i - b * HW;
    float v[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) v[c] = c < C ? static_cast<float>(__ldg(src + (b * C + c) * HW + pix)) * scale : 0.f;
    u32x8 o;
#pragma unroll
    for (int k = 0; k < 8; ++k) o.v[k] = pack_bf16x2(v[2 * k], v[2 * k + 1]);
    stg256(dst + i * 16, o);
  }
}

// ------------------------------------------------------------------------------------------------
// 2x2 stride-2 max pool, NHWC bf16 (model/unet_parts.py:26).  One thread: one output pixel x 16 channels.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t max_bf16x2(uint32_t a, uint32_t b) {
  const float lo = fmaxf(bf16lo(a), bf16lo(b));
  const float hi = fmaxf(bf16hi(a), bf16hi(b));
  return (__float_as_uint(lo) >> 16) | (__float_as_uint(hi) & 0xffff0000u);
}
__global__ void maxpool2x2_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, int B, int H, int W, int C) {
  grid_dep_sync();
  const int Ho = H / 2, Wo = W / 2, G = C / 16;
  const long long total = static_cast<long long>(B) * Ho * Wo * G;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % G);
    long long r = i / G;
    const int xo = static_cast<int>(r % Wo);
    r /= Wo;
    const int yo = static_cast<int>(r % Ho);
    const int b = static_cast<int>(r / Ho);
    const bf16* p00 = x + ((static_cast<long long>(b) * H + 2 * yo) * W + 2 * xo) * C + g * 16;
    const u32x8 a = ldg256(p00), bb = ldg256(p00 + C), c = ldg256(p00 + static_cast<long long>(W) * C),
                d = ldg256(p00 + static_cast<long long>(W) * C + C);
    u32x8 o;
#pragma unroll
    for (int k = 0; k < 8; ++k) o.v[k] = max_bf16x2(max_bf16x2(a.v[k], bb.v[k]), max_bf16x2(c.v[k], d.v[k]));
    stg256(y + i * 16, o);
  }
}

// ------------------------------------------------------------------------------------------------
// Backward of [ReLU -> (skip branch, 2x2 max-pool branch)]:
//   dz = ( d_skip + route(d_pool -> first arg-max of the 2x2 window) ) * (act > 0)
// `act` is the saved post-ReLU activation.  d_skip may be NULL (no skip consumer) and may alias dz.
// Replaces MaxPool2DWithIndicesBackward0 + the gradient add + ReluBackward0 of autograd.
// ------------------------------------------------------------------------------------------------
__global__ void pool_bwd_fused_kernel(const bf16* __restrict__ act, const bf16* __restrict__ d_pool,
                                      const bf16* d_skip, bf16* dz, int B, int H, int W, int C, int relu_mask) {
  grid_dep_sync();
  const int Ho = H / 2, Wo = W / 2, G = C / 16;
  const long long total = static_cast<long long>(B) * Ho * Wo * G;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % G);
    long long r = i / G;
    const int xo = static_cast<int>(r % Wo);
    r /= Wo;
    const int yo = static_cast<int>(r % Ho);
    const int b = static_cast<int>(r / Ho);
    const long long o00 = ((static_cast<long long>(b) * H + 2 * yo) * W + 2 * xo) * C + g * 16;
    const long long offs[4] = {o00, o00 + C, o00 + static_cast<long long>(W) * C, o00 + static_cast<long long>(W) * C + C};
    u32x8 a[4], ds[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) a[k] = ldg256(act + offs[k]);
    if (d_skip != nullptr) {
#pragma unroll
      for (int k = 0; k < 4; ++k) ds[k] = ldg256(d_skip + offs[k]);
    }
    const u32x8 dp = ldg256(d_pool + i * 16);
    u32x8 out[4];
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      float res[4][2];
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = half ? bf16hi(a[k].v[w]) : bf16lo(a[k].v[w]);
        int arg = 0;
        float best = v[0];
#pragma unroll
        for (int k = 1; k < 4; ++k)
          if (v[k] > best) {
            best = v[k];
            arg = k;
          }
        const float gp = half ? bf16hi(dp.v[w]) : bf16lo(dp.v[w]);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float gsum = (k == arg) ? gp : 0.f;
          if (d_skip != nullptr) gsum += half ? bf16hi(ds[k].v[w]) : bf16lo(ds[k].v[w]);
          if (relu_mask && !(v[k] > 0.f)) gsum = 0.f;
          res[k][half] = gsum;
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) out[k].v[w] = pack_bf16x2(res[k][0], res[k][1]);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) stg256(dz + offs[k], out[k]);
  }
}

// ------------------------------------------------------------------------------------------------
// Per-channel sum over pixels of an NHWC bf16 tensor (bias gradients), deterministic two-stage.
// ------------------------------------------------------------------------------------------------
constexpr int kCsumBlocks = 296;
__global__ void channel_sum_stage1(const bf16* __restrict__ x, float* __restrict__ partial, long long P, int C) {
  grid_dep_sync();
  // block handles a contiguous pixel range; thread t owns channel pair (t % (C/2)) and strides over pixels
  const int pairs = C / 2;
  const int lanes_per_pix = pairs;
  const int pix_per_iter = blockDim.x / lanes_per_pix;
  const int cp = threadIdx.x % lanes_per_pix;
  const int psub = threadIdx.x / lanes_per_pix;
  const long long per_block = (P + gridDim.x - 1) / gridDim.x;
  const long long p0 = blockIdx.x * per_block;
  long long p1 = p0 + per_block;
  if (p1 > P) p1 = P;
  float s0 = 0.f, s1 = 0.f;
  if (psub < pix_per_iter) {
    const uint32_t* x32 = reinterpret_cast<const uint32_t*>(x);
    for (long long p = p0 + psub; p < p1; p += pix_per_iter) {
      const uint32_t u = __ldg(x32 + p * pairs + cp);
      s0 += bf16lo(u);
      s1 += bf16hi(u);
    }
  }
  extern __shared__ float sh[];  // [blockDim.x][2]
  sh[2 * threadIdx.x] = s0;
  sh[2 * threadIdx.x + 1] = s1;
  __syncthreads();
  if (threadIdx.x < lanes_per_pix) {
    float a0 = 0.f, a1 = 0.f;
    for (int k = 0; k < pix_per_iter; ++k) {
      a0 += sh[2 * (k * lanes_per_pix + threadIdx.x)];
      a1 += sh[2 * (k * lanes_per_pix + threadIdx.x) + 1];
    }
    partial[static_cast<size_t>(blockIdx.x) * C + 2 * threadIdx.x] = a0;
    partial[static_cast<size_t>(blockIdx.x) * C + 2 * threadIdx.x + 1] = a1;
  }
}
// Same partial sums with 256-bit loads (C a multiple of 16): C/16 threads per pixel, 16 fp32 accumulators per
// thread, four pixels in flight per thread.  The strided-scalar version above reached 0.6 TB/s; this one is
// bound by HBM.
__global__ void __launch_bounds__(256)
channel_sum_stage1_v8(const bf16* __restrict__ x, float* __restrict__ partial, long long P, int C) {
  grid_dep_sync();
  __shared__ float sh[256][17];
  const int lanes = C / 16;
  const int ppi = 256 / lanes;
  const int lane = threadIdx.x % lanes;
  const int psub = threadIdx.x / lanes;
  const long long per_block = (P + gridDim.x - 1) / gridDim.x;
  const long long p0 = blockIdx.x * per_block;
  long long p1 = p0 + per_block;
  if (p1 > P) p1 = P;
  float acc[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) acc[j] = 0.f;
  if (psub < ppi) {
    const bf16* base = x + lane * 16;
    long long p = p0 + psub;
    for (; p + 3LL * ppi < p1; p += 4LL * ppi) {
      u32x8 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = ldg256(base + (p + static_cast<long long>(u) * ppi) * C);
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          acc[2 * k] += bf16lo(v[u].v[k]);
          acc[2 * k + 1] += bf16hi(v[u].v[k]);
        }
    }
    for (; p < p1; p += ppi) {
      const u32x8 v = ldg256(base + p * C);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        acc[2 * k] += bf16lo(v.v[k]);
        acc[2 * k + 1] += bf16hi(v.v[k]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 16; ++j) sh[threadIdx.x][j] = acc[j];
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += 256) {
    const int ln = c >> 4, j = c & 15;
    float a = 0.f;
    for (int k = 0; k < ppi; ++k) a += sh[k * lanes + ln][j];
    partial[static_cast<size_t>(blockIdx.x) * C + c] = a;
  }
}

static int launch_channel_sum_stage1(const bf16* x, float* scratch, long long pixels, int c, cudaStream_t st) {
  if (c % 16 == 0 && c <= 4096) {
    DPT_CUDA_OK(launch_chained(channel_sum_stage1_v8, dim3(kCsumBlocks), dim3(256), 0, st, x, scratch, pixels, c));
  } else {
    int threads = 256;
    while (threads < c / 2) threads <<= 1;
    DPT_CUDA_OK(launch_chained(channel_sum_stage1, dim3(kCsumBlocks), dim3(threads), threads * 2 * sizeof(float), st, x, scratch, pixels, c));
  }
  return 0;
}

// Stage 2: one block per 32 channels, 8 threads per channel walk the per-block partials with stride 8 (the partials
// of one channel are C floats apart, so a single thread walking all of them serialises on L2 latency: that was a
// ~20 us floor per bias gradient), then a fixed-order 8-way sum in shared memory: bit-reproducible.
constexpr int kCsum2Lanes = 8;
__global__ void __launch_bounds__(32 * kCsum2Lanes)
channel_sum_stage2(const float* __restrict__ partial, float* __restrict__ out, int nblocks, int C, float scale,
                   int accumulate = 0) {
  grid_dep_sync();
  __shared__ float sh[kCsum2Lanes][33];
  const int o = threadIdx.x & 31, sl = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + o;
  float acc = 0.f;
  if (c < C) {
#pragma unroll 8
    for (int b = sl; b < nblocks; b += kCsum2Lanes) acc += partial[static_cast<size_t>(b) * C + c];
  }
  sh[sl][o] = acc;
  __syncthreads();
  if (sl == 0 && c < C) {
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < kCsum2Lanes; ++k) v += sh[k][o];
    out[c] = accumulate ? out[c] + v * scale : v * scale;
  }
}

// ------------------------------------------------------------------------------------------------
// Kernel-side weight layouts from the fp32 master parameters (PyTorch layouts), table driven: one launch
// re-lays every layer.
//   kind 0: Conv2d 3x3   w[co][ci][3][3] -> wf[t][co][ci_pad] (forward), wd[8-t][ci][co] (backward-data)
//   kind 1: ConvTranspose2d 2x2 s2  w[ci][co][2][2] -> wf[0][q*Cout+co][ci], wd[q][ci][co], bias4[q*Cout+co]
// ------------------------------------------------------------------------------------------------
// One launch walks the whole table.  Work units (a 32x32x9 tile, or 256 (co,ci) pairs of an untiled entry) are numbered
// across all entries and dealt round-robin to the blocks, so the three layers that hold 70 % of the parameters do not
// leave most of the GPU idle (a block-per-entry grid took 0.085 ms for 62 MB of traffic).
// Every block walks the whole table, so the table is copied to shared memory first (one load latency instead of one per
// entry: 23 dependent 64-byte reads were ~15 us of a block's life), and a tile's source floats are fetched as 128-bit loads
// that are all in flight before the first conversion (36 dependent scalar loads per thread before: the kernel took 68 us
// for 62 MB of traffic with the memory system 6 % busy).
constexpr int kPrepTableSmem = 48;
static_assert(sizeof(dpt_prep_entry) % 16 == 0, "the table is copied as uint4");
__global__ void __launch_bounds__(256, 3) prep_weights_kernel(const dpt_prep_entry* __restrict__ table, int n_entries) {
  grid_dep_sync();
  __shared__ bf16 sh[32][290];  // tiled path; row pitch 145 words (odd): the transposed read is conflict-free
  __shared__ uint4 tab_raw[kPrepTableSmem * sizeof(dpt_prep_entry) / 16];
  const dpt_prep_entry* tab = table;
  if (n_entries <= kPrepTableSmem) {
    const int n16 = n_entries * static_cast<int>(sizeof(dpt_prep_entry) / 16);
    for (int i = threadIdx.x; i < n16; i += 256) tab_raw[i] = reinterpret_cast<const uint4*>(table)[i];
    __syncthreads();
    tab = reinterpret_cast<const dpt_prep_entry*>(tab_raw);
  }
  const int G = static_cast<int>(gridDim.x);
  long long offset = 0;  // global index of this entry's first work unit
  for (int ei = 0; ei < n_entries; ++ei) {
    const dpt_prep_entry e = tab[ei];
    const long long pairs = static_cast<long long>(e.cout) * e.cin_pad;
    bf16* wf = static_cast<bf16*>(e.wf);
    bf16* wd = static_cast<bf16*>(e.wd);
    const int first = static_cast<int>((static_cast<long long>(blockIdx.x) - offset % G + G) % G);
    const bool w_vec = (reinterpret_cast<uintptr_t>(e.w) & 15) == 0;  // the tiled paths read the source as float4
    if (e.kind == 0 && e.cout % 32 == 0 && e.cin % 32 == 0 && e.cin == e.cin_pad && w_vec) {
      // Tiled path: a 32(co) x 32(ci) x 9 tile goes through shared memory so that the source rows (288 contiguous
      // floats per co) and both destinations (ci-contiguous wf, co-contiguous wd) are accessed in full sectors.  The
      // thread-per-(co,ci) form below writes wd with a stride of cout elements.
      const int tiles_ci = e.cin / 32;
      const int tiles = (e.cout / 32) * tiles_ci;
      const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
      for (int tile = first; tile < tiles; tile += G) {
        const int co0 = (tile / tiles_ci) * 32, ci0 = (tile % tiles_ci) * 32;
        __syncthreads();
        {  // 32 rows of 288 contiguous floats = 72 float4 each (16-byte aligned: ci0 * 9 floats is a multiple of 288)
          float4 v[9];
#pragma unroll
          for (int k = 0; k < 9; ++k) {
            const int idx = threadIdx.x + k * 256, r = idx / 72, c4 = idx - r * 72;
            v[k] = *reinterpret_cast<const float4*>(e.w + (static_cast<long long>(co0 + r) * e.cin + ci0) * 9 + 4 * c4);
          }
#pragma unroll
          for (int k = 0; k < 9; ++k) {
            const int idx = threadIdx.x + k * 256, r = idx / 72, c4 = idx - r * 72;
            const float os = e.oscale != nullptr ? e.oscale[co0 + r] : 1.f;
            __nv_bfloat162* d = reinterpret_cast<__nv_bfloat162*>(&sh[r][4 * c4]);
            d[0] = __floats2bfloat162_rn(os * v[k].x, os * v[k].y);
            d[1] = __floats2bfloat162_rn(os * v[k].z, os * v[k].w);
          }
        }
        __syncthreads();
        for (int pr = wid; pr < 288; pr += 8) {
          const int t = pr >> 5, l = pr & 31;
          wf[(static_cast<long long>(t) * e.cout + co0 + l) * e.cin_pad + ci0 + lane] = sh[l][lane * 9 + t];
          if (wd != nullptr) wd[(static_cast<long long>(8 - t) * e.cin + ci0 + l) * e.cout + co0 + lane] = sh[lane][l * 9 + t];
        }
      }
      offset += tiles;
      continue;
    }
    if (e.kind == 1 && e.cout % 32 == 0 && e.cin % 32 == 0 && w_vec) {
      // ConvTranspose weights [ci][co][2][2]: a 32(ci) x 32(co) x 4 tile; source rows are 128 contiguous floats per ci,
      // wf rows are ci-contiguous, wd rows co-contiguous (untiled, both the reads and the wd writes are fully scattered).
      const int tiles_co = e.cout / 32;
      const int tiles = (e.cin / 32) * tiles_co;
      const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
      for (int tile = first; tile < tiles; tile += G) {
        const int ci0 = (tile / tiles_co) * 32, co0 = (tile % tiles_co) * 32;
        __syncthreads();
        {  // 32 rows (ci) of 128 contiguous floats (co * 4 + q) = 32 float4 each
          float4 v[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int idx = threadIdx.x + k * 256, r = idx >> 5, c4 = idx & 31;
            v[k] = *reinterpret_cast<const float4*>(e.w + (static_cast<long long>(ci0 + r) * e.cout + co0) * 4 + 4 * c4);
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int idx = threadIdx.x + k * 256, r = idx >> 5, c4 = idx & 31;
            __nv_bfloat162* d = reinterpret_cast<__nv_bfloat162*>(&sh[r][4 * c4]);
            d[0] = __floats2bfloat162_rn(v[k].x, v[k].y);
            d[1] = __floats2bfloat162_rn(v[k].z, v[k].w);
          }
        }
        __syncthreads();
        for (int pr = wid; pr < 128; pr += 8) {
          const int q = pr >> 5, l = pr & 31;
          wf[(static_cast<long long>(q) * e.cout + co0 + l) * e.cin + ci0 + lane] = sh[lane][l * 4 + q];
          if (wd != nullptr) wd[(static_cast<long long>(q) * e.cin + ci0 + l) * e.cout + co0 + lane] = sh[l][lane * 4 + q];
        }
        if (ci0 == 0 && e.bias4 != nullptr && e.bias != nullptr && threadIdx.x < 128)
          e.bias4[(threadIdx.x >> 5) * e.cout + co0 + lane] = e.bias[co0 + lane];
      }
      offset += tiles;
      continue;
    }
    const long long units = (pairs + 255) / 256;
    for (long long u = first; u < units; u += G) {
      const long long i = u * 256 + threadIdx.x;
      if (i >= pairs) continue;
      const int ci = static_cast<int>(i % e.cin_pad);
      const int co = static_cast<int>(i / e.cin_pad);
      if (e.kind == 0) {
        const float os = e.oscale != nullptr ? e.oscale[co] : 1.f;
        for (int t = 0; t < 9; ++t) {
          const float v = ci < e.cin ? os * e.w[(static_cast<long long>(co) * e.cin + ci) * 9 + t] : 0.f;
          wf[(static_cast<long long>(t) * e.cout + co) * e.cin_pad + ci] = __float2bfloat16(v);
          if (wd != nullptr && ci < e.cin)
            wd[(static_cast<long long>(8 - t) * e.cin + ci) * e.cout + co] = __float2bfloat16(v);
        }
      } else {
        for (int q = 0; q < 4; ++q) {
          const float v = e.w[(static_cast<long long>(ci) * e.cout + co) * 4 + q];
          wf[(static_cast<long long>(q) * e.cout + co) * e.cin + ci] = __float2bfloat16(v);
          if (wd != nullptr) wd[(static_cast<long long>(q) * e.cin + ci) * e.cout + co] = __float2bfloat16(v);
        }
        if (ci == 0 && e.bias4 != nullptr && e.bias != nullptr) {
          for (int q = 0; q < 4; ++q) e.bias4[q * e.cout + co] = e.bias[co];
        }
      }
    }
    offset += units;
  }
}

}  // namespace dpt

using namespace dpt;

extern "C" const char* dpt_last_error(void) { return last_error_ref().c_str(); }
extern "C" int dpt_version(void) { return 100; }
extern "C" long long dpt_launch_count(void) { return g_launch_count.load(std::memory_order_relaxed); }

extern "C" int dpt_pack_input_ex(const void* src, int src_dtype, float scale, void* dst, int batch, int channels, int h, int w,
                                 void* stream) {
  if (channels < 1 || channels > 16) return fail("pack_input: channels=%d must be in 1..16", channels);
  const long long HW = static_cast<long long>(h) * w;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (src_dtype == 0) {
    DPT_CUDA_OK(launch_chained(pack_input_kernel<float>, dim3(grid_for(batch * HW, 256)), dim3(256), 0, st, static_cast<const float*>(src), static_cast<bf16*>(dst),
                                                                         batch, channels, HW, scale));
  } else if (src_dtype == 1) {
    DPT_CUDA_OK(launch_chained(pack_input_kernel<uint8_t>, dim3(grid_for(batch * HW, 256)), dim3(256), 0, st, static_cast<const uint8_t*>(src),
                                                                           static_cast<bf16*>(dst), batch, channels, HW, scale));
  } else {
    return fail("pack_input: src_dtype %d (0 = float32, 1 = uint8)", src_dtype);
  }
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}
extern "C" int dpt_pack_input(const float* src, void* dst, int batch, int channels, int h, int w, void* stream) {
  return dpt_pack_input_ex(src, 0, 1.f, dst, batch, channels, h, w, stream);
}

extern "C" int dpt_maxpool2x2(const void* x, void* y, int batch, int h, int w, int c, void* stream) {
  // odd h / w: the last row / column is dropped, as nn.MaxPool2d(2, 2) does (floor)
  if (c % 16 || h < 2 || w < 2) return fail("maxpool2x2: need c%%16==0 and h,w >= 2 (got %d,%d,%d)", c, h, w);
  const long long work = static_cast<long long>(batch) * (h / 2) * (w / 2) * (c / 16);
  DPT_CUDA_OK(launch_chained(maxpool2x2_kernel, dim3(grid_for(work, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream), static_cast<const bf16*>(x), static_cast<bf16*>(y), batch, h, w, c));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_pool_bwd_fused(const void* act, const void* d_pool, const void* d_skip, void* dz, int batch, int h,
                                  int w, int c, int relu_mask, void* stream) {
  if (c % 16 || h % 2 || w % 2) return fail("pool_bwd: need c%%16==0 and even h,w (got %d,%d,%d)", c, h, w);
  const long long work = static_cast<long long>(batch) * (h / 2) * (w / 2) * (c / 16);
  DPT_CUDA_OK(launch_chained(pool_bwd_fused_kernel, dim3(grid_for(work, 128)), dim3(128), 0, static_cast<cudaStream_t>(stream), static_cast<const bf16*>(act), static_cast<const bf16*>(d_pool), static_cast<const bf16*>(d_skip),
      static_cast<bf16*>(dz), batch, h, w, c, relu_mask));
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" size_t dpt_channel_sum_scratch_bytes(int c) { return static_cast<size_t>(kCsumBlocks) * c * sizeof(float); }

extern "C" int dpt_channel_sum_ex(const void* x, float* out, long long pixels, int c, float scale, int accumulate,
                                  void* scratch, size_t scratch_bytes, void* stream);
extern "C" int dpt_channel_sum(const void* x, float* out, long long pixels, int c, float scale, void* scratch,
                               size_t scratch_bytes, void* stream) {
  return dpt_channel_sum_ex(x, out, pixels, c, scale, 0, scratch, scratch_bytes, stream);
}
extern "C" int dpt_channel_sum_ex(const void* x, float* out, long long pixels, int c, float scale, int accumulate,
                                  void* scratch, size_t scratch_bytes, void* stream) {
  if (c % 2 || c > 2048) return fail("channel_sum: unsupported c=%d", c);
  if (scratch_bytes < dpt_channel_sum_scratch_bytes(c)) return fail("channel_sum: scratch too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (launch_channel_sum_stage1(static_cast<const bf16*>(x), static_cast<float*>(scratch), pixels, c, st)) return -1;
  DPT_CUDA_OK(launch_chained(channel_sum_stage2, dim3((c + 31) / 32), dim3(32 * kCsum2Lanes), 0, st, static_cast<const float*>(scratch), out, kCsumBlocks, c, scale,
                                                      accumulate));
  g_launch_count.fetch_add(2, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_prep_weights(const dpt_prep_entry* device_table, int n_entries, void* stream) {
  if (n_entries <= 0) return 0;
  DPT_CUDA_OK(launch_chained(prep_weights_kernel, dim3(6 * sm_count()), dim3(256), 0, static_cast<cudaStream_t>(stream), device_table, n_entries));  // latency-bound per block: keep every SM full
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Per-channel sum with library-owned scratch: the same two kernels as dpt_channel_sum, for hosts that are
// not allowed to allocate device memory themselves.
//
// Who needs this.  dpt_channel_sum takes its partial-sum buffer from the caller, which presumes the caller
// can get device memory from somewhere.  Some hosts cannot: an inference or control process whose device
// memory was carved up at start-up and whose policy forbids any later cudaMalloc (a fixed memory budget shared
// with other tenants of the GPU), a plug-in that runs inside somebody else's framework and has no access
// to that framework's allocator, or a latency-critical loop where the device-wide synchronisation a
// cudaMalloc / cudaFree implies would stall every other stream.  For them the partial sums live in the
// array below, which is part of the library image: every GPU that loads libdpt_b200.so carries these
// DPT_CHANNEL_SUM_STATIC_MAX_C * 296 floats (592 KiB) from load until the process exits.
//
// What the array's name means where.  In device code `g_csum_static` is the array.  In host code the same
// name is only a handle the CUDA runtime uses to find the symbol; passing it to a kernel as if it were a
// pointer hands the kernel a host address (a fault, or on machines where the GPU can reach host memory a
// silent write to the wrong place).  Host code must ask for the device address with cudaGetSymbolAddress,
// once per device with that device current, and pass THAT pointer.  This entry point does so on its first
// call on each device and remembers the answer; make that first call during set-up, before any stream
// capture is begun on the device.
//
// Sharing.  There is one array per GPU and every call on that GPU uses it.  It still holds the previous
// call's partial sums when the next call starts; that is harmless here because stage 1 overwrites every
// slot stage 2 reads (nothing has to start at zero, so nothing is cleared).  Two calls that are in flight
// on the same GPU at the same time - different streams, or different host threads - write the same slots
// and race: keep calls on one GPU ordered on one stream (or ordered by events).
// No cudaMalloc, cudaFree or other allocating routine is called here.
// ------------------------------------------------------------------------------------------------
namespace dpt {
static __device__ __align__(256) float g_csum_static[kCsumBlocks * DPT_CHANNEL_SUM_STATIC_MAX_C];

constexpr int kCsumMaxDevices = 64;
static float* g_csum_static_addr[kCsumMaxDevices] = {nullptr};
static std::mutex g_csum_static_mu;

// Device address of g_csum_static on the CURRENT device (looked up once per device).  The stream the caller passes must
// belong to that device - a launch on another device's stream fails anyway - and that is checked here, so the array of
// one GPU is never handed to a kernel that runs on another.
static int csum_static_scratch(float** out, cudaStream_t stream) {
  int dev = -1;
  DPT_CUDA_OK(cudaGetDevice(&dev));
  if (stream != nullptr && stream != cudaStreamLegacy && stream != cudaStreamPerThread) {
    int sdev = -1;
    if (cudaStreamGetDevice(stream, &sdev) == cudaSuccess && sdev != dev)
      return fail("channel_sum_static: the stream belongs to device %d but device %d is current; make the stream's device "
                  "current before the call", sdev, dev);
    cudaGetLastError();
  }
  if (dev < 0 || dev >= kCsumMaxDevices) return fail("channel_sum_static: device index %d unsupported", dev);
  std::lock_guard<std::mutex> lock(g_csum_static_mu);
  if (g_csum_static_addr[dev] == nullptr) {
    float* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_csum_static);
    if (e != cudaSuccess || ptr == nullptr)
      return fail("channel_sum_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
    g_csum_static_addr[dev] = ptr;
  }
  *out = g_csum_static_addr[dev];
  return 0;
}
}  // namespace dpt

extern "C" int dpt_channel_sum_static(const void* x, float* out, long long pixels, int c, float scale, void* stream) {
  if (c % 2 || c < 2) return fail("channel_sum_static: c=%d must be a positive even number", c);
  if (c > DPT_CHANNEL_SUM_STATIC_MAX_C)
    return fail("channel_sum_static: c=%d exceeds the built-in scratch (at most %d channels); use dpt_channel_sum "
                "with caller-provided scratch for wider tensors", c, DPT_CHANNEL_SUM_STATIC_MAX_C);
  if (pixels < 0) return fail("channel_sum_static: negative pixel count");
  float* scratch = nullptr;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (csum_static_scratch(&scratch, st)) return -1;
  if (launch_channel_sum_stage1(static_cast<const bf16*>(x), scratch, pixels, c, st)) return -1;
  DPT_CUDA_OK(launch_chained(channel_sum_stage2, dim3((c + 31) / 32), dim3(32 * kCsum2Lanes), 0, st, scratch, out, kCsumBlocks, c, scale, 0));
  g_launch_count.fetch_add(2, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}
