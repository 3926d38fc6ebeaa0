// Weight-gradient implicit GEMM on tcgen05 / TMEM (sm_100a), split-K over pixel tiles with a deterministic
// second-stage reduction.  Replaces the cuDNN backward-filter kernels autograd reaches for
// utils/train_utils.py:69 of the reference (23 ConvolutionBackward0 nodes per step).
//
//   dW[m][n][tap] = sum_pixels Z[pixel][m] * X[pixel + tap][n]
//
// Z is the un-shifted operand (3x3 conv: the output gradient dZ, m = C_out; ConvTranspose: the layer
// input, m = C_in) and X the shifted one (3x3 conv: the layer input, one 18x18 halo box serves all nine
// taps through shifted UMMA descriptors; ConvTranspose: the output gradient read through a stride-2 view,
// one box per tap).  The reduction dimension is the pixel index, which is the slow dimension of NHWC, so both
// operands are MN-major UMMA operands read in place from TMA-written boxes: no transposes anywhere.
//
// A CTA owns one (128 x 32) slab of (m, n) for all taps: taps*32 fp32 TMEM columns (288 of 512 for 3x3), and
// walks a strided subset of the 16x16-pixel tiles (split-K).  Its partial result goes to scratch memory;
// dpt_wgrad_reduce sums the partials in a fixed order (bit-reproducible) and scatters them into the fp32
// gradient in PyTorch's parameter layout.
//
// Warps: 0 = TMA producer, 1 = TMEM owner + MMA issuer, 2..5 = epilogue (idle until the last tile retires).
#include <atomic>
#include <memory>
#include <new>

#include "../../include/dpt_b200.h"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dpt {

extern std::atomic<long long> g_launch_count;

constexpr int kWgThreads = 192;
constexpr int kWgTile = 16;  // pixel tile is 16 x 16
constexpr int kWgMaxStages = 3;
constexpr int kWgSmemBudget = 216 * 1024;

enum { WG_HALO = DPT_CONV_3X3, WG_TAPBOX = DPT_CONV_3X3_TAPBOX, WG_STRIDED = DPT_CONV_2X2S2 };

struct WgradParams {
  CUtensorMap mapZ;
  CUtensorMap mapX[2];
  int mode;
  int B, PH, PW;
  int tiles_x, tiles_y, total_ptiles;
  int m_tiles, n_tiles, ksplit;
  int z_box_ch, z_boxes, z_row_bytes, z_box_bytes;  // z_boxes boxes of z_box_ch channels per M tile
  int z_lbo;                                         // byte stride between consecutive M atoms (0 = alias)
  int x_ch, x_row_bytes, x_box_bytes, x_box_tx;      // x_ch = N tile (32 or 16)
  int x_tiles0;                                      // N tiles that come from source 0
  int taps, nboxes, taps_per_box, pitch;
  int stage_bytes, stage_tx, nstages;
  int tmem_cols;
  float* partial;  // [ksplit][n_tiles][m_tiles][128][taps * x_ch]
};

template <int MODE>
__global__ void __launch_bounds__(kWgThreads, 1) igemm_wgrad_kernel(const __grid_constant__ WgradParams p) {
  constexpr int TAPS = MODE == WG_STRIDED ? 4 : 9;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full[kWgMaxStages], empty[kWgMaxStages];
  __shared__ uint64_t acc_full;
  __shared__ uint32_t tmem_base_slot;

  const int warp = warp_idx_sync();
  const int lane = threadIdx.x & 31;
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t s_addr0 = (raw_addr + 1023u) & ~1023u;
  uint8_t* const s_ptr0 = smem_raw + (s_addr0 - raw_addr);

  // CTA -> (m tile, n tile, split)
  int id = blockIdx.x;
  const int m_tile = id % p.m_tiles;
  id /= p.m_tiles;
  const int n_tile = id % p.n_tiles;
  const int split = id / p.n_tiles;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kWgMaxStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(&acc_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(&tmem_base_slot, static_cast<uint32_t>(p.tmem_cols));
    tmem_relinquish();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.mapZ);
    tma_prefetch_desc(&p.mapX[n_tile >= p.x_tiles0 ? 1 : 0]);
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_slot, 0);
  const int my_tiles = split < p.total_ptiles ? (p.total_ptiles - split + p.ksplit - 1) / p.ksplit : 0;

  if (warp == 0) {
    {
      const int xsrc = n_tile >= p.x_tiles0 ? 1 : 0;
      const int xc0 = (n_tile - (xsrc ? p.x_tiles0 : 0)) * p.x_ch;
      const int zc0 = m_tile * 128;
      uint32_t it = 0;
      for (int pt = split; pt < p.total_ptiles; pt += p.ksplit, ++it) {
        int t = pt;
        const int tx = t % p.tiles_x;
        t /= p.tiles_x;
        const int ty = t % p.tiles_y;
        const int img = t / p.tiles_y;
        const int x0 = tx * kWgTile, y0 = ty * kWgTile;
        const uint32_t slot = it % static_cast<uint32_t>(p.nstages);
        const uint32_t ph = (it / static_cast<uint32_t>(p.nstages)) & 1u;
        mbar_wait(&empty[slot], ph ^ 1u, 11);
        if (elect_one_sync()) {
          mbar_arrive_expect_tx(&full[slot], static_cast<uint32_t>(p.stage_tx));
          uint8_t* st = s_ptr0 + slot * p.stage_bytes;
          for (int zb = 0; zb < p.z_boxes; ++zb)
            tma_load_5d(st + zb * p.z_box_bytes, &p.mapZ, &full[slot], zc0 + zb * p.z_box_ch, x0, y0, img, 0);
          uint8_t* xs = st + p.z_boxes * p.z_box_bytes;
          for (int bx = 0; bx < p.nboxes; ++bx) {
            uint8_t* dst = xs + bx * p.x_box_bytes;
            if (p.mode == WG_HALO) {
              tma_load_5d(dst, &p.mapX[xsrc], &full[slot], xc0, x0 - 1, y0 - 1, img, 0);
            } else if (p.mode == WG_TAPBOX) {
              tma_load_5d(dst, &p.mapX[xsrc], &full[slot], xc0, x0 + (bx % 3) - 1, y0 + (bx / 3) - 1, img, 0);
            } else {  // strided 2x2: dims (C, dx, x, dy, b*PH + y)
              tma_load_5d(dst, &p.mapX[xsrc], &full[slot], xc0, bx & 1, x0, bx >> 1, img * p.PH + y0);
            }
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // MMA issuer: warp-uniform loop, one elected lane issues.  The tap loop is unrolled (TAPS is a template
    // parameter) and every descriptor is a running value plus a precomputed per-tap offset, so the uniform
    // datapath spends ~2 instructions per tcgen05.mma.
    {
      const uint32_t idesc = umma_idesc_bf16(p.x_ch, 1, 1);  // both operands MN-major
      const uint64_t z_desc0 = umma_smem_desc(0u, static_cast<uint32_t>(p.z_lbo), static_cast<uint32_t>(8 * p.z_row_bytes),
                                              umma_layout_type(p.z_row_bytes));
      const uint64_t x_desc0 =
          umma_smem_desc(0u, 16u, static_cast<uint32_t>(8 * p.x_row_bytes), umma_layout_type(p.x_row_bytes));
      uint32_t tap_off[TAPS];  // 16-byte units
#pragma unroll
      for (int t = 0; t < TAPS; ++t)
        tap_off[t] = MODE == WG_HALO ? static_cast<uint32_t>(((t / 3) * p.pitch + (t % 3)) * p.x_row_bytes) >> 4
                                     : static_cast<uint32_t>(t * p.x_box_bytes) >> 4;
      const uint32_t z_step = static_cast<uint32_t>(kWgTile * p.z_row_bytes) >> 4;
      const uint32_t x_step = static_cast<uint32_t>(p.pitch * p.x_row_bytes) >> 4;
      const uint32_t xch = static_cast<uint32_t>(p.x_ch);
      uint32_t slot = 0, ph = 0;
      bool fresh = true;
      for (int pt = split; pt < p.total_ptiles; pt += p.ksplit) {
        mbar_wait(&full[slot], ph, 12);
        tc_fence_after_sync();
        const uint32_t z_base = s_addr0 + slot * static_cast<uint32_t>(p.stage_bytes);
        const uint32_t x_base = z_base + static_cast<uint32_t>(p.z_boxes * p.z_box_bytes);
        if (elect_one_sync()) {
          uint64_t dz = z_desc0 + (z_base >> 4);
          uint64_t dx = x_desc0 + (x_base >> 4);
          int r = 0;
          if (fresh) {  // very first K step of this CTA overwrites the accumulators
#pragma unroll
            for (int t = 0; t < TAPS; ++t) umma_bf16(tmem_base + t * xch, dz, dx + tap_off[t], idesc, 0u);
            dz += z_step;
            dx += x_step;
            r = 1;
          }
#pragma unroll 1
          for (; r < kWgTile; ++r) {  // one image row of 16 pixels = one K=16 step
#pragma unroll
            for (int t = 0; t < TAPS; ++t) umma_bf16(tmem_base + t * xch, dz, dx + tap_off[t], idesc, 1u);
            dz += z_step;
            dx += x_step;
          }
          umma_commit(&empty[slot]);
        }
        __syncwarp();
        fresh = false;
        if (++slot == static_cast<uint32_t>(p.nstages)) {
          slot = 0;
          ph ^= 1u;
        }
      }
      if (elect_one_sync()) umma_commit(&acc_full);
      __syncwarp();
    }
  } else {
    // epilogue: TMEM -> fp32 partial slab in scratch
    const int q = warp & 3;
    const int row = q * 32 + lane;
    if (my_tiles > 0) {
      mbar_wait(&acc_full, 0, 13);
      tc_fence_after_sync();
    }
    const int cols = p.taps * p.x_ch;
    float* out = p.partial + ((static_cast<size_t>(split) * p.n_tiles + n_tile) * p.m_tiles + m_tile) * 128 * cols +
                 static_cast<size_t>(row) * cols;
    for (int c = 0; c < cols; c += 16) {
      uint32_t v[16];
      if (my_tiles > 0) {
        tmem_ld_32x16(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(c), v);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = 0u;
      }
      float4* o4 = reinterpret_cast<float4*>(out + c);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        o4[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]), __uint_as_float(v[4 * i + 2]),
                            __uint_as_float(v[4 * i + 3]));
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc(tmem_base, static_cast<uint32_t>(p.tmem_cols));
  }
}

// Second stage: dW[m*sm + n*sn + t*st] = scale * sum_s partial[s][n_tile][m_tile][m%128][t*x_ch + n%x_ch]
struct WgradReduceParams {
  const float* partial;
  float* dw;
  int M, N, taps, x_ch, m_tiles, n_tiles, ksplit;
  long long sm, sn, st;
  float scale;
  int accumulate;
};
__global__ void wgrad_reduce_kernel(const WgradReduceParams p) {
  const long long total = static_cast<long long>(p.M) * p.n_tiles * p.x_ch * p.taps;
  const int cols = p.taps * p.x_ch;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    // order the index so neighbouring threads read neighbouring partial columns: (m, n_tile, tap, n_in)
    long long r = i;
    const int n_in = static_cast<int>(r % p.x_ch);
    r /= p.x_ch;
    const int tap = static_cast<int>(r % p.taps);
    r /= p.taps;
    const int n_tile = static_cast<int>(r % p.n_tiles);
    const int m = static_cast<int>(r / p.n_tiles);
    const int n = n_tile * p.x_ch + n_in;
    if (n >= p.N) continue;
    const int m_tile = m >> 7, m_in = m & 127;
    const size_t slab = static_cast<size_t>(128) * cols;
    const float* src = p.partial + (static_cast<size_t>(n_tile) * p.m_tiles + m_tile) * slab +
                       static_cast<size_t>(m_in) * cols + tap * p.x_ch + n_in;
    const size_t split_stride = static_cast<size_t>(p.n_tiles) * p.m_tiles * slab;
    float acc = 0.f;
    for (int s = 0; s < p.ksplit; ++s) acc += src[s * split_stride];
    float* dst = p.dw + (m * p.sm + n * p.sn + tap * p.st);
    *dst = p.accumulate ? *dst + acc * p.scale : acc * p.scale;
  }
}

}  // namespace dpt

using WgradKernelFn = void (*)(const dpt::WgradParams);

struct dpt_wgrad_plan {
  WgradKernelFn fn;
  dpt::WgradParams p;
  dpt::WgradReduceParams r;
  int grid;
  size_t smem;
  size_t scratch_bytes;
};

extern "C" int dpt_wgrad_plan_create(const dpt_wgrad_desc* d, dpt_wgrad_plan** out) {
  using namespace dpt;
  if (!d || !out) return fail("dpt_wgrad_plan_create: null argument");
  if (d->mode != WG_HALO && d->mode != WG_TAPBOX && d->mode != WG_STRIDED) return fail("wgrad: bad mode %d", d->mode);
  if (d->num_x < 1 || d->num_x > 2) return fail("wgrad: num_x must be 1 or 2");
  const int xc0 = d->x_c[0], xc1 = d->num_x == 2 ? d->x_c[1] : 0;
  if (xc0 % 16 || xc1 % 16 || d->z_c % 32) return fail("wgrad: channels x=(%d,%d) z=%d unsupported", xc0, xc1, d->z_c);
  const int x_ch = (xc0 % 32 == 0 && xc1 % 32 == 0) ? 32 : 16;

  std::unique_ptr<dpt_wgrad_plan> owner(new (std::nothrow) dpt_wgrad_plan());
  dpt_wgrad_plan* plan = owner.get();
  if (!plan) return fail("wgrad: out of host memory");
  WgradParams& p = plan->p;
  memset(&p, 0, sizeof(p));
  p.mode = d->mode;
  p.B = d->batch;
  p.PH = d->h;
  p.PW = d->w;
  p.tiles_x = ceil_div(d->w, kWgTile);
  p.tiles_y = ceil_div(d->h, kWgTile);
  p.total_ptiles = d->batch * p.tiles_x * p.tiles_y;
  p.taps = d->mode == WG_STRIDED ? 4 : 9;
  p.nboxes = d->mode == WG_HALO ? 1 : p.taps;
  p.taps_per_box = p.taps / p.nboxes;
  p.pitch = d->mode == WG_HALO ? kWgTile + 2 : kWgTile;
  p.x_ch = x_ch;
  p.x_row_bytes = x_ch * 2;
  p.x_box_tx = (d->mode == WG_HALO ? (kWgTile + 2) * (kWgTile + 2) : kWgTile * kWgTile) * p.x_row_bytes;
  p.x_box_bytes = (p.x_box_tx + 1023) & ~1023;
  p.x_tiles0 = xc0 / x_ch;
  p.n_tiles = (xc0 + xc1) / x_ch;
  p.m_tiles = ceil_div(d->z_c, 128);
  p.z_box_ch = d->z_c >= 64 ? 64 : 32;
  p.z_row_bytes = p.z_box_ch * 2;
  p.z_boxes = d->z_c >= 128 ? 2 : 1;
  p.z_box_bytes = kWgTile * kWgTile * p.z_row_bytes;  // 256 rows: already a multiple of 1024
  p.z_lbo = p.z_boxes == 2 ? p.z_box_bytes : 0;
  p.stage_tx = p.z_boxes * p.z_box_bytes + p.nboxes * p.x_box_tx;
  p.stage_bytes = p.z_boxes * p.z_box_bytes + p.nboxes * p.x_box_bytes;
  p.nstages = kWgSmemBudget / p.stage_bytes;
  if (p.nstages > kWgMaxStages) p.nstages = kWgMaxStages;
  if (p.nstages < 1) return fail("wgrad: stage of %d bytes does not fit", p.stage_bytes);
  int cols = 32;
  while (cols < p.taps * x_ch) cols <<= 1;
  p.tmem_cols = cols;

  const int sms = sm_count();
  int ksplit = d->ksplit > 0 ? d->ksplit : ceil_div(2 * sms, p.m_tiles * p.n_tiles);
  if (ksplit > p.total_ptiles) ksplit = p.total_ptiles;
  if (ksplit < 1) ksplit = 1;
  p.ksplit = ksplit;
  plan->grid = p.m_tiles * p.n_tiles * ksplit;
  plan->scratch_bytes = static_cast<size_t>(plan->grid) * 128 * p.taps * x_ch * sizeof(float);
  if (d->scratch == nullptr || d->scratch_bytes < plan->scratch_bytes) {
    const size_t need = plan->scratch_bytes;
    return fail("wgrad: scratch of %zu bytes needed, %zu given", need, d->scratch_bytes);
  }
  p.partial = static_cast<float*>(d->scratch);

  // tensor maps
  {
    const uint64_t C = d->z_c;
    uint64_t dims[5] = {C, (uint64_t)d->w, (uint64_t)d->h, (uint64_t)d->batch, 1};
    uint64_t str[4] = {C * 2, C * 2 * d->w, C * 2 * d->w * d->h, C * 2 * d->w * d->h * d->batch};
    uint32_t box[5] = {(uint32_t)p.z_box_ch, kWgTile, kWgTile, 1, 1};
    if (make_tmap_bf16(&p.mapZ, d->z, 5, dims, str, box, p.z_row_bytes)) return -1;
  }
  for (int i = 0; i < d->num_x; ++i) {
    const uint64_t C = d->x_c[i];
    uint64_t dims[5], str[4];
    uint32_t box[5];
    if (d->mode == WG_STRIDED) {
      dims[0] = C; dims[1] = 2; dims[2] = d->w; dims[3] = 2; dims[4] = (uint64_t)d->batch * d->h;
      str[0] = C * 2; str[1] = 2 * C * 2; str[2] = 2ull * d->w * C * 2; str[3] = 2 * str[2];
      box[0] = x_ch; box[1] = 1; box[2] = kWgTile; box[3] = 1; box[4] = kWgTile;
    } else {
      dims[0] = C; dims[1] = d->w; dims[2] = d->h; dims[3] = d->batch; dims[4] = 1;
      str[0] = C * 2; str[1] = str[0] * d->w; str[2] = str[1] * d->h; str[3] = str[2] * d->batch;
      box[0] = x_ch; box[1] = p.pitch; box[2] = p.pitch; box[3] = 1; box[4] = 1;
    }
    if (make_tmap_bf16(&p.mapX[i], d->x[i], 5, dims, str, box, p.x_row_bytes)) return -1;
  }

  WgradReduceParams& r = plan->r;
  r.partial = p.partial;
  r.dw = d->dw;
  r.M = d->z_c;
  r.N = d->n_valid > 0 ? d->n_valid : xc0 + xc1;
  r.taps = p.taps;
  r.x_ch = x_ch;
  r.m_tiles = p.m_tiles;
  r.n_tiles = p.n_tiles;
  r.ksplit = ksplit;
  r.sm = d->dw_stride_m;
  r.sn = d->dw_stride_n;
  r.st = d->dw_stride_t;
  r.scale = d->scale == 0.f ? 1.f : d->scale;
  r.accumulate = d->accumulate;
  plan->smem = static_cast<size_t>(p.nstages) * p.stage_bytes + 1024;
  plan->fn = d->mode == WG_HALO ? igemm_wgrad_kernel<WG_HALO>
             : d->mode == WG_TAPBOX ? igemm_wgrad_kernel<WG_TAPBOX> : igemm_wgrad_kernel<WG_STRIDED>;
  const cudaError_t e = cudaFuncSetAttribute(plan->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
  if (e != cudaSuccess) {
    return fail("cudaFuncSetAttribute(wgrad smem) failed: %s", cudaGetErrorString(e));
  }
  *out = owner.release();
  return 0;
}

extern "C" size_t dpt_wgrad_scratch_bytes(const dpt_wgrad_desc* d) {
  using namespace dpt;
  if (!d) return 0;
  const int xc0 = d->x_c[0], xc1 = d->num_x == 2 ? d->x_c[1] : 0;
  const int x_ch = (xc0 % 32 == 0 && xc1 % 32 == 0) ? 32 : 16;
  const int taps = d->mode == WG_STRIDED ? 4 : 9;
  const int n_tiles = (xc0 + xc1) / x_ch;
  const int m_tiles = ceil_div(d->z_c, 128);
  const int total_ptiles = d->batch * ceil_div(d->w, kWgTile) * ceil_div(d->h, kWgTile);
  int ksplit = d->ksplit > 0 ? d->ksplit : ceil_div(2 * sm_count(), m_tiles * n_tiles);
  if (ksplit > total_ptiles) ksplit = total_ptiles;
  if (ksplit < 1) ksplit = 1;
  return static_cast<size_t>(m_tiles) * n_tiles * ksplit * 128 * taps * x_ch * sizeof(float);
}

extern "C" int dpt_wgrad_plan_launch(const dpt_wgrad_plan* plan, void* stream) {
  using namespace dpt;
  if (!plan) return fail("dpt_wgrad_plan_launch: null plan");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  plan->fn<<<plan->grid, kWgThreads, plan->smem, st>>>(plan->p);
  DPT_CUDA_OK(cudaGetLastError());
  const long long total = static_cast<long long>(plan->r.M) * plan->r.n_tiles * plan->r.x_ch * plan->r.taps;
  int blocks = static_cast<int>((total + 255) / 256);
  if (blocks > 4 * sm_count()) blocks = 4 * sm_count();
  wgrad_reduce_kernel<<<blocks, 256, 0, st>>>(plan->r);
  g_launch_count.fetch_add(2, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" void dpt_wgrad_plan_destroy(dpt_wgrad_plan* plan) { delete plan; }
