// Weight-gradient implicit GEMM on tcgen05 / TMEM (sm_100a), split-K over pixel tiles with a deterministic
// second-stage reduction.  Replaces the cuDNN backward-filter kernels autograd reaches for
// utils/train_utils.py:69 of the reference (23 ConvolutionBackward0 nodes per step).
//
//   dW[m][n][tap] = sum_pixels Z[pixel][m] * X[pixel + tap][n]
//
// Z is the un-shifted ("plain") operand (3x3 conv: the output gradient dZ, m = C_out; ConvTranspose: the layer
// input, m = C_in) and X the shifted ("stacked") one (3x3 conv: the layer input; ConvTranspose: the output
// gradient read through a stride-2 view, one box per tap).  The reduction dimension is the pixel index, the slow
// dimension of NHWC, so both operands are MN-major UMMA operands read in place from TMA-written boxes.
//
// Tap stacking.  X is cut into 32-channel tiles (64-byte rows, SWIZZLE_64B).  An MN-major UMMA operand of
// M = 128 is four 32-channel atoms LBO bytes apart - and nothing says the four atoms must be different channels:
//   * 3x3: LBO = one pixel (64 bytes), so atom j is the SAME 32 channels shifted j pixels to the right: one MMA
//     covers the taps kw = 0,1,2 (atom 3 is a harmless extra); the three kh are three MMAs on three accumulators;
//   * ConvTranspose 2x2: LBO = one box, so atom j is tap j's box: one MMA covers all four taps.
// The accumulator is D[row = tap_in_group*32 + x_channel][col = z_channel], N = up to 128 z channels per CTA.
// Versus one MMA per tap with N = 32 this issues 3x (3x3) / 4x (convT) fewer MMAs and uses 96 / 128 of the M
// rows even for 32-channel layers (the un-stacked form padded M from 32 to 128 and was shared-memory bound).
//
// Nine-tap form (3x3 layers whose output channel count is 32 or a multiple of 64).  The Z operand can be stacked the
// same way along N: its atoms (nt <= 64 channels) are LBO = one image row of the tile apart, so atom j is the Z tile
// shifted by j rows.  Row step r = -1..16 pairs X box row r+1 with Z rows r-1, r, r+1 and ONE tcgen05.mma
// (M = 128 = kw x channels, N = 3 nt = kh x channels) covers all nine taps; Z rows outside the tile must read as zero,
// so each stage keeps two zeroed image rows on both sides of where TMA writes the tile.  18 MMAs per tile instead
// of 48, 7 KB instead of 15 KB of operand reads per row on the 32-channel layers (a tcgen05.mma costs
// (128 + N) x 32 B / 128 B per cycle below N = 128, profiles/mma_cost_b200.jsonl), and 57 KB stages (three in flight)
// instead of 85 KB ones on the deep layers.
//
// A CTA owns one (x tile, z tile) pair and walks a strided subset of the 16x16-pixel tiles (split-K, sized to ONE wave:
// one CTA per SM); partial accumulators go to scratch, a second kernel sums them in a fixed order (bit-reproducible)
// and scatters them into the fp32 gradient in PyTorch's parameter layout (optionally accumulating, for micro-batches).
//
// Bias gradient.  db[c] = sum over pixels of dZ[pixel][c] (ConvTranspose: of the X operand).  The CTAs of x tile 0
// (ConvTranspose: z tile 0) see every such tile once and their four epilogue warps idle until the last MMA retires:
// they add up the columns of each stage straight from the TMA-written shared memory (the swizzle is undone by giving
// each thread rows of one swizzle class), write per-split partials next to the weight partials, and the reduce kernel
// finishes both.  No kernel re-reads dZ from HBM for the bias.
//
// Warps: 0 = TMA producer, 1 = TMEM owner + MMA issuer, 2..5 = bias column sums during the main loop, then epilogue.
#include <atomic>
#include <cstdlib>
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
constexpr int kWgSmemBudget = 206 * 1024;  // + 8.5 KB static for the fused bias-gradient reduction

enum { WG_HALO = DPT_CONV_3X3, WG_STRIDED = DPT_CONV_2X2S2 };

struct WgradParams {
  CUtensorMap mapZ;
  CUtensorMap mapX[2];
  int mode;
  int B, PH, PW;
  int tiles_x, tiles_y, total_ptiles;
  int m_tiles, n_tiles, ksplit;                      // m: z-channel tiles (GEMM N), n: x-channel tiles (GEMM M)
  int nt;                                            // z channels per CTA (GEMM N): 32..128
  int z_box_ch, z_boxes, z_row_bytes, z_box_bytes;   // z_boxes boxes of z_box_ch channels per z tile
  int x_ch, x_row_bytes, x_box_bytes, x_box_tx;      // x_ch = 32 (16 for the zero-padded first layer)
  int x_lbo;                                         // byte distance between stacked atoms
  int x_tiles0;                                      // x tiles that come from source 0
  int groups, nboxes, pitch;                         // accumulator groups (3x3: kh = 3, convT: 1)
  int stage_bytes, stage_tx, nstages;
  int tmem_cols;
  int z_pad_bytes;     // 9-tap form: two zeroed image rows before and after the Z tile inside its slot
  int z_region_bytes;  // Z part of a stage (boxes + pads); the X boxes follow
  float* partial;  // [ksplit][n_tiles][m_tiles][128][groups * nt]
  float* bias_partial;  // [ksplit][z channels] column sums of Z (the bias gradient of a 3x3 conv), or NULL
};

template <int GROUPS>
__global__ void __launch_bounds__(kWgThreads, 1) igemm_wgrad_kernel(const __grid_constant__ WgradParams p) {
  grid_dep_launch();  // the next kernel's prologue may overlap this grid's tail (ptx.cuh)
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full[kWgMaxStages], empty[kWgMaxStages];
  __shared__ uint64_t acc_full;
  __shared__ uint32_t tmem_base_slot;
  __shared__ float bred[128][17];  // fused bias gradient: per-thread column sums before the fixed-order reduction

  const int warp = warp_idx_sync();
  const int lane = threadIdx.x & 31;
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t s_addr0 = (raw_addr + 1023u) & ~1023u;
  uint8_t* const s_ptr0 = smem_raw + (s_addr0 - raw_addr);

  // CTA -> (z tile, x tile, split)
  int id = blockIdx.x;
  const int m_tile = id % p.m_tiles;
  id /= p.m_tiles;
  const int n_tile = id % p.n_tiles;
  const int split = id / p.n_tiles;

  // Fused bias gradient: db[c] = sum over pixels of Z[pixel][c].  The CTAs of x tile 0 see every Z tile of their
  // (z tile, split) exactly once, and their four epilogue warps idle until the last MMA retires: they add up the
  // columns of each Z stage straight from the TMA-written shared memory (+7 % shared-memory reads on the 32-channel
  // layers) instead of a separate kernel re-reading dZ from HBM (0.53 ms of the 5.2 ms step).
  // ConvTranspose form: the bias gradient is the column sum of the X operand (the output gradient, one box per tap), and
  // the CTAs of z tile 0 see every X tile of their (x tile, split) once.
  const bool bias_cta = p.bias_partial != nullptr && (GROUPS == 1 ? m_tile == 0 : n_tile == 0);
  if (threadIdx.x == 0) {
    for (int i = 0; i < kWgMaxStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], bias_cta ? 5 : 1);  // the MMA commit, plus one arrival per column-summing warp
    }
    mbar_init(&acc_full, 1);
    fence_mbar_init();
  }
  if (GROUPS == 9) {
    // 9-tap form: the Z operand of row step r is the three image rows r-1, r, r+1 of the tile; rows outside the tile
    // must read as zero, so every stage keeps two zero rows on each side of where TMA writes the tile.
    const int n16 = p.z_pad_bytes >> 4;
    for (int st = 0; st < p.nstages; ++st) {
      uint4* head = reinterpret_cast<uint4*>(s_ptr0 + st * p.stage_bytes);
      uint4* tail = reinterpret_cast<uint4*>(s_ptr0 + st * p.stage_bytes + p.z_pad_bytes + p.z_boxes * p.z_box_bytes);
      for (int i = threadIdx.x; i < n16; i += kWgThreads) {
        head[i] = make_uint4(0u, 0u, 0u, 0u);
        tail[i] = make_uint4(0u, 0u, 0u, 0u);
      }
    }
    fence_proxy_async_smem();  // the tensor core reads shared memory through the async proxy
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
  grid_dep_wait();  // barriers, TMEM and the zeroed pad rows are set up: now the previous kernel's output is needed
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_slot, 0);
  const int my_tiles = split < p.total_ptiles ? (p.total_ptiles - split + p.ksplit - 1) / p.ksplit : 0;

  if (warp == 0) {
    // ===================================== TMA producer ==========================================
    const int xsrc = n_tile >= p.x_tiles0 ? 1 : 0;
    const int xc0 = (n_tile - (xsrc ? p.x_tiles0 : 0)) * p.x_ch;
    const int zc0 = m_tile * p.nt;
    uint32_t slot = 0, ph = 0;
    for (int pt = split; pt < p.total_ptiles; pt += p.ksplit) {
      int t = pt;
      const int tx = t % p.tiles_x;
      t /= p.tiles_x;
      const int ty = t % p.tiles_y;
      const int img = t / p.tiles_y;
      const int x0 = tx * kWgTile, y0 = ty * kWgTile;
      mbar_wait(&empty[slot], ph ^ 1u, 11);
      if (elect_one_sync()) {
        mbar_arrive_expect_tx(&full[slot], static_cast<uint32_t>(p.stage_tx));
        uint8_t* st = s_ptr0 + slot * p.stage_bytes;
        for (int zb = 0; zb < p.z_boxes; ++zb)
          tma_load_5d(st + p.z_pad_bytes + zb * p.z_box_bytes, &p.mapZ, &full[slot], zc0 + zb * p.z_box_ch, x0, y0, img, 0);
        uint8_t* xs = st + p.z_region_bytes;
        if (p.mode == WG_HALO) {
          tma_load_5d(xs, &p.mapX[xsrc], &full[slot], xc0, x0 - 1, y0 - 1, img, 0);
        } else {  // strided 2x2: dims (C, dx, x, dy, b*PH + y), one box per tap
          for (int bx = 0; bx < 4; ++bx)
            tma_load_5d(xs + bx * p.x_box_bytes, &p.mapX[xsrc], &full[slot], xc0, bx & 1, x0, bx >> 1, img * p.PH + y0);
        }
      }
      __syncwarp();
      if (++slot == static_cast<uint32_t>(p.nstages)) {
        slot = 0;
        ph ^= 1u;
      }
    }
  } else if (warp == 1) {
    // ===================================== MMA issuer ============================================
    // Warp-uniform loop, one elected lane issues; the group loop is unrolled and every descriptor is a running
    // value plus a constant, so the uniform datapath spends ~2 instructions per tcgen05.mma.
    // both operands MN-major, M = 128; N = nt, or 3 * nt in the 9-tap form (three row-shifted copies of the Z tile)
    const uint32_t idesc = umma_idesc_bf16(GROUPS == 9 ? 3 * p.nt : p.nt, 1, 1);
    const uint64_t x_desc0 = umma_smem_desc(0u, static_cast<uint32_t>(p.x_lbo), static_cast<uint32_t>(8 * p.x_row_bytes),
                                            umma_layout_type(p.x_row_bytes));
    // Z atoms (nt-channel groups along N): the next 64-channel box, or - 9-tap form - the next image row of the tile
    const uint64_t z_desc0 = umma_smem_desc(0u, static_cast<uint32_t>(GROUPS == 9 ? kWgTile * p.z_row_bytes : p.z_box_bytes),
                                            static_cast<uint32_t>(8 * p.z_row_bytes), umma_layout_type(p.z_row_bytes));
    const uint32_t z_step = static_cast<uint32_t>(kWgTile * p.z_row_bytes) >> 4;   // next image row of the tile
    const uint32_t x_step = static_cast<uint32_t>(p.pitch * p.x_row_bytes) >> 4;   // next image row of the box
    const uint32_t nt = static_cast<uint32_t>(p.nt);
    uint32_t slot = 0, ph = 0;
    bool fresh = true;
    for (int pt = split; pt < p.total_ptiles; pt += p.ksplit) {
      mbar_wait(&full[slot], ph, 12);
      tc_fence_after_sync();
      const uint32_t z_base = s_addr0 + slot * static_cast<uint32_t>(p.stage_bytes);
      const uint32_t x_base = z_base + static_cast<uint32_t>(p.z_region_bytes);
      if (elect_one_sync()) {
        uint64_t dz = z_desc0 + (z_base >> 4);
        uint64_t dx = x_desc0 + (x_base >> 4);
        if (GROUPS == 9) {
          // Row step r = -1..16 pairs X box row r+1 (atoms = kw shifts) with Z rows r-1, r, r+1 (atoms, zero outside the
          // tile): ONE tcgen05.mma per step covers all nine taps (column group j <-> kh = 2 - j).  18 MMAs of
          // M128 x N(3 nt) per tile instead of 48 of M128 x N(nt): 7 KB instead of 15 KB of shared-memory operand
          // reads per step on the 32-channel layers, which were bound by exactly that.
          umma_bf16(tmem_base, dx, dz, idesc, fresh ? 0u : 1u);
#pragma unroll 1
          for (int r = 1; r < kWgTile + 2; ++r) {
            dz += z_step;
            dx += x_step;
            umma_bf16(tmem_base, dx, dz, idesc, 1u);
          }
        } else {
          int r = 0;
          if (fresh) {  // very first K step of this CTA overwrites the accumulators
#pragma unroll
            for (int g = 0; g < GROUPS; ++g) umma_bf16(tmem_base + g * nt, dx + g * x_step, dz, idesc, 0u);
            dz += z_step;
            dx += x_step;
            r = 1;
          }
#pragma unroll 1
          for (; r < kWgTile; ++r) {  // one image row of 16 pixels = one K=16 step
#pragma unroll
            for (int g = 0; g < GROUPS; ++g) umma_bf16(tmem_base + g * nt, dx + g * x_step, dz, idesc, 1u);
            dz += z_step;
            dx += x_step;
          }
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
  } else {
    // ===================================== epilogue: TMEM -> fp32 partial slab ====================
    if (bias_cta) {
      // Thread -> (16-byte chunk position, row class): the TMA swizzle XORs the chunk index with row bits, so a thread
      // that only visits rows of one class reads the same logical channels every time (8 fp32 accumulators per box).
      const int et = static_cast<int>(threadIdx.x) - 64;  // 0..127
      const int row_bytes = GROUPS == 1 ? p.x_row_bytes : p.z_row_bytes;
      const bool wide = row_bytes == 128;                  // SWIZZLE_128B (64-channel boxes) vs SWIZZLE_64B (32)
      const int chunkpos = wide ? (et & 7) : (et & 3);
      const int rr = wide ? ((et >> 3) & 7) : ((et >> 2) & 7);   // row & 7
      const int k0 = wide ? (et >> 6) : (et >> 5);
      const int kstep = wide ? 2 : 4;
      constexpr int kBoxes = GROUPS == 1 ? 4 : 2;          // convT: the four tap boxes all add into one sum
      const int nb = GROUPS == 1 ? 4 : p.z_boxes;
      const int box_stride = GROUPS == 1 ? p.x_box_bytes : p.z_box_bytes;
      const int base_off = GROUPS == 1 ? p.z_region_bytes : p.z_pad_bytes;
      float bacc[2][8];
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int e = 0; e < 8; ++e) bacc[b][e] = 0.f;
      uint32_t slot = 0, ph = 0;
      for (int pt = split; pt < p.total_ptiles; pt += p.ksplit) {
        mbar_wait(&full[slot], ph, 14);
        // ConvTranspose form: the X map merges (batch, y) into one dimension, so a tile that hangs over the bottom edge
        // holds rows of the NEXT image there (harmless for the GEMM: the matching Z rows are zero-filled) - skip them.
        int k_end = 32;  // box row = ly * 16 + lx = rr + 8 k  ->  ly = k >> 1
        if (GROUPS == 1) {
          const int ty = (pt / p.tiles_x) % p.tiles_y;
          const int rows_left = p.PH - ty * kWgTile;
          if (rows_left < kWgTile) k_end = 2 * rows_left;
        }
        const uint8_t* zs = s_ptr0 + slot * p.stage_bytes + base_off + chunkpos * 16;
#pragma unroll
        for (int b = 0; b < kBoxes; ++b) {
          if (b < nb) {
            constexpr int kAccMask = GROUPS == 1 ? 0 : 1;
            float(&acc)[8] = bacc[b & kAccMask];
            const uint8_t* zb = zs + b * box_stride;
            for (int k = k0; k < k_end; k += kstep) {
              const uint4 u = *reinterpret_cast<const uint4*>(zb + (rr + 8 * k) * row_bytes);
              acc[0] += bf16lo(u.x);
              acc[1] += bf16hi(u.x);
              acc[2] += bf16lo(u.y);
              acc[3] += bf16hi(u.y);
              acc[4] += bf16lo(u.z);
              acc[5] += bf16hi(u.z);
              acc[6] += bf16lo(u.w);
              acc[7] += bf16hi(u.w);
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
        if (++slot == static_cast<uint32_t>(p.nstages)) {
          slot = 0;
          ph ^= 1u;
        }
      }
#pragma unroll
      for (int b = 0; b < 2; ++b)
#pragma unroll
        for (int e = 0; e < 8; ++e) bred[et][b * 8 + e] = bacc[b][e];
      asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps only
      const int n_ch = GROUPS == 1 ? p.x_ch : p.nt;        // channels this CTA sums
      if (et < n_ch) {
        const int box_ch = GROUPS == 1 ? p.x_ch : p.z_box_ch;
        const int b = et / box_ch, cc = et - b * box_ch;
        const int lc = cc >> 3, e = cc & 7;
        float acc = 0.f;
        if (wide) {
          for (int r8 = 0; r8 < 8; ++r8)
            for (int kk = 0; kk < 2; ++kk) acc += bred[(lc ^ r8) | (r8 << 3) | (kk << 6)][b * 8 + e];
        } else {
          for (int r8 = 0; r8 < 8; ++r8)
            for (int kk = 0; kk < 4; ++kk) acc += bred[(lc ^ (r8 >> 1)) | (r8 << 2) | (kk << 5)][b * 8 + e];
        }
        const int tile = GROUPS == 1 ? n_tile : m_tile, tiles = GROUPS == 1 ? p.n_tiles : p.m_tiles;
        p.bias_partial[(static_cast<size_t>(split) * tiles + tile) * n_ch + et] = acc;
      }
    }
    const int q = warp & 3;
    const int row = q * 32 + lane;
    if (my_tiles > 0) {
      mbar_wait(&acc_full, 0, 13);
      tc_fence_after_sync();
    }
    const int cols = p.groups * p.nt;
    float* out = p.partial + ((static_cast<size_t>(split) * p.n_tiles + n_tile) * p.m_tiles + m_tile) * 128 * cols +
                 static_cast<size_t>(row) * cols;
    for (int c = 0; c < cols; c += 32) {
      uint32_t v[32];
      if (my_tiles > 0) {
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(c), v);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0u;
      }
      float4* o4 = reinterpret_cast<float4*>(out + c);
#pragma unroll
      for (int i = 0; i < 8; ++i)
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

// Second stage: dW[m*sm + n*sn + t*st] (+)= scale * sum_s partial[s][n/x_ch][m/nt][row][col]
//   3x3 : t = kh*3 + kw,  row = kw*x_ch + n%x_ch,  col = kh*nt + m%nt
//   convT: t = q,         row = q*x_ch + n%x_ch,   col = m%nt
struct WgradReduceParams {
  const float* partial;
  float* dw;
  int M, N, n_valid, taps, groups, x_ch, nt, m_tiles, n_tiles, ksplit;
  long long sm, sn, st;
  float scale;
  int accumulate;
  const float* bias_partial;  // [ksplit][nbias] or NULL
  float* dbias;
  int nbias;  // 3x3: M (z channels); ConvTranspose: N (x channels)
  int flip_kh;  // 9-tap form: column group j holds kh = 2 - j
};
// Few splits (deep layers: many outputs, ksplit <= 8): one thread per output, splits walked serially.
__global__ void wgrad_reduce_flat_kernel(const WgradReduceParams p) {
  grid_dep_sync();
  const long long total = static_cast<long long>(p.n_valid) * p.taps * p.M;
  const long long total_b = total + (p.dbias != nullptr ? p.nbias : 0);
  const int cols = p.groups * p.nt;
  const size_t slab = static_cast<size_t>(128) * cols;
  const size_t split_stride = static_cast<size_t>(p.n_tiles) * p.m_tiles * slab;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total_b;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    if (i >= total) {  // fused bias gradient: column sums of Z per split
      const int m = static_cast<int>(i - total);
      float acc = 0.f;
      for (int s = 0; s < p.ksplit; ++s) acc += p.bias_partial[static_cast<size_t>(s) * p.nbias + m];
      p.dbias[m] = p.accumulate ? p.dbias[m] + acc * p.scale : acc * p.scale;
      continue;
    }
    long long r = i;
    const int m = static_cast<int>(r % p.M);
    r /= p.M;
    const int t = static_cast<int>(r % p.taps);
    const int n = static_cast<int>(r / p.taps);
    const int n_tile = n / p.x_ch, n_in = n - n_tile * p.x_ch;
    const int m_tile = m / p.nt, m_in = m - m_tile * p.nt;
    int row, col;
    if (p.groups == 3) {
      row = (t % 3) * p.x_ch + n_in;
      col = (p.flip_kh ? 2 - t / 3 : t / 3) * p.nt + m_in;
    } else {
      row = t * p.x_ch + n_in;
      col = m_in;
    }
    const float* src = p.partial + (static_cast<size_t>(n_tile) * p.m_tiles + m_tile) * slab +
                       static_cast<size_t>(row) * cols + col;
    float acc = 0.f;
    for (int s = 0; s < p.ksplit; ++s) acc += src[s * split_stride];
    float* dst = p.dw + (m * p.sm + n * p.sn + t * p.st);
    *dst = p.accumulate ? *dst + acc * p.scale : acc * p.scale;
  }
}

// Few splits, tiled: the flat form above reads the partials coalesced (m fastest) but then every thread stores 4 bytes
// PyTorch's weight layout apart ([m][n][t]: consecutive m are 9*C_in floats apart): 2.4 M scattered stores took 28 us
// for the 512x512 layer.  Here a block owns 32 z channels (m) x 32 x channels (n) x all taps, sums the splits while
// reading 128-byte runs along m, turns the tile in shared memory and writes runs of taps*32 consecutive floats per m.
// Needs x_ch == 32, M % 32 == 0, n_valid % 32 == 0 and (sn, st) == (taps, 1).
constexpr int kRedTileThreads = 256;
__global__ void __launch_bounds__(kRedTileThreads) wgrad_reduce_tiled_kernel(const WgradReduceParams p) {
  grid_dep_sync();
  __shared__ float sh[32 * 9][33];  // [(n_in, t)][m_in]
  const int cols = p.groups * p.nt;
  const size_t slab = static_cast<size_t>(128) * cols;
  const size_t split_stride = static_cast<size_t>(p.n_tiles) * p.m_tiles * slab;
  const int m_groups = p.M / 32, n_groups = p.n_valid / 32;
  const int tiles = m_groups * n_groups;
  const int run = 32 * p.taps;  // consecutive floats per m in the destination
  for (int tile = blockIdx.x; tile < tiles + (p.dbias != nullptr ? 1 : 0); tile += gridDim.x) {
    if (tile == tiles) {  // fused bias gradient: column sums of Z per split
      for (int m = threadIdx.x; m < p.nbias; m += kRedTileThreads) {
        float acc = 0.f;
        for (int s = 0; s < p.ksplit; ++s) acc += p.bias_partial[static_cast<size_t>(s) * p.nbias + m];
        p.dbias[m] = p.accumulate ? p.dbias[m] + acc * p.scale : acc * p.scale;
      }
      continue;
    }
    const int mg = tile % m_groups, n_tile = tile / m_groups;  // x_ch == 32: an n group is an x tile
    const int m0 = mg * 32;
    const int m_tile = m0 / p.nt, m_in0 = m0 - m_tile * p.nt;
    const float* base = p.partial + (static_cast<size_t>(n_tile) * p.m_tiles + m_tile) * slab;
    __syncthreads();
    for (int idx = threadIdx.x; idx < p.taps * 1024; idx += kRedTileThreads) {
      const int m_in = idx & 31, n_in = (idx >> 5) & 31, t = idx >> 10;
      int row, col;
      if (p.groups == 3) {
        row = (t % 3) * 32 + n_in;
        col = (p.flip_kh ? 2 - t / 3 : t / 3) * p.nt + m_in0 + m_in;
      } else {
        row = t * 32 + n_in;
        col = m_in0 + m_in;
      }
      const float* src = base + static_cast<size_t>(row) * cols + col;
      float acc = 0.f;
      for (int sp = 0; sp < p.ksplit; ++sp) acc += src[sp * split_stride];
      sh[n_in * p.taps + t][m_in] = acc * p.scale;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < 32 * run; idx += kRedTileThreads) {
      const int m_in = idx / run, r = idx - m_in * run;  // r = n_in * taps + t
      float* dst = p.dw + (static_cast<long long>(m0 + m_in) * p.sm + static_cast<long long>(n_tile) * 32 * p.sn + r);
      const float v = sh[r][m_in];
      *dst = p.accumulate ? *dst + v : v;
    }
  }
}

// Many splits (shallow layers: few outputs, hundreds of splits).
// One block reduces 32 consecutive z channels of one (x channel, tap): 8 threads per output walk the splits with
// stride 8 (eight loads in flight per output instead of one), then a fixed-order 8-way sum in shared memory.
constexpr int kRedLanes = 8;
__global__ void __launch_bounds__(32 * kRedLanes) wgrad_reduce_kernel(const WgradReduceParams p) {
  grid_dep_sync();
  __shared__ float sh[kRedLanes][33];
  const int cols = p.groups * p.nt;
  const size_t slab = static_cast<size_t>(128) * cols;
  const size_t split_stride = static_cast<size_t>(p.n_tiles) * p.m_tiles * slab;
  const int m_groups = p.M / 32;
  const long long total = static_cast<long long>(p.n_valid) * p.taps * m_groups;
  const long long total_b = total + (p.dbias != nullptr ? p.nbias / 32 : 0);
  const int o = threadIdx.x & 31, sl = threadIdx.x >> 5;
  for (long long i = blockIdx.x; i < total_b; i += gridDim.x) {
    if (i >= total) {  // fused bias gradient (block-uniform branch)
      const int m = static_cast<int>(i - total) * 32 + o;
      float acc = 0.f;
      for (int s = sl; s < p.ksplit; s += kRedLanes) acc += p.bias_partial[static_cast<size_t>(s) * p.nbias + m];
      __syncthreads();
      sh[sl][o] = acc;
      __syncthreads();
      if (sl == 0) {
        float v = 0.f;
#pragma unroll
        for (int k = 0; k < kRedLanes; ++k) v += sh[k][o];
        p.dbias[m] = p.accumulate ? p.dbias[m] + v * p.scale : v * p.scale;
      }
      continue;
    }
    long long r = i;
    const int mg = static_cast<int>(r % m_groups);
    r /= m_groups;
    const int t = static_cast<int>(r % p.taps);
    const int n = static_cast<int>(r / p.taps);
    const int m = mg * 32 + o;
    const int n_tile = n / p.x_ch, n_in = n - n_tile * p.x_ch;
    const int m_tile = m / p.nt, m_in = m - m_tile * p.nt;
    int row, col;
    if (p.groups == 3) {
      row = (t % 3) * p.x_ch + n_in;
      col = (p.flip_kh ? 2 - t / 3 : t / 3) * p.nt + m_in;
    } else {
      row = t * p.x_ch + n_in;
      col = m_in;
    }
    const float* src = p.partial + (static_cast<size_t>(n_tile) * p.m_tiles + m_tile) * slab +
                       static_cast<size_t>(row) * cols + col;
    float acc = 0.f;
    for (int s = sl; s < p.ksplit; s += kRedLanes) acc += src[s * split_stride];
    __syncthreads();
    sh[sl][o] = acc;
    __syncthreads();
    if (sl == 0) {
      float v = 0.f;
#pragma unroll
      for (int k = 0; k < kRedLanes; ++k) v += sh[k][o];
      float* dst = p.dw + (m * p.sm + n * p.sn + t * p.st);
      *dst = p.accumulate ? *dst + v * p.scale : v * p.scale;
    }
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

namespace {

struct WgGeometry {
  int x_ch, nt, groups, taps, n_tiles, m_tiles, total_ptiles, ksplit, nine;
  size_t partial_bytes, scratch_bytes;
};

int wg_geometry(const dpt_wgrad_desc* d, WgGeometry* g) {
  using namespace dpt;
  const int xc0 = d->x_c[0], xc1 = d->num_x == 2 ? d->x_c[1] : 0;
  if (d->num_x < 1 || d->num_x > 2) return fail("wgrad: num_x must be 1 or 2");
  if (xc0 <= 0 || xc0 % 16 || xc1 % 16 || d->z_c <= 0 || d->z_c % 32)
    return fail("wgrad: channels x=(%d,%d) z=%d unsupported", xc0, xc1, d->z_c);
  const bool strided = d->mode == WG_STRIDED;
  g->x_ch = (xc0 % 32 == 0 && xc1 % 32 == 0) ? 32 : 16;
  if (strided && g->x_ch != 32) return fail("wgrad: the strided (ConvTranspose) form needs x channels in multiples of 32");
  g->taps = strided ? 4 : 9;
  g->groups = strided ? 1 : 3;
  // 9-tap form (one MMA per row step covers all nine taps, N = 3 nt <= 192, so nt <= 64): every 3x3 layer whose output
  // channel count is 32 or a multiple of 64.  On the shallow layers it halves the shared-memory operand reads they
  // are bound by; on the deep ones the 64-channel z tile makes a 57 KB stage (three in flight instead of two 85 KB ones)
  // and measured 5-35 % faster as well (sum over the 18 layers 1.417 -> 1.343 ms).  DPT_WGRAD_NINE=0 selects the
  // one-MMA-per-kh form (N = 128) for comparison.
  static const int nine_env = [] {
    const char* e = getenv("DPT_WGRAD_NINE");
    return e ? atoi(e) : 1;
  }();
  g->nine = !strided && nine_env && (d->z_c == 32 || d->z_c % 64 == 0);
  const int cap = strided || (g->nine && d->z_c > 64) ? 64 : 128;  // z channels per CTA (GEMM N)
  g->nt = d->z_c < cap ? d->z_c : cap;
  if (d->z_c % g->nt) g->nt = 32;
  g->n_tiles = (xc0 + xc1) / g->x_ch;
  g->m_tiles = d->z_c / g->nt;
  g->total_ptiles = d->batch * ceil_div(d->w, kWgTile) * ceil_div(d->h, kWgTile);
  // Split-K sizing.  One CTA is resident per SM (shared memory + up to 512 TMEM columns), so the grid is one wave:
  // measured over the 18 layers of the benchmark UNet, 148 CTAs total 1.58 ms, 296 1.92 ms, 444 2.24 ms (fewer partial
  // slabs to write and re-read, no second-wave prologue).  DPT_WGRAD_TARGET_CTAS overrides the budget for tuning runs.
  static const int target_env = [] {
    const char* e = getenv("DPT_WGRAD_TARGET_CTAS");
    return e ? atoi(e) : 0;
  }();
  const int pairs = g->m_tiles * g->n_tiles;
  int ksplit = d->ksplit > 0 ? d->ksplit : (target_env > 0 ? target_env : sm_count()) / pairs;
  if (ksplit > g->total_ptiles) ksplit = g->total_ptiles;
  if (ksplit < 1) ksplit = 1;
  g->ksplit = ksplit;
  g->partial_bytes = static_cast<size_t>(g->m_tiles) * g->n_tiles * ksplit * 128 * g->groups * g->nt * sizeof(float);
  // always reserve room for the fused bias-gradient partials [ksplit][z_c] (ConvTranspose: [ksplit][x_c])
  g->scratch_bytes = g->partial_bytes + static_cast<size_t>(ksplit) * (strided ? xc0 : d->z_c) * sizeof(float);
  return 0;
}

}  // namespace

extern "C" size_t dpt_wgrad_scratch_bytes(const dpt_wgrad_desc* d) {
  WgGeometry g;
  if (!d || wg_geometry(d, &g)) return 0;
  return g.scratch_bytes;
}

extern "C" int dpt_wgrad_plan_create(const dpt_wgrad_desc* d, dpt_wgrad_plan** out) {
  using namespace dpt;
  if (!d || !out) return fail("dpt_wgrad_plan_create: null argument");
  if (d->mode != DPT_CONV_3X3 && d->mode != DPT_CONV_3X3_TAPBOX && d->mode != DPT_CONV_2X2S2)
    return fail("wgrad: bad mode %d", d->mode);
  const bool strided = d->mode == WG_STRIDED;  // DPT_CONV_3X3_TAPBOX is served by the halo form as well
  WgGeometry g;
  if (wg_geometry(d, &g)) return -1;
  const int xc0 = d->x_c[0], xc1 = d->num_x == 2 ? d->x_c[1] : 0;

  std::unique_ptr<dpt_wgrad_plan> owner(new (std::nothrow) dpt_wgrad_plan());
  dpt_wgrad_plan* plan = owner.get();
  if (!plan) return fail("wgrad: out of host memory");
  WgradParams& p = plan->p;
  memset(&p, 0, sizeof(p));
  p.mode = strided ? WG_STRIDED : WG_HALO;
  p.B = d->batch;
  p.PH = d->h;
  p.PW = d->w;
  p.tiles_x = ceil_div(d->w, kWgTile);
  p.tiles_y = ceil_div(d->h, kWgTile);
  p.total_ptiles = g.total_ptiles;
  p.groups = g.groups;
  p.nboxes = strided ? 4 : 1;
  p.pitch = strided ? kWgTile : kWgTile + 2;
  p.x_ch = g.x_ch;
  p.x_row_bytes = g.x_ch * 2;
  p.x_box_tx = (strided ? kWgTile * kWgTile : (kWgTile + 2) * (kWgTile + 2)) * p.x_row_bytes;
  p.x_box_bytes = (p.x_box_tx + 1023) & ~1023;
  // stacked atoms: 3x3 -> the next pixel of the same box; convT -> the next tap's box
  p.x_lbo = strided ? p.x_box_bytes : p.x_row_bytes;
  p.x_tiles0 = xc0 / g.x_ch;
  p.n_tiles = g.n_tiles;
  p.m_tiles = g.m_tiles;
  p.nt = g.nt;
  p.z_box_ch = g.nt >= 64 ? 64 : 32;
  p.z_row_bytes = p.z_box_ch * 2;
  p.z_boxes = g.nt / p.z_box_ch;
  p.z_box_bytes = kWgTile * kWgTile * p.z_row_bytes;  // 256 rows: a multiple of 1024
  p.z_pad_bytes = g.nine ? 2 * kWgTile * p.z_row_bytes : 0;  // two zero image rows on each side (2 or 4 KB)
  p.z_region_bytes = p.z_boxes * p.z_box_bytes + 2 * p.z_pad_bytes;
  p.stage_tx = p.z_boxes * p.z_box_bytes + p.nboxes * p.x_box_tx;
  // the halo form reads a few pixels past the last tap row of the box (the extra atoms of the last rows): pad the slot
  p.stage_bytes = p.z_region_bytes + p.nboxes * p.x_box_bytes + 1024;
  p.nstages = kWgSmemBudget / p.stage_bytes;
  if (p.nstages > kWgMaxStages) p.nstages = kWgMaxStages;
  if (p.nstages < 1) return fail("wgrad: stage of %d bytes does not fit", p.stage_bytes);
  int cols = 32;
  while (cols < g.groups * g.nt) cols <<= 1;
  p.tmem_cols = cols;
  p.ksplit = g.ksplit;
  plan->grid = p.m_tiles * p.n_tiles * g.ksplit;
  plan->scratch_bytes = g.scratch_bytes;
  if (d->scratch == nullptr || d->scratch_bytes < plan->scratch_bytes)
    return fail("wgrad: scratch of %zu bytes needed, %zu given", plan->scratch_bytes, d->scratch_bytes);
  p.partial = static_cast<float*>(d->scratch);
  if (d->dbias != nullptr) {
    p.bias_partial = reinterpret_cast<float*>(static_cast<uint8_t*>(d->scratch) + g.partial_bytes);
  }

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
    if (strided) {
      dims[0] = C; dims[1] = 2; dims[2] = d->w; dims[3] = 2; dims[4] = (uint64_t)d->batch * d->h;
      str[0] = C * 2; str[1] = 2 * C * 2; str[2] = 2ull * d->w * C * 2; str[3] = 2 * str[2];
      box[0] = g.x_ch; box[1] = 1; box[2] = kWgTile; box[3] = 1; box[4] = kWgTile;
    } else {
      dims[0] = C; dims[1] = d->w; dims[2] = d->h; dims[3] = d->batch; dims[4] = 1;
      str[0] = C * 2; str[1] = str[0] * d->w; str[2] = str[1] * d->h; str[3] = str[2] * d->batch;
      box[0] = g.x_ch; box[1] = p.pitch; box[2] = p.pitch; box[3] = 1; box[4] = 1;
    }
    if (make_tmap_bf16(&p.mapX[i], d->x[i], 5, dims, str, box, p.x_row_bytes)) return -1;
  }

  WgradReduceParams& r = plan->r;
  r.partial = p.partial;
  r.dw = d->dw;
  r.M = d->z_c;
  r.N = xc0 + xc1;
  r.n_valid = d->n_valid > 0 ? d->n_valid : xc0 + xc1;
  r.taps = g.taps;
  r.groups = g.groups;
  r.x_ch = g.x_ch;
  r.nt = g.nt;
  r.m_tiles = p.m_tiles;
  r.n_tiles = p.n_tiles;
  r.ksplit = g.ksplit;
  r.sm = d->dw_stride_m;
  r.sn = d->dw_stride_n;
  r.st = d->dw_stride_t;
  r.scale = d->scale == 0.f ? 1.f : d->scale;
  r.accumulate = d->accumulate;
  r.bias_partial = p.bias_partial;
  r.dbias = p.bias_partial != nullptr ? d->dbias : nullptr;
  r.nbias = strided ? xc0 : d->z_c;
  plan->smem = static_cast<size_t>(p.nstages) * p.stage_bytes + 1024;
  r.flip_kh = g.nine;
  plan->fn = strided ? igemm_wgrad_kernel<1> : (g.nine ? igemm_wgrad_kernel<9> : igemm_wgrad_kernel<3>);
  const cudaError_t e = cudaFuncSetAttribute(plan->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024);  // + 8.7 KB static <= 227 KB
  if (e != cudaSuccess) return fail("cudaFuncSetAttribute(wgrad smem) failed: %s", cudaGetErrorString(e));
  *out = owner.release();
  return 0;
}

static bool tiled_reduce_enabled() {
  static int v = [] {
    const char* e = getenv("DPT_WGRAD_TILED_REDUCE");
    return e ? atoi(e) : 0;
  }();
  return v != 0;
}

extern "C" int dpt_wgrad_plan_launch(const dpt_wgrad_plan* plan, void* stream) {
  using namespace dpt;
  if (!plan) return fail("dpt_wgrad_plan_launch: null plan");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DPT_CUDA_OK(launch_chained(plan->fn, dim3(plan->grid), dim3(kWgThreads), plan->smem, st, plan->p));
  DPT_CUDA_OK(cudaGetLastError());
  if (plan->r.ksplit >= 2 * kRedLanes) {
    const long long total = static_cast<long long>(plan->r.M / 32) * plan->r.n_valid * plan->r.taps + plan->r.nbias / 32;
    const int blocks = total > 16LL * sm_count() ? 16 * sm_count() : static_cast<int>(total);
    DPT_CUDA_OK(launch_chained(wgrad_reduce_kernel, dim3(blocks), dim3(32 * kRedLanes), 0, st, plan->r));
  } else if (plan->r.x_ch == 32 && plan->r.M % 32 == 0 && plan->r.n_valid % 32 == 0 && plan->r.st == 1 &&
             plan->r.sn == plan->r.taps && plan->r.taps <= 9 && tiled_reduce_enabled()) {
    const int tiles = (plan->r.M / 32) * (plan->r.n_valid / 32) + (plan->r.dbias != nullptr ? 1 : 0);
    DPT_CUDA_OK(launch_chained(wgrad_reduce_tiled_kernel, dim3(tiles < 8 * sm_count() ? tiles : 8 * sm_count()), dim3(kRedTileThreads), 0, st, plan->r));
  } else {
    const long long total = static_cast<long long>(plan->r.M) * plan->r.n_valid * plan->r.taps + plan->r.nbias;
    long long blocks = (total + 255) / 256;
    if (blocks > 8LL * sm_count()) blocks = 8LL * sm_count();
    DPT_CUDA_OK(launch_chained(wgrad_reduce_flat_kernel, dim3(static_cast<int>(blocks)), dim3(256), 0, st, plan->r));
  }
  g_launch_count.fetch_add(2, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" void dpt_wgrad_plan_destroy(dpt_wgrad_plan* plan) { delete plan; }
