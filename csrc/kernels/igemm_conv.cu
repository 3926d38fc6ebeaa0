// Implicit-GEMM convolution on tcgen05 / TMEM fed by TMA (sm_100a).
//
// One persistent, warp-specialised kernel covers every "pixels x channels" GEMM of the UNet:
//   * 3x3 pad-1 convolution forward and backward-data (model/unet_parts.py:10-12 of the reference, and the
//     ConvolutionBackward0 nodes autograd creates for utils/train_utils.py:69),
//   * ConvTranspose2d(k=2,s=2) forward (1x1 GEMM + pixel-shuffle store, model/unet_parts.py:51-54) and its
//     backward-data (2x2 stride-2 gather),
// with bias / ReLU / ReLU-backward mask / BatchNorm statistics fused in the epilogue, the channel concat of
// model/unet_parts.py:59 fused into the K loop (two source tensor maps), and the concat backward fused into
// the store (two destinations).
//
// Geometry.  A CTA tile is 16 rows x 8*S pixels of one image and NT output channels.  Each of the S
// sub-tiles (16 rows x 8 pixels = 128 GEMM rows) is one UMMA M=128 accumulator of NT fp32 columns in TMEM.
// For the 3x3 case the producer TMA-loads ONE halo box (18 x (8S+2) pixels x <=64 channels, zero-filled
// outside the image = the padding) per K chunk; the nine filter taps are nine UMMA shared-memory
// descriptors that point at different starting pixels of that same box: rows of a sub-tile are 8
// consecutive pixels (one swizzle atom's worth of rows), and consecutive image rows are a constant
// stride apart (SBO = box pitch), so a tap is just a different start address.  The 128B/64B/32B swizzle is
// a function of the absolute shared-memory address, identical for the TMA write and the UMMA read, so no
// re-layout is ever needed.  L2->smem traffic is (18*(8S+2))/(16*8S) = 1.20x (S=4) of the activation bytes
// instead of the 9x of a tap-by-tap im2col.
//
// Warps: 0 = A (activation) TMA producer, 1 = B (weight) TMA producer, 2 = TMEM owner + single-thread MMA
// issuer, 3..10 = epilogue, two warps per TMEM lane quadrant (TMEM -> registers -> bias/ReLU/mask -> bf16 -> 256-bit stores).
// Pipelines: A ring (a_full/a_empty), B ring (b_full/b_empty) or resident B (whole filter in smem when it
// fits), double-buffered accumulators (acc_full/acc_empty) so the epilogue of tile i overlaps the MMAs of
// tile i+1.
#include <atomic>
#include <memory>
#include <new>

#include "../../include/dpt_b200.h"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dpt {

extern std::atomic<long long> g_launch_count;

constexpr int kTileRows = 16;
constexpr int kIgemmThreads = 352;  // 11 warps: A producer, B producer, MMA issuer, 8 epilogue
constexpr int kMaxAStages = 4;
constexpr int kMaxBStages = 12;
constexpr int kSmemBudget = 216 * 1024;  // of the 227 KB a CTA may own; the rest is slack for alignment

enum { MODE_HALO = DPT_CONV_3X3, MODE_TAPBOX = DPT_CONV_3X3_TAPBOX, MODE_STRIDED = DPT_CONV_2X2S2,
       MODE_POINT = DPT_CONV_1X1 };

struct IgemmParams {
  CUtensorMap mapA[2];
  CUtensorMap mapW;
  int mode, out_mode;
  int B, PH, PW;
  int tiles_x, tiles_y, n_tiles_n, total_tiles;
  int S, NT;
  int chunk_ch, chunk_bytes;
  int chunks0, chunks_total;
  int taps, nboxes, taps_per_box;
  int box_pitch;       // pixels per row of an A box
  int a_box_tx_bytes;  // bytes one A box delivers
  int a_slot_bytes;    // slot stride (1024-aligned)
  int b_tile_bytes;
  int nA, nB, b_resident;
  int acc_bufs, tmem_cols;
  const float* bias;
  int relu;
  bf16* dst[2];
  int dst_c[2];
  const bf16* mask[2];
  int split_n;
  int cluster;       // CTAs per cluster (1, 2 or 4): they work on `cluster` pixel tiles of one N tile and share the filter loads
  int pixel_tiles;   // total_tiles / n_tiles_n
  bf16* pool;    // optional second output: nn.MaxPool2d(2, 2) of the (bias + ReLU) result, [B, PH/2, PW/2, cout]
  float* stats;  // BatchNorm statistics: per-CTA partial sums [grid][2][cout] (each CTA writes the NT columns it owns)
  int cout;
};

struct TileCoord {
  int n_idx, img, y0, x0;
};
__device__ __forceinline__ TileCoord decode_tile(const IgemmParams& p, int tile) {
  TileCoord t;
  t.n_idx = tile % p.n_tiles_n;
  int pt = tile / p.n_tiles_n;
  const int tx = pt % p.tiles_x;
  pt /= p.tiles_x;
  const int ty = pt % p.tiles_y;
  t.img = pt / p.tiles_y;
  t.y0 = ty * kTileRows;
  t.x0 = tx * 8 * p.S;
  return t;
}

// Which tile this CTA works on in its `it`-th round, and how many rounds it has.  cluster == 1: tile = blockIdx.x + it *
// grid.  cluster > 1: the clusters walk "groups" = (N tile, `cluster` consecutive pixel tiles); rank r of the cluster
// takes the r-th pixel tile of the group.  A group at the end may be short: the CTAs without a tile of their own redo the
// last pixel tile (they must keep consuming the shared filter stream) and store nothing (`store` = false).
__device__ __forceinline__ int cta_rounds(const IgemmParams& p) {
  const int cl = p.cluster;
  const int units = cl == 1 ? p.total_tiles : p.n_tiles_n * ((p.pixel_tiles + cl - 1) / cl);
  const int me = static_cast<int>(blockIdx.x) / cl, n = static_cast<int>(gridDim.x) / cl;
  return me < units ? (units - me + n - 1) / n : 0;
}
__device__ __forceinline__ int cta_tile(const IgemmParams& p, int it, bool& store) {
  const int cl = p.cluster;
  store = true;
  if (cl == 1) return static_cast<int>(blockIdx.x) + it * static_cast<int>(gridDim.x);
  const int group = static_cast<int>(blockIdx.x) / cl + it * (static_cast<int>(gridDim.x) / cl);
  const int n_idx = group % p.n_tiles_n;
  int pix = (group / p.n_tiles_n) * cl + static_cast<int>(blockIdx.x) % cl;
  if (pix >= p.pixel_tiles) {
    pix = p.pixel_tiles - 1;
    store = false;
  }
  return n_idx + p.n_tiles_n * pix;
}

// S sub-tiles x KSTEPS K=16 steps of one filter tap: S*KSTEPS tcgen05.mma with immediate descriptor offsets.
template <int S, int KSTEPS, bool kFirst>
__device__ __forceinline__ void issue_tap(uint32_t d_base, uint32_t nt, uint64_t a_desc, uint64_t b_desc,
                                          uint32_t idesc) {
  constexpr uint32_t kChunkBytes = KSTEPS * 32;
#pragma unroll
  for (int s = 0; s < S; ++s) {
#pragma unroll
    for (int k = 0; k < KSTEPS; ++k) {
      umma_bf16(d_base + static_cast<uint32_t>(s) * nt, a_desc + ((s * 8 * kChunkBytes + 32 * k) >> 4),
                b_desc + ((32 * k) >> 4), idesc, (kFirst && k == 0) ? 0u : 1u);
    }
  }
}

// VAR_HALO_STATIC: streamed filter with a ring of exactly nine slots, slot == tap.  The nine barriers and descriptors of
// the unrolled tap loop are then compile-time offsets and the ring needs no index arithmetic: ncu showed the ONE thread
// that issues the MMAs spending ~40 dependent instructions (R2UR moves of the slot index, an IMAD per barrier address, a
// constant-bank load of the ring size) per tap, ~490 cycles for 256 cycles of tensor work, on every wide layer - while
// the filter tiles it waited for had already arrived 99 % of the time (profiles/ncu_r2_step_ref.md).
enum { VAR_HALO_RESIDENT = 0, VAR_HALO_STREAM = 1, VAR_OTHER = 2, VAR_HALO_STATIC = 3 };

// Element offset of (pixel gy,gx; channel n_glob) in the destination it belongs to (d = 0/1 for the split outputs of
// a concat backward; SHUFFLE2 = ConvTranspose pixel shuffle).  The ReLU-backward mask shares the layout.
__device__ __forceinline__ size_t epi_offset(const IgemmParams& p, const TileCoord& t, int gy, int gx, int n_glob, int& d) {
  d = n_glob >= p.split_n ? 1 : 0;
  const int cbase = n_glob - (d ? p.split_n : 0);
  if (p.out_mode == DPT_OUT_SHUFFLE2) {
    const int cq = p.dst_c[0];
    const int quad = cbase / cq;
    const int co = cbase - quad * cq;
    const size_t opix = (static_cast<size_t>(t.img) * (2 * p.PH) + (2 * gy + (quad >> 1))) * (2 * p.PW) + (2 * gx + (quad & 1));
    return opix * cq + co;
  }
  const size_t pix = (static_cast<size_t>(t.img) * p.PH + gy) * p.PW + gx;
  return pix * p.dst_c[d] + cbase;
}

// One 32-column chunk of the epilogue: registers (already loaded from TMEM) -> bias / ReLU -> bf16x2 -> ReLU-backward
// gate as a packed AND -> two 256-bit stores.  The flags are compile-time so the chunk is ~250 straight-line
// instructions: the epilogue loop has to live in the ~6 KB L0 instruction cache (see the note at the loop).
template <bool BIAS, bool RELU, bool MASK>
__device__ __forceinline__ void epi_chunk_fast(const IgemmParams& p, const uint32_t (&v)[32], int n_glob, bool valid, int d,
                                               size_t off, const u32x8& m0, const u32x8& m1) {
  float f[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);
  if (BIAS) {
    const float4* b4 = reinterpret_cast<const float4*>(p.bias + n_glob);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 b = __ldg(b4 + i);
      f[4 * i + 0] += b.x;
      f[4 * i + 1] += b.y;
      f[4 * i + 2] += b.z;
      f[4 * i + 3] += b.w;
    }
  }
  if (RELU) {
#pragma unroll
    for (int i = 0; i < 32; ++i) f[i] = fmaxf(f[i], 0.f);
  }
  u32x8 o0, o1;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    o0.v[i] = pack_bf16x2(f[2 * i], f[2 * i + 1]);
    o1.v[i] = pack_bf16x2(f[16 + 2 * i], f[16 + 2 * i + 1]);
  }
  if (MASK) {
    if (p.mask[d] != nullptr) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        o0.v[i] &= bf16x2_gt0_mask(m0.v[i]);
        o1.v[i] &= bf16x2_gt0_mask(m1.v[i]);
      }
    }
  }
  if (valid) {
    stg256(p.dst[d] + off, o0);
    stg256(p.dst[d] + off + 16, o1);
  }
}

// bias + ReLU + store, and nn.MaxPool2d(2, 2) of the same values as a second, quarter-size store (the encoder's pooling,
// model/unet_parts.py:26,30-39, without a kernel that re-reads the full-resolution tensor).  A warp's 32 rows are 4 image
// rows x 8 pixels, so a 2x2 window lives in lanes {l, l^1, l^8, l^9}: two shuffles + two packed maxima per register on the
// packed bf16x2 words, stored by the window's top-left lane.
__device__ __forceinline__ void epi_chunk_pool(const IgemmParams& p, const uint32_t (&v)[32], int n_glob, bool valid, size_t off,
                                               bool pool_valid, size_t poff) {
  float f[32];
  const float4* b4 = reinterpret_cast<const float4*>(p.bias + n_glob);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float4 b = __ldg(b4 + i);
    f[4 * i + 0] = fmaxf(__uint_as_float(v[4 * i + 0]) + b.x, 0.f);
    f[4 * i + 1] = fmaxf(__uint_as_float(v[4 * i + 1]) + b.y, 0.f);
    f[4 * i + 2] = fmaxf(__uint_as_float(v[4 * i + 2]) + b.z, 0.f);
    f[4 * i + 3] = fmaxf(__uint_as_float(v[4 * i + 3]) + b.w, 0.f);
  }
  u32x8 o0, o1;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    o0.v[i] = pack_bf16x2(f[2 * i], f[2 * i + 1]);
    o1.v[i] = pack_bf16x2(f[16 + 2 * i], f[16 + 2 * i + 1]);
  }
  if (valid) {
    stg256(p.dst[0] + off, o0);
    stg256(p.dst[0] + off + 16, o1);
  }
  // rows outside the image hold garbage, but a window that is stored lies entirely inside (pool_valid)
  auto mx = [](uint32_t a, uint32_t b) {
    __nv_bfloat162 x, y;
    *reinterpret_cast<uint32_t*>(&x) = a;
    *reinterpret_cast<uint32_t*>(&y) = b;
    const __nv_bfloat162 m = __hmax2(x, y);
    return *reinterpret_cast<const uint32_t*>(&m);
  };
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    uint32_t a = o0.v[i], b = o1.v[i];
    a = mx(a, __shfl_xor_sync(0xffffffffu, a, 1));
    b = mx(b, __shfl_xor_sync(0xffffffffu, b, 1));
    a = mx(a, __shfl_xor_sync(0xffffffffu, a, 8));
    b = mx(b, __shfl_xor_sync(0xffffffffu, b, 8));
    o0.v[i] = a;
    o1.v[i] = b;
  }
  if (pool_valid) {
    stg256(p.pool + poff, o0);
    stg256(p.pool + poff + 16, o1);
  }
}

// BatchNorm batch statistics from the fp32 accumulators (north-star variant: Conv3x3(bias=False) -> BatchNorm -> ReLU).
// No atomics and no cross-lane traffic per chunk: every lane owns one GEMM row (pixel) and keeps a running sum and sum of
// squares for each of the 32 columns of "its" column chunk (a CTA owns ONE N tile for its whole life when the grid is a
// multiple of the N-tile count, and a warp sees one column chunk for NT <= 64), i.e. 64 FADD/FFMA per chunk.  The 32-lane
// transpose-reduction below runs once per kernel (NT <= 64) or once per chunk (NT = 128, two column chunks per warp, where
// the tensor pipe, not the epilogue, is the limit).  Per-CTA totals go to global memory in a fixed order: bit-reproducible.
__device__ __forceinline__ float warp_transpose_sum(float (&x)[32], int lane) {
  // after the step with offset o, lanes with bit o set hold the upper half of the surviving columns: lane L ends up with
  // the sum over all 32 lanes of column L
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int k = 0; k < off; ++k) {
      const float send = up ? x[k] : x[k + off];
      const float keep = up ? x[k + off] : x[k];
      x[k] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return x[0];
}

__device__ __forceinline__ void epi_store_plain(const IgemmParams& p, const float (&f)[32], bool valid, int d, size_t off) {
  if (valid) {
    u32x8 o0, o1;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      o0.v[i] = pack_bf16x2(f[2 * i], f[2 * i + 1]);
      o1.v[i] = pack_bf16x2(f[16 + 2 * i], f[16 + 2 * i + 1]);
    }
    stg256(p.dst[d] + off, o0);
    stg256(p.dst[d] + off + 16, o1);
  }
}

// Epilogue role (8 warps: two per TMEM lane quadrant, `half` 0/1 takes the even/odd 32-column chunks of every tile).
//
// Why it looks like this.  On the N <= 64 layers at full and half resolution (the expensive ones) the epilogue, not
// the tensor pipe, bounded the kernel: ncu showed the four epilogue warps busy 100 % of the time with half of their
// stall samples on `no_inst` (instruction fetch): the chunk body was unrolled four times with every flag tested at
// run time, ~23 KB of code walked once per tile through a ~6 KB L0 / 32 KB L1.5 instruction cache shared with the
// unrolled MMA-issue code.  Now the chunk body is specialised at compile time (~4 KB), NOT unrolled, and two warps
// per scheduler cover each other's tcgen05.ld / mask-load / fetch latency.  The ReLU-backward mask words are
// requested two chunks ahead of their use, across tile boundaries, and rotated through registers.
// STATS: 0 = none, 1 = per-lane running sums (NT <= 64), 2 = per-chunk transpose-reduction (NT = 128),
//        3 = no statistics: bias + ReLU with the fused 2x2 max-pool output (epi_chunk_pool)
template <int STATS, bool BIAS, bool RELU, bool MASK>
__device__ __forceinline__ void epilogue_role(const IgemmParams& p, uint32_t tmem_base, uint64_t* acc_full, uint64_t* acc_empty,
                                              int q, int half, int lane, float* stat_sm) {
  const int row = q * 32 + lane;   // GEMM row inside a sub-tile
  const int ry = row >> 3;         // image row inside the tile
  const int rx = row & 7;          // pixel inside the 8-pixel group
  const int jn = p.NT >> 5;        // 32-column chunks per sub-tile: 1, 2 or 4
  const int jsh = jn == 4 ? 2 : (jn == 2 ? 1 : 0);
  const int nchunks = p.S * jn;    // 1, 2, 4, 8 or 16
  const bool idle = nchunks == 1 && half == 1;
  const int cpw = nchunks > 1 ? nchunks >> 1 : 1;  // chunks per warp per tile
  const int cpw_sh = cpw == 8 ? 3 : (cpw == 4 ? 2 : (cpw == 2 ? 1 : 0));
  const int my_tiles = cta_rounds(p);
  const int n_total = my_tiles * cpw;

  struct Cursor {
    int itile;
    bool store;
    TileCoord t;
  };
  // where this warp's n-th chunk lands (valid / destination d / first channel / element offset)
  auto locate = [&](Cursor& cur, int n, bool& valid, int& d, int& n_glob, int& c) -> size_t {
    const int it = n >> cpw_sh;
    if (it != cur.itile) {
      cur.itile = it;
      cur.t = decode_tile(p, cta_tile(p, it, cur.store));
    }
    c = half + 2 * (n & (cpw - 1));
    const int s = c >> jsh, j = c & (jn - 1);
    const int gy = cur.t.y0 + ry;
    const int gx = cur.t.x0 + 8 * s + rx;
    valid = gy < p.PH && gx < p.PW && cur.store;
    n_glob = cur.t.n_idx * p.NT + 32 * j;
    return epi_offset(p, cur.t, gy, gx, n_glob, d);
  };
  Cursor cm{-1, true, {}}, cp{-1, true, {}};
  u32x8 mc0, mc1, mn0, mn1, mf0, mf1;  // mask words of chunk n (current), n+1 (next), n+2 (far)
  auto fetch_mask = [&](int n, u32x8& a, u32x8& b) {
    if (n < n_total) {
      bool valid;
      int d, n_glob, c;
      const size_t off = locate(cp, n, valid, d, n_glob, c);
      if (valid && p.mask[d] != nullptr) {
        a = ldg256(p.mask[d] + off);
        b = ldg256(p.mask[d] + off + 16);
      }
    }
  };
  if (MASK && !idle) {
    fetch_mask(0, mc0, mc1);
    fetch_mask(1, mn0, mn1);
  }
  constexpr int kAcc = STATS == 1 ? 32 : 1;
  float sa[kAcc], sq[kAcc];  // STATS 1: this lane's running column sums / sums of squares
#pragma unroll
  for (int i = 0; i < kAcc; ++i) sa[i] = sq[i] = 0.f;
  float t0a = 0.f, t1a = 0.f, t0b = 0.f, t1b = 0.f;  // STATS 2: column `lane` of column-chunk slot a / b
  uint32_t buf = 0;
  for (int n = 0; n < n_total; ++n) {
    const int k = n & (cpw - 1);
    if (k == 0) {
      const uint32_t itile = static_cast<uint32_t>(n >> cpw_sh);
      buf = itile % static_cast<uint32_t>(p.acc_bufs);
      const uint32_t accph = (itile / static_cast<uint32_t>(p.acc_bufs)) & 1u;
      mbar_wait(&acc_full[buf], accph, 7);
      tc_fence_after_sync();
    }
    if (!idle) {
      if (MASK) fetch_mask(n + 2, mf0, mf1);
      bool valid;
      int d, n_glob, c;
      const size_t off = locate(cm, n, valid, d, n_glob, c);
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * static_cast<uint32_t>(p.S * p.NT) +
                             static_cast<uint32_t>(32 * c);
      if (STATS == 3) {
        uint32_t v[32];
        tmem_ld_32x32(taddr, v);
        tmem_ld_wait();
        const int gy = cm.t.y0 + ry, gx = cm.t.x0 + 8 * (c >> jsh) + rx;
        const bool pv = (lane & 9) == 0 && gy + 1 < p.PH && gx + 1 < p.PW && cm.store;
        const size_t poff = ((static_cast<size_t>(cm.t.img) * (p.PH >> 1) + (gy >> 1)) * (p.PW >> 1) + (gx >> 1)) * p.cout + n_glob;
        epi_chunk_pool(p, v, n_glob, valid, off, pv, poff);
      } else if (STATS) {
        uint32_t v[32];
        tmem_ld_32x32(taddr, v);
        tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);
        if (STATS == 1) {
          if (valid) {
#pragma unroll
            for (int i = 0; i < kAcc; ++i) {
              sa[i] += f[i];
              sq[i] = fmaf(f[i], f[i], sq[i]);
            }
          }
        } else {
          float a[32], b[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            a[i] = valid ? f[i] : 0.f;
            b[i] = a[i] * a[i];
          }
          const float r0 = warp_transpose_sum(a, lane), r1 = warp_transpose_sum(b, lane);
          if (((c & (jn - 1)) >> 1) == 0) {
            t0a += r0;
            t1a += r1;
          } else {
            t0b += r0;
            t1b += r1;
          }
        }
        epi_store_plain(p, f, valid, d, off);
      } else {
        uint32_t v[32];
        tmem_ld_32x32(taddr, v);
        tmem_ld_wait();
        epi_chunk_fast<BIAS, RELU, MASK>(p, v, n_glob, valid, d, off, mc0, mc1);
      }
      if (MASK) {
        mc0 = mn0;
        mc1 = mn1;
        mn0 = mf0;
        mn1 = mf1;
      }
    }
    if (k == cpw - 1) {
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
    }
  }
  if (STATS == 1 || STATS == 2) {
    // CTA totals: [half][q][slot][sum|sumsq][32 columns] in shared memory, then a fixed-order sum over the warps that
    // covered each column chunk (jn = 1: both halves; 2: half == chunk; 4: half == chunk & 1, slot == chunk >> 1)
    if constexpr (STATS == 1) {
      t0a = warp_transpose_sum(sa, lane);
      t1a = warp_transpose_sum(sq, lane);
    }
    float* mine = stat_sm + (half * 4 + q) * 128;
    mine[lane] = t0a;
    mine[32 + lane] = t1a;
    mine[64 + lane] = t0b;
    mine[96 + lane] = t1b;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    const int t = (half * 4 + q) * 32 + lane;
    if (t < p.NT) {
      const int j = t >> 5, i = t & 31;
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int qq = 0; qq < 4; ++qq) {
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
#pragma unroll
          for (int ss = 0; ss < 2; ++ss) {
            const bool contrib = jn == 1 ? ss == 0 : (jn == 2 ? (ss == 0 && hh == j) : (hh == (j & 1) && ss == (j >> 1)));
            if (contrib) {
              s0 += stat_sm[(hh * 4 + qq) * 128 + ss * 64 + i];
              s1 += stat_sm[(hh * 4 + qq) * 128 + ss * 64 + 32 + i];
            }
          }
        }
      }
      // row r of the statistics array belongs to N tile r % n_tiles_n (what dpt_bn_finalize assumes); with clusters the
      // CTAs of cluster c = q * n_tiles_n + n_idx take rows (q * cluster + rank) * n_tiles_n + n_idx
      const int cid = static_cast<int>(blockIdx.x) / p.cluster, rank = static_cast<int>(blockIdx.x) % p.cluster;
      const int n_idx = cid % p.n_tiles_n;
      const int r = ((cid / p.n_tiles_n) * p.cluster + rank) * p.n_tiles_n + n_idx;
      float* row = p.stats + static_cast<size_t>(r) * 2 * p.cout + n_idx * p.NT + t;
      row[0] = s0;
      row[p.cout] = s1;
    }
  }
}

// CL > 1 (cluster launch, streamed 3x3 filters only): the CL CTAs of a cluster run CL pixel tiles of the SAME N tile in
// lock step on the filter stream.  Each CTA fetches 1/CL of every filter tile and TMA-multicasts it into the same slot of
// all CL CTAs, so the L2 -> SM traffic of the operand every pixel tile re-reads is divided by CL (the deep layers moved
// 1.2 MB of filter per 128x128 tile against 9.4 us of MMAs: 85 % of the chip's TMA rate with all SMs busy).  A slot is
// free when the MMAs of ALL CL CTAs have read it: b_empty counts CL multicast commits.
template <int S, int KSTEPS, int VARIANT, int CL>
__global__ void __launch_bounds__(kIgemmThreads, 1) igemm_conv_kernel(const __grid_constant__ IgemmParams p) {
  constexpr uint32_t kChunkBytes = KSTEPS * 32;
  constexpr uint16_t kClusterMask = static_cast<uint16_t>((1u << CL) - 1u);
  grid_dep_launch();  // the next kernel's prologue may overlap this grid's tail (ptx.cuh)
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t a_full[kMaxAStages], a_empty[kMaxAStages];
  __shared__ uint64_t b_full[kMaxBStages], b_empty[kMaxBStages];
  __shared__ uint64_t acc_full[2], acc_empty[2];
  __shared__ uint64_t b_all;
  __shared__ uint32_t tmem_base_slot;
  __shared__ float stat_sm[8 * 128];  // BatchNorm statistics hand-over between the epilogue warps

  const int warp = warp_idx_sync();
  const int lane = threadIdx.x & 31;

  // 1024-byte aligned carve-up: [A slots][B tiles]
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t a_addr0 = (raw_addr + 1023u) & ~1023u;
  uint8_t* const a_ptr0 = smem_raw + (a_addr0 - raw_addr);
  const uint32_t b_addr0 = a_addr0 + static_cast<uint32_t>(p.nA * p.a_slot_bytes);
  uint8_t* const b_ptr0 = a_ptr0 + p.nA * p.a_slot_bytes;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kMaxAStages; ++i) {
      mbar_init(&a_full[i], 1);
      mbar_init(&a_empty[i], 1);
    }
    for (int i = 0; i < kMaxBStages; ++i) {
      mbar_init(&b_full[i], 1);
      mbar_init(&b_empty[i], CL);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], 8);  // one arrival per epilogue warp
    }
    mbar_init(&b_all, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(&tmem_base_slot, static_cast<uint32_t>(p.tmem_cols));
    tmem_relinquish();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.mapA[0]);
    if (p.chunks_total > p.chunks0) tma_prefetch_desc(&p.mapA[1]);
  }
  if (warp == 1 && lane == 0) tma_prefetch_desc(&p.mapW);
  tc_fence_before_sync();
  __syncthreads();
  if constexpr (CL > 1) cluster_sync_all();  // the peers' barriers are initialised before anything arrives on them
  tc_fence_after_sync();
  grid_dep_wait();  // barriers and TMEM are set up: from here on the previous kernel's output is read
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_slot, 0);
  const int rounds = cta_rounds(p);

  if (warp == 0) {
    // ================================ A producer: activation boxes ================================
    {
      uint32_t it = 0;
      for (int round = 0; round < rounds; ++round) {
        bool store;
        const TileCoord t = decode_tile(p, cta_tile(p, round, store));
        for (int ck = 0; ck < p.chunks_total; ++ck) {
          const int src = ck >= p.chunks0 ? 1 : 0;
          const int c0 = (ck - (src ? p.chunks0 : 0)) * p.chunk_ch;
          for (int bx = 0; bx < p.nboxes; ++bx, ++it) {
            const uint32_t slot = it % static_cast<uint32_t>(p.nA);
            const uint32_t ph = (it / static_cast<uint32_t>(p.nA)) & 1u;
            mbar_wait(&a_empty[slot], ph ^ 1u, 1);
            // box coordinates in the 5-D tensor map (see dpt_conv_plan_create for the two views)
            int k1, k2, k3, k4;
            if (p.mode == MODE_HALO) {
              k1 = t.x0 - 1, k2 = t.y0 - 1, k3 = t.img, k4 = 0;
            } else if (p.mode == MODE_TAPBOX) {
              k1 = t.x0 + (bx % 3) - 1, k2 = t.y0 + (bx / 3) - 1, k3 = t.img, k4 = 0;
            } else if (p.mode == MODE_STRIDED) {  // dims (C, dx:2, x:PW, dy:2, b*PH + y)
              k1 = bx & 1, k2 = t.x0, k3 = bx >> 1, k4 = t.img * p.PH + t.y0;
            } else {
              k1 = t.x0, k2 = t.y0, k3 = t.img, k4 = 0;
            }
            if (elect_one_sync()) {
              mbar_arrive_expect_tx(&a_full[slot], static_cast<uint32_t>(p.a_box_tx_bytes));
              tma_load_5d(a_ptr0 + slot * p.a_slot_bytes, &p.mapA[src], &a_full[slot], c0, k1, k2, k3, k4);
            }
            __syncwarp();
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ B producer: filter tiles ====================================
    {
      if (p.b_resident) {
        if (rounds > 0 && elect_one_sync()) {
          const uint32_t total = static_cast<uint32_t>(p.chunks_total * p.taps * p.b_tile_bytes);
          mbar_arrive_expect_tx(&b_all, total);
          for (int ck = 0; ck < p.chunks_total; ++ck)
            for (int tap = 0; tap < p.taps; ++tap)
              tma_load_3d(b_ptr0 + (ck * p.taps + tap) * p.b_tile_bytes, &p.mapW, &b_all, ck * p.chunk_ch, 0, tap);
        }
      } else {
        uint32_t it = 0;
        const int crank = CL > 1 ? static_cast<int>(cluster_ctarank()) : 0;
        const int part_rows = p.NT / CL;  // filter rows this CTA fetches (and multicasts) of every tile
        if constexpr (VARIANT == VAR_HALO_STATIC) {
          uint32_t pass = 0;  // one pass over the nine slots per K chunk
          for (int round = 0; round < rounds; ++round) {
            bool store;
            const int n0 = (cta_tile(p, round, store) % p.n_tiles_n) * p.NT;
            for (int ck = 0; ck < p.chunks_total; ++ck, ++pass) {
              const uint32_t ph = pass & 1u;
#pragma unroll
              for (int tap = 0; tap < 9; ++tap) {
                mbar_wait(&b_empty[tap], ph ^ 1u, 2);
                if (elect_one_sync()) {
                  mbar_arrive_expect_tx(&b_full[tap], static_cast<uint32_t>(p.b_tile_bytes));
                  tma_load_3d(b_ptr0 + tap * p.b_tile_bytes, &p.mapW, &b_full[tap], ck * p.chunk_ch, n0, tap);
                }
              }
            }
          }
        } else
        for (int round = 0; round < rounds; ++round) {
          bool store;
          const int n0 = (cta_tile(p, round, store) % p.n_tiles_n) * p.NT;
          for (int ck = 0; ck < p.chunks_total; ++ck) {
            for (int tap = 0; tap < p.taps; ++tap, ++it) {
              const uint32_t slot = it % static_cast<uint32_t>(p.nB);
              const uint32_t ph = (it / static_cast<uint32_t>(p.nB)) & 1u;
              mbar_wait(&b_empty[slot], ph ^ 1u, 2);
              if (elect_one_sync()) {
                mbar_arrive_expect_tx(&b_full[slot], static_cast<uint32_t>(p.b_tile_bytes));
                if constexpr (CL > 1) {
                  tma_load_3d_mcast(b_ptr0 + slot * p.b_tile_bytes + crank * part_rows * p.chunk_bytes, &p.mapW, &b_full[slot],
                                    ck * p.chunk_ch, n0 + crank * part_rows, tap, kClusterMask);
                } else {
                  tma_load_3d(b_ptr0 + slot * p.b_tile_bytes, &p.mapW, &b_full[slot], ck * p.chunk_ch, n0, tap);
                }
              }
            }
          }
        }
      }
    }
  } else if (warp == 2) {
    // ================================ MMA issuer ===================================================
    // ONE elected lane runs the whole role (the other 31 lanes go straight to the final barrier), so ptxas keeps
    // everything on the uniform datapath.  That datapath retires one dependent instruction every ~4-7 cycles
    // and there is no other warp to hide it, so the loop is written for instruction count: S, KSTEPS and the
    // tap structure (VARIANT) are template parameters, every descriptor is "previous + immediate", kernel
    // parameters are read into registers once (ncu: the generic loop spent ~700 cycles per tap for 8 MMAs).
    if (elect_one_sync()) {
      const uint32_t idesc = umma_idesc_bf16(p.NT, 0, 0);
      const uint32_t layout = umma_layout_type(kChunkBytes);
      const uint32_t pitch = static_cast<uint32_t>(p.box_pitch);
      const uint64_t a_desc0 = umma_smem_desc(0u, 16u, pitch * kChunkBytes, layout);
      const uint64_t b_desc0 = umma_smem_desc(0u, 16u, 8u * kChunkBytes, layout);
      const uint32_t nt = static_cast<uint32_t>(p.NT);
      const uint32_t nA = static_cast<uint32_t>(p.nA), nB = static_cast<uint32_t>(p.nB);
      const uint32_t a_slot_units = static_cast<uint32_t>(p.a_slot_bytes) >> 4;
      const uint32_t b_tile_units = static_cast<uint32_t>(p.b_tile_bytes) >> 4;
      const uint32_t acc_bufs = static_cast<uint32_t>(p.acc_bufs);
      const int chunks_total = p.chunks_total, nboxes = p.nboxes;
      constexpr bool kHalo = VARIANT != VAR_OTHER;  // RESIDENT, STREAM and STATIC are the 3x3 halo forms
      constexpr bool kResident = VARIANT == VAR_HALO_RESIDENT;
      constexpr bool kStatic = VARIANT == VAR_HALO_STATIC;
      const uint64_t b_static0 = b_desc0 + (b_addr0 >> 4);  // slot 0 of the static ring; slot t is t * b_tile_units further
      const uint32_t b_full0 = smem_u32(b_full), b_empty0 = smem_u32(b_empty);  // static ring: barrier t is 8 t further
      const bool resident_rt = !kHalo && p.b_resident;  // non-3x3 modes keep the runtime choice
      // tap offsets inside a halo box, in 16-byte units (kh rows down, kw pixels right): loop-invariant
      const uint32_t row_units = (pitch * kChunkBytes) >> 4;
      constexpr uint32_t kPixUnits = kChunkBytes >> 4;
      uint32_t sa = 0, pha = 0, sb = 0, phb = 0, itile = 0;
      bool b_ok = false;  // static ring: result of the early poll of the next tap's filter barrier
      if ((kResident || resident_rt) && rounds > 0) {
        mbar_wait(&b_all, 0, 3);
        tc_fence_after_sync();
      }
      for (int round = 0; round < rounds; ++round, ++itile) {
        const uint32_t buf = acc_bufs == 2 ? (itile & 1u) : 0u;
        const uint32_t accph = acc_bufs == 2 ? ((itile >> 1) & 1u) : (itile & 1u);
        mbar_wait(&acc_empty[buf], accph ^ 1u, 4);
        tc_fence_after_sync();
        const uint32_t d_base = tmem_base + buf * static_cast<uint32_t>(S) * nt;
        uint64_t b_res = b_desc0 + (b_addr0 >> 4);  // resident filter: walks (chunk, tap) in order
        bool first = true;
        for (int ck = 0; ck < chunks_total; ++ck) {
          for (int bx = 0; bx < nboxes; ++bx) {
            mbar_wait(&a_full[sa], pha, 5);
            tc_fence_after_sync();
            const uint64_t a_slot = a_desc0 + ((a_addr0 >> 4) + sa * a_slot_units);
            if constexpr (kHalo) {
#pragma unroll
              for (int tp = 0; tp < 9; ++tp) {
                const uint64_t a_desc = a_slot + (static_cast<uint32_t>(tp / 3) * row_units + (tp % 3) * kPixUnits);
                uint64_t b_desc;
                if constexpr (kResident) {
                  b_desc = b_res;
                  b_res += b_tile_units;
                } else if constexpr (kStatic) {
                  // the barrier of THIS tap was polled before the previous tap's MMAs were issued (b_ok): the ~30-cycle
                  // round trip of the poll hides behind that issue instead of standing in front of this one
                  if (!b_ok) mbar_wait_addr(b_full0 + 8u * tp, phb, 6);
                  tc_fence_after_sync();
                  b_ok = tp < 8 ? mbar_try_wait_addr(b_full0 + 8u * (tp + 1), phb) : mbar_try_wait_addr(b_full0, phb ^ 1u);
                  b_desc = b_static0 + static_cast<uint32_t>(tp) * b_tile_units;
                } else {
                  mbar_wait(&b_full[sb], phb, 6);
                  tc_fence_after_sync();
                  b_desc = b_desc0 + ((b_addr0 >> 4) + sb * b_tile_units);
                }
                if (tp == 0 && first) {
                  issue_tap<S, KSTEPS, true>(d_base, nt, a_desc, b_desc, idesc);
                } else {
                  issue_tap<S, KSTEPS, false>(d_base, nt, a_desc, b_desc, idesc);
                }
                if constexpr (kStatic) {
                  umma_commit_addr(b_empty0 + 8u * tp);
                } else if constexpr (!kResident) {
                  if constexpr (CL > 1) {
                    umma_commit_mcast(&b_empty[sb], kClusterMask);
                  } else {
                    umma_commit(&b_empty[sb]);
                  }
                  if (++sb == nB) {
                    sb = 0;
                    phb ^= 1u;
                  }
                }
              }
              if constexpr (kStatic) phb ^= 1u;
            } else {  // one tap per box (1x1, strided 2x2, tap-box 3x3)
              uint64_t b_desc;
              if (resident_rt) {
                b_desc = b_res;
                b_res += b_tile_units;
              } else {
                mbar_wait(&b_full[sb], phb, 6);
                tc_fence_after_sync();
                b_desc = b_desc0 + ((b_addr0 >> 4) + sb * b_tile_units);
              }
              if (first) {
                issue_tap<S, KSTEPS, true>(d_base, nt, a_slot, b_desc, idesc);
              } else {
                issue_tap<S, KSTEPS, false>(d_base, nt, a_slot, b_desc, idesc);
              }
              if (!resident_rt) {
                umma_commit(&b_empty[sb]);
                if (++sb == nB) {
                  sb = 0;
                  phb ^= 1u;
                }
              }
            }
            first = false;
            umma_commit(&a_empty[sa]);
            if (++sa == nA) {
              sa = 0;
              pha ^= 1u;
            }
          }
        }
        umma_commit(&acc_full[buf]);
      }
    }
  } else {
    // ================================ epilogue warps 3..10 =======================================
    const int q = warp & 3;           // TMEM lane quadrant this warp may read
    const int half = (warp - 3) >> 2;  // which of the quadrant's two warps
    const bool has_mask = p.mask[0] != nullptr || p.mask[1] != nullptr;
    if (p.pool != nullptr) {
      epilogue_role<3, true, true, false>(p, tmem_base, acc_full, acc_empty, q, half, lane, stat_sm);
    } else if (p.stats != nullptr) {
      if (p.NT <= 64) {
        epilogue_role<1, false, false, false>(p, tmem_base, acc_full, acc_empty, q, half, lane, stat_sm);
      } else {
        epilogue_role<2, false, false, false>(p, tmem_base, acc_full, acc_empty, q, half, lane, stat_sm);
      }
    } else {
      switch ((p.bias != nullptr ? 1 : 0) | (p.relu ? 2 : 0) | (has_mask ? 4 : 0)) {
        case 0: epilogue_role<0, false, false, false>(p, tmem_base, acc_full, acc_empty, q, half, lane, stat_sm); break;
        case 1: epilogue_role<0, true, false, false>(p, tmem_base, acc_full, acc_empty, q, half, lane, stat_sm); break;
        case 2: epilogue_role<0, false, true, false>(p, tmem_base, acc_full, acc_empty, q, half, lane, stat_sm); break;
        case 3: epilogue_role<0, true, true, false>(p, tmem_base, acc_full, acc_empty, q, half, lane, stat_sm); break;
        case 4: epilogue_role<0, false, false, true>(p, tmem_base, acc_full, acc_empty, q, half, lane, stat_sm); break;
        case 5: epilogue_role<0, true, false, true>(p, tmem_base, acc_full, acc_empty, q, half, lane, stat_sm); break;
        case 6: epilogue_role<0, false, true, true>(p, tmem_base, acc_full, acc_empty, q, half, lane, stat_sm); break;
        default: epilogue_role<0, true, true, true>(p, tmem_base, acc_full, acc_empty, q, half, lane, stat_sm); break;
      }
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if constexpr (CL > 1) cluster_sync_all();  // no CTA leaves while a peer may still multicast into it or arrive on its barriers
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc(tmem_base, static_cast<uint32_t>(p.tmem_cols));
  }
}

}  // namespace dpt

// ======================================= host side ============================================
using IgemmKernelFn = void (*)(const dpt::IgemmParams);

struct dpt_conv_plan {
  dpt::IgemmParams p;
  IgemmKernelFn fn;
  int grid;
  size_t smem;
  int cluster;
};

namespace {

template <int VARIANT>
IgemmKernelFn pick_kernel_v(int S, int ksteps) {
  using namespace dpt;
  switch (S * 10 + ksteps) {
    case 11: return igemm_conv_kernel<1, 1, VARIANT, 1>;
    case 12: return igemm_conv_kernel<1, 2, VARIANT, 1>;
    case 14: return igemm_conv_kernel<1, 4, VARIANT, 1>;
    case 21: return igemm_conv_kernel<2, 1, VARIANT, 1>;
    case 22: return igemm_conv_kernel<2, 2, VARIANT, 1>;
    case 24: return igemm_conv_kernel<2, 4, VARIANT, 1>;
    case 41: return igemm_conv_kernel<4, 1, VARIANT, 1>;
    case 42: return igemm_conv_kernel<4, 2, VARIANT, 1>;
    case 44: return igemm_conv_kernel<4, 4, VARIANT, 1>;
    default: return nullptr;
  }
}
// cluster variants exist for the shapes of the wide layers: 64-channel K chunks, one or two sub-tiles, streamed filters
IgemmKernelFn pick_cluster_kernel(int S, int ksteps, int cluster) {
  using namespace dpt;
  if (ksteps != 4) return nullptr;
  if (cluster == 2) return S == 1 ? igemm_conv_kernel<1, 4, VAR_HALO_STREAM, 2> : S == 2 ? igemm_conv_kernel<2, 4, VAR_HALO_STREAM, 2> : nullptr;
  if (cluster == 4) return S == 1 ? igemm_conv_kernel<1, 4, VAR_HALO_STREAM, 4> : nullptr;
  return nullptr;
}
// DPT_CONV_STATIC_RING=0 falls back to the indexed ring (A/B measurements)
bool static_ring() {
  static int v = [] {
    const char* e = getenv("DPT_CONV_STATIC_RING");
    return e ? atoi(e) : 1;
  }();
  return v != 0;
}
IgemmKernelFn pick_kernel(int S, int ksteps, int mode, int resident, int nB) {
  if (mode != dpt::MODE_HALO) return pick_kernel_v<dpt::VAR_OTHER>(S, ksteps);
  if (resident) return pick_kernel_v<dpt::VAR_HALO_RESIDENT>(S, ksteps);
  return nB == 9 && static_ring() ? pick_kernel_v<dpt::VAR_HALO_STATIC>(S, ksteps) : pick_kernel_v<dpt::VAR_HALO_STREAM>(S, ksteps);
}

// True if `ptr` is ordinary device memory of the CURRENT device.  TMA multicast loads here read only this GPU's memory,
// since on B200 a TMA multicast load from another GPU's memory hung the GPU: anything else (a peer's buffer, an imported
// IPC / VMM mapping, managed or host memory) keeps the plain per-CTA loads.
bool is_local_device_memory(const void* ptr) {
  cudaPointerAttributes attr{};
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  if (cudaPointerGetAttributes(&attr, ptr) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return attr.type == cudaMemoryTypeDevice && attr.device == dev;
}

// ring depth caps (sweeps: DPT_CONV_MAX_NB / DPT_CONV_MAX_NA)
int max_nb() {
  static int v = [] {
    const char* e = getenv("DPT_CONV_MAX_NB");
    const int n = e ? atoi(e) : 8;
    return n < 2 ? 2 : (n > dpt::kMaxBStages ? dpt::kMaxBStages : n);
  }();
  return v;
}
int max_na() {
  static int v = [] {
    const char* e = getenv("DPT_CONV_MAX_NA");
    const int n = e ? atoi(e) : dpt::kMaxAStages;
    return n < 2 ? 2 : (n > dpt::kMaxAStages ? dpt::kMaxAStages : n);
  }();
  return v;
}

int cluster_request() {
  static int v = [] {
    // measured (profiles/layers_cluster_r2.md): with the filter ring 8 slots deep the wide layers are bound by the latency
    // of the ring, not by L2 bandwidth, and sharing the tiles gains nothing (-4 % with clusters of 2): off by default
    const char* e = getenv("DPT_CONV_CLUSTER");
    return e ? atoi(e) : 1;
  }();
  return v;
}

int pow2_cols(int need) {
  int c = 32;
  while (c < need) c <<= 1;
  return c;
}

}  // namespace

extern "C" int dpt_conv_plan_create(const dpt_conv_desc* d, dpt_conv_plan** out) {
  using namespace dpt;
  if (!d || !out) return fail("dpt_conv_plan_create: null argument");
  if (d->mode < 0 || d->mode > 3) return fail("conv: bad mode %d", d->mode);
  if (d->num_src < 1 || d->num_src > 2 || d->num_dst < 1 || d->num_dst > 2)
    return fail("conv: num_src/num_dst must be 1 or 2");
  if (d->cout % 32 != 0 || d->cout <= 0) return fail("conv: cout=%d must be a positive multiple of 32", d->cout);
  const int c0 = d->src_c[0], c1 = d->num_src == 2 ? d->src_c[1] : 0;
  if (c0 <= 0 || c0 % 16 || c1 % 16) return fail("conv: source channels (%d,%d) must be multiples of 16", c0, c1);
  auto divides = [&](int ch) { return c0 % ch == 0 && c1 % ch == 0; };
  const int chunk_ch = divides(64) ? 64 : divides(32) ? 32 : 16;

  std::unique_ptr<dpt_conv_plan> owner(new (std::nothrow) dpt_conv_plan());
  dpt_conv_plan* plan = owner.get();
  if (!plan) return fail("conv: out of host memory");
  IgemmParams& p = plan->p;
  memset(&p, 0, sizeof(p));
  p.mode = d->mode;
  p.out_mode = d->out_mode;
  p.B = d->batch;
  p.PH = d->h;
  p.PW = d->w;
  p.chunk_ch = chunk_ch;
  p.chunk_bytes = chunk_ch * 2;
  p.chunks0 = c0 / chunk_ch;
  p.chunks_total = (c0 + c1) / chunk_ch;
  p.taps = d->mode == MODE_HALO || d->mode == MODE_TAPBOX ? 9 : d->mode == MODE_STRIDED ? 4 : 1;
  p.nboxes = d->mode == MODE_TAPBOX ? 9 : d->mode == MODE_STRIDED ? 4 : 1;
  p.taps_per_box = p.taps / p.nboxes;
  p.cout = d->cout;
  p.bias = d->bias;
  p.relu = d->relu;
  p.stats = d->stats;
  p.pool = static_cast<bf16*>(d->pool_dst);
  if (p.pool != nullptr && (d->bias == nullptr || !d->relu || d->stats != nullptr || d->num_dst != 1 || d->mask[0] != nullptr ||
                            d->out_mode != DPT_OUT_NHWC || (d->h & 1) || (d->w & 1)))
    return fail("conv: the fused max-pool output needs bias + ReLU, one NHWC destination, no mask / statistics, even h and w");
  for (int i = 0; i < 2; ++i) {
    p.dst[i] = i < d->num_dst ? static_cast<bf16*>(d->dst[i]) : nullptr;
    p.dst_c[i] = i < d->num_dst ? d->dst_c[i] : 0;
    p.mask[i] = i < d->num_dst ? static_cast<const bf16*>(d->mask[i]) : nullptr;
  }
  if (d->out_mode == DPT_OUT_SHUFFLE2) {
    if (d->num_dst != 1 || d->dst_c[0] * 4 != d->cout || d->dst_c[0] % 32)
      return fail("conv: pixel-shuffle store needs one dst with cout/4 channels (multiple of 32)");
    p.split_n = d->cout;
  } else {
    const int sum = d->dst_c[0] + (d->num_dst == 2 ? d->dst_c[1] : 0);
    if (sum != d->cout) return fail("conv: dst channels (%d) != cout (%d)", sum, d->cout);
    p.split_n = d->num_dst == 2 ? d->dst_c[0] : d->cout;
  }

  // ---- tile shape ----
  int NT = d->tile_n > 0 ? d->tile_n : (d->cout >= 128 ? 128 : d->cout);
  if (NT % 32 || NT > 256 || d->cout % NT) return fail("conv: bad N tile %d for cout %d", NT, d->cout);
  if (d->num_dst == 2) {
    // the destination split must fall on an N-tile boundary: the widest tile that does (128 -> 64 -> 32)
    while (NT > 32 && (p.split_n % NT || d->cout % NT)) NT >>= 1;
    if (p.split_n % NT) return fail("conv: dst split %d is not a multiple of 32", p.split_n);
  }
  const int n_tiles_n = d->cout / NT;
  const int sms = sm_count();
  int S = d->tile_s;
  auto pixel_tiles = [&](int s) { return d->batch * ceil_div(d->h, kTileRows) * ceil_div(d->w, 8 * s); };
  if (S <= 0) {
    // measured sweep over the benchmark UNet's layers (scripts/bench_layers.py --tile-n/--tile-s): 32-channel tiles want
    // four sub-tiles per CTA tile, 64-channel tiles two, 128-channel tiles one (smaller A boxes, more stages in flight)
    S = NT >= 128 ? 1 : (NT == 64 ? 2 : 4);
    while (S > 1 && (2 * S * NT > 512 || pixel_tiles(S) * n_tiles_n < 4 * sms)) S >>= 1;
  }
  if (S != 1 && S != 2 && S != 4) return fail("conv: tile_s must be 1, 2 or 4");
  if (S * NT > 512) return fail("conv: tile_s*tile_n = %d exceeds the 512 TMEM columns", S * NT);

  // ---- shared-memory budget: A ring, B ring (or resident filter) ----
  int nA = 0, nB = 0, resident = 0, a_slot = 0, a_tx = 0, pitch = 0;
  const int b_tile = NT * p.chunk_bytes;
  for (;; S >>= 1) {
    const bool halo = d->mode == MODE_HALO;
    pitch = 8 * S + (halo ? 2 : 0);
    const int rows = kTileRows + (halo ? 2 : 0);
    a_tx = rows * pitch * p.chunk_bytes;
    a_slot = (a_tx + 1023) & ~1023;
    const int all_b = p.taps * p.chunks_total * b_tile;
    resident = (n_tiles_n == 1 && all_b <= 100 * 1024 && kSmemBudget - all_b >= 2 * a_slot) ? 1 : 0;
    if (resident) {
      nB = 0;
      nA = (kSmemBudget - all_b) / a_slot;
    } else {
      // keep two A slots (one in use, one in flight), give the rest to the B ring: the filter tiles are the
      // stream that is consumed fastest and whose TMA latency must be covered
      nB = (kSmemBudget - 2 * a_slot) / b_tile;
      // nine slots (one per tap, static addressing in the issue loop) whenever they fit beside two A slots
      if (halo && nB >= 9 && static_ring()) nB = 9;
      else if (nB > max_nb()) nB = max_nb();
      if (nB < 2) nB = 2;
      nA = (kSmemBudget - nB * b_tile) / a_slot;
    }
    if (nA > max_na()) nA = max_na();
    if (nA >= 2 || S == 1) break;
  }
  if (nA < 1) return fail("conv: tile does not fit in shared memory");
  p.S = S;
  p.NT = NT;
  p.n_tiles_n = n_tiles_n;
  p.tiles_x = ceil_div(d->w, 8 * S);
  p.tiles_y = ceil_div(d->h, kTileRows);
  p.total_tiles = d->batch * p.tiles_x * p.tiles_y * n_tiles_n;
  p.box_pitch = pitch;
  p.a_box_tx_bytes = a_tx;
  p.a_slot_bytes = a_slot;
  p.b_tile_bytes = b_tile;
  p.nA = nA;
  p.nB = nB;
  p.b_resident = resident;
  p.acc_bufs = 2 * S * NT <= 512 ? 2 : 1;
  p.tmem_cols = pow2_cols(p.acc_bufs * S * NT);
  p.pixel_tiles = d->batch * p.tiles_x * p.tiles_y;

  // ---- cluster multicast of the streamed filter tiles (wide 3x3 layers) ----
  int cluster = d->cluster > 0 ? d->cluster : cluster_request();
  IgemmKernelFn cluster_fn = nullptr;
  if (cluster > 1 && d->mode == MODE_HALO && !resident && NT == 128 && p.pixel_tiles >= 2 * cluster &&
      ((NT / cluster) * p.chunk_bytes) % 1024 == 0 && is_local_device_memory(d->weight)) {
    cluster_fn = pick_cluster_kernel(S, chunk_ch / 16, cluster);
  }
  if (!cluster_fn) cluster = 1;
  p.cluster = cluster;

  // ---- tensor maps ----
  const bool halo = d->mode == MODE_HALO;
  for (int i = 0; i < d->num_src; ++i) {
    const uint64_t C = static_cast<uint64_t>(d->src_c[i]);
    uint64_t dims[5], str[4];
    uint32_t box[5];
    const void* base = d->src[i];
    if (d->mode == MODE_STRIDED) {
      // source is [B, 2H, 2W, C]; view as (C, dx, x, dy, b*H + y)
      dims[0] = C; dims[1] = 2; dims[2] = d->w; dims[3] = 2; dims[4] = static_cast<uint64_t>(d->batch) * d->h;
      str[0] = C * 2;                       // dx
      str[1] = 2 * C * 2;                   // x
      str[2] = 2ull * d->w * C * 2;         // dy
      str[3] = 2 * str[2];                  // (b, y)
      box[0] = chunk_ch; box[1] = 1; box[2] = 8 * S; box[3] = 1; box[4] = kTileRows;
    } else {
      // the window [h, w] at (src_y0, src_x0) of a [batch, src_h, src_w, C] tensor: base moved to the window's corner,
      // the tensor's own pitches, the window's extents (so the halo reads zeros around the WINDOW)
      const uint64_t sh = d->src_h[i] > 0 ? d->src_h[i] : d->h, sw = d->src_w[i] > 0 ? d->src_w[i] : d->w;
      if (d->src_y0[i] < 0 || d->src_x0[i] < 0 || d->src_y0[i] + d->h > static_cast<int>(sh) ||
          d->src_x0[i] + d->w > static_cast<int>(sw))
        return fail("conv: source %d window (%d,%d)+(%d,%d) leaves the %llux%llu tensor", i, d->src_y0[i], d->src_x0[i], d->h,
                    d->w, (unsigned long long)sh, (unsigned long long)sw);
      dims[0] = C; dims[1] = d->w; dims[2] = d->h; dims[3] = d->batch; dims[4] = 1;
      str[0] = C * 2;
      str[1] = str[0] * sw;
      str[2] = str[1] * sh;
      str[3] = str[2] * d->batch;
      box[0] = chunk_ch; box[1] = pitch; box[2] = kTileRows + (halo ? 2 : 0); box[3] = 1; box[4] = 1;
      base = static_cast<const char*>(base) + (static_cast<uint64_t>(d->src_y0[i]) * sw + d->src_x0[i]) * C * 2;
    }
    if (make_tmap_bf16(&p.mapA[i], base, 5, dims, str, box, p.chunk_bytes)) return -1;
  }
  {
    const uint64_t K = static_cast<uint64_t>(c0 + c1);
    uint64_t dims[3] = {K, static_cast<uint64_t>(d->cout), static_cast<uint64_t>(p.taps)};
    uint64_t str[2] = {K * 2, K * 2 * d->cout};
    uint32_t box[3] = {static_cast<uint32_t>(chunk_ch), static_cast<uint32_t>(NT / cluster), 1};
    if (make_tmap_bf16(&p.mapW, d->weight, 3, dims, str, box, p.chunk_bytes)) return -1;
  }

  plan->smem = static_cast<size_t>(nA) * a_slot +
               static_cast<size_t>(resident ? p.taps * p.chunks_total : nB) * b_tile + 1024;
  // The attribute is per device: set it for the current device whenever a plan is created (cheap).
  plan->fn = cluster > 1 ? cluster_fn : pick_kernel(S, chunk_ch / 16, d->mode, resident, nB);
  if (!plan->fn) return fail("conv: no kernel for tile_s=%d chunk=%d", S, chunk_ch);
  const cudaError_t attr_err =
      cudaFuncSetAttribute(plan->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget + 1024);  // + ~4.5 KB static
  if (attr_err != cudaSuccess) {
    return fail("cudaFuncSetAttribute(max dynamic smem) failed: %s", cudaGetErrorString(attr_err));
  }
  plan->cluster = cluster;
  // work units: tiles, or (N tile, `cluster` pixel tiles) groups; one unit per CTA / cluster at a time, all co-resident
  const int units = cluster == 1 ? p.total_tiles : n_tiles_n * ceil_div(p.pixel_tiles, cluster);
  int slots = sms / cluster;
  if (cluster > 1) {
    cudaLaunchConfig_t cfg{};
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cluster;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.gridDim = dim3(sms / cluster * cluster);
    cfg.blockDim = dim3(kIgemmThreads);
    cfg.dynamicSmemBytes = plan->smem;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int max_clusters = 0;
    const cudaError_t oe = cudaOccupancyMaxActiveClusters(&max_clusters, reinterpret_cast<const void*>(plan->fn), &cfg);
    if (oe != cudaSuccess || max_clusters < 1) {
      cudaGetLastError();
      return fail("conv: cudaOccupancyMaxActiveClusters(cluster %d) failed: %s", cluster, cudaGetErrorString(oe));
    }
    if (max_clusters < slots) slots = max_clusters;  // a persistent grid must be resident in ONE wave
  }
  int groups = units < slots ? units : slots;
  if (p.stats != nullptr) {
    // a CTA must own ONE N tile for its whole life (unit -> N tile is unit % n_tiles_n and the unit stride is `groups`)
    if (NT != 32 && NT != 64 && NT != 128) return fail("conv: BatchNorm statistics need an N tile of 32, 64 or 128 (got %d)", NT);
    if (d->bias != nullptr || d->relu || d->num_dst != 1 || d->mask[0] != nullptr || d->out_mode != DPT_OUT_NHWC)
      return fail("conv: the BatchNorm-statistics epilogue takes no bias / ReLU / mask / split or shuffled store");
    groups = groups / n_tiles_n * n_tiles_n;
    if (groups < n_tiles_n) return fail("conv: fewer SMs than N tiles");
  }
  plan->grid = groups * cluster;
  *out = owner.release();
  return 0;
}

extern "C" int dpt_conv_plan_launch(const dpt_conv_plan* plan, void* stream) {
  using namespace dpt;
  if (!plan) return fail("dpt_conv_plan_launch: null plan");
  if (plan->cluster > 1) {
    cudaLaunchConfig_t cfg{};
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = plan->cluster;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.gridDim = dim3(plan->grid);
    cfg.blockDim = dim3(kIgemmThreads);
    cfg.dynamicSmemBytes = plan->smem;
    cfg.stream = static_cast<cudaStream_t>(stream);
    cfg.attrs = at;
    cfg.numAttrs = 1;
    DPT_CUDA_OK(cudaLaunchKernelEx(&cfg, plan->fn, plan->p));
  } else {
    DPT_CUDA_OK(launch_chained(plan->fn, dim3(plan->grid), dim3(kIgemmThreads), plan->smem, static_cast<cudaStream_t>(stream), plan->p));
  }
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" void dpt_conv_plan_destroy(dpt_conv_plan* plan) { delete plan; }

extern "C" int dpt_conv_plan_cluster(const dpt_conv_plan* plan) { return plan ? plan->cluster : 0; }

extern "C" int dpt_conv_plan_info(const dpt_conv_plan* plan, int info[9]) {
  if (!plan || !info) return dpt::fail("dpt_conv_plan_info: null argument");
  const dpt::IgemmParams& p = plan->p;
  info[0] = plan->grid;
  info[1] = static_cast<int>(plan->smem);
  info[2] = p.S;
  info[3] = p.NT;
  info[4] = p.nA;
  info[5] = p.nB;
  info[6] = p.b_resident;
  info[7] = p.acc_bufs;
  info[8] = p.chunk_ch;
  return 0;
}
