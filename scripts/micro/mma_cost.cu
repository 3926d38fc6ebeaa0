// Cycles per tcgen05.mma (kind::f16, bf16, M = 128, K = 16, cta_group::1) as a function of N, operands in shared
// memory: the number every tile-shape decision in csrc/kernels hangs on.  One CTA per SM, one elected thread issues
// ITER back-to-back MMAs on the same accumulator and commits; clock64 around issue..retire.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Iinclude -o build/mma_cost scripts/micro/mma_cost.cu
#include <cstdio>
#include <cstdlib>

#include "../../csrc/common/ptx.cuh"

using namespace dpt;

// ROTATE: every MMA reads a different A tile (8 of them, 4 KB apart) and B tile (9 of them), like the conv kernels do;
// without it all MMAs name the same descriptors.
template <int A_MN_MAJOR, int ROTATE, int ROWB = 32>
__global__ void __launch_bounds__(128, 1) mma_cost_kernel(int N, int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t done;
  __shared__ uint32_t tmem_slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(&done, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = tmem_slot;
  if (warp == 1 && elect_one_sync()) {
    // K-major 32-byte rows (SWIZZLE_32B): A = 128 rows, B = N rows; or MN-major 64-byte rows for both
    const uint32_t idesc = umma_idesc_bf16(N, A_MN_MAJOR, A_MN_MAJOR);
    const uint64_t a = A_MN_MAJOR ? umma_smem_desc(base, 64u, 512u, umma_layout_type(64))
                                  : umma_smem_desc(base, 16u, 8u * ROWB, umma_layout_type(ROWB));
    const uint64_t b = A_MN_MAJOR ? umma_smem_desc(base + 32768u, 1024u, 512u, umma_layout_type(64))
                                  : umma_smem_desc(base + 65536u, 16u, 8u * ROWB, umma_layout_type(ROWB));
    const long long t0 = clock64();
    if (ROTATE) {
      for (int i = 0; i < iters; i += 8) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
          // 4 sub-tiles x (ROWB / 32) K slices of 32 bytes inside ROWB-byte rows, nine filter taps
          umma_bf16(tmem + (u & 3) * 32, a + ((((u & 3) * 128u * ROWB) + ((u >> 2) % (ROWB / 32)) * 32u) >> 4),
                    b + (((((i >> 3) % 9) * 128u * ROWB) + ((u >> 2) % (ROWB / 32)) * 32u) >> 4), idesc, i ? 1u : 0u);
      }
    } else {
      for (int i = 0; i < iters; ++i) umma_bf16(tmem, a, b, idesc, i ? 1u : 0u);
    }
    umma_commit(&done);
    mbar_wait(&done, 0, 1);
    const long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after_sync();
    tmem_dealloc(tmem, 512);
  }
}

// The conv kernel's exact A addressing: one 18 x PITCH pixel halo box of 64-byte rows (32 channels, SWIZZLE_64B); a
// sub-tile's 16 image rows are 8-pixel groups PITCH*64 bytes apart (SBO); tap (kh, kw) starts kh rows down and kw
// pixels right; 4 sub-tiles, 2 K slices, 9 resident filter tiles of N x 32 channels.
// LD_PER_TILE > 0: warps 0..3 (one per TMEM lane quadrant) read the OTHER accumulator buffer with tcgen05.ld while
// the MMAs run, LD_PER_TILE x 32 columns each per tile: what an epilogue does.  Does reading TMEM slow the tensor pipe?
template <int PITCH>
__global__ void __launch_bounds__(160, 1) mma_halo_kernel(int N, int tiles, int ld_per_tile, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t done;
  __shared__ uint32_t tmem_slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(&done, 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = tmem_slot;
  if (warp < 4 && ld_per_tile > 0) {
    uint32_t sink = 0;
    for (int i = 0; i < tiles * ld_per_tile; ++i) {
      uint32_t v[32];
      tmem_ld_32x32(tmem + (static_cast<uint32_t>(warp * 32) << 16) + 256u + static_cast<uint32_t>(i & 3) * 32u, v);
      tmem_ld_wait();
#pragma unroll
      for (int k = 0; k < 32; ++k) sink ^= v[k];
    }
    if (sink == 0x12345678u) out[200] = sink;
  }
  if (warp == 4 && elect_one_sync()) {
    const uint32_t idesc = umma_idesc_bf16(N, 0, 0);
    const uint64_t a0 = umma_smem_desc(base, 16u, PITCH * 64u, umma_layout_type(64));
    const uint64_t b0 = umma_smem_desc(base + 64u * 1024u, 16u, 512u, umma_layout_type(64));
    const long long t0 = clock64();
    for (int t = 0; t < tiles; ++t) {
#pragma unroll
      for (int tp = 0; tp < 9; ++tp) {
        const uint64_t a = a0 + (((tp / 3) * PITCH * 64u + (tp % 3) * 64u) >> 4);
        const uint64_t b = b0 + ((tp * N * 64u) >> 4);
#pragma unroll
        for (int sub = 0; sub < 4; ++sub)
#pragma unroll
          for (int k = 0; k < 2; ++k)
            umma_bf16(tmem + sub * N, a + ((sub * 8 * 64u + 32u * k) >> 4), b + ((32u * k) >> 4), idesc,
                      (t | tp | k) ? 1u : 0u);
      }
    }
    umma_commit(&done);
    mbar_wait(&done, 0, 1);
    out[blockIdx.x] = clock64() - t0;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after_sync();
    tmem_dealloc(tmem, 512);
  }
}

template <int PITCH>
void run_halo(long long* d, int smem) {
  cudaFuncSetAttribute(mma_halo_kernel<PITCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int ld : {0, 16, 64, 256})
  for (int N : {32, 64}) {
    const int tiles = 64;
    for (int rep = 0; rep < 2; ++rep) {
      mma_halo_kernel<PITCH><<<148, 160, smem>>>(N, tiles, ld, d);
      if (cudaDeviceSynchronize() != cudaSuccess) {
        printf("error in halo kernel\n");
        return;
      }
    }
    long long h[148];
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("{\"operands\": \"conv halo box, pitch %d pixels (SBO %d bytes), 9 taps x 4 sub-tiles x 2 K slices\", \"M\": 128, "
           "\"N\": %d, \"K\": 16, \"concurrent_tcgen05_ld_x32_per_warp_per_tile\": %d, \"cycles_per_mma\": %.1f}\n", PITCH, PITCH * 64,
           N, ld, double(mx) / (tiles * 72));
  }
}

int main() {
  long long* d = nullptr;
  cudaMalloc(&d, 256 * sizeof(long long));
  const int iters = 4096;
  const int smem = 220 * 1024;
  cudaFuncSetAttribute(mma_cost_kernel<0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(mma_cost_kernel<1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(mma_cost_kernel<0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(mma_cost_kernel<0, 1, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(mma_cost_kernel<0, 1, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int mn = 0; mn < 5; ++mn)
    for (int N : {16, 32, 64, 96, 128, 192, 256}) {
      if (mn >= 2 && N > 128) continue;  // rotating accumulators use 4 x 32..128 columns
      for (int rep = 0; rep < 2; ++rep) {
        if (mn == 1)
          mma_cost_kernel<1, 0><<<148, 128, smem>>>(N, iters, d);
        else if (mn == 0)
          mma_cost_kernel<0, 0><<<148, 128, smem>>>(N, iters, d);
        else if (mn == 2)
          mma_cost_kernel<0, 1><<<148, 128, smem>>>(N, iters, d);
        else if (mn == 3)
          mma_cost_kernel<0, 1, 64><<<148, 128, smem>>>(N, iters, d);
        else
          mma_cost_kernel<0, 1, 128><<<148, 128, smem>>>(N, iters, d);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
          printf("error: %s\n", cudaGetErrorString(e));
          return 1;
        }
      }
      long long h[148];
      cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
      long long mx = 0;
      for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
      const double cyc = double(mx) / iters;
      printf("{\"operands\": \"%s\", \"M\": 128, \"N\": %d, \"K\": 16, \"cycles_per_mma\": %.1f, \"flop_per_cycle\": %.0f, "
             "\"operand_bytes_per_cycle\": %.1f}\n",
             mn == 1 ? "MN-major" : mn == 0 ? "K-major" : mn == 2 ? "K-major, rotating A/B tiles" : mn == 3 ? "K-major 64-byte rows (SWIZZLE_64B), rotating" : "K-major 128-byte rows (SWIZZLE_128B), rotating", N, cyc, 2.0 * 128 * N * 16 / cyc, (128 + N) * 32.0 / cyc);
    }
  run_halo<34>(d, smem);
  return 0;
}
