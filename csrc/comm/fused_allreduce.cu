// Fused gradient collectives over NVLink 5 / NVSwitch peer memory (sm_100a).
//
// What torch's DistributedDataParallel does with a C++ Reducer + NCCL all-reduce + a separate multi-tensor
// Adam (utils/train_utils.py:196,224-225 of the reference), and what DataParallel does with
// ReduceAddCoalesced + a root-only optimizer + Broadcast every forward (utils/train_utils.py:98,138,144-145),
// is ONE kernel here: every rank owns 1/world of the flat gradient arena, pulls that slice from all peers
// (P2P loads, or one multimem.ld_reduce that lets the NVSwitch do the sum), scales it, applies Adam /
// RMSprop to its slice of the fp32 master weights (optimizer state is therefore sharded for free), and
// pushes the updated weights back into every peer's parameter arena (P2P stores or one multimem.st).
// A reduce-scatter, the optimizer and an all-gather, with 2*(world-1)/world*bytes over the wire and no
// separate optimizer pass over HBM.
//
// Variants, chosen per bucket size by the caller:
//   DPT_AR_TWO_SHOT  : sharded as described (large buckets)
//   DPT_AR_ONE_SHOT  : every rank reads every peer and updates its whole replica (small buckets: one
//                      barrier-to-barrier latency, no write-back)
//   DPT_AR_MULTIMEM  : two-shot with NVLS multicast (multimem.ld_reduce / multimem.st)
//
// Cross-GPU synchronisation: per-block flag slots in a symmetric "signal pad" (pad[block][peer]); a block
// publishes an epoch number to its slot in every peer's pad with st.release.sys and spins on its own pad
// with ld.acquire.sys.  Spins are bounded by a watchdog: a dead peer becomes an error, not a hang.
#include <atomic>

#include "../../include/dpt_b200.h"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dpt {

extern std::atomic<long long> g_launch_count;

constexpr int kCommMaxBlocks = DPT_COMM_MAX_BLOCKS;
constexpr int kCommThreads = 512;

struct Team {
  int rank, world;
  uint32_t* pads[DPT_MAX_PEERS];  // pads[r]: rank r's signal pad as mapped here, [kCommMaxBlocks][DPT_MAX_PEERS]
  uint32_t* epoch;                // local, [kCommMaxBlocks]
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Block-level barrier across all ranks of the team.  Every thread of the block must call it.
__device__ __forceinline__ void team_barrier(const Team& t, uint32_t epoch, int tag) {
  __syncthreads();
  if (threadIdx.x < t.world) {
    const int peer = threadIdx.x;
    st_release_sys(t.pads[peer] + blockIdx.x * DPT_MAX_PEERS + t.rank, epoch);
    const uint32_t* mine = t.pads[t.rank] + blockIdx.x * DPT_MAX_PEERS + peer;
    const uint64_t t0 = globaltimer_ns();
    uint32_t spins = 0;
    while (static_cast<int32_t>(ld_acquire_sys(mine) - epoch) < 0) {
      if ((++spins & 0xff) == 0 && globaltimer_ns() - t0 > 5 * DPT_WATCHDOG_NS) {
        printf("[dpt watchdog] team barrier timed out: rank %d block %d waiting for rank %d (epoch %u, tag %d)\n",
               t.rank, (int)blockIdx.x, peer, epoch, tag);
        __trap();
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float4 ld_f4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st_f4(float* p, const float4& v) { *reinterpret_cast<float4*>(p) = v; }

__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

struct FusedArParams {
  Team team;
  float* grads[DPT_MAX_PEERS];
  float* params[DPT_MAX_PEERS];
  float* mc_grads;
  float* mc_params;
  float* m;  // Adam: exp_avg      | RMSprop: square_avg
  float* v;  // Adam: exp_avg_sq   | RMSprop: momentum buffer (or NULL)
  const float* opt;
  long long n4;  // float4 count
  int algo, optimizer;
  float eps, wd, gscale, alpha, momentum;
};

__device__ __forceinline__ void opt_apply(const FusedArParams& p, float& w, float g, float& m, float& v, float lr,
                                          float b1, float b2, float bc1, float bc2) {
  g = g * p.gscale + p.wd * w;
  if (p.optimizer == DPT_OPT_ADAM) {
    m = b1 * m + (1.f - b1) * g;
    v = b2 * v + (1.f - b2) * g * g;
    w -= (lr / bc1) * (m / (sqrtf(v) / sqrtf(bc2) + p.eps));
  } else {  // RMSprop
    m = p.alpha * m + (1.f - p.alpha) * g * g;
    const float avg = sqrtf(m) + p.eps;
    if (p.momentum > 0.f) {
      v = p.momentum * v + g / avg;
      w -= lr * v;
    } else {
      w -= lr * g / avg;
    }
  }
}

// W = peers unrolled (>= world), U = elements per thread per iteration: U * W = 8 independent 16-byte peer loads are in
// flight per thread before the first add, whatever the team size (NVLink latency ~2 us needs the depth).
template <int U, int W>
__global__ void __launch_bounds__(kCommThreads) fused_allreduce_opt_kernel(const FusedArParams p) {
  const Team& t = p.team;
  const uint32_t epoch0 = t.epoch[blockIdx.x];
  team_barrier(t, epoch0 + 1, 21);  // every rank's gradients are complete, every rank is done reading weights

  const float lr = p.opt[0], bc1 = p.opt[2], bc2 = p.opt[3], b1 = p.opt[4], b2 = p.opt[5];
  const bool has_v = p.v != nullptr;
  long long lo = 0, hi = p.n4;
  if (p.algo != DPT_AR_ONE_SHOT) {
    const long long per = (p.n4 + t.world - 1) / t.world;
    lo = per * t.rank;
    hi = lo + per < p.n4 ? lo + per : p.n4;
  }
  float* my_params = p.params[t.rank];
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i0 = lo + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i0 < hi; i0 += U * stride) {
    float4 g[U];
    if (p.algo == DPT_AR_MULTIMEM) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + u * stride;
        if (i < hi) g[u] = multimem_ld_reduce_f4(p.mc_grads + 4 * i);
      }
    } else {
      float4 part[U][W];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + u * stride;
#pragma unroll
        for (int r = 0; r < W; ++r)
          if (r < t.world && i < hi) part[u][r] = ld_f4(p.grads[r] + 4 * i);  // all loads in flight before the first add
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < W; ++r)  // fixed rank order: identical bits on every rank
          if (r < t.world) {
            g[u].x += part[u][r].x;
            g[u].y += part[u][r].y;
            g[u].z += part[u][r].z;
            g[u].w += part[u][r].w;
          }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i >= hi) break;
      float4 w = ld_f4(my_params + 4 * i);
      float4 m = ld_f4(p.m + 4 * i);
      float4 v = has_v ? ld_f4(p.v + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
      opt_apply(p, w.x, g[u].x, m.x, v.x, lr, b1, b2, bc1, bc2);
      opt_apply(p, w.y, g[u].y, m.y, v.y, lr, b1, b2, bc1, bc2);
      opt_apply(p, w.z, g[u].z, m.z, v.z, lr, b1, b2, bc1, bc2);
      opt_apply(p, w.w, g[u].w, m.w, v.w, lr, b1, b2, bc1, bc2);
      st_f4(p.m + 4 * i, m);
      if (has_v) st_f4(p.v + 4 * i, v);
      if (p.algo == DPT_AR_ONE_SHOT) {
        st_f4(my_params + 4 * i, w);
      } else if (p.algo == DPT_AR_MULTIMEM) {
        multimem_st_f4(p.mc_params + 4 * i, w);
      } else {
#pragma unroll
        for (int r = 0; r < W; ++r)
          if (r < t.world) st_f4(p.params[r] + 4 * i, w);
      }
    }
  }
  __threadfence_system();
  team_barrier(t, epoch0 + 2, 22);  // updated weights are visible everywhere; gradients may be overwritten
  if (threadIdx.x == 0) t.epoch[blockIdx.x] = epoch0 + 2;
}

// Pull-broadcast: every rank copies `n4` float4 from root's buffer into its own (DDP construction-time
// parameter sync, utils/train_utils.py:196; BatchNorm buffer sync each step for the north-star variant).
struct BcastParams {
  Team team;
  float* bufs[DPT_MAX_PEERS];
  long long n4;
  int root;
};
__global__ void __launch_bounds__(kCommThreads) bcast_kernel(const BcastParams p) {
  const Team& t = p.team;
  const uint32_t epoch0 = t.epoch[blockIdx.x];
  team_barrier(t, epoch0 + 1, 23);
  if (t.rank != p.root) {
    const float* src = p.bufs[p.root];
    float* dst = p.bufs[t.rank];
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < p.n4;
         i += static_cast<long long>(gridDim.x) * blockDim.x)
      st_f4(dst + 4 * i, ld_f4(src + 4 * i));
  }
  __threadfence_system();
  team_barrier(t, epoch0 + 2, 24);
  if (threadIdx.x == 0) t.epoch[blockIdx.x] = epoch0 + 2;
}

// The same with ONE store stream: the root writes through the NVLS multicast address and the NVSwitch replicates every
// 16-byte store into all ranks' copies (multimem.st), instead of world-1 ranks each pulling the buffer over their own
// link.  `mc` is the multicast mapping of the buffers in `bufs` (MulticastBuffer.mc_ptr).
struct BcastMcParams {
  Team team;
  const float* src;  // the root's own copy (unicast address); unused on the other ranks
  float* mc;
  long long n4;
  int root;
};
__global__ void __launch_bounds__(kCommThreads) bcast_multimem_kernel(const BcastMcParams p) {
  const Team& t = p.team;
  const uint32_t epoch0 = t.epoch[blockIdx.x];
  team_barrier(t, epoch0 + 1, 25);  // nobody still reads the old contents
  if (t.rank == p.root) {
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < p.n4;
         i += static_cast<long long>(gridDim.x) * blockDim.x)
      multimem_st_f4(p.mc + 4 * i, ld_f4(p.src + 4 * i));
  }
  __threadfence_system();
  team_barrier(t, epoch0 + 2, 26);
  if (threadIdx.x == 0) t.epoch[blockIdx.x] = epoch0 + 2;
}

// Plain sum all-reduce in place (two-shot P2P), for metrics / BatchNorm statistics / tests.
struct SumParams {
  Team team;
  float* bufs[DPT_MAX_PEERS];
  long long n4;
  float scale;
};
__global__ void __launch_bounds__(kCommThreads) allreduce_sum_kernel(const SumParams p) {
  const Team& t = p.team;
  const uint32_t epoch0 = t.epoch[blockIdx.x];
  team_barrier(t, epoch0 + 1, 25);
  const long long per = (p.n4 + t.world - 1) / t.world;
  const long long lo = per * t.rank;
  const long long hi = lo + per < p.n4 ? lo + per : p.n4;
  for (long long i = lo + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < hi;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 part[DPT_MAX_PEERS];
#pragma unroll
    for (int r = 0; r < DPT_MAX_PEERS; ++r)
      if (r < t.world) part[r] = ld_f4(p.bufs[r] + 4 * i);
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int r = 0; r < DPT_MAX_PEERS; ++r)
      if (r < t.world) {
        g.x += part[r].x;
        g.y += part[r].y;
        g.z += part[r].z;
        g.w += part[r].w;
      }
    g.x *= p.scale;
    g.y *= p.scale;
    g.z *= p.scale;
    g.w *= p.scale;
    // every rank's copy of element i is only read by its owner (this rank), so writing in place is safe
    // once the owner has its sum in registers.
#pragma unroll
    for (int r = 0; r < DPT_MAX_PEERS; ++r)
      if (r < t.world) st_f4(p.bufs[r] + 4 * i, g);
  }
  __threadfence_system();
  team_barrier(t, epoch0 + 2, 26);
  if (threadIdx.x == 0) t.epoch[blockIdx.x] = epoch0 + 2;
}

static int fill_team(Team* t, const dpt_team* team) {
  if (!team) return fail("comm: null team");
  if (team->world < 1 || team->world > DPT_MAX_PEERS || team->rank < 0 || team->rank >= team->world)
    return fail("comm: bad team rank %d / world %d", team->rank, team->world);
  t->rank = team->rank;
  t->world = team->world;
  for (int r = 0; r < DPT_MAX_PEERS; ++r) t->pads[r] = r < team->world ? static_cast<uint32_t*>(team->pads[r]) : nullptr;
  t->epoch = static_cast<uint32_t*>(team->epoch);
  if (!t->epoch) return fail("comm: team has no epoch buffer");
  for (int r = 0; r < team->world; ++r)
    if (!t->pads[r]) return fail("comm: team pad %d is null", r);
  return 0;
}

static int comm_blocks(int requested) {
  int b = requested > 0 ? requested : 128;
  if (b > kCommMaxBlocks) b = kCommMaxBlocks;
  const int sms = sm_count();
  if (b > sms) b = sms;  // blocks spin on peers: they must all be resident
  return b;
}

}  // namespace dpt

using namespace dpt;

extern "C" int dpt_fused_allreduce_opt(const dpt_team* team, const dpt_fused_ar_desc* d, void* stream) {
  if (!d) return fail("fused_allreduce: null desc");
  FusedArParams p;
  memset(&p, 0, sizeof(p));
  if (fill_team(&p.team, team)) return -1;
  if (d->n % 4) return fail("fused_allreduce: n=%lld must be a multiple of 4", d->n);
  for (int r = 0; r < team->world; ++r) {
    p.grads[r] = static_cast<float*>(d->grads[r]);
    p.params[r] = static_cast<float*>(d->params[r]);
    if (!p.grads[r] || !p.params[r]) return fail("fused_allreduce: null peer buffer for rank %d", r);
  }
  p.mc_grads = static_cast<float*>(d->mc_grads);
  p.mc_params = static_cast<float*>(d->mc_params);
  if (d->algo == DPT_AR_MULTIMEM && (!p.mc_grads || !p.mc_params))
    return fail("fused_allreduce: multimem variant needs multicast addresses");
  p.m = d->state0;
  p.v = d->state1;
  if (!p.m || (d->optimizer == DPT_OPT_ADAM && !p.v)) return fail("fused_allreduce: optimizer state missing");
  p.opt = d->opt_state;
  p.n4 = d->n / 4;
  p.algo = d->algo;
  p.optimizer = d->optimizer;
  p.eps = d->eps;
  p.wd = d->weight_decay;
  p.gscale = d->grad_scale;
  p.alpha = d->alpha;
  p.momentum = d->momentum;
  const int nblk = comm_blocks(d->blocks);
  cudaStream_t cs = static_cast<cudaStream_t>(stream);
  if (team->world <= 2) {
    fused_allreduce_opt_kernel<4, 2><<<nblk, kCommThreads, 0, cs>>>(p);
  } else if (team->world <= 4) {
    fused_allreduce_opt_kernel<2, 4><<<nblk, kCommThreads, 0, cs>>>(p);
  } else {
    fused_allreduce_opt_kernel<1, 8><<<nblk, kCommThreads, 0, cs>>>(p);
  }
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_team_bcast(const dpt_team* team, void* const* bufs, long long n, int root, int blocks,
                              void* stream) {
  BcastParams p;
  memset(&p, 0, sizeof(p));
  if (fill_team(&p.team, team)) return -1;
  if (n % 4) return fail("bcast: n=%lld must be a multiple of 4", n);
  for (int r = 0; r < team->world; ++r) p.bufs[r] = static_cast<float*>(bufs[r]);
  p.n4 = n / 4;
  p.root = root;
  bcast_kernel<<<comm_blocks(blocks), kCommThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_team_bcast_multimem(const dpt_team* team, const void* root_src, void* mc, long long n, int root, int blocks,
                                       void* stream) {
  BcastMcParams p;
  memset(&p, 0, sizeof(p));
  if (fill_team(&p.team, team)) return -1;
  if (n % 4) return fail("bcast_multimem: n=%lld must be a multiple of 4", n);
  if (!mc) return fail("bcast_multimem: needs the multicast address of the buffer");
  p.src = static_cast<const float*>(root_src);
  p.mc = static_cast<float*>(mc);
  p.n4 = n / 4;
  p.root = root;
  bcast_multimem_kernel<<<comm_blocks(blocks), kCommThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_team_allreduce_sum(const dpt_team* team, void* const* bufs, long long n, float scale, int blocks,
                                      void* stream) {
  SumParams p;
  memset(&p, 0, sizeof(p));
  if (fill_team(&p.team, team)) return -1;
  if (n % 4) return fail("allreduce_sum: n=%lld must be a multiple of 4", n);
  for (int r = 0; r < team->world; ++r) p.bufs[r] = static_cast<float*>(bufs[r]);
  p.n4 = n / 4;
  p.scale = scale;
  allreduce_sum_kernel<<<comm_blocks(blocks), kCommThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  DPT_CUDA_OK(cudaGetLastError());
  return 0;
}

extern "C" int dpt_enable_peer_access(int device, int peer) {
  int cur = 0;
  DPT_CUDA_OK(cudaGetDevice(&cur));
  DPT_CUDA_OK(cudaSetDevice(device));
  int can = 0;
  cudaError_t e = cudaDeviceCanAccessPeer(&can, device, peer);
  if (e != cudaSuccess || !can) {
    cudaSetDevice(cur);
    return fail("device %d cannot access peer %d", device, peer);
  }
  e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    e = cudaSuccess;
  }
  cudaSetDevice(cur);
  if (e != cudaSuccess) return fail("cudaDeviceEnablePeerAccess(%d->%d): %s", device, peer, cudaGetErrorString(e));
  return 0;
}
