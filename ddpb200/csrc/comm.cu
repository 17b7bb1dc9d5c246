// K10-K14: gradient synchronisation over NVLink peer memory, no NCCL.
//
// Every rank owns one "symmetric" buffer: [signal pad | gradient arena], allocated so
// that all peers can map it (CUDA IPC or torch symmetric memory, see
// ddpb200/parallel/comm.py) and optionally bound to an NVLS multicast address.
// The kernels below replace the reference's collectives:
//
//   part2b  dist.all_reduce(grad, SUM); grad /= W  per parameter
//           (/root/reference/part2/part2b/main.py:97-103)        -> allreduce_avg_kernel
//   part3   DDP's bucketed all-reduce (/root/reference/part3/main.py:174)
//                                                                  -> same kernel, per bucket,
//                                                                     on a side stream
//   part2a  gather -> mean on rank 0 -> scatter
//           (/root/reference/part2/part2a/main.py:97-115)         -> reduce_root_bcast_kernel
//   DDP constructor's parameter broadcast (C7)                    -> broadcast_kernel
//
// All of them are ONE launch that (1) waits for every peer, (2) moves/reduces data with
// P2P loads/stores (or multimem.ld_reduce / multimem.st through the switch), scaling by
// 1/W on the fly, (3) waits again and (4) optionally applies the SGD update to the whole
// range, so "all-reduce + divide + optimizer step" is a single kernel.
//
// allreduce is a two-shot reduce-scatter + all-gather done IN PLACE: rank r reduces only
// slice r (reading it from every peer) and writes the average back into slice r of every
// peer.  Nobody else touches slice r in between, so no staging copy is needed, and all
// ranks end up with bitwise identical gradients (they all receive rank r's result).
#include "common.cuh"

namespace ddpb {

constexpr int kMaxRanks = 8;
constexpr int kSignalPadBytes = 64 * 1024;          // 16384 flags: 2048 slots x 8 ranks
constexpr int kCommThreads = 512;

struct CommCtx {
  float* buf[kMaxRanks];   // peer-mapped base of each rank's symmetric buffer (after the signal pad)
  uint32_t* sig[kMaxRanks];
  float* mc;               // multicast alias of buf (null: not available)
  int rank, world;
};

struct SgdArgs {
  float* p;   // parameter arena (local), null = no fused update
  float* m;   // momentum arena (local)
  __nv_bfloat16* shadow;  // bf16 copy of the parameters (conv operand), may be null
  float lr, mu, wd;
  long long off, n4;      // range the update covers (may be wider than the synchronised range)
};

#ifndef DDPB_COMM_TIMEOUT_NS
#define DDPB_COMM_TIMEOUT_NS 60000000000ull
#endif

__device__ __forceinline__ uint32_t cas_release_sys(uint32_t* addr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.cas.release.sys.global.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t cas_acquire_sys(uint32_t* addr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.cas.acquire.sys.global.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val) : "memory");
  return old;
}

__device__ __forceinline__ void st_release_sys_u32(uint32_t* addr, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* addr) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}

// Cross-GPU barrier for one CTA.  Word [slot][q] of a rank's signal pad is written only by rank q; word
// [slot][own rank] is the rank's own barrier count for that slot.  Barrier number e of a slot: every rank
// stores e into "its" word on every peer (a posted release store: no round trip, unlike the CAS toggle of
// round 1, whose remote atomics cost ~12 us per barrier) and waits until every peer's word in its own pad
// has reached e.  The counts only grow, so a rank that runs ahead into barrier e + 1 cannot disturb a peer
// still waiting for e; all ranks issue the same barrier sequence per slot, and the state lives in the pad,
// so CUDA-graph replays continue the count.  (Wrap-around after 2^32 barriers per slot is handled by the
// signed difference.)
__device__ __forceinline__ void peer_barrier(const CommCtx& c, int slot) {
  __syncthreads();
  if (threadIdx.x < c.world) {
    const int peer = threadIdx.x;
    uint32_t* mine = c.sig[c.rank] + slot * kMaxRanks;
    const uint32_t epoch = mine[c.rank] + 1u;      // same value in every participating thread
    __syncwarp(__activemask());
    __threadfence_system();                        // this CTA's earlier writes (ordered by the barrier above)
    if (peer == c.rank) {
      mine[c.rank] = epoch;
    } else {
      st_release_sys_u32(c.sig[peer] + slot * kMaxRanks + c.rank, epoch);
      const unsigned long long t0 = global_timer_ns();
      uint32_t spins = 0;
      while (static_cast<int32_t>(ld_acquire_sys_u32(mine + peer) - epoch) < 0) {
        if ((++spins & 0x3FFu) == 0 && global_timer_ns() - t0 > DDPB_COMM_TIMEOUT_NS) {
          printf("ddpb200: rank %d timed out waiting for rank %d (slot %d, barrier %u)\n", c.rank, peer, slot, epoch);
          __trap();
        }
      }
    }
    __threadfence_system();
  }
  __syncthreads();
}

// All CTAs of THIS kernel on THIS GPU meet (sense-reversing, self-resetting: safe across
// launches and graph replays).  peer_barrier only pairs CTA b with CTA b of the peers, so
// "my peers' CTA b finished" says nothing about their CTA b'; the fused SGD reads the whole
// range and therefore must wait until every local CTA has passed its peer barrier.  The two
// words live at the end of the local signal pad.  Requires all CTAs co-resident (grid <= SMs;
// used only by launches that own the GPU: the last reduction of an iteration).
__device__ __forceinline__ void local_grid_barrier(const CommCtx& c) {
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* count = c.sig[c.rank] + (kSignalPadBytes / 4 - 2);
    uint32_t* gen = count + 1;
    uint32_t g;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(g) : "l"(gen) : "memory");
    __threadfence();
    if (atomicAdd(count, 1u) == gridDim.x - 1) {
      *count = 0;
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(gen), "r"(g + 1) : "memory");
    } else {
      const unsigned long long t0 = global_timer_ns();
      uint32_t now, spins = 0;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(now) : "l"(gen) : "memory");
        if ((++spins & 0x3FFu) == 0 && global_timer_ns() - t0 > DDPB_COMM_TIMEOUT_NS) {
          printf("ddpb200: rank %d local grid barrier timed out (block %d)\n", c.rank, blockIdx.x);
          __trap();
        }
      } while (now == g);
    }
    __threadfence();
  }
  __syncthreads();
}

__device__ __forceinline__ float4 ld_sys(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_sys(float* p, const float4& v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
__device__ __forceinline__ float4 mc_ld_reduce(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void mc_st(float* p, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

__device__ __forceinline__ void sgd_range(const CommCtx& c, const SgdArgs& s, long long off, long long n4) {
  const float* g = c.buf[c.rank] + off;
  float* p = s.p + off;
  float* m = s.m + off;
  constexpr int kU = 2;   // iterations in flight per thread (the stores of one would otherwise fence the next loads)
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * kU;
  for (long long base = blockIdx.x * static_cast<long long>(blockDim.x) * kU + threadIdx.x; base < n4; base += stride) {
    float4 gv[kU], pv[kU], mv[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = base + static_cast<long long>(u) * blockDim.x;
      if (i < n4) {
        gv[u] = ld_sys(g + 4 * i);   // peers wrote this: do not trust L1
        pv[u] = reinterpret_cast<const float4*>(p)[i];
        mv[u] = reinterpret_cast<const float4*>(m)[i];
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = base + static_cast<long long>(u) * blockDim.x;
      if (i >= n4) continue;
      float4 gg;
      gg.x = fmaf(s.wd, pv[u].x, gv[u].x);
      gg.y = fmaf(s.wd, pv[u].y, gv[u].y);
      gg.z = fmaf(s.wd, pv[u].z, gv[u].z);
      gg.w = fmaf(s.wd, pv[u].w, gv[u].w);
      mv[u].x = fmaf(s.mu, mv[u].x, gg.x);
      mv[u].y = fmaf(s.mu, mv[u].y, gg.y);
      mv[u].z = fmaf(s.mu, mv[u].z, gg.z);
      mv[u].w = fmaf(s.mu, mv[u].w, gg.w);
      pv[u].x = fmaf(-s.lr, mv[u].x, pv[u].x);
      pv[u].y = fmaf(-s.lr, mv[u].y, pv[u].y);
      pv[u].z = fmaf(-s.lr, mv[u].z, pv[u].z);
      pv[u].w = fmaf(-s.lr, mv[u].w, pv[u].w);
      reinterpret_cast<float4*>(p)[i] = pv[u];
      reinterpret_cast<float4*>(m)[i] = mv[u];
      if (s.shadow != nullptr)
        reinterpret_cast<uint2*>(s.shadow + off)[i] =
            make_uint2(pack_bf16x2(pv[u].x, pv[u].y), pack_bf16x2(pv[u].z, pv[u].w));
    }
  }
}

// kMc: use the NVLS multicast address (in-switch reduction and broadcast).
template <bool kMc>
__global__ void __launch_bounds__(kCommThreads)
allreduce_avg_kernel(const CommCtx c, long long off, long long n4, int slot_base, SgdArgs sgd) {
  const int slot = slot_base + blockIdx.x;
  peer_barrier(c, slot);  // every rank's gradients for this range are complete
  const float inv_w = 1.f / static_cast<float>(c.world);
  const long long lo = n4 * c.rank / c.world, hi = n4 * (c.rank + 1) / c.world;
  // Every remote access is a round trip over NVLink (through the switch for multimem): a thread
  // keeps kU independent 16-byte requests per peer in flight instead of one.  With one, the 32-CTA
  // reductions that overlap backward moved 256 KB per round trip and took longer than the
  // backward pass they were meant to hide behind.
  // (P2P path: the loads from all `world` peers are already independent; kU = 1 keeps the CTA at
  // half the register file so that it can share an SM)
  constexpr int kU = kMc ? 4 : 1;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * kU;
  for (long long base = lo + blockIdx.x * static_cast<long long>(blockDim.x) * kU + threadIdx.x; base < hi;
       base += stride) {
    float4 acc[kU];
    if (kMc) {
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long i = base + static_cast<long long>(u) * blockDim.x;
        if (i < hi) acc[u] = mc_ld_reduce(c.mc + off + 4 * i);
      }
    } else {
      float4 v[kU][kMaxRanks];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long i = base + static_cast<long long>(u) * blockDim.x;
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
          if (i < hi && r < c.world) v[u][r] = ld_sys(c.buf[r] + off + 4 * i);
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        acc[u] = v[u][0];
#pragma unroll
        for (int r = 1; r < kMaxRanks; ++r) {   // fixed order: identical rounding on every rank
          if (r < c.world) {
            acc[u].x += v[u][r].x;
            acc[u].y += v[u][r].y;
            acc[u].z += v[u][r].z;
            acc[u].w += v[u][r].w;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = base + static_cast<long long>(u) * blockDim.x;
      if (i >= hi) continue;
      const long long e = off + 4 * i;
      acc[u].x *= inv_w;
      acc[u].y *= inv_w;
      acc[u].z *= inv_w;
      acc[u].w *= inv_w;
      if (kMc) {
        mc_st(c.mc + e, acc[u]);
      } else {
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
          if (r < c.world) st_sys(c.buf[r] + e, acc[u]);
      }
    }
  }
  peer_barrier(c, slot);  // every slice has landed everywhere
  if (sgd.p != nullptr) {
    local_grid_barrier(c);   // ... as seen by ALL of this GPU's CTAs, not just this one
    sgd_range(c, sgd, sgd.off, sgd.n4);
  }
}

// ------------------------------------------------------------------------------------------
// K10 + batched table (SURVEY.md par. 8): MANY ranges in ONE launch with ONE pair of cross-GPU
// barriers.  The reference's part2b issues 34 blocking all-reduces per iteration, 32 of them <= 2 KB
// (/root/reference/part2/part2b/main.py:101-103); launched one by one here, each pays two ~12 us
// barriers (1.6 ms per step at 2 GPUs, profiles/r2_call10_*).  This kernel takes the whole list:
//   * "small" ranges (the host puts them first; at most kHold float4 per thread in total) are reduced
//     ONE-SHOT: every rank reduces every element itself -- one multimem.ld_reduce through the switch,
//     or W peer loads summed in rank order -- keeps the mean in registers across the closing barrier,
//     then writes it to its OWN buffer and applies the SGD update from the same registers.  No remote
//     stores, no local grid barrier, no second pass over the data;
//   * the other ranges use the two-shot slice scheme of allreduce_avg_kernel.
// ------------------------------------------------------------------------------------------
constexpr int kMaxRanges = 64;
constexpr int kHold = 4;
struct RangeTable {
  long long off[kMaxRanges];   // floats, multiple of 4
  int n4[kMaxRanges];          // float4 per range
  int pre4[kMaxRanges + 1];    // prefix sums of n4 over the small ranges
  int count, small_count;
};

__device__ __forceinline__ void sgd_apply4(const SgdArgs& s, long long e, const float4& g) {
  float4 pv = *reinterpret_cast<const float4*>(s.p + e);
  float4 mv = *reinterpret_cast<const float4*>(s.m + e);
  float4 gg;
  gg.x = fmaf(s.wd, pv.x, g.x);
  gg.y = fmaf(s.wd, pv.y, g.y);
  gg.z = fmaf(s.wd, pv.z, g.z);
  gg.w = fmaf(s.wd, pv.w, g.w);
  mv.x = fmaf(s.mu, mv.x, gg.x);
  mv.y = fmaf(s.mu, mv.y, gg.y);
  mv.z = fmaf(s.mu, mv.z, gg.z);
  mv.w = fmaf(s.mu, mv.w, gg.w);
  pv.x = fmaf(-s.lr, mv.x, pv.x);
  pv.y = fmaf(-s.lr, mv.y, pv.y);
  pv.z = fmaf(-s.lr, mv.z, pv.z);
  pv.w = fmaf(-s.lr, mv.w, pv.w);
  *reinterpret_cast<float4*>(s.p + e) = pv;
  *reinterpret_cast<float4*>(s.m + e) = mv;
  if (s.shadow != nullptr)
    *reinterpret_cast<uint2*>(s.shadow + e) = make_uint2(pack_bf16x2(pv.x, pv.y), pack_bf16x2(pv.z, pv.w));
}

template <bool kMc>
__global__ void __launch_bounds__(kCommThreads)
allreduce_multi_kernel(const CommCtx c, const __grid_constant__ RangeTable tab, int slot_base, SgdArgs sgd) {
  const int slot = slot_base + blockIdx.x;
  const float inv_w = 1.f / static_cast<float>(c.world);
  const long long T = static_cast<long long>(gridDim.x) * blockDim.x;
  const long long g = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  const int small_total4 = tab.pre4[tab.small_count];
  peer_barrier(c, slot);   // every rank's gradients for all ranges are complete
  // ---- one-shot part: chunk idx of the concatenated small ranges -> this thread's registers
  float4 held[kHold];
  long long held_e[kHold];
#pragma unroll
  for (int h = 0; h < kHold; ++h) {
    const long long idx = g + h * T;
    held_e[h] = -1;
    if (idx < small_total4) {
      int r = 0;
      while (idx >= tab.pre4[r + 1]) ++r;
      const long long e = tab.off[r] + 4 * (idx - tab.pre4[r]);
      float4 acc;
      if (kMc) {
        acc = mc_ld_reduce(c.mc + e);
      } else {
        acc = ld_sys(c.buf[0] + e);
        for (int q = 1; q < c.world; ++q) {   // rank order: identical rounding on every rank
          const float4 v = ld_sys(c.buf[q] + e);
          acc.x += v.x;
          acc.y += v.y;
          acc.z += v.z;
          acc.w += v.w;
        }
      }
      acc.x *= inv_w;
      acc.y *= inv_w;
      acc.z *= inv_w;
      acc.w *= inv_w;
      held[h] = acc;
      held_e[h] = e;
    }
  }
  // ---- two-shot part: rank r owns slice r of every large range
  for (int r = tab.small_count; r < tab.count; ++r) {
    const long long off = tab.off[r], n4 = tab.n4[r];
    const long long lo = n4 * c.rank / c.world, hi = n4 * (c.rank + 1) / c.world;
    for (long long i = lo + g; i < hi; i += T) {
      const long long e = off + 4 * i;
      float4 acc;
      if (kMc) {
        acc = mc_ld_reduce(c.mc + e);
      } else {
        acc = ld_sys(c.buf[0] + e);
        for (int q = 1; q < c.world; ++q) {
          const float4 v = ld_sys(c.buf[q] + e);
          acc.x += v.x;
          acc.y += v.y;
          acc.z += v.z;
          acc.w += v.w;
        }
      }
      acc.x *= inv_w;
      acc.y *= inv_w;
      acc.z *= inv_w;
      acc.w *= inv_w;
      if (kMc) {
        mc_st(c.mc + e, acc);
      } else {
        for (int q = 0; q < c.world; ++q) st_sys(c.buf[q] + e, acc);
      }
    }
  }
  peer_barrier(c, slot);   // everyone has read the small ranges; every large slice has landed everywhere
#pragma unroll
  for (int h = 0; h < kHold; ++h) {
    if (held_e[h] >= 0) {
      *reinterpret_cast<float4*>(c.buf[c.rank] + held_e[h]) = held[h];
      if (sgd.p != nullptr) sgd_apply4(sgd, held_e[h], held[h]);
    }
  }
  if (sgd.p != nullptr && tab.small_count < tab.count) {
    local_grid_barrier(c);
    for (int r = tab.small_count; r < tab.count; ++r) sgd_range(c, sgd, tab.off[r], tab.n4[r]);
  }
}

// part2a: the root pulls every rank's gradients, averages, pushes the mean to everyone.
template <bool kMc>
__global__ void __launch_bounds__(kCommThreads)
reduce_root_bcast_kernel(const CommCtx c, long long off, long long n4, int slot_base, int root, SgdArgs sgd) {
  const int slot = slot_base + blockIdx.x;
  peer_barrier(c, slot);
  if (c.rank == root && kMc && c.world > 2) {
    // NVLS: the switch sums all ranks' copies on the way to the root (one inbound stream of S bytes
    // instead of W-1 peer reads) and fans the mean out again with one multicast store per element.
    // With two ranks there is only one peer to read and the plain loads are faster (2 GPUs: 0.839 vs
    // 0.878 ms per step, profiles/r2_call11_*)
    const float inv_w = 1.f / static_cast<float>(c.world);
    constexpr int kU = 4;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * kU;
    for (long long base = blockIdx.x * static_cast<long long>(blockDim.x) * kU + threadIdx.x; base < n4;
         base += stride) {
      float4 acc[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long i = base + static_cast<long long>(u) * blockDim.x;
        if (i < n4) acc[u] = mc_ld_reduce(c.mc + off + 4 * i);
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long i = base + static_cast<long long>(u) * blockDim.x;
        if (i >= n4) continue;
        acc[u].x *= inv_w;
        acc[u].y *= inv_w;
        acc[u].z *= inv_w;
        acc[u].w *= inv_w;
        mc_st(c.mc + off + 4 * i, acc[u]);
      }
    }
  } else if (c.rank == root) {
    const float inv_w = 1.f / static_cast<float>(c.world);
    constexpr int kU = 1;   // the loads from all peers are issued together (see allreduce_avg_kernel)
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * kU;
    for (long long base = blockIdx.x * static_cast<long long>(blockDim.x) * kU + threadIdx.x; base < n4;
         base += stride) {
      float4 v[kU][kMaxRanks];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long i = base + static_cast<long long>(u) * blockDim.x;
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
          if (i < n4 && r < c.world) v[u][r] = ld_sys(c.buf[r] + off + 4 * i);
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long i = base + static_cast<long long>(u) * blockDim.x;
        if (i >= n4) continue;
        const long long e = off + 4 * i;
        float4 acc = v[u][0];
#pragma unroll
        for (int r = 1; r < kMaxRanks; ++r) {
          if (r < c.world) {
            acc.x += v[u][r].x;
            acc.y += v[u][r].y;
            acc.z += v[u][r].z;
            acc.w += v[u][r].w;
          }
        }
        acc.x *= inv_w;
        acc.y *= inv_w;
        acc.z *= inv_w;
        acc.w *= inv_w;
        if (kMc) {
          mc_st(c.mc + e, acc);
        } else {
#pragma unroll
          for (int r = 0; r < kMaxRanks; ++r)
            if (r < c.world) st_sys(c.buf[r] + e, acc);
        }
      }
    }
  }
  peer_barrier(c, slot);
  if (sgd.p != nullptr) {
    local_grid_barrier(c);
    sgd_range(c, sgd, sgd.off, sgd.n4);
  }
}

// Root's range is copied into every peer's buffer.
template <bool kMc>
__global__ void __launch_bounds__(kCommThreads)
broadcast_kernel(const CommCtx c, long long off, long long n4, int slot_base, int root) {
  const int slot = slot_base + blockIdx.x;
  peer_barrier(c, slot);
  if (c.rank == root) {
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
      const long long e = off + 4 * i;
      const float4 v = *reinterpret_cast<const float4*>(c.buf[root] + e);
      if (kMc) {
        mc_st(c.mc + e, v);
      } else {
        for (int r = 0; r < c.world; ++r)
          if (r != root) st_sys(c.buf[r] + e, v);
      }
    }
  }
  peer_barrier(c, slot);
}

__global__ void __launch_bounds__(kCommThreads) barrier_kernel(const CommCtx c, int slot_base) {
  peer_barrier(c, slot_base + blockIdx.x);
}

static int fill_ctx(CommCtx* c, const void* const* bufs, const void* mc, int rank, int world) {
  DDPB_REQUIRE(world >= 1 && world <= kMaxRanks, "comm: world size must be 1..%d (got %d)", kMaxRanks, world);
  DDPB_REQUIRE(rank >= 0 && rank < world, "comm: bad rank %d for world %d", rank, world);
  for (int r = 0; r < kMaxRanks; ++r) {
    c->buf[r] = nullptr;
    c->sig[r] = nullptr;
  }
  for (int r = 0; r < world; ++r) {
    DDPB_REQUIRE(bufs[r] != nullptr, "comm: peer %d buffer not mapped", r);
    uint8_t* base = static_cast<uint8_t*>(const_cast<void*>(bufs[r]));
    c->sig[r] = reinterpret_cast<uint32_t*>(base);
    c->buf[r] = reinterpret_cast<float*>(base + kSignalPadBytes);
  }
  c->mc = (mc != nullptr)
              ? reinterpret_cast<float*>(static_cast<uint8_t*>(const_cast<void*>(mc)) + kSignalPadBytes)
              : nullptr;
  c->rank = rank;
  c->world = world;
  return 0;
}

static int check_range(long long off, long long n, int blocks, int slot_base) {
  DDPB_REQUIRE(off >= 0 && off % 4 == 0 && n >= 0 && n % 4 == 0,
               "comm: offset/count must be multiples of 4 floats (got %lld/%lld)", off, n);
  DDPB_REQUIRE(blocks >= 1 && slot_base >= 0 && (slot_base + blocks) * kMaxRanks * 4 <= kSignalPadBytes - 64,
               "comm: %d blocks from slot %d do not fit the signal pad", blocks, slot_base);
  return 0;
}

}  // namespace ddpb

using namespace ddpb;

extern "C" int ddpb_comm_signal_pad_bytes() { return kSignalPadBytes; }
extern "C" int ddpb_comm_max_ranks() { return kMaxRanks; }

// bufs: `world` base pointers of the symmetric buffers as mapped in THIS process
// (bufs[rank] is the local one); mc: multicast mapping of the same buffer or null.
// off/n: range inside the gradient arena in floats.  p/m: local parameter and momentum
// arenas for the fused SGD step (null = only synchronise).  use_mc selects NVLS.
// sgd_off/sgd_n: range the fused update covers; sgd_n < 0 means "the synchronised range".  A
// wider range lets the LAST reduction of an iteration also apply the update for gradients
// that earlier (overlapped, SGD-free) reductions already averaged.
extern "C" int ddpb_allreduce_avg(const void* const* bufs, const void* mc, int rank, int world, long long off,
                                  long long n, int blocks, int slot_base, int use_mc, float* p, float* m, void* shadow,
                                  float lr, float momentum, float weight_decay, long long sgd_off, long long sgd_n,
                                  int threads, cudaStream_t stream) {
  // threads per CTA: 512 (default, 0) when the launch owns the GPU; 256 for reductions that
  // overlap backward, so that a CTA (256 x 40 registers) fits on an SM next to a persistent
  // convolution CTA (320 x 152 registers) instead of displacing it.
  if (threads <= 0) threads = kCommThreads;
  DDPB_REQUIRE(threads % 32 == 0 && threads >= 64 && threads <= kCommThreads, "comm: bad CTA size %d", threads);
  CommCtx c;
  if (int rc = fill_ctx(&c, bufs, mc, rank, world)) return rc;
  if (int rc = check_range(off, n, blocks, slot_base)) return rc;
  DDPB_REQUIRE(!use_mc || c.mc != nullptr, "comm: multicast requested but no multicast mapping");
  if (sgd_n < 0) {   // default: update exactly the synchronised range
    sgd_off = off;
    sgd_n = n;
  }
  DDPB_REQUIRE(sgd_off >= 0 && sgd_off % 4 == 0 && sgd_n % 4 == 0, "comm: SGD range must be multiples of 4 floats");
  SgdArgs s{p, m, static_cast<__nv_bfloat16*>(shadow), lr, momentum, weight_decay, sgd_off, sgd_n / 4};
  if (use_mc)
    allreduce_avg_kernel<true><<<blocks, threads, 0, stream>>>(c, off, n / 4, slot_base, s);
  else
    allreduce_avg_kernel<false><<<blocks, threads, 0, stream>>>(c, off, n / 4, slot_base, s);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Many ranges, one launch (see allreduce_multi_kernel).  offs / ns: `count` ranges in floats (multiples
// of 4, pairwise disjoint).  Ranges of at most small_max floats are reduced one-shot as long as their
// total fits in the threads' registers (4 float4 per thread); the rest two-shot.  p / m / shadow as for
// ddpb_allreduce_avg: the SGD update covers exactly the listed ranges.
extern "C" int ddpb_allreduce_multi(const void* const* bufs, const void* mc, int rank, int world,
                                    const long long* offs, const long long* ns, int count, long long small_max,
                                    int blocks, int slot_base, int use_mc, float* p, float* m, void* shadow, float lr,
                                    float momentum, float weight_decay, cudaStream_t stream) {
  CommCtx c;
  if (int rc = fill_ctx(&c, bufs, mc, rank, world)) return rc;
  DDPB_REQUIRE(count >= 1 && count <= kMaxRanges, "comm: 1..%d ranges per launch (got %d)", kMaxRanges, count);
  DDPB_REQUIRE(!use_mc || c.mc != nullptr, "comm: multicast requested but no multicast mapping");
  if (int rc = check_range(0, 0, blocks, slot_base)) return rc;
  RangeTable tab;
  const long long hold_cap = static_cast<long long>(blocks) * kCommThreads * kHold;   // float4
  long long small4 = 0;
  int ns_small = 0;
  bool is_small[kMaxRanges];
  for (int i = 0; i < count; ++i) {
    if (int rc = check_range(offs[i], ns[i], blocks, slot_base)) return rc;
    DDPB_REQUIRE(ns[i] / 4 < (1LL << 31), "comm: range too long");
    is_small[i] = ns[i] <= small_max && small4 + ns[i] / 4 <= hold_cap;
    if (is_small[i]) {
      small4 += ns[i] / 4;
      ++ns_small;
    }
  }
  int k = 0;
  tab.pre4[0] = 0;
  for (int pass = 0; pass < 2; ++pass)
    for (int i = 0; i < count; ++i)
      if (is_small[i] == (pass == 0)) {
        tab.off[k] = offs[i];
        tab.n4[k] = static_cast<int>(ns[i] / 4);
        if (pass == 0) tab.pre4[k + 1] = tab.pre4[k] + tab.n4[k];
        ++k;
      }
  for (int i = ns_small; i < kMaxRanges; ++i) tab.pre4[i + 1] = tab.pre4[ns_small];
  tab.count = count;
  tab.small_count = ns_small;
  SgdArgs s{p, m, static_cast<__nv_bfloat16*>(shadow), lr, momentum, weight_decay, 0, 0};
  if (use_mc)
    allreduce_multi_kernel<true><<<blocks, kCommThreads, 0, stream>>>(c, tab, slot_base, s);
  else
    allreduce_multi_kernel<false><<<blocks, kCommThreads, 0, stream>>>(c, tab, slot_base, s);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int ddpb_reduce_root_bcast(const void* const* bufs, const void* mc, int rank, int world, int root,
                                      long long off, long long n, int blocks, int slot_base, int use_mc, float* p,
                                      float* m, void* shadow, float lr, float momentum, float weight_decay,
                                      long long sgd_off, long long sgd_n, cudaStream_t stream) {
  CommCtx c;
  if (int rc = fill_ctx(&c, bufs, mc, rank, world)) return rc;
  if (int rc = check_range(off, n, blocks, slot_base)) return rc;
  DDPB_REQUIRE(root >= 0 && root < world, "comm: bad root %d", root);
  DDPB_REQUIRE(!use_mc || c.mc != nullptr, "comm: multicast requested but no multicast mapping");
  if (sgd_n < 0) {   // default: update exactly the synchronised range
    sgd_off = off;
    sgd_n = n;
  }
  DDPB_REQUIRE(sgd_off >= 0 && sgd_off % 4 == 0 && sgd_n % 4 == 0, "comm: SGD range must be multiples of 4 floats");
  SgdArgs s{p, m, static_cast<__nv_bfloat16*>(shadow), lr, momentum, weight_decay, sgd_off, sgd_n / 4};
  if (use_mc)
    reduce_root_bcast_kernel<true><<<blocks, kCommThreads, 0, stream>>>(c, off, n / 4, slot_base, root, s);
  else
    reduce_root_bcast_kernel<false><<<blocks, kCommThreads, 0, stream>>>(c, off, n / 4, slot_base, root, s);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int ddpb_broadcast(const void* const* bufs, const void* mc, int rank, int world, int root, long long off,
                              long long n, int blocks, int slot_base, int use_mc, cudaStream_t stream) {
  CommCtx c;
  if (int rc = fill_ctx(&c, bufs, mc, rank, world)) return rc;
  if (int rc = check_range(off, n, blocks, slot_base)) return rc;
  DDPB_REQUIRE(root >= 0 && root < world, "comm: bad root %d", root);
  DDPB_REQUIRE(!use_mc || c.mc != nullptr, "comm: multicast requested but no multicast mapping");
  if (use_mc)
    broadcast_kernel<true><<<blocks, kCommThreads, 0, stream>>>(c, off, n / 4, slot_base, root);
  else
    broadcast_kernel<false><<<blocks, kCommThreads, 0, stream>>>(c, off, n / 4, slot_base, root);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int ddpb_peer_barrier(const void* const* bufs, int rank, int world, int blocks, int slot_base,
                                 cudaStream_t stream) {
  CommCtx c;
  if (int rc = fill_ctx(&c, bufs, nullptr, rank, world)) return rc;
  if (int rc = check_range(0, 0, blocks, slot_base)) return rc;
  barrier_kernel<<<blocks, kCommThreads, 0, stream>>>(c, slot_base);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---- CUDA IPC bootstrap (used when torch symmetric memory is unavailable) ----------------
// Rank-local allocation whose handle is shipped to the peers through the rendezvous store
// (the same TCPStore the reference uses for its process group, part2/part2b/main.py:56-58).
extern "C" int ddpb_ipc_alloc(long long bytes, void** ptr, void* handle64) {
  DDPB_CHECK_CUDA(cudaMalloc(ptr, static_cast<size_t>(bytes)));
  DDPB_CHECK_CUDA(cudaMemset(*ptr, 0, static_cast<size_t>(bytes)));
  cudaIpcMemHandle_t h;
  DDPB_CHECK_CUDA(cudaIpcGetMemHandle(&h, *ptr));
  static_assert(sizeof(h) == 64, "IPC handle size");
  memcpy(handle64, &h, sizeof(h));
  return 0;
}
extern "C" int ddpb_ipc_open(const void* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  DDPB_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
extern "C" int ddpb_ipc_close(void* ptr) {
  DDPB_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}
extern "C" int ddpb_ipc_free(void* ptr) {
  DDPB_CHECK_CUDA(cudaFree(ptr));
  return 0;
}
