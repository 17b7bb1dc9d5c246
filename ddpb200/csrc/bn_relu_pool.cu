// K3/K4: BatchNorm2d(track_running_stats=False) + ReLU (+ MaxPool2d(2,2))
// forward and backward as fused NHWC bf16 kernels.  Reference modules:
// /root/reference/part1/model.py:15-16 (MaxPool2d), :24 (BatchNorm2d, batch statistics in
// train AND eval), :25 (ReLU).  Three ATen ops per direction in the reference; here
//   forward : one pass   raw conv output -> normalised, rectified, (pooled) activation
//   backward: two passes (per-channel sums, then dx), recomputing the normalised value,
//             the ReLU mask and the pool argmax from the saved raw conv output instead
//             of storing masks / indices.
// Per-channel sums of the forward come from the conv epilogue (conv_fprop.cu) or from
// ddpb_channel_stats below (stand-alone variant, also the oracle in the unit tests).
#include "common.cuh"

namespace ddpb {

constexpr int kBnThreads = 256;

struct BnGeom {
  int N, H, W, C;   // pre-pool geometry
  int pool;         // 1: 2x2/2 max-pool follows the ReLU
  float eps;
  // a stride of 32 / 256 pooled output pixels as {step % Wo, step / Wo}, for the kernels that walk the pooled pixels
  // incrementally (kernel parameters: no registers)
  int d32[2], d256[2];
};

static BnGeom make_geom(int N, int H, int W, int C, int pool, float eps) {
  BnGeom g{N, H, W, C, pool, eps, {0, 0}, {0, 0}};
  const int Wo = W / 2;
  if (pool && Wo > 0) {
    g.d32[0] = 32 % Wo;
    g.d32[1] = 32 / Wo;
    g.d256[0] = 256 % Wo;
    g.d256[1] = 256 / Wo;
  }
  return g;
}

// scale/shift per channel in shared memory: y = x*scale + shift.
__device__ __forceinline__ void bn_coeffs_to_smem(const float* __restrict__ stats, const float* __restrict__ gamma,
                                                  const float* __restrict__ beta, int C, float count, float eps,
                                                  float* s_scale, float* s_shift, float* s_mean, float* s_rstd) {
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float mean = stats[c] / count;
    float var = stats[C + c] / count - mean * mean;   // biased variance, as nn.BatchNorm2d normalises
    var = fmaxf(var, 0.f);
    const float rstd = rsqrtf(var + eps);
    const float sc = gamma[c] * rstd;
    s_scale[c] = sc;
    s_shift[c] = beta[c] - mean * sc;
    if (s_mean) {
      s_mean[c] = mean;
      s_rstd[c] = rstd;
    }
  }
}

// ------------------------------------------------------------------------------------------
// stand-alone per-channel sum / sum of squares (accumulates into stats[2][C])
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBnThreads)
channel_stats_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ stats, long long P, int C) {
  extern __shared__ float s_acc[];  // [2][C]
  pdl_launch_dependents();
  pdl_wait();
  for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) s_acc[c] = 0.f;
  __syncthreads();
  const int groups = C / 8;
  const long long items = P * groups;
  float a1[8], a2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) a1[j] = a2[j] = 0.f;
  int my_group = -1;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < items; i += stride) {
    const int g = static_cast<int>(i % groups);
    if (g != my_group) {
      if (my_group >= 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          atomicAdd(&s_acc[my_group * 8 + j], a1[j]);
          atomicAdd(&s_acc[C + my_group * 8 + j], a2[j]);
          a1[j] = a2[j] = 0.f;
        }
      }
      my_group = g;
    }
    float v[8];
    load8(x + i * 8, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      a1[j] += v[j];
      a2[j] += v[j] * v[j];
    }
  }
  if (my_group >= 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      atomicAdd(&s_acc[my_group * 8 + j], a1[j]);
      atomicAdd(&s_acc[C + my_group * 8 + j], a2[j]);
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) atomicAdd(&stats[c], s_acc[c]);
}

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBnThreads)
bn_relu_pool_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ stats,
                        const float* __restrict__ gamma, const float* __restrict__ beta,
                        __nv_bfloat16* __restrict__ y, float* __restrict__ saved /* [2][C] mean, rstd */,
                        BnGeom g) {
  extern __shared__ float s_coef[];  // scale[C], shift[C], mean[C], rstd[C]
  pdl_launch_dependents();
  pdl_wait();
  float* s_scale = s_coef;
  float* s_shift = s_coef + g.C;
  float* s_mean = s_coef + 2 * g.C;
  float* s_rstd = s_coef + 3 * g.C;
  const float count = static_cast<float>(static_cast<long long>(g.N) * g.H * g.W);
  bn_coeffs_to_smem(stats, gamma, beta, g.C, count, g.eps, s_scale, s_shift, s_mean, s_rstd);
  __syncthreads();
  if (blockIdx.x == 0 && saved != nullptr) {
    for (int c = threadIdx.x; c < g.C; c += blockDim.x) {
      saved[c] = s_mean[c];
      saved[g.C + c] = s_rstd[c];
    }
  }
  const int groups = g.C / 8;
  const int Ho = g.pool ? g.H / 2 : g.H, Wo = g.pool ? g.W / 2 : g.W;
  const long long items = static_cast<long long>(g.N) * Ho * Wo * groups;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < items; i += stride) {
    const int cg = static_cast<int>(i % groups);
    const long long opix = i / groups;
    float sc[8], sh[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sc[j] = s_scale[cg * 8 + j];
      sh[j] = s_shift[cg * 8 + j];
    }
    float out[8];
    if (!g.pool) {
      float v[8];
      load8(x + opix * g.C + cg * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) out[j] = fmaxf(fmaf(v[j], sc[j], sh[j]), 0.f);
    } else {
      const int wo = static_cast<int>(opix % Wo);
      const long long t = opix / Wo;
      const int ho = static_cast<int>(t % Ho);
      const long long n = t / Ho;
      const long long base = ((n * g.H + 2 * ho) * g.W + 2 * wo);
#pragma unroll
      for (int j = 0; j < 8; ++j) out[j] = 0.f;   // ReLU floor
#pragma unroll
      for (int d = 0; d < 4; ++d) {
        const long long ipix = base + (d >> 1) * g.W + (d & 1);
        float v[8];
        load8(x + ipix * g.C + cg * 8, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) out[j] = fmaxf(out[j], fmaf(v[j], sc[j], sh[j]));
      }
    }
    store8(y + opix * g.C + cg * 8, out);
  }
}

// ------------------------------------------------------------------------------------------
// backward.  Work item = one OUTPUT position (a 2x2 window when pooling) x 8 channels, so
// every input element is loaded exactly once per pass and the pool argmax is found once
// per window (first maximum wins, like ATen; a window whose maximum is <= 0 passes nothing).
//   pass 1: sums[0][c] = sum g, sums[1][c] = sum g*xhat   (g = dy routed through pool+ReLU)
//           block 0 also clears the conv-bias gradient slot that pass 2 accumulates into.
//   pass 2: dx = gamma*rstd*(g - mean(g) - xhat*mean(g*xhat)); block 0 publishes
//           dgamma = sum g*xhat, dbeta = sum g; every block adds its share of sum(dx) to dbias.
// ------------------------------------------------------------------------------------------
struct BwdItem {
  long long ipix[4];   // input pixel index of the window positions (only [0] without pooling)
  long long opix;
  int cg;
};

__device__ __forceinline__ BwdItem bwd_item(const BnGeom& g, long long i, int groups) {
  BwdItem it;
  it.cg = static_cast<int>(i % groups);
  it.opix = i / groups;
  if (!g.pool) {
    it.ipix[0] = it.opix;
    return it;
  }
  const int Ho = g.H >> 1, Wo = g.W >> 1;
  const int wo = static_cast<int>(it.opix % Wo);
  const long long t = it.opix / Wo;
  const int ho = static_cast<int>(t % Ho);
  const long long n = t / Ho;
  const long long base = (n * g.H + 2 * ho) * g.W + 2 * wo;
  it.ipix[0] = base;
  it.ipix[1] = base + 1;
  it.ipix[2] = base + g.W;
  it.ipix[3] = base + g.W + 1;
  return it;
}

// Adds v (one value per lane; lanes l and l+groups hold the same channel) into dst[...] with
// as few shared-memory atomics as possible.
__device__ __forceinline__ void group_accumulate(float* dst, float v, int groups, int cg, int j) {
  for (int off = 16; off >= groups; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  if (groups >= 32 || static_cast<int>(lane_id()) < groups) atomicAdd(&dst[cg * 8 + j], v);
}

// element j (0..7) of a packed row of 8 bf16
__device__ __forceinline__ float bf16_at(const uint4& u, int j) {
  const uint32_t w = (j < 2) ? u.x : (j < 4) ? u.y : (j < 6) ? u.z : u.w;
  return __uint_as_float((j & 1) ? (w & 0xffff0000u) : (w << 16));
}
__device__ __forceinline__ uint4 ldg16(const __nv_bfloat16* p) { return *reinterpret_cast<const uint4*>(p); }

// Pass 1.  sums[0][c] += sum g ; sums[1][c] += sum g*x (RAW x: pass 2 turns it into
// sum g*xhat = rstd*(sum g*x - mean*sum g), which keeps mean/rstd out of this loop's registers).
template <bool kPool>
__global__ void __launch_bounds__(kBnThreads, 3)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                     const float* __restrict__ saved, const float* __restrict__ gamma,
                     const float* __restrict__ beta, float* __restrict__ sums /* [2][C] accumulated */,
                     float* __restrict__ dbias_clear /* [C] or null */, BnGeom g) {
  extern __shared__ float s_mem[];  // scale[C], shift[C], acc[2][C]
  pdl_launch_dependents();
  pdl_wait();
  float* s_scale = s_mem;
  float* s_shift = s_mem + g.C;
  float* s_acc = s_mem + 2 * g.C;
  for (int c = threadIdx.x; c < g.C; c += blockDim.x) {
    const float mean = saved[c], rstd = saved[g.C + c];
    const float sc = gamma[c] * rstd;
    s_scale[c] = sc;
    s_shift[c] = beta[c] - mean * sc;
    s_acc[c] = 0.f;
    s_acc[g.C + c] = 0.f;
    if (blockIdx.x == 0 && dbias_clear != nullptr) dbias_clear[c] = 0.f;
  }
  __syncthreads();
  const int groups = g.C / 8;
  const long long out_pix = static_cast<long long>(g.N) * (kPool ? g.H / 2 : g.H) * (kPool ? g.W / 2 : g.W);
  const long long items = out_pix * groups;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  // blockDim.x is a multiple of `groups` (host checks), so a thread keeps its channel group
  const int cg = static_cast<int>((blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) % groups);
  float sc[8], sh[8], a1[8], a2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = s_scale[cg * 8 + j];
    sh[j] = s_shift[cg * 8 + j];
    a1[j] = a2[j] = 0.f;
  }
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < items; i += stride) {
    const BwdItem it = bwd_item(g, i, groups);
    const uint4 d = ldg16(dy + it.opix * g.C + cg * 8);
    if (!kPool) {
      const uint4 v = ldg16(x + it.ipix[0] * g.C + cg * 8);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xv = bf16_at(v, j);
        const float gr = (fmaf(xv, sc[j], sh[j]) > 0.f) ? bf16_at(d, j) : 0.f;
        a1[j] += gr;
        a2[j] = fmaf(gr, xv, a2[j]);
      }
    } else {
      uint4 v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = ldg16(x + it.ipix[q] * g.C + cg * 8);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float best = 0.f, xa = 0.f;
        bool hit = false;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float xv = bf16_at(v[q], j);
          const float z = fmaf(xv, sc[j], sh[j]);
          if (z > best) {
            best = z;
            xa = xv;
            hit = true;
          }
        }
        const float gr = hit ? bf16_at(d, j) : 0.f;
        a1[j] += gr;
        a2[j] = fmaf(gr, xa, a2[j]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    group_accumulate(s_acc, a1[j], groups, cg, j);
    group_accumulate(s_acc + g.C, a2[j], groups, cg, j);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < 2 * g.C; c += blockDim.x) atomicAdd(&sums[c], s_acc[c]);
}

// Pass 2.  With A = -scale*rstd*mean(g*xhat) and Bc = -scale*mean(g) - A*mean:
//   dx = scale*g + A*x + Bc     (one fma chain per element, four coefficient vectors)
template <bool kPool>
__global__ void __launch_bounds__(kBnThreads, 2)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                    const float* __restrict__ saved, const float* __restrict__ gamma,
                    const float* __restrict__ beta, const float* __restrict__ sums,
                    __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta,
                    float* __restrict__ dbias, BnGeom g) {
  extern __shared__ float s_mem[];  // scale, shift, A, Bc, acc : 5*C
  pdl_launch_dependents();
  pdl_wait();
  float* s_scale = s_mem;
  float* s_shift = s_mem + g.C;
  float* s_A = s_mem + 2 * g.C;
  float* s_B = s_mem + 3 * g.C;
  float* s_acc = s_mem + 4 * g.C;
  const float inv_count = 1.f / static_cast<float>(static_cast<long long>(g.N) * g.H * g.W);
  for (int c = threadIdx.x; c < g.C; c += blockDim.x) {
    const float mean = saved[c], rstd = saved[g.C + c];
    const float sc = gamma[c] * rstd;
    const float sum_g = sums[c];
    const float sum_gxhat = rstd * (sums[g.C + c] - mean * sum_g);
    const float A = -sc * rstd * sum_gxhat * inv_count;
    s_scale[c] = sc;
    s_shift[c] = beta[c] - mean * sc;
    s_A[c] = A;
    s_B[c] = -sc * sum_g * inv_count - A * mean;
    s_acc[c] = 0.f;
    if (blockIdx.x == 0) {
      if (dbeta) dbeta[c] = sum_g;
      if (dgamma) dgamma[c] = sum_gxhat;
    }
  }
  __syncthreads();
  const int groups = g.C / 8;
  const long long out_pix = static_cast<long long>(g.N) * (kPool ? g.H / 2 : g.H) * (kPool ? g.W / 2 : g.W);
  const long long items = out_pix * groups;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const int cg = static_cast<int>((blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) % groups);
  float sc[8], sh[8], cA[8], cB[8], ab[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = cg * 8 + j;
    sc[j] = s_scale[c];
    sh[j] = s_shift[c];
    cA[j] = s_A[c];
    cB[j] = s_B[c];
    ab[j] = 0.f;
  }
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < items; i += stride) {
    const BwdItem it = bwd_item(g, i, groups);
    const uint4 d = ldg16(dy + it.opix * g.C + cg * 8);
    if (!kPool) {
      const uint4 v = ldg16(x + it.ipix[0] * g.C + cg * 8);
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xv = bf16_at(v, j);
        const float gr = (fmaf(xv, sc[j], sh[j]) > 0.f) ? bf16_at(d, j) : 0.f;
        o[j] = fmaf(sc[j], gr, fmaf(cA[j], xv, cB[j]));
        ab[j] += o[j];
      }
      store8(dx + it.ipix[0] * g.C + cg * 8, o);
    } else {
      uint4 v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = ldg16(x + it.ipix[q] * g.C + cg * 8);
      int arg[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float best = 0.f;
        arg[j] = -1;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float z = fmaf(bf16_at(v[q], j), sc[j], sh[j]);
          if (z > best) {
            best = z;
            arg[j] = q;
          }
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float gr = (arg[j] == q) ? bf16_at(d, j) : 0.f;
          o[j] = fmaf(sc[j], gr, fmaf(cA[j], bf16_at(v[q], j), cB[j]));
          ab[j] += o[j];
        }
        store8(dx + it.ipix[q] * g.C + cg * 8, o);
      }
    }
  }
  if (dbias != nullptr) {
#pragma unroll
    for (int j = 0; j < 8; ++j) group_accumulate(s_acc, ab[j], groups, cg, j);
    __syncthreads();
    for (int c = threadIdx.x; c < g.C; c += blockDim.x) atomicAdd(&dbias[c], s_acc[c]);
  }
}


// ------------------------------------------------------------------------------------------
// Fused backward (one launch, one read of x and dy from memory).  Every layer but the first
// is small enough that each of ~148 CTAs can keep ITS slice of (x, dy) in shared memory:
// phase 1 streams the slice in, accumulates the per-channel sums and parks the data in smem;
// a grid-wide barrier waits for the global sums; phase 2 produces dx from smem.  Launched
// cooperatively (all CTAs co-resident); the barrier counter is one word the caller zeroes.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void grid_barrier_once(unsigned int* counter, unsigned int expected) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(counter, 1u);
    const unsigned long long t0 = global_timer_ns();
    unsigned int seen, spins = 0;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(counter) : "memory");
      if ((++spins & 0x3FFu) == 0 && global_timer_ns() - t0 > DDPB_WAIT_TIMEOUT_NS) {
        printf("ddpb200: bn_bwd_fused grid barrier timed out (block %d: %u of %u)\n", blockIdx.x, seen, expected);
        __trap();
      }
    } while (seen < expected);
    __threadfence();
  }
  __syncthreads();
}

// 256 threads x <=128 registers: small enough to share an SM with a gradient-sync CTA (512
// threads, ~40 registers).  A cooperative launch needs ALL its CTAs resident at once, so a
// fatter CTA would make every fused BatchNorm backward wait for the overlapped reductions of
// mode 'ddp' to drain (measured: per-layer buckets got slower; numbers in DESIGN.md par. 7).
constexpr int kBnFusedThreads = 256;

constexpr int kSliceGroups = 8;   // a CTA owns 8 channel groups = 64 channels ...

// ... and a range of output pixels.  Slicing the channels keeps the per-CTA fixed work
// (coefficients, atomics: 2 x 64 instead of 2 x C) proportional to what the CTA processes;
// with whole-C CTAs the 512-channel layers spent most of their time in that overhead.
template <bool kPool>
__global__ void __launch_bounds__(kBnFusedThreads, 2)
bn_bwd_fused_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                    const float* __restrict__ saved, const float* __restrict__ gamma,
                    const float* __restrict__ beta, float* __restrict__ sums /* [2][C] zero, then counter */,
                    __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta,
                    float* __restrict__ dbias, BnGeom g, int pix_per_cta, int n_ranges) {
  constexpr int kSlot = kPool ? 5 : 2;          // uint4 words parked per item: window of x + dy
  constexpr int SC = kSliceGroups * 8;          // channels per CTA
  __shared__ float s_scale[SC], s_shift[SC], s_A[SC], s_B[SC], s_acc[2 * SC];
  extern __shared__ uint4 s_data[];
  pdl_launch_dependents();
  pdl_wait();
  unsigned int* counter = reinterpret_cast<unsigned int*>(sums + 2 * g.C);
  const int groups = g.C / 8;
  const int slice = blockIdx.x / n_ranges, range = blockIdx.x % n_ranges;
  const int c0 = slice * SC;
  const long long out_pix = static_cast<long long>(g.N) * (kPool ? g.H / 2 : g.H) * (kPool ? g.W / 2 : g.W);
  const long long p_begin = static_cast<long long>(range) * pix_per_cta;
  long long p_end = p_begin + pix_per_cta;
  if (p_end > out_pix) p_end = out_pix;
  const int n_local = p_end > p_begin ? static_cast<int>(p_end - p_begin) * kSliceGroups : 0;
  const int lg = threadIdx.x % kSliceGroups;            // my channel group inside the slice (fixed)
  const int cg = slice * kSliceGroups + lg;             // ... and in the tensor
  // ---- phase 1a: issue EVERY global load of this CTA's slice (x window + dy) and park the data
  // in shared memory.  Nothing here depends on the coefficients, so these loads are in flight
  // together with the coefficient loads below: the kernel is a chain of memory round trips
  // (ncu: 8% of DRAM bandwidth, time in scoreboard waits), and this removes one of them while
  // putting all of a thread's loads in flight at once.
  for (int i = threadIdx.x; i < n_local; i += blockDim.x) {
    const long long opix = p_begin + i / kSliceGroups;
    const BwdItem it = bwd_item(g, opix * groups + cg, groups);
    uint4* slot = s_data + static_cast<long long>(i) * kSlot;
    slot[kSlot - 1] = ldg16(dy + it.opix * g.C + cg * 8);
#pragma unroll
    for (int q = 0; q < kSlot - 1; ++q) slot[q] = ldg16(x + it.ipix[q] * g.C + cg * 8);
  }
  for (int c = threadIdx.x; c < SC; c += blockDim.x) {
    const float mean = saved[c0 + c], rstd = saved[g.C + c0 + c];
    const float sc = gamma[c0 + c] * rstd;
    s_scale[c] = sc;
    s_shift[c] = beta[c0 + c] - mean * sc;
    s_acc[c] = 0.f;
    s_acc[SC + c] = 0.f;
    if (range == 0 && dbias != nullptr) dbias[c0 + c] = 0.f;
  }
  __syncthreads();
  float sc[8], sh[8], a1[8], a2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = s_scale[lg * 8 + j];
    sh[j] = s_shift[lg * 8 + j];
    a1[j] = a2[j] = 0.f;
  }
  // ---- phase 1b: per-channel sums from the parked data (each thread re-reads its own slots)
  for (int i = threadIdx.x; i < n_local; i += blockDim.x) {
    const uint4* slot = s_data + static_cast<long long>(i) * kSlot;
    const uint4 d = slot[kSlot - 1];
    if (!kPool) {
      const uint4 v = slot[0];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xv = bf16_at(v, j);
        const float gr = (fmaf(xv, sc[j], sh[j]) > 0.f) ? bf16_at(d, j) : 0.f;
        a1[j] += gr;
        a2[j] = fmaf(gr, xv, a2[j]);
      }
    } else {
      uint4 v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = slot[q];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float best = 0.f, xa = 0.f;
        bool hit = false;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float xv = bf16_at(v[q], j);
          const float z = fmaf(xv, sc[j], sh[j]);
          if (z > best) {
            best = z;
            xa = xv;
            hit = true;
          }
        }
        const float gr = hit ? bf16_at(d, j) : 0.f;
        a1[j] += gr;
        a2[j] = fmaf(gr, xa, a2[j]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    group_accumulate(s_acc, a1[j], kSliceGroups, lg, j);
    group_accumulate(s_acc + SC, a2[j], kSliceGroups, lg, j);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < 2 * SC; c += blockDim.x)
    atomicAdd(&sums[(c / SC) * g.C + c0 + (c % SC)], s_acc[c]);
  grid_barrier_once(counter, gridDim.x);
  // ---- phase 2: coefficients from the global sums, dx from the parked slice
  const float inv_count = 1.f / static_cast<float>(static_cast<long long>(g.N) * g.H * g.W);
  for (int c = threadIdx.x; c < SC; c += blockDim.x) {
    const float mean = saved[c0 + c], rstd = saved[g.C + c0 + c];
    const float sum_g = __ldcg(sums + c0 + c);
    const float sum_gxhat = rstd * (__ldcg(sums + g.C + c0 + c) - mean * sum_g);
    const float A = -s_scale[c] * rstd * sum_gxhat * inv_count;
    s_A[c] = A;
    s_B[c] = -s_scale[c] * sum_g * inv_count - A * mean;
    s_acc[c] = 0.f;
    if (range == 0) {
      if (dbeta) dbeta[c0 + c] = sum_g;
      if (dgamma) dgamma[c0 + c] = sum_gxhat;
    }
  }
  __syncthreads();
  float cA[8], cB[8], ab[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    cA[j] = s_A[lg * 8 + j];
    cB[j] = s_B[lg * 8 + j];
    ab[j] = 0.f;
  }
  for (int i = threadIdx.x; i < n_local; i += blockDim.x) {
    const long long opix = p_begin + i / kSliceGroups;
    const BwdItem it = bwd_item(g, opix * groups + cg, groups);
    const uint4* slot = s_data + static_cast<long long>(i) * kSlot;
    const uint4 d = slot[kSlot - 1];
    if (!kPool) {
      const uint4 v = slot[0];
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xv = bf16_at(v, j);
        const float gr = (fmaf(xv, sc[j], sh[j]) > 0.f) ? bf16_at(d, j) : 0.f;
        o[j] = fmaf(sc[j], gr, fmaf(cA[j], xv, cB[j]));
        ab[j] += o[j];
      }
      store8(dx + it.ipix[0] * g.C + cg * 8, o);
    } else {
      uint4 v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = slot[q];
      int arg[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float best = 0.f;
        arg[j] = -1;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float z = fmaf(bf16_at(v[q], j), sc[j], sh[j]);
          if (z > best) {
            best = z;
            arg[j] = q;
          }
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float gr = (arg[j] == q) ? bf16_at(d, j) : 0.f;
          o[j] = fmaf(sc[j], gr, fmaf(cA[j], bf16_at(v[q], j), cB[j]));
          ab[j] += o[j];
        }
        store8(dx + it.ipix[q] * g.C + cg * 8, o);
      }
    }
  }
  if (dbias != nullptr) {
#pragma unroll
    for (int j = 0; j < 8; ++j) group_accumulate(s_acc, ab[j], kSliceGroups, lg, j);
    __syncthreads();
    for (int c = threadIdx.x; c < SC; c += blockDim.x) atomicAdd(&dbias[c0 + c], s_acc[c]);
  }
}

// ------------------------------------------------------------------------------------------
// Deterministic two-kernel backward (default, mode 3).  The cooperative kernel above needs every one
// of its CTAs resident at once, so inside the training step it could not start while CTAs of the
// weight-gradient branch (or of an overlapped gradient reduction / SGD update) still held SMs: the
// branches of the backward graph serialised on it (profiles/r1_call40_*, r2_call04_*).  These two
// kernels are ordinary launches with a small footprint (256 threads, 3 KB of shared memory) that
// co-run with tensor-core CTAs, use no float atomics -- every sum is taken in a fixed order, so
// dgamma / dbeta are bitwise reproducible -- and keep several 16-byte loads per thread in flight.
//
//   reduce: CTA (slice of 64 channels, pixel range) -> partial (sum g, sum g*x) per channel into
//           partials[slice][range][2][64]; the LAST CTA of a slice to finish (ticket counter) adds the
//           slice's partials in range order into sums[2][C] and resets the counter for the next launch.
//   apply : coefficients from sums, dx = scale*g + A*x + Bc streamed out; CTA (slice, 0) publishes
//           dgamma, dbeta and the conv-bias gradient.  The latter is sum(dx) = scale*sum(g) +
//           A*sum(x) + Bc*P evaluated from the sums: analytically zero (BatchNorm removes the channel
//           mean; SURVEY.md par. 2.5), numerically fp32 rounding noise like the reference's value.
// ------------------------------------------------------------------------------------------
// Pooling window of 4 x (8 channels, packed bf16): the element that wins the max-pool AFTER BatchNorm is the
// largest x where gamma * rstd > 0 and the smallest where it is < 0, so it can be found on the raw bf16 values
// with packed max instructions (sign bit flipped for the channels with a negative scale) instead of four
// fp32 fma + compare + select chains per channel.  flip: 0x8000 in the halves whose scale is negative.
__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
  const __nv_bfloat162 r = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
  return *reinterpret_cast<const uint32_t*>(&r);
}
__device__ __forceinline__ uint4 window_best(const uint4 (&v)[4], const uint4& flip) {
  uint4 r;
  r.x = bf16x2_max(bf16x2_max(v[0].x ^ flip.x, v[1].x ^ flip.x), bf16x2_max(v[2].x ^ flip.x, v[3].x ^ flip.x)) ^ flip.x;
  r.y = bf16x2_max(bf16x2_max(v[0].y ^ flip.y, v[1].y ^ flip.y), bf16x2_max(v[2].y ^ flip.y, v[3].y ^ flip.y)) ^ flip.y;
  r.z = bf16x2_max(bf16x2_max(v[0].z ^ flip.z, v[1].z ^ flip.z), bf16x2_max(v[2].z ^ flip.z, v[3].z ^ flip.z)) ^ flip.z;
  r.w = bf16x2_max(bf16x2_max(v[0].w ^ flip.w, v[1].w ^ flip.w), bf16x2_max(v[2].w ^ flip.w, v[3].w ^ flip.w)) ^ flip.w;
  return r;
}
__device__ __forceinline__ uint4 sign_flip_mask(const float (&sc)[8]) {
  uint4 f;
  f.x = (sc[0] < 0.f ? 0x8000u : 0u) | (sc[1] < 0.f ? 0x80000000u : 0u);
  f.y = (sc[2] < 0.f ? 0x8000u : 0u) | (sc[3] < 0.f ? 0x80000000u : 0u);
  f.z = (sc[4] < 0.f ? 0x8000u : 0u) | (sc[5] < 0.f ? 0x80000000u : 0u);
  f.w = (sc[6] < 0.f ? 0x8000u : 0u) | (sc[7] < 0.f ? 0x80000000u : 0u);
  return f;
}

// CTAs per SM the two kernels are compiled for / resident waves of the launch (measured: 3 / 2 / 2 is best,
// more CTAs per SM spill and run 1.5-1.8x slower: call 18, numbers in DESIGN.md par. 7)
#ifndef DDPB_BN2_REDUCE_MINB
#define DDPB_BN2_REDUCE_MINB 3
#endif
#ifndef DDPB_BN2_APPLY_MINB
#define DDPB_BN2_APPLY_MINB 2
#endif
#ifndef DDPB_BN2_WAVES
#define DDPB_BN2_WAVES 2
#endif
constexpr int kBn2Threads = 256;
constexpr int kBn2Rows = kBn2Threads / kSliceGroups;   // 32 output pixels per CTA pass

struct Bn2Plan {
  int n_slices, n_ranges, pix_per_cta;
};

static Bn2Plan bn2_plan(long long out_pix, int C) {
  Bn2Plan pl;
  pl.n_slices = C / (kSliceGroups * 8);
  // one resident wave (two CTAs per SM over all slices), at least two 32-pixel passes per CTA
  long long ranges = (static_cast<long long>(DDPB_BN2_WAVES) * sm_count()) / pl.n_slices;
  const long long max_ranges = (out_pix + 2 * kBn2Rows - 1) / (2 * kBn2Rows);
  if (ranges > max_ranges) ranges = max_ranges;
  if (ranges < 1) ranges = 1;
  long long ppc = (out_pix + ranges - 1) / ranges;
  ppc = (ppc + kBn2Rows - 1) / kBn2Rows * kBn2Rows;
  pl.pix_per_cta = static_cast<int>(ppc);
  pl.n_ranges = static_cast<int>((out_pix + ppc - 1) / ppc);
  return pl;
}

// Loads one work item: dy of output pixel `opix` and the matching x window (1 pixel, or the 2x2 pooling
// window whose first pixel index is returned).  32-bit pixel arithmetic (the host checks N*H*W*C < 2^31):
// the 64-bit divisions of the older kernels cost more instructions than the rest of the item.
// Position of a pooled output pixel, kept incrementally: a thread's items are kBn2Rows output pixels apart, so
// (wo, ho, n) advance by constant amounts with two carries instead of three integer divisions per item (the
// divisions were ~40% of the instructions of the pooled kernels, and the compute phase between two rounds of
// loads, not the memory system, is what bounds them).
struct Bn2Pos {
  int wo, row;          // current output pixel: column, and row counted over all images (n * Ho + ho)
};
__device__ __forceinline__ Bn2Pos bn2_pos(const BnGeom& g, int opix) {
  const int Wo = g.W >> 1;
  Bn2Pos p;
  p.wo = opix % Wo;
  p.row = opix / Wo;
  return p;
}
// step: g.d32 or g.d256 (the thread's stride in pooled pixels as {step % Wo, step / Wo}, from the host)
__device__ __forceinline__ void bn2_advance(Bn2Pos& p, const BnGeom& g, const int (&step)[2]) {
  const int Wo = g.W >> 1;
  p.wo += step[0];
  const int carry = p.wo >= Wo ? 1 : 0;
  p.wo -= carry ? Wo : 0;
  p.row += step[1] + carry;
}

template <bool kPool>
__device__ __forceinline__ int bn2_load(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                        const BnGeom& g, int opix, const Bn2Pos& pos, int cg,
                                        uint4 (&v)[kPool ? 4 : 1], uint4& d) {
  int base = opix;
  if (kPool) base = 2 * (pos.row * g.W + pos.wo);      // H is even: input row = 2 * (n * Ho + ho)
  d = ldg16(dy + static_cast<size_t>(opix) * g.C + cg * 8);
#pragma unroll
  for (int q = 0; q < (kPool ? 4 : 1); ++q)
    v[q] = ldg16(x + static_cast<size_t>(base + (q >> 1) * g.W + (q & 1)) * g.C + cg * 8);
  return base;
}

template <bool kPool>
__global__ void __launch_bounds__(kBn2Threads, DDPB_BN2_REDUCE_MINB)
bn_bwd2_reduce_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                      const float* __restrict__ saved, const float* __restrict__ gamma,
                      const float* __restrict__ beta, float* __restrict__ sums, float* __restrict__ partials,
                      BnGeom g, int pix_per_cta, int n_ranges) {
  constexpr int SC = kSliceGroups * 8;
  constexpr int kWin = kPool ? 4 : 1;
  constexpr int kU = kPool ? 2 : 4;      // items in flight per thread
  __shared__ float s_part[kBn2Threads / 32][2 * SC];
  __shared__ unsigned int s_last;
  pdl_launch_dependents();
  const int slice = blockIdx.x / n_ranges, range = blockIdx.x % n_ranges;
  const int lg = threadIdx.x % kSliceGroups, prow = threadIdx.x / kSliceGroups;
  const int cg = slice * kSliceGroups + lg;
  const int c0 = slice * SC;
  const int out_pix = g.N * (kPool ? g.H / 2 : g.H) * (kPool ? g.W / 2 : g.W);
  const int p_begin = range * pix_per_cta;
  int p_end = p_begin + pix_per_cta;
  if (p_end > out_pix) p_end = out_pix;
  pdl_wait();
  float sc[8], sh[8], a1[8], a2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = cg * 8 + j;
    const float mean = __ldg(saved + c), rstd = __ldg(saved + g.C + c);
    sc[j] = __ldg(gamma + c) * rstd;
    sh[j] = __ldg(beta + c) - mean * sc[j];
    a1[j] = a2[j] = 0.f;
  }
  const uint4 flip = sign_flip_mask(sc);
  (void)flip;
  Bn2Pos pos = {};
  if (kPool) pos = bn2_pos(g, p_begin + prow);
  for (int base = p_begin + prow; base < p_end; base += kBn2Rows * kU) {
    uint4 v[kU][kWin], d[kU];
    int ipix[kU];
    bool ok[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int opix = base + u * kBn2Rows;
      ok[u] = opix < p_end;
      if (ok[u]) ipix[u] = bn2_load<kPool>(x, dy, g, opix, pos, cg, v[u], d[u]);
      if (kPool) bn2_advance(pos, g, g.d32);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      if (!ok[u]) continue;
      uint4 xb;                       // the window element that survives pool + ReLU, 8 channels packed
      if constexpr (kPool) xb = window_best(v[u], flip); else xb = v[u][0];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xa = bf16_at(xb, j);
        const float gr = (fmaf(xa, sc[j], sh[j]) > 0.f) ? bf16_at(d[u], j) : 0.f;
        a1[j] += gr;
        a2[j] = fmaf(gr, xa, a2[j]);
      }
    }
  }
  // fixed-order tree: lanes l, l^8, l^16 (same channel group), then the 8 warps, then the ranges
  const int warp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    a1[j] += __shfl_xor_sync(0xffffffffu, a1[j], 8);
    a2[j] += __shfl_xor_sync(0xffffffffu, a2[j], 8);
    a1[j] += __shfl_xor_sync(0xffffffffu, a1[j], 16);
    a2[j] += __shfl_xor_sync(0xffffffffu, a2[j], 16);
  }
  if (lane_id() < static_cast<uint32_t>(kSliceGroups)) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s_part[warp][lg * 8 + j] = a1[j];
      s_part[warp][SC + lg * 8 + j] = a2[j];
    }
  }
  __syncthreads();
  float* my_partial = partials + (static_cast<long long>(slice) * n_ranges + range) * (2 * SC);
  if (threadIdx.x < 2 * SC) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kBn2Threads / 32; ++w) t += s_part[w][threadIdx.x];
    __stcg(my_partial + threadIdx.x, t);
  }
  // ticket: the last CTA of this slice to get here folds the slice's partials (in range order)
  unsigned int* counter = reinterpret_cast<unsigned int*>(sums + 2 * g.C) + slice;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(counter, 1u) == static_cast<unsigned int>(n_ranges - 1)) ? 1u : 0u;
  __syncthreads();
  if (s_last == 0u) return;
  __threadfence();
  // 256 threads: thread t < 128 owns value t over the even ranges, t >= 128 value t-128 over the odd ones
  {
    const int val = threadIdx.x % (2 * SC), par = threadIdx.x / (2 * SC);
    const float* src = partials + static_cast<long long>(slice) * n_ranges * (2 * SC) + val;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    int r = par;
    for (; r + 14 < n_ranges; r += 16) {
      float t[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) t[k] = __ldcg(src + static_cast<long long>(r + 2 * k) * (2 * SC));
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += t[k];
    }
    for (; r < n_ranges; r += 2) acc[0] += __ldcg(src + static_cast<long long>(r) * (2 * SC));
    acc[0] += acc[4]; acc[1] += acc[5]; acc[2] += acc[6]; acc[3] += acc[7];
    s_part[par][val] = (acc[0] + acc[1]) + (acc[2] + acc[3]);
  }
  __syncthreads();
  if (threadIdx.x < 2 * SC) {
    const int which = threadIdx.x / SC, c = threadIdx.x % SC;
    sums[which * g.C + c0 + c] = s_part[0][threadIdx.x] + s_part[1][threadIdx.x];
  }
  if (threadIdx.x == 0) *counter = 0u;   // ready for the next launch (graph replay)
}

template <bool kPool>
__global__ void __launch_bounds__(kBn2Threads, DDPB_BN2_APPLY_MINB)
bn_bwd2_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                     const float* __restrict__ saved, const float* __restrict__ gamma,
                     const float* __restrict__ beta, const float* __restrict__ sums, __nv_bfloat16* __restrict__ dx,
                     float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dbias, BnGeom g,
                     int pix_per_cta, int n_ranges) {
  constexpr int kWin = kPool ? 4 : 1;
  constexpr int kU = kPool ? 2 : 4;
  pdl_launch_dependents();
  const int slice = blockIdx.x / n_ranges, range = blockIdx.x % n_ranges;
  const int lg = threadIdx.x % kSliceGroups, prow = threadIdx.x / kSliceGroups;
  const int cg = slice * kSliceGroups + lg;
  const int out_pix = g.N * (kPool ? g.H / 2 : g.H) * (kPool ? g.W / 2 : g.W);
  const int p_begin = range * pix_per_cta;
  int p_end = p_begin + pix_per_cta;
  if (p_end > out_pix) p_end = out_pix;
  const float count = static_cast<float>(static_cast<long long>(g.N) * g.H * g.W);
  const float inv_count = 1.f / count;
  pdl_wait();
  float sc[8], sh[8], cA[8], cB[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = cg * 8 + j;
    const float mean = __ldg(saved + c), rstd = __ldg(saved + g.C + c);
    sc[j] = __ldg(gamma + c) * rstd;
    sh[j] = __ldg(beta + c) - mean * sc[j];
    const float sum_g = __ldcg(sums + c);
    const float sum_gxhat = rstd * (__ldcg(sums + g.C + c) - mean * sum_g);
    cA[j] = -sc[j] * rstd * sum_gxhat * inv_count;
    cB[j] = -sc[j] * sum_g * inv_count - cA[j] * mean;
    if (range == 0 && prow == 0) {
      if (dbeta != nullptr) dbeta[c] = sum_g;
      if (dgamma != nullptr) dgamma[c] = sum_gxhat;
      // sum over all pixels of dx, from the sums (sum x = mean * count)
      if (dbias != nullptr) dbias[c] = fmaf(sc[j], sum_g, fmaf(cA[j], mean * count, cB[j] * count));
    }
  }
  const uint4 flip = sign_flip_mask(sc);
  (void)flip;
  Bn2Pos pos = {};
  if (kPool) pos = bn2_pos(g, p_begin + prow);
  for (int base = p_begin + prow; base < p_end; base += kBn2Rows * kU) {
    uint4 v[kU][kWin], d[kU];
    int ipix[kU];
    bool ok[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int opix = base + u * kBn2Rows;
      ok[u] = opix < p_end;
      if (ok[u]) ipix[u] = bn2_load<kPool>(x, dy, g, opix, pos, cg, v[u], d[u]);
      if (kPool) bn2_advance(pos, g, g.d32);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      if (!ok[u]) continue;
      if constexpr (!kPool) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xv = bf16_at(v[u][0], j);
          const float gr = (fmaf(xv, sc[j], sh[j]) > 0.f) ? bf16_at(d[u], j) : 0.f;
          o[j] = fmaf(sc[j], gr, fmaf(cA[j], xv, cB[j]));
        }
        store8(dx + static_cast<size_t>(ipix[u]) * g.C + cg * 8, o);
      } else {
        // route dy to the window element that won the max-pool: packed compare against the winner, first match wins
        const uint4 xb = window_best(v[u], flip);
        float gsel[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) gsel[j] = (fmaf(bf16_at(xb, j), sc[j], sh[j]) > 0.f) ? bf16_at(d[u], j) : 0.f;
        uint32_t found[4] = {0u, 0u, 0u, 0u};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t vw[4] = {v[u][q].x, v[u][q].y, v[u][q].z, v[u][q].w};
          const uint32_t bw[4] = {xb.x, xb.y, xb.z, xb.w};
          uint32_t m[4];
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            const uint32_t eq = __heq2_mask(*reinterpret_cast<const __nv_bfloat162*>(&vw[w]),
                                            *reinterpret_cast<const __nv_bfloat162*>(&bw[w]));
            m[w] = eq & ~found[w];
            found[w] |= eq;
          }
          float o[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const bool sel = ((m[j >> 1] >> ((j & 1) * 16)) & 1u) != 0u;
            o[j] = fmaf(sc[j], sel ? gsel[j] : 0.f, fmaf(cA[j], bf16_at(v[u][q], j), cB[j]));
          }
          store8(dx + static_cast<size_t>(ipix[u] + (q >> 1) * g.W + (q & 1)) * g.C + cg * 8, o);
        }
      }
    }
  }
}

template <bool kPool>
static int launch_bn_bwd2(const __nv_bfloat16* x, const __nv_bfloat16* dy, const float* saved, const float* gamma,
                          const float* beta, float* sums, float* partials, long long partial_elems,
                          __nv_bfloat16* dx, float* dgamma, float* dbeta, float* dbias, BnGeom g, long long out_pix,
                          cudaStream_t stream) {
  const Bn2Plan pl = bn2_plan(out_pix, g.C);
  const long long need = static_cast<long long>(pl.n_slices) * pl.n_ranges * 2 * kSliceGroups * 8;
  DDPB_REQUIRE(partials != nullptr && partial_elems >= need,
               "bn_relu_pool_bwd: the partial-sum buffer holds %lld floats, this shape needs %lld "
               "(ddpb_bn_bwd_partials_elems)", partial_elems, need);
  DDPB_REQUIRE(pl.n_slices <= 32, "bn_relu_pool_bwd: at most 2048 channels");
  DDPB_REQUIRE(static_cast<long long>(g.N) * g.H * g.W * g.C < (1LL << 31), "bn_relu_pool_bwd: tensor too large");
  const int grid = pl.n_slices * pl.n_ranges;
  const float* sums_c = sums;
  DDPB_CHECK_CUDA(launch_pdl(bn_bwd2_reduce_kernel<kPool>, grid, kBn2Threads, 0, stream, x, dy, saved, gamma, beta,
                             sums, partials, g, pl.pix_per_cta, pl.n_ranges));
  DDPB_CHECK_CUDA(launch_pdl(bn_bwd2_apply_kernel<kPool>, grid, kBn2Threads, 0, stream, x, dy, saved, gamma, beta,
                             sums_c, dx, dgamma, dbeta, dbias, g, pl.pix_per_cta, pl.n_ranges));
  return 0;
}

// ------------------------------------------------------------------------------------------
// Channel-owner backward for the small late layers (default path when the layer fits): ONE ordinary
// launch, no grid-wide synchronisation at all.  A CTA owns 8 channels and ALL pixels of the layer, so
// the per-channel sums are complete inside the CTA: it streams its (x window, dy) items into shared
// memory with every load in flight at once, reduces in a fixed order (registers -> warp shuffles ->
// the 8 warps), and produces dx from the parked data.  Used when a CTA's slice is at most 48 KB (batch 256:
// layers 6 and 7; smaller batches: more layers), where two latency-bound launches become one.
// Deterministic like the two-kernel path.
// ------------------------------------------------------------------------------------------
constexpr int kBnOwnThreads = 256;

template <bool kPool>
__global__ void __launch_bounds__(kBnOwnThreads, 1)
bn_bwd_owner_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                    const float* __restrict__ saved, const float* __restrict__ gamma,
                    const float* __restrict__ beta, __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma,
                    float* __restrict__ dbeta, float* __restrict__ dbias, BnGeom g) {
  constexpr int kSlot = kPool ? 5 : 2;   // uint4 words parked per item: x window + dy
  constexpr int kWin = kPool ? 4 : 1;
  extern __shared__ uint4 s_data[];
  __shared__ float s_coef[4][8];                        // scale, shift, A, Bc
  __shared__ float s_red[kBnOwnThreads / 32][16];
  __shared__ float s_tot[16];
  pdl_launch_dependents();
  const int cg = blockIdx.x;                            // this CTA's 8 channels
  const int out_pix = g.N * (kPool ? g.H / 2 : g.H) * (kPool ? g.W / 2 : g.W);
  pdl_wait();
  // ---- phase 1a: every global load of the CTA, straight into shared memory
  Bn2Pos pos = {};
  static_assert(kBnOwnThreads == 256 && kBn2Rows == 32, "BnGeom carries the strides 32 and 256");
  if (kPool) pos = bn2_pos(g, threadIdx.x);
  for (int i = threadIdx.x; i < out_pix; i += kBnOwnThreads) {
    uint4 v[kWin], d;
    bn2_load<kPool>(x, dy, g, i, pos, cg, v, d);
    if (kPool) bn2_advance(pos, g, g.d256);
    uint4* slot = s_data + static_cast<size_t>(i) * kSlot;
#pragma unroll
    for (int q = 0; q < kWin; ++q) slot[q] = v[q];
    slot[kSlot - 1] = d;
  }
  if (threadIdx.x < 8) {
    const int c = cg * 8 + threadIdx.x;
    const float mean = saved[c], rstd = saved[g.C + c];
    const float sc = gamma[c] * rstd;
    s_coef[0][threadIdx.x] = sc;
    s_coef[1][threadIdx.x] = beta[c] - mean * sc;
  }
  __syncthreads();
  float sc[8], sh[8], a1[8], a2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = s_coef[0][j];
    sh[j] = s_coef[1][j];
    a1[j] = a2[j] = 0.f;
  }
  // ---- phase 1b: per-channel sums from this thread's own slots
  for (int i = threadIdx.x; i < out_pix; i += kBnOwnThreads) {
    const uint4* slot = s_data + static_cast<size_t>(i) * kSlot;
    const uint4 d = slot[kSlot - 1];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float gr, xa;
      if (!kPool) {
        xa = bf16_at(slot[0], j);
        gr = (fmaf(xa, sc[j], sh[j]) > 0.f) ? bf16_at(d, j) : 0.f;
      } else {
        float best = 0.f;
        xa = 0.f;
        bool hit = false;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float xv = bf16_at(slot[q], j);
          const float z = fmaf(xv, sc[j], sh[j]);
          if (z > best) {
            best = z;
            xa = xv;
            hit = true;
          }
        }
        gr = hit ? bf16_at(d, j) : 0.f;
      }
      a1[j] += gr;
      a2[j] = fmaf(gr, xa, a2[j]);
    }
  }
  // fixed-order block reduction
#pragma unroll
  for (int j = 0; j < 8; ++j) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
      a1[j] += __shfl_xor_sync(0xffffffffu, a1[j], off);
      a2[j] += __shfl_xor_sync(0xffffffffu, a2[j], off);
    }
  }
  if (lane_id() == 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s_red[threadIdx.x >> 5][j] = a1[j];
      s_red[threadIdx.x >> 5][8 + j] = a2[j];
    }
  }
  __syncthreads();
  if (threadIdx.x < 16) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kBnOwnThreads / 32; ++w) t += s_red[w][threadIdx.x];
    s_tot[threadIdx.x] = t;
  }
  __syncthreads();
  if (threadIdx.x < 8) {
    const int c = cg * 8 + threadIdx.x;
    const float count = static_cast<float>(static_cast<long long>(g.N) * g.H * g.W);
    const float mean = saved[c], rstd = saved[g.C + c];
    const float scale = s_coef[0][threadIdx.x];
    const float sum_g = s_tot[threadIdx.x];
    const float sum_gxhat = rstd * (s_tot[8 + threadIdx.x] - mean * sum_g);
    const float A = -scale * rstd * sum_gxhat / count;
    const float Bc = -scale * sum_g / count - A * mean;
    s_coef[2][threadIdx.x] = A;
    s_coef[3][threadIdx.x] = Bc;
    if (dbeta != nullptr) dbeta[c] = sum_g;
    if (dgamma != nullptr) dgamma[c] = sum_gxhat;
    if (dbias != nullptr) dbias[c] = fmaf(scale, sum_g, fmaf(A, mean * count, Bc * count));
  }
  __syncthreads();
  float cA[8], cB[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    cA[j] = s_coef[2][j];
    cB[j] = s_coef[3][j];
  }
  // ---- phase 2: dx from the parked data
  for (int i = threadIdx.x; i < out_pix; i += kBnOwnThreads) {
    const uint4* slot = s_data + static_cast<size_t>(i) * kSlot;
    const uint4 d = slot[kSlot - 1];
    int base = i;
    if (kPool) {
      const int Ho = g.H >> 1, Wo = g.W >> 1;
      const int wo = i % Wo, t = i / Wo;
      base = ((t / Ho) * g.H + 2 * (t % Ho)) * g.W + 2 * wo;
    }
    if (!kPool) {
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xv = bf16_at(slot[0], j);
        const float gr = (fmaf(xv, sc[j], sh[j]) > 0.f) ? bf16_at(d, j) : 0.f;
        o[j] = fmaf(sc[j], gr, fmaf(cA[j], xv, cB[j]));
      }
      store8(dx + static_cast<size_t>(base) * g.C + cg * 8, o);
    } else {
      int arg[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float best = 0.f;
        arg[j] = -1;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float z = fmaf(bf16_at(slot[q], j), sc[j], sh[j]);
          if (z > best) {
            best = z;
            arg[j] = q;
          }
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float gr = (arg[j] == q) ? bf16_at(d, j) : 0.f;
          o[j] = fmaf(sc[j], gr, fmaf(cA[j], bf16_at(slot[q], j), cB[j]));
        }
        store8(dx + static_cast<size_t>(base + (q >> 1) * g.W + (q & 1)) * g.C + cg * 8, o);
      }
    }
  }
}

// Measured at batch 256 (profiles/r2_call16_*): layers 6 / 7 (32 / 20 KB per CTA) 9.4 -> 6.1 and 10.2 -> 5.8 us,
// but layers 4 / 5 (128 / 80 KB, 16 items per thread on only 64 CTAs) got 2 us SLOWER than the two-kernel path.
constexpr int kBnOwnMaxSmem = 48 * 1024;

// Returns -1 when the layer is not a channel-owner case (too many pixels for one CTA's shared memory,
// or too few channel groups to be worth it).
template <bool kPool>
static int launch_bn_bwd_owner(const __nv_bfloat16* x, const __nv_bfloat16* dy, const float* saved, const float* gamma,
                               const float* beta, __nv_bfloat16* dx, float* dgamma, float* dbeta, float* dbias,
                               BnGeom g, long long out_pix, cudaStream_t stream) {
  const size_t smem = static_cast<size_t>(out_pix) * (kPool ? 5 : 2) * 16;
  if (smem > static_cast<size_t>(kBnOwnMaxSmem) || g.C / 8 < 32 || out_pix < 1) return -1;
  if (static_cast<long long>(g.N) * g.H * g.W * g.C >= (1LL << 31)) return -1;
  static bool configured = false;
  if (!configured) {
    DDPB_CHECK_CUDA(cudaFuncSetAttribute(bn_bwd_owner_kernel<kPool>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kBnOwnMaxSmem));
    configured = true;
  }
  DDPB_CHECK_CUDA(launch_pdl(bn_bwd_owner_kernel<kPool>, g.C / 8, kBnOwnThreads, smem, stream, x, dy, saved, gamma,
                             beta, dx, dgamma, dbeta, dbias, g));
  return 0;
}

// 0: two-kernel path with float atomics (round 1); 1: fused cooperative kernel, one CTA per SM; 2: fused
// cooperative, up to two CTAs per SM; 3 (default): deterministic two-kernel path above -- falls back to 2
// when the caller gives no partial-sum buffer
static int g_bn_fused = 3;

// Tries the fused cooperative kernel; returns -1 if the shape does not fit in shared memory.
template <bool kPool>
static int launch_bn_bwd_fused(const __nv_bfloat16* x, const __nv_bfloat16* dy, const float* saved,
                               const float* gamma, const float* beta, float* sums, __nv_bfloat16* dx,
                               float* dgamma, float* dbeta, float* dbias, BnGeom g, long long items,
                               cudaStream_t stream) {
  const int groups = g.C / 8;
  if (groups % kSliceGroups != 0) return -1;
  const int n_slices = groups / kSliceGroups;
  const int sms = sm_count();
  const long long out_pix = items / groups;
  // Two CTAs per SM when their slices fit (2 x <=100 KB): twice the loads in flight for a kernel
  // that is bound by memory latency, not bandwidth.  Otherwise one CTA per SM with <=200 KB.
  long long ppc = 0;
  int n_ranges = 0;
  size_t smem = 0;
  for (int per_sm = (g_bn_fused >= 2 ? 2 : 1); per_sm >= 1; --per_sm) {
    n_ranges = (per_sm * sms) / n_slices;
    if (n_ranges < 1) return -1;
    if (n_ranges > out_pix) n_ranges = static_cast<int>(out_pix);
    ppc = (out_pix + n_ranges - 1) / n_ranges;
    n_ranges = static_cast<int>((out_pix + ppc - 1) / ppc);
    smem = static_cast<size_t>(ppc) * kSliceGroups * (kPool ? 5 : 2) * 16;
    if (smem <= static_cast<size_t>(per_sm == 2 ? 100 : 200) * 1024) break;
    if (per_sm == 1) return -1;
  }
  static bool configured = false;
  if (!configured) {
    DDPB_CHECK_CUDA(cudaFuncSetAttribute(bn_bwd_fused_kernel<kPool>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         200 * 1024));
    configured = true;
  }
  int ppc_i = static_cast<int>(ppc);
  void* args[] = {&x, &dy, &saved, &gamma, &beta, &sums, &dx, &dgamma, &dbeta, &dbias, &g, &ppc_i, &n_ranges};
  DDPB_CHECK_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(bn_bwd_fused_kernel<kPool>),
                                              dim3(n_slices * n_ranges), dim3(kBnFusedThreads), args, smem, stream));
  return 0;
}



static int bn_grid(long long items) {
  long long blocks = (items + kBnThreads - 1) / kBnThreads;
  const long long cap = 4LL * sm_count();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace ddpb

using namespace ddpb;

#define DDPB_BN_CHECK_GEOM(name)                                                                   \
  DDPB_REQUIRE(C > 0 && C % 8 == 0 && C <= 2048, name ": C must be a multiple of 8, <= 2048 (got %d)", C); \
  DDPB_REQUIRE(!pool || (H % 2 == 0 && W % 2 == 0), name ": pooling needs even H, W (got %dx%d)", H, W)

// stats[2][C] += (sum x, sum x^2) over the P rows of x[P][C] (bf16).
extern "C" int ddpb_channel_stats(const void* x, float* stats, long long P, int C, cudaStream_t stream) {
  DDPB_REQUIRE(C > 0 && C % 8 == 0 && C <= 2048, "channel_stats: bad C %d", C);
  const long long items = P * (C / 8);
  DDPB_CHECK_CUDA(launch_pdl(channel_stats_kernel, bn_grid(items), kBnThreads, 2 * C * sizeof(float), stream,
                             static_cast<const __nv_bfloat16*>(x), stats, P, C));
  return 0;
}

// y = maxpool2x2?(relu(batchnorm(x)));  x: [N*H*W][C] bf16 raw conv output, stats: [2][C]
// sums over those rows, saved: [2][C] receives (mean, rstd) for the backward pass.
int bn_relu_pool_fwd_launch(const void* x, const float* stats, const float* gamma, const float* beta, void* y,
                            float* saved, int N, int H, int W, int C, int pool, float eps, cudaStream_t stream);

extern "C" int ddpb_bn_relu_pool_fwd(const void* x, const float* stats, const float* gamma, const float* beta,
                                     void* y, float* saved, int N, int H, int W, int C, int pool, float eps,
                                     cudaStream_t stream) {
  return bn_relu_pool_fwd_launch(x, stats, gamma, beta, y, saved, N, H, W, C, pool, eps, stream);
}

int bn_relu_pool_fwd_launch(const void* x, const float* stats, const float* gamma, const float* beta, void* y,
                            float* saved, int N, int H, int W, int C, int pool, float eps, cudaStream_t stream) {
  DDPB_BN_CHECK_GEOM("bn_relu_pool_fwd");
  BnGeom g = make_geom(N, H, W, C, pool, eps);
  const long long items = static_cast<long long>(N) * (pool ? H / 2 : H) * (pool ? W / 2 : W) * (C / 8);
  DDPB_CHECK_CUDA(launch_pdl(bn_relu_pool_fwd_kernel, bn_grid(items), kBnThreads, 4 * C * sizeof(float), stream,
                             static_cast<const __nv_bfloat16*>(x), stats, gamma, beta, static_cast<__nv_bfloat16*>(y),
                             saved, g));
  return 0;
}

extern "C" void ddpb_bn_bwd_set_fused(int enabled) { g_bn_fused = enabled; }

// Backward of the fused block.  dy: gradient w.r.t. the (pooled) output, bf16.
// sums: fp32 scratch of 2*C + 32 floats that MUST be zero on entry (per-channel sums in the first
// 2*C; the words after them are ticket counters / the cooperative kernel's grid barrier.  The default
// path leaves the counters zero again, the older paths do not).
// partials / partial_elems: scratch of ddpb_bn_bwd_partials_elems() floats for the default
// deterministic path (no zeroing needed); null selects the cooperative kernel of round 1.
// dx: [N*H*W][C] bf16 gradient w.r.t. the raw conv output.  dgamma/dbeta/dbias: fp32 [C]
// (dbias may be null).
// Floats of partial-sum scratch the default (deterministic two-kernel) backward wants for this shape.
extern "C" long long ddpb_bn_bwd_partials_elems(int N, int H, int W, int C, int pool) {
  if (C <= 0 || C % (kSliceGroups * 8) != 0) return 0;
  const long long out_pix = static_cast<long long>(N) * (pool ? H / 2 : H) * (pool ? W / 2 : W);
  const Bn2Plan pl = bn2_plan(out_pix, C);
  return static_cast<long long>(pl.n_slices) * pl.n_ranges * 2 * kSliceGroups * 8;
}

extern "C" int ddpb_bn_relu_pool_bwd(const void* x, const void* dy, const float* saved, const float* gamma,
                                     const float* beta, float* sums, void* dx, float* dgamma, float* dbeta,
                                     float* dbias, float* partials, long long partial_elems, int N, int H, int W,
                                     int C, int pool, float eps, cudaStream_t stream) {
  DDPB_BN_CHECK_GEOM("bn_relu_pool_bwd");
  BnGeom g = make_geom(N, H, W, C, pool, eps);
  DDPB_REQUIRE((C / 8) <= kBnThreads && kBnThreads % (C / 8) == 0,
               "bn_relu_pool_bwd: C/8 must divide %d (got C=%d)", kBnThreads, C);
  const long long items = static_cast<long long>(N) * (pool ? H / 2 : H) * (pool ? W / 2 : W) * (C / 8);
  const int grid = bn_grid(items);
  const __nv_bfloat16* xb = static_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* dyb = static_cast<const __nv_bfloat16*>(dy);
  __nv_bfloat16* dxb = static_cast<__nv_bfloat16*>(dx);
  if (g_bn_fused >= 3) {
    // small late layers: one launch, a CTA owns 8 channels and every pixel
    const long long out_pix = items / (C / 8);
    const int rc = pool ? launch_bn_bwd_owner<true>(xb, dyb, saved, gamma, beta, dxb, dgamma, dbeta, dbias, g, out_pix,
                                                    stream)
                        : launch_bn_bwd_owner<false>(xb, dyb, saved, gamma, beta, dxb, dgamma, dbeta, dbias, g, out_pix,
                                                     stream);
    if (rc >= 0) return rc;
  }
  if (g_bn_fused >= 3 && partials != nullptr && C % (kSliceGroups * 8) == 0) {
    const long long out_pix = items / (C / 8);
    return pool ? launch_bn_bwd2<true>(xb, dyb, saved, gamma, beta, sums, partials, partial_elems, dxb, dgamma, dbeta,
                                       dbias, g, out_pix, stream)
                : launch_bn_bwd2<false>(xb, dyb, saved, gamma, beta, sums, partials, partial_elems, dxb, dgamma, dbeta,
                                        dbias, g, out_pix, stream);
  }
  if (g_bn_fused) {
    const int rc = pool ? launch_bn_bwd_fused<true>(xb, dyb, saved, gamma, beta, sums, dxb, dgamma, dbeta, dbias, g,
                                                    items, stream)
                        : launch_bn_bwd_fused<false>(xb, dyb, saved, gamma, beta, sums, dxb, dgamma, dbeta, dbias, g,
                                                     items, stream);
    if (rc >= 0) return rc;   // -1: slice too large for shared memory, use the two-pass kernels
  }
  const float* sums_c = sums;
  if (pool) {
    DDPB_CHECK_CUDA(launch_pdl(bn_bwd_reduce_kernel<true>, grid, kBnThreads, 4 * C * sizeof(float), stream, xb, dyb,
                               saved, gamma, beta, sums, dbias, g));
    DDPB_CHECK_CUDA(launch_pdl(bn_bwd_apply_kernel<true>, grid, kBnThreads, 5 * C * sizeof(float), stream, xb, dyb,
                               saved, gamma, beta, sums_c, dxb, dgamma, dbeta, dbias, g));
  } else {
    DDPB_CHECK_CUDA(launch_pdl(bn_bwd_reduce_kernel<false>, grid, kBnThreads, 4 * C * sizeof(float), stream, xb, dyb,
                               saved, gamma, beta, sums, dbias, g));
    DDPB_CHECK_CUDA(launch_pdl(bn_bwd_apply_kernel<false>, grid, kBnThreads, 5 * C * sizeof(float), stream, xb, dyb,
                               saved, gamma, beta, sums_c, dxb, dgamma, dbeta, dbias, g));
  }
  return 0;
}
