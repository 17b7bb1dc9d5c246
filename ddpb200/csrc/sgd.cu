// K9: SGD with momentum and weight decay over the flat parameter arena
// (reference: optim.SGD(lr=0.1, momentum=0.9, weight_decay=1e-4) at
// /root/reference/part1/main.py:124-125, stepped at :77), followed by the repack of the
// updated fp32 master filters into the two bf16 operand layouts the conv kernels read.
//
// Update rule (torch.optim.SGD, dampening 0, no Nesterov):
//   g <- grad_scale * g + wd * p ;  m <- mu * m + g ;  p <- p - lr * m
// A momentum buffer that starts at zero gives m = g on the first step, which is what
// torch does when it creates the buffer, so no "first step" flag is needed.
#include "common.cuh"

namespace ddpb {

__device__ __forceinline__ float4 sgd_update4(float4 p, float4 g, float4& m, float lr, float mu, float wd,
                                              float gs) {
  g.x = fmaf(gs, g.x, wd * p.x);
  g.y = fmaf(gs, g.y, wd * p.y);
  g.z = fmaf(gs, g.z, wd * p.z);
  g.w = fmaf(gs, g.w, wd * p.w);
  m.x = fmaf(mu, m.x, g.x);
  m.y = fmaf(mu, m.y, g.y);
  m.z = fmaf(mu, m.z, g.z);
  m.w = fmaf(mu, m.w, g.w);
  p.x = fmaf(-lr, m.x, p.x);
  p.y = fmaf(-lr, m.y, p.y);
  p.z = fmaf(-lr, m.z, p.z);
  p.w = fmaf(-lr, m.w, p.w);
  return p;
}

__global__ void __launch_bounds__(256)
sgd_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
           __nv_bfloat16* __restrict__ shadow, long long n, float lr, float mu, float wd, float gs) {
  pdl_launch_dependents();
  pdl_wait();
  const long long n4 = n >> 2;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = reinterpret_cast<const float4*>(g)[i];
    float4 mv = reinterpret_cast<float4*>(m)[i];
    pv = sgd_update4(pv, gv, mv, lr, mu, wd, gs);
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    if (shadow != nullptr)   // bf16 copy in the same flat order = the conv kernels' filter operand
      reinterpret_cast<uint2*>(shadow)[i] = make_uint2(pack_bf16x2(pv.x, pv.y), pack_bf16x2(pv.z, pv.w));
  }
  // tail (arena sizes are padded to 4, but stay correct if they are not)
  for (long long i = (n4 << 2) + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += stride) {
    const float gg = fmaf(gs, g[i], wd * p[i]);
    const float mm = fmaf(mu, m[i], gg);
    m[i] = mm;
    const float pn = fmaf(-lr, mm, p[i]);
    p[i] = pn;
    if (shadow != nullptr) shadow[i] = __float2bfloat16_rn(pn);
  }
}

// ---- filter repack ---------------------------------------------------------------------
// master   : fp32 [K][taps][C]          (a (K,C,3,3) torch tensor in channels_last memory format)
// fprop    : bf16 [K][taps*Cpad]        (Cpad >= C, zero padded; layer 0: taps*C=27 -> one row of 64)
// dgrad    : bf16 [C][taps][K] with the taps rotated by 180 degrees: Wd[c][t][k] = W[k][taps-1-t][c]
struct PackLayer {
  const float* master;
  __nv_bfloat16* fprop;
  __nv_bfloat16* dgrad;  // null: no dgrad pack (first layer)
  int K, C, taps;
  int fprop_ld;          // row length of the fprop pack in elements
  int flat;              // 1: the (taps, C) axes are flattened and zero padded to fprop_ld (first layer)
  int block_begin;       // first block of this layer in the fused launch
  int tiles_k, tiles_c;
};
constexpr int kMaxPackLayers = 16;
struct PackTable {
  PackLayer layer[kMaxPackLayers];
  int count;
};

__global__ void __launch_bounds__(256)
pack_weights_kernel(const __grid_constant__ PackTable tab) {
  __shared__ float s_tile[32][33];
  int li = 0;
  while (li + 1 < tab.count && static_cast<int>(blockIdx.x) >= tab.layer[li + 1].block_begin) ++li;
  const PackLayer& L = tab.layer[li];
  int t = blockIdx.x - L.block_begin;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  if (L.flat) {
    // one block per 32 rows; row length taps*C (tiny)
    const int row_len = L.taps * L.C;
    for (int r = ty; r < 32; r += 8) {
      const int k = t * 32 + r;
      if (k >= L.K) continue;
      for (int c = tx; c < L.fprop_ld; c += 32) {
        const float v = (c < row_len) ? L.master[static_cast<long long>(k) * row_len + c] : 0.f;
        L.fprop[static_cast<long long>(k) * L.fprop_ld + c] = __float2bfloat16_rn(v);
      }
    }
    return;
  }
  const int ct = t % L.tiles_c;
  t /= L.tiles_c;
  const int kt = t % L.tiles_k;
  const int tap = t / L.tiles_k;
  const int k0 = kt * 32, c0 = ct * 32;
  // read 32(k) x 32(c), coalesced along c
  for (int r = ty; r < 32; r += 8) {
    const int k = k0 + r, c = c0 + tx;
    float v = 0.f;
    if (k < L.K && c < L.C) v = L.master[(static_cast<long long>(k) * L.taps + tap) * L.C + c];
    s_tile[r][tx] = v;
    if (k < L.K && c < L.C)
      L.fprop[static_cast<long long>(k) * L.fprop_ld + tap * L.C + c] = __float2bfloat16_rn(v);
  }
  if (L.dgrad == nullptr) return;
  __syncthreads();
  // write 32(c) x 32(k), coalesced along k
  const int tap_r = L.taps - 1 - tap;
  for (int r = ty; r < 32; r += 8) {
    const int c = c0 + r, k = k0 + tx;
    if (c < L.C && k < L.K)
      L.dgrad[(static_cast<long long>(c) * L.taps + tap_r) * L.K + k] = __float2bfloat16_rn(s_tile[tx][r]);
  }
}

}  // namespace ddpb

using namespace ddpb;

// In-place SGD step over n contiguous fp32 elements.  grad_scale multiplies the gradient
// first (1/world_size when the caller holds a SUM of per-rank gradients).  shadow (bf16, n
// elements, may be null) receives the rounded new parameters: because conv filters are
// stored [C_out][3][3][C_in] in the arena, that copy IS the tensor-core operand of fprop
// and dgrad, so no separate repack pass is needed.
extern "C" int ddpb_sgd_step(float* p, const float* g, float* m, void* shadow, long long n, float lr,
                             float momentum, float weight_decay, float grad_scale, cudaStream_t stream) {
  DDPB_REQUIRE((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m)) % 16 == 0,
               "sgd_step: arenas must be 16-byte aligned");
  long long blocks = ((n >> 2) + 255) / 256;
  const long long cap = 8LL * sm_count();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  DDPB_CHECK_CUDA(launch_pdl(sgd_kernel, static_cast<int>(blocks), 256, 0, stream, p, g, m,
                             static_cast<__nv_bfloat16*>(shadow), n, lr, momentum, weight_decay, grad_scale));
  return 0;
}

// Repack `count` conv layers in ONE launch.  Arrays are per layer; dgrad[i] may be null;
// flat[i] = 1 selects the first-layer form (row of taps*C values zero padded to fprop_ld).
extern "C" int ddpb_pack_weights(int count, const float* const* master, void* const* fprop, void* const* dgrad,
                                 const int* K, const int* C, const int* taps, const int* fprop_ld,
                                 const int* flat, cudaStream_t stream) {
  DDPB_REQUIRE(count >= 1 && count <= kMaxPackLayers, "pack_weights: 1..%d layers (got %d)", kMaxPackLayers,
               count);
  PackTable tab;
  tab.count = count;
  int blocks = 0;
  for (int i = 0; i < count; ++i) {
    PackLayer& L = tab.layer[i];
    L.master = master[i];
    L.fprop = static_cast<__nv_bfloat16*>(fprop[i]);
    L.dgrad = static_cast<__nv_bfloat16*>(dgrad[i]);
    L.K = K[i];
    L.C = C[i];
    L.taps = taps[i];
    L.fprop_ld = fprop_ld[i];
    L.flat = flat[i];
    L.block_begin = blocks;
    L.tiles_k = (K[i] + 31) / 32;
    L.tiles_c = (C[i] + 31) / 32;
    if (L.flat) {
      DDPB_REQUIRE(L.fprop_ld >= L.taps * L.C, "pack_weights: layer %d: fprop_ld %d < taps*C %d", i, L.fprop_ld,
                   L.taps * L.C);
      blocks += L.tiles_k;
    } else {
      DDPB_REQUIRE(L.fprop_ld == L.taps * L.C, "pack_weights: layer %d: fprop_ld must equal taps*C", i);
      blocks += L.taps * L.tiles_k * L.tiles_c;
    }
  }
  pack_weights_kernel<<<blocks, 256, 0, stream>>>(tab);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
