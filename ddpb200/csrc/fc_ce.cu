// K7/K8/K15: classifier Linear(512,10) + CrossEntropyLoss(mean), forward and
// backward, plus top-1 accuracy.  Reference: fc1 at /root/reference/part1/model.py:40,45,
// criterion at /root/reference/part1/main.py:75,119, accuracy at part1/main.py:105-106.
// The GEMM is 10 columns wide, so it lives on CUDA cores: one warp per sample in the
// forward; the backward splits the feature dimension over CTAs so dW needs no cross-CTA
// reduction and the result is deterministic.
#include "common.cuh"

namespace ddpb {

constexpr int kMaxClasses = 16;

// kK: compile-time class count (10 for the reference model) or kMaxClasses with the real
// count checked at run time.
template <int kK>
__device__ __forceinline__ bool class_ok(int j, int K) { return kK == kMaxClasses ? j < K : true; }

// One warp per row; the K x F weight matrix is staged in shared memory once per CTA.  The
// row's activations are requested BEFORE the staging barrier so both latencies overlap.
// Writes logits, per-row loss, dlogits = (softmax - onehot)/B and a per-row "correct" flag.
constexpr int kFwdMaxF = 1024;   // features per row held in registers: 32 lanes x 32

template <int kK>
__global__ void __launch_bounds__(256)
fc_ce_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                 const long long* __restrict__ labels, float* __restrict__ logits, float* __restrict__ row_loss,
                 float* __restrict__ dlogits, int* __restrict__ row_correct, int B, int F, int K) {
  extern __shared__ float s_w[];  // [K][F]
  pdl_launch_dependents();
  pdl_wait();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const bool row_ok = warp < B;
  const __nv_bfloat16* xr = x + static_cast<long long>(row_ok ? warp : 0) * F;
  // lane owns feature pairs {2*lane + 64*t, +1}: 4-byte loads, conflict-free float2 reads of s_w
  constexpr int kPairs = kFwdMaxF / 64;
  uint32_t xraw[kPairs];
#pragma unroll
  for (int t = 0; t < kPairs; ++t) {
    const int i = 2 * lane + 64 * t;
    xraw[t] = (row_ok && i < F) ? __ldg(reinterpret_cast<const uint32_t*>(xr + i)) : 0u;
  }
  const int label = row_ok ? static_cast<int>(labels[warp]) : 0;
  {
    // 16-byte loads, several in flight per thread: the staging is pure latency otherwise
    const int n4 = (K * F) >> 2;
    const float4* w4 = reinterpret_cast<const float4*>(w);
    float4* s4 = reinterpret_cast<float4*>(s_w);
#pragma unroll 8
    for (int i = threadIdx.x; i < n4; i += blockDim.x) s4[i] = __ldg(w4 + i);
  }
  __syncthreads();
  if (!row_ok) return;
  float acc[kK];
#pragma unroll
  for (int j = 0; j < kK; ++j) acc[j] = 0.f;
#pragma unroll
  for (int t = 0; t < kPairs; ++t) {
    const int i = 2 * lane + 64 * t;
    if (i < F) {
      const float x0 = __uint_as_float(xraw[t] << 16), x1 = __uint_as_float(xraw[t] & 0xffff0000u);
#pragma unroll
      for (int j = 0; j < kK; ++j)
        if (class_ok<kK>(j, K)) {
          const float2 wv = *reinterpret_cast<const float2*>(s_w + j * F + i);
          acc[j] = fmaf(x1, wv.y, fmaf(x0, wv.x, acc[j]));
        }
    }
  }
#pragma unroll
  for (int j = 0; j < kK; ++j) {
    if (class_ok<kK>(j, K)) {
#pragma unroll
      for (int s = 16; s >= 1; s >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], s);
      acc[j] += b[j];
    }
  }
  // every lane now holds all K logits
  float mx = -INFINITY;
  int arg = 0;
#pragma unroll
  for (int j = 0; j < kK; ++j)
    if (class_ok<kK>(j, K) && acc[j] > mx) {
      mx = acc[j];
      arg = j;
    }
  float se = 0.f;
#pragma unroll
  for (int j = 0; j < kK; ++j)
    if (class_ok<kK>(j, K)) se += __expf(acc[j] - mx);
  const float lse = mx + __logf(se);
  if (lane == 0) {
    float picked = 0.f;
#pragma unroll
    for (int j = 0; j < kK; ++j)
      if (j == label) picked = acc[j];
    row_loss[warp] = lse - picked;
    if (row_correct) row_correct[warp] = (arg == label) ? 1 : 0;
  }
  const float invB = 1.f / static_cast<float>(B);
#pragma unroll
  for (int j = 0; j < kK; ++j) {
    if (class_ok<kK>(j, K) && lane == j) {
      if (logits) logits[warp * K + j] = acc[j];
      if (dlogits) dlogits[warp * K + j] = (__expf(acc[j] - lse) - (j == label ? 1.f : 0.f)) * invB;
    }
  }
}

// Deterministic mean loss and correct-count by one CTA of 256 threads (fixed summation order).
__device__ __forceinline__ void loss_reduce_block(const float* __restrict__ row_loss,
                                                  const int* __restrict__ row_correct, float* __restrict__ loss_out,
                                                  int* __restrict__ correct_out, int B, float* s_l, int* s_c) {
  float l = 0.f;
  int c = 0;
  for (int i = threadIdx.x; i < B; i += 256) {
    l += row_loss[i];
    if (row_correct) c += row_correct[i];
  }
  s_l[threadIdx.x] = l;
  s_c[threadIdx.x] = c;
  __syncthreads();
  for (int s = 128; s >= 1; s >>= 1) {
    if (threadIdx.x < s) {
      s_l[threadIdx.x] += s_l[threadIdx.x + s];
      s_c[threadIdx.x] += s_c[threadIdx.x + s];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    if (loss_out) *loss_out = s_l[0] / static_cast<float>(B);
    if (correct_out) *correct_out = s_c[0];
  }
}

__global__ void __launch_bounds__(256)
loss_reduce_kernel(const float* __restrict__ row_loss, const int* __restrict__ row_correct,
                   float* __restrict__ loss_out, int* __restrict__ correct_out, int B) {
  __shared__ float s_l[256];
  __shared__ int s_c[256];
  pdl_launch_dependents();
  pdl_wait();
  loss_reduce_block(row_loss, row_correct, loss_out, correct_out, B, s_l, s_c);
}

// Backward.  CTA c owns kFcFeat features: dW[:, own] and dx[:, own], so dW needs no cross-CTA
// reduction and is deterministic.  256 threads = 8 feature pairs x 32 row slices; each thread
// first requests ALL of its rows of x (kFcRows/32 independent 4-byte loads), then runs the
// FMAs against dlogits staged in shared memory.  The CTA after the last feature CTA computes
// db and, when asked, the mean loss / correct count of the forward (saves a launch).
constexpr int kFcRows = 512;   // rows per pass
constexpr int kFcFeat = 16;    // features per CTA
constexpr int kFcRowsPerThread = kFcRows / 32;

template <int kK>
__global__ void __launch_bounds__(256)
fc_bwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ dlogits,
              float* __restrict__ dw, float* __restrict__ db, __nv_bfloat16* __restrict__ dx, int B, int F, int K,
              const float* __restrict__ row_loss, const int* __restrict__ row_correct, float* __restrict__ loss_out,
              int* __restrict__ correct_out) {
  extern __shared__ float s_dl[];  // [kFcRows][K]; reused for the cross-warp reduction
  pdl_launch_dependents();
  pdl_wait();
  if (blockIdx.x == gridDim.x - 1) {
    // ---- db (+ loss): 16 threads per class-group walk the rows
    __shared__ float s_db[16][kMaxClasses];
    const int j = threadIdx.x & 15;
    float dbv = 0.f;
    if (j < K)
#pragma unroll 8
      for (int r = threadIdx.x >> 4; r < B; r += 16) dbv += dlogits[static_cast<long long>(r) * K + j];
    s_db[threadIdx.x >> 4][j] = dbv;
    __syncthreads();
    if (threadIdx.x < K) {
      float t = 0.f;
#pragma unroll
      for (int g = 0; g < 16; ++g) t += s_db[g][threadIdx.x];
      db[threadIdx.x] = t;
    }
    if (row_loss != nullptr && (loss_out != nullptr || correct_out != nullptr)) {
      float* s_l = s_dl;
      int* s_c = reinterpret_cast<int*>(s_dl + 256);
      loss_reduce_block(row_loss, row_correct, loss_out, correct_out, B, s_l, s_c);
    }
    return;
  }
  const int pair = threadIdx.x & 7;
  const int slice = threadIdx.x >> 3;   // 0..31
  const int f = blockIdx.x * kFcFeat + 2 * pair;
  const bool fok = f < F;               // F is even (checked by the host)
  float w0[kK], w1[kK], a0[kK], a1[kK];
#pragma unroll
  for (int j = 0; j < kK; ++j) {
    const bool ok = class_ok<kK>(j, K) && fok;
    const float2 wv = ok ? __ldg(reinterpret_cast<const float2*>(w + j * F + f)) : make_float2(0.f, 0.f);
    w0[j] = wv.x;
    w1[j] = wv.y;
    a0[j] = a1[j] = 0.f;
  }
  for (int r0 = 0; r0 < B; r0 += kFcRows) {
    const int rows = (B - r0) < kFcRows ? (B - r0) : kFcRows;
    uint32_t xraw[kFcRowsPerThread];
#pragma unroll
    for (int u = 0; u < kFcRowsPerThread; ++u) {
      const int r = slice + 32 * u;
      xraw[u] = (fok && r < rows) ? __ldg(reinterpret_cast<const uint32_t*>(x + static_cast<long long>(r0 + r) * F + f)) : 0u;
    }
    if (r0 > 0) __syncthreads();   // previous pass is done with s_dl
#pragma unroll 4
    for (int i = threadIdx.x; i < rows * K; i += 256) s_dl[i] = __ldg(dlogits + static_cast<long long>(r0) * K + i);
    __syncthreads();
#pragma unroll
    for (int u = 0; u < kFcRowsPerThread; ++u) {
      const int r = slice + 32 * u;
      if (r < rows) {
        const float x0 = __uint_as_float(xraw[u] << 16), x1 = __uint_as_float(xraw[u] & 0xffff0000u);
        float d0 = 0.f, d1 = 0.f;
#pragma unroll
        for (int j = 0; j < kK; ++j) {
          if (class_ok<kK>(j, K)) {
            const float d = s_dl[r * K + j];
            a0[j] = fmaf(d, x0, a0[j]);
            a1[j] = fmaf(d, x1, a1[j]);
            d0 = fmaf(d, w0[j], d0);
            d1 = fmaf(d, w1[j], d1);
          }
        }
        if (dx != nullptr && fok)
          *reinterpret_cast<uint32_t*>(dx + static_cast<long long>(r0 + r) * F + f) = pack_bf16x2(d0, d1);
      }
    }
  }
  // reduce over the 32 row slices: 4 slices live in each warp (lane bits 3,4), 8 warps via smem
#pragma unroll
  for (int j = 0; j < kK; ++j) {
    a0[j] += __shfl_xor_sync(0xffffffffu, a0[j], 8);
    a0[j] += __shfl_xor_sync(0xffffffffu, a0[j], 16);
    a1[j] += __shfl_xor_sync(0xffffffffu, a1[j], 8);
    a1[j] += __shfl_xor_sync(0xffffffffu, a1[j], 16);
  }
  __syncthreads();   // s_dl is free
  float* s_red = s_dl;   // [8 warps][kK][kFcFeat]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane < 8) {
#pragma unroll
    for (int j = 0; j < kK; ++j) {
      s_red[(warp * kK + j) * kFcFeat + 2 * lane] = a0[j];
      s_red[(warp * kK + j) * kFcFeat + 2 * lane + 1] = a1[j];
    }
  }
  __syncthreads();
  for (int o = threadIdx.x; o < kK * kFcFeat; o += 256) {
    const int j = o / kFcFeat, fl = o % kFcFeat;
    const int ff = blockIdx.x * kFcFeat + fl;
    if (class_ok<kK>(j, K) && ff < F) {
      float t = 0.f;
#pragma unroll
      for (int wv = 0; wv < 8; ++wv) t += s_red[(wv * kK + j) * kFcFeat + fl];
      dw[j * F + ff] = t;
    }
  }
}

}  // namespace ddpb

using namespace ddpb;

// x: [B][F] bf16, w: [K][F] fp32, b: [K] fp32, labels: [B] int64.
// Outputs (any may be null except row_loss): logits [B][K], row_loss [B], dlogits [B][K]
// (already divided by B, i.e. the gradient of the MEAN loss), row_correct [B],
// loss (scalar mean), correct (scalar count).
extern "C" int ddpb_fc_ce_fwd(const void* x, const float* w, const float* b, const long long* labels,
                              float* logits, float* row_loss, float* dlogits, int* row_correct, float* loss,
                              int* correct, int B, int F, int K, cudaStream_t stream) {
  DDPB_REQUIRE(K >= 1 && K <= kMaxClasses, "fc_ce_fwd: at most %d classes (got %d)", kMaxClasses, K);
  DDPB_REQUIRE(row_loss != nullptr, "fc_ce_fwd: row_loss buffer is required");
  DDPB_REQUIRE(F % 4 == 0 && F <= kFwdMaxF, "fc_ce_fwd: features must be a multiple of 4 and <= %d (got %d)", kFwdMaxF, F);
  const int warps_per_block = 8;
  const int blocks = (B + warps_per_block - 1) / warps_per_block;
  const size_t smem = static_cast<size_t>(K) * F * sizeof(float);
  DDPB_REQUIRE(smem <= 48 * 1024, "fc_ce_fwd: K*F = %d exceeds the 48 KB weight stage", K * F);
  const __nv_bfloat16* xb = static_cast<const __nv_bfloat16*>(x);
  if (K == 10)
    DDPB_CHECK_CUDA(launch_pdl(fc_ce_fwd_kernel<10>, blocks, 256, smem, stream, xb, w, b, labels, logits, row_loss,
                               dlogits, row_correct, B, F, K));
  else
    DDPB_CHECK_CUDA(launch_pdl(fc_ce_fwd_kernel<kMaxClasses>, blocks, 256, smem, stream, xb, w, b, labels, logits,
                               row_loss, dlogits, row_correct, B, F, K));
  if (loss != nullptr || correct != nullptr)
    DDPB_CHECK_CUDA(launch_pdl(loss_reduce_kernel, 1, 256, 0, stream, row_loss, row_correct, loss, correct, B));
  return 0;
}

// dw: [K][F] fp32, db: [K] fp32, dx: [B][F] bf16 (may be null).  row_loss / row_correct / loss /
// correct (all may be null): when given, the kernel also writes the mean loss and the correct
// count of the preceding ddpb_fc_ce_fwd (called with loss = correct = null).
extern "C" int ddpb_fc_bwd(const void* x, const float* w, const float* dlogits, float* dw, float* db, void* dx,
                           int B, int F, int K, const float* row_loss, const int* row_correct, float* loss,
                           int* correct, cudaStream_t stream) {
  DDPB_REQUIRE(K >= 1 && K <= kMaxClasses, "fc_bwd: at most %d classes (got %d)", kMaxClasses, K);
  DDPB_REQUIRE(F % 2 == 0, "fc_bwd: the feature count must be even (got %d)", F);
  const int blocks = (F + kFcFeat - 1) / kFcFeat + 1;
  size_t smem = static_cast<size_t>(kFcRows) * K * sizeof(float);
  const size_t red = static_cast<size_t>(8) * kMaxClasses * kFcFeat * sizeof(float);
  if (smem < red) smem = red;
  const __nv_bfloat16* xb = static_cast<const __nv_bfloat16*>(x);
  __nv_bfloat16* dxb = static_cast<__nv_bfloat16*>(dx);
  if (K == 10)
    DDPB_CHECK_CUDA(launch_pdl(fc_bwd_kernel<10>, blocks, 256, smem, stream, xb, w, dlogits, dw, db, dxb, B, F, K,
                               row_loss, row_correct, loss, correct));
  else
    DDPB_CHECK_CUDA(launch_pdl(fc_bwd_kernel<kMaxClasses>, blocks, 256, smem, stream, xb, w, dlogits, dw, db, dxb, B,
                               F, K, row_loss, row_correct, loss, correct));
  return 0;
}
