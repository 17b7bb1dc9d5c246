// K7/K8/K15: classifier Linear(512,10) + CrossEntropyLoss(mean), forward and
// backward, plus top-1 accuracy.  Reference: fc1 at /root/reference/part1/model.py:40,45,
// criterion at /root/reference/part1/main.py:75,119, accuracy at part1/main.py:105-106.
// The GEMM is 10 columns wide, so it lives on CUDA cores: one warp per sample in the
// forward; the backward splits the feature dimension over CTAs so dW needs no cross-CTA
// reduction and the result is deterministic.
#include "common.cuh"

namespace ddpb {

constexpr int kMaxClasses = 16;

// One warp per row; the K x F weight matrix is staged in shared memory once per CTA.
// Writes logits, per-row loss, dlogits = (softmax - onehot)/B and a per-row "correct" flag.
__global__ void __launch_bounds__(256)
fc_ce_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                 const long long* __restrict__ labels, float* __restrict__ logits, float* __restrict__ row_loss,
                 float* __restrict__ dlogits, int* __restrict__ row_correct, int B, int F, int K) {
  extern __shared__ float s_w[];  // [K][F]
  {
    // 16-byte loads, several in flight per thread: the staging is pure latency otherwise
    const int n4 = (K * F) >> 2;
    const float4* w4 = reinterpret_cast<const float4*>(w);
    float4* s4 = reinterpret_cast<float4*>(s_w);
#pragma unroll 8
    for (int i = threadIdx.x; i < n4; i += blockDim.x) s4[i] = __ldg(w4 + i);
    for (int i = (n4 << 2) + threadIdx.x; i < K * F; i += blockDim.x) s_w[i] = w[i];
  }
  __syncthreads();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= B) return;
  float acc[kMaxClasses];
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j) acc[j] = 0.f;
  const __nv_bfloat16* xr = x + static_cast<long long>(warp) * F;
#pragma unroll 16
  for (int i = lane; i < F; i += 32) {
    const float xv = __bfloat162float(xr[i]);
#pragma unroll
    for (int j = 0; j < kMaxClasses; ++j)
      if (j < K) acc[j] = fmaf(xv, s_w[j * F + i], acc[j]);   // lanes hit consecutive banks
  }
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j) {
    if (j < K) {
#pragma unroll
      for (int s = 16; s >= 1; s >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], s);
      acc[j] += b[j];
    }
  }
  // every lane now holds all K logits
  float mx = -INFINITY;
  int arg = 0;
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j)
    if (j < K && acc[j] > mx) {
      mx = acc[j];
      arg = j;
    }
  float se = 0.f;
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j)
    if (j < K) se += __expf(acc[j] - mx);
  const float lse = mx + __logf(se);
  const int label = static_cast<int>(labels[warp]);
  if (lane == 0) {
    float picked = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxClasses; ++j)
      if (j == label) picked = acc[j];
    row_loss[warp] = lse - picked;
    if (row_correct) row_correct[warp] = (arg == label) ? 1 : 0;
  }
  const float invB = 1.f / static_cast<float>(B);
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j) {
    if (j < K && lane == j) {
      if (logits) logits[warp * K + j] = acc[j];
      if (dlogits) dlogits[warp * K + j] = (__expf(acc[j] - lse) - (j == label ? 1.f : 0.f)) * invB;
    }
  }
}

// Deterministic mean loss and correct-count (single block).
__global__ void __launch_bounds__(256)
loss_reduce_kernel(const float* __restrict__ row_loss, const int* __restrict__ row_correct,
                   float* __restrict__ loss_out, int* __restrict__ correct_out, int B) {
  __shared__ float s_l[256];
  __shared__ int s_c[256];
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

// Backward.  CTA c owns features [c*32, c*32+32): dW[:, own], dx[:, own]; CTA 0 also db.
// 256 threads = 32 features x 8 row slices; dlogits are staged through shared memory in
// chunks of kFcRows rows so the inner loop only issues the (independent) loads of x.
constexpr int kFcRows = 512;

__global__ void __launch_bounds__(256)
fc_bwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ w, const float* __restrict__ dlogits,
              float* __restrict__ dw, float* __restrict__ db, __nv_bfloat16* __restrict__ dx, int B, int F, int K) {
  __shared__ float s_part[8][kMaxClasses][33];
  extern __shared__ float s_dl[];  // [kFcRows][K]
  const int f_local = threadIdx.x & 31;
  const int slice = threadIdx.x >> 5;
  const int f = blockIdx.x * 32 + f_local;
  const bool fok = f < F;
  float wcol[kMaxClasses], acc[kMaxClasses], dbv = 0.f;
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j) {
    wcol[j] = (j < K && fok) ? w[j * F + f] : 0.f;
    acc[j] = 0.f;
  }
  for (int r0 = 0; r0 < B; r0 += kFcRows) {
    const int rows = (B - r0) < kFcRows ? (B - r0) : kFcRows;
#pragma unroll 8
    for (int i = threadIdx.x; i < rows * K; i += blockDim.x) s_dl[i] = __ldg(dlogits + static_cast<long long>(r0) * K + i);
    __syncthreads();
#pragma unroll 16
    for (int r = slice; r < rows; r += 8) {
      const float xv = fok ? __bfloat162float(x[static_cast<long long>(r0 + r) * F + f]) : 0.f;
      float dxv = 0.f;
#pragma unroll
      for (int j = 0; j < kMaxClasses; ++j) {
        if (j < K) {
          const float d = s_dl[r * K + j];   // same address across the warp: broadcast
          acc[j] = fmaf(d, xv, acc[j]);
          dxv = fmaf(d, wcol[j], dxv);
        }
      }
      if (dx != nullptr && fok) dx[static_cast<long long>(r0 + r) * F + f] = __float2bfloat16_rn(dxv);
    }
    if (blockIdx.x == 0) {
      // db: every thread sums a strided share of the rows for class (tid % 16), then the 16
      // threads that share a class are combined below
      const int j = threadIdx.x & 15;
      if (j < K)
        for (int r = threadIdx.x >> 4; r < rows; r += 16) dbv += s_dl[r * K + j];
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < kMaxClasses; ++j) s_part[slice][j][f_local] = acc[j];
  __syncthreads();
  if (slice == 0 && fok) {
#pragma unroll
    for (int j = 0; j < kMaxClasses; ++j) {
      if (j < K) {
        float t = 0.f;
#pragma unroll
        for (int sl = 0; sl < 8; ++sl) t += s_part[sl][j][f_local];
        dw[j * F + f] = t;
      }
    }
  }
  if (blockIdx.x == 0) {
    __syncthreads();                      // s_part is free again
    float* s_db = &s_part[0][0][0];       // [16 row groups][16 classes]
    s_db[(threadIdx.x >> 4) * 16 + (threadIdx.x & 15)] = dbv;
    __syncthreads();
    if (threadIdx.x < K) {
      float t = 0.f;
#pragma unroll
      for (int g = 0; g < 16; ++g) t += s_db[g * 16 + threadIdx.x];
      db[threadIdx.x] = t;
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
  const int warps_per_block = 8;
  const int blocks = (B + warps_per_block - 1) / warps_per_block;
  const size_t smem = static_cast<size_t>(K) * F * sizeof(float);
  DDPB_REQUIRE(smem <= 48 * 1024, "fc_ce_fwd: K*F = %d exceeds the 48 KB weight stage", K * F);
  fc_ce_fwd_kernel<<<blocks, 256, smem, stream>>>(static_cast<const __nv_bfloat16*>(x), w, b, labels, logits,
                                                   row_loss, dlogits, row_correct, B, F, K);
  DDPB_CHECK_CUDA(cudaGetLastError());
  if (loss != nullptr || correct != nullptr) {
    loss_reduce_kernel<<<1, 256, 0, stream>>>(row_loss, row_correct, loss, correct, B);
    DDPB_CHECK_CUDA(cudaGetLastError());
  }
  return 0;
}

// dw: [K][F] fp32, db: [K] fp32, dx: [B][F] bf16 (may be null).
extern "C" int ddpb_fc_bwd(const void* x, const float* w, const float* dlogits, float* dw, float* db, void* dx,
                           int B, int F, int K, cudaStream_t stream) {
  DDPB_REQUIRE(K >= 1 && K <= kMaxClasses, "fc_bwd: at most %d classes (got %d)", kMaxClasses, K);
  const int blocks = (F + 31) / 32;
  fc_bwd_kernel<<<blocks, 256, static_cast<size_t>(kFcRows) * K * sizeof(float), stream>>>(
      static_cast<const __nv_bfloat16*>(x), w, dlogits, dw, db, static_cast<__nv_bfloat16*>(dx), B, F, K);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
