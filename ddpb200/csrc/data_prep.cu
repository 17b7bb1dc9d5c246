// K16: input pipeline kernels.  The reference augments on CPU workers with
// torchvision (RandomCrop(32, padding=4), RandomHorizontalFlip, ToTensor, Normalize:
// /root/reference/part1/main.py:20-28); here the raw uint8 batch is copied to the GPU and
// one kernel crops, flips, normalises AND lays the image out as the first conv layer's
// GEMM operand: a [N*H*W][64] bf16 matrix whose row for pixel (h,w) holds the 3x3x3
// neighbourhood (column = tap*3 + channel, zero beyond 27).  With that operand layer 0
// runs on the same tcgen05 kernel as the others (K2).
#include "common.cuh"

namespace ddpb {

struct PrepParams {
  int N, H, W;
  float mean[3], inv_std[3];
  int pad;  // RandomCrop padding (4) when aug != null
};

// aug: per image {crop offset y, crop offset x, flip, unused} or null (no augmentation).
template <bool kUint8>
__device__ __forceinline__ float src_value(const void* __restrict__ src, const PrepParams& p,
                                           const signed char* __restrict__ aug, int n, int h, int w, int c) {
  if (h < 0 || h >= p.H || w < 0 || w >= p.W) return 0.f;  // conv zero padding
  if (kUint8) {
    int sh = h, sw = w;
    if (aug != nullptr) {
      const signed char* a = aug + n * 4;
      if (a[2]) sw = p.W - 1 - sw;                 // flip applies to the cropped image
      sh = sh + a[0] - p.pad;
      sw = sw + a[1] - p.pad;
    }
    float raw = 0.f;                                // torchvision pads the uint8 image with 0
    if (sh >= 0 && sh < p.H && sw >= 0 && sw < p.W)
      raw = static_cast<float>(static_cast<const unsigned char*>(src)[((static_cast<long long>(n) * p.H + sh) * p.W + sw) * 3 + c]);
    return (raw * (1.f / 255.f) - p.mean[c]) * p.inv_std[c];
  } else {
    return static_cast<const float*>(src)[((static_cast<long long>(n) * 3 + c) * p.H + h) * p.W + w];
  }
}

// One thread gathers the 27 neighbourhood values of one pixel (taps unrolled at compile
// time), the block then writes its 256 rows of 128 bytes fully coalesced from shared memory.
template <bool kUint8>
__global__ void __launch_bounds__(256)
first_layer_operand_kernel(const void* __restrict__ src, const signed char* __restrict__ aug,
                           __nv_bfloat16* __restrict__ out, PrepParams p) {
  __shared__ uint32_t s_rows[256][17];   // 32 bf16 columns per pixel (+1 word of padding)
  const long long P = static_cast<long long>(p.N) * p.H * p.W;
  const long long pix = blockIdx.x * 256LL + threadIdx.x;
  if (pix < P) {
    const int w = static_cast<int>(pix % p.W);
    const long long t = pix / p.W;
    const int h = static_cast<int>(t % p.H);
    const int n = static_cast<int>(t / p.H);
    float v[32];
#pragma unroll
    for (int col = 0; col < 32; ++col) {
      if (col < 27) {
        const int tap = col / 3, c = col % 3;
        v[col] = src_value<kUint8>(src, p, aug, n, h + tap / 3 - 1, w + tap % 3 - 1, c);
      } else {
        v[col] = 0.f;
      }
    }
#pragma unroll
    for (int j = 0; j < 16; ++j) s_rows[threadIdx.x][j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
  }
  __syncthreads();
  uint4* dst = reinterpret_cast<uint4*>(out) + blockIdx.x * 256LL * 8;
  const long long rows_here = (P - blockIdx.x * 256LL) < 256 ? (P - blockIdx.x * 256LL) : 256;
  for (int i = threadIdx.x; i < rows_here * 8; i += 256) {
    const int r = i >> 3, chunk = i & 7;
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (chunk < 4) o = make_uint4(s_rows[r][chunk * 4], s_rows[r][chunk * 4 + 1], s_rows[r][chunk * 4 + 2],
                                  s_rows[r][chunk * 4 + 3]);
    dst[i] = o;
  }
}

static int prep_grid(long long pixels) {
  long long blocks = (pixels + 255) / 256;   // one block per 256 pixels (not grid-stride)
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace ddpb

using namespace ddpb;

// src: fp32 NCHW [N][3][H][W], already normalised.  out: bf16 [N*H*W][64].
extern "C" int ddpb_first_layer_operand_f32(const float* src, void* out, int N, int H, int W,
                                            cudaStream_t stream) {
  PrepParams p{};
  p.N = N;
  p.H = H;
  p.W = W;
  first_layer_operand_kernel<false><<<prep_grid(static_cast<long long>(N) * H * W), 256, 0, stream>>>(
      src, nullptr, static_cast<__nv_bfloat16*>(out), p);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// src: uint8 NHWC [N][H][W][3] raw pixels.  aug: int8 [N][4] = {crop_y, crop_x, flip, 0} with crop
// offsets in [0, 2*pad], or null for the evaluation transform (normalise only).
extern "C" int ddpb_first_layer_operand_u8(const unsigned char* src, const signed char* aug, void* out, int N,
                                           int H, int W, const float* mean3, const float* std3, int pad,
                                           cudaStream_t stream) {
  PrepParams p{};
  p.N = N;
  p.H = H;
  p.W = W;
  p.pad = pad;
  for (int c = 0; c < 3; ++c) {
    DDPB_REQUIRE(std3[c] > 0.f, "first_layer_operand_u8: std must be positive");
    p.mean[c] = mean3[c];
    p.inv_std[c] = 1.f / std3[c];
  }
  first_layer_operand_kernel<true><<<prep_grid(static_cast<long long>(N) * H * W), 256, 0, stream>>>(
      src, aug, static_cast<__nv_bfloat16*>(out), p);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
