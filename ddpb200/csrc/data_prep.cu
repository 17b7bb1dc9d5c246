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

template <bool kUint8>
__global__ void __launch_bounds__(256)
first_layer_operand_kernel(const void* __restrict__ src, const signed char* __restrict__ aug,
                           __nv_bfloat16* __restrict__ out, PrepParams p) {
  const long long P = static_cast<long long>(p.N) * p.H * p.W;
  // 8 threads per pixel, each writes one 16-byte chunk (8 columns) of the 128-byte row
  const long long items = P * 8;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < items; i += stride) {
    const int chunk = static_cast<int>(i & 7);
    const long long pix = i >> 3;
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (chunk < 4) {
      const int w = static_cast<int>(pix % p.W);
      const long long t = pix / p.W;
      const int h = static_cast<int>(t % p.H);
      const int n = static_cast<int>(t / p.H);
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int col = chunk * 8 + j;
        float x = 0.f;
        if (col < 27) {
          const int tap = col / 3, c = col - tap * 3;
          x = src_value<kUint8>(src, p, aug, n, h + tap / 3 - 1, w + tap % 3 - 1, c);
        }
        v[j] = x;
      }
      o.x = pack_bf16x2(v[0], v[1]);
      o.y = pack_bf16x2(v[2], v[3]);
      o.z = pack_bf16x2(v[4], v[5]);
      o.w = pack_bf16x2(v[6], v[7]);
    }
    reinterpret_cast<uint4*>(out)[i] = o;
  }
}

static int prep_grid(long long items) {
  long long blocks = (items + 255) / 256;
  const long long cap = 8LL * sm_count();
  if (blocks > cap) blocks = cap;
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
  first_layer_operand_kernel<false><<<prep_grid(static_cast<long long>(N) * H * W * 8), 256, 0, stream>>>(
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
  first_layer_operand_kernel<true><<<prep_grid(static_cast<long long>(N) * H * W * 8), 256, 0, stream>>>(
      src, aug, static_cast<__nv_bfloat16*>(out), p);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
