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
    // selects, not p.mean[c]: a run-time index would move the parameter arrays to local memory
    const float m = c == 0 ? p.mean[0] : (c == 1 ? p.mean[1] : p.mean[2]);
    const float is = c == 0 ? p.inv_std[0] : (c == 1 ? p.inv_std[1] : p.inv_std[2]);
    return (raw * (1.f / 255.f) - m) * is;
  } else {
    return static_cast<const float*>(src)[((static_cast<long long>(n) * 3 + c) * p.H + h) * p.W + w];
  }
}

// Zero-padded bf16 pack of the first layer's filters, [C_out][27] fp32 -> [C_out][64] bf16, done
// by one extra CTA of the operand kernel (the filters change every step, the pack is 8 KB).
struct PackArgs {
  const float* master;      // null: no pack
  __nv_bfloat16* packed;
  int c_out;
};

__device__ __forceinline__ void pack_first_filters(const PackArgs& k) {
  for (int i = threadIdx.x; i < k.c_out * 64; i += blockDim.x) {
    const int row = i >> 6, col = i & 63;
    k.packed[i] = __float2bfloat16_rn(col < 27 ? k.master[row * 27 + col] : 0.f);
  }
}

__device__ __forceinline__ void write_rows(const uint32_t (*s_rows)[17], __nv_bfloat16* __restrict__ out,
                                           long long P) {
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

// General shapes: one thread gathers the 27 neighbourhood values of one pixel from global
// memory, the block then writes its 256 rows of 128 bytes coalesced from shared memory.
template <bool kUint8>
__global__ void __launch_bounds__(256)
first_layer_operand_kernel(const void* __restrict__ src, const signed char* __restrict__ aug,
                           __nv_bfloat16* __restrict__ out, PrepParams p, PackArgs k) {
  __shared__ uint32_t s_rows[256][17];   // 32 bf16 columns per pixel (+1 word of padding)
  pdl_launch_dependents();
  pdl_wait();
  const long long P = static_cast<long long>(p.N) * p.H * p.W;
  if (blockIdx.x * 256LL >= P) {
    pack_first_filters(k);
    return;
  }
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
  write_rows(s_rows, out, P);
}

// Tiled variant for images whose rows divide a CTA's 256 pixels (CIFAR: 8 rows of 32).  The CTA
// first stages the (rows+2) x (W+2) x 3 source patch -- cropped, flipped, normalised, conv
// padding included -- in shared memory with coalesced loads; a pixel's 27 values are then three
// runs of 9 consecutive floats, read with compile-time offsets (lane stride 3 words: no bank
// conflicts).  Every source byte is fetched once instead of nine times.
constexpr int kPatchMax = 2400;   // (256/W + 2) * (W + 2) * 3 floats for any W dividing 256

template <bool kUint8>
__global__ void __launch_bounds__(256)
first_layer_operand_tile_kernel(const void* __restrict__ src, const signed char* __restrict__ aug,
                                __nv_bfloat16* __restrict__ out, PrepParams p, PackArgs k) {
  __shared__ uint32_t s_rows[256][17];
  __shared__ float s_patch[kPatchMax];
  pdl_launch_dependents();
  pdl_wait();
  const long long P = static_cast<long long>(p.N) * p.H * p.W;
  if (blockIdx.x * 256LL >= P) {
    pack_first_filters(k);
    return;
  }
  const int rows = 256 / p.W;                       // image rows of this CTA
  const long long pix0 = blockIdx.x * 256LL;
  const int n = static_cast<int>(pix0 / (p.H * p.W));
  const int h0 = static_cast<int>((pix0 / p.W) % p.H);
  const int pw = (p.W + 2) * 3;                     // patch row pitch in floats
  const int total = (rows + 2) * pw;
  for (int i = threadIdx.x; i < total; i += 256) {
    const int pr = i / pw, rem = i - pr * pw;
    const int pc = rem / 3, c = rem - pc * 3;
    s_patch[i] = src_value<kUint8>(src, p, aug, n, h0 + pr - 1, pc - 1, c);
  }
  __syncthreads();
  {
    const int hl = threadIdx.x / p.W, w = threadIdx.x - hl * p.W;
    const float* base = s_patch + hl * pw + w * 3;  // patch (hl, w) = source (h-1, w-1)
    float v[28];
#pragma unroll
    for (int dy = 0; dy < 3; ++dy)
#pragma unroll
      for (int e = 0; e < 9; ++e) v[dy * 9 + e] = base[dy * pw + e];
    v[27] = 0.f;
#pragma unroll
    for (int j = 0; j < 14; ++j) s_rows[threadIdx.x][j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
    s_rows[threadIdx.x][14] = 0u;
    s_rows[threadIdx.x][15] = 0u;
  }
  __syncthreads();
  write_rows(s_rows, out, P);
}

static bool tile_shape(int H, int W) {
  return W >= 1 && W <= 256 && 256 % W == 0 && (static_cast<long long>(H) * W) % 256 == 0;
}

// one block per 256 pixels (not grid-stride) plus one for the filter pack when asked
static int prep_grid(long long pixels, const PackArgs& k) {
  long long blocks = (pixels + 255) / 256;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks) + (k.master != nullptr ? 1 : 0);
}

template <bool kUint8>
static cudaError_t launch_prep(const void* src, const signed char* aug, void* out, const PrepParams& p, const PackArgs& k,
                        cudaStream_t stream) {
  const int grid = prep_grid(static_cast<long long>(p.N) * p.H * p.W, k);
  if (tile_shape(p.H, p.W))
    return launch_pdl(first_layer_operand_tile_kernel<kUint8>, grid, 256, 0, stream, src, aug,
                      static_cast<__nv_bfloat16*>(out), p, k);
  return launch_pdl(first_layer_operand_kernel<kUint8>, grid, 256, 0, stream, src, aug,
                    static_cast<__nv_bfloat16*>(out), p, k);
}

}  // namespace ddpb

using namespace ddpb;

// src: fp32 NCHW [N][3][H][W], already normalised.  out: bf16 [N*H*W][64].
// w0_master / w0_packed (both or neither null): the first layer's fp32 filters [c_out0][27] and
// their zero-padded bf16 GEMM operand [c_out0][64], refreshed by the same launch.
extern "C" int ddpb_first_layer_operand_f32(const float* src, void* out, int N, int H, int W, const float* w0_master,
                                            void* w0_packed, int c_out0, cudaStream_t stream) {
  PrepParams p{};
  p.N = N;
  p.H = H;
  p.W = W;
  DDPB_REQUIRE((w0_master == nullptr) == (w0_packed == nullptr), "first_layer_operand: give both filter pointers or none");
  PackArgs k{w0_master, static_cast<__nv_bfloat16*>(w0_packed), c_out0};
  DDPB_CHECK_CUDA(launch_prep<false>(src, nullptr, out, p, k, stream));
  return 0;
}

// src: uint8 NHWC [N][H][W][3] raw pixels.  aug: int8 [N][4] = {crop_y, crop_x, flip, 0} with crop
// offsets in [0, 2*pad], or null for the evaluation transform (normalise only).
extern "C" int ddpb_first_layer_operand_u8(const unsigned char* src, const signed char* aug, void* out, int N,
                                           int H, int W, const float* mean3, const float* std3, int pad,
                                           const float* w0_master, void* w0_packed, int c_out0,
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
  DDPB_REQUIRE((w0_master == nullptr) == (w0_packed == nullptr), "first_layer_operand: give both filter pointers or none");
  PackArgs k{w0_master, static_cast<__nv_bfloat16*>(w0_packed), c_out0};
  DDPB_CHECK_CUDA(launch_prep<true>(src, aug, out, p, k, stream));
  return 0;
}

// ------------------------------------------------------------------------------------------
// Device-side random numbers for the input pipeline (SURVEY.md par. 8, K16): Philox4x32-10, a
// counter-based generator -- value = f(seed, step, image, word) -- so a captured step that reads its
// step number from a device counter draws fresh numbers on every graph replay without any host input.
// The reference draws on CPU workers (torchvision RandomCrop / RandomHorizontalFlip,
// /root/reference/part1/main.py:23-26).
// ------------------------------------------------------------------------------------------
namespace ddpb {

__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  constexpr uint32_t kM0 = 0xD2511F53u, kM1 = 0xCD9E8D57u, kW0 = 0x9E3779B9u, kW1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(kM0, ctr.x), lo0 = kM0 * ctr.x;
    const uint32_t hi1 = __umulhi(kM1, ctr.z), lo1 = kM1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += kW0;
    key.y += kW1;
  }
  return ctr;
}

// aug[n] = {crop offset y in [0, 2 pad], crop offset x in [0, 2 pad], flip in {0, 1}, 0}; one block, any N.
// *step is read by every thread first and bumped by one thread at the end: the next launch (or graph
// replay) draws the next set.
__global__ void __launch_bounds__(256)
draw_augmentation_kernel(signed char* __restrict__ aug, int N, int pad, unsigned long long seed,
                         unsigned int* __restrict__ step) {
  pdl_launch_dependents();
  pdl_wait();
  const unsigned int s = *step;
  __syncthreads();
  const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
  const uint32_t span = 2u * static_cast<uint32_t>(pad) + 1u;
  for (int n = threadIdx.x; n < N; n += blockDim.x) {
    const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(n), s, 0x41554721u /* "AUG!" stream */, 0u), key);
    char4 a;
    a.x = static_cast<signed char>(__umulhi(r.x, span));   // unbiased enough: span << 2^32
    a.y = static_cast<signed char>(__umulhi(r.y, span));
    a.z = static_cast<signed char>(r.z >> 31);
    a.w = 0;
    reinterpret_cast<char4*>(aug)[n] = a;
  }
  __syncthreads();
  if (threadIdx.x == 0) *step = s + 1u;
}

// Synthetic CIFAR-shaped batch generated on the device: label = uniform class, image = a smooth per-class
// colour template (a 4x4x3 grid drawn from the seed, bilinearly interpolated) + Gaussian noise of sigma 0.25,
// clamped to [0, 1] and quantised to uint8 -- the construction of ddpb200.data.synthetic_cifar10 with its own
// random stream.  One thread per pixel.
__global__ void __launch_bounds__(256)
synthetic_batch_kernel(unsigned char* __restrict__ img, long long* __restrict__ labels, int N, int H, int W,
                       int classes, unsigned long long seed, unsigned int* __restrict__ step) {
  const unsigned int s = *step;
  const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
  const long long pix = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (pix >= static_cast<long long>(N) * H * W) return;
  const int n = static_cast<int>(pix / (H * W)), hw = static_cast<int>(pix % (H * W));
  const int h = hw / W, w = hw % W;
  const uint4 rl = philox4x32_10(make_uint4(static_cast<uint32_t>(n), s, 0x4C41424Cu /* labels */, 0u), key);
  const int label = static_cast<int>(__umulhi(rl.x, static_cast<uint32_t>(classes)));
  if (hw == 0) labels[n] = label;
  // template value at (h, w): bilinear over a 4x4 grid whose nodes depend on (seed, class) only
  const float gy = (h + 0.5f) * 4.f / H - 0.5f, gx = (w + 0.5f) * 4.f / W - 0.5f;
  const int y0 = max(0, min(3, static_cast<int>(floorf(gy)))), x0 = max(0, min(3, static_cast<int>(floorf(gx))));
  const int y1 = min(3, y0 + 1), x1 = min(3, x0 + 1);
  const float fy = fminf(fmaxf(gy - y0, 0.f), 1.f), fx = fminf(fmaxf(gx - x0, 0.f), 1.f);
  auto node = [&](int y, int x) {   // three channel values in [0, 1)
    const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(label), static_cast<uint32_t>(y * 4 + x), 0x54454D50u, 0u), key);
    return make_float3(r.x * 2.3283064e-10f, r.y * 2.3283064e-10f, r.z * 2.3283064e-10f);
  };
  const float3 a = node(y0, x0), b = node(y0, x1), c = node(y1, x0), d = node(y1, x1);
  const uint4 rn = philox4x32_10(make_uint4(static_cast<uint32_t>(pix), static_cast<uint32_t>(pix >> 32) ^ s, 0x4E4F4953u, 1u), key);
  // Box-Muller: two uniforms -> two normals; the third channel takes one of a second pair
  const float u0 = (rn.x + 1.f) * 2.3283064e-10f, u1 = rn.y * 2.3283064e-10f;
  const float u2 = (rn.z + 1.f) * 2.3283064e-10f, u3 = rn.w * 2.3283064e-10f;
  const float m0 = sqrtf(-2.f * __logf(u0)), m1 = sqrtf(-2.f * __logf(u2));
  const float nz[3] = {m0 * __cosf(6.2831853f * u1), m0 * __sinf(6.2831853f * u1), m1 * __cosf(6.2831853f * u3)};
  const float t[3] = {(a.x * (1 - fx) + b.x * fx) * (1 - fy) + (c.x * (1 - fx) + d.x * fx) * fy,
                      (a.y * (1 - fx) + b.y * fx) * (1 - fy) + (c.y * (1 - fx) + d.y * fx) * fy,
                      (a.z * (1 - fx) + b.z * fx) * (1 - fy) + (c.z * (1 - fx) + d.z * fx) * fy};
#pragma unroll
  for (int ch = 0; ch < 3; ++ch) {
    const float v = fminf(fmaxf(t[ch] + 0.25f * nz[ch], 0.f), 1.f);
    img[pix * 3 + ch] = static_cast<unsigned char>(v * 255.f);
  }
}

__global__ void bump_counter_kernel(unsigned int* step) { *step += 1u; }

}  // namespace ddpb

// O16 random draws on the device.  aug: int8 [N][4] (crop y, crop x, flip, 0); step: device counter (starts
// wherever the caller set it) that this launch reads and increments.
extern "C" int ddpb_draw_augmentation(signed char* aug, int N, int pad, unsigned long long seed, unsigned int* step,
                                      cudaStream_t stream) {
  DDPB_REQUIRE(aug != nullptr && step != nullptr && N >= 1 && pad >= 0 && pad <= 60, "draw_augmentation: bad arguments");
  DDPB_CHECK_CUDA(launch_pdl(ddpb::draw_augmentation_kernel, 1, 256, 0, stream, aug, N, pad, seed, step));
  return 0;
}

// Synthetic uint8 NHWC batch + int64 labels generated on the device (no host data at all); increments *step.
extern "C" int ddpb_synthetic_batch(unsigned char* img, long long* labels, int N, int H, int W, int classes,
                                    unsigned long long seed, unsigned int* step, cudaStream_t stream) {
  DDPB_REQUIRE(img != nullptr && labels != nullptr && step != nullptr && N >= 1 && classes >= 1, "synthetic_batch: bad arguments");
  const long long pixels = static_cast<long long>(N) * H * W;
  ddpb::synthetic_batch_kernel<<<static_cast<int>((pixels + 255) / 256), 256, 0, stream>>>(img, labels, N, H, W, classes,
                                                                                           seed, step);
  DDPB_CHECK_CUDA(cudaGetLastError());
  ddpb::bump_counter_kernel<<<1, 1, 0, stream>>>(step);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
