// K1/K2/K5: 3x3 stride-1 pad-1 convolution as an implicit GEMM on the
// 5th-generation tensor cores.  One kernel serves
//   * fprop  (reference op O1: nn.Conv2d built at /root/reference/part1/model.py:18-23),
//   * dgrad  (O2, autograd of the same module) by feeding dY and the rotated/transposed
//            filter bank, and
//   * the first layer (C_in = 3) as a 1-tap GEMM over an im2col'ed input (taps == 1).
//
// Layout: activations NHWC bf16, filters [C_out][taps][C_in] bf16 (K-major), fp32
// accumulation in TMEM.  GEMM view: M = N*H*W output pixels (128 per CTA),
// N = C_out (BLOCK_N per CTA), K = taps*C_in in chunks of 64.
//
// The A operand never materialises an im2col matrix: a 4-D TMA box
// (64 ch, W, bh, bn) whose origin is shifted by the tap offset lands in shared
// memory as 128 rows x 128 bytes, already in the SWIZZLE_128B K-major layout
// tcgen05.mma wants, and TMA zero-fills the halo.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA
// issuer (one elected lane), warps 2-5 = epilogue (TMEM -> registers -> global),
// which also adds the bias and accumulates the per-channel sum / sum of squares
// that BatchNorm needs (O4), so the conv output is not re-read for statistics.
#include "common.cuh"

namespace ddpb {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;  // bf16 elements = one 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kConvThreads = 192;

template <int BLOCK_N>
struct FpropCfg {
  static constexpr int kStages = (BLOCK_N >= 256) ? 4 : 3;
  static constexpr int kABytes = kBlockM * kBlockK * 2;   // 16 KB
  static constexpr int kBBytes = BLOCK_N * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStatsBytes = 4 * 2 * BLOCK_N * 4;
  // 1024 B of slack so the tile ring can be aligned by hand
  static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kStatsBytes + 256;
};

struct FpropParams {
  __nv_bfloat16* out;    // [P][C_out]
  const float* bias;     // [C_out] or nullptr
  float* stats;          // [2][C_out] (sum, sum of squares) or nullptr
  int P;                 // N*H*W
  int C_in, C_out;
  int taps;              // 9 (3x3) or 1 (pre-gathered input)
  int bh, bn, hblocks;   // TMA box geometry: box = (64, W, bh, bn); hblocks = H / bh
};

template <int BLOCK_N>
__global__ void __launch_bounds__(kConvThreads)
conv_fprop_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                  const FpropParams p) {
  using Cfg = FpropCfg<BLOCK_N>;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operand tiles need 1024-byte alignment.
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* tiles = smem;
  float* s_stats = reinterpret_cast<float*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes + Cfg::kStatsBytes);
  uint64_t* full_bar = bars;                       // [kStages]
  uint64_t* empty_bar = bars + Cfg::kStages;       // [kStages]
  uint64_t* tmem_full_bar = bars + 2 * Cfg::kStages;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * Cfg::kStages + 1);

  const int warp = threadIdx.x >> 5;
  const int m_tile = blockIdx.x;
  const int n_tile = blockIdx.y;
  const int cchunks = p.C_in / kBlockK;
  const int num_k = p.taps * cchunks;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_w);
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int i = 0; i < Cfg::kStages; ++i) {
        mbar_init(&full_bar[i], 1);
        mbar_init(&empty_bar[i], 1);
      }
      mbar_init(tmem_full_bar, 1);
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_ptr_smem, BLOCK_N);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      const int n_img = (m_tile / p.hblocks) * p.bn;
      const int h0 = (m_tile % p.hblocks) * p.bh;
      int stage = 0;
      uint32_t phase = 0;
      for (int k = 0; k < num_k; ++k) {
        const int tap = k / cchunks;
        const int cc = k - tap * cchunks;
        int dr = 0, ds = 0;
        if (p.taps == 9) {
          dr = tap / 3 - 1;
          ds = tap % 3 - 1;
        }
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* a_dst = tiles + stage * Cfg::kStageBytes;
        uint8_t* b_dst = a_dst + Cfg::kABytes;
        mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
        tma_load_4d(a_dst, &tmap_x, &full_bar[stage], cc * kBlockK, ds, h0 + dr, n_img);
        tma_load_2d(b_dst, &tmap_w, &full_bar[stage], tap * p.C_in + cc * kBlockK, n_tile * BLOCK_N);
        if (++stage == Cfg::kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(kBlockM, BLOCK_N, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      for (int k = 0; k < num_k; ++k) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(tiles + stage * Cfg::kStageBytes);
        const uint32_t b_addr = a_addr + Cfg::kABytes;
#pragma unroll
        for (int kk = 0; kk < kBlockK / kUmmaK; ++kk) {
          const uint64_t adesc = make_smem_desc_sw128(a_addr + kk * kUmmaK * 2, 16, 1024);
          const uint64_t bdesc = make_smem_desc_sw128(b_addr + kk * kUmmaK * 2, 16, 1024);
          umma_f16(tmem_base, adesc, bdesc, idesc, (k | kk) != 0);
        }
        umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs retire
        if (++stage == Cfg::kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma_commit(tmem_full_bar);  // accumulator complete
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    const int row = q * 32 + lane_id();
    const long long pix = static_cast<long long>(m_tile) * kBlockM + row;
    const bool valid = pix < p.P;
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    __nv_bfloat16* out_row = p.out + pix * p.C_out + n_tile * BLOCK_N;
#pragma unroll 1
    for (int chunk = 0; chunk < BLOCK_N / 32; ++chunk) {
      uint32_t raw[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + chunk * 32, raw);
      tmem_ld_wait();
      float v[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float x = __uint_as_float(raw[j]);
        if (p.bias != nullptr) x += __ldg(p.bias + n_tile * BLOCK_N + chunk * 32 + j);
        v[j] = valid ? bf16_round(x) : 0.f;
      }
      if (valid) {
        uint4* dst = reinterpret_cast<uint4*>(out_row + chunk * 32);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 o;
          o.x = pack_bf16x2(v[8 * j + 0], v[8 * j + 1]);
          o.y = pack_bf16x2(v[8 * j + 2], v[8 * j + 3]);
          o.z = pack_bf16x2(v[8 * j + 4], v[8 * j + 5]);
          o.w = pack_bf16x2(v[8 * j + 6], v[8 * j + 7]);
          dst[j] = o;
        }
      }
      if (p.stats != nullptr) {
        float sq[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) sq[j] = v[j] * v[j];
        const float s1 = warp_transpose_reduce(v);
        const float s2 = warp_transpose_reduce(sq);
        s_stats[(q * 2 + 0) * BLOCK_N + chunk * 32 + lane_id()] = s1;
        s_stats[(q * 2 + 1) * BLOCK_N + chunk * 32 + lane_id()] = s2;
      }
    }
    if (p.stats != nullptr) {
      // only the four epilogue warps meet here (named barrier 1, 128 threads)
      asm volatile("bar.sync 1, 128;" ::: "memory");
      const int t = threadIdx.x - 64;
      for (int c = t; c < 2 * BLOCK_N; c += 128) {
        const int which = c / BLOCK_N, col = c - which * BLOCK_N;
        float acc = 0.f;
#pragma unroll
        for (int w = 0; w < 4; ++w) acc += s_stats[(w * 2 + which) * BLOCK_N + col];
        atomicAdd(p.stats + which * p.C_out + n_tile * BLOCK_N + col, acc);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BLOCK_N);
  }
}

template <int BLOCK_N>
static int launch_fprop(const CUtensorMap& tx, const CUtensorMap& tw, const FpropParams& p, int m_tiles,
                        cudaStream_t stream) {
  using Cfg = FpropCfg<BLOCK_N>;
  static bool configured = false;
  if (!configured) {
    DDPB_CHECK_CUDA(cudaFuncSetAttribute(conv_fprop_kernel<BLOCK_N>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    configured = true;
  }
  dim3 grid(m_tiles, p.C_out / BLOCK_N);
  conv_fprop_kernel<BLOCK_N><<<grid, kConvThreads, Cfg::kSmemBytes, stream>>>(tx, tw, p);
  DDPB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace ddpb

using namespace ddpb;

// x: [N][H][W][C_in] bf16; w: [C_out][taps*C_in] bf16; y: [N*H*W][C_out] bf16.
// bias (fp32, may be null) is added before rounding; stats (fp32 [2][C_out], may be
// null) is ACCUMULATED into (caller zeroes it).  block_n: 0 = choose automatically.
extern "C" int ddpb_conv_fprop(const void* x, const void* w, const float* bias, void* y, float* stats,
                               int N, int H, int W, int C_in, int C_out, int taps, int block_n,
                               cudaStream_t stream) {
  DDPB_REQUIRE(taps == 1 || taps == 9, "conv_fprop: taps must be 1 or 9 (got %d)", taps);
  DDPB_REQUIRE(C_in % 64 == 0 && C_in > 0, "conv_fprop: C_in must be a multiple of 64 (got %d)", C_in);
  DDPB_REQUIRE(C_out % 64 == 0 && C_out > 0, "conv_fprop: C_out must be a multiple of 64 (got %d)", C_out);
  int bh, bn, hblocks;
  DDPB_REQUIRE(pixel_box(H, W, kBlockM, &bh, &bn, &hblocks),
               "conv_fprop: unsupported spatial size %dx%d for 128-pixel tiles", H, W);
  if (block_n == 0) block_n = (C_out % 128 == 0) ? 128 : 64;
  DDPB_REQUIRE(block_n == 64 || block_n == 128 || block_n == 256, "conv_fprop: bad block_n %d", block_n);
  DDPB_REQUIRE(C_out % block_n == 0, "conv_fprop: C_out %d not divisible by block_n %d", C_out, block_n);

  CUtensorMap tx, tw;
  {
    uint64_t dims[4] = {(uint64_t)C_in, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    uint64_t strides[3] = {(uint64_t)C_in * 2, (uint64_t)W * C_in * 2, (uint64_t)H * W * C_in * 2};
    uint32_t box[4] = {64, (uint32_t)W, (uint32_t)bh, (uint32_t)bn};
    if (encode_tmap(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x, dims, strides, box,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B))
      return 3;
  }
  {
    uint64_t dims[2] = {(uint64_t)taps * C_in, (uint64_t)C_out};
    uint64_t strides[1] = {(uint64_t)taps * C_in * 2};
    uint32_t box[2] = {64, (uint32_t)block_n};
    if (encode_tmap(&tw, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, w, dims, strides, box,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B))
      return 3;
  }
  FpropParams p;
  p.out = static_cast<__nv_bfloat16*>(y);
  p.bias = bias;
  p.stats = stats;
  p.P = N * H * W;
  p.C_in = C_in;
  p.C_out = C_out;
  p.taps = taps;
  p.bh = bh;
  p.bn = bn;
  p.hblocks = hblocks;
  const int m_tiles = (p.P + kBlockM - 1) / kBlockM;
  switch (block_n) {
    case 64: return launch_fprop<64>(tx, tw, p, m_tiles, stream);
    case 128: return launch_fprop<128>(tx, tw, p, m_tiles, stream);
    default: return launch_fprop<256>(tx, tw, p, m_tiles, stream);
  }
}
