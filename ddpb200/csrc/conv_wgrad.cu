// K6: convolution weight gradient (reference op O3, produced by autograd
// for the nn.Conv2d at /root/reference/part1/model.py:18-23) on tcgen05 tensor cores.
//
//   dW[k][tap][c] = sum over pixels p of dY[p][k] * X[p shifted by tap][c]
//
// GEMM view per filter tap: M = C_out (128 per CTA), N = C_in (64 or 128 per CTA),
// reduction over the N*H*W pixels.  Both operands are stored pixel-major with the
// channel contiguous, i.e. they are "MN-major" for this GEMM, which tcgen05.mma
// consumes directly (a_major = b_major = 1): no transposes are materialised.
//   A = dY  : 2-D TMA boxes of (64 channels, 64 pixels)
//   B = X   : 4-D TMA boxes (64 ch, W, bh, bn) shifted by the tap offset, halo zero-filled
// Each box is 64 rows of 128 bytes with SWIZZLE_128B; a 128-wide M (or N) is two boxes
// LBO = 8 KB apart, and 8-pixel groups are SBO = 1 KB apart.
//
// The pixel reduction is split over gridDim.y CTAs; partial tiles go to a scratch
// buffer [split][C_out][taps*C_in] that ddpb_wgrad_reduce sums in a fixed order
// (deterministic, unlike atomics).  With one split the tile is written straight to
// its final place (for the training engine: the flat gradient arena).
#include "common.cuh"
#include <stdlib.h>

namespace ddpb {

constexpr int kWgThreads = 192;
constexpr int kWgPix = 64;                    // pixels per pipeline stage
constexpr int kWgBoxBytes = kWgPix * 128;     // one (64 ch x 64 px) box = 8 KB

// kDeep = false: <= 97 KB of shared memory, two CTAs per SM share the tensor pipe (the ring of one CTA is
// latency bound: ~1.1 us TMA round trip over 3 stages).  kDeep = true, for grids of at most one CTA per SM
// (nobody to share with): the ring is twice as deep.
template <int BLOCK_N, bool kDeep = false>
struct WgradCfg {
  static constexpr int kStages = ((BLOCK_N >= 128) ? 3 : 4) * (kDeep ? 2 : 1);
  static constexpr int kABytes = 2 * kWgBoxBytes;                 // M = 128 channels of dY
  static constexpr int kBBytes = (BLOCK_N / 64) * kWgBoxBytes;    // N channels of X
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + 256;
};

struct WgradParams {
  float* out;          // final gradient or scratch (see out_split_stride)
  long long out_split_stride;  // elements between splits (0 when gridDim.y == 1)
  int out_ld;          // row stride of `out` in elements (taps * C_in)
  int P;               // pixels
  int C_in, C_out;
  int taps;
  int m_tiles, n_tiles;
  int bh, bn, hblocks;  // 64-pixel box geometry of X
  int dup_m;            // C_out == 64: both halves of M read the same 64 channels
  int rows_per_split;   // rows of `out` one split owns (the output tensor map stacks the splits along its rows)
};

template <int BLOCK_N, bool kDeep>
__global__ void __launch_bounds__(kWgThreads)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x,
                  const __grid_constant__ CUtensorMap tmap_out, const WgradParams p) {
  using Cfg = WgradCfg<BLOCK_N, kDeep>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1k(smem_raw);
  uint8_t* tiles = smem;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + Cfg::kStages;
  uint64_t* tmem_full_bar = bars + 2 * Cfg::kStages;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * Cfg::kStages + 1);

  const int warp = threadIdx.x >> 5;
  // blockIdx.x -> (tap, m_tile, n_tile)
  int t = blockIdx.x;
  const int n_tile = t % p.n_tiles;
  t /= p.n_tiles;
  const int m_tile = t % p.m_tiles;
  const int tap = t / p.m_tiles;
  const int split = blockIdx.y, splits = gridDim.y;
  const int pix_blocks = (p.P + kWgPix - 1) / kWgPix;
  const int pb_begin = static_cast<int>(static_cast<long long>(pix_blocks) * split / splits);
  const int pb_end = static_cast<int>(static_cast<long long>(pix_blocks) * (split + 1) / splits);
  const int num_k = pb_end - pb_begin;

  pdl_launch_dependents();
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmap_dy);
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_out);
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
  pdl_wait();   // the prologue above overlapped the previous kernel's tail
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    if (elect_one()) {
      int dr = 0, ds = 0;
      if (p.taps == 9) {
        dr = tap / 3 - 1;
        ds = tap % 3 - 1;
      }
      const int k0 = m_tile * 128;
      const int k1 = p.dup_m ? k0 : k0 + 64;
      int stage = 0;
      uint32_t phase = 0;
      for (int pb = pb_begin; pb < pb_end; ++pb) {
        const int n_img = (pb / p.hblocks) * p.bn;
        const int h0 = (pb % p.hblocks) * p.bh;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* a_dst = tiles + stage * Cfg::kStageBytes;
        uint8_t* b_dst = a_dst + Cfg::kABytes;
        mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
        tma_load_2d(a_dst, &tmap_dy, &full_bar[stage], k0, pb * kWgPix);
        tma_load_2d(a_dst + kWgBoxBytes, &tmap_dy, &full_bar[stage], k1, pb * kWgPix);
#pragma unroll
        for (int j = 0; j < BLOCK_N / 64; ++j)
          tma_load_4d(b_dst + j * kWgBoxBytes, &tmap_x, &full_bar[stage], n_tile * BLOCK_N + j * 64, ds,
                      h0 + dr, n_img);
        if (++stage == Cfg::kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(128, BLOCK_N, 1, 1);
      int stage = 0;
      uint32_t phase = 0;
      for (int k = 0; k < num_k; ++k) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(tiles + stage * Cfg::kStageBytes);
        const uint32_t b_addr = a_addr + Cfg::kABytes;
#pragma unroll
        for (int kk = 0; kk < kWgPix / 16; ++kk) {
          // 16 pixels = 16 rows of 128 bytes further down the box
          const uint64_t adesc = make_smem_desc_sw128(a_addr + kk * 16 * 128, kWgBoxBytes, 1024);
          const uint64_t bdesc = make_smem_desc_sw128(b_addr + kk * 16 * 128, kWgBoxBytes, 1024);
          umma_f16(tmem_base, adesc, bdesc, idesc, (k | kk) != 0);
        }
        umma_commit(&empty_bar[stage]);
        if (++stage == Cfg::kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma_commit(tmem_full_bar);
    }
  } else {
    // Epilogue through shared memory and TMA stores: a lane holds one output row (channel), so direct stores
    // were 32 lanes x 16 bytes to 32 different rows per instruction (partial lines; the same pattern cost the
    // fprop epilogue 2 us per 32-column chunk and warp).  The operand ring is idle once the accumulator is
    // complete: each warp stages 32 rows x 32 fp32 columns (one 128-byte swizzle row per lane, conflict free)
    // in two alternating 4 KB boxes of it and hands them to the TMA engine.
    const int q = warp & 3;
    const int r = static_cast<int>(lane_id());
    if (num_k > 0) {
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
    }
    const uint32_t my_boxes = smem_u32(tiles) + static_cast<uint32_t>(q) * 8192u;
    const int out_row = split * p.rows_per_split + m_tile * 128 + q * 32;
    const int out_col = tap * p.C_in + n_tile * BLOCK_N;
#pragma unroll 1
    for (int chunk = 0; chunk < BLOCK_N / 32; ++chunk) {
      uint32_t raw[32];
      if (num_k > 0) {
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + chunk * 32, raw);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) raw[j] = 0u;
      }
      const uint32_t box = my_boxes + (chunk & 1) * 4096u;
      if (lane_id() == 0) bulk_wait_read<1>();     // the store issued from this box two chunks ago has read it
      __syncwarp();
#pragma unroll
      for (int k = 0; k < 8; ++k)
        st_shared_v4(box + r * 128 + ((k ^ (r & 7)) << 4), raw[4 * k], raw[4 * k + 1], raw[4 * k + 2], raw[4 * k + 3]);
      fence_proxy_async();
      __syncwarp();
      if (lane_id() == 0) {
        // (C_out == 64: rows 64..127 of the tile are duplicates and land in the unused half of the split's scratch)
        tma_store_2d_saddr(&tmap_out, box, out_col + chunk * 32, out_row);
        bulk_commit();
      }
    }
    if (lane_id() == 0) bulk_wait<0>();     // shared memory must outlive the reads; the writes are done at kernel end
    __syncwarp();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BLOCK_N);
  }
}

// Second stage of the split reduction, fixed summation order (deterministic).
// Dense form: the partial tiles have exactly the output's shape -> flat float4 streaming sum.
__global__ void __launch_bounds__(256)
wgrad_reduce_dense_kernel(const float4* __restrict__ ws, float4* __restrict__ out, int splits, long long stride4,
                          long long n4) {
  pdl_launch_dependents();
  pdl_wait();
  const long long step = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4; i += step) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    int sp = 0;
    for (; sp + 4 <= splits; sp += 4) {   // four independent loads in flight
      const float4 a = ws[i + (sp + 0) * stride4], b = ws[i + (sp + 1) * stride4];
      const float4 c = ws[i + (sp + 2) * stride4], d = ws[i + (sp + 3) * stride4];
      acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
      acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w;
      acc.x += c.x; acc.y += c.y; acc.z += c.z; acc.w += c.w;
      acc.x += d.x; acc.y += d.y; acc.z += d.z; acc.w += d.w;
    }
    for (; sp < splits; ++sp) {
      const float4 a = ws[i + sp * stride4];
      acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w;
    }
    out[i] = acc;
  }
}

// Compact form (first layer: 27 of 64 columns, hundreds of splits): one block per output row,
// 32 columns x 8 split lanes, then a fixed-order combine of the 8 lanes.
__global__ void __launch_bounds__(256)
wgrad_reduce_rows_kernel(const float* __restrict__ ws, float* __restrict__ out, int splits, long long split_stride,
                         int cols_valid, int ws_ld, int out_ld) {
  __shared__ float s_part[8][33];
  pdl_launch_dependents();
  pdl_wait();
  const int r = blockIdx.x;
  const int lane_s = threadIdx.x >> 5;
  for (int c0 = 0; c0 < cols_valid; c0 += 32) {
    const int c = c0 + (threadIdx.x & 31);
    float acc = 0.f;
    if (c < cols_valid) {
      const float* src = ws + static_cast<long long>(r) * ws_ld + c;
      for (int sp = lane_s; sp < splits; sp += 8) acc += src[sp * split_stride];
    }
    s_part[lane_s][threadIdx.x & 31] = acc;
    __syncthreads();
    if (lane_s == 0 && c < cols_valid) {
      float t = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) t += s_part[j][threadIdx.x & 31];
      out[static_cast<long long>(r) * out_ld + c] = t;
    }
    __syncthreads();
  }
}

template <int BLOCK_N, bool kDeep>
static int launch_wgrad_cfg(const CUtensorMap& tdy, const CUtensorMap& tx, const CUtensorMap& to, const WgradParams& p,
                            int splits, cudaStream_t stream) {
  using Cfg = WgradCfg<BLOCK_N, kDeep>;
  static bool configured = false;
  if (!configured) {
    DDPB_CHECK_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel<BLOCK_N, kDeep>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes));
    configured = true;
  }
  dim3 grid(p.taps * p.m_tiles * p.n_tiles, splits);
  DDPB_CHECK_CUDA(launch_pdl(conv_wgrad_kernel<BLOCK_N, kDeep>, grid, kWgThreads, Cfg::kSmemBytes, stream, tdy, tx, to, p));
  return 0;
}

// Deep ring for single-wave grids: 159 -> 147 us for the eight layers, 0.681 -> 0.669 ms per step
// (profiles/r2_call22_*).  Round 1 had measured it slower inside the step, because a 193 KB CTA kept the
// then-cooperative BatchNorm backward from starting; that kernel is an ordinary launch now.
static int g_wgrad_deep = -1;   // -1: environment DDPB_WGRAD_DEEP (default on), 0 / 1

template <int BLOCK_N>
static int launch_wgrad(const CUtensorMap& tdy, const CUtensorMap& tx, const CUtensorMap& to, const WgradParams& p,
                        int splits, cudaStream_t stream) {
  if (g_wgrad_deep < 0) {
    const char* e = getenv("DDPB_WGRAD_DEEP");
    g_wgrad_deep = e ? atoi(e) : 1;
  }
  const long long ctas = static_cast<long long>(p.taps) * p.m_tiles * p.n_tiles * splits;
  if (g_wgrad_deep && ctas <= sm_count()) return launch_wgrad_cfg<BLOCK_N, true>(tdy, tx, to, p, splits, stream);
  return launch_wgrad_cfg<BLOCK_N, false>(tdy, tx, to, p, splits, stream);
}

// Launch plan: the one-tap kernel with `groups` pixel splits.
//
// Tried and removed (measured slower at the training shapes, batch 256; the logs stay under profiles/):
//   * three taps per CTA with the split reduction over distributed shared memory inside a cluster:
//     187-197 us for the eight layers against 159 us (r1_call43_wgrad_tap_fused.txt);
//   * a cta_group::2 kernel (256 output channels x 128/256 input channels per CTA pair): 197 us against
//     160 us (r2_call01_pair_modes_layer_times.txt) -- it doubles the partial-tile volume.
struct WgPlan {
  int block_n;
  int groups;
};

static int legacy_splits(int P, int tiles) {
  const int pix_blocks = (P + kWgPix - 1) / kWgPix;
  // two CTAs fit per SM: aim for ONE full wave (rounding down avoids a nearly empty second wave)
  int target = 2 * sm_count();
  if (target <= 0) target = 296;
  int splits = target / tiles;
  // keep at least 8 pipeline steps (512 pixels) per CTA so the prologue amortises
  int max_splits = pix_blocks / 8;
  if (max_splits < 1) max_splits = 1;
  if (splits > max_splits) splits = max_splits;
  if (splits > 256) splits = 256;
  if (splits < 1) splits = 1;
  // with about one tile per SM already, write dw directly instead of splitting in two: measured
  // 20 us per step faster for the three 512-channel layers (profiles/r1_call37_*)
  if (splits == 2 && tiles >= sm_count() - 8) splits = 1;
  return splits;
}

static WgPlan plan_wgrad(int N, int H, int W, int C_in, int C_out, int taps) {
  WgPlan pl;
  pl.block_n = (C_in % 128 == 0) ? 128 : 64;
  const int m_tiles = (C_out + 127) / 128;
  const int n_tiles = C_in / pl.block_n;
  pl.groups = legacy_splits(N * H * W, taps * m_tiles * n_tiles);
  return pl;
}

}  // namespace ddpb

using namespace ddpb;

// Number of pixel splits the launcher will use for this shape (so callers can size scratch).
extern "C" int ddpb_conv_wgrad_splits(int N, int H, int W, int C_in, int C_out, int taps) {
  return plan_wgrad(N, H, W, C_in, C_out, taps).groups;
}

// Floats of scratch a launch with this plan writes.  `dense`: dw takes every column with no row padding
// (dw_cols_valid == dw_ld == taps * C_in); any other dw layout goes through scratch even with one split.
static long long wgrad_scratch_need(const WgPlan& pl, int C_in, int C_out, int taps, bool dense) {
  const long long rows = (C_out < 128) ? 128 : C_out;   // dup_m tiles still own 128 rows of scratch
  const bool direct = pl.groups == 1 && C_out >= 128 && dense;
  return direct ? 0 : static_cast<long long>(pl.groups) * rows * taps * C_in;
}

// Scratch floats needed by ddpb_conv_wgrad for this shape with a dense dw (0 if it writes dw directly).
extern "C" long long ddpb_conv_wgrad_scratch_elems(int N, int H, int W, int C_in, int C_out, int taps) {
  return wgrad_scratch_need(plan_wgrad(N, H, W, C_in, C_out, taps), C_in, C_out, taps, true);
}

// Same for an arbitrary dw layout: a dw with padded rows (dw_ld > taps * C_in) or fewer valid columns
// is always assembled from scratch, also when the reduction is not split.
extern "C" long long ddpb_conv_wgrad_scratch_elems_ld(int N, int H, int W, int C_in, int C_out, int taps,
                                                      int dw_cols_valid, int dw_ld) {
  const bool dense = dw_cols_valid == taps * C_in && dw_ld == taps * C_in;
  return wgrad_scratch_need(plan_wgrad(N, H, W, C_in, C_out, taps), C_in, C_out, taps, dense);
}

// Shared implementation: `scratch` may be any device buffer of at least
// ddpb_conv_wgrad_scratch_elems floats.
static int wgrad_impl(const void* dy, const void* x, float* dw, float* scratch, long long scratch_elems, int N, int H,
                      int W, int C_in, int C_out, int taps, int dw_cols_valid, int dw_ld, cudaStream_t stream) {
  DDPB_REQUIRE(taps == 1 || taps == 9, "conv_wgrad: taps must be 1 or 9 (got %d)", taps);
  DDPB_REQUIRE(C_in % 64 == 0 && C_in > 0, "conv_wgrad: C_in must be a multiple of 64 (got %d)", C_in);
  DDPB_REQUIRE(C_out == 64 || C_out % 128 == 0, "conv_wgrad: C_out must be 64 or a multiple of 128 (got %d)",
               C_out);
  int bh, bn, hblocks;
  DDPB_REQUIRE(pixel_box(H, W, kWgPix, &bh, &bn, &hblocks),
               "conv_wgrad: unsupported spatial size %dx%d for 64-pixel boxes", H, W);
  const WgPlan pl = plan_wgrad(N, H, W, C_in, C_out, taps);
  const int block_n = pl.block_n;
  const int splits = pl.groups;   // partial tiles in global memory
  const bool direct = (splits == 1 && C_out >= 128 && dw_cols_valid == taps * C_in && dw_ld == taps * C_in);
  DDPB_REQUIRE(dw_cols_valid > 0 && dw_cols_valid <= taps * C_in && dw_ld >= dw_cols_valid,
               "conv_wgrad: bad dw layout (cols_valid %d, ld %d, taps*C_in %d)", dw_cols_valid, dw_ld, taps * C_in);
  {
    // what this launch will really write: also one split goes through scratch when dw is not dense
    const long long rows_need = (C_out < 128) ? 128 : C_out;
    const long long need = direct ? 0 : static_cast<long long>(splits) * rows_need * taps * C_in;
    DDPB_REQUIRE(direct || scratch != nullptr, "conv_wgrad: this shape needs %lld floats of scratch memory", need);
    DDPB_REQUIRE(need <= scratch_elems || direct,
                 "conv_wgrad: this shape and dw layout need %lld floats of scratch but the buffer holds %lld", need,
                 scratch_elems);
  }

  CUtensorMap tdy, tx;
  {
    uint64_t dims[2] = {(uint64_t)C_out, (uint64_t)N * H * W};
    uint64_t strides[1] = {(uint64_t)C_out * 2};
    uint32_t box[2] = {64, (uint32_t)kWgPix};
    if (encode_tmap(&tdy, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dy, dims, strides, box,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B))
      return 3;
  }
  {
    uint64_t dims[4] = {(uint64_t)C_in, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    uint64_t strides[3] = {(uint64_t)C_in * 2, (uint64_t)W * C_in * 2, (uint64_t)H * W * C_in * 2};
    uint32_t box[4] = {64, (uint32_t)W, (uint32_t)bh, (uint32_t)bn};
    if (encode_tmap(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x, dims, strides, box,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B))
      return 3;
  }
  WgradParams p;
  p.P = N * H * W;
  p.C_in = C_in;
  p.C_out = C_out;
  p.taps = taps;
  p.m_tiles = (C_out + 127) / 128;
  p.n_tiles = C_in / block_n;
  p.bh = bh;
  p.bn = bn;
  p.hblocks = hblocks;
  p.dup_m = (C_out == 64);
  p.out_ld = taps * C_in;
  const long long rows = (C_out < 128) ? 128 : C_out;
  if (direct) {
    p.out = dw;
    p.out_ld = dw_ld;
    p.out_split_stride = 0;
  } else {
    p.out = scratch;
    p.out_split_stride = rows * taps * C_in;
  }
  p.rows_per_split = direct ? 0 : static_cast<int>(rows);
  // output (dw itself, or the splits' partial tiles stacked along the rows): fp32 [rows][taps * C_in], written in
  // boxes of 32 rows x 32 columns (128-byte swizzle rows)
  CUtensorMap to;
  {
    DDPB_REQUIRE((reinterpret_cast<uintptr_t>(p.out) & 15u) == 0, "conv_wgrad: dw / scratch must be 16-byte aligned");
    uint64_t dims[2] = {(uint64_t)taps * C_in, direct ? (uint64_t)C_out : (uint64_t)splits * rows};
    uint64_t strides[1] = {(uint64_t)taps * C_in * 4};
    uint32_t box[2] = {32, 32};
    if (encode_tmap(&to, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, p.out, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_128B))
      return 3;
  }
  const int rc = (block_n == 128) ? launch_wgrad<128>(tdy, tx, to, p, splits, stream)
                                  : launch_wgrad<64>(tdy, tx, to, p, splits, stream);
  if (rc) return rc;
  if (!direct) {
    const bool dense = (dw_cols_valid == taps * C_in && dw_ld == taps * C_in && C_out >= 128);
    if (dense) {
      const long long n4 = static_cast<long long>(C_out) * taps * C_in / 4;
      long long blocks = (n4 + 255) / 256;
      if (blocks > 8LL * sm_count()) blocks = 8LL * sm_count();
      DDPB_CHECK_CUDA(launch_pdl(wgrad_reduce_dense_kernel, static_cast<int>(blocks), 256, 0, stream,
                                 reinterpret_cast<const float4*>(scratch), reinterpret_cast<float4*>(dw), splits,
                                 static_cast<long long>(p.out_split_stride / 4), n4));
    } else {
      DDPB_CHECK_CUDA(launch_pdl(wgrad_reduce_rows_kernel, C_out, 256, 0, stream, static_cast<const float*>(scratch),
                                 dw, splits, static_cast<long long>(p.out_split_stride), dw_cols_valid, taps * C_in,
                                 dw_ld));
    }
  }
  return 0;
}

// dy: [N*H*W][C_out] bf16; x: [N][H][W][C_in] bf16; dw: fp32 [C_out][dw_ld], of which the
// first dw_cols_valid (<= taps*C_in) columns of every row are written.  scratch: device
// buffer of scratch_elems floats; ddpb_conv_wgrad_scratch_elems_ld() says how many this shape
// and dw layout need (null / 0 allowed when that is 0).  A buffer that is too small is an error,
// never an out-of-bounds write.
extern "C" int ddpb_conv_wgrad(const void* dy, const void* x, float* dw, float* scratch, long long scratch_elems,
                               int N, int H, int W, int C_in, int C_out, int taps, int dw_cols_valid, int dw_ld,
                               cudaStream_t stream) {
  return wgrad_impl(dy, x, dw, scratch, scratch != nullptr ? scratch_elems : 0, N, H, W, C_in, C_out, taps,
                    dw_cols_valid, dw_ld, stream);
}

// ------------------------------------------------------------------------------------------
// ddpb_conv_wgrad_static: the same weight-gradient computation for callers that may not
// allocate device memory.  (No counterpart in the reference.)
//
// Who needs this: code that runs where a cudaMalloc is off limits -- for example a training
// or serving loop that must never hit cudaMalloc's implicit device-wide synchronisation, a
// process whose GPU memory was handed over in full to another allocator at start-up so that a
// later driver allocation could fail, or a host application that embeds this library and
// forbids it to touch the allocator.  ddpb_conv_wgrad makes such callers provide a scratch
// buffer of a shape-dependent size; here the split-reduction scratch is a fixed array that
// lives in the library image instead, so the caller passes only inputs and outputs.
//
// Cost and limits:
//   * the array (kWgradStaticMaxFloats floats = 16 MiB) is resident on EVERY GPU that loads
//     the library, from load to process exit, whether or not this entry point is ever used;
//   * shapes whose split reduction needs more than that are rejected with an error -- use
//     ddpb_conv_wgrad with your own buffer for those.  On a 148-SM B200 the planner never asks
//     for more than ~4.85 M floats; 4 Mi covers e.g. VGG-11's layers 0-2, 4 and 5 at batch 256
//     and every layer at batch <= 128, but not the 4.7 M-float cases (layers 3, 6, 7 at 256);
//   * the first call on each device looks up the array's device address
//     (cudaGetSymbolAddress) and caches it; make that first call with the right device
//     current, before any stream capture begins, and from ONE thread (the cache is written
//     without a lock; later calls may come from any thread);
//   * all calls share the one array and it keeps whatever the previous call left in it.
//     The kernel overwrites every partial sum that the second stage reads, so nothing has to be
//     cleared between calls -- but two calls whose kernels overlap on the GPU (different
//     streams, or one device used from two threads) race on it.  Serialise them.
// It performs no cudaMalloc / cudaFree and calls no allocating library routine.
// ------------------------------------------------------------------------------------------
constexpr long long kWgradStaticMaxFloats = 4LL * 1024 * 1024;

static __device__ __align__(256) float g_wgrad_static_scratch[kWgradStaticMaxFloats];

// In host code `g_wgrad_static_scratch` is only a handle for the runtime, not a device
// address; the address a kernel can use has to be asked for, once per device.
static int wgrad_static_scratch(float** out) {
  constexpr int kMaxDevices = 64;
  static float* per_device[kMaxDevices] = {};
  int dev = -1;
  DDPB_CHECK_CUDA(cudaGetDevice(&dev));
  DDPB_REQUIRE(dev >= 0 && dev < kMaxDevices, "conv_wgrad_static: device ordinal %d out of range", dev);
  if (per_device[dev] == nullptr) {
    void* ptr = nullptr;
    DDPB_CHECK_CUDA(cudaGetSymbolAddress(&ptr, g_wgrad_static_scratch));
    DDPB_REQUIRE(ptr != nullptr, "conv_wgrad_static: no device address for the built-in scratch array");
    per_device[dev] = static_cast<float*>(ptr);
  }
  *out = per_device[dev];
  return 0;
}

// Capacity of the built-in scratch array in floats.
extern "C" long long ddpb_conv_wgrad_static_max_elems() { return kWgradStaticMaxFloats; }

// Same arguments and result as ddpb_conv_wgrad minus `scratch`.
extern "C" int ddpb_conv_wgrad_static(const void* dy, const void* x, float* dw, int N, int H, int W, int C_in,
                                      int C_out, int taps, int dw_cols_valid, int dw_ld, cudaStream_t stream) {
  const long long need = ddpb_conv_wgrad_scratch_elems_ld(N, H, W, C_in, C_out, taps, dw_cols_valid, dw_ld);
  DDPB_REQUIRE(need <= kWgradStaticMaxFloats,
               "conv_wgrad_static: this shape needs %lld floats of scratch but the built-in array holds %lld; "
               "call ddpb_conv_wgrad with a caller-provided buffer instead",
               need, kWgradStaticMaxFloats);
  float* scratch = nullptr;
  if (int rc = wgrad_static_scratch(&scratch)) return rc;
  return wgrad_impl(dy, x, dw, scratch, kWgradStaticMaxFloats, N, H, W, C_in, C_out, taps, dw_cols_valid, dw_ld,
                    stream);
}
