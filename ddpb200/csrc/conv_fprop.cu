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
#include <stdlib.h>
#include <cstdlib>
#include <type_traits>

namespace ddpb {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;  // bf16 elements = one 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kConvThreads = 192;

template <int BLOCK_N>
struct FpropCfg {
#ifdef DDPB_FPROP_STAGES
  static constexpr int kStages = DDPB_FPROP_STAGES;
#else
  static constexpr int kStages = (BLOCK_N >= 256) ? 4 : 3;
#endif
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
  float* partial;        // split-K: fp32 [gridDim.z][P][C_out] partial sums (null when gridDim.z == 1)
  unsigned long long* dbg;  // optional per-CTA timestamps, 8 slots per CTA: {start, after set-up, main loop done,
                            // epilogue done / exit, [persistent kernels: 4 = epilogues done, 5 = past the grid barrier]}
  int dbg_mode;             // timing probes (results are WRONG): bit 0 = skip the B loads, bit 1 = skip the A loads
                            // once the ring is primed
  // fused BatchNorm + ReLU (+2x2 max-pool) tail of the persistent kernel (act == null: off)
  const float* gamma;
  const float* beta;
  __nv_bfloat16* act;    // [P or P/4][C_out]
  float* saved;          // [2][C_out] mean, rstd for the backward pass
  unsigned int* grid_sync;  // one zeroed word: all CTAs meet here once the statistics are complete
  int H, W, pool;
  float eps;
  // Split-K summed inside the persistent kernel (split_sync != null): every epilogue warp takes a ticket per
  // (tile, warp); the warp that draws the last ticket waits for the other splits' slices, adds them to its own
  // accumulator in split order (so the result does not depend on who was last) and runs the normal epilogue.
  // Two words per (tile, warp): tickets drawn, slices written; the last warp leaves both at zero.
  unsigned int* split_sync;
  int tail_tmem;         // persistent kernel: the BatchNorm tail re-reads the accumulators from TMEM (every tile of a CTA
                         // still has its own accumulator buffer) instead of the bf16 tiles from L2
};

__device__ __forceinline__ void dbg_stamp(const FpropParams& p, int slot) {
  if (p.dbg != nullptr) {
    const long long cta = blockIdx.x + gridDim.x * (blockIdx.y + static_cast<long long>(gridDim.y) * blockIdx.z);
    p.dbg[cta * 8 + slot] = global_timer_ns();
  }
}

// Epilogue shared by the 1-CTA and CTA-pair kernels: this warp's 32 accumulator rows ->
// (+bias, bf16, per-channel statistics) -> global, or raw fp32 partials under split-K.
template <int BLOCK_N>
__device__ __forceinline__ void conv_epilogue(const FpropParams& p, uint32_t tmem_base, uint64_t* tmem_full_bar,
                                              float* s_stats, int warp, int m_tile, int n_tile) {
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    const int row = q * 32 + lane_id();
    const long long pix = static_cast<long long>(m_tile) * kBlockM + row;
    const bool valid = pix < p.P;
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    __nv_bfloat16* out_row = p.out + pix * p.C_out + n_tile * BLOCK_N;
    if (p.partial != nullptr) {
      // split-K: hand the raw fp32 tile to the finalize kernel (bias, rounding, statistics there)
      float* prow = p.partial + (static_cast<long long>(blockIdx.z) * p.P + pix) * p.C_out + n_tile * BLOCK_N;
#pragma unroll 1
      for (int chunk = 0; chunk < BLOCK_N / 32; ++chunk) {
        uint32_t raw[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + chunk * 32, raw);
        tmem_ld_wait();
        if (valid) {
          uint4* dst = reinterpret_cast<uint4*>(prow + chunk * 32);
#pragma unroll
          for (int j = 0; j < 8; ++j) dst[j] = make_uint4(raw[4 * j], raw[4 * j + 1], raw[4 * j + 2], raw[4 * j + 3]);
        }
      }
    } else {
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
    }  // direct (non split-K) epilogue
  }

// kDgrad: the B operand is read straight from the FPROP filter pack [C_out][taps][C_in]
// as an MN-major tile (rows = reduction channel k, 64 output channels c contiguous), with the
// taps visited in reverse -- dgrad needs no transposed / rotated copy of the filters.
// In that mode p.C_in is the reduction width (the layer's C_out) and p.C_out the produced
// width (the layer's C_in).
template <int BLOCK_N, bool kDgrad>
__global__ void __launch_bounds__(kConvThreads)
conv_fprop_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                  const FpropParams p) {
  using Cfg = FpropCfg<BLOCK_N>;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operand tiles need 1024-byte alignment.
  uint8_t* smem = align_smem_1k(smem_raw);
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
  // split-K: blockIdx.z owns k-steps [k_begin, k_end) of the taps*C_in/64 reduction steps
  const int total_k = p.taps * cchunks;
  const int k_begin = static_cast<int>(static_cast<long long>(total_k) * blockIdx.z / gridDim.z);
  const int k_end = static_cast<int>(static_cast<long long>(total_k) * (blockIdx.z + 1) / gridDim.z);
  const int num_k = k_end - k_begin;

  pdl_launch_dependents();
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
  if (threadIdx.x == 0) dbg_stamp(p, 0);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();   // everything above overlapped the previous kernel's tail; global memory from here on
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (threadIdx.x == 0) dbg_stamp(p, 1);

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      const int n_img = (m_tile / p.hblocks) * p.bn;
      const int h0 = (m_tile % p.hblocks) * p.bh;
      int stage = 0;
      uint32_t phase = 0;
      for (int k = k_begin; k < k_end; ++k) {
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
#if defined(DDPB_DBG_SKIP_A) || defined(DDPB_DBG_SKIP_B)
        // timing experiments only (results are wrong): after the ring is primed, skip one operand
        const bool primed = (k - k_begin) >= Cfg::kStages;
#ifdef DDPB_DBG_SKIP_A
        const bool do_a = !primed, do_b = true;
#else
        const bool do_a = true, do_b = !primed;
#endif
        mbar_expect_tx(&full_bar[stage], (do_a ? Cfg::kABytes : 0) + (do_b ? Cfg::kBBytes : 0));
        if (do_a) tma_load_4d(a_dst, &tmap_x, &full_bar[stage], cc * kBlockK, ds, h0 + dr, n_img);
        if (!do_b) {
        } else
#else
        mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
        tma_load_4d(a_dst, &tmap_x, &full_bar[stage], cc * kBlockK, ds, h0 + dr, n_img);
#endif
        if (!kDgrad) {
          tma_load_2d(b_dst, &tmap_w, &full_bar[stage], tap * p.C_in + cc * kBlockK, n_tile * BLOCK_N);
        } else {
#pragma unroll
          for (int j = 0; j < BLOCK_N / 64; ++j)
            tma_load_2d(b_dst + j * 8192, &tmap_w, &full_bar[stage],
                        (p.taps - 1 - tap) * p.C_out + n_tile * BLOCK_N + j * 64, cc * kBlockK);
        }
        if (++stage == Cfg::kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(kBlockM, BLOCK_N, 0, kDgrad ? 1 : 0);
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
          // K-major B: 32 bytes further along the 128-byte row; MN-major B (dgrad): 16 rows down,
          // 64-channel column blocks 8 KB apart
          const uint64_t bdesc = kDgrad ? make_smem_desc_sw128(b_addr + kk * kUmmaK * 128, 8192, 1024)
                                        : make_smem_desc_sw128(b_addr + kk * kUmmaK * 2, 16, 1024);
          umma_f16(tmem_base, adesc, bdesc, idesc, (k | kk) != 0);
        }
        umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs retire
        if (++stage == Cfg::kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma_commit(tmem_full_bar);  // accumulator complete
      if (p.dbg != nullptr) {
        mbar_wait(tmem_full_bar, 0);
        dbg_stamp(p, 2);
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    conv_epilogue<BLOCK_N>(p, tmem_base, tmem_full_bar, s_stats, warp, m_tile, n_tile);
  }

  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) dbg_stamp(p, 3);
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BLOCK_N);
  }
}

// Split-K finalize: y = bf16(sum_s partial[s] + bias); stats += (sum y, sum y^2) per channel.
// One thread = one pixel x 8 channels; 256 % (C_out/8) == 0 keeps a thread on its channels.
__global__ void __launch_bounds__(256)
splitk_finalize_kernel(const float* __restrict__ partial, const float* __restrict__ bias,
                       __nv_bfloat16* __restrict__ out, float* __restrict__ stats, int splits, long long P,
                       int C_out) {
  extern __shared__ float s_acc[];  // [2][C_out]
  pdl_launch_dependents();
  for (int c = threadIdx.x; c < 2 * C_out; c += blockDim.x) s_acc[c] = 0.f;
  __syncthreads();
  pdl_wait();
  const int groups = C_out / 8;
  const long long items = P * groups;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const int cg = static_cast<int>((blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) % groups);
  float b[8], a1[8], a2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    b[j] = bias != nullptr ? bias[cg * 8 + j] : 0.f;
    a1[j] = a2[j] = 0.f;
  }
  const long long split_stride = P * C_out;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < items; i += stride) {
    const float* src = partial + i * 8;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = b[j];
    for (int sp = 0; sp < splits; ++sp) {
      const float4 lo = *reinterpret_cast<const float4*>(src + sp * split_stride);
      const float4 hi = *reinterpret_cast<const float4*>(src + sp * split_stride + 4);
      v[0] += lo.x; v[1] += lo.y; v[2] += lo.z; v[3] += lo.w;
      v[4] += hi.x; v[5] += hi.y; v[6] += hi.z; v[7] += hi.w;
    }
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]);
    o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]);
    o.w = pack_bf16x2(v[6], v[7]);
    reinterpret_cast<uint4*>(out)[i] = o;
    if (stats != nullptr) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float r = bf16_round(v[j]);
        a1[j] += r;
        a2[j] += r * r;
      }
    }
  }
  if (stats != nullptr) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float x1 = a1[j], x2 = a2[j];
      for (int off = 16; off >= groups; off >>= 1) {
        x1 += __shfl_xor_sync(0xffffffffu, x1, off);
        x2 += __shfl_xor_sync(0xffffffffu, x2, off);
      }
      if (groups >= 32 || static_cast<int>(lane_id()) < groups) {
        atomicAdd(&s_acc[cg * 8 + j], x1);
        atomicAdd(&s_acc[C_out + cg * 8 + j], x2);
      }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < 2 * C_out; c += blockDim.x) atomicAdd(&stats[c], s_acc[c]);
  }
}

#include "conv_persist.inl"
#include "conv_pair_persist.inl"

// ------------------------------------------------------------------------------------------
// Launch planning.  Measured on B200 (profiles/): these kernels are bound by operand ingest
// per SM (~45 B/clk) unless the tile is large, and by fixed per-CTA cost when the reduction is
// short.  The planner estimates every legal (CTA pair?, BLOCK_N, split-K) choice with that
// model and takes the cheapest.
// ------------------------------------------------------------------------------------------
struct ConvPlan {
  int pair;      // kernel: 0 = one tile per CTA, 2 = persistent, 3 = persistent CTA pair (cta_group::2)
  int block_n;
  int splits;
  bool fused;    // split-K partial sums added up inside the persistent kernel (needs the caller's sync words)
};

static int g_force_pair = -1, g_force_block_n = -1, g_force_splits = -1;
static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return (e != nullptr && e[0] != 0) ? atoi(e) : dflt;
}
// split-K summed inside the persistent kernel when the caller gives sync words (0: second kernel); environment
// DDPB_FUSE_SPLIT / DDPB_TAIL_TMEM set the start values (A/B runs of bench.py), the hooks below change them
static int g_fuse_split = env_int("DDPB_FUSE_SPLIT", 1);
constexpr int kSplitSyncWords = 4096;   // two words per (output tile, epilogue warp): up to 256 tiles
static int g_tail_tmem = env_int("DDPB_TAIL_TMEM", 1);   // BatchNorm tail of the persistent kernel from TMEM where possible (0: always from L2)
static unsigned long long* g_dbg = nullptr;
static int g_dbg_mode = 0;

// Cycle estimates for the persistent kernel, calibrated on B200 with the in-kernel timeline
// hook (profiles/r1_call11_conv_timeline_persistent.txt): a 64-deep k-step costs max(1.35 x tensor time,
// stage bytes / 55 B per clk of operand ingest); draining a tile costs ~30 cycles per column
// with eight epilogue warps and overlaps the next tile's main loop.
static double plan_cost(int bn, int splits, long long P, int C_out, int total_k, bool fused, bool with_tail) {
  const int m_tiles = static_cast<int>((P + kBlockM - 1) / kBlockM);
  const int items = m_tiles * (C_out / bn) * splits;
  const int sms = sm_count();
  const double k_iters = static_cast<double>(total_k) / splits;
  const double stage_bytes = 16384.0 + bn * 128.0;
  const double tensor = 2.0 * bn * 1.35, ingest = stage_bytes / 55.0;
  const double mainloop = k_iters * (tensor > ingest ? tensor : ingest);
  const double epi = 30.0 * bn * (splits > 1 ? 0.6 : 1.0);
  const int per_sm = (items + sms - 1) / sms;
  double t = per_sm * (mainloop > epi ? mainloop : epi) + (mainloop > epi ? epi : mainloop) + 2500.0;
  // split-K finalize: re-reads splits x P x C_out floats (L2-resident at these sizes) + a launch
  // summed inside the kernel: the warp that draws the last ticket reads the other splits' slices of its tile
  // (measured, profiles/r2_call34_*: ticket + wait for the slowest split + one L2 round trip per split and
  // 32-column chunk, about 7 us at 4 splits: worth it where it keeps the BatchNorm tail inside the kernel)
  if (splits > 1 && fused) t += 8000.0 + (splits - 1) * 2700.0 + (bn == 64 ? 4500.0 : 0.0);
  // (64-wide: the model above is ~2 us optimistic for the fused 64 x2 plan of layers 6 / 7, measured 22.1 / 22.9 us
  // against 20.5 / 21.5 us for 128 x4, profiles/r2_call37_*)
  else if (splits > 1) t += 6000.0 + splits * static_cast<double>(P) * C_out * 4.0 / 3000.0;
  // a BatchNorm tail that cannot run inside the kernel (split-K summed by the second kernel) is a third launch
  if (splits > 1 && !fused && with_tail) t += 8000.0;
  return t;
}

// Automatic choice: the persistent kernel with the cheapest (BLOCK_N, split-K); the
// one-tile-per-CTA and CTA-pair kernels stay available through ddpb_conv_force (mode 0 / 1).
static ConvPlan plan_conv_mode(int mode, long long P, int C_out, int total_k, int block_n_hint, bool allow_split,
                               bool fuse_ok, bool with_tail, bool* found) {
  ConvPlan best{mode, (C_out % 128 == 0) ? 128 : 64, 1, false};
  const long long m_tiles = (P + kBlockM - 1) / kBlockM;
  double best_t = 1e30;
  const int bns[3] = {64, 128, 256};
  for (int bi = 0; bi < 3; ++bi) {
    const int bn = bns[bi];
    if (C_out % bn != 0) continue;
    if (mode == 3 && bn < 128) continue;   // a CTA pair splits B in halves of >= 64 rows
    if (block_n_hint != 0 && bn != block_n_hint) continue;
    if (g_force_block_n >= 0 && bn != g_force_block_n) continue;
    for (int splits = 1; splits <= 8; ++splits) {
      if (g_force_splits >= 0 && splits != g_force_splits) continue;
      if (mode == 3 && splits > 1) continue;              // the persistent CTA-pair kernel does not split K
      if (splits > 1 && (!allow_split || total_k / splits < 4)) continue;
      // in-kernel sum: persistent kernel, sync words for every (tile, warp); else the finalize kernel's thread mapping
      // and one item per CTA: a warp waiting for the other splits' slices then only ever waits for warps whose
      // single accumulator is already being produced (with several items per CTA, waits could chain through
      // recycled accumulator buffers)
      const bool fused = splits > 1 && fuse_ok && mode == 2 && m_tiles * (C_out / bn) * 16 <= kSplitSyncWords &&
                         m_tiles * (C_out / bn) * splits <= sm_count();
      if (splits > 1 && !fused && (C_out / 8 > 256 || 256 % (C_out / 8) != 0)) continue;
      const double t = plan_cost(bn, splits, P, C_out, total_k, fused, with_tail);
      if (t < best_t) {
        best_t = t;
        best = ConvPlan{mode, bn, splits, fused};
      }
    }
  }
  if (found != nullptr) *found = best_t < 1e30;
  return best;
}

static ConvPlan plan_conv(long long P, int C_out, int total_k, int block_n_hint, bool allow_split, bool fuse_ok,
                          bool with_tail = false) {
  const int mode = (g_force_pair == 0 || g_force_pair == 3) ? g_force_pair : 2;
  fuse_ok = fuse_ok && g_fuse_split != 0;
  bool found = false;
  const ConvPlan pl = plan_conv_mode(mode, P, C_out, total_k, block_n_hint, allow_split, fuse_ok, with_tail, &found);
  // mode 3 (persistent CTA pair) has no 64-wide tile and no split-K: such shapes use mode 2
  if (mode == 3 && !found) return plan_conv_mode(2, P, C_out, total_k, block_n_hint, allow_split, fuse_ok, with_tail, nullptr);
  return pl;
}

// floats of split-K scratch a plan needs: whole 128-row tiles per split (the in-kernel sum stores tile slices)
static long long plan_scratch(const ConvPlan& pl, long long P, int C_out) {
  return pl.splits > 1 ? pl.splits * ((P + kBlockM - 1) / kBlockM) * kBlockM * C_out : 0;
}

template <int BLOCK_N, bool kDgrad>
static int launch_fprop(const CUtensorMap& tx, const CUtensorMap& tw, const CUtensorMap& /*to*/, const FpropParams& p,
                        int m_tiles, int splits, cudaStream_t stream) {
  using Cfg = FpropCfg<BLOCK_N>;
  static bool configured = false;
  if (!configured) {
    DDPB_CHECK_CUDA(cudaFuncSetAttribute(conv_fprop_kernel<BLOCK_N, kDgrad>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    configured = true;
  }
  dim3 grid(m_tiles, p.C_out / BLOCK_N, splits);
  DDPB_CHECK_CUDA(launch_pdl(conv_fprop_kernel<BLOCK_N, kDgrad>, grid, kConvThreads, Cfg::kSmemBytes, stream, tx, tw, p));
  return 0;
}

}  // namespace ddpb

using namespace ddpb;

// Testing / tuning hook: force the planner's choice (-1 = automatic) for all later launches.
extern "C" void ddpb_conv_force(int pair, int block_n, int splits) {
  g_force_pair = pair;
  g_force_block_n = block_n;
  g_force_splits = splits;
}

// Words (uint32) of the `split_sync` buffer the conv entry points take: zero it ONCE; the kernels leave it zero.
// Launches that may run concurrently need separate buffers.
extern "C" int ddpb_conv_split_sync_words() { return kSplitSyncWords; }

// Tuning hook: 0 = split-K partial sums always go through the second (finalize) kernel even with sync words.
extern "C" void ddpb_conv_fuse_split(int on) { g_fuse_split = on; }

// Tuning hook: 1 (default) = the fused BatchNorm tail reads the accumulators back from TMEM where that is
// faster (one tile per CTA), 2 = wherever the shape allows it, 0 = it always re-reads the stored bf16 tiles
// from L2.  All give the same output.
extern "C" void ddpb_conv_tail_from_tmem(int on) { g_tail_tmem = on; }

// Profiling hook: device buffer of 8 x (number of CTAs) uint64 that every conv CTA stamps with
// %globaltimer at start / after set-up / when its accumulator is complete / at exit (null = off).
extern "C" void ddpb_conv_debug_timing(unsigned long long* buf) { g_dbg = buf; }

// Timing probes for the persistent kernels (the results are numerically wrong): bit 0 = the producer
// stops loading B once the ring is primed, bit 1 = same for A.  0 = normal operation.
extern "C" void ddpb_conv_debug_mode(int mode) { g_dbg_mode = mode; }

// What the planner picks for this shape: out3 = {pair, block_n, splits}.
extern "C" void ddpb_conv_plan(int N, int H, int W, int C_in, int C_out, int taps, int block_n, int* out3) {
  const ConvPlan pl = plan_conv(static_cast<long long>(N) * H * W, C_out, taps * (C_in / 64), block_n, true, true);
  out3[0] = pl.pair;
  out3[1] = pl.block_n;
  out3[2] = pl.splits;
}

// The same with the caller's situation spelled out: with_tail = the fused conv + BatchNorm entry point,
// with_sync = sync words are passed.  out4 = {pair, block_n, splits, split-K summed inside the kernel}.
extern "C" void ddpb_conv_plan_ex(int N, int H, int W, int C_in, int C_out, int taps, int block_n, int with_tail,
                                  int with_sync, int* out4) {
  const ConvPlan pl = plan_conv(static_cast<long long>(N) * H * W, C_out, taps * (C_in / 64), block_n, true,
                                with_sync != 0, with_tail != 0);
  out4[0] = pl.pair;
  out4[1] = pl.block_n;
  out4[2] = pl.splits;
  out4[3] = pl.fused ? 1 : 0;
}

// Floats of scratch ddpb_conv_fprop wants for this shape (0: it will not split K).
extern "C" long long ddpb_conv_fprop_scratch_elems(int N, int H, int W, int C_in, int C_out, int taps,
                                                   int block_n) {
  const long long P = static_cast<long long>(N) * H * W;
  if (C_out % 64 != 0 || C_in % 64 != 0) return 0;
  // enough for the plan with and the plan without the in-kernel sum (the caller may or may not pass sync words)
  // and with / without a BatchNorm tail
  long long need = 0;
  for (int v = 0; v < 4; ++v) {
    const long long a = plan_scratch(plan_conv(P, C_out, taps * (C_in / 64), block_n, true, (v & 1) != 0, (v & 2) != 0), P, C_out);
    need = a > need ? a : need;
  }
  return need;
}

// Shared host path.  dgrad == 0: y[P][C_out] = conv(x[..C_in], w[C_out][taps*C_in]).
// dgrad == 1: x is dY [..C_in := layer C_out], w is the layer's FPROP pack
// [layer C_out][taps][layer C_in], y = dX [P][C_out := layer C_in].
struct BnTail {
  const float* gamma;
  const float* beta;
  void* act;
  float* saved;
  int pool;
  float eps;
};

// Returns 0 on success; *fused_bn (if given) tells whether the BatchNorm tail ran inside the conv kernel.
static int conv_impl(int dgrad, const void* x, const void* w, const float* bias, void* y, float* stats, int N, int H,
                     int W, int C_in, int C_out, int taps, int block_n, float* scratch, long long scratch_elems,
                     unsigned int* split_sync, cudaStream_t stream, const BnTail* bn_tail = nullptr,
                     int* fused_bn = nullptr) {
  const char* name = dgrad ? "conv_dgrad" : "conv_fprop";
  DDPB_REQUIRE(taps == 1 || taps == 9, "%s: taps must be 1 or 9 (got %d)", name, taps);
  DDPB_REQUIRE(C_in % 64 == 0 && C_in > 0, "%s: reduction channels must be a multiple of 64 (got %d)", name, C_in);
  DDPB_REQUIRE(C_out % 64 == 0 && C_out > 0, "%s: output channels must be a multiple of 64 (got %d)", name, C_out);
  int bh, bn, hblocks;
  DDPB_REQUIRE(pixel_box(H, W, kBlockM, &bh, &bn, &hblocks), "%s: unsupported spatial size %dx%d for 128-pixel tiles",
               name, H, W);
  DDPB_REQUIRE(block_n == 0 || block_n == 64 || block_n == 128 || block_n == 256, "%s: bad block_n %d", name, block_n);
  DDPB_REQUIRE(block_n == 0 || C_out % block_n == 0, "%s: output channels %d not divisible by block_n %d", name,
               C_out, block_n);
  const long long P = static_cast<long long>(N) * H * W;
  const bool with_tail = bn_tail != nullptr && stats != nullptr;
  ConvPlan pl = plan_conv(P, C_out, taps * (C_in / 64), block_n, scratch != nullptr, split_sync != nullptr, with_tail);
  // The split-K choice is cost based, so a smaller batch can want MORE scratch than a larger one.  A plan
  // whose partial tiles do not fit in the caller's buffer is never launched: plan again without split-K.
  if (pl.splits > 1 && plan_scratch(pl, P, C_out) > scratch_elems) {
    DDPB_REQUIRE(g_force_splits < 0,
                 "%s: split-K x%d was forced but the scratch buffer holds %lld floats (needs %lld)", name, pl.splits,
                 scratch_elems, plan_scratch(pl, P, C_out));
    pl = plan_conv(P, C_out, taps * (C_in / 64), block_n, false, false);
  }
  block_n = pl.block_n;
  const int b_rows = (pl.pair == 3) ? block_n / 2 : block_n;   // filter rows one CTA loads per k-step

  CUtensorMap tx, tw;
  {
    uint64_t dims[4] = {(uint64_t)C_in, (uint64_t)W, (uint64_t)H, (uint64_t)N};
    uint64_t strides[3] = {(uint64_t)C_in * 2, (uint64_t)W * C_in * 2, (uint64_t)H * W * C_in * 2};
    uint32_t box[4] = {64, (uint32_t)W, (uint32_t)bh, (uint32_t)bn};
    if (encode_tmap(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x, dims, strides, box,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B))
      return 3;
  }
  if (!dgrad) {
    uint64_t dims[2] = {(uint64_t)taps * C_in, (uint64_t)C_out};
    uint64_t strides[1] = {(uint64_t)taps * C_in * 2};
    uint32_t box[2] = {64, (uint32_t)b_rows};
    if (encode_tmap(&tw, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, w, dims, strides, box,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B))
      return 3;
  } else {
    // fprop pack seen as rows = layer C_out (our reduction), columns = taps * layer C_in
    uint64_t dims[2] = {(uint64_t)taps * C_out, (uint64_t)C_in};
    uint64_t strides[1] = {(uint64_t)taps * C_out * 2};
    uint32_t box[2] = {64, 64};
    if (encode_tmap(&tw, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, w, dims, strides, box,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B))
      return 3;
  }
  CUtensorMap to;
  {
    // output tiles leave through TMA stores: boxes of 32 rows x 64 (or, for 64-wide tiles, 32) columns
    const uint32_t box_cols = block_n >= 128 ? 64 : 32;
    uint64_t dims[2] = {(uint64_t)C_out, (uint64_t)P};
    uint64_t strides[1] = {(uint64_t)C_out * 2};
    uint32_t box[2] = {box_cols, 32};
    if (encode_tmap(&to, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, y, dims, strides, box,
                    box_cols == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_128B))
      return 3;
  }
  if (pl.pair == 2 && pl.splits > 1 && !pl.fused) {
    // persistent kernel, split-K summed by the second kernel: no bf16 tile is stored by this launch, and the
    // same parameter slot carries the fp32 map of the partial buffer [split][P][C_out], boxes of 32 x 32
    uint64_t dims[3] = {(uint64_t)C_out, (uint64_t)P, (uint64_t)pl.splits};
    uint64_t strides[2] = {(uint64_t)C_out * 4, (uint64_t)P * C_out * 4};
    uint32_t box[3] = {32, 32, 1};
    if (encode_tmap(&to, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, scratch, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_128B))
      return 3;
  }
  FpropParams p;
  p.out = static_cast<__nv_bfloat16*>(y);
  p.bias = bias;
  p.stats = stats;
  p.P = static_cast<int>(P);
  p.C_in = C_in;
  p.C_out = C_out;
  p.taps = taps;
  p.bh = bh;
  p.bn = bn;
  p.hblocks = hblocks;
  p.dbg = g_dbg;
  p.dbg_mode = g_dbg_mode;
  p.gamma = p.beta = nullptr;
  p.act = nullptr;
  p.saved = nullptr;
  p.grid_sync = nullptr;
  p.H = H;
  p.W = W;
  p.pool = 0;
  p.eps = 0.f;
  p.tail_tmem = 0;
  p.split_sync = pl.fused ? split_sync : nullptr;
  if (fused_bn != nullptr) *fused_bn = 0;
  // BatchNorm tail inside the kernel: persistent kernel, whole 2x2 windows per tile
  if (bn_tail != nullptr && stats != nullptr && (pl.pair == 2 || pl.pair == 3) && (pl.splits == 1 || pl.fused) &&
      (!bn_tail->pool || bh % 2 == 0)) {
    p.gamma = bn_tail->gamma;
    p.beta = bn_tail->beta;
    p.act = static_cast<__nv_bfloat16*>(bn_tail->act);
    p.saved = bn_tail->saved;
    p.grid_sync = reinterpret_cast<unsigned int*>(stats + 2 * C_out);
    p.pool = bn_tail->pool;
    p.eps = bn_tail->eps;
    if (fused_bn != nullptr) *fused_bn = 1;
    // Tail from TMEM: possible when no accumulator buffer is recycled (tiles per CTA <= buffers) and, with
    // pooling, a warp's 32 accumulator rows hold whole 2x2 windows (W a power of two <= 16: the window is
    // lanes l, l^1, l^W, l^W^1).  Measured per layer at batch 256 (profiles/r2_call32_*): with one tile per CTA
    // the tail takes 4.9 / 3.6 / 3.6 / 3.1 us from TMEM against 8.5 / 4.1 / 4.3 / 3.4 us from L2 (layers 2-5);
    // with four tiles per CTA (layer 1) 6.2 against 5.4 us: the pass is bound by the shared-memory pipe
    // (per-column coefficients are broadcast loads) and uses 8 of the 10 warps.  g_tail_tmem = 2 forces it
    // wherever it is possible (tests).
    const long long items = static_cast<long long>((p.P + kBlockM - 1) / kBlockM) * (C_out / block_n);
    const long long grid = items < sm_count() ? items : sm_count();
    const int acc_bufs = (512 / block_n) < 4 ? (512 / block_n) : 4;
    const long long per_cta = (items + grid - 1) / grid;
    const bool pool_ok = !bn_tail->pool || (W >= 2 && W <= 16 && (W & (W - 1)) == 0 && H % 2 == 0);
    if (pl.pair == 2 && pl.splits == 1 && pool_ok &&
        ((g_tail_tmem == 1 && per_cta == 1) || (g_tail_tmem == 2 && per_cta <= acc_bufs)))
      p.tail_tmem = 1;
  }
  const int m_tiles = (p.P + kBlockM - 1) / kBlockM;
  const int splits = pl.splits;
  p.partial = splits > 1 ? scratch : nullptr;
  int rc = 0;
#define DDPB_CONV_DISPATCH(FN)                                                     \
  if (!dgrad) {                                                                    \
    switch (block_n) {                                                             \
      case 64: rc = (pl.pair == 2 ? launch_persist<64, false>(tx, tw, to, p, m_tiles, splits, stream)          \
                                    : launch_fprop<64, false>(tx, tw, to, p, m_tiles, splits, stream)); break; \
      case 128: rc = FN<128, false>(tx, tw, to, p, m_tiles, splits, stream); break;    \
      default: rc = FN<256, false>(tx, tw, to, p, m_tiles, splits, stream); break;     \
    }                                                                              \
  } else {                                                                         \
    switch (block_n) {                                                             \
      case 64: rc = (pl.pair == 2 ? launch_persist<64, true>(tx, tw, to, p, m_tiles, splits, stream)           \
                                    : launch_fprop<64, true>(tx, tw, to, p, m_tiles, splits, stream)); break;  \
      case 128: rc = FN<128, true>(tx, tw, to, p, m_tiles, splits, stream); break;     \
      default: rc = FN<256, true>(tx, tw, to, p, m_tiles, splits, stream); break;      \
    }                                                                              \
  }
  if (pl.pair == 2) {
    DDPB_CONV_DISPATCH(launch_persist)
  } else if (pl.pair == 3) {
    DDPB_CONV_DISPATCH(launch_pair_persist)
  } else {
    DDPB_CONV_DISPATCH(launch_fprop)
  }
#undef DDPB_CONV_DISPATCH
  if (rc != 0 || splits == 1 || pl.fused) return rc;
  // (one channel-owner launch doing finalize + statistics + BatchNorm for the small split-K layers was tried and
  // removed: 64 CTAs pulling the fp32 partial tiles in 32-byte pieces took 27-28 us against 25 us for the
  // finalize + BatchNorm kernels, profiles/r2_call16_*)
  const long long items = static_cast<long long>(p.P) * (C_out / 8);
  long long blocks = (items + 255) / 256;
  if (blocks > 4LL * sm_count()) blocks = 4LL * sm_count();
  const float* partial = scratch;
  DDPB_CHECK_CUDA(launch_pdl(splitk_finalize_kernel, static_cast<int>(blocks), 256, 2 * C_out * sizeof(float), stream,
                             partial, bias, p.out, stats, splits, static_cast<long long>(p.P), C_out));
  return 0;
}

// x: [N][H][W][C_in] bf16; w: [C_out][taps*C_in] bf16; y: [N*H*W][C_out] bf16.
// bias (fp32, may be null) is added before rounding; stats (fp32 [2][C_out], may be
// null) is ACCUMULATED into (caller zeroes it).  block_n: 0 = choose automatically.
// scratch: device buffer of scratch_elems floats (ddpb_conv_fprop_scratch_elems() says how many the
// planner would like), or null to forbid split-K (small problems then run on few SMs).  A plan
// that does not fit in scratch_elems is replaced by one without split-K, never launched.
extern "C" int ddpb_conv_fprop(const void* x, const void* w, const float* bias, void* y, float* stats,
                               int N, int H, int W, int C_in, int C_out, int taps, int block_n,
                               float* scratch, long long scratch_elems, unsigned int* split_sync, cudaStream_t stream) {
  return conv_impl(0, x, w, bias, y, stats, N, H, W, C_in, C_out, taps, block_n, scratch,
                   scratch != nullptr ? scratch_elems : 0, split_sync, stream);
}

int bn_relu_pool_fwd_launch(const void* x, const float* stats, const float* gamma, const float* beta, void* y,
                            float* saved, int N, int H, int W, int C, int pool, float eps, cudaStream_t stream);

// Convolution + BatchNorm(batch statistics) + ReLU (+ 2x2 max-pool) forward: O1+O4+O6+O7.
// conv_out ([N*H*W][C_out] bf16) is still written (the backward pass needs it); act is the
// block's output.  stats: fp32 scratch of 2*C_out + 32 floats, ZERO on entry.  When the
// persistent kernel runs without split-K the normalisation happens inside it after a
// grid-wide barrier; otherwise the separate BatchNorm kernel is launched.
extern "C" int ddpb_conv_bn_relu_pool_fwd(const void* x, const void* w, const float* bias, void* conv_out,
                                          float* stats, const float* gamma, const float* beta, void* act,
                                          float* saved, int N, int H, int W, int C_in, int C_out, int taps,
                                          int pool, float eps, float* scratch, long long scratch_elems,
                                          unsigned int* split_sync, cudaStream_t stream) {
  BnTail bn{gamma, beta, act, saved, pool, eps};
  int fused = 0;
  // (summing the split-K partials inside the persistent kernel after a second grid barrier was tried and
  // removed: 8 us per step slower than separate kernels with programmatic launch overlap, profiles/r1_call40_*)
  if (int rc = conv_impl(0, x, w, bias, conv_out, stats, N, H, W, C_in, C_out, taps, 0, scratch,
                         scratch != nullptr ? scratch_elems : 0, split_sync, stream, &bn, &fused))
    return rc;
  if (fused) return 0;
  return bn_relu_pool_fwd_launch(conv_out, stats, gamma, beta, act, saved, N, H, W, C_out, pool, eps, stream);
}

// Data gradient of the 3x3 convolution (O2).  dy: [N][H][W][C_out] bf16; w: the SAME bf16 pack
// fprop uses, [C_out][9][C_in]; dx: [N*H*W][C_in] bf16.  scratch: as for fprop, sized with
// ddpb_conv_fprop_scratch_elems(N, H, W, C_out, C_in, 9, block_n).
extern "C" int ddpb_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int C_in, int C_out,
                               int block_n, float* scratch, long long scratch_elems, unsigned int* split_sync,
                               cudaStream_t stream) {
  return conv_impl(1, dy, w, nullptr, dx, nullptr, N, H, W, C_out, C_in, 9, block_n, scratch,
                   scratch != nullptr ? scratch_elems : 0, split_sync, stream);
}
