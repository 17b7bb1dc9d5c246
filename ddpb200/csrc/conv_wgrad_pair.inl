// CTA-pair weight-gradient kernel (included by conv_wgrad.cu) -- ddpb_conv_wgrad_set_mode(2).
// NOT YET RUN ON HARDWARE (written at the end of round 1 without GPU time left; its tests are
// gated by DDPB_TEST_EXPERIMENTAL=1, see docs/NEXT.md).
//
// Same GEMM as conv_wgrad_kernel (A = dY and B = X both MN-major, 64-pixel k-steps), but the two
// CTAs of a (2,1,1) cluster issue ONE tcgen05.mma with cta_group::2: M = 256 output channels
// (each CTA supplies its own 128 channels of dY) x N = BLOCK_N input channels (each CTA supplies
// HALF of them).  For BLOCK_N = 256 a CTA still loads 32 KB per k-step, as the one-CTA kernel
// does for a 128x128 tile, but that now feeds a 128x256 slice of the accumulator: twice the math
// per byte written to and read from shared memory, which is what bounds the one-CTA kernel
// (DESIGN.md par. 7).  Protocol as in conv_pair_kernel: both producers signal the LEADER's full
// barrier, the leader's MMA thread multicasts "slot free" / "accumulator ready" to both CTAs.
// One tile per pair, pixel range split over gridDim.y, partial tiles summed by the existing
// deterministic reduction kernels.

template <int BLOCK_N>
struct WgradPairCfg {
  static_assert(BLOCK_N == 128 || BLOCK_N == 256, "each CTA of the pair supplies BLOCK_N / 2 >= 64 channels of X");
  static constexpr int kStages = 3;                                    // <= 97 KB: two CTAs per SM
  static constexpr int kABytes = 2 * kWgBoxBytes;                      // this CTA's 128 channels of dY
  static constexpr int kBBytes = (BLOCK_N / 128) * kWgBoxBytes;        // this CTA's half of the X channels
  static constexpr int kStageBytes = kABytes + kBBytes;                // 24 / 32 KB
  static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + 256;
};

// gridDim.x = 2 x (taps x m_pairs x n_tiles): consecutive CTAs form a cluster; gridDim.y = splits.
// p.m_tiles = number of 256-row PAIR tiles along C_out, p.n_tiles = C_in / BLOCK_N.
template <int BLOCK_N>
__global__ void __launch_bounds__(kWgThreads)
conv_wgrad_pair_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x,
                       const WgradParams p) {
  using Cfg = WgradPairCfg<BLOCK_N>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* tiles = smem;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* full_bar = bars;                       // used in the leader CTA
  uint64_t* empty_bar = bars + Cfg::kStages;       // each CTA waits on its own
  uint64_t* tmem_full_bar = bars + 2 * Cfg::kStages;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * Cfg::kStages + 1);

  const int warp = threadIdx.x >> 5;
  const uint32_t rank = cluster_ctarank();         // 0 = leader
  int t = blockIdx.x >> 1;                         // pair tile -> (tap, m_pair, n_tile)
  const int n_tile = t % p.n_tiles;
  t /= p.n_tiles;
  const int m_pair = t % p.m_tiles;
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
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int i = 0; i < Cfg::kStages; ++i) {
        mbar_init(&full_bar[i], 1);    // the leader's producer arrives with the byte count of BOTH CTAs
        mbar_init(&empty_bar[i], 1);   // one multicast commit from the leader's MMA thread
      }
      mbar_init(tmem_full_bar, 1);
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc_pair(tmem_ptr_smem, BLOCK_N);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();   // both CTAs' barriers exist before anything is signalled across the pair
  tc_fence_after();
  pdl_wait();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (elect_one()) {
      int dr = 0, ds = 0;
      if (p.taps == 9) {
        dr = tap / 3 - 1;
        ds = tap % 3 - 1;
      }
      const int k0 = (2 * m_pair + static_cast<int>(rank)) * 128;               // this CTA's output channels
      const int c0 = n_tile * BLOCK_N + static_cast<int>(rank) * (BLOCK_N / 2); // this CTA's input channels
      int stage = 0;
      uint32_t phase = 0;
      for (int pb = pb_begin; pb < pb_end; ++pb) {
        const int n_img = (pb / p.hblocks) * p.bn;
        const int h0 = (pb % p.hblocks) * p.bh;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* a_dst = tiles + stage * Cfg::kStageBytes;
        uint8_t* b_dst = a_dst + Cfg::kABytes;
        if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * Cfg::kStageBytes);
        tma_load_2d_pair(a_dst, &tmap_dy, &full_bar[stage], k0, pb * kWgPix);
        tma_load_2d_pair(a_dst + kWgBoxBytes, &tmap_dy, &full_bar[stage], k0 + 64, pb * kWgPix);
#pragma unroll
        for (int j = 0; j < BLOCK_N / 128; ++j)
          tma_load_4d_pair(b_dst + j * kWgBoxBytes, &tmap_x, &full_bar[stage], c0 + j * 64, ds, h0 + dr, n_img);
        if (++stage == Cfg::kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(256, BLOCK_N, 1, 1);
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
          umma_f16_pair(tmem_base, adesc, bdesc, idesc, (k | kk) != 0);
        }
        umma_commit_pair(&empty_bar[stage], 0x3);   // slot free in BOTH CTAs
        if (++stage == Cfg::kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma_commit_pair(tmem_full_bar, 0x3);         // accumulators ready in BOTH CTAs
    }
  } else {
    // ===================== epilogue (both CTAs: own 128 rows, all BLOCK_N columns) =====================
    const int q = warp & 3;
    const int row = q * 32 + lane_id();
    const int k_ch = (2 * m_pair + static_cast<int>(rank)) * 128 + row;
    const bool valid = k_ch < p.C_out;
    float* out_row = p.out + static_cast<long long>(split) * p.out_split_stride +
                     static_cast<long long>(k_ch) * p.out_ld + tap * p.C_in + n_tile * BLOCK_N;
    if (num_k > 0) {
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
    }
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
      if (valid) {
        uint4* dst = reinterpret_cast<uint4*>(out_row + chunk * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j) dst[j] = make_uint4(raw[4 * j], raw[4 * j + 1], raw[4 * j + 2], raw[4 * j + 3]);
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();   // nobody frees TMEM or exits while the peer may still signal / read
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, BLOCK_N);
  }
}

// p.m_tiles / p.n_tiles must already be in PAIR-tile units (C_out / 256, C_in / BLOCK_N).
template <int BLOCK_N>
static int launch_wgrad_pair(const CUtensorMap& tdy, const CUtensorMap& tx, const WgradParams& p, int splits,
                             cudaStream_t stream) {
  using Cfg = WgradPairCfg<BLOCK_N>;
  static bool configured = false;
  if (!configured) {
    DDPB_CHECK_CUDA(cudaFuncSetAttribute(conv_wgrad_pair_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * p.taps * p.m_tiles * p.n_tiles, splits, 1);
  cfg.blockDim = dim3(kWgThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  DDPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, conv_wgrad_pair_kernel<BLOCK_N>, tdy, tx, p));
  return 0;
}
