// Tap-fused weight-gradient kernel with an on-chip split reduction (included by conv_wgrad.cu).
//
// The one-tap kernel above is bound by operand traffic from L2, not by the tensor pipe: every
// k-step brings 32 KB for one 128x128x64 MMA block, and the dY tile is fetched again for each
// of the nine taps.  Its pixel splits also meet in global memory: up to 19 MB of fp32 partial
// tiles per layer are written, read back and summed by a second kernel.  This kernel changes both:
//   * one CTA computes the THREE taps of a filter row (dr fixed, ds = -1, 0, +1) from one dY
//     tile: 64 KB per k-step feed three MMA blocks (21 KB each instead of 32 KB); the three
//     accumulators sit side by side in TMEM (3 x BLOCK_N columns);
//   * the pixel range is split over the CTAs of a thread-block cluster (1, S, 1).  When the main
//     loop is done every CTA parks its accumulators in its own shared memory (the operand ring
//     is idle by then), the cluster synchronises, and CTA r sums rows [128 r / S, 128 (r+1) / S)
//     of all S copies through distributed shared memory in a fixed order -- deterministic -- and
//     writes them straight to dW.  Global partial tiles are only needed for the G > 1 groups of
//     clusters that very long pixel ranges are divided into (S x G splits in total).
// Used for 3x3 layers with C_out a multiple of 128; the first layer (one tap, 64 filters) keeps
// the kernel above.

constexpr int kWg3Taps = 3;

template <int BLOCK_N>
struct Wgrad3Cfg {
  static constexpr int kABytes = 2 * kWgBoxBytes;                   // 128 channels of dY
  static constexpr int kBTapBytes = (BLOCK_N / 64) * kWgBoxBytes;   // BLOCK_N channels of X, one tap
  static constexpr int kStageBytes = kABytes + kWg3Taps * kBTapBytes;   // 64 KB / 40 KB
  static constexpr int kStages = (BLOCK_N >= 128) ? 3 : 4;
  static constexpr int kAccCols = kWg3Taps * BLOCK_N;               // 384 / 192 TMEM columns in use
  static constexpr int kTmemCols = (kAccCols > 256) ? 512 : 256;
  static constexpr int kParkBytes = 128 * kAccCols * 4;             // accumulators parked as fp32
  static_assert(kParkBytes <= kStages * kStageBytes, "the parked accumulators must fit in the operand ring");
  static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + 256;
};

__device__ __forceinline__ uint32_t map_to_cta(uint32_t smem_addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(cta));
  return r;
}

__device__ __forceinline__ float4 ld_dsmem_f4(uint32_t cluster_addr) {
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(cluster_addr));
  return v;
}

// gridDim.x = 3 tap rows x m_tiles x n_tiles, gridDim.y = S x G pixel splits, cluster = (1, S, 1).
template <int BLOCK_N>
__global__ void __launch_bounds__(kWgThreads, 1)
conv_wgrad3_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x,
                   const WgradParams p, int cluster_size) {
  using Cfg = Wgrad3Cfg<BLOCK_N>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* tiles = smem;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + Cfg::kStages;
  uint64_t* tmem_full_bar = bars + 2 * Cfg::kStages;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 2 * Cfg::kStages + 1);

  const int warp = threadIdx.x >> 5;
  int t = blockIdx.x;
  const int n_tile = t % p.n_tiles;
  t /= p.n_tiles;
  const int m_tile = t % p.m_tiles;
  const int tap_row = t / p.m_tiles;                 // 0..2: dr = tap_row - 1
  const int split = blockIdx.y, splits = gridDim.y;
  const int rank = split % cluster_size;             // == %cluster_ctarank for a (1, S, 1) cluster
  const int group = split / cluster_size;
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
        mbar_init(&full_bar[i], 1);
        mbar_init(&empty_bar[i], 1);
      }
      mbar_init(tmem_full_bar, 1);
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_ptr_smem, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      const int dr = tap_row - 1;
      const int k0 = m_tile * 128;
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
        tma_load_2d(a_dst + kWgBoxBytes, &tmap_dy, &full_bar[stage], k0 + 64, pb * kWgPix);
#pragma unroll
        for (int tt = 0; tt < kWg3Taps; ++tt)
#pragma unroll
          for (int j = 0; j < BLOCK_N / 64; ++j)
            tma_load_4d(b_dst + tt * Cfg::kBTapBytes + j * kWgBoxBytes, &tmap_x, &full_bar[stage],
                        n_tile * BLOCK_N + j * 64, tt - 1, h0 + dr, n_img);
        if (++stage == Cfg::kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
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
        for (int tt = 0; tt < kWg3Taps; ++tt) {
#pragma unroll
          for (int kk = 0; kk < kWgPix / 16; ++kk) {
            const uint64_t adesc = make_smem_desc_sw128(a_addr + kk * 16 * 128, kWgBoxBytes, 1024);
            const uint64_t bdesc =
                make_smem_desc_sw128(b_addr + tt * Cfg::kBTapBytes + kk * 16 * 128, kWgBoxBytes, 1024);
            umma_f16(tmem_base + tt * BLOCK_N, adesc, bdesc, idesc, (k | kk) != 0);
          }
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
    // ===================== epilogue warps: TMEM -> dW (one CTA) or -> shared memory (cluster) =====
    const int q = warp & 3;
    const int row = q * 32 + lane_id();       // output channel within the M tile
    if (num_k > 0) {
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
    }
    float* out_row = p.out + static_cast<long long>(group) * p.out_split_stride +
                     static_cast<long long>(m_tile * 128 + row) * p.out_ld + (tap_row * kWg3Taps) * p.C_in +
                     n_tile * BLOCK_N;
    float4* park_row = reinterpret_cast<float4*>(tiles) + static_cast<long long>(row) * (Cfg::kAccCols / 4);
#pragma unroll 1
    for (int chunk = 0; chunk < Cfg::kAccCols / 32; ++chunk) {
      uint32_t raw[32];
      if (num_k > 0) {
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + chunk * 32, raw);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) raw[j] = 0u;
      }
      if (cluster_size == 1) {
        const int tt = (chunk * 32) / BLOCK_N, c = (chunk * 32) % BLOCK_N;
        uint4* dst = reinterpret_cast<uint4*>(out_row + tt * p.C_in + c);
#pragma unroll
        for (int j = 0; j < 8; ++j) dst[j] = make_uint4(raw[4 * j], raw[4 * j + 1], raw[4 * j + 2], raw[4 * j + 3]);
      } else {
        // 16-byte pieces of a 128-byte group are XOR-swizzled with the row so that the 32 rows a
        // warp writes at once (row pitch = a multiple of 128 bytes) spread over all banks
#pragma unroll
        for (int j = 0; j < 8; ++j)
          park_row[chunk * 8 + (j ^ (row & 7))] =
              make_float4(__uint_as_float(raw[4 * j]), __uint_as_float(raw[4 * j + 1]),
                          __uint_as_float(raw[4 * j + 2]), __uint_as_float(raw[4 * j + 3]));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (cluster_size > 1) {
    cluster_sync_all();   // every CTA's accumulators are parked
    const int row_begin = 128 * rank / cluster_size, row_end = 128 * (rank + 1) / cluster_size;
    constexpr int kRowF4 = Cfg::kAccCols / 4;
    const int items = (row_end - row_begin) * kRowF4;
    const uint32_t park_base = smem_u32(tiles);
    float* out_base = p.out + static_cast<long long>(group) * p.out_split_stride +
                      static_cast<long long>(m_tile * 128) * p.out_ld + (tap_row * kWg3Taps) * p.C_in +
                      n_tile * BLOCK_N;
    // every remote read is a round trip through the SM-to-SM network: keep 4 items x S loads in
    // flight per thread (the first version issued them one at a time and was slower than the
    // global-memory reduction it replaces)
    constexpr int kMaxCluster = 4, kItems = 4;
    uint32_t remote[kMaxCluster];
#pragma unroll
    for (int s = 0; s < kMaxCluster; ++s)
      remote[s] = map_to_cta(park_base, static_cast<uint32_t>(s < cluster_size ? s : 0));
    for (int i0 = threadIdx.x; i0 < items; i0 += blockDim.x * kItems) {
      float4 v[kItems][kMaxCluster];
      uint32_t off[kItems];
#pragma unroll
      for (int u = 0; u < kItems; ++u) {
        const int i = i0 + u * blockDim.x;
        const int row = row_begin + i / kRowF4;
        const int c16 = i % kRowF4;
        off[u] = (static_cast<uint32_t>(row) * kRowF4 + ((c16 & ~7) | ((c16 & 7) ^ (row & 7)))) * 16u;
        if (i < items) {
#pragma unroll
          for (int s = 0; s < kMaxCluster; ++s)
            if (s < cluster_size) v[u][s] = ld_dsmem_f4(remote[s] + off[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < kItems; ++u) {
        const int i = i0 + u * blockDim.x;
        if (i >= items) continue;
        float4 acc = v[u][0];
#pragma unroll
        for (int s = 1; s < kMaxCluster; ++s) {   // fixed order: deterministic
          if (s < cluster_size) {
            acc.x += v[u][s].x;
            acc.y += v[u][s].y;
            acc.z += v[u][s].z;
            acc.w += v[u][s].w;
          }
        }
        const int row = row_begin + i / kRowF4;
        const int col = (i % kRowF4) * 4;
        const int tt = col / BLOCK_N, c = col % BLOCK_N;
        *reinterpret_cast<float4*>(out_base + static_cast<long long>(row) * p.out_ld + tt * p.C_in + c) = acc;
      }
    }
    cluster_sync_all();   // nobody leaves while a peer may still read its shared memory
  }
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

template <int BLOCK_N>
static int launch_wgrad3(const CUtensorMap& tdy, const CUtensorMap& tx, const WgradParams& p, int cluster_size,
                         int groups, cudaStream_t stream) {
  using Cfg = Wgrad3Cfg<BLOCK_N>;
  static bool configured = false;
  if (!configured) {
    DDPB_CHECK_CUDA(cudaFuncSetAttribute(conv_wgrad3_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(kWg3Taps * p.m_tiles * p.n_tiles, cluster_size * groups, 1);
  cfg.blockDim = dim3(kWgThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1;
  attr[0].val.clusterDim.y = cluster_size;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  DDPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, conv_wgrad3_kernel<BLOCK_N>, tdy, tx, p, cluster_size));
  return 0;
}
