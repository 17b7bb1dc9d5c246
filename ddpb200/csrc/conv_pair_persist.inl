// Persistent CTA-pair convolution kernel (included by conv_fprop.cu) -- selectable with
// ddpb_conv_force(3, ...); NOT the default: see the status note in DESIGN.md par. 7.
//
// Why: the main loop of the persistent kernel (conv_persist.inl) runs at about half the tensor
// rate, and the numbers fit a shared-memory bandwidth bound: per 64-deep k-step a 128x128 tile
// writes 32 KB of operands into shared memory (TMA) and reads them back (tcgen05.mma).  With
// cta_group::2 the two CTAs of a cluster issue ONE M = 256 MMA; each CTA loads its own 128 pixel
// rows of A but only HALF of the filter tile B, so a 256x256 pair tile costs each SM 32 KB per
// k-step for twice the math.  The non-persistent CTA-pair kernel has this operand split but only
// a 5-stage ring and one tile per CTA (set-up and epilogue exposed); this kernel combines the
// split with the persistent structure:
//   * one cluster of two CTAs per SM pair, looping over 256-row x BLOCK_N pair tiles;
//   * the whole shared memory of each CTA is one operand ring (6 x 32 KB for BLOCK_N = 256);
//   * 2 (BLOCK_N = 256) or 4 (BLOCK_N = 128) TMEM accumulators, so the epilogues of both CTAs
//     drain pair tile i while the leader's MMA thread already works on i + 1;
//   * barriers: every CTA's producer waits on its OWN empty barrier (the leader's commit is
//     multicast to both), TMA completions of BOTH CTAs count on the LEADER's full barrier, and
//     "accumulator drained" is 16 arrivals (8 epilogue warps x 2 CTAs) on the LEADER's barrier,
//     the peer's warps arriving remotely through the cluster address space.
// Split-K is not implemented here (such shapes use the persistent one-CTA kernel).  The fused
// BatchNorm tail at the end mirrors conv_persist.inl.

template <int BLOCK_N>
struct PairPersistCfg {
  static_assert(BLOCK_N == 128 || BLOCK_N == 256, "a CTA pair splits B in halves of at least 64 rows");
  static constexpr int kABytes = kBlockM * kBlockK * 2;            // 16 KB: this CTA's pixels
  static constexpr int kBBytes = (BLOCK_N / 2) * kBlockK * 2;      // this CTA's half of the filters
  static constexpr int kStageBytes = kABytes + kBBytes;            // 24 / 32 KB
  static constexpr int kStages = kRingBytes / kStageBytes;         // 8 / 6
  static constexpr int kStatsBytes = 4 * 2 * BLOCK_N * 4;
  static constexpr int kBiasBytes = 8 * (BLOCK_N / 2) * 4;
  static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kStatsBytes + kBiasBytes + 512;
  static constexpr int kAccBufs = (512 / BLOCK_N) < 4 ? (512 / BLOCK_N) : 4;
  static constexpr int kTmemCols = kAccBufs * BLOCK_N;
};

__device__ __forceinline__ uint32_t pp_map_to_cta(uint32_t smem_addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(cta));
  return r;
}

// One arrival on an mbarrier that lives in another CTA of the cluster.
__device__ __forceinline__ void pp_arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}

template <int BLOCK_N, bool kDgrad>
__global__ void __launch_bounds__(512, 1)   // 320 threads are launched; 512 caps the registers at 128
conv_pair_persist_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                         const FpropParams p, int m_tiles, int n_tiles) {
  using Cfg = PairPersistCfg<BLOCK_N>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1k(smem_raw);
  uint8_t* tiles = smem;
  float* s_stats = reinterpret_cast<float*>(smem + Cfg::kStages * Cfg::kStageBytes);
  float* s_bias = reinterpret_cast<float*>(smem + Cfg::kStages * Cfg::kStageBytes + Cfg::kStatsBytes);
  uint64_t* bars =
      reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes + Cfg::kStatsBytes + Cfg::kBiasBytes);
  uint64_t* full_bar = bars;                         // [kStages]  (used in the leader CTA)
  uint64_t* empty_bar = bars + Cfg::kStages;         // [kStages]  (each CTA waits on its own)
  uint64_t* tmem_full = bars + 2 * Cfg::kStages;     // [kAccBufs] (each CTA waits on its own)
  uint64_t* tmem_empty = tmem_full + Cfg::kAccBufs;  // [kAccBufs] (used in the leader CTA)
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + Cfg::kAccBufs);

  const int warp = threadIdx.x >> 5;
  const uint32_t rank = cluster_ctarank();           // 0 = leader
  const int pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int m_pairs = (m_tiles + 1) / 2;             // pair tile mp covers m-tiles 2 mp and 2 mp + 1
  const int n_items = m_pairs * n_tiles;
  const int cchunks = p.C_in / kBlockK;
  const int total_k = p.taps * cchunks;

  pdl_launch_dependents();
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_w);
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int i = 0; i < Cfg::kStages; ++i) {
        mbar_init(&full_bar[i], 1);    // the leader's producer arrives with the byte count of BOTH CTAs
        mbar_init(&empty_bar[i], 1);   // one multicast commit from the leader's MMA thread
      }
      for (int i = 0; i < Cfg::kAccBufs; ++i) {
        mbar_init(&tmem_full[i], 1);   // one multicast commit
        mbar_init(&tmem_empty[i], 16); // 8 epilogue warps of each of the two CTAs
      }
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc_pair(tmem_ptr_smem, Cfg::kTmemCols);
    tmem_relinquish_pair();
  }
  if (threadIdx.x == 0) dbg_stamp(p, 0);
  tc_fence_before();
  cluster_sync_all();   // both CTAs' barriers exist before anything is signalled across the pair
  tc_fence_after();
  pdl_wait();
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (threadIdx.x == 0) dbg_stamp(p, 1);

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (elect_one()) {
      int stage = 0, issued = 0;
      uint32_t phase = 0;
      for (int item = pair_id; item < n_items; item += n_pairs) {
        const int mp = item % m_pairs, n_tile = item / m_pairs;
        const int m_tile = 2 * mp + static_cast<int>(rank);   // past the end for an odd m_tiles: TMA zero-fills
        const int n_img = (m_tile / p.hblocks) * p.bn;
        const int h0 = (m_tile % p.hblocks) * p.bh;
        const int b_row0 = n_tile * BLOCK_N + static_cast<int>(rank) * (BLOCK_N / 2);
        for (int k = 0; k < total_k; ++k) {
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
          const bool primed = issued >= Cfg::kStages;
          ++issued;
          const bool do_a = !(primed && (p.dbg_mode & 2)), do_b = !(primed && (p.dbg_mode & 1));
          if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * ((do_a ? Cfg::kABytes : 0) + (do_b ? Cfg::kBBytes : 0)));
          if (do_a) tma_load_4d_pair(a_dst, &tmap_x, &full_bar[stage], cc * kBlockK, ds, h0 + dr, n_img);
          if (!do_b) {
          } else if (!kDgrad) {
            tma_load_2d_pair(b_dst, &tmap_w, &full_bar[stage], tap * p.C_in + cc * kBlockK, b_row0);
          } else {
#pragma unroll
            for (int j = 0; j < BLOCK_N / 128; ++j)
              tma_load_2d_pair(b_dst + j * 8192, &tmap_w, &full_bar[stage],
                               (p.taps - 1 - tap) * p.C_out + b_row0 + j * 64, cc * kBlockK);
          }
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * kBlockM, BLOCK_N, 0, kDgrad ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int item = pair_id; item < n_items; item += n_pairs, ++local) {
        const int buf = local % Cfg::kAccBufs;
        const uint32_t use = static_cast<uint32_t>(local / Cfg::kAccBufs);
        mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);   // both CTAs drained this accumulator (free at first use)
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BLOCK_N;
        for (int k = 0; k < total_k; ++k) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(tiles + stage * Cfg::kStageBytes);
          const uint32_t b_addr = a_addr + Cfg::kABytes;
#pragma unroll
          for (int kk = 0; kk < kBlockK / kUmmaK; ++kk) {
            const uint64_t adesc = make_smem_desc_sw128(a_addr + kk * kUmmaK * 2, 16, 1024);
            const uint64_t bdesc = kDgrad ? make_smem_desc_sw128(b_addr + kk * kUmmaK * 128, 8192, 1024)
                                          : make_smem_desc_sw128(b_addr + kk * kUmmaK * 2, 16, 1024);
            umma_f16_pair(d_tmem, adesc, bdesc, idesc, (k | kk) != 0);
          }
          umma_commit_pair(&empty_bar[stage], 0x3);   // slot free in BOTH CTAs
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_pair(&tmem_full[buf], 0x3);       // accumulators ready in BOTH CTAs
        if (p.dbg != nullptr && item + n_pairs >= n_items) {
          mbar_wait(&tmem_full[buf], use & 1);
          dbg_stamp(p, 2);
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2..9, both CTAs, own 128 TMEM lanes) =====================
    const int ew = warp - 2;                 // 0..7
    const int q = warp & 3;                  // TMEM lane quarter
    const int half = ew >> 2;                // which half of the tile's columns
    constexpr int kChunks = BLOCK_N / 32;
    constexpr int kMyChunks = kChunks / 2;
    const int chunk0 = half * kMyChunks;
    float* my_bias = s_bias + ew * (BLOCK_N / 2);
    const int row = q * 32 + lane_id();
    const bool do_stats = (p.stats != nullptr);
    auto flush_stats = [&](int n_tile_done) {
      asm volatile("bar.sync 1, 256;" ::: "memory");   // the eight epilogue warps
      for (int c = threadIdx.x - 64; c < 2 * BLOCK_N; c += 256) {
        const int which = c / BLOCK_N, col = c - which * BLOCK_N;
        float acc = 0.f;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          acc += s_stats[(w * 2 + which) * BLOCK_N + col];
          s_stats[(w * 2 + which) * BLOCK_N + col] = 0.f;
        }
        atomicAdd(p.stats + which * p.C_out + n_tile_done * BLOCK_N + col, acc);
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
    };
    if (do_stats) {
      for (int c = threadIdx.x - 64; c < 8 * BLOCK_N; c += 256) s_stats[c] = 0.f;
      asm volatile("bar.sync 1, 256;" ::: "memory");
    }
    const uint32_t leader_empty0 = pp_map_to_cta(smem_u32(&tmem_empty[0]), 0);
    int cur_n = -1;
    int local = 0;
    for (int item = pair_id; item < n_items; item += n_pairs, ++local) {
      const int mp = item % m_pairs, n_tile = item / m_pairs;
      const int m_tile = 2 * mp + static_cast<int>(rank);
      if (do_stats && n_tile != cur_n) {
        if (cur_n >= 0) flush_stats(cur_n);
        cur_n = n_tile;
      }
      const int buf = local % Cfg::kAccBufs;
      const uint32_t use = static_cast<uint32_t>(local / Cfg::kAccBufs);
      const long long pix = static_cast<long long>(m_tile) * kBlockM + row;
      const bool valid = pix < p.P;
      if (p.bias != nullptr) {
#pragma unroll
        for (int c = 0; c < kMyChunks; ++c)
          my_bias[c * 32 + lane_id()] = __ldg(p.bias + n_tile * BLOCK_N + (chunk0 + c) * 32 + lane_id());
        __syncwarp();
      }
      mbar_wait(&tmem_full[buf], use & 1);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + buf * BLOCK_N + (static_cast<uint32_t>(q * 32) << 16);
      __nv_bfloat16* out_row = p.out + pix * p.C_out + n_tile * BLOCK_N;
#pragma unroll 1
      for (int c = 0; c < kMyChunks; ++c) {
        uint32_t raw[32];
        tmem_ld_32x32(t_addr + (chunk0 + c) * 32, raw);
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          float x = __uint_as_float(raw[j]);
          if (p.bias != nullptr) x += my_bias[c * 32 + j];
          v[j] = valid ? bf16_round(x) : 0.f;
        }
        if (valid) {
          uint4* dst = reinterpret_cast<uint4*>(out_row + (chunk0 + c) * 32);
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
        if (do_stats) {
          float sq[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) sq[j] = v[j] * v[j];
          const float s1 = warp_transpose_reduce(v);
          const float s2 = warp_transpose_reduce(sq);
          s_stats[(q * 2 + 0) * BLOCK_N + (chunk0 + c) * 32 + lane_id()] += s1;
          s_stats[(q * 2 + 1) * BLOCK_N + (chunk0 + c) * 32 + lane_id()] += s2;
        }
      }
      // accumulator fully read by this warp: tell the leader's MMA thread
      tc_fence_before();
      __syncwarp();
      if (lane_id() == 0) {
        if (rank == 0)
          mbar_arrive(&tmem_empty[buf]);
        else
          pp_arrive_remote(leader_empty0 + static_cast<uint32_t>(buf) * 8u);
      }
    }
    if (do_stats && cur_n >= 0) flush_stats(cur_n);
    if (threadIdx.x == 64) dbg_stamp(p, 4);
  }

  tc_fence_before();
  cluster_sync_all();   // nobody frees TMEM or exits while the peer may still signal / be signalled
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
  }

  // ===================== fused BatchNorm + ReLU (+ pool) =====================
  // Same scheme as in conv_persist.inl: all CTAs meet at a grid-wide barrier once the statistics
  // are complete (one CTA per SM, all co-resident), then every CTA normalises the 128-pixel tiles
  // IT wrote (m-tile 2 mp + rank of each of its cluster's pair tiles), which hold whole 2x2
  // pooling windows.  NOT YET RUN ON HARDWARE (the numerics tests of round 1 covered act == null).
  if (p.act != nullptr) {
    persist_grid_barrier(p.grid_sync, "pair conv+bn");
    if (threadIdx.x == 0) dbg_stamp(p, 5);
    float* s_scale = reinterpret_cast<float*>(tiles);   // the operand ring is idle now
    float* s_shift = s_scale + p.C_out;
    const float inv_count = 1.f / static_cast<float>(p.P);
    for (int c = threadIdx.x; c < p.C_out; c += blockDim.x) {
      const float mean = __ldcg(p.stats + c) * inv_count;
      float var = __ldcg(p.stats + p.C_out + c) * inv_count - mean * mean;
      var = fmaxf(var, 0.f);
      const float rstd = rsqrtf(var + p.eps);
      const float sc = p.gamma[c] * rstd;
      s_scale[c] = sc;
      s_shift[c] = p.beta[c] - mean * sc;
      if (blockIdx.x == 0 && p.saved != nullptr) {
        p.saved[c] = mean;
        p.saved[p.C_out + c] = rstd;
      }
    }
    __syncthreads();
    constexpr int kGroups = BLOCK_N / 8;                 // 16-byte channel groups per tile row
    constexpr int kUnroll = 4;                           // items per thread in flight
    const int Ho = p.pool ? p.H / 2 : p.H, Wo = p.pool ? p.W / 2 : p.W;
    const int out_rows = p.pool ? kBlockM / 4 : kBlockM; // output pixels per tile
    const int per_tile = out_rows * kGroups;
    const int nwin = p.pool ? 4 : 1;
    const int my_tiles = (n_items - pair_id + n_pairs - 1) / n_pairs;
    const long long my_total = static_cast<long long>(my_tiles) * per_tile;
    for (long long base = threadIdx.x; base < my_total; base += static_cast<long long>(blockDim.x) * kUnroll) {
      uint4 raw[kUnroll][4];
      long long opix[kUnroll];
      int chv[kUnroll];
      bool ok[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long flat = base + static_cast<long long>(u) * blockDim.x;
        ok[u] = flat < my_total;
        const int tile_local = static_cast<int>(flat / per_tile);
        const int w_item = static_cast<int>(flat - static_cast<long long>(tile_local) * per_tile);
        const int item = pair_id + tile_local * n_pairs;
        const int mp = item % m_pairs, n_tile = item / m_pairs;
        const int m_tile = 2 * mp + static_cast<int>(rank);
        const int n_img0 = (m_tile / p.hblocks) * p.bn;
        const int h0 = (m_tile % p.hblocks) * p.bh;
        const int cgp = w_item % kGroups;
        const int orow = w_item / kGroups;
        chv[u] = n_tile * BLOCK_N + cgp * 8;
        long long in00;
        if (!p.pool) {
          in00 = static_cast<long long>(m_tile) * kBlockM + orow;
          opix[u] = in00;
        } else {
          const int wo = orow % Wo;
          const int t2 = orow / Wo;
          const int ho = t2 % (p.bh / 2);
          const int n_img = n_img0 + t2 / (p.bh / 2);
          in00 = (static_cast<long long>(n_img) * p.H + h0 + 2 * ho) * p.W + 2 * wo;
          opix[u] = (static_cast<long long>(n_img) * Ho + (h0 / 2 + ho)) * Wo + wo;
        }
        ok[u] = ok[u] && in00 < p.P;
        if (ok[u]) {
#pragma unroll
          for (int d = 0; d < 4; ++d)
            if (d < nwin)
              raw[u][d] = __ldcg(reinterpret_cast<const uint4*>(p.out + (in00 + (d >> 1) * p.W + (d & 1)) * p.C_out + chv[u]));
        }
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        if (!ok[u]) continue;
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = 0.f;   // ReLU floor
#pragma unroll
        for (int d = 0; d < 4; ++d) {
          if (d < nwin) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint32_t wv = (j < 2) ? raw[u][d].x : (j < 4) ? raw[u][d].y : (j < 6) ? raw[u][d].z : raw[u][d].w;
              const float xv = __uint_as_float((j & 1) ? (wv & 0xffff0000u) : (wv << 16));
              o[j] = fmaxf(o[j], fmaf(xv, s_scale[chv[u] + j], s_shift[chv[u] + j]));
            }
          }
        }
        store8(p.act + opix[u] * p.C_out + chv[u], o);
      }
    }
  }
  if (threadIdx.x == 0) dbg_stamp(p, 3);
}

template <int BLOCK_N, bool kDgrad>
static int launch_pair_persist(const CUtensorMap& tx, const CUtensorMap& tw, const CUtensorMap& /*to*/,
                               const FpropParams& p, int m_tiles, int splits, cudaStream_t stream) {
  using Cfg = PairPersistCfg<BLOCK_N>;
  DDPB_REQUIRE(splits == 1, "conv: the persistent CTA-pair kernel does not split K");
  static bool configured = false;
  if (!configured) {
    DDPB_CHECK_CUDA(cudaFuncSetAttribute(conv_pair_persist_kernel<BLOCK_N, kDgrad>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    configured = true;
  }
  const int n_tiles = p.C_out / BLOCK_N;
  const int items = (m_tiles + 1) / 2 * n_tiles;
  int pairs = sm_count() / 2;
  if (pairs > items) pairs = items;
  if (pairs < 1) pairs = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs, 1, 1);
  cfg.blockDim = dim3(kPersistThreads);
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
  DDPB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, conv_pair_persist_kernel<BLOCK_N, kDgrad>, tx, tw, p, m_tiles, n_tiles));
  return 0;
}
