// Persistent convolution kernel (included by conv_fprop.cu).
//
// Measured on B200 (profiles/r1_call10_conv_timeline.txt): in the one-tile-per-CTA kernel a
// 128x128 tile spends 13.7 us in its main loop and 7.2 us in its epilogue, and a single CTA's
// TMA -> MMA ring is latency bound (~1.1 us round trip, 3 stages).  This kernel fixes both:
//   * one CTA per SM loops over tiles, so barrier / TMEM set-up happens once;
//   * the whole shared memory is one deep operand ring (~192 KB in flight per SM);
//   * two TMEM accumulators: while the eight epilogue warps drain tile i, the producer and
//     the MMA warp already run tile i+1;
//   * the epilogue reads the bias from shared memory, not from global.
// Warp roles (320 threads): 0 = TMA producer, 1 = MMA issuer + TMEM owner, 2..9 = epilogue;
// warps w and w+4 share a TMEM lane quarter and split the tile's columns between them.

constexpr int kPersistThreads = 320;
constexpr int kRingBytes = 192 * 1024;

template <int BLOCK_N>
struct PersistCfg {
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = BLOCK_N * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  // Epilogue staging: each of the 8 epilogue warps owns two swizzled boxes of 32 rows x kBoxCols bf16
  // columns.  A drained accumulator chunk goes registers -> box (16-byte shared stores, conflict
  // free under the TMA swizzle) -> one TMA store per box, instead of 32 lanes each storing 16 bytes
  // to 32 different rows of global memory (the epilogue was bound by those partial-line stores:
  // 2 us per 32-column chunk per warp, profiles/r2_call02_*).  The channel statistics are column
  // sums read back from the box (32 conflict-free shared loads per lane) instead of two 31-shuffle
  // transposes per chunk.
  static constexpr int kBoxCols = (BLOCK_N >= 128) ? 64 : 32;          // columns per box (128 / 64 byte rows)
  static constexpr int kBoxBytes = 32 * kBoxCols * 2;                  // 4 KB / 2 KB
  static constexpr int kStagingBytes = 8 * 2 * kBoxBytes;              // 64 KB / 32 KB
  static constexpr int kStatsBytes = 4 * 2 * BLOCK_N * 4;      // [quarter][sum, sumsq][column]
  static constexpr int kBiasBytes = 8 * (BLOCK_N / 2) * 4;     // warp-private copies of the bias slice
  static constexpr int kFixedBytes = 1024 + kStagingBytes + kStatsBytes + kBiasBytes + 512;
#ifdef DDPB_PERSIST_STAGES
  static constexpr int kStages = DDPB_PERSIST_STAGES;
#else
  // the rest of the 227 KB is the operand ring: 7 / 4 / 3 stages for BLOCK_N = 64 / 128 / 256.  The main
  // loop is bound by the tensor pipe's operand fetch, not by the loads (skipping them changes nothing,
  // profiles/r2_call02_*), so a shallower ring than round 1's 192 KB costs nothing.
  static constexpr int kStages = (227 * 1024 - kFixedBytes) / kStageBytes;
#endif
  static constexpr int kSmemBytes = kFixedBytes + kStages * kStageBytes;
  // accumulators in TMEM: as many as fit in 512 columns, at most 4.  With short reductions
  // (first layers: 1 or 9 k-steps per tile) the MMA -> epilogue -> MMA hand-off latency, not the
  // work, bounds a 2-deep ring; 4 buffers let the MMA warp run three tiles ahead.
  static constexpr int kAccBufs = (512 / BLOCK_N) < 4 ? (512 / BLOCK_N) : 4;
  static constexpr int kTmemCols = kAccBufs * BLOCK_N;         // 256 / 512 / 512 columns
};

struct PersistItem {
  int m_tile, n_tile, split;
};

__device__ __forceinline__ PersistItem persist_item(int item, int m_tiles, int n_tiles) {
  PersistItem it;
  it.m_tile = item % m_tiles;
  const int r = item / m_tiles;
  it.n_tile = r % n_tiles;
  it.split = r / n_tiles;
  return it;
}

// Grid-wide barrier for the persistent kernel (all CTAs are co-resident: at most one per SM).
// `word` must be zero when the kernel starts; each barrier of a launch needs its own word.
__device__ __forceinline__ void persist_grid_barrier(unsigned int* word, const char* what) {
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    atomicAdd(word, 1u);
    const unsigned long long t0 = global_timer_ns();
    unsigned int seen, spins = 0;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(word) : "memory");
      if ((++spins & 0x3FFu) == 0 && global_timer_ns() - t0 > DDPB_WAIT_TIMEOUT_NS) {
        printf("ddpb200: %s grid barrier timed out (block %d: %u of %u)\n", what, blockIdx.x, seen, gridDim.x);
        __trap();
      }
    } while (seen < gridDim.x);
    __threadfence();
  }
  __syncthreads();
}

// kFuseSplit: split-K partial sums are added up inside the kernel (one item per CTA; see FpropParams::split_sync).
// A separate instantiation: the ticket logic in the epilogue loop cost the first layer (14 short tiles per
// CTA, epilogue bound) 4 us when it was a run-time branch.
template <int BLOCK_N, bool kDgrad, bool kFuseSplit>
__global__ void __launch_bounds__(512, 1)   // 512, not 320: caps the kernel at 128 registers per thread
conv_persist_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
                    const __grid_constant__ CUtensorMap tmap_out, const FpropParams p, int m_tiles, int n_tiles,
                    int splits) {
  using Cfg = PersistCfg<BLOCK_N>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem_1k(smem_raw);
  uint8_t* tiles = smem;
  uint8_t* staging = smem + Cfg::kStages * Cfg::kStageBytes;    // 1024-byte aligned (stages are multiples of 8 KB)
  float* s_stats = reinterpret_cast<float*>(staging + Cfg::kStagingBytes);
  float* s_bias = reinterpret_cast<float*>(staging + Cfg::kStagingBytes + Cfg::kStatsBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + Cfg::kStagingBytes + Cfg::kStatsBytes + Cfg::kBiasBytes);
  uint64_t* full_bar = bars;                         // [kStages]
  uint64_t* empty_bar = bars + Cfg::kStages;         // [kStages]
  uint64_t* tmem_full = bars + 2 * Cfg::kStages;     // [kAccBufs]
  uint64_t* tmem_empty = tmem_full + Cfg::kAccBufs;  // [kAccBufs]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + Cfg::kAccBufs);

  const int warp = threadIdx.x >> 5;
  const int n_items = m_tiles * n_tiles * splits;
  const int cchunks = p.C_in / kBlockK;
  const int total_k = p.taps * cchunks;

  pdl_launch_dependents();
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_w);
    tma_prefetch_desc(&tmap_out);
  }
  if (warp == 1) {
    if (elect_one()) {
      for (int i = 0; i < Cfg::kStages; ++i) {
        mbar_init(&full_bar[i], 1);
        mbar_init(&empty_bar[i], 1);
      }
      for (int i = 0; i < Cfg::kAccBufs; ++i) {
        mbar_init(&tmem_full[i], 1);
        mbar_init(&tmem_empty[i], 8);   // one arrival per epilogue warp
      }
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_ptr_smem, Cfg::kTmemCols);
    tmem_relinquish();
  }
  if (threadIdx.x == 0) dbg_stamp(p, 0);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();   // barrier init / TMEM allocation above overlapped the previous kernel's tail
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (threadIdx.x == 0) dbg_stamp(p, 1);

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int stage = 0, issued = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const PersistItem it = persist_item(item, m_tiles, n_tiles);
        const int n_img = (it.m_tile / p.hblocks) * p.bn;
        const int h0 = (it.m_tile % p.hblocks) * p.bh;
        const int k_begin = static_cast<int>(static_cast<long long>(total_k) * it.split / splits);
        const int k_end = static_cast<int>(static_cast<long long>(total_k) * (it.split + 1) / splits);
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
          const bool primed = issued >= Cfg::kStages;
          ++issued;
          const bool do_a = !(primed && (p.dbg_mode & 2)), do_b = !(primed && (p.dbg_mode & 1));
          mbar_expect_tx(&full_bar[stage], (do_a ? Cfg::kABytes : 0) + (do_b ? Cfg::kBBytes : 0));
          if (do_a) tma_load_4d(a_dst, &tmap_x, &full_bar[stage], cc * kBlockK, ds, h0 + dr, n_img);
          if (!do_b) {
          } else if (!kDgrad) {
            tma_load_2d(b_dst, &tmap_w, &full_bar[stage], tap * p.C_in + cc * kBlockK, it.n_tile * BLOCK_N);
          } else {
#pragma unroll
            for (int j = 0; j < BLOCK_N / 64; ++j)
              tma_load_2d(b_dst + j * 8192, &tmap_w, &full_bar[stage],
                          (p.taps - 1 - tap) * p.C_out + it.n_tile * BLOCK_N + j * 64, cc * kBlockK);
          }
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(kBlockM, BLOCK_N, 0, kDgrad ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int local = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++local) {
        const PersistItem it = persist_item(item, m_tiles, n_tiles);
        const int k_begin = static_cast<int>(static_cast<long long>(total_k) * it.split / splits);
        const int k_end = static_cast<int>(static_cast<long long>(total_k) * (it.split + 1) / splits);
        const int buf = local % Cfg::kAccBufs;
        const uint32_t use = static_cast<uint32_t>(local / Cfg::kAccBufs);
        mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);   // epilogue drained this accumulator (free at first use)
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BLOCK_N;
        for (int k = k_begin; k < k_end; ++k) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(tiles + stage * Cfg::kStageBytes);
          const uint32_t b_addr = a_addr + Cfg::kABytes;
#pragma unroll
          for (int kk = 0; kk < kBlockK / kUmmaK; ++kk) {
            const uint64_t adesc = make_smem_desc_sw128(a_addr + kk * kUmmaK * 2, 16, 1024);
            const uint64_t bdesc = kDgrad ? make_smem_desc_sw128(b_addr + kk * kUmmaK * 128, 8192, 1024)
                                          : make_smem_desc_sw128(b_addr + kk * kUmmaK * 2, 16, 1024);
            umma_f16(d_tmem, adesc, bdesc, idesc, (k > k_begin) || (kk > 0));
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full[buf]);
        if (p.dbg != nullptr && item + static_cast<int>(gridDim.x) >= n_items) {
          mbar_wait(&tmem_full[buf], use & 1);
          dbg_stamp(p, 2);
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    const int ew = warp - 2;                 // 0..7
    const int q = warp & 3;                  // TMEM lane quarter
    const int half = ew >> 2;                // which half of the tile's columns
    constexpr int kChunks = BLOCK_N / 32;
    constexpr int kMyChunks = kChunks / 2;   // >= 1 (BLOCK_N >= 64)
    const int chunk0 = half * kMyChunks;
    float* my_bias = s_bias + ew * (BLOCK_N / 2);
    const int row = q * 32 + lane_id();
    // Per-channel statistics: every lane keeps running sums of "its" columns of each box in registers
    // across this CTA's tiles; they go to shared memory and on to global atomics only when the column
    // block changes (or at the end): with one flush per tile the 2*C_out atomic addresses were hit by
    // every tile of the layer and serialised in L2 (first layer: 2048 tiles onto 128 addresses).
    constexpr bool fuse_split = kFuseSplit;
    const bool do_stats = (p.partial == nullptr || fuse_split) && (p.stats != nullptr);
    constexpr int kSliceVec = kMyChunks * 8 * 32;             // 16-byte words of one warp's fp32 slice of a tile
    constexpr int kBoxCols = Cfg::kBoxCols;
    constexpr int kBoxes = kMyChunks * 32 / kBoxCols;        // boxes per tile and warp: 1 / 1 / 2
    constexpr int kChunksPerBox = kBoxCols / 32;             // 1 / 2 / 2
    constexpr int kColsPerLane = kBoxCols / 32;              // columns of a box whose sums a lane keeps
    constexpr int kRowBytes = kBoxCols * 2;                  // 64 / 128: the TMA swizzle span
    float racc1[kBoxes][kColsPerLane], racc2[kBoxes][kColsPerLane];
#pragma unroll
    for (int b = 0; b < kBoxes; ++b)
#pragma unroll
      for (int j = 0; j < kColsPerLane; ++j) racc1[b][j] = racc2[b][j] = 0.f;
    const uint32_t my_boxes = smem_u32(staging) + static_cast<uint32_t>(ew) * 2u * Cfg::kBoxBytes;
    // byte offset of 16-byte chunk k of row r inside a box, as the TMA swizzle lays it out
    auto box_off = [&](int r, int k) -> uint32_t {
      const int sw = (kRowBytes == 128) ? (r & 7) : ((r >> 1) & 3);
      return static_cast<uint32_t>(r * kRowBytes + ((k ^ sw) << 4));
    };
    uint32_t box_count = 0;
    auto flush_stats = [&](int n_tile_done) {
#pragma unroll
      for (int b = 0; b < kBoxes; ++b)
#pragma unroll
        for (int j = 0; j < kColsPerLane; ++j) {
          const int col = chunk0 * 32 + b * kBoxCols + static_cast<int>(lane_id()) * kColsPerLane + j;
          s_stats[(q * 2 + 0) * BLOCK_N + col] += racc1[b][j];
          s_stats[(q * 2 + 1) * BLOCK_N + col] += racc2[b][j];
          racc1[b][j] = racc2[b][j] = 0.f;
        }
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
    int cur_n = -1;
    int local = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++local) {
      const PersistItem it = persist_item(item, m_tiles, n_tiles);
      if (do_stats && it.n_tile != cur_n) {
        if (cur_n >= 0) flush_stats(cur_n);
        cur_n = it.n_tile;
      }
      const int buf = local % Cfg::kAccBufs;
      const uint32_t use = static_cast<uint32_t>(local / Cfg::kAccBufs);
      const long long pix = static_cast<long long>(it.m_tile) * kBlockM + row;
      const bool valid = pix < p.P;
      bool direct = (p.partial == nullptr);
      unsigned int* ticket = nullptr;
      const int out_tile = it.n_tile * m_tiles + it.m_tile;
      if (fuse_split) {
        ticket = p.split_sync + (static_cast<long long>(out_tile) * 8 + ew) * 2;
        unsigned int order = 0;
        if (lane_id() == 0) order = atomicAdd(ticket, 1u);
        order = __shfl_sync(0xffffffffu, order, 0);
        direct = (order == static_cast<unsigned int>(splits - 1));     // last of this (tile, warp): sums and finishes it
      }
      const bool reduce_here = fuse_split && direct;
      // stage this warp's slice of the bias while the accumulator is still being produced
      if (direct && p.bias != nullptr) {
#pragma unroll
        for (int c = 0; c < kMyChunks; ++c)
          my_bias[c * 32 + lane_id()] = __ldg(p.bias + it.n_tile * BLOCK_N + (chunk0 + c) * 32 + lane_id());
        __syncwarp();
      }
      mbar_wait(&tmem_full[buf], use & 1);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + buf * BLOCK_N + (static_cast<uint32_t>(q * 32) << 16);
      if (!direct && fuse_split) {
        // this warp's 32 rows x BLOCK_N/2 columns, laid out [chunk][16-byte word][lane]: every store (and every load of
        // the warp that sums) is one contiguous 512-byte run
        uint4* dst = reinterpret_cast<uint4*>(p.partial) +
                     ((static_cast<long long>(it.split) * (m_tiles * n_tiles) + out_tile) * 8 + ew) * kSliceVec + lane_id();
#pragma unroll 1
        for (int c = 0; c < kMyChunks; ++c) {
          uint32_t raw[32];
          tmem_ld_32x32(t_addr + (chunk0 + c) * 32, raw);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            dst[(c * 8 + j) * 32] = make_uint4(raw[4 * j], raw[4 * j + 1], raw[4 * j + 2], raw[4 * j + 3]);
        }
        __threadfence();          // every lane's slice stores before the count below
        __syncwarp();
        if (lane_id() == 0) atomicAdd(ticket + 1, 1u);
      } else if (!direct) {
        // two-kernel split-K: fp32 partial tiles [split][P][C_out], written as 32 x 32 boxes through shared memory
        // and TMA stores (tmap_out is the 3-d fp32 map of the partial buffer in this mode).  Direct stores were
        // 32 lanes x 16 bytes to 32 different rows per instruction, as in the old wgrad epilogue.
        constexpr bool kTwoBufs = Cfg::kBoxBytes >= 4096;     // a 32 x 32 fp32 box is 4 KB: two fit only for BLOCK_N >= 128
        const int r = static_cast<int>(lane_id());
#pragma unroll 1
        for (int c = 0; c < kMyChunks; ++c) {
          uint32_t raw[32];
          tmem_ld_32x32(t_addr + (chunk0 + c) * 32, raw);
          tmem_ld_wait();
          const uint32_t box = my_boxes + (kTwoBufs ? (box_count & 1u) * 4096u : 0u);
          ++box_count;
          if (lane_id() == 0) {
            if (kTwoBufs) bulk_wait_read<1>(); else bulk_wait_read<0>();
          }
          __syncwarp();
#pragma unroll
          for (int k = 0; k < 8; ++k)
            st_shared_v4(box + r * 128 + ((k ^ (r & 7)) << 4), raw[4 * k], raw[4 * k + 1], raw[4 * k + 2], raw[4 * k + 3]);
          fence_proxy_async();
          __syncwarp();
          if (lane_id() == 0) {
            // rows past P are clipped by the map (per split: the split index is its own dimension)
            tma_store_3d_saddr(&tmap_out, box, it.n_tile * BLOCK_N + (chunk0 + c) * 32, it.m_tile * kBlockM + q * 32, it.split);
            bulk_commit();
          }
        }
      } else {
        if (reduce_here) {
          if (lane_id() == 0) {
            const unsigned long long t0 = global_timer_ns();
            unsigned int seen, spins = 0;
            do {
              asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(ticket + 1) : "memory");
              if ((++spins & 0x3FFu) == 0 && global_timer_ns() - t0 > DDPB_WAIT_TIMEOUT_NS) {
                printf("ddpb200: split-K reduction timed out (block %d, tile %d: %u of %d slices)\n", blockIdx.x, out_tile,
                       seen, splits - 1);
                __trap();
              }
            } while (seen < static_cast<unsigned int>(splits - 1));
            ticket[0] = 0u;       // nobody else touches the pair any more in this launch: leave it zero for the next
            ticket[1] = 0u;
          }
          __syncwarp();
          __threadfence();
        }
        const uint4* slices = reinterpret_cast<const uint4*>(p.partial) + (static_cast<long long>(out_tile) * 8 + ew) * kSliceVec + lane_id();
        const long long slice_stride = static_cast<long long>(m_tiles) * n_tiles * 8 * kSliceVec;    // between splits
#pragma unroll 1
        for (int b = 0; b < kBoxes; ++b) {
          const uint32_t box = my_boxes + (box_count & 1u) * Cfg::kBoxBytes;
          ++box_count;
          // the TMA store issued from this buffer two boxes ago must have finished READING it
          if (lane_id() == 0) bulk_wait_read<1>();
          __syncwarp();
          const int r = static_cast<int>(lane_id());
#pragma unroll
          for (int cc = 0; cc < kChunksPerBox; ++cc) {
            const int c = b * kChunksPerBox + cc;
            uint32_t raw[32];
            if (!reduce_here) {
              tmem_ld_32x32(t_addr + (chunk0 + c) * 32, raw);
              tmem_ld_wait();
            } else {
              // sum over the splits in split order, this CTA's own accumulator at its position
              float acc[32];
#pragma unroll
              for (int j = 0; j < 32; ++j) acc[j] = 0.f;
              auto add_slice = [&](int sp, float (&dst)[32]) {
                const uint4* src = slices + sp * slice_stride + c * 8 * 32;
                uint4 v[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = __ldcg(src + j * 32);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  dst[4 * j] += __uint_as_float(v[j].x);
                  dst[4 * j + 1] += __uint_as_float(v[j].y);
                  dst[4 * j + 2] += __uint_as_float(v[j].z);
                  dst[4 * j + 3] += __uint_as_float(v[j].w);
                }
              };
#pragma unroll 1
              for (int sp = 0; sp < it.split; ++sp) add_slice(sp, acc);
              tmem_ld_32x32(t_addr + (chunk0 + c) * 32, raw);
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) acc[j] += __uint_as_float(raw[j]);
#pragma unroll 1
              for (int sp = it.split + 1; sp < splits; ++sp) add_slice(sp, acc);
#pragma unroll
              for (int j = 0; j < 32; ++j) raw[j] = __float_as_uint(acc[j]);
            }
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              float x0 = __uint_as_float(raw[2 * j]), x1 = __uint_as_float(raw[2 * j + 1]);
              if (p.bias != nullptr) {
                const float2 bb = *reinterpret_cast<const float2*>(my_bias + c * 32 + 2 * j);   // one broadcast load
                x0 += bb.x;
                x1 += bb.y;
              }
              pk[j] = valid ? pack_bf16x2(x0, x1) : 0u;   // rows past the end add zeros to the channel sums
            }
#pragma unroll
            for (int k = 0; k < 4; ++k)
              st_shared_v4(box + box_off(r, cc * 4 + k), pk[4 * k], pk[4 * k + 1], pk[4 * k + 2], pk[4 * k + 3]);
          }
          fence_proxy_async();      // this lane's generic-proxy writes -> visible to the TMA engine
          __syncwarp();
          if (lane_id() == 0) {
            // rows past P are clipped by the tensor map, so a ragged last tile needs no masking
            tma_store_2d_saddr(&tmap_out, box, it.n_tile * BLOCK_N + chunk0 * 32 + b * kBoxCols,
                               it.m_tile * kBlockM + q * 32);
            bulk_commit();
          }
          if (do_stats) {
            // column sums of the box just written (the stored, rounded values: what BatchNorm will see)
            const int word = (kColsPerLane == 2) ? r : (r >> 1);      // 4-byte word of the row this lane reads
            const int k16 = word >> 2, w4 = (word & 3) << 2;
#pragma unroll 8
            for (int rr = 0; rr < 32; ++rr) {
              const uint32_t wv = ld_shared_u32(box + box_off(rr, k16) + w4);
              if (kColsPerLane == 2) {
                const float lo = __uint_as_float(wv << 16), hi = __uint_as_float(wv & 0xffff0000u);
                racc1[b][0] += lo;
                racc2[b][0] = fmaf(lo, lo, racc2[b][0]);
                racc1[b][kColsPerLane - 1] += hi;
                racc2[b][kColsPerLane - 1] = fmaf(hi, hi, racc2[b][kColsPerLane - 1]);
              } else {
                const float x = __uint_as_float((r & 1) ? (wv & 0xffff0000u) : (wv << 16));
                racc1[b][0] += x;
                racc2[b][0] = fmaf(x, x, racc2[b][0]);
              }
            }
          }
        }
      }
      // accumulator fully read: hand it back to the MMA warp before the (slower) statistics flush
      tc_fence_before();
      __syncwarp();
      if (lane_id() == 0) mbar_arrive(&tmem_empty[buf]);
    }
    // every box this warp handed to the TMA engine is in global memory before anyone (the BatchNorm
    // tail after the grid barrier, or the next kernel) reads the tiles back
    if (lane_id() == 0) {
      bulk_wait<0>();
      fence_proxy_async_all();
    }
    __syncwarp();
    if (do_stats && cur_n >= 0) flush_stats(cur_n);
    if (threadIdx.x == 64) dbg_stamp(p, 4);
  }

  const bool tail_tmem = (p.act != nullptr) && (p.tail_tmem != 0);   // the accumulators are read once more below
  tc_fence_before();
  __syncthreads();
  if (warp == 1 && !tail_tmem) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }

  // ===================== fused BatchNorm + ReLU (+ pool) =====================
  // The conv output of this layer is only ever consumed by its BatchNorm, whose statistics
  // were accumulated by the epilogues above.  Instead of a second kernel, all CTAs meet at a
  // grid-wide barrier (they are co-resident: one per SM), turn the global sums into
  // scale/shift and then normalise the tiles THEY just wrote (still in L2).  Every 128-pixel
  // tile holds whole 2x2 pooling windows, so no CTA needs another CTA's pixels.
  if (p.act != nullptr) {
    persist_grid_barrier(p.grid_sync, "conv+bn");
    if (threadIdx.x == 0) dbg_stamp(p, 5);
    // the operand ring is idle now: reuse it for the per-channel coefficients
    float* s_scale = reinterpret_cast<float*>(tiles);
    float* s_shift = s_scale + p.C_out;
    const float inv_count = 1.f / static_cast<float>(p.P);
    for (int c = threadIdx.x; c < p.C_out; c += blockDim.x) {
      const float mean = __ldcg(p.stats + c) * inv_count;
      float var = __ldcg(p.stats + p.C_out + c) * inv_count - mean * mean;
      var = fmaxf(var, 0.f);
      const float rstd = rsqrtf(var + p.eps);
      const float sc = p.gamma[c] * rstd;
      if (tail_tmem) {     // {scale, shift} of a column pair in one 16-byte word (one broadcast load per pair)
        s_scale[(c >> 1) * 4 + (c & 1) * 2] = sc;
        s_scale[(c >> 1) * 4 + (c & 1) * 2 + 1] = p.beta[c] - mean * sc;
      } else {
        s_scale[c] = sc;
        s_shift[c] = p.beta[c] - mean * sc;
      }
      if (blockIdx.x == 0 && p.saved != nullptr) {
        p.saved[c] = mean;
        p.saved[p.C_out + c] = rstd;
      }
    }
    __syncthreads();
    constexpr int kGroups = BLOCK_N / 8;                 // 16-byte channel groups per tile row
    const int Ho = p.pool ? p.H / 2 : p.H, Wo = p.pool ? p.W / 2 : p.W;
    const int out_rows = p.pool ? kBlockM / 4 : kBlockM; // output pixels per tile
    const int per_tile = out_rows * kGroups;
    // flat index over ALL of this CTA's tiles, so a thread always has kUnroll independent items
    // in flight (a single tile is only 256..1024 items: fewer than the block has threads)
    const int out_tiles = m_tiles * n_tiles;             // (with split-K, several items share a tile)
    const int my_tiles = (out_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
    // without split-K a CTA normalises the tiles it wrote; with it (several CTAs per tile, all tiles final after the
    // grid barrier) the flat item range of the whole layer is cut into one contiguous piece per CTA
    constexpr bool spread = kFuseSplit;   // (the tail only runs with one split or with the in-kernel sum)
    const long long all_total = static_cast<long long>(out_tiles) * per_tile;
    const long long flat_lo = spread ? all_total * blockIdx.x / gridDim.x : 0;
    const long long my_total = spread ? all_total * (blockIdx.x + 1) / gridDim.x - flat_lo : static_cast<long long>(my_tiles) * per_tile;
    // The pass is latency bound (the tiles come back from L2), so what counts is 16-byte loads in flight per
    // thread: 4 items x 4 window loads with pooling, 8 items x 1 load without (measured: the un-pooled
    // layers ran this pass at 1.8 TB/s with 4 loads in flight, profiles/r2_call15_*)
    auto bn_tail_pass = [&](auto unroll_tag, auto win_tag) {
      constexpr int kUnroll = decltype(unroll_tag)::value;   // items per thread in flight
      constexpr int kWinLoads = decltype(win_tag)::value;    // loads per item
      for (long long base = threadIdx.x; base < my_total; base += static_cast<long long>(blockDim.x) * kUnroll) {
        uint4 raw[kUnroll][kWinLoads];
        long long opix[kUnroll];
        int chv[kUnroll];
        bool ok[kUnroll];
        // 1) addresses + all loads of up to kUnroll items
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const long long flat = base + static_cast<long long>(u) * blockDim.x;
          ok[u] = flat < my_total;
          const int tile_local = static_cast<int>((flat_lo + flat) / per_tile);
          const int w_item = static_cast<int>(flat_lo + flat - static_cast<long long>(tile_local) * per_tile);
          const PersistItem it = persist_item(spread ? tile_local : blockIdx.x + tile_local * gridDim.x, m_tiles, n_tiles);
          const int n_img0 = (it.m_tile / p.hblocks) * p.bn;
          const int h0 = (it.m_tile % p.hblocks) * p.bh;
          const int cgp = w_item % kGroups;
          const int orow = w_item / kGroups;
          chv[u] = it.n_tile * BLOCK_N + cgp * 8;
          long long in00;
          if (!p.pool) {
            in00 = static_cast<long long>(it.m_tile) * kBlockM + orow;
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
            for (int d = 0; d < kWinLoads; ++d)
              raw[u][d] = __ldcg(reinterpret_cast<const uint4*>(p.out + (in00 + (d >> 1) * p.W + (d & 1)) * p.C_out + chv[u]));
          }
        }
        // 2) normalise, rectify, pool, store
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          if (!ok[u]) continue;
          float o[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = 0.f;   // ReLU floor
#pragma unroll
          for (int d = 0; d < kWinLoads; ++d) {
            {
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
    };
    if (tail_tmem) {
      // Every tile of this CTA still sits in its own TMEM buffer: the eight epilogue warps read their rows
      // again (+bias, rounded to bf16 exactly as stored), normalise, rectify and pool across lanes -- the 32
      // rows of a warp are whole 2x2 windows: lanes l, l^1 (next column), l^W (next image row).  Same values
      // as the pass over the stored tiles below, without the trip through L2.
      if (warp >= 2) {
        tc_fence_after();
        const int ew = warp - 2, q = warp & 3, half = ew >> 2;
        constexpr int kMyChunks = BLOCK_N / 64;
        const int chunk0 = half * kMyChunks;
        const int lane = static_cast<int>(lane_id());
        const int row = q * 32 + lane;
        const bool writer = !p.pool || ((lane & 1) == 0 && (lane & p.W) == 0);
        const int Ho = p.H / 2, Wo = p.W / 2;
        float* my_bias = s_bias + ew * (BLOCK_N / 2);
        int local = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++local) {
          const PersistItem it = persist_item(item, m_tiles, n_tiles);
          const int pix = it.m_tile * kBlockM + row;      // (P is an int)
          const bool valid = pix < p.P;
          long long opix = pix;
          if (p.pool) {
            const int w = pix % p.W, t = pix / p.W;
            const int h = t % p.H;
            opix = (static_cast<long long>(t / p.H) * Ho + h / 2) * Wo + w / 2;
          }
          __syncwarp();
#pragma unroll
          for (int c = 0; c < kMyChunks; ++c)
            my_bias[c * 32 + lane] = (p.bias != nullptr) ? __ldg(p.bias + it.n_tile * BLOCK_N + (chunk0 + c) * 32 + lane) : 0.f;
          __syncwarp();
          const uint32_t t_addr = tmem_base + local * BLOCK_N + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
          for (int c = 0; c < kMyChunks; ++c) {
            const int col0 = it.n_tile * BLOCK_N + (chunk0 + c) * 32;
            uint32_t raw[32];
            tmem_ld_32x32(t_addr + (chunk0 + c) * 32, raw);
            tmem_ld_wait();
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              // the shared-memory pipe is what bounds this pass: one 8-byte and one 16-byte broadcast load per pair
              const float2 bb = *reinterpret_cast<const float2*>(my_bias + c * 32 + 2 * j);
              const float4 cf = *reinterpret_cast<const float4*>(s_scale + (col0 + 2 * j) * 2);
              const uint32_t xb = pack_bf16x2(__uint_as_float(raw[2 * j]) + bb.x, __uint_as_float(raw[2 * j + 1]) + bb.y);
              const float y0 = fmaxf(fmaf(__uint_as_float(xb << 16), cf.x, cf.y), 0.f);
              const float y1 = fmaxf(fmaf(__uint_as_float(xb & 0xffff0000u), cf.z, cf.w), 0.f);
              pk[j] = pack_bf16x2(y0, y1);
            }
            if (p.pool) {        // rounding is monotonic: the max of the rounded values is the rounded max
#pragma unroll
              for (int j = 0; j < 16; ++j) pk[j] = bf16x2_max_u32(pk[j], __shfl_xor_sync(0xffffffffu, pk[j], 1));
#pragma unroll
              for (int j = 0; j < 16; ++j) pk[j] = bf16x2_max_u32(pk[j], __shfl_xor_sync(0xffffffffu, pk[j], p.W));
            }
            if (valid && writer) {
              uint4* dst = reinterpret_cast<uint4*>(p.act + opix * p.C_out + col0);
#pragma unroll
              for (int k = 0; k < 4; ++k) dst[k] = make_uint4(pk[4 * k], pk[4 * k + 1], pk[4 * k + 2], pk[4 * k + 3]);
            }
          }
        }
        tc_fence_before();
      }
      __syncthreads();
      if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, Cfg::kTmemCols);
      }
    } else if (p.pool)
      bn_tail_pass(std::integral_constant<int, 4>{}, std::integral_constant<int, 4>{});
    else
      bn_tail_pass(std::integral_constant<int, 8>{}, std::integral_constant<int, 1>{});
  }
  if (threadIdx.x == 0) dbg_stamp(p, 3);
}

template <int BLOCK_N, bool kDgrad, bool kFuseSplit>
static int launch_persist_impl(const CUtensorMap& tx, const CUtensorMap& tw, const CUtensorMap& to, const FpropParams& p,
                               int m_tiles, int splits, cudaStream_t stream) {
  using Cfg = PersistCfg<BLOCK_N>;
  static bool configured = false;
  if (!configured) {
    DDPB_CHECK_CUDA(cudaFuncSetAttribute(conv_persist_kernel<BLOCK_N, kDgrad, kFuseSplit>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    configured = true;
  }
  const int n_tiles = p.C_out / BLOCK_N;
  const int items = m_tiles * n_tiles * splits;
  int grid = sm_count();
  if (grid > items) grid = items;
  DDPB_REQUIRE(!kFuseSplit || (grid == items && p.partial != nullptr && p.split_sync != nullptr),
               "conv: the in-kernel split-K sum needs one item per CTA, scratch and sync words");
  DDPB_CHECK_CUDA(launch_pdl(conv_persist_kernel<BLOCK_N, kDgrad, kFuseSplit>, grid, kPersistThreads, Cfg::kSmemBytes,
                             stream, tx, tw, to, p, m_tiles, n_tiles, splits));
  return 0;
}

template <int BLOCK_N, bool kDgrad>
static int launch_persist(const CUtensorMap& tx, const CUtensorMap& tw, const CUtensorMap& to, const FpropParams& p,
                          int m_tiles, int splits, cudaStream_t stream) {
  if (p.split_sync != nullptr && splits > 1)
    return launch_persist_impl<BLOCK_N, kDgrad, true>(tx, tw, to, p, m_tiles, splits, stream);
  return launch_persist_impl<BLOCK_N, kDgrad, false>(tx, tw, to, p, m_tiles, splits, stream);
}
