// binning_buckets.cu -- MSD-bucket binning: one global partition pass on the top tile bits, then
// every bucket is finished by ONE CTA.
//
// Same result as duplicateWithKeys + cub::DeviceRadixSort::SortPairs + identifyTileRanges of the
// reference (DGR/cuda_rasterizer/rasterizer_impl.cu:70-138, 290-319): the unique ordering by
// (tile, depth bits, Gaussian index) -- see binning_packed.cu for why that is what a stable sort
// of the reference's key list produces.
//
// Why: at the sizes of this path (R ~ 1e7 instances -- L2-resident) an LSD radix sort is not
// bandwidth-bound but serialised by its passes: every onesweep pass chains a few thousand CTAs
// through a decoupled look-back and all passes go through global memory. Here only the FIRST digit
// (the top <= 10 bits of the tile id) is a global pass. It cuts the list into <= 1024 contiguous
// buckets of a few adjacent tiles each, and one CTA finishes a bucket.
//
// Two forms of the instance word:
//   presorted   the Gaussians were put in front-to-back order first (launch_depth_order), the
//               instances are emitted in that order and every pass is stable, so a word is just
//               tile | index -- 32 bits when they fit (1M Gaussians at 1280x720: 12 + 20), else 64
//               -- and a bucket only needs a stable split on the <= 7 tile bits below the bucket
//               digit (bucket_split_kernel, words in registers). No limit on the number of
//               Gaussians: the depth never has to share the word with the index.
//   full        tile | depth bits | index in 64 bits (when they fit); bucket_sort_kernel orders the
//               remaining 32 + few bits of a bucket with 9-bit LSD passes in shared memory. Used for
//               small frames, where sorting the Gaussians first costs more than it saves, and when
//               the presort gave up (a depth outside its 30-bit key range).
// Either way the CTA that finishes a bucket also writes its tiles' ranges (identifyTileRanges
// disappears) and the bare index list. The sorted 64-bit keys themselves are not written: no kernel
// reads them (launch_binning_keys rebuilds them for the parity tests).
#include "common.cuh"

namespace egs {

constexpr int MSD_BITS_MAX = 10;
constexpr int MSD_BINS = 1 << MSD_BITS_MAX;        // 1024
constexpr int MSD_THREADS = 256;
constexpr int MSD_WARPS = MSD_THREADS / 32;
constexpr int MSD_DPT = MSD_BINS / MSD_THREADS;    // digits owned by one thread (4)
// words per thread of the partition kernel: 32 KB of exchange buffer per CTA either way
template <typename WORD>
struct MsdCfg {
  static constexpr int ITEMS = sizeof(WORD) == 8 ? 16 : 32;
  static constexpr int TILE = MSD_THREADS * ITEMS;   // 4096 64-bit or 8192 32-bit words per CTA
};
static_assert(MsdCfg<uint64_t>::TILE == SORT_TILE, "status sizing");
static_assert(MSD_BINS <= SORT_STATUS_WORDS, "status sizing");

constexpr int BK_THREADS = 1024;
constexpr int BK_WARPS = BK_THREADS / 32;
constexpr int BK_BITS = 9;
constexpr int BK_RADIX = 1 << BK_BITS;             // 512
constexpr int BK_CAP = 12288;                      // words per bucket held in shared memory (2 buffers)

// ---------------------------------------------------------------------------------------------
// duplicate, then bucket histogram
// ---------------------------------------------------------------------------------------------
// One packed word (tile | depth bits | index) per instance, in Gaussian-index order
// (duplicateWithKeys, rasterizer_impl.cu:70-111), and the bucket histogram of those words.
//
// Warp-cooperative: a warp takes 32 consecutive Gaussians. Their instances form one contiguous run
// of the output (the offsets are a prefix sum), and the lanes share that run evenly: lane l writes
// output slots base + 32 c + l. A slot finds its Gaussian by a 5-step binary search over the 32
// start offsets (shuffles) and its tile from the slot's position inside that Gaussian's rectangle.
// Compared with one thread looping over its own rectangle this removes the divergence between
// small and large rectangles and makes every store coalesced; the bucket histogram is counted on the
// way (count_bucket below), so it needs no pass of its own.
//
// words of GeomPtrs::sort_aux (depth presort): three digit histograms, their exclusive scans, tickets,
// meta (2: P for the partition kernel, 4: count of set-aside Gaussians), then three status areas
constexpr int DO_HIST = 0, DO_START = 3 * 1024, DO_TICKET = 6 * 1024, DO_META = DO_TICKET + 8,
              DO_STATUS = DO_META + 8;
static_assert(DO_STATUS == DEPTH_ORDER_AUX_WORDS, "aux layout");
constexpr int DUP_THREADS = 256;
// CTAs per SM of the duplicate kernel: the owner search is a chain of dependent shuffles, so the kernel
// wants every warp slot (8 x 256 threads; 4 CTAs measured 0.096 against 0.090 ms at C3)
constexpr int DUP_CTAS_PER_SM = 8;
// A Gaussian covering more tiles than this is left to duplicate_big_kernel (front-to-back order only):
// the nearest Gaussians are the largest on screen, they all sit in the first groups of that order, and
// a warp that owns 32 of them would emit over a hundred thousand words alone (measured: the kernel
// went from 0.08 to 0.18 ms on two of the eight bench views).
constexpr uint32_t DUP_BIG = 512;

// One shared-memory reduction per word into the CTA's bucket histogram. No match.any to count equal
// lanes once: on sm_100 a match.any occupies its SM for about two cycles per DISTINCT value in the warp
// (63 cycles for 32 distinct 10-bit values, 15 for 8; tools/micro/match_bench.cu), several times what
// the atomics it saved cost (duplicate_words_kernel at C3: 0.077 -> 0.047 ms).
__device__ __forceinline__ void count_bucket(uint32_t* s_hist, bool live, uint32_t bucket) {
  if (live) atomicAdd(&s_hist[bucket], 1u);
}

// floor(n / d) for n * d < 2^32 with m = ceil(2^32 / d) (d >= 2; d == 1 is passed as m == 0)
__device__ __forceinline__ uint32_t div_by(uint32_t n, uint32_t m) { return m ? __umulhi(n, m) : n; }
__device__ __forceinline__ uint32_t div_magic(uint32_t d) { return d > 1u ? 0xffffffffu / d + 1u : 0u; }

// DEPTH: tile | depth bits | index (64-bit words only); otherwise tile | index
template <typename WORD, bool DEPTH>
__device__ __forceinline__ WORD make_word(uint32_t tile, uint32_t dbits, uint32_t idx, int idx_bits) {
  if (DEPTH) return (WORD)(((uint64_t)tile << (32 + idx_bits)) | ((uint64_t)dbits << idx_bits) | idx);
  return (WORD)(((WORD)tile << idx_bits) | (WORD)idx);
}

template <typename WORD, bool DEPTH>
__global__ void __launch_bounds__(DUP_THREADS)
duplicate_words_kernel(int P, const float2* __restrict__ means2D, const float* __restrict__ depths,
                       const uint32_t* __restrict__ offsets, const int* __restrict__ radii,
                       const uint32_t* __restrict__ order /* NULL: slot j of the scan is Gaussian j */,
                       WORD* __restrict__ words, int grid_x, int grid_y, int idx_bits,
                       InstanceCount count, int bucket_shift, uint32_t* __restrict__ bucket_hist,
                       uint32_t* __restrict__ big_list /* with order: scan slots left to duplicate_big_kernel */,
                       uint32_t* __restrict__ big_count) {
  static_assert(!DEPTH || sizeof(WORD) == 8, "depth bits need the 64-bit word");
  uint32_t n;
  if (!count.get(n)) return;   // more instances than the buffer holds: write nothing
  __shared__ uint32_t s_hist[MSD_BINS];
  for (int i = threadIdx.x; i < MSD_BINS; i += DUP_THREADS) s_hist[i] = 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int groups = (P + 31) / 32;
  for (int group = blockIdx.x * (DUP_THREADS / 32) + warp; group < groups; group += gridDim.x * (DUP_THREADS / 32)) {
    const int slot = group * 32 + (int)lane;
    const bool valid = slot < P;
    const int idx = !valid ? 0 : order ? (int)order[slot] : slot;
    const int r = valid ? radii[idx] : 0;
    // the rectangle as: first tile id, tiles per row (as the multiplier of div_by), tile-id step from one
    // row to the next minus the row length
    uint32_t first = 0, magic = 0, skip = 0, cnt = 0, dbits = 0;
    if (r > 0) {
      uint2 lo, hi;
      tile_rect(means2D[idx], r, grid_x, grid_y, lo, hi);
      const uint32_t w = hi.x - lo.x;
      first = lo.y * (uint32_t)grid_x + lo.x;
      magic = div_magic(w);
      skip = (uint32_t)grid_x - w;
      cnt = w * (hi.y - lo.y);
      if (DEPTH) dbits = __float_as_uint(depths[idx]);
    }
    // first output slot of this Gaussian; the warp's own runs, big ones taken out, are numbered
    // contiguously (`mine`) and shared evenly between the lanes
    const uint32_t start = (valid ? offsets[slot] : 0u) - cnt;
    if (big_list != nullptr && cnt > DUP_BIG) {
      big_list[atomicAdd(big_count, 1u)] = (uint32_t)slot;
      cnt = 0;
    }
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    const uint32_t mine = incl - cnt;
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    for (uint32_t o = lane; o - lane < total; o += 32) {   // warp-uniform trip count
      // owner = the last lane whose run starts at or before o (empty runs sort before it)
      int L = 0;
#pragma unroll
      for (int step = 16; step > 0; step >>= 1) {
        const uint32_t s_mid = __shfl_sync(0xffffffffu, mine, L + step);
        if (s_mid <= o) L += step;
      }
      const uint32_t o_mine = __shfl_sync(0xffffffffu, mine, L);
      const uint32_t o_start = __shfl_sync(0xffffffffu, start, L);
      const uint32_t o_first = __shfl_sync(0xffffffffu, first, L);
      const uint32_t o_magic = __shfl_sync(0xffffffffu, magic, L);
      const uint32_t o_skip = __shfl_sync(0xffffffffu, skip, L);
      const uint32_t o_d = DEPTH ? __shfl_sync(0xffffffffu, dbits, L) : 0u;
      const uint32_t o_idx = __shfl_sync(0xffffffffu, (uint32_t)idx, L);
      const bool live = o < total;
      const uint32_t local = live ? o - o_mine : 0u;
      // row dy, column local - dy w of the rectangle: tile = first + dy grid_x + (local - dy w)
      const uint32_t tile = o_first + local + div_by(local, o_magic) * o_skip;
      if (live) words[o_start + local] = make_word<WORD, DEPTH>(tile, o_d, o_idx, idx_bits);
      count_bucket(s_hist, live, tile >> bucket_shift);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < MSD_BINS; i += DUP_THREADS) {
    const uint32_t c = s_hist[i];
    if (c) atomicAdd(&bucket_hist[i], c);
  }
}

// The Gaussians duplicate_words_kernel set aside: one CTA per Gaussian at a time, its words shared
// between the 256 threads.
template <typename WORD, bool DEPTH>
__global__ void __launch_bounds__(DUP_THREADS)
duplicate_big_kernel(const float2* __restrict__ means2D, const float* __restrict__ depths,
                     const uint32_t* __restrict__ offsets, const int* __restrict__ radii,
                     const uint32_t* __restrict__ order, WORD* __restrict__ words, int grid_x, int grid_y,
                     int idx_bits, InstanceCount count, int bucket_shift, uint32_t* __restrict__ bucket_hist,
                     const uint32_t* __restrict__ big_list, const uint32_t* __restrict__ big_count) {
  uint32_t n;
  if (!count.get(n)) return;
  const uint32_t n_big = *big_count;
  if (blockIdx.x >= n_big) return;
  __shared__ uint32_t s_hist[MSD_BINS];
  for (int i = threadIdx.x; i < MSD_BINS; i += DUP_THREADS) s_hist[i] = 0;
  __syncthreads();
  for (uint32_t b = blockIdx.x; b < n_big; b += gridDim.x) {
    const uint32_t slot = big_list[b];
    const uint32_t idx = order[slot];
    uint2 lo, hi;
    tile_rect(means2D[idx], radii[idx], grid_x, grid_y, lo, hi);
    const uint32_t w = hi.x - lo.x, cnt = w * (hi.y - lo.y);
    const uint32_t first = lo.y * (uint32_t)grid_x + lo.x, magic = div_magic(w), skip = (uint32_t)grid_x - w;
    const uint32_t start = offsets[slot] - cnt;
    const uint32_t dbits = DEPTH ? __float_as_uint(depths[idx]) : 0u;
    for (uint32_t t = threadIdx.x; t - threadIdx.x < cnt; t += DUP_THREADS) {   // CTA-uniform trip count
      const bool live = t < cnt;
      const uint32_t local = live ? t : 0u;
      const uint32_t tile = first + local + div_by(local, magic) * skip;
      if (live) words[start + t] = make_word<WORD, DEPTH>(tile, dbits, idx, idx_bits);
      count_bucket(s_hist, live, tile >> bucket_shift);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < MSD_BINS; i += DUP_THREADS) {
    const uint32_t c = s_hist[i];
    if (c) atomicAdd(&bucket_hist[i], c);
  }
}

// ---------------------------------------------------------------------------------------------
// global stable partition by bucket (one onesweep pass, 10-bit digit)
// ---------------------------------------------------------------------------------------------
constexpr uint32_t MSD_FLAG_AGG = 1u << 30;
constexpr uint32_t MSD_FLAG_PREFIX = 1u << 31;
constexpr uint32_t MSD_VALUE_MASK = (1u << 30) - 1u;

template <typename WORD>
struct MsdSmem {
  uint32_t warp_hist[MSD_WARPS][MSD_BINS];   // 32 KB: per-warp digit counters (shared-memory atomics)
  WORD exch[MsdCfg<WORD>::TILE];             // 32 KB
  uint16_t digit_local[MSD_BINS];            // first slot of each digit in exch (< TILE)
  uint32_t digit_global[MSD_BINS];
  uint32_t warp_sums[MSD_WARPS];
  uint32_t hist_sums[MSD_WARPS];
  uint32_t tile;
};

// four adjacent status words in one request (each word is read atomically; the vector as a whole
// need not be)
__device__ __forceinline__ uint4 ld_relaxed_v4(const uint32_t* p) {
  uint4 v;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_v4(uint32_t* p, uint4 v) {
  asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// 256 threads x 16 (64-bit) or 32 (32-bit) words: the fixed per-CTA work (clearing and scanning the
// per-warp counters, the look-back) is shared by many words per thread (the first version ran 1024
// threads x 4 words and spent more instructions on that bookkeeping than on the words), and the
// 32-bit form halves both the bytes and the length of the look-back chain. A thread owns four
// adjacent digits and walks their four look-back chains together.
// The digit is (word >> shift) & mask: the top bits of the word for the bucket partition (mask all
// ones), any 10-bit field for the stable LSD passes of the depth presort. EMIT_IDX: the pass writes
// only the low index bits of every word, as 32-bit values (last pass of the presort).
// Lanes of the warp holding the same 10-bit digit as this one. match.any costs its SM ~2 cycles per
// distinct value among the lanes (tools/micro/match_bench.cu): cheap for the instance words, whose
// neighbours share tiles (~12 distinct digits in 32 words), 63 cycles for the depth keys of the presort,
// whose digits are as good as random. For those, ten ballots (2.3 cycles each) and a few logic
// operations give the same mask.
template <bool BALLOTS>
__device__ __forceinline__ uint32_t digit_peers(uint32_t d, bool ok, unsigned lane) {
  if (!BALLOTS) return __match_any_sync(0xffffffffu, ok ? d : (0x10000u + lane));
  uint32_t p = __ballot_sync(0xffffffffu, ok);
#pragma unroll
  for (int k = 0; k < MSD_BITS_MAX; ++k) {
    const bool bit = (d >> k) & 1u;
    const uint32_t b = __ballot_sync(0xffffffffu, bit);
    p &= bit ? b : ~b;
  }
  return ok ? p : (1u << lane);
}

template <typename WORD, bool EMIT_IDX, bool BALLOTS = false>
__global__ void __launch_bounds__(MSD_THREADS, 3)
msd_partition_kernel(const WORD* __restrict__ in, WORD* __restrict__ out, InstanceCount count,
                     int shift, uint32_t mask, const uint32_t* __restrict__ digit_hist /* [MSD_BINS] counts of the whole input */,
                     uint32_t* __restrict__ start_out /* NULL, or [MSD_BINS + 1]: their exclusive scan, for later kernels */,
                     uint32_t* __restrict__ ticket, uint32_t* __restrict__ status_raw,
                     uint32_t* __restrict__ out_idx, uint32_t idx_mask) {
  constexpr int ITEMS = MsdCfg<WORD>::ITEMS;
  constexpr int TILE = MsdCfg<WORD>::TILE;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  MsdSmem<WORD>& S = *reinterpret_cast<MsdSmem<WORD>*>(smem_raw);
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // Where every digit starts in the output: the exclusive scan of the histogram the producer counted.
  // Every CTA forms it for itself (1024 values, a few instructions per thread: no scan kernel in
  // between, no launch), CTA 0 also leaves it in memory for the kernels that finish the buckets. A frame
  // that is being skipped has an all-zero histogram, so they see empty buckets.
  uint32_t hcount[MSD_DPT], hstart;
  {
    const uint4 h = *reinterpret_cast<const uint4*>(digit_hist + tid * MSD_DPT);
    hcount[0] = h.x; hcount[1] = h.y; hcount[2] = h.z; hcount[3] = h.w;
    const uint32_t mine = h.x + h.y + h.z + h.w;
    uint32_t incl = mine;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, s);
      if (lane >= s) incl += t;
    }
    if (lane == 31) S.hist_sums[warp] = incl;
    hstart = incl - mine;
  }
  uint32_t n;
  const bool run = count.get(n);   // false: the guessed capacity was too small, forward re-runs this stage
  if (run && tid == 0) S.tile = atomicAdd(ticket, 1u);
  {
    uint4* z = reinterpret_cast<uint4*>(&S.warp_hist[0][0]);
    constexpr int VEC = (int)(sizeof(S.warp_hist) / sizeof(uint4));
#pragma unroll
    for (int i = 0; i < VEC / MSD_THREADS; ++i) z[i * MSD_THREADS + tid] = make_uint4(0u, 0u, 0u, 0u);
  }
  __syncthreads();
  for (unsigned w = 0; w < warp; ++w) hstart += S.hist_sums[w];
  if (blockIdx.x == 0 && start_out != nullptr) {
    uint32_t run_start = hstart;
#pragma unroll
    for (int k = 0; k < MSD_DPT; ++k) {
      start_out[tid * MSD_DPT + k] = run_start;
      run_start += hcount[k];
    }
    if (tid == MSD_THREADS - 1) start_out[MSD_BINS] = run_start;
  }
  if (!run) return;
  const uint32_t tile = S.tile;
  const unsigned lt_mask = (1u << lane) - 1u;
  const size_t block_base = (size_t)tile * TILE;
  // the grid is sized for the capacity; CTAs whose tile lies beyond the real count have nothing to
  // do and nobody looks back at them (look-backs only walk towards tile 0)
  if (block_base >= (size_t)n) return;
  const size_t warp_base = block_base + (size_t)warp * (32 * ITEMS);

  WORD word[ITEMS];
  uint32_t rank2[ITEMS / 2];   // two 16-bit ranks per register
  // words of this warp that exist (the last tile is ragged)
  const uint32_t lim = (size_t)n > warp_base ? (uint32_t)min((size_t)(32 * ITEMS), (size_t)n - warp_base) : 0u;
  {
    const WORD* src = in + warp_base + lane;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) word[i] = (uint32_t)(i * 32) + lane < lim ? src[i * 32] : ~(WORD)0;
  }
  // Rank of every word among the warp's words of the same digit, in load order. The lanes holding a
  // digit are found with match.any; the lowest of them adds their number to the warp's counter of
  // that digit with ONE shared-memory atomic and hands the old value to the others. The counter is
  // only touched by atomics of this warp, which take effect in issue order, so nothing has to wait
  // between two words: the rounds of a batch overlap instead of forming a chain of
  // read - synchronise - write - synchronise (32 rounds of ~130 cycles each, which was most of the
  // kernel's time).
  constexpr int BATCH = 8;
  uint32_t* my_hist = &S.warp_hist[warp][0];
#pragma unroll
  for (int i0 = 0; i0 < ITEMS; i0 += BATCH) {
    uint32_t peers[BATCH], before[BATCH];
#pragma unroll
    for (int k = 0; k < BATCH; ++k) {
      const bool ok = (uint32_t)((i0 + k) * 32) + lane < lim;
      const uint32_t d = (uint32_t)(word[i0 + k] >> shift) & mask;
      peers[k] = digit_peers<BALLOTS>(d, ok, lane);
    }
#pragma unroll
    for (int k = 0; k < BATCH; ++k) {
      const bool ok = (uint32_t)((i0 + k) * 32) + lane < lim;
      const uint32_t d = (uint32_t)(word[i0 + k] >> shift) & mask;
      before[k] = 0u;
      if (ok && (peers[k] & lt_mask) == 0u) before[k] = atomicAdd(&my_hist[d], (uint32_t)__popc(peers[k]));
    }
#pragma unroll
    for (int k = 0; k < BATCH; ++k) {
      const uint32_t rk = __shfl_sync(0xffffffffu, before[k], __ffs((int)peers[k]) - 1) + __popc(peers[k] & lt_mask);
      const int i = i0 + k;
      if (i & 1) rank2[i / 2] |= rk << 16; else rank2[i / 2] = rk;
    }
  }
  __syncthreads();
  {
    // digits d0 .. d0 + 3 of this thread: per-warp counts -> exclusive warp offsets, block counts
    const uint32_t d0 = tid * MSD_DPT;
    uint32_t block_count[MSD_DPT] = {0u, 0u, 0u, 0u};
    {
      // the thread's four counters of one warp are one 16-byte shared access
      uint4 c[MSD_WARPS];
#pragma unroll
      for (int w = 0; w < MSD_WARPS; ++w) c[w] = *reinterpret_cast<const uint4*>(&S.warp_hist[w][d0]);
#pragma unroll
      for (int w = 0; w < MSD_WARPS; ++w) {
        *reinterpret_cast<uint4*>(&S.warp_hist[w][d0]) = make_uint4(block_count[0], block_count[1], block_count[2], block_count[3]);
        block_count[0] += c[w].x; block_count[1] += c[w].y; block_count[2] += c[w].z; block_count[3] += c[w].w;
      }
    }
    const uint32_t mine = block_count[0] + block_count[1] + block_count[2] + block_count[3];
    uint32_t* st = status_raw + (size_t)tile * MSD_BINS + d0;
    static_assert(MSD_DPT == 4, "a thread's digits are one 16-byte status vector");
    if (tile != 0)
      st_relaxed_v4(st, make_uint4(MSD_FLAG_AGG | block_count[0], MSD_FLAG_AGG | block_count[1],
                                   MSD_FLAG_AGG | block_count[2], MSD_FLAG_AGG | block_count[3]));
    uint32_t incl = mine;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, s);
      if (lane >= s) incl += t;
    }
    if (lane == 31) S.warp_sums[warp] = incl;
    __syncthreads();
    uint32_t start = incl - mine;   // first slot of digit d0 in the CTA's exchange buffer
    for (unsigned w = 0; w < warp; ++w) start += S.warp_sums[w];
    uint32_t local_start[MSD_DPT];
#pragma unroll
    for (int k = 0; k < MSD_DPT; ++k) {
      local_start[k] = start;
      S.digit_local[d0 + k] = (uint16_t)start;
      start += block_count[k];
    }
    uint32_t excl[MSD_DPT];
#pragma unroll
    for (int k = 0; k < MSD_DPT; ++k) excl[k] = 0;
    if (tile != 0) {
      // A predecessor publishes its four words together (first all aggregates, later all prefixes),
      // so the four chains move in step: one 16-byte read per predecessor tile. A word that is still
      // empty means that tile has not ranked yet: read it again.
      const uint32_t* prev = st - MSD_BINS;
      unsigned open = (1u << MSD_DPT) - 1u;
      while (open) {
        const uint4 v = ld_relaxed_v4(prev);
        const uint32_t s[MSD_DPT] = {v.x, v.y, v.z, v.w};
        bool all_seen = true;
#pragma unroll
        for (int k = 0; k < MSD_DPT; ++k) all_seen = all_seen && (s[k] & (MSD_FLAG_AGG | MSD_FLAG_PREFIX)) != 0u;
        if (!all_seen) continue;
#pragma unroll
        for (int k = 0; k < MSD_DPT; ++k) {
          if ((open >> k) & 1u) {
            excl[k] += s[k] & MSD_VALUE_MASK;
            if (s[k] & MSD_FLAG_PREFIX) open &= ~(1u << k);
          }
        }
        prev -= MSD_BINS;
      }
    }
    st_relaxed_v4(st, make_uint4(MSD_FLAG_PREFIX | (excl[0] + block_count[0]), MSD_FLAG_PREFIX | (excl[1] + block_count[1]),
                                 MSD_FLAG_PREFIX | (excl[2] + block_count[2]), MSD_FLAG_PREFIX | (excl[3] + block_count[3])));
    uint32_t global_start = hstart;
#pragma unroll
    for (int k = 0; k < MSD_DPT; ++k) {
      S.digit_global[d0 + k] = global_start + excl[k] - local_start[k];
      global_start += hcount[k];
    }
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    if ((uint32_t)(i * 32) + lane < lim) {
      const uint32_t d = (uint32_t)(word[i] >> shift) & mask;
      const uint32_t rk = (i & 1) ? rank2[i / 2] >> 16 : rank2[i / 2] & 0xffffu;
      S.exch[S.digit_local[d] + my_hist[d] + rk] = word[i];
    }
  }
  __syncthreads();
  const uint32_t valid = (uint32_t)min((size_t)TILE, (size_t)n - block_base);
#pragma unroll
  for (int m = 0; m < ITEMS; ++m) {
    const uint32_t k = m * MSD_THREADS + tid;
    if (k < valid) {
      const WORD w = S.exch[k];
      const uint32_t dst = S.digit_global[(uint32_t)(w >> shift) & mask] + k;
      if (EMIT_IDX)
        out_idx[dst] = (uint32_t)w & idx_mask;
      else
        out[dst] = w;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// one CTA sorts one bucket on the remaining key bits, emits index list, keys and tile ranges
// ---------------------------------------------------------------------------------------------
struct BucketSmem {
  uint64_t buf[2][BK_CAP];                   // 2 x 96 KB
  uint16_t warp_hist[BK_WARPS][BK_RADIX];    // 32 KB: per-warp digit counts, then exclusive offsets
  uint32_t digit_start[BK_RADIX];
  uint32_t warp_sums[BK_RADIX / 32];
};
static_assert(sizeof(BucketSmem) <= 227 * 1024, "bucket sorter exceeds the 227 KB a CTA can have");
static_assert(sizeof(uint32_t) * BK_WARPS * BK_RADIX <= sizeof(uint64_t) * BK_CAP, "overflow counters alias buf[0]");

// Stable counting pass for a bucket held in shared memory (the common case). Each warp owns a
// contiguous chunk of at most BK_ROUNDS * 32 items of A and ranks them once: an item's rank among
// the warp's earlier items with the same digit stays in a register while the per-warp counts are
// scanned, then the item is placed at digit start + warp offset + rank.
constexpr int BK_ROUNDS = ((BK_CAP + BK_WARPS - 1) / BK_WARPS + 31) / 32;   // 12

__device__ __forceinline__ void bucket_pass_smem(BucketSmem& S, int cur /* source buffer 0 | 1 */, uint32_t n,
                                                 int shift, uint32_t mask) {
  // indexed through S so that the compiler knows both buffers are shared memory (LDS / STS instead
  // of generic loads and stores through a pointer that may also name an overflow bucket)
  const uint64_t* A = S.buf[cur];
  uint64_t* B = S.buf[cur ^ 1];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned lt_mask = (1u << lane) - 1u;
  uint16_t* hist = &S.warp_hist[0][0];
  {
    uint32_t* z = reinterpret_cast<uint32_t*>(hist);
    for (int i = tid; i < BK_WARPS * BK_RADIX / 2; i += BK_THREADS) z[i] = 0;
  }
  __syncthreads();
  const uint32_t chunk = ((n + BK_WARPS - 1) / BK_WARPS + 31u) & ~31u;
  const uint32_t begin = warp * chunk;
  const uint32_t end = min(n, begin + chunk);
  uint16_t rank[BK_ROUNDS];
#pragma unroll
  for (int r = 0; r < BK_ROUNDS; ++r) {
    const uint32_t row = begin + r * 32;
    if (row < end) {   // warp-uniform
      const uint32_t i = row + lane;
      const bool ok = i < end;
      const uint32_t d = ok ? (uint32_t)(A[i] >> shift) & mask : 0u;
      // the sorter is issue-bound: match.any measured 12 % faster than one ballot per digit bit
      const uint32_t peers = __match_any_sync(0xffffffffu, ok ? d : (0x10000u + lane));
      const uint32_t before = ok ? hist[warp * BK_RADIX + d] : 0u;
      __syncwarp();
      if (ok && (peers & lt_mask) == 0u) hist[warp * BK_RADIX + d] = (uint16_t)(before + __popc(peers));
      __syncwarp();
      rank[r] = (uint16_t)(before + __popc(peers & lt_mask));
    }
  }
  __syncthreads();
  if (tid < BK_RADIX) {
    uint32_t run = 0;
#pragma unroll 8
    for (int w = 0; w < BK_WARPS; ++w) {
      const uint32_t c = hist[w * BK_RADIX + tid];
      hist[w * BK_RADIX + tid] = (uint16_t)run;
      run += c;
    }
    uint32_t incl = run;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, s);
      if (lane >= s) incl += t;
    }
    if (lane == 31) S.warp_sums[warp] = incl;
    S.digit_start[tid] = incl - run;
  }
  __syncthreads();
  if (tid < BK_RADIX) {
    uint32_t woff = 0;
    for (unsigned w = 0; w < warp; ++w) woff += S.warp_sums[w];
    S.digit_start[tid] += woff;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < BK_ROUNDS; ++r) {
    const uint32_t i = begin + r * 32 + lane;
    if (i < end) {
      const uint64_t w = A[i];
      const uint32_t d = (uint32_t)(w >> shift) & mask;
      B[S.digit_start[d] + hist[warp * BK_RADIX + d] + rank[r]] = w;
    }
  }
  __syncthreads();
}

// General form of the same pass, for overflow buckets that stay in global memory: each warp owns a
// contiguous chunk of A; sweep 1 counts per (warp, digit), sweep 2 re-walks the chunk and places every item
// at digit_start + warp offset + running count, which keeps the pass stable.
// `hist` is [BK_WARPS][BK_RADIX]: 16-bit counters in S.warp_hist for buckets held in shared memory
// (a warp's chunk is at most 384 items), 32-bit counters aliased on the unused S.buf[0] for
// overflow buckets that live in global memory.
template <typename CNT>
__device__ __forceinline__ void bucket_pass(BucketSmem& S, CNT* hist, const uint64_t* A, uint64_t* B,
                                            uint32_t n, int shift, uint32_t mask) {
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned lt_mask = (1u << lane) - 1u;
  for (int i = tid; i < BK_WARPS * BK_RADIX; i += BK_THREADS) hist[i] = 0;
  __syncthreads();
  const uint32_t chunk = ((n + BK_WARPS - 1) / BK_WARPS + 31u) & ~31u;   // per warp, multiple of 32
  const uint32_t begin = warp * chunk;
  const uint32_t end = min(n, begin + chunk);
  for (uint32_t i = begin + lane; i < begin + chunk && begin < end; i += 32) {
    const bool ok = i < end;
    const uint32_t d = ok ? (uint32_t)(A[i] >> shift) & mask : 0u;
    const uint32_t peers = match_digit<BK_BITS>(d, ok);
    if (ok && (peers & lt_mask) == 0u) hist[warp * BK_RADIX + d] += (CNT)__popc(peers);   // only this warp writes its row
    __syncwarp();
    if (i - lane + 32 >= end) break;   // warp-uniform: last round of this chunk
  }
  __syncthreads();
  if (tid < BK_RADIX) {
    uint32_t run = 0;
#pragma unroll 8
    for (int w = 0; w < BK_WARPS; ++w) {
      const uint32_t c = hist[w * BK_RADIX + tid];
      hist[w * BK_RADIX + tid] = (CNT)run;
      run += c;
    }
    uint32_t incl = run;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, s);
      if (lane >= s) incl += t;
    }
    if (lane == 31) S.warp_sums[warp] = incl;
    S.digit_start[tid] = incl - run;   // exclusive within the warp of digits; warp offset added below
  }
  __syncthreads();
  if (tid < BK_RADIX) {
    uint32_t woff = 0;
    for (unsigned w = 0; w < warp; ++w) woff += S.warp_sums[w];
    S.digit_start[tid] += woff;
  }
  __syncthreads();
  for (uint32_t i = begin + lane; i < begin + chunk && begin < end; i += 32) {
    const bool ok = i < end;
    const uint64_t w = ok ? A[i] : 0ull;
    const uint32_t d = ok ? (uint32_t)(w >> shift) & mask : 0u;
    const uint32_t peers = match_digit<BK_BITS>(d, ok);
    if (ok) {
      const uint32_t before = hist[warp * BK_RADIX + d];
      B[S.digit_start[d] + before + __popc(peers & lt_mask)] = w;
    }
    __syncwarp();
    if (ok && (peers & lt_mask) == 0u) hist[warp * BK_RADIX + d] += (CNT)__popc(peers);
    __syncwarp();
    if (i - lane + 32 >= end) break;
  }
  __syncthreads();
}

__device__ __forceinline__ void sort_one_bucket(BucketSmem& S, const uint32_t bucket, uint64_t* __restrict__ words,
                                                uint64_t* __restrict__ scratch, const uint32_t* __restrict__ bucket_start,
                                                uint32_t* __restrict__ point_list, uint2* __restrict__ ranges, int n_tiles,
                                                int idx_bits, int bucket_shift, int rem_bits) {
  const uint32_t seg = bucket_start[bucket];
  const uint32_t n = bucket_start[bucket + 1] - seg;
  const unsigned tid = threadIdx.x;

  // this bucket's tiles: (0, 0) unless found below (the reference zero-fills, rasterizer_impl.cu:311)
  const uint32_t tile_lo = bucket << bucket_shift;
  const uint32_t tile_hi = min((uint32_t)n_tiles, (bucket + 1u) << bucket_shift);
  for (uint32_t t = tile_lo + tid; t < tile_hi; t += BK_THREADS) ranges[t] = make_uint2(0u, 0u);
  if (n == 0) return;

  const bool in_smem = n <= (uint32_t)BK_CAP;
  uint64_t* A;
  uint64_t* B;
  if (in_smem) {
    A = S.buf[0];
    B = S.buf[1];
    for (uint32_t i = tid; i < n; i += BK_THREADS) A[i] = words[seg + i];
  } else {
    A = words + seg;
    B = scratch + seg;
  }
  __syncthreads();

  const int passes = (rem_bits + BK_BITS - 1) / BK_BITS;
  for (int p = 0; p < passes; ++p) {
    const int shift = idx_bits + p * BK_BITS;
    const int width = min(BK_BITS, rem_bits - p * BK_BITS);
    if (in_smem)
      bucket_pass_smem(S, p & 1, n, shift, (1u << width) - 1u);
    else
      bucket_pass<uint32_t>(S, reinterpret_cast<uint32_t*>(&S.buf[0][0]), A, B, n, shift, (1u << width) - 1u);
    uint64_t* t = A;
    A = B;
    B = t;
  }

  // A holds the bucket sorted by (tile, depth bits, index)
  const uint64_t idx_mask = (1ull << idx_bits) - 1ull;
  const int tile_shift = idx_bits + 32;
  for (uint32_t i = tid; i < n; i += BK_THREADS) {
    const uint64_t w = A[i];
    const uint32_t tile = (uint32_t)(w >> tile_shift);
    point_list[seg + i] = (uint32_t)(w & idx_mask);
    if (i == 0) {
      ranges[tile].x = seg;
    } else {
      const uint32_t prev = (uint32_t)(A[i - 1] >> tile_shift);
      if (prev != tile) {
        ranges[prev].y = seg + i;
        ranges[tile].x = seg + i;
      }
    }
    if (i == n - 1) ranges[tile].y = seg + n;
  }
}

// Full words (tile | depth bits | index): CTA b sorts bucket b on every key bit below the bucket digit.
__global__ void __launch_bounds__(BK_THREADS)
bucket_sort_kernel(uint64_t* __restrict__ words /* partitioned */,
                   uint64_t* __restrict__ scratch /* other ping-pong buffer (overflow buckets) */,
                   const uint32_t* __restrict__ bucket_start, uint32_t* __restrict__ point_list,
                   uint2* __restrict__ ranges, int n_tiles, int idx_bits, int bucket_shift,
                   int rem_bits /* key bits below the bucket digit */) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BucketSmem& S = *reinterpret_cast<BucketSmem*>(smem_raw);
  // (a frame whose instances did not fit, or whose presort gave up, has an all-zero bucket histogram:
  // every bucket is empty and the tile ranges come out (0, 0), so the blend that was enqueued behind
  // this kernel draws the background and forward re-runs the stage)
  sort_one_bucket(S, blockIdx.x, words, scratch, bucket_start, point_list, ranges, n_tiles, idx_bits, bucket_shift,
                  rem_bits);
}

// ---------------------------------------------------------------------------------------------
// presorted buckets: a stable split on the tile bits, words in registers
// ---------------------------------------------------------------------------------------------
// After the depth presort a bucket only has to be split by the BITS tile bits below the bucket digit.
// That needs no staging in shared memory: a thread keeps its (at most 24) words in registers, a warp
// ranks them (one ballot per bin for up to 8 bins, match.any with per-warp counters beyond that), the
// per-warp counts are scanned in a few hundred bytes of shared memory, and index list and tile ranges
// go straight to global memory. Two CTAs fit an SM, where bucket_sort_kernel's two 96 KB buffers
// allow one (0.090 -> 0.061 ms at C3).
// A bucket longer than one chunk (SplitCfg) is taken in chunks: a counting sweep first gives the bin sizes,
// then every chunk is ranked as above on top of the running bin offsets. So there is no limit on the
// bucket size (R = 25M instances at 6M Gaussians make every bucket exceed one chunk).
constexpr int BS_THREADS = 512;
constexpr int BS_WARPS = BS_THREADS / 32;
// words a thread holds per chunk: 24 32-bit or 16 64-bit ones (64 registers per thread at two CTAs per SM)
template <typename WORD>
struct SplitCfg {
  static constexpr int ROUNDS = sizeof(WORD) == 8 ? 16 : 24;
  static constexpr int CHUNK = BS_THREADS * ROUNDS;   // 12288 or 8192
};
constexpr int BS_MAX_BITS = 7;                      // up to 2^17 tiles (8192 x 4096 pixels)

template <typename WORD, int BITS>
__global__ void __launch_bounds__(BS_THREADS, 2)
bucket_split_kernel(const WORD* __restrict__ words /* partitioned */, const uint32_t* __restrict__ bucket_start,
                    uint32_t* __restrict__ point_list, uint2* __restrict__ ranges, int n_tiles, int idx_bits) {
  constexpr int BINS = 1 << BITS;
  constexpr bool BALLOT = BITS <= 3;
  constexpr int BS_ROUNDS = SplitCfg<WORD>::ROUNDS;
  constexpr int BS_CHUNK = SplitCfg<WORD>::CHUNK;
  __shared__ uint32_t s_warp[BS_WARPS][BINS];   // per-warp counts, then exclusive offsets inside the bin
  __shared__ uint32_t s_bin[BINS + 1];          // first slot of every bin in the bucket
  __shared__ uint32_t s_tot[BINS];              // size of every bin
  __shared__ uint32_t s_run[BINS];              // words of earlier chunks in every bin
  // (skipped frames arrive here with an all-zero bucket histogram: see bucket_sort_kernel)
  const uint32_t bucket = blockIdx.x;
  const uint32_t seg = bucket_start[bucket];
  const uint32_t n = bucket_start[bucket + 1] - seg;
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned lt_mask = (1u << lane) - 1u;
  const uint32_t idx_mask = (uint32_t)((1ull << idx_bits) - 1ull);
  for (int i = tid; i < BINS; i += BS_THREADS) {
    s_tot[i] = 0u;
    s_run[i] = 0u;
  }
  __syncthreads();
  const uint32_t chunks = (n + BS_CHUNK - 1) / BS_CHUNK;
  if (chunks > 1) {   // counting sweep (CTA-uniform)
    uint32_t c = 0;
    for (uint32_t base = warp * 32; base < n; base += BS_THREADS) {   // warp-uniform trip count
      const uint32_t i = base + lane;
      const bool ok = i < n;
      const uint32_t d = ok ? (uint32_t)(words[seg + i] >> idx_bits) & (BINS - 1) : 0u;
      if (BALLOT) {
#pragma unroll
        for (int b = 0; b < BINS; ++b) {
          const uint32_t m = __ballot_sync(0xffffffffu, ok && d == (uint32_t)b);
          if (lane == (unsigned)b) c += __popc(m);
        }
      } else if (ok) {
        atomicAdd(&s_tot[d], 1u);
      }
    }
    if (BALLOT && lane < BINS && c) atomicAdd(&s_tot[lane], c);
    __syncthreads();
  }
  if (chunks != 1) {   // bin starts known up front (no words at all: all zero)
    if (tid == 0) {
      uint32_t run = 0;
      for (int b = 0; b < BINS; ++b) {
        s_bin[b] = run;
        run += s_tot[b];
      }
      s_bin[BINS] = run;
    }
    __syncthreads();
  }

  for (uint32_t ch = 0; ch < chunks; ++ch) {
    const uint32_t c0 = ch * BS_CHUNK;
    const uint32_t nc = min((uint32_t)BS_CHUNK, n - c0);
    const uint32_t per_warp = ((nc + BS_WARPS - 1) / BS_WARPS + 31u) & ~31u;   // multiple of 32, <= 32 * ROUNDS
    const uint32_t begin = warp * per_warp;
    const uint32_t end = min(nc, begin + per_warp);
    const WORD* src = words + seg + c0;

    WORD w[BS_ROUNDS];
#pragma unroll
    for (int r = 0; r < BS_ROUNDS; ++r) {
      const uint32_t i = begin + r * 32 + lane;
      w[r] = i < end ? src[i] : ~(WORD)0;
    }
    uint32_t rank2[BS_ROUNDS / 2];   // two 16-bit ranks per register
    uint32_t cnt[BALLOT ? BINS : 1];
    if (BALLOT) {
#pragma unroll
      for (int b = 0; b < BINS; ++b) cnt[b] = 0;
    } else {
      for (int b = lane; b < BINS; b += 32) s_warp[warp][b] = 0u;
      __syncwarp();
    }
#pragma unroll
    for (int r = 0; r < BS_ROUNDS; ++r) {
      uint32_t rk = 0;
      if (begin + r * 32 < end) {   // warp-uniform
        const bool ok = begin + r * 32 + lane < end;
        const uint32_t d = (uint32_t)(w[r] >> idx_bits) & (BINS - 1);
        if (BALLOT) {
#pragma unroll
          for (int b = 0; b < BINS; ++b) {
            const uint32_t m = __ballot_sync(0xffffffffu, ok && d == (uint32_t)b);
            if (d == (uint32_t)b) rk = cnt[b] + __popc(m & lt_mask);
            cnt[b] += __popc(m);
          }
        } else {
          const uint32_t peers = __match_any_sync(0xffffffffu, ok ? d : (0x10000u + lane));
          const uint32_t before = ok ? s_warp[warp][d] : 0u;
          __syncwarp();
          if (ok && (peers & lt_mask) == 0u) s_warp[warp][d] = before + __popc(peers);
          __syncwarp();
          rk = before + __popc(peers & lt_mask);
        }
      }
      if (r & 1) rank2[r / 2] |= rk << 16; else rank2[r / 2] = rk;
    }
    if (BALLOT) {
#pragma unroll
      for (int b = 0; b < BINS; ++b)
        if (lane == (unsigned)b) s_warp[warp][b] = cnt[b];
    }
    __syncthreads();   // also: the previous chunk's scatter is over before s_run moves
    uint32_t chunk_total = 0;
    if (tid < BINS) {
      uint32_t run = 0;
#pragma unroll
      for (int q = 0; q < BS_WARPS; ++q) {
        const uint32_t c = s_warp[q][tid];
        s_warp[q][tid] = run;
        run += c;
      }
      chunk_total = run;
      if (chunks == 1) s_tot[tid] = run;
    }
    __syncthreads();
    if (chunks == 1) {
      if (tid == 0) {
        uint32_t run = 0;
        for (int b = 0; b < BINS; ++b) {
          s_bin[b] = run;
          run += s_tot[b];
        }
        s_bin[BINS] = run;
      }
      __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < BS_ROUNDS; ++r) {
      const uint32_t i = begin + r * 32 + lane;
      if (i < end) {
        const uint32_t d = (uint32_t)(w[r] >> idx_bits) & (BINS - 1);
        const uint32_t rk = (r & 1) ? rank2[r / 2] >> 16 : rank2[r / 2] & 0xffffu;
        point_list[seg + s_bin[d] + s_run[d] + s_warp[warp][d] + rk] = (uint32_t)w[r] & idx_mask;
      }
    }
    __syncthreads();
    if (tid < BINS) s_run[tid] += chunk_total;
  }
  // bin b is tile (bucket << BITS) | b; empty tiles read (0, 0) like the reference's zero fill
  if (tid < BINS) {
    const uint32_t tile = (bucket << BITS) | tid;
    if (tile < (uint32_t)n_tiles) {
      const uint32_t lo = s_bin[tid], hi = s_bin[tid + 1];
      ranges[tile] = hi > lo ? make_uint2(seg + lo, seg + hi) : make_uint2(0u, 0u);
    }
  }
}

// keys[i] = tile << 32 | depth bits of the Gaussian at position i of the sorted list: what the
// reference's sorted key array holds (rasterizer_impl.cu:301-309). Nothing on the path reads it.
__global__ void __launch_bounds__(256)
binning_keys_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                    const float* __restrict__ depths, uint64_t* __restrict__ keys, int n_tiles) {
  for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const uint2 r = ranges[t];
    for (uint32_t i = r.x + threadIdx.x; i < r.y; i += 256)
      keys[i] = ((uint64_t)t << 32) | __float_as_uint(depths[point_list[i]]);
  }
}

// ---------------------------------------------------------------------------------------------
// front-to-back order of the Gaussians
// ---------------------------------------------------------------------------------------------
// The per-tile lists are ordered by (tile, depth bits, index). Sorting the ~1M Gaussians by depth once
// and emitting the ~10M instances in that order leaves the instance sorter only the tile bits to do
// (its passes are stable): one pass per bucket instead of four. A visible Gaussian lies beyond the
// near plane (depth > 0.2, preprocess), so depth bits - bits(0.2f) fits 30 bits for every depth below
// 6.8e37: three stable 10-bit passes of the partition kernel above, the last one writing the bare
// indices. Culled Gaussians get the largest key and end up behind the visible ones (they emit nothing).
// A frame with a depth beyond that raises the veto flag instead, and the bucket sorter then orders all
// the bits itself as if there had been no presort: the order is then merely some permutation, which is
// all the scan and the duplication need.
constexpr uint32_t DEPTH_KEY_BASE = 0x3E4CCCCDu;   // bits of 0.2f
constexpr uint32_t DEPTH_KEY_MAX = (1u << 30) - 1u;

__global__ void __launch_bounds__(256)
depth_words_kernel(int P, const float* __restrict__ depths, const int* __restrict__ radii, int idx_bits,
                   uint64_t* __restrict__ words, uint32_t* __restrict__ hist /* [3][MSD_BINS] */,
                   uint32_t* __restrict__ meta, uint32_t* __restrict__ veto_flag,
                   bool scramble /* test: keys unrelated to depth + veto */) {
  __shared__ uint32_t s_hist[3 * MSD_BINS];
  for (int i = threadIdx.x; i < 3 * MSD_BINS; i += 256) s_hist[i] = 0;
  __syncthreads();
  bool veto = false;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < P; i += gridDim.x * 256) {
    uint32_t key = DEPTH_KEY_MAX;
    if (radii[i] > 0) {
      const uint32_t b = __float_as_uint(depths[i]);
      key = b - DEPTH_KEY_BASE;
      if (scramble) key *= 2654435761u;   // a bijection of the 30-bit keys: equal depths still tie, as in a real veto
      if (scramble || b < DEPTH_KEY_BASE || key >= DEPTH_KEY_MAX) {
        veto = true;
        key &= DEPTH_KEY_MAX;
      }
    }
    words[i] = ((uint64_t)key << idx_bits) | (uint32_t)i;
    atomicAdd(&s_hist[key & (MSD_BINS - 1)], 1u);
    atomicAdd(&s_hist[MSD_BINS + ((key >> 10) & (MSD_BINS - 1))], 1u);
    atomicAdd(&s_hist[2 * MSD_BINS + (key >> 20)], 1u);
  }
  if (veto) *veto_flag = 1u;
  if (blockIdx.x == 0 && threadIdx.x == 0) meta[2] = (uint32_t)P;
  __syncthreads();
  for (int i = threadIdx.x; i < 3 * MSD_BINS; i += 256) {
    const uint32_t c = s_hist[i];
    if (c) atomicAdd(&hist[i], c);
  }
}

const uint32_t* presort_veto_flag(const GeomPtrs& g) { return g.readback + 1; }

int launch_depth_order(const GeomPtrs& g, int P, const int* radii, cudaStream_t stream, bool debug) {
  StageTimer timer(stream, ST_DEPTH_ORDER);
  uint32_t* aux = g.sort_aux;
  const int ib = tile_id_bits((uint32_t)(P > 1 ? P - 1 : 1));
  // the aux words were cleared with the rest of the zeroed run when the forward began
  const int blocks = P < 148 * 4 * 256 ? (P + 255) / 256 : 148 * 4;
  // test hook (egs_debug_option "presort_scramble"): an order unrelated to depth + the veto raised
  const bool scramble = debug_options().presort_scramble != 0;
  depth_words_kernel<<<blocks, 256, 0, stream>>>(P, g.depths, radii, ib, g.sort_keys[0], aux + DO_HIST, aux + DO_META,
                                                 g.readback + 1, scramble);
  EGS_AFTER_LAUNCH(stream, debug);
  constexpr int SMEM = (int)sizeof(MsdSmem<uint64_t>);
  EGS_CUDA_TRY(cudaFuncSetAttribute(msd_partition_kernel<uint64_t, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
  EGS_CUDA_TRY(cudaFuncSetAttribute(msd_partition_kernel<uint64_t, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
  const InstanceCount count{aux + DO_META + 2, (uint32_t)P, nullptr};
  const int grid = (int)g.sort_blocks;
  const size_t status_stride = g.sort_blocks * MSD_BINS;
  msd_partition_kernel<uint64_t, false, true><<<grid, MSD_THREADS, SMEM, stream>>>(
      g.sort_keys[0], g.sort_keys[1], count, ib, MSD_BINS - 1, aux + DO_HIST, nullptr, aux + DO_TICKET, aux + DO_STATUS,
      nullptr, 0u);
  EGS_AFTER_LAUNCH(stream, debug);
  msd_partition_kernel<uint64_t, false, true><<<grid, MSD_THREADS, SMEM, stream>>>(
      g.sort_keys[1], g.sort_keys[0], count, ib + 10, MSD_BINS - 1, aux + DO_HIST + MSD_BINS, nullptr, aux + DO_TICKET + 1,
      aux + DO_STATUS + status_stride, nullptr, 0u);
  EGS_AFTER_LAUNCH(stream, debug);
  msd_partition_kernel<uint64_t, true, true><<<grid, MSD_THREADS, SMEM, stream>>>(
      g.sort_keys[0], nullptr, count, ib + 20, MSD_BINS - 1, aux + DO_HIST + 2 * MSD_BINS, nullptr, aux + DO_TICKET + 2,
      aux + DO_STATUS + 2 * status_stride, g.sort_vals[0], (uint32_t)((1ull << ib) - 1ull));
  EGS_AFTER_LAUNCH(stream, debug);
  return EGS_OK;
}

// full words: tile | depth | index must fit 64 bits and most buckets the sorter's shared memory;
// otherwise the pure LSD paths are the better choice
bool bucket_path_ok(int P, int R, int W, int H) {
  if (!packed_path_fits(P, W, H)) return false;
  return (long long)R <= (long long)MSD_BINS * BK_CAP * 8 / 10;
}

// presorted words (tile | index, 32 or 64 bits): any P; the split handles up to BS_MAX_BITS tile bits
// below the 10-bit bucket digit
bool presorted_path_ok(int P, int W, int H) {
  const int tb = tile_id_bits((uint32_t)(tiles_x(W) * tiles_y(H)));
  return P >= 1 && tb - MSD_BITS_MAX <= BS_MAX_BITS;
}

template <typename WORD, bool DEPTH>
static int launch_duplicate(const GeomPtrs& g, WORD* words, const int* radii, const uint32_t* gauss_order, int P, int gx,
                            int gy, int ib, InstanceCount count, int bucket_shift, uint32_t* hist, cudaStream_t stream,
                            bool debug) {
  // each warp strides over groups of 32 Gaussians; the bucket histogram comes out of the
  // same kernel. With the presort, sort_keys[1] is free again and holds the list of set-aside
  // Gaussians, counted in meta[4] (zeroed with the rest of the aux words when the forward began).
  uint32_t* big_list = gauss_order ? reinterpret_cast<uint32_t*>(g.sort_keys[1]) : nullptr;
  uint32_t* big_count = gauss_order ? g.sort_aux + DO_META + 4 : nullptr;
  duplicate_words_kernel<WORD, DEPTH><<<148 * DUP_CTAS_PER_SM, DUP_THREADS, 0, stream>>>(
      P, g.means2D, g.depths, g.point_offsets, radii, gauss_order, words, gx, gy, ib, count, bucket_shift, hist, big_list,
      big_count);
  EGS_AFTER_LAUNCH(stream, debug);
  if (gauss_order) {
    duplicate_big_kernel<WORD, DEPTH><<<148 * 4, DUP_THREADS, 0, stream>>>(
        g.means2D, g.depths, g.point_offsets, radii, gauss_order, words, gx, gy, ib, count, bucket_shift, hist, big_list,
        big_count);
    EGS_AFTER_LAUNCH(stream, debug);
  }
  return EGS_OK;
}

template <typename WORD>
static int launch_partition(const WORD* in, WORD* out, int capacity, InstanceCount count, int shift, const uint32_t* hist,
                            uint32_t* start, uint32_t* ticket, uint32_t* status, cudaStream_t stream, bool debug) {
  constexpr int SMEM = (int)sizeof(MsdSmem<WORD>);
  EGS_CUDA_TRY(cudaFuncSetAttribute(msd_partition_kernel<WORD, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
  const int grid = (capacity + MsdCfg<WORD>::TILE - 1) / MsdCfg<WORD>::TILE;
  msd_partition_kernel<WORD, false><<<grid, MSD_THREADS, SMEM, stream>>>(in, out, count, shift, 0xffffffffu, hist, start,
                                                                        ticket, status, nullptr, 0u);
  EGS_AFTER_LAUNCH(stream, debug);
  return EGS_OK;
}

template <typename WORD>
static int launch_split(const WORD* words, const uint32_t* start, uint32_t* point_list, uint2* ranges, int n_tiles, int ib,
                        int bits, cudaStream_t stream, bool debug) {
#define EGS_SPLIT_CASE(B)                                                                                         \
  case B:                                                                                                         \
    bucket_split_kernel<WORD, B><<<MSD_BINS, BS_THREADS, 0, stream>>>(words, start, point_list, ranges, n_tiles, ib); \
    break;
  switch (bits) {
    EGS_SPLIT_CASE(0) EGS_SPLIT_CASE(1) EGS_SPLIT_CASE(2) EGS_SPLIT_CASE(3) EGS_SPLIT_CASE(4) EGS_SPLIT_CASE(5)
    EGS_SPLIT_CASE(6) EGS_SPLIT_CASE(7)
    default:
      set_error("binning: %d tile bits below the bucket digit (image too large for the presorted path)", bits);
      return EGS_ERR_TOO_LARGE;
  }
#undef EGS_SPLIT_CASE
  EGS_AFTER_LAUNCH(stream, debug);
  return EGS_OK;
}

// histogram, ticket and the status words of the single partition pass (not the whole area the
// five-pass paths need)
int zero_binning_buckets(const BinningPtrs& b, cudaStream_t stream) {
  const size_t zero = reinterpret_cast<char*>(b.status + b.sort_blocks * MSD_BINS) - reinterpret_cast<char*>(b.hist);
  EGS_CUDA_TRY(cudaMemsetAsync(b.hist, 0, zero, stream));
  return EGS_OK;
}

// `capacity` sizes the grids and the workspace `b`; the real number of instances is read on the
// device through `count` (it may still be in flight when this is enqueued). capacity == 0 means
// the count is known to be 0. `zeroed`: the caller already ran zero_binning_buckets on this stream.
int launch_binning_buckets(const GeomPtrs& g, const BinningPtrs& b, const ImagePtrs& im, const int* radii,
                           const uint32_t* gauss_order, bool presorted, int P, int capacity, InstanceCount count, int W,
                           int H, cudaStream_t stream, bool debug, bool zeroed) {
  const int gx = tiles_x(W), gy = tiles_y(H);
  const int n_tiles = gx * gy;
  if (capacity == 0) {
    EGS_CUDA_TRY(cudaMemsetAsync(im.ranges, 0, (size_t)n_tiles * sizeof(uint2), stream));
    return EGS_OK;
  }
  const int ib = tile_id_bits((uint32_t)(P > 1 ? P - 1 : 1));
  const int tb = tile_id_bits((uint32_t)n_tiles);
  const int mb = tb < MSD_BITS_MAX ? tb : MSD_BITS_MAX;   // bucket digit = top mb bits of the tile id
  const int bucket_shift = tb - mb;
  uint32_t* hist = b.hist;                 // [MSD_BINS]
  uint32_t* start = b.hist + MSD_BINS;     // [MSD_BINS + 1]
  static_assert(2 * MSD_BINS + 1 <= SORT_HIST_WORDS, "histogram area too small");
  const bool narrow = presorted && tb + ib <= 32;   // tile | index fits a 32-bit word
  {
    StageTimer timer(stream, ST_DUPLICATE);
    if (!zeroed)
      if (int rc0 = zero_binning_buckets(b, stream)) return rc0;
    int rc;
    if (narrow)
      rc = launch_duplicate<uint32_t, false>(g, reinterpret_cast<uint32_t*>(b.keys[1]), radii, gauss_order, P, gx, gy, ib,
                                             count, bucket_shift, hist, stream, debug);
    else if (presorted)
      rc = launch_duplicate<uint64_t, false>(g, b.keys[1], radii, gauss_order, P, gx, gy, ib, count, bucket_shift, hist,
                                             stream, debug);
    else
      rc = launch_duplicate<uint64_t, true>(g, b.keys[1], radii, gauss_order, P, gx, gy, ib, count, bucket_shift, hist,
                                            stream, debug);
    if (rc) return rc;
  }
  {
    StageTimer timer(stream, ST_SORT_HIST);   // reported as the global partition pass
    const int shift = ib + (presorted ? 0 : 32) + bucket_shift;
    int rc;
    if (narrow)
      rc = launch_partition<uint32_t>(reinterpret_cast<const uint32_t*>(b.keys[1]), reinterpret_cast<uint32_t*>(b.keys[0]),
                                      capacity, count, shift, hist, start, b.ticket, b.status, stream, debug);
    else
      rc = launch_partition<uint64_t>(b.keys[1], b.keys[0], capacity, count, shift, hist, start, b.ticket, b.status, stream,
                                      debug);
    if (rc) return rc;
  }
  {
    StageTimer timer(stream, ST_SORT_PASSES);   // reported as the per-bucket stage
    if (presorted) {
      // emitted front to back, the stable partition left every bucket ordered by depth: only the tile
      // bits below the bucket digit remain
      int rc;
      if (narrow)
        rc = launch_split<uint32_t>(reinterpret_cast<const uint32_t*>(b.keys[0]), start, b.vals[0], im.ranges, n_tiles, ib,
                                    bucket_shift, stream, debug);
      else
        rc = launch_split<uint64_t>(b.keys[0], start, b.vals[0], im.ranges, n_tiles, ib, bucket_shift, stream, debug);
      if (rc) return rc;
    } else {
      EGS_CUDA_TRY(cudaFuncSetAttribute(bucket_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)sizeof(BucketSmem)));
      bucket_sort_kernel<<<MSD_BINS, BK_THREADS, sizeof(BucketSmem), stream>>>(
          b.keys[0], b.keys[1], start, b.vals[0], im.ranges, n_tiles, ib, bucket_shift, bucket_shift + 32);
      EGS_AFTER_LAUNCH(stream, debug);
    }
  }
  return EGS_OK;
}

int launch_binning_keys(const GeomPtrs& g, const ImagePtrs& im, const uint32_t* point_list, uint64_t* keys, int W, int H,
                        cudaStream_t stream) {
  const int n_tiles = tiles_x(W) * tiles_y(H);
  binning_keys_kernel<<<n_tiles < 148 * 8 ? n_tiles : 148 * 8, 256, 0, stream>>>(im.ranges, point_list, g.depths, keys,
                                                                               n_tiles);
  EGS_AFTER_LAUNCH(stream, false);
  return EGS_OK;
}

}  // namespace egs
