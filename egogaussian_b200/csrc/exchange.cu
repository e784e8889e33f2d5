// exchange.cu -- the gradient exchange of view-data-parallel training as ONE kernel per rank over
// peer-mapped memory (NVLink 5 / NVSwitch), with or without the Adam update inside it.
//
// Every rank holds the full Gaussian set and back-propagates another view; the flat gradient
// buffers of all ranks are mapped into every rank's address space. Rank r owns slice r:
//
//   barrier in   every rank has finished its backward (a flag per peer, release/acquire at system scope)
//   reduce       load slice r of all `world` buffers with 16-byte peer loads, add in rank order
//   [update]     egs_exchange_adam_step: Adam on slice r of the parameters, moments sharded by rank
//   broadcast    store the sums (or the new parameters) into slice r of every rank's buffer
//   barrier out  every rank's stores have landed before anyone's next kernel starts
//
// so each link carries (world - 1) / world of the buffer once in each direction and nothing is staged
// or copied: the same volume as a ring all-reduce, at full NVSwitch bandwidth from the first byte,
// one launch, and with the update fused the optimiser's 28 B/float pass shrinks to the rank's own slice.
// Sums are formed in rank order by exactly one rank per element, so all ranks see the same bits.
#include "common.cuh"
#include "adam_math.cuh"

#include <math.h>
#include <atomic>
#include <map>
#include <mutex>
#include <utility>

namespace egs {

constexpr int EX_THREADS = 512;
constexpr int EX_SLOT_IN = 0;                                   // [0, 8): "rank q entered exchange e"
constexpr int EX_SLOT_OUT = EGS_MAX_PEERS;                      // [8, 16): "rank q finished exchange e"
constexpr int EX_SLOT_COUNT = 2 * EGS_MAX_PEERS;                // CTAs of this rank that finished
constexpr int EX_SLOT_ERROR = EGS_EXCHANGE_SIGNAL_WORDS - 1;    // set when a peer never showed up
// how long a kernel waits for a peer before it gives up (egs_exchange_set_timeout_ms)
static std::atomic<unsigned long long> g_timeout_ns{20ull * 1000 * 1000 * 1000};

struct ExSegmentDev {
  long long begin4, end4;   // in 4-float chunks
  float neg_step_a, neg_step_b;
  int period, split;
};

struct ExLaunch {
  int world, rank;
  int n_src;                        // buffers summed: one per rank (src == grads), or one per view of a batch
  const float* src[EGS_MAX_PEERS];
  float* grads[EGS_MAX_PEERS];      // where the sums go: every rank's buffer
  float* params[EGS_MAX_PEERS];
  uint32_t* signals[EGS_MAX_PEERS];
  float* grads_mc;    // multicast mappings of the same buffers, or NULL
  float* params_mc;
  long long lo4, hi4;   // this rank's slice, in 4-float chunks
  float scale;
  uint32_t epoch;
  unsigned long long timeout_ns;
  // Adam
  float* exp_avg;
  float* exp_avg_sq;
  ExSegmentDev seg[EGS_EXCHANGE_MAX_SEGMENTS];
  int n_segments;
  float one_minus_b1, b2, one_minus_b2, eps, bc2_sqrt;
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// peer data changes between launches under the same addresses: keep it out of the L1
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ float ld_peer_f32(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_f4(float* p, float4 v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// threads [0, world) each wait for one peer's flag to reach `epoch`. Returns false (to every thread
// of the CTA) when a peer did not show up in time; the error word of this rank is then set.
__device__ __forceinline__ bool wait_peer_flags(const ExLaunch& L, int slot, int& s_timed_out) {
  if (threadIdx.x == 0) s_timed_out = 0;
  __syncthreads();
  if ((int)threadIdx.x < L.world) {
    const uint32_t* flag = L.signals[L.rank] + slot + threadIdx.x;
    const unsigned long long t0 = global_ns();
    while ((int)(ld_acquire_sys(flag) - L.epoch) < 0) {
      if (global_ns() - t0 > L.timeout_ns) {
        L.signals[L.rank][EX_SLOT_ERROR] = 1u;
        s_timed_out = 1;
        break;
      }
    }
  }
  __syncthreads();
  return s_timed_out == 0;
}

// my backward is complete (this kernel follows it in stream order); wait until every peer's is.
// false: some peer never arrived. The caller then skips the data phase -- a straggler's buffer still
// holds the previous iteration, and summing it (or, worse, stepping the optimiser on it and
// broadcasting the result) would corrupt every rank -- but still runs barrier_out, so that peers
// which did arrive are released.
__device__ __forceinline__ bool barrier_in(const ExLaunch& L, int& s_timed_out) {
  if (blockIdx.x == 0 && (int)threadIdx.x < L.world)
    st_release_sys(L.signals[threadIdx.x] + EX_SLOT_IN + L.rank, L.epoch);
  return wait_peer_flags(L, EX_SLOT_IN, s_timed_out);
}

// all my stores are visible everywhere, then all peers' stores are visible here; the last CTA of the
// grid speaks for the rank
__device__ __forceinline__ void barrier_out(const ExLaunch& L, bool& s_last, int& s_timed_out) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* count = L.signals[L.rank] + EX_SLOT_COUNT;
    const uint32_t done = atomicAdd(count, 1u);
    __threadfence_system();
    s_last = done == gridDim.x - 1;
    if (s_last) *count = 0u;   // every other CTA of this launch is past it
  }
  __syncthreads();
  if (!s_last) return;
  if ((int)threadIdx.x < L.world) st_release_sys(L.signals[threadIdx.x] + EX_SLOT_OUT + L.rank, L.epoch);
  wait_peer_flags(L, EX_SLOT_OUT, s_timed_out);
}

// multicast addresses (NVLink SHARP): one load returns the sum over all ranks, formed in the switch;
// one store lands in every rank's buffer
__device__ __forceinline__ float4 ld_reduce_mc_f4(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_mc_f4(float* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

__device__ __forceinline__ void store_to_all(float* const* bufs, int world, long long c, float4 v) {
#pragma unroll
  for (int q = 0; q < EGS_MAX_PEERS; ++q)
    if (q < world) st_peer_f4(bufs[q] + 4 * c, v);
}

// Adam on chunk c of this rank's slice; returns the new parameters
__device__ __forceinline__ float4 adam_chunk(const ExLaunch& L, const ExSegmentDev& S, long long c, const float4 g) {
  float ns0 = S.neg_step_a, ns1 = S.neg_step_a, ns2 = S.neg_step_a, ns3 = S.neg_step_a;
  if (S.period > 0) {
    const int r0 = (int)(((c - S.begin4) * 4) % S.period);
    auto pick = [&](int r) { return r % S.period < S.split ? S.neg_step_a : S.neg_step_b; };   // any period >= 1
    ns0 = pick(r0); ns1 = pick(r0 + 1); ns2 = pick(r0 + 2); ns3 = pick(r0 + 3);
  }
  const long long k4 = (c - L.lo4) * 4;
  float4 p = *reinterpret_cast<const float4*>(L.params[L.rank] + 4 * c);
  float4 m = *reinterpret_cast<const float4*>(L.exp_avg + k4);
  float4 v = *reinterpret_cast<const float4*>(L.exp_avg_sq + k4);
  adam_elem(p.x, g.x, m.x, v.x, L.one_minus_b1, L.b2, L.one_minus_b2, L.eps, ns0, L.bc2_sqrt);
  adam_elem(p.y, g.y, m.y, v.y, L.one_minus_b1, L.b2, L.one_minus_b2, L.eps, ns1, L.bc2_sqrt);
  adam_elem(p.z, g.z, m.z, v.z, L.one_minus_b1, L.b2, L.one_minus_b2, L.eps, ns2, L.bc2_sqrt);
  adam_elem(p.w, g.w, m.w, v.w, L.one_minus_b1, L.b2, L.one_minus_b2, L.eps, ns3, L.bc2_sqrt);
  *reinterpret_cast<float4*>(L.exp_avg + k4) = m;
  *reinterpret_cast<float4*>(L.exp_avg_sq + k4) = v;
  return p;
}

__device__ __forceinline__ int segment_of(const ExLaunch& L, long long c) {
  int si = -1;
#pragma unroll 1
  for (int k = 0; k < L.n_segments; ++k)
    if (c >= L.seg[k].begin4 && c < L.seg[k].end4) si = k;
  return si;
}

// Data phase over peer loads / stores. WMAX: the smallest of 2, 4, 8 that holds `world`; a thread keeps 8
// peer loads of 16 bytes in flight (8 / WMAX chunks from every rank), which is what it takes to fill an
// NVLink with 1024 threads per SM.
// `tid` / `threads`: index and number of the CTA's threads that take part (all of them, or the few warps
// the factored exchange sets aside for this phase)
template <bool ADAM, int WMAX>
__device__ __forceinline__ void sum_slice_peer(const ExLaunch& L, int tid = (int)threadIdx.x, const int EX_THREADS = egs::EX_THREADS) {
  constexpr int U = 8 / WMAX;
  const long long stride = (long long)gridDim.x * EX_THREADS * U;
  for (long long c0 = L.lo4 + (long long)blockIdx.x * EX_THREADS * U + tid; c0 < L.hi4; c0 += stride) {
    float4 part[U][WMAX];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int q = 0; q < WMAX; ++q)
        if (q < L.n_src && c0 + u * EX_THREADS < L.hi4) part[u][q] = ld_peer_f4(L.src[q] + 4 * (c0 + u * EX_THREADS));
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long c = c0 + u * EX_THREADS;
      if (c >= L.hi4) break;
      float4 g = part[u][0];   // rank (view) order, whoever adds
#pragma unroll
      for (int q = 1; q < WMAX; ++q)
        if (q < L.n_src) {
          g.x += part[u][q].x; g.y += part[u][q].y; g.z += part[u][q].z; g.w += part[u][q].w;
        }
      g.x *= L.scale; g.y *= L.scale; g.z *= L.scale; g.w *= L.scale;
      if (ADAM) {
        const int si = segment_of(L, c);
        if (si >= 0) {
          store_to_all(L.params, L.world, c, adam_chunk(L, L.seg[si], c, g));
          continue;
        }
      }
      store_to_all(L.grads, L.world, c, g);
    }
  }
}

// The same through the switch: (world - 1) / world of the buffer crosses each link once per direction
// instead of twice. Four reductions of 16 bytes in flight per thread.
template <bool ADAM, int U>
__device__ __forceinline__ void sum_slice_mc(const ExLaunch& L, int tid = (int)threadIdx.x, const int EX_THREADS = egs::EX_THREADS) {
  const long long stride = (long long)gridDim.x * EX_THREADS * U;
  for (long long c0 = L.lo4 + (long long)blockIdx.x * EX_THREADS * U + tid; c0 < L.hi4; c0 += stride) {
    float4 g[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (c0 + u * EX_THREADS < L.hi4) g[u] = ld_reduce_mc_f4(L.grads_mc + 4 * (c0 + u * EX_THREADS));
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long c = c0 + u * EX_THREADS;
      if (c >= L.hi4) break;
      g[u].x *= L.scale; g[u].y *= L.scale; g[u].z *= L.scale; g[u].w *= L.scale;
      if (ADAM) {
        const int si = segment_of(L, c);
        if (si >= 0) {
          st_mc_f4(L.params_mc + 4 * c, adam_chunk(L, L.seg[si], c, g[u]));
          continue;
        }
      }
      st_mc_f4(L.grads_mc + 4 * c, g[u]);
    }
  }
}

template <bool ADAM, int WMAX>
__global__ void __launch_bounds__(EX_THREADS, 2) exchange_kernel(const __grid_constant__ ExLaunch L) {
  __shared__ bool s_last;
  __shared__ int s_timed_out;
  if (barrier_in(L, s_timed_out)) sum_slice_peer<ADAM, WMAX>(L);
  barrier_out(L, s_last, s_timed_out);
}

template <bool ADAM>
__global__ void __launch_bounds__(EX_THREADS, 2) exchange_mc_kernel(const __grid_constant__ ExLaunch L) {
  __shared__ bool s_last;
  __shared__ int s_timed_out;
  if (barrier_in(L, s_timed_out)) sum_slice_mc<ADAM, 4>(L);
  barrier_out(L, s_last, s_timed_out);
}

// ---------------------------------------------------------------------------------------------
// factored SH gradient
// ---------------------------------------------------------------------------------------------
// dL/dsh is 12 M of the 4 (14 + 3 M) bytes a Gaussian's gradient takes (192 of 248 at SH degree 3), but
// per view it has rank one: dL/dsh[k][c] = basis_k(direction from that view's camera) * dL/dRGB[c]. So
// the backward of a view-data-parallel step writes the three floats of dL/dRGB (EgsBackwardArgs.dL_drgb)
// and stores them, from inside its per-Gaussian kernel, into that view's slot of EVERY rank's gather
// buffer (peer or multicast stores: fire and forget, overlapped with the kernel's own HBM traffic; pulling
// them here with loads instead measured 0.5 ms at 8 GPUs, because remote loads need ~20 MB in flight to
// fill the links and a load-then-compute loop has a fraction of that). Here every rank reads all views'
// 12 bytes per Gaussian from its own memory, evaluates each view's basis from the (replicated) positions
// and that view's camera centre, and adds the products in view order into its own dL/dsh tensor. Per GPU
// (views - 1) * 12 bytes per Gaussian cross the link instead of 2 * 192 * (world - 1) / world; the
// products and the order of the additions are those of summing the expanded rows view by view, so the
// result is the same bits. The other 14 floats per Gaussian go through the ordinary slice exchange
// first, inside the same pair of barriers.
struct ShFactorLaunch {
  int P, D, M;
  int n;   // views: one per rank, or all views of a batch in their global order
  int phases;   // timing hook: 0 both, 1 slice sum only, 2 expansion only
  const float* means3D;
  const float* drgb[EGS_MAX_PEERS];   // every view's slot (rows + camera centre) in THIS rank's memory
  float* dsh;
};
constexpr int SF_PITCH = 49;                         // odd: a lane walking its own row is conflict-free
constexpr int SF_WARPS = EX_THREADS / 32;
constexpr int SF_STAGE = EGS_MAX_PEERS * 96;          // floats: a warp's 32 dL/dRGB triples of every view
constexpr int SF_WARP_FLOATS = 32 * SF_PITCH + 2 * SF_STAGE;
constexpr int SF_SMEM = (SF_WARPS * SF_WARP_FLOATS + EGS_MAX_PEERS * 4) * 4;

// 16 bytes global -> shared without a register in between, cached at L2 only: the slots are written by
// other GPUs through NVLink, and L2 is where this GPU's memory is coherent
__device__ __forceinline__ void cp_async_16(float* smem_dst, const float* gmem_src) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Camera centres of all views into shared memory (every thread of the CTA calls this).
__device__ __forceinline__ void load_view_cameras(const ShFactorLaunch& F, float* smem) {
  float* s_cam = smem + SF_WARPS * SF_WARP_FLOATS;             // [views][4]
  if (threadIdx.x < (unsigned)F.n * 3) {
    // the slots were written (by this rank's and the peers' backward kernels) before their ranks' flags:
    // ordered by the acquire in barrier_in; sys-scope loads keep stale L1 lines out
    const int q = threadIdx.x / 3, c = threadIdx.x % 3;
    s_cam[4 * q + c] = ld_peer_f32(F.drgb[q] + ((size_t)F.P * 3 + 3) / 4 * 4 + c);
  }
  __syncthreads();
}

// A warp takes 32 Gaussians at a time: their triples of every view (96 floats = 24 x 16 bytes per view)
// arrive in a double-buffered stage by cp.async, one group ahead of the arithmetic; a lane accumulates
// its Gaussian's 48 sums over the views in registers, the rows go through a tile and leave as 16-byte
// stores. The CTA owns a contiguous range of groups and its warps draw from it through a shared counter,
// so the warps that first ran the slice sum join in when they are done.
__device__ __forceinline__ void expand_sh_sum(const ExLaunch& L, const ShFactorLaunch& F, float* smem, uint32_t* s_next) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* tile = smem + warp * SF_WARP_FLOATS;
  float* stage = tile + 32 * SF_PITCH;                         // [2][views][96]
  const float* s_cam = smem + SF_WARPS * SF_WARP_FLOATS;       // [views][4]
  const int groups = (F.P + 31) / 32;
  const int per_cta = (groups + (int)gridDim.x - 1) / (int)gridDim.x;
  const int g_lo = (int)blockIdx.x * per_cta, g_hi = min(groups, g_lo + per_cta);
  const int row_floats = 3 * F.M;
  auto draw = [&]() -> int {   // the next group of this CTA, the same value on every lane
    uint32_t v = 0;
    if (lane == 0) v = atomicAdd(s_next, 1u);
    return g_lo + (int)__shfl_sync(0xffffffffu, v, 0);
  };
  auto fetch = [&](int grp, int buf) {
    const size_t at = (size_t)grp * 96 + lane * 4;
    if (grp < g_hi && lane < 24 && at < (size_t)F.P * 3) {   // the last 16 bytes may run into the slot's padding
#pragma unroll
      for (int q = 0; q < EGS_MAX_PEERS; ++q)
        if (q < F.n) cp_async_16(stage + buf * SF_STAGE + q * 96 + lane * 4, F.drgb[q] + at);
    }
    cp_async_commit();
  };
  auto position = [&](int g) -> float3 {   // this lane's Gaussian of group g
    const int i = g * 32 + (int)lane;
    return (g < g_hi && i < F.P) ? make_float3(F.means3D[3 * i], F.means3D[3 * i + 1], F.means3D[3 * i + 2])
                                 : make_float3(0.f, 0.f, 0.f);
  };
  int grp = draw();
  fetch(grp, 0);
  float3 pos_next = position(grp);
  for (int it = 0; grp < g_hi; ++it) {
    const int buf = it & 1;
    const int next = draw();
    fetch(next, buf ^ 1);
    const int first = grp * 32;
    const int rows = min(32, F.P - first);
    const float3 pos = pos_next;
    pos_next = position(next);   // a group ahead, like the triples: the load is not waited for here
    cp_async_wait<1>();   // this group's triples have landed (the next group's may still be in flight)
    __syncwarp();
    const float* st = stage + buf * SF_STAGE;
    float acc[48];
#pragma unroll
    for (int k = 0; k < 48; ++k) acc[k] = 0.f;
#pragma unroll
    for (int q = 0; q < EGS_MAX_PEERS; ++q)
      if (q < F.n) {
        const float dr = st[q * 96 + 3 * lane], dg = st[q * 96 + 3 * lane + 1], db = st[q * 96 + 3 * lane + 2];
        // same expressions as sh_backward (preprocess_bwd.cu)
        const float3 dir = make_float3(pos.x - s_cam[4 * q], pos.y - s_cam[4 * q + 1], pos.z - s_cam[4 * q + 2]);
        const float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
        const float x = dir.x / len, y = dir.y / len, z = dir.z / len;
        float b[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) b[k] = 0.f;
        sh_basis_values(F.D, x, y, z, b);
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          // product and sum rounded separately: the sum of the views' expanded rows, not an FMA chain
          const float pr = __fmul_rn(b[k], dr), pg = __fmul_rn(b[k], dg), pb = __fmul_rn(b[k], db);
          if (q == 0) {
            acc[3 * k] = pr; acc[3 * k + 1] = pg; acc[3 * k + 2] = pb;
          } else {
            acc[3 * k] = __fadd_rn(acc[3 * k], pr);
            acc[3 * k + 1] = __fadd_rn(acc[3 * k + 1], pg);
            acc[3 * k + 2] = __fadd_rn(acc[3 * k + 2], pb);
          }
        }
      }
#pragma unroll
    for (int k = 0; k < 48; ++k)
      if (k < row_floats) tile[lane * SF_PITCH + k] = L.scale == 1.f ? acc[k] : acc[k] * L.scale;
    __syncwarp();
    float* dst = F.dsh + (size_t)first * row_floats;
    if (F.M == 16) {
      for (int f = lane; f < rows * 12; f += 32) {
        const int row = f / 12, k = (f % 12) * 4;
        const float* srow = tile + row * SF_PITCH + k;
        reinterpret_cast<float4*>(dst)[f] = make_float4(srow[0], srow[1], srow[2], srow[3]);
      }
    } else {
      for (int e = lane; e < rows * row_floats; e += 32) dst[e] = tile[(e / row_floats) * SF_PITCH + e % row_floats];
    }
    __syncwarp();   // the tile and this stage buffer are free for the group after next
    grp = next;
  }
  cp_async_wait<0>();
}

// Warps of a CTA that run the slice sum of the small gradients (NVLink-bound: a few warps with many
// (A push form of this phase -- every rank stores its contributions into the owners' staging areas, the
// owners add their slices from local memory -- was built, bit-identical, and measured slower: 0.133
// against 0.117 ms at 2 ranks, 0.195 against 0.162 at 4. The phase is bound by the bytes per link, about
// 560 GB/s per direction either way, not by loads in flight, and the push form adds a barrier.)
// requests in flight each) while the others already expand the SH sums (local, arithmetic and latency);
// they join the expansion when their part of the slice is done.
// Through the switch the sum is bound by the switch's reduction rate whatever the number of warps (0.2 ms
// for 56 MB at 8 ranks with 4 or with 16 warps); peer loads want more requests in flight.
constexpr int SF_SUM_WARPS_MC = 4;
constexpr int SF_SUM_WARPS_PEER = 8;

template <bool MC, int WMAX>
__global__ void __launch_bounds__(EX_THREADS, 1)
exchange_factored_kernel(const __grid_constant__ ExLaunch L, const __grid_constant__ ShFactorLaunch F) {
  extern __shared__ __align__(16) float sf_smem[];
  __shared__ bool s_last;
  __shared__ int s_timed_out;
  __shared__ uint32_t s_next;
  if (threadIdx.x == 0) s_next = 0u;
  if (barrier_in(L, s_timed_out)) {
    load_view_cameras(F, sf_smem);
    constexpr int SUM_THREADS = (MC ? SF_SUM_WARPS_MC : SF_SUM_WARPS_PEER) * 32;
    if (F.phases != 2 && threadIdx.x < SUM_THREADS) {
      if (MC) sum_slice_mc<false, 16>(L, (int)threadIdx.x, SUM_THREADS);
      else sum_slice_peer<false, WMAX>(L, (int)threadIdx.x, SUM_THREADS);
    }
    if (F.phases != 1) expand_sh_sum(L, F, sf_smem, &s_next);
  }
  barrier_out(L, s_last, s_timed_out);
}

static int fill_group(const char* who, const EgsPeerGroup* g, long long n, ExLaunch& L) {
  if (!g || g->world < 1 || g->world > EGS_MAX_PEERS || g->rank < 0 || g->rank >= g->world || n < 0 || (n & 3)) {
    set_error("%s: world must be in [1, %d], rank in [0, world), n_floats a non-negative multiple of 4", who,
              EGS_MAX_PEERS);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  for (int q = 0; q < g->world; ++q) {
    if ((n > 0 && !g->grads[q]) || !g->signals[q] || (reinterpret_cast<uintptr_t>(g->grads[q]) & 15u)) {
      set_error("%s: gradient buffer or signal words of rank %d missing or not 16-byte aligned", who, q);
      return EGS_ERR_INVALID_ARGUMENT;
    }
    L.grads[q] = g->grads[q];
    L.src[q] = g->grads[q];
    L.signals[q] = g->signals[q];
  }
  L.n_src = g->world;
  if (reinterpret_cast<uintptr_t>(g->grads_multicast) & 15u) {
    set_error("%s: multicast address not 16-byte aligned", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  L.grads_mc = g->grads_multicast;
  L.timeout_ns = g_timeout_ns.load(std::memory_order_relaxed);
  L.world = g->world;
  L.rank = g->rank;
  const long long n4 = n / 4, per = (n4 + g->world - 1) / g->world;
  L.lo4 = per * g->rank < n4 ? per * g->rank : n4;
  L.hi4 = L.lo4 + per < n4 ? L.lo4 + per : n4;
  return EGS_OK;
}

// The flag barriers need every CTA of the grid on an SM at the same time (a CTA that is not resident
// cannot signal, and the resident ones spin for it). How many fit is a property of the device and of
// the kernel's resources: SM count x CTAs per SM from the occupancy calculator, looked up once per
// device and kernel. A caller's max_ctas beyond that is refused at once instead of timing out on the
// GPU. (Other kernels holding SMs on other streams only delay the exchange: their CTAs retire.)
template <typename K>
static int resident_cta_limit(K kernel, int smem, int* limit) {
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, int> cached;   // (kernel, device) -> CTAs
  int dev = 0;
  EGS_CUDA_TRY(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(mu);
  const auto key = std::make_pair(reinterpret_cast<const void*>(kernel), dev);
  auto it = cached.find(key);
  if (it == cached.end()) {
    int sms = 0, per_sm = 0;
    EGS_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (smem > 0) EGS_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    EGS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, EX_THREADS, smem));
    if (sms * per_sm <= 0) {
      set_error("exchange: the kernel does not fit on device %d", dev);
      return EGS_ERR_CUDA;
    }
    it = cached.emplace(key, sms * per_sm).first;
  }
  *limit = it->second;
  return EGS_OK;
}

template <typename K>
static int exchange_grid(const char* who, K kernel, int smem, const ExLaunch& L, int max_ctas, long long min_ctas,
                         unsigned* grid) {
  int limit = 0;
  if (int rc = resident_cta_limit(kernel, smem, &limit)) return rc;
  if (max_ctas > limit) {
    set_error("%s: max_ctas %d exceeds the %d CTAs of this kernel that can be resident at once on this device", who,
              max_ctas, limit);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  long long ctas = (L.hi4 - L.lo4 + EX_THREADS - 1) / EX_THREADS;   // (more than the unrolled loop needs: harmless)
  if (ctas < min_ctas) ctas = min_ctas;
  const long long cap = max_ctas > 0 ? max_ctas : limit;
  if (ctas > cap) ctas = cap;
  *grid = (unsigned)(ctas < 1 ? 1 : ctas);   // an empty slice still takes part in the barriers
  return EGS_OK;
}

template <typename K>
static int launch_exchange(const char* who, K kernel, const ExLaunch& L, int max_ctas, cudaStream_t st) {
  unsigned grid = 0;
  if (int rc = exchange_grid(who, kernel, 0, L, max_ctas, 0, &grid)) return rc;
  kernel<<<grid, EX_THREADS, 0, st>>>(L);
  EGS_AFTER_LAUNCH(st, false);
  return EGS_OK;
}

}  // namespace egs

extern "C" long long egs_exchange_shard_floats(long long n_floats, int world) {
  if (world < 1 || n_floats < 0) return 0;
  const long long n4 = (n_floats + 3) / 4;
  return (n4 + world - 1) / world * 4;
}

extern "C" int egs_exchange_sum(const EgsPeerGroup* group, long long n_floats, float scale, uint32_t epoch,
                                int max_ctas, void* stream) {
  using namespace egs;
  ExLaunch L = {};
  if (int rc = fill_group("egs_exchange_sum", group, n_floats, L)) return rc;
  L.scale = scale;
  L.epoch = epoch;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  static const char* who = "egs_exchange_sum";
  if (L.grads_mc) return launch_exchange(who, exchange_mc_kernel<false>, L, max_ctas, st);
  if (L.world <= 2) return launch_exchange(who, exchange_kernel<false, 2>, L, max_ctas, st);
  if (L.world <= 4) return launch_exchange(who, exchange_kernel<false, 4>, L, max_ctas, st);
  return launch_exchange(who, exchange_kernel<false, 8>, L, max_ctas, st);
}

template <typename K>
static int launch_factored(const char* who, K kernel, const egs::ExLaunch& L, const egs::ShFactorLaunch& F, int max_ctas,
                           cudaStream_t st) {
  using namespace egs;
  unsigned grid = 0;
  const long long want = ((long long)F.P + 32 * SF_WARPS - 1) / (32 * SF_WARPS);   // one group of 32 Gaussians per warp
  if (int rc = exchange_grid(who, kernel, SF_SMEM, L, max_ctas, want, &grid)) return rc;
  kernel<<<grid, EX_THREADS, SF_SMEM, st>>>(L, F);
  EGS_AFTER_LAUNCH(st, false);
  return EGS_OK;
}

extern "C" int egs_exchange_sum_factored(const EgsPeerGroup* group, long long n_floats, const EgsShFactors* sh,
                                         float scale, uint32_t epoch, int max_ctas, void* stream) {
  using namespace egs;
  static const char* who = "egs_exchange_sum_factored";
  ExLaunch L = {};
  if (int rc = fill_group(who, group, n_floats, L)) return rc;
  if (!sh || sh->P < 0 || sh->sh_degree < 0 || sh->sh_degree > 3 || sh->sh_coeffs < (sh->sh_degree + 1) * (sh->sh_degree + 1) ||
      sh->sh_coeffs > 16 || (sh->P > 0 && (!sh->means3D || !sh->dL_dsh)) ||
      (reinterpret_cast<uintptr_t>(sh->dL_dsh) & 15u)) {
    set_error("%s: bad SH description (degree 0..3, (degree+1)^2 <= sh_coeffs <= 16, means3D, 16-byte aligned dL_dsh)", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  ShFactorLaunch F = {};
  F.P = sh->P; F.D = sh->sh_degree; F.M = sh->sh_coeffs;
  F.means3D = sh->means3D;
  F.dsh = sh->dL_dsh;
  if (sh->n_views < 0 || sh->n_views > EGS_MAX_PEERS) {
    set_error("%s: n_views must be in [0, %d]", who, EGS_MAX_PEERS);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  F.n = sh->n_views > 0 ? sh->n_views : L.world;
  F.phases = debug_options().exchange_phases;
  for (int q = 0; q < F.n; ++q) {
    if (!sh->drgb[q] || (reinterpret_cast<uintptr_t>(sh->drgb[q]) & 15u)) {
      set_error("%s: dL/dRGB slot of view %d missing or not 16-byte aligned", who, q);
      return EGS_ERR_INVALID_ARGUMENT;
    }
    F.drgb[q] = sh->drgb[q];
  }
  if (sh->n_views > 0) {   // a batch of views: their flat buffers are the sources, group->grads the destinations
    for (int q = 0; q < F.n; ++q) {
      if ((n_floats > 0 && !sh->view_flat[q]) || (reinterpret_cast<uintptr_t>(sh->view_flat[q]) & 15u)) {
        set_error("%s: flat buffer of view %d missing or not 16-byte aligned", who, q);
        return EGS_ERR_INVALID_ARGUMENT;
      }
      L.src[q] = sh->view_flat[q];
    }
    L.n_src = F.n;
    bool per_rank = F.n == L.world;   // one view per rank, written into the group's own buffers?
    for (int q = 0; q < F.n && per_rank; ++q) per_rank = sh->view_flat[q] == group->grads[q];
    if (!per_rank) L.grads_mc = nullptr;   // the switch reduces one buffer per rank only
  }
  L.scale = scale;
  L.epoch = epoch;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (L.grads_mc) return launch_factored(who, exchange_factored_kernel<true, 2>, L, F, max_ctas, st);
  if (L.n_src <= 2) return launch_factored(who, exchange_factored_kernel<false, 2>, L, F, max_ctas, st);
  if (L.n_src <= 4) return launch_factored(who, exchange_factored_kernel<false, 4>, L, F, max_ctas, st);
  return launch_factored(who, exchange_factored_kernel<false, 8>, L, F, max_ctas, st);
}

extern "C" int egs_exchange_set_timeout_ms(unsigned long long ms) {
  using namespace egs;
  if (ms == 0) {
    set_error("egs_exchange_set_timeout_ms: the timeout must be positive");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  g_timeout_ns.store(ms * 1000000ull, std::memory_order_relaxed);
  return EGS_OK;
}

extern "C" int egs_exchange_adam_step(const EgsPeerGroup* group, const EgsExchangeAdamArgs* adam, long long n_floats,
                                      float scale, uint32_t epoch, int max_ctas, void* stream) {
  using namespace egs;
  static const char* who = "egs_exchange_adam_step";
  ExLaunch L = {};
  if (int rc = fill_group(who, group, n_floats, L)) return rc;
  if (!adam || (n_floats > 0 && (!adam->exp_avg || !adam->exp_avg_sq)) || adam->step < 1 || adam->n_segments < 0 ||
      adam->n_segments > EGS_EXCHANGE_MAX_SEGMENTS || (adam->n_segments > 0 && !adam->segments)) {
    set_error("%s: moments missing, step < 1 or more than %d segments", who, EGS_EXCHANGE_MAX_SEGMENTS);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  for (int q = 0; q < L.world; ++q) {
    if ((n_floats > 0 && !adam->params[q]) || (reinterpret_cast<uintptr_t>(adam->params[q]) & 15u)) {
      set_error("%s: parameter buffer of rank %d missing or not 16-byte aligned", who, q);
      return EGS_ERR_INVALID_ARGUMENT;
    }
    L.params[q] = adam->params[q];
  }
  if ((reinterpret_cast<uintptr_t>(adam->exp_avg) | reinterpret_cast<uintptr_t>(adam->exp_avg_sq)) & 15u) {
    set_error("%s: moments must be 16-byte aligned", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  // scalar prologue in double like torch's python-side bias corrections (see optim.cu)
  const double bc1 = 1.0 - pow(adam->beta1, (double)adam->step);
  const double bc2 = 1.0 - pow(adam->beta2, (double)adam->step);
  long long prev_end = 0;
  for (int k = 0; k < adam->n_segments; ++k) {
    const EgsExchangeSegment& s = adam->segments[k];
    if (s.begin < prev_end || s.end < s.begin || s.end > n_floats || (s.begin & 3) || (s.end & 3) || s.period < 0 ||
        (s.period > 0 && (s.split < 0 || s.split > s.period))) {
      set_error("%s: segment %d is not an ascending 4-float-aligned range inside the buffer", who, k);
      return EGS_ERR_INVALID_ARGUMENT;
    }
    prev_end = s.end;
    ExSegmentDev& d = L.seg[k];
    d.begin4 = s.begin / 4;
    d.end4 = s.end / 4;
    d.neg_step_a = (float)(-(s.lr_a / bc1));
    d.neg_step_b = (float)(-(s.lr_b / bc1));
    d.period = s.period;
    d.split = s.split;
  }
  L.n_segments = adam->n_segments;
  L.exp_avg = adam->exp_avg;
  L.exp_avg_sq = adam->exp_avg_sq;
  L.one_minus_b1 = (float)(1.0 - adam->beta1);
  L.b2 = (float)adam->beta2;
  L.one_minus_b2 = (float)(1.0 - adam->beta2);
  L.eps = (float)adam->eps;
  L.bc2_sqrt = (float)sqrt(bc2);
  L.scale = scale;
  L.epoch = epoch;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // through the switch only when both buffers have a multicast mapping
  L.params_mc = adam->params_multicast;
  if (!L.params_mc || (reinterpret_cast<uintptr_t>(L.params_mc) & 15u)) L.grads_mc = L.params_mc = nullptr;
  if (L.grads_mc) return launch_exchange(who, exchange_mc_kernel<true>, L, max_ctas, st);
  if (L.world <= 2) return launch_exchange(who, exchange_kernel<true, 2>, L, max_ctas, st);
  if (L.world <= 4) return launch_exchange(who, exchange_kernel<true, 4>, L, max_ctas, st);
  return launch_exchange(who, exchange_kernel<true, 8>, L, max_ctas, st);
}
