// abi.cu -- the extern "C" entry points of include/egs_rasterizer.h: argument validation,
// workspace carving and stage orchestration on the caller's stream.
//
// Orchestration mirrors CudaRasterizer::Rasterizer::forward / backward
// (DGR/cuda_rasterizer/rasterizer_impl.cu:197-339, 343-447): preprocess -> scan -> one D2H read
// of num_rendered -> binning workspace -> duplicate -> sort -> ranges -> blend; and for backward
// tile blend -> per-Gaussian chain. Unlike the reference everything is enqueued on the stream
// the caller passes (the reference uses the legacy default stream).
#include "common.cuh"

#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <mutex>
#include <vector>

namespace egs {

static thread_local char g_error[512] = "";
static std::atomic<unsigned long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}
void count_launch(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ---- stage timing ---------------------------------------------------------------------------
struct ProfRecord {
  int stage;
  int device;   // events belong to the device they were created on
  cudaEvent_t a, b;
  unsigned long long launches_before;
  unsigned long long launches;
};
static std::atomic<bool> g_prof_on{false};
static std::mutex g_prof_mu;
static std::vector<ProfRecord*> g_prof_pending;
static std::vector<ProfRecord*> g_prof_free;

bool profiling_enabled() { return g_prof_on.load(std::memory_order_relaxed); }

void profile_begin(cudaStream_t stream, int stage, void** token) {
  ProfRecord* r = nullptr;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    *token = nullptr;
    return;
  }
  {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    for (size_t i = g_prof_free.size(); i-- > 0;) {   // a pooled pair is reused only on its own device
      if (g_prof_free[i]->device == dev) {
        r = g_prof_free[i];
        g_prof_free.erase(g_prof_free.begin() + (long)i);
        break;
      }
    }
  }
  if (!r) {
    r = new ProfRecord();
    r->device = dev;
    if (cudaEventCreate(&r->a) != cudaSuccess || cudaEventCreate(&r->b) != cudaSuccess) {
      delete r;
      *token = nullptr;
      return;
    }
  }
  r->stage = stage;
  r->launches_before = g_launches.load(std::memory_order_relaxed);
  cudaEventRecord(r->a, stream);
  *token = r;
}

void profile_end(cudaStream_t stream, void* token) {
  ProfRecord* r = static_cast<ProfRecord*>(token);
  cudaEventRecord(r->b, stream);
  r->launches = g_launches.load(std::memory_order_relaxed) - r->launches_before;
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_pending.push_back(r);
}

static const char* kStageNames[ST_COUNT] = {"preprocess_fwd", "scan", "duplicate_keys", "sort_histogram",
                                            "sort_passes", "tile_ranges", "blend_fwd", "bwd_zero",
                                            "blend_bwd", "preprocess_bwd", "adam", "se3",
                                            "model_fwd", "model_bwd", "image_loss", "depth_order"};

// Pinned read-back slot of a forward call (word 0: instance count, 1: presort veto, 2: prefilter
// fault), one per host thread AND device: the event that marks the arrival belongs to the device it
// was created on, so a thread that renders on cuda:0 and later on cuda:1 needs one of each. Created
// on first use with that device current and kept for the life of the process.
constexpr int kMaxDevices = 64;
struct HostCountSlot {
  uint32_t* host = nullptr;
  cudaEvent_t scanned = nullptr, arrived = nullptr;
  cudaStream_t side = nullptr;   // the read-back copy runs beside the caller's stream, not in it
};
static HostCountSlot* host_count_slot() {
  static thread_local HostCountSlot slots[kMaxDevices];
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return nullptr;
  HostCountSlot& slot = slots[dev];
  if (!slot.host) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, 64, cudaHostAllocDefault) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&slot.arrived, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&slot.scanned, cudaEventDisableTiming) != cudaSuccess ||
        cudaStreamCreateWithFlags(&slot.side, cudaStreamNonBlocking) != cudaSuccess) {
      if (slot.arrived) cudaEventDestroy(slot.arrived);
      if (slot.scanned) cudaEventDestroy(slot.scanned);
      slot.arrived = slot.scanned = nullptr;
      cudaFreeHost(p);
      return nullptr;
    }
    slot.host = static_cast<uint32_t*>(p);
  }
  return &slot;
}

// ---- test / A-B switches (egs_debug_option) ---------------------------------------------------
static std::atomic<int> g_dbg_sort{SORT_AUTO}, g_dbg_presort{-1}, g_dbg_scramble{0}, g_dbg_fwd_v2{0}, g_dbg_xphases{0}, g_dbg_bwd_ppl{0};
DebugOptions debug_options() {
  DebugOptions o;
  o.sort = g_dbg_sort.load(std::memory_order_relaxed);
  o.presort = g_dbg_presort.load(std::memory_order_relaxed);
  o.presort_scramble = g_dbg_scramble.load(std::memory_order_relaxed);
  o.fwd_v2 = g_dbg_fwd_v2.load(std::memory_order_relaxed);
  o.exchange_phases = g_dbg_xphases.load(std::memory_order_relaxed);
  o.bwd_ppl = g_dbg_bwd_ppl.load(std::memory_order_relaxed);
  return o;
}

static int check_pairs(const char* who, const void* shs, const void* colors, const void* scales,
                       const void* rotations, const void* cov) {
  if ((shs == nullptr) == (colors == nullptr)) {
    set_error("%s: provide exactly one of shs / colors_precomp", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  const bool sr = scales != nullptr && rotations != nullptr;
  const bool sr_any = scales != nullptr || rotations != nullptr;
  if ((!sr && cov == nullptr) || (sr_any && cov != nullptr)) {
    set_error("%s: provide exactly one of the scale/rotation pair or cov3D_precomp", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  return EGS_OK;
}

}  // namespace egs

using namespace egs;

extern "C" size_t egs_geom_bytes(int P) { return GeomLayout((size_t)(P < 0 ? 0 : P)).total; }
extern "C" size_t egs_binning_bytes(int R, int W, int H) {
  (void)W; (void)H;
  return BinningLayout((size_t)(R < 0 ? 0 : R)).total;
}
extern "C" size_t egs_image_bytes(int W, int H) { return ImageLayout(W, H).total; }
// inverse of egs_binning_bytes on capacities (multiples of SORT_TILE): the size is strictly
// increasing in the number of sort blocks, so a binary search recovers it
extern "C" int egs_binning_capacity(size_t bytes) {
  size_t lo = 0, hi = (size_t)1 << 18;   // blocks; 2^18 * 4096 = 2^30 instances
  while (lo < hi) {
    const size_t mid = (lo + hi) / 2;
    if (BinningLayout(mid * SORT_TILE).total < bytes) lo = mid + 1; else hi = mid;
  }
  return (int)(lo * SORT_TILE);
}
extern "C" size_t egs_backward_scratch_bytes(int P) { return BackwardScratch::bytes((size_t)(P < 0 ? 0 : P)); }

// Raw-parameter mode: the five derived inputs must be absent and their raw sources present.
static int check_raw_model(const char* who, const EgsModelArgs* m, int P, int sh_degree, const void* shs,
                           const void* colors, const void* opacities, const void* scales, const void* rotations,
                           const void* cov) {
  if (shs || opacities || scales || rotations || cov) {
    set_error("%s: opacities / shs / scales / rotations / cov3D_precomp must be NULL (they come from the model)", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (m->P != P || !m->scaling_raw || !m->rotation_raw || !m->opacity_raw || m->row0_copies < 0 ||
      (reinterpret_cast<uintptr_t>(m->rotation_raw) & 15u) != 0) {
    set_error("%s: model needs P == args->P, scaling_raw, rotation_raw (16-byte aligned) and opacity_raw", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (!colors) {
    if (!m->features_dc || m->sh_coeffs < 1 || m->sh_coeffs > 16 || (m->sh_coeffs > 1 && !m->features_rest) ||
        m->sh_coeffs < (sh_degree + 1) * (sh_degree + 1)) {
      set_error("%s: model features missing or sh_coeffs %d unusable for degree %d", who, m->sh_coeffs, sh_degree);
      return EGS_ERR_INVALID_ARGUMENT;
    }
  }
  return EGS_OK;
}

static int forward_impl(const char* who, const EgsForwardArgs* args, const EgsModelArgs* model, void* stream_v,
                        int* num_rendered) {
  if (!args || !num_rendered) {
    set_error("%s: NULL args", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  const EgsForwardArgs& a = *args;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  *num_rendered = 0;
  if (a.P < 0 || a.width <= 0 || a.height <= 0 || a.sh_degree < 0 || a.sh_degree > 3) {
    set_error("%s: bad P / image size / sh_degree", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (!a.out_color || !a.out_depth || !a.out_alpha || !a.background) {
    set_error("%s: NULL output or background", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  const size_t N = (size_t)a.width * a.height;
  if (a.P == 0) {
    // rasterize_points.cu:84 skips the whole pipeline and returns the zero-filled outputs
    EGS_CUDA_TRY(cudaMemsetAsync(a.out_color, 0, 3 * N * sizeof(float), stream));
    EGS_CUDA_TRY(cudaMemsetAsync(a.out_depth, 0, N * sizeof(float), stream));
    EGS_CUDA_TRY(cudaMemsetAsync(a.out_alpha, 0, N * sizeof(float), stream));
    return EGS_OK;
  }
  if (model != nullptr) {
    if (int rc = check_raw_model(who, model, a.P, a.sh_degree, a.shs, a.colors_precomp, a.opacities, a.scales,
                                 a.rotations, a.cov3D_precomp))
      return rc;
  } else if (int rc = check_pairs(who, a.shs, a.colors_precomp, a.scales, a.rotations, a.cov3D_precomp)) {
    return rc;
  }
  if (!a.means3D || (!a.opacities && !model) || !a.viewmatrix || !a.projmatrix || !a.campos || !a.radii ||
      !a.geom_alloc || !a.binning_alloc || !a.image_alloc) {
    set_error("%s: NULL required input", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (a.shs && a.sh_coeffs < (a.sh_degree + 1) * (a.sh_degree + 1)) {
    set_error("%s: sh_coeffs %d too small for degree %d", who, a.sh_coeffs, a.sh_degree);
    return EGS_ERR_INVALID_ARGUMENT;
  }

  void* geom_mem = a.geom_alloc(a.geom_ctx, egs_geom_bytes(a.P));
  void* img_mem = a.image_alloc(a.image_ctx, egs_image_bytes(a.width, a.height));
  if (!geom_mem || !img_mem) {
    set_error("%s: workspace callback returned NULL", who);
    return EGS_ERR_ALLOC;
  }
  const GeomPtrs g(geom_mem, (size_t)a.P);
  const ImagePtrs im(img_mem, a.width, a.height);

  const DebugOptions dbg = debug_options();
  const bool fits = packed_path_fits(a.P, a.width, a.height);   // tile | depth | index in one 64-bit word
  const bool forced = dbg.sort != SORT_AUTO;
  // Front-to-back presort of the Gaussians (binning_buckets.cu): the instances are then emitted in
  // depth order, carry no depth bits and a bucket only needs a stable split on a few tile bits.
  // Sorting the Gaussians costs ~0.06 ms plus 0.04 ms per million and saves ~0.02 ms per million
  // INSTANCES, so it is used from 3M expected instances on (the caller's guess; without one, from 300k
  // Gaussians on), and whenever the full word does not fit (more than 2^(32 - tile bits) Gaussians).
  // egs_debug_option("presort", "0" / "1") forces it off / on (A/B, tests); the forced sort paths
  // never use it.
  const bool presort_wanted = dbg.presort == 0   ? false
                              : dbg.presort == 1 ? a.P >= 64
                              : !fits            ? a.P >= 64
                              : a.num_rendered_hint > 0 ? a.num_rendered_hint >= 3000000
                                                        : a.P >= 300000;
  const bool presort = presort_wanted && !forced && presorted_path_ok(a.P, a.width, a.height);
  const uint32_t* order = presort ? g.sort_vals[0] : nullptr;

  // Speculative run: with a caller-provided guess the binning workspace is sized from the guess and
  // binning + blend are enqueued before the count is known; the kernels read the count (and the veto)
  // on the device and skip themselves if it does not fit. Only the bucket path takes part.
  const bool speculate = a.num_rendered_hint > 0 && !forced && !a.debug &&
                         (presort || bucket_path_ok(a.P, a.num_rendered_hint, a.width, a.height));
  int capacity = 0;
  void* bin_mem = nullptr;
  if (speculate) {
    capacity = a.num_rendered_hint;
    bin_mem = a.binning_alloc(a.binning_ctx, egs_binning_bytes(capacity, a.width, a.height));
    if (!bin_mem) {
      set_error("%s: binning workspace callback returned NULL", who);
      return EGS_ERR_ALLOC;
    }
  }

  // Everything that must start from zero is cleared here, in front of the first kernel: a memset
  // between two kernels costs the stream ~4 us each time.
  EGS_CUDA_TRY(cudaMemsetAsync(g.zero_begin, 0, g.zero_bytes, stream));
  if (speculate)
    if (int rc = zero_binning_buckets(BinningPtrs(bin_mem, (size_t)capacity), stream)) return rc;

  if (int rc = launch_preprocess_forward(a, model, g, stream)) return rc;
  if (presort)
    if (int rc = launch_depth_order(g, a.P, a.radii, stream, a.debug)) return rc;
  if (int rc = launch_scan_tiles(g, a.P, order, stream, a.debug)) return rc;

  // The one device->host read per forward (the reference's is rasterizer_impl.cu:281-282): the
  // instance count, the presort's veto and the prefilter fault are one 12-byte record; it is copied
  // to a pinned slot on a side stream that waits for the scan, so that the caller's stream runs on
  // into the binning kernels instead of idling through the copy.
  HostCountSlot* slot = host_count_slot();
  if (!slot) {
    set_error("%s: cannot allocate the pinned read-back slot", who);
    return EGS_ERR_CUDA;
  }
  const uint32_t* n_dev = g.readback;
  const uint32_t* veto_dev = presort ? presort_veto_flag(g) : nullptr;
  EGS_CUDA_TRY(cudaEventRecord(slot->scanned, stream));
  EGS_CUDA_TRY(cudaStreamWaitEvent(slot->side, slot->scanned, 0));
  EGS_CUDA_TRY(cudaMemcpyAsync(slot->host, g.readback, 3 * sizeof(uint32_t), cudaMemcpyDeviceToHost, slot->side));
  EGS_CUDA_TRY(cudaEventRecord(slot->arrived, slot->side));

  bool enqueued = false;
  if (speculate) {
    const BinningPtrs b(bin_mem, (size_t)capacity);
    const InstanceCount count{n_dev, (uint32_t)capacity, veto_dev};
    if (int rc = launch_binning_buckets(g, b, im, a.radii, order, presort, a.P, capacity, count, a.width, a.height, stream,
                                        false, /*zeroed=*/true))
      return rc;
    if (int rc = launch_blend_forward(a, g, b.vals[0], im, stream)) return rc;
    enqueued = true;
  }

  // wait for the count only (it arrived right after the scan; the GPU keeps working on what was
  // enqueued above)
  EGS_CUDA_TRY(cudaEventSynchronize(slot->arrived));
  const uint32_t R_host = slot->host[0];
  const bool vetoed = presort && slot->host[1] != 0u;
  if (a.prefiltered && slot->host[2] != 0u) {
    // the reference prints this and traps the kernel (auxiliary.h:154-161); here it is an error code
    set_error("%s: Point is filtered although prefiltered is set. This shouldn't happen!", who);
    return EGS_ERR_PREFILTERED;
  }
  if (R_host >= (1u << 30)) {
    set_error("%s: num_rendered %u exceeds the supported 2^30", who, R_host);
    return EGS_ERR_TOO_LARGE;
  }
  const int R = (int)R_host;
  *num_rendered = R;
  if (enqueued && R <= capacity && !vetoed) return EGS_OK;

  // exact run (no guess, guess too small, presort vetoed, forced path, or a path that does not speculate)
  bin_mem = a.binning_alloc(a.binning_ctx, egs_binning_bytes(R, a.width, a.height));
  if (!bin_mem) {
    set_error("%s: binning workspace callback returned NULL", who);
    return EGS_ERR_ALLOC;
  }
  const BinningPtrs b(bin_mem, (size_t)R);
  // Binning paths, all with identical results:
  //   buckets, presorted   tile | index words, one global partition + a register-resident split per bucket
  //   buckets, full words  tile | depth | index words, one global partition + one CTA sort per bucket
  //   packed   full words, five global onesweep passes (R too large for the bucket sorter)
  //   pairs    keys + value array through the same onesweep kernel (the full word does not fit 64 bits
  //            and the presort is off or gave up)
  //   pairs8   the first, plain 8-bit pair sort of binning.cu, kept as an independent cross-check
  // `order` stays what the scan ran over in every case (any permutation serves the unsorted paths).
  const InstanceCount count{n_dev, (uint32_t)R, nullptr};
  if (dbg.sort == SORT_PAIRS8) {
    if (int rc = launch_binning(g, b, im, a.radii, a.P, R, a.width, a.height, stream, a.debug)) return rc;
  } else if (presort && !vetoed) {
    if (int rc = launch_binning_buckets(g, b, im, a.radii, order, true, a.P, R, count, a.width, a.height, stream, a.debug,
                                        false))
      return rc;
  } else if (!fits || dbg.sort == SORT_PAIRS) {
    if (int rc = launch_binning_packed(g, b, im, a.radii, order, a.P, R, a.width, a.height, true, stream, a.debug)) return rc;
  } else if (bucket_path_ok(a.P, R, a.width, a.height) && dbg.sort != SORT_PACKED) {
    if (int rc = launch_binning_buckets(g, b, im, a.radii, order, false, a.P, R, count, a.width, a.height, stream, a.debug,
                                        false))
      return rc;
  } else {
    if (int rc = launch_binning_packed(g, b, im, a.radii, order, a.P, R, a.width, a.height, false, stream, a.debug)) return rc;
  }
  return launch_blend_forward(a, g, b.vals[0], im, stream);
}

extern "C" int egs_forward(const EgsForwardArgs* args, void* stream, int* num_rendered) {
  return forward_impl("egs_forward", args, nullptr, stream, num_rendered);
}

extern "C" int egs_forward_raw(const EgsForwardArgs* args, const EgsModelArgs* model, void* stream,
                               int* num_rendered) {
  if (!model) {
    set_error("egs_forward_raw: NULL model");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  return forward_impl("egs_forward_raw", args, model, stream, num_rendered);
}

static int backward_impl(const char* who, const EgsBackwardArgs* args, const EgsModelArgs* model,
                         const EgsModelGrads* grads, void* stream_v) {
  if (!args) {
    set_error("%s: NULL args", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  const EgsBackwardArgs& a = *args;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  if (a.P < 0 || a.width <= 0 || a.height <= 0 || a.num_rendered < 0) {
    set_error("%s: bad sizes", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (a.P == 0) return EGS_OK;
  if (model != nullptr) {
    if (int rc = check_raw_model(who, model, a.P, a.sh_degree, a.shs, a.colors_precomp, nullptr, a.scales,
                                 a.rotations, a.cov3D_precomp))
      return rc;
    if (!grads || !grads->d_scaling_raw || !grads->d_rotation_raw || !grads->d_opacity_raw ||
        (reinterpret_cast<uintptr_t>(grads->d_rotation_raw) & 15u) != 0 ||
        (!a.colors_precomp && !a.dL_drgb && (!grads->d_features_dc || (model->sh_coeffs > 1 && !grads->d_features_rest)))) {
      set_error("%s: missing raw-parameter gradient output (d_rotation_raw must be 16-byte aligned)", who);
      return EGS_ERR_INVALID_ARGUMENT;
    }
  } else if (int rc = check_pairs(who, a.shs, a.colors_precomp, a.scales, a.rotations, a.cov3D_precomp)) {
    return rc;
  }
  if (!a.means3D || !a.radii || !a.out_alpha || !a.geom_buffer || !a.image_buffer ||
      (a.num_rendered > 0 && !a.binning_buffer) || !a.dL_dout_color || !a.scratch ||
      (!a.colors_only && (!a.dL_dmeans2D || (!a.dL_dopacity && !model) || !a.dL_dmeans3D)) ||
      !a.background || !a.viewmatrix || !a.projmatrix || !a.campos) {
    set_error("%s: NULL required argument", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (a.dL_drgb && (a.colors_precomp || a.colors_only)) {
    set_error("%s: dL_drgb (factored SH gradient) needs SH colours", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (a.drgb_n_remote < 0 || a.drgb_n_remote > EGS_MAX_PEERS || ((a.drgb_n_remote > 0 || a.drgb_multicast) && !a.dL_drgb) ||
      (reinterpret_cast<uintptr_t>(a.dL_drgb) & 15u) || (reinterpret_cast<uintptr_t>(a.drgb_multicast) & 15u)) {
    set_error("%s: bad dL_drgb destinations (16-byte aligned slots, at most %d remote ones, dL_drgb itself set)", who,
              EGS_MAX_PEERS);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  for (int q = 0; q < a.drgb_n_remote; ++q)
    if (!a.drgb_remote[q] || (reinterpret_cast<uintptr_t>(a.drgb_remote[q]) & 15u)) {
      set_error("%s: drgb_remote[%d] missing or not 16-byte aligned", who, q);
      return EGS_ERR_INVALID_ARGUMENT;
    }
  if (a.colors_only && (!a.colors_precomp || model != nullptr)) {
    set_error("%s: colors_only needs colors_precomp (and is not a raw-parameter mode)", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if ((a.colors_precomp && !a.dL_dcolors) ||
      (!a.colors_only && ((a.cov3D_precomp && !a.dL_dcov3D) || (a.shs && !a.dL_dsh && !a.dL_drgb) ||
                          (a.scales && (!a.dL_dscales || !a.dL_drotations))))) {
    set_error("%s: missing gradient output for a provided input", who);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  const GeomPtrs g(const_cast<void*>(a.geom_buffer), (size_t)a.P);
  const ImagePtrs im(const_cast<void*>(a.image_buffer), a.width, a.height);
  const BackwardScratch s(a.scratch);

  {
    StageTimer t(stream, ST_BWD_ZERO);
    EGS_CUDA_TRY(cudaMemsetAsync(s.acc, 0, (size_t)a.P * 48, stream));
  }
  if (a.num_rendered > 0) {
    const size_t layout_for = a.binning_bytes ? (size_t)egs_binning_capacity(a.binning_bytes) : (size_t)a.num_rendered;
    const BinningPtrs b(const_cast<void*>(a.binning_buffer), layout_for);
    const uint32_t* point_list = b.vals[0];
    if (int rc = launch_blend_backward(a, g, point_list, im, s, stream)) return rc;
  }
  return launch_preprocess_backward(a, model, grads, g, s, stream);
}

extern "C" int egs_backward(const EgsBackwardArgs* args, void* stream) {
  return backward_impl("egs_backward", args, nullptr, nullptr, stream);
}

extern "C" int egs_backward_raw(const EgsBackwardArgs* args, const EgsModelArgs* model, const EgsModelGrads* grads,
                                void* stream) {
  if (!model) {
    set_error("egs_backward_raw: NULL model");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  return backward_impl("egs_backward_raw", args, model, grads, stream);
}

// ---- backward with the scratch inside the library image --------------------------------------
// For hosts that may not allocate on the device (see the header). The accumulator records sit
// in a file-scope device array; host code may not use that name as an address, so its device
// address is looked up once per device and handed to the ordinary backward path as `scratch`.
static __device__ __align__(256) float4 g_static_backward_records[(size_t)EGS_STATIC_SCRATCH_MAX_GAUSSIANS * 3];

namespace {
std::mutex g_static_scratch_mu;
void* g_static_scratch_addr[kMaxDevices] = {};   // per device, filled on that device's first call
}  // namespace

extern "C" int egs_backward_static_scratch(const EgsBackwardArgs* args, void* stream) {
  if (!args) {
    set_error("egs_backward_static_scratch: NULL args");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (args->P > EGS_STATIC_SCRATCH_MAX_GAUSSIANS) {
    set_error("egs_backward_static_scratch: P = %d exceeds the built-in scratch of %d Gaussians "
              "(EGS_STATIC_SCRATCH_MAX_GAUSSIANS); use egs_backward with a caller-provided scratch",
              args->P, EGS_STATIC_SCRATCH_MAX_GAUSSIANS);
    return EGS_ERR_TOO_LARGE;
  }
  int dev = -1;
  EGS_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("egs_backward_static_scratch: device ordinal %d not supported", dev);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  void* records = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_static_scratch_mu);
    if (g_static_scratch_addr[dev] == nullptr) {
      // one-time set-up for this device (it is the current one): where does the array live here?
      void* p = nullptr;
      EGS_CUDA_TRY(cudaGetSymbolAddress(&p, g_static_backward_records));
      g_static_scratch_addr[dev] = p;
    }
    records = g_static_scratch_addr[dev];
  }
  EgsBackwardArgs with_scratch = *args;   // plain struct copy on the host stack; nothing is allocated
  with_scratch.scratch = records;
  // egs_backward clears the P records it uses before accumulating into them
  return egs_backward(&with_scratch, stream);
}

extern "C" int egs_mark_visible(int P, const float* means3D, const float* viewmatrix,
                                const float* projmatrix, unsigned char* present, void* stream) {
  if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) {
    set_error("egs_mark_visible: NULL argument");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (P == 0) return EGS_OK;
  return launch_mark_visible(P, means3D, viewmatrix, projmatrix, present, static_cast<cudaStream_t>(stream));
}

extern "C" int egs_geom_view(const void* geom_buffer, int P, EgsGeomView* out) {
  if (!geom_buffer || !out || P < 0) return EGS_ERR_INVALID_ARGUMENT;
  const GeomPtrs g(const_cast<void*>(geom_buffer), (size_t)P);
  out->depths = g.depths;
  out->means2D = reinterpret_cast<const float*>(g.means2D);
  out->conic_opacity = reinterpret_cast<const float*>(g.conic_opacity);
  out->rgb_depth = reinterpret_cast<const float*>(g.rgb_depth);
  out->clamped = g.clamped;
  out->tiles_touched = g.tiles_touched;
  out->point_offsets = g.point_offsets;
  return EGS_OK;
}

extern "C" int egs_binning_view(const void* binning_buffer, int R, int W, int H, EgsBinningView* out) {
  if (!binning_buffer || !out || R < 0) return EGS_ERR_INVALID_ARGUMENT;
  const BinningPtrs b(const_cast<void*>(binning_buffer), (size_t)R);
  (void)W; (void)H;
  out->keys_sorted = b.keys[0];
  out->values_sorted = b.vals[0];
  return EGS_OK;
}

extern "C" int egs_binning_keys(const void* geom_buffer, int P, void* binning_buffer, int R, const void* image_buffer,
                                int W, int H, void* stream) {
  if (!geom_buffer || !binning_buffer || !image_buffer || P < 0 || R < 0 || W <= 0 || H <= 0) {
    set_error("egs_binning_keys: NULL blob or bad size");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  if (P == 0 || R == 0) return EGS_OK;
  const GeomPtrs g(const_cast<void*>(geom_buffer), (size_t)P);
  const BinningPtrs b(binning_buffer, (size_t)R);
  const ImagePtrs im(const_cast<void*>(image_buffer), W, H);
  return launch_binning_keys(g, im, b.vals[0], b.keys[0], W, H, static_cast<cudaStream_t>(stream));
}

extern "C" int egs_debug_option(const char* name, const char* value) {
  if (!name) {
    set_error("egs_debug_option: NULL name");
    return EGS_ERR_INVALID_ARGUMENT;
  }
  const bool unset = value == nullptr || value[0] == 0;
  if (strcmp(name, "sort") == 0) {
    int v = SORT_AUTO;
    if (!unset) {
      if (strcmp(value, "packed") == 0) v = SORT_PACKED;
      else if (strcmp(value, "pairs") == 0) v = SORT_PAIRS;
      else if (strcmp(value, "pairs8") == 0) v = SORT_PAIRS8;
      else {
        set_error("egs_debug_option: sort = packed | pairs | pairs8, got '%s'", value);
        return EGS_ERR_INVALID_ARGUMENT;
      }
    }
    g_dbg_sort.store(v, std::memory_order_relaxed);
  } else if (strcmp(name, "presort") == 0) {
    g_dbg_presort.store(unset ? -1 : (value[0] == '0' ? 0 : 1), std::memory_order_relaxed);
  } else if (strcmp(name, "presort_scramble") == 0) {
    g_dbg_scramble.store(unset || value[0] == '0' ? 0 : 1, std::memory_order_relaxed);
  } else if (strcmp(name, "blend_forward") == 0) {
    g_dbg_fwd_v2.store(!unset && strcmp(value, "v2") == 0 ? 1 : 0, std::memory_order_relaxed);
  } else if (strcmp(name, "blend_backward") == 0) {
    g_dbg_bwd_ppl.store(unset ? 0 : value[0] == '2' ? 2 : 4, std::memory_order_relaxed);
  } else if (strcmp(name, "exchange_phases") == 0) {
    g_dbg_xphases.store(unset ? 0 : (value[0] == 'a' ? 1 : 2), std::memory_order_relaxed);
  } else {
    set_error("egs_debug_option: unknown option '%s'", name);
    return EGS_ERR_INVALID_ARGUMENT;
  }
  return EGS_OK;
}

extern "C" int egs_image_view(const void* image_buffer, int W, int H, EgsImageView* out) {
  if (!image_buffer || !out) return EGS_ERR_INVALID_ARGUMENT;
  const ImagePtrs im(const_cast<void*>(image_buffer), W, H);
  out->n_contrib = im.n_contrib;
  out->ranges = reinterpret_cast<const uint32_t*>(im.ranges);
  return EGS_OK;
}

extern "C" int egs_profile_enable(int on) {
  g_prof_on.store(on != 0, std::memory_order_relaxed);
  return EGS_OK;
}

extern "C" int egs_profile_read(double* ms, unsigned long long* count) {
  if (!ms || !count) return EGS_ERR_INVALID_ARGUMENT;
  std::vector<ProfRecord*> take;
  {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    take.swap(g_prof_pending);
  }
  int rc = EGS_OK;
  for (ProfRecord* r : take) {
    float t = 0.f;
    if (cudaEventSynchronize(r->b) == cudaSuccess && cudaEventElapsedTime(&t, r->a, r->b) == cudaSuccess) {
      ms[r->stage] += t;
      count[r->stage] += r->launches;
    } else {
      rc = EGS_ERR_CUDA;
    }
  }
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (ProfRecord* r : take) g_prof_free.push_back(r);
  return rc;
}

extern "C" const char* egs_profile_stage_name(int stage) {
  return (stage >= 0 && stage < ST_COUNT) ? kStageNames[stage] : "";
}

extern "C" unsigned long long egs_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
extern "C" const char* egs_last_error(void) { return g_error; }
extern "C" int egs_abi_version(void) { return EGS_ABI_VERSION; }
