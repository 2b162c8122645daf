// pcseg_api.cu -- C ABI (include/pcseg_b200.h) over the sm_100a kernels.
//
// Host-side plumbing only: argument checks that mirror the reference's CHECKs,
// device buffers, kernel launches on the context's stream and result
// materialisation.  No CPU implementation of any hot-path stage lives here:
// if CUDA is unavailable pcseg_create fails and nothing else can be called.
#include "../../include/pcseg_b200.h"

#include <algorithm>
#include <chrono>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iterator>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "grow_cluster.cuh"
#include "planar_kernel.cuh"
#include "meanshift_grow.cuh"
#include "meanshift_wave.cuh"
#include "normals.cuh"
#include "seeds.cuh"
#include "results_pack.cuh"
#include "plane_estimator.cuh"
#include "msort.cuh"
#include "voxel_cluster.cuh"
#include "grow_cluster_wave.cuh"
#include "grow_cluster_comp.cuh"

using namespace pcseg;

namespace {

constexpr int kGrowBlock = 256;
#ifndef PCSEG_CLUSTER_BLOCK
#define PCSEG_CLUSTER_BLOCK 512
#endif
constexpr int kClusterBlock = PCSEG_CLUSTER_BLOCK;  // queue entries evaluated per step of cluster growth
constexpr int kPipeBlock = 32 * kPipeWarps;  // planar pipeline path (grow_planar_pipe.cuh)
#ifndef PCSEG_WAVE_BLOCK
#define PCSEG_WAVE_BLOCK 512
#endif
#ifndef PCSEG_WAVE2_BLOCK
#define PCSEG_WAVE2_BLOCK 288
#endif
constexpr int kWave2Block = PCSEG_WAVE2_BLOCK;  // planar fast path (grow_planar_wave2.cuh): W = kWave2Block - 64
constexpr int kWaveBlock = PCSEG_WAVE_BLOCK;  // planar fast path (grow_planar_wave.cuh): W = kWaveBlock - 32 entries per window
constexpr int kInvCap = 4096;

// Device memory of the calls' temporaries comes from a per-device cache: the single-cloud entry
// points (mean shift, unorganized clustering, sort) take a dozen buffers per call, and
// cudaMalloc / cudaFree of hundreds of MB cost milliseconds each (measured on C5: 20 - 60 ms of a
// 110 ms call, varying from box to box).  A returned block is kept (up to PCSEG_POOL_MB, default
// 8192, per device) and handed to the next request of about its size.  Returning a block
// synchronises the device first -- what cudaFree did implicitly -- so a block is never reused
// while work that reads it is in flight.  The cache is emptied when the last context of the device
// is destroyed.
struct DevPool {
  static constexpr int kMaxDev = 64;
  std::mutex mu;
  std::multimap<size_t, void*> blocks[kMaxDev];
  size_t held[kMaxDev] = {};
  int live_ctx[kMaxDev] = {};
  static size_t limit() {
    static const size_t v = [] {
      const char* e = getenv("PCSEG_POOL_MB");
      return (size_t)(e ? std::max(0ll, atoll(e)) : 8192ll) << 20;
    }();
    return v;
  }
  static int device() {
    int d = 0;
    if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= kMaxDev) { cudaGetLastError(); return -1; }
    return d;
  }
  void trim(int d, size_t keep) {   // mu held
    while (held[d] > keep && !blocks[d].empty()) {
      auto it = std::prev(blocks[d].end());
      held[d] -= it->first;
      cudaFree(it->second);
      blocks[d].erase(it);
    }
  }
  cudaError_t get(size_t need, void** p, size_t* got) {
    need = need < ((size_t)1 << 20) ? (need + 511) & ~(size_t)511 : (need + (((size_t)1 << 21) - 1)) & ~(((size_t)1 << 21) - 1);
    const int d = device();
    if (d >= 0) {
      std::lock_guard<std::mutex> lock(mu);
      auto it = blocks[d].lower_bound(need);
      if (it != blocks[d].end() && it->first <= need + need / 4 + ((size_t)1 << 20)) {
        *p = it->second; *got = it->first;
        held[d] -= it->first;
        blocks[d].erase(it);
        return cudaSuccess;
      }
    }
    cudaError_t e = cudaMalloc(p, need);
    if (e == cudaErrorMemoryAllocation && d >= 0) {   // give the cached blocks back and try again
      cudaGetLastError();
      { std::lock_guard<std::mutex> lock(mu); trim(d, 0); }
      e = cudaMalloc(p, need);
    }
    *got = need;
    return e;
  }
  void put(void* p, size_t bytes) {
    const int d = device();
    cudaDeviceSynchronize();
    std::lock_guard<std::mutex> lock(mu);
    if (d < 0 || live_ctx[d] <= 0 || bytes > limit() / 2) { cudaFree(p); return; }
    blocks[d].emplace(bytes, p);
    held[d] += bytes;
    trim(d, limit());
  }
  void context_created(int d) { std::lock_guard<std::mutex> lock(mu); if (d >= 0 && d < kMaxDev) ++live_ctx[d]; }
  void context_destroyed(int d) {
    std::lock_guard<std::mutex> lock(mu);
    if (d < 0 || d >= kMaxDev) return;
    if (--live_ctx[d] <= 0) { live_ctx[d] = 0; trim(d, 0); }
  }
};
DevPool g_pool;

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  cudaError_t ensure(size_t need) {
    if (need <= bytes) return cudaSuccess;
    release();
    cudaError_t e = g_pool.get(need, &p, &bytes);
    if (e != cudaSuccess) { p = nullptr; bytes = 0; }
    return e;
  }
  void release() {
    if (p) g_pool.put(p, bytes);
    p = nullptr;
    bytes = 0;
  }
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { release(); }
  template <typename T>
  T* as() const { return static_cast<T*>(p); }
};

}  // namespace

struct pcseg_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string last_error;
  long long launches = 0;
  pcseg_batch* scratch = nullptr;  // cached single-frame workspace
  DevBuf tmp_a, tmp_b;
  // host<->device copies of the batched pipeline run on their own streams so that they
  // overlap the kernels (chunked H2D under the normals stage, normals D2H under growth)
  cudaStream_t copy_in = nullptr, copy_out = nullptr;
};

struct pcseg_regions {
  std::vector<int> inlier_offsets{0};
  std::vector<int> inlier_indices;
  std::vector<int> label_ids;
  std::vector<int> seed_indices;
  std::vector<float> planes, centroids, areas, estimators;
  std::vector<int> boundary_offsets{0};
  std::vector<int> boundary_indices;
  std::vector<int> disc_offsets{0};
  std::vector<int> disc_indices;
  std::vector<int> hull_offsets{0};
  std::vector<float> hull_points;
  std::vector<float> seed_points;
  std::vector<float> bboxes;   // 6 per region (min xyz, max xyz): pcseg_cluster_unorganized
};

struct pcseg_batch {
  pcseg_ctx* ctx = nullptr;
  int n_frames = 0, rows = 0, cols = 0;
  size_t N = 0;
  int qcap = 0, bcap = 0, rcap_planar = 0, rcap_cluster = 0;
  DevBuf xyz, normals, labels, queue, claim, inl_planar, inl_cluster, boundary, bflags;
  DevBuf meta_planar, meta_cluster, stats, inv_list, inv_count, seeds, n_seeds;
  const float* xyz_dev = nullptr;      // where this run's points live on the device
  const float* normals_dev = nullptr;  // where this run's normals live on the device
  bool claim_ready = false;
  // Workspaces made by pcseg_batch_create own a stream set, so that two of them on one context
  // overlap (pcseg_batch_submit / pcseg_batch_wait: the copies of one run under the kernels of
  // the other).  The context's cached single-frame workspace has none and uses the context's.
  cudaStream_t own_stream = nullptr, own_copy_in = nullptr, own_copy_out = nullptr;
  // The normals kernel of a run goes to a LOW-priority stream of its own, the growth stages stay on
  // the (high-priority) main stream: with several workspaces in flight the CTA scheduler then hands
  // freed SM slots to the long, latency-bound growth CTAs of one workspace before the hundreds of
  // thousands of short normals CTAs of another, which fill the issue slots the growth leaves idle
  // instead of running in front of it.  cur_stream: what S() returns while the normals stage is enqueued.
  cudaStream_t own_normals = nullptr, cur_stream = nullptr;
  cudaEvent_t ev_nrm[2] = {nullptr, nullptr};
  bool in_flight = false, wait_copy_out = false;
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  std::vector<cudaEvent_t> chunk_ev;  // 2 per chunk (H2D done, normals done) + 2
  DevBuf prof;
  // region results packed over the batch (results_pack.cuh); allocated on first use
  DevBuf pk_offsets, pk_planar, pk_pinl, pk_pinl_xyz, pk_bnd, pk_bnd_xyz, pk_bdisc, pk_disc_idx, pk_disc_pos,
      pk_hull, pk_cluster, pk_cinl, pk_cinl_xyz, pk_err;
  int* h_meta = nullptr;               // pinned: [5][F+1] offsets, [F][8] stats, 1 pack error word
  // what the run in flight was asked for (pcseg_batch_wait finishes it)
  pcseg_pipeline_params run_params{};
  std::vector<int> run_seeds;          // PCSEG_SEEDS_EXPLICIT list (host copy for a re-run)
  bool run_valid = false;
  int* run_labels_out = nullptr;
  pcseg_batch_results* run_results = nullptr;
  bool cluster_ran = false;
};

namespace {

#define CU_TRY(ctx, expr)                                                        \
  do {                                                                           \
    cudaError_t _e = (expr);                                                     \
    if (_e != cudaSuccess) {                                                     \
      (ctx)->last_error = std::string(#expr) + ": " + cudaGetErrorString(_e);    \
      return PCSEG_ERR_CUDA;                                                     \
    }                                                                            \
  } while (0)

bool is_device_ptr(const void* p) {
  if (!p) return false;
  cudaPointerAttributes at;
  cudaError_t e = cudaPointerGetAttributes(&at, p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

// Brings `bytes` at src (host or device) to the device; returns the device view.
int to_device(pcseg_ctx* c, const void* src, size_t bytes, DevBuf* staging, const void** out, cudaStream_t stream = nullptr) {
  if (is_device_ptr(src)) {
    *out = src;
    return PCSEG_OK;
  }
  CU_TRY(c, staging->ensure(bytes));
  CU_TRY(c, cudaMemcpyAsync(staging->p, src, bytes, cudaMemcpyHostToDevice, stream ? stream : c->stream));
  *out = staging->p;
  return PCSEG_OK;
}

int from_device(pcseg_ctx* c, void* dst, const void* src_dev, size_t bytes) {
  if (dst == src_dev) return PCSEG_OK;
  cudaMemcpyKind k = is_device_ptr(dst) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
  CU_TRY(c, cudaMemcpyAsync(dst, src_dev, bytes, k, c->stream));
  return PCSEG_OK;
}

__global__ void fill_int_kernel(int* p, size_t n, int v) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) p[i] = v;
}

// *flag |= 1 if any label lies outside [lo, hi): the single-cloud fast paths borrow label values outside
// that range for their temporary codes and must not meet them in the caller's grid
__global__ void label_range_kernel(const int* __restrict__ p, size_t n, int lo, int hi, int* flag) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  bool bad = false;
  for (; i < n; i += stride) { const int v = p[i]; bad |= v < lo || v >= hi; }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

// The stream set a workspace enqueues on: its own (pcseg_batch_create) or the context's.
inline cudaStream_t S(const pcseg_batch* b) { return b->cur_stream ? b->cur_stream : b->own_stream ? b->own_stream : b->ctx->stream; }
inline cudaStream_t SIn(const pcseg_batch* b) { return b->own_copy_in ? b->own_copy_in : b->ctx->copy_in; }
inline cudaStream_t SOut(const pcseg_batch* b) { return b->own_copy_out ? b->own_copy_out : b->ctx->copy_out; }

int fill_int(pcseg_ctx* c, int* p, size_t n, int v, cudaStream_t stream = nullptr) {
  if (n == 0) return PCSEG_OK;
  if (!stream) stream = c->stream;
  int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
  fill_int_kernel<<<blocks, 256, 0, stream>>>(p, n, v);
  ++c->launches;
  CU_TRY(c, cudaGetLastError());
  return PCSEG_OK;
}

int check_planar_cfg(const pcseg_planar_config* cfg) {
  // quirk Q8: size() < int and size() % int promote to size_t in the reference;
  // period <= 0 is UB, min_region_inliers < 1 makes NaN seeds emit stale regions.
  if (cfg->plane_model_reestimation_period <= 0 || cfg->min_region_inliers < 1) return PCSEG_ERR_CONFIG;
  return PCSEG_OK;
}
int planar_region_bound(const pcseg_batch* b, const pcseg_planar_config* cfg) {
  // An accepted region has a boundary that spans more than 3 rows and 3 columns
  // (ps/planar_region.cc:91-106) and is 4-connected, so it has at least 5 + 5 - 1 cells.
  return (int)(b->N / (size_t)std::max(cfg->min_region_inliers, 9)) + 1;
}
int check_cluster_cfg(const pcseg_cluster_config* cfg) {
  if (cfg->min_region_inliers < 1 || cfg->half_search_window < 0) return PCSEG_ERR_CONFIG;
  if (cfg->half_search_window > 7) return PCSEG_ERR_UNSUPPORTED;   // 224 window slots; the reference's default is 1
  return PCSEG_OK;
}

int batch_alloc(pcseg_batch* b) {
  pcseg_ctx* c = b->ctx;
  const size_t F = b->n_frames, N = b->N;
  b->qcap = 1024;  // power of two >= N + 1024: queue positions wrap with a mask
  while ((size_t)b->qcap < N + 1024) b->qcap <<= 1;
  b->bcap = (int)(2 * N) + 1024;
  b->rcap_planar = (int)std::min<size_t>(N / 4 + 16, 4096);
  b->rcap_cluster = (int)N + 1;
  CU_TRY(c, b->labels.ensure(F * N * sizeof(int)));
  CU_TRY(c, b->queue.ensure(F * (size_t)b->qcap * 2 * sizeof(int)));
  CU_TRY(c, b->claim.ensure(F * N * sizeof(int)));
  CU_TRY(c, b->inl_planar.ensure(F * N * sizeof(int)));
  CU_TRY(c, b->inl_cluster.ensure(F * N * sizeof(int)));
  CU_TRY(c, b->boundary.ensure(F * (size_t)b->bcap * sizeof(int)));
  CU_TRY(c, b->bflags.ensure(F * (size_t)b->bcap));
  CU_TRY(c, b->meta_planar.ensure(F * (size_t)b->rcap_planar * sizeof(PlanarMeta)));
  CU_TRY(c, b->stats.ensure(F * sizeof(FrameStats)));
  CU_TRY(c, b->inv_list.ensure(F * (size_t)kInvCap * sizeof(int)));
  CU_TRY(c, b->inv_count.ensure(F * sizeof(int)));
#ifdef PCSEG_PROFILE
  CU_TRY(c, b->prof.ensure(F * 16 * sizeof(long long)));
  CU_TRY(c, cudaMemset(b->prof.p, 0, F * 16 * sizeof(long long)));
#endif
  int st = fill_int(c, b->claim.as<int>(), F * N, INT_MAX);
  if (st) return st;
  b->claim_ready = true;
  return PCSEG_OK;
}

}  // namespace

// ===========================================================================
// Stage launchers (all asynchronous on ctx->stream)
// ===========================================================================
namespace {

// Frames [f0, f0 + nf) of the batch (nf < 0: all of them).
int run_normals(pcseg_batch* b, const float* xyz_dev, float* normals_dev, float* curvature_dev,
                const pcseg_normals_params* p, const double sensor_t[3], int r0, int r1, int c0, int c1,
                int f0 = 0, int nf = -1) {
  pcseg_ctx* c = b->ctx;
  cudaStream_t stream = S(b);
  if (r1 <= r0 || c1 <= c0) return PCSEG_OK;
  if (nf < 0) nf = b->n_frames - f0;
  if (nf <= 0) return PCSEG_OK;
  NormalsArgs a;
  a.xyz = xyz_dev + (size_t)f0 * b->N * 3;
  a.normals = normals_dev + (size_t)f0 * b->N * 3;
  a.curvature = curvature_dev ? curvature_dev + (size_t)f0 * b->N : nullptr;
  a.rows = b->rows; a.cols = b->cols; a.n_frames = nf;
  a.r0 = r0; a.r1 = r1; a.c0 = c0; a.c1 = c1;
  a.min2 = p->min_neighbor_distance * p->min_neighbor_distance;  // ps/algorithms.h:111-114
  a.max2 = p->max_neighbor_distance * p->max_neighbor_distance;
  a.diag = p->include_diagonal_neighbors;
  a.min_support = p->min_num_support_neighbors;
  a.sx = (float)sensor_t[0]; a.sy = (float)sensor_t[1]; a.sz = (float)sensor_t[2];
  a.inv_list = b->inv_list.as<int>() + (size_t)f0 * kInvCap;
  a.inv_count = b->inv_count.as<int>() + f0;
  a.inv_cap = kInvCap;
  CU_TRY(c, cudaMemsetAsync(a.inv_count, 0, nf * sizeof(int), stream));
  const int work = (r1 - r0) * (c1 - c0);
  dim3 grid((work + kNormalsBlock - 1) / kNormalsBlock, nf);
  if (a.max2 <= FLT_MAX) normals_kernel<false><<<grid, kNormalsBlock, 0, stream>>>(a);
  else normals_kernel<true><<<grid, kNormalsBlock, 0, stream>>>(a);  // max distance inf/NaN: explicit allFinite() test
  normals_fixup_kernel<<<nf, 256, 0, stream>>>(a);
  c->launches += 2;
  CU_TRY(c, cudaGetLastError());
  return PCSEG_OK;
}

int make_cluster_args(pcseg_batch* b, const float* xyz_dev, const int* seeds_dev, int n_seeds,
                      const pcseg_cluster_config* cfg, int id_offset, int offset_from_planar, ClusterArgs* out) {
  pcseg_ctx* c = b->ctx;
  b->rcap_cluster = (int)(b->N / (size_t)cfg->min_region_inliers) + 1;
  CU_TRY(c, b->meta_cluster.ensure((size_t)b->n_frames * b->rcap_cluster * sizeof(ClusterMeta)));
  ClusterArgs a;
  a.xyz = xyz_dev; a.labels = b->labels.as<int>();
  a.rows = b->rows; a.cols = b->cols;
  a.seeds = seeds_dev; a.n_seeds = seeds_dev ? n_seeds : (int)b->N;
  a.min_inliers = cfg->min_region_inliers; a.sq_thr = cfg->squared_distance_threshold;
  a.half_window = cfg->half_search_window;
  a.id_offset = id_offset; a.id_offset_from_planar = offset_from_planar;
  a.queue = b->queue.as<int>(); a.qcap = b->qcap;
  a.claim = b->claim.as<int>(); a.inliers = b->inl_cluster.as<int>();
  a.meta = b->meta_cluster.as<ClusterMeta>(); a.rcap = b->rcap_cluster;
  a.stats = b->stats.as<FrameStats>();
  *out = a;
  return PCSEG_OK;
}

// fuse_cluster (nullable): cluster stage to run behind each frame's planar stage in the same
// CTA; *fused tells the caller whether that happened (shared-memory fast path only).
int run_planar(pcseg_batch* b, const float* xyz_dev, const float* normals_dev, const int* seeds_dev,
               int n_seeds, int seeds_stride, const int* n_seeds_dev, const pcseg_planar_config* cfg,
               const float q[4], int id_offset, const ClusterArgs* fuse_cluster = nullptr, bool* fused = nullptr) {
  pcseg_ctx* c = b->ctx;
  cudaStream_t stream = S(b);
  if (fused) *fused = false;
  PlanarArgs a;
  a.xyz = xyz_dev; a.normals = normals_dev; a.labels = b->labels.as<int>();
  a.rows = b->rows; a.cols = b->cols;
  a.seeds = seeds_dev; a.n_seeds = n_seeds; a.seeds_stride = seeds_stride; a.n_seeds_dev = n_seeds_dev;
  a.max_plane_distance = cfg->max_plane_distance; a.min_region_area = cfg->min_region_area;
  a.min_inliers = cfg->min_region_inliers; a.period = cfg->plane_model_reestimation_period;
  a.disc_min_range = cfg->discontinuity_min_range; a.disc_max_range = cfg->discontinuity_max_range;
  a.disc_angle = cfg->discontinuity_normal_angle_diff; a.disc_z_diff = cfg->discontinuity_z_diff;
  a.disc_z_ratio = cfg->discontinuity_z_ratio;
  a.qx = q[0]; a.qy = q[1]; a.qz = q[2]; a.qw = q[3];
  a.id_offset = id_offset;
  a.queue = b->queue.as<int>(); a.qcap = b->qcap;
  a.claim = b->claim.as<int>(); a.inliers = b->inl_planar.as<int>();
  a.boundary = b->boundary.as<int>(); a.bflags = b->bflags.as<unsigned char>(); a.bcap = b->bcap;
  a.meta = b->meta_planar.as<PlanarMeta>(); a.rcap = b->rcap_planar;
  a.stats = b->stats.as<FrameStats>();
  a.prof = b->prof.as<long long>();
  a.mask_words = (int)((b->N + 31) / 32);
  const char* mode_env = getenv("PCSEG_PLANAR_MODE");  // diagnostics: "pipe" / "global" force the older paths
  const bool force_global = getenv("PCSEG_FORCE_GLOBAL_PATH") || (mode_env && !strcmp(mode_env, "global"));
  const bool want_pipe = mode_env && !strcmp(mode_env, "pipe");
  cudaFuncAttributes fattr;
  const bool want_wave1 = mode_env && !strcmp(mode_env, "wave1");
  a.nstage = 0;
  // ---- window-parallel fast path with asynchronous re-fits (grow_planar_wave2.cuh)
  if (!force_global && !want_pipe && !want_wave1 && b->N < (size_t)kW2MaxCells) {
    CU_TRY(c, cudaFuncGetAttributes(&fattr, planar_grow_kernel<kWave2Block, 3, false>));
    const size_t st = fattr.sharedSizeBytes;
#ifdef PCSEG_W2_SMALL
    const int nstage = 256;
#else
    const int nstage = 512;
#endif
    const size_t fixed = (size_t)nstage * 40 + kW2Hash * 4 + kW2Rej * 4 + 16;
    const size_t smem_max = 227 * 1024 - st;
    // as many frames per SM as the mask allows (228 KB per SM, 1 KB reserved per CTA): 3, 2 or 1
    int ring = 0;
    for (int per_sm = 3; per_sm >= 1 && !ring; --per_sm) {
      const size_t lim = per_sm == 1 ? smem_max : (228 * 1024) / per_sm - 1024 - st;
      for (int r = (per_sm == 3 ? 256 : 2048); r >= 256 && !ring; r /= 2)
        if ((size_t)a.mask_words * 4 + (size_t)r * 20 + fixed <= lim) ring = r;
    }
    if (ring) {
      const size_t smem = (size_t)a.mask_words * 4 + (size_t)ring * 20 + fixed;
      a.ring = ring;
      a.nstage = nstage;
      CU_TRY(c, cudaFuncSetAttribute(planar_grow_kernel<kWave2Block, 3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      planar_grow_kernel<kWave2Block, 3, false><<<b->n_frames, kWave2Block, smem, stream>>>(a, ClusterArgs{});
      ++c->launches;
      CU_TRY(c, cudaGetLastError());
      return PCSEG_OK;
    }
  }
  // ---- window-parallel path with in-window re-fits (grow_planar_wave.cuh) when mask + ring + tables fit
  if (!force_global && !want_pipe && b->N < (size_t)kWaveMaxCells) {
    CU_TRY(c, cudaFuncGetAttributes(&fattr, planar_grow_kernel<kWaveBlock, 2, false>));
    const size_t st = fattr.sharedSizeBytes;
    const size_t fixed = (size_t)(kWaveBlock - 32) * 40 + kWaveHash * 4 + kWaveRej * 4 + 16;
    const size_t smem_two = (228 * 1024) / 2 - 1024 - st;   // two frames per SM (1 KB reserved per CTA)
    const size_t smem_max = 227 * 1024 - st;
    int ring = 4096;
    while (ring > 512 && (size_t)a.mask_words * 4 + (size_t)ring * 20 + fixed > smem_two) ring /= 2;
    if ((size_t)a.mask_words * 4 + (size_t)ring * 20 + fixed > smem_two) {  // one frame per SM
      ring = 4096;
      while (ring > 512 && (size_t)a.mask_words * 4 + (size_t)ring * 20 + fixed > smem_max) ring /= 2;
    }
    const size_t smem = (size_t)a.mask_words * 4 + (size_t)ring * 20 + fixed;
    if (smem <= smem_max) {
      a.ring = ring;
      CU_TRY(c, cudaFuncSetAttribute(planar_grow_kernel<kWaveBlock, 2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      planar_grow_kernel<kWaveBlock, 2, false><<<b->n_frames, kWaveBlock, smem, stream>>>(a, ClusterArgs{});
      ++c->launches;
      CU_TRY(c, cudaGetLastError());
      return PCSEG_OK;
    }
  }
  // ---- warp-specialised pipeline path when the 1-bit/cell mask and the FIFO ring fit
  a.ring = 2048;
  const size_t smem_fixed = 32 * 11 * sizeof(float) + kHashSize * sizeof(unsigned);
  size_t smem = (size_t)a.mask_words * 4 + (size_t)a.ring * 20 + smem_fixed;
  CU_TRY(c, cudaFuncGetAttributes(&fattr, planar_grow_kernel<kPipeBlock, 1, false>));
  const size_t smem_static = fattr.sharedSizeBytes;
  // two frames per SM when they fit (228 KB per SM, 1 KB reserved per CTA), else one
  const size_t smem_two = (228 * 1024) / 2 - 1024 - smem_static;
  const size_t smem_max = 227 * 1024 - smem_static;
  bool use_smem = b->N < (1u << 23) && (size_t)a.mask_words * 4 + 256 * 20 + smem_fixed <= smem_max && !force_global;
  if (use_smem) {
    if (smem > smem_two && (size_t)a.mask_words * 4 + 1024 * 20 + smem_fixed <= smem_two) a.ring = 1024;
    smem = (size_t)a.mask_words * 4 + (size_t)a.ring * 20 + smem_fixed;
    while (smem > smem_max) { a.ring /= 2; smem = (size_t)a.mask_words * 4 + (size_t)a.ring * 20 + smem_fixed; }
    if (fuse_cluster && getenv("PCSEG_FUSE")) {  // opt-in: measured slower (the extra code thrashes the instruction cache)
      CU_TRY(c, cudaFuncSetAttribute(planar_grow_kernel<kPipeBlock, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      planar_grow_kernel<kPipeBlock, 1, true><<<b->n_frames, kPipeBlock, smem, stream>>>(a, *fuse_cluster);
      if (fused) *fused = true;
    } else {
      CU_TRY(c, cudaFuncSetAttribute(planar_grow_kernel<kPipeBlock, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      planar_grow_kernel<kPipeBlock, 1, false><<<b->n_frames, kPipeBlock, smem, stream>>>(a, ClusterArgs{});
    }
  } else {
    planar_grow_kernel<kGrowBlock, 0, false><<<b->n_frames, kGrowBlock, 0, stream>>>(a, ClusterArgs{});
  }
  ++c->launches;
  CU_TRY(c, cudaGetLastError());
  return PCSEG_OK;
}

int run_cluster(pcseg_batch* b, const float* xyz_dev, const int* seeds_dev, int n_seeds,
                const pcseg_cluster_config* cfg, int id_offset, int offset_from_planar) {
  pcseg_ctx* c = b->ctx;
  cudaStream_t stream = S(b);
  ClusterArgs a;
  int st = make_cluster_args(b, xyz_dev, seeds_dev, n_seeds, cfg, id_offset, offset_from_planar, &a);
  if (st) return st;
  // h == 1 with the frame's label mask in shared memory (grow_cluster_wave.cuh): opt-in
  // (PCSEG_CLUSTER_MODE=wave).  Parity-green, but measured slower than the L2 path on the C2 batch
  // (6.5 vs 6.0 ms; a 256 k-point floor cluster alone 5.1 vs 3.8 ms): its 256-entry windows are half the
  // 512-entry steps of cluster_grow_kernel and the 8-neighbour frontier is wide.
  const char* cmode = getenv("PCSEG_CLUSTER_MODE");
  if (a.half_window == 1 && b->N < (size_t)kCwMaxCells && cmode && !strcmp(cmode, "wave")) {
    cudaFuncAttributes fattr;
    CU_TRY(c, cudaFuncGetAttributes(&fattr, cluster_wave_kernel));
    const int mask_words = (int)((b->N + 31) / 32);
    const size_t fixed = (size_t)kCwHash * 4 + kCwRej * 4;
    const size_t smem_two = (228 * 1024) / 2 - 1024 - fattr.sharedSizeBytes;
    const size_t smem_max = 227 * 1024 - fattr.sharedSizeBytes;
    int ring = 2048;
    while (ring > 256 && (size_t)mask_words * 4 + (size_t)ring * 32 + fixed > smem_two) ring /= 2;
    if ((size_t)mask_words * 4 + (size_t)ring * 32 + fixed > smem_two) {
      ring = 2048;
      while (ring > 256 && (size_t)mask_words * 4 + (size_t)ring * 32 + fixed > smem_max) ring /= 2;
    }
    const size_t smem = (size_t)mask_words * 4 + (size_t)ring * 32 + fixed;
    if (smem <= smem_max) {
      CU_TRY(c, cudaFuncSetAttribute(cluster_wave_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      cluster_wave_kernel<<<b->n_frames, kCwBlock, smem, stream>>>(a, mask_words, ring);
      ++c->launches;
      CU_TRY(c, cudaGetLastError());
      return PCSEG_OK;
    }
  }
  if (a.half_window == 1) cluster_grow_kernel<kClusterBlock, true><<<b->n_frames, kClusterBlock, 0, stream>>>(a);
  else cluster_grow_kernel<kClusterBlock, false><<<b->n_frames, kClusterBlock, 0, stream>>>(a);
  ++c->launches;
  CU_TRY(c, cudaGetLastError());
  return PCSEG_OK;
}

// One large frame with all-descending seeds: bands of columns separated by >= h columns without an
// unlabeled finite cell are independent (grow_cluster.cuh: ClusterBand) and are grown by one CTA each.
// *used = false when the frame has fewer than two bands (the caller then takes the one-CTA path).
int run_cluster_banded(pcseg_batch* b, const float* xyz_dev, const pcseg_cluster_config* cfg, int id_offset, bool* used) {
  pcseg_ctx* c = b->ctx;
  cudaStream_t stream = S(b);
  *used = false;
  const int rows = b->rows, cols = b->cols, h = cfg->half_search_window;
  ClusterArgs a;
  int st = make_cluster_args(b, xyz_dev, nullptr, 0, cfg, id_offset, 0, &a);
  if (st) return st;
  DevBuf d_flag;
  CU_TRY(c, d_flag.ensure((size_t)cols * sizeof(int)));
  cluster_colflag_kernel<<<(unsigned)(((size_t)cols * 32 + 255) / 256), 256, 0, stream>>>(xyz_dev, a.labels, rows, cols, d_flag.as<int>());
  ++c->launches;
  std::vector<int> flag(cols);
  CU_TRY(c, cudaMemcpyAsync(flag.data(), d_flag.p, (size_t)cols * sizeof(int), cudaMemcpyDeviceToHost, stream));
  CU_TRY(c, cudaStreamSynchronize(stream));
  // bands in ascending columns: a run of >= max(h, 1) inactive columns separates two bands
  std::vector<int> cb, ce;
  const int gap = std::max(h, 1);
  int col = 0;
  while (col < cols) {
    while (col < cols && !flag[col]) ++col;
    if (col >= cols) break;
    int begin = col, last_active = col, idle = 0;
    while (col < cols && idle < gap) {
      if (flag[col]) { last_active = col; idle = 0; } else ++idle;
      ++col;
    }
    cb.push_back(begin);
    ce.push_back(last_active + 1);
  }
  const int nb = (int)cb.size();
  if (nb < 2) return PCSEG_OK;
  // tables, bands in ASCENDING column order: the list "all indices descending" is consumed back to front,
  // i.e. in ascending index order (ps/segmentation.h:254-256), so lower columns are reached first
  std::vector<int> h_cb(nb), h_ce(nb), h_qcap(nb), h_moff(nb);
  std::vector<long long> h_qoff(nb);
  long long qtot = 0;
  int mtot = 0;
  for (int k = 0; k < nb; ++k) {
    const int i = k;
    h_cb[k] = cb[i]; h_ce[k] = ce[i];
    const long long cells = (long long)(ce[i] - cb[i]) * rows;
    int qc = 1024;
    while (qc < cells) qc <<= 1;
    h_qcap[k] = qc; h_qoff[k] = qtot; qtot += qc;
    h_moff[k] = mtot; mtot += (int)(cells / std::max(cfg->min_region_inliers, 1)) + 1;
  }
  DevBuf d_tab, d_queue, d_meta, d_out;
  const size_t tab_bytes = (size_t)nb * (4 * sizeof(int) + sizeof(long long));
  CU_TRY(c, d_tab.ensure(tab_bytes));
  CU_TRY(c, d_queue.ensure((size_t)qtot * 2 * sizeof(int)));
  CU_TRY(c, d_meta.ensure((size_t)mtot * sizeof(ClusterMeta)));
  CU_TRY(c, d_out.ensure((size_t)nb * 6 * sizeof(int)));
  char* tp = d_tab.as<char>();
  long long* p_qoff = reinterpret_cast<long long*>(tp);
  int* p_cb = reinterpret_cast<int*>(tp + (size_t)nb * sizeof(long long));
  int* p_ce = p_cb + nb; int* p_qcap = p_ce + nb; int* p_moff = p_qcap + nb;
  CU_TRY(c, cudaMemcpyAsync(p_qoff, h_qoff.data(), (size_t)nb * sizeof(long long), cudaMemcpyHostToDevice, stream));
  CU_TRY(c, cudaMemcpyAsync(p_cb, h_cb.data(), (size_t)nb * sizeof(int), cudaMemcpyHostToDevice, stream));
  CU_TRY(c, cudaMemcpyAsync(p_ce, h_ce.data(), (size_t)nb * sizeof(int), cudaMemcpyHostToDevice, stream));
  CU_TRY(c, cudaMemcpyAsync(p_qcap, h_qcap.data(), (size_t)nb * sizeof(int), cudaMemcpyHostToDevice, stream));
  CU_TRY(c, cudaMemcpyAsync(p_moff, h_moff.data(), (size_t)nb * sizeof(int), cudaMemcpyHostToDevice, stream));
  ClusterBandTable t;
  t.c_begin = p_cb; t.c_end = p_ce; t.q_off = p_qoff; t.q_cap = p_qcap; t.m_off = p_moff;
  t.queue = d_queue.as<int>(); t.meta = d_meta.as<ClusterMeta>(); t.out = d_out.as<int>(); t.rcap_total = mtot;
  if (a.half_window == 1) cluster_band_kernel<kClusterBlock, true><<<nb, kClusterBlock, 0, stream>>>(a, t);
  else cluster_band_kernel<kClusterBlock, false><<<nb, kClusterBlock, 0, stream>>>(a, t);
  int* region_base = d_out.as<int>() + 4 * nb;
  cluster_band_merge_kernel<<<1, 256, 0, stream>>>(a, t, nb, region_base);
  // the band queues are idle now: scratch for packing the inlier lists in region order
  cluster_band_compact_kernel<<<nb, 256, 0, stream>>>(a, t, nb, region_base, d_queue.as<int>());
  cluster_band_copyback_kernel<<<296, 256, 0, stream>>>(a, d_queue.as<int>());
  cluster_band_relabel_kernel<<<296, 256, 0, stream>>>(a);
  c->launches += 5;
  CU_TRY(c, cudaGetLastError());
  CU_TRY(c, cudaStreamSynchronize(stream));   // the tables above die with this scope
  *used = true;
  return PCSEG_OK;
}

// One large frame, one CTA per connected component (grow_cluster_comp.cuh).  *used = false: not
// applicable (fewer than two components that can yield a region) -- the caller takes another path.
int run_cluster_components(pcseg_batch* b, const float* xyz_dev, const pcseg_cluster_config* cfg, int id_offset, bool* used) {
  pcseg_ctx* c = b->ctx;
  cudaStream_t stream = S(b);
  *used = false;
  const int rows = b->rows, cols = b->cols, h = cfg->half_search_window;
  const size_t N = b->N;
  if (N >= ((size_t)1 << 30)) return PCSEG_OK;
  ClusterArgs a;
  int st = make_cluster_args(b, xyz_dev, nullptr, 0, cfg, id_offset, 0, &a);
  if (st) return st;
  const int min_inl = std::max(cfg->min_region_inliers, 1);
  DevBuf d_parent, d_count, d_hi, d_list, d_nlist;
  {   // the component codes live below kCompUnlabeledBase + 2^30: a caller grid with such values takes another path
    CU_TRY(c, d_nlist.ensure(sizeof(int)));
    CU_TRY(c, cudaMemsetAsync(d_nlist.p, 0, sizeof(int), stream));
    label_range_kernel<<<148 * 8, 256, 0, stream>>>(a.labels, N, kCompUnlabeledBase + (1 << 30), INT_MAX, d_nlist.as<int>());
    int bad = 0;
    CU_TRY(c, cudaMemcpyAsync(&bad, d_nlist.p, sizeof(int), cudaMemcpyDeviceToHost, stream));
    CU_TRY(c, cudaStreamSynchronize(stream));
    ++c->launches;
    if (bad) return PCSEG_OK;
  }
  const int list_cap = (int)std::min<size_t>(N / (size_t)min_inl + 1, (size_t)1 << 22);
  CU_TRY(c, d_parent.ensure(N * sizeof(int)));
  CU_TRY(c, d_count.ensure(N * sizeof(int)));
  CU_TRY(c, d_hi.ensure(N * sizeof(int)));
  CU_TRY(c, d_list.ensure((size_t)list_cap * 3 * sizeof(int)));
  CU_TRY(c, d_nlist.ensure(sizeof(int)));
  CU_TRY(c, cudaMemsetAsync(d_nlist.p, 0, sizeof(int), stream));
  const unsigned grid = (unsigned)((N + 255) / 256);
  cluster_comp_init_kernel<<<grid, 256, 0, stream>>>(xyz_dev, a.labels, (int)N, d_parent.as<int>(), d_count.as<int>(), d_hi.as<int>());
  cluster_comp_link_kernel<<<grid, 256, 0, stream>>>(xyz_dev, rows, cols, h, a.sq_thr, d_parent.as<int>());
  cluster_comp_flatten_kernel<<<grid, 256, 0, stream>>>((int)N, d_parent.as<int>(), a.labels, d_count.as<int>(), d_hi.as<int>());
  cluster_comp_list_kernel<<<grid, 256, 0, stream>>>((int)N, d_parent.as<int>(), d_count.as<int>(), d_hi.as<int>(), min_inl,
                                                     d_list.as<int>(), list_cap, d_nlist.as<int>());
  c->launches += 4;
  const unsigned rgrid = (unsigned)std::min<size_t>((N + 255) / 256, (size_t)148 * 16);
  int nc = 0;
  CU_TRY(c, cudaMemcpyAsync(&nc, d_nlist.p, sizeof(int), cudaMemcpyDeviceToHost, stream));
  CU_TRY(c, cudaStreamSynchronize(stream));
  if (nc > list_cap) nc = list_cap;   // cannot happen: a listed component has >= min_inl cells
  struct Comp { int root, count, hi; };
  std::vector<Comp> comps((size_t)nc);
  if (nc > 0) CU_TRY(c, cudaMemcpy(comps.data(), d_list.p, (size_t)nc * 3 * sizeof(int), cudaMemcpyDeviceToHost));
  std::sort(comps.begin(), comps.end(), [](const Comp& x, const Comp& y) { return x.root < y.root; });
  // tables
  std::vector<int> h_root(nc), h_hi(nc), h_qcap(nc), h_ioff(nc), h_moff(nc);
  std::vector<long long> h_qoff(nc);
  long long qtot = 0;
  int mtot = 0, itot = 0;
  for (int k = 0; k < nc; ++k) {
    h_root[k] = comps[k].root; h_hi[k] = comps[k].hi;
    int qc = 1024;
    while (qc < comps[k].count) qc <<= 1;
    h_qcap[k] = qc; h_qoff[k] = qtot; qtot += qc;
    h_ioff[k] = itot; itot += comps[k].count;
    h_moff[k] = mtot; mtot += comps[k].count / min_inl + 1;
  }
  int n_regions = 0, n_inliers = 0, err = 0, steps = 0;
  if (nc > 0) {
    DevBuf d_tab, d_queue, d_meta, d_out, d_mv;
    const size_t tab_bytes = (size_t)nc * (5 * sizeof(int) + sizeof(long long));
    CU_TRY(c, d_tab.ensure(tab_bytes));
    CU_TRY(c, d_queue.ensure(std::max((size_t)qtot * 2, N) * sizeof(int)));   // later: scratch for the packed inlier lists
    CU_TRY(c, d_meta.ensure((size_t)mtot * sizeof(ClusterMeta)));
    CU_TRY(c, d_out.ensure((size_t)nc * 4 * sizeof(int)));
    char* tp = d_tab.as<char>();
    long long* p_qoff = reinterpret_cast<long long*>(tp);
    int* p_root = reinterpret_cast<int*>(tp + (size_t)nc * sizeof(long long));
    int* p_hi = p_root + nc; int* p_qcap = p_hi + nc; int* p_ioff = p_qcap + nc; int* p_moff = p_ioff + nc;
    CU_TRY(c, cudaMemcpyAsync(p_qoff, h_qoff.data(), (size_t)nc * sizeof(long long), cudaMemcpyHostToDevice, stream));
    CU_TRY(c, cudaMemcpyAsync(p_root, h_root.data(), (size_t)nc * sizeof(int), cudaMemcpyHostToDevice, stream));
    CU_TRY(c, cudaMemcpyAsync(p_hi, h_hi.data(), (size_t)nc * sizeof(int), cudaMemcpyHostToDevice, stream));
    CU_TRY(c, cudaMemcpyAsync(p_qcap, h_qcap.data(), (size_t)nc * sizeof(int), cudaMemcpyHostToDevice, stream));
    CU_TRY(c, cudaMemcpyAsync(p_ioff, h_ioff.data(), (size_t)nc * sizeof(int), cudaMemcpyHostToDevice, stream));
    CU_TRY(c, cudaMemcpyAsync(p_moff, h_moff.data(), (size_t)nc * sizeof(int), cudaMemcpyHostToDevice, stream));
    ClusterCompTable t;
    t.root = p_root; t.hi = p_hi; t.q_off = p_qoff; t.q_cap = p_qcap; t.inl_off = p_ioff; t.m_off = p_moff;
    t.queue = d_queue.as<int>(); t.meta = d_meta.as<ClusterMeta>(); t.out = d_out.as<int>(); t.rcap_total = mtot;
    t.comp = d_parent.as<int>();
    if (a.half_window == 1) cluster_comp_kernel<kClusterBlock, true><<<nc, kClusterBlock, 0, stream>>>(a, t);
    else cluster_comp_kernel<kClusterBlock, false><<<nc, kClusterBlock, 0, stream>>>(a, t);
    ++c->launches;
    CU_TRY(c, cudaGetLastError());
    // the regions of all components in the reference's order: ascending seed index
    std::vector<int> h_out((size_t)nc * 4);
    std::vector<ClusterMeta> h_meta((size_t)mtot);
    CU_TRY(c, cudaMemcpyAsync(h_out.data(), d_out.p, h_out.size() * sizeof(int), cudaMemcpyDeviceToHost, stream));
    CU_TRY(c, cudaMemcpyAsync(h_meta.data(), d_meta.p, h_meta.size() * sizeof(ClusterMeta), cudaMemcpyDeviceToHost, stream));
    CU_TRY(c, cudaStreamSynchronize(stream));
    struct Reg { int seed, src, count; };
    std::vector<Reg> regs;
    for (int k = 0; k < nc; ++k) {
      err |= h_out[4 * k + 2]; steps += h_out[4 * k + 3];
      for (int r = 0; r < h_out[4 * k]; ++r) {
        const ClusterMeta& m = h_meta[(size_t)h_moff[k] + r];
        regs.push_back(Reg{m.label_id, m.inlier_begin, m.inlier_count});   // label_id carries the seed in component mode
      }
    }
    std::sort(regs.begin(), regs.end(), [](const Reg& x, const Reg& y) { return x.seed < y.seed; });
    if ((int)regs.size() > a.rcap) { err |= kErrRegionCap; regs.resize((size_t)a.rcap); }
    n_regions = (int)regs.size();
    std::vector<ClusterCompMove> mv((size_t)n_regions);
    for (int r = 0; r < n_regions; ++r) {
      mv[r] = ClusterCompMove{regs[r].src, regs[r].count, n_inliers, id_offset + r};
      n_inliers += regs[r].count;
    }
    if (n_regions > 0) {
      CU_TRY(c, d_mv.ensure((size_t)n_regions * sizeof(ClusterCompMove)));
      CU_TRY(c, cudaMemcpyAsync(d_mv.p, mv.data(), (size_t)n_regions * sizeof(ClusterCompMove), cudaMemcpyHostToDevice, stream));
      cluster_comp_pack_kernel<<<(unsigned)std::min(n_regions, 148 * 8), 256, 0, stream>>>(a, d_mv.as<ClusterCompMove>(), n_regions, d_queue.as<int>());
      ++c->launches;
    }
    cluster_comp_stats_kernel<<<1, 1, 0, stream>>>(a, n_regions, n_inliers, err, steps);
    if (n_regions > 0) cluster_band_copyback_kernel<<<296, 256, 0, stream>>>(a, d_queue.as<int>());
    cluster_comp_restore_kernel<<<rgrid, 256, 0, stream>>>(a.labels, (int)N);
    c->launches += 3;
    CU_TRY(c, cudaGetLastError());
    CU_TRY(c, cudaStreamSynchronize(stream));   // the tables above die with this scope
  } else {
    cluster_comp_stats_kernel<<<1, 1, 0, stream>>>(a, 0, 0, 0, 0);
    cluster_comp_restore_kernel<<<rgrid, 256, 0, stream>>>(a.labels, (int)N);
    c->launches += 2;
    CU_TRY(c, cudaStreamSynchronize(stream));
  }
  *used = true;
  return PCSEG_OK;
}

int map_error_flags(pcseg_ctx* c, int flags) {
  if (flags == 0) return PCSEG_OK;
  c->last_error = "per-frame capacity/trace error flags = " + std::to_string(flags);
  return PCSEG_ERR_CAPACITY;
}

}  // namespace

// ===========================================================================
// Region materialisation
// ===========================================================================
namespace {

// Packs the region results of frames [f0, f0 + nf) on the device (results_pack.cuh) and queues the
// copy of the frame offsets / stats / pack error word into the workspace's pinned staging block.
// Enqueued on the workspace's main stream behind the growth kernels.
struct PackWant {
  bool inlier_xyz = false, boundary_xyz = false, hull = false, cluster_stats = false, cluster = false;
};

int run_pack(pcseg_batch* b, int f0, int nf, const PackWant& w) {
  pcseg_ctx* c = b->ctx;
  cudaStream_t stream = S(b);
  const size_t N = b->N, F = (size_t)nf;
  const size_t ecap = F * (size_t)b->bcap;
  CU_TRY(c, b->pk_offsets.ensure(5 * ((size_t)b->n_frames + 1) * sizeof(int)));
  CU_TRY(c, b->pk_planar.ensure(F * (size_t)b->rcap_planar * sizeof(pcseg_planar_record)));
  CU_TRY(c, b->pk_pinl.ensure(F * N * sizeof(int)));
  if (w.inlier_xyz) CU_TRY(c, b->pk_pinl_xyz.ensure(F * N * 3 * sizeof(float)));
  CU_TRY(c, b->pk_bnd.ensure(ecap * sizeof(int)));
  if (w.boundary_xyz) CU_TRY(c, b->pk_bnd_xyz.ensure(ecap * 3 * sizeof(float)));
  CU_TRY(c, b->pk_bdisc.ensure(ecap));
  CU_TRY(c, b->pk_disc_idx.ensure(ecap * sizeof(int)));
  CU_TRY(c, b->pk_disc_pos.ensure(ecap * sizeof(int)));
  if (w.hull) CU_TRY(c, b->pk_hull.ensure(ecap * 3 * sizeof(float)));
  if (w.cluster) {
    CU_TRY(c, b->pk_cluster.ensure(F * (size_t)b->rcap_cluster * sizeof(pcseg_cluster_record)));
    CU_TRY(c, b->pk_cinl.ensure(F * N * sizeof(int)));
    if (w.inlier_xyz) CU_TRY(c, b->pk_cinl_xyz.ensure(F * N * 3 * sizeof(float)));
  }
  CU_TRY(c, b->pk_err.ensure(sizeof(int)));
  if (!b->h_meta) CU_TRY(c, cudaMallocHost((void**)&b->h_meta, (5 * ((size_t)b->n_frames + 1) + 8 * (size_t)b->n_frames + 1) * sizeof(int)));
  PackArgs a;
  a.F = nf; a.N = (int)N; a.rows = b->rows;
  a.xyz = b->xyz_dev + (size_t)f0 * N * 3;
  a.stats = b->stats.as<FrameStats>() + f0;
  a.meta_planar = b->meta_planar.as<PlanarMeta>() + (size_t)f0 * b->rcap_planar; a.rcap_planar = b->rcap_planar;
  a.meta_cluster = w.cluster ? b->meta_cluster.as<ClusterMeta>() + (size_t)f0 * b->rcap_cluster : nullptr;
  a.rcap_cluster = b->rcap_cluster;
  a.inl_planar = b->inl_planar.as<int>() + (size_t)f0 * N;
  a.inl_cluster = b->inl_cluster.as<int>() + (size_t)f0 * N;
  a.boundary = b->boundary.as<int>() + (size_t)f0 * b->bcap;
  a.bflags = b->bflags.as<unsigned char>() + (size_t)f0 * b->bcap; a.bcap = b->bcap;
  a.scratch = b->queue.as<int>() + (size_t)f0 * b->qcap * 2; a.scratch_stride = (size_t)b->qcap * 2;
  a.offsets = b->pk_offsets.as<int>();
  a.planar = b->pk_planar.as<pcseg_planar_record>();
  a.p_inl = b->pk_pinl.as<int>(); a.p_inl_xyz = w.inlier_xyz ? b->pk_pinl_xyz.as<float>() : nullptr;
  a.bnd = b->pk_bnd.as<int>(); a.bnd_xyz = w.boundary_xyz ? b->pk_bnd_xyz.as<float>() : nullptr;
  a.bnd_disc = b->pk_bdisc.as<unsigned char>();
  a.disc_idx = b->pk_disc_idx.as<int>(); a.disc_pos = b->pk_disc_pos.as<int>();
  a.hull = w.hull ? b->pk_hull.as<float>() : nullptr;
  a.cluster = b->pk_cluster.as<pcseg_cluster_record>();
  a.c_inl = b->pk_cinl.as<int>(); a.c_inl_xyz = (w.cluster && w.inlier_xyz) ? b->pk_cinl_xyz.as<float>() : nullptr;
  a.want_cluster_stats = w.cluster_stats ? 1 : 0;
  a.err = b->pk_err.as<int>();
  CU_TRY(c, cudaMemsetAsync(a.err, 0, sizeof(int), stream));
  pack_offsets_kernel<<<1, 1024, 0, stream>>>(a);
  const int slices = std::max(1, std::min(64, (148 * 8 + nf - 1) / nf));
  pack_bulk_kernel<<<dim3(slices, nf), 256, 0, stream>>>(a);
  CU_TRY(c, cudaFuncSetAttribute(pack_region_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPackRegionSmem));
  pack_region_kernel<<<nf, kPackBlock, kPackRegionSmem, stream>>>(a);
  c->launches += 3;
  CU_TRY(c, cudaGetLastError());
  int* h = b->h_meta;
  CU_TRY(c, cudaMemcpyAsync(h, a.offsets, 5 * (F + 1) * sizeof(int), cudaMemcpyDeviceToHost, stream));
  CU_TRY(c, cudaMemcpyAsync(h + 5 * (F + 1), a.stats, F * sizeof(FrameStats), cudaMemcpyDeviceToHost, stream));
  CU_TRY(c, cudaMemcpyAsync(h + 5 * (F + 1) + 8 * F, a.err, sizeof(int), cudaMemcpyDeviceToHost, stream));
  return PCSEG_OK;
}

// Host view of the staging block after the stream has been synchronised.
struct PackTotals {
  const int* off[5];
  const int* stats;
  int n_planar, n_pinl, n_bnd, n_cluster, n_cinl, err;
};
PackTotals pack_totals(const pcseg_batch* b, int nf) {
  PackTotals t;
  const int F1 = nf + 1;
  for (int k = 0; k < 5; ++k) t.off[k] = b->h_meta + k * F1;
  t.stats = b->h_meta + 5 * F1;
  t.n_planar = t.off[0][nf]; t.n_pinl = t.off[1][nf]; t.n_bnd = t.off[2][nf];
  t.n_cluster = t.off[3][nf]; t.n_cinl = t.off[4][nf];
  t.err = b->h_meta[5 * F1 + 8 * nf];
  for (int f = 0; f < nf; ++f) t.err |= t.stats[8 * f + 5];
  return t;
}

// dst (host or device, nullable) <- count elements of src_dev; false when the capacity is too small.
template <typename T>
bool copy_result(pcseg_batch* b, T* dst, long long cap, const T* src_dev, long long count, cudaError_t* e, int mult = 1) {
  if (!dst || count <= 0) return true;
  if (count > cap) return false;
  cudaMemcpyKind k = is_device_ptr(dst) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
  cudaError_t r = cudaMemcpyAsync(dst, src_dev, (size_t)count * mult * sizeof(T), k, S(b));
  if (r != cudaSuccess && *e == cudaSuccess) *e = r;
  return true;
}

// One frame's regions through the packed path -> the CSR vectors of pcseg_regions.
int collect_planar(pcseg_batch* b, int frame, pcseg_regions* r) {
  pcseg_ctx* c = b->ctx;
  cudaStream_t stream = S(b);
  PackWant w;
  w.hull = true;
  int st = run_pack(b, frame, 1, w);
  if (st) return st;
  CU_TRY(c, cudaStreamSynchronize(stream));
  const PackTotals t = pack_totals(b, 1);
  const int n = t.n_planar;
  std::vector<pcseg_planar_record> rec(n);
  std::vector<int> inl(t.n_pinl), bnd(t.n_bnd), didx(t.n_bnd);
  std::vector<float> hull(3 * (size_t)t.n_bnd);
  if (n) CU_TRY(c, cudaMemcpyAsync(rec.data(), b->pk_planar.p, n * sizeof(pcseg_planar_record), cudaMemcpyDeviceToHost, stream));
  if (t.n_pinl) CU_TRY(c, cudaMemcpyAsync(inl.data(), b->pk_pinl.p, inl.size() * sizeof(int), cudaMemcpyDeviceToHost, stream));
  if (t.n_bnd) {
    CU_TRY(c, cudaMemcpyAsync(bnd.data(), b->pk_bnd.p, bnd.size() * sizeof(int), cudaMemcpyDeviceToHost, stream));
    CU_TRY(c, cudaMemcpyAsync(didx.data(), b->pk_disc_idx.p, didx.size() * sizeof(int), cudaMemcpyDeviceToHost, stream));
    CU_TRY(c, cudaMemcpyAsync(hull.data(), b->pk_hull.p, hull.size() * sizeof(float), cudaMemcpyDeviceToHost, stream));
  }
  CU_TRY(c, cudaStreamSynchronize(stream));
  r->inlier_indices = std::move(inl);
  r->boundary_indices = std::move(bnd);
  for (int i = 0; i < n; ++i) {
    const pcseg_planar_record& m = rec[i];
    r->inlier_offsets.push_back(m.inlier_begin + m.inlier_count);
    r->label_ids.push_back(m.label_id);
    r->seed_indices.push_back(m.seed_index);
    r->planes.insert(r->planes.end(), m.plane, m.plane + 4);
    r->centroids.insert(r->centroids.end(), m.centroid, m.centroid + 3);
    r->areas.push_back(m.area);
    r->estimators.insert(r->estimators.end(), m.estimator, m.estimator + 10);
    r->boundary_offsets.push_back(m.boundary_begin + m.boundary_count);
    r->disc_indices.insert(r->disc_indices.end(), didx.begin() + m.discontinuous_begin,
                           didx.begin() + m.discontinuous_begin + m.discontinuous_count);
    r->disc_offsets.push_back((int)r->disc_indices.size());
    r->hull_points.insert(r->hull_points.end(), hull.begin() + 3 * (size_t)m.hull_begin,
                          hull.begin() + 3 * (size_t)(m.hull_begin + m.hull_count));
    r->hull_offsets.push_back((int)(r->hull_points.size() / 3));
  }
  return map_error_flags(c, t.err);
}

int collect_cluster(pcseg_batch* b, int frame, pcseg_regions* r) {
  pcseg_ctx* c = b->ctx;
  cudaStream_t stream = S(b);
  PackWant w;
  w.cluster = true;
  int st = run_pack(b, frame, 1, w);
  if (st) return st;
  CU_TRY(c, cudaStreamSynchronize(stream));
  const PackTotals t = pack_totals(b, 1);
  const int n = t.n_cluster;
  std::vector<pcseg_cluster_record> rec(n);
  std::vector<int> inl(t.n_cinl);
  if (n) CU_TRY(c, cudaMemcpyAsync(rec.data(), b->pk_cluster.p, n * sizeof(pcseg_cluster_record), cudaMemcpyDeviceToHost, stream));
  if (t.n_cinl) CU_TRY(c, cudaMemcpyAsync(inl.data(), b->pk_cinl.p, inl.size() * sizeof(int), cudaMemcpyDeviceToHost, stream));
  CU_TRY(c, cudaStreamSynchronize(stream));
  r->inlier_indices = std::move(inl);
  for (int i = 0; i < n; ++i) {
    r->inlier_offsets.push_back(rec[i].inlier_begin + rec[i].inlier_count);
    r->label_ids.push_back(rec[i].label_id);
  }
  return map_error_flags(c, t.err);
}

int get_scratch(pcseg_ctx* c, int rows, int cols, pcseg_batch** out) {
  if (c->scratch && (c->scratch->rows != rows || c->scratch->cols != cols)) {
    pcseg_batch_destroy(c->scratch);
    c->scratch = nullptr;
  }
  if (!c->scratch) {
    int st = pcseg_batch_create(c, 1, rows, cols, &c->scratch);
    if (st) return st;
    // the single-frame entry points issue their copies on the context's stream: the cached
    // workspace must launch there too (S / SIn / SOut fall back to the context's streams)
    pcseg_batch* b = c->scratch;
    for (cudaStream_t* s : {&b->own_stream, &b->own_copy_in, &b->own_copy_out, &b->own_normals})
      if (*s) { cudaStreamSynchronize(*s); cudaStreamDestroy(*s); *s = nullptr; }
  }
  *out = c->scratch;
  return PCSEG_OK;
}

int check_seeds_host(const int* seeds_host_or_dev, int n, size_t N, pcseg_ctx* c, DevBuf* staging, const int** dev,
                     cudaStream_t stream = nullptr) {
  // Validates the range on the host copy (the reference would read out of
  // bounds); device-resident seeds are trusted.
  if (n == 0) { *dev = (const int*)staging->p; return PCSEG_OK; }
  if (!is_device_ptr(seeds_host_or_dev)) {
    for (int i = 0; i < n; ++i)
      if (seeds_host_or_dev[i] < 0 || (size_t)seeds_host_or_dev[i] >= N) return PCSEG_ERR_SEED_RANGE;
  }
  const void* d;
  int st = to_device(c, seeds_host_or_dev, (size_t)n * sizeof(int), staging, &d, stream);
  *dev = (const int*)d;
  return st;
}

}  // namespace

// ===========================================================================
// extern "C" entry points
// ===========================================================================
extern "C" {

void pcseg_default_normals_params(pcseg_normals_params* p) {
  p->min_neighbor_distance = 0.1f;  // ps/algorithms.h:313-322 (float <- double literal)
  p->max_neighbor_distance = 1.0f;
  p->include_diagonal_neighbors = 1;
  p->min_num_support_neighbors = 4;
}
void pcseg_default_planar_config(pcseg_planar_config* c) {
  c->max_plane_distance = 0.05f;  // ps/planar_region.h:93-121
  c->min_region_area = 0.05f;
  c->min_region_inliers = 5;
  c->plane_model_reestimation_period = 30;
  c->discontinuity_min_range = 1.2f;
  c->discontinuity_max_range = 4.0f;
  c->discontinuity_normal_angle_diff = 5.0f;
  c->discontinuity_z_diff = 0.05f;
  c->discontinuity_z_ratio = 0.7f;
}
void pcseg_default_cluster_config(pcseg_cluster_config* c) {
  c->min_region_inliers = 7;  // ps/cluster_region.h:53-63
  c->squared_distance_threshold = 1.0f;
  c->half_search_window = 1;
}

int pcseg_create(int device, pcseg_ctx** out) {
  if (!out) return PCSEG_ERR_NULL;
  *out = nullptr;
  // Workspaces overlap through their own stream sets; with the default 8 hardware queues
  // independent streams share a queue and serialise (4 workspaces x 3 streams: 46 -> 35 ms per
  // C2 step).  Only effective if no CUDA context exists yet; an explicit setting wins.
  setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) {
    cudaGetLastError();
    return PCSEG_ERR_CUDA;  // no CPU fallback
  }
  if (cudaSetDevice(device) != cudaSuccess) return PCSEG_ERR_CUDA;
  pcseg_ctx* c = new pcseg_ctx();
  c->device = device;
  if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&c->copy_in, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&c->copy_out, cudaStreamNonBlocking) != cudaSuccess) {
    delete c;
    return PCSEG_ERR_CUDA;
  }
  g_pool.context_created(device);
  *out = c;
  return PCSEG_OK;
}

void pcseg_destroy(pcseg_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->scratch) pcseg_batch_destroy(c->scratch);
  c->tmp_a.release();
  c->tmp_b.release();
  if (c->copy_in) cudaStreamDestroy(c->copy_in);
  if (c->copy_out) cudaStreamDestroy(c->copy_out);
  if (c->stream) cudaStreamDestroy(c->stream);
  g_pool.context_destroyed(c->device);
  delete c;
}

const char* pcseg_status_string(int s) {
  switch (s) {
    case PCSEG_OK: return "ok";
    case PCSEG_ERR_NULL: return "null pointer (reference: CHECK(ptr))";
    case PCSEG_ERR_DIMS: return "dimension mismatch or rows <= 1 (reference: CHECK_EQ / CHECK_GT(rows, 1))";
    case PCSEG_ERR_CONFIG: return "config value that is undefined behaviour in the reference";
    case PCSEG_ERR_CUDA: return "CUDA error";
    case PCSEG_ERR_CAPACITY: return "internal per-frame buffer overflow";
    case PCSEG_ERR_UNSUPPORTED: return "unsupported parameter (documented limit)";
    case PCSEG_ERR_SEED_RANGE: return "seed index out of range";
  }
  return "unknown";
}
const char* pcseg_last_error(const pcseg_ctx* c) { return c ? c->last_error.c_str() : ""; }
long long pcseg_launch_count(const pcseg_ctx* c) { return c ? c->launches : 0; }
void* pcseg_stream(const pcseg_ctx* c) { return c ? (void*)c->stream : nullptr; }

int pcseg_prepare_unlabeled_labels(pcseg_ctx* c, int* labels, int n) {
  if (!c || !labels) return PCSEG_ERR_NULL;
  if (n <= 0) return PCSEG_OK;
  cudaSetDevice(c->device);
  if (is_device_ptr(labels)) {
    int st = fill_int(c, labels, (size_t)n, kUnlabeled);
    if (st) return st;
  } else {
    CU_TRY(c, c->tmp_a.ensure((size_t)n * sizeof(int)));
    int st = fill_int(c, c->tmp_a.as<int>(), (size_t)n, kUnlabeled);
    if (st) return st;
    CU_TRY(c, cudaMemcpyAsync(labels, c->tmp_a.p, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  }
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  return PCSEG_OK;
}

int pcseg_normals_organized(pcseg_ctx* c, const float* xyz, int rows, int cols, const double sensor_t[3],
                            const pcseg_normals_params* params, int row_begin, int row_end, int col_begin,
                            int col_end, float* normals, float* curvature) {
  if (!c || !xyz || !normals || !params || !sensor_t) return PCSEG_ERR_NULL;
  if (rows <= 0 || cols <= 0) return PCSEG_ERR_DIMS;
  if (row_begin < 0 || col_begin < 0 || row_end > rows || col_end > cols) return PCSEG_ERR_DIMS;
  cudaSetDevice(c->device);
  pcseg_batch* b;
  int st = get_scratch(c, rows, cols, &b);
  if (st) return st;
  const size_t N = b->N;
  const void* xyz_dev;
  if ((st = to_device(c, xyz, N * 3 * sizeof(float), &b->xyz, &xyz_dev))) return st;
  // normals are in/out: cells outside the range keep the caller's values
  float* nrm_dev;
  const bool n_dev = is_device_ptr(normals);
  const bool full = row_begin == 0 && col_begin == 0 && row_end == rows && col_end == cols;
  if (n_dev) nrm_dev = normals;
  else {
    CU_TRY(c, b->normals.ensure(N * 3 * sizeof(float)));
    nrm_dev = b->normals.as<float>();
    if (!full) CU_TRY(c, cudaMemcpyAsync(nrm_dev, normals, N * 3 * sizeof(float), cudaMemcpyHostToDevice, c->stream));
  }
  float* curv_dev = nullptr;
  const bool cv_dev = curvature && is_device_ptr(curvature);
  if (curvature) {
    if (cv_dev) curv_dev = curvature;
    else {
      CU_TRY(c, c->tmp_b.ensure(N * sizeof(float)));
      curv_dev = c->tmp_b.as<float>();
      if (!full) CU_TRY(c, cudaMemcpyAsync(curv_dev, curvature, N * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    }
  }
  if ((st = run_normals(b, (const float*)xyz_dev, nrm_dev, curv_dev, params, sensor_t, row_begin, row_end, col_begin, col_end))) return st;
  if (!n_dev && (st = from_device(c, normals, nrm_dev, N * 3 * sizeof(float)))) return st;
  if (curvature && !cv_dev && (st = from_device(c, curvature, curv_dev, N * sizeof(float)))) return st;
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  return PCSEG_OK;
}

int pcseg_segment_planar(pcseg_ctx* c, const float* xyz, const float* normals, int rows, int cols,
                         const float robot_q_xyzw[4], const int* seeds, int n_seeds,
                         const pcseg_planar_config* cfg, int id_offset, int* labels, pcseg_regions** out) {
  if (!c || !xyz || !normals || !cfg || !labels || !out || !robot_q_xyzw) return PCSEG_ERR_NULL;  // CHECK(regions), CHECK(labels)
  if (n_seeds > 0 && !seeds) return PCSEG_ERR_NULL;
  if (rows <= 1 || cols <= 0) return PCSEG_ERR_DIMS;  // CHECK_GT(points.Rows(), 1), ps/planar_region.h:264
  int st = check_planar_cfg(cfg);
  if (st) return st;
  cudaSetDevice(c->device);
  pcseg_batch* b;
  if ((st = get_scratch(c, rows, cols, &b))) return st;
  const size_t N = b->N;
  const void *xyz_dev, *nrm_dev;
  if ((st = to_device(c, xyz, N * 3 * sizeof(float), &b->xyz, &xyz_dev))) return st;
  if ((st = to_device(c, normals, N * 3 * sizeof(float), &b->normals, &nrm_dev))) return st;
  b->xyz_dev = (const float*)xyz_dev;
  b->normals_dev = (const float*)nrm_dev;
  pcseg_regions* r = nullptr;
  for (int attempt = 0;; ++attempt) {
    CU_TRY(c, cudaMemcpyAsync(b->labels.p, labels, N * sizeof(int),
                              is_device_ptr(labels) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
    CU_TRY(c, cudaMemsetAsync(b->stats.p, 0, sizeof(FrameStats), c->stream));
    if (n_seeds > 0) {
      const int* seeds_dev;
      if ((st = check_seeds_host(seeds, n_seeds, N, c, &b->seeds, &seeds_dev))) return st;
      if ((st = run_planar(b, b->xyz_dev, b->normals_dev, seeds_dev, n_seeds, 0, nullptr, cfg, robot_q_xyzw, id_offset))) return st;
    }
    r = new pcseg_regions();
    st = collect_planar(b, 0, r);
    if (st == PCSEG_OK) break;
    delete r;
    // more planes than the region table holds (the reference has no limit): grow it and repeat
    const int bound = planar_region_bound(b, cfg);
    if (st != PCSEG_ERR_CAPACITY || attempt > 0 || b->rcap_planar >= bound) return st;
    b->rcap_planar = bound;
    CU_TRY(c, b->meta_planar.ensure((size_t)b->rcap_planar * sizeof(PlanarMeta)));
  }
  if ((st = from_device(c, labels, b->labels.p, N * sizeof(int)))) { delete r; return st; }
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  *out = r;
  return PCSEG_OK;
}

int pcseg_segment_cluster(pcseg_ctx* c, const float* xyz, int rows, int cols, const int* seeds, int n_seeds,
                          const pcseg_cluster_config* cfg, int id_offset, int* labels, pcseg_regions** out) {
  if (!c || !xyz || !cfg || !labels || !out) return PCSEG_ERR_NULL;
  if (n_seeds > 0 && !seeds) return PCSEG_ERR_NULL;
  if (rows <= 1 || cols <= 0) return PCSEG_ERR_DIMS;  // CHECK_GT(points.Rows(), 1), ps/cluster_region.h:154
  int st = check_cluster_cfg(cfg);
  if (st) return st;
  cudaSetDevice(c->device);
  pcseg_batch* b;
  if ((st = get_scratch(c, rows, cols, &b))) return st;
  const size_t N = b->N;
  const void* xyz_dev;
  if ((st = to_device(c, xyz, N * 3 * sizeof(float), &b->xyz, &xyz_dev))) return st;
  b->xyz_dev = (const float*)xyz_dev;
  CU_TRY(c, cudaMemcpyAsync(b->labels.p, labels, N * sizeof(int),
                            is_device_ptr(labels) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
  CU_TRY(c, cudaMemsetAsync(b->stats.p, 0, sizeof(FrameStats), c->stream));
  if (n_seeds > 0) {
    const int* seeds_dev;
    // one large frame with the seed list "all indices descending" (ps/clustering_test.cc:57-59): independent
    // parts are grown concurrently -- connected components, else column bands (PCSEG_CLUSTER_BANDS=0 turns
    // both off)
    bool banded = false;
    const char* be = getenv("PCSEG_CLUSTER_BANDS");
    if (!(be && be[0] == '0') && N >= ((size_t)1 << 18) && (size_t)n_seeds == N && !is_device_ptr(seeds)) {
      bool alldesc = true;
      for (size_t k = 0; k < N && alldesc; ++k) alldesc = seeds[k] == (int)(N - 1 - k);
      // first choice: one CTA per connected component (PCSEG_CLUSTER_COMPONENTS=0: column bands only)
      const char* ce = getenv("PCSEG_CLUSTER_COMPONENTS");
      if (alldesc && !(ce && ce[0] == '0') && (st = run_cluster_components(b, b->xyz_dev, cfg, id_offset, &banded))) return st;
      if (alldesc && !banded && (st = run_cluster_banded(b, b->xyz_dev, cfg, id_offset, &banded))) return st;
    }
    if (!banded) {
      if ((st = check_seeds_host(seeds, n_seeds, N, c, &b->seeds, &seeds_dev))) return st;
      if ((st = run_cluster(b, b->xyz_dev, seeds_dev, n_seeds, cfg, id_offset, 0))) return st;
    }
  }
  if ((st = from_device(c, labels, b->labels.p, N * sizeof(int)))) return st;
  pcseg_regions* r = new pcseg_regions();
  st = collect_cluster(b, 0, r);
  if (st) { delete r; return st; }
  *out = r;
  return PCSEG_OK;
}

}  // extern "C"

namespace {

int run_seed_finder(pcseg_batch* b, const float* xyz_dev, const float* nrm_dev, int nbhd, float max_d,
                    int min_support) {
  pcseg_ctx* c = b->ctx;
  cudaStream_t stream = S(b);
  if (b->rows > b->cols) return PCSEG_ERR_UNSUPPORTED;  // transposed reads leave the buffer (UB in the reference)
  if (nbhd < 1 || nbhd * nbhd >= kSeedMaxKeys) return PCSEG_ERR_UNSUPPORTED;
  const size_t F = b->n_frames, N = b->N;
  CU_TRY(c, b->seeds.ensure(F * N * sizeof(int)));
  CU_TRY(c, b->n_seeds.ensure(F * sizeof(int)));
  SeedArgs a;
  a.xyz = xyz_dev; a.normals = nrm_dev; a.rows = b->rows; a.cols = b->cols;
  a.half = nbhd / 2; a.nbhd = nbhd; a.max_plane_distance = max_d; a.min_support = min_support;
  a.counts = b->queue.as<int>();  // scratch: the growth queue is idle here (F*qcap*2 >= F*N ints)
  a.seeds = b->seeds.as<int>(); a.stride = (int)N; a.n_seeds = b->n_seeds.as<int>();
  // counts are laid out [F][N] inside the queue buffer
  dim3 grid((unsigned)((N + 255) / 256), (unsigned)F);
  seed_support_count_kernel<<<grid, 256, 0, stream>>>(a);
  {
    const int nkeys = nbhd * nbhd + 1;
    const size_t smem = (size_t)kSeedWarps * nkeys * sizeof(int);
    if (smem > 48 * 1024)
      CU_TRY(c, cudaFuncSetAttribute(seed_order_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    seed_order_kernel<<<(unsigned)F, 32 * kSeedWarps, smem, stream>>>(a, nkeys);
  }
  c->launches += 2;
  CU_TRY(c, cudaGetLastError());
  return PCSEG_OK;
}

}  // namespace

extern "C" {

int pcseg_seeds_plane_support(pcseg_ctx* c, const float* xyz, const float* normals, int rows, int cols,
                              int neighborhood_size, float max_plane_distance, int min_num_support,
                              int* seeds_out, int cap, int* n_out) {
  if (!c || !xyz || !normals || !seeds_out || !n_out) return PCSEG_ERR_NULL;
  if (rows <= 1 || cols <= 0) return PCSEG_ERR_DIMS;  // CHECK_GT(points.Rows(), 1), ps/segmentation.h:197
  cudaSetDevice(c->device);
  pcseg_batch* b;
  int st = get_scratch(c, rows, cols, &b);
  if (st) return st;
  const size_t N = b->N;
  const void *xyz_dev, *nrm_dev;
  if ((st = to_device(c, xyz, N * 3 * sizeof(float), &b->xyz, &xyz_dev))) return st;
  if ((st = to_device(c, normals, N * 3 * sizeof(float), &b->normals, &nrm_dev))) return st;
  if ((st = run_seed_finder(b, (const float*)xyz_dev, (const float*)nrm_dev, neighborhood_size,
                            max_plane_distance, min_num_support))) return st;
  int n = 0;
  CU_TRY(c, cudaMemcpyAsync(&n, b->n_seeds.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  *n_out = n;
  const int ncopy = std::min(n, cap);
  if (ncopy > 0 && (st = from_device(c, seeds_out, b->seeds.p, (size_t)ncopy * sizeof(int)))) return st;
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  return PCSEG_OK;
}

int pcseg_seeds_average_normals(pcseg_ctx* c, const float* xyz, const float* normals, int rows, int cols,
                                int neighborhood_size, int min_num_valid_normals, float min_avg_normal_length,
                                int* seeds_out, int cap, int* n_out) {
  if (!c || !xyz || !normals || !seeds_out || !n_out) return PCSEG_ERR_NULL;
  if (rows <= 1 || cols <= 0) return PCSEG_ERR_DIMS;  // CHECK_GT(points.Rows(), 1), ps/segmentation.h:143
  if (rows > cols) return PCSEG_ERR_UNSUPPORTED;       // transposed reads leave the buffer (UB in the reference)
  if (neighborhood_size < 1 || neighborhood_size > kAvgMaxNbhd) return PCSEG_ERR_UNSUPPORTED;
  *n_out = 0;
  if (rows < neighborhood_size || cols < neighborhood_size) return PCSEG_OK;  // :150-152
  cudaSetDevice(c->device);
  pcseg_batch* b;
  int st = get_scratch(c, rows, cols, &b);
  if (st) return st;
  const size_t N = b->N;
  const void* nrm_dev;
  if ((st = to_device(c, normals, N * 3 * sizeof(float), &b->normals, &nrm_dev))) return st;
  CU_TRY(c, b->seeds.ensure(N * sizeof(int)));
  CU_TRY(c, b->n_seeds.ensure(sizeof(int)));
  AvgSeedArgs a;
  a.normals = (const float*)nrm_dev; a.rows = rows; a.cols = cols;
  a.nbhd = neighborhood_size; a.half = neighborhood_size / 2; a.min_valid = min_num_valid_normals;
  a.sq_min_len = min_avg_normal_length * min_avg_normal_length;  // :146-147
  a.flags = b->queue.as<unsigned char>(); a.flag_stride = N;      // scratch: the idle growth queue
  a.seeds = b->seeds.as<int>(); a.stride = (int)N; a.n_seeds = b->n_seeds.as<int>();
  const int nrows = rows - 2 * a.half;
  seed_avg_normals_flag_kernel<<<dim3((nrows + 63) / 64, 1), 64, 0, c->stream>>>(a);
  seed_avg_normals_order_kernel<<<1, 1024, 0, c->stream>>>(a);
  c->launches += 2;
  CU_TRY(c, cudaGetLastError());
  int n = 0;
  CU_TRY(c, cudaMemcpyAsync(&n, b->n_seeds.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  *n_out = n;
  const int ncopy = std::min(n, cap);
  if (ncopy > 0 && (st = from_device(c, seeds_out, b->seeds.p, (size_t)ncopy * sizeof(int)))) return st;
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  return PCSEG_OK;
}

// ---------------------------------------------------------------------------
// pcseg_seeds_average_normals_noalloc: the seed finder above for callers that cannot call
// cudaMalloc (no counterpart in the reference).
// Scratch = the flag byte per window position, in a file-scope device array instead of the
// workspace's queue buffer.  g_noalloc_flags names the array: in HOST code that name is a
// host-side stand-in, not a device address, so the address is asked for once per device with
// cudaGetSymbolAddress (that device current) and only that pointer is handed to the kernels.
// ---------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_flags[PCSEG_NOALLOC_MAX_CELLS];

namespace {
constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_flags_dev[kNoallocMaxDevices] = {};

// Device address of g_noalloc_flags on `device` (which must be current); looked up once per device.
cudaError_t noalloc_flags_address(int device, unsigned char** out) {
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_flags_dev[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_flags);
    if (e != cudaSuccess) return e;
    g_noalloc_flags_dev[device] = static_cast<unsigned char*>(p);
  }
  *out = g_noalloc_flags_dev[device];
  return cudaSuccess;
}
}  // namespace

int pcseg_seeds_average_normals_noalloc(pcseg_ctx* c, const float* normals_dev, int rows, int cols,
                                        int neighborhood_size, int min_num_valid_normals,
                                        float min_avg_normal_length, int* seeds_out_dev, int cap, int* n_out_dev) {
  if (!c || !normals_dev || !seeds_out_dev || !n_out_dev) return PCSEG_ERR_NULL;
  if (rows <= 1 || cols <= 0) return PCSEG_ERR_DIMS;   // CHECK_GT(points.Rows(), 1), ps/segmentation.h:143
  if (rows > cols) return PCSEG_ERR_UNSUPPORTED;        // as pcseg_seeds_average_normals
  if (neighborhood_size < 1 || neighborhood_size > kAvgMaxNbhd) return PCSEG_ERR_UNSUPPORTED;
  if (c->device < 0 || c->device >= kNoallocMaxDevices) return PCSEG_ERR_UNSUPPORTED;
  const long long N = (long long)rows * cols;
  if (N > (long long)PCSEG_NOALLOC_MAX_CELLS) {
    c->last_error = "pcseg_seeds_average_normals_noalloc: rows * cols = " + std::to_string(N) +
                    " exceeds the static scratch of PCSEG_NOALLOC_MAX_CELLS = " +
                    std::to_string((long long)PCSEG_NOALLOC_MAX_CELLS) + " cells; use pcseg_seeds_average_normals";
    return PCSEG_ERR_CAPACITY;
  }
  CU_TRY(c, cudaSetDevice(c->device));
  if (!is_device_ptr(normals_dev) || !is_device_ptr(seeds_out_dev) || !is_device_ptr(n_out_dev)) {
    c->last_error = "pcseg_seeds_average_normals_noalloc: normals, seeds_out and n_out must be device pointers "
                    "(a host pointer would need a staging allocation)";
    return PCSEG_ERR_UNSUPPORTED;
  }
  const int half = neighborhood_size / 2;
  if (rows < neighborhood_size || cols < neighborhood_size) {  // ps/segmentation.h:150-152: no seeds
    CU_TRY(c, cudaMemsetAsync(n_out_dev, 0, sizeof(int), c->stream));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
    return PCSEG_OK;
  }
  const int nrows = rows - 2 * half;
  const long long positions = (long long)nrows * cols;
  if ((long long)cap < positions) {
    c->last_error = "pcseg_seeds_average_normals_noalloc: cap = " + std::to_string(cap) +
                    " is below the number of window positions " + std::to_string(positions);
    return PCSEG_ERR_CAPACITY;
  }
  unsigned char* flags = nullptr;
  CU_TRY(c, noalloc_flags_address(c->device, &flags));
  AvgSeedArgs a;
  a.normals = normals_dev; a.rows = rows; a.cols = cols;
  a.nbhd = neighborhood_size; a.half = half; a.min_valid = min_num_valid_normals;
  a.sq_min_len = min_avg_normal_length * min_avg_normal_length;  // ps/segmentation.h:146-147
  a.flags = flags; a.flag_stride = (size_t)PCSEG_NOALLOC_MAX_CELLS;
  a.seeds = seeds_out_dev; a.stride = cap; a.n_seeds = n_out_dev;
  // every flag in [0, positions) is written by the first kernel before the second reads it:
  // the previous call's contents of the shared array need no clearing
  seed_avg_normals_flag_kernel<<<dim3((nrows + 63) / 64, 1), 64, 0, c->stream>>>(a);
  seed_avg_normals_order_kernel<<<1, 1024, 0, c->stream>>>(a);
  c->launches += 2;
  CU_TRY(c, cudaGetLastError());
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  return PCSEG_OK;
}

int pcseg_seeds_last_planar_regions(pcseg_ctx* c, const float* xyz, const float* normals, int n_points, int n_prev,
                                    const int* prev_inlier_counts, const float* prev_centroids,
                                    const float* prev_plane_normals, const float q_xyzw[4], const float translation[3],
                                    float max_distance, float max_normal_difference_angle, int* seeds_inout, int cap,
                                    int* n_inout) {
  if (!c || !xyz || !normals || !q_xyzw || !translation || !seeds_inout || !n_inout) return PCSEG_ERR_NULL;
  if (n_prev > 0 && (!prev_inlier_counts || !prev_centroids || !prev_plane_normals)) return PCSEG_ERR_NULL;
  if (n_points < 0 || n_prev < 0 || *n_inout < 0) return PCSEG_ERR_DIMS;
  if (n_prev == 0 || n_points == 0) return PCSEG_OK;
  cudaSetDevice(c->device);
  // std::multimap<int, SeedPoint>: ascending inlier count, insertion order within a count (:490-498)
  std::vector<int> order(n_prev);
  for (int i = 0; i < n_prev; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int l, int r) { return prev_inlier_counts[l] < prev_inlier_counts[r]; });
  std::vector<float> cen(3 * (size_t)n_prev), nrm(3 * (size_t)n_prev);
  for (int i = 0; i < n_prev; ++i)
    for (int k = 0; k < 3; ++k) {
      cen[3 * i + k] = prev_centroids[3 * order[i] + k];
      nrm[3 * i + k] = prev_plane_normals[3 * order[i] + k];
    }
  DevBuf sx, sn, sc, sp, so;
  const void *xyz_dev, *nrm_dev;
  int st;
  if ((st = to_device(c, xyz, (size_t)n_points * 3 * sizeof(float), &sx, &xyz_dev))) return st;
  if ((st = to_device(c, normals, (size_t)n_points * 3 * sizeof(float), &sn, &nrm_dev))) return st;
  CU_TRY(c, sc.ensure(cen.size() * sizeof(float)));
  CU_TRY(c, sp.ensure(nrm.size() * sizeof(float)));
  CU_TRY(c, so.ensure((size_t)n_prev * sizeof(int)));
  CU_TRY(c, cudaMemcpyAsync(sc.p, cen.data(), cen.size() * sizeof(float), cudaMemcpyHostToDevice, c->stream));
  CU_TRY(c, cudaMemcpyAsync(sp.p, nrm.data(), nrm.size() * sizeof(float), cudaMemcpyHostToDevice, c->stream));
  LastSeedArgs a;
  a.xyz = (const float*)xyz_dev; a.normals = (const float*)nrm_dev; a.n_points = n_points;
  a.centroids = sc.as<float>(); a.pnormals = sp.as<float>();
  a.qx = q_xyzw[0]; a.qy = q_xyzw[1]; a.qz = q_xyzw[2]; a.qw = q_xyzw[3];
  a.tx = translation[0]; a.ty = translation[1]; a.tz = translation[2];
  a.max_sq = max_distance * max_distance;                                  // :485
  a.cos_max = (float)std::cos((double)max_normal_difference_angle);        // cosf, :486 (libm policy)
  a.out = so.as<int>();
  seed_last_regions_kernel<<<n_prev, 512, 0, c->stream>>>(a);
  ++c->launches;
  CU_TRY(c, cudaGetLastError());
  std::vector<int> found(n_prev);
  CU_TRY(c, cudaMemcpyAsync(found.data(), so.p, (size_t)n_prev * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  int n = *n_inout;
  bool fits = true;
  for (int i = 0; i < n_prev; ++i)
    if (found[i] >= 0) {
      if (n < cap) seeds_inout[n] = found[i];
      else fits = false;
      ++n;
    }
  *n_inout = n;
  return fits ? PCSEG_OK : PCSEG_ERR_CAPACITY;
}

int pcseg_fit_planes(pcseg_ctx* c, const float* xyz, int n_points, const int* offsets, const int* indices,
                     int n_regions, const float* weights, const float* hints, float* query_out, float* state_out) {
  if (!c || !xyz || !offsets || !query_out) return PCSEG_ERR_NULL;
  if (n_points < 0 || n_regions < 0) return PCSEG_ERR_DIMS;
  if (n_regions == 0) return PCSEG_OK;
  cudaSetDevice(c->device);
  // the CSR must be well-formed (AtUnsafe would read out of bounds): host-resident lists are checked
  int total = 0;
  if (!is_device_ptr(offsets)) {
    for (int r = 0; r < n_regions; ++r)
      if (offsets[r] < 0 || offsets[r + 1] < offsets[r]) return PCSEG_ERR_DIMS;
    total = offsets[n_regions];
  } else {
    CU_TRY(c, cudaMemcpy(&total, offsets + n_regions, sizeof(int), cudaMemcpyDeviceToHost));
  }
  if (total > 0 && !indices) return PCSEG_ERR_NULL;
  if (total > 0 && !is_device_ptr(indices))
    for (int i = 0; i < total; ++i)
      if (indices[i] < 0 || indices[i] >= n_points) return PCSEG_ERR_SEED_RANGE;
  DevBuf sx, so, si, sw, sh, sq, ss;
  const void *xyz_dev, *off_dev, *idx_dev = nullptr, *w_dev = nullptr, *h_dev = nullptr;
  int st;
  if ((st = to_device(c, xyz, (size_t)n_points * 3 * sizeof(float), &sx, &xyz_dev))) return st;
  if ((st = to_device(c, offsets, ((size_t)n_regions + 1) * sizeof(int), &so, &off_dev))) return st;
  if (total > 0 && (st = to_device(c, indices, (size_t)total * sizeof(int), &si, &idx_dev))) return st;
  if (weights && total > 0 && (st = to_device(c, weights, (size_t)total * sizeof(float), &sw, &w_dev))) return st;
  if (hints && (st = to_device(c, hints, (size_t)n_regions * 3 * sizeof(float), &sh, &h_dev))) return st;
  const bool q_dev = is_device_ptr(query_out), s_dev = state_out && is_device_ptr(state_out);
  FitArgs a;
  a.xyz = (const float*)xyz_dev; a.offsets = (const int*)off_dev; a.indices = (const int*)idx_dev;
  a.weights = (const float*)w_dev; a.hints = (const float*)h_dev; a.n_regions = n_regions;
  a.query = query_out; a.state = state_out;
  if (!q_dev) { CU_TRY(c, sq.ensure((size_t)n_regions * 10 * sizeof(float))); a.query = sq.as<float>(); }
  if (state_out && !s_dev) { CU_TRY(c, ss.ensure((size_t)n_regions * 10 * sizeof(float))); a.state = ss.as<float>(); }
  const int blocks = std::min((n_regions + kFitWarps - 1) / kFitWarps, 148 * 8);
  fit_planes_csr_kernel<<<blocks, 32 * kFitWarps, 0, c->stream>>>(a);
  ++c->launches;
  CU_TRY(c, cudaGetLastError());
  if (!q_dev) CU_TRY(c, cudaMemcpyAsync(query_out, a.query, (size_t)n_regions * 10 * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
  if (state_out && !s_dev) CU_TRY(c, cudaMemcpyAsync(state_out, a.state, (size_t)n_regions * 10 * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  return PCSEG_OK;
}

int pcseg_plane_estimator_replay(pcseg_ctx* c, const int* ops, const float* args, int n_ops, const double* poses,
                                 int n_poses, float* query_out, int cap, int* n_queries) {
  if (!c || !n_queries || (n_ops > 0 && (!ops || !args)) || (cap > 0 && !query_out)) return PCSEG_ERR_NULL;
  if (n_ops < 0 || n_poses < 0 || cap < 0 || (n_poses > 0 && !poses)) return PCSEG_ERR_DIMS;
  *n_queries = 0;
  if (n_ops == 0) return PCSEG_OK;
  cudaSetDevice(c->device);
  DevBuf so, sa, sp, sq, sr;
  const void *ops_dev, *args_dev, *poses_dev = nullptr;
  int st;
  if ((st = to_device(c, ops, (size_t)n_ops * sizeof(int), &so, &ops_dev))) return st;
  if ((st = to_device(c, args, (size_t)n_ops * 4 * sizeof(float), &sa, &args_dev))) return st;
  if (n_poses > 0 && (st = to_device(c, poses, (size_t)n_poses * 7 * sizeof(double), &sp, &poses_dev))) return st;
  CU_TRY(c, sq.ensure(std::max<size_t>(1, (size_t)cap) * 10 * sizeof(float)));
  CU_TRY(c, sr.ensure(sizeof(int)));
  ReplayArgs a;
  a.ops = (const int*)ops_dev; a.args = (const float*)args_dev; a.n_ops = n_ops;
  a.poses = (const double*)poses_dev; a.n_poses = n_poses;
  a.out = sq.as<float>(); a.cap = cap; a.result = sr.as<int>();
  estimator_replay_kernel<<<1, 32, 0, c->stream>>>(a);
  ++c->launches;
  CU_TRY(c, cudaGetLastError());
  int res = 0;
  CU_TRY(c, cudaMemcpyAsync(&res, sr.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  if (res == -1) return PCSEG_ERR_CAPACITY;
  if (res == -2) return PCSEG_ERR_UNSUPPORTED;
  if (res < 0) return PCSEG_ERR_CONFIG;  // CHECK_GT(cumulative_weights_, 0.0f)
  *n_queries = res;
  if (res > 0) {
    cudaMemcpyKind k = is_device_ptr(query_out) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    CU_TRY(c, cudaMemcpyAsync(query_out, sq.p, (size_t)res * 10 * sizeof(float), k, c->stream));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
  }
  return PCSEG_OK;
}

namespace {
int estimators_apply(pcseg_ctx* c, float* dst, const float* src, int n, const double* poses, int per_state, bool merge) {
  if (!c || !dst || (merge && !src) || (!merge && !poses)) return PCSEG_ERR_NULL;
  if (n < 0) return PCSEG_ERR_DIMS;
  if (n == 0) return PCSEG_OK;
  cudaSetDevice(c->device);
  DevBuf sd, ss, sp, se;
  const size_t bytes = (size_t)n * 10 * sizeof(float);
  const void *dst_dev, *src_dev = nullptr, *poses_dev = nullptr;
  int st;
  const bool d_dev = is_device_ptr(dst);
  if ((st = to_device(c, dst, bytes, &sd, &dst_dev))) return st;
  if (merge && (st = to_device(c, src, bytes, &ss, &src_dev))) return st;
  if (poses && (st = to_device(c, poses, (size_t)(per_state ? n : 1) * 7 * sizeof(double), &sp, &poses_dev))) return st;
  CU_TRY(c, se.ensure(sizeof(int)));
  CU_TRY(c, cudaMemsetAsync(se.p, 0, sizeof(int), c->stream));
  const int blocks = (n + 127) / 128, stride = per_state ? 7 : 0;
  if (merge) estimator_merge_kernel<<<blocks, 128, 0, c->stream>>>((float*)dst_dev, (const float*)src_dev, n, (const double*)poses_dev, stride, se.as<int>());
  else estimator_transform_kernel<<<blocks, 128, 0, c->stream>>>((float*)dst_dev, n, (const double*)poses_dev, stride, se.as<int>());
  ++c->launches;
  CU_TRY(c, cudaGetLastError());
  int err = 0;
  CU_TRY(c, cudaMemcpyAsync(&err, se.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  if (!d_dev) CU_TRY(c, cudaMemcpyAsync(dst, dst_dev, bytes, cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  return err ? PCSEG_ERR_CONFIG : PCSEG_OK;
}
}  // namespace

int pcseg_plane_estimators_transform(pcseg_ctx* c, float* states, int n, const double* poses, int pose_per_state) {
  return estimators_apply(c, states, nullptr, n, poses, pose_per_state, false);
}
int pcseg_plane_estimators_merge(pcseg_ctx* c, float* dst, const float* src, int n, const double* poses, int pose_per_state) {
  return estimators_apply(c, dst, src, n, poses, pose_per_state, true);
}

int pcseg_mean_shift_modes_range(pcseg_ctx* c, const float* xyz, int rows, int cols, const int* labels,
                                 int iterations, int cell_begin, int cell_end, float* modes_out) {
  if (!c || !xyz || !labels || !modes_out) return PCSEG_ERR_NULL;  // CHECK_NE(labels, nullptr)
  if (rows <= 0 || cols <= 0) return PCSEG_ERR_DIMS;
  const size_t N = (size_t)rows * cols;
  if (cell_begin < 0 || cell_end > (long long)N || cell_begin > cell_end) return PCSEG_ERR_DIMS;
  const size_t M = (size_t)(cell_end - cell_begin);
  if (M == 0) return PCSEG_OK;
  cudaSetDevice(c->device);
  const void *xyz_dev, *lab_dev;
  int st;
  DevBuf sx, sl;  // transient staging (big clouds: not worth caching)
  if ((st = to_device(c, xyz, N * 3 * sizeof(float), &sx, &xyz_dev))) return st;
  if ((st = to_device(c, labels, N * sizeof(int), &sl, &lab_dev))) { sx.release(); return st; }
  const bool out_dev = is_device_ptr(modes_out);
  DevBuf so;
  ModeRecord* modes_dev = (ModeRecord*)modes_out;
  if (!out_dev) {
    if (so.ensure(M * sizeof(ModeRecord)) != cudaSuccess) { sx.release(); sl.release(); return PCSEG_ERR_CUDA; }
    modes_dev = so.as<ModeRecord>();
  }
  MeanShiftArgs a;
  a.xyz = (const float*)xyz_dev; a.labels = (const int*)lab_dev; a.rows = rows; a.cols = cols;
  a.iterations = iterations; a.cell_begin = cell_begin; a.cell_end = cell_end; a.modes = modes_dev;
  DevBuf spk;
  {
    const size_t Nall = (size_t)rows * cols;
    if (spk.ensure(Nall * sizeof(float4)) != cudaSuccess) { sx.release(); sl.release(); so.release(); return PCSEG_ERR_CUDA; }
    a.packed = spk.as<float4>();
    meanshift_pack_kernel<<<(unsigned)((Nall + 255) / 256), 256, 0, c->stream>>>(a.xyz, a.labels, (int)Nall, spk.as<float4>());
  }
  meanshift_modes_kernel<<<(unsigned)((M + 255) / 256), 256, 0, c->stream>>>(a);
  c->launches += 2;
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && !out_dev)
    e = cudaMemcpyAsync(modes_out, modes_dev, M * sizeof(ModeRecord), cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  sx.release(); sl.release(); so.release(); spk.release();
  if (e != cudaSuccess) { c->last_error = cudaGetErrorString(e); return PCSEG_ERR_CUDA; }
  return PCSEG_OK;
}

int pcseg_mean_shift_modes(pcseg_ctx* c, const float* xyz, int rows, int cols, const int* labels,
                           int iterations, float* modes_out) {
  if (rows <= 0 || cols <= 0) return c ? PCSEG_ERR_DIMS : PCSEG_ERR_NULL;
  return pcseg_mean_shift_modes_range(c, xyz, rows, cols, labels, iterations, 0, rows * cols, modes_out);
}

namespace {
// Device-wide inclusive scan of n packed 64-bit counters (msort.cuh); tiles = scratch of ceil(n / kScanTile) words.
int scan64(pcseg_ctx* c, const unsigned long long* v, int n, unsigned long long* out, DevBuf* tiles) {
  if (n <= 0) return PCSEG_OK;
  const int n_tiles = (n + kScanTile - 1) / kScanTile;
  CU_TRY(c, tiles->ensure((size_t)n_tiles * sizeof(unsigned long long)));
  scan64_reduce_kernel<<<n_tiles, kScanBlock, 0, c->stream>>>(v, n, tiles->as<unsigned long long>());
  scan64_tiles_kernel<<<1, kScanBlock, 0, c->stream>>>(tiles->as<unsigned long long>(), n_tiles);
  scan64_apply_kernel<<<n_tiles, kScanBlock, 0, c->stream>>>(v, n, tiles->as<unsigned long long>(), out);
  c->launches += 3;
  CU_TRY(c, cudaGetLastError());
  return PCSEG_OK;
}

// PCSEG_MS_DEBUG=1: wall time of the phases of a mean-shift call on stderr (synchronises at every mark)
struct MsPhaseClock {
  bool on = getenv("PCSEG_MS_DEBUG") != nullptr;
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  void mark(pcseg_ctx* c, const char* what) {
    if (!on) return;
    cudaStreamSynchronize(c->stream);
    const auto t1 = std::chrono::steady_clock::now();
    fprintf(stderr, "ms phase %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
};

// The sort of ps/mean_shift_segmentation.h:262-280 on the device (msort.cuh): intensities of the n
// cells (< 0: not a valid seed) -> cell ids of the valid seeds in the order the reference's
// std::sort (sort_mode 0, libstdc++ introsort replayed level by level) or std::stable_sort
// (sort_mode 1) leaves them.  *S_out = number of valid seeds.
int ms_sort_device(pcseg_ctx* c, const float* d_intensity, int n, int sort_mode, DevBuf* d_cells, int* S_out) {
  *S_out = 0;
  if (n <= 0) return PCSEG_OK;
  cudaStream_t st = c->stream;
  DevBuf d_v, d_incl, d_tiles, d_key, d_idx, d_seg, d_posL, d_posR, d_segs[2], d_cnt, d_hist, d_hincl;
  CU_TRY(c, d_v.ensure((size_t)n * 8));
  CU_TRY(c, d_incl.ensure((size_t)n * 8));
  CU_TRY(c, d_cnt.ensure(4 * sizeof(int)));  // [0] next segment count, [1] overflow / bad-key flags
  CU_TRY(c, cudaMemsetAsync(d_cnt.p, 0, 4 * sizeof(int), st));
  unsigned long long* v = d_v.as<unsigned long long>();
  unsigned long long* incl = d_incl.as<unsigned long long>();
  int* cnt = d_cnt.as<int>();
  const int gb = (n + 255) / 256;
  msort_valid_flag_kernel<<<gb, 256, 0, st>>>(d_intensity, n, v, cnt + 1);
  ++c->launches;
  int rc = scan64(c, v, n, incl, &d_tiles);
  if (rc) return rc;
  unsigned long long total = 0;
  int h_cnt[4];
  CU_TRY(c, cudaMemcpyAsync(&total, incl + (n - 1), 8, cudaMemcpyDeviceToHost, st));
  CU_TRY(c, cudaMemcpyAsync(h_cnt, cnt, sizeof(h_cnt), cudaMemcpyDeviceToHost, st));
  CU_TRY(c, cudaStreamSynchronize(st));
  if (h_cnt[1]) {
    c->last_error = "mean-shift intensities are not integers in [0, 255] (window counts expected)";
    return PCSEG_ERR_UNSUPPORTED;
  }
  const int S = (int)total;
  *S_out = S;
  if (S == 0) return PCSEG_OK;
  CU_TRY(c, d_key.ensure((size_t)S));
  CU_TRY(c, d_idx.ensure((size_t)S * sizeof(int)));
  CU_TRY(c, d_seg.ensure((size_t)S * sizeof(int)));
  CU_TRY(c, d_cells->ensure((size_t)S * sizeof(int)));
  unsigned char* key = d_key.as<unsigned char>();
  int* idx = d_idx.as<int>();
  int* seg = d_seg.as<int>();
  msort_compact_kernel<<<gb, 256, 0, st>>>(d_intensity, n, incl, key, idx, seg);
  ++c->launches;
  if (sort_mode == 0 && S > kSortThreshold) {
    // __introsort_loop, one level per iteration (stl_algo.h:1918-1936)
    const int seg_cap = S / (kSortThreshold + 1) + 2;
    CU_TRY(c, d_segs[0].ensure((size_t)seg_cap * sizeof(SortSeg)));
    CU_TRY(c, d_segs[1].ensure((size_t)seg_cap * sizeof(SortSeg)));
    CU_TRY(c, d_posL.ensure(((size_t)S + 1) * sizeof(int)));
    CU_TRY(c, d_posR.ensure(((size_t)S + 1) * sizeof(int)));
    SortSeg first{};
    first.lo = 0; first.hi = S;
    CU_TRY(c, cudaMemcpyAsync(d_segs[0].p, &first, sizeof(SortSeg), cudaMemcpyHostToDevice, st));
    int depth_limit = 0;
    while ((2 << depth_limit) <= S) ++depth_limit;  // std::__lg(n)
    depth_limit *= 2;
    int n_cur = 1, cur = 0;
    const int sb = (S + 255) / 256;
    for (int level = 0; n_cur > 0; ++level) {
      if (level >= depth_limit) {
        c->last_error = "std::sort would fall back to heapsort on this input (introsort depth limit): not replayed";
        return PCSEG_ERR_UNSUPPORTED;
      }
      SortSeg* segs = d_segs[cur].as<SortSeg>();
      SortSeg* next = d_segs[cur ^ 1].as<SortSeg>();
      const int gs = (n_cur + 255) / 256;
      msort_pivot_kernel<<<gs, 256, 0, st>>>(segs, n_cur, key, idx);
      msort_flag_kernel<<<sb, 256, 0, st>>>(segs, seg, key, S, v);
      c->launches += 2;
      if ((rc = scan64(c, v, S, incl, &d_tiles))) return rc;
      msort_pos_kernel<<<sb, 256, 0, st>>>(v, incl, S, d_posL.as<int>(), d_posR.as<int>());
      CU_TRY(c, cudaMemsetAsync(cnt, 0, sizeof(int), st));
      msort_cut_kernel<<<gs, 256, 0, st>>>(segs, n_cur, incl, d_posL.as<int>(), d_posR.as<int>(), next, cnt, seg_cap, cnt + 1);
      msort_swap_kernel<<<sb, 256, 0, st>>>(segs, seg, v, incl, d_posR.as<int>(), S, key, idx);
      c->launches += 3;
      CU_TRY(c, cudaGetLastError());
      CU_TRY(c, cudaMemcpyAsync(h_cnt, cnt, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
      CU_TRY(c, cudaStreamSynchronize(st));
      if (h_cnt[1]) { c->last_error = "sort segment table overflow"; return PCSEG_ERR_CAPACITY; }
      n_cur = h_cnt[0];
      cur ^= 1;
    }
  }
  // __final_insertion_sort == the stable sort of what the loop left (sort_mode 1: of the seed order)
  const int n_chunks = (S + kCountChunk - 1) / kCountChunk;
  const int hn = 256 * n_chunks;
  CU_TRY(c, d_hist.ensure((size_t)hn * 8));
  CU_TRY(c, d_hincl.ensure((size_t)hn * 8));
  const int hb = (n_chunks + kCountWarps - 1) / kCountWarps;
  msort_hist_kernel<<<hb, 32 * kCountWarps, 0, st>>>(key, S, n_chunks, d_hist.as<unsigned long long>());
  ++c->launches;
  if ((rc = scan64(c, d_hist.as<unsigned long long>(), hn, d_hincl.as<unsigned long long>(), &d_tiles))) return rc;
  msort_scatter_kernel<<<hb, 32 * kCountWarps, 0, st>>>(key, idx, S, n_chunks, d_hist.as<unsigned long long>(),
                                                        d_hincl.as<unsigned long long>(), d_cells->as<int>());
  ++c->launches;
  CU_TRY(c, cudaGetLastError());
  CU_TRY(c, cudaStreamSynchronize(st));  // the temporaries die with this scope
  return PCSEG_OK;
}

// Sort + growth phase on device-resident mode records (ps/mean_shift_segmentation.h:262-329).
int mean_shift_finish(pcseg_ctx* c, const float* xyz_dev, int rows, int cols, const ModeRecord* modes_dev,
                      const pcseg_cluster_config* cfg, int id_offset, int sort_mode, int* labels_dev,
                      pcseg_regions** out) {
  const size_t N = (size_t)rows * cols;
  MsPhaseClock clk;
  clk.mark(c, "before finish (modes)");
  // (1) intensities of the valid seeds, in seed order (ascending cloud index, :219-229,267-274)
  DevBuf d_int, d_cells;
  CU_TRY(c, d_int.ensure(N * sizeof(float)));
  ms_intensity_kernel<<<(unsigned)((N + 255) / 256), 256, 0, c->stream>>>(modes_dev, (int)N, d_int.as<float>());
  ++c->launches;
  // (2) the reference's sort call, :276-280, replayed on the device (msort.cuh; quirk Q7).
  // sort_mode 1 = stable (documented patch P2).
  int S = 0;
  {
    const int rc = ms_sort_device(c, d_int.as<float>(), (int)N, sort_mode, &d_cells, &S);
    if (rc) return rc;
  }
  d_int.release();
  clk.mark(c, "intensity + sort");
  pcseg_regions* res = new pcseg_regions();
  if (S == 0) { *out = res; return PCSEG_OK; }
  struct ResGuard {  // the result reaches the caller only on success (no live handle beside an error status)
    pcseg_regions* r; pcseg_regions** out; bool ok = false;
    ~ResGuard() { if (ok) *out = r; else delete r; }
  } guard{res, out};
  // (3) growth on the device
  const int rcap = (int)std::min<size_t>(N / (size_t)std::max(cfg->min_region_inliers, 1) + 1, (size_t)1 << 22);
  DevBuf d_seeds, d_claim, d_inl, d_meta, d_out;
  CU_TRY(c, d_seeds.ensure((size_t)S * sizeof(MsSeedRec)));
  CU_TRY(c, d_claim.ensure(N * sizeof(int)));
  CU_TRY(c, d_inl.ensure(N * sizeof(int)));
  CU_TRY(c, d_meta.ensure((size_t)rcap * sizeof(MsRegionMeta)));
  CU_TRY(c, d_out.ensure(4 * sizeof(int)));
  CU_TRY(c, cudaMemsetAsync(d_out.p, 0, 4 * sizeof(int), c->stream));
  ms_gather_kernel<<<(unsigned)((S + 255) / 256), 256, 0, c->stream>>>(modes_dev, d_cells.as<int>(), S, rows,
                                                                        d_seeds.as<MsSeedRec>());
  fill_int_kernel<<<(unsigned)std::min<size_t>((N + 255) / 256, 65535), 256, 0, c->stream>>>(d_claim.as<int>(), N, INT_MAX);
  MsGrowArgs g;
  g.xyz = xyz_dev; g.labels = labels_dev; g.claim = d_claim.as<int>(); g.seeds = d_seeds.as<MsSeedRec>();
  g.n_seeds = S; g.rows = rows; g.cols = cols; g.half_window = cfg->half_search_window;
  g.min_inliers = cfg->min_region_inliers; g.id_offset = id_offset; g.inliers = d_inl.as<int>();
  g.meta = d_meta.as<MsRegionMeta>(); g.rcap = rcap; g.out = d_out.as<int>(); g.first_alive = nullptr;
  DevBuf d_alive;
  CU_TRY(c, d_alive.ensure((size_t)S));
  CU_TRY(c, cudaMemsetAsync(d_alive.p, 1, (size_t)S, c->stream));
  g.alive = d_alive.as<unsigned char>();
  c->launches += 2;
  int h_out[4] = {0, 0, 0, 0};
  clk.mark(c, "growth buffers + gather");
  const char* ms_mode = getenv("PCSEG_MS_GROWTH");
  if (ms_mode && strcmp(ms_mode, "serial") == 0) {
    // growth in rounds (meanshift_grow.cuh): the serial CTA accepts a few regions, the filter kernel
    // marks the modes they killed on the whole GPU, the CTA resumes behind the last batch it finished
    int pos = 0, n_reg = 0, n_inl = 0, r_applied = 0;
    while (pos < S) {
      if (n_reg > r_applied) {
        const unsigned fgrid = (unsigned)std::min<size_t>(((size_t)(S - pos) + 255) / 256, (size_t)148 * 8);
        ms_filter_kernel<<<fgrid, 256, 0, c->stream>>>(g, pos, r_applied, n_reg);
        r_applied = n_reg;
        ++c->launches;
      }
      ms_grow_kernel<<<1, kMsGrowBlock, 0, c->stream>>>(g, pos, r_applied, n_inl);
      ++c->launches;
      CU_TRY(c, cudaGetLastError());
      CU_TRY(c, cudaMemcpyAsync(h_out, d_out.p, sizeof(h_out), cudaMemcpyDeviceToHost, c->stream));
      CU_TRY(c, cudaStreamSynchronize(c->stream));
      if (h_out[1]) break;
      if (h_out[3] <= pos && h_out[0] == n_reg) { c->last_error = "mean-shift growth made no progress"; d_alive.release(); return PCSEG_ERR_CUDA; }
      pos = h_out[3]; n_reg = h_out[0]; n_inl = h_out[2];
    }
  } else {
    // waves of regions grown at the same time and resolved in order (meanshift_wave.cuh)
    {   // their temporary labels start at kMsTmpBase: the caller's grid must stay below
      CU_TRY(c, cudaMemsetAsync(d_out.as<int>() + 1, 0, sizeof(int), c->stream));
      label_range_kernel<<<148 * 8, 256, 0, c->stream>>>(labels_dev, N, INT_MIN, kMsTmpBase, d_out.as<int>() + 1);
      int bad = 0;
      CU_TRY(c, cudaMemcpyAsync(&bad, d_out.as<int>() + 1, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
      CU_TRY(c, cudaStreamSynchronize(c->stream));
      CU_TRY(c, cudaMemsetAsync(d_out.as<int>() + 1, 0, sizeof(int), c->stream));
      ++c->launches;
      if (bad) {
        c->last_error = "mean shift: label values >= 0x40000000 in the caller's grid (reserved for temporary labels); "
                        "PCSEG_MS_GROWTH=serial has no such limit";
        return PCSEG_ERR_UNSUPPORTED;
      }
    }
    const int kmax = (int)std::min<size_t>(kMsWaveMax, (size_t)S);
    size_t cap_cells = (size_t)1 << 18;   // ranks above 0: 1 MB each (rank 0 has room for every cell)
    if (const char* e = getenv("PCSEG_MS_SLOT_CAP")) cap_cells = (size_t)std::max(1, atoi(e));   // tests: force slot overflows
    const int slot_cap = (int)std::min<size_t>(N, cap_cells);
    DevBuf d_slots, d_wave;
    CU_TRY(c, d_slots.ensure((N + (size_t)(kmax - 1) * slot_cap) * sizeof(int)));
    CU_TRY(c, d_wave.ensure(kWvInts * sizeof(int)));
    CU_TRY(c, cudaMemsetAsync(d_wave.p, 0, kWvInts * sizeof(int), c->stream));
    MsWaveArgs w;
    g.first_alive = d_wave.as<int>() + kWvFirstAlive;
    w.g = g; w.slot_inliers = d_slots.as<int>(); w.slot0 = N; w.slot_cap = slot_cap; w.wave = d_wave.as<int>();
    int h_wave[4] = {0, 0, 0, 0};
    int pos = 0, n_reg = 0, n_inl = 0, r_applied = 0, K = kmax;
    bool first = true;
    while (pos < S) {
      if (first || n_reg > r_applied) {   // dead-mode flags up to date before the candidates are picked
        const unsigned fgrid = (unsigned)std::min<size_t>(((size_t)(S - pos) + 255) / 256, (size_t)148 * 8);
        CU_TRY(c, cudaMemsetAsync(w.g.first_alive, 0x7f, sizeof(int), c->stream));   // "none" until the filter finds one
        ms_filter_kernel<<<fgrid, 256, 0, c->stream>>>(w.g, pos, r_applied, n_reg);
        r_applied = n_reg;
        first = false;
        ++c->launches;
      }
      ms_wave_select_kernel<<<1, 1024, 0, c->stream>>>(w, pos, K);
      ms_wave_grow_kernel<<<K, kMsGrowBlock, 0, c->stream>>>(w);
      ms_wave_check_kernel<<<dim3(64, K), 256, 0, c->stream>>>(w);
      ms_wave_plan_kernel<<<1, kMsWaveMax, 0, c->stream>>>(w, n_reg, n_inl);
      ms_wave_undo_kernel<<<dim3(16, K), 256, 0, c->stream>>>(w);
      ms_wave_dead_kernel<<<(K + 7) / 8, 256, 0, c->stream>>>(w);
      ms_wave_commit_kernel<<<dim3(16, K), 256, 0, c->stream>>>(w);
      c->launches += 7;
      CU_TRY(c, cudaGetLastError());
      CU_TRY(c, cudaMemcpyAsync(h_out, d_out.p, sizeof(h_out), cudaMemcpyDeviceToHost, c->stream));
      CU_TRY(c, cudaMemcpyAsync(h_wave, d_wave.p, sizeof(h_wave), cudaMemcpyDeviceToHost, c->stream));
      CU_TRY(c, cudaStreamSynchronize(c->stream));
      if (h_out[1]) break;
      const int n_cand = h_wave[kWvNCand], stop = h_wave[kWvStop];
      if (n_cand == 0) break;   // no live mode left
      if (h_out[3] <= pos && stop == 0) { c->last_error = "mean-shift growth made no progress"; return PCSEG_ERR_CUDA; }
      if (getenv("PCSEG_MS_DEBUG"))
        fprintf(stderr, "ms wave: pos %d K %d cand %d stop %d accepted %d regions %d\n", pos, K, n_cand, stop,
                h_wave[kWvNAcc], h_out[0]);
      pos = h_out[3]; n_reg = h_out[0]; n_inl = h_out[2];
      K = stop == n_cand ? std::min(2 * K, kmax) : std::max(K / 2, std::min(4, kmax));
    }
  }
  clk.mark(c, "growth");
  d_alive.release();
  const int R = h_out[0], total = h_out[2];
  std::vector<MsRegionMeta> meta(R);
  res->inlier_indices.resize(total);
  if (R > 0) CU_TRY(c, cudaMemcpyAsync(meta.data(), d_meta.p, (size_t)R * sizeof(MsRegionMeta), cudaMemcpyDeviceToHost, c->stream));
  if (total > 0) CU_TRY(c, cudaMemcpyAsync(res->inlier_indices.data(), d_inl.p, (size_t)total * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  for (int r = 0; r < R; ++r) {
    res->inlier_offsets.push_back(meta[r].inlier_begin + meta[r].inlier_count);
    res->label_ids.push_back(meta[r].label_id);
    res->seed_points.insert(res->seed_points.end(), meta[r].seed, meta[r].seed + 3);
  }
  clk.mark(c, "results to host");
  d_cells.release(); d_seeds.release(); d_claim.release(); d_inl.release(); d_meta.release(); d_out.release();
  clk.mark(c, "free");
  if (h_out[1]) { c->last_error = "mean-shift region table overflow"; return PCSEG_ERR_CAPACITY; }
  guard.ok = true;
  return PCSEG_OK;
}
}  // namespace

int pcseg_mean_shift_sort_order(pcseg_ctx* c, const float* intensities, int n, int sort_mode, int* order_out,
                                int* n_out) {
  if (!c || !intensities || !order_out || !n_out) return PCSEG_ERR_NULL;
  if (n < 0) return PCSEG_ERR_DIMS;
  *n_out = 0;
  if (n == 0) return PCSEG_OK;
  cudaSetDevice(c->device);
  DevBuf si, d_cells;
  const void* int_dev;
  int st = to_device(c, intensities, (size_t)n * sizeof(float), &si, &int_dev);
  if (st) return st;
  int S = 0;
  if ((st = ms_sort_device(c, (const float*)int_dev, n, sort_mode, &d_cells, &S))) return st;
  *n_out = S;
  if (S > 0) {
    cudaMemcpyKind k = is_device_ptr(order_out) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    CU_TRY(c, cudaMemcpyAsync(order_out, d_cells.p, (size_t)S * sizeof(int), k, c->stream));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
  }
  return PCSEG_OK;
}

int pcseg_mean_shift_from_modes(pcseg_ctx* c, const float* xyz, int rows, int cols, const float* modes,
                                const pcseg_cluster_config* cfg, int id_offset, int sort_mode, int* labels,
                                pcseg_regions** out) {
  if (!c || !xyz || !modes || !cfg || !labels || !out) return PCSEG_ERR_NULL;
  if (rows <= 0 || cols <= 0) return PCSEG_ERR_DIMS;
  int st = check_cluster_cfg(cfg);
  if (st) return st;
  cudaSetDevice(c->device);
  const size_t N = (size_t)rows * cols;
  DevBuf sx, sl, sm;
  const void *xyz_dev, *modes_dev;
  if ((st = to_device(c, xyz, N * 3 * sizeof(float), &sx, &xyz_dev))) return st;
  if ((st = to_device(c, modes, N * sizeof(ModeRecord), &sm, &modes_dev))) { sx.release(); return st; }
  int* lab_dev = labels;
  if (!is_device_ptr(labels)) {
    const void* tmp;
    if ((st = to_device(c, labels, N * sizeof(int), &sl, &tmp))) { sx.release(); sm.release(); return st; }
    lab_dev = (int*)sl.p;
  }
  st = mean_shift_finish(c, (const float*)xyz_dev, rows, cols, (const ModeRecord*)modes_dev, cfg, id_offset,
                         sort_mode, lab_dev, out);
  if (st == PCSEG_OK && lab_dev != labels) {
    st = from_device(c, labels, lab_dev, N * sizeof(int));
    if (st == PCSEG_OK && cudaStreamSynchronize(c->stream) != cudaSuccess) st = PCSEG_ERR_CUDA;
  }
  sx.release(); sl.release(); sm.release();
  return st;
}

int pcseg_mean_shift(pcseg_ctx* c, const float* xyz, int rows, int cols, const pcseg_cluster_config* cfg,
                     int iterations, int id_offset, int sort_mode, int* labels, pcseg_regions** out) {
  if (!c || !xyz || !cfg || !labels || !out) return PCSEG_ERR_NULL;  // CHECK_NE(labels/regions, nullptr), :212-213
  if (rows <= 0 || cols <= 0) return PCSEG_ERR_DIMS;
  int st = check_cluster_cfg(cfg);
  if (st) return st;
  cudaSetDevice(c->device);
  const size_t N = (size_t)rows * cols;
  MsPhaseClock clk;
  DevBuf sx, sl, sm;
  const void* xyz_dev;
  if ((st = to_device(c, xyz, N * 3 * sizeof(float), &sx, &xyz_dev))) return st;
  int* lab_dev = labels;
  if (!is_device_ptr(labels)) {
    const void* tmp;
    if ((st = to_device(c, labels, N * sizeof(int), &sl, &tmp))) { sx.release(); return st; }
    lab_dev = (int*)sl.p;
  }
  if (sm.ensure(N * sizeof(ModeRecord)) != cudaSuccess) { sx.release(); sl.release(); return PCSEG_ERR_CUDA; }
  clk.mark(c, "inputs to device + alloc");
  MeanShiftArgs a;
  a.xyz = (const float*)xyz_dev; a.labels = lab_dev; a.rows = rows; a.cols = cols;
  a.iterations = iterations; a.cell_begin = 0; a.cell_end = (int)N; a.modes = sm.as<ModeRecord>();
  DevBuf spk;
  if (spk.ensure(N * sizeof(float4)) != cudaSuccess) { sx.release(); sl.release(); sm.release(); return PCSEG_ERR_CUDA; }
  a.packed = spk.as<float4>();
  meanshift_pack_kernel<<<(unsigned)((N + 255) / 256), 256, 0, c->stream>>>(a.xyz, a.labels, (int)N, spk.as<float4>());
  meanshift_modes_kernel<<<(unsigned)((N + 255) / 256), 256, 0, c->stream>>>(a);
  c->launches += 2;
  st = mean_shift_finish(c, (const float*)xyz_dev, rows, cols, sm.as<ModeRecord>(), cfg, id_offset, sort_mode,
                         lab_dev, out);
  if (st == PCSEG_OK && lab_dev != labels) {
    st = from_device(c, labels, lab_dev, N * sizeof(int));
    if (st == PCSEG_OK && cudaStreamSynchronize(c->stream) != cudaSuccess) st = PCSEG_ERR_CUDA;
  }
  clk.mark(c, "labels to host");
  sx.release(); sl.release(); sm.release();
  clk.mark(c, "free inputs");
  return st;
}

// ---- unorganized clouds: voxel-hash Euclidean clustering (voxel_cluster.cuh) ----------------------
int pcseg_cluster_unorganized(pcseg_ctx* c, const float* xyz, int n, const pcseg_voxel_cluster_config* cfg,
                              int initial_id_offset, int* labels, pcseg_regions** out) {
  if (!c || !xyz || !cfg || !labels || !out) return PCSEG_ERR_NULL;
  if (n <= 0) return PCSEG_ERR_DIMS;
  if (!(cfg->radius > 0.0f) || !(cfg->radius <= FLT_MAX) || cfg->min_region_inliers < 1) return PCSEG_ERR_CONFIG;
  cudaSetDevice(c->device);
  cudaStream_t st = c->stream;
  DevBuf sx, sl;
  const void* xyz_dev;
  int rc = to_device(c, xyz, (size_t)n * 3 * sizeof(float), &sx, &xyz_dev);
  if (rc) return rc;
  int* lab_dev = labels;
  if (!is_device_ptr(labels)) {
    const void* tmp;
    if ((rc = to_device(c, labels, (size_t)n * sizeof(int), &sl, &tmp))) { sx.release(); return rc; }
    lab_dev = (int*)sl.p;
  }
  size_t cap = 1024;
  while (cap < 2 * (size_t)n) cap <<= 1;
  DevBuf d_keys, d_head, d_next, d_slot, d_parent, d_count, d_v, d_incl, d_tiles, d_ord, d_rb, d_rc, d_rr,
      d_idx[2], d_key, d_hist, d_hincl, d_cen, d_box, d_sv, d_sincl, d_cs, d_cb, d_cc, d_cbox;
  DevBuf* all[] = {&sx, &sl, &d_keys, &d_head, &d_next, &d_slot, &d_parent, &d_count, &d_v, &d_incl, &d_tiles, &d_ord,
                   &d_rb, &d_rc, &d_rr, &d_idx[0], &d_idx[1], &d_key, &d_hist, &d_hincl, &d_cen, &d_box,
                   &d_sv, &d_sincl, &d_cs, &d_cb, &d_cc};
  struct Cleanup { DevBuf** b; int k; ~Cleanup() { for (int i = 0; i < k; ++i) b[i]->release(); } } cleanup{all, (int)(sizeof(all) / sizeof(all[0]))};
  CU_TRY(c, d_keys.ensure(cap * 8));
  CU_TRY(c, d_head.ensure(cap * 4));
  CU_TRY(c, d_next.ensure((size_t)n * 4));
  CU_TRY(c, d_slot.ensure((size_t)n * 4));
  CU_TRY(c, d_parent.ensure((size_t)n * 4));
  CU_TRY(c, d_count.ensure((size_t)n * 4));
  CU_TRY(c, d_v.ensure((size_t)n * 8));
  CU_TRY(c, d_incl.ensure((size_t)n * 8));
  CU_TRY(c, cudaMemsetAsync(d_keys.p, 0xff, cap * 8, st));
  CU_TRY(c, cudaMemsetAsync(d_head.p, 0, cap * 4, st));
  VoxArgs a;
  a.xyz = (const float*)xyz_dev; a.labels = lab_dev; a.n = n;
  a.radius = cfg->radius; a.r2 = cfg->radius * cfg->radius; a.cell = 0.57f * cfg->radius;
  a.min_inliers = cfg->min_region_inliers; a.id_offset = initial_id_offset;
  a.keys = d_keys.as<unsigned long long>(); a.head = d_head.as<int>(); a.cap_mask = (int)(cap - 1);
  a.next = d_next.as<int>(); a.slot_of = d_slot.as<int>(); a.parent = d_parent.as<int>(); a.count = d_count.as<int>();
  const unsigned grid = (unsigned)((n + 255) / 256);
  vox_insert_kernel<<<grid, 256, 0, st>>>(a);
  // cells -> CSR of their points
  CU_TRY(c, d_sv.ensure(cap * 8));
  CU_TRY(c, d_sincl.ensure(cap * 8));
  const unsigned sgrid = (unsigned)((cap + 255) / 256);
  vox_slotflag_kernel<<<sgrid, 256, 0, st>>>(a, (int)cap, d_sv.as<unsigned long long>());
  c->launches += 2;
  if ((rc = scan64(c, d_sv.as<unsigned long long>(), (int)cap, d_sincl.as<unsigned long long>(), &d_tiles))) return rc;
  unsigned long long slast = 0;
  CU_TRY(c, cudaMemcpyAsync(&slast, d_sincl.as<unsigned long long>() + (cap - 1), 8, cudaMemcpyDeviceToHost, st));
  CU_TRY(c, cudaStreamSynchronize(st));
  const int n_cells = (int)(slast & 0xffffffffull);
  if (n_cells > 0) {
    CU_TRY(c, d_cs.ensure((size_t)n_cells * 4));
    CU_TRY(c, d_cb.ensure((size_t)n_cells * 4));
    CU_TRY(c, d_cc.ensure((size_t)n_cells * 4));
    vox_cells_kernel<<<sgrid, 256, 0, st>>>(a, (int)cap, d_sincl.as<unsigned long long>(), d_cs.as<int>(), d_cb.as<int>(), d_cc.as<int>());
    vox_scatter_kernel<<<grid, 256, 0, st>>>(a, d_sincl.as<unsigned long long>(), d_cb.as<int>());
    vox_union_cell_kernel<<<grid, 256, 0, st>>>(a, d_sincl.as<unsigned long long>(), d_cb.as<int>());
    CU_TRY(c, d_cbox.ensure((size_t)n_cells * 24));
    const unsigned wgrid = (unsigned)(((size_t)n_cells * 32 + 255) / 256);
    vox_cellbox_kernel<<<wgrid, 256, 0, st>>>(a, d_cb.as<int>(), d_cc.as<int>(), n_cells, d_cbox.as<float>());
    vox_union_pairs_kernel<true><<<wgrid, 256, 0, st>>>(a, d_sincl.as<unsigned long long>(), d_cs.as<int>(), d_cb.as<int>(),
                                                        d_cc.as<int>(), n_cells, d_cbox.as<float>());
    vox_union_pairs_kernel<false><<<wgrid, 256, 0, st>>>(a, d_sincl.as<unsigned long long>(), d_cs.as<int>(), d_cb.as<int>(),
                                                         d_cc.as<int>(), n_cells, d_cbox.as<float>());
    c->launches += 6;
  }
  vox_flatten_kernel<<<grid, 256, 0, st>>>(a);
  vox_rootflag_kernel<<<grid, 256, 0, st>>>(a, d_v.as<unsigned long long>());
  c->launches += 2;
  if ((rc = scan64(c, d_v.as<unsigned long long>(), n, d_incl.as<unsigned long long>(), &d_tiles))) return rc;
  unsigned long long last = 0;
  CU_TRY(c, cudaMemcpyAsync(&last, d_incl.as<unsigned long long>() + (n - 1), 8, cudaMemcpyDeviceToHost, st));
  CU_TRY(c, cudaStreamSynchronize(st));
  const int R = (int)(last & 0xffffffffull), total = (int)(last >> 32);
  pcseg_regions* res = new pcseg_regions();
  if (R > 0) {
    CU_TRY(c, d_ord.ensure((size_t)n * 4));
    CU_TRY(c, d_rb.ensure((size_t)R * 4));
    CU_TRY(c, d_rc.ensure((size_t)R * 4));
    CU_TRY(c, d_rr.ensure((size_t)R * 4));
    vox_label_kernel<<<grid, 256, 0, st>>>(a, d_incl.as<unsigned long long>(), R, d_ord.as<int>(), d_rb.as<int>(),
                                           d_rc.as<int>(), d_rr.as<int>());
    ++c->launches;
    // inlier lists: stable LSD radix sort of the point indices by region ordinal (points outside every
    // region carry the ordinal R and end up behind the lists); ascending index inside a region
    CU_TRY(c, d_idx[0].ensure((size_t)n * 4));
    CU_TRY(c, d_idx[1].ensure((size_t)n * 4));
    CU_TRY(c, d_key.ensure((size_t)n));
    const int n_chunks = (n + kCountChunk - 1) / kCountChunk;
    const int hn = 256 * n_chunks;
    CU_TRY(c, d_hist.ensure((size_t)hn * 8));
    CU_TRY(c, d_hincl.ensure((size_t)hn * 8));
    const int hb = (n_chunks + kCountWarps - 1) / kCountWarps;
    vox_iota_kernel<<<grid, 256, 0, st>>>(d_idx[0].as<int>(), n);
    ++c->launches;
    int bits = 1;
    while ((R >> bits) != 0) ++bits;
    int cur = 0;
    for (int shift = 0; shift < bits; shift += 8) {
      vox_digit_kernel<<<grid, 256, 0, st>>>(d_ord.as<int>(), d_idx[cur].as<int>(), n, shift, d_key.as<unsigned char>());
      msort_hist_kernel<<<hb, 32 * kCountWarps, 0, st>>>(d_key.as<unsigned char>(), n, n_chunks, d_hist.as<unsigned long long>());
      c->launches += 2;
      if ((rc = scan64(c, d_hist.as<unsigned long long>(), hn, d_hincl.as<unsigned long long>(), &d_tiles))) { delete res; return rc; }
      msort_scatter_kernel<<<hb, 32 * kCountWarps, 0, st>>>(d_key.as<unsigned char>(), d_idx[cur].as<int>(), n, n_chunks,
                                                            d_hist.as<unsigned long long>(), d_hincl.as<unsigned long long>(),
                                                            d_idx[cur ^ 1].as<int>());
      ++c->launches;
      cur ^= 1;
    }
    CU_TRY(c, d_cen.ensure((size_t)R * 12));
    CU_TRY(c, d_box.ensure((size_t)R * 24));
    vox_stats_kernel<<<(unsigned)R, kVoxStatsBlock, 0, st>>>((const float*)xyz_dev, d_idx[cur].as<int>(),
                                                                              d_rb.as<int>(), d_rc.as<int>(), R,
                                                                              d_cen.as<float>(), d_box.as<float>());
    ++c->launches;
    if (cudaGetLastError() != cudaSuccess) { delete res; c->last_error = "voxel clustering launch failed"; return PCSEG_ERR_CUDA; }
    std::vector<int> begin(R), count(R);
    res->inlier_indices.resize(total);
    res->centroids.resize(3 * (size_t)R);
    res->bboxes.resize(6 * (size_t)R);
    cudaMemcpyAsync(begin.data(), d_rb.p, (size_t)R * 4, cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(count.data(), d_rc.p, (size_t)R * 4, cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(res->inlier_indices.data(), d_idx[cur].p, (size_t)total * 4, cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(res->centroids.data(), d_cen.p, (size_t)R * 12, cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(res->bboxes.data(), d_box.p, (size_t)R * 24, cudaMemcpyDeviceToHost, st);
    if (cudaStreamSynchronize(st) != cudaSuccess) { delete res; c->last_error = cudaGetErrorString(cudaGetLastError()); return PCSEG_ERR_CUDA; }
    for (int r = 0; r < R; ++r) {
      res->inlier_offsets.push_back(begin[r] + count[r]);
      res->label_ids.push_back(initial_id_offset + r);
    }
  }
  if (lab_dev != labels) {
    rc = from_device(c, labels, lab_dev, (size_t)n * sizeof(int));
    if (rc == PCSEG_OK && cudaStreamSynchronize(st) != cudaSuccess) rc = PCSEG_ERR_CUDA;
    if (rc) { delete res; return rc; }
  }
  *out = res;
  return PCSEG_OK;
}

// ---- batch pipeline --------------------------------------------------------
int pcseg_batch_create(pcseg_ctx* c, int n_frames, int rows, int cols, pcseg_batch** out) {
  if (!c || !out) return PCSEG_ERR_NULL;
  if (n_frames <= 0 || rows <= 0 || cols <= 0) return PCSEG_ERR_DIMS;
  if ((size_t)n_frames * rows * cols > (size_t)INT_MAX) return PCSEG_ERR_UNSUPPORTED;  // packed result offsets are int32
  cudaSetDevice(c->device);
  pcseg_batch* b = new pcseg_batch();
  b->ctx = c; b->n_frames = n_frames; b->rows = rows; b->cols = cols; b->N = (size_t)rows * cols;
  int st = batch_alloc(b);
  if (st) { pcseg_batch_destroy(b); return st; }
  for (auto& e : b->ev) cudaEventCreate(&e);
  int prio_lo = 0, prio_hi = 0;   // numerically lower = higher priority
  cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
  const char* pe = getenv("PCSEG_STREAM_PRIORITIES");
  const bool prio = !(pe && pe[0] == '0') && prio_hi < prio_lo;
  if (cudaStreamCreateWithPriority(&b->own_stream, cudaStreamNonBlocking, prio ? prio_hi : prio_lo) != cudaSuccess ||
      cudaStreamCreateWithFlags(&b->own_copy_in, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&b->own_copy_out, cudaStreamNonBlocking) != cudaSuccess ||
      (prio && (cudaStreamCreateWithPriority(&b->own_normals, cudaStreamNonBlocking, prio_lo) != cudaSuccess ||
                cudaEventCreateWithFlags(&b->ev_nrm[0], cudaEventDisableTiming) != cudaSuccess ||
                cudaEventCreateWithFlags(&b->ev_nrm[1], cudaEventDisableTiming) != cudaSuccess))) {
    pcseg_batch_destroy(b);
    return PCSEG_ERR_CUDA;
  }
  *out = b;
  return PCSEG_OK;
}

void pcseg_batch_destroy(pcseg_batch* b) {
  if (!b) return;
  cudaSetDevice(b->ctx->device);
  cudaStreamSynchronize(b->ctx->stream);
  for (cudaStream_t s : {b->own_stream, b->own_copy_in, b->own_copy_out, b->own_normals})
    if (s) { cudaStreamSynchronize(s); cudaStreamDestroy(s); }
  for (auto& e : b->ev_nrm) if (e) cudaEventDestroy(e);
  DevBuf* all[] = {&b->xyz, &b->normals, &b->labels, &b->queue, &b->claim, &b->inl_planar, &b->inl_cluster,
                   &b->boundary, &b->bflags, &b->meta_planar, &b->meta_cluster, &b->stats, &b->inv_list,
                   &b->inv_count, &b->seeds, &b->n_seeds, &b->pk_offsets, &b->pk_planar, &b->pk_pinl,
                   &b->pk_pinl_xyz, &b->pk_bnd, &b->pk_bnd_xyz, &b->pk_bdisc, &b->pk_disc_idx, &b->pk_disc_pos,
                   &b->pk_hull, &b->pk_cluster, &b->pk_cinl, &b->pk_cinl_xyz, &b->pk_err};
  for (DevBuf* d : all) d->release();
  b->prof.release();
  if (b->h_meta) cudaFreeHost(b->h_meta);
  for (auto& e : b->ev) if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : b->chunk_ev) cudaEventDestroy(e);
  delete b;
}

namespace {
int enqueue_normals_stage(pcseg_batch* b, const float* xyz, float* normals_out);
int enqueue_growth_stage(pcseg_batch* b);
}  // namespace

int pcseg_batch_submit_regions(pcseg_batch* b, const float* xyz, const pcseg_pipeline_params* p, float* normals_out,
                               int* labels_out, pcseg_batch_results* results) {
  if (!b || !xyz || !p) return PCSEG_ERR_NULL;
  if (b->in_flight) {
    int st = pcseg_batch_wait(b);
    if (st) return st;
  }
  if (b->rows <= 1) return PCSEG_ERR_DIMS;
  int st = check_planar_cfg(&p->planar);
  if (st) return st;
  if (p->run_cluster_stage && (st = check_cluster_cfg(&p->cluster))) return st;
  if (p->planar_seed_mode == PCSEG_SEEDS_EXPLICIT && p->n_planar_seeds > 0 && !p->planar_seeds) return PCSEG_ERR_NULL;
  // the run's parameters are kept: pcseg_batch_wait may have to repeat the growth stages with a
  // larger region table (the reference has no region limit)
  b->run_params = *p;
  b->run_seeds.clear();
  if (p->planar_seed_mode == PCSEG_SEEDS_EXPLICIT && p->n_planar_seeds > 0) {
    if (is_device_ptr(p->planar_seeds)) b->run_params.planar_seeds = p->planar_seeds;
    else {
      b->run_seeds.assign(p->planar_seeds, p->planar_seeds + p->n_planar_seeds);
      b->run_params.planar_seeds = b->run_seeds.data();
    }
  }
  b->run_labels_out = labels_out;
  b->run_results = results;
  b->run_valid = true;
  cudaSetDevice(b->ctx->device);
  if ((st = enqueue_normals_stage(b, xyz, normals_out))) return st;
  if ((st = enqueue_growth_stage(b))) return st;
  b->in_flight = true;
  return PCSEG_OK;
}

int pcseg_batch_submit(pcseg_batch* b, const float* xyz, const pcseg_pipeline_params* p, float* normals_out,
                       int* labels_out) {
  return pcseg_batch_submit_regions(b, xyz, p, normals_out, labels_out, nullptr);
}

int pcseg_batch_wait(pcseg_batch* b) {
  if (!b) return PCSEG_ERR_NULL;
  pcseg_ctx* c = b->ctx;
  cudaSetDevice(c->device);
  if (!b->in_flight) return PCSEG_OK;
  b->in_flight = false;
  cudaStream_t stream = S(b);
  CU_TRY(c, cudaStreamSynchronize(stream));
  const int F = b->n_frames;
  PackTotals t = pack_totals(b, F);
  if ((t.err & kErrRegionCap) && b->rcap_planar < planar_region_bound(b, &b->run_params.planar)) {
    // more planes than the region table holds: repeat the growth stages with the full bound
    b->rcap_planar = planar_region_bound(b, &b->run_params.planar);
    CU_TRY(c, b->meta_planar.ensure((size_t)F * b->rcap_planar * sizeof(PlanarMeta)));
    int st = enqueue_growth_stage(b);
    if (st) return st;
    CU_TRY(c, cudaStreamSynchronize(stream));
    t = pack_totals(b, F);
  }
  bool fits = true;
  if (pcseg_batch_results* r = b->run_results) {
    r->n_planar = t.n_planar; r->n_planar_inliers = t.n_pinl; r->n_boundary = t.n_bnd;
    r->n_cluster = t.n_cluster; r->n_cluster_inliers = t.n_cinl; r->error_flags = t.err;
    // one copy per array, exact sizes (now known), on the workspace's stream
    cudaError_t e = cudaSuccess;
    const int F1 = F + 1;
    fits &= copy_result(b, r->frame_offsets, 5 * F1, b->pk_offsets.as<int>(), 5 * F1, &e);
    fits &= copy_result(b, r->frame_stats, 8 * F, b->stats.as<int>(), 8 * F, &e);
    fits &= copy_result(b, r->planar, r->planar_cap, b->pk_planar.as<pcseg_planar_record>(), t.n_planar, &e);
    fits &= copy_result(b, r->planar_inlier_indices, r->planar_inlier_cap, b->pk_pinl.as<int>(), t.n_pinl, &e);
    if (r->want_inlier_xyz)
      fits &= copy_result(b, r->planar_inlier_xyz, r->planar_inlier_cap, b->pk_pinl_xyz.as<float>(), t.n_pinl, &e, 3);
    fits &= copy_result(b, r->boundary_indices, r->boundary_cap, b->pk_bnd.as<int>(), t.n_bnd, &e);
    if (r->want_boundary_xyz)
      fits &= copy_result(b, r->boundary_xyz, r->boundary_cap, b->pk_bnd_xyz.as<float>(), t.n_bnd, &e, 3);
    fits &= copy_result(b, r->boundary_discontinuous, r->boundary_cap, b->pk_bdisc.as<unsigned char>(), t.n_bnd, &e);
    fits &= copy_result(b, r->discontinuous_indices, r->boundary_cap, b->pk_disc_idx.as<int>(), t.n_bnd, &e);
    fits &= copy_result(b, r->discontinuous_positions, r->boundary_cap, b->pk_disc_pos.as<int>(), t.n_bnd, &e);
    if (r->want_hull)
      fits &= copy_result(b, r->hull_points, r->boundary_cap, b->pk_hull.as<float>(), t.n_bnd, &e, 3);
    if (b->cluster_ran) {
      fits &= copy_result(b, r->cluster, r->cluster_cap, b->pk_cluster.as<pcseg_cluster_record>(), t.n_cluster, &e);
      fits &= copy_result(b, r->cluster_inlier_indices, r->cluster_inlier_cap, b->pk_cinl.as<int>(), t.n_cinl, &e);
      if (r->want_inlier_xyz)
        fits &= copy_result(b, r->cluster_inlier_xyz, r->cluster_inlier_cap, b->pk_cinl_xyz.as<float>(), t.n_cinl, &e, 3);
    }
    CU_TRY(c, e);
    CU_TRY(c, cudaStreamSynchronize(stream));
  }
  if (b->wait_copy_out) CU_TRY(c, cudaStreamSynchronize(SOut(b)));
  if (t.err) return map_error_flags(c, t.err);   // truncated label grid / region list: never silently OK
  if (!fits) {
    c->last_error = "a pcseg_batch_results buffer is smaller than the result (see the n_* totals)";
    return PCSEG_ERR_CAPACITY;
  }
  return PCSEG_OK;
}

int pcseg_batch_run(pcseg_batch* b, const float* xyz, const pcseg_pipeline_params* p, float* normals_out,
                    int* labels_out) {
  int st = pcseg_batch_submit(b, xyz, p, normals_out, labels_out);
  if (st) return st;
  return pcseg_batch_wait(b);
}

namespace {
// Normals of the whole batch.  Host buffers are processed in chunks of frames: chunk k's H2D
// (copy-in stream) overlaps the normals kernel of chunk k-1, and its normals D2H (copy-out
// stream) overlaps everything after.
int enqueue_normals_stage(pcseg_batch* b, const float* xyz, float* normals_out) {
  pcseg_ctx* c = b->ctx;
  const pcseg_pipeline_params* p = &b->run_params;
  cudaStream_t main_stream = S(b);
  if (b->own_normals) {   // behind everything already queued on the main stream
    CU_TRY(c, cudaEventRecord(b->ev_nrm[0], main_stream));
    CU_TRY(c, cudaStreamWaitEvent(b->own_normals, b->ev_nrm[0], 0));
    b->cur_stream = b->own_normals;
  }
  struct Restore { pcseg_batch* b; ~Restore() { b->cur_stream = nullptr; } } restore{b};
  cudaStream_t stream = S(b), s_in = SIn(b), s_out = SOut(b);
  const size_t F = b->n_frames, N = b->N;
  const bool xyz_host = !is_device_ptr(xyz);
  const bool nrm_host = normals_out && !is_device_ptr(normals_out);
  if (xyz_host) CU_TRY(c, b->xyz.ensure(F * N * 3 * sizeof(float)));
  b->xyz_dev = xyz_host ? b->xyz.as<float>() : xyz;
  float* nrm_dev;
  if (normals_out && !nrm_host) nrm_dev = normals_out;
  else {
    CU_TRY(c, b->normals.ensure(F * N * 3 * sizeof(float)));
    nrm_dev = b->normals.as<float>();
  }
  b->normals_dev = nrm_dev;
  b->wait_copy_out = false;
  cudaEventRecord(b->ev[0], stream);
  int st;
  if (!xyz_host && !nrm_host) {
    if ((st = run_normals(b, b->xyz_dev, nrm_dev, nullptr, &p->normals, p->sensor_translation, 0, b->rows, 0, b->cols))) return st;
  } else {
    const size_t chunk_frames = std::max<size_t>(1, std::min<size_t>(F, (size_t)(96u << 20) / (N * 12)));
    const size_t n_chunks = (F + chunk_frames - 1) / chunk_frames;
    while (b->chunk_ev.size() < 2 * n_chunks + 2) {
      cudaEvent_t e;
      CU_TRY(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      b->chunk_ev.push_back(e);
    }
    // the copy streams start after everything already queued on the main stream
    cudaEvent_t start = b->chunk_ev[2 * n_chunks];
    CU_TRY(c, cudaEventRecord(start, stream));
    CU_TRY(c, cudaStreamWaitEvent(s_in, start, 0));
    CU_TRY(c, cudaStreamWaitEvent(s_out, start, 0));
    for (size_t k = 0; k < n_chunks; ++k) {
      const size_t f0 = k * chunk_frames, nf = std::min(chunk_frames, F - f0);
      const size_t off = f0 * N * 3, cnt = nf * N * 3 * sizeof(float);
      if (xyz_host) {
        CU_TRY(c, cudaMemcpyAsync(b->xyz.as<float>() + off, xyz + off, cnt, cudaMemcpyHostToDevice, s_in));
        CU_TRY(c, cudaEventRecord(b->chunk_ev[2 * k], s_in));
        CU_TRY(c, cudaStreamWaitEvent(stream, b->chunk_ev[2 * k], 0));
      }
      if ((st = run_normals(b, b->xyz_dev, nrm_dev, nullptr, &p->normals, p->sensor_translation, 0, b->rows, 0,
                            b->cols, (int)f0, (int)nf))) return st;
      if (nrm_host) {
        CU_TRY(c, cudaEventRecord(b->chunk_ev[2 * k + 1], stream));
        CU_TRY(c, cudaStreamWaitEvent(s_out, b->chunk_ev[2 * k + 1], 0));
        CU_TRY(c, cudaMemcpyAsync(normals_out + off, nrm_dev + off, cnt, cudaMemcpyDeviceToHost, s_out));
        b->wait_copy_out = true;
      }
    }
  }
  cudaEventRecord(b->ev[1], stream);
  if (stream != main_stream) {   // the growth stages wait for the normals
    CU_TRY(c, cudaEventRecord(b->ev_nrm[1], stream));
    CU_TRY(c, cudaStreamWaitEvent(main_stream, b->ev_nrm[1], 0));
  }
  return PCSEG_OK;
}

// PrepareUnlabeledLabels -> seeds -> planar growth -> cluster growth -> labels out -> region pack.
// Everything after the normals; pcseg_batch_wait repeats it when the region table was too small.
int enqueue_growth_stage(pcseg_batch* b) {
  pcseg_ctx* c = b->ctx;
  const pcseg_pipeline_params* p = &b->run_params;
  cudaStream_t stream = S(b);
  const size_t F = b->n_frames, N = b->N;
  float* nrm_dev = const_cast<float*>(b->normals_dev);
  int st;
  if ((st = fill_int(c, b->labels.as<int>(), F * N, kUnlabeled, stream))) return st;  // PrepareUnlabeledLabels
  CU_TRY(c, cudaMemsetAsync(b->stats.p, 0, F * sizeof(FrameStats), stream));
  const int* seeds_dev = nullptr;
  int n_seeds = (int)N, stride = 0;
  const int* n_seeds_dev = nullptr;
  if (p->planar_seed_mode == PCSEG_SEEDS_PLANE_SUPPORT) {
    if ((st = run_seed_finder(b, b->xyz_dev, nrm_dev, 9, 0.05f, 12))) return st;  // ps/segmentation.h:193-194 defaults
    seeds_dev = b->seeds.as<int>(); stride = (int)N; n_seeds_dev = b->n_seeds.as<int>();
  } else if (p->planar_seed_mode == PCSEG_SEEDS_EXPLICIT) {
    if (p->n_planar_seeds > 0) {
      if ((st = check_seeds_host(p->planar_seeds, p->n_planar_seeds, N, c, &b->seeds, &seeds_dev, stream))) return st;
    }
    n_seeds = p->n_planar_seeds;
  }
  cudaEventRecord(b->ev[2], stream);
  bool fused = false;
  if (!(p->planar_seed_mode == PCSEG_SEEDS_EXPLICIT && n_seeds == 0)) {
    ClusterArgs ca;
    if (p->run_cluster_stage && (st = make_cluster_args(b, b->xyz_dev, nullptr, (int)N, &p->cluster, 0, 1, &ca))) return st;
    if ((st = run_planar(b, b->xyz_dev, nrm_dev, seeds_dev, n_seeds, stride, n_seeds_dev, &p->planar,
                         p->robot_q_xyzw, 0, p->run_cluster_stage ? &ca : nullptr, &fused))) return st;
  }
  cudaEventRecord(b->ev[3], stream);
  if (p->run_cluster_stage && !fused) {
    if ((st = run_cluster(b, b->xyz_dev, nullptr, (int)N, &p->cluster, 0, 1))) return st;
  }
  b->cluster_ran = p->run_cluster_stage != 0;
  cudaEventRecord(b->ev[4], stream);
  if (b->run_labels_out) {
    // label grid out on the copy-out stream, beside the region pack
    cudaStream_t s_out = SOut(b);
    while (b->chunk_ev.size() < 1) {
      cudaEvent_t e;
      CU_TRY(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      b->chunk_ev.push_back(e);
    }
    cudaEvent_t grown = b->chunk_ev.back();
    CU_TRY(c, cudaEventRecord(grown, stream));
    CU_TRY(c, cudaStreamWaitEvent(s_out, grown, 0));
    cudaMemcpyKind k = is_device_ptr(b->run_labels_out) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    CU_TRY(c, cudaMemcpyAsync(b->run_labels_out, b->labels.p, F * N * sizeof(int), k, s_out));
    b->wait_copy_out = true;
  }
  if (b->run_results) {
    PackWant w;
    w.inlier_xyz = b->run_results->want_inlier_xyz != 0;
    w.boundary_xyz = b->run_results->want_boundary_xyz != 0;
    w.hull = b->run_results->want_hull != 0;
    w.cluster_stats = b->run_results->want_cluster_stats != 0;
    w.cluster = b->cluster_ran;
    if ((st = run_pack(b, 0, (int)F, w))) return st;
  } else {
    // error flags travel with every run (pcseg_batch_wait must not return OK on a truncated result)
    if (!b->h_meta) CU_TRY(c, cudaMallocHost((void**)&b->h_meta, (5 * (F + 1) + 8 * F + 1) * sizeof(int)));
    std::memset(b->h_meta, 0, (5 * (F + 1) + 1) * sizeof(int));
    b->h_meta[5 * (F + 1) + 8 * F] = 0;
    CU_TRY(c, cudaMemcpyAsync(b->h_meta + 5 * (F + 1), b->stats.p, F * sizeof(FrameStats), cudaMemcpyDeviceToHost, stream));
  }
  cudaEventRecord(b->ev[5], stream);
  return PCSEG_OK;
}
}  // namespace

int pcseg_batch_frame_stats(pcseg_batch* b, int* stats8) {
  if (!b || !stats8) return PCSEG_ERR_NULL;
  pcseg_ctx* c = b->ctx;
  cudaSetDevice(c->device);
  if (int w = pcseg_batch_wait(b)) return w;
  static_assert(sizeof(FrameStats) == 8 * sizeof(int), "stats layout");
  CU_TRY(c, cudaMemcpyAsync(stats8, b->stats.p, (size_t)b->n_frames * sizeof(FrameStats), cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  return PCSEG_OK;
}

int pcseg_batch_regions(pcseg_batch* b, int frame, int kind, pcseg_regions** out) {
  if (!b || !out) return PCSEG_ERR_NULL;
  if (frame < 0 || frame >= b->n_frames) return PCSEG_ERR_DIMS;
  cudaSetDevice(b->ctx->device);
  if (int w = pcseg_batch_wait(b)) return w;
  pcseg_regions* r = new pcseg_regions();
  int st = kind == 0 ? collect_planar(b, frame, r) : collect_cluster(b, frame, r);
  if (st) { delete r; return st; }
  *out = r;
  return PCSEG_OK;
}

int pcseg_batch_stage_times(pcseg_batch* b, float* ms5) {
  if (!b || !ms5) return PCSEG_ERR_NULL;
  pcseg_ctx* c = b->ctx;
  cudaSetDevice(c->device);
  if (int w = pcseg_batch_wait(b)) return w;
  CU_TRY(c, cudaEventSynchronize(b->ev[5]));
  // normals (+fix-up) | label init + seed finder | planar growth | cluster growth | region pack
  for (int i = 0; i < 5; ++i) CU_TRY(c, cudaEventElapsedTime(ms5 + i, b->ev[i], b->ev[i + 1]));
  return PCSEG_OK;
}

int pcseg_batch_profile(pcseg_batch* b, long long* out16_per_frame) {
  if (!b || !out16_per_frame) return PCSEG_ERR_NULL;
  pcseg_ctx* c = b->ctx;
  if (!b->prof.p) return PCSEG_ERR_UNSUPPORTED;
  CU_TRY(c, cudaMemcpy(out16_per_frame, b->prof.p, (size_t)b->n_frames * 16 * sizeof(long long), cudaMemcpyDeviceToHost));
  CU_TRY(c, cudaMemset(b->prof.p, 0, (size_t)b->n_frames * 16 * sizeof(long long)));
  return PCSEG_OK;
}

void* pcseg_batch_stream(const pcseg_batch* b) { return b ? (void*)(b->own_stream ? b->own_stream : b->ctx->stream) : nullptr; }
int* pcseg_batch_device_labels(pcseg_batch* b) { return b ? b->labels.as<int>() : nullptr; }
float* pcseg_batch_device_normals(pcseg_batch* b) { return b ? const_cast<float*>(b->normals_dev) : nullptr; }

// ---- result export ------------------------------------------
int pcseg_transform_in_place(pcseg_ctx* c, float* points_xyz, float* normals_xyz, int n, const double q_xyzw[4],
                             const double translation[3]) {
  if (!c || !q_xyzw || !translation) return PCSEG_ERR_NULL;
  if (n < 0) return PCSEG_ERR_DIMS;
  if (n == 0 || (!points_xyz && !normals_xyz)) return PCSEG_OK;
  cudaSetDevice(c->device);
  const size_t bytes = (size_t)n * 3 * sizeof(float);
  TransformArgs a;
  a.n = n;
  // new_pose_current.cast<float>(), ps/multichannel_cloud.cc:55
  a.qx = (float)q_xyzw[0]; a.qy = (float)q_xyzw[1]; a.qz = (float)q_xyzw[2]; a.qw = (float)q_xyzw[3];
  a.tx = (float)translation[0]; a.ty = (float)translation[1]; a.tz = (float)translation[2];
  const bool p_dev = points_xyz && is_device_ptr(points_xyz), n_dev = normals_xyz && is_device_ptr(normals_xyz);
  a.points = points_xyz; a.normals = normals_xyz;
  if (points_xyz && !p_dev) {
    CU_TRY(c, c->tmp_a.ensure(bytes));
    CU_TRY(c, cudaMemcpyAsync(c->tmp_a.p, points_xyz, bytes, cudaMemcpyHostToDevice, c->stream));
    a.points = c->tmp_a.as<float>();
  }
  if (normals_xyz && !n_dev) {
    CU_TRY(c, c->tmp_b.ensure(bytes));
    CU_TRY(c, cudaMemcpyAsync(c->tmp_b.p, normals_xyz, bytes, cudaMemcpyHostToDevice, c->stream));
    a.normals = c->tmp_b.as<float>();
  }
  const int blocks = (int)std::min<size_t>(((size_t)n + 255) / 256, 148 * 16);
  transform_in_place_kernel<<<blocks, 256, 0, c->stream>>>(a);
  ++c->launches;
  CU_TRY(c, cudaGetLastError());
  if (points_xyz && !p_dev) CU_TRY(c, cudaMemcpyAsync(points_xyz, a.points, bytes, cudaMemcpyDeviceToHost, c->stream));
  if (normals_xyz && !n_dev) CU_TRY(c, cudaMemcpyAsync(normals_xyz, a.normals, bytes, cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  return PCSEG_OK;
}

int pcseg_gather_points(pcseg_ctx* c, const float* xyz, int n_points, const int* indices, int n_indices,
                        float* out_xyz, const int* discontinuous, int n_discontinuous, int* positions_out,
                        int* n_positions) {
  if (!c || !xyz || (n_indices > 0 && (!indices || !out_xyz))) return PCSEG_ERR_NULL;
  if (n_discontinuous > 0 && (!discontinuous || !positions_out || !n_positions)) return PCSEG_ERR_NULL;
  if (n_points < 0 || n_indices < 0 || n_discontinuous < 0) return PCSEG_ERR_DIMS;
  if (n_positions) *n_positions = 0;
  if (n_indices == 0) return PCSEG_OK;
  cudaSetDevice(c->device);
  if (!is_device_ptr(indices))
    for (int i = 0; i < n_indices; ++i)
      if (indices[i] < 0 || indices[i] >= n_points) return PCSEG_ERR_SEED_RANGE;  // AtUnsafe would read out of bounds
  if (n_discontinuous > 0 && !is_device_ptr(discontinuous))
    for (int i = 0; i < n_discontinuous; ++i)
      if (discontinuous[i] < 0 || discontinuous[i] >= n_points) return PCSEG_ERR_SEED_RANGE;
  DevBuf sx, si, so, sd, sm, sp, sn;
  const void *xyz_dev, *idx_dev;
  int st;
  if ((st = to_device(c, xyz, (size_t)n_points * 3 * sizeof(float), &sx, &xyz_dev))) return st;
  if ((st = to_device(c, indices, (size_t)n_indices * sizeof(int), &si, &idx_dev))) return st;
  const bool out_dev = is_device_ptr(out_xyz);
  float* out_d = out_xyz;
  if (!out_dev) {
    CU_TRY(c, so.ensure((size_t)n_indices * 3 * sizeof(float)));
    out_d = so.as<float>();
  }
  const int blocks = std::min((n_indices + 255) / 256, 148 * 16);
  gather_points_kernel<<<blocks, 256, 0, c->stream>>>((const float*)xyz_dev, (const int*)idx_dev, n_indices, out_d);
  ++c->launches;
  if (!out_dev) CU_TRY(c, cudaMemcpyAsync(out_xyz, out_d, (size_t)n_indices * 3 * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
  if (n_discontinuous > 0) {
    const void* disc_dev;
    if ((st = to_device(c, discontinuous, (size_t)n_discontinuous * sizeof(int), &sd, &disc_dev))) return st;
    CU_TRY(c, sm.ensure((size_t)n_points));
    CU_TRY(c, cudaMemsetAsync(sm.p, 0, (size_t)n_points, c->stream));
    CU_TRY(c, sp.ensure((size_t)n_indices * sizeof(int)));
    CU_TRY(c, sn.ensure(sizeof(int)));
    mark_cells_kernel<<<std::min((n_discontinuous + 255) / 256, 148 * 16), 256, 0, c->stream>>>(
        (const int*)disc_dev, n_discontinuous, sm.as<unsigned char>(), 1);
    positions_of_marked_kernel<<<1, 1024, 0, c->stream>>>((const int*)idx_dev, n_indices, sm.as<unsigned char>(),
                                                          sp.as<int>(), sn.as<int>());
    c->launches += 2;
    int np = 0;
    CU_TRY(c, cudaMemcpyAsync(&np, sn.p, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
    *n_positions = np;
    if (np > 0) {
      cudaMemcpyKind k = is_device_ptr(positions_out) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
      CU_TRY(c, cudaMemcpyAsync(positions_out, sp.p, (size_t)np * sizeof(int), k, c->stream));
    }
  }
  CU_TRY(c, cudaGetLastError());
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  return PCSEG_OK;
}

// ---- accessors ---------------------------------------------------------------
int pcseg_regions_count(const pcseg_regions* r) { return (int)r->label_ids.size(); }
const int* pcseg_regions_inlier_offsets(const pcseg_regions* r) { return r->inlier_offsets.data(); }
const int* pcseg_regions_inlier_indices(const pcseg_regions* r) { return r->inlier_indices.data(); }
const int* pcseg_regions_label_ids(const pcseg_regions* r) { return r->label_ids.data(); }
const int* pcseg_regions_seed_indices(const pcseg_regions* r) { return r->seed_indices.data(); }
const float* pcseg_regions_planes(const pcseg_regions* r) { return r->planes.data(); }
const float* pcseg_regions_centroids(const pcseg_regions* r) { return r->centroids.data(); }
const float* pcseg_regions_areas(const pcseg_regions* r) { return r->areas.data(); }
const float* pcseg_regions_estimators(const pcseg_regions* r) { return r->estimators.data(); }
const int* pcseg_regions_boundary_offsets(const pcseg_regions* r) { return r->boundary_offsets.data(); }
const int* pcseg_regions_boundary_indices(const pcseg_regions* r) { return r->boundary_indices.data(); }
const int* pcseg_regions_discontinuous_offsets(const pcseg_regions* r) { return r->disc_offsets.data(); }
const int* pcseg_regions_discontinuous_indices(const pcseg_regions* r) { return r->disc_indices.data(); }
const int* pcseg_regions_hull_offsets(const pcseg_regions* r) { return r->hull_offsets.data(); }
const float* pcseg_regions_hull_points(const pcseg_regions* r) { return r->hull_points.data(); }
const float* pcseg_regions_seed_points(const pcseg_regions* r) { return r->seed_points.data(); }
const float* pcseg_regions_bboxes(const pcseg_regions* r) { return r->bboxes.data(); }
void pcseg_regions_free(pcseg_regions* r) { delete r; }

}  // extern "C"
