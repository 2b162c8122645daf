// pcseg_api.cu -- C ABI (include/pcseg_b200.h) over the sm_100a kernels.
//
// Host-side plumbing only: argument checks that mirror the reference's CHECKs,
// device buffers, kernel launches on the context's stream and result
// materialisation.  No CPU implementation of any hot-path stage lives here:
// if CUDA is unavailable pcseg_create fails and nothing else can be called.
#include "../../include/pcseg_b200.h"

#include <algorithm>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "grow_cluster.cuh"
#include "planar_kernel.cuh"
#include "meanshift_grow.cuh"
#include "normals.cuh"
#include "seeds.cuh"

using namespace pcseg;

namespace {

constexpr int kGrowBlock = 256;
#ifndef PCSEG_CLUSTER_BLOCK
#define PCSEG_CLUSTER_BLOCK 512
#endif
constexpr int kClusterBlock = PCSEG_CLUSTER_BLOCK;  // queue entries evaluated per step of cluster growth
constexpr int kPipeBlock = 32 * kPipeWarps;  // planar fast path (grow_planar_pipe.cuh)
constexpr int kInvCap = 4096;

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  cudaError_t ensure(size_t need) {
    if (need <= bytes) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
    cudaError_t e = cudaMalloc(&p, need);
    if (e == cudaSuccess) bytes = need;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
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
  std::vector<cudaEvent_t> chunk_ev;  // 2 per chunk: H2D done, normals done
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
  bool in_flight = false, wait_copy_out = false;
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  DevBuf prof;
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
int to_device(pcseg_ctx* c, const void* src, size_t bytes, DevBuf* staging, const void** out) {
  if (is_device_ptr(src)) {
    *out = src;
    return PCSEG_OK;
  }
  CU_TRY(c, staging->ensure(bytes));
  CU_TRY(c, cudaMemcpyAsync(staging->p, src, bytes, cudaMemcpyHostToDevice, c->stream));
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

int fill_int(pcseg_ctx* c, int* p, size_t n, int v) {
  if (n == 0) return PCSEG_OK;
  int blocks = (int)std::min<size_t>((n + 255) / 256, 148 * 8);
  fill_int_kernel<<<blocks, 256, 0, c->stream>>>(p, n, v);
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
int check_cluster_cfg(const pcseg_cluster_config* cfg) {
  if (cfg->min_region_inliers < 1 || cfg->half_search_window < 0) return PCSEG_ERR_CONFIG;
  if (cfg->half_search_window > 3) return PCSEG_ERR_UNSUPPORTED;
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
  CU_TRY(c, cudaMemsetAsync(a.inv_count, 0, nf * sizeof(int), c->stream));
  const int work = (r1 - r0) * (c1 - c0);
  dim3 grid((work + kNormalsBlock - 1) / kNormalsBlock, nf);
  if (a.max2 <= FLT_MAX) normals_kernel<false><<<grid, kNormalsBlock, 0, c->stream>>>(a);
  else normals_kernel<true><<<grid, kNormalsBlock, 0, c->stream>>>(a);  // max distance inf/NaN: explicit allFinite() test
  normals_fixup_kernel<<<nf, 256, 0, c->stream>>>(a);
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
  // shared-memory fast path when the 1-bit/cell mask and the FIFO ring fit
  a.mask_words = (int)((b->N + 31) / 32);
  a.ring = 2048;
  const size_t smem_fixed = 32 * 11 * sizeof(float) + kHashSize * sizeof(unsigned);
  size_t smem = (size_t)a.mask_words * 4 + (size_t)a.ring * 20 + smem_fixed;
  cudaFuncAttributes fattr;
  CU_TRY(c, cudaFuncGetAttributes(&fattr, planar_grow_kernel<kPipeBlock, true, false>));
  const size_t smem_static = fattr.sharedSizeBytes;
  // two frames per SM when they fit (228 KB per SM, 1 KB reserved per CTA), else one
  const size_t smem_two = (228 * 1024) / 2 - 1024 - smem_static;
  const size_t smem_max = 227 * 1024 - smem_static;
  bool use_smem = b->N < (1u << 24) && (size_t)a.mask_words * 4 + 256 * 20 + smem_fixed <= smem_max &&
                  !getenv("PCSEG_FORCE_GLOBAL_PATH");
  if (use_smem) {
    if (smem > smem_two && (size_t)a.mask_words * 4 + 1024 * 20 + smem_fixed <= smem_two) a.ring = 1024;
    smem = (size_t)a.mask_words * 4 + (size_t)a.ring * 20 + smem_fixed;
    while (smem > smem_max) { a.ring /= 2; smem = (size_t)a.mask_words * 4 + (size_t)a.ring * 20 + smem_fixed; }
    if (fuse_cluster && getenv("PCSEG_FUSE")) {  // opt-in: measured slower (the extra code thrashes the instruction cache)
      CU_TRY(c, cudaFuncSetAttribute(planar_grow_kernel<kPipeBlock, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      planar_grow_kernel<kPipeBlock, true, true><<<b->n_frames, kPipeBlock, smem, c->stream>>>(a, *fuse_cluster);
      if (fused) *fused = true;
    } else {
      CU_TRY(c, cudaFuncSetAttribute(planar_grow_kernel<kPipeBlock, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      planar_grow_kernel<kPipeBlock, true, false><<<b->n_frames, kPipeBlock, smem, c->stream>>>(a, ClusterArgs{});
    }
  } else {
    planar_grow_kernel<kGrowBlock, false, false><<<b->n_frames, kGrowBlock, 0, c->stream>>>(a, ClusterArgs{});
  }
  ++c->launches;
  CU_TRY(c, cudaGetLastError());
  return PCSEG_OK;
}

int run_cluster(pcseg_batch* b, const float* xyz_dev, const int* seeds_dev, int n_seeds,
                const pcseg_cluster_config* cfg, int id_offset, int offset_from_planar) {
  pcseg_ctx* c = b->ctx;
  ClusterArgs a;
  int st = make_cluster_args(b, xyz_dev, seeds_dev, n_seeds, cfg, id_offset, offset_from_planar, &a);
  if (st) return st;
  if (a.half_window == 1) cluster_grow_kernel<kClusterBlock, true><<<b->n_frames, kClusterBlock, 0, c->stream>>>(a);
  else cluster_grow_kernel<kClusterBlock, false><<<b->n_frames, kClusterBlock, 0, c->stream>>>(a);
  ++c->launches;
  CU_TRY(c, cudaGetLastError());
  return PCSEG_OK;
}

int map_error_flags(pcseg_ctx* c, int flags) {
  if (flags == 0) return PCSEG_OK;
  c->last_error = "per-frame capacity/trace error flags = " + std::to_string(flags);
  return PCSEG_ERR_CAPACITY;
}

// Host-side PlanarConvexHull output for one accepted region (result formatting:
// the acceptance gate "hull has >= 3 vertices" already ran on the device).
// ps/algorithms.h:526-549 with patch P1; PoseFromPlane / ConvexHull restated
// (third-party, absent from the reference tree) as in the kernel's plane frame.
struct HV3 { float x, y, z; };
inline HV3 hcross(HV3 a, HV3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline HV3 hadd(HV3 a, HV3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline float hdot(HV3 a, HV3 b) { return a.x * b.x + (a.y * b.y + a.z * b.z); }
inline HV3 hrot(float qx, float qy, float qz, float qw, HV3 v) {
  HV3 qv{qx, qy, qz};
  HV3 uv = hcross(qv, v);
  uv = hadd(uv, uv);
  HV3 t{v.x + qw * uv.x, v.y + qw * uv.y, v.z + qw * uv.z};
  return hadd(t, hcross(qv, uv));
}
void host_planar_hull(const float* boundary_xyz, int nb, const float plane[4], std::vector<float>* out) {
  HV3 n{plane[0], plane[1], plane[2]};
  HV3 v1 = n;
  float z = hdot(n, n);
  if (z > 0.0f) { float s = sqrtf(z); v1 = {n.x / s, n.y / s, n.z / s}; }
  HV3 v0{0.0f, 0.0f, 1.0f};
  float c = hdot(v1, v0);
  float qx, qy, qz, qw;
  if (c < -1.0f + FLT_EPSILON) { qx = 1.0f; qy = qz = qw = 0.0f; }
  else {
    HV3 axis = hcross(v0, v1);
    float s = sqrtf((1.0f + c) * 2.0f);
    float invs = 1.0f / s;
    qx = axis.x * invs; qy = axis.y * invs; qz = axis.z * invs; qw = s * 0.5f;
  }
  HV3 t{-plane[3] * n.x, -plane[3] * n.y, -plane[3] * n.z};
  HV3 rt = hrot(-qx, -qy, -qz, qw, t);
  HV3 ti{-rt.x, -rt.y, -rt.z};
  struct P2 { double x, y; };
  std::vector<P2> pts(nb);
  for (int i = 0; i < nb; ++i) {
    HV3 p{boundary_xyz[3 * i], boundary_xyz[3 * i + 1], boundary_xyz[3 * i + 2]};
    HV3 pp = hadd(hrot(-qx, -qy, -qz, qw, p), ti);
    pts[i] = {pp.x, pp.y};
  }
  std::sort(pts.begin(), pts.end(), [](const P2& a, const P2& b) { return a.x < b.x || (a.x == b.x && a.y < b.y); });
  pts.erase(std::unique(pts.begin(), pts.end(), [](const P2& a, const P2& b) { return a.x == b.x && a.y == b.y; }), pts.end());
  std::vector<P2> h;
  int m = (int)pts.size();
  if (m < 3) h = pts;
  else {
    h.resize(2 * m);
    int k = 0;
    auto cr = [](const P2& o, const P2& a, const P2& b) { return (a.x - o.x) * (b.y - o.y) - (a.y - o.y) * (b.x - o.x); };
    for (int i = 0; i < m; ++i) { while (k >= 2 && cr(h[k - 2], h[k - 1], pts[i]) <= 0) --k; h[k++] = pts[i]; }
    for (int i = m - 2, t0 = k + 1; i >= 0; --i) { while (k >= t0 && cr(h[k - 2], h[k - 1], pts[i]) <= 0) --k; h[k++] = pts[i]; }
    h.resize(k - 1);
  }
  for (const P2& p : h) {
    HV3 pp{(float)p.x, (float)p.y, 0.0f};
    HV3 w = hadd(hrot(qx, qy, qz, qw, pp), t);
    out->push_back(w.x); out->push_back(w.y); out->push_back(w.z);
  }
}

}  // namespace

// ===========================================================================
// Region materialisation
// ===========================================================================
namespace {

int collect_planar(pcseg_batch* b, int frame, pcseg_regions* r) {
  pcseg_ctx* c = b->ctx;
  FrameStats st;
  CU_TRY(c, cudaMemcpyAsync(&st, b->stats.as<FrameStats>() + frame, sizeof(st), cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  const int n = st.n_planar;
  std::vector<PlanarMeta> meta(n);
  std::vector<int> inl(st.planar_inliers), bnd(st.boundary_len);
  std::vector<unsigned char> bfl(st.boundary_len);
  if (n) CU_TRY(c, cudaMemcpyAsync(meta.data(), b->meta_planar.as<PlanarMeta>() + (size_t)frame * b->rcap_planar,
                                   n * sizeof(PlanarMeta), cudaMemcpyDeviceToHost, c->stream));
  if (!inl.empty()) CU_TRY(c, cudaMemcpyAsync(inl.data(), b->inl_planar.as<int>() + (size_t)frame * b->N,
                                              inl.size() * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  if (!bnd.empty()) {
    CU_TRY(c, cudaMemcpyAsync(bnd.data(), b->boundary.as<int>() + (size_t)frame * b->bcap,
                              bnd.size() * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CU_TRY(c, cudaMemcpyAsync(bfl.data(), b->bflags.as<unsigned char>() + (size_t)frame * b->bcap, bfl.size(),
                              cudaMemcpyDeviceToHost, c->stream));
  }
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  // boundary xyz for the hull output: gather on the host from a device->host copy
  // of just the boundary points (indexed copy kernel avoided: tiny data).
  std::vector<float> bxyz(3 * bnd.size());
  {
    // one strided gather via cudaMemcpy2D is not possible for arbitrary indices;
    // fetch the frame's points once if they are not host-visible.
    std::vector<float> frame_xyz(3 * b->N);
    CU_TRY(c, cudaMemcpyAsync(frame_xyz.data(), b->xyz_dev + (size_t)frame * b->N * 3, frame_xyz.size() * sizeof(float),
                              cudaMemcpyDeviceToHost, c->stream));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
    for (size_t i = 0; i < bnd.size(); ++i)
      for (int k = 0; k < 3; ++k) bxyz[3 * i + k] = frame_xyz[3 * (size_t)bnd[i] + k];
  }
  for (int i = 0; i < n; ++i) {
    const PlanarMeta& m = meta[i];
    r->inlier_indices.insert(r->inlier_indices.end(), inl.begin() + m.inlier_begin,
                             inl.begin() + m.inlier_begin + m.inlier_count);
    r->inlier_offsets.push_back((int)r->inlier_indices.size());
    r->label_ids.push_back(m.label_id);
    r->seed_indices.push_back(m.seed_index);
    r->planes.insert(r->planes.end(), m.plane, m.plane + 4);
    r->centroids.insert(r->centroids.end(), m.centroid, m.centroid + 3);
    r->areas.push_back(m.area);
    r->estimators.insert(r->estimators.end(), m.est, m.est + 10);
    r->boundary_indices.insert(r->boundary_indices.end(), bnd.begin() + m.boundary_begin,
                               bnd.begin() + m.boundary_begin + m.boundary_count);
    r->boundary_offsets.push_back((int)r->boundary_indices.size());
    std::vector<int> disc;
    for (int k = 0; k < m.boundary_count; ++k)
      if (bfl[m.boundary_begin + k]) disc.push_back(bnd[m.boundary_begin + k]);
    std::sort(disc.begin(), disc.end());
    disc.erase(std::unique(disc.begin(), disc.end()), disc.end());
    r->disc_indices.insert(r->disc_indices.end(), disc.begin(), disc.end());
    r->disc_offsets.push_back((int)r->disc_indices.size());
    host_planar_hull(bxyz.data() + 3 * (size_t)m.boundary_begin, m.boundary_count, m.plane, &r->hull_points);
    r->hull_offsets.push_back((int)(r->hull_points.size() / 3));
  }
  return map_error_flags(c, st.error_flags);
}

int collect_cluster(pcseg_batch* b, int frame, pcseg_regions* r) {
  pcseg_ctx* c = b->ctx;
  FrameStats st;
  CU_TRY(c, cudaMemcpyAsync(&st, b->stats.as<FrameStats>() + frame, sizeof(st), cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  const int n = st.n_cluster;
  std::vector<ClusterMeta> meta(n);
  std::vector<int> inl(st.cluster_inliers);
  if (n) CU_TRY(c, cudaMemcpyAsync(meta.data(), b->meta_cluster.as<ClusterMeta>() + (size_t)frame * b->rcap_cluster,
                                   n * sizeof(ClusterMeta), cudaMemcpyDeviceToHost, c->stream));
  if (!inl.empty()) CU_TRY(c, cudaMemcpyAsync(inl.data(), b->inl_cluster.as<int>() + (size_t)frame * b->N,
                                              inl.size() * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  for (int i = 0; i < n; ++i) {
    const ClusterMeta& m = meta[i];
    r->inlier_indices.insert(r->inlier_indices.end(), inl.begin() + m.inlier_begin,
                             inl.begin() + m.inlier_begin + m.inlier_count);
    r->inlier_offsets.push_back((int)r->inlier_indices.size());
    r->label_ids.push_back(m.label_id);
  }
  return map_error_flags(c, st.error_flags);
}

int get_scratch(pcseg_ctx* c, int rows, int cols, pcseg_batch** out) {
  if (c->scratch && (c->scratch->rows != rows || c->scratch->cols != cols)) {
    pcseg_batch_destroy(c->scratch);
    c->scratch = nullptr;
  }
  if (!c->scratch) {
    int st = pcseg_batch_create(c, 1, rows, cols, &c->scratch);
    if (st) return st;
  }
  *out = c->scratch;
  return PCSEG_OK;
}

int check_seeds_host(const int* seeds_host_or_dev, int n, size_t N, pcseg_ctx* c, DevBuf* staging, const int** dev) {
  // Validates the range on the host copy (the reference would read out of
  // bounds); device-resident seeds are trusted.
  if (n == 0) { *dev = (const int*)staging->p; return PCSEG_OK; }
  if (!is_device_ptr(seeds_host_or_dev)) {
    for (int i = 0; i < n; ++i)
      if (seeds_host_or_dev[i] < 0 || (size_t)seeds_host_or_dev[i] >= N) return PCSEG_ERR_SEED_RANGE;
  }
  const void* d;
  int st = to_device(c, seeds_host_or_dev, (size_t)n * sizeof(int), staging, &d);
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
  *out = c;
  return PCSEG_OK;
}

void pcseg_destroy(pcseg_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->scratch) pcseg_batch_destroy(c->scratch);
  c->tmp_a.release();
  c->tmp_b.release();
  for (cudaEvent_t e : c->chunk_ev) cudaEventDestroy(e);
  if (c->copy_in) cudaStreamDestroy(c->copy_in);
  if (c->copy_out) cudaStreamDestroy(c->copy_out);
  if (c->stream) cudaStreamDestroy(c->stream);
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
  CU_TRY(c, cudaMemcpyAsync(b->labels.p, labels, N * sizeof(int),
                            is_device_ptr(labels) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, c->stream));
  CU_TRY(c, cudaMemsetAsync(b->stats.p, 0, sizeof(FrameStats), c->stream));
  if (n_seeds > 0) {
    const int* seeds_dev;
    if ((st = check_seeds_host(seeds, n_seeds, N, c, &b->seeds, &seeds_dev))) return st;
    if ((st = run_planar(b, b->xyz_dev, b->normals_dev, seeds_dev, n_seeds, 0, nullptr, cfg, robot_q_xyzw, id_offset))) return st;
  }
  if ((st = from_device(c, labels, b->labels.p, N * sizeof(int)))) return st;
  pcseg_regions* r = new pcseg_regions();
  st = collect_planar(b, 0, r);
  if (st) { delete r; return st; }
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
    if ((st = check_seeds_host(seeds, n_seeds, N, c, &b->seeds, &seeds_dev))) return st;
    if ((st = run_cluster(b, b->xyz_dev, seeds_dev, n_seeds, cfg, id_offset, 0))) return st;
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
  seed_support_count_kernel<<<grid, 256, 0, c->stream>>>(a);
  {
    const int nkeys = nbhd * nbhd + 1;
    const size_t smem = (size_t)kSeedWarps * nkeys * sizeof(int);
    if (smem > 48 * 1024)
      CU_TRY(c, cudaFuncSetAttribute(seed_order_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    seed_order_kernel<<<(unsigned)F, 32 * kSeedWarps, smem, c->stream>>>(a, nkeys);
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
  meanshift_modes_kernel<<<(unsigned)((M + 255) / 256), 256, 0, c->stream>>>(a);
  ++c->launches;
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && !out_dev)
    e = cudaMemcpyAsync(modes_out, modes_dev, M * sizeof(ModeRecord), cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  sx.release(); sl.release(); so.release();
  if (e != cudaSuccess) { c->last_error = cudaGetErrorString(e); return PCSEG_ERR_CUDA; }
  return PCSEG_OK;
}

int pcseg_mean_shift_modes(pcseg_ctx* c, const float* xyz, int rows, int cols, const int* labels,
                           int iterations, float* modes_out) {
  if (rows <= 0 || cols <= 0) return c ? PCSEG_ERR_DIMS : PCSEG_ERR_NULL;
  return pcseg_mean_shift_modes_range(c, xyz, rows, cols, labels, iterations, 0, rows * cols, modes_out);
}

namespace {
// Sort + growth phase on device-resident mode records (ps/mean_shift_segmentation.h:262-329).
int mean_shift_finish(pcseg_ctx* c, const float* xyz_dev, int rows, int cols, const ModeRecord* modes_dev,
                      const pcseg_cluster_config* cfg, int id_offset, int sort_mode, int* labels_dev,
                      pcseg_regions** out) {
  const size_t N = (size_t)rows * cols;
  // (1) intensities of the valid seeds, in seed order (ascending cloud index, :219-229,267-274)
  DevBuf d_int;
  CU_TRY(c, d_int.ensure(N * sizeof(float)));
  ms_intensity_kernel<<<(unsigned)((N + 255) / 256), 256, 0, c->stream>>>(modes_dev, (int)N, d_int.as<float>());
  ++c->launches;
  std::vector<float> h_int(N);
  CU_TRY(c, cudaMemcpyAsync(h_int.data(), d_int.p, N * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  d_int.release();
  // (2) the reference's own sort call on (intensity) keys, :276-280.  The permutation of an
  // unstable sort depends only on the comparison outcomes, so sorting (cell, intensity) pairs
  // with the same comparator reproduces the permutation of the reference's records (for the
  // same standard library; quirk Q7).  sort_mode 1 = stable (documented patch P2).
  std::vector<std::pair<int, double>> keys;
  keys.reserve(N);
  for (size_t i = 0; i < N; ++i)
    if (h_int[i] >= 0.0f) keys.emplace_back((int)i, (double)h_int[i]);
  auto cmp = [](const std::pair<int, double>& l, const std::pair<int, double>& r) { return l.second < r.second; };
  if (sort_mode == 0) std::sort(keys.begin(), keys.end(), cmp);
  else std::stable_sort(keys.begin(), keys.end(), cmp);
  const int S = (int)keys.size();
  pcseg_regions* res = new pcseg_regions();
  *out = res;
  if (S == 0) return PCSEG_OK;
  std::vector<int> h_cells(S);
  for (int i = 0; i < S; ++i) h_cells[i] = keys[i].first;
  // (3) growth on the device
  const int rcap = (int)std::min<size_t>(N / (size_t)std::max(cfg->min_region_inliers, 1) + 1, (size_t)1 << 22);
  DevBuf d_cells, d_seeds, d_claim, d_inl, d_meta, d_out;
  CU_TRY(c, d_cells.ensure((size_t)S * sizeof(int)));
  CU_TRY(c, d_seeds.ensure((size_t)S * sizeof(MsSeedRec)));
  CU_TRY(c, d_claim.ensure(N * sizeof(int)));
  CU_TRY(c, d_inl.ensure(N * sizeof(int)));
  CU_TRY(c, d_meta.ensure((size_t)rcap * sizeof(MsRegionMeta)));
  CU_TRY(c, d_out.ensure(4 * sizeof(int)));
  CU_TRY(c, cudaMemcpyAsync(d_cells.p, h_cells.data(), (size_t)S * sizeof(int), cudaMemcpyHostToDevice, c->stream));
  CU_TRY(c, cudaMemsetAsync(d_out.p, 0, 4 * sizeof(int), c->stream));
  ms_gather_kernel<<<(unsigned)((S + 255) / 256), 256, 0, c->stream>>>(modes_dev, d_cells.as<int>(), S, rows,
                                                                        d_seeds.as<MsSeedRec>());
  fill_int_kernel<<<(unsigned)std::min<size_t>((N + 255) / 256, 65535), 256, 0, c->stream>>>(d_claim.as<int>(), N, INT_MAX);
  MsGrowArgs g;
  g.xyz = xyz_dev; g.labels = labels_dev; g.claim = d_claim.as<int>(); g.seeds = d_seeds.as<MsSeedRec>();
  g.n_seeds = S; g.rows = rows; g.cols = cols; g.half_window = cfg->half_search_window;
  g.min_inliers = cfg->min_region_inliers; g.id_offset = id_offset; g.inliers = d_inl.as<int>();
  g.meta = d_meta.as<MsRegionMeta>(); g.rcap = rcap; g.out = d_out.as<int>();
  ms_grow_kernel<<<1, kMsGrowBlock, 0, c->stream>>>(g);
  c->launches += 3;
  CU_TRY(c, cudaGetLastError());
  int h_out[4];
  CU_TRY(c, cudaMemcpyAsync(h_out, d_out.p, sizeof(h_out), cudaMemcpyDeviceToHost, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
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
  d_cells.release(); d_seeds.release(); d_claim.release(); d_inl.release(); d_meta.release(); d_out.release();
  if (h_out[1]) { c->last_error = "mean-shift region table overflow"; return PCSEG_ERR_CAPACITY; }
  return PCSEG_OK;
}
}  // namespace

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
  MeanShiftArgs a;
  a.xyz = (const float*)xyz_dev; a.labels = lab_dev; a.rows = rows; a.cols = cols;
  a.iterations = iterations; a.cell_begin = 0; a.cell_end = (int)N; a.modes = sm.as<ModeRecord>();
  meanshift_modes_kernel<<<(unsigned)((N + 255) / 256), 256, 0, c->stream>>>(a);
  ++c->launches;
  st = mean_shift_finish(c, (const float*)xyz_dev, rows, cols, sm.as<ModeRecord>(), cfg, id_offset, sort_mode,
                         lab_dev, out);
  if (st == PCSEG_OK && lab_dev != labels) {
    st = from_device(c, labels, lab_dev, N * sizeof(int));
    if (st == PCSEG_OK && cudaStreamSynchronize(c->stream) != cudaSuccess) st = PCSEG_ERR_CUDA;
  }
  sx.release(); sl.release(); sm.release();
  return st;
}

// ---- batch pipeline --------------------------------------------------------
int pcseg_batch_create(pcseg_ctx* c, int n_frames, int rows, int cols, pcseg_batch** out) {
  if (!c || !out) return PCSEG_ERR_NULL;
  if (n_frames <= 0 || rows <= 0 || cols <= 0) return PCSEG_ERR_DIMS;
  cudaSetDevice(c->device);
  pcseg_batch* b = new pcseg_batch();
  b->ctx = c; b->n_frames = n_frames; b->rows = rows; b->cols = cols; b->N = (size_t)rows * cols;
  int st = batch_alloc(b);
  if (st) { pcseg_batch_destroy(b); return st; }
  for (auto& e : b->ev) cudaEventCreate(&e);
  if (cudaStreamCreateWithFlags(&b->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&b->own_copy_in, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&b->own_copy_out, cudaStreamNonBlocking) != cudaSuccess) {
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
  for (cudaStream_t s : {b->own_stream, b->own_copy_in, b->own_copy_out})
    if (s) { cudaStreamSynchronize(s); cudaStreamDestroy(s); }
  DevBuf* all[] = {&b->xyz, &b->normals, &b->labels, &b->queue, &b->claim, &b->inl_planar, &b->inl_cluster,
                   &b->boundary, &b->bflags, &b->meta_planar, &b->meta_cluster, &b->stats, &b->inv_list,
                   &b->inv_count, &b->seeds, &b->n_seeds};
  for (DevBuf* d : all) d->release();
  b->prof.release();
  for (auto& e : b->ev) if (e) cudaEventDestroy(e);
  delete b;
}

namespace {
// Routes everything a run enqueues to the workspace's own streams (calls on one context are
// serialised by the caller, so swapping the context's current streams is safe).
struct StreamScope {
  pcseg_ctx* c;
  cudaStream_t s, ci, co;
  StreamScope(pcseg_ctx* ctx, pcseg_batch* b) : c(ctx), s(ctx->stream), ci(ctx->copy_in), co(ctx->copy_out) {
    if (b->own_stream) { c->stream = b->own_stream; c->copy_in = b->own_copy_in; c->copy_out = b->own_copy_out; }
  }
  ~StreamScope() { c->stream = s; c->copy_in = ci; c->copy_out = co; }
};
int batch_enqueue(pcseg_batch* b, const float* xyz, const pcseg_pipeline_params* p, float* normals_out,
                  int* labels_out);
}  // namespace

int pcseg_batch_submit(pcseg_batch* b, const float* xyz, const pcseg_pipeline_params* p, float* normals_out,
                       int* labels_out) {
  if (!b || !xyz || !p) return PCSEG_ERR_NULL;
  if (b->in_flight) {
    int st = pcseg_batch_wait(b);
    if (st) return st;
  }
  StreamScope scope(b->ctx, b);
  int st = batch_enqueue(b, xyz, p, normals_out, labels_out);
  if (st == PCSEG_OK) b->in_flight = true;
  return st;
}

int pcseg_batch_wait(pcseg_batch* b) {
  if (!b) return PCSEG_ERR_NULL;
  pcseg_ctx* c = b->ctx;
  cudaSetDevice(c->device);
  if (!b->in_flight) return PCSEG_OK;
  b->in_flight = false;
  CU_TRY(c, cudaStreamSynchronize(b->own_stream ? b->own_stream : c->stream));
  if (b->wait_copy_out) CU_TRY(c, cudaStreamSynchronize(b->own_copy_out ? b->own_copy_out : c->copy_out));
  return PCSEG_OK;
}

int pcseg_batch_run(pcseg_batch* b, const float* xyz, const pcseg_pipeline_params* p, float* normals_out,
                    int* labels_out) {
  int st = pcseg_batch_submit(b, xyz, p, normals_out, labels_out);
  if (st) return st;
  return pcseg_batch_wait(b);
}

namespace {
int batch_enqueue(pcseg_batch* b, const float* xyz, const pcseg_pipeline_params* p, float* normals_out,
                  int* labels_out) {
  pcseg_ctx* c = b->ctx;
  if (b->rows <= 1) return PCSEG_ERR_DIMS;
  int st = check_planar_cfg(&p->planar);
  if (st) return st;
  if (p->run_cluster_stage && (st = check_cluster_cfg(&p->cluster))) return st;
  cudaSetDevice(c->device);
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
  cudaEventRecord(b->ev[0], c->stream);
  if (!xyz_host && !nrm_host) {
    if ((st = run_normals(b, b->xyz_dev, nrm_dev, nullptr, &p->normals, p->sensor_translation, 0, b->rows, 0, b->cols))) return st;
  } else {
    // Host buffers: chunks of frames; chunk k's H2D (copy_in stream) overlaps the normals
    // kernel of chunk k-1, and its normals D2H (copy_out stream) overlaps everything after.
    const size_t chunk_frames = std::max<size_t>(1, std::min<size_t>(F, (size_t)(96u << 20) / (N * 12) ));
    const size_t n_chunks = (F + chunk_frames - 1) / chunk_frames;
    while (c->chunk_ev.size() < 2 * n_chunks + 1) {
      cudaEvent_t e;
      CU_TRY(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      c->chunk_ev.push_back(e);
    }
    // the copy streams start after everything already queued on the main stream
    cudaEvent_t start = c->chunk_ev[2 * n_chunks];
    CU_TRY(c, cudaEventRecord(start, c->stream));
    CU_TRY(c, cudaStreamWaitEvent(c->copy_in, start, 0));
    CU_TRY(c, cudaStreamWaitEvent(c->copy_out, start, 0));
    for (size_t k = 0; k < n_chunks; ++k) {
      const size_t f0 = k * chunk_frames, nf = std::min(chunk_frames, F - f0);
      const size_t off = f0 * N * 3, cnt = nf * N * 3 * sizeof(float);
      if (xyz_host) {
        CU_TRY(c, cudaMemcpyAsync(b->xyz.as<float>() + off, xyz + off, cnt, cudaMemcpyHostToDevice, c->copy_in));
        CU_TRY(c, cudaEventRecord(c->chunk_ev[2 * k], c->copy_in));
        CU_TRY(c, cudaStreamWaitEvent(c->stream, c->chunk_ev[2 * k], 0));
      }
      if ((st = run_normals(b, b->xyz_dev, nrm_dev, nullptr, &p->normals, p->sensor_translation, 0, b->rows, 0,
                            b->cols, (int)f0, (int)nf))) return st;
      if (nrm_host) {
        CU_TRY(c, cudaEventRecord(c->chunk_ev[2 * k + 1], c->stream));
        CU_TRY(c, cudaStreamWaitEvent(c->copy_out, c->chunk_ev[2 * k + 1], 0));
        CU_TRY(c, cudaMemcpyAsync(normals_out + off, nrm_dev + off, cnt, cudaMemcpyDeviceToHost, c->copy_out));
      }
    }
  }
  cudaEventRecord(b->ev[1], c->stream);
  if ((st = fill_int(c, b->labels.as<int>(), F * N, kUnlabeled))) return st;  // PrepareUnlabeledLabels
  CU_TRY(c, cudaMemsetAsync(b->stats.p, 0, F * sizeof(FrameStats), c->stream));
  const int* seeds_dev = nullptr;
  int n_seeds = (int)N, stride = 0;
  const int* n_seeds_dev = nullptr;
  if (p->planar_seed_mode == PCSEG_SEEDS_PLANE_SUPPORT) {
    if ((st = run_seed_finder(b, b->xyz_dev, nrm_dev, 9, 0.05f, 12))) return st;  // ps/segmentation.h:193-194 defaults
    seeds_dev = b->seeds.as<int>(); stride = (int)N; n_seeds_dev = b->n_seeds.as<int>();
  } else if (p->planar_seed_mode == PCSEG_SEEDS_EXPLICIT) {
    if (p->n_planar_seeds > 0 && !p->planar_seeds) return PCSEG_ERR_NULL;
    if (p->n_planar_seeds > 0) {
      if ((st = check_seeds_host(p->planar_seeds, p->n_planar_seeds, N, c, &b->seeds, &seeds_dev))) return st;
    }
    n_seeds = p->n_planar_seeds;
  }
  cudaEventRecord(b->ev[2], c->stream);
  bool fused = false;
  if (!(p->planar_seed_mode == PCSEG_SEEDS_EXPLICIT && n_seeds == 0)) {
    ClusterArgs ca;
    if (p->run_cluster_stage && (st = make_cluster_args(b, b->xyz_dev, nullptr, (int)N, &p->cluster, 0, 1, &ca))) return st;
    if ((st = run_planar(b, b->xyz_dev, nrm_dev, seeds_dev, n_seeds, stride, n_seeds_dev, &p->planar,
                         p->robot_q_xyzw, 0, p->run_cluster_stage ? &ca : nullptr, &fused))) return st;
  }
  cudaEventRecord(b->ev[3], c->stream);
  if (p->run_cluster_stage && !fused) {
    if ((st = run_cluster(b, b->xyz_dev, nullptr, (int)N, &p->cluster, 0, 1))) return st;
  }
  cudaEventRecord(b->ev[4], c->stream);
  if (labels_out && (st = from_device(c, labels_out, b->labels.p, F * N * sizeof(int)))) return st;
  b->wait_copy_out = nrm_host;
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

int pcseg_batch_stage_times(pcseg_batch* b, float* ms4) {
  if (!b || !ms4) return PCSEG_ERR_NULL;
  pcseg_ctx* c = b->ctx;
  cudaSetDevice(c->device);
  if (int w = pcseg_batch_wait(b)) return w;
  CU_TRY(c, cudaEventSynchronize(b->ev[4]));
  // normals (+fix-up) | label init + seed finder | planar growth | cluster growth
  for (int i = 0; i < 4; ++i) CU_TRY(c, cudaEventElapsedTime(ms4 + i, b->ev[i], b->ev[i + 1]));
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
void pcseg_regions_free(pcseg_regions* r) { delete r; }

}  // extern "C"
