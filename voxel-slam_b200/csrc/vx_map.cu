// vx_map.cu — the sliding-window voxel map on the device:
//   cut_voxel_multi / OctoTree::allocate / push / Bf_var   voxel_map.hpp:1964-2096, :1204-1237, :1105-1143, :106-121
//   cut_voxel (fixed points) / allocate_fix / push_fix*    voxel_map.hpp:2108-2152, :1239-1264, :1149-1178, :1270-1299
//   OctoTree::recut / subdivide / plane_judge               voxel_map.hpp:1396-1456, :1307-1338, :1185-1195
//   OctoTree::tras_opt / LidarFactor::push_voxel            voxel_map.hpp:1605-1638, :139-147
//   OctoTree::margi / plane_update                          voxel_map.hpp:1465-1598, :1344-1388
//   drivers multi_recut / multi_margi                       voxelslam.cpp:1682-1737, :1590-1679
//
// Data layout (HBM).  Nodes are rows of a growing table; a root-voxel hash (packed key -> node) sits on top:
//   nodes[]   vx_map_node (408 B)  what OctoTree::match reads: centre, quater_length, octo_state, children, plane
//   meta[]    NodeMeta             root, path, layer, isexist, sw != nullptr, slide-set membership, last_num, opt_state
//   add[] fix[]  Mom (80 B)        pcr_add / pcr_fix as 6 unique second moments + 3 first moments + N
//   loc[node*W + slot]  Mom        SlideWindow::pcrs_local
//   cov[node*81]        double     cov_add (9x9)
//   eig[node*12]        double     eig_value (3) + eig_vector (9, column-major)
// The raw points the reference keeps in per-leaf vectors (SlideWindow::points, OctoTree::point_fix) stay where they
// arrived: one scan buffer per physical window slot and one append-only pool of fixed points, each with an int32 leaf id
// per point.  Insertion = key -> stable sort -> segment -> ordered sums; subdivision = re-label + the same segmented sum;
// marginalisation = a per-leaf kernel + a compaction of the oldest slot into the fixed pool.  Node ids are assigned by
// prefix sums over sorted segments, so a run is reproducible bit for bit; every per-leaf sum runs over its points in the
// reference's own order (fixed points first, then window slots oldest to newest, scan order inside a slot).
#include "vx_factor.cuh"
#include "vx_math.cuh"
#include "vx_planemap.cuh"
#include "vx_voxelize.cuh"
#include "vx_frontend.cuh"
#include <cub/cub.cuh>
#include <algorithm>

using namespace vx;

namespace {

struct Mom { double P[6]; double v[3]; int32_t N; int32_t pad; };   // P: xx xy xz yy yz zz
static_assert(sizeof(Mom) == 80, "Mom layout");

struct NodeMeta {
  uint64_t root_key;
  int32_t root, path, layer;
  int32_t isexist, has_sw, in_slide, last_num, opt_state;
  double jour;
};

constexpr uint64_t CODE_CREATE = 1ull << 40;   // target codes: [0, 2^40) existing leaf id; CREATE | (parent*8 + child)
constexpr uint64_t CODE_SKIP = ~0ull;
constexpr int MAX_W = 32;
constexpr int MAX_MG = 4;                      // frames marginalised per call (the reference uses 1)

struct PoseSet { double R[MAX_W][9]; double p[MAX_W][3]; };   // column-major R, kernel argument (<= 4 KB)
struct SlotSet {                                               // the window's scan buffers in reference order i = 0..n-1
  const vx_point_var *pts[MAX_W];
  int32_t *leaf[MAX_W];
  int64_t beg[MAX_W + 1];                                      // running offsets in the concatenated index space
  int32_t phys[MAX_W];                                         // mp[i]
  int n;
};

}  // namespace

struct vx_map {
  vx_ctx *ctx = nullptr;
  vx_map_params prm;
  int W = 0;
  std::vector<int> mp;
  // root hash
  DevBuf hkeys, hvals;
  uint64_t hcap = 0;
  int64_t n_roots = 0;
  // node table
  int64_t n_nodes = 0, cap_nodes = 0;
  DevBuf nodes, meta, add, fix, loc, cov, eig, flag_a, flag_b;
  // window slots (physical)
  std::vector<DevBuf> slot_pts, slot_leaf;
  std::vector<int64_t> slot_n;
  // fixed-point pool
  DevBuf fix_pts, fix_leaf;
  int64_t n_fix = 0, cap_fix = 0;
  // work buffers
  DevBuf w_key0, w_key1, w_val0, w_val1, w_flag, w_scan, w_seg, w_misc, w_cnt, w_part, w_pw, w_xyz;
  vx_planemap view;
};

namespace {

int grow_keep(vx_ctx *ctx, DevBuf &b, size_t used, size_t want) {
  if (want <= b.cap) return VX_OK;
  if (used == 0) return b.reserve(want + want / 2);
  DevBuf nb;
  VX_TRY(nb.reserve(want));
  if (used) VX_CUDA(cudaMemcpyAsync(nb.p, b.p, used, cudaMemcpyDeviceToDevice, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  b.release();
  b = nb;
  return VX_OK;
}

int ensure_nodes(vx_map *m, int64_t need) {
  if (need <= m->cap_nodes) return VX_OK;
  vx_ctx *ctx = m->ctx;
  const int64_t nc = std::max<int64_t>(need + need / 2, 1024), n = m->n_nodes;
  const int W = m->W;
  VX_TRY(grow_keep(ctx, m->nodes, sizeof(vx_map_node) * n, sizeof(vx_map_node) * nc));
  VX_TRY(grow_keep(ctx, m->meta, sizeof(NodeMeta) * n, sizeof(NodeMeta) * nc));
  VX_TRY(grow_keep(ctx, m->add, sizeof(Mom) * n, sizeof(Mom) * nc));
  VX_TRY(grow_keep(ctx, m->fix, sizeof(Mom) * n, sizeof(Mom) * nc));
  VX_TRY(grow_keep(ctx, m->loc, sizeof(Mom) * n * W, sizeof(Mom) * nc * W));
  VX_TRY(grow_keep(ctx, m->cov, sizeof(double) * 81 * n, sizeof(double) * 81 * nc));
  VX_TRY(grow_keep(ctx, m->eig, sizeof(double) * 12 * n, sizeof(double) * 12 * nc));
  VX_TRY(grow_keep(ctx, m->flag_a, sizeof(int32_t) * n, sizeof(int32_t) * (nc + 1)));   // split / fold flags survive a growth
  VX_TRY(grow_keep(ctx, m->flag_b, sizeof(int32_t) * n, sizeof(int32_t) * (nc + 1)));
  m->cap_nodes = (int64_t)std::min<size_t>({m->nodes.cap / sizeof(vx_map_node), m->meta.cap / sizeof(NodeMeta), m->add.cap / sizeof(Mom),
                                            m->fix.cap / sizeof(Mom), m->loc.cap / (sizeof(Mom) * W), m->cov.cap / (81 * sizeof(double)),
                                            m->eig.cap / (12 * sizeof(double)), m->flag_a.cap / 4 - 1, m->flag_b.cap / 4 - 1});
  return VX_OK;
}

int ensure_fix(vx_map *m, int64_t need) {
  if (need <= m->cap_fix) return VX_OK;
  const int64_t nc = std::max<int64_t>(need + need / 2, 4096);
  VX_TRY(grow_keep(m->ctx, m->fix_pts, sizeof(vx_point_var) * m->n_fix, sizeof(vx_point_var) * nc));
  VX_TRY(grow_keep(m->ctx, m->fix_leaf, sizeof(int32_t) * m->n_fix, sizeof(int32_t) * nc));
  m->cap_fix = (int64_t)std::min(m->fix_pts.cap / sizeof(vx_point_var), m->fix_leaf.cap / sizeof(int32_t));
  return VX_OK;
}

// stable sort of (key, val) pairs by key into (key1, val1); n < 2^31
int sort_pairs(vx_map *m, int64_t n, int end_bit = 64) {
  vx_ctx *ctx = m->ctx;
  size_t tb = 0;
  VX_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, m->w_key0.as<uint64_t>(), m->w_key1.as<uint64_t>(), m->w_val0.as<int32_t>(), m->w_val1.as<int32_t>(),
                                          (int)n, 0, end_bit, ctx->stream));
  VX_TRY(ctx->scratch.reserve(tb));
  VX_CUDA(cub::DeviceRadixSort::SortPairs(ctx->scratch.p, tb, m->w_key0.as<uint64_t>(), m->w_key1.as<uint64_t>(), m->w_val0.as<int32_t>(),
                                          m->w_val1.as<int32_t>(), (int)n, 0, end_bit, ctx->stream));
  ctx->launches += 4;
  return VX_OK;
}
int scan_incl(vx_map *m, const int32_t *in, int32_t *out, int64_t n) {
  vx_ctx *ctx = m->ctx;
  size_t tb = 0;
  VX_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tb, in, out, (int)n, ctx->stream));
  VX_TRY(ctx->scratch.reserve(tb));
  VX_CUDA(cub::DeviceScan::InclusiveSum(ctx->scratch.p, tb, in, out, (int)n, ctx->stream));
  ctx->launches += 2;
  return VX_OK;
}
int work_reserve(vx_map *m, int64_t n) {
  n = std::max<int64_t>(n, 1);
  VX_TRY(m->w_key0.reserve(8 * n)); VX_TRY(m->w_key1.reserve(8 * n)); VX_TRY(m->w_val0.reserve(4 * n)); VX_TRY(m->w_val1.reserve(4 * n));
  VX_TRY(m->w_flag.reserve(4 * (n + 1))); VX_TRY(m->w_scan.reserve(4 * (n + 1))); VX_TRY(m->w_seg.reserve(4 * (n + 2)));
  VX_TRY(m->w_misc.reserve(8 * (n + 2)));
  VX_TRY(m->w_cnt.reserve(64));
  return VX_OK;
}
int read_i32(vx_ctx *ctx, const int32_t *d, int32_t *h, int n = 1) {
  VX_CUDA(cudaMemcpyAsync(h, d, 4 * n, cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  return VX_OK;
}

}  // namespace

// =====================================================================================================================
// kernels shared by insertion, fixed-point insertion and subdivision
// =====================================================================================================================
namespace {

__device__ __forceinline__ double fma_off(double acc, double a, double b) { return __dadd_rn(acc, __dmul_rn(a, b)); }

// key of a world point held at xyz[i*stride .. +3)
__global__ void k_pt_rootkey(const double *__restrict__ xyz, int stride, int64_t n, double vs, uint64_t *__restrict__ keys, int32_t *__restrict__ idx,
                             int *__restrict__ bad) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double *p = xyz + i * stride;
  uint64_t k;
  if (!pack_key(key_axis_d(p[0], vs), key_axis_d(p[1], vs), key_axis_d(p[2], vs), k)) *bad = 1;
  keys[i] = k;
  idx[i] = (int32_t)i;
}
__global__ void k_heads(const uint64_t *__restrict__ keys, int64_t n, int32_t *__restrict__ flag) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  flag[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}
// seg_start[s] = first sorted position of segment s, seg_start[nseg] = n; a trailing run of CODE_SKIP keys is cut off
__global__ void k_seg_starts(const uint64_t *__restrict__ keys, const int32_t *__restrict__ flag, const int32_t *__restrict__ incl, int64_t n,
                             int32_t *__restrict__ seg_start, int32_t *__restrict__ nseg_out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (flag[i]) {
    seg_start[incl[i] - 1] = (int32_t)i;
    if (keys[i] == CODE_SKIP) *nseg_out = incl[i] - 1;          // only one such head can exist (sorted last)
  }
  if (i == n - 1) { seg_start[incl[i]] = (int32_t)n; if (keys[i] != CODE_SKIP) *nseg_out = incl[i]; }
}

__global__ void k_root_find(const uint64_t *__restrict__ skeys, const int32_t *__restrict__ seg_start, int nseg, const uint64_t *__restrict__ hkeys,
                            const int32_t *__restrict__ hvals, uint64_t hmask, int32_t *__restrict__ seg_root, int32_t *__restrict__ isnew) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nseg) return;
  const int32_t r = hmask ? root_lookup(hkeys, hvals, hmask, skeys[seg_start[s]]) : -1;
  seg_root[s] = r;
  isnew[s] = r < 0 ? 1 : 0;
}
__global__ void k_rehash(const uint64_t *__restrict__ ok, const int32_t *__restrict__ ov, uint64_t ocap, uint64_t *__restrict__ nk, int32_t *__restrict__ nv,
                         uint64_t nmask) {
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i >= ocap) return;
  const uint64_t key = ok[i];
  if (key == ~0ull) return;
  uint64_t h = mix64(key) & nmask;
  for (;;) {
    const unsigned long long prev = atomicCAS((unsigned long long *)&nk[h], ~0ull, (unsigned long long)key);
    if (prev == ~0ull) { nv[h] = ov[i]; return; }
    h = (h + 1) & nmask;
  }
}

__device__ __forceinline__ void node_defaults(vx_map_node &nd, NodeMeta &mt) {
  nd.octo_state = 0; nd.is_plane = 0; nd.radius = 0.0f;
  for (int k = 0; k < 8; k++) nd.child[k] = -1;
  for (int k = 0; k < 3; k++) { nd.p_center[k] = 0; nd.p_normal[k] = 0; }
  for (int k = 0; k < 36; k++) nd.plane_var[k] = 0;
  mt.isexist = 0; mt.has_sw = 0; mt.in_slide = 0; mt.last_num = 0; mt.opt_state = -1; mt.jour = 0;
}

// new_root (voxel_map.hpp:1935-1943) for the segments that missed the hash; marks of cut_voxel_multi (:1990-2010) for all.
// The moment / cov / eig rows of the new id range were zeroed by the host before this launch.
__global__ void k_root_create(const uint64_t *__restrict__ skeys, const int32_t *__restrict__ seg_start, int nseg, int32_t *__restrict__ seg_root,
                              const int32_t *__restrict__ isnew, const int32_t *__restrict__ isnew_incl, int32_t base, uint64_t *__restrict__ hkeys,
                              int32_t *__restrict__ hvals, uint64_t hmask, vx_map_node *__restrict__ nodes, NodeMeta *__restrict__ meta, double vs,
                              int mark_slide, double jour_new) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nseg) return;
  int32_t id = seg_root[s];
  if (isnew[s]) {
    id = base + isnew_incl[s] - 1;
    const uint64_t key = skeys[seg_start[s]];
    int64_t kx, ky, kz;
    unpack_key(key, kx, ky, kz);
    vx_map_node &nd = nodes[id];
    NodeMeta &mt = meta[id];
    node_defaults(nd, mt);
    nd.center[0] = __dmul_rn(__dadd_rn(0.5, (double)kx), vs);
    nd.center[1] = __dmul_rn(__dadd_rn(0.5, (double)ky), vs);
    nd.center[2] = __dmul_rn(__dadd_rn(0.5, (double)kz), vs);
    nd.quater_length = (float)(vs / 4.0);
    mt.root_key = key; mt.root = id; mt.path = 0; mt.layer = 0;
    mt.jour = jour_new;
    uint64_t h = mix64(key) & hmask;
    for (;;) {
      const unsigned long long prev = atomicCAS((unsigned long long *)&hkeys[h], ~0ull, (unsigned long long)key);
      if (prev == ~0ull) { hvals[h] = id; break; }
      h = (h + 1) & hmask;
    }
    seg_root[s] = id;
    // the scan is walked point by point: the second point of a brand-new cell already finds the root and marks it (:1994)
    if (mark_slide && seg_start[s + 1] - seg_start[s] >= 2) mt.isexist = 1;
  } else if (mark_slide) {
    meta[id].isexist = 1;                       // iter->second->isexist = true (:1994)
  }
  if (mark_slide) meta[id].in_slide = 1;         // feat_tem_map[position] = ... (:1996 / :2008)
}

// OctoTree::allocate (voxel_map.hpp:1204-1237) / allocate_fix (:1239-1264) without the push: where does the point end?
__global__ void k_descend(const double *__restrict__ xyz, int stride, int64_t n, const int32_t *__restrict__ sidx, const int32_t *__restrict__ seg_of,
                          const int32_t *__restrict__ seg_root, const vx_map_node *__restrict__ nodes, int max_layer, int fix_mode,
                          uint64_t *__restrict__ code, int32_t *__restrict__ val) {
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= n) return;
  const int32_t i = sidx[j];
  const double *p = xyz + (int64_t)i * stride;
  int32_t cur = seg_root[seg_of[j] - 1];
  int layer = 0;
  uint64_t c;
  for (;;) {
    const vx_map_node &nd = nodes[cur];
    if (nd.octo_state == 0) { c = (uint64_t)cur; break; }
    if (fix_mode && layer >= max_layer) { c = CODE_SKIP; break; }
    const int ci = 4 * (p[0] > nd.center[0] ? 1 : 0) + 2 * (p[1] > nd.center[1] ? 1 : 0) + (p[2] > nd.center[2] ? 1 : 0);
    const int32_t nx = nd.child[ci];
    if (nx < 0) { c = CODE_CREATE | ((uint64_t)cur * 8 + ci); break; }
    cur = nx; layer++;
  }
  code[i] = c;
  val[i] = i;
}

__global__ void k_seg_isnew(const uint64_t *__restrict__ skeys, const int32_t *__restrict__ seg_start, int nseg, int32_t *__restrict__ isnew) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nseg) return;
  isnew[s] = (skeys[seg_start[s]] & CODE_CREATE) ? 1 : 0;
}

// ensure_leaf (voxel_map.hpp:1222-1232) for every CREATE segment; seg_leaf = the node each segment's points are pushed into
__global__ void k_child_create(const uint64_t *__restrict__ skeys, const int32_t *__restrict__ seg_start, int nseg, const int32_t *__restrict__ isnew,
                               const int32_t *__restrict__ isnew_incl, int32_t base, vx_map_node *__restrict__ nodes, NodeMeta *__restrict__ meta,
                               int32_t *__restrict__ seg_leaf) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nseg) return;
  const uint64_t code = skeys[seg_start[s]];
  if (!isnew[s]) { seg_leaf[s] = (int32_t)code; return; }
  const int32_t id = base + isnew_incl[s] - 1;
  const uint64_t pc = code & (CODE_CREATE - 1);
  const int32_t parent = (int32_t)(pc >> 3);
  const int ci = (int)(pc & 7);
  const int xyz[3] = {(ci >> 2) & 1, (ci >> 1) & 1, ci & 1};
  vx_map_node &pn = nodes[parent];
  const NodeMeta &pm = meta[parent];
  vx_map_node &nd = nodes[id];
  NodeMeta &mt = meta[id];
  node_defaults(nd, mt);
  for (int k = 0; k < 3; k++) nd.center[k] = __dadd_rn(pn.center[k], (double)__fmul_rn((float)(2 * xyz[k] - 1), pn.quater_length));
  nd.quater_length = pn.quater_length / 2;
  mt.root_key = pm.root_key; mt.root = pm.root; mt.layer = pm.layer + 1; mt.path = pm.path | (ci << (3 * pm.layer));
  pn.child[ci] = id;
  seg_leaf[s] = id;
}

// ---- Bf_var (voxel_map.hpp:106-121), one entry at a time --------------------------------------------------------------
__device__ __forceinline__ double bi_entry(int a, int j, const double *v) {
  // rows: 0:(2x,0,0) 1:(y,x,0) 2:(z,0,x) 3:(0,2y,0) 4:(0,z,y) 5:(0,0,2z)
  switch (a * 3 + j) {
    case 0: return __dmul_rn(2.0, v[0]);
    case 3: return v[1];
    case 4: return v[0];
    case 6: return v[2];
    case 8: return v[0];
    case 10: return __dmul_rn(2.0, v[1]);
    case 13: return v[2];
    case 14: return v[1];
    case 17: return __dmul_rn(2.0, v[2]);
    default: return 0.0;
  }
}
// (Bi * var)(a, k), var column-major
__device__ __forceinline__ double biup_entry(int a, int k, const double *var, const double *v) {
  double s = __dmul_rn(bi_entry(a, 0, v), var[k * 3 + 0]);
  s = fma_off(s, bi_entry(a, 1, v), var[k * 3 + 1]);
  s = fma_off(s, bi_entry(a, 2, v), var[k * 3 + 2]);
  return s;
}
__device__ __forceinline__ double bf_entry(int r, int c, const double *var, const double *v) {
  if (r < 6 && c < 6) {
    double s = __dmul_rn(biup_entry(r, 0, var, v), bi_entry(c, 0, v));
    s = fma_off(s, biup_entry(r, 1, var, v), bi_entry(c, 1, v));
    s = fma_off(s, biup_entry(r, 2, var, v), bi_entry(c, 2, v));
    return s;
  }
  if (r < 6) return biup_entry(r, c - 6, var, v);
  if (c < 6) return biup_entry(c, r - 6, var, v);
  return var[(c - 6) * 3 + (r - 6)];
}
// second/first moment component c (0..8) of a point: xx xy xz yy yz zz x y z
__device__ __forceinline__ double mom_entry(int c, const double *p) {
  switch (c) {
    case 0: return __dmul_rn(p[0], p[0]);
    case 1: return __dmul_rn(p[0], p[1]);
    case 2: return __dmul_rn(p[0], p[2]);
    case 3: return __dmul_rn(p[1], p[1]);
    case 4: return __dmul_rn(p[1], p[2]);
    case 5: return __dmul_rn(p[2], p[2]);
    default: return p[c - 6];
  }
}

struct AccArgs {
  SlotSet ss;                       // window points: g in [n_fix_range + ss.beg[i], n_fix_range + ss.beg[i+1]) is slot i
  int64_t n_fix_range;              // fixed points: g in [0, n_fix_range)
  const vx_point_var *fix_pts;      // fixed source (pool, or the caller's array for cut_voxel_fix)
  int32_t *fix_leaf;                // label array of the fixed source (may be NULL)
  const double *pw_explicit;        // world points of slot 0 given by the caller (insertion), else R p + t from `poses`
  const PoseSet *poses;
  int W;
  int fix_novar;                    // push_fix_novar (no covariance term) instead of push_fix
};

// Pushing the sorted points into their target nodes, in two steps so that one huge leaf (a wall right in front of the
// sensor collects thousands of points of a sweep) does not serialise the pass:
//   window point: OctoTree::push (:1105-1143)  pcrs_local[mp[ord]] += local, pcr_add += world, cov_add += Bf_var(pv, pw)
//   fixed point : push_fix (:1149-1162) / push_fix_novar (:1168-1178)   pcr_fix += p, pcr_add += p[, cov_add += Bf_var(pv, p)]
// (1) k_acc_items: the sorted array is cut into ITEMS = runs of at most ACC_ITEM points of one segment and one source
//     (pcr_fix or one window slot).  One warp per item: 32 points at a time are gathered into a shared-memory tile (all
//     gathers in flight together); per point lanes 0..17 form Bi (6x3) and Bi*var once, then lane l adds its components
//     l, l+32, l+64 of {pcr_add P6 v3 | cov_add 81} and lanes 0..8 the 9 numbers of the source cluster.  Sums run in scan
//     order inside an item and are written as a 100-double partial.
// (2) k_acc_final: one warp per segment adds the item partials, in order, onto the stored clusters / cov_add.
// Everything is a fixed-order sum: reproducible bit for bit, within 1e-15 relative of a strictly sequential sum.
constexpr int ACC_WARPS = 4;
constexpr int ACC_ROW = 17;        // pnt 3 | var 9 | pw 3 | pad 2  (odd stride: conflict-free row reads)
constexpr int ACC_ITEM = 128;      // points per item
constexpr int ACC_PART = 100;      // pcr_add 9 | cov_add 81 | source cluster 9 | count

__device__ __forceinline__ int slot_of(const AccArgs &A, int64_t g, int64_t &k) {
  if (g < A.n_fix_range) { k = g; return -1; }
  const int64_t gw = g - A.n_fix_range;
  int slot = 0;
  while (slot + 1 < A.ss.n && gw >= A.ss.beg[slot + 1]) slot++;
  k = gw - A.ss.beg[slot];
  return slot;
}
// item heads: segment start, source change, or every ACC_ITEM points of a segment
__global__ void k_item_heads(const int32_t *__restrict__ sval, const int32_t *__restrict__ seg_incl, const int32_t *__restrict__ seg_start, int64_t n_valid,
                             AccArgs A, int32_t *__restrict__ flag) {
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= n_valid) return;
  const int32_t s0 = seg_start[seg_incl[j] - 1];
  int f = (j == s0 || ((j - s0) % ACC_ITEM) == 0) ? 1 : 0;
  if (!f) {
    int64_t k;
    f = slot_of(A, sval[j], k) != slot_of(A, sval[j - 1], k) ? 1 : 0;
  }
  flag[j] = f;
}
__global__ void k_item_starts(const int32_t *__restrict__ flag, const int32_t *__restrict__ incl, int64_t n_valid, int32_t *__restrict__ item_start) {
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= n_valid) return;
  if (flag[j]) item_start[incl[j] - 1] = (int32_t)j;
  if (j == n_valid - 1) item_start[incl[j]] = (int32_t)n_valid;
}

// Every accumulated number is  sum_{m<3} u_m * w_m  with u, w taken from the point's row {pnt 3 | var 9 | pw 3 | 1 | 0} or from
// the staged {Bi 18 | Bi*var 18}: a 6x6 entry is a 3-term dot product, a copied entry is u*1 + 0*0 + 0*0 (exact), a moment is
// pw_i*pw_j (or pw_k*1).  The operand indices are per-lane constants, so the inner loop has no divergent control flow.
struct AccOp { int16_t u[3], w[3]; };     // index < 64: into the row, >= 64: into the staging buffer (index - 64)
__device__ __forceinline__ AccOp acc_op_for(int c) {
  constexpr int ONE = 15, ZERO = 16, PW = 12, VAR = 3, BI = 64, BIUP = 64 + 18;
  AccOp o;
  for (int m = 0; m < 3; m++) { o.u[m] = ZERO; o.w[m] = ZERO; }
  if (c < 6) { const int ix[6][2] = {{0, 0}, {0, 1}, {0, 2}, {1, 1}, {1, 2}, {2, 2}}; o.u[0] = PW + ix[c][0]; o.w[0] = PW + ix[c][1]; }
  else if (c < 9) { o.u[0] = PW + (c - 6); o.w[0] = ONE; }
  else if (c < 90) {
    const int q = c - 9, r = q % 9, cc = q / 9;
    if (r < 6 && cc < 6) { for (int m = 0; m < 3; m++) { o.u[m] = BIUP + r * 3 + m; o.w[m] = BI + cc * 3 + m; } }
    else if (r < 6) { o.u[0] = BIUP + r * 3 + (cc - 6); o.w[0] = ONE; }
    else if (cc < 6) { o.u[0] = BIUP + cc * 3 + (r - 6); o.w[0] = ONE; }
    else { o.u[0] = VAR + (cc - 6) * 3 + (r - 6); o.w[0] = ONE; }
  }
  return o;
}

__global__ void __launch_bounds__(32 * ACC_WARPS) k_acc_items(const int32_t *__restrict__ sval, const int32_t *__restrict__ item_start, int nitems,
                                                              const int32_t *__restrict__ seg_incl, const int32_t *__restrict__ seg_leaf, AccArgs A,
                                                              double *__restrict__ part) {
  __shared__ double tile_s[ACC_WARPS][32 * ACC_ROW];
  __shared__ double bi_s[ACC_WARPS][2][36];      // [parity][Bi 18 | Bi*var 18]
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int item = blockIdx.x * ACC_WARPS + wib;
  if (item >= nitems) return;
  double *tile = tile_s[wib];
  const int32_t b = item_start[item], e = item_start[item + 1];
  const int32_t leaf = seg_leaf[seg_incl[b] - 1];
  int64_t k0;
  const int slot = slot_of(A, sval[b], k0);          // one source per item
  double a[3] = {0.0, 0.0, 0.0}, run = 0.0;
  AccOp op[3];
#pragma unroll
  for (int t = 0; t < 3; t++) op[t] = acc_op_for(lane + 32 * t);
  // source-cluster moment of lanes 0..8: local pnt for window points, world pnt (= pnt) for fixed points
  const AccOp rop = acc_op_for(lane < 9 ? lane : 0);
  const int roff = slot < 0 ? 0 : -12;                // window: read pnt (row[0..2]) instead of pw (row[12..14])
  // staging roles of lanes 0..17: Bi(aa, j) = cf * pw[ix] and (Bi var)(aa, j) = sum_m Bi(aa, m) var(m, j)
  const int aa = (lane < 18 ? lane : 0) / 3, jj = (lane < 18 ? lane : 0) % 3;
  double cf[3]; int ix[3];
#pragma unroll
  for (int m = 0; m < 3; m++) {
    // rows of Bi: 0:(2x,0,0) 1:(y,x,0) 2:(z,0,x) 3:(0,2y,0) 4:(0,z,y) 5:(0,0,2z)   (voxel_map.hpp:108-113)
    const int code = aa * 3 + m;
    cf[m] = (code == 0 || code == 10 || code == 17) ? 2.0 : ((code == 3 || code == 4 || code == 6 || code == 8 || code == 13 || code == 14) ? 1.0 : 0.0);
    ix[m] = (code == 0 || code == 4 || code == 8) ? 0 : ((code == 3 || code == 10 || code == 14) ? 1 : 2);
  }
  const bool with_cov = !(slot < 0 && A.fix_novar);
  int par = 0;
  for (int32_t base = b; base < e; base += 32) {
    const int cnt = min(32, e - base);
    __syncwarp();
    if (lane < cnt) {                                   // gather: one point per lane
      int64_t k;
      const int64_t g = sval[base + lane];
      slot_of(A, g, k);
      const vx_point_var *pv;
      double pw[3];
      if (slot < 0) {
        pv = A.fix_pts + k;
        pw[0] = pv->pnt[0]; pw[1] = pv->pnt[1]; pw[2] = pv->pnt[2];
        if (A.fix_leaf) A.fix_leaf[k] = leaf;          // the point now lives in `leaf`
      } else {
        pv = A.ss.pts[slot] + k;
        A.ss.leaf[slot][k] = leaf;
        if (A.pw_explicit) { pw[0] = A.pw_explicit[3 * k]; pw[1] = A.pw_explicit[3 * k + 1]; pw[2] = A.pw_explicit[3 * k + 2]; }
        else {
          const double *R = A.poses->R[slot], *t = A.poses->p[slot];
#pragma unroll
          for (int r = 0; r < 3; r++)
            pw[r] = __dadd_rn(fma_off(fma_off(__dmul_rn(R[0 * 3 + r], pv->pnt[0]), R[1 * 3 + r], pv->pnt[1]), R[2 * 3 + r], pv->pnt[2]), t[r]);
        }
      }
      double *row = tile + lane * ACC_ROW;
#pragma unroll
      for (int q = 0; q < 3; q++) row[q] = pv->pnt[q];
#pragma unroll
      for (int q = 0; q < 9; q++) row[3 + q] = pv->var[q];
#pragma unroll
      for (int q = 0; q < 3; q++) row[12 + q] = pw[q];
      row[15] = 1.0; row[16] = 0.0;
    }
    __syncwarp();
    for (int q = 0; q < cnt; q++) {
      const double *row = tile + q * ACC_ROW;
      double *bs = bi_s[wib][par];
      if (with_cov && lane < 18) {
        const double b0 = __dmul_rn(cf[0], row[12 + ix[0]]), b1 = __dmul_rn(cf[1], row[12 + ix[1]]), b2 = __dmul_rn(cf[2], row[12 + ix[2]]);
        bs[lane] = jj == 0 ? b0 : (jj == 1 ? b1 : b2);
        bs[18 + lane] = fma_off(fma_off(__dmul_rn(b0, row[3 + jj * 3 + 0]), b1, row[3 + jj * 3 + 1]), b2, row[3 + jj * 3 + 2]);
      }
      if (lane < 9) run = __dadd_rn(run, __dmul_rn(row[rop.u[0] + roff], row[rop.w[0] == 15 ? 15 : rop.w[0] + roff]));
      __syncwarp();
#pragma unroll
      for (int t = 0; t < 3; t++) {
        if (t < 2 || lane < 26) {                       // components 0..89
          if (with_cov || lane + 32 * t < 9) {
            const double *pu0 = op[t].u[0] < 64 ? row + op[t].u[0] : bs + (op[t].u[0] - 64);
            const double *pu1 = op[t].u[1] < 64 ? row + op[t].u[1] : bs + (op[t].u[1] - 64);
            const double *pu2 = op[t].u[2] < 64 ? row + op[t].u[2] : bs + (op[t].u[2] - 64);
            const double *pw0 = op[t].w[0] < 64 ? row + op[t].w[0] : bs + (op[t].w[0] - 64);
            const double *pw1 = op[t].w[1] < 64 ? row + op[t].w[1] : bs + (op[t].w[1] - 64);
            const double *pw2 = op[t].w[2] < 64 ? row + op[t].w[2] : bs + (op[t].w[2] - 64);
            a[t] = __dadd_rn(a[t], fma_off(fma_off(__dmul_rn(*pu0, *pw0), *pu1, *pw1), *pu2, *pw2));
          }
        }
      }
      par ^= 1;                                          // the next point writes the other staging buffer
    }
  }
  double *p = part + (int64_t)item * ACC_PART;
#pragma unroll
  for (int t = 0; t < 3; t++) if (lane + 32 * t < 90) p[lane + 32 * t] = a[t];
  if (lane < 9) p[90 + lane] = run;
  if (lane == 0) p[99] = (double)(e - b);
}

__global__ void __launch_bounds__(32 * ACC_WARPS) k_acc_final(const int32_t *__restrict__ sval, const int32_t *__restrict__ seg_start,
                                                              const int32_t *__restrict__ seg_leaf, int nseg, const int32_t *__restrict__ item_incl,
                                                              const int32_t *__restrict__ item_start, int nitems, const double *__restrict__ part, AccArgs A,
                                                              NodeMeta *__restrict__ meta, Mom *__restrict__ add, Mom *__restrict__ fix, Mom *__restrict__ loc,
                                                              double *__restrict__ cov) {
  const int lane = threadIdx.x & 31;
  const int seg = blockIdx.x * ACC_WARPS + (threadIdx.x >> 5);
  if (seg >= nseg) return;
  const int32_t leaf = seg_leaf[seg];
  const int it0 = item_incl[seg_start[seg]] - 1;
  const int it1 = seg + 1 < nseg ? item_incl[seg_start[seg + 1]] - 1 : nitems;
  double *addv = reinterpret_cast<double *>(&add[leaf]);
  double *covv = cov + (int64_t)leaf * 81;
  double a[3];
#pragma unroll
  for (int t = 0; t < 3; t++) { const int c = lane + 32 * t; a[t] = c < 9 ? addv[c] : (c < 90 ? covv[c - 9] : 0.0); }
  int n_add = 0, n_win = 0;
  for (int it = it0; it < it1; it++) {
    const double *p = part + (int64_t)it * ACC_PART;
#pragma unroll
    for (int t = 0; t < 3; t++) if (lane + 32 * t < 90) a[t] = __dadd_rn(a[t], p[lane + 32 * t]);
    int64_t k;
    const int slot = slot_of(A, sval[item_start[it]], k);
    const int n = (int)p[99];
    Mom &tgt = slot < 0 ? fix[leaf] : loc[(int64_t)leaf * A.W + A.ss.phys[slot]];
    if (lane < 9) { double *tv = reinterpret_cast<double *>(&tgt); tv[lane] = __dadd_rn(tv[lane], p[90 + lane]); }
    if (lane == 0) tgt.N += n;
    __syncwarp();
    n_add += n;
    if (slot >= 0) n_win += n;
  }
#pragma unroll
  for (int t = 0; t < 3; t++) {
    const int c = lane + 32 * t;
    if (c < 9) addv[c] = a[t];
    else if (c < 90) covv[c - 9] = a[t];
  }
  if (lane == 0) {
    add[leaf].N += n_add;
    if (n_win > 0) { meta[leaf].has_sw = 1; meta[leaf].isexist = 1; }   // push: sw allocated, isexist = true (:1108-1125)
  }
}

}  // namespace

// =====================================================================================================================
// host drivers: root resolution + descent + child creation (shared by window and fixed insertion)
// =====================================================================================================================
namespace {

int zero_new_rows(vx_map *m, int64_t base, int64_t cnt) {
  if (cnt <= 0) return VX_OK;
  cudaStream_t st = m->ctx->stream;
  const int W = m->W;
  VX_CUDA(cudaMemsetAsync(m->add.as<Mom>() + base, 0, sizeof(Mom) * cnt, st));
  VX_CUDA(cudaMemsetAsync(m->fix.as<Mom>() + base, 0, sizeof(Mom) * cnt, st));
  VX_CUDA(cudaMemsetAsync(m->loc.as<Mom>() + base * W, 0, sizeof(Mom) * cnt * W, st));
  VX_CUDA(cudaMemsetAsync(m->cov.as<double>() + base * 81, 0, sizeof(double) * 81 * cnt, st));
  VX_CUDA(cudaMemsetAsync(m->eig.as<double>() + base * 12, 0, sizeof(double) * 12 * cnt, st));
  return VX_OK;
}

int ensure_hash(vx_map *m, int64_t roots_after) {
  vx_ctx *ctx = m->ctx;
  uint64_t want = std::max<uint64_t>(m->hcap, 1024);
  while (want < (uint64_t)roots_after * 2 + 2) want <<= 1;
  if (want == m->hcap) return VX_OK;
  DevBuf nk, nv;
  VX_TRY(nk.reserve(8 * want)); VX_TRY(nv.reserve(4 * want));
  VX_CUDA(cudaMemsetAsync(nk.p, 0xff, 8 * want, ctx->stream));
  VX_CUDA(cudaMemsetAsync(nv.p, 0xff, 4 * want, ctx->stream));
  if (m->hcap)
    VX_KERNEL(ctx, "k_rehash", k_rehash<<<blocks_for((int64_t)m->hcap, 256), 256, 0, ctx->stream>>>(m->hkeys.as<uint64_t>(), m->hvals.as<int32_t>(), m->hcap,
                                                                                                  nk.as<uint64_t>(), nv.as<int32_t>(), want - 1));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  m->hkeys.release(); m->hvals.release();
  m->hkeys = nk; m->hvals = nv; m->hcap = want;
  return VX_OK;
}

// int32 work arrays carved from w_misc (4 x (n + 2))
struct Misc { int32_t *a, *b, *c, *d; };
Misc misc_of(vx_map *m, int64_t n) { int32_t *p = m->w_misc.as<int32_t>(); return Misc{p, p + (n + 2), p + 2 * (n + 2), p + 3 * (n + 2)}; }

// After this call: w_key1 / w_val1 hold (target code, point index) sorted by code, w_seg the segment starts, misc.c the
// node of each segment (children created), *nseg the number of non-skipped segments, *n_root_segs the number of distinct
// root voxels touched.  early_out_below > 0 reproduces voxel_map.hpp:2044-2045 (return after the roots were made).
int place_points(vx_map *m, const double *d_xyz, int stride, int64_t n, bool fix_mode, double jour_new, int early_out_below, int *n_root_segs, int *nseg_out,
                 bool *early_out) {
  vx_ctx *ctx = m->ctx;
  cudaStream_t st = ctx->stream;
  *early_out = false; *nseg_out = 0; *n_root_segs = 0;
  if (n == 0) return VX_OK;
  VX_ARG(n < ((int64_t)1 << 31));
  VX_TRY(work_reserve(m, 2 * n + 4));
  Misc mi = misc_of(m, n);
  int32_t *cnt = m->w_cnt.as<int32_t>();
  VX_CUDA(cudaMemsetAsync(cnt, 0, 64, st));
  VX_KERNEL(ctx, "k_pt_rootkey", k_pt_rootkey<<<blocks_for(n, 256), 256, 0, st>>>(d_xyz, stride, n, m->prm.voxel_size, m->w_key0.as<uint64_t>(), m->w_val0.as<int32_t>(), cnt));
  VX_TRY(sort_pairs(m, n, 63));
  VX_KERNEL(ctx, "k_heads", k_heads<<<blocks_for(n, 256), 256, 0, st>>>(m->w_key1.as<uint64_t>(), n, m->w_flag.as<int32_t>()));
  VX_TRY(scan_incl(m, m->w_flag.as<int32_t>(), m->w_scan.as<int32_t>(), n));
  VX_KERNEL(ctx, "k_seg_starts", k_seg_starts<<<blocks_for(n, 256), 256, 0, st>>>(m->w_key1.as<uint64_t>(), m->w_flag.as<int32_t>(), m->w_scan.as<int32_t>(), n,
                                                                               m->w_seg.as<int32_t>(), cnt + 1));
  int32_t h[2];
  VX_TRY(read_i32(ctx, cnt, h, 2));
  if (h[0]) { set_error("voxel map: a cell index does not fit the 21-bit packed key"); return VX_ERR_ARG; }
  const int nroot = h[1];
  *n_root_segs = nroot;
  // roots: find, count the new ones, make room, create
  VX_KERNEL(ctx, "k_root_find", k_root_find<<<blocks_for(nroot, 256), 256, 0, st>>>(m->w_key1.as<uint64_t>(), m->w_seg.as<int32_t>(), nroot, m->hkeys.as<uint64_t>(),
                                                                                 m->hvals.as<int32_t>(), m->hcap ? m->hcap - 1 : 0, mi.a, mi.b));
  VX_TRY(scan_incl(m, mi.b, mi.d, nroot));
  int32_t n_new = 0;
  VX_TRY(read_i32(ctx, mi.d + (nroot - 1), &n_new));
  VX_TRY(ensure_hash(m, m->n_roots + n_new));
  VX_TRY(ensure_nodes(m, m->n_nodes + n_new));
  VX_TRY(zero_new_rows(m, m->n_nodes, n_new));
  VX_KERNEL(ctx, "k_root_create", k_root_create<<<blocks_for(nroot, 256), 256, 0, st>>>(m->w_key1.as<uint64_t>(), m->w_seg.as<int32_t>(), nroot, mi.a, mi.b, mi.d,
                                                                                     (int32_t)m->n_nodes, m->hkeys.as<uint64_t>(), m->hvals.as<int32_t>(), m->hcap - 1,
                                                                                     m->nodes.as<vx_map_node>(), m->meta.as<NodeMeta>(), m->prm.voxel_size,
                                                                                     fix_mode ? 0 : 1, jour_new));
  m->n_nodes += n_new; m->n_roots += n_new;
  if (early_out_below > 0 && nroot < early_out_below) { *early_out = true; return VX_OK; }
  // descent -> target codes (by original index), second sort groups the points of one target in scan order
  VX_KERNEL(ctx, "k_descend", k_descend<<<blocks_for(n, 256), 256, 0, st>>>(d_xyz, stride, n, m->w_val1.as<int32_t>(), m->w_scan.as<int32_t>(), mi.a,
                                                                         m->nodes.as<vx_map_node>(), m->prm.max_layer, fix_mode ? 1 : 0,
                                                                         m->w_key0.as<uint64_t>(), m->w_val0.as<int32_t>()));
  VX_TRY(sort_pairs(m, n, 42));
  VX_KERNEL(ctx, "k_heads", k_heads<<<blocks_for(n, 256), 256, 0, st>>>(m->w_key1.as<uint64_t>(), n, m->w_flag.as<int32_t>()));
  VX_TRY(scan_incl(m, m->w_flag.as<int32_t>(), m->w_scan.as<int32_t>(), n));
  VX_CUDA(cudaMemsetAsync(cnt, 0, 64, st));
  VX_KERNEL(ctx, "k_seg_starts", k_seg_starts<<<blocks_for(n, 256), 256, 0, st>>>(m->w_key1.as<uint64_t>(), m->w_flag.as<int32_t>(), m->w_scan.as<int32_t>(), n,
                                                                               m->w_seg.as<int32_t>(), cnt + 1));
  VX_TRY(read_i32(ctx, cnt + 1, h, 1));
  const int nseg = h[0];
  *nseg_out = nseg;
  if (nseg == 0) return VX_OK;
  VX_KERNEL(ctx, "k_seg_isnew", k_seg_isnew<<<blocks_for(nseg, 256), 256, 0, st>>>(m->w_key1.as<uint64_t>(), m->w_seg.as<int32_t>(), nseg, mi.b));
  VX_TRY(scan_incl(m, mi.b, mi.d, nseg));
  int32_t n_child = 0;
  VX_TRY(read_i32(ctx, mi.d + (nseg - 1), &n_child));
  VX_TRY(ensure_nodes(m, m->n_nodes + n_child));
  VX_TRY(zero_new_rows(m, m->n_nodes, n_child));
  VX_KERNEL(ctx, "k_child_create", k_child_create<<<blocks_for(nseg, 256), 256, 0, st>>>(m->w_key1.as<uint64_t>(), m->w_seg.as<int32_t>(), nseg, mi.b, mi.d,
                                                                                       (int32_t)m->n_nodes, m->nodes.as<vx_map_node>(), m->meta.as<NodeMeta>(), mi.c));
  m->n_nodes += n_child;
  return VX_OK;
}

int launch_accumulate(vx_map *m, int nseg, const int32_t *seg_leaf, const AccArgs &A) {
  if (nseg == 0) return VX_OK;
  vx_ctx *ctx = m->ctx;
  cudaStream_t st = ctx->stream;
  // sorted layout at this point: w_val1 = point refs, w_seg = segment starts (nseg + 1), w_scan = inclusive segment id per position
  int32_t n_valid = 0;
  VX_TRY(read_i32(ctx, m->w_seg.as<int32_t>() + nseg, &n_valid));
  if (n_valid == 0) return VX_OK;
  int32_t *item_flag = m->w_flag.as<int32_t>();
  int32_t *item_incl = reinterpret_cast<int32_t *>(m->w_key0.p);                // w_key0 is free once the sort is done
  int32_t *item_start = item_incl + (n_valid + 2);
  VX_KERNEL(ctx, "k_item_heads", k_item_heads<<<blocks_for(n_valid, 256), 256, 0, st>>>(m->w_val1.as<int32_t>(), m->w_scan.as<int32_t>(), m->w_seg.as<int32_t>(), n_valid, A, item_flag));
  VX_TRY(scan_incl(m, item_flag, item_incl, n_valid));
  int32_t nitems = 0;
  VX_TRY(read_i32(ctx, item_incl + (n_valid - 1), &nitems));
  VX_KERNEL(ctx, "k_item_starts", k_item_starts<<<blocks_for(n_valid, 256), 256, 0, st>>>(item_flag, item_incl, n_valid, item_start));
  VX_TRY(m->w_part.reserve(sizeof(double) * ACC_PART * (size_t)nitems));
  VX_KERNEL(ctx, "k_acc_items", k_acc_items<<<blocks_for(nitems, ACC_WARPS), 32 * ACC_WARPS, 0, st>>>(m->w_val1.as<int32_t>(), item_start, nitems, m->w_scan.as<int32_t>(), seg_leaf, A,
                                                                                                   m->w_part.as<double>()));
  VX_KERNEL(ctx, "k_acc_final", k_acc_final<<<blocks_for(nseg, ACC_WARPS), 32 * ACC_WARPS, 0, st>>>(m->w_val1.as<int32_t>(), m->w_seg.as<int32_t>(), seg_leaf, nseg, item_incl, item_start,
                                                                                                 nitems, m->w_part.as<double>(), A, m->meta.as<NodeMeta>(), m->add.as<Mom>(),
                                                                                                 m->fix.as<Mom>(), m->loc.as<Mom>(), m->cov.as<double>()));
  return VX_OK;
}

int upload_poses(vx_map *m, const vx_pose *poses, int n, DevBuf &d) {
  static_assert(sizeof(PoseSet) <= 4096, "PoseSet size");
  PoseSet ps;
  std::memset(&ps, 0, sizeof(ps));
  for (int i = 0; i < n; i++) { std::copy(poses[i].R, poses[i].R + 9, ps.R[i]); std::copy(poses[i].p, poses[i].p + 3, ps.p[i]); }
  VX_TRY(d.reserve(sizeof(PoseSet)));
  VX_CUDA(cudaMemcpyAsync(d.p, &ps, sizeof(PoseSet), cudaMemcpyHostToDevice, m->ctx->stream));
  VX_CUDA(cudaStreamSynchronize(m->ctx->stream));   // ps is a stack object
  return VX_OK;
}

SlotSet window_slots(vx_map *m, int win_count) {
  SlotSet ss;
  std::memset(&ss, 0, sizeof(ss));
  ss.n = win_count;
  int64_t run = 0;
  for (int i = 0; i < win_count; i++) {
    const int s = m->mp[i];
    ss.pts[i] = m->slot_pts[s].as<vx_point_var>();
    ss.leaf[i] = m->slot_leaf[s].as<int32_t>();
    ss.beg[i] = run;
    ss.phys[i] = s;
    run += m->slot_n[s];
  }
  ss.beg[win_count] = run;
  return ss;
}

}  // namespace

// =====================================================================================================================
// recut (A9), tras_opt (A11), margi (A12)
// =====================================================================================================================
namespace {

struct MapPrm { double min_eig; double thre[4]; double min_point[4]; int max_layer, max_points, W; };

__device__ __forceinline__ s33 mom_cov(const Mom &c) {   // PointCluster::cov (tools.hpp:333-337)
  const double n = (double)c.N;
  const double cx = c.v[0] / n, cy = c.v[1] / n, cz = c.v[2] / n;
  s33 A;
  A.xx = c.P[0] / n - cx * cx; A.xy = c.P[1] / n - cx * cy; A.xz = c.P[2] / n - cx * cz;
  A.yy = c.P[3] / n - cy * cy; A.yz = c.P[4] / n - cy * cz; A.zz = c.P[5] / n - cz * cz;
  return A;
}

// OctoTree::recut at a leaf of layer L (voxel_map.hpp:1396-1431): decide plane / split.
__global__ void k_recut_leaf(int64_t n_nodes, int L, MapPrm P, vx_map_node *__restrict__ nodes, NodeMeta *__restrict__ meta, const Mom *__restrict__ add,
                             double *__restrict__ eig, int32_t *__restrict__ split, int32_t *__restrict__ n_split) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  split[i] = 0;
  NodeMeta &mt = meta[i];
  vx_map_node &nd = nodes[i];
  if (mt.layer != L || nd.octo_state != 0 || !meta[mt.root].in_slide) return;
  mt.opt_state = -1;
  if ((double)add[i].N <= P.min_point[L]) { nd.is_plane = 0; return; }
  if (!mt.isexist || !mt.has_sw) return;
  double lam[3], U[9];
  eig3_jacobi(mom_cov(add[i]), lam, U);
  double *e = eig + i * 12;
  for (int k = 0; k < 3; k++) e[k] = lam[k];
  for (int k = 0; k < 9; k++) e[3 + k] = U[k];
  const bool plane = lam[0] < P.min_eig && (lam[0] / lam[2]) < P.thre[L];     // plane_judge (:1185-1195)
  nd.is_plane = plane ? 1 : 0;
  if (plane || L >= P.max_layer) return;
  split[i] = 1;
  atomicAdd(n_split, 1);
}

// fix_divide (:1270-1299) + subdivide (:1307-1338): target child of every point that lives in a leaf being split
__global__ void k_split_codes(int64_t G, int64_t n_fix, const vx_point_var *__restrict__ fix_pts, const int32_t *__restrict__ fix_leaf, SlotSet ss,
                              const PoseSet *__restrict__ poses, const int32_t *__restrict__ split, const vx_map_node *__restrict__ nodes,
                              uint64_t *__restrict__ code, int32_t *__restrict__ val) {
  const int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (g >= G) return;
  val[g] = (int32_t)g;
  int32_t leaf;
  double pw[3];
  if (g < n_fix) {
    leaf = fix_leaf[g];
    if (leaf < 0 || !split[leaf]) { code[g] = CODE_SKIP; return; }
    pw[0] = fix_pts[g].pnt[0]; pw[1] = fix_pts[g].pnt[1]; pw[2] = fix_pts[g].pnt[2];
  } else {
    const int64_t gw = g - n_fix;
    int slot = 0;
    while (slot + 1 < ss.n && gw >= ss.beg[slot + 1]) slot++;
    const int64_t k = gw - ss.beg[slot];
    leaf = ss.leaf[slot][k];
    if (leaf < 0 || !split[leaf]) { code[g] = CODE_SKIP; return; }
    const double *p = ss.pts[slot][k].pnt, *R = poses->R[slot], *t = poses->p[slot];
#pragma unroll
    for (int r = 0; r < 3; r++) pw[r] = __dadd_rn(fma_off(fma_off(__dmul_rn(R[0 * 3 + r], p[0]), R[1 * 3 + r], p[1]), R[2 * 3 + r], p[2]), t[r]);
  }
  const vx_map_node &nd = nodes[leaf];
  const int ci = 4 * (pw[0] > nd.center[0] ? 1 : 0) + 2 * (pw[1] > nd.center[1] ? 1 : 0) + (pw[2] > nd.center[2] ? 1 : 0);
  code[g] = CODE_CREATE | ((uint64_t)leaf * 8 + ci);
}

// end of the split branch of recut (:1433-1452): the SlideWindow goes back to the pool, the node becomes internal
__global__ void k_finish_split(int64_t n_nodes, const int32_t *__restrict__ split, vx_map_node *__restrict__ nodes, NodeMeta *__restrict__ meta,
                               Mom *__restrict__ loc, int W) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_nodes || !split[i]) return;
  nodes[i].octo_state = 1;
  meta[i].has_sw = 0;
  double *z = reinterpret_cast<double *>(loc + i * W);
  for (int k = 0; k < W * 10; k++) z[k] = 0.0;
}

// tras_opt (voxel_map.hpp:1605-1638): which leaves go into the LidarFactor
__global__ void k_tras_select(int64_t n_nodes, const vx_map_node *__restrict__ nodes, const NodeMeta *__restrict__ meta, const double *__restrict__ eig,
                              int32_t *__restrict__ sel) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  const NodeMeta &mt = meta[i];
  const vx_map_node &nd = nodes[i];
  int s = 0;
  if (nd.octo_state == 0 && meta[mt.root].in_slide && mt.isexist && nd.is_plane && mt.has_sw) {
    const double *e = eig + i * 12;
    if (!(e[0] / e[1] > 0.12)) s = 1;
  }
  sel[i] = s;
}
__global__ void k_tras_count(int64_t n_nodes, const int32_t *__restrict__ sel, const int32_t *__restrict__ incl, NodeMeta *__restrict__ meta,
                             const Mom *__restrict__ loc, int W, int32_t *__restrict__ vnode, int32_t *__restrict__ vcnt) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_nodes || !sel[i]) return;
  const int v = incl[i] - 1;
  meta[i].opt_state = v;                                   // :1626
  vnode[v] = (int32_t)i;
  int c = 0;
  for (int s = 0; s < W; s++) c += loc[i * W + s].N != 0 ? 1 : 0;
  vcnt[v] = c;
}
struct MpSet { int32_t mp[MAX_W]; };
// LidarFactor::push_voxel (voxel_map.hpp:139-147) straight into the device factor's SoA planes
__global__ void k_tras_fill(int64_t V, int64_t nnz, const int32_t *__restrict__ vnode, const int32_t *__restrict__ vcnt, const int32_t *__restrict__ vincl,
                            MpSet M, int W, const Mom *__restrict__ loc, const Mom *__restrict__ fixm, const Mom *__restrict__ add,
                            const double *__restrict__ eig, int32_t *__restrict__ off, int32_t *__restrict__ frame, double *__restrict__ mom,
                            int32_t *__restrict__ cnt, double *__restrict__ fix, double *__restrict__ coe, double *__restrict__ lam, double *__restrict__ U,
                            double *__restrict__ mrg, int32_t *__restrict__ win) {
  const int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  const int32_t node = vnode[v];
  int32_t e = vincl[v] - vcnt[v];
  off[v] = e;
  if (v == V - 1) off[V] = (int32_t)nnz;
  for (int i = 0; i < W; i++) {
    const Mom &c = loc[(int64_t)node * W + M.mp[i]];      // pcrs[i] = sw->pcrs_local[mp[i]] (:1623-1624)
    if (c.N == 0) continue;
    frame[e] = i;
    for (int k = 0; k < 6; k++) mom[k * nnz + e] = c.P[k];
    for (int k = 0; k < 3; k++) mom[(6 + k) * nnz + e] = c.v[k];
    cnt[e] = c.N;
    e++;
  }
  const Mom &f = fixm[node], &a = add[node];
  for (int k = 0; k < 6; k++) { fix[k * V + v] = f.P[k]; mrg[k * V + v] = a.P[k]; }
  for (int k = 0; k < 3; k++) { fix[(6 + k) * V + v] = f.v[k]; mrg[(6 + k) * V + v] = a.v[k]; }
  fix[9 * V + v] = (double)f.N; mrg[9 * V + v] = (double)a.N;
  coe[v] = 1.0;
  const double *eg = eig + (int64_t)node * 12;
  for (int k = 0; k < 3; k++) lam[k * V + v] = eg[k];
  for (int k = 0; k < 9; k++) U[k * V + v] = eg[3 + k];
  win[v] = 0;
}

__global__ void k_count_slide(int64_t n_nodes, const NodeMeta *__restrict__ meta, int32_t *__restrict__ out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  if (meta[i].layer == 0 && meta[i].in_slide) atomicAdd(out, 1);
}

}  // namespace

namespace {

// PointCluster::transform (tools.hpp:357-363) on the 6+3+N form; returns the symmetric world cluster
__device__ __forceinline__ Mom mom_transform(const Mom &s, const double *R, const double *p) {
  Mom o;
  const double n = (double)s.N;
  double Rv[3];
  for (int r = 0; r < 3; r++) Rv[r] = R[0 * 3 + r] * s.v[0] + R[1 * 3 + r] * s.v[1] + R[2 * 3 + r] * s.v[2];
  for (int r = 0; r < 3; r++) o.v[r] = Rv[r] + n * p[r];
  const double Pf[9] = {s.P[0], s.P[1], s.P[2], s.P[1], s.P[3], s.P[4], s.P[2], s.P[4], s.P[5]};   // symmetric, row r col c = Pf[r*3+c]
  double T[9];   // T = R * P   (T[r*3+c])
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) T[r * 3 + c] = R[0 * 3 + r] * Pf[0 * 3 + c] + R[1 * 3 + r] * Pf[1 * 3 + c] + R[2 * 3 + r] * Pf[2 * 3 + c];
  const int ix[6][2] = {{0, 0}, {0, 1}, {0, 2}, {1, 1}, {1, 2}, {2, 2}};
  for (int k = 0; k < 6; k++) {
    const int r = ix[k][0], c = ix[k][1];
    const double rpr = T[r * 3 + 0] * R[0 * 3 + c] + T[r * 3 + 1] * R[1 * 3 + c] + T[r * 3 + 2] * R[2 * 3 + c];   // (R P R^T)(r,c)
    o.P[k] = rpr + Rv[r] * p[c] + Rv[c] * p[r] + n * (p[r] * p[c]);
  }
  o.N = s.N; o.pad = 0;
  return o;
}
__device__ __forceinline__ void mom_add(Mom &a, const Mom &b) { for (int k = 0; k < 6; k++) a.P[k] += b.P[k]; for (int k = 0; k < 3; k++) a.v[k] += b.v[k]; a.N += b.N; }
__device__ __forceinline__ void mom_sub(Mom &a, const Mom &b) { for (int k = 0; k < 6; k++) a.P[k] -= b.P[k]; for (int k = 0; k < 3; k++) a.v[k] -= b.v[k]; a.N -= b.N; }

// OctoTree::plane_update (voxel_map.hpp:1344-1388)
__device__ void plane_update(vx_map_node &nd, const Mom &add, const double *eg, const double *cov) {
  const double nv = 1.0 / (double)add.N;
  double ctr[3] = {add.v[0] / (double)add.N, add.v[1] / (double)add.N, add.v[2] / (double)add.N};
  const double *lam = eg, *U = eg + 3;   // U column-major: column k = U[3k..3k+2]
  double uc[27];                           // u_c, 3 x 9, row-major uc[r*9 + c]
  for (int k = 0; k < 27; k++) uc[k] = 0.0;
  const double *ul = U;                    // l = 0
  for (int k = 1; k < 3; k++) {
    const double *uk = U + 3 * k;
    double f[9];
    // ukl = uk * ul^T ; fkl = [ukl00, ukl10+ukl01, ukl20+ukl02, ukl11, ukl12+ukl21, ukl22, -(uk.c ul + ul.c uk)]
    f[0] = uk[0] * ul[0];
    f[1] = uk[1] * ul[0] + uk[0] * ul[1];
    f[2] = uk[2] * ul[0] + uk[0] * ul[2];
    f[3] = uk[1] * ul[1];
    f[4] = uk[1] * ul[2] + uk[2] * ul[1];
    f[5] = uk[2] * ul[2];
    const double dk = uk[0] * ctr[0] + uk[1] * ctr[1] + uk[2] * ctr[2], dl = ul[0] * ctr[0] + ul[1] * ctr[1] + ul[2] * ctr[2];
    for (int r = 0; r < 3; r++) f[6 + r] = -(dk * ul[r] + dl * uk[r]);
    const double w = nv / (lam[0] - lam[k]);
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 9; c++) uc[r * 9 + c] += w * uk[r] * f[c];
  }
  double Jc[27];                           // Jc = u_c * cov_add (cov column-major 9x9)
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 9; c++) {
      double s = 0;
      for (int k = 0; k < 9; k++) s += uc[r * 9 + k] * cov[c * 9 + k];
      Jc[r * 9 + c] = s;
    }
  double *pv = nd.plane_var;               // 6x6 column-major
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) {
      double s = 0;
      for (int k = 0; k < 9; k++) s += Jc[r * 9 + k] * uc[c * 9 + k];
      pv[c * 6 + r] = s;
      const double jn = nv * Jc[r * 9 + 6 + c];
      pv[(3 + c) * 6 + r] = jn;
      pv[r * 6 + (3 + c)] = jn;
      pv[(3 + c) * 6 + (3 + r)] = nv * nv * cov[(6 + c) * 9 + (6 + r)];
    }
  for (int k = 0; k < 3; k++) { nd.p_center[k] = ctr[k]; nd.p_normal[k] = ul[k]; }
  nd.radius = (float)lam[2];
}

// OctoTree::margi at a leaf (voxel_map.hpp:1468-1575).  fold[i] = 1: the leaf's points of the marginalised slots join
// point_fix; clearfix[i] = 1: point_fix is dropped (:1560-1565).
__global__ void k_margi_leaf(int64_t n_nodes, MapPrm P, MpSet M, int win_count, int mgsize, const PoseSet *__restrict__ poses, vx_map_node *__restrict__ nodes,
                             NodeMeta *__restrict__ meta, Mom *__restrict__ add, Mom *__restrict__ fix, Mom *__restrict__ loc, const double *__restrict__ cov,
                             double *__restrict__ eig, int64_t V, const double *__restrict__ f_lam, const double *__restrict__ f_U, const double *__restrict__ f_mrg,
                             int32_t *__restrict__ fold, int32_t *__restrict__ clearfix, int32_t *__restrict__ err) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  fold[i] = 0; clearfix[i] = 0;
  NodeMeta &mt = meta[i];
  vx_map_node &nd = nodes[i];
  if (nd.octo_state != 0 || !meta[mt.root].in_slide) return;
  if (!mt.isexist || !mt.has_sw) return;
  if (mt.opt_state >= V) { atomicExch(err, 1); return; }          // "Error: opt_state" exit (:1488-1492)
  const int W = P.W;
  Mom world[MAX_MG];
  for (int k = 0; k < mgsize; k++) world[k].N = 0;
  Mom a;
  double *eg = eig + i * 12;
  if (mt.opt_state >= 0) {                                           // optimizer cache (:1495-1507)
    const int64_t v = mt.opt_state;
    for (int k = 0; k < 6; k++) a.P[k] = f_mrg[k * V + v];
    for (int k = 0; k < 3; k++) a.v[k] = f_mrg[(6 + k) * V + v];
    a.N = (int32_t)f_mrg[9 * V + v]; a.pad = 0;
    for (int k = 0; k < 3; k++) eg[k] = f_lam[k * V + v];
    for (int k = 0; k < 9; k++) eg[3 + k] = f_U[k * V + v];
    mt.opt_state = -1;
    for (int k = 0; k < mgsize; k++) {
      const Mom &c = loc[i * W + M.mp[k]];
      if (c.N != 0) world[k] = mom_transform(c, poses->R[k], poses->p[k]);
    }
  } else {                                                           // recompute (:1509-1528)
    a = fix[i];
    for (int k = 0; k < win_count; k++) {
      const Mom &c = loc[i * W + M.mp[k]];
      if (c.N != 0) {
        const Mom w = mom_transform(c, poses->R[k], poses->p[k]);
        if (k < mgsize) world[k] = w;
        mom_add(a, w);
      }
    }
    if (nd.is_plane) {
      double lam[3], U[9];
      eig3_jacobi(mom_cov(a), lam, U);
      for (int k = 0; k < 3; k++) eg[k] = lam[k];
      for (int k = 0; k < 9; k++) eg[3 + k] = U[k];
    }
  }
  Mom fx = fix[i];
  if (fx.N < P.max_points && nd.is_plane)
    if (a.N - mt.last_num >= 5 || mt.last_num <= 10) { plane_update(nd, a, eg, cov + i * 81); mt.last_num = a.N; }
  if (fx.N < P.max_points) {
    bool any = false;
    for (int k = 0; k < mgsize; k++) if (world[k].N != 0) { mom_add(fx, world[k]); any = true; }
    if (any && mt.layer < P.max_layer) fold[i] = 1;                  // sw->points is only filled below max_layer (:1131-1132)
  } else {
    for (int k = 0; k < mgsize; k++) if (world[k].N != 0) mom_sub(a, world[k]);
    clearfix[i] = 1;
  }
  for (int k = 0; k < mgsize; k++) {
    Mom &c = loc[i * W + M.mp[k]];
    if (c.N != 0) { double *z = reinterpret_cast<double *>(&c); for (int q = 0; q < 10; q++) z[q] = 0.0; }
  }
  add[i] = a;
  fix[i] = fx;
  mt.isexist = (fx.N >= a.N) ? 0 : 1;
}

// internal nodes, bottom-up: isexist = OR over children (voxel_map.hpp:1582-1597)
__global__ void k_margi_internal(int64_t n_nodes, int L, const vx_map_node *__restrict__ nodes, NodeMeta *__restrict__ meta) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  NodeMeta &mt = meta[i];
  if (mt.layer != L || nodes[i].octo_state != 1 || !meta[mt.root].in_slide) return;
  int ex = 0;
  for (int k = 0; k < 8; k++) { const int32_t c = nodes[i].child[k]; if (c >= 0 && meta[c].isexist) ex = 1; }
  mt.isexist = ex;
}

// the marginalised slots' points of folding leaves -> point_fix, in world coordinates (:1547-1556)
__global__ void k_fold_flags(const int32_t *__restrict__ leaf, int64_t n, const int32_t *__restrict__ fold, int32_t *__restrict__ flag) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int32_t l = leaf[k];
  flag[k] = (l >= 0 && fold[l]) ? 1 : 0;
}
__global__ void k_fold_scatter(const vx_point_var *__restrict__ pts, const int32_t *__restrict__ leaf, int64_t n, const int32_t *__restrict__ flag,
                               const int32_t *__restrict__ incl, const PoseSet *__restrict__ poses, int pose_idx, vx_point_var *__restrict__ fix_pts,
                               int32_t *__restrict__ fix_leaf, int64_t base) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n || !flag[k]) return;
  const int64_t d = base + incl[k] - 1;
  vx_point_var o = pts[k];
  const double *R = poses->R[pose_idx], *t = poses->p[pose_idx];
  const double p[3] = {o.pnt[0], o.pnt[1], o.pnt[2]};
  for (int r = 0; r < 3; r++) o.pnt[r] = __dadd_rn(fma_off(fma_off(__dmul_rn(R[0 * 3 + r], p[0]), R[1 * 3 + r], p[1]), R[2 * 3 + r], p[2]), t[r]);
  fix_pts[d] = o;
  fix_leaf[d] = leaf[k];
}
__global__ void k_clear_fix(int64_t n_fix, int32_t *__restrict__ fix_leaf, const int32_t *__restrict__ clearfix, int32_t *__restrict__ n_dead) {
  const int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (g >= n_fix) return;
  const int32_t l = fix_leaf[g];
  if (l >= 0 && clearfix[l]) { fix_leaf[g] = -1; atomicAdd(n_dead, 1); }
}
// slide-set erase (voxelslam.cpp:1665-1674) + clear_slwd (voxel_map.hpp:1856-1880)
__global__ void k_slide_jour(int64_t n_nodes, NodeMeta *__restrict__ meta, double jour) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  if (meta[i].layer == 0 && meta[i].in_slide) meta[i].jour = jour;
}
__global__ void k_erase_mark(int64_t n_nodes, const NodeMeta *__restrict__ meta, int32_t *__restrict__ erased) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  erased[i] = (meta[i].layer == 0 && meta[i].in_slide && !meta[i].isexist) ? 1 : 0;
}
__global__ void k_erase_apply(int64_t n_nodes, NodeMeta *__restrict__ meta, const int32_t *__restrict__ erased, Mom *__restrict__ loc, int W) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  NodeMeta &mt = meta[i];
  if (!erased[mt.root]) return;
  if (mt.has_sw) {
    mt.has_sw = 0;
    double *z = reinterpret_cast<double *>(loc + i * W);
    for (int k = 0; k < W * 10; k++) z[k] = 0.0;
  }
}
__global__ void k_erase_finish(int64_t n_nodes, NodeMeta *__restrict__ meta, const int32_t *__restrict__ erased) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  if (erased[i]) meta[i].in_slide = 0;
}
// compaction of the fixed pool (dead = label -1)
__global__ void k_live_flags(const int32_t *__restrict__ fix_leaf, int64_t n, int32_t *__restrict__ flag) {
  const int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (g < n) flag[g] = fix_leaf[g] >= 0 ? 1 : 0;
}
__global__ void k_compact_fix(const vx_point_var *__restrict__ src, const int32_t *__restrict__ src_leaf, int64_t n, const int32_t *__restrict__ flag,
                              const int32_t *__restrict__ incl, vx_point_var *__restrict__ dst, int32_t *__restrict__ dst_leaf) {
  const int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (g >= n || !flag[g]) return;
  dst[incl[g] - 1] = src[g];
  dst_leaf[incl[g] - 1] = src_leaf[g];
}

__global__ void k_export(int64_t n_nodes, const vx_map_node *__restrict__ nodes, const NodeMeta *__restrict__ meta, const Mom *__restrict__ add,
                         const Mom *__restrict__ fix, const double *__restrict__ eig, const double *__restrict__ cov, vx_map_node_state *__restrict__ out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_nodes) return;
  vx_map_node_state s;
  const NodeMeta &mt = meta[i];
  unpack_key(mt.root_key, s.root_key[0], s.root_key[1], s.root_key[2]);
  s.layer = mt.layer; s.path = mt.path; s.octo_state = nodes[i].octo_state; s.isexist = mt.isexist; s.has_sw = mt.has_sw;
  s.is_plane = nodes[i].is_plane; s.last_num = mt.last_num; s.opt_state = mt.opt_state; s.in_slide = meta[mt.root].in_slide;
  s.n_window_pts = 0; s.n_fix_pts = 0; s._pad = 0;
  const Mom *src[2] = {&add[i], &fix[i]};
  vx_cluster *dst[2] = {&s.pcr_add, &s.pcr_fix};
  for (int q = 0; q < 2; q++) {
    const Mom &c = *src[q];
    vx_cluster &o = *dst[q];
    o.P[0] = c.P[0]; o.P[1] = c.P[1]; o.P[2] = c.P[2]; o.P[3] = c.P[1]; o.P[4] = c.P[3]; o.P[5] = c.P[4]; o.P[6] = c.P[2]; o.P[7] = c.P[4]; o.P[8] = c.P[5];
    o.v[0] = c.v[0]; o.v[1] = c.v[1]; o.v[2] = c.v[2]; o.N = c.N; o._pad = 0;
  }
  for (int k = 0; k < 3; k++) s.eig_value[k] = eig[i * 12 + k];
  double tr = 0;
  for (int k = 0; k < 9; k++) tr += cov[i * 81 + k * 10];
  s.cov_add_trace = tr;
  out[i] = s;
}
__global__ void k_count_labels(const int32_t *__restrict__ leaf, int64_t n, int32_t *__restrict__ counts) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int32_t l = leaf[k];
  if (l >= 0) atomicAdd(&counts[l], 1);
}

}  // namespace

// =====================================================================================================================
// C ABI
// =====================================================================================================================
namespace {
MapPrm map_prm(const vx_map *m) {
  MapPrm P;
  P.min_eig = m->prm.min_eigen_value;
  for (int k = 0; k < 4; k++) { P.thre[k] = m->prm.plane_eigen_value_thre[k]; P.min_point[k] = m->prm.min_point[k]; }
  P.max_layer = m->prm.max_layer; P.max_points = m->prm.max_points; P.W = m->W;
  return P;
}
MpSet mp_set(const vx_map *m) { MpSet M; std::memset(&M, 0, sizeof(M)); for (int i = 0; i < m->W; i++) M.mp[i] = m->mp[i]; return M; }
int count_slide(vx_map *m, int32_t *out) {
  vx_ctx *ctx = m->ctx;
  VX_TRY(m->w_cnt.reserve(64));
  VX_CUDA(cudaMemsetAsync(m->w_cnt.p, 0, 64, ctx->stream));
  if (m->n_nodes) VX_KERNEL(ctx, "k_count_slide", k_count_slide<<<blocks_for(m->n_nodes, 256), 256, 0, ctx->stream>>>(m->n_nodes, m->meta.as<NodeMeta>(), m->w_cnt.as<int32_t>()));
  return read_i32(ctx, m->w_cnt.as<int32_t>(), out);
}
}  // namespace

namespace {
__global__ void k_keep_flags(const int32_t *__restrict__ lab, int64_t n, const NodeMeta *__restrict__ meta, int max_layer, int32_t *__restrict__ flag) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int32_t l = lab[k];
  flag[k] = (l >= 0 && meta[l].layer < max_layer) ? 1 : 0;
}
__global__ void k_append_scatter(const vx_point_var *__restrict__ pts, const int32_t *__restrict__ lab, int64_t n, const int32_t *__restrict__ flag,
                                 const int32_t *__restrict__ incl, vx_point_var *__restrict__ fix_pts, int32_t *__restrict__ fix_leaf, int64_t base) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n || !flag[k]) return;
  fix_pts[base + incl[k] - 1] = pts[k];
  fix_leaf[base + incl[k] - 1] = lab[k];
}
// appends the labelled points whose leaf keeps raw points to the fixed pool
int fix_append(vx_map *m, const vx_point_var *d_pts, const int32_t *d_lab, int64_t n) {
  vx_ctx *ctx = m->ctx;
  cudaStream_t st = ctx->stream;
  VX_TRY(work_reserve(m, 2 * n + 4));
  VX_KERNEL(ctx, "k_keep_flags", k_keep_flags<<<blocks_for(n, 256), 256, 0, st>>>(d_lab, n, m->meta.as<NodeMeta>(), m->prm.max_layer, m->w_flag.as<int32_t>()));
  VX_TRY(scan_incl(m, m->w_flag.as<int32_t>(), m->w_scan.as<int32_t>(), n));
  int32_t keep = 0;
  VX_TRY(read_i32(ctx, m->w_scan.as<int32_t>() + (n - 1), &keep));
  if (keep == 0) return VX_OK;
  VX_TRY(ensure_fix(m, m->n_fix + keep));
  VX_KERNEL(ctx, "k_append_scatter", k_append_scatter<<<blocks_for(n, 256), 256, 0, st>>>(d_pts, d_lab, n, m->w_flag.as<int32_t>(), m->w_scan.as<int32_t>(),
                                                                                       m->fix_pts.as<vx_point_var>(), m->fix_leaf.as<int32_t>(), m->n_fix));
  m->n_fix += keep;
  return VX_OK;
}
}  // namespace

// shared body of the two window-insertion entry points: `fill` puts the n pointVar records at d_pts and the world points at d_pw
template <class Fill>
static int cut_voxel_impl(vx_ctx *ctx, vx_map *m, int64_t n, int ord, Fill fill) {
  cudaStream_t st = ctx->stream;
  const int s = m->mp[ord];
  const int64_t base = m->slot_n[s];
  // the scan joins its physical slot (appended: a slot that multi_margi skipped keeps its stale points, voxelslam.cpp:1616-1617)
  VX_TRY(grow_keep(ctx, m->slot_pts[s], sizeof(vx_point_var) * base, sizeof(vx_point_var) * (base + n)));
  VX_TRY(grow_keep(ctx, m->slot_leaf[s], sizeof(int32_t) * base, sizeof(int32_t) * (base + n)));
  vx_point_var *d_pts = m->slot_pts[s].as<vx_point_var>() + base;
  int32_t *d_leaf = m->slot_leaf[s].as<int32_t>() + base;
  VX_TRY(m->w_pw.reserve(sizeof(double) * 3 * n));
  double *d_pw = m->w_pw.as<double>();
  VX_TRY(fill(d_pts, d_pw));
  VX_CUDA(cudaMemsetAsync(d_leaf, 0xff, sizeof(int32_t) * n, st));
  int nroot = 0, nseg = 0; bool early = false;
  VX_TRY(place_points(m, d_pw, 3, n, false, 0.0, m->prm.thread_num, &nroot, &nseg, &early));
  if (early) return VX_OK;                     // voxel_map.hpp:2044-2045: nothing is inserted
  AccArgs A;
  std::memset(&A, 0, sizeof(A));
  A.ss.n = 1; A.ss.pts[0] = d_pts; A.ss.leaf[0] = d_leaf; A.ss.beg[0] = 0; A.ss.beg[1] = n; A.ss.phys[0] = s;
  A.n_fix_range = 0; A.pw_explicit = d_pw; A.W = m->W;
  VX_TRY(launch_accumulate(m, nseg, misc_of(m, n).c, A));
  m->slot_n[s] = base + n;
  return VX_OK;
}

extern "C" {

int vx_map_create(vx_ctx *ctx, const vx_map_params *params, vx_map **out) {
  VX_ARG(ctx && params && out);
  VX_ARG(params->voxel_size > 0 && params->win_size >= 1 && params->win_size <= MAX_W && params->max_layer >= 0 && params->max_layer <= 3);
  VX_TRY(bind(ctx));
  vx_map *m = new vx_map();
  m->ctx = ctx; m->prm = *params; m->W = params->win_size;
  m->mp.resize(m->W);
  for (int i = 0; i < m->W; i++) m->mp[i] = i;             // voxelslam.cpp:3158-3160
  m->slot_pts.resize(m->W); m->slot_leaf.resize(m->W); m->slot_n.assign(m->W, 0);
  m->view.owns = false;
  *out = m;
  return VX_OK;
}

int vx_map_free(vx_map *m) {
  if (!m) return VX_OK;
  cudaSetDevice(m->ctx->device);
  cudaStreamSynchronize(m->ctx->stream);
  for (DevBuf *b : {&m->hkeys, &m->hvals, &m->nodes, &m->meta, &m->add, &m->fix, &m->loc, &m->cov, &m->eig, &m->flag_a, &m->flag_b, &m->fix_pts, &m->fix_leaf,
                    &m->w_key0, &m->w_key1, &m->w_val0, &m->w_val1, &m->w_flag, &m->w_scan, &m->w_seg, &m->w_misc, &m->w_cnt, &m->w_part, &m->w_pw, &m->w_xyz})
    b->release();
  for (auto &b : m->slot_pts) b.release();
  for (auto &b : m->slot_leaf) b.release();
  m->view.hkeys.p = m->view.hvals.p = m->view.nodes.p = nullptr;
  delete m;
  return VX_OK;
}

int64_t vx_map_num_nodes(const vx_map *m) { return m ? m->n_nodes : 0; }
int64_t vx_map_num_roots(const vx_map *m) { return m ? m->n_roots : 0; }

int vx_map_shift_window(vx_map *m, int mgsize) {
  VX_ARG(m && mgsize >= 0 && mgsize <= m->W);
  for (int i = 0; i < m->W; i++) { m->mp[i] += mgsize; if (m->mp[i] >= m->W) m->mp[i] -= m->W; }
  return VX_OK;
}

int vx_map_planemap(vx_map *m, vx_planemap **view) {
  VX_ARG(m && view);
  vx_planemap &v = m->view;
  v.ctx = m->ctx; v.owns = false;
  v.hkeys.p = m->hkeys.p; v.hkeys.cap = 0; v.hvals.p = m->hvals.p; v.hvals.cap = 0; v.nodes.p = m->nodes.p; v.nodes.cap = 0;
  v.hmask = m->hcap ? m->hcap - 1 : 0; v.n_nodes = m->n_nodes; v.n_roots = m->n_roots; v.voxel_size = m->prm.voxel_size;
  if (!m->hcap) { VX_TRY(ensure_hash(m, 0)); v.hkeys.p = m->hkeys.p; v.hvals.p = m->hvals.p; v.hmask = m->hcap - 1; }
  *view = &v;
  return VX_OK;
}

int vx_map_cut_voxel(vx_ctx *ctx, vx_map *m, const vx_point_var *pvec, const double *pwld, int64_t n, int ord) {
  VX_ARG(ctx && m && m->ctx == ctx && n >= 0 && ord >= 0 && ord < m->W && (n == 0 || (pvec && pwld)));
  VX_TRY(bind(ctx));
  if (n == 0) return VX_OK;
  cudaStream_t st = ctx->stream;
  const int rc = cut_voxel_impl(ctx, m, n, ord, [&](vx_point_var *d_pts, double *d_pw) -> int {
    VX_CUDA(cudaMemcpyAsync(d_pts, pvec, sizeof(vx_point_var) * n, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemcpyAsync(d_pw, pwld, sizeof(double) * 3 * n, cudaMemcpyHostToDevice, st));
    return VX_OK;
  });
  cudaStreamSynchronize(st);
  return rc;
}

int vx_map_cut_voxel_xyz(vx_ctx *ctx, vx_map *m, const float *xyz, int64_t n, const vx_pose *ext, double dept_err, double beam_err, const vx_pose *x,
                         const double *rot_var9, const double *tsl_var9, int ord) {
  VX_ARG(ctx && m && m->ctx == ctx && n >= 0 && ord >= 0 && ord < m->W && ext && x && rot_var9 && tsl_var9 && (n == 0 || xyz));
  VX_TRY(bind(ctx));
  if (n == 0) return VX_OK;
  cudaStream_t st = ctx->stream;
  const int rc = cut_voxel_impl(ctx, m, n, ord, [&](vx_point_var *d_pts, double *d_pw) -> int {
    VX_TRY(m->w_xyz.reserve(sizeof(float) * 3 * n));
    VX_CUDA(cudaMemcpyAsync(m->w_xyz.p, xyz, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, st));
    VX_TRY(var_init_device(ctx, m->w_xyz.as<float>(), n, *ext, dept_err, beam_err, d_pts));        // voxelslam.cpp:1896
    VX_TRY(pvec_update_device(ctx, d_pts, n, *x, rot_var9, tsl_var9, d_pts, d_pw));                 // voxelslam.cpp:1903
    return VX_OK;
  });
  cudaStreamSynchronize(st);
  return rc;
}

int vx_map_cut_voxel_fix(vx_ctx *ctx, vx_map *m, const vx_point_var *pvec, int64_t n, double jour) {
  VX_ARG(ctx && m && m->ctx == ctx && n >= 0 && (n == 0 || pvec));
  VX_TRY(bind(ctx));
  if (n == 0) return VX_OK;
  cudaStream_t st = ctx->stream;
  DevBuf d_in, d_lab;
  auto body = [&]() -> int {
    VX_TRY(d_in.reserve(sizeof(vx_point_var) * n)); VX_TRY(d_lab.reserve(sizeof(int32_t) * n));
    VX_CUDA(cudaMemcpyAsync(d_in.p, pvec, sizeof(vx_point_var) * n, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemsetAsync(d_lab.p, 0xff, sizeof(int32_t) * n, st));
    int nroot = 0, nseg = 0; bool early = false;
    VX_TRY(place_points(m, reinterpret_cast<const double *>(d_in.p), 12, n, true, jour, 0, &nroot, &nseg, &early));
    AccArgs A;
    std::memset(&A, 0, sizeof(A));
    A.n_fix_range = n; A.fix_pts = d_in.as<vx_point_var>(); A.fix_leaf = d_lab.as<int32_t>(); A.W = m->W; A.fix_novar = 1;
    VX_TRY(launch_accumulate(m, nseg, misc_of(m, n).c, A));
    // point_fix keeps the raw points of leaves below max_layer (:1170-1171): append those, in input order
    VX_TRY(fix_append(m, d_in.as<vx_point_var>(), d_lab.as<int32_t>(), n));
    return VX_OK;
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  d_in.release(); d_lab.release();
  return rc;
}

}  // extern "C"

extern "C" {

int vx_map_recut(vx_ctx *ctx, vx_map *m, const vx_pose *poses, int win_count, vx_factor *f) {
  VX_ARG(ctx && m && m->ctx == ctx && poses && f && f->ctx == ctx && win_count >= 1 && win_count <= m->W);
  VX_ARG(f->W == m->W && f->n_win == 1);
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  // voxhess.clear() precedes the call in the reference (voxelslam.cpp:1916-1917): the factor is rebuilt from scratch
  VX_TRY(vx_factor_clear(f));
  int32_t n_slide = 0;
  VX_TRY(count_slide(m, &n_slide));
  if (n_slide < m->prm.thread_num) return VX_OK;                     // voxelslam.cpp:1693-1694
  DevBuf d_poses;
  const MapPrm P = map_prm(m);
  auto body = [&]() -> int {
    VX_TRY(upload_poses(m, poses, win_count, d_poses));
    int32_t *cnt = m->w_cnt.as<int32_t>();
    for (int L = 0; L <= m->prm.max_layer; L++) {
      VX_CUDA(cudaMemsetAsync(cnt, 0, 64, st));
      VX_KERNEL(ctx, "k_recut_leaf", k_recut_leaf<<<blocks_for(m->n_nodes, 128), 128, 0, st>>>(m->n_nodes, L, P, m->nodes.as<vx_map_node>(), m->meta.as<NodeMeta>(),
                                                                                            m->add.as<Mom>(), m->eig.as<double>(), m->flag_a.as<int32_t>(), cnt));
      int32_t n_split = 0;
      VX_TRY(read_i32(ctx, cnt, &n_split));
      if (n_split == 0) continue;
      // every point of a split leaf moves one level down: fixed points first (fix_divide), then the window slots
      SlotSet ss = window_slots(m, win_count);
      const int64_t G = m->n_fix + ss.beg[win_count];
      VX_ARG(G < ((int64_t)1 << 31));
      const int64_t n_before = m->n_nodes;
      if (G > 0) {
        VX_TRY(work_reserve(m, 2 * G + 4));
        VX_KERNEL(ctx, "k_split_codes", k_split_codes<<<blocks_for(G, 256), 256, 0, st>>>(G, m->n_fix, m->fix_pts.as<vx_point_var>(), m->fix_leaf.as<int32_t>(), ss,
                                                                                       d_poses.as<PoseSet>(), m->flag_a.as<int32_t>(), m->nodes.as<vx_map_node>(),
                                                                                       m->w_key0.as<uint64_t>(), m->w_val0.as<int32_t>()));
        VX_TRY(sort_pairs(m, G, 42));
        VX_KERNEL(ctx, "k_heads", k_heads<<<blocks_for(G, 256), 256, 0, st>>>(m->w_key1.as<uint64_t>(), G, m->w_flag.as<int32_t>()));
        VX_TRY(scan_incl(m, m->w_flag.as<int32_t>(), m->w_scan.as<int32_t>(), G));
        VX_CUDA(cudaMemsetAsync(cnt, 0, 64, st));
        VX_KERNEL(ctx, "k_seg_starts", k_seg_starts<<<blocks_for(G, 256), 256, 0, st>>>(m->w_key1.as<uint64_t>(), m->w_flag.as<int32_t>(), m->w_scan.as<int32_t>(), G,
                                                                                     m->w_seg.as<int32_t>(), cnt + 1));
        int32_t nseg = 0;
        VX_TRY(read_i32(ctx, cnt + 1, &nseg));
        if (nseg > 0) {
          Misc mi = misc_of(m, G);
          VX_KERNEL(ctx, "k_seg_isnew", k_seg_isnew<<<blocks_for(nseg, 256), 256, 0, st>>>(m->w_key1.as<uint64_t>(), m->w_seg.as<int32_t>(), nseg, mi.b));
          VX_TRY(scan_incl(m, mi.b, mi.d, nseg));
          VX_TRY(ensure_nodes(m, m->n_nodes + nseg));
          VX_TRY(zero_new_rows(m, m->n_nodes, nseg));
          VX_KERNEL(ctx, "k_child_create", k_child_create<<<blocks_for(nseg, 256), 256, 0, st>>>(m->w_key1.as<uint64_t>(), m->w_seg.as<int32_t>(), nseg, mi.b, mi.d,
                                                                                               (int32_t)m->n_nodes, m->nodes.as<vx_map_node>(), m->meta.as<NodeMeta>(), mi.c));
          m->n_nodes += nseg;
          AccArgs A;
          std::memset(&A, 0, sizeof(A));
          A.ss = ss; A.n_fix_range = m->n_fix; A.fix_pts = m->fix_pts.as<vx_point_var>(); A.fix_leaf = m->fix_leaf.as<int32_t>();
          A.poses = d_poses.as<PoseSet>(); A.W = m->W; A.fix_novar = 0;
          VX_TRY(launch_accumulate(m, nseg, mi.c, A));
        }
      }
      // split flags were computed for the first n_before nodes
      VX_KERNEL(ctx, "k_finish_split", k_finish_split<<<blocks_for(n_before, 256), 256, 0, st>>>(n_before, m->flag_a.as<int32_t>(), m->nodes.as<vx_map_node>(),
                                                                                              m->meta.as<NodeMeta>(), m->loc.as<Mom>(), m->W));
    }
    // tras_opt over the slide set (voxelslam.cpp:1735-1736)
    const int64_t N = m->n_nodes;
    VX_TRY(work_reserve(m, 2 * N + 4));
    VX_KERNEL(ctx, "k_tras_select", k_tras_select<<<blocks_for(N, 256), 256, 0, st>>>(N, m->nodes.as<vx_map_node>(), m->meta.as<NodeMeta>(), m->eig.as<double>(),
                                                                                   m->flag_a.as<int32_t>()));
    VX_TRY(scan_incl(m, m->flag_a.as<int32_t>(), m->flag_b.as<int32_t>(), N));
    int32_t V = 0;
    VX_TRY(read_i32(ctx, m->flag_b.as<int32_t>() + (N - 1), &V));
    if (V == 0) return VX_OK;
    Misc mi = misc_of(m, N);
    VX_KERNEL(ctx, "k_tras_count", k_tras_count<<<blocks_for(N, 256), 256, 0, st>>>(N, m->flag_a.as<int32_t>(), m->flag_b.as<int32_t>(), m->meta.as<NodeMeta>(),
                                                                                 m->loc.as<Mom>(), m->W, mi.a, mi.b));
    VX_TRY(scan_incl(m, mi.b, mi.d, V));
    int32_t nnz = 0;
    VX_TRY(read_i32(ctx, mi.d + (V - 1), &nnz));
    VX_TRY(factor_alloc(f, V, nnz));
    VX_KERNEL(ctx, "k_tras_fill", k_tras_fill<<<blocks_for(V, 128), 128, 0, st>>>(V, nnz, mi.a, mi.b, mi.d, mp_set(m), m->W, m->loc.as<Mom>(), m->fix.as<Mom>(),
                                                                               m->add.as<Mom>(), m->eig.as<double>(), f->off.as<int32_t>(), f->frame.as<int32_t>(),
                                                                               f->mom.as<double>(), f->cnt.as<int32_t>(), f->fix.as<double>(), f->coe.as<double>(),
                                                                               f->lam.as<double>(), f->U.as<double>(), f->mrg.as<double>(), f->win.as<int32_t>()));
    const int32_t wv[2] = {0, V};
    VX_CUDA(cudaMemcpyAsync(f->wv_off.p, wv, sizeof(wv), cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaStreamSynchronize(st));
    return factor_finish_upload(f);
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  d_poses.release();
  return rc;
}

int vx_map_margi(vx_ctx *ctx, vx_map *m, const vx_pose *poses, int win_count, int mgsize, vx_factor *f, double jour) {
  VX_ARG(ctx && m && m->ctx == ctx && poses && f && win_count >= 1 && win_count <= m->W && mgsize >= 1 && mgsize <= MAX_MG && mgsize <= win_count);
  VX_ARG(f->W == m->W && f->n_win == 1);
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  int32_t n_slide = 0;
  VX_TRY(count_slide(m, &n_slide));
  if (n_slide < m->prm.thread_num) return VX_OK;                     // voxelslam.cpp:1616-1617
  DevBuf d_poses;
  const MapPrm P = map_prm(m);
  const int64_t N = m->n_nodes;
  auto body = [&]() -> int {
    VX_TRY(upload_poses(m, poses, win_count, d_poses));
    int32_t *cnt = m->w_cnt.as<int32_t>();
    VX_CUDA(cudaMemsetAsync(cnt, 0, 64, st));
    VX_KERNEL(ctx, "k_slide_jour", k_slide_jour<<<blocks_for(N, 256), 256, 0, st>>>(N, m->meta.as<NodeMeta>(), jour));
    VX_KERNEL(ctx, "k_margi_leaf", k_margi_leaf<<<blocks_for(N, 64), 64, 0, st>>>(N, P, mp_set(m), win_count, mgsize, d_poses.as<PoseSet>(), m->nodes.as<vx_map_node>(),
                                                                               m->meta.as<NodeMeta>(), m->add.as<Mom>(), m->fix.as<Mom>(), m->loc.as<Mom>(),
                                                                               m->cov.as<double>(), m->eig.as<double>(), f->V, f->lam.as<double>(), f->U.as<double>(),
                                                                               f->mrg.as<double>(), m->flag_a.as<int32_t>(), m->flag_b.as<int32_t>(), cnt));
    int32_t err = 0;
    VX_TRY(read_i32(ctx, cnt, &err));
    if (err) { set_error("Error: opt_state: a leaf points past the end of the LidarFactor (voxel_map.hpp:1488-1492)"); return VX_ERR_OPT_STATE; }
    for (int L = m->prm.max_layer - 1; L >= 0; L--)
      VX_KERNEL(ctx, "k_margi_internal", k_margi_internal<<<blocks_for(N, 256), 256, 0, st>>>(N, L, m->nodes.as<vx_map_node>(), m->meta.as<NodeMeta>()));
    // point_fix bookkeeping: drop, then append the marginalised slots' points of the folding leaves
    if (m->n_fix) VX_KERNEL(ctx, "k_clear_fix", k_clear_fix<<<blocks_for(m->n_fix, 256), 256, 0, st>>>(m->n_fix, m->fix_leaf.as<int32_t>(), m->flag_b.as<int32_t>(), cnt + 2));
    for (int k = 0; k < mgsize; k++) {
      const int s = m->mp[k];
      const int64_t n = m->slot_n[s];
      if (n == 0) continue;
      VX_TRY(work_reserve(m, 2 * n + 4));
      VX_KERNEL(ctx, "k_fold_flags", k_fold_flags<<<blocks_for(n, 256), 256, 0, st>>>(m->slot_leaf[s].as<int32_t>(), n, m->flag_a.as<int32_t>(), m->w_flag.as<int32_t>()));
      VX_TRY(scan_incl(m, m->w_flag.as<int32_t>(), m->w_scan.as<int32_t>(), n));
      int32_t keep = 0;
      VX_TRY(read_i32(ctx, m->w_scan.as<int32_t>() + (n - 1), &keep));
      if (keep) {
        VX_TRY(ensure_fix(m, m->n_fix + keep));
        VX_KERNEL(ctx, "k_fold_scatter", k_fold_scatter<<<blocks_for(n, 256), 256, 0, st>>>(m->slot_pts[s].as<vx_point_var>(), m->slot_leaf[s].as<int32_t>(), n,
                                                                                         m->w_flag.as<int32_t>(), m->w_scan.as<int32_t>(), d_poses.as<PoseSet>(), k,
                                                                                         m->fix_pts.as<vx_point_var>(), m->fix_leaf.as<int32_t>(), m->n_fix));
        m->n_fix += keep;
      }
      m->slot_n[s] = 0;                            // sw->points[mp[i]].clear() on every leaf that held points of the slot
    }
    // dead roots leave the slide set and give their SlideWindows back (voxelslam.cpp:1665-1674)
    VX_KERNEL(ctx, "k_erase_mark", k_erase_mark<<<blocks_for(N, 256), 256, 0, st>>>(N, m->meta.as<NodeMeta>(), m->flag_a.as<int32_t>()));
    VX_KERNEL(ctx, "k_erase_apply", k_erase_apply<<<blocks_for(N, 256), 256, 0, st>>>(N, m->meta.as<NodeMeta>(), m->flag_a.as<int32_t>(), m->loc.as<Mom>(), m->W));
    VX_KERNEL(ctx, "k_erase_finish", k_erase_finish<<<blocks_for(N, 256), 256, 0, st>>>(N, m->meta.as<NodeMeta>(), m->flag_a.as<int32_t>()));
    // compact the fixed pool when a quarter of it is dead
    int32_t dead = 0;
    VX_TRY(read_i32(ctx, cnt + 2, &dead));
    static_assert(sizeof(int32_t) == 4, "");
    if (dead > 0 && (int64_t)dead * 4 >= m->n_fix) {
      const int64_t n = m->n_fix;
      VX_TRY(work_reserve(m, 2 * n + 4));
      VX_KERNEL(ctx, "k_live_flags", k_live_flags<<<blocks_for(n, 256), 256, 0, st>>>(m->fix_leaf.as<int32_t>(), n, m->w_flag.as<int32_t>()));
      VX_TRY(scan_incl(m, m->w_flag.as<int32_t>(), m->w_scan.as<int32_t>(), n));
      int32_t live = 0;
      VX_TRY(read_i32(ctx, m->w_scan.as<int32_t>() + (n - 1), &live));
      DevBuf np, nl;
      VX_TRY(np.reserve(sizeof(vx_point_var) * std::max<int64_t>(live, 1) * 3 / 2)); VX_TRY(nl.reserve(sizeof(int32_t) * std::max<int64_t>(live, 1) * 3 / 2));
      VX_KERNEL(ctx, "k_compact_fix", k_compact_fix<<<blocks_for(n, 256), 256, 0, st>>>(m->fix_pts.as<vx_point_var>(), m->fix_leaf.as<int32_t>(), n, m->w_flag.as<int32_t>(),
                                                                                     m->w_scan.as<int32_t>(), np.as<vx_point_var>(), nl.as<int32_t>()));
      VX_CUDA(cudaStreamSynchronize(st));
      m->fix_pts.release(); m->fix_leaf.release();
      m->fix_pts = np; m->fix_leaf = nl; m->n_fix = live;
      m->cap_fix = (int64_t)std::min(np.cap / sizeof(vx_point_var), nl.cap / sizeof(int32_t));
    }
    return VX_OK;
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  d_poses.release();
  return rc;
}

int vx_map_export(vx_map *m, vx_map_node_state *states, vx_map_node *nodes) {
  VX_ARG(m != nullptr);
  vx_ctx *ctx = m->ctx;
  VX_TRY(bind(ctx));
  const int64_t N = m->n_nodes;
  if (N == 0) return VX_OK;
  cudaStream_t st = ctx->stream;
  if (nodes) VX_CUDA(cudaMemcpyAsync(nodes, m->nodes.p, sizeof(vx_map_node) * N, cudaMemcpyDeviceToHost, st));
  if (states) {
    VX_TRY(ctx->scratch2.reserve(sizeof(vx_map_node_state) * N));
    VX_KERNEL(ctx, "k_export", k_export<<<blocks_for(N, 128), 128, 0, st>>>(N, m->nodes.as<vx_map_node>(), m->meta.as<NodeMeta>(), m->add.as<Mom>(), m->fix.as<Mom>(),
                                                                         m->eig.as<double>(), m->cov.as<double>(), ctx->scratch2.as<vx_map_node_state>()));
    VX_CUDA(cudaMemcpyAsync(states, ctx->scratch2.p, sizeof(vx_map_node_state) * N, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    // raw-point counts per node
    std::vector<int32_t> cw(N, 0), cf(N, 0);
    VX_TRY(ctx->scratch.reserve(sizeof(int32_t) * N));
    int32_t *d_c = ctx->scratch.as<int32_t>();
    VX_CUDA(cudaMemsetAsync(d_c, 0, sizeof(int32_t) * N, st));
    for (int s = 0; s < m->W; s++)
      if (m->slot_n[s]) VX_KERNEL(ctx, "k_count_labels", k_count_labels<<<blocks_for(m->slot_n[s], 256), 256, 0, st>>>(m->slot_leaf[s].as<int32_t>(), m->slot_n[s], d_c));
    VX_CUDA(cudaMemcpyAsync(cw.data(), d_c, sizeof(int32_t) * N, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    VX_CUDA(cudaMemsetAsync(d_c, 0, sizeof(int32_t) * N, st));
    if (m->n_fix) VX_KERNEL(ctx, "k_count_labels", k_count_labels<<<blocks_for(m->n_fix, 256), 256, 0, st>>>(m->fix_leaf.as<int32_t>(), m->n_fix, d_c));
    VX_CUDA(cudaMemcpyAsync(cf.data(), d_c, sizeof(int32_t) * N, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    // the reference keeps raw points only below max_layer (voxel_map.hpp:1131-1132, :1151, :1170); deeper labels are bookkeeping
    for (int64_t i = 0; i < N; i++) {
      const bool keeps = states[i].layer < m->prm.max_layer;
      states[i].n_window_pts = keeps ? cw[i] : 0; states[i].n_fix_pts = keeps ? cf[i] : 0;
    }
  }
  VX_CUDA(cudaStreamSynchronize(st));
  return VX_OK;
}

}  // extern "C"
