// vx_lio.cu — LIO ESIKF point-to-plane update over a flattened voxel-plane map:
//   match(feat_map, wld, ...) / OctoTree::match / OctoTree::inside   voxel_map.hpp:2167-2205, :1649-1721, :1836-1848
//   VOXEL_SLAM::lio_state_estimation                                 voxelslam.cpp:962-1098
// The map is what OctoTree::match reads, flattened: an open-addressing hash (packed root key -> root
// node) and a node table {centre, quater_length, children, plane}.  One thread per point: covariance to
// world (voxelslam.cpp:1011), root lookup / cached-leaf fast path (:1020-1028), descent, the float-narrowed
// gates, residual + Jacobian, and a deterministic two-stage reduction of HTH / HTz / n n^T / match_num.
// Everything that decides an ASSOCIATION (wld, var_world, sigma_l, the gates) is evaluated in the
// reference's operation order with explicit round-to-nearest multiplies and adds (no FMA contraction),
// so point->plane association is bit-exact; only the sums over points are re-ordered.
#include "vx_common.cuh"
#include "vx_voxelize.cuh"
#include "vx_frontend.cuh"
#include "vx_planemap.cuh"
#include <cmath>
#include <algorithm>
#include <vector>

using namespace vx;

namespace {

struct M3d { double m[9]; };   // row-major m[r*3+c]
__device__ __forceinline__ double mad_rn(double acc, double a, double b) { return __dadd_rn(acc, __dmul_rn(a, b)); }
// out(r,c) = (A(r,0)B(0,c) + A(r,1)B(1,c)) + A(r,2)B(2,c)   — the straight-line order of a coefficient-wise product
__device__ __forceinline__ M3d mul33(const M3d &A, const M3d &B) {
  M3d o;
#pragma unroll
  for (int r = 0; r < 3; r++)
#pragma unroll
    for (int c = 0; c < 3; c++)
      o.m[r * 3 + c] = mad_rn(mad_rn(__dmul_rn(A.m[r * 3], B.m[c]), A.m[r * 3 + 1], B.m[3 + c]), A.m[r * 3 + 2], B.m[6 + c]);
  return o;
}
__device__ __forceinline__ M3d tr33(const M3d &A) {
  M3d o;
#pragma unroll
  for (int r = 0; r < 3; r++)
#pragma unroll
    for (int c = 0; c < 3; c++) o.m[r * 3 + c] = A.m[c * 3 + r];
  return o;
}
__device__ __forceinline__ void mulv3(const M3d &A, const double *v, double *o) {
#pragma unroll
  for (int r = 0; r < 3; r++) o[r] = mad_rn(mad_rn(__dmul_rn(A.m[r * 3], v[0]), A.m[r * 3 + 1], v[1]), A.m[r * 3 + 2], v[2]);
}
__device__ __forceinline__ M3d hat3(const double *v) {
  M3d o;
  o.m[0] = 0.0; o.m[1] = -v[2]; o.m[2] = v[1];
  o.m[3] = v[2]; o.m[4] = 0.0; o.m[5] = -v[0];
  o.m[6] = -v[1]; o.m[7] = v[0]; o.m[8] = 0.0;
  return o;
}
__device__ __forceinline__ M3d from_colmajor(const double *p) {
  M3d o;
#pragma unroll
  for (int r = 0; r < 3; r++)
#pragma unroll
    for (int c = 0; c < 3; c++) o.m[r * 3 + c] = p[c * 3 + r];
  return o;
}

struct LioPose { M3d R, Rt, rot_var, tsl_var; double p[3]; };

constexpr int NACC = 34;   // HTH upper 21 | HTz 6 | nnt upper 6 | match 1
constexpr int LIO_THREADS = 128;

// OctoTree::match at a leaf (voxel_map.hpp:1652-1691).  Returns 1 and sigma_d on association.
__device__ __forceinline__ int leaf_match(const vx_map_node &nd, const double *wld, const M3d &var_wld, double &sigma_d) {
  if (!nd.is_plane) return 0;
  double d[3] = {__dsub_rn(wld[0], nd.p_center[0]), __dsub_rn(wld[1], nd.p_center[1]), __dsub_rn(wld[2], nd.p_center[2])};
  const double dot = mad_rn(mad_rn(__dmul_rn(nd.p_normal[0], d[0]), nd.p_normal[1], d[1]), nd.p_normal[2], d[2]);
  const float dis_to_plane = (float)fabs(dot);
  const double e0 = __dsub_rn(nd.p_center[0], wld[0]), e1 = __dsub_rn(nd.p_center[1], wld[1]), e2 = __dsub_rn(nd.p_center[2], wld[2]);
  const float dis_to_center = (float)mad_rn(mad_rn(__dmul_rn(e0, e0), e1, e1), e2, e2);
  const float range_dis = __fsub_rn(dis_to_center, __fmul_rn(dis_to_plane, dis_to_plane));
  if (!(range_dis <= __fmul_rn(9.0f, nd.radius))) return 0;
  // J_nq = [ (wld - center)^T , -normal^T ];  sigma_l = J plane_var J^T + n^T var_wld n   (plane_var column-major 6x6)
  const double J[6] = {d[0], d[1], d[2], -nd.p_normal[0], -nd.p_normal[1], -nd.p_normal[2]};
  double t[6];
#pragma unroll
  for (int c = 0; c < 6; c++) {
    double s = __dmul_rn(J[0], nd.plane_var[c * 6 + 0]);
#pragma unroll
    for (int k = 1; k < 6; k++) s = mad_rn(s, J[k], nd.plane_var[c * 6 + k]);
    t[c] = s;
  }
  double sigma_l = __dmul_rn(t[0], J[0]);
#pragma unroll
  for (int k = 1; k < 6; k++) sigma_l = mad_rn(sigma_l, t[k], J[k]);
  double nv[3];
#pragma unroll
  for (int c = 0; c < 3; c++) nv[c] = mad_rn(mad_rn(__dmul_rn(nd.p_normal[0], var_wld.m[c]), nd.p_normal[1], var_wld.m[3 + c]), nd.p_normal[2], var_wld.m[6 + c]);
  sigma_l = __dadd_rn(sigma_l, mad_rn(mad_rn(__dmul_rn(nv[0], nd.p_normal[0]), nv[1], nd.p_normal[1]), nv[2], nd.p_normal[2]));
  if (!((double)dis_to_plane < __dmul_rn(3.0, __dsqrt_rn(sigma_l)))) return 0;
  sigma_d = sigma_l;
  return 1;
}

__global__ void __launch_bounds__(LIO_THREADS)
k_lio_pass(const vx_point_var *__restrict__ pts, int64_t n, LioPose X, const uint64_t *__restrict__ hkeys, const int32_t *__restrict__ hvals,
           uint64_t hmask, const vx_map_node *__restrict__ nodes, double voxel_size, int32_t *__restrict__ leaf_cache,
           int32_t *__restrict__ flags, double *__restrict__ partial) {
  __shared__ double sh[(LIO_THREADS / 32) * NACC];
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  double acc[NACC];
#pragma unroll
  for (int k = 0; k < NACC; k++) acc[k] = 0.0;
  if (i < n) {
    const vx_point_var pv = pts[i];
    const M3d var = from_colmajor(pv.var);
    const M3d phat = hat3(pv.pnt);
    // var_world = R var R^T + phat rot_var phat^T + tsl_var   (voxelslam.cpp:1011)
    const M3d A = mul33(mul33(X.R, var), X.Rt);
    const M3d B = mul33(mul33(phat, X.rot_var), tr33(phat));
    M3d vw;
#pragma unroll
    for (int k = 0; k < 9; k++) vw.m[k] = __dadd_rn(__dadd_rn(A.m[k], B.m[k]), X.tsl_var.m[k]);
    double wld[3];
    mulv3(X.R, pv.pnt, wld);
#pragma unroll
    for (int k = 0; k < 3; k++) wld[k] = __dadd_rn(wld[k], X.p[k]);
    int flag = 0, leaf = -1;
    double sigma_d = 0.0;
    int32_t cached = leaf_cache ? leaf_cache[i] : -1;
    bool use_cache = false;
    if (cached >= 0) {   // OctoTree::inside (voxel_map.hpp:1836-1848)
      const vx_map_node &c = nodes[cached];
      const double hl = (double)__fmul_rn(c.quater_length, 2.0f);
      use_cache = wld[0] >= __dsub_rn(c.center[0], hl) && wld[0] <= __dadd_rn(c.center[0], hl) && wld[1] >= __dsub_rn(c.center[1], hl) &&
                  wld[1] <= __dadd_rn(c.center[1], hl) && wld[2] >= __dsub_rn(c.center[2], hl) && wld[2] <= __dadd_rn(c.center[2], hl);
    }
    int32_t node = -1;
    if (use_cache) node = cached;
    else {
      uint64_t key;
      if (pack_key(key_axis_d(wld[0], voxel_size), key_axis_d(wld[1], voxel_size), key_axis_d(wld[2], voxel_size), key)) {
        uint64_t h = mix64(key) & hmask;
        while (true) {
          const uint64_t k = hkeys[h];
          if (k == key) { node = hvals[h]; break; }
          if (k == ~0ull) break;
          h = (h + 1) & hmask;
        }
      }
    }
    while (node >= 0) {
      const vx_map_node &nd = nodes[node];
      if (nd.octo_state == 0) {
        flag = leaf_match(nd, wld, vw, sigma_d);
        if (flag) leaf = node;
        break;
      }
      const int leafnum = 4 * (wld[0] > nd.center[0]) + 2 * (wld[1] > nd.center[1]) + (wld[2] > nd.center[2]);   // voxel_map.hpp:1695-1712
      node = nd.child[leafnum];
    }
    if (flags) flags[i] = flag;
    if (flag) {
      if (leaf_cache) leaf_cache[i] = leaf;
      const vx_map_node &nd = nodes[leaf];
      const double R_inv = 1.0 / (0.0005 + sigma_d);
      const double d0 = wld[0] - nd.p_center[0], d1 = wld[1] - nd.p_center[1], d2 = wld[2] - nd.p_center[2];
      const double resi = nd.p_normal[0] * d0 + nd.p_normal[1] * d1 + nd.p_normal[2] * d2;
      double jac[6], tmp[3];
      const M3d PR = mul33(phat, X.Rt);
      mulv3(PR, nd.p_normal, tmp);
      jac[0] = tmp[0]; jac[1] = tmp[1]; jac[2] = tmp[2]; jac[3] = nd.p_normal[0]; jac[4] = nd.p_normal[1]; jac[5] = nd.p_normal[2];
      int q = 0;
#pragma unroll
      for (int r = 0; r < 6; r++)
#pragma unroll
        for (int c = r; c < 6; c++) acc[q++] = R_inv * (jac[r] * jac[c]);
#pragma unroll
      for (int r = 0; r < 6; r++) acc[21 + r] = -(R_inv * jac[r] * resi);
      acc[27] = jac[3] * jac[3]; acc[28] = jac[3] * jac[4]; acc[29] = jac[3] * jac[5];
      acc[30] = jac[4] * jac[4]; acc[31] = jac[4] * jac[5]; acc[32] = jac[5] * jac[5];
      acc[33] = 1.0;
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < NACC; k++) {
    double x = acc[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0) sh[warp * NACC + k] = x;
  }
  __syncthreads();
  if (threadIdx.x < NACC) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < LIO_THREADS / 32; w++) s += sh[w * NACC + threadIdx.x];
    partial[(size_t)blockIdx.x * NACC + threadIdx.x] = s;
  }
}
// int match(feat_map, wld, pla, var_wld, sigma_d, oc)  voxel_map.hpp:2167-2205 for a batch of WORLD points with their world
// covariances: root lookup, descent (OctoTree::match :1693-1720), the leaf gates.  flag / leaf node / sigma_d per point.
__global__ void k_match_points(const double *__restrict__ wld3, const double *__restrict__ var9, int64_t n, const uint64_t *__restrict__ hkeys,
                               const int32_t *__restrict__ hvals, uint64_t hmask, const vx_map_node *__restrict__ nodes, double voxel_size,
                               int32_t *__restrict__ flags, int32_t *__restrict__ leaf, double *__restrict__ sigma) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double wld[3] = {wld3[3 * i], wld3[3 * i + 1], wld3[3 * i + 2]};
  const M3d vw = from_colmajor(var9 + 9 * i);
  int flag = 0, lf = -1;
  double sigma_d = 0.0;
  int32_t node = -1;
  uint64_t key;
  if (pack_key(key_axis_d(wld[0], voxel_size), key_axis_d(wld[1], voxel_size), key_axis_d(wld[2], voxel_size), key)) node = root_lookup(hkeys, hvals, hmask, key);
  while (node >= 0) {
    const vx_map_node &nd = nodes[node];
    if (nd.octo_state == 0) {
      flag = leaf_match(nd, wld, vw, sigma_d);
      if (flag) lf = node;
      break;
    }
    node = nd.child[4 * (wld[0] > nd.center[0]) + 2 * (wld[1] > nd.center[1]) + (wld[2] > nd.center[2])];
  }
  flags[i] = flag; leaf[i] = lf; sigma[i] = flag ? sigma_d : 0.0;
}

// fixed-order sum of the block partials// fixed-order sum of the block partials
__global__ void __launch_bounds__(256) k_lio_fin(const double *__restrict__ partial, int nblocks, double *__restrict__ out) {
  __shared__ double sh[256];
  const int k = blockIdx.x;
  double s = 0.0;
  for (int b = threadIdx.x; b < nblocks; b += 256) s += partial[(size_t)b * NACC + k];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) out[k] = sh[0];
}

}  // namespace

namespace {

LioPose make_pose(const vx_pose *pose, const double *rot_var9, const double *tsl_var9) {
  LioPose X;
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) {
      X.R.m[r * 3 + c] = pose->R[c * 3 + r];
      X.Rt.m[r * 3 + c] = pose->R[r * 3 + c];
      X.rot_var.m[r * 3 + c] = rot_var9[c * 3 + r];
      X.tsl_var.m[r * 3 + c] = tsl_var9[c * 3 + r];
    }
  for (int k = 0; k < 3; k++) X.p[k] = pose->p[k];
  return X;
}

// device-resident pass: d_pts, d_cache, d_flags on the device; result -> host out[NACC]
int lio_pass_device(vx_ctx *ctx, vx_planemap *map, const vx_point_var *d_pts, int64_t n, const LioPose &X, int32_t *d_cache, int32_t *d_flags,
                    double *out34) {
  const int nblocks = blocks_for(n, LIO_THREADS);
  VX_TRY(ctx->scratch.reserve(sizeof(double) * ((size_t)nblocks + 1) * NACC));
  double *partial = ctx->scratch.as<double>();
  double *fin = partial + (size_t)nblocks * NACC;
  VX_KERNEL(ctx, "k_lio_pass", k_lio_pass<<<nblocks, LIO_THREADS, 0, ctx->stream>>>(d_pts, n, X, map->hkeys.as<uint64_t>(), map->hvals.as<int32_t>(), map->hmask,
                                                                                   map->nodes.as<vx_map_node>(), map->voxel_size, d_cache, d_flags, partial));
  VX_KERNEL(ctx, "k_lio_fin", k_lio_fin<<<NACC, 256, 0, ctx->stream>>>(partial, nblocks, fin));
  // through the context's pinned staging area: a copy into pageable memory is staged by the driver and costs a second wait
  static_assert(sizeof(double) * NACC <= (1 << 16), "pinned staging area too small");
  VX_CUDA(cudaMemcpyAsync(ctx->pinned, fin, sizeof(double) * NACC, cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  memcpy(out34, ctx->pinned, sizeof(double) * NACC);
  return VX_OK;
}
void unpack_acc(const double *a, double *HTH36, double *HTz6, double *nnt9, int32_t *match_num) {
  int q = 0;
  for (int r = 0; r < 6; r++)
    for (int c = r; c < 6; c++) { HTH36[c * 6 + r] = a[q]; HTH36[r * 6 + c] = a[q]; q++; }
  for (int r = 0; r < 6; r++) HTz6[r] = a[21 + r];
  const int ix[6][2] = {{0, 0}, {0, 1}, {0, 2}, {1, 1}, {1, 2}, {2, 2}};
  for (int k = 0; k < 6; k++) { nnt9[ix[k][1] * 3 + ix[k][0]] = a[27 + k]; nnt9[ix[k][0] * 3 + ix[k][1]] = a[27 + k]; }
  *match_num = (int32_t)llround(a[33]);
}

// ---- 15x15 host algebra of lio_state_estimation (column-major) ------------------------------------------
struct M15h { double d[225]; double &operator()(int r, int c) { return d[c * 15 + r]; } double operator()(int r, int c) const { return d[c * 15 + r]; } };
// inverse by partial-pivot LU (the path Eigen takes for fixed sizes > 4, voxelslam.cpp:987 / :1054)
M15h inverse15(const M15h &Ain) {
  const int N = 15;
  M15h A = Ain, inv;
  int perm[N];
  for (int i = 0; i < N; i++) perm[i] = i;
  for (int k = 0; k < N; k++) {
    int p = k; double best = std::fabs(A(k, k));
    for (int i = k + 1; i < N; i++) if (std::fabs(A(i, k)) > best) { best = std::fabs(A(i, k)); p = i; }
    if (p != k) { for (int c = 0; c < N; c++) std::swap(A(k, c), A(p, c)); std::swap(perm[k], perm[p]); }
    const double piv = A(k, k);
    for (int i = k + 1; i < N; i++) A(i, k) /= piv;
    for (int c = k + 1; c < N; c++) { const double akc = A(k, c); for (int i = k + 1; i < N; i++) A(i, c) -= A(i, k) * akc; }
  }
  for (int c = 0; c < N; c++) {
    double y[N];
    for (int i = 0; i < N; i++) y[i] = (perm[i] == c) ? 1.0 : 0.0;
    for (int i = 0; i < N; i++) for (int j = 0; j < i; j++) y[i] -= A(i, j) * y[j];
    for (int i = N - 1; i >= 0; i--) { for (int j = i + 1; j < N; j++) y[i] -= A(i, j) * y[j]; y[i] /= A(i, i); }
    for (int i = 0; i < N; i++) inv(i, c) = y[i];
  }
  return inv;
}
void mat3_mul(const double *A, const double *B, double *C) {   // column-major 3x3
  for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) { double s = 0; for (int k = 0; k < 3; k++) s += A[k * 3 + r] * B[c * 3 + k]; C[c * 3 + r] = s; }
}
void so3_exp(const double *w, double *E) {   // tools.hpp:51-66
  const double n = std::sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  for (int k = 0; k < 9; k++) E[k] = (k % 4 == 0) ? 1.0 : 0.0;
  if (n >= 1e-11) {
    const double x = w[0] / n, y = w[1] / n, z = w[2] / n;
    const double K[9] = {0, z, -y, -z, 0, x, y, -x, 0};
    double KK[9];
    mat3_mul(K, K, KK);
    const double s = std::sin(n), c1 = 1.0 - std::cos(n);
    for (int k = 0; k < 9; k++) E[k] += s * K[k] + c1 * KK[k];
  }
}
void so3_log(const double *R, double *w) {   // tools.hpp:86-91
  const double tr = R[0] + R[4] + R[8];
  const double theta = (tr > 3.0 - 1e-6) ? 0.0 : std::acos(0.5 * (tr - 1));
  const double K[3] = {R[5] - R[7], R[6] - R[2], R[1] - R[3]};   // R(2,1)-R(1,2), R(0,2)-R(2,0), R(1,0)-R(0,1)
  const double f = (std::fabs(theta) < 0.001) ? 0.5 : 0.5 * theta / std::sin(theta);
  for (int k = 0; k < 3; k++) w[k] = f * K[k];
}
// smallest eigenvalue of a symmetric 3x3 (host Jacobi; voxelslam.cpp:1090-1097 only compares it with 14)
double min_eig3(const double *A9) {
  double a[3][3] = {{A9[0], A9[3], A9[6]}, {A9[1], A9[4], A9[7]}, {A9[2], A9[5], A9[8]}};
  for (int sweep = 0; sweep < 30; sweep++) {
    const double off = std::fabs(a[0][1]) + std::fabs(a[0][2]) + std::fabs(a[1][2]);
    if (off < 1e-300) break;
    for (int p = 0; p < 2; p++)
      for (int q = p + 1; q < 3; q++) {
        if (a[p][q] == 0.0) continue;
        const double th = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
        double t = 1.0 / (std::fabs(th) + std::sqrt(th * th + 1.0));
        if (th < 0) t = -t;
        const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < 3; k++) { const double x = a[k][p], y = a[k][q]; a[k][p] = c * x - s * y; a[k][q] = s * x + c * y; }
        for (int k = 0; k < 3; k++) { const double x = a[p][k], y = a[q][k]; a[p][k] = c * x - s * y; a[q][k] = s * x + c * y; }
      }
  }
  return std::min(a[0][0], std::min(a[1][1], a[2][2]));
}

}  // namespace

extern "C" {

int vx_planemap_create(vx_ctx *ctx, const int64_t *root_keys, const int32_t *root_node, int64_t n_roots, const vx_map_node *nodes, int64_t n_nodes,
                       double voxel_size, vx_planemap **out) {
  VX_ARG(out != nullptr && n_roots >= 0 && n_nodes >= 0 && voxel_size > 0);
  VX_ARG(n_roots == 0 || (root_keys && root_node && nodes));
  *out = nullptr;
  VX_TRY(bind(ctx));
  uint64_t cap = 16;
  while (cap < (uint64_t)n_roots * 2 + 2) cap <<= 1;
  std::vector<uint64_t> hk(cap, ~0ull);
  std::vector<int32_t> hv(cap, -1);
  for (int64_t r = 0; r < n_roots; r++) {
    uint64_t key;
    if (!pack_key(root_keys[3 * r], root_keys[3 * r + 1], root_keys[3 * r + 2], key)) { set_error("root voxel %lld does not fit the 21-bit packed key", (long long)r); return VX_ERR_ARG; }
    VX_ARG(root_node[r] >= 0 && root_node[r] < n_nodes);
    uint64_t h = mix64(key) & (cap - 1);
    while (hk[h] != ~0ull) { if (hk[h] == key) { set_error("duplicate root voxel"); return VX_ERR_ARG; } h = (h + 1) & (cap - 1); }
    hk[h] = key; hv[h] = root_node[r];
  }
  vx_planemap *m = new vx_planemap();
  m->ctx = ctx; m->hmask = cap - 1; m->n_nodes = n_nodes; m->n_roots = n_roots; m->voxel_size = voxel_size;
  auto body = [&]() -> int {
    VX_TRY(m->hkeys.reserve(8 * cap)); VX_TRY(m->hvals.reserve(4 * cap)); VX_TRY(m->nodes.reserve(sizeof(vx_map_node) * std::max<int64_t>(n_nodes, 1)));
    VX_CUDA(cudaMemcpyAsync(m->hkeys.p, hk.data(), 8 * cap, cudaMemcpyHostToDevice, ctx->stream));
    VX_CUDA(cudaMemcpyAsync(m->hvals.p, hv.data(), 4 * cap, cudaMemcpyHostToDevice, ctx->stream));
    if (n_nodes) VX_CUDA(cudaMemcpyAsync(m->nodes.p, nodes, sizeof(vx_map_node) * n_nodes, cudaMemcpyHostToDevice, ctx->stream));
    VX_CUDA(cudaStreamSynchronize(ctx->stream));
    return VX_OK;
  };
  const int s = body();
  if (s != VX_OK) { m->hkeys.release(); m->hvals.release(); m->nodes.release(); delete m; return s; }
  *out = m;
  return VX_OK;
}

int vx_planemap_free(vx_planemap *m) {
  if (!m) return VX_OK;
  if (!m->owns) { set_error("vx_planemap_free: this handle is a view owned by a vx_map"); return VX_ERR_ARG; }
  cudaSetDevice(m->ctx->device);
  cudaStreamSynchronize(m->ctx->stream);
  m->hkeys.release(); m->hvals.release(); m->nodes.release();
  delete m;
  return VX_OK;
}

int vx_planemap_match(vx_ctx *ctx, vx_planemap *map, const double *wld3, const double *var_wld9, int64_t n, int32_t *flags, int32_t *leaf, double *sigma_d,
                      vx_map_node *planes) {
  VX_ARG(map != nullptr && n >= 0 && (n == 0 || (wld3 && var_wld9 && flags)));
  VX_TRY(bind(ctx));
  if (n == 0) return VX_OK;
  cudaStream_t st = ctx->stream;
  DevBuf d_w, d_v, d_f, d_l, d_s;
  std::vector<int32_t> lf(n);
  auto body = [&]() -> int {
    VX_TRY(d_w.reserve(24 * n)); VX_TRY(d_v.reserve(72 * n)); VX_TRY(d_f.reserve(4 * n)); VX_TRY(d_l.reserve(4 * n)); VX_TRY(d_s.reserve(8 * n));
    VX_CUDA(cudaMemcpyAsync(d_w.p, wld3, 24 * n, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemcpyAsync(d_v.p, var_wld9, 72 * n, cudaMemcpyHostToDevice, st));
    VX_KERNEL(ctx, "k_match_points", k_match_points<<<blocks_for(n, 128), 128, 0, st>>>(d_w.as<double>(), d_v.as<double>(), n, map->hkeys.as<uint64_t>(), map->hvals.as<int32_t>(),
                                                                                     map->hmask, map->nodes.as<vx_map_node>(), map->voxel_size, d_f.as<int32_t>(),
                                                                                     d_l.as<int32_t>(), d_s.as<double>()));
    VX_CUDA(cudaMemcpyAsync(flags, d_f.p, 4 * n, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaMemcpyAsync(lf.data(), d_l.p, 4 * n, cudaMemcpyDeviceToHost, st));
    if (sigma_d) VX_CUDA(cudaMemcpyAsync(sigma_d, d_s.p, 8 * n, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    if (planes)        // the Plane the reference hands back through `pla` (a copy of the matched leaf's record)
      for (int64_t k = 0; k < n; k++)
        if (lf[k] >= 0) VX_CUDA(cudaMemcpyAsync(&planes[k], map->nodes.as<vx_map_node>() + lf[k], sizeof(vx_map_node), cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    if (leaf) std::copy(lf.begin(), lf.end(), leaf);
    return VX_OK;
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  for (DevBuf *b : {&d_w, &d_v, &d_f, &d_l, &d_s}) b->release();
  return rc;
}

int vx_lio_match_accumulate(vx_ctx *ctx, vx_planemap *map, const vx_point_var *pts, int64_t n, const vx_pose *pose, const double *rot_var9,
                            const double *tsl_var9, int32_t *leaf_cache, double *HTH36, double *HTz6, double *nnt9, int32_t *match_num, int32_t *flags) {
  VX_ARG(map != nullptr && pose != nullptr && rot_var9 && tsl_var9 && HTH36 && HTz6 && nnt9 && match_num && n >= 0 && (n == 0 || pts));
  VX_TRY(bind(ctx));
  double acc[NACC] = {0};
  if (n > 0) {
    DevBuf d_pts, d_cache, d_flags;
    auto body = [&]() -> int {
      VX_TRY(d_pts.reserve(sizeof(vx_point_var) * n)); VX_TRY(d_cache.reserve(4 * n)); VX_TRY(d_flags.reserve(4 * n));
      VX_CUDA(cudaMemcpyAsync(d_pts.p, pts, sizeof(vx_point_var) * n, cudaMemcpyHostToDevice, ctx->stream));
      if (leaf_cache) VX_CUDA(cudaMemcpyAsync(d_cache.p, leaf_cache, 4 * n, cudaMemcpyHostToDevice, ctx->stream));
      else VX_CUDA(cudaMemsetAsync(d_cache.p, 0xff, 4 * n, ctx->stream));
      VX_TRY(lio_pass_device(ctx, map, d_pts.as<vx_point_var>(), n, make_pose(pose, rot_var9, tsl_var9), d_cache.as<int32_t>(), d_flags.as<int32_t>(), acc));
      if (leaf_cache) VX_CUDA(cudaMemcpyAsync(leaf_cache, d_cache.p, 4 * n, cudaMemcpyDeviceToHost, ctx->stream));
      if (flags) VX_CUDA(cudaMemcpyAsync(flags, d_flags.p, 4 * n, cudaMemcpyDeviceToHost, ctx->stream));
      VX_CUDA(cudaStreamSynchronize(ctx->stream));
      return VX_OK;
    };
    const int s = body();
    cudaStreamSynchronize(ctx->stream);
    d_pts.release(); d_cache.release(); d_flags.release();
    VX_TRY(s);
  }
  unpack_acc(acc, HTH36, HTz6, nnt9, match_num);
  return VX_OK;
}

}  // extern "C"

namespace {
// pts != NULL: points with covariances from the host; else raw xyz + var_init on the device
int lio_state_estimation_impl(vx_ctx *ctx, vx_planemap *map, const vx_point_var *pts, const float *xyz, const vx_pose *ext, double dept_err,
                              double beam_err, int64_t n, vx_lio_state *x, int32_t *healthy, int32_t *iters, int32_t *match_nums4) {
  VX_ARG(map != nullptr && x != nullptr && n >= 0 && (n == 0 || pts || (xyz && ext)));
  VX_TRY(bind(ctx));
  DevBuf d_pts, d_cache, d_xyz;
  int rc = VX_OK;
  double nnt[9] = {0};
  int n_it = 0;
  auto body = [&]() -> int {
    VX_TRY(d_pts.reserve(sizeof(vx_point_var) * std::max<int64_t>(n, 1))); VX_TRY(d_cache.reserve(4 * std::max<int64_t>(n, 1)));
    if (n && pts) VX_CUDA(cudaMemcpyAsync(d_pts.p, pts, sizeof(vx_point_var) * n, cudaMemcpyHostToDevice, ctx->stream));
    if (n && !pts) {
      VX_TRY(d_xyz.reserve(sizeof(float) * 3 * n));
      VX_CUDA(cudaMemcpyAsync(d_xyz.p, xyz, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, ctx->stream));
      VX_TRY(var_init_device(ctx, d_xyz.as<float>(), n, *ext, dept_err, beam_err, d_pts.as<vx_point_var>()));
    }
    VX_CUDA(cudaMemsetAsync(d_cache.p, 0xff, 4 * std::max<int64_t>(n, 1), ctx->stream));
    const vx_lio_state x_prop = *x;
    const int num_max_iter = 4;   // voxelslam.cpp:968
    int rematch_num = 0;
    M15h cov, cov_inv, G, HTH15;
    std::copy(x->cov, x->cov + 225, cov.d);
    cov_inv = inverse15(cov);
    std::fill(G.d, G.d + 225, 0.0);
    std::fill(HTH15.d, HTH15.d + 225, 0.0);
    for (int it = 0; it < num_max_iter; it++) {
      double acc[NACC] = {0}, HTH[36], HTz[6];
      int32_t mn = 0;
      vx_pose ps;
      std::copy(x->R, x->R + 9, ps.R); std::copy(x->p, x->p + 3, ps.p);
      double rot_var[9], tsl_var[9];
      for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) { rot_var[c * 3 + r] = x->cov[c * 15 + r]; tsl_var[c * 3 + r] = x->cov[(3 + c) * 15 + (3 + r)]; }
      if (n) VX_TRY(lio_pass_device(ctx, map, d_pts.as<vx_point_var>(), n, make_pose(&ps, rot_var, tsl_var), d_cache.as<int32_t>(), nullptr, acc));
      unpack_acc(acc, HTH, HTz, nnt, &mn);
      if (match_nums4) match_nums4[it] = mn;
      n_it = it + 1;
      for (int c = 0; c < 6; c++) for (int r = 0; r < 6; r++) HTH15(r, c) = HTH[c * 6 + r];
      M15h S;
      for (int k = 0; k < 225; k++) S.d[k] = HTH15.d[k] + cov_inv.d[k];
      const M15h K1 = inverse15(S);
      // G.block<15,6>(0,0) = K_1.block<15,6>(0,0) * HTH   (voxelslam.cpp:1055)
      for (int c = 0; c < 6; c++) for (int r = 0; r < 15; r++) { double s = 0; for (int k = 0; k < 6; k++) s += K1(r, k) * HTH[c * 6 + k]; G(r, c) = s; }
      // vec = x_prop - x_curr   (IMUST::operator-, tools.hpp:164-173)
      double vec[15], RtR[9], Rt[9];
      for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) Rt[c * 3 + r] = x->R[r * 3 + c];
      mat3_mul(Rt, x_prop.R, RtR);
      so3_log(RtR, vec);
      for (int k = 0; k < 3; k++) { vec[3 + k] = x_prop.p[k] - x->p[k]; vec[6 + k] = x_prop.v[k] - x->v[k]; vec[9 + k] = x_prop.bg[k] - x->bg[k]; vec[12 + k] = x_prop.ba[k] - x->ba[k]; }
      // solution = K_1(:,0:6) HTz + vec - G(:,0:6) vec(0:6)
      double sol[15];
      for (int r = 0; r < 15; r++) {
        double a = 0, b = 0;
        for (int k = 0; k < 6; k++) a += K1(r, k) * HTz[k];
        for (int k = 0; k < 6; k++) b += G(r, k) * vec[k];
        sol[r] = a + vec[r] - b;
      }
      // x_curr += solution   (IMUST::operator+=, tools.hpp:154-162)
      double E[9], Rn[9];
      so3_exp(sol, E);
      mat3_mul(x->R, E, Rn);
      std::copy(Rn, Rn + 9, x->R);
      for (int k = 0; k < 3; k++) { x->p[k] += sol[3 + k]; x->v[k] += sol[6 + k]; x->bg[k] += sol[9 + k]; x->ba[k] += sol[12 + k]; }
      const double rn = std::sqrt(sol[0] * sol[0] + sol[1] * sol[1] + sol[2] * sol[2]);
      const double tn = std::sqrt(sol[3] * sol[3] + sol[4] * sol[4] + sol[5] * sol[5]);
      const bool converged = (rn * 57.3 < 0.01) && (tn * 100 < 0.015);   // voxelslam.cpp:1069
      if (converged || (rematch_num == 0 && it == num_max_iter - 2)) rematch_num++;
      if (rematch_num >= 2 || it == num_max_iter - 1) {
        // x_curr.cov = (I - G) * x_curr.cov   (voxelslam.cpp:1081)
        M15h IG, nc;
        for (int c = 0; c < 15; c++) for (int r = 0; r < 15; r++) IG(r, c) = (r == c ? 1.0 : 0.0) - G(r, c);
        for (int c = 0; c < 15; c++) for (int r = 0; r < 15; r++) { double s = 0; for (int k = 0; k < 15; k++) s += IG(r, k) * cov(k, c); nc(r, c) = s; }
        std::copy(nc.d, nc.d + 225, x->cov);
        break;
      }
    }
    return VX_OK;
  };
  rc = body();
  cudaStreamSynchronize(ctx->stream);
  d_pts.release(); d_cache.release(); d_xyz.release();
  VX_TRY(rc);
  if (iters) *iters = n_it;
  if (healthy) *healthy = !(min_eig3(nnt) < 14) ? 1 : 0;   // voxelslam.cpp:1090-1097
  return VX_OK;
}
}  // namespace

extern "C" {

int vx_lio_state_estimation(vx_ctx *ctx, vx_planemap *map, const vx_point_var *pts, int64_t n, vx_lio_state *x, int32_t *healthy, int32_t *iters,
                            int32_t *match_nums4) {
  VX_ARG(n == 0 || pts);
  return lio_state_estimation_impl(ctx, map, pts, nullptr, nullptr, 0, 0, n, x, healthy, iters, match_nums4);
}

int vx_lio_state_estimation_xyz(vx_ctx *ctx, vx_planemap *map, const float *xyz, int64_t n, const vx_pose *ext, double dept_err, double beam_err,
                                vx_lio_state *x, int32_t *healthy, int32_t *iters, int32_t *match_nums4) {
  VX_ARG(ext != nullptr && (n == 0 || xyz));
  return lio_state_estimation_impl(ctx, map, nullptr, xyz, ext, dept_err, beam_err, n, x, healthy, iters, match_nums4);
}

}  // extern "C"
