// ORACLE — TEST INFRASTRUCTURE ONLY (see vmath.hpp header). PARITY UNPINNED.
//
// capi.cpp — flat C entry points over the oracle so tests/ and bench.py's cpu_baseline leg can
// drive it through ctypes.  POD layouts deliberately equal the product's (include/voxelba.h) so the
// same numpy arrays feed both sides; the oracle does NOT include the product header.
#include "ba.hpp"
#include "li_ba.hpp"
#include "octree.hpp"
#include "gba.hpp"
#include "lio.hpp"
#include "icp.hpp"
#include "session.hpp"
#include "btc.hpp"
#include "deskew.hpp"
#include <chrono>

using namespace orc;

template <class A> static void cp_in(A &dst, const double *src, int n) { for (int k = 0; k < n; k++) dst[k] = src[k]; }
template <class A> static void cp_out(const A &src, double *dst, int n) { for (int k = 0; k < n; k++) dst[k] = src[k]; }

extern "C" {

struct oc_cluster { double P[9]; double v[3]; int32_t N; int32_t _pad; };
struct oc_pose { double R[9]; double p[3]; };

static PointCluster from_c(const oc_cluster &c) {
  PointCluster r;
  for (int i = 0; i < 9; i++) r.P[i] = c.P[i];
  for (int i = 0; i < 3; i++) r.v[i] = c.v[i];
  r.N = c.N;
  return r;
}
static void to_c(const PointCluster &r, oc_cluster &c) {
  for (int i = 0; i < 9; i++) c.P[i] = r.P[i];
  for (int i = 0; i < 3; i++) c.v[i] = r.v[i];
  c.N = r.N; c._pad = 0;
}
static std::vector<IMUST> poses_in(const oc_pose *p, int n) {
  std::vector<IMUST> xs(n);
  for (int i = 0; i < n; i++) {
    for (int k = 0; k < 9; k++) xs[i].R[k] = p[i].R[k];
    for (int k = 0; k < 3; k++) xs[i].p[k] = p[i].p[k];
  }
  return xs;
}
static void poses_out(const std::vector<IMUST> &xs, oc_pose *p) {
  for (size_t i = 0; i < xs.size(); i++) {
    for (int k = 0; k < 9; k++) p[i].R[k] = xs[i].R[k];
    for (int k = 0; k < 3; k++) p[i].p[k] = xs[i].p[k];
  }
}

// ---- A1 -------------------------------------------------------------------------------------
void orc_keys_f64(const double *xyz, int64_t n, double voxel_size, int64_t *out) {
  for (int64_t i = 0; i < 3 * n; i++) out[i] = key_axis_d(xyz[i], voxel_size);
}
void orc_keys_f32(const float *xyz, int64_t n, double voxel_size, int64_t *out) {
  for (int64_t i = 0; i < 3 * n; i++) out[i] = key_axis_f(xyz[i], voxel_size);
}

// ---- linear algebra -------------------------------------------------------------------------
void orc_eig3(const double *A, int64_t n, double *w, double *U) {
  for (int64_t i = 0; i < n; i++) {
    M3 a; for (int k = 0; k < 9; k++) a[k] = A[9 * i + k];
    V3 l; M3 u;
    eig3_sym(a, l, u);
    for (int k = 0; k < 3; k++) w[3 * i + k] = l[k];
    for (int k = 0; k < 9; k++) U[9 * i + k] = u[k];
  }
}
void orc_ldlt_solve(const double *A, const double *b, int n, double *x) {
  MatX m(n, n); VecX vb(n), vx(n);
  std::memcpy(m.d.data(), A, sizeof(double) * n * n);
  for (int i = 0; i < n; i++) vb[i] = b[i];
  ldlt_solve(m, vb, vx);
  for (int i = 0; i < n; i++) x[i] = vx[i];
}
void orc_inverse15(const double *A, double *out) {
  M15 a; for (int k = 0; k < 225; k++) a[k] = A[k];
  M15 r = inverse_lu(a);
  for (int k = 0; k < 225; k++) out[k] = r[k];
}
void orc_exp(const double *w, double *R) { M3 r = Exp(vec3(w[0], w[1], w[2])); for (int k = 0; k < 9; k++) R[k] = r[k]; }
void orc_log(const double *R, double *w) { M3 r; for (int k = 0; k < 9; k++) r[k] = R[k]; V3 l = Log(r); for (int k = 0; k < 3; k++) w[k] = l[k]; }
void orc_cluster_transform(const oc_cluster *in, const oc_pose *pose, oc_cluster *out) {
  std::vector<IMUST> xs = poses_in(pose, 1);
  PointCluster r; r.transform(from_c(*in), xs[0]);
  to_c(r, *out);
}

// ---- B0..B3 ---------------------------------------------------------------------------------
void *orc_factor_create(int win_size) { return new LidarFactor(win_size); }
void orc_factor_free(void *h) { delete (LidarFactor *)h; }
int64_t orc_factor_num_voxels(void *h) { return (int64_t)((LidarFactor *)h)->plvec_voxels.size(); }
int orc_factor_win_size(void *h) { return ((LidarFactor *)h)->win_size; }
void orc_factor_set_dense(void *h, int64_t V, const oc_cluster *plvec, const oc_cluster *sig, const double *coe,
                          const double *eigval, const double *eigvec, const oc_cluster *pcr_adds) {
  LidarFactor &f = *(LidarFactor *)h;
  f.clear();
  const int W = f.win_size;
  for (int64_t v = 0; v < V; v++) {
    std::vector<PointCluster> pcrs(W);
    for (int i = 0; i < W; i++) pcrs[i] = from_c(plvec[v * W + i]);
    V3 ev; M3 evec;
    for (int k = 0; k < 3; k++) ev[k] = eigval[3 * v + k];
    for (int k = 0; k < 9; k++) evec[k] = eigvec[9 * v + k];
    f.push_voxel(pcrs, from_c(sig[v]), coe[v], ev, evec, from_c(pcr_adds[v]));
  }
}
void orc_factor_get_dense(void *h, oc_cluster *plvec, oc_cluster *sig, double *coe, double *eigval, double *eigvec, oc_cluster *pcr_adds) {
  LidarFactor &f = *(LidarFactor *)h;
  const int W = f.win_size;
  for (size_t v = 0; v < f.plvec_voxels.size(); v++) {
    if (plvec) for (int i = 0; i < W; i++) to_c(f.plvec_voxels[v][i], plvec[v * W + i]);
    if (sig) to_c(f.sig_vecs[v], sig[v]);
    if (coe) coe[v] = f.coeffs[v];
    if (eigval) for (int k = 0; k < 3; k++) eigval[3 * v + k] = f.eig_values[v][k];
    if (eigvec) for (int k = 0; k < 9; k++) eigvec[9 * v + k] = f.eig_vectors[v][k];
    if (pcr_adds) to_c(f.pcr_adds[v], pcr_adds[v]);
  }
}
// number of (voxel, frame) pairs with N != 0
int64_t orc_factor_num_voxel_frames(void *h) {
  LidarFactor &f = *(LidarFactor *)h;
  int64_t n = 0;
  for (auto &pv : f.plvec_voxels) for (auto &c : pv) if (c.N != 0) n++;
  return n;
}
void orc_factor_hessian(void *h, const oc_pose *poses, int64_t head, int64_t end, double *Hess, double *JacT, double *residual) {
  LidarFactor &f = *(LidarFactor *)h;
  std::vector<IMUST> xs = poses_in(poses, f.win_size);
  int jl = 6 * f.win_size;
  MatX H(jl, jl); VecX J(jl); double r = 0;
  f.acc_evaluate2(xs, (int)head, (int)end, H, J, r);
  std::memcpy(Hess, H.d.data(), sizeof(double) * jl * jl);
  for (int i = 0; i < jl; i++) JacT[i] = J[i];
  *residual = r;
}
void orc_factor_residual(void *h, const oc_pose *poses, int64_t head, int64_t end, double *residual) {
  LidarFactor &f = *(LidarFactor *)h;
  std::vector<IMUST> xs = poses_in(poses, f.win_size);
  double r = 0;
  f.evaluate_only_residual(xs, (int)head, (int)end, r);
  *residual = r;
}
// Lidar_BA_Optimizer::divide_thread with `thd_num` ranges, optionally on real threads (CPU baseline)
double orc_ba_divide_thread(void *h, const oc_pose *poses, int thd_num, int use_threads, double *Hess, double *JacT) {
  LidarFactor &f = *(LidarFactor *)h;
  std::vector<IMUST> xs = poses_in(poses, f.win_size);
  Lidar_BA_Optimizer opt;
  opt.win_size = f.win_size; opt.jac_leng = 6 * f.win_size; opt.thd_num = thd_num; opt.use_threads = use_threads != 0;
  MatX H(opt.jac_leng, opt.jac_leng); VecX J(opt.jac_leng);
  double r = opt.divide_thread(xs, f, H, J);
  if (Hess) std::memcpy(Hess, H.d.data(), sizeof(double) * opt.jac_leng * opt.jac_leng);
  if (JacT) for (int i = 0; i < opt.jac_leng; i++) JacT[i] = J[i];
  return r;
}
// returns 1 converged, 0 not, -3 where the reference exits ("Too Less Voxel")
int orc_ba_lidar(void *h, oc_pose *poses, double *hess_out, double *resis2, int max_iter, int thd_num, int use_threads,
                 int *accept_trace, int *n_trace) {
  LidarFactor &f = *(LidarFactor *)h;
  std::vector<IMUST> xs = poses_in(poses, f.win_size);
  Lidar_BA_Optimizer opt;
  opt.thd_num = thd_num; opt.use_threads = use_threads != 0;
  MatX hess; std::vector<double> resis;
  bool fatal = false;
  bool conv = opt.damping_iter(xs, f, &hess, resis, max_iter, false, &fatal);
  if (fatal) return -3;
  poses_out(xs, poses);
  if (hess_out) std::memcpy(hess_out, hess.d.data(), sizeof(double) * hess.d.size());
  if (resis2) { resis2[0] = resis[0]; resis2[1] = resis[1]; }
  if (n_trace) {
    *n_trace = (int)opt.accept_trace.size();
    if (accept_trace) for (size_t i = 0; i < opt.accept_trace.size(); i++) accept_trace[i] = opt.accept_trace[i];
  }
  return conv ? 1 : 0;
}

// ---- A2 ------------------------------------------------------------------------------------
int64_t orc_downsample_close(const PointXYZIN *pts, int64_t n, double voxel_size, PointXYZIN *out) {
  std::vector<PointXYZIN> pl(pts, pts + n);
  down_sampling_close(pl, voxel_size);
  std::copy(pl.begin(), pl.end(), out);
  return (int64_t)pl.size();
}
int64_t orc_downsample_pvec(const double *pv, int64_t n, double voxel_size, PointXYZIN *out) {
  PVec pvec(n);
  for (int64_t i = 0; i < n; i++) { for (int k = 0; k < 3; k++) pvec[i].pnt[k] = pv[12 * i + k]; for (int k = 0; k < 9; k++) pvec[i].var[k] = pv[12 * i + 3 + k]; }
  std::vector<PointXYZIN> keep;
  down_sampling_pvec(pvec, voxel_size, keep);
  std::copy(keep.begin(), keep.end(), out);
  return (int64_t)keep.size();
}
int64_t orc_downsample_voxel(const PointXYZIN *pts, int64_t n, double voxel_size, PointXYZIN *out) {
  std::vector<PointXYZIN> pl(pts, pts + n);
  down_sampling_voxel(pl, voxel_size);
  for (size_t i = 0; i < pl.size(); i++) out[i] = pl[i];
  return (int64_t)pl.size();
}

// ---- A8 + A10: OctreeGBA::cut_voxel per keyframe + OctreeGBA_multi_recut -> a new LidarFactor --
// gba_params: {gba_voxel_size, gba_min_eigen_value, eigen_value_array[4] (stored = 1/yaml), max_layer}
void *orc_gba_build(const float *xyz, const int64_t *cloud_off, const oc_pose *poses, int W, double gba_voxel_size,
                    double gba_min_eigen_value, const double *eig_array4, int max_layer) {
  Params prm;
  prm.gba_voxel_size = gba_voxel_size;
  prm.gba_min_eigen_value = gba_min_eigen_value;
  for (int k = 0; k < 4; k++) prm.gba_eigen_value_array[k] = eig_array4[k];
  prm.max_layer = max_layer;
  std::vector<IMUST> xs = poses_in(poses, W);
  GbaMap oct_map;
  for (int i = 0; i < W; i++)
    gba_cut_voxel(oct_map, xs[i], xyz + 3 * cloud_off[i], (int)(cloud_off[i + 1] - cloud_off[i]), i, W, &prm);
  LidarFactor *f = new LidarFactor(W);
  OctreeGBA_multi_recut(oct_map, *f, 1);
  return f;
}

// ---- C1h: HBA_add_edge for one window.  local3 = {voxel_size, min_eigen_value} + plane_eigen_value_thre[4];
// gba = {gba_voxel_size, gba_min_eigen_value} + eig_array[4]; returns number of edges, -3 when fatal.
struct oc_edge { int32_t m1, m2, id1, id2; double rot[9]; double tra[3]; double cov[6]; };
int64_t orc_hba_add_edge(const float *xyz, const int64_t *cloud_off, const int32_t *ids, const int32_t *mps, oc_pose *poses, int W,
                         double voxel_size, double min_eigen_value, const double *plane_thre4, double gba_voxel_size,
                         double gba_min_eigen_value, const double *gba_eig4, int max_layer, int max_iter, int thread_num,
                         oc_edge *edges, double *hess_out, double *resis_first, double *resis_last, int32_t *n_voxels, int *n_outer,
                         int want_cloud, PointXYZIN *merged, int64_t *merged_n) {
  Params prm;
  prm.voxel_size = voxel_size; prm.min_eigen_value = min_eigen_value; prm.max_layer = max_layer;
  prm.gba_voxel_size = gba_voxel_size; prm.gba_min_eigen_value = gba_min_eigen_value;
  for (int k = 0; k < 4; k++) { prm.plane_eigen_value_thre[k] = plane_thre4[k]; prm.gba_eigen_value_array[k] = gba_eig4[k]; }
  std::vector<IMUST> xs = poses_in(poses, W);
  std::vector<KeyframeCloud> smps(W);
  std::vector<int> maps;
  for (int i = 0; i < W; i++) {
    smps[i].xyz = xyz + 3 * cloud_off[i]; smps[i].n = (int)(cloud_off[i + 1] - cloud_off[i]);
    smps[i].id = ids ? ids[i] : i; smps[i].mp = mps ? mps[i] : 0;
    bool has = false; for (int m : maps) if (m == smps[i].mp) has = true;
    if (!has) maps.push_back(smps[i].mp);
  }
  HbaResult out;
  HBA_add_edge(xs, smps, maps, max_iter, thread_num, want_cloud != 0, prm, out, false);
  if (n_outer) *n_outer = (int)out.n_voxels.size();
  for (size_t k = 0; k < out.n_voxels.size(); k++) if (n_voxels) n_voxels[k] = out.n_voxels[k];
  for (size_t k = 0; k < out.resis_first.size(); k++) { if (resis_first) resis_first[k] = out.resis_first[k]; if (resis_last) resis_last[k] = out.resis_last[k]; }
  if (out.fatal) return -3;
  poses_out(out.xs, poses);
  if (hess_out) std::memcpy(hess_out, out.hess.d.data(), sizeof(double) * out.hess.d.size());
  for (size_t k = 0; k < out.edges.size(); k++) {
    oc_edge &e = edges[k]; const PgoEdge &s = out.edges[k];
    e.m1 = s.m1; e.m2 = s.m2; e.id1 = s.id1; e.id2 = s.id2;
    for (int q = 0; q < 9; q++) e.rot[q] = s.rot[q];
    for (int q = 0; q < 3; q++) e.tra[q] = s.tra[q];
    for (int q = 0; q < 6; q++) e.cov[q] = s.cov[q];
  }
  if (want_cloud && merged) { for (size_t k = 0; k < out.merged.size(); k++) merged[k] = out.merged[k]; if (merged_n) *merged_n = (int64_t)out.merged.size(); }
  return (int64_t)out.edges.size();
}

// ---- local sliding-window map session (voxelslam.cpp:1888-2043 without the LIO/BA steps: poses are given) ----
struct LocalSession {
  MapCtx ctx;
  SurfMap surf_map, surf_map_slide;
  std::vector<std::vector<SlideWindow *>> sws;
  std::vector<IMUST> x_buf;
  std::vector<PVecPtr> pvec_buf;
  LidarFactor voxhess{1};
  int win_size = 10, win_count = 0, thread_num = 5;
  double dept_err = 0.02, beam_err = 0.05, jour = 0;
  std::vector<OctoTree *> flat;   // last export order
};
void *orc_local_create(int win_size, double voxel_size, double min_eigen_value, const double *thre4, int max_layer, int thread_num,
                       double dept_err, double beam_err) {
  LocalSession *s = new LocalSession();
  s->win_size = win_size; s->thread_num = thread_num; s->dept_err = dept_err; s->beam_err = beam_err;
  s->ctx.prm.voxel_size = voxel_size; s->ctx.prm.min_eigen_value = min_eigen_value; s->ctx.prm.max_layer = max_layer;
  for (int k = 0; k < 4; k++) s->ctx.prm.plane_eigen_value_thre[k] = thre4[k];
  s->ctx.mp.resize(win_size);
  for (int i = 0; i < win_size; i++) s->ctx.mp[i] = i;    // voxelslam.cpp:3158-3160
  s->sws.resize(thread_num);
  s->voxhess.win_size = win_size;
  return s;
}
static PVec pvec_in(const double *pv, int n) {
  PVec v(n);
  for (int i = 0; i < n; i++) { for (int k = 0; k < 3; k++) v[i].pnt[k] = pv[12 * i + k]; for (int k = 0; k < 9; k++) v[i].var[k] = pv[12 * i + 3 + k]; }
  return v;
}
static IMUST state_in(const oc_pose *pose, const double *cov225) {
  IMUST x;
  for (int k = 0; k < 9; k++) x.R[k] = pose->R[k];
  for (int k = 0; k < 3; k++) x.p[k] = pose->p[k];
  if (cov225) for (int k = 0; k < 225; k++) x.cov[k] = cov225[k];
  return x;
}
// points: float xyz in the body frame (extrinsic = identity).  Returns the pointVar records (pnt body frame,
// var AFTER pvec_update = world-frame covariance, voxelslam.hpp:242-265) when pv_out != null, the world points when pw_out != null.
// voxelslam.cpp:1905-1917: window bookkeeping + voxhess.clear() + cut_voxel_multi
void orc_local_insert(void *h, const float *xyz, int n, const oc_pose *pose, const double *cov225, double *pv_out /* n x 12 */, double *pw_out /* n x 3 */) {
  LocalSession &S = *(LocalSession *)h;
  std::vector<PointXYZIN> pl(n);
  for (int i = 0; i < n; i++) { std::memset(&pl[i], 0, sizeof(PointXYZIN)); pl[i].x = xyz[3 * i]; pl[i].y = xyz[3 * i + 1]; pl[i].z = xyz[3 * i + 2]; }
  IMUST ext;   // identity extrinsic
  PVecPtr pptr(new PVec());
  var_init(ext, pl, *pptr, S.dept_err, S.beam_err);
  IMUST x_curr = state_in(pose, cov225);
  std::vector<V3> pwld;
  pvec_update(*pptr, x_curr, pwld);
  for (int i = 0; i < n; i++) {
    if (pv_out) { for (int k = 0; k < 3; k++) pv_out[12 * i + k] = (*pptr)[i].pnt[k]; for (int k = 0; k < 9; k++) pv_out[12 * i + 3 + k] = (*pptr)[i].var[k]; }
    if (pw_out) for (int k = 0; k < 3; k++) pw_out[3 * i + k] = pwld[i][k];
  }
  S.win_count++;
  S.x_buf.push_back(x_curr);
  S.pvec_buf.push_back(pptr);
  S.voxhess.clear();
  S.voxhess.win_size = S.win_size;
  cut_voxel_multi(S.surf_map, *S.pvec_buf[S.win_count - 1], S.win_count - 1, S.surf_map_slide, S.win_size, pwld, S.sws, &S.ctx);
}
void orc_local_recut(void *h) {   // voxelslam.cpp:1919
  LocalSession &S = *(LocalSession *)h;
  multi_recut(S.surf_map_slide, S.win_count, S.x_buf, S.voxhess, S.sws, S.thread_num);
}
// voxelslam.cpp:1966-2040 without the optimizer: multi_margi, mp rotation, buffer shift.  Returns 1 when the window was full,
// -1 where the reference prints "Error: opt_state" and exits.
int orc_local_margi(void *h) {
  LocalSession &S = *(LocalSession *)h;
  if (S.win_count < S.win_size) return 0;
  const int mgsize = 1;
  bool ok = multi_margi(S.surf_map_slide, S.jour, S.win_count, S.x_buf, S.voxhess, S.sws[0], S.thread_num);
  for (int i = 0; i < S.win_size; i++) { S.ctx.mp[i] += mgsize; if (S.ctx.mp[i] >= S.win_size) S.ctx.mp[i] -= S.win_size; }
  for (int i = mgsize; i < S.win_count; i++) { S.x_buf[i - mgsize] = S.x_buf[i]; std::swap(S.pvec_buf[i - mgsize], S.pvec_buf[i]); }
  for (int i = S.win_count - mgsize; i < S.win_count; i++) { S.x_buf.pop_back(); S.pvec_buf.pop_back(); }
  S.win_count -= mgsize;
  return ok ? 1 : -1;
}
void orc_local_push(void *h, const float *xyz, int n, const oc_pose *pose, const double *cov225, double *pv_out /* n x 12 */) {
  orc_local_insert(h, xyz, n, pose, cov225, pv_out, nullptr);
  orc_local_recut(h);
  orc_local_margi(h);
}
int orc_local_win_count(void *h) { return ((LocalSession *)h)->win_count; }
void orc_local_get_poses(void *h, oc_pose *poses) { LocalSession &S = *(LocalSession *)h; poses_out(S.x_buf, poses); }
void orc_local_set_poses(void *h, const oc_pose *poses) {   // the optimizer's write-back into x_buf (R, p only)
  LocalSession &S = *(LocalSession *)h;
  for (int i = 0; i < S.win_count; i++) { for (int k = 0; k < 9; k++) S.x_buf[i].R[k] = poses[i].R[k]; for (int k = 0; k < 3; k++) S.x_buf[i].p[k] = poses[i].p[k]; }
}
void *orc_local_factor(void *h) { return &((LocalSession *)h)->voxhess; }   // borrowed: do not orc_factor_free
void orc_local_set_max_points(void *h, int max_points) { ((LocalSession *)h)->ctx.prm.max_points = max_points; }
void orc_local_set_jour(void *h, double jour) { ((LocalSession *)h)->jour = jour; }
// cut_voxel(feat_map, pvec, wdsize, jour)  voxel_map.hpp:2108-2152; pv rows: world pnt 3 + var 9
void orc_local_insert_fix(void *h, const double *pv, int n, double jour) {
  LocalSession &S = *(LocalSession *)h;
  PVec pvec = pvec_in(pv, n);
  cut_voxel_fix(S.surf_map, pvec, S.win_size, jour, &S.ctx);
}
int64_t orc_local_factor_voxels(void *h) { return (int64_t)((LocalSession *)h)->voxhess.plvec_voxels.size(); }
// flattened export of surf_map: node record = what OctoTree::match / inside read (voxel_map.hpp:1649-1721, :1836-1848)
struct oc_map_node {
  double center[3];
  float quater_length; int32_t octo_state;
  int32_t child[8];
  int32_t is_plane; float radius;
  double p_center[3], p_normal[3];
  double plane_var[36];
};
static void flatten(OctoTree *t, std::vector<OctoTree *> &out, std::vector<int> &idx_of_parent_slot) {
  (void)idx_of_parent_slot;
  out.push_back(t);
  if (t->octo_state == 1) for (int i = 0; i < 8; i++) if (t->leaves[i]) flatten(t->leaves[i], out, idx_of_parent_slot);
}
int64_t orc_local_export_count(void *h, int64_t *n_roots) {
  LocalSession &S = *(LocalSession *)h;
  S.flat.clear();
  std::vector<int> dummy;
  for (auto &kv : S.surf_map) flatten(kv.second, S.flat, dummy);
  *n_roots = (int64_t)S.surf_map.size();
  return (int64_t)S.flat.size();
}
void orc_local_export(void *h, int64_t *root_keys /* n_roots x 3 */, int32_t *root_node, oc_map_node *nodes) {
  LocalSession &S = *(LocalSession *)h;
  std::unordered_map<OctoTree *, int> id;
  for (size_t i = 0; i < S.flat.size(); i++) id[S.flat[i]] = (int)i;
  int64_t r = 0;
  for (auto &kv : S.surf_map) {
    root_keys[3 * r] = kv.first.x; root_keys[3 * r + 1] = kv.first.y; root_keys[3 * r + 2] = kv.first.z;
    root_node[r] = id[kv.second];
    r++;
  }
  for (size_t i = 0; i < S.flat.size(); i++) {
    OctoTree *t = S.flat[i];
    oc_map_node &nd = nodes[i];
    std::memset(&nd, 0, sizeof(nd));
    for (int k = 0; k < 3; k++) nd.center[k] = t->voxel_center[k];
    nd.quater_length = t->quater_length; nd.octo_state = t->octo_state;
    for (int k = 0; k < 8; k++) nd.child[k] = (t->octo_state == 1 && t->leaves[k]) ? id[t->leaves[k]] : -1;
    nd.is_plane = t->plane.is_plane ? 1 : 0; nd.radius = t->plane.radius;
    for (int k = 0; k < 3; k++) { nd.p_center[k] = t->plane.center[k]; nd.p_normal[k] = t->plane.normal[k]; }
    for (int k = 0; k < 36; k++) nd.plane_var[k] = t->plane.plane_var[k];
  }
}
// extended per-node record in the order of the last orc_local_export_count (layout = vx_map_node_state of the product header)
struct oc_node_state {
  int64_t root_key[3];
  int32_t layer, path, octo_state, isexist, has_sw, is_plane, last_num, opt_state, in_slide, n_window_pts, n_fix_pts, _pad;
  oc_cluster pcr_add, pcr_fix;
  double eig_value[3];
  double cov_add_trace;
};
static void state_walk(LocalSession &S, OctoTree *t, const VOXEL_LOC &key, int path, bool in_slide, std::unordered_map<OctoTree *, oc_node_state> &out) {
  oc_node_state st;
  std::memset(&st, 0, sizeof(st));
  st.root_key[0] = key.x; st.root_key[1] = key.y; st.root_key[2] = key.z;
  st.layer = t->layer; st.path = path; st.octo_state = t->octo_state; st.isexist = t->isexist ? 1 : 0; st.has_sw = t->sw != nullptr ? 1 : 0;
  st.is_plane = t->plane.is_plane ? 1 : 0; st.last_num = t->last_num; st.opt_state = t->opt_state; st.in_slide = in_slide ? 1 : 0;
  if (t->sw) for (auto &v : t->sw->points) st.n_window_pts += (int)v.size();
  st.n_fix_pts = (int)t->point_fix.size();
  to_c(t->pcr_add, st.pcr_add); to_c(t->pcr_fix, st.pcr_fix);
  for (int k = 0; k < 3; k++) st.eig_value[k] = t->eig_value[k];
  for (int k = 0; k < 9; k++) st.cov_add_trace += t->cov_add(k, k);
  out[t] = st;
  if (t->octo_state == 1) for (int i = 0; i < 8; i++) if (t->leaves[i]) state_walk(S, t->leaves[i], key, path | (i << (3 * t->layer)), in_slide, out);
}
void orc_local_export_states(void *h, oc_node_state *states) {
  LocalSession &S = *(LocalSession *)h;
  std::unordered_map<OctoTree *, oc_node_state> all;
  for (auto &kv : S.surf_map) state_walk(S, kv.second, kv.first, 0, S.surf_map_slide.find(kv.first) != S.surf_map_slide.end(), all);
  for (size_t i = 0; i < S.flat.size(); i++) states[i] = all[S.flat[i]];
}
void orc_local_free(void *h) { delete (LocalSession *)h; }

// ---- D1 / D2 over the session's surf_map ---------------------------------------------------------------
// pv: n x 12 (pnt 3 + var 9 col-major, var = body-frame covariance as var_init leaves it).
void orc_var_init(const float *xyz, int n, double dept_err, double beam_err, double *pv_out) {
  std::vector<PointXYZIN> pl(n);
  for (int i = 0; i < n; i++) { std::memset(&pl[i], 0, sizeof(PointXYZIN)); pl[i].x = xyz[3 * i]; pl[i].y = xyz[3 * i + 1]; pl[i].z = xyz[3 * i + 2]; }
  IMUST ext; PVec pvec;
  var_init(ext, pl, pvec, dept_err, beam_err);
  for (int i = 0; i < n; i++) { for (int k = 0; k < 3; k++) pv_out[12 * i + k] = pvec[i].pnt[k]; for (int k = 0; k < 9; k++) pv_out[12 * i + 3 + k] = pvec[i].var[k]; }
}
void orc_var_init_ext(const float *xyz, int n, const oc_pose *ext_pose, double dept_err, double beam_err, double *pv_out) {
  std::vector<PointXYZIN> pl(n);
  for (int i = 0; i < n; i++) { std::memset(&pl[i], 0, sizeof(PointXYZIN)); pl[i].x = xyz[3 * i]; pl[i].y = xyz[3 * i + 1]; pl[i].z = xyz[3 * i + 2]; }
  IMUST ext = poses_in(ext_pose, 1)[0]; PVec pvec;
  var_init(ext, pl, pvec, dept_err, beam_err);
  for (int i = 0; i < n; i++) { for (int k = 0; k < 3; k++) pv_out[12 * i + k] = pvec[i].pnt[k]; for (int k = 0; k < 9; k++) pv_out[12 * i + 3 + k] = pvec[i].var[k]; }
}
void orc_pvec_update(const double *pv, int n, const oc_pose *pose, const double *cov225, double *pv_out, double *pwld) {
  PVec pvec = pvec_in(pv, n);
  IMUST x = state_in(pose, cov225);
  std::vector<V3> w;
  pvec_update(pvec, x, w);
  for (int i = 0; i < n; i++) {
    for (int k = 0; k < 3; k++) pv_out[12 * i + k] = pvec[i].pnt[k];
    for (int k = 0; k < 9; k++) pv_out[12 * i + 3 + k] = pvec[i].var[k];
    for (int k = 0; k < 3; k++) pwld[3 * i + k] = w[i][k];
  }
}
// one pass (voxelslam.cpp:1005-1050); leaf cache in/out as indices into the last export (-1 = none)
void orc_lio_pass(void *h, const double *pv, int n, const oc_pose *pose, const double *cov225, int32_t *leaf_cache, double *HTH36, double *HTz6,
                  double *nnt9, int *match_num, int32_t *flags) {
  LocalSession &S = *(LocalSession *)h;
  PVec pvec = pvec_in(pv, n);
  IMUST x = state_in(pose, cov225);
  std::unordered_map<OctoTree *, int> id;
  for (size_t i = 0; i < S.flat.size(); i++) id[S.flat[i]] = (int)i;
  std::vector<OctoTree *> octos(n, nullptr);
  for (int i = 0; i < n; i++) if (leaf_cache && leaf_cache[i] >= 0) octos[i] = S.flat[leaf_cache[i]];
  LioPassOut o; std::vector<int> fl;
  lio_pass(S.surf_map, pvec, x, octos, &S.ctx, o, &fl);
  for (int k = 0; k < 36; k++) HTH36[k] = o.HTH[k];
  for (int k = 0; k < 6; k++) HTz6[k] = o.HTz[k];
  for (int k = 0; k < 9; k++) nnt9[k] = o.nnt[k];
  *match_num = o.match_num;
  for (int i = 0; i < n; i++) { if (flags) flags[i] = fl[i]; if (leaf_cache) leaf_cache[i] = octos[i] ? id[octos[i]] : -1; }
}
// full lio_state_estimation (voxelslam.cpp:962-1098); state15 in/out = [R(9) p(3) v(3) bg(3) ba(3)], cov in/out
int orc_lio_state_estimation(void *h, const double *pv, int n, double *state21, double *cov225, int *iters, int *match_nums4) {
  LocalSession &S = *(LocalSession *)h;
  PVec pvec = pvec_in(pv, n);
  IMUST x;
  for (int k = 0; k < 9; k++) x.R[k] = state21[k];
  for (int k = 0; k < 3; k++) { x.p[k] = state21[9 + k]; x.v[k] = state21[12 + k]; x.bg[k] = state21[15 + k]; x.ba[k] = state21[18 + k]; }
  for (int k = 0; k < 225; k++) x.cov[k] = cov225[k];
  LioTrace tr;
  bool ok = lio_state_estimation(S.surf_map, pvec, x, &S.ctx, &tr);
  for (int k = 0; k < 9; k++) state21[k] = x.R[k];
  for (int k = 0; k < 3; k++) { state21[9 + k] = x.p[k]; state21[12 + k] = x.v[k]; state21[15 + k] = x.bg[k]; state21[18 + k] = x.ba[k]; }
  for (int k = 0; k < 225; k++) cov225[k] = x.cov[k];
  if (iters) *iters = tr.iters;
  if (match_nums4) for (int k = 0; k < tr.iters && k < 4; k++) match_nums4[k] = tr.passes[k].match_num;
  return ok ? 1 : 0;
}

// ---- B6 / B4 / B5 ---------------------------------------------------------------------------
struct oc_imu_pre {
  double R_delta[9], p_delta[3], v_delta[3], bg[3], ba[3];
  double R_bg[9], p_bg[9], p_ba[9], v_bg[9], v_ba[9];
  double dtime;
  double dbg[3], dba[3], dbg_buf[3], dba_buf[3];
  double cov[225];
};
struct oc_imu_noise { double scale_gravity; double noise_meas[36]; double noise_walk[36]; };
struct oc_imu_sample { double t, gyr[3], acc[3]; };
struct oc_imu_state { double R[9], p[3], v[3], bg[3], ba[3], g[3]; };

static ImuNoise noise_in(const oc_imu_noise *n) {
  ImuNoise z; z.scale_gravity = n->scale_gravity;
  for (int k = 0; k < 36; k++) { z.noiseMeas[k] = n->noise_meas[k]; z.noiseWalk[k] = n->noise_walk[k]; }
  return z;
}
static void pre_in(const oc_imu_pre *p, IMU_PRE &q) {
  cp_in(q.R_delta, p->R_delta, 9); cp_in(q.p_delta, p->p_delta, 3); cp_in(q.v_delta, p->v_delta, 3); cp_in(q.bg, p->bg, 3); cp_in(q.ba, p->ba, 3);
  cp_in(q.R_bg, p->R_bg, 9); cp_in(q.p_bg, p->p_bg, 9); cp_in(q.p_ba, p->p_ba, 9); cp_in(q.v_bg, p->v_bg, 9); cp_in(q.v_ba, p->v_ba, 9);
  q.dtime = p->dtime;
  cp_in(q.dbg, p->dbg, 3); cp_in(q.dba, p->dba, 3); cp_in(q.dbg_buf, p->dbg_buf, 3); cp_in(q.dba_buf, p->dba_buf, 3);
  cp_in(q.cov, p->cov, 225);
}
static void pre_out(const IMU_PRE &q, oc_imu_pre *p) {
  cp_out(q.R_delta, p->R_delta, 9); cp_out(q.p_delta, p->p_delta, 3); cp_out(q.v_delta, p->v_delta, 3); cp_out(q.bg, p->bg, 3); cp_out(q.ba, p->ba, 3);
  cp_out(q.R_bg, p->R_bg, 9); cp_out(q.p_bg, p->p_bg, 9); cp_out(q.p_ba, p->p_ba, 9); cp_out(q.v_bg, p->v_bg, 9); cp_out(q.v_ba, p->v_ba, 9);
  p->dtime = q.dtime;
  cp_out(q.dbg, p->dbg, 3); cp_out(q.dba, p->dba, 3); cp_out(q.dbg_buf, p->dbg_buf, 3); cp_out(q.dba_buf, p->dba_buf, 3);
  cp_out(q.cov, p->cov, 225);
}
static IMUST imu_state_in(const oc_imu_state &s) {
  IMUST x;
  cp_in(x.R, s.R, 9); cp_in(x.p, s.p, 3); cp_in(x.v, s.v, 3); cp_in(x.bg, s.bg, 3); cp_in(x.ba, s.ba, 3); cp_in(x.g, s.g, 3);
  return x;
}
static void state_out(const IMUST &x, oc_imu_state &s) {
  cp_out(x.R, s.R, 9); cp_out(x.p, s.p, 3); cp_out(x.v, s.v, 3); cp_out(x.bg, s.bg, 3); cp_out(x.ba, s.ba, 3); cp_out(x.g, s.g, 3);
}

void orc_imu_pre_init(oc_imu_pre *pre, const double *bg, const double *ba) {
  ImuNoise nz;
  IMU_PRE q(&nz, vec3(bg[0], bg[1], bg[2]), vec3(ba[0], ba[1], ba[2]));
  pre_out(q, pre);
}
void orc_imu_pre_push(oc_imu_pre *pre, const oc_imu_noise *noise, const oc_imu_sample *s, int n) {
  ImuNoise nz = noise_in(noise);
  IMU_PRE q(&nz);
  pre_in(pre, q);
  std::vector<ImuSample> v(n);
  for (int k = 0; k < n; k++) { v[k].t = s[k].t; cp_in(v[k].gyr, s[k].gyr, 3); cp_in(v[k].acc, s[k].acc, 3); }
  q.push_imu(v);
  pre_out(q, pre);
}
void orc_imu_pre_evaluate(const oc_imu_pre *pre, const oc_imu_state *st1, const oc_imu_state *st2, int with_gravity, double *jtj, double *gg,
                          double *cost) {
  ImuNoise nz;
  IMU_PRE q(&nz);
  pre_in(pre, q);
  IMUST a = imu_state_in(*st1), b = imu_state_in(*st2);
  const bool jac = jtj && gg;
  if (with_gravity) {
    Mat<33, 33> h; Mat<33, 1> g;
    *cost = q.evaluate<33>(a, b, h, g, jac);
    if (jac) { cp_out(h, jtj, 33 * 33); cp_out(g, gg, 33); }
  } else {
    Mat<30, 30> h; Mat<30, 1> g;
    *cost = q.evaluate<30>(a, b, h, g, jac);
    if (jac) { cp_out(h, jtj, 30 * 30); cp_out(g, gg, 30); }
  }
}
// LI_BA_Optimizer (gravity = 0, max_iter ignored: 3) / LI_BA_OptimizerGravity (gravity = 1)
void orc_ba_li(void *h, oc_imu_state *xs, oc_imu_pre *pres, double imu_coef, int gravity, int max_iter, int use_threads, double *resis2,
               double *hess_out, int *accept_trace, int *n_trace) {
  LidarFactor &f = *(LidarFactor *)h;
  const int W = f.win_size;
  ImuNoise nz;
  std::vector<IMUST> x(W);
  for (int i = 0; i < W; i++) x[i] = imu_state_in(xs[i]);
  std::deque<IMU_PRE *> imus;
  for (int i = 0; i + 1 < W; i++) { IMU_PRE *q = new IMU_PRE(&nz); pre_in(&pres[i], *q); imus.push_back(q); }
  MatX hess; std::vector<double> resis;
  std::vector<int> trace;
  if (gravity) {
    LI_BA_OptimizerGravity opt; opt.imu_coef = imu_coef; opt.use_threads = use_threads != 0;
    opt.damping_iter(x, f, imus, resis, &hess, max_iter);
    trace = opt.accept_trace;
  } else {
    LI_BA_Optimizer opt; opt.imu_coef = imu_coef; opt.use_threads = use_threads != 0;
    opt.damping_iter(x, f, imus, &hess);
    trace = opt.accept_trace;
  }
  for (int i = 0; i < W; i++) state_out(x[i], xs[i]);
  for (int i = 0; i + 1 < W; i++) { pre_out(*imus[i], &pres[i]); delete imus[i]; }
  if (hess_out) std::memcpy(hess_out, hess.d.data(), sizeof(double) * hess.d.size());
  if (resis2 && resis.size() >= 2) { resis2[0] = resis[0]; resis2[1] = resis[1]; }
  if (n_trace) *n_trace = (int)trace.size();
  if (accept_trace) for (size_t i = 0; i < trace.size(); i++) accept_trace[i] = trace[i];
}

// icp_normal (loop_refine.hpp:47-145).  pose: R (column-major 9) + p (3), in/out.  match_nums / resis: up to 20 entries.
int orc_icp_normal(const float *src_xyz, const float *src_nrm, int ns, const float *tar_xyz, const float *tar_nrm, int nt, oc_pose *pose,
                   double icp_eigval, int *iters, int *match_nums, double *resis, double *eig3, int *is_converge) {
  M3 R; V3 p;
  for (int k = 0; k < 9; k++) R[k] = pose->R[k];
  for (int k = 0; k < 3; k++) p[k] = pose->p[k];
  IcpTrace tr;
  const bool ok = icp_normal(src_xyz, src_nrm, ns, tar_xyz, tar_nrm, nt, p, R, icp_eigval, &tr);
  for (int k = 0; k < 9; k++) pose->R[k] = R[k];
  for (int k = 0; k < 3; k++) pose->p[k] = p[k];
  if (iters) *iters = tr.iters;
  for (int k = 0; k < tr.iters; k++) { if (match_nums) match_nums[k] = tr.match_num[k]; if (resis) resis[k] = tr.resi[k]; }
  if (eig3) for (int k = 0; k < 3; k++) eig3[k] = tr.eig[k];
  if (is_converge) *is_converge = tr.is_converge;
  return ok ? 1 : 0;
}

// ---- stored sessions ------------------------------------------------------------------------
struct oc_scan_state { double t, p[3], R[9], v[3], bg[3], ba[3], g[3], v6[6]; };
static_assert(sizeof(oc_scan_state) == sizeof(ScanState), "scan state layout");

int orc_write_lidarstate(const char *path, const oc_scan_state *xs, int64_t n) {
  std::vector<ScanState> v(n);
  if (n) std::memcpy(v.data(), xs, sizeof(ScanState) * n);
  return write_lidarstate(path, v) ? 0 : -1;
}
int orc_read_lidarstate(const char *path, oc_scan_state *out, int64_t cap, int64_t *n) {
  std::vector<ScanState> v;
  if (!read_lidarstate(path, v)) return -1;
  *n = (int64_t)v.size();
  if ((int64_t)v.size() > cap) return -2;
  if (!v.empty()) std::memcpy(out, v.data(), sizeof(ScanState) * v.size());
  return 0;
}
int orc_write_pcd_xyzi(const char *path, const float *xyzi, int64_t n) { return write_pcd_xyzi(path, xyzi, n) ? 0 : -1; }
int orc_write_edges(const char *path, const char *const *n1, const char *const *n2, const int32_t *id1, const int32_t *id2, const double *tra3,
                    const double *quat4, int64_t n) {
  std::vector<EdgeLine> es(n);
  for (int64_t i = 0; i < n; i++) {
    es[i].n1 = n1[i]; es[i].n2 = n2[i]; es[i].id1 = id1[i]; es[i].id2 = id2[i];
    for (int k = 0; k < 3; k++) es[i].tra[k] = tra3[3 * i + k];
    for (int k = 0; k < 4; k++) es[i].q[k] = quat4[4 * i + k];
  }
  return write_edges(path, es) ? 0 : -1;
}
void orc_quat_to_R(const double *q_xyzw, double *R9) { quat_to_R(q_xyzw[3], q_xyzw[0], q_xyzw[1], q_xyzw[2], R9); }
void orc_R_to_quat(const double *R9, double *q_xyzw) { R_to_quat(R9, q_xyzw); }
// voxelslam.cpp:345-378; out: capacity = total number of points
void orc_keyframe_assemble(const float *xyz, const int64_t *cloud_off, int win, const oc_pose *poses, double voxel_size, PointXYZIN *out, int64_t *n_out) {
  std::vector<IMUST> xs = poses_in(poses, win);
  std::vector<const float *> ptr(win); std::vector<int64_t> n(win);
  for (int j = 0; j < win; j++) { ptr[j] = xyz + 3 * cloud_off[j]; n[j] = cloud_off[j + 1] - cloud_off[j]; }
  std::vector<PointXYZIN> pl = keyframe_assemble(ptr, n, xs, voxel_size);
  for (size_t k = 0; k < pl.size(); k++) out[k] = pl[k];
  *n_out = (int64_t)pl.size();
}

// BTC plane voxels (BTC.cpp:279-334): out = 10 doubles per plane (center 3 | normal 3 | min_eigen_value | radius | d | points_size)
void orc_btc_plane_cloud(const float *xyz, int64_t n, double voxel_size, double plane_thre, int init_num, double *out, int64_t cap, int64_t *n_out) {
  std::vector<BtcPlane> pl = btc_plane_cloud(xyz, n, voxel_size, plane_thre, init_num);
  *n_out = (int64_t)pl.size();
  for (size_t k = 0; k < pl.size() && (int64_t)k < cap; k++) std::memcpy(out + 10 * k, &pl[k], sizeof(BtcPlane));
}

// IMUEKF::motion_blur, per-point half (ekf_imu.hpp:138-163).  pts: n x {x, y, z, curvature} floats (in place); poses: n_poses x 22 doubles
// {t, R[9] column-major, p[3], v[3], angvel[3], acc[3]}; xc / lidar_to_imu: {R[9] column-major, p[3]}.
void orc_deskew(float *pts, int64_t n, const double *poses, int n_poses, const oc_pose *xc, const oc_pose *lidar_to_imu, int32_t *assign) {
  std::vector<ImuPose> ip(n_poses);
  for (int i = 0; i < n_poses; i++) {
    const double *q = poses + 22 * (size_t)i;
    ip[i].t = q[0];
    for (int k = 0; k < 9; k++) ip[i].R[k] = q[1 + k];
    for (int k = 0; k < 3; k++) { ip[i].p[k] = q[10 + k]; ip[i].v[k] = q[13 + k]; ip[i].angvel[k] = q[16 + k]; ip[i].acc[k] = q[19 + k]; }
  }
  M3 xR, lR; V3 xp, lp;
  for (int k = 0; k < 9; k++) { xR[k] = xc->R[k]; lR[k] = lidar_to_imu->R[k]; }
  for (int k = 0; k < 3; k++) { xp[k] = xc->p[k]; lp[k] = lidar_to_imu->p[k]; }
  motion_blur_points(reinterpret_cast<SkewPoint *>(pts), (long)n, ip, xR, xp, lR, lp, assign);
}

// voxel-frames evaluated by acc_evaluate2 since the last reset (bench.py's CPU arm); reset != 0 clears after reading
int64_t orc_vf_counter(int reset) {
  const int64_t v = g_voxel_frames_evaluated.load();
  if (reset) g_voxel_frames_evaluated.store(0);
  return v;
}
double orc_now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // extern "C"
