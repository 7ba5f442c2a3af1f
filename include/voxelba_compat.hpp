// voxelba_compat.hpp — header-only C++14 adaptor that re-creates the reference's class surface for the BA path on top
// of the C ABI in voxelba.h, so call sites of VoxelSLAM/src/voxelslam.cpp compile unchanged:
//
//   class LidarFactor               voxel_map.hpp:124-339   push_voxel / acc_evaluate2 / evaluate_only_residual / clear,
//                                                           public member vectors kept (margi reads them, :1495-1501)
//   class Lidar_BA_Optimizer        voxel_map.hpp:342-498   divide_thread / only_residual / damping_iter
//   class LI_BA_Optimizer           voxel_map.hpp:504-714   damping_iter(x_stats, voxhess, imus_factor, hess*)
//   class LI_BA_OptimizerGravity    voxel_map.hpp:717-976   damping_iter(x_stats, voxhess, imus_factor, resis, hess*, max_iter)
//   down_sampling_voxel(cloud, voxel_size)   tools.hpp:201-238 (any cloud with a contiguous `points` vector of 48-byte points)
//
// How a maintainer uses it: keep tools.hpp (IMUST, PointCluster, Exp/Log ...) and preintegration.hpp (IMU_PRE), delete the
// four classes above from voxel_map.hpp and `#include "voxelba_compat.hpp"` after tools.hpp.  The adaptor is written against
// the member NAMES of the reference's own types (IMUST::R/p/v/bg/ba/g, PointCluster::P/v/N, IMU_PRE::R_delta ... cov) and
// against Eigen's interface subset {data(), rows(), cols(), resize(r,c), setZero()}; with Eigen in scope those are the
// reference's types, byte-compatible with the POD structs of voxelba.h (column-major doubles).
// Without Eigen (this repository's own tests) define VOXELBA_COMPAT_STANDALONE: minimal stand-ins of exactly that subset are
// provided in namespace Eigen together with IMUST / PointCluster / IMU_PRE mirrors.
//
// Error behaviour follows the reference: where it prints and exit(0)s ("Too Less Voxel", voxel_map.hpp:399-403) so does
// the adaptor; any other library error is fatal too (the reference has no error channel on this path).
#pragma once
#include "voxelba.h"

#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <deque>
#include <vector>

#ifdef VOXELBA_COMPAT_STANDALONE
// ---------------------------------------------------------------------------------------------------------------------
// stand-ins (tests only): the interface subset of Eigen the adaptor touches + mirrors of the reference's POD-like types
// ---------------------------------------------------------------------------------------------------------------------
namespace Eigen {
template <int R, int C> struct FixedMat {
  double d[R * C];
  FixedMat() { std::memset(d, 0, sizeof(d)); }
  double *data() { return d; }
  const double *data() const { return d; }
  double &operator()(int r, int c) { return d[c * R + r]; }
  double operator()(int r, int c) const { return d[c * R + r]; }
  double &operator[](int i) { return d[i]; }
  double operator[](int i) const { return d[i]; }
  void setZero() { std::memset(d, 0, sizeof(d)); }
  void setIdentity() { setZero(); for (int i = 0; i < (R < C ? R : C); i++) d[i * R + i] = 1.0; }
};
using Matrix3d = FixedMat<3, 3>;
using Vector3d = FixedMat<3, 1>;
struct MatrixXd {
  std::vector<double> d; int r = 0, c = 0;
  void resize(int rr, int cc) { r = rr; c = cc; d.assign((size_t)rr * cc, 0.0); }
  void setZero() { std::fill(d.begin(), d.end(), 0.0); }
  int rows() const { return r; }
  int cols() const { return c; }
  double *data() { return d.data(); }
  const double *data() const { return d.data(); }
  double &operator()(int i, int j) { return d[(size_t)j * r + i]; }
  double operator()(int i, int j) const { return d[(size_t)j * r + i]; }
};
struct VectorXd {
  std::vector<double> d;
  void resize(int n) { d.assign(n, 0.0); }
  void setZero() { std::fill(d.begin(), d.end(), 0.0); }
  int size() const { return (int)d.size(); }
  int rows() const { return (int)d.size(); }
  double *data() { return d.data(); }
  const double *data() const { return d.data(); }
  double &operator[](int i) { return d[i]; }
  double operator[](int i) const { return d[i]; }
};
}  // namespace Eigen

struct IMUST {                       // tools.hpp:135-199 (members only)
  double t = 0;
  Eigen::Matrix3d R; Eigen::Vector3d p, v, bg, ba, g;
  Eigen::FixedMat<15, 15> cov;
  IMUST() { R.setIdentity(); }
};
struct PointCluster {                // tools.hpp:304-365 (members only)
  Eigen::Matrix3d P; Eigen::Vector3d v; int N = 0;
  void clear() { P.setZero(); v.setZero(); N = 0; }
};
struct IMU_PRE {                     // preintegration.hpp:11-48 (members only)
  Eigen::Matrix3d R_delta; Eigen::Vector3d p_delta, v_delta, bg, ba;
  Eigen::Matrix3d R_bg, p_bg, p_ba, v_bg, v_ba;
  double dtime = 0;
  Eigen::Vector3d dbg, dba, dbg_buf, dba_buf;
  Eigen::FixedMat<15, 15> cov;
  IMU_PRE() { R_delta.setIdentity(); }
};
#ifndef PLV
#define PLV(a) std::vector<Eigen::FixedMat<a, 1>>
#define PLM(a) std::vector<Eigen::FixedMat<a, a>>
#endif
#endif  // VOXELBA_COMPAT_STANDALONE

namespace vxc {

inline void die(const char *where, int code) {
  std::printf("libvoxelba: %s failed (%d): %s\n", where, code, vx_last_error());
  std::exit(0);   // the reference's only error exit on this path is printf + exit(0) (voxel_map.hpp:399-403)
}
#define VXC_CHECK(call) do { int _c = (call); if (_c != VX_OK) ::vxc::die(#call, _c); } while (0)

// one context per calling thread (the reference's three threads work on different maps).  An object of this
// adaptor (LidarFactor, the map sentinels) belongs to the thread that first used it: its device handle was created on that
// thread's context, and every later call must come from the same thread — as in the reference, where a voxhess / surf_map is only
// ever touched by the thread that owns it.  The public member vectors of LidarFactor may be READ freely; change them only through
// push_voxel / clear (direct edits are not seen by the device copy).
inline vx_ctx *context() {
  static thread_local vx_ctx *ctx = nullptr;
  if (!ctx) VXC_CHECK(vx_ctx_create(0, &ctx));
  return ctx;
}

template <class PC> inline vx_cluster to_c(const PC &s) {
  vx_cluster c;
  std::memcpy(c.P, s.P.data(), sizeof(c.P)); std::memcpy(c.v, s.v.data(), sizeof(c.v));
  c.N = s.N; c._pad = 0;
  return c;
}
template <class PC> inline void from_c(const vx_cluster &c, PC &s) {
  std::memcpy(s.P.data(), c.P, sizeof(c.P)); std::memcpy(s.v.data(), c.v, sizeof(c.v)); s.N = c.N;
}
template <class ST> inline vx_pose pose_of(const ST &x) {
  vx_pose p;
  std::memcpy(p.R, x.R.data(), sizeof(p.R)); std::memcpy(p.p, x.p.data(), sizeof(p.p));
  return p;
}
template <class ST> inline vx_imu_state state_of(const ST &x) {
  vx_imu_state s;
  std::memcpy(s.R, x.R.data(), 72); std::memcpy(s.p, x.p.data(), 24); std::memcpy(s.v, x.v.data(), 24);
  std::memcpy(s.bg, x.bg.data(), 24); std::memcpy(s.ba, x.ba.data(), 24); std::memcpy(s.g, x.g.data(), 24);
  return s;
}
template <class ST> inline void state_to(const vx_imu_state &s, ST &x) {
  std::memcpy(x.R.data(), s.R, 72); std::memcpy(x.p.data(), s.p, 24); std::memcpy(x.v.data(), s.v, 24);
  std::memcpy(x.bg.data(), s.bg, 24); std::memcpy(x.ba.data(), s.ba, 24); std::memcpy(x.g.data(), s.g, 24);
}
template <class PRE> inline vx_imu_pre pre_of(const PRE &q) {
  vx_imu_pre p;
  std::memcpy(p.R_delta, q.R_delta.data(), 72); std::memcpy(p.p_delta, q.p_delta.data(), 24); std::memcpy(p.v_delta, q.v_delta.data(), 24);
  std::memcpy(p.bg, q.bg.data(), 24); std::memcpy(p.ba, q.ba.data(), 24);
  std::memcpy(p.R_bg, q.R_bg.data(), 72); std::memcpy(p.p_bg, q.p_bg.data(), 72); std::memcpy(p.p_ba, q.p_ba.data(), 72);
  std::memcpy(p.v_bg, q.v_bg.data(), 72); std::memcpy(p.v_ba, q.v_ba.data(), 72);
  p.dtime = q.dtime;
  std::memcpy(p.dbg, q.dbg.data(), 24); std::memcpy(p.dba, q.dba.data(), 24); std::memcpy(p.dbg_buf, q.dbg_buf.data(), 24); std::memcpy(p.dba_buf, q.dba_buf.data(), 24);
  std::memcpy(p.cov, q.cov.data(), sizeof(p.cov));
  return p;
}
template <class PRE> inline void pre_back(const vx_imu_pre &p, PRE &q) {   // damping_iter only moves the bias linearisation point
  std::memcpy(q.dbg.data(), p.dbg, 24); std::memcpy(q.dba.data(), p.dba, 24);
  std::memcpy(q.dbg_buf.data(), p.dbg_buf, 24); std::memcpy(q.dba_buf.data(), p.dba_buf, 24);
}

// pcl::PointCloud keeps width / height beside points; any cloud type without them is left alone
template <class Cloud> inline auto set_cloud_dims(Cloud &c, uint32_t n, int) -> decltype(c.width = n, void()) { c.width = n; c.height = 1; }
template <class Cloud> inline void set_cloud_dims(Cloud &, uint32_t, long) {}
#ifdef VOXELBA_COMPAT_STANDALONE
inline Eigen::FixedMat<6, 1> vec6() { return Eigen::FixedMat<6, 1>(); }
#else
inline Eigen::Matrix<double, 6, 1> vec6() { return Eigen::Matrix<double, 6, 1>::Zero(); }
#endif
}  // namespace vxc

// ---------------------------------------------------------------------------------------------------------------------
// class LidarFactor (voxel_map.hpp:124-339)
// ---------------------------------------------------------------------------------------------------------------------
class LidarFactor {
public:
  std::vector<PointCluster> sig_vecs;
  std::vector<std::vector<PointCluster>> plvec_voxels;
  std::vector<double> coeffs;
  PLV(3) eig_values;
  PLM(3) eig_vectors;
  std::vector<PointCluster> pcr_adds;
  int win_size;

  explicit LidarFactor(int _w) : win_size(_w) {}
  LidarFactor(const LidarFactor &) = delete;
  LidarFactor &operator=(const LidarFactor &) = delete;
  ~LidarFactor() { if (dev_) vx_factor_free(dev_); }

  // voxel_map.hpp:139-147
  void push_voxel(std::vector<PointCluster> &vec_orig, PointCluster &fix, double coe, Eigen::Vector3d &eig_value, Eigen::Matrix3d &eig_vector,
                  PointCluster &pcr_add) {
    on_device_ = false;
    plvec_voxels.push_back(vec_orig);
    sig_vecs.push_back(fix);
    coeffs.push_back(coe);
    eig_values.push_back(eig_value);
    eig_vectors.push_back(eig_vector);
    pcr_adds.push_back(pcr_add);
    dirty_ = true;
  }

  // voxel_map.hpp:150-282
  void acc_evaluate2(const std::vector<IMUST> &xs, int head, int end, Eigen::MatrixXd &Hess, Eigen::VectorXd &JacT, double &residual) {
    sync_device();
    const int jl = 6 * win_size;
    Hess.resize(jl, jl); JacT.resize(jl);           // the reference calls setZero() on caller-sized objects (:152-153)
    Hess.setZero(); JacT.setZero(); residual = 0;
    std::vector<vx_pose> ps = poses(xs);
    VXC_CHECK(vx_factor_hessian(vxc::context(), dev_, ps.data(), head, end, Hess.data(), JacT.data(), &residual));
  }

  // voxel_map.hpp:285-325 — refreshes eig_values / eig_vectors / pcr_adds of the voxels in [head, end) (:317-319)
  void evaluate_only_residual(const std::vector<IMUST> &xs, int head, int end, double &residual) {
    sync_device();
    residual = 0;
    std::vector<vx_pose> ps = poses(xs);
    VXC_CHECK(vx_factor_residual(vxc::context(), dev_, ps.data(), head, end, &residual));
    pull_cache(head, end);
  }

  // voxel_map.hpp:328-336
  void clear() {
    sig_vecs.clear(); plvec_voxels.clear(); eig_values.clear(); eig_vectors.clear(); pcr_adds.clear(); coeffs.clear();
    dirty_ = true; on_device_ = false;
  }

  // ---- adaptor plumbing ----
  vx_factor *device() { sync_device(); return dev_; }
  // the handle for a producer that fills the factor ON THE DEVICE (multi_recut's tras_opt, OctreeGBA_multi_recut): the public member
  // vectors stay empty (nothing on this path reads them: margi takes the cache from the device handle), pull_host() fills them on demand
  vx_factor *device_for_fill() {
    if (dev_ && dev_w_ != win_size) { vx_factor_free(dev_); dev_ = nullptr; }
    if (!dev_) { VXC_CHECK(vx_factor_create(vxc::context(), win_size, 1, &dev_)); dev_w_ = win_size; }
    sig_vecs.clear(); plvec_voxels.clear(); eig_values.clear(); eig_vectors.clear(); pcr_adds.clear(); coeffs.clear();
    dirty_ = false; on_device_ = true;
    return dev_;
  }
  int64_t num_voxels() const { return on_device_ && dev_ ? vx_factor_num_voxels(dev_) : (int64_t)plvec_voxels.size(); }
  // device -> public member vectors (plvec_voxels dense per window slot, coeffs, cache), for callers that read them
  void pull_host() {
    if (!on_device_ || !dev_) return;
    const int64_t V = vx_factor_num_voxels(dev_), nnz = vx_factor_num_voxel_frames(dev_);
    std::vector<int64_t> off(V + 1); std::vector<int32_t> fr(nnz); std::vector<vx_cluster> cl(nnz), mrg(V); std::vector<double> coe(V), ev(3 * V), evec(9 * V);
    VXC_CHECK(vx_factor_download_csr(dev_, off.data(), fr.data(), cl.data(), coe.data()));
    VXC_CHECK(vx_factor_download_cache(dev_, ev.data(), evec.data(), mrg.data()));
    plvec_voxels.assign(V, std::vector<PointCluster>(win_size)); sig_vecs.assign(V, PointCluster()); coeffs = coe;
    eig_values.resize(V); eig_vectors.resize(V); pcr_adds.resize(V);
    for (int64_t a = 0; a < V; a++) {
      for (int64_t e = off[a]; e < off[a + 1]; e++) vxc::from_c(cl[e], plvec_voxels[a][fr[e]]);
      std::memcpy(eig_values[a].data(), &ev[3 * a], 24); std::memcpy(eig_vectors[a].data(), &evec[9 * a], 72); vxc::from_c(mrg[a], pcr_adds[a]);
    }
  }
  // after an optimizer ran on the device handle: bring the cache the reference would have left in the member vectors
  void pull_cache(int head, int end) {
    const size_t V = plvec_voxels.size();
    if (V == 0 || on_device_) return;
    std::vector<double> ev(3 * V), evec(9 * V);
    std::vector<vx_cluster> mrg(V);
    VXC_CHECK(vx_factor_download_cache(dev_, ev.data(), evec.data(), mrg.data()));
    if (end < 0 || end > (int)V) end = (int)V;
    for (int a = head; a < end; a++) {
      std::memcpy(eig_values[a].data(), &ev[3 * a], 24);
      std::memcpy(eig_vectors[a].data(), &evec[9 * a], 72);
      vxc::from_c(mrg[a], pcr_adds[a]);
    }
  }

private:
  vx_factor *dev_ = nullptr;
  int dev_w_ = -1;
  bool dirty_ = true;
  bool on_device_ = false;     // filled by a device producer: the member vectors are not authoritative

  std::vector<vx_pose> poses(const std::vector<IMUST> &xs) const {
    std::vector<vx_pose> ps(win_size);
    for (int i = 0; i < win_size; i++) {
      if (i < (int)xs.size()) ps[i] = vxc::pose_of(xs[i]);
      else { std::memset(&ps[i], 0, sizeof(vx_pose)); ps[i].R[0] = ps[i].R[4] = ps[i].R[8] = 1.0; }
    }
    return ps;
  }
  void sync_device() {
    if (dev_ && dev_w_ != win_size) { vx_factor_free(dev_); dev_ = nullptr; }
    if (!dev_) { VXC_CHECK(vx_factor_create(vxc::context(), win_size, 1, &dev_)); dev_w_ = win_size; dirty_ = true; }
    if (!dirty_) return;
    const size_t V = plvec_voxels.size();
    std::vector<vx_cluster> pl(V * win_size), sig(V), add(V);
    std::vector<double> ev(3 * V), evec(9 * V);
    for (size_t a = 0; a < V; a++) {
      for (int i = 0; i < win_size; i++) pl[a * win_size + i] = vxc::to_c(plvec_voxels[a][i]);
      sig[a] = vxc::to_c(sig_vecs[a]); add[a] = vxc::to_c(pcr_adds[a]);
      std::memcpy(&ev[3 * a], eig_values[a].data(), 24); std::memcpy(&evec[9 * a], eig_vectors[a].data(), 72);
    }
    VXC_CHECK(vx_factor_upload_dense(dev_, (int64_t)V, pl.data(), sig.data(), coeffs.data(), ev.data(), evec.data(), add.data()));
    dirty_ = false;
  }
};

// ---------------------------------------------------------------------------------------------------------------------
// class Lidar_BA_Optimizer (voxel_map.hpp:342-498)
// ---------------------------------------------------------------------------------------------------------------------
class Lidar_BA_Optimizer {
public:
  int win_size, jac_leng, thd_num = 2;

  // voxel_map.hpp:347-389 (the worker threads are the device; the split into thd_num ranges only re-orders the sum)
  double divide_thread(std::vector<IMUST> &x_stats, LidarFactor &voxhess, Eigen::MatrixXd &Hess, Eigen::VectorXd &JacT) {
    double residual = 0;
    voxhess.acc_evaluate2(x_stats, 0, (int)voxhess.num_voxels(), Hess, JacT, residual);
    return residual;
  }
  // voxel_map.hpp:391-420, including the exit of :399-403
  double only_residual(std::vector<IMUST> &x_stats, LidarFactor &voxhess) {
    const int g_size = (int)voxhess.num_voxels();
    if (g_size < thd_num) { std::printf("Too Less Voxel"); std::exit(0); }
    double residual = 0;
    voxhess.evaluate_only_residual(x_stats, 0, g_size, residual);
    return residual;
  }
  // voxel_map.hpp:422-497
  bool damping_iter(std::vector<IMUST> &x_stats, LidarFactor &voxhess, Eigen::MatrixXd *hess, std::vector<double> &resis, int max_iter = 3,
                    bool is_display = false) {
    win_size = voxhess.win_size;
    jac_leng = win_size * 6;
    std::vector<vx_pose> ps(win_size);
    for (int i = 0; i < win_size; i++) ps[i] = vxc::pose_of(x_stats[i]);
    hess->resize(jac_leng, jac_leng);
    double r2[2] = {0, 0};
    int conv = 0;
    const int rc = vx_ba_lidar(vxc::context(), voxhess.device(), ps.data(), hess->data(), r2, max_iter, thd_num, &conv, nullptr, nullptr);
    if (rc == VX_ERR_TOO_FEW) { std::printf("Too Less Voxel"); std::exit(0); }
    if (rc != VX_OK) vxc::die("vx_ba_lidar", rc);
    for (int i = 0; i < win_size; i++) { std::memcpy(x_stats[i].R.data(), ps[i].R, 72); std::memcpy(x_stats[i].p.data(), ps[i].p, 24); }
    resis.push_back(r2[0]); resis.push_back(r2[1]);
    voxhess.pull_cache(0, -1);
    if (is_display) std::printf("LiDAR BA: %lf -> %lf\n", r2[0], r2[1]);
    return conv != 0;
  }
};

// ---------------------------------------------------------------------------------------------------------------------
// class LI_BA_Optimizer (voxel_map.hpp:504-714) and class LI_BA_OptimizerGravity (voxel_map.hpp:717-976)
// ---------------------------------------------------------------------------------------------------------------------
#ifndef VOXELBA_COMPAT_IMU_COEF
extern double imu_coef;   // voxel_map.hpp:500 — the reference's global stays where it is (voxelslam.cpp:927 sets it)
#define VOXELBA_COMPAT_IMU_COEF imu_coef
#endif

namespace vxc {
template <bool GRAVITY>
inline void li_damping(std::vector<IMUST> &x_stats, LidarFactor &voxhess, std::deque<IMU_PRE *> &imus_factor, std::vector<double> *resis,
                       Eigen::MatrixXd *hess, int max_iter) {
  const int W = voxhess.win_size, n = 15 * W + (GRAVITY ? 3 : 0);
  std::vector<vx_imu_state> xs(W);
  for (int i = 0; i < W; i++) xs[i] = state_of(x_stats[i]);
  std::vector<vx_imu_pre> pres(W > 1 ? W - 1 : 0);
  for (int i = 0; i + 1 < W; i++) pres[i] = pre_of(*imus_factor[i]);
  hess->resize(n, n);
  double r2[2] = {0, 0};
  if (GRAVITY) VXC_CHECK(vx_ba_li_gravity(context(), voxhess.device(), xs.data(), pres.data(), VOXELBA_COMPAT_IMU_COEF, r2, hess->data(), max_iter, nullptr, nullptr));
  else VXC_CHECK(vx_ba_li(context(), voxhess.device(), xs.data(), pres.data(), VOXELBA_COMPAT_IMU_COEF, hess->data(), nullptr, nullptr));
  for (int i = 0; i < W; i++) state_to(xs[i], x_stats[i]);
  for (int i = 0; i + 1 < W; i++) pre_back(pres[i], *imus_factor[i]);
  if (GRAVITY && resis) { resis->push_back(r2[0]); resis->push_back(r2[1]); }
  voxhess.pull_cache(0, -1);
}
}  // namespace vxc

class LI_BA_Optimizer {
public:
  int win_size, jac_leng, imu_leng;
  void damping_iter(std::vector<IMUST> &x_stats, LidarFactor &voxhess, std::deque<IMU_PRE *> &imus_factor, Eigen::MatrixXd *hess) {
    win_size = voxhess.win_size; jac_leng = win_size * 6; imu_leng = win_size * 15;
    vxc::li_damping<false>(x_stats, voxhess, imus_factor, nullptr, hess, 3);
  }
};

class LI_BA_OptimizerGravity {
public:
  int win_size, jac_leng, imu_leng;
  void damping_iter(std::vector<IMUST> &x_stats, LidarFactor &voxhess, std::deque<IMU_PRE *> &imus_factor, std::vector<double> &resis,
                    Eigen::MatrixXd *hess, int max_iter = 2) {
    win_size = voxhess.win_size; jac_leng = win_size * 6; imu_leng = win_size * 15 + 3;
    vxc::li_damping<true>(x_stats, voxhess, imus_factor, &resis, hess, max_iter);
  }
};

// ---------------------------------------------------------------------------------------------------------------------
// down_sampling_voxel (tools.hpp:201-238) for any cloud type with a contiguous `points` vector of 48-byte
// pcl::PointXYZINormal-layout records (pcl::PointCloud<PointType> qualifies).  Output order: ascending cell key.
// ---------------------------------------------------------------------------------------------------------------------
template <class Cloud>
inline void down_sampling_voxel(Cloud &pl_feat, double voxel_size) {
  static_assert(sizeof(pl_feat.points[0]) == sizeof(vx_point48), "48-byte PointXYZINormal layout expected");
  const int64_t n = (int64_t)pl_feat.points.size();
  if (n == 0) return;
  std::vector<vx_point48> out(n);
  int64_t m = 0;
  VXC_CHECK(vx_downsample_voxel(vxc::context(), reinterpret_cast<const vx_point48 *>(pl_feat.points.data()), n, voxel_size, out.data(), &m));
  pl_feat.points.resize(m);
  std::memcpy(pl_feat.points.data(), out.data(), sizeof(vx_point48) * m);
  vxc::set_cloud_dims(pl_feat, (uint32_t)m, 0);     // the reference clear()s and push_back()s: width follows the size, height 1
}

using VOX_HESS = LidarFactor;      // north_star's name for the factor

#include "voxelba_compat_map.hpp"
