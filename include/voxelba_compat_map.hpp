// voxelba_compat_map.hpp — the MAP side of the reference's class surface on top of the C ABI (included by voxelba_compat.hpp):
//
//   unordered_map<VOXEL_LOC, OctoTree*> surf_map, surf_map_slide     voxelslam.hpp (members of VOXEL_SLAM)
//   cut_voxel(surf_map, pvec, win_count, surf_map_slide, wdsize, pwld, sws)          voxel_map.hpp:1896-1960
//   cut_voxel_multi(surf_map, pvec, win_count, surf_map_slide, wdsize, pwld, sws)    voxel_map.hpp:1964-2096
//   cut_voxel(surf_map, pvec, wdsize, jour)                                           voxel_map.hpp:2108-2152
//   multi_recut(surf_map_slide, win_count, x_buf, voxhess, sws)                       voxelslam.cpp:1682-1737 (a VOXEL_SLAM member)
//   multi_margi(surf_map_slide, jour, win_count, x_buf, voxhess, sws[0])              voxelslam.cpp:1590-1679 (a VOXEL_SLAM member)
//   match(surf_map, wld, pla, var_wld, sigma_d, oc)                                   voxel_map.hpp:2167-2205
//   lio_state_estimation(surf_map, pptr, x_curr)                                      voxelslam.cpp:962-1098 (member; map + state explicit here)
//   unordered_map<VOXEL_LOC, OctreeGBA*> oct_map; OctreeGBA::cut_voxel(oct_map, x, plptr, i, wdsize)   loop_refine.hpp:446-479
//   OctreeGBA_multi_recut(oct_map, voxhess, thread_num)                               loop_refine.hpp:483-537
//   HBA_add_edge(p_xs, p_smps, gba_edges, maps, max_iter, thread_num, plptr)          voxelslam.cpp:2822-3015
//
// The reference's maps are std::unordered_map of node pointers.  Here the map lives on the device (vx_map / the GBA build), and
// the std::unordered_map the caller declares carries ONE sentinel entry whose OctoTree* / OctreeGBA* owns the device handle — so
// the reference's declarations and call lines compile unchanged.  What the caller must still change (documented in
// INTEGRATION.md): the `mp` ring rotation (voxelslam.cpp:2014-2019) becomes vxc::shift_window(surf_map, mgsize); map clean-up
// (delete of every OctoTree*) becomes vxc::release(surf_map).
// The reference's globals are read where it reads them: voxel_size, min_eigen_value, max_layer, max_points, min_point,
// plane_eigen_value_thre (voxel_map.hpp:98-104), gba_voxel_size, gba_min_eigen_value, gba_eigen_value_array (loop_refine.hpp:269-271).
#pragma once

#include <memory>
#include <unordered_map>

#ifdef VOXELBA_COMPAT_STANDALONE
// ---- stand-ins (tests only) of what tools.hpp / voxel_map.hpp / loop_refine.hpp define around the path -------------------------
struct VOXEL_LOC {                         // tools.hpp:24-35
  int64_t x, y, z;
  VOXEL_LOC(int64_t vx = 0, int64_t vy = 0, int64_t vz = 0) : x(vx), y(vy), z(vz) {}
  bool operator==(const VOXEL_LOC &o) const { return x == o.x && y == o.y && z == o.z; }
};
namespace std {
template <> struct hash<VOXEL_LOC> {      // tools.hpp:37-49
  size_t operator()(const VOXEL_LOC &s) const { return (size_t)(((s.z * 116101) % 10000000000LL + s.y) * 116101 % 10000000000LL + s.x); }
};
}  // namespace std
struct pointVar { Eigen::Vector3d pnt; Eigen::Matrix3d var; };      // voxel_map.hpp:18-34
typedef std::vector<pointVar> PVec;
typedef std::shared_ptr<std::vector<pointVar>> PVecPtr;
struct Plane {                             // voxel_map.hpp:83-96 (members match() hands out)
  Eigen::Vector3d center, normal;
  Eigen::FixedMat<6, 6> plane_var;
  float radius = 0;
  bool is_plane = false;
};
class SlideWindow {};                      // voxel_map.hpp:1009-1042: recycled by the reference; only its container sizes matter here
extern double voxel_size, min_eigen_value;
extern int max_layer, max_points;
extern Eigen::FixedMat<4, 1> min_point;
extern std::vector<double> plane_eigen_value_thre;
extern double gba_voxel_size, gba_min_eigen_value;
extern std::vector<double> gba_eigen_value_array;
struct PointType { float x, y, z, _p0, normal_x, normal_y, normal_z, _p1, intensity, curvature, _p2, _p3; };   // pcl::PointXYZINormal
namespace pcl {
template <class P> struct PointCloud {
  typedef std::shared_ptr<PointCloud<P>> Ptr;
  std::vector<P> points; uint32_t width = 0, height = 1; bool is_dense = true;
  size_t size() const { return points.size(); }
  void clear() { points.clear(); width = 0; height = 1; }
  void reserve(size_t n) { points.reserve(n); }
  void push_back(const P &p) { points.push_back(p); width = (uint32_t)points.size(); height = 1; }
};
}  // namespace pcl
struct Keyframe {                          // loop_refine.hpp (members HBA_add_edge touches)
  IMUST x0; pcl::PointCloud<PointType>::Ptr plptr; int id = 0, mp = 0;
  explicit Keyframe(const IMUST &x) : x0(x), plptr(new pcl::PointCloud<PointType>()) {}
};
struct PGO_Edge_Record { int m1, m2, id1, id2; Eigen::Matrix3d rot; Eigen::Vector3d tra; Eigen::FixedMat<6, 1> cov; };
struct PGO_Edges {                         // loop_refine.hpp:206-235 flattened: one record per push
  std::vector<PGO_Edge_Record> records;
  void push(int m1, int m2, int id1, int id2, Eigen::Matrix3d &rot, Eigen::Vector3d &tra, Eigen::FixedMat<6, 1> &v6) { records.push_back({m1, m2, id1, id2, rot, tra, v6}); }
};
#endif  // VOXELBA_COMPAT_STANDALONE

// ---- sentinel node types: the only OctoTree / OctreeGBA objects that ever exist own the device handles ----------------------------
class OctoTree {
public:
  vx_map *map = nullptr;                   // the sliding-window map (surf_map and surf_map_slide share it)
  bool owner = false;
  int win_size = 0;
  Plane plane;                             // what `pla` of match() points at
  ~OctoTree() { if (owner && map) vx_map_free(map); }
};
class OctreeGBA {
public:
  std::vector<float> xyz; std::vector<int64_t> off{0}; std::vector<vx_pose> poses; int wdsize = 0;
  static void cut_voxel(std::unordered_map<VOXEL_LOC, OctreeGBA *> &feat_map, IMUST &xx, pcl::PointCloud<PointType>::Ptr plptr, int fnum, int wdsize_) {
    OctreeGBA *&acc = feat_map[VOXEL_LOC(0, 0, 0)];
    if (!acc) { acc = new OctreeGBA(); acc->wdsize = wdsize_; }
    if ((int)acc->poses.size() != fnum) { std::printf("OctreeGBA::cut_voxel: frames must arrive in order\n"); std::exit(0); }
    for (const PointType &ap : plptr->points) { acc->xyz.push_back(ap.x); acc->xyz.push_back(ap.y); acc->xyz.push_back(ap.z); }
    acc->off.push_back((int64_t)acc->xyz.size() / 3);
    acc->poses.push_back(vxc::pose_of(xx));
  }
};

namespace vxc {
typedef std::unordered_map<VOXEL_LOC, OctoTree *> SurfMap;
inline OctoTree *sentinel(SurfMap &m) { auto it = m.find(VOXEL_LOC(0, 0, 0)); return it == m.end() ? nullptr : it->second; }
inline vx_map_params map_params(int wdsize, int thread_num) {
  vx_map_params p;
  p.voxel_size = voxel_size; p.min_eigen_value = min_eigen_value;
  for (int k = 0; k < 4; k++) { p.plane_eigen_value_thre[k] = plane_eigen_value_thre[k]; p.min_point[k] = min_point[k]; }
  p.max_layer = max_layer; p.max_points = max_points; p.win_size = wdsize; p.thread_num = thread_num;
  return p;
}
// the device map behind (surf_map, surf_map_slide); created by the first insertion
inline vx_map *device_map(SurfMap &feat_map, SurfMap *slide, int wdsize, int thread_num) {
  OctoTree *s = sentinel(feat_map);
  if (!s) {
    s = new OctoTree(); s->owner = true; s->win_size = wdsize;
    const vx_map_params p = map_params(wdsize, thread_num);
    VXC_CHECK(vx_map_create(context(), &p, &s->map));
    feat_map[VOXEL_LOC(0, 0, 0)] = s;
  }
  if (slide && !sentinel(*slide)) { OctoTree *v = new OctoTree(); v->map = s->map; v->win_size = s->win_size; (*slide)[VOXEL_LOC(0, 0, 0)] = v; }
  return s->map;
}
inline vx_map *device_map(SurfMap &m) { OctoTree *s = sentinel(m); if (!s) { std::printf("libvoxelba: map used before the first cut_voxel\n"); std::exit(0); } return s->map; }
inline void shift_window(SurfMap &m, int mgsize) { VXC_CHECK(vx_map_shift_window(device_map(m), mgsize)); }   // the `mp` rotation, voxelslam.cpp:2014-2019
inline void release(SurfMap &m) { for (auto &kv : m) delete kv.second; m.clear(); }
inline std::vector<vx_pose> poses_of(const std::vector<IMUST> &xs, int n) {
  std::vector<vx_pose> ps(n);
  for (int i = 0; i < n; i++) ps[i] = pose_of(xs[i]);
  return ps;
}
}  // namespace vxc

// voxel_map.hpp:1896-1960 / :1964-2096 (the two differ only in how the reference threads the insertion)
inline void cut_voxel(vxc::SurfMap &feat_map, PVecPtr pvec, int win_count, vxc::SurfMap &feat_tem_map, int wdsize, PLV(3) &pwld,
                      std::vector<SlideWindow *> &sws) {
  (void)sws;
  static_assert(sizeof(pointVar) == sizeof(vx_point_var), "pointVar layout");
  vx_map *m = vxc::device_map(feat_map, &feat_tem_map, wdsize, 1);
  VXC_CHECK(vx_map_cut_voxel(vxc::context(), m, reinterpret_cast<const vx_point_var *>(pvec->data()), pwld.empty() ? nullptr : pwld[0].data(), (int64_t)pvec->size(), win_count));
}
inline void cut_voxel_multi(vxc::SurfMap &feat_map, PVecPtr pvec, int win_count, vxc::SurfMap &feat_tem_map, int wdsize, PLV(3) &pwld,
                            std::vector<std::vector<SlideWindow *>> &sws) {
  vx_map *m = vxc::device_map(feat_map, &feat_tem_map, wdsize, (int)sws.size());
  VXC_CHECK(vx_map_cut_voxel(vxc::context(), m, reinterpret_cast<const vx_point_var *>(pvec->data()), pwld.empty() ? nullptr : pwld[0].data(), (int64_t)pvec->size(), win_count));
}
// voxel_map.hpp:2108-2152
inline void cut_voxel(vxc::SurfMap &feat_map, PVec &pvec, int wdsize, double jour) {
  vx_map *m = vxc::device_map(feat_map, nullptr, wdsize, 1);
  VXC_CHECK(vx_map_cut_voxel_fix(vxc::context(), m, reinterpret_cast<const vx_point_var *>(pvec.data()), (int64_t)pvec.size(), jour));
}
// voxelslam.cpp:1682-1737: recut of every root of the slide set + tras_opt; the LidarFactor is filled ON THE DEVICE
inline void multi_recut(vxc::SurfMap &feat_map, int win_count, std::vector<IMUST> &xs, LidarFactor &voxopt, std::vector<std::vector<SlideWindow *>> &sws) {
  (void)sws;
  const std::vector<vx_pose> ps = vxc::poses_of(xs, win_count);
  VXC_CHECK(vx_map_recut(vxc::context(), vxc::device_map(feat_map), ps.data(), win_count, voxopt.device_for_fill()));
}
// voxelslam.cpp:1590-1679
inline void multi_margi(vxc::SurfMap &feat_map, double jour, int win_count, std::vector<IMUST> &xs, LidarFactor &voxopt, std::vector<SlideWindow *> &sw, int mgsize = 1) {
  (void)sw;
  const std::vector<vx_pose> ps = vxc::poses_of(xs, win_count);
  const int rc = vx_map_margi(vxc::context(), vxc::device_map(feat_map), ps.data(), win_count, mgsize, voxopt.device(), jour);
  if (rc == VX_ERR_OPT_STATE) { std::printf("Error: opt_state\n"); std::exit(0); }     // voxel_map.hpp:1488-1492
  VXC_CHECK(rc);
}
// voxel_map.hpp:2167-2205.  `pla` points at the Plane copy kept in the map's sentinel (valid until the next match on this map);
// `oc` is set to the sentinel (non-null) on a match: the reference only tests and re-uses it as a cache hint.
inline int match(vxc::SurfMap &feat_map, Eigen::Vector3d &wld, Plane *&pla, Eigen::Matrix3d &var_wld, double &sigma_d, OctoTree *&oc) {
  OctoTree *s = vxc::sentinel(feat_map);
  if (!s) return 0;
  vx_planemap *view = nullptr;
  VXC_CHECK(vx_map_planemap(s->map, &view));
  int32_t flag = 0, leaf = -1; double sg = 0; vx_map_node nd;
  VXC_CHECK(vx_planemap_match(vxc::context(), view, wld.data(), var_wld.data(), 1, &flag, &leaf, &sg, &nd));
  if (flag) {
    std::memcpy(s->plane.center.data(), nd.p_center, 24); std::memcpy(s->plane.normal.data(), nd.p_normal, 24);
    std::memcpy(s->plane.plane_var.data(), nd.plane_var, sizeof(nd.plane_var));
    s->plane.radius = nd.radius; s->plane.is_plane = nd.is_plane != 0;
    pla = &s->plane; sigma_d = sg; oc = s;
  }
  return flag;
}
// voxelslam.cpp:962-1098 with the map and the state explicit (a member of VOXEL_SLAM in the reference): x_curr's R, p, v, bg, ba, cov
// are updated; returns the reference's bool.
inline bool lio_state_estimation(vxc::SurfMap &surf_map, PVecPtr pptr, IMUST &x_curr) {
  vx_planemap *view = nullptr;
  VXC_CHECK(vx_map_planemap(vxc::device_map(surf_map), &view));
  vx_lio_state st;
  std::memcpy(st.R, x_curr.R.data(), 72); std::memcpy(st.p, x_curr.p.data(), 24); std::memcpy(st.v, x_curr.v.data(), 24);
  std::memcpy(st.bg, x_curr.bg.data(), 24); std::memcpy(st.ba, x_curr.ba.data(), 24); std::memcpy(st.cov, x_curr.cov.data(), sizeof(st.cov));
  int32_t healthy = 0, iters = 0, mn[4] = {0, 0, 0, 0};
  VXC_CHECK(vx_lio_state_estimation(vxc::context(), view, reinterpret_cast<const vx_point_var *>(pptr->data()), (int64_t)pptr->size(), &st, &healthy, &iters, mn));
  std::memcpy(x_curr.R.data(), st.R, 72); std::memcpy(x_curr.p.data(), st.p, 24); std::memcpy(x_curr.v.data(), st.v, 24);
  std::memcpy(x_curr.bg.data(), st.bg, 24); std::memcpy(x_curr.ba.data(), st.ba, 24); std::memcpy(x_curr.cov.data(), st.cov, sizeof(st.cov));
  return healthy != 0;
}

// loop_refine.hpp:483-537: OctreeGBA::recut of every root + push_voxel of every accepted plane, on the device; the octrees are consumed
inline void OctreeGBA_multi_recut(std::unordered_map<VOXEL_LOC, OctreeGBA *> &oct_map, LidarFactor &voxhess, int thd_num) {
  (void)thd_num;
  auto it = oct_map.find(VOXEL_LOC(0, 0, 0));
  if (it == oct_map.end() || !it->second) return;
  OctreeGBA *acc = it->second;
  vx_gba_params gp;
  gp.voxel_size = gba_voxel_size; gp.min_eigen_value = gba_min_eigen_value; gp.max_layer = max_layer; gp._pad = 0;
  for (int k = 0; k < 4; k++) gp.eigen_value_array[k] = gba_eigen_value_array[k];
  voxhess.win_size = acc->wdsize;
  VXC_CHECK(vx_gba_build(vxc::context(), acc->xyz.data(), acc->off.data(), (int)acc->poses.size(), nullptr, acc->poses.data(), &gp, voxhess.device_for_fill()));
  delete acc;
  oct_map.clear();
}

// voxelslam.cpp:2822-3015 as one library call (session filter :2836-2857 included): xs <- optimized poses of the connected keyframes
// are written back into p_xs, edges pushed into gba_edges, the merged submap cloud into plptr when given.
template <class KF, class EDGES>
inline void HBA_add_edge(std::vector<IMUST> &p_xs, std::vector<KF *> &p_smps, EDGES &gba_edges, std::vector<int> &maps, int max_iter, int thread_num,
                         pcl::PointCloud<PointType>::Ptr plptr = nullptr) {
  std::vector<int> sel;
  int last_mp = -1, isCnct = 0;
  for (int i = 0; i < (int)p_smps.size(); i++) {
    KF *smp = p_smps[i];
    if (smp->mp != last_mp) { isCnct = 0; for (int &m : maps) if (smp->mp == m) { isCnct = 1; break; } last_mp = smp->mp; }
    if (isCnct) sel.push_back(i);
  }
  const int wdsize = (int)sel.size();
  if (wdsize < 2) return;
  std::vector<float> xyz; std::vector<int64_t> off{0}; std::vector<int32_t> ids(wdsize), mps(wdsize); std::vector<vx_pose> ps(wdsize);
  for (int a = 0; a < wdsize; a++) {
    KF *kf = p_smps[sel[a]];
    for (const PointType &p : kf->plptr->points) { xyz.push_back(p.x); xyz.push_back(p.y); xyz.push_back(p.z); }
    off.push_back((int64_t)xyz.size() / 3); ids[a] = kf->id; mps[a] = kf->mp; ps[a] = vxc::pose_of(p_xs[sel[a]]);
  }
  vx_hba_params prm;
  prm.gba.voxel_size = gba_voxel_size; prm.gba.min_eigen_value = gba_min_eigen_value; prm.gba.max_layer = max_layer; prm.gba._pad = 0;
  for (int k = 0; k < 4; k++) { prm.gba.eigen_value_array[k] = gba_eigen_value_array[k]; prm.plane_eigen_value_thre[k] = plane_eigen_value_thre[k]; }
  prm.voxel_size = voxel_size; prm.min_eigen_value = min_eigen_value;
  std::vector<vx_pgo_edge> edges((size_t)wdsize * (wdsize - 1) / 2);
  int64_t n_edges = 0; int32_t status = 0;
  VXC_CHECK(vx_hba_add_edge(vxc::context(), xyz.data(), off.data(), wdsize, nullptr, ids.data(), mps.data(), ps.data(), wdsize, 1, &prm, max_iter, thread_num,
                            edges.data(), &n_edges, nullptr, nullptr, nullptr, &status));
  if (status == VX_ERR_TOO_FEW) { std::printf("Too Less Voxel"); std::exit(0); }      // voxel_map.hpp:399-403
  for (int a = 0; a < wdsize; a++) { std::memcpy(p_xs[sel[a]].R.data(), ps[a].R, 72); std::memcpy(p_xs[sel[a]].p.data(), ps[a].p, 24); }
  for (int64_t e = 0; e < n_edges; e++) {
    Eigen::Matrix3d rot; Eigen::Vector3d tra; decltype(vxc::vec6()) v6 = vxc::vec6();
    std::memcpy(rot.data(), edges[e].rot, 72); std::memcpy(tra.data(), edges[e].tra, 24); std::memcpy(v6.data(), edges[e].cov, 48);
    gba_edges.push(edges[e].m1, edges[e].m2, edges[e].id1, edges[e].id2, rot, tra, v6);
  }
  if (plptr != nullptr) {            // voxelslam.cpp:2954-2985
    std::vector<vx_point48> cells(std::max<size_t>(xyz.size() / 3, 1));
    int64_t wo[2] = {0, 0};
    VXC_CHECK(vx_hba_merge_clouds(vxc::context(), xyz.data(), off.data(), wdsize, nullptr, mps.data(), ps.data(), wdsize, 1, voxel_size, cells.data(), (int64_t)cells.size(), wo));
    plptr->clear(); plptr->reserve((size_t)wo[1]);
    for (int64_t k = 0; k < wo[1]; k++) { PointType ap; std::memcpy(&ap, &cells[k], sizeof(ap)); plptr->push_back(ap); }
  }
}
