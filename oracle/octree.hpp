// ORACLE — TEST INFRASTRUCTURE ONLY (see vmath.hpp header). PARITY UNPINNED.
//
// octree.hpp — restatement of the sliding-window voxel map:
//   /root/reference/VoxelSLAM/src/voxel_map.hpp:39-81      down_sampling_pvec
//   /root/reference/VoxelSLAM/src/voxel_map.hpp:106-121    Bf_var
//   /root/reference/VoxelSLAM/src/voxel_map.hpp:1009-1042  SlideWindow
//   /root/reference/VoxelSLAM/src/voxel_map.hpp:1047-1881  OctoTree
//   /root/reference/VoxelSLAM/src/voxel_map.hpp:1896-2205  cut_voxel x2, cut_voxel_multi, match
//   /root/reference/VoxelSLAM/src/voxelslam.cpp:1590-1737  multi_margi, multi_recut
//   /root/reference/VoxelSLAM/src/tools.hpp:201-302        down_sampling_voxel / _close
// The reference's globals (`mp`, voxel_size, ...) are carried in a MapCtx passed explicitly.
// Mutexes are dropped: the oracle runs the per-root work serially in the reference's own
// unordered_map iteration order, which is one of the orders the threaded reference can produce
// for everything that is order-sensitive (per-leaf sums are private to a root voxel).
#pragma once
#include "ba.hpp"

namespace orc {

struct PointXYZIN {  // pcl::PointXYZINormal, 48 B: data[4], normal[4], intensity, curvature, pad[2]
  float x, y, z, _p0;
  float normal_x, normal_y, normal_z, _p1;
  float intensity, curvature, _p2, _p3;
};
static_assert(sizeof(PointXYZIN) == 48, "PointXYZINormal layout");

struct MapCtx {
  Params prm;
  std::vector<int> mp;   // voxel_map.hpp:1046 ring index, voxelslam.cpp:3158-3160 / :2014-2019
};

// voxel_map.hpp:106-121
inline void Bf_var(const pointVar &pv, Mat<9, 9> &bcov, const V3 &vec) {
  Mat<6, 3> Bi;
  Bi(0, 0) = 2 * vec[0];
  Bi(1, 0) = vec[1]; Bi(1, 1) = vec[0];
  Bi(2, 0) = vec[2]; Bi(2, 2) = vec[0];
  Bi(3, 1) = 2 * vec[1];
  Bi(4, 1) = vec[2]; Bi(4, 2) = vec[1];
  Bi(5, 2) = 2 * vec[2];
  Mat<6, 3> Biup = Bi * pv.var;
  bcov.setBlock(0, 0, Biup * Bi.transpose());
  bcov.setBlock(0, 6, Biup);
  bcov.setBlock(6, 0, Biup.transpose());
  bcov.setBlock(6, 6, pv.var);
}

class SlideWindow {  // voxel_map.hpp:1009-1042
public:
  std::vector<PVec> points;
  std::vector<PointCluster> pcrs_local;
  explicit SlideWindow(int wdsize) { pcrs_local.resize(wdsize); points.resize(wdsize); }
  void resize(int wdsize) { if ((int)points.size() != wdsize) { points.resize(wdsize); pcrs_local.resize(wdsize); } }
  void clear() { for (size_t i = 0; i < points.size(); i++) { points[i].clear(); pcrs_local[i].clear(); } }
};

class OctoTree {
public:
  SlideWindow *sw = nullptr;
  PointCluster pcr_add;
  Mat<9, 9> cov_add;
  PointCluster pcr_fix;
  PVec point_fix;
  int layer, octo_state, wdsize;
  OctoTree *leaves[8];
  double voxel_center[3];
  double jour = 0;
  float quater_length;
  Plane plane;
  bool isexist = false;
  V3 eig_value; M3 eig_vector;
  int last_num = 0, opt_state = -1;
  MapCtx *ctx;

  OctoTree(int _l, int _w, MapCtx *c) : layer(_l), octo_state(0), wdsize(_w), ctx(c) {
    for (int i = 0; i < 8; i++) leaves[i] = nullptr;
  }
  ~OctoTree() {}

  // voxel_map.hpp:1105-1143
  void push(int ord, const pointVar &pv, const V3 &pw, std::vector<SlideWindow *> &sws) {
    if (sw == nullptr) {
      if (!sws.empty()) { sw = sws.back(); sws.pop_back(); sw->resize(wdsize); }
      else sw = new SlideWindow(wdsize);
    }
    if (!isexist) isexist = true;
    int mord = ctx->mp[ord];
    if (layer < ctx->prm.max_layer) sw->points[mord].push_back(pv);
    sw->pcrs_local[mord].push(pv.pnt);
    pcr_add.push(pw);
    Mat<9, 9> Bi;
    Bf_var(pv, Bi, pw);
    cov_add += Bi;
  }
  // voxel_map.hpp:1149-1162
  void push_fix(const pointVar &pv) {
    if (layer < ctx->prm.max_layer) point_fix.push_back(pv);
    pcr_fix.push(pv.pnt);
    pcr_add.push(pv.pnt);
    Mat<9, 9> Bi;
    Bf_var(pv, Bi, pv.pnt);
    cov_add += Bi;
  }
  // voxel_map.hpp:1168-1178
  void push_fix_novar(const pointVar &pv) {
    if (layer < ctx->prm.max_layer) point_fix.push_back(pv);
    pcr_fix.push(pv.pnt);
    pcr_add.push(pv.pnt);
  }
  // voxel_map.hpp:1185-1195
  bool plane_judge(const V3 &ev) const {
    return (ev[0] < ctx->prm.min_eigen_value && (ev[0] / ev[2]) < ctx->prm.plane_eigen_value_thre[layer]);
  }
  int child_of(const V3 &pw, int xyz[3]) const {  // voxel_map.hpp:1214-1219
    for (int k = 0; k < 3; k++) xyz[k] = (pw[k] > voxel_center[k]) ? 1 : 0;
    return 4 * xyz[0] + 2 * xyz[1] + xyz[2];
  }
  OctoTree *ensure_leaf(int leafnum, const int xyz[3]) {  // voxel_map.hpp:1222-1232
    if (leaves[leafnum] == nullptr) {
      OctoTree *c = new OctoTree(layer + 1, wdsize, ctx);
      c->voxel_center[0] = voxel_center[0] + (2 * xyz[0] - 1) * quater_length;
      c->voxel_center[1] = voxel_center[1] + (2 * xyz[1] - 1) * quater_length;
      c->voxel_center[2] = voxel_center[2] + (2 * xyz[2] - 1) * quater_length;
      c->quater_length = quater_length / 2;
      leaves[leafnum] = c;
    }
    return leaves[leafnum];
  }
  // voxel_map.hpp:1204-1237
  void allocate(int ord, const pointVar &pv, const V3 &pw, std::vector<SlideWindow *> &sws) {
    if (octo_state == 0) { push(ord, pv, pw, sws); return; }
    int xyz[3]; int leafnum = child_of(pw, xyz);
    ensure_leaf(leafnum, xyz)->allocate(ord, pv, pw, sws);
  }
  // voxel_map.hpp:1239-1264
  void allocate_fix(const pointVar &pv) {
    if (octo_state == 0) { push_fix_novar(pv); }
    else if (layer < ctx->prm.max_layer) {
      int xyz[3]; int leafnum = child_of(pv.pnt, xyz);
      ensure_leaf(leafnum, xyz)->allocate_fix(pv);
    }
  }
  // voxel_map.hpp:1270-1299
  void fix_divide(std::vector<SlideWindow *> &) {
    for (pointVar &pv : point_fix) {
      int xyz[3]; int leafnum = child_of(pv.pnt, xyz);
      ensure_leaf(leafnum, xyz)->push_fix(pv);
    }
  }
  // voxel_map.hpp:1307-1338
  void subdivide(int si, const IMUST &xx, std::vector<SlideWindow *> &sws) {
    for (pointVar &pv : sw->points[ctx->mp[si]]) {
      V3 pw = xx.R * pv.pnt + xx.p;
      int xyz[3]; int leafnum = child_of(pw, xyz);
      ensure_leaf(leafnum, xyz)->push(si, pv, pw, sws);
    }
  }
  // voxel_map.hpp:1344-1388
  void plane_update() {
    plane.center = pcr_add.v / (double)pcr_add.N;
    const int l = 0;
    V3 u[3] = {eig_vector.col(0), eig_vector.col(1), eig_vector.col(2)};
    double nv = 1.0 / pcr_add.N;
    Mat<3, 9> u_c;
    for (int k = 0; k < 3; k++)
      if (k != l) {
        M3 ukl = u[k] * u[l].transpose();
        Mat<1, 9> fkl;
        fkl[0] = ukl(0, 0); fkl[1] = ukl(1, 0) + ukl(0, 1); fkl[2] = ukl(2, 0) + ukl(0, 2);
        fkl[3] = ukl(1, 1); fkl[4] = ukl(1, 2) + ukl(2, 1); fkl[5] = ukl(2, 2);
        V3 tail = -(u[k].dot(plane.center) * u[l] + u[l].dot(plane.center) * u[k]);
        fkl[6] = tail[0]; fkl[7] = tail[1]; fkl[8] = tail[2];
        u_c += nv / (eig_value[l] - eig_value[k]) * (u[k] * fkl);
      }
    Mat<3, 9> Jc = u_c * cov_add;
    plane.plane_var.setBlock(0, 0, Jc * u_c.transpose());
    M3 Jc_N = nv * Jc.block<3, 3>(0, 6);
    plane.plane_var.setBlock(0, 3, Jc_N);
    plane.plane_var.setBlock(3, 0, Jc_N.transpose());
    plane.plane_var.setBlock(3, 3, nv * nv * cov_add.block<3, 3>(6, 6));
    plane.normal = u[0];
    plane.radius = (float)eig_value[2];
  }
  // voxel_map.hpp:1396-1456
  void recut(int win_count, std::vector<IMUST> &x_buf, std::vector<SlideWindow *> &sws) {
    if (octo_state == 0) {
      if (layer >= 0) {
        opt_state = -1;
        if (pcr_add.N <= ctx->prm.min_point[layer]) { plane.is_plane = false; return; }
        if (!isexist || sw == nullptr) return;
        eig3_sym(pcr_add.cov(), eig_value, eig_vector);
        plane.is_plane = plane_judge(eig_value);
        if (plane.is_plane) return;
        else if (layer >= ctx->prm.max_layer) return;
      }
      if (pcr_fix.N != 0) { fix_divide(sws); PVec().swap(point_fix); }
      for (int i = 0; i < win_count; i++) subdivide(i, x_buf[i], sws);
      sw->clear(); sws.push_back(sw); sw = nullptr;
      octo_state = 1;
    }
    for (int i = 0; i < 8; i++) if (leaves[i] != nullptr) leaves[i]->recut(win_count, x_buf, sws);
  }
  // voxel_map.hpp:1465-1598.  Returns false where the reference would printf+exit(0) (:1488-1492).
  bool margi(int win_count, int mgsize, std::vector<IMUST> &x_buf, const LidarFactor &vox_opt) {
    if (octo_state == 0 && layer >= 0) {
      if (!isexist || sw == nullptr) return true;
      std::vector<PointCluster> pcrs_world(wdsize);
      if (opt_state >= (int)vox_opt.pcr_adds.size()) return false;
      const std::vector<int> &mp = ctx->mp;
      if (opt_state >= 0) {
        pcr_add = vox_opt.pcr_adds[opt_state];
        eig_value = vox_opt.eig_values[opt_state];
        eig_vector = vox_opt.eig_vectors[opt_state];
        opt_state = -1;
        for (int i = 0; i < mgsize; i++)
          if (sw->pcrs_local[mp[i]].N != 0) pcrs_world[i].transform(sw->pcrs_local[mp[i]], x_buf[i]);
      } else {
        pcr_add = pcr_fix;
        for (int i = 0; i < win_count; i++)
          if (sw->pcrs_local[mp[i]].N != 0) {
            pcrs_world[i].transform(sw->pcrs_local[mp[i]], x_buf[i]);
            pcr_add += pcrs_world[i];
          }
        if (plane.is_plane) eig3_sym(pcr_add.cov(), eig_value, eig_vector);
      }
      if (pcr_fix.N < ctx->prm.max_points && plane.is_plane)
        if (pcr_add.N - last_num >= 5 || last_num <= 10) { plane_update(); last_num = pcr_add.N; }

      if (pcr_fix.N < ctx->prm.max_points) {
        for (int i = 0; i < mgsize; i++)
          if (pcrs_world[i].N != 0) {
            pcr_fix += pcrs_world[i];
            for (pointVar pv : sw->points[mp[i]]) {
              pv.pnt = x_buf[i].R * pv.pnt + x_buf[i].p;
              point_fix.push_back(pv);
            }
          }
      } else {
        for (int i = 0; i < mgsize; i++) if (pcrs_world[i].N != 0) pcr_add -= pcrs_world[i];
        if (point_fix.size() != 0) PVec().swap(point_fix);
      }
      for (int i = 0; i < mgsize; i++)
        if (sw->pcrs_local[mp[i]].N != 0) { sw->pcrs_local[mp[i]].clear(); sw->points[mp[i]].clear(); }
      isexist = !(pcr_fix.N >= pcr_add.N);
      return true;
    } else {
      isexist = false;
      bool ok = true;
      for (int i = 0; i < 8; i++)
        if (leaves[i] != nullptr) {
          ok = leaves[i]->margi(win_count, mgsize, x_buf, vox_opt) && ok;
          isexist = isexist || leaves[i]->isexist;
        }
      return ok;
    }
  }
  // voxel_map.hpp:1605-1638
  void tras_opt(LidarFactor &vox_opt) {
    if (octo_state == 0) {
      if (layer >= 0 && isexist && plane.is_plane && sw != nullptr) {
        if (eig_value[0] / eig_value[1] > 0.12) return;
        double coe = 1;
        std::vector<PointCluster> pcrs(wdsize);
        for (int i = 0; i < wdsize; i++) pcrs[i] = sw->pcrs_local[ctx->mp[i]];
        opt_state = (int)vox_opt.plvec_voxels.size();
        vox_opt.push_voxel(pcrs, pcr_fix, coe, eig_value, eig_vector, pcr_add);
      }
    } else {
      for (int i = 0; i < 8; i++) if (leaves[i] != nullptr) leaves[i]->tras_opt(vox_opt);
    }
  }
  // voxel_map.hpp:1649-1721
  int match(const V3 &wld, Plane *&pla, double &, const M3 &var_wld, double &sigma_d, OctoTree *&oc) {
    int flag = 0;
    if (octo_state == 0) {
      if (plane.is_plane) {
        float dis_to_plane = (float)std::fabs(plane.normal.dot(wld - plane.center));
        float dis_to_center = (float)(plane.center - wld).squaredNorm();
        float range_dis = (dis_to_center - dis_to_plane * dis_to_plane);
        if (range_dis <= 3 * 3 * plane.radius) {
          Mat<1, 6> J_nq;
          V3 d = wld - plane.center;
          for (int k = 0; k < 3; k++) { J_nq[k] = d[k]; J_nq[3 + k] = -plane.normal[k]; }
          double sigma_l = (J_nq * plane.plane_var * J_nq.transpose())[0];
          sigma_l += (plane.normal.transpose() * var_wld * plane.normal)[0];
          if (dis_to_plane < 3 * std::sqrt(sigma_l)) {
            oc = this; sigma_d = sigma_l; pla = &plane;
            flag = 1;
          }
        }
      }
    } else {
      int xyz[3]; int leafnum = child_of(wld, xyz);
      if (leaves[leafnum] != nullptr) { double mpb = 0; flag = leaves[leafnum]->match(wld, pla, mpb, var_wld, sigma_d, oc); }
    }
    return flag;
  }
  // voxel_map.hpp:1729-1745
  void tras_ptr(std::vector<OctoTree *> &out) {
    if (octo_state == 1)
      for (int i = 0; i < 8; i++) if (leaves[i] != nullptr) { out.push_back(leaves[i]); leaves[i]->tras_ptr(out); }
  }
  // voxel_map.hpp:1836-1848
  bool inside(const V3 &wld) const {
    double hl = quater_length * 2;
    return (wld[0] >= voxel_center[0] - hl && wld[0] <= voxel_center[0] + hl &&
            wld[1] >= voxel_center[1] - hl && wld[1] <= voxel_center[1] + hl &&
            wld[2] >= voxel_center[2] - hl && wld[2] <= voxel_center[2] + hl);
  }
  // voxel_map.hpp:1856-1880
  void clear_slwd(std::vector<SlideWindow *> &sws) {
    if (octo_state != 0) for (int i = 0; i < 8; i++) if (leaves[i] != nullptr) leaves[i]->clear_slwd(sws);
    if (sw != nullptr) { sw->clear(); sws.push_back(sw); sw = nullptr; }
  }
};

using SurfMap = VoxelMap<OctoTree *>;

inline OctoTree *new_root(const VOXEL_LOC &position, int wdsize, MapCtx *ctx) {  // voxel_map.hpp:1935-1943
  OctoTree *ot = new OctoTree(0, wdsize, ctx);
  ot->voxel_center[0] = (0.5 + position.x) * ctx->prm.voxel_size;
  ot->voxel_center[1] = (0.5 + position.y) * ctx->prm.voxel_size;
  ot->voxel_center[2] = (0.5 + position.z) * ctx->prm.voxel_size;
  ot->quater_length = (float)(ctx->prm.voxel_size / 4.0);
  return ot;
}

// voxel_map.hpp:1896-1949 (single-thread window insertion)
inline void cut_voxel(SurfMap &feat_map, PVec &pvec, int win_count, SurfMap &feat_tem_map, int wdsize,
                      std::vector<V3> &pwld, std::vector<SlideWindow *> &sws, MapCtx *ctx) {
  int plsize = (int)pvec.size();
  for (int i = 0; i < plsize; i++) {
    pointVar &pv = pvec[i];
    V3 &pw = pwld[i];
    VOXEL_LOC position = voxel_of(pw, ctx->prm.voxel_size);
    auto iter = feat_map.find(position);
    if (iter != feat_map.end()) {
      iter->second->allocate(win_count, pv, pw, sws);
      iter->second->isexist = true;
      if (feat_tem_map.find(position) == feat_tem_map.end()) feat_tem_map[position] = iter->second;
    } else {
      // the reference allocates before setting the centre; a fresh root is a leaf so order is moot
      OctoTree *ot = new_root(position, wdsize, ctx);
      ot->allocate(win_count, pv, pw, sws);
      feat_map[position] = ot;
      feat_tem_map[position] = ot;
    }
  }
}

// voxel_map.hpp:1964-2096.  `thd_num` = sws.size() in the reference; the early-out at :2044-2045
// (fewer touched roots than threads => nothing is inserted although the maps were updated) is kept.
inline void cut_voxel_multi(SurfMap &feat_map, PVec &pvec, int win_count, SurfMap &feat_tem_map, int wdsize,
                            std::vector<V3> &pwld, std::vector<std::vector<SlideWindow *>> &sws, MapCtx *ctx) {
  std::unordered_map<OctoTree *, std::vector<int>> map_pvec;
  int plsize = (int)pvec.size();
  for (int i = 0; i < plsize; i++) {
    V3 &pw = pwld[i];
    VOXEL_LOC position = voxel_of(pw, ctx->prm.voxel_size);
    auto iter = feat_map.find(position);
    OctoTree *ot = nullptr;
    if (iter != feat_map.end()) {
      iter->second->isexist = true;
      if (feat_tem_map.find(position) == feat_tem_map.end()) feat_tem_map[position] = iter->second;
      ot = iter->second;
    } else {
      ot = new_root(position, wdsize, ctx);
      feat_map[position] = ot;
      feat_tem_map[position] = ot;
    }
    map_pvec[ot].push_back(i);
  }
  int thd_num = (int)sws.size();
  int g_size = (int)map_pvec.size();
  if (g_size < thd_num) return;
  // per-root insertion order = ascending point index; roots are independent, so one pass suffices
  for (auto &kv : map_pvec)
    for (int k : kv.second) kv.first->allocate(win_count, pvec[k], pwld[k], sws[0]);
}

// voxel_map.hpp:2108-2152 (fixed, already-world-frame points)
inline void cut_voxel_fix(SurfMap &feat_map, PVec &pvec, int wdsize, double jour, MapCtx *ctx) {
  for (pointVar &pv : pvec) {
    VOXEL_LOC position = voxel_of(pv.pnt, ctx->prm.voxel_size);
    auto iter = feat_map.find(position);
    if (iter != feat_map.end()) iter->second->allocate_fix(pv);
    else {
      OctoTree *ot = new_root(position, wdsize, ctx);
      ot->push_fix_novar(pv);
      ot->jour = jour;
      feat_map[position] = ot;
    }
  }
}

// voxel_map.hpp:2167-2205
inline int match(SurfMap &feat_map, const V3 &wld, Plane *&pla, const M3 &var_wld, double &sigma_d, OctoTree *&oc, MapCtx *ctx) {
  int flag = 0;
  VOXEL_LOC position = voxel_of(wld, ctx->prm.voxel_size);
  auto iter = feat_map.find(position);
  if (iter != feat_map.end()) {
    double max_prob = 0;
    flag = iter->second->match(wld, pla, max_prob, var_wld, sigma_d, oc);
  }
  return flag;
}

// voxelslam.cpp:1682-1737 (early-out :1693-1694 kept; serial in map iteration order)
inline void multi_recut(SurfMap &feat_map, int win_count, std::vector<IMUST> &xs, LidarFactor &voxopt,
                        std::vector<std::vector<SlideWindow *>> &sws, int thread_num) {
  int g_size = (int)feat_map.size();
  if (g_size < thread_num) return;
  for (auto iter = feat_map.begin(); iter != feat_map.end(); iter++) iter->second->recut(win_count, xs, sws[0]);
  for (size_t i = 1; i < sws.size(); i++) { sws[0].insert(sws[0].end(), sws[i].begin(), sws[i].end()); sws[i].clear(); }
  for (auto iter = feat_map.begin(); iter != feat_map.end(); iter++) iter->second->tras_opt(voxopt);
}

// voxelslam.cpp:1590-1679 (early-out :1616-1617 kept)
inline bool multi_margi(SurfMap &feat_map, double jour, int win_count, std::vector<IMUST> &xs, LidarFactor &voxopt,
                        std::vector<SlideWindow *> &sw, int thread_num) {
  int g_size = (int)feat_map.size();
  if (g_size < thread_num) return true;
  bool ok = true;
  for (auto iter = feat_map.begin(); iter != feat_map.end(); iter++) {
    iter->second->jour = jour;
    ok = iter->second->margi(win_count, 1, xs, voxopt) && ok;
  }
  for (auto iter = feat_map.begin(); iter != feat_map.end();) {
    if (iter->second->isexist) iter++;
    else { iter->second->clear_slwd(sw); feat_map.erase(iter++); }
  }
  return ok;
}

// tools.hpp:201-238 — sequential float running mean; output in hash-map iteration order.
inline void down_sampling_voxel(std::vector<PointXYZIN> &pl_feat, double voxel_size) {
  if (voxel_size < 0.001) return;
  VoxelMap<PointXYZIN> feat_map;
  for (PointXYZIN &p_c : pl_feat) {
    const float *data = &p_c.x;
    VOXEL_LOC position(key_axis_f(data[0], voxel_size), key_axis_f(data[1], voxel_size), key_axis_f(data[2], voxel_size));
    auto iter = feat_map.find(position);
    if (iter == feat_map.end()) {
      PointXYZIN pp = p_c;
      pp.curvature = 1;
      feat_map[position] = pp;
    } else {
      PointXYZIN &pp = iter->second;
      pp.x = (pp.x * pp.curvature + p_c.x) / (pp.curvature + 1);
      pp.y = (pp.y * pp.curvature + p_c.y) / (pp.curvature + 1);
      pp.z = (pp.z * pp.curvature + p_c.z) / (pp.curvature + 1);
      pp.curvature += 1;
    }
  }
  pl_feat.clear();
  for (auto iter = feat_map.begin(); iter != feat_map.end(); ++iter) pl_feat.push_back(iter->second);
}

// tools.hpp:240-302 — keep the input point nearest the (float) cell centroid.
inline void down_sampling_close(std::vector<PointXYZIN> &pl_feat, double voxel_size) {
  if (voxel_size < 0.001) return;
  VoxelMap<std::vector<PointXYZIN>> feat_map;
  for (PointXYZIN &p_c : pl_feat) {
    const float *data = &p_c.x;
    VOXEL_LOC position(key_axis_f(data[0], voxel_size), key_axis_f(data[1], voxel_size), key_axis_f(data[2], voxel_size));
    feat_map[position].push_back(p_c);
  }
  pl_feat.clear();
  for (auto iter = feat_map.begin(); iter != feat_map.end(); ++iter) {
    std::vector<PointXYZIN> &pl = iter->second;
    PointXYZIN pb = pl[0];
    int plsize = (int)pl.size();
    for (int i = 1; i < plsize; i++) { pb.x += pl[i].x; pb.y += pl[i].y; pb.z += pl[i].z; }
    pb.x /= plsize; pb.y /= plsize; pb.z /= plsize;
    double ndis = 100; int mnum = 0;
    for (int i = 0; i < plsize; i++) {
      double xx = pb.x - pl[i].x, yy = pb.y - pl[i].y, zz = pb.z - pl[i].z;
      double dis = xx * xx + yy * yy + zz * zz;
      if (dis < ndis) { mnum = i; ndis = dis; }
    }
    pl_feat.push_back(pl[mnum]);
  }
}

// voxel_map.hpp:39-81 — double running mean of point and covariance; diag(var) in normal_xyz.
inline void down_sampling_pvec(PVec &pvec, double voxel_size, std::vector<PointXYZIN> &pl_keep) {
  VoxelMap<std::pair<pointVar, int>> feat_map;
  for (pointVar &pv : pvec) {
    VOXEL_LOC position = voxel_of(pv.pnt, voxel_size);
    auto iter = feat_map.find(position);
    if (iter == feat_map.end()) feat_map[position] = std::make_pair(pv, 1);
    else {
      auto &pp = iter->second;
      pp.first.pnt = (pp.first.pnt * (double)pp.second + pv.pnt) / (double)(pp.second + 1);
      pp.first.var = (pp.first.var * (double)pp.second + pv.var) / (double)(pp.second + 1);
      pp.second += 1;
    }
  }
  pl_keep.clear();
  pl_keep.reserve(feat_map.size());
  PointXYZIN ap;
  std::memset(&ap, 0, sizeof(ap));
  (&ap.x)[3] = 1.0f;   // pcl::PointXYZINormal() (PCL 1.10 point_types.hpp) sets data[3] = 1
  for (auto iter = feat_map.begin(); iter != feat_map.end(); ++iter) {
    pointVar &pv = iter->second.first;
    ap.x = (float)pv.pnt[0]; ap.y = (float)pv.pnt[1]; ap.z = (float)pv.pnt[2];
    ap.normal_x = (float)pv.var(0, 0); ap.normal_y = (float)pv.var(1, 1); ap.normal_z = (float)pv.var(2, 2);
    pl_keep.push_back(ap);
  }
}

}  // namespace orc
