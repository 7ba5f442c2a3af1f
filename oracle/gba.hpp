// ORACLE — TEST INFRASTRUCTURE ONLY (see vmath.hpp header). PARITY UNPINNED.
//
// gba.hpp — restatement of the global-BA voxel map and the hierarchical edge builder:
//   /root/reference/VoxelSLAM/src/loop_refine.hpp:273-481  class OctreeGBA
//   /root/reference/VoxelSLAM/src/loop_refine.hpp:483-537  OctreeGBA_multi_recut
//   /root/reference/VoxelSLAM/src/voxelslam.cpp:2822-2951  VOXEL_SLAM::HBA_add_edge (BA + edges)
//   /root/reference/VoxelSLAM/src/voxelslam.cpp:2954-2985  merged submap cloud
//   /root/reference/VoxelSLAM/src/voxelslam.cpp:3028-3139  thd_globalmapping window schedule
#pragma once
#include "octree.hpp"

namespace orc {

class OctreeGBA {
public:
  std::vector<std::vector<V3>> locals, worlds;
  PointCluster pcr_add;
  int layer, octo_state, wdsize;
  OctreeGBA *leaves[8];
  double voxel_center[3];
  float quater_length;
  bool is_plane;
  const Params *prm;

  OctreeGBA(int _l, int _w, const Params *p) : layer(_l), octo_state(0), wdsize(_w), is_plane(false), prm(p) {
    locals.resize(wdsize); worlds.resize(wdsize);
    for (int i = 0; i < 8; i++) leaves[i] = nullptr;
  }
  ~OctreeGBA() { for (int i = 0; i < 8; i++) delete leaves[i]; }

  bool plane_judge(const V3 &ev) const {  // loop_refine.hpp:310-314
    return (ev[0] < prm->gba_min_eigen_value && (ev[0] / ev[2]) < prm->gba_eigen_value_array[layer]);
  }
  void push(int ord, const V3 &local, const V3 &world) {  // loop_refine.hpp:316-321
    locals[ord].push_back(local);
    worlds[ord].push_back(world);
    pcr_add.push(world);
  }
  void subdivide() {  // loop_refine.hpp:323-356 (node recycling through oct_buf dropped: allocation only)
    for (int i = 0; i < wdsize; i++)
      for (size_t j = 0; j < locals[i].size(); j++) {
        V3 &pl = locals[i][j];
        V3 &pw = worlds[i][j];
        int xyz[3] = {0, 0, 0};
        for (int k = 0; k < 3; k++) if (pw[k] > voxel_center[k]) xyz[k] = 1;
        int leafnum = 4 * xyz[0] + 2 * xyz[1] + xyz[2];
        if (leaves[leafnum] == nullptr) {
          OctreeGBA *c = new OctreeGBA(layer + 1, wdsize, prm);
          c->voxel_center[0] = voxel_center[0] + (2 * xyz[0] - 1) * quater_length;
          c->voxel_center[1] = voxel_center[1] + (2 * xyz[1] - 1) * quater_length;
          c->voxel_center[2] = voxel_center[2] + (2 * xyz[2] - 1) * quater_length;
          c->quater_length = quater_length / 2;
          leaves[leafnum] = c;
        }
        leaves[leafnum]->push(i, pl, pw);
      }
  }
  // loop_refine.hpp:358-405
  void recut(LidarFactor &vox_opt) {
    if (pcr_add.N <= 10) return;
    V3 eig_value; M3 eig_vector;
    eig3_sym(pcr_add.cov(), eig_value, eig_vector);
    is_plane = plane_judge(eig_value);
    if (is_plane) {
      if (pcr_add.N < 10) return;
      int exi = 0;
      for (int i = 0; i < wdsize; i++) if (locals[i].size() != 0) exi++;
      if (exi <= 1) return;
      if (eig_value[0] / eig_value[1] > 0.12) return;
      double coe = 1.0;
      std::vector<PointCluster> pcrs(wdsize);
      for (int i = 0; i < wdsize; i++) for (V3 &v : locals[i]) pcrs[i].push(v);
      PointCluster pcr_fix;
      vox_opt.push_voxel(pcrs, pcr_fix, coe, eig_value, eig_vector, pcr_add);
      return;
    } else if (layer >= prm->max_layer) {
      return;
    } else {
      subdivide();
      octo_state = 1;
    }
    for (int i = 0; i < 8; i++) if (leaves[i] != nullptr) leaves[i]->recut(vox_opt);
  }
};

using GbaMap = VoxelMap<OctreeGBA *>;

// loop_refine.hpp:446-479.  Cloud = float xyz triples (PointXYZINormal's x,y,z).
inline void gba_cut_voxel(GbaMap &feat_map, const IMUST &xc, const float *xyz, int npts, int win_count, int wdsize, const Params *prm) {
  for (int n = 0; n < npts; n++) {
    V3 local = vec3(xyz[3 * n + 0], xyz[3 * n + 1], xyz[3 * n + 2]);
    V3 world = xc.R * local + xc.p;
    VOXEL_LOC position = voxel_of(world, prm->gba_voxel_size);
    auto iter = feat_map.find(position);
    if (iter != feat_map.end()) iter->second->push(win_count, local, world);
    else {
      OctreeGBA *ot = new OctreeGBA(0, wdsize, prm);
      ot->push(win_count, local, world);
      ot->voxel_center[0] = (0.5 + position.x) * prm->gba_voxel_size;
      ot->voxel_center[1] = (0.5 + position.y) * prm->gba_voxel_size;
      ot->voxel_center[2] = (0.5 + position.z) * prm->gba_voxel_size;
      ot->quater_length = (float)(prm->gba_voxel_size / 4.0);
      feat_map[position] = ot;
    }
  }
}

// loop_refine.hpp:483-537.  The voxel order of the resulting factor is: map iteration order split
// into thd_num consecutive buckets (bucket boundary rule :491-496), bucket 0 first, then the others
// appended (:529-534) — i.e. plain map iteration order.  The roots are freed as the reference does.
inline void OctreeGBA_multi_recut(GbaMap &feat_map, LidarFactor &voxhess, int /*thd_num*/) {
  for (auto iter = feat_map.begin(); iter != feat_map.end(); iter++) {
    iter->second->recut(voxhess);
    delete iter->second;
    iter->second = nullptr;
  }
}

struct KeyframeCloud {      // the part of `struct Keyframe` (voxel_map.hpp:979-1006) the path uses
  const float *xyz; int n;  // local-frame points
  int id, mp;
};
struct PgoEdge { int m1, m2, id1, id2; M3 rot; V3 tra; V6 cov; };

struct HbaResult {
  std::vector<IMUST> xs;            // optimised poses of the selected keyframes
  MatX hess;                        // last pre-gauge Hessian snapshot
  std::vector<PgoEdge> edges;       // voxelslam.cpp:2926-2951
  std::vector<PointXYZIN> merged;   // voxelslam.cpp:2954-2985 (only if want_cloud)
  std::vector<double> resis_first, resis_last;   // per outer iteration (test introspection)
  std::vector<int> n_voxels;                      // per outer iteration
  bool fatal = false;
};

// voxelslam.cpp:2822-3015.  `base` carries the LocalBA thresholds (voxel_size, min_eigen_value,
// plane_eigen_value_thre, max_layer) AND the GBA ones; the coarse->fine swap (:2873-2883) copies
// the former over the latter for the final / converged pass.
inline void HBA_add_edge(const std::vector<IMUST> &p_xs, const std::vector<KeyframeCloud> &p_smps, const std::vector<int> &maps,
                         int max_iter, int thread_num, bool want_cloud, const Params &base, HbaResult &out, bool use_threads = false) {
  std::vector<KeyframeCloud> smps;
  std::vector<IMUST> xs;
  int last_mp = -1, isCnct = 0;
  for (size_t i = 0; i < p_smps.size(); i++) {
    const KeyframeCloud &smp = p_smps[i];
    if (smp.mp != last_mp) {
      isCnct = 0;
      for (int m : maps) if (smp.mp == m) { isCnct = 1; break; }
      last_mp = smp.mp;
    }
    if (isCnct) { smps.push_back(smp); xs.push_back(p_xs[i]); }
  }
  int wdsize = (int)smps.size();
  MatX hess;
  Params prm = base;
  int up = 4;
  int converge_flag = 0;
  double converge_thre = 0.05;

  for (int iterCnt = 0; iterCnt < max_iter; iterCnt++) {
    if (converge_flag == 1 || iterCnt == max_iter - 1) {
      prm.gba_voxel_size = base.voxel_size;
      for (int k = 0; k < 4; k++) prm.gba_eigen_value_array[k] = base.plane_eigen_value_thre[k];
      prm.gba_min_eigen_value = base.min_eigen_value;
    }
    GbaMap oct_map;
    for (int i = 0; i < wdsize; i++) gba_cut_voxel(oct_map, xs[i], smps[i].xyz, smps[i].n, i, wdsize, &prm);
    LidarFactor voxhess(wdsize);
    OctreeGBA_multi_recut(oct_map, voxhess, thread_num);

    Lidar_BA_Optimizer opt_lsv;
    opt_lsv.thd_num = thread_num;
    opt_lsv.use_threads = use_threads;
    std::vector<double> resis;
    bool fatal = false;
    bool is_converge = opt_lsv.damping_iter(xs, voxhess, &hess, resis, up, false, &fatal);
    out.n_voxels.push_back((int)voxhess.plvec_voxels.size());
    if (fatal) { out.fatal = true; break; }
    out.resis_first.push_back(resis[0]);
    out.resis_last.push_back(resis[1]);

    if ((std::fabs(resis[0] - resis[1]) / resis[0] < converge_thre && is_converge) || (iterCnt == max_iter - 2 && converge_flag == 0)) {
      converge_thre = 0.01;
      if (converge_flag == 0) converge_flag = 1;
      else if (converge_flag == 1) break;
    }
  }

  out.xs = xs;
  out.hess = hess;
  if (out.fatal) return;
  for (int i = 0; i < wdsize - 1; i++)
    for (int j = i + 1; j < wdsize; j++) {
      bool isAdd = true;
      V6 v6;
      for (int k = 0; k < 6; k++) {
        double hc = std::fabs(hess(6 * i + k, 6 * j + k));
        if (hc < 1e-6) { isAdd = false; break; }
        v6[k] = 1.0 / hc;
      }
      if (isAdd) {
        PgoEdge e;
        e.m1 = smps[i].mp; e.m2 = smps[j].mp; e.id1 = smps[i].id; e.id2 = smps[j].id;
        e.tra = xs[i].R.transpose() * (xs[j].p - xs[i].p);
        e.rot = xs[i].R.transpose() * xs[j].R;
        e.cov = v6;
        out.edges.push_back(e);
      }
    }

  if (want_cloud) {
    std::vector<PointXYZIN> pl;
    const IMUST &xc = xs[0];
    for (int i = 0; i < wdsize; i++) {
      V3 dp = xc.R.transpose() * (xs[i].p - xc.p);
      M3 dR = xc.R.transpose() * xs[i].R;
      for (int n = 0; n < smps[i].n; n++) {
        PointXYZIN ap;
        std::memset(&ap, 0, sizeof(ap));
        V3 v3 = vec3(smps[i].xyz[3 * n], smps[i].xyz[3 * n + 1], smps[i].xyz[3 * n + 2]);
        v3 = dR * v3 + dp;
        ap.x = (float)v3[0]; ap.y = (float)v3[1]; ap.z = (float)v3[2];
        ap.intensity = (float)smps[i].mp;
        pl.push_back(ap);
      }
    }
    down_sampling_voxel(pl, base.voxel_size / 8);
    out.merged = pl;
  }
}

// voxelslam.cpp:3033-3034, :3066-3077, :3137-3138 — bottom-level window schedule for one session
// of n keyframes: windows of wdsize keyframes, stride mgsize.  Returns the first index of each.
inline std::vector<int> gba_window_starts(int n_keyframes, int wdsize = 10, int mgsize = 5) {
  std::vector<int> starts;
  for (int s = 0; s + wdsize <= n_keyframes; s += mgsize) starts.push_back(s);
  return starts;
}

}  // namespace orc
