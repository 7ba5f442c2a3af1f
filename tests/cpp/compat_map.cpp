// compat_map.cpp — the reference's own call lines of the local-mapping loop (voxelslam.cpp:1905-1991) and of HBA_add_edge's body
// (voxelslam.cpp:2887-2899), compiled against include/voxelba_compat.hpp (stand-alone type mirrors) and run on the GPU.
// tests/test_compat_map_gpu.py feeds it the oracle session's inputs and compares what comes back with the oracle's results.
//   usage: compat_map <in.bin> <out.bin>
#define VOXELBA_COMPAT_STANDALONE
#include "voxelba_compat.hpp"

#include <cmath>
#include <cstdint>
#include <fstream>
#include <iostream>

using namespace std;

// the reference's globals (voxel_map.hpp:98-104, :500; loop_refine.hpp:269-271), set by the test input
double imu_coef = 1e-4;
double voxel_size = 1.0, min_eigen_value = 0.0025;
int max_layer = 2, max_points = 100;
Eigen::FixedMat<4, 1> min_point;
vector<double> plane_eigen_value_thre;
double gba_voxel_size = 1.0, gba_min_eigen_value = 0.01;
vector<double> gba_eigen_value_array;

template <class T> static void rd(ifstream &f, T *p, size_t n) { f.read(reinterpret_cast<char *>(p), sizeof(T) * n); }
template <class T> static void wr(ofstream &f, const T *p, size_t n) { f.write(reinterpret_cast<const char *>(p), sizeof(T) * n); }

static IMUST state(const vx_imu_state &s) {
  IMUST x;
  memcpy(x.R.data(), s.R, 72); memcpy(x.p.data(), s.p, 24); memcpy(x.v.data(), s.v, 24); memcpy(x.bg.data(), s.bg, 24); memcpy(x.ba.data(), s.ba, 24); memcpy(x.g.data(), s.g, 24);
  return x;
}
static IMU_PRE *imu_pre(const vx_imu_pre &p) {
  IMU_PRE *q = new IMU_PRE();
  memcpy(q->R_delta.data(), p.R_delta, 72); memcpy(q->p_delta.data(), p.p_delta, 24); memcpy(q->v_delta.data(), p.v_delta, 24);
  memcpy(q->bg.data(), p.bg, 24); memcpy(q->ba.data(), p.ba, 24);
  memcpy(q->R_bg.data(), p.R_bg, 72); memcpy(q->p_bg.data(), p.p_bg, 72); memcpy(q->p_ba.data(), p.p_ba, 72); memcpy(q->v_bg.data(), p.v_bg, 72); memcpy(q->v_ba.data(), p.v_ba, 72);
  q->dtime = p.dtime;
  memcpy(q->dbg.data(), p.dbg, 24); memcpy(q->dba.data(), p.dba, 24); memcpy(q->dbg_buf.data(), p.dbg_buf, 24); memcpy(q->dba_buf.data(), p.dba_buf, 24);
  memcpy(q->cov.data(), p.cov, sizeof(p.cov));
  return q;
}

int main(int argc, char **argv) {
  if (argc < 3) { cerr << "usage: compat_map in.bin out.bin\n"; return 2; }
  ifstream in(argv[1], ios::binary);
  ofstream out(argv[2], ios::binary);
  int32_t win_size = 0, thread_num = 0;
  rd(in, &win_size, 1); rd(in, &thread_num, 1);
  double thre[4], gthre[4];
  rd(in, &voxel_size, 1); rd(in, &min_eigen_value, 1); rd(in, thre, 4); rd(in, &gba_voxel_size, 1); rd(in, &gba_min_eigen_value, 1); rd(in, gthre, 4);
  plane_eigen_value_thre.assign(thre, thre + 4); gba_eigen_value_array.assign(gthre, gthre + 4);
  for (int k = 0; k < 4; k++) min_point[k] = 5;                                   // voxelslam.cpp:917

  // ================= part A: the local-mapping loop, voxelslam.cpp:1905-1991 =================
  vector<vx_imu_state> st(win_size); vector<vx_imu_pre> pres(win_size - 1);
  rd(in, st.data(), win_size); rd(in, pres.data(), win_size - 1);
  unordered_map<VOXEL_LOC, OctoTree *> surf_map, surf_map_slide;                 // voxelslam.hpp
  vector<vector<SlideWindow *>> sws(thread_num);
  LidarFactor voxhess(win_size);
  vector<IMUST> x_buf; vector<PVecPtr> pvec_buf; deque<IMU_PRE *> imu_pre_buf;
  Eigen::MatrixXd hess;
  int win_count = 0;
  double jour = 0;
  for (int k = 0; k < win_size; k++) {
    int32_t n = 0; rd(in, &n, 1);
    PVecPtr pptr(new PVec(n)); PLV(3) pwld(n);
    rd(in, reinterpret_cast<double *>(pptr->data()), (size_t)12 * n); rd(in, reinterpret_cast<double *>(pwld.data()), (size_t)3 * n);
    IMUST x_curr = state(st[k]);
    // ---- voxelslam.cpp:1905-1919, verbatim (keyframe_loading, the timers and the degrade branch are not on the path)
    win_count++;
    x_buf.push_back(x_curr);
    pvec_buf.push_back(pptr);
    if (win_count > 1)
    {
      imu_pre_buf.push_back(imu_pre(pres[win_count - 2]));                        // new IMU_PRE(...)->push_imu(imus): the factor is a test input
    }
    voxhess.clear();
    voxhess.win_size = win_size;

    cut_voxel_multi(surf_map, pvec_buf[win_count - 1], win_count - 1, surf_map_slide, win_size, pwld, sws);

    multi_recut(surf_map_slide, win_count, x_buf, voxhess, sws);
  }
  const double n_vox = (double)voxhess.num_voxels();
  wr(out, &n_vox, 1);
  if (win_count >= win_size)
  {
    // ---- voxelslam.cpp:1956-1961, verbatim
    LI_BA_OptimizerGravity opt_lsv;
    vector<double> resis;
    opt_lsv.damping_iter(x_buf, voxhess, imu_pre_buf, resis, &hess, 5);
    for (int i = 0; i < win_size; i++) { vx_imu_state s = vxc::state_of(x_buf[i]); wr(out, &s, 1); }
    wr(out, resis.data(), 2);
    double v6[6];
    for (int i = 0; i < 6; i++) v6[i] = 1.0 / fabs(hess(i, 15 + i));             // bl->v6 = hess.block<6, 6>(0, DIM).diagonal(), :1971-1973
    wr(out, v6, 6);

    // ---- voxelslam.cpp:1991, verbatim
    multi_margi(surf_map_slide, jour, win_count, x_buf, voxhess, sws[0]);
    vxc::shift_window(surf_map, 1);                                               // the `mp` rotation of :2014-2019
  }
  {   // match() on a world point of the last scan and lio_state_estimation of a fresh sweep against the map (voxelslam.cpp:1027, :962)
    int32_t n = 0; rd(in, &n, 1);
    PVecPtr pptr(new PVec(n));
    rd(in, reinterpret_cast<double *>(pptr->data()), (size_t)12 * n);
    vx_imu_state s0; rd(in, &s0, 1);
    IMUST x_curr = state(s0);
    rd(in, x_curr.cov.data(), 225);
    int matched = 0; double sig_sum = 0;
    for (int i = 0; i < n; i += 50) {
      Eigen::Vector3d wld; Eigen::Matrix3d var_wld = (*pptr)[i].var;
      for (int r = 0; r < 3; r++) wld[r] = x_curr.R(r, 0) * (*pptr)[i].pnt[0] + x_curr.R(r, 1) * (*pptr)[i].pnt[1] + x_curr.R(r, 2) * (*pptr)[i].pnt[2] + x_curr.p[r];
      Plane *pla = nullptr; double sigma_d = 0; OctoTree *oc = nullptr;
      int flag = match(surf_map, wld, pla, var_wld, sigma_d, oc);                 // voxelslam.cpp:1027
      if (flag) { matched++; sig_sum += sigma_d; if (!pla || !oc || !pla->is_plane) return 5; }
    }
    const bool ok = lio_state_estimation(surf_map, pptr, x_curr);
    vx_imu_state s = vxc::state_of(x_curr);
    const double m = matched, o = ok ? 1.0 : 0.0;
    wr(out, &m, 1); wr(out, &sig_sum, 1); wr(out, &o, 1); wr(out, &s, 1);
  }
  vxc::release(surf_map); vxc::release(surf_map_slide);

  // ================= part B: HBA_add_edge's body, voxelslam.cpp:2887-2899 =================
  int32_t wdsize = 0; rd(in, &wdsize, 1);
  vector<IMUST> xs(wdsize); vector<Keyframe *> smps(wdsize);
  for (int i = 0; i < wdsize; i++) {
    vx_pose p; rd(in, &p, 1);
    memcpy(xs[i].R.data(), p.R, 72); memcpy(xs[i].p.data(), p.p, 24);
    smps[i] = new Keyframe(xs[i]); smps[i]->id = 100 + i; smps[i]->mp = 0;
    int32_t n = 0; rd(in, &n, 1);
    vector<float> c((size_t)3 * n); rd(in, c.data(), c.size());
    for (int k = 0; k < n; k++) { PointType ap = {}; ap.x = c[3 * k]; ap.y = c[3 * k + 1]; ap.z = c[3 * k + 2]; smps[i]->plptr->push_back(ap); }
  }
  if (!in) { cerr << "short input\n"; return 2; }
  vector<IMUST> p_xs = xs;
  {
    int up = 4; bool is_display = false; thread_num = 2;
    // ---- voxelslam.cpp:2887-2899, verbatim
    unordered_map<VOXEL_LOC, OctreeGBA *> oct_map;
    for (int i = 0; i < wdsize; i++)
      OctreeGBA::cut_voxel(oct_map, xs[i], smps[i]->plptr, i, wdsize);

    LidarFactor voxhess(wdsize);
    OctreeGBA_multi_recut(oct_map, voxhess, thread_num);

    Lidar_BA_Optimizer opt_lsv;
    opt_lsv.thd_num = thread_num;
    vector<double> resis;
    bool is_converge = opt_lsv.damping_iter(xs, voxhess, &hess, resis, up, is_display);
    // ----
    const double nv = (double)voxhess.num_voxels(), c = is_converge ? 1.0 : 0.0;
    wr(out, &nv, 1); wr(out, resis.data(), 2); wr(out, &c, 1);
    for (int i = 0; i < wdsize; i++) { vx_pose p = vxc::pose_of(xs[i]); wr(out, &p, 1); }
  }
  {   // the whole HBA_add_edge as thd_globalmapping calls it (voxelslam.cpp:3096): max_iter 1, thread_num 2, submap cloud out
    PGO_Edges gba_edges1; vector<int> mps{0};
    pcl::PointCloud<PointType>::Ptr plptr(new pcl::PointCloud<PointType>());
    VOX_HESS alias_check(wdsize); (void)alias_check;
    HBA_add_edge(p_xs, smps, gba_edges1, mps, 1, 2, plptr);
    const double ne = (double)gba_edges1.records.size(), nc = (double)plptr->size();
    wr(out, &ne, 1); wr(out, &nc, 1);
    for (auto &e : gba_edges1.records) { const double ids[2] = {(double)e.id1, (double)e.id2}; wr(out, ids, 2); wr(out, e.tra.data(), 3); wr(out, e.cov.data(), 6); }
    for (int i = 0; i < wdsize; i++) { vx_pose p = vxc::pose_of(p_xs[i]); wr(out, &p, 1); }
  }
  return out ? 0 : 4;
}
