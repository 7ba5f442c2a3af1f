// compat_smoke.cpp — reference-style call sites compiled against include/voxelba_compat.hpp (stand-alone type mirrors).
// Reads a raw dump of the tests/golden/factor.npz problem, drives it exactly the way voxelslam.cpp drives the reference
// classes (push_voxel -> acc_evaluate2 / evaluate_only_residual -> Lidar_BA_Optimizer::damping_iter) and writes the results
// for tests/test_compat_gpu.py to compare with the committed fixture.
//   usage: compat_smoke <in.bin> <out.bin>
#define VOXELBA_COMPAT_STANDALONE
#include "voxelba_compat.hpp"

#include <cstdint>
#include <fstream>
#include <iostream>

double imu_coef = 1e-4;   // voxel_map.hpp:500

template <class T> static void rd(std::ifstream &f, T *p, size_t n) { f.read(reinterpret_cast<char *>(p), sizeof(T) * n); }
template <class T> static void wr(std::ofstream &f, const T *p, size_t n) { f.write(reinterpret_cast<const char *>(p), sizeof(T) * n); }

static PointCluster cluster(const vx_cluster &c) { PointCluster s; vxc::from_c(c, s); return s; }

int main(int argc, char **argv) {
  if (argc < 3) { std::cerr << "usage: compat_smoke in.bin out.bin\n"; return 2; }
  std::ifstream in(argv[1], std::ios::binary);
  int32_t V = 0, W = 0;
  rd(in, &V, 1); rd(in, &W, 1);
  std::vector<vx_cluster> plvec((size_t)V * W), sig(V), merged(V);
  std::vector<double> coe(V), eigval(3 * (size_t)V), eigvec(9 * (size_t)V);
  std::vector<vx_pose> poses(W), start(W);
  rd(in, plvec.data(), plvec.size()); rd(in, sig.data(), V); rd(in, coe.data(), V); rd(in, eigval.data(), eigval.size());
  rd(in, eigvec.data(), eigvec.size()); rd(in, merged.data(), V); rd(in, poses.data(), W); rd(in, start.data(), W);
  if (!in) { std::cerr << "short input\n"; return 2; }

  auto fill = [&](LidarFactor &voxhess) {
    for (int a = 0; a < V; a++) {
      std::vector<PointCluster> pcrs(W);
      for (int i = 0; i < W; i++) pcrs[i] = cluster(plvec[(size_t)a * W + i]);
      PointCluster fix = cluster(sig[a]), add = cluster(merged[a]);
      Eigen::Vector3d ev; Eigen::Matrix3d evec;
      std::memcpy(ev.data(), &eigval[3 * a], 24); std::memcpy(evec.data(), &eigvec[9 * a], 72);
      voxhess.push_voxel(pcrs, fix, coe[a], ev, evec, add);        // tras_opt -> push_voxel, voxel_map.hpp:1626-1627
    }
  };
  auto states = [&](const std::vector<vx_pose> &ps) {
    std::vector<IMUST> xs(W);
    for (int i = 0; i < W; i++) { std::memcpy(xs[i].R.data(), ps[i].R, 72); std::memcpy(xs[i].p.data(), ps[i].p, 24); }
    return xs;
  };

  std::ofstream out(argv[2], std::ios::binary);
  {
    LidarFactor voxhess(W);
    fill(voxhess);
    std::vector<IMUST> xs = states(poses);
    double residual = 0;
    voxhess.evaluate_only_residual(xs, 0, V, residual);             // voxel_map.hpp:411
    Eigen::MatrixXd Hess; Eigen::VectorXd JacT;
    Hess.resize(6 * W, 6 * W); JacT.resize(6 * W);
    double cost = 0;
    voxhess.acc_evaluate2(xs, 0, V, Hess, JacT, cost);              // voxel_map.hpp:376
    wr(out, &residual, 1); wr(out, &cost, 1); wr(out, Hess.data(), (size_t)36 * W * W); wr(out, JacT.data(), (size_t)6 * W);
    wr(out, voxhess.eig_values[0].data(), 3);                       // the cache came back into the public members (:317-319)
    voxhess.clear();
    if (!voxhess.plvec_voxels.empty()) return 3;
  }
  {
    LidarFactor voxhess(W);
    fill(voxhess);
    std::vector<IMUST> xs = states(start);
    Lidar_BA_Optimizer opt;
    opt.only_residual(xs, voxhess);
    Eigen::MatrixXd hess; std::vector<double> resis;
    const bool conv = opt.damping_iter(xs, voxhess, &hess, resis, 4);   // voxelslam.cpp:2895-2899 (up = 4)
    std::vector<vx_pose> ps(W);
    for (int i = 0; i < W; i++) ps[i] = vxc::pose_of(xs[i]);
    const double c = conv ? 1.0 : 0.0;
    wr(out, ps.data(), W); wr(out, resis.data(), 2); wr(out, &c, 1); wr(out, hess.data(), (size_t)36 * W * W);
  }
  return out ? 0 : 4;
}
