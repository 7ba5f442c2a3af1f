// ORACLE — TEST INFRASTRUCTURE ONLY (see vmath.hpp header). PARITY UNPINNED.
//
// ba.hpp — restatement of the LiDAR BA factor and the LiDAR-only LM optimizer:
//   /root/reference/VoxelSLAM/src/voxel_map.hpp:124-339  class LidarFactor  (north_star: VOX_HESS)
//   /root/reference/VoxelSLAM/src/voxel_map.hpp:342-498  class Lidar_BA_Optimizer
// Thread fan-out is restated as "evaluate each range into its own private Hess/JacT, then sum
// in range order" which is arithmetically what the reference's join loop does
// (voxel_map.hpp:376-386); `use_threads` only decides whether the ranges run concurrently.
#pragma once
#include <atomic>
#include "types.hpp"
#include <thread>
#include <cstdio>
#include <cstdlib>

namespace orc {

// bench bookkeeping: voxel-frames that went through acc_evaluate2 since the last reset (orc_vf_counter)
inline std::atomic<int64_t> g_voxel_frames_evaluated{0};

class LidarFactor {
public:
  std::vector<PointCluster> sig_vecs;
  std::vector<std::vector<PointCluster>> plvec_voxels;
  std::vector<double> coeffs;
  std::vector<V3> eig_values;
  std::vector<M3> eig_vectors;
  std::vector<PointCluster> pcr_adds;
  int win_size;

  explicit LidarFactor(int w) : win_size(w) {}

  // voxel_map.hpp:139-147
  void push_voxel(const std::vector<PointCluster> &vec_orig, const PointCluster &fix, double coe,
                  const V3 &eig_value, const M3 &eig_vector, const PointCluster &pcr_add) {
    plvec_voxels.push_back(vec_orig);
    sig_vecs.push_back(fix);
    coeffs.push_back(coe);
    eig_values.push_back(eig_value);
    eig_vectors.push_back(eig_vector);
    pcr_adds.push_back(pcr_add);
  }

  // voxel_map.hpp:150-282
  void acc_evaluate2(const std::vector<IMUST> &xs, int head, int end, MatX &Hess, VecX &JacT, double &residual) const {
    Hess.setZero(); JacT.setZero(); residual = 0;
    const int kk = 0;
    std::vector<V3> viRiTuk(win_size);
    std::vector<M3> viRiTukukT(win_size);
    std::vector<M36> Auk(win_size);
    M3 umumT;
    const M3 I33 = M3::Identity();

    for (int a = head; a < end; a++) {
      const std::vector<PointCluster> &sig_orig = plvec_voxels[a];
      double coe = coeffs[a];
      {   // bench bookkeeping only (no reference counterpart): voxel-frames this call goes through, for the CPU arm's metric
        int64_t k = 0;
        for (const PointCluster &pc : sig_orig) k += pc.N != 0;
        g_voxel_frames_evaluated.fetch_add(k, std::memory_order_relaxed);
      }

      V3 lmbd = eig_values[a];
      M3 U = eig_vectors[a];
      int NN = pcr_adds[a].N;
      V3 vBar = pcr_adds[a].v / (double)NN;

      V3 u[3] = {U.col(0), U.col(1), U.col(2)};
      const V3 &uk = u[kk];
      M3 ukukT = uk * uk.transpose();
      umumT.setZero();
      for (int i = 0; i < 3; i++)
        if (i != kk) umumT += 2.0 / (lmbd[kk] - lmbd[i]) * (u[i] * u[i].transpose());

      for (int i = 0; i < win_size; i++)
        if (sig_orig[i].N != 0) {
          M3 Pi = sig_orig[i].P;
          V3 vi = sig_orig[i].v;
          M3 Ri = xs[i].R;
          double ni = sig_orig[i].N;

          M3 vihat = hat(vi);
          V3 RiTuk = Ri.transpose() * uk;
          M3 RiTukhat = hat(RiTuk);

          V3 PiRiTuk = Pi * RiTuk;
          viRiTuk[i] = vihat * RiTuk;
          viRiTukukT[i] = viRiTuk[i] * uk.transpose();

          V3 ti_v = xs[i].p - vBar;
          double ukTti_v = uk.dot(ti_v);

          M3 combo1 = hat(PiRiTuk) + vihat * ukTti_v;
          V3 combo2 = Ri * vi + ni * ti_v;
          M3 left = (Ri * Pi + ti_v * vi.transpose()) * RiTukhat - Ri * combo1;
          M3 right = combo2 * uk.transpose() + combo2.dot(uk) * I33;
          Auk[i].setBlock(0, 0, left);
          Auk[i].setBlock(0, 3, right);
          Auk[i] /= (double)NN;

          V6 jjt = Auk[i].transpose() * uk;
          for (int r = 0; r < 6; r++) JacT[6 * i + r] += coe * jjt[r];

          M3 HRt = 2.0 / NN * (1.0 - ni / NN) * viRiTukukT[i];
          M6 Hb = Auk[i].transpose() * umumT * Auk[i];
          M3 rr = 2.0 / NN * ((combo1 - RiTukhat * Pi) * RiTukhat)
                  - 2.0 / NN / NN * (viRiTuk[i] * viRiTuk[i].transpose())
                  - 0.5 * hat(jjt.block<3, 1>(0, 0));
          Hb.addBlock(0, 0, rr);
          Hb.addBlock(0, 3, HRt);
          Hb.addBlock(3, 0, HRt.transpose());
          Hb.addBlock(3, 3, 2.0 / NN * (ni - ni * ni / NN) * ukukT);

          Hess.addBlock(6 * i, 6 * i, coe * Hb);
        }

      for (int i = 0; i < win_size - 1; i++)
        if (sig_orig[i].N != 0) {
          double ni = sig_orig[i].N;
          for (int j = i + 1; j < win_size; j++)
            if (sig_orig[j].N != 0) {
              double nj = sig_orig[j].N;
              M6 Hb = Auk[i].transpose() * umumT * Auk[j];
              Hb.addBlock(0, 0, -2.0 / NN / NN * (viRiTuk[i] * viRiTuk[j].transpose()));
              Hb.addBlock(0, 3, -2.0 * nj / NN / NN * viRiTukukT[i]);
              Hb.addBlock(3, 0, -2.0 * ni / NN / NN * viRiTukukT[j].transpose());
              Hb.addBlock(3, 3, -2.0 * ni * nj / NN / NN * ukukT);
              Hess.addBlock(6 * i, 6 * j, coe * Hb);
            }
        }

      residual += coe * lmbd[kk];
    }

    for (int i = 1; i < win_size; i++)
      for (int j = 0; j < i; j++)
        for (int r = 0; r < 6; r++)
          for (int c = 0; c < 6; c++)
            Hess(6 * i + r, 6 * j + c) = Hess(6 * j + c, 6 * i + r);
  }

  // voxel_map.hpp:285-325 — NOTE the write-back of eig_values/eig_vectors/pcr_adds (:317-319)
  void evaluate_only_residual(const std::vector<IMUST> &xs, int head, int end, double &residual) {
    residual = 0;
    const int kk = 0;
    PointCluster pcr;
    for (int a = head; a < end; a++) {
      const std::vector<PointCluster> &sig_orig = plvec_voxels[a];
      PointCluster sig = sig_vecs[a];
      for (int i = 0; i < win_size; i++)
        if (sig_orig[i].N != 0) {
          pcr.transform(sig_orig[i], xs[i]);
          sig += pcr;
        }
      V3 vBar = sig.v / (double)sig.N;
      M3 cm = sig.P / (double)sig.N - vBar * vBar.transpose();
      V3 lmbd; M3 U;
      eig3_sym(cm, lmbd, U);
      eig_values[a] = lmbd;
      eig_vectors[a] = U;
      pcr_adds[a] = sig;
      residual += coeffs[a] * lmbd[kk];
    }
  }

  void clear() {  // voxel_map.hpp:328-336
    sig_vecs.clear(); plvec_voxels.clear(); eig_values.clear();
    eig_vectors.clear(); pcr_adds.clear(); coeffs.clear();
  }
};

using VOX_HESS = LidarFactor;  // the BALM2/HBA name used by BASELINE.json:north_star

// Runs f(i) for i in [0,n): i==0 on the caller, the rest on worker threads (or serially).
template <class F>
inline void fan_out(int n, bool use_threads, F f) {
  if (!use_threads || n <= 1) { for (int i = 0; i < n; i++) f(i); return; }
  std::vector<std::thread> th;
  for (int i = 1; i < n; i++) th.emplace_back(f, i);
  f(0);
  for (auto &t : th) t.join();
}

class Lidar_BA_Optimizer {
public:
  int win_size = 0, jac_leng = 0, thd_num = 2;
  bool use_threads = false;
  // trace of the LM accept/reject sequence and damping values (test introspection only)
  std::vector<int> accept_trace;
  std::vector<double> u_trace;

  // voxel_map.hpp:347-389
  double divide_thread(std::vector<IMUST> &x_stats, LidarFactor &voxhess, MatX &Hess, VecX &JacT) {
    double residual = 0;
    Hess.setZero(); JacT.setZero();
    int tthd_num = thd_num;
    int g_size = (int)voxhess.plvec_voxels.size();
    if (g_size < tthd_num) tthd_num = 1;
    std::vector<MatX> hessians(tthd_num);
    std::vector<VecX> jacobins(tthd_num);
    std::vector<double> resis(tthd_num, 0.0);
    for (int i = 0; i < tthd_num; i++) { hessians[i].resize(jac_leng, jac_leng); jacobins[i].resize(jac_leng); }
    double part = 1.0 * g_size / tthd_num;
    fan_out(tthd_num, use_threads, [&](int i) {
      int head = (i == 0) ? 0 : (int)(part * i);
      int end = (int)(part * (i + 1));
      voxhess.acc_evaluate2(x_stats, head, end, hessians[i], jacobins[i], resis[i]);
    });
    for (int i = 0; i < tthd_num; i++) { Hess += hessians[i]; JacT += jacobins[i]; residual += resis[i]; }
    return residual;
  }

  // voxel_map.hpp:391-420 — returns false (reference: printf + exit(0)) when V < thd_num
  bool only_residual(std::vector<IMUST> &x_stats, LidarFactor &voxhess, double &residual1) {
    residual1 = 0;
    int g_size = (int)voxhess.plvec_voxels.size();
    if (g_size < thd_num) return false;
    std::vector<double> residuals(thd_num, 0.0);
    double part = 1.0 * g_size / thd_num;
    fan_out(thd_num, use_threads, [&](int i) {
      voxhess.evaluate_only_residual(x_stats, (int)(part * i), (int)(part * (i + 1)), residuals[i]);
    });
    for (int i = 0; i < thd_num; i++) residual1 += residuals[i];
    return true;
  }

  // voxel_map.hpp:422-497.  Returns is_converge; `fatal` is set where the reference would exit(0).
  bool damping_iter(std::vector<IMUST> &x_stats, LidarFactor &voxhess, MatX *hess, std::vector<double> &resis,
                    int max_iter = 3, bool is_display = false, bool *fatal = nullptr) {
    win_size = voxhess.win_size;
    jac_leng = win_size * 6;
    double u = 0.01, v = 2;
    MatX Hess(jac_leng, jac_leng);
    VecX JacT(jac_leng), dxi(jac_leng), Ddiag(jac_leng);
    hess->resize(jac_leng, jac_leng);
    double residual1 = 0, residual2 = 0, q;
    bool is_calc_hess = true;
    std::vector<IMUST> x_stats_temp = x_stats;
    bool is_converge = true;
    if (fatal) *fatal = false;

    for (int i = 0; i < max_iter; i++) {
      if (is_calc_hess) {
        residual1 = divide_thread(x_stats, voxhess, Hess, JacT);
        *hess = Hess;
      }
      if (i == 0) resis.push_back(residual1);

      for (int c = 0; c < jac_leng; c++) for (int r = 0; r < 6; r++) Hess(r, c) = 0.0;
      for (int c = 0; c < 6; c++) for (int r = 0; r < jac_leng; r++) Hess(r, c) = 0.0;
      for (int r = 0; r < 6; r++) Hess(r, r) = 1.0;
      for (int r = 0; r < 6; r++) JacT[r] = 0.0;

      for (int r = 0; r < jac_leng; r++) Ddiag[r] = Hess(r, r);
      MatX A = Hess;
      for (int r = 0; r < jac_leng; r++) A(r, r) += u * Ddiag[r];
      VecX nb(jac_leng);
      for (int r = 0; r < jac_leng; r++) nb[r] = -JacT[r];
      ldlt_solve(A, nb, dxi);

      for (int j = 0; j < win_size; j++) {
        x_stats_temp[j].R = x_stats[j].R * Exp(dxi.seg<3>(6 * j));
        x_stats_temp[j].p = x_stats[j].p + dxi.seg<3>(6 * j + 3);
      }
      double q1 = 0;
      for (int r = 0; r < jac_leng; r++) q1 += dxi[r] * (u * Ddiag[r] * dxi[r] - JacT[r]);
      q1 *= 0.5;

      if (!only_residual(x_stats_temp, voxhess, residual2)) {
        if (fatal) { *fatal = true; return false; }
        printf("Too Less Voxel"); exit(0);
      }

      q = (residual1 - residual2);
      if (is_display)
        printf("iter%d: (%lf %lf) u: %lf v: %.1lf q: %.2lf %lf %lf\n", i, residual1, residual2, u, v, q / q1, q1, q);

      if (q > 0) {
        x_stats = x_stats_temp;
        double one_three = 1.0 / 3;
        q = q / q1;
        v = 2;
        q = 1 - std::pow(2 * q - 1, 3);
        u *= (q < one_three ? one_three : q);
        is_calc_hess = true;
        accept_trace.push_back(1);
      } else {
        u = u * v;
        v = 2 * v;
        is_calc_hess = false;
        is_converge = false;
        accept_trace.push_back(0);
      }
      u_trace.push_back(u);

      if (std::fabs((residual1 - residual2) / residual1) < 1e-6) break;
    }
    resis.push_back(residual2);
    return is_converge;
  }
};

}  // namespace orc
