// ORACLE — TEST INFRASTRUCTURE ONLY (see vmath.hpp header). PARITY UNPINNED.
//
// li_ba.hpp — restatement of the IMU preintegration factor and the two LiDAR-inertial LM
// optimizers:
//   /root/reference/VoxelSLAM/src/preintegration.hpp:11-303  class IMU_PRE
//   /root/reference/VoxelSLAM/src/voxel_map.hpp:504-714      class LI_BA_Optimizer
//   /root/reference/VoxelSLAM/src/voxel_map.hpp:717-976      class LI_BA_OptimizerGravity
// ROS message plumbing (push_imu's sensor_msgs::Imu deque, preintegration.hpp:50-73) is replaced
// by a plain array of (t, gyr, acc) samples with the identical mid-point rule.
#pragma once
#include "ba.hpp"
#include <deque>

namespace orc {

struct ImuSample { double t; V3 gyr, acc; };

struct ImuNoise {                 // preintegration.hpp:8-9 globals
  double scale_gravity = 1.0;
  M6 noiseMeas, noiseWalk;
};

class IMU_PRE {
public:
  M3 R_delta; V3 p_delta, v_delta, bg, ba;
  M3 R_bg, p_bg, p_ba, v_bg, v_ba;
  double dtime = 0;
  V3 dbg, dba, dbg_buf, dba_buf;
  M15 cov;
  const ImuNoise *noise;

  IMU_PRE(const ImuNoise *nz, const V3 &bg1 = V3(), const V3 &ba1 = V3()) : noise(nz) {  // preintegration.hpp:32-48
    bg = bg1; ba = ba1;
    R_delta.setIdentity();
  }

  // preintegration.hpp:50-73
  void push_imu(const std::vector<ImuSample> &imus) {
    for (size_t k = 1; k < imus.size(); k++) {
      const ImuSample &i1 = imus[k - 1], &i2 = imus[k];
      double dt = i2.t - i1.t;
      V3 cur_gyr = 0.5 * (i1.gyr + i2.gyr);
      V3 cur_acc = 0.5 * (i1.acc + i2.acc);
      cur_gyr = cur_gyr - bg;
      cur_acc = cur_acc * noise->scale_gravity - ba;
      add_imu(cur_gyr, cur_acc, dt);
    }
  }

  // preintegration.hpp:75-135
  void add_imu(const V3 &cur_gyr, const V3 &cur_acc, double dt) {
    dtime += dt;
    M3 R_inc = Exp(cur_gyr, dt);
    M3 R_jr = jr(cur_gyr * dt);
    M3 R_dt = dt * R_delta;
    M3 R_dt2_2 = 0.5 * dt * dt * R_delta;
    M3 acc_skew = hat(cur_acc);

    p_ba = p_ba + v_ba * dt - R_dt2_2;
    p_bg = p_bg + v_bg * dt - R_dt2_2 * acc_skew * R_bg;
    v_ba = v_ba - R_dt;
    v_bg = v_bg - R_dt * acc_skew * R_bg;
    R_bg = R_inc.transpose() * R_bg - R_jr * dt;

    Mat<9, 9> A; Mat<9, 6> B;
    A.setIdentity();
    A.setBlock(0, 0, R_inc.transpose());
    A.setBlock(3, 0, -(R_dt2_2 * acc_skew));
    A.setBlock(3, 6, M3::Identity() * dt);
    A.setBlock(6, 0, -(R_dt * acc_skew));
    B.setBlock(0, 0, R_jr * dt);
    B.setBlock(3, 3, R_dt2_2);
    B.setBlock(6, 3, R_dt);

    Mat<9, 9> c99 = cov.block<9, 9>(0, 0);
    c99 = A * c99 * A.transpose() + B * noise->noiseMeas * B.transpose();
    cov.setBlock(0, 0, c99);
    Mat<6, 6> c66 = cov.block<6, 6>(9, 9);
    c66 += noise->noiseWalk * dt;
    cov.setBlock(9, 9, c66);

    p_delta += v_delta * dt + R_dt2_2 * cur_acc;
    v_delta += R_dt * cur_acc;
    R_delta = R_delta * R_inc;
  }

  // Shared body of give_evaluate (preintegration.hpp:137-212) and give_evaluate_g (:214-294).
  // jtj/gg are (30 or 33)-sized, column-major, caller-zeroed; written only if jac_enable.
  template <int NC>
  double evaluate(const IMUST &st1, const IMUST &st2, Mat<NC, NC> &jtj, Mat<NC, 1> &gg, bool jac_enable) const {
    static_assert(NC == 2 * DIM || NC == 2 * DIM + 3, "30 or 33 columns");
    Mat<DIM, DIM> joca, jocb;
    V15 rr;

    M3 R_correct = R_delta * Exp(R_bg * dbg);
    V3 t_correct = p_delta + p_bg * dbg + p_ba * dba;
    V3 v_correct = v_delta + v_bg * dbg + v_ba * dba;

    M3 res_r = R_correct.transpose() * st1.R.transpose() * st2.R;
    V3 exp_v = st1.R.transpose() * (st2.v - st1.v - dtime * st1.g);
    V3 res_v = exp_v - v_correct;
    V3 exp_t = st1.R.transpose() * (st2.p - st1.p - st1.v * dtime - 0.5 * dtime * dtime * st1.g);
    V3 res_t = exp_t - t_correct;
    V3 res_bg = st2.bg - st1.bg;
    V3 res_ba = st2.ba - st1.ba;
    double b_wei = 1;

    rr.setBlock(0, 0, Log(res_r));
    rr.setBlock(3, 0, res_t);
    rr.setBlock(6, 0, res_v);
    rr.setBlock(9, 0, res_bg * b_wei);
    rr.setBlock(12, 0, res_ba * b_wei);

    M15 cov_inv = inverse_lu(cov);

    if (jac_enable) {
      const M3 I33 = M3::Identity();
      M3 JR_inv = jr_inv(res_r);
      joca.setBlock(0, 0, -(JR_inv * st2.R.transpose() * st1.R));
      jocb.setBlock(0, 0, JR_inv);
      joca.setBlock(0, 9, -(JR_inv * res_r.transpose() * jr(R_bg * dbg) * R_bg));

      joca.setBlock(3, 0, hat(exp_t));
      joca.setBlock(3, 3, -st1.R.transpose());
      joca.setBlock(3, 6, -(st1.R.transpose() * dtime));
      joca.setBlock(3, 9, -p_bg);
      joca.setBlock(3, 12, -p_ba);
      jocb.setBlock(3, 3, st1.R.transpose());

      joca.setBlock(6, 0, hat(exp_v));
      joca.setBlock(6, 6, -st1.R.transpose());
      joca.setBlock(6, 9, -v_bg);
      joca.setBlock(6, 12, -v_ba);
      jocb.setBlock(6, 6, st1.R.transpose());

      joca.setBlock(9, 9, -(I33 * b_wei));
      joca.setBlock(12, 12, -(I33 * b_wei));
      jocb.setBlock(9, 9, I33 * b_wei);
      jocb.setBlock(12, 12, I33 * b_wei);

      Mat<DIM, NC> joc;
      joc.setBlock(0, 0, joca);
      joc.setBlock(0, DIM, jocb);
      if (NC == 2 * DIM + 3) {
        Mat<DIM, 3> jocg;
        jocg.setBlock(3, 0, st1.R.transpose() * (-0.5 * dtime * dtime));
        jocg.setBlock(6, 0, st1.R.transpose() * (-dtime));
        for (int r = 0; r < DIM; r++) for (int c = 0; c < 3; c++) joc(r, 2 * DIM + c) = jocg(r, c);
      }
      Mat<NC, DIM> jt_ci = joc.transpose() * cov_inv;
      jtj = jt_ci * joc;
      gg = jt_ci * rr;
    }
    return rr.dot(cov_inv * rr);
  }

  void update_state(const V15 &dxi) {  // preintegration.hpp:296-303
    dbg_buf = dbg; dba_buf = dba;
    dbg += dxi.block<3, 1>(9, 0);
    dba += dxi.block<3, 1>(12, 0);
  }
};

// ---------------------------------------------------------------------------------------
// Common LM driver for LI_BA_Optimizer (gravity=false) and LI_BA_OptimizerGravity (true).
// The two reference classes differ only in: unknown count (15W vs 15W+3), factor evaluate
// (give_evaluate vs give_evaluate_g + gravity scatter, voxel_map.hpp:785-795), gauge (15 vs 6
// fixed, voxel_map.hpp:653-656 vs :906-909), iteration count (3 fixed vs max_iter), resis
// recording (Gravity only) and the g retraction (voxel_map.hpp:921-930).
// ---------------------------------------------------------------------------------------
class LI_BA_Common {
public:
  int win_size = 0, jac_leng = 0, imu_leng = 0;
  bool use_threads = false;
  double imu_coef = 1e-4;
  std::vector<int> accept_trace;

  // voxel_map.hpp:509-517 / :726-738
  void hess_plus(MatX &Hess, VecX &JacT, const MatX &hs, const VecX &js) {
    for (int i = 0; i < win_size; i++) {
      for (int r = 0; r < DVEL; r++) JacT[i * DIM + r] += js[i * DVEL + r];
      for (int j = 0; j < win_size; j++)
        for (int c = 0; c < DVEL; c++) for (int r = 0; r < DVEL; r++)
          Hess(i * DIM + r, j * DIM + c) += hs(i * DVEL + r, j * DVEL + c);
    }
  }

  // voxel_map.hpp:519-584 / :746-825
  double divide_thread(bool gravity, std::vector<IMUST> &x_stats, LidarFactor &voxhess, std::deque<IMU_PRE *> &imus_factor, MatX &Hess, VecX &JacT) {
    int thd_num = 5;
    double residual = 0;
    Hess.setZero(); JacT.setZero();
    int tthd_num = thd_num;
    int g_size = (int)voxhess.plvec_voxels.size();
    if (g_size < tthd_num) tthd_num = 1;
    double part = 1.0 * g_size / tthd_num;
    std::vector<MatX> hessians(tthd_num);
    std::vector<VecX> jacobins(tthd_num);
    std::vector<double> resis(tthd_num, 0.0);
    for (int i = 0; i < tthd_num; i++) { hessians[i].resize(jac_leng, jac_leng); jacobins[i].resize(jac_leng); }

    // the caller thread evaluates the IMU chain while workers run (same arithmetic either way)
    auto imu_part = [&]() {
      for (int i = 0; i < win_size - 1; i++) {
        if (!gravity) {
          Mat<30, 30> jtj; Mat<30, 1> gg;
          residual += imus_factor[i]->evaluate<30>(x_stats[i], x_stats[i + 1], jtj, gg, true);
          Hess.addBlock(i * DIM, i * DIM, jtj);
          for (int r = 0; r < 30; r++) JacT[i * DIM + r] += gg[r];
        } else {
          Mat<33, 33> jtj; Mat<33, 1> gg;
          residual += imus_factor[i]->evaluate<33>(x_stats[i], x_stats[i + 1], jtj, gg, true);
          Hess.addBlock(i * DIM, i * DIM, jtj.block<30, 30>(0, 0));
          Hess.addBlock(i * DIM, imu_leng - 3, jtj.block<30, 3>(0, 30));
          Hess.addBlock(imu_leng - 3, i * DIM, jtj.block<3, 30>(30, 0));
          Hess.addBlock(imu_leng - 3, imu_leng - 3, jtj.block<3, 3>(30, 30));
          for (int r = 0; r < 30; r++) JacT[i * DIM + r] += gg[r];
          for (int r = 0; r < 3; r++) JacT[imu_leng - 3 + r] += gg[30 + r];
        }
      }
      Hess *= imu_coef;
      JacT *= imu_coef;
      residual *= (imu_coef * 0.5);
    };
    auto lidar_part = [&](int i) {
      int head = (i == 0) ? 0 : (int)(part * i);
      int end = (int)(part * (i + 1));
      voxhess.acc_evaluate2(x_stats, head, end, hessians[i], jacobins[i], resis[i]);
    };
    if (use_threads && tthd_num > 1) {
      std::vector<std::thread> th;
      for (int i = 1; i < tthd_num; i++) th.emplace_back(lidar_part, i);
      imu_part();
      lidar_part(0);
      for (auto &t : th) t.join();
    } else {
      imu_part();
      for (int i = 0; i < tthd_num; i++) lidar_part(i);
    }
    for (int i = 0; i < tthd_num; i++) { hess_plus(Hess, JacT, hessians[i], jacobins[i]); residual += resis[i]; }
    return residual;
  }

  // voxel_map.hpp:586-622 / :831-870
  double only_residual(bool gravity, std::vector<IMUST> &x_stats, LidarFactor &voxhess, std::deque<IMU_PRE *> &imus_factor) {
    double residual1 = 0, residual2 = 0;
    int thd_num = 5;
    int g_size = (int)voxhess.plvec_voxels.size();
    if (g_size < thd_num) thd_num = 1;
    std::vector<double> residuals(thd_num, 0.0);
    double part = 1.0 * g_size / thd_num;
    for (int i = 0; i < win_size - 1; i++) {
      if (!gravity) { Mat<30, 30> jtj; Mat<30, 1> gg; residual1 += imus_factor[i]->evaluate<30>(x_stats[i], x_stats[i + 1], jtj, gg, false); }
      else { Mat<33, 33> jtj; Mat<33, 1> gg; residual1 += imus_factor[i]->evaluate<33>(x_stats[i], x_stats[i + 1], jtj, gg, false); }
    }
    residual1 *= (imu_coef * 0.5);
    fan_out(thd_num, use_threads, [&](int i) {
      voxhess.evaluate_only_residual(x_stats, (int)(part * i), (int)(part * (i + 1)), residuals[i]);
    });
    for (int i = 0; i < thd_num; i++) residual2 += residuals[i];
    return residual1 + residual2;
  }

  // voxel_map.hpp:624-713 (gravity=false: 3 fixed iterations, resis untouched) and
  // voxel_map.hpp:878-975 (gravity=true).
  void damping_iter_impl(bool gravity, std::vector<IMUST> &x_stats, LidarFactor &voxhess, std::deque<IMU_PRE *> &imus_factor,
                         std::vector<double> *resis, MatX *hess, int max_iter) {
    win_size = voxhess.win_size;
    jac_leng = win_size * 6;
    imu_leng = win_size * DIM + (gravity ? 3 : 0);
    const int nfix = gravity ? 6 : DIM;
    double u = 0.01, v = 2;
    MatX Hess(imu_leng, imu_leng);
    VecX JacT(imu_leng), dxi(imu_leng), Ddiag(imu_leng);
    hess->resize(imu_leng, imu_leng);
    double residual1 = 0, residual2 = 0, q;
    bool is_calc_hess = true;
    std::vector<IMUST> x_stats_temp = x_stats;

    for (int i = 0; i < max_iter; i++) {
      if (is_calc_hess) {
        residual1 = divide_thread(gravity, x_stats, voxhess, imus_factor, Hess, JacT);
        *hess = Hess;
      }
      if (gravity && i == 0 && resis) resis->push_back(residual1);

      for (int c = 0; c < imu_leng; c++) for (int r = 0; r < nfix; r++) Hess(r, c) = 0.0;
      for (int c = 0; c < nfix; c++) for (int r = 0; r < imu_leng; r++) Hess(r, c) = 0.0;
      for (int r = 0; r < nfix; r++) Hess(r, r) = 1.0;
      for (int r = 0; r < nfix; r++) JacT[r] = 0.0;

      for (int r = 0; r < imu_leng; r++) Ddiag[r] = Hess(r, r);
      MatX A = Hess;
      for (int r = 0; r < imu_leng; r++) A(r, r) += u * Ddiag[r];
      VecX nb(imu_leng);
      for (int r = 0; r < imu_leng; r++) nb[r] = -JacT[r];
      ldlt_solve(A, nb, dxi);

      if (gravity) x_stats_temp[0].g += dxi.seg<3>(imu_leng - 3);
      for (int j = 0; j < win_size; j++) {
        x_stats_temp[j].R = x_stats[j].R * Exp(dxi.seg<3>(DIM * j));
        x_stats_temp[j].p = x_stats[j].p + dxi.seg<3>(DIM * j + 3);
        x_stats_temp[j].v = x_stats[j].v + dxi.seg<3>(DIM * j + 6);
        x_stats_temp[j].bg = x_stats[j].bg + dxi.seg<3>(DIM * j + 9);
        x_stats_temp[j].ba = x_stats[j].ba + dxi.seg<3>(DIM * j + 12);
        if (gravity) x_stats_temp[j].g = x_stats_temp[0].g;
      }
      for (int j = 0; j < win_size - 1; j++) imus_factor[j]->update_state(dxi.seg<DIM>(DIM * j));

      double q1 = 0;
      for (int r = 0; r < imu_leng; r++) q1 += dxi[r] * (u * Ddiag[r] * dxi[r] - JacT[r]);
      q1 *= 0.5;

      residual2 = only_residual(gravity, x_stats_temp, voxhess, imus_factor);
      q = (residual1 - residual2);

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
        for (int j = 0; j < win_size - 1; j++) {
          imus_factor[j]->dbg = imus_factor[j]->dbg_buf;
          imus_factor[j]->dba = imus_factor[j]->dba_buf;
        }
        accept_trace.push_back(0);
      }
      if (std::fabs((residual1 - residual2) / residual1) < 1e-6) break;
    }
    if (gravity && resis) resis->push_back(residual2);
  }
};

class LI_BA_Optimizer : public LI_BA_Common {
public:
  void damping_iter(std::vector<IMUST> &x_stats, LidarFactor &voxhess, std::deque<IMU_PRE *> &imus_factor, MatX *hess) {
    damping_iter_impl(false, x_stats, voxhess, imus_factor, nullptr, hess, 3);
  }
};

class LI_BA_OptimizerGravity : public LI_BA_Common {
public:
  void damping_iter(std::vector<IMUST> &x_stats, LidarFactor &voxhess, std::deque<IMU_PRE *> &imus_factor,
                    std::vector<double> &resis, MatX *hess, int max_iter = 2) {
    damping_iter_impl(true, x_stats, voxhess, imus_factor, &resis, hess, max_iter);
  }
};

}  // namespace orc
