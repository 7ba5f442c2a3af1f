// oracle/deskew.hpp — TEST INFRASTRUCTURE ONLY (CPU restatement; parity unpinned: the reference ships no vectors).
// The per-point half of IMUEKF::motion_blur, /root/reference/VoxelSLAM/src/ekf_imu.hpp:138-163: every point of a sweep is moved
// into the sensor frame at the END of the sweep with the IMU-propagated pose that precedes its time stamp (PointType::curvature).
// The IMU propagation that produces imu_poses (ekf_imu.hpp:57-120) is the EKF and stays with the caller (out of scope).
#pragma once
#include "vmath.hpp"
#include <vector>

namespace orc {
struct ImuPose { double t; M3 R; V3 p, v, angvel, acc; };   // IMUST as emplace_back(offt, R_imu, pos_imu, vel_imu, angvel_avr, acc_imu) fills it
struct SkewPoint { float x, y, z, curvature; };

// The loop of ekf_imu.hpp:138-163, statement for statement, including that the walk stops at the first point from the back whose
// stamp does not exceed the pose's time, and that pcl_in.begin() is compensated again by every earlier pose once it is reached.
// assign (optional): index of the LAST pose applied to each point, -1 = untouched.
inline void motion_blur_points(SkewPoint *pts, long n, const std::vector<ImuPose> &imu_poses, const M3 &xcR, const V3 &xcp, const M3 &Lid_rot_to_IMU,
                               const V3 &Lid_offset_to_IMU, int *assign) {
  if (n <= 0) return;
  if (assign) for (long k = 0; k < n; k++) assign[k] = -1;
  long it = n - 1;
  for (int i = (int)imu_poses.size() - 1; i >= 0; i--) {
    const ImuPose &head = imu_poses[i];
    const M3 R_imu = head.R;
    const V3 acc_imu = head.acc, vel_imu = head.v, pos_imu = head.p, angvel_avr = head.angvel;
    for (; (double)pts[it].curvature > head.t; it--) {
      const double dt = (double)pts[it].curvature - head.t;
      const M3 R_i = R_imu * Exp(angvel_avr, dt);
      const V3 T_ei = pos_imu + vel_imu * dt + 0.5 * acc_imu * dt * dt - xcp;
      const V3 P_i = vec3(pts[it].x, pts[it].y, pts[it].z);
      const V3 P_compensate = Lid_rot_to_IMU.transpose() * (xcR.transpose() * (R_i * (Lid_rot_to_IMU * P_i + Lid_offset_to_IMU) + T_ei) - Lid_offset_to_IMU);
      pts[it].x = (float)P_compensate[0];
      pts[it].y = (float)P_compensate[1];
      pts[it].z = (float)P_compensate[2];
      if (assign) assign[it] = i;
      if (it == 0) break;
    }
  }
}
}  // namespace orc
