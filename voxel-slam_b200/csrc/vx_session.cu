// vx_session.cu — the stored-session formats either side of the path, host code:
//   alidarState.txt   save_pose voxelslam.cpp:181-204 / read_lidarstate voxelslam.hpp:268-306
//   <i>.pcd           save_pcd voxelslam.cpp:166-179 (pcl::io::savePCDFileBinary of PointXYZI) / loadPCDFile :340
//   edge.txt          pgo_edges_io voxelslam.cpp:207-279
// so that recorded sessions can be fed to the device path (vx_keyframe_assemble, vx_hba_add_edge, vx_gba_build).  Number
// formatting is the reference's iostream formatting restated with printf conversions: `fixed << setprecision(6)` = "%.6f",
// the sticky `setprecision(7)` after it = "%.7f", a fresh stream with `setprecision(7)` = "%.7g".  Quaternion <-> matrix follow
// Eigen's Quaterniond(Matrix3d) and Quaterniond::matrix() operation for operation.
#include "vx_common.cuh"
#include <cerrno>
#include <cmath>
#include <cstdlib>
#include <string>
#include <vector>

using namespace vx;

namespace {

void quat_to_R(double w, double x, double y, double z, double *R) {   // QuaternionBase::toRotationMatrix, R column-major
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[3] = txy - twz;       R[6] = txz + twy;
  R[1] = txy + twz;       R[4] = 1 - (txx + tzz); R[7] = tyz - twx;
  R[2] = txz - twy;       R[5] = tyz + twx;       R[8] = 1 - (txx + tyy);
}
void R_to_quat(const double *R, double *q) {                           // quaternionbase_assign_impl<Matrix3d>; q = x y z w
#define M_(r, c) R[(c) * 3 + (r)]
  double t = M_(0, 0) + M_(1, 1) + M_(2, 2);
  if (t > 0) {
    t = std::sqrt(t + 1.0);
    q[3] = 0.5 * t;
    t = 0.5 / t;
    q[0] = (M_(2, 1) - M_(1, 2)) * t; q[1] = (M_(0, 2) - M_(2, 0)) * t; q[2] = (M_(1, 0) - M_(0, 1)) * t;
  } else {
    int i = 0;
    if (M_(1, 1) > M_(0, 0)) i = 1;
    if (M_(2, 2) > M_(i, i)) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(M_(i, i) - M_(j, j) - M_(k, k) + 1.0);
    q[i] = 0.5 * t;
    t = 0.5 / t;
    q[3] = (M_(k, j) - M_(j, k)) * t;
    q[j] = (M_(j, i) + M_(i, j)) * t;
    q[k] = (M_(k, i) + M_(i, k)) * t;
  }
#undef M_
}

struct File {
  FILE *f = nullptr;
  File(const char *path, const char *mode) { f = fopen(path, mode); }
  ~File() { if (f) fclose(f); }
};

bool read_line(FILE *f, std::string &line) {
  line.clear();
  int c;
  bool any = false;
  while ((c = fgetc(f)) != EOF) { any = true; if (c == '\n') break; line.push_back((char)c); }
  return any;
}

}  // namespace

extern "C" {

void vx_quat_to_R(const double *q_xyzw, double *R9) { quat_to_R(q_xyzw[3], q_xyzw[0], q_xyzw[1], q_xyzw[2], R9); }
void vx_R_to_quat(const double *R9, double *q_xyzw) { R_to_quat(R9, q_xyzw); }

int vx_session_write_lidarstate(const char *path, const vx_scan_state *xs, int64_t n) {
  VX_ARG(path && n >= 0 && (n == 0 || xs));
  File fp(path, "w");
  if (!fp.f) { set_error("cannot open %s for writing", path); return VX_ERR_ARG; }
  for (int64_t i = 0; i < n; i++) {
    const vx_scan_state &x = xs[i];
    double q[4];
    R_to_quat(x.R, q);
    fprintf(fp.f, "%.6f ", x.t);
    fprintf(fp.f, "%.7f %.7f %.7f ", x.p[0], x.p[1], x.p[2]);
    fprintf(fp.f, "%.7f %.7f %.7f %.7f", q[0], q[1], q[2], q[3]);
    const double *blocks[4] = {x.v, x.bg, x.ba, x.g};
    for (const double *b : blocks) fprintf(fp.f, " %.7f %.7f %.7f", b[0], b[1], b[2]);
    for (int j = 0; j < 6; j++) fprintf(fp.f, " %.7f", x.v6[j]);
    fputc('\n', fp.f);
  }
  return VX_OK;
}

int vx_session_read_lidarstate(const char *path, vx_scan_state *out, int64_t cap, int64_t *n) {
  VX_ARG(path && n && cap >= 0 && (cap == 0 || out));
  File fp(path, "r");
  if (!fp.f) { set_error("Error: %s not found", path); return VX_ERR_ARG; }   // the reference prints this and exits (:271-275)
  std::string line;
  std::vector<double> nums;
  int64_t cnt = 0;
  while (read_line(fp.f, line)) {
    nums.clear();
    // getline(ss, str, ' ') + stod(str): single-space separated fields; an empty field is an error there (stod throws)
    size_t a = 0;
    while (a <= line.size()) {
      size_t b = line.find(' ', a);
      if (b == std::string::npos) b = line.size();
      if (b == a && a == line.size()) break;              // getline on an exhausted stream yields nothing
      const std::string tok = line.substr(a, b - a);
      char *end = nullptr;
      errno = 0;
      const double v = strtod(tok.c_str(), &end);
      if (end == tok.c_str()) { set_error("%s: line %lld: field '%s' is not a number", path, (long long)cnt + 1, tok.c_str()); return VX_ERR_ARG; }
      nums.push_back(v);
      a = b + 1;
    }
    if (nums.size() < 8) { set_error("%s: line %lld has %zu fields (< 8)", path, (long long)cnt + 1, nums.size()); return VX_ERR_ARG; }
    if (cnt < cap) {
      vx_scan_state x;
      memset(&x, 0, sizeof(x));
      x.R[0] = x.R[4] = x.R[8] = 1.0; x.g[2] = -9.8;      // IMUST::setZero (tools.hpp:188-197)
      x.t = nums[0];
      x.p[0] = nums[1]; x.p[1] = nums[2]; x.p[2] = nums[3];
      quat_to_R(nums[7], nums[4], nums[5], nums[6], x.R);
      if (nums.size() >= 20)
        for (int k = 0; k < 3; k++) { x.v[k] = nums[8 + k]; x.bg[k] = nums[11 + k]; x.ba[k] = nums[14 + k]; x.g[k] = nums[17 + k]; }
      if (nums.size() >= 26)
        for (int k = 0; k < 6; k++) x.v6[k] = nums[20 + k];
      out[cnt] = x;
    }
    cnt++;
  }
  *n = cnt;
  if (cnt > cap) { set_error("%s holds %lld scans, capacity %lld", path, (long long)cnt, (long long)cap); return VX_ERR_ARG; }
  return VX_OK;
}

int vx_session_write_pcd_xyzi(const char *path, const float *xyzi, int64_t n) {
  VX_ARG(path && n >= 0 && (n == 0 || xyzi));
  File fp(path, "wb");
  if (!fp.f) { set_error("cannot open %s for writing", path); return VX_ERR_ARG; }
  fprintf(fp.f, "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z intensity\nSIZE 4 4 4 4\nTYPE F F F F\nCOUNT 1 1 1 1\n"
                "WIDTH %lld\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %lld\nDATA binary\n", (long long)n, (long long)n);
  if (n && fwrite(xyzi, 16, (size_t)n, fp.f) != (size_t)n) { set_error("short write to %s", path); return VX_ERR_ARG; }
  return VX_OK;
}

// pcl::io::loadPCDFile into PointCloud<PointXYZI> (voxelslam.cpp:340): x, y, z and intensity by field NAME from a v0.7 file,
// DATA ascii or binary (what savePCDFileBinary / savePCDFileASCII write); a missing intensity field reads as 0.
int vx_session_read_pcd_xyzi(const char *path, float *xyzi, int64_t cap, int64_t *n) {
  VX_ARG(path && n && cap >= 0 && (cap == 0 || xyzi));
  File fp(path, "rb");
  if (!fp.f) { set_error("cannot open %s", path); return VX_ERR_ARG; }
  std::vector<std::string> fields; std::vector<int> sizes, counts; std::vector<char> types;
  int64_t points = -1, width = -1, height = 1;
  std::string line, data;
  auto split = [](const std::string &l) { std::vector<std::string> t; size_t a = 0; while (a < l.size()) { while (a < l.size() && isspace((unsigned char)l[a])) a++; size_t b = a; while (b < l.size() && !isspace((unsigned char)l[b])) b++; if (b > a) t.push_back(l.substr(a, b - a)); a = b; } return t; };
  while (read_line(fp.f, line)) {
    if (!line.empty() && line[0] == '#') continue;
    const std::vector<std::string> t = split(line);
    if (t.empty()) continue;
    if (t[0] == "FIELDS") fields.assign(t.begin() + 1, t.end());
    else if (t[0] == "SIZE") for (size_t k = 1; k < t.size(); k++) sizes.push_back(atoi(t[k].c_str()));
    else if (t[0] == "TYPE") for (size_t k = 1; k < t.size(); k++) types.push_back(t[k][0]);
    else if (t[0] == "COUNT") for (size_t k = 1; k < t.size(); k++) counts.push_back(atoi(t[k].c_str()));
    else if (t[0] == "WIDTH" && t.size() > 1) width = atoll(t[1].c_str());
    else if (t[0] == "HEIGHT" && t.size() > 1) height = atoll(t[1].c_str());
    else if (t[0] == "POINTS" && t.size() > 1) points = atoll(t[1].c_str());
    else if (t[0] == "DATA" && t.size() > 1) { data = t[1]; break; }
  }
  if (points < 0 && width >= 0) points = width * height;
  if (counts.empty()) counts.assign(fields.size(), 1);
  if (fields.empty() || sizes.size() != fields.size() || types.size() != fields.size() || counts.size() != fields.size() || points < 0 ||
      (data != "binary" && data != "ascii")) {
    set_error("%s: unsupported or malformed PCD header (DATA '%s')", path, data.c_str());
    return VX_ERR_ARG;
  }
  int off[4] = {-1, -1, -1, -1}, col[4] = {-1, -1, -1, -1}, stride = 0, ncol = 0;
  const char *want[4] = {"x", "y", "z", "intensity"};
  for (size_t k = 0; k < fields.size(); k++) {
    for (int w = 0; w < 4; w++)
      if (fields[k] == want[w]) {
        if (sizes[k] != 4 || types[k] != 'F' || counts[k] != 1) { set_error("%s: field %s is not a float32", path, want[w]); return VX_ERR_ARG; }
        off[w] = stride; col[w] = ncol;
      }
    stride += sizes[k] * counts[k]; ncol += counts[k];
  }
  if (off[0] < 0 || off[1] < 0 || off[2] < 0) { set_error("%s: no x / y / z fields", path); return VX_ERR_ARG; }
  *n = points;
  if (points > cap) { set_error("%s holds %lld points, capacity %lld", path, (long long)points, (long long)cap); return VX_ERR_ARG; }
  if (data == "binary") {
    std::vector<unsigned char> rec((size_t)stride);
    for (int64_t i = 0; i < points; i++) {
      if (fread(rec.data(), 1, (size_t)stride, fp.f) != (size_t)stride) { set_error("%s: truncated at point %lld", path, (long long)i); return VX_ERR_ARG; }
      for (int w = 0; w < 4; w++) { float v = 0.f; if (off[w] >= 0) memcpy(&v, rec.data() + off[w], 4); xyzi[4 * i + w] = v; }
    }
  } else {
    for (int64_t i = 0; i < points; i++) {
      if (!read_line(fp.f, line)) { set_error("%s: truncated at point %lld", path, (long long)i); return VX_ERR_ARG; }
      const std::vector<std::string> t = split(line);
      if ((int)t.size() < ncol) { set_error("%s: point %lld has %zu columns", path, (long long)i, t.size()); return VX_ERR_ARG; }
      for (int w = 0; w < 4; w++) xyzi[4 * i + w] = col[w] >= 0 ? strtof(t[col[w]].c_str(), nullptr) : 0.f;
    }
  }
  return VX_OK;
}

// edge.txt, write side (voxelslam.cpp:258-276): "name1 name2 id1 id2 tx ty tz qx qy qz qw", one line per stored relative pose;
// `absent` (may be NULL) = the verbatim lines of sessions not loaded, written first (:260-261).
int vx_session_write_edges(const char *path, const char *const *fnames, int n_names, const vx_pgo_edge *edges, int64_t n, const char *absent) {
  VX_ARG(path && n >= 0 && (n == 0 || (edges && fnames)));
  File fp(path, "w");
  if (!fp.f) { set_error("cannot open %s for writing", path); return VX_ERR_ARG; }
  if (absent) fputs(absent, fp.f);
  for (int64_t i = 0; i < n; i++) {
    const vx_pgo_edge &e = edges[i];
    VX_ARG(e.m1 >= 0 && e.m1 < n_names && e.m2 >= 0 && e.m2 < n_names);
    double q[4];
    R_to_quat(e.rot, q);
    fprintf(fp.f, "%s %s %d %d %.7g %.7g %.7g %.7g %.7g %.7g %.7g\n", fnames[e.m1], fnames[e.m2], e.id1, e.id2, e.tra[0], e.tra[1], e.tra[2],
            q[0], q[1], q[2], q[3]);
  }
  return VX_OK;
}

// edge.txt, read side (voxelslam.cpp:216-255).  Lines whose two session names are both in fnames become edges (cov = 1e-6
// each, :213), ordered so that m1 <= m2 — a reversed line is inverted: tra <- -R^T tra, R <- R^T, ids swapped (:244-250).
// Other lines that do not mention `bagname` are appended verbatim to absent_buf (:252-255) for the next write.
int vx_session_read_edges(const char *path, const char *const *fnames, int n_names, const char *bagname, vx_pgo_edge *edges, int64_t cap,
                          int64_t *n, char *absent_buf, int64_t absent_cap, int64_t *absent_len) {
  VX_ARG(path && n && cap >= 0 && (cap == 0 || edges) && (n_names == 0 || fnames));
  *n = 0;
  if (absent_len) *absent_len = 0;
  File fp(path, "r");
  if (!fp.f) return VX_OK;                                   // ifstream on a missing file: the read loop never runs
  std::string line, absent;
  int64_t cnt = 0;
  while (read_line(fp.f, line)) {
    std::vector<std::string> sts;
    { size_t a = 0; while (a < line.size()) { while (a < line.size() && isspace((unsigned char)line[a])) a++; size_t b = a; while (b < line.size() && !isspace((unsigned char)line[b])) b++; if (b > a) sts.push_back(line.substr(a, b - a)); a = b; } }
    if (sts.size() < 2) continue;
    int mp[2] = {-1, -1};
    for (int i = 0; i < 2; i++)
      for (int j = 0; j < n_names; j++)
        if (sts[i] == fnames[j]) { mp[i] = j; break; }
    if (mp[0] != -1 && mp[1] != -1) {
      if (sts.size() < 11) { set_error("%s: edge line with %zu fields", path, sts.size()); return VX_ERR_ARG; }
      const int id1 = atoi(sts[2].c_str()), id2 = atoi(sts[3].c_str());
      double v3[3] = {strtod(sts[4].c_str(), nullptr), strtod(sts[5].c_str(), nullptr), strtod(sts[6].c_str(), nullptr)};
      double rot[9];
      quat_to_R(strtod(sts[10].c_str(), nullptr), strtod(sts[7].c_str(), nullptr), strtod(sts[8].c_str(), nullptr), strtod(sts[9].c_str(), nullptr), rot);
      vx_pgo_edge e;
      for (int k = 0; k < 6; k++) e.cov[k] = 1e-6;
      if (mp[0] <= mp[1]) {
        e.m1 = mp[0]; e.m2 = mp[1]; e.id1 = id1; e.id2 = id2;
        for (int k = 0; k < 9; k++) e.rot[k] = rot[k];
        for (int k = 0; k < 3; k++) e.tra[k] = v3[k];
      } else {
        e.m1 = mp[1]; e.m2 = mp[0]; e.id1 = id2; e.id2 = id1;
        for (int r = 0; r < 3; r++) {           // -rot^T v3: row r of rot^T = column r of rot; Eigen negates the matrix first
          double sacc = 0.0;
          for (int k = 0; k < 3; k++) sacc += (-rot[r * 3 + k]) * v3[k];
          e.tra[r] = sacc;
          for (int c = 0; c < 3; c++) e.rot[c * 3 + r] = rot[r * 3 + c];
        }
      }
      if (cnt < cap) edges[cnt] = e;
      cnt++;
    } else if (!bagname || (sts[0] != bagname && sts[1] != bagname)) {
      absent += line; absent += "\n";
    }
  }
  *n = cnt;
  if (cnt > cap) { set_error("%s holds %lld edges, capacity %lld", path, (long long)cnt, (long long)cap); return VX_ERR_ARG; }
  if (absent_len) *absent_len = (int64_t)absent.size();
  if (absent_buf) {
    if ((int64_t)absent.size() + 1 > absent_cap) { set_error("absent-line buffer too small (%zu bytes needed)", absent.size() + 1); return VX_ERR_ARG; }
    memcpy(absent_buf, absent.c_str(), absent.size() + 1);
  }
  return VX_OK;
}

}  // extern "C"
