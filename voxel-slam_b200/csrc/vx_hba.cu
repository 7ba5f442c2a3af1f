// vx_hba.cu — VOXEL_SLAM::HBA_add_edge (voxelslam.cpp:2822-3015) over the device path:
//   per outer iteration: OctreeGBA::cut_voxel + OctreeGBA_multi_recut (vx_gba.cu) -> Lidar_BA_Optimizer::damping_iter
//   with up = 4 (vx_ba.cu); coarse -> fine threshold swap (:2873-2883); convergence logic (:2906-2917);
//   pose-graph edges from the pre-gauge Hessian snapshot (:2926-2951).
// A call may carry a BATCH of independent windows (the 399 bottom-level windows of a 2,000-keyframe
// session, voxelslam.cpp:3096): each window keeps its own converge_flag / threshold state; windows that
// want the coarse and the fine thresholds in the same outer iteration are built as two sub-batches.
#include "vx_factor.cuh"
#include "vx_voxelize.cuh"
#include "vx_hba.cuh"
#include <cmath>
#include <algorithm>
#include <memory>

using namespace vx;

namespace vx {
// One HBA_add_edge (voxelslam.cpp:2860-2951) over NW independent windows of W frames each.  `build` makes the LidarFactor of
// the windows `ws` at the poses `ps` (ns * W entries) with the thresholds `gp`: the single-GPU map build, or the sharded one
// (then NW == 1 and `comm` sums every evaluation over the ranks).
int hba_core(vx_ctx *ctx, vx_comm *comm, const HbaBuildFn &build, const int32_t *frame_id, const int32_t *frame_mp, vx_pose *poses, int W, int NW,
             const vx_hba_params *prm, int max_iter, int thread_num, vx_pgo_edge *edges, int64_t *n_edges, double *hess_out, double *resis_out,
             int32_t *n_voxels_out, int32_t *status_out, int64_t *n_voxel_frames_evaluated, vx_factor **keep) {
  const int jl = 6 * W;
  const size_t hsz = (size_t)jl * jl;
  const int64_t cap = (int64_t)W * (W - 1) / 2;   // edge slots per window
  // the full snapshot only when the caller wants it; the edges need 6 numbers per frame pair (k_edge_diag)
  // ediag / ed are written in full before they are read (k_edge_diag fills every slot of every window that reports VX_OK): no zero
  // fill — at the top level (W = 400) each is 3.8 MB of fresh pages, and the fill showed as 0.6-1.3 ms per outer iteration
  std::vector<double> hess(hess_out ? hsz * NW : 0, 0.0);
  std::unique_ptr<double[]> ediag(new double[(size_t)NW * cap * 6]);
  std::vector<int> cflag(NW, 0), done(NW, 0), status(NW, VX_OK), nvox(NW, 0);
  std::vector<double> cthre(NW, 0.05);
  const int up = 4;   // voxelslam.cpp:2867
  PhaseTimer pt("hba_core", ctx->stream);
  for (int iterCnt = 0; iterCnt < max_iter; iterCnt++) {
    // which thresholds does each live window use in this outer iteration?
    std::vector<int> group[2];
    for (int w = 0; w < NW; w++) {
      if (done[w]) continue;
      const bool fine = cflag[w] == 1 || iterCnt == max_iter - 1;   // voxelslam.cpp:2873
      group[fine ? 1 : 0].push_back(w);
    }
    if (group[0].empty() && group[1].empty()) break;
    for (int gi = 0; gi < 2; gi++) {
      const std::vector<int> &ws = group[gi];
      if (ws.empty()) continue;
      const int ns = (int)ws.size();
      vx_gba_params gp = prm->gba;
      if (gi == 1) {   // gba_* <- LocalBA thresholds (voxelslam.cpp:2875-2882)
        gp.voxel_size = prm->voxel_size; gp.min_eigen_value = prm->min_eigen_value;
        for (int k = 0; k < 4; k++) gp.eigen_value_array[k] = prm->plane_eigen_value_thre[k];
      }
      std::vector<vx_pose> ps((size_t)ns * W);
      for (int a = 0; a < ns; a++)
        for (int i = 0; i < W; i++) ps[(size_t)a * W + i] = poses[ws[a] * W + i];
      vx_factor *fac = nullptr;
      VX_TRY(vx_factor_create(ctx, W, ns, &fac));
      std::vector<double> hs(hess_out ? hsz * ns : 0), rs(2 * (size_t)ns);
      // the edge diagonals of an outer iteration are only read for windows that END with it: the last iteration, or a window that
      // entered it with cflag == 1 (voxelslam.cpp:2906-2917 — only such a window can become done).  Otherwise skip the kernel + copy
      // (0.6 ms per top-level iteration at W = 400).
      bool need_ed = iterCnt == max_iter - 1;
      for (int w : ws) need_ed = need_ed || cflag[w] == 1;
      std::unique_ptr<double[]> ed(need_ed ? new double[(size_t)ns * cap * 6] : nullptr);
      std::vector<int> conv(ns), wst(ns), nit(ns, 0);
      pt.lap("iteration setup");
      int s = build(ws, ps.data(), gp, fac);
      pt.lap("build");
      if (s == VX_OK) s = ba_lidar_impl(ctx, fac, ps.data(), hess_out ? hs.data() : nullptr, rs.data(), up, thread_num, conv.data(), nullptr, nit.data(), wst.data(), comm, ed.get());
      std::vector<int32_t> wv(ns + 1, 0);
      if (s == VX_OK && fac->V > 0) {
        cudaMemcpyAsync(wv.data(), fac->wv_off.p, sizeof(int32_t) * (ns + 1), cudaMemcpyDeviceToHost, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
      }
      pt.lap("damping_iter");
      if (n_voxel_frames_evaluated) {   // [0] voxel-frames through acc_evaluate2, [1] through evaluate_only_residual, [2] / [3] passes of each
        n_voxel_frames_evaluated[0] += fac->nnz * fac->n_hess_evals; n_voxel_frames_evaluated[1] += fac->nnz * fac->n_res_evals;
        n_voxel_frames_evaluated[2] += fac->n_hess_evals; n_voxel_frames_evaluated[3] += fac->n_res_evals;
      }
      if (keep && s == VX_OK) { if (*keep) vx_factor_free(*keep); *keep = fac; }   // the latest factor stays with the caller
      else vx_factor_free(fac);
      pt.lap("factor_free");
      VX_TRY(s);
      for (int a = 0; a < ns; a++) {
        const int w = ws[a];
        nvox[w] = wv[a + 1] - wv[a];
        if (wst[a] != VX_OK) { status[w] = wst[a]; done[w] = 1; continue; }   // reference: "Too Less Voxel", exit(0)
        for (int i = 0; i < W; i++) poses[w * W + i] = ps[(size_t)a * W + i];
        if (hess_out) std::copy(hs.begin() + hsz * a, hs.begin() + hsz * (a + 1), hess.begin() + hsz * w);
        if (need_ed) std::copy(ed.get() + (size_t)a * cap * 6, ed.get() + (size_t)(a + 1) * cap * 6, ediag.get() + (size_t)w * cap * 6);
        const double r0 = rs[2 * a], r1 = rs[2 * a + 1];
        if (resis_out) { resis_out[2 * w] = r0; resis_out[2 * w + 1] = r1; }
        // voxelslam.cpp:2906-2917
        if ((std::fabs(r0 - r1) / r0 < cthre[w] && conv[a]) || (iterCnt == max_iter - 2 && cflag[w] == 0)) {
          cthre[w] = 0.01;
          if (cflag[w] == 0) cflag[w] = 1;
          else if (cflag[w] == 1) done[w] = 1;
        }
      }
    }
  }
  pt.lap("iteration close");
  // pose-graph edges (voxelslam.cpp:2926-2951): cov_k = 1/|hess(6i+k, 6j+k)| of the pre-gauge snapshot
  for (int w = 0; w < NW; w++) {
    int64_t ne = 0;
    if (status[w] == VX_OK && edges) {
      const double *D = ediag.get() + (size_t)w * cap * 6;
      int64_t pr = 0;
      for (int i = 0; i < W - 1; i++)
        for (int j = i + 1; j < W; j++, pr++) {
          vx_pgo_edge e;
          bool add = true;
          for (int k = 0; k < 6; k++) {
            const double hc = std::fabs(D[pr * 6 + k]);
            if (hc < 1e-6) { add = false; break; }
            e.cov[k] = 1.0 / hc;
          }
          if (!add) continue;
          const vx_pose &a = poses[w * W + i], &b = poses[w * W + j];
          e.m1 = frame_mp ? frame_mp[w * W + i] : 0; e.m2 = frame_mp ? frame_mp[w * W + j] : 0;
          e.id1 = frame_id ? frame_id[w * W + i] : i; e.id2 = frame_id ? frame_id[w * W + j] : j;
          const double d[3] = {b.p[0] - a.p[0], b.p[1] - a.p[1], b.p[2] - a.p[2]};
          for (int r = 0; r < 3; r++) {   // R_i^T (p_j - p_i), R_i^T R_j  (column-major)
            double sacc = 0.0;
            for (int k = 0; k < 3; k++) sacc += a.R[r * 3 + k] * d[k];
            e.tra[r] = sacc;
            for (int c = 0; c < 3; c++) {
              double m = 0.0;
              for (int k = 0; k < 3; k++) m += a.R[r * 3 + k] * b.R[c * 3 + k];
              e.rot[c * 3 + r] = m;
            }
          }
          edges[(size_t)w * cap + ne++] = e;
        }
    }
    if (n_edges) n_edges[w] = ne;
    if (status_out) status_out[w] = status[w];
    if (n_voxels_out) n_voxels_out[w] = nvox[w];
  }
  if (hess_out) std::copy(hess.begin(), hess.end(), hess_out);
  pt.lap("edges");
  return VX_OK;
}
}  // namespace vx

extern "C" int vx_hba_add_edge(vx_ctx *ctx, const float *xyz, const int64_t *cloud_off, int n_clouds, const int32_t *frame_cloud,
                               const int32_t *frame_id, const int32_t *frame_mp, vx_pose *poses, int W, int n_windows,
                               const vx_hba_params *prm, int max_iter, int thread_num, vx_pgo_edge *edges, int64_t *n_edges,
                               double *hess_out, double *resis_out, int32_t *n_voxels_out, int32_t *status_out) {
  VX_ARG(prm != nullptr && poses != nullptr && cloud_off != nullptr && W >= 2 && n_windows >= 1 && max_iter >= 0 && n_clouds >= 1);
  VX_TRY(bind(ctx));
  const int64_t NP = cloud_off[n_clouds];
  VX_ARG(NP >= 0 && (NP == 0 || xyz != nullptr));
  // clouds go to the device once for the whole call
  DevBuf d_xyz, d_coff;
  auto body = [&]() -> int {
    VX_TRY(d_xyz.reserve(sizeof(float) * 3 * std::max<int64_t>(NP, 1)));
    VX_TRY(d_coff.reserve(8 * (size_t)(n_clouds + 1)));
    if (NP) VX_CUDA(cudaMemcpyAsync(d_xyz.p, xyz, sizeof(float) * 3 * NP, cudaMemcpyHostToDevice, ctx->stream));
    VX_CUDA(cudaMemcpyAsync(d_coff.p, cloud_off, 8 * (size_t)(n_clouds + 1), cudaMemcpyHostToDevice, ctx->stream));
    HbaBuildFn build = [&](const std::vector<int> &ws, const vx_pose *ps, const vx_gba_params &gp, vx_factor *fac) -> int {
      std::vector<int32_t> fc(ws.size() * (size_t)W);
      for (size_t a = 0; a < ws.size(); a++)
        for (int i = 0; i < W; i++) { const int g = ws[a] * W + i; fc[a * W + i] = frame_cloud ? frame_cloud[g] : g; }
      return gba_build_impl(ctx, d_xyz.as<float>(), d_coff.as<int64_t>(), cloud_off, n_clouds, fc.data(), ps, &gp, fac, nullptr, 0);
    };
    return hba_core(ctx, nullptr, build, frame_id, frame_mp, poses, W, n_windows, prm, max_iter, thread_num, edges, n_edges, hess_out, resis_out,
                    n_voxels_out, status_out, nullptr);
  };
  const int rc = body();
  cudaStreamSynchronize(ctx->stream);
  d_xyz.release(); d_coff.release();
  return rc;
}

// ---- merged submap cloud of HBA_add_edge (voxelslam.cpp:2954-2985) ---------------------------------------------------
namespace {
struct RelPose { double dR[9]; double dp[3]; float mp; };   // dR row-major
// every point of keyframe i into the frame of keyframe 0: v3 = dR * v3 + dp (double), stored back as float (:2963-2969)
__global__ void k_merge_transform(const float *__restrict__ xyz, int64_t beg, int64_t n, RelPose rp, vx_point48 *__restrict__ out) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= n) return;
  const float *s = xyz + 3 * (beg + k);
  const double x = (double)s[0], y = (double)s[1], z = (double)s[2];
  vx_point48 ap;
  ap._p0 = 0.0f; ap.normal_x = ap.normal_y = ap.normal_z = ap._p1 = 0.0f; ap.curvature = 0.0f; ap._p2 = ap._p3 = 0.0f;
  double v[3];
#pragma unroll
  for (int r = 0; r < 3; r++)
    v[r] = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(rp.dR[r * 3], x), __dmul_rn(rp.dR[r * 3 + 1], y)), __dmul_rn(rp.dR[r * 3 + 2], z)), rp.dp[r]);
  ap.x = (float)v[0]; ap.y = (float)v[1]; ap.z = (float)v[2];
  ap.intensity = rp.mp;
  out[k] = ap;
}
}  // namespace

// Clouds cidx[0..W) moved into the frame of pose `ref` (dp = x_ref.R^T (x_i.p - x_ref.p), dR = x_ref.R^T x_i.R, v3 = dR v3 + dp
// in doubles, narrowed to float), intensity = mp[i], then down_sampling_voxel(leaf).  Result: ncell records in d_out.
static int assemble_window(vx_ctx *ctx, const float *d_xyz, const int64_t *cloud_off, const int *cidx, const vx_pose *xs, const int32_t *mp,
                           int W, int ref, double leaf, DevBuf &d_in, DevBuf &d_out, int32_t *ncell) {
  cudaStream_t st = ctx->stream;
  int64_t total = 0;
  for (int i = 0; i < W; i++) total += cloud_off[cidx[i] + 1] - cloud_off[cidx[i]];
  *ncell = 0;
  if (total == 0) return VX_OK;
  VX_TRY(d_in.reserve(sizeof(vx_point48) * total)); VX_TRY(d_out.reserve(sizeof(vx_point48) * total));
  const vx_pose &xc = xs[ref];
  int64_t at = 0;
  for (int i = 0; i < W; i++) {
    const int c = cidx[i];
    const int64_t n = cloud_off[c + 1] - cloud_off[c];
    if (n == 0) continue;
    const vx_pose &xi = xs[i];
    RelPose rp;
    // R column-major: R(r,c) = R[c*3+r]
    double d[3] = {xi.p[0] - xc.p[0], xi.p[1] - xc.p[1], xi.p[2] - xc.p[2]};
    for (int r = 0; r < 3; r++) {
      double sacc = 0; for (int k = 0; k < 3; k++) sacc += xc.R[r * 3 + k] * d[k];
      rp.dp[r] = sacc;
      for (int cc = 0; cc < 3; cc++) { double t = 0; for (int k = 0; k < 3; k++) t += xc.R[r * 3 + k] * xi.R[cc * 3 + k]; rp.dR[r * 3 + cc] = t; }
    }
    rp.mp = (float)(mp ? mp[i] : 0);
    VX_KERNEL(ctx, "k_merge_transform", k_merge_transform<<<blocks_for(n, 256), 256, 0, st>>>(d_xyz, cloud_off[c], n, rp, d_in.as<vx_point48>() + at));
    at += n;
  }
  return downsample_device(ctx, d_in.p, total, leaf, DS_MEAN, d_out.as<vx_point48>(), ncell);
}

// ---- all windows at once -----------------------------------------------------------------------------------------------------
#include <cub/cub.cuh>
namespace {
struct RelPoseD { double dR[9]; double dp[3]; };   // dR row-major, per frame

__device__ __forceinline__ int frame_of_inst(const int64_t *__restrict__ inst_off, int F, int64_t j) {
  int lo = 0, hi = F;
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (inst_off[mid] <= j) lo = mid; else hi = mid; }
  return lo;
}
// Key layout of the batched merge: window | x | y | z, each axis re-biased by the smallest cell any window holds and given
// exactly the bits its range needs (42 bits = 6 radix passes on the 2,000-keyframe job; the fixed 3 x 18 + 9 = 63-bit layout ran 8).
// The range comes from the clouds' bounding boxes (vx::cloud_bboxes): per frame the cells of the 8 transformed corners, one cell of margin.
struct MergeLayout { int lo[3]; int bx, by, bz; };
struct MergeRange { int lo[3], hi[3]; };
__global__ void k_merge_cells(const float *__restrict__ bbox, const int32_t *__restrict__ frame_cloud, int F, const RelPoseD *__restrict__ rel, double leaf,
                              MergeRange *__restrict__ range) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= F) return;
  const float *bb = bbox + 6 * (size_t)frame_cloud[g];
  if (!(bb[0] <= bb[3])) return;                       // empty cloud
  const RelPoseD &rp = rel[g];
  int lo[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff}, hi[3] = {-0x7fffffff, -0x7fffffff, -0x7fffffff};
  for (int c = 0; c < 8; c++) {
    const double x = (double)bb[(c & 1) ? 3 : 0], y = (double)bb[(c & 2) ? 4 : 1], z = (double)bb[(c & 4) ? 5 : 2];
#pragma unroll
    for (int r = 0; r < 3; r++) {
      const float v = (float)__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(rp.dR[r * 3], x), __dmul_rn(rp.dR[r * 3 + 1], y)), __dmul_rn(rp.dR[r * 3 + 2], z)), rp.dp[r]);
      const int64_t k = key_axis_f(v, leaf);
      const int q = (int)max((int64_t)-0x7ffffff0, min((int64_t)0x7ffffff0, k));
      lo[r] = min(lo[r], q - 1); hi[r] = max(hi[r], q + 1);
    }
  }
#pragma unroll
  for (int a = 0; a < 3; a++) { atomicMin(&range->lo[a], lo[a]); atomicMax(&range->hi[a], hi[a]); }
}
// frame of instance j given g0 <= frame(j): a walk of 0 or 1 steps for real clouds
__device__ __forceinline__ int frame_walk_inst(const int64_t *__restrict__ inst_off, int F, int g0, int64_t j) {
  int g = g0;
  while (g + 1 < F && inst_off[g + 1] <= j) g++;
  return g;
}
// every point instance of every window into its window's frame (:2958-2969, the arithmetic of k_merge_transform) + its batched key
__global__ void k_merge_all(const float *__restrict__ xyz, const int64_t *__restrict__ coff, const int32_t *__restrict__ frame_cloud,
                            const int64_t *__restrict__ inst_off, int F, const RelPoseD *__restrict__ rel, const int32_t *__restrict__ frame_win, double leaf,
                            MergeLayout ml, int64_t NI, float *__restrict__ out, uint64_t *__restrict__ keys, int32_t *__restrict__ idx, int *__restrict__ bad) {
  __shared__ int g0s;
  if (threadIdx.x == 0) g0s = frame_of_inst(inst_off, F, blockIdx.x * (int64_t)blockDim.x);
  __syncthreads();
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= NI) return;
  const int g = frame_walk_inst(inst_off, F, g0s, j);
  const float *s = xyz + 3 * (coff[frame_cloud[g]] + (j - inst_off[g]));
  const double x = (double)s[0], y = (double)s[1], z = (double)s[2];
  const RelPoseD &rp = rel[g];
  float v[3];
#pragma unroll
  for (int r = 0; r < 3; r++)
    v[r] = (float)__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(rp.dR[r * 3], x), __dmul_rn(rp.dR[r * 3 + 1], y)), __dmul_rn(rp.dR[r * 3 + 2], z)), rp.dp[r]);
  out[3 * j] = v[0]; out[3 * j + 1] = v[1]; out[3 * j + 2] = v[2];
  const int64_t kx = key_axis_f(v[0], leaf) - ml.lo[0], ky = key_axis_f(v[1], leaf) - ml.lo[1], kz = key_axis_f(v[2], leaf) - ml.lo[2];
  if (kx < 0 || ky < 0 || kz < 0 || kx >= ((int64_t)1 << ml.bx) || ky >= ((int64_t)1 << ml.by) || kz >= ((int64_t)1 << ml.bz)) *bad = 1;   // never, by construction
  keys[j] = ((uint64_t)frame_win[g] << (ml.bx + ml.by + ml.bz)) | ((uint64_t)kx << (ml.by + ml.bz)) | ((uint64_t)ky << ml.bz) | (uint64_t)kz;
  idx[j] = (int32_t)j;
}
__global__ void k_mheads(const uint64_t *__restrict__ keys, int64_t n, int32_t *__restrict__ flag) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) flag[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}
__global__ void k_mstarts(const int32_t *__restrict__ flag, const int32_t *__restrict__ incl, int64_t n, int32_t *__restrict__ start) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (flag[i]) start[incl[i] - 1] = (int32_t)i;
  if (i == n - 1) start[incl[i]] = (int32_t)n;
}
// one thread per cell: the reference's sequential float running mean (tools.hpp:227-230) over the cell's points in input order
__global__ void k_mmean(const float *__restrict__ pts, const int32_t *__restrict__ idx, const int32_t *__restrict__ start, int64_t ncell,
                        const int64_t *__restrict__ inst_off, int F, const int32_t *__restrict__ frame_mp, float *__restrict__ out, vx_point48 *__restrict__ out48) {
  const int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= ncell) return;
  const int32_t b = start[c], e = start[c + 1];
  const int32_t j0 = idx[b];
  float x = pts[3 * (int64_t)j0], y = pts[3 * (int64_t)j0 + 1], z = pts[3 * (int64_t)j0 + 2], cnt = 1.0f;
  for (int32_t k = b + 1; k < e; k++) {
    const float *q = pts + 3 * (int64_t)idx[k];
    const float c1 = __fadd_rn(cnt, 1.0f);
    x = __fdiv_rn(__fadd_rn(__fmul_rn(x, cnt), q[0]), c1);
    y = __fdiv_rn(__fadd_rn(__fmul_rn(y, cnt), q[1]), c1);
    z = __fdiv_rn(__fadd_rn(__fmul_rn(z, cnt), q[2]), c1);
    cnt = c1;
  }
  out[3 * c] = x; out[3 * c + 1] = y; out[3 * c + 2] = z;
  if (out48) {
    vx_point48 ap;
    ap.x = x; ap.y = y; ap.z = z; ap._p0 = 0.0f; ap.normal_x = ap.normal_y = ap.normal_z = ap._p1 = 0.0f;
    ap.intensity = frame_mp ? (float)frame_mp[frame_of_inst(inst_off, F, j0)] : 0.0f;
    ap.curvature = cnt; ap._p2 = ap._p3 = 0.0f;
    out48[c] = ap;
  }
}
// win_off[w] = first cell whose window id >= w
__global__ void k_mwin_off(const uint64_t *__restrict__ keys, const int32_t *__restrict__ start, int64_t ncell, int NW, int wshift, int64_t *__restrict__ win_off) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w > NW) return;
  int64_t lo = 0, hi = ncell;
  while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if ((int)(keys[start[mid]] >> wshift) < w) lo = mid + 1; else hi = mid; }
  win_off[w] = lo;
}
}  // namespace

namespace vx {
int merge_clouds_batched(vx_ctx *ctx, const float *d_xyz, const int64_t *cloud_off, int n_clouds, const int32_t *frame_cloud, const int32_t *frame_mp,
                         const vx_pose *poses, const int32_t *wf_off, int n_windows, double leaf, DevBuf &d_out_xyz, int64_t *win_off, DevBuf *d_out48,
                         int *overflow) {
  cudaStream_t st = ctx->stream;
  const int NW = n_windows, F = wf_off[NW];
  if (overflow) *overflow = 0;
  VX_ARG(NW >= 1 && F >= NW);
  std::vector<int64_t> inst_off(F + 1, 0);
  std::vector<int32_t> fc(F), fwin(F);
  std::vector<RelPoseD> rel(F);
  for (int w = 0; w < NW; w++) {
    const vx_pose &xc = poses[wf_off[w]];
    for (int g = wf_off[w]; g < wf_off[w + 1]; g++) {
      fc[g] = frame_cloud ? frame_cloud[g] : g;
      VX_ARG(fc[g] >= 0 && fc[g] < n_clouds);
      fwin[g] = w;
      inst_off[g + 1] = inst_off[g] + (cloud_off[fc[g] + 1] - cloud_off[fc[g]]);
      const vx_pose &xi = poses[g];
      // dp = x_0.R^T (x_i.p - x_0.p), dR = x_0.R^T x_i.R  (voxelslam.cpp:2960-2961); R column-major: R(r,c) = R[c*3+r]
      const double d[3] = {xi.p[0] - xc.p[0], xi.p[1] - xc.p[1], xi.p[2] - xc.p[2]};
      for (int r = 0; r < 3; r++) {
        double sacc = 0; for (int k = 0; k < 3; k++) sacc += xc.R[r * 3 + k] * d[k];
        rel[g].dp[r] = sacc;
        for (int cc = 0; cc < 3; cc++) { double t = 0; for (int k = 0; k < 3; k++) t += xc.R[r * 3 + k] * xi.R[cc * 3 + k]; rel[g].dR[r * 3 + cc] = t; }
      }
    }
  }
  const int64_t NI = inst_off[F];
  VX_ARG(NI < ((int64_t)1 << 31));
  for (int w = 0; w <= NW; w++) win_off[w] = 0;
  if (NI == 0) return VX_OK;
  DevBuf d_coff, d_fc, d_io, d_rel, d_fw, d_mp, d_pts, d_k0, d_k1, d_i0, d_i1, d_flag, d_incl, d_start, d_bad, d_woff, d_bbox, d_rng;
  std::vector<DevBuf *> all = {&d_coff, &d_fc, &d_io, &d_rel, &d_fw, &d_mp, &d_pts, &d_k0, &d_k1, &d_i0, &d_i1, &d_flag, &d_incl, &d_start, &d_bad, &d_woff, &d_bbox, &d_rng};
  auto body = [&]() -> int {
    VX_TRY(d_coff.reserve(8 * (size_t)(n_clouds + 1))); VX_TRY(d_fc.reserve(4 * (size_t)F)); VX_TRY(d_io.reserve(8 * (size_t)(F + 1)));
    VX_TRY(d_rel.reserve(sizeof(RelPoseD) * (size_t)F)); VX_TRY(d_fw.reserve(4 * (size_t)F)); VX_TRY(d_mp.reserve(4 * (size_t)F));
    VX_TRY(d_pts.reserve(12 * (size_t)NI)); VX_TRY(d_k0.reserve(8 * (size_t)NI)); VX_TRY(d_k1.reserve(8 * (size_t)NI));
    VX_TRY(d_i0.reserve(4 * (size_t)NI)); VX_TRY(d_i1.reserve(4 * (size_t)NI)); VX_TRY(d_flag.reserve(4 * (size_t)NI)); VX_TRY(d_incl.reserve(4 * (size_t)NI));
    VX_TRY(d_start.reserve(4 * ((size_t)NI + 1))); VX_TRY(d_bad.reserve(4)); VX_TRY(d_woff.reserve(8 * (size_t)(NW + 1)));
    VX_CUDA(cudaMemcpyAsync(d_coff.p, cloud_off, 8 * (size_t)(n_clouds + 1), cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemcpyAsync(d_fc.p, fc.data(), 4 * (size_t)F, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemcpyAsync(d_io.p, inst_off.data(), 8 * (size_t)(F + 1), cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemcpyAsync(d_rel.p, rel.data(), sizeof(RelPoseD) * (size_t)F, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemcpyAsync(d_fw.p, fwin.data(), 4 * (size_t)F, cudaMemcpyHostToDevice, st));
    if (frame_mp) VX_CUDA(cudaMemcpyAsync(d_mp.p, frame_mp, 4 * (size_t)F, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemsetAsync(d_bad.p, 0, 4, st));
    // cell range of all windows -> key layout
    MergeRange r0;
    for (int a = 0; a < 3; a++) { r0.lo[a] = 0x7fffffff; r0.hi[a] = -0x7fffffff; }
    VX_TRY(d_bbox.reserve(sizeof(float) * 6 * (size_t)n_clouds)); VX_TRY(d_rng.reserve(sizeof(MergeRange)));
    VX_CUDA(cudaMemcpyAsync(d_rng.p, &r0, sizeof(r0), cudaMemcpyHostToDevice, st));
    VX_TRY(cloud_bboxes(ctx, d_xyz, d_coff.as<int64_t>(), n_clouds, d_bbox.as<float>()));
    VX_KERNEL(ctx, "k_merge_cells", k_merge_cells<<<blocks_for(F, 128), 128, 0, st>>>(d_bbox.as<float>(), d_fc.as<int32_t>(), F, d_rel.as<RelPoseD>(), leaf, d_rng.as<MergeRange>()));
    VX_CUDA(cudaMemcpyAsync(&r0, d_rng.p, sizeof(r0), cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    if (r0.lo[0] > r0.hi[0]) { for (int a = 0; a < 3; a++) { r0.lo[a] = 0; r0.hi[a] = 0; } }
    auto bits_of = [](int64_t n) { int b = 1; while (((int64_t)1 << b) < n) b++; return b; };
    MergeLayout ml;
    for (int a = 0; a < 3; a++) ml.lo[a] = r0.lo[a];
    ml.bx = bits_of((int64_t)r0.hi[0] - r0.lo[0] + 1); ml.by = bits_of((int64_t)r0.hi[1] - r0.lo[1] + 1); ml.bz = bits_of((int64_t)r0.hi[2] - r0.lo[2] + 1);
    int wbits = 1; while ((1 << wbits) < NW) wbits++;
    const int cbits = ml.bx + ml.by + ml.bz;
    if (cbits + wbits > 64) { if (overflow) *overflow = 1; set_error("merged submap clouds: the cell range needs %d + %d key bits", cbits, wbits); return VX_ERR_ARG; }
    VX_KERNEL(ctx, "k_merge_all", k_merge_all<<<blocks_for(NI, 256), 256, 0, st>>>(d_xyz, d_coff.as<int64_t>(), d_fc.as<int32_t>(), d_io.as<int64_t>(), F, d_rel.as<RelPoseD>(),
                                                                                d_fw.as<int32_t>(), leaf, ml, NI, d_pts.as<float>(), d_k0.as<uint64_t>(), d_i0.as<int32_t>(), d_bad.as<int>()));
    size_t tb = 0;
    VX_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, d_k0.as<uint64_t>(), d_k1.as<uint64_t>(), d_i0.as<int32_t>(), d_i1.as<int32_t>(), (int)NI, 0, cbits + wbits, st));
    VX_TRY(ctx->scratch.reserve(tb));
    VX_CUDA(cub::DeviceRadixSort::SortPairs(ctx->scratch.p, tb, d_k0.as<uint64_t>(), d_k1.as<uint64_t>(), d_i0.as<int32_t>(), d_i1.as<int32_t>(), (int)NI, 0, cbits + wbits, st));
    ctx->launches += 4;
    VX_KERNEL(ctx, "k_mheads", k_mheads<<<blocks_for(NI, 256), 256, 0, st>>>(d_k1.as<uint64_t>(), NI, d_flag.as<int32_t>()));
    VX_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tb, d_flag.as<int32_t>(), d_incl.as<int32_t>(), (int)NI, st));
    VX_TRY(ctx->scratch.reserve(tb));
    VX_CUDA(cub::DeviceScan::InclusiveSum(ctx->scratch.p, tb, d_flag.as<int32_t>(), d_incl.as<int32_t>(), (int)NI, st));
    ctx->launches += 2;
    VX_KERNEL(ctx, "k_mstarts", k_mstarts<<<blocks_for(NI, 256), 256, 0, st>>>(d_flag.as<int32_t>(), d_incl.as<int32_t>(), NI, d_start.as<int32_t>()));
    int32_t ncell = 0; int bad = 0;
    VX_CUDA(cudaMemcpyAsync(&ncell, d_incl.as<int32_t>() + (NI - 1), 4, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaMemcpyAsync(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    if (bad) { if (overflow) *overflow = 1; set_error("merged submap clouds: a cell index left the range its cloud's bounding box gave"); return VX_ERR_ARG; }
    VX_TRY(d_out_xyz.reserve(12 * (size_t)ncell));
    if (d_out48) VX_TRY(d_out48->reserve(sizeof(vx_point48) * (size_t)ncell));
    VX_KERNEL(ctx, "k_mmean", k_mmean<<<blocks_for(ncell, 128), 128, 0, st>>>(d_pts.as<float>(), d_i1.as<int32_t>(), d_start.as<int32_t>(), ncell, d_io.as<int64_t>(), F,
                                                                           frame_mp ? d_mp.as<int32_t>() : nullptr, d_out_xyz.as<float>(),
                                                                           d_out48 ? d_out48->as<vx_point48>() : nullptr));
    VX_KERNEL(ctx, "k_mwin_off", k_mwin_off<<<blocks_for(NW + 1, 128), 128, 0, st>>>(d_k1.as<uint64_t>(), d_start.as<int32_t>(), ncell, NW, cbits, d_woff.as<int64_t>()));
    VX_CUDA(cudaMemcpyAsync(win_off, d_woff.p, 8 * (size_t)(NW + 1), cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    return VX_OK;
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  for (DevBuf *b : all) b->release();
  return rc;
}
}  // namespace vx

extern "C" int vx_hba_merge_clouds(vx_ctx *ctx, const float *xyz, const int64_t *cloud_off, int n_clouds, const int32_t *frame_cloud,
                                   const int32_t *frame_mp, const vx_pose *poses, int W, int n_windows, double voxel_size, vx_point48 *out,
                                   int64_t out_cap, int64_t *win_off) {
  VX_ARG(ctx && cloud_off && poses && win_off && W >= 1 && n_windows >= 1 && n_clouds >= 1 && voxel_size > 0 && out_cap >= 0);
  VX_TRY(bind(ctx));
  const int64_t NP = cloud_off[n_clouds];
  VX_ARG(NP >= 0 && (NP == 0 || xyz != nullptr));
  cudaStream_t st = ctx->stream;
  DevBuf d_xyz, d_in, d_out;
  auto body = [&]() -> int {
    VX_TRY(d_xyz.reserve(sizeof(float) * 3 * std::max<int64_t>(NP, 1)));
    if (NP) VX_CUDA(cudaMemcpyAsync(d_xyz.p, xyz, sizeof(float) * 3 * NP, cudaMemcpyHostToDevice, st));
    {   // all windows in one sort / segment pass; a scene too wide for the batched key falls through to the per-window loop
      std::vector<int32_t> wf(n_windows + 1);
      for (int w = 0; w <= n_windows; w++) wf[w] = w * W;
      DevBuf d_cells, d_cells48;
      int overflow = 0;
      int rcb = merge_clouds_batched(ctx, d_xyz.as<float>(), cloud_off, n_clouds, frame_cloud, frame_mp, poses, wf.data(), n_windows, voxel_size / 8, d_cells, win_off,
                                     &d_cells48, &overflow);
      if (rcb == VX_OK) {
        if (win_off[n_windows] > out_cap) { rcb = VX_ERR_ARG; set_error("vx_hba_merge_clouds: output capacity %lld too small", (long long)out_cap); }
        else if (win_off[n_windows] > 0) {
          const cudaError_t ce = cudaMemcpyAsync(out, d_cells48.p, sizeof(vx_point48) * (size_t)win_off[n_windows], cudaMemcpyDeviceToHost, st);
          cudaStreamSynchronize(st);
          if (ce != cudaSuccess) rcb = VX_ERR_CUDA;
        }
      }
      d_cells.release(); d_cells48.release();
      if (rcb == VX_OK || !overflow) return rcb;
    }
    int64_t written = 0;
    win_off[0] = 0;
    for (int w = 0; w < n_windows; w++) {
      std::vector<int> cidx(W);
      for (int i = 0; i < W; i++) { const int c = frame_cloud ? frame_cloud[w * W + i] : w * W + i; VX_ARG(c >= 0 && c < n_clouds); cidx[i] = c; }
      int32_t ncell = 0;
      // every keyframe cloud into the frame of keyframe 0 (:2958-2969), down_sampling_voxel(pl, voxel_size / 8) (:2978)
      VX_TRY(assemble_window(ctx, d_xyz.as<float>(), cloud_off, cidx.data(), poses + (size_t)w * W, frame_mp ? frame_mp + (size_t)w * W : nullptr, W, 0,
                             voxel_size / 8, d_in, d_out, &ncell));
      if (ncell > 0) {
        if (written + ncell > out_cap) { set_error("vx_hba_merge_clouds: output capacity %lld too small", (long long)out_cap); return VX_ERR_ARG; }
        VX_CUDA(cudaMemcpyAsync(out + written, d_out.p, sizeof(vx_point48) * ncell, cudaMemcpyDeviceToHost, st));
        written += ncell;
      }
      win_off[w + 1] = written;
    }
    VX_CUDA(cudaStreamSynchronize(st));
    return VX_OK;
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  d_xyz.release(); d_in.release(); d_out.release();
  return rc;
}

// Keyframe assembly of a stored session (voxelslam.cpp:345-378; the online twin :2355-2385 uses the same relative poses): the
// `win` scans of a keyframe moved into the frame of the LAST scan's pose, then down_sampling_voxel(pl_lc, voxel_size / 10).
extern "C" int vx_keyframe_assemble(vx_ctx *ctx, const float *xyz, const int64_t *cloud_off, int win, const vx_pose *poses, double voxel_size,
                                    vx_point48 *out, int64_t out_cap, int64_t *n_out) {
  VX_ARG(ctx && cloud_off && poses && n_out && win >= 1 && voxel_size > 0 && out_cap >= 0);
  VX_TRY(bind(ctx));
  const int64_t NP = cloud_off[win];
  VX_ARG(NP >= 0 && (NP == 0 || xyz != nullptr));
  cudaStream_t st = ctx->stream;
  DevBuf d_xyz, d_in, d_out;
  *n_out = 0;
  auto body = [&]() -> int {
    VX_TRY(d_xyz.reserve(sizeof(float) * 3 * std::max<int64_t>(NP, 1)));
    if (NP) VX_CUDA(cudaMemcpyAsync(d_xyz.p, xyz, sizeof(float) * 3 * NP, cudaMemcpyHostToDevice, st));
    std::vector<int> cidx(win);
    for (int i = 0; i < win; i++) cidx[i] = i;
    int32_t ncell = 0;
    VX_TRY(assemble_window(ctx, d_xyz.as<float>(), cloud_off, cidx.data(), poses, nullptr, win, win - 1, voxel_size / 10, d_in, d_out, &ncell));
    if (ncell > out_cap) { set_error("vx_keyframe_assemble: output capacity %lld too small for %d cells", (long long)out_cap, ncell); return VX_ERR_ARG; }
    if (ncell > 0) VX_CUDA(cudaMemcpyAsync(out, d_out.p, sizeof(vx_point48) * ncell, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    *n_out = ncell;
    return VX_OK;
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  d_xyz.release(); d_in.release(); d_out.release();
  return rc;
}
