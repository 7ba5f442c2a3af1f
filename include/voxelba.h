/*
 * voxelba.h — C ABI of libvoxelba: the B200 (sm_100a) implementation of Voxel-SLAM's voxel-map
 * bundle-adjustment / point-to-plane hot path.
 *
 * The reference has no FFI or plugin API for this path: it is reached by direct
 * C++ calls on header-defined classes.  Each entry point below therefore names the reference
 * class member / free function it replaces (file:line into /root/reference/VoxelSLAM/src/).
 * `include/voxelba_compat.hpp` re-creates the reference's C++ names on top of this ABI.
 *
 * Conventions
 *   - all matrices are COLUMN-MAJOR doubles (Eigen's default, which the reference uses);
 *   - POD structs mirror the reference's own structs byte-for-byte where one exists;
 *   - every function returns VX_OK (0) or a negative VX_ERR_* code; vx_last_error() gives text;
 *   - calls are synchronous unless the name ends in `_async`;
 *   - there is NO CPU fallback: without a CUDA device every compute call returns VX_ERR_CUDA.
 */
#ifndef VOXELBA_H
#define VOXELBA_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VX_OK 0
#define VX_ERR_CUDA (-1)      /* CUDA runtime failure or no device */
#define VX_ERR_ARG (-2)       /* bad argument */
#define VX_ERR_TOO_FEW (-3)   /* reference: printf("Too Less Voxel"); exit(0)   voxel_map.hpp:399-403 */
#define VX_ERR_OPT_STATE (-4) /* reference: printf("Error: opt_state"); exit(0)  voxel_map.hpp:1488-1492 */
#define VX_ERR_NOMEM (-5)
#define VX_ERR_NCCL (-6)      /* NCCL missing or a collective failed (multi-GPU entry points only) */

typedef struct vx_ctx vx_ctx;       /* one per host thread: device, stream, scratch */
typedef struct vx_comm vx_comm;     /* one per (process, GPU): NCCL communicator over the GPUs that share a problem */
typedef struct vx_factor vx_factor; /* class LidarFactor (a.k.a. VOX_HESS), voxel_map.hpp:124-339 */

/* class PointCluster {Matrix3d P; Vector3d v; int N;}  tools.hpp:304-310 (104 bytes, 8-aligned) */
typedef struct vx_cluster {
  double P[9];
  double v[3];
  int32_t N;
  int32_t _pad;
} vx_cluster;

/* the part of struct IMUST the LiDAR factor reads: R (tools.hpp:139) and p (tools.hpp:140) */
typedef struct vx_pose {
  double R[9];
  double p[3];
} vx_pose;

/* pcl::PointXYZINormal as the reference stores it (48 bytes: data[4], normal[4], intensity, curvature, pad) */
typedef struct vx_point48 {
  float x, y, z, _p0;
  float normal_x, normal_y, normal_z, _p1;
  float intensity, curvature, _p2, _p3;
} vx_point48;
/* struct pointVar  voxel_map.hpp:18-34: point + 3x3 covariance (column-major) */
typedef struct vx_point_var { double pnt[3]; double var[9]; } vx_point_var;

/* ---- context ------------------------------------------------------------------------------ */
int vx_ctx_create(int device, vx_ctx **out);
int vx_ctx_destroy(vx_ctx *ctx);
int vx_ctx_sync(vx_ctx *ctx);
void *vx_ctx_stream(vx_ctx *ctx);          /* the cudaStream_t every launch of this ctx uses */
const char *vx_last_error(void);           /* thread-local text of the last failure */
int vx_version(void);
/* number of kernels this library launched on `ctx` since creation (bench.py: gpu_launches) */
int64_t vx_ctx_launch_count(vx_ctx *ctx);

/* Per-kernel device timing for bench.py's roofline leg (the reference brackets the same stages with
 * ros::Time, voxel_map.hpp:639-650).  While enabled, every kernel this library launches on `ctx` is
 * bracketed by two CUDA events on the ctx stream.  profile_read synchronises, then returns one
 * entry per kernel name: newline-separated names, summed milliseconds and launch counts. */
int vx_ctx_profile(vx_ctx *ctx, int enable);   /* also clears what was recorded */
int vx_ctx_profile_read(vx_ctx *ctx, char *names, size_t names_cap, double *ms, int64_t *counts,
                        int max_entries, int *n_out);

/* FP64 pipe peak of the device, measured in place (MEASURED_PEAKS.json has no FP64 figure):
 * mode 0 = DFMA on the CUDA cores, mode 1 = mma.sync m8n8k4 f64 on the tensor pipe, mode 2 = half the warps of each
 * (do the two share a datapath?).  TFLOP/s. */
int vx_bench_fp64(vx_ctx *ctx, int mode, double *tflops);
/* One warp, `ilp` (1, 2, 4, 8) independent dependent chains of DFMA (mma = 0) or DMMA m8n8k4 (mma = 1): clock cycles per
 * instruction issued; ilp = 1 is the dependent-issue latency that sizes the parallelism phase A of B1 needs. */
int vx_bench_fp64_latency(vx_ctx *ctx, int mma, int ilp, double *cycles_per_instr);

/* ---- B0: LidarFactor storage -------------------------------------------------------------- */
/* LidarFactor::LidarFactor(int _w)  voxel_map.hpp:136.  `n_windows` > 1 makes a BATCH of
 * independent factors that share W (bottom-level GBA windows, voxelslam.cpp:3096): frame f of
 * window b is global frame b*W+f, outputs are n_windows separate Hess/JacT/cost. */
int vx_factor_create(vx_ctx *ctx, int win_size, int n_windows, vx_factor **out);
int vx_factor_free(vx_factor *f);
/* LidarFactor::clear()  voxel_map.hpp:328-336 */
int vx_factor_clear(vx_factor *f);
/* LidarFactor::push_voxel() x V  voxel_map.hpp:139-147.  Dense form = the reference's own layout:
 * plvec[V][W] (N==0 marks an absent frame), sig_vecs[V], coeffs[V], eig_values[V][3],
 * eig_vectors[V][9] col-major, pcr_adds[V].  Replaces the factor's content. */
int vx_factor_upload_dense(vx_factor *f, int64_t V, const vx_cluster *plvec, const vx_cluster *sig_vecs,
                           const double *coeffs, const double *eig_values, const double *eig_vectors,
                           const vx_cluster *pcr_adds);
/* Same content in compressed form for large W (top-level GBA, W~399: >95% of plvec is N==0):
 * voxel a owns entries [offsets[a], offsets[a+1]) with ascending `frames` (global frame ids)
 * and their clusters. */
int vx_factor_upload_csr(vx_factor *f, int64_t V, const int64_t *offsets, const int32_t *frames,
                         const vx_cluster *clusters, const vx_cluster *sig_vecs, const double *coeffs,
                         const double *eig_values, const double *eig_vectors, const vx_cluster *pcr_adds);
/* read back what evaluate_only_residual caches  voxel_map.hpp:317-319 */
int vx_factor_download_cache(vx_factor *f, double *eig_values, double *eig_vectors, vx_cluster *pcr_adds);
/* read back the factor content in the vx_factor_upload_csr form (offsets V+1, frames/clusters nnz, coeffs V);
 * any pointer may be NULL.  The reference reads these members directly (plvec_voxels, coeffs). */
int vx_factor_download_csr(vx_factor *f, int64_t *offsets, int32_t *frames, vx_cluster *clusters, double *coeffs);
/* Evaluation path selection.  acc_evaluate2 has two device formulations: a fused tensor-core kernel for
 * the window sizes the reference runs (W = 10, 20) and a generic pair-list path for any W (top-level
 * GBA, W ~ 399).  on != 0 forces the generic path (tests compare the two). */
int vx_factor_force_generic(vx_factor *f, int on);
int64_t vx_factor_num_voxels(const vx_factor *f);
int64_t vx_factor_num_voxel_frames(const vx_factor *f); /* sum over voxels of frames with N != 0 */
int vx_factor_win_size(const vx_factor *f);
int vx_factor_num_windows(const vx_factor *f);

/* ---- B1 / B2: evaluation -------------------------------------------------------------------- */
/* LidarFactor::acc_evaluate2(xs, head, end, Hess, JacT, residual)  voxel_map.hpp:150-282.
 * poses: n_windows*W entries.  Hess: n_windows x (6W x 6W) col-major, JacT: n_windows x 6W,
 * residual: n_windows.  Reads the cached eig_values / eig_vectors / pcr_adds (voxel_map.hpp:187-190). */
int vx_factor_hessian(vx_ctx *ctx, vx_factor *f, const vx_pose *poses, int64_t head, int64_t end,
                      double *Hess, double *JacT, double *residual);
/* LidarFactor::evaluate_only_residual(xs, head, end, residual)  voxel_map.hpp:285-325.
 * Refreshes the cache for voxels in [head,end) (voxel_map.hpp:317-319). */
int vx_factor_residual(vx_ctx *ctx, vx_factor *f, const vx_pose *poses, int64_t head, int64_t end,
                       double *residual);

/* Device-resident variants: no host copies, no synchronisation.  Results stay in the factor's
 * device buffers (vx_factor_device_buffers) so a collective or the next kernel can consume them. */
int vx_factor_set_poses(vx_ctx *ctx, vx_factor *f, const vx_pose *poses);        /* H2D, async */
int vx_factor_hessian_async(vx_ctx *ctx, vx_factor *f, int64_t head, int64_t end);
int vx_factor_residual_async(vx_ctx *ctx, vx_factor *f, int64_t head, int64_t end);
/* device pointers: Hess (n_windows*(6W)^2), JacT (n_windows*6W), cost_hess (n_windows),
 * cost_resid (n_windows); any out pointer may be NULL */
int vx_factor_device_buffers(vx_factor *f, void **Hess, void **JacT, void **cost_hess, void **cost_resid);

/* ---- A1: the voxel key rule, inlined 7x in the reference (voxel_map.hpp:1907-1918, :1977-1990, :2114-2125,
 * :2173-2184, :42-52; tools.hpp:206-216; loop_refine.hpp:452-459).  xyz: n x 3 doubles, or floats when
 * is_float_input (tools.hpp:211 promotes a float coordinate).  out_keys: n x 3 int64 (VOXEL_LOC x,y,z). */
int vx_keys(vx_ctx *ctx, const void *xyz, int64_t n, double voxel_size, int is_float_input, int64_t *out_keys);

/* vx_keys that allocates nothing.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernel, same
 * arguments and results as vx_keys, but the device staging (coordinates in, keys out) is a fixed file-scope device array of
 * VX_KEYS_STATIC_MAX_POINTS points (12 MiB, held by every GPU that loads the library until exit) instead of the context's
 * grow-on-demand scratch; no cudaMalloc / cudaFree / allocating library call is made.  n > VX_KEYS_STATIC_MAX_POINTS ->
 * VX_ERR_ARG (vx_last_error names the limit).  All calls on one GPU share the array, which keeps the last call's contents:
 * calls that overlap on the same GPU (two contexts / threads) race — serialise them. */
#define VX_KEYS_STATIC_MAX_POINTS 262144
int vx_keys_static(vx_ctx *ctx, const void *xyz, int64_t n, double voxel_size, int is_float_input, int64_t *out_keys);

/* ---- A2: void down_sampling_voxel(pcl::PointCloud<PointType>&, double)  tools.hpp:201-238.
 * out must hold n points; *out_n = occupied cells.  Per cell the reference's sequential float running
 * mean is replayed exactly (bit-identical x,y,z and curvature = count; other fields of the cell's first
 * point kept); cells come out in ascending packed-key order (the reference: unordered_map order). */
int vx_downsample_voxel(vx_ctx *ctx, const vx_point48 *pts, int64_t n, double voxel_size, vx_point48 *out, int64_t *out_n);
/* A4 down_sampling_close(pl, voxel_size)  tools.hpp:240-302: per cell the float centroid (sequential sums), then the INPUT
 * point nearest to it is kept whole (timestamps in `curvature` survive for de-skew).  Same ordering note as above. */
int vx_downsample_close(vx_ctx *ctx, const vx_point48 *pts, int64_t n, double voxel_size, vx_point48 *out, int64_t *out_n);
/* A3 down_sampling_pvec(pvec, voxel_size, pl_keep)  voxel_map.hpp:39-81: cells keyed on the DOUBLE point, running double mean
 * (x*c + p)/(c+1) of point and covariance; out = centroid as float xyz, diag(var) in normal_xyz, other fields as a
 * default-constructed pcl::PointXYZINormal (data[3] = 1, rest 0). */
int vx_downsample_pvec(vx_ctx *ctx, const vx_point_var *pvec, int64_t n, double voxel_size, vx_point48 *out, int64_t *out_n);

/* ---- A8 + A10: OctreeGBA::cut_voxel per keyframe (loop_refine.hpp:446-479) + OctreeGBA_multi_recut
 * (loop_refine.hpp:483-537) -> LidarFactor, as HBA_add_edge does at voxelslam.cpp:2887-2892.
 * The thresholds are the reference's globals gba_voxel_size / gba_min_eigen_value / gba_eigen_value_array
 * (loop_refine.hpp:269-271; the array is stored as 1/yaml, voxelslam.cpp:3023-3024) and max_layer. */
typedef struct vx_gba_params {
  double voxel_size;
  double min_eigen_value;
  double eigen_value_array[4];
  int32_t max_layer;   /* 0..3 */
  int32_t _pad;
} vx_gba_params;
/* xyz: local-frame keyframe clouds, float x,y,z triples, cloud c = points [cloud_off[c], cloud_off[c+1]).
 * Frame g (0 <= g < n_windows*W of `f`, window-major) uses cloud frame_cloud[g] (NULL: g itself) and
 * poses[g]; overlapping windows (stride 5, size 10: voxelslam.cpp:3137-3138) reference one cloud twice.
 * Replaces the content of `f` (cache = the recut-time eig_value / eig_vector / pcr_add, as tras_opt /
 * push_voxel leave it).  Voxels are ordered by (window, root cell, octree path). */
int vx_gba_build(vx_ctx *ctx, const float *xyz, const int64_t *cloud_off, int n_clouds, const int32_t *frame_cloud,
                 const vx_pose *poses, const vx_gba_params *params, vx_factor *f);

/* Multi-GPU voxelization (one exchange step): every rank keys its slice of points and routes each to the rank that owns its
 * ROOT voxel, owner = mix64(packed root key) mod n_ranks, so that a root voxel — and therefore every (leaf, frame) cluster
 * under it — lives on exactly one rank; sort / segment / moments / recut are then local and Hess is a sum over ranks.
 * d_* are DEVICE pointers (the all-to-all between the two calls is NCCL's, on the same buffers); poses / cloud_off are host. */
int vx_gba_point_owner(vx_ctx *ctx, const float *d_xyz, const int32_t *d_frame, int64_t n, const vx_pose *poses, int n_frames,
                       double voxel_size, int n_ranks, int32_t *d_owner);
int vx_gba_build_device(vx_ctx *ctx, const float *d_xyz, const int64_t *cloud_off, int n_clouds, const int32_t *frame_cloud,
                        const vx_pose *poses, const vx_gba_params *params, vx_factor *f);

/* ---- multi-GPU: one process per GPU, NCCL over NVLink / NVSwitch ------------------------------------------
 * The reference parallelises acc_evaluate2 / evaluate_only_residual over host threads and adds the per-thread partial
 * Hess / JacT / residual (voxel_map.hpp:373-386); across GPUs the same sum is ONE all-reduce of
 * [lower triangle of Hess | 15 W upper entries of the diagonal blocks | JacT | cost] per evaluation.  Voxelization across GPUs
 * is one exchange: each point goes to the rank that owns its ROOT voxel (vx_gba_build_sharded).
 * Bootstrap: rank 0 calls vx_comm_unique_id and hands the 128 bytes to the other ranks by whatever the host program has
 * (MPI_Bcast, a file, torch.distributed's store); every rank then calls vx_comm_create on its own context.  NCCL is bound at
 * run time (libnccl.so.2), so single-GPU use needs no NCCL at all. */
#define VX_COMM_ID_BYTES 128
int vx_comm_unique_id(void *id128);                                               /* ncclGetUniqueId */
int vx_comm_create(vx_ctx *ctx, const void *id128, int rank, int n_ranks, vx_comm **out);   /* ncclCommInitRank; n_ranks == 1 needs no id */
int vx_comm_free(vx_comm *comm);
int vx_comm_rank(const vx_comm *comm);
int vx_comm_size(const vx_comm *comm);
int vx_comm_nccl_version(void);                                                   /* 0 when NCCL is unavailable */
int vx_comm_allreduce_f64(vx_comm *comm, void *d_buf, int64_t n);                 /* in-place SUM of n device doubles, async on the ctx stream */
/* bytes all-reduced / exchanged point-to-point and collectives issued on this communicator so far (bench.py reports them) */
int vx_comm_stats(const vx_comm *comm, int64_t *bytes_allreduced, int64_t *bytes_exchanged, int64_t *n_collectives);

/* OctreeGBA::cut_voxel + OctreeGBA_multi_recut (loop_refine.hpp:446-537) with the ROOT voxels partitioned over the ranks:
 * xyz / cloud_off describe THIS rank's clouds (host), cloud c being the points of global frame cloud_frame[c] (one frame per
 * cloud); poses holds ALL F = n_windows * W frames of `f` on every rank.  Each point is keyed at its pose, routed to
 * owner = mix64(packed root key) mod n_ranks by a stable bucket partition + one grouped ncclSend/ncclRecv round (16 B/point),
 * and the receiver builds the LidarFactor of the root voxels it owns: per (leaf, frame) moments are bitwise the single-GPU
 * ones; Hess / JacT / cost of the whole problem are sums over ranks (vx_ba_lidar_sharded).  *n_points_owned: points received. */
int vx_gba_build_sharded(vx_ctx *ctx, vx_comm *comm, const float *xyz, const int64_t *cloud_off, int n_clouds, const int32_t *cloud_frame,
                         const vx_pose *poses, const vx_gba_params *params, vx_factor *f, int64_t *n_points_owned);
/* vx_ba_lidar on a factor whose voxels are sharded over the ranks (voxel_map.hpp:422-497 with the thread sum :382-384 taken
 * across GPUs): every rank passes the same poses and gets the same result; the LM solve is replicated. */
int vx_ba_lidar_sharded(vx_ctx *ctx, vx_comm *comm, vx_factor *f, vx_pose *poses, double *hess_out, double *resis, int max_iter, int thd_num,
                        int *converged, int *accept_trace, int *n_iters);

/* ---- B3: Lidar_BA_Optimizer::damping_iter(x_stats, voxhess, hess*, resis, max_iter, is_display)
 * voxel_map.hpp:422-497.  Levenberg-Marquardt on 6W unknowns per window: u=0.01, v=2 (:427), *hess
 * snapshot before the gauge fix (:446), pose 0 fixed (:452-455), (Hess + u*diag(Hess)).ldlt().solve(-JacT)
 * on the LOWER triangle (:457-458), retraction R*Exp(dtheta), p+dp (:460-464), gain ratio and damping
 * update (:465-490), stop at |dr/r| < 1e-6 (:492).  The factor's cache follows the reference's stateful
 * semantics (after a rejected step it holds the rejected trial).
 *   poses      in/out, n_windows*W
 *   hess_out   n_windows x (6W x 6W) col-major, may be NULL        (the reference's MatrixXd* hess)
 *   resis      n_windows x 2: first and last cost                    (the reference's vector<double>& resis)
 *   thd_num    only reproduces the reference's "Too Less Voxel" exit (:399-403) -> VX_ERR_TOO_FEW
 *   converged  n_windows: the reference's bool return (false if any step was rejected), may be NULL
 *   accept_trace  n_windows x max_iter, 1 accepted / 0 rejected / untouched when not run, may be NULL
 *   n_iters    n_windows: LM iterations actually run, may be NULL */
int vx_ba_lidar(vx_ctx *ctx, vx_factor *f, vx_pose *poses, double *hess_out, double *resis, int max_iter,
                int thd_num, int *converged, int *accept_trace, int *n_iters);

/* ---- B6: class IMU_PRE  preintegration.hpp:11-303 — host-side, a POD the caller owns.
 * 3x3 / 15x15 members are column-major.  noise = the reference's globals imupre_scale_gravity, noiseMeas, noiseWalk
 * (preintegration.hpp:8-9; filled at voxelslam.cpp:934-939). */
typedef struct vx_imu_pre {
  double R_delta[9], p_delta[3], v_delta[3], bg[3], ba[3];
  double R_bg[9], p_bg[9], p_ba[9], v_bg[9], v_ba[9];
  double dtime;
  double dbg[3], dba[3], dbg_buf[3], dba_buf[3];
  double cov[225];
} vx_imu_pre;
typedef struct vx_imu_noise { double scale_gravity; double noise_meas[36]; double noise_walk[36]; } vx_imu_noise;
typedef struct vx_imu_sample { double t, gyr[3], acc[3]; } vx_imu_sample;
/* the IMUST members the LI optimizers read and write (tools.hpp:135-199): R, p, v, bg, ba, g */
typedef struct vx_imu_state { double R[9], p[3], v[3], bg[3], ba[3], g[3]; } vx_imu_state;

int vx_imu_pre_init(vx_imu_pre *pre, const double *bg3, const double *ba3);                 /* ctor, preintegration.hpp:32-48 */
int vx_imu_pre_push(vx_imu_pre *pre, const vx_imu_noise *nz, const vx_imu_sample *s, int n); /* push_imu :50-73 (mid-point rule) */
int vx_imu_pre_add(vx_imu_pre *pre, const vx_imu_noise *nz, const double *gyr3, const double *acc3, double dt); /* add_imu :75-135 */
/* give_evaluate (:137-212, with_gravity=0: 30 columns) / give_evaluate_g (:214-294, with_gravity=1: 33 columns).
 * jtj (nc x nc col-major) and gg (nc) are written when both are non-NULL (the reference's jac_enable); *cost = rr^T cov^-1 rr. */
int vx_imu_pre_evaluate(const vx_imu_pre *pre, const vx_imu_state *st1, const vx_imu_state *st2, int with_gravity, double *jtj,
                        double *gg, double *cost);
/* The same factor evaluation on the device for a CHAIN of n_factors factors (factor i between states[i] and states[i+1]), the
 * kernel vx_ba_li / vx_ba_li_gravity evaluate the IMU chain with: jtj n_factors x (nc x nc), gg n_factors x nc
 * (both may be NULL: costs only), cost n_factors; nc = 30 or 33 (with_gravity). */
int vx_imu_pre_evaluate_device(vx_ctx *ctx, const vx_imu_pre *pres, const vx_imu_state *states, int n_factors, int with_gravity, double *jtj,
                               double *gg, double *cost);
int vx_imu_pre_update_state(vx_imu_pre *pre, const double *dxi15);                          /* :296-303 */

/* ---- B4: LI_BA_Optimizer::damping_iter(x_stats, voxhess, imus_factor, hess*)  voxel_map.hpp:624-713.
 * 15W unknowns [theta p v bg ba]; IMU chain 30x30 blocks on the host, scaled by imu_coef (:565-567); LiDAR 6x6 blocks
 * from the device factor scattered stride 6 -> 15 (hess_plus :509-517); gauge fixes all 15 of pose 0 (:653-656); 3 fixed
 * iterations (:643); bias linearisation shifted by update_state and rolled back on reject (:670-671, :701-705).
 *   x          in/out, W states              imus   W-1 factors, in/out (dbg/dba move)
 *   hess_out   (15W)^2 col-major, may be NULL      accept_trace  3 ints, may be NULL     n_iters may be NULL
 * The factor must be a single window (n_windows == 1). */
int vx_ba_li(vx_ctx *ctx, vx_factor *f, vx_imu_state *x, vx_imu_pre *imus, double imu_coef, double *hess_out,
             int *accept_trace, int *n_iters);
/* ---- B5: LI_BA_OptimizerGravity::damping_iter(x_stats, voxhess, imus_factor, resis, hess*, max_iter=2)
 * voxel_map.hpp:878-975.  15W+3 unknowns (gravity last, :921); 33x33 factor blocks scattered into pose-pose / pose-g /
 * g-pose / g-g (:785-795); gauge fixes the first 6 only (:906-909); g += dxi.tail(3) copied to every frame (:921-930).
 *   resis2: first and last cost;  hess_out (15W+3)^2. */
int vx_ba_li_gravity(vx_ctx *ctx, vx_factor *f, vx_imu_state *x, vx_imu_pre *imus, double imu_coef, double *resis2,
                     double *hess_out, int max_iter, int *accept_trace, int *n_iters);

/* (Hess + u*D).ldlt().solve(b) the way vx_ba_lidar solves large windows (>= 384 unknowns): Eigen's diagonal-pivot order from
 * the input diagonal, blocked right-looking LDL^T of the permuted LOWER triangle and the substitution on the device.
 * A: n x n column-major host matrix (only its lower triangle is read). */
int vx_ldlt_solve(vx_ctx *ctx, const double *A, int n, const double *b, double *x);

/* ---- stored sessions (host code, no device needed) and the keyframe assembly ----------------------
 * One scan of alidarState.txt: IMUST t, p, R (column-major), v, bg, ba, g and ScanPose::v6 (voxelslam.cpp:181-204). */
typedef struct vx_scan_state { double t, p[3], R[9], v[3], bg[3], ba[3], g[3], v6[6]; } vx_scan_state;
/* Eigen's Quaterniond::matrix() and Quaterniond(Matrix3d) as the formats use them; q = x y z w. */
void vx_quat_to_R(const double *q_xyzw, double *R9);
void vx_R_to_quat(const double *R9, double *q_xyzw);
/* FileReaderWriter::save_pose voxelslam.cpp:181-204 (the caller keeps the reference's "fewer than 100 scans: no file" rule) */
int vx_session_write_lidarstate(const char *path, const vx_scan_state *xs, int64_t n);
/* read_lidarstate voxelslam.hpp:268-306: 8, 20 or 26 single-space separated numbers per line; *n = scans in the file */
int vx_session_read_lidarstate(const char *path, vx_scan_state *out, int64_t cap, int64_t *n);
/* FileReaderWriter::save_pcd voxelslam.cpp:166-179: PCD v0.7, DATA binary, FIELDS x y z intensity; xyzi: n x 4 floats */
int vx_session_write_pcd_xyzi(const char *path, const float *xyzi, int64_t n);
/* pcl::io::loadPCDFile into PointXYZI (voxelslam.cpp:340): fields by name, DATA ascii | binary; *n = points in the file */
int vx_session_read_pcd_xyzi(const char *path, float *xyzi, int64_t cap, int64_t *n);
/* vx_pgo_edge is declared further down */
struct vx_pgo_edge;
/* FileReaderWriter::pgo_edges_io voxelslam.cpp:207-279, write (io = 1) and read (io = 0) sides */
int vx_session_write_edges(const char *path, const char *const *fnames, int n_names, const struct vx_pgo_edge *edges, int64_t n, const char *absent);
int vx_session_read_edges(const char *path, const char *const *fnames, int n_names, const char *bagname, struct vx_pgo_edge *edges, int64_t cap,
                          int64_t *n, char *absent_buf, int64_t absent_cap, int64_t *absent_len);
/* Keyframe assembly of a stored session (voxelslam.cpp:345-378; online twin :2355-2385): the `win` scans (cloud_off[win + 1]
 * offsets into xyz) moved into the frame of the LAST scan's pose, then down_sampling_voxel(pl_lc, voxel_size / 10), on the device.
 * out: cells in ascending packed-key order; *n_out = cells. */
int vx_keyframe_assemble(vx_ctx *ctx, const float *xyz, const int64_t *cloud_off, int win, const vx_pose *poses, double voxel_size,
                         vx_point48 *out, int64_t out_cap, int64_t *n_out);

/* ---- bool icp_normal(pl_src, pl_tar, pose, icp_eigval)   loop_refine.hpp:47-145
 * Point-to-plane ICP with normal-compatibility gates, the geometric check of a loop-closure candidate.  src / tar: n x 3 float
 * points and n x 3 float normals (PointXYZINormal's x,y,z / normal_x,y,z).  pose (R, p) is the reference's pair<Vector3d,
 * Matrix3d>, updated in place.  Nearest neighbour = exact 1-NN under FLANN's float squared distance (lowest target index on
 * ties).  *accepted = the reference's return value (smallest eigenvalue of sum n n^T > icp_eigval and converged).
 * match_nums20 / resis20: per-iteration match count and 0.5 * sum r^2 (<= 20 entries, may be NULL); eig3: ascending. */
int vx_icp_normal(vx_ctx *ctx, const float *src_xyz, const float *src_nrm, int64_t ns, const float *tar_xyz, const float *tar_nrm,
                  int64_t nt, vx_pose *pose, double icp_eigval, int32_t *accepted, int32_t *iters, int32_t *match_nums20,
                  double *resis20, double *eig3, int32_t *is_converge);

/* STDescManager::init_voxel_map + BTCOctoTree::init_plane + STDescManager::get_plane (BTC.cpp:279-334, :90-143): the plane cloud
 * the loop-closure thread hands to icp_normal.  xyz: n x 3 floats.  A voxel (BTC's own key rule: double quotient, truncated)
 * with more than init_num points whose smallest covariance eigenvalue is below plane_thre yields one PointXYZINormal (center,
 * normal; the sign of the normal is as arbitrary as Eigen::EigenSolver's).  aux (may be NULL): per plane center 3 | normal 3 |
 * min_eigen_value | radius | d | points_size (BTCPlane).  Planes in ascending voxel-key order (the reference: hash-map order). */
int vx_btc_plane_cloud(vx_ctx *ctx, const float *xyz, int64_t n, double voxel_size, double plane_thre, int init_num, vx_point48 *out,
                       double *aux, int64_t out_cap, int64_t *n_out);

/* ---- C1h: VOXEL_SLAM::HBA_add_edge(p_xs, p_smps, gba_edges, maps, max_iter, thread_num, plptr)
 * voxelslam.cpp:2822-3015 (the session filter :2836-2857 stays with the caller: pass only connected frames).
 * For every window: up to max_iter outer iterations of {rebuild the GBA voxel map at the current poses,
 * damping_iter with up = 4}; the last / first-converged pass swaps the GBA thresholds for the LocalBA ones
 * (:2873-2883); then pose-graph edges for all i < j with cov_k = 1/|hess(6i+k, 6j+k)|, dropped when any
 * |.| < 1e-6 (:2926-2951).  n_windows > 1 = independent windows in one call (bottom level, :3096). */
typedef struct vx_hba_params {
  vx_gba_params gba;                 /* GBA/voxel_size, GBA/min_eigen_value, 1/GBA/eigen_value_array, LocalBA/max_layer */
  double voxel_size;                 /* Odometry/voxel_size          (fine pass) */
  double min_eigen_value;            /* Odometry/min_eigen_value     (fine pass) */
  double plane_eigen_value_thre[4];  /* 1/LocalBA/plane_eigen_value_thre (fine pass) */
} vx_hba_params;
/* PGO_Edge payload pushed by PGO_Edges::push (loop_refine.hpp:163-235) */
typedef struct vx_pgo_edge {
  int32_t m1, m2, id1, id2;
  double rot[9];   /* R_i^T R_j, column-major */
  double tra[3];   /* R_i^T (p_j - p_i) */
  double cov[6];
} vx_pgo_edge;
/* frame_id / frame_mp: Keyframe::id / Keyframe::mp per frame (may be NULL).  edges: n_windows * W(W-1)/2 slots,
 * window b writes n_edges[b] edges from slot b*W(W-1)/2.  hess_out: n_windows x (6W)^2 (may be NULL),
 * resis_out: n_windows x 2 of the last damping_iter, n_voxels_out / status_out per window (status
 * VX_ERR_TOO_FEW where the reference prints "Too Less Voxel" and exits). */
int vx_hba_add_edge(vx_ctx *ctx, const float *xyz, const int64_t *cloud_off, int n_clouds, const int32_t *frame_cloud,
                    const int32_t *frame_id, const int32_t *frame_mp, vx_pose *poses, int W, int n_windows,
                    const vx_hba_params *prm, int max_iter, int thread_num, vx_pgo_edge *edges, int64_t *n_edges,
                    double *hess_out, double *resis_out, int32_t *n_voxels_out, int32_t *status_out);

/* The merged submap cloud HBA_add_edge returns through plptr (voxelslam.cpp:2954-2985): every keyframe cloud of a window moved
 * into the frame of the window's first keyframe with the (optimized) poses, intensity = the keyframe's session id (frame_mp),
 * then down_sampling_voxel(pl, voxel_size / 8).  Windows are written back to back, window w = out[win_off[w], win_off[w+1]);
 * cells in ascending packed-key order inside a window. */
int vx_hba_merge_clouds(vx_ctx *ctx, const float *xyz, const int64_t *cloud_off, int n_clouds, const int32_t *frame_cloud,
                        const int32_t *frame_mp, const vx_pose *poses, int W, int n_windows, double voxel_size, vx_point48 *out,
                        int64_t out_cap, int64_t *win_off);

/* ---- C2h: the bottom-up schedule of VOXEL_SLAM::thd_globalmapping (voxelslam.cpp:3018-3141) --------------------------------
 * Keyframes arrive session by session; a bottom window opens every `mgsize` keyframes and holds `wdsize` of them (:3033-3034,
 * :3076-3082, :3135-3138).  When a session ends (:3127-3133) or the global pass fires (:3054-3058) the keyframes still in the
 * local queue (mgsize .. wdsize-1 of them) form one more, shorter window.  session_sizes: keyframes per session.
 * Outputs (cap entries each): window w covers keyframes [win_start[w], win_start[w] + win_size[w]) of session win_session[w]
 * (win_start counts inside the session). */
int vx_hba_schedule(const int32_t *session_sizes, int n_sessions, int wdsize, int mgsize, int32_t *win_session, int32_t *win_start,
                    int32_t *win_size, int cap, int32_t *n_windows);

/* Which windows of the schedule rank `rank` of `n_ranks` optimises in vx_hba_global: the contiguous run [*lo, *hi), balanced by
 * cost (a window shorter than wdsize is a batch of its own and weighs as much as 16 full ones).  The reference optimises its
 * windows one after another on one host thread (voxelslam.cpp:3094-3096); this is the multi-GPU partition of that loop. */
int vx_hba_window_range(const int32_t *win_size, int n_windows, int wdsize, int rank, int n_ranks, int32_t *lo, int32_t *hi);

/* The whole hierarchical global BA of thd_globalmapping for stored sessions, bottom level + top level (voxelslam.cpp:3084-3127):
 *   bottom  every window of the schedule through HBA_add_edge(xs = Keyframe::x0, max_iter = bottom_max_iter (1), thread_num = 2,
 *           plptr) -> edges `edges1` (gba_edges1) and one submap per window: pose Keyframe::x0 of its first keyframe, id / mp of
 *           that keyframe, cloud = the window's keyframe clouds merged with the optimized poses, down-sampled at voxel_size / 8
 *           (:3094-3099, :2954-2985)
 *   top     HBA_add_edge over all submaps whose session is in cnct_map (:2836-2857) with xs = kf_x of the submap's keyframe
 *           (multimap_scanPoses[mp][id]->x, :3110), max_iter = total_max_iter, thread_num = 5 -> edges `edges2` (gba_edges2)
 * With a communicator the bottom windows are block-distributed over the ranks (independent problems, no collective; edges are
 * gathered), the submap clouds stay on the GPU that made them, and the top level is voxel-sharded: points exchanged by
 * root-voxel owner at every outer iteration, [Hess | JacT | cost] all-reduced per evaluation, LM solve replicated.
 * Every rank passes the SAME host inputs and receives the SAME outputs.
 *   xyz / cloud_off     all keyframe clouds (local frame, float triples), n_kf + 1 offsets
 *   kf_session          session (Keyframe::mp) of every keyframe, ascending; kf_id: Keyframe::id
 *   kf_x0 / kf_x        Keyframe::x0 and the current scan pose of the keyframe's id (kf_x == NULL: x0)
 *   edges1 / edges2     capacity cap1 / cap2 edges; *n1 / *n2 written
 *   top_poses / top_kf  optimized pose and first keyframe of every top-level frame (n_top written; capacity = number of windows)
 *   stats (16 doubles, may be NULL): [0] windows, [1] top-level frames, [2] voxel-frames through acc_evaluate2 in all LM passes
 *           (this rank), [3..8] ms upload / bottom / merge / gather / top / total, [9] points in the top-level map,
 *           [10] bottom voxel-frame evaluations, [11] top voxel-frame evaluations (this rank), [12] / [13] voxel-frames through
 *           evaluate_only_residual at the bottom / top, [14] / [15] acc_evaluate2 / evaluate_only_residual passes at the top level */
typedef struct vx_hba_global_params {
  vx_hba_params hba;
  int32_t wdsize, mgsize;            /* 10, 5     voxelslam.cpp:3033-3034 */
  int32_t bottom_max_iter;           /* 1         :3096 */
  int32_t bottom_thread_num;         /* 2         :3096 */
  int32_t total_max_iter;            /* GBA/total_max_iter :3025 */
  int32_t top_thread_num;            /* 5         :3035 */
  int32_t xyz_on_device;             /* != 0: `xyz` is a DEVICE pointer to ALL keyframe clouds (resident keyframes: no upload) */
  int32_t keep_factors;              /* != 0: the LidarFactors of the last outer iteration are kept for vx_hba_global_take_factor */
} vx_hba_global_params;
int vx_hba_global(vx_ctx *ctx, vx_comm *comm, const float *xyz, const int64_t *cloud_off, int n_kf, const int32_t *kf_session,
                  const int32_t *kf_id, const vx_pose *kf_x0, const vx_pose *kf_x, const int32_t *cnct_map, int n_cnct,
                  const vx_hba_global_params *prm, vx_pgo_edge *edges1, int64_t cap1, int64_t *n1, vx_pgo_edge *edges2, int64_t cap2,
                  int64_t *n2, vx_pose *top_poses, int32_t *top_kf, int32_t *n_top, double *stats);
/* The LidarFactor (this rank's share) a vx_hba_global call with keep_factors left behind: which = 0 the batch of full bottom
 * windows, which = 1 the top level; poses (may be NULL, n_windows * W entries) = the poses it was last evaluated at.  The caller
 * owns the handle (vx_factor_free).  *out = NULL when nothing was kept. */
int vx_hba_global_take_factor(vx_ctx *ctx, int which, vx_factor **out);
/* acc_evaluate2 on a factor whose voxels are sharded over the ranks, results summed over the ranks on the device (one packed
 * all-reduce, voxel_map.hpp:382-384 across GPUs); asynchronous on the ctx stream, results in vx_factor_device_buffers. */
int vx_factor_hessian_sharded_async(vx_ctx *ctx, vx_comm *comm, vx_factor *f);

/* ---- D1 / D2: LIO ESIKF point-to-plane update -----------------------------------------------------------
 * The voxel map as OctoTree::match reads it (voxel_map.hpp:1649-1721), flattened by the caller: one record per
 * OctoTree node; root voxels (the unordered_map<VOXEL_LOC, OctoTree*> of voxel_map.hpp:2167-2205) by key. */
typedef struct vx_map_node {
  double center[3];        /* OctoTree::voxel_center */
  float quater_length;     /* OctoTree::quater_length (float, voxel_map.hpp:1067) */
  int32_t octo_state;      /* 0 leaf, 1 subdivided */
  int32_t child[8];        /* node index of leaves[4x+2y+z] or -1 */
  int32_t is_plane;        /* Plane::is_plane */
  float radius;            /* Plane::radius (float, voxel_map.hpp:89) */
  double p_center[3], p_normal[3];
  double plane_var[36];    /* Plane::plane_var 6x6 column-major */
} vx_map_node;
typedef struct vx_planemap vx_planemap;
/* struct pointVar {Vector3d pnt; Matrix3d var;}  voxel_map.hpp:18-34 (var column-major) */
/* the part of IMUST the update touches (tools.hpp:135-199): R, p, v, bg, ba and the 15x15 covariance, column-major */
typedef struct vx_lio_state { double R[9], p[3], v[3], bg[3], ba[3], cov[225]; } vx_lio_state;

int vx_planemap_create(vx_ctx *ctx, const int64_t *root_keys /* n_roots x 3 */, const int32_t *root_node, int64_t n_roots,
                       const vx_map_node *nodes, int64_t n_nodes, double voxel_size, vx_planemap **out);
int vx_planemap_free(vx_planemap *map);
/* One pass of the per-point loop of lio_state_estimation (voxelslam.cpp:1005-1050): var_world, wld, cached-leaf fast
 * path / match(), residual and Jacobian, HTH (6x6) / HTz (6) / sum n n^T (3x3) / match_num.  leaf_cache: in/out node
 * index per point (-1 none; NULL = no cache), the reference's `octos` vector (voxelslam.cpp:1020-1028).  flags: per
 * point association flag (may be NULL).  rot_var / tsl_var: x_curr.cov blocks (0,0) and (3,3), column-major. */
int vx_lio_match_accumulate(vx_ctx *ctx, vx_planemap *map, const vx_point_var *pts, int64_t n, const vx_pose *pose,
                            const double *rot_var9, const double *tsl_var9, int32_t *leaf_cache, double *HTH36, double *HTz6,
                            double *nnt9, int32_t *match_num, int32_t *flags);
/* int match(feat_map, wld, pla, var_wld, sigma_d, oc)  voxel_map.hpp:2167-2205 for a batch of n WORLD points (wld3: n x 3) with their
 * world covariances (var_wld9: n x 9 column-major): flags (the reference's return value), leaf (node index of `oc`, -1), sigma_d, and
 * — when planes != NULL — a copy of the matched leaf's record (the Plane behind `pla`).  leaf / sigma_d / planes may be NULL. */
int vx_planemap_match(vx_ctx *ctx, vx_planemap *map, const double *wld3, const double *var_wld9, int64_t n, int32_t *flags, int32_t *leaf,
                      double *sigma_d, vx_map_node *planes);
/* bool VOXEL_SLAM::lio_state_estimation(PVecPtr pptr)  voxelslam.cpp:962-1098: <= 4 iterations, 15x15 gain on the host,
 * points and leaf cache stay on the device.  *healthy = the reference's return (lambda_min(sum n n^T) >= 14). */
int vx_lio_state_estimation(vx_ctx *ctx, vx_planemap *map, const vx_point_var *pts, int64_t n, vx_lio_state *x, int32_t *healthy,
                            int32_t *iters, int32_t *match_nums4);

/* ---- A6 / A7 / A9 / A11 / A12: the sliding-window voxel map on the device ------------------------------------------
 * class OctoTree voxel_map.hpp:1047-1881, class SlideWindow :1009-1042, cut_voxel / cut_voxel_multi :1896-2096,
 * cut_voxel (fixed points) :2108-2152, drivers multi_recut voxelslam.cpp:1682-1737 and multi_margi :1590-1679.
 * The map lives in HBM as a node table (the vx_map_node match view + moments, 9x9 cov_add, per-slot clusters) under a
 * root-voxel hash; the window's raw points stay in per-slot scan buffers with one leaf id per point, so "subdivide" is a
 * re-labelling + segmented sum instead of list surgery.  The reference's globals (voxel_map.hpp:98-104, `mp` :1046) are
 * fields of vx_map_params / internal state. */
typedef struct vx_map vx_map;
typedef struct vx_map_params {
  double voxel_size, min_eigen_value;
  double plane_eigen_value_thre[4];   /* stored inverted, voxelslam.cpp:930-931 */
  double min_point[4];                /* voxelslam.cpp:917 */
  int32_t max_layer, max_points;      /* voxel_map.hpp:98-104 */
  int32_t win_size, thread_num;       /* thread_num only reproduces the "fewer roots than threads" early-outs */
} vx_map_params;
/* inspection record, one per node (tests / debugging); keyed by (root cell, layer, path), path = 3 bits per level */
typedef struct vx_map_node_state {
  int64_t root_key[3];
  int32_t layer, path, octo_state, isexist, has_sw, is_plane, last_num, opt_state, in_slide, n_window_pts, n_fix_pts, _pad;
  vx_cluster pcr_add, pcr_fix;
  double eig_value[3];
  double cov_add_trace;
} vx_map_node_state;

int vx_map_create(vx_ctx *ctx, const vx_map_params *params, vx_map **out);
int vx_map_free(vx_map *map);
/* cut_voxel_multi(surf_map, pvec, win_count, surf_map_slide, win_size, pwld, sws)  voxel_map.hpp:1964-2096.
 * pvec[i].pnt is in the scan-local frame, pvec[i].var the world covariance (after pvec_update), pwld the world points;
 * `ord` = the frame's position in the window (win_count - 1).  Keeps the ":2044-2045" quirk: when the scan touches fewer
 * root voxels than thread_num the roots are created / marked but no point is inserted. */
int vx_map_cut_voxel(vx_ctx *ctx, vx_map *map, const vx_point_var *pvec, const double *pwld, int64_t n, int ord);
/* The same insertion fed with the raw scan: var_init (voxelslam.cpp:1896), pvec_update (:1903) and cut_voxel_multi (:1917) run
 * back to back on the device; 12 B/point cross PCIe instead of 120.  rot_var / tsl_var: x_curr.cov blocks (0,0), (3,3). */
int vx_map_cut_voxel_xyz(vx_ctx *ctx, vx_map *map, const float *xyz, int64_t n, const vx_pose *ext, double dept_err, double beam_err,
                         const vx_pose *x, const double *rot_var9, const double *tsl_var9, int ord);
/* cut_voxel(feat_map, pvec, wdsize, jour)  voxel_map.hpp:2108-2152: already-world-frame points into pcr_fix / pcr_add only */
int vx_map_cut_voxel_fix(vx_ctx *ctx, vx_map *map, const vx_point_var *pvec, int64_t n, double jour);
/* multi_recut(surf_map_slide, win_count, x_buf, voxhess, sws)  voxelslam.cpp:1682-1737: OctoTree::recut on every root of the
 * slide set (eigen-solve, plane test, subdivision with the CURRENT poses), then tras_opt: the LidarFactor of all live
 * plane leaves is written into `f` on the device (f: win_size frames, single window).  poses: win_count entries. */
int vx_map_recut(vx_ctx *ctx, vx_map *map, const vx_pose *poses, int win_count, vx_factor *f);
/* multi_margi(surf_map_slide, jour, win_count, x_buf, voxhess, sws[0])  voxelslam.cpp:1590-1679 with mgsize frames folded
 * into pcr_fix; reads the optimizer cache of `f` (voxel_map.hpp:1495-1501); erases dead roots from the slide set.
 * VX_ERR_OPT_STATE where the reference prints and exits (voxel_map.hpp:1488-1492). */
int vx_map_margi(vx_ctx *ctx, vx_map *map, const vx_pose *poses, int win_count, int mgsize, vx_factor *f, double jour);
/* the `mp` ring rotation of voxelslam.cpp:2014-2019 */
int vx_map_shift_window(vx_map *map, int mgsize);
/* non-owning match view for vx_lio_* (valid until the next map-mutating call) */
int vx_map_planemap(vx_map *map, vx_planemap **view);
int64_t vx_map_num_nodes(const vx_map *map);
int64_t vx_map_num_roots(const vx_map *map);
int vx_map_export(vx_map *map, vx_map_node_state *states, vx_map_node *nodes);   /* num_nodes records each; either may be NULL */

/* ---- A5: calcBodyVar voxelslam.hpp:180-200, var_init :210-234, pvec_update :242-265.
 * vx_var_init: per-point 3x3 measurement covariance from range (dept_err) and bearing (beam_err, degrees) noise, point and
 * covariance rotated by the extrinsic `ext`; float xyz triples in (12 B/point), vx_point_var out (96 B/point).
 * vx_pvec_update: var <- R var R^T + hat(p) rot_var hat(p)^T + tsl_var and pwld = R p + t (pwld may be NULL); pv_out may be
 * pv_in.  rot_var / tsl_var: x_curr.cov blocks (0,0) and (3,3), column-major. */
int vx_var_init(vx_ctx *ctx, const float *xyz, int64_t n, const vx_pose *ext, double dept_err, double beam_err, vx_point_var *out);
int vx_pvec_update(vx_ctx *ctx, const vx_point_var *pv_in, int64_t n, const vx_pose *x, const double *rot_var9, const double *tsl_var9,
                   vx_point_var *pv_out, double *pwld);
/* ---- IMU de-skew — the per-point half of IMUEKF::motion_blur (ekf_imu.hpp:138-163): every point of a sweep
 * is moved into the sensor frame at the END of the sweep with the IMU-propagated pose that precedes its time stamp
 * (PointType::curvature).  The propagation that produces `poses` (ekf_imu.hpp:57-120) is the EKF and stays with the caller.
 *   pts      in/out host, x / y / z / curvature are read, x / y / z rewritten (float, as the reference stores them)
 *   poses    the reference's imu_poses, ascending t: IMUST(offt, R_imu, pos_imu, vel_imu, angvel_avr, acc_imu)
 *   xc       state at the end of the sweep (R, p);  lidar_to_imu: Lid_rot_to_IMU, Lid_offset_to_IMU
 *   assign   (may be NULL) index of the pose applied to each point, -1 = untouched (stamp not after any pose)
 * The reference's walk is reproduced: a point takes the latest pose whose time is below the MINIMUM stamp from that point to the
 * end of the sweep (for time-sorted sweeps: below its own stamp), and the first point is compensated again by every earlier pose. */
typedef struct vx_imu_pose { double t, R[9], p[3], v[3], angvel[3], acc[3]; } vx_imu_pose;
int vx_deskew(vx_ctx *ctx, vx_point48 *pts, int64_t n, const vx_imu_pose *poses, int n_poses, const vx_pose *xc, const vx_pose *lidar_to_imu,
              int32_t *assign);

/* lio_state_estimation fed with the raw scan: xyz is uploaded (12 B/point instead of 96), var_init runs on the device
 * (voxelslam.cpp:1880-1890 call order: var_init(extrin_para, pl, pptr, dept_err, beam_err) then lio_state_estimation(pptr)). */
int vx_lio_state_estimation_xyz(vx_ctx *ctx, vx_planemap *map, const float *xyz, int64_t n, const vx_pose *ext, double dept_err,
                                double beam_err, vx_lio_state *x, int32_t *healthy, int32_t *iters, int32_t *match_nums4);

#ifdef __cplusplus
}
#endif
#endif /* VOXELBA_H */
