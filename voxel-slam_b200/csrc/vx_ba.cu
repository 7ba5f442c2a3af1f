// vx_ba.cu — Lidar_BA_Optimizer::damping_iter (voxel_map.hpp:422-497) over the device LidarFactor.
//
// acc_evaluate2 / evaluate_only_residual run on the device; so do the gauge fix and the (6W x 6W) LDL^T solve of every
// window with up to 128 unknowns (k_lm_solve below: W <= 21, i.e. every local BA and every bottom-level GBA window).  The
// LM control flow and the pose retraction (6W numbers per window) stay on the host.  Larger systems (top-level GBA) take
// the host sweep or the blocked device factorisation of vx_ldlt.cu.  A factor created with n_windows > 1 is a BATCH of independent problems (bottom-level GBA
// windows, voxelslam.cpp:3096): every window keeps its own damping state (u, v), accept/reject
// sequence and stop test, and each device pass evaluates all windows at once.
#include "vx_factor.cuh"
#include "vx_comm.cuh"
#include "vx_hostmat.hpp"
#include <cmath>
#include <limits>
#include <algorithm>
#include <cstdlib>

using namespace vx;

namespace {

// R <- R * Exp(w)   (tools.hpp:51-66: Rodrigues, identity below 1e-11)
void retract_rot(const double *R, const double *w, double *out) {
  const double n = std::sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  double E[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};   // column-major
  if (n >= 1e-11) {
    const double x = w[0] / n, y = w[1] / n, z = w[2] / n;
    const double K[9] = {0, z, -y, -z, 0, x, y, -x, 0};   // column-major hat(axis)
    double KK[9];
    for (int c = 0; c < 3; c++)
      for (int r = 0; r < 3; r++) KK[c * 3 + r] = K[0 * 3 + r] * K[c * 3 + 0] + K[1 * 3 + r] * K[c * 3 + 1] + K[2 * 3 + r] * K[c * 3 + 2];
    const double s = std::sin(n), c1 = 1.0 - std::cos(n);
    for (int k = 0; k < 9; k++) E[k] += s * K[k] + c1 * KK[k];
  }
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) out[c * 3 + r] = R[0 * 3 + r] * E[c * 3 + 0] + R[1 * 3 + r] * E[c * 3 + 1] + R[2 * 3 + r] * E[c * 3 + 2];
}


// ---- batched LM step on the device (6W <= LM_MAXN unknowns per window) ------------------------------------------
// One CTA per window, the system in shared memory, one thread per row.  The arithmetic is the host sweep's
// (hm::ldlt_lower_solve, vx_hostmat.hpp: Eigen's unblocked diagonal-pivot LDL^T on the lower triangle) operation for
// operation — explicit round-to-nearest multiplies and adds, no FMA contraction, sums in the same order — so dx, and with
// it every accept / reject decision of the LM loop, is bitwise what the host solve gives.  What it removes is the traffic:
// the 6W x 6W Hessians stay in HBM (11.7 MB per LM iteration at 399 windows), only dx and q1 (6W + 1 doubles per window)
// cross PCIe, and 399 solves run side by side instead of one after another on a host core.
constexpr int LM_MAXN = 128;

// windows with mask[w] != 0 take the new evaluation; the others keep their old Hess / JacT (voxel_map.hpp:486-489)
__global__ void k_lm_keep(const double *__restrict__ H, const double *__restrict__ J, double *__restrict__ Hk, double *__restrict__ Jk,
                          const int32_t *__restrict__ mask, int n) {
  const int w = blockIdx.x;
  if (!mask[w]) return;
  const size_t nn = (size_t)n * n;
  const size_t t0 = blockIdx.y * (size_t)blockDim.x + threadIdx.x, step = (size_t)gridDim.y * blockDim.x;
  if (nn % 2 == 0) {
    const double2 *src = reinterpret_cast<const double2 *>(H + nn * w);
    double2 *dst = reinterpret_cast<double2 *>(Hk + nn * w);
    for (size_t k = t0; k < nn / 2; k += step) dst[k] = src[k];
  } else {
    for (size_t k = t0; k < nn; k += step) Hk[nn * w + k] = H[nn * w + k];
  }
  for (size_t k = t0; k < (size_t)n; k += step) Jk[(size_t)n * w + k] = J[(size_t)n * w + k];
}

// Operation order.  The host sweep is LEFT-looking with interchanges: at step k it swaps k <-> pivot, then forms column k as
//   A(t,k) - sum_{j<k, tmp_j != 0} L(t,j) * (D_j L(k,j))   (j ascending, product and sum each rounded),  A(k,k) - acc_k,
//   acc_k = sum_{j<k} L(k,j) * (D_j L(k,j)) from 0.0, then divides the column by the new A(k,k).
// Here the same arithmetic runs RIGHT-looking on the matrix permuted up front: (1) the pivot sequence is a function of the
// input diagonal alone (when step k picks, positions >= k still hold input values), so one warp runs the host's selection on
// a copy of the diagonal; (2) the interchanges only move data, so the matrix is loaded as P A P^T (lower storage:
// entry (i,j) <- A(max(p_i,p_j), min(p_i,p_j))); (3) once column j is final, EVERY later entry (t,k) takes its j-th term —
// the terms still arrive in ascending j, on the same operands, with the same two roundings, so each entry goes through the
// very sequence of values it has on the host, but a step is (n-j)^2/2 independent multiply-adds over the CTA instead of a
// chain of k dependent ones per thread: 0.21 -> 0.05 ms per launch, and dx stays bitwise the host's.
template <int NT>
__global__ void __launch_bounds__(NT)
k_lm_solve(const double *__restrict__ Hk, const double *__restrict__ Jk, const double *__restrict__ u_w, const int32_t *__restrict__ act,
           int n, double *__restrict__ dx, double *__restrict__ q1) {
  extern __shared__ double lm_sm[];
  double *A = lm_sm;                 // [n][n] column-major, lower triangle used, pivot order
  double *wv = A + (size_t)n * n;    // [n]  D_j * L(k,j) of the current column j
  double *x = wv + n;                // [n]
  double *D = x + n;                 // [n]  gauged diagonal, input order
  double *Jg = D + n;                // [n]  gauged gradient, input order
  __shared__ int s_pos[LM_MAXN];     // s_pos[i] = input index at position i
  __shared__ double s_d[LM_MAXN];
  __shared__ double s_acc[LM_MAXN];
  const int w = blockIdx.x, t = threadIdx.x, lane = t & 31, wp = t >> 5;
  if (!act[w]) return;
  const double u = u_w[w];
  const double *Hw = Hk + (size_t)n * n * w;
#define LM_AT(r, c) A[(size_t)(c) * n + (r)]
  // gauge: pose 0 fixed (voxel_map.hpp:452-455); damping on the gauged diagonal (:457)
  if (t < n) {
    const double d0 = t < 6 ? 1.0 : Hw[(size_t)t * n + t];
    D[t] = d0;
    s_d[t] = fabs(__dadd_rn(d0, __dmul_rn(u, d0)));
    Jg[t] = t < 6 ? 0.0 : Jk[(size_t)n * w + t];
    s_pos[t] = t;
    s_acc[t] = 0.0;
  }
  __syncthreads();
  // the host's selection: first index of the largest |diagonal| in (k, n) if it beats |A(k,k)|; comparisons with NaN are false
  // (keys: the bit pattern of a non-negative double orders like the value; +1 keeps 0 for "no candidate / NaN"; the warp maximum
  // is three redux.sync — high word, low word among the high-word winners, then the smallest index among the winners)
  if (wp == 0) {
    for (int k = 0; k < n; k++) {
      unsigned long long key = 0ull; int bi = n;
      for (int i = k + 1 + lane; i < n; i += 32) {
        const double v = s_d[i];
        const unsigned long long kv = v == v ? (unsigned long long)__double_as_longlong(v) + 1ull : 0ull;
        if (kv > key) { key = kv; bi = i; }
      }
      const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
      const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
      const unsigned mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? lo : 0u);
      const int best = (int)__reduce_min_sync(0xffffffffu, (hi == mhi && lo == mlo && key != 0ull) ? (unsigned)bi : (unsigned)n);
      const double dk = s_d[k];
      const unsigned long long kk = dk == dk ? (unsigned long long)__double_as_longlong(dk) + 1ull : ~0ull;   // NaN at k: keep k
      const unsigned long long mk = ((unsigned long long)mhi << 32) | mlo;
      const int big = (mk > kk && best < n) ? best : k;
      __syncwarp();
      if (lane == 0 && big != k) {
        s_d[k] = s_d[big]; s_d[big] = dk;
        const int pk = s_pos[k]; s_pos[k] = s_pos[big]; s_pos[big] = pk;
      }
      __syncwarp();
    }
  }
  __syncthreads();
  for (int c = wp; c < n; c += NT / 32) {
    const int pc = s_pos[c];
    for (int r = c + lane; r < n; r += 32) {
      const int pr = s_pos[r];
      const int hi = max(pr, pc), lo = min(pr, pc);
      double v;
      if (hi < 6 || lo < 6) v = hi == lo ? 1.0 : 0.0;
      else v = Hw[(size_t)lo * n + hi];
      if (r == c) v = __dadd_rn(v, __dmul_rn(u, v));
      LM_AT(r, c) = v;
    }
  }
  if (t < n) x[t] = -Jg[s_pos[t]];
  __syncthreads();
  // thread = (row r, half h): row r's entries (r,k), j < k < r, alternate between its two threads; L(r,j) stays in a register.
  // Four columns are loaded, updated and stored as a batch (the compiler cannot tell A(r,k) and A(r,k+2) apart, so the batch is
  // what gives the loop its memory-level parallelism); the host's skip of a zero tmp is a select.
  constexpr int RW = NT / 2;
  const int r = t % RW, h = t / RW;
  for (int j = 0; j < n; j++) {
    // column j is complete but for its own diagonal term and the division
    const double ajj = __dadd_rn(LM_AT(j, j), -s_acc[j]);
    double l = 0.0;
    const bool live = r > j && r < n;
    if (live) {
      l = LM_AT(r, j);
      if (fabs(ajj) > 0.0) l = l / ajj;
      if (h == 0) { LM_AT(r, j) = l; wv[r] = __dmul_rn(ajj, l); }
    }
    __syncthreads();
    if (t == j) LM_AT(j, j) = ajj;
    if (live) {
      if (h == 0) s_acc[r] = __dadd_rn(s_acc[r], __dmul_rn(l, wv[r]));
      for (int k = j + 1 + h; k < r; k += 8) {
        double a[4], wk[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const int kq = k + 2 * q;
          wk[q] = kq < r ? wv[kq] : 0.0;
          a[q] = kq < r ? LM_AT(r, kq) : 0.0;
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const double na = __dadd_rn(a[q], -__dmul_rn(l, wk[q]));
          a[q] = wk[q] != 0.0 ? na : a[q];
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const int kq = k + 2 * q;
          if (kq < r) LM_AT(r, kq) = a[q];
        }
      }
    }
    __syncthreads();
  }
  // solve: L, D, L^T on the permuted right-hand side
  for (int j = 0; j < n; j++) {
    const double xj = x[j];
    if (xj != 0.0 && t > j && t < n) x[t] = __dadd_rn(x[t], -__dmul_rn(LM_AT(t, j), xj));
    __syncthreads();
  }
  if (t < n) {
    const double tol = 1.0 / 1.7976931348623157e308;
    x[t] = fabs(LM_AT(t, t)) > tol ? x[t] / LM_AT(t, t) : 0.0;
  }
  __syncthreads();
  // backward: x_j - sum_{i>j} L(i,j) x_i with i ascending is one dependent chain of n(n-1)/2 additions whatever is done; the
  // products are not on it — as soon as x_i is final the rows above park L(i,j) x_i in the unused upper triangle (A(j,i)),
  // and the chain of row i-1 only adds
  // (one barrier per row: the chain of row i-1 forms its newest term L(i,i-1) x_i itself while the other threads park theirs)
  for (int i = n - 1; i >= 1; i--) {
    if (t >= 1 && t <= i - 1) LM_AT(t - 1, i) = __dmul_rn(LM_AT(i, t - 1), x[i]);      // rows 0 .. i-2, by threads 1 .. i-1
    if (t == 0) {
      double s = __dadd_rn(x[i - 1], -__dmul_rn(LM_AT(i, i - 1), x[i]));
#pragma unroll 4
      for (int ii = i + 1; ii < n; ii++) s = __dadd_rn(s, -LM_AT(i - 1, ii));
      x[i - 1] = s;
    }
    __syncthreads();
  }
  if (t < n) wv[s_pos[t]] = x[t];        // back to input order
  __syncthreads();
  if (t == 0) {
    // q1 = 0.5 dx^T (u D dx - JacT)   (voxel_map.hpp:470)
    double q = 0.0;
    for (int r = 0; r < n; r++) q = __dadd_rn(q, __dmul_rn(wv[r], __dadd_rn(__dmul_rn(__dmul_rn(u, D[r]), wv[r]), -Jg[r])));
    q1[w] = 0.5 * q;
  }
  if (t < n) dx[(size_t)n * w + t] = wv[t];
#undef LM_AT
}

// The numbers HBA_add_edge reads off the Hessian snapshot (voxelslam.cpp:2926-2951): hess(6i+k, 6j+k), i < j, k < 6, in the
// order the edges are made; 6 W (W-1)/2 doubles per window instead of the (6W)^2 matrix (3.8 MB instead of 46 MB at W = 400).
__global__ void k_edge_diag(const double *__restrict__ H, int W, double *__restrict__ out) {
  const int n = 6 * W, w = blockIdx.y;
  const int64_t npair = (int64_t)W * (W - 1) / 2;
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= npair * 6) return;
  const int64_t p = t / 6; const int k = (int)(t % 6);
  // pair p -> (i, j): p = i (2W - i - 1) / 2 + (j - i - 1)
  int i = (int)((2.0 * W - 1.0 - sqrt((2.0 * W - 1.0) * (2.0 * W - 1.0) - 8.0 * (double)p)) * 0.5);
  while ((int64_t)i * (2 * W - i - 1) / 2 > p) i--;
  while ((int64_t)(i + 1) * (2 * W - i - 2) / 2 <= p) i++;
  const int j = (int)(p - (int64_t)i * (2 * W - i - 1) / 2) + i + 1;
  out[(size_t)w * npair * 6 + t] = H[(size_t)n * n * w + (size_t)(6 * j + k) * n + (6 * i + k)];
}

// ---- voxel-sharded evaluation: [lower triangle of Hess | JacT | cost] of every window, packed for ONE all-reduce ------------
// (the reference adds its per-thread partial Hess / JacT / residual the same way, voxel_map.hpp:382-384).  Hess is symmetric
// block-wise (acc_evaluate2 mirrors the upper blocks, :279-281) except for the antisymmetric -0.5*hat(jjt) term inside the
// diagonal blocks (:242), so the lower triangle alone is not the matrix: the strict upper part of each 6x6 DIAGONAL block
// travels too (15 numbers per frame).  Packed length per window: n(n+1)/2 + 15 W + n + 1.
__device__ __forceinline__ int64_t tri_col_start(int64_t c, int64_t n) { return c * n - c * (c - 1) / 2; }   // packed lower triangle, column-major
__global__ void k_pack_eval(const double *__restrict__ H, const double *__restrict__ J, const double *__restrict__ cost, int n, int64_t plen,
                            double *__restrict__ pack) {
  const int w = blockIdx.y;
  const double *Hw = H + (size_t)n * n * w;
  double *pw = pack + plen * w;
  const int64_t ntri = (int64_t)n * (n + 1) / 2;
  const int c = blockIdx.x;                       // one CTA per column
  if (c < n) {
    const int64_t cs = tri_col_start(c, n);
    for (int r = c + threadIdx.x; r < n; r += blockDim.x) pw[cs + (r - c)] = Hw[(size_t)c * n + r];
    // strict upper part of the diagonal 6x6 block: rows 6i .. c-1 of column c
    const int i = c / 6, cc = c % 6;
    if ((int)threadIdx.x < cc) pw[ntri + 15 * i + cc * (cc - 1) / 2 + threadIdx.x] = Hw[(size_t)c * n + 6 * i + threadIdx.x];
    if (threadIdx.x == 0) pw[ntri + 15 * (n / 6) + c] = J[(size_t)n * w + c];
    if (c == 0 && threadIdx.x == 0) pw[ntri + 15 * (n / 6) + n] = cost[w];
  }
}
__global__ void k_unpack_eval(const double *__restrict__ pack, int n, int64_t plen, double *__restrict__ H, double *__restrict__ J, double *__restrict__ cost) {
  const int w = blockIdx.y;
  double *Hw = H + (size_t)n * n * w;
  const double *pw = pack + plen * w;
  const int64_t ntri = (int64_t)n * (n + 1) / 2;
  const int c = blockIdx.x;
  if (c < n) {
    const int64_t cs = tri_col_start(c, n);
    const int dlim = 6 * (c / 6) + 6;               // rows below the diagonal block are mirrored into the upper triangle
    for (int r = c + threadIdx.x; r < n; r += blockDim.x) {
      const double v = pw[cs + (r - c)];
      Hw[(size_t)c * n + r] = v;
      if (r >= dlim) Hw[(size_t)r * n + c] = v;
    }
    const int i = c / 6, cc = c % 6;
    if ((int)threadIdx.x < cc) Hw[(size_t)c * n + 6 * i + threadIdx.x] = pw[ntri + 15 * i + cc * (cc - 1) / 2 + threadIdx.x];
    if (threadIdx.x == 0) J[(size_t)n * w + c] = pw[ntri + 15 * (n / 6) + c];
    if (c == 0 && threadIdx.x == 0) cost[w] = pw[ntri + 15 * (n / 6) + n];
  }
}

}  // namespace

namespace vx {
// Sum of the evaluation a voxel-sharded factor just produced (f->H, f->J, f->cost_h) over the ranks of `comm`, in place.
int allreduce_evaluation(vx_ctx *ctx, vx_comm *comm, vx_factor *f) {
  if (!comm || comm->size == 1) return VX_OK;
  const int n = 6 * f->W, NW = f->n_win;
  const int64_t plen = (int64_t)n * (n + 1) / 2 + 15 * f->W + n + 1;
  VX_TRY(comm->pack.reserve(sizeof(double) * (size_t)plen * NW));
  cudaStream_t st = ctx->stream;
  VX_KERNEL(ctx, "k_pack_eval", k_pack_eval<<<dim3(n, NW), 128, 0, st>>>(f->H.as<double>(), f->J.as<double>(), f->cost_h.as<double>(), n, plen, comm->pack.as<double>()));
  VX_TRY(comm_allreduce_f64(comm, comm->pack.as<double>(), plen * NW));
  VX_KERNEL(ctx, "k_unpack_eval", k_unpack_eval<<<dim3(n, NW), 128, 0, st>>>(comm->pack.as<double>(), n, plen, f->H.as<double>(), f->J.as<double>(), f->cost_h.as<double>()));
  return VX_OK;
}

int ldlt_lower_solve_device(vx_ctx *ctx, const double *A, int n, const double *b, double *x);   // vx_ldlt.cu
int lm_solve_large_device(vx_ctx *ctx, const double *dH, const double *dJ, double u, int n, double *dx, double *q1, int nfix = 6);   // vx_ldlt.cu
// win_status (optional, n_windows): when given, a window with fewer voxels than thd_num is flagged
// VX_ERR_TOO_FEW and left untouched instead of failing the whole batch.
// comm (optional): the factor holds only this rank's voxels of every window; Hess / JacT / cost are all-reduced after each
// evaluation, the LM step is replicated (every rank sees bitwise the same sums, so every rank takes the same decisions).
// edge_diag (optional, n_windows x W(W-1)/2 x 6): hess(6i+k, 6j+k) of the snapshot *hess would hold, without moving the matrix.
int ba_lidar_impl(vx_ctx *ctx, vx_factor *f, vx_pose *poses, double *hess_out, double *resis, int max_iter, int thd_num,
                  int *converged, int *accept_trace, int *n_iters, int *win_status, vx_comm *comm, double *edge_diag) {
  VX_ARG(f != nullptr && poses != nullptr && max_iter >= 0);
  VX_TRY(bind(ctx));
  const int W = f->W, NW = f->n_win, jl = 6 * W, F = W * NW;
  const bool sharded = comm != nullptr && comm->size > 1;
  int64_t Vglobal = f->V;
  std::vector<char> few;
  // Lidar_BA_Optimizer::only_residual exits the process when a problem has fewer voxels than threads
  // (voxel_map.hpp:399-403); the drop-in reports it instead.
  {
    std::vector<int32_t> wv(NW + 1);
    VX_CUDA(cudaMemcpyAsync(wv.data(), f->wv_off.p, sizeof(int32_t) * (NW + 1), cudaMemcpyDeviceToHost, ctx->stream));
    VX_CUDA(cudaStreamSynchronize(ctx->stream));
    std::vector<int64_t> nv(NW + 1);
    for (int w = 0; w < NW; w++) nv[w] = f->V > 0 ? wv[w + 1] - wv[w] : 0;
    nv[NW] = f->V;
    if (sharded) {   // voxel counts of the whole problem
      VX_TRY(comm->counts.reserve(8 * (size_t)(NW + 1)));
      VX_CUDA(cudaMemcpyAsync(comm->counts.p, nv.data(), 8 * (size_t)(NW + 1), cudaMemcpyHostToDevice, ctx->stream));
      VX_TRY(comm_allreduce_i64(comm, comm->counts.as<int64_t>(), NW + 1));
      VX_CUDA(cudaMemcpyAsync(nv.data(), comm->counts.p, 8 * (size_t)(NW + 1), cudaMemcpyDeviceToHost, ctx->stream));
      VX_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    Vglobal = nv[NW];
    few.assign(NW, 0);
    for (int w = 0; w < NW; w++) {
      if (win_status) win_status[w] = VX_OK;
      if (nv[w] < thd_num) {
        if (!win_status) { set_error("Too Less Voxel: window %d has %lld voxels < thd_num %d", w, (long long)nv[w], thd_num); return VX_ERR_TOO_FEW; }
        win_status[w] = VX_ERR_TOO_FEW; few[w] = 1;
      }
    }
  }
  static const int ldlt_device_min = [] { const char *e = std::getenv("VX_LDLT_DEVICE_MIN"); return e ? std::atoi(e) : 129; }();
  const size_t hsz = (size_t)jl * jl;
  // device LM step: k_lm_keep + k_lm_solve for windows of up to LM_MAXN unknowns, the blocked factorisation of vx_ldlt.cu from
  // ldlt_device_min unknowns on (top-level GBA); Hess and JacT never leave the device.  In between (129...383 unknowns) and
  // under VX_LM_HOST the host sweep.
  const bool small = jl <= LM_MAXN;
  const bool dev = (small || jl >= ldlt_device_min || sharded) && Vglobal > 0 && (sharded || f->nnz > 0) && getenv("VX_LM_HOST") == nullptr;
  const bool host_mats = !dev;
  std::vector<double> Hess(host_mats ? hsz * NW : 0), JacT(host_mats ? (size_t)jl * NW : 0), Hnew(Hess.size()), Jnew(JacT.size()), r_new(NW), r1(NW, 0.0), r2(NW, 0.0);
  std::vector<double> u(NW, 0.01), v(NW, 2.0), A(host_mats ? hsz : 0), dxi(jl), D(jl), nb(jl), tmp, q1s(NW, 0.0);
  std::vector<char> calc(NW, 1), active(NW, 1), conv(NW, 1);
  std::vector<int> iters(NW, 0), tr;
  std::vector<vx_pose> x(poses, poses + F), xt(poses, poses + F);
  if (hess_out && (!dev || max_iter == 0)) std::fill(hess_out, hess_out + hsz * NW, 0.0);   // the device path writes it whole at the end
  if (resis) std::fill(resis, resis + 2 * (size_t)NW, 0.0);
  for (int w = 0; w < NW; w++) if (few[w]) { active[w] = 0; conv[w] = 0; }

  std::vector<int32_t> mask(NW);
  std::vector<double> dx_all, Hsnap;
  if (edge_diag && host_mats) Hsnap.assign(hsz * NW, 0.0);
  cudaStream_t st = ctx->stream;
  if (dev) {
    VX_TRY(f->lm_H.reserve(sizeof(double) * hsz * NW)); VX_TRY(f->lm_J.reserve(sizeof(double) * jl * NW));
    VX_TRY(f->lm_dx.reserve(sizeof(double) * jl * NW)); VX_TRY(f->lm_q1.reserve(sizeof(double) * NW));
    VX_TRY(f->lm_u.reserve(sizeof(double) * NW)); VX_TRY(f->lm_mask.reserve(sizeof(int32_t) * NW));
    dx_all.resize((size_t)jl * NW);
    const size_t smem = sizeof(double) * ((size_t)jl * jl + 4 * jl);
    if (!small) {}
    else if (jl <= 64) VX_CUDA(cudaFuncSetAttribute(k_lm_solve<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else VX_CUDA(cudaFuncSetAttribute(k_lm_solve<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  bool evaluated = false;

  PhaseTimer pt("ba_lidar", ctx->stream);
  for (int it = 0; it < max_iter; it++) {
    bool any_active = false, any_calc = false;
    for (int w = 0; w < NW; w++) { any_active |= active[w] != 0; any_calc |= (active[w] && calc[w]); }
    if (!any_active) break;
    if (any_calc && dev) {
      VX_TRY(vx_factor_set_poses(ctx, f, x.data()));
      VX_TRY(vx_factor_hessian_async(ctx, f, 0, -1));
      VX_TRY(allreduce_evaluation(ctx, comm, f));
      for (int w = 0; w < NW; w++) mask[w] = active[w] && calc[w];
      VX_CUDA(cudaMemcpyAsync(f->lm_mask.p, mask.data(), sizeof(int32_t) * NW, cudaMemcpyHostToDevice, st));
      VX_KERNEL(ctx, "k_lm_keep", k_lm_keep<<<dim3(NW, (unsigned)std::max<size_t>(1, std::min<size_t>(hsz / 4096, 1024))), 256, 0, st>>>(f->H.as<double>(), f->J.as<double>(), f->lm_H.as<double>(), f->lm_J.as<double>(),
                                                              f->lm_mask.as<int32_t>(), jl));
      VX_CUDA(cudaMemcpyAsync(r_new.data(), f->cost_h.p, sizeof(double) * NW, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaStreamSynchronize(st));
      for (int w = 0; w < NW; w++) if (mask[w]) r1[w] = r_new[w];
      evaluated = true;
      pt.lap("hessian");
    } else if (any_calc) {
      VX_TRY(vx_factor_hessian(ctx, f, x.data(), 0, -1, Hnew.data(), Jnew.data(), r_new.data()));
      for (int w = 0; w < NW; w++)
        if (active[w] && calc[w]) {
          std::copy(Hnew.begin() + hsz * w, Hnew.begin() + hsz * (w + 1), Hess.begin() + hsz * w);
          std::copy(Jnew.begin() + (size_t)jl * w, Jnew.begin() + (size_t)jl * (w + 1), JacT.begin() + (size_t)jl * w);
          r1[w] = r_new[w];
          if (hess_out) std::copy(Hnew.begin() + hsz * w, Hnew.begin() + hsz * (w + 1), hess_out + hsz * w);   // *hess = Hess, before the gauge fix (:446)
          if (!Hsnap.empty()) std::copy(Hnew.begin() + hsz * w, Hnew.begin() + hsz * (w + 1), Hsnap.begin() + hsz * w);
        }
    }
    if (dev && !small) {
      for (int w = 0; w < NW; w++)
        if (active[w])
          VX_TRY(lm_solve_large_device(ctx, f->lm_H.as<double>() + hsz * w, f->lm_J.as<double>() + (size_t)jl * w, u[w], jl,
                                       dx_all.data() + (size_t)jl * w, &q1s[w]));
    } else if (dev) {
      for (int w = 0; w < NW; w++) mask[w] = active[w];
      VX_CUDA(cudaMemcpyAsync(f->lm_mask.p, mask.data(), sizeof(int32_t) * NW, cudaMemcpyHostToDevice, st));
      VX_CUDA(cudaMemcpyAsync(f->lm_u.p, u.data(), sizeof(double) * NW, cudaMemcpyHostToDevice, st));
      const size_t smem = sizeof(double) * ((size_t)jl * jl + 4 * jl);
      if (jl <= 64) {
        VX_KERNEL(ctx, "k_lm_solve", k_lm_solve<128><<<NW, 128, smem, st>>>(f->lm_H.as<double>(), f->lm_J.as<double>(), f->lm_u.as<double>(),
                                                                        f->lm_mask.as<int32_t>(), jl, f->lm_dx.as<double>(), f->lm_q1.as<double>()));
      } else {
        VX_KERNEL(ctx, "k_lm_solve", k_lm_solve<256><<<NW, 256, smem, st>>>(f->lm_H.as<double>(), f->lm_J.as<double>(), f->lm_u.as<double>(),
                                                                          f->lm_mask.as<int32_t>(), jl, f->lm_dx.as<double>(), f->lm_q1.as<double>()));
      }
      VX_CUDA(cudaMemcpyAsync(dx_all.data(), f->lm_dx.p, sizeof(double) * jl * NW, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaMemcpyAsync(q1s.data(), f->lm_q1.p, sizeof(double) * NW, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaStreamSynchronize(st));
    }
    pt.lap("solve");
    for (int w = 0; w < NW; w++) {
      if (!active[w]) continue;
      if (it == 0 && resis) resis[2 * w] = r1[w];
      const double *dxw = dxi.data();
      if (dev) {
        dxw = dx_all.data() + (size_t)jl * w;
      } else {
        double *H = Hess.data() + hsz * w, *J = JacT.data() + (size_t)jl * w;
        // gauge: pose 0 fixed (voxel_map.hpp:452-455)
        for (int c = 0; c < jl; c++) for (int r = 0; r < 6; r++) H[(size_t)c * jl + r] = 0.0;
        for (int c = 0; c < 6; c++) for (int r = 0; r < jl; r++) H[(size_t)c * jl + r] = 0.0;
        for (int r = 0; r < 6; r++) { H[(size_t)r * jl + r] = 1.0; J[r] = 0.0; }
        for (int r = 0; r < jl; r++) D[r] = H[(size_t)r * jl + r];
        std::copy(H, H + hsz, A.begin());
        for (int r = 0; r < jl; r++) { A[(size_t)r * jl + r] += u[w] * D[r]; nb[r] = -J[r]; }
        // the host sweep, or for the top-level GBA window (W ~ 399, 2394 unknowns, 4.6 GFLOP) the blocked device
        // factorisation (vx_ldlt.cu)
        if (jl >= ldlt_device_min) VX_TRY(ldlt_lower_solve_device(ctx, A.data(), jl, nb.data(), dxi.data()));
        else hm::ldlt_lower_solve(A.data(), jl, nb.data(), dxi.data(), tr, tmp);
        double q1 = 0.0;
        for (int r = 0; r < jl; r++) q1 += dxi[r] * (u[w] * D[r] * dxi[r] - J[r]);
        q1s[w] = 0.5 * q1;
      }
      for (int j = 0; j < W; j++) {
        const vx_pose &s = x[(size_t)w * W + j];
        vx_pose &d = xt[(size_t)w * W + j];
        retract_rot(s.R, &dxw[6 * j], d.R);
        for (int k = 0; k < 3; k++) d.p[k] = s.p[k] + dxw[6 * j + 3 + k];
      }
    }
    if (sharded) {
      VX_TRY(vx_factor_set_poses(ctx, f, xt.data()));
      VX_TRY(vx_factor_residual_async(ctx, f, 0, -1));
      VX_TRY(comm_allreduce_f64(comm, f->cost_r.as<double>(), NW));
      VX_CUDA(cudaMemcpyAsync(r_new.data(), f->cost_r.p, sizeof(double) * NW, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaStreamSynchronize(st));
    } else {
      VX_TRY(vx_factor_residual(ctx, f, xt.data(), 0, -1, r_new.data()));
    }
    pt.lap("residual");
    for (int w = 0; w < NW; w++) {
      if (!active[w]) continue;
      r2[w] = r_new[w];
      double q = r1[w] - r2[w];
      const double q1 = q1s[w];
      iters[w]++;
      if (q > 0) {
        for (int j = 0; j < W; j++) x[(size_t)w * W + j] = xt[(size_t)w * W + j];
        q = q / q1;
        v[w] = 2;
        q = 1 - std::pow(2 * q - 1, 3);
        u[w] *= (q < 1.0 / 3 ? 1.0 / 3 : q);
        calc[w] = 1;
        if (accept_trace) accept_trace[(size_t)w * max_iter + it] = 1;
      } else {
        u[w] = u[w] * v[w];
        v[w] = 2 * v[w];
        calc[w] = 0;
        conv[w] = 0;
        if (accept_trace) accept_trace[(size_t)w * max_iter + it] = 0;
        // keep feeding this window's rejected trial to later residual passes until a new one replaces it
      }
      if (std::fabs((r1[w] - r2[w]) / r1[w]) < 1e-6) {
        active[w] = 0;
      }
    }
    // windows that stopped keep the cache of their LAST trial (voxel_map.hpp:317-319): pin xt to it
  }
  if (dev && hess_out && evaluated) {   // *hess = Hess of the last accepted evaluation, before the gauge fix (:446)
    VX_CUDA(cudaMemcpyAsync(hess_out, f->lm_H.p, sizeof(double) * hsz * NW, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    for (int w = 0; w < NW; w++) if (few[w]) std::fill(hess_out + hsz * w, hess_out + hsz * (w + 1), 0.0);
  } else if (dev && hess_out && max_iter > 0) {
    std::fill(hess_out, hess_out + hsz * NW, 0.0);
  }
  if (edge_diag) {
    const int64_t npair = (int64_t)W * (W - 1) / 2;
    if (!(dev && evaluated)) std::fill(edge_diag, edge_diag + (size_t)NW * npair * 6, 0.0);      // the device path overwrites every slot
    if (dev && evaluated) {
      VX_TRY(ctx->scratch2.reserve(sizeof(double) * (size_t)NW * npair * 6));
      VX_KERNEL(ctx, "k_edge_diag", k_edge_diag<<<dim3((unsigned)blocks_for(npair * 6, 256), NW), 256, 0, st>>>(f->lm_H.as<double>(), W, ctx->scratch2.as<double>()));
      VX_CUDA(cudaMemcpyAsync(edge_diag, ctx->scratch2.p, sizeof(double) * (size_t)NW * npair * 6, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaStreamSynchronize(st));
      for (int w = 0; w < NW; w++) if (few[w]) std::fill(edge_diag + (size_t)w * npair * 6, edge_diag + (size_t)(w + 1) * npair * 6, 0.0);
    } else if (!dev && !Hsnap.empty()) {
      for (int w = 0; w < NW; w++) {
        int64_t p = 0;
        for (int i = 0; i < W - 1; i++)
          for (int j = i + 1; j < W; j++, p++)
            for (int k = 0; k < 6; k++) edge_diag[((size_t)w * npair + p) * 6 + k] = Hsnap[hsz * w + (size_t)(6 * j + k) * jl + (6 * i + k)];
      }
    }
  }
  pt.lap("hess_out");
  for (int w = 0; w < NW; w++) {
    if (resis) resis[2 * w + 1] = r2[w];
    if (converged) converged[w] = conv[w];
    if (n_iters) n_iters[w] = iters[w];
  }
  std::copy(x.begin(), x.end(), poses);
  return VX_OK;
}
}  // namespace vx

extern "C" int vx_ba_lidar(vx_ctx *ctx, vx_factor *f, vx_pose *poses, double *hess_out, double *resis, int max_iter, int thd_num,
                           int *converged, int *accept_trace, int *n_iters) {
  return vx::ba_lidar_impl(ctx, f, poses, hess_out, resis, max_iter, thd_num, converged, accept_trace, n_iters, nullptr, nullptr, nullptr);
}

// Lidar_BA_Optimizer::damping_iter on a factor whose voxels are sharded over the ranks of `comm` (each rank's `f` holds the
// voxels it owns, e.g. from vx_gba_build_sharded): one all-reduce of [Hess | JacT | cost] per evaluation — the sum the reference
// takes over its worker threads (voxel_map.hpp:382-384) — and one of the cost per trial step; every rank returns the same result.
extern "C" int vx_ba_lidar_sharded(vx_ctx *ctx, vx_comm *comm, vx_factor *f, vx_pose *poses, double *hess_out, double *resis, int max_iter,
                                   int thd_num, int *converged, int *accept_trace, int *n_iters) {
  VX_ARG(comm != nullptr && comm->ctx == ctx);
  return vx::ba_lidar_impl(ctx, f, poses, hess_out, resis, max_iter, thd_num, converged, accept_trace, n_iters, nullptr, comm, nullptr);
}
