// vx_ldlt.cu — (Hess + u*D).ldlt().solve(-JacT) for LARGE windows (voxel_map.hpp:457-458 at the top level of the
// hierarchical GBA, W ~ 399 submaps -> 2394 unknowns, 4.6 GFLOP per solve: SURVEY §8a B3, §8f rank 2).
//
// Eigen's LDLT is the unblocked LEFT-looking algorithm with "diagonal pivoting": at step k it picks the largest |A(i,i)|,
// i >= k.  In a left-looking sweep the diagonal entries of the not-yet-processed columns still hold their ORIGINAL values,
// so the whole pivot sequence is a function of the input diagonal alone.  That makes the factorisation
//     P A P^T = L D L^T   (no pivoting inside),   P = the selection order of |diag(A)|,
// which this file evaluates as a right-looking blocked sweep on the device:
//   host    pivot sequence from the diagonal (n^2/2 comparisons), exactly Eigen's transposition rule
//   k_permute_lower   B(i,j) = A(max(p_i,p_j), min(p_i,p_j)), i >= j  — only the LOWER triangle of the (not exactly
//                     symmetric) input is ever read, as Eigen does
//   per 32-column panel: k_ldlt_diag (1 CTA: the 32x32 diagonal block), k_ldlt_rows (one thread per row below it),
//                        k_ldlt_trailing (64x64 tiles of the trailing lower triangle -= L21 D L21^T)
//   k_ldlt_substitute  forward / diagonal / backward substitution, one CTA, the vector in shared memory, 32-column panels
//   The factor never leaves the device; vx_ba_lidar hands over the device Hessian (gauge fix and damping applied while it is
//   permuted, k_permute_gauged), so per LM iteration 4 n doubles cross PCIe instead of 4 n^2.
// Sums are re-associated relative to the sequential sweep; results agree to rounding (tests: dxi <= 1e-9 relative,
// identical LM accept/reject traces).
#include "vx_common.cuh"
#include <cmath>
#include <limits>
#include <vector>
#include <algorithm>

using namespace vx;

namespace {

constexpr int NB = 32;     // panel width
constexpr int TR = 64;     // trailing tile

__global__ void k_permute_lower(const double *__restrict__ A, int n, const int32_t *__restrict__ pos, double *__restrict__ B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
  if (i >= n || i < j) return;
  const int pi = pos[i], pj = pos[j];
  const int r = max(pi, pj), c = min(pi, pj);
  B[(size_t)j * n + i] = A[(size_t)c * n + r];
}

// diagonal block [c0, c0+nb): in-place L11 (unit lower, strictly-lower part stored) and D (diagonal).  Right-looking with one
// thread per entry: column k is scaled by its pivot, then every remaining entry takes its rank-1 update — 3 barriers per
// column instead of a dependent chain of k multiply-adds on 32 threads (25 us -> 5 us per panel).
__global__ void __launch_bounds__(NB * NB) k_ldlt_diag(double *__restrict__ B, int n, int c0, int nb) {
  __shared__ double S[NB][NB + 1];
  const int i = threadIdx.x % NB, j = threadIdx.x / NB;
  const bool in = i < nb && j < nb && i >= j;
  if (in) S[i][j] = B[(size_t)(c0 + j) * n + c0 + i];
  __syncthreads();
  for (int k = 0; k < nb; k++) {
    const double d = S[k][k];
    __syncthreads();
    if (j == k && i > k && i < nb && fabs(d) > 0.0) S[i][k] /= d;
    __syncthreads();
    if (in && j > k) S[i][j] -= S[i][k] * d * S[j][k];
    __syncthreads();
  }
  if (in) B[(size_t)(c0 + j) * n + c0 + i] = S[i][j];
}

// rows below the diagonal block: L21(r,j) = (A(r,j) - sum_{t<j} L21(r,t) D_t L11(j,t)) / D_j.  One warp per row, lane = column
// of the panel: column j is final once the updates of columns < j have been applied, its value is broadcast and the lanes to its
// right take their update — 32 short steps instead of one thread walking a 496-term triangle (19 us -> 3 us per panel).  A CTA
// moves a 32-row tile through shared memory so that the global accesses stay coalesced along the rows.
__global__ void __launch_bounds__(NB * 32) k_ldlt_rows(double *__restrict__ B, int n, int c0, int nb) {
  __shared__ double L11[NB][NB + 1];
  __shared__ double D[NB];
  __shared__ double T[32][NB + 1];    // [row in tile][column of the panel]
  const int tx = threadIdx.x % 32, ty = threadIdx.x / 32;
  {
    const int i = tx, j = ty;
    if (i < nb && j < nb && i >= j) { const double v = B[(size_t)(c0 + j) * n + c0 + i]; if (i == j) D[i] = v; else L11[i][j] = v; }
  }
  const int r0 = c0 + nb + blockIdx.x * 32;
  if (ty < nb && r0 + tx < n) T[tx][ty] = B[(size_t)(c0 + ty) * n + r0 + tx];
  __syncthreads();
  {
    const int row = ty, lane = tx;      // warp ty <-> row r0 + ty
    if (r0 + row < n) {
      double a = lane < nb ? T[row][lane] : 0.0;
      for (int j = 0; j < nb; j++) {
        double l = __shfl_sync(0xffffffffu, a, j);
        const double dj = D[j];
        if (fabs(dj) > 0.0) l /= dj;
        if (lane == j) a = l;
        else if (lane > j && lane < nb) a -= l * dj * L11[lane][j];
      }
      if (lane < nb) T[row][lane] = a;
    }
  }
  __syncthreads();
  if (ty < nb && r0 + tx < n) B[(size_t)(c0 + ty) * n + r0 + tx] = T[tx][ty];
}

// trailing update of the lower triangle: C(I,J) -= L(I,:) D L(J,:)^T over the panel's nb columns
__global__ void __launch_bounds__(256) k_ldlt_trailing(double *__restrict__ B, int n, int c0, int nb) {
  const int t0 = c0 + nb;
  const int bi = blockIdx.x, bj = blockIdx.y;
  if (bj > bi) return;
  __shared__ double Li[NB][TR + 1], Lj[NB][TR + 1];    // [t][row]; Lj pre-scaled by D_t
  const int r0 = t0 + bi * TR, q0 = t0 + bj * TR;
  for (int k = threadIdx.x; k < NB * TR; k += blockDim.x) {
    const int rr = k % TR, t = k / TR;
    double a = 0.0, b = 0.0;
    if (t < nb) {
      const double d = B[(size_t)(c0 + t) * n + c0 + t];
      if (r0 + rr < n) a = B[(size_t)(c0 + t) * n + r0 + rr];
      if (q0 + rr < n) b = B[(size_t)(c0 + t) * n + q0 + rr] * d;
    }
    Li[t][rr] = a; Lj[t][rr] = b;
  }
  __syncthreads();
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;     // 16 x 16 threads, 4 x 4 outputs each
  double acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; a++)
#pragma unroll
    for (int b = 0; b < 4; b++) acc[a][b] = 0.0;
#pragma unroll 8
  for (int t = 0; t < NB; t++) {
    double x[4], y[4];
#pragma unroll
    for (int a = 0; a < 4; a++) { x[a] = Li[t][tx + 16 * a]; y[a] = Lj[t][ty + 16 * a]; }
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
      for (int b = 0; b < 4; b++) acc[a][b] += x[a] * y[b];
  }
#pragma unroll
  for (int a = 0; a < 4; a++)
#pragma unroll
    for (int b = 0; b < 4; b++) {
      const int r = r0 + tx + 16 * a, c = q0 + ty + 16 * b;
      if (r < n && c < n && r >= c) B[(size_t)c * n + r] -= acc[a][b];
    }
}


// gauged diagonal of the un-gauged device Hessian: D[r] = 1 for the six gauge rows (voxel_map.hpp:452-455), else H(r,r)
__global__ void k_diag_gauged(const double *__restrict__ H, int n, double *__restrict__ D) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n) D[r] = r < 6 ? 1.0 : H[(size_t)r * n + r];
}

// B(i,j) = Ag(max(p_i,p_j), min(p_i,p_j)), i >= j, where Ag = gauge-fixed H + u * diag (voxel_map.hpp:452-458) formed on the fly
__global__ void k_permute_gauged(const double *__restrict__ H, int n, const int32_t *__restrict__ pos, double u, double *__restrict__ B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
  if (i >= n || i < j) return;
  const int pi = pos[i], pj = pos[j];
  const int r = max(pi, pj), c = min(pi, pj);
  double v;
  if (r < 6 || c < 6) v = r == c ? 1.0 : 0.0;
  else v = H[(size_t)c * n + r];
  if (r == c) v = __dadd_rn(v, __dmul_rn(u, v));     // A(r,r) += u * D[r], D = the gauged diagonal
  B[(size_t)j * n + i] = v;
}

// x <- L^-T D^-1 L^-1 x on the factor in B (unit lower L below the diagonal, D on it).  One CTA of 1024 threads, x in shared
// memory; per 32-column panel the diagonal block is solved by warp 0 (shuffles) and the rest is a panel-wide update (forward:
// one thread per row, 32 independent loads in flight; backward: one warp per column, coalesced dot products).
__global__ void __launch_bounds__(1024) k_ldlt_substitute(const double *__restrict__ B, int n, double *__restrict__ x) {
  extern __shared__ double xs[];
  __shared__ double L11[NB][NB + 1];
  __shared__ double part[NB];
  const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
  for (int i = tid; i < n; i += 1024) xs[i] = x[i];
  __syncthreads();
  for (int c0 = 0; c0 < n; c0 += NB) {
    const int nb = min(NB, n - c0);
    { const int i = tid % NB, j = tid / NB; if (i < nb && j < nb && i > j) L11[i][j] = B[(size_t)(c0 + j) * n + c0 + i]; }
    __syncthreads();
    if (wp == 0) {
      double xi = lane < nb ? xs[c0 + lane] : 0.0;
      for (int t = 0; t < nb; t++) {
        const double xt = __shfl_sync(0xffffffffu, xi, t);
        if (lane > t && lane < nb) xi -= L11[lane][t] * xt;
      }
      if (lane < nb) xs[c0 + lane] = xi;
    }
    __syncthreads();
    for (int r = c0 + nb + tid; r < n; r += 1024) {
      double acc = xs[r];
#pragma unroll 8
      for (int t = 0; t < nb; t++) acc -= B[(size_t)(c0 + t) * n + r] * xs[c0 + t];
      xs[r] = acc;
    }
    __syncthreads();
  }
  {
    const double tol = 1.0 / 1.7976931348623157e308;
    for (int i = tid; i < n; i += 1024) { const double d = B[(size_t)i * n + i]; xs[i] = fabs(d) > tol ? xs[i] / d : 0.0; }
  }
  __syncthreads();
  for (int c0 = ((n - 1) / NB) * NB; c0 >= 0; c0 -= NB) {
    const int nb = min(NB, n - c0);
    { const int i = tid % NB, j = tid / NB; if (i < nb && j < nb && i > j) L11[i][j] = B[(size_t)(c0 + j) * n + c0 + i]; }
    if (wp < nb) {
      const double *col = B + (size_t)(c0 + wp) * n;
      double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
      int i = c0 + nb + lane;
      for (; i + 96 < n; i += 128) { s0 += col[i] * xs[i]; s1 += col[i + 32] * xs[i + 32]; s2 += col[i + 64] * xs[i + 64]; s3 += col[i + 96] * xs[i + 96]; }
      for (; i < n; i += 32) s0 += col[i] * xs[i];
      double sacc = (s0 + s1) + (s2 + s3);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) sacc += __shfl_down_sync(0xffffffffu, sacc, o);
      if (lane == 0) part[wp] = sacc;
    }
    __syncthreads();
    if (wp == 0) {
      double xj = lane < nb ? xs[c0 + lane] - part[lane] : 0.0;
      for (int t = nb - 1; t >= 0; t--) {
        const double xt = __shfl_sync(0xffffffffu, xj, t);
        if (lane < t) xj -= L11[t][lane] * xt;
      }
      if (lane < nb) xs[c0 + lane] = xj;
    }
    __syncthreads();
  }
  for (int i = tid; i < n; i += 1024) x[i] = xs[i];
}

}  // namespace

namespace vx {

// Eigen's pivot rule on the damped diagonal d (non-negative): at step k the largest d_i, i >= k, first index on ties; the swap
// moves the displaced entry to the winner's old position.  pos[k] = original index that ends up at position k, tr[k] = the
// transposition of step k.  With distinct values the order is simply "descending" and the transpositions follow from an
// index map in O(n log n); ties take the literal O(n^2) selection.
static void pivot_order(std::vector<double> d, std::vector<int32_t> &pos, std::vector<int32_t> &tr) {
  const int n = (int)d.size();
  pos.resize(n); tr.resize(n);
  std::vector<int32_t> ord(n);
  for (int i = 0; i < n; i++) ord[i] = i;
  bool distinct = true;
  for (int i = 0; i < n; i++) if (!(d[i] == d[i])) distinct = false;    // NaN: no ordering to sort by
  if (distinct)
    std::sort(ord.begin(), ord.end(), [&](int a, int b) { return d[a] > d[b] || (d[a] == d[b] && a < b); });
  for (int k = 0; distinct && k + 1 < n; k++) if (!(d[ord[k]] > d[ord[k + 1]])) distinct = false;   // ties: the literal rule
  if (distinct) {
    std::vector<int32_t> at(n), where(n);   // at[p] = original index now at position p; where[i] = position of original index i
    for (int i = 0; i < n; i++) { at[i] = i; where[i] = i; }
    for (int k = 0; k < n; k++) {
      const int win = ord[k], big = where[win];
      tr[k] = big;
      if (big != k) { const int lose = at[k]; at[k] = win; at[big] = lose; where[win] = k; where[lose] = big; }
    }
    for (int k = 0; k < n; k++) pos[k] = at[k];
    return;
  }
  for (int i = 0; i < n; i++) pos[i] = i;
  for (int k = 0; k < n; k++) {
    int big = k; double best = d[k];
    for (int i = k + 1; i < n; i++) if (d[i] > best) { best = d[i]; big = i; }
    tr[k] = big;
    if (big != k) { std::swap(d[k], d[big]); std::swap(pos[k], pos[big]); }
  }
}

// blocked factorisation of the permuted lower triangle in dB, then x <- (L D L^T)^-1 x on the device; x: host, n doubles
static int factor_and_substitute(vx_ctx *ctx, double *dB, int n, double *x) {
  cudaStream_t st = ctx->stream;
  for (int c0 = 0; c0 < n; c0 += NB) {
    const int nb = std::min(NB, n - c0);
    VX_KERNEL(ctx, "k_ldlt_diag", k_ldlt_diag<<<1, NB * NB, 0, st>>>(dB, n, c0, nb));
    const int below = n - c0 - nb;
    if (below > 0) {
      VX_KERNEL(ctx, "k_ldlt_rows", k_ldlt_rows<<<blocks_for(below, 32), NB * 32, 0, st>>>(dB, n, c0, nb));
      const int nt = (below + TR - 1) / TR;
      VX_KERNEL(ctx, "k_ldlt_trailing", k_ldlt_trailing<<<dim3(nt, nt), 256, 0, st>>>(dB, n, c0, nb));
    }
  }
  VX_TRY(ctx->scratch2.reserve(sizeof(double) * n));
  double *dx = ctx->scratch2.as<double>();
  VX_CUDA(cudaMemcpyAsync(dx, x, sizeof(double) * n, cudaMemcpyHostToDevice, st));
  const size_t smem = sizeof(double) * n;
  VX_ARG(smem <= 200 * 1024);     // 25600 unknowns
  static size_t attr = 0;
  if (smem > 48 * 1024 && smem > attr) { VX_CUDA(cudaFuncSetAttribute(k_ldlt_substitute, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = smem; }
  VX_KERNEL(ctx, "k_ldlt_substitute", k_ldlt_substitute<<<1, 1024, smem, st>>>(dB, n, dx));
  VX_CUDA(cudaMemcpyAsync(x, dx, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
  VX_CUDA(cudaStreamSynchronize(st));
  return VX_OK;
}

// A: n x n column-major on the HOST (only the lower triangle is read); solves A x = b as Eigen's LDLT would.
int ldlt_lower_solve_device(vx_ctx *ctx, const double *A, int n, const double *b, double *x) {
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  std::vector<double> d(n);
  std::vector<int32_t> pos, tr;
  for (int i = 0; i < n; i++) d[i] = std::fabs(A[(size_t)i * n + i]);
  pivot_order(d, pos, tr);
  const size_t bytes = sizeof(double) * (size_t)n * n;
  DevBuf dA, dB, dP;
  for (int i = 0; i < n; i++) x[i] = b[i];
  for (int k = 0; k < n; k++) if (tr[k] != k) std::swap(x[k], x[tr[k]]);
  auto body = [&]() -> int {
    VX_TRY(dA.reserve(bytes)); VX_TRY(dB.reserve(bytes)); VX_TRY(dP.reserve(sizeof(int32_t) * n));
    VX_CUDA(cudaMemcpyAsync(dA.p, A, bytes, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemcpyAsync(dP.p, pos.data(), sizeof(int32_t) * n, cudaMemcpyHostToDevice, st));
    VX_KERNEL(ctx, "k_permute_lower", k_permute_lower<<<dim3(blocks_for(n, 128), n), 128, 0, st>>>(dA.as<double>(), n, dP.as<int32_t>(), dB.as<double>()));
    return factor_and_substitute(ctx, dB.as<double>(), n, x);
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  dA.release(); dB.release(); dP.release();
  VX_TRY(rc);
  for (int k = n - 1; k >= 0; k--) if (tr[k] != k) std::swap(x[k], x[tr[k]]);
  return VX_OK;
}

// The LM step of one LARGE window on the device Hessian: dH (n x n, un-gauged, column-major) and dJ (n) stay where they are.
// dx = (gauge(H) + u diag(gauge(H))).ldlt().solve(-gauge(JacT)),  q1 = 0.5 dx^T (u D dx - JacT)   (voxel_map.hpp:452-470)
int lm_solve_large_device(vx_ctx *ctx, const double *dH, const double *dJ, double u, int n, double *dx, double *q1) {
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  std::vector<double> D(n), J(n), d(n);
  std::vector<int32_t> pos, tr;
  DevBuf dB, dP, dD;
  auto body = [&]() -> int {
    VX_TRY(dB.reserve(sizeof(double) * (size_t)n * n)); VX_TRY(dP.reserve(sizeof(int32_t) * n)); VX_TRY(dD.reserve(sizeof(double) * n));
    VX_KERNEL(ctx, "k_diag_gauged", k_diag_gauged<<<blocks_for(n, 256), 256, 0, st>>>(dH, n, dD.as<double>()));
    VX_CUDA(cudaMemcpyAsync(D.data(), dD.p, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaMemcpyAsync(J.data(), dJ, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    for (int r = 0; r < 6 && r < n; r++) J[r] = 0.0;
    for (int r = 0; r < n; r++) d[r] = std::fabs(D[r] + u * D[r]);
    pivot_order(d, pos, tr);
    for (int i = 0; i < n; i++) dx[i] = -J[i];
    for (int k = 0; k < n; k++) if (tr[k] != k) std::swap(dx[k], dx[tr[k]]);
    VX_CUDA(cudaMemcpyAsync(dP.p, pos.data(), sizeof(int32_t) * n, cudaMemcpyHostToDevice, st));
    VX_KERNEL(ctx, "k_permute_gauged", k_permute_gauged<<<dim3(blocks_for(n, 128), n), 128, 0, st>>>(dH, n, dP.as<int32_t>(), u, dB.as<double>()));
    return factor_and_substitute(ctx, dB.as<double>(), n, dx);
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  dB.release(); dP.release(); dD.release();
  VX_TRY(rc);
  for (int k = n - 1; k >= 0; k--) if (tr[k] != k) std::swap(dx[k], dx[tr[k]]);
  double q = 0.0;
  for (int r = 0; r < n; r++) q += dx[r] * (u * D[r] * dx[r] - J[r]);
  *q1 = 0.5 * q;
  return VX_OK;
}

}  // namespace vx

// test / bench entry: solve one system through the device path
extern "C" int vx_ldlt_solve(vx_ctx *ctx, const double *A, int n, const double *b, double *x) {
  VX_ARG(ctx && A && b && x && n >= 1);
  return vx::ldlt_lower_solve_device(ctx, A, n, b, x);
}
