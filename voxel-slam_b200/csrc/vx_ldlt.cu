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
//   host    forward / diagonal / backward substitution on the downloaded factor (n^2 work)
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

// diagonal block [c0, c0+nb): in-place L11 (unit lower, strictly-lower part stored) and D (diagonal)
__global__ void __launch_bounds__(NB * NB) k_ldlt_diag(double *__restrict__ B, int n, int c0, int nb) {
  __shared__ double S[NB][NB + 1];
  const int i = threadIdx.x % NB, j = threadIdx.x / NB;
  if (i < nb && j < nb && i >= j) S[i][j] = B[(size_t)(c0 + j) * n + c0 + i];
  __syncthreads();
  for (int k = 0; k < nb; k++) {
    // column k: subtract the contributions of columns t < k, then divide by the pivot
    if (j == 0 && i >= k && i < nb) {
      double s = S[i][k];
      for (int t = 0; t < k; t++) s -= S[i][t] * S[t][t] * S[k][t];
      S[i][k] = s;
    }
    __syncthreads();
    if (j == 0 && i > k && i < nb) { const double d = S[k][k]; if (fabs(d) > 0.0) S[i][k] /= d; }
    __syncthreads();
  }
  if (i < nb && j < nb && i >= j) B[(size_t)(c0 + j) * n + c0 + i] = S[i][j];
}

// rows below the diagonal block: L21 = (A21 - sum_t L21(:,t) D_t L11(j,t)) / D_j, one thread per row
__global__ void __launch_bounds__(128) k_ldlt_rows(double *__restrict__ B, int n, int c0, int nb) {
  __shared__ double L11[NB][NB + 1];
  __shared__ double D[NB];
  for (int k = threadIdx.x; k < NB * NB; k += blockDim.x) {
    const int i = k % NB, j = k / NB;
    if (i < nb && j < nb && i >= j) { const double v = B[(size_t)(c0 + j) * n + c0 + i]; if (i == j) D[i] = v; else L11[i][j] = v; }
  }
  __syncthreads();
  const int r = c0 + nb + blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  double row[NB];
#pragma unroll
  for (int j = 0; j < NB; j++) row[j] = j < nb ? B[(size_t)(c0 + j) * n + r] : 0.0;
#pragma unroll
  for (int j = 0; j < NB; j++) {
    if (j < nb) {
      double s = row[j];
#pragma unroll
      for (int t = 0; t < NB; t++) if (t < j) s -= row[t] * D[t] * L11[j][t];
      if (fabs(D[j]) > 0.0) s /= D[j];
      row[j] = s;
    }
  }
#pragma unroll
  for (int j = 0; j < NB; j++) if (j < nb) B[(size_t)(c0 + j) * n + r] = row[j];
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

}  // namespace

namespace vx {

// A: n x n column-major on the HOST (only the lower triangle is read); solves A x = b as Eigen's LDLT would.
int ldlt_lower_solve_device(vx_ctx *ctx, const double *A, int n, const double *b, double *x) {
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  // 1. Eigen's pivot rule on the input diagonal: at step k the largest |d_i|, i >= k, first index on ties; the swap moves
  //    the displaced entry to the winner's old position.
  std::vector<double> d(n);
  std::vector<int32_t> pos(n), tr(n);
  for (int i = 0; i < n; i++) { d[i] = std::fabs(A[(size_t)i * n + i]); pos[i] = i; }
  for (int k = 0; k < n; k++) {
    int big = k; double best = d[k];
    for (int i = k + 1; i < n; i++) if (d[i] > best) { best = d[i]; big = i; }
    tr[k] = big;
    if (big != k) { std::swap(d[k], d[big]); std::swap(pos[k], pos[big]); }
  }
  const size_t bytes = sizeof(double) * (size_t)n * n;
  DevBuf dA, dB, dP;
  std::vector<double> F((size_t)n * n);
  auto body = [&]() -> int {
    VX_TRY(dA.reserve(bytes)); VX_TRY(dB.reserve(bytes)); VX_TRY(dP.reserve(sizeof(int32_t) * n));
    VX_CUDA(cudaMemcpyAsync(dA.p, A, bytes, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemcpyAsync(dP.p, pos.data(), sizeof(int32_t) * n, cudaMemcpyHostToDevice, st));
    VX_CUDA(cudaMemsetAsync(dB.p, 0, bytes, st));
    VX_KERNEL(ctx, "k_permute_lower", k_permute_lower<<<dim3(blocks_for(n, 128), n), 128, 0, st>>>(dA.as<double>(), n, dP.as<int32_t>(), dB.as<double>()));
    for (int c0 = 0; c0 < n; c0 += NB) {
      const int nb = std::min(NB, n - c0);
      VX_KERNEL(ctx, "k_ldlt_diag", k_ldlt_diag<<<1, NB * NB, 0, st>>>(dB.as<double>(), n, c0, nb));
      const int below = n - c0 - nb;
      if (below > 0) {
        VX_KERNEL(ctx, "k_ldlt_rows", k_ldlt_rows<<<blocks_for(below, 128), 128, 0, st>>>(dB.as<double>(), n, c0, nb));
        const int nt = (below + TR - 1) / TR;
        VX_KERNEL(ctx, "k_ldlt_trailing", k_ldlt_trailing<<<dim3(nt, nt), 256, 0, st>>>(dB.as<double>(), n, c0, nb));
      }
    }
    VX_CUDA(cudaMemcpyAsync(F.data(), dB.p, bytes, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    return VX_OK;
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  dA.release(); dB.release(); dP.release();
  VX_TRY(rc);
  // 2. substitution on the host: x = P^T L^-T D^-1 L^-1 P b   (Eigen: LDLT::_solve_impl)
#define F_(r, c) F[(size_t)(c) * n + (r)]
  for (int i = 0; i < n; i++) x[i] = b[i];
  for (int k = 0; k < n; k++) if (tr[k] != k) std::swap(x[k], x[tr[k]]);
  for (int j = 0; j < n; j++) { const double xj = x[j]; if (xj != 0.0) for (int i = j + 1; i < n; i++) x[i] -= F_(i, j) * xj; }
  const double tol = 1.0 / std::numeric_limits<double>::max();
  for (int i = 0; i < n; i++) x[i] = std::fabs(F_(i, i)) > tol ? x[i] / F_(i, i) : 0.0;
  for (int j = n - 1; j >= 0; j--) { double s = x[j]; for (int i = j + 1; i < n; i++) s -= F_(i, j) * x[i]; x[j] = s; }
  for (int k = n - 1; k >= 0; k--) if (tr[k] != k) std::swap(x[k], x[tr[k]]);
#undef F_
  return VX_OK;
}

}  // namespace vx

// test / bench entry: solve one system through the device path
extern "C" int vx_ldlt_solve(vx_ctx *ctx, const double *A, int n, const double *b, double *x) {
  VX_ARG(ctx && A && b && x && n >= 1);
  return vx::ldlt_lower_solve_device(ctx, A, n, b, x);
}
