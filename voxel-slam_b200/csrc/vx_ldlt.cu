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
//   per 32-column panel: k_ldlt_panel (one warp per row below the diagonal block),
//                        k_ldlt_trailing (64x64 tiles of the trailing lower triangle -= L21 D L21^T; tile (0,0) goes on to
//                        factor the next panel's 32x32 diagonal block, so that step is off the panel's dependent chain)
//   substitution: forward + diagonal ride in k_ldlt_panel; backward = k_ldlt_back, one multi-CTA launch per panel
//   The factor never leaves the device; vx_ba_lidar hands over the device Hessian (gauge fix and damping applied while it is
//   permuted, k_permute_gauged), so per LM iteration 4 n doubles cross PCIe instead of 4 n^2.
// Sums are re-associated relative to the sequential sweep; results agree to rounding (tests: dxi <= 1e-9 relative,
// identical LM accept/reject traces).
#include "vx_common.cuh"
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <limits>
#include <vector>
#include <algorithm>

using namespace vx;

namespace {

// Panel width 32.  Tried and removed (round 2): 64-column panels (the 64 x 64 diagonal block as a 2 x 2 arrangement of the 32 x 32
// routine, a lane holding two columns in the panel / backward kernels, the trailing K loop run twice over the same strips) halve
// the steps to 38 — and every kernel took twice as long per launch (serialised: trailing 29 us against 16.8, panel 15.7 against
// 9.4, backward 14 against 8.9), so one LM iteration at 2,394 unknowns came out at 2.37 ms against 2.29.  What a step costs is not
// the kernel boundary but work proportional to the columns it retires: the diagonal block's pivot chain (~135 cycles per column),
// the 8 / 32 KB of L11 every panel and backward CTA loads, the strips of the update.
// Also tried and removed: the trailing update FUSED with the next panel's work (tile (0,0) factors the next diagonal block, runs
// the forward substitution and publishes both behind a release flag; the tiles of tile column 0 spin on the flag and apply the
// panel's triangular solve to their own rows in place — bit-identical factor, one launch per step instead of two): 2.78 ms per LM
// iteration against 2.29.  After the flag a waiting tile still pays a global round trip for L11, two barriers, the 32-step solve
// and a store — the ~4.5 us the separate panel kernel costs — and the column-0 tiles of the later waves lengthen the kernel's tail.
// Together with the programmatic-launch result: kernel boundaries are NOT what the sweep spends its time on; its chain is
// update (~2.5 us) -> diagonal block (~4.5 us) -> panel solve (~4.5 us) per 32 columns, each a load / compute / store round.
constexpr int NB = 32;     // panel width
constexpr int TR = 64;     // trailing tile

// Programmatic dependent launch (sm_90+), opt-in (VX_LDLT_PDL=1) and a measured dead end: launched with the programmatic-
// serialisation attribute a kernel's CTAs may be placed before its predecessor has drained and block in pdl_wait() until the
// predecessor's writes are visible; the edges survive stream capture.  On the 2,394-unknown sweep (scripts/top_wall.py, one LM
// iteration): 2.27 ms with the edges, 2.28 ms without; with an early griddepcontrol.launch_dependents in every kernel 6.70 ms
// against 6.34 for three iterations (the waiting CTAs take slots from the predecessor's tail).  Inside a replayed graph the
// dependent launches are already back to back; what remains between two kernels is the drain of the predecessor, which a
// dependent that needs ALL of its output cannot skip.  Every kernel of the sweep calls pdl_wait() first, so the edges are safe.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
template <class... KArgs, class... Args>
cudaError_t launch_chain(bool pdl, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

__global__ void k_permute_lower(const double *__restrict__ A, int n, const int32_t *__restrict__ pos, double *__restrict__ B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
  if (i >= n || i < j) return;
  const int pi = pos[i], pj = pos[j];
  const int r = max(pi, pj), c = min(pi, pj);
  B[(size_t)j * n + i] = A[(size_t)c * n + r];
}

// In-place L11 (unit lower, strictly-lower part) and D (diagonal) of a 32x32 block in shared memory (zero-padded lower triangle),
// NT threads, 8 columns at a time: warp 0 (lane = row) takes the 8 columns through all rows with the 8 entries of its row in
// registers (28 rank-1 steps, unrolled; operands by shuffle), then every thread applies the rank-8 update to what is left of
// the block.  Straight-line code for all 496 steps is instruction-fetch bound (a CTA runs it once), a rolled 32-column loop is a
// chain of 496 dependent shared-memory round trips; this form has 32 divisions + 8 barriers on its chain.
// 1/d for a normal-range d: hardware seed (rcp.approx.ftz.f64, ~20 bits) + two Newton steps, relative error ~1e-16 (not the
// correctly rounded quotient: the factor agrees with the sequential sweep to rounding, as the re-associated sums already do)
__device__ __forceinline__ double rcp_newton(double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  r = fma(fma(-d, r, 1.0), r, r);
  r = fma(fma(-d, r, 1.0), r, r);
  return r;
}

template <int NT>
__device__ __forceinline__ void ldlt_diag_block(double (*L11)[NB + 1], int tid) {
  const int tx = tid % 32, ty = tid / 32;
  for (int kb = 0; kb < NB; kb += 8) {
    if (ty == 0) {
      const int r = kb + tx;             // rows kb .. NB-1
      double a[8];
#pragma unroll
      for (int c = 0; c < 8; c++) a[c] = r < NB ? L11[r][kb + c] : 0.0;
#pragma unroll
      for (int k = 0; k < 8; k++) {
        const double d = __shfl_sync(0xffffffffu, a[k], k);
        double l = a[k];
        const double ad = fabs(d);
        if (ad > 1e-290 && ad < 1e290) l *= rcp_newton(d);       // the 32 pivots are the block's dependent chain: ~55 cycles, not ~300
        else if (ad > 0.0) l /= d;
        if (tx > k) a[k] = l;
        const double ld = l * d;
#pragma unroll
        for (int j = k + 1; j < 8; j++) {
          const double ljk = __shfl_sync(0xffffffffu, l, j);
          if (tx >= j) a[j] -= ld * ljk;
        }
      }
      if (r < NB) {
#pragma unroll
        for (int c = 0; c < 8; c++) if (tx >= c) L11[r][kb + c] = a[c];
      }
    }
    __syncthreads();
    const int m = NB - kb - 8;           // rows / columns left
    for (int e = tid; e < m * m; e += NT) {
      const int i = kb + 8 + e % m, j = kb + 8 + e / m;
      if (i >= j) {
        double acc = L11[i][j];
#pragma unroll
        for (int c = 0; c < 8; c++) acc -= L11[i][kb + c] * L11[kb + c][kb + c] * L11[j][kb + c];
        L11[i][j] = acc;
      }
    }
    __syncthreads();
  }
}

// the first diagonal block; the others are factored by tile (0,0) of the trailing update that completes them
__global__ void __launch_bounds__(256) k_ldlt_diag0(const double *__restrict__ B, double *__restrict__ F, int n, int nb) {
  __shared__ double L11[NB][NB + 1];
  pdl_wait();
  for (int e = threadIdx.x; e < NB * NB; e += 256) {
    const int i = e % NB, j = e / NB;
    L11[i][j] = (i < nb && j < nb && i >= j) ? B[(size_t)j * n + i] : 0.0;
  }
  __syncthreads();
  ldlt_diag_block<256>(L11, threadIdx.x);
  for (int e = threadIdx.x; e < NB * NB; e += 256) F[e] = L11[e % NB][e / NB];
}

// Panel [c0, c0+nb): the rows below the (already factored, F[panel][col][row]) diagonal block:
// L21(r,j) = (A(r,j) - sum_{t<j} L21(r,t) D_t L11(j,t)) / D_j, one warp per row, lane = column of the panel; the entry of column
// j BEFORE its division is L21(r,j) D_j, so it is broadcast as it stands and the lanes to its right take their update — the
// division stays off the dependent chain.  The forward substitution rides along.
constexpr int PR = 16;     // rows per CTA
__global__ void __launch_bounds__(PR * 32) k_ldlt_panel(double *__restrict__ B, const double *__restrict__ F, int n, int c0, int nb, double *__restrict__ x,
                                                       double *__restrict__ z) {
  __shared__ double L11[NB][NB + 1];
  __shared__ double D[NB], Y[NB];
  __shared__ double T[PR][NB + 1];    // [row in tile][column of the panel]
  const int tid = threadIdx.x, tx = tid % 32, ty = tid / 32;
  pdl_wait();
  for (int e = tid; e < NB * NB; e += PR * 32) L11[e % NB][e / NB] = F[(size_t)(c0 / NB) * NB * NB + e];
  const int r0 = c0 + nb + blockIdx.x * PR;
  for (int e = tid; e < PR * NB; e += PR * 32) {
    const int row = e % PR, col = e / PR;
    if (col < nb && r0 + row < n) T[row][col] = B[(size_t)(c0 + col) * n + r0 + row];
  }
  __syncthreads();
  if (ty == 0) {
    D[tx] = L11[tx][tx];
    __syncwarp();
    // forward substitution rides along: y_p = L11^-1 x_p (x_p already carries the updates of the panels to the left)
    double xi = tx < nb ? x[c0 + tx] : 0.0;
    for (int t = 0; t < nb; t++) {
      const double xt = __shfl_sync(0xffffffffu, xi, t);
      if (tx > t && tx < nb) xi -= L11[tx][t] * xt;
    }
    Y[tx] = xi;
    if (blockIdx.x == 0 && tx < nb) {       // z = D^-1 L^-1 x, Eigen's rule for tiny pivots (LDLT::_solve_impl)
      const double tol = 1.0 / 1.7976931348623157e308;
      const double d = D[tx];
      z[c0 + tx] = fabs(d) > tol ? xi / d : 0.0;
    }
  }
  __syncthreads();
  {
    const int row = ty, lane = tx;      // warp ty <-> row r0 + ty
    if (r0 + row < n) {
      double a = lane < nb ? T[row][lane] : 0.0;
      for (int j = 0; j < nb; j++) {       // lanes right of j take L(r,j) D_j L11(lane,j) = (entry before its division) * L11(lane,j):
        const double pj = __shfl_sync(0xffffffffu, a, j);      // the division stays off the dependent chain
        if (lane > j && lane < nb) a -= pj * L11[lane][j];
      }
      if (lane < nb) { const double dj = D[lane]; if (fabs(dj) > 0.0) a /= dj; T[row][lane] = a; }
      double sx = lane < nb ? a * Y[lane] : 0.0;       // x_r -= L21(r,:) y_p
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) sx += __shfl_xor_sync(0xffffffffu, sx, o);
      if (lane == 0) x[r0 + row] -= sx;
    }
  }
  __syncthreads();
  for (int e = tid; e < PR * NB; e += PR * 32) {
    const int row = e % PR, col = e / PR;
    if (col < nb && r0 + row < n) B[(size_t)(c0 + col) * n + r0 + row] = T[row][col];
  }
}

// The next panel's diagonal block on its own: C(I,I) -= L(I,:) D L(I,:)^T for the 32 x 32 block at t0 = c0 + nb — the sums of
// k_ldlt_trailing's tile (0,0), operation for operation — then its factorisation.  One CTA, launched on a second branch of the sweep
// beside the trailing update (which then leaves that block alone), so the block's ~4.5 us pivot chain no longer rides behind an update.
__global__ void __launch_bounds__(256) k_ldlt_dnext(const double *__restrict__ B, double *__restrict__ F, int n, int c0, int nb) {
  __shared__ double S[NB][NB + 1];       // S[t][i] = L(t0 + i, c0 + t)
  __shared__ double Dk[NB];
  __shared__ double L11[NB][NB + 1];
  const int t0 = c0 + nb, tid = threadIdx.x;
  const double *Fd = F + (size_t)(c0 / NB) * NB * NB;
  for (int e = tid; e < NB * NB; e += 256) {
    const int i = e % NB, t = e / NB;
    S[t][i] = (t < nb && t0 + i < n) ? B[(size_t)(c0 + t) * n + t0 + i] : 0.0;
  }
  if (tid < NB) Dk[tid] = tid < nb ? Fd[tid * NB + tid] : 0.0;
  __syncthreads();
  for (int e = tid; e < NB * NB; e += 256) {
    const int i = e % NB, j = e / NB;
    double v = 0.0;
    if (i >= j && t0 + i < n && t0 + j < n) {
      const double cc = B[(size_t)(t0 + j) * n + t0 + i];
      double acc = 0.0;
#pragma unroll 8
      for (int t = 0; t < NB; t++) acc = fma(S[t][i], S[t][j] * Dk[t], acc);
      v = cc - acc;
    }
    L11[i][j] = v;
  }
  __syncthreads();
  ldlt_diag_block<256>(L11, tid);
  for (int e = tid; e < NB * NB; e += 256) F[(size_t)(t0 / NB) * NB * NB + e] = L11[e % NB][e / NB];
}

// trailing update of the lower triangle: C(I,J) -= L(I,:) D L(J,:)^T over the panel's nb columns.  OWN_DIAG: tile (0,0) also
// factors the next diagonal block (single-branch sweep); otherwise it leaves that block to k_ldlt_dnext and does not write it.
template <bool OWN_DIAG>
__global__ void __launch_bounds__(256) k_ldlt_trailing(double *__restrict__ B, double *__restrict__ F, int n, int c0, int nb) {
  const int t0 = c0 + nb;
  int bi = (int)((sqrtf(8.0f * (float)blockIdx.x + 1.0f) - 1.0f) * 0.5f);      // lower-triangular tiles only: blockIdx.x = bi (bi + 1) / 2 + bj
  while (bi * (bi + 1) / 2 > (int)blockIdx.x) bi--;
  while ((bi + 1) * (bi + 2) / 2 <= (int)blockIdx.x) bi++;
  const int bj = (int)blockIdx.x - bi * (bi + 1) / 2;
  __shared__ double Li[NB][TR + 1], Lj[NB][TR + 1];    // [t][row]; Lj pre-scaled by D_t
  const int r0 = t0 + bi * TR, q0 = t0 + bj * TR;
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;     // 16 x 16 threads, 4 x 4 outputs each
  pdl_wait();
  // every global load of the tile is issued before anything waits: the C entries first (used last), then the two L strips
  double cc[4][4];
#pragma unroll
  for (int a = 0; a < 4; a++)
#pragma unroll
    for (int b = 0; b < 4; b++) {
      const int r = r0 + tx + 16 * a, c = q0 + ty + 16 * b;
      cc[a][b] = (r < n && c < n && r >= c) ? B[(size_t)c * n + r] : 0.0;
    }
  double la[NB * TR / 256], lb[NB * TR / 256];
#pragma unroll
  for (int q = 0; q < NB * TR / 256; q++) {
    const int k = threadIdx.x + 256 * q;
    const int rr = k % TR, t = k / TR;
    la[q] = (t < nb && r0 + rr < n) ? B[(size_t)(c0 + t) * n + r0 + rr] : 0.0;
    lb[q] = (t < nb && q0 + rr < n) ? B[(size_t)(c0 + t) * n + q0 + rr] * F[(size_t)(c0 / NB) * NB * NB + t * NB + t] : 0.0;
  }
#pragma unroll
  for (int q = 0; q < NB * TR / 256; q++) {
    const int k = threadIdx.x + 256 * q;
    Li[k / TR][k % TR] = la[q]; Lj[k / TR][k % TR] = lb[q];
  }
  __syncthreads();
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
      const bool dnext = !OWN_DIAG && r < t0 + NB && c < t0 + NB;      // the block k_ldlt_dnext is reading right now
      if (r < n && c < n && r >= c && !dnext) B[(size_t)c * n + r] = cc[a][b] - acc[a][b];
    }
  if (OWN_DIAG && bi == 0 && bj == 0) {   // this tile completes the next panel's diagonal block: factor it here, off the panel's chain
    __syncthreads();                      // everybody is done with Li / Lj
    double (*L11)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(&Li[0][0]);
#pragma unroll
    for (int a = 0; a < 2; a++)
#pragma unroll
      for (int b = 0; b < 2; b++) {
        const int i = tx + 16 * a, j = ty + 16 * b;
        L11[i][j] = (t0 + i < n && t0 + j < n && i >= j) ? cc[a][b] - acc[a][b] : 0.0;
      }
    __syncthreads();
    ldlt_diag_block<256>(L11, threadIdx.x);
    for (int e = threadIdx.x; e < NB * NB; e += 256) F[(size_t)(t0 / NB) * NB * NB + e] = L11[e % NB][e / NB];
  }
}


// The same update on the FP64 tensor pipe (mma.sync.m8n8k4.f64, SASS DMMA.8x8x4): 128 x 128 tiles of the lower triangle, 8 warps,
// each warp a 32 x 64 sub-tile = 4 x 8 accumulator tiles held in registers and INITIALISED WITH C (the panel strip is stored negated,
// so D = (-L_I) (L_J D)^T + C is the updated entry: one pass, no separate subtraction).  Operand strips live in shared memory as
// [k][row] with leading dimension 132 (= 4 mod 16): fragment loads and the transposing stores are both conflict-free.
// Per k-step a warp issues 12 fragment loads for 32 DMMAs (the 4 x 4-per-thread DFMA tile needed 8 loads per 16 FMAs).
// The north_star's "tensor cores only if the solve proves to be a dense contraction": this is that contraction (4.6 GFLOP at W = 400).
constexpr int TT = 128, LDT = TT + 4;
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__global__ void __launch_bounds__(256, 1) k_ldlt_trailing_mma(double *__restrict__ B, double *__restrict__ F, int n, int c0, int nb) {
  extern __shared__ double tsm[];
  double *Li = tsm, *Lj = tsm + NB * LDT;          // [t][row]; Li negated, Lj pre-scaled by D_t
  const int t0 = c0 + nb;
  int bi = (int)((sqrtf(8.0f * (float)blockIdx.x + 1.0f) - 1.0f) * 0.5f);      // lower-triangular tiles only: blockIdx.x = bi (bi + 1) / 2 + bj
  while (bi * (bi + 1) / 2 > (int)blockIdx.x) bi--;
  while ((bi + 1) * (bi + 2) / 2 <= (int)blockIdx.x) bi++;
  const int bj = (int)blockIdx.x - bi * (bi + 1) / 2;
  const int r0 = t0 + bi * TT, q0 = t0 + bj * TT;
  const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
  const int wm = wp & 3, wn = wp >> 2;
  const int lr = lane >> 2, lc = lane & 3;
  pdl_wait();
  // accumulators <- C (issued first: the longest latency), then the strips
  double acc[4][8][2];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const int r = r0 + wm * 32 + i * 8 + lr, c = q0 + wn * 64 + j * 8 + 2 * lc;
      const bool rin = r < n;
      acc[i][j][0] = (rin && c < n) ? B[(size_t)c * n + r] : 0.0;
      acc[i][j][1] = (rin && c + 1 < n) ? B[(size_t)(c + 1) * n + r] : 0.0;
    }
  const double *Fd = F + (size_t)(c0 / NB) * NB * NB;
#pragma unroll 4
  for (int q = 0; q < NB * TT / 256; q++) {
    const int k = tid + 256 * q;
    const int rr = k % TT, t = k / TT;
    const bool tin = t < nb;
    Li[t * LDT + rr] = (tin && r0 + rr < n) ? -B[(size_t)(c0 + t) * n + r0 + rr] : 0.0;
    Lj[t * LDT + rr] = (tin && q0 + rr < n) ? B[(size_t)(c0 + t) * n + q0 + rr] * Fd[t * NB + t] : 0.0;
  }
  __syncthreads();
#pragma unroll
  for (int ks = 0; ks < NB / 4; ks++) {
    const double *la = Li + (4 * ks + lc) * LDT + wm * 32 + lr;
    const double *lb = Lj + (4 * ks + lc) * LDT + wn * 64 + lr;
    double a[4], b[8];
#pragma unroll
    for (int i = 0; i < 4; i++) a[i] = la[8 * i];
#pragma unroll
    for (int j = 0; j < 8; j++) b[j] = lb[8 * j];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 8; j++) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
  }
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const int r = r0 + wm * 32 + i * 8 + lr, c = q0 + wn * 64 + j * 8 + 2 * lc;
      if (r < n && c < n && r >= c) B[(size_t)c * n + r] = acc[i][j][0];
      if (r < n && c + 1 < n && r >= c + 1) B[(size_t)(c + 1) * n + r] = acc[i][j][1];
    }
  if (bi == 0 && bj == 0) {               // this tile completes the next panel's diagonal block: factor it here, off the panel's chain
    __syncthreads();                      // everybody is done with the strips
    double (*L11)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(tsm);
    if (wp == 0) {                        // warp 0 holds rows 0..31 x columns 0..63 of the tile
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++)
#pragma unroll
          for (int e = 0; e < 2; e++) {
            const int rr = i * 8 + lr, cc = j * 8 + 2 * lc + e;
            L11[rr][cc] = (t0 + rr < n && t0 + cc < n && rr >= cc) ? acc[i][j][e] : 0.0;
          }
    }
    __syncthreads();
    ldlt_diag_block<256>(L11, tid);
    for (int e = tid; e < NB * NB; e += 256) F[(size_t)(t0 / NB) * NB * NB + e] = L11[e % NB][e / NB];
  }
}

// gauged diagonal of the un-gauged device Hessian: D[r] = 1 for the six gauge rows (voxel_map.hpp:452-455), else H(r,r)
__global__ void k_diag_gauged(const double *__restrict__ H, int n, int nfix, double *__restrict__ D) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n) D[r] = r < nfix ? 1.0 : H[(size_t)r * n + r];
}

// B(i,j) = Ag(max(p_i,p_j), min(p_i,p_j)), i >= j, where Ag = gauge-fixed H + u * diag (voxel_map.hpp:452-458) formed on the fly
__global__ void k_permute_gauged(const double *__restrict__ H, int n, int nfix, const int32_t *__restrict__ pos, double u, double *__restrict__ B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
  if (i >= n || i < j) return;
  const int pi = pos[i], pj = pos[j];
  const int r = max(pi, pj), c = min(pi, pj);
  double v;
  if (r < nfix || c < nfix) v = r == c ? 1.0 : 0.0;
  else v = H[(size_t)c * n + r];
  if (r == c) v = __dadd_rn(v, __dmul_rn(u, v));     // A(r,r) += u * D[r], D = the gauged diagonal
  B[(size_t)j * n + i] = v;
}

// Backward substitution, one launch per panel from the last to the first: every CTA solves the panel's 32 unknowns
// L11^T x_p = z_p in warp 0 (shuffles; z_p already carries the updates of the panels to the right), CTA 0 stores them, and each
// warp then takes BC columns i < c0 of the factor: z_i -= L(c0.., i)^T x_p, lane = row of the panel (256 contiguous bytes per
// column).  x_p goes to its own array: the other CTAs of the launch are still reading z_p.
constexpr int BC = 4;
__global__ void __launch_bounds__(256) k_ldlt_back(const double *__restrict__ B, const double *__restrict__ F, int n, int c0, int nb, double *__restrict__ z,
                                                   double *__restrict__ xo) {
  __shared__ double L11[NB][NB + 1];
  __shared__ double X[NB];
  const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
  pdl_wait();
  for (int e = tid; e < NB * NB; e += 256) L11[e % NB][e / NB] = F[(size_t)(c0 / NB) * NB * NB + e];
  __syncthreads();
  if (wp == 0) {
    double xj = lane < nb ? z[c0 + lane] : 0.0;
    for (int t = nb - 1; t >= 0; t--) {
      const double xt = __shfl_sync(0xffffffffu, xj, t);
      if (lane < t) xj -= L11[t][lane] * xt;
    }
    X[lane] = xj;
    if (blockIdx.x == 0 && lane < nb) xo[c0 + lane] = xj;
  }
  __syncthreads();
  const double xr = lane < nb ? X[lane] : 0.0;
  const int i0 = (blockIdx.x * 8 + wp) * BC;
  double v[BC];
#pragma unroll
  for (int q = 0; q < BC; q++) v[q] = (i0 + q < c0 && lane < nb) ? B[(size_t)(i0 + q) * n + c0 + lane] * xr : 0.0;
#pragma unroll
  for (int q = 0; q < BC; q++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
    if (lane == 0 && i0 + q < c0) z[i0 + q] -= v[q];
  }
}

}  // namespace

namespace vx {

// Eigen's pivot rule on the damped diagonal d (non-negative): at step k the largest d_i among the positions i >= k, the first
// POSITION on ties; the swap moves the displaced entry to the winner's old position.  pos[k] = original index that ends up at
// position k, tr[k] = the transposition of step k.  Evaluated in O(n log n): entries sorted by value; within a run of equal
// values the members are taken in the order of their positions when the run starts — a member still waiting is never the one
// displaced (the displaced entry sits at position k, the smallest position left: were it a member it would be the winner),
// so those positions stay put while the run is consumed.  (The gauge rows make ties the rule, not the exception: six 1 + u.)
// NaN on the diagonal: the literal O(n^2) selection.
static void pivot_order(const std::vector<double> &d_in, std::vector<int32_t> &pos, std::vector<int32_t> &tr) {
  const std::vector<double> &d = d_in;
  const int n = (int)d.size();
  pos.resize(n); tr.resize(n);
  bool ordered = true;
  for (int i = 0; i < n; i++) if (!(d[i] == d[i])) ordered = false;
  if (ordered) {
    std::vector<int32_t> ord(n), at(n), where(n);   // at[p] = original index now at position p; where[i] = position of index i
    {   // sort (value, index) records, not indices through a comparator that chases them: 0.15 -> 0.03 ms at n = 2394, once per LM solve
      struct Rec { double d; int32_t i; };
      std::vector<Rec> rec(n);
      for (int i = 0; i < n; i++) { rec[i].d = d[i]; rec[i].i = i; at[i] = i; where[i] = i; }
      std::sort(rec.begin(), rec.end(), [](const Rec &a, const Rec &b) { return a.d > b.d || (a.d == b.d && a.i < b.i); });
      for (int i = 0; i < n; i++) ord[i] = rec[i].i;
    }
    int k = 0;
    while (k < n) {
      int e = k + 1;
      while (e < n && d[ord[e]] == d[ord[k]]) e++;
      if (e - k > 1) std::sort(ord.begin() + k, ord.begin() + e, [&](int a, int b) { return where[a] < where[b]; });
      for (; k < e; k++) {
        const int win = ord[k], big = where[win];
        tr[k] = big;
        if (big != k) { const int lose = at[k]; at[k] = win; at[big] = lose; where[win] = k; where[lose] = big; }
      }
    }
    for (int i = 0; i < n; i++) pos[i] = at[i];
    return;
  }
  std::vector<double> w(d);
  for (int i = 0; i < n; i++) pos[i] = i;
  for (int k = 0; k < n; k++) {
    int big = k; double best = w[k];
    for (int i = k + 1; i < n; i++) if (w[i] > best) { best = w[i]; big = i; }
    tr[k] = big;
    if (big != k) { std::swap(w[k], w[big]); std::swap(pos[k], pos[big]); }
  }
}

// the sweep: panel + trailing update per 32-column panel, left to right (the forward substitution rides in the panel kernel),
// then the backward substitution right to left: 3 launches per panel, 225 at n = 2394
constexpr size_t TRAIL_SMEM = sizeof(double) * 2 * NB * LDT;
static int enqueue_sweep(vx_ctx *ctx, double *dB, double *dF, int n, double *dx, double *dz, double *dxo) {
  cudaStream_t st = ctx->stream;
  {   // > 48 KB of dynamic shared memory: opt in once per device
    static bool attr_set[64] = {};
    if (!attr_set[ctx->device & 63]) {
      VX_CUDA(cudaFuncSetAttribute(k_ldlt_trailing_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TRAIL_SMEM));
      attr_set[ctx->device & 63] = true;
    }
  }
  static const bool pdl = std::getenv("VX_LDLT_PDL") != nullptr;
  static const bool two_branch = std::getenv("VX_LDLT_ONE_BRANCH") == nullptr && std::getenv("VX_LDLT_DMMA") == nullptr;
  if (two_branch && !ctx->ldlt_aux) {
    VX_CUDA(cudaStreamCreateWithFlags(&ctx->ldlt_aux, cudaStreamNonBlocking));
    for (int k = 0; k < 2; k++) VX_CUDA(cudaEventCreateWithFlags(&ctx->ldlt_ev[k], cudaEventDisableTiming));
  }
  VX_KERNEL(ctx, "k_ldlt_diag0", launch_chain(false, k_ldlt_diag0, dim3(1), dim3(256), 0, st, dB, dF, n, std::min(NB, n)));   // head of the chain: a full dependency on what wrote dB
  for (int c0 = 0; c0 < n; c0 += NB) {
    const int nb = std::min(NB, n - c0);
    const int below = n - c0 - nb;
    VX_KERNEL(ctx, "k_ldlt_panel", launch_chain(pdl, k_ldlt_panel, dim3(std::max(1, blocks_for(below, PR))), dim3(PR * 32), 0, st, dB, dF, n, c0, nb, dx, dz));
    if (below > 0) {
      // Measured on B200 at n = 2400 (scripts/ldlt_time.py): 64 x 64 DFMA tiles 1.25 ms per solve (3.7 TFLOP/s), 128 x 128 DMMA tiles
      // 1.80 ms (2.6 TFLOP/s).  With a 32-column panel the update moves 16 B of C per 64 flop and every launch is a short wave of
      // tiles: it is bound by L2 latency and wave quantisation (190 tiles on 148 SMs at one CTA per SM), not by the FP64 pipe, so
      // the tensor-pipe formulation buys nothing here.  VX_LDLT_DMMA=1 selects it for the A/B record.
      static const bool dfma_tiles = std::getenv("VX_LDLT_DMMA") == nullptr;
      if (dfma_tiles) {
        const int nt = (below + TR - 1) / TR;
        if (two_branch) {
          // fork: the next diagonal block on the aux stream beside the update of everything else; join before the next panel
          cudaStream_t sd = ctx->profiling ? st : ctx->ldlt_aux;
          if (sd != st) { VX_CUDA(cudaEventRecord(ctx->ldlt_ev[0], st)); VX_CUDA(cudaStreamWaitEvent(sd, ctx->ldlt_ev[0], 0)); }
          if (ctx->profiling) { VX_KERNEL(ctx, "k_ldlt_dnext", k_ldlt_dnext<<<1, 256, 0, sd>>>(dB, dF, n, c0, nb)); }
          else { k_ldlt_dnext<<<1, 256, 0, sd>>>(dB, dF, n, c0, nb); ctx->launches++; VX_CUDA(cudaGetLastError()); }
          if (sd != st) VX_CUDA(cudaEventRecord(ctx->ldlt_ev[1], sd));
          VX_KERNEL(ctx, "k_ldlt_trailing", launch_chain(false, k_ldlt_trailing<false>, dim3(nt * (nt + 1) / 2), dim3(256), 0, st, dB, dF, n, c0, nb));
          if (sd != st) VX_CUDA(cudaStreamWaitEvent(st, ctx->ldlt_ev[1], 0));
        } else {
          VX_KERNEL(ctx, "k_ldlt_trailing", launch_chain(pdl, k_ldlt_trailing<true>, dim3(nt * (nt + 1) / 2), dim3(256), 0, st, dB, dF, n, c0, nb));
        }
      } else {
        const int nt = (below + TT - 1) / TT;
        VX_KERNEL(ctx, "k_ldlt_trailing_mma", launch_chain(pdl, k_ldlt_trailing_mma, dim3(nt * (nt + 1) / 2), dim3(256), TRAIL_SMEM, st, dB, dF, n, c0, nb));
      }
    }
  }
  for (int c0 = ((n - 1) / NB) * NB; c0 >= 0; c0 -= NB) {
    const int nb = std::min(NB, n - c0);
    VX_KERNEL(ctx, "k_ldlt_back", launch_chain(pdl, k_ldlt_back, dim3(std::max(1, blocks_for(c0, 8 * BC))), dim3(256), 0, st, (const double *)dB, (const double *)dF, n, c0, nb, dz, dxo));
  }
  return VX_OK;
}

// blocked factorisation of the permuted lower triangle in dB and x <- (L D L^T)^-1 x on the device; x: host, n doubles.
// The sweep is launch-bound towards its end (the trailing matrix shrinks to a few tiles), so it is captured once per (n, buffers)
// into a CUDA graph and replayed; under the kernel profiler the launches are issued one by one so that they can be timed.
static int factor_and_substitute(vx_ctx *ctx, double *dB, int n, double *x) {
  cudaStream_t st = ctx->stream;
  const int np = (n + NB - 1) / NB;
  VX_TRY(ctx->ldlt_side.reserve(sizeof(double) * ((size_t)NB * NB * np + 3 * (size_t)n)));
  double *dF = ctx->ldlt_side.as<double>();
  double *dx = dF + (size_t)NB * NB * np, *dz = dx + n, *dxo = dz + n;
  VX_CUDA(cudaMemcpyAsync(dx, x, sizeof(double) * n, cudaMemcpyHostToDevice, st));
  static const bool use_graph = std::getenv("VX_LDLT_NO_GRAPH") == nullptr;
  if (ctx->profiling || !use_graph) {
    VX_TRY(enqueue_sweep(ctx, dB, dF, n, dx, dz, dxo));
  } else {
    if (!ctx->ldlt_exec || ctx->ldlt_n != n || ctx->ldlt_B != dB || ctx->ldlt_F != dF) {
      if (ctx->ldlt_exec) { cudaGraphExecDestroy(ctx->ldlt_exec); ctx->ldlt_exec = nullptr; }
      cudaGraph_t g = nullptr;
      const int64_t before = ctx->launches;
      VX_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
      const int rc = enqueue_sweep(ctx, dB, dF, n, dx, dz, dxo);
      const cudaError_t ce = cudaStreamEndCapture(st, &g);
      ctx->ldlt_launches = ctx->launches - before;
      ctx->launches = before;
      if (rc != VX_OK) { if (g) cudaGraphDestroy(g); return rc; }
      VX_CUDA(ce);
      const cudaError_t ci = cudaGraphInstantiate(&ctx->ldlt_exec, g, 0);
      cudaGraphDestroy(g);
      VX_CUDA(ci);
      ctx->ldlt_n = n; ctx->ldlt_B = dB; ctx->ldlt_F = dF;
    }
    VX_CUDA(cudaGraphLaunch(ctx->ldlt_exec, st));
    ctx->launches += ctx->ldlt_launches;
  }
  VX_CUDA(cudaMemcpyAsync(x, dxo, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
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
  DevBuf dA, dP;
  DevBuf &dB = ctx->ldlt_mat;       // kept across calls: the captured sweep is replayed while its address stays
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
  dA.release(); dP.release();
  VX_TRY(rc);
  for (int k = n - 1; k >= 0; k--) if (tr[k] != k) std::swap(x[k], x[tr[k]]);
  return VX_OK;
}

// The LM step of one LARGE window on the device Hessian: dH (n x n, un-gauged, column-major) and dJ (n) stay where they are.
// dx = (gauge(H) + u diag(gauge(H))).ldlt().solve(-gauge(JacT)),  q1 = 0.5 dx^T (u D dx - JacT)   (voxel_map.hpp:452-470)
// nfix: leading unknowns held by the gauge (6 = pose 0 of the LiDAR-only / gravity optimizers, 15 = state 0 of LI_BA_Optimizer).
int lm_solve_large_device(vx_ctx *ctx, const double *dH, const double *dJ, double u, int n, double *dx, double *q1, int nfix) {
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  std::vector<double> D(n), J(n), d(n);
  std::vector<int32_t> pos, tr;
  DevBuf dP, dD;
  DevBuf &dB = ctx->ldlt_mat;
  auto body = [&]() -> int {
    VX_TRY(dB.reserve(sizeof(double) * (size_t)n * n)); VX_TRY(dP.reserve(sizeof(int32_t) * n)); VX_TRY(dD.reserve(sizeof(double) * n));
    auto T0 = std::chrono::steady_clock::now();
    auto lap = [&](const char *w) { if (getenv("VX_LDLT_TIMING")) { cudaStreamSynchronize(st); auto t = std::chrono::steady_clock::now(); fprintf(stderr, "[ldlt] %s %.3f ms\n", w, std::chrono::duration<double, std::milli>(t - T0).count()); T0 = t; } };
    VX_KERNEL(ctx, "k_diag_gauged", k_diag_gauged<<<blocks_for(n, 256), 256, 0, st>>>(dH, n, nfix, dD.as<double>()));
    VX_CUDA(cudaMemcpyAsync(D.data(), dD.p, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaMemcpyAsync(J.data(), dJ, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    for (int r = 0; r < nfix && r < n; r++) J[r] = 0.0;
    for (int r = 0; r < n; r++) d[r] = std::fabs(D[r] + u * D[r]);
    lap("diag");
    pivot_order(d, pos, tr);
    lap("pivot");
    for (int i = 0; i < n; i++) dx[i] = -J[i];
    for (int k = 0; k < n; k++) if (tr[k] != k) std::swap(dx[k], dx[tr[k]]);
    VX_CUDA(cudaMemcpyAsync(dP.p, pos.data(), sizeof(int32_t) * n, cudaMemcpyHostToDevice, st));
    VX_KERNEL(ctx, "k_permute_gauged", k_permute_gauged<<<dim3(blocks_for(n, 128), n), 128, 0, st>>>(dH, n, nfix, dP.as<int32_t>(), u, dB.as<double>()));
    lap("permute");
    const int r = factor_and_substitute(ctx, dB.as<double>(), n, dx);
    lap("factor");
    return r;
  };
  const int rc = body();
  cudaStreamSynchronize(st);
  dP.release(); dD.release();
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
