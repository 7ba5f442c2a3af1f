// vx_hess_small.cu — fused LidarFactor::acc_evaluate2 (voxel_map.hpp:150-282) for small windows
// (W = 10: LocalBA/win_size and the GBA wdsize, voxelslam.cpp:919 / :3033;  W = 20: config C3).
//
// One CTA evaluates one ITEM = a contiguous voxel range of one window and keeps that item's share
// of the window's 6W x 6W Hessian ON CHIP for the whole item:
//   phase A  one thread per voxel-frame: closed-form g1, g2, w (18 doubles) and the diagonal-block
//            corrections + gradient (30 doubles) -> shared memory, field-major, odd stride
//   phase B  the CTA's warps are ROLES; role r owns the block tiles t = r, r+NROLES, ... of the upper
//            block triangle.  For every voxel of the chunk (warp-uniform control flow) and every
//            owned tile (i,j) whose two frames are occupied, ONE FP64 tensor-core instruction
//            mma.sync.m8n8k4 adds the rank-3 block  sum_m S_m g_m,i g_m,j^T  (k = m, 3 of 4 used;
//            6 of 8 rows/cols used) into the tile's accumulator fragment, which stays in registers.
//            A- and B-fragment are the SAME shared-memory element (g_m[r] for lane (r,m)), so a voxel
//            costs k shared loads per warp; the 36-loads-per-108-FMA operand traffic that bounds a
//            CUDA-core formulation is gone.
//   phase C  fixed owner threads add the per-frame corrections / gradient in voxel order.
// The item's partial blocks go to HBM once (16 KB at W = 10); k_fin_small adds the items of a window
// in ascending order and mirrors the upper blocks (voxel_map.hpp:279-281).  No atomics anywhere:
// the result is bitwise reproducible.
#include "vx_factor.cuh"
#include "vx_math.cuh"
#include "vx_terms.cuh"
#include <algorithm>

using namespace vx;

namespace {

constexpr int NF_G = 18;   // g1[6] g2[6] w[6]
constexpr int NF_E = 30;   // rot-rot 9 | rot-tra 9 | tra-tra 6 (upper) | gradient 6
constexpr int NF = NF_G + NF_E;
constexpr int CHV = 32;    // max voxels per chunk

template <int W> struct TileTab {
  int i[W * (W + 1) / 2], j[W * (W + 1) / 2];
  constexpr TileTab() : i(), j() {
    int t = 0;
    for (int a = 0; a < W; a++)
      for (int b = a; b < W; b++) { i[t] = a; j[t] = b; t++; }
  }
};

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int W, int NROLES> struct SmallCfg {
  static constexpr int T = W * (W + 1) / 2;
  static constexpr int TPR = (T + NROLES - 1) / NROLES;   // tiles per role
  static constexpr int NT = 32 * NROLES;                  // threads per CTA = voxel-frames per chunk
  static constexpr int STRIDE = NT + 1;                   // odd: field-major rows land on distinct banks
  static constexpr int NDQ = (W * NF_E + NT - 1) / NT;    // (frame, component) sums owned per thread
  static constexpr size_t SMEM = sizeof(double) * ((size_t)NF * STRIDE + 4 * CHV) + sizeof(uint32_t) * 2 * CHV;
};

// phase B for one role: all owned tiles of one voxel
template <int W, int NROLES, int ROLE>
__device__ __forceinline__ void role_voxel(double (&C)[SmallCfg<W, NROLES>::TPR][2], const double (&a)[W], double sm, uint32_t mask) {
  constexpr TileTab<W> tab{};
  constexpr int T = SmallCfg<W, NROLES>::T, TPR = SmallCfg<W, NROLES>::TPR;
#pragma unroll
  for (int s = 0; s < TPR; s++) {
    const int t = s * NROLES + ROLE;
    if (t < T) {
      const int i = tab.i[t < T ? t : 0], j = tab.j[t < T ? t : 0];
      if (((mask >> i) & (mask >> j) & 1u) != 0u) dmma884(C[s][0], C[s][1], a[i], a[j] * sm);
    }
  }
}

template <int W, int NROLES, int ROLE>
__device__ __forceinline__ void role_store(const double (&C)[SmallCfg<W, NROLES>::TPR][2], double *__restrict__ part, int lane) {
  constexpr int T = SmallCfg<W, NROLES>::T, TPR = SmallCfg<W, NROLES>::TPR;
  const int row = lane >> 2, col = (lane & 3) * 2;
#pragma unroll
  for (int s = 0; s < TPR; s++) {
    const int t = s * NROLES + ROLE;
    if (t < T && row < 6 && col < 6) {
      part[(size_t)t * 36 + col * 6 + row] = C[s][0];
      part[(size_t)t * 36 + (col + 1) * 6 + row] = C[s][1];
    }
  }
}

template <int W, int NROLES, int ROLE>
__device__ __forceinline__ void role_chunk(double (&C)[SmallCfg<W, NROLES>::TPR][2], const double *__restrict__ Gs, const double *__restrict__ Ss,
                                           const uint32_t *__restrict__ mask_s, const uint32_t *__restrict__ base_s, int nv, int lane) {
  constexpr int STRIDE = SmallCfg<W, NROLES>::STRIDE;
  const int m = min(lane & 3, 2), r = min(lane >> 2, 5);
  const double *gl = Gs + (m * 6 + r) * STRIDE;
  for (int lv = 0; lv < nv; lv++) {
    const uint32_t mask = mask_s[lv];
    const uint32_t base = base_s[lv];
    const double sm = Ss[lv * 4 + (lane & 3)];
    double a[W];
#pragma unroll
    for (int f = 0; f < W; f++) {
      a[f] = 0.0;
      if ((mask >> f) & 1u) a[f] = gl[base + __popc(mask & ((1u << f) - 1u))];
    }
    role_voxel<W, NROLES, ROLE>(C, a, sm, mask);
  }
}

#define VX_ROLE_CASE(R, CALL) case R: if constexpr (R < NROLES) { CALL; } break;
#define VX_ROLE_SWITCH(role, FN, ...)                                   \
  switch (role) {                                                       \
    VX_ROLE_CASE(0, (FN<W, NROLES, 0>(__VA_ARGS__)))                     \
    VX_ROLE_CASE(1, (FN<W, NROLES, (1 < NROLES ? 1 : 0)>(__VA_ARGS__)))  \
    VX_ROLE_CASE(2, (FN<W, NROLES, (2 < NROLES ? 2 : 0)>(__VA_ARGS__)))  \
    VX_ROLE_CASE(3, (FN<W, NROLES, (3 < NROLES ? 3 : 0)>(__VA_ARGS__)))  \
    VX_ROLE_CASE(4, (FN<W, NROLES, (4 < NROLES ? 4 : 0)>(__VA_ARGS__)))  \
    VX_ROLE_CASE(5, (FN<W, NROLES, (5 < NROLES ? 5 : 0)>(__VA_ARGS__)))  \
    VX_ROLE_CASE(6, (FN<W, NROLES, (6 < NROLES ? 6 : 0)>(__VA_ARGS__)))  \
    VX_ROLE_CASE(7, (FN<W, NROLES, (7 < NROLES ? 7 : 0)>(__VA_ARGS__)))  \
    default: break;                                                     \
  }

// One CTA per item.  item -> (window w, split s): voxels [lo, hi) = the s-th of `splits` equal parts of
// [max(head, wv_off[w]), min(end, wv_off[w+1])).
template <int W, int NROLES>
__global__ void __launch_bounds__(32 * NROLES)
k_hess_small(const int32_t *__restrict__ off, const int32_t *__restrict__ frame, const int32_t *__restrict__ vox,
             const double *__restrict__ mom, const int32_t *__restrict__ cnt, int64_t nnz, const double *__restrict__ lam,
             const double *__restrict__ U, const double *__restrict__ mrg, const double *__restrict__ coe,
             const int32_t *__restrict__ wv_off, const double *__restrict__ poses, double *__restrict__ part_t,
             double *__restrict__ part_d, int64_t V, int64_t head, int64_t end, int splits) {
  using Cfg = SmallCfg<W, NROLES>;
  constexpr int NT = Cfg::NT, STRIDE = Cfg::STRIDE, T = Cfg::T, TPR = Cfg::TPR, NDQ = Cfg::NDQ;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *Gs = reinterpret_cast<double *>(smem_raw);            // [NF][STRIDE]
  double *Ss = Gs + (size_t)NF * STRIDE;                        // [CHV][4]  coe*{2/(l0-l1), 2/(l0-l2), -2/NN^2, 0}
  uint32_t *mask_s = reinterpret_cast<uint32_t *>(Ss + 4 * CHV);  // [CHV] occupied-frame bit mask
  uint32_t *base_s = mask_s + CHV;                              // [CHV] first record of the voxel in the chunk
  __shared__ int s_nv;

  const int tid = threadIdx.x, lane = tid & 31, role = tid >> 5;
  const int item = blockIdx.x;
  const int w = item / splits, sp = item - w * splits;
  int64_t lo = wv_off[w], hi = wv_off[w + 1];
  if (lo < head) lo = head;
  if (hi > end) hi = end;
  if (hi < lo) hi = lo;
  {
    const int64_t n = hi - lo;
    const int64_t a = lo + n * sp / splits, b = lo + n * (sp + 1) / splits;
    lo = a; hi = b;
  }
  const double *ps_w = poses + 12 * (int64_t)w * W;

  double C[TPR][2];
#pragma unroll
  for (int s = 0; s < TPR; s++) { C[s][0] = 0.0; C[s][1] = 0.0; }
  double dacc[NDQ];
#pragma unroll
  for (int q = 0; q < NDQ; q++) dacc[q] = 0.0;

  int64_t v0 = lo;
  while (v0 < hi) {
    // ---- chunk = the longest run of voxels from v0 with <= NT voxel-frames and <= CHV voxels
    const int32_t e0 = off[v0];
    {
      const int64_t cand = v0 + 1 + tid;
      const bool ok = tid < CHV && cand <= hi && off[cand] - e0 <= NT;
      const int nvv = __syncthreads_count(ok);
      if (tid == 0) s_nv = nvv;
    }
    if (tid < CHV) { mask_s[tid] = 0u; }
    __syncthreads();
    const int nv = s_nv;
    const int32_t e1 = off[v0 + nv];
    if (tid < nv) base_s[tid] = (uint32_t)(off[v0 + tid] - e0);
    // ---- phase A
    const int32_t e = e0 + tid;
    if (e < e1) {
      const int64_t v = vox[e];
      const int lv = (int)(v - v0);
      const int fr = frame[e] - w * W;
      atomicOr(&mask_s[lv], 1u << fr);
      const double *ps = ps_w + 12 * fr;
      const rot R = load_rot(ps);
      const d3 p = mk3(ps[9], ps[10], ps[11]);
      s33 P;
      P.xx = mom[0 * nnz + e]; P.xy = mom[1 * nnz + e]; P.xz = mom[2 * nnz + e];
      P.yy = mom[3 * nnz + e]; P.yz = mom[4 * nnz + e]; P.zz = mom[5 * nnz + e];
      const d3 vi = mk3(mom[6 * nnz + e], mom[7 * nnz + e], mom[8 * nnz + e]);
      const double n = (double)cnt[e];
      const d3 uk = mk3(U[0 * V + v], U[1 * V + v], U[2 * V + v]);
      const d3 u1 = mk3(U[3 * V + v], U[4 * V + v], U[5 * V + v]);
      const d3 u2 = mk3(U[6 * V + v], U[7 * V + v], U[8 * V + v]);
      const double NN = mrg[9 * V + v];
      const d3 vbar = mk3(mrg[6 * V + v] / NN, mrg[7 * V + v] / NN, mrg[8 * V + v] / NN);
      const double c = coe[v];
      if (e == off[v]) {   // first voxel-frame of the voxel also publishes the voxel's S (voxel_map.hpp:198-201, :265-268)
        const double l0 = lam[v], l1 = lam[V + v], l2 = lam[2 * V + v];
        Ss[lv * 4 + 0] = c * (2.0 / (l0 - l1));
        Ss[lv * 4 + 1] = c * (2.0 / (l0 - l2));
        Ss[lv * 4 + 2] = c * (-2.0 / NN / NN);
        Ss[lv * 4 + 3] = 0.0;
      }
      VfTerms t;
      vf_terms(R, p, P, vi, n, uk, u1, u2, vbar, NN, t);
      double *g = Gs + tid;
#pragma unroll
      for (int k = 0; k < NF_G; k++) g[k * STRIDE] = t.g[k];
      double *ee = g + NF_G * STRIDE;
#pragma unroll
      for (int cc = 0; cc < 3; cc++)
#pragma unroll
        for (int r = 0; r < 3; r++) {
          ee[(cc * 3 + r) * STRIDE] = c * t.E[cc * 6 + r];              // rot-rot (r, cc)
          ee[(9 + cc * 3 + r) * STRIDE] = c * t.E[(3 + cc) * 6 + r];    // rot-tra (r, 3+cc) == (3+cc, r)
        }
      ee[18 * STRIDE] = c * t.E[3 * 6 + 3]; ee[19 * STRIDE] = c * t.E[4 * 6 + 3]; ee[20 * STRIDE] = c * t.E[5 * 6 + 3];
      ee[21 * STRIDE] = c * t.E[4 * 6 + 4]; ee[22 * STRIDE] = c * t.E[5 * 6 + 4]; ee[23 * STRIDE] = c * t.E[5 * 6 + 5];
#pragma unroll
      for (int r = 0; r < 6; r++) ee[(24 + r) * STRIDE] = c * t.jj[r];
    }
    __syncthreads();
    // ---- phase B: tensor-core rank-3 block updates, one role per warp
    VX_ROLE_SWITCH(role, role_chunk, C, Gs, Ss, mask_s, base_s, nv, lane)
    // ---- phase C: per-frame corrections + gradient, fixed owner threads, voxel order
#pragma unroll
    for (int q = 0; q < NDQ; q++) {
      const int id = tid + q * NT;
      if (id < W * NF_E) {
        const int fr = id / NF_E, comp = id - fr * NF_E;
        const double *src = Gs + (NF_G + comp) * STRIDE;
        double acc = dacc[q];
        for (int lv = 0; lv < nv; lv++) {
          const uint32_t mask = mask_s[lv];
          if ((mask >> fr) & 1u) acc += src[base_s[lv] + __popc(mask & ((1u << fr) - 1u))];
        }
        dacc[q] = acc;
      }
    }
    __syncthreads();
    v0 += nv;
  }
  // ---- item partials
  double *pt = part_t + (size_t)item * T * 36;
  VX_ROLE_SWITCH(role, role_store, C, pt, lane)
  double *pd = part_d + (size_t)item * W * NF_E;
#pragma unroll
  for (int q = 0; q < NDQ; q++) {
    const int id = tid + q * NT;
    if (id < W * NF_E) pd[id] = dacc[q];
  }
}

// Adds the items of each window in ascending order and writes Hess (upper blocks + mirrored lower
// blocks, voxel_map.hpp:279-281) and JacT.  grid = (T, n_win), 64 threads.
template <int W>
__global__ void k_fin_small(const double *__restrict__ part_t, const double *__restrict__ part_d, double *__restrict__ H,
                            double *__restrict__ J, int splits) {
  constexpr int T = W * (W + 1) / 2;
  constexpr TileTab<W> tab{};
  __shared__ int ti[T], tj[T];
  for (int k = threadIdx.x; k < T; k += blockDim.x) { ti[k] = tab.i[k]; tj[k] = tab.j[k]; }
  __syncthreads();
  const int t = blockIdx.x, w = blockIdx.y, x = threadIdx.x;
  const int i = ti[t], j = tj[t];
  const int64_t jl = 6 * W;
  double *Hw = H + (size_t)w * jl * jl;
  if (x < 36) {
    const int r = x % 6, c = x / 6;
    double s = 0.0;
    for (int sp = 0; sp < splits; sp++) s += part_t[((size_t)(w * splits + sp) * T + t) * 36 + x];
    if (i == j) {
      // corrections: rot-rot 9 | rot-tra 9 (placed at (r,3+c) and (3+c,r)) | tra-tra upper 6
      int comp;
      if (r < 3 && c < 3) comp = c * 3 + r;
      else if (r < 3) comp = 9 + (c - 3) * 3 + r;
      else if (c < 3) comp = 9 + (r - 3) * 3 + c;
      else {
        const int a = min(r, c) - 3, b = max(r, c) - 3;   // upper index (a <= b): 00 01 02 11 12 22
        comp = 18 + (a == 0 ? b : (a == 1 ? 2 + b : 5));
      }
      double e = 0.0;
      for (int sp = 0; sp < splits; sp++) e += part_d[(size_t)(w * splits + sp) * W * NF_E + i * NF_E + comp];
      Hw[(6 * i + c) * jl + (6 * i + r)] = s + e;
    } else {
      Hw[(6 * j + c) * jl + (6 * i + r)] = s;
      Hw[(6 * i + r) * jl + (6 * j + c)] = s;
    }
  } else if (x < 42 && i == j) {
    const int r = x - 36;
    double g = 0.0;
    for (int sp = 0; sp < splits; sp++) g += part_d[(size_t)(w * splits + sp) * W * NF_E + i * NF_E + 24 + r];
    J[(size_t)w * jl + 6 * i + r] = g;
  }
}

template <int W, int NROLES>
int launch_small(vx_ctx *ctx, vx_factor *f, int64_t head, int64_t end) {
  using Cfg = SmallCfg<W, NROLES>;
  cudaStream_t st = ctx->stream;
  // items: enough CTAs to fill the machine several times over, but >= ~96 voxels per item
  int splits = 1;
  {
    const int64_t want_items = (int64_t)ctx->sm_count * 8;
    const int64_t vox_per_win = std::max<int64_t>(1, f->V / f->n_win);
    int64_t s = (want_items + f->n_win - 1) / f->n_win;
    s = std::min<int64_t>(s, std::max<int64_t>(1, vox_per_win / 96));
    splits = (int)std::max<int64_t>(1, s);
  }
  const int64_t items = (int64_t)f->n_win * splits;
  VX_TRY(f->part_t.reserve(sizeof(double) * items * Cfg::T * 36));
  VX_TRY(f->part_d.reserve(sizeof(double) * items * W * NF_E));
  static bool attr_set = false;
  if (!attr_set) {
    VX_CUDA(cudaFuncSetAttribute(k_hess_small<W, NROLES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM));
    attr_set = true;
  }
  VX_KERNEL(ctx, "k_hess_small", k_hess_small<W, NROLES><<<(unsigned)items, Cfg::NT, Cfg::SMEM, st>>>(
      f->off.as<int32_t>(), f->frame.as<int32_t>(), f->vox.as<int32_t>(), f->mom.as<double>(), f->cnt.as<int32_t>(), f->nnz,
      f->lam.as<double>(), f->U.as<double>(), f->mrg.as<double>(), f->coe.as<double>(), f->wv_off.as<int32_t>(),
      f->poses.as<double>(), f->part_t.as<double>(), f->part_d.as<double>(), f->V, head, end, splits));
  VX_KERNEL(ctx, "k_fin_small", k_fin_small<W><<<dim3(Cfg::T, f->n_win), 64, 0, st>>>(f->part_t.as<double>(), f->part_d.as<double>(),
                                                                                 f->H.as<double>(), f->J.as<double>(), splits));
  return VX_OK;
}

}  // namespace

namespace vx {
// Returns VX_OK and sets *handled when W has a fused instantiation.  Hess/JacT are fully overwritten.
int hessian_small(vx_ctx *ctx, vx_factor *f, int64_t head, int64_t end, bool *handled) {
  *handled = true;
  switch (f->W) {
    case 10: return launch_small<10, 4>(ctx, f, head, end);
    case 20: return launch_small<20, 8>(ctx, f, head, end);
    default: *handled = false; return VX_OK;
  }
}
}  // namespace vx
