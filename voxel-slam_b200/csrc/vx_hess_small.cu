// vx_hess_small.cu — fused LidarFactor::acc_evaluate2 (voxel_map.hpp:150-282) for small windows
// (W = 10: LocalBA/win_size and the GBA wdsize, voxelslam.cpp:919 / :3033;  W = 20: config C3).
//
// The rank-3 part of every block of the window Hessian is  Hess(i,j) = sum_v sum_m S_m g_m,i g_m,j^T
// (vx_factor.cu header).  Stack the rows g_m of all voxels of a window into one (3V) x (6W) matrix A,
// absent frames zero: the whole upper block triangle is the symmetric rank-3V update  A^T diag(S) A,
// a tall-skinny SYRK.  One CTA evaluates one ITEM (a contiguous voxel range of one window) and streams
// it through shared memory in chunks:
//   phase A  one thread per voxel-frame: closed-form g1, g2, w (18 doubles) -> rows 3v..3v+2 of the
//            chunk's A panel in shared memory; diagonal-block corrections + gradient (30 doubles) ->
//            a compact field-major record
//   phase B  FP64 tensor cores: the 6W x 6W accumulator is cut into 8x8 tiles, each warp (ROLE) owns a
//            contiguous run of upper tiles and keeps them in registers for the whole item.  Per 4 panel
//            rows a warp loads one fragment per tile row/column it needs (conflict-free, leading
//            dimension = 4 mod 16) and issues one mma.sync.m8n8k4.f64 per owned tile: no per-voxel
//            control flow, all of m, n and k used.
//   phase C  fixed owner threads add the per-frame corrections / gradient in voxel order.
// The item's partial tiles go to HBM once; k_fin_small adds the items of a window in ascending order,
// writes the upper blocks and mirrors them (voxel_map.hpp:279-281).  No atomics: bitwise reproducible.
#include "vx_factor.cuh"
#include "vx_math.cuh"
#include "vx_terms.cuh"
#include <algorithm>
#include <cstdlib>

using namespace vx;

namespace {

constexpr int NF_E = 30;   // rot-rot 9 | rot-tra 9 | tra-tra 6 (upper) | gradient 6
constexpr int CHV = 21;    // max voxels per chunk (3 * 21 = 63 panel rows <= ROWS)
constexpr int ROWS = 64;   // panel rows per chunk

template <int W, int NROLES> struct SmallCfg {
  static constexpr int NC = 6 * W;
  static constexpr int NT8 = (NC + 7) / 8;                 // 8-wide tile columns
  static constexpr int LD = NT8 * 8 + 4;                   // leading dimension: 4 or 12 mod 16 -> conflict-free fragments
  static constexpr int T8 = NT8 * (NT8 + 1) / 2;           // upper tiles
  static constexpr int TPR = (T8 + NROLES - 1) / NROLES;   // tiles per role
  static constexpr int NT = 32 * NROLES;                   // threads per CTA = voxel-frames per chunk
  static constexpr int STRIDE = NT + 1;                    // record NT is an all-zero record
  static constexpr int NDQ = (W * NF_E + NT - 1) / NT;     // (frame, component) sums owned per thread
  // record table: per frame 24 entries (CHV = 21 chunk-local voxels, padded to whole 64-bit words) holding the record index of
  // (frame, voxel) or NT = absent; one byte per entry when NT <= 128, two otherwise
  static constexpr int RBYTES = NT <= 128 ? 1 : 2;
  static constexpr int RB = 24 * RBYTES;                   // bytes per frame
  static constexpr uint32_t RFILL = NT <= 128 ? (uint32_t)NT * 0x01010101u : (uint32_t)NT * 0x00010001u;
  static constexpr int NVF = 17;                            // per-voxel fields staged per chunk: U 9 | S 3, vBar 3, 1/NN | coe
  static constexpr size_t SMEM = sizeof(double) * ((size_t)ROWS * LD + (size_t)NF_E * STRIDE + ROWS + 12 * W + NVF * CHV) + (size_t)W * RB + 16;
};

template <int NT8> struct TileTab8 {
  int I[NT8 * (NT8 + 1) / 2], J[NT8 * (NT8 + 1) / 2];
  constexpr TileTab8() : I(), J() {
    int t = 0;
    for (int a = 0; a < NT8; a++)
      for (int b = a; b < NT8; b++) { I[t] = a; J[t] = b; t++; }
  }
};
// does role r need the fragment of tile column b?
template <int W, int NROLES> struct NeedTab {
  bool need[NROLES][SmallCfg<W, NROLES>::NT8];
  constexpr NeedTab() : need() {
    constexpr int NT8 = SmallCfg<W, NROLES>::NT8, T8 = SmallCfg<W, NROLES>::T8, TPR = SmallCfg<W, NROLES>::TPR;
    TileTab8<NT8> tab{};
    for (int r = 0; r < NROLES; r++)
      for (int s = 0; s < TPR; s++) {
        int t = r * TPR + s;
        if (t < T8) { need[r][tab.I[t]] = true; need[r][tab.J[t]] = true; }
      }
  }
};

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// phase B of one role over the chunk's panel (nks k-steps of 4 rows)
template <int W, int NROLES, int ROLE>
__device__ __forceinline__ void role_chunk(double (&C)[SmallCfg<W, NROLES>::TPR][2], const double *__restrict__ As,
                                           const double *__restrict__ Srow, int nks, int lane) {
  using Cfg = SmallCfg<W, NROLES>;
  constexpr int NT8 = Cfg::NT8, T8 = Cfg::T8, TPR = Cfg::TPR, LD = Cfg::LD;
  constexpr TileTab8<NT8> tab{};
  constexpr NeedTab<W, NROLES> nd{};
  const double *ap = As + (lane & 3) * LD + (lane >> 2);
  const double *sp = Srow + (lane & 3);
#pragma unroll 2
  for (int ks = 0; ks < nks; ks++) {
    const double s = sp[4 * ks];
    double a[NT8], b[NT8];
#pragma unroll
    for (int x = 0; x < NT8; x++) {
      a[x] = 0.0; b[x] = 0.0;
      if (nd.need[ROLE][x]) { a[x] = ap[(size_t)ks * 4 * LD + 8 * x]; b[x] = a[x] * s; }
    }
#pragma unroll
    for (int q = 0; q < TPR; q++) {
      const int t = ROLE * TPR + q;
      if (t < T8) dmma884(C[q][0], C[q][1], a[tab.I[t < T8 ? t : 0]], b[tab.J[t < T8 ? t : 0]]);
    }
  }
}

// accumulator fragments -> the item's partial tiles (8x8 row-major each)
template <int W, int NROLES, int ROLE>
__device__ __forceinline__ void role_store(const double (&C)[SmallCfg<W, NROLES>::TPR][2], double *__restrict__ part, int lane) {
  constexpr int T8 = SmallCfg<W, NROLES>::T8, TPR = SmallCfg<W, NROLES>::TPR;
#pragma unroll
  for (int q = 0; q < TPR; q++) {
    const int t = ROLE * TPR + q;
    if (t < T8) *reinterpret_cast<double2 *>(part + (size_t)t * 64 + (lane >> 2) * 8 + (lane & 3) * 2) = make_double2(C[q][0], C[q][1]);
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

// Phase A of one voxel-frame: vf_terms (vx_terms.cuh) with every result stored where phases B / C read it, the common factors
// coe * 2 / NN folded once (cs2) and nothing kept alive longer than needed.  g: panel rows 3lv..3lv+2 at columns 6fr..;
// ee: this record's column of Es.
template <int LD, int STRIDE>
__device__ __forceinline__ void vf_emit(const rot &R, d3 p, const s33 &P, d3 v, double n, d3 uk, d3 u1, d3 u2, d3 vbar, double iNN,
                                        double cc, double *__restrict__ g, double *__restrict__ ee) {
  const double cs2 = cc * (2.0 * iNN);
  const d3 a = rtmul(R, uk), b1 = rtmul(R, u1), b2 = rtmul(R, u2);
  const d3 Pa = smul(P, a);
  const d3 t = p - vbar;
  const double c = dot(uk, t);
  const d3 combo2 = axpy(n, t, rmul(R, v));
  const d3 va = cross(v, a);                       // viRiTuk
  const double c2uk = dot(combo2, uk);
  // gradient coe * Auk^T uk (voxel_map.hpp:231-237) and half of its rotation part (the antisymmetric correction below)
  const d3 gj = cs2 * axpy(c, va, cross(Pa, a));
  const d3 hj = 0.5 * gj;
  const double ct = cs2 * c2uk;
  ee[24 * STRIDE] = gj.x; ee[25 * STRIDE] = gj.y; ee[26 * STRIDE] = gj.z;
  ee[27 * STRIDE] = ct * uk.x; ee[28 * STRIDE] = ct * uk.y; ee[29 * STRIDE] = ct * uk.z;
  g[2 * LD + 0] = va.x; g[2 * LD + 1] = va.y; g[2 * LD + 2] = va.z;
  g[2 * LD + 3] = n * uk.x; g[2 * LD + 4] = n * uk.y; g[2 * LD + 5] = n * uk.z;
  {
    const d3 q = axpy(dot(t, u1), v, smul(P, b1));
    const d3 gr = iNN * (cross(Pa, b1) - cross(a, q) + c * cross(v, b1));
    const d3 gt = iNN * axpy(c2uk, u1, dot(combo2, u1) * uk);
    g[0] = gr.x; g[1] = gr.y; g[2] = gr.z; g[3] = gt.x; g[4] = gt.y; g[5] = gt.z;
  }
  {
    const d3 q = axpy(dot(t, u2), v, smul(P, b2));
    const d3 gr = iNN * (cross(Pa, b2) - cross(a, q) + c * cross(v, b2));
    const d3 gt = iNN * axpy(c2uk, u2, dot(combo2, u2) * uk);
    g[LD + 0] = gr.x; g[LD + 1] = gr.y; g[LD + 2] = gr.z; g[LD + 3] = gt.x; g[LD + 4] = gt.y; g[LD + 5] = gt.z;
  }
  // diagonal-block corrections (vx_terms.cuh: E), times coe.  rot-rot: component cc*3 + r <-> E(r, cc)
  {
    const d3 q = axpy(c, v, Pa);
    const double qa = dot(q, a);
    const d3 p0 = mk3(P.xx, P.xy, P.xz), p1 = mk3(P.xy, P.yy, P.yz), p2 = mk3(P.xz, P.yz, P.zz);
    const d3 c0 = cross(a, p0), c1 = cross(a, p1), c2 = cross(a, p2);
    const d3 m0 = mk3(c0.x, c1.x, c2.x), m1 = mk3(c0.y, c1.y, c2.y), m2 = mk3(c0.z, c1.z, c2.z);
    const d3 r0 = cross(m0, a), r1 = cross(m1, a), r2 = cross(m2, a);
    ee[0 * STRIDE] = cs2 * (a.x * q.x - qa - r0.x);
    ee[1 * STRIDE] = fma(cs2, a.y * q.x - r1.x, -hj.z);
    ee[2 * STRIDE] = fma(cs2, a.z * q.x - r2.x, hj.y);
    ee[3 * STRIDE] = fma(cs2, a.x * q.y - r0.y, hj.z);
    ee[4 * STRIDE] = cs2 * (a.y * q.y - qa - r1.y);
    ee[5 * STRIDE] = fma(cs2, a.z * q.y - r2.y, -hj.x);
    ee[6 * STRIDE] = fma(cs2, a.x * q.z - r0.z, -hj.y);
    ee[7 * STRIDE] = fma(cs2, a.y * q.z - r1.z, hj.x);
    ee[8 * STRIDE] = cs2 * (a.z * q.z - qa - r2.z);
  }
  // rot-tra: component 9 + cc*3 + r <-> E(r, 3+cc) = (2/NN) va_r uk_cc
  {
    const d3 sva = cs2 * va;
    ee[9 * STRIDE] = sva.x * uk.x;  ee[10 * STRIDE] = sva.y * uk.x; ee[11 * STRIDE] = sva.z * uk.x;
    ee[12 * STRIDE] = sva.x * uk.y; ee[13 * STRIDE] = sva.y * uk.y; ee[14 * STRIDE] = sva.z * uk.y;
    ee[15 * STRIDE] = sva.x * uk.z; ee[16 * STRIDE] = sva.y * uk.z; ee[17 * STRIDE] = sva.z * uk.z;
  }
  // tra-tra upper 00 01 02 11 12 22: (2/NN) n uk_r uk_c
  {
    const d3 su = (cs2 * n) * uk;
    ee[18 * STRIDE] = su.x * uk.x; ee[19 * STRIDE] = su.x * uk.y; ee[20 * STRIDE] = su.x * uk.z;
    ee[21 * STRIDE] = su.y * uk.y; ee[22 * STRIDE] = su.y * uk.z; ee[23 * STRIDE] = su.z * uk.z;
  }
}

// per-voxel-frame operands of phase A (76 B of the cluster record)
struct VfRow {
  s33 P; d3 vi; int32_t cnt, frame;
  __device__ __forceinline__ void load(const int32_t *__restrict__ frame_, const double *__restrict__ mom, const int32_t *__restrict__ cnt_,
                                       int64_t nnz, int64_t e) {
    frame = frame_[e]; cnt = cnt_[e];
    P.xx = mom[0 * nnz + e]; P.xy = mom[1 * nnz + e]; P.xz = mom[2 * nnz + e];
    P.yy = mom[3 * nnz + e]; P.yz = mom[4 * nnz + e]; P.zz = mom[5 * nnz + e];
    vi = mk3(mom[6 * nnz + e], mom[7 * nnz + e], mom[8 * nnz + e]);
  }
};

// cp.async the NVF per-voxel fields of voxels [vbeg, vbeg + CHV) (clamped to the last voxel) into s_vox[NVF][CHV]
template <int NVF, int NT>
__device__ __forceinline__ void stage_voxels(double *s_vox, const double *__restrict__ U, const double *__restrict__ prep,
                                             const double *__restrict__ coe, int64_t V, int64_t vbeg, int tid) {
  for (int idx = tid; idx < NVF * CHV; idx += NT) {
    const int fld = idx / CHV, j = idx - fld * CHV;
    const double *base = fld < 9 ? U + (int64_t)fld * V : (fld < 16 ? prep + (int64_t)(fld - 9) * V : coe);
    int64_t v = vbeg + j;
    if (v >= V) v = V - 1;
    const unsigned dst = (unsigned)__cvta_generic_to_shared(s_vox + idx);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(base + v) : "memory");
  }
}

// One CTA per item.  item -> (window w, split s): voxels [lo, hi) = the s-th of `splits` equal parts of
// [max(head, wv_off[w]), min(end, wv_off[w+1])).
template <int W, int NROLES, int MINB>
__global__ void __launch_bounds__(32 * NROLES, MINB)
k_hess_small(const int32_t *__restrict__ off, const int32_t *__restrict__ frame, const double *__restrict__ mom,
             const int32_t *__restrict__ cnt, int64_t nnz, const double *__restrict__ prep, const double *__restrict__ U,
             const double *__restrict__ coe, const int32_t *__restrict__ wv_off, const double *__restrict__ poses,
             double *__restrict__ part_t, double *__restrict__ part_d, int64_t V, int64_t head, int64_t end, int splits, int w_base) {
  using Cfg = SmallCfg<W, NROLES>;
  constexpr int NT = Cfg::NT, STRIDE = Cfg::STRIDE, LD = Cfg::LD, T8 = Cfg::T8, TPR = Cfg::TPR, NDQ = Cfg::NDQ, RB = Cfg::RB;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double *As = reinterpret_cast<double *>(smem_raw);        // [ROWS][LD] the chunk's panel of A
  double *Es = As + (size_t)ROWS * LD;                      // [NF_E][STRIDE] corrections + gradient per record
  double *Srow = Es + (size_t)NF_E * STRIDE;                // [ROWS] coe*{2/(l0-l1), 2/(l0-l2), -2/NN^2} per panel row
  double *s_pose = Srow + ROWS;                             // [W][12] the window's poses
  double *s_vox = s_pose + 12 * W;                          // [NVF][CHV] per-voxel operands of the chunk (cp.async, one chunk ahead)
  uint8_t *recb = reinterpret_cast<uint8_t *>(s_vox + Cfg::NVF * CHV);  // [W][24] record of (frame, chunk-local voxel), NT = absent
  __shared__ int s_off[CHV + 1];     // off[v0 + 1 + t] of the chunk
  __shared__ uint8_t s_lv[NT];       // chunk-local voxel of each voxel-frame

  const int tid = threadIdx.x, lane = tid & 31, role = tid >> 5;
  const int item = w_base * splits + blockIdx.x;
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
  for (int k = tid; k < 12 * W; k += NT) s_pose[k] = poses[12 * (int64_t)w * W + k];
  if (tid < NF_E) Es[tid * STRIDE + NT] = 0.0;   // the all-zero record

  double C[TPR][2];
#pragma unroll
  for (int s = 0; s < TPR; s++) { C[s][0] = 0.0; C[s][1] = 0.0; }
  double dacc[NDQ];
#pragma unroll
  for (int q = 0; q < NDQ; q++) dacc[q] = 0.0;

  // chunk = the longest run of voxels from v0 with <= NT voxel-frames and <= CHV voxels.  Everything chunk n+1 needs is on its
  // way while chunk n is multiplied: its voxel run (offsets loaded at the top of chunk n, counted at the closing barrier), its
  // per-voxel-frame operands (registers pf, issued after phase A) and its per-voxel operands (cp.async into s_vox, ditto) —
  // rows e1.. and voxels v0+nv.. follow the chunk contiguously in every field array.
  int64_t v0 = lo;
  int32_t e0 = 0;
  int nv = 0;
  const int32_t e_last = off[hi] - 1;        // last voxel-frame of the item (clamp for look-ahead rows)
  VfRow pf;
  if (v0 < hi) {
    e0 = off[v0];
    const int64_t cand = v0 + 1 + tid;
    bool ok = false;
    if (tid < CHV && cand <= hi) { const int32_t o = off[cand]; ok = o - e0 <= NT; s_off[tid] = o; }
    pf.load(frame, mom, cnt, nnz, min(e0 + tid, e_last));
    stage_voxels<Cfg::NVF, NT>(s_vox, U, prep, coe, V, v0, tid);
    nv = __syncthreads_count(ok);
  }
  while (v0 < hi) {
    const int32_t e1 = s_off[nv - 1];
    const int64_t v0n = v0 + nv, candn = v0n + 1 + tid;
    int32_t on = 0;
    if (tid < CHV && candn <= hi) on = off[candn];
    {
      double2 *z = reinterpret_cast<double2 *>(As);
      for (int k = tid; k < ROWS * LD / 2; k += NT) z[k] = make_double2(0.0, 0.0);
      if (tid < ROWS) Srow[tid] = 0.0;
      for (int k = tid; k < W * RB / 4; k += NT) reinterpret_cast<uint32_t *>(recb)[k] = Cfg::RFILL;
      if (tid < nv) {   // voxel tid of the chunk owns voxel-frames [off[v0+tid], off[v0+tid+1]) - e0
        const int a = tid == 0 ? 0 : s_off[tid - 1] - e0, b = s_off[tid] - e0;
        for (int j = a; j < b; j++) s_lv[j] = (uint8_t)tid;
      }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();
    // ---- phase A: one thread per voxel-frame
    const int32_t e = e0 + tid;
    if (e < e1) {
      const int lv = s_lv[tid];
      const int fr = pf.frame - w * W;
      const double *sv = s_vox + lv;
      const d3 uk = mk3(sv[0 * CHV], sv[1 * CHV], sv[2 * CHV]);
      const d3 u1 = mk3(sv[3 * CHV], sv[4 * CHV], sv[5 * CHV]);
      const d3 u2 = mk3(sv[6 * CHV], sv[7 * CHV], sv[8 * CHV]);
      const d3 vbar = mk3(sv[12 * CHV], sv[13 * CHV], sv[14 * CHV]);
      const double iNN = sv[15 * CHV];
      const double c = sv[16 * CHV];
      const double *ps = s_pose + 12 * fr;
      const rot R = load_rot(ps);
      const d3 p = mk3(ps[9], ps[10], ps[11]);
      if (Cfg::RBYTES == 1) recb[fr * RB + lv] = (uint8_t)tid;
      else reinterpret_cast<uint16_t *>(recb)[fr * 24 + lv] = (uint16_t)tid;
      // every voxel-frame of the voxel publishes the voxel's S (identical values; voxel_map.hpp:198-201, :265-268)
      Srow[3 * lv + 0] = sv[9 * CHV]; Srow[3 * lv + 1] = sv[10 * CHV]; Srow[3 * lv + 2] = sv[11 * CHV];
      vf_emit<LD, STRIDE>(R, p, pf.P, pf.vi, (double)pf.cnt, uk, u1, u2, vbar, iNN, c, As + (size_t)(3 * lv) * LD + 6 * fr, Es + tid);
    }
    __syncthreads();
    // ---- look ahead: operands of chunk n+1
    if (v0n < hi) {
      pf.load(frame, mom, cnt, nnz, min(e1 + tid, e_last));
      stage_voxels<Cfg::NVF, NT>(s_vox, U, prep, coe, V, v0n, tid);
    }
    // ---- phase B: tensor-core SYRK of the panel, one role per warp
    const int nks = (3 * nv + 3) >> 2;
    VX_ROLE_SWITCH(role, role_chunk, C, As, Srow, nks, lane)
    // ---- phase C: per-frame corrections + gradient.  The owner thread of (frame, component) reads the frame's CHV record
    //      bytes (absent -> the zero record) and adds the records in four interleaved chains.
#pragma unroll
    for (int q = 0; q < NDQ; q++) {
      const int id = tid + q * NT;
      if (id < W * NF_E) {
        const int fr = id / NF_E, comp = id - fr * NF_E;
        const double *src = Es + comp * STRIDE;
        const unsigned long long *rw = reinterpret_cast<const unsigned long long *>(recb + fr * RB);
        constexpr int EPW = 8 / Cfg::RBYTES, NWD = 24 / EPW;     // entries per word, words per frame
        unsigned long long wd[NWD];
#pragma unroll
        for (int k = 0; k < NWD; k++) wd[k] = rw[k];
        double sacc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
        for (int k = 0; k < CHV; k++) {
          const unsigned idx = (unsigned)(wd[k / EPW] >> (8 * Cfg::RBYTES * (k % EPW))) & (Cfg::RBYTES == 1 ? 0xffu : 0xffffu);
          sacc[k & 3] += src[idx];
        }
        dacc[q] += (sacc[0] + sacc[1]) + (sacc[2] + sacc[3]);
      }
    }
    // ---- close the chunk and fix the next one (s_off was last read before the previous barrier)
    {
      bool ok = false;
      if (tid < CHV && candn <= hi) { ok = on - e1 <= NT; s_off[tid] = on; }
      nv = __syncthreads_count(ok);
      v0 = v0n; e0 = e1;
    }
  }
  // ---- item partials
  double *pt = part_t + (size_t)item * T8 * 64;
  VX_ROLE_SWITCH(role, role_store, C, pt, lane)
  double *pd = part_d + (size_t)item * W * NF_E;
#pragma unroll
  for (int q = 0; q < NDQ; q++) {
    const int id = tid + q * NT;
    if (id < W * NF_E) pd[id] = dacc[q];
  }
}

// Adds the items of each window in ascending order and writes Hess (upper blocks + mirrored lower
// blocks, voxel_map.hpp:279-281) and JacT.  One CTA per window.
// When `cost` is given, the y == 1 CTA of a window also reduces the window's cost sum coe * lambda_0 over [head, end) — the same
// 256-thread strided sums and tree as k_cost (vx_factor.cu), so the value is bitwise k_cost's — which saves that launch.
template <int W, int NT8>
__global__ void __launch_bounds__(256) k_fin_small(const double *__restrict__ part_t, const double *__restrict__ part_d,
                                                   double *__restrict__ H, double *__restrict__ J, int splits, int w_base,
                                                   const double *__restrict__ lam, const double *__restrict__ coe,
                                                   const int32_t *__restrict__ wv_off, double *__restrict__ cost, int64_t head, int64_t end) {
  constexpr int T = W * (W + 1) / 2, T8 = NT8 * (NT8 + 1) / 2;
  const int w = w_base + blockIdx.x;
  const int64_t jl = 6 * W;
  double *Hw = H + (size_t)w * jl * jl;
  const double *pt = part_t + (size_t)w * splits * T8 * 64;
  const double *pd = part_d + (size_t)w * splits * W * NF_E;
  for (int idx = blockIdx.y * blockDim.x + threadIdx.x; idx < T * 36; idx += blockDim.x * gridDim.y) {
    const int t = idx / 36, x = idx - t * 36;
    int i = 0, rem = t;                      // t-th upper block (i, j), row-major
    while (rem >= W - i) { rem -= W - i; i++; }
    const int j = i + rem;
    const int r = x % 6, c = x / 6;
    // element (R, Cc) of the symmetric rank-3 sum; only upper 8x8 tiles are stored
    int R = 6 * i + r, Cc = 6 * j + c;
    if ((R >> 3) > (Cc >> 3)) { const int tmp = R; R = Cc; Cc = tmp; }
    const int I = R >> 3, Jt = Cc >> 3;
    const int t8 = I * NT8 - I * (I - 1) / 2 + (Jt - I);
    const size_t eo = (size_t)t8 * 64 + (R & 7) * 8 + (Cc & 7);
    double s = 0.0;
    for (int sp = 0; sp < splits; sp++) s += pt[(size_t)sp * T8 * 64 + eo];
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
      for (int sp = 0; sp < splits; sp++) e += pd[(size_t)sp * W * NF_E + i * NF_E + comp];
      Hw[(6 * i + c) * jl + (6 * i + r)] = s + e;
    } else {
      Hw[(6 * j + c) * jl + (6 * i + r)] = s;
      Hw[(6 * i + r) * jl + (6 * j + c)] = s;
    }
  }
  if (blockIdx.y == 1 && cost != nullptr) {
    __shared__ double sh[256];
    int64_t b = wv_off[w], e = wv_off[w + 1];
    if (b < head) b = head;
    if (e > end) e = end;
    double acc = 0.0;
    for (int64_t v = b + threadIdx.x; v < e; v += 256) acc += coe[v] * lam[v];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
      if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
      __syncthreads();
    }
    if (threadIdx.x == 0) cost[w] = sh[0];
  }
  if (blockIdx.y != 0) return;
  for (int idx = threadIdx.x; idx < 6 * W; idx += blockDim.x) {
    const int i = idx / 6, r = idx - i * 6;
    double g = 0.0;
    for (int sp = 0; sp < splits; sp++) g += pd[(size_t)sp * W * NF_E + i * NF_E + 24 + r];
    J[(size_t)w * jl + idx] = g;
  }
}


template <int W, int NROLES, int MINB>
int launch_small(vx_ctx *ctx, vx_factor *f, int64_t head, int64_t end, int w_lo, int w_hi, bool with_cost) {
  using Cfg = SmallCfg<W, NROLES>;
  cudaStream_t st = ctx->stream;
  // items: enough CTAs to fill the machine several times over, but >= ~96 voxels per item
  int splits = 1;
  {
    // ~3.6 items per resident CTA slot (3 CTAs per SM): measured best on the bench workload (splits 3 -> 4: -7 %), the last
    // round of CTAs then ends nearly together
    const int64_t want_items = (int64_t)(ctx->sm_count * 3 * 3.6);
    const int64_t vox_per_win = std::max<int64_t>(1, f->V / f->n_win);
    int64_t s = (want_items + f->n_win / 2) / f->n_win;
    s = std::min<int64_t>(s, std::max<int64_t>(1, vox_per_win / 96));
    splits = (int)std::max<int64_t>(1, s);
    // Batched regime: a FIXED number of items per window.  The items' partial tiles are added in item order, so the split decides
    // how a window's sum is associated; with the adaptive rule above a window's Hessian changed in the last bits with the number
    // of windows that happened to share its launch (400 on one GPU, 50 per rank on eight), and the top level of a hierarchical
    // job turns such bits into a handful of edges on the other side of its |hess| < 1e-6 test.  4 is what the rule gives for the
    // 400-window bottom level (measured best there).
    if (f->n_win >= 4) splits = 4;      // one to three windows (local BA): the adaptive rule above, for latency
    static const int forced = getenv("VX_HS_SPLITS") ? atoi(getenv("VX_HS_SPLITS")) : 0;
    if (forced > 0) splits = forced;
  }
  const int64_t items = (int64_t)f->n_win * splits;     // partials are indexed by the global item
  VX_TRY(f->part_t.reserve(sizeof(double) * items * Cfg::T8 * 64));
  VX_TRY(f->part_d.reserve(sizeof(double) * items * W * NF_E));
  {   // the opt-in is per DEVICE, not per process: one flag per device ordinal
    static bool attr_set[64] = {};
    const int d = ctx->device & 63;
    if (!attr_set[d]) {
      VX_CUDA(cudaFuncSetAttribute(k_hess_small<W, NROLES, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM));
      attr_set[d] = true;
    }
  }
  VX_KERNEL(ctx, "k_hess_small", k_hess_small<W, NROLES, MINB><<<(unsigned)((w_hi - w_lo) * splits), Cfg::NT, Cfg::SMEM, st>>>(
      f->off.as<int32_t>(), f->frame.as<int32_t>(), f->mom.as<double>(), f->cnt.as<int32_t>(), f->nnz,
      f->cS.as<double>(), f->U.as<double>(), f->coe.as<double>(), f->wv_off.as<int32_t>(),
      f->poses.as<double>(), f->part_t.as<double>(), f->part_d.as<double>(), f->V, head, end, splits, w_lo));
  VX_KERNEL(ctx, "k_fin_small", k_fin_small<W, Cfg::NT8><<<dim3(w_hi - w_lo, 4), 256, 0, st>>>(f->part_t.as<double>(), f->part_d.as<double>(),
                                                                                   f->H.as<double>(), f->J.as<double>(), splits, w_lo,
                                                                                   f->lam.as<double>(), f->coe.as<double>(), f->wv_off.as<int32_t>(),
                                                                                   with_cost ? f->cost_h.as<double>() : nullptr, head, end));
  return VX_OK;
}

}  // namespace

namespace vx {
// Returns VX_OK and sets *handled when W has a fused instantiation.  Hess/JacT of windows [w_lo, w_hi) (all when w_hi < 0) are
// fully overwritten.
int hessian_small(vx_ctx *ctx, vx_factor *f, int64_t head, int64_t end, bool *handled, int w_lo, int w_hi, bool with_cost) {
  *handled = true;
  if (w_hi < 0) { w_lo = 0; w_hi = f->n_win; }
  if (w_hi <= w_lo) return VX_OK;
  switch (f->W) {
    case 10: {
      static const int minb = getenv("VX_HS_MINB") ? atoi(getenv("VX_HS_MINB")) : 3;
      if (minb == 4) return launch_small<10, 4, 4>(ctx, f, head, end, w_lo, w_hi, with_cost);
      if (minb == 2) return launch_small<10, 4, 2>(ctx, f, head, end, w_lo, w_hi, with_cost);
      return launch_small<10, 4, 3>(ctx, f, head, end, w_lo, w_hi, with_cost);
    }
    case 20: return launch_small<20, 8, 1>(ctx, f, head, end, w_lo, w_hi, with_cost);
    // other window sizes a caller may configure (win_size / wdsize are ROS parameters, voxelslam.cpp:919, :3033); anything else
    // takes the pair-list formulation of vx_factor.cu
    case 5: return launch_small<5, 4, 3>(ctx, f, head, end, w_lo, w_hi, with_cost);
    case 8: return launch_small<8, 4, 3>(ctx, f, head, end, w_lo, w_hi, with_cost);
    case 12: return launch_small<12, 4, 2>(ctx, f, head, end, w_lo, w_hi, with_cost);
    case 15: return launch_small<15, 8, 1>(ctx, f, head, end, w_lo, w_hi, with_cost);
    case 16: return launch_small<16, 8, 1>(ctx, f, head, end, w_lo, w_hi, with_cost);
    default: *handled = false; return VX_OK;
  }
}
}  // namespace vx
