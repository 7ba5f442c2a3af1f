// vx_factor.cu — LidarFactor on device: storage, reduction plans, and the two BA evaluation passes
//   B2  evaluate_only_residual  (voxel_map.hpp:285-325)  -> k_residual
//   B1  acc_evaluate2           (voxel_map.hpp:150-282)  -> k_voxel_prep, k_diag, k_offdiag, k_fin_*
//
// B1 formulation.  With uk=u0, S1=2/(l0-l1), S2=2/(l0-l2), S3=-2/NN^2 and, per voxel-frame i,
//   g_m,i = Auk_i^T u_m (m=1,2)      w_i = [ v_i x (R_i^T uk) ; n_i uk ]
// the reference's blocks are exactly
//   Hess(i,j) += coe * sum_m S_m g_m,i g_m,j^T                    (i<j, voxel_map.hpp:264-270)
//   Hess(i,i) += coe * ( sum_m S_m g_m,i g_m,i^T + E_i )          (voxel_map.hpp:239-251)
// where E_i holds the closed-form corrections that are not part of the rank-3 form.  The sum over
// voxels is done without atomics: voxel-frames are visited in frame-major order (diagonal tiles) and
// voxel-frame pairs in tile-major order (off-diagonal tiles), each tile's list is cut into fixed
// segments, one block reduces one segment in registers + a fixed-shape tree, and a second pass adds
// the segment partials of a tile in ascending order.  The result is bitwise reproducible.
#include "vx_factor.cuh"
#include "vx_math.cuh"
#include "vx_terms.cuh"
#include <cub/cub.cuh>
#include <algorithm>
#include <cstdlib>

using namespace vx;

// ----------------------------------------------------------------------------------------------
// allocation / upload
// ----------------------------------------------------------------------------------------------
namespace vx {

int factor_alloc(vx_factor *f, int64_t V, int64_t nnz) {
  f->V = V; f->nnz = nnz; f->plan_ready = false; f->prep_lo = f->prep_hi = 0;
  const int F = f->F();
  VX_TRY(f->off.reserve(sizeof(int32_t) * (V + 1)));
  VX_TRY(f->frame.reserve(sizeof(int32_t) * std::max<int64_t>(nnz, 1)));
  VX_TRY(f->vox.reserve(sizeof(int32_t) * std::max<int64_t>(nnz, 1)));
  VX_TRY(f->mom.reserve(sizeof(double) * 9 * std::max<int64_t>(nnz, 1)));
  VX_TRY(f->cnt.reserve(sizeof(int32_t) * std::max<int64_t>(nnz, 1)));
  VX_TRY(f->fix.reserve(sizeof(double) * 10 * std::max<int64_t>(V, 1)));
  VX_TRY(f->coe.reserve(sizeof(double) * std::max<int64_t>(V, 1)));
  VX_TRY(f->lam.reserve(sizeof(double) * 3 * std::max<int64_t>(V, 1)));
  VX_TRY(f->U.reserve(sizeof(double) * 9 * std::max<int64_t>(V, 1)));
  VX_TRY(f->mrg.reserve(sizeof(double) * 10 * std::max<int64_t>(V, 1)));
  VX_TRY(f->win.reserve(sizeof(int32_t) * std::max<int64_t>(V, 1)));
  VX_TRY(f->wv_off.reserve(sizeof(int32_t) * (f->n_win + 1)));
  VX_TRY(f->cS.reserve(sizeof(double) * 7 * std::max<int64_t>(V, 1)));
  VX_TRY(f->res_v.reserve(sizeof(double) * std::max<int64_t>(V, 1)));
  VX_TRY(f->G.reserve(sizeof(double) * 18 * std::max<int64_t>(nnz, 1)));
  VX_TRY(f->poses.reserve(sizeof(double) * 12 * F));
  const size_t jl = (size_t)6 * f->W;
  VX_TRY(f->H.reserve(sizeof(double) * jl * jl * f->n_win));
  VX_TRY(f->J.reserve(sizeof(double) * jl * f->n_win));
  VX_TRY(f->cost_h.reserve(sizeof(double) * f->n_win));
  VX_TRY(f->cost_r.reserve(sizeof(double) * f->n_win));
  return VX_OK;
}

// ---- plan kernels ----------------------------------------------------------------------------
__global__ void k_fill_vox(const int32_t *__restrict__ off, int32_t *__restrict__ vox, int64_t V) {
  int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  for (int32_t e = off[v]; e < off[v + 1]; e++) vox[e] = (int32_t)v;
}
__global__ void k_iota(int32_t *p, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) p[i] = (int32_t)i;
}
// off[f] = first index in sorted keys with key >= f, f in [0, F]
__global__ void k_lower_bound(const int32_t *__restrict__ keys, int64_t n, int32_t *__restrict__ out, int F) {
  int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f > F) return;
  int64_t lo = 0, hi = n;
  while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (keys[mid] < f) lo = mid + 1; else hi = mid; }
  out[f] = (int32_t)lo;
}
__global__ void k_pair_count(const int32_t *__restrict__ off, int64_t *__restrict__ cnt, int64_t V) {
  int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  int64_t k = off[v + 1] - off[v];
  cnt[v] = k * (k - 1) / 2;
}
__global__ void k_pair_write(const int32_t *__restrict__ off, const int32_t *__restrict__ frame, const int64_t *__restrict__ poff,
                             int64_t *__restrict__ keys, int64_t *__restrict__ vals, int64_t V, int64_t F) {
  int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  int64_t o = poff[v];
  int32_t b = off[v], e = off[v + 1];
  for (int32_t a = b; a < e; a++)
    for (int32_t c = a + 1; c < e; c++) {
      keys[o] = (int64_t)frame[a] * F + frame[c];
      vals[o] = ((int64_t)a << 32) | (uint32_t)c;
      o++;
    }
}
__global__ void k_seg_count(const int32_t *__restrict__ toff, int32_t *__restrict__ nseg, int64_t nt, int seg) {
  int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= nt) return;
  int32_t len = toff[t + 1] - toff[t];
  nseg[t] = (len + seg - 1) / seg;
}
__global__ void k_seg_fill(const int32_t *__restrict__ toff, const int32_t *__restrict__ soff, int32_t *__restrict__ seg_tile,
                           int32_t *__restrict__ seg_beg, int32_t *__restrict__ seg_end, int64_t nt, int seg) {
  int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= nt) return;
  int32_t b = toff[t], e = toff[t + 1];
  int32_t s = soff[t];
  for (int32_t x = b; x < e; x += seg, s++) { seg_tile[s] = (int32_t)t; seg_beg[s] = x; seg_end[s] = min(x + seg, e); }
}
__global__ void k_counts_to_i32(const int64_t *__restrict__ in, int32_t *__restrict__ out, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) out[i] = (int32_t)in[i];
}

static int read_i32(vx_ctx *ctx, const int32_t *d, int32_t *h) {
  VX_CUDA(cudaMemcpyAsync(h, d, sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  return VX_OK;
}
static int read_i64(vx_ctx *ctx, const int64_t *d, int64_t *h) {
  VX_CUDA(cudaMemcpyAsync(h, d, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  return VX_OK;
}

// exclusive scan of n int32 values into out[0..n] (out[n] = total)
static int scan_i32(vx_ctx *ctx, const int32_t *in, int32_t *out, int64_t n) {
  size_t tb = 0;
  VX_CUDA(cudaMemsetAsync(out, 0, sizeof(int32_t), ctx->stream));
  VX_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tb, in, out + 1, (int)n, ctx->stream));
  VX_TRY(ctx->scratch.reserve(tb));
  VX_CUDA(cub::DeviceScan::InclusiveSum(ctx->scratch.p, tb, in, out + 1, (int)n, ctx->stream));
  ctx->launches += 1;
  return VX_OK;
}

// vox[e] = owning voxel of voxel-frame e (needed by every evaluation path); the pair-list plans are built
// lazily by the first generic-path Hessian (the fused small-window path does not need them).
int factor_finish_upload(vx_factor *f) {
  vx_ctx *ctx = f->ctx;
  VX_TRY(bind(ctx));
  f->plan_ready = false; f->prep_lo = f->prep_hi = 0;
  f->n_pairs = f->n_tiles = f->n_dseg = f->n_oseg = 0;
  if (f->V == 0 || f->nnz == 0) { f->plan_ready = true; return VX_OK; }
  k_fill_vox<<<blocks_for(f->V, 256), 256, 0, ctx->stream>>>(f->off.as<int32_t>(), f->vox.as<int32_t>(), f->V);
  VX_LAUNCH_CHECK(ctx);
  return VX_OK;
}

int factor_build_plan(vx_factor *f) {
  vx_ctx *ctx = f->ctx;
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  const int64_t V = f->V, nnz = f->nnz;
  const int F = f->F();
  f->plan_ready = false; f->prep_lo = f->prep_hi = 0;
  f->n_pairs = f->n_tiles = f->n_dseg = f->n_oseg = 0;
  if (V == 0 || nnz == 0) { f->plan_ready = true; return VX_OK; }
  VX_ARG(nnz < (int64_t)1 << 31);

  // segment length: aim at >= 2 blocks per SM, bounded so a thread reduces at most 32 entries
  {
    int64_t want = nnz / (2 * (int64_t)ctx->sm_count);
    int seg = 128;
    while (seg < want && seg < 4096) seg <<= 1;
    f->seg_len = seg;
  }
  const int seg = f->seg_len;

  k_fill_vox<<<blocks_for(V, 256), 256, 0, st>>>(f->off.as<int32_t>(), f->vox.as<int32_t>(), V);
  VX_LAUNCH_CHECK(ctx);

  // ---- frame-major order: stable sort of voxel-frame ids by frame
  VX_TRY(f->fm_order.reserve(sizeof(int32_t) * nnz));
  VX_TRY(f->fm_off.reserve(sizeof(int32_t) * (F + 1)));
  {
    DevBuf &tmp = ctx->scratch2;
    VX_TRY(tmp.reserve(sizeof(int32_t) * nnz * 2));
    int32_t *iota = tmp.as<int32_t>();
    int32_t *keys_out = iota + nnz;
    k_iota<<<blocks_for(nnz, 256), 256, 0, st>>>(iota, nnz);
    VX_LAUNCH_CHECK(ctx);
    int bits = 1; while ((1 << bits) < F) bits++;
    size_t tb = 0;
    VX_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, f->frame.as<int32_t>(), keys_out, iota, f->fm_order.as<int32_t>(), (int)nnz, 0, bits, st));
    VX_TRY(ctx->scratch.reserve(tb));
    VX_CUDA(cub::DeviceRadixSort::SortPairs(ctx->scratch.p, tb, f->frame.as<int32_t>(), keys_out, iota, f->fm_order.as<int32_t>(), (int)nnz, 0, bits, st));
    ctx->launches += 3;
    k_lower_bound<<<blocks_for(F + 1, 128), 128, 0, st>>>(keys_out, nnz, f->fm_off.as<int32_t>(), F);
    VX_LAUNCH_CHECK(ctx);
  }
  // diagonal segments
  {
    DevBuf &tmp = ctx->scratch2;
    VX_TRY(tmp.reserve(sizeof(int32_t) * (F + 1)));
    int32_t *nseg = tmp.as<int32_t>();
    VX_TRY(f->dseg_off.reserve(sizeof(int32_t) * (F + 1)));
    k_seg_count<<<blocks_for(F, 128), 128, 0, st>>>(f->fm_off.as<int32_t>(), nseg, F, seg);
    VX_LAUNCH_CHECK(ctx);
    VX_TRY(scan_i32(ctx, nseg, f->dseg_off.as<int32_t>(), F));
    int32_t total = 0;
    VX_TRY(read_i32(ctx, f->dseg_off.as<int32_t>() + F, &total));
    f->n_dseg = total;
    VX_TRY(f->dseg_frame.reserve(sizeof(int32_t) * std::max(total, 1)));
    VX_TRY(f->dseg_beg.reserve(sizeof(int32_t) * std::max(total, 1)));
    VX_TRY(f->dseg_end.reserve(sizeof(int32_t) * std::max(total, 1)));
    k_seg_fill<<<blocks_for(F, 128), 128, 0, st>>>(f->fm_off.as<int32_t>(), f->dseg_off.as<int32_t>(), f->dseg_frame.as<int32_t>(),
                                                 f->dseg_beg.as<int32_t>(), f->dseg_end.as<int32_t>(), F, seg);
    VX_LAUNCH_CHECK(ctx);
    VX_TRY(f->dpart.reserve(sizeof(double) * 42 * std::max(total, 1)));
  }

  // ---- pairs, sorted by tile
  {
    DevBuf pc, po;
    VX_TRY(pc.reserve(sizeof(int64_t) * (V + 1)));
    VX_TRY(po.reserve(sizeof(int64_t) * (V + 1)));
    k_pair_count<<<blocks_for(V, 256), 256, 0, st>>>(f->off.as<int32_t>(), pc.as<int64_t>(), V);
    VX_LAUNCH_CHECK(ctx);
    VX_CUDA(cudaMemsetAsync(po.p, 0, sizeof(int64_t), st));
    size_t tb = 0;
    VX_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tb, pc.as<int64_t>(), po.as<int64_t>() + 1, (int)V, st));
    VX_TRY(ctx->scratch.reserve(tb));
    VX_CUDA(cub::DeviceScan::InclusiveSum(ctx->scratch.p, tb, pc.as<int64_t>(), po.as<int64_t>() + 1, (int)V, st));
    ctx->launches += 1;
    int64_t np = 0;
    VX_TRY(read_i64(ctx, po.as<int64_t>() + V, &np));
    f->n_pairs = np;
    VX_ARG(np < (int64_t)1 << 31);
    if (np > 0) {
      DevBuf k0, v0, k1;
      VX_TRY(k0.reserve(sizeof(int64_t) * np));
      VX_TRY(v0.reserve(sizeof(int64_t) * np));
      VX_TRY(k1.reserve(sizeof(int64_t) * np));
      VX_TRY(f->pair_ab.reserve(sizeof(int64_t) * np));
      k_pair_write<<<blocks_for(V, 128), 128, 0, st>>>(f->off.as<int32_t>(), f->frame.as<int32_t>(), po.as<int64_t>(), k0.as<int64_t>(), v0.as<int64_t>(), V, F);
      VX_LAUNCH_CHECK(ctx);
      int bits = 1; while (((int64_t)1 << bits) < (int64_t)F * F) bits++;
      VX_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, k0.as<int64_t>(), k1.as<int64_t>(), v0.as<int64_t>(), f->pair_ab.as<int64_t>(), (int)np, 0, bits, st));
      VX_TRY(ctx->scratch.reserve(tb));
      VX_CUDA(cub::DeviceRadixSort::SortPairs(ctx->scratch.p, tb, k0.as<int64_t>(), k1.as<int64_t>(), v0.as<int64_t>(), f->pair_ab.as<int64_t>(), (int)np, 0, bits, st));
      ctx->launches += 3;
      // tiles = runs of equal key
      int64_t max_tiles = std::min<int64_t>(np, (int64_t)F * (f->W - 1) / 2 + 1);
      DevBuf runs, nruns;
      VX_TRY(f->tile_key.reserve(sizeof(int64_t) * max_tiles));
      VX_TRY(runs.reserve(sizeof(int32_t) * (max_tiles + 1)));
      VX_TRY(nruns.reserve(sizeof(int32_t)));
      VX_CUDA(cub::DeviceRunLengthEncode::Encode(nullptr, tb, k1.as<int64_t>(), f->tile_key.as<int64_t>(), runs.as<int32_t>(), nruns.as<int32_t>(), (int)np, st));
      VX_TRY(ctx->scratch.reserve(tb));
      VX_CUDA(cub::DeviceRunLengthEncode::Encode(ctx->scratch.p, tb, k1.as<int64_t>(), f->tile_key.as<int64_t>(), runs.as<int32_t>(), nruns.as<int32_t>(), (int)np, st));
      ctx->launches += 2;
      int32_t nt = 0;
      VX_TRY(read_i32(ctx, nruns.as<int32_t>(), &nt));
      f->n_tiles = nt;
      VX_TRY(f->tile_off.reserve(sizeof(int32_t) * (nt + 1)));
      VX_TRY(scan_i32(ctx, runs.as<int32_t>(), f->tile_off.as<int32_t>(), nt));
      // off-diagonal segments
      DevBuf nseg;
      VX_TRY(nseg.reserve(sizeof(int32_t) * (nt + 1)));
      VX_TRY(f->oseg_off.reserve(sizeof(int32_t) * (nt + 1)));
      k_seg_count<<<blocks_for(nt, 128), 128, 0, st>>>(f->tile_off.as<int32_t>(), nseg.as<int32_t>(), nt, seg);
      VX_LAUNCH_CHECK(ctx);
      VX_TRY(scan_i32(ctx, nseg.as<int32_t>(), f->oseg_off.as<int32_t>(), nt));
      int32_t total = 0;
      VX_TRY(read_i32(ctx, f->oseg_off.as<int32_t>() + nt, &total));
      f->n_oseg = total;
      VX_TRY(f->oseg_tile.reserve(sizeof(int32_t) * std::max(total, 1)));
      VX_TRY(f->oseg_beg.reserve(sizeof(int32_t) * std::max(total, 1)));
      VX_TRY(f->oseg_end.reserve(sizeof(int32_t) * std::max(total, 1)));
      k_seg_fill<<<blocks_for(nt, 128), 128, 0, st>>>(f->tile_off.as<int32_t>(), f->oseg_off.as<int32_t>(), f->oseg_tile.as<int32_t>(),
                                                    f->oseg_beg.as<int32_t>(), f->oseg_end.as<int32_t>(), nt, seg);
      VX_LAUNCH_CHECK(ctx);
      VX_TRY(f->opart.reserve(sizeof(double) * 36 * std::max(total, 1)));
      VX_CUDA(cudaStreamSynchronize(st));
      k0.release(); v0.release(); k1.release(); runs.release(); nruns.release(); nseg.release();
    }
    VX_CUDA(cudaStreamSynchronize(st));
    pc.release(); po.release();
  }
  f->plan_ready = true;
  return VX_OK;
}

}  // namespace vx

// ----------------------------------------------------------------------------------------------
// evaluation kernels
// ----------------------------------------------------------------------------------------------
namespace {

constexpr int RES_THREADS = 128;

// B2: one thread per voxel.  sig = fix + sum_i transform(cluster_i, pose_i) (tools.hpp:357-363),
// eigen-solve of sig.P/N - vbar vbar^T (voxel_map.hpp:312), cache write-back (voxel_map.hpp:317-319).
template <int MINB>
__global__ void __launch_bounds__(RES_THREADS, MINB)
k_residual(const int32_t *__restrict__ off, const int32_t *__restrict__ frame, const double *__restrict__ mom,
           const int32_t *__restrict__ cnt, int64_t nnz, const double *__restrict__ fix, const double *__restrict__ coe,
           const double *__restrict__ poses, double *__restrict__ lam, double *__restrict__ U, double *__restrict__ mrg,
           double *__restrict__ res_v, double *__restrict__ prep, int64_t V, int64_t head, int64_t end) {
  int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (v >= V || v < head || v >= end) return;
  s33 P;
  P.xx = fix[0 * V + v]; P.xy = fix[1 * V + v]; P.xz = fix[2 * V + v];
  P.yy = fix[3 * V + v]; P.yz = fix[4 * V + v]; P.zz = fix[5 * V + v];
  d3 sv = mk3(fix[6 * V + v], fix[7 * V + v], fix[8 * V + v]);
  double N = fix[9 * V + v];
  const int32_t b = off[v], e = off[v + 1];
  for (int32_t x = b; x < e; x++) {
    const double *ps = poses + 12 * (int64_t)frame[x];
    rot R = load_rot(ps);
    d3 p = mk3(ps[9], ps[10], ps[11]);
    s33 Pi;
    Pi.xx = mom[0 * nnz + x]; Pi.xy = mom[1 * nnz + x]; Pi.xz = mom[2 * nnz + x];
    Pi.yy = mom[3 * nnz + x]; Pi.yz = mom[4 * nnz + x]; Pi.zz = mom[5 * nnz + x];
    d3 vi = mk3(mom[6 * nnz + x], mom[7 * nnz + x], mom[8 * nnz + x]);
    double n = (double)cnt[x];
    d3 Rv = rmul(R, vi);
    s33 RPR = rprt(R, Pi);
    // P' = R P R^T + Rv p^T + p Rv^T + n p p^T
    P.xx += RPR.xx + 2.0 * Rv.x * p.x + n * p.x * p.x;
    P.xy += RPR.xy + Rv.x * p.y + p.x * Rv.y + n * p.x * p.y;
    P.xz += RPR.xz + Rv.x * p.z + p.x * Rv.z + n * p.x * p.z;
    P.yy += RPR.yy + 2.0 * Rv.y * p.y + n * p.y * p.y;
    P.yz += RPR.yz + Rv.y * p.z + p.y * Rv.z + n * p.y * p.z;
    P.zz += RPR.zz + 2.0 * Rv.z * p.z + n * p.z * p.z;
    sv = sv + Rv + n * p;
    N += n;
  }
  double invN = 1.0 / N;
  d3 vb = mk3(sv.x / N, sv.y / N, sv.z / N);
  s33 C;
  C.xx = P.xx / N - vb.x * vb.x; C.xy = P.xy / N - vb.x * vb.y; C.xz = P.xz / N - vb.x * vb.z;
  C.yy = P.yy / N - vb.y * vb.y; C.yz = P.yz / N - vb.y * vb.z; C.zz = P.zz / N - vb.z * vb.z;
  double l[3], u[9];
  eig3_jacobi(C, l, u);
  lam[0 * V + v] = l[0]; lam[1 * V + v] = l[1]; lam[2 * V + v] = l[2];
#pragma unroll
  for (int k = 0; k < 9; k++) U[k * V + v] = u[k];
  mrg[0 * V + v] = P.xx; mrg[1 * V + v] = P.xy; mrg[2 * V + v] = P.xz;
  mrg[3 * V + v] = P.yy; mrg[4 * V + v] = P.yz; mrg[5 * V + v] = P.zz;
  mrg[6 * V + v] = sv.x; mrg[7 * V + v] = sv.y; mrg[8 * V + v] = sv.z;
  mrg[9 * V + v] = N;
  const double cf = coe[v];
  res_v[v] = cf * l[0];
  // the divisions the next acc_evaluate2 needs per voxel (k_cost_prep below states them)
  prep[v] = cf * (2.0 / (l[0] - l[1]));
  prep[V + v] = cf * (2.0 / (l[0] - l[2]));
  prep[2 * V + v] = cf * (-2.0 / N / N);
  prep[3 * V + v] = vb.x; prep[4 * V + v] = vb.y; prep[5 * V + v] = vb.z;
  prep[6 * V + v] = invN;
}

// B2, staged form (opt-in, see launch_residual for the measurement).  The thread-per-voxel loop above walks each voxel's records with a stride of one voxel's record count between
// neighbouring lanes and one record in flight per lane: latency-bound (long scoreboard 55 %, r01n capture).  Here a CTA owns RES_THREADS
// consecutive voxels = ONE contiguous record range; its threads stream that range lane-per-record (every plane load a full coalesced
// line, RS_PER records in flight per thread), transform each record by its pose and park the 10 transformed moments in shared memory;
// then thread v adds the parked records of voxel v in record order (the order of the loop above; the compiler contracts the two
// forms differently, so the caches agree to rounding, not bit for bit) and does the eigen-solve and the cache write-back.
constexpr int RS_PER = 2;                          // records per thread per chunk
constexpr int RS_CHUNK = RES_THREADS * RS_PER;     // records parked at a time: 10 x 256 doubles = 20 KB
template <int MINB>
__global__ void __launch_bounds__(RES_THREADS, MINB)
k_residual_staged(const int32_t *__restrict__ off, const int32_t *__restrict__ frame, const double *__restrict__ mom,
                  const int32_t *__restrict__ cnt, int64_t nnz, const double *__restrict__ fix, const double *__restrict__ coe,
                  const double *__restrict__ poses, double *__restrict__ lam, double *__restrict__ U, double *__restrict__ mrg,
                  double *__restrict__ res_v, double *__restrict__ prep, int64_t V, int64_t head, int64_t end) {
  __shared__ double park[10][RS_CHUNK];
  const int64_t v0 = blockIdx.x * (int64_t)RES_THREADS;
  const int64_t vlo = max(v0, max(head, (int64_t)0)), vhi = min(v0 + RES_THREADS, min(end, V));
  if (vlo >= vhi) return;                                         // uniform per CTA
  const int64_t v = v0 + threadIdx.x;
  const bool mine = v >= vlo && v < vhi;
  const int32_t xb = off[vlo], xe = off[vhi];
  const int32_t b = mine ? off[v] : 0, e = mine ? off[v + 1] : 0;
  s33 P; d3 sv; double N = 0.0;
  P.xx = P.xy = P.xz = P.yy = P.yz = P.zz = 0.0; sv = mk3(0.0, 0.0, 0.0);
  if (mine) {
    P.xx = fix[0 * V + v]; P.xy = fix[1 * V + v]; P.xz = fix[2 * V + v];
    P.yy = fix[3 * V + v]; P.yz = fix[4 * V + v]; P.zz = fix[5 * V + v];
    sv = mk3(fix[6 * V + v], fix[7 * V + v], fix[8 * V + v]);
    N = fix[9 * V + v];
  }
  for (int32_t c0 = xb; c0 < xe; c0 += RS_CHUNK) {
    // ---- stream: every load of the chunk is issued before the first transform
    double m[RS_PER][9]; int32_t fr[RS_PER], cn[RS_PER];
#pragma unroll
    for (int q = 0; q < RS_PER; q++) {
      const int32_t x = c0 + q * RES_THREADS + threadIdx.x;
      const bool in = x < xe;
      fr[q] = in ? frame[x] : 0; cn[q] = in ? cnt[x] : 0;
#pragma unroll
      for (int k = 0; k < 9; k++) m[q][k] = in ? mom[(int64_t)k * nnz + x] : 0.0;
    }
#pragma unroll
    for (int q = 0; q < RS_PER; q++) {
      const int slot = q * RES_THREADS + threadIdx.x;
      const double *ps = poses + 12 * (int64_t)fr[q];
      const rot R = load_rot(ps);
      const d3 p = mk3(ps[9], ps[10], ps[11]);
      s33 Pi;
      Pi.xx = m[q][0]; Pi.xy = m[q][1]; Pi.xz = m[q][2]; Pi.yy = m[q][3]; Pi.yz = m[q][4]; Pi.zz = m[q][5];
      const d3 vi = mk3(m[q][6], m[q][7], m[q][8]);
      const double n = (double)cn[q];
      const d3 Rv = rmul(R, vi);
      const s33 RPR = rprt(R, Pi);
      // the record's whole contribution, summed exactly as the loop of k_residual sums it into P (left to right)
      park[0][slot] = RPR.xx + 2.0 * Rv.x * p.x + n * p.x * p.x;
      park[1][slot] = RPR.xy + Rv.x * p.y + p.x * Rv.y + n * p.x * p.y;
      park[2][slot] = RPR.xz + Rv.x * p.z + p.x * Rv.z + n * p.x * p.z;
      park[3][slot] = RPR.yy + 2.0 * Rv.y * p.y + n * p.y * p.y;
      park[4][slot] = RPR.yz + Rv.y * p.z + p.y * Rv.z + n * p.y * p.z;
      park[5][slot] = RPR.zz + 2.0 * Rv.z * p.z + n * p.z * p.z;
      park[6][slot] = Rv.x; park[7][slot] = Rv.y; park[8][slot] = Rv.z;
      park[9][slot] = n;
    }
    __syncthreads();
    // ---- gather: voxel v takes its records of this chunk in record order
    const int32_t lo = max(b, c0), hi = min(e, c0 + RS_CHUNK);
    for (int32_t x = lo; x < hi; x++) {
      const int s = x - c0;
      const double n = park[9][s];
      const d3 p = mk3(poses[12 * (int64_t)frame[x] + 9], poses[12 * (int64_t)frame[x] + 10], poses[12 * (int64_t)frame[x] + 11]);
      P.xx += park[0][s]; P.xy += park[1][s]; P.xz += park[2][s]; P.yy += park[3][s]; P.yz += park[4][s]; P.zz += park[5][s];
      sv = sv + mk3(park[6][s], park[7][s], park[8][s]) + n * p;
      N += n;
    }
    __syncthreads();
  }
  if (!mine) return;
  double invN = 1.0 / N;
  d3 vb = mk3(sv.x / N, sv.y / N, sv.z / N);
  s33 C;
  C.xx = P.xx / N - vb.x * vb.x; C.xy = P.xy / N - vb.x * vb.y; C.xz = P.xz / N - vb.x * vb.z;
  C.yy = P.yy / N - vb.y * vb.y; C.yz = P.yz / N - vb.y * vb.z; C.zz = P.zz / N - vb.z * vb.z;
  double l[3], u[9];
  eig3_jacobi(C, l, u);
  lam[0 * V + v] = l[0]; lam[1 * V + v] = l[1]; lam[2 * V + v] = l[2];
#pragma unroll
  for (int k = 0; k < 9; k++) U[k * V + v] = u[k];
  mrg[0 * V + v] = P.xx; mrg[1 * V + v] = P.xy; mrg[2 * V + v] = P.xz;
  mrg[3 * V + v] = P.yy; mrg[4 * V + v] = P.yz; mrg[5 * V + v] = P.zz;
  mrg[6 * V + v] = sv.x; mrg[7 * V + v] = sv.y; mrg[8 * V + v] = sv.z;
  mrg[9 * V + v] = N;
  const double cf = coe[v];
  res_v[v] = cf * l[0];
  prep[v] = cf * (2.0 / (l[0] - l[1]));
  prep[V + v] = cf * (2.0 / (l[0] - l[2]));
  prep[2 * V + v] = cf * (-2.0 / N / N);
  prep[3 * V + v] = vb.x; prep[4 * V + v] = vb.y; prep[5 * V + v] = vb.z;
  prep[6 * V + v] = invN;
}

// cost[w] = sum over voxels of window w (restricted to [head,end)) of coe*lambda_0, fixed-shape tree.
template <bool FROM_LAM>
__global__ void __launch_bounds__(256)
k_cost(const double *__restrict__ src, const double *__restrict__ coe, const int32_t *__restrict__ wv_off,
       double *__restrict__ cost, int64_t head, int64_t end) {
  __shared__ double sh[256];
  const int w = blockIdx.x;
  int64_t b = wv_off[w], e = wv_off[w + 1];
  if (b < head) b = head;
  if (e > end) e = end;
  double acc = 0.0;
  for (int64_t v = b + threadIdx.x; v < e; v += 256) acc += FROM_LAM ? coe[v] * src[v] : src[v];
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) cost[w] = sh[0];
}

// The fused small-window path wants every per-voxel division done once per voxel, not once per voxel-frame: this is k_cost
// (window cost from the cached eigenvalues) that also writes, per voxel of [head, end),
//   prep[0..2] = coe * {2/(l0-l1), 2/(l0-l2), -2/NN^2}   (voxel_map.hpp:198-201, :265-268)
//   prep[3..5] = vBar = v / NN,  prep[6] = 1 / NN          (voxel_map.hpp:193-195)
__global__ void __launch_bounds__(1024)
k_cost_prep(const double *__restrict__ lam, const double *__restrict__ coe, const double *__restrict__ mrg, const int32_t *__restrict__ wv_off,
            double *__restrict__ cost, double *__restrict__ prep, int64_t V, int64_t head, int64_t end) {
  __shared__ double sh[1024];
  const int w = blockIdx.x;
  int64_t b = wv_off[w], e = wv_off[w + 1];
  if (b < head) b = head;
  if (e > end) e = end;
  double acc = 0.0;
  for (int64_t v = b + threadIdx.x; v < e; v += 1024) {
    const double l0 = lam[v], l1 = lam[V + v], l2 = lam[2 * V + v], c = coe[v], NN = mrg[9 * V + v];
    acc += c * l0;
    prep[v] = c * (2.0 / (l0 - l1));
    prep[V + v] = c * (2.0 / (l0 - l2));
    prep[2 * V + v] = c * (-2.0 / NN / NN);
    prep[3 * V + v] = mrg[6 * V + v] / NN;
    prep[4 * V + v] = mrg[7 * V + v] / NN;
    prep[5 * V + v] = mrg[8 * V + v] / NN;
    prep[6 * V + v] = 1.0 / NN;
  }
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 512; s > 0; s >>= 1) {
    if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) cost[w] = sh[0];
}

// per voxel: cS = coe * {2/(l0-l1), 2/(l0-l2), -2/NN^2}   (voxel_map.hpp:198-201, :265-268)
__global__ void k_voxel_prep(const double *__restrict__ lam, const double *__restrict__ mrg, const double *__restrict__ coe,
                             double *__restrict__ cS, int64_t V) {
  int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  double l0 = lam[v], l1 = lam[V + v], l2 = lam[2 * V + v];
  double NN = mrg[9 * V + v];
  double c = coe[v];
  cS[v] = c * (2.0 / (l0 - l1));
  cS[V + v] = c * (2.0 / (l0 - l2));
  cS[2 * V + v] = c * (-2.0 / NN / NN);
}

constexpr int DIAG_THREADS = 128;
constexpr int DIAG_WARPS = DIAG_THREADS / 32;

// fixed-shape block reduction of NV doubles per thread; result for value k lands in out[k]
template <int NV, int THREADS>
__device__ __forceinline__ void block_reduce_store(double (&acc)[NV], double *sh /* [THREADS/32][NV] */, double *out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < NV; k++) {
    double x = acc[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0) sh[warp * NV + k] = x;
  }
  __syncthreads();
  if (threadIdx.x < NV) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < THREADS / 32; w++) s += sh[w * NV + threadIdx.x];
    out[threadIdx.x] = s;
  }
}

// B1 diagonal tiles + gradient: one block per segment of the frame-major voxel-frame list.
__global__ void __launch_bounds__(DIAG_THREADS)
k_diag(const int32_t *__restrict__ seg_frame, const int32_t *__restrict__ seg_beg, const int32_t *__restrict__ seg_end,
       const int32_t *__restrict__ fm_order, const int32_t *__restrict__ vox, const double *__restrict__ mom,
       const int32_t *__restrict__ cnt, int64_t nnz, const double *__restrict__ lam, const double *__restrict__ U,
       const double *__restrict__ mrg, const double *__restrict__ coe, const double *__restrict__ cS,
       const double *__restrict__ poses, double *__restrict__ G, double *__restrict__ dpart, int64_t V,
       int64_t head, int64_t end) {
  __shared__ double sh[DIAG_WARPS * 42];
  const int s = blockIdx.x;
  const int fr = seg_frame[s];
  const double *ps = poses + 12 * (int64_t)fr;
  const rot R = load_rot(ps);
  const d3 p = mk3(ps[9], ps[10], ps[11]);
  double acc[42];
#pragma unroll
  for (int k = 0; k < 42; k++) acc[k] = 0.0;
  for (int32_t x = seg_beg[s] + threadIdx.x; x < seg_end[s]; x += DIAG_THREADS) {
    const int32_t e = fm_order[x];
    const int64_t v = vox[e];
    if (v < head || v >= end) continue;
    s33 P;
    P.xx = mom[0 * nnz + e]; P.xy = mom[1 * nnz + e]; P.xz = mom[2 * nnz + e];
    P.yy = mom[3 * nnz + e]; P.yz = mom[4 * nnz + e]; P.zz = mom[5 * nnz + e];
    d3 vi = mk3(mom[6 * nnz + e], mom[7 * nnz + e], mom[8 * nnz + e]);
    double n = (double)cnt[e];
    d3 uk = mk3(U[0 * V + v], U[1 * V + v], U[2 * V + v]);
    d3 u1 = mk3(U[3 * V + v], U[4 * V + v], U[5 * V + v]);
    d3 u2 = mk3(U[6 * V + v], U[7 * V + v], U[8 * V + v]);
    double NN = mrg[9 * V + v];
    d3 vbar = mk3(mrg[6 * V + v] / NN, mrg[7 * V + v] / NN, mrg[8 * V + v] / NN);
    VfTerms t;
    vf_terms(R, p, P, vi, n, uk, u1, u2, vbar, 1.0 / NN, t);
    double *g = G + 18 * (int64_t)e;
#pragma unroll
    for (int k = 0; k < 18; k += 2) *reinterpret_cast<double2 *>(g + k) = make_double2(t.g[k], t.g[k + 1]);
    const double c = coe[v];
    const double s1 = cS[v], s2 = cS[V + v], s3 = cS[2 * V + v];
#pragma unroll
    for (int cc = 0; cc < 6; cc++) {
      const double x1 = s1 * t.g[cc], x2 = s2 * t.g[6 + cc], x3 = s3 * t.g[12 + cc];
#pragma unroll
      for (int r = 0; r < 6; r++)
        acc[cc * 6 + r] += fma(t.g[r], x1, fma(t.g[6 + r], x2, fma(t.g[12 + r], x3, c * t.E[cc * 6 + r])));
    }
#pragma unroll
    for (int r = 0; r < 6; r++) acc[36 + r] = fma(c, t.jj[r], acc[36 + r]);
  }
  block_reduce_store<42, DIAG_THREADS>(acc, sh, dpart + 42 * (int64_t)s);
}

// B1 off-diagonal tiles: one block per segment of the tile-major pair list.
__global__ void __launch_bounds__(DIAG_THREADS)
k_offdiag(const int32_t *__restrict__ seg_beg, const int32_t *__restrict__ seg_end, const int64_t *__restrict__ pair_ab,
          const int32_t *__restrict__ vox, const double *__restrict__ cS, const double *__restrict__ G,
          double *__restrict__ opart, int64_t V, int64_t head, int64_t end) {
  __shared__ double sh[DIAG_WARPS * 36];
  const int s = blockIdx.x;
  double acc[36];
#pragma unroll
  for (int k = 0; k < 36; k++) acc[k] = 0.0;
  for (int32_t x = seg_beg[s] + threadIdx.x; x < seg_end[s]; x += DIAG_THREADS) {
    const int64_t ab = pair_ab[x];
    const int32_t a = (int32_t)(ab >> 32), b = (int32_t)(ab & 0xffffffff);
    const int64_t v = vox[a];
    if (v < head || v >= end) continue;
    double ga[18], gb[18];
    const double2 *pa = reinterpret_cast<const double2 *>(G + 18 * (int64_t)a);
    const double2 *pb = reinterpret_cast<const double2 *>(G + 18 * (int64_t)b);
#pragma unroll
    for (int k = 0; k < 9; k++) { double2 t = pa[k]; ga[2 * k] = t.x; ga[2 * k + 1] = t.y; }
#pragma unroll
    for (int k = 0; k < 9; k++) { double2 t = pb[k]; gb[2 * k] = t.x; gb[2 * k + 1] = t.y; }
    const double s1 = cS[v], s2 = cS[V + v], s3 = cS[2 * V + v];
#pragma unroll
    for (int cc = 0; cc < 6; cc++) {
      const double x1 = s1 * gb[cc], x2 = s2 * gb[6 + cc], x3 = s3 * gb[12 + cc];
#pragma unroll
      for (int r = 0; r < 6; r++)
        acc[cc * 6 + r] += fma(ga[r], x1, fma(ga[6 + r], x2, ga[12 + r] * x3));
    }
  }
  block_reduce_store<36, DIAG_THREADS>(acc, sh, opart + 36 * (int64_t)s);
}

// The same for SPARSE tile lists (top-level GBA: W ~ 399, ~4 frames per voxel -> 1-2 pairs per tile, tens of thousands of
// tiles): one WARP per segment, 8 segments per CTA, no shared memory or barrier; the shuffle tree is only as deep as the
// segment is long (lanes past its end hold exact zeros, so the result equals the full tree's).  A CTA per 1-pair segment spent
// its time being scheduled: 0.56 ms per evaluation at 82 k voxel-frames.
__global__ void __launch_bounds__(256)
k_offdiag_warp(const int32_t *__restrict__ seg_beg, const int32_t *__restrict__ seg_end, const int64_t *__restrict__ pair_ab,
               const int32_t *__restrict__ vox, const double *__restrict__ cS, const double *__restrict__ G,
               double *__restrict__ opart, int64_t V, int64_t head, int64_t end, int64_t n_seg) {
  const int64_t s = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (s >= n_seg) return;
  const int32_t beg = seg_beg[s], fin = seg_end[s];
  double acc[36];
#pragma unroll
  for (int k = 0; k < 36; k++) acc[k] = 0.0;
  for (int32_t x = beg + lane; x < fin; x += 32) {
    const int64_t ab = pair_ab[x];
    const int32_t a = (int32_t)(ab >> 32), b = (int32_t)(ab & 0xffffffff);
    const int64_t v = vox[a];
    if (v < head || v >= end) continue;
    double ga[18], gb[18];
    const double2 *pa = reinterpret_cast<const double2 *>(G + 18 * (int64_t)a);
    const double2 *pb = reinterpret_cast<const double2 *>(G + 18 * (int64_t)b);
#pragma unroll
    for (int k = 0; k < 9; k++) { double2 t = pa[k]; ga[2 * k] = t.x; ga[2 * k + 1] = t.y; }
#pragma unroll
    for (int k = 0; k < 9; k++) { double2 t = pb[k]; gb[2 * k] = t.x; gb[2 * k + 1] = t.y; }
    const double s1 = cS[v], s2 = cS[V + v], s3 = cS[2 * V + v];
#pragma unroll
    for (int cc = 0; cc < 6; cc++) {
      const double x1 = s1 * gb[cc], x2 = s2 * gb[6 + cc], x3 = s3 * gb[12 + cc];
#pragma unroll
      for (int r = 0; r < 6; r++)
        acc[cc * 6 + r] += fma(ga[r], x1, fma(ga[6 + r], x2, ga[12 + r] * x3));
    }
  }
  const int len = fin - beg;
  for (int o = 16; o > 0; o >>= 1) {
    if (o < len) {                        // warp-uniform
#pragma unroll
      for (int k = 0; k < 36; k++) acc[k] += __shfl_down_sync(0xffffffffu, acc[k], o);
    }
  }
  if (lane == 0) {
    double2 *out = reinterpret_cast<double2 *>(opart + 36 * s);
#pragma unroll
    for (int k = 0; k < 18; k++) out[k] = make_double2(acc[2 * k], acc[2 * k + 1]);
  }
}

// second pass: add segment partials of each frame in ascending order; write H(i,i) and JacT(i)
__global__ void k_fin_diag(const int32_t *__restrict__ dseg_off, const double *__restrict__ dpart, double *__restrict__ H,
                           double *__restrict__ J, int W) {
  const int fr = blockIdx.x;
  const int t = threadIdx.x;
  if (t >= 42) return;
  double s = 0.0;
  for (int32_t x = dseg_off[fr]; x < dseg_off[fr + 1]; x++) s += dpart[42 * (int64_t)x + t];
  const int w = fr / W, i = fr % W;
  const int64_t jl = 6 * (int64_t)W;
  if (t < 36) {
    const int r = t % 6, c = t / 6;
    H[w * jl * jl + (6 * i + c) * jl + (6 * i + r)] = s;
  } else {
    J[w * jl + 6 * i + (t - 36)] = s;
  }
}
// second pass for off-diagonal tiles; mirrors the block into the lower triangle (voxel_map.hpp:279-281)
__global__ void k_fin_off(const int64_t *__restrict__ tile_key, const int32_t *__restrict__ oseg_off,
                          const double *__restrict__ opart, double *__restrict__ H, int W, int64_t F) {
  const int tl = blockIdx.x;
  const int t = threadIdx.x;
  if (t >= 36) return;
  double s = 0.0;
  for (int32_t x = oseg_off[tl]; x < oseg_off[tl + 1]; x++) s += opart[36 * (int64_t)x + t];
  const int64_t key = tile_key[tl];
  const int fi = (int)(key / F), fj = (int)(key % F);
  const int w = fi / W, i = fi % W, j = fj % W;
  const int64_t jl = 6 * (int64_t)W;
  const int r = t % 6, c = t / 6;
  double *Hw = H + w * jl * jl;
  Hw[(6 * j + c) * jl + (6 * i + r)] = s;
  Hw[(6 * i + r) * jl + (6 * j + c)] = s;
}

}  // namespace

// ----------------------------------------------------------------------------------------------
// host-side glue
// ----------------------------------------------------------------------------------------------
static int launch_residual(vx_ctx *ctx, vx_factor *f, int64_t head, int64_t end) {
  cudaStream_t st = ctx->stream;
  f->n_res_evals++;
  if (f->V > 0) {
    static const int res_minb = [] { const char *e = std::getenv("VX_RES_MINB"); return e ? std::atoi(e) : 6; }();   // 80 registers, 24 warps per SM: measured best
#define VX_RES_LAUNCH(MB) VX_KERNEL(ctx, "k_residual", k_residual<MB><<<blocks_for(f->V, RES_THREADS), RES_THREADS, 0, st>>>( \
        f->off.as<int32_t>(), f->frame.as<int32_t>(), f->mom.as<double>(), f->cnt.as<int32_t>(), f->nnz, f->fix.as<double>(), \
        f->coe.as<double>(), f->poses.as<double>(), f->lam.as<double>(), f->U.as<double>(), f->mrg.as<double>(), \
        f->res_v.as<double>(), f->cS.as<double>(), f->V, head, end))
    // Measured on B200, round-1 workload (399,000 voxels, 2.39 M records; scripts/b12_prof.py): thread-per-voxel loop 0.103 ms per
    // launch, staged form 0.112 (6 CTAs/SM) ... 0.131 (3 CTAs/SM).  The pass issues ~2,800 instructions per voxel, two thirds of
    // them the FP64 eigen-solve: the two barriers per chunk cost more than the coalescing returns.  VX_RES_STAGED=1 selects it.
    static const bool staged = std::getenv("VX_RES_STAGED") != nullptr;
    static const int stg_minb = [] { const char *e = std::getenv("VX_RES_STAGED_MINB"); return e ? std::atoi(e) : 6; }();
#define VX_RES_LAUNCH_STAGED(MB) VX_KERNEL(ctx, "k_residual", k_residual_staged<MB><<<blocks_for(f->V, RES_THREADS), RES_THREADS, 0, st>>>( \
        f->off.as<int32_t>(), f->frame.as<int32_t>(), f->mom.as<double>(), f->cnt.as<int32_t>(), f->nnz, f->fix.as<double>(), \
        f->coe.as<double>(), f->poses.as<double>(), f->lam.as<double>(), f->U.as<double>(), f->mrg.as<double>(), \
        f->res_v.as<double>(), f->cS.as<double>(), f->V, head, end))
    if (staged) {
      if (stg_minb >= 6) VX_RES_LAUNCH_STAGED(6);
      else if (stg_minb >= 5) VX_RES_LAUNCH_STAGED(5);
      else if (stg_minb >= 4) VX_RES_LAUNCH_STAGED(4);
      else if (stg_minb >= 3) VX_RES_LAUNCH_STAGED(3);
      else VX_RES_LAUNCH_STAGED(0);
    }
    else if (res_minb >= 8) VX_RES_LAUNCH(8);
    else if (res_minb >= 7) VX_RES_LAUNCH(7);
    else if (res_minb >= 6) VX_RES_LAUNCH(6);
    else VX_RES_LAUNCH(0);
#undef VX_RES_LAUNCH
#undef VX_RES_LAUNCH_STAGED
    f->prep_lo = std::max<int64_t>(head, 0); f->prep_hi = std::min<int64_t>(end, f->V);
  }
  VX_KERNEL(ctx, "k_cost", k_cost<false><<<f->n_win, 256, 0, st>>>(f->res_v.as<double>(), nullptr, f->wv_off.as<int32_t>(), f->cost_r.as<double>(), head, end));
  return VX_OK;
}

// host destinations of a grouped evaluation (vx_factor_hessian on a batched factor)
struct GroupOut { double *Hess, *JacT; int groups; };

static int launch_hessian(vx_ctx *ctx, vx_factor *f, int64_t head, int64_t end, const GroupOut *group_out = nullptr) {
  cudaStream_t st = ctx->stream;
  f->n_hess_evals++;
  const size_t jl = (size_t)6 * f->W;
  bool fused = false;
  if (f->V > 0 && f->nnz > 0 && !f->force_pairlist && hessian_small_handles(f->W)) {
    // cost (+ the per-voxel divisions unless the last residual pass left them), then the fused kernel (overwrites all of H, J)
    const int64_t lo = std::max<int64_t>(head, 0), hi = std::min<int64_t>(end, f->V);
    // covered: the cost comes out of the finaliser (k_fin_small, same sums as k_cost); uncovered: k_cost_prep writes both
    const bool covered = f->prep_lo <= lo && hi <= f->prep_hi;
    if (!covered) {
      VX_KERNEL(ctx, "k_cost_prep", k_cost_prep<<<f->n_win, 1024, 0, st>>>(f->lam.as<double>(), f->coe.as<double>(), f->mrg.as<double>(), f->wv_off.as<int32_t>(),
                                                                       f->cost_h.as<double>(), f->cS.as<double>(), f->V, head, end));
      f->prep_lo = lo; f->prep_hi = hi;
    }
    if (!group_out) {
      VX_TRY(hessian_small(ctx, f, head, end, &fused, 0, -1, covered));
      return VX_OK;
    }
    // host output of a batch: the windows are evaluated in groups, each on its own stream with its Hess/JacT copy behind its
    // last kernel.  The 11.7 MB of Hess at the bench workload take as long over PCIe as two thirds of the evaluation; with
    // concurrent groups the kernels keep the SMs as full as one launch would (a later group's CTAs take the slots an earlier
    // one's last wave leaves idle) and only the last group's copy is exposed.  The group streams carry descending priorities
    // (vx_ctx.cu): without them the block scheduler interleaves the groups, all finish together and no copy starts early
    // (measured: 0.61 ms without priorities, 0.48 ms with, 0.557 ms for groups run back to back on one stream).  Groups that
    // shrink towards the end (4 4 4 4 4 3 2 1) did not help: 0.49 ms.
    const size_t jl6 = (size_t)6 * f->W;
    const int G = group_out->groups;
    VX_CUDA(cudaEventRecord(ctx->fork_ev, st));
    for (int gi = 0; gi < G; gi++) {
      const int w_lo = (int)((int64_t)f->n_win * gi / G), w_hi = (int)((int64_t)f->n_win * (gi + 1) / G);
      if (w_hi <= w_lo) continue;
      cudaStream_t gs = ctx->group_stream[gi];
      VX_CUDA(cudaStreamWaitEvent(gs, ctx->fork_ev, 0));
      ctx->stream = gs;                                   // hessian_small launches (and times) on the context stream
      const int rc = hessian_small(ctx, f, head, end, &fused, w_lo, w_hi, covered);
      ctx->stream = st;
      VX_TRY(rc);
      if (group_out->Hess) VX_CUDA(cudaMemcpyAsync(group_out->Hess + (size_t)w_lo * jl6 * jl6, f->H.as<double>() + (size_t)w_lo * jl6 * jl6,
                                                   sizeof(double) * jl6 * jl6 * (w_hi - w_lo), cudaMemcpyDeviceToHost, gs));
      if (group_out->JacT) VX_CUDA(cudaMemcpyAsync(group_out->JacT + (size_t)w_lo * jl6, f->J.as<double>() + (size_t)w_lo * jl6,
                                                   sizeof(double) * jl6 * (w_hi - w_lo), cudaMemcpyDeviceToHost, gs));
      VX_CUDA(cudaEventRecord(ctx->group_ev[gi], gs));
      VX_CUDA(cudaStreamWaitEvent(st, ctx->group_ev[gi], 0));   // the main stream is complete when every group is
    }
    return VX_OK;
  }
  if (!fused) {
    if (!f->plan_ready) VX_TRY(factor_build_plan(f));
    VX_CUDA(cudaMemsetAsync(f->H.p, 0, sizeof(double) * jl * jl * f->n_win, st));
    VX_CUDA(cudaMemsetAsync(f->J.p, 0, sizeof(double) * jl * f->n_win, st));
  }
  if (!fused && f->V > 0 && f->nnz > 0) {
    VX_KERNEL(ctx, "k_voxel_prep", k_voxel_prep<<<blocks_for(f->V, 256), 256, 0, st>>>(f->lam.as<double>(), f->mrg.as<double>(), f->coe.as<double>(), f->cS.as<double>(), f->V));
    if (f->n_dseg > 0) {
      VX_KERNEL(ctx, "k_diag", k_diag<<<(int)f->n_dseg, DIAG_THREADS, 0, st>>>(
          f->dseg_frame.as<int32_t>(), f->dseg_beg.as<int32_t>(), f->dseg_end.as<int32_t>(), f->fm_order.as<int32_t>(),
          f->vox.as<int32_t>(), f->mom.as<double>(), f->cnt.as<int32_t>(), f->nnz, f->lam.as<double>(), f->U.as<double>(),
          f->mrg.as<double>(), f->coe.as<double>(), f->cS.as<double>(), f->poses.as<double>(), f->G.as<double>(),
          f->dpart.as<double>(), f->V, head, end));
      VX_KERNEL(ctx, "k_fin_diag", k_fin_diag<<<f->F(), 64, 0, st>>>(f->dseg_off.as<int32_t>(), f->dpart.as<double>(), f->H.as<double>(), f->J.as<double>(), f->W));
    }
    if (f->n_oseg > 0) {
      if (f->n_pairs < 32 * f->n_oseg)     // short segments on average: a warp each
        VX_KERNEL(ctx, "k_offdiag_warp", k_offdiag_warp<<<blocks_for(f->n_oseg, 8), 256, 0, st>>>(
            f->oseg_beg.as<int32_t>(), f->oseg_end.as<int32_t>(), f->pair_ab.as<int64_t>(), f->vox.as<int32_t>(),
            f->cS.as<double>(), f->G.as<double>(), f->opart.as<double>(), f->V, head, end, f->n_oseg));
      else
        VX_KERNEL(ctx, "k_offdiag", k_offdiag<<<(int)f->n_oseg, DIAG_THREADS, 0, st>>>(
            f->oseg_beg.as<int32_t>(), f->oseg_end.as<int32_t>(), f->pair_ab.as<int64_t>(), f->vox.as<int32_t>(),
            f->cS.as<double>(), f->G.as<double>(), f->opart.as<double>(), f->V, head, end));
      VX_KERNEL(ctx, "k_fin_off", k_fin_off<<<(int)f->n_tiles, 64, 0, st>>>(f->tile_key.as<int64_t>(), f->oseg_off.as<int32_t>(), f->opart.as<double>(),
                                             f->H.as<double>(), f->W, f->F()));
    }
  }
  VX_KERNEL(ctx, "k_cost", k_cost<true><<<f->n_win, 256, 0, st>>>(f->lam.as<double>(), f->coe.as<double>(), f->wv_off.as<int32_t>(), f->cost_h.as<double>(), head, end));
  return VX_OK;
}

// ---- upload: AoS host records -> SoA planes, transposed on the device ---------------------------
// One thread per voxel-frame: reads the 104-byte reference PointCluster record (tools.hpp:308-310)
// src[idx ? idx[e] : e] and scatters P's upper triangle, v and N into the component planes.
__global__ void k_unpack_vf(const vx_cluster *__restrict__ src, const int64_t *__restrict__ idx, double *__restrict__ mom,
                            int32_t *__restrict__ cnt, int64_t nnz, int *__restrict__ err) {
  int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (e >= nnz) return;
  const vx_cluster &c = src[idx ? idx[e] : e];
  if (c.N == 0) atomicMax(err, 4);
  mom[0 * nnz + e] = c.P[0]; mom[1 * nnz + e] = c.P[3]; mom[2 * nnz + e] = c.P[6];
  mom[3 * nnz + e] = c.P[4]; mom[4 * nnz + e] = c.P[7]; mom[5 * nnz + e] = c.P[8];
  mom[6 * nnz + e] = c.v[0]; mom[7 * nnz + e] = c.v[1]; mom[8 * nnz + e] = c.v[2];
  cnt[e] = c.N;
}
// per voxel: sig_vecs -> fix planes, pcr_adds -> mrg planes, eig_values/eig_vectors -> lam/U planes
__global__ void k_unpack_voxel(const vx_cluster *__restrict__ sig, const vx_cluster *__restrict__ add,
                               const double *__restrict__ ev, const double *__restrict__ evec, double *__restrict__ fix,
                               double *__restrict__ mrg, double *__restrict__ lam, double *__restrict__ U, int64_t V) {
  int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  {
    const vx_cluster &c = sig[v];
    fix[0 * V + v] = c.P[0]; fix[1 * V + v] = c.P[3]; fix[2 * V + v] = c.P[6];
    fix[3 * V + v] = c.P[4]; fix[4 * V + v] = c.P[7]; fix[5 * V + v] = c.P[8];
    fix[6 * V + v] = c.v[0]; fix[7 * V + v] = c.v[1]; fix[8 * V + v] = c.v[2];
    fix[9 * V + v] = (double)c.N;
  }
  {
    const vx_cluster &c = add[v];
    mrg[0 * V + v] = c.P[0]; mrg[1 * V + v] = c.P[3]; mrg[2 * V + v] = c.P[6];
    mrg[3 * V + v] = c.P[4]; mrg[4 * V + v] = c.P[7]; mrg[5 * V + v] = c.P[8];
    mrg[6 * V + v] = c.v[0]; mrg[7 * V + v] = c.v[1]; mrg[8 * V + v] = c.v[2];
    mrg[9 * V + v] = (double)c.N;
  }
#pragma unroll
  for (int k = 0; k < 3; k++) lam[k * V + v] = ev[3 * v + k];
#pragma unroll
  for (int k = 0; k < 9; k++) U[k * V + v] = evec[9 * v + k];
}

// Device-side validation of the CSR structure + window of each voxel.  err: 1 bad frame id / voxel spans windows,
// 2 frames not strictly ascending, 3 voxels not grouped by ascending window, 5 empty voxel in a batched factor.
__global__ void k_validate_vox(const int32_t *__restrict__ off, const int32_t *__restrict__ frame, int64_t V, int W, int NW, int32_t *__restrict__ win,
                               int *__restrict__ err) {
  const int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (v >= V) return;
  const int32_t b = off[v], e = off[v + 1];
  if (e < b) { atomicMax(err, 1); win[v] = 0; return; }
  if (e == b) { if (NW > 1) atomicMax(err, 5); win[v] = 0; return; }
  const int32_t w = frame[b] / W;
  for (int32_t x = b; x < e; x++) {
    const int32_t fr = frame[x];
    if (fr < 0 || fr >= W * NW || fr / W != w) atomicMax(err, 1);
    if (x > b && fr <= frame[x - 1]) atomicMax(err, 2);
  }
  win[v] = w;
  if (v > 0 && off[v] > off[v - 1] && frame[off[v - 1]] / W > w) atomicMax(err, 3);
}
__global__ void k_win_offsets(const int32_t *__restrict__ win, int64_t V, int NW, int32_t *__restrict__ wv_off) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w > NW) return;
  int64_t lo = 0, hi = V;
  while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (win[mid] < w) lo = mid + 1; else hi = mid; }
  wv_off[w] = (int32_t)lo;
}

// `cl_host` holds n_src records; voxel-frame e is record idx[e] (idx may be null = identity).  `frames`: nnz ids.
static int upload_common(vx_factor *f, int64_t V, const std::vector<int32_t> &off, const int32_t *frames, int64_t nnz,
                         const vx_cluster *cl_host, int64_t n_src, const std::vector<int64_t> *idx, const vx_cluster *sig_vecs,
                         const double *coeffs, const double *eig_values, const double *eig_vectors, const vx_cluster *pcr_adds) {
  vx_ctx *ctx = f->ctx;
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  const int W = f->W, NW = f->n_win;
  VX_TRY(factor_alloc(f, V, nnz));
#define VX_H2D(dst, src, bytes) VX_CUDA(cudaMemcpyAsync((dst), (src), (bytes), cudaMemcpyHostToDevice, st))
  VX_H2D(f->off.p, off.data(), sizeof(int32_t) * (V + 1));
  // staging area for the raw AoS records
  DevBuf &raw = ctx->scratch2;
  const size_t b_cl = sizeof(vx_cluster) * (size_t)n_src, b_idx = idx ? sizeof(int64_t) * (size_t)nnz : 0;
  const size_t b_vc = sizeof(vx_cluster) * (size_t)V, b_ev = sizeof(double) * 3 * (size_t)V, b_evec = sizeof(double) * 9 * (size_t)V;
  VX_TRY(raw.reserve(b_cl + b_idx + 2 * b_vc + b_ev + b_evec + 64));
  char *base = raw.as<char>();
  vx_cluster *d_cl = (vx_cluster *)base;
  int64_t *d_idx = (int64_t *)(base + b_cl);
  vx_cluster *d_sig = (vx_cluster *)(base + b_cl + b_idx);
  vx_cluster *d_add = (vx_cluster *)(base + b_cl + b_idx + b_vc);
  double *d_ev = (double *)(base + b_cl + b_idx + 2 * b_vc);
  double *d_evec = (double *)(base + b_cl + b_idx + 2 * b_vc + b_ev);
  int *d_err = (int *)(base + b_cl + b_idx + 2 * b_vc + b_ev + b_evec);
  VX_CUDA(cudaMemsetAsync(d_err, 0, sizeof(int), st));
  if (nnz > 0) {
    VX_H2D(f->frame.p, frames, sizeof(int32_t) * nnz);
    VX_H2D(d_cl, cl_host, b_cl);
    if (idx) VX_H2D(d_idx, idx->data(), b_idx);
    VX_KERNEL(ctx, "k_unpack_vf", k_unpack_vf<<<blocks_for(nnz, 256), 256, 0, st>>>(d_cl, idx ? d_idx : nullptr, f->mom.as<double>(), f->cnt.as<int32_t>(), nnz, d_err));
  }
  if (V > 0) {
    VX_H2D(d_sig, sig_vecs, b_vc);
    VX_H2D(d_add, pcr_adds, b_vc);
    VX_H2D(d_ev, eig_values, b_ev);
    VX_H2D(d_evec, eig_vectors, b_evec);
    VX_H2D(f->coe.p, coeffs, sizeof(double) * V);
    VX_KERNEL(ctx, "k_unpack_voxel", k_unpack_voxel<<<blocks_for(V, 256), 256, 0, st>>>(d_sig, d_add, d_ev, d_evec, f->fix.as<double>(), f->mrg.as<double>(),
                                                                                     f->lam.as<double>(), f->U.as<double>(), V));
    VX_KERNEL(ctx, "k_validate_vox", k_validate_vox<<<blocks_for(V, 256), 256, 0, st>>>(f->off.as<int32_t>(), f->frame.as<int32_t>(), V, W, NW, f->win.as<int32_t>(), d_err));
  }
  VX_KERNEL(ctx, "k_win_offsets", k_win_offsets<<<blocks_for(NW + 1, 128), 128, 0, st>>>(f->win.as<int32_t>(), V, NW, f->wv_off.as<int32_t>()));
#undef VX_H2D
  int err = 0;
  VX_CUDA(cudaMemcpyAsync(&err, d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
  VX_CUDA(cudaStreamSynchronize(st));   // host staging goes out of scope
  if (err) {
    const char *what[] = {"", "a voxel spans windows or has a bad frame id", "frames of a voxel are not strictly ascending",
                          "voxels must be grouped by ascending window", "a voxel-frame cluster has N == 0", "a batched factor has a voxel without frames"};
    set_error("factor upload: %s", what[err < 6 ? err : 1]);
    f->V = 0; f->nnz = 0;
    return VX_ERR_ARG;
  }
  return factor_finish_upload(f);
}

extern "C" {

int vx_factor_create(vx_ctx *ctx, int win_size, int n_windows, vx_factor **out) {
  VX_ARG(out != nullptr);
  *out = nullptr;
  VX_TRY(bind(ctx));
  VX_ARG(win_size >= 1 && n_windows >= 1);
  VX_ARG((int64_t)win_size * n_windows < (1 << 30));
  vx_factor *f = new vx_factor();
  f->ctx = ctx; f->W = win_size; f->n_win = n_windows;
  int s = factor_alloc(f, 0, 0);
  if (s != VX_OK) { delete f; return s; }
  std::vector<int32_t> z(n_windows + 2, 0);
  VX_CUDA(cudaMemcpyAsync(f->wv_off.p, z.data(), sizeof(int32_t) * (n_windows + 1), cudaMemcpyHostToDevice, ctx->stream));
  VX_CUDA(cudaMemcpyAsync(f->off.p, z.data(), sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  f->plan_ready = true;
  *out = f;
  return VX_OK;
}

int vx_factor_free(vx_factor *f) {
  if (!f) return VX_OK;
  cudaSetDevice(f->ctx->device);
  cudaStreamSynchronize(f->ctx->stream);
  vx::DevBuf *all[] = {&f->off, &f->frame, &f->vox, &f->mom, &f->cnt, &f->fix, &f->coe, &f->lam, &f->U, &f->mrg, &f->win, &f->wv_off,
                       &f->poses, &f->H, &f->J, &f->cost_h, &f->cost_r, &f->cS, &f->res_v, &f->G, &f->fm_order, &f->fm_off,
                       &f->dseg_frame, &f->dseg_beg, &f->dseg_end, &f->dseg_off, &f->pair_ab, &f->tile_key, &f->tile_off,
                       &f->oseg_tile, &f->oseg_beg, &f->oseg_end, &f->oseg_off, &f->dpart, &f->opart, &f->part_t, &f->part_d,
                       &f->lm_H, &f->lm_J, &f->lm_dx, &f->lm_q1, &f->lm_u, &f->lm_mask};
  for (auto *b : all) b->release();
  delete f;
  return VX_OK;
}

int vx_factor_clear(vx_factor *f) {
  VX_ARG(f != nullptr);
  VX_TRY(bind(f->ctx));
  f->V = 0; f->nnz = 0; f->n_pairs = f->n_tiles = f->n_dseg = f->n_oseg = 0;
  std::vector<int32_t> z(f->n_win + 2, 0);
  VX_CUDA(cudaMemcpyAsync(f->wv_off.p, z.data(), sizeof(int32_t) * (f->n_win + 1), cudaMemcpyHostToDevice, f->ctx->stream));
  VX_CUDA(cudaStreamSynchronize(f->ctx->stream));
  f->plan_ready = true;
  return VX_OK;
}

int vx_factor_upload_dense(vx_factor *f, int64_t V, const vx_cluster *plvec, const vx_cluster *sig_vecs, const double *coeffs,
                           const double *eig_values, const double *eig_vectors, const vx_cluster *pcr_adds) {
  VX_ARG(f != nullptr && V >= 0);
  VX_ARG(V == 0 || (plvec && sig_vecs && coeffs && eig_values && eig_vectors && pcr_adds));
  VX_ARG(f->n_win == 1);
  const int W = f->W;
  std::vector<int32_t> off(V + 1, 0), frames;
  std::vector<int64_t> idx;
  for (int64_t v = 0; v < V; v++) {
    for (int i = 0; i < W; i++) {
      const vx_cluster *c = plvec + v * W + i;
      if (c->N != 0) { frames.push_back(i); idx.push_back(v * W + i); }
    }
    VX_ARG(frames.size() < ((size_t)1 << 31));
    off[v + 1] = (int32_t)frames.size();
  }
  return upload_common(f, V, off, frames.data(), (int64_t)frames.size(), plvec, V * W, &idx, sig_vecs, coeffs, eig_values, eig_vectors, pcr_adds);
}

int vx_factor_upload_csr(vx_factor *f, int64_t V, const int64_t *offsets, const int32_t *frames_in, const vx_cluster *clusters,
                         const vx_cluster *sig_vecs, const double *coeffs, const double *eig_values, const double *eig_vectors,
                         const vx_cluster *pcr_adds) {
  VX_ARG(f != nullptr && V >= 0 && offsets != nullptr);
  const int64_t nnz = offsets[V];
  VX_ARG(nnz >= 0 && nnz < ((int64_t)1 << 31));
  VX_ARG(V == 0 || (sig_vecs && coeffs && eig_values && eig_vectors && pcr_adds));
  VX_ARG(nnz == 0 || (frames_in && clusters));
  std::vector<int32_t> off(V + 1);
  for (int64_t v = 0; v <= V; v++) { VX_ARG(offsets[v] >= 0 && offsets[v] <= nnz && (v == 0 || offsets[v] >= offsets[v - 1])); off[v] = (int32_t)offsets[v]; }
  return upload_common(f, V, off, frames_in, nnz, clusters, nnz, nullptr, sig_vecs, coeffs, eig_values, eig_vectors, pcr_adds);
}

int vx_factor_download_cache(vx_factor *f, double *eig_values, double *eig_vectors, vx_cluster *pcr_adds) {
  VX_ARG(f != nullptr);
  vx_ctx *ctx = f->ctx;
  VX_TRY(bind(ctx));
  const int64_t V = f->V;
  if (V == 0) return VX_OK;
  std::vector<double> lam(3 * V), U(9 * V), mrg(10 * V);
  VX_CUDA(cudaMemcpyAsync(lam.data(), f->lam.p, sizeof(double) * 3 * V, cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaMemcpyAsync(U.data(), f->U.p, sizeof(double) * 9 * V, cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaMemcpyAsync(mrg.data(), f->mrg.p, sizeof(double) * 10 * V, cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  for (int64_t v = 0; v < V; v++) {
    if (eig_values) for (int k = 0; k < 3; k++) eig_values[3 * v + k] = lam[k * V + v];
    if (eig_vectors) for (int k = 0; k < 9; k++) eig_vectors[9 * v + k] = U[k * V + v];
    if (pcr_adds) {
      vx_cluster &c = pcr_adds[v];
      c.P[0] = mrg[0 * V + v]; c.P[1] = c.P[3] = mrg[1 * V + v]; c.P[2] = c.P[6] = mrg[2 * V + v];
      c.P[4] = mrg[3 * V + v]; c.P[5] = c.P[7] = mrg[4 * V + v]; c.P[8] = mrg[5 * V + v];
      c.v[0] = mrg[6 * V + v]; c.v[1] = mrg[7 * V + v]; c.v[2] = mrg[8 * V + v];
      c.N = (int32_t)llround(mrg[9 * V + v]); c._pad = 0;
    }
  }
  return VX_OK;
}

int vx_factor_download_csr(vx_factor *f, int64_t *offsets, int32_t *frames, vx_cluster *clusters, double *coeffs) {
  VX_ARG(f != nullptr);
  vx_ctx *ctx = f->ctx;
  VX_TRY(bind(ctx));
  const int64_t V = f->V, nnz = f->nnz;
  cudaStream_t st = ctx->stream;
  if (offsets) {
    std::vector<int32_t> off(V + 1, 0);
    if (V > 0) VX_CUDA(cudaMemcpyAsync(off.data(), f->off.p, sizeof(int32_t) * (V + 1), cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    for (int64_t v = 0; v <= V; v++) offsets[v] = off[v];
  }
  if (coeffs && V > 0) VX_CUDA(cudaMemcpyAsync(coeffs, f->coe.p, sizeof(double) * V, cudaMemcpyDeviceToHost, st));
  if (frames && nnz > 0) VX_CUDA(cudaMemcpyAsync(frames, f->frame.p, sizeof(int32_t) * nnz, cudaMemcpyDeviceToHost, st));
  if (clusters && nnz > 0) {
    std::vector<double> mom(9 * (size_t)nnz);
    std::vector<int32_t> cnt(nnz);
    VX_CUDA(cudaMemcpyAsync(mom.data(), f->mom.p, sizeof(double) * 9 * nnz, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaMemcpyAsync(cnt.data(), f->cnt.p, sizeof(int32_t) * nnz, cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    for (int64_t e = 0; e < nnz; e++) {
      vx_cluster &c = clusters[e];
      c.P[0] = mom[0 * nnz + e]; c.P[1] = c.P[3] = mom[1 * nnz + e]; c.P[2] = c.P[6] = mom[2 * nnz + e];
      c.P[4] = mom[3 * nnz + e]; c.P[5] = c.P[7] = mom[4 * nnz + e]; c.P[8] = mom[5 * nnz + e];
      c.v[0] = mom[6 * nnz + e]; c.v[1] = mom[7 * nnz + e]; c.v[2] = mom[8 * nnz + e];
      c.N = cnt[e]; c._pad = 0;
    }
  }
  VX_CUDA(cudaStreamSynchronize(st));
  return VX_OK;
}

int vx_factor_force_generic(vx_factor *f, int on) {
  VX_ARG(f != nullptr);
  f->force_pairlist = on != 0;
  return VX_OK;
}

int64_t vx_factor_num_voxels(const vx_factor *f) { return f ? f->V : 0; }
int64_t vx_factor_num_voxel_frames(const vx_factor *f) { return f ? f->nnz : 0; }
int vx_factor_win_size(const vx_factor *f) { return f ? f->W : 0; }
int vx_factor_num_windows(const vx_factor *f) { return f ? f->n_win : 0; }

int vx_factor_set_poses(vx_ctx *ctx, vx_factor *f, const vx_pose *poses) {
  VX_ARG(f != nullptr && poses != nullptr);
  VX_TRY(bind(ctx));
  static_assert(sizeof(vx_pose) == 12 * sizeof(double), "vx_pose layout");
  VX_CUDA(cudaMemcpyAsync(f->poses.p, poses, sizeof(vx_pose) * f->F(), cudaMemcpyHostToDevice, ctx->stream));
  return VX_OK;
}

static void clamp_range(const vx_factor *f, int64_t &head, int64_t &end) {
  if (head < 0) head = 0;
  if (end < 0 || end > f->V) end = f->V;
}

int vx_factor_hessian_async(vx_ctx *ctx, vx_factor *f, int64_t head, int64_t end) {
  VX_ARG(f != nullptr);
  VX_TRY(bind(ctx));
  clamp_range(f, head, end);
  return launch_hessian(ctx, f, head, end);
}
int vx_factor_residual_async(vx_ctx *ctx, vx_factor *f, int64_t head, int64_t end) {
  VX_ARG(f != nullptr);
  VX_TRY(bind(ctx));
  clamp_range(f, head, end);
  return launch_residual(ctx, f, head, end);
}

int vx_factor_device_buffers(vx_factor *f, void **Hess, void **JacT, void **cost_hess, void **cost_resid) {
  VX_ARG(f != nullptr);
  if (Hess) *Hess = f->H.p;
  if (JacT) *JacT = f->J.p;
  if (cost_hess) *cost_hess = f->cost_h.p;
  if (cost_resid) *cost_resid = f->cost_r.p;
  return VX_OK;
}

int vx_factor_hessian(vx_ctx *ctx, vx_factor *f, const vx_pose *poses, int64_t head, int64_t end, double *Hess, double *JacT, double *residual) {
  VX_ARG(f != nullptr && poses != nullptr);
  VX_TRY(vx_factor_set_poses(ctx, f, poses));
  const size_t jl = (size_t)6 * f->W;
  cudaStream_t st = ctx->stream;
  const bool grouped = f->n_win >= 32 && (Hess || JacT) && f->V > 0 && f->nnz > 0 && !f->force_pairlist && hessian_small_handles(f->W);
  if (grouped) {
    static const int groups = getenv("VX_HESS_GROUPS") ? std::max(1, std::min(8, atoi(getenv("VX_HESS_GROUPS")))) : 8;
    GroupOut go{Hess, JacT, groups};
    clamp_range(f, head, end);
    VX_TRY(launch_hessian(ctx, f, head, end, &go));
  } else {
    VX_TRY(vx_factor_hessian_async(ctx, f, head, end));
    if (Hess) VX_CUDA(cudaMemcpyAsync(Hess, f->H.p, sizeof(double) * jl * jl * f->n_win, cudaMemcpyDeviceToHost, st));
    if (JacT) VX_CUDA(cudaMemcpyAsync(JacT, f->J.p, sizeof(double) * jl * f->n_win, cudaMemcpyDeviceToHost, st));
  }
  if (residual) VX_CUDA(cudaMemcpyAsync(residual, f->cost_h.p, sizeof(double) * f->n_win, cudaMemcpyDeviceToHost, st));
  VX_CUDA(cudaStreamSynchronize(st));
  return VX_OK;
}

int vx_factor_residual(vx_ctx *ctx, vx_factor *f, const vx_pose *poses, int64_t head, int64_t end, double *residual) {
  VX_ARG(f != nullptr && poses != nullptr);
  VX_TRY(vx_factor_set_poses(ctx, f, poses));
  VX_TRY(vx_factor_residual_async(ctx, f, head, end));
  if (residual) VX_CUDA(cudaMemcpyAsync(residual, f->cost_r.p, sizeof(double) * f->n_win, cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  return VX_OK;
}

}  // extern "C"
