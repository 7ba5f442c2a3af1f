// vx_gba.cu — the global-BA voxel map as a sort/segment pass (SURVEY §8a rows A8 + A10):
//   OctreeGBA::cut_voxel          loop_refine.hpp:446-479   world = R*local + p, root key, per-frame point lists
//   OctreeGBA::recut / subdivide  loop_refine.hpp:323-405   eigen-solve, plane test, split to max_layer
//   OctreeGBA_multi_recut         loop_refine.hpp:483-537   -> LidarFactor::push_voxel for every accepted plane
// The reference grows a pointer octree per root voxel.  Here the octree is a KEY: every point's path
// through the (up to max_layer) subdivisions is a pure function of the point and its root cell —
// child = 4*[x>cx] + 2*[y>cy] + [z>cz] against centres built by +-quater_length (loop_refine.hpp:331-347) —
// so a node of layer l is a run of equal (window, root key, first 3l path bits) in the sorted point
// array, and "recut" is a segmented reduction + a per-node decision, evaluated for all layers at once:
//   1. per point instance: world point (no FMA contraction, so cells are bit-exact), packed root key, path
//   2. stable radix sorts -> order (window, root key, path, frame, original order)
//   3. one head-bit byte per sorted instance, one fused scan of the per-layer / per-segment head counters -> COMPACT index arrays
//      (segment starts, and per node: start, first segment, first leaf, first child, parent)
//   4. finest layer: world clusters straight from the sorted points (per frame, then summed: the reference's association)
//   5. coarser layers from their children; per node 3x3 eigen-solve, plane / accept / subdivide flags, top-down liveness
//   6. accepted planes -> voxels (ordered by position), segments -> (voxel, frame) scan-local clusters -> device LidarFactor
#include <cstdlib>
#include <cstdio>
#include "vx_factor.cuh"
#include "vx_math.cuh"
#include "vx_voxelize.cuh"
#include <cub/cub.cuh>
#include <algorithm>

using namespace vx;

namespace {

constexpr int MAXL = 3;          // max_layer supported (reference default 2)
constexpr int FRAME_BITS = 22;

struct GbaP {
  double vs, inv_vs, min_eig, thr[4];
  int L;
};

// world = R*local + p exactly as a straight-line double evaluation without contraction:
// ((R(r,0)*l0 + R(r,1)*l1) + R(r,2)*l2) + p[r]
__device__ __forceinline__ d3 world_of(const double *__restrict__ ps, double lx, double ly, double lz) {
  d3 w;
  w.x = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(ps[0], lx), __dmul_rn(ps[3], ly)), __dmul_rn(ps[6], lz)), ps[9]);
  w.y = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(ps[1], lx), __dmul_rn(ps[4], ly)), __dmul_rn(ps[7], lz)), ps[10]);
  w.z = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(ps[2], lx), __dmul_rn(ps[5], ly)), __dmul_rn(ps[8], lz)), ps[11]);
  return w;
}

__device__ __forceinline__ int frame_of(const int64_t *__restrict__ inst_off, int F, int64_t j) {
  int lo = 0, hi = F;   // largest g with inst_off[g] <= j
  while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (inst_off[mid] <= j) lo = mid; else hi = mid; }
  return lo;
}

// frame of instance j given g0 <= frame(j) (the frame of the CTA's first instance): a walk of 0 or 1 steps for real clouds
__device__ __forceinline__ int frame_walk(const int64_t *__restrict__ inst_off, int F, int g0, int64_t j) {
  int g = g0;
  while (g + 1 < F && inst_off[g + 1] <= j) g++;
  return g;
}
// key_axis_d without the FP64 division on the common path: q = coord * (1 / voxel_size) is within 2.3e-16 |q| of the exact quotient
// and the correctly rounded quotient within 1.2e-16 of it; when both ends of that interval round to the SAME float, so does the
// reference's quotient (rounding is monotone).  Otherwise (a quotient within ~1e-16 of a float rounding boundary) the division.
__device__ __forceinline__ int64_t key_axis_fast(double coord, double vs, double inv_vs) {
  const double q = coord * inv_vs;
  const double m = fabs(q) * 6e-16;
  const float lo = __double2float_rn(q - m), hi = __double2float_rn(q + m);
  float loc = lo;
  if (lo != hi) loc = __double2float_rn(__ddiv_rn(coord, vs));
  if (loc < 0.0f) loc = __fadd_rn(loc, -1.0f);
  return (int64_t)loc;
}

// 1. keys
__global__ void k_gba_point(const float *__restrict__ xyz, const int64_t *__restrict__ cloud_off, const int32_t *__restrict__ frame_cloud,
                            const int64_t *__restrict__ inst_off, const int32_t *__restrict__ pt_frame, int F, const double *__restrict__ poses, GbaP prm,
                            int64_t NI, uint32_t *__restrict__ key1, uint64_t *__restrict__ key2, int32_t *__restrict__ idx, int *__restrict__ bad) {
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= NI) return;
  // cloud mode: instance j is a point of the cloud of frame g; explicit mode (pt_frame, the routed points of a sharded
  // build): instance j is point j and carries its frame
  const int g = pt_frame ? pt_frame[j] : frame_of(inst_off, F, j);
  const int64_t pt = pt_frame ? j : cloud_off[frame_cloud[g]] + (j - inst_off[g]);
  const double lx = (double)xyz[3 * pt], ly = (double)xyz[3 * pt + 1], lz = (double)xyz[3 * pt + 2];
  const d3 w = world_of(poses + 12 * (int64_t)g, lx, ly, lz);
  const int64_t kx = key_axis_d(w.x, prm.vs), ky = key_axis_d(w.y, prm.vs), kz = key_axis_d(w.z, prm.vs);
  uint64_t k2;
  if (!pack_key(kx, ky, kz, k2)) *bad = 1;
  // path: centre of the root = (0.5 + key) * voxel_size (loop_refine.hpp:469-471), quater_length = voxel_size/4 as float (:472)
  double cx = __dmul_rn(__dadd_rn(0.5, (double)kx), prm.vs), cy = __dmul_rn(__dadd_rn(0.5, (double)ky), prm.vs), cz = __dmul_rn(__dadd_rn(0.5, (double)kz), prm.vs);
  float q = (float)(prm.vs / 4.0);
  uint32_t path = 0;
  for (int l = 0; l < prm.L; l++) {
    const int bx = w.x > cx, by = w.y > cy, bz = w.z > cz;
    path = (path << 3) | (uint32_t)(4 * bx + 2 * by + bz);
    const double dq = (double)q;
    cx = __dadd_rn(cx, bx ? dq : -dq); cy = __dadd_rn(cy, by ? dq : -dq); cz = __dadd_rn(cz, bz ? dq : -dq);
    q = q / 2.0f;
  }
  key1[j] = (path << FRAME_BITS) | (uint32_t)g;
  key2[j] = k2;
  idx[j] = (int32_t)j;
}

template <class T>
__global__ void k_gather(const T *__restrict__ src, const int32_t *__restrict__ idx, T *__restrict__ dst, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[idx[i]];
}
__global__ void k_gather_win(const uint32_t *__restrict__ key1, const int32_t *__restrict__ idx, int W, uint32_t *__restrict__ win, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) win[i] = (key1[idx[i]] & ((1u << FRAME_BITS) - 1u)) / (uint32_t)W;
}

// ---- single-key path: (window | root cell re-biased by the scene's minimum cell | path | frame in window) in ONE 64-bit key,
// one radix sort over exactly the bits in use (43 bits for the 399-window bottom level of C4, 39 for its top level) instead of
// three stable sorts over 28 + 63 + 9 bits with gathers in between.  Cells are re-biased per axis by the minimum over all point
// instances (monotone per axis, so the voxel order (window, root cell x/y/z, path) is the multi-sort path's).
struct CellRange { int lo[3], hi[3]; };
__device__ __forceinline__ void inst_point(const float *__restrict__ xyz, const int64_t *__restrict__ cloud_off, const int32_t *__restrict__ frame_cloud,
                                           const int64_t *__restrict__ inst_off, const int32_t *__restrict__ pt_frame, int F, const double *__restrict__ poses,
                                           int64_t j, int &g, d3 &w) {
  g = pt_frame ? pt_frame[j] : frame_of(inst_off, F, j);
  const int64_t pt = pt_frame ? j : cloud_off[frame_cloud[g]] + (j - inst_off[g]);
  w = world_of(poses + 12 * (int64_t)g, (double)xyz[3 * pt], (double)xyz[3 * pt + 1], (double)xyz[3 * pt + 2]);
}
__global__ void __launch_bounds__(256)
k_gba_cell_range(const float *__restrict__ xyz, const int64_t *__restrict__ cloud_off, const int32_t *__restrict__ frame_cloud,
                 const int64_t *__restrict__ inst_off, const int32_t *__restrict__ pt_frame, int F, const double *__restrict__ poses, double vs, int64_t NI,
                 CellRange *__restrict__ range, int *__restrict__ bad) {
  int lo[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff}, hi[3] = {-0x7fffffff, -0x7fffffff, -0x7fffffff};
  for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < NI; j += (int64_t)gridDim.x * blockDim.x) {
    int g; d3 w;
    inst_point(xyz, cloud_off, frame_cloud, inst_off, pt_frame, F, poses, j, g, w);
    const int64_t k[3] = {key_axis_d(w.x, vs), key_axis_d(w.y, vs), key_axis_d(w.z, vs)};
#pragma unroll
    for (int a = 0; a < 3; a++) {
      if (k[a] < -(int64_t)KEY_BIAS || k[a] >= (int64_t)KEY_BIAS) *bad = 1;
      const int c = (int)max((int64_t)-0x7ffffff0, min((int64_t)0x7ffffff0, k[a]));
      lo[a] = min(lo[a], c); hi[a] = max(hi[a], c);
    }
  }
#pragma unroll
  for (int a = 0; a < 3; a++) {
    lo[a] = __reduce_min_sync(0xffffffffu, lo[a]); hi[a] = __reduce_max_sync(0xffffffffu, hi[a]);
    if ((threadIdx.x & 31) == 0) { atomicMin(&range->lo[a], lo[a]); atomicMax(&range->hi[a], hi[a]); }
  }
}
// Cloud mode: the cell range only sizes the key fields, so a superset will do — per cloud the bounding box of its (scan-local)
// points, read ONCE per unique point (the bottom level references every cloud from two windows), per frame instance the cells of
// the 8 transformed corners (world = R l + p is affine in l: the extremes sit at corners), one cell of margin for the rounding of
// interior points.  Replaces a pass that transformed and divided every point instance (1.5 of the 21 ms of the bottom-level build).
__global__ void __launch_bounds__(256) k_cloud_bbox(const float *__restrict__ xyz, const int64_t *__restrict__ cloud_off, float *__restrict__ bbox) {
  __shared__ float sh[6][8];
  const int c = blockIdx.x;
  const int64_t b = cloud_off[c], e = cloud_off[c + 1];
  float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
#pragma unroll 4
  for (int64_t i = b + threadIdx.x; i < e; i += 256) {
    const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    lo[0] = fminf(lo[0], x); hi[0] = fmaxf(hi[0], x); lo[1] = fminf(lo[1], y); hi[1] = fmaxf(hi[1], y); lo[2] = fminf(lo[2], z); hi[2] = fmaxf(hi[2], z);
  }
#pragma unroll
  for (int a = 0; a < 3; a++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o)); hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o)); }
    if ((threadIdx.x & 31) == 0) { sh[a][threadIdx.x >> 5] = lo[a]; sh[3 + a][threadIdx.x >> 5] = hi[a]; }
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    float v = sh[threadIdx.x][0];
    for (int k = 1; k < 8; k++) v = threadIdx.x < 3 ? fminf(v, sh[threadIdx.x][k]) : fmaxf(v, sh[threadIdx.x][k]);
    bbox[6 * (size_t)c + threadIdx.x] = v;
  }
}
__global__ void k_frame_cells(const float *__restrict__ bbox, const int32_t *__restrict__ frame_cloud, int F, const double *__restrict__ poses, double vs,
                              CellRange *__restrict__ range) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= F) return;
  const float *bb = bbox + 6 * (size_t)frame_cloud[g];
  if (!(bb[0] <= bb[3])) return;                       // empty cloud
  int lo[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff}, hi[3] = {-0x7fffffff, -0x7fffffff, -0x7fffffff};
  for (int c = 0; c < 8; c++) {
    const d3 w = world_of(poses + 12 * (int64_t)g, (double)bb[(c & 1) ? 3 : 0], (double)bb[(c & 2) ? 4 : 1], (double)bb[(c & 4) ? 5 : 2]);
    const int64_t k[3] = {key_axis_d(w.x, vs), key_axis_d(w.y, vs), key_axis_d(w.z, vs)};
#pragma unroll
    for (int a = 0; a < 3; a++) {
      const int q = (int)max((int64_t)-0x7ffffff0, min((int64_t)0x7ffffff0, k[a]));
      lo[a] = min(lo[a], q - 1); hi[a] = max(hi[a], q + 1);
    }
  }
#pragma unroll
  for (int a = 0; a < 3; a++) { atomicMin(&range->lo[a], lo[a]); atomicMax(&range->hi[a], hi[a]); }
}
struct KeyLayout { int lo[3]; int bx, by, bz, pbits, fbits, W; };
__global__ void __launch_bounds__(256)
k_gba_key64(const float *__restrict__ xyz, const int64_t *__restrict__ cloud_off, const int32_t *__restrict__ frame_cloud,
            const int64_t *__restrict__ inst_off, const int32_t *__restrict__ pt_frame, int F, const double *__restrict__ poses, GbaP prm, KeyLayout kl,
            int64_t NI, uint64_t *__restrict__ key, int32_t *__restrict__ idx, int *__restrict__ bad) {
  __shared__ int g0s;
  if (threadIdx.x == 0) g0s = pt_frame ? 0 : frame_of(inst_off, F, blockIdx.x * (int64_t)blockDim.x);
  __syncthreads();
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= NI) return;
  const int g = pt_frame ? pt_frame[j] : frame_walk(inst_off, F, g0s, j);
  const int64_t pt = pt_frame ? j : cloud_off[frame_cloud[g]] + (j - inst_off[g]);
  const d3 w = world_of(poses + 12 * (int64_t)g, (double)xyz[3 * pt], (double)xyz[3 * pt + 1], (double)xyz[3 * pt + 2]);
  const int64_t kx = key_axis_fast(w.x, prm.vs, prm.inv_vs), ky = key_axis_fast(w.y, prm.vs, prm.inv_vs), kz = key_axis_fast(w.z, prm.vs, prm.inv_vs);
  if (kx < -KEY_BIAS || kx >= KEY_BIAS || ky < -KEY_BIAS || ky >= KEY_BIAS || kz < -KEY_BIAS || kz >= KEY_BIAS) *bad = 1;   // pack_key's rule
  double cx = __dmul_rn(__dadd_rn(0.5, (double)kx), prm.vs), cy = __dmul_rn(__dadd_rn(0.5, (double)ky), prm.vs), cz = __dmul_rn(__dadd_rn(0.5, (double)kz), prm.vs);
  float q = (float)(prm.vs / 4.0);
  uint32_t path = 0;
  for (int l = 0; l < prm.L; l++) {      // the descent of k_gba_point
    const int bx = w.x > cx, by = w.y > cy, bz = w.z > cz;
    path = (path << 3) | (uint32_t)(4 * bx + 2 * by + bz);
    const double dq = (double)q;
    cx = __dadd_rn(cx, bx ? dq : -dq); cy = __dadd_rn(cy, by ? dq : -dq); cz = __dadd_rn(cz, bz ? dq : -dq);
    q = q / 2.0f;
  }
  const uint64_t win = (uint64_t)(g / kl.W), fiw = (uint64_t)(g % kl.W);
  uint64_t k = win;
  k = (k << kl.bx) | (uint64_t)(kx - kl.lo[0]);
  k = (k << kl.by) | (uint64_t)(ky - kl.lo[1]);
  k = (k << kl.bz) | (uint64_t)(kz - kl.lo[2]);
  k = (k << kl.pbits) | path;
  k = (k << kl.fbits) | fiw;
  key[j] = k;
  idx[j] = (int32_t)pt;             // the sort carries the POINT, so nothing downstream looks a frame's cloud up again
}
// multi-sort path: its order array holds instance numbers; -> point numbers (cloud mode)
__global__ void k_inst_to_point(int32_t *__restrict__ sidx, const uint32_t *__restrict__ k1, const int64_t *__restrict__ cloud_off, const int32_t *__restrict__ frame_cloud,
                                const int64_t *__restrict__ inst_off, int64_t n) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int g = (int)(k1[i] & ((1u << FRAME_BITS) - 1u));
  sidx[i] = (int32_t)(cloud_off[frame_cloud[g]] + ((int64_t)sidx[i] - inst_off[g]));
}
// sorted single key -> the (key1, key2) pair the head / segment kernels read: key1 = path << FRAME_BITS | global frame,
// key2 = window | re-biased root cell (only compared for equality)
__global__ void k_gba_unpack64(const uint64_t *__restrict__ key, KeyLayout kl, int64_t NI, uint32_t *__restrict__ k1, uint64_t *__restrict__ k2) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= NI) return;
  const uint64_t k = key[i];
  const uint32_t fiw = (uint32_t)(k & ((1ull << kl.fbits) - 1ull));
  const uint32_t path = (uint32_t)((k >> kl.fbits) & ((1ull << kl.pbits) - 1ull));
  const uint64_t root = k >> (kl.fbits + kl.pbits);
  const uint32_t win = (uint32_t)(root >> (kl.bx + kl.by + kl.bz));
  k1[i] = (path << FRAME_BITS) | (win * (uint32_t)kl.W + fiw);
  k2[i] = root;
}

// 3. head flags, ONE byte per sorted instance: bit l (0..L) = first instance of a layer-l node, bit 7 = first instance of a
// (finest leaf, frame) segment.  Heads nest: a layer-l head is a head of every finer layer and of a segment.
constexpr int HB_SEG = 0x80;
__global__ void k_gba_headbits(const uint32_t *__restrict__ k1, const uint64_t *__restrict__ k2, int W, int L, int64_t n, int64_t n_pad, uint8_t *__restrict__ hb) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_pad) return;
  if (i >= n) { hb[i] = 0; return; }                 // padding up to a whole 16-byte word
  if (i == 0) { hb[0] = (uint8_t)(HB_SEG | ((1u << (L + 1)) - 1u)); return; }
  const uint32_t fm = (1u << FRAME_BITS) - 1u;
  const uint32_t a1 = k1[i], b1 = k1[i - 1];
  const bool root_new = k2[i] != k2[i - 1] || (a1 & fm) / (uint32_t)W != (b1 & fm) / (uint32_t)W;
  const uint32_t pa = a1 >> FRAME_BITS, pb = b1 >> FRAME_BITS;
  uint32_t m = 0;
  for (int l = 0; l <= L; l++) {
    const int sh = 3 * (L - l);
    if (root_new || (pa >> sh) != (pb >> sh)) m |= 1u << l;
  }
  if (((m >> L) & 1u) || (a1 & fm) != (b1 & fm)) m |= HB_SEG;
  hb[i] = (uint8_t)m;
}

// The L + 2 head counters (layers 0..L, then segments) are scanned together, 4096 instances per CTA (16 bytes per thread):
//   k_hb_count   per-CTA totals                      [NCNT][n_blk]
//   k_hb_blocks  exclusive scan over the CTAs, grand totals (the only values the host reads)
//   k_hb_emit    per-CTA exclusive scan + the COMPACT index arrays, written at the heads only:
//       seg_start[s], seg_leaf[s]                                   per (leaf, frame) segment
//       nd.start / first_seg / first_leaf / first_child / parent    per node of every layer
// (the round-2 form wrote five int32 flag arrays + five inclusive scans over ALL instances and looked nodes up through them:
//  100 B of traffic per instance against 2 B + 8 B per segment + 20 B per node here).
constexpr int NCNT = MAXL + 2;
constexpr int HB_ITEMS = 16, HB_THREADS = 256;      // 4096 instances per CTA
struct NodeArrays {
  int32_t *start[MAXL + 1], *first_seg[MAXL + 1], *first_leaf[MAXL + 1], *first_child[MAXL + 1], *parent[MAXL + 1];
};
__device__ __forceinline__ int hb_popc(const uint4 &w, int bit) {
  const uint32_t m = 0x01010101u;
  return __popc((w.x >> bit) & m) + __popc((w.y >> bit) & m) + __popc((w.z >> bit) & m) + __popc((w.w >> bit) & m);
}
template <int L>
__global__ void __launch_bounds__(HB_THREADS) k_hb_count(const uint4 *__restrict__ hb16, int64_t n16, int64_t n_blk, int32_t *__restrict__ blk_cnt) {
  __shared__ int32_t sh[NCNT][HB_THREADS / 32];
  const int64_t q = blockIdx.x * (int64_t)HB_THREADS + threadIdx.x;
  uint4 w = make_uint4(0, 0, 0, 0);
  if (q < n16) w = hb16[q];
#pragma unroll
  for (int c = 0; c < NCNT; c++) {
    int v = (c <= L) ? hb_popc(w, c) : (c == L + 1 ? hb_popc(w, 7) : 0);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[c][threadIdx.x >> 5] = v;
  }
  __syncthreads();
  if (threadIdx.x < NCNT) {
    int t = 0;
#pragma unroll
    for (int k = 0; k < HB_THREADS / 32; k++) t += sh[threadIdx.x][k];
    blk_cnt[(size_t)threadIdx.x * n_blk + blockIdx.x] = t;
  }
}
// one CTA per counter: exclusive scan of blk_cnt[c][0..n_blk) in place, total -> tot[c]
__global__ void __launch_bounds__(1024) k_hb_blocks(int32_t *__restrict__ blk_cnt, int64_t n_blk, int32_t *__restrict__ tot) {
  __shared__ int32_t sh[1024];
  int32_t *a = blk_cnt + (size_t)blockIdx.x * n_blk;
  const int64_t per = (n_blk + 1023) / 1024;
  const int64_t b = threadIdx.x * per, e = min(b + per, n_blk);
  int32_t t = 0;
  for (int64_t i = b; i < e; i++) t += a[i];
  sh[threadIdx.x] = t;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {
    const int32_t v = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
    __syncthreads();
    sh[threadIdx.x] += v;
    __syncthreads();
  }
  int32_t run = sh[threadIdx.x] - t;
  for (int64_t i = b; i < e; i++) { const int32_t v = a[i]; a[i] = run; run += v; }
  if (threadIdx.x == 1023) tot[blockIdx.x] = sh[1023];
}
template <int L>
__global__ void __launch_bounds__(HB_THREADS) k_hb_emit(const uint4 *__restrict__ hb16, int64_t n16, int64_t n_blk, const int32_t *__restrict__ blk_base,
                                                       int32_t *__restrict__ seg_start, int32_t *__restrict__ seg_leaf, NodeArrays nd) {
  __shared__ int32_t sh[NCNT][HB_THREADS / 32];
  const int64_t q = blockIdx.x * (int64_t)HB_THREADS + threadIdx.x;
  uint4 w = make_uint4(0, 0, 0, 0);
  if (q < n16) w = hb16[q];
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  int32_t cnt[NCNT];        // -> number of heads of counter c before this thread's first instance
#pragma unroll
  for (int c = 0; c < NCNT; c++) {
    const int mine = (c <= L) ? hb_popc(w, c) : (c == L + 1 ? hb_popc(w, 7) : 0);
    int v = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += u; }
    if (lane == 31) sh[c][wp] = v;
    cnt[c] = v - mine;
  }
  __syncthreads();
#pragma unroll
  for (int c = 0; c < NCNT; c++) {
    int base = (c <= L + 1) ? blk_base[(size_t)c * n_blk + blockIdx.x] : 0;
    for (int k = 0; k < wp; k++) base += sh[c][k];
    cnt[c] += base;
  }
  if (q >= n16) return;
  const uint32_t words[4] = {w.x, w.y, w.z, w.w};
  const int64_t i0 = q * HB_ITEMS;
#pragma unroll
  for (int b = 0; b < HB_ITEMS; b++) {
    const uint32_t m = (words[b >> 2] >> (8 * (b & 3))) & 0xffu;
    if (!m) continue;
    const int32_t i = (int32_t)(i0 + b);
    const int32_t s = cnt[L + 1];            // a node head is a segment head: this instance opens segment s
#pragma unroll
    for (int l = 0; l <= L; l++) {
      if (!((m >> l) & 1u)) continue;
      const int32_t k = cnt[l];
      nd.start[l][k] = i;
      nd.first_seg[l][k] = s;
      nd.first_leaf[l][k] = cnt[L];          // heads nest, so the leaf that opens here has not been counted yet
      if (l < L) nd.first_child[l][k] = cnt[l + 1];
      if (l > 0) nd.parent[l][k] = ((m >> (l - 1)) & 1u) ? cnt[l - 1] : cnt[l - 1] - 1;   // opens here too, or was opened earlier
    }
#pragma unroll
    for (int l = 0; l <= L; l++) if ((m >> l) & 1u) cnt[l]++;
    seg_start[s] = i;
    seg_leaf[s] = cnt[L] - 1;
    cnt[L + 1]++;
  }
}
// the closing entries of the compact arrays
__global__ void k_hb_close(int L, int32_t n, int32_t nseg, NodeArrays nd, int32_t nn0, int32_t nn1, int32_t nn2, int32_t nn3, int32_t *__restrict__ seg_start) {
  const int32_t nn[4] = {nn0, nn1, nn2, nn3};
  if (threadIdx.x == 0) seg_start[nseg] = n;
  const int l = threadIdx.x;
  if (l <= L) {
    nd.start[l][nn[l]] = n;
    nd.first_seg[l][nn[l]] = nseg;
    nd.first_leaf[l][nn[l]] = nn[L];
    if (l < L) nd.first_child[l][nn[l]] = nn[l + 1];
  }
}
// start[id] = position of the head with inclusive-scan value id+1; start[total] = n
__global__ void k_starts(const int32_t *__restrict__ flag, const int32_t *__restrict__ incl, int64_t n, int32_t *__restrict__ start) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (flag[i]) start[incl[i] - 1] = (int32_t)i;
  if (i == n - 1) start[incl[i]] = (int32_t)n;
}

// 4 + 5a. node statistics.  nodeS: [10][nn] world P6 v3 N ; nodeE: [12][nn] lam3 U9 ; fspan: [2][nn] min / max frame ;
// flags bit0 big(N>10) bit1 plane bit2 accept
__device__ __forceinline__ void node_decide(const double a[9], int64_t N, int fmin, int fmax, const GbaP &prm, int layer, int64_t nn, int64_t k,
                                            double *__restrict__ nodeS, double *__restrict__ nodeE, int32_t *__restrict__ fspan, int32_t *__restrict__ flags) {
  fspan[k] = fmin; fspan[nn + k] = fmax;
#pragma unroll
  for (int c = 0; c < 9; c++) nodeS[(size_t)c * nn + k] = a[c];
  nodeS[(size_t)9 * nn + k] = (double)N;
  int fl = 0;
  double l[3] = {0, 0, 0}, u[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  if (N > 10) {   // loop_refine.hpp:360
    fl |= 1;
    const double dn = (double)N;
    const double cx = a[6] / dn, cy = a[7] / dn, cz = a[8] / dn;   // PointCluster::cov, tools.hpp:333-337
    s33 C;
    C.xx = a[0] / dn - cx * cx; C.xy = a[1] / dn - cx * cy; C.xz = a[2] / dn - cx * cz;
    C.yy = a[3] / dn - cy * cy; C.yz = a[4] / dn - cy * cz; C.zz = a[5] / dn - cz * cz;
    eig3_jacobi(C, l, u);
    const bool plane = l[0] < prm.min_eig && (l[0] / l[2]) < prm.thr[layer];   // plane_judge, loop_refine.hpp:310-314
    if (plane) {
      fl |= 2;
      // accept: >= 2 occupied frames (:372-376) and lambda0/lambda1 <= 0.12 (:378)
      if (fmin != fmax && !(l[0] / l[1] > 0.12)) fl |= 4;
    }
  }
#pragma unroll
  for (int c = 0; c < 3; c++) nodeE[(size_t)c * nn + k] = l[c];
#pragma unroll
  for (int c = 0; c < 9; c++) nodeE[(size_t)(3 + c) * nn + k] = u[c];
  flags[k] = fl;
}

// The finest layer, straight from the sorted points: GL lanes per leaf node, one (leaf, frame) segment per lane at a time.  Each
// frame's world cluster is summed on its own, in scan order, and the clusters are then added to the node's in frame order
// (pcr_add[frame], then their sum: loop_refine.hpp:320, :355-358) — for ANY GL, so the lane count is a pure scheduling choice:
// 1 where a leaf holds a point or two (bottom level: 0.0625 m leaves of single scans), 8 where it holds dozens (merged submaps).
// The round-2 form parked 152 B per segment in HBM between two kernels (6.9 GB written + 3.3 GB read back at the bottom level of
// the 2,000-keyframe job: profiles/r02d_top_kernels_ncu.txt).
template <int GL, bool DECIDE>
__global__ void __launch_bounds__(128)
k_leaf_stats(const float *__restrict__ xyz, const double *__restrict__ poses, const int32_t *__restrict__ spt, const uint32_t *__restrict__ k1,
             const int32_t *__restrict__ first_seg, const int32_t *__restrict__ seg_start, int64_t nn, GbaP prm, int layer,
             double *__restrict__ nodeS, double *__restrict__ nodeE, int32_t *__restrict__ fspan, int32_t *__restrict__ flags) {
  const int64_t k = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / GL;
  const int lane = threadIdx.x & (GL - 1);
  const unsigned gmask = GL == 32 ? 0xffffffffu : (((1u << GL) - 1u) << ((threadIdx.x & 31) & ~(GL - 1)));
  const bool on = k < nn;
  const int32_t s0 = on ? first_seg[k] : 0, s1 = on ? first_seg[k + 1] : 0;
  double a[9];
#pragma unroll
  for (int c = 0; c < 9; c++) a[c] = 0.0;
  int fmin = 0x7fffffff, fmax = -1;
  int32_t N = 0;
  for (int32_t sb = s0; sb < s1; sb += GL) {          // uniform within the group
    const int32_t s = sb + lane;
    double f[9];
#pragma unroll
    for (int c = 0; c < 9; c++) f[c] = 0.0;
    int g = 0; int32_t np = 0;
    if (s < s1) {
      const int32_t p0 = seg_start[s], p1 = seg_start[s + 1];
      g = (int)(k1[p0] & ((1u << FRAME_BITS) - 1u));
      const double *ps = poses + 12 * (int64_t)g;
      for (int32_t j = p0; j < p1; j++) {
        const int64_t pt = spt[j];
        const d3 w = world_of(ps, (double)xyz[3 * pt], (double)xyz[3 * pt + 1], (double)xyz[3 * pt + 2]);
        f[0] += w.x * w.x; f[1] += w.x * w.y; f[2] += w.x * w.z; f[3] += w.y * w.y; f[4] += w.y * w.z; f[5] += w.z * w.z;
        f[6] += w.x; f[7] += w.y; f[8] += w.z;
      }
      np = p1 - p0;
    }
#pragma unroll
    for (int q = 0; q < GL; q++) {
      if (sb + q < s1) {                              // uniform within the group
#pragma unroll
        for (int c = 0; c < 9; c++) a[c] += GL == 1 ? f[c] : __shfl_sync(gmask, f[c], q, GL);
        const int gq = GL == 1 ? g : __shfl_sync(gmask, g, q, GL);
        fmin = min(fmin, gq); fmax = max(fmax, gq);
        N += GL == 1 ? np : __shfl_sync(gmask, np, q, GL);
      }
    }
  }
  if (!on || lane != 0) return;
  if (DECIDE) { node_decide(a, (int64_t)N, fmin, fmax, prm, layer, nn, k, nodeS, nodeE, fspan, flags); return; }
  fspan[k] = fmin; fspan[nn + k] = fmax;
#pragma unroll
  for (int c = 0; c < 9; c++) nodeS[(size_t)c * nn + k] = a[c];
  nodeS[(size_t)9 * nn + k] = (double)N;
}
// the decisions of the leaves whose sums a multi-lane k_leaf_stats left in nodeS / fspan (the eigen-solve at one lane in GL would idle the rest)
__global__ void __launch_bounds__(128) k_leaf_decide(int64_t nn, GbaP prm, int layer, double *__restrict__ nodeS, double *__restrict__ nodeE, int32_t *__restrict__ fspan,
                                                    int32_t *__restrict__ flags) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= nn) return;
  double a[9];
#pragma unroll
  for (int c = 0; c < 9; c++) a[c] = nodeS[(size_t)c * nn + k];
  node_decide(a, (int64_t)nodeS[(size_t)9 * nn + k], fspan[k], fspan[nn + k], prm, layer, nn, k, nodeS, nodeE, fspan, flags);
}

// A coarser node = the sum of its (up to 8) children of the next layer, in order: every point is read once instead of once per layer.
__global__ void k_node_stats(const int32_t *__restrict__ first_child, int64_t nn, GbaP prm, int layer, int64_t nn_child, const double *__restrict__ childS,
                             const int32_t *__restrict__ child_fspan, double *__restrict__ nodeS, double *__restrict__ nodeE, int32_t *__restrict__ fspan,
                             int32_t *__restrict__ flags) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= nn) return;
  double a[9];
#pragma unroll
  for (int c = 0; c < 9; c++) a[c] = 0.0;
  int64_t N = 0;
  int fmin = 0x7fffffff, fmax = -1;
  const int32_t c0 = first_child[k], c1 = first_child[k + 1];
  for (int32_t ch = c0; ch < c1; ch++) {
#pragma unroll
    for (int c = 0; c < 9; c++) a[c] += childS[(size_t)c * nn_child + ch];
    N += (int64_t)childS[(size_t)9 * nn_child + ch];
    fmin = min(fmin, child_fspan[ch]); fmax = max(fmax, child_fspan[nn_child + ch]);
  }
  node_decide(a, N, fmin, fmax, prm, layer, nn, k, nodeS, nodeE, fspan, flags);
}

// 5b. top-down liveness: a node exists iff every ancestor was subdivided (N>10, not a plane, layer < max_layer)
// flags bit3 alive, bit4 emit.  emit_at is indexed by the node's first LEAF (every node opens a leaf; emitted nodes are disjoint).
__global__ void k_node_alive(const int32_t *__restrict__ first_leaf, int64_t nn, int layer, const int32_t *__restrict__ parent,
                             const int32_t *__restrict__ parent_flags, int32_t *__restrict__ flags, int32_t *__restrict__ emit_at) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= nn) return;
  int fl = flags[k];
  bool alive = true;
  if (layer > 0) {
    const int pf = parent_flags[parent[k]];
    alive = (pf & 8) && (pf & 1) && !(pf & 2);   // parent alive and subdivided (parent layer < L by construction)
  }
  if (alive) fl |= 8;
  if (alive && (fl & 4)) { fl |= 16; emit_at[first_leaf[k]] = 1; }
  flags[k] = fl;
}

// 6a. voxel records of the emitted nodes of one layer
__global__ void k_emit_voxels(const int32_t *__restrict__ first_leaf, const int32_t *__restrict__ flags, int64_t nn, const int32_t *__restrict__ vid_excl,
                              const double *__restrict__ nodeS, const double *__restrict__ nodeE, int64_t V, double *__restrict__ lam,
                              double *__restrict__ U, double *__restrict__ mrg, double *__restrict__ fix, double *__restrict__ coe) {
  const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= nn || !(flags[k] & 16)) return;
  const int64_t v = vid_excl[first_leaf[k]];
#pragma unroll
  for (int c = 0; c < 10; c++) { mrg[(size_t)c * V + v] = nodeS[(size_t)c * nn + k]; fix[(size_t)c * V + v] = 0.0; }
#pragma unroll
  for (int c = 0; c < 3; c++) lam[(size_t)c * V + v] = nodeE[(size_t)c * nn + k];
#pragma unroll
  for (int c = 0; c < 9; c++) U[(size_t)c * V + v] = nodeE[(size_t)(3 + c) * nn + k];
  coe[v] = 1.0;   // loop_refine.hpp:380
}

struct LevelPtrs {
  const int32_t *first_leaf[MAXL + 1];
  const int32_t *parent[MAXL + 1];
  const int32_t *flags[MAXL + 1];
};
// 6b. segment -> key (voxel << fb | frame) of the emitted ancestor (walk up from the segment's leaf); segments under no emitted
// node get voxel = V (they sort last).  The key uses exactly bits_for(V + 1) + fb bits.
__global__ void k_seg_voxel(const int32_t *__restrict__ seg_start, const int32_t *__restrict__ seg_leaf, const uint32_t *__restrict__ k1, int64_t nseg, LevelPtrs lp,
                            int L, const int32_t *__restrict__ vid_excl, int64_t V, int fb, uint64_t *__restrict__ key, int32_t *__restrict__ val,
                            int32_t *__restrict__ n_valid) {
  const int64_t s = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool valid = false;
  if (s < nseg) {
    uint64_t kk = (uint64_t)V << fb;
    int32_t k = seg_leaf[s];
    for (int l = L; l >= 0; l--) {
      if (lp.flags[l][k] & 16) {
        kk = ((uint64_t)(uint32_t)vid_excl[lp.first_leaf[l][k]] << fb) | (k1[seg_start[s]] & ((1u << FRAME_BITS) - 1u));
        valid = true;
        break;
      }
      if (l > 0) k = lp.parent[l][k];
    }
    key[s] = kk;
    val[s] = (int32_t)s;
  }
  const int c = __syncthreads_count(valid);                  // segments of emitted voxels: they sort before the rest
  if (threadIdx.x == 0 && c) atomicAdd(n_valid, c);
}
__global__ void k_vf_heads(const uint64_t *__restrict__ key, int64_t n, int64_t V, int fb, int32_t *__restrict__ flag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  flag[i] = ((int64_t)(key[i] >> fb) != V && (i == 0 || key[i] != key[i - 1])) ? 1 : 0;
}
// 6c. GL lanes per voxel-frame: the scan-local cluster (what LidarFactor stores per voxel-frame, loop_refine.hpp:382-385) of each of
// its segments, summed from the points in scan order (one segment per lane at a time), then the segments in order — for any GL
template <int GL>
__global__ void __launch_bounds__(128)
k_vf_fill(const uint64_t *__restrict__ key, const int32_t *__restrict__ sval, const int32_t *__restrict__ vf_start, int64_t nnz, int64_t nvalid,
          const int32_t *__restrict__ seg_start, const int32_t *__restrict__ spt, const float *__restrict__ xyz, int fb, double *__restrict__ mom,
          int32_t *__restrict__ cnt, int32_t *__restrict__ frame, int32_t *__restrict__ vox) {
  const int64_t e = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / GL;
  const int lane = threadIdx.x & (GL - 1);
  const unsigned gmask = GL == 32 ? 0xffffffffu : (((1u << GL) - 1u) << ((threadIdx.x & 31) & ~(GL - 1)));
  const bool on = e < nnz;
  const int32_t b = on ? vf_start[e] : 0, en = on ? ((e + 1 < nnz) ? vf_start[e + 1] : (int32_t)nvalid) : 0;
  double a[9];
#pragma unroll
  for (int c = 0; c < 9; c++) a[c] = 0.0;
  int32_t N = 0;
  for (int32_t ib = b; ib < en; ib += GL) {
    const int32_t i = ib + lane;
    double f[9];
#pragma unroll
    for (int c = 0; c < 9; c++) f[c] = 0.0;
    int32_t np = 0;
    if (i < en) {
      const int32_t s = sval[i];
      const int32_t p0 = seg_start[s], p1 = seg_start[s + 1];
      for (int32_t j = p0; j < p1; j++) {
        const int64_t pt = spt[j];
        const double lx = (double)xyz[3 * pt], ly = (double)xyz[3 * pt + 1], lz = (double)xyz[3 * pt + 2];
        f[0] += lx * lx; f[1] += lx * ly; f[2] += lx * lz; f[3] += ly * ly; f[4] += ly * lz; f[5] += lz * lz;
        f[6] += lx; f[7] += ly; f[8] += lz;
      }
      np = p1 - p0;
    }
#pragma unroll
    for (int q = 0; q < GL; q++) {
      if (ib + q < en) {
#pragma unroll
        for (int c = 0; c < 9; c++) a[c] += GL == 1 ? f[c] : __shfl_sync(gmask, f[c], q, GL);
        N += GL == 1 ? np : __shfl_sync(gmask, np, q, GL);
      }
    }
  }
  if (!on || lane != 0) return;
#pragma unroll
  for (int c = 0; c < 9; c++) mom[(size_t)c * nnz + e] = a[c];
  cnt[e] = N;
  frame[e] = (int32_t)(key[b] & ((1ull << fb) - 1ull));
  vox[e] = (int32_t)(key[b] >> fb);
}
// off[v] = first voxel-frame of voxel v (vox[] ascending), off[V] = nnz ; win[v], wv_off
__global__ void k_vox_off(const int32_t *__restrict__ vox, int64_t nnz, int64_t V, int32_t *__restrict__ off) {
  const int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (v > V) return;
  int64_t lo = 0, hi = nnz;
  while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (vox[mid] < v) lo = mid + 1; else hi = mid; }
  off[v] = (int32_t)lo;
}
__global__ void k_vox_win(const int32_t *__restrict__ off, const int32_t *__restrict__ frame, int64_t V, int W, int32_t *__restrict__ win) {
  const int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (v < V) win[v] = frame[off[v]] / W;
}
__global__ void k_win_off(const int32_t *__restrict__ win, int64_t V, int NW, int32_t *__restrict__ wv_off) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w > NW) return;
  int64_t lo = 0, hi = V;
  while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (win[mid] < w) lo = mid + 1; else hi = mid; }
  wv_off[w] = (int32_t)lo;
}

}  // namespace

namespace {

struct Tmp {   // RAII set of temporary device buffers
  std::vector<DevBuf *> all;
  ~Tmp() { for (DevBuf *b : all) { b->release(); delete b; } }
  DevBuf &make() { all.push_back(new DevBuf()); return *all.back(); }
};

int incl_scan(vx_ctx *ctx, const int32_t *in, int32_t *out, int64_t n) {
  size_t tb = 0;
  VX_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tb, in, out, (int)n, ctx->stream));
  VX_TRY(ctx->scratch.reserve(tb));
  VX_CUDA(cub::DeviceScan::InclusiveSum(ctx->scratch.p, tb, in, out, (int)n, ctx->stream));
  ctx->launches += 2;
  return VX_OK;
}
int excl_scan(vx_ctx *ctx, const int32_t *in, int32_t *out, int64_t n) {
  size_t tb = 0;
  VX_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, in, out, (int)n, ctx->stream));
  VX_TRY(ctx->scratch.reserve(tb));
  VX_CUDA(cub::DeviceScan::ExclusiveSum(ctx->scratch.p, tb, in, out, (int)n, ctx->stream));
  ctx->launches += 2;
  return VX_OK;
}
template <class K>
int sort_pairs(vx_ctx *ctx, const K *kin, K *kout, const int32_t *vin, int32_t *vout, int64_t n, int begin_bit, int end_bit) {
  size_t tb = 0;
  VX_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, kin, kout, vin, vout, (int)n, begin_bit, end_bit, ctx->stream));
  VX_TRY(ctx->scratch.reserve(tb));
  VX_CUDA(cub::DeviceRadixSort::SortPairs(ctx->scratch.p, tb, kin, kout, vin, vout, (int)n, begin_bit, end_bit, ctx->stream));
  ctx->launches += 4;
  return VX_OK;
}
int read_last(vx_ctx *ctx, const int32_t *d, int64_t n, int32_t *h) {
  VX_CUDA(cudaMemcpyAsync(h, d + (n - 1), sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  return VX_OK;
}
int bits_for(int64_t n) { int b = 1; while (((int64_t)1 << b) < n) b++; return b; }

}  // namespace

namespace {
struct View { const void *p; template <class X> const X *as() const { return (const X *)p; } };
}
namespace vx {
// bbox[c] = {min x, min y, min z, max x, max y, max z} of cloud c (scan-local floats); an empty cloud gets min > max
int cloud_bboxes(vx_ctx *ctx, const float *d_xyz, const int64_t *d_coff, int n_clouds, float *d_bbox) {
  VX_KERNEL(ctx, "k_cloud_bbox", k_cloud_bbox<<<n_clouds, 256, 0, ctx->stream>>>(d_xyz, d_coff, d_bbox));
  return VX_OK;
}
// d_xyz / d_coff: the clouds already on the device (uploaded once per HBA_add_edge call); cloud_off: host copy.
// explicit mode (d_pt_frame != nullptr): d_xyz_in holds n_pts routed points, point j belongs to frame d_pt_frame[j]; points of
// one frame must appear in their original cloud order (cloud_off / frame_cloud are then unused).
int gba_build_impl(vx_ctx *ctx, const float *d_xyz_in, const int64_t *d_coff_in, const int64_t *cloud_off, int n_clouds, const int32_t *frame_cloud,
                   const vx_pose *poses, const vx_gba_params *params, vx_factor *f, const int32_t *d_pt_frame, int64_t n_pts) {
  const bool explicit_pts = d_pt_frame != nullptr;
  VX_ARG(f != nullptr && params != nullptr && poses != nullptr && (explicit_pts || (cloud_off != nullptr && n_clouds >= 1)));
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  const int W = f->W, NW = f->n_win, F = W * NW;
  VX_ARG(params->max_layer >= 0 && params->max_layer <= MAXL && params->voxel_size > 0);
  VX_ARG(F < (1 << FRAME_BITS));
  VX_ARG(explicit_pts || frame_cloud != nullptr || n_clouds == F);
  GbaP prm;
  prm.vs = params->voxel_size; prm.inv_vs = 1.0 / params->voxel_size; prm.min_eig = params->min_eigen_value; prm.L = params->max_layer;
  for (int k = 0; k < 4; k++) prm.thr[k] = params->eigen_value_array[k];
  const int L = prm.L;
  const int64_t NP = explicit_pts ? n_pts : cloud_off[n_clouds];
  std::vector<int32_t> fc(F, 0);
  std::vector<int64_t> inst_off(F + 1, 0);
  if (!explicit_pts)
    for (int g = 0; g < F; g++) {
      fc[g] = frame_cloud ? frame_cloud[g] : g;
      VX_ARG(fc[g] >= 0 && fc[g] < n_clouds);
      inst_off[g + 1] = inst_off[g] + (cloud_off[fc[g] + 1] - cloud_off[fc[g]]);
    }
  const int64_t NI = explicit_pts ? n_pts : inst_off[F];
  VX_ARG(NI < ((int64_t)1 << 31) && NP >= 0 && NP < ((int64_t)1 << 31));
  if (NI == 0) return vx_factor_clear(f);
  VX_ARG(d_xyz_in != nullptr && (explicit_pts || d_coff_in != nullptr));

  Tmp T;
  View d_xyz{d_xyz_in}, d_coff{d_coff_in};
  DevBuf &d_fc = T.make(), &d_ioff = T.make(), &d_poses = T.make(), &d_bad = T.make();
  VX_TRY(d_fc.reserve(4 * F));
  VX_TRY(d_ioff.reserve(8 * (F + 1))); VX_TRY(d_poses.reserve(sizeof(vx_pose) * F)); VX_TRY(d_bad.reserve(4));
  VX_CUDA(cudaMemcpyAsync(d_fc.p, fc.data(), 4 * F, cudaMemcpyHostToDevice, st));
  VX_CUDA(cudaMemcpyAsync(d_ioff.p, inst_off.data(), 8 * (F + 1), cudaMemcpyHostToDevice, st));
  VX_CUDA(cudaMemcpyAsync(d_poses.p, poses, sizeof(vx_pose) * F, cudaMemcpyHostToDevice, st));
  VX_CUDA(cudaMemsetAsync(d_bad.p, 0, 4, st));

  PhaseTimer pt("gba_build", st);
  // ---- 1 + 2. keys and order.  Single-key path when (window | re-biased root cell | path | frame in window) fits 64 bits.
  DevBuf &k1t = T.make(), &k2u = T.make(), &ia = T.make(), &ib = T.make();
  VX_TRY(k1t.reserve(4 * NI)); VX_TRY(k2u.reserve(8 * NI)); VX_TRY(ia.reserve(4 * NI)); VX_TRY(ib.reserve(4 * NI));
  int32_t *sidx = nullptr;
  uint32_t *k1s = k1t.as<uint32_t>();
  uint64_t *k2s = k2u.as<uint64_t>();
  bool single_key = getenv("VX_GBA_MULTISORT") == nullptr;
  KeyLayout kl;
  if (single_key) {
    DevBuf &d_rng = T.make();
    VX_TRY(d_rng.reserve(sizeof(CellRange)));
    CellRange r0;
    for (int a = 0; a < 3; a++) { r0.lo[a] = 0x7fffffff; r0.hi[a] = -0x7fffffff; }
    VX_CUDA(cudaMemcpyAsync(d_rng.p, &r0, sizeof(r0), cudaMemcpyHostToDevice, st));
    int bad = 0;
    if (explicit_pts) {
      const int grid = (int)std::min<int64_t>(blocks_for(NI, 256), (int64_t)ctx->sm_count * 8);
      VX_KERNEL(ctx, "k_gba_cell_range", k_gba_cell_range<<<grid, 256, 0, st>>>(d_xyz.as<float>(), d_coff.as<int64_t>(), d_fc.as<int32_t>(), d_ioff.as<int64_t>(), d_pt_frame, F,
                                                                               d_poses.as<double>(), prm.vs, NI, d_rng.as<CellRange>(), d_bad.as<int>()));
      VX_CUDA(cudaMemcpyAsync(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost, st));
    } else {
      DevBuf &d_bbox = T.make();
      VX_TRY(d_bbox.reserve(sizeof(float) * 6 * (size_t)n_clouds));
      VX_TRY(cloud_bboxes(ctx, d_xyz.as<float>(), d_coff.as<int64_t>(), n_clouds, d_bbox.as<float>()));
      VX_KERNEL(ctx, "k_frame_cells", k_frame_cells<<<blocks_for(F, 128), 128, 0, st>>>(d_bbox.as<float>(), d_fc.as<int32_t>(), F, d_poses.as<double>(), prm.vs,
                                                                                   d_rng.as<CellRange>()));
    }
    VX_CUDA(cudaMemcpyAsync(&r0, d_rng.p, sizeof(r0), cudaMemcpyDeviceToHost, st));
    VX_CUDA(cudaStreamSynchronize(st));
    if (bad) { set_error("OctreeGBA::cut_voxel: a root cell index does not fit the 21-bit packed key"); return VX_ERR_ARG; }
    if (r0.lo[0] > r0.hi[0]) { for (int a = 0; a < 3; a++) { r0.lo[a] = 0; r0.hi[a] = 0; } }     // no finite point at all
    for (int a = 0; a < 3; a++) kl.lo[a] = r0.lo[a];
    kl.bx = bits_for((int64_t)r0.hi[0] - r0.lo[0] + 1); kl.by = bits_for((int64_t)r0.hi[1] - r0.lo[1] + 1); kl.bz = bits_for((int64_t)r0.hi[2] - r0.lo[2] + 1);
    kl.pbits = std::max(1, 3 * L); kl.fbits = bits_for(W); kl.W = W;
    const int total = bits_for(NW) + kl.bx + kl.by + kl.bz + kl.pbits + kl.fbits;
    if (total > 64) single_key = false;
    else {
      DevBuf &ka = T.make(), &kb = T.make();
      VX_TRY(ka.reserve(8 * NI)); VX_TRY(kb.reserve(8 * NI));
      VX_KERNEL(ctx, "k_gba_key64", k_gba_key64<<<blocks_for(NI, 256), 256, 0, st>>>(d_xyz.as<float>(), d_coff.as<int64_t>(), d_fc.as<int32_t>(), d_ioff.as<int64_t>(), d_pt_frame, F,
                                                                                  d_poses.as<double>(), prm, kl, NI, ka.as<uint64_t>(), ia.as<int32_t>(), d_bad.as<int>()));
      // Cloud mode: instances are generated in (window, frame, scan) order, and the radix sort is stable — so among equal
      // (window, root cell, path) the frames are ALREADY ascending and the frame field (the low fbits) need not be sorted:
      // 39 bits = 5 passes instead of 43 = 6 at the bottom level, 30 = 4 instead of 39 = 5 at the top level (W = 400: 9 frame bits).
      // Routed points (explicit mode) promise scan order within a frame only, so they sort every bit.
      const int begin_bit = explicit_pts ? 0 : kl.fbits;
      VX_TRY(sort_pairs<uint64_t>(ctx, ka.as<uint64_t>(), kb.as<uint64_t>(), ia.as<int32_t>(), ib.as<int32_t>(), NI, begin_bit, total));
      VX_KERNEL(ctx, "k_gba_unpack64", k_gba_unpack64<<<blocks_for(NI, 256), 256, 0, st>>>(kb.as<uint64_t>(), kl, NI, k1s, k2s));
      sidx = ib.as<int32_t>();
      VX_CUDA(cudaMemcpyAsync(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaStreamSynchronize(st));
      ka.release(); kb.release();
      if (bad) { set_error("OctreeGBA::cut_voxel: a root cell index does not fit the 21-bit packed key"); return VX_ERR_ARG; }
    }
  }
  if (!single_key) {
    // three stable sorts: (path, frame), then root key, then window (any scene extent)
    DevBuf &k1i = T.make(), &k2i = T.make(), &k2t = T.make();
    VX_TRY(k1i.reserve(4 * NI)); VX_TRY(k2i.reserve(8 * NI)); VX_TRY(k2t.reserve(8 * NI));
    VX_CUDA(cudaMemsetAsync(d_bad.p, 0, 4, st));
    VX_KERNEL(ctx, "k_gba_point", k_gba_point<<<blocks_for(NI, 256), 256, 0, st>>>(d_xyz.as<float>(), d_coff.as<int64_t>(), d_fc.as<int32_t>(), d_ioff.as<int64_t>(), d_pt_frame, F,
                                                                                  d_poses.as<double>(), prm, NI, k1i.as<uint32_t>(), k2i.as<uint64_t>(), ia.as<int32_t>(), d_bad.as<int>()));
    VX_TRY(sort_pairs<uint32_t>(ctx, k1i.as<uint32_t>(), k1t.as<uint32_t>(), ia.as<int32_t>(), ib.as<int32_t>(), NI, 0, FRAME_BITS + 3 * L));
    VX_KERNEL(ctx, "k_gather", k_gather<uint64_t><<<blocks_for(NI, 256), 256, 0, st>>>(k2i.as<uint64_t>(), ib.as<int32_t>(), k2t.as<uint64_t>(), NI));
    VX_TRY(sort_pairs<uint64_t>(ctx, k2t.as<uint64_t>(), k2u.as<uint64_t>(), ib.as<int32_t>(), ia.as<int32_t>(), NI, 0, 63));
    sidx = ia.as<int32_t>();
    if (NW > 1) {
      VX_KERNEL(ctx, "k_gather_win", k_gather_win<<<blocks_for(NI, 256), 256, 0, st>>>(k1i.as<uint32_t>(), ia.as<int32_t>(), W, k1t.as<uint32_t>(), NI));
      VX_TRY(sort_pairs<uint32_t>(ctx, k1t.as<uint32_t>(), reinterpret_cast<uint32_t *>(k2t.p), ia.as<int32_t>(), ib.as<int32_t>(), NI, 0, bits_for(NW)));
      sidx = ib.as<int32_t>();
    }
    VX_KERNEL(ctx, "k_gather", k_gather<uint32_t><<<blocks_for(NI, 256), 256, 0, st>>>(k1i.as<uint32_t>(), sidx, k1s, NI));
    VX_KERNEL(ctx, "k_gather", k_gather<uint64_t><<<blocks_for(NI, 256), 256, 0, st>>>(k2i.as<uint64_t>(), sidx, k2s, NI));
    {
      int bad = 0;
      VX_CUDA(cudaMemcpyAsync(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaStreamSynchronize(st));
      if (bad) { set_error("OctreeGBA::cut_voxel: a root cell index does not fit the 21-bit packed key"); return VX_ERR_ARG; }
    }
    k1i.release(); k2i.release(); k2t.release();
    if (!explicit_pts)
      VX_KERNEL(ctx, "k_inst_to_point", k_inst_to_point<<<blocks_for(NI, 256), 256, 0, st>>>(sidx, k1s, d_coff.as<int64_t>(), d_fc.as<int32_t>(), d_ioff.as<int64_t>(), NI));
  }

  pt.lap("keys+sort");
  // ---- 3. head bits, one fused scan of the L + 2 counters, compact index arrays
  const int64_t n16 = (NI + HB_ITEMS - 1) / HB_ITEMS, n_pad = n16 * HB_ITEMS, n_blk = (n16 + HB_THREADS - 1) / HB_THREADS;
  DevBuf &hb = T.make(), &blk = T.make(), &d_tot = T.make();
  VX_TRY(hb.reserve((size_t)n_pad)); VX_TRY(blk.reserve(sizeof(int32_t) * NCNT * (size_t)n_blk)); VX_TRY(d_tot.reserve(sizeof(int32_t) * NCNT));
  VX_KERNEL(ctx, "k_gba_headbits", k_gba_headbits<<<blocks_for(n_pad, 256), 256, 0, st>>>(k1s, k2s, W, L, NI, n_pad, hb.as<uint8_t>()));
#define VX_HB_L(CALL) do { if (L == 0) { CALL(0); } else if (L == 1) { CALL(1); } else if (L == 2) { CALL(2); } else { CALL(3); } } while (0)
#define VX_HB_COUNT(LL) VX_KERNEL(ctx, "k_hb_count", k_hb_count<LL><<<(unsigned)n_blk, HB_THREADS, 0, st>>>(hb.as<uint4>(), n16, n_blk, blk.as<int32_t>()))
  VX_HB_L(VX_HB_COUNT);
#undef VX_HB_COUNT
  VX_KERNEL(ctx, "k_hb_blocks", k_hb_blocks<<<L + 2, 1024, 0, st>>>(blk.as<int32_t>(), n_blk, d_tot.as<int32_t>()));
  int32_t tot[NCNT] = {0, 0, 0, 0, 0};
  VX_CUDA(cudaMemcpyAsync(tot, d_tot.p, sizeof(int32_t) * (L + 2), cudaMemcpyDeviceToHost, st));
  VX_CUDA(cudaStreamSynchronize(st));
  int32_t nn[MAXL + 1] = {0, 0, 0, 0};
  for (int l = 0; l <= L; l++) nn[l] = tot[l];
  const int32_t nseg = tot[L + 1];
  DevBuf &seg_start = T.make(), &seg_leaf = T.make();
  VX_TRY(seg_start.reserve(4 * ((size_t)nseg + 1))); VX_TRY(seg_leaf.reserve(4 * (size_t)nseg));
  NodeArrays nd;
  for (int l = 0; l <= MAXL; l++) {
    nd.start[l] = nd.first_seg[l] = nd.first_leaf[l] = nd.first_child[l] = nd.parent[l] = nullptr;
    if (l > L) continue;
    DevBuf &x = T.make();                           // the five per-node arrays of a layer in one allocation
    const size_t m = (size_t)nn[l] + 1;
    VX_TRY(x.reserve(4 * 5 * m));
    nd.start[l] = x.as<int32_t>(); nd.first_seg[l] = nd.start[l] + m; nd.first_leaf[l] = nd.first_seg[l] + m;
    nd.first_child[l] = nd.first_leaf[l] + m; nd.parent[l] = nd.first_child[l] + m;
  }
#define VX_HB_EMIT(LL) VX_KERNEL(ctx, "k_hb_emit", k_hb_emit<LL><<<(unsigned)n_blk, HB_THREADS, 0, st>>>(hb.as<uint4>(), n16, n_blk, blk.as<int32_t>(), \
                                                                                                   seg_start.as<int32_t>(), seg_leaf.as<int32_t>(), nd))
  VX_HB_L(VX_HB_EMIT);
#undef VX_HB_EMIT
#undef VX_HB_L
  VX_KERNEL(ctx, "k_hb_close", k_hb_close<<<1, 32, 0, st>>>(L, (int32_t)NI, nseg, nd, nn[0], nn[1], nn[2], nn[3], seg_start.as<int32_t>()));

  pt.lap("heads+scans");
  // ---- 4 + 5. node statistics (finest layer from the points, coarser layers from their children), decisions, liveness, emission marks
  DevBuf *nodeS[MAXL + 1], *nodeE[MAXL + 1], *nflag[MAXL + 1], *nspan[MAXL + 1];
  DevBuf &emit_at = T.make(), &vid = T.make();
  VX_TRY(emit_at.reserve(4 * (size_t)nn[L])); VX_TRY(vid.reserve(4 * (size_t)nn[L]));
  VX_CUDA(cudaMemsetAsync(emit_at.p, 0, 4 * (size_t)nn[L], st));
  for (int l = 0; l <= L; l++) {
    nodeS[l] = &T.make(); nodeE[l] = &T.make(); nflag[l] = &T.make(); nspan[l] = &T.make();
    VX_TRY(nodeS[l]->reserve(sizeof(double) * 10 * (size_t)nn[l])); VX_TRY(nodeE[l]->reserve(sizeof(double) * 12 * (size_t)nn[l]));
    VX_TRY(nflag[l]->reserve(4 * (size_t)nn[l])); VX_TRY(nspan[l]->reserve(8 * (size_t)nn[l]));
  }
  // lanes per leaf / per voxel-frame from the mean number of POINTS each holds (any choice gives the same bits).  Measured on the
  // 2,000-keyframe job: bottom level (5 points per leaf, 11 per voxel-frame) 1 lane 2.4 / 2.5 ms, 2 lanes 3.1 / 2.9, 8 lanes 6.2 / 3.4;
  // top level (72 per leaf, 61 per voxel-frame) 1 lane 1.08 / 0.80 ms, 4 lanes 0.79 / 0.57, 8 lanes 0.82 / 0.57.
  auto lanes_for = [](double mean_points) { return mean_points < 16.0 ? 1 : mean_points < 32.0 ? 2 : mean_points < 64.0 ? 4 : 8; };
  static const int gl_env = getenv("VX_SEG_LANES") ? atoi(getenv("VX_SEG_LANES")) : 0;
  {
    const int gl = gl_env ? gl_env : lanes_for((double)NI / std::max(1, nn[L]));
    if (getenv("VX_TIMING")) fprintf(stderr, "[gba_build] NI %lld nseg %d leaves %d (%.2f segments per leaf) -> %d lanes\n", (long long)NI, nseg, nn[L], (double)nseg / std::max(1, nn[L]), gl);
#define VX_LEAF(GL, DEC) VX_KERNEL(ctx, "k_leaf_stats", (k_leaf_stats<GL, DEC><<<blocks_for((int64_t)nn[L] * GL, 128), 128, 0, st>>>(                        \
        d_xyz.as<float>(), d_poses.as<double>(), sidx, k1s, nd.first_seg[L], seg_start.as<int32_t>(), nn[L], prm, L, nodeS[L]->as<double>(),                  \
        nodeE[L]->as<double>(), nspan[L]->as<int32_t>(), nflag[L]->as<int32_t>())))
    if (gl <= 1) { VX_LEAF(1, true); }
    else {
      if (gl == 2) { VX_LEAF(2, false); } else if (gl == 4) { VX_LEAF(4, false); } else { VX_LEAF(8, false); }
      VX_KERNEL(ctx, "k_leaf_decide", k_leaf_decide<<<blocks_for(nn[L], 128), 128, 0, st>>>(nn[L], prm, L, nodeS[L]->as<double>(), nodeE[L]->as<double>(),
                                                                                       nspan[L]->as<int32_t>(), nflag[L]->as<int32_t>()));
    }
#undef VX_LEAF
  }
  pt.lap("leaf_stats");
  for (int l = L - 1; l >= 0; l--)
    VX_KERNEL(ctx, "k_node_stats", k_node_stats<<<blocks_for(nn[l], 128), 128, 0, st>>>(nd.first_child[l], nn[l], prm, l, nn[l + 1], nodeS[l + 1]->as<double>(),
                                                                                     nspan[l + 1]->as<int32_t>(), nodeS[l]->as<double>(), nodeE[l]->as<double>(),
                                                                                     nspan[l]->as<int32_t>(), nflag[l]->as<int32_t>()));
  for (int l = 0; l <= L; l++)        // top-down liveness
    VX_KERNEL(ctx, "k_node_alive", k_node_alive<<<blocks_for(nn[l], 256), 256, 0, st>>>(nd.first_leaf[l], nn[l], l, nd.parent[l], l > 0 ? nflag[l - 1]->as<int32_t>() : nullptr,
                                                                                     nflag[l]->as<int32_t>(), emit_at.as<int32_t>()));
  VX_TRY(excl_scan(ctx, emit_at.as<int32_t>(), vid.as<int32_t>(), nn[L]));
  int32_t V = 0;
  {
    int32_t last_e = 0, last_f = 0;
    VX_TRY(read_last(ctx, vid.as<int32_t>(), nn[L], &last_e));
    VX_TRY(read_last(ctx, emit_at.as<int32_t>(), nn[L], &last_f));
    V = last_e + last_f;
  }
  if (V == 0) return vx_factor_clear(f);

  pt.lap("node_stats");
  // ---- 6. segments -> (voxel, frame) entries
  DevBuf &sk = T.make(), &sk2 = T.make(), &sv = T.make(), &sv2 = T.make(), &vfh = T.make(), &vfi = T.make(), &vfs = T.make();
  VX_TRY(sk.reserve(8 * (size_t)nseg)); VX_TRY(sk2.reserve(8 * (size_t)nseg)); VX_TRY(sv.reserve(4 * (size_t)nseg)); VX_TRY(sv2.reserve(4 * (size_t)nseg));
  VX_TRY(vfh.reserve(4 * (size_t)nseg)); VX_TRY(vfi.reserve(4 * (size_t)nseg));
  LevelPtrs lp;
  for (int l = 0; l <= MAXL; l++) {
    lp.first_leaf[l] = l <= L ? nd.first_leaf[l] : nullptr;
    lp.parent[l] = l <= L ? nd.parent[l] : nullptr;
    lp.flags[l] = l <= L ? nflag[l]->as<int32_t>() : nullptr;
  }
  const int fb = bits_for(F);
  DevBuf &d_nvalid = T.make();
  VX_TRY(d_nvalid.reserve(4));
  VX_CUDA(cudaMemsetAsync(d_nvalid.p, 0, 4, st));
  VX_KERNEL(ctx, "k_seg_voxel", k_seg_voxel<<<blocks_for(nseg, 256), 256, 0, st>>>(seg_start.as<int32_t>(), seg_leaf.as<int32_t>(), k1s, nseg, lp, L, vid.as<int32_t>(), V, fb,
                                                                                sk.as<uint64_t>(), sv.as<int32_t>(), d_nvalid.as<int32_t>()));
  VX_TRY(sort_pairs<uint64_t>(ctx, sk.as<uint64_t>(), sk2.as<uint64_t>(), sv.as<int32_t>(), sv2.as<int32_t>(), nseg, 0, fb + bits_for((int64_t)V + 1)));
  VX_KERNEL(ctx, "k_vf_heads", k_vf_heads<<<blocks_for(nseg, 256), 256, 0, st>>>(sk2.as<uint64_t>(), nseg, V, fb, vfh.as<int32_t>()));
  VX_TRY(incl_scan(ctx, vfh.as<int32_t>(), vfi.as<int32_t>(), nseg));
  int32_t nnz = 0;
  VX_TRY(read_last(ctx, vfi.as<int32_t>(), nseg, &nnz));
  VX_TRY(vfs.reserve(4 * ((size_t)nnz + 1)));
  VX_KERNEL(ctx, "k_starts", k_starts<<<blocks_for(nseg, 256), 256, 0, st>>>(vfh.as<int32_t>(), vfi.as<int32_t>(), nseg, vfs.as<int32_t>()));
  int32_t nvalid32 = 0;
  VX_TRY(read_last(ctx, d_nvalid.as<int32_t>(), 1, &nvalid32));
  const int64_t nvalid = nvalid32;

  pt.lap("seg->vf sort");
  // ---- 7. the device LidarFactor
  VX_TRY(factor_alloc(f, V, nnz));
  {
    const int gl = gl_env ? gl_env : lanes_for((double)nvalid / std::max(1, nnz) * ((double)NI / std::max(1, nseg)));
    if (getenv("VX_TIMING")) fprintf(stderr, "[gba_build] V %d nnz %d valid segments %lld (%.2f per voxel-frame) -> %d lanes\n", V, nnz, (long long)nvalid, (double)nvalid / std::max(1, nnz), gl);
#define VX_VFF(GL) VX_KERNEL(ctx, "k_vf_fill", k_vf_fill<GL><<<blocks_for((int64_t)nnz * GL, 128), 128, 0, st>>>(                                              \
        sk2.as<uint64_t>(), sv2.as<int32_t>(), vfs.as<int32_t>(), nnz, nvalid, seg_start.as<int32_t>(), sidx, d_xyz.as<float>(), fb, f->mom.as<double>(),   \
        f->cnt.as<int32_t>(), f->frame.as<int32_t>(), f->vox.as<int32_t>()))
    if (gl <= 1) { VX_VFF(1); } else if (gl == 2) { VX_VFF(2); } else if (gl == 4) { VX_VFF(4); } else { VX_VFF(8); }
#undef VX_VFF
  }
  for (int l = 0; l <= L; l++)
    VX_KERNEL(ctx, "k_emit_voxels", k_emit_voxels<<<blocks_for(nn[l], 256), 256, 0, st>>>(nd.first_leaf[l], nflag[l]->as<int32_t>(), nn[l], vid.as<int32_t>(),
                                                                                       nodeS[l]->as<double>(), nodeE[l]->as<double>(), V, f->lam.as<double>(), f->U.as<double>(),
                                                                                       f->mrg.as<double>(), f->fix.as<double>(), f->coe.as<double>()));
  VX_KERNEL(ctx, "k_vox_off", k_vox_off<<<blocks_for((int64_t)V + 1, 256), 256, 0, st>>>(f->vox.as<int32_t>(), nnz, V, f->off.as<int32_t>()));
  VX_KERNEL(ctx, "k_vox_win", k_vox_win<<<blocks_for(V, 256), 256, 0, st>>>(f->off.as<int32_t>(), f->frame.as<int32_t>(), V, W, f->win.as<int32_t>()));
  VX_KERNEL(ctx, "k_win_off", k_win_off<<<blocks_for(NW + 1, 128), 128, 0, st>>>(f->win.as<int32_t>(), V, NW, f->wv_off.as<int32_t>()));
  VX_CUDA(cudaStreamSynchronize(st));
  pt.lap("emit");
  return factor_finish_upload(f);
}
}  // namespace vx

extern "C" int vx_gba_build(vx_ctx *ctx, const float *xyz, const int64_t *cloud_off, int n_clouds, const int32_t *frame_cloud,
                            const vx_pose *poses, const vx_gba_params *params, vx_factor *f) {
  VX_ARG(cloud_off != nullptr && n_clouds >= 1);
  VX_TRY(bind(ctx));
  const int64_t NP = cloud_off[n_clouds];
  VX_ARG(NP >= 0 && (NP == 0 || xyz != nullptr));
  DevBuf d_xyz, d_coff;
  int s = VX_OK;
  auto body = [&]() -> int {
    VX_TRY(d_xyz.reserve(sizeof(float) * 3 * std::max<int64_t>(NP, 1)));
    VX_TRY(d_coff.reserve(8 * (size_t)(n_clouds + 1)));
    if (NP) VX_CUDA(cudaMemcpyAsync(d_xyz.p, xyz, sizeof(float) * 3 * NP, cudaMemcpyHostToDevice, ctx->stream));
    VX_CUDA(cudaMemcpyAsync(d_coff.p, cloud_off, 8 * (size_t)(n_clouds + 1), cudaMemcpyHostToDevice, ctx->stream));
    return gba_build_impl(ctx, d_xyz.as<float>(), d_coff.as<int64_t>(), cloud_off, n_clouds, frame_cloud, poses, params, f, nullptr, 0);
  };
  s = body();
  cudaStreamSynchronize(ctx->stream);
  d_xyz.release(); d_coff.release();
  return s;
}

// ---- multi-GPU voxelization: key every point, route it to the rank that owns its ROOT voxel ----------
#include "vx_planemap.cuh"   // mix64
namespace {
__global__ void k_point_owner(const float *__restrict__ xyz, const int32_t *__restrict__ frame, int64_t n, const double *__restrict__ poses, double vs,
                              int n_ranks, int32_t *__restrict__ owner, int *__restrict__ bad) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const d3 w = world_of(poses + 12 * (int64_t)frame[i], (double)xyz[3 * i], (double)xyz[3 * i + 1], (double)xyz[3 * i + 2]);
  uint64_t k;
  if (!pack_key(key_axis_d(w.x, vs), key_axis_d(w.y, vs), key_axis_d(w.z, vs), k)) *bad = 1;
  owner[i] = (int32_t)(mix64(k) % (uint64_t)n_ranks);
}
}  // namespace

// All pointers except `poses` are DEVICE pointers (the exchange that follows is an NCCL all-to-all on those buffers).
// owner[i] = mix64(packed root key of R_f p_i + t_f at `voxel_size`) mod n_ranks: every root voxel lives on exactly one rank.
extern "C" int vx_gba_point_owner(vx_ctx *ctx, const float *d_xyz, const int32_t *d_frame, int64_t n, const vx_pose *poses, int n_frames,
                                  double voxel_size, int n_ranks, int32_t *d_owner) {
  VX_ARG(ctx && poses && n >= 0 && n_frames >= 1 && n_ranks >= 1 && voxel_size > 0 && (n == 0 || (d_xyz && d_frame && d_owner)));
  VX_TRY(bind(ctx));
  if (n == 0) return VX_OK;
  VX_TRY(ctx->scratch.reserve(sizeof(double) * 12 * n_frames + 64));
  double *d_poses = ctx->scratch.as<double>();
  int *d_bad = reinterpret_cast<int *>(d_poses + 12 * (size_t)n_frames);
  VX_CUDA(cudaMemcpyAsync(d_poses, poses, sizeof(vx_pose) * n_frames, cudaMemcpyHostToDevice, ctx->stream));
  VX_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), ctx->stream));
  VX_KERNEL(ctx, "k_point_owner", k_point_owner<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(d_xyz, d_frame, n, d_poses, voxel_size, n_ranks, d_owner, d_bad));
  int bad = 0;
  VX_CUDA(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  VX_CUDA(cudaStreamSynchronize(ctx->stream));
  if (bad) { set_error("vx_gba_point_owner: a cell index does not fit the 21-bit packed key"); return VX_ERR_ARG; }
  return VX_OK;
}

// vx_gba_build on DEVICE-resident clouds (d_xyz: float triples; cloud_off: HOST offsets, n_clouds + 1).
extern "C" int vx_gba_build_device(vx_ctx *ctx, const float *d_xyz, const int64_t *cloud_off, int n_clouds, const int32_t *frame_cloud,
                                   const vx_pose *poses, const vx_gba_params *params, vx_factor *f) {
  VX_ARG(cloud_off != nullptr && n_clouds >= 1);
  VX_TRY(bind(ctx));
  DevBuf d_coff;
  auto body = [&]() -> int {
    VX_TRY(d_coff.reserve(8 * (size_t)(n_clouds + 1)));
    VX_CUDA(cudaMemcpyAsync(d_coff.p, cloud_off, 8 * (size_t)(n_clouds + 1), cudaMemcpyHostToDevice, ctx->stream));
    return gba_build_impl(ctx, d_xyz, d_coff.as<int64_t>(), cloud_off, n_clouds, frame_cloud, poses, params, f, nullptr, 0);
  };
  const int s = body();
  cudaStreamSynchronize(ctx->stream);
  d_coff.release();
  return s;
}

// ---- sharded build: the exchange inside the library ---------------------------------
// Every rank holds some of the clouds (one frame each).  k_owner_frame keys every local point at the CURRENT poses and names the
// rank that owns its root voxel; a stable bucket partition (k_part_count -> scan -> k_part_scatter: points of one destination keep
// their order, so a frame's points arrive in cloud order and every (leaf, frame) moment is summed in the single-GPU order, bit for
// bit) lays them out by destination; one grouped ncclSend / ncclRecv round moves 16 B per point (float xyz + frame id); the
// receiver runs the ordinary sort / segment / recut on what it owns.  Hess / JacT / cost are then sums over ranks.
#include "vx_comm.cuh"
namespace {
constexpr int PART_THREADS = 256;
constexpr int PART_TILE = 2048;      // points per block, walked in order in chunks of PART_THREADS
constexpr int PART_MAXR = 64;        // ranks

__global__ void k_owner_frame(const float *__restrict__ xyz, const int64_t *__restrict__ coff, const int32_t *__restrict__ cloud_frame, int n_clouds, int64_t n,
                              const double *__restrict__ poses, double vs, int n_ranks, uint8_t *__restrict__ owner, int32_t *__restrict__ frame,
                              int *__restrict__ bad) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int lo = 0, hi = n_clouds;   // largest c with coff[c] <= i
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (coff[mid] <= i) lo = mid; else hi = mid; }
  const int g = cloud_frame[lo];
  const d3 w = world_of(poses + 12 * (int64_t)g, (double)xyz[3 * i], (double)xyz[3 * i + 1], (double)xyz[3 * i + 2]);
  uint64_t k;
  if (!pack_key(key_axis_d(w.x, vs), key_axis_d(w.y, vs), key_axis_d(w.z, vs), k)) *bad = 1;
  owner[i] = (uint8_t)(mix64(k) % (uint64_t)n_ranks);
  frame[i] = g;
}
// cnt[o * n_blocks + b] = points of tile b that go to rank o
__global__ void __launch_bounds__(PART_THREADS)
k_part_count(const uint8_t *__restrict__ owner, int64_t n, int n_ranks, int n_blocks, int32_t *__restrict__ cnt) {
  __shared__ int sc[PART_MAXR];
  if (threadIdx.x < PART_MAXR) sc[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * PART_TILE;
  for (int k = threadIdx.x; k < PART_TILE; k += PART_THREADS) {
    const int64_t i = base + k;
    if (i < n) atomicAdd(&sc[owner[i]], 1);
  }
  __syncthreads();
  if (threadIdx.x < n_ranks) cnt[(size_t)threadIdx.x * n_blocks + blockIdx.x] = sc[threadIdx.x];
}
// stable scatter: point i of tile b goes to off[owner * n_blocks + b] + (its rank among the tile's earlier points of that owner)
__global__ void __launch_bounds__(PART_THREADS)
k_part_scatter(const float *__restrict__ xyz, const uint8_t *__restrict__ owner, const int32_t *__restrict__ frame, int64_t n, int n_ranks, int n_blocks,
               const int32_t *__restrict__ off, float *__restrict__ sxyz, int32_t *__restrict__ sframe) {
  __shared__ int run[PART_MAXR];                       // points of each owner already placed by earlier chunks
  __shared__ int wc[PART_THREADS / 32][PART_MAXR];     // per-warp counts of the current chunk
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  if (threadIdx.x < PART_MAXR) run[threadIdx.x] = 0;
  for (int k = threadIdx.x; k < (PART_THREADS / 32) * PART_MAXR; k += PART_THREADS) (&wc[0][0])[k] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * PART_TILE;
  for (int c = 0; c < PART_TILE; c += PART_THREADS) {
    const int64_t i = base + c + threadIdx.x;
    const int o = i < n ? (int)owner[i] : 255;
    const unsigned m = __match_any_sync(0xffffffffu, o);
    const int rw = __popc(m & ((1u << lane) - 1u));
    if (o != 255 && rw == 0) wc[wp][o] = __popc(m);
    __syncthreads();
    if (o != 255) {
      int pre = run[o];
      for (int w2 = 0; w2 < wp; w2++) pre += wc[w2][o];
      const int64_t pos = (int64_t)off[(size_t)o * n_blocks + blockIdx.x] + pre + rw;
      sxyz[3 * pos] = xyz[3 * i]; sxyz[3 * pos + 1] = xyz[3 * i + 1]; sxyz[3 * pos + 2] = xyz[3 * i + 2];
      sframe[pos] = frame[i];
    }
    __syncthreads();
    if (threadIdx.x < n_ranks) {
      int s = 0;
      for (int w2 = 0; w2 < PART_THREADS / 32; w2++) { s += wc[w2][threadIdx.x]; wc[w2][threadIdx.x] = 0; }
      run[threadIdx.x] += s;
    }
    __syncthreads();
  }
}
}  // namespace

namespace vx {
// d_xyz: this rank's clouds on the device (cloud c = points [cloud_off[c], cloud_off[c+1]), all of frame cloud_frame[c]);
// poses: ALL F frames (replicated).  *n_owned = points this rank received.  stats (may be null): [0] ms partition, [1] ms exchange.
int gba_build_sharded_impl(vx_ctx *ctx, vx_comm *comm, const float *d_xyz, const int64_t *cloud_off, int n_clouds, const int32_t *cloud_frame,
                           const vx_pose *poses, const vx_gba_params *params, vx_factor *f, int64_t *n_owned) {
  VX_ARG(ctx && comm && comm->ctx == ctx && f && params && poses && cloud_off && n_clouds >= 0 && comm->size <= PART_MAXR);
  VX_TRY(bind(ctx));
  cudaStream_t st = ctx->stream;
  const int G = comm->size, F = f->W * f->n_win;
  const int64_t n = n_clouds > 0 ? cloud_off[n_clouds] : 0;
  VX_ARG(n < ((int64_t)1 << 31) && (n == 0 || (d_xyz && cloud_frame)));
  for (int c = 0; c < n_clouds; c++) VX_ARG(cloud_frame[c] >= 0 && cloud_frame[c] < F);
  const int nb = (int)std::max<int64_t>(1, (n + PART_TILE - 1) / PART_TILE);
  DevBuf d_coff, d_cf, d_poses, d_owner, d_frame, d_cnt, d_off, d_sxyz, d_sframe, d_rxyz, d_rframe, d_bad;
  std::vector<DevBuf *> all = {&d_coff, &d_cf, &d_poses, &d_owner, &d_frame, &d_cnt, &d_off, &d_sxyz, &d_sframe, &d_rxyz, &d_rframe, &d_bad};
  auto body = [&]() -> int {
    std::vector<int64_t> send_cnt(G, 0), send_off(G, 0), recv_cnt(G, 0), recv_off(G, 0);
    if (n > 0) {
      VX_TRY(d_coff.reserve(8 * (size_t)(n_clouds + 1))); VX_TRY(d_cf.reserve(4 * (size_t)n_clouds)); VX_TRY(d_poses.reserve(sizeof(vx_pose) * (size_t)F));
      VX_TRY(d_owner.reserve(n)); VX_TRY(d_frame.reserve(4 * n)); VX_TRY(d_cnt.reserve(4 * ((size_t)G * nb + 1))); VX_TRY(d_off.reserve(4 * ((size_t)G * nb + 1)));
      VX_TRY(d_sxyz.reserve(12 * n)); VX_TRY(d_sframe.reserve(4 * n)); VX_TRY(d_bad.reserve(4));
      VX_CUDA(cudaMemcpyAsync(d_coff.p, cloud_off, 8 * (size_t)(n_clouds + 1), cudaMemcpyHostToDevice, st));
      VX_CUDA(cudaMemcpyAsync(d_cf.p, cloud_frame, 4 * (size_t)n_clouds, cudaMemcpyHostToDevice, st));
      VX_CUDA(cudaMemcpyAsync(d_poses.p, poses, sizeof(vx_pose) * (size_t)F, cudaMemcpyHostToDevice, st));
      VX_CUDA(cudaMemsetAsync(d_bad.p, 0, 4, st));
      VX_KERNEL(ctx, "k_owner_frame", k_owner_frame<<<blocks_for(n, 256), 256, 0, st>>>(d_xyz, d_coff.as<int64_t>(), d_cf.as<int32_t>(), n_clouds, n, d_poses.as<double>(),
                                                                                       params->voxel_size, G, d_owner.as<uint8_t>(), d_frame.as<int32_t>(), d_bad.as<int>()));
      VX_KERNEL(ctx, "k_part_count", k_part_count<<<nb, PART_THREADS, 0, st>>>(d_owner.as<uint8_t>(), n, G, nb, d_cnt.as<int32_t>()));
      VX_TRY(excl_scan(ctx, d_cnt.as<int32_t>(), d_off.as<int32_t>(), (int64_t)G * nb));
      VX_KERNEL(ctx, "k_part_scatter", k_part_scatter<<<nb, PART_THREADS, 0, st>>>(d_xyz, d_owner.as<uint8_t>(), d_frame.as<int32_t>(), n, G, nb, d_off.as<int32_t>(),
                                                                                 d_sxyz.as<float>(), d_sframe.as<int32_t>()));
      // first slot of every destination = off[o * nb]
      std::vector<int32_t> first(G);
      int bad = 0;
      VX_CUDA(cudaMemcpy2DAsync(first.data(), 4, d_off.p, 4 * (size_t)nb, 4, G, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaMemcpyAsync(&bad, d_bad.p, 4, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaStreamSynchronize(st));
      if (bad) { set_error("sharded GBA build: a root cell index does not fit the 21-bit packed key"); return VX_ERR_ARG; }
      for (int o = 0; o < G; o++) { send_off[o] = first[o]; send_cnt[o] = (o + 1 < G ? first[o + 1] : n) - first[o]; }
    }
    // counts: every rank learns the whole matrix M[s][d]
    std::vector<int64_t> M((size_t)G * G, 0);
    if (G > 1) {
      VX_TRY(comm->counts.reserve(8 * (size_t)G * (G + 1)));
      int64_t *d_m = comm->counts.as<int64_t>();
      VX_CUDA(cudaMemcpyAsync(d_m + (size_t)G * G, send_cnt.data(), 8 * (size_t)G, cudaMemcpyHostToDevice, st));
      VX_TRY(comm_allgather_bytes(comm, d_m + (size_t)G * G, d_m, 8 * (size_t)G));
      VX_CUDA(cudaMemcpyAsync(M.data(), d_m, 8 * (size_t)G * G, cudaMemcpyDeviceToHost, st));
      VX_CUDA(cudaStreamSynchronize(st));
    } else {
      M[0] = send_cnt[0];
    }
    int64_t nrecv = 0;
    for (int s = 0; s < G; s++) { recv_cnt[s] = M[(size_t)s * G + comm->rank]; recv_off[s] = nrecv; nrecv += recv_cnt[s]; }
    VX_ARG(nrecv < ((int64_t)1 << 31));
    if (n_owned) *n_owned = nrecv;
    VX_TRY(d_rxyz.reserve(12 * (size_t)std::max<int64_t>(nrecv, 1))); VX_TRY(d_rframe.reserve(4 * (size_t)std::max<int64_t>(nrecv, 1)));
    VX_TRY(comm_alltoallv(comm, d_sxyz.p, send_off.data(), send_cnt.data(), d_rxyz.p, recv_off.data(), recv_cnt.data(), 12));
    VX_TRY(comm_alltoallv(comm, d_sframe.p, send_off.data(), send_cnt.data(), d_rframe.p, recv_off.data(), recv_cnt.data(), 4));
    d_owner.release(); d_frame.release(); d_cnt.release(); d_off.release();
    if (nrecv == 0) return vx_factor_clear(f);
    return gba_build_impl(ctx, d_rxyz.as<float>(), nullptr, nullptr, 0, nullptr, poses, params, f, d_rframe.as<int32_t>(), nrecv);
  };
  const int s = body();
  cudaStreamSynchronize(st);
  for (DevBuf *b : all) b->release();
  return s;
}
}  // namespace vx

// OctreeGBA::cut_voxel + OctreeGBA_multi_recut (loop_refine.hpp:446-537) with the ROOT voxels partitioned over the ranks of `comm`.
extern "C" int vx_gba_build_sharded(vx_ctx *ctx, vx_comm *comm, const float *xyz, const int64_t *cloud_off, int n_clouds, const int32_t *cloud_frame,
                                    const vx_pose *poses, const vx_gba_params *params, vx_factor *f, int64_t *n_points_owned) {
  VX_ARG(ctx && comm && cloud_off && n_clouds >= 0);
  VX_TRY(bind(ctx));
  const int64_t NP = n_clouds > 0 ? cloud_off[n_clouds] : 0;
  VX_ARG(NP >= 0 && (NP == 0 || xyz != nullptr));
  DevBuf d_xyz;
  auto body = [&]() -> int {
    VX_TRY(d_xyz.reserve(12 * (size_t)std::max<int64_t>(NP, 1)));
    if (NP) VX_CUDA(cudaMemcpyAsync(d_xyz.p, xyz, 12 * (size_t)NP, cudaMemcpyHostToDevice, ctx->stream));
    return gba_build_sharded_impl(ctx, comm, d_xyz.as<float>(), cloud_off, n_clouds, cloud_frame, poses, params, f, n_points_owned);
  };
  const int s = body();
  cudaStreamSynchronize(ctx->stream);
  d_xyz.release();
  return s;
}
