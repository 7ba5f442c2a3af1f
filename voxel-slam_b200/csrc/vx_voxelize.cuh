// vx_voxelize.cuh — voxel key rule shared by every voxelization kernel.
#pragma once
#include "vx_common.cuh"

namespace vx {

// The reference's key rule, inlined 7x (voxel_map.hpp:1907-1918, :1977-1990, :2114-2125, :2173-2184,
// :42-52; loop_refine.hpp:452-459):  float loc = coord / voxel_size;  if (loc < 0) loc -= 1.0;  (int64_t)loc.
// The divide is done in double, ROUNDED TO FLOAT, the -1 is applied to the float, and the conversion
// truncates toward zero: (-1,0) -> -1, exactly -1.0 -> -2, -0.0 -> 0.  Not floor().
__host__ __device__ __forceinline__ int64_t key_axis_d(double coord, double voxel_size) {
#ifdef __CUDA_ARCH__
  float loc = __double2float_rn(__ddiv_rn(coord, voxel_size));
  if (loc < 0.0f) loc = __fadd_rn(loc, -1.0f);
#else
  float loc = (float)(coord / voxel_size);
  if (loc < 0.0f) loc = loc - 1.0f;
#endif
  return (int64_t)loc;
}
// tools.hpp:211 / :250 — the coordinate is a float promoted to double; "loc -= 1.0" goes through double
// and back to float, which rounds to the same float as the float subtraction.
__host__ __device__ __forceinline__ int64_t key_axis_f(float coord, double voxel_size) { return key_axis_d((double)coord, voxel_size); }

// Packed key: 21 bits per axis, biased by 2^20.  Returns false when a cell index does not fit.
constexpr int64_t KEY_BIAS = (int64_t)1 << 20;
__host__ __device__ __forceinline__ bool pack_key(int64_t x, int64_t y, int64_t z, uint64_t &out) {
  const int64_t bx = x + KEY_BIAS, by = y + KEY_BIAS, bz = z + KEY_BIAS;
  const int64_t lim = (int64_t)1 << 21;
  if (bx < 0 || by < 0 || bz < 0 || bx >= lim || by >= lim || bz >= lim) { out = ~0ull; return false; }
  out = ((uint64_t)bx << 42) | ((uint64_t)by << 21) | (uint64_t)bz;
  return true;
}
__host__ __device__ __forceinline__ void unpack_key(uint64_t k, int64_t &x, int64_t &y, int64_t &z) {
  x = (int64_t)((k >> 42) & 0x1fffff) - KEY_BIAS;
  y = (int64_t)((k >> 21) & 0x1fffff) - KEY_BIAS;
  z = (int64_t)(k & 0x1fffff) - KEY_BIAS;
}

enum DsMode { DS_MEAN, DS_CLOSE, DS_PVEC };   // A2 down_sampling_voxel, A4 down_sampling_close, A3 down_sampling_pvec
// down-sampling of device-resident points (vx_point48, or vx_point_var for DS_PVEC) into d_out (room for n records)
int downsample_device(vx_ctx *ctx, const void *d_pts, int64_t n, double voxel_size, DsMode mode, vx_point48 *d_out, int32_t *ncell_out);

}  // namespace vx
