// vx_planemap.cuh — the flattened voxel-plane map OctoTree::match reads (voxel_map.hpp:1649-1721, :2167-2205):
// an open-addressing hash (packed root key -> root node, linear probing, ~0 = empty) over a vx_map_node table.
// Owned either by vx_planemap_create (uploaded from the host) or by a vx_map (device-resident sliding-window map,
// vx_map.cu), which hands out a non-owning view.
#pragma once
#include "vx_common.cuh"

struct vx_planemap {
  vx_ctx *ctx = nullptr;
  vx::DevBuf hkeys, hvals, nodes;
  uint64_t hmask = 0;
  int64_t n_nodes = 0, n_roots = 0;
  double voxel_size = 1.0;
  bool owns = true;
};

namespace vx {
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 27; x *= 0x94d049bb133111ebull; x ^= x >> 31;
  return x;
}
// node index of the root voxel with this packed key, -1 when absent
__device__ __forceinline__ int32_t root_lookup(const uint64_t *__restrict__ hkeys, const int32_t *__restrict__ hvals, uint64_t hmask, uint64_t key) {
  uint64_t h = mix64(key) & hmask;
  for (;;) {
    const uint64_t k = hkeys[h];
    if (k == key) return hvals[h];
    if (k == ~0ull) return -1;
    h = (h + 1) & hmask;
  }
}
}  // namespace vx
