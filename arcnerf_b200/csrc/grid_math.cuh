// Level geometry and cell arithmetic of the multiresolution hash grid, shared by the stand-alone encoding kernels
// (hashgrid.cu) and the fused field kernels (field_fused.cuh).  Semantics: tiny-cuda-nn encodings/grid.h:111-204 and
// common_device.h:402-443 (see the header of hashgrid.cu).
#pragma once
#include <stdint.h>

namespace gridmath {

constexpr int kMaxLevels = 32;

struct GridLevel {
    uint32_t offset;   // first entry of this level in the flat table
    uint32_t size;     // entries in this level
    uint32_t res;      // vertices per axis
    float scale;
    uint32_t hashed;   // 1: coherent-prime hash, 0: dense stride index
    uint32_t mask;     // size-1 when size is a power of two, else 0
};
struct GridGeom {
    GridLevel lv[kMaxLevels];
    uint32_t n_levels;
};

__device__ __forceinline__ uint32_t entry_index(const GridLevel &g, uint32_t x, uint32_t y, uint32_t z) {
    uint32_t idx;
    if (g.hashed) idx = x ^ (y * 2654435761u) ^ (z * 805459861u);
    else idx = x + y * g.res + z * (g.res * g.res);
    if (g.mask) return idx & g.mask;
    return idx < g.size ? idx : idx % g.size;
}

struct Cell {
    uint32_t gx, gy, gz;
    float fx, fy, fz;
};
__device__ __forceinline__ Cell locate(const GridLevel &g, float x, float y, float z) {
    Cell c;
    float px = fmaf(x, g.scale, 0.5f), py = fmaf(y, g.scale, 0.5f), pz = fmaf(z, g.scale, 0.5f);
    const float flx = floorf(px), fly = floorf(py), flz = floorf(pz);
    c.gx = (uint32_t)(int)flx; c.gy = (uint32_t)(int)fly; c.gz = (uint32_t)(int)flz;
    c.fx = px - flx; c.fy = py - fly; c.fz = pz - flz;
    return c;
}


// host: per-level records from the arrays arcb200_grid_geometry produced; the reference's index loop (grid.h:170-187) decides
// dense vs hashed: dense while the running stride fits the level
static inline int pack_geom(GridGeom &geom, int n_levels, const uint32_t *offsets, const float *scales, const uint32_t *resolutions) {
    if (n_levels < 1 || n_levels > kMaxLevels) return -1;
    geom.n_levels = (uint32_t)n_levels;
    for (int l = 0; l < n_levels; ++l) {
        GridLevel &g = geom.lv[l];
        g.offset = offsets[l];
        g.size = offsets[l + 1] - offsets[l];
        g.res = resolutions[l];
        g.scale = scales[l];
        uint32_t stride = 1;
        for (int dim = 0; dim < 3 && stride <= g.size; ++dim) stride *= g.res;
        g.hashed = g.size < stride ? 1u : 0u;
        g.mask = (g.size & (g.size - 1)) == 0 ? g.size - 1 : 0u;
        if (g.size == 0) return -1;
    }
    return 0;
}

}  // namespace gridmath
