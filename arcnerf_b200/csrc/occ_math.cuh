// Device arithmetic shared by the ray/occupancy kernels (march.cu, occupancy.cu).  Every user of this
// header is compiled with -fmad=false: results must equal the reference's HOST_DEVICE functions evaluated
// one rounded fp32 operation at a time (arcnerf/ops/include/volume_func.h), which is what oracle/ pins.
#pragma once
#include "common.cuh"

namespace occ {

struct F3 { float x, y, z; };
struct Box { F3 lo, hi; };

__device__ __forceinline__ F3 load3(const float *p) { return F3{p[0], p[1], p[2]}; }
__device__ __forceinline__ float min3f(float a, float b, float c) { float m = a < b ? a : b; return m < c ? m : c; }
__device__ __forceinline__ float max3f(float a, float b, float c) { float m = a > b ? a : b; return m > c ? m : c; }
__device__ __forceinline__ float clampf(float v, float lo, float hi) { return v < lo ? lo : (hi < v ? hi : v); }
__device__ __forceinline__ F3 ray_at(F3 o, F3 d, float t) { return F3{o.x + d.x * t, o.y + d.y * t, o.z + d.z * t}; }
__device__ __forceinline__ bool inside(F3 p, const Box &b) {
    return p.x >= b.lo.x && p.y >= b.lo.y && p.z >= b.lo.z && p.x <= b.hi.x && p.y <= b.hi.y && p.z <= b.hi.z;
}

// Slab test with the reference's ordering of comparisons (volume_func.h:17-56): per-axis swap, early
// reject between axes, (-1,-1) for a miss.
__device__ __forceinline__ void slab_test(F3 o, F3 d, F3 lo, F3 hi, float &tn, float &tf) {
    float tmin = (lo.x - o.x) / d.x, tmax = (hi.x - o.x) / d.x, s;
    if (tmin > tmax) { s = tmin; tmin = tmax; tmax = s; }
    float tymin = (lo.y - o.y) / d.y, tymax = (hi.y - o.y) / d.y;
    if (tymin > tymax) { s = tymin; tymin = tymax; tymax = s; }
    if (tmin > tymax || tymin > tmax) { tn = -1.0f; tf = -1.0f; return; }
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    float tzmin = (lo.z - o.z) / d.z, tzmax = (hi.z - o.z) / d.z;
    if (tzmin > tzmax) { s = tzmin; tzmin = tzmax; tzmax = s; }
    if (tmin > tzmax || tzmin > tmax) { tn = -1.0f; tf = -1.0f; return; }
    if (tzmin > tmin) tmin = tzmin;
    if (tzmax < tmax) tmax = tzmax;
    tn = tmin; tf = tmax;
}

// Bool volume, flat index x*n*n + y*n + z; outside the grid reads as empty (volume_func.h:59-88).
__device__ __forceinline__ bool occupied_flat(F3 p, const uint8_t *__restrict__ vol, const Box &b, uint32_t n) {
    const float fn = (float)n;
    const float vx = (p.x - b.lo.x) / ((b.hi.x - b.lo.x) / fn);
    const float vy = (p.y - b.lo.y) / ((b.hi.y - b.lo.y) / fn);
    const float vz = (p.z - b.lo.z) / ((b.hi.z - b.lo.z) / fn);
    if (min3f(vx, vy, vz) < 0 || max3f(vx, vy, vz) >= fn) return false;
    const uint32_t x = (uint32_t)floorf(vx), y = (uint32_t)floorf(vy), z = (uint32_t)floorf(vz);
    return __ldg(vol + (x * (n * n) + y * n + z)) != 0;
}

__device__ __forceinline__ uint32_t spread3(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__device__ __forceinline__ uint32_t morton3(uint32_t x, uint32_t y, uint32_t z) {
    return spread3(x) | (spread3(y) << 1) | (spread3(z) << 2);
}
__device__ __forceinline__ uint32_t unspread3(uint32_t x) {
    x &= 0x49249249u;
    x = (x | (x >> 2)) & 0xc30c30c3u;
    x = (x | (x >> 4)) & 0x0f00f00fu;
    x = (x | (x >> 8)) & 0xff0000ffu;
    x = (x | (x >> 16)) & 0x0000ffffu;
    return x;
}

// Truncating cast, clamp into the grid, Morton code (volume_func.h:169-181).
__device__ __forceinline__ uint32_t cell_morton(F3 p, const Box &b, uint32_t n) {
    const float fn = (float)n;
    const int ix = (int)((p.x - b.lo.x) / ((b.hi.x - b.lo.x) / fn));
    const int iy = (int)((p.y - b.lo.y) / ((b.hi.y - b.lo.y) / fn));
    const int iz = (int)((p.z - b.lo.z) / ((b.hi.z - b.lo.z) / fn));
    const float top = (float)((int)n - 1);
    return morton3((uint32_t)clampf((float)ix, 0.f, top), (uint32_t)clampf((float)iy, 0.f, top),
                   (uint32_t)clampf((float)iz, 0.f, top));
}
__device__ __forceinline__ bool occupied_bit(F3 p, const uint8_t *__restrict__ bits, const Box &b, uint32_t n) {
    const uint32_t idx = cell_morton(p, b, n);
    return (__ldg(bits + (idx >> 3)) & (1u << (idx & 7u))) != 0;
}

// Distance (in t) to the next voxel boundary, reference form (volume_func.h:100-118): positions are
// scaled by n_grid whatever the box size, the nearest axis crossing wins, never negative.
__device__ __forceinline__ float next_voxel_distance(F3 p, F3 d, const Box &b, uint32_t n) {
    const float fn = (float)n;
    const float cx = (b.lo.x + b.hi.x) / 2.0f, cy = (b.lo.y + b.hi.y) / 2.0f, cz = (b.lo.z + b.hi.z) / 2.0f;
    const float hx = (b.hi.x - b.lo.x) / 2.0f, hy = (b.hi.y - b.lo.y) / 2.0f, hz = (b.hi.z - b.lo.z) / 2.0f;
    const float ix = 1.0f / d.x, iy = 1.0f / d.y, iz = 1.0f / d.z;
    const float px = fn * p.x, py = fn * p.y, pz = fn * p.z;
    const float tx = (floorf((px + cx) + hx * copysignf(1.0f, d.x)) - px) * ix;
    const float ty = (floorf((py + cy) + hy * copysignf(1.0f, d.y)) - py) * iy;
    const float tz = (floorf((pz + cz) + hz * copysignf(1.0f, d.z)) - pz) * iz;
    return fmaxf(min3f(tx, ty, tz) / fn, 0.0f);
}

__device__ __forceinline__ Box load_box(const float *__restrict__ a) {
    return Box{F3{__ldg(a), __ldg(a + 1), __ldg(a + 2)}, F3{__ldg(a + 3), __ldg(a + 4), __ldg(a + 5)}};
}

}  // namespace occ
