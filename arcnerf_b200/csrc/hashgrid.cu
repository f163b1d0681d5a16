// Multiresolution hash-grid encoding (instant-ngp / tiny-cuda-nn "HashGrid", 3-D inputs, linear
// interpolation, coherent-prime hash): forward (+ optional dy/dx), gradient w.r.t. the table, gradient w.r.t.
// the input, and the double-backward trio the eikonal loss needs.
//
// Semantics follow tiny-cuda-nn (vendored at /root/reference/include/tiny-cuda-nn, the `backend: tcnn`
// path of ArcNerf's HashGridEmbedder, hashgrid_encoder.py:102-113,182-184):
//   level geometry   encodings/grid.h:195-204,982-1012   scale = exp2f(l*log2(s))*base - 1, res = ceil(scale)+1,
//                                                        entries rounded up to 8 and capped at 2^log2_T
//   position         common_device.h:402-443             pos = x*scale + 0.5; cell = floor(pos); frac = pos - cell
//   index            encodings/grid.h:111-187            dense x + y*res + z*res^2, or x ^ y*2654435761 ^ z*805459861,
//                                                        both modulo the level's entry count
//   table layout     flat [level][entry][feature], levels at offset_table (grid.h:1001-1012)
// Differences from the reference kernels, by design:
//   * the table is read as fp16 (the reference's parameter precision) but blended in fp32 and rounded once
//     (the reference rounds to fp16 after every corner, grid.h:339-341): results are closer to the fp32 math;
//   * table gradients are accumulated in fp32 with vector atomics (reference: half2 atomics + loss scale);
//   * output is written row-major (N, L*F) directly through a shared-memory tile (reference: SoA + transpose);
//   * level geometry is computed once on the host (arcb200_grid_geometry) instead of per thread.
//
// Roofline: a pure gather - per sample 8 corners x L levels x F fp16 (512 B at L=16, F=2) out of a table that
// is L2-resident on B200 (24.4 MB fp16 << 126 MB L2).  HBM sees only 12 B in + 2*L*F B out per sample.
#include "common.cuh"
#include "grid_math.cuh"
#include <math.h>

namespace {

using namespace gridmath;

template <int F>
__device__ __forceinline__ void load_entry(const __half *__restrict__ level_table, uint32_t idx, float (&out)[F]);
template <> __device__ __forceinline__ void load_entry<1>(const __half *__restrict__ level_table, uint32_t idx, float (&out)[1]) {
    out[0] = __half2float(__ldg(level_table + idx));
}
template <> __device__ __forceinline__ void load_entry<2>(const __half *__restrict__ level_table, uint32_t idx, float (&out)[2]) {
    const float2 v = __half22float2(__ldg(reinterpret_cast<const __half2 *>(level_table) + idx));
    out[0] = v.x; out[1] = v.y;
}
template <> __device__ __forceinline__ void load_entry<4>(const __half *__restrict__ level_table, uint32_t idx, float (&out)[4]) {
    const uint2 r = __ldg(reinterpret_cast<const uint2 *>(level_table) + idx);
    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&r.x));
    const float2 b = __half22float2(*reinterpret_cast<const __half2 *>(&r.y));
    out[0] = a.x; out[1] = a.y; out[2] = b.x; out[3] = b.y;
}
template <> __device__ __forceinline__ void load_entry<8>(const __half *__restrict__ level_table, uint32_t idx, float (&out)[8]) {
    const uint4 r = __ldg(reinterpret_cast<const uint4 *>(level_table) + idx);
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
    #pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&w[k]));
        out[2 * k] = a.x; out[2 * k + 1] = a.y;
    }
}

constexpr int kTileSamples = 32;   // samples per CTA tile (one per lane)
constexpr int kFwdWarps = 8;

// Forward.  CTA = 8 warps x 32 lanes; lane = sample of the tile, warp w takes levels w, w+8, ...
// The 32 x (L*F) fp16 tile is assembled in shared memory and stored with fully coalesced 4-byte words.
// DYDX: also writes dy/dx (N, L*F, 3) fp32 for the input gradient / normals.
template <int F, bool DYDX>
__global__ void __launch_bounds__(kTileSamples * kFwdWarps)
grid_fwd_kernel(uint32_t n, const int32_t *__restrict__ n_dev, const float *__restrict__ xyz,
                const __half *__restrict__ table, GridGeom geom, __half *__restrict__ out, float *__restrict__ dy_dx) {
    arc_pdl_sync();
    extern __shared__ __half tile[];   // [32][L*F + 2]
    if (n_dev) n = min(n, (uint32_t)max(*n_dev, 0));
    const uint32_t L = geom.n_levels;
    const uint32_t row = L * F;
    const uint32_t pitch = row + 2;
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    for (uint32_t base = blockIdx.x * kTileSamples; base < n; base += gridDim.x * kTileSamples) {
    const uint32_t i = base + lane;
    const bool valid = i < n;
    float x = 0.f, y = 0.f, z = 0.f;
    if (valid) { x = __ldg(xyz + 3 * (size_t)i); y = __ldg(xyz + 3 * (size_t)i + 1); z = __ldg(xyz + 3 * (size_t)i + 2); }

    for (uint32_t level = w; level < L; level += kFwdWarps) {
        const GridLevel g = geom.lv[level];
        const __half *lt = table + (size_t)g.offset * F;
        const Cell c = locate(g, x, y, z);
        float v[8][F];
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint32_t idx = entry_index(g, c.gx + (k & 1), c.gy + ((k >> 1) & 1), c.gz + ((k >> 2) & 1));
            load_entry<F>(lt, idx, v[k]);
        }
        const float wx[2] = {1.0f - c.fx, c.fx}, wy[2] = {1.0f - c.fy, c.fy}, wz[2] = {1.0f - c.fz, c.fz};
        float acc[F];
        #pragma unroll
        for (int f = 0; f < F; ++f) acc[f] = 0.f;
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float wgt = (wx[k & 1] * wy[(k >> 1) & 1]) * wz[(k >> 2) & 1];
            #pragma unroll
            for (int f = 0; f < F; ++f) acc[f] = fmaf(wgt, v[k][f], acc[f]);
        }
        #pragma unroll
        for (int f = 0; f < F; ++f) tile[lane * pitch + level * F + f] = __float2half_rn(acc[f]);

        if (DYDX && valid) {
            // d/dx: +-1 along the differentiated axis, bilinear weights on the other two, times scale
            float *dst = dy_dx + ((size_t)i * row + level * F) * 3;
            #pragma unroll
            for (int f = 0; f < F; ++f) {
                float gx = 0.f, gy = 0.f, gz = 0.f;
                #pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int cx = k & 1, cy = (k >> 1) & 1, cz = (k >> 2) & 1;
                    gx += (cx ? 1.f : -1.f) * (wy[cy] * wz[cz]) * v[k][f];
                    gy += (cy ? 1.f : -1.f) * (wx[cx] * wz[cz]) * v[k][f];
                    gz += (cz ? 1.f : -1.f) * (wx[cx] * wy[cy]) * v[k][f];
                }
                dst[3 * f] = g.scale * gx; dst[3 * f + 1] = g.scale * gy; dst[3 * f + 2] = g.scale * gz;
            }
        }
    }
    __syncthreads();
    // coalesced store: the tile is 32 rows of `row` halves; rows are contiguous in `out`
    const uint32_t n_rows = min((uint32_t)kTileSamples, n - base);
    if ((row & 1u) == 0) {
        const uint32_t words_per_row = row >> 1;
        __half2 *o2 = reinterpret_cast<__half2 *>(out + (size_t)base * row);
        for (uint32_t k = threadIdx.x; k < n_rows * words_per_row; k += blockDim.x) {
            const uint32_t r = k / words_per_row, cidx = k - r * words_per_row;
            o2[k] = *reinterpret_cast<const __half2 *>(&tile[r * pitch + 2 * cidx]);
        }
    } else {
        for (uint32_t k = threadIdx.x; k < n_rows * row; k += blockDim.x) {
            const uint32_t r = k / row, cidx = k - r * row;
            out[(size_t)base * row + k] = tile[r * pitch + cidx];
        }
    }
    __syncthreads();   // the tile is reused by the next iteration
    }
}

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }

template <int F>
__device__ __forceinline__ void add_entry(float *__restrict__ level_grad, uint32_t idx, const float (&g)[F], float wgt) {
    if (F == 2) {
        atomicAdd(reinterpret_cast<float2 *>(level_grad) + idx, make_float2(g[0] * wgt, g[1] * wgt));
    } else if (F == 4) {
        atomicAdd(reinterpret_cast<float4 *>(level_grad) + idx, make_float4(g[0] * wgt, g[1] * wgt, g[2] * wgt, g[3] * wgt));
    } else {
        #pragma unroll
        for (int f = 0; f < F; ++f) atomicAdd(level_grad + (size_t)idx * F + f, g[f] * wgt);
    }
}

// Table gradient.  Same tiling as the forward; dL/dy (N, L*F) of type T is staged through shared memory
// (coalesced), every (sample, level) then scatters 8 weighted copies with fp32 vector atomics
// (red.global.add.v2.f32 for F=2).  `gscale` multiplies dL/dy on the fly (undo of a loss scale).
// SECOND: instead of the interpolation weights use sum_d v_d * scale * d(w)/d(f_d) with v = dL/d(dL/dx)
// (N,3): the table term of the double backward (reference: kernel_grid_backward_input_backward_grid).
template <int F, typename T, bool SECOND>
__global__ void __launch_bounds__(kTileSamples * kFwdWarps)
grid_bwd_kernel(uint32_t n, const int32_t *__restrict__ n_dev, const float *__restrict__ xyz,
                const T *__restrict__ dL_dy, GridGeom geom, float *__restrict__ grad_table, float gscale,
                const float *__restrict__ v_in) {
    arc_pdl_sync();
    extern __shared__ float gtile[];   // [32][L*F + 1]
    if (n_dev) n = min(n, (uint32_t)max(*n_dev, 0));
    const uint32_t L = geom.n_levels;
    const uint32_t row = L * F;
    const uint32_t pitch = row + 1;
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    for (uint32_t base = blockIdx.x * kTileSamples; base < n; base += gridDim.x * kTileSamples) {
    const uint32_t i = base + lane;
    const bool valid = i < n;
    const uint32_t n_rows = min((uint32_t)kTileSamples, n - base);
    for (uint32_t k = threadIdx.x; k < n_rows * row; k += blockDim.x) {
        const uint32_t r = k / row, cidx = k - r * row;
        gtile[r * pitch + cidx] = to_f32<T>(dL_dy[(size_t)base * row + k]) * gscale;
    }
    float x = 0.f, y = 0.f, z = 0.f, vx = 0.f, vy = 0.f, vz = 0.f;
    if (valid) {
        x = __ldg(xyz + 3 * (size_t)i); y = __ldg(xyz + 3 * (size_t)i + 1); z = __ldg(xyz + 3 * (size_t)i + 2);
        if (SECOND) { vx = __ldg(v_in + 3 * (size_t)i); vy = __ldg(v_in + 3 * (size_t)i + 1); vz = __ldg(v_in + 3 * (size_t)i + 2); }
    }
    __syncthreads();
    for (uint32_t level = w; valid && level < L; level += kFwdWarps) {
        const GridLevel g = geom.lv[level];
        float *lg = grad_table + (size_t)g.offset * F;
        const Cell c = locate(g, x, y, z);
        float gr[F];
        bool any = false;
        #pragma unroll
        for (int f = 0; f < F; ++f) { gr[f] = gtile[lane * pitch + level * F + f]; any |= (gr[f] != 0.f); }
        if (!any) continue;
        const float wx[2] = {1.0f - c.fx, c.fx}, wy[2] = {1.0f - c.fy, c.fy}, wz[2] = {1.0f - c.fz, c.fz};
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int cx = k & 1, cy = (k >> 1) & 1, cz = (k >> 2) & 1;
            float wgt;
            if (SECOND) {
                wgt = g.scale * (vx * (cx ? 1.f : -1.f) * (wy[cy] * wz[cz]) + vy * (cy ? 1.f : -1.f) * (wx[cx] * wz[cz]) +
                                 vz * (cz ? 1.f : -1.f) * (wx[cx] * wy[cy]));
            } else {
                wgt = (wx[cx] * wy[cy]) * wz[cz];
            }
            const uint32_t idx = entry_index(g, c.gx + cx, c.gy + cy, c.gz + cz);
            add_entry<F>(lg, idx, gr, wgt);
        }
    }
    __syncthreads();   // the staged gradients are reused by the next iteration
    }
}

// dL/dx (N,3) = sum_k dL/dy_k * dy_k/dx     (reference: kernel_grid_backward_input, grid.h:549-576)
template <typename T>
__global__ void grid_bwd_input_kernel(uint32_t n, uint32_t row, const T *__restrict__ dL_dy,
                                      const float *__restrict__ dy_dx, float gscale, float *__restrict__ dL_dx) {
    arc_pdl_sync();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float ax = 0.f, ay = 0.f, az = 0.f;
    const T *g = dL_dy + (size_t)i * row;
    const float *d = dy_dx + (size_t)i * row * 3;
    for (uint32_t k = 0; k < row; ++k) {
        const float gk = to_f32<T>(g[k]);
        ax = fmaf(gk, d[3 * k], ax); ay = fmaf(gk, d[3 * k + 1], ay); az = fmaf(gk, d[3 * k + 2], az);
    }
    dL_dx[3 * (size_t)i] = ax * gscale; dL_dx[3 * (size_t)i + 1] = ay * gscale; dL_dx[3 * (size_t)i + 2] = az * gscale;
}

// Double backward, the two non-table terms (reference: kernel_grid_backward_input_backward_input,
// grid.h:701-868, and ..._backward_dLdoutput, grid.h:870-895).  With v = dL2/d(dL/dx) (N,3):
//   d_dLdy[k] = sum_d v_d * dy_k/dx_d
//   d_x[e]    = scale^2 * sum_k dL/dy_k * sum_{d != e} v_d * sum_c s_d s_e w_c^{(other axis)} T_c,k
// One thread per (sample, level); d_x is accumulated over levels with atomics into a zeroed (N,3) buffer.
template <int F, typename T>
__global__ void grid_bwd_bwd_input_kernel(uint32_t n, const float *__restrict__ xyz, const __half *__restrict__ table,
                                          GridGeom geom, const T *__restrict__ dL_dy, float gscale,
                                          const float *__restrict__ v_in, float *__restrict__ d_dLdy,
                                          float *__restrict__ d_x) {
    arc_pdl_sync();
    const uint32_t L = geom.n_levels;
    const uint32_t row = L * F;
    const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t i = gid / L, level = gid - i * L;
    if (i >= n) return;
    const GridLevel g = geom.lv[level];
    const __half *lt = table + (size_t)g.offset * F;
    const float x = __ldg(xyz + 3 * (size_t)i), y = __ldg(xyz + 3 * (size_t)i + 1), z = __ldg(xyz + 3 * (size_t)i + 2);
    const float vx = __ldg(v_in + 3 * (size_t)i), vy = __ldg(v_in + 3 * (size_t)i + 1), vz = __ldg(v_in + 3 * (size_t)i + 2);
    const Cell c = locate(g, x, y, z);
    float v[8][F];
    #pragma unroll
    for (int k = 0; k < 8; ++k) {
        const uint32_t idx = entry_index(g, c.gx + (k & 1), c.gy + ((k >> 1) & 1), c.gz + ((k >> 2) & 1));
        load_entry<F>(lt, idx, v[k]);
    }
    const float wx[2] = {1.0f - c.fx, c.fx}, wy[2] = {1.0f - c.fy, c.fy}, wz[2] = {1.0f - c.fz, c.fz};
    float ex = 0.f, ey = 0.f, ez = 0.f;
    #pragma unroll
    for (int f = 0; f < F; ++f) {
        float gx = 0.f, gy = 0.f, gz = 0.f;      // first derivatives
        float hxy = 0.f, hxz = 0.f, hyz = 0.f;   // mixed second derivatives (pure ones vanish)
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int cx = k & 1, cy = (k >> 1) & 1, cz = (k >> 2) & 1;
            const float sx = cx ? 1.f : -1.f, sy = cy ? 1.f : -1.f, sz = cz ? 1.f : -1.f;
            gx += sx * (wy[cy] * wz[cz]) * v[k][f];
            gy += sy * (wx[cx] * wz[cz]) * v[k][f];
            gz += sz * (wx[cx] * wy[cy]) * v[k][f];
            hxy += sx * sy * wz[cz] * v[k][f];
            hxz += sx * sz * wy[cy] * v[k][f];
            hyz += sy * sz * wx[cx] * v[k][f];
        }
        const size_t o = (size_t)i * row + level * F + f;
        if (d_dLdy) d_dLdy[o] = g.scale * (vx * gx + vy * gy + vz * gz);
        if (d_x) {
            const float gk = to_f32<T>(dL_dy[o]) * gscale;
            const float s2 = g.scale * g.scale;
            ex += gk * s2 * (vy * hxy + vz * hxz);
            ey += gk * s2 * (vx * hxy + vz * hyz);
            ez += gk * s2 * (vx * hxz + vy * hyz);
        }
    }
    if (d_x) {
        atomicAdd(d_x + 3 * (size_t)i, ex);
        atomicAdd(d_x + 3 * (size_t)i + 1, ey);
        atomicAdd(d_x + 3 * (size_t)i + 2, ez);
    }
}

__global__ void cast_f32_f16_kernel(size_t n, const float *__restrict__ src, __half *__restrict__ dst) {
    arc_pdl_sync();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t n4 = n / 4;
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n4; k += stride) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(src) + k);
        __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t *>(&a);
        pk.y = *reinterpret_cast<uint32_t *>(&b);
        reinterpret_cast<uint2 *>(dst)[k] = pk;
    }
    for (size_t k = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) dst[k] = __float2half_rn(src[k]);
}

unsigned tile_blocks(int64_t n) {
    long long want = (n + kTileSamples - 1) / kTileSamples;
    const long long cap = (long long)arc_sm_count() * 16;
    return (unsigned)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace

extern "C" {

// Host-side level geometry, following the reference's constructor arithmetic (grid.h:195-204,982-1012).
// offsets[n_levels+1] (entries, not floats), scales[n_levels], resolutions[n_levels] are HOST arrays.
// Returns the total number of table entries (>0) or a negative error.
int64_t arcb200_grid_geometry(int n_levels, int log2_hashmap_size, int base_resolution, float per_level_scale,
                              uint32_t *offsets, float *scales, uint32_t *resolutions) {
    if (n_levels < 1 || n_levels > kMaxLevels || !offsets || !scales || !resolutions || log2_hashmap_size < 1 ||
        log2_hashmap_size > 30 || base_resolution < 1) {
        arcb200_set_error("grid_geometry: invalid argument (1 <= n_levels <= %d)", kMaxLevels);
        return ARCB200_ERR_ARG;
    }
    const float log2_pls = log2f(per_level_scale);
    uint32_t offset = 0;
    for (int l = 0; l < n_levels; ++l) {
        const float scale = exp2f((float)l * log2_pls) * (float)base_resolution - 1.0f;
        const uint32_t res = (uint32_t)ceilf(scale) + 1u;
        const uint32_t max_params = 0xFFFFFFFFu / 2;
        uint32_t params = powf((float)res, 3.0f) > (float)max_params ? max_params : res * res * res;
        params = (params + 7u) / 8u * 8u;
        const uint32_t cap = 1u << log2_hashmap_size;
        if (params > cap) params = cap;
        offsets[l] = offset;
        scales[l] = scale;
        resolutions[l] = res;
        offset += params;
    }
    offsets[n_levels] = offset;
    return (int64_t)offset;
}

// fp32 master parameters -> fp16 working copy (the reference casts its flat `params` tensor on every
// forward, tinycudann/modules.py:156; callers here do it only when the parameters changed).
int arcb200_cast_f32_to_f16(const float *src, void *dst_half, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0, "cast_f32_to_f16: negative n");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(src && dst_half, "cast_f32_to_f16: null pointer");
    ARC_REQUIRE(((uintptr_t)src & 15u) == 0 && ((uintptr_t)dst_half & 7u) == 0, "cast_f32_to_f16: misaligned buffer");
    long long blocks = (n / 4 + 255) / 256;
    const long long cap = (long long)arc_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    arc_launch(cast_f32_f16_kernel, (unsigned)blocks, 256, 0, (cudaStream_t)stream, (size_t)n, src, (__half *)dst_half);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

#define GRID_DISPATCH_F(F_, ...)                                             \
    switch (F_) {                                                            \
        case 1: { constexpr int F = 1; __VA_ARGS__; break; }                 \
        case 2: { constexpr int F = 2; __VA_ARGS__; break; }                 \
        case 4: { constexpr int F = 4; __VA_ARGS__; break; }                 \
        case 8: { constexpr int F = 8; __VA_ARGS__; break; }                 \
        default: arcb200_set_error("hash grid: n_features_per_level must be 1, 2, 4 or 8"); return ARCB200_ERR_ARG; \
    }

// Forward.  xyz (n,3) fp32 in [0,1]; table: fp16, flat [level][entry][feature]; out (n, L*F) fp16 row-major;
// dy_dx (n, L*F, 3) fp32 or NULL.  offsets/scales/resolutions: HOST arrays from arcb200_grid_geometry.
// n_dev (device int32, optional): only min(n, *n_dev) rows are processed - the row count of a packed sample list
// can stay on the device (no host synchronisation between the sampler and the encoder).
// Replaces tcnn Module.fwd for a HashGrid encoding (bindings/torch/tinycudann/bindings.cpp:78-109 ->
// GridEncodingTemplated::forward_impl, grid.h:1034-1105).
int arcb200_grid_fwd(const float *xyz, const void *table_half, int n_levels, int n_features, const uint32_t *offsets,
                     const float *scales, const uint32_t *resolutions, void *out_half, float *dy_dx, int64_t n,
                     const int32_t *n_dev, void *stream) {
    ARC_REQUIRE(n >= 0 && n < (1ll << 31) - 64, "grid_fwd: n out of range");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(xyz && table_half && out_half && offsets && scales && resolutions, "grid_fwd: null pointer");
    GridGeom geom;
    ARC_REQUIRE(pack_geom(geom, n_levels, offsets, scales, resolutions) == 0, "grid_fwd: bad level geometry");
    const unsigned blocks = tile_blocks(n);
    const size_t smem = (size_t)kTileSamples * (n_levels * n_features + 2) * sizeof(__half);
    cudaStream_t st = (cudaStream_t)stream;
    GRID_DISPATCH_F(n_features,
        if (dy_dx) arc_launch(grid_fwd_kernel<F, true>, blocks, kTileSamples * kFwdWarps, smem, st, 
            (uint32_t)n, n_dev, xyz, (const __half *)table_half, geom, (__half *)out_half, dy_dx);
        else arc_launch(grid_fwd_kernel<F, false>, blocks, kTileSamples * kFwdWarps, smem, st, 
            (uint32_t)n, n_dev, xyz, (const __half *)table_half, geom, (__half *)out_half, nullptr));
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Table gradient: grad_table (fp32, same flat layout as the table) += sum over samples.  The caller zeroes
// grad_table.  dL_dy (n, L*F) is fp16 (dy_is_half != 0) or fp32, multiplied by gscale on the fly.
// Replaces the grid part of tcnn Module.bwd (bindings.cpp:111-170 -> kernel_grid_backward, grid.h:394-517).
int arcb200_grid_bwd(const float *xyz, const void *dL_dy, int dy_is_half, float gscale, int n_levels, int n_features,
                     const uint32_t *offsets, const float *scales, const uint32_t *resolutions, float *grad_table,
                     int64_t n, const int32_t *n_dev, void *stream) {
    ARC_REQUIRE(n >= 0 && n < (1ll << 31) - 64, "grid_bwd: n out of range");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(xyz && dL_dy && grad_table && offsets && scales && resolutions, "grid_bwd: null pointer");
    GridGeom geom;
    ARC_REQUIRE(pack_geom(geom, n_levels, offsets, scales, resolutions) == 0, "grid_bwd: bad level geometry");
    const unsigned blocks = tile_blocks(n);
    const size_t smem = (size_t)kTileSamples * (n_levels * n_features + 1) * sizeof(float);
    cudaStream_t st = (cudaStream_t)stream;
    GRID_DISPATCH_F(n_features,
        if (dy_is_half) arc_launch(grid_bwd_kernel<F, __half, false>, blocks, kTileSamples * kFwdWarps, smem, st, 
            (uint32_t)n, n_dev, xyz, (const __half *)dL_dy, geom, grad_table, gscale, nullptr);
        else arc_launch(grid_bwd_kernel<F, float, false>, blocks, kTileSamples * kFwdWarps, smem, st, 
            (uint32_t)n, n_dev, xyz, (const float *)dL_dy, geom, grad_table, gscale, nullptr));
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// dL_dx (n,3) = dL_dy . dy_dx   (reference: kernel_grid_backward_input, grid.h:549-576)
int arcb200_grid_bwd_input(const void *dL_dy, int dy_is_half, float gscale, const float *dy_dx, int n_levels,
                           int n_features, float *dL_dx, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0 && n < (1ll << 31), "grid_bwd_input: n out of range");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(dL_dy && dy_dx && dL_dx && n_levels > 0 && n_features > 0, "grid_bwd_input: null/invalid argument");
    const uint32_t row = (uint32_t)(n_levels * n_features);
    if (dy_is_half) arc_launch(grid_bwd_input_kernel<__half>, arc_blocks(n, 128), 128, 0, (cudaStream_t)stream, 
        (uint32_t)n, row, (const __half *)dL_dy, dy_dx, gscale, dL_dx);
    else arc_launch(grid_bwd_input_kernel<float>, arc_blocks(n, 128), 128, 0, (cudaStream_t)stream, 
        (uint32_t)n, row, (const float *)dL_dy, dy_dx, gscale, dL_dx);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Double backward of the input gradient (eikonal loss).  v = dL2/d(dL/dx) (n,3).  Outputs (each optional):
//   grad_table  += table term (caller zeroes / accumulates)      [grid.h:578-699]
//   d_dLdy (n, L*F) fp32                                         [grid.h:870-895]
//   d_x (n,3) fp32, ZEROED by this call then accumulated         [grid.h:701-868]
// Replaces tcnn Module.bwd_bwd_input for a HashGrid encoding (bindings.cpp:172-241).
int arcb200_grid_bwd_bwd_input(const float *xyz, const void *table_half, const void *dL_dy, int dy_is_half,
                               float gscale, const float *v, int n_levels, int n_features, const uint32_t *offsets,
                               const float *scales, const uint32_t *resolutions, float *grad_table, float *d_dLdy,
                               float *d_x, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0 && n * (int64_t)(n_levels > 0 ? n_levels : 1) < (1ll << 31), "grid_bwd_bwd_input: n out of range");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(xyz && table_half && dL_dy && v && offsets && scales && resolutions, "grid_bwd_bwd_input: null pointer");
    GridGeom geom;
    ARC_REQUIRE(pack_geom(geom, n_levels, offsets, scales, resolutions) == 0, "grid_bwd_bwd_input: bad level geometry");
    cudaStream_t st = (cudaStream_t)stream;
    if (grad_table) {
        const unsigned blocks = tile_blocks(n);
        const size_t smem = (size_t)kTileSamples * (n_levels * n_features + 1) * sizeof(float);
        GRID_DISPATCH_F(n_features,
            if (dy_is_half) arc_launch(grid_bwd_kernel<F, __half, true>, blocks, kTileSamples * kFwdWarps, smem, st, 
                (uint32_t)n, nullptr, xyz, (const __half *)dL_dy, geom, grad_table, gscale, v);
            else arc_launch(grid_bwd_kernel<F, float, true>, blocks, kTileSamples * kFwdWarps, smem, st, 
                (uint32_t)n, nullptr, xyz, (const float *)dL_dy, geom, grad_table, gscale, v));
        ARC_LAUNCH_CHECK();
    }
    if (d_dLdy || d_x) {
        if (d_x) ARC_CUDA(cudaMemsetAsync(d_x, 0, (size_t)n * 3 * sizeof(float), st));
        const unsigned blocks = arc_blocks(n * n_levels, 128);
        GRID_DISPATCH_F(n_features,
            if (dy_is_half) arc_launch(grid_bwd_bwd_input_kernel<F, __half>, blocks, 128, 0, st, 
                (uint32_t)n, xyz, (const __half *)table_half, geom, (const __half *)dL_dy, gscale, v, d_dLdy, d_x);
            else arc_launch(grid_bwd_bwd_input_kernel<F, float>, blocks, 128, 0, st, 
                (uint32_t)n, xyz, (const __half *)table_half, geom, (const float *)dL_dy, gscale, v, d_dLdy, d_x));
        ARC_LAUNCH_CHECK();
    }
    return ARCB200_OK;
}

}  // extern "C"
