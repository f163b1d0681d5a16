// Occupancy-grid maintenance kernels (run every `epoch_optim` steps): candidate-cell sampling, max-splat,
// EMA decay, float grid -> bitfield, set-bit count, and the grouped max used by VolumeBound.optimize.
// Compiled with -fmad=false like march.cu (sample positions are compared bit-for-bit with the oracle).
//
// These are HBM-bound integer/byte passes over 128^3 cells (8 MB of floats, 256 KB of bits); each is a
// single coalesced sweep, grid sized from the SM count.
#include "common.cuh"
#include "pcg32.cuh"
#include "occ_math.cuh"

using namespace occ;

namespace {

// max over non-negative floats through their bit patterns (reference: atomicMax on float-as-uint,
// volume_func_kernel.cu:297-309 and bitfield_func_kernel.cu:217-234).
__global__ void group_max_kernel(uint32_t n, const float *__restrict__ vals, const int64_t *__restrict__ group,
                                 float *__restrict__ out) {
    arc_pdl_sync();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    atomicMax((uint32_t *)&out[group[i]], __float_as_uint(vals[i]));
}

__global__ void splat_max_kernel(uint32_t n, const uint32_t *__restrict__ idx, const float *__restrict__ vals,
                                 float *__restrict__ grid_tmp) {
    arc_pdl_sync();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    atomicMax((uint32_t *)&grid_tmp[idx[i]], __float_as_uint(vals[i]));
}

// Candidate cells: up to 10 LCG-hashed tries for a cell above `thresh`, then a jittered position in the
// cell (bitfield_func_kernel.cu:143-185; multivol_func_kernel.cu:177-240 when n_cascades > 0).
template <bool CASCADED>
__global__ void grid_samples_kernel(uint32_t n_elements, Pcg32 rng, uint32_t step, const float *__restrict__ grid,
                                    float *__restrict__ pos_out, uint32_t *__restrict__ idx_out, uint32_t n_grid,
                                    float thresh, const float *__restrict__ aabb, uint32_t n_cascades,
                                    int inclusive) {
    arc_pdl_sync();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_elements) return;
    rng.advance((uint64_t)(int64_t)(i * 4u));
    const uint32_t n3 = n_grid * n_grid * n_grid;
    uint32_t level = 0;
    if (CASCADED) {
        if (inclusive) level = (uint32_t)(rng.next_f32() * n_cascades) % n_cascades;
        else while (level == 0) level = (uint32_t)(rng.next_f32() * n_cascades) % n_cascades;
    }
    uint32_t idx = 0;
    for (uint32_t j = 0; j < 10; ++j) {
        idx = ((i + step * n_elements) * 56924617u + j * 19349663u + 96925573u) % n3;
        if (CASCADED) idx += (inclusive ? level : level - 1) * n3;
        if (grid[idx] > thresh) break;
    }
    const uint32_t c = idx % n3;
    const float x = (float)unspread3(c), y = (float)unspread3(c >> 1), z = (float)unspread3(c >> 2);
    const float u0 = rng.next_f32(), u1 = rng.next_f32(), u2 = rng.next_f32();
    const float fn = (float)n_grid;
    float px = (x + u0) / fn, py = (y + u1) / fn, pz = (z + u2) / fn;
    if (CASCADED) {
        const Box b = load_box(aabb);
        const float s = scalbnf(1.0f, (int)level);
        px = ((px - 0.5f) * (b.hi.x - b.lo.x)) * s + (b.lo.x + b.hi.x) / 2.0f;
        py = ((py - 0.5f) * (b.hi.y - b.lo.y)) * s + (b.lo.y + b.hi.y) / 2.0f;
        pz = ((pz - 0.5f) * (b.hi.z - b.lo.z)) * s + (b.lo.z + b.hi.z) / 2.0f;
    }
    pos_out[3 * (size_t)i] = px;
    pos_out[3 * (size_t)i + 1] = py;
    pos_out[3 * (size_t)i + 2] = pz;
    idx_out[i] = idx;
}

// grid = grid < 0 ? grid : max(grid*decay, tmp)   (bitfield_func_kernel.cu:260-275)
__global__ void ema_kernel(uint32_t n, float decay, float *__restrict__ grid, const float *__restrict__ tmp) {
    arc_pdl_sync();
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float prev = grid[i];
        grid[i] = prev < 0.f ? prev : fmaxf(prev * decay, tmp[i]);
    }
}

// 8 cells -> 1 byte, threshold min(opa_thres, mean) (bitfield_func_kernel.cu:302-324).  Two float4 loads
// per byte; `mean` may come from device memory (no host round trip for the mean).
__global__ void to_bitfield_kernel(uint32_t n_bytes, const float *__restrict__ grid, uint8_t *__restrict__ bits,
                                   float mean_host, const float *__restrict__ mean_dev, float opa_thres) {
    arc_pdl_sync();
    const float mean = mean_dev ? __ldg(mean_dev) : mean_host;
    const float thresh = opa_thres < mean ? opa_thres : mean;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_bytes; i += gridDim.x * blockDim.x) {
        const float4 a = __ldg((const float4 *)grid + 2 * (size_t)i);
        const float4 b = __ldg((const float4 *)grid + 2 * (size_t)i + 1);
        uint32_t v = 0;
        v |= a.x > thresh ? 1u : 0u;  v |= a.y > thresh ? 2u : 0u;
        v |= a.z > thresh ? 4u : 0u;  v |= a.w > thresh ? 8u : 0u;
        v |= b.x > thresh ? 16u : 0u; v |= b.y > thresh ? 32u : 0u;
        v |= b.z > thresh ? 64u : 0u; v |= b.w > thresh ? 128u : 0u;
        bits[i] = (uint8_t)v;
    }
}

// counter += number of set bits.  faithful != 0 reproduces the reference's `byte && (1 << j)` test
// (bitfield_func_kernel.cu:365), which adds 8 for every non-zero byte.
__global__ void count_bits_kernel(uint32_t n_bytes, const uint8_t *__restrict__ bits, float *__restrict__ counter,
                                  int faithful) {
    arc_pdl_sync();
    uint32_t c = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_bytes; i += gridDim.x * blockDim.x) {
        const uint32_t b = bits[i];
        c += faithful ? (b ? 8u : 0u) : (uint32_t)__popc(b);
    }
    #pragma unroll
    for (int s = 16; s > 0; s >>= 1) c += __shfl_xor_sync(ARC_FULL_MASK, c, s);
    __shared__ uint32_t warp_c[32];
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    if (lane == 0) warp_c[w] = c;
    __syncthreads();
    if (w == 0) {
        c = lane < (blockDim.x >> 5) ? warp_c[lane] : 0u;
        #pragma unroll
        for (int s = 16; s > 0; s >>= 1) c += __shfl_xor_sync(ARC_FULL_MASK, c, s);
        if (lane == 0 && c) atomicAdd(counter, (float)c);
    }
}

unsigned int sweep_blocks(long long n, int threads) {
    long long want = (n + threads - 1) / threads;
    long long cap = (long long)arc_sm_count() * 8;
    return (unsigned int)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace

extern "C" {

int arcb200_tensor_reduce_max(const float *full_tensor, const int64_t *group_idx, int n_group, float *uni_tensor,
                              int64_t n, void *stream) {
    (void)n_group;
    ARC_REQUIRE(n >= 0 && n < (1ll << 32), "tensor_reduce_max: bad n");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(full_tensor && group_idx && uni_tensor, "tensor_reduce_max: null pointer");
    arc_launch(group_max_kernel, arc_blocks(n, 256), 256, 0, (cudaStream_t)stream, (uint32_t)n, full_tensor, group_idx, uni_tensor);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_generate_grid_samples(const float *density_grid, int ema_step, int64_t n_elements, int n_grid,
                                  float thresh, float *positions, uint32_t *indices, uint64_t rng_state,
                                  uint64_t rng_inc, void *stream) {
    ARC_REQUIRE(n_elements >= 0 && n_elements < (1ll << 32), "generate_grid_samples: bad n_elements");
    if (n_elements == 0) return ARCB200_OK;
    ARC_REQUIRE(density_grid && positions && indices && n_grid > 0, "generate_grid_samples: null/invalid argument");
    Pcg32 rng;
    rng.state = rng_state;
    rng.inc = rng_inc;
    arc_launch(grid_samples_kernel<false>, arc_blocks(n_elements, 256), 256, 0, (cudaStream_t)stream, 
        (uint32_t)n_elements, rng, (uint32_t)ema_step, density_grid, positions, indices, (uint32_t)n_grid, thresh,
        nullptr, 0u, 0);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_generate_grid_samples_multivol(const float *density_grid, int ema_step, int64_t n_elements,
                                           const float *aabb, int n_cascade, int n_grid, float thresh,
                                           int inclusive, float *positions, uint32_t *indices,
                                           uint64_t rng_state, uint64_t rng_inc, void *stream) {
    ARC_REQUIRE(n_elements >= 0 && n_elements < (1ll << 32), "generate_grid_samples_multivol: bad n_elements");
    if (n_elements == 0) return ARCB200_OK;
    ARC_REQUIRE(density_grid && positions && indices && aabb && n_grid > 0, "generate_grid_samples_multivol: null/invalid argument");
    ARC_REQUIRE(n_cascade > (inclusive ? 0 : 1), "generate_grid_samples_multivol: need n_cascade >= 1 (>= 2 when not inclusive)");
    Pcg32 rng;
    rng.state = rng_state;
    rng.inc = rng_inc;
    arc_launch(grid_samples_kernel<true>, arc_blocks(n_elements, 256), 256, 0, (cudaStream_t)stream, 
        (uint32_t)n_elements, rng, (uint32_t)ema_step, density_grid, positions, indices, (uint32_t)n_grid, thresh,
        aabb, (uint32_t)n_cascade, inclusive);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_splat_grid_samples(const float *density, const uint32_t *indices, int64_t n_samples,
                               float *density_grid_tmp, void *stream) {
    ARC_REQUIRE(n_samples >= 0 && n_samples < (1ll << 32), "splat_grid_samples: bad n");
    if (n_samples == 0) return ARCB200_OK;
    ARC_REQUIRE(density && indices && density_grid_tmp, "splat_grid_samples: null pointer");
    arc_launch(splat_max_kernel, arc_blocks(n_samples, 256), 256, 0, (cudaStream_t)stream, (uint32_t)n_samples, indices,
                                                                                  density, density_grid_tmp);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_ema_grid_samples_nerf(const float *density_grid_tmp, int64_t n_elements, float decay,
                                  float *density_grid, void *stream) {
    ARC_REQUIRE(n_elements >= 0 && n_elements < (1ll << 32), "ema_grid_samples_nerf: bad n");
    if (n_elements == 0) return ARCB200_OK;
    ARC_REQUIRE(density_grid_tmp && density_grid, "ema_grid_samples_nerf: null pointer");
    arc_launch(ema_kernel, sweep_blocks(n_elements, 256), 256, 0, (cudaStream_t)stream, (uint32_t)n_elements, decay,
                                                                               density_grid, density_grid_tmp);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// n_cells = number of float cells covered (n_grid^3, times the cascade count for the multi-volume form).
// mean_dev, when not null, is read on the device instead of mean_host.
int arcb200_update_bitfield(const float *density_grid, float mean_host, const float *mean_dev, uint8_t *bitfield,
                            float opa_thres, int64_t n_cells, void *stream) {
    ARC_REQUIRE(n_cells >= 0 && n_cells % 8 == 0 && n_cells < (1ll << 35), "update_bitfield: n_cells must be a multiple of 8");
    if (n_cells == 0) return ARCB200_OK;
    ARC_REQUIRE(density_grid && bitfield, "update_bitfield: null pointer");
    ARC_REQUIRE(((uintptr_t)density_grid & 15u) == 0, "update_bitfield: density grid must be 16-byte aligned");
    arc_launch(to_bitfield_kernel, sweep_blocks(n_cells / 8, 256), 256, 0, (cudaStream_t)stream, 
        (uint32_t)(n_cells / 8), density_grid, bitfield, mean_host, mean_dev, opa_thres);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_count_bitfield(const uint8_t *bitfield, float *counter, int64_t n_cells, int faithful, void *stream) {
    ARC_REQUIRE(n_cells >= 0 && n_cells % 8 == 0 && n_cells < (1ll << 35), "count_bitfield: n_cells must be a multiple of 8");
    if (n_cells == 0) return ARCB200_OK;
    ARC_REQUIRE(bitfield && counter, "count_bitfield: null pointer");
    arc_launch(count_bits_kernel, sweep_blocks(n_cells / 8, 256), 256, 0, (cudaStream_t)stream, (uint32_t)(n_cells / 8), bitfield,
                                                                                       counter, faithful);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
