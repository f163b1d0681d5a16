// Occupancy refresh of the volume bound on the device.
//
// Reference: VolumeBound.optimize (arcnerf/models/base_modules/obj_bound/volume_bound.py:160-212) with
// Volume.update_opafield_by_voxel_idx / update_bitfield_by_opafield (arcnerf/geometry/volume.py:983-1018).  The
// reference builds the refresh from torch.randperm(n^3), torch.where(bitfield), torch.unique(dim=0) and indexed
// scatters ("where the time goes").  Here the same rule runs as a handful of streaming kernels with no
// host synchronisation:
//   1. candidates: n_uniform DISTINCT voxels drawn by a keyed Feistel permutation of the flat voxel index (sampling
//      without replacement, like randperm(n^3)[:n_uniform]) + the first n_occ occupied voxels in flat order (compaction
//      by block counts, one-block scan, ordered write; exactly occ_voxel_idx[:n_sample]); one jittered point per
//      candidate (pcg32), unused candidate slots are marked -1;
//   2. (caller) opacity of the points = density * dt through the field kernels;
//   3. per-voxel max by atomicMax into a scratch field preset to -1, then for every voxel that received a sample
//      opa = opa >= 0 ? max(opa * ema, new) : opa  (voxels without a sample are left untouched, unlike the instant-ngp
//      rule in occupancy.cu) while the same pass sums max(opa, 0);
//   4. occupancy byte = opa >= min(mean, threshold).
#include "common.cuh"
#include "pcg32.cuh"

namespace {

constexpr int kOccBlock = 1024;   // voxels per block of the compaction kernels (256 threads x 4)

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
// 4-round balanced Feistel network on 2*half_bits bits: a bijection of [0, 2^(2*half_bits))
__device__ __forceinline__ uint32_t feistel(uint32_t v, int half_bits, uint32_t key) {
    const uint32_t mask = (1u << half_bits) - 1u;
    uint32_t l = v >> half_bits, r = v & mask;
    #pragma unroll
    for (int round = 0; round < 4; ++round) {
        const uint32_t f = mix32(r ^ (key + 0x9e3779b9u * (uint32_t)(round + 1))) & mask;
        const uint32_t t = l ^ f;
        l = r; r = t;
    }
    return (l << half_bits) | r;
}

__global__ void occ_count_kernel(uint32_t n_voxels, const uint8_t *__restrict__ volume, uint32_t *__restrict__ block_counts) {
    arc_pdl_sync();
    __shared__ uint32_t warp_sum[8];
    const uint32_t base = blockIdx.x * kOccBlock + threadIdx.x * 4;
    uint32_t c = 0;
    #pragma unroll
    for (int k = 0; k < 4; ++k) c += (base + k < n_voxels && volume[base + k]) ? 1u : 0u;
    #pragma unroll
    for (int s = 16; s > 0; s >>= 1) c += __shfl_xor_sync(ARC_FULL_MASK, c, s);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < 8; ++w) t += warp_sum[w];
        block_counts[blockIdx.x] = t;
    }
}

// exclusive scan of block_counts in place (one CTA); total -> n_occ[0]
__global__ void occ_scan_kernel(uint32_t n_blocks, uint32_t *__restrict__ block_counts, uint32_t *__restrict__ n_occ) {
    arc_pdl_sync();
    __shared__ uint32_t part[1024];
    const uint32_t per = (n_blocks + blockDim.x - 1) / blockDim.x;
    const uint32_t lo = threadIdx.x * per, hi = min(n_blocks, lo + per);
    uint32_t s = 0;
    for (uint32_t i = lo; i < hi; ++i) s += block_counts[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (uint32_t off = 1; off < blockDim.x; off <<= 1) {
        const uint32_t v = threadIdx.x >= off ? part[threadIdx.x - off] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t run = part[threadIdx.x] - s;
    for (uint32_t i = lo; i < hi; ++i) { const uint32_t c = block_counts[i]; block_counts[i] = run; run += c; }
    if (threadIdx.x == blockDim.x - 1) n_occ[0] = part[threadIdx.x];
}

// candidate i < n_uniform: permuted voxel; candidate n_uniform + j: j-th occupied voxel (flat order) or -1
__global__ void refresh_candidates_kernel(uint32_t n_voxels, uint32_t n_grid, uint32_t n_uniform, uint32_t n_occ_max,
                                          const uint8_t *__restrict__ volume, const uint32_t *__restrict__ block_offsets,
                                          uint32_t key, int half_bits, Pcg32 rng, const float *__restrict__ aabb,
                                          int32_t *__restrict__ idx_out, float *__restrict__ pts_out) {
    arc_pdl_sync();
    __shared__ uint32_t warp_off[8];
    const uint32_t n_blocks_occ = (n_voxels + kOccBlock - 1) / kOccBlock;
    if (blockIdx.x < n_blocks_occ) {
        // ordered write of this block's occupied voxels
        const uint32_t base = blockIdx.x * kOccBlock + threadIdx.x * 4;
        uint32_t flags = 0;
        #pragma unroll
        for (int k = 0; k < 4; ++k) flags |= (base + k < n_voxels && volume[base + k]) ? (1u << k) : 0u;
        const uint32_t mine = __popc(flags);
        uint32_t incl = mine;
        const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
        #pragma unroll
        for (int s = 1; s < 32; s <<= 1) { const uint32_t v = __shfl_up_sync(ARC_FULL_MASK, incl, s); if (lane >= (uint32_t)s) incl += v; }
        if (lane == 31) warp_off[warp] = incl;
        __syncthreads();
        uint32_t before = block_offsets[blockIdx.x] + incl - mine;
        for (uint32_t w = 0; w < warp; ++w) before += warp_off[w];
        #pragma unroll
        for (int k = 0; k < 4; ++k)
            if (flags & (1u << k)) { if (before < n_occ_max) idx_out[n_uniform + before] = (int32_t)(base + k); ++before; }
    } else {
        const uint32_t i = (blockIdx.x - n_blocks_occ) * blockDim.x + threadIdx.x;
        if (i < n_uniform) {
            uint32_t v = i;
            do { v = feistel(v, half_bits, key); } while (v >= n_voxels);      // cycle walking keeps it a bijection of [0, n_voxels)
            idx_out[i] = (int32_t)v;
        }
    }
}

// jittered point inside every candidate voxel (flat index x*n*n + y*n + z); slot without a candidate -> box centre
__global__ void refresh_points_kernel(uint32_t n_cand, uint32_t n_grid, Pcg32 rng, const float *__restrict__ aabb,
                                      const int32_t *__restrict__ idx, float *__restrict__ pts) {
    arc_pdl_sync();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_cand) return;
    const int32_t v = idx[i];
    float p[3];
    if (v < 0) {
        #pragma unroll
        for (int a = 0; a < 3; ++a) p[a] = 0.5f * (aabb[a] + aabb[3 + a]);
    } else {
        // per-candidate generator state from a 64-bit mix of (call state, candidate) - a jump-ahead of i*3 draws would cost
        // a 64-step multiply chain per thread, and the jitter only has to be uniform inside the voxel
        uint64_t zz = rng.state + 0x9e3779b97f4a7c15ull * (uint64_t)(i + 1);
        zz = (zz ^ (zz >> 30)) * 0xbf58476d1ce4e5b9ull;
        zz = (zz ^ (zz >> 27)) * 0x94d049bb133111ebull;
        rng.state = zz ^ (zz >> 31);
        const uint32_t c[3] = {(uint32_t)v / (n_grid * n_grid), ((uint32_t)v / n_grid) % n_grid, (uint32_t)v % n_grid};
        #pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float vs = (aabb[3 + a] - aabb[a]) / (float)n_grid;
            p[a] = aabb[a] + ((float)c[a] + rng.next_f32()) * vs;          // centre + (u - 0.5) * voxel size
        }
    }
    pts[3 * i] = p[0]; pts[3 * i + 1] = p[1]; pts[3 * i + 2] = p[2];
}

__global__ void refresh_splat_kernel(uint32_t n, uint32_t n_voxels, const int32_t *__restrict__ idx,
                                     const float *__restrict__ opacity, float *__restrict__ tmp) {
    arc_pdl_sync();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t v = idx[i];
    if (v < 0 || (uint32_t)v >= n_voxels) return;       // unused slot (or an index outside the volume)
    // tmp is preset to -1 (sign bit set): for non-negative floats the signed-int order is the float order
    atomicMax(reinterpret_cast<int *>(tmp) + v, __float_as_int(fmaxf(opacity[i], 0.0f)));
}

__global__ void refresh_fill_kernel(uint32_t n, float value, float *__restrict__ tmp, float *__restrict__ sum) {
    arc_pdl_sync();
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) tmp[i] = value;
    if (blockIdx.x == 0 && threadIdx.x == 0 && sum) sum[0] = 0.0f;
}

__global__ void refresh_ema_kernel(uint32_t n, float ema, int use_ema, const float *__restrict__ tmp, float *__restrict__ opa,
                                   float *__restrict__ sum) {
    arc_pdl_sync();
    float acc = 0.0f;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float o = opa[i];
        const float t = tmp[i];
        if (t >= 0.0f && o >= 0.0f) { o = use_ema ? fmaxf(o * ema, t) : t; opa[i] = o; }
        acc += fmaxf(o, 0.0f);
    }
    #pragma unroll
    for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(ARC_FULL_MASK, acc, s);
    __shared__ float ws[32];
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        acc = threadIdx.x < (blockDim.x >> 5) ? ws[threadIdx.x] : 0.0f;
        #pragma unroll
        for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(ARC_FULL_MASK, acc, s);
        if (threadIdx.x == 0) atomicAdd(sum, acc);
    }
}

__global__ void refresh_threshold_kernel(uint32_t n, const float *__restrict__ opa, const float *__restrict__ sum,
                                         float threshold, uint8_t *__restrict__ volume) {
    arc_pdl_sync();
    const float thres = fminf(sum[0] / (float)n, threshold);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        volume[i] = opa[i] >= thres ? 1 : 0;
}

}  // namespace

extern "C" {

// Bytes of scratch for arcb200_volume_refresh_candidates / _apply on an n_grid^3 volume.
int64_t arcb200_volume_refresh_scratch_bytes(int n_grid) {
    const int64_t n = (int64_t)n_grid * n_grid * n_grid;
    const int64_t blocks = (n + kOccBlock - 1) / kOccBlock;
    return n * 4 /* per-voxel max */ + (blocks + 64) * 4 /* block counts */ + 256 /* n_occ, sum */;
}

// Step 1.  volume: (n_grid^3) occupancy bytes (x-major flat order); idx_out (n_uniform + n_occ_max) int32, -1 where
// fewer than n_occ_max voxels are occupied; pts_out (n_uniform + n_occ_max, 3) world positions.  aabb: (2,3) min; max.
// key selects the permutation (use the refresh counter); rng_state/rng_inc: pcg32 for the jitter (caller advances by
// 3 * (n_uniform + n_occ_max) afterwards, or by 2^32 like the other samplers).
int arcb200_volume_refresh_candidates(const uint8_t *volume, int n_grid, int64_t n_uniform, int64_t n_occ_max,
                                      uint32_t key, uint64_t rng_state, uint64_t rng_inc, const float *aabb, void *scratch,
                                      int32_t *idx_out, float *pts_out, void *stream) {
    ARC_REQUIRE(n_grid >= 1 && n_grid <= 1024, "volume_refresh: n_grid out of range");
    const int64_t n = (int64_t)n_grid * n_grid * n_grid;
    ARC_REQUIRE(n_uniform >= 0 && n_uniform <= n && n_occ_max >= 0 && n_occ_max <= n, "volume_refresh: candidate counts out of range");
    ARC_REQUIRE(n_uniform + n_occ_max < (1ll << 31), "volume_refresh: too many candidates");
    if (n_uniform + n_occ_max == 0) return ARCB200_OK;
    ARC_REQUIRE(volume && aabb && scratch && idx_out && pts_out, "volume_refresh: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t blocks_occ = (uint32_t)((n + kOccBlock - 1) / kOccBlock);
    uint32_t *counts = reinterpret_cast<uint32_t *>(reinterpret_cast<uint8_t *>(scratch) + n * 4);
    uint32_t *n_occ = counts + blocks_occ + 32;
    int bits = 2;
    while ((1ll << bits) < n) bits += 2;                                    // even number of bits for the balanced network
    ARC_CUDA(cudaMemsetAsync(idx_out + n_uniform, 0xff, (size_t)n_occ_max * 4, st));
    arc_launch(occ_count_kernel, blocks_occ, 256, 0, st, (uint32_t)n, volume, counts);
    arc_launch(occ_scan_kernel, 1, 1024, 0, st, blocks_occ, counts, n_occ);
    Pcg32 rng;
    rng.state = rng_state; rng.inc = rng_inc;
    const uint32_t blocks_uni = (uint32_t)((n_uniform + 255) / 256);
    arc_launch(refresh_candidates_kernel, blocks_occ + blocks_uni, 256, 0, st, (uint32_t)n, (uint32_t)n_grid, (uint32_t)n_uniform,
               (uint32_t)n_occ_max, volume, (const uint32_t *)counts, key, bits / 2, rng, aabb, idx_out, pts_out);
    arc_launch(refresh_points_kernel, arc_blocks(n_uniform + n_occ_max, 256), 256, 0, st, (uint32_t)(n_uniform + n_occ_max),
               (uint32_t)n_grid, rng, aabb, (const int32_t *)idx_out, pts_out);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Steps 3 and 4.  opacity (n_cand) of the candidate points; opafield (n_grid^3) updated in place; volume (n_grid^3)
// occupancy bytes overwritten.  ema < 0: no moving average (new value replaces the old one).
int arcb200_volume_refresh_apply(const int32_t *idx, const float *opacity, int64_t n_cand, float ema, float threshold,
                                 int n_grid, void *scratch, float *opafield, uint8_t *volume, void *stream) {
    ARC_REQUIRE(n_grid >= 1 && n_grid <= 1024 && n_cand >= 0 && n_cand < (1ll << 31), "volume_refresh_apply: sizes out of range");
    ARC_REQUIRE(idx && opacity && scratch && opafield && volume, "volume_refresh_apply: null pointer");
    const int64_t n = (int64_t)n_grid * n_grid * n_grid;
    cudaStream_t st = (cudaStream_t)stream;
    float *tmp = reinterpret_cast<float *>(scratch);
    const uint32_t blocks_occ = (uint32_t)((n + kOccBlock - 1) / kOccBlock);
    float *sum = reinterpret_cast<float *>(reinterpret_cast<uint8_t *>(scratch) + n * 4) + blocks_occ + 48;
    const unsigned sweep = (unsigned)((n / 4 + 255) / 256 < (int64_t)arc_sm_count() * 8 ? (n / 4 + 255) / 256 + 1 : arc_sm_count() * 8);
    arc_launch(refresh_fill_kernel, sweep, 256, 0, st, (uint32_t)n, -1.0f, tmp, sum);
    if (n_cand > 0) arc_launch(refresh_splat_kernel, arc_blocks(n_cand, 256), 256, 0, st, (uint32_t)n_cand, (uint32_t)n, idx, opacity, tmp);
    arc_launch(refresh_ema_kernel, sweep, 256, 0, st, (uint32_t)n, ema, ema >= 0.0f ? 1 : 0, (const float *)tmp, opafield, sum);
    arc_launch(refresh_threshold_kernel, sweep, 256, 0, st, (uint32_t)n, (const float *)opafield, (const float *)sum, threshold, volume);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
