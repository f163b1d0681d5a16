// Ray / occupancy kernels of the hot path: slab test, occupancy marching (bool volume, Morton bitfield,
// cascaded bitfield), point-in-occupied-voxel test, and the packed (ragged) sample layout.
//
// THIS TRANSLATION UNIT IS COMPILED WITH -fmad=false.  Sample positions, counts and masks must be
// bit-identical to the reference's arithmetic evaluated one rounded fp32 operation at a time (the
// semantics of its HOST_DEVICE functions, arcnerf/ops/include/volume_func.h, on the host - which is what
// oracle/ pins).  A contracted FMA in `o + d*t` or in the distance-to-next-voxel expression moves a
// sample by 1 ulp and can flip a voxel decision.
//
// B200 design (not the reference's one-thread-per-ray loop):
//  * One WARP marches one ray.  The reference's loop only ever visits values of the chain
//    c_0 = t_start, c_{k+1} = fl(c_k + dt): an occupied sample advances by one link, an empty one by
//    the smallest number of links that reaches the next-voxel distance.  So the 32 lanes evaluate 32
//    consecutive chain links at once (position, box test, occupancy lookup, skip target - 32 lookups in
//    flight instead of one dependent lookup per step) and the visit order is then resolved with ballots:
//    runs of occupied links are emitted together (warp-level compaction: rank inside the run + running
//    count = output slot, written coalesced), an empty link jumps to the first lane whose chain value
//    reaches its skip target.
//  * Grids are sized in warps; the 2 MB bool volume / 256 KB bitfield stay L2/L1 resident.
#include "common.cuh"
#include "pcg32.cuh"
#include "occ_math.cuh"
#include <math_constants.h>

using namespace occ;

namespace {

// ------------------------------------------------------------------------------------------------
// Slab test kernel.  One thread per (ray, volume).  near/far = 0 and mask = false unless the near root
// is strictly positive (a ray that starts inside the box is a miss, as in the reference kernel).
__global__ void aabb_kernel(uint32_t n_rays, uint32_t n_v, const float *__restrict__ rays_o,
                            const float *__restrict__ rays_d, const float *__restrict__ aabb,
                            float *__restrict__ near, float *__restrict__ far, float *__restrict__ pts,
                            uint8_t *__restrict__ mask) {
    arc_pdl_sync();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rays * n_v) return;
    const uint32_t r = i / n_v, v = i % n_v;
    const F3 o = load3(rays_o + 3 * (size_t)r), d = load3(rays_d + 3 * (size_t)r);
    float tn, tf;
    slab_test(o, d, load3(aabb + 6 * v), load3(aabb + 6 * v + 3), tn, tf);
    const bool hit = tn > 0;
    const float a = hit ? tn : 0.0f, b = hit ? tf : 0.0f;
    near[i] = a;
    far[i] = b;
    mask[i] = hit ? 1 : 0;
    if (pts) {
        float *p = pts + 6 * (size_t)i;
        p[0] = o.x + a * d.x; p[1] = o.y + a * d.y; p[2] = o.z + a * d.z;
        p[3] = o.x + b * d.x; p[4] = o.y + b * d.y; p[5] = o.z + b * d.z;
    }
}

__global__ void occ_points_kernel(uint32_t n, const float *__restrict__ xyz, const uint8_t *__restrict__ vol,
                                  const float *__restrict__ aabb, uint32_t n_grid, uint8_t *__restrict__ out) {
    arc_pdl_sync();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Box box = load_box(aabb);
    out[i] = occupied_flat(load3(xyz + 3 * (size_t)i), vol, box, n_grid) ? 1 : 0;
}

// ------------------------------------------------------------------------------------------------
// Warp-per-ray occupancy march, fixed step.  BIT = 0: bool volume; BIT = 1: Morton bitfield.
// Writes, for ray i: zvals[i*stride + 0..count) = sample distances (and mask = 1 there);
//   fill_tail != 0 -> the rest of the row as the reference leaves it (repeat of the last distance with
//   mask 0; an empty ray is all zeros);  counts[i] = count (if counts != nullptr).
template <int BIT>
__global__ void __launch_bounds__(256)
march_warp_kernel(uint32_t n_rays, const float *__restrict__ rays_o, const float *__restrict__ rays_d,
                  const float *__restrict__ near, const float *__restrict__ far, const float *__restrict__ aabb,
                  const uint8_t *__restrict__ field, uint32_t n_grid, uint32_t n_pts, float dt,
                  float near_distance, uint64_t rng_state, uint64_t rng_inc, float *__restrict__ zvals,
                  uint8_t *__restrict__ mask, int32_t *__restrict__ counts, int fill_tail) {
    arc_pdl_sync();
    const uint32_t ray = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31u;
    if (ray >= n_rays) return;

    const Box box = load_box(aabb);
    const F3 o = load3(rays_o + 3 * (size_t)ray), d = load3(rays_d + 3 * (size_t)ray);
    const float t_end = far[ray];
    float chain = fmaxf(near[ray], near_distance);
    {
        Pcg32 rng;
        rng.state = rng_state;
        rng.inc = rng_inc;
        rng.advance((uint64_t)(int64_t)(ray * 8u));
        chain = chain + dt * rng.next_f32();
    }

    float *z_row = zvals + (size_t)ray * n_pts;
    uint8_t *m_row = mask ? mask + (size_t)ray * n_pts : nullptr;

    uint32_t emitted = 0;       // samples written so far (the reference's j)
    uint32_t first_lane = 0;    // next visited link is the first lane >= first_lane ...
    float reach = -CUDART_INF_F;  // ... whose chain value is >= reach
    float last_t = 0.0f;
    bool done = false;
    const uint32_t lane_lt = (1u << lane) - 1u;

    uint32_t chunks = 0;
    while (!done) {
        // a step too small to move t (t + dt == t) would never terminate, here as in the reference: give up after
        // 4 Mi links instead of hanging the device
        if (++chunks > (1u << 17)) break;
        // 32 consecutive links of the chain; every lane runs the same 32 dependent adds.
        float mine = chain;
        #pragma unroll
        for (uint32_t s = 0; s < 32; ++s) {
            if (s == lane) mine = chain;
            chain = chain + dt;
        }
        const F3 p = ray_at(o, d, mine);
        const bool live = (mine <= t_end) && inside(p, box);
        bool occ = false;
        if (live) occ = BIT ? occupied_bit(p, field, box, n_grid) : occupied_flat(p, field, box, n_grid);
        const float skip_to = mine + next_voxel_distance(p, d, box, n_grid);

        const uint32_t live_mask = __ballot_sync(ARC_FULL_MASK, live);
        const uint32_t occ_mask = __ballot_sync(ARC_FULL_MASK, occ);

        // entry link of this chunk
        // (`!(c < reach)` rather than `c >= reach`: a NaN target ends the reference's do-while after one step)
        const uint32_t cand = __ballot_sync(ARC_FULL_MASK, lane >= first_lane && !(mine < reach));
        if (cand == 0) { first_lane = 0; continue; }                 // the next visit lies in a later chunk
        const uint32_t entry = __ffs(cand) - 1;

        // successor of every link inside the chunk (32 = leaves the chunk): an occupied link steps to the next
        // one, an empty link to the first link whose chain value reaches its skip target (at least one step,
        // the reference's do-while).  The chain is non-decreasing over lanes: lower bound by 5 shuffles.
        uint32_t hop;
        {
            uint32_t lo = 0;
            #pragma unroll
            for (uint32_t step = 16; step >= 1; step >>= 1) {
                const float probe = __shfl_sync(ARC_FULL_MASK, mine, lo + step - 1);
                if (probe < skip_to) lo += step;
            }
            const float c31 = __shfl_sync(ARC_FULL_MASK, mine, 31);
            if (lo == 31 && c31 < skip_to) lo = 32;
            hop = occ ? lane + 1 : max(lane + 1, lo);
        }
        // links visited from the entry: reachability by pointer doubling (5 rounds cover 31 hops)
        uint32_t visited = 1u << entry;
        #pragma unroll
        for (int r = 0; r < 5; ++r) {
            const uint32_t add = (((visited >> lane) & 1u) && hop < 32) ? (1u << hop) : 0u;
            visited |= __reduce_or_sync(ARC_FULL_MASK, add);
            const uint32_t hop2 = __shfl_sync(ARC_FULL_MASK, hop, hop & 31u);
            hop = hop < 32 ? hop2 : 32u;
        }
        // the walk ends at the first visited link that is past `far` or outside the box
        const uint32_t stop = visited & ~live_mask;
        uint32_t valid = visited;
        if (stop) { valid &= (1u << (__ffs(stop) - 1)) - 1u; done = true; }
        // warp-level compaction of the occupied visited links: rank inside the set + running count = slot
        const uint32_t emit = valid & occ_mask;
        const uint32_t room = n_pts - emitted;
        const uint32_t rank = __popc(emit & lane_lt);
        const bool write = ((emit >> lane) & 1u) && rank < room;
        if (write) {
            z_row[emitted + rank] = mine;
            if (m_row) m_row[emitted + rank] = 1;
        }
        const uint32_t wrote = __ballot_sync(ARC_FULL_MASK, write);
        if (wrote) {
            last_t = __shfl_sync(ARC_FULL_MASK, mine, 31 - __clz(wrote));
            emitted += __popc(wrote);
        }
        if (emitted >= n_pts) done = true;
        if (!done) {
            // leave the chunk through the last visited link: an occupied one continues with the very next link,
            // an empty one carries its skip target into the next chunk
            const uint32_t last = 31 - __clz(valid);
            const float carry = __shfl_sync(ARC_FULL_MASK, skip_to, last);
            reach = ((occ_mask >> last) & 1u) ? -CUDART_INF_F : carry;
            first_lane = 0;
        }
    }

    if (counts && lane == 0) counts[ray] = (int32_t)emitted;
    if (fill_tail) {
        const float pad = emitted > 0 ? last_t : 0.0f;
        for (uint32_t k = emitted + lane; k < n_pts; k += 32) {
            z_row[k] = pad;
            if (m_row) m_row[k] = 0;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Cascaded (multi-volume) bitfield march with cone stepping (multivol_func_kernel.cu:19-106); the kernel is below the helpers.
__device__ __forceinline__ int mip_level(F3 p, const Box &inner, uint32_t n_cascades) {
    const float cx = (inner.lo.x + inner.hi.x) / 2.0f, cy = (inner.lo.y + inner.hi.y) / 2.0f,
                cz = (inner.lo.z + inner.hi.z) / 2.0f;
    const float hx = (inner.hi.x - inner.lo.x) / 2.0f, hy = (inner.hi.y - inner.lo.y) / 2.0f,
                hz = (inner.hi.z - inner.lo.z) / 2.0f;
    int ex, ey, ez;
    frexpf(fabsf(p.x - cx) * (1.0f / hx), &ex);
    frexpf(fabsf(p.y - cy) * (1.0f / hy), &ey);
    frexpf(fabsf(p.z - cz) * (1.0f / hz), &ez);
    int e = max(max(ex, ey), ez);
    return min((int)n_cascades - 1, max(0, e));
}
__device__ __forceinline__ bool occupied_cascade(F3 p, const uint8_t *__restrict__ bits, uint32_t mip,
                                                 const Box &inner, uint32_t n, bool inclusive) {
    const float cx = (inner.lo.x + inner.hi.x) / 2.0f, cy = (inner.lo.y + inner.hi.y) / 2.0f,
                cz = (inner.lo.z + inner.hi.z) / 2.0f;
    const float s = scalbnf(1.0f, -(int)mip);
    const F3 q{(p.x - cx) * s + cx, (p.y - cy) * s + cy, (p.z - cz) * s + cz};
    const uint32_t idx = cell_morton(q, inner, n);
    const uint32_t level = inclusive ? mip : mip - 1;
    const uint32_t off = (n * n * n) * level;
    return (__ldg(bits + (idx >> 3) + (off >> 3)) & (1u << (idx & 7u))) != 0;
}
__device__ __forceinline__ float cone_step(float t, float cone, float lo, float hi) { return clampf(t * cone, lo, hi); }

// One warp per ray.  Between two visits of the inner volume the reference's loop only ever stands on links of the chain
// c_0 = t_start, c_{k+1} = fl(c_k + clamp(c_k * cone, smin, smax)): an occupied link steps to the next one, an empty link walks
// the same chain up to its skip target.  So, as in march_warp_kernel, the 32 lanes evaluate 32 consecutive links at once (32
// cascade look-ups in flight), ballots and pointer doubling resolve which links the reference would have visited, and occupied
// visited links are emitted with warp-level compaction.  A visited link inside the inner volume (exclusive mode) leaves the
// chain - the reference then steps with the FIXED step of that link - so the chunk ends there, the sample count is reset and
// the chain is re-based on the point that loop reaches.
__global__ void __launch_bounds__(256)
march_multivol_kernel(uint32_t n_rays, const float *__restrict__ rays_o, const float *__restrict__ rays_d,
                      const float *__restrict__ near, const float *__restrict__ far, const float *__restrict__ min_aabb,
                      const float *__restrict__ aabb, const uint8_t *__restrict__ bits, uint32_t n_grid, uint32_t n_cascade,
                      uint32_t n_pts, float cone, float smin, float smax, float near_distance, int inclusive,
                      uint64_t rng_state, uint64_t rng_inc, float *__restrict__ zvals, uint8_t *__restrict__ mask) {
    arc_pdl_sync();
    const uint32_t ray = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31u;
    if (ray >= n_rays) return;
    const Box inner = load_box(min_aabb), outer = load_box(aabb);
    const F3 o = load3(rays_o + 3 * (size_t)ray), d = load3(rays_d + 3 * (size_t)ray);
    const float t_end = far[ray];
    float chain = fmaxf(near[ray], near_distance);
    {
        Pcg32 rng;
        rng.state = rng_state;
        rng.inc = rng_inc;
        rng.advance((uint64_t)(int64_t)(ray * 8u));
        chain = chain + cone_step(chain, cone, smin, smax) * rng.next_f32();
    }
    float *z_row = zvals + (size_t)ray * n_pts;
    uint8_t *m_row = mask + (size_t)ray * n_pts;
    uint32_t emitted = 0;
    float reach = -CUDART_INF_F, last_t = 0.0f;
    bool done = false;
    const uint32_t lane_lt = (1u << lane) - 1u;
    uint32_t chunks = 0;
    while (!done) {
        if (++chunks > (1u << 17)) break;                 // a step too small to move t never terminates in the reference either
        // While the next visit is the very link the chain stands on and that link lies in the inner volume (exclusive mode), the
        // reference resets and re-bases link after link (a camera inside the foreground box crosses about a hundred voxels this
        // way): pure arithmetic, done by all lanes alike without building a 32-link chunk for each of them.
        if (!inclusive) {
            uint32_t guard = 0;
            while (!(chain < reach) && chain <= t_end && ++guard < (1u << 20)) {
                const F3 p0 = ray_at(o, d, chain);
                if (!inside(p0, outer) || mip_level(p0, inner, n_cascade) != 0) break;
                const float target = chain + next_voxel_distance(p0, d, inner, n_grid);
                const float dt0 = cone_step(chain, cone, smin, smax);
                float t = chain;
                uint32_t g2 = 0;
                do { t = t + dt0; } while (t < target && ++g2 < (1u << 22));
                emitted = 0;
                chain = t;
                reach = -CUDART_INF_F;
            }
        }
        float mine = chain;
        #pragma unroll
        for (uint32_t s = 0; s < 32; ++s) {
            if (s == lane) mine = chain;
            chain = chain + cone_step(chain, cone, smin, smax);
        }
        const F3 p = ray_at(o, d, mine);
        const bool live = (mine <= t_end) && inside(p, outer);
        bool occ = false, reset = false;
        if (live) {
            const uint32_t mip = (uint32_t)mip_level(p, inner, n_cascade);
            if (mip == 0 && !inclusive) reset = true;
            else occ = occupied_cascade(p, bits, mip, inner, n_grid, inclusive != 0);
        }
        const float skip_to = mine + next_voxel_distance(p, d, inner, n_grid);
        const uint32_t live_mask = __ballot_sync(ARC_FULL_MASK, live);
        const uint32_t occ_mask = __ballot_sync(ARC_FULL_MASK, occ);
        const uint32_t reset_mask = __ballot_sync(ARC_FULL_MASK, reset);
        const uint32_t cand = __ballot_sync(ARC_FULL_MASK, !(mine < reach));
        if (cand == 0) continue;                          // the next visit lies in a later chunk
        const uint32_t entry = __ffs(cand) - 1;
        uint32_t hop;
        {
            uint32_t lo = 0;
            #pragma unroll
            for (uint32_t step = 16; step >= 1; step >>= 1) {
                const float probe = __shfl_sync(ARC_FULL_MASK, mine, lo + step - 1);
                if (probe < skip_to) lo += step;
            }
            const float c31 = __shfl_sync(ARC_FULL_MASK, mine, 31);
            if (lo == 31 && c31 < skip_to) lo = 32;
            hop = occ ? lane + 1 : max(lane + 1, lo);
            if (reset || !live) hop = 32;                 // the walk ends on such a link
        }
        uint32_t visited = 1u << entry;
        #pragma unroll
        for (int r = 0; r < 5; ++r) {
            const uint32_t add = (((visited >> lane) & 1u) && hop < 32) ? (1u << hop) : 0u;
            visited |= __reduce_or_sync(ARC_FULL_MASK, add);
            const uint32_t hop2 = __shfl_sync(ARC_FULL_MASK, hop, hop & 31u);
            hop = hop < 32 ? hop2 : 32u;
        }
        // the walk inside this chunk ends at the first visited link that is dead (past `far` / outside) or inside the inner volume
        const uint32_t stop = visited & (~live_mask | reset_mask);
        uint32_t valid = visited;
        int stop_lane = -1;
        if (stop) { stop_lane = __ffs(stop) - 1; valid &= (1u << stop_lane) - 1u; }
        const uint32_t emit = valid & occ_mask;
        const uint32_t room = n_pts - emitted;
        const uint32_t rank = __popc(emit & lane_lt);
        const bool write = ((emit >> lane) & 1u) && rank < room;
        if (write) z_row[emitted + rank] = mine;
        const uint32_t wrote = __ballot_sync(ARC_FULL_MASK, write);
        if (wrote) {
            last_t = __shfl_sync(ARC_FULL_MASK, mine, 31 - __clz(wrote));
            emitted += __popc(wrote);
        }
        if (emitted >= n_pts) { done = true; continue; }  // the reference's `j < n_pts` ends the loop before the next link is looked at
        if (stop_lane >= 0) {
            if (!((live_mask >> stop_lane) & 1u)) { done = true; continue; }
            // inner volume: samples so far are dropped, then fixed steps of this link's size up to its skip target
            const float ti = __shfl_sync(ARC_FULL_MASK, mine, stop_lane);
            const float target = __shfl_sync(ARC_FULL_MASK, skip_to, stop_lane);
            const float dti = cone_step(ti, cone, smin, smax);
            float t = ti;
            uint32_t guard = 0;
            do { t = t + dti; } while (t < target && ++guard < (1u << 22));
            emitted = 0;
            chain = t;
            reach = -CUDART_INF_F;
            continue;
        }
        const uint32_t last = 31 - __clz(valid);
        const float carry = __shfl_sync(ARC_FULL_MASK, skip_to, last);
        reach = ((occ_mask >> last) & 1u) ? -CUDART_INF_F : carry;
    }
    const float pad = emitted > 0 ? last_t : 0.0f;
    for (uint32_t k = lane; k < n_pts; k += 32) {
        if (k >= emitted) z_row[k] = pad;
        m_row[k] = k < emitted ? 1 : 0;
    }
}

// ------------------------------------------------------------------------------------------------
// Packed (ragged) layout.  counts[n_rays] -> offsets[n_rays+1] (exclusive scan), total and max count.
// Three tiny kernels; tiles of 1024 rays.
constexpr int kScanTile = 1024;

__global__ void __launch_bounds__(kScanTile)
scan_tiles_kernel(uint32_t n, const int32_t *__restrict__ counts, int32_t *__restrict__ offsets,
                  int32_t *__restrict__ tile_sums, int32_t *__restrict__ max_count) {
    arc_pdl_sync();
    __shared__ int32_t warp_sums[32];
    const uint32_t i = blockIdx.x * kScanTile + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const int32_t c = i < n ? counts[i] : 0;
    int32_t incl = c;
    #pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
        const int32_t up = __shfl_up_sync(ARC_FULL_MASK, incl, s);
        if ((int)lane >= s) incl += up;
    }
    int32_t mx = c;
    #pragma unroll
    for (int s = 16; s > 0; s >>= 1) mx = max(mx, __shfl_xor_sync(ARC_FULL_MASK, mx, s));
    if (lane == 31) warp_sums[w] = incl;
    if (lane == 0 && mx > 0) atomicMax(max_count, mx);
    __syncthreads();
    if (w == 0) {
        int32_t ws = warp_sums[lane];
        int32_t wi = ws;
        #pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const int32_t up = __shfl_up_sync(ARC_FULL_MASK, wi, s);
            if ((int)lane >= s) wi += up;
        }
        warp_sums[lane] = wi - ws;                 // exclusive warp prefix
        if (lane == 31) tile_sums[blockIdx.x] = wi;  // tile total
    }
    __syncthreads();
    if (i < n) offsets[i] = warp_sums[w] + incl - c;   // exclusive within the tile
}

__global__ void __launch_bounds__(1024)
scan_tile_sums_kernel(uint32_t n_tiles, int32_t *__restrict__ tile_sums, int32_t *__restrict__ total) {
    arc_pdl_sync();
    // single block: exclusive scan of tile_sums in place, total to *total
    __shared__ int32_t warp_sums[32];
    __shared__ int32_t carry_s;
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_tiles; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const int32_t c = i < n_tiles ? tile_sums[i] : 0;
        int32_t incl = c;
        #pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const int32_t up = __shfl_up_sync(ARC_FULL_MASK, incl, s);
            if ((int)lane >= s) incl += up;
        }
        if (lane == 31) warp_sums[w] = incl;
        __syncthreads();
        if (w == 0) {
            const int32_t ws = warp_sums[lane];
            int32_t wi = ws;
            #pragma unroll
            for (int s = 1; s < 32; s <<= 1) {
                const int32_t up = __shfl_up_sync(ARC_FULL_MASK, wi, s);
                if ((int)lane >= s) wi += up;
            }
            warp_sums[lane] = wi - ws;
        }
        __syncthreads();
        const int32_t carry = carry_s;
        if (i < n_tiles) tile_sums[i] = carry + warp_sums[w] + incl - c;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + warp_sums[w] + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry_s;
}

// Warp per ray: offsets[ray] += tile prefix; copy the ray's samples from the row-strided scratch to the
// packed arrays and derive what the encoder and the compositor need.
__global__ void __launch_bounds__(256)
pack_samples_kernel(uint32_t n_rays, uint32_t n_pts, const float *__restrict__ rays_o,
                    const float *__restrict__ rays_d, const float *__restrict__ zrows,
                    const int32_t *__restrict__ counts, int32_t *__restrict__ offsets,
                    const int32_t *__restrict__ tile_prefix, const int32_t *__restrict__ total,
                    int64_t capacity, float *__restrict__ t_out, float *__restrict__ xyz_out,
                    int32_t *__restrict__ ray_id_out, const float *__restrict__ aabb, float *__restrict__ unit_out) {
    arc_pdl_sync();
    const uint32_t ray = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31u;
    if (ray > n_rays) return;
    if (ray == n_rays) {   // one extra warp writes the end sentinel
        if (lane == 0) offsets[n_rays] = *total;
        return;
    }
    const int32_t base = offsets[ray] + tile_prefix[ray / kScanTile];
    __syncwarp();
    if (lane == 0) offsets[ray] = base;
    const int32_t cnt = counts[ray];
    const F3 o = load3(rays_o + 3 * (size_t)ray), d = load3(rays_d + 3 * (size_t)ray);
    const float *z = zrows + (size_t)ray * n_pts;
    for (int32_t k = lane; k < cnt; k += 32) {
        const int64_t s = (int64_t)base + k;
        if (s >= capacity) break;    // caller-provided capacity exceeded: samples beyond it are dropped
        const float t = z[k];
        t_out[s] = t;
        const F3 p = ray_at(o, d, t);
        if (xyz_out) { xyz_out[3 * s] = p.x; xyz_out[3 * s + 1] = p.y; xyz_out[3 * s + 2] = p.z; }
        if (unit_out) {   // (p - min) / (max - min): the normalisation HashGridEmbedder applies (hashgrid_encoder.py:183)
            const Box b = load_box(aabb);
            unit_out[3 * s] = (p.x - b.lo.x) / (b.hi.x - b.lo.x);
            unit_out[3 * s + 1] = (p.y - b.lo.y) / (b.hi.y - b.lo.y);
            unit_out[3 * s + 2] = (p.z - b.lo.z) / (b.hi.z - b.lo.z);
        }
        if (ray_id_out) ray_id_out[s] = (int32_t)ray;
    }
}

}  // namespace

// ================================================================================================
extern "C" {

int arcb200_aabb_intersection(const float *rays_o, const float *rays_d, const float *aabb, float *near,
                              float *far, float *pts, uint8_t *mask, int64_t n_rays, int64_t n_v, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_v >= 0, "aabb_intersection: negative size");
    if (n_rays * n_v == 0) return ARCB200_OK;
    ARC_REQUIRE(rays_o && rays_d && aabb && near && far && mask, "aabb_intersection: null pointer");
    ARC_REQUIRE(n_rays * n_v < (1ll << 32), "aabb_intersection: n_rays*n_v must fit 32 bits");
    arc_launch(aabb_kernel, arc_blocks(n_rays * n_v, 256), 256, 0, (cudaStream_t)stream, 
        (uint32_t)n_rays, (uint32_t)n_v, rays_o, rays_d, aabb, near, far, pts, mask);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_check_pts_in_occ_voxel(const float *xyz, const uint8_t *bitfield, const float *aabb,
                                   int n_grid, uint8_t *out, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0 && n < (1ll << 32), "check_pts_in_occ_voxel: bad n");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(xyz && bitfield && aabb && out && n_grid > 0, "check_pts_in_occ_voxel: null/invalid argument");
    arc_launch(occ_points_kernel, arc_blocks(n, 256), 256, 0, (cudaStream_t)stream, (uint32_t)n, xyz, bitfield,
                                                                          aabb, (uint32_t)n_grid, out);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

static int launch_march(int bit, const float *rays_o, const float *rays_d, const float *near, const float *far,
                        int n_pts, float dt, const float *aabb, int n_grid, const uint8_t *field,
                        float near_distance, float *zvals, uint8_t *mask, int32_t *counts, int fill_tail,
                        int64_t n_rays, uint64_t rng_state, uint64_t rng_inc, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_rays < (1ll << 27), "sparse_volume_sampling: n_rays out of range");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(rays_o && rays_d && near && far && aabb && field && zvals, "sparse_volume_sampling: null pointer");
    ARC_REQUIRE(n_pts > 0 && n_grid > 0 && dt > 0.0f, "sparse_volume_sampling: n_pts, n_grid and dt must be positive");
    const unsigned int blocks = arc_blocks(n_rays * 32, 256);
    if (bit)
        arc_launch(march_warp_kernel<1>, blocks, 256, 0, (cudaStream_t)stream, 
            (uint32_t)n_rays, rays_o, rays_d, near, far, aabb, field, (uint32_t)n_grid, (uint32_t)n_pts, dt,
            near_distance, rng_state, rng_inc, zvals, mask, counts, fill_tail);
    else
        arc_launch(march_warp_kernel<0>, blocks, 256, 0, (cudaStream_t)stream, 
            (uint32_t)n_rays, rays_o, rays_d, near, far, aabb, field, (uint32_t)n_grid, (uint32_t)n_pts, dt,
            near_distance, rng_state, rng_inc, zvals, mask, counts, fill_tail);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_sparse_volume_sampling(const float *rays_o, const float *rays_d, const float *near, const float *far,
                                   int n_pts, float dt, const float *aabb, int n_grid,
                                   const uint8_t *bitfield, float near_distance, float *zvals, uint8_t *mask,
                                   int64_t n_rays, uint64_t rng_state, uint64_t rng_inc, void *stream) {
    ARC_REQUIRE(mask != nullptr || n_rays == 0, "sparse_volume_sampling: null mask");
    return launch_march(0, rays_o, rays_d, near, far, n_pts, dt, aabb, n_grid, bitfield, near_distance, zvals,
                        mask, nullptr, 1, n_rays, rng_state, rng_inc, stream);
}

int arcb200_sparse_volume_sampling_bit(const float *rays_o, const float *rays_d, const float *near,
                                       const float *far, int n_pts, float dt, const float *aabb, int n_grid,
                                       const uint8_t *bitfield, float near_distance, float *zvals, uint8_t *mask,
                                       int64_t n_rays, uint64_t rng_state, uint64_t rng_inc, void *stream) {
    ARC_REQUIRE(mask != nullptr || n_rays == 0, "sparse_volume_sampling_bit: null mask");
    return launch_march(1, rays_o, rays_d, near, far, n_pts, dt, aabb, n_grid, bitfield, near_distance, zvals,
                        mask, nullptr, 1, n_rays, rng_state, rng_inc, stream);
}

int arcb200_sparse_sampling_in_multivol_bitfield(const float *rays_o, const float *rays_d, const float *near,
                                                 const float *far, int n_pts, float cone_angle, float min_step,
                                                 float max_step, const float *min_aabb,
                                                 const float *aabb, int n_grid, int n_cascade,
                                                 const uint8_t *bitfield, float near_distance, int inclusive,
                                                 float *zvals, uint8_t *mask, int64_t n_rays, uint64_t rng_state,
                                                 uint64_t rng_inc, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_rays < (1ll << 31), "multivol sampling: n_rays out of range");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(rays_o && rays_d && near && far && min_aabb && aabb && bitfield && zvals && mask,
                "multivol sampling: null pointer");
    ARC_REQUIRE(n_pts > 0 && n_grid > 0 && n_cascade > 0 && min_step > 0.0f, "multivol sampling: invalid sizes/steps");
    arc_launch(march_multivol_kernel, arc_blocks(n_rays * 32, 256), 256, 0, (cudaStream_t)stream, 
        (uint32_t)n_rays, rays_o, rays_d, near, far, min_aabb, aabb, bitfield,
        (uint32_t)n_grid, (uint32_t)n_cascade, (uint32_t)n_pts, cone_angle, min_step, max_step, near_distance,
        inclusive, rng_state, rng_inc, zvals, mask);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Bytes of scratch arcb200_march_packed needs for n_rays rays of at most n_pts samples.
int64_t arcb200_march_packed_scratch_bytes(int64_t n_rays, int n_pts) {
    const int64_t n_tiles = (n_rays + kScanTile - 1) / kScanTile;
    int64_t b = n_rays * (int64_t)n_pts * 4;             // row-strided sample distances
    b = (b + 255) / 256 * 256;
    b += ((n_tiles + 1) * 4 + 255) / 256 * 256;          // tile sums
    return b;
}

// March + pack.  Outputs: counts[n_rays], offsets[n_rays+1] (exclusive prefix; offsets[n_rays] = total),
// meta[0] = total samples, meta[1] = max samples on one ray, and the packed per-sample arrays
// t[capacity], xyz[capacity*3] (world position, optional), ray_id[capacity] (optional), unit[capacity*3] (position
// normalised to the box, (p-min)/(max-min), optional: the hash-grid encoder's input).  Samples of a ray
// are contiguous and rays appear in index order.  `scratch` must hold arcb200_march_packed_scratch_bytes().
int arcb200_march_packed(int bit, const float *rays_o, const float *rays_d, const float *near, const float *far,
                         int n_pts, float dt, const float *aabb, int n_grid, const uint8_t *field,
                         float near_distance, int64_t n_rays, uint64_t rng_state, uint64_t rng_inc,
                         void *scratch, int32_t *counts, int32_t *offsets, int32_t *meta, int64_t capacity,
                         float *t_out, float *xyz_out, int32_t *ray_id_out, float *unit_out, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_rays < (1ll << 27), "march_packed: n_rays out of range");
    ARC_REQUIRE(counts && offsets && meta, "march_packed: null bookkeeping pointer");
    cudaStream_t st = (cudaStream_t)stream;
    ARC_CUDA(cudaMemsetAsync(meta, 0, 2 * sizeof(int32_t), st));
    if (n_rays == 0) {
        ARC_CUDA(cudaMemsetAsync(offsets, 0, sizeof(int32_t), st));
        return ARCB200_OK;
    }
    ARC_REQUIRE(scratch && t_out && capacity >= 0, "march_packed: null scratch/output");
    float *zrows = (float *)scratch;
    int64_t zbytes = (n_rays * (int64_t)n_pts * 4 + 255) / 256 * 256;
    int32_t *tile_sums = (int32_t *)((char *)scratch + zbytes);
    const uint32_t n_tiles = (uint32_t)((n_rays + kScanTile - 1) / kScanTile);
    int rc = launch_march(bit, rays_o, rays_d, near, far, n_pts, dt, aabb, n_grid, field, near_distance, zrows,
                          nullptr, counts, 0, n_rays, rng_state, rng_inc, stream);
    if (rc != ARCB200_OK) return rc;
    arc_launch(scan_tiles_kernel, n_tiles, kScanTile, 0, st, (uint32_t)n_rays, counts, offsets, tile_sums, meta + 1);
    ARC_LAUNCH_CHECK();
    arc_launch(scan_tile_sums_kernel, 1, 1024, 0, st, n_tiles, tile_sums, meta);
    ARC_LAUNCH_CHECK();
    arc_launch(pack_samples_kernel, arc_blocks((n_rays + 1) * 32, 256), 256, 0, st, 
        (uint32_t)n_rays, (uint32_t)n_pts, rays_o, rays_d, zrows, counts, offsets, tile_sums, meta, capacity,
        t_out, xyz_out, ray_id_out, aabb, unit_out);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
