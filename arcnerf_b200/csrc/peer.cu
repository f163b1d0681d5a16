// Data-parallel optimiser step over NVLink peer memory (one process per GPU, same node).
//
// The reference trains data-parallel through torch DDP: every rank all-reduces the full fp32 gradient and then
// runs the same Adam on the full parameter vector (common/trainer/basic_trainer.py:192-202,
// common/trainer/optimizer.py).  For the NGP hot path that is 48.8 MB of hash-table gradient per step, produced by
// the LAST backward kernel, so the exchange cannot hide behind the backward.  Here the exchange and the update are
// one kernel: rank r owns a contiguous slice of the parameter vector, loads that slice of every rank's gradient
// straight from the peers' HBM (NVLink loads, summed in rank order), applies Adam to its slice and stores the fp16
// working copy of the slice into every rank's table (NVLink stores).  Per rank and step that moves (N-1)/N * 4 B
// in and (N-1)/N * 2 B out per parameter over NVLink and 1/N of the Adam HBM stream - no NCCL on the data path.
// Small replicated vectors (the two MLP heads) use the same kernel with the full range on every rank: each rank
// sums all ranks' gradients in the same order, so the replicas stay bit-identical.
//
// Ordering between ranks is two flag barriers per step (arcb200_peer_barrier), each a 1-block kernel on the
// step's stream: A) "my backward is complete, my gradient may be read" before the update, B) "I no longer read your
// gradient nor write your table" after it.  Flags are epoch counters that only grow, kept in the shared arena.
#include "adam.cuh"
#include <string.h>
#include <stdlib.h>

namespace {

constexpr int kMaxWorld = 16;

struct PeerList {
    const float *grad[kMaxWorld];   // gradient vectors of all ranks (index = rank), same layout on every rank
    __half *half_out[kMaxWorld];    // fp16 working copies of all ranks, or all null
};

struct FlagList {
    unsigned long long *flags[kMaxWorld];   // flags[w] = base of rank w's flag array (kMaxWorld entries)
};

__device__ __forceinline__ float4 ld_peer4(const float *p, size_t k4) {
    // L1-bypassing 128-bit load: peer lines are never held in this SM's L1 across kernels, the home L2 is coherent
    return __ldcg(reinterpret_cast<const float4 *>(p) + k4);
}

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    a = fminf(fmaxf(a, -65504.0f), 65504.0f);          // saturate instead of overflowing to inf
    b = fminf(fmaxf(b, -65504.0f), 65504.0f);
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}

__device__ __forceinline__ void acc4(float4 &a, const float4 &b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }

// W > 0: world size known at compile time (all peer loads of one element group are in flight together);
// W == 0: run-time world size, sequential accumulation.
template <int W>
__global__ void adam_peer_kernel(size_t lo, size_t hi, float *__restrict__ p, float *__restrict__ m,
                                 float *__restrict__ v, const __grid_constant__ PeerList peers, int world,
                                 AdamHyper h, const int32_t *__restrict__ step_dev) {
    arc_pdl_sync();
    if (step_dev) bias_from_step(h, max(*step_dev, 1));
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t k_lo = lo / 4, k_hi = hi / 4;            // lo is a multiple of 4
    const bool want_half = peers.half_out[0] != nullptr;
    // two element groups per trip: every load of both groups (peer and local) is issued before the first store
    auto load_grad = [&](size_t k) {
        float4 gg;
        if (W > 0) {
            float4 part[W > 0 ? W : 1];
            #pragma unroll
            for (int w = 0; w < W; ++w) part[w] = ld_peer4(peers.grad[w], k);
            gg = part[0];
            #pragma unroll
            for (int w = 1; w < W; ++w) acc4(gg, part[w]);
        } else {
            gg = ld_peer4(peers.grad[0], k);
            for (int w = 1; w < world; ++w) acc4(gg, ld_peer4(peers.grad[w], k));
        }
        return gg;
    };
    auto update = [&](size_t k, float4 gg, float4 pp, float4 mm, float4 vv) {
        adam_one(pp.x, mm.x, vv.x, gg.x, h); adam_one(pp.y, mm.y, vv.y, gg.y, h);
        adam_one(pp.z, mm.z, vv.z, gg.z, h); adam_one(pp.w, mm.w, vv.w, gg.w, h);
        reinterpret_cast<float4 *>(p)[k] = pp;
        reinterpret_cast<float4 *>(m)[k] = mm;
        reinterpret_cast<float4 *>(v)[k] = vv;
        if (want_half) {
            __half2 a = __floats2half2_rn(pp.x, pp.y), b = __floats2half2_rn(pp.z, pp.w);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t *>(&a);
            pk.y = *reinterpret_cast<uint32_t *>(&b);
            const int n = W > 0 ? W : world;
            #pragma unroll
            for (int w = 0; w < (W > 0 ? W : kMaxWorld); ++w)
                if (w < n) reinterpret_cast<uint2 *>(peers.half_out[w])[k] = pk;
        }
    };
    for (size_t k = k_lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < k_hi; k += 2 * stride) {
        const size_t k2 = k + stride;
        const bool two = k2 < k_hi;
        const float4 g1 = load_grad(k);
        const float4 p1 = reinterpret_cast<const float4 *>(p)[k], m1 = reinterpret_cast<const float4 *>(m)[k],
                     v1 = reinterpret_cast<const float4 *>(v)[k];
        float4 g2, p2, m2, v2;
        if (two) {
            g2 = load_grad(k2);
            p2 = reinterpret_cast<const float4 *>(p)[k2]; m2 = reinterpret_cast<const float4 *>(m)[k2];
            v2 = reinterpret_cast<const float4 *>(v)[k2];
        }
        update(k, g1, p1, m1, v1);
        if (two) update(k2, g2, p2, m2, v2);
    }
    // scalar tail of the range (hi not a multiple of 4)
    for (size_t k = k_hi * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < hi; k += stride) {
        float g = __ldcg(peers.grad[0] + k);
        for (int w = 1; w < world; ++w) g += __ldcg(peers.grad[w] + k);
        float pp = p[k], mm = m[k], vv = v[k];
        adam_one(pp, mm, vv, g, h);
        p[k] = pp; m[k] = mm; v[k] = vv;
        if (want_half)
            for (int w = 0; w < world; ++w) peers.half_out[w][k] = __float2half_rn(pp);
    }
}

// ---- half-precision gradient exchange ---------------------------------------------------------------------------
// The reference's hash-grid gradients are fp16 to begin with (tcnn accumulates dL/dparams in __half with a loss scale of
// 128 and hands torch a float copy, tinycudann/modules.py:93-96,141), so what DDP averages has fp16 precision per rank.
// Here every rank keeps its fp32-accumulated gradient, but what TRAVELS is a loss-scaled fp16 copy of the part other
// ranks own: half the NVLink bytes of the fp32 exchange.  The owner adds its own fp32 slice and the peers' fp16 slices
// in fp32.

// g16 = half(g32 * scale) outside [keep_lo, keep_hi) (the slice this rank owns and reads in fp32); g32 is cleared there.
__global__ void grad_to_half_kernel(size_t n, size_t keep_lo, size_t keep_hi, float *__restrict__ g32,
                                    __half *__restrict__ g16, float scale) {
    arc_pdl_sync();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t n8 = n / 8;
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n8; k += stride) {
        const size_t e = k * 8;
        if (e >= keep_lo && e < keep_hi) continue;                  // keep_lo / keep_hi are multiples of 8
        float4 a = __ldcs(reinterpret_cast<const float4 *>(g32) + 2 * k), b = __ldcs(reinterpret_cast<const float4 *>(g32) + 2 * k + 1);
        uint4 w;
        w.x = pack_h2(a.x * scale, a.y * scale); w.y = pack_h2(a.z * scale, a.w * scale);
        w.z = pack_h2(b.x * scale, b.y * scale); w.w = pack_h2(b.z * scale, b.w * scale);
        reinterpret_cast<uint4 *>(g16)[k] = w;
        reinterpret_cast<float4 *>(g32)[2 * k] = make_float4(0.f, 0.f, 0.f, 0.f);
        reinterpret_cast<float4 *>(g32)[2 * k + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (size_t e = n8 * 8 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += stride) {
        if (e >= keep_lo && e < keep_hi) continue;
        g16[e] = __float2half_rn(fminf(fmaxf(g32[e] * scale, -65504.0f), 65504.0f));
        g32[e] = 0.f;
    }
}

struct Peer16List {
    const __half *grad16[kMaxWorld];   // every rank's fp16 gradient copy (entry `rank` is unused)
    __half *half_out[kMaxWorld];
};

// Adam over [lo, hi) on  own fp32 gradient + sum over the other ranks of their fp16 gradient / scale;  the own gradient
// slice is cleared, the updated fp16 working copy goes to every rank.  Eight elements per thread and trip.
template <int W>
__global__ void adam_peer16_kernel(size_t lo, size_t hi, float *__restrict__ p, float *__restrict__ m, float *__restrict__ v,
                                   float *__restrict__ g_own, const __grid_constant__ Peer16List peers, int world, int rank,
                                   float inv_scale, AdamHyper h, const int32_t *__restrict__ step_dev) {
    arc_pdl_sync();
    if (step_dev) bias_from_step(h, max(*step_dev, 1));
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t k_lo = lo / 8, k_hi = hi / 8;
    const int nw = W > 0 ? W : world;
    for (size_t k = k_lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < k_hi; k += stride) {
        uint4 part[W > 0 ? W : kMaxWorld];
        #pragma unroll
        for (int w = 0; w < (W > 0 ? W : kMaxWorld); ++w)
            if (w < nw && w != rank) part[w] = __ldcg(reinterpret_cast<const uint4 *>(peers.grad16[w]) + k);
        float g[8];
        {
            const float4 a = reinterpret_cast<const float4 *>(g_own)[2 * k], b = reinterpret_cast<const float4 *>(g_own)[2 * k + 1];
            g[0] = a.x; g[1] = a.y; g[2] = a.z; g[3] = a.w; g[4] = b.x; g[5] = b.y; g[6] = b.z; g[7] = b.w;
        }
        float pv[8], mv[8], vv[8];
        #pragma unroll
        for (int q = 0; q < 2; ++q) {
            const float4 a = reinterpret_cast<const float4 *>(p)[2 * k + q], b = reinterpret_cast<const float4 *>(m)[2 * k + q],
                         c = reinterpret_cast<const float4 *>(v)[2 * k + q];
            pv[4 * q] = a.x; pv[4 * q + 1] = a.y; pv[4 * q + 2] = a.z; pv[4 * q + 3] = a.w;
            mv[4 * q] = b.x; mv[4 * q + 1] = b.y; mv[4 * q + 2] = b.z; mv[4 * q + 3] = b.w;
            vv[4 * q] = c.x; vv[4 * q + 1] = c.y; vv[4 * q + 2] = c.z; vv[4 * q + 3] = c.w;
        }
        float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        #pragma unroll
        for (int w = 0; w < (W > 0 ? W : kMaxWorld); ++w) {
            if (w >= nw || w == rank) continue;
            const uint32_t wd[4] = {part[w].x, part[w].y, part[w].z, part[w].w};
            #pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&wd[q]));
                s[2 * q] += f.x; s[2 * q + 1] += f.y;
            }
        }
        uint32_t hw[4];
        #pragma unroll
        for (int q = 0; q < 8; ++q) {
            adam_one(pv[q], mv[q], vv[q], g[q] + s[q] * inv_scale, h);
        }
        #pragma unroll
        for (int q = 0; q < 4; ++q) hw[q] = pack_h2(pv[2 * q], pv[2 * q + 1]);
        #pragma unroll
        for (int q = 0; q < 2; ++q) {
            reinterpret_cast<float4 *>(p)[2 * k + q] = make_float4(pv[4 * q], pv[4 * q + 1], pv[4 * q + 2], pv[4 * q + 3]);
            reinterpret_cast<float4 *>(m)[2 * k + q] = make_float4(mv[4 * q], mv[4 * q + 1], mv[4 * q + 2], mv[4 * q + 3]);
            reinterpret_cast<float4 *>(v)[2 * k + q] = make_float4(vv[4 * q], vv[4 * q + 1], vv[4 * q + 2], vv[4 * q + 3]);
            reinterpret_cast<float4 *>(g_own)[2 * k + q] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        const uint4 pk = make_uint4(hw[0], hw[1], hw[2], hw[3]);
        #pragma unroll
        for (int w = 0; w < (W > 0 ? W : kMaxWorld); ++w)
            if (w < nw) reinterpret_cast<uint4 *>(peers.half_out[w])[k] = pk;
    }
    for (size_t e = k_hi * 8 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < hi; e += stride) {
        float s = 0.f;
        for (int w = 0; w < world; ++w)
            if (w != rank) s += __half2float(__ldcg(peers.grad16[w] + e));
        float pp = p[e], mm = m[e], vv = v[e];
        adam_one(pp, mm, vv, g_own[e] + s * inv_scale, h);
        p[e] = pp; m[e] = mm; v[e] = vv; g_own[e] = 0.f;
        for (int w = 0; w < world; ++w) peers.half_out[w][e] = __float2half_rn(pp);
    }
}

template <int W>
void launch_adam_peer16(unsigned blocks, cudaStream_t st, size_t lo, size_t hi, float *p, float *m, float *v, float *g_own,
                        const Peer16List &pl, int world, int rank, float inv_scale, const AdamHyper &h,
                        const int32_t *step_dev) {
    arc_launch(adam_peer16_kernel<W>, blocks, 256, 0, st, lo, hi, p, m, v, g_own, pl, world, rank, inv_scale, h, step_dev);
}

// ---- the exchange done by the switch (NVLS): the arena is also mapped through a multicast object (torch symmetric memory sets
// it up, arcnerf_b200/peer.py SymmArena); one multimem.ld_reduce returns the sum over ALL ranks' fp16 gradient copies of 8
// entries (added in fp32 inside the fabric, returned as f16x2), one multimem.st writes the updated fp16 entries into every
// rank's working table.  Per element a GPU issues one load and one store instead of world-1 loads and world stores, and each
// NVLink carries the reduced / broadcast data once.
__device__ __forceinline__ uint4 mc_ld_reduce_h8(const void *mc) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(mc) : "memory");
    return r;
}
__device__ __forceinline__ void mc_st_h8(void *mc, const uint4 &v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f16x2 [%0], {%1, %2, %3, %4};"
                 :: "l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// Adam over [lo, hi) (multiples of 8) on  sum over all ranks of grad16 / scale;  mc_grad16 / mc_half_out are MULTICAST addresses of
// the fp16 gradient copy and of the fp16 working table.
__global__ void adam_mc16_kernel(size_t lo, size_t hi, float *__restrict__ p, float *__restrict__ m, float *__restrict__ v,
                                 const __half *mc_grad16, __half *mc_half_out, float inv_scale, AdamHyper h,
                                 const int32_t *__restrict__ step_dev) {
    arc_pdl_sync();
    if (step_dev) bias_from_step(h, max(*step_dev, 1));
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    constexpr int U = 1;                                  // reductions in flight per thread (more did not help: N = 2 measured 85 vs 76 us at U = 4)
    for (size_t k0 = lo / 8 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; k0 < hi / 8; k0 += stride * U) {
        uint4 gsu[U];
        #pragma unroll
        for (int u = 0; u < U; ++u)
            if (k0 + u * stride < hi / 8) gsu[u] = mc_ld_reduce_h8(reinterpret_cast<const uint4 *>(mc_grad16) + k0 + u * stride);
        #pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t k = k0 + u * stride;
            if (k >= hi / 8) break;
            const uint4 gs = gsu[u];
            float pv[8], mv[8], vv[8], g[8];
            #pragma unroll
            for (int q = 0; q < 2; ++q) {
                const float4 a = reinterpret_cast<const float4 *>(p)[2 * k + q], b = reinterpret_cast<const float4 *>(m)[2 * k + q],
                             c = reinterpret_cast<const float4 *>(v)[2 * k + q];
                pv[4 * q] = a.x; pv[4 * q + 1] = a.y; pv[4 * q + 2] = a.z; pv[4 * q + 3] = a.w;
                mv[4 * q] = b.x; mv[4 * q + 1] = b.y; mv[4 * q + 2] = b.z; mv[4 * q + 3] = b.w;
                vv[4 * q] = c.x; vv[4 * q + 1] = c.y; vv[4 * q + 2] = c.z; vv[4 * q + 3] = c.w;
            }
            const uint32_t wd[4] = {gs.x, gs.y, gs.z, gs.w};
            #pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&wd[q]));
                g[2 * q] = f.x * inv_scale; g[2 * q + 1] = f.y * inv_scale;
            }
            #pragma unroll
            for (int q = 0; q < 8; ++q) adam_one(pv[q], mv[q], vv[q], g[q], h);
            #pragma unroll
            for (int q = 0; q < 2; ++q) {
                reinterpret_cast<float4 *>(p)[2 * k + q] = make_float4(pv[4 * q], pv[4 * q + 1], pv[4 * q + 2], pv[4 * q + 3]);
                reinterpret_cast<float4 *>(m)[2 * k + q] = make_float4(mv[4 * q], mv[4 * q + 1], mv[4 * q + 2], mv[4 * q + 3]);
                reinterpret_cast<float4 *>(v)[2 * k + q] = make_float4(vv[4 * q], vv[4 * q + 1], vv[4 * q + 2], vv[4 * q + 3]);
            }
            mc_st_h8(reinterpret_cast<uint4 *>(mc_half_out) + k,
                     make_uint4(pack_h2(pv[0], pv[1]), pack_h2(pv[2], pv[3]), pack_h2(pv[4], pv[5]), pack_h2(pv[6], pv[7])));
        }
    }
}

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// One block, one thread per rank.  epoch[0] += 1; publish it into slot `rank` of every rank's flag array, then wait
// until every slot of the own array has reached it.  A wait longer than timeout_ns raises status[0] and - unless the
// caller asked for the probing mode (fatal == 0, tests) - traps: the step that follows would sum gradients a peer has not
// finished and update the replicas differently on every rank, so a rank that fell out of step must end the job, not
// continue (the trap surfaces as a launch failure at the next synchronisation; the GPU is not left spinning).
__global__ void peer_barrier_kernel(const __grid_constant__ FlagList fl, int world, int rank,
                                    unsigned long long *epoch, int32_t *status, unsigned long long timeout_ns, int fatal) {
    arc_pdl_sync();
    __shared__ unsigned long long target;
    if (threadIdx.x == 0) {
        target = *epoch + 1ull;
        *epoch = target;
        __threadfence_system();          // everything earlier on this stream (incl. stores to peers) before the flags
    }
    __syncthreads();
    const unsigned long long e = target;
    const int w = threadIdx.x;
    if (w < world) {
        st_release_sys(fl.flags[w] + rank, e);
        const unsigned long long *mine = fl.flags[rank] + w;
        const unsigned long long t0 = global_ns();
        while (ld_acquire_sys(mine) < e) {
            if (global_ns() - t0 > timeout_ns) {
                atomicExch(status, 1);
                __threadfence_system();
                if (fatal) __trap();
                break;
            }
            __nanosleep(64);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) __threadfence_system();
}

template <int W>
void launch_adam_peer(unsigned blocks, cudaStream_t st, size_t lo, size_t hi, float *p, float *m, float *v,
                      const PeerList &pl, int world, const AdamHyper &h, const int32_t *step_dev) {
    arc_launch(adam_peer_kernel<W>, blocks, 256, 0, st, lo, hi, p, m, v, pl, world, h, step_dev);
}

}  // namespace

extern "C" {

// Device memory that other processes on the node can map: cudaMalloc + IPC handle (64 bytes, to be sent to the
// peers by any host channel).  No counterpart in the reference (DDP/NCCL own their buffers).
int arcb200_peer_alloc(int64_t bytes, void **ptr_out, unsigned char *handle_out) {
    ARC_REQUIRE(bytes > 0 && ptr_out && handle_out, "peer_alloc: bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    void *p = nullptr;
    ARC_CUDA(cudaMalloc(&p, (size_t)bytes));
    cudaIpcMemHandle_t hd;
    cudaError_t e = cudaIpcGetMemHandle(&hd, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        arcb200_set_error("peer_alloc: cudaIpcGetMemHandle -> %s", cudaGetErrorString(e));
        return ARCB200_ERR_CUDA;
    }
    ARC_CUDA(cudaMemset(p, 0, (size_t)bytes));
    memcpy(handle_out, &hd, 64);
    *ptr_out = p;
    return ARCB200_OK;
}

// Map a peer's arena into this process (current device must be able to reach the owner over NVLink/PCIe P2P).
int arcb200_peer_open(const unsigned char *handle, void **ptr_out) {
    ARC_REQUIRE(handle && ptr_out, "peer_open: null pointer");
    cudaIpcMemHandle_t hd;
    memcpy(&hd, handle, 64);
    void *p = nullptr;
    ARC_CUDA(cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess));
    *ptr_out = p;
    return ARCB200_OK;
}

int arcb200_peer_close(void *ptr) {
    if (ptr) ARC_CUDA(cudaIpcCloseMemHandle(ptr));
    return ARCB200_OK;
}

int arcb200_peer_free(void *ptr) {
    if (ptr) ARC_CUDA(cudaFree(ptr));
    return ARCB200_OK;
}

// Flag barrier between the ranks of a node on `stream` (see the file header).  flags[w]: rank w's flag array
// (>= 16 uint64, zero-initialised, in peer-mapped memory), as mapped in THIS process; epoch: this rank's device
// counter (uint64, starts at 0, private); status: device int32 raised to 1 when a wait exceeded the timeout.
// timeout_ms > 0: that many milliseconds, FATAL (the kernel traps after raising status: a step computed from gradients a peer
// has not finished must not be applied); timeout_ms == 0: the default, ARCB200_PEER_TIMEOUT_MS from the environment or
// 600 000 ms (host work of minutes between steps - checkpoints, evaluation - is normal), fatal; timeout_ms < 0: -timeout_ms
// milliseconds, NOT fatal (status only; for probing).
int arcb200_peer_barrier(void *const *flags, int world, int rank, void *epoch, int32_t *status, int timeout_ms,
                         void *stream) {
    ARC_REQUIRE(flags && epoch && status, "peer_barrier: null pointer");
    ARC_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, "peer_barrier: world %d rank %d out of range (max %d ranks)",
                world, rank, kMaxWorld);
    FlagList fl;
    for (int w = 0; w < kMaxWorld; ++w) fl.flags[w] = w < world ? (unsigned long long *)flags[w] : nullptr;
    for (int w = 0; w < world; ++w) ARC_REQUIRE(fl.flags[w] != nullptr, "peer_barrier: flags[%d] is null", w);
    int fatal = 1;
    if (timeout_ms < 0) { fatal = 0; timeout_ms = -timeout_ms; }
    if (timeout_ms == 0) {
        const char *e = getenv("ARCB200_PEER_TIMEOUT_MS");
        const long v = e ? atol(e) : 0;
        timeout_ms = v > 0 && v < 2000000000L ? (int)v : 600000;
    }
    arc_launch(peer_barrier_kernel, 1, 32, 0, (cudaStream_t)stream, fl, world, rank, (unsigned long long *)epoch, status,
                                                           (unsigned long long)timeout_ms * 1000000ull, fatal);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Fused gradient exchange + Adam + parameter broadcast over elements [lo, hi) of a flat parameter vector.
//   grads[w]    : rank w's gradient vector (fp32, full length, peer-mapped); summed over w = 0..world-1 in that order
//   p, m, v     : this rank's fp32 parameters and Adam moments (full-length vectors; only [lo, hi) is touched)
//   half_out[w] : rank w's fp16 working copy (full length) - the updated slice is stored into every one of them -
//                 or half_out == NULL for no fp16 copy
// Gradients are NOT cleared (peers may still be reading them: clear after the closing barrier).  lo % 4 == 0.
// Arithmetic = arcb200_adam_step on the summed gradient with grad_scale (1/world for DDP's mean).
int arcb200_adam_step_peer(const void *const *grads, void *const *half_out, int world, int64_t lo, int64_t hi,
                           float *p, float *m, float *v, int step, float lr, float beta1, float beta2, float eps,
                           float weight_decay, float grad_scale, const int32_t *step_dev, void *stream) {
    ARC_REQUIRE(world >= 1 && world <= kMaxWorld, "adam_step_peer: world %d out of range (max %d ranks)", world, kMaxWorld);
    ARC_REQUIRE(lo >= 0 && hi >= lo && (lo & 3) == 0, "adam_step_peer: bad range [%lld, %lld)", (long long)lo, (long long)hi);
    if (hi == lo) return ARCB200_OK;
    ARC_REQUIRE(grads && p && m && v && (step >= 1 || step_dev), "adam_step_peer: null pointer or step < 1");
    if (step < 1) step = 1;
    ARC_REQUIRE((((uintptr_t)p | (uintptr_t)m | (uintptr_t)v) & 15u) == 0, "adam_step_peer: misaligned buffer");
    PeerList pl;
    for (int w = 0; w < kMaxWorld; ++w) {
        pl.grad[w] = w < world ? (const float *)grads[w] : nullptr;
        pl.half_out[w] = (w < world && half_out) ? (__half *)half_out[w] : nullptr;
    }
    for (int w = 0; w < world; ++w) {
        ARC_REQUIRE(pl.grad[w] && ((uintptr_t)pl.grad[w] & 15u) == 0, "adam_step_peer: grads[%d] null or misaligned", w);
        ARC_REQUIRE(!half_out || (pl.half_out[w] && ((uintptr_t)pl.half_out[w] & 7u) == 0),
                    "adam_step_peer: half_out[%d] null or misaligned", w);
    }
    const AdamHyper h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, grad_scale, step);
    const int64_t n = hi - lo;
    long long blocks = (n / 4 + 255) / 256;
    const long long cap = (long long)arc_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    cudaStream_t st = (cudaStream_t)stream;
    switch (world) {
        case 1: launch_adam_peer<1>((unsigned)blocks, st, lo, hi, p, m, v, pl, world, h, step_dev); break;
        case 2: launch_adam_peer<2>((unsigned)blocks, st, lo, hi, p, m, v, pl, world, h, step_dev); break;
        case 4: launch_adam_peer<4>((unsigned)blocks, st, lo, hi, p, m, v, pl, world, h, step_dev); break;
        case 8: launch_adam_peer<8>((unsigned)blocks, st, lo, hi, p, m, v, pl, world, h, step_dev); break;
        default: launch_adam_peer<0>((unsigned)blocks, st, lo, hi, p, m, v, pl, world, h, step_dev); break;
    }
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Loss-scaled fp16 copy of the gradient for the exchange: g16[i] = half(g32[i] * scale) for i outside [keep_lo, keep_hi)
// (multiples of 8: the slice this rank owns, which it reads in fp32 itself); g32 is cleared where it was converted.
int arcb200_grad_to_half(float *g32, void *g16, int64_t n, int64_t keep_lo, int64_t keep_hi, float scale, void *stream) {
    ARC_REQUIRE(n >= 0 && keep_lo >= 0 && keep_hi >= keep_lo && (keep_lo & 7) == 0 && ((keep_hi & 7) == 0 || keep_hi >= n),
                "grad_to_half: bad range");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(g32 && g16 && (((uintptr_t)g32 | (uintptr_t)g16) & 15u) == 0, "grad_to_half: null or misaligned pointer");
    long long blocks = (n / 8 + 255) / 256;
    const long long cap = (long long)arc_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    arc_launch(grad_to_half_kernel, (unsigned)blocks, 256, 0, (cudaStream_t)stream, (size_t)n, (size_t)keep_lo, (size_t)keep_hi, g32,
               (__half *)g16, scale);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// arcb200_adam_step_peer with the half-precision exchange: the gradient of [lo, hi) (lo % 8 == 0) is g_own (this rank's
// fp32 vector, cleared over the range afterwards) + sum over the OTHER ranks w of grads16[w] / scale (their fp16 copies from
// arcb200_grad_to_half, peer loads), times grad_scale.  Everything else as arcb200_adam_step_peer.
int arcb200_adam_step_peer16(float *g_own, const void *const *grads16, void *const *half_out, int world, int rank,
                             int64_t lo, int64_t hi, float *p, float *m, float *v, float scale, int step, float lr,
                             float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                             const int32_t *step_dev, void *stream) {
    ARC_REQUIRE(world >= 1 && world <= kMaxWorld && rank >= 0 && rank < world, "adam_step_peer16: world %d rank %d out of range", world, rank);
    ARC_REQUIRE(lo >= 0 && hi >= lo && (lo & 7) == 0, "adam_step_peer16: bad range [%lld, %lld)", (long long)lo, (long long)hi);
    if (hi == lo) return ARCB200_OK;
    ARC_REQUIRE(g_own && grads16 && half_out && p && m && v && scale > 0.f && (step >= 1 || step_dev), "adam_step_peer16: null pointer, scale or step");
    if (step < 1) step = 1;
    ARC_REQUIRE((((uintptr_t)p | (uintptr_t)m | (uintptr_t)v | (uintptr_t)g_own) & 15u) == 0, "adam_step_peer16: misaligned buffer");
    Peer16List pl;
    for (int w = 0; w < kMaxWorld; ++w) {
        pl.grad16[w] = (w < world && w != rank) ? (const __half *)grads16[w] : nullptr;
        pl.half_out[w] = w < world ? (__half *)half_out[w] : nullptr;
    }
    for (int w = 0; w < world; ++w) {
        ARC_REQUIRE(w == rank || (pl.grad16[w] && ((uintptr_t)pl.grad16[w] & 15u) == 0), "adam_step_peer16: grads16[%d] null or misaligned", w);
        ARC_REQUIRE(pl.half_out[w] && ((uintptr_t)pl.half_out[w] & 15u) == 0, "adam_step_peer16: half_out[%d] null or misaligned", w);
    }
    const AdamHyper h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, grad_scale, step);
    long long blocks = ((hi - lo) / 8 + 255) / 256;
    const long long cap = (long long)arc_sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    cudaStream_t st = (cudaStream_t)stream;
    const float inv = 1.0f / scale;
    switch (world) {
        case 2: launch_adam_peer16<2>((unsigned)blocks, st, lo, hi, p, m, v, g_own, pl, world, rank, inv, h, step_dev); break;
        case 4: launch_adam_peer16<4>((unsigned)blocks, st, lo, hi, p, m, v, g_own, pl, world, rank, inv, h, step_dev); break;
        case 8: launch_adam_peer16<8>((unsigned)blocks, st, lo, hi, p, m, v, g_own, pl, world, rank, inv, h, step_dev); break;
        default: launch_adam_peer16<0>((unsigned)blocks, st, lo, hi, p, m, v, g_own, pl, world, rank, inv, h, step_dev); break;
    }
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// The same step with the reduce and the broadcast done in the NVSwitch: mc_grad16 / mc_half_out are multicast addresses (one
// mapping that reaches the same offset of every rank's arena) of the fp16 gradient copies (arcb200_grad_to_half over the WHOLE
// vector: keep_lo = keep_hi = 0, so the own gradient takes the same fp16 rounding as the peers') and of the fp16 working table.
// lo and hi are multiples of 8.  Same ordering contract as arcb200_adam_step_peer: a barrier before and after.
int arcb200_adam_step_mc16(const void *mc_grad16, void *mc_half_out, int64_t lo, int64_t hi, float *p, float *m, float *v,
                           float scale, int step, float lr, float beta1, float beta2, float eps, float weight_decay,
                           float grad_scale, const int32_t *step_dev, void *stream) {
    ARC_REQUIRE(lo >= 0 && hi >= lo && (lo & 7) == 0 && (hi & 7) == 0, "adam_step_mc16: bad range [%lld, %lld)", (long long)lo, (long long)hi);
    if (hi == lo) return ARCB200_OK;
    ARC_REQUIRE(mc_grad16 && mc_half_out && p && m && v && scale > 0.f && (step >= 1 || step_dev), "adam_step_mc16: null pointer, scale or step");
    if (step < 1) step = 1;
    ARC_REQUIRE((((uintptr_t)p | (uintptr_t)m | (uintptr_t)v | (uintptr_t)mc_grad16 | (uintptr_t)mc_half_out) & 15u) == 0,
                "adam_step_mc16: misaligned buffer");
    const AdamHyper h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, grad_scale, step);
    long long blocks = ((hi - lo) / 8 + 255) / 256;
    const long long cap = (long long)arc_sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    arc_launch(adam_mc16_kernel, (unsigned)blocks, 256, 0, (cudaStream_t)stream, (size_t)lo, (size_t)hi, p, m, v,
               (const __half *)mc_grad16, (__half *)mc_half_out, 1.0f / scale, h, step_dev);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
