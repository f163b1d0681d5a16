// Measurement aid, not on the hot path: random-gather bandwidth of an L2-resident table, the roof the hash-grid encode
// kernels are measured against ("L2 bandwidth to be measured on the box with a gather micro-benchmark").
// No counterpart in the reference.
//
// Every thread issues `unroll` independent loads of `req_bytes` (4, 8 or 16) at pseudo-random entries of the table per trip
// (a 32-bit mix of the request number, so that neighbouring lanes hit unrelated sectors like hashed grid levels do) and
// folds them into a checksum that is written once per thread, so the loads cannot be optimised away.
#include "common.cuh"

namespace {

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

template <typename V> __device__ __forceinline__ uint32_t fold(const V &v);
template <> __device__ __forceinline__ uint32_t fold<uint32_t>(const uint32_t &v) { return v; }
template <> __device__ __forceinline__ uint32_t fold<uint2>(const uint2 &v) { return v.x ^ v.y; }
template <> __device__ __forceinline__ uint32_t fold<uint4>(const uint4 &v) { return v.x ^ v.y ^ v.z ^ v.w; }

template <typename V, int UNROLL>
__global__ void __launch_bounds__(256) probe_gather_kernel(const V *__restrict__ table, uint32_t n_entries, int trips,
                                                            uint32_t seed, uint32_t *__restrict__ sink) {
    arc_pdl_sync();
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t acc = 0;
    uint32_t req = (tid * (uint32_t)trips) * UNROLL + seed;
    for (int t = 0; t < trips; ++t) {
        V v[UNROLL];
        #pragma unroll
        for (int k = 0; k < UNROLL; ++k) {
            const uint32_t idx = __umulhi(mix32(req + k), n_entries);      // uniform in [0, n_entries)
            v[k] = __ldg(table + idx);
        }
        #pragma unroll
        for (int k = 0; k < UNROLL; ++k) acc ^= fold<V>(v[k]);
        req += UNROLL;
    }
    sink[tid] = acc;
}

}  // namespace

extern "C" {

// table: n_bytes of device memory (any content); req_bytes in {4, 8, 16}; n_threads * trips * 8 requests are issued.
// sink: n_threads uint32.  Returns the number of requests issued (negative on error).
long long arcb200_probe_gather(const void *table, int64_t n_bytes, int req_bytes, int n_threads, int trips, uint32_t seed,
                               uint32_t *sink, void *stream) {
    ARC_REQUIRE(table && sink && n_bytes >= 1024 && n_threads > 0 && trips > 0, "probe_gather: bad argument");
    ARC_REQUIRE(req_bytes == 4 || req_bytes == 8 || req_bytes == 16, "probe_gather: req_bytes must be 4, 8 or 16");
    ARC_REQUIRE(n_bytes / req_bytes < 0xffffffffLL, "probe_gather: table too large");
    n_threads = (n_threads + 255) / 256 * 256;
    const uint32_t n_entries = (uint32_t)(n_bytes / req_bytes);
    const dim3 grid(n_threads / 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (req_bytes == 4) arc_launch(probe_gather_kernel<uint32_t, 8>, grid, 256, 0, st, (const uint32_t *)table, n_entries, trips, seed, sink);
    else if (req_bytes == 8) arc_launch(probe_gather_kernel<uint2, 8>, grid, 256, 0, st, (const uint2 *)table, n_entries, trips, seed, sink);
    else arc_launch(probe_gather_kernel<uint4, 8>, grid, 256, 0, st, (const uint4 *)table, n_entries, trips, seed, sink);
    ARC_LAUNCH_CHECK();
    return (long long)n_threads * trips * 8;
}

}  // extern "C"
