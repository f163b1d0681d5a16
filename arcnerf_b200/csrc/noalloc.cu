// arcb200_march_packed_static: the packed occupancy march for callers that have nowhere to get scratch memory from.
//
// arcb200_march_packed needs a scratch buffer (one row of candidate sample distances per ray plus the tile sums of
// the offset scan) and, like every entry point of this library, expects the caller to bring it.  Two kinds of host
// cannot do that:
//   * a sampler called from inside another engine's per-frame callback, where a cudaMalloc is off the table because
//     its implicit device-wide synchronisation would stall every other stream the engine is running;
//   * a host that owns all device memory through its own arena and hands a plug-in nothing but the input and output
//     buffers of the call - there is no hook through which the plug-in side could obtain a raw device allocation.
// For them the scratch lives here, in a fixed file-scope device array that is part of the library image: it costs
// kStaticScratchBytes of device memory on every GPU that loads the library, from load until the process exits, and
// it bounds the problem size (see the limits below).  The call itself allocates nothing: no cudaMalloc/cudaFree and
// no allocating library routine is reached from it.
//
// All calls on a device share the one array, and it keeps whatever the previous call left in it.  The march does
// not depend on stale contents (every scratch word it reads was written earlier in the same call; nothing has to
// start at zero), but two calls whose kernels overlap on the GPU - different streams, or different host threads
// feeding different streams - would write the same rows and race.  Keep calls on one device ordered.
//
// The first call on each device looks the array's device address up (cudaGetSymbolAddress, with that device
// current) and caches it; in host code the array's name is only a handle for that lookup, not an address a kernel
// could use.  Make that first call outside any stream capture.
#include "common.cuh"
#include <mutex>

extern "C" int64_t arcb200_march_packed_scratch_bytes(int64_t n_rays, int n_pts);
extern "C" int arcb200_march_packed(int bit, const float *rays_o, const float *rays_d, const float *near, const float *far,
                                    int n_pts, float dt, const float *aabb, int n_grid, const uint8_t *field,
                                    float near_distance, int64_t n_rays, uint64_t rng_state, uint64_t rng_inc,
                                    void *scratch, int32_t *counts, int32_t *offsets, int32_t *meta, int64_t capacity,
                                    float *t_out, float *xyz_out, int32_t *ray_id_out, float *unit_out, void *stream);

namespace {

// Limits: n_rays * n_pts <= 4 Mi candidate slots (e.g. 4096 rays x 1024 samples, the NGP training batch) and
// n_rays <= 256 Ki.  16 MiB of rows + 4 KiB for the scan's tile sums.
constexpr int64_t kMaxSlots = 4ll << 20;
constexpr int64_t kMaxRays = 256ll << 10;
constexpr size_t kStaticScratchBytes = (size_t)kMaxSlots * 4 + 4096;

static __device__ __align__(256) unsigned char g_march_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
void *g_scratch_addr[kMaxDevices] = {nullptr};
std::mutex g_scratch_mutex;

// Device address of g_march_scratch on the current device, looked up once per device.
int scratch_address(void **out) {
    int dev = -1;
    ARC_CUDA(cudaGetDevice(&dev));
    ARC_REQUIRE(dev >= 0 && dev < kMaxDevices, "march_packed_static: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (g_scratch_addr[dev] == nullptr) {
        void *p = nullptr;
        ARC_CUDA(cudaGetSymbolAddress(&p, g_march_scratch));
        ARC_REQUIRE(p != nullptr, "march_packed_static: cudaGetSymbolAddress returned a null address");
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return ARCB200_OK;
}

}  // namespace

extern "C" {

// Largest n_rays * n_pts the static scratch can serve (also: n_rays <= 262144).
int64_t arcb200_march_packed_static_max_slots(void) { return kMaxSlots; }

// Same march, same outputs as arcb200_march_packed, scratch taken from the library's static device array.
int arcb200_march_packed_static(int bit, const float *rays_o, const float *rays_d, const float *near, const float *far,
                                int n_pts, float dt, const float *aabb, int n_grid, const uint8_t *field,
                                float near_distance, int64_t n_rays, uint64_t rng_state, uint64_t rng_inc,
                                int32_t *counts, int32_t *offsets, int32_t *meta, int64_t capacity, float *t_out,
                                float *xyz_out, int32_t *ray_id_out, float *unit_out, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_pts > 0, "march_packed_static: n_rays must be >= 0 and n_pts positive");
    if (n_rays > kMaxRays || n_rays * (int64_t)n_pts > kMaxSlots ||
        (size_t)arcb200_march_packed_scratch_bytes(n_rays, n_pts) > kStaticScratchBytes) {
        arcb200_set_error("march_packed_static: %lld rays x %d samples exceeds the static scratch (at most %lld rays and "
                          "%lld ray*sample slots); use arcb200_march_packed with caller-provided scratch",
                          (long long)n_rays, n_pts, (long long)kMaxRays, (long long)kMaxSlots);
        return ARCB200_ERR_LIMIT;
    }
    void *scratch = nullptr;
    const int rc = scratch_address(&scratch);
    if (rc != ARCB200_OK) return rc;
    return arcb200_march_packed(bit, rays_o, rays_d, near, far, n_pts, dt, aabb, n_grid, field, near_distance, n_rays,
                                rng_state, rng_inc, scratch, counts, offsets, meta, capacity, t_out, xyz_out, ray_id_out,
                                unit_out, stream);
}

}  // extern "C"
