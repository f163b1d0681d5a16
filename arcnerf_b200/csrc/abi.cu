// libarcb200: process-level pieces of the C ABI (error text, version, device queries, pcg32 host helpers).
#include "common.cuh"
#include <stdlib.h>
#include "pcg32.cuh"
#include <stdarg.h>
#include <string.h>

static thread_local char g_err[512] = "";

void arcb200_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int arc_sm_count() {
    static int sm[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (sm[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        sm[dev] = v;
    }
    return sm[dev];
}

bool arc_pdl_enabled() {
    static int on = -1;
    if (on < 0) {
        const char *e = getenv("ARCB200_PDL");
        on = (e && e[0] == '0') ? 0 : 1;
    }
    return on != 0;
}

extern "C" {

const char *arcb200_last_error(void) { return g_err; }

int arcb200_version(void) { return 100; }

// Number of CUDA devices the library can see; < 0 when the runtime cannot initialise (no driver / no GPU).
int arcb200_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        arcb200_set_error("cudaGetDeviceCount -> %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return ARCB200_ERR_CUDA;
    }
    return n;
}

// Host-side pcg32 bookkeeping so that callers can keep the per-module generator of the reference
// (`static pcg32 rng{9121}`, arcnerf/ops/include/common.h:21-23) as two plain integers.
void arcb200_rng_init(uint64_t seed, uint64_t *state, uint64_t *inc) {
    Pcg32 r(seed);
    *state = r.state;
    *inc = r.inc;
}

void arcb200_rng_advance(uint64_t *state, uint64_t inc, int64_t delta) {
    Pcg32 r;
    r.state = *state;
    r.inc = inc;
    r.advance((uint64_t)delta);
    *state = r.state;
}

}  // extern "C"
