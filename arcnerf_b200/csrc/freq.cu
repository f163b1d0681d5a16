// Frequency (positional) encoding of the vanilla-NeRF configuration, forward and input gradient.
// Reference: arcnerf/models/base_modules/encoding/freq_encoder.py:10-88 - out = [x | sin(x f_0) | cos(x f_0) | sin(x f_1) |
// ...] with f_k = 2^k (log sampling) or linspace(1, 2^(n-1), n); 3 inputs x 10 bands -> 63 columns for positions, 4 bands
// -> 27 for view directions.  The reference builds it from 2 n torch ops and a concatenation (2 n + 1 launches and as many
// (N, d) temporaries); here one streaming kernel writes the row once: 4 d B in, 4 d (2 n + 1) B out per point.
// The frequency bands come from the caller (device array), so both sampling rules of the reference are covered.
#include "common.cuh"

namespace {

constexpr int kMaxFreqs = 32;
struct FreqBands { float f[kMaxFreqs]; };

// one thread per (point, input channel): its 2 n + 1 outputs are d-strided in the row
__global__ void freq_fwd_kernel(size_t n_elem, int d, const float *__restrict__ x, const __grid_constant__ FreqBands fb,
                                int n_freqs, int include_input, float *__restrict__ out, int64_t ld_out) {
    arc_pdl_sync();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_elem; i += (size_t)gridDim.x * blockDim.x) {
        const size_t row = i / d;
        const int c = (int)(i - row * d);
        const float v = x[i];
        float *o = out + row * ld_out + c;
        if (include_input) { *o = v; o += d; }
        for (int k = 0; k < n_freqs; ++k) {
            float s, co;
            sincosf(v * fb.f[k], &s, &co);
            o[0] = s;
            o[d] = co;
            o += 2 * d;
        }
    }
}

// g_x = g[x part] + sum_k f_k (cos(x f_k) g[sin part] - sin(x f_k) g[cos part])
__global__ void freq_bwd_kernel(size_t n_elem, int d, const float *__restrict__ x, const __grid_constant__ FreqBands fb,
                                int n_freqs, int include_input, const float *__restrict__ g_out, int64_t ld_g,
                                float *__restrict__ g_x) {
    arc_pdl_sync();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_elem; i += (size_t)gridDim.x * blockDim.x) {
        const size_t row = i / d;
        const int c = (int)(i - row * d);
        const float v = x[i];
        const float *g = g_out + row * ld_g + c;
        float acc = 0.0f;
        if (include_input) { acc = *g; g += d; }
        for (int k = 0; k < n_freqs; ++k) {
            float s, co;
            sincosf(v * fb.f[k], &s, &co);
            acc += fb.f[k] * (co * g[0] - s * g[d]);
            g += 2 * d;
        }
        g_x[i] = acc;
    }
}

unsigned freq_blocks(size_t n_elem) {
    long long b = (long long)((n_elem + 255) / 256);
    const long long cap = (long long)arc_sm_count() * 16;
    return (unsigned)(b < cap ? (b < 1 ? 1 : b) : cap);
}

}  // namespace

extern "C" {

// x (n, d) fp32 contiguous; freqs: HOST array of n_freqs (<= 32) bands; out (n, >= d*(2*n_freqs+include_input)) fp32 with
// leading dimension ld_out.  Replaces FreqEmbedder.forward (freq_encoder.py:66-88), periodic functions (sin, cos).
int arcb200_freq_embed_fwd(const float *x, int d, const float *freqs, int n_freqs, int include_input, float *out,
                           int64_t ld_out, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0 && d >= 1, "freq_embed_fwd: sizes out of range");
    ARC_REQUIRE(n_freqs >= 0 && n_freqs <= kMaxFreqs, "freq_embed_fwd: at most %d frequency bands", kMaxFreqs);
    ARC_REQUIRE(n_freqs > 0 || include_input, "freq_embed_fwd: nothing to write (no bands and no input copy)");
    ARC_REQUIRE(ld_out >= (int64_t)d * (2 * n_freqs + (include_input ? 1 : 0)), "freq_embed_fwd: leading dimension too small");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(x && out && (freqs || n_freqs == 0), "freq_embed_fwd: null pointer");
    FreqBands fb;
    for (int k = 0; k < kMaxFreqs; ++k) fb.f[k] = k < n_freqs ? freqs[k] : 0.0f;
    arc_launch(freq_fwd_kernel, freq_blocks((size_t)n * d), 256, 0, (cudaStream_t)stream, (size_t)n * d, d, x, fb, n_freqs,
               include_input, out, ld_out);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Gradient of the above w.r.t. x (what autograd derives for the reference; needed for surface normals of SDF models).
int arcb200_freq_embed_bwd(const float *x, int d, const float *freqs, int n_freqs, int include_input, const float *g_out,
                           int64_t ld_g, float *g_x, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0 && d >= 1, "freq_embed_bwd: sizes out of range");
    ARC_REQUIRE(n_freqs >= 0 && n_freqs <= kMaxFreqs, "freq_embed_bwd: at most %d frequency bands", kMaxFreqs);
    ARC_REQUIRE(ld_g >= (int64_t)d * (2 * n_freqs + (include_input ? 1 : 0)), "freq_embed_bwd: leading dimension too small");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(x && g_out && g_x && (freqs || n_freqs == 0), "freq_embed_bwd: null pointer");
    FreqBands fb;
    for (int k = 0; k < kMaxFreqs; ++k) fb.f[k] = k < n_freqs ? freqs[k] : 0.0f;
    arc_launch(freq_bwd_kernel, freq_blocks((size_t)n * d), 256, 0, (cudaStream_t)stream, (size_t)n * d, d, x, fb, n_freqs,
               include_input, g_out, ld_g, g_x);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
