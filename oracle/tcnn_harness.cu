// TEST INFRASTRUCTURE ONLY -- never linked into or called from the product path.
//
// C entry points around the REFERENCE's own tiny-cuda-nn (vendored submodule of ArcNerf,
// /root/reference/include/tiny-cuda-nn @ 8e6e242f), built for sm_100 from the sources where they lie
// (oracle/Makefile target `tcnn`; nothing is copied into this repository).  The calls below are the ones
// the reference's torch binding makes (bindings/torch/tinycudann/bindings.cpp:78-241 Module::fwd / bwd /
// bwd_bwd_input / initial_params) on tcnn::cpp::Module (include/tiny-cuda-nn/cpp_api.h:72-100), with raw
// device pointers in place of torch tensors.  Used by tests/golden/make_golden_tcnn.py on the GPU box to
// write the golden vectors that pin oracle/fields.py and the CUDA path (hash grid, SH, FullyFusedMLP), and
// to time the reference's kernels on the same B200.
#include <tiny-cuda-nn/cpp_api.h>
#include <tiny-cuda-nn/encodings/grid.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

using namespace tcnn::cpp;

__global__ void device_scales_kernel(int n_levels, float log2_per_level_scale, uint32_t base_resolution, float *out) {
    const int l = threadIdx.x;
    if (l < n_levels) out[l] = tcnn::grid_scale((uint32_t)l, log2_per_level_scale, base_resolution);
}

namespace {
thread_local std::string g_err;
template <class F> int guarded(F f) {
    try { f(); return 0; }
    catch (const std::exception &e) { g_err = e.what(); return -1; }
    catch (...) { g_err = "unknown exception"; return -1; }
}
}  // namespace

extern "C" {

const char *tcnnref_last_error() { return g_err.c_str(); }

// bindings.cpp:318-330 create_encoding / create_network
void *tcnnref_create_encoding(int n_input_dims, const char *encoding_json, int fp16) {
    Module *m = nullptr;
    if (guarded([&] { m = create_encoding((uint32_t)n_input_dims, json::parse(encoding_json), fp16 ? Fp16 : Fp32); })) return nullptr;
    return m;
}
void *tcnnref_create_network(int n_input_dims, int n_output_dims, const char *network_json) {
    Module *m = nullptr;
    if (guarded([&] { m = create_network((uint32_t)n_input_dims, (uint32_t)n_output_dims, json::parse(network_json)); })) return nullptr;
    return m;
}
void tcnnref_destroy(void *m) { delete (Module *)m; }

long long tcnnref_n_params(void *m) { return (long long)((Module *)m)->n_params(); }
int tcnnref_n_input_dims(void *m) { return (int)((Module *)m)->n_input_dims(); }
int tcnnref_n_output_dims(void *m) { return (int)((Module *)m)->n_output_dims(); }
int tcnnref_param_is_fp16(void *m) { return ((Module *)m)->param_precision() == Fp16; }
int tcnnref_output_is_fp16(void *m) { return ((Module *)m)->output_precision() == Fp16; }
int tcnnref_batch_granularity() { return (int)batch_size_granularity(); }

// bindings.cpp:243-247: fp32 initial parameters from tcnn's own pcg32(seed) stream
int tcnnref_initial_params(void *m, unsigned long long seed, float *params_fp32) {
    return guarded([&] { ((Module *)m)->initialize_params((size_t)seed, params_fp32); });
}

int tcnnref_inference(void *m, void *stream, unsigned n, const float *input, void *output, void *params) {
    return guarded([&] { ((Module *)m)->inference((cudaStream_t)stream, n, input, output, params); });
}

// forward that keeps a context for the backward passes (bindings.cpp:102-106); returns an owning handle
void *tcnnref_forward(void *m, void *stream, unsigned n, const float *input, void *output, void *params,
                      int prepare_input_gradients) {
    Context *ctx = new Context();
    if (guarded([&] { *ctx = ((Module *)m)->forward((cudaStream_t)stream, n, input, output, params, prepare_input_gradients != 0); })) {
        delete ctx;
        return nullptr;
    }
    return ctx;
}
void tcnnref_free_context(void *ctx) { delete (Context *)ctx; }

// bindings.cpp:111-170
int tcnnref_backward(void *m, void *stream, void *ctx, unsigned n, float *dL_dinput, const void *dL_doutput,
                     void *dL_dparams, const float *input, const void *output, const void *params) {
    return guarded([&] {
        ((Module *)m)->backward((cudaStream_t)stream, *(Context *)ctx, n, dL_dinput, dL_doutput, dL_dparams, input, output, params);
    });
}

// bindings.cpp:172-241
int tcnnref_backward_backward_input(void *m, void *stream, void *ctx, unsigned n, const float *dL_ddLdinput,
                                    const float *input, const void *dL_doutput, void *dL_dparams, void *dL_ddLdoutput,
                                    float *dL_dinput, const void *params) {
    return guarded([&] {
        ((Module *)m)->backward_backward_input((cudaStream_t)stream, *(Context *)ctx, n, dL_ddLdinput, input, dL_doutput,
                                               dL_dparams, dL_ddLdoutput, dL_dinput, params);
    });
}

void tcnnref_free_temporary_memory() { free_temporary_memory(); }

// The level scales as kernel_grid computes them per thread (grid.h:270-271 -> grid_scale, :195-200): the device's exp2f is not
// bit-identical to the host's, so the golden file records what the kernels actually used.
void tcnnref_device_grid_scales(int n_levels, float log2_per_level_scale, int base_resolution, float *out_dev) {
    device_scales_kernel<<<1, 64>>>(n_levels, log2_per_level_scale, (uint32_t)base_resolution, out_dev);
}

}  // extern "C"
