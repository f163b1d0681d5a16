// Small fully-fused MLP (tiny-cuda-nn "FullyFusedMLP": no biases, 16-multiple widths) on the Blackwell tensor
// cores: tcgen05.mma with fp16 operands in shared memory, fp32 accumulators in tensor memory (TMEM), weights
// staged once per CTA by the TMA engine (1-D bulk copy), activations never leaving the SM between layers.
//
// Semantics: /root/reference/include/tiny-cuda-nn/src/fully_fused_mlp.cu (kernel_mlp_fused :499-557, layer
// order and weight layout :658-671, backward :150-259 + split-k weight gradients :781-822), as driven by
// ArcNerf's FusedMLPGeoNet / FusedMLPRadianceNet (tcnn_fusedmlp_module.py:66-77,162-173):
//   y = act_out( W_last . act( ... act( W_0 . x ) ) ), weights row-major (out x in), input layer first,
//   the last matrix padded to 16 rows; inputs narrower than a multiple of 16 are padded with ONES
//   (encodings/identity.h:63-64), so the pad columns of W_0 act as trainable biases.
// Precision: fp16 weights and activations like the reference, but fp32 accumulation (the reference accumulates
// in fp16 inside wmma, fully_fused_mlp.cu:68).
//
// Kernel shape: one CTA = 128 threads = one 128-row tile at a time (thread r owns row r = TMEM lane r);
// persistent grid-stride loop over tiles.  Per layer: one elected thread issues K/16 tcgen05.mma
// (M=128, N=layer width), commits to an mbarrier; all threads then pull their accumulator row out of TMEM
// (tcgen05.ld 32x32b), apply the activation and write the fp16 row back into the shared-memory operand tile of
// the next layer.
// Backward, two kernels:
//   * mlp_bwd_v2_kernel (mlp_bwd_v2.cuh) - the training path of the NGP step: fp16 inputs, ReLU hidden layers of width 64 and
//     hidden activations SAVED by the forward (`act_save`, reloaded into the operand tiles by bulk copies, nothing is
//     re-computed).  One CTA per SM with four tile slots and a control warp that issues every MMA; weight-gradient
//     accumulators shared by all slots in TMEM.  arcb200_mlp_bwd_ex selects it whenever those conditions hold.
//   * mlp_bwd_kernel (below) - every other configuration reachable through the plugin API (fp32 inputs, other widths or
//     activations, no saved activations): one tile per CTA, re-computes the forward activations on the tensor cores unless
//     they were saved, chains the data gradients the same way and accumulates the weight gradients dW = G^T . H over ALL
//     tiles of the CTA in TMEM (batch = contraction dimension, operands read MN-major from the very same tiles), flushing
//     them once per CTA with fp32 atomics.
#include "common.cuh"
#include "tc05.cuh"
#include "adam.cuh"
#include "sh.cuh"
#include "grid_math.cuh"
#include <string.h>
#include <stdlib.h>

namespace {

constexpr int kMaxMM = 4;            // matmuls per network (hidden layers + 1)
constexpr int kRows = 128;           // rows per tile
constexpr uint32_t kTileLBO = kRows * 16;   // column-group stride of a 128-row tile (bytes)

enum Act { ACT_NONE = 0, ACT_RELU = 1, ACT_SIGMOID = 2, ACT_SOFTPLUS = 3, ACT_EXP = 4 };

struct MlpCfg {
    int n_mm;              // number of matmuls
    int in_dim, in_pad;    // real and 16-padded input width
    int width;             // hidden width (64 or 128 for training; 16..128 forward)
    int out_dim, out_pad;  // real and 16-padded output width
    int hidden_act, out_act;
    int K[kMaxMM], N[kMaxMM];
    uint32_t w_off[kMaxMM];   // byte offset of layer l's weight tile in the packed blob
    uint32_t w_bytes;         // blob size
    uint32_t p_off[kMaxMM];   // element offset of layer l in the flat parameter vector
};

// 16 values at a time with the activation chosen once (uniform branch), not per element
__device__ __forceinline__ void act16_fwd(int a, const uint32_t (&r)[16], float (&v)[16]) {
    switch (a) {
        case ACT_RELU:
            #pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = fmaxf(__uint_as_float(r[k]), 0.f);
            break;
        case ACT_SIGMOID:
            #pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = 1.0f / (1.0f + __expf(-__uint_as_float(r[k])));
            break;
        case ACT_SOFTPLUS:
            #pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = __logf(__expf(__uint_as_float(r[k]) * 10.0f) + 1.0f) / 10.0f;
            break;
        case ACT_EXP:
            #pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = __expf(__uint_as_float(r[k]));
            break;
        default:
            #pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = __uint_as_float(r[k]);
    }
}
// v[k] *= act'(.) expressed through the activation output y[k]
__device__ __forceinline__ void act16_bwd_mul(int a, float (&v)[16], const float (&y)[16]) {
    switch (a) {
        case ACT_RELU:
            #pragma unroll
            for (int k = 0; k < 16; ++k) v[k] = y[k] > 0.f ? v[k] : 0.f;
            break;
        case ACT_SIGMOID:
            #pragma unroll
            for (int k = 0; k < 16; ++k) v[k] *= y[k] * (1.0f - y[k]);
            break;
        case ACT_SOFTPLUS:
            #pragma unroll
            for (int k = 0; k < 16; ++k) v[k] *= 1.0f - __expf(-y[k] * 10.0f);
            break;
        case ACT_EXP:
            #pragma unroll
            for (int k = 0; k < 16; ++k) v[k] *= y[k];
            break;
        default: break;
    }
}

// Visit this thread's accumulator row in 16-column blocks: fn(block index, 16 raw fp32 words).  Four blocks are
// handled per trip; the TMEM loads of the second pair are in flight while the first pair is being processed.
template <class Fn>
__device__ __forceinline__ void for_each_acc_block(uint32_t taddr, int nblk, Fn fn) {
    for (int cb0 = 0; cb0 < nblk; cb0 += 4) {
        const int nb = min(4, nblk - cb0);
        uint32_t r0[16], r1[16], r2[16], r3[16];
        const uint32_t a = taddr + (uint32_t)cb0 * 16;
        tc05::tmem_ld16(a, r0);
        if (nb > 1) tc05::tmem_ld16(a + 16, r1);
        tc05::tmem_ld_wait();
        if (nb > 2) tc05::tmem_ld16(a + 32, r2);
        if (nb > 3) tc05::tmem_ld16(a + 48, r3);
        fn(cb0, r0);
        if (nb > 1) fn(cb0 + 1, r1);
        if (nb > 2) {
            tc05::tmem_ld_wait();
            fn(cb0 + 2, r2);
            if (nb > 3) fn(cb0 + 3, r3);
        }
    }
}

template <typename T> __device__ __forceinline__ float ld_f32(const T *p);
template <> __device__ __forceinline__ float ld_f32<float>(const float *p) { return __ldg(p); }
template <> __device__ __forceinline__ float ld_f32<__half>(const __half *p) { return __half2float(__ldg(p)); }
template <typename T> __device__ __forceinline__ void st_val(T *p, float v);
template <> __device__ __forceinline__ void st_val<float>(float *p, float v) { *p = v; }
template <> __device__ __forceinline__ void st_val<__half>(__half *p, float v) { *p = __float2half_rn(v); }

// 16-byte word of row `r`, column group `cg` in a 128-row core-matrix tile
__device__ __forceinline__ uint4 *tile_word(uint8_t *tile, int r, int cg) {
    return reinterpret_cast<uint4 *>(tile + (size_t)cg * kTileLBO + (size_t)r * 16);
}

// Rows [row0, row0+128) of X (fp32 or fp16, leading dimension ldx) -> fp16 operand tile.  Columns >= in_dim are 1.0
// (the reference's identity-encoding padding), rows beyond n are zero.  The whole CTA cooperates: 16-byte chunks
// are visited in (row, column-group) order, i.e. in the order they lie in global memory, so a warp reads
// consecutive addresses; `vec` (set by the host: in_dim and ldx multiples of 8, 16-byte aligned base) selects the
// vector path, otherwise every element is loaded on its own.
template <typename TIN>
__device__ __forceinline__ void load_input_tile(uint8_t *tile, const TIN *__restrict__ X, int64_t ldx, int64_t row0,
                                                int64_t n, int in_dim, int in_pad, bool vec, int tid) {
    const int ncg = in_pad >> 3;
    for (int q = tid; q < kRows * ncg; q += kRows) {
        const int r = q / ncg, cg = q - r * ncg;
        const int64_t grow = row0 + r;
        const int c0 = cg * 8;
        uint4 w;
        if (grow >= n) {
            w = make_uint4(0u, 0u, 0u, 0u);
        } else if (c0 >= in_dim) {
            w = make_uint4(0x3C003C00u, 0x3C003C00u, 0x3C003C00u, 0x3C003C00u);   // eight fp16 ones
        } else if (vec) {
            if (sizeof(TIN) == 2) {
                w = __ldg(reinterpret_cast<const uint4 *>(X + grow * ldx + c0));
            } else {
                const float4 a = __ldg(reinterpret_cast<const float4 *>(X + grow * ldx + c0));
                const float4 b = __ldg(reinterpret_cast<const float4 *>(X + grow * ldx + c0) + 1);
                w.x = tc05::pack_half2(a.x, a.y); w.y = tc05::pack_half2(a.z, a.w);
                w.z = tc05::pack_half2(b.x, b.y); w.w = tc05::pack_half2(b.z, b.w);
            }
        } else {
            float v[8];
            #pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = (c0 + k < in_dim) ? ld_f32<TIN>(X + grow * ldx + c0 + k) : 1.0f;
            w.x = tc05::pack_half2(v[0], v[1]); w.y = tc05::pack_half2(v[2], v[3]);
            w.z = tc05::pack_half2(v[4], v[5]); w.w = tc05::pack_half2(v[6], v[7]);
        }
        *tile_word(tile, r, cg) = w;
    }
}

// 16 consecutive outputs of one row -> global memory (fp32 or fp16), vector stores when the host vouches for
// alignment (`vec`: ld and column offset keep 16-byte alignment), masked scalar stores otherwise.
template <typename TOUT>
__device__ __forceinline__ void store_out16(TOUT *__restrict__ dst, const float (&v)[16], int c0, int n_cols, bool vec) {
    if (vec && c0 + 16 <= n_cols) {
        if (sizeof(TOUT) == 2) {
            uint4 w0, w1;
            w0.x = tc05::pack_half2(v[0], v[1]);   w0.y = tc05::pack_half2(v[2], v[3]);
            w0.z = tc05::pack_half2(v[4], v[5]);   w0.w = tc05::pack_half2(v[6], v[7]);
            w1.x = tc05::pack_half2(v[8], v[9]);   w1.y = tc05::pack_half2(v[10], v[11]);
            w1.z = tc05::pack_half2(v[12], v[13]); w1.w = tc05::pack_half2(v[14], v[15]);
            reinterpret_cast<uint4 *>(dst + c0)[0] = w0;
            reinterpret_cast<uint4 *>(dst + c0)[1] = w1;
        } else {
            #pragma unroll
            for (int k = 0; k < 4; ++k)
                reinterpret_cast<float4 *>(dst + c0)[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
        }
    } else {
        #pragma unroll
        for (int k = 0; k < 16; ++k)
            if (c0 + k < n_cols) st_val<TOUT>(dst + c0 + k, v[k]);
    }
}

// A (128 x K, K-major tile) . B^T with B = weight tile (N rows x K cols, K-major)  ->  TMEM (128 x N)
__device__ __forceinline__ void issue_forward_mm(uint32_t tmem_d, uint32_t a_tile, uint32_t w_tile, int K, int N) {
    const uint32_t idesc = tc05::make_idesc(128, (uint32_t)N, 0, 0);
    const uint32_t w_lbo = (uint32_t)N * 16;
    for (int k = 0; k < K / 16; ++k) {
        const uint64_t da = tc05::make_desc(a_tile + (uint32_t)k * 2 * kTileLBO, kTileLBO, 128);
        const uint64_t db = tc05::make_desc(w_tile + (uint32_t)k * 2 * w_lbo, w_lbo, 128);
        tc05::mma_f16_ss(tmem_d, da, db, idesc, k > 0);
    }
}
// data gradient: G (128 x N, K-major tile) . W (N rows x K cols, read MN-major: "N" index = input column) -> (128 x K)
__device__ __forceinline__ void issue_dgrad_mm(uint32_t tmem_d, uint32_t g_tile, uint32_t w_tile, int K, int N) {
    const uint32_t idesc = tc05::make_idesc(128, (uint32_t)K, 0, 1);
    const uint32_t w_sbo = (uint32_t)N * 16;     // stride between 8-column groups of the weight tile
    for (int k = 0; k < N / 16; ++k) {
        const uint64_t da = tc05::make_desc(g_tile + (uint32_t)k * 2 * kTileLBO, kTileLBO, 128);
        const uint64_t db = tc05::make_desc(w_tile + (uint32_t)k * 256, 128, w_sbo);
        tc05::mma_f16_ss(tmem_d, da, db, idesc, k > 0);
    }
}
// weight gradient: P^T . Q over the 128 rows of the tile, P (128 x M) and Q (128 x N) both read MN-major -> (M x N)
__device__ __forceinline__ void issue_wgrad_mm(uint32_t tmem_d, uint32_t p_tile, uint32_t q_tile, int M, int N, bool accumulate) {
    const uint32_t idesc = tc05::make_idesc((uint32_t)M, (uint32_t)N, 1, 1);
    for (int k = 0; k < kRows / 16; ++k) {
        const uint64_t da = tc05::make_desc(p_tile + (uint32_t)k * 256, 128, kTileLBO);
        const uint64_t db = tc05::make_desc(q_tile + (uint32_t)k * 256, 128, kTileLBO);
        tc05::mma_f16_ss(tmem_d, da, db, idesc, (accumulate || k > 0) ? 1u : 0u);
    }
}

struct Shared {
    uint64_t mbar_w;
    uint64_t mbar_mma;
    uint64_t mbar_h;      // saved hidden activations of the current tile have landed (backward, bulk copies)
    uint32_t tmem_base;
    uint32_t pad;
};

// Optional extras of the forward: hidden activations saved for the backward (operand-tile layout, per 128-row tile)
// and exp() of output channel 0 (the density head's TruncExp).
struct FwdExtra {
    uint8_t *act_save;
    float *exp0_out;
    // spherical-harmonics encoding of a per-row direction written into columns [sh_col, sh_col + 16) of the (fp16) output
    // row: the geometry head then leaves the complete input row of the radiance head behind (no separate SH launch)
    const float *sh_dirs;
    const int32_t *sh_index;
    int sh_col;
    int sh_remap;
    float *zero_scalar;      // optional: one float cleared by this launch (the step's loss accumulator)
};
// Optional extras of the backward: saved hidden activations / outputs instead of re-computation, and an extra
// gradient on output channel 0 that arrives through exp(): g0[row] * exp(clamp(x0[row], +-clip)).
struct BwdExtra {
    const uint8_t *act_saved;
    const void *y_saved;
    int64_t ldys;
    int y_is_half;
    const float *g0;
    const void *x0;
    int64_t ldx0;
    int x0_is_half;
    float clip;
};

// ---------------------------------------------------------------------------------------------- forward
template <typename TIN, typename TOUT, uint32_t TMEM_COLS>
__global__ void __launch_bounds__(kRows, TMEM_COLS == 64 ? 5 : 4)
mlp_fwd_kernel(int64_t n, const int32_t *__restrict__ n_dev, const TIN *__restrict__ X, int64_t ldx,
               const uint8_t *__restrict__ blob, const __grid_constant__ MlpCfg cfg, TOUT *__restrict__ Y, int64_t ldy,
               int io_vec, FwdExtra ex) {
    extern __shared__ __align__(128) uint8_t smem[];
    Shared *sh = reinterpret_cast<Shared *>(smem);
    uint8_t *w_s = smem + 128;
    uint8_t *a_s = w_s + ((cfg.w_bytes + 127u) & ~127u);
    const int tid = threadIdx.x, warp = tid >> 5;

    if (tid == 32) {
        tc05::mbar_init(&sh->mbar_w, 1);
        tc05::mbar_init(&sh->mbar_mma, 1);
        tc05::mbar_init(&sh->mbar_h, 1);
        tc05::fence_mbar_init();
    }
    if (warp == 0) tc05::tmem_alloc<TMEM_COLS>(&sh->tmem_base);
    tc05::fence_before_sync();
    __syncthreads();
    tc05::fence_after_sync();
    arc_pdl_sync();      // barriers and tensor memory are set up; from here on inputs of earlier kernels are read
    if (n_dev) { const int64_t nd = *n_dev; n = nd < n ? (nd > 0 ? nd : 0) : n; }
    if (ex.zero_scalar && blockIdx.x == 0 && tid == 0) *ex.zero_scalar = 0.0f;
    const uint32_t tmem = sh->tmem_base;
    if (tid == 0) {
        tc05::mbar_expect_tx(&sh->mbar_w, cfg.w_bytes);
        tc05::bulk_g2s(w_s, blob, cfg.w_bytes, &sh->mbar_w);
    }
    const uint32_t a_addr = tc05::smem_u32(a_s), w_addr = tc05::smem_u32(w_s);
    const uint32_t my_tmem = tmem + ((uint32_t)(warp * 32) << 16);
    uint32_t phase = 0;
    bool weights_ready = false;

    const int64_t n_tiles = (n + kRows - 1) / kRows;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t grow = tile * kRows + tid;
        load_input_tile<TIN>(a_s, X, ldx, tile * kRows, n, cfg.in_dim, cfg.in_pad, (io_vec & 1) != 0, tid);
        if (!weights_ready) { tc05::mbar_wait(&sh->mbar_w, 0); weights_ready = true; }
        for (int l = 0; l < cfg.n_mm; ++l) {
            tc05::fence_proxy_async();
            tc05::fence_before_sync();
            __syncthreads();
            if (tid == 0) {
                tc05::fence_after_sync();
                issue_forward_mm(tmem, a_addr, w_addr + cfg.w_off[l], cfg.K[l], cfg.N[l]);
                tc05::mma_commit(&sh->mbar_mma);
            }
            if (l == 0 && ex.sh_dirs && sizeof(TOUT) == 2 && grow < n) {
                // while the first layer's MMAs run: degree-4 SH of this row's direction -> 16 fp16 values
                const size_t src = ex.sh_index ? (size_t)__ldg(ex.sh_index + grow) : (size_t)grow;
                float x = __ldg(ex.sh_dirs + 3 * src), y = __ldg(ex.sh_dirs + 3 * src + 1), z = __ldg(ex.sh_dirs + 3 * src + 2);
                sh_input(x, y, z, ex.sh_remap);
                float o[25];
                sh_basis(x, y, z, 4, o);
                uint4 w[2];
                uint32_t *pw = reinterpret_cast<uint32_t *>(w);
                #pragma unroll
                for (int k = 0; k < 8; ++k) pw[k] = tc05::pack_half2(o[2 * k], o[2 * k + 1]);
                uint4 *dst = reinterpret_cast<uint4 *>(reinterpret_cast<__half *>(Y) + grow * ldy + ex.sh_col);
                dst[0] = w[0];
                dst[1] = w[1];
            }
            tc05::mbar_wait(&sh->mbar_mma, phase);
            phase ^= 1u;
            tc05::fence_after_sync();
            const bool last = (l == cfg.n_mm - 1);
            const int act = last ? cfg.out_act : cfg.hidden_act;
            for_each_acc_block(my_tmem, cfg.N[l] / 16, [&](int cb, const uint32_t (&rr)[16]) {
                float v[16];
                act16_fwd(act, rr, v);
                if (!last) {
                    uint4 w0, w1;
                    w0.x = tc05::pack_half2(v[0], v[1]);   w0.y = tc05::pack_half2(v[2], v[3]);
                    w0.z = tc05::pack_half2(v[4], v[5]);   w0.w = tc05::pack_half2(v[6], v[7]);
                    w1.x = tc05::pack_half2(v[8], v[9]);   w1.y = tc05::pack_half2(v[10], v[11]);
                    w1.z = tc05::pack_half2(v[12], v[13]); w1.w = tc05::pack_half2(v[14], v[15]);
                    *tile_word(a_s, tid, 2 * cb) = w0;
                    *tile_word(a_s, tid, 2 * cb + 1) = w1;
                    if (ex.act_save) {   // same tile layout in global memory: consecutive threads, consecutive 16-byte words
                        uint8_t *gt = ex.act_save + ((size_t)tile * (cfg.n_mm - 1) + l) * ((size_t)kRows * cfg.width * 2);
                        *tile_word(gt, tid, 2 * cb) = w0;
                        *tile_word(gt, tid, 2 * cb + 1) = w1;
                    }
                } else if (grow < n) {
                    store_out16<TOUT>(Y + grow * ldy, v, cb * 16, cfg.out_dim, (io_vec & 2) != 0);
                    if (ex.exp0_out && cb == 0)
                        ex.exp0_out[grow] = expf(sizeof(TOUT) == 2 ? __half2float(__float2half_rn(v[0])) : v[0]);
                }
            });
        }
    }
    tc05::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc05::tmem_dealloc<TMEM_COLS>(tmem);
}

// ---------------------------------------------------------------------------------------------- backward
// TMEM columns: [0, act_cols) data tile of the layer in flight; then one weight-gradient accumulator per layer.
struct BwdPlan {
    int wg_m[kMaxMM], wg_n[kMaxMM];   // accumulator shape; wg_t = 1: it holds dW^T (M = input columns)
    int wg_t[kMaxMM];
    uint32_t wg_col[kMaxMM];
    uint32_t act_cols;
};

template <typename TIN, typename TGX, uint32_t TMEM_COLS>
__global__ void __launch_bounds__(kRows)
mlp_bwd_kernel(int64_t n, const int32_t *__restrict__ n_dev, const TIN *__restrict__ X, int64_t ldx,
               const uint8_t *__restrict__ blob, const __grid_constant__ MlpCfg cfg, const __grid_constant__ BwdPlan plan,
               const float *__restrict__ dY, int64_t ldy, float loss_scale, TGX *__restrict__ dX, int64_t ldgx,
               float *__restrict__ dW, int io_vec, BwdExtra ex) {
    extern __shared__ __align__(128) uint8_t smem[];
    Shared *sh = reinterpret_cast<Shared *>(smem);
    uint8_t *w_s = smem + 128;
    uint8_t *x_s = w_s + ((cfg.w_bytes + 127u) & ~127u);                  // 128 x in_pad
    uint8_t *h_s = x_s + (size_t)kRows * cfg.in_pad * 2;                  // (n_mm-1) tiles of 128 x width
    uint8_t *g_s = h_s + (size_t)(cfg.n_mm - 1) * kRows * cfg.width * 2;  // 128 x max(width, out_pad)
    const size_t h_bytes = (size_t)kRows * cfg.width * 2;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 32) {
        tc05::mbar_init(&sh->mbar_w, 1);
        tc05::mbar_init(&sh->mbar_mma, 1);
        tc05::mbar_init(&sh->mbar_h, 1);
        tc05::fence_mbar_init();
    }
    if (warp == 0) tc05::tmem_alloc<TMEM_COLS>(&sh->tmem_base);
    tc05::fence_before_sync();
    __syncthreads();
    tc05::fence_after_sync();
    arc_pdl_sync();      // barriers and tensor memory are set up; from here on inputs of earlier kernels are read
    if (n_dev) { const int64_t nd = *n_dev; n = nd < n ? (nd > 0 ? nd : 0) : n; }
    const uint32_t tmem = sh->tmem_base;
    if (tid == 0) {
        tc05::mbar_expect_tx(&sh->mbar_w, cfg.w_bytes);
        tc05::bulk_g2s(w_s, blob, cfg.w_bytes, &sh->mbar_w);
    }
    const uint32_t w_addr = tc05::smem_u32(w_s), x_addr = tc05::smem_u32(x_s), h_addr = tc05::smem_u32(h_s),
                   g_addr = tc05::smem_u32(g_s);
    const uint32_t my_tmem = tmem + ((uint32_t)(warp * 32) << 16);
    uint32_t phase = 0, hphase = 0;
    bool weights_ready = false, first_tile = true;
    const float inv_scale = 1.0f / loss_scale;

    auto sync_issue_wait = [&](auto issue) {
        tc05::fence_proxy_async();
        tc05::fence_before_sync();
        __syncthreads();
        if (tid == 0) {
            tc05::fence_after_sync();
            issue();
            tc05::mma_commit(&sh->mbar_mma);
        }
        tc05::mbar_wait(&sh->mbar_mma, phase);
        phase ^= 1u;
        tc05::fence_after_sync();
    };
    auto store16 = [&](uint8_t *tile, int cb, const float (&v)[16]) {
        uint4 w0, w1;
        w0.x = tc05::pack_half2(v[0], v[1]);   w0.y = tc05::pack_half2(v[2], v[3]);
        w0.z = tc05::pack_half2(v[4], v[5]);   w0.w = tc05::pack_half2(v[6], v[7]);
        w1.x = tc05::pack_half2(v[8], v[9]);   w1.y = tc05::pack_half2(v[10], v[11]);
        w1.z = tc05::pack_half2(v[12], v[13]); w1.w = tc05::pack_half2(v[14], v[15]);
        *tile_word(tile, tid, 2 * cb) = w0;
        *tile_word(tile, tid, 2 * cb + 1) = w1;
    };

    const int64_t n_tiles = (n + kRows - 1) / kRows;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t grow = tile * kRows + tid;
        load_input_tile<TIN>(x_s, X, ldx, tile * kRows, n, cfg.in_dim, cfg.in_pad, (io_vec & 1) != 0, tid);
        if (!weights_ready) { tc05::mbar_wait(&sh->mbar_w, 0); weights_ready = true; }

        // gradient at the network output for 16 columns of this thread's row (+ the exp-channel extra on column 0)
        auto load_dy16 = [&](int cb, float (&g)[16]) {
            if ((io_vec & 4) && grow < n && cb * 16 + 16 <= cfg.out_dim) {
                #pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float4 q = __ldg(reinterpret_cast<const float4 *>(dY + grow * ldy + cb * 16) + k);
                    g[4 * k] = q.x; g[4 * k + 1] = q.y; g[4 * k + 2] = q.z; g[4 * k + 3] = q.w;
                }
            } else {
                #pragma unroll
                for (int k = 0; k < 16; ++k) {
                    const int c = cb * 16 + k;
                    g[k] = (grow < n && c < cfg.out_dim) ? __ldg(dY + grow * ldy + c) : 0.0f;
                }
            }
            if (ex.g0 && cb == 0 && grow < n) {
                const float x0 = ex.x0_is_half ? __half2float(__ldg(reinterpret_cast<const __half *>(ex.x0) + grow * ex.ldx0))
                                               : __ldg(reinterpret_cast<const float *>(ex.x0) + grow * ex.ldx0);
                g[0] += __ldg(ex.g0 + grow) * expf(fminf(fmaxf(x0, -ex.clip), ex.clip));
            }
        };

        if (ex.act_saved) {
            // ---- saved activations: bulk copies (TMA engine) straight into the operand tiles, no re-computation
            __syncthreads();                               // every reader of the previous tile's h_s / g_s is done
            const uint32_t hb = (uint32_t)h_bytes * (uint32_t)(cfg.n_mm - 1);
            if (tid == 0 && hb) {
                tc05::mbar_expect_tx(&sh->mbar_h, hb);
                for (int l = 0; l + 1 < cfg.n_mm; ++l)
                    tc05::bulk_g2s(h_s + (size_t)l * h_bytes,
                                   ex.act_saved + ((size_t)tile * (cfg.n_mm - 1) + l) * h_bytes, (uint32_t)h_bytes, &sh->mbar_h);
            }
            for (int cb = 0; cb < cfg.out_pad / 16; ++cb) {
                float g[16], v[16];
                load_dy16(cb, g);
                #pragma unroll
                for (int k = 0; k < 16; ++k) v[k] = g[k] * loss_scale;
                if (cfg.out_act != ACT_NONE) {
                    float y[16];
                    #pragma unroll
                    for (int k = 0; k < 16; ++k) {
                        const int c = cb * 16 + k;
                        y[k] = 0.f;
                        if (grow < n && c < cfg.out_dim)
                            y[k] = ex.y_is_half ? __half2float(__ldg(reinterpret_cast<const __half *>(ex.y_saved) + grow * ex.ldys + c))
                                                : __ldg(reinterpret_cast<const float *>(ex.y_saved) + grow * ex.ldys + c);
                    }
                    act16_bwd_mul(cfg.out_act, v, y);
                }
                store16(g_s, cb, v);
            }
            if (hb) tc05::mbar_wait(&sh->mbar_h, hphase);
            hphase ^= 1u;
        } else {
        // ---- forward re-computation; hidden activations stay in shared memory
        for (int l = 0; l < cfg.n_mm; ++l) {
            const uint32_t a_tile = l == 0 ? x_addr : h_addr + (uint32_t)((l - 1) * h_bytes);
            sync_issue_wait([&] { issue_forward_mm(tmem, a_tile, w_addr + cfg.w_off[l], cfg.K[l], cfg.N[l]); });
            const bool last = (l == cfg.n_mm - 1);
            for_each_acc_block(my_tmem, cfg.N[l] / 16, [&](int cb, const uint32_t (&rr)[16]) {
                float v[16];
                if (!last) {
                    act16_fwd(cfg.hidden_act, rr, v);
                    store16(h_s + (size_t)l * h_bytes, cb, v);
                } else {
                    // gradient w.r.t. the last pre-activation, scaled; padded outputs and rows carry none
                    float g[16], y[16];
                    load_dy16(cb, g);
                    act16_fwd(cfg.out_act, rr, y);
                    #pragma unroll
                    for (int k = 0; k < 16; ++k) v[k] = g[k] * loss_scale;
                    act16_bwd_mul(cfg.out_act, v, y);
                    store16(g_s, cb, v);
                }
            });
        }
        }

        // ---- backward chain
        for (int l = cfg.n_mm - 1; l >= 0; --l) {
            const uint32_t in_tile = l == 0 ? x_addr : h_addr + (uint32_t)((l - 1) * h_bytes);
            const bool need_dgrad = (l > 0) || (dX != nullptr);
            sync_issue_wait([&] {
                if (plan.wg_t[l]) issue_wgrad_mm(tmem + plan.wg_col[l], in_tile, g_addr, plan.wg_m[l], plan.wg_n[l], !first_tile);
                else issue_wgrad_mm(tmem + plan.wg_col[l], g_addr, in_tile, plan.wg_m[l], plan.wg_n[l], !first_tile);
                if (need_dgrad) issue_dgrad_mm(tmem, g_addr, w_addr + cfg.w_off[l], cfg.K[l], cfg.N[l]);
            });
            if (!need_dgrad) continue;
            for_each_acc_block(my_tmem, cfg.K[l] / 16, [&](int cb, const uint32_t (&rr)[16]) {
                float v[16];
                if (l > 0) {
                    // through the hidden activation of layer l-1 (derivative from its stored output)
                    uint8_t *hp = h_s + (size_t)(l - 1) * h_bytes;
                    const uint4 a = *tile_word(hp, tid, 2 * cb);
                    const uint4 b = *tile_word(hp, tid, 2 * cb + 1);
                    const uint32_t hw[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
                    float y[16];
                    #pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const float2 h2 = __half22float2(*reinterpret_cast<const __half2 *>(&hw[k]));
                        y[2 * k] = h2.x; y[2 * k + 1] = h2.y;
                        v[2 * k] = __uint_as_float(rr[2 * k]); v[2 * k + 1] = __uint_as_float(rr[2 * k + 1]);
                    }
                    act16_bwd_mul(cfg.hidden_act, v, y);
                    store16(g_s, cb, v);
                } else if (grow < n) {
                    #pragma unroll
                    for (int k = 0; k < 16; ++k) v[k] = __uint_as_float(rr[k]) * inv_scale;
                    store_out16<TGX>(dX + grow * ldgx, v, cb * 16, cfg.in_dim, (io_vec & 2) != 0);
                }
            });
        }
        first_tile = false;
    }

    // ---- flush the weight-gradient accumulators of this CTA
    tc05::fence_before_sync();
    __syncthreads();
    tc05::fence_after_sync();
    if (!first_tile) {
        for (int l = 0; l < cfg.n_mm; ++l) {
            const int M = plan.wg_m[l], Nn = plan.wg_n[l];
            // M = 128: row = TMEM lane; M = 64: row 16*w + lane lives in lane 32*w + lane (lanes 0..15 of each warp)
            const int row = (M == 128) ? tid : (warp * 16 + lane);
            const bool owner = (M == 128) || (lane < 16);
            float *dst = dW + cfg.p_off[l];
            for (int cb = 0; cb < Nn / 16; ++cb) {
                uint32_t r[16];
                tc05::tmem_ld16(my_tmem + plan.wg_col[l] + (uint32_t)cb * 16, r);
                tc05::tmem_ld_wait();
                if (!owner) continue;
                if (!plan.wg_t[l]) {
                    // acc[out = row][in = col]: 16 consecutive parameters -> four 16-byte vector atomics
                    float4 *d4 = reinterpret_cast<float4 *>(dst + (size_t)row * cfg.K[l] + cb * 16);
                    #pragma unroll
                    for (int k = 0; k < 4; ++k)
                        atomicAdd(d4 + k, make_float4(__uint_as_float(r[4 * k]) * inv_scale, __uint_as_float(r[4 * k + 1]) * inv_scale,
                                                      __uint_as_float(r[4 * k + 2]) * inv_scale, __uint_as_float(r[4 * k + 3]) * inv_scale));
                } else {
                    // acc[in = row][out = col]; parameters are (out x in) row-major
                    #pragma unroll
                    for (int k = 0; k < 16; ++k) {
                        const float g = __uint_as_float(r[k]) * inv_scale;
                        if (g != 0.0f) atomicAdd(dst + (size_t)(cb * 16 + k) * cfg.K[l] + row, g);
                    }
                }
            }
        }
    }
    tc05::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc05::tmem_dealloc<TMEM_COLS>(tmem);
}

// fp32 flat parameters (out x in row-major per layer) -> fp16 core-matrix tiles (see tc05.cuh)
__global__ void pack_weights_kernel(MlpCfg cfg, const float *__restrict__ params, uint8_t *__restrict__ blob) {
    arc_pdl_sync();
    const int l = blockIdx.y;
    if (l >= cfg.n_mm) return;
    const int K = cfg.K[l], N = cfg.N[l];
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < K * N; e += gridDim.x * blockDim.x) {
        const int r = e / K, c = e - r * K;
        __half *tile = reinterpret_cast<__half *>(blob + cfg.w_off[l]);
        tile[(size_t)(c >> 3) * (N * 8) + (size_t)r * 8 + (c & 7)] = __float2half_rn(params[cfg.p_off[l] + e]);
    }
}

// ---- optimiser step of the (small) networks: Adam on the flat fp32 parameters and the fp16 operand tiles of the
// updated weights in the same pass, for up to kMaxJobs networks per launch (blockIdx.y = network).  The gradient is
// the sum over `world` vectors in index order (1 = this GPU's own; several = every rank's copy through peer memory).
constexpr int kMaxJobs = 4;
constexpr int kMaxGradSrc = 16;

struct AdamPackJob {
    MlpCfg cfg;
    int n_params;
    float *p, *m, *v;
    uint8_t *blob;
    float *grad[kMaxGradSrc];
};
struct AdamPackJobs { AdamPackJob job[kMaxJobs]; };

__global__ void mlp_adam_pack_kernel(const __grid_constant__ AdamPackJobs jobs, int world, AdamHyper h, int zero_grad,
                                     const int32_t *__restrict__ step_dev) {
    arc_pdl_sync();
    const AdamPackJob &J = jobs.job[blockIdx.y];
    if (step_dev) bias_from_step(h, max(*step_dev, 1));
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < J.n_params; i += gridDim.x * blockDim.x) {
        float g = __ldcg(J.grad[0] + i);
        for (int w = 1; w < world; ++w) g += __ldcg(J.grad[w] + i);
        float pp = J.p[i], mm = J.m[i], vv = J.v[i];
        adam_one(pp, mm, vv, g, h);
        J.p[i] = pp; J.m[i] = mm; J.v[i] = vv;
        if (zero_grad) J.grad[0][i] = 0.f;
        int l = 0;
        #pragma unroll
        for (int q = 1; q < kMaxMM; ++q)
            if (q < J.cfg.n_mm && (uint32_t)i >= J.cfg.p_off[q]) l = q;
        const int K = J.cfg.K[l], N = J.cfg.N[l];
        const int e = i - (int)J.cfg.p_off[l];
        const int r = e / K, c = e - r * K;
        __half *tile = reinterpret_cast<__half *>(J.blob + J.cfg.w_off[l]);
        tile[(size_t)(c >> 3) * (N * 8) + (size_t)r * 8 + (c & 7)] = __float2half_rn(pp);
    }
}

int make_cfg(MlpCfg &c, int in_dim, int width, int n_hidden, int out_dim, int hidden_act, int out_act) {
    if (n_hidden < 1 || n_hidden + 1 > kMaxMM) return -1;
    if (width % 16 != 0 || width < 16 || width > 128) return -1;
    if (in_dim < 1 || out_dim < 1) return -1;
    c.n_mm = n_hidden + 1;
    c.in_dim = in_dim;
    c.in_pad = (in_dim + 15) / 16 * 16;
    c.width = width;
    c.out_dim = out_dim;
    c.out_pad = (out_dim + 15) / 16 * 16;
    if (c.in_pad > 128 || c.out_pad > 128) return -1;
    c.hidden_act = hidden_act;
    c.out_act = out_act;
    uint32_t off = 0, poff = 0;
    for (int l = 0; l < c.n_mm; ++l) {
        c.K[l] = l == 0 ? c.in_pad : width;
        c.N[l] = l == c.n_mm - 1 ? c.out_pad : width;
        c.w_off[l] = off;
        c.p_off[l] = poff;
        off += (uint32_t)c.K[l] * c.N[l] * 2;
        poff += (uint32_t)c.K[l] * c.N[l];
    }
    c.w_bytes = off;
    return (int)poff;
}

int make_plan(const MlpCfg &c, BwdPlan &p) {
    uint32_t act_cols = 16;
    for (int l = 0; l < c.n_mm; ++l) { act_cols = max(act_cols, (uint32_t)c.N[l]); act_cols = max(act_cols, (uint32_t)c.K[l]); }
    p.act_cols = act_cols;
    uint32_t col = act_cols;
    for (int l = 0; l < c.n_mm; ++l) {
        if (c.N[l] == 64 || c.N[l] == 128) { p.wg_t[l] = 0; p.wg_m[l] = c.N[l]; p.wg_n[l] = c.K[l]; }
        else if (c.K[l] == 64 || c.K[l] == 128) { p.wg_t[l] = 1; p.wg_m[l] = c.K[l]; p.wg_n[l] = c.N[l]; }
        else return -1;
        p.wg_col[l] = col;
        col += (uint32_t)p.wg_n[l];
    }
    return (int)col;
}

static size_t fwd_smem(const MlpCfg &c) {
    const int maxk = c.in_pad > c.width ? c.in_pad : c.width;
    return 128 + ((c.w_bytes + 127u) & ~127u) + (size_t)kRows * maxk * 2;
}
static size_t bwd_smem(const MlpCfg &c) {
    const int gw = c.width > c.out_pad ? c.width : c.out_pad;
    return 128 + ((c.w_bytes + 127u) & ~127u) + (size_t)kRows * c.in_pad * 2 + (size_t)(c.n_mm - 1) * kRows * c.width * 2 +
           (size_t)kRows * gw * 2;
}

template <typename K>
static int set_smem(K kernel, size_t bytes) {
    if (bytes > 227 * 1024) { arcb200_set_error("mlp: %zu bytes of shared memory needed, 227 KB available", bytes); return ARCB200_ERR_LIMIT; }
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) { arcb200_set_error("cudaFuncSetAttribute -> %s", cudaGetErrorString(e)); return ARCB200_ERR_CUDA; }
    return ARCB200_OK;
}

// Persistent grid: resident CTAs per SM (registers and shared memory from the occupancy calculator, TMEM columns on
// top) times the SM count; a grid larger than what is resident would run its last CTAs as a second, mostly idle wave.
template <typename K>
static unsigned mlp_grid(K kernel, int64_t n, size_t smem_bytes, uint32_t tmem_cols) {
    const int64_t tiles = (n + kRows - 1) / kRows;
    static thread_local size_t cached_smem = (size_t)-1;
    static thread_local int cached_per_sm = 0;
    if (cached_smem != smem_bytes) {
        int per_sm = (int)((227 * 1024) / (smem_bytes + 1024));
        cudaFuncAttributes fa;
        if (cudaFuncGetAttributes(&fa, kernel) == cudaSuccess && fa.numRegs > 0) {
            const int regs_per_cta = ((fa.numRegs + 7) / 8 * 8) * kRows;      // allocation granularity: 8 registers per thread
            const int by_regs = 65536 / regs_per_cta;
            if (per_sm > by_regs) per_sm = by_regs;
        } else {
            cudaGetLastError();
        }
        if (getenv("ARCB200_MLP_DEBUG")) fprintf(stderr, "[mlp_grid] smem %zu tmem %u regs %d -> %d CTAs/SM before the TMEM limit\n", smem_bytes, tmem_cols, fa.numRegs, per_sm);
        const int by_tmem = (int)(512 / tmem_cols);
        if (per_sm > by_tmem) per_sm = by_tmem;
        if (per_sm > 8) per_sm = 8;
        if (per_sm < 1) per_sm = 1;
        cached_per_sm = per_sm;
        cached_smem = smem_bytes;
    }
    int per_sm = cached_per_sm;
    if (const char *e = getenv("ARCB200_MLP_CTAS_PER_SM")) { const int v = atoi(e); if (v > 0 && v < per_sm) per_sm = v; }   // experiments
    const int64_t cap = (int64_t)arc_sm_count() * per_sm;
    return (unsigned)(tiles < cap ? tiles : cap);
}

#include "mlp_bwd_v2.cuh"
#include "field_fused.cuh"

}  // namespace

extern "C" {

// Number of parameters (elements of the flat fp32 vector) of the network, tcnn layout; negative on a bad shape.
int arcb200_mlp_n_params(int in_dim, int width, int n_hidden, int out_dim) {
    MlpCfg c;
    const int n = make_cfg(c, in_dim, width, n_hidden, out_dim, 0, 0);
    if (n < 0) arcb200_set_error("mlp: unsupported shape (width multiple of 16 in [16,128], 1..3 hidden layers, dims <= 128)");
    return n;
}

// Bytes of the packed fp16 weight blob produced by arcb200_mlp_pack_weights.
int arcb200_mlp_blob_bytes(int in_dim, int width, int n_hidden, int out_dim) {
    MlpCfg c;
    if (make_cfg(c, in_dim, width, n_hidden, out_dim, 0, 0) < 0) return ARCB200_ERR_ARG;
    return (int)c.w_bytes;
}

// params: flat fp32, tcnn order (input layer first, each matrix (out x in) row-major, in = padded input width,
// last matrix padded to 16 rows).  blob: device buffer of arcb200_mlp_blob_bytes() bytes, 16-byte aligned.
int arcb200_mlp_pack_weights(const float *params, void *blob, int in_dim, int width, int n_hidden, int out_dim,
                             void *stream) {
    MlpCfg c;
    ARC_REQUIRE(make_cfg(c, in_dim, width, n_hidden, out_dim, 0, 0) >= 0, "mlp_pack_weights: unsupported shape");
    ARC_REQUIRE(params && blob && ((uintptr_t)blob & 15u) == 0, "mlp_pack_weights: null or misaligned pointer");
    arc_launch(pack_weights_kernel, dim3(8, c.n_mm), 256, 0, (cudaStream_t)stream, c, params, (uint8_t *)blob);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Adam + re-packing of n_mlps (<= 4) small networks in ONE launch: replaces, per network, arcb200_adam_step followed by
// arcb200_mlp_pack_weights (the reference: torch.optim.Adam on tcnn's `params`, common/trainer/optimizer.py, and
// tcnn's own fp32->fp16 cast of the parameters at the next forward, tinycudann/modules.py:156).
//   shapes : host int[n_mlps*4] = in_dim, width, n_hidden, out_dim of each network
//   p, m, v, blobs : host arrays of n_mlps device pointers
//   grads  : host array of n_mlps*world device pointers, grads[j*world + w] = rank w's gradient of network j (world = 1:
//            the local gradient).  Summed in index order, scaled by grad_scale.  zero_grad clears grads[j*world + 0].
int arcb200_mlp_adam_pack(int n_mlps, const int *shapes, float *const *p, float *const *m, float *const *v,
                          void *const *grads, int world, void *const *blobs, int step, float lr, float beta1,
                          float beta2, float eps, float weight_decay, float grad_scale, int zero_grad,
                          const int32_t *step_dev, void *stream) {
    ARC_REQUIRE(n_mlps >= 1 && n_mlps <= kMaxJobs, "mlp_adam_pack: 1..%d networks per call", kMaxJobs);
    ARC_REQUIRE(world >= 1 && world <= kMaxGradSrc, "mlp_adam_pack: world %d out of range (max %d)", world, kMaxGradSrc);
    ARC_REQUIRE(shapes && p && m && v && grads && blobs && (step >= 1 || step_dev), "mlp_adam_pack: null pointer or step < 1");
    AdamPackJobs jobs;
    memset(&jobs, 0, sizeof(jobs));
    int n_max = 0;
    for (int j = 0; j < n_mlps; ++j) {
        AdamPackJob &J = jobs.job[j];
        J.n_params = make_cfg(J.cfg, shapes[4 * j], shapes[4 * j + 1], shapes[4 * j + 2], shapes[4 * j + 3], 0, 0);
        ARC_REQUIRE(J.n_params > 0, "mlp_adam_pack: unsupported shape of network %d", j);
        J.p = p[j]; J.m = m[j]; J.v = v[j]; J.blob = (uint8_t *)blobs[j];
        ARC_REQUIRE(J.p && J.m && J.v && J.blob, "mlp_adam_pack: null buffer of network %d", j);
        for (int w = 0; w < world; ++w) {
            J.grad[w] = (float *)grads[j * world + w];
            ARC_REQUIRE(J.grad[w] != nullptr, "mlp_adam_pack: null gradient %d of network %d", w, j);
        }
        n_max = J.n_params > n_max ? J.n_params : n_max;
    }
    const AdamHyper h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, grad_scale, step < 1 ? 1 : step);
    arc_launch(mlp_adam_pack_kernel, dim3((n_max + 255) / 256, n_mlps), 256, 0, (cudaStream_t)stream, jobs, world, h, zero_grad, step_dev);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Forward.  X (n, in_dim) fp32 or fp16 (x_is_half), leading dimension ldx elements; Y (n, out_dim) fp32 or fp16,
// leading dimension ldy.  Activation codes: 0 none, 1 ReLU, 2 sigmoid, 3 softplus (tcnn, beta 10), 4 exp.
// Replaces tcnn Module.fwd for a FullyFusedMLP network (bindings.cpp:78-109 -> fully_fused_mlp.cu:499-700).
// Bytes of the activation-save buffer of arcb200_mlp_fwd_ex for n rows (hidden activations of every 128-row tile).
int64_t arcb200_mlp_act_save_bytes(int64_t n, int width, int n_hidden) {
    if (n < 0 || width < 16 || n_hidden < 1) return ARCB200_ERR_ARG;
    const int64_t tiles = (n + kRows - 1) / kRows;
    return tiles * n_hidden * (int64_t)kRows * width * 2;
}

int arcb200_mlp_fwd_ex(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                       int out_dim, int hidden_act, int out_act, void *Y, int y_is_half, int64_t ldy, void *act_save,
                       float *exp0_out, int64_t n, const int32_t *n_dev, void *stream);

int arcb200_mlp_fwd(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                    int out_dim, int hidden_act, int out_act, void *Y, int y_is_half, int64_t ldy, int64_t n,
                    const int32_t *n_dev, void *stream) {
    return arcb200_mlp_fwd_ex(X, x_is_half, ldx, blob, in_dim, width, n_hidden, out_dim, hidden_act, out_act, Y, y_is_half,
                              ldy, nullptr, nullptr, n, n_dev, stream);
}

// As arcb200_mlp_fwd, plus: act_save (optional, arcb200_mlp_act_save_bytes(n,...) bytes, 16-byte aligned) receives the
// hidden activations for arcb200_mlp_bwd_ex; exp0_out (optional, n floats) receives exp(output channel 0), taken from
// the value as stored in Y (the density head's TruncExp forward, arcnerf/ops/trunc_exp.py:10-24).
static int mlp_fwd_impl(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                        int out_dim, int hidden_act, int out_act, void *Y, int y_is_half, int64_t ldy, FwdExtra fex, int64_t n,
                        const int32_t *n_dev, void *stream) {
    ARC_REQUIRE(n >= 0, "mlp_fwd: negative n");
    ARC_REQUIRE(((uintptr_t)fex.act_save & 15u) == 0, "mlp_fwd: act_save must be 16-byte aligned");
    if (fex.sh_dirs) {
        ARC_REQUIRE(y_is_half && fex.sh_col >= out_dim && fex.sh_col + 16 <= ldy && (fex.sh_col * 2) % 16 == 0 &&
                    (ldy * 2) % 16 == 0 && ((uintptr_t)Y & 15u) == 0,
                    "mlp_fwd_sh: needs an fp16 output whose rows are 16-byte aligned and have room for 16 more columns at sh_col");
    }
    if (n == 0) return ARCB200_OK;
    MlpCfg c;
    ARC_REQUIRE(make_cfg(c, in_dim, width, n_hidden, out_dim, hidden_act, out_act) >= 0, "mlp_fwd: unsupported shape");
    ARC_REQUIRE(X && blob && Y, "mlp_fwd: null pointer");
    ARC_REQUIRE(((uintptr_t)blob & 15u) == 0, "mlp_fwd: weight blob must be 16-byte aligned");
    ARC_REQUIRE(ldx >= in_dim && ldy >= out_dim, "mlp_fwd: leading dimension too small");
    const size_t smem = fwd_smem(c);
    cudaStream_t st = (cudaStream_t)stream;
    // vector I/O when rows keep 16-byte alignment: bit 0 input, bit 1 output
    int io_vec = 0;
    if (in_dim % 8 == 0 && ldx % 8 == 0 && ((uintptr_t)X & 15u) == 0) io_vec |= 1;
    {
        const int esz = y_is_half ? 2 : 4;
        if ((ldy * esz) % 16 == 0 && ((uintptr_t)Y & 15u) == 0) io_vec |= 2;
    }
    // accumulator columns: the widest layer output; 64 columns let eight CTAs share an SM's tensor memory
    const bool narrow = c.width <= 64 && c.out_pad <= 64;
#define LAUNCH_FWD2(TI, TO, COLS)                                                                           \
    do {                                                                                                    \
        int rc = set_smem(mlp_fwd_kernel<TI, TO, COLS>, smem);                                              \
        if (rc != ARCB200_OK) return rc;                                                                    \
        const unsigned grid = mlp_grid(mlp_fwd_kernel<TI, TO, COLS>, n, smem, COLS);                         \
        arc_launch(mlp_fwd_kernel<TI, TO, COLS>, grid, kRows, smem, st, n, n_dev, (const TI *)X, ldx, (const uint8_t *)blob, c, (TO *)Y, ldy, io_vec, fex); \
    } while (0)
#define LAUNCH_FWD(TI, TO) do { if (narrow) LAUNCH_FWD2(TI, TO, 64); else LAUNCH_FWD2(TI, TO, 128); } while (0)
    if (x_is_half) { if (y_is_half) LAUNCH_FWD(__half, __half); else LAUNCH_FWD(__half, float); }
    else { if (y_is_half) LAUNCH_FWD(float, __half); else LAUNCH_FWD(float, float); }
#undef LAUNCH_FWD
#undef LAUNCH_FWD2
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_mlp_fwd_ex(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                       int out_dim, int hidden_act, int out_act, void *Y, int y_is_half, int64_t ldy, void *act_save,
                       float *exp0_out, int64_t n, const int32_t *n_dev, void *stream) {
    FwdExtra fex;
    memset(&fex, 0, sizeof(fex));
    fex.act_save = (uint8_t *)act_save;
    fex.exp0_out = exp0_out;
    return mlp_fwd_impl(X, x_is_half, ldx, blob, in_dim, width, n_hidden, out_dim, hidden_act, out_act, Y, y_is_half, ldy, fex, n,
                        n_dev, stream);
}

// As arcb200_mlp_fwd_ex with an fp16 output, plus the degree-4 spherical-harmonics encoding of a per-row direction
// (dirs (n_dirs,3) fp32, row i uses dirs[index[i]] or dirs[i]) written to output columns [sh_col, sh_col+16) (and,
// optionally, one float cleared: the loss accumulator that a later kernel of the step adds into): what
// FusedMLPGeoNet followed by SHEmbedder and the `[feat, sh]` concatenation of fuse_radiance_inputs (mode 'fv',
// encoder_mlp_network.py:93-118) leave behind, in one launch.  Same values as arcb200_sh_fwd(degree 4, remap).
int arcb200_mlp_fwd_sh(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                       int out_dim, int hidden_act, int out_act, void *Y, int64_t ldy, void *act_save, float *exp0_out,
                       const float *sh_dirs, const int32_t *sh_index, int sh_col, int sh_remap, float *zero_scalar,
                       int64_t n, const int32_t *n_dev, void *stream) {
    ARC_REQUIRE(sh_dirs != nullptr, "mlp_fwd_sh: null direction pointer");
    FwdExtra fex;
    memset(&fex, 0, sizeof(fex));
    fex.act_save = (uint8_t *)act_save;
    fex.exp0_out = exp0_out;
    fex.sh_dirs = sh_dirs; fex.sh_index = sh_index; fex.sh_col = sh_col; fex.sh_remap = sh_remap;
    fex.zero_scalar = zero_scalar;
    return mlp_fwd_impl(X, x_is_half, ldx, blob, in_dim, width, n_hidden, out_dim, hidden_act, out_act, Y, 1, ldy, fex, n, n_dev,
                        stream);
}

// Backward.  dY (n, out_dim) fp32 (leading dimension ldy); dX (n, in_dim) fp32 or fp16 or NULL; dW: flat fp32
// gradient in the parameter layout, ACCUMULATED (caller zeroes it).  Gradients flow through fp16 tiles scaled by
// loss_scale (128 in the reference, tinycudann/modules.py:141) and are unscaled on the way out.
// width must be 64 or 128.  Replaces tcnn Module.bwd for a FullyFusedMLP (bindings.cpp:111-170 ->
// fully_fused_mlp.cu:702-857).
int arcb200_mlp_bwd_ex(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                       int out_dim, int hidden_act, int out_act, const float *dY, int64_t ldy, float loss_scale,
                       void *dX, int dx_is_half, int64_t ldgx, float *dW, const void *act_saved, const void *y_saved,
                       int ys_is_half, int64_t ldys, const float *g0, const void *x0, int x0_is_half, int64_t ldx0,
                       float clip, int64_t n, const int32_t *n_dev, void *stream);

int arcb200_mlp_bwd(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                    int out_dim, int hidden_act, int out_act, const float *dY, int64_t ldy, float loss_scale,
                    void *dX, int dx_is_half, int64_t ldgx, float *dW, int64_t n, const int32_t *n_dev, void *stream) {
    return arcb200_mlp_bwd_ex(X, x_is_half, ldx, blob, in_dim, width, n_hidden, out_dim, hidden_act, out_act, dY, ldy,
                              loss_scale, dX, dx_is_half, ldgx, dW, nullptr, nullptr, 0, 0, nullptr, nullptr, 0, 0, 0.f, n,
                              n_dev, stream);
}

// As arcb200_mlp_bwd, plus:
//   act_saved : hidden activations written by arcb200_mlp_fwd_ex for the same X / weights / n (skips the forward
//               re-computation); then y_saved (n,out_dim; fp16 or fp32, leading dimension ldys) must hold the network
//               outputs whenever the output activation is not "none";
//   g0, x0    : extra gradient on output channel 0 arriving through exp(): dY[:,0] += g0[row]*exp(clamp(x0[row*ldx0],
//               +-clip)) - the density head (TruncExp backward, arcnerf/ops/trunc_exp.py:26-37) without a separate pass.
int arcb200_mlp_bwd_ex(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                       int out_dim, int hidden_act, int out_act, const float *dY, int64_t ldy, float loss_scale,
                       void *dX, int dx_is_half, int64_t ldgx, float *dW, const void *act_saved, const void *y_saved,
                       int ys_is_half, int64_t ldys, const float *g0, const void *x0, int x0_is_half, int64_t ldx0,
                       float clip, int64_t n, const int32_t *n_dev, void *stream) {
    ARC_REQUIRE(n >= 0, "mlp_bwd: negative n");
    ARC_REQUIRE(!act_saved || out_act == 0 || y_saved, "mlp_bwd: saved activations need the saved outputs when the output activation is not none");
    ARC_REQUIRE(((uintptr_t)act_saved & 15u) == 0, "mlp_bwd: act_saved must be 16-byte aligned");
    ARC_REQUIRE(!g0 || x0, "mlp_bwd: g0 given without x0");
    BwdExtra bex;
    bex.act_saved = (const uint8_t *)act_saved; bex.y_saved = y_saved; bex.ldys = ldys; bex.y_is_half = ys_is_half;
    bex.g0 = g0; bex.x0 = x0; bex.ldx0 = ldx0; bex.x0_is_half = x0_is_half; bex.clip = clip;
    if (n == 0) return ARCB200_OK;
    MlpCfg c;
    ARC_REQUIRE(make_cfg(c, in_dim, width, n_hidden, out_dim, hidden_act, out_act) >= 0, "mlp_bwd: unsupported shape");
    BwdPlan plan;
    const int cols = make_plan(c, plan);
    ARC_REQUIRE(cols > 0, "mlp_bwd: hidden width must be 64 or 128");
    ARC_REQUIRE(X && blob && dY && dW, "mlp_bwd: null pointer");
    ARC_REQUIRE(loss_scale > 0.0f, "mlp_bwd: loss_scale must be positive");
    ARC_REQUIRE(ldx >= in_dim && ldy >= out_dim && (!dX || ldgx >= in_dim), "mlp_bwd: leading dimension too small");
    if (cols > 512) { arcb200_set_error("mlp_bwd: %d TMEM columns needed, 512 available", cols); return ARCB200_ERR_LIMIT; }
    const size_t smem = bwd_smem(c);
    cudaStream_t st = (cudaStream_t)stream;
    // vector I/O flags: bit 0 input X, bit 1 output dX, bit 2 dY
    int io_vec = 0;
    if (in_dim % 8 == 0 && ldx % 8 == 0 && ((uintptr_t)X & 15u) == 0) io_vec |= 1;
    if (dX && (ldgx * (dx_is_half ? 2 : 4)) % 16 == 0 && ((uintptr_t)dX & 15u) == 0) io_vec |= 2;
    if (ldy % 4 == 0 && ((uintptr_t)dY & 15u) == 0) io_vec |= 4;
    if (bwd_v2_applies(c, x_is_half, dx_is_half, dX, io_vec, bex)) {
        int nwg = 4;
        if (const char *e = getenv("ARCB200_MLP_BWD_SLOTS")) nwg = atoi(e);
        int rc;
        switch (nwg) {
            case 2: rc = launch_bwd_v2<2>(c, plan, n, n_dev, X, ldx, blob, dY, ldy, loss_scale, dX, ldgx, dW, io_vec, bex, st); break;
            case 3: rc = launch_bwd_v2<3>(c, plan, n, n_dev, X, ldx, blob, dY, ldy, loss_scale, dX, ldgx, dW, io_vec, bex, st); break;
            case 5: rc = launch_bwd_v2<5>(c, plan, n, n_dev, X, ldx, blob, dY, ldy, loss_scale, dX, ldgx, dW, io_vec, bex, st); break;
            case 6: rc = launch_bwd_v2<6>(c, plan, n, n_dev, X, ldx, blob, dY, ldy, loss_scale, dX, ldgx, dW, io_vec, bex, st); break;
            default: rc = launch_bwd_v2<4>(c, plan, n, n_dev, X, ldx, blob, dY, ldy, loss_scale, dX, ldgx, dW, io_vec, bex, st); break;
        }
        if (rc < 0) return rc;
        if (rc == 0) { ARC_LAUNCH_CHECK(); return ARCB200_OK; }
    }
#define LAUNCH_BWD(TI, TG, COLS)                                                                            \
    do {                                                                                                    \
        int rc = set_smem(mlp_bwd_kernel<TI, TG, COLS>, smem);                                              \
        if (rc != ARCB200_OK) return rc;                                                                    \
        const unsigned grid = mlp_grid(mlp_bwd_kernel<TI, TG, COLS>, n, smem, COLS);                         \
        arc_launch(mlp_bwd_kernel<TI, TG, COLS>, grid, kRows, smem, st, n, n_dev, (const TI *)X, ldx, (const uint8_t *)blob, c, plan, dY, \
                                                               ldy, loss_scale, (TG *)dX, ldgx, dW, io_vec, bex); \
    } while (0)
#define LAUNCH_BWD_T(COLS)                                                                                  \
    do {                                                                                                    \
        if (x_is_half) { if (dx_is_half) LAUNCH_BWD(__half, __half, COLS); else LAUNCH_BWD(__half, float, COLS); } \
        else { if (dx_is_half) LAUNCH_BWD(float, __half, COLS); else LAUNCH_BWD(float, float, COLS); }      \
    } while (0)
    if (cols <= 128) LAUNCH_BWD_T(128);
    else if (cols <= 256) LAUNCH_BWD_T(256);
    else LAUNCH_BWD_T(512);
#undef LAUNCH_BWD_T
#undef LAUNCH_BWD
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Fused field forward of the NGP foreground model (csrc/field_fused.cuh): positions in the unit cube -> density and colour
// of every sample in ONE launch: hash-grid encode (16 levels x 2 features) -> geometry head 32-64-16 with the TruncExp
// density on channel 0 -> [feature(16) | SH degree 4 of the view direction] -> radiance head 32-64-64-3 with a sigmoid.
// Same values as arcb200_grid_fwd + arcb200_mlp_fwd_sh + arcb200_mlp_fwd_ex run one after the other.  Replaces
// Base3dModel._forward_pts_dir for the NGP configuration (arcnerf/models/base_3d_model.py:233-254: geo_net(pts) then
// radiance_net(pts, view_dirs, None, feat), i.e. tcnn_fusedmlp_module.py:81-111,177-192).
//   unit (n,3) fp32; table_half + level geometry as for arcb200_grid_fwd (must be 16 levels x 2 features);
//   geo_blob / rad_blob: packed weights of the 32-64-16 (ReLU, none) and 32-64-64-3 (ReLU, sigmoid) networks;
//   dirs (n_rays,3) fp32 with ray_id (n) int32, or ray_id NULL and dirs (n,3); sh_remap as in arcb200_sh_fwd;
//   sigma (n) fp32, rgb (n,3) fp32: outputs.  Optional outputs for the backward kernels (all or none may be NULL):
//   enc_out (n,32) fp16, rad_in_out (n,32) fp16, geo_act / rad_act (arcb200_mlp_act_save_bytes each), zero_scalar (one
//   float cleared by the launch).  n_dev: optional device row count.
int arcb200_field_fwd(const float *unit, const void *table_half, int n_levels, int n_features, const uint32_t *offsets,
                      const float *scales, const uint32_t *resolutions, const void *geo_blob, const void *rad_blob,
                      const float *dirs, const int32_t *ray_id, int sh_remap, void *enc_out, void *rad_in_out, void *geo_act,
                      void *rad_act, float *sigma, float *rgb, float *zero_scalar, int64_t n, const int32_t *n_dev,
                      void *stream) {
    ARC_REQUIRE(n >= 0 && n < (1ll << 31) - 256, "field_fwd: n out of range");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(unit && table_half && offsets && scales && resolutions && geo_blob && rad_blob && dirs && sigma && rgb,
                "field_fwd: null pointer");
    ARC_REQUIRE(n_levels == 16 && n_features == 2, "field_fwd: the fused kernel is built for 16 levels x 2 features (got %d x %d)",
                n_levels, n_features);
    ARC_REQUIRE((((uintptr_t)geo_blob | (uintptr_t)rad_blob | (uintptr_t)enc_out | (uintptr_t)rad_in_out | (uintptr_t)geo_act |
                  (uintptr_t)rad_act) & 15u) == 0, "field_fwd: blobs and saved tensors must be 16-byte aligned");
    gridmath::GridGeom geom;
    ARC_REQUIRE(gridmath::pack_geom(geom, n_levels, offsets, scales, resolutions) == 0, "field_fwd: bad level geometry");
    MlpCfg geo, rad;
    make_cfg(geo, 32, 64, 1, 16, ACT_RELU, ACT_NONE);
    make_cfg(rad, 32, 64, 2, 3, ACT_RELU, ACT_SIGMOID);
    FieldFwdArgs a;
    a.unit = unit; a.table = (const __half *)table_half; a.geo_blob = (const uint8_t *)geo_blob; a.rad_blob = (const uint8_t *)rad_blob;
    a.dirs = dirs; a.ray_id = ray_id; a.sh_remap = sh_remap; a.enc_out = (__half *)enc_out; a.rad_in_out = (__half *)rad_in_out;
    a.geo_act = (uint8_t *)geo_act; a.rad_act = (uint8_t *)rad_act; a.sigma = sigma; a.rgb = rgb; a.zero_scalar = zero_scalar;
    const size_t smem = field_fwd_smem(geo, rad);
    int rc = set_smem(field_fwd_kernel, smem);
    if (rc != ARCB200_OK) return rc;
    // one persistent CTA per SM (512 threads: 8 layer warps + 8 gather warps)
    const int64_t tiles = (n + kRows - 1) / kRows, cap = (int64_t)arc_sm_count();
    arc_launch(field_fwd_kernel, (unsigned)(tiles < cap ? tiles : cap), kFieldThreads, smem, (cudaStream_t)stream, n, n_dev, geom, geo,
               rad, a);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

#ifdef ARCB200_PROF
// debug builds: read and clear the phase clocks of the fused forward (16 uint64)
int arcb200_debug_field_prof(unsigned long long *out16) {
    ARC_CUDA(cudaMemcpyFromSymbol(out16, g_field_prof, sizeof(unsigned long long) * 16));
    unsigned long long z[16] = {0};
    ARC_CUDA(cudaMemcpyToSymbol(g_field_prof, z, sizeof(z)));
    return ARCB200_OK;
}
#endif

}  // extern "C"
