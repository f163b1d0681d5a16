// The vanilla-NeRF networks as ONE chain of dense layers on the Blackwell tensor cores.
//
// Reference: arcnerf/models/base_modules/geo_rad_model/linear_network_module.py - GeoNet (:17-197: D hidden layers of
// width W with ReLU, the input embedding concatenated again after layer `skip`, a last linear layer with 1 + W_feat
// outputs) followed by RadianceNet (:234-335, mode 'vf': [view embedding, feature] -> W_rad -> 3, sigmoid), both fed by
// FreqEmbedder (encoding/freq_encoder.py:10-88).  The reference runs every layer as a torch.nn.Linear (one cuBLAS GEMM +
// bias + activation kernels per layer, activations through HBM in between); NeRF (nerf_model.py:52-91) and the NeRF++
// background model (nerfpp_bkg_model.py:51-114, 4-D inputs) both use this pair of networks.
//
// Here a tile of 128 points goes through ALL layers inside one CTA:
//   * the activations of the tile live in shared memory as the K-major fp16 operand of the next layer (columns 0..255 the
//     hidden state, columns 256.. the frequency embedding, computed in the kernel from the raw inputs and kept for the skip
//     connection; the view embedding takes its place once the skip layer is done);
//   * weights stream from L2 through a ring of 16 KB chunks (all output rows x 32 input columns) filled by 1-D bulk copies
//     (TMA engine) of one producer thread; one MMA thread issues tcgen05.mma (M = 128, N = layer width, K = 16) into
//     tensor memory; a CTA carries TWO tiles (2 x 256 accumulator columns = all of TMEM) that use every chunk one after
//     the other, so that a chunk fetched from L2 feeds 256 rows;
//   * eight epilogue warps (four per tile) read the accumulators (tcgen05.ld), add the bias, apply the activation and
//     write the next operand tile; the epilogue of one tile runs while the other tile's MMAs of the layer are still going.
// The same kernel runs the data-gradient chain of the backward pass (weights transposed at packing time, the epilogue
// multiplies by the ReLU mask of the saved activations); weight and bias gradients are a second kernel (dense_wgrad_kernel:
// batch-contraction GEMMs dW = dZ^T . H with both operands read MN-major from row-major HBM tensors).
// Precision: fp16 operands and saved activations, fp32 accumulation, fp32 biases and outputs (the reference is fp32).
#include "common.cuh"
#include "tc05.cuh"
#include <math.h>
#include <string.h>
#include <stdlib.h>

namespace {

constexpr int kRows = 128;
constexpr int kMaxLayers = 12;
constexpr int kMaxChunks = 12;
constexpr int kChunkK = 32;                 // input columns per weight chunk
constexpr int kHCols = 256;                 // hidden region of the operand tile
constexpr int kEpiPerTile = 256;            // warps 0-7: tile 0 of the pair, warps 8-15: tile 1; two warps per TMEM row quadrant
constexpr int kProducerWarp = 16, kMmaWarp = 17;
constexpr int kThreads = 576;
constexpr int kMaxSlots = 6;
constexpr uint32_t kSlotBytes = 256 * kChunkK * 2;

enum { ACT_NONE = 0, ACT_RELU = 1, ACT_SIGMOID = 2 };
enum { F_MASK = 1, F_RANK1 = 2, F_SIGMA = 4, F_VIEW = 8, F_OUT_RGB = 16 };

struct Layer {
    uint32_t w_off;            // byte offset of the first weight chunk in the blob
    int32_t bias_off;          // float offset into `vec`, -1: no bias
    int32_t save_col;          // column offset of the saved output (fp16, row stride N) in the save buffer, -1: not saved
    int32_t mask_col;          // backward: column offset (forward save buffer) of the activations whose ReLU gates the output
    int32_t vec_off;           // F_SIGMA: sigma weights (256) + bias; F_RANK1: row vector added as d_sigma * vec
    uint16_t N;                // output columns, multiple of 16
    uint8_t n_chunks, act, flags, pad0, pad1, pad2;
    uint8_t src_cg[kMaxChunks];   // first 8-column group of each 32-column K chunk in the operand tile
};
struct Program {
    Layer layer[kMaxLayers];
    int32_t n_layers;
    int32_t aux_cols;          // 64 or 96: embedding region after the 256 hidden columns
    int32_t d_x, n_freq_x, embed_x;
    int32_t n_freq_v, embed_v;
    int32_t save_cols;         // total columns of the save buffer of this pass
    int32_t xe_col, ve_col;    // forward: columns of the saved position / view embeddings
    int32_t n_slots;
    uint32_t blob_bytes;
};

struct ChainArgs {
    const float *x;            // (n, d_x) raw inputs                       [forward]
    const float *dirs;         // (n, 3) view directions (any length)       [forward]
    const uint8_t *blob;       // packed weight chunks
    const float *vec;          // biases and row vectors
    __half *save;              // this pass's saved tensors, column-blocked: region c starts at c * n_cap, row stride = N
    const __half *fwd_save;    // backward: the forward's save buffer (ReLU masks)
    float *sigma;              // forward out (n)
    float *rgb;                // forward out (n,3); backward in: the forward's rgb
    const float *d_sigma;      // backward in (n)
    const float *d_rgb;        // backward in (n,3)
    __half *dsig16;            // backward out (n,16): column 0 = d_sigma (operand of the sigma-row weight gradient)
    int dbg_skip;              // profiling builds only: bit 0 no bias add, bit 1 no operand-tile stores, bit 2 no TMEM loads
    float loss_scale;          // backward: gradients travel through the fp16 tiles multiplied by this (undone by the weight-gradient flush)
    int64_t n, n_cap;
};

struct Ctl {
    uint64_t w_full[kMaxSlots], w_empty[kMaxSlots];
    uint64_t acc_full[2];      // all MMAs of the current layer on tile t have completed
    uint64_t h_ready[2];       // 128 arrivals: the operand tile t of the next layer is in shared memory
    uint32_t tmem_base, pad;
};

__device__ __forceinline__ void mbar_arrive(uint64_t *mbar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(tc05::smem_u32(mbar)) : "memory");
}
// Wait with back-off: the roles of this kernel wait for each other most of the time, and 100+ threads polling an mbarrier
// keep the shared-memory pipeline busy with their probes while the other warps need it for their loads and stores.
__device__ __forceinline__ void mbar_wait_sleep(uint64_t *mbar, uint32_t parity, unsigned ns) {
    while (!tc05::mbar_try_wait(mbar, parity)) __nanosleep(ns);
}
__device__ __forceinline__ uint4 *tile_word(uint8_t *tile, int r, int cg) {
    return reinterpret_cast<uint4 *>(tile + (size_t)cg * (kRows * 16) + (size_t)r * 16);
}
__device__ __forceinline__ uint4 pack8(const float *v) {
    uint4 w;
    w.x = tc05::pack_half2(v[0], v[1]); w.y = tc05::pack_half2(v[2], v[3]);
    w.z = tc05::pack_half2(v[4], v[5]); w.w = tc05::pack_half2(v[6], v[7]);
    return w;
}
__device__ __forceinline__ void unpack8(const uint4 &w, float *v) {
    const uint32_t u[4] = {w.x, w.y, w.z, w.w};
    #pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float2 f = __half22float2(*reinterpret_cast<const __half2 *>(&u[k]));
        v[2 * k] = f.x; v[2 * k + 1] = f.y;
    }
}

// Frequency embedding of a d-vector, [x | sin(x 2^0) | cos(x 2^0) | sin(x 2^1) | ...] (freq_encoder.py:78-84), written as fp16
// elements into row r of the operand tile starting at column group cg0; columns [embed, pad_cols) are zeroed.  One sincosf per
// (band, channel).
__device__ __forceinline__ void write_embedding(uint8_t *tile, int r, int cg0, const float *x, int d, int n_freq, int pad_cols, bool live) {
    auto put = [&](int c, float v) {
        *reinterpret_cast<__half *>(tile + (size_t)(cg0 + (c >> 3)) * (kRows * 16) + (size_t)r * 16 + (c & 7) * 2) = __float2half_rn(v);
    };
    const int embed = d * (2 * n_freq + 1);
    for (int ch = 0; ch < d; ++ch) {
        const float xv = live ? x[ch] : 0.f;
        put(ch, xv);
        float f = 1.0f;
        for (int k = 0; k < n_freq; ++k, f *= 2.0f) {
            float sn, cs;
            sincosf(xv * f, &sn, &cs);
            put((1 + 2 * k) * d + ch, live ? sn : 0.f);
            put((2 + 2 * k) * d + ch, live ? cs : 0.f);
        }
    }
    for (int c = embed; c < pad_cols; ++c) put(c, 0.f);
}

// Saved tensors (activations of the forward, gradients of the backward) keep the operand-tile layout in global memory: a
// tensor of width N is a sequence of 128-row tiles of N * 256 bytes, each the image of the shared-memory tile (16-byte word of
// row r and 8-column group cg at (cg * 128 + r) * 16).  A tile leaves shared memory as ONE bulk copy issued by a single
// thread, the backward reads its masks as consecutive 16-byte words across a warp, and a 64-row half of a tile is exactly the
// MN-major operand of the weight-gradient kernel.
__device__ __forceinline__ size_t tiled_word(int64_t row, int N, int cg) {          // offset in halfs
    return (size_t)(row >> 7) * (size_t)N * kRows + ((size_t)cg * kRows + (size_t)(row & 127)) * 8;
}
__device__ __forceinline__ void bulk_s2g(void *gmem_dst, const void *smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(gmem_dst), "r"(tc05::smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

#ifdef ARCB200_PROF
// debug builds: cycles spent by the MMA thread waiting for weights [0] / for operand tiles [1] / in total [2], by the epilogue
// thread 0 waiting for accumulators [3] / working [4], producer waiting for a free slot [5]; [6] pairs
__device__ unsigned long long g_chain_prof[8];
__device__ unsigned long long g_chain_layer_prof[16];   // epilogue work cycles of thread 0 per layer
__device__ unsigned long long g_chain_load_prof[2];     // load phase of thread 0
#define CHAIN_PROF(slot, expr) do { const long long t0__ = clock64(); expr; atomicAdd(&g_chain_prof[slot], (unsigned long long)(clock64() - t0__)); } while (0)
#else
#define CHAIN_PROF(slot, expr) do { expr; } while (0)
#endif

template <bool BWD>
__global__ void __launch_bounds__(kThreads, 1)
dense_chain_kernel(const __grid_constant__ Program prog, const ChainArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    Ctl *ctl = reinterpret_cast<Ctl *>(smem);
    const uint32_t tile_bytes = (uint32_t)kRows * (kHCols + prog.aux_cols) * 2;
    uint8_t *tiles = smem + 256;
    uint8_t *ring = tiles + 2 * tile_bytes;
    float *vec_s = reinterpret_cast<float *>(ring + (size_t)prog.n_slots * kSlotBytes);     // 2 x 256 floats: the layer's bias / row vector per tile
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int S = prog.n_slots;

    if (tid == 0) {
        for (int s = 0; s < S; ++s) { tc05::mbar_init(&ctl->w_full[s], 1); tc05::mbar_init(&ctl->w_empty[s], 1); }
        for (int t = 0; t < 2; ++t) { tc05::mbar_init(&ctl->acc_full[t], 1); tc05::mbar_init(&ctl->h_ready[t], kEpiPerTile); }
        tc05::fence_mbar_init();
    }
    if (warp == kMmaWarp) tc05::tmem_alloc<512>(&ctl->tmem_base);
    tc05::fence_before_sync();
    __syncthreads();
    tc05::fence_after_sync();
    arc_pdl_sync();
    const uint32_t tmem = ctl->tmem_base;
    const int64_t n = a.n;
    const int64_t n_pairs = (n + 2 * kRows - 1) / (2 * kRows);
    const int L = prog.n_layers;

    if (warp == kProducerWarp) {
        // ================================================================ weight producer: ring of chunks, program order
        if (lane == 0) {
            uint32_t i = 0;
            for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
                for (int l = 0; l < L; ++l) {
                    const Layer &ly = prog.layer[l];
                    const uint32_t bytes = (uint32_t)ly.N * kChunkK * 2;
                    for (int c = 0; c < ly.n_chunks; ++c, ++i) {
                        const int s = (int)(i % (uint32_t)S);
                        CHAIN_PROF(5, mbar_wait_sleep(&ctl->w_empty[s], ((i / (uint32_t)S) & 1u) ^ 1u, 200));
                        tc05::mbar_expect_tx(&ctl->w_full[s], bytes);
                        tc05::bulk_g2s(ring + (size_t)s * kSlotBytes, a.blob + ly.w_off + (size_t)c * bytes, bytes, &ctl->w_full[s]);
                    }
                }
            }
        }
    } else if (warp == kMmaWarp) {
        // ================================================================ MMA issue: both tiles use every chunk
        if (lane == 0) {
            uint32_t i = 0, hph[2] = {0, 0};
#ifdef ARCB200_PROF
            const long long mma_t0 = clock64();
#endif
            const uint32_t t_addr[2] = {tc05::smem_u32(tiles), tc05::smem_u32(tiles) + tile_bytes};
            const uint32_t ring_a = tc05::smem_u32(ring);
            for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
                for (int l = 0; l < L; ++l) {
                    const Layer &ly = prog.layer[l];
                    const uint32_t N = ly.N, idesc = tc05::make_idesc(128, N, 0, 0), w_lbo = N * 16;
                    for (int c = 0; c < ly.n_chunks; ++c, ++i) {
                        const int s = (int)(i % (uint32_t)S);
                        CHAIN_PROF(0, mbar_wait_sleep(&ctl->w_full[s], (i / (uint32_t)S) & 1u, 40));
                        const uint32_t w_a = ring_a + (uint32_t)s * kSlotBytes;
                        #pragma unroll
                        for (int t = 0; t < 2; ++t) {
                            if (c == 0) {                     // the tile's operand for this layer has been written
                                CHAIN_PROF(1, mbar_wait_sleep(&ctl->h_ready[t], hph[t], 40));
                                hph[t] ^= 1u;
                                if (a.save) {
                                    // what the epilogue warps have just finished goes to global memory as whole tiles: the output of the
                                    // previous layer (columns 0..N-1), the position embedding after the load phase, the view embedding
                                    // after the layer that wrote it
                                    const int64_t tile_idx = pair * 2 + t;
                                    uint8_t *tile_s = tiles + (size_t)t * tile_bytes;
                                    if (tile_idx * kRows < n) {
                                        if (l == 0) {
                                            if (!BWD) {
                                                const int xe_pad = ((prog.embed_x + 31) / 32) * 32;
                                                bulk_s2g(a.save + (size_t)prog.xe_col * a.n_cap + (size_t)tile_idx * xe_pad * kRows,
                                                         tile_s + (size_t)(kHCols / 8) * (kRows * 16), (uint32_t)xe_pad * kRows * 2);
                                            }
                                        } else {
                                            const Layer &pl = prog.layer[l - 1];
                                            if (pl.save_col >= 0)
                                                bulk_s2g(a.save + (size_t)pl.save_col * a.n_cap + (size_t)tile_idx * pl.N * kRows, tile_s,
                                                         (uint32_t)pl.N * kRows * 2);
                                            if (!BWD && (pl.flags & F_VIEW))
                                                bulk_s2g(a.save + (size_t)prog.ve_col * a.n_cap + (size_t)tile_idx * 32 * kRows,
                                                         tile_s + (size_t)(kHCols / 8) * (kRows * 16), 32u * kRows * 2);
                                        }
                                    }
                                    bulk_commit();
                                }
                            }
                            tc05::fence_after_sync();
                            #pragma unroll
                            for (int k = 0; k < kChunkK / 16; ++k) {
                                const uint64_t da = tc05::make_desc(t_addr[t] + (uint32_t)(ly.src_cg[c] + 2 * k) * (kRows * 16), kRows * 16, 128);
                                const uint64_t db = tc05::make_desc(w_a + (uint32_t)k * 2 * w_lbo, w_lbo, 128);
                                tc05::mma_f16_ss(tmem + (uint32_t)t * 256, da, db, idesc, (c > 0 || k > 0) ? 1u : 0u);
                            }
                            if (c == ly.n_chunks - 1) {
                                // the epilogue of this layer overwrites the tile: the stores issued above must have read it
                                if (a.save) bulk_wait_read_all();
                                tc05::mma_commit(&ctl->acc_full[t]);
                            }
                        }
                        tc05::mma_commit(&ctl->w_empty[s]);      // the chunk may be overwritten once these MMAs are done
                    }
                }
                if (BWD && a.save) {
                    const Layer &pl = prog.layer[L - 1];
                    for (int t = 0; t < 2; ++t) {
                        mbar_wait_sleep(&ctl->h_ready[t], hph[t], 40);
                        hph[t] ^= 1u;
                        const int64_t tile_idx = pair * 2 + t;
                        if (tile_idx * kRows < n && pl.save_col >= 0)
                            bulk_s2g(a.save + (size_t)pl.save_col * a.n_cap + (size_t)tile_idx * pl.N * kRows, tiles + (size_t)t * tile_bytes,
                                     (uint32_t)pl.N * kRows * 2);
                    }
                    bulk_commit();
                    bulk_wait_read_all();
                    mbar_arrive(&ctl->acc_full[0]);
                    mbar_arrive(&ctl->acc_full[1]);
                }
#ifdef ARCB200_PROF
                atomicAdd(&g_chain_prof[6], 1ull);
#endif
            }
#ifdef ARCB200_PROF
            atomicAdd(&g_chain_prof[2], (unsigned long long)(clock64() - mma_t0));
#endif
            if (a.save) bulk_wait_all();
        }
    } else {
        // ================================================================ epilogue warps: thread = one row of one tile
        // row r of tile t; the two warps of a row quadrant (half 0 / 1) split the columns of every layer
        const int t = warp >> 3, half = (warp >> 2) & 1;
        const int r = (warp & 3) * 32 + lane;
        uint8_t *tile = tiles + (size_t)t * tile_bytes;
        const uint32_t acc = tmem + (uint32_t)t * 256 + ((uint32_t)((warp & 3) * 32) << 16);
        const int aux_cg0 = kHCols / 8;
        uint32_t accph = 0;
        for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
            const int64_t grow = (pair * 2 + t) * kRows + r;
            const bool live = grow < n;
            float dsig = 0.f;
#ifdef ARCB200_PROF
            const long long ld_t0 = clock64();
#endif
            // ---- the tile's first operand (written by the half-0 thread of the row)
            if (half != 0) {
                if (BWD && live) dsig = __ldg(a.d_sigma + grow) * a.loss_scale;
            } else if (!BWD) {
                float x[4] = {0.f, 0.f, 0.f, 0.f};
                if (live) {
                    #pragma unroll
                    for (int k = 0; k < 4; ++k) if (k < prog.d_x) x[k] = __ldg(a.x + grow * prog.d_x + k);
                }
                write_embedding(tile, r, aux_cg0, x, prog.d_x, prog.n_freq_x, prog.aux_cols, live);
            } else {
                // gradient at the colour output through the sigmoid, padded to one 32-column chunk
                float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (live) {
                    #pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        const float y = __ldg(a.rgb + grow * 3 + k);
                        v[k] = __ldg(a.d_rgb + grow * 3 + k) * y * (1.0f - y) * a.loss_scale;
                    }
                    dsig = __ldg(a.d_sigma + grow) * a.loss_scale;
                }
                const uint4 w = pack8(v), z = make_uint4(0u, 0u, 0u, 0u);
                *tile_word(tile, r, 0) = w;
                *tile_word(tile, r, 1) = z; *tile_word(tile, r, 2) = z; *tile_word(tile, r, 3) = z;
                if (live) {
                    // region 0 of the backward save (width 16) and the d_sigma operand (width 16), tiled like every saved tensor
                    *reinterpret_cast<uint4 *>(a.save + tiled_word(grow, 16, 0)) = w;
                    *reinterpret_cast<uint4 *>(a.save + tiled_word(grow, 16, 1)) = z;
                    float s8[8] = {dsig, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                    *reinterpret_cast<uint4 *>(a.dsig16 + tiled_word(grow, 16, 0)) = pack8(s8);
                    *reinterpret_cast<uint4 *>(a.dsig16 + tiled_word(grow, 16, 1)) = z;
                }
            }
            tc05::fence_proxy_async();
            tc05::fence_before_sync();
            mbar_arrive(&ctl->h_ready[t]);
#ifdef ARCB200_PROF
            if (tid == 0) atomicAdd(&g_chain_load_prof[0], (unsigned long long)(clock64() - ld_t0));
#endif

            for (int l = 0; l < L; ++l) {
                const Layer &ly = prog.layer[l];
                // every field the epilogue needs, read once: the inline-asm TMEM instructions are memory barriers for the
                // compiler, which would otherwise fetch them again (indexed constant loads) in front of every 16 columns
                const int N = ly.N, act = ly.act, flags = ly.flags, bias_off = ly.bias_off, vec_off = ly.vec_off,
                          save_col = ly.save_col, mask_col = ly.mask_col;
                float sigma_v = 0.f;
                if ((flags & F_SIGMA) && half == 0) {
                    // density row of the last geometry layer on the CUDA cores while that layer's MMAs run: the operand tile is
                    // read-only until the accumulators are complete
                    const float *sw = a.vec + vec_off;
                    float wl[8];                                   // lane j holds weights j, j + 32, ...: one coalesced read per warp
                    #pragma unroll
                    for (int q8 = 0; q8 < 8; ++q8) wl[q8] = __ldg(sw + q8 * 32 + lane);
                    float acc_s = __ldg(sw + kHCols);
                    #pragma unroll
                    for (int q8 = 0; q8 < 8; ++q8) {
                        #pragma unroll
                        for (int g4 = 0; g4 < 4; ++g4) {
                            float h[8];
                            unpack8(*tile_word(tile, r, q8 * 4 + g4), h);
                            #pragma unroll
                            for (int k = 0; k < 8; ++k) acc_s = fmaf(h[k], __shfl_sync(0xffffffffu, wl[q8], g4 * 8 + k), acc_s);
                        }
                    }
                    sigma_v = acc_s;
                }
                // the layer's bias (forward) or row vector (backward) into shared memory while the MMAs of the layer run: one
                // coalesced read per layer instead of a global load in front of every 16 columns of the epilogue
                float *lv = vec_s + t * 256;
                const int lv_off = bias_off >= 0 ? bias_off : ((BWD && (flags & F_RANK1)) ? vec_off : -1);
                if (lv_off >= 0) {
                    asm volatile("bar.sync %0, 256;" :: "r"(1 + t) : "memory");          // everyone is done with the previous layer's vector
                    lv[r + 128 * half] = r + 128 * half < N ? __ldg(a.vec + lv_off + r + 128 * half) : 0.f;
                    asm volatile("bar.sync %0, 256;" :: "r"(1 + t) : "memory");
                }
#ifdef ARCB200_PROF
                const long long ep_t0 = clock64();
#endif
                if (lane == 0) mbar_wait_sleep(&ctl->acc_full[t], accph, 100);      // one lane polls for the warp
                __syncwarp();
                accph ^= 1u;
#ifdef ARCB200_PROF
                const long long ep_t1 = clock64();
                if (tid == 0) atomicAdd(&g_chain_prof[3], (unsigned long long)(ep_t1 - ep_t0));
#endif
                tc05::fence_after_sync();
                if ((flags & F_SIGMA) && half == 0 && live) a.sigma[grow] = sigma_v;
                const float *bias = bias_off >= 0 ? lv : nullptr;
                const __half *mask = (BWD && (flags & F_MASK)) ? a.fwd_save + (size_t)mask_col * a.n_cap : nullptr;
                const float *r1 = (BWD && (flags & F_RANK1)) ? lv : nullptr;
                // 64 columns per trip: four TMEM loads and (backward) the eight mask words of the trip are in flight together
                // this half's 16-column blocks: [cb_lo, cb_hi); two blocks (32 columns) per trip, their TMEM loads and (backward)
                // mask words in flight together
                const int nblk = N / 16, cb_lo = half * ((nblk + 1) / 2), cb_hi = half ? nblk : (nblk + 1) / 2;
                for (int cb = cb_lo; cb < cb_hi; cb += 2) {
#ifdef ARCB200_PROF
                    if (a.dbg_skip & 8) break;
#endif
                    const int nb = (cb_hi - cb) < 2 ? (cb_hi - cb) : 2;
                    uint32_t q[2][16];
                    uint4 mw[4];
#ifdef ARCB200_PROF
                    if (a.dbg_skip & 4) {
                        #pragma unroll
                        for (int k = 0; k < 16; ++k) { q[0][k] = (uint32_t)(k + cb); q[1][k] = (uint32_t)(k * 3 + cb); }
                    } else
#endif
                    {
                    tc05::tmem_ld16(acc + (uint32_t)cb * 16, q[0]);
                    if (nb > 1) tc05::tmem_ld16(acc + (uint32_t)cb * 16 + 16, q[1]);
                    }
                    if (mask) {
                        #pragma unroll
                        for (int w = 0; w < 4; ++w)
                            mw[w] = (live && w < 2 * nb) ? __ldg(reinterpret_cast<const uint4 *>(mask + tiled_word(grow, N, 2 * cb + w))) : make_uint4(0u, 0u, 0u, 0u);
                    }
#ifdef ARCB200_PROF
                    const long long tp0 = clock64();
#endif
                    tc05::tmem_ld_wait();
#ifdef ARCB200_PROF
                    const long long tp1 = clock64();
                    if (tid == 0) atomicAdd(&g_chain_load_prof[1], (unsigned long long)(tp1 - tp0));
#endif
                    #pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        if (u >= nb) break;
                        const int c0 = (cb + u) * 16;
                        float v[16];
                        #pragma unroll
                        for (int k = 0; k < 16; ++k) v[k] = __uint_as_float(q[u][k]);
                        if (bias
#ifdef ARCB200_PROF
                            && !(a.dbg_skip & 1)
#endif
                        ) {
                            float4 b[4];
                            #pragma unroll
                            for (int k4 = 0; k4 < 4; ++k4) b[k4] = reinterpret_cast<const float4 *>(bias + c0)[k4];
                            #pragma unroll
                            for (int k4 = 0; k4 < 4; ++k4) {
                                v[4 * k4] += b[k4].x; v[4 * k4 + 1] += b[k4].y; v[4 * k4 + 2] += b[k4].z; v[4 * k4 + 3] += b[k4].w;
                            }
                        }
                        if (r1) {
                            #pragma unroll
                            for (int k4 = 0; k4 < 4; ++k4) {
                                const float4 b = reinterpret_cast<const float4 *>(r1 + c0)[k4];
                                v[4 * k4] = fmaf(dsig, b.x, v[4 * k4]); v[4 * k4 + 1] = fmaf(dsig, b.y, v[4 * k4 + 1]);
                                v[4 * k4 + 2] = fmaf(dsig, b.z, v[4 * k4 + 2]); v[4 * k4 + 3] = fmaf(dsig, b.w, v[4 * k4 + 3]);
                            }
                        }
                        if (act == ACT_RELU) {
                            #pragma unroll
                            for (int k = 0; k < 16; ++k) v[k] = fmaxf(v[k], 0.f);
                        } else if (act == ACT_SIGMOID) {
                            #pragma unroll
                            for (int k = 0; k < 16; ++k) v[k] = 1.0f / (1.0f + __expf(-v[k]));
                        }
                        if (mask) {
                            float m[16];
                            unpack8(mw[2 * u], m);
                            unpack8(mw[2 * u + 1], m + 8);
                            #pragma unroll
                            for (int k = 0; k < 16; ++k) v[k] = m[k] > 0.f ? v[k] : 0.f;
                        }
                        if (flags & F_OUT_RGB) {
                            if (live && c0 == 0) { a.rgb[grow * 3] = v[0]; a.rgb[grow * 3 + 1] = v[1]; a.rgb[grow * 3 + 2] = v[2]; }
                        } else {
                            const uint4 w0 = pack8(v), w1 = pack8(v + 8);
#ifdef ARCB200_PROF
                            if (!(a.dbg_skip & 2) || w0.x == 0x12345678u)
#endif
                            {
                            *tile_word(tile, r, 2 * (cb + u)) = w0;
                            *tile_word(tile, r, 2 * (cb + u) + 1) = w1;
                            }
                        }
                    }
                }
                if (!BWD && (flags & F_VIEW) && half == 1) {
                    // the position embedding has served its last layer: the view embedding takes its columns
                    float d[3] = {0.f, 0.f, 1.f};
                    if (live) {
                        d[0] = __ldg(a.dirs + grow * 3); d[1] = __ldg(a.dirs + grow * 3 + 1); d[2] = __ldg(a.dirs + grow * 3 + 2);
                        const float inv = 1.0f / sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
                        d[0] *= inv; d[1] *= inv; d[2] *= inv;
                    }
                    write_embedding(tile, r, aux_cg0, d, 3, prog.n_freq_v, 32, live);
                }
                if (l + 1 < L || (BWD && a.save)) {
#ifdef ARCB200_PROF
                    if (!(a.dbg_skip & 16))
#endif
                    tc05::fence_proxy_async();
                    tc05::fence_before_sync();
                    mbar_arrive(&ctl->h_ready[t]);
                }
#ifdef ARCB200_PROF
                if (tid == 0) { atomicAdd(&g_chain_prof[4], (unsigned long long)(clock64() - ep_t1)); atomicAdd(&g_chain_layer_prof[l], (unsigned long long)(clock64() - ep_t1)); }
#endif
            }
            if (BWD && a.save) {
                // the MMA thread stores the last gradient tile and says when the tile may be overwritten by the next pair
                if (lane == 0) mbar_wait_sleep(&ctl->acc_full[t], accph, 100);
                __syncwarp();
                accph ^= 1u;
            }
            tc05::fence_before_sync();      // the next pair's first MMA must not overtake these accumulator reads
        }
    }
    tc05::fence_before_sync();
    __syncthreads();
    if (warp == kMmaWarp) tc05::tmem_dealloc<512>(tmem);
}

// ---------------------------------------------------------------------------------------------- weight gradients
// dW[0:N, 0:K] += dZ^T . A over the batch, db[0:N] += column sums of dZ.  dZ (n, ldz) and A (n, lda) fp16 row-major.
// CTA: a contiguous range of rows, staged 64 at a time (cp.async, three stages) as MN-major operands: a 16-byte word of
// a row-major row IS a core-matrix row, so staging is a plain copy.  Accumulators: output rows 0..127 in TMEM columns
// [0, K), rows 128..255 in [256, 256 + K).  One flush per CTA with vector atomics.
constexpr int kWgRows = 64;
constexpr int kWgThreads = 256;
constexpr int kWgStages = 3;

struct WgArgs {
    const __half *dz; int64_t ldz; int N;        // N: multiple of 16 (<= 256); n_out_valid rows of dW are written
    const __half *act; int64_t lda; int K;       // K: multiple of 16 (<= 256); k_valid columns are written
    int n_out_valid, k_valid;
    float *dW; int64_t ldw;                      // fp32 (N x K) block inside a larger matrix
    float *db;                                   // fp32 (N) or null
    float out_scale;                             // multiplies everything that is flushed (1 / loss scale)
    int64_t n;
};

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src, bool pred) {
    const uint32_t sz = pred ? 16u : 0u;         // src-size 0: the destination is zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(tc05::smem_u32(smem_dst)), "l"(gmem_src), "r"(sz) : "memory");
}

// All weight-gradient GEMMs of a backward pass run as ONE launch: job j owns CTAs [first[j], first[j+1]), which split the
// batch rows between them.  A job with few CTAs flushes few partial sums (the flush is atomics on the same dW entries), and
// the jobs together fill the machine.
#ifdef ARCB200_PROF
__device__ unsigned long long g_wg_prof[8];   // thread 0 of big jobs: [0] wait stage free, [1] issue loads, [2] wait loads, [3] column sums, [4] sync + MMA issue, [5] iterations
#define WG_PROF(slot, expr) do { const long long t0__ = clock64(); expr; if (tid == 0 && a.N == 256 && a.K == 256) atomicAdd(&g_wg_prof[slot], (unsigned long long)(clock64() - t0__)); } while (0)
#else
#define WG_PROF(slot, expr) do { expr; } while (0)
#endif
constexpr int kMaxWgJobs = 20;
struct WgJobs {
    WgArgs job[kMaxWgJobs];
    int32_t first[kMaxWgJobs + 1];
    int32_t n_jobs;
};

__global__ void __launch_bounds__(kWgThreads, 1) dense_wgrad_kernel(const __grid_constant__ WgJobs jobs) {
    int ji = 0;
    while (ji + 1 < jobs.n_jobs && (int)blockIdx.x >= jobs.first[ji + 1]) ++ji;
    const WgArgs &a = jobs.job[ji];
    const int cta = (int)blockIdx.x - jobs.first[ji], n_cta = jobs.first[ji + 1] - jobs.first[ji];
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *mbar = reinterpret_cast<uint64_t *>(smem);          // [kWgStages] stage consumed by the tensor core
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem + 64);
    float *db_s = reinterpret_cast<float *>(smem + 128);          // 256 floats
    uint8_t *stage0 = smem + 128 + 1024;
    const int Mp = (a.N + 127) / 128 * 128;                       // dZ tile columns (padded to whole 128-row halves of dW)
    const uint32_t dz_bytes = (uint32_t)kWgRows * Mp * 2, act_bytes = (uint32_t)kWgRows * a.K * 2;
    const uint32_t stage_bytes = dz_bytes + act_bytes;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < kWgStages; ++s) tc05::mbar_init(&mbar[s], 1);
        tc05::fence_mbar_init();
    }
    db_s[tid] = 0.f;
    if (warp == 0) tc05::tmem_alloc<512>(tmem_slot);
    tc05::fence_before_sync();
    __syncthreads();
    tc05::fence_after_sync();
    arc_pdl_sync();
    const uint32_t tmem = *tmem_slot;
    // rows of this CTA
    const int64_t blocks = (a.n + kWgRows - 1) / kWgRows;
    const int64_t per = (blocks + n_cta - 1) / n_cta;
    const int64_t b0 = (int64_t)cta * per, b1 = b0 + per < blocks ? b0 + per : blocks;
    const int n_iter = b1 > b0 ? (int)(b1 - b0) : 0;
    const int dz_words = kWgRows * (a.N / 8), act_words = kWgRows * (a.K / 8);
    const int n_half = Mp / 128;
    float colsum[8][8];                  // bias gradient: partial column sums of the dZ words this thread stages
    #pragma unroll
    for (int q = 0; q < 8; ++q)
        #pragma unroll
        for (int k = 0; k < 8; ++k) colsum[q][k] = 0.f;

    auto stage_load = [&](int it) {
        uint8_t *st = stage0 + (size_t)(it % kWgStages) * stage_bytes;
        const int64_t row0 = (b0 + it) * kWgRows;
        // a word = 8 columns of one row; word w -> (row = w / groups, group = w % groups): consecutive threads read consecutive
        // 16-byte pieces of a row (coalesced), and write the MN-major tile: group * (64 * 16) + row * 16
        // sources are in the tiled layout (see tiled_word): word w -> (group = w / 64, row = w % 64), so that consecutive threads
        // read consecutive 16-byte words of a tile's column group and write consecutive words of the stage
        for (int w = tid; w < dz_words; w += kWgThreads) {
            const int cg = w >> 6, rr = w & 63;
            cp_async16(st + (size_t)cg * (kWgRows * 16) + (size_t)rr * 16, a.dz + tiled_word(row0 + rr, (int)a.ldz, cg), row0 + rr < a.n);
        }
        for (int w = tid; w < act_words; w += kWgThreads) {
            const int cg = w >> 6, rr = w & 63;
            cp_async16(st + dz_bytes + (size_t)cg * (kWgRows * 16) + (size_t)rr * 16, a.act + tiled_word(row0 + rr, (int)a.lda, cg), row0 + rr < a.n);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    // zero the padding columns of the dZ tiles once (N < Mp): stale shared memory must not reach the tensor core as NaN
    if (a.N < Mp) {
        for (int s = 0; s < kWgStages; ++s)
            for (int w = tid; w < kWgRows * (Mp - a.N) / 8; w += kWgThreads)
                *reinterpret_cast<uint4 *>(stage0 + (size_t)s * stage_bytes + (size_t)(a.N / 8) * (kWgRows * 16) + (size_t)w * 16) = make_uint4(0u, 0u, 0u, 0u);
    }
    for (int it = 0; it < kWgStages - 1; ++it) {
        if (it < n_iter) stage_load(it); else asm volatile("cp.async.commit_group;" ::: "memory");
    }
    uint32_t ph[kWgStages] = {0, 0, 0};
    int commits[kWgStages] = {0, 0, 0}, waits[kWgStages] = {0, 0, 0};
    const uint32_t idesc = tc05::make_idesc(128, (uint32_t)a.K, 1, 1);
    for (int it = 0; it < n_iter; ++it) {
        const int s = it % kWgStages;
        // the stage that the next load overwrites was consumed by the MMAs of iteration it-1
        if (it + kWgStages - 1 < n_iter) {
            const int sn = (it + kWgStages - 1) % kWgStages;
            if (it >= 1) { WG_PROF(0, tc05::mbar_wait(&mbar[sn], ph[sn])); ph[sn] ^= 1u; ++waits[sn]; }
            WG_PROF(1, stage_load(it + kWgStages - 1));
        } else {
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
        WG_PROF(2, asm volatile("cp.async.wait_group %0;" :: "n"(kWgStages - 1) : "memory"));
#ifdef ARCB200_PROF
        const long long cs_t0 = clock64();
#endif
        // bias gradient: column sums of the dZ words this thread staged itself (visible to it after its own wait)
        if (a.db) {
            uint8_t *st = stage0 + (size_t)s * stage_bytes;
            // word tid + 256 q of every stage is column group (tid >> 6) + 4 q: one register accumulator set per q
            #pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int w = tid + q * kWgThreads;
                if (w < dz_words) {
                    float v[8];
                    unpack8(*reinterpret_cast<const uint4 *>(st + (size_t)(w >> 6) * (kWgRows * 16) + (size_t)(w & 63) * 16), v);
                    #pragma unroll
                    for (int k = 0; k < 8; ++k) colsum[q][k] += v[k];
                }
            }
        }
#ifdef ARCB200_PROF
        const long long cs_t1 = clock64();
        if (tid == 0 && a.N == 256 && a.K == 256) { atomicAdd(&g_wg_prof[3], (unsigned long long)(cs_t1 - cs_t0)); atomicAdd(&g_wg_prof[5], 1ull); }
#endif
        tc05::fence_proxy_async();
        tc05::fence_before_sync();
        __syncthreads();
        if (tid == 0) {
            tc05::fence_after_sync();
            const uint32_t st_a = tc05::smem_u32(stage0 + (size_t)s * stage_bytes);
            for (int h = 0; h < n_half; ++h) {
                #pragma unroll
                for (int k = 0; k < kWgRows / 16; ++k) {
                    const uint64_t da = tc05::make_desc(st_a + (uint32_t)h * 16 * (kWgRows * 16) + (uint32_t)k * 256, 128, kWgRows * 16);
                    const uint64_t db = tc05::make_desc(st_a + dz_bytes + (uint32_t)k * 256, 128, kWgRows * 16);
                    tc05::mma_f16_ss(tmem + (uint32_t)h * 256, da, db, idesc, (it > 0 || k > 0) ? 1u : 0u);
                }
            }
            tc05::mma_commit(&mbar[s]);
        }
#ifdef ARCB200_PROF
        if (tid == 0 && a.N == 256 && a.K == 256) atomicAdd(&g_wg_prof[4], (unsigned long long)(clock64() - cs_t1));
#endif
        ++commits[s];
    }
    // drain: every commit that has not been waited for yet (all threads followed the same schedule, so they agree on the counts)
    for (int s = 0; s < kWgStages; ++s)
        for (; waits[s] < commits[s]; ++waits[s]) { tc05::mbar_wait(&mbar[s], ph[s]); ph[s] ^= 1u; }
    tc05::fence_after_sync();
    if (n_iter > 0) {
        // flush: warp w reads TMEM lanes 32 (w % 4); warps 0-3 take half 0, warps 4-7 half 1
        const int h = warp >> 2, row = h * 128 + (warp & 3) * 32 + lane;
        if (h < n_half) {
            const uint32_t acc = tmem + (uint32_t)h * 256 + ((uint32_t)((warp & 3) * 32) << 16);
            for (int cb = 0; cb < a.K / 16; ++cb) {
                uint32_t q[16];
                tc05::tmem_ld16(acc + (uint32_t)cb * 16, q);
                tc05::tmem_ld_wait();
                if (row < a.n_out_valid) {
                    float *dst = a.dW + (size_t)row * a.ldw + cb * 16;
                    if ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0 && cb * 16 + 16 <= a.k_valid) {
                        // 16-byte vector reductions: a quarter of the L2 operations of the scalar form
                        #pragma unroll
                        for (int k4 = 0; k4 < 4; ++k4)
                            atomicAdd(reinterpret_cast<float4 *>(dst) + k4,
                                      make_float4(__uint_as_float(q[4 * k4]) * a.out_scale, __uint_as_float(q[4 * k4 + 1]) * a.out_scale,
                                                  __uint_as_float(q[4 * k4 + 2]) * a.out_scale, __uint_as_float(q[4 * k4 + 3]) * a.out_scale));
                    } else {
                        #pragma unroll
                        for (int k = 0; k < 16; ++k)
                            if (cb * 16 + k < a.k_valid) atomicAdd(dst + k, __uint_as_float(q[k]) * a.out_scale);
                    }
                }
            }
        }
        if (a.db) {
            #pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int w = tid + q * kWgThreads;
                if (w < dz_words) {
                    #pragma unroll
                    for (int k = 0; k < 8; ++k) atomicAdd(&db_s[(w >> 6) * 8 + k], colsum[q][k]);
                }
            }
            __syncthreads();
            if (tid < a.N && tid < a.n_out_valid) atomicAdd(a.db + tid, db_s[tid] * a.out_scale);
        }
    }
    tc05::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc05::tmem_dealloc<512>(tmem);
}

// ---------------------------------------------------------------------------------------------- packing
// One descriptor per layer of a program: chunk c, tile row r, chunk column j  <-  params[base[c] + r*stride_r + j*stride_j]
// for r < r_valid and j < j_valid[c], zero otherwise.
struct PackLayer {
    uint32_t w_off;
    int32_t N, n_chunks, r_valid;
    int64_t stride_r, stride_j;
    int64_t base[kMaxChunks];
    int32_t j_valid[kMaxChunks];
};
struct PackPlan { PackLayer layer[kMaxLayers]; int32_t n_layers; };

__global__ void chain_pack_kernel(const __grid_constant__ PackPlan plan, const float *__restrict__ params, uint8_t *__restrict__ blob) {
    arc_pdl_sync();
    const PackLayer &pl = plan.layer[blockIdx.y];
    const int per_chunk = pl.N * kChunkK;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < per_chunk * pl.n_chunks; e += gridDim.x * blockDim.x) {
        const int c = e / per_chunk, q = e - c * per_chunk;
        const int r = q / kChunkK, j = q - r * kChunkK;
        float v = 0.f;
        if (r < pl.r_valid && j < pl.j_valid[c]) v = params[pl.base[c] + r * pl.stride_r + j * pl.stride_j];
        __half *chunk = reinterpret_cast<__half *>(blob + pl.w_off + (size_t)c * per_chunk * 2);
        chunk[(size_t)(j >> 3) * (pl.N * 8) + (size_t)r * 8 + (j & 7)] = __float2half_rn(v);
    }
}

struct VecCopy { int64_t src; int32_t dst, count; };
struct VecPlan { VecCopy c[2 * kMaxLayers + 4]; int32_t n; };
__global__ void chain_vec_kernel(const __grid_constant__ VecPlan plan, const float *__restrict__ params, float *__restrict__ vec) {
    arc_pdl_sync();
    const VecCopy &c = plan.c[blockIdx.x];
    for (int k = threadIdx.x; k < c.count; k += blockDim.x) vec[c.dst + k] = c.src >= 0 ? params[c.src + k] : 0.f;
}

// ---------------------------------------------------------------------------------------------- host: the NeRF architecture
struct NerfArch {
    int d_x, n_freq_x, n_freq_v, W, D, skip, W_rad;
    int embed_x, embed_v, xe_pad, aux_cols;
    // flat parameter layout: geometry layers 0..D (weight (out x in) row-major, then bias), radiance layers 0..1
    int64_t w_off[kMaxLayers], b_off[kMaxLayers];
    int in_dim[kMaxLayers], out_dim[kMaxLayers];
    int n_lin;              // D + 1 + 2
    int64_t n_params;
};

int make_arch(NerfArch &A, int d_x, int n_freq_x, int n_freq_v, int W, int D, int skip, int W_rad) {
    if (W != 256 || D < 2 || D > 8 || skip > D - 2 || (W_rad != 128 && W_rad != 64 && W_rad != 256)) return -1;
    if (d_x < 1 || d_x > 4 || n_freq_x < 0 || n_freq_x > 15 || n_freq_v < 0 || n_freq_v > 4) return -1;
    A.d_x = d_x; A.n_freq_x = n_freq_x; A.n_freq_v = n_freq_v; A.W = W; A.D = D; A.skip = skip; A.W_rad = W_rad;
    A.embed_x = d_x * (2 * n_freq_x + 1);
    A.embed_v = 3 * (2 * n_freq_v + 1);
    if (A.embed_x > 96 || A.embed_v > 32) return -1;
    A.xe_pad = (A.embed_x + 31) / 32 * 32;
    A.aux_cols = A.xe_pad < 32 ? 32 : A.xe_pad;
    int64_t off = 0;
    int li = 0;
    for (int i = 0; i <= D; ++i, ++li) {
        A.in_dim[li] = i == 0 ? A.embed_x : ((skip >= 0 && i - 1 == skip) ? W + A.embed_x : W);
        A.out_dim[li] = i == D ? 1 + W : W;
        A.w_off[li] = off; off += (int64_t)A.in_dim[li] * A.out_dim[li];
        A.b_off[li] = off; off += A.out_dim[li];
    }
    A.in_dim[li] = A.embed_v + W; A.out_dim[li] = W_rad; A.w_off[li] = off; off += (int64_t)A.in_dim[li] * W_rad; A.b_off[li] = off; off += W_rad; ++li;
    A.in_dim[li] = W_rad; A.out_dim[li] = 3; A.w_off[li] = off; off += (int64_t)W_rad * 3; A.b_off[li] = off; off += 3; ++li;
    A.n_lin = li;
    A.n_params = off;
    return 0;
}

// vec layout: [biases of every linear layer, in order, each padded to 16][sigma weights W + sigma bias (padded to W+16)]
struct VecLayout { int32_t bias[kMaxLayers]; int32_t sigma; int32_t total; };
void make_vec_layout(const NerfArch &A, VecLayout &V) {
    int off = 0;
    for (int l = 0; l < A.n_lin; ++l) { V.bias[l] = off; off += (A.out_dim[l] + 15) / 16 * 16 + 16; }
    V.sigma = off; off += A.W + 16;
    V.total = off;
}

static int n_slots_for(const NerfArch &A) {
    const size_t tile = (size_t)kRows * (kHCols + A.aux_cols) * 2;
    const long long left = 227ll * 1024 - 256 - 2 * 256 * 4 - 2 * (long long)tile;
    int s = (int)(left / kSlotBytes);
    return s > kMaxSlots ? kMaxSlots : s;
}

// forward program + its packing plan.  Save buffer columns (forward): [x embedding xe_pad | view embedding 32 | H_0 .. H_{D-1}
// (W each) | feature W | radiance hidden W_rad]
void build_forward(const NerfArch &A, const VecLayout &V, Program &P, PackPlan &K, VecPlan &C) {
    memset(&P, 0, sizeof(P)); memset(&K, 0, sizeof(K)); memset(&C, 0, sizeof(C));
    P.aux_cols = A.aux_cols; P.d_x = A.d_x; P.n_freq_x = A.n_freq_x; P.embed_x = A.embed_x; P.n_freq_v = A.n_freq_v; P.embed_v = A.embed_v;
    P.n_slots = n_slots_for(A);
    int col = 0;
    P.xe_col = col; col += A.xe_pad;
    P.ve_col = col; col += 32;
    uint32_t w_off = 0;
    const int aux_cg0 = kHCols / 8;
    int L = 0;
    auto add = [&](int lin, int N, int act, int flags, int save_cols, int row0, int r_valid, int n_h_chunks, int h_col0, int n_aux_chunks,
                   int aux_col0, int aux_valid) {
        Layer &ly = P.layer[L]; PackLayer &pk = K.layer[L];
        ly.w_off = w_off; ly.N = (uint16_t)N; ly.act = (uint8_t)act; ly.flags = (uint8_t)flags;
        ly.bias_off = V.bias[lin] + (row0 ? 0 : 0); ly.mask_col = -1; ly.vec_off = -1;
        ly.save_col = save_cols > 0 ? col : -1;
        if (save_cols > 0) col += save_cols;
        pk.w_off = w_off; pk.N = N; pk.r_valid = r_valid; pk.stride_r = A.in_dim[lin]; pk.stride_j = 1;
        int c = 0;
        for (int q = 0; q < n_h_chunks; ++q, ++c) {
            ly.src_cg[c] = (uint8_t)(q * 4);
            pk.base[c] = A.w_off[lin] + (int64_t)row0 * A.in_dim[lin] + h_col0 + q * kChunkK;
            pk.j_valid[c] = kChunkK;
        }
        for (int q = 0; q < n_aux_chunks; ++q, ++c) {
            ly.src_cg[c] = (uint8_t)(aux_cg0 + q * 4);
            pk.base[c] = A.w_off[lin] + (int64_t)row0 * A.in_dim[lin] + aux_col0 + q * kChunkK;
            const int left = aux_valid - q * kChunkK;
            pk.j_valid[c] = left > kChunkK ? kChunkK : (left > 0 ? left : 0);
        }
        ly.n_chunks = (uint8_t)c; pk.n_chunks = c;
        w_off += (uint32_t)c * N * kChunkK * 2;
        ++L;
    };
    const int W = A.W, xe_chunks = A.xe_pad / kChunkK;
    for (int i = 0; i < A.D; ++i) {
        if (i == 0) add(0, W, ACT_RELU, 0, W, 0, W, 0, 0, xe_chunks, 0, A.embed_x);
        else if (A.skip >= 0 && i - 1 == A.skip) add(i, W, ACT_RELU, 0, W, 0, W, W / kChunkK, 0, xe_chunks, W, A.embed_x);   // [h | embedding]
        else add(i, W, ACT_RELU, 0, W, 0, W, W / kChunkK, 0, 0, 0, 0);
    }
    // last geometry layer: rows 1..W = feature (tensor core), row 0 = density (CUDA cores, F_SIGMA); then the view embedding
    add(A.D, W, ACT_NONE, F_SIGMA | F_VIEW, W, 1, W, W / kChunkK, 0, 0, 0, 0);
    P.layer[L - 1].bias_off = V.bias[A.D] + 16;          // feature biases are stored after a 16-float slot holding the sigma bias
    P.layer[L - 1].vec_off = V.sigma;
    // radiance: input [view embedding (embed_v) | feature (W)] -> W_rad -> 3
    add(A.D + 1, A.W_rad, ACT_RELU, 0, A.W_rad, 0, A.W_rad, W / kChunkK, A.embed_v, 1, 0, A.embed_v);
    add(A.D + 2, 16, ACT_SIGMOID, F_OUT_RGB, 0, 0, 3, A.W_rad / kChunkK, 0, 0, 0, 0);
    P.n_layers = L; K.n_layers = L;
    P.save_cols = col;
    P.blob_bytes = w_off;
    // vectors: biases (the last geometry layer's bias is split: [sigma bias | 15 pad | feature biases]), sigma row
    int n = 0;
    for (int l = 0; l < A.n_lin; ++l) {
        if (l == A.D) {
            C.c[n++] = VecCopy{A.b_off[l], V.bias[l], 1};
            C.c[n++] = VecCopy{A.b_off[l] + 1, V.bias[l] + 16, A.W};
        } else {
            C.c[n++] = VecCopy{A.b_off[l], V.bias[l], A.out_dim[l]};
            const int pad = (A.out_dim[l] + 15) / 16 * 16 - A.out_dim[l];
            if (pad) C.c[n++] = VecCopy{-1, V.bias[l] + A.out_dim[l], pad};
        }
    }
    C.c[n++] = VecCopy{A.w_off[A.D], V.sigma, A.W};             // row 0 of the last geometry layer
    C.c[n++] = VecCopy{A.b_off[A.D], V.sigma + A.W, 1};
    C.n = n;
}

// backward (data-gradient) program.  Save buffer columns (backward): [dZ of the colour layer 16 | dZ radiance hidden W_rad |
// dZ feature W | dZ_{D-1} .. dZ_0 (W each)]; the forward's save columns give the ReLU masks.
void build_backward(const NerfArch &A, const VecLayout &V, const Program &F, Program &P, PackPlan &K) {
    memset(&P, 0, sizeof(P)); memset(&K, 0, sizeof(K));
    P.aux_cols = A.aux_cols; P.d_x = A.d_x; P.embed_x = A.embed_x; P.embed_v = A.embed_v;
    P.n_slots = n_slots_for(A);
    const int W = A.W;
    int col = 16;            // region 0: dZ of the colour layer, written by the load phase
    uint32_t w_off = 0;
    int L = 0;
    // tile row r <- weight column (c0 + r); chunk column j <- weight row (row0 + 32 c + j)
    auto add = [&](int lin, int N, int c0, int row0, int k_rows, int flags, int mask_col, int save_cols) {
        Layer &ly = P.layer[L]; PackLayer &pk = K.layer[L];
        ly.w_off = w_off; ly.N = (uint16_t)N; ly.act = ACT_NONE; ly.flags = (uint8_t)flags; ly.bias_off = -1;
        ly.mask_col = mask_col; ly.vec_off = (flags & F_RANK1) ? V.sigma : -1;
        ly.save_col = col; col += save_cols;
        pk.w_off = w_off; pk.N = N; pk.r_valid = N; pk.stride_r = 1; pk.stride_j = A.in_dim[lin];
        const int nc = (k_rows + kChunkK - 1) / kChunkK;
        for (int c = 0; c < nc; ++c) {
            ly.src_cg[c] = (uint8_t)(c * 4);
            pk.base[c] = A.w_off[lin] + (int64_t)(row0 + c * kChunkK) * A.in_dim[lin] + c0;
            const int left = k_rows - c * kChunkK;
            pk.j_valid[c] = left > kChunkK ? kChunkK : left;
        }
        ly.n_chunks = (uint8_t)nc; pk.n_chunks = nc;
        w_off += (uint32_t)nc * N * kChunkK * 2;
        ++L;
    };
    // forward save columns of H_i
    auto h_col = [&](int i) { return F.layer[i].save_col; };
    const int lin_r1 = A.D + 2, lin_r0 = A.D + 1;
    add(lin_r1, A.W_rad, 0, 0, 3, F_MASK, F.layer[A.D + 1].save_col, A.W_rad);                 // -> dZ radiance hidden
    add(lin_r0, W, A.embed_v, 0, A.W_rad, 0, -1, W);                                           // -> dZ feature (linear)
    add(A.D, W, 0, 1, W, F_MASK | F_RANK1, h_col(A.D - 1), W);                                 // -> dZ_{D-1}
    for (int i = A.D - 1; i >= 1; --i) add(i, W, 0, 0, W, F_MASK, h_col(i - 1), W);            // layer i transposed -> dZ_{i-1}
    P.n_layers = L; K.n_layers = L;
    P.save_cols = col;
    P.blob_bytes = w_off;
}

static size_t chain_smem(const Program &P) {
    return 256 + 2 * (size_t)kRows * (kHCols + P.aux_cols) * 2 + (size_t)P.n_slots * kSlotBytes + 2 * 256 * sizeof(float);
}

template <typename K>
static int set_max_smem(K kernel, size_t bytes) {
    if (bytes > 227 * 1024) { arcb200_set_error("dense chain: %zu bytes of shared memory needed, 227 KB available", bytes); return ARCB200_ERR_LIMIT; }
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) { arcb200_set_error("cudaFuncSetAttribute -> %s", cudaGetErrorString(e)); return ARCB200_ERR_CUDA; }
    return ARCB200_OK;
}

struct Built { NerfArch A; VecLayout V; Program F, B; PackPlan KF, KB; VecPlan C; };
static int build_all(Built &b, int d_x, int n_freq_x, int n_freq_v, int W, int D, int skip, int W_rad) {
    if (make_arch(b.A, d_x, n_freq_x, n_freq_v, W, D, skip, W_rad) != 0) return -1;
    make_vec_layout(b.A, b.V);
    build_forward(b.A, b.V, b.F, b.KF, b.C);
    build_backward(b.A, b.V, b.F, b.B, b.KB);
    if (b.F.n_slots < 2) return -1;
    return 0;
}

static void add_wgrad(WgJobs &J, const __half *dz, int64_t ldz, int N, int n_out_valid, const __half *act, int64_t lda, int K, int k_valid,
                      float *dW, int64_t ldw, float *db, float out_scale, int64_t n) {
    WgArgs &w = J.job[J.n_jobs++];
    w.dz = dz; w.ldz = ldz; w.N = N; w.act = act; w.lda = lda; w.K = K; w.n_out_valid = n_out_valid; w.k_valid = k_valid;
    w.dW = dW; w.ldw = ldw; w.db = db; w.out_scale = out_scale; w.n = n;
}

// CTAs per job in proportion to the cost of one of its 64-row iterations, which the staging of the two operands and the
// fixed per-iteration work dominate (measured: about 3100 cycles for 256 + 256 columns, of which 1000 are tensor-core time),
// at least one each, all SMs used
static int launch_wgrad_jobs(WgJobs &J, cudaStream_t st) {
    const int sms = arc_sm_count();
    double work[kMaxWgJobs], total = 0.0;
    size_t smem = 0;
    for (int j = 0; j < J.n_jobs; ++j) {
        const int Mp = (J.job[j].N + 127) / 128 * 128;
        work[j] = (double)(Mp + J.job[j].K) + 128.0;
        total += work[j];
        const size_t need = 128 + 1024 + (size_t)kWgStages * kWgRows * (Mp + J.job[j].K) * 2;
        smem = need > smem ? need : smem;
    }
    const int64_t blocks = (J.job[0].n + kWgRows - 1) / kWgRows;
    int used = 0;
    J.first[0] = 0;
    for (int j = 0; j < J.n_jobs; ++j) {
        int c = (int)(work[j] / total * (sms - J.n_jobs)) + 1;
        if (c > blocks) c = (int)blocks;
        used += c;
        J.first[j + 1] = used;
    }
    int rc = set_max_smem(dense_wgrad_kernel, smem);
    if (rc != ARCB200_OK) return rc;
    arc_launch(dense_wgrad_kernel, (unsigned)used, kWgThreads, smem, st, J);
    return ARCB200_OK;
}

}  // namespace

extern "C" {

#define NERF_ARCH_ARGS int d_x, int n_freq_x, int n_freq_v, int W, int D, int skip, int W_rad
#define NERF_ARCH_PASS d_x, n_freq_x, n_freq_v, W, D, skip, W_rad

// Sizes of everything a caller allocates for the vanilla-NeRF network pair of the given architecture:
// out[0] parameters (floats, flat: geometry layers 0..D as weight (out x in) then bias, radiance layers likewise),
// out[1] forward blob bytes, out[2] backward blob bytes, out[3] vector floats, out[4] forward save columns (fp16 per point),
// out[5] backward save columns.  Returns 0, or a negative code for an unsupported architecture.
int arcb200_nerf_mlp_sizes(NERF_ARCH_ARGS, int64_t *out) {
    Built b;
    ARC_REQUIRE(out != nullptr, "nerf_mlp_sizes: null pointer");
    ARC_REQUIRE(build_all(b, NERF_ARCH_PASS) == 0,
                "nerf_mlp: unsupported architecture (W = 256, 2 <= D <= 8, one skip or none, W_rad in {64,128,256}, embeddings <= 96 / 32 columns)");
    out[0] = b.A.n_params; out[1] = b.F.blob_bytes; out[2] = b.B.blob_bytes; out[3] = b.V.total; out[4] = b.F.save_cols; out[5] = b.B.save_cols;
    return ARCB200_OK;
}

// params (flat fp32, layout above) -> operand chunks of the forward chain (blob_fwd), of the data-gradient chain (blob_bwd,
// may be NULL for inference) and the bias / row vectors (vec).  Run after every parameter update.
int arcb200_nerf_mlp_pack(NERF_ARCH_ARGS, const float *params, void *blob_fwd, void *blob_bwd, float *vec, void *stream) {
    Built b;
    ARC_REQUIRE(build_all(b, NERF_ARCH_PASS) == 0, "nerf_mlp_pack: unsupported architecture");
    ARC_REQUIRE(params && blob_fwd && vec && (((uintptr_t)blob_fwd | (uintptr_t)blob_bwd | (uintptr_t)vec) & 15u) == 0,
                "nerf_mlp_pack: null or misaligned pointer");
    cudaStream_t st = (cudaStream_t)stream;
    arc_launch(chain_pack_kernel, dim3(64, b.KF.n_layers), 256, 0, st, b.KF, params, (uint8_t *)blob_fwd);
    if (blob_bwd) arc_launch(chain_pack_kernel, dim3(64, b.KB.n_layers), 256, 0, st, b.KB, params, (uint8_t *)blob_bwd);
    arc_launch(chain_vec_kernel, b.C.n, 128, 0, st, b.C, params, vec);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Forward of GeoNet + RadianceNet with the frequency embeddings fused in: x (n, d_x) fp32 raw inputs, dirs (n,3) fp32 (any
// length, normalised inside like RadianceNet does) -> sigma (n) fp32 raw density output, rgb (n,3) fp32 after the sigmoid.
// save: NULL (inference) or fp16 buffer of (forward save columns) x n_cap elements kept for arcb200_nerf_mlp_bwd.
// Replaces GeoNet.forward + RadianceNet.forward (linear_network_module.py:174-197,317-335) incl. FreqEmbedder.forward.
int arcb200_nerf_mlp_fwd(NERF_ARCH_ARGS, const float *x, const float *dirs, const void *blob_fwd, const float *vec, void *save,
                         int64_t n_cap, float *sigma, float *rgb, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0 && n_cap >= n, "nerf_mlp_fwd: bad sizes");
    if (n == 0) return ARCB200_OK;
    Built b;
    ARC_REQUIRE(build_all(b, NERF_ARCH_PASS) == 0, "nerf_mlp_fwd: unsupported architecture");
    ARC_REQUIRE(x && dirs && blob_fwd && vec && sigma && rgb, "nerf_mlp_fwd: null pointer");
    ARC_REQUIRE((((uintptr_t)blob_fwd | (uintptr_t)vec | (uintptr_t)save) & 15u) == 0 && (n_cap % 128) == 0,
                "nerf_mlp_fwd: blob / vec / save must be 16-byte aligned and n_cap a multiple of 128");
    ChainArgs a;
    memset(&a, 0, sizeof(a));
    a.x = x; a.dirs = dirs; a.blob = (const uint8_t *)blob_fwd; a.vec = vec; a.save = (__half *)save; a.sigma = sigma; a.rgb = rgb;
    a.n = n; a.n_cap = n_cap;
    if (const char *e = getenv("ARCB200_CHAIN_DBG")) a.dbg_skip = atoi(e);
    const size_t smem = chain_smem(b.F);
    int rc = set_max_smem(dense_chain_kernel<false>, smem);
    if (rc != ARCB200_OK) return rc;
    const int64_t pairs = (n + 2 * kRows - 1) / (2 * kRows), cap = arc_sm_count();
    arc_launch(dense_chain_kernel<false>, (unsigned)(pairs < cap ? pairs : cap), kThreads, smem, (cudaStream_t)stream, b.F, a);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Backward: d_sigma (n), d_rgb (n,3) fp32 -> d_params (flat fp32, ACCUMULATED; the caller zeroes it).  fwd_save / rgb: what
// arcb200_nerf_mlp_fwd left behind for the same inputs; bwd_save: fp16 scratch of (backward save columns) x n_cap; dsig16:
// fp16 scratch (n_cap, 16).  loss_scale: the gradients pass through fp16 tiles multiplied by it and are divided by it on the way
// out (128 is what the reference's tcnn path uses; values below 6e-5 would otherwise lose bits).  No gradient w.r.t. the inputs
// is produced (NeRF does not need it).
// Replaces autograd through the torch Linear stacks of the reference.
int arcb200_nerf_mlp_bwd(NERF_ARCH_ARGS, const void *blob_bwd, const float *vec, const void *fwd_save, const float *rgb,
                         const float *d_sigma, const float *d_rgb, float loss_scale, void *bwd_save, void *dsig16, int64_t n_cap,
                         float *d_params, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0 && n_cap >= n && loss_scale > 0.f, "nerf_mlp_bwd: bad sizes or loss scale");
    if (n == 0) return ARCB200_OK;
    Built b;
    ARC_REQUIRE(build_all(b, NERF_ARCH_PASS) == 0, "nerf_mlp_bwd: unsupported architecture");
    ARC_REQUIRE(blob_bwd && vec && fwd_save && rgb && d_sigma && d_rgb && bwd_save && dsig16 && d_params, "nerf_mlp_bwd: null pointer");
    ARC_REQUIRE((((uintptr_t)blob_bwd | (uintptr_t)vec | (uintptr_t)fwd_save | (uintptr_t)bwd_save | (uintptr_t)dsig16) & 15u) == 0 &&
                (n_cap % 128) == 0, "nerf_mlp_bwd: buffers must be 16-byte aligned and n_cap a multiple of 128");
    cudaStream_t st = (cudaStream_t)stream;
    ChainArgs a;
    memset(&a, 0, sizeof(a));
    a.blob = (const uint8_t *)blob_bwd; a.vec = vec; a.save = (__half *)bwd_save; a.fwd_save = (const __half *)fwd_save;
    a.rgb = const_cast<float *>(rgb); a.d_sigma = d_sigma; a.d_rgb = d_rgb; a.dsig16 = (__half *)dsig16; a.loss_scale = loss_scale; a.n = n; a.n_cap = n_cap;
    const size_t smem = chain_smem(b.B);
    int rc = set_max_smem(dense_chain_kernel<true>, smem);
    if (rc != ARCB200_OK) return rc;
    const int64_t pairs = (n + 2 * kRows - 1) / (2 * kRows), cap = arc_sm_count();
    const char *parts_env = getenv("ARCB200_NERF_BWD_PARTS");           // profiling aid: 1 = data-gradient chain only, 2 = weight gradients only
    const int parts = parts_env ? atoi(parts_env) : 3;
    if (parts & 1) arc_launch(dense_chain_kernel<true>, (unsigned)(pairs < cap ? pairs : cap), kThreads, smem, st, b.B, a);
    ARC_LAUNCH_CHECK();
    if (!(parts & 2)) return ARCB200_OK;

    // weight and bias gradients, one batch-contraction GEMM per (layer, input source)
    const NerfArch &A = b.A;
    const __half *fs = (const __half *)fwd_save, *bs = (const __half *)bwd_save;
    auto fcol = [&](int c) { return fs + (size_t)c * n_cap; };
    auto bcol = [&](int c) { return bs + (size_t)c * n_cap; };
    const __half *xe = fcol(b.F.xe_col), *ve = fcol(b.F.ve_col);
    auto H = [&](int i) { return fcol(b.F.layer[i].save_col); };            // forward layer i output (i = D: feature, D+1: radiance hidden)
    // backward save regions: 0: dZ colour (ld 16) | 1: dZ radiance hidden | 2: dZ feature | 3..: dZ_{D-1} .. dZ_0
    const __half *dz_rgb = bcol(0), *dz_rh = bcol(b.B.layer[0].save_col), *dz_feat = bcol(b.B.layer[1].save_col);
    auto dZ = [&](int i) { return bcol(b.B.layer[2 + (D - 1 - i)].save_col); };
    float *g = d_params;
    const int l_r0 = D + 1, l_r1 = D + 2;
    WgJobs J;
    memset(&J, 0, sizeof(J));
#define WG(...) add_wgrad(J, __VA_ARGS__, 1.0f / loss_scale, n)
    // colour layer (3 x W_rad)
    WG(dz_rgb, 16, 16, 3, H(D + 1), A.W_rad, A.W_rad, A.W_rad, g + A.w_off[l_r1], A.W_rad, g + A.b_off[l_r1]);
    // radiance hidden layer (W_rad x (embed_v + W)): view columns, then feature columns
    WG(dz_rh, A.W_rad, A.W_rad, A.W_rad, ve, 32, 32, A.embed_v, g + A.w_off[l_r0], A.in_dim[l_r0], g + A.b_off[l_r0]);
    WG(dz_rh, A.W_rad, A.W_rad, A.W_rad, H(D), W, W, W, g + A.w_off[l_r0] + A.embed_v, A.in_dim[l_r0], nullptr);
    // last geometry layer: rows 1..W from dZ feature, row 0 from d_sigma
    WG(dz_feat, W, W, W, H(D - 1), W, W, W, g + A.w_off[D] + W, W, g + A.b_off[D] + 1);
    WG((const __half *)dsig16, 16, 16, 1, H(D - 1), W, W, W, g + A.w_off[D], W, g + A.b_off[D]);
    for (int i = D - 1; i >= 0; --i) {
        if (i == 0) {
            WG(dZ(0), W, W, W, xe, A.xe_pad, A.xe_pad, A.embed_x, g + A.w_off[0], A.in_dim[0], g + A.b_off[0]);
        } else {
            WG(dZ(i), W, W, W, H(i - 1), W, W, W, g + A.w_off[i], A.in_dim[i], g + A.b_off[i]);
            if (A.skip >= 0 && i - 1 == A.skip)
                WG(dZ(i), W, W, W, xe, A.xe_pad, A.xe_pad, A.embed_x, g + A.w_off[i] + W, A.in_dim[i], nullptr);
        }
    }
#undef WG
    rc = launch_wgrad_jobs(J, st);
    if (rc != ARCB200_OK) return rc;
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

#ifdef ARCB200_PROF
int arcb200_debug_wg_prof(unsigned long long *out8) {
    ARC_CUDA(cudaMemcpyFromSymbol(out8, g_wg_prof, sizeof(unsigned long long) * 8));
    unsigned long long z[8] = {0};
    ARC_CUDA(cudaMemcpyToSymbol(g_wg_prof, z, sizeof(z)));
    return ARCB200_OK;
}
int arcb200_debug_chain_prof(unsigned long long *out8) {
    ARC_CUDA(cudaMemcpyFromSymbol(out8, g_chain_prof, sizeof(unsigned long long) * 8));
    ARC_CUDA(cudaMemcpyFromSymbol(out8 + 8, g_chain_layer_prof, sizeof(unsigned long long) * 16));
    ARC_CUDA(cudaMemcpyFromSymbol(out8 + 24, g_chain_load_prof, sizeof(unsigned long long) * 2));
    unsigned long long z[16] = {0};
    ARC_CUDA(cudaMemcpyToSymbol(g_chain_prof, z, sizeof(unsigned long long) * 8));
    ARC_CUDA(cudaMemcpyToSymbol(g_chain_layer_prof, z, sizeof(z)));
    ARC_CUDA(cudaMemcpyToSymbol(g_chain_load_prof, z, sizeof(unsigned long long) * 2));
    return ARCB200_OK;
}
#endif

}  // extern "C"
