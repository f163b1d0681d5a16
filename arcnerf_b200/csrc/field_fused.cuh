// Fused field forward of the NGP foreground model: hash-grid encode -> geometry head (32-64-16, TruncExp density) ->
// [feature | SH(view)] -> radiance head (32-64-64-3, sigmoid), one launch, nothing between the stages touches HBM.
// Included by mlp_tc.cu (it uses that file's tile / MMA / activation helpers).
//
// Reference: what FgModel.get_sigma_radiance_by_mask_pts -> Base3dModel._forward_pts_dir runs per chunk of samples
// (arcnerf/models/fg_model.py:264-318, base_3d_model.py:233-254): FusedMLPGeoNet.forward (HashGridEmbedder -> tcnn.Network,
// TruncExp on channel 0; tcnn_fusedmlp_module.py:81-111), then FusedMLPRadianceNet.forward with mode 'fv'
// (SHEmbedder of the normalised view direction, concatenation [geo_feat, sh], tcnn.Network with a sigmoid output; :177-192).
// Arithmetic is that of the stand-alone kernels (grid_fwd_kernel, mlp_fwd_kernel): fp16 table blended in fp32 and rounded
// once, fp16 activations, fp32 accumulation in tensor memory - the fused launch returns the same values.
//
// Shape of the kernel.  One persistent CTA per SM, 512 threads in two roles, 128-sample tiles:
//   gather warps (8)  warp g encodes levels 2g, 2g+1 for samples lane, lane+32, lane+64, lane+96 (32 independent 4-byte
//            gathers in flight per thread and level) straight into the K-major operand tile of the first layer, double
//            buffered in shared memory, and run ahead of the layers by one tile (full / empty mbarriers).  The gathers are
//            what bounds the kernel: the SM's L2 sector rate (measured: one tile per ~6.6k cycles whatever else runs), so
//            these warps must never wait for the layers.
//   layer warps (8)   one elected thread issues the tcgen05.mma of a layer (M=128, N=64 or 16, K=16 per instruction) and
//            commits to an mbarrier; the eight warps then read their accumulator rows out of TMEM (row = 32*(warp%4)+lane,
//            the two warps of a row quadrant split the columns), apply the activation and write the next layer's fp16
//            operand tile.  The SH encoding of the view direction is computed while the first layer's MMAs run.  The five
//            layers of a tile take less than one tile's gathers, so they hide behind them.
// Training additionally stores what the backward kernels read: the encoding (n,32) fp16, the radiance input (n,32) fp16 and
// the hidden activations of both heads in operand-tile layout (same buffers the unfused path hands over).

constexpr int kFieldLayerThreads = 256;     // warps 0..7: the layers (TMEM row quadrant = warp % 4)
constexpr int kFieldGatherWarps = 8;        // warps 8..15: the gathers
constexpr int kFieldThreads = kFieldLayerThreads + 32 * kFieldGatherWarps;

#ifdef ARCB200_PROF
// phase clocks of the fused forward (debug builds only): [0] tiles, [1] wait for the gathered tile, [3..7] the five layers,
// [8] layer side total, [9] gather side: wait for a free buffer, [10] gather, [11] gather side total
__device__ unsigned long long g_field_prof[16];
#define FIELD_PROF_MARK(slot) do { if (tid == 0) { const long long t__ = clock64(); atomicAdd(&g_field_prof[slot], (unsigned long long)(t__ - prof_t)); prof_t = t__; } } while (0)
#define FIELD_PROF_MARK_G(slot) do { if (tid == kFieldLayerThreads) { const long long t__ = clock64(); atomicAdd(&g_field_prof[slot], (unsigned long long)(t__ - prof_t)); prof_t = t__; } } while (0)
#else
#define FIELD_PROF_MARK(slot) do { } while (0)
#define FIELD_PROF_MARK_G(slot) do { } while (0)
#endif

struct FieldFwdArgs {
    const float *unit;          // (n,3) positions in the unit cube
    const __half *table;        // fp16 hash table
    const uint8_t *geo_blob, *rad_blob;
    const float *dirs;          // (n_rays,3) view directions
    const int32_t *ray_id;      // (n) ray of each sample, or null: dirs is per sample
    int sh_remap;
    __half *enc_out;            // (n,32) or null
    __half *rad_in_out;         // (n,32) or null
    uint8_t *geo_act, *rad_act; // saved hidden activations or null
    float *sigma;               // (n)
    float *rgb;                 // (n,3)
    float *zero_scalar;         // optional float cleared by the launch
};

struct FieldShared {
    uint64_t mbar_w;
    uint64_t mbar_mma;
    uint64_t full[2];           // gathered tile b is complete (one arrival per gather warp)
    uint64_t empty[2];          // the layers no longer read tile b
    uint32_t tmem_base;
    uint32_t pad;
};

// fp16 pair of 16 floats -> two 16-byte words
__device__ __forceinline__ void pack16(const float (&v)[16], uint4 &w0, uint4 &w1) {
    w0.x = tc05::pack_half2(v[0], v[1]);   w0.y = tc05::pack_half2(v[2], v[3]);
    w0.z = tc05::pack_half2(v[4], v[5]);   w0.w = tc05::pack_half2(v[6], v[7]);
    w1.x = tc05::pack_half2(v[8], v[9]);   w1.y = tc05::pack_half2(v[10], v[11]);
    w1.z = tc05::pack_half2(v[12], v[13]); w1.w = tc05::pack_half2(v[14], v[15]);
}

// barrier among the layer warps only (the gather warps never join it)
__device__ __forceinline__ void layer_bar() { asm volatile("bar.sync 1, %0;" :: "n"(kFieldLayerThreads) : "memory"); }

__global__ void __launch_bounds__(kFieldThreads, 1)
field_fwd_kernel(int64_t n, const int32_t *__restrict__ n_dev, const __grid_constant__ gridmath::GridGeom geom,
                 const __grid_constant__ MlpCfg geo, const __grid_constant__ MlpCfg rad, const FieldFwdArgs a) {
    using namespace gridmath;
    extern __shared__ __align__(128) uint8_t smem[];
    FieldShared *sh = reinterpret_cast<FieldShared *>(smem);
    uint8_t *wg_s = smem + 128;                                            // geometry-head weights
    uint8_t *wr_s = wg_s + ((geo.w_bytes + 127u) & ~127u);                 // radiance-head weights
    uint8_t *enc_s = wr_s + ((rad.w_bytes + 127u) & ~127u);                // 2 x (128 x 32) operand tiles: encoding
    uint8_t *rin_s = enc_s + 2 * kRows * 32 * 2;                           // 128 x 32 operand tile: [feat | SH]
    uint8_t *h_s = rin_s + kRows * 32 * 2;                                 // 128 x 64 operand tile: hidden activations
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 32) {
        tc05::mbar_init(&sh->mbar_w, 1);
        tc05::mbar_init(&sh->mbar_mma, 1);
        tc05::mbar_init(&sh->full[0], kFieldGatherWarps);
        tc05::mbar_init(&sh->full[1], kFieldGatherWarps);
        tc05::mbar_init(&sh->empty[0], 1);
        tc05::mbar_init(&sh->empty[1], 1);
        tc05::fence_mbar_init();
    }
    if (warp == 0) tc05::tmem_alloc<64>(&sh->tmem_base);
    tc05::fence_before_sync();
    __syncthreads();
    tc05::fence_after_sync();
    arc_pdl_sync();
    if (n_dev) { const int64_t nd = *n_dev; n = nd < n ? (nd > 0 ? nd : 0) : n; }
    const int64_t n_tiles = (n + kRows - 1) / kRows;

    if (warp >= kFieldLayerThreads / 32) {
        // =============================================================== gather warps
        const int gw = warp - kFieldLayerThreads / 32;
        int it = 0;
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
            const int b = it & 1;
#ifdef ARCB200_PROF
            long long prof_t = clock64();
            const long long prof_t0 = prof_t;
#endif
            tc05::mbar_wait(&sh->empty[b], (((uint32_t)it >> 1) & 1u) ^ 1u);      // passes at once on the first two tiles
            FIELD_PROF_MARK_G(9);
            uint8_t *dst_tile = enc_s + (size_t)b * (kRows * 32 * 2);
            const int64_t row0 = tile * kRows;
            float px[4], py[4], pz[4];
            #pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t r = row0 + lane + 32 * j;
                const bool ok = r < n;
                px[j] = ok ? __ldg(a.unit + 3 * r) : 0.0f;
                py[j] = ok ? __ldg(a.unit + 3 * r + 1) : 0.0f;
                pz[j] = ok ? __ldg(a.unit + 3 * r + 2) : 0.0f;
            }
            #pragma unroll 1
            for (int li = 0; li < 2; ++li) {
                const int level = 2 * gw + li;
                const GridLevel g = geom.lv[level];
                const __half2 *lt = reinterpret_cast<const __half2 *>(a.table) + g.offset;
                __half2 v[4][8];
                float fx[4], fy[4], fz[4];
                #pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const Cell c = locate(g, px[j], py[j], pz[j]);
                    fx[j] = c.fx; fy[j] = c.fy; fz[j] = c.fz;
                    #pragma unroll
                    for (int k = 0; k < 8; ++k)
                        v[j][k] = __ldg(lt + entry_index(g, c.gx + (k & 1), c.gy + ((k >> 1) & 1), c.gz + ((k >> 2) & 1)));
                }
                #pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float wx[2] = {1.0f - fx[j], fx[j]}, wy[2] = {1.0f - fy[j], fy[j]}, wz[2] = {1.0f - fz[j], fz[j]};
                    float a0 = 0.f, a1 = 0.f;
                    #pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const float wgt = (wx[k & 1] * wy[(k >> 1) & 1]) * wz[(k >> 2) & 1];
                        const float2 f = __half22float2(v[j][k]);
                        a0 = fmaf(wgt, f.x, a0); a1 = fmaf(wgt, f.y, a1);
                    }
                    // row s, columns 2*level, 2*level+1 of the K-major tile: 16-byte word (s, level/4), 4-byte slot level%4
                    uint32_t *word = reinterpret_cast<uint32_t *>(tile_word(dst_tile, lane + 32 * j, level >> 2));
                    word[level & 3] = tc05::pack_half2(a0, a1);
                }
            }
            tc05::fence_proxy_async();            // these shared-memory writes are read by the tensor core
            __syncwarp();
            if (lane == 0) mbar_arrive(&sh->full[b]);
            FIELD_PROF_MARK_G(10);
#ifdef ARCB200_PROF
            if (tid == kFieldLayerThreads) atomicAdd(&g_field_prof[11], (unsigned long long)(clock64() - prof_t0));
#endif
        }
    } else {
        // =============================================================== layer warps
        const int quad = warp & 3, half = warp >> 2;
        const int row = quad * 32 + lane;                                      // accumulator row of this thread
        if (a.zero_scalar && blockIdx.x == 0 && tid == 0) *a.zero_scalar = 0.0f;
        const uint32_t tmem = sh->tmem_base;
        if (tid == 0) {
            tc05::mbar_expect_tx(&sh->mbar_w, geo.w_bytes + rad.w_bytes);
            tc05::bulk_g2s(wg_s, a.geo_blob, geo.w_bytes, &sh->mbar_w);
            tc05::bulk_g2s(wr_s, a.rad_blob, rad.w_bytes, &sh->mbar_w);
        }
        const uint32_t enc_a = tc05::smem_u32(enc_s), rin_a = tc05::smem_u32(rin_s), h_a = tc05::smem_u32(h_s);
        const uint32_t wg_a = tc05::smem_u32(wg_s), wr_a = tc05::smem_u32(wr_s);
        const uint32_t my_tmem = tmem + ((uint32_t)(quad * 32) << 16);
        uint32_t phase = 0;
        bool weights_ready = false;

        // one layer: make the operand tile visible to the tensor core, issue, wait for the accumulators
        auto run_layer = [&](uint32_t a_tile, uint32_t w_tile, int K, int N) {
            tc05::fence_proxy_async();
            tc05::fence_before_sync();
            layer_bar();
            if (tid == 0) {
                tc05::fence_after_sync();
                issue_forward_mm(tmem, a_tile, w_tile, K, N);
                tc05::mma_commit(&sh->mbar_mma);
            }
        };
        auto wait_layer = [&]() {
            tc05::mbar_wait(&sh->mbar_mma, phase);
            phase ^= 1u;
            tc05::fence_after_sync();
        };
        // hidden layer epilogue: 64 accumulator columns -> ReLU -> fp16 operand tile (+ saved copy); warps of a quadrant
        // split the four 16-column blocks
        auto hidden_epilogue = [&](uint8_t *save_tile) {
            uint32_t r0[16], r1[16];
            const int cb0 = half * 2;
            tc05::tmem_ld16(my_tmem + (uint32_t)cb0 * 16, r0);
            tc05::tmem_ld16(my_tmem + (uint32_t)cb0 * 16 + 16, r1);
            tc05::tmem_ld_wait();
            float v[16];
            uint4 w0, w1;
            act16_fwd(ACT_RELU, r0, v);
            pack16(v, w0, w1);
            *tile_word(h_s, row, 2 * cb0) = w0;
            *tile_word(h_s, row, 2 * cb0 + 1) = w1;
            if (save_tile) { *tile_word(save_tile, row, 2 * cb0) = w0; *tile_word(save_tile, row, 2 * cb0 + 1) = w1; }
            act16_fwd(ACT_RELU, r1, v);
            pack16(v, w0, w1);
            *tile_word(h_s, row, 2 * cb0 + 2) = w0;
            *tile_word(h_s, row, 2 * cb0 + 3) = w1;
            if (save_tile) { *tile_word(save_tile, row, 2 * cb0 + 2) = w0; *tile_word(save_tile, row, 2 * cb0 + 3) = w1; }
        };

        int it = 0;
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
            const int b = it & 1;
            const int64_t grow = tile * kRows + row;
            uint8_t *enc_tile = enc_s + (size_t)b * (kRows * 32 * 2);
#ifdef ARCB200_PROF
            long long prof_t = clock64();
            const long long prof_t0 = prof_t;
            if (tid == 0) atomicAdd(&g_field_prof[0], 1ull);
#endif
            tc05::mbar_wait(&sh->full[b], ((uint32_t)it >> 1) & 1u);
            if (!weights_ready) { tc05::mbar_wait(&sh->mbar_w, 0); weights_ready = true; }
            FIELD_PROF_MARK(1);

            // ---- geometry head, layer 0 (32 -> 64); meanwhile SH of the view direction and the saved encoding
            run_layer(enc_a + (uint32_t)b * (kRows * 32 * 2), wg_a + geo.w_off[0], 32, 64);
            if (half == 1) {
                if (grow < n) {
                    const size_t src = a.ray_id ? (size_t)__ldg(a.ray_id + grow) : (size_t)grow;
                    float x = __ldg(a.dirs + 3 * src), y = __ldg(a.dirs + 3 * src + 1), z = __ldg(a.dirs + 3 * src + 2);
                    sh_input(x, y, z, a.sh_remap);
                    float o[25];
                    sh_basis(x, y, z, 4, o);
                    float v[16];
                    #pragma unroll
                    for (int k = 0; k < 16; ++k) v[k] = o[k];
                    uint4 w0, w1;
                    pack16(v, w0, w1);
                    *tile_word(rin_s, row, 2) = w0;
                    *tile_word(rin_s, row, 3) = w1;
                    if (a.rad_in_out) {
                        uint4 *dst = reinterpret_cast<uint4 *>(a.rad_in_out + grow * 32 + 16);
                        dst[0] = w0; dst[1] = w1;
                    }
                } else {
                    *tile_word(rin_s, row, 2) = make_uint4(0u, 0u, 0u, 0u);
                    *tile_word(rin_s, row, 3) = make_uint4(0u, 0u, 0u, 0u);
                }
            } else if (a.enc_out && grow < n) {
                uint4 *dst = reinterpret_cast<uint4 *>(a.enc_out + grow * 32);
                #pragma unroll
                for (int cg = 0; cg < 4; ++cg) dst[cg] = *tile_word(enc_tile, row, cg);
            }
            wait_layer();
            hidden_epilogue(a.geo_act ? a.geo_act + (size_t)tile * ((size_t)kRows * 64 * 2) : nullptr);
            FIELD_PROF_MARK(3);

            // ---- geometry head, layer 1 (64 -> 16): feature row + density
            run_layer(h_a, wg_a + geo.w_off[1], 64, 16);
            // every layer warp is past the barrier above: the tensor core and the copies are done with the gathered tile
            if (tid == 0) mbar_arrive(&sh->empty[b]);
            wait_layer();
            if (half == 0) {
                uint32_t r[16];
                tc05::tmem_ld16(my_tmem, r);
                tc05::tmem_ld_wait();
                float v[16];
                #pragma unroll
                for (int k = 0; k < 16; ++k) v[k] = __uint_as_float(r[k]);
                uint4 w0, w1;
                pack16(v, w0, w1);
                *tile_word(rin_s, row, 0) = w0;
                *tile_word(rin_s, row, 1) = w1;
                if (grow < n) {
                    if (a.rad_in_out) {
                        uint4 *dst = reinterpret_cast<uint4 *>(a.rad_in_out + grow * 32);
                        dst[0] = w0; dst[1] = w1;
                    }
                    a.sigma[grow] = expf(__half2float(__float2half_rn(v[0])));     // TruncExp forward on the stored value
                }
            }
            FIELD_PROF_MARK(4);

            // ---- radiance head: 32 -> 64 -> 64 -> 3 (padded to 16), sigmoid
            run_layer(rin_a, wr_a + rad.w_off[0], 32, 64);
            wait_layer();
            hidden_epilogue(a.rad_act ? a.rad_act + ((size_t)tile * 2) * ((size_t)kRows * 64 * 2) : nullptr);
            FIELD_PROF_MARK(5);
            run_layer(h_a, wr_a + rad.w_off[1], 64, 64);
            wait_layer();
            hidden_epilogue(a.rad_act ? a.rad_act + ((size_t)tile * 2 + 1) * ((size_t)kRows * 64 * 2) : nullptr);
            FIELD_PROF_MARK(6);
            run_layer(h_a, wr_a + rad.w_off[2], 64, 16);
            wait_layer();
            if (half == 0) {                 // warp-uniform: tcgen05.ld is a .sync.aligned instruction, every lane must execute it
                uint32_t r[16];
                tc05::tmem_ld16(my_tmem, r);
                tc05::tmem_ld_wait();
                if (grow < n) {
                    #pragma unroll
                    for (int k = 0; k < 3; ++k) a.rgb[grow * 3 + k] = 1.0f / (1.0f + __expf(-__uint_as_float(r[k])));
                }
            }
            FIELD_PROF_MARK(7);
#ifdef ARCB200_PROF
            if (tid == 0) atomicAdd(&g_field_prof[8], (unsigned long long)(clock64() - prof_t0));
#endif
            tc05::fence_before_sync();      // the next tile's first MMA must not overtake these accumulator reads
        }
        if (!weights_ready && tid == 0) tc05::mbar_wait(&sh->mbar_w, 0);      // a CTA without tiles: let the weight copy land before exit
    }
    tc05::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc05::tmem_dealloc<64>(sh->tmem_base);
}

static size_t field_fwd_smem(const MlpCfg &geo, const MlpCfg &rad) {
    return 128 + ((geo.w_bytes + 127u) & ~127u) + ((rad.w_bytes + 127u) & ~127u) + (size_t)kRows * (2 * 32 + 32 + 64) * 2;
}
