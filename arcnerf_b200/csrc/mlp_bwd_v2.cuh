// Backward of the fused MLP for the training hot path (ReLU hidden layers of width 64, fp16 inputs, hidden
// activations saved by the forward): several 128-row tiles in flight per SM.
//
// Included by mlp_tc.cu inside its anonymous namespace (uses MlpCfg, BwdPlan, BwdExtra, issue_*_mm, tile_word).
//
// Why a second kernel: the chain of a tile (gradient tile -> MMA -> TMEM read-back -> next gradient tile -> ...) is a
// sequence of short dependent phases, so one tile keeps neither the tensor pipe nor the load/store units busy; the
// throughput of an SM is the number of tiles it has in flight divided by the chain latency.  The generic kernel holds
// one tile per CTA and is limited to 2-3 CTAs per SM by tensor memory (every CTA owns a full set of weight-gradient
// accumulators).  Here ONE CTA per SM runs NWG warpgroups, each owning a tile slot (operand tiles in shared memory +
// 64 TMEM columns for the data gradient), and a control warp whose elected lane issues every tcgen05.mma of the CTA.
// The weight-gradient accumulators are shared by all slots (one issuing thread -> the accumulations are ordered), the
// packed weights are staged once per SM, and the final atomic flush happens once per SM instead of once per tile slot.
//
// Shared memory per slot: X tile (128 x in_pad fp16), the saved hidden activations H_0..H_{L-2} (128 x 64 fp16 each,
// bulk copies from the forward's save buffer) and the 16-column gradient tile at the network output.  The gradient
// tile of a hidden layer is written IN PLACE over that layer's activation tile (row r by thread r) once the MMAs that
// read the activations have completed, so no further buffer is needed.
//
// Per tile and slot (warpgroup view): loads (cp.async for X, bulk copies for H) -> build G_L (dY * loss_scale * act')
// -> arrive g_ready -> for l = L-1..0: wait mma_done; read the data gradient out of TMEM, multiply by ReLU'(H_{l-1}),
// store it as fp16 over H_{l-1} (or, l = 0, un-scale and store dX) -> arrive g_ready.  Control lane: poll the slots'
// g_ready barriers; for the slot that is ready issue dW_l += G_l^T . H_{l-1} and dG_{l-1} = G_l . W_l, commit to the
// slot's mma_done barrier.

struct SlotBars {
    uint64_t h_full;      // bulk copies of the saved activations have landed (tx count)
    uint64_t g_ready;     // 128 arrivals: the slot's gradient tile (and, first round, its loads) are in shared memory
    uint64_t mma_done;    // tcgen05.commit of the round's MMAs
    int32_t tile;         // CTA-local tile number taken by the slot (written by the slot's thread 0)
    int32_t done;         // slot has run out of tiles
};
constexpr int kMaxSlots = 6;
struct SharedV2 {
    uint64_t mbar_w;
    uint32_t tmem_base;
    int32_t next_tile;    // CTA-local tile counter (slots take tiles blockIdx.x + k * gridDim.x in arrival order)
    uint32_t started;     // bit l: weight-gradient accumulator l was written (set by the control lane at the end)
    uint32_t pad;
    SlotBars slot[kMaxSlots];
};

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(tc05::smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *mbar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(tc05::smem_u32(mbar)) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
    asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(threads) : "memory");
}

// Accumulator row of this thread in 16-column blocks, two blocks per trip (32 live registers; with several tiles in
// flight per SM the deeper batching of for_each_acc_block is not needed, and registers are what limits the slot count).
template <class Fn>
__device__ __forceinline__ void for_each_acc_pair(uint32_t taddr, int nblk, Fn fn) {
    for (int cb0 = 0; cb0 < nblk; cb0 += 2) {
        uint32_t r0[16], r1[16];
        tc05::tmem_ld16(taddr + (uint32_t)cb0 * 16, r0);
        if (cb0 + 1 < nblk) tc05::tmem_ld16(taddr + (uint32_t)cb0 * 16 + 16, r1);
        tc05::tmem_ld_wait();
        fn(cb0, r0);
        if (cb0 + 1 < nblk) fn(cb0 + 1, r1);
    }
}

static size_t bwd_v2_slot_bytes(const MlpCfg &c) {
    return (size_t)kRows * c.in_pad * 2 + (size_t)(c.n_mm - 1) * kRows * c.width * 2 + (size_t)kRows * 16 * 2;
}
static size_t bwd_v2_smem(const MlpCfg &c, int nwg) {
    return 512 + ((c.w_bytes + 127u) & ~127u) + (size_t)nwg * bwd_v2_slot_bytes(c);
}

// wg_col[l]: TMEM column of layer l's weight-gradient accumulator (after the NWG data-gradient regions).
struct V2Cols { uint32_t wg_col[kMaxMM]; };

template <int NWG>
__global__ void __launch_bounds__(NWG * 128 + 32, 1)
mlp_bwd_v2_kernel(int64_t n, const int32_t *__restrict__ n_dev, const __half *__restrict__ X, int64_t ldx,
                  const uint8_t *__restrict__ blob, const __grid_constant__ MlpCfg cfg,
                  const __grid_constant__ BwdPlan plan, const __grid_constant__ V2Cols cols,
                  const float *__restrict__ dY, int64_t ldy, float loss_scale, float *__restrict__ dX, int64_t ldgx,
                  float *__restrict__ dW, int io_vec, BwdExtra ex) {
    extern __shared__ __align__(128) uint8_t smem[];
    SharedV2 *sh = reinterpret_cast<SharedV2 *>(smem);
    uint8_t *w_s = smem + 512;
    const size_t x_bytes = (size_t)kRows * cfg.in_pad * 2, h_bytes = (size_t)kRows * cfg.width * 2;
    const uint32_t hb = (uint32_t)(h_bytes * (cfg.n_mm - 1));
    const size_t slot_bytes = x_bytes + hb + (size_t)kRows * 32;
    uint8_t *slots = w_s + ((cfg.w_bytes + 127u) & ~127u);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wg = tid >> 7, wtid = tid & 127;
    constexpr int kCtlWarp = NWG * 4;
    constexpr uint32_t TMEM_COLS = 512;

    if (tid == 0) {
        tc05::mbar_init(&sh->mbar_w, 1);
        for (int s = 0; s < NWG; ++s) {
            tc05::mbar_init(&sh->slot[s].h_full, 1);
            tc05::mbar_init(&sh->slot[s].g_ready, 128);
            tc05::mbar_init(&sh->slot[s].mma_done, 1);
            sh->slot[s].done = 0;
        }
        sh->next_tile = 0;
        sh->started = 0;
        tc05::fence_mbar_init();
    }
    if (warp == kCtlWarp) tc05::tmem_alloc<TMEM_COLS>(&sh->tmem_base);
    tc05::fence_before_sync();
    __syncthreads();
    tc05::fence_after_sync();
    arc_pdl_sync();      // barriers and tensor memory are set up; from here on inputs of earlier kernels are read
    if (n_dev) { const int64_t nd = *n_dev; n = nd < n ? (nd > 0 ? nd : 0) : n; }
    const uint32_t tmem = sh->tmem_base;
    const int64_t n_tiles = (n + kRows - 1) / kRows;
    const float inv_scale = 1.0f / loss_scale;

    if (warp == kCtlWarp) {
        // ------------------------------------------------------------------ control warp: every MMA of the CTA
        if (lane == 0) {
            tc05::mbar_expect_tx(&sh->mbar_w, cfg.w_bytes);
            tc05::bulk_g2s(w_s, blob, cfg.w_bytes, &sh->mbar_w);
            tc05::mbar_wait(&sh->mbar_w, 0);
            const uint32_t w_addr = tc05::smem_u32(w_s);
            int layer[NWG];
            uint32_t gph[NWG];
            bool active[NWG];
            #pragma unroll
            for (int s = 0; s < NWG; ++s) { layer[s] = cfg.n_mm - 1; gph[s] = 0; active[s] = true; }
            uint32_t started = 0;                 // bit l: accumulator of layer l holds a partial sum already
            int n_active = NWG;
            while (n_active > 0) {
                #pragma unroll
                for (int s = 0; s < NWG; ++s) {
                    if (!active[s] || !tc05::mbar_test_wait(&sh->slot[s].g_ready, gph[s])) continue;
                    gph[s] ^= 1u;
                    if (*reinterpret_cast<volatile int32_t *>(&sh->slot[s].done)) { active[s] = false; --n_active; continue; }
                    tc05::fence_after_sync();
                    const int l = layer[s];
                    const uint32_t x_addr = tc05::smem_u32(slots + (size_t)s * slot_bytes);
                    const uint32_t h_addr = x_addr + (uint32_t)x_bytes;
                    const uint32_t g_addr = h_addr + hb;
                    const uint32_t g_tile = (l == cfg.n_mm - 1) ? g_addr : h_addr + (uint32_t)(l * h_bytes);
                    const uint32_t in_tile = (l == 0) ? x_addr : h_addr + (uint32_t)((l - 1) * h_bytes);
                    const bool acc = (started >> l) & 1u;
                    if (plan.wg_t[l]) issue_wgrad_mm(tmem + cols.wg_col[l], in_tile, g_tile, plan.wg_m[l], plan.wg_n[l], acc);
                    else issue_wgrad_mm(tmem + cols.wg_col[l], g_tile, in_tile, plan.wg_m[l], plan.wg_n[l], acc);
                    started |= 1u << l;
                    if (l > 0 || dX != nullptr)
                        issue_dgrad_mm(tmem + (uint32_t)s * 64, g_tile, w_addr + cfg.w_off[l], cfg.K[l], cfg.N[l]);
                    tc05::mma_commit(&sh->slot[s].mma_done);
                    layer[s] = (l == 0) ? cfg.n_mm - 1 : l - 1;
                }
            }
            sh->started = started;                // tells the flush which accumulators were ever written
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------------ tile slot `wg`
        SlotBars *sb = &sh->slot[wg];
        uint8_t *x_s = slots + (size_t)wg * slot_bytes;
        uint8_t *h_s = x_s + x_bytes;
        uint8_t *g_s = h_s + hb;
        const uint32_t act_tmem = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)wg * 64;
        const int ncg = cfg.in_pad >> 3;
        uint32_t mph = 0, hph = 0;
        while (true) {
            if (wtid == 0) sb->tile = atomicAdd(&sh->next_tile, 1);
            named_bar_sync(1 + wg, 128);
            const int64_t tile = (int64_t)blockIdx.x + (int64_t)sb->tile * gridDim.x;
            if (tile >= n_tiles) {
                if (wtid == 0) sb->done = 1;
                mbar_arrive(&sb->g_ready);
                break;
            }
            const int64_t grow = tile * kRows + wtid;
            const bool valid = grow < n;
            // ---- loads of this tile: X row (cp.async), saved activations (bulk copies)
            for (int cg = 0; cg < ncg; ++cg) {
                uint4 *dst = tile_word(x_s, wtid, cg);
                if (valid) cp_async16(dst, X + grow * ldx + cg * 8);
                else *dst = make_uint4(0u, 0u, 0u, 0u);
            }
            if (wtid == 0) {
                tc05::mbar_expect_tx(&sb->h_full, hb);
                for (int l = 0; l + 1 < cfg.n_mm; ++l)
                    tc05::bulk_g2s(h_s + (size_t)l * h_bytes, ex.act_saved + ((size_t)tile * (cfg.n_mm - 1) + l) * h_bytes,
                                   (uint32_t)h_bytes, &sb->h_full);
            }
            // ---- gradient at the network output (16 padded columns) for this thread's row
            {
                float g[16];
                if ((io_vec & 4) && valid && cfg.out_dim == 16) {
                    #pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float4 q = __ldg(reinterpret_cast<const float4 *>(dY + grow * ldy) + k);
                        g[4 * k] = q.x; g[4 * k + 1] = q.y; g[4 * k + 2] = q.z; g[4 * k + 3] = q.w;
                    }
                } else {
                    #pragma unroll
                    for (int k = 0; k < 16; ++k) g[k] = (valid && k < cfg.out_dim) ? __ldg(dY + grow * ldy + k) : 0.0f;
                }
                if (ex.g0 && valid) {
                    const float x0 = ex.x0_is_half ? __half2float(__ldg(reinterpret_cast<const __half *>(ex.x0) + grow * ex.ldx0))
                                                   : __ldg(reinterpret_cast<const float *>(ex.x0) + grow * ex.ldx0);
                    g[0] += __ldg(ex.g0 + grow) * expf(fminf(fmaxf(x0, -ex.clip), ex.clip));
                }
                #pragma unroll
                for (int k = 0; k < 16; ++k) g[k] *= loss_scale;
                if (cfg.out_act != ACT_NONE) {
                    float y[16];
                    #pragma unroll
                    for (int k = 0; k < 16; ++k) {
                        y[k] = 0.f;
                        if (valid && k < cfg.out_dim)
                            y[k] = ex.y_is_half ? __half2float(__ldg(reinterpret_cast<const __half *>(ex.y_saved) + grow * ex.ldys + k))
                                                : __ldg(reinterpret_cast<const float *>(ex.y_saved) + grow * ex.ldys + k);
                    }
                    act16_bwd_mul(cfg.out_act, g, y);
                }
                uint4 w0, w1;
                w0.x = tc05::pack_half2(g[0], g[1]);   w0.y = tc05::pack_half2(g[2], g[3]);
                w0.z = tc05::pack_half2(g[4], g[5]);   w0.w = tc05::pack_half2(g[6], g[7]);
                w1.x = tc05::pack_half2(g[8], g[9]);   w1.y = tc05::pack_half2(g[10], g[11]);
                w1.z = tc05::pack_half2(g[12], g[13]); w1.w = tc05::pack_half2(g[14], g[15]);
                *tile_word(g_s, wtid, 0) = w0;
                *tile_word(g_s, wtid, 1) = w1;
            }
            cp_async_wait_all();
            if (hb) tc05::mbar_wait(&sb->h_full, hph);
            hph ^= 1u;
            tc05::fence_proxy_async();
            tc05::fence_before_sync();          // TMEM reads of the previous tile precede the arrival as well
            mbar_arrive(&sb->g_ready);

            for (int l = cfg.n_mm - 1; l >= 0; --l) {
                tc05::mbar_wait(&sb->mma_done, mph);
                mph ^= 1u;
                tc05::fence_after_sync();
                if (l > 0) {
                    // dG through ReLU of layer l-1, written over that layer's activations (this thread's row)
                    uint8_t *hp = h_s + (size_t)(l - 1) * h_bytes;
                    for_each_acc_pair(act_tmem, cfg.K[l] / 16, [&](int cb, const uint32_t (&rr)[16]) {
                        uint4 *p0 = tile_word(hp, wtid, 2 * cb), *p1 = tile_word(hp, wtid, 2 * cb + 1);
                        const uint4 a = *p0, b = *p1;
                        const uint32_t hw[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
                        uint32_t o[8];
                        #pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const float2 h2 = __half22float2(*reinterpret_cast<const __half2 *>(&hw[k]));
                            o[k] = tc05::pack_half2(h2.x > 0.f ? __uint_as_float(rr[2 * k]) : 0.f,
                                                    h2.y > 0.f ? __uint_as_float(rr[2 * k + 1]) : 0.f);
                        }
                        *p0 = make_uint4(o[0], o[1], o[2], o[3]);
                        *p1 = make_uint4(o[4], o[5], o[6], o[7]);
                    });
                    tc05::fence_proxy_async();
                    tc05::fence_before_sync();
                    mbar_arrive(&sb->g_ready);
                } else if (dX != nullptr) {
                    for_each_acc_pair(act_tmem, cfg.K[0] / 16, [&](int cb, const uint32_t (&rr)[16]) {
                        if (!valid) return;
                        float v[16];
                        #pragma unroll
                        for (int k = 0; k < 16; ++k) v[k] = __uint_as_float(rr[k]) * inv_scale;
                        store_out16<float>(dX + grow * ldgx, v, cb * 16, cfg.in_dim, (io_vec & 2) != 0);
                    });
                }
            }
        }
    }

    // ---- flush the weight-gradient accumulators of this CTA (layer l by warpgroup l % NWG)
    tc05::fence_before_sync();
    __syncthreads();
    tc05::fence_after_sync();
    const uint32_t started = sh->started;
    if (warp != kCtlWarp) {
        const uint32_t my_tmem = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        const int w4 = warp & 3;
        for (int l = 0; l < cfg.n_mm; ++l) {
            if ((l % NWG) != wg || !((started >> l) & 1u)) continue;
            const int M = plan.wg_m[l], Nn = plan.wg_n[l];
            const int row = (M == 128) ? wtid : (w4 * 16 + lane);
            const bool owner = (M == 128) || (lane < 16);
            float *dst = dW + cfg.p_off[l];
            for (int cb = 0; cb < Nn / 16; ++cb) {
                uint32_t r[16];
                tc05::tmem_ld16(my_tmem + cols.wg_col[l] + (uint32_t)cb * 16, r);
                tc05::tmem_ld_wait();
                if (!owner) continue;
                if (!plan.wg_t[l]) {
                    float4 *d4 = reinterpret_cast<float4 *>(dst + (size_t)row * cfg.K[l] + cb * 16);
                    #pragma unroll
                    for (int k = 0; k < 4; ++k)
                        atomicAdd(d4 + k, make_float4(__uint_as_float(r[4 * k]) * inv_scale, __uint_as_float(r[4 * k + 1]) * inv_scale,
                                                      __uint_as_float(r[4 * k + 2]) * inv_scale, __uint_as_float(r[4 * k + 3]) * inv_scale));
                } else {
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
    if (warp == kCtlWarp) tc05::tmem_dealloc<TMEM_COLS>(tmem);
}

// Host side: can the v2 kernel take this call?  (fp16 X on the vector path without ones-padding, width 64, ReLU hidden
// layers, linear or sigmoid output of at most 16 channels, saved activations, fp32 dX or none)
static bool bwd_v2_applies(const MlpCfg &c, int x_is_half, int dx_is_half, const void *dX, int io_vec, const BwdExtra &ex) {
    if (getenv("ARCB200_MLP_BWD_V1")) return false;
    return x_is_half && (io_vec & 1) && c.in_dim == c.in_pad && c.width == 64 && c.hidden_act == ACT_RELU &&
           (c.out_act == ACT_NONE || c.out_act == ACT_SIGMOID) && c.out_pad == 16 && ex.act_saved != nullptr &&
           (!dX || (!dx_is_half && (io_vec & 2))) && c.n_mm >= 2;
}

template <int NWG>
static int launch_bwd_v2(const MlpCfg &c, const BwdPlan &plan, int64_t n, const int32_t *n_dev, const void *X, int64_t ldx,
                         const void *blob, const float *dY, int64_t ldy, float loss_scale, void *dX, int64_t ldgx, float *dW,
                         int io_vec, const BwdExtra &ex, cudaStream_t st) {
    V2Cols cols;
    uint32_t col = NWG * 64;
    for (int l = 0; l < kMaxMM; ++l) {
        cols.wg_col[l] = col;
        if (l < c.n_mm) col += (uint32_t)plan.wg_n[l];
    }
    if (col > 512) return 1;                                   // does not fit: caller falls back
    const size_t smem = bwd_v2_smem(c, NWG);
    if (smem > 227 * 1024) return 1;
    static thread_local size_t smem_set = 0;
    if (smem_set < smem) {
        cudaError_t e = cudaFuncSetAttribute(mlp_bwd_v2_kernel<NWG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { arcb200_set_error("cudaFuncSetAttribute -> %s", cudaGetErrorString(e)); return ARCB200_ERR_CUDA; }
        smem_set = smem;
    }
    const int64_t tiles = (n + kRows - 1) / kRows;
    const int64_t sms = arc_sm_count();
    const unsigned grid = (unsigned)(tiles < sms ? tiles : sms);
    arc_launch(mlp_bwd_v2_kernel<NWG>, grid, NWG * 128 + 32, smem, st, n, n_dev, (const __half *)X, ldx, (const uint8_t *)blob, c, plan,
                                                              cols, dY, ldy, loss_scale, (float *)dX, ldgx, dW, io_vec, ex);
    return 0;
}
