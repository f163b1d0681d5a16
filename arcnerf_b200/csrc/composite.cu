// Alpha compositing along rays (volume-rendering quadrature) as a segmented per-ray transmittance scan,
// forward and backward, plus NeuS's SDF->alpha map and inverse-CDF resampling.
//
// Semantics: arcnerf/render/ray_helper.py:476-620 (ray_marching + alpha_to_weights):
//   delta_i = t_{i+1} - t_i, |delta| < 1e-5 -> 0;  alpha_i = 1 - exp(-relu(sigma_i + noise_i) * delta_i)
//   T_i = prod_{k<i} (1 - alpha_k + 1e-10);  w_i = alpha_i * T_i
//   rgb = sum w_i c_i (+ T_last * bkg | + 1 - sum w_i);  depth = sum w_i t_i;  mask = sum w_i
// where, without add_inf_z, the last sample of a ray only closes the last interval (it carries no weight),
// and T_last is the reference's trans_shift[:, -1], i.e. the transmittance BEFORE the last weighted entry.
//
// Layout: a ray is a segment [start, start+n) of flat per-sample arrays.  Dense (R, P) tensors are the
// special case start = r*P, n = P; the packed layout produced by arcb200_march_packed passes its offsets.
// For packed rays the kernel reproduces what the reference computes on the padded (R, P_max) matrix that
// FgModel builds from the same samples (fg_model.py:311-316: the tail repeats the last sample, so padded
// intervals have delta = 0 and weight 0): only a ray that owns P_max samples sees its last interval left out
// of T_last.
//
// B200 design: one warp per ray, 32 intervals per step, transmittance by a warp-shuffle product scan with a
// running carry; per-ray sums by shuffle reductions.  The backward walks the ray's 32-wide chunks in reverse
// with a suffix-sum scan (no prefix-subtraction cancellation), re-deriving T from per-chunk carries kept in
// registers.  HBM-bound: 20 B read per sample forward (sigma, rgb, t), 20 B read + 16 B written backward.
#include "common.cuh"

namespace {

constexpr int kMaxChunks = 128;   // 128 chunks x 32 = 4096 samples per ray at most

struct Seg { int64_t start; int32_t n; };

__device__ __forceinline__ Seg segment(uint32_t ray, const int32_t *__restrict__ offsets, int32_t dense_p, int64_t capacity) {
    Seg s;
    if (offsets) {
        s.start = offsets[ray];
        int64_t end = offsets[ray + 1];
        if (end > capacity) end = capacity;     // samples beyond the packed capacity were dropped by the sampler
        s.n = end > s.start ? (int32_t)(end - s.start) : 0;
    } else {
        s.start = (int64_t)ray * dense_p;
        s.n = dense_p;
    }
    return s;
}

// What one lane needs for one weighted entry.  Fetched one 32-wide chunk AHEAD of its use: the loads of chunk k+1
// are in flight while chunk k goes through the shuffle scan, so a long ray pays the memory latency once, not per chunk.
struct Item {
    float geo, t0, t1, nz, r, g, b, gw;
};

template <bool IS_ALPHA, bool WITH_RAD>
__device__ __forceinline__ Item fetch_item(const float *__restrict__ geo, const float *__restrict__ t,
                                           const float *__restrict__ noise, const float *__restrict__ rad,
                                           const float *__restrict__ g_w, int64_t start, int32_t i, int32_t n, int32_t m) {
    Item it;
    it.geo = 0.f; it.t0 = 0.f; it.t1 = 0.f; it.nz = 0.f; it.r = 0.f; it.g = 0.f; it.b = 0.f; it.gw = 0.f;
    if (i >= 0 && i < m) {
        const int64_t s = start + i;
        it.geo = geo[s];
        it.t0 = t[s];
        if (!IS_ALPHA) {
            it.t1 = (i < n - 1) ? t[s + 1] : 0.f;
            if (noise) it.nz = noise[s];
        }
        if (WITH_RAD && rad) { it.r = rad[3 * s]; it.g = rad[3 * s + 1]; it.b = rad[3 * s + 2]; }
        if (g_w) it.gw = g_w[s];
    }
    return it;
}

template <bool IS_ALPHA>
__device__ __forceinline__ float item_alpha(const Item &it, int32_t i, int32_t n, float &delta, float &dens) {
    if (IS_ALPHA) { delta = 0.f; dens = 0.f; return it.geo; }
    delta = (i < n - 1) ? (it.t1 - it.t0) : 1e10f;
    if (fabsf(delta) < 1e-5f) delta = 0.0f;
    dens = fmaxf(it.geo + it.nz, 0.0f);
    return 1.0f - expf(-dens * delta);
}

__device__ __forceinline__ float warp_prod_inclusive(float f, uint32_t lane) {
    #pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
        const float up = __shfl_up_sync(ARC_FULL_MASK, f, s);
        if ((int)lane >= s) f *= up;
    }
    return f;
}
__device__ __forceinline__ float warp_sum(float v) {
    #pragma unroll
    for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(ARC_FULL_MASK, v, s);
    return v;
}

// ------------------------------------------------------------------------------------------------ forward
template <bool IS_ALPHA>
__global__ void __launch_bounds__(256)
composite_fwd_kernel(uint32_t n_rays, const int32_t *__restrict__ offsets, int32_t dense_p, int64_t capacity,
                     const int32_t *__restrict__ pmax_dev, const float *__restrict__ geo,
                     const float *__restrict__ rad, const float *__restrict__ t, const float *__restrict__ noise,
                     int add_inf_z, int white_bkg, const float *__restrict__ bkg, int bkg_per_ray,
                     float *__restrict__ rgb, float *__restrict__ depth, float *__restrict__ acc_out,
                     float *__restrict__ w_out, float *__restrict__ trans_out, float *__restrict__ alpha_out,
                     float *__restrict__ t_last_out) {
    arc_pdl_sync();
    const uint32_t ray = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31u;
    if (ray >= n_rays) return;
    const Seg sg = segment(ray, offsets, dense_p, capacity);
    const int32_t n = sg.n;
    const int32_t m = (IS_ALPHA || add_inf_z) ? n : (n > 0 ? n - 1 : 0);   // weighted entries
    const int32_t pmax = offsets ? (pmax_dev ? *pmax_dev : -1) : dense_p;
    // entries whose factor enters T_last (see header): all but the last one when the ray fills the padded row
    const int32_t k_bkg = (n == pmax || IS_ALPHA || add_inf_z || !offsets) ? (m > 0 ? m - 1 : 0) : m;

    float carry = 1.0f, t_bkg = 1.0f;
    float ar = 0.f, ag = 0.f, ab = 0.f, ad = 0.f, aw = 0.f;
    Item cur = fetch_item<IS_ALPHA, true>(geo, t, noise, rad, nullptr, sg.start, (int32_t)lane, n, m);
    for (int32_t base = 0; base < m; base += 32) {
        const int32_t i = base + (int32_t)lane;
        const bool on = i < m;
        const int64_t s = sg.start + i;
        const Item nxt = fetch_item<IS_ALPHA, true>(geo, t, noise, rad, nullptr, sg.start, i + 32, n, m);
        float delta, dens;
        const float alpha = on ? item_alpha<IS_ALPHA>(cur, i, n, delta, dens) : 0.0f;
        const float f = on ? (1.0f - alpha + 1e-10f) : 1.0f;
        const float incl = warp_prod_inclusive(f, lane);
        float excl = __shfl_up_sync(ARC_FULL_MASK, incl, 1);
        if (lane == 0) excl = 1.0f;
        const float T = carry * excl;
        const float w = alpha * T;
        if (on) {
            if (rad) { ar = fmaf(w, cur.r, ar); ag = fmaf(w, cur.g, ag); ab = fmaf(w, cur.b, ab); }
            ad = fmaf(w, cur.t0, ad);
            aw += w;
            if (w_out) w_out[s] = w;
            if (trans_out) trans_out[s] = T;
            if (alpha_out) alpha_out[s] = alpha;
        }
        // transmittance in front of entry k_bkg (it may be the one-past-the-end entry)
        if (k_bkg >= base && k_bkg < base + 32) t_bkg = __shfl_sync(ARC_FULL_MASK, T, k_bkg - base);
        carry *= __shfl_sync(ARC_FULL_MASK, incl, 31);
        if (k_bkg == base + 32) t_bkg = carry;
        cur = nxt;
    }
    // the unweighted closing sample(s) of the segment
    for (int32_t i = m + (int32_t)lane; i < n; i += 32) {
        const int64_t s = sg.start + i;
        if (w_out) w_out[s] = 0.f;
        if (trans_out) trans_out[s] = 0.f;
        if (alpha_out) alpha_out[s] = 0.f;
    }
    ar = warp_sum(ar); ag = warp_sum(ag); ab = warp_sum(ab); ad = warp_sum(ad); aw = warp_sum(aw);
    if (lane == 0) {
        if (rgb) {
            if (bkg) {
                const float *b = bkg + (bkg_per_ray ? 3 * (size_t)ray : 0);
                ar += t_bkg * b[0]; ag += t_bkg * b[1]; ab += t_bkg * b[2];
            } else if (white_bkg) {
                ar += 1.0f - aw; ag += 1.0f - aw; ab += 1.0f - aw;
            }
            rgb[3 * (size_t)ray] = ar; rgb[3 * (size_t)ray + 1] = ag; rgb[3 * (size_t)ray + 2] = ab;
        }
        if (depth) depth[ray] = ad;
        if (acc_out) acc_out[ray] = aw;
        if (t_last_out) t_last_out[ray] = t_bkg;
    }
}

// ------------------------------------------------------------------------------------------------ backward
template <bool IS_ALPHA>
__global__ void __launch_bounds__(256)
composite_bwd_kernel(uint32_t n_rays, const int32_t *__restrict__ offsets, int32_t dense_p, int64_t capacity,
                     const int32_t *__restrict__ pmax_dev, const float *__restrict__ geo,
                     const float *__restrict__ rad, const float *__restrict__ t, const float *__restrict__ noise,
                     int add_inf_z, int white_bkg, const float *__restrict__ bkg, int bkg_per_ray,
                     const float *__restrict__ g_rgb, const float *__restrict__ g_depth,
                     const float *__restrict__ g_acc, const float *__restrict__ g_w,
                     float *__restrict__ d_geo, float *__restrict__ d_rad) {
    arc_pdl_sync();
    const uint32_t ray = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31u;
    if (ray >= n_rays) return;
    const Seg sg = segment(ray, offsets, dense_p, capacity);
    const int32_t n = sg.n;
    const int32_t m = (IS_ALPHA || add_inf_z) ? n : (n > 0 ? n - 1 : 0);
    const int32_t pmax = offsets ? (pmax_dev ? *pmax_dev : -1) : dense_p;
    const int32_t k_bkg = (n == pmax || IS_ALPHA || add_inf_z || !offsets) ? (m > 0 ? m - 1 : 0) : m;

    float gr = 0.f, gg = 0.f, gb = 0.f;
    if (g_rgb && rad) { gr = g_rgb[3 * (size_t)ray]; gg = g_rgb[3 * (size_t)ray + 1]; gb = g_rgb[3 * (size_t)ray + 2]; }
    const float gd = g_depth ? g_depth[ray] : 0.f;
    float ga = g_acc ? g_acc[ray] : 0.f;
    float v_bkg = 0.f;
    if (g_rgb && rad) {
        if (bkg) {
            const float *b = bkg + (bkg_per_ray ? 3 * (size_t)ray : 0);
            v_bkg = gr * b[0] + gg * b[1] + gb * b[2];
        } else if (white_bkg) {
            ga -= (gr + gg + gb);     // rgb += 1 - sum(w)
        }
    }

    // pass 1 (forward): transmittance carry in front of every 32-wide chunk; lane c of register q keeps chunk 32q+c
    const int32_t n_chunks = (m + 31) >> 5;
    float carries[kMaxChunks / 32];
    #pragma unroll
    for (int q = 0; q < kMaxChunks / 32; ++q) carries[q] = 1.0f;
    float carry = 1.0f, t_bkg = 1.0f;
    Item cur = fetch_item<IS_ALPHA, false>(geo, t, noise, nullptr, nullptr, sg.start, (int32_t)lane, n, m);
    for (int32_t c = 0; c < n_chunks; ++c) {
        const int32_t base = c << 5;
        const int32_t i = base + (int32_t)lane;
        const bool on = i < m;
        const Item nxt = fetch_item<IS_ALPHA, false>(geo, t, noise, nullptr, nullptr, sg.start, i + 32, n, m);
        float delta, dens;
        const float alpha = on ? item_alpha<IS_ALPHA>(cur, i, n, delta, dens) : 0.0f;
        const float f = on ? (1.0f - alpha + 1e-10f) : 1.0f;
        const float incl = warp_prod_inclusive(f, lane);
        #pragma unroll
        for (int q = 0; q < kMaxChunks / 32; ++q)
            if ((c >> 5) == q && (int32_t)lane == (c & 31)) carries[q] = carry;
        if (k_bkg >= base && k_bkg < base + 32) {
            float excl = __shfl_up_sync(ARC_FULL_MASK, incl, 1);
            if (lane == 0) excl = 1.0f;
            t_bkg = __shfl_sync(ARC_FULL_MASK, carry * excl, k_bkg - base);
        }
        carry *= __shfl_sync(ARC_FULL_MASK, incl, 31);
        if (k_bkg == base + 32) t_bkg = carry;
        cur = nxt;
    }
    const float bkg_term = t_bkg * v_bkg;

    // pass 2 (reverse): suffix sums of w_j * v_j
    float suffix = 0.0f;
    cur = fetch_item<IS_ALPHA, true>(geo, t, noise, rad, g_w, sg.start, ((n_chunks - 1) << 5) + (int32_t)lane, n, m);
    for (int32_t c = n_chunks - 1; c >= 0; --c) {
        const int32_t base = c << 5;
        const int32_t i = base + (int32_t)lane;
        const bool on = i < m;
        const int64_t s = sg.start + i;
        const Item nxt = fetch_item<IS_ALPHA, true>(geo, t, noise, rad, g_w, sg.start, i - 32, n, m);
        float delta = 0.f, dens = 0.f;
        const float alpha = on ? item_alpha<IS_ALPHA>(cur, i, n, delta, dens) : 0.0f;
        const float f = on ? (1.0f - alpha + 1e-10f) : 1.0f;
        const float incl = warp_prod_inclusive(f, lane);
        float excl = __shfl_up_sync(ARC_FULL_MASK, incl, 1);
        if (lane == 0) excl = 1.0f;
        float c0 = 1.0f;
        #pragma unroll
        for (int q = 0; q < kMaxChunks / 32; ++q)
            if ((c >> 5) == q) c0 = __shfl_sync(ARC_FULL_MASK, carries[q], c & 31);
        const float T = c0 * excl;
        const float w = alpha * T;
        float v = 0.f;
        if (on) {
            if (rad) v = gr * cur.r + gg * cur.g + gb * cur.b;
            v += gd * cur.t0 + ga + cur.gw;
        }
        // exclusive suffix sum inside the chunk
        float wv = w * v, sfx = wv;
        #pragma unroll
        for (int st = 1; st < 32; st <<= 1) {
            const float dn = __shfl_down_sync(ARC_FULL_MASK, sfx, st);
            if ((int)lane + st < 32) sfx += dn;
        }
        const float later = suffix + (sfx - wv);
        if (on) {
            const float tail = later + (i < k_bkg ? bkg_term : 0.0f);
            const float d_alpha = T * v - tail / f;
            if (IS_ALPHA) d_geo[s] = d_alpha;
            else d_geo[s] = dens > 0.0f ? d_alpha * delta * (1.0f - alpha) : 0.0f;
            if (d_rad) { d_rad[3 * s] = w * gr; d_rad[3 * s + 1] = w * gg; d_rad[3 * s + 2] = w * gb; }
        }
        suffix += __shfl_sync(ARC_FULL_MASK, sfx, 0);
        cur = nxt;
    }
    for (int32_t i = m + (int32_t)lane; i < n; i += 32) {
        const int64_t s = sg.start + i;
        d_geo[s] = 0.f;
        if (d_rad) { d_rad[3 * s] = 0.f; d_rad[3 * s + 1] = 0.f; d_rad[3 * s + 2] = 0.f; }
    }
}


// ------------------------------------------------------------------------------------------------ fused training pass
// Forward compositing, the Huber image loss of the ray and the backward in ONE kernel (density input, radiance given):
// the step's composite_fwd + huber_loss + composite_bwd.  Pass 1 is the forward kernel's loop (same accumulation order,
// so rgb is bit-identical) and also records the chunk carries; the loss gradient of the ray's colour is formed in
// registers; pass 2 is the backward kernel's reverse loop.  Semantics: ray_marching (ray_helper.py:476-593) ->
// ImgLoss/HuberLoss (arcnerf/loss/img_loss.py:48-98) -> autograd.
__global__ void __launch_bounds__(256)
composite_train_kernel(uint32_t n_rays, const int32_t *__restrict__ offsets, int32_t dense_p, int64_t capacity,
                       const int32_t *__restrict__ pmax_dev, const float *__restrict__ geo, const float *__restrict__ rad,
                       const float *__restrict__ t, const float *__restrict__ noise, int add_inf_z, int white_bkg,
                       const float *__restrict__ bkg, int bkg_per_ray, const float *__restrict__ target, float delta_h,
                       float scale, float *__restrict__ rgb, float *__restrict__ g_rgb_out, float *__restrict__ loss_sum,
                       float *__restrict__ d_geo, float *__restrict__ d_rad, int32_t *__restrict__ step_counter) {
    arc_pdl_sync();
    if (step_counter && blockIdx.x == 0 && threadIdx.x == 0) *step_counter += 1;    // device-resident optimiser step count
    const uint32_t ray = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31u;
    __shared__ float block_loss[8];
    float my_loss = 0.0f;
    if (ray < n_rays) {
        const Seg sg = segment(ray, offsets, dense_p, capacity);
        const int32_t n = sg.n;
        const int32_t m = add_inf_z ? n : (n > 0 ? n - 1 : 0);
        const int32_t pmax = offsets ? (pmax_dev ? *pmax_dev : -1) : dense_p;
        const int32_t k_bkg = (n == pmax || add_inf_z || !offsets) ? (m > 0 ? m - 1 : 0) : m;

        // pass 1: forward
        const int32_t n_chunks = (m + 31) >> 5;
        float carries[kMaxChunks / 32];
        #pragma unroll
        for (int q = 0; q < kMaxChunks / 32; ++q) carries[q] = 1.0f;
        float carry = 1.0f, t_bkg = 1.0f;
        float ar = 0.f, ag = 0.f, ab = 0.f, aw = 0.f;
        // loads run three chunks ahead of their use: a long ray (hundreds of samples) is what the kernel's duration hangs on,
        // and with one chunk of look-ahead every trip would wait most of an L2 round trip
        Item cur = fetch_item<false, true>(geo, t, noise, rad, nullptr, sg.start, (int32_t)lane, n, m);
        Item n1 = fetch_item<false, true>(geo, t, noise, rad, nullptr, sg.start, (int32_t)lane + 32, n, m);
        Item n2 = fetch_item<false, true>(geo, t, noise, rad, nullptr, sg.start, (int32_t)lane + 64, n, m);
        for (int32_t c = 0; c < n_chunks; ++c) {
            const int32_t base = c << 5;
            const int32_t i = base + (int32_t)lane;
            const bool on = i < m;
            const Item n3 = fetch_item<false, true>(geo, t, noise, rad, nullptr, sg.start, i + 96, n, m);
            float delta, dens;
            const float alpha = on ? item_alpha<false>(cur, i, n, delta, dens) : 0.0f;
            const float f = on ? (1.0f - alpha + 1e-10f) : 1.0f;
            const float incl = warp_prod_inclusive(f, lane);
            float excl = __shfl_up_sync(ARC_FULL_MASK, incl, 1);
            if (lane == 0) excl = 1.0f;
            const float T = carry * excl;
            const float w = alpha * T;
            if (on) { ar = fmaf(w, cur.r, ar); ag = fmaf(w, cur.g, ag); ab = fmaf(w, cur.b, ab); aw += w; }
            #pragma unroll
            for (int q = 0; q < kMaxChunks / 32; ++q)
                if ((c >> 5) == q && (int32_t)lane == (c & 31)) carries[q] = carry;
            if (k_bkg >= base && k_bkg < base + 32) t_bkg = __shfl_sync(ARC_FULL_MASK, T, k_bkg - base);
            carry *= __shfl_sync(ARC_FULL_MASK, incl, 31);
            if (k_bkg == base + 32) t_bkg = carry;
            cur = n1; n1 = n2; n2 = n3;
        }
        ar = warp_sum(ar); ag = warp_sum(ag); ab = warp_sum(ab); aw = warp_sum(aw);
        float b0 = 0.f, b1 = 0.f, b2 = 0.f;
        if (bkg) {
            const float *b = bkg + (bkg_per_ray ? 3 * (size_t)ray : 0);
            b0 = b[0]; b1 = b[1]; b2 = b[2];
            ar += t_bkg * b0; ag += t_bkg * b1; ab += t_bkg * b2;
        } else if (white_bkg) {
            ar += 1.0f - aw; ag += 1.0f - aw; ab += 1.0f - aw;
        }
        // Huber loss of this ray's colour and its gradient (every lane holds the same values)
        float gr, gg, gb;
        {
            const float col[3] = {ar, ag, ab};
            float gcol[3];
            #pragma unroll
            for (int k = 0; k < 3; ++k) {
                const float d = col[k] - target[3 * (size_t)ray + k];
                const float a = fabsf(d);
                if (a < delta_h) { my_loss += 0.5f / delta_h * (a * a); gcol[k] = scale * d / delta_h; }
                else { my_loss += a - 0.5f * delta_h; gcol[k] = scale * (d > 0.f ? 1.f : -1.f); }
            }
            gr = gcol[0]; gg = gcol[1]; gb = gcol[2];
        }
        if (lane == 0) {
            if (rgb) { rgb[3 * (size_t)ray] = ar; rgb[3 * (size_t)ray + 1] = ag; rgb[3 * (size_t)ray + 2] = ab; }
            if (g_rgb_out) { g_rgb_out[3 * (size_t)ray] = gr; g_rgb_out[3 * (size_t)ray + 1] = gg; g_rgb_out[3 * (size_t)ray + 2] = gb; }
        }
        float ga = 0.f, v_bkg = 0.f;
        if (bkg) v_bkg = gr * b0 + gg * b1 + gb * b2;
        else if (white_bkg) ga -= (gr + gg + gb);
        const float bkg_term = t_bkg * v_bkg;

        // pass 2: reverse, suffix sums of w_j * v_j
        float suffix = 0.0f;
        const int32_t last_i = ((n_chunks - 1) << 5) + (int32_t)lane;
        cur = fetch_item<false, true>(geo, t, noise, rad, nullptr, sg.start, last_i, n, m);
        n1 = fetch_item<false, true>(geo, t, noise, rad, nullptr, sg.start, last_i - 32, n, m);
        n2 = fetch_item<false, true>(geo, t, noise, rad, nullptr, sg.start, last_i - 64, n, m);
        for (int32_t c = n_chunks - 1; c >= 0; --c) {
            const int32_t base = c << 5;
            const int32_t i = base + (int32_t)lane;
            const bool on = i < m;
            const int64_t s = sg.start + i;
            const Item n3 = fetch_item<false, true>(geo, t, noise, rad, nullptr, sg.start, i - 96, n, m);
            float delta = 0.f, dens = 0.f;
            const float alpha = on ? item_alpha<false>(cur, i, n, delta, dens) : 0.0f;
            const float f = on ? (1.0f - alpha + 1e-10f) : 1.0f;
            const float incl = warp_prod_inclusive(f, lane);
            float excl = __shfl_up_sync(ARC_FULL_MASK, incl, 1);
            if (lane == 0) excl = 1.0f;
            float c0 = 1.0f;
            #pragma unroll
            for (int q = 0; q < kMaxChunks / 32; ++q)
                if ((c >> 5) == q) c0 = __shfl_sync(ARC_FULL_MASK, carries[q], c & 31);
            const float T = c0 * excl;
            const float w = alpha * T;
            float v = 0.f;
            if (on) { v = gr * cur.r + gg * cur.g + gb * cur.b; v += 0.0f * cur.t0 + ga + 0.0f; }
            float wv = w * v, sfx = wv;
            #pragma unroll
            for (int st = 1; st < 32; st <<= 1) {
                const float dn = __shfl_down_sync(ARC_FULL_MASK, sfx, st);
                if ((int)lane + st < 32) sfx += dn;
            }
            const float later = suffix + (sfx - wv);
            if (on) {
                const float tail = later + (i < k_bkg ? bkg_term : 0.0f);
                const float d_alpha = T * v - tail / f;
                d_geo[s] = dens > 0.0f ? d_alpha * delta * (1.0f - alpha) : 0.0f;
                if (d_rad) { d_rad[3 * s] = w * gr; d_rad[3 * s + 1] = w * gg; d_rad[3 * s + 2] = w * gb; }
            }
            suffix += __shfl_sync(ARC_FULL_MASK, sfx, 0);
            cur = n1; n1 = n2; n2 = n3;
        }
        for (int32_t i = m + (int32_t)lane; i < n; i += 32) {
            const int64_t s = sg.start + i;
            d_geo[s] = 0.f;
            if (d_rad) { d_rad[3 * s] = 0.f; d_rad[3 * s + 1] = 0.f; d_rad[3 * s + 2] = 0.f; }
        }
    }
    // one atomic per CTA for the loss
    if (lane == 0) block_loss[threadIdx.x >> 5] = my_loss;
    __syncthreads();
    if (threadIdx.x == 0 && loss_sum) {
        float l = 0.f;
        for (int w = 0; w < 8; ++w) l += block_loss[w];
        if (l != 0.f) atomicAdd(loss_sum, l * scale);
    }
}

// NeuS: alpha from SDF at interval mid points (arcnerf/models/neus_model.py:242-265), dense (R, P) inputs with
// zvals (R, P+1):  a = (Phi(s*(f - slope*d/2)) - Phi(s*(f + slope*d/2)) + 1e-5) / (Phi(prev) + 1e-5), clipped.
__global__ void sdf_to_alpha_kernel(int64_t n, int32_t p, const float *__restrict__ sdf,
                                    const float *__restrict__ zvals, const float *__restrict__ slope,
                                    const float *__restrict__ s_dev, float s_host, int clip, float *__restrict__ alpha) {
    arc_pdl_sync();
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const float sc = s_dev ? __ldg(s_dev) : s_host;
    const int64_t r = k / p, c = k - r * p;
    const float *z = zvals + r * (p + 1) + c;
    const float dist = z[1] - z[0];
    const float half_step = slope[k] * dist * 0.5f;
    const float prev = 1.0f / (1.0f + expf(-(sdf[k] - half_step) * sc));
    const float next = 1.0f / (1.0f + expf(-(sdf[k] + half_step) * sc));
    float a = (prev - next + 1e-5f) / (prev + 1e-5f);
    if (clip) a = fminf(fmaxf(a, 0.0f), 1.0f);
    alpha[k] = a;
}

// backward of the above w.r.t. sdf, slope and the scale s (sum over all entries, atomically into d_s[0])
__global__ void sdf_to_alpha_bwd_kernel(int64_t n, int32_t p, const float *__restrict__ sdf,
                                        const float *__restrict__ zvals, const float *__restrict__ slope,
                                        const float *__restrict__ s_dev, float s_host, int clip,
                                        const float *__restrict__ g_alpha, float *__restrict__ d_sdf,
                                        float *__restrict__ d_slope, float *__restrict__ d_s) {
    arc_pdl_sync();
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    float ds = 0.f;
    if (k < n) {
        const float sc = s_dev ? __ldg(s_dev) : s_host;
        const int64_t r = k / p, c = k - r * p;
        const float *z = zvals + r * (p + 1) + c;
        const float dist = z[1] - z[0];
        const float hs = slope[k] * dist * 0.5f;
        const float xp = sdf[k] - hs, xn = sdf[k] + hs;
        const float prev = 1.0f / (1.0f + expf(-xp * sc)), next = 1.0f / (1.0f + expf(-xn * sc));
        const float den = prev + 1e-5f;
        const float a = (prev - next + 1e-5f) / den;
        float g = g_alpha[k];
        if (clip && (a < 0.0f || a > 1.0f)) g = 0.f;
        // a = 1 - (next) / den  + ...:  da/dprev = (next) / den^2 ... written out:
        const float da_dprev = (den - (prev - next + 1e-5f)) / (den * den);
        const float da_dnext = -1.0f / den;
        const float dprev = prev * (1.0f - prev), dnext = next * (1.0f - next);   // sigmoid'
        const float gp = g * da_dprev * dprev, gn = g * da_dnext * dnext;
        if (d_sdf) d_sdf[k] = (gp + gn) * sc;
        if (d_slope) d_slope[k] = (gn - gp) * sc * dist * 0.5f;
        ds = gp * xp + gn * xn;
    }
    if (d_s) {
        ds = warp_sum(ds);
        if ((threadIdx.x & 31) == 0 && ds != 0.f) atomicAdd(d_s, ds);
    }
}

// Inverse-CDF resampling (ray_helper.py:410-473): one thread per (ray, new sample).
// bins (R, nb); cdf (R, nb) (non-decreasing, cdf[0] = 0); u (R, ns) in [0,1]; out (R, ns) UNSORTED
// (the caller sorts, as the reference does).  searchsorted(right=True) + linear interpolation.
__global__ void sample_cdf_kernel(int64_t total, int32_t nb, int32_t ns, const float *__restrict__ bins,
                                  const float *__restrict__ cdf, const float *__restrict__ u, int u_per_ray,
                                  float eps, float *__restrict__ out) {
    arc_pdl_sync();
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= total) return;
    const int64_t r = k / ns;
    const int32_t j = (int32_t)(k - r * ns);
    const float uu = u_per_ray ? u[k] : u[j];
    const float *c = cdf + r * nb, *b = bins + r * nb;
    int32_t lo = 0, hi = nb;               // first index with c[idx] > uu
    while (lo < hi) {
        const int32_t mid = (lo + hi) >> 1;
        if (c[mid] > uu) hi = mid; else lo = mid + 1;
    }
    const int32_t below = min(max(lo - 1, 0), nb - 1), above = min(max(lo, 0), nb - 1);
    float denom = c[above] - c[below];
    if (denom < eps) denom = 1.0f;
    const float tt = (uu - c[below]) / denom;
    out[k] = b[below] + tt * (b[above] - b[below]);
}

}  // namespace

extern "C" {

// Forward compositing.  See the header comment for the layout.  geo = sigma (is_alpha == 0) or alpha.
//   offsets (n_rays+1) int32 or NULL (dense: dense_p samples per ray); capacity = length of the per-sample arrays
//   pmax_dev: device int32 holding the padded row length P_max of the equivalent dense call (packed only;
//             NULL: never treat a ray as row-filling)
//   rad (N,3) or NULL; noise (N) or NULL; bkg (n_rays,3) if bkg_per_ray else (1,3), or NULL
//   outputs: rgb (n_rays,3) (NULL allowed when rad is NULL), depth, acc (n_rays); per-sample w/trans/alpha or NULL
// Replaces ray_marching + alpha_to_weights (arcnerf/render/ray_helper.py:476-620).
//   arcb200_composite_fwd_ex: + t_last_out (n_rays) or NULL = the transmittance the background is weighted with, the
//   reference's trans_shift[:, -1] (what FullModel.blend_bkg_rgb multiplies the background model's colour and depth by)
int arcb200_composite_fwd_ex(const int32_t *offsets, int dense_p, int64_t capacity, const int32_t *pmax_dev,
                             const float *geo, int is_alpha, const float *rad, const float *t, const float *noise,
                             int add_inf_z, int white_bkg, const float *bkg, int bkg_per_ray, float *rgb, float *depth,
                             float *acc, float *w_out, float *trans_out, float *alpha_out, float *t_last_out, int64_t n_rays,
                             void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_rays < (1ll << 27), "composite_fwd: n_rays out of range");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(geo && t, "composite_fwd: null input");
    ARC_REQUIRE(offsets || dense_p > 0, "composite_fwd: need offsets or a positive dense_p");
    ARC_REQUIRE(!rad || rgb, "composite_fwd: radiance given but rgb output is null");
    const unsigned blocks = arc_blocks(n_rays * 32, 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (is_alpha)
        arc_launch(composite_fwd_kernel<true>, blocks, 256, 0, st, (uint32_t)n_rays, offsets, dense_p, capacity, pmax_dev, geo, rad,
                                                           t, noise, add_inf_z, white_bkg, bkg, bkg_per_ray, rgb, depth,
                                                           acc, w_out, trans_out, alpha_out, t_last_out);
    else
        arc_launch(composite_fwd_kernel<false>, blocks, 256, 0, st, (uint32_t)n_rays, offsets, dense_p, capacity, pmax_dev, geo, rad,
                                                            t, noise, add_inf_z, white_bkg, bkg, bkg_per_ray, rgb, depth,
                                                            acc, w_out, trans_out, alpha_out, t_last_out);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_composite_fwd(const int32_t *offsets, int dense_p, int64_t capacity, const int32_t *pmax_dev,
                          const float *geo, int is_alpha, const float *rad, const float *t, const float *noise,
                          int add_inf_z, int white_bkg, const float *bkg, int bkg_per_ray, float *rgb, float *depth,
                          float *acc, float *w_out, float *trans_out, float *alpha_out, int64_t n_rays,
                          void *stream) {
    return arcb200_composite_fwd_ex(offsets, dense_p, capacity, pmax_dev, geo, is_alpha, rad, t, noise, add_inf_z, white_bkg, bkg,
                                    bkg_per_ray, rgb, depth, acc, w_out, trans_out, alpha_out, nullptr, n_rays, stream);
}

// Backward: d_geo (N) and d_rad (N,3 or NULL) from g_rgb (n_rays,3), g_depth, g_acc (n_rays) and g_w (N), each
// optional.  Every entry of d_geo / d_rad inside the segments is written.  Rays longer than 4096 samples are
// refused.  (The reference differentiates ray_marching with autograd; no zvals gradient is produced there
// either: samples come from a no_grad sampler.)
int arcb200_composite_bwd(const int32_t *offsets, int dense_p, int64_t capacity, const int32_t *pmax_dev,
                          const float *geo, int is_alpha, const float *rad, const float *t, const float *noise,
                          int add_inf_z, int white_bkg, const float *bkg, int bkg_per_ray, const float *g_rgb,
                          const float *g_depth, const float *g_acc, const float *g_w, float *d_geo, float *d_rad,
                          int max_samples_per_ray, int64_t n_rays, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_rays < (1ll << 27), "composite_bwd: n_rays out of range");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(geo && t && d_geo, "composite_bwd: null input");
    ARC_REQUIRE(offsets || dense_p > 0, "composite_bwd: need offsets or a positive dense_p");
    if ((offsets ? max_samples_per_ray : dense_p) > kMaxChunks * 32) {
        arcb200_set_error("composite_bwd: at most %d samples per ray are supported", kMaxChunks * 32);
        return ARCB200_ERR_LIMIT;
    }
    const unsigned blocks = arc_blocks(n_rays * 32, 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (is_alpha)
        arc_launch(composite_bwd_kernel<true>, blocks, 256, 0, st, (uint32_t)n_rays, offsets, dense_p, capacity, pmax_dev, geo, rad,
                                                           t, noise, add_inf_z, white_bkg, bkg, bkg_per_ray, g_rgb,
                                                           g_depth, g_acc, g_w, d_geo, d_rad);
    else
        arc_launch(composite_bwd_kernel<false>, blocks, 256, 0, st, (uint32_t)n_rays, offsets, dense_p, capacity, pmax_dev, geo, rad,
                                                            t, noise, add_inf_z, white_bkg, bkg, bkg_per_ray, g_rgb,
                                                            g_depth, g_acc, g_w, d_geo, d_rad);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Fused training pass over the rays: rgb = composite(sigma, radiance, t) as arcb200_composite_fwd, loss_sum[0] += weight *
// mean(huber_delta(rgb - target)) as arcb200_huber_loss, and d_geo / d_rad as arcb200_composite_bwd fed with that loss
// gradient - one kernel instead of three (density input; rgb, g_rgb_out optional outputs).  step_counter (optional device
// int32) is incremented once: the optimiser step count that arcb200_adam_step reads through step_dev.
int arcb200_composite_huber_train(const int32_t *offsets, int dense_p, int64_t capacity, const int32_t *pmax_dev,
                                  const float *geo, const float *rad, const float *t, const float *noise, int add_inf_z,
                                  int white_bkg, const float *bkg, int bkg_per_ray, const float *target, float delta,
                                  float weight, float *rgb, float *g_rgb_out, float *loss_sum, float *d_geo, float *d_rad,
                                  int32_t *step_counter, int max_samples_per_ray, int64_t n_rays, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_rays < (1ll << 27), "composite_huber_train: n_rays out of range");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(geo && rad && t && target && d_geo && delta > 0.f, "composite_huber_train: null input or non-positive delta");
    ARC_REQUIRE(offsets || dense_p > 0, "composite_huber_train: need offsets or a positive dense_p");
    if ((offsets ? max_samples_per_ray : dense_p) > kMaxChunks * 32) {
        arcb200_set_error("composite_huber_train: at most %d samples per ray are supported", kMaxChunks * 32);
        return ARCB200_ERR_LIMIT;
    }
    arc_launch(composite_train_kernel, arc_blocks(n_rays * 32, 256), 256, 0, (cudaStream_t)stream, (uint32_t)n_rays, offsets,
               dense_p, capacity, pmax_dev, geo, rad, t, noise, add_inf_z, white_bkg, bkg, bkg_per_ray, target, delta,
               weight / (float)(n_rays * 3), rgb, g_rgb_out, loss_sum, d_geo, d_rad, step_counter);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Replaces sdf_to_alpha (arcnerf/models/neus_model.py:242-265).  sdf, slope, alpha (R,P); zvals (R,P+1);
// the scale s from device memory (s_dev != NULL) or s_host.
int arcb200_sdf_to_alpha(const float *sdf, const float *zvals, const float *slope, const float *s_dev, float s_host,
                         int clip, float *alpha, int64_t n_rays, int p, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && p >= 0, "sdf_to_alpha: negative size");
    const int64_t n = n_rays * p;
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(sdf && zvals && slope && alpha, "sdf_to_alpha: null pointer");
    arc_launch(sdf_to_alpha_kernel, arc_blocks(n, 256), 256, 0, (cudaStream_t)stream, n, p, sdf, zvals, slope, s_dev, s_host, clip, alpha);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// d_s (1 float, optional) is accumulated atomically; zero it before the call.
int arcb200_sdf_to_alpha_bwd(const float *sdf, const float *zvals, const float *slope, const float *s_dev,
                             float s_host, int clip, const float *g_alpha, float *d_sdf, float *d_slope, float *d_s,
                             int64_t n_rays, int p, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && p >= 0, "sdf_to_alpha_bwd: negative size");
    const int64_t n = n_rays * p;
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(sdf && zvals && slope && g_alpha, "sdf_to_alpha_bwd: null pointer");
    arc_launch(sdf_to_alpha_bwd_kernel, arc_blocks(n, 256), 256, 0, (cudaStream_t)stream, n, p, sdf, zvals, slope, s_dev, s_host, clip,
                                                                              g_alpha, d_sdf, d_slope, d_s);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Replaces the gather/interpolate core of sample_cdf (arcnerf/render/ray_helper.py:434-473); the caller builds the
// cdf (cumsum) and sorts the result, as the reference does with torch ops.
int arcb200_sample_cdf(const float *bins, const float *cdf, const float *u, int u_per_ray, float eps, float *out,
                       int64_t n_rays, int n_bins, int n_sample, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_bins > 0 && n_sample >= 0, "sample_cdf: invalid sizes");
    const int64_t total = n_rays * n_sample;
    if (total == 0) return ARCB200_OK;
    ARC_REQUIRE(bins && cdf && u && out, "sample_cdf: null pointer");
    arc_launch(sample_cdf_kernel, arc_blocks(total, 256), 256, 0, (cudaStream_t)stream, total, n_bins, n_sample, bins, cdf, u,
                                                                             u_per_ray, eps, out);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
