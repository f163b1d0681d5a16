/*
 * TEST INFRASTRUCTURE ONLY.  CPU oracle for the ray/occupancy part of the hot path.
 *
 * A plain-C restatement of the arithmetic of ArcNerf's `arcnerf/ops` CUDA kernels, used by
 * tests/ (parity checker), __graft_entry__.smoke() and bench.py's cpu_baseline leg.  The
 * product path (arcnerf_b200/) never links, loads or calls this file.
 *
 * Canonical floating-point semantics: every fp32 operation is individually rounded (no FMA
 * contraction) - the same semantics the reference's HOST_DEVICE functions have when built for
 * the host (oracle/_ref), which is what this file is pinned against (tests/test_oracle_sampler.py).
 * The CUDA kernels in arcnerf_b200/csrc/march.cu are compiled with -fmad=false to match.
 *
 * Reference lines followed (all under /root/reference/arcnerf/ops):
 *   include/pcg32.h:52-161                pcg32 seed / next_uint / next_float / advance
 *   include/volume_func.h:17-56           slab test
 *   include/volume_func.h:59-88           flat-index occupancy lookup (bool volume)
 *   include/volume_func.h:100-134         distance / advance to the next voxel
 *   include/volume_func.h:142-194         Morton code, bitfield occupancy lookup
 *   include/volume_func.h:201-297         cascaded (multi-volume) helpers
 *   src/volume_func/volume_func_kernel.cu:16-37,74-123,174-236,297-309
 *   src/bitfield_func/bitfield_func_kernel.cu:20-82,143-185,217-234,260-275,302-324,352-370
 *   src/multivol_func/multivol_func_kernel.cu:19-106,177-240,278-300
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include <string.h>

/* ------------------------------------------------------------------ pcg32 (pcg32.h:52-161) */
#define PCG_MULT 0x5851f42d4c957f2dULL

typedef struct { uint64_t state, inc; } rng_t;

static uint32_t rng_u32(rng_t *r) {
    uint64_t old = r->state;
    r->state = old * PCG_MULT + r->inc;
    uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((~rot + 1u) & 31));
}
static float rng_f32(rng_t *r) {
    union { uint32_t u; float f; } x;
    x.u = (rng_u32(r) >> 9) | 0x3f800000u;
    return x.f - 1.0f;
}
static void rng_jump(rng_t *r, uint64_t delta) {
    uint64_t cm = PCG_MULT, cp = r->inc, am = 1u, ap = 0u;
    while (delta > 0) {
        if (delta & 1) { am *= cm; ap = ap * cm + cp; }
        cp = (cm + 1) * cp;
        cm *= cm;
        delta >>= 1;
    }
    r->state = am * r->state + ap;
}
void orc_rng_init(uint64_t seed, uint64_t *state, uint64_t *inc) {
    rng_t r; r.state = 0u; r.inc = (1ull << 1u) | 1u;   /* initseq = 1 */
    rng_u32(&r); r.state += seed; rng_u32(&r);
    *state = r.state; *inc = r.inc;
}
void orc_rng_advance(uint64_t *state, uint64_t inc, int64_t delta) {
    rng_t r = { *state, inc }; rng_jump(&r, (uint64_t)delta); *state = r.state;
}
float orc_rng_next_float(uint64_t *state, uint64_t inc) {
    rng_t r = { *state, inc }; float f = rng_f32(&r); *state = r.state; return f;
}


/* ------------------------------------------------------------------ tiny pthread parallel-for
 * (the image's gcc has no libgomp).  orc_set_threads(n) selects the worker count used by the
 * per-ray / per-element loops below; 1 (default) runs inline. */
#include <pthread.h>
#include <stdatomic.h>
static int g_threads = 1;
void orc_set_threads(int n) { g_threads = n < 1 ? 1 : (n > 256 ? 256 : n); }
typedef void (*body_fn)(int i, void *ctx);
typedef struct { body_fn fn; void *ctx; int n, chunk; atomic_int next; } pf_t;
static void *pf_worker(void *arg) {
    pf_t *p = (pf_t *)arg;
    for (;;) {
        int b = atomic_fetch_add(&p->next, p->chunk);
        if (b >= p->n) break;
        int e = b + p->chunk < p->n ? b + p->chunk : p->n;
        for (int i = b; i < e; ++i) p->fn(i, p->ctx);
    }
    return NULL;
}
static void parallel_for(int n, int chunk, body_fn fn, void *ctx) {
    if (g_threads <= 1 || n <= chunk) { for (int i = 0; i < n; ++i) fn(i, ctx); return; }
    pf_t p; p.fn = fn; p.ctx = ctx; p.n = n; p.chunk = chunk; atomic_init(&p.next, 0);
    pthread_t th[256];
    int nt = g_threads;
    for (int k = 1; k < nt; ++k) pthread_create(&th[k], NULL, pf_worker, &p);
    pf_worker(&p);
    for (int k = 1; k < nt; ++k) pthread_join(th[k], NULL);
}

/* ------------------------------------------------------------------ small helpers */
typedef struct { float x, y, z; } v3;
static inline v3 ld3(const float *p) { v3 v = { p[0], p[1], p[2] }; return v; }
static inline float min3(float a, float b, float c) { float m = a < b ? a : b; return m < c ? m : c; }
static inline float max3(float a, float b, float c) { float m = a > b ? a : b; return m > c ? m : c; }
static inline float clampf(float v, float lo, float hi) { return v < lo ? lo : (hi < v ? hi : v); }
static inline v3 ray_at(v3 o, v3 d, float t) {
    v3 p = { o.x + d.x * t, o.y + d.y * t, o.z + d.z * t }; return p;
}
static inline int in_aabb(v3 p, v3 lo, v3 hi) {
    return p.x >= lo.x && p.y >= lo.y && p.z >= lo.z && p.x <= hi.x && p.y <= hi.y && p.z <= hi.z;
}

/* ------------------------------------------------------------------ slab test (volume_func.h:17-56) */
static void slab(v3 o, v3 d, v3 lo, v3 hi, float *tn, float *tf) {
    float tmin = (lo.x - o.x) / d.x, tmax = (hi.x - o.x) / d.x, s;
    if (tmin > tmax) { s = tmin; tmin = tmax; tmax = s; }
    float tymin = (lo.y - o.y) / d.y, tymax = (hi.y - o.y) / d.y;
    if (tymin > tymax) { s = tymin; tymin = tymax; tymax = s; }
    if (tmin > tymax || tymin > tmax) { *tn = -1.0f; *tf = -1.0f; return; }
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    float tzmin = (lo.z - o.z) / d.z, tzmax = (hi.z - o.z) / d.z;
    if (tzmin > tzmax) { s = tzmin; tzmin = tzmax; tzmax = s; }
    if (tmin > tzmax || tzmin > tmax) { *tn = -1.0f; *tf = -1.0f; return; }
    if (tzmin > tmin) tmin = tzmin;
    if (tzmax < tmax) tmax = tzmax;
    *tn = tmin; *tf = tmax;
}

/* volume_func_kernel.cu:74-123.  aabb is (N_v,2,3): min xyz then max xyz.  A ray whose near
 * root is not strictly positive (origin inside or behind) reports a miss. */
void orc_aabb_intersection(int n_rays, int n_v, const float *o, const float *d, const float *aabb,
                           float *near, float *far, float *pts, uint8_t *mask) {
    for (int i = 0; i < n_rays * n_v; ++i) {
        int r = i / n_v, v = i % n_v;
        v3 ro = ld3(o + 3 * r), rd = ld3(d + 3 * r);
        float tn, tf;
        slab(ro, rd, ld3(aabb + 6 * v), ld3(aabb + 6 * v + 3), &tn, &tf);
        if (tn > 0) { near[i] = tn; far[i] = tf; mask[i] = 1; }
        else { near[i] = 0.0f; far[i] = 0.0f; mask[i] = 0; }
        /* near*d + o, each product rounded before the add (Eigen: o + near*d) */
        float *p = pts + 6 * (size_t)i;
        p[0] = ro.x + near[i] * rd.x; p[1] = ro.y + near[i] * rd.y; p[2] = ro.z + near[i] * rd.z;
        p[3] = ro.x + far[i] * rd.x;  p[4] = ro.y + far[i] * rd.y;  p[5] = ro.z + far[i] * rd.z;
    }
}

/* ------------------------------------------------------------------ occupancy lookups */
/* volume_func.h:59-88: bool volume, flat index x*n*n + y*n + z */
static int occ_flat(v3 p, const uint8_t *vol, v3 lo, v3 hi, uint32_t n) {
    float fn = (float)n;
    float vx = (p.x - lo.x) / ((hi.x - lo.x) / fn);
    float vy = (p.y - lo.y) / ((hi.y - lo.y) / fn);
    float vz = (p.z - lo.z) / ((hi.z - lo.z) / fn);
    if (min3(vx, vy, vz) < 0 || max3(vx, vy, vz) >= fn) return 0;
    uint32_t x = (uint32_t)floorf(vx), y = (uint32_t)floorf(vy), z = (uint32_t)floorf(vz);
    return vol[x * (n * n) + y * n + z] != 0;
}

/* volume_func.h:142-157 */
static uint32_t spread3(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
static uint32_t morton(uint32_t x, uint32_t y, uint32_t z) { return spread3(x) | (spread3(y) << 1) | (spread3(z) << 2); }
/* volume_func.h:159-167 */
static uint32_t unspread3(uint32_t x) {
    x &= 0x49249249u;
    x = (x | (x >> 2)) & 0xc30c30c3u;
    x = (x | (x >> 4)) & 0x0f00f00fu;
    x = (x | (x >> 8)) & 0xff0000ffu;
    x = (x | (x >> 16)) & 0x0000ffffu;
    return x;
}
uint32_t orc_morton3D(uint32_t x, uint32_t y, uint32_t z) { return morton(x, y, z); }
uint32_t orc_morton3D_invert(uint32_t x) { return unspread3(x); }

/* volume_func.h:169-194: truncating int cast, clamp to the grid, Morton bit lookup */
static uint32_t cell_morton(v3 p, v3 lo, v3 hi, uint32_t n) {
    float fn = (float)n;
    int ix = (int)((p.x - lo.x) / ((hi.x - lo.x) / fn));
    int iy = (int)((p.y - lo.y) / ((hi.y - lo.y) / fn));
    int iz = (int)((p.z - lo.z) / ((hi.z - lo.z) / fn));
    float top = (float)((int)n - 1);
    return morton((uint32_t)clampf((float)ix, 0.f, top), (uint32_t)clampf((float)iy, 0.f, top),
                  (uint32_t)clampf((float)iz, 0.f, top));
}
static int occ_bit(v3 p, const uint8_t *bits, v3 lo, v3 hi, uint32_t n) {
    uint32_t idx = cell_morton(p, lo, hi, n);
    return (bits[idx / 8] & (1 << (idx % 8))) != 0;
}

/* volume_func_kernel.cu:16-37 */
void orc_check_pts_in_occ_voxel(int n_pts, const float *xyz, const uint8_t *vol, const float *aabb,
                                int n_grid, uint8_t *out) {
    v3 lo = ld3(aabb), hi = ld3(aabb + 3);
    for (int i = 0; i < n_pts; ++i) out[i] = (uint8_t)occ_flat(ld3(xyz + 3 * (size_t)i), vol, lo, hi, (uint32_t)n_grid);
}

/* ------------------------------------------------------------------ voxel skipping */
/* volume_func.h:100-118.  Note the reference scales by n_grid*pos whatever the box size. */
static float dist_next_voxel(v3 p, v3 d, v3 lo, v3 hi, uint32_t n) {
    float fn = (float)n;
    float cx = (lo.x + hi.x) / 2.0f, cy = (lo.y + hi.y) / 2.0f, cz = (lo.z + hi.z) / 2.0f;
    float hx = (hi.x - lo.x) / 2.0f, hy = (hi.y - lo.y) / 2.0f, hz = (hi.z - lo.z) / 2.0f;
    float ix = 1.0f / d.x, iy = 1.0f / d.y, iz = 1.0f / d.z;
    float px = fn * p.x, py = fn * p.y, pz = fn * p.z;
    float tx = (floorf((px + cx) + hx * copysignf(1.0f, d.x)) - px) * ix;
    float ty = (floorf((py + cy) + hy * copysignf(1.0f, d.y)) - py) * iy;
    float tz = (floorf((pz + cz) + hz * copysignf(1.0f, d.z)) - pz) * iz;
    return fmaxf(min3(tx, ty, tz) / fn, 0.0f);
}
/* volume_func.h:121-134 */
static float skip_fixed(float t, float dt, v3 p, v3 d, v3 lo, v3 hi, uint32_t n) {
    float target = t + dist_next_voxel(p, d, lo, hi, n);
    do { t += dt; } while (t < target);
    return t;
}

static void pad_tail(float *z, uint32_t j, uint32_t n_pts) {
    if (j > 0 && j < n_pts) { float last = z[j - 1]; for (; j < n_pts; ++j) z[j] = last; }
}

/* volume_func_kernel.cu:174-236 (bit == 0, bool volume) / bitfield_func_kernel.cu:20-82
 * (bit == 1, Morton bitfield).  zvals and mask must be zero-filled by the caller, exactly as
 * volume_func.py:100-101 does.  (rng_state, rng_inc) is the module generator as passed by value
 * to the launch; ray i advances its private copy by i*8 (32-bit product).
 * With zvals == NULL only the per-ray sample count is produced (packed-layout tests). */
typedef struct {
    int bit; const float *o, *d, *near, *far; v3 lo, hi; const uint8_t *field; uint32_t n_grid, n_pts;
    float dt, near_distance; uint64_t rng_state, rng_inc; float *zvals; uint8_t *mask; int32_t *counts;
} march_ctx;

static void march_ray(int i, void *vctx) {
    const march_ctx *c = (const march_ctx *)vctx;
    rng_t r = { c->rng_state, c->rng_inc };
    rng_jump(&r, (uint64_t)(int64_t)((uint32_t)i * 8u));
    v3 ro = ld3(c->o + 3 * (size_t)i), rd = ld3(c->d + 3 * (size_t)i);
    float t = fmaxf(c->near[i], c->near_distance);
    float t_end = c->far[i];
    t += c->dt * rng_f32(&r);
    uint32_t j = 0;
    float *z = c->zvals ? c->zvals + (size_t)i * c->n_pts : NULL;
    uint8_t *m = c->mask ? c->mask + (size_t)i * c->n_pts : NULL;
    while (t <= t_end && j < c->n_pts) {
        v3 p = ray_at(ro, rd, t);
        if (!in_aabb(p, c->lo, c->hi)) break;
        int occ = c->bit ? occ_bit(p, c->field, c->lo, c->hi, c->n_grid) : occ_flat(p, c->field, c->lo, c->hi, c->n_grid);
        if (occ) { if (z) { z[j] = t; m[j] = 1; } ++j; t += c->dt; }
        else t = skip_fixed(t, c->dt, p, rd, c->lo, c->hi, c->n_grid);
    }
    if (c->counts) c->counts[i] = (int32_t)j;
    if (z) pad_tail(z, j, c->n_pts);
}

void orc_sparse_volume_sampling(int bit, int n_rays, const float *o, const float *d, const float *near,
                                const float *far, const float *aabb, const uint8_t *field, int n_grid,
                                int n_pts, float dt, float near_distance, uint64_t rng_state,
                                uint64_t rng_inc, float *zvals, uint8_t *mask) {
    march_ctx c = { bit, o, d, near, far, ld3(aabb), ld3(aabb + 3), field, (uint32_t)n_grid, (uint32_t)n_pts,
                    dt, near_distance, rng_state, rng_inc, zvals, mask, NULL };
    parallel_for(n_rays, 64, march_ray, &c);
}

void orc_sparse_volume_count(int bit, int n_rays, const float *o, const float *d, const float *near,
                             const float *far, const float *aabb, const uint8_t *field, int n_grid,
                             int n_pts, float dt, float near_distance, uint64_t rng_state,
                             uint64_t rng_inc, int32_t *counts) {
    march_ctx c = { bit, o, d, near, far, ld3(aabb), ld3(aabb + 3), field, (uint32_t)n_grid, (uint32_t)n_pts,
                    dt, near_distance, rng_state, rng_inc, NULL, NULL, counts };
    parallel_for(n_rays, 64, march_ray, &c);
}

/* ------------------------------------------------------------------ cascaded volumes */
/* volume_func.h:201-223 */
static int mip_of(v3 p, v3 lo, v3 hi, uint32_t n_cascades) {
    float cx = (lo.x + hi.x) / 2.0f, cy = (lo.y + hi.y) / 2.0f, cz = (lo.z + hi.z) / 2.0f;
    float hx = (hi.x - lo.x) / 2.0f, hy = (hi.y - lo.y) / 2.0f, hz = (hi.z - lo.z) / 2.0f;
    int ex, ey, ez;
    frexpf(fabsf(p.x - cx) * (1.0f / hx), &ex);
    frexpf(fabsf(p.y - cy) * (1.0f / hy), &ey);
    frexpf(fabsf(p.z - cz) * (1.0f / hz), &ez);
    int e = ex > ey ? ex : ey; if (ez > e) e = ez;
    if (e < 0) e = 0;
    int top = (int)n_cascades - 1;
    return e < top ? e : top;
}
/* volume_func.h:226-263 */
static int occ_cascade(v3 p, const uint8_t *bits, uint32_t mip, v3 lo, v3 hi, uint32_t n, int inclusive) {
    float cx = (lo.x + hi.x) / 2.0f, cy = (lo.y + hi.y) / 2.0f, cz = (lo.z + hi.z) / 2.0f;
    float s = scalbnf(1.0f, -(int)mip);
    v3 q = { (p.x - cx) * s + cx, (p.y - cy) * s + cy, (p.z - cz) * s + cz };
    uint32_t idx = cell_morton(q, lo, hi, n);
    uint32_t level = inclusive ? mip : mip - 1;   /* uint32 wrap for mip==0 never reached: caller resets */
    uint32_t off = (n * n * n) * level;
    return (bits[idx / 8 + off / 8] & (1 << (idx % 8))) != 0;
}
/* volume_func.h:266-268 */
static float cone_dt(float t, float cone, float lo, float hi) { return clampf(t * cone, lo, hi); }
/* volume_func.h:272-297 */
static float skip_cone(float t, float cone, float smin, float smax, v3 p, v3 d, v3 lo, v3 hi, uint32_t n) {
    float target = t + dist_next_voxel(p, d, lo, hi, n);
    do { t += cone_dt(t, cone, smin, smax); } while (t < target);
    return t;
}

/* multivol_func_kernel.cu:19-106 */
void orc_sparse_sampling_in_multivol_bitfield(int n_rays, const float *o, const float *d, const float *near,
        const float *far, const float *min_aabb, const float *aabb, const uint8_t *bits, int n_grid,
        int n_cascade, int n_pts, float cone, float smin, float smax, float near_distance, int inclusive,
        uint64_t rng_state, uint64_t rng_inc, float *zvals, uint8_t *mask) {
    v3 lo = ld3(aabb), hi = ld3(aabb + 3), ilo = ld3(min_aabb), ihi = ld3(min_aabb + 3);
    for (int i = 0; i < n_rays; ++i) {
        rng_t r = { rng_state, rng_inc };
        rng_jump(&r, (uint64_t)(int64_t)((uint32_t)i * 8u));
        v3 ro = ld3(o + 3 * (size_t)i), rd = ld3(d + 3 * (size_t)i);
        float t = fmaxf(near[i], near_distance), t_end = far[i];
        t += cone_dt(t, cone, smin, smax) * rng_f32(&r);
        uint32_t j = 0;
        float *z = zvals + (size_t)i * n_pts; uint8_t *m = mask + (size_t)i * n_pts;
        while (t <= t_end && j < (uint32_t)n_pts) {
            v3 p = ray_at(ro, rd, t);
            if (!in_aabb(p, lo, hi)) break;
            float dt = cone_dt(t, cone, smin, smax);
            uint32_t mip = (uint32_t)mip_of(p, ilo, ihi, (uint32_t)n_cascade);
            if (mip == 0 && !inclusive) {
                /* the ray re-entered the inner box: drop everything sampled so far */
                for (uint32_t k = 0; k <= j; ++k) { z[k] = 0.0f; m[k] = 0; }
                j = 0;
                t = skip_fixed(t, dt, p, rd, ilo, ihi, (uint32_t)n_grid);
            } else if (occ_cascade(p, bits, mip, ilo, ihi, (uint32_t)n_grid, inclusive)) {
                z[j] = t; m[j] = 1; ++j; t += dt;
            } else {
                t = skip_cone(t, cone, smin, smax, p, rd, ilo, ihi, (uint32_t)n_grid);
            }
        }
        pad_tail(z, j, (uint32_t)n_pts);
    }
}

/* ------------------------------------------------------------------ occupancy-grid maintenance */
/* volume_func_kernel.cu:297-309: max of non-negative floats per group through their bit patterns */
void orc_tensor_reduce_max(int n, const float *full, const int64_t *group, int n_group, float *uni) {
    (void)n_group;
    for (int i = 0; i < n; ++i) {
        uint32_t a, b; memcpy(&a, &full[i], 4); memcpy(&b, &uni[group[i]], 4);
        if (a > b) memcpy(&uni[group[i]], &a, 4);
    }
}

/* bitfield_func_kernel.cu:143-185 */
void orc_generate_grid_samples(int n_elements, uint64_t rng_state, uint64_t rng_inc, uint32_t step,
                               const float *grid, float *pos, uint32_t *indices, uint32_t n_grid, float thresh) {
    uint32_t n3 = n_grid * n_grid * n_grid;
    for (int ii = 0; ii < n_elements; ++ii) {
        uint32_t i = (uint32_t)ii;
        rng_t r = { rng_state, rng_inc };
        rng_jump(&r, (uint64_t)(int64_t)(i * 4u));
        uint32_t idx = 0;
        for (uint32_t j = 0; j < 10; ++j) {
            idx = ((i + step * (uint32_t)n_elements) * 56924617u + j * 19349663u + 96925573u) % n3;
            if (grid[idx] > thresh) break;
        }
        uint32_t c = idx % n3;
        float x = (float)unspread3(c), y = (float)unspread3(c >> 1), z = (float)unspread3(c >> 2);
        float u0 = rng_f32(&r), u1 = rng_f32(&r), u2 = rng_f32(&r);
        float fn = (float)n_grid;
        pos[3 * (size_t)i] = (x + u0) / fn; pos[3 * (size_t)i + 1] = (y + u1) / fn; pos[3 * (size_t)i + 2] = (z + u2) / fn;
        indices[i] = idx;
    }
}

/* multivol_func_kernel.cu:177-240 */
void orc_generate_grid_samples_multivol(int n_elements, const float *aabb, uint64_t rng_state, uint64_t rng_inc,
        uint32_t step, const float *grid, float *pos, uint32_t *indices, uint32_t n_cascades, uint32_t n_grid,
        float thresh, int inclusive) {
    uint32_t n3 = n_grid * n_grid * n_grid;
    v3 lo = ld3(aabb), hi = ld3(aabb + 3);
    for (int ii = 0; ii < n_elements; ++ii) {
        uint32_t i = (uint32_t)ii;
        rng_t r = { rng_state, rng_inc };
        rng_jump(&r, (uint64_t)(int64_t)(i * 4u));
        uint32_t level = 0;
        if (inclusive) level = (uint32_t)(rng_f32(&r) * n_cascades) % n_cascades;
        else while (level == 0) level = (uint32_t)(rng_f32(&r) * n_cascades) % n_cascades;
        uint32_t idx = 0;
        for (uint32_t j = 0; j < 10; ++j) {
            idx = ((i + step * (uint32_t)n_elements) * 56924617u + j * 19349663u + 96925573u) % n3;
            idx += (inclusive ? level : level - 1) * n3;
            if (grid[idx] > thresh) break;
        }
        uint32_t c = idx % n3;
        float x = (float)unspread3(c), y = (float)unspread3(c >> 1), z = (float)unspread3(c >> 2);
        float u0 = rng_f32(&r), u1 = rng_f32(&r), u2 = rng_f32(&r);
        float fn = (float)n_grid, s = scalbnf(1.0f, (int)level);
        float cx = (lo.x + hi.x) / 2.0f, cy = (lo.y + hi.y) / 2.0f, cz = (lo.z + hi.z) / 2.0f;
        pos[3 * (size_t)i]     = (((x + u0) / fn - 0.5f) * (hi.x - lo.x)) * s + cx;
        pos[3 * (size_t)i + 1] = (((y + u1) / fn - 0.5f) * (hi.y - lo.y)) * s + cy;
        pos[3 * (size_t)i + 2] = (((z + u2) / fn - 0.5f) * (hi.z - lo.z)) * s + cz;
        indices[i] = idx;
    }
}

/* bitfield_func_kernel.cu:217-234 */
void orc_splat_grid_samples(int n, const uint32_t *indices, const float *density, float *grid_tmp) {
    for (int i = 0; i < n; ++i) {
        uint32_t a, b; memcpy(&a, &density[i], 4); memcpy(&b, &grid_tmp[indices[i]], 4);
        if (a > b) memcpy(&grid_tmp[indices[i]], &a, 4);
    }
}
/* bitfield_func_kernel.cu:260-275 */
void orc_ema_grid_samples_nerf(int n, float decay, float *grid, const float *grid_tmp) {
    for (int i = 0; i < n; ++i) {
        float prev = grid[i];
        grid[i] = (prev < 0.f) ? prev : fmaxf(prev * decay, grid_tmp[i]);
    }
}
/* bitfield_func_kernel.cu:302-324 (and its copy multivol_func_kernel.cu:278-300); n_bytes = cells/8 */
void orc_grid_to_bitfield(int n_bytes, const float *grid, uint8_t *bits, float mean_density, float opa_thres) {
    float thresh = opa_thres < mean_density ? opa_thres : mean_density;
    for (int i = 0; i < n_bytes; ++i) {
        uint8_t b = 0;
        for (int j = 0; j < 8; ++j) b |= grid[(size_t)i * 8 + j] > thresh ? (uint8_t)(1 << j) : 0;
        bits[i] = b;
    }
}
/* bitfield_func_kernel.cu:352-370.  The reference tests `byte && (1<<j)` (logical and), so every
 * non-zero byte adds 8; reproduced as is (faithful != 0) or corrected (faithful == 0). */
void orc_count_bitfield(int n_bytes, const uint8_t *bits, float *counter, int faithful) {
    float c = counter[0];
    for (int i = 0; i < n_bytes; ++i)
        for (int j = 0; j < 8; ++j)
            if (faithful ? (bits[i] != 0) : ((bits[i] >> j) & 1)) c += 1.0f;
    counter[0] = c;
}

/* ------------------------------------------------------------------ hash-grid level geometry
 * tiny-cuda-nn include/tiny-cuda-nn/encodings/grid.h:195-204 (grid_scale / grid_resolution, evaluated by the
 * HOST in the encoding's constructor, :982-1012): scale = exp2f(level * log2f(per_level_scale)) * base - 1,
 * resolution = ceilf(scale) + 1, entries = min(roundup8(res^3), 2^log2_T).  libm's exp2f/log2f as the
 * reference's host code uses them (numpy's differ by an ulp on some levels). */
int64_t orc_grid_geometry(int n_levels, int log2_hashmap_size, int base_resolution, float per_level_scale,
                          uint32_t *offsets, float *scales, uint32_t *resolutions) {
    const float l2 = log2f(per_level_scale);
    uint32_t off = 0;
    for (int l = 0; l < n_levels; ++l) {
        float s = exp2f((float)l * l2) * (float)base_resolution - 1.0f;
        uint32_t r = (uint32_t)ceilf(s) + 1u;
        uint64_t dense = (uint64_t)r * r * r;
        if (dense > 0xFFFFFFFFull / 2) dense = 0xFFFFFFFFull / 2;
        dense = (dense + 7) / 8 * 8;
        uint64_t cap = 1ull << log2_hashmap_size;
        if (dense > cap) dense = cap;
        offsets[l] = off; scales[l] = s; resolutions[l] = r;
        off += (uint32_t)dense;
    }
    offsets[n_levels] = off;
    return (int64_t)off;
}
