// Mesh extraction on the GPU: voxel-centre lattice points and marching cubes over a dense scalar volume.
//
// Replaces, for the inference path arcnerf/eval/infer_func.py:448-500 (run_infer_volume):
//   * Volume.cal_volume_pts (arcnerf/geometry/volume.py:225-232): the (n^3, 3) lattice of voxel centres - generated here slab by
//     slab on the fly, so the 512^3 grid of config 4 never exists as a 1.6 GB point tensor;
//   * extract_mesh (arcnerf/geometry/mesh.py:12-47): skimage.measure.marching_cubes on the host -> welded vertices, faces and
//     vertex normals produced on the device in four passes:
//       A  per lattice point: how many of its three +axis edges cross the level (block sums)           \  exclusive scans of the
//       B  per lattice point: vertex base index; writes vertex positions and normals                    >  block sums by one CTA
//       C  per cell: triangle count from the 256-case table (block sums)                               /
//       D  per cell: faces (vertex indices looked up through the per-point base indices)
//     A vertex lives on a lattice edge and is computed once, from the edge's lower end point, so vertices are welded by
//     construction (no sort/unique pass) and the mesh is watertight wherever the surface does not leave the volume.
// The case table is generated (tools/gen_mc_table.py); skimage's Lewiner tables are not in this image, so vertex/face ORDER and
// the resolution of ambiguous cases differ from skimage's - positions of vertices on the same lattice edges are the same linear
// interpolation.  Compiled with -fmad=false: tests compare with a numpy restatement bit for bit.
#include <cuda_runtime.h>
#include <stdint.h>

#include "arcb200.h"
#include "mc_table.cuh"

namespace {

constexpr int kBlock = 256;

// ------------------------------------------------------------------------------------------------ lattice points
// torch.linspace(lo + vs / 2, hi - vs / 2, n)[i]: symmetric evaluation (from the start for the first half, from the end for the
// second), step = (end - start) / (n - 1), each value one fused multiply-add (what ATen's CPU kernel evaluates: checked bit for
// bit against torch.linspace, tests/golden/make_golden_volume.py)
__device__ __forceinline__ float linspace_at(float start, float end, float step, int n, int i) {
    return i < n / 2 ? fmaf(step, (float)i, start) : fmaf(-step, (float)(n - 1 - i), end);
}

__global__ void volume_pts_kernel(float3 start, float3 end, float3 step, int n, int x0, int64_t n_pts, float* __restrict__ pts) {
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_pts) return;
    int k = (int)(idx % n);
    int j = (int)((idx / n) % n);
    int i = (int)(idx / ((int64_t)n * n)) + x0;
    pts[3 * idx + 0] = linspace_at(start.x, end.x, step.x, n, i);
    pts[3 * idx + 1] = linspace_at(start.y, end.y, step.y, n, j);
    pts[3 * idx + 2] = linspace_at(start.z, end.z, step.z, n, k);
}

// ------------------------------------------------------------------------------------------------ marching cubes
struct Vol {
    const float* v;
    int nx, ny, nz;
    float level;
    int ascent;
    __device__ __forceinline__ float at(int i, int j, int k) const { return v[((int64_t)i * ny + j) * nz + k]; }
    __device__ __forceinline__ bool inside(float f) const { return ascent ? f < level : f > level; }
};

__device__ __forceinline__ int block_exclusive_scan(int x, int* total) {
    // 256 threads: warp scans + scan of the 8 warp sums
    __shared__ int warp_sum[kBlock / 32];
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += y;
    }
    __syncthreads();                                      // protects warp_sum between successive calls
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    int base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kBlock / 32; ++w) {
        int s = warp_sum[w];
        if (w < warp) base += s;
        tot += s;
    }
    *total = tot;
    return base + incl - x;
}

// crossings of the three +axis edges anchored at lattice point (i, j, k): bit a set when edge a crosses
__device__ __forceinline__ int point_crossings(const Vol& V, int i, int j, int k, float f0, float* f1) {
    bool in0 = V.inside(f0);
    int m = 0;
    if (i + 1 < V.nx) { f1[0] = V.at(i + 1, j, k); m |= (V.inside(f1[0]) != in0) << 0; }
    if (j + 1 < V.ny) { f1[1] = V.at(i, j + 1, k); m |= (V.inside(f1[1]) != in0) << 1; }
    if (k + 1 < V.nz) { f1[2] = V.at(i, j, k + 1); m |= (V.inside(f1[2]) != in0) << 2; }
    return m;
}

__global__ void mc_point_count_kernel(Vol V, int64_t n_pts, int32_t* __restrict__ block_sums) {
    int64_t p = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    int c = 0;
    if (p < n_pts) {
        int k = (int)(p % V.nz), j = (int)((p / V.nz) % V.ny), i = (int)(p / ((int64_t)V.nz * V.ny));
        float f1[3];
        c = __popc(point_crossings(V, i, j, k, V.v[p], f1));
    }
    int tot;
    block_exclusive_scan(c, &tot);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

// central difference of the volume along `axis` at a lattice point (one-sided on the border), in index units
__device__ __forceinline__ float lattice_grad(const Vol& V, int i, int j, int k, int axis) {
    int n = axis == 0 ? V.nx : axis == 1 ? V.ny : V.nz;
    int c = axis == 0 ? i : axis == 1 ? j : k;
    int lo = c > 0 ? c - 1 : c, hi = c + 1 < n ? c + 1 : c;
    float a = axis == 0 ? V.at(lo, j, k) : axis == 1 ? V.at(i, lo, k) : V.at(i, j, lo);
    float b = axis == 0 ? V.at(hi, j, k) : axis == 1 ? V.at(i, hi, k) : V.at(i, j, hi);
    return (b - a) / (float)(hi - lo);
}

__global__ void mc_point_emit_kernel(Vol V, int64_t n_pts, const int32_t* __restrict__ block_base, float3 spacing, float3 offset,
                                     int32_t* __restrict__ vbase, float* __restrict__ verts, float* __restrict__ normals,
                                     int64_t v_cap) {
    int64_t p = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    int m = 0, i = 0, j = 0, k = 0;
    float f0 = 0.f, f1[3] = {0.f, 0.f, 0.f};
    if (p < n_pts) {
        k = (int)(p % V.nz), j = (int)((p / V.nz) % V.ny), i = (int)(p / ((int64_t)V.nz * V.ny));
        f0 = V.v[p];
        m = point_crossings(V, i, j, k, f0, f1);
    }
    int tot;
    int base = block_base[blockIdx.x] + block_exclusive_scan(__popc(m), &tot);
    if (p >= n_pts) return;
    vbase[p] = base;
    if (!m) return;
    float g0[3];
    if (normals) {
#pragma unroll
        for (int a = 0; a < 3; ++a) g0[a] = lattice_grad(V, i, j, k, a);
    }
    int r = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        if (!((m >> a) & 1)) continue;
        int64_t vi = base + r++;
        if (vi >= v_cap) return;
        float t = (V.level - f0) / (f1[a] - f0);
        float pi = (float)i + (a == 0 ? t : 0.f), pj = (float)j + (a == 1 ? t : 0.f), pk = (float)k + (a == 2 ? t : 0.f);
        verts[3 * vi + 0] = pi * spacing.x - offset.x;
        verts[3 * vi + 1] = pj * spacing.y - offset.y;
        verts[3 * vi + 2] = pk * spacing.z - offset.z;
        if (normals) {
            int i1 = i + (a == 0), j1 = j + (a == 1), k1 = k + (a == 2);
            float n[3];
#pragma unroll
            for (int b = 0; b < 3; ++b) {
                float g1 = lattice_grad(V, i1, j1, k1, b);
                n[b] = g0[b] + t * (g1 - g0[b]);
            }
            // gradient in world units; outside is where the value falls ('descent') or rises ('ascent')
            float sgn = V.ascent ? 1.f : -1.f;
            float nx = sgn * (n[0] / spacing.x), ny = sgn * (n[1] / spacing.y), nz = sgn * (n[2] / spacing.z);
            float len = sqrtf(nx * nx + ny * ny + nz * nz) + 1e-8f;              // transformation.py:11-25 normalize
            normals[3 * vi + 0] = nx / len;
            normals[3 * vi + 1] = ny / len;
            normals[3 * vi + 2] = nz / len;
        }
    }
}

__device__ __forceinline__ int cell_case(const Vol& V, int i, int j, int k) {
    int c = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) c |= (int)V.inside(V.at(i + (q & 1), j + ((q >> 1) & 1), k + ((q >> 2) & 1))) << q;
    return c;
}

__global__ void mc_cell_count_kernel(Vol V, int64_t n_cells, int32_t* __restrict__ block_sums) {
    int64_t c = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    int nt = 0;
    if (c < n_cells) {
        int cz = V.nz - 1, cy = V.ny - 1;
        int k = (int)(c % cz), j = (int)((c / cz) % cy), i = (int)(c / ((int64_t)cz * cy));
        nt = kMcNumTris[cell_case(V, i, j, k)];
    }
    int tot;
    block_exclusive_scan(nt, &tot);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

__global__ void mc_cell_emit_kernel(Vol V, int64_t n_cells, const int32_t* __restrict__ block_base,
                                    const int32_t* __restrict__ vbase, int32_t* __restrict__ faces, int64_t f_cap) {
    int64_t c = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    int cs = 0, i = 0, j = 0, k = 0;
    if (c < n_cells) {
        int cz = V.nz - 1, cy = V.ny - 1;
        k = (int)(c % cz), j = (int)((c / cz) % cy), i = (int)(c / ((int64_t)cz * cy));
        cs = cell_case(V, i, j, k);
    }
    int nt = kMcNumTris[cs], tot;
    int64_t base = block_base[blockIdx.x] + block_exclusive_scan(nt, &tot);
    if (!nt) return;
    // vertex index of each of the 12 cube edges that carries one: base of the anchor point + rank of the axis among that point's
    // crossing edges (lower axes first, the order pass B writes them in)
    int vid[12];
#pragma unroll
    for (int e = 0; e < 12; ++e) {
        vid[e] = -1;
        int a = kMcEdges[e][0], b = kMcEdges[e][1];
        if (((cs >> a) & 1) == ((cs >> b) & 1)) continue;
        int axis = kMcEdges[e][2];
        int pi = i + kMcEdges[e][3], pj = j + kMcEdges[e][4], pk = k + kMcEdges[e][5];
        int64_t p = ((int64_t)pi * V.ny + pj) * V.nz + pk;
        int r = 0;
        if (axis > 0) {
            bool in0 = V.inside(V.v[p]);
            if (pi + 1 < V.nx) r += V.inside(V.at(pi + 1, pj, pk)) != in0;
            if (axis > 1 && pj + 1 < V.ny) r += V.inside(V.at(pi, pj + 1, pk)) != in0;
        }
        vid[e] = vbase[p] + r;
    }
    for (int t = 0; t < nt; ++t) {
        int64_t f = base + t;
        if (f >= f_cap) return;
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            int e = kMcTriTable[cs][3 * t + q], v = 0;
#pragma unroll
            for (int ee = 0; ee < 12; ++ee)
                if (ee == e) v = vid[ee];                 // keeps vid[] in registers
            faces[3 * f + q] = v;
        }
    }
}

// exclusive scan of n block sums in place by one CTA (n is in the 10^5 range: 16 items per thread and pass), total -> sums[n]
__global__ void mc_scan_kernel(int32_t* __restrict__ sums, int64_t n) {
    constexpr int kItems = 16;
    __shared__ int carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int64_t start = 0; start < n; start += (int64_t)kBlock * kItems) {
        int64_t b = start + (int64_t)threadIdx.x * kItems;
        int x[kItems], s = 0;
#pragma unroll
        for (int q = 0; q < kItems; ++q) {
            x[q] = b + q < n ? sums[b + q] : 0;
            s += x[q];
        }
        int tot;
        int base = carry_s + block_exclusive_scan(s, &tot);
#pragma unroll
        for (int q = 0; q < kItems; ++q) {
            if (b + q < n) sums[b + q] = base;
            base += x[q];
        }
        __syncthreads();
        if (threadIdx.x == 0) carry_s += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) sums[n] = carry_s;
}

// ------------------------------------------------------------------------------------------------ mesh components
// get_normals / get_face_centers (arcnerf/geometry/mesh.py:88-125; trimesh face normals + mean_vertex_normals): unit face normal by
// the right-hand rule, face centre, and per vertex the sum of the unit normals of its faces (normalised by the second kernel)
__global__ void mesh_face_kernel(const float* __restrict__ verts, const int32_t* __restrict__ faces, int64_t n_f,
                                 float* __restrict__ face_normals, float* __restrict__ face_centers, float* __restrict__ vert_acc) {
    int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_f) return;
    int ia = faces[3 * f], ib = faces[3 * f + 1], ic = faces[3 * f + 2];
    float a[3], b[3], c[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) { a[k] = verts[3 * (int64_t)ia + k]; b[k] = verts[3 * (int64_t)ib + k]; c[k] = verts[3 * (int64_t)ic + k]; }
    float u[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, w[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
    float n[3] = {u[1] * w[2] - u[2] * w[1], u[2] * w[0] - u[0] * w[2], u[0] * w[1] - u[1] * w[0]};
    float len = sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    float inv = len > 0.f ? 1.f / len : 0.f;               // degenerate faces contribute nothing (trimesh drops them too)
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        n[k] *= inv;
        if (face_normals) face_normals[3 * f + k] = n[k];
        if (face_centers) face_centers[3 * f + k] = (a[k] + b[k] + c[k]) / 3.0f;
    }
    if (vert_acc) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            atomicAdd(vert_acc + 3 * (int64_t)ia + k, n[k]);
            atomicAdd(vert_acc + 3 * (int64_t)ib + k, n[k]);
            atomicAdd(vert_acc + 3 * (int64_t)ic + k, n[k]);
        }
    }
}

__global__ void mesh_vert_normalize_kernel(float* __restrict__ vn, int64_t n_v) {
    int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n_v) return;
    float x = vn[3 * v], y = vn[3 * v + 1], z = vn[3 * v + 2];
    float len = sqrtf(x * x + y * y + z * z) + 1e-8f;      // transformation.py:11-25 normalize
    vn[3 * v] = x / len; vn[3 * v + 1] = y / len; vn[3 * v + 2] = z / len;
}

inline int64_t blocks_of(int64_t n) { return (n + kBlock - 1) / kBlock; }

inline int check_vol(const float* vol, int nx, int ny, int nz) {
    if (!vol || nx < 2 || ny < 2 || nz < 2) return ARCB200_ERR_ARG;
    if ((int64_t)nx * ny * nz > ((int64_t)1 << 33)) return ARCB200_ERR_ARG;
    return 0;
}

}  // namespace

extern "C" {

int arcb200_volume_pts(const float* lo, const float* hi, int n_grid, int x0, int x1, float* pts, void* stream) {
    if (!lo || !hi || !pts || n_grid < 1 || x0 < 0 || x1 > n_grid || x0 > x1) return ARCB200_ERR_ARG;
    if (x0 == x1) return ARCB200_OK;
    float s[3], e[3], st[3];
    for (int a = 0; a < 3; ++a) {
        // volume.py:225-231: v = (hi - lo) / n in fp32 (a torch float tensor), the ends in Python doubles rounded to fp32
        float v = (hi[a] - lo[a]) / (float)n_grid;
        s[a] = (float)((double)lo[a] + 0.5 * (double)v);
        e[a] = (float)((double)hi[a] - 0.5 * (double)v);
        st[a] = n_grid > 1 ? (e[a] - s[a]) / (float)(n_grid - 1) : 0.f;
    }
    int64_t n_pts = (int64_t)(x1 - x0) * n_grid * n_grid;
    volume_pts_kernel<<<(unsigned)blocks_of(n_pts), kBlock, 0, (cudaStream_t)stream>>>(
        make_float3(s[0], s[1], s[2]), make_float3(e[0], e[1], e[2]), make_float3(st[0], st[1], st[2]), n_grid, x0, n_pts, pts);
    return cudaGetLastError() == cudaSuccess ? ARCB200_OK : ARCB200_ERR_CUDA;
}

int arcb200_mc_scratch_sizes(int nx, int ny, int nz, int64_t* out) {
    if (!out || nx < 2 || ny < 2 || nz < 2) return ARCB200_ERR_ARG;
    int64_t n_pts = (int64_t)nx * ny * nz, n_cells = (int64_t)(nx - 1) * (ny - 1) * (nz - 1);
    out[0] = blocks_of(n_pts) + 1;                        // int32 entries of point_blocks
    out[1] = blocks_of(n_cells) + 1;                      // int32 entries of cell_blocks
    out[2] = n_pts;                                       // int32 entries of vbase
    return ARCB200_OK;
}

int arcb200_mc_count(const float* vol, int nx, int ny, int nz, float level, int ascent, int32_t* point_blocks, int32_t* cell_blocks,
                     void* stream) {
    if (int rc = check_vol(vol, nx, ny, nz)) return rc;
    if (!point_blocks || !cell_blocks) return ARCB200_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    Vol V{vol, nx, ny, nz, level, ascent};
    int64_t n_pts = (int64_t)nx * ny * nz, n_cells = (int64_t)(nx - 1) * (ny - 1) * (nz - 1);
    mc_point_count_kernel<<<(unsigned)blocks_of(n_pts), kBlock, 0, st>>>(V, n_pts, point_blocks);
    mc_scan_kernel<<<1, kBlock, 0, st>>>(point_blocks, blocks_of(n_pts));
    mc_cell_count_kernel<<<(unsigned)blocks_of(n_cells), kBlock, 0, st>>>(V, n_cells, cell_blocks);
    mc_scan_kernel<<<1, kBlock, 0, st>>>(cell_blocks, blocks_of(n_cells));
    return cudaGetLastError() == cudaSuccess ? ARCB200_OK : ARCB200_ERR_CUDA;
}

int arcb200_mc_emit(const float* vol, int nx, int ny, int nz, float level, int ascent, const float* spacing, const float* offset,
                    const int32_t* point_blocks, const int32_t* cell_blocks, int32_t* vbase, float* verts, float* normals,
                    int64_t v_cap, int32_t* faces, int64_t f_cap, void* stream) {
    if (int rc = check_vol(vol, nx, ny, nz)) return rc;
    if (!spacing || !offset || !point_blocks || !cell_blocks || !vbase || v_cap < 0 || f_cap < 0) return ARCB200_ERR_ARG;
    if ((v_cap > 0 && !verts) || (f_cap > 0 && !faces)) return ARCB200_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    Vol V{vol, nx, ny, nz, level, ascent};
    int64_t n_pts = (int64_t)nx * ny * nz, n_cells = (int64_t)(nx - 1) * (ny - 1) * (nz - 1);
    mc_point_emit_kernel<<<(unsigned)blocks_of(n_pts), kBlock, 0, st>>>(
        V, n_pts, point_blocks, make_float3(spacing[0], spacing[1], spacing[2]), make_float3(offset[0], offset[1], offset[2]), vbase,
        verts, normals, v_cap);
    mc_cell_emit_kernel<<<(unsigned)blocks_of(n_cells), kBlock, 0, st>>>(V, n_cells, cell_blocks, vbase, faces, f_cap);
    return cudaGetLastError() == cudaSuccess ? ARCB200_OK : ARCB200_ERR_CUDA;
}

int arcb200_mesh_components(const float* verts, const int32_t* faces, int64_t n_v, int64_t n_f, float* face_normals,
                            float* face_centers, float* vert_normals, void* stream) {
    if (n_v < 0 || n_f < 0 || (n_f > 0 && (!verts || !faces))) return ARCB200_ERR_ARG;
    cudaStream_t st = (cudaStream_t)stream;
    if (vert_normals && n_v > 0 && cudaMemsetAsync(vert_normals, 0, sizeof(float) * 3 * n_v, st) != cudaSuccess) return ARCB200_ERR_CUDA;
    if (n_f > 0) mesh_face_kernel<<<(unsigned)blocks_of(n_f), kBlock, 0, st>>>(verts, faces, n_f, face_normals, face_centers, vert_normals);
    if (vert_normals && n_v > 0) mesh_vert_normalize_kernel<<<(unsigned)blocks_of(n_v), kBlock, 0, st>>>(vert_normals, n_v);
    return cudaGetLastError() == cudaSuccess ? ARCB200_OK : ARCB200_ERR_CUDA;
}

}  // extern "C"
