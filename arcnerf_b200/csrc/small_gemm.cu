// fp32 products of a tall matrix (n rows, n ~ 10^5..10^6) with a small one (at most 128 x 128): the dense layers of the
// NeuS-NGP SDF network and their first and second derivatives.
// Reference: arcnerf/models/base_modules/linear.py DenseLayer (torch.nn.Linear -> cuBLAS sgemm) inside GeoNet
// (geo_rad_model/linear_network_module.py:174-197) with W = 64, D = 1 on top of the hash grid
// (configs/expr/NeRF/lego/nerf_lego_neus_ngp.yaml: 32 -> 64 softplus -> 1 + 16), and torch autograd for
// BaseGeoNet.forward_with_grad (base_network.py:30-44, create_graph=True) and for the eikonal term's double backward.
// The layers are far too narrow for the tensor cores to matter (2 k-steps), and the second-order passes need fp32, so these are
// CUDA-core kernels; both are bound by streaming the tall operand through HBM once.
//   gemm_nt : C (n x N)  = A (n x K) . B^T        B (N x K) small      -> forward y = x W^T, data gradient dx = dy W
//   gemm_tn : C (N x K) += A^T . B                A (n x N), B (n x K) -> weight gradient dW = dy^T x (atomics, caller zeroes C)
#include "common.cuh"

namespace {

constexpr int kRowsPerBlock = 128;
constexpr int kMaxDim = 128;

// one thread per row: its K inputs in registers; the small matrix sits in shared memory with rows padded to a multiple of four
// and is read as broadcast float4s, four output columns at a time (16 FMAs per four 16-byte shared loads)
template <int KMAX>
__global__ void __launch_bounds__(kRowsPerBlock) gemm_nt_kernel(int64_t n, int N, int K, const float *__restrict__ A, int64_t lda,
                                                                 const float *__restrict__ B, int64_t ldb, const float *__restrict__ bias,
                                                                 float *__restrict__ C, int64_t ldc) {
    arc_pdl_sync();
    extern __shared__ __align__(16) float sm[];
    const int Kp = (K + 3) & ~3, Np = (N + 3) & ~3;
    float *Bs = sm;                                   // Np x Kp, zero padded
    float *tile = sm + Np * Kp;                       // 128 x (max(K, N) + 1): the input rows, then the output rows
    const int pitch = (K > N ? K : N) + 1;
    for (int e = threadIdx.x; e < Np * Kp; e += blockDim.x) {
        const int j = e / Kp, k = e - j * Kp;
        Bs[e] = (j < N && k < K) ? B[(size_t)j * ldb + k] : 0.f;
    }
    for (int64_t row0 = (int64_t)blockIdx.x * kRowsPerBlock; row0 < n; row0 += (int64_t)gridDim.x * kRowsPerBlock) {
        __syncthreads();
        const int rows = (int)((n - row0) < kRowsPerBlock ? (n - row0) : kRowsPerBlock);
        for (int e = threadIdx.x; e < rows * K; e += blockDim.x) {            // coalesced when lda == K
            const int r = e / K, c = e - r * K;
            tile[r * pitch + c] = A[(row0 + r) * lda + c];
        }
        __syncthreads();
        float a[KMAX];
        const int r = threadIdx.x;
        #pragma unroll
        for (int k = 0; k < KMAX; ++k) a[k] = (k < K && r < rows) ? tile[r * pitch + k] : 0.f;
        __syncthreads();
        for (int j = 0; j < Np; j += 4) {
            float acc[4];
            #pragma unroll
            for (int q = 0; q < 4; ++q) acc[q] = (bias && j + q < N) ? bias[j + q] : 0.f;
            #pragma unroll
            for (int k = 0; k < KMAX; k += 4) {
                if (k < Kp) {
                    #pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float4 w = *reinterpret_cast<const float4 *>(Bs + (j + q) * Kp + k);
                        acc[q] = fmaf(a[k], w.x, acc[q]); acc[q] = fmaf(a[k + 1], w.y, acc[q]);
                        acc[q] = fmaf(a[k + 2], w.z, acc[q]); acc[q] = fmaf(a[k + 3], w.w, acc[q]);
                    }
                }
            }
            #pragma unroll
            for (int q = 0; q < 4; ++q)
                if (j + q < N) tile[r * pitch + j + q] = acc[q];
        }
        __syncthreads();
        for (int e = threadIdx.x; e < rows * N; e += blockDim.x) {
            const int rr = e / N, c = e - rr * N;
            C[(row0 + rr) * ldc + c] = tile[rr * pitch + c];
        }
    }
}

// the small result is cut into 4 x 4 register tiles; the 256 threads form G groups of (tiles) threads, group g takes the rows
// r = g (mod G) of the block's 64-row slab (two 16-byte shared loads per 16 FMAs); one atomic flush per thread at the end
__global__ void __launch_bounds__(256) gemm_tn_kernel(int64_t n, int N, int K, const float *__restrict__ A, int64_t lda,
                                                       const float *__restrict__ B, int64_t ldb, float *__restrict__ C, int64_t ldc) {
    arc_pdl_sync();
    extern __shared__ __align__(16) float sm[];
    constexpr int kSlab = 64;
    const int Np = (N + 3) & ~3, Kp = (K + 3) & ~3;
    float *As = sm;                      // 64 x Np, zero padded
    float *Bs = sm + kSlab * Np;         // 64 x Kp
    const int tiles_k = Kp / 4, tiles = (Np / 4) * tiles_k;          // <= 256 (N * K <= 4096)
    const int G = 256 / tiles > 0 ? 256 / tiles : 1;
    const int tile_id = threadIdx.x % tiles, g = threadIdx.x / tiles;
    const bool active = g < G;
    const int i0 = (tile_id / tiles_k) * 4, j0 = (tile_id % tiles_k) * 4;
    float acc[4][4];
    #pragma unroll
    for (int p = 0; p < 4; ++p)
        #pragma unroll
        for (int q = 0; q < 4; ++q) acc[p][q] = 0.f;
    const int64_t slabs = (n + kSlab - 1) / kSlab;
    for (int64_t s = blockIdx.x; s < slabs; s += gridDim.x) {
        const int64_t row0 = s * kSlab;
        const int rows = (int)((n - row0) < kSlab ? (n - row0) : kSlab);
        __syncthreads();
        for (int e = threadIdx.x; e < kSlab * Np; e += blockDim.x) {
            const int r = e / Np, c = e - r * Np;
            As[e] = (r < rows && c < N) ? A[(row0 + r) * lda + c] : 0.f;
        }
        for (int e = threadIdx.x; e < kSlab * Kp; e += blockDim.x) {
            const int r = e / Kp, c = e - r * Kp;
            Bs[e] = (r < rows && c < K) ? B[(row0 + r) * ldb + c] : 0.f;
        }
        __syncthreads();
        if (active) {
            for (int r = g; r < kSlab; r += G) {
                const float4 av = *reinterpret_cast<const float4 *>(As + r * Np + i0);
                const float4 bv = *reinterpret_cast<const float4 *>(Bs + r * Kp + j0);
                const float aa[4] = {av.x, av.y, av.z, av.w}, bb[4] = {bv.x, bv.y, bv.z, bv.w};
                #pragma unroll
                for (int p = 0; p < 4; ++p)
                    #pragma unroll
                    for (int q = 0; q < 4; ++q) acc[p][q] = fmaf(aa[p], bb[q], acc[p][q]);
            }
        }
    }
    if (active) {
        #pragma unroll
        for (int p = 0; p < 4; ++p)
            #pragma unroll
            for (int q = 0; q < 4; ++q)
                if (i0 + p < N && j0 + q < K && acc[p][q] != 0.f) atomicAdd(C + (size_t)(i0 + p) * ldc + (j0 + q), acc[p][q]);
    }
}

// softplus with slope beta (torch.nn.Softplus(beta), threshold 20) and its first two derivatives
__global__ void softplus_kernel(size_t n, const float *__restrict__ x, float beta, int order, float *__restrict__ y) {
    arc_pdl_sync();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float z = x[i] * beta;
        float out;
        if (order == 0) {
            out = z > 20.f ? x[i] : log1pf(expf(z)) / beta;
        } else {
            const float s = z > 20.f ? 1.f : 1.f / (1.f + expf(-z));      // torch's backward of the thresholded branch is exactly 1
            out = order == 1 ? s : (z > 20.f ? 0.f : beta * s * (1.f - s));
        }
        y[i] = out;
    }
}

}  // namespace

extern "C" {

// C (n x N, leading dimension ldc) = A (n x K, lda) . B^T with B (N x K, ldb), plus bias (N) when given.  N, K <= 128.
// Replaces torch.nn.functional.linear on tall inputs (DenseLayer / nn.Linear, arcnerf/models/base_modules/linear.py).
int arcb200_gemm_nt(const float *A, int64_t lda, const float *B, int64_t ldb, const float *bias, float *C, int64_t ldc, int64_t n,
                    int N, int K, void *stream) {
    ARC_REQUIRE(n >= 0 && N >= 1 && K >= 1 && N <= kMaxDim && K <= kMaxDim, "gemm_nt: the small dimensions must lie in [1, %d]", kMaxDim);
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(A && B && C && lda >= K && ldb >= K && ldc >= N, "gemm_nt: null pointer or leading dimension too small");
    const size_t smem = ((size_t)((N + 3) & ~3) * ((K + 3) & ~3) + (size_t)kRowsPerBlock * ((K > N ? K : N) + 1)) * sizeof(float);
    long long blocks = (n + kRowsPerBlock - 1) / kRowsPerBlock;
    const long long cap = (long long)arc_sm_count() * 4;
    const unsigned grid = (unsigned)(blocks < cap ? blocks : cap);
    cudaStream_t st = (cudaStream_t)stream;
#define NT(KM) do { \
        if (smem > 48 * 1024) ARC_CUDA(cudaFuncSetAttribute(gemm_nt_kernel<KM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        arc_launch(gemm_nt_kernel<KM>, grid, kRowsPerBlock, smem, st, n, N, K, A, lda, B, ldb, bias, C, ldc); } while (0)
    if (K <= 32) NT(32); else if (K <= 64) NT(64); else NT(128);
#undef NT
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// C (N x K, ldc) += A^T . B with A (n x N, lda), B (n x K, ldb); the caller zeroes C.  N, K <= 128; a result of more than 256
// register tiles (4 x 4) is produced in several launches over row ranges of C.
// Replaces the weight-gradient GEMM autograd runs for nn.Linear.
int arcb200_gemm_tn(const float *A, int64_t lda, const float *B, int64_t ldb, float *C, int64_t ldc, int64_t n, int N, int K,
                    void *stream) {
    ARC_REQUIRE(n >= 0 && N >= 1 && K >= 1 && N <= kMaxDim && K <= kMaxDim, "gemm_tn: the small dimensions must lie in [1, %d]", kMaxDim);
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(A && B && C && lda >= N && ldb >= K && ldc >= K, "gemm_tn: null pointer or leading dimension too small");
    const int tiles_k = (K + 3) / 4;
    const int rows_per_launch = (256 / tiles_k) * 4;                 // rows of C whose tiles fit one CTA
    long long slabs = (n + 63) / 64;
    const long long cap = (long long)arc_sm_count() * 4;
    for (int c0 = 0; c0 < N; c0 += rows_per_launch) {
        const int Nc = N - c0 < rows_per_launch ? N - c0 : rows_per_launch;
        const size_t smem = (size_t)64 * (((Nc + 3) & ~3) + ((K + 3) & ~3)) * sizeof(float);
        if (smem > 48 * 1024) ARC_CUDA(cudaFuncSetAttribute(gemm_tn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        arc_launch(gemm_tn_kernel, (unsigned)(slabs < cap ? slabs : cap), 256, smem, (cudaStream_t)stream, n, Nc, K, A + c0, lda, B, ldb,
                   C + (size_t)c0 * ldc, ldc);
        ARC_LAUNCH_CHECK();
    }
    return ARCB200_OK;
}

// y = d^order softplus_beta(x) / dx^order for order 0, 1, 2 (torch.nn.Softplus(beta), threshold 20: arcnerf/models/base_modules/
// activation.py) - the activation of the SDF network and what its first and second derivative passes need.
int arcb200_softplus(const float *x, float beta, int order, float *y, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0 && order >= 0 && order <= 2 && beta > 0.f, "softplus: bad argument");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(x && y, "softplus: null pointer");
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)arc_sm_count() * 16;
    arc_launch(softplus_kernel, (unsigned)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream, (size_t)n, x, beta, order, y);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
