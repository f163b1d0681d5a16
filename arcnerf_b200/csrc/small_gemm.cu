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

// one thread per row: its K inputs in registers, the small matrix in shared memory (read as a broadcast)
template <int KMAX>
__global__ void __launch_bounds__(kRowsPerBlock) gemm_nt_kernel(int64_t n, int N, int K, const float *__restrict__ A, int64_t lda,
                                                                 const float *__restrict__ B, int64_t ldb, const float *__restrict__ bias,
                                                                 float *__restrict__ C, int64_t ldc) {
    arc_pdl_sync();
    extern __shared__ float sm[];
    float *Bs = sm;                                   // N x K
    float *tile = sm + N * K;                         // 128 x (max(K, N) + 1): the input rows, then the output rows
    const int pitch = (K > N ? K : N) + 1;
    for (int e = threadIdx.x; e < N * K; e += blockDim.x) Bs[e] = B[(size_t)(e / K) * ldb + (e % K)];
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
        for (int j = 0; j < N; ++j) {
            float acc = bias ? bias[j] : 0.f;
            #pragma unroll
            for (int k = 0; k < KMAX; ++k)
                if (k < K) acc = fmaf(a[k], Bs[j * K + k], acc);
            tile[r * pitch + j] = acc;
        }
        __syncthreads();
        for (int e = threadIdx.x; e < rows * N; e += blockDim.x) {
            const int rr = e / N, c = e - rr * N;
            C[(row0 + rr) * ldc + c] = tile[rr * pitch + c];
        }
    }
}

// every thread owns up to 16 entries of the small result and walks the rows of the block's slab through shared memory
__global__ void __launch_bounds__(256) gemm_tn_kernel(int64_t n, int N, int K, const float *__restrict__ A, int64_t lda,
                                                       const float *__restrict__ B, int64_t ldb, float *__restrict__ C, int64_t ldc) {
    arc_pdl_sync();
    extern __shared__ float sm[];
    constexpr int kSlab = 64;
    float *As = sm;                      // 64 x N
    float *Bs = sm + kSlab * N;          // 64 x K
    float acc[16];
    #pragma unroll
    for (int q = 0; q < 16; ++q) acc[q] = 0.f;
    const int64_t slabs = (n + kSlab - 1) / kSlab;
    for (int64_t s = blockIdx.x; s < slabs; s += gridDim.x) {
        const int64_t row0 = s * kSlab;
        const int rows = (int)((n - row0) < kSlab ? (n - row0) : kSlab);
        __syncthreads();
        for (int e = threadIdx.x; e < kSlab * N; e += blockDim.x) {
            const int r = e / N, c = e - r * N;
            As[e] = r < rows ? A[(row0 + r) * lda + c] : 0.f;
        }
        for (int e = threadIdx.x; e < kSlab * K; e += blockDim.x) {
            const int r = e / K, c = e - r * K;
            Bs[e] = r < rows ? B[(row0 + r) * ldb + c] : 0.f;
        }
        __syncthreads();
        #pragma unroll
        for (int q = 0; q < 16; ++q) {
            const int e = threadIdx.x + q * 256;
            if (e < N * K) {
                const int i = e / K, j = e - i * K;
                float v = acc[q];
                for (int r = 0; r < kSlab; ++r) v = fmaf(As[r * N + i], Bs[r * K + j], v);
                acc[q] = v;
            }
        }
    }
    #pragma unroll
    for (int q = 0; q < 16; ++q) {
        const int e = threadIdx.x + q * 256;
        if (e < N * K && acc[q] != 0.f) atomicAdd(C + (size_t)(e / K) * ldc + (e % K), acc[q]);
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
    const size_t smem = ((size_t)N * K + (size_t)kRowsPerBlock * ((K > N ? K : N) + 1)) * sizeof(float);
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

// C (N x K, ldc) += A^T . B with A (n x N, lda), B (n x K, ldb); the caller zeroes C.  N * K <= 4096.
// Replaces the weight-gradient GEMM autograd runs for nn.Linear.
int arcb200_gemm_tn(const float *A, int64_t lda, const float *B, int64_t ldb, float *C, int64_t ldc, int64_t n, int N, int K,
                    void *stream) {
    ARC_REQUIRE(n >= 0 && N >= 1 && K >= 1 && N <= kMaxDim && K <= kMaxDim && N * K <= 4096, "gemm_tn: small dimensions out of range (N * K <= 4096)");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(A && B && C && lda >= N && ldb >= K && ldc >= K, "gemm_tn: null pointer or leading dimension too small");
    const size_t smem = (size_t)64 * (N + K) * sizeof(float);
    long long slabs = (n + 63) / 64;
    const long long cap = (long long)arc_sm_count() * 4;
    if (smem > 48 * 1024) ARC_CUDA(cudaFuncSetAttribute(gemm_tn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    arc_launch(gemm_tn_kernel, (unsigned)(slabs < cap ? slabs : cap), 256, smem, (cudaStream_t)stream, n, N, K, A, lda, B, ldb, C, ldc);
    ARC_LAUNCH_CHECK();
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
