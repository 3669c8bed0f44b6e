// K3 (FP32 path): register-blocked SIMT SGEMM with fused bias / ReLU / dropout / mask epilogues.
// This is the tight-tolerance (1e-5) path the parity tests run; the throughput path is the
// tcgen05 TF32 kernel in mlp_tcgen05.cu.
#include "ctr_common.cuh"

namespace {

constexpr int BM = 128, BN = 128, BK = 16, PADS = 4;

struct Epilogue {
    const float* bias;     // [N] or null
    int relu;
    float keep;            // dropout keep prob (1 = off)
    uint64_t seed;
    const float* mask;     // [M, ldmask] or null : out *= mask > 0 ? mask_scale : 0
    int64_t ldmask;
    float mask_scale;
};

__device__ __forceinline__ float dropout_u01(uint64_t seed, uint64_t idx) {
    uint64_t z = seed + idx * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (float) (z >> 40) * (1.0f / 16777216.0f);
}

// C[M,N] (+split partials) = A'[M,K] * B'[K,N];  TA: A stored [K,M]; TB: B stored [N,K]
template <bool TA, bool TB>
__global__ void __launch_bounds__(256)
sgemm_kernel(int M, int N, int K, const float* __restrict__ A, int64_t lda, const float* __restrict__ B, int64_t ldb,
             float* __restrict__ C, int64_t ldc, int k_per_split, Epilogue ep) {
    __shared__ __align__(16) float As[2][BK][BM + PADS];
    __shared__ __align__(16) float Bs[2][BK][BN + PADS];
    const int t = threadIdx.x;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int kbeg = blockIdx.z * k_per_split;
    const int kend = (kbeg + k_per_split < K) ? kbeg + k_per_split : K;
    const int ty = t / 16, tx = t % 16;

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    float ra[8], rb[8];
    auto load_tiles = [&](int k0) {
        if (!TA) {   // A[m, k], k contiguous: thread -> (m = t/2, 8 consecutive k)
            const int m = m0 + t / 2, kq = k0 + (t % 2) * 8;
#pragma unroll
            for (int i = 0; i < 8; ++i) ra[i] = (m < M && kq + i < kend) ? A[(int64_t) m * lda + kq + i] : 0.f;
        } else {     // A[k, m], m contiguous: thread -> (k = t/16, 8 consecutive m)
            const int k = k0 + t / 16, mq = m0 + (t % 16) * 8;
#pragma unroll
            for (int i = 0; i < 8; ++i) ra[i] = (k < kend && mq + i < M) ? A[(int64_t) k * lda + mq + i] : 0.f;
        }
        if (!TB) {   // B[k, n], n contiguous
            const int k = k0 + t / 16, nq = n0 + (t % 16) * 8;
#pragma unroll
            for (int i = 0; i < 8; ++i) rb[i] = (k < kend && nq + i < N) ? B[(int64_t) k * ldb + nq + i] : 0.f;
        } else {     // B[n, k], k contiguous
            const int n = n0 + t / 2, kq = k0 + (t % 2) * 8;
#pragma unroll
            for (int i = 0; i < 8; ++i) rb[i] = (n < N && kq + i < kend) ? B[(int64_t) n * ldb + kq + i] : 0.f;
        }
    };
    auto store_tiles = [&](int buf) {
        if (!TA) {
            const int m = t / 2, kq = (t % 2) * 8;
#pragma unroll
            for (int i = 0; i < 8; ++i) As[buf][kq + i][m] = ra[i];
        } else {
            const int k = t / 16, mq = (t % 16) * 8;
#pragma unroll
            for (int i = 0; i < 8; ++i) As[buf][k][mq + i] = ra[i];
        }
        if (!TB) {
            const int k = t / 16, nq = (t % 16) * 8;
#pragma unroll
            for (int i = 0; i < 8; ++i) Bs[buf][k][nq + i] = rb[i];
        } else {
            const int n = t / 2, kq = (t % 2) * 8;
#pragma unroll
            for (int i = 0; i < 8; ++i) Bs[buf][kq + i][n] = rb[i];
        }
    };

    int buf = 0;
    if (kbeg < kend) {
        load_tiles(kbeg);
        store_tiles(0);
    }
    __syncthreads();
    for (int k0 = kbeg; k0 < kend; k0 += BK) {
        const bool more = k0 + BK < kend;
        if (more) load_tiles(k0 + BK);
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
            const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        if (more) {
            store_tiles(buf ^ 1);
            __syncthreads();
            buf ^= 1;
        }
    }

    float* Cz = C + (int64_t) blockIdx.z * M * ldc;   // split partials are stacked [z][M][ldc]
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
            if (n >= N) continue;
            float v = acc[i][j];
            if (ep.bias) v += ep.bias[n];
            if (ep.relu) v = v > 0.f ? v : 0.f;
            if (ep.keep < 1.f) v = dropout_u01(ep.seed, (uint64_t) m * (uint64_t) N + n) < ep.keep ? v / ep.keep : 0.f;
            if (ep.mask) v = ep.mask[(int64_t) m * ep.ldmask + n] > 0.f ? v * ep.mask_scale : 0.f;
            Cz[(int64_t) m * ldc + n] = v;
        }
    }
}

__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ ws, int splits, int64_t n, float* __restrict__ out) {
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        float a = 0.f;
        for (int z = 0; z < splits; ++z) a += ws[(int64_t) z * n + i];
        out[i] = a;
    }
}

int launch_sgemm(ctr_handle* h, cudaStream_t st, bool ta, bool tb, int M, int N, int K, const float* A, int64_t lda,
                 const float* B, int64_t ldb, float* C, int64_t ldc, int splits, const Epilogue& ep) {
    dim3 grid((unsigned) ctr_cdiv(N, BN), (unsigned) ctr_cdiv(M, BM), (unsigned) splits);
    int kps = (int) ctr_align(ctr_cdiv(K, splits), BK);
    if (!ta && !tb) sgemm_kernel<false, false><<<grid, 256, 0, st>>>(M, N, K, A, lda, B, ldb, C, ldc, kps, ep);
    else if (!ta && tb) sgemm_kernel<false, true><<<grid, 256, 0, st>>>(M, N, K, A, lda, B, ldb, C, ldc, kps, ep);
    else if (ta && !tb) sgemm_kernel<true, false><<<grid, 256, 0, st>>>(M, N, K, A, lda, B, ldb, C, ldc, kps, ep);
    else sgemm_kernel<true, true><<<grid, 256, 0, st>>>(M, N, K, A, lda, B, ldb, C, ldc, kps, ep);
    CTR_CHECK_LAUNCH(h, "sgemm");
    return CTR_OK;
}

int wgrad_splits(ctr_handle* h, int B, int K, int N) {
    int64_t tiles = ctr_cdiv(K, BM) * ctr_cdiv(N, BN);
    int64_t want = ctr_cdiv((int64_t) h->sm_count * 2, tiles);
    int64_t maxs = ctr_cdiv(B, 512);   // keep at least 512 reduction rows per split
    if (want > maxs) want = maxs;
    if (want < 1) want = 1;
    return (int) want;
}

}  // namespace

// TF32 tcgen05 path (mlp_tcgen05.cu)
int ctr_tc_mlp_fwd(ctr_handle* h, cudaStream_t st, int B, int K, int N, const float* X, int64_t ldx, const float* Wt,
                   const float* bias, int relu, float keep, uint64_t seed, float* Y, int64_t ldy, uint32_t* relu_bits);
int ctr_tc_mlp_bwd_data(ctr_handle* h, cudaStream_t st, int B, int K, int N, const float* dZ, int64_t lddz,
                        const float* W, const float* mask, int64_t ldmask, float mask_scale, float* dX, int64_t lddx,
                        const uint32_t* relu_bits);
int64_t ctr_tc_relu_bits_bytes(int B, int N);
int64_t ctr_tc_bwd_weight_ws_bytes(ctr_handle* h, int B, int K, int N);
int ctr_tc_mlp_bwd_weight(ctr_handle* h, cudaStream_t st, int B, int K, int N, const float* X, int64_t ldx,
                          const float* dZ, int64_t lddz, float* dW, float* dbias, void* ws, int64_t ws_bytes);

extern "C" int ctr_mlp_fwd(ctr_handle* h, ctr_stream s, int precision, int B, int K, int N, const float* X, int64_t ldx,
                           const float* W, const float* Wt, const float* bias, int relu, float keep, uint64_t seed,
                           float* Y, int64_t ldy) {
    CTR_REQUIRE(h, X && Y && B >= 0 && K > 0 && N > 0 && ldx >= K && ldy >= N, "mlp_fwd: bad args");
    CTR_REQUIRE(h, keep > 0.f && keep <= 1.f, "mlp_fwd: keep must be in (0,1]");
    if (B == 0) return CTR_OK;
    if (precision == CTR_PREC_TF32) {
        CTR_REQUIRE(h, Wt != nullptr, "mlp_fwd: TF32 path needs the K-major copy Wt");
        return ctr_tc_mlp_fwd(h, ctr_st(s), B, K, N, X, ldx, Wt, bias, relu, keep, seed, Y, ldy, nullptr);
    }
    CTR_REQUIRE(h, precision == CTR_PREC_FP32 && W, "mlp_fwd: bad precision / W null");
    Epilogue ep{bias, relu, keep, seed, nullptr, 0, 1.f};
    return launch_sgemm(h, ctr_st(s), false, false, B, N, K, X, ldx, W, N, Y, ldy, 1, ep);
}

extern "C" int ctr_mlp_bwd_data(ctr_handle* h, ctr_stream s, int precision, int B, int K, int N, const float* dZ,
                                int64_t lddz, const float* W, const float* mask, int64_t ldmask, float mask_scale,
                                float* dX, int64_t lddx) {
    CTR_REQUIRE(h, dZ && W && dX && B >= 0 && K > 0 && N > 0 && lddz >= N && lddx >= K, "mlp_bwd_data: bad args");
    if (B == 0) return CTR_OK;
    if (precision == CTR_PREC_TF32)
        return ctr_tc_mlp_bwd_data(h, ctr_st(s), B, K, N, dZ, lddz, W, mask, ldmask, mask_scale, dX, lddx, nullptr);
    CTR_REQUIRE(h, precision == CTR_PREC_FP32, "mlp_bwd_data: bad precision");
    // dX[B,K] = dZ[B,N] @ W[K,N]^T : "B" operand stored [K_out, N_red] = TB layout
    Epilogue ep{nullptr, 0, 1.f, 0, mask, ldmask, mask_scale};
    return launch_sgemm(h, ctr_st(s), false, true, B, K, N, dZ, lddz, W, N, dX, lddx, 1, ep);
}

// ---- TF32 only: ReLU sign bits instead of the fp32 activation as the data gradient's mask (include/ctr_b200.h)
extern "C" int64_t ctr_mlp_relu_bits_bytes(ctr_handle* h, int B, int N) {
    if (!h || B <= 0 || N <= 0) return 0;
    return ctr_tc_relu_bits_bytes(B, N);
}

extern "C" int ctr_mlp_fwd_bits(ctr_handle* h, ctr_stream s, int B, int K, int N, const float* X, int64_t ldx, const float* Wt,
                                const float* bias, float keep, uint64_t seed, float* Y, int64_t ldy, uint32_t* relu_bits) {
    CTR_REQUIRE(h, X && Y && Wt && bias && relu_bits && B >= 0 && K > 0 && N > 0 && ldx >= K && ldy >= N, "mlp_fwd_bits: bad args");
    CTR_REQUIRE(h, keep > 0.f && keep <= 1.f, "mlp_fwd_bits: keep must be in (0,1]");
    if (B == 0) return CTR_OK;
    return ctr_tc_mlp_fwd(h, ctr_st(s), B, K, N, X, ldx, Wt, bias, 1, keep, seed, Y, ldy, relu_bits);
}

extern "C" int ctr_mlp_bwd_data_bits(ctr_handle* h, ctr_stream s, int B, int K, int N, const float* dZ, int64_t lddz,
                                     const float* W, const uint32_t* relu_bits, float mask_scale, float* dX, int64_t lddx) {
    CTR_REQUIRE(h, dZ && W && dX && relu_bits && B >= 0 && K > 0 && N > 0 && lddz >= N && lddx >= K, "mlp_bwd_data_bits: bad args");
    if (B == 0) return CTR_OK;
    return ctr_tc_mlp_bwd_data(h, ctr_st(s), B, K, N, dZ, lddz, W, nullptr, 0, mask_scale, dX, lddx, relu_bits);
}

extern "C" int64_t ctr_mlp_bwd_weight_ws_bytes(ctr_handle* h, int precision, int B, int K, int N) {
    if (!h) return 0;
    int64_t cs = ctr_colsum_ws_bytes(h, B, N);
    if (precision == CTR_PREC_TF32) return ctr_tc_bwd_weight_ws_bytes(h, B, K, N) + cs;
    return ctr_align((int64_t) wgrad_splits(h, B, K, N) * K * N * 4, 256) + cs;
}

extern "C" int ctr_mlp_bwd_weight(ctr_handle* h, ctr_stream s, int precision, int B, int K, int N, const float* X,
                                  int64_t ldx, const float* dZ, int64_t lddz, float* dW, float* dbias, void* ws,
                                  int64_t ws_bytes, int x_ones_col) {
    CTR_REQUIRE(h, X && dZ && dW && ws && B > 0 && K > 0 && N > 0 && ldx >= K && lddz >= N, "mlp_bwd_weight: bad args");
    if (ws_bytes < ctr_mlp_bwd_weight_ws_bytes(h, precision, B, K, N))
        return ctr_fail(h, CTR_E_WORKSPACE, "mlp_bwd_weight: ws too small");
    cudaStream_t st = ctr_st(s);
    char* p = (char*) ws;
    int64_t gemm_ws;
    int rc;
    if (precision == CTR_PREC_TF32) {
        gemm_ws = ctr_tc_bwd_weight_ws_bytes(h, B, K, N);
        // X[:, K] == 1 (caller's promise): the bias gradient is the extra row of the same tensor-core GEMM
        const bool fused_bias = dbias && x_ones_col && ldx > K;
        rc = ctr_tc_mlp_bwd_weight(h, st, B, K, N, X, ldx, dZ, lddz, dW, fused_bias ? dbias : nullptr, p, gemm_ws);
        if (rc) return rc;
        if (fused_bias) return CTR_OK;
    } else {
        CTR_REQUIRE(h, precision == CTR_PREC_FP32, "mlp_bwd_weight: bad precision");
        const int splits = wgrad_splits(h, B, K, N);
        gemm_ws = ctr_align((int64_t) splits * K * N * 4, 256);
        Epilogue ep{nullptr, 0, 1.f, 0, nullptr, 0, 1.f};
        // dW[K,N] = X[B,K]^T @ dZ[B,N] : A stored [red, M] = TA layout, reduction over B
        if (splits == 1) {
            rc = launch_sgemm(h, st, true, false, K, N, B, X, ldx, dZ, lddz, dW, N, 1, ep);
            if (rc) return rc;
        } else {
            rc = launch_sgemm(h, st, true, false, K, N, B, X, ldx, dZ, lddz, (float*) p, N, splits, ep);
            if (rc) return rc;
            int64_t n = (int64_t) K * N;
            splitk_reduce_kernel<<<(unsigned) ctr_cdiv(n, 256), 256, 0, st>>>((const float*) p, splits, n, dW);
            CTR_CHECK_LAUNCH(h, "splitk_reduce");
        }
    }
    if (dbias) return ctr_colsum(h, s, dZ, lddz, B, N, nullptr, dbias, p + gemm_ws, ws_bytes - gemm_ws);
    return CTR_OK;
}
