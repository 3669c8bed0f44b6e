// Cross network of DCN (reference: test_model/DCN.py:73-88).
//
//   x_{l+1} = x0 * (x_l . w_l) + b_l + x_l          l = 0 .. n_layers-1,   x_0 = x0 = the dense feature row
//
// (the reference writes it as tensordot(matmul(x0, x_l^T), w_l, 1): [B,D,1] x [B,1,D] -> [B,D,D], contracted with w_l
// over the last axis = x0_i * sum_j x_l_j w_j; the D x D outer product is never formed here).
// A row is tens to ~100 floats, so this is per-sample register work: one warp per sample, lane l owns columns
// l, l+32, ...; the only cross-lane step is the dot product.  HBM-bound (one read of x0, one write of x_L).
// Backward recomputes the x_l of a sample in registers (cheaper than storing n_layers rows per sample), walks the
// layers in reverse and accumulates dw_l, db_l per lane over the warp's samples; block partials go to a workspace and a
// second pass sums them in block order: deterministic.
#include "ctr_common.cuh"

namespace {

constexpr int CROSS_MAXT = 4;       // columns per lane: D <= 128 (the head kernel's n_fm limit as well)
constexpr int CROSS_MAXL = 4;       // cross layers (the reference uses 2, test_model/DCN.py:28)
constexpr int CROSS_WARPS = 8;

struct CrossPtrs {
    const float* w[CROSS_MAXL];
    const float* b[CROSS_MAXL];
};

__global__ void __launch_bounds__(CROSS_WARPS * 32)
cross_fwd_kernel(int B, int D, int L, const float* __restrict__ x0, int64_t ldx, const CrossPtrs P, float* __restrict__ out,
                 int64_t ldo) {
    const int lane = threadIdx.x & 31;
    const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
    float w[CROSS_MAXL][CROSS_MAXT], bb[CROSS_MAXL][CROSS_MAXT];
#pragma unroll
    for (int l = 0; l < CROSS_MAXL; ++l)
#pragma unroll
        for (int t = 0; t < CROSS_MAXT; ++t) {
            const int j = lane + 32 * t;
            const bool ok = l < L && j < D;
            w[l][t] = ok ? P.w[l][j] : 0.f;
            bb[l][t] = ok ? P.b[l][j] : 0.f;
        }
    for (int64_t b = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < B; b += nw) {
        float x0v[CROSS_MAXT], x[CROSS_MAXT];
#pragma unroll
        for (int t = 0; t < CROSS_MAXT; ++t) {
            const int j = lane + 32 * t;
            x0v[t] = j < D ? x0[b * ldx + j] : 0.f;
            x[t] = x0v[t];
        }
#pragma unroll
        for (int l = 0; l < CROSS_MAXL; ++l) {
            if (l < L) {
                float s = 0.f;
#pragma unroll
                for (int t = 0; t < CROSS_MAXT; ++t) s = fmaf(x[t], w[l][t], s);
                s = ctr_warp_sum(s);
#pragma unroll
                for (int t = 0; t < CROSS_MAXT; ++t) x[t] = fmaf(x0v[t], s, bb[l][t]) + x[t];
            }
        }
#pragma unroll
        for (int t = 0; t < CROSS_MAXT; ++t) {
            const int j = lane + 32 * t;
            if (j < D) out[b * ldo + j] = x[t];
        }
    }
}

// gradient of the cross network's parameters.  g_out[b, :] = dlogit[b] * w_proj[:]  (the head's projection of x_L).
// For l = L-1 .. 0 with g = dL/dx_{l+1}:   t = g . x0 ;  db_l += g ;  dw_l += t * x_l ;  dL/dx_l = g + t * w_l
// (dL/dx0 is not needed: the features are inputs).  ws: [blocks][L][2][D] partial sums.
__global__ void __launch_bounds__(CROSS_WARPS * 32)
cross_bwd_pass1(int B, int D, int L, const float* __restrict__ x0, int64_t ldx, const CrossPtrs P,
                const float* __restrict__ dlogit, const float* __restrict__ w_proj, float* __restrict__ ws) {
    __shared__ float sm[CROSS_WARPS][CROSS_MAXL * 2 * 32 * CROSS_MAXT];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t nw = ((int64_t) gridDim.x * blockDim.x) >> 5;
    float w[CROSS_MAXL][CROSS_MAXT], bb[CROSS_MAXL][CROSS_MAXT], wp[CROSS_MAXT];
    float gw[CROSS_MAXL][CROSS_MAXT], gb[CROSS_MAXL][CROSS_MAXT];
#pragma unroll
    for (int l = 0; l < CROSS_MAXL; ++l)
#pragma unroll
        for (int t = 0; t < CROSS_MAXT; ++t) {
            const int j = lane + 32 * t;
            const bool ok = l < L && j < D;
            w[l][t] = ok ? P.w[l][j] : 0.f;
            bb[l][t] = ok ? P.b[l][j] : 0.f;
            gw[l][t] = 0.f;
            gb[l][t] = 0.f;
        }
#pragma unroll
    for (int t = 0; t < CROSS_MAXT; ++t) wp[t] = lane + 32 * t < D ? w_proj[lane + 32 * t] : 0.f;
    // fixed sample -> warp assignment and a fixed order inside the warp: deterministic sums
    for (int64_t b = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < B; b += nw) {
        float x0v[CROSS_MAXT], xs[CROSS_MAXL][CROSS_MAXT], x[CROSS_MAXT];
#pragma unroll
        for (int t = 0; t < CROSS_MAXT; ++t) {
            const int j = lane + 32 * t;
            x0v[t] = j < D ? x0[b * ldx + j] : 0.f;
            x[t] = x0v[t];
        }
#pragma unroll
        for (int l = 0; l < CROSS_MAXL; ++l) {            // forward again, keeping x_l
            if (l < L) {
                float s = 0.f;
#pragma unroll
                for (int t = 0; t < CROSS_MAXT; ++t) {
                    xs[l][t] = x[t];
                    s = fmaf(x[t], w[l][t], s);
                }
                s = ctr_warp_sum(s);
#pragma unroll
                for (int t = 0; t < CROSS_MAXT; ++t) x[t] = fmaf(x0v[t], s, bb[l][t]) + x[t];
            }
        }
        const float dl = dlogit[b];
        float g[CROSS_MAXT];
#pragma unroll
        for (int t = 0; t < CROSS_MAXT; ++t) g[t] = dl * wp[t];
#pragma unroll
        for (int l = CROSS_MAXL - 1; l >= 0; --l) {
            if (l < L) {
                float tt = 0.f;
#pragma unroll
                for (int t = 0; t < CROSS_MAXT; ++t) tt = fmaf(g[t], x0v[t], tt);
                tt = ctr_warp_sum(tt);
#pragma unroll
                for (int t = 0; t < CROSS_MAXT; ++t) {
                    gb[l][t] += g[t];
                    gw[l][t] = fmaf(tt, xs[l][t], gw[l][t]);
                    g[t] = fmaf(tt, w[l][t], g[t]);
                }
            }
        }
    }
    // warps of the block in fixed order
#pragma unroll
    for (int l = 0; l < CROSS_MAXL; ++l)
#pragma unroll
        for (int t = 0; t < CROSS_MAXT; ++t) {
            sm[wid][(l * 2 + 0) * 32 * CROSS_MAXT + lane + 32 * t] = gw[l][t];
            sm[wid][(l * 2 + 1) * 32 * CROSS_MAXT + lane + 32 * t] = gb[l][t];
        }
    __syncthreads();
    const int per = L * 2 * D;
    for (int i = threadIdx.x; i < per; i += blockDim.x) {
        const int l2 = i / D, j = i - l2 * D;              // l2 = l * 2 + (0: w, 1: b)
        float a = 0.f;
#pragma unroll
        for (int ww = 0; ww < CROSS_WARPS; ++ww) a += sm[ww][l2 * 32 * CROSS_MAXT + j];
        ws[(int64_t) blockIdx.x * per + i] = a;
    }
}

struct CrossGrads {
    float* gw[CROSS_MAXL];
    float* gb[CROSS_MAXL];
};

__global__ void cross_bwd_pass2(const float* __restrict__ ws, int nblk, int D, int L, const CrossGrads G) {
    const int per = L * 2 * D;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= per) return;
    float a = 0.f;
    for (int k = 0; k < nblk; ++k) a += ws[(int64_t) k * per + i];      // block order: deterministic
    const int l2 = i / D, j = i - l2 * D;
    ((l2 & 1) ? G.gb[l2 >> 1] : G.gw[l2 >> 1])[j] = a;
}

int cross_blocks(ctr_handle* h, int B) {
    int64_t want = ctr_cdiv((int64_t) B, CROSS_WARPS * 4);      // >= 4 samples per warp
    int64_t cap = (int64_t) h->sm_count * 4;
    if (want < 1) want = 1;
    return (int) (want < cap ? want : cap);
}

int check_cross(ctr_handle* h, const ctr_cross_cfg* c) {
    CTR_REQUIRE(h, c != nullptr, "cross cfg is null");
    CTR_REQUIRE(h, c->B >= 0 && c->D >= 1 && c->D <= 32 * CROSS_MAXT, "cross: D must be in [1, %d] (got %d)", 32 * CROSS_MAXT, c->D);
    CTR_REQUIRE(h, c->n_layers >= 1 && c->n_layers <= CROSS_MAXL, "cross: n_layers must be in [1, %d] (got %d)", CROSS_MAXL, c->n_layers);
    CTR_REQUIRE(h, c->ldx >= c->D, "cross: ldx < D");
    for (int l = 0; l < c->n_layers; ++l) CTR_REQUIRE(h, c->w[l] && c->b[l], "cross: null parameters of layer %d", l);
    return CTR_OK;
}

}  // namespace

extern "C" int64_t ctr_cross_bwd_ws_bytes(ctr_handle* h, const ctr_cross_cfg* cfg) {
    if (!h || !cfg) return 0;
    return (int64_t) cross_blocks(h, cfg->B) * cfg->n_layers * 2 * cfg->D * 4 + 256;
}

extern "C" int ctr_cross_fwd(ctr_handle* h, ctr_stream s, const ctr_cross_cfg* cfg, const float* x0, float* out, int64_t ldo) {
    int rc = check_cross(h, cfg);
    if (rc) return rc;
    CTR_REQUIRE(h, x0 && out && ldo >= cfg->D, "cross_fwd: bad args");
    if (cfg->B == 0) return CTR_OK;
    CrossPtrs P;
    for (int l = 0; l < CROSS_MAXL; ++l) {
        P.w[l] = l < cfg->n_layers ? cfg->w[l] : nullptr;
        P.b[l] = l < cfg->n_layers ? cfg->b[l] : nullptr;
    }
    cross_fwd_kernel<<<cross_blocks(h, cfg->B), CROSS_WARPS * 32, 0, ctr_st(s)>>>(cfg->B, cfg->D, cfg->n_layers, x0, cfg->ldx, P, out, ldo);
    CTR_CHECK_LAUNCH(h, "cross_fwd");
    return CTR_OK;
}

extern "C" int ctr_cross_bwd(ctr_handle* h, ctr_stream s, const ctr_cross_cfg* cfg, const float* x0, const float* dlogit,
                             const float* w_proj, void* ws, int64_t ws_bytes) {
    int rc = check_cross(h, cfg);
    if (rc) return rc;
    CTR_REQUIRE(h, x0 && dlogit && w_proj && ws, "cross_bwd: null pointer");
    for (int l = 0; l < cfg->n_layers; ++l) CTR_REQUIRE(h, cfg->gw[l] && cfg->gb[l], "cross_bwd: null gradient buffers of layer %d", l);
    if (ws_bytes < ctr_cross_bwd_ws_bytes(h, cfg)) return ctr_fail(h, CTR_E_WORKSPACE, "cross_bwd: ws too small");
    CrossPtrs P;
    CrossGrads G;
    for (int l = 0; l < CROSS_MAXL; ++l) {
        const bool on = l < cfg->n_layers;
        P.w[l] = on ? cfg->w[l] : nullptr;
        P.b[l] = on ? cfg->b[l] : nullptr;
        G.gw[l] = on ? cfg->gw[l] : nullptr;
        G.gb[l] = on ? cfg->gb[l] : nullptr;
    }
    const int nblk = cross_blocks(h, cfg->B);
    cross_bwd_pass1<<<nblk, CROSS_WARPS * 32, 0, ctr_st(s)>>>(cfg->B, cfg->D, cfg->n_layers, x0, cfg->ldx, P, dlogit, w_proj, (float*) ws);
    CTR_CHECK_LAUNCH(h, "cross_bwd pass1");
    const int per = cfg->n_layers * 2 * cfg->D;
    cross_bwd_pass2<<<(unsigned) ctr_cdiv(per, 128), 128, 0, ctr_st(s)>>>((const float*) ws, nblk, cfg->D, cfg->n_layers, G);
    CTR_CHECK_LAUNCH(h, "cross_bwd pass2");
    return CTR_OK;
}
