// Handle management and small layout utilities of libctr_b200.so.
#include "ctr_common.cuh"

extern "C" int ctr_abi_version(void) { return CTR_ABI_VERSION; }

// SHA-256 of the sources this library was built from (deep_learning_b200/build.py passes it in; the marker in front lets
// the build script read it out of the file without loading it).  The loader compares it with the sources next to it: a
// stale binary with fresh timestamps does not go unnoticed.
#ifndef CTR_SOURCE_HASH
#define CTR_SOURCE_HASH "unknown"
#endif
static const char g_source_hash[] = "CTR_SOURCE_HASH=" CTR_SOURCE_HASH;
extern "C" const char* ctr_source_hash(void) { return g_source_hash + 16; }
extern "C" int ctr_sizeof_embed_cfg(void) { return (int) sizeof(ctr_embed_cfg); }

// Scratch of ctr_reduce_static: the block partials of one reduction, in device memory that exists from the moment the
// library is loaded (every GPU that loads it holds these 4 KB until the process exits) -- so that entry point never
// allocates.  NB in HOST code `g_reduce_scratch` is only a stand-in for the symbol, not a device address: the address a
// kernel can use comes from cudaGetSymbolAddress, taken per device with that device current (resolve_reduce_scratch).
static __device__ __align__(256) float g_reduce_scratch[CTR_REDUCE_STATIC_MAX_PARTIALS];

// once per handle (= per device), at set-up: the handle's device is made current for the lookup and the caller's
// current device is restored.  No allocation: the array is part of the loaded module.
static void resolve_reduce_scratch(ctr_handle* h) {
    int cur = -1;
    cudaError_t e = cudaGetDevice(&cur);
    if (e == cudaSuccess && cur != h->device) e = cudaSetDevice(h->device);
    void* p = nullptr;
    if (e == cudaSuccess) e = cudaGetSymbolAddress(&p, g_reduce_scratch);
    if (cur >= 0 && cur != h->device) cudaSetDevice(cur);
    h->reduce_scratch = e == cudaSuccess ? static_cast<float*>(p) : nullptr;
    h->reduce_scratch_rc = (int) e;
    if (e != cudaSuccess) cudaGetLastError();      // a failed lookup must not poison the next launch check
}

extern "C" int ctr_create(int device, ctr_handle** out) {
    if (!out) return CTR_E_INVALID;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return CTR_E_CUDA;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return CTR_E_CUDA;
    ctr_handle* h = new ctr_handle();
    h->device = device;
    h->sm_count = prop.multiProcessorCount;
    h->cc_major = prop.major;
    h->cc_minor = prop.minor;
    h->err[0] = 0;
    h->launches = 0;
    resolve_reduce_scratch(h);     // set-up for ctr_reduce_static only; a failure is reported by that entry point, not here
    *out = h;
    return CTR_OK;
}

extern "C" int ctr_destroy(ctr_handle* h) {
    delete h;
    return CTR_OK;
}

extern "C" const char* ctr_last_error(ctr_handle* h) { return h ? h->err : "null handle"; }
extern "C" int ctr_sm_count(ctr_handle* h) { return h ? h->sm_count : 0; }
extern "C" int64_t ctr_launch_count(ctr_handle* h) { return h ? (int64_t) h->launches : 0; }
extern "C" int ctr_has_tcgen05(ctr_handle* h) { return h && h->cc_major == 10 ? 1 : 0; }

// ------------------------------------------------------------------ table pack / unpack
__global__ void table_pack_kernel(const float* __restrict__ emb, const float* __restrict__ first, int64_t R,
                                  int k, int stride, float* __restrict__ table) {
    int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    int64_t total = R * stride;
    if (i >= total) return;
    int64_t r = i / stride;
    int j = (int) (i - r * stride);
    float v = 0.f;
    if (j < k) v = emb[r * k + j];
    else if (j == k && first) v = first[r];
    table[i] = v;
}

__global__ void table_unpack_kernel(const float* __restrict__ table, int64_t R, int k, int stride,
                                    float* __restrict__ emb, float* __restrict__ first) {
    int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    int64_t total = R * stride;
    if (i >= total) return;
    int64_t r = i / stride;
    int j = (int) (i - r * stride);
    if (j < k) emb[r * k + j] = table[i];
    else if (j == k && first) first[r] = table[i];
}

extern "C" int ctr_table_pack(ctr_handle* h, ctr_stream s, const float* emb, const float* first, int64_t R, int k,
                              int table_stride, float* table) {
    CTR_REQUIRE(h, emb && table && R > 0 && k > 0 && table_stride >= k + (first ? 1 : 0), "table_pack: bad args");
    int64_t total = R * table_stride;
    table_pack_kernel<<<(unsigned) ctr_cdiv(total, 256), 256, 0, ctr_st(s)>>>(emb, first, R, k, table_stride, table);
    CTR_CHECK_LAUNCH(h, "table_pack");
    return CTR_OK;
}

extern "C" int ctr_table_unpack(ctr_handle* h, ctr_stream s, const float* table, int64_t R, int k, int table_stride,
                                float* emb, float* first) {
    CTR_REQUIRE(h, emb && table && R > 0 && k > 0 && table_stride >= k + (first ? 1 : 0), "table_unpack: bad args");
    int64_t total = R * table_stride;
    table_unpack_kernel<<<(unsigned) ctr_cdiv(total, 256), 256, 0, ctr_st(s)>>>(table, R, k, table_stride, emb, first);
    CTR_CHECK_LAUNCH(h, "table_unpack");
    return CTR_OK;
}

// ------------------------------------------------------------------ transpose [K,N] -> [N,K]
__global__ void transpose_kernel(const float* __restrict__ in, int K, int N, float* __restrict__ out, int64_t ld_out, int rnd) {
    __shared__ float tile[32][33];
    int n0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        int kk = k0 + i, nn = n0 + threadIdx.x;
        tile[i][threadIdx.x] = (kk < K && nn < N) ? in[(int64_t) kk * N + nn] : 0.f;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        int nn = n0 + i, kk = k0 + threadIdx.x;
        if (nn < N && kk < K) out[(int64_t) nn * ld_out + kk] = rnd ? ctr_round_tf32(tile[threadIdx.x][i]) : tile[threadIdx.x][i];
    }
}

__global__ void round_tf32_kernel(const float* __restrict__ in, float* __restrict__ out, int64_t n) {
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x)
        out[i] = ctr_round_tf32(in[i]);
}

extern "C" int ctr_round_tf32_copy(ctr_handle* h, ctr_stream s, const float* in, float* out, int64_t n) {
    CTR_REQUIRE(h, in && out && n >= 0, "round_tf32_copy: bad args");
    if (n == 0) return CTR_OK;
    int64_t blocks = ctr_cdiv(n, 256);
    if (blocks > (int64_t) h->sm_count * 8) blocks = (int64_t) h->sm_count * 8;
    round_tf32_kernel<<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(in, out, n);
    CTR_CHECK_LAUNCH(h, "round_tf32_copy");
    return CTR_OK;
}

extern "C" int ctr_transpose(ctr_handle* h, ctr_stream s, const float* in, int K, int N, float* out, int64_t ld_out,
                             int round_tf32) {
    CTR_REQUIRE(h, in && out && K > 0 && N > 0 && ld_out >= K, "transpose: bad args");
    dim3 grid((unsigned) ctr_cdiv(N, 32), (unsigned) ctr_cdiv(K, 32)), block(32, 8);
    transpose_kernel<<<grid, block, 0, ctr_st(s)>>>(in, K, N, out, ld_out, round_tf32);
    CTR_CHECK_LAUNCH(h, "transpose");
    return CTR_OK;
}

// ------------------------------------------------------------------ row gather / perm inversion
__global__ void gather_rows_kernel(const float* __restrict__ src, int src_stride, const uint32_t* __restrict__ idx,
                                   const uint32_t* __restrict__ n_rows, int64_t n_max, uint32_t mask, int width,
                                   float* __restrict__ dst, int dst_stride) {
    int64_t n = n_rows ? (int64_t) *n_rows : n_max;
    if (n > n_max) n = n_max;
    int64_t total = n * width;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
        int64_t r = i / width;
        int j = (int) (i - r * width);
        uint32_t key = idx[r];
        float v = 0.f;
        if (key != CTR_PAD_KEY) v = src[(int64_t) (key & mask) * src_stride + j];
        dst[r * dst_stride + j] = v;
    }
}

extern "C" int ctr_gather_rows(ctr_handle* h, ctr_stream s, const float* src, int src_stride, const uint32_t* idx,
                               const uint32_t* n_rows, int64_t n_max, int idx_mask_bits, int width, float* dst,
                               int dst_stride) {
    CTR_REQUIRE(h, src && idx && dst && width > 0 && n_max >= 0, "gather_rows: bad args");
    if (n_max == 0) return CTR_OK;
    uint32_t mask = (idx_mask_bits <= 0 || idx_mask_bits >= 32) ? 0xFFFFFFFFu : ((1u << idx_mask_bits) - 1u);
    int64_t total = n_max * width;
    int64_t blocks = ctr_cdiv(total, 256);
    int64_t cap = (int64_t) h->sm_count * 32;
    if (blocks > cap) blocks = cap;
    gather_rows_kernel<<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(src, src_stride, idx, n_rows, n_max, mask, width, dst,
                                                                 dst_stride);
    CTR_CHECK_LAUNCH(h, "gather_rows");
    return CTR_OK;
}

__global__ void invert_perm_kernel(const uint32_t* __restrict__ perm, const uint32_t* __restrict__ uidx, int64_t n,
                                   uint32_t* __restrict__ inv) {
    int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) inv[perm[i]] = uidx[i];
}

extern "C" int ctr_invert_perm(ctr_handle* h, ctr_stream s, const uint32_t* perm, const uint32_t* uidx, int64_t n,
                               uint32_t* inv) {
    CTR_REQUIRE(h, perm && uidx && inv && n >= 0, "invert_perm: bad args");
    if (n == 0) return CTR_OK;
    invert_perm_kernel<<<(unsigned) ctr_cdiv(n, 256), 256, 0, ctr_st(s)>>>(perm, uidx, n, inv);
    CTR_CHECK_LAUNCH(h, "invert_perm");
    return CTR_OK;
}

// ------------------------------------------------------------------ deterministic scalar reduction
template <int MODE>
__device__ __forceinline__ float red_f(float x) {
    if (MODE == 1) return x * x;
    if (MODE == 2) return fabsf(x);
    return x;
}

template <int MODE>
__global__ void reduce_pass1(const float* __restrict__ x, int64_t n, float* __restrict__ part) {
    __shared__ float sm[32];
    // fixed contiguous slice per block -> result independent of scheduling
    int64_t per = (n + gridDim.x - 1) / gridDim.x;
    int64_t lo = per * blockIdx.x, hi = lo + per < n ? lo + per : n;
    float acc = 0.f;
    for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) acc += red_f<MODE>(x[i]);
    float r = ctr_block_sum(acc, sm);
    if (threadIdx.x == 0) part[blockIdx.x] = r;
}

__global__ void reduce_pass2(const float* __restrict__ part, int nparts, float* __restrict__ out) {
    __shared__ float sm[32];
    float acc = 0.f;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) acc += part[i];
    float r = ctr_block_sum(acc, sm);
    if (threadIdx.x == 0) out[0] = r;
}

extern "C" int ctr_reduce(ctr_handle* h, ctr_stream s, const float* x, int64_t n, int mode, float* out, void* ws,
                          int64_t ws_bytes) {
    CTR_REQUIRE(h, x && out && ws && n >= 0 && mode >= 0 && mode <= 2, "reduce: bad args");
    int64_t blocks = ctr_cdiv(n > 0 ? n : 1, 4096);
    if (blocks > 1024) blocks = 1024;
    if (ws_bytes < (int64_t) (blocks * sizeof(float))) return ctr_fail(h, CTR_E_WORKSPACE, "reduce: ws too small");
    float* part = (float*) ws;
    if (mode == 0) reduce_pass1<0><<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(x, n, part);
    else if (mode == 1) reduce_pass1<1><<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(x, n, part);
    else reduce_pass1<2><<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(x, n, part);
    CTR_CHECK_LAUNCH(h, "reduce_pass1");
    reduce_pass2<<<1, 256, 0, ctr_st(s)>>>(part, (int) blocks, out);
    CTR_CHECK_LAUNCH(h, "reduce_pass2");
    return CTR_OK;
}

// ctr_reduce for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc): same two kernels, same block slices, so the result is bit-identical to ctr_reduce -- the block partials go
// to the file-scope static array instead of a caller workspace.  Calls no cudaMalloc / cudaFree / allocating routine.
// Nothing in the scratch has to start at zero: pass 1 overwrites partials [0, blocks) and pass 2 reads exactly those, so
// whatever the previous call left behind is never read.  ALL calls on a device share the one array: two calls that
// overlap on the GPU (different streams, no ordering between them) race; calls on one stream are ordered and safe.
extern "C" int ctr_reduce_static(ctr_handle* h, ctr_stream s, const float* x, int64_t n, int mode, float* out) {
    CTR_REQUIRE(h, x && out && n >= 0 && mode >= 0 && mode <= 2, "reduce_static: bad args");
    CTR_REQUIRE(h, n <= CTR_REDUCE_STATIC_MAX_N, "reduce_static: n = %lld exceeds the static scratch's limit of %lld elements; "
                "use ctr_reduce with a caller workspace", (long long) n, (long long) CTR_REDUCE_STATIC_MAX_N);
    if (!h->reduce_scratch)
        return ctr_fail(h, CTR_E_CUDA, "reduce_static: cudaGetSymbolAddress of the scratch failed on device %d at ctr_create: %s",
                        h->device, cudaGetErrorString((cudaError_t) h->reduce_scratch_rc));
    int64_t blocks = ctr_cdiv(n > 0 ? n : 1, 4096);
    if (blocks > CTR_REDUCE_STATIC_MAX_PARTIALS) blocks = CTR_REDUCE_STATIC_MAX_PARTIALS;
    float* part = h->reduce_scratch;
    if (mode == 0) reduce_pass1<0><<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(x, n, part);
    else if (mode == 1) reduce_pass1<1><<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(x, n, part);
    else reduce_pass1<2><<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(x, n, part);
    CTR_CHECK_LAUNCH(h, "reduce_static pass1");
    reduce_pass2<<<1, 256, 0, ctr_st(s)>>>(part, (int) blocks, out);
    CTR_CHECK_LAUNCH(h, "reduce_static pass2");
    return CTR_OK;
}
