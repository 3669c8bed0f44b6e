// Peer-memory exchange over NVLink 5 / NVSwitch: slab export/import (CUDA IPC), the all-ranks barrier and the fused
// owner-gather-and-push kernel (include/ctr_b200.h, "peer memory").  The fused gradient push lives in embed_bwd.cu
// next to the finalize kernel it extends.
#include "ctr_common.cuh"
#include "peer_route.cuh"

extern "C" int ctr_peer_alloc(ctr_handle* h, int64_t bytes, void** slab, void* ipc_handle) {
    CTR_REQUIRE(h, bytes > 0 && slab && ipc_handle, "peer_alloc: bad args");
    static_assert(sizeof(cudaIpcMemHandle_t) == CTR_IPC_HANDLE_BYTES, "IPC handle size");
    void* p = nullptr;
    CTR_CUDA(h, cudaMalloc(&p, (size_t) bytes));
    cudaError_t e = cudaMemset(p, 0, (size_t) bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaIpcMemHandle_t hd;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&hd, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return ctr_fail(h, CTR_E_CUDA, "peer_alloc: %s", cudaGetErrorString(e));
    }
    memcpy(ipc_handle, &hd, sizeof(hd));
    *slab = p;
    return CTR_OK;
}

extern "C" int ctr_peer_open(ctr_handle* h, const void* ipc_handle, void** slab) {
    CTR_REQUIRE(h, ipc_handle && slab, "peer_open: bad args");
    cudaIpcMemHandle_t hd;
    memcpy(&hd, ipc_handle, sizeof(hd));
    void* p = nullptr;
    CTR_CUDA(h, cudaIpcOpenMemHandle(&p, hd, cudaIpcMemLazyEnablePeerAccess));
    *slab = p;
    return CTR_OK;
}

extern "C" int ctr_peer_close(ctr_handle* h, void* slab) {
    if (slab) CTR_CUDA(h, cudaIpcCloseMemHandle(slab));
    return CTR_OK;
}

extern "C" int ctr_peer_free(ctr_handle* h, void* slab) {
    if (slab) CTR_CUDA(h, cudaFree(slab));
    return CTR_OK;
}

// ------------------------------------------------------------------------------------------------ barrier
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// lane p: tell peer p "rank arrived at `epoch`" (word [rank] of p's flags), then wait for p's word in the own flags.
// Stores of earlier kernels on this stream are complete at this kernel's start; the system-scope fence + release
// store order them before the arrival word as seen from the peer.
__global__ void __launch_bounds__(32) peer_barrier_kernel(const PeerRoute rt, int64_t flags_off, uint32_t epoch,
                                                          uint32_t* __restrict__ status, unsigned long long timeout_ns) {
    const int p = threadIdx.x;
    __threadfence_system();
    if (p < rt.world) {
        uint32_t* theirs = reinterpret_cast<uint32_t*>(static_cast<char*>(rt.slab[p]) + flags_off) + rt.rank;
        st_release_sys(theirs, epoch);
        const uint32_t* mine = reinterpret_cast<const uint32_t*>(static_cast<const char*>(rt.slab[rt.rank]) + flags_off) + p;
        const unsigned long long t0 = global_ns();
        unsigned spins = 0;
        while ((int32_t) (ld_acquire_sys(mine) - epoch) < 0) {
            if ((++spins & 1023u) == 0 && global_ns() - t0 > timeout_ns) {
                if (status) *status = 1u + (uint32_t) p;
                break;
            }
        }
    }
    __syncwarp();
    __threadfence_system();
}

extern "C" int ctr_peer_barrier(ctr_handle* h, ctr_stream s, const ctr_peer_route* rt, int64_t flags_off, uint32_t epoch,
                                uint32_t* status, int timeout_ms) {
    CTR_REQUIRE(h, rt && rt->world >= 1 && rt->world <= CTR_MAX_PEERS && rt->rank >= 0 && rt->rank < rt->world &&
                       flags_off >= 0 && flags_off % 4 == 0, "peer_barrier: bad args");
    if (timeout_ms <= 0) timeout_ms = 10000;
    peer_barrier_kernel<<<1, 32, 0, ctr_st(s)>>>(make_route(rt), flags_off, epoch, status,
                                                 (unsigned long long) timeout_ms * 1000000ull);
    CTR_CHECK_LAUNCH(h, "peer_barrier");
    return CTR_OK;
}

// ------------------------------------------------------------------------------------------------ gather + push
// one thread per (row, 4 columns): 128-bit table load (L1-bypassing: every row is wanted once), 128-bit store into the
// requester's staging buffer -- consecutive rows of a segment are consecutive at the destination, so the NVLink
// stores of a warp are contiguous
constexpr int GP_UN = 4;       // rows in flight per thread: the id -> row -> NVLink store chain is latency, not bandwidth
__global__ void __launch_bounds__(256)
gather_rows_push_kernel(const float* __restrict__ table, int table_stride, const uint32_t* __restrict__ idx, int64_t n,
                        uint32_t mask, int width, int ds, const PeerRoute rt) {
    const int nq = ds >> 2;
    n = route_total(rt, n);
    const int64_t total = n * nq;
    const int64_t nthreads = (int64_t) gridDim.x * blockDim.x;
    for (int64_t t0 = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t0 < total; t0 += nthreads * GP_UN) {
        float4 v[GP_UN];
        int64_t ii[GP_UN];
        int cc[GP_UN];
        uint32_t key[GP_UN];
#pragma unroll
        for (int k = 0; k < GP_UN; ++k) {        // item k of this thread is nthreads apart: a warp still reads consecutive items
            const int64_t t = t0 + k * nthreads;
            ii[k] = t < total ? t / nq : -1;
            cc[k] = t < total ? (int) (t - ii[k] * nq) * 4 : 0;
            key[k] = ii[k] >= 0 ? __ldg(idx + ii[k]) : CTR_PAD_KEY;
        }
#pragma unroll
        for (int k = 0; k < GP_UN; ++k) {
            v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            const int c = cc[k];
            if (key[k] != CTR_PAD_KEY && c < width) {
                const float* src = table + (int64_t) (key[k] & mask) * table_stride + c;
                if (c + 4 <= table_stride) {
                    v[k] = ctr_ldg_f4_stream(src);
                } else {
                    v[k].x = src[0];
                    if (c + 1 < table_stride) v[k].y = src[1];
                    if (c + 2 < table_stride) v[k].z = src[2];
                }
                if (c + 1 >= width) v[k].y = 0.f;
                if (c + 2 >= width) v[k].z = 0.f;
                if (c + 3 >= width) v[k].w = 0.f;
            }
        }
#pragma unroll
        for (int k = 0; k < GP_UN; ++k) {
            if (ii[k] >= 0) {
                float* dst = route_row(rt, (uint32_t) ii[k], ds);
                *reinterpret_cast<float4*>(dst + cc[k]) = v[k];
            }
        }
    }
}

extern "C" int ctr_gather_rows_push(ctr_handle* h, ctr_stream s, const float* table, int table_stride, const uint32_t* idx,
                                    int64_t n_rows, int idx_mask_bits, int width, int dst_stride, const ctr_peer_route* rt) {
    CTR_REQUIRE(h, table && idx && rt && width > 0 && dst_stride >= width && n_rows >= 0 && n_rows < (1ll << 32),
                "gather_rows_push: bad args");
    CTR_REQUIRE(h, rt->world >= 1 && rt->world <= CTR_MAX_PEERS && dst_stride % 4 == 0 && table_stride % 4 == 0 &&
                       rt->dst_off % 16 == 0 && (reinterpret_cast<uintptr_t>(table) & 15) == 0,
                "gather_rows_push: strides must be multiples of 4 floats, buffers 16-byte aligned");
    CTR_REQUIRE(h, rt->dyn || (int64_t) rt->seg_end[rt->world - 1] == n_rows, "gather_rows_push: route does not cover the request list");
    if (n_rows == 0) return CTR_OK;      // (with a device route block n_rows is the capacity of the request list)
    uint32_t mask = (idx_mask_bits <= 0 || idx_mask_bits >= 32) ? 0xFFFFFFFFu : ((1u << idx_mask_bits) - 1u);
    int64_t blocks = ctr_cdiv(n_rows * (dst_stride / 4), 256 * GP_UN);
    int64_t cap = (int64_t) h->sm_count * 8;
    if (blocks > cap) blocks = cap;
    gather_rows_push_kernel<<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(table, table_stride, idx, n_rows, mask, width,
                                                                      dst_stride, make_route(rt));
    CTR_CHECK_LAUNCH(h, "gather_rows_push");
    return CTR_OK;
}

// ------------------------------------------------------------------------------------------------ plain pushes + reduction
// rows of 32-bit words -> their routed position in the owning peer's buffer (the id-plan key exchange: width 1)
__global__ void __launch_bounds__(256)
push_rows_kernel(const uint32_t* __restrict__ src, int64_t n, int width, const PeerRoute rt) {
    n = route_total(rt, n);
    const int64_t total = n * width;
    for (int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t) gridDim.x * blockDim.x) {
        const int64_t i = t / width;
        const int c = (int) (t - i * width);
        reinterpret_cast<uint32_t*>(route_row(rt, (uint32_t) i, width))[c] = src[t];
    }
}

// the same n words into EVERY peer's slab at the route's byte offset (the per-owner unique counts of the id plan)
__global__ void __launch_bounds__(256)
bcast_words_kernel(const uint32_t* __restrict__ src, int n, const PeerRoute rt) {
    const int total = n * rt.world;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gridDim.x * blockDim.x) {
        const int p = t / n, i = t - p * n;
        reinterpret_cast<uint32_t*>(rt.buf[p])[i] = src[i];
    }
}

// out[i] = sum over ranks, in rank order, of the buffer at the route's byte offset in every slab: the data-parallel sum of
// the dense gradients read straight out of peer memory.  Same order on every rank => bitwise identical replicas.
__global__ void __launch_bounds__(256)
peer_reduce_sum_kernel(const PeerRoute rt, int64_t n4, float4* __restrict__ out) {
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x) {
        float4 v[CTR_MAX_PEERS];
#pragma unroll
        for (int p = 0; p < CTR_MAX_PEERS; ++p)
            if (p < rt.world) v[p] = *(reinterpret_cast<const float4*>(rt.buf[p]) + i);      // all loads in flight together
        float4 a = v[0];
#pragma unroll
        for (int p = 1; p < CTR_MAX_PEERS; ++p)
            if (p < rt.world) { a.x += v[p].x; a.y += v[p].y; a.z += v[p].z; a.w += v[p].w; }
        out[i] = a;
    }
}

static inline bool route_ok(const ctr_peer_route* rt) {
    return rt && rt->world >= 1 && rt->world <= CTR_MAX_PEERS && rt->rank >= 0 && rt->rank < rt->world && rt->dst_off >= 0;
}

extern "C" int ctr_peer_push_rows(ctr_handle* h, ctr_stream s, const uint32_t* src, int64_t n_rows, int width,
                                  const ctr_peer_route* rt) {
    CTR_REQUIRE(h, src && route_ok(rt) && width > 0 && n_rows >= 0 && n_rows < (1ll << 32) && rt->dst_off % 4 == 0,
                "peer_push_rows: bad args");
    CTR_REQUIRE(h, rt->dyn || (int64_t) rt->seg_end[rt->world - 1] == n_rows, "peer_push_rows: route does not cover the list");
    if (n_rows == 0) return CTR_OK;
    int64_t blocks = ctr_cdiv(n_rows * width, 256);
    const int64_t cap = (int64_t) h->sm_count * 4;
    if (blocks > cap) blocks = cap;
    push_rows_kernel<<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(src, n_rows, width, make_route(rt));
    CTR_CHECK_LAUNCH(h, "peer_push_rows");
    return CTR_OK;
}

extern "C" int ctr_peer_bcast(ctr_handle* h, ctr_stream s, const uint32_t* src, int n_words, const ctr_peer_route* rt) {
    CTR_REQUIRE(h, src && route_ok(rt) && n_words > 0 && n_words <= (1 << 20) && rt->dst_off % 4 == 0, "peer_bcast: bad args");
    bcast_words_kernel<<<(unsigned) ctr_cdiv((int64_t) n_words * rt->world, 256), 256, 0, ctr_st(s)>>>(src, n_words, make_route(rt));
    CTR_CHECK_LAUNCH(h, "peer_bcast");
    return CTR_OK;
}

extern "C" int ctr_peer_reduce_sum(ctr_handle* h, ctr_stream s, const ctr_peer_route* rt, int64_t n, float* out) {
    CTR_REQUIRE(h, out && route_ok(rt) && n >= 0 && n % 4 == 0 && rt->dst_off % 16 == 0 &&
                       (reinterpret_cast<uintptr_t>(out) & 15) == 0, "peer_reduce_sum: n must be a multiple of 4, buffers 16-byte aligned");
    if (n == 0) return CTR_OK;
    int64_t blocks = ctr_cdiv(n / 4, 256);
    const int64_t cap = (int64_t) h->sm_count * 2;
    if (blocks > cap) blocks = cap;
    peer_reduce_sum_kernel<<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(make_route(rt), n / 4, reinterpret_cast<float4*>(out));
    CTR_CHECK_LAUNCH(h, "peer_reduce_sum");
    return CTR_OK;
}

// ------------------------------------------------------------------------------------------------ device-side routing
// one warp: lane p / lane o walks the (tiny) count matrix.  See ctr_peer_plan_routes in the header for the layout.
__global__ void __launch_bounds__(32)
plan_routes_kernel(const uint32_t* __restrict__ mat, int world, int rank, int64_t req_cap, int64_t grad_cap,
                   uint32_t* __restrict__ plan, uint32_t* __restrict__ status) {
    const int l = threadIdx.x;
    uint32_t recv = 0, send = 0, v_dst = 0, g_dst = 0, col = 0;
    if (l < world) {
        recv = mat[l * world + rank];                      // rows peer l requests of me
        send = mat[rank * world + l];                      // rows I request of owner l
        for (int o = 0; o < rank; ++o) v_dst += mat[l * world + o];       // where my rows start in peer l's unique list
        for (int p = 0; p < rank; ++p) g_dst += mat[p * world + l];       // where my rows start in owner l's receive buffer
        for (int p = 0; p < world; ++p) col += mat[p * world + l];        // everything owner l will receive
    }
    uint32_t r_inc = recv, s_inc = send;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, r_inc, o), b = __shfl_up_sync(0xffffffffu, s_inc, o);
        if (l >= o) { r_inc += a; s_inc += b; }
    }
    const uint32_t n_req = __shfl_sync(0xffffffffu, r_inc, world - 1), U = __shfl_sync(0xffffffffu, s_inc, world - 1);
    if (l < CTR_MAX_PEERS) {
        plan[CTR_PLAN_VEC + l] = l < world ? r_inc : n_req;
        plan[CTR_PLAN_VEC + 16 + l] = v_dst;
        plan[CTR_PLAN_GRAD + l] = l < world ? s_inc : U;
        plan[CTR_PLAN_GRAD + 16 + l] = g_dst;
    }
    if (l == 0) {
        plan[CTR_PLAN_VEC + 32] = n_req;
        plan[CTR_PLAN_GRAD + 32] = U;
        plan[CTR_PLAN_NREQ] = n_req;
        plan[CTR_PLAN_U] = U;
    }
    // the same matrix on every rank => the same verdict on every rank
    const bool over = l < world && ((int64_t) col > grad_cap || (int64_t) col > req_cap);
    const uint32_t who = __ballot_sync(0xffffffffu, over);
    if (l == 0 && who && status) status[CTR_STATUS_EXCHANGE] = (uint32_t) __ffs(who);
}

extern "C" int ctr_peer_plan_routes(ctr_handle* h, ctr_stream s, const uint32_t* mat, int world, int rank, int64_t req_cap,
                                    int64_t grad_cap, uint32_t* plan, uint32_t* status) {
    CTR_REQUIRE(h, mat && plan && world >= 1 && world <= CTR_MAX_PEERS && rank >= 0 && rank < world && req_cap >= 0 && grad_cap >= 0,
                "peer_plan_routes: bad args");
    plan_routes_kernel<<<1, 32, 0, ctr_st(s)>>>(mat, world, rank, req_cap, grad_cap, plan, status);
    CTR_CHECK_LAUNCH(h, "peer_plan_routes");
    return CTR_OK;
}

__global__ void zero_row_at_kernel(float* __restrict__ buf, int stride, int width, const uint32_t* __restrict__ row_index) {
    float* row = buf + (int64_t) *row_index * stride;
    for (int j = threadIdx.x; j < width; j += blockDim.x) row[j] = 0.f;
}

extern "C" int ctr_zero_row_at(ctr_handle* h, ctr_stream s, float* buf, int stride, int width, const uint32_t* row_index) {
    CTR_REQUIRE(h, buf && row_index && width > 0 && stride >= width, "zero_row_at: bad args");
    zero_row_at_kernel<<<1, 64, 0, ctr_st(s)>>>(buf, stride, width, row_index);
    CTR_CHECK_LAUNCH(h, "zero_row_at");
    return CTR_OK;
}

// ------------------------------------------------------------------------------------------------ merge of sorted segments
struct SegEnds {
    uint32_t end[CTR_MAX_PEERS];
};

__global__ void __launch_bounds__(256)
merge_segments_kernel(const uint32_t* __restrict__ keys, uint32_t n, SegEnds se, int world,
                      uint32_t* __restrict__ out_keys, uint32_t* __restrict__ out_idx, const uint32_t* __restrict__ se_dev) {
    if (se_dev) {                 // split sizes known to the device only
        for (int q = 0; q < CTR_MAX_PEERS; ++q) se.end[q] = __ldg(se_dev + (q < world ? q : world - 1));
        n = se.end[world - 1];
    }
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        int p = 0;
        uint32_t pb = 0;
        for (int q = 0; q < world - 1; ++q) {
            const bool past = j >= se.end[q];
            p += past ? 1 : 0;
            pb = past ? se.end[q] : pb;
        }
        const uint32_t key = __ldg(keys + j);
        uint32_t pos = j - pb;                       // entries of the own segment before this one
        uint32_t qb = 0;
        for (int q = 0; q < world; ++q) {
            const uint32_t qe = se.end[q];
            if (q != p) {                            // entries of segment q that sort before (q < p: before or equal)
                uint32_t lo = qb, hi = qe;
                while (lo < hi) {
                    const uint32_t mid = (lo + hi) >> 1;
                    const uint32_t v = __ldg(keys + mid);
                    const bool before = q < p ? v <= key : v < key;
                    lo = before ? mid + 1 : lo;
                    hi = before ? hi : mid;
                }
                pos += lo - qb;
            }
            qb = qe;
        }
        out_keys[pos] = key;
        out_idx[pos] = j;
    }
}

extern "C" int ctr_merge_sorted_segments(ctr_handle* h, ctr_stream s, const uint32_t* keys, int64_t n, const uint32_t* seg_end,
                                         int world, uint32_t* out_keys, uint32_t* out_idx, const uint32_t* seg_end_dev) {
    CTR_REQUIRE(h, keys && (seg_end || seg_end_dev) && out_keys && out_idx && world >= 1 && world <= CTR_MAX_PEERS && n >= 0 &&
                       n < (1ll << 32), "merge_sorted_segments: bad args");
    CTR_REQUIRE(h, seg_end_dev || (int64_t) seg_end[world - 1] == n, "merge_sorted_segments: segments do not cover the list");
    if (n == 0) return CTR_OK;
    SegEnds se;
    for (int p = 0; p < CTR_MAX_PEERS; ++p) se.end[p] = seg_end_dev ? 0u : (p < world ? seg_end[p] : (uint32_t) n);
    int64_t blocks = ctr_cdiv(n, 256);
    int64_t cap = (int64_t) h->sm_count * 8;
    if (blocks > cap) blocks = cap;
    merge_segments_kernel<<<(unsigned) blocks, 256, 0, ctr_st(s)>>>(keys, (uint32_t) n, se, world, out_keys, out_idx, seg_end_dev);
    CTR_CHECK_LAUNCH(h, "merge_sorted_segments");
    return CTR_OK;
}
