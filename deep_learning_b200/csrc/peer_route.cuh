// Device-side form of ctr_peer_route (include/ctr_b200.h): destination base pointers with the buffer offset applied,
// passed to kernels by value.
#pragma once
#include "ctr_common.cuh"

struct PeerRoute {
    int world, rank;
    void* slab[CTR_MAX_PEERS];
    float* buf[CTR_MAX_PEERS];          // slab[p] + dst_off + dst_row[p] * 0 (row offset applied in route_row)
    uint32_t seg_end[CTR_MAX_PEERS];
    uint32_t dst_row[CTR_MAX_PEERS];
    const uint32_t* dyn;                // device route block (ctr_peer_plan_routes) or nullptr: overrides seg_end / dst_row
};

static inline PeerRoute make_route(const ctr_peer_route* rt) {
    PeerRoute r;
    r.world = rt->world;
    r.rank = rt->rank;
    for (int p = 0; p < CTR_MAX_PEERS; ++p) {
        const bool on = p < rt->world;
        r.slab[p] = on ? rt->slab[p] : nullptr;
        r.buf[p] = on ? reinterpret_cast<float*>(static_cast<char*>(rt->slab[p]) + rt->dst_off) : nullptr;
        r.seg_end[p] = on ? rt->seg_end[p] : 0xFFFFFFFFu;
        r.dst_row[p] = on ? rt->dst_row[p] : 0u;
    }
    r.dyn = rt->dyn;
    return r;
}

#ifdef __CUDACC__
// rows of the routed list: known to the host, or only to the device (dyn)
__device__ __forceinline__ int64_t route_total(const PeerRoute& rt, int64_t n_host) {
    return rt.dyn ? (int64_t) __ldg(rt.dyn + 32) : n_host;
}
// row i of the source list -> its destination row (row stride ds floats) in the owning peer's buffer
__device__ __forceinline__ float* route_row(const PeerRoute& rt, uint32_t i, int ds) {
    int p = 0;
    uint32_t begin = 0;
    if (rt.dyn) {               // the route block is a few L1-resident words
        for (int j = 0; j < rt.world - 1; ++j) {
            const uint32_t se = __ldg(rt.dyn + j);
            const bool past = i >= se;
            p += past ? 1 : 0;
            begin = past ? se : begin;
        }
        return rt.buf[p] + (int64_t) (__ldg(rt.dyn + 16 + p) + (i - begin)) * ds;
    }
#pragma unroll
    for (int j = 0; j < CTR_MAX_PEERS - 1; ++j) {
        const bool past = j < rt.world - 1 && i >= rt.seg_end[j];
        p += past ? 1 : 0;
        begin = past ? rt.seg_end[j] : begin;
    }
    return rt.buf[p] + (int64_t) (rt.dst_row[p] + (i - begin)) * ds;
}
#endif
