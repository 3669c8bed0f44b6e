// Stable LSD radix sort (9-bit digits), flag scans and run-length structure of sorted keys.
// These are the integer half of K4 (the deterministic replacement of TF's unique +
// unsorted_segment_sum) and of the AUC; all results are bit-exact and order-deterministic.
#include "ctr_common.cuh"

namespace {

constexpr int RADIX_BITS = 9;                            // 27-bit row keys (1e8 rows, any shard count) sort in 3 passes
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 8;                            // keys per thread (16: 95 registers, 2 CTAs/SM, 1.4 waves at 1.7 M keys)
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;     // 2048 keys per block
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int WARP_SPAN = SORT_TILE / SORT_WARPS;        // 256 consecutive keys per warp
constexpr int DIGITS_PER_THREAD = RADIX / SORT_THREADS;
static_assert(DIGITS_PER_THREAD * SORT_THREADS == RADIX, "whole digits per thread");

// ---- pass 1: per-block digit histogram, stored digit-major hist[d * nblk + blk] ---------------
__global__ void __launch_bounds__(SORT_THREADS)
radix_hist_kernel(const uint32_t* __restrict__ keys, int64_t n, int shift, uint32_t* __restrict__ hist, int nblk) {
    __shared__ uint32_t sh[RADIX];
    for (int i = threadIdx.x; i < RADIX; i += SORT_THREADS) sh[i] = 0;
    __syncthreads();
    int64_t base = (int64_t) blockIdx.x * SORT_TILE;
#pragma unroll 4
    for (int it = 0; it < SORT_ITEMS; ++it) {
        int64_t i = base + it * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&sh[(keys[i] >> shift) & (RADIX - 1)], 1u);
    }
    __syncthreads();
    for (int d = threadIdx.x; d < RADIX; d += SORT_THREADS) hist[(int64_t) d * nblk + blockIdx.x] = sh[d];
}

// block-wide exclusive scan of one value per thread (blockDim.x == 1024 or 256); returns exclusive
// prefix, *total gets the block total.  smem: 33 uint32
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* sm, uint32_t* total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) sm[w] = inc;
    __syncthreads();
    if (w == 0) {
        int nw = blockDim.x >> 5;
        uint32_t x = lane < nw ? sm[lane] : 0u;
        uint32_t xi = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, xi, o);
            if (lane >= o) xi += t;
        }
        sm[lane] = xi - x;               // exclusive warp offsets
        if (lane == 31) sm[32] = xi;     // block total
    }
    __syncthreads();
    uint32_t res = sm[w] + inc - v;
    *total = sm[32];
    __syncthreads();
    return res;
}

// ---- pass 2: one block per digit scans its row of per-block counts; row totals to tot[d] --------
__global__ void __launch_bounds__(1024)
radix_scan_rows_kernel(uint32_t* __restrict__ hist, int nblk, uint32_t* __restrict__ tot) {
    __shared__ uint32_t sm[33];
    uint32_t* row = hist + (int64_t) blockIdx.x * nblk;
    uint32_t carry = 0;
    for (int i0 = 0; i0 < nblk; i0 += 1024) {
        int i = i0 + threadIdx.x;
        uint32_t v = i < nblk ? row[i] : 0u;
        uint32_t total;
        uint32_t ex = block_excl_scan(v, sm, &total);
        if (i < nblk) row[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) tot[blockIdx.x] = carry;
}

// ---- pass 3: stable scatter ------------------------------------------------------------------
// Warp w owns keys [w*256, (w+1)*256) of the tile and walks them in 8 rounds of 32 consecutive keys;
// a ballot-built peer mask ranks equal digits inside a round in lane (= index) order, a per-warp counter
// table carries ranks across rounds, a cross-warp exclusive sum and a scan over the digits finish the tile-local
// rank.  The tile is then ordered in shared memory and written out from there: consecutive threads store consecutive
// positions of a digit's run, so a warp store touches a few sectors instead of 32.
// 6 CTAs per SM (40 registers, 36 KB of shared memory each): the 833 tiles of a 64K x 26 batch are then ONE wave; at the 62
// registers the compiler takes unasked it was 4 per SM and a second, 40 %-full wave.  SORT_MINB is a build-time knob.
#ifndef SORT_MINB
#define SORT_MINB 6
#endif
__global__ void __launch_bounds__(SORT_THREADS, SORT_MINB)
radix_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                     uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int64_t n, int shift,
                     const uint32_t* __restrict__ hist, const uint32_t* __restrict__ tot, int nblk) {
    __shared__ uint32_t wcnt[SORT_WARPS][RADIX];   // per-warp digit counts -> per-warp bases
    __shared__ uint32_t dbase[RADIX];              // global position of the digit's first key of this tile, minus lstart
    __shared__ uint32_t lstart[RADIX];             // tile-local start of each digit
    __shared__ uint32_t skey[SORT_TILE], sval[SORT_TILE];
    __shared__ uint32_t sm[33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int d0 = threadIdx.x * DIGITS_PER_THREAD;        // this thread's digits in the digit-indexed steps
    for (int i = threadIdx.x; i < SORT_WARPS * RADIX; i += SORT_THREADS) (&wcnt[0][0])[i] = 0;
    // digit bases: exclusive scan of the row totals + this block's offset inside the row
    {
        uint32_t t[DIGITS_PER_THREAD], sum = 0;
#pragma unroll
        for (int j = 0; j < DIGITS_PER_THREAD; ++j) {
            t[j] = tot[d0 + j];
            sum += t[j];
        }
        uint32_t total;
        uint32_t ex = block_excl_scan(sum, sm, &total);   // includes the __syncthreads for wcnt zeroing
#pragma unroll
        for (int j = 0; j < DIGITS_PER_THREAD; ++j) {
            dbase[d0 + j] = ex + hist[(int64_t) (d0 + j) * nblk + blockIdx.x];
            ex += t[j];
        }
    }
    __syncthreads();

    const int64_t tbase = (int64_t) blockIdx.x * SORT_TILE;
    const int64_t wbase = tbase + (int64_t) w * WARP_SPAN;
    uint32_t key[SORT_ITEMS];
    uint32_t off[SORT_ITEMS];   // rank among equal digits inside this warp's span
#pragma unroll
    for (int it = 0; it < SORT_ITEMS; ++it) {
        int64_t i = wbase + it * 32 + lane;
        key[it] = i < n ? keys_in[i] : 0xFFFFFFFFu;
    }
#pragma unroll
    for (int it = 0; it < SORT_ITEMS; ++it) {
        int64_t i = wbase + it * 32 + lane;
        bool valid = i < n;
        uint32_t d = (key[it] >> shift) & (RADIX - 1);
        // invalid lanes use a digit slot of their own (RADIX-1 of PAD keys still must not mix):
        // match on (valid, d) so padding lanes never rank against real keys
        // lanes with the same digit, from RADIX_BITS + 1 ballots.  (__match_any_sync does this in one instruction, but the
        // hardware walks the distinct values of the warp one by one: with 9-bit digits nearly every lane is its own
        // group, and MATCH.ANY was what this kernel's 29 us per pass were made of.)
        uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int b = 0; b < RADIX_BITS; ++b) {
            const uint32_t m = __ballot_sync(0xffffffffu, (d >> b) & 1u);
            peers &= ((d >> b) & 1u) ? m : ~m;
        }
        if (!valid) peers = 1u << lane;
        uint32_t rank = __popc(peers & ((1u << lane) - 1u));
        int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (valid && lane == leader) {
            old = wcnt[w][d];
            wcnt[w][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        off[it] = old + rank;
        __syncwarp();
    }
    __syncthreads();
    // cross-warp exclusive sum per digit, then the tile-local start of each digit
    {
        uint32_t cnt[DIGITS_PER_THREAD], sum = 0;
#pragma unroll
        for (int j = 0; j < DIGITS_PER_THREAD; ++j) {
            uint32_t run = 0;
#pragma unroll
            for (int ww = 0; ww < SORT_WARPS; ++ww) {
                uint32_t c = wcnt[ww][d0 + j];
                wcnt[ww][d0 + j] = run;
                run += c;
            }
            cnt[j] = run;
            sum += run;
        }
        uint32_t total;
        uint32_t ex = block_excl_scan(sum, sm, &total);
#pragma unroll
        for (int j = 0; j < DIGITS_PER_THREAD; ++j) {
            lstart[d0 + j] = ex;
            dbase[d0 + j] -= ex;
            ex += cnt[j];
        }
    }
    __syncthreads();
#pragma unroll
    for (int it = 0; it < SORT_ITEMS; ++it) {
        int64_t i = wbase + it * 32 + lane;
        if (i < n) {
            uint32_t d = (key[it] >> shift) & (RADIX - 1);
            uint32_t lp = lstart[d] + wcnt[w][d] + off[it];
            skey[lp] = key[it];
            sval[lp] = vals_in[i];
        }
    }
    __syncthreads();
    const int nvalid = (int) (n - tbase < (int64_t) SORT_TILE ? n - tbase : (int64_t) SORT_TILE);
#pragma unroll
    for (int it = 0; it < SORT_ITEMS; ++it) {
        int idx = it * SORT_THREADS + threadIdx.x;
        if (idx < nvalid) {
            uint32_t k = skey[idx];
            uint32_t pos = dbase[(k >> shift) & (RADIX - 1)] + (uint32_t) idx;
            keys_out[pos] = k;
            vals_out[pos] = sval[idx];
        }
    }
}

__global__ void copy_pairs_kernel(const uint32_t* __restrict__ ka, const uint32_t* __restrict__ va,
                                  uint32_t* __restrict__ kb, uint32_t* __restrict__ vb, int64_t n) {
    int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        kb[i] = ka[i];
        vb[i] = va[i];
    }
}

// ids -> (key, flat position) pairs, fused with the per-block digit histogram of the FIRST radix pass (the pairs are
// produced tile by tile in exactly the order radix_scatter_kernel will read them, so its histogram pass is free)
__global__ void __launch_bounds__(SORT_THREADS)
build_keys_kernel(const void* __restrict__ ids, int id_bytes, int64_t rows, int64_t ld, int col0, int ncols,
                  int64_t id_offset, int drop_zero, int64_t extra_base, int n_extra, int n_shards, int local_bits,
                  int64_t id_limit, uint32_t* __restrict__ status, uint32_t* __restrict__ keys, uint32_t* __restrict__ perm,
                  uint32_t* __restrict__ hist, int nblk) {
    __shared__ uint32_t sh[RADIX];
    if (hist) {
        for (int i = threadIdx.x; i < RADIX; i += SORT_THREADS) sh[i] = 0;
        __syncthreads();
    }
    const int64_t n = rows * ncols;
    const int64_t base = (int64_t) blockIdx.x * SORT_TILE;
#pragma unroll 4
    for (int it = 0; it < SORT_ITEMS; ++it) {
        const int64_t i = base + it * SORT_THREADS + threadIdx.x;
        if (i >= n + n_extra) break;
        int64_t p, id;
        bool pad = false;
        if (i < n) {
            int64_t r = i / ncols;
            int c = (int) (i - r * ncols);
            p = r * ld + col0 + c;
            id = ctr_load_id(ids, id_bytes, p) + id_offset;
            pad = drop_zero && id == 0;
            if (id_limit > 0 && (uint64_t) id >= (uint64_t) id_limit) {      // no gradient for a lookup that read nothing
                ctr_flag_bad_id(status, col0 + c);
                pad = true;
            }
        } else {
            p = rows * ld + (i - n);
            id = extra_base + (i - n);
        }
        uint32_t key;
        if (pad) key = CTR_PAD_KEY;
        else if (n_shards > 1) key = ((uint32_t) (id % n_shards) << local_bits) | (uint32_t) (id / n_shards);
        else key = (uint32_t) id;
        keys[i] = key;
        perm[i] = (uint32_t) p;
        if (hist) atomicAdd(&sh[key & (RADIX - 1)], 1u);
    }
    if (hist) {
        __syncthreads();
        for (int d = threadIdx.x; d < RADIX; d += SORT_THREADS) hist[(int64_t) d * nblk + blockIdx.x] = sh[d];
    }
}

// ---- generic 3-phase flag scan ------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;   // 2048

struct PosLabel {   // 1 where the label of the i-th sorted element is positive
    const float* labels;
    const uint32_t* perm;
    __device__ __forceinline__ uint32_t operator()(int64_t i) const { return labels[perm[i]] > 0.5f ? 1u : 0u; }
};

template <class F>
__global__ void __launch_bounds__(SCAN_THREADS) scan_phase1(F f, int64_t n, uint32_t* __restrict__ blk) {
    __shared__ uint32_t sm[33];
    int64_t base = (int64_t) blockIdx.x * SCAN_TILE + (int64_t) threadIdx.x * SCAN_ITEMS;
    uint32_t c = 0;
#pragma unroll
    for (int it = 0; it < SCAN_ITEMS; ++it)
        if (base + it < n) c += f(base + it);
    uint32_t total;
    block_excl_scan(c, sm, &total);
    if (threadIdx.x == 0) blk[blockIdx.x] = total;
}

// single block: exclusive scan of blk[0..nblk) in place, grand total to *total_out (and blk[nblk])
__global__ void __maxnreg__(24) scan_phase2(uint32_t* __restrict__ blk, int nblk, uint32_t* __restrict__ total_out) {
    __shared__ uint32_t sm[33];
    uint32_t carry = 0;
    for (int i0 = 0; i0 < nblk; i0 += 1024) {
        int i = i0 + threadIdx.x;
        uint32_t v = i < nblk ? blk[i] : 0u;
        uint32_t total;
        uint32_t ex = block_excl_scan(v, sm, &total);
        if (i < nblk) blk[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) {
        blk[nblk] = carry;
        if (total_out) *total_out = carry;
    }
}

struct ExclOut {   // out[i] = exclusive prefix
    uint32_t* out;
    __device__ __forceinline__ void operator()(int64_t i, uint32_t incl, uint32_t flag) const { out[i] = incl - flag; }
};

template <class F, class O>
__global__ void __launch_bounds__(SCAN_THREADS) scan_phase3(F f, int64_t n, const uint32_t* __restrict__ blk, O out) {
    __shared__ uint32_t sm[33];
    int64_t base = (int64_t) blockIdx.x * SCAN_TILE + (int64_t) threadIdx.x * SCAN_ITEMS;
    uint32_t fl[SCAN_ITEMS];
    uint32_t c = 0;
#pragma unroll
    for (int it = 0; it < SCAN_ITEMS; ++it) {
        fl[it] = (base + it < n) ? f(base + it) : 0u;
        c += fl[it];
    }
    uint32_t total;
    uint32_t run = block_excl_scan(c, sm, &total) + blk[blockIdx.x];
#pragma unroll
    for (int it = 0; it < SCAN_ITEMS; ++it) {
        if (base + it < n) {
            run += fl[it];
            out(base + it, run, fl[it]);
        }
    }
}

__global__ void shard_counts_kernel(const uint32_t* __restrict__ uniq, const uint32_t* __restrict__ n_unique,
                                    int n_shards, int local_bits, uint32_t* __restrict__ counts) {
    int o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= n_shards) return;
    uint32_t U = *n_unique;
    auto lower = [&](uint32_t key) {
        uint32_t lo = 0, hi = U;
        while (lo < hi) {
            uint32_t mid = (lo + hi) >> 1;
            if (uniq[mid] < key) lo = mid + 1; else hi = mid;
        }
        return lo;
    };
    uint32_t a = lower((uint32_t) o << local_bits);
    uint32_t b = (o + 1 < n_shards) ? lower((uint32_t) (o + 1) << local_bits) : lower(CTR_PAD_KEY);
    counts[o] = b - a;
}

inline int sort_nblk(int64_t n) { return (int) ctr_cdiv(n > 0 ? n : 1, SORT_TILE); }

}  // namespace

extern "C" int64_t ctr_sort_ws_bytes(ctr_handle*, int64_t n) {
    int64_t nb = sort_nblk(n);
    return ctr_align(n * 4, 256) * 2 + ctr_align((int64_t) RADIX * nb * 4, 256) + RADIX * 4 + 256;
}

// `passes` stable LSD passes from (ka, va), ping-ponging with (kb, vb); the result is in (ka, va) when passes is even,
// else in (kb, vb).  first_hist_ready: hist already holds the per-block histogram of pass 0 (build_keys_kernel).
static int sort_passes(ctr_handle* h, cudaStream_t st, uint32_t* ka, uint32_t* va, uint32_t* kb, uint32_t* vb, int64_t n,
                       int passes, uint32_t* hist, uint32_t* tot, bool first_hist_ready) {
    const int nblk = sort_nblk(n);
    for (int pass = 0; pass < passes; ++pass) {
        int shift = pass * RADIX_BITS;
        if (!(pass == 0 && first_hist_ready)) {
            radix_hist_kernel<<<nblk, SORT_THREADS, 0, st>>>(ka, n, shift, hist, nblk);
            h->launches += 1;
        }
        radix_scan_rows_kernel<<<RADIX, 1024, 0, st>>>(hist, nblk, tot);
        radix_scatter_kernel<<<nblk, SORT_THREADS, 0, st>>>(ka, va, kb, vb, n, shift, hist, tot, nblk);
        h->launches += 1;
        uint32_t* t;
        t = ka; ka = kb; kb = t;
        t = va; va = vb; vb = t;
        CTR_CHECK_LAUNCH(h, "sort pass");
    }
    return CTR_OK;
}

struct SortWs {
    uint32_t *k2, *v2, *hist, *tot;
};
static SortWs carve_sort_ws(void* ws, int64_t n) {
    const int nblk = sort_nblk(n);
    char* p = (char*) ws;
    SortWs w;
    w.k2 = (uint32_t*) p; p += ctr_align(n * 4, 256);
    w.v2 = (uint32_t*) p; p += ctr_align(n * 4, 256);
    w.hist = (uint32_t*) p; p += ctr_align((int64_t) RADIX * nblk * 4, 256);
    w.tot = (uint32_t*) p;
    return w;
}

extern "C" int ctr_sort_pairs(ctr_handle* h, ctr_stream s, uint32_t* keys, uint32_t* vals, int64_t n, int key_bits,
                              void* ws, int64_t ws_bytes) {
    CTR_REQUIRE(h, keys && vals && ws && n >= 0 && key_bits >= 1 && key_bits <= 32, "sort_pairs: bad args");
    CTR_REQUIRE(h, n < (int64_t) 1 << 31, "sort_pairs: n too large");
    if (n == 0) return CTR_OK;
    if (ws_bytes < ctr_sort_ws_bytes(h, n)) return ctr_fail(h, CTR_E_WORKSPACE, "sort_pairs: ws too small");
    const SortWs w = carve_sort_ws(ws, n);
    const int passes = (key_bits + RADIX_BITS - 1) / RADIX_BITS;
    cudaStream_t st = ctr_st(s);
    int rc = sort_passes(h, st, keys, vals, w.k2, w.v2, n, passes, w.hist, w.tot, false);
    if (rc) return rc;
    if (passes & 1) {        // in-place contract: an odd pass count ends in the workspace
        copy_pairs_kernel<<<(unsigned) ctr_cdiv(n, 256), 256, 0, st>>>(w.k2, w.v2, keys, vals, n);
        CTR_CHECK_LAUNCH(h, "sort_pairs copy");
    }
    return CTR_OK;
}

extern "C" int ctr_sort_ids(ctr_handle* h, ctr_stream s, const void* ids, int id_bytes, int64_t rows, int64_t ld,
                            int col0, int ncols, int64_t id_offset, int drop_zero, int64_t extra_base, int n_extra,
                            int n_shards, int local_bits,
                            int key_bits, int64_t id_limit, uint32_t* status, uint32_t* keys_out, uint32_t* perm_out,
                            void* ws, int64_t ws_bytes) {
    CTR_REQUIRE(h, ids && keys_out && perm_out && (id_bytes == 4 || id_bytes == 8), "sort_ids: bad pointers");
    CTR_REQUIRE(h, rows >= 0 && ncols >= 0 && col0 >= 0 && col0 + ncols <= ld, "sort_ids: bad shape");
    CTR_REQUIRE(h, n_shards >= 1 && (n_shards == 1 || (local_bits > 0 && local_bits < 31)), "sort_ids: bad sharding");
    CTR_REQUIRE(h, n_extra >= 0 && rows * ld + n_extra < ((int64_t) 1 << 32), "sort_ids: flat positions exceed 32 bits");
    CTR_REQUIRE(h, key_bits >= 1 && key_bits <= 32, "sort_ids: bad key_bits");
    int64_t n = rows * ncols + n_extra;
    if (n == 0) return CTR_OK;
    CTR_REQUIRE(h, ws && n < (int64_t) 1 << 31, "sort_ids: n too large / no workspace");
    if (ws_bytes < ctr_sort_ws_bytes(h, n)) return ctr_fail(h, CTR_E_WORKSPACE, "sort_ids: ws too small");
    const SortWs w = carve_sort_ws(ws, n);
    // padding keys (0xFFFFFFFF) must sort behind every real key: all 32 bits when ids can be dropped.  (Ids outside
    // id_limit also become padding keys, but that batch is a fault: the status word stops the optimizer, the order of its
    // key list no longer matters, and the healthy path does not pay a fourth pass for it.)
    const int passes = ((drop_zero ? 32 : key_bits) + RADIX_BITS - 1) / RADIX_BITS;
    // the pairs are built where an odd / even number of ping-pong passes ends in (keys_out, perm_out): no final copy
    uint32_t *k0 = (passes & 1) ? w.k2 : keys_out, *v0 = (passes & 1) ? w.v2 : perm_out;
    uint32_t *k1 = (passes & 1) ? keys_out : w.k2, *v1 = (passes & 1) ? perm_out : w.v2;
    const int nblk = sort_nblk(n);
    build_keys_kernel<<<nblk, SORT_THREADS, 0, ctr_st(s)>>>(ids, id_bytes, rows, ld, col0, ncols, id_offset, drop_zero,
                                                          extra_base, n_extra, n_shards, local_bits, id_limit, status,
                                                          k0, v0, w.hist, nblk);
    CTR_CHECK_LAUNCH(h, "build_keys");
    return sort_passes(h, ctr_st(s), k0, v0, k1, v1, n, passes, w.hist, w.tot, true);
}

namespace {
// ---- run-length structure of sorted keys in two launches --------------------------------------------------------------
// Warp w of a block owns 256 consecutive keys and walks them in 8 rounds of 32 (coalesced); a round's head flags are one
// ballot, so ranks inside the warp are popcounts of register-held ballots.  Pass 1 writes the flag count of each
// 2048-key tile; pass 2 recomputes the flags (the keys are L2-resident), lets every block sum the counts of the tiles
// before it (a few loads per thread) and emits uidx / uniq / seg_start.  The last block also writes *n_unique and
// seg_start[U] = n.
constexpr int UQ_ROUNDS = 8;
constexpr int UQ_TILE = SCAN_THREADS * UQ_ROUNDS;       // 2048
__device__ __forceinline__ uint32_t uq_flags(const uint32_t* __restrict__ keys, int64_t n, int64_t wbase, int lane,
                                             uint32_t (&bal)[UQ_ROUNDS], uint32_t (&k)[UQ_ROUNDS]) {
    uint32_t prev = (lane == 0 && wbase > 0 && wbase < n) ? keys[wbase - 1] : 0u;
#pragma unroll
    for (int r = 0; r < UQ_ROUNDS; ++r) {
        const int64_t i = wbase + r * 32 + lane;
        k[r] = i < n ? keys[i] : 0u;
    }
    uint32_t total = 0;
#pragma unroll
    for (int r = 0; r < UQ_ROUNDS; ++r) {
        const int64_t i = wbase + r * 32 + lane;
        uint32_t up = __shfl_up_sync(0xffffffffu, k[r], 1);
        if (lane == 0) up = prev;
        const bool flag = i < n && (i == 0 || k[r] != up);
        bal[r] = __ballot_sync(0xffffffffu, flag);
        total += __popc(bal[r]);
        prev = __shfl_sync(0xffffffffu, k[r], 31);
    }
    return total;
}

__global__ void __launch_bounds__(SCAN_THREADS)
unique_count_kernel(const uint32_t* __restrict__ keys, int64_t n, uint32_t* __restrict__ blk, const uint32_t* __restrict__ n_dev) {
    __shared__ uint32_t wtot[SCAN_THREADS / 32];
    if (n_dev) n = (int64_t) *n_dev < n ? (int64_t) *n_dev : n;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t bal[UQ_ROUNDS], k[UQ_ROUNDS];
    const uint32_t t = uq_flags(keys, n, (int64_t) blockIdx.x * UQ_TILE + w * (32 * UQ_ROUNDS), lane, bal, k);
    if (lane == 0) wtot[w] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t s = 0;
#pragma unroll
        for (int i = 0; i < SCAN_THREADS / 32; ++i) s += wtot[i];
        blk[blockIdx.x] = s;
    }
}

__global__ void __launch_bounds__(SCAN_THREADS)
unique_emit_kernel(const uint32_t* __restrict__ keys, int64_t n, const uint32_t* __restrict__ blk, uint32_t* __restrict__ uniq,
                   uint32_t* __restrict__ seg_start, uint32_t* __restrict__ uidx, uint32_t* __restrict__ n_unique,
                   const uint32_t* __restrict__ n_dev) {
    __shared__ uint32_t wtot[SCAN_THREADS / 32], red[SCAN_THREADS / 32];
    if (n_dev) n = (int64_t) *n_dev < n ? (int64_t) *n_dev : n;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    // flags of the tiles before this one
    uint32_t before = 0;
    for (int j = threadIdx.x; j < (int) blockIdx.x; j += SCAN_THREADS) before += blk[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) before += __shfl_xor_sync(0xffffffffu, before, o);
    if (lane == 0) red[w] = before;
    const int64_t wbase = (int64_t) blockIdx.x * UQ_TILE + w * (32 * UQ_ROUNDS);
    uint32_t bal[UQ_ROUNDS], k[UQ_ROUNDS];
    const uint32_t t = uq_flags(keys, n, wbase, lane, bal, k);
    if (lane == 0) wtot[w] = t;
    __syncthreads();
    uint32_t run = 0, all = 0;
#pragma unroll
    for (int i = 0; i < SCAN_THREADS / 32; ++i) {
        run += red[i];
        if (i < w) run += wtot[i];
        all += wtot[i];
    }
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < UQ_ROUNDS; ++r) {
        const int64_t i = wbase + r * 32 + lane;
        const bool flag = (bal[r] >> lane) & 1u;
        const uint32_t incl = run + __popc(bal[r] & lt) + (flag ? 1u : 0u);
        if (i < n) {
            const uint32_t u = incl - 1u;
            uidx[i] = u;
            if (flag) {
                uniq[u] = k[r];
                seg_start[u] = (uint32_t) i;
            }
        }
        run += __popc(bal[r]);
    }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
        uint32_t tot_before = 0;
#pragma unroll
        for (int i = 0; i < SCAN_THREADS / 32; ++i) tot_before += red[i];
        const uint32_t U = tot_before + all;
        *n_unique = U;
        seg_start[U] = (uint32_t) n;
    }
}

}  // namespace

extern "C" int ctr_unique_segments(ctr_handle* h, ctr_stream s, const uint32_t* sorted_keys, int64_t n, uint32_t* uniq,
                                   uint32_t* seg_start, uint32_t* uidx, uint32_t* n_unique, void* ws, int64_t ws_bytes,
                                   const uint32_t* n_dev) {
    CTR_REQUIRE(h, sorted_keys && uniq && seg_start && uidx && n_unique && ws && n >= 0, "unique_segments: bad args");
    cudaStream_t st = ctr_st(s);
    if (n == 0) {
        CTR_CUDA(h, cudaMemsetAsync(n_unique, 0, 4, st));
        CTR_CUDA(h, cudaMemsetAsync(seg_start, 0, 4, st));
        return CTR_OK;
    }
    int nblk = (int) ctr_cdiv(n, UQ_TILE);
    if (ws_bytes < (int64_t) (nblk + 2) * 4) return ctr_fail(h, CTR_E_WORKSPACE, "unique_segments: ws too small");
    uint32_t* blk = (uint32_t*) ws;
    unique_count_kernel<<<nblk, SCAN_THREADS, 0, st>>>(sorted_keys, n, blk, n_dev);
    h->launches += 1;
    unique_emit_kernel<<<nblk, SCAN_THREADS, 0, st>>>(sorted_keys, n, blk, uniq, seg_start, uidx, n_unique, n_dev);
    CTR_CHECK_LAUNCH(h, "unique_segments");
    return CTR_OK;
}

extern "C" int ctr_shard_counts(ctr_handle* h, ctr_stream s, const uint32_t* uniq, const uint32_t* n_unique,
                                int n_shards, int local_bits, uint32_t* counts) {
    CTR_REQUIRE(h, uniq && n_unique && counts && n_shards >= 1 && local_bits > 0 && local_bits < 32,
                "shard_counts: bad args");
    shard_counts_kernel<<<(unsigned) ctr_cdiv(n_shards, 64), 64, 0, ctr_st(s)>>>(uniq, n_unique, n_shards, local_bits,
                                                                                 counts);
    CTR_CHECK_LAUNCH(h, "shard_counts");
    return CTR_OK;
}

// exclusive prefix count of positive labels in sorted order (used by ctr_auc): out[i], out[n] = total
int ctr_scan_pos_labels(ctr_handle* h, cudaStream_t st, const float* labels, const uint32_t* perm, int64_t n,
                        uint32_t* out, uint32_t* blk_ws) {
    int nblk = (int) ctr_cdiv(n, SCAN_TILE);
    PosLabel f{labels, perm};
    scan_phase1<PosLabel><<<nblk, SCAN_THREADS, 0, st>>>(f, n, blk_ws);
    scan_phase2<<<1, 1024, 0, st>>>(blk_ws, nblk, out + n);
    scan_phase3<PosLabel, ExclOut><<<nblk, SCAN_THREADS, 0, st>>>(f, n, blk_ws, ExclOut{out});
    h->launches += 2;
    CTR_CHECK_LAUNCH(h, "scan_pos_labels");
    return CTR_OK;
}
