// Host-side input decoding for the CTR hot path: text lines and TFRecord/tf.Example bytes ->
// row-major arrays in caller-owned (pinned) host memory, ready for one cudaMemcpyAsync per tensor.
//
// Replaces the per-line Python of utils/data_loader_load.py:94-139 (str.split + list appends + pickle) and the
// tf.data graphs of utils/data_loader_load.py:187-243, utils/data_loader_textline.py:4-36 and
// utils/data_loader.py:7-40 (TFRecordDataset + parse_single_example).  No device code in this file.
#include <cerrno>
#include <charconv>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <thread>
#include <vector>

#include "ctr_b200.h"

namespace {

// ---- text lines --------------------------------------------------------------------------------------------------
struct LineCfg {
    int n_cont, n_cate, third_kind, n_third, id_bytes;
};

inline void store_id(void* base, int id_bytes, int64_t pos, long long v) {
    if (id_bytes == 8) static_cast<int64_t*>(base)[pos] = (int64_t) v;
    else static_cast<int32_t*>(base)[pos] = (int32_t) v;
}

// Python's float(s) then a cast to float32 (what feeding str lists to a float32 placeholder does,
// utils/data_loader_load.py:24-25 -> models/deepfm.py:174-180): strtod + round-to-nearest narrowing.
// Clinger's fast path: [-]digits[.digits] with a mantissa below 2^53 and at most 22 fractional digits is
// mantissa / 10^frac with both operands exact in binary64, so ONE correctly rounded division gives strtod's result.
const double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                           1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

// eight ASCII digits at once (little-endian SWAR, the well-known fast_float / simdjson arithmetic)
inline uint64_t load8(const char* p) {
    uint64_t v;
    memcpy(&v, p, 8);
    return v;
}
inline bool eight_digits(uint64_t v) {
    return ((v & 0xF0F0F0F0F0F0F0F0ull) | (((v + 0x0606060606060606ull) & 0xF0F0F0F0F0F0F0F0ull) >> 4)) == 0x3333333333333333ull;
}
inline uint32_t parse8(uint64_t v) {
    const uint64_t mask = 0x000000FF000000FFull;
    const uint64_t mul1 = 0x000F424000000064ull;      // 100 + (1000000 << 32)
    const uint64_t mul2 = 0x0000271000000001ull;      // 1 + (10000 << 32)
    v -= 0x3030303030303030ull;
    v = (v * 10) + (v >> 8);
    v = (((v & mask) * mul1) + (((v >> 16) & mask) * mul2)) >> 32;
    return (uint32_t) v;
}

// scans [-+]digits[.digits] from p and stops at the first other character; returns that position, or nullptr when the
// spelling is not of this form or the value is outside the exact fast path (the caller then takes the whole token
// through parse_float_slow)
inline const char* scan_decimal(const char* p, const char* e, double* out) {
    bool neg = false;
    if (p < e && (*p == '-' || *p == '+')) neg = *p++ == '-';
    uint64_t m = 0;
    int nd = 0, frac = 0;
    for (; p < e && *p >= '0' && *p <= '9'; ++p, ++nd) {
        if (nd >= 19) return nullptr;
        m = m * 10 + (uint64_t) (*p - '0');
    }
    if (p < e && *p == '.') {
        ++p;
        while (e - p >= 8 && nd + 8 <= 19) {
            const uint64_t w = load8(p);
            if (!eight_digits(w)) break;
            m = m * 100000000ull + parse8(w);
            p += 8; nd += 8; frac += 8;
        }
        for (; p < e && *p >= '0' && *p <= '9'; ++p, ++nd, ++frac) {
            if (nd >= 19) return nullptr;
            m = m * 10 + (uint64_t) (*p - '0');
        }
    }
    if (nd == 0 || frac > 22 || m >= ((uint64_t) 1 << 53)) return nullptr;
    const double d = (double) m / kPow10[frac];
    *out = neg ? -d : d;
    return p;
}

// everything the fast path leaves: long mantissas and exponents through std::from_chars (correctly rounded, = strtod),
// the remaining spellings Python's float() accepts (leading '+', inf, nan, surrounding blanks) through strtod
inline bool parse_float_slow(const char* p, const char* e, float* out) {
    const size_t n = (size_t) (e - p);
    if (n == 0) return false;
    const char* f = (*p == '-') ? p + 1 : p;
    if (f < e && ((*f >= '0' && *f <= '9') || *f == '.')) {
        double d;
        const std::from_chars_result r = std::from_chars(p, e, d);
        if (r.ec == std::errc() && r.ptr == e) {
            *out = (float) d;
            return true;
        }
    }
    char tmp[64];
    if (n >= sizeof(tmp)) return false;
    memcpy(tmp, p, n);
    tmp[n] = 0;
    if (memchr(tmp, 'x', n) || memchr(tmp, 'X', n)) return false;     // strtod reads hex floats, Python's float() does not
    char* end = nullptr;
    const double d = strtod(tmp, &end);
    if (end == tmp) return false;
    while (*end == ' ' || *end == '\t' || *end == '\r') ++end;
    if (*end != 0) return false;
    *out = (float) d;
    return true;
}

inline bool parse_float_tok(const char* p, const char* e, float* out) {
    double fd;
    if (scan_decimal(p, e, &fd) == e) {
        *out = (float) fd;
        return true;
    }
    return parse_float_slow(p, e, out);
}

// one float of a `sep`-separated list inside [a, q): value stored, *next = end of the token (at `sep` or at q)
inline bool next_float(const char* a, const char* q, char sep, float* out, const char** next) {
    double fd;
    const char* b = scan_decimal(a, q, &fd);
    if (b && (b == q || *b == sep)) {
        *out = (float) fd;
        *next = b;
        return true;
    }
    b = (const char*) memchr(a, sep, (size_t) (q - a));
    if (!b) b = q;
    *next = b;
    return parse_float_slow(a, b, out);
}

inline bool parse_int_tok(const char* p, const char* e, long long* out);
// one non-negative decimal id of a `sep`-separated list; anything else goes through parse_int_tok
inline bool next_int(const char* a, const char* q, char sep, long long* out, const char** next) {
    const char* b = a;
    long long v = 0;
    for (; b < q && *b >= '0' && *b <= '9' && b - a < 18; ++b) v = v * 10 + (*b - '0');
    if (b > a && (b == q || *b == sep)) {
        *out = v;
        *next = b;
        return true;
    }
    b = (const char*) memchr(a, sep, (size_t) (q - a));
    if (!b) b = q;
    *next = b;
    return parse_int_tok(a, b, out);
}

inline bool parse_int_tok(const char* p, const char* e, long long* out) {
    if (e - p > 0 && e - p <= 18 && *p >= '0' && *p <= '9') {      // plain non-negative decimal: no strtoll
        long long v = 0;
        const char* q = p;
        for (; q < e && *q >= '0' && *q <= '9'; ++q) v = v * 10 + (*q - '0');
        if (q == e) {
            *out = v;
            return true;
        }
    }
    char tmp[32];
    const size_t n = (size_t) (e - p);
    if (n == 0 || n >= sizeof(tmp)) return false;
    memcpy(tmp, p, n);
    tmp[n] = 0;
    char* end = nullptr;
    errno = 0;
    const long long v = strtoll(tmp, &end, 10);
    if (end == tmp || errno) return false;
    while (*end == ' ' || *end == '\t' || *end == '\r') ++end;
    if (*end != 0) return false;
    *out = v;
    return true;
}

// next token of [p, e) split on the single character `sep` (str.split(sep) semantics: empty tokens count)
inline const char* tok_end(const char* p, const char* e, char sep) {
    const char* q = (const char*) memchr(p, sep, (size_t) (e - p));
    return q ? q : e;
}

bool parse_line(const char* p, const char* e, const LineCfg& c, int64_t row, float* label, float* cont, void* cate,
                void* third_ids, float* third_vals) {
    while (e > p && (e[-1] == '\r' || e[-1] == '\n')) --e;
    // column 0: label
    const char* q = tok_end(p, e, ',');
    if (!parse_float_tok(p, q, label + row)) return false;
    if (q == e) return false;
    p = q + 1;
    // column 1: n_cont floats separated by one space
    q = tok_end(p, e, ',');
    {
        const char* a = p;
        for (int j = 0; j < c.n_cont; ++j) {
            const char* b;
            if (!next_float(a, q, ' ', cont + row * (int64_t) c.n_cont + j, &b)) return false;
            if (j + 1 < c.n_cont) {
                if (b == q) return false;
                a = b + 1;
            } else if (b != q) {
                return false;   // more values than the model's cont_field_size (utils/data_loader_load.py:68-71)
            }
        }
    }
    if (q == e) return c.n_cate == 0 && c.third_kind == 0;
    p = q + 1;
    // column 2: n_cate integer ids
    q = tok_end(p, e, ',');
    {
        const char* a = p;
        for (int j = 0; j < c.n_cate; ++j) {
            const char* b;
            long long v;
            if (!next_int(a, q, ' ', &v, &b)) return false;
            store_id(cate, c.id_bytes, row * (int64_t) c.n_cate + j, v);
            if (j + 1 < c.n_cate) {
                if (b == q) return false;
                a = b + 1;
            } else if (b != q) {
                return false;
            }
        }
    }
    if (c.third_kind == 0) return true;
    if (q == e) return false;
    p = q + 1;
    q = tok_end(p, e, ',');
    const char* a = p;
    for (int j = 0; j < c.n_third; ++j) {
        const char* b;
        long long v;
        if (c.third_kind == 1) {           // wide ids (utils/data_loader_load.py:125-127)
            if (!next_int(a, q, ' ', &v, &b)) return false;
        } else {                           // id=value pairs (get_libsvm_value, utils/data_loader_load.py:12-21)
            b = tok_end(a, q, ' ');
            const char* eq = tok_end(a, b, '=');
            if (eq == b || !parse_int_tok(a, eq, &v)) return false;
            if (!parse_float_tok(eq + 1, b, third_vals + row * (int64_t) c.n_third + j)) return false;
        }
        store_id(third_ids, c.id_bytes, row * (int64_t) c.n_third + j, v);
        if (j + 1 < c.n_third) {
            if (b == q) return false;
            a = b + 1;
        } else if (b != q) {
            return false;
        }
    }
    return true;
}

// ---- TFRecord + tf.Example ------------------------------------------------------------------------------------------
uint32_t g_crc_table[8][256];
std::atomic<int> g_crc_ready{0};

void crc_init() {
    if (g_crc_ready.load(std::memory_order_acquire)) return;
    for (uint32_t i = 0; i < 256; ++i) {
        uint32_t c = i;
        for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;   // CRC-32C (Castagnoli), reflected
        g_crc_table[0][i] = c;
    }
    for (uint32_t i = 0; i < 256; ++i)
        for (int t = 1; t < 8; ++t) g_crc_table[t][i] = (g_crc_table[t - 1][i] >> 8) ^ g_crc_table[0][g_crc_table[t - 1][i] & 0xff];
    g_crc_ready.store(1, std::memory_order_release);
}

uint32_t crc32c(const uint8_t* p, size_t n) {
    uint32_t c = 0xffffffffu;
    while (n >= 8) {
        uint32_t lo, hi;
        memcpy(&lo, p, 4);
        memcpy(&hi, p + 4, 4);
        lo ^= c;
        c = g_crc_table[7][lo & 0xff] ^ g_crc_table[6][(lo >> 8) & 0xff] ^ g_crc_table[5][(lo >> 16) & 0xff] ^
            g_crc_table[4][lo >> 24] ^ g_crc_table[3][hi & 0xff] ^ g_crc_table[2][(hi >> 8) & 0xff] ^
            g_crc_table[1][(hi >> 16) & 0xff] ^ g_crc_table[0][hi >> 24];
        p += 8;
        n -= 8;
    }
    while (n--) c = (c >> 8) ^ g_crc_table[0][(c ^ *p++) & 0xff];
    return c ^ 0xffffffffu;
}

inline uint32_t masked_crc(const uint8_t* p, size_t n) {
    const uint32_t c = crc32c(p, n);
    return ((c >> 15) | (c << 17)) + 0xa282ead8u;
}

inline bool varint(const uint8_t*& p, const uint8_t* e, uint64_t* v) {
    uint64_t r = 0;
    for (int sh = 0; sh < 64 && p < e; sh += 7) {
        const uint8_t b = *p++;
        r |= (uint64_t) (b & 0x7f) << sh;
        if (!(b & 0x80)) {
            *v = r;
            return true;
        }
    }
    return false;
}

inline bool skip_field(const uint8_t*& p, const uint8_t* e, int wt) {
    uint64_t n;
    switch (wt) {
        case 0: return varint(p, e, &n);
        case 1: if (e - p < 8) return false; p += 8; return true;
        case 2: if (!varint(p, e, &n) || (uint64_t) (e - p) < n) return false; p += n; return true;
        case 5: if (e - p < 4) return false; p += 4; return true;
        default: return false;
    }
}

// FloatList / Int64List: `repeated T value = 1`, packed (wire type 2) or one element per tag
bool read_floats(const uint8_t* p, const uint8_t* e, float* out, int want) {
    int got = 0;
    while (p < e) {
        uint64_t tag;
        if (!varint(p, e, &tag)) return false;
        const int fn = (int) (tag >> 3), wt = (int) (tag & 7);
        if (fn == 1 && wt == 2) {
            uint64_t n;
            if (!varint(p, e, &n) || (uint64_t) (e - p) < n || (n & 3)) return false;
            if (got + (int) (n / 4) > want) return false;
            memcpy(out + got, p, n);
            got += (int) (n / 4);
            p += n;
        } else if (fn == 1 && wt == 5) {
            if (e - p < 4 || got >= want) return false;
            memcpy(out + got++, p, 4);
            p += 4;
        } else if (!skip_field(p, e, wt)) {
            return false;
        }
    }
    return got == want;
}

bool read_int64s(const uint8_t* p, const uint8_t* e, int64_t* out, int want) {
    int got = 0;
    while (p < e) {
        uint64_t tag;
        if (!varint(p, e, &tag)) return false;
        const int fn = (int) (tag >> 3), wt = (int) (tag & 7);
        if (fn == 1 && wt == 2) {
            uint64_t n;
            if (!varint(p, e, &n) || (uint64_t) (e - p) < n) return false;
            const uint8_t* pe = p + n;
            while (p < pe) {
                uint64_t v;
                if (got >= want || !varint(p, pe, &v)) return false;
                out[got++] = (int64_t) v;
            }
        } else if (fn == 1 && wt == 0) {
            uint64_t v;
            if (got >= want || !varint(p, e, &v)) return false;
            out[got++] = (int64_t) v;
        } else if (!skip_field(p, e, wt)) {
            return false;
        }
    }
    return got == want;
}

struct ExampleOut {
    float* label;
    float* cont;
    float* vec;
    int64_t* cate;
    int n_cont, n_vec, n_cate;
};

// Feature { oneof kind { BytesList bytes_list = 1; FloatList float_list = 2; Int64List int64_list = 3; } }
bool read_feature(const uint8_t* p, const uint8_t* e, const char* key, size_t klen, const ExampleOut& o, int64_t row, unsigned* seen) {
    auto is = [&](const char* s) { return klen == strlen(s) && memcmp(key, s, klen) == 0; };
    int which = is("label") ? 0 : is("cont_feats") ? 1 : is("vector_feats") ? 2 : is("cate_feats") ? 3 : -1;
    if (which < 0) return true;     // parse_single_example ignores keys the spec does not name
    while (p < e) {
        uint64_t tag, n;
        if (!varint(p, e, &tag)) return false;
        const int fn = (int) (tag >> 3), wt = (int) (tag & 7);
        if (wt != 2) {
            if (!skip_field(p, e, wt)) return false;
            continue;
        }
        if (!varint(p, e, &n) || (uint64_t) (e - p) < n) return false;
        const uint8_t* pe = p + n;
        bool ok = true;
        if (fn == 2 && which == 0) ok = read_floats(p, pe, o.label + row, 1);
        else if (fn == 2 && which == 1) ok = o.n_cont == 0 || read_floats(p, pe, o.cont + row * (int64_t) o.n_cont, o.n_cont);
        else if (fn == 2 && which == 2) ok = o.n_vec == 0 ? (p == pe || true) : read_floats(p, pe, o.vec + row * (int64_t) o.n_vec, o.n_vec);
        else if (fn == 3 && which == 3) ok = read_int64s(p, pe, o.cate + row * (int64_t) o.n_cate, o.n_cate);
        else ok = false;            // wrong list kind for a FixedLenFeature
        if (!ok) return false;
        *seen |= 1u << which;
        p = pe;
    }
    return true;
}

// Example { Features features = 1; }  Features { map<string, Feature> feature = 1; }  map entry { key = 1; value = 2; }
bool read_example(const uint8_t* p, const uint8_t* e, const ExampleOut& o, int64_t row) {
    unsigned seen = 0;
    while (p < e) {
        uint64_t tag, n;
        if (!varint(p, e, &tag)) return false;
        if ((tag >> 3) != 1 || (tag & 7) != 2) {
            if (!skip_field(p, e, (int) (tag & 7))) return false;
            continue;
        }
        if (!varint(p, e, &n) || (uint64_t) (e - p) < n) return false;
        const uint8_t* fe = p + n;               // Features
        while (p < fe) {
            uint64_t t2, n2;
            if (!varint(p, fe, &t2)) return false;
            if ((t2 >> 3) != 1 || (t2 & 7) != 2) {
                if (!skip_field(p, fe, (int) (t2 & 7))) return false;
                continue;
            }
            if (!varint(p, fe, &n2) || (uint64_t) (fe - p) < n2) return false;
            const uint8_t* me = p + n2;          // one map entry
            const char* key = nullptr;
            size_t klen = 0;
            const uint8_t* val = nullptr;
            const uint8_t* vale = nullptr;
            while (p < me) {
                uint64_t t3, n3;
                if (!varint(p, me, &t3)) return false;
                if ((t3 & 7) != 2) {
                    if (!skip_field(p, me, (int) (t3 & 7))) return false;
                    continue;
                }
                if (!varint(p, me, &n3) || (uint64_t) (me - p) < n3) return false;
                if ((t3 >> 3) == 1) { key = (const char*) p; klen = n3; }
                else if ((t3 >> 3) == 2) { val = p; vale = p + n3; }
                p += n3;
            }
            if (key && val && !read_feature(val, vale, key, klen, o, row, &seen)) return false;
        }
    }
    // FixedLenFeature without a default: a missing key is an error (utils/data_loader.py:12-24)
    unsigned need = 1u | 8u | (o.n_cont ? 2u : 0u) | (o.n_vec ? 4u : 0u);
    return (seen & need) == need;
}

}  // namespace

extern "C" int ctr_parse_lines(const char* host_buf, int64_t len, int n_cont, int n_cate, int third_kind, int n_third,
                               int id_bytes, int64_t max_rows, int n_threads, float* host_label, float* host_cont,
                               void* host_cate, void* host_third_ids, float* host_third_vals, int64_t* host_n_rows,
                               int64_t* host_consumed, int64_t* host_bad_line) {
    if (!host_buf || len < 0 || n_cont < 0 || n_cate < 0 || third_kind < 0 || third_kind > 2 || n_third < 0 ||
        (id_bytes != 4 && id_bytes != 8) || max_rows < 0 || !host_label || !host_n_rows || !host_consumed || !host_bad_line ||
        (n_cont && !host_cont) || (n_cate && !host_cate) || (third_kind && n_third && !host_third_ids) ||
        (third_kind == 2 && n_third && !host_third_vals))
        return CTR_E_INVALID;
    *host_n_rows = 0;
    *host_consumed = 0;
    *host_bad_line = -1;
    // pass 1: complete, non-blank lines (a trailing piece without '\n' is left for the next chunk)
    std::vector<std::pair<int64_t, int64_t>> lines;
    int64_t pos = 0;
    while (pos < len && (int64_t) lines.size() < max_rows) {
        const char* nl = (const char*) memchr(host_buf + pos, '\n', (size_t) (len - pos));
        if (!nl) break;
        const int64_t end = nl - host_buf;
        int64_t e2 = end;
        while (e2 > pos && (host_buf[e2 - 1] == '\r' || host_buf[e2 - 1] == ' ')) --e2;
        if (e2 > pos) lines.emplace_back(pos, end);
        pos = end + 1;
    }
    const int64_t n = (int64_t) lines.size();
    const LineCfg cfg{n_cont, n_cate, third_kind, n_third, id_bytes};
    std::atomic<int64_t> bad{INT64_MAX};
    auto work = [&](int64_t lo, int64_t hi) {
        for (int64_t r = lo; r < hi; ++r) {
            if (!parse_line(host_buf + lines[r].first, host_buf + lines[r].second, cfg, r, host_label, host_cont, host_cate,
                            host_third_ids, host_third_vals)) {
                int64_t cur = bad.load();
                while (r < cur && !bad.compare_exchange_weak(cur, r)) {}
                return;
            }
        }
    };
    int T = n_threads > 0 ? n_threads : (int) std::thread::hardware_concurrency();
    if (T < 1) T = 1;
    if ((int64_t) T > n / 256 + 1) T = (int) (n / 256 + 1);
    if (T == 1) {
        work(0, n);
    } else {
        std::vector<std::thread> th;
        const int64_t per = (n + T - 1) / T;
        for (int t = 0; t < T; ++t) th.emplace_back(work, std::min<int64_t>(n, t * per), std::min<int64_t>(n, (t + 1) * per));
        for (auto& x : th) x.join();
    }
    if (bad.load() != INT64_MAX) {
        *host_bad_line = bad.load();
        return CTR_E_INVALID;
    }
    *host_n_rows = n;
    *host_consumed = pos;
    return CTR_OK;
}

extern "C" int ctr_parse_tfrecords(const char* host_buf, int64_t len, int n_cont, int n_vec, int n_cate, int verify_crc,
                                   int64_t max_rows, float* host_label, float* host_cont, float* host_vec,
                                   int64_t* host_cate, int64_t* host_n_rows, int64_t* host_consumed) {
    if (!host_buf || len < 0 || n_cont < 0 || n_vec < 0 || n_cate <= 0 || max_rows < 0 || !host_label || !host_cate ||
        !host_n_rows || !host_consumed || (n_cont && !host_cont) || (n_vec && !host_vec))
        return CTR_E_INVALID;
    if (verify_crc) crc_init();
    const uint8_t* b = (const uint8_t*) host_buf;
    const ExampleOut o{host_label, host_cont, host_vec, host_cate, n_cont, n_vec, n_cate};
    int64_t pos = 0, row = 0;
    *host_n_rows = 0;
    *host_consumed = 0;
    // record = uint64 length | uint32 masked_crc32c(length) | data | uint32 masked_crc32c(data), little endian
    while (row < max_rows && len - pos >= 12) {
        uint64_t dl;
        memcpy(&dl, b + pos, 8);
        if (dl > (uint64_t) INT64_MAX / 2) return CTR_E_INVALID;
        if ((uint64_t) (len - pos) < 16 + dl) break;          // incomplete record: wait for more bytes
        if (verify_crc) {
            uint32_t c0, c1;
            memcpy(&c0, b + pos + 8, 4);
            memcpy(&c1, b + pos + 12 + dl, 4);
            if (c0 != masked_crc(b + pos, 8) || c1 != masked_crc(b + pos + 12, (size_t) dl)) return CTR_E_INVALID;
        }
        if (!read_example(b + pos + 12, b + pos + 12 + dl, o, row)) return CTR_E_INVALID;
        pos += 16 + (int64_t) dl;
        ++row;
        *host_n_rows = row;
        *host_consumed = pos;
    }
    return CTR_OK;
}
