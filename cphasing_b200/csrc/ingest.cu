// Host-side table decode for the extractors: the reference parses Pore-C tables / 4DN pairs with the
// multi-threaded polars CSV reader (/root/reference/cphasing/hypergraph.py:1706-1717, 904-909); this is the equivalent
// runtime piece here -- a multi-threaded TSV field parser that turns one block of text into the columnar
// u32 / u32 / i8 arrays the device kernels take, so that the Python layer can decode block k + 1 while block k crosses
// PCIe and is grouped on the GPU (cphasing_b200/ingest.py).  Host code only (no kernels); C-ABI like the rest.
//
//   fields: up to 8 selected columns per line, each of kind 0 = unsigned integer -> u32, 1 = contig NAME -> index through an
//           open-addressing table keyed by FNV-1a (names outside the table -> n_names, the inner join of :1435 drops them),
//           2 = integer -> i8 (saturating; MAPQ)
//   lines that start with '#' and empty lines are skipped; a line with too few columns or a non-numeric number counts as
//   malformed (*n_bad) and is skipped.
#include "common.cuh"
#include <string.h>
#include <thread>
#include <vector>

namespace {

constexpr int TSV_MAX_FIELDS = 8;

inline u64 fnv1a(const char* p, size_t n) {
    u64 h = 1469598103934665603ull;
    for (size_t i = 0; i < n; ++i) {
        h ^= (unsigned char)p[i];
        h *= 1099511628211ull;
    }
    return h;
}

struct NameTable {
    const char* blob;
    const u64* off;      // [n + 1]
    u32 n;
    const u32* slots;    // [size] name index + 1, 0 = empty; size is a power of two
    u32 mask;
    u32 lookup(const char* p, size_t len) const {
        u32 s = (u32)fnv1a(p, len) & mask;
        for (;;) {
            const u32 v = slots[s];
            if (v == 0) return n;
            const u32 i = v - 1;
            if (off[i + 1] - off[i] == len && memcmp(blob + off[i], p, len) == 0) return i;
            s = (s + 1) & mask;
        }
    }
};

struct Spec {
    int nf;
    int col[TSV_MAX_FIELDS];
    int kind[TSV_MAX_FIELDS];
    void* out[TSV_MAX_FIELDS];
    int last_col;
};

inline bool parse_uint(const char* p, const char* e, u64* v) {
    if (p == e) return false;
    u64 x = 0;
    for (; p < e; ++p) {
        const unsigned d = (unsigned)(*p - '0');
        if (d > 9) return false;
        x = x * 10 + d;
    }
    *v = x;
    return true;
}
inline bool parse_int(const char* p, const char* e, long long* v) {
    bool neg = false;
    if (p < e && (*p == '-' || *p == '+')) {
        neg = *p == '-';
        ++p;
    }
    u64 x;
    if (!parse_uint(p, e, &x)) return false;
    *v = neg ? -(long long)x : (long long)x;
    return true;
}

inline bool is_data_line(const char* p, const char* e) { return p < e && *p != '#' && !(e - p == 1 && *p == '\r'); }

// lines in [a, b): b is either the end of the buffer or just past a '\n'
size_t count_lines(const char* a, const char* b) {
    size_t n = 0;
    while (a < b) {
        const char* nl = (const char*)memchr(a, '\n', (size_t)(b - a));
        const char* e = nl ? nl : b;
        if (is_data_line(a, e)) ++n;
        a = nl ? nl + 1 : b;
    }
    return n;
}

// parses the lines of [a, b) into rows first.. of the outputs; returns the number of rows written, counts malformed lines
size_t parse_range(const char* a, const char* b, const Spec& sp, const NameTable& nt, size_t first, size_t* bad) {
    size_t row = first;
    while (a < b) {
        const char* nl = (const char*)memchr(a, '\n', (size_t)(b - a));
        const char* e = nl ? nl : b;
        const char* next = nl ? nl + 1 : b;
        if (e > a && e[-1] == '\r') --e;
        if (is_data_line(a, e)) {
            bool ok = true;
            int col = 0, f = 0;
            const char* p = a;
            while (f < sp.nf) {
                const char* t = (const char*)memchr(p, '\t', (size_t)(e - p));
                const char* fe = t ? t : e;
                if (col == sp.col[f]) {
                    if (sp.kind[f] == 0) {
                        u64 v;
                        ok = parse_uint(p, fe, &v) && v <= 0xffffffffull;
                        if (ok) ((u32*)sp.out[f])[row] = (u32)v;
                    } else if (sp.kind[f] == 1) {
                        ((u32*)sp.out[f])[row] = nt.lookup(p, (size_t)(fe - p));
                    } else {
                        long long v;
                        ok = parse_int(p, fe, &v);
                        if (ok) ((int8_t*)sp.out[f])[row] = (int8_t)(v > 127 ? 127 : (v < -128 ? -128 : v));
                    }
                    ++f;
                    if (!ok) break;
                }
                if (!t) break;
                p = t + 1;
                ++col;
            }
            if (ok && f == sp.nf) ++row;
            else ++*bad;
        }
        a = next;
    }
    return row - first;
}

}  // namespace

extern "C" u64 cph_fnv1a(const char* p, u64 len) { return fnv1a(p, (size_t)len); }

// Decodes one block of TSV text (whole lines; a final line without '\n' is taken as complete).
//   n_fields selected columns, in increasing column order: field_col[k], field_kind[k] (0 u32 / 1 name / 2 i8), out[k] with
//   room for `cap` rows (cph_tsv_count_rows gives an upper bound).  names_*: the contig-name table for kind 1 (may be null
//   when no field is a name).  *n_rows / *n_bad receive the rows written and the malformed lines skipped.
extern "C" int cph_tsv_count_rows(const char* buf, u64 len, int n_threads, u64* n_rows) {
    ARG_CHECK(n_rows && (len == 0 || buf));
    if (n_threads < 1) n_threads = 1;
    if ((u64)n_threads > len / (1 << 20) + 1) n_threads = (int)(len / (1 << 20) + 1);
    std::vector<const char*> cut(n_threads + 1);
    cut[0] = buf;
    cut[n_threads] = buf + len;
    for (int t = 1; t < n_threads; ++t) {
        const char* p = buf + len / n_threads * t;
        if (p < cut[t - 1]) p = cut[t - 1];
        const char* nl = (const char*)memchr(p, '\n', (size_t)(buf + len - p));
        cut[t] = nl ? nl + 1 : buf + len;
    }
    std::vector<size_t> cnt(n_threads, 0);
    std::vector<std::thread> th;
    for (int t = 1; t < n_threads; ++t) th.emplace_back([&, t] { cnt[t] = count_lines(cut[t], cut[t + 1]); });
    cnt[0] = count_lines(cut[0], cut[1]);
    for (auto& x : th) x.join();
    u64 tot = 0;
    for (size_t c : cnt) tot += c;
    *n_rows = tot;
    return CPH_OK;
}

extern "C" int cph_tsv_parse(const char* buf, u64 len, int n_threads, int n_fields, const int* field_col, const int* field_kind,
                             void* const* out, u64 cap, const char* names_blob, const u64* names_off, u32 n_names,
                             const u32* name_slots, u32 n_slots, u64* n_rows, u64* n_bad) {
    ARG_CHECK(n_rows && n_bad && (len == 0 || buf));
    ARG_CHECK(n_fields >= 1 && n_fields <= TSV_MAX_FIELDS && field_col && field_kind && out);
    Spec sp;
    sp.nf = n_fields;
    bool need_names = false;
    for (int k = 0; k < n_fields; ++k) {
        ARG_CHECK(field_kind[k] >= 0 && field_kind[k] <= 2 && out[k]);
        ARG_CHECK(k == 0 || field_col[k] > field_col[k - 1]);
        sp.col[k] = field_col[k];
        sp.kind[k] = field_kind[k];
        sp.out[k] = out[k];
        need_names |= field_kind[k] == 1;
    }
    NameTable nt{names_blob, names_off, n_names, name_slots, n_slots ? n_slots - 1 : 0};
    if (need_names) ARG_CHECK(names_blob && names_off && name_slots && n_slots >= 2 && (n_slots & (n_slots - 1)) == 0);
    if (n_threads < 1) n_threads = 1;
    if ((u64)n_threads > len / (1 << 20) + 1) n_threads = (int)(len / (1 << 20) + 1);
    std::vector<const char*> cut(n_threads + 1);
    cut[0] = buf;
    cut[n_threads] = buf + len;
    for (int t = 1; t < n_threads; ++t) {
        const char* p = buf + len / n_threads * t;
        if (p < cut[t - 1]) p = cut[t - 1];
        const char* nl = (const char*)memchr(p, '\n', (size_t)(buf + len - p));
        cut[t] = nl ? nl + 1 : buf + len;
    }
    // pass 1: data lines per range -> first output row of every range (an upper bound: malformed lines leave gaps that
    // pass 3 closes)
    std::vector<size_t> cnt(n_threads, 0), wrote(n_threads, 0), bad(n_threads, 0), first(n_threads + 1, 0);
    {
        std::vector<std::thread> th;
        for (int t = 1; t < n_threads; ++t) th.emplace_back([&, t] { cnt[t] = count_lines(cut[t], cut[t + 1]); });
        cnt[0] = count_lines(cut[0], cut[1]);
        for (auto& x : th) x.join();
    }
    for (int t = 0; t < n_threads; ++t) first[t + 1] = first[t] + cnt[t];
    if (first[n_threads] > cap) {
        cph_set_error(__FILE__, __LINE__, "cph_tsv_parse: output arrays too small for the lines of this block");
        return CPH_ERR_WORKSPACE;
    }
    {
        std::vector<std::thread> th;
        for (int t = 1; t < n_threads; ++t)
            th.emplace_back([&, t] { wrote[t] = parse_range(cut[t], cut[t + 1], sp, nt, first[t], &bad[t]); });
        wrote[0] = parse_range(cut[0], cut[1], sp, nt, first[0], &bad[0]);
        for (auto& x : th) x.join();
    }
    // pass 3 (only when some range skipped malformed lines): close the gaps, keeping the file order
    size_t pos = 0, nbad = 0;
    for (int t = 0; t < n_threads; ++t) {
        if (pos != first[t] && wrote[t]) {
            for (int k = 0; k < n_fields; ++k) {
                const size_t w = sp.kind[k] == 2 ? 1 : 4;
                memmove((char*)sp.out[k] + pos * w, (char*)sp.out[k] + first[t] * w, wrote[t] * w);
            }
        }
        pos += wrote[t];
        nbad += bad[t];
    }
    *n_rows = pos;
    *n_bad = nbad;
    return CPH_OK;
}
