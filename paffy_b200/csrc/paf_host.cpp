// paf_host.cpp - host side of the add_mismatches path above the batch C ABI: PAF text in, PAF text out.
//
// Mirrors, for this path only, what the reference's command mains do around paf_encode_mismatches
// (impl/paf_add_mismatches.c:106-140, impl/paf_view.c:148-192, impl/paf_trim.c:113-127): read records,
// look the contigs up, encode (on the GPU, as one batch), paf_check, write - with the reference's record
// order, tag normalisation, failure ordering (records before a bad one are still written) and exit codes.
// Parsing restates paf_parse / cigar_parse (impl/paf.c:70-111, 137-209) and printing restates
// paf_write_to_buffer (impl/paf.c:317-389); the trim pass restates impl/paf.c:811-953. Nothing is copied.
// There is no CPU implementation of the base comparison here: every M op goes through the CUDA kernels.
#include "../../include/paffy_b200.h"

#include <algorithm>
#include <atomic>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace {

constexpr uint64_t LENM = (1ull << 56) - 1;
enum { OPM = 0, OPI = 1, OPD = 2, OPEQ = 3, OPX = 4 };
inline int op_code(uint64_t w) { return (int)(int8_t)(w >> 56); }
inline int64_t op_len(uint64_t w) { return (int64_t)(w << 8) >> 8; }
inline uint64_t mk_op(int64_t len, int code) { return ((uint64_t)len & LENM) | ((uint64_t)(uint8_t)code << 56); }

// one parsed record; names point into the input buffer
struct Rec {
    const char *qname = nullptr, *tname = nullptr;
    uint32_t qname_len = 0, tname_len = 0;
    int64_t qlen = 0, qs = 0, qe = 0, tlen = 0, ts = 0, te = 0, nmatch = 0, nbases = 0, mapq = 0;
    int64_t score = 0, tile_level = -1, chain_id = -1, chain_score = -1;
    bool same_strand = true, has_cigar = false;
    char type = 0;
    int8_t bad = 0;         // parse failure: 1 too few fields, 2 strand, 3 cigar letter, 4 tp value
    char bad_char = 0;
    int64_t op_begin = 0, op_end = 0;  // into the owning chunk's op vector, later into the global one
};

inline int64_t to_i64(const char *s, const char *e) {  // impl/paf.c:36-48: optional '-', digits, nothing checked
    bool neg = s < e && *s == '-';
    if (neg) s++;
    int64_t v = 0;
    for (; s < e && *s >= '0' && *s <= '9'; s++) v = v * 10 + (*s - '0');
    return neg ? -v : v;
}

// CIGAR text -> op words (impl/paf.c:70-111). Returns false on a letter other than MID=X.
bool parse_cigar(const char *s, const char *e, std::vector<uint64_t> &ops, char *bad) {
    while (s < e) {
        int64_t n = 0;
        for (; s < e && *s >= '0' && *s <= '9'; s++) n = n * 10 + (*s - '0');
        if (s == e) { *bad = 0; return false; }  // digits without a letter: the reference reads the NUL as a letter
        int code;
        switch (*s) {
            case 'M': code = OPM; break;
            case 'I': code = OPI; break;
            case 'D': code = OPD; break;
            case '=': code = OPEQ; break;
            case 'X': code = OPX; break;
            default: *bad = *s; return false;
        }
        ops.push_back(mk_op(n, code));
        s++;
    }
    return true;
}

// One line (no newline, '\r' already removed) -> Rec. Tabs delimit like strtok_r: empty fields do not exist.
void parse_line(const char *b, const char *e, Rec &r, std::vector<uint64_t> &ops) {
    const char *tok[12], *tok_e[12];
    const char *p = b;
    int nf = 0;
    auto next = [&](const char *&s0, const char *&s1) -> bool {
        while (p < e && *p == '\t') p++;
        if (p >= e) return false;
        s0 = p;
        while (p < e && *p != '\t') p++;
        s1 = p;
        return true;
    };
    for (; nf < 12; nf++)
        if (!next(tok[nf], tok_e[nf])) break;
    if (nf < 12) { r.bad = 1; return; }
    r.qname = tok[0]; r.qname_len = (uint32_t)(tok_e[0] - tok[0]);
    r.qlen = to_i64(tok[1], tok_e[1]); r.qs = to_i64(tok[2], tok_e[2]); r.qe = to_i64(tok[3], tok_e[3]);
    const char strand = tok[4][0];
    if (strand != '+' && strand != '-') { r.bad = 2; r.bad_char = strand; return; }
    r.same_strand = strand == '+';
    r.tname = tok[5]; r.tname_len = (uint32_t)(tok_e[5] - tok[5]);
    r.tlen = to_i64(tok[6], tok_e[6]); r.ts = to_i64(tok[7], tok_e[7]); r.te = to_i64(tok[8], tok_e[8]);
    r.nmatch = to_i64(tok[9], tok_e[9]); r.nbases = to_i64(tok[10], tok_e[10]); r.mapq = to_i64(tok[11], tok_e[11]);
    r.op_begin = r.op_end = (int64_t)ops.size();
    const char *s0, *s1;
    while (next(s0, s1)) {
        if (s1 - s0 < 5 || s0[2] != ':' || s0[4] != ':') continue;  // not shaped XX:T:value
        const char a = s0[0], c = s0[1];
        const char *v = s0 + 5;
        if (a == 't' && c == 'p') {
            r.type = v < s1 ? *v : 0;
            if (r.type != 'P' && r.type != 'S' && r.type != 'I') { r.bad = 4; return; }  // the reference asserts
        } else if (a == 'A' && c == 'S') r.score = to_i64(v, s1);
        else if (a == 't' && c == 'l') r.tile_level = to_i64(v, s1);
        else if (a == 'c' && c == 'n') r.chain_id = to_i64(v, s1);
        else if (a == 's' && c == '1') r.chain_score = to_i64(v, s1);
        else if (a == 'c' && c == 'g') {
            ops.resize((size_t)r.op_begin);  // a repeated cg tag replaces the earlier one
            r.has_cigar = v < s1;            // an empty value parses to no cigar at all (impl/paf.c:71)
            if (!parse_cigar(v, s1, ops, &r.bad_char)) { r.bad = 3; return; }
            r.op_end = (int64_t)ops.size();
        }
    }
}

inline char *put_i64(char *w, int64_t v) {  // impl/paf.c:10-34
    char tmp[24];
    int n = 0;
    uint64_t u = v < 0 ? 0 - (uint64_t)v : (uint64_t)v;
    if (v < 0) *w++ = '-';
    do { tmp[n++] = (char)('0' + u % 10); u /= 10; } while (u);
    while (n) *w++ = tmp[--n];
    return w;
}

// Everything paf_write_to_buffer prints before the CIGAR ops (impl/paf.c:320-368), "\tcg:Z:" included when
// the record has a cigar. Appends to `out`.
void render_prefix(const Rec &r, std::string &out) {
    char buf[512];
    char *w = buf;
    auto tab_i64 = [&](int64_t v) { *w++ = '\t'; w = put_i64(w, v); };
    out.append(r.qname, r.qname_len);
    tab_i64(r.qlen); tab_i64(r.qs); tab_i64(r.qe);
    *w++ = '\t'; *w++ = r.same_strand ? '+' : '-'; *w++ = '\t';
    out.append(buf, (size_t)(w - buf));
    out.append(r.tname, r.tname_len);
    w = buf;
    tab_i64(r.tlen); tab_i64(r.ts); tab_i64(r.te); tab_i64(r.nmatch); tab_i64(r.nbases); tab_i64(r.mapq);
    if (r.type != 0 || r.tile_level != -1) {
        memcpy(w, "\ttp:A:", 6); w += 6;
        *w++ = r.type ? r.type : (r.tile_level > 1 ? 'S' : 'P');
    }
    auto tag = [&](const char *name, int64_t v) { *w++ = '\t'; memcpy(w, name, 5); w += 5; w = put_i64(w, v); };
    if (r.score != INT_MAX) tag("AS:i:", r.score);  // AS:i:0 appears when the input had none
    if (r.tile_level != -1) tag("tl:i:", r.tile_level);
    if (r.chain_id != -1) tag("cn:i:", r.chain_id);
    if (r.chain_score != -1) tag("s1:i:", r.chain_score);
    if (r.has_cigar) { memcpy(w, "\tcg:Z:", 6); w += 6; }
    out.append(buf, (size_t)(w - buf));
}

void render_ops(const uint64_t *ops, int64_t n, std::string &out) {  // impl/paf.c:367-380
    char buf[32];
    for (int64_t k = 0; k < n; k++) {
        char *w = put_i64(buf, op_len(ops[k]));
        const int c = op_code(ops[k]);
        *w++ = (c >= 0 && c <= 4) ? "MID=X"[c] : 'N';
        out.append(buf, (size_t)(w - buf));
    }
}

// paf_check (impl/paf.c:427-461): 0 when valid, else the failing check in reference order and the start of its
// st_errAbort message (the printed record follows it).
int check_record(const Rec &r, const uint64_t *ops, int64_t n, std::string *msg) {
    char buf[160];
    auto fail = [&](int code, const char *text) { if (msg) *msg = text; return code; };
    if (r.qs < 0 || r.qs >= r.qlen) return fail(1, "Paf query start coordinates are invalid,");
    if (r.qs > r.qe || r.qe > r.qlen) return fail(2, "Paf query end coordinates are invalid,");
    if (r.ts < 0 || r.ts >= r.tlen) return fail(3, "Paf target start coordinates are invalid,");
    if (r.ts > r.te || r.te > r.tlen) return fail(4, "Paf target end coordinates are invalid,");
    if (r.has_cigar) {
        int64_t q = 0, t = 0;
        for (int64_t k = 0; k < n; k++) {
            if (op_code(ops[k]) != OPD) q += op_len(ops[k]);
            if (op_code(ops[k]) != OPI) t += op_len(ops[k]);
        }
        if (q != r.qe - r.qs) {
            snprintf(buf, sizeof buf, "Paf cigar alignment does not match query length: %lld vs. %lld", (long long)q, (long long)(r.qe - r.qs));
            return fail(5, buf);
        }
        if (t != r.te - r.ts) {
            snprintf(buf, sizeof buf, "Paf cigar alignment does not match target length: %lld vs. %lld", (long long)t, (long long)(r.te - r.ts));
            return fail(6, buf);
        }
    }
    return 0;
}

// ---- trim (impl/paf.c:811-953). An alignment view: the record's coordinates plus a window [lo, hi) of its ops,
// so dropping ops at either end is O(1) like the reference's cigar->start. -------------------------------------
struct Aln {
    Rec *r;
    uint64_t *ops;
    int64_t lo, hi;
    bool flipped;  // inverted view: query/target swapped, I/D swapped, op order reversed on '-' strands
};
inline bool is_match_code(int c) { return c == OPM || c == OPEQ; }
// identity as the reference computes it: float32 quotient of a float32-rounded numerator, widened to double
inline double ident32(int64_t m, int64_t mm) { return (double)((float)m / (float)(m + mm)); }

// k-th op of the view in walking order, with I/D swapped when flipped
inline uint64_t view_op(const Aln &a, int64_t k) {
    const bool backwards = a.flipped && !a.r->same_strand;
    const uint64_t w = a.ops[backwards ? a.hi - 1 - k : a.lo + k];
    if (!a.flipped) return w;
    const int c = op_code(w);
    return c == OPI ? mk_op(op_len(w), OPD) : c == OPD ? mk_op(op_len(w), OPI) : w;
}

// drop the first `count` ops of the view, moving the start coordinates (impl/paf.c:842-861)
void drop_front(Aln &a, int64_t count) {
    Rec &r = *a.r;
    for (int64_t k = 0; k < count; k++) {
        const uint64_t w = view_op(a, k);
        const int c = op_code(w);
        const int64_t n = op_len(w);
        // in the flipped view "target" is the record's query and vice versa
        if (!a.flipped) {
            if (c != OPI) r.ts += n;
            if (c != OPD) { if (r.same_strand) r.qs += n; else r.qe -= n; }
        } else {
            if (c != OPI) r.qs += n;                                          // flipped target = query
            if (c != OPD) { if (r.same_strand) r.ts += n; else r.te -= n; }   // flipped query = target
        }
    }
    if (a.flipped && !r.same_strand) a.hi -= count; else a.lo += count;
}

// impl/paf.c:863-904 with the scan of :811-840
void trim_prefix(Aln &a, float threshold_f, float identity_f, int64_t max_trim) {
    const double threshold = threshold_f, identity = identity_f;
    const int64_t n = a.hi - a.lo;
    int64_t m = 0, mm = 0, last_low = -1;
    for (int64_t k = 0; k < n; k++) {
        const uint64_t w = view_op(a, k);
        if (is_match_code(op_code(w))) m += op_len(w); else mm += op_len(w);
        if (max_trim >= 0 && m + mm > max_trim) break;
        if (ident32(m, mm) < threshold) last_low = k;
    }
    if (last_low < 0) return;
    int64_t sm = 0, smm = 0, keep_from = -1;
    for (int64_t k = last_low; k >= 0; k--) {
        const uint64_t w = view_op(a, k);
        if (is_match_code(op_code(w))) sm += op_len(w); else smm += op_len(w);
        if (ident32(sm, smm) >= identity) keep_from = k;
    }
    const int64_t count = keep_from >= 0 ? keep_from : last_low + 1;
    if (count > 0) drop_front(a, count);
}

// paf_trim_unreliable_tails (impl/paf.c:906-953). Returns false where the reference's final assert would fire.
bool trim_unreliable_tails(Rec &r, uint64_t *ops, int64_t &lo, int64_t &hi, float score_fraction, float max_fraction) {
    int64_t m = 0, mm = 0;
    for (int64_t k = lo; k < hi; k++) {
        if (is_match_code(op_code(ops[k]))) m += op_len(ops[k]); else mm += op_len(ops[k]);
    }
    const double identity = ident32(m, mm);
    const double threshold = identity - (identity * score_fraction);
    const int64_t max_trim = (int64_t)((m + mm) * max_fraction);
    Aln a{&r, ops, lo, hi, false};
    trim_prefix(a, (float)threshold, (float)identity, max_trim);
    a.flipped = true;
    trim_prefix(a, (float)threshold, (float)identity, max_trim);
    lo = a.lo; hi = a.hi;
    int64_t fm = 0, fmm = 0;
    for (int64_t k = lo; k < hi; k++) {
        if (is_match_code(op_code(ops[k]))) fm += op_len(ops[k]); else fmm += op_len(ops[k]);
    }
    return ident32(fm, fmm) >= identity;
}

// ---- reading: the whole input, split into lines the way the reference's line reader does (sonLibFile.c:42-71:
// '\r' dropped everywhere, a last line without '\n' still counts, a trailing empty line ends the input) ----------
bool slurp(FILE *fh, std::string &buf) {
    char tmp[1 << 16];
    size_t got;
    while ((got = fread(tmp, 1, sizeof tmp, fh)) > 0) buf.append(tmp, got);
    return !ferror(fh);
}

struct Parsed {
    std::vector<Rec> recs;
    std::vector<uint64_t> ops;
};

// parse lines [begin, end) of `text` (already free of '\r') on `threads` workers, keeping input order
void parse_all(const std::string &text, int threads, Parsed &out) {
    const size_t n = text.size();
    threads = std::max(1, std::min<int>(threads, (int)(n / (1 << 16)) + 1));
    std::vector<size_t> cut(threads + 1, n);
    cut[0] = 0;
    for (int t = 1; t < threads; t++) {
        size_t p = n / threads * t;
        const void *nl = p < n ? memchr(text.data() + p, '\n', n - p) : nullptr;
        cut[t] = nl ? (size_t)((const char *)nl - text.data()) + 1 : n;
    }
    for (int t = 1; t <= threads; t++) cut[t] = std::max(cut[t], cut[t - 1]);
    std::vector<Parsed> part(threads);
    auto work = [&](int t) {
        Parsed &P = part[t];
        const char *p = text.data() + cut[t], *e = text.data() + cut[t + 1];
        while (p < e) {
            const char *nl = (const char *)memchr(p, '\n', (size_t)(e - p));
            const char *le = nl ? nl : e;
            if (le == p && !nl) break;
            Rec r;
            parse_line(p, le, r, P.ops);
            P.recs.push_back(r);
            p = nl ? nl + 1 : e;
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < threads; t++) th.emplace_back(work, t);
    work(0);
    for (auto &x : th) x.join();
    size_t nr = 0, no = 0;
    for (auto &P : part) { nr += P.recs.size(); no += P.ops.size(); }
    out.recs.reserve(nr);
    out.ops.reserve(no);
    for (auto &P : part) {
        const int64_t shift = (int64_t)out.ops.size();
        for (Rec &r : P.recs) { r.op_begin += shift; r.op_end += shift; out.recs.push_back(r); }
        out.ops.insert(out.ops.end(), P.ops.begin(), P.ops.end());
    }
}

int hw_threads() {
    if (const char *e = getenv("PAFFY_THREADS")) { const int t = atoi(e); if (t > 0) return t; }
    const unsigned n = std::thread::hardware_concurrency();
    return n ? (int)std::min(n, 32u) : 4;
}

template <class F>
void parallel_for(int64_t n, int threads, F f) {  // f(begin, end, thread)
    threads = (int)std::max<int64_t>(1, std::min<int64_t>(threads, n / 4096 + 1));
    std::vector<std::thread> th;
    for (int t = 1; t < threads; t++) th.emplace_back(f, n * t / threads, n * (t + 1) / threads, t);
    f(0, n / threads, 0);
    for (auto &x : th) x.join();
}

}  // namespace

