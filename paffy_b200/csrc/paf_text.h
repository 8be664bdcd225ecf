// paf_text.h - PAF text <-> record helpers shared by the file filter (paf_host.cpp) and the paf.h library surface
// (paf_api.cpp). Restates paf_parse / cigar_parse (impl/paf.c:70-111, 137-209), paf_write_to_buffer (impl/paf.c:317-389),
// paf_check (impl/paf.c:427-461) and paf_remove_mismatches (impl/paf.c:786-809) on 64-bit op words (length in bits 0..55,
// op in bits 56..63: the bit layout of the reference's CigarRecord, inc/paf.h:61-65). Internal, not part of the ABI.
#pragma once
#include <climits>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace paftext {

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
    const char *cg = nullptr;          // value of the (last) cg tag, when the caller asked for the text only
    uint32_t cg_len = 0;
};

inline int64_t to_i64(const char *s, const char *e) {  // impl/paf.c:36-48: optional '-', digits, nothing checked
    bool neg = s < e && *s == '-';
    if (neg) s++;
    int64_t v = 0;
    for (; s < e && *s >= '0' && *s <= '9'; s++) v = v * 10 + (*s - '0');
    return neg ? -v : v;
}

// CIGAR text -> op words (impl/paf.c:70-111). Returns false on a letter other than MID=X.
inline bool parse_cigar(const char *s, const char *e, std::vector<uint64_t> &ops, char *bad) {
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
// parse_ops = false keeps the cg value as text (paf_parse's parse_cigar_string = 0, impl/paf.c:194-198).
inline void parse_line(const char *b, const char *e, Rec &r, std::vector<uint64_t> &ops, bool parse_ops = true) {
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
            if (!parse_ops) { r.cg = v; r.cg_len = (uint32_t)(s1 - v); continue; }
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
inline void render_prefix(const Rec &r, std::string &out) {
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

inline void render_ops(const uint64_t *ops, int64_t n, std::string &out) {  // impl/paf.c:367-380
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
inline int check_record(const Rec &r, const uint64_t *ops, int64_t n, std::string *msg) {
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

// paf_remove_mismatches (impl/paf.c:786-809): consecutive M / = / X ops become one M, across original op boundaries.
// Compacts ops[lo, hi) in place and returns the new hi.
inline int64_t merge_matches(uint64_t *ops, int64_t lo, int64_t hi) {
    int64_t w = lo;
    for (int64_t k = lo; k < hi; k++) {
        const int c = op_code(ops[k]);
        const bool diag = c == OPM || c == OPEQ || c == OPX;
        if (diag && w > lo && op_code(ops[w - 1]) == OPM) ops[w - 1] = mk_op(op_len(ops[w - 1]) + op_len(ops[k]), OPM);
        else ops[w++] = diag ? mk_op(op_len(ops[k]), OPM) : ops[k];
    }
    return w;
}

}  // namespace paftext
