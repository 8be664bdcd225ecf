// paf_api.cpp - the reference's paf.h library surface (inc/paf.h:81-269) on the reference's own structs, exported under
// the reference's names so that stPaf.a's users link against libpaffy_b200.so unchanged (include/paf_b200_compat.h).
//
// Everything here is O(#ops) / O(line) host work in the reference too; the base comparison (paf_encode_mismatches)
// is NOT here: it runs on the GPU (paf_compat.cpp -> pafgpu_encode_one). The functions are restated over 64-bit op
// words - a CigarRecord array IS an array of such words (length bits 0..55, op bits 56..63) - with the text helpers the
// file filter uses (paf_text.h); nothing is copied from impl/paf.c. Fatal errors follow st_errAbort
// (sonLibCommon.c:152-161): "ERROR: ...\n" on stderr, exit(1); the reference's live asserts abort().
#include "../../include/paf_b200_compat.h"
#include "paf_text.h"

#include <algorithm>
#include <cstdarg>
#include <cstdlib>

using namespace paftext;

namespace {

[[noreturn]] void die(const char *fmt, ...) {
    fputs("ERROR: ", stderr);
    va_list ap;
    va_start(ap, fmt);
    vfprintf(stderr, fmt, ap);
    va_end(ap);
    fputc('\n', stderr);
    exit(1);
}

[[noreturn]] void assert_failed(const char *what) {
    fprintf(stderr, "paffy_b200: Assertion `%s' failed.\n", what);
    abort();
}

void *xmalloc(size_t n) {
    void *p = malloc(n ? n : 1);
    if (!p) die("out of memory (%zu bytes)", n);
    return p;
}

char *copy_text(const char *s, size_t n) {
    char *d = (char *)xmalloc(n + 1);
    memcpy(d, s, n);
    d[n] = 0;
    return d;
}

// the live ops of a cigar as op words
inline uint64_t *words_of(Cigar *c) { return c ? reinterpret_cast<uint64_t *>(c->recs + c->start) : nullptr; }
inline int64_t count_of(const Cigar *c) { return c ? c->length : 0; }

Cigar *cigar_from_words(const uint64_t *w, size_t n) {
    Cigar *c = (Cigar *)xmalloc(sizeof(Cigar));
    c->recs = (CigarRecord *)xmalloc(n * sizeof(CigarRecord));
    if (n) memcpy(c->recs, w, n * sizeof(uint64_t));
    c->length = c->capacity = (int64_t)n;
    c->start = 0;
    return c;
}

void parse_failure(const Rec &r) {
    switch (r.bad) {
        case 2: die("Got an unexpected strand character (%c) in a paf string\n", r.bad_char);
        case 3: die("Got an unexpected character paf cigar string: %c\n", r.bad_char);
        case 4: assert_failed("paf->type == 'P' || paf->type == 'S' || paf->type == 'I'");
        default: die("Too few fields in a paf line (the reference dereferences NULL here)");
    }
}

// Paf fields -> the text record the renderer works on (names are borrowed)
Rec rec_of(const Paf *p) {
    Rec r;
    r.qname = p->query_name; r.qname_len = (uint32_t)strlen(p->query_name);
    r.tname = p->target_name; r.tname_len = (uint32_t)strlen(p->target_name);
    r.qlen = p->query_length; r.qs = p->query_start; r.qe = p->query_end;
    r.tlen = p->target_length; r.ts = p->target_start; r.te = p->target_end;
    r.nmatch = p->num_matches; r.nbases = p->num_bases; r.mapq = p->mapping_quality;
    r.score = p->score; r.tile_level = p->tile_level; r.chain_id = p->chain_id; r.chain_score = p->chain_score;
    r.same_strand = p->same_strand; r.type = p->type;
    r.has_cigar = p->cigar != nullptr || p->cigar_string != nullptr;
    return r;
}

// the line paf_write_to_buffer prints (impl/paf.c:317-389), newline included
void render(Paf *p, std::string &out) {
    out.clear();
    render_prefix(rec_of(p), out);
    if (p->cigar) render_ops(words_of(p->cigar), count_of(p->cigar), out);
    else if (p->cigar_string) out.append(p->cigar_string);
    out.push_back('\n');
}

// paf_estimate_buffer_size (impl/paf.c:391-396): the size callers' buffers are grown from
int64_t estimate(const Paf *p) {
    const int64_t cg = p->cigar_string ? (int64_t)strlen(p->cigar_string) : 12 * count_of(p->cigar);
    return cg + 300 + (int64_t)strlen(p->query_name) + (int64_t)strlen(p->target_name);
}

thread_local std::string t_line;

inline bool is_aligned_op(int c) { return c == OPM || c == OPEQ || c == OPX; }

// Identity of m matches among m + mm columns the way the reference computes it everywhere: float32 quotient of the
// float32-rounded numerator by the int64 sum converted to float32, widened to double (impl/paf.c:832, 886, 923).
inline double identity_of(int64_t m, int64_t mm) { return (double)((float)m / (float)(m + mm)); }

// Sums of matching (M, =) and other (X, I, D: "indels are treated as mismatches", impl/paf.c:826-828) op lengths
void trim_sums(const uint64_t *w, int64_t n, int64_t &m, int64_t &mm) {
    m = mm = 0;
    for (int64_t k = 0; k < n; k++) {
        const int c = op_code(w[k]);
        if (c == OPM || c == OPEQ) m += op_len(w[k]); else mm += op_len(w[k]);
    }
}

// Number of leading ops paf_trim_unreliable_prefix removes (impl/paf.c:863-897 with the scan of :811-840): the longest
// prefix whose running identity is below `threshold` (scan stops once more than max_trim columns were seen), shortened
// by its longest suffix with identity >= `identity`.
int64_t unreliable_prefix_ops(const uint64_t *w, int64_t n, double threshold, double identity, int64_t max_trim) {
    int64_t m = 0, mm = 0, last_low = -1;
    for (int64_t k = 0; k < n; k++) {
        const int c = op_code(w[k]);
        if (c == OPM || c == OPEQ) m += op_len(w[k]); else mm += op_len(w[k]);
        if (max_trim >= 0 && m + mm > max_trim) break;
        if (identity_of(m, mm) < threshold) last_low = k;
    }
    if (last_low < 0) return 0;
    int64_t sm = 0, smm = 0, keep_from = -1;
    for (int64_t k = last_low; k >= 0; k--) {
        const int c = op_code(w[k]);
        if (c == OPM || c == OPEQ) sm += op_len(w[k]); else smm += op_len(w[k]);
        if (identity_of(sm, smm) >= identity) keep_from = k;
    }
    return keep_from >= 0 ? keep_from : last_low + 1;
}

// drop the first `count` ops, moving the coordinates they covered (impl/paf.c:842-861)
void drop_leading_ops(Paf *p, int64_t count) {
    const uint64_t *w = words_of(p->cigar);
    for (int64_t k = 0; k < count; k++) {
        const int c = op_code(w[k]);
        const int64_t len = op_len(w[k]);
        if (c != OPI) p->target_start += len;
        if (c != OPD) { if (p->same_strand) p->query_start += len; else p->query_end -= len; }
    }
    p->cigar->start += count;
    p->cigar->length -= count;
}

// One end of the fixed trim (cigar_trim / cigar_trim_back, impl/paf.c:518-576): `bases` aligned columns leave the front
// (back = false) or the back of the cigar together with every insert / delete met on the way; the op the cut falls in is
// shortened in place.
void trim_end(Cigar *c, bool back, int64_t bases, int64_t *qc, int64_t *tc, int q_sign, int t_sign) {
    int64_t done = 0;
    while (c->length > 0) {
        uint64_t *w = reinterpret_cast<uint64_t *>(c->recs + c->start + (back ? c->length - 1 : 0));
        const int code = op_code(*w);
        const int64_t len = op_len(*w);
        if (is_aligned_op(code)) {
            if (done >= bases) break;
            if (done + len > bases) {
                const int64_t cut = bases - done;
                *w = mk_op(len - cut, code);
                *qc += q_sign * cut;
                *tc += t_sign * cut;
                if (!(len - cut > 0)) assert_failed("r->length > 0");
                break;
            }
            done += len;
            *qc += q_sign * len;
            *tc += t_sign * len;
        } else if (code == OPI) {
            *qc += q_sign * len;
        } else {
            if (code != OPD) assert_failed("r->op == query_delete");
            *tc += t_sign * len;
        }
        if (!back) c->start++;
        c->length--;
    }
}

}  // namespace

extern "C" {

void cigar_destruct(Cigar *c) {
    if (!c) return;
    free(c->recs);
    free(c);
}

void paf_destruct(Paf *paf) {
    free(paf->cigar_string);
    cigar_destruct(paf->cigar);
    free(paf->query_name);
    free(paf->target_name);
    free(paf);
}

Cigar *cigar_parse(char *cigar_string) {
    if (cigar_string[0] == 0) return nullptr;
    std::vector<uint64_t> ops;
    char bad = 0;
    if (!parse_cigar(cigar_string, cigar_string + strlen(cigar_string), ops, &bad))
        die("Got an unexpected character paf cigar string: %c\n", bad);
    return cigar_from_words(ops.data(), ops.size());
}

Paf *paf_parse(char *paf_string, bool parse_cigar_string) {
    Rec r;
    std::vector<uint64_t> ops;
    parse_line(paf_string, paf_string + strlen(paf_string), r, ops, parse_cigar_string);
    if (r.bad) parse_failure(r);
    Paf *p = (Paf *)calloc(1, sizeof(Paf));
    if (!p) die("out of memory");
    p->query_name = copy_text(r.qname, r.qname_len);
    p->target_name = copy_text(r.tname, r.tname_len);
    p->query_length = r.qlen; p->query_start = r.qs; p->query_end = r.qe;
    p->target_length = r.tlen; p->target_start = r.ts; p->target_end = r.te;
    p->num_matches = r.nmatch; p->num_bases = r.nbases; p->mapping_quality = r.mapq;
    p->score = r.score; p->tile_level = r.tile_level; p->chain_id = r.chain_id; p->chain_score = r.chain_score;
    p->same_strand = r.same_strand; p->type = r.type;
    if (parse_cigar_string) {
        if (r.has_cigar) p->cigar = cigar_from_words(ops.data() + r.op_begin, (size_t)(r.op_end - r.op_begin));
    } else if (r.cg) {
        p->cigar_string = copy_text(r.cg, r.cg_len);
    }
    return p;
}

// One line through the rules of the reference's reader (sonLibFile.c:42-71): a '\r' is replaced by the byte after it,
// '\n' or EOF ends the line, byte 0xff reads as EOF (the reader narrows getc to char), the buffer grows to 2n+1.
// Returns -1 at end of file, else the line length.
static int64_t read_line(FILE *fh, char **buf, int64_t *cap) {
    int64_t i = 0;
    for (;;) {
        int ch = (int)(char)getc_unlocked(fh);
        if (ch == '\r') ch = getc_unlocked(fh);
        if (i == *cap - 1) {
            *cap = 2 * *cap + 1;
            *buf = (char *)realloc(*buf, (size_t)*cap + 1);
            if (!*buf) die("out of memory");
        }
        if (ch == '\n' || ch == EOF) {
            (*buf)[i] = 0;
            return feof_unlocked(fh) ? -1 : i;
        }
        (*buf)[i++] = (char)ch;
    }
}

Paf *paf_read_with_buffer(FILE *fh, bool parse_cigar_string, char **paf_buffer, int64_t *paf_length_buffer) {
    const int64_t n = read_line(fh, paf_buffer, paf_length_buffer);
    if (n == -1 && (*paf_buffer)[0] == 0) return nullptr;
    return paf_parse(*paf_buffer, parse_cigar_string);
}

Paf *paf_read(FILE *fh, bool parse_cigar_string) {
    int64_t cap = 4096;
    char *buf = (char *)xmalloc((size_t)cap);
    Paf *p = paf_read_with_buffer(fh, parse_cigar_string, &buf, &cap);
    free(buf);
    return p;
}

Paf *paf_read2(FILE *fh) { return paf_read(fh, true); }

int64_t cigar_number_of_records(Paf *paf) { return count_of(paf->cigar); }

char *paf_print(Paf *paf) {
    render(paf, t_line);
    char *s = copy_text(t_line.data(), t_line.size());
    s[t_line.size() - 1] = 0;  // the newline
    return s;
}

void paf_write(Paf *paf, FILE *fh) {
    render(paf, t_line);
    fwrite(t_line.data(), 1, t_line.size(), fh);
}

void paf_write_with_buffer(Paf *paf, FILE *fh, char **paf_buffer, int64_t *paf_length_buffer) {
    render(paf, t_line);
    // the caller's buffer is grown by the reference's rule (impl/paf.c:399-403), and to the line if that is longer
    const int64_t need = std::max<int64_t>(estimate(paf), (int64_t)t_line.size());
    if (need > *paf_length_buffer) {
        *paf_length_buffer = need * 2 + 1;
        *paf_buffer = (char *)realloc(*paf_buffer, (size_t)*paf_length_buffer);
        if (!*paf_buffer) die("out of memory");
    }
    memcpy(*paf_buffer, t_line.data(), t_line.size());
    fwrite(*paf_buffer, 1, t_line.size(), fh);
}

void paf_stats_calc(Paf *paf, int64_t *matches, int64_t *mismatches, int64_t *query_inserts, int64_t *query_deletes,
                    int64_t *query_insert_bases, int64_t *query_delete_bases, bool zero_counts) {
    if (zero_counts) *matches = *mismatches = *query_inserts = *query_deletes = *query_insert_bases = *query_delete_bases = 0;
    const uint64_t *w = words_of(paf->cigar);
    for (int64_t k = 0, n = count_of(paf->cigar); k < n; k++) {
        const int64_t len = op_len(w[k]);
        switch (op_code(w[k])) {
            case OPM: case OPEQ: *matches += len; break;
            case OPX: *mismatches += len; break;
            case OPI: ++*query_inserts; *query_insert_bases += len; break;
            case OPD: ++*query_deletes; *query_delete_bases += len; break;
            default: assert_failed("c->op == query_delete");
        }
    }
}

void paf_pretty_print(Paf *paf, char *query_seq, char *target_seq, FILE *fh, bool include_alignment) {
    int64_t s[6];
    paf_stats_calc(paf, &s[0], &s[1], &s[2], &s[3], &s[4], &s[5], true);
    fprintf(fh, "Query:%s\tQ-start:%lld\tQ-length:%lld\tTarget:%s\tT-start:%lld\tT-length:%lld\tSame-strand:%i\tScore:%lld\t"
                "Identity:%f\tIdentity-with-gaps%f\tAligned-bases:%lld\tQuery-inserts:%lld\tQuery-deletes:%lld\n",
            paf->query_name, (long long)paf->query_start, (long long)(paf->query_end - paf->query_start), paf->target_name,
            (long long)paf->target_start, (long long)(paf->target_end - paf->target_start), (int)paf->same_strand, (long long)paf->score,
            (float)s[0] / (s[0] + s[1]), (float)s[0] / (s[0] + s[1] + s[4] + s[5]), (long long)(s[0] + s[1]), (long long)s[2],
            (long long)s[3]);
    if (!include_alignment) return;
    // three rows (target, query, '*' where the upper-cased bytes agree), 150 columns per block (impl/paf.c:282-314)
    const uint64_t *w = words_of(paf->cigar);
    const int64_t n = count_of(paf->cigar);
    std::string rows[3];
    int64_t q_used = 0, t_at = paf->target_start;
    auto complement = [](char ch) -> char {  // stString_reverseComplementChar (sonLibString.c:184-242): case kept
        const bool lower = ch >= 'a' && ch <= 'z';
        const char u = lower ? (char)(ch - 32) : ch;
        char v = u;
        switch (u) {
            case 'A': v = 'T'; break; case 'T': v = 'A'; break; case 'C': v = 'G'; break; case 'G': v = 'C'; break;
            case 'M': v = 'K'; break; case 'K': v = 'M'; break; case 'R': v = 'Y'; break; case 'Y': v = 'R'; break;
            case 'B': v = 'V'; break; case 'V': v = 'B'; break; case 'D': v = 'H'; break; case 'H': v = 'D'; break;
            default: break;
        }
        return lower ? (char)(v + 32) : v;
    };
    auto up = [](char ch) { return (ch >= 'a' && ch <= 'z') ? (char)(ch - 32) : ch; };
    for (int64_t k = 0; k < n; k++) {
        const int c = op_code(w[k]);
        for (int64_t l = 0, len = op_len(w[k]); l < len; l++) {
            char t = '-', q = '-';
            if (c != OPI) t = target_seq[t_at++];
            if (c != OPD) {
                q = paf->same_strand ? query_seq[paf->query_start + q_used] : complement(query_seq[paf->query_end - (q_used + 1)]);
                q_used++;
            }
            rows[0].push_back(t);
            rows[1].push_back(q);
            rows[2].push_back(up(t) == up(q) ? '*' : ' ');
        }
    }
    for (size_t at = 0; at < rows[0].size(); at += 150)
        for (const std::string &row : rows) {
            // (the reference prints with %s: a NUL byte inside a row would end it early; sequences hold letters)
            const size_t len = std::min<size_t>(150, row.size() - at);
            fwrite(row.data() + at, 1, strnlen(row.data() + at, len), fh);
            fputc('\n', fh);
        }
}

void paf_check(Paf *paf) {
    Rec r = rec_of(paf);
    r.has_cigar = paf->cigar != nullptr;
    std::string why;
    if (check_record(r, words_of(paf->cigar), count_of(paf->cigar), &why)) {
        char *line = paf_print(paf);
        die("%s %s", why.c_str(), line);
    }
}

void paf_invert(Paf *paf) {
    std::swap(paf->query_start, paf->target_start);
    std::swap(paf->query_end, paf->target_end);
    std::swap(paf->query_length, paf->target_length);
    std::swap(paf->query_name, paf->target_name);
    uint64_t *w = words_of(paf->cigar);
    const int64_t n = count_of(paf->cigar);
    for (int64_t k = 0; k < n; k++) {
        const int c = op_code(w[k]);
        if (c == OPI || c == OPD) w[k] = mk_op(op_len(w[k]), c == OPI ? OPD : OPI);
    }
    if (!paf->same_strand && n > 1) std::reverse(w, w + n);
}

int64_t paf_get_number_of_aligned_bases(Paf *paf) {
    const uint64_t *w = words_of(paf->cigar);
    int64_t total = 0;
    for (int64_t k = 0, n = count_of(paf->cigar); k < n; k++)
        if (is_aligned_op(op_code(w[k]))) total += op_len(w[k]);
    return total;
}

void paf_trim_ends(Paf *paf, int64_t end_bases_to_trim) {
    // the cigar is ordered along the target; on the '-' strand its front is the query's end
    if (paf->same_strand) {
        trim_end(paf->cigar, false, end_bases_to_trim, &paf->query_start, &paf->target_start, 1, 1);
        trim_end(paf->cigar, true, end_bases_to_trim, &paf->query_end, &paf->target_end, -1, -1);
    } else {
        trim_end(paf->cigar, false, end_bases_to_trim, &paf->query_end, &paf->target_start, -1, 1);
        trim_end(paf->cigar, true, end_bases_to_trim, &paf->query_start, &paf->target_end, 1, -1);
    }
}

void paf_trim_end_fraction(Paf *paf, float percentage) {
    if (!(percentage >= 0 && percentage <= 1.0)) assert_failed("percentage >= 0 && percentage <= 1.0");
    // int64 * float -> float, / 2.0 -> double, truncated (impl/paf.c:593)
    const float scaled = (float)paf_get_number_of_aligned_bases(paf) * percentage;
    paf_trim_ends(paf, (int64_t)((double)scaled / 2.0));
}

Paf *paf_shatter2(Paf *paf, int64_t query_start, int64_t target_start, int64_t length) {
    Paf *s = (Paf *)calloc(1, sizeof(Paf));
    if (!s) die("out of memory");
    s->query_name = copy_text(paf->query_name, strlen(paf->query_name));
    s->target_name = copy_text(paf->target_name, strlen(paf->target_name));
    s->query_length = paf->query_length; s->target_length = paf->target_length;
    s->query_start = query_start; s->query_end = query_start + length;
    s->target_start = target_start; s->target_end = target_start + length;
    s->same_strand = paf->same_strand;
    const uint64_t one = mk_op(length, OPM);
    s->cigar = cigar_from_words(&one, 1);
    s->score = paf->score; s->mapping_quality = paf->mapping_quality;
    s->num_matches = s->num_bases = length;
    s->tile_level = paf->tile_level; s->type = paf->type; s->chain_id = paf->chain_id;  // chain_score stays 0 (calloc), as in the reference
    paf_check(s);
    return s;
}

void paf_remove_mismatches(Paf *paf) {
    if (!paf->cigar) return;
    paf->cigar->length = merge_matches(words_of(paf->cigar), 0, paf->cigar->length);
}

void paf_trim_unreliable_prefix(Paf *paf, float identity_threshold, float identity, int64_t max_trim) {
    const int64_t count = unreliable_prefix_ops(words_of(paf->cigar), count_of(paf->cigar), (double)identity_threshold, (double)identity, max_trim);
    if (count > 0) drop_leading_ops(paf, count);
}

void paf_trim_unreliable_tails(Paf *paf, float score_fraction, float max_fraction_to_trim) {
    int64_t m, mm;
    trim_sums(words_of(paf->cigar), count_of(paf->cigar), m, mm);
    const double identity = identity_of(m, mm);
    const double threshold = identity - identity * score_fraction;  // double arithmetic on the promoted float (impl/paf.c:924)
    const int64_t max_trim = (int64_t)((float)(m + mm) * max_fraction_to_trim);
    // both values travel on as float parameters (impl/paf.c:863, 928)
    paf_trim_unreliable_prefix(paf, (float)threshold, (float)identity, max_trim);
    paf_invert(paf);  // the suffix is the prefix of the inverted record
    paf_trim_unreliable_prefix(paf, (float)threshold, (float)identity, max_trim);
    paf_invert(paf);
    int64_t fm, fmm;
    trim_sums(words_of(paf->cigar), count_of(paf->cigar), fm, fmm);
    if (!(identity_of(fm, fmm) >= identity)) assert_failed("final_identity >= identity");
}

}  // extern "C"
