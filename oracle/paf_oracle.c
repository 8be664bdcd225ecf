/*
 * paf_oracle.c - CPU restatement of the paffy add_mismatches hot path (see paf_oracle.h).
 *
 * TEST INFRASTRUCTURE ONLY: the checker for paffy_b200's CUDA kernels, never part of the product
 * path. Written from the behavioural description in SURVEY.md section 8a; each function cites the
 * reference lines whose behaviour it restates. Parity status: pinned (see header).
 */
#define _GNU_SOURCE
#include "paf_oracle.h"
#include <ctype.h>
#include <limits.h>
#include <stdlib.h>
#include <string.h>

#define LEN_MASK ((1ULL << 56) - 1)
static inline int64_t op_len(uint64_t w) { return (int64_t)(w << 8) >> 8; } /* signed 56-bit field */
static inline int op_code(uint64_t w) { return (int)(int8_t)(w >> 56); }
static inline uint64_t mk_op(int64_t len, int code) {
    return ((uint64_t)len & LEN_MASK) | ((uint64_t)(uint8_t)code << 56);
}

/* ---- complement predicate: sonLibString.c:184-242 ------------------------------------------------
 * Pairs A/T C/G M/K R/Y B/V D/H swap, everything else is a fixed point, case is kept. */
char oracle_revcomp_char(char c) {
    static const char from[] = "ACGTMKRYBVDH", to[] = "TGCAKMYRVBHD";
    int up = toupper((unsigned char)c);
    const char *hit = up ? strchr(from, up) : NULL;
    if (!hit) return c;
    char r = to[hit - from];
    return (c == up) ? r : (char)tolower((unsigned char)r);
}

/* ---- a1: impl/paf.c:739-784 -----------------------------------------------------------------------
 * M ops become alternating =/X runs, one run list per op (runs never join across ops, a zero-length M
 * emits nothing); every other op is copied and only advances the cursors. */
typedef struct { uint64_t *v; int64_t n, cap; } OpVec;
static void push(OpVec *o, uint64_t w) {
    if (o->n == o->cap) { o->cap = o->cap ? o->cap * 2 : 64; o->v = realloc(o->v, o->cap * sizeof(uint64_t)); }
    o->v[o->n++] = w;
}

int64_t oracle_encode_mismatches(const uint64_t *ops, int64_t n_ops, int same_strand,
                                 int64_t query_start, int64_t query_end, int64_t target_start,
                                 const char *query_seq, const char *target_seq, uint64_t **out) {
    OpVec o = {0};
    int64_t q_used = 0, t_at = target_start;
    for (int64_t k = 0; k < n_ops; k++) {
        int64_t n = op_len(ops[k]);
        int code = op_code(ops[k]);
        if (code != O_M) {
            push(&o, ops[k]);
            if (code != O_D) q_used += n;
            if (code != O_I) t_at += n;
            continue;
        }
        int64_t run = 0;
        int run_is_match = 0;
        for (int64_t i = 0; i < n; i++) {
            char tb = target_seq[t_at + i];
            char qb = same_strand ? query_seq[query_start + q_used + i]
                                  : oracle_revcomp_char(query_seq[query_end - 1 - q_used - i]);
            int same = toupper((unsigned char)tb) == toupper((unsigned char)qb);
            if (run && same == run_is_match) { run++; continue; }
            if (run) push(&o, mk_op(run, run_is_match ? O_EQ : O_X));
            run = 1;
            run_is_match = same;
        }
        if (run) push(&o, mk_op(run, run_is_match ? O_EQ : O_X));
        q_used += n;
        t_at += n;
    }
    if (!o.v) o.v = malloc(sizeof(uint64_t));
    *out = o.v;
    return o.n;
}

/* ---- a9: impl/paf.c:786-809 ---------------------------------------------------------------------- */
int64_t oracle_remove_mismatches(uint64_t *ops, int64_t n_ops) {
    int64_t w = 0;
    for (int64_t r = 0; r < n_ops; r++) {
        int code = op_code(ops[r]);
        int diag = code == O_M || code == O_EQ || code == O_X;
        if (diag && w > 0 && op_code(ops[w - 1]) == O_M)
            ops[w - 1] = mk_op(op_len(ops[w - 1]) + op_len(ops[r]), O_M);
        else
            ops[w++] = diag ? mk_op(op_len(ops[r]), O_M) : ops[r];
    }
    return w;
}

/* ---- a7: impl/paf.c:236-260 ---------------------------------------------------------------------- */
void oracle_stats(const uint64_t *ops, int64_t n_ops, int64_t c[6]) {
    for (int64_t k = 0; k < n_ops; k++) {
        int64_t n = op_len(ops[k]);
        switch (op_code(ops[k])) {
            case O_M: case O_EQ: c[0] += n; break;
            case O_X: c[1] += n; break;
            case O_I: c[2]++; c[4] += n; break;
            default:  c[3]++; c[5] += n; break; /* reference asserts this is D */
        }
    }
}

int64_t oracle_aligned_bases(const uint64_t *ops, int64_t n_ops) { /* impl/paf.c:507-516 */
    int64_t a = 0;
    for (int64_t k = 0; k < n_ops; k++) {
        int code = op_code(ops[k]);
        if (code == O_M || code == O_EQ || code == O_X) a += op_len(ops[k]);
    }
    return a;
}

/* ---- a4: impl/paf.c:70-111 ------------------------------------------------------------------------
 * Every byte that is not a digit must be one of MID=X; a missing number means length 0. */
int64_t oracle_cigar_parse(const char *s, uint64_t **out) {
    *out = NULL;
    if (!*s) return 0;
    OpVec o = {0};
    while (*s) {
        int64_t n = 0;
        for (; *s >= '0' && *s <= '9'; s++) n = n * 10 + (*s - '0');
        int code;
        switch (*s) {
            case 'M': code = O_M; break;
            case 'I': code = O_I; break;
            case 'D': code = O_D; break;
            case '=': code = O_EQ; break;
            case 'X': code = O_X; break;
            default: free(o.v); return -1;
        }
        push(&o, mk_op(n, code));
        s++;
    }
    *out = o.v;
    return o.n;
}

static int64_t parse_i64(const char *s) { /* impl/paf.c:36-48: optional '-', digits, no checking */
    int neg = *s == '-';
    int64_t v = 0;
    for (s += neg; *s >= '0' && *s <= '9'; s++) v = v * 10 + (*s - '0');
    return neg ? -v : v;
}

/* impl/paf.c:137-209. strtok_r semantics: runs of tabs collapse, so empty fields do not exist. */
int oracle_paf_parse(const char *line, OPaf *p) {
    memset(p, 0, sizeof *p);
    p->tile_level = p->chain_id = p->chain_score = -1;
    char *copy = strdup(line), *save = NULL;
    char *f[12];
    for (int i = 0; i < 12; i++) {
        f[i] = strtok_r(i ? NULL : copy, "\t", &save);
        if (!f[i]) { free(copy); return -1; } /* the reference has undefined behaviour here */
    }
    if (f[4][0] != '+' && f[4][0] != '-') { free(copy); return -1; }
    p->query_name = strdup(f[0]);
    p->query_length = parse_i64(f[1]); p->query_start = parse_i64(f[2]); p->query_end = parse_i64(f[3]);
    p->same_strand = f[4][0] == '+';
    p->target_name = strdup(f[5]);
    p->target_length = parse_i64(f[6]); p->target_start = parse_i64(f[7]); p->target_end = parse_i64(f[8]);
    p->num_matches = parse_i64(f[9]); p->num_bases = parse_i64(f[10]); p->mapping_quality = parse_i64(f[11]);
    int rc = 0;
    for (char *t; (t = strtok_r(NULL, "\t", &save));) {
        size_t tl = strlen(t);
        if (tl < 5 || t[2] != ':' || t[4] != ':') continue; /* reference reads t[2], t[4] unguarded */
        const char *v = t + 5;
        if (t[0] == 't' && t[1] == 'p') p->type = v[0];
        else if (t[0] == 'A' && t[1] == 'S') p->score = parse_i64(v);
        else if (t[0] == 't' && t[1] == 'l') p->tile_level = parse_i64(v);
        else if (t[0] == 'c' && t[1] == 'n') p->chain_id = parse_i64(v);
        else if (t[0] == 's' && t[1] == '1') p->chain_score = parse_i64(v);
        else if (t[0] == 'c' && t[1] == 'g') {
            free(p->ops);
            p->n_ops = oracle_cigar_parse(v, &p->ops);
            if (p->n_ops < 0) { rc = -1; p->n_ops = 0; }
            p->has_cigar = *v != '\0';
        }
    }
    free(copy);
    return rc;
}

void oracle_paf_free(OPaf *p) {
    free(p->query_name); free(p->target_name); free(p->ops);
    memset(p, 0, sizeof *p);
}

/* ---- a5: impl/paf.c:10-34, 317-389 --------------------------------------------------------------- */
static char *put_i64(char *w, int64_t v) {
    char tmp[24];
    int n = 0;
    uint64_t u = v < 0 ? 0 - (uint64_t)v : (uint64_t)v;
    if (v < 0) *w++ = '-';
    do { tmp[n++] = (char)('0' + u % 10); u /= 10; } while (u);
    while (n) *w++ = tmp[--n];
    return w;
}
static char *put_tag(char *w, const char *tag, int64_t v) {
    *w++ = '\t'; memcpy(w, tag, 5); w += 5;
    return put_i64(w, v);
}

int64_t oracle_paf_print(const OPaf *p, char **out) {
    size_t cap = 400 + strlen(p->query_name) + strlen(p->target_name) + 22 * (size_t)p->n_ops;
    char *b = malloc(cap), *w = b;
    const int64_t head[] = {p->query_length, p->query_start, p->query_end};
    const int64_t tail[] = {p->target_length, p->target_start, p->target_end, p->num_matches, p->num_bases,
                            p->mapping_quality};
    w = stpcpy(w, p->query_name);
    for (int i = 0; i < 3; i++) { *w++ = '\t'; w = put_i64(w, head[i]); }
    *w++ = '\t'; *w++ = p->same_strand ? '+' : '-'; *w++ = '\t';
    w = stpcpy(w, p->target_name);
    for (int i = 0; i < 6; i++) { *w++ = '\t'; w = put_i64(w, tail[i]); }
    if (p->type || p->tile_level != -1) {
        memcpy(w, "\ttp:A:", 6); w += 6;
        *w++ = p->type ? p->type : (p->tile_level > 1 ? 'S' : 'P');
    }
    if (p->score != INT_MAX) w = put_tag(w, "AS:i:", p->score); /* so AS:i:0 appears when absent */
    if (p->tile_level != -1) w = put_tag(w, "tl:i:", p->tile_level);
    if (p->chain_id != -1) w = put_tag(w, "cn:i:", p->chain_id);
    if (p->chain_score != -1) w = put_tag(w, "s1:i:", p->chain_score);
    if (p->has_cigar) {
        memcpy(w, "\tcg:Z:", 6); w += 6;
        for (int64_t k = 0; k < p->n_ops; k++) {
            w = put_i64(w, op_len(p->ops[k]));
            int c = op_code(p->ops[k]);
            *w++ = (c >= 0 && c <= 4) ? "MID=X"[c] : 'N';
        }
    }
    *w++ = '\n';
    *out = b;
    return w - b;
}

/* ---- a6: impl/paf.c:427-461 ---------------------------------------------------------------------- */
int oracle_paf_check(const OPaf *p) {
    if (p->query_start < 0 || p->query_start >= p->query_length) return 1;
    if (p->query_start > p->query_end || p->query_end > p->query_length) return 2;
    if (p->target_start < 0 || p->target_start >= p->target_length) return 3;
    if (p->target_start > p->target_end || p->target_end > p->target_length) return 4;
    if (p->has_cigar) {
        int64_t q = 0, t = 0;
        for (int64_t k = 0; k < p->n_ops; k++) {
            if (op_code(p->ops[k]) != O_D) q += op_len(p->ops[k]);
            if (op_code(p->ops[k]) != O_I) t += op_len(p->ops[k]);
        }
        if (q != p->query_end - p->query_start) return 5;
        if (t != p->target_end - p->target_start) return 6;
    }
    return 0;
}

/* ---- impl/paf.c:469-490 -------------------------------------------------------------------------- */
void oracle_paf_invert(OPaf *p) {
    int64_t t;
    char *s;
    t = p->query_start; p->query_start = p->target_start; p->target_start = t;
    t = p->query_end; p->query_end = p->target_end; p->target_end = t;
    t = p->query_length; p->query_length = p->target_length; p->target_length = t;
    s = p->query_name; p->query_name = p->target_name; p->target_name = s;
    for (int64_t k = 0; k < p->n_ops; k++) {
        int c = op_code(p->ops[k]);
        if (c == O_I || c == O_D) p->ops[k] = mk_op(op_len(p->ops[k]), c == O_I ? O_D : O_I);
    }
    if (!p->same_strand)
        for (int64_t a = 0, b = p->n_ops - 1; a < b; a++, b--) {
            uint64_t w = p->ops[a]; p->ops[a] = p->ops[b]; p->ops[b] = w;
        }
}

/* ---- a8: impl/paf.c:811-953 -----------------------------------------------------------------------
 * All identities are float32 quotients widened to double, exactly as the reference computes them. */
static inline double ident(int64_t m, int64_t mm) { return (double)((float)m / (float)(m + mm)); }
static inline int is_match_op(uint64_t w) { return op_code(w) == O_M || op_code(w) == O_EQ; }

/* drop the first `count` ops, moving the start coordinates (impl/paf.c:842-861) */
static void drop_front(OPaf *p, int64_t count) {
    for (int64_t k = 0; k < count; k++) {
        int c = op_code(p->ops[k]);
        int64_t n = op_len(p->ops[k]);
        if (c != O_I) p->target_start += n;
        if (c != O_D) { if (p->same_strand) p->query_start += n; else p->query_end -= n; }
    }
    memmove(p->ops, p->ops + count, (size_t)(p->n_ops - count) * sizeof(uint64_t));
    p->n_ops -= count;
}

/* impl/paf.c:863-904 with the scan of :811-840 inlined */
static void trim_prefix(OPaf *p, float threshold_f, float identity_f, int64_t max_trim) {
    double threshold = threshold_f, identity = identity_f; /* float params promoted on compare */
    int64_t m = 0, mm = 0, last_low = -1;
    for (int64_t k = 0; k < p->n_ops; k++) {
        if (is_match_op(p->ops[k])) m += op_len(p->ops[k]); else mm += op_len(p->ops[k]);
        if (max_trim >= 0 && m + mm > max_trim) break;
        if (ident(m, mm) < threshold) last_low = k;
    }
    if (last_low < 0) return;
    int64_t sm = 0, smm = 0, keep_from = -1;
    for (int64_t k = last_low; k >= 0; k--) {
        if (is_match_op(p->ops[k])) sm += op_len(p->ops[k]); else smm += op_len(p->ops[k]);
        if (ident(sm, smm) >= identity) keep_from = k;
    }
    int64_t count = keep_from >= 0 ? keep_from : last_low + 1;
    if (count > 0) drop_front(p, count);
}

void oracle_trim_unreliable_tails(OPaf *p, float score_fraction, float max_fraction_to_trim) {
    int64_t m = 0, mm = 0;
    for (int64_t k = 0; k < p->n_ops; k++) {
        if (is_match_op(p->ops[k])) m += op_len(p->ops[k]); else mm += op_len(p->ops[k]);
    }
    double identity = ident(m, mm);
    double threshold = identity - (identity * score_fraction);
    int64_t max_trim = (int64_t)((m + mm) * max_fraction_to_trim); /* int64 -> float multiply, truncated */
    /* the callee takes float parameters: both doubles are narrowed at the call (impl/paf.c:863,928) */
    trim_prefix(p, (float)threshold, (float)identity, max_trim);
    oracle_paf_invert(p);
    trim_prefix(p, (float)threshold, (float)identity, max_trim);
    oracle_paf_invert(p);
}

/* impl/paf.c:518-587: trim `n` aligned bases off each end, splitting an op when needed, and always
 * strip leading / trailing indels. */
static void trim_side(OPaf *p, int64_t n, int front, int64_t *qc, int qs, int64_t *tc, int ts) {
    int64_t done = 0;
    while (p->n_ops > 0) {
        uint64_t *r = &p->ops[front ? 0 : p->n_ops - 1];
        int c = op_code(*r);
        int diag = c == O_M || c == O_EQ || c == O_X;
        if (diag && done >= n) break;
        int64_t len = op_len(*r);
        if (diag) {
            if (done + len > n) {
                int64_t cut = n - done;
                *r = mk_op(len - cut, c);
                *qc += qs * cut; *tc += ts * cut;
                break;
            }
            done += len;
            *qc += qs * len; *tc += ts * len;
        } else if (c == O_I) *qc += qs * len;
        else *tc += ts * len;
        if (front) memmove(p->ops, p->ops + 1, (size_t)(p->n_ops - 1) * sizeof(uint64_t));
        p->n_ops--;
    }
}

void oracle_trim_ends(OPaf *p, int64_t n) {
    if (p->same_strand) {
        trim_side(p, n, 1, &p->query_start, 1, &p->target_start, 1);
        trim_side(p, n, 0, &p->query_end, -1, &p->target_end, -1);
    } else {
        trim_side(p, n, 1, &p->query_end, -1, &p->target_start, 1);
        trim_side(p, n, 0, &p->query_start, 1, &p->target_end, -1);
    }
}

void oracle_trim_end_fraction(OPaf *p, float fraction) { /* impl/paf.c:589-598 */
    int64_t aligned = oracle_aligned_bases(p->ops, p->n_ops);
    int64_t n = (int64_t)((double)((float)aligned * fraction) / 2.0);
    oracle_trim_ends(p, n);
}

/* ---- a3: bioioC.c:155-232, sonLibHash.c:68-73 ---------------------------------------------------- */
typedef struct { char *name, *seq; int64_t len; } OContig;
struct OFasta { OContig *c; int64_t n, cap; };

OFasta *oracle_fasta_new(void) { return calloc(1, sizeof(OFasta)); }
int64_t oracle_fasta_count(const OFasta *f) { return f->n; }

int oracle_fasta_add(OFasta *f, const char *name, const char *seq, int64_t len) {
    char *copy = malloc((size_t)len + 1);
    memcpy(copy, seq, (size_t)len);
    copy[len] = 0;
    for (int64_t i = 0; i < f->n; i++)
        if (!strcmp(f->c[i].name, name)) { /* last duplicate wins */
            free(f->c[i].seq); f->c[i].seq = copy; f->c[i].len = len;
            return 0;
        }
    if (f->n == f->cap) { f->cap = f->cap ? 2 * f->cap : 16; f->c = realloc(f->c, f->cap * sizeof(OContig)); }
    f->c[f->n++] = (OContig){strdup(name), copy, len};
    return 0;
}

int oracle_fasta_load(OFasta *f, const char *path) {
    FILE *fh = fopen(path, "r");
    if (!fh) return -1;
    fseek(fh, 0, SEEK_END);
    long sz = ftell(fh);
    rewind(fh);
    char *buf = malloc((size_t)sz + 1);
    if (fread(buf, 1, (size_t)sz, fh) != (size_t)sz) { fclose(fh); free(buf); return -1; }
    fclose(fh);
    char *seq = malloc((size_t)sz + 1);
    long i = 0;
    int rc = 0;
    while (i < sz && buf[i] != '>') i++; /* anything before the first '>' is ignored */
    while (i < sz && rc == 0) {
        long h0 = ++i; /* header = the whole rest of the line */
        while (i < sz && buf[i] != '\n') i++;
        char *name = strndup(buf + h0, (size_t)(i - h0));
        int64_t n = 0;
        for (i += (i < sz); i < sz && buf[i] != '>'; i++) { /* a '>' anywhere starts the next record */
            char ch = buf[i];
            if (ch == '\n' || ch == ' ' || ch == '\t') continue;
            if (!isalpha((unsigned char)ch) && ch != '-') { rc = -2; break; } /* includes '\r' */
            seq[n++] = ch;
        }
        if (rc == 0) oracle_fasta_add(f, name, seq, n);
        free(name);
    }
    free(seq);
    free(buf);
    return rc;
}

const char *oracle_fasta_get(const OFasta *f, const char *name, int64_t *len) {
    for (int64_t i = 0; i < f->n; i++)
        if (!strcmp(f->c[i].name, name)) { if (len) *len = f->c[i].len; return f->c[i].seq; }
    return NULL;
}

void oracle_fasta_free(OFasta *f) {
    for (int64_t i = 0; i < f->n; i++) { free(f->c[i].name); free(f->c[i].seq); }
    free(f->c);
    free(f);
}

/* ---- impl/paf_add_mismatches.c:106-140 (+ impl/paf_trim.c:116-127, impl/paf_view.c:176-178) ------ */
int oracle_add_mismatches_file(const OFasta *f, const char *in_path, const char *out_path, int mode,
                               float trim_r, float trim_t, int64_t stats_out[7], int64_t *aligned_out) {
    FILE *in = fopen(in_path, "r"), *out = fopen(out_path, "w");
    if (!in || !out) return 2;
    char *line = NULL;
    size_t cap = 0;
    ssize_t got;
    int rc = 0;
    int64_t aligned = 0;
    if (stats_out) memset(stats_out, 0, 7 * sizeof(int64_t));
    while (rc == 0 && (got = getline(&line, &cap, in)) >= 0) {
        /* line reader: sonLibFile.c:42-71 drops every '\r' and the final '\n' */
        ssize_t w = 0;
        for (ssize_t r = 0; r < got; r++) if (line[r] != '\r' && line[r] != '\n') line[w++] = line[r];
        line[w] = 0;
        if (w == 0 && feof(in)) break;
        OPaf p;
        if (oracle_paf_parse(line, &p)) { rc = 1; oracle_paf_free(&p); break; }
        if (mode == 1) {
            if (p.has_cigar) p.n_ops = oracle_remove_mismatches(p.ops, p.n_ops);
        } else {
            const char *q = oracle_fasta_get(f, p.query_name, NULL);
            const char *t = q ? oracle_fasta_get(f, p.target_name, NULL) : NULL;
            if (!q || !t) { rc = 1; oracle_paf_free(&p); break; }
            if (p.has_cigar) {
                for (int64_t k = 0; k < p.n_ops; k++) if (op_code(p.ops[k]) == O_M) aligned += op_len(p.ops[k]);
                uint64_t *enc;
                int64_t n = oracle_encode_mismatches(p.ops, p.n_ops, p.same_strand, p.query_start, p.query_end,
                                                     p.target_start, q, t, &enc);
                free(p.ops);
                p.ops = enc;
                p.n_ops = n;
            }
            if (stats_out) { oracle_stats(p.ops, p.n_ops, stats_out); stats_out[6]++; }
        }
        if (oracle_paf_check(&p)) rc = 1;
        if (rc == 0 && mode == 2) {
            oracle_trim_unreliable_tails(&p, trim_r, trim_t);
            if (oracle_paf_check(&p)) rc = 1;
        }
        if (rc == 0) {
            char *txt;
            int64_t n = oracle_paf_print(&p, &txt);
            fwrite(txt, 1, (size_t)n, out);
            free(txt);
        }
        oracle_paf_free(&p);
    }
    free(line);
    fclose(in);
    fclose(out);
    if (aligned_out) *aligned_out = aligned;
    return rc;
}
