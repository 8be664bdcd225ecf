/* paf_sonlib.c - the functions of the reference's paf.h that traffic in sonLib containers (inc/paf.h:177-182, 209,
 * 223-253; impl/paf.c:492-505, 629-737), over the library surface of libpaffy_b200.so. Shipped as the static archive
 * libpaffy_b200_stpaf.a and linked into the HOST program, which brings its own sonLib (stList / stHash / bioioC), exactly
 * as stPaf.a expects (Makefile:30-34 of the reference). Out of the hot path (needed only so that
 * the full paffy CLI links without impl/paf.c); restated, not copied.
 *
 * Only the prototypes of the sonLib functions used are declared here (sonLibList.h:40-85, sonLibHash.h:64-69,
 * bioioC.h:49-51), so the file builds without the sonLib headers; compile it with -DPAF_SONLIB_WITH_HEADERS and sonLib's
 * include path to check them against the real declarations.
 */
#ifdef PAF_SONLIB_WITH_HEADERS
#include "paf.h" /* the reference's own header (types, sonLib.h) */
#include "bioioC.h"
#else
#include <stdint.h>
typedef struct _stList stList;
typedef struct _stHash stHash;
stList *stList_construct3(int64_t size, void (*destructElement)(void *));
void stList_destruct(stList *list);
int64_t stList_length(stList *list);
void *stList_get(stList *list, int64_t index);
void stList_append(stList *list, void *item);
void *stList_peek(stList *list);
void *stList_pop(stList *list);
void stHash_insert(stHash *hash, void *key, void *value);
void *stHash_search(stHash *hash, void *key);
stList *fastaDecodeHeader(const char *fastaHeader);
char *fastaEncodeHeader(stList *attributes);
#endif
#include "../../include/paf_b200_compat.h"

#include <inttypes.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* inc/paf.h:214-218, 238-241: declared here because this header stands in for paf.h */
#ifndef ST_PAF_H_
typedef struct _sequenceCountArray {
    char *name;
    int64_t length;
    uint16_t *counts;
} SequenceCountArray;
typedef struct _interval {
    char *name;
    int64_t start, end, length;
} Interval;
#endif

static void live_assert(int ok, const char *what) {
    if (!ok) {
        fprintf(stderr, "paffy_b200: Assertion `%s' failed.\n", what);
        abort();
    }
}

static void destruct_paf(void *p) { paf_destruct((Paf *)p); }

/* inc/paf.h:177, impl/paf.c:492-499 */
stList *read_pafs(FILE *fh, bool parse_cigar_string) {
    stList *all = stList_construct3(0, destruct_paf);
    int64_t cap = 4096;
    char *line = malloc((size_t)cap);
    for (Paf *p; (p = paf_read_with_buffer(fh, parse_cigar_string, &line, &cap)) != NULL;) stList_append(all, p);
    free(line);
    return all;
}

/* inc/paf.h:182, impl/paf.c:501-505 */
void write_pafs(FILE *fh, stList *pafs) {
    int64_t cap = 4096;
    char *line = malloc((size_t)cap);
    const int64_t n = stList_length(pafs);
    for (int64_t i = 0; i < n; i++) paf_write_with_buffer((Paf *)stList_get(pafs, i), fh, &line, &cap);
    free(line);
}

/* inc/paf.h:209, impl/paf.c:629-663: one single-M record per M op of the cigar (target order; on the '-' strand the
 * query coordinate walks down from query_end) */
stList *paf_shatter(Paf *paf) {
    stList *parts = stList_construct3(0, destruct_paf);
    const int64_t n = cigar_count(paf->cigar);
    const int64_t step = paf->same_strand ? 1 : -1;
    int64_t q = paf->same_strand ? paf->query_start : paf->query_end, t = paf->target_start;
    for (int64_t k = 0; k < n; k++) {
        const CigarRecord *r = cigar_get(paf->cigar, k);
        const int64_t len = r->length;
        live_assert(len >= 1, "p->length >= 1");
        if (r->op == match) {
            /* the block's lowest query coordinate: where the walk stands ('+') or where it arrives ('-') */
            const int64_t q_lo = paf->same_strand ? q : q - len;
            stList_append(parts, paf_shatter2(paf, q_lo, t, len));
            q += step * len;
            t += len;
        } else if (r->op == query_insert) {
            q += step * len;
        } else {
            live_assert(r->op == query_delete, "p->op == query_delete");
            t += len;
        }
    }
    live_assert(t == paf->target_end, "target_coordinate == paf->target_end");
    live_assert(q == (paf->same_strand ? paf->query_end : paf->query_start), "query_coordinate == paf->query_end / query_start");
    return parts;
}

/* inc/paf.h:223, impl/paf.c:669-673 */
void sequenceCountArray_destruct(SequenceCountArray *a) {
    free(a->name);
    free(a->counts);
    free(a);
}

/* inc/paf.h:228, impl/paf.c:675-688 */
SequenceCountArray *get_alignment_count_array(stHash *by_name, Paf *paf) {
    SequenceCountArray *a = stHash_search(by_name, paf->query_name);
    if (a) {
        live_assert(a->length == paf->query_length, "seq_count_array->length == paf->query_length");
        return a;
    }
    a = calloc(1, sizeof *a);
    a->name = strdup(paf->query_name);
    a->length = paf->query_length;
    a->counts = calloc((size_t)(paf->query_length > 0 ? paf->query_length : 1), sizeof(uint16_t));
    stHash_insert(by_name, a->name, a);
    return a;
}

/* inc/paf.h:233, impl/paf.c:690-709: +1 (saturating below INT16_MAX) on every query base under an aligned column */
void increase_alignment_level_counts(SequenceCountArray *a, Paf *paf) {
    int64_t at = paf->query_start;
    const int64_t n = cigar_count(paf->cigar);
    for (int64_t k = 0; k < n; k++) {
        const CigarRecord *r = cigar_get(paf->cigar, k);
        if (r->op == query_delete) continue;
        const int64_t len = r->length;
        if (r->op != query_insert) {
            live_assert(r->op == match || r->op == sequence_match || r->op == sequence_mismatch, "aligned op");
            live_assert(at >= 0 && at + len <= paf->query_end && at + len <= paf->query_length && at + len <= a->length,
                        "aligned block inside the query");
            for (uint16_t *c = a->counts + at, *e = c + len; c < e; c++)
                if (*c < INT16_MAX - 1) ++*c;
        }
        at += len;
    }
    live_assert(at == paf->query_end, "i == paf->query_end");
}

/* inc/paf.h:243, impl/paf.c:711-714 */
void interval_destruct(Interval *i) {
    free(i->name);
    free(i);
}

/* inc/paf.h:248, impl/paf.c:716-731: "name|...|length|start" -> the last two attributes are numbers, the rest is the name */
Interval *decode_fasta_header(char *fasta_header) {
    Interval *iv = calloc(1, sizeof *iv);
    stList *attrs = fastaDecodeHeader(fasta_header);
    char *last = stList_pop(attrs);  /* the last attribute is the start, the one before it the length */
    live_assert(sscanf(last, "%" SCNi64, &iv->start) == 1, "j == 1");
    free(last);
    last = stList_pop(attrs);
    live_assert(sscanf(last, "%" SCNi64, &iv->length) == 1, "j == 1");
    free(last);
    iv->name = fastaEncodeHeader(attrs);
    stList_destruct(attrs);
    return iv;
}

/* inc/paf.h:253, impl/paf.c:733-737 */
int cmp_intervals(const void *a, const void *b) {
    const Interval *x = a, *y = b;
    const int by_name = strcmp(x->name, y->name);
    if (by_name) return by_name;
    return (x->start > y->start) - (x->start < y->start);
}
