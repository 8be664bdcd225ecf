/*
 * paf_b200_compat.h - the reference's `paf.h` library surface (inc/paf.h:52-269), exported by libpaffy_b200.so under the
 * reference's own names, struct layouts and ownership rules, so that code written against paffy's stPaf.a (the paffy
 * CLI mains, Cactus) links against this library unchanged. Each declaration cites the reference interface it replaces.
 *
 * Include EITHER the reference's inc/paf.h (which needs sonLib.h) OR this header: the type definitions are skipped
 * when inc/paf.h has been included first (its guard is ST_PAF_H_).
 *
 * What runs where: paf_encode_mismatches compares bases on the GPU (no CPU fallback: without a CUDA device it is fatal
 * like st_errAbort). Everything else here is O(#ops) or O(line) host work in the reference as well (SURVEY.md section 2
 * note) and is host C++ in this library. Errors are fatal the reference's way: "ERROR: <message>\n" on stderr and
 * exit(1) (sonLibCommon.c:152-161); failed asserts abort().
 *
 * The functions of paf.h that traffic in sonLib containers (read_pafs, write_pafs, paf_shatter, the coverage-array and
 * interval helpers: inc/paf.h:177-182, 209, 223-253) are in paffy_b200/csrc/paf_sonlib.c, shipped as the static
 * archive libpaffy_b200_stpaf.a: they need the host program's sonLib (stList / stHash), exactly like stPaf.a does.
 * paf_chain (inc/paf.h:187) lives in the reference's impl/chaining.c, which compiles against this surface as it is.
 */
#ifndef PAF_B200_COMPAT_H_
#define PAF_B200_COMPAT_H_
#include <stdbool.h>
#include <stdint.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef ST_PAF_H_ /* the reference header defines the same names */
/* inc/paf.h:52-58 */
typedef enum _cigarOp { match = 0, query_insert = 1, query_delete = 2, sequence_match = 3, sequence_mismatch = 4 } CigarOp;
/* inc/paf.h:61-64: one 64-bit word, length in bits 0..55, op in bits 56..63 (x86-64 / GCC bit-field order) */
typedef struct _cigar_record {
    int64_t length : 56;
    int64_t op : 8;
} CigarRecord;
/* inc/paf.h:67-73 */
typedef struct _cigar {
    CigarRecord *recs;
    int64_t length, start, capacity;
} Cigar;
/* inc/paf.h:75-76 */
static inline int64_t cigar_count(Cigar *c) { return c ? c->length : 0; }
static inline CigarRecord *cigar_get(Cigar *c, int64_t i) { return &c->recs[c->start + i]; }
/* inc/paf.h:88-109 */
typedef struct _paf {
    char *query_name;
    int64_t query_length, query_start, query_end;
    char *target_name;
    int64_t target_length, target_start, target_end;
    Cigar *cigar;
    char *cigar_string;
    int64_t score, mapping_quality, num_matches, num_bases, tile_level, chain_id, chain_score;
    bool same_strand;
    char type;
} Paf;

/* ---- text <-> record (inc/paf.h:81-161; impl/paf.c:50-230, 317-425) ------------------------------------------------ */
Cigar *cigar_parse(char *cigar_string);                  /* inc/paf.h:81, impl/paf.c:70-111; "" -> NULL */
void cigar_destruct(Cigar *cigar);                        /* inc/paf.h:86, impl/paf.c:50-55 */
void paf_destruct(Paf *paf);                              /* inc/paf.h:114, impl/paf.c:57-68 */
Paf *paf_parse(char *paf_string, bool parse_cigar_string); /* inc/paf.h:119, impl/paf.c:137-209 */
Paf *paf_read(FILE *fh, bool parse_cigar_string);         /* inc/paf.h:124, impl/paf.c:220-226 */
Paf *paf_read2(FILE *fh);                                 /* inc/paf.h:129, impl/paf.c:228-230 */
/* inc/paf.h:135, impl/paf.c:211-218 over the line reader sonLibFile.c:42-71 (may realloc *paf_buffer) */
Paf *paf_read_with_buffer(FILE *fh, bool parse_cigar_string, char **paf_buffer, int64_t *paf_length_buffer);
char *paf_print(Paf *paf);                                /* inc/paf.h:140, impl/paf.c:419-425; malloc'd, caller frees */
void paf_write(Paf *paf, FILE *fh);                       /* inc/paf.h:156, impl/paf.c:408-417 */
void paf_write_with_buffer(Paf *paf, FILE *fh, char **paf_buffer, int64_t *paf_length_buffer); /* inc/paf.h:161, impl/paf.c:398-406 */

/* ---- per-record arithmetic on the CIGAR (host, O(#ops)) ------------------------------------------------------------ */
void paf_stats_calc(Paf *paf, int64_t *matches, int64_t *mismatches, int64_t *query_inserts, int64_t *query_deletes,
                    int64_t *query_insert_bases, int64_t *query_delete_bases, bool zero_counts); /* inc/paf.h:145, impl/paf.c:236-260 */
void paf_pretty_print(Paf *paf, char *query_seq, char *target_seq, FILE *fh, bool include_alignment); /* inc/paf.h:151, impl/paf.c:269-315 */
void paf_check(Paf *paf);                                 /* inc/paf.h:167, impl/paf.c:427-461 */
void paf_invert(Paf *paf);                                /* inc/paf.h:172, impl/paf.c:469-490 */
int64_t paf_get_number_of_aligned_bases(Paf *paf);        /* inc/paf.h:194, impl/paf.c:507-516 */
void paf_trim_ends(Paf *paf, int64_t end_bases_to_trim);  /* inc/paf.h:199, impl/paf.c:578-587 */
void paf_trim_end_fraction(Paf *paf, float percentage);   /* inc/paf.h:204, impl/paf.c:589-598 */
void paf_remove_mismatches(Paf *paf);                     /* inc/paf.h:263, impl/paf.c:786-809 */
void paf_trim_unreliable_tails(Paf *paf, float score_fraction, float max_fraction_to_trim); /* inc/paf.h:269, impl/paf.c:906-953 */
#endif /* ST_PAF_H_ */

/* externally linked in the reference although not in its header */
int64_t cigar_number_of_records(Paf *paf);                /* impl/paf.c:232-234 */
Paf *paf_shatter2(Paf *paf, int64_t query_start, int64_t target_start, int64_t length); /* impl/paf.c:600-627 */
void paf_trim_unreliable_prefix(Paf *paf, float identity_threshold, float identity, int64_t max_trim); /* impl/paf.c:863-904 */

/* Replaces paf_encode_mismatches (inc/paf.h:258, impl/paf.c:739-784): rewrites every M op of paf->cigar into =/X
 * runs by comparing the two whole-contig sequences on the GPU. Frees the old recs array and installs a malloc'd one
 * (start = 0), exactly like the reference, so callers may keep using cigar_get / paf_destruct. cigar == NULL is a
 * no-op. A failure (no CUDA device, coordinates outside the sequences) is fatal like st_errAbort: message on
 * stderr, exit(1). Thread-safe (one process-wide context, serialised). Whole contigs are packed once per
 * (pointer, length, content fingerprint) and stay resident in HBM, so that the loop of `paffy view` / Cactus over the
 * records of a few contigs uploads each contig once. */
void paf_encode_mismatches(Paf *paf, char *query_seq, char *target_seq);

/* Same with the mismatch count of the record's M ops returned (what paf_stats_calc would add to `mismatches`). */
int64_t pafgpu_paf_encode_mismatches(Paf *paf, const char *query_seq, const char *target_seq);

#ifdef __cplusplus
}
#endif
#endif
