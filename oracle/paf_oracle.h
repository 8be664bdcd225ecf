/*
 * paf_oracle.h - CPU restatement of the paffy add_mismatches hot path.
 *
 * TEST INFRASTRUCTURE ONLY. This is the parity checker for paffy_b200's CUDA path. Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it. The
 * product (paffy_b200/) never links, imports or executes anything in oracle/.
 *
 * Parity status: PINNED. Every function here is checked (tests/test_oracle.py) against
 *   - the reference's own known-answer unit tests (tests/paf_unit_test.c, values restated in
 *     tests/golden/reference_unit_vectors.json), and
 *   - outputs of the unmodified reference compiled from /root/reference (oracle/_ref/paffy),
 *     stored as fixtures under tests/golden/ by tests/golden/make_golden.py.
 *
 * A CIGAR op is one 64-bit word laid out like the reference's CigarRecord bit-field
 * (inc/paf.h:61-65): bits 0..55 length, bits 56..63 op code, codes per inc/paf.h:52-58.
 */
#ifndef PAF_ORACLE_H_
#define PAF_ORACLE_H_
#include <stdint.h>
#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { O_M = 0, O_I = 1, O_D = 2, O_EQ = 3, O_X = 4 };

typedef struct {
    char *query_name, *target_name;
    int64_t query_length, query_start, query_end;
    int64_t target_length, target_start, target_end;
    int64_t num_matches, num_bases, mapping_quality;
    int64_t score, tile_level, chain_id, chain_score;
    int same_strand; /* 1 for '+', 0 for '-' */
    char type;       /* 'P','S','I' or 0 when the line had no tp tag */
    int has_cigar;   /* a cg tag with a non-empty value was present */
    uint64_t *ops;   /* has_cigar: n_ops words */
    int64_t n_ops;
} OPaf;

/* sonLibString.c:184-242 */
char oracle_revcomp_char(char c);

/* impl/paf.c:739-784. Returns the number of output ops written to *out (malloc'd, caller frees). */
int64_t oracle_encode_mismatches(const uint64_t *ops, int64_t n_ops, int same_strand,
                                 int64_t query_start, int64_t query_end, int64_t target_start,
                                 const char *query_seq, const char *target_seq, uint64_t **out);

/* impl/paf.c:786-809. In place; returns new op count. */
int64_t oracle_remove_mismatches(uint64_t *ops, int64_t n_ops);

/* impl/paf.c:236-260. counts = {matches, mismatches, query_inserts, query_deletes,
 * query_insert_bases, query_delete_bases}, accumulated (caller zeroes). */
void oracle_stats(const uint64_t *ops, int64_t n_ops, int64_t counts[6]);

/* impl/paf.c:507-516 */
int64_t oracle_aligned_bases(const uint64_t *ops, int64_t n_ops);

/* impl/paf.c:70-111. Returns op count, -1 on an unexpected op letter; *out malloc'd (NULL if empty). */
int64_t oracle_cigar_parse(const char *text, uint64_t **out);

/* impl/paf.c:137-209. Parses one PAF line (no trailing newline) into *p. Returns 0, or -1 on a bad
 * strand / cigar. The line is not modified. */
int oracle_paf_parse(const char *line, OPaf *p);
void oracle_paf_free(OPaf *p);

/* impl/paf.c:317-389. Writes the record + '\n' into a malloc'd buffer; returns its length. */
int64_t oracle_paf_print(const OPaf *p, char **out);

/* impl/paf.c:427-461. Returns 0 if valid, else 1..6 for the failing check in reference order. */
int oracle_paf_check(const OPaf *p);

/* impl/paf.c:469-490 */
void oracle_paf_invert(OPaf *p);

/* impl/paf.c:906-953 (and 811-904). Modifies coordinates and op array in place. */
void oracle_trim_unreliable_tails(OPaf *p, float score_fraction, float max_fraction_to_trim);

/* impl/paf.c:578-598 */
void oracle_trim_ends(OPaf *p, int64_t end_bases_to_trim);
void oracle_trim_end_fraction(OPaf *p, float fraction);

/* bioioC.c:155-232 + sonLibHash.c:68-73: load FASTA files into a name->sequence table (whole header
 * line is the name; later duplicates replace earlier ones). Returns NULL on an inadmissible byte. */
typedef struct OFasta OFasta;
OFasta *oracle_fasta_new(void);
int oracle_fasta_load(OFasta *f, const char *path);
int oracle_fasta_add(OFasta *f, const char *name, const char *seq, int64_t len);
const char *oracle_fasta_get(const OFasta *f, const char *name, int64_t *len);
int64_t oracle_fasta_count(const OFasta *f);
void oracle_fasta_free(OFasta *f);

/* impl/paf_add_mismatches.c:106-140 as a file-to-file filter. mode 0: add mismatches, 1: remove (-a),
 * 2: add mismatches then trim_unreliable_tails(trim_r, trim_t) (== `add_mismatches | trim`).
 * If stats_out != NULL the six paf_stats_calc counters summed over all (encoded) records plus the
 * record count are stored (7 values) as `view -s` would accumulate them (impl/paf_view.c:176-178).
 * Returns 0, 1 when a record fails paf_check or names an unknown contig (earlier records are
 * already written, like the reference), 2 on I/O trouble. aligned_out: sum of input M lengths. */
int oracle_add_mismatches_file(const OFasta *f, const char *in_path, const char *out_path, int mode,
                               float trim_r, float trim_t, int64_t stats_out[7], int64_t *aligned_out);

#ifdef __cplusplus
}
#endif
#endif
