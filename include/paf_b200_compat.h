/*
 * paf_b200_compat.h - the slice of the reference's paf.h surface (inc/paf.h) that the add_mismatches path touches,
 * re-declared with byte-identical layouts so that objects compiled against the reference header interoperate with
 * libpaffy_b200.so. Code that already includes the reference's own paf.h does NOT need this header: it only has to
 * link libpaffy_b200.so in place of impl/paf.c's paf_encode_mismatches.
 *
 * Layouts (x86-64, GCC; checked by static asserts in paf_compat.cpp and by tests/test_abi.py):
 *   CigarRecord  8 bytes: one int64 holding length in bits 0..55 and op in bits 56..63   (inc/paf.h:61-65)
 *   Cigar       32 bytes: recs, length, start, capacity                                   (inc/paf.h:68-74)
 *   Paf        144 bytes: field order of inc/paf.h:88-109 (cigar at 64, score at 80, same_strand at 136, type at 137)
 */
#ifndef PAF_B200_COMPAT_H_
#define PAF_B200_COMPAT_H_
#include <stdbool.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef PAF_H_ /* the reference header defines the same names */
typedef enum _cigarOp { match = 0, query_insert = 1, query_delete = 2, sequence_match = 3, sequence_mismatch = 4 } CigarOp;
typedef struct _cigar_record {
    int64_t length : 56;
    int64_t op : 8;
} CigarRecord;
typedef struct _cigar {
    CigarRecord *recs;
    int64_t length, start, capacity;
} Cigar;
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
#endif

/* Replaces paf_encode_mismatches (inc/paf.h:258, impl/paf.c:739-784): rewrites every M op of paf->cigar into =/X
 * runs by comparing the two whole-contig sequences on the GPU. Frees the old recs array and installs a malloc'd one
 * (start = 0), exactly like the reference, so callers may keep using cigar_get / paf_destruct. cigar == NULL is a
 * no-op. A failure (no CUDA device, coordinates outside the sequences) is fatal like st_errAbort: message on
 * stderr, exit(1). Thread-safe (one process-wide context, serialised). */
void paf_encode_mismatches(Paf *paf, char *query_seq, char *target_seq);

/* Same with the mismatch count of the record's M ops returned (what paf_stats_calc would add to `mismatches`). */
int64_t pafgpu_paf_encode_mismatches(Paf *paf, const char *query_seq, const char *target_seq);

#ifdef __cplusplus
}
#endif
#endif
