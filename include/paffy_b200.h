/*
 * paffy_b200.h - C ABI of the B200-native add_mismatches path (libpaffy_b200.so).
 *
 * Plain pointers and sizes only; no torch / C++ types cross this boundary. Every entry point cites
 * the reference interface it replaces (paths relative to the paffy source tree). The reference-side
 * binding a maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *   - A CIGAR op is one 64-bit word with the bit layout of the reference's CigarRecord
 *     (inc/paf.h:61-65): bits 0..55 length, bits 56..63 op; op codes per inc/paf.h:52-58
 *     (0 M, 1 I, 2 D, 3 =, 4 X). Host arrays of these words ARE CigarRecord arrays.
 *   - All functions return 0 on success or a negative PAFGPU_E* code; pafgpu_last_error() gives the
 *     message. There is no CPU fallback: without a CUDA device every call fails with
 *     PAFGPU_ENODEVICE.
 *   - "device" buffers are plain CUDA device pointers on the context's device.
 */
#ifndef PAFFY_B200_H_
#define PAFFY_B200_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PAFGPU_ABI_VERSION 2

enum {
    PAFGPU_OK = 0,
    PAFGPU_ENODEVICE = -1, /* no usable CUDA device / CUDA runtime error */
    PAFGPU_EINVAL = -2,    /* bad argument */
    PAFGPU_ENOCONTIG = -3, /* unknown contig name / id */
    PAFGPU_EBOUNDS = -4,   /* a record's CIGAR walks outside its contig (reference: undefined behaviour) */
    PAFGPU_EFASTA = -5,    /* inadmissible byte in FASTA (reference: st_errAbort, bioioC.c:196-199) */
    PAFGPU_EIO = -6,
    PAFGPU_ESTATE = -7     /* call out of order (e.g. fetch before submit) */
};

typedef struct pafgpu_ctx pafgpu_ctx;

int pafgpu_abi_version(void);
/* Number of CUDA devices visible (0 if none / no driver). */
int pafgpu_device_count(void);
/* Create a context bound to one CUDA device (one context per GPU; multi-GPU = one per rank). */
int pafgpu_create(int device, pafgpu_ctx **out);
void pafgpu_destroy(pafgpu_ctx *ctx);
/* A second context on the same GPU that SHARES the parent's packed genome (no copy): its own stream and work buffers, so
 * that the H2D copy of one batch, the kernels of another and the D2H copy of a third overlap on one device. The genome
 * must not be modified while siblings encode. (The reference has no counterpart: it is single-threaded.) */
int pafgpu_create_sibling(pafgpu_ctx *parent, pafgpu_ctx **out);
int pafgpu_device_of(const pafgpu_ctx *ctx);
/* The k-th helper context of `parent` on `device`: created on first use, owned by the parent (destroyed with it), its
 * genome kept equal to the parent's (shared on the same GPU, cloned GPU to GPU elsewhere). The file pipeline uses these
 * so that repeated calls do not pay context, stream and buffer creation again. */
int pafgpu_helper(pafgpu_ctx *parent, int device, int k, pafgpu_ctx **out);
const char *pafgpu_last_error(const pafgpu_ctx *ctx); /* ctx may be NULL: last creation error */

/* ---- genome store: replaces the name -> char* hash the reference fills with
 *      fastaReadToFunction + fastaRead_readToMapFunction (sonLib bioioC.c:155-232;
 *      impl/paf_add_mismatches.c:89-96) and queries with stHash_search (:116,:123). -------------------
 * Contigs are packed to 2 bits/base in HBM; bytes other than ACGT (any case) are kept exactly in a
 * sparse exception plane so the byte-compare semantics of impl/paf.c:755-757 are preserved. */
int pafgpu_genome_reserve(pafgpu_ctx *ctx, int64_t total_bases);
/* Add one contig from host ASCII (not NUL-terminated needed). A repeated name replaces the earlier
 * contig (sonLibHash.c:68-73). Returns the contig id (>=0) or a negative error. */
int64_t pafgpu_genome_add_contig(pafgpu_ctx *ctx, const char *name, const char *seq, int64_t len);
/* Same, sequence bytes already in device memory (synthetic genomes generated on the GPU). */
int64_t pafgpu_genome_add_contig_device(pafgpu_ctx *ctx, const char *name, const void *seq_dev, int64_t len);
/* Read a FASTA file with the reference reader's admission rules (whole header line is the name; '>'
 * anywhere ends a record; newline/space/tab skipped; any other non-letter except '-' is an error). */
int pafgpu_genome_load_fasta(pafgpu_ctx *ctx, const char *path);
int64_t pafgpu_genome_find(const pafgpu_ctx *ctx, const char *name); /* id or PAFGPU_ENOCONTIG */
int64_t pafgpu_genome_contig_length(const pafgpu_ctx *ctx, int64_t contig);
int64_t pafgpu_genome_num_contigs(const pafgpu_ctx *ctx);
/* Read bases [start, start+len) of a contig back as the upper-cased bytes the kernel compares. */
int pafgpu_genome_fetch(pafgpu_ctx *ctx, int64_t contig, int64_t start, int64_t len, char *out);

/* Replicate the packed genome of `src` into `dst` (another GPU of the box) with a device-to-device copy over NVLink /
 * NVSwitch (cudaMemcpyPeerAsync) instead of loading the FASTA once per GPU: the reference's multi-process recipe
 * (tests/paf_pipeline_test.sh:42-67) re-reads both FASTA files in every shard process. Same device: shares the store. */
int pafgpu_genome_clone(pafgpu_ctx *dst, pafgpu_ctx *src);
/* Persisted packed genome: write / read the packed arena + contig table, so that the next process of a
 * pipeline skips fastaReadToFunction (bioioC.c:155-232) altogether. `key` (may be NULL) is stored in the file and must
 * be equal at load time (callers put the FASTA paths, sizes and mtimes there); a differing key or a damaged file makes
 * the load fail (PAFGPU_ESTATE / PAFGPU_EIO) and the caller reads the FASTA instead. */
int pafgpu_genome_save(pafgpu_ctx *ctx, const char *path, const char *key);
int pafgpu_genome_load_packed(pafgpu_ctx *ctx, const char *path, const char *key);

/* ---- batch description: replaces the per-record fields paf_encode_mismatches reads from Paf
 *      (impl/paf.c:746,751-752: target_start, query_start, query_end, same_strand) and the
 *      Cigar array (inc/paf.h:67-73), flattened for a whole batch of records. ---------------------- */
typedef struct {
    int64_t n_records;
    int64_t n_ops;            /* == op_begin[n_records] */
    const int64_t *op_begin;  /* n_records + 1 offsets into ops; every record has >= 1 op */
    const uint64_t *ops;      /* n_ops CigarRecord words, records back to back; no zero-length M ops */
    const int32_t *q_contig;  /* per record: contig ids from pafgpu_genome_find */
    const int32_t *t_contig;
    const int64_t *q_start;   /* Paf.query_start */
    const int64_t *q_end;     /* Paf.query_end */
    const int64_t *t_start;   /* Paf.target_start */
    const uint8_t *same_strand; /* 1 for '+', 0 for '-' */
} pafgpu_batch;

typedef struct {
    int64_t n_out_ops;      /* total output CigarRecord words */
    int64_t aligned_bases;  /* sum of input M lengths (the unit of the throughput metric) */
    int64_t n_windows;      /* work windows (128 bases) scheduled */
    int64_t n_tiles;        /* warp tiles scheduled */
    int64_t retried;        /* 1 when the output buffer guess was too small and the encode pass ran twice */
    float ms_ingest;        /* CUDA-event time of the CIGAR ingest kernels */
    float ms_encode;        /* CUDA-event time of the whole encode stage: resets + compare + scan + emit + stitch */
    float ms_kernel;        /* CUDA-event time of encode_kernel alone (reads the sequences, writes the output ops: the roofline kernel) */
    float ms_emit;          /* CUDA-event time of stitch_kernel (closes runs left open across warp tiles) */
} pafgpu_result;

/* Run CIGAR ingest + the base-comparison kernel on one batch (a1 + a7 of SURVEY.md section 8).
 * _host: arrays are host memory; they are copied H2D (pinned staging) inside the call.
 * _device: arrays are device memory already (inputs resident in HBM).
 * Results stay on the device until fetched. Synchronous on return. */
int pafgpu_encode_batch_host(pafgpu_ctx *ctx, const pafgpu_batch *b, pafgpu_result *res);
int pafgpu_encode_batch_device(pafgpu_ctx *ctx, const pafgpu_batch *b, pafgpu_result *res);

/* Copy results of the last batch to host memory. Any pointer may be NULL to skip it.
 *   out_ops       n_out_ops words           (the =/X CIGARs, records back to back)
 *   out_op_begin  n_records + 1 offsets into out_ops
 *   rec_mismatch  n_records: number of mismatching aligned bases found in each record's M ops
 *                 (matches = sum of M lengths - this; with the I/D/=/X totals of the input CIGAR
 *                 this is everything paf_stats_calc, impl/paf.c:236-260, returns) */
int pafgpu_fetch_ops(pafgpu_ctx *ctx, uint64_t *out_ops, int64_t *out_op_begin, int64_t *rec_mismatch);
/* Device pointers to the same results (valid until the next batch on this context). */
int pafgpu_result_device_ptrs(pafgpu_ctx *ctx, const uint64_t **out_ops, const int64_t **out_op_begin,
                              const int64_t **rec_mismatch);

/* ---- device serialiser: replaces paf_write_to_buffer's CIGAR loop (impl/paf.c:367-380) and
 *      int64_to_str (:10-34) for the whole batch. The caller supplies, per record, the text that
 *      precedes the CIGAR ("...\tcg:Z:", fields 1-12 and tags rendered per :320-365); the device
 *      appends <len><op>... and '\n'. ------------------------------------------------------------- */
int pafgpu_serialise(pafgpu_ctx *ctx, const char *prefix_bytes, const int64_t *prefix_begin /* n_records+1 */,
                     int64_t *text_bytes_out, float *ms_out);
/* Copy the serialised text (text_bytes_out bytes) and optionally per-record line offsets (n_records+1). */
int pafgpu_fetch_text(pafgpu_ctx *ctx, char *text, int64_t *line_begin);

/* The same serialiser for callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc). This entry point has NO counterpart in the reference.
 * It needs no context and calls no cudaMalloc / cudaFree / allocating routine: its scratch (look-back
 * words, ticket counter, one record index per 2048 ops) is a fixed device array of 393 240 bytes that every GPU loading
 * the library holds until exit.
 *   limit   n_ops <= PAFGPU_NO_ALLOC_MAX_OPS (2^26 = 67 108 864 output ops per call); more -> PAFGPU_EINVAL with a message
 *           (pafgpu_last_error(NULL)). n_records < 2^31.
 *   inputs  (device memory of `device`) ops_dev[n_ops] packed CIGAR words, op_begin_dev[n_records + 1] offsets into them,
 *           prefix_dev / prefix_begin_dev[n_records + 1]: the text before each record's CIGAR, as for pafgpu_serialise
 *   outputs (device memory) text_dev[text_cap]: the lines; line_begin_dev[n_records + 1]: offset of every line, the last
 *           entry = bytes of the whole text. Nothing is written beyond text_cap.
 *   total_bytes  host pointer or NULL. Non-NULL: the call waits for `stream`, stores the byte total there and returns
 *           PAFGPU_EINVAL if it exceeds text_cap. NULL: nothing but asynchronous work on `stream` (a cudaStream_t, NULL =
 *           default stream); the caller reads line_begin_dev[n_records] itself.
 * All calls share the one scratch array (the part a call uses is cleared on its stream first): calls that overlap on
 * the same GPU - two streams, two threads - race. Serialise them. */
#define PAFGPU_NO_ALLOC_MAX_OPS (1ll << 26)
int pafgpu_serialise_no_alloc(int device, void *stream, const uint64_t *ops_dev, int64_t n_ops, const int64_t *op_begin_dev,
                              int64_t n_records, const char *prefix_dev, const int64_t *prefix_begin_dev, char *text_dev,
                              int64_t text_cap, int64_t *line_begin_dev, int64_t *total_bytes);

/* ---- PAF text in, PAF text out: replaces, for a chunk of whole lines, paf_read_with_buffer /
 *      paf_parse / cigar_parse (impl/paf.c:70-111, 137-218), the contig lookups (impl/paf_add_mismatches.c:116-127),
 *      paf_encode_mismatches, paf_check (impl/paf.c:427-461) and paf_write_to_buffer (impl/paf.c:317-389) of
 *      `paffy add_mismatches`: the raw bytes go to the device, are split into lines and fields, parsed, looked up, encoded
 *      and rendered there; the finished text is fetched with pafgpu_fetch_text. `text` is free of '\r'. A record the
 *      reference would abort on is only located (first_bad): the records before it are processed, the caller prints
 *      the reference's message for that line (bytes [bad_line_begin, bad_line_end) of `text`). ------------------------ */
typedef struct {
    int64_t records;        /* records processed (= lines before first_bad, or all) */
    int64_t records_in;     /* lines of the chunk */
    int64_t n_ops, n_out_ops, aligned_bases;
    int64_t text_bytes;     /* bytes of output text waiting on the device */
    int64_t first_bad;      /* -1, or the index of the first record the reference would abort on */
    int64_t bad_line_begin, bad_line_end;
    int64_t bad_kind;       /* 0: before the encode (parse, unknown contig, paf_check, walks outside the contig); after a trim: 1 = the
                               moved coordinates fail paf_check, 2 = the closing assert of paf_trim_unreliable_tails (impl/paf.c:952) */
    int64_t bad_coords[4];  /* bad_kind 1 / 2: query_start, query_end, target_start, target_end of that record after the trim */
    int64_t stats[7];       /* as pafgpu_file_report.stats, for the processed records */
    float ms_parse, ms_ingest, ms_kernel, ms_encode, ms_render; /* CUDA-event ms: text -> SoA, ingest, encode_kernel, encode stage, SoA -> text */
} pafgpu_text_result;
int pafgpu_add_mismatches_text(pafgpu_ctx *ctx, const char *text, int64_t n_bytes, pafgpu_text_result *res);
/* The same with the pass that follows the encode in the pipeline chosen by `mode`: PAFGPU_MODE_ADD (plain), PAFGPU_MODE_ADD_TRIM
 * (`| paffy trim -r trim_score_fraction -t trim_max_fraction`, or `-f -t trim_max_fraction` when trim_score_fraction < 0: the
 * trim and the paf_check behind it run on the device, the lines are printed with the moved coordinates and the kept ops),
 * PAFGPU_MODE_VIEW (`| paffy view -s -t`: only the totals in res->stats, no text). */
typedef struct {
    int mode;
    float trim_score_fraction, trim_max_fraction;
} pafgpu_text_opts;
int pafgpu_add_mismatches_text_opts(pafgpu_ctx *ctx, const char *text, int64_t n_bytes, const pafgpu_text_opts *opts,
                                    pafgpu_text_result *res);

/* ---- device trim: replaces paf_trim_unreliable_tails (inc/paf.h:269, impl/paf.c:906-953) for the whole batch, on the
 *      encoded ops still resident on the device. score_fraction / max_fraction are `paffy trim -r / -t`. t_end: host
 *      array of Paf.target_end per record. coords receives 4 * n_records values: query_start[], query_end[],
 *      target_start[], target_end[] after trimming. kept_ops (optional): ops left per record. failed (optional):
 *      1 where the reference's final assert (impl/paf.c:952) would fire. A following pafgpu_serialise prints only the
 *      kept ops; pafgpu_fetch_ops still returns all of them. ------------------------------------------------------- */
int pafgpu_trim(pafgpu_ctx *ctx, float score_fraction, float max_fraction, const int64_t *t_end, int64_t *coords,
                int64_t *kept_ops, uint8_t *failed);

/* Fixed trim, `paffy trim -f -t end_fraction` (paf_trim_end_fraction, inc/paf.h:204, impl/paf.c:589-598 with cigar_trim /
 * cigar_trim_back :518-576) for the whole batch on the device: (int64)((float)aligned * end_fraction / 2.0) aligned
 * bases leave each end together with the inserts / deletes met on the way; the op a cut falls in is shortened in place
 * on the device. Same coords / kept_ops outputs as pafgpu_trim; the following pafgpu_serialise prints the kept ops. */
int pafgpu_trim_fixed(pafgpu_ctx *ctx, float end_fraction, const int64_t *t_end, int64_t *coords, int64_t *kept_ops);

/* Stage a batch whose records are ALREADY encoded (the input of `paffy trim`, impl/paf_trim.c:113-127, which has no
 * FASTA): ops, op_begin, q_start, q_end, t_start, same_strand of `batch` (host memory; q_contig / t_contig ignored) go
 * to the device as if an encode call had produced them, so that pafgpu_trim / pafgpu_trim_fixed, pafgpu_serialise and
 * pafgpu_fetch_* work on them. No base is compared; per-record mismatch counts read 0. */
int pafgpu_stage_ops_host(pafgpu_ctx *ctx, const pafgpu_batch *batch);

/* ---- one-record convenience used by the paf.h-compatible paf_encode_mismatches
 *      (impl/paf.c:739-784): sequences are whole-contig host strings, indexed by absolute coordinate.
 *      Packs only the spans the record touches. out_ops is malloc'd (caller frees). ----------------- */
int pafgpu_encode_one(pafgpu_ctx *ctx, const uint64_t *ops, int64_t n_ops, int same_strand, int64_t query_start,
                      int64_t query_end, int64_t target_start, const char *query_seq, const char *target_seq,
                      uint64_t **out_ops, int64_t *n_out, int64_t *mismatches);

/* The same for a record whose contigs are already packed in the context's genome (ids from pafgpu_genome_add_contig /
 * pafgpu_genome_find): nothing but the record travels. This is what paf_encode_mismatches does after the first call with
 * a given sequence pointer. */
int pafgpu_encode_one_resident(pafgpu_ctx *ctx, const uint64_t *ops, int64_t n_ops, int same_strand, int64_t query_start,
                               int64_t query_end, int64_t target_start, int64_t q_contig, int64_t t_contig, uint64_t **out_ops,
                               int64_t *n_out, int64_t *mismatches);
/* Forget every contig of the genome store (the arena keeps its capacity); arena bases in use (both copies and pads). */
int pafgpu_genome_clear(pafgpu_ctx *ctx);
int64_t pafgpu_genome_total_bases(const pafgpu_ctx *ctx);

/* Kernel self-timing for bench.py: re-run the encode stage of the last submitted batch `reps` times
 * (inputs and genome resident in HBM). ms_stage: mean CUDA-event ms of resets + kernel + stitch;
 * ms_kernel: mean CUDA-event ms of the base-comparison kernel alone. */
int pafgpu_time_encode(pafgpu_ctx *ctx, int reps, float *ms_stage, float *ms_kernel);

/* Load balance of the last encode_kernel launch: the busiest SM's busy time and the mean over the SMs used (ms). */
int pafgpu_encode_balance(pafgpu_ctx *ctx, double *max_ms, double *mean_ms, int *sms_used);

/* Index (within the batch) of the record that made the last encode call fail with PAFGPU_EBOUNDS, else -1. */
int64_t pafgpu_last_bad_record(const pafgpu_ctx *ctx);

/* Pinned host memory for batch arrays and fetched text (cudaHostAlloc / cudaFreeHost): the _host entry points copy
 * at full PCIe/C2C speed only from such memory. */
void *pafgpu_host_alloc(int64_t bytes);
void pafgpu_host_free(void *p);

/* ---- whole-file filter: replaces the record loop of paffy_add_mismatches_main (impl/paf_add_mismatches.c:106-140)
 *      and, for mode 2, the loop of paffy_trim_main behind it (impl/paf_trim.c:113-127). Reads PAF text, runs every M
 *      op through the CUDA kernels in batches, writes PAF text in input order. A record that the reference would
 *      abort on (unknown contig, paf_check failure, bad strand / CIGAR letter) ends the run: every record before it
 *      has been written, the reference's message is printed to stderr and kept in pafgpu_last_error, and 1 is
 *      returned (the reference's exit code). Negative returns are library errors. ------------------------------- */
enum { PAFGPU_MODE_ADD = 0, PAFGPU_MODE_REMOVE = 1, PAFGPU_MODE_ADD_TRIM = 2,
       /* the record loop of paffy_trim_main alone (impl/paf_trim.c:113-127): no FASTA, no encode; the parsed ops are staged
        * on the device, trimmed there (identity trim, or the fixed trim when trim_score_fraction < 0: `-f -t
        * trim_max_fraction`), paf_check'ed and serialised */
       PAFGPU_MODE_TRIM = 3,
       /* the record loop of paffy_view_main (impl/paf_view.c:152-192): encode every record, then print per-alignment
        * stats lines (paf_pretty_print without the base-level alignment, impl/paf.c:270-279) and / or the aggregate line.
        * mode = PAFGPU_MODE_VIEW | PAFGPU_VIEW_PER_ALIGNMENT | PAFGPU_VIEW_AGGREGATE. No paf_check, like the reference. */
       PAFGPU_MODE_VIEW = 4, PAFGPU_VIEW_PER_ALIGNMENT = 1, PAFGPU_VIEW_AGGREGATE = 2 };
typedef struct {
    int mode;                  /* PAFGPU_MODE_*: add_mismatches, add_mismatches -a, add_mismatches | trim, view */
    float trim_score_fraction; /* paffy trim -r (reference default 0.05); < 0 selects the fixed trim `-f` (modes 2 and 3) */
    float trim_max_fraction;   /* paffy trim -t (reference default 1.0) */
    int threads;               /* host threads for parsing / rendering, 0 = all */
    int64_t batch_bases;       /* aligned bases per GPU batch, 0 = default */
    int gpus;                  /* GPUs of the box the job is sharded over (the context's own first), 0 = $PAFFY_GPUS (a count,
                                  "all", or a comma-separated device list) or 1. The packed genome is cloned GPU to GPU. */
    int streams;               /* contexts (stream + work buffers) per GPU so that copies and kernels overlap, 0 = $PAFFY_STREAMS or 2 */
    int64_t chunk_bytes;       /* bytes of PAF text per chunk (the unit that is read, dealt to a GPU and written in order), 0 = 48 MiB */
} pafgpu_file_opts;
typedef struct {
    int64_t records;           /* records written */
    int64_t aligned_bases;     /* sum of input M lengths of the written records */
    int64_t batches;
    int64_t failed_record;     /* input index of the record that ended the run, -1 if none */
    int64_t stats[7];          /* paf_stats_calc totals over the encoded records + record count, as `paffy view -s`
                                  accumulates them (impl/paf_view.c:176-178): matches, mismatches, query_inserts,
                                  query_deletes, query_insert_bases, query_delete_bases, records */
    double s_read, s_parse, s_gpu, s_host, s_write; /* seconds per stage, summed over the worker threads (they overlap) */
    double s_total;            /* wall seconds of the whole call */
    int64_t chunks, gpus, workers; /* chunks processed; GPUs and worker contexts used */
} pafgpu_file_report;
/* in_path / out_path NULL = stdin / stdout. The call streams: the input is read in chunks of whole lines, chunks are
 * processed concurrently (opts->streams contexts on each of opts->gpus GPUs) and written in input order, so memory is
 * bounded by the chunks in flight and output starts before the input ends, like the reference's record loop. */
int pafgpu_add_mismatches_file(pafgpu_ctx *ctx, const char *in_path, const char *out_path, const pafgpu_file_opts *opts,
                               pafgpu_file_report *report);

/* `paffy add_mismatches ... | paffy filter ...` in one pass (impl/paf_filter.c:26-176): the test of
 * paffy_filter_main (:123-134) is applied to every record behind the encode, on paf_stats_calc's counts taken from the
 * input op sums and the kernel's per-record mismatch tallies (float32 identity quotients, double thresholds); lines
 * that fail are not written (with `invert`, -x, only those are). Fields = the reference's options and defaults:
 * -s min_chain_score (-1), -t min_alignment_score (-1), -u min_identity (-1.0), -v min_identity_with_gaps (-1.0),
 * -w max_tile_level (-1 = off), -x invert (0). filter NULL = pafgpu_add_mismatches_file. mode must be PAFGPU_MODE_ADD.
 * report->stats still covers every encoded record. */
typedef struct {
    int64_t min_chain_score, min_alignment_score;
    double min_identity, min_identity_with_gaps;
    int64_t max_tile_level;
    int invert;
} pafgpu_filter;
int pafgpu_add_mismatches_file_filtered(pafgpu_ctx *ctx, const char *in_path, const char *out_path, const pafgpu_file_opts *opts,
                                        const pafgpu_filter *filter, pafgpu_file_report *report);

/* Drop-in for the reference's subcommand entry point (impl/paf_add_mismatches.c:26; paffy_main.c:52-77 dispatches to
 * it by name): same options (-i/--inputFile, -o/--outputFile, -a/--removeMismatches, -l/--logLevel, -h/--help,
 * FASTA files as positionals), same exit codes. Uses CUDA device $PAFFY_B200_DEVICE (default 0). */
int paffy_add_mismatches_main(int argc, char *argv[]);

/* Drop-in for paffy_view_main (impl/paf_view.c:39-215): options -i/--inputFile, -o/--outputFile, -s/--printAggregateStats,
 * -t/--noPerAlignmentStats, -u/--errorIfIdentityLowerThanX, -v/--errorIfAlignedBasesLowerThanX, -l/--logLevel, -h/--help,
 * at least one FASTA positional (else exit 1). Every record is encoded on the GPU exactly as add_mismatches would and
 * its stats printed from the kernel's tallies. -a/--includeAlignment (base-level rendering of the original-case FASTA
 * bytes) is refused with exit 1: the device keeps case-folded 2-bit genomes. */
int paffy_view_main(int argc, char *argv[]);

/* Drop-in for paffy_trim_main (impl/paf_trim.c:38-140): options -i/--inputFile, -o/--outputFile, -r/--trimIdentity
 * (default 0.05), -t/--trimFraction (default 1.0), -f/--fixedTrim, -l/--logLevel, -h/--help; no FASTA. Each batch of
 * parsed records is staged on the device (pafgpu_stage_ops_host), trimmed there (pafgpu_trim / pafgpu_trim_fixed),
 * checked like paf_check and printed by the device serialiser. */
int paffy_trim_main(int argc, char *argv[]);

/* ---- multi-GPU sharding: cut n input-ordered records with the given weights (aligned bases) into
 *      `parts` contiguous ranges of near-equal weight. cuts has parts + 1 entries, cuts[0] = 0, cuts[parts] = n. */
int pafgpu_partition(const int64_t *weights, int64_t n, int parts, int64_t *cuts);

#ifdef __cplusplus
}
#endif
#endif /* PAFFY_B200_H_ */
