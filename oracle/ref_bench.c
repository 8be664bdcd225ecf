/*
 * ref_bench.c - timing driver for the UNMODIFIED reference paf_encode_mismatches (impl/paf.c:739-784).
 *
 * TEST / BENCH INFRASTRUCTURE ONLY (bench.py cpu_baseline and --impl reference). Links the reference's
 * own objects (oracle/Makefile `ref` target); everything measured is reference code: FASTA load
 * through fastaReadToFunction, paf_read, then paf_encode_mismatches on T pthreads (the function is
 * re-entrant; records are dealt round-robin). Prints one JSON line.
 *
 *   [REF_BENCH_REPS=n] ref_bench <threads> <in.paf> <fasta>... [--write out.paf]
 *   Each rep re-reads the PAF (untimed) and times the encode of every record.
 */
#include <pthread.h>
#include <time.h>
#include "paf.h"
#include "bioioC.h"

typedef struct { stList *pafs; stHash *seqs; int tid, nthreads; int64_t out_ops; } Job;

static double now(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

static void *worker(void *arg) {
    Job *j = arg;
    for (int64_t i = j->tid; i < stList_length(j->pafs); i += j->nthreads) {
        Paf *p = stList_get(j->pafs, i);
        char *q = stHash_search(j->seqs, p->query_name), *t = stHash_search(j->seqs, p->target_name);
        if (!q || !t) { fprintf(stderr, "ref_bench: missing contig for record %lld\n", (long long)i); exit(1); }
        paf_encode_mismatches(p, q, t);
        j->out_ops += cigar_count(p->cigar);
    }
    return NULL;
}

int main(int argc, char **argv) {
    if (argc < 4) { fprintf(stderr, "usage: ref_bench <threads> <in.paf> <fasta>... [--write out.paf]\n"); return 2; }
    int nthreads = atoi(argv[1]);
    if (nthreads < 1) nthreads = 1;
    const char *write_path = NULL;
    double t0 = now();
    stHash *seqs = stHash_construct3(stHash_stringKey, stHash_stringEqualKey, free, free);
    for (int a = 3; a < argc; a++) {
        if (!strcmp(argv[a], "--write") && a + 1 < argc) { write_path = argv[++a]; continue; }
        FILE *fh = fopen(argv[a], "r");
        if (!fh) { fprintf(stderr, "ref_bench: cannot open %s\n", argv[a]); return 2; }
        fastaReadToFunction(fh, seqs, fastaRead_readToMapFunction);
        fclose(fh);
    }
    double t_load = now() - t0;
    int reps = 1;
    const char *reps_env = getenv("REF_BENCH_REPS");
    if (reps_env) reps = atoi(reps_env) > 0 ? atoi(reps_env) : 1;
    pthread_t *th = malloc(sizeof(pthread_t) * nthreads);
    Job *jobs = calloc(nthreads, sizeof(Job));
    printf("{\"threads\": %d, \"load_s\": %.6f, \"reps\": [", nthreads, t_load);
    for (int rep = 0; rep < reps; rep++) {
        t0 = now();
        FILE *in = fopen(argv[2], "r");
        if (!in) { fprintf(stderr, "ref_bench: cannot open %s\n", argv[2]); return 2; }
        stList *pafs = read_pafs(in, 1);
        fclose(in);
        double t_parse = now() - t0;
        int64_t aligned = 0;
        for (int64_t i = 0; i < stList_length(pafs); i++) {
            Paf *p = stList_get(pafs, i);
            for (int64_t k = 0; k < cigar_count(p->cigar); k++)
                if (cigar_get(p->cigar, k)->op == match) aligned += cigar_get(p->cigar, k)->length;
        }
        t0 = now();
        for (int t = 0; t < nthreads; t++) {
            jobs[t] = (Job){pafs, seqs, t, nthreads, 0};
            pthread_create(&th[t], NULL, worker, &jobs[t]);
        }
        int64_t out_ops = 0;
        for (int t = 0; t < nthreads; t++) { pthread_join(th[t], NULL); out_ops += jobs[t].out_ops; }
        double t_encode = now() - t0;
        t0 = now();
        if (write_path && rep == 0) {
            FILE *out = fopen(write_path, "w");
            for (int64_t i = 0; i < stList_length(pafs); i++) { paf_check(stList_get(pafs, i)); paf_write(stList_get(pafs, i), out); }
            fclose(out);
        }
        double t_write = now() - t0;
        printf("%s{\"records\": %lld, \"aligned_bp\": %lld, \"out_ops\": %lld, \"parse_s\": %.6f, \"encode_s\": %.6f, \"write_s\": %.6f}",
               rep ? ", " : "", (long long)stList_length(pafs), (long long)aligned, (long long)out_ops, t_parse, t_encode, t_write);
        stList_destruct(pafs);
    }
    printf("]}\n");
    return 0;
}
