// paf_compat.cpp - paf_encode_mismatches with the reference's signature and ownership rules (impl/paf.c:739-784),
// running on the CUDA path through pafgpu_encode_one. No CPU comparison of bases happens here.
#include "../../include/paf_b200_compat.h"
#include "../../include/paffy_b200.h"

#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>

static_assert(sizeof(CigarRecord) == 8, "CigarRecord must be one 64-bit word");
static_assert(sizeof(Cigar) == 32, "Cigar layout");
static_assert(sizeof(Paf) == 144, "Paf layout");
static_assert(offsetof(Paf, cigar) == 64 && offsetof(Paf, cigar_string) == 72 && offsetof(Paf, score) == 80, "Paf layout");
static_assert(offsetof(Paf, same_strand) == 136 && offsetof(Paf, type) == 137, "Paf layout");

namespace {
std::mutex g_mu;
pafgpu_ctx *g_ctx = nullptr;

pafgpu_ctx *context() {
    if (!g_ctx) {
        const char *dev = getenv("PAFFY_B200_DEVICE");
        if (pafgpu_create(dev ? atoi(dev) : 0, &g_ctx)) {
            fprintf(stderr, "ERROR: %s\n", pafgpu_last_error(nullptr));
            exit(1);
        }
    }
    return g_ctx;
}
}  // namespace

namespace {

// Whole contigs stay packed in HBM between calls, keyed by the caller's sequence pointer: `paffy view` and Cactus call
// paf_encode_mismatches record after record with the same few whole-contig strings (impl/paf_view.c:151-183), so only
// the first call with a pointer pays strlen + upload + pack; later calls upload the record alone. A fingerprint of the
// string's head guards against a pointer that was reused for another sequence.
struct CachedSeq {
    int64_t contig, len;
    uint64_t head;
};
std::unordered_map<const char *, CachedSeq> g_seqs;
int64_t g_serial = 0;
constexpr int64_t CACHE_BASES = (int64_t)24 << 30;  // arena bases (both strands + pads) before the cache starts over

uint64_t head_print(const char *s) {
    uint64_t h = 1469598103934665603ull;
    for (int i = 0; i < 256 && s[i]; i++) { h ^= (unsigned char)s[i]; h *= 1099511628211ull; }
    return h;
}

int64_t resident_contig(pafgpu_ctx *ctx, const char *seq) {
    const uint64_t head = head_print(seq);
    auto it = g_seqs.find(seq);
    if (it != g_seqs.end() && it->second.head == head) return it->second.contig;
    const int64_t len = (int64_t)strlen(seq);
    if (pafgpu_genome_total_bases(ctx) + 2 * len + 8192 > CACHE_BASES) {
        if (pafgpu_genome_clear(ctx)) return -1;
        g_seqs.clear();
    }
    char name[48];
    snprintf(name, sizeof name, "seq#%lld", (long long)g_serial++);
    const int64_t id = pafgpu_genome_add_contig(ctx, name, seq, len);
    if (id < 0) return id;
    g_seqs[seq] = CachedSeq{id, len, head};
    return id;
}

}  // namespace

extern "C" int64_t pafgpu_paf_encode_mismatches(Paf *paf, const char *query_seq, const char *target_seq) {
    if (!paf || !paf->cigar) return 0;
    Cigar *cg = paf->cigar;
    std::lock_guard<std::mutex> lk(g_mu);
    pafgpu_ctx *ctx = context();
    uint64_t *out = nullptr;
    int64_t n_out = 0, mism = 0;
    // a CigarRecord array is an array of op words (length in bits 0..55, op in bits 56..63)
    const uint64_t *in = reinterpret_cast<const uint64_t *>(cg->recs + cg->start);
    const int64_t tc = resident_contig(ctx, target_seq), qc = tc < 0 ? -1 : resident_contig(ctx, query_seq);
    // (the second lookup may have started the cache over: the target is looked up again)
    const int64_t tc2 = qc < 0 ? -1 : resident_contig(ctx, target_seq);
    const int rc = (tc < 0 || qc < 0 || tc2 < 0) ? -1
                                                 : pafgpu_encode_one_resident(ctx, in, cg->length, paf->same_strand ? 1 : 0, paf->query_start,
                                                                              paf->query_end, paf->target_start, qc, tc2, &out, &n_out, &mism);
    if (rc) {
        fprintf(stderr, "ERROR: paf_encode_mismatches: %s\n", pafgpu_last_error(ctx));
        exit(1);
    }
    free(cg->recs);
    cg->recs = reinterpret_cast<CigarRecord *>(out);
    cg->start = 0;
    cg->length = n_out;
    cg->capacity = n_out > 0 ? n_out : 1;
    return mism;
}

extern "C" void paf_encode_mismatches(Paf *paf, char *query_seq, char *target_seq) {
    pafgpu_paf_encode_mismatches(paf, query_seq, target_seq);
}
