// paf_compat.cpp - paf_encode_mismatches with the reference's signature and ownership rules (impl/paf.c:739-784),
// running on the CUDA path through pafgpu_encode_one. No CPU comparison of bases happens here.
#include "../../include/paf_b200_compat.h"
#include "../../include/paffy_b200.h"

#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

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

extern "C" int64_t pafgpu_paf_encode_mismatches(Paf *paf, const char *query_seq, const char *target_seq) {
    if (!paf || !paf->cigar) return 0;
    Cigar *cg = paf->cigar;
    std::lock_guard<std::mutex> lk(g_mu);
    pafgpu_ctx *ctx = context();
    uint64_t *out = nullptr;
    int64_t n_out = 0, mism = 0;
    // a CigarRecord array is an array of op words (length in bits 0..55, op in bits 56..63)
    const uint64_t *in = reinterpret_cast<const uint64_t *>(cg->recs + cg->start);
    const int rc = pafgpu_encode_one(ctx, in, cg->length, paf->same_strand ? 1 : 0, paf->query_start, paf->query_end,
                                     paf->target_start, query_seq, target_seq, &out, &n_out, &mism);
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
