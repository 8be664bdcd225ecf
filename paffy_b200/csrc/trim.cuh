// trim.cuh - paf_trim_unreliable_tails on the device (a8 of SURVEY.md section 8; impl/paf.c:811-953 restated, nothing
// copied): one thread per record walks the record's =/X/I/D ops with the reference's exact float32 arithmetic
// (float32 quotient of float32-rounded int64 values widened to double, explicit round-to-nearest intrinsics so that
// no FMA contraction changes a comparison) and leaves the kept op window [lo, hi) plus the moved coordinates.
// Trimming is op-granular and never splits an op, so the serialiser only has to skip the dropped ops.
#pragma once
#include "kernels.cuh"

namespace pafgpu {

struct TrimParams {
    int64_t n_records;
    const uint64_t *out;           // encoded ops of the batch
    const int64_t *rec_out_begin;  // n_records + 1
    const uint8_t *same_strand;
    int64_t *qs, *qe, *ts, *te;    // per record, updated in place
    int64_t *lo, *hi;              // kept ops: absolute indices into out
    uint8_t *failed;               // 1 where the reference's final assert (identity did not drop) would fire
    float score_fraction, max_fraction;
};

__device__ __forceinline__ double trim_identity(int64_t m, int64_t mm) {
    return (double)__fdiv_rn(__ll2float_rn(m), __ll2float_rn(m + mm));
}
__device__ __forceinline__ bool trim_is_match(uint64_t w) {
    const int c = (int)(w >> 56);
    return c == OP_M || c == OP_EQ;
}
__device__ __forceinline__ int64_t trim_len(uint64_t w) { return (int64_t)(w << 8) >> 8; }

// k-th op of the view in walking order; the inverted view (query and target swapped, impl/paf.c:469-490) walks a
// '-' strand record backwards and swaps I and D
__device__ __forceinline__ uint64_t trim_view_op(const uint64_t *ops, int64_t lo, int64_t hi, bool flipped, bool same_strand, int64_t k) {
    const uint64_t w = __ldg(ops + ((flipped && !same_strand) ? hi - 1 - k : lo + k));
    if (!flipped) return w;
    const int c = (int)(w >> 56);
    if (c == OP_I) return (w & LEN_MASK) | ((uint64_t)OP_D << 56);
    if (c == OP_D) return (w & LEN_MASK) | ((uint64_t)OP_I << 56);
    return w;
}

struct TrimRec {
    int64_t qs, qe, ts, te, lo, hi;
    bool same_strand;
};

// impl/paf.c:863-904 with the prefix scan of :811-840 and the coordinate moves of :842-861
__device__ __forceinline__ void trim_prefix(const uint64_t *ops, TrimRec &r, bool flipped, double threshold, double identity, int64_t max_trim) {
    const int64_t n = r.hi - r.lo;
    int64_t m = 0, mm = 0, last_low = -1;
    for (int64_t k = 0; k < n; k++) {
        const uint64_t w = trim_view_op(ops, r.lo, r.hi, flipped, r.same_strand, k);
        if (trim_is_match(w)) m += trim_len(w); else mm += trim_len(w);
        if (max_trim >= 0 && m + mm > max_trim) break;
        if (trim_identity(m, mm) < threshold) last_low = k;
    }
    if (last_low < 0) return;
    int64_t sm = 0, smm = 0, keep_from = -1;
    for (int64_t k = last_low; k >= 0; k--) {
        const uint64_t w = trim_view_op(ops, r.lo, r.hi, flipped, r.same_strand, k);
        if (trim_is_match(w)) sm += trim_len(w); else smm += trim_len(w);
        if (trim_identity(sm, smm) >= identity) keep_from = k;
    }
    const int64_t count = keep_from >= 0 ? keep_from : last_low + 1;
    for (int64_t k = 0; k < count; k++) {
        const uint64_t w = trim_view_op(ops, r.lo, r.hi, flipped, r.same_strand, k);
        const int c = (int)(w >> 56);
        const int64_t len = trim_len(w);
        if (!flipped) {
            if (c != OP_I) r.ts += len;
            if (c != OP_D) { if (r.same_strand) r.qs += len; else r.qe -= len; }
        } else {  // in the inverted view "target" is the record's query and vice versa
            if (c != OP_I) r.qs += len;
            if (c != OP_D) { if (r.same_strand) r.ts += len; else r.te -= len; }
        }
    }
    if (flipped && !r.same_strand) r.hi -= count; else r.lo += count;
}

__global__ void __launch_bounds__(128) trim_kernel(TrimParams P) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.n_records) return;
    TrimRec r;
    r.qs = P.qs[i]; r.qe = P.qe[i]; r.ts = P.ts[i]; r.te = P.te[i];
    r.lo = P.rec_out_begin[i]; r.hi = P.rec_out_begin[i + 1];
    r.same_strand = P.same_strand[i] != 0;
    bool failed = false;
    if (r.hi > r.lo) {
        int64_t m = 0, mm = 0;
        for (int64_t k = r.lo; k < r.hi; k++) {
            const uint64_t w = __ldg(P.out + k);
            if (trim_is_match(w)) m += trim_len(w); else mm += trim_len(w);
        }
        const double identity = trim_identity(m, mm);
        const double threshold = __dsub_rn(identity, __dmul_rn(identity, (double)P.score_fraction));
        const int64_t max_trim = __float2ll_rz(__fmul_rn(__ll2float_rn(m + mm), P.max_fraction));
        // the reference hands both values on as float parameters
        const double thr = (double)__double2float_rn(threshold), idn = (double)__double2float_rn(identity);
        trim_prefix(P.out, r, false, thr, idn, max_trim);
        trim_prefix(P.out, r, true, thr, idn, max_trim);
        int64_t fm = 0, fmm = 0;
        for (int64_t k = r.lo; k < r.hi; k++) {
            const uint64_t w = __ldg(P.out + k);
            if (trim_is_match(w)) fm += trim_len(w); else fmm += trim_len(w);
        }
        failed = !(trim_identity(fm, fmm) >= identity);
    }
    P.qs[i] = r.qs; P.qe[i] = r.qe; P.ts[i] = r.ts; P.te[i] = r.te;
    P.lo[i] = r.lo; P.hi[i] = r.hi;
    P.failed[i] = failed ? 1 : 0;
}

}  // namespace pafgpu
