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

// ---- the same trim with ONE WARP per record, for long CIGARs (config C3: ~14 000 ops per record would be walked five
// times by a single thread): the running sums become warp scans over 32 ops at a time, the reference's "last index
// whose prefix identity is low" / "lowest index whose suffix identity is high enough" become ballots. The sums are
// exact integers either way, so every float32 quotient and every comparison is the one trim_kernel makes.
__device__ __forceinline__ void trim_scan2(int64_t &a, int64_t &b, unsigned lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int64_t ta = __shfl_up_sync(FULL, a, o), tb = __shfl_up_sync(FULL, b, o);
        if (lane >= (unsigned)o) { a += ta; b += tb; }
    }
}

__device__ __forceinline__ void trim_prefix_warp(const uint64_t *ops, TrimRec &r, bool flipped, double threshold, double identity, int64_t max_trim,
                                                 unsigned lane) {
    const int64_t n = r.hi - r.lo;
    int64_t cm = 0, cmm = 0, last_low = -1;
    for (int64_t base = 0; base < n; base += 32) {
        const int64_t k = base + lane;
        const bool valid = k < n;
        const uint64_t w = valid ? trim_view_op(ops, r.lo, r.hi, flipped, r.same_strand, k) : 0ull;
        int64_t a = valid && trim_is_match(w) ? trim_len(w) : 0, b = valid && !trim_is_match(w) ? trim_len(w) : 0;
        trim_scan2(a, b, lane);
        a += cm; b += cmm;
        const unsigned over = __ballot_sync(FULL, valid && max_trim >= 0 && a + b > max_trim);
        const unsigned first_over = over ? (unsigned)__ffs(over) - 1u : 32u;
        const unsigned low = __ballot_sync(FULL, valid && lane < first_over && trim_identity(a, b) < threshold);
        if (low) last_low = base + (31 - __clz(low));
        if (over) break;
        cm = __shfl_sync(FULL, a, 31); cmm = __shfl_sync(FULL, b, 31);
    }
    if (last_low < 0) return;
    int64_t keep_from = -1;
    cm = cmm = 0;
    for (int64_t top = last_low; top >= 0; top -= 32) {
        const int64_t k = top - lane;
        const bool valid = k >= 0;
        const uint64_t w = valid ? trim_view_op(ops, r.lo, r.hi, flipped, r.same_strand, k) : 0ull;
        int64_t a = valid && trim_is_match(w) ? trim_len(w) : 0, b = valid && !trim_is_match(w) ? trim_len(w) : 0;
        trim_scan2(a, b, lane);
        a += cm; b += cmm;
        const unsigned hit = __ballot_sync(FULL, valid && trim_identity(a, b) >= identity);
        if (hit) keep_from = top - (31 - __clz(hit));  // the highest lane is the lowest index
        cm = __shfl_sync(FULL, a, 31); cmm = __shfl_sync(FULL, b, 31);
    }
    const int64_t count = keep_from >= 0 ? keep_from : last_low + 1;
    int64_t d_t = 0, d_q = 0;  // bases the dropped ops cover in the view's target and query
    for (int64_t k = lane; k < count; k += 32) {
        const uint64_t w = trim_view_op(ops, r.lo, r.hi, flipped, r.same_strand, k);
        const int c = (int)(w >> 56);
        if (c != OP_I) d_t += trim_len(w);
        if (c != OP_D) d_q += trim_len(w);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) { d_t += __shfl_xor_sync(FULL, d_t, o); d_q += __shfl_xor_sync(FULL, d_q, o); }
    if (!flipped) {
        r.ts += d_t;
        if (r.same_strand) r.qs += d_q; else r.qe -= d_q;
    } else {  // in the inverted view "target" is the record's query and vice versa
        r.qs += d_t;
        if (r.same_strand) r.ts += d_q; else r.te -= d_q;
    }
    if (flipped && !r.same_strand) r.hi -= count; else r.lo += count;
}

__device__ __forceinline__ void trim_sums_warp(const uint64_t *ops, int64_t lo, int64_t hi, unsigned lane, int64_t &m, int64_t &mm) {
    m = mm = 0;
    for (int64_t k = lo + lane; k < hi; k += 32) {
        const uint64_t w = __ldg(ops + k);
        if (trim_is_match(w)) m += trim_len(w); else mm += trim_len(w);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) { m += __shfl_xor_sync(FULL, m, o); mm += __shfl_xor_sync(FULL, mm, o); }
}

__global__ void __launch_bounds__(128) trim_warp_kernel(TrimParams P) {
    const unsigned lane = threadIdx.x & 31;
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (i >= P.n_records) return;
    TrimRec r;
    r.qs = P.qs[i]; r.qe = P.qe[i]; r.ts = P.ts[i]; r.te = P.te[i];
    r.lo = P.rec_out_begin[i]; r.hi = P.rec_out_begin[i + 1];
    r.same_strand = P.same_strand[i] != 0;
    bool failed = false;
    if (r.hi > r.lo) {
        int64_t m, mm;
        trim_sums_warp(P.out, r.lo, r.hi, lane, m, mm);
        const double identity = trim_identity(m, mm);
        const double threshold = __dsub_rn(identity, __dmul_rn(identity, (double)P.score_fraction));
        const int64_t max_trim = __float2ll_rz(__fmul_rn(__ll2float_rn(m + mm), P.max_fraction));
        const double thr = (double)__double2float_rn(threshold), idn = (double)__double2float_rn(identity);
        trim_prefix_warp(P.out, r, false, thr, idn, max_trim, lane);
        trim_prefix_warp(P.out, r, true, thr, idn, max_trim, lane);
        int64_t fm, fmm;
        trim_sums_warp(P.out, r.lo, r.hi, lane, fm, fmm);
        failed = !(trim_identity(fm, fmm) >= identity);
    }
    if (lane == 0) {
        P.qs[i] = r.qs; P.qe[i] = r.qe; P.ts[i] = r.ts; P.te[i] = r.te;
        P.lo[i] = r.lo; P.hi[i] = r.hi;
        P.failed[i] = failed ? 1 : 0;
    }
}

// Fixed trim, `paffy trim -f` (paf_trim_end_fraction, impl/paf.c:589-598 -> paf_trim_ends :578-587 -> cigar_trim /
// cigar_trim_back :518-576): end = (int64)((float)aligned * fraction / 2.0) aligned bases go from each end, together with
// every insert / delete met on the way; the op the cut falls in is shortened IN PLACE in `ops` (the only trim that splits
// an op). One thread per record; plain loads, because the back walk may re-read the op the front walk shortened.
__device__ __forceinline__ bool trim_is_aligned(int c) { return c == OP_M || c == OP_EQ || c == OP_X; }

__device__ __forceinline__ void trim_fixed_walk(uint64_t *ops, int64_t &lo, int64_t &hi, bool back, int64_t end_bases, int64_t &qc, int64_t &tc,
                                                int64_t q_sign, int64_t t_sign) {
    int64_t trimmed = 0;
    while (hi > lo) {
        uint64_t *p = ops + (back ? hi - 1 : lo);
        const uint64_t w = *p;
        const int c = (int)(w >> 56);
        const int64_t len = trim_len(w);
        const bool al = trim_is_aligned(c);
        if (al && trimmed >= end_bases) break;
        if (al) {
            if (trimmed + len > end_bases) {
                const int64_t i = end_bases - trimmed;
                *p = ((uint64_t)(len - i) & LEN_MASK) | ((uint64_t)c << 56);
                qc += q_sign * i;
                tc += t_sign * i;
                break;
            }
            trimmed += len;
            qc += q_sign * len;
            tc += t_sign * len;
        } else if (c == OP_I) {
            qc += q_sign * len;
        } else {
            tc += t_sign * len;
        }
        if (back) hi--; else lo++;
    }
}

__global__ void __launch_bounds__(128) trim_fixed_kernel(TrimParams P, uint64_t *ops) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.n_records) return;
    int64_t qs = P.qs[i], qe = P.qe[i], ts = P.ts[i], te = P.te[i];
    int64_t lo = P.rec_out_begin[i], hi = P.rec_out_begin[i + 1];
    int64_t aligned = 0;
    for (int64_t k = lo; k < hi; k++) {
        const uint64_t w = ops[k];
        if (trim_is_aligned((int)(w >> 56))) aligned += trim_len(w);
    }
    // int64 * float -> float, / 2.0 -> double, truncated (impl/paf.c:593)
    const int64_t end_bases = __double2ll_rz(__ddiv_rn((double)__fmul_rn(__ll2float_rn(aligned), P.max_fraction), 2.0));
    if (P.same_strand[i]) {
        trim_fixed_walk(ops, lo, hi, false, end_bases, qs, ts, 1, 1);
        trim_fixed_walk(ops, lo, hi, true, end_bases, qe, te, -1, -1);
    } else {
        trim_fixed_walk(ops, lo, hi, false, end_bases, qe, ts, -1, 1);
        trim_fixed_walk(ops, lo, hi, true, end_bases, qs, te, 1, -1);
    }
    P.qs[i] = qs; P.qe[i] = qe; P.ts[i] = ts; P.te[i] = te;
    P.lo[i] = lo; P.hi[i] = hi;
    P.failed[i] = 0;
}

}  // namespace pafgpu
