// serialise.cuh - device serialiser: =/X CIGAR words -> PAF text (a5 of SURVEY.md section 8).
// Restates the CIGAR loop of paf_write_to_buffer (impl/paf.c:367-380) and int64_to_str (:10-34):
// each op is <signed decimal length><op letter>, letters "MID=X", 'N' for any other code.
// The host renders the per-record text up to and including "\tcg:Z:"; the device lays out the lines
// (prefix + ops + '\n') with prefix sums and writes every byte.
#pragma once
#include "kernels.cuh"

namespace pafgpu {

constexpr int SER_THREADS = 256, SER_OPS_PER_BLOCK = 2048;

struct SerialiseParams {
    int64_t n_records, n_out;
    const uint64_t *out;           // output CIGAR words
    const int64_t *rec_out_begin;  // n_records + 1
    const uint64_t *op_toff;       // n_out + 1: exclusive scan of per-op text lengths
    const char *prefix;
    const int64_t *prefix_begin;   // n_records + 1
    const int64_t *line_begin;     // n_records + 1
    char *text;
};

__device__ __forceinline__ int64_t op_signed_len(uint64_t w) { return (int64_t)(w << 8) >> 8; }

__device__ __forceinline__ uint32_t dec_digits(uint64_t v) {
    uint32_t n = 1;
    if (v >= 100000000ull) { v /= 100000000ull; n += 8; }
    if (v >= 10000ull) { v /= 10000ull; n += 4; }
    if (v >= 100ull) { v /= 100ull; n += 2; }
    if (v >= 10ull) n += 1;
    return n;
}

// largest r in [lo, hi) with rec_out_begin[r] <= k
__device__ __forceinline__ int64_t record_of_op(const int64_t *__restrict__ rec_out_begin, int64_t lo, int64_t hi, int64_t k) {
    while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (__ldg(rec_out_begin + mid) <= k) lo = mid; else hi = mid;
    }
    return lo;
}

// text bytes of every op; ops outside their record's kept window [keep_lo[r], keep_hi[r]) (set by trim_kernel; null =
// keep everything) get length 0 and so vanish from the line
__global__ void __launch_bounds__(SER_THREADS) op_text_len_kernel(const uint64_t *__restrict__ out, int64_t n, uint32_t *__restrict__ tl,
                                                                  const int64_t *__restrict__ rec_out_begin, int64_t n_records,
                                                                  const int64_t *__restrict__ keep_lo, const int64_t *__restrict__ keep_hi) {
    __shared__ int64_t r_range[2];
    const int64_t k0 = (int64_t)blockIdx.x * SER_OPS_PER_BLOCK;
    const int64_t k1 = min(k0 + (int64_t)SER_OPS_PER_BLOCK, n);
    if (keep_lo && threadIdx.x < 2) r_range[threadIdx.x] = record_of_op(rec_out_begin, 0, n_records, threadIdx.x == 0 ? k0 : k1 - 1);
    __syncthreads();
    for (int64_t k = k0 + threadIdx.x; k < k1; k += SER_THREADS) {
        const int64_t len = op_signed_len(out[k]);
        const uint64_t mag = len < 0 ? 0 - (uint64_t)len : (uint64_t)len;
        uint32_t d = dec_digits(mag);
        if (mag >= 10000000000000000ull) d = 17;  // dec_digits folds at most 8+4+2+1 extra digits
        uint32_t bytes = d + (len < 0) + 1;
        if (keep_lo) {
            const int64_t r = record_of_op(rec_out_begin, r_range[0], r_range[1] + 1, k);
            if (k < __ldg(keep_lo + r) || k >= __ldg(keep_hi + r)) bytes = 0;
        }
        tl[k] = bytes;
    }
}

// line_begin[r] = bytes of all earlier prefixes + earlier CIGAR text + one newline per earlier record
__global__ void line_begin_kernel(int64_t n_records, const int64_t *prefix_begin, const int64_t *rec_out_begin,
                                  const uint64_t *op_toff, int64_t *line_begin) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r > n_records) return;
    line_begin[r] = prefix_begin[r] + (int64_t)op_toff[rec_out_begin[r]] + r;
}

// one warp per record: copy the host-rendered prefix and terminate the line
__global__ void prefix_copy_kernel(SerialiseParams S) {
    const unsigned lane = threadIdx.x & 31;
    const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (r >= S.n_records) return;
    const int64_t p0 = S.prefix_begin[r], p1 = S.prefix_begin[r + 1], dst = S.line_begin[r];
    for (int64_t i = lane; i < p1 - p0; i += 32) S.text[dst + i] = S.prefix[p0 + i];
    if (lane == 0) S.text[S.line_begin[r + 1] - 1] = '\n';
}

// one thread per output op; the block narrows the record search to the records it overlaps
__global__ void __launch_bounds__(SER_THREADS) op_text_kernel(SerialiseParams S) {
    __shared__ int64_t r_range[2];
    const int64_t k0 = (int64_t)blockIdx.x * SER_OPS_PER_BLOCK;
    const int64_t k1 = min(k0 + (int64_t)SER_OPS_PER_BLOCK, S.n_out);
    if (threadIdx.x < 2) r_range[threadIdx.x] = record_of_op(S.rec_out_begin, 0, S.n_records, threadIdx.x == 0 ? k0 : k1 - 1);
    __syncthreads();
    for (int64_t k = k0 + threadIdx.x; k < k1; k += SER_THREADS) {
        const uint64_t t0 = __ldg(S.op_toff + k), t1 = __ldg(S.op_toff + k + 1);
        if (t1 == t0) continue;  // trimmed away
        const int64_t r = record_of_op(S.rec_out_begin, r_range[0], r_range[1] + 1, k);
        const uint64_t w = __ldg(S.out + k);
        const int64_t len = op_signed_len(w);
        uint64_t mag = len < 0 ? 0 - (uint64_t)len : (uint64_t)len;
        char *p = S.text + (int64_t)t0 + __ldg(S.prefix_begin + r + 1) + r;
        char *e = S.text + (int64_t)t1 + __ldg(S.prefix_begin + r + 1) + r;
        const unsigned code = (unsigned)(w >> 56);
        *--e = code <= 4 ? "MID=X"[code] : 'N';
        do { *--e = (char)('0' + mag % 10); mag /= 10; } while (mag);
        if (len < 0) *p = '-';
    }
}

}  // namespace pafgpu
