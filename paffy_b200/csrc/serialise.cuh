// serialise.cuh - device serialiser: =/X CIGAR words -> PAF text (a5 of SURVEY.md section 8).
// Restates the CIGAR loop of paf_write_to_buffer (impl/paf.c:367-380) and int64_to_str (:10-34):
// each op is <signed decimal length><op letter>, letters "MID=X", 'N' for any other code.
// The host renders the per-record text up to and including "\tcg:Z:"; the device lays out the lines
// (prefix + ops + '\n') with prefix sums and writes every byte.
#pragma once
#include "kernels.cuh"

namespace pafgpu {

constexpr int SER_THREADS = 256, SER_PER_THREAD = 8, SER_OPS_PER_BLOCK = SER_THREADS * SER_PER_THREAD;
#ifndef SER_MINB
#define SER_MINB 5
#endif
constexpr int SER_ROB = SER_OPS_PER_BLOCK + 8;  // record offsets a block keeps in shared memory
constexpr int SER_STAGE = 16384;  // bytes of CIGAR text a block stages in shared memory (more than that: written straight out)

struct SerialiseParams {
    int64_t n_records, n_out;
    const uint64_t *out;              // output CIGAR words
    const int64_t *rec_out_begin;     // n_records + 1
    const uint32_t *block_rec;        // per block of SER_OPS_PER_BLOCK ops: the record that owns its first op
    unsigned long long *status;       // one look-back word per block, zero on entry
    const char *prefix;
    const int64_t *prefix_begin;      // n_records + 1
    int64_t *line_begin;              // n_records + 1 (written by ser_write_kernel; [n_records] = bytes of the whole text)
    char *text;
    int64_t text_cap;                 // bytes `text` holds: nothing is written beyond (the host then grows it and runs again)
    const int64_t *keep_lo, *keep_hi; // per record, set by the trim kernels; null = every op is kept
};

__device__ __forceinline__ int64_t op_signed_len(uint64_t w) { return (int64_t)(w << 8) >> 8; }

// bytes of "<signed decimal length><op letter>" (impl/paf.c:367-380), any length
__device__ __noinline__ uint32_t op_text_bytes_any(uint64_t w) {
    const int64_t len = op_signed_len(w);
    uint64_t v = len < 0 ? 0 - (uint64_t)len : (uint64_t)len;
    uint32_t n = 2 + (len < 0);
    while (v >= 10) { v /= 10; n++; }
    return n;
}
__device__ __forceinline__ char op_letter(uint64_t w) {  // "MID=X", 'N' for any other code
    const unsigned code = (unsigned)(w >> 56);
    return code < 4 ? (char)((0x3d44494du >> (8 * code)) & 0xffu) : code == 4 ? 'X' : 'N';
}
// the text of any op, in the `bytes` bytes at p
__device__ __noinline__ void op_text_put_any(char *p, uint32_t bytes, uint64_t w) {
    char *e = p + bytes;
    const int64_t len = op_signed_len(w);
    uint64_t m = len < 0 ? 0 - (uint64_t)len : (uint64_t)len;
    *--e = op_letter(w);
    do { *--e = (char)('0' + m % 10); m /= 10; } while (m);
    if (len < 0) *p = '-';
}
// nearly every op is shorter than 10000 bases: 1-4 digits and the letter, measured and written without a loop
__device__ __forceinline__ bool op_is_short(uint64_t w) { return ((uint32_t)(w >> 32) & 0xffffffu) == 0 && (uint32_t)w < 10000u; }
__device__ __forceinline__ uint32_t op_text_bytes(uint64_t w) {
    const uint32_t v = (uint32_t)w;
    return op_is_short(w) ? 2u + (v >= 10u) + (v >= 100u) + (v >= 1000u) : op_text_bytes_any(w);
}
// the digits of v < 10000 as characters, most significant first, in the low `nd` bytes
__device__ __forceinline__ uint32_t short_digits(uint32_t v, uint32_t nd) {
    const uint32_t q1 = v / 10u, q2 = v / 100u, q3 = v / 1000u;
    const uint32_t packed = (q3 | ((q2 - 10u * q3) << 8) | ((q1 - 10u * q2) << 16) | ((v - 10u * q1) << 24)) + 0x30303030u;
    return packed >> (8u * (4u - nd));
}

// largest r in [lo, hi) with rec_out_begin[r] <= k
__device__ __forceinline__ int64_t record_of_op(const int64_t *__restrict__ rec_out_begin, int64_t lo, int64_t hi, int64_t k) {
    while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (__ldg(rec_out_begin + mid) <= k) lo = mid; else hi = mid;
    }
    return lo;
}

// block_rec[b] = the record that owns op b * SER_OPS_PER_BLOCK: one thread per record fills in the blocks that begin
// inside its ops (a block's threads then search a handful of records instead of all of them)
__global__ void ser_block_records_kernel(int64_t n_records, const int64_t *__restrict__ rec_out_begin, uint32_t *__restrict__ block_rec) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_records) return;
    const int64_t a = rec_out_begin[r], b = rec_out_begin[r + 1];
    for (int64_t blk = (a + SER_OPS_PER_BLOCK - 1) / SER_OPS_PER_BLOCK; blk * SER_OPS_PER_BLOCK < b; blk++) block_rec[blk] = (uint32_t)r;
}

// no output ops at all: every line is its prefix and a newline
__global__ void line_begin_no_ops_kernel(int64_t n_records, const int64_t *prefix_begin, int64_t *line_begin) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r <= n_records) line_begin[r] = prefix_begin[r] + r;
}

// one warp per record: copy the host-rendered prefix and terminate the line
__global__ void prefix_copy_kernel(SerialiseParams S) {
    const unsigned lane = threadIdx.x & 31;
    const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (r >= S.n_records) return;
    const int64_t p0 = S.prefix_begin[r], p1 = S.prefix_begin[r + 1], dst = S.line_begin[r], end = S.line_begin[r + 1];
    if (end > S.text_cap) return;
    for (int64_t i = lane; i < p1 - p0; i += 32) S.text[dst + i] = S.prefix[p0 + i];
    if (lane == 0) S.text[end - 1] = '\n';
}


// The warp copies n bytes from shared memory to global memory, both at any alignment: the destination's whole 4-byte words
// are stored as words (each put together from the two shared words it straddles), the ragged ends as bytes. The shared
// buffer must be readable up to the 4-byte word behind src + n.
__device__ __forceinline__ void warp_copy_out(char *dst, const uint8_t *src, uint32_t n, unsigned lane) {
    uint32_t head = (4u - (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 3u)) & 3u;
    if (head > n) head = n;
    if (lane < head) dst[lane] = (char)src[lane];
    const uint32_t body = (n - head) >> 2;
    const uint8_t *s = src + head;
    const uint32_t mis = (uint32_t)(__cvta_generic_to_shared(s) & 3u);
    const uint32_t *sw = reinterpret_cast<const uint32_t *>(s - mis);
    uint32_t *d = reinterpret_cast<uint32_t *>(dst + head);
    for (uint32_t j = lane; j < body; j += 32) d[j] = __funnelshift_r(sw[j], sw[j + 1], 8u * mis);
    const uint32_t done = head + 4u * body;
    if (lane < n - done) dst[done + lane] = (char)src[done + lane];
}


// ONE pass over the output ops: block b takes ops [b, b + 1) * SER_OPS_PER_BLOCK, thread t loads ops
// [8 t, 8 t + 8), measures their text, the block scans the lengths and publishes their sum; the CIGAR bytes of all
// earlier blocks come from decoupled look-back (kernels.cuh), asked for only after the block has rendered its ops into
// shared memory, when the blocks before it have long published theirs. An op of record r lands at
//   text + (CIGAR bytes before it) + prefix_begin[r + 1] + r
// (the prefixes of records 0..r and one newline per earlier record come before it), so the staged text leaves in one
// piece per record, copied by the warps with aligned word stores (warp_copy_out). A block whose text does not fit the
// staging buffer writes straight to global memory. line_begin of a record is written where its first op is handled -
// together with the empty records right before it; the last block does the same for what sits at the op total, so
// line_begin[n_records] = bytes of the whole text.
__global__ void __launch_bounds__(SER_THREADS, SER_MINB) ser_write_kernel(SerialiseParams S) {
    __shared__ __align__(16) uint8_t s_text[SER_STAGE + 16];
    __shared__ uint16_t s_seg_off[SER_OPS_PER_BLOCK + 2], s_seg_r[SER_OPS_PER_BLOCK + 2];
    __shared__ int32_t s_rob[SER_ROB];
    __shared__ uint32_t s_warp[SER_THREADS / 32];
    __shared__ uint64_t s_tb;
    const unsigned tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    // tiles in block-index order: blocks are dispatched in that order, so every earlier tile belongs to a block that is
    // running or done (the assumption every single-pass scan keyed by blockIdx makes)
    const int64_t tile = blockIdx.x;
    const int64_t k0 = tile * SER_OPS_PER_BLOCK;
    const int64_t k1 = min(k0 + (int64_t)SER_OPS_PER_BLOCK, S.n_out);
    const int64_t *__restrict__ rob = S.rec_out_begin;
    const bool keep = S.keep_lo != nullptr;

    const int32_t jf = (int32_t)(tid * SER_PER_THREAD), j1 = (int32_t)(k1 - k0);  // block-relative op indices
    uint64_t w[SER_PER_THREAD];
    if (jf + SER_PER_THREAD <= j1) {
        const uint4 *src = reinterpret_cast<const uint4 *>(S.out + k0 + jf);
#pragma unroll
        for (int q = 0; q < SER_PER_THREAD / 2; q++) {
            const uint4 v = __ldg(src + q);
            w[2 * q] = (uint64_t)v.x | ((uint64_t)v.y << 32);
            w[2 * q + 1] = (uint64_t)v.z | ((uint64_t)v.w << 32);
        }
    } else {
#pragma unroll
        for (int j = 0; j < SER_PER_THREAD; j++) w[j] = jf + j < j1 ? __ldg(S.out + k0 + jf + j) : 0ull;
    }
    // the records the block's ops belong to: [r_lo, r_hi]; their op offsets relative to the block go to shared memory
    // (clamped to [-1, 2 * SER_OPS_PER_BLOCK]: only offsets inside the block are ever compared for equality)
    const int64_t r_lo = __ldg(S.block_rec + tile);
    const int64_t r_hi = k1 < S.n_out ? (int64_t)__ldg(S.block_rec + tile + 1) : S.n_records - 1;
    const bool rob_shared = r_hi - r_lo + 2 <= SER_ROB;
    auto rel = [&](int64_t v) -> int32_t { return (int32_t)max((int64_t)-1, min(v - k0, (int64_t)2 * SER_OPS_PER_BLOCK)); };
    if (rob_shared)
        for (int64_t i = tid; i < r_hi - r_lo + 2; i += SER_THREADS) s_rob[i] = rel(__ldg(rob + r_lo + i));
    __syncthreads();
    auto rob_at = [&](int32_t i) -> int32_t { return rob_shared ? s_rob[i] : rel(__ldg(rob + r_lo + i)); };  // op offset of record r_lo + i

    // first walk: the record of every op, the text bytes of every op, which ops start a record
    int32_t i_first = 0, cur0 = 0, next0 = 0;  // (records as indices relative to r_lo)
    if (jf < j1) {
        int32_t lo = 0, hi = (int32_t)min(r_hi - r_lo + 1, (int64_t)0x7fffffff);  // largest i in [lo, hi) whose first op is not behind jf
        while (hi - lo > 1) {
            const int32_t mid = (lo + hi) >> 1;
            if (rob_at(mid) <= jf) lo = mid; else hi = mid;
        }
        i_first = lo;
        cur0 = rob_at(i_first);
        next0 = rob_at(i_first + 1);
    }
    uint32_t lens = 0, lens_hi = 0;  // text bytes of the 8 ops, one byte each
    uint32_t starts = 0, mine = 0;
    {
        int32_t i = i_first, cur = cur0, next = next0;
        int64_t klo = 0, khi = 0;
        if (keep && jf < j1) { klo = __ldg(S.keep_lo + r_lo + i); khi = __ldg(S.keep_hi + r_lo + i); }
#pragma unroll
        for (int j = 0; j < SER_PER_THREAD; j++) {
            const int32_t k = jf + j;
            if (k < j1) {
                if (k >= next) {
                    do { i++; cur = next; next = rob_at(i + 1); } while (k >= next);
                    if (keep) { klo = __ldg(S.keep_lo + r_lo + i); khi = __ldg(S.keep_hi + r_lo + i); }
                }
                if (k == cur) starts |= 1u << j;
                uint32_t b = op_text_bytes(w[j]);
                if (keep && (k0 + k < klo || k0 + k >= khi)) b = 0;  // trimmed away
                if (j < 4) lens |= b << (8 * j); else lens_hi |= b << (8 * (j - 4));
                mine += b;
            }
        }
    }
    const uint32_t true_starts = starts;
    if (tid == 0) starts |= 1u;  // the block's first op opens its first piece, whoever's record it is
    // block scan of (text bytes, pieces): both fit 16 bits (2048 ops x 19 bytes at most)
    const uint32_t packed = mine | ((uint32_t)__popc(starts) << 16);
    uint32_t inc = packed;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL, inc, o);
        if (lane >= (unsigned)o) inc += t;
    }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    uint32_t base = inc - packed, tot = 0;
#pragma unroll
    for (unsigned q = 0; q < SER_THREADS / 32; q++) {
        const uint32_t v = s_warp[q];
        if (q < wid) base += v;
        tot += v;
    }
    const uint32_t tot_bytes = tot & 0xffffu, n_seg = tot >> 16;
    const bool staged = tot_bytes <= (uint32_t)SER_STAGE && r_hi - r_lo < 65536;
    // the block's sum goes out and the sum of the earlier blocks is fetched by one warp while the others render; only a
    // block that writes straight to global memory has to wait for it first
    if (wid == SER_THREADS / 32 - 1) {
        const uint64_t excl = lookback_exclusive(S.status, tile, tot_bytes, lane);
        if (lane == 0) s_tb = excl;
    }
    if (!staged) __syncthreads();

    // second walk: render the ops (into the staging buffer, or straight out) and note where the pieces begin
    {
        uint32_t off = base & 0xffffu, sidx = base >> 16;
        int32_t i = i_first, next = next0;
#pragma unroll
        for (int j = 0; j < SER_PER_THREAD; j++) {
            const int32_t k = jf + j;
            if (k < j1) {
                while (k >= next) { i++; next = rob_at(i + 1); }
                const uint32_t b = ((j < 4 ? lens : lens_hi) >> (8 * (j & 3))) & 0xffu;
                if (staged) {
                    if ((starts >> j) & 1u) { s_seg_off[sidx] = (uint16_t)off; s_seg_r[sidx] = (uint16_t)i; sidx++; }
                    if (b) {
                        uint8_t *p = s_text + off;
                        if (op_is_short(w[j])) {
                            const uint32_t nd = b - 1u, t = short_digits((uint32_t)w[j], nd);
                            p[0] = (uint8_t)t;
                            if (nd > 1) p[1] = (uint8_t)(t >> 8);
                            if (nd > 2) p[2] = (uint8_t)(t >> 16);
                            if (nd > 3) p[3] = (uint8_t)(t >> 24);
                            p[nd] = (uint8_t)op_letter(w[j]);
                        } else {
                            op_text_put_any(reinterpret_cast<char *>(p), b, w[j]);
                        }
                    }
                } else {
                    const int64_t r = r_lo + i, tb = (int64_t)s_tb;
                    if ((true_starts >> j) & 1u)
                        for (int64_t q = r; q >= 0 && __ldg(rob + q) == k0 + k; q--) S.line_begin[q] = __ldg(S.prefix_begin + q) + tb + off + q;
                    const int64_t pos = tb + off + __ldg(S.prefix_begin + r + 1) + r;
                    if (b && pos + b <= S.text_cap) op_text_put_any(S.text + pos, b, w[j]);
                }
                off += b;
            }
        }
    }
    if (staged) {
        if (tid == 0) s_seg_off[n_seg] = (uint16_t)tot_bytes;
        __syncthreads();
    }
    const int64_t tb = (int64_t)s_tb;
    if (k1 == S.n_out && tid == 0)  // what sits at the op total: trailing empty records and the end of the text
        for (int64_t q = S.n_records; q >= 0 && __ldg(rob + q) == S.n_out; q--) S.line_begin[q] = __ldg(S.prefix_begin + q) + tb + tot_bytes + q;
    if (!staged) return;
    // line_begin of the records that start in this block (piece 0 only if the block's first op starts its record)
    for (uint32_t seg = tid; seg < n_seg; seg += SER_THREADS) {
        const int64_t r = r_lo + s_seg_r[seg];
        const int64_t kk = __ldg(rob + r);
        if (seg == 0 && kk != k0) continue;
        for (int64_t q = r; q >= 0 && __ldg(rob + q) == kk; q--) S.line_begin[q] = __ldg(S.prefix_begin + q) + tb + s_seg_off[seg] + q;
    }
    // copy out, piece by piece: many pieces are dealt to the warps whole, few (long records) in 512-byte slices
    if (n_seg >= 2 * SER_THREADS / 32) {
        for (uint32_t seg = wid; seg < n_seg; seg += SER_THREADS / 32) {
            const uint32_t o0 = s_seg_off[seg], o1 = s_seg_off[seg + 1];
            if (o1 == o0) continue;
            const int64_t r = r_lo + s_seg_r[seg];
            const int64_t shift = tb + __ldg(S.prefix_begin + r + 1) + r;
            if (shift + o1 <= S.text_cap) warp_copy_out(S.text + shift + o0, s_text + o0, o1 - o0, lane);
        }
    } else {
        uint32_t slice = 0;
        for (uint32_t seg = 0; seg < n_seg; seg++) {
            const uint32_t o0 = s_seg_off[seg], o1 = s_seg_off[seg + 1];
            const uint32_t mine_from = (wid - slice) & (SER_THREADS / 32 - 1);  // first slice of this piece that is this warp's
            slice += (o1 - o0 + 511u) >> 9;
            if (o0 + 512u * mine_from >= o1) continue;
            const int64_t r = r_lo + s_seg_r[seg];
            const int64_t shift = tb + __ldg(S.prefix_begin + r + 1) + r;
            for (uint32_t a = o0 + 512u * mine_from; a < o1; a += 512u * (SER_THREADS / 32)) {
                const uint32_t n = min(512u, o1 - a);
                if (shift + a + n <= S.text_cap) warp_copy_out(S.text + shift + a, s_text + a, n, lane);
            }
        }
    }
}

}  // namespace pafgpu
