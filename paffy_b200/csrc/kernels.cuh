// kernels.cuh - sm_100a kernels of the add_mismatches path (genome packing, CIGAR ingest, the
// base-comparison scan, run stitching, device serialiser). Included by pafgpu.cu only.
//
// Reference behaviour restated (never copied): paf_encode_mismatches impl/paf.c:739-784,
// stString_reverseComplementChar sonLibString.c:184-242, paf_write_to_buffer impl/paf.c:367-380.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pafgpu {

constexpr uint64_t LEN_MASK = (1ull << 56) - 1;
constexpr uint64_t POS_MASK = (1ull << 48) - 1;
constexpr uint64_t QFLAG_EXC = 1ull << 62;   // qpos: the op's target or query span touches a non-ACGT word
constexpr uint64_t TFLAG_HEAD = 1ull << 63;  // tpos: first op of a record
constexpr uint64_t EVEN_BITS = 0x5555555555555555ull;
constexpr int OP_M = 0, OP_I = 1, OP_D = 2, OP_EQ = 3, OP_X = 4;
constexpr unsigned FULL = 0xffffffffu;
constexpr uint32_t INF32 = 0xffffffffu;
constexpr uint64_t INF64 = ~0ull;
constexpr int ARENA_ALIGN = 1024;  // bases; contig offsets and pads are multiples of this
#ifndef ENC_WIN_WORDS
#define ENC_WIN_WORDS 4
#endif
constexpr int WINDOW_BASES = 32 * ENC_WIN_WORDS;  // bases per work window of the encode kernel (encode.cuh)

// ------------------------------------------------------------------------------------------------
// Genome arena: 2 bits/base (A0 C1 G2 T3), bit-sliced per 32 bases: one 64-bit word holds the low
// code bits of 32 consecutive bases in bits 0..31 (base k at bit k) and their high code bits in bits
// 32..63. A 32-base window is then two 32-bit planes, so alignment is one funnel shift per plane,
// reverse is one BREV per plane, complement is NOT and the mismatch mask has one bit per base.
// Every other byte lives in an exception plane: one flag bit per word, a rank directory (exclusive
// count of flagged words per 32-word group) and 32 upper-cased bytes per flagged word.
// ------------------------------------------------------------------------------------------------
struct Genome {
    const uint64_t *words;
    const uint32_t *flags;  // bit (w & 31) of flags[w >> 5]
    const uint32_t *rank;   // flagged words before group w >> 5
    const uint8_t *exc;     // 32 bytes per flagged word
};

__device__ __forceinline__ uint64_t ldg64(const uint64_t *p) { return __ldg(p); }

__device__ __forceinline__ uint32_t flagged_before(const Genome &G, uint64_t w) {
    const uint32_t f = __ldg(G.flags + (w >> 5));
    return __ldg(G.rank + (w >> 5)) + __popc(f & ((1u << (w & 31)) - 1u));
}
__device__ __forceinline__ bool word_flagged(const Genome &G, uint64_t w) {
    return (__ldg(G.flags + (w >> 5)) >> (w & 31)) & 1u;
}
// true when any word overlapping bases [lo, hi] holds a non-ACGT byte
__device__ __forceinline__ bool span_has_exception(const Genome &G, uint64_t lo, uint64_t hi) {
    return flagged_before(G, (hi >> 5) + 1) != flagged_before(G, lo >> 5);
}
// The upper-cased byte the reference would see at arena position p.
__device__ __forceinline__ uint8_t byte_at(const Genome &G, uint64_t p) {
    const uint64_t w = p >> 5;
    if (word_flagged(G, w)) return __ldg(G.exc + (uint64_t)flagged_before(G, w) * 32 + (p & 31));
    const uint64_t word = ldg64(G.words + w);
    const unsigned k = (unsigned)(p & 31);
    const unsigned code = ((unsigned)(word >> k) & 1u) | (((unsigned)(word >> (32 + k)) & 1u) << 1);
    return (uint8_t)((0x54474341u >> (code * 8)) & 0xffu);  // "ACGT"
}
// Complement of an upper-cased byte (sonLibString.c:188-242): six swapped pairs, all else fixed.
__device__ __forceinline__ uint8_t complement_upper(uint8_t c) {
    switch (c) {
        case 'A': return 'T'; case 'T': return 'A'; case 'C': return 'G'; case 'G': return 'C';
        case 'M': return 'K'; case 'K': return 'M'; case 'R': return 'Y'; case 'Y': return 'R';
        case 'B': return 'V'; case 'V': return 'B'; case 'D': return 'H'; case 'H': return 'D';
        default: return c;
    }
}
__device__ __forceinline__ uint8_t upper(uint8_t c) { return (c >= 'a' && c <= 'z') ? c - 32 : c; }

// ---- pack: ASCII -> 2-bit words + flag bits. One thread per word, one warp per flag word. ------------
// rc = true packs the reverse complement of the contig (base i = complement of source base len-1-i), so that '-'
// strand alignments read the query ascending like '+' ones: no reversal or complement is left in the hot kernel.
__device__ __forceinline__ uint8_t source_byte(const uint8_t *__restrict__ src, int64_t len, int64_t i, bool rc) {
    if (i >= len) return (uint8_t)'A';
    const uint8_t u = upper(__ldg(src + (rc ? len - 1 - i : i)));
    return rc ? complement_upper(u) : u;
}
__global__ void pack_kernel(const uint8_t *__restrict__ src, int64_t len, uint64_t word0, uint64_t *words,
                            uint32_t *flags, int64_t n_words /* multiple of 32 */, bool rc) {
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    const int64_t b0 = w * 32;
    uint64_t packed = 0;
    bool exc = false;
    uint8_t c[32];
    if (!rc && b0 + 32 <= len && ((uintptr_t)(src + b0) & 15) == 0) {
        const uint4 a = __ldg(reinterpret_cast<const uint4 *>(src + b0));
        const uint4 b = __ldg(reinterpret_cast<const uint4 *>(src + b0) + 1);
        const uint32_t v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < 32; k++) c[k] = upper((uint8_t)(v[k >> 2] >> ((k & 3) * 8)));
    } else {
#pragma unroll
        for (int k = 0; k < 32; k++) c[k] = source_byte(src, len, b0 + k, rc);
    }
#pragma unroll
    for (int k = 0; k < 32; k++) {
        const uint8_t u = c[k];
        const bool acgt = (u == 'A') | (u == 'C') | (u == 'G') | (u == 'T');
        exc |= !acgt;
        const uint64_t code = acgt ? (uint64_t)(((u >> 1) ^ (u >> 2)) & 3u) : 0ull;
        packed |= ((code & 1) << k) | ((code >> 1) << (32 + k));
    }
    words[word0 + w] = packed;
    const unsigned ballot = __ballot_sync(FULL, exc);
    if ((threadIdx.x & 31) == 0) flags[(word0 + w) >> 5] = ballot;
}

// popcount of each flag word (input of the rank scan)
__global__ void flag_popc_kernel(const uint32_t *flags, int64_t n, uint32_t *out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __popc(flags[i]);
}
// rank[i] = base + scanned[i]
__global__ void rank_fill_kernel(const uint64_t *scanned, int64_t n, uint32_t base, uint32_t *rank) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) rank[i] = base + (uint32_t)scanned[i];
}
// store the 32 upper-cased bytes of every flagged word
__global__ void exc_fill_kernel(const uint8_t *__restrict__ src, int64_t len, uint64_t word0, const uint32_t *flags,
                                const uint32_t *rank, uint8_t *exc, int64_t n_words, bool rc) {
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    const uint64_t gw = word0 + w;
    const uint32_t f = flags[gw >> 5];
    if (!((f >> (gw & 31)) & 1u)) return;
    const uint64_t r = rank[gw >> 5] + __popc(f & ((1u << (gw & 31)) - 1u));
    for (int k = 0; k < 32; k++) {
        const int64_t b = w * 32 + k;
        exc[r * 32 + k] = source_byte(src, len, b, rc);
    }
}
// read back bases as the bytes the comparison sees (tests / pafgpu_genome_fetch)
__global__ void unpack_kernel(Genome G, uint64_t p0, int64_t len, uint8_t *out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < len) out[i] = byte_at(G, p0 + i);
}

// ------------------------------------------------------------------------------------------------
// Exclusive scan u32 -> u64, out has n + 1 entries.
// ------------------------------------------------------------------------------------------------
#ifndef SCAN_APPLY_MINB
#define SCAN_APPLY_MINB 4
#endif
constexpr int SCAN_THREADS = 256, SCAN_ITEMS = 16, SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ uint64_t warp_incl_scan(uint64_t v, unsigned lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint64_t t = __shfl_up_sync(FULL, v, o);
        if (lane >= (unsigned)o) v += t;
    }
    return v;
}
// inclusive block scan for SCAN_THREADS threads; returns inclusive value, *total = block sum
__device__ __forceinline__ uint64_t block_incl_scan(uint64_t v, uint64_t *total) {
    __shared__ uint64_t wsum[SCAN_THREADS / 32];
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint64_t inc = warp_incl_scan(v, lane);
    __syncthreads();
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    uint64_t off = 0, tot = 0;
#pragma unroll
    for (int k = 0; k < SCAN_THREADS / 32; k++) {
        const uint64_t s = wsum[k];
        if (k < (int)wid) off += s;
        tot += s;
    }
    *total = tot;
    return inc + off;
}

// Decoupled look-back, run by ONE warp of the block that owns `tile`: publishes the tile's sum `tot` and returns the sum of
// all earlier tiles, read from their status words (2 flag bits + 62 value bits: aggregate / inclusive prefix). The tile
// order must come from a ticket counter, so that every earlier tile belongs to a block that is running or done.
__device__ __forceinline__ uint64_t lookback_exclusive(unsigned long long *status, int64_t tile, uint64_t tot, unsigned lane) {
    volatile unsigned long long *st = status;
    if (lane == 0) atomicMax(status + tile, (tile == 0 ? (2ull << 62) : (1ull << 62)) | tot);
    uint64_t excl = 0;
    if (tile > 0) {
        int64_t j = tile - 1 - (int64_t)lane;
        for (;;) {
            unsigned long long w = j >= 0 ? st[j] : (2ull << 62);
            unsigned has_incl;
            for (;;) {
                has_incl = __ballot_sync(FULL, (w >> 62) == 2);
                const unsigned need = has_incl ? has_incl ^ (has_incl - 1u) : FULL;  // the tiles back to the nearest inclusive prefix
                if (!(__ballot_sync(FULL, (w >> 62) == 0) & need)) break;
                __nanosleep(32);
                w = j >= 0 ? st[j] : (2ull << 62);
            }
            const unsigned upto = has_incl ? (unsigned)__ffs(has_incl) - 1u : 31u;
            unsigned long long c = lane <= upto ? (w & ((1ull << 62) - 1)) : 0ull;
#pragma unroll
            for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(FULL, c, o);
            excl += c;
            if (has_incl) break;
            j -= 32;
        }
        if (lane == 0) atomicMax(status + tile, (2ull << 62) | (excl + tot));
    }
    return excl;
}

// ONE pass with decoupled look-back: a block takes the next tile from a ticket counter (so the tiles before it belong to
// blocks that are running or done), scans its 4096 items, publishes its sum, and obtains the sum of all earlier tiles
// from their status words (2 flag bits + 62 value bits: aggregate / inclusive prefix). One read of the input and one
// write of the output (round 1 ran three kernels: block sums, a single-block spine, apply). `status` (n_tiles words) and
// `ticket` must be zero on entry. Thread t owns SCAN_ITEMS consecutive items; a warp's 32 x SCAN_ITEMS items travel between
// global memory (coalesced: item = 32 * k + lane) and their owners through shared memory (owner o keeps its items at
// o * (SCAN_ITEMS + 1) ..: the odd stride keeps both views off each other's banks).
// tile_first_op (optional): in[] are windows per op; every warp tile of tile_windows (a power of two) windows learns the op that owns its
// first window (op j owns windows [out[j], out[j + 1])).
constexpr unsigned long long SCAN_AGG = 1ull << 62, SCAN_INCL = 2ull << 62, SCAN_VAL = (1ull << 62) - 1;
__global__ void __launch_bounds__(SCAN_THREADS, SCAN_APPLY_MINB) scan_onepass_kernel(const uint32_t *__restrict__ in, int64_t n,
                                                                   uint64_t *__restrict__ out, unsigned long long *status, unsigned int *ticket,
                                                                   uint64_t *out_total, uint32_t *__restrict__ tile_first_op,
                                                                   uint32_t tile_windows, uint64_t tile_cap) {
    static_assert(SCAN_ITEMS == 16, "index arithmetic below assumes 16 items per thread");
    __shared__ uint64_t stage[SCAN_THREADS / 32][32 * (SCAN_ITEMS + 1)];
    __shared__ unsigned s_tile;
    __shared__ uint64_t s_excl;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    const unsigned tile_shift = 31u - (unsigned)__clz((int)tile_windows);  // tile_windows is a power of two: shifts, not 64-bit divisions
    __syncthreads();
    const int64_t tile = s_tile;
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint64_t *st64 = stage[wid];
    uint32_t *st32 = reinterpret_cast<uint32_t *>(st64);
    const int64_t wbase = tile * SCAN_TILE + (int64_t)wid * 32 * SCAN_ITEMS;
    const int64_t base = wbase + (int64_t)lane * SCAN_ITEMS;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const unsigned p = 32u * k + lane;
        const int64_t i = wbase + p;
        st32[(p >> 4) * (SCAN_ITEMS + 1) + (p & 15u)] = i < n ? in[i] : 0u;
    }
    __syncwarp();
    uint32_t v[SCAN_ITEMS];
    uint64_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        v[k] = st32[lane * (SCAN_ITEMS + 1) + k];
        s += v[k];
    }
    uint64_t tot;
    const uint64_t incl = block_incl_scan(s, &tot);  // (its barriers also fence the staging buffer)
    if (wid == 0) {
        const uint64_t excl = lookback_exclusive(status, tile, tot, lane);
        if (lane == 0) s_excl = excl;
    }
    __syncthreads();
    uint64_t run = s_excl + incl - s;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        st64[lane * (SCAN_ITEMS + 1) + k] = run;
        if (base + k < n && tile_first_op && v[k])
            for (uint64_t t = (run + tile_windows - 1) >> tile_shift; (t << tile_shift) < run + v[k] && t < tile_cap; t++)
                tile_first_op[t] = (uint32_t)(base + k);
        run += v[k];
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const unsigned p = 32u * k + lane;
        const int64_t i = wbase + p;
        if (i < n) out[i] = st64[(p >> 4) * (SCAN_ITEMS + 1) + (p & 15u)];
    }
    if (base <= n - 1 && n - 1 < base + SCAN_ITEMS) {
        out[n] = s_excl + tot;
        if (out_total) *out_total = s_excl + tot;
    }
}

// ------------------------------------------------------------------------------------------------
// CIGAR ingest (a4 -> SoA): one warp per record walks its ops in strides of 32, turning the
// reference's running cursors (qi, tj of impl/paf.c:746-775) into warp prefix sums. Per op it
// writes the absolute arena position of its first target and query base, the owning record and the
// number of work windows of WINDOW_BASES bases (1 for a pass-through op).
// ------------------------------------------------------------------------------------------------
struct IngestParams {
    int64_t n_records, n_ops;
    const int64_t *op_begin;
    const uint64_t *ops;
    const int32_t *q_contig, *t_contig;
    const int64_t *q_start, *q_end, *t_start;
    const uint8_t *same_strand;
    const int64_t *contig_off, *contig_off_rc, *contig_len;  // forward copy, reverse-complement copy, length
    int64_t n_contigs;
    Genome G;
    uint64_t *tpos, *qpos;
    uint32_t *op_rec, *nwin;
    unsigned long long *bad_record;  // min index of a record that fails validation (init ~0)
    unsigned long long *aligned;     // sum of M lengths
    int has_exceptions;              // 0: the genome holds only ACGT, no op can touch the exception plane
    // text mode (batches parsed on the device, paftext.cuh): a record without ops is legal and is left alone; a record with
    // a cigar must also pass paf_check's op sums (impl/paf.c:440-459): sum of non-D lengths = query span, non-I = target span
    const int64_t *t_end;            // null outside text mode
    const uint8_t *has_cg;
};

// G lanes per record (8, 16 or 32, chosen by the host from the mean ops per record) so that short CIGARs do not leave
// most of a warp idle; every lane stays in the loop until the longest record of the warp is done (the scans shuffle).
template <int G>
__global__ void __launch_bounds__(256) ingest_kernel(IngestParams P) {
    __shared__ unsigned long long block_aligned;
    if (threadIdx.x == 0) block_aligned = 0;
    __syncthreads();
    const unsigned lane = threadIdx.x & 31, gl = lane & (G - 1);
    const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
    const bool live = r < P.n_records;
    int64_t j0 = 0, j1 = 0, qc = -1, tc = -1, qs = 0, qe = 0, ts = 0;
    bool rev = false;
    if (live) {
        j0 = P.op_begin[r]; j1 = P.op_begin[r + 1];
        qc = P.q_contig[r]; tc = P.t_contig[r];
        rev = P.same_strand[r] == 0;
        qs = P.q_start[r]; qe = P.q_end[r]; ts = P.t_start[r];
    }
    const bool text_mode = P.t_end != nullptr;
    const bool exempt = text_mode && live && j1 == j0 && !P.has_cg[r];  // no cigar: nothing is dereferenced or checked
    bool ok = live && qc >= 0 && qc < P.n_contigs && tc >= 0 && tc < P.n_contigs && (j1 > j0 || text_mode) && qs >= 0 && ts >= 0 && qs <= qe;
    int64_t q_len = 0, t_len = 0, q_off = 0, t_off = 0;
    if (ok) {
        q_len = P.contig_len[qc]; t_len = P.contig_len[tc];
        q_off = rev ? P.contig_off_rc[qc] : P.contig_off[qc]; t_off = P.contig_off[tc];
        ok = qe <= q_len && ts <= t_len;
    }
    const uint64_t t_base = (uint64_t)(t_off + ts);
    // '-' strand: the query walks down from query_end - 1, i.e. up from q_len - query_end in the reverse-complement copy
    const uint64_t q_base = (uint64_t)(rev ? q_off + (q_len - qe) : q_off + qs);
    uint64_t t_used = 0, q_used = 0, m_sum = 0;
    bool overflow = false;
    uint64_t w_next = (j0 + gl < j1) ? __ldg(P.ops + j0 + gl) : 0;  // one iteration ahead: two loads in flight per lane
    for (int64_t base = j0; __any_sync(FULL, base < j1); base += G) {
        const int64_t j = base + gl;
        const bool have = j < j1;
        const uint64_t w = w_next;
        w_next = (j + G < j1) ? __ldg(P.ops + j + G) : 0;
        const int code = (int)(w >> 56);
        const uint64_t len = w & LEN_MASK;
        if (have && ((len >> 47) || (code == OP_M && len == 0))) overflow = true;  // negative / absurd length, or a 0M (callers drop those)
        const uint64_t tl = (have && code != OP_I) ? len : 0, ql = (have && code != OP_D) ? len : 0;
        uint64_t t_inc = tl, q_inc = ql;
#pragma unroll
        for (int o = 1; o < G; o <<= 1) {
            const uint64_t a = __shfl_up_sync(FULL, t_inc, o, G), b = __shfl_up_sync(FULL, q_inc, o, G);
            if (gl >= (unsigned)o) { t_inc += a; q_inc += b; }
        }
        if (have) {
            uint64_t tp = (uint64_t)ARENA_ALIGN / 2, qp = (uint64_t)ARENA_ALIGN / 2, qflags = 0;
            uint32_t nw = 1;
            if (ok) {  // a record that fails validation still gets one harmless window per op so downstream kernels stay in bounds
                tp = t_base + t_used + (t_inc - tl);
                const uint64_t qd = q_used + (q_inc - ql);
                qp = q_base + qd;
                if (code == OP_M) {
                    nw = (uint32_t)((len + (WINDOW_BASES - 1)) / WINDOW_BASES);
                    m_sum += len;
                    // bounds of this op decide whether its spans may be dereferenced at all
                    const bool inside = (ts + (int64_t)(t_used + t_inc) <= t_len) &&
                                        (rev ? (qe - (int64_t)(qd + len) >= 0) : (qs + (int64_t)(qd + len) <= q_len));
                    if (P.has_exceptions && len > 0 && inside && !overflow) {
                        if (span_has_exception(P.G, tp, tp + len - 1) || span_has_exception(P.G, qp, qp + len - 1))
                            qflags |= QFLAG_EXC;
                    }
                }
            }
            P.tpos[j] = (tp & POS_MASK) | (j == j0 ? TFLAG_HEAD : 0);
            P.qpos[j] = (qp & POS_MASK) | qflags;
            P.op_rec[j] = (uint32_t)r | (j == j0 ? 0x80000000u : 0u);  // top bit: first op of the record
            P.nwin[j] = nw;
        }
        t_used += __shfl_sync(FULL, t_inc, G - 1, G);
        q_used += __shfl_sync(FULL, q_inc, G - 1, G);
    }
    // memory safety: the walk must stay inside both contigs (the reference has UB here, a1-edge vii)
    const unsigned group_mask = G == 32 ? FULL : ((1u << G) - 1u) << (lane & ~(unsigned)(G - 1));
    overflow = (__ballot_sync(FULL, overflow) & group_mask) != 0;
    bool inside = ok && !overflow && ts + (int64_t)t_used <= t_len && (rev ? qe - (int64_t)q_used >= 0 : qs + (int64_t)q_used <= q_len);
    if (text_mode && inside && ((int64_t)q_used != qe - qs || (int64_t)t_used != P.t_end[r] - ts)) inside = false;  // paf_check's sums
    if (live && !inside && !exempt && gl == 0) atomicMin(P.bad_record, (unsigned long long)r);
    if (!inside) m_sum = 0;
#pragma unroll
    for (int o = 16; o; o >>= 1) m_sum += __shfl_xor_sync(FULL, m_sum, o);
    if (lane == 0 && m_sum) atomicAdd(&block_aligned, (unsigned long long)m_sum);
    __syncthreads();
    if (threadIdx.x == 0 && block_aligned) atomicAdd(P.aligned, block_aligned);  // one global atomic per block, not per record
}

// ---- FASTA sequence text -> contiguous sequence bytes (the admission rules of fastaReadToFunction,
// bioioC.c:186-203): newline, space and tab are dropped, every other byte must be a letter or '-'. raw holds the bytes
// between a header line and the next '>'. Two passes over tiles of FA_TILE bytes: kept bytes per tile (+ the position of
// the first inadmissible byte), an exclusive scan of the tile counts, then each tile re-derives its in-tile offsets
// (thread-owned runs of FA_PER_THREAD bytes, block scan through shared memory) and copies its kept bytes out.
constexpr int FA_THREADS = 256, FA_PER_THREAD = 16, FA_TILE = FA_THREADS * FA_PER_THREAD;

__device__ __forceinline__ bool fa_dropped(unsigned char ch) { return ch == '\n' || ch == ' ' || ch == '\t'; }
__device__ __forceinline__ bool fa_admissible(unsigned char ch) {
    return (ch >= 'A' && ch <= 'Z') || (ch >= 'a' && ch <= 'z') || ch == '-';  // isalpha in the C locale, or '-'
}

__global__ void __launch_bounds__(FA_THREADS) fasta_count_kernel(const unsigned char *__restrict__ raw, int64_t n, uint32_t *tile_kept,
                                                                  unsigned long long *first_bad) {
    __shared__ uint32_t warp_sum[FA_THREADS / 32];
    const int64_t b0 = (int64_t)blockIdx.x * FA_TILE + (int64_t)threadIdx.x * FA_PER_THREAD;
    uint32_t kept = 0;
#pragma unroll
    for (int k = 0; k < FA_PER_THREAD; k++) {
        const int64_t i = b0 + k;
        if (i < n) {
            const unsigned char ch = raw[i];
            if (!fa_dropped(ch)) {
                kept++;
                if (!fa_admissible(ch)) atomicMin(first_bad, (unsigned long long)i);
            }
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) kept += __shfl_xor_sync(FULL, kept, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = kept;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < FA_THREADS / 32; w++) t += warp_sum[w];
        tile_kept[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(FA_THREADS) fasta_compact_kernel(const unsigned char *__restrict__ raw, int64_t n,
                                                                    const uint64_t *__restrict__ tile_off, unsigned char *out) {
    __shared__ uint32_t warp_sum[FA_THREADS / 32];
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t b0 = (int64_t)blockIdx.x * FA_TILE + (int64_t)threadIdx.x * FA_PER_THREAD;
    unsigned char keep[FA_PER_THREAD];
    uint32_t kept = 0;
#pragma unroll
    for (int k = 0; k < FA_PER_THREAD; k++) {
        const int64_t i = b0 + k;
        const unsigned char ch = i < n ? raw[i] : (unsigned char)'\n';
        keep[k] = ch;
        kept += !fa_dropped(ch);
    }
    uint32_t incl = kept;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(FULL, incl, o);
        if (lane >= (unsigned)o) incl += v;
    }
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    uint32_t before = incl - kept;
    for (unsigned w = 0; w < warp; w++) before += warp_sum[w];
    unsigned char *dst = out + tile_off[blockIdx.x] + before;
#pragma unroll
    for (int k = 0; k < FA_PER_THREAD; k++)
        if (!fa_dropped(keep[k])) *dst++ = keep[k];
}

// first op of each warp tile: op j owns windows [woff[j], woff[j+1])
__global__ void tile_first_op_kernel(const uint64_t *__restrict__ woff, int64_t n_ops, uint32_t tile_windows,
                                     uint32_t *tile_first_op) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_ops) return;
    const uint64_t w0 = woff[j], w1 = woff[j + 1];
    if (w1 == w0) return;
    for (uint64_t t = (w0 + tile_windows - 1) / tile_windows; t * tile_windows < w1; t++) tile_first_op[t] = (uint32_t)j;
}

}  // namespace pafgpu
