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
constexpr uint64_t QFLAG_REV = 1ull << 63;   // qpos: '-' strand, query walks downwards
constexpr uint64_t QFLAG_EXC = 1ull << 62;   // qpos: the op's target or query span touches a non-ACGT word
constexpr uint64_t TFLAG_HEAD = 1ull << 63;  // tpos: first op of a record
constexpr uint64_t EVEN_BITS = 0x5555555555555555ull;
constexpr int OP_M = 0, OP_I = 1, OP_D = 2, OP_EQ = 3, OP_X = 4;
constexpr unsigned FULL = 0xffffffffu;
constexpr uint32_t INF32 = 0xffffffffu;
constexpr uint64_t INF64 = ~0ull;
constexpr int ARENA_ALIGN = 1024;  // bases; contig offsets and pads are multiples of this

// ------------------------------------------------------------------------------------------------
// Genome arena: 2 bits/base (A0 C1 G2 T3) in 64-bit words of 32 bases, plus an exception plane for
// every other byte: one flag bit per word, a rank directory (exclusive count of flagged words per
// 32-word group) and 32 upper-cased bytes per flagged word.
// ------------------------------------------------------------------------------------------------
struct Genome {
    const uint64_t *words;
    const uint32_t *flags;  // bit (w & 31) of flags[w >> 5]
    const uint32_t *rank;   // flagged words before group w >> 5
    const uint8_t *exc;     // 32 bytes per flagged word
};

__device__ __forceinline__ uint64_t ldg64(const uint64_t *p) { return __ldg(p); }

// 32 consecutive bases starting at arena position p; base k in bits [2k, 2k+1].
__device__ __forceinline__ uint64_t fetch32(const uint64_t *__restrict__ g, uint64_t p) {
    const uint64_t w = p >> 5;
    const unsigned sh = (unsigned)(p & 31) * 2;
    const uint64_t lo = ldg64(g + w), hi = ldg64(g + w + 1);
    return sh ? (lo >> sh) | (hi << (64 - sh)) : lo;
}

// Reverse the order of the 32 two-bit codes and complement them (A<->T, C<->G is bitwise NOT).
__device__ __forceinline__ uint64_t revcomp32(uint64_t v) {
    const uint64_t y = __brevll(v);
    return ~(((y >> 1) & EVEN_BITS) | ((y & EVEN_BITS) << 1));
}

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
    const unsigned code = (unsigned)(ldg64(G.words + w) >> ((p & 31) * 2)) & 3u;
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

// ---- pack: ASCII -> 2-bit words + flag bits. One thread per word, one warp per flag word. --------
__global__ void pack_kernel(const uint8_t *__restrict__ src, int64_t len, uint64_t word0, uint64_t *words,
                            uint32_t *flags, int64_t n_words /* multiple of 32 */) {
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    const int64_t b0 = w * 32;
    uint64_t packed = 0;
    bool exc = false;
    uint8_t c[32];
    if (b0 + 32 <= len && ((uintptr_t)(src + b0) & 15) == 0) {
        const uint4 a = __ldg(reinterpret_cast<const uint4 *>(src + b0));
        const uint4 b = __ldg(reinterpret_cast<const uint4 *>(src + b0) + 1);
        const uint32_t v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < 32; k++) c[k] = (uint8_t)(v[k >> 2] >> ((k & 3) * 8));
    } else {
#pragma unroll
        for (int k = 0; k < 32; k++) c[k] = (b0 + k < len) ? __ldg(src + b0 + k) : (uint8_t)'A';
    }
#pragma unroll
    for (int k = 0; k < 32; k++) {
        const uint8_t u = upper(c[k]);
        const bool acgt = (u == 'A') | (u == 'C') | (u == 'G') | (u == 'T');
        exc |= !acgt;
        const uint64_t code = acgt ? (uint64_t)(((u >> 1) ^ (u >> 2)) & 3u) : 0ull;
        packed |= code << (2 * k);
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
                                const uint32_t *rank, uint8_t *exc, int64_t n_words) {
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_words) return;
    const uint64_t gw = word0 + w;
    const uint32_t f = flags[gw >> 5];
    if (!((f >> (gw & 31)) & 1u)) return;
    const uint64_t r = rank[gw >> 5] + __popc(f & ((1u << (gw & 31)) - 1u));
    for (int k = 0; k < 32; k++) {
        const int64_t b = w * 32 + k;
        exc[r * 32 + k] = b < len ? upper(src[b]) : (uint8_t)'A';
    }
}
// read back bases as the bytes the comparison sees (tests / pafgpu_genome_fetch)
__global__ void unpack_kernel(Genome G, uint64_t p0, int64_t len, uint8_t *out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < len) out[i] = byte_at(G, p0 + i);
}

// ------------------------------------------------------------------------------------------------
// Exclusive scan u32 -> u64, three phases (block sums, scan of sums, apply). out has n + 1 entries.
// ------------------------------------------------------------------------------------------------
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

__global__ void __launch_bounds__(SCAN_THREADS) scan_sums_kernel(const uint32_t *__restrict__ in, int64_t n,
                                                                  uint64_t *block_sums) {
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
    uint64_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const int64_t i = base + (int64_t)k * SCAN_THREADS + threadIdx.x;
        if (i < n) s += in[i];
    }
    uint64_t tot;
    block_incl_scan(s, &tot);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}
// single block: block_sums -> exclusive, total to out_total
__global__ void __launch_bounds__(SCAN_THREADS) scan_spine_kernel(uint64_t *block_sums, int64_t nb,
                                                                   uint64_t *out_total) {
    uint64_t carry = 0;
    for (int64_t base = 0; base < nb; base += SCAN_THREADS) {
        const int64_t i = base + threadIdx.x;
        const uint64_t v = i < nb ? block_sums[i] : 0;
        uint64_t tot;
        const uint64_t inc = block_incl_scan(v, &tot);
        if (i < nb) block_sums[i] = carry + inc - v;
        carry += tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) *out_total = carry;
}
__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(const uint32_t *__restrict__ in, int64_t n,
                                                                   const uint64_t *__restrict__ block_sums,
                                                                   uint64_t *__restrict__ out) {
    // thread t owns SCAN_ITEMS consecutive items so the block scan runs once
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    uint64_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        v[k] = (base + k < n) ? in[base + k] : 0u;
        s += v[k];
    }
    uint64_t tot;
    uint64_t run = block_incl_scan(s, &tot) - s + block_sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (base + k < n) out[base + k] = run;
        run += v[k];
    }
    if (base <= n - 1 && n - 1 < base + SCAN_ITEMS) out[n] = block_sums[blockIdx.x] + tot;
}

// ------------------------------------------------------------------------------------------------
// CIGAR ingest (a4 -> SoA): one warp per record walks its ops in strides of 32, turning the
// reference's running cursors (qi, tj of impl/paf.c:746-775) into warp prefix sums. Per op it
// writes the absolute arena position of its first target and query base, the owning record and the
// number of 32-base work windows (1 for a pass-through op).
// ------------------------------------------------------------------------------------------------
struct IngestParams {
    int64_t n_records, n_ops;
    const int64_t *op_begin;
    const uint64_t *ops;
    const int32_t *q_contig, *t_contig;
    const int64_t *q_start, *q_end, *t_start;
    const uint8_t *same_strand;
    const int64_t *contig_off, *contig_len;
    int64_t n_contigs;
    Genome G;
    uint64_t *tpos, *qpos;
    uint32_t *op_rec, *nwin;
    unsigned long long *bad_record;  // min index of a record that fails validation (init ~0)
    unsigned long long *aligned;     // sum of M lengths
};

__global__ void __launch_bounds__(256) ingest_kernel(IngestParams P) {
    const unsigned lane = threadIdx.x & 31;
    const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (r >= P.n_records) return;
    const int64_t j0 = P.op_begin[r], j1 = P.op_begin[r + 1];
    const int64_t qc = P.q_contig[r], tc = P.t_contig[r];
    const bool rev = P.same_strand[r] == 0;
    const int64_t qs = P.q_start[r], qe = P.q_end[r], ts = P.t_start[r];
    bool ok = qc >= 0 && qc < P.n_contigs && tc >= 0 && tc < P.n_contigs && j1 > j0 && qs >= 0 && ts >= 0 && qs <= qe;
    int64_t q_len = 0, t_len = 0, q_off = 0, t_off = 0;
    if (ok) {
        q_len = P.contig_len[qc]; t_len = P.contig_len[tc];
        q_off = P.contig_off[qc]; t_off = P.contig_off[tc];
        ok = qe <= q_len && ts <= t_len;
    }
    if (!ok) {
        if (lane == 0) atomicMin(P.bad_record, (unsigned long long)r);
        // still give every op a harmless window so downstream kernels stay in bounds
        for (int64_t j = j0 + lane; j < j1; j += 32) {
            P.tpos[j] = (j == j0 ? TFLAG_HEAD : 0) | (uint64_t)ARENA_ALIGN / 2;
            P.qpos[j] = (uint64_t)ARENA_ALIGN / 2;
            P.op_rec[j] = (uint32_t)r;
            P.nwin[j] = 1;
        }
        return;
    }
    const uint64_t t_base = (uint64_t)(t_off + ts);
    const uint64_t q_base = (uint64_t)(rev ? q_off + qe - 1 : q_off + qs);
    uint64_t t_used = 0, q_used = 0, m_sum = 0;
    bool overflow = false;
    for (int64_t base = j0; base < j1; base += 32) {
        const int64_t j = base + lane;
        const bool have = j < j1;
        const uint64_t w = have ? P.ops[j] : 0;
        const int code = (int)(w >> 56);
        const uint64_t len = w & LEN_MASK;
        if (have && ((len >> 47) || (code == OP_M && len == 0))) overflow = true;  // negative / absurd length, or a 0M (callers drop those)
        const uint64_t tl = (have && code != OP_I) ? len : 0, ql = (have && code != OP_D) ? len : 0;
        const uint64_t t_inc = warp_incl_scan(tl, lane), q_inc = warp_incl_scan(ql, lane);
        if (have) {
            const uint64_t tp = t_base + t_used + (t_inc - tl);
            const uint64_t qd = q_used + (q_inc - ql);
            const uint64_t qp = rev ? q_base - qd : q_base + qd;
            uint64_t qflags = rev ? QFLAG_REV : 0;
            uint32_t nw = 1;
            if (code == OP_M) {
                nw = (uint32_t)((len + 31) >> 5);
                m_sum += len;
                // bounds of this op decide whether its spans may be dereferenced at all
                const bool inside = (ts + (int64_t)(t_used + t_inc) <= t_len) &&
                                    (rev ? (qe - (int64_t)(qd + len) >= 0) : (qs + (int64_t)(qd + len) <= q_len));
                if (len > 0 && inside && !overflow) {
                    const uint64_t qlo = rev ? qp - (len - 1) : qp;
                    if (span_has_exception(P.G, tp, tp + len - 1) || span_has_exception(P.G, qlo, qlo + len - 1))
                        qflags |= QFLAG_EXC;
                }
            }
            P.tpos[j] = (tp & POS_MASK) | (j == j0 ? TFLAG_HEAD : 0);
            P.qpos[j] = (qp & POS_MASK) | qflags;
            P.op_rec[j] = (uint32_t)r;
            P.nwin[j] = nw;
        }
        t_used += __shfl_sync(FULL, t_inc, 31);
        q_used += __shfl_sync(FULL, q_inc, 31);
    }
    overflow = __any_sync(FULL, overflow);
    // memory safety: the walk must stay inside both contigs (the reference has UB here, a1-edge vii)
    const bool inside = !overflow && ts + (int64_t)t_used <= t_len &&
                        (rev ? qe - (int64_t)q_used >= 0 : qs + (int64_t)q_used <= q_len);
    if (!inside && lane == 0) atomicMin(P.bad_record, (unsigned long long)r);
#pragma unroll
    for (int o = 16; o; o >>= 1) m_sum += __shfl_xor_sync(FULL, m_sum, o);
    if (lane == 0 && m_sum) atomicAdd(P.aligned, (unsigned long long)m_sum);
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

// ------------------------------------------------------------------------------------------------
// The base-comparison scan (a1 + a7). Work axis = 32-base windows of all ops of the batch
// (pass-through ops own one window), cut into fixed tiles of 32*ITERS windows; one warp per tile,
// tiles handed out by an atomic ticket so megabase CIGARs spread over the whole grid. Per window a
// lane loads 2x64-bit target and query words, funnel-shifts them into phase, reverse-complements the
// query word in registers on '-' strands, XORs, folds to one mismatch bit per base and derives the
// run-start bits. Output slots come from a warp scan + a decoupled look-back over tile aggregates
// (single pass over the sequence data). Run lengths are differences of run starts; a run still open
// at the end of a tile is closed by stitch_kernel.
// ------------------------------------------------------------------------------------------------
struct EncodeParams {
    Genome G;
    const uint64_t *ops, *woff, *tpos, *qpos;
    const uint32_t *op_rec, *tile_first_op;
    int64_t n_ops;
    uint64_t n_win;
    int64_t n_tiles;
    uint64_t *out;
    int64_t out_cap;
    unsigned long long *tile_desc;  // look-back over CTA tiles: status << 62 | count
    uint64_t *tile_excl;            // exclusive output rank of every warp tile
    int64_t n_btiles;               // CTA tiles = ceil(n_tiles / warps per CTA)
    uint64_t *tile_first_closer;    // absolute window-axis position of the first run start / op end
    int64_t *tile_open_slot;        // output slot of a run left open, -1 if none (pre-set)
    int64_t *rec_out_begin;
    unsigned long long *rec_mism;
    unsigned int *ticket;
    unsigned long long *total_out;
    int debug;  // experiments only (PAFGPU_DEBUG): 1 skip look-back, 2 skip emission, 4 skip tallies
};

constexpr unsigned long long DESC_AGG = 1ull << 62, DESC_PREFIX = 2ull << 62, DESC_VAL = (1ull << 62) - 1;

__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// mismatch bit of one aligned base of an op, through the byte plane (slow, exact)
__device__ __forceinline__ unsigned slow_mismatch(const Genome &G, uint64_t tp, uint64_t qp, bool rev) {
    const uint8_t t = byte_at(G, tp);
    uint8_t q = byte_at(G, qp);
    if (rev) q = complement_upper(q);
    return t != q;
}

// sign-extend the low 56 bits
__device__ __forceinline__ int64_t sext56(uint64_t v) { return (int64_t)(v << 8) >> 8; }
// 2-bit code (A0 C1 G2 T3) at arena position p of a plain word
__device__ __forceinline__ unsigned code_of(uint64_t word, uint64_t p) { return (unsigned)(word >> ((p & 31) * 2)) & 3u; }

// table flags (top byte of the tB entry)
constexpr uint64_t EF_HEAD = 1ull << 56, EF_PASS = 1ull << 57, EF_REV = 1ull << 58, EF_EXC = 1ull << 59;
// META bits
constexpr uint32_t MB_PASS = 1u << 6, MB_LAST = 1u << 7, MB_ACTIVE = 1u << 8, MB_WI0 = 1u << 9, MB_REV = 1u << 10,
                   MB_SLOW = 1u << 11, MB_PM = 1u << 12;
constexpr int MS_SHT = 13, MS_SHQ = 18, MS_IDX = 23;

template <int ITERS, int WARPS>
__global__ void __launch_bounds__(32 * WARPS) encode_kernel(EncodeParams P) {
    constexpr uint32_t TW = 32u * ITERS;
    constexpr int TAB = (int)TW + 32;  // a tile of TW windows overlaps at most TW + 1 ops
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint64_t *__restrict__ gw = P.G.words;
    __shared__ unsigned s_ticket;
    __shared__ uint32_t s_cnt[WARPS];
    __shared__ uint64_t s_excl;
    // per-warp op table of the current tile: everything a window needs to know about its op
    __shared__ uint64_t t_tb[WARPS][TAB];   // flags << 56 | (arena position of the op's base under tile window 0) & 2^56-1
    __shared__ int64_t t_qb[WARPS][TAB];    // same for the query (walks down on '-' strands)
    __shared__ uint64_t t_rem[WARPS][TAB];  // M: op bases left at tile window 0 (signed); pass-through: the op word itself
    __shared__ uint32_t t_rec[WARPS][TAB];
    __shared__ uint32_t s_bm[WARPS][ITERS];  // bit p: an op starts at tile window p

    unsigned nxt = 0;
    if (threadIdx.x == 0) nxt = atomicAdd(P.ticket, 1u);
    for (;;) {
        // one ticket per CTA tile (WARPS consecutive warp tiles): one look-back link per CTA. The next ticket
        // is drawn a tile ahead so its round trip hides behind the work.
        if (threadIdx.x == 0) {
            s_ticket = nxt;
            nxt = atomicAdd(P.ticket, 1u);
        }
        __syncthreads();
        const int64_t btile = s_ticket;
        if (btile >= P.n_btiles) return;
        const int64_t tile = btile * WARPS + wid;
        const bool live = tile < P.n_tiles;
        const uint64_t s0 = (uint64_t)tile * TW;
        const uint64_t tile_end = min(s0 + TW, P.n_win);

        // ---- op table: lane k of group g owns op cur + 32 g + k (all loads independent: one round trip)
        unsigned adj = 0;
        if (lane < ITERS) s_bm[wid][lane] = 0;
        __syncwarp();
        if (live) {
            const int64_t cur = P.tile_first_op[tile];
            bool more = true;
            for (int g = 0; more; g++) {
                const int64_t k = cur + 32 * g + lane;
                const bool valid = k < P.n_ops;
                uint64_t wo = INF64, opw = 0, tpw = 0, qpw = 0;
                uint32_t rec = 0;
                if (valid) {
                    wo = __ldg(P.woff + k); opw = __ldg(P.ops + k); tpw = __ldg(P.tpos + k); qpw = __ldg(P.qpos + k);
                    rec = __ldg(P.op_rec + k);
                }
                const bool in = valid && wo < tile_end;
                if (in) {
                    const int e = 32 * g + (int)lane;
                    const int64_t back = 32 * (int64_t)(s0 - wo);  // op bases before tile window 0 (negative: op starts inside)
                    const bool rev = (qpw & QFLAG_REV) != 0, pass = (int)(opw >> 56) != OP_M;
                    uint64_t fl = (tpw & TFLAG_HEAD) ? EF_HEAD : 0;
                    if (pass) fl |= EF_PASS;
                    if (rev) fl |= EF_REV;
                    if (qpw & QFLAG_EXC) fl |= EF_EXC;
                    t_tb[wid][e] = fl | ((uint64_t)((int64_t)(tpw & POS_MASK) + back) & LEN_MASK);
                    t_qb[wid][e] = rev ? (int64_t)(qpw & POS_MASK) - back : (int64_t)(qpw & POS_MASK) + back;
                    t_rem[wid][e] = pass ? opw : (uint64_t)((int64_t)(opw & LEN_MASK) - back);
                    t_rec[wid][e] = rec;
                    if (wo >= s0) atomicOr(&s_bm[wid][(wo - s0) >> 5], 1u << ((wo - s0) & 31));
                }
                if (g == 0) adj = __shfl_sync(FULL, (unsigned)(wo == s0), 0);
                more = __all_sync(FULL, in);
            }
        }
        __syncwarp();

        uint64_t D[ITERS], MM[ITERS];
        uint32_t META[ITERS];
        uint32_t FC[ITERS];  // first closer (tile-relative window-axis position), later: my closer
        uint32_t EX[ITERS];  // exclusive output rank inside the tile
        uint32_t tile_count = 0;
        unsigned carry_pm = 0;
        uint32_t acc_rec = INF32, acc_mm = 0;

        // ---- phase A: window -> op through the start bitmap, then issue every sequence load of the tile
        uint64_t TL[ITERS], TH[ITERS], QL[ITERS], QH[ITERS];
        uint64_t pm_t = 0, pm_q = 0;  // lane 0, first window only: words holding the base before the window
        uint32_t pre = 0;
#pragma unroll
        for (int it = 0; it < ITERS; it++) {
            const uint32_t srel = 32u * it + lane;
            const bool active = live && s0 + srel < P.n_win;
            const uint32_t w = s_bm[wid][it];
            const uint32_t idx = pre + __popc(w & (0xffffffffu >> (31 - lane))) - adj;
            pre += __popc(w);
            uint32_t meta = 0;
            TL[it] = TH[it] = QL[it] = QH[it] = 0;
            if (active) {
                const uint64_t F = t_tb[wid][idx];
                meta = MB_ACTIVE | (idx << MS_IDX) | (((w >> lane) & 1u) ? MB_WI0 : 0u);
                if (F & EF_PASS) {
                    meta |= MB_PASS;
                } else {
                    const int64_t back = 32 * (int64_t)srel;
                    const bool rev = (F & EF_REV) != 0;
                    const uint64_t tp = (uint64_t)(sext56(F) + back);
                    const int64_t qb = t_qb[wid][idx];
                    const uint64_t qp = (uint64_t)(rev ? qb - back : qb + back);
                    const int64_t rem = (int64_t)t_rem[wid][idx] - back;
                    const uint32_t n = rem < 32 ? (uint32_t)rem : 32u;
                    if (rem <= 32) meta |= MB_LAST;
                    if (rev) meta |= MB_REV;
                    const uint64_t qlo = rev ? qp - 31 : qp;  // first base of the 32 the query word is cut from
                    bool slow = false;
                    if (F & EF_EXC) {
                        const uint64_t ql = rev ? qp - (n - 1) : qp;
                        slow = span_has_exception(P.G, tp, tp + n - 1) || span_has_exception(P.G, ql, ql + n - 1);
                    }
                    if (slow) meta |= MB_SLOW;
                    meta |= n | ((uint32_t)(tp & 31) << MS_SHT) | ((uint32_t)(qlo & 31) << MS_SHQ);
                    TL[it] = ldg64(gw + (tp >> 5)); TH[it] = ldg64(gw + (tp >> 5) + 1);
                    QL[it] = ldg64(gw + (qlo >> 5)); QH[it] = ldg64(gw + (qlo >> 5) + 1);
                    if (it == 0 && lane == 0 && !(meta & MB_WI0)) {
                        // mismatch state of the base just before the tile (same op)
                        const uint64_t qprev = rev ? qp + 1 : qp - 1;
                        if (F & EF_EXC) {
                            if (slow_mismatch(P.G, tp - 1, qprev, rev)) meta |= MB_PM;
                        } else {
                            pm_t = ldg64(gw + ((tp - 1) >> 5));
                            pm_q = ldg64(gw + (qprev >> 5));
                        }
                    }
                }
            }
            META[it] = meta;
        }

        // ---- phase B: mismatch bits, run starts, output ranks
#pragma unroll
        for (int it = 0; it < ITERS; it++) {
            uint32_t meta = META[it];
            const bool active = (meta & MB_ACTIVE) != 0;
            const uint32_t idx = meta >> MS_IDX;
            const uint32_t n = meta & 63u;
            const bool wi0 = (meta & MB_WI0) != 0;
            uint64_t m = 0, d = 0;
            uint32_t fc = INF32, rec = INF32;
            bool head = false;
            if (active) {
                rec = t_rec[wid][idx];
                head = wi0 && (t_tb[wid][idx] & EF_HEAD);
                if (meta & MB_PASS) {
                    m = t_rem[wid][idx];
                    d = 1;
                } else {
                    const bool rev = (meta & MB_REV) != 0;
                    const uint64_t valid = n == 32 ? ~0ull : ((1ull << (2 * n)) - 1);
                    if (!(meta & MB_SLOW)) {
                        const unsigned sht = ((meta >> MS_SHT) & 31u) * 2, shq = ((meta >> MS_SHQ) & 31u) * 2;
                        const uint64_t t = sht ? (TL[it] >> sht) | (TH[it] << (64 - sht)) : TL[it];
                        uint64_t q = shq ? (QL[it] >> shq) | (QH[it] << (64 - shq)) : QL[it];
                        if (rev) q = revcomp32(q);
                        const uint64_t x = t ^ q;
                        m = (x | (x >> 1)) & EVEN_BITS & valid;
                    } else {
                        const int64_t back = 32 * (int64_t)(32u * it + lane);
                        const uint64_t tp = (uint64_t)(sext56(t_tb[wid][idx]) + back);
                        const int64_t qb = t_qb[wid][idx];
                        const uint64_t qp = (uint64_t)(rev ? qb - back : qb + back);
                        for (uint32_t k = 0; k < n; k++)
                            m |= (uint64_t)slow_mismatch(P.G, tp + k, rev ? qp - k : qp + k, rev) << (2 * k);
                    }
                    if (it == 0 && lane == 0 && !wi0 && !(t_tb[wid][idx] & EF_EXC)) {
                        const int64_t qb = t_qb[wid][idx];
                        const uint64_t tp = (uint64_t)sext56(t_tb[wid][idx]);
                        const uint64_t qprev = (uint64_t)(rev ? qb + 1 : qb - 1);
                        const unsigned tc = code_of(pm_t, tp - 1);
                        unsigned qc = code_of(pm_q, qprev);
                        if (rev) qc ^= 3u;
                        if (tc != qc) meta |= MB_PM;
                    }
                    // per-record mismatch tally, flushed when the lane moves to another record
                    const uint32_t mm = __popcll(m);
                    if (rec != acc_rec && !(P.debug & 4)) {
                        if (acc_mm) atomicAdd(P.rec_mism + acc_rec, (unsigned long long)acc_mm);
                        acc_rec = rec;
                        acc_mm = 0;
                    }
                    acc_mm += mm;
                }
            }
            // previous base's mismatch bit: lane-1 (same iteration) or lane 31 of the last iteration
            {
                const unsigned hi = (unsigned)(m >> 62) & 1u;
                unsigned pm = __shfl_up_sync(FULL, hi, 1);
                if (lane == 0) pm = it == 0 ? ((meta & MB_PM) ? 1u : 0u) : carry_pm;
                carry_pm = __shfl_sync(FULL, hi, 31);
                if (active && !(meta & MB_PASS)) {
                    const uint64_t valid = n == 32 ? ~0ull : ((1ull << (2 * n)) - 1);
                    d = ((m ^ ((m << 2) | pm)) & EVEN_BITS & valid) | (wi0 && n ? 1ull : 0ull);
                    const uint32_t xrel = 32u * (32u * it + lane);
                    if (d) fc = xrel + ((uint32_t)(__ffsll((long long)d) - 1) >> 1);
                    else if (meta & MB_LAST) fc = xrel + n;
                }
            }
            const uint32_t cnt = __popcll(d);
            uint32_t inc = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, inc, o);
                if (lane >= (unsigned)o) inc += t;
            }
            EX[it] = tile_count + inc - cnt;
            tile_count += __shfl_sync(FULL, inc, 31);
            // first output op of a record, tile-relative for now (rec_out_fix_kernel adds the tile's rank)
            if (head) P.rec_out_begin[rec] = (int64_t)(((uint64_t)tile << 16) | EX[it]);
            D[it] = d; MM[it] = m; META[it] = meta; FC[it] = fc;
        }

        // ---- per-record mismatch tallies
        {
            const uint32_t r0 = __shfl_sync(FULL, acc_rec, 0);
            if (__all_sync(FULL, acc_rec == r0)) {
                const uint32_t sum = __reduce_add_sync(FULL, acc_mm);
                if (lane == 0 && sum && r0 != INF32) atomicAdd(P.rec_mism + r0, (unsigned long long)sum);
            } else if (acc_mm && acc_rec != INF32) {
                atomicAdd(P.rec_mism + acc_rec, (unsigned long long)acc_mm);
            }
        }

        // ---- output rank of this warp tile: CTA-level sum, then one decoupled look-back per CTA (warp 0)
        if (lane == 0) s_cnt[wid] = tile_count;
        __syncthreads();
        uint32_t warp_off = 0, cta_count = 0;
#pragma unroll
        for (int k = 0; k < WARPS; k++) {
            const uint32_t c = s_cnt[k];
            if (k < (int)wid) warp_off += c;
            cta_count += c;
        }
        if (wid == 0) {
            uint64_t bex = 0;
            if (btile > 0 && !(P.debug & 1)) {
                if (lane == 0) st_relaxed(P.tile_desc + btile, DESC_AGG | cta_count);
                int64_t look = btile - 1;
                for (;;) {
                    const int64_t i = look - lane;
                    unsigned long long dsc = DESC_PREFIX;
                    if (i >= 0) {
                        unsigned ns = 32;
                        while (((dsc = ld_relaxed(P.tile_desc + i)) >> 62) == 0) {
                            __nanosleep(ns);
                            if (ns < 256) ns <<= 1;
                        }
                    }
                    const unsigned has_prefix = __ballot_sync(FULL, (dsc >> 62) == 2);
                    uint64_t v = dsc & DESC_VAL;
                    if (has_prefix) {
                        const unsigned first = __ffs(has_prefix) - 1;
                        if (lane > first) v = 0;
                    }
#pragma unroll
                    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
                    bex += v;
                    if (has_prefix) break;
                    look -= 32;
                }
            }
            if (lane == 0) {
                st_relaxed(P.tile_desc + btile, DESC_PREFIX | (bex + cta_count));
                if (btile == P.n_btiles - 1) *P.total_out = bex + cta_count;
                s_excl = bex;
            }
        }
        __syncthreads();
        const uint64_t excl = s_excl + warp_off;
        if (!live || (P.debug & 2)) continue;
        if (lane == 0) P.tile_excl[tile] = excl;

        // ---- emission, last iteration first so each lane knows where its trailing run closes
        uint32_t later_min = INF32;
#pragma unroll
        for (int it = ITERS - 1; it >= 0; it--) {
            uint32_t sfx = FC[it];  // inclusive suffix min over lanes
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_down_sync(FULL, sfx, o);
                if (lane + o < 32) sfx = min(sfx, t);
            }
            uint32_t closer = __shfl_down_sync(FULL, sfx, 1);
            if (lane == 31) closer = INF32;
            closer = min(closer, later_min);
            later_min = min(later_min, __shfl_sync(FULL, sfx, 0));

            const uint32_t meta = META[it];
            if (!(meta & (1u << 8))) continue;
            uint64_t slot = excl + EX[it];
            if (meta & (1u << 6)) {
                if ((int64_t)slot < P.out_cap) P.out[slot] = MM[it];
            } else {
                uint64_t d = D[it];
                const uint64_t m = MM[it];
                const uint32_t n = meta & 63u;
                const uint32_t xrel = 32u * (32u * it + lane);
                while (d) {
                    const uint32_t b2 = (uint32_t)__ffsll((long long)d) - 1;
                    d &= d - 1;
                    const uint64_t typ = (uint64_t)(((m >> b2) & 1) ? OP_X : OP_EQ) << 56;
                    uint64_t len;
                    if (d) len = ((uint32_t)__ffsll((long long)d) - 1 - b2) >> 1;
                    else if (meta & (1u << 7)) len = n - (b2 >> 1);
                    else if (closer != INF32) len = closer - (xrel + (b2 >> 1));
                    else {
                        len = 32ull * s0 + xrel + (b2 >> 1);  // provisional: absolute start, fixed by stitch_kernel
                        P.tile_open_slot[tile] = (int64_t)slot;
                    }
                    if ((int64_t)slot < P.out_cap) P.out[slot] = len | typ;
                    slot++;
                }
            }
        }
        if (lane == 0) P.tile_first_closer[tile] = later_min == INF32 ? INF64 : 32ull * s0 + later_min;
    }
}

// A run left open at the end of tile t closes at the first run start / op end of a later tile.
__global__ void stitch_kernel(int64_t n_tiles, const int64_t *tile_open_slot, const uint64_t *tile_first_closer,
                              uint64_t *out, int64_t out_cap) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tiles) return;
    const int64_t slot = tile_open_slot[t];
    if (slot < 0 || slot >= out_cap) return;
    int64_t u = t + 1;
    while (u < n_tiles && tile_first_closer[u] == INF64) u++;
    if (u >= n_tiles) return;  // cannot happen: the op's end is always a closer
    const uint64_t rec = out[slot];
    out[slot] = (tile_first_closer[u] - (rec & LEN_MASK)) | (rec & ~LEN_MASK);
}

// rec_out_begin holds (warp tile << 16 | rank inside the tile); add the tile's exclusive rank
__global__ void rec_out_fix_kernel(int64_t n_records, int64_t *rec_out_begin, const uint64_t *tile_excl,
                                   const unsigned long long *total_out) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r > n_records) return;
    if (r == n_records) { rec_out_begin[r] = (int64_t)*total_out; return; }
    const uint64_t v = (uint64_t)rec_out_begin[r];
    rec_out_begin[r] = (int64_t)(tile_excl[v >> 16] + (v & 0xffff));
}

}  // namespace pafgpu
