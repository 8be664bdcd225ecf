// encode.cuh - the base-comparison scan (a1 + a7 of SURVEY.md section 8): M ops -> =/X runs.
// Restates paf_encode_mismatches (impl/paf.c:739-784); nothing is copied from it.
//
// Work axis = 64-base windows of all ops of the batch (an M op of length n owns ceil(n/64) windows,
// every other op owns one), cut into warp tiles of 256 windows, so megabase CIGARs spread over the
// whole grid and no CTA ever waits for another. Two kernels, one pass over the sequences:
//   compare_kernel  per tile: op lanes load the tile's ops once and leave, per op, everything a window
//                   needs in shared memory (word index + bit phase of target and query, bases left,
//                   flags) plus a bitmap of the windows at which an op starts, so a window finds its op
//                   with one popcount. Every lane then issues the 6 x 64-bit loads of several windows
//                   at once, funnel-shifts the bit planes into phase, BREV + NOT the query on '-'
//                   strands, one LOP3 per plane -> one mismatch bit per base. It stores the 64-bit
//                   mismatch mask of every window, counts run starts per tile and tallies mismatches
//                   per record.
//   (scan of the per-tile counts -> output rank of every tile)
//   emit_kernel     per tile: re-reads the masks (1/4 of the sequence bytes), scatters run starts into a
//                   per-warp list in shared memory; a run's length is the distance to the next entry,
//                   so the list is turned into CigarRecord words and stored with coalesced 8-byte
//                   stores. A run still open at the end of a tile is closed by stitch_kernel.
// An earlier single-kernel version (decoupled look-back between tiles) lost more time waiting on
// predecessor tiles than the mask round trip costs; see DESIGN.md.
#pragma once
#include "kernels.cuh"

namespace pafgpu {

struct EncodeParams {
    Genome G;
    const uint64_t *ops, *woff, *tpos, *qpos;
    const uint32_t *op_rec, *tile_first_op;
    int64_t n_ops;
    uint64_t n_win;
    int64_t n_tiles;
    uint64_t *out;
    int64_t out_cap;
    uint64_t *masks;                // n_win mismatch masks (scratch between the two kernels)
    uint32_t *tile_count;           // run starts + pass-through ops per warp tile
    const uint64_t *tile_excl;      // exclusive scan of tile_count: output rank of every warp tile
    uint64_t *tile_first_closer;    // absolute window-axis coordinate (64 * window + base) of the tile's first run start / op end
    int64_t *tile_open_slot;        // output slot of a run left open, -1 if none (pre-set)
    int64_t *rec_out_begin;
    unsigned long long *rec_mism;
    int debug;  // experiments only (PAFGPU_DEBUG): 4 skip tallies
};

constexpr int WIN_BASES = 64;
constexpr int TILE_ITERS = 8;                 // windows per lane per warp tile
constexpr int TILE_WIN = 32 * TILE_ITERS;     // windows per warp tile
constexpr int TILE_TAB = TILE_WIN + 2;        // a tile overlaps at most TILE_WIN + 1 ops

// mismatch bit of one aligned base of an op, through the byte plane (slow, exact)
__device__ __forceinline__ unsigned slow_mismatch(const Genome &G, uint64_t tp, uint64_t qp, bool rev) {
    const uint8_t t = byte_at(G, tp);
    uint8_t q = byte_at(G, qp);
    if (rev) q = complement_upper(q);
    return t != q;
}


// per-op table entry of compare_kernel (16 bytes, read with one LDS.128)
struct __align__(16) OpEntry {
    int32_t tw;      // word index of the target base under tile window 0 (may be negative for ops starting later)
    int32_t qw;      // word index of the lowest query base of tile window 0's span
    int32_t rem;     // M: op bases left at tile window 0, clamped to 2^30
    uint32_t misc;   // bits 0-4 target phase, 5-9 query phase, then flags
};
constexpr uint32_t EF_HEAD = 1u << 10, EF_PASS = 1u << 11, EF_REV = 1u << 12, EF_EXC = 1u << 13;

// META bits of one window
constexpr uint32_t MB_N = 127u, MB_PASS = 1u << 7, MB_LAST = 1u << 8, MB_ACTIVE = 1u << 9, MB_WI0 = 1u << 10,
                   MB_REV = 1u << 11, MB_SLOW = 1u << 12, MB_PM = 1u << 13, MB_HEAD = 1u << 14;
constexpr int MS_IDX = 16;  // table index, 9 bits
constexpr int MS_GAP = 25;  // op-start windows: 64 - bases in the previous window, 6 bits

template <int WARPS>
struct CompareSmem {
    OpEntry tab[WARPS][TILE_TAB];
    uint32_t rec[WARPS][TILE_TAB];
    uint32_t bm[WARPS][TILE_ITERS];
};

// index of the op owning tile window `srel`, from the bitmap of op starts (bit p of bm: an op starts at window p)
__device__ __forceinline__ uint32_t op_index(const uint32_t *bm, int it, unsigned lane, uint32_t &pre, unsigned adj, uint32_t &w) {
    w = bm[it];
    const uint32_t idx = pre + __popc(w & (0xffffffffu >> (31 - lane))) - adj;
    pre += __popc(w);
    return idx;
}

template <int WARPS, int BATCH>
__global__ void __launch_bounds__(32 * WARPS) compare_kernel(EncodeParams P) {
    static_assert(TILE_ITERS % BATCH == 0, "BATCH must divide the iterations of a tile");
    __shared__ CompareSmem<WARPS> S;
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint64_t *__restrict__ gw = P.G.words;

    for (int64_t tile = (int64_t)blockIdx.x * WARPS + wid; tile < P.n_tiles; tile += (int64_t)gridDim.x * WARPS) {
        const uint64_t s0 = (uint64_t)tile * TILE_WIN;
        const uint64_t tile_end = min(s0 + (uint64_t)TILE_WIN, P.n_win);
        const uint32_t n_act = (uint32_t)(tile_end - s0);

        // ---- op table: lane k of group g owns op cur + 32 g + k (all loads independent: one round trip)
        unsigned adj = 0;
        __syncwarp();
        if (lane < TILE_ITERS) S.bm[wid][lane] = 0;
        __syncwarp();
        {
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
                    const int64_t back = WIN_BASES * (int64_t)(s0 - wo);  // op bases before tile window 0 (negative: op starts inside)
                    const bool rev = (qpw & QFLAG_REV) != 0, pass = (int)(opw >> 56) != OP_M;
                    const int64_t tb = (int64_t)(tpw & POS_MASK) + back;
                    // lowest query base of tile window 0's span: '+' walks up from it, '-' starts 63 below the top base
                    const int64_t qb = rev ? (int64_t)(qpw & POS_MASK) - back - (WIN_BASES - 1) : (int64_t)(qpw & POS_MASK) + back;
                    int64_t rem = (int64_t)(opw & LEN_MASK) - back;
                    if (rem > (1ll << 30)) rem = 1ll << 30;
                    OpEntry en;
                    en.tw = (int32_t)(tb >> 5);
                    en.qw = (int32_t)(qb >> 5);
                    en.rem = (int32_t)rem;
                    en.misc = (uint32_t)(tb & 31) | ((uint32_t)(qb & 31) << 5) | (pass ? EF_PASS : 0u) | (rev ? EF_REV : 0u) |
                              ((qpw & QFLAG_EXC) ? EF_EXC : 0u);
                    S.tab[wid][e] = en;
                    S.rec[wid][e] = rec & 0x7fffffffu;  // top bit: first op of its record
                    if (wo >= s0) atomicOr(&S.bm[wid][(wo - s0) >> 5], 1u << ((wo - s0) & 31));
                }
                if (g == 0) adj = __shfl_sync(FULL, (unsigned)(wo == s0), 0);
                more = __all_sync(FULL, in);
            }
        }
        __syncwarp();

        uint32_t tile_count = 0, pre = 0;
        uint32_t carry_m = 0;
        uint32_t acc_rec = INF32, acc_mm = 0;
#pragma unroll
        for (int hb = 0; hb < TILE_ITERS; hb += BATCH) {
            // ---- phase A: window -> op, then issue every sequence load of the batch
            uint32_t META[BATCH];
            uint64_t T0[BATCH], T1[BATCH], T2[BATCH], Q0[BATCH], Q1[BATCH], Q2[BATCH];
            uint64_t pm_t = 0, pm_q = 0;  // lane 0, first window only: words holding the base before the window
#pragma unroll
            for (int b = 0; b < BATCH; b++) {
                const int it = hb + b;
                const uint32_t srel = 32u * it + lane;
                uint32_t w;
                const uint32_t idx = op_index(S.bm[wid], it, lane, pre, adj, w);
                uint32_t meta = 0;
                T0[b] = T1[b] = T2[b] = Q0[b] = Q1[b] = Q2[b] = 0;
                if (srel < n_act) {
                    const OpEntry en = S.tab[wid][idx];
                    meta = MB_ACTIVE | (idx << MS_IDX) | (((w >> lane) & 1u) ? MB_WI0 : 0u);
                    if (en.misc & EF_PASS) {
                        meta |= MB_PASS;
                    } else {
                        const bool rev = (en.misc & EF_REV) != 0;
                        const int32_t rem = en.rem - WIN_BASES * (int32_t)srel;
                        const uint32_t n = rem < WIN_BASES ? (uint32_t)rem : (uint32_t)WIN_BASES;
                        if (rev) meta |= MB_REV;
                        const int32_t tw = en.tw + 2 * (int32_t)srel;
                        const int32_t qw = rev ? en.qw - 2 * (int32_t)srel : en.qw + 2 * (int32_t)srel;
                        bool slow = false;
                        if (en.misc & EF_EXC) {
                            const uint64_t tp = ((uint64_t)(uint32_t)tw << 5) + (en.misc & 31u);
                            const uint64_t ql = ((uint64_t)(uint32_t)qw << 5) + ((en.misc >> 5) & 31u) + (rev ? WIN_BASES - n : 0u);
                            slow = span_has_exception(P.G, tp, tp + n - 1) || span_has_exception(P.G, ql, ql + n - 1);
                        }
                        if (slow) meta |= MB_SLOW;
                        meta |= n;
                        const uint64_t *tpn = gw + tw, *qpn = gw + qw;
                        T0[b] = ldg64(tpn); T1[b] = ldg64(tpn + 1); T2[b] = ldg64(tpn + 2);
                        Q0[b] = ldg64(qpn); Q1[b] = ldg64(qpn + 1); Q2[b] = ldg64(qpn + 2);
                        if (it == 0 && lane == 0 && !(meta & MB_WI0)) {
                            // mismatch state of the base just before the tile (same op)
                            const uint64_t tp = ((uint64_t)(uint32_t)tw << 5) + (en.misc & 31u);
                            const uint64_t qtop = ((uint64_t)(uint32_t)qw << 5) + ((en.misc >> 5) & 31u) + (rev ? WIN_BASES - 1 : 0);
                            const uint64_t qprev = rev ? qtop + 1 : qtop - 1;
                            if (en.misc & EF_EXC) {
                                if (slow_mismatch(P.G, tp - 1, qprev, rev)) meta |= MB_PM;
                            } else {
                                pm_t = ldg64(gw + ((tp - 1) >> 5));
                                pm_q = ldg64(gw + (qprev >> 5));
                            }
                        }
                    }
                }
                META[b] = meta;
            }
            // ---- phase B: mismatch bits (stored for emit_kernel), run-start count, per-record tallies
#pragma unroll
            for (int b = 0; b < BATCH; b++) {
                const int it = hb + b;
                uint32_t meta = META[b];
                const uint32_t srel = 32u * it + lane;
                const bool active = (meta & MB_ACTIVE) != 0;
                const uint32_t idx = (meta >> MS_IDX) & 511u;
                const uint32_t n = meta & MB_N;
                const bool wi0 = (meta & MB_WI0) != 0;
                const bool isM = active && !(meta & MB_PASS);
                uint32_t m0 = 0, m1 = 0;
                if (isM) {
                    const uint32_t rec = S.rec[wid][idx];
                    const bool rev = (meta & MB_REV) != 0;
                    const uint32_t misc = S.tab[wid][idx].misc;
                    if (!(meta & MB_SLOW)) {
                        const unsigned sht = misc & 31u, shq = (misc >> 5) & 31u;
                        const uint32_t tl0 = __funnelshift_r((uint32_t)T0[b], (uint32_t)T1[b], sht);
                        const uint32_t tl1 = __funnelshift_r((uint32_t)T1[b], (uint32_t)T2[b], sht);
                        const uint32_t th0 = __funnelshift_r((uint32_t)(T0[b] >> 32), (uint32_t)(T1[b] >> 32), sht);
                        const uint32_t th1 = __funnelshift_r((uint32_t)(T1[b] >> 32), (uint32_t)(T2[b] >> 32), sht);
                        uint32_t ql0 = __funnelshift_r((uint32_t)Q0[b], (uint32_t)Q1[b], shq);
                        uint32_t ql1 = __funnelshift_r((uint32_t)Q1[b], (uint32_t)Q2[b], shq);
                        uint32_t qh0 = __funnelshift_r((uint32_t)(Q0[b] >> 32), (uint32_t)(Q1[b] >> 32), shq);
                        uint32_t qh1 = __funnelshift_r((uint32_t)(Q1[b] >> 32), (uint32_t)(Q2[b] >> 32), shq);
                        if (rev) {  // reverse the 64 bases (swap + BREV) and complement (NOT of both planes)
                            const uint32_t a = ~__brev(ql1), c = ~__brev(ql0), d = ~__brev(qh1), e = ~__brev(qh0);
                            ql0 = a; ql1 = c; qh0 = d; qh1 = e;
                        }
                        m0 = (tl0 ^ ql0) | (th0 ^ qh0);
                        m1 = (tl1 ^ ql1) | (th1 ^ qh1);
                    } else {
                        const OpEntry en = S.tab[wid][idx];
                        const uint64_t tp = ((uint64_t)(uint32_t)(en.tw + 2 * (int32_t)srel) << 5) + (misc & 31u);
                        const uint64_t qlo = ((uint64_t)(uint32_t)(rev ? en.qw - 2 * (int32_t)srel : en.qw + 2 * (int32_t)srel) << 5) + ((misc >> 5) & 31u);
                        for (uint32_t k = 0; k < n; k++) {
                            const unsigned mm = slow_mismatch(P.G, tp + k, rev ? qlo + (WIN_BASES - 1) - k : qlo + k, rev);
                            if (k < 32) m0 |= mm << k; else m1 |= mm << (k - 32);
                        }
                    }
                    // bases beyond the op's end do not exist
                    const uint64_t valid = ~0ull >> (WIN_BASES - n);
                    m0 &= (uint32_t)valid;
                    m1 &= (uint32_t)(valid >> 32);
                    P.masks[s0 + srel] = ((uint64_t)m1 << 32) | m0;
                    if (it == 0 && lane == 0 && !wi0 && !(misc & EF_EXC)) {
                        const OpEntry en = S.tab[wid][idx];
                        const uint64_t tp = ((uint64_t)(uint32_t)en.tw << 5) + (misc & 31u);
                        const uint64_t qtop = ((uint64_t)(uint32_t)en.qw << 5) + ((misc >> 5) & 31u) + (rev ? WIN_BASES - 1 : 0);
                        const uint64_t qprev = rev ? qtop + 1 : qtop - 1;
                        const unsigned kt = (unsigned)((tp - 1) & 31), kq = (unsigned)(qprev & 31);
                        const unsigned tc = ((unsigned)(pm_t >> kt) & 1u) | (((unsigned)(pm_t >> (32 + kt)) & 1u) << 1);
                        unsigned qc = ((unsigned)(pm_q >> kq) & 1u) | (((unsigned)(pm_q >> (32 + kq)) & 1u) << 1);
                        if (rev) qc ^= 3u;
                        if (tc != qc) meta |= MB_PM;
                    }
                    // per-record mismatch tally, flushed when the lane moves to another record
                    if (rec != acc_rec && !(P.debug & 4)) {
                        if (acc_mm) atomicAdd(P.rec_mism + acc_rec, (unsigned long long)acc_mm);
                        acc_rec = rec;
                        acc_mm = 0;
                    }
                    acc_mm += __popc(m0) + __popc(m1);
                }
                // mismatch state of the previous base: top bit of the previous window
                uint32_t pm = __shfl_up_sync(FULL, m1, 1);
                if (lane == 0) pm = it == 0 ? ((meta & MB_PM) ? 0x80000000u : 0u) : carry_m;
                carry_m = __shfl_sync(FULL, m1, 31);
                uint32_t cnt = active ? 1u : 0u;  // pass-through ops are one output each
                if (isM) {
                    const uint64_t valid = ~0ull >> (WIN_BASES - n);
                    uint32_t d0 = (m0 ^ __funnelshift_l(pm, m0, 1)) & (uint32_t)valid;
                    const uint32_t d1 = (m1 ^ __funnelshift_l(m0, m1, 1)) & (uint32_t)(valid >> 32);
                    if (wi0) d0 |= 1u;
                    cnt = __popc(d0) + __popc(d1);
                }
                tile_count += cnt;
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) tile_count += __shfl_xor_sync(FULL, tile_count, o);
        if (lane == 0) P.tile_count[tile] = tile_count;

        // ---- per-record mismatch tallies
        if (!(P.debug & 4)) {
            const uint32_t r0 = __shfl_sync(FULL, acc_rec, 0);
            if (__all_sync(FULL, acc_rec == r0)) {
                const uint32_t sum = __reduce_add_sync(FULL, acc_mm);
                if (lane == 0 && sum && r0 != INF32) atomicAdd(P.rec_mism + r0, (unsigned long long)sum);
            } else if (acc_mm && acc_rec != INF32) {
                atomicAdd(P.rec_mism + acc_rec, (unsigned long long)acc_mm);
            }
        }
    }
}

// ---- emit_kernel ------------------------------------------------------------------------------------
// list entry: bits 0-13 start coordinate inside the tile (64 * window + base), 14-19 gap back to the end of the
// previous op (op-start entries only: 64 - bases in the previous window), 20 X run, 21 pass-through, 22-30 table index
constexpr uint32_t LE_START = 0x3fffu, LE_X = 1u << 20, LE_PASS = 1u << 21;
constexpr int LS_GAP = 14, LS_IDX = 22;
constexpr uint32_t TERM_OPEN = 0xffffffffu;
constexpr int LIST_CAP = 8 * 64 + 32;  // 8 lanes can start 512 runs in one iteration

struct __align__(16) EmitEntry {
    uint64_t opw;   // the op word (copied through when it is not an M op)
    int32_t rem;    // M: op bases left at tile window 0, clamped to 2^30
    uint32_t rec;   // record, top bit: first op of the record
};

template <int WARPS>
struct EmitSmem {
    EmitEntry tab[WARPS][TILE_TAB];
    uint32_t list[WARPS][LIST_CAP];
    uint32_t bm[WARPS][TILE_ITERS];
    uint32_t itc[WARPS][TILE_ITERS];
};

template <int WARPS>
__global__ void __launch_bounds__(32 * WARPS) emit_kernel(EncodeParams P) {
    constexpr int ITERS = TILE_ITERS;
    __shared__ EmitSmem<WARPS> S;
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;

    for (int64_t tile = (int64_t)blockIdx.x * WARPS + wid; tile < P.n_tiles; tile += (int64_t)gridDim.x * WARPS) {
        const uint64_t s0 = (uint64_t)tile * TILE_WIN;
        const uint64_t tile_end = min(s0 + (uint64_t)TILE_WIN, P.n_win);
        const uint32_t n_act = (uint32_t)(tile_end - s0);
        const uint64_t excl = __ldg(P.tile_excl + tile);
        const uint32_t tile_count = __ldg(P.tile_count + tile);

        // the masks of the whole tile: independent loads, issued before anything else
        uint64_t MK[ITERS];
#pragma unroll
        for (int it = 0; it < ITERS; it++) MK[it] = 32u * it + lane < n_act ? __ldg(P.masks + s0 + 32u * it + lane) : 0ull;
        uint32_t pm0 = 0;  // lane 0: top bit of the window before the tile
        if (lane == 0 && s0 > 0) pm0 = (uint32_t)(__ldg(P.masks + s0 - 1) >> 32);

        // ---- op table
        unsigned adj = 0;
        __syncwarp();
        if (lane < ITERS) S.bm[wid][lane] = 0;
        __syncwarp();
        {
            const int64_t cur = P.tile_first_op[tile];
            bool more = true;
            for (int g = 0; more; g++) {
                const int64_t k = cur + 32 * g + lane;
                const bool valid = k < P.n_ops;
                uint64_t wo = INF64, opw = 0;
                uint32_t rec = 0;
                if (valid) { wo = __ldg(P.woff + k); opw = __ldg(P.ops + k); rec = __ldg(P.op_rec + k); }
                const bool in = valid && wo < tile_end;
                if (in) {
                    const int e = 32 * g + (int)lane;
                    int64_t rem = (int64_t)(opw & LEN_MASK) - WIN_BASES * (int64_t)(s0 - wo);
                    if (rem > (1ll << 30)) rem = 1ll << 30;
                    EmitEntry en;
                    en.opw = opw; en.rem = (int32_t)rem; en.rec = rec;
                    S.tab[wid][e] = en;
                    if (wo >= s0) atomicOr(&S.bm[wid][(wo - s0) >> 5], 1u << ((wo - s0) & 31));
                }
                if (g == 0) adj = __shfl_sync(FULL, (unsigned)(wo == s0), 0);
                more = __all_sync(FULL, in);
            }
        }
        __syncwarp();

        // ---- run starts and their ranks inside the tile
        uint32_t D0[ITERS], D1[ITERS], EX[ITERS], META[ITERS];
        uint32_t run = 0, pre = 0, term_local = TERM_OPEN;
        uint32_t carry_m = 0, carry_n = WIN_BASES;
#pragma unroll
        for (int it = 0; it < ITERS; it++) {
            const uint32_t srel = 32u * it + lane;
            const bool active = srel < n_act;
            uint32_t w;
            const uint32_t idx = op_index(S.bm[wid], it, lane, pre, adj, w);
            const uint32_t m0 = (uint32_t)MK[it], m1 = (uint32_t)(MK[it] >> 32);
            uint32_t meta = 0, n = WIN_BASES, d0 = 0, d1 = 0, cnt = 0;
            bool isM = false;
            if (active) {
                const EmitEntry en = S.tab[wid][idx];
                const bool wi0 = (w >> lane) & 1u;
                meta = MB_ACTIVE | (idx << MS_IDX) | (wi0 ? MB_WI0 : 0u);
                if ((int)(en.opw >> 56) != OP_M) {
                    meta |= MB_PASS;
                    cnt = 1;
                } else {
                    isM = true;
                    const int32_t rem = en.rem - WIN_BASES * (int32_t)srel;
                    n = rem < WIN_BASES ? (uint32_t)rem : (uint32_t)WIN_BASES;
                    if (rem <= WIN_BASES) {
                        meta |= MB_LAST;
                        if (srel == n_act - 1) term_local = WIN_BASES * srel + n;
                    }
                }
                if (wi0 && (en.rec >> 31)) meta |= MB_HEAD;
            }
            // neighbours: mismatch state of the previous base and size of the previous window
            const uint32_t mm1 = isM ? m1 : 0u;
            uint32_t pm = __shfl_up_sync(FULL, mm1, 1);
            uint32_t pn = __shfl_up_sync(FULL, n, 1);
            if (lane == 0) {
                pm = it == 0 ? pm0 : carry_m;
                pn = it == 0 ? (uint32_t)WIN_BASES : carry_n;
            }
            carry_m = __shfl_sync(FULL, mm1, 31);
            carry_n = __shfl_sync(FULL, n, 31);
            if (isM) {
                const uint64_t valid = ~0ull >> (WIN_BASES - n);
                d0 = (m0 ^ __funnelshift_l(pm, m0, 1)) & (uint32_t)valid;
                d1 = (m1 ^ __funnelshift_l(m0, m1, 1)) & (uint32_t)(valid >> 32);
                if (meta & MB_WI0) d0 |= 1u;
                cnt = __popc(d0) + __popc(d1);
            }
            if (meta & MB_WI0) meta |= (WIN_BASES - pn) << MS_GAP;
            uint32_t inc = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, inc, o);
                if (lane >= (unsigned)o) inc += t;
            }
            EX[it] = run + inc - cnt;
            run += __shfl_sync(FULL, inc, 31);
            if (lane == 0) S.itc[wid][it] = run;
            if (meta & MB_HEAD) P.rec_out_begin[S.tab[wid][idx].rec & 0x7fffffffu] = (int64_t)(excl + EX[it]);
            D0[it] = d0; D1[it] = d1; META[it] = meta;
        }
        const uint32_t term = __shfl_sync(FULL, term_local, (n_act - 1) & 31);
        if (lane == 0 && tile_count == 0) P.tile_first_closer[tile] = term == TERM_OPEN ? INF64 : WIN_BASES * s0 + term;

        // ---- emission: scatter run starts into the list, then turn it into CigarRecord words (coalesced stores)
        uint32_t *L = S.list[wid];
        uint32_t flushed = 0;  // list entries already written out
        auto flush = [&](uint32_t fill, bool final) {
            __syncwarp();
            const uint32_t n_emit = final ? fill : fill - 1;
            for (uint32_t i = lane; i < n_emit; i += 32) {
                const uint32_t e = L[i];
                const uint64_t slot = excl + flushed + i;
                uint64_t word;
                if (e & LE_PASS) {
                    word = S.tab[wid][(e >> LS_IDX) & 511u].opw;
                } else {
                    uint32_t close = term;
                    if (i + 1 < fill) {
                        const uint32_t nx = L[i + 1];
                        close = (nx & LE_START) - ((nx >> LS_GAP) & 63u);
                    }
                    const uint64_t typ = (uint64_t)((e & LE_X) ? OP_X : OP_EQ) << 56;
                    if (close == TERM_OPEN) {
                        word = typ | (WIN_BASES * s0 + (e & LE_START));  // provisional: absolute start, fixed by stitch_kernel
                        P.tile_open_slot[tile] = (int64_t)slot;
                    } else {
                        word = typ | (uint64_t)(close - (e & LE_START));
                    }
                }
                P.out[slot] = word;
                if (flushed + i == 0) P.tile_first_closer[tile] = WIN_BASES * s0 + (e & LE_START) - ((e >> LS_GAP) & 63u);
            }
            __syncwarp();
            if (!final) {
                const uint32_t last = L[fill - 1];
                __syncwarp();
                if (lane == 0) L[0] = last;
                flushed += fill - 1;
            }
        };
        // one lane's run starts of one window -> list
        auto scatter = [&](int it) {
            const uint32_t meta = META[it];
            if (!(meta & MB_ACTIVE)) return;
            uint32_t r = EX[it] - flushed;
            const uint32_t x0 = WIN_BASES * (32u * it + lane);
            const uint32_t gap = ((meta >> MS_GAP) & 63u) << LS_GAP;
            if (meta & MB_PASS) {
                L[r] = x0 | gap | LE_PASS | (((meta >> MS_IDX) & 511u) << LS_IDX);
                return;
            }
            const uint32_t m0 = (uint32_t)MK[it], m1 = (uint32_t)(MK[it] >> 32);
            uint32_t d = D0[it];
            bool first = true;
            while (d) {
                const uint32_t b = __ffs(d) - 1;
                d &= d - 1;
                L[r++] = (x0 + b) | (first ? gap : 0u) | (((m0 >> b) & 1u) ? LE_X : 0u);
                first = false;
            }
            d = D1[it];
            while (d) {
                const uint32_t b = __ffs(d) - 1;
                d &= d - 1;
                L[r++] = (x0 + 32 + b) | (((m1 >> b) & 1u) ? LE_X : 0u);
            }
        };
#pragma unroll
        for (int it = 0; it < ITERS; it++) {
            const uint32_t before = it ? S.itc[wid][it - 1] : 0u, after = S.itc[wid][it];
            if (after - flushed <= (uint32_t)LIST_CAP) {
                scatter(it);  // the usual case: the whole iteration fits behind what is already listed
            } else {
                // dense run starts: 8 lanes at a time (at most 512 entries), flushing in between
#pragma unroll 1
                for (int g = 0; g < 4; g++) {
                    const uint32_t g_begin = __shfl_sync(FULL, EX[it], 8 * g);
                    const uint32_t g_end = g < 3 ? __shfl_sync(FULL, EX[it], (8 * g + 8) & 31) : after;
                    if (g_end - flushed > (uint32_t)LIST_CAP && g_begin > flushed) flush(g_begin - flushed, false);
                    if ((lane >> 3) == (unsigned)g) scatter(it);
                }
            }
            (void)before;
        }
        if (tile_count > flushed) flush(tile_count - flushed, true);
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

}  // namespace pafgpu
