// encode.cuh - the base-comparison scan (a1 + a7 of SURVEY.md section 8): M ops -> =/X runs.
// Restates paf_encode_mismatches (impl/paf.c:739-784); nothing is copied from it.
//
// Work axis = 64-base windows of all ops of the batch (an M op of length n owns ceil(n/64) windows,
// every other op owns one), cut into warp tiles of 256 windows; a CTA takes the next CTA tile (one warp
// tile per warp) from a ticket counter, so megabase CIGARs spread over the whole grid. ONE kernel, one
// pass over the sequences, no scratch round trip through HBM:
//   A  op table     op lanes load the tile's ops once and leave, per op, everything a window needs in
//                   shared memory (word index + bit phase of target and query, bases left, flags) plus a
//                   bitmap of the windows at which an op starts, so a window finds its op with a popcount.
//   B  compare      striped (window = 32 * iteration + lane, coalesced): every lane issues the 6 x 64-bit
//                   loads of several windows at once, funnel-shifts the bit planes into phase, BREV + NOT
//                   the query on '-' strands, one LOP3 per plane -> one mismatch bit per base. The 64-bit
//                   mask of every window goes to shared memory; mismatches are tallied per record.
//   C  count        blocked (lane owns 8 consecutive windows): run starts per lane, warp scan.
//   D  look-back    the CTA publishes its number of output ops and obtains the number before it with a
//                   decoupled look-back over the tiles ticketed earlier (those are already running, so the
//                   wait is short and cannot deadlock).
//   E  emit         blocked: a lane walks its windows, a run's length is the distance to the next run start
//                   (in-lane; the last run of a lane closes at the first event of a later lane, found with
//                   a suffix-min over lanes). CigarRecord words are staged in shared memory and stored with
//                   coalesced 8-byte stores. A run still open at the end of a warp tile is closed by
//                   stitch_kernel.
#pragma once
#include "kernels.cuh"

namespace pafgpu {

struct EncodeParams {
    Genome G;
    const uint64_t *ops, *woff, *tpos, *qpos;
    const uint32_t *op_rec, *tile_first_op;
    int64_t n_ops;
    uint64_t n_win;
    int64_t n_tiles;                // warp tiles
    int64_t n_cta_tiles;
    uint64_t *out;
    int64_t out_cap;                // words; stores beyond it are dropped (the host grows the buffer and re-runs)
    unsigned long long *status;     // per CTA tile: look-back word (2 flag bits + 62 value bits), pre-set to 0
    unsigned int *ticket;           // pre-set to 0
    unsigned long long *total_out;  // number of output ops of the whole batch
    uint64_t *tile_first_closer;    // absolute window-axis coordinate (64 * window + base) of the warp tile's first run start / op end
    int32_t *tile_open_rank;        // rank inside the warp tile of a run left open, -1 if none (pre-set)
    uint64_t *tile_slot0;           // output slot of the warp tile's first event
    int64_t *rec_out_begin;
    unsigned long long *rec_mism;
};

constexpr int WIN_BASES = 64;
constexpr int TILE_ITERS = 8;                 // windows per lane per warp tile
constexpr int TILE_WIN = 32 * TILE_ITERS;     // windows per warp tile
constexpr int TILE_TAB = TILE_WIN + 2;        // a tile overlaps at most TILE_WIN + 1 ops
constexpr int MASK_PAD = TILE_WIN + TILE_WIN / 8;  // mask of window p sits at p + (p >> 3): conflict-free blocked reads
constexpr int LIST_CAP = 1024;                // events converted per round
constexpr uint32_t COORD_OPEN = 0xffffffffu;
constexpr unsigned long long ST_AGG = 1ull << 62, ST_INCL = 2ull << 62, ST_VAL = (1ull << 62) - 1;

// mismatch bit of one aligned base of an op, through the byte plane (slow, exact)
__device__ __forceinline__ unsigned slow_mismatch(const Genome &G, uint64_t tp, uint64_t qp, bool rev) {
    const uint8_t t = byte_at(G, tp);
    uint8_t q = byte_at(G, qp);
    if (rev) q = complement_upper(q);
    return t != q;
}

// per-op table entry (16 bytes, read with one LDS.128)
//   M op:          tw / qw = word index of the target base / lowest query base under tile window 0,
//                  rem = op bases left at tile window 0 = tile-relative coordinate of the op's end
//   any other op:  tw / qw = low / high half of the op word (copied through)
struct __align__(16) OpEntry {
    int32_t tw, qw, rem;
    uint32_t misc;   // bits 0-4 target phase, 5-9 query phase, then flags
};
constexpr uint32_t EF_PASS = 1u << 11, EF_REV = 1u << 12, EF_EXC = 1u << 13;
constexpr int32_t REM_CLAMP = 1 << 30;

struct WarpSmem {
    OpEntry tab[TILE_TAB];
    uint32_t rec[TILE_TAB];   // owning record, top bit: first op of the record
    union {
        uint64_t mask[MASK_PAD];
        uint32_t list[LIST_CAP + 2];   // 32-bit event entries of one round (+ successor of the last)
    } u;
    uint32_t bm[TILE_ITERS];  // bit p: an op starts at tile window p
};

// ---- A: op table of one warp tile. Returns adj (1 when table entry 0 starts exactly at tile window 0) and
// whether any op of the tile touches a non-ACGT word.
__device__ __forceinline__ void build_table(const EncodeParams &P, WarpSmem &S, int64_t tile, uint64_t s0, uint64_t tile_end,
                                            unsigned lane, unsigned &adj, bool &any_exc) {
    __syncwarp();
    if (lane < TILE_ITERS) S.bm[lane] = 0;
    __syncwarp();
    const int64_t cur = P.tile_first_op[tile];
    bool more = true, exc = false;
    adj = 0;
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
            OpEntry en;
            if ((int)(opw >> 56) != OP_M) {
                en.tw = (int32_t)(uint32_t)opw; en.qw = (int32_t)(uint32_t)(opw >> 32); en.rem = 0; en.misc = EF_PASS;
            } else {
                const int64_t back = WIN_BASES * (int64_t)(s0 - wo);  // op bases before tile window 0 (negative: op starts inside)
                const bool rev = (qpw & QFLAG_REV) != 0;
                const int64_t tb = (int64_t)(tpw & POS_MASK) + back;
                // lowest query base of tile window 0's span: '+' walks up from it, '-' starts 63 below the top base
                const int64_t qb = rev ? (int64_t)(qpw & POS_MASK) - back - (WIN_BASES - 1) : (int64_t)(qpw & POS_MASK) + back;
                int64_t rem = (int64_t)(opw & LEN_MASK) - back;
                if (rem > REM_CLAMP) rem = REM_CLAMP;
                en.tw = (int32_t)(tb >> 5);
                en.qw = (int32_t)(qb >> 5);
                en.rem = (int32_t)rem;
                en.misc = (uint32_t)(tb & 31) | ((uint32_t)(qb & 31) << 5) | (rev ? EF_REV : 0u) | ((qpw & QFLAG_EXC) ? EF_EXC : 0u);
                exc |= (qpw & QFLAG_EXC) != 0;
            }
            S.tab[e] = en;
            S.rec[e] = rec;
            if (wo >= s0) atomicOr(&S.bm[(wo - s0) >> 5], 1u << ((wo - s0) & 31));
        }
        if (g == 0) adj = __shfl_sync(FULL, (unsigned)(wo == s0), 0);
        more = __all_sync(FULL, in);
    }
    any_exc = __any_sync(FULL, exc);
    __syncwarp();
}

// ---- B: compare. Striped windows; masks to shared memory; per-record mismatch tallies. Returns (lane 0) the
// mismatch state of the base just before the tile when tile window 0 continues an op.
template <int BATCH, bool EXC>
__device__ __forceinline__ uint32_t compare_tile(const EncodeParams &P, WarpSmem &S, uint32_t n_act, unsigned lane, unsigned adj) {
    static_assert(TILE_ITERS % BATCH == 0, "BATCH must divide the iterations of a tile");
    const uint64_t *__restrict__ gw = P.G.words;
    uint32_t pre = 0, pm0 = 0;
    uint32_t acc_rec = INF32, acc_mm = 0;
#pragma unroll
    for (int hb = 0; hb < TILE_ITERS; hb += BATCH) {
        // issue every sequence load of the batch (inactive and pass-through windows read arena word 0: no branches)
        uint32_t IDX[BATCH];
        uint64_t T0[BATCH], T1[BATCH], T2[BATCH], Q0[BATCH], Q1[BATCH], Q2[BATCH];
        uint64_t pm_t = 0, pm_q = 0;  // lane 0, first window only: words holding the base before the window
        uint32_t pm_k = 0;
#pragma unroll
        for (int b = 0; b < BATCH; b++) {
            const int it = hb + b;
            const uint32_t srel = 32u * it + lane;
            const uint32_t w = S.bm[it];
            uint32_t idx = pre + __popc(w & (0xffffffffu >> (31 - lane))) - adj;
            pre += __popc(w);
            if (srel >= n_act) idx = 0;
            IDX[b] = idx;
            const OpEntry en = S.tab[idx];
            const bool isM = srel < n_act && !(en.misc & EF_PASS);
            const bool rev = (en.misc & EF_REV) != 0;
            const int32_t tw = isM ? en.tw + 2 * (int32_t)srel : 0;
            const int32_t qw = isM ? (rev ? en.qw - 2 * (int32_t)srel : en.qw + 2 * (int32_t)srel) : 0;
            const uint64_t *tpn = gw + tw, *qpn = gw + qw;
            T0[b] = ldg64(tpn); T1[b] = ldg64(tpn + 1); T2[b] = ldg64(tpn + 2);
            Q0[b] = ldg64(qpn); Q1[b] = ldg64(qpn + 1); Q2[b] = ldg64(qpn + 2);
            if (it == 0 && lane == 0 && isM && !(w & 1u)) {
                // tile window 0 continues an op: mismatch state of the base just before the tile
                const uint64_t tp = ((uint64_t)(uint32_t)tw << 5) + (en.misc & 31u);
                const uint64_t qtop = ((uint64_t)(uint32_t)qw << 5) + ((en.misc >> 5) & 31u) + (rev ? WIN_BASES - 1 : 0);
                const uint64_t qprev = rev ? qtop + 1 : qtop - 1;
                if (EXC && (en.misc & EF_EXC)) {
                    pm0 = slow_mismatch(P.G, tp - 1, qprev, rev) ? 0x80000000u : 0u;
                } else {
                    pm_t = ldg64(gw + ((tp - 1) >> 5));
                    pm_q = ldg64(gw + (qprev >> 5));
                    pm_k = (uint32_t)((tp - 1) & 31) | ((uint32_t)(qprev & 31) << 5) | (rev ? 1u << 10 : 0u) | (1u << 11);
                }
            }
        }
        if (hb == 0 && (pm_k >> 11)) {  // lane 0 only; consumed after the batch's loads were issued
            const unsigned kt = pm_k & 31u, kq = (pm_k >> 5) & 31u;
            const unsigned tc = ((unsigned)(pm_t >> kt) & 1u) | (((unsigned)(pm_t >> (32 + kt)) & 1u) << 1);
            unsigned qc = ((unsigned)(pm_q >> kq) & 1u) | (((unsigned)(pm_q >> (32 + kq)) & 1u) << 1);
            if (pm_k & (1u << 10)) qc ^= 3u;
            pm0 = tc != qc ? 0x80000000u : 0u;
        }
        // mismatch bits
#pragma unroll
        for (int b = 0; b < BATCH; b++) {
            const int it = hb + b;
            const uint32_t srel = 32u * it + lane;
            const uint32_t idx = IDX[b];
            const OpEntry en = S.tab[idx];
            const bool isM = srel < n_act && !(en.misc & EF_PASS);
            const bool rev = (en.misc & EF_REV) != 0;
            const int32_t rem = en.rem - WIN_BASES * (int32_t)srel;
            const uint32_t n = !isM ? 0u : rem < WIN_BASES ? (uint32_t)rem : (uint32_t)WIN_BASES;
            const uint32_t misc = en.misc;
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
            uint32_t m0 = (tl0 ^ ql0) | (th0 ^ qh0);
            uint32_t m1 = (tl1 ^ ql1) | (th1 ^ qh1);
            if (EXC && isM && (misc & EF_EXC)) {
                const uint64_t tp = ((uint64_t)(uint32_t)(en.tw + 2 * (int32_t)srel) << 5) + sht;
                const uint64_t qlo = ((uint64_t)(uint32_t)(rev ? en.qw - 2 * (int32_t)srel : en.qw + 2 * (int32_t)srel) << 5) + shq;
                const uint64_t ql = qlo + (rev ? WIN_BASES - n : 0u);
                if (span_has_exception(P.G, tp, tp + n - 1) || span_has_exception(P.G, ql, ql + n - 1)) {
                    m0 = m1 = 0;
                    for (uint32_t k = 0; k < n; k++) {
                        const unsigned mm = slow_mismatch(P.G, tp + k, rev ? qlo + (WIN_BASES - 1) - k : qlo + k, rev);
                        if (k < 32) m0 |= mm << k; else m1 |= mm << (k - 32);
                    }
                }
            }
            // bases beyond the op's end do not exist
            const uint64_t valid = n ? ~0ull >> (WIN_BASES - n) : 0ull;
            m0 &= (uint32_t)valid;
            m1 &= (uint32_t)(valid >> 32);
            S.u.mask[srel + (srel >> 3)] = ((uint64_t)m1 << 32) | m0;
            // per-record mismatch tally, flushed when the lane moves to another record
            if (isM) {
                const uint32_t rec = S.rec[idx] & 0x7fffffffu;
                if (rec != acc_rec) {
                    if (acc_mm) atomicAdd(P.rec_mism + acc_rec, (unsigned long long)acc_mm);
                    acc_rec = rec;
                    acc_mm = 0;
                }
                acc_mm += __popc(m0) + __popc(m1);
            }
        }
    }
    // flush the tallies: one atomic per warp when the whole tile is one record
    const uint32_t r0 = __shfl_sync(FULL, acc_rec, 0);
    if (__all_sync(FULL, acc_rec == r0)) {
        const uint32_t sum = __reduce_add_sync(FULL, acc_mm);
        if (lane == 0 && sum && r0 != INF32) atomicAdd(P.rec_mism + r0, (unsigned long long)sum);
    } else if (acc_mm && acc_rec != INF32) {
        atomicAdd(P.rec_mism + acc_rec, (unsigned long long)acc_mm);
    }
    return pm0;
}

// ---- C / E: blocked walk. Lane owns tile windows 8 * lane .. 8 * lane + 7. An "event" is an output op: the
// start of an =/X run or a pass-through op. Coordinates are tile-relative (64 * window + base).
// Events are first listed as 32-bit entries in tile order:
//   bits 0-13  run start coordinate (pass-through op: its table index)
//   bits 14-28 coordinate at which the run before this event closes: the event's own coordinate inside an op,
//              the previous op's end at an op start
//   bit 29 X run, bit 30 pass-through op
// so that entry i's length is close(entry i + 1) - start(entry i), computed by any lane.
constexpr uint32_t LE_START = 0x3fffu, LE_CLOSE = 0x7fffu, LE_X = 1u << 29, LE_PASS = 1u << 30;
constexpr int LS_CLOSE = 14;

struct LaneWalk {
    uint64_t MK[TILE_ITERS];  // masks of the lane's windows
    uint32_t byte;            // op-start bits of the lane's windows
    int cur0;                 // table index of the op owning the window before the lane's first
    uint32_t pm;              // top bit: mismatch state of the base before the lane's first window
    uint32_t n_win;           // active windows of the lane
};

// run-start bits of one M window (n >= 1 bases); the op-start bit is handled by the caller
__device__ __forceinline__ void run_starts(uint32_t m0, uint32_t m1, uint32_t pm, uint32_t n, uint32_t &d0, uint32_t &d1) {
    d0 = m0 ^ ((m0 << 1) | (pm >> 31));
    d1 = m1 ^ __funnelshift_l(m0, m1, 1);
    const uint32_t v0 = n >= 32 ? 0xffffffffu : (1u << n) - 1u;
    const uint32_t v1 = n > 32 ? 0xffffffffu >> (64 - n) : 0u;
    d0 &= v0;
    d1 &= v1;
}

// C: number of events of the lane; end_closer (meaningful on the lane owning the tile's last active window): end
// of that window's M op when it ends inside the tile, else COORD_OPEN
__device__ __forceinline__ void lane_count(const WarpSmem &S, const LaneWalk &W, unsigned lane, uint32_t &cnt, uint32_t &end_closer) {
    cnt = 0;
    int cur = W.cur0;
    OpEntry en = S.tab[cur < 0 ? 0 : cur];
    uint32_t pm = W.pm;
#pragma unroll
    for (int j = 0; j < TILE_ITERS; j++) {
        if ((uint32_t)j < W.n_win) {
            const uint32_t p = 8u * lane + j;
            const bool start = (W.byte >> j) & 1u;
            if (start) en = S.tab[++cur];
            if (en.misc & EF_PASS) {
                cnt++;
                pm = 0;
            } else {
                const int32_t rem = en.rem - WIN_BASES * (int32_t)p;
                const uint32_t n = rem < WIN_BASES ? (uint32_t)rem : (uint32_t)WIN_BASES;
                const uint32_t m0 = (uint32_t)W.MK[j], m1 = (uint32_t)(W.MK[j] >> 32);
                uint32_t d0, d1;
                run_starts(m0, m1, pm, n, d0, d1);
                if (start) d0 |= 1u;
                cnt += __popc(d0) + __popc(d1);
                pm = m1;
            }
        }
    }
    end_closer = (!(en.misc & EF_PASS) && en.rem <= (int32_t)(WIN_BASES * (8u * lane + W.n_win))) ? (uint32_t)en.rem : COORD_OPEN;
}

// E1: list the lane's events, entries base .. base + LIST_CAP of the tile (one more than a round converts, so the
// last converted entry sees its successor). ex = rank of the lane's first event inside the tile. Also leaves the
// tile-relative output rank of every record's first op in rec_out_begin (rec_begin_kernel adds the tile's slot).
__device__ __forceinline__ void lane_scatter(const EncodeParams &P, WarpSmem &S, const LaneWalk &W, unsigned lane, uint32_t ex, uint32_t base) {
    uint32_t *L = S.u.list;
    uint32_t r = ex - base;  // wraps for events before the round; comes back to 0 exactly at entry `base`
    int cur = W.cur0;
    OpEntry en = S.tab[cur < 0 ? 0 : cur];
    uint32_t pm = W.pm;
#pragma unroll
    for (int j = 0; j < TILE_ITERS; j++) {
        if ((uint32_t)j < W.n_win) {
            const uint32_t p = 8u * lane + j;
            const uint32_t x0 = WIN_BASES * p;
            const bool start = (W.byte >> j) & 1u;
            uint32_t close0 = x0;
            if (start) {
                // the previous op ends here: an M op at its end coordinate, a pass-through op at the window boundary
                if (cur >= 0 && !(en.misc & EF_PASS)) close0 = (uint32_t)en.rem;
                en = S.tab[++cur];
                const uint32_t recw = S.rec[cur];
                if ((recw >> 31) && r <= (uint32_t)LIST_CAP) P.rec_out_begin[recw & 0x7fffffffu] = (int64_t)(r + base);
            }
            if (en.misc & EF_PASS) {
                if (r <= (uint32_t)LIST_CAP) L[r] = (uint32_t)cur | (close0 << LS_CLOSE) | LE_PASS;
                r++;
                pm = 0;
            } else {
                const int32_t rem = en.rem - WIN_BASES * (int32_t)p;
                const uint32_t n = rem < WIN_BASES ? (uint32_t)rem : (uint32_t)WIN_BASES;
                const uint32_t m0 = (uint32_t)W.MK[j], m1 = (uint32_t)(W.MK[j] >> 32);
                uint32_t d0, d1;
                run_starts(m0, m1, pm, n, d0, d1);
                if (start) {
                    if (r <= (uint32_t)LIST_CAP) L[r] = x0 | (close0 << LS_CLOSE) | ((m0 & 1u) ? LE_X : 0u);
                    r++;
                    d0 &= ~1u;
                }
                while (d0) {
                    const uint32_t b = (uint32_t)__ffs((int)d0) - 1u;
                    d0 &= d0 - 1u;
                    if (r <= (uint32_t)LIST_CAP) L[r] = (x0 + b) * ((1u << LS_CLOSE) + 1u) | (((m0 >> b) & 1u) ? LE_X : 0u);
                    r++;
                }
                while (d1) {
                    const uint32_t b = (uint32_t)__ffs((int)d1) - 1u;
                    d1 &= d1 - 1u;
                    if (r <= (uint32_t)LIST_CAP) L[r] = (x0 + 32u + b) * ((1u << LS_CLOSE) + 1u) | (((m1 >> b) & 1u) ? LE_X : 0u);
                    r++;
                }
                pm = m1;
            }
        }
    }
}

template <int WARPS, int BATCH>
__global__ void __launch_bounds__(32 * WARPS) encode_kernel(EncodeParams P) {
    __shared__ WarpSmem SM[WARPS];
    __shared__ uint32_t s_wcount[WARPS];
    __shared__ unsigned long long s_excl;
    __shared__ int64_t s_ticket;
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    WarpSmem &S = SM[wid];

    for (;;) {
        if (threadIdx.x == 0) s_ticket = (int64_t)atomicAdd(P.ticket, 1u);
        __syncthreads();
        const int64_t ct = s_ticket;
        if (ct >= P.n_cta_tiles) break;
        const int64_t tile = ct * WARPS + wid;
        const uint64_t s0 = (uint64_t)tile * TILE_WIN;
        const uint32_t n_act = s0 < P.n_win ? (uint32_t)min((uint64_t)TILE_WIN, P.n_win - s0) : 0u;

        LaneWalk W;
        uint32_t cnt = 0, end_closer = COORD_OPEN, inc = 0;
        if (n_act) {
            unsigned adj;
            bool any_exc;
            build_table(P, S, tile, s0, s0 + n_act, lane, adj, any_exc);
            uint32_t pm0 = any_exc ? compare_tile<BATCH, true>(P, S, n_act, lane, adj) : compare_tile<BATCH, false>(P, S, n_act, lane, adj);
            __syncwarp();
            // blocked view of the tile
            W.byte = (S.bm[lane >> 2] >> (8 * (lane & 3))) & 0xffu;
            W.n_win = 8u * lane >= n_act ? 0u : min(8u, n_act - 8u * lane);
#pragma unroll
            for (int j = 0; j < TILE_ITERS; j++) W.MK[j] = S.u.mask[9 * lane + j];
            uint32_t oinc = __popc(W.byte);
            const uint32_t oc = oinc;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, oinc, o);
                if (lane >= (unsigned)o) oinc += t;
            }
            W.cur0 = (int)(oinc - oc) - (int)adj;
            const uint32_t up = __shfl_up_sync(FULL, (uint32_t)(W.MK[TILE_ITERS - 1] >> 32), 1);
            W.pm = lane ? up : pm0;
            lane_count(S, W, lane, cnt, end_closer);
            inc = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, inc, o);
                if (lane >= (unsigned)o) inc += t;
            }
        }
        const uint32_t tile_count = __shfl_sync(FULL, inc, 31);
        const uint32_t ex = inc - cnt;
        if (lane == 0) s_wcount[wid] = tile_count;
        __syncthreads();

        // ---- D (first half): publish the CTA tile's number of events, then list round 0 before anybody waits
        volatile unsigned long long *st = P.status;
        unsigned long long agg = 0;
        if (wid == 0) {
#pragma unroll
            for (int w = 0; w < WARPS; w++) agg += s_wcount[w];
            if (lane == 0) st[ct] = (ct == 0 ? ST_INCL : ST_AGG) | agg;
        }
        const uint32_t term = n_act ? __shfl_sync(FULL, end_closer, (n_act - 1) >> 3) : COORD_OPEN;
        if (n_act) {
            __syncwarp();  // every lane holds its masks in registers: the list may overwrite them
            if (cnt && ex <= (uint32_t)LIST_CAP) lane_scatter(P, S, W, lane, ex, 0u);
            if (lane == 0 && tile_count <= (uint32_t)LIST_CAP) S.u.list[tile_count] = (term & LE_CLOSE) << LS_CLOSE;
        }
        // ---- D (second half): decoupled look-back over CTA tiles (warp 0)
        if (wid == 0) {
            unsigned long long excl = 0;
            if (ct > 0) {
                int64_t j = ct - 1 - (int64_t)lane;
                for (;;) {
                    unsigned long long v = j >= 0 ? st[j] : (2ull << 62);
                    while (__any_sync(FULL, (v >> 62) == 0)) {
                        __nanosleep(64);
                        v = j >= 0 ? st[j] : (2ull << 62);
                    }
                    const unsigned incl = __ballot_sync(FULL, (v >> 62) == 2);
                    const unsigned upto = incl ? (unsigned)__ffs(incl) - 1u : 31u;  // nearest tile holding an inclusive prefix
                    unsigned long long c = lane <= upto ? (v & ST_VAL) : 0ull;
#pragma unroll
                    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(FULL, c, o);
                    excl += c;
                    if (incl) break;
                    j -= 32;
                }
                if (lane == 0) st[ct] = ST_INCL | (excl + agg);
            }
            if (lane == 0) {
                s_excl = excl;
                if (ct == P.n_cta_tiles - 1) *P.total_out = excl + agg;
            }
        }
        __syncthreads();

        // ---- E2: entries -> CigarRecord words, coalesced stores
        if (n_act) {
            unsigned long long slot0 = s_excl;
#pragma unroll
            for (int w = 0; w < WARPS; w++) slot0 += (unsigned)w < wid ? s_wcount[w] : 0u;
            if (lane == 0) {
                P.tile_slot0[tile] = slot0;
                P.tile_first_closer[tile] = tile_count ? WIN_BASES * s0 + ((S.u.list[0] >> LS_CLOSE) & LE_CLOSE)
                                                       : (term == COORD_OPEN ? INF64 : WIN_BASES * s0 + term);
            }
            for (uint32_t base = 0; base < tile_count; base += LIST_CAP) {
                if (base) {  // dense tiles: further rounds
                    __syncwarp();
                    if (cnt && ex <= base + LIST_CAP && ex + cnt > base) lane_scatter(P, S, W, lane, ex, base);
                    if (lane == 0 && tile_count - base <= (uint32_t)LIST_CAP) S.u.list[tile_count - base] = (term & LE_CLOSE) << LS_CLOSE;
                }
                __syncwarp();
                const uint32_t n_round = min((uint32_t)LIST_CAP, tile_count - base);
                for (uint32_t i = lane; i < n_round; i += 32) {
                    const uint32_t e = S.u.list[i], nx = S.u.list[i + 1];
                    uint32_t lo, hi;
                    if (e & LE_PASS) {
                        const OpEntry en = S.tab[e & 0x1ffu];
                        lo = (uint32_t)en.tw; hi = (uint32_t)en.qw;
                    } else {
                        const uint32_t start = e & LE_START;
                        hi = (uint32_t)((e & LE_X) ? OP_X : OP_EQ) << 24;
                        lo = ((nx >> LS_CLOSE) & LE_CLOSE) - start;
                        if (base + i == tile_count - 1 && term == COORD_OPEN) {
                            // run still open at the tile's end: provisional absolute start, turned into a length by stitch_kernel
                            const uint64_t abs_start = WIN_BASES * s0 + start;
                            lo = (uint32_t)abs_start; hi |= (uint32_t)(abs_start >> 32);
                            P.tile_open_rank[tile] = (int32_t)(base + i);
                        }
                    }
                    const unsigned long long slot = slot0 + base + i;
                    if (slot < (unsigned long long)P.out_cap) P.out[slot] = ((uint64_t)hi << 32) | lo;
                }
            }
        }
    }
}

// A run left open at the end of warp tile t closes at the first run start / op end of a later tile.
__global__ void stitch_kernel(int64_t n_tiles, const int32_t *tile_open_rank, const uint64_t *tile_slot0, const uint64_t *tile_first_closer,
                              uint64_t *out, int64_t out_cap) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tiles) return;
    const int32_t rank = tile_open_rank[t];
    if (rank < 0) return;
    const int64_t slot = (int64_t)tile_slot0[t] + rank;
    if (slot >= out_cap) return;
    int64_t u = t + 1;
    while (u < n_tiles && tile_first_closer[u] == INF64) u++;
    if (u >= n_tiles) return;  // cannot happen: the op's end is always a closer
    const uint64_t rec = out[slot];
    out[slot] = (tile_first_closer[u] - (rec & LEN_MASK)) | (rec & ~LEN_MASK);
}

// rec_out_begin[r] holds the rank of the record's first output op inside its warp tile: add the tile's first slot.
__global__ void rec_begin_kernel(int64_t n_records, const int64_t *op_begin, const uint64_t *woff, const uint64_t *tile_slot0,
                                 int64_t *rec_out_begin) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_records) return;
    const uint64_t tile = woff[op_begin[r]] / TILE_WIN;
    rec_out_begin[r] += (int64_t)tile_slot0[tile];
}

}  // namespace pafgpu
