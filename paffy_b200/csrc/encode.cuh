// encode.cuh - the base-comparison scan (a1 + a7 of SURVEY.md section 8): M ops -> =/X runs.
// Restates paf_encode_mismatches (impl/paf.c:739-784); nothing is copied from it.
//
// Work axis = windows of WIN_BASES bases (128 by default: 4 bit-sliced words) of all ops of the batch (an M op of length
// n owns ceil(n / WIN_BASES) windows, every other op owns one), cut into warp tiles of 16 384 window-axis bases (512
// words: 16 per lane); a CTA takes the next CTA tile (one warp
// tile per warp) from a ticket counter, so megabase CIGARs spread over the whole grid. ONE kernel, one
// pass over the sequences, no scratch round trip through HBM:
//   A  op table     op lanes load the tile's ops once and leave, per op, everything a window needs in
//                   shared memory (word index + bit phase of target and query, bases left, flags) plus a
//                   bitmap of the windows at which an op starts, so a window finds its op with a popcount.
//   B  compare      striped (window = 32 * iteration + lane, coalesced): every lane issues the WIN_WORDS + 1 64-bit
//                   loads per stream of a window (or of several windows) at once, funnel-shifts the bit planes into phase ('-' strand
//                   queries are read from the reverse-complement copy of the contig), one LOP3 per plane -> one mismatch bit per base. The 64-bit
//                   mask of every window goes to shared memory; mismatches are tallied per record.
//   C  count        blocked (lane owns 16 consecutive words): run starts per lane, warp scan.
//   D  look-back    the CTA publishes its number of output ops and obtains the number before it with a
//                   decoupled look-back over the tiles ticketed earlier (those are already running, so the
//                   wait is short and cannot deadlock).
//   E  emit         the tile's events are dealt to the lanes in equal shares by rank; a lane walks its share
//                   last to first, a run's length is the distance to the event handled one trip earlier, and the
//                   32-bit list entry in shared memory is the output word squeezed (code + length). The op pass
//                   shortens the run before every op start and writes pass-through ops in; the list then goes
//                   out as CigarRecord words with coalesced 8-byte stores. A run still open at the end of a warp
//                   tile is closed by stitch_kernel.
#pragma once
#include "kernels.cuh"

#ifndef ENC_NO_PREFETCH
#define ENC_NO_PREFETCH 0
#endif
#ifndef ENC_SHFL_T2
#define ENC_SHFL_T2 0  // 1: the third word of a window comes from the next lane's first word by shuffle when both windows
                       // belong to the same op (4 instead of 6 sequence loads per window on the common path)
#endif
#ifndef ENC_LDG128
#define ENC_LDG128 1   // 1: the WIN_WORDS + 1 words of a window come from aligned 128-bit loads (the pair holding the first word
                       // and the pairs behind it), the word parity of the first one picked with selects: fewer, wider L1 requests
#endif
#ifndef ENC_ROLL_HB
#define ENC_ROLL_HB 0  // 1: the batch loop of the compare phase is not unrolled (smaller code)
#endif

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
    int64_t *tile_open_slot;        // output slot of a run left open, -1 if none (pre-set)
    int64_t *rec_out_begin;
    unsigned long long *rec_mism;
    struct OpEntry *tab_overflow;   // TILE_TAB entries per resident warp: table entries TAB_SMEM.. of tiles with many short ops
    unsigned long long *cta_span;   // per CTA: {ns from its first tile to its last, SM id}: the load-balance figure of SURVEY 8d
};

#ifndef ENC_WIN_WORDS
#define ENC_WIN_WORDS 4
#endif
constexpr int WIN_WORDS = ENC_WIN_WORDS;      // 32-base words per window (2 or 4)
constexpr int WIN_BASES = 32 * WIN_WORDS;
constexpr int LANE_WORDS = 16;                // words a lane owns in the blocked view (512 bases)
constexpr int TILE_ITERS = LANE_WORDS / WIN_WORDS;  // windows per lane per warp tile
constexpr int TILE_WIN = 32 * TILE_ITERS;     // windows per warp tile
constexpr int TILE_WORDS = 32 * LANE_WORDS;   // 32-base words per warp tile
static_assert(WIN_WORDS == 2 || WIN_WORDS == 4, "windows of 64 or 128 bases");
constexpr int TILE_TAB = TILE_WIN + 2;        // a tile overlaps at most TILE_WIN + 1 ops
#ifndef ENC_TAB_SMEM
#define ENC_TAB_SMEM 80
#endif
#ifndef ENC_LIST_CAP
#define ENC_LIST_CAP 384
#endif
constexpr int TAB_SMEM = ENC_TAB_SMEM;                 // table entries kept in shared memory; denser tiles spill the rest to HBM
constexpr int MASK_WORDS = 2 + TILE_WORDS + 2 * (TILE_WORDS / 32);  // mismatch bits of word k sit at 32-bit index 2 + k + 2 * (k >> 5):
                                                     // striped 8-byte stores and blocked 8-byte loads (lane stride 64 bytes, one more
                                                     // 8-byte slot every other lane) are both free of bank conflicts
constexpr int LIST_CAP = ENC_LIST_CAP;                 // events converted per round
constexpr uint32_t COORD_OPEN = 0xffffffffu;
constexpr unsigned long long ST_AGG = 1ull << 62, ST_INCL = 2ull << 62, ST_VAL = (1ull << 62) - 1;

// mismatch bit of one aligned base of an op, through the byte plane (slow, exact)
__device__ __forceinline__ unsigned slow_mismatch(const Genome &G, uint64_t tp, uint64_t qp) {
    return byte_at(G, tp) != byte_at(G, qp);  // '-' strand queries live reverse-complemented in the arena
}

// per-op table entry (16 bytes, read with one LDS.128)
//   M op:          tw / qw = word index of the target base / lowest query base under tile window 0,
//                  rem (low 16 bits) = op bases left at tile window 0 = tile-relative coordinate of the op's end,
//                  clamped to 0xffff (anything beyond the tile's 16 384 bases only has to stay "beyond")
//   every op:      rem (high 16 bits) = owning record, relative to the record of table entry 0
//   any other op:  tw / qw = low / high half of the op word (copied through)
struct __align__(16) OpEntry {
    int32_t tw, qw, rem;
    uint32_t misc;   // bits 0-4 target phase, 5-9 query phase, then EF_* flags and ES_* fields
};
constexpr uint32_t EF_PASS = 1u << 11, EF_EXC = 1u << 13;
// ops that start inside the tile: bits 14-21 the tile window they start at, 22-28 bases between the previous M op's
// end and that window boundary (WIN_BASES - bases in its last window), 29 first op of its record
constexpr int ES_WIN = 14, ES_GAP = 22;
constexpr uint32_t EF_HEAD = 1u << 29;
constexpr int32_t REM_CLAMP = 0xffff;
__device__ __forceinline__ int32_t entry_rem(const struct OpEntry &en);
__device__ __forceinline__ uint32_t entry_rec(const struct OpEntry &en);

__device__ __forceinline__ int32_t entry_rem(const OpEntry &en) { return en.rem & 0xffff; }
__device__ __forceinline__ uint32_t entry_rec(const OpEntry &en) { return (uint32_t)en.rem >> 16; }

struct WarpSmem {
    OpEntry tab[TAB_SMEM];
    alignas(8) uint32_t mask[MASK_WORDS];  // mismatch bits of word k at 2 + k + 2 * (k >> 5); the word before a lane's 16
                                      // words is the word before them (a pad word filled in for even lanes)
    uint32_t list[LIST_CAP + 1];      // events of one round (+ sink), see LE_*
    alignas(16) uint16_t prel[TILE_WIN];  // events of the owning lane before window p
    uint16_t lanex[32];               // events of the tile before lane l
    uint32_t bm[TILE_ITERS];          // bit p: an op starts at tile window p
    uint32_t bmpre[TILE_ITERS];       // op starts before word k of bm
    uint2 nf[TILE_WORDS / 32];        // per 32 words: .x those holding events, .y those an op starts at
};

// table entry i of the warp's tile: shared memory, or the warp's overflow area in HBM for the rare tile with more
// than TAB_SMEM ops (written and read by the same warp, L1 bypassed)
struct TileTab {
    WarpSmem &S;
    OpEntry *otab;
    __device__ __forceinline__ OpEntry get(uint32_t i) const {
        if (i < (uint32_t)TAB_SMEM) return S.tab[i];
        const int4 v = __ldcg(reinterpret_cast<const int4 *>(otab + i));
        OpEntry en;
        en.tw = v.x; en.qw = v.y; en.rem = v.z; en.misc = (uint32_t)v.w;
        return en;
    }
    __device__ __forceinline__ void put(uint32_t i, const OpEntry &en) const {
        if (i < (uint32_t)TAB_SMEM) S.tab[i] = en;
        else otab[i] = en;
    }
};

// ---- A: op table of one warp tile. adj = 1 when table entry 0 starts exactly at tile window 0; n_tab = entries;
// rec_first = record of entry 0; any_exc = some op of the tile touches a non-ACGT word.
__device__ __forceinline__ void build_table(const EncodeParams &P, WarpSmem &S, const TileTab &T, int64_t cur, uint64_t s0, uint64_t tile_end,
                                            unsigned lane, unsigned &adj, uint32_t &n_tab, uint32_t &rec_first, bool &any_exc) {
    __syncwarp();
    if (lane < TILE_ITERS) S.bm[lane] = 0;
    __syncwarp();
    bool more = true, exc = false;
    adj = 0;
    n_tab = 0;
    rec_first = 0;
    uint32_t carry_end = 0;  // end coordinate of the last op of the previous group if it is an M op, else 0
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
        if (g == 0) {
            adj = __shfl_sync(FULL, (unsigned)(wo == s0), 0);
            rec_first = __shfl_sync(FULL, rec, 0) & 0x7fffffffu;
        }
        OpEntry en;
        en.tw = en.qw = en.rem = 0;
        en.misc = EF_PASS;
        uint32_t my_end = 0;
        if (in) {
            if ((int)(opw >> 56) != OP_M) {
                en.tw = (int32_t)(uint32_t)opw; en.qw = (int32_t)(uint32_t)(opw >> 32);
            } else {
                const int64_t back = WIN_BASES * (int64_t)(s0 - wo);  // op bases before tile window 0 (negative: op starts inside)
                const int64_t tb = (int64_t)(tpw & POS_MASK) + back;
                const int64_t qb = (int64_t)(qpw & POS_MASK) + back;
                int64_t rem = (int64_t)(opw & LEN_MASK) - back;
                if (rem > REM_CLAMP) rem = REM_CLAMP;
                en.tw = (int32_t)(tb >> 5);
                en.qw = (int32_t)(qb >> 5);
                en.rem = (int32_t)rem;
                en.misc = (uint32_t)(tb & 31) | ((uint32_t)(qb & 31) << 5) | ((qpw & QFLAG_EXC) ? EF_EXC : 0u);
                exc |= (qpw & QFLAG_EXC) != 0;
                my_end = (uint32_t)rem;
            }
        }
        // end of the op before mine (0: none in the table, or a pass-through op)
        uint32_t prev_end = __shfl_up_sync(FULL, my_end, 1);
        if (lane == 0) prev_end = carry_end;
        carry_end = __shfl_sync(FULL, my_end, 31);
        if (in) {
            if (wo >= s0) {
                const uint32_t w = (uint32_t)(wo - s0);
                const uint32_t gap = prev_end ? WIN_BASES * w - prev_end : 0u;
                en.misc |= (w << ES_WIN) | ((gap & (uint32_t)(WIN_BASES - 1)) << ES_GAP) | ((rec >> 31) ? EF_HEAD : 0u);
                atomicOr(&S.bm[w >> 5], 1u << (w & 31));
            }
            en.rem |= (int32_t)(((rec & 0x7fffffffu) - rec_first) << 16);
            T.put(32u * g + lane, en);
        }
        const unsigned inb = __ballot_sync(FULL, in);
        n_tab += __popc(inb);
        more = inb == FULL;
    }
    any_exc = __any_sync(FULL, exc);
    __syncwarp();
    {
        // op starts before word k of the bitmap (all lanes take part in the shuffles; TILE_ITERS <= 8 words)
        const uint32_t mine = lane < TILE_ITERS ? (uint32_t)__popc(S.bm[lane]) : 0u;
        uint32_t inc = mine;
#pragma unroll
        for (int o = 1; o < TILE_ITERS; o <<= 1) {
            const uint32_t t = __shfl_up_sync(FULL, inc, o);
            if (lane >= (unsigned)o) inc += t;
        }
        if (lane < TILE_ITERS) S.bmpre[lane] = inc - mine;
    }
    __syncwarp();
}

// ---- B: compare. Striped windows; masks to shared memory; per-record mismatch tallies. Returns (lane 0) the
// mismatch state of the base just before the tile when tile window 0 continues an op.
// Software-pipelined: the loads of window k + 1 are issued as soon as the words of window k have been turned into
// mismatch bits (their registers are free again), so they fly while window k's tail handling, tally and store run.
struct WinLoads {  // one window in flight
    OpEntry en;
#if ENC_LDG128
    uint4 tp[(WIN_WORDS + 2) / 2], qp[(WIN_WORDS + 2) / 2];  // the aligned pairs covering WIN_WORDS + 1 words at either parity
    bool t_odd, q_odd;
#else
    uint64_t tw[WIN_WORDS + 1], qw[WIN_WORDS + 1];
#endif
};

template <bool EXC>
__device__ __forceinline__ void window_issue(const EncodeParams &P, const WarpSmem &S, const TileTab &T, int it, uint32_t &pre, uint32_t n_act,
                                             unsigned lane, unsigned adj, WinLoads &L) {
    constexpr int WW = WIN_WORDS;
    const uint64_t *__restrict__ gw = P.G.words;
    const uint32_t srel = 32u * it + lane;
    const uint32_t w = S.bm[it];
    const uint32_t idx = pre + __popc(w & (0xffffffffu >> (31 - lane))) - adj;
    pre += __popc(w);
    L.en = EXC ? T.get(idx) : S.tab[idx];  // (the slow variant also serves tiles whose table spills)
    const bool isM = srel < n_act && !(L.en.misc & EF_PASS);
    // inactive and pass-through windows read arena word 0: no branches
    const int32_t tw = isM ? L.en.tw + WW * (int32_t)srel : 0;
    const int32_t qw = isM ? L.en.qw + WW * (int32_t)srel : 0;
#if ENC_LDG128
    const uint4 *tp4 = reinterpret_cast<const uint4 *>(gw + (tw & ~1)), *qp4 = reinterpret_cast<const uint4 *>(gw + (qw & ~1));
    L.t_odd = tw & 1; L.q_odd = qw & 1;
#pragma unroll
    for (int k = 0; k < (WW + 2) / 2; k++) L.tp[k] = __ldg(tp4 + k);
#pragma unroll
    for (int k = 0; k < (WW + 2) / 2; k++) L.qp[k] = __ldg(qp4 + k);
#else
#pragma unroll
    for (int k = 0; k <= WW; k++) L.tw[k] = ldg64(gw + tw + k);
#pragma unroll
    for (int k = 0; k <= WW; k++) L.qw[k] = ldg64(gw + qw + k);
#endif
}

template <int BATCH, bool EXC>
__device__ __forceinline__ uint32_t compare_tile(const EncodeParams &P, WarpSmem &S, const TileTab &T, uint32_t n_act, unsigned lane, unsigned adj, uint32_t rec_first) {
    constexpr int WW = WIN_WORDS;
    const uint64_t *__restrict__ gw = P.G.words;
    uint32_t pre = 0, pm0 = 0;
    uint32_t cur = 0, acc = 0;  // record (relative to rec_first) being tallied, and the lane's share of its mismatches
    WinLoads L;
    window_issue<EXC>(P, S, T, 0, pre, n_act, lane, adj, L);
    if (!ENC_NO_PREFETCH) {
        // the first window's loads are in flight: pull the sequence of every later window of the tile into L2 now, so
        // that only this first one waits for DRAM
        uint32_t pre2 = pre;
#pragma unroll
        for (int it = 1; it < TILE_ITERS; it++) {
            const uint32_t srel = 32u * it + lane;
            const uint32_t w = S.bm[it];
            const uint32_t idx = pre2 + __popc(w & (0xffffffffu >> (31 - lane))) - adj;
            pre2 += __popc(w);
            if (srel < n_act) {
                const OpEntry en = EXC ? T.get(idx) : S.tab[idx];
                if (!(en.misc & EF_PASS)) {
                    const uint64_t *tpn = gw + (en.tw + WW * (int32_t)srel), *qpn = gw + (en.qw + WW * (int32_t)srel);
                    // a window spans WW + 1 words, the next lane's window starts WW words on: the middle word of every
                    // window touches every 32-byte sector of the op's span
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(tpn + WW / 2));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(qpn + WW / 2));
                }
            }
        }
    }
#pragma unroll
    for (int it = 0; it < TILE_ITERS; it++) {
        const uint32_t srel = 32u * it + lane;
        const OpEntry en = L.en;
        const bool isM = srel < n_act && !(en.misc & EF_PASS);
        const int32_t rem = entry_rem(en) - WIN_BASES * (int32_t)srel;
        const uint32_t n = !isM ? 0u : rem < WIN_BASES ? (uint32_t)rem : (uint32_t)WIN_BASES;
        const uint32_t misc = en.misc;
        const unsigned sht = misc & 31u, shq = (misc >> 5) & 31u;
        // mismatch bits of the window
        uint32_t m[WW];
        {
            uint64_t TW[WW + 1], QW[WW + 1];
#if ENC_LDG128
            // word j of the loaded pairs: .x/.y of pair j / 2 when j is even, .z/.w when it is odd
            auto word = [](const uint4 *pp, int j) -> uint64_t {
                const uint4 v = pp[j >> 1];
                return (j & 1) ? ((uint64_t)v.w << 32) | v.z : ((uint64_t)v.y << 32) | v.x;
            };
#pragma unroll
            for (int k = 0; k <= WW; k++) {
                TW[k] = L.t_odd ? word(L.tp, k + 1) : word(L.tp, k);
                QW[k] = L.q_odd ? word(L.qp, k + 1) : word(L.qp, k);
            }
#else
#pragma unroll
            for (int k = 0; k <= WW; k++) { TW[k] = L.tw[k]; QW[k] = L.qw[k]; }
#endif
#pragma unroll
            for (int k = 0; k < WW; k++) {
                const uint32_t tl = __funnelshift_r((uint32_t)TW[k], (uint32_t)TW[k + 1], sht);
                const uint32_t th = __funnelshift_r((uint32_t)(TW[k] >> 32), (uint32_t)(TW[k + 1] >> 32), sht);
                const uint32_t ql = __funnelshift_r((uint32_t)QW[k], (uint32_t)QW[k + 1], shq);
                const uint32_t qh = __funnelshift_r((uint32_t)(QW[k] >> 32), (uint32_t)(QW[k + 1] >> 32), shq);
                m[k] = (tl ^ ql) | (th ^ qh);
            }
            if (it == 0 && lane == 0 && isM && !(S.bm[0] & 1u)) {
                // tile window 0 continues an op: mismatch state of the base just before the tile. That base sits one bit
                // down in the window's first word, or (phase 0) on top of the word before - which the aligned pair holds
                // too when the first word is the odd one. Only phase 0 at an even word costs a load of its own.
                if (EXC && (misc & EF_EXC)) {
                    const uint64_t tp = ((uint64_t)(uint32_t)en.tw << 5) + sht, qp = ((uint64_t)(uint32_t)en.qw << 5) + shq;
                    pm0 = slow_mismatch(P.G, tp - 1, qp - 1) ? 0x80000000u : 0u;
                } else {
#if ENC_LDG128
                    const uint64_t pm_t = sht ? TW[0] : L.t_odd ? word(L.tp, 0) : ldg64(gw + (en.tw - 1));
                    const uint64_t pm_q = shq ? QW[0] : L.q_odd ? word(L.qp, 0) : ldg64(gw + (en.qw - 1));
#else
                    const uint64_t pm_t = sht ? TW[0] : ldg64(gw + (en.tw - 1));
                    const uint64_t pm_q = shq ? QW[0] : ldg64(gw + (en.qw - 1));
#endif
                    const unsigned kt = (sht - 1u) & 31u, kq = (shq - 1u) & 31u;
                    const unsigned tc = ((unsigned)(pm_t >> kt) & 1u) | (((unsigned)(pm_t >> (32 + kt)) & 1u) << 1);
                    const unsigned qc = ((unsigned)(pm_q >> kq) & 1u) | (((unsigned)(pm_q >> (32 + kq)) & 1u) << 1);
                    pm0 = tc != qc ? 0x80000000u : 0u;
                }
            }
        }
        // the loaded words are consumed: the next window's loads go out now and fly during the rest of this one
        if (it + 1 < TILE_ITERS) window_issue<EXC>(P, S, T, it + 1, pre, n_act, lane, adj, L);
        if (EXC && isM && (misc & EF_EXC)) {
            const uint64_t tp = ((uint64_t)(uint32_t)(en.tw + WW * (int32_t)srel) << 5) + sht;
            const uint64_t qlo = ((uint64_t)(uint32_t)(en.qw + WW * (int32_t)srel) << 5) + shq;
            if (span_has_exception(P.G, tp, tp + n - 1) || span_has_exception(P.G, qlo, qlo + n - 1)) {
#pragma unroll
                for (int k = 0; k < WW; k++) {
                    uint32_t mk = 0;
                    for (uint32_t j = 32u * k; j < n && j < 32u * k + 32u; j++) mk |= slow_mismatch(P.G, tp + j, qlo + j) << (j & 31u);
                    m[k] = mk;
                }
            }
        }
        // bases beyond the op's end do not exist: shifting them out leaves the op's last base on top, the tally
        // counts what is left, and the arithmetic shift back copies that last base over the tail, so that no run
        // appears to start behind the end (the next window starts another op). 64 bases at a time; a half that lies
        // wholly behind the end repeats the top bit of the half before it.
        uint32_t mm_count = 0;
        uint64_t below = 0;
        const uint32_t wdx = 2u + WW * srel + 2u * ((WW * srel) >> 5);
#pragma unroll
        for (int h = 0; h < WW / 2; h++) {
            const uint64_t raw = isM ? ((uint64_t)m[2 * h + 1] << 32) | m[2 * h] : 0ull;
            const bool some = n > 64u * h;  // this half holds bases of the op (always true for the first half of an M window)
            const uint32_t nh = n - 64u * h;
            const unsigned sh = (isM && some && nh < 64u) ? 64u - nh : 0u;
            const uint64_t src = (h == 0 || some) ? raw : (uint64_t)((int64_t)below >> 63);
            const uint64_t up = src << sh;
            mm_count += (h == 0 || some) ? (uint32_t)__popcll(up) : 0u;
            const uint64_t mx = (uint64_t)((int64_t)up >> sh);
            below = mx;
            *reinterpret_cast<uint64_t *>(&S.mask[wdx + 2 * h]) = mx;
        }
        // per-record mismatch tally. Records do not decrease along the window axis, so the warp tallies one
        // record at a time and moves on together: one reduction and one atomic per record and tile.
        {
            const uint32_t rec = isM ? entry_rec(en) : cur;
            for (;;) {
                acc += rec == cur ? mm_count : 0u;
                if (__all_sync(FULL, rec <= cur)) break;
                const uint32_t sum = __reduce_add_sync(FULL, acc);
                if (lane == 0 && sum) atomicAdd(P.rec_mism + rec_first + cur, (unsigned long long)sum);
                acc = 0;
                cur = __reduce_min_sync(FULL, rec > cur ? rec : 0xffffffffu);
            }
        }
    }
    {
        const uint32_t sum = __reduce_add_sync(FULL, acc);
        if (lane == 0 && sum) atomicAdd(P.rec_mism + rec_first + cur, (unsigned long long)sum);
    }
    return pm0;
}

// ---- C / E: blocked view. Lane owns tile words 16 * lane .. 16 * lane + 15 (32 bases each) = TILE_ITERS windows.
// An "event" is an output op: the start of an =/X run, or a pass-through op (its window holds no mismatch bits and
// starts an op, so it shows up as one forced start bit). Coordinates are tile-relative (32 * word + base).
// List entry of an event (32 bits) = its output word squeezed: bits 24-31 the op code, bits 0-23 the length. A run's
// length is the distance to the next event, known to the lane that walks its events last to first (the last event of
// a lane closes at the first event of the next lane that has any, or where the tile's last op ends); the op pass
// shortens the run before an op start by the bases between the previous op's end and the window boundary, and writes
// pass-through ops in (one whose length does not fit is stored whole after the copy-out, escape_pass).
constexpr uint32_t LE_LEN = 0xffffffu;
constexpr uint32_t COORD_END = 32u * TILE_WORDS;  // stand-in closing coordinate of a run left open at the tile's end

// run-start bits of one 32-base word: m = its mismatch bits, prev = the word before it (top bit used), forced = an op
// starts at bit 0
__device__ __forceinline__ uint32_t start_bits(uint32_t m, uint32_t prev, uint32_t forced) {
    return (m ^ ((m << 1) | (prev >> 31))) | forced;
}

// table index of the op owning tile window w
__device__ __forceinline__ uint32_t window_op(const WarpSmem &S, uint32_t w, unsigned adj) {
    return S.bmpre[w >> 5] + __popc(S.bm[w >> 5] & (0xffffffffu >> (31 - (w & 31)))) - adj;
}

// the lane's op-start windows (bit j = its window j) as a bitmap over its 16 words: bit j -> bit WIN_WORDS * j
__device__ __forceinline__ uint32_t spread_bits(uint32_t starts) {
    uint32_t x = starts;
    if (WIN_WORDS == 2) {
        x = (x | (x << 4)) & 0x0f0fu;
        x = (x | (x << 2)) & 0x3333u;
        x = (x | (x << 1)) & 0x5555u;
    } else {
        x = (x | (x << 6)) & 0x0303u;
        x = (x | (x << 3)) & 0x1111u;
    }
    return x;
}

// C: events of the lane, the bitmap of its words that hold any, and the lane-relative event prefix of its windows
__device__ __forceinline__ void lane_count(WarpSmem &S, unsigned lane, uint32_t n_win, uint32_t starts, uint32_t pm, uint32_t &cnt,
                                           uint32_t &nz) {
    uint32_t *M = S.mask + 2 + 16 * lane + 2 * (lane >> 1);
    if (!(lane & 1)) M[-1] = pm;  // pad word: the listing reads "the word before" without a special case
    cnt = 0;
    nz = 0;
    uint32_t prev = pm;
    uint32_t pk[TILE_ITERS / 2];
#pragma unroll
    for (int j = 0; j < TILE_ITERS; j++) {
        if (j & 1) pk[j >> 1] |= cnt << 16; else pk[j >> 1] = cnt;
        const bool live = (uint32_t)j < n_win;
#pragma unroll
        for (int h = 0; h < WIN_WORDS / 2; h++) {
            const uint2 m = *reinterpret_cast<const uint2 *>(M + WIN_WORDS * j + 2 * h);
            if (live) {
                const uint32_t d0 = start_bits(m.x, prev, h == 0 ? (starts >> j) & 1u : 0u), d1 = start_bits(m.y, m.x, 0u);
                cnt += __popc(d0) + __popc(d1);
                nz |= ((d0 ? 1u : 0u) | (d1 ? 2u : 0u)) << (WIN_WORDS * j + 2 * h);
            }
            prev = m.y;
        }
    }
    static_assert(TILE_ITERS == 8 || TILE_ITERS == 4, "the prefixes of a lane's windows go out as one store");
    if (TILE_ITERS == 8) *reinterpret_cast<uint4 *>(&S.prel[TILE_ITERS * lane]) = make_uint4(pk[0], pk[1], pk[TILE_ITERS / 2 - 2], pk[TILE_ITERS / 2 - 1]);
    else *reinterpret_cast<uint2 *>(&S.prel[TILE_ITERS * lane]) = make_uint2(pk[0], pk[1]);
    uint16_t *nf16 = reinterpret_cast<uint16_t *>(S.nf);
    nf16[4 * (lane >> 1) + (lane & 1)] = (uint16_t)nz;
    nf16[4 * (lane >> 1) + 2 + (lane & 1)] = (uint16_t)spread_bits(starts);
}

static_assert(OP_X == OP_EQ + 1, "the X flag of an event is added to OP_EQ");
constexpr uint32_t LIST_SINK = LIST_CAP;  // entry nobody reads: events outside the round land there

__device__ __forceinline__ uint32_t top_bit(uint32_t x) {
    uint32_t r;
    asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(x));
    return r;
}

// E1: the events of one round, ranks base .. base + n_walk - 1 of the tile, dealt to the lanes in equal shares BY RANK
// (not by position: a warp is as slow as its busiest lane, and run starts cluster). A lane finds the window holding
// the last event of its share (the event prefixes of the count phase, two short binary searches), drops the events of
// that window ranked above its share, then walks its share last to first, one event per trip whatever word it comes
// from; words without events are skipped through the tile's bitmap of words holding events (the current group of 32
// words stays in registers). A run closes where the next event starts: in-lane that is the event handled one trip
// earlier; the last event of a share gets the first event of the next lane's share afterwards (closing = where the
// tile's last run closes; only used when the round holds the tile's last event). The entry of rank base + LIST_CAP
// (first event of the next round, walked only for its coordinate) lands in the sink. Returns the coordinate of the
// lane's first event.
__device__ __forceinline__ uint32_t list_round(WarpSmem &S, unsigned lane, uint32_t base, uint32_t n_walk, uint32_t closing) {
    const uint32_t *M = S.mask;
    const uint32_t q = (n_walk + 31u) >> 5;
    const uint32_t lo_r = lane * q, hi_r = min(lo_r + q, n_walk);
    uint32_t next_c = 0;
    if (lo_r < hi_r) {
        const uint32_t R = base + hi_r - 1u;  // tile rank of the last event of the share
        uint32_t t = 0;
#pragma unroll
        for (uint32_t step = 16; step; step >>= 1)
            if ((uint32_t)S.lanex[t + step] <= R) t += step;
        const uint32_t Rl = R - S.lanex[t];
        uint32_t W = (uint32_t)TILE_ITERS * t;
#pragma unroll
        for (uint32_t step = TILE_ITERS / 2; step; step >>= 1)
            if ((uint32_t)S.prel[W + step] <= Rl) W += step;
        // run starts of window W, word by word: the word that holds the event of lane rank Rl; its events ranked above
        // belong to the next lane
        uint32_t wi = (uint32_t)WIN_WORDS * W;
        const uint32_t ix0 = 2u + wi + 2u * (wi >> 5);
        uint32_t d = 0, m = 0, skip = 0;
        {
            const uint32_t forced0 = (S.nf[wi >> 5].y >> (wi & 31u)) & 1u;
            uint32_t cum = S.prel[W], prevw = M[ix0 - 1], sel = 0;
#pragma unroll
            for (int k = 0; k < WIN_WORDS; k++) {
                const uint32_t mk = M[ix0 + k];
                const uint32_t dk = start_bits(mk, prevw, k == 0 ? forced0 : 0u);
                const uint32_t ck = __popc(dk);
                if (cum <= Rl && Rl < cum + ck) { d = dk; m = mk; sel = (uint32_t)k; skip = cum + ck - 1u - Rl; }
                cum += ck;
                prevw = mk;
            }
            wi += sel;
        }
        for (skip = min(skip, 31u); skip; skip--) d &= ~(0x80000000u >> __clz((int)d));
        // the 32 words below wi as a sliding bitmap (bit i <-> word wb + i; words below 0 do not exist): a share seldom
        // reaches further down, so the walk hardly ever reloads it
        int32_t wb = (int32_t)wi - 32;
        uint32_t avail, fy;
        auto load_bitmap = [&]() {
            const int32_t g = wb >> 5;
            const uint2 lo = g >= 0 ? S.nf[g] : make_uint2(0u, 0u), hi = g >= -1 ? S.nf[(g + 1) & (TILE_WORDS / 32 - 1)] : make_uint2(0u, 0u);
            avail = __funnelshift_r(lo.x, hi.x, (uint32_t)wb & 31u);
            fy = __funnelshift_r(lo.y, hi.y, (uint32_t)wb & 31u);
        };
        load_bitmap();
        uint32_t nc = (uint32_t)OP_EQ << 24;  // closing coordinate of the run being listed, + the '=' code
        for (uint32_t r = hi_r - 1u;; r--) {
            if (d == 0) {
                if (avail == 0) {  // bounded: never spins
                    do {
                        wb -= 32;
                        load_bitmap();
                    } while (avail == 0 && wb > 0);
                }
                const uint32_t bit = top_bit(avail);
                avail ^= 1u << bit;
                wi = (uint32_t)(wb + (int32_t)bit) & (uint32_t)(TILE_WORDS - 1);  // (the mask only matters if the counts lied)
                const uint32_t jx = 2u + wi + 2u * (wi >> 5);
                m = M[jx];
                d = start_bits(m, M[jx - 1], (fy >> bit) & 1u);
            }
            const uint32_t b = top_bit(d);
            d ^= 1u << b;
            const uint32_t c = 32u * wi + b;
            S.list[r] = (((m >> b) & 1u) << 24) + (nc - c);
            nc = c + ((uint32_t)OP_EQ << 24);
            if (r == lo_r) break;
        }
        next_c = nc - ((uint32_t)OP_EQ << 24);
    }
    // the last run of the share closes at the first event of the next share
    const uint32_t up_c = __shfl_down_sync(FULL, next_c, 1);
    if (lo_r < hi_r) S.list[hi_r - 1u] += (lane < 31 && lo_r + q < n_walk) ? up_c : closing;
    return next_c;
}

// rounds after the first (tiles with more than LIST_CAP events): kept out of line, the hot path stays small
__device__ __noinline__ void list_round_dense(WarpSmem &S, unsigned lane, uint32_t base, uint32_t n_walk, uint32_t closing) {
    list_round(S, lane, base, n_walk, closing);
}

// E1': op pass. Every op that starts inside the tile patches the list: the run before it closes at the previous op's
// end, a pass-through op is copied through, the first op of a record leaves the record's first output slot.
// Returns whether some pass-through op of the round does not fit a list entry (escape_pass stores those).
__device__ __forceinline__ bool op_pass(const EncodeParams &P, WarpSmem &S, const TileTab &T, unsigned lane, unsigned adj, uint32_t n_tab, uint32_t rec_first,
                                        uint32_t base, unsigned long long slot0, int64_t tile, uint64_t s0) {
    bool escaped = false;
    for (uint32_t e = lane + (adj ? 0u : 1u); e < n_tab; e += 32) {
        const OpEntry en = T.get(e);
        const uint32_t misc = en.misc;
        const uint32_t w = (misc >> ES_WIN) & 0xffu;
        const uint32_t r = (uint32_t)S.lanex[w / (uint32_t)TILE_ITERS] + S.prel[w];
        const uint32_t k = r - base;
        const uint32_t gap = (misc >> ES_GAP) & (uint32_t)(WIN_BASES - 1);
        if (k == 0) {
            // the run before this op lives in an earlier round (already stored) or an earlier tile (stitch_kernel)
            if (base == 0) P.tile_first_closer[tile] = WIN_BASES * s0 + WIN_BASES * w - gap;
            else if (gap && slot0 + r - 1 < (unsigned long long)P.out_cap) P.out[slot0 + r - 1] -= gap;
        } else if (k < (uint32_t)LIST_CAP && gap) {  // k == LIST_CAP: first event of the next round, patched there
            S.list[k - 1] -= gap;
        }
        if (k < (uint32_t)LIST_CAP) {
            if (misc & EF_PASS) {
                const uint32_t len = (uint32_t)en.tw, top = (uint32_t)en.qw;
                const bool fits = len <= LE_LEN && !(top & LE_LEN);
                S.list[k] = top | (len & LE_LEN);
                escaped |= !fits;
            }
            if (misc & EF_HEAD) P.rec_out_begin[rec_first + entry_rec(en)] = (int64_t)(slot0 + r);
        }
    }
    return __any_sync(FULL, escaped);
}

// pass-through ops of the round whose length needs more than the 24 bits of a list entry: stored whole, after the
// copy-out (out of line: a 16 Mb insertion or deletion is rare)
__device__ __noinline__ void escape_pass(const EncodeParams &P, const WarpSmem &S, const TileTab &T, unsigned lane, unsigned adj, uint32_t n_tab,
                                         uint32_t base, unsigned long long slot0) {
    for (uint32_t e = lane + (adj ? 0u : 1u); e < n_tab; e += 32) {
        const OpEntry en = T.get(e);
        const uint32_t w = (en.misc >> ES_WIN) & 0xffu;
        const uint32_t r = (uint32_t)S.lanex[w / (uint32_t)TILE_ITERS] + S.prel[w];
        const uint32_t len = (uint32_t)en.tw, top = (uint32_t)en.qw;
        if ((en.misc & EF_PASS) && r - base < (uint32_t)LIST_CAP && (len > LE_LEN || (top & LE_LEN)) && slot0 + r < (unsigned long long)P.out_cap)
            P.out[slot0 + r] = ((uint64_t)top << 32) | len;
    }
}

#ifndef ENC_MINB
#define ENC_MINB 8
#endif
#ifndef ENC_DIAG
#define ENC_DIAG 0  // experiments only: bit 0 no conversion pass, 1 no listing, 3 no compare, 4 no table, 5 no look-back
#endif
#ifndef ENC_SLEEP
#define ENC_SLEEP 250  // ns between polls of the look-back (20 ns ... 1 us measure the same; fewer polls leave the issue slots to others)
#endif
#ifndef ENC_WAIT_ALL
#define ENC_WAIT_ALL 0
#endif
template <int WARPS, int BATCH>
__global__ void __launch_bounds__(32 * WARPS, ENC_MINB * 4 / WARPS) encode_kernel(EncodeParams P) {
    // up to 4 warps the tiles' working sets fit static shared memory; the 8-warp variant takes them from dynamic memory
    extern __shared__ __align__(16) unsigned char enc_dyn_smem[];
    __shared__ WarpSmem SM_static[WARPS <= 4 ? WARPS : 1];
    WarpSmem *SM = WARPS <= 4 ? SM_static : reinterpret_cast<WarpSmem *>(enc_dyn_smem);
    __shared__ uint32_t s_wcounts[2][WARPS];  // by iteration parity: a fast warp's next count must not overwrite what a slow one still reads
    __shared__ unsigned long long s_excl;
    __shared__ int64_t s_ticket;
    __shared__ uint32_t s_claim;  // finished listings, all iterations: the first warp of an iteration to finish does the look-back
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    WarpSmem &S = SM[wid];
    const TileTab T{S, P.tab_overflow + ((size_t)blockIdx.x * WARPS + wid) * TILE_TAB};

    // the ticket of the next CTA tile is taken while the current one is finished (its latency hides behind the
    // conversion pass and the barrier that publishes the look-back result also publishes it)
    if (threadIdx.x == 0) { s_ticket = (int64_t)atomicAdd(P.ticket, 1u); s_claim = 0; }
    __syncthreads();
    int64_t ct = s_ticket;
    uint32_t iter = 0;
    if (threadIdx.x == 0) {  // (kept in memory, not in a register, across the loop)
        unsigned long long t_begin;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_begin));
        P.cta_span[2 * blockIdx.x] = t_begin;
    }
    // first op of this warp's tile; for every later tile it is fetched (and the tile's op arrays pulled into L2) while
    // the previous tile is converted, so a tile does not start with a chain of DRAM round trips
    uint32_t first_op = ct * WARPS + wid < P.n_tiles ? __ldg(P.tile_first_op + ct * WARPS + wid) : 0u;
    while (ct < P.n_cta_tiles) {
        uint32_t *s_wcount = s_wcounts[iter & 1u];
        const int64_t tile = ct * WARPS + wid;
        const uint64_t s0 = (uint64_t)tile * TILE_WIN;
        const uint32_t n_act = s0 < P.n_win ? (uint32_t)min((uint64_t)TILE_WIN, P.n_win - s0) : 0u;

        unsigned adj = 0;
        uint32_t cnt = 0, nz = 0, inc = 0, starts = 0, pm = 0, term = COORD_OPEN, n_tab = 0, rec_first = 0;
        if (n_act) {
            bool any_exc;
            if (!(ENC_DIAG & 16)) build_table(P, S, T, (int64_t)first_op, s0, s0 + n_act, lane, adj, n_tab, rec_first, any_exc);
            else { any_exc = false; if (lane < TILE_ITERS) S.bm[lane] = S.bmpre[lane] = 0; }
            uint32_t pm0 = 0;
            if (ENC_DIAG & 24) { for (int q = lane; q < MASK_WORDS; q += 32) S.mask[q] = 0x100u * (q & 3); }
            else pm0 = (any_exc || n_tab > (uint32_t)TAB_SMEM) ? compare_tile<BATCH, true>(P, S, T, n_act, lane, adj, rec_first)
                                         : compare_tile<BATCH, false>(P, S, T, n_act, lane, adj, rec_first);
            __syncwarp();
            // blocked view of the tile
            starts = (S.bm[((uint32_t)TILE_ITERS * lane) >> 5] >> (((uint32_t)TILE_ITERS * lane) & 31u)) & ((1u << TILE_ITERS) - 1u);
            const uint32_t n_win = (uint32_t)TILE_ITERS * lane >= n_act ? 0u : min((uint32_t)TILE_ITERS, n_act - (uint32_t)TILE_ITERS * lane);
            // last word of the lane before (odd lanes: directly before the lane's words; even lanes: before the pad)
            pm = lane ? S.mask[2 + 16 * lane + 2 * (lane >> 1) - ((lane & 1) ? 1 : 3)] : pm0;
            lane_count(S, lane, n_win, starts, pm, cnt, nz);
            inc = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, inc, o);
                if (lane >= (unsigned)o) inc += t;
            }
            S.lanex[lane] = (uint16_t)(inc - cnt);
            // where a run still open at the tile's last window closes: the end of that window's op if it lies inside the tile
            const OpEntry en = T.get(window_op(S, n_act - 1, adj));
            term = (en.misc & EF_PASS) ? WIN_BASES * n_act : entry_rem(en) <= (int32_t)(WIN_BASES * n_act) ? (uint32_t)entry_rem(en) : COORD_OPEN;
        }
        const uint32_t tile_count = __shfl_sync(FULL, inc, 31);
        if (lane == 0) s_wcount[wid] = tile_count;
        __syncthreads();

        // ---- D (first half): publish the CTA tile's number of events, then list round 0 before anybody waits.
        // (atomicMax: the inclusive word written later by another warp must win whatever order the two stores land in;
        // its flag bits make it the larger value)
        volatile unsigned long long *st = P.status;
        unsigned long long agg = 0;
#pragma unroll
        for (int w = 0; w < WARPS; w++) agg += s_wcount[w];
        if (wid == 0 && lane == 0) atomicMax(P.status + ct, (ct == 0 ? ST_INCL : ST_AGG) | agg);
        // round 0 of the event list, before anybody waits for the look-back
        const uint32_t closing = term == COORD_OPEN ? COORD_END : term;
        if (!(ENC_DIAG & 2) && tile_count) {
            const uint32_t first_c = list_round(S, lane, 0u, min(tile_count, (uint32_t)LIST_CAP + 1u), closing);
            // a run left open by an earlier tile closes at this tile's first event (the op pass corrects it when that
            // event is an op start)
            if (lane == 0) P.tile_first_closer[tile] = WIN_BASES * s0 + first_c;
        }
        // ---- D (second half): decoupled look-back over CTA tiles, by the first warp of the CTA that has finished its
        // listing (it would only wait for the others at the barrier below)
        uint32_t claim = 0;
        if (lane == 0) claim = atomicAdd(&s_claim, 1u);
        claim = __shfl_sync(FULL, claim, 0);
        if (claim == iter * (uint32_t)WARPS) {
            unsigned long long excl = 0;
            if (ct > 0 && !(ENC_DIAG & 32)) {
                int64_t j = ct - 1 - (int64_t)lane;
                for (;;) {
                    unsigned long long v = j >= 0 ? st[j] : (2ull << 62);
                    unsigned incl;
                    for (;;) {
                        incl = __ballot_sync(FULL, (v >> 62) == 2);
#if ENC_WAIT_ALL
                        const unsigned need = FULL;
#else
                        // only the tiles back to the nearest inclusive prefix have to be published
                        const unsigned need = incl ? incl ^ (incl - 1u) : FULL;
#endif
                        if (!(__ballot_sync(FULL, (v >> 62) == 0) & need)) break;
                        __nanosleep(ENC_SLEEP);
                        v = j >= 0 ? st[j] : (2ull << 62);
                    }
                    const unsigned upto = incl ? (unsigned)__ffs(incl) - 1u : 31u;  // nearest tile holding an inclusive prefix
                    unsigned long long c = lane <= upto ? (v & ST_VAL) : 0ull;
#pragma unroll
                    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(FULL, c, o);
                    excl += c;
                    if (incl) break;
                    j -= 32;
                }
                if (lane == 0) atomicMax(P.status + ct, ST_INCL | (excl + agg));
            }
            if (lane == 0) {
                s_excl = excl;
                if (ct == P.n_cta_tiles - 1) *P.total_out = excl + agg;
                s_ticket = (int64_t)atomicAdd(P.ticket, 1u);
            }
        }
        __syncthreads();
        const int64_t next_ct = s_ticket;
        const int64_t next_tile = next_ct * WARPS + wid;
        const uint32_t next_first_op = next_tile < P.n_tiles ? __ldg(P.tile_first_op + next_tile) : 0u;

        // ---- E2: op pass, then the list goes out as CigarRecord words with coalesced stores
        if (n_act && !(ENC_DIAG & 1)) {
            unsigned long long slot0 = s_excl;
#pragma unroll
            for (int w = 0; w < WARPS; w++) slot0 += (unsigned)w < wid ? s_wcount[w] : 0u;
            if (lane == 0 && tile_count == 0) P.tile_first_closer[tile] = term == COORD_OPEN ? INF64 : WIN_BASES * s0 + term;
            // events that may be stored (the host re-runs the pass with a larger buffer when some were dropped)
            uint32_t storable = (unsigned long long)P.out_cap > slot0 ? (uint32_t)min((unsigned long long)tile_count, (unsigned long long)P.out_cap - slot0) : 0u;
            asm volatile("" : "+r"(storable));  // keep the loop's bound check 32 bits wide
            for (uint32_t base = 0; base < tile_count; base += LIST_CAP) {
                __syncwarp();
                if (base) list_round_dense(S, lane, base, min(tile_count - base, (uint32_t)LIST_CAP + 1u), closing);  // dense tiles
                __syncwarp();
                const bool escaped = op_pass(P, S, T, lane, adj, n_tab, rec_first, base, slot0, tile, s0);
                __syncwarp();
                const uint32_t round_end = min(base + (uint32_t)LIST_CAP, tile_count);
                const uint32_t lim = min(round_end, storable);
                const uint32_t *lp = S.list + lane;
                uint64_t *po = P.out + slot0 + base + lane;
                for (uint32_t i = base + lane; i < lim; i += 32, lp += 32, po += 32) {
                    const uint32_t e = *lp;
                    *po = ((uint64_t)(e & ~LE_LEN) << 32) | (e & LE_LEN);
                }
                if (escaped) {
                    __syncwarp();
                    escape_pass(P, S, T, lane, adj, n_tab, base, slot0);
                }
                if (term == COORD_OPEN && round_end == tile_count) {
                    // the tile's last run is still open: provisional absolute start, turned into a length by stitch_kernel
                    __syncwarp();
                    if (lane == 0 && tile_count - 1 < storable) {
                        const uint32_t e = S.list[tile_count - 1 - base];
                        const uint64_t abs_start = WIN_BASES * s0 + (COORD_END - (e & LE_LEN));
                        P.out[slot0 + tile_count - 1] = ((uint64_t)(e & ~LE_LEN) << 32) | abs_start;
                        P.tile_open_slot[tile] = (int64_t)(slot0 + tile_count - 1);
                    }
                }
            }
        }
        if (next_tile < P.n_tiles) {
            const int64_t k = (int64_t)next_first_op + lane;
            if (k < P.n_ops) {
                asm volatile("prefetch.global.L2 [%0];" ::"l"(P.woff + k));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(P.ops + k));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(P.tpos + k));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(P.qpos + k));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(P.op_rec + k));
            }
        }
        first_op = next_first_op;
        ct = next_ct;
        iter++;
    }
    if (threadIdx.x == 0) {
        unsigned long long t_end;
        unsigned smid;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
        asm("mov.u32 %0, %%smid;" : "=r"(smid));
        P.cta_span[2 * blockIdx.x] = iter ? t_end - P.cta_span[2 * blockIdx.x] : 0ull;
        P.cta_span[2 * blockIdx.x + 1] = smid;
    }
}

// A run left open at the end of warp tile t closes at the first run start / op end of a later tile.
__global__ void stitch_kernel(int64_t n_tiles, const int64_t *tile_open_slot, const uint64_t *tile_first_closer, uint64_t *out,
                              int64_t out_cap) {
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
