// paftext.cuh - PAF text on the device: raw PAF bytes -> the SoA batch the encode path
// consumes, and the parsed columns -> the text of fields 1-12 + tags in front of every CIGAR. Restates paf_parse /
// cigar_parse (impl/paf.c:70-111, 137-209) and the head of paf_write_to_buffer (impl/paf.c:317-366) as data-parallel
// passes; nothing is copied.
//
//   delimiters   per 4 KB tile of text: newlines and tabs counted, scanned, their positions scattered (nl_pos, tab_pos)
//                with, per line, the index of its first tab - a record then finds its fields by indexing, not scanning
//   records      one thread per line: fields 1-12 (strtok semantics: empty fields do not exist), the tags the reference
//                keeps (tp AS tl cn s1 cg), contig ids through a name hash table, the byte range of the cg value
//   ops          byte-parallel over the text: every M I D = X inside a cg value closes one op whose length is the decimal
//                number before it (two passes: count -> scan -> write); zero-length M ops are dropped here
//                (SURVEY a1-edge ii), anything else inside a cg value marks the record as one the reference aborts on
//   prefix       per record the bytes before the CIGAR are measured and later rendered straight into the output text
// A record the reference would abort on is only LOCATED here (smallest record index, atomicMin); the host renders the
// reference's message for that one line and the run ends there.
#pragma once
#include "kernels.cuh"
#include "serialise.cuh"

namespace pafgpu {

constexpr int TX_THREADS = 256, TX_PER_THREAD = 16, TX_TILE = TX_THREADS * TX_PER_THREAD;

struct NameEntry {      // open-addressing table: name hash -> contig id
    uint64_t hash;      // 0 = empty slot (a real hash of 0 is stored as 1)
    uint32_t name_off, name_len;
    int32_t contig;
    int32_t pad;
};

struct TextRecords {    // parsed columns, one entry per line
    uint32_t *line_begin;  // byte offset of the line; line r ends (its '\n') at nl_pos[r]
    uint32_t *qn_off, *qn_len, *tn_off, *tn_len;
    int64_t *qlen, *qs, *qe, *tlen, *ts, *te, *nmatch, *nbases, *mapq, *score, *tl, *cn, *s1;
    uint8_t *type, *strand, *has_cg;
    uint32_t *cg_begin, *cg_end;
    int32_t *q_contig, *t_contig;
};

__device__ __forceinline__ uint64_t name_hash(const char *s, uint32_t n) {  // FNV-1a
    uint64_t h = 1469598103934665603ull;
    for (uint32_t i = 0; i < n; i++) { h ^= (uint8_t)s[i]; h *= 1099511628211ull; }
    return h ? h : 1;
}

// ---- delimiters ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TX_THREADS) tx_count_delims_kernel(const char *__restrict__ text, int64_t n, uint32_t *tile_nl, uint32_t *tile_tab) {
    __shared__ uint32_t s_nl, s_tab;
    if (threadIdx.x == 0) s_nl = s_tab = 0;
    __syncthreads();
    const int64_t b0 = (int64_t)blockIdx.x * TX_TILE + (int64_t)threadIdx.x * TX_PER_THREAD;
    uint32_t nl = 0, tab = 0;
    if (b0 + TX_PER_THREAD <= n) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(text + b0));
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 16; k++) {
            const uint32_t ch = (w[k >> 2] >> (8 * (k & 3))) & 0xffu;
            nl += ch == '\n';
            tab += ch == '\t';
        }
    } else {
        for (int k = 0; k < TX_PER_THREAD && b0 + k < n; k++) { nl += text[b0 + k] == '\n'; tab += text[b0 + k] == '\t'; }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) { nl += __shfl_xor_sync(FULL, nl, o); tab += __shfl_xor_sync(FULL, tab, o); }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&s_nl, nl); atomicAdd(&s_tab, tab); }
    __syncthreads();
    if (threadIdx.x == 0) { tile_nl[blockIdx.x] = s_nl; tile_tab[blockIdx.x] = s_tab; }
}

// exclusive prefix of (a, b) over the block's threads; totals are not needed here
__device__ __forceinline__ void tx_block_excl2(uint32_t a, uint32_t b, uint32_t &ea, uint32_t &eb) {
    __shared__ uint32_t wa[TX_THREADS / 32], wb[TX_THREADS / 32];
    const unsigned lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t ia = a, ib = b;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t ta = __shfl_up_sync(FULL, ia, o), tb = __shfl_up_sync(FULL, ib, o);
        if (lane >= (unsigned)o) { ia += ta; ib += tb; }
    }
    __syncthreads();
    if (lane == 31) { wa[wid] = ia; wb[wid] = ib; }
    __syncthreads();
    uint32_t oa = 0, ob = 0;
    for (unsigned k = 0; k < wid; k++) { oa += wa[k]; ob += wb[k]; }
    ea = oa + ia - a;
    eb = ob + ib - b;
}

// nl_pos[k] = byte of the k-th newline, tab_pos[k] likewise; line_tab0[k + 1] = tabs before the k-th newline = index of
// the first tab of line k + 1 (line_tab0[0] = 0 is set by the host)
__global__ void __launch_bounds__(TX_THREADS) tx_scatter_delims_kernel(const char *__restrict__ text, int64_t n, const uint64_t *__restrict__ nl_off,
                                                                      const uint64_t *__restrict__ tab_off, uint32_t *nl_pos, uint32_t *tab_pos,
                                                                      uint32_t *line_tab0) {
    const int64_t b0 = (int64_t)blockIdx.x * TX_TILE + (int64_t)threadIdx.x * TX_PER_THREAD;
    char c[TX_PER_THREAD];
    uint32_t nl = 0, tab = 0;
#pragma unroll
    for (int k = 0; k < TX_PER_THREAD; k++) {
        c[k] = b0 + k < n ? text[b0 + k] : (char)0;
        nl += c[k] == '\n';
        tab += c[k] == '\t';
    }
    uint32_t en, et;
    tx_block_excl2(nl, tab, en, et);
    uint64_t in = nl_off[blockIdx.x] + en, it = tab_off[blockIdx.x] + et;
#pragma unroll
    for (int k = 0; k < TX_PER_THREAD; k++) {
        if (c[k] == '\n') { nl_pos[in] = (uint32_t)(b0 + k); line_tab0[in + 1] = (uint32_t)it; in++; }
        else if (c[k] == '\t') tab_pos[it++] = (uint32_t)(b0 + k);
    }
}

// ---- records ------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int64_t tx_to_i64(const char *s, const char *e) {  // impl/paf.c:36-48
    const bool neg = s < e && *s == '-';
    if (neg) s++;
    int64_t v = 0;
    for (; s < e && *s >= '0' && *s <= '9'; s++) v = v * 10 + (*s - '0');
    return neg ? -v : v;
}

struct ParseParams {
    const char *text;
    int64_t n_records;
    const uint32_t *nl_pos, *tab_pos, *line_tab0;
    TextRecords R;
    const NameEntry *names;
    uint32_t name_mask;       // table size - 1
    const char *name_blob;
    unsigned long long *first_bad;  // smallest index of a record the reference would abort on
};

__device__ __forceinline__ int32_t tx_find_contig(const ParseParams &P, const char *s, uint32_t n) {
    const uint64_t h = name_hash(s, n);
    for (uint32_t slot = (uint32_t)h & P.name_mask;; slot = (slot + 1) & P.name_mask) {
        const NameEntry e = P.names[slot];
        if (e.hash == 0) return -1;
        if (e.hash == h && e.name_len == n) {
            bool same = true;
            for (uint32_t i = 0; i < n && same; i++) same = P.name_blob[e.name_off + i] == s[i];
            if (same) return e.contig;
        }
    }
}

__global__ void __launch_bounds__(128) tx_parse_records_kernel(ParseParams P) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P.n_records) return;
    const uint32_t lb = r ? P.nl_pos[r - 1] + 1 : 0u, le = P.nl_pos[r];
    const char *T = P.text;
    uint32_t ti = P.line_tab0[r];  // next tab of the line (tabs of later lines follow: bounded by le)
    uint32_t at = lb;
    // next token [s0, s1): maximal run without tabs, empty ones skipped (strtok_r)
    auto next = [&](uint32_t &s0, uint32_t &s1) -> bool {
        for (;;) {
            if (at >= le) return false;
            const uint32_t tp = P.tab_pos[ti];  // (tab_pos carries a sentinel behind the last tab)
            const uint32_t end = tp < le ? tp : le;
            if (tp < le) ti++;
            s0 = at; s1 = end;
            at = end + 1;
            if (s1 > s0) return true;
        }
    };
    TextRecords &R = P.R;
    R.line_begin[r] = lb;
    uint32_t f0[12], f1[12];
    int nf = 0;
    for (; nf < 12; nf++)
        if (!next(f0[nf], f1[nf])) break;
    bool bad = nf < 12;
    int64_t score = 0, tl = -1, cn = -1, s1v = -1;
    uint8_t type = 0, has_cg = 0, strand = 1;
    uint32_t cgb = 0, cge = 0;
    int32_t qc = -1, tc = -1;
    if (!bad) {
        R.qn_off[r] = f0[0]; R.qn_len[r] = f1[0] - f0[0];
        R.qlen[r] = tx_to_i64(T + f0[1], T + f1[1]); R.qs[r] = tx_to_i64(T + f0[2], T + f1[2]); R.qe[r] = tx_to_i64(T + f0[3], T + f1[3]);
        const char sc = T[f0[4]];
        if (sc != '+' && sc != '-') bad = true;
        strand = sc == '+';
        R.tn_off[r] = f0[5]; R.tn_len[r] = f1[5] - f0[5];
        R.tlen[r] = tx_to_i64(T + f0[6], T + f1[6]); R.ts[r] = tx_to_i64(T + f0[7], T + f1[7]); R.te[r] = tx_to_i64(T + f0[8], T + f1[8]);
        R.nmatch[r] = tx_to_i64(T + f0[9], T + f1[9]); R.nbases[r] = tx_to_i64(T + f0[10], T + f1[10]); R.mapq[r] = tx_to_i64(T + f0[11], T + f1[11]);
        uint32_t s0, s1;
        while (!bad && next(s0, s1)) {
            if (s1 - s0 < 5 || T[s0 + 2] != ':' || T[s0 + 4] != ':') continue;  // not shaped XX:T:value
            const char a = T[s0], c = T[s0 + 1];
            const char *v = T + s0 + 5, *ve = T + s1;
            if (a == 't' && c == 'p') {
                type = v < ve ? (uint8_t)*v : 0;
                if (type != 'P' && type != 'S' && type != 'I') bad = true;  // the reference asserts
            } else if (a == 'A' && c == 'S') score = tx_to_i64(v, ve);
            else if (a == 't' && c == 'l') tl = tx_to_i64(v, ve);
            else if (a == 'c' && c == 'n') cn = tx_to_i64(v, ve);
            else if (a == 's' && c == '1') s1v = tx_to_i64(v, ve);
            else if (a == 'c' && c == 'g') { cgb = s0 + 5; cge = s1; has_cg = cge > cgb; }  // a repeated cg tag replaces the earlier one
        }
        if (!bad) {
            qc = tx_find_contig(P, T + f0[0], f1[0] - f0[0]);
            tc = tx_find_contig(P, T + f0[5], f1[5] - f0[5]);
            if (qc < 0 || tc < 0) bad = true;  // impl/paf_add_mismatches.c:116-127
            // the coordinate tests of paf_check (impl/paf.c:428-439); its op sums are tested by the ingest kernel
            const int64_t ql = R.qlen[r], qs = R.qs[r], qe = R.qe[r], tln = R.tlen[r], ts = R.ts[r], te = R.te[r];
            if (qs < 0 || qs >= ql || qs > qe || qe > ql || ts < 0 || ts >= tln || ts > te || te > tln) bad = true;
        }
    }
    R.score[r] = score; R.tl[r] = tl; R.cn[r] = cn; R.s1[r] = s1v;
    R.type[r] = type; R.strand[r] = strand; R.has_cg[r] = has_cg;
    R.cg_begin[r] = has_cg ? cgb : 0u; R.cg_end[r] = has_cg ? cge : 0u;
    R.q_contig[r] = qc; R.t_contig[r] = tc;
    if (bad) atomicMin(P.first_bad, (unsigned long long)r);
}

// ---- ops ----------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int tx_op_code(char ch) {
    return ch == 'M' ? OP_M : ch == 'I' ? OP_I : ch == 'D' ? OP_D : ch == '=' ? OP_EQ : ch == 'X' ? OP_X : -1;
}
// decimal number ending just before byte i (not reaching below `lo`); digits: how many there were
__device__ __forceinline__ uint64_t tx_number_before(const char *__restrict__ text, uint32_t i, uint32_t lo, uint32_t &digits) {
    uint64_t v = 0, scale = 1;
    digits = 0;
    while (i > lo) {
        const char ch = text[i - 1];
        if (ch < '0' || ch > '9') break;
        v += scale * (uint64_t)(ch - '0');
        scale *= 10;
        i--;
        digits++;
    }
    return v;
}

struct OpsParams {
    const char *text;
    int64_t n;
    int64_t n_records;
    const uint64_t *nl_off;              // newlines before each tile = record of the tile's first byte
    const uint32_t *cg_begin, *cg_end;   // per record
    uint32_t *tile_ops;                  // pass 1: ops per tile
    uint8_t *thread_ops;                 // pass 1 -> pass 2: ops per thread (16 bytes of text hold at most 16)
    const uint64_t *op_off;              // pass 2: ops before each tile
    uint64_t *ops;                       // pass 2
    int64_t *op_begin;                   // pass 2: n_records + 1 op offsets (written at every newline; [0] = 0 is preset)
    unsigned long long *first_bad;
};

// movemask of a per-byte compare result (0xff / 0x00 bytes): bit k = byte k
__device__ __forceinline__ uint32_t tx_mask4(uint32_t m) { return ((m & 0x01010101u) * 0x01020408u) >> 24; }

// WRITE = false: count the ops of the tile. WRITE = true: write them. Thread t owns bytes [16 t, 16 t + 16) (one 16-byte
// load) and classifies them four at a time (byte-wise SIMD compares -> 16-bit masks: newline, digit, op letter, 'M', '0').
// The newline mask cuts the window into line segments; inside a segment the bytes of the record's cg value are a bit
// range, the op letters inside it are the ops (one popc counts them) and everything else there must be a digit. Only a
// letter is ever looked at again: its length is the run of digit bits below it, and a run that reaches the window's
// first byte continues in the bytes before it (one look back per thread at most). A zero-length M vanishes
// (SURVEY a1-edge ii): only an M whose last digit is '0' (or missing) can be one, and those are parsed in the count pass.
template <bool WRITE>
__global__ void __launch_bounds__(TX_THREADS) tx_ops_kernel(OpsParams P) {
    __shared__ uint32_t s_total;
    if (threadIdx.x == 0) s_total = 0;
    const int64_t b0 = (int64_t)blockIdx.x * TX_TILE + (int64_t)threadIdx.x * TX_PER_THREAD;
    uint32_t w[4] = {0u, 0u, 0u, 0u};
    if (b0 + TX_PER_THREAD <= P.n) {  // (the text buffer is 16-byte aligned)
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(P.text + b0));
        w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
    } else {
        for (int k = 0; k < TX_PER_THREAD && b0 + k < P.n; k++) w[k >> 2] |= (uint32_t)(uint8_t)P.text[b0 + k] << (8 * (k & 3));
    }
    uint32_t NL = 0, DG = 0, LT = 0, MM = 0, ZZ = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const uint32_t x = w[j];
        const uint32_t m = __vcmpeq4(x, 0x4d4d4d4du);                                                       // M
        const uint32_t l = m | __vcmpeq4(x, 0x49494949u) | __vcmpeq4(x, 0x44444444u) | __vcmpeq4(x, 0x3d3d3d3du) | __vcmpeq4(x, 0x58585858u);  // I D = X
        NL |= tx_mask4(__vcmpeq4(x, 0x0a0a0a0au)) << (4 * j);
        DG |= tx_mask4(__vcmpgeu4(x, 0x30303030u) & __vcmpleu4(x, 0x39393939u)) << (4 * j);
        LT |= tx_mask4(l) << (4 * j);
        MM |= tx_mask4(m) << (4 * j);
        ZZ |= tx_mask4(__vcmpeq4(x, 0x30303030u)) << (4 * j);
    }
    uint32_t dummy, en;
    tx_block_excl2((uint32_t)__popc(NL), 0u, en, dummy);
    const int64_t rec0 = (int64_t)P.nl_off[blockIdx.x] + en;
    uint32_t cgb0 = 0, cge0 = 0;
    if (rec0 < P.n_records) { cgb0 = P.cg_begin[rec0]; cge0 = P.cg_end[rec0]; }

    auto byte_at = [&](int j) -> uint32_t {
        const uint32_t x = j < 8 ? (j < 4 ? w[0] : w[1]) : (j < 12 ? w[2] : w[3]);
        return (x >> (8 * (j & 3))) & 0xffu;
    };
    // length of the op whose letter is byte p of the window; s = first byte of the cg value inside the window, b = its
    // first byte in the text
    auto op_len = [&](int p, int s, uint32_t b) -> uint64_t {
        const uint32_t below = ~DG & ((1u << p) - 1u);
        int d0 = below ? 32 - __clz(below) : 0;  // first digit of the run that ends at p
        if (d0 < s) d0 = s;
        uint64_t v = 0;
        const int nd = p - d0;
        if (nd <= 8) {
            if (nd > 0) {
                // the eight bytes below the letter as one little-endian word (bytes before the window read as nothing), the
                // ones that are no digits of this number blanked; eight decimal digits -> their value with three multiplies
                const uint64_t lo8 = (uint64_t)w[0] | ((uint64_t)w[1] << 32), hi8 = (uint64_t)w[2] | ((uint64_t)w[3] << 32);
                uint64_t x;
                if (p >= 8) {
                    const unsigned sh = 8u * (unsigned)(p - 8);
                    x = sh == 0 ? lo8 : sh == 64 ? hi8 : (lo8 >> sh) | (hi8 << (64u - sh));
                } else {
                    x = lo8 << (8u * (unsigned)(8 - p));
                }
                const uint64_t keep = ~0ull << (8u * (unsigned)(8 - nd));
                x = (x & keep) - (0x3030303030303030ull & keep);
                x = (x * 10) + (x >> 8);
                v = (((x & 0x000000FF000000FFull) * 0x000F424000000064ull) + (((x >> 16) & 0x000000FF000000FFull) * 0x0000271000000001ull)) >> 32;
            }
        } else {
            for (int j = d0; j < p; j++) v = v * 10 + (uint64_t)(byte_at(j) - '0');
        }
        if (d0 == 0 && (uint32_t)b0 > b) {        // the digits before the window come first
            uint32_t dg;
            uint64_t head = tx_number_before(P.text, (uint32_t)b0, b, dg);
            for (int q = 0; q < p; q++) head *= 10;
            v += head;
        }
        return v;
    };
    // the walk over the line segments of the window; seg(rr, b, ops, s) sees the op letters of one record, line_end(rr)
    // every newline
    auto walk = [&](auto seg, auto line_end) {
        int64_t rr = rec0;
        uint32_t b = cgb0, e = cge0, nlm = NL;
        int lo = 0;
        for (;;) {
            const int hi = nlm ? __ffs(nlm) - 1 : TX_PER_THREAD;
            const int64_t rb = (int64_t)b - b0, re = (int64_t)e - b0;
            const int s = rb > lo ? (int)(rb < hi ? rb : hi) : lo, t = re < hi ? (int)(re > lo ? re : lo) : hi;
            if (s < t) {
                const uint32_t reg = ((1u << t) - 1u) & ~((1u << s) - 1u);
                if (!WRITE) {
                    // impl/paf.c:102 (a byte that is no digit and no op letter), and digits without a letter at the end of
                    // the value (the reference reads the NUL as one)
                    if ((reg & ~(DG | LT)) || (re == t && ((DG >> (t - 1)) & 1u))) atomicMin(P.first_bad, (unsigned long long)rr);
                }
                seg(rr, b, LT & reg, s);
            }
            if (!nlm) break;
            rr++;
            line_end(rr);
            b = e = 0;
            if (rr < P.n_records) { b = P.cg_begin[rr]; e = P.cg_end[rr]; }
            lo = hi + 1;
            nlm &= nlm - 1;
        }
    };

    uint32_t mine = 0;
    const size_t me = (size_t)blockIdx.x * TX_THREADS + threadIdx.x;
    if (!WRITE) {
        walk([&](int64_t, uint32_t b, uint32_t ops, int s) {
                 mine += (uint32_t)__popc(ops);
                 for (uint32_t cand = ops & MM & (((~DG | ZZ) << 1) | 1u); cand; cand &= cand - 1)
                     if (op_len(__ffs(cand) - 1, s, b) == 0) mine--;
             },
             [&](int64_t) {});
        P.thread_ops[me] = (uint8_t)mine;  // the write pass does not count again
#pragma unroll
        for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(FULL, mine, o);
        __syncthreads();
        if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&s_total, mine);
        __syncthreads();
        if (threadIdx.x == 0) P.tile_ops[blockIdx.x] = s_total;
        return;
    }
    mine = P.thread_ops[me];
    uint32_t eo;
    tx_block_excl2(mine, 0u, eo, dummy);
    uint64_t at = P.op_off[blockIdx.x] + eo;
    walk([&](int64_t, uint32_t b, uint32_t ops, int s) {
             for (; ops; ops &= ops - 1) {
                 const int p = __ffs(ops) - 1;
                 const int code = tx_op_code((char)byte_at(p));
                 const uint64_t len = op_len(p, s, b);
                 if (!(code == OP_M && len == 0)) P.ops[at++] = (len & LEN_MASK) | ((uint64_t)code << 56);
             }
         },
         [&](int64_t rr) { P.op_begin[rr] = (int64_t)at; });  // ops before the line that starts here (the last newline leaves the total)
}

// ---- prefix: the bytes before the CIGAR ops ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t tx_i64_chars(int64_t v) {
    uint64_t u = v < 0 ? 0 - (uint64_t)v : (uint64_t)v;
    uint32_t n = v < 0 ? 2 : 1;
    while (u >> 32) { u /= 10; n++; }  // (rare: the 32-bit tail below is the usual path)
    const uint32_t x = (uint32_t)u;
    return n + (x >= 10u) + (x >= 100u) + (x >= 1000u) + (x >= 10000u) + (x >= 100000u) + (x >= 1000000u) + (x >= 10000000u) + (x >= 100000000u) +
           (x >= 1000000000u);
}
__device__ __forceinline__ char *tx_put_i64(char *w, int64_t v) {
    const uint32_t n = tx_i64_chars(v);
    uint64_t u = v < 0 ? 0 - (uint64_t)v : (uint64_t)v;
    if (v < 0) *w = '-';
    char *e = w + n;
    while (u >> 32) { *--e = (char)('0' + u % 10); u /= 10; }
    uint32_t x = (uint32_t)u;
    do { *--e = (char)('0' + x % 10u); x /= 10u; } while (x);
    return w + n;
}

struct PrefixParams {
    const char *text;
    int64_t n_records;
    TextRecords R;
    uint32_t *prefix_len;            // measure pass (read again by the render pass)
    const int64_t *line_begin;       // render pass: where each output line starts / ends
    char *out;
    int64_t out_cap;                 // render pass: bytes `out` holds (a line that would end beyond is not written)
};

constexpr int TXP_THREADS = 128, TXP_SLAB = 196;  // slab: bytes a thread stages its line head in (odd word stride: no bank conflicts)

// everything paf_write_to_buffer prints before the ops (impl/paf.c:320-368), "\tcg:Z:" included when the record has a cigar.
// RENDER: every thread renders its record's head into its slab of shared memory, then the warp copies the 32 slabs out
// one after the other with aligned word stores - 32 threads each storing single bytes into their own line cost a
// memory sector per byte. A head longer than the slab is written straight out by its thread.
template <bool RENDER>
__global__ void __launch_bounds__(TXP_THREADS) tx_prefix_kernel(PrefixParams P) {
    __shared__ __align__(16) uint8_t s_slab[RENDER ? TXP_THREADS * TXP_SLAB + 16 : 16];
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = r < P.n_records;
    const TextRecords &R = P.R;
    uint32_t n = 0;
    int64_t begin = 0;
    bool in_slab = false;
    if (live) {
        const int64_t v12[9] = {R.qlen[r], R.qs[r], R.qe[r], R.tlen[r], R.ts[r], R.te[r], R.nmatch[r], R.nbases[r], R.mapq[r]};
        const int64_t score = R.score[r], tl = R.tl[r], cn = R.cn[r], s1 = R.s1[r];
        const uint8_t type = R.type[r];
        const bool tp = type != 0 || tl != -1;
        if (!RENDER) {
            // q \t len \t start \t end \t strand \t t \t len \t start \t end \t matches \t bases \t mapq: 11 tabs + the strand byte
            n = R.qn_len[r] + R.tn_len[r] + 11 + 1;
#pragma unroll
            for (int k = 0; k < 9; k++) n += tx_i64_chars(v12[k]);
            if (tp) n += 7;
            if (score != 2147483647ll) n += 6 + tx_i64_chars(score);
            if (tl != -1) n += 6 + tx_i64_chars(tl);
            if (cn != -1) n += 6 + tx_i64_chars(cn);
            if (s1 != -1) n += 6 + tx_i64_chars(s1);
            if (R.has_cg[r]) n += 6;
            P.prefix_len[r] = n;
            return;
        }
        n = P.prefix_len[r];
        begin = P.line_begin[r];
        const int64_t end = P.line_begin[r + 1];
        if (end > P.out_cap) {
            n = 0;  // the buffer is too small: the host grows it and runs again
        } else {
            in_slab = n <= (uint32_t)(TXP_SLAB - 4);
            char *w = in_slab ? reinterpret_cast<char *>(s_slab) + threadIdx.x * TXP_SLAB : P.out + begin;
            for (uint32_t i = 0, m = R.qn_len[r]; i < m; i++) *w++ = P.text[R.qn_off[r] + i];
            for (int k = 0; k < 3; k++) { *w++ = '\t'; w = tx_put_i64(w, v12[k]); }
            *w++ = '\t'; *w++ = R.strand[r] ? '+' : '-'; *w++ = '\t';
            for (uint32_t i = 0, m = R.tn_len[r]; i < m; i++) *w++ = P.text[R.tn_off[r] + i];
            for (int k = 3; k < 9; k++) { *w++ = '\t'; w = tx_put_i64(w, v12[k]); }
            auto tag = [&](char a, char b, char t) { *w++ = '\t'; *w++ = a; *w++ = b; *w++ = ':'; *w++ = t; *w++ = ':'; };
            if (tp) { tag('t', 'p', 'A'); *w++ = type ? (char)type : (tl > 1 ? 'S' : 'P'); }
            if (score != 2147483647ll) { tag('A', 'S', 'i'); w = tx_put_i64(w, score); }
            if (tl != -1) { tag('t', 'l', 'i'); w = tx_put_i64(w, tl); }
            if (cn != -1) { tag('c', 'n', 'i'); w = tx_put_i64(w, cn); }
            if (s1 != -1) { tag('s', '1', 'i'); w = tx_put_i64(w, s1); }
            if (R.has_cg[r]) tag('c', 'g', 'Z');
            P.out[end - 1] = '\n';
        }
    }
    if (!RENDER) return;
    __syncwarp();
    const unsigned lane = threadIdx.x & 31, w0 = threadIdx.x & ~31u;
    const uint32_t len = in_slab ? n : 0u;
#pragma unroll 1
    for (int i = 0; i < 32; i++) {
        const uint32_t ni = __shfl_sync(FULL, len, i);
        const int64_t bi = __shfl_sync(FULL, begin, i);
        if (ni) warp_copy_out(P.out + bi, s_slab + (w0 + i) * TXP_SLAB, ni, lane);
    }
}

// ---- totals `paffy view -s` accumulates (impl/paf_view.c:176-178), for the report: sums over the input ops by kind ---
__global__ void __launch_bounds__(256) tx_op_totals_kernel(const uint64_t *__restrict__ ops, int64_t n, unsigned long long *tot /* 8 */) {
    unsigned long long t[7] = {0, 0, 0, 0, 0, 0, 0};  // M bases, = bases, X bases, I count, I bases, D count, D bases
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t w = ops[k];
        const uint64_t len = w & LEN_MASK;
        switch ((int)(w >> 56)) {
            case OP_M: t[0] += len; break;
            case OP_EQ: t[1] += len; break;
            case OP_X: t[2] += len; break;
            case OP_I: t[3]++; t[4] += len; break;
            case OP_D: t[5]++; t[6] += len; break;
        }
    }
#pragma unroll
    for (int q = 0; q < 7; q++) {
#pragma unroll
        for (int o = 16; o; o >>= 1) t[q] += __shfl_xor_sync(FULL, t[q], o);
        if ((threadIdx.x & 31) == 0 && t[q]) atomicAdd(tot + q, t[q]);
    }
}
__global__ void __launch_bounds__(256) tx_sum_kernel(const unsigned long long *__restrict__ v, int64_t n, unsigned long long *out) {
    unsigned long long s = 0;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) s += v[k];
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
    if ((threadIdx.x & 31) == 0 && s) atomicAdd(out, s);
}

// paf_check after a trim (impl/paf_trim.c:124): the coordinate tests on the moved coordinates (the op sums follow them by
// construction of the trim), and the reference's closing assert of paf_trim_unreliable_tails (`failed`). Locates the first
// record that fails.
__global__ void tx_trim_check_kernel(int64_t n_records, const int64_t *__restrict__ qs, const int64_t *__restrict__ qe, const int64_t *__restrict__ ts,
                                     const int64_t *__restrict__ te, const int64_t *__restrict__ qlen, const int64_t *__restrict__ tlen,
                                     const int64_t *__restrict__ op_begin, const uint8_t *__restrict__ failed, unsigned long long *first_bad) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_records) return;
    // a record without ops (no cg tag, or nothing but zero-length M ops): the reference's trim sees 0 / 0 = NaN and its
    // closing assert fires (impl/paf.c:952); the run ends there
    const bool bad = op_begin[r + 1] == op_begin[r] || failed[r] || qs[r] < 0 || qs[r] >= qlen[r] || qs[r] > qe[r] || qe[r] > qlen[r] || ts[r] < 0 || ts[r] >= tlen[r] ||
                     ts[r] > te[r] || te[r] > tlen[r];
    if (bad) atomicMin(first_bad, (unsigned long long)r);
}

// rec_out_begin of records without ops (no cg tag, or only zero-length M ops): the encode kernel never meets them; their
// empty output range starts where the next record with ops starts = the record owning input op op_begin[r].
__global__ void tx_empty_out_begin_kernel(int64_t n_records, const int64_t *__restrict__ op_begin, const uint32_t *__restrict__ op_rec,
                                          int64_t n_ops, const unsigned long long *__restrict__ total_out, int64_t *rec_out_begin) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_records) return;
    const int64_t j = op_begin[r];
    if (op_begin[r + 1] > j) return;
    rec_out_begin[r] = j < n_ops ? rec_out_begin[op_rec[j] & 0x7fffffffu] : (int64_t)*total_out;
}

}  // namespace pafgpu
