// pafgpu.cu - C ABI (include/paffy_b200.h) over the sm_100a kernels: genome store, CIGAR ingest,
// batch encode, result fetch. One context per GPU; no CPU fallback anywhere in this file.
#include "../../include/paffy_b200.h"
#include "kernels.cuh"
#include "encode.cuh"
#include "serialise.cuh"
#include "trim.cuh"
#include "paftext.cuh"
#include "internal.h"

#include <algorithm>
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

using namespace pafgpu;

namespace {

thread_local std::string g_create_error;

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    // grow to at least `bytes`; keep the first `keep` bytes. The zero fill and the copy of the kept bytes are ordered on
    // `stream` (the context's stream: kernels that follow on it see them) and have finished when this returns.
    cudaError_t ensure(size_t bytes, size_t keep = 0, bool zero_new = false, cudaStream_t stream = nullptr) {
        if (bytes <= cap) return cudaSuccess;
        size_t want = std::max(bytes, cap + cap / 2);
        void *q = nullptr;
        cudaError_t e = cudaMalloc(&q, want);
        if (e != cudaSuccess) {  // retry with the exact size before giving up
            want = bytes;
            e = cudaMalloc(&q, want);
            if (e != cudaSuccess) return e;
        }
        if (zero_new) e = cudaMemsetAsync(q, 0, want, stream);
        if (e == cudaSuccess && keep && p) e = cudaMemcpyAsync(q, p, std::min(keep, cap), cudaMemcpyDeviceToDevice, stream);
        if (e == cudaSuccess && (zero_new || (keep && p))) e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) { cudaFree(q); return e; }
        if (p) cudaFree(p);
        p = q;
        cap = want;
        return cudaSuccess;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T> T *as() const { return static_cast<T *>(p); }
};

struct GenomeStore {
    DevBuf words, flags, rank, exc;
    int64_t cap_bases = 0;                // multiple of ARENA_ALIGN
    int64_t next_off = ARENA_ALIGN;       // first free arena position (front pad of one block)
    uint32_t exc_total = 0;               // flagged words so far
    std::vector<std::string> names;
    std::vector<int64_t> off, off_rc, len;  // arena offset of the forward copy, of the reverse-complement copy, length
    std::unordered_map<std::string, int64_t> index;
    DevBuf d_off, d_off_rc, d_len;
    bool table_dirty = true;
    DevBuf d_names, d_name_blob;  // name -> contig id hash table for the device PAF parser (paftext.cuh)
    uint32_t name_mask = 0;
    bool names_dirty = true;
    uint64_t generation = 0;  // bumped by every change of the contig set
    int device = 0;
    std::mutex mu;  // sibling contexts share one store (pafgpu_create_sibling): the lazy table upload is serialised

    Genome view() const { return Genome{words.as<uint64_t>(), flags.as<uint32_t>(), rank.as<uint32_t>(), exc.as<uint8_t>()}; }
    void reset() {
        next_off = ARENA_ALIGN;
        exc_total = 0;
        names.clear(); off.clear(); off_rc.clear(); len.clear(); index.clear();
        table_dirty = names_dirty = true;
    }
    void release() {
        words.release(); flags.release(); rank.release(); exc.release(); d_off.release(); d_off_rc.release(); d_len.release();
        d_names.release(); d_name_blob.release();
    }
    ~GenomeStore() {
        int cur = 0;
        cudaGetDevice(&cur);
        cudaSetDevice(device);
        release();
        cudaSetDevice(cur);
    }
};

}  // namespace

struct pafgpu_ctx {
    int device = 0;
    int sm_count = 148;
    int enc_ctas_per_sm = 8, enc_warps = 4;  // 4 warps per CTA, 8 CTAs per SM: measured best (C2 2.065 ms; 8 warps 2.115, 2 warps 2.164, 16 warps 2.273)
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[8] = {};
    cudaEvent_t ev_tx[4] = {};  // text path: parse begin / end, render begin / end
    std::string err;
    uint64_t host_total = 0;
    cudaEvent_t ev_block = nullptr;  // PAFGPU_BLOCKING_SYNC=1: host threads sleep in stream_sync instead of spinning
    uint64_t *h_words = nullptr;  // pinned + mapped: kernels leave the few words the host has to read here, so that
                                  // those reads never queue behind bulk copies on a copy engine
    std::mutex mu;  // library callers may encode from several threads (SURVEY 8b threading row)
    // helper contexts of the file pipeline (extra streams on this GPU, primaries on the other GPUs), kept for later calls
    struct Helper { int device, k; pafgpu_ctx *ctx; const void *genome_of; uint64_t generation; };
    std::vector<Helper> helpers;

    std::shared_ptr<GenomeStore> genome_ref = std::make_shared<GenomeStore>();  // shared by sibling contexts on the same device
    GenomeStore &G() const { return *genome_ref; }
    GenomeStore scratch;

    // staging for host batches
    DevBuf in_op_begin, in_ops, in_qc, in_tc, in_qs, in_qe, in_ts, in_strand, in_seq;
    // per-batch work buffers
    DevBuf tpos, qpos, op_rec, nwin, woff, scan_sums, tile_first_op, status, tile_first_closer, tile_open_slot, tab_overflow;
    DevBuf out, rec_out_begin, rec_mism, scalars, cta_span;
    unsigned last_grid = 0;
    // serialiser
    DevBuf ser_prefix, ser_prefix_begin, ser_oplen, ser_optoff, ser_line_begin, ser_text;
    // trim
    DevBuf trim_coords, trim_lo, trim_hi, trim_failed;
    // device PAF text ingest (paftext.cuh): chunk text, delimiter positions, parsed columns, parsed ops
    DevBuf tx_text, tx_tile_nl, tx_tile_tab, tx_nl_off, tx_tab_off, tx_nl_pos, tx_tab_pos, tx_line_tab0, tx_cols, tx_tile_ops, tx_op_off, tx_ops,
        tx_scal;
    const int64_t *ingest_t_end = nullptr;   // text mode: paf_check's op sums are tested by the ingest kernel, records without
    const uint8_t *ingest_has_cg = nullptr;  // ops are legal
    // FASTA reader: raw sequence text of one record, kept bytes per tile and their scan, {first bad byte, kept total}
    DevBuf fa_raw, fa_tile_kept, fa_tile_off, fa_scal;
    bool trimmed = false;
    const uint8_t *last_strand = nullptr;
    const int64_t *last_qs = nullptr, *last_qe = nullptr, *last_ts = nullptr;
    int64_t ser_text_bytes = 0;

    // last batch
    bool have_batch = false;
    bool staged_only = false;  // the batch was uploaded already encoded (pafgpu_stage_ops_host): nothing to re-time
    int64_t bad_record = -1;
    int64_t n_records = 0, n_ops = 0, n_out = 0, n_tiles = 0, aligned = 0, retried = 0;
    uint64_t n_win = 0;
    EncodeParams last{};
};

namespace {

constexpr uint32_t TILE_WINDOWS = (uint32_t)TILE_WIN;
enum { SC_BAD = 0, SC_ALIGNED = 1, SC_TOTAL = 2, SC_TICKET = 3, SC_NWIN = 4, SC_TMP = 5, SC_COUNT = 8 };

// wait for the context's stream; with ev_block the thread sleeps (many ranks x contexts on few host cores)
static cudaError_t stream_sync(pafgpu_ctx *c) {
    if (!c->ev_block) return cudaStreamSynchronize(c->stream);
    cudaError_t e = cudaEventRecord(c->ev_block, c->stream);
    return e != cudaSuccess ? e : cudaEventSynchronize(c->ev_block);
}

__global__ void words_out_kernel(volatile uint64_t *host_dst, const uint64_t *src, int n, uint64_t *dev_dst, const uint64_t *dev_src) {
    if ((int)threadIdx.x < n) host_dst[threadIdx.x] = src[threadIdx.x];
    if (dev_dst && threadIdx.x == 0) *dev_dst = *dev_src;
}
__global__ void scalars_init_kernel(uint64_t *sc) { sc[threadIdx.x] = threadIdx.x == SC_BAD ? ~0ull : 0ull; }

// host_dst[0..n) = device words src[0..n) (n <= 32), through the mapped page; optionally *dev_dst = *dev_src on the way.
// Synchronises the stream.
static cudaError_t read_words(pafgpu_ctx *c, uint64_t *host_dst, const uint64_t *src, int n, uint64_t *dev_dst = nullptr,
                              const uint64_t *dev_src = nullptr) {
    words_out_kernel<<<1, 32, 0, c->stream>>>(c->h_words, src, n, dev_dst, dev_src);
    cudaError_t e = stream_sync(c);
    if (e != cudaSuccess) return e;
    for (int i = 0; i < n; i++) host_dst[i] = ((volatile uint64_t *)c->h_words)[i];
    return cudaSuccess;
}

int fail(pafgpu_ctx *c, int code, const std::string &msg) {
    if (c) c->err = msg; else g_create_error = msg;
    return code;
}
#define CU(call)                                                                                    \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess)                                                                      \
            return fail(c, PAFGPU_ENODEVICE, std::string(#call) + ": " + cudaGetErrorString(e_));   \
    } while (0)

inline int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }
inline unsigned blocks_for(int64_t n, int per) { return (unsigned)std::max<int64_t>(1, (n + per - 1) / per); }

// exclusive scan of n u32 values into out[0..n] (u64); total also lands in *total_dev if non-null
int scan_u32(pafgpu_ctx *c, const uint32_t *in, int64_t n, uint64_t *out, uint64_t *total_dev, uint32_t *tile_first_op = nullptr,
             uint32_t tile_windows = 1, int64_t tile_cap = 0) {
    if (n == 0) {
        CU(cudaMemsetAsync(out, 0, sizeof(uint64_t), c->stream));
        if (total_dev) CU(cudaMemsetAsync(total_dev, 0, sizeof(uint64_t), c->stream));
        return 0;
    }
    const int64_t nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    CU(c->scan_sums.ensure((size_t)(nb + 1) * 8));
    uint64_t *tot = total_dev ? total_dev : c->scalars.as<uint64_t>() + SC_TMP;
    {
        // one pass with decoupled look-back: status words of the tiles + the ticket counter behind them, zeroed
        CU(cudaMemsetAsync(c->scan_sums.p, 0, (size_t)(nb + 1) * 8, c->stream));
        scan_onepass_kernel<<<(unsigned)nb, SCAN_THREADS, 0, c->stream>>>(in, n, out, c->scan_sums.as<unsigned long long>(),
                                                                         reinterpret_cast<unsigned int *>(c->scan_sums.as<uint64_t>() + nb), tot,
                                                                         tile_first_op, tile_windows, (uint64_t)tile_cap);
    }
    CU(cudaGetLastError());
    return 0;
}

int genome_grow(pafgpu_ctx *c, GenomeStore &g, int64_t need_bases) {
    if (need_bases <= g.cap_bases) return 0;
    int64_t cap = std::max<int64_t>(need_bases, g.cap_bases + g.cap_bases / 2);
    cap = align_up(cap, ARENA_ALIGN);
    CU(stream_sync(c));
    // +4 words / +2 groups of slack so a window's third word and rank lookups one past the end stay in bounds
    if (cap >= (1ll << 36)) return fail(c, PAFGPU_EINVAL, "genome arena limited to 2^36 bases per GPU (32-bit word indices)");
    CU(g.words.ensure((size_t)(cap / 32 + 4) * 8, (size_t)(g.cap_bases / 32 + 4) * 8, true, c->stream));
    CU(g.flags.ensure((size_t)(cap / 1024 + 2) * 4, (size_t)(g.cap_bases / 1024 + 2) * 4, true, c->stream));
    CU(g.rank.ensure((size_t)(cap / 1024 + 2) * 4, (size_t)(g.cap_bases / 1024 + 2) * 4, true, c->stream));
    g.cap_bases = cap;
    return 0;
}

// pack one copy (forward or reverse complement) of `len` ASCII bytes at device pointer src at arena offset `off`
int genome_pack_copy(pafgpu_ctx *c, GenomeStore &g, const uint8_t *src, int64_t len, int64_t off, int64_t span, bool rc) {
    const int64_t n_words = span / 32, n_groups = span / 1024;
    if (len > 0) {
        pack_kernel<<<blocks_for(n_words, 256), 256, 0, c->stream>>>(src, len, (uint64_t)off / 32, g.words.as<uint64_t>(),
                                                                      g.flags.as<uint32_t>(), n_words, rc);
    }
    // rank directory for the new groups
    CU(c->nwin.ensure((size_t)n_groups * 4));
    CU(c->woff.ensure((size_t)(n_groups + 1) * 8));
    flag_popc_kernel<<<blocks_for(n_groups, 256), 256, 0, c->stream>>>(g.flags.as<uint32_t>() + off / 1024, n_groups,
                                                                        c->nwin.as<uint32_t>());
    if (int rc2 = scan_u32(c, c->nwin.as<uint32_t>(), n_groups, c->woff.as<uint64_t>(), nullptr)) return rc2;
    rank_fill_kernel<<<blocks_for(n_groups, 256), 256, 0, c->stream>>>(c->woff.as<uint64_t>(), n_groups, g.exc_total,
                                                                        g.rank.as<uint32_t>() + off / 1024);
    uint64_t n_exc = 0;
    CU(cudaMemcpyAsync(&n_exc, c->woff.as<uint64_t>() + n_groups, 8, cudaMemcpyDeviceToHost, c->stream));
    CU(stream_sync(c));
    if (n_exc) {
        CU(g.exc.ensure((size_t)(g.exc_total + n_exc) * 32, (size_t)g.exc_total * 32, false, c->stream));
        exc_fill_kernel<<<blocks_for(n_words, 256), 256, 0, c->stream>>>(src, len, (uint64_t)off / 32, g.flags.as<uint32_t>(),
                                                                          g.rank.as<uint32_t>(), g.exc.as<uint8_t>(), n_words, rc);
        g.exc_total += (uint32_t)n_exc;
    } else if (!g.exc.p) {
        CU(g.exc.ensure(32));
    }
    // groups after this copy (up to capacity) must report the new running total
    {
        const int64_t g0 = (off + span) / 1024, g1 = g.cap_bases / 1024 + 2;
        std::vector<uint32_t> fill((size_t)std::min<int64_t>(g1 - g0, 4), g.exc_total);
        // only the next pad / next contig start is ever consulted before being rewritten; set a few groups
        CU(cudaMemcpyAsync(g.rank.as<uint32_t>() + g0, fill.data(), fill.size() * 4, cudaMemcpyHostToDevice, c->stream));
        CU(stream_sync(c));
    }
    CU(cudaGetLastError());
    return 0;
}

// pack `len` ASCII bytes at device pointer src as a new contig: forward copy, then its reverse complement
int64_t genome_add_device(pafgpu_ctx *c, GenomeStore &g, const char *name, const uint8_t *src, int64_t len) {
    if (len < 0 || !name) return fail(c, PAFGPU_EINVAL, "bad contig");
    const int64_t off = g.next_off;
    const int64_t span = align_up(len, ARENA_ALIGN) + ARENA_ALIGN;  // contig + trailing pad block
    if (int rc = genome_grow(c, g, off + 2 * span + ARENA_ALIGN)) return rc;
    if (int rc = genome_pack_copy(c, g, src, len, off, span, false)) return rc;
    if (int rc = genome_pack_copy(c, g, src, len, off + span, span, true)) return rc;
    const int64_t id = (int64_t)g.names.size();
    g.names.emplace_back(name);
    g.off.push_back(off);
    g.off_rc.push_back(off + span);
    g.len.push_back(len);
    g.index[g.names.back()] = id;  // a repeated name points at the newest contig (sonLibHash.c:68-73)
    g.next_off = off + 2 * span;
    g.table_dirty = g.names_dirty = true;
    g.generation++;
    return id;
}

int64_t genome_add_host(pafgpu_ctx *c, GenomeStore &g, const char *name, const char *seq, int64_t len) {
    CU(c->in_seq.ensure((size_t)std::max<int64_t>(len, 16)));
    if (len > 0) CU(cudaMemcpyAsync(c->in_seq.p, seq, (size_t)len, cudaMemcpyHostToDevice, c->stream));
    return genome_add_device(c, g, name, c->in_seq.as<uint8_t>(), len);
}

int genome_sync_table(pafgpu_ctx *c, GenomeStore &g) {
    std::lock_guard<std::mutex> lk(g.mu);
    if (!g.table_dirty) return 0;
    const size_t n = g.off.size();
    CU(g.d_off.ensure(std::max<size_t>(n, 1) * 8));
    CU(g.d_len.ensure(std::max<size_t>(n, 1) * 8));
    CU(g.d_off_rc.ensure(std::max<size_t>(n, 1) * 8));
    if (n) {
        CU(cudaMemcpyAsync(g.d_off.p, g.off.data(), n * 8, cudaMemcpyHostToDevice, c->stream));
        CU(cudaMemcpyAsync(g.d_off_rc.p, g.off_rc.data(), n * 8, cudaMemcpyHostToDevice, c->stream));
        CU(cudaMemcpyAsync(g.d_len.p, g.len.data(), n * 8, cudaMemcpyHostToDevice, c->stream));
        CU(stream_sync(c));
    }
    g.table_dirty = false;
    return 0;
}

#ifndef ENC_BATCH
#define ENC_BATCH (ENC_WIN_WORDS == 4 ? 1 : 2)  // windows whose loads a lane issues together: 10 / 12 64-bit loads in flight
#endif
constexpr int CMP_BATCH = ENC_BATCH;

// one pass of encode_kernel + stitch_kernel for c->last; leaves the output op count in c->n_out
int launch_encode_once(pafgpu_ctx *c) {
    EncodeParams &P = c->last;
    uint64_t *sc = c->scalars.as<uint64_t>();
    const int64_t T = c->n_tiles;
    const int W = c->enc_warps;
    P.n_cta_tiles = (T + W - 1) / W;
    const int64_t CT = P.n_cta_tiles;
    CU(cudaMemsetAsync(c->tile_open_slot.p, 0xff, (size_t)std::max<int64_t>(T, 1) * 8, c->stream));
    CU(cudaMemsetAsync(c->rec_mism.p, 0, (size_t)std::max<int64_t>(c->n_records, 1) * 8, c->stream));
    CU(cudaMemsetAsync(c->status.p, 0, (size_t)std::max<int64_t>(CT, 1) * 8, c->stream));
    CU(cudaMemsetAsync(sc + SC_TOTAL, 0, 16, c->stream));  // SC_TOTAL and SC_TICKET
    P.out = c->out.as<uint64_t>();
    P.out_cap = (int64_t)(c->out.cap / 8);
    {
        const size_t warps = (size_t)c->sm_count * c->enc_ctas_per_sm * W;
        CU(c->tab_overflow.ensure(warps * TILE_TAB * sizeof(OpEntry)));
        P.tab_overflow = c->tab_overflow.as<OpEntry>();
    }
    CU(c->cta_span.ensure((size_t)c->sm_count * c->enc_ctas_per_sm * 16));
    P.cta_span = c->cta_span.as<unsigned long long>();
    c->last_grid = 0;
    CU(cudaEventRecord(c->ev[4], c->stream));
    if (CT > 0) {
        const unsigned grid = (unsigned)std::min<int64_t>(CT, (int64_t)c->sm_count * c->enc_ctas_per_sm);
        c->last_grid = grid;
        if (W == 1) encode_kernel<1, CMP_BATCH><<<grid, 32, 0, c->stream>>>(P);
        else if (W == 2) encode_kernel<2, CMP_BATCH><<<grid, 64, 0, c->stream>>>(P);
        else if (W == 8) encode_kernel<8, CMP_BATCH><<<grid, 256, 8 * sizeof(WarpSmem), c->stream>>>(P);
        else if (W == 16) encode_kernel<16, CMP_BATCH><<<grid, 512, 16 * sizeof(WarpSmem), c->stream>>>(P);
        else encode_kernel<4, CMP_BATCH><<<grid, 128, 0, c->stream>>>(P);
    }
    CU(cudaEventRecord(c->ev[5], c->stream));
    if (T > 0) stitch_kernel<<<blocks_for(T, 256), 256, 0, c->stream>>>(T, P.tile_open_slot, P.tile_first_closer, P.out, P.out_cap);
    CU(cudaEventRecord(c->ev[7], c->stream));
    CU(read_words(c, &c->host_total, sc + SC_TOTAL, 1, (uint64_t *)(P.rec_out_begin + c->n_records), sc + SC_TOTAL));
    CU(cudaGetLastError());
    c->n_out = (int64_t)c->host_total;
    return 0;
}

// The number of output ops is only known after the pass: the buffer is sized from a guess (or the last batch) and,
// when the guess was too small, grown to the exact size and the pass repeated (stores beyond capacity are dropped).
int launch_encode(pafgpu_ctx *c, int64_t aligned) {
    const int64_t guess = c->n_ops + aligned / 32 + 1024;
    CU(c->out.ensure((size_t)guess * 8));
    c->retried = 0;
    if (int rc = launch_encode_once(c)) return rc;
    if (c->n_out > (int64_t)(c->out.cap / 8)) {
        CU(c->out.ensure((size_t)c->n_out * 8));
        c->retried = 1;
        if (int rc = launch_encode_once(c)) return rc;
    }
    return 0;
}

// all arrays of *b are device pointers here
int encode_device(pafgpu_ctx *c, GenomeStore &g, const pafgpu_batch *b, pafgpu_result *res) {
    if (!b || b->n_records < 0 || b->n_ops < 0) return fail(c, PAFGPU_EINVAL, "bad batch");
    if (b->n_ops >= (1ll << 32) || b->n_records >= (1ll << 31)) return fail(c, PAFGPU_EINVAL, "batch too large: split it");
    if (int rc = genome_sync_table(c, g)) return rc;
    c->have_batch = false;
    c->trimmed = false;
    c->bad_record = -1;
    const int64_t R = b->n_records, N = b->n_ops;
    CU(c->scalars.ensure(SC_COUNT * 8));
    CU(c->tpos.ensure((size_t)std::max<int64_t>(N, 1) * 8));
    CU(c->qpos.ensure((size_t)std::max<int64_t>(N, 1) * 8));
    CU(c->op_rec.ensure((size_t)std::max<int64_t>(N, 1) * 4));
    CU(c->nwin.ensure((size_t)std::max<int64_t>(N, 1) * 4));
    CU(c->woff.ensure((size_t)(N + 1) * 8));
    CU(c->rec_out_begin.ensure((size_t)(R + 1) * 8));
    CU(c->rec_mism.ensure((size_t)std::max<int64_t>(R, 1) * 8));
    uint64_t *sc = c->scalars.as<uint64_t>();
    scalars_init_kernel<<<1, SC_COUNT, 0, c->stream>>>(sc);

    CU(cudaEventRecord(c->ev[0], c->stream));
    IngestParams I{};
    I.n_records = R; I.n_ops = N;
    I.op_begin = b->op_begin; I.ops = b->ops;
    I.q_contig = b->q_contig; I.t_contig = b->t_contig;
    I.q_start = b->q_start; I.q_end = b->q_end; I.t_start = b->t_start; I.same_strand = b->same_strand;
    I.contig_off = g.d_off.as<int64_t>(); I.contig_off_rc = g.d_off_rc.as<int64_t>(); I.contig_len = g.d_len.as<int64_t>(); I.n_contigs = (int64_t)g.off.size();
    I.G = g.view();
    I.has_exceptions = g.exc_total > 0;
    I.t_end = c->ingest_t_end; I.has_cg = c->ingest_has_cg;
    I.tpos = c->tpos.as<uint64_t>(); I.qpos = c->qpos.as<uint64_t>();
    I.op_rec = c->op_rec.as<uint32_t>(); I.nwin = c->nwin.as<uint32_t>();
    I.bad_record = (unsigned long long *)(sc + SC_BAD);
    I.aligned = (unsigned long long *)(sc + SC_ALIGNED);
    if (R > 0) {
        // lanes per record: fewer lanes = more records in flight per SM, which hides the chain of dependent loads
        // per record better than finishing a record in one stride (measured at C2, 20 ops per record: 4 lanes 0.53 ms,
        // 8 lanes 0.58 ms, 16 lanes 0.69 ms, 32 lanes 0.90 ms for the ingest stage)
        int64_t mean_ops = N / R;
        // (round 2: 2 lanes 0.60 ms, 1 lane 0.93 ms, 8 lanes 0.43 ms against 0.39 ms with 4)
        if (const char *e = getenv("PAFGPU_INGEST_G")) mean_ops = atoi(e) <= 4 ? 1 : atoi(e) == 8 ? 100 : atoi(e) == 16 ? 1000 : 100000;  // experiments
        if (mean_ops <= 64) ingest_kernel<4><<<blocks_for(R * 4, 256), 256, 0, c->stream>>>(I);
        else if (mean_ops <= 256) ingest_kernel<8><<<blocks_for(R * 8, 256), 256, 0, c->stream>>>(I);
        else if (mean_ops <= 2048) ingest_kernel<16><<<blocks_for(R * 16, 256), 256, 0, c->stream>>>(I);
        else ingest_kernel<32><<<blocks_for(R * 32, 256), 256, 0, c->stream>>>(I);
    }
    // the scan of the windows per op also leaves the first op of every warp tile; the tile count is only known after it,
    // so the array is sized from a guess and the separate kernel runs in the rare case that was too small
    CU(c->tile_first_op.ensure((size_t)std::max<int64_t>(std::max<int64_t>(c->n_tiles + c->n_tiles / 2, N / 8), 1024) * 4));
    const int64_t tfo_cap = (int64_t)(c->tile_first_op.cap / 4);
    if (int rc = scan_u32(c, c->nwin.as<uint32_t>(), N, c->woff.as<uint64_t>(), sc + SC_NWIN, c->tile_first_op.as<uint32_t>(), TILE_WINDOWS,
                          tfo_cap))
        return rc;
    uint64_t host_sc[SC_COUNT];
    CU(read_words(c, host_sc, sc, SC_COUNT));
    if (host_sc[SC_BAD] != ~0ull) {
        char msg[160];
        snprintf(msg, sizeof msg, "record %llu: CIGAR/coordinates walk outside the contig (or unknown contig id)",
                 (unsigned long long)host_sc[SC_BAD]);
        c->bad_record = (int64_t)host_sc[SC_BAD];
        return fail(c, PAFGPU_EBOUNDS, msg);
    }
    const uint64_t n_win = host_sc[SC_NWIN];
    const int64_t aligned = (int64_t)host_sc[SC_ALIGNED];
    const int64_t n_tiles = (int64_t)((n_win + TILE_WINDOWS - 1) / TILE_WINDOWS);
    const bool tfo_done = n_tiles <= tfo_cap;
    CU(c->tile_first_op.ensure((size_t)std::max<int64_t>(n_tiles, 1) * 4));
    CU(c->status.ensure((size_t)std::max<int64_t>(n_tiles, 1) * 8));  // one word per CTA tile; a CTA tile is >= 1 warp tile
    CU(c->tile_first_closer.ensure((size_t)std::max<int64_t>(n_tiles, 1) * 8));
    CU(c->tile_open_slot.ensure((size_t)std::max<int64_t>(n_tiles, 1) * 8));
    if (N > 0 && !tfo_done)
        tile_first_op_kernel<<<blocks_for(N, 256), 256, 0, c->stream>>>(c->woff.as<uint64_t>(), N, TILE_WINDOWS,
                                                                        c->tile_first_op.as<uint32_t>());
    CU(cudaEventRecord(c->ev[1], c->stream));

    EncodeParams &P = c->last;
    P.G = g.view();
    P.ops = b->ops; P.woff = c->woff.as<uint64_t>(); P.tpos = c->tpos.as<uint64_t>(); P.qpos = c->qpos.as<uint64_t>();
    P.op_rec = c->op_rec.as<uint32_t>(); P.tile_first_op = c->tile_first_op.as<uint32_t>();
    P.n_ops = N; P.n_win = n_win; P.n_tiles = n_tiles;
    P.out = c->out.as<uint64_t>(); P.out_cap = (int64_t)(c->out.cap / 8);
    P.status = c->status.as<unsigned long long>();
    P.ticket = (unsigned int *)(sc + SC_TICKET);
    P.total_out = (unsigned long long *)(sc + SC_TOTAL);
    P.tile_first_closer = c->tile_first_closer.as<uint64_t>();
    P.tile_open_slot = c->tile_open_slot.as<int64_t>();
    c->last_strand = b->same_strand; c->last_qs = b->q_start; c->last_qe = b->q_end; c->last_ts = b->t_start;
    P.rec_out_begin = c->rec_out_begin.as<int64_t>();
    P.rec_mism = c->rec_mism.as<unsigned long long>();
    c->n_records = R; c->n_ops = N; c->n_tiles = n_tiles; c->n_win = n_win;

    CU(cudaEventRecord(c->ev[2], c->stream));
    c->aligned = aligned;
    if (int rc = launch_encode(c, aligned)) return rc;
    CU(cudaEventRecord(c->ev[3], c->stream));
    CU(stream_sync(c));
    c->have_batch = true;
    c->staged_only = false;
    if (res) {
        memset(res, 0, sizeof *res);
        res->n_out_ops = c->n_out;
        res->aligned_bases = aligned;
        res->n_windows = (int64_t)n_win;
        res->n_tiles = n_tiles;
        res->retried = c->retried;
        cudaEventElapsedTime(&res->ms_ingest, c->ev[0], c->ev[1]);
        cudaEventElapsedTime(&res->ms_encode, c->ev[2], c->ev[3]);
        cudaEventElapsedTime(&res->ms_kernel, c->ev[4], c->ev[5]);
        cudaEventElapsedTime(&res->ms_emit, c->ev[5], c->ev[7]);
    }
    return 0;
}

int encode_host(pafgpu_ctx *c, GenomeStore &g, const pafgpu_batch *b, pafgpu_result *res) {
    if (!b) return fail(c, PAFGPU_EINVAL, "null batch");
    const int64_t R = b->n_records, N = b->n_ops;
    auto up = [&](DevBuf &d, const void *src, size_t bytes) -> int {
        CU(d.ensure(std::max<size_t>(bytes, 8)));
        if (bytes) CU(cudaMemcpyAsync(d.p, src, bytes, cudaMemcpyHostToDevice, c->stream));
        return 0;
    };
    int rc;
    if ((rc = up(c->in_op_begin, b->op_begin, (size_t)(R + 1) * 8))) return rc;
    if ((rc = up(c->in_ops, b->ops, (size_t)N * 8))) return rc;
    if ((rc = up(c->in_qc, b->q_contig, (size_t)R * 4))) return rc;
    if ((rc = up(c->in_tc, b->t_contig, (size_t)R * 4))) return rc;
    if ((rc = up(c->in_qs, b->q_start, (size_t)R * 8))) return rc;
    if ((rc = up(c->in_qe, b->q_end, (size_t)R * 8))) return rc;
    if ((rc = up(c->in_ts, b->t_start, (size_t)R * 8))) return rc;
    if ((rc = up(c->in_strand, b->same_strand, (size_t)R))) return rc;
    pafgpu_batch d = *b;
    d.op_begin = c->in_op_begin.as<int64_t>(); d.ops = c->in_ops.as<uint64_t>();
    d.q_contig = c->in_qc.as<int32_t>(); d.t_contig = c->in_tc.as<int32_t>();
    d.q_start = c->in_qs.as<int64_t>(); d.q_end = c->in_qe.as<int64_t>(); d.t_start = c->in_ts.as<int64_t>();
    d.same_strand = c->in_strand.as<uint8_t>();
    return encode_device(c, g, &d, res);
}

int fetch_ops(pafgpu_ctx *c, uint64_t *out_ops, int64_t *out_op_begin, int64_t *rec_mismatch) {
    if (!c->have_batch) return fail(c, PAFGPU_ESTATE, "no batch has been encoded");
    if (out_ops && c->n_out) CU(cudaMemcpyAsync(out_ops, c->out.p, (size_t)c->n_out * 8, cudaMemcpyDeviceToHost, c->stream));
    if (out_op_begin)
        CU(cudaMemcpyAsync(out_op_begin, c->rec_out_begin.p, (size_t)(c->n_records + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
    if (rec_mismatch && c->n_records)
        CU(cudaMemcpyAsync(rec_mismatch, c->rec_mism.p, (size_t)c->n_records * 8, cudaMemcpyDeviceToHost, c->stream));
    CU(stream_sync(c));
    return 0;
}

}  // namespace

// ================================================================================================
extern "C" {

int pafgpu_abi_version(void) { return PAFGPU_ABI_VERSION; }

int pafgpu_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int pafgpu_create(int device, pafgpu_ctx **out) {
    pafgpu_ctx *c = nullptr;
    if (!out) return fail(c, PAFGPU_EINVAL, "null out");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(c, PAFGPU_ENODEVICE, std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "count is 0") +
                                             " (paffy_b200 has no CPU fallback)");
    if (device < 0 || device >= n) return fail(c, PAFGPU_EINVAL, "device index out of range");
    CU(cudaSetDevice(device));
    pafgpu_ctx *ctx = new pafgpu_ctx();
    ctx->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) ctx->sm_count = prop.multiProcessorCount;
    c = ctx;
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete ctx;
        c = nullptr;
        return fail(c, PAFGPU_ENODEVICE, "cudaStreamCreate failed");
    }
    for (auto &ev : ctx->ev) cudaEventCreate(&ev);
    for (auto &ev : ctx->ev_tx) cudaEventCreate(&ev);
    if (const char *e = getenv("PAFGPU_BLOCKING_SYNC"))
        if (atoi(e)) cudaEventCreateWithFlags(&ctx->ev_block, cudaEventBlockingSync | cudaEventDisableTiming);
    if (cudaHostAlloc((void **)&ctx->h_words, 64 * sizeof(uint64_t), cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        cudaStreamDestroy(ctx->stream);
        delete ctx;
        c = nullptr;
        return fail(c, PAFGPU_ENODEVICE, "cudaHostAlloc failed");
    }
    {
        int per_sm = 0;
        if (const char *e = getenv("PAFGPU_WARPS")) ctx->enc_warps = atoi(e) == 1 ? 1 : atoi(e) == 2 ? 2 : atoi(e) == 8 ? 8 : atoi(e) == 16 ? 16 : 4;
        if (ctx->enc_warps == 8)
            cudaFuncSetAttribute(encode_kernel<8, CMP_BATCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(8 * sizeof(WarpSmem)));
        if (ctx->enc_warps == 16)
            cudaFuncSetAttribute(encode_kernel<16, CMP_BATCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(16 * sizeof(WarpSmem)));
        cudaError_t oe = ctx->enc_warps == 16  ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, encode_kernel<16, CMP_BATCH>, 512, 16 * sizeof(WarpSmem))
                         : ctx->enc_warps == 8   ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, encode_kernel<8, CMP_BATCH>, 256, 8 * sizeof(WarpSmem))
                         : ctx->enc_warps == 1   ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, encode_kernel<1, CMP_BATCH>, 32, 0)
                         : ctx->enc_warps == 2 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, encode_kernel<2, CMP_BATCH>, 64, 0)
                                               : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, encode_kernel<4, CMP_BATCH>, 128, 0);
        if (oe == cudaSuccess && per_sm > 0) ctx->enc_ctas_per_sm = per_sm;
    }
    ctx->scalars.ensure(SC_COUNT * 8);
    ctx->G().device = device;
    *out = ctx;
    return 0;
}

int pafgpu_create_sibling(pafgpu_ctx *parent, pafgpu_ctx **out) {
    if (!parent || !out) return PAFGPU_EINVAL;
    const int rc = pafgpu_create(parent->device, out);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(parent->mu);
    (*out)->genome_ref = parent->genome_ref;  // the store the new context was created with is dropped (it owns nothing yet)
    return 0;
}

int pafgpu_device_of(const pafgpu_ctx *c) { return c ? c->device : PAFGPU_EINVAL; }

void pafgpu_destroy(pafgpu_ctx *c) {
    if (!c) return;
    for (auto &h : c->helpers) pafgpu_destroy(h.ctx);
    c->helpers.clear();
    cudaSetDevice(c->device);
    stream_sync(c);
    c->scratch.release();  // (the shared genome store goes with its last context)
    DevBuf *bufs[] = {&c->in_op_begin, &c->in_ops, &c->in_qc, &c->in_tc, &c->in_qs, &c->in_qe, &c->in_ts, &c->in_strand,
                      &c->in_seq, &c->tpos, &c->qpos, &c->op_rec, &c->nwin, &c->woff, &c->scan_sums, &c->tile_first_op,
                      &c->status, &c->tile_first_closer, &c->tile_open_slot, &c->tab_overflow, &c->out, &c->rec_out_begin, &c->rec_mism,
                      &c->scalars, &c->cta_span, &c->trim_coords, &c->trim_lo, &c->trim_hi, &c->trim_failed, &c->ser_prefix, &c->ser_prefix_begin, &c->ser_oplen, &c->ser_optoff, &c->ser_line_begin,
                      &c->ser_text, &c->tx_text, &c->tx_tile_nl, &c->tx_tile_tab, &c->tx_nl_off, &c->tx_tab_off, &c->tx_nl_pos, &c->tx_tab_pos,
                      &c->tx_line_tab0, &c->tx_cols, &c->tx_tile_ops, &c->tx_op_off, &c->tx_ops, &c->tx_scal, &c->fa_raw, &c->fa_tile_kept,
                      &c->fa_tile_off, &c->fa_scal};
    for (DevBuf *b : bufs) b->release();
    for (auto &ev : c->ev) if (ev) cudaEventDestroy(ev);
    for (auto &ev : c->ev_tx) if (ev) cudaEventDestroy(ev);
    if (c->ev_block) cudaEventDestroy(c->ev_block);
    if (c->h_words) cudaFreeHost(c->h_words);
    cudaStreamDestroy(c->stream);
    delete c;
}

const char *pafgpu_last_error(const pafgpu_ctx *c) { return c ? c->err.c_str() : g_create_error.c_str(); }

int64_t pafgpu_last_bad_record(const pafgpu_ctx *c) { return c ? c->bad_record : -1; }

void *pafgpu_host_alloc(int64_t bytes) {
    void *p = nullptr;
    if (bytes <= 0 || cudaHostAlloc(&p, (size_t)bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void pafgpu_host_free(void *p) { if (p) cudaFreeHost(p); }

// ---- genome store ------------------------------------------------------------------------------
int pafgpu_genome_reserve(pafgpu_ctx *c, int64_t total_bases) {
    if (!c) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    return genome_grow(c, c->G(), c->G().next_off + 2 * total_bases + 4 * ARENA_ALIGN);  // forward + reverse-complement copies
}

int64_t pafgpu_genome_add_contig(pafgpu_ctx *c, const char *name, const char *seq, int64_t len) {
    if (!c || !name || (!seq && len)) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    return genome_add_host(c, c->G(), name, seq, len);
}

int64_t pafgpu_genome_add_contig_device(pafgpu_ctx *c, const char *name, const void *seq_dev, int64_t len) {
    if (!c || !name || (!seq_dev && len)) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    CU(cudaDeviceSynchronize());  // the producer of seq_dev may be on another stream
    return genome_add_device(c, c->G(), name, static_cast<const uint8_t *>(seq_dev), len);
}

// One FASTA record whose sequence text (everything between the header line and the next '>') is raw[0..n): the text goes
// to the device as it is, fasta_count_kernel / fasta_compact_kernel drop newline, space and tab and test every other
// byte, and the contiguous sequence is packed straight from device memory.
static int fasta_add_record(pafgpu_ctx *c, const std::string &name, const char *raw, int64_t n) {
    CU(c->in_seq.ensure((size_t)std::max<int64_t>(n, 16)));
    if (n == 0) {
        const int64_t id = genome_add_device(c, c->G(), name.c_str(), c->in_seq.as<uint8_t>(), 0);
        return id < 0 ? (int)id : 0;
    }
    const int64_t nt = (n + FA_TILE - 1) / FA_TILE;
    CU(c->fa_raw.ensure((size_t)n));
    CU(c->fa_tile_kept.ensure((size_t)nt * 4));
    CU(c->fa_tile_off.ensure((size_t)(nt + 1) * 8));
    CU(c->fa_scal.ensure(16));
    CU(cudaMemcpyAsync(c->fa_raw.p, raw, (size_t)n, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemsetAsync(c->fa_scal.p, 0xff, 8, c->stream));
    uint64_t *scal = c->fa_scal.as<uint64_t>();
    fasta_count_kernel<<<(unsigned)nt, FA_THREADS, 0, c->stream>>>(c->fa_raw.as<unsigned char>(), n, c->fa_tile_kept.as<uint32_t>(),
                                                                  (unsigned long long *)scal);
    if (int rc = scan_u32(c, c->fa_tile_kept.as<uint32_t>(), nt, c->fa_tile_off.as<uint64_t>(), scal + 1)) return rc;
    uint64_t host[2] = {0, 0};
    CU(read_words(c, host, scal, 2));
    if (host[0] != ~0ull) {  // bioioC.c:196-199
        char msg[96];
        snprintf(msg, sizeof msg, "!!Got an unexpected character in input fasta sequence: '%c' ", raw[host[0]]);
        return fail(c, PAFGPU_EFASTA, msg);
    }
    fasta_compact_kernel<<<(unsigned)nt, FA_THREADS, 0, c->stream>>>(c->fa_raw.as<unsigned char>(), n, c->fa_tile_off.as<uint64_t>(),
                                                                    c->in_seq.as<unsigned char>());
    CU(cudaGetLastError());
    const int64_t id = genome_add_device(c, c->G(), name.c_str(), c->in_seq.as<uint8_t>(), (int64_t)host[1]);
    return id < 0 ? (int)id : 0;
}

int pafgpu_genome_load_fasta(pafgpu_ctx *c, const char *path) {
    if (!c || !path) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    // the file as one byte range: mapped when it is a regular file, read whole otherwise (pipes)
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return fail(c, PAFGPU_EIO, std::string("cannot open ") + path);
    struct stat sb;
    const char *d = nullptr;
    size_t n = 0;
    void *map = MAP_FAILED;
    std::vector<char> whole;
    if (fstat(fd, &sb) == 0 && S_ISREG(sb.st_mode) && sb.st_size > 0) {
        map = mmap(nullptr, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (map != MAP_FAILED) { d = (const char *)map; n = (size_t)sb.st_size; }
    }
    if (map == MAP_FAILED) {
        char tmp[1 << 16];
        ssize_t got;
        while ((got = read(fd, tmp, sizeof tmp)) > 0) whole.insert(whole.end(), tmp, tmp + got);
        d = whole.data(); n = whole.size();
    }
    close(fd);
    // Record boundaries with the rules of bioioC.c:155-209: everything before the first '>' is skipped; the header is the
    // whole line (a '>' inside it belongs to the name); a '>' anywhere after it starts the next record; EOF right after
    // a header still yields an (empty) contig.
    int rc = 0;
    const char *p = n ? (const char *)memchr(d, '>', n) : nullptr;
    const char *end = d + n;
    while (p && rc == 0) {
        const char *hs = p + 1;
        const char *nl = hs < end ? (const char *)memchr(hs, '\n', (size_t)(end - hs)) : nullptr;
        const std::string name(hs, nl ? nl : end);
        const char *sb0 = nl ? nl + 1 : end;
        const char *nx = sb0 < end ? (const char *)memchr(sb0, '>', (size_t)(end - sb0)) : nullptr;
        rc = fasta_add_record(c, name, sb0, (int64_t)((nx ? nx : end) - sb0));
        p = nx;
    }
    const cudaError_t se = stream_sync(c);  // the mapping goes away below; every copy out of it must have happened
    if (map != MAP_FAILED) munmap(map, (size_t)sb.st_size);
    if (rc == 0 && se != cudaSuccess) return fail(c, PAFGPU_ENODEVICE, std::string("FASTA load: ") + cudaGetErrorString(se));
    return rc;
}

int64_t pafgpu_genome_find(const pafgpu_ctx *c, const char *name) {
    if (!c || !name) return PAFGPU_EINVAL;
    auto it = c->G().index.find(name);
    return it == c->G().index.end() ? (int64_t)PAFGPU_ENOCONTIG : it->second;
}

int64_t pafgpu_genome_contig_length(const pafgpu_ctx *c, int64_t id) {
    if (!c || id < 0 || id >= (int64_t)c->G().len.size()) return PAFGPU_ENOCONTIG;
    return c->G().len[(size_t)id];
}

int64_t pafgpu_genome_num_contigs(const pafgpu_ctx *c) { return c ? (int64_t)c->G().index.size() : PAFGPU_EINVAL; }

int pafgpu_genome_fetch(pafgpu_ctx *c, int64_t id, int64_t start, int64_t len, char *out) {
    if (!c || !out) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    if (id < 0 || id >= (int64_t)c->G().len.size()) return fail(c, PAFGPU_ENOCONTIG, "bad contig id");
    if (start < 0 || len < 0 || start + len > c->G().len[(size_t)id]) return fail(c, PAFGPU_EBOUNDS, "range outside contig");
    if (len == 0) return 0;
    CU(cudaSetDevice(c->device));
    CU(c->in_seq.ensure((size_t)len));
    unpack_kernel<<<blocks_for(len, 256), 256, 0, c->stream>>>(c->G().view(), (uint64_t)(c->G().off[(size_t)id] + start), len,
                                                               c->in_seq.as<uint8_t>());
    CU(cudaMemcpyAsync(out, c->in_seq.p, (size_t)len, cudaMemcpyDeviceToHost, c->stream));
    CU(stream_sync(c));
    return 0;
}


// ---- genome replication and persistence ----------------------------------------------------------------------------
// Bytes of each arena array in use (everything behind next_off is zero / not consulted).
namespace {
struct ArenaBytes { size_t words, flags, rank, exc; };
ArenaBytes arena_bytes(const GenomeStore &g) {
    return ArenaBytes{(size_t)(g.next_off / 32 + 4) * 8, (size_t)(g.next_off / 1024 + 2) * 4, (size_t)(g.next_off / 1024 + 2) * 4,
                      std::max<size_t>((size_t)g.exc_total * 32, 32)};
}
// a fresh store on c's device holding the host-side tables of src and zeroed arrays of the same capacity
int genome_shell(pafgpu_ctx *c, const GenomeStore &src, std::shared_ptr<GenomeStore> &out) {
    auto g = std::make_shared<GenomeStore>();
    g->device = c->device;
    g->cap_bases = 0;
    g->next_off = src.next_off; g->exc_total = src.exc_total;
    g->names = src.names; g->off = src.off; g->off_rc = src.off_rc; g->len = src.len; g->index = src.index;
    g->table_dirty = true;
    if (int rc = genome_grow(c, *g, std::max<int64_t>(src.next_off + ARENA_ALIGN, ARENA_ALIGN))) return rc;
    CU(g->exc.ensure(arena_bytes(src).exc));
    out = g;
    return 0;
}
}  // namespace

// The packed genome of `src` copied GPU to GPU into `dst` (cudaMemcpyPeerAsync: NVLink / NVSwitch when peer access is
// available, staged through the host otherwise): multi-GPU runs pack the FASTA once, not once per GPU.
int pafgpu_genome_clone(pafgpu_ctx *dst, pafgpu_ctx *src) {
    if (!dst || !src) return PAFGPU_EINVAL;
    if (dst == src || dst->genome_ref == src->genome_ref) return 0;
    pafgpu_ctx *c = dst;
    std::lock_guard<std::mutex> lk(dst->mu);
    std::shared_ptr<GenomeStore> from = src->genome_ref;
    CU(cudaSetDevice(dst->device));
    if (dst->device == from->device) { dst->genome_ref = from; return 0; }  // same GPU: share, do not copy
    if (cudaDeviceEnablePeerAccess(from->device, 0) != cudaSuccess) cudaGetLastError();  // already enabled / not possible: the copy still works
    std::shared_ptr<GenomeStore> g;
    if (int rc = genome_shell(dst, *from, g)) return rc;
    const ArenaBytes nb = arena_bytes(*from);
    CU(cudaMemcpyPeerAsync(g->words.p, dst->device, from->words.p, from->device, nb.words, dst->stream));
    CU(cudaMemcpyPeerAsync(g->flags.p, dst->device, from->flags.p, from->device, nb.flags, dst->stream));
    CU(cudaMemcpyPeerAsync(g->rank.p, dst->device, from->rank.p, from->device, nb.rank, dst->stream));
    if (from->exc.p) CU(cudaMemcpyPeerAsync(g->exc.p, dst->device, from->exc.p, from->device, std::min(nb.exc, from->exc.cap), dst->stream));
    CU(stream_sync(dst));
    dst->genome_ref = g;
    return 0;
}

// k-th helper context of `parent` on `device` (k >= 1 on the parent's own device, k >= 0 elsewhere), created on first use
// and kept until the parent is destroyed; its genome is the parent's (shared on the same GPU, cloned GPU to GPU - again
// only when the parent's contig set changed - on another).
int pafgpu_helper(pafgpu_ctx *parent, int device, int k, pafgpu_ctx **out) {
    if (!parent || !out) return PAFGPU_EINVAL;
    *out = nullptr;
    pafgpu_ctx::Helper *h = nullptr;
    for (auto &x : parent->helpers)
        if (x.device == device && x.k == k) h = &x;
    if (!h) {
        pafgpu_ctx *c = nullptr;
        if (int rc = pafgpu_create(device, &c)) return rc;
        parent->helpers.push_back(pafgpu_ctx::Helper{device, k, c, nullptr, 0});
        h = &parent->helpers.back();
    }
    const GenomeStore *want = parent->genome_ref.get();
    if (device == parent->device) {
        h->ctx->genome_ref = parent->genome_ref;
    } else if (h->genome_of != want || h->generation != want->generation) {
        // another helper on that device may hold the clone already
        pafgpu_ctx *holder = nullptr;
        for (auto &x : parent->helpers)
            if (x.device == device && x.ctx != h->ctx && x.genome_of == want && x.generation == want->generation) holder = x.ctx;
        if (holder) h->ctx->genome_ref = holder->genome_ref;
        else if (int rc = pafgpu_genome_clone(h->ctx, parent)) {
            pafgpu_internal_fail(parent, rc, pafgpu_last_error(h->ctx));
            return rc;
        }
    }
    h->genome_of = want;
    h->generation = want->generation;
    *out = h->ctx;
    return 0;
}

// Packed-genome file (the FASTA load is repeated by every process of a pipeline; the packed arena is
// half the size of the FASTA text and loads at copy speed). Layout: magic, key, counters, contig table, the four arrays.
namespace {
const char GENOME_MAGIC[8] = {'P', 'A', 'F', 'B', '2', 'G', '0', '1'};
bool put(FILE *f, const void *p, size_t n) { return n == 0 || fwrite(p, 1, n, f) == n; }
bool get(FILE *f, void *p, size_t n) { return n == 0 || fread(p, 1, n, f) == n; }
}  // namespace

int pafgpu_genome_save(pafgpu_ctx *c, const char *path, const char *key) {
    if (!c || !path) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    GenomeStore &g = c->G();
    const std::string tmp = std::string(path) + ".tmp";
    FILE *f = fopen(tmp.c_str(), "wb");
    if (!f) return fail(c, PAFGPU_EIO, std::string("cannot write ") + tmp);
    const std::string k = key ? key : "";
    const ArenaBytes nb = arena_bytes(g);
    const uint64_t head[8] = {(uint64_t)k.size(), (uint64_t)g.next_off, g.exc_total, (uint64_t)g.names.size(), nb.words, nb.flags, nb.rank, nb.exc};
    bool ok = put(f, GENOME_MAGIC, 8) && put(f, head, sizeof head) && put(f, k.data(), k.size());
    for (size_t i = 0; ok && i < g.names.size(); i++) {
        const uint64_t row[4] = {(uint64_t)g.names[i].size(), (uint64_t)g.off[i], (uint64_t)g.off_rc[i], (uint64_t)g.len[i]};
        ok = put(f, row, sizeof row) && put(f, g.names[i].data(), g.names[i].size());
    }
    // arrays: device -> pinned staging -> file, a slice at a time
    const size_t slice = (size_t)64 << 20;
    void *stage = pafgpu_host_alloc((int64_t)slice);
    if (!stage) { fclose(f); return fail(c, PAFGPU_ENODEVICE, "pinned host allocation failed"); }
    const DevBuf *arr[4] = {&g.words, &g.flags, &g.rank, &g.exc};
    const size_t bytes[4] = {nb.words, nb.flags, nb.rank, nb.exc};
    cudaError_t ce = cudaSuccess;
    for (int a = 0; ok && a < 4 && ce == cudaSuccess; a++)
        for (size_t at = 0; ok && at < bytes[a] && ce == cudaSuccess; at += slice) {
            const size_t n = std::min(slice, bytes[a] - at);
            if (arr[a]->p && at + n <= arr[a]->cap) {
                ce = cudaMemcpyAsync(stage, (const char *)arr[a]->p + at, n, cudaMemcpyDeviceToHost, c->stream);
                if (ce == cudaSuccess) ce = stream_sync(c);
            } else {
                memset(stage, 0, n);
            }
            ok = put(f, stage, n);
        }
    pafgpu_host_free(stage);
    ok = (fclose(f) == 0) && ok;
    if (ce != cudaSuccess) { remove(tmp.c_str()); return fail(c, PAFGPU_ENODEVICE, std::string("genome save: ") + cudaGetErrorString(ce)); }
    if (!ok || rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); return fail(c, PAFGPU_EIO, std::string("cannot write ") + path); }
    return 0;
}

int pafgpu_genome_load_packed(pafgpu_ctx *c, const char *path, const char *key) {
    if (!c || !path) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    FILE *f = fopen(path, "rb");
    if (!f) return fail(c, PAFGPU_EIO, std::string("cannot open ") + path);
    char magic[8];
    uint64_t head[8];
    const std::string want = key ? key : "";
    std::string k;
    bool ok = get(f, magic, 8) && memcmp(magic, GENOME_MAGIC, 8) == 0 && get(f, head, sizeof head) && head[0] < (1u << 20);
    if (ok) { k.resize((size_t)head[0]); ok = get(f, &k[0], k.size()); }
    if (!ok || k != want) { fclose(f); return fail(c, PAFGPU_ESTATE, ok ? "packed genome was made from other inputs (key differs)" : "not a packed genome file"); }
    GenomeStore src;  // host tables only
    src.device = c->device;
    src.next_off = (int64_t)head[1]; src.exc_total = (uint32_t)head[2];
    for (uint64_t i = 0; ok && i < head[3]; i++) {
        uint64_t row[4];
        ok = get(f, row, sizeof row) && row[0] < (1u << 20);
        if (!ok) break;
        std::string name((size_t)row[0], '\0');
        ok = get(f, &name[0], name.size());
        src.names.push_back(name); src.off.push_back((int64_t)row[1]); src.off_rc.push_back((int64_t)row[2]); src.len.push_back((int64_t)row[3]);
        src.index[name] = (int64_t)i;  // a repeated name points at the newest contig
    }
    std::shared_ptr<GenomeStore> g;
    if (ok) {
        const ArenaBytes nb = arena_bytes(src);
        ok = nb.words == head[4] && nb.flags == head[5] && nb.rank == head[6] && nb.exc == head[7];
    }
    if (!ok) { fclose(f); return fail(c, PAFGPU_EIO, "packed genome file is damaged"); }
    if (int rc = genome_shell(c, src, g)) { fclose(f); return rc; }
    const size_t slice = (size_t)64 << 20;
    char *stage[2] = {(char *)pafgpu_host_alloc((int64_t)slice), (char *)pafgpu_host_alloc((int64_t)slice)};
    cudaEvent_t done[2] = {nullptr, nullptr};
    cudaError_t ce = (stage[0] && stage[1]) ? cudaSuccess : cudaErrorMemoryAllocation;
    for (int i = 0; i < 2 && ce == cudaSuccess; i++) ce = cudaEventCreateWithFlags(&done[i], cudaEventDisableTiming);
    DevBuf *arr[4] = {&g->words, &g->flags, &g->rank, &g->exc};
    int turn = 0;
    for (int a = 0; ok && a < 4 && ce == cudaSuccess; a++)
        for (size_t at = 0; ok && at < head[4 + a] && ce == cudaSuccess; at += slice, turn ^= 1) {
            const size_t n = std::min<size_t>(slice, head[4 + a] - at);
            ce = cudaEventSynchronize(done[turn]);  // the copy that last used this staging buffer
            if (ce != cudaSuccess) break;
            ok = get(f, stage[turn], n);
            if (!ok) break;
            ce = cudaMemcpyAsync((char *)arr[a]->p + at, stage[turn], n, cudaMemcpyHostToDevice, c->stream);
            if (ce == cudaSuccess) ce = cudaEventRecord(done[turn], c->stream);
        }
    if (ce == cudaSuccess) ce = stream_sync(c);
    for (int i = 0; i < 2; i++) { if (done[i]) cudaEventDestroy(done[i]); pafgpu_host_free(stage[i]); }
    fclose(f);
    if (ce != cudaSuccess) return fail(c, PAFGPU_ENODEVICE, std::string("packed genome load: ") + cudaGetErrorString(ce));
    if (!ok) return fail(c, PAFGPU_EIO, "packed genome file is truncated");
    c->genome_ref = g;
    return 0;
}

// ---- batch encode ------------------------------------------------------------------------------
int pafgpu_encode_batch_host(pafgpu_ctx *c, const pafgpu_batch *b, pafgpu_result *res) {
    if (!c) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    return encode_host(c, c->G(), b, res);
}

int pafgpu_encode_batch_device(pafgpu_ctx *c, const pafgpu_batch *b, pafgpu_result *res) {
    if (!c) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    CU(cudaDeviceSynchronize());  // inputs may have been produced on another stream
    return encode_device(c, c->G(), b, res);
}

int pafgpu_fetch_ops(pafgpu_ctx *c, uint64_t *out_ops, int64_t *out_op_begin, int64_t *rec_mismatch) {
    if (!c) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    return fetch_ops(c, out_ops, out_op_begin, rec_mismatch);
}

int pafgpu_result_device_ptrs(pafgpu_ctx *c, const uint64_t **out_ops, const int64_t **out_op_begin, const int64_t **rec_mismatch) {
    if (!c) return PAFGPU_EINVAL;
    if (!c->have_batch) return fail(c, PAFGPU_ESTATE, "no batch has been encoded");
    if (out_ops) *out_ops = c->out.as<uint64_t>();
    if (out_op_begin) *out_op_begin = c->rec_out_begin.as<int64_t>();
    if (rec_mismatch) *rec_mismatch = c->rec_mism.as<int64_t>();
    return 0;
}

int pafgpu_time_encode(pafgpu_ctx *c, int reps, float *ms_stage, float *ms_kernel) {
    if (!c || reps <= 0) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->have_batch || c->staged_only) return fail(c, PAFGPU_ESTATE, "no batch has been encoded");
    CU(cudaSetDevice(c->device));
    double sum = 0, ksum = 0;
    for (int i = 0; i < reps; i++) {
        CU(cudaEventRecord(c->ev[2], c->stream));
        if (int rc = launch_encode(c, c->aligned)) return rc;
        CU(cudaEventRecord(c->ev[3], c->stream));
        CU(stream_sync(c));
        float ms = 0;
        cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]);
        sum += ms;
        cudaEventElapsedTime(&ms, c->ev[4], c->ev[5]); ksum += ms;
    }
    if (ms_stage) *ms_stage = (float)(sum / reps);
    if (ms_kernel) *ms_kernel = (float)(ksum / reps);
    return 0;
}

// Load balance of the last encode_kernel launch ("max / mean per-SM busy time"): every persistent CTA leaves the
// nanoseconds between its first and its last tile and the SM it ran on; an SM's busy time is the longest of its CTAs.
int pafgpu_encode_balance(pafgpu_ctx *c, double *max_ms, double *mean_ms, int *sms_used) {
    if (!c || !max_ms || !mean_ms) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->have_batch || c->last_grid == 0) return fail(c, PAFGPU_ESTATE, "no batch has been encoded");
    CU(cudaSetDevice(c->device));
    std::vector<unsigned long long> span((size_t)c->last_grid * 2);
    CU(cudaMemcpyAsync(span.data(), c->cta_span.p, span.size() * 8, cudaMemcpyDeviceToHost, c->stream));
    CU(stream_sync(c));
    std::unordered_map<unsigned long long, unsigned long long> per_sm;
    for (unsigned b = 0; b < c->last_grid; b++) per_sm[span[2 * b + 1]] = std::max(per_sm[span[2 * b + 1]], span[2 * b]);
    double mx = 0, sum = 0;
    for (auto &kv : per_sm) { mx = std::max(mx, (double)kv.second); sum += (double)kv.second; }
    *max_ms = mx * 1e-6;
    *mean_ms = per_sm.empty() ? 0 : sum / (double)per_sm.size() * 1e-6;
    if (sms_used) *sms_used = (int)per_sm.size();
    return 0;
}

// ---- one record with raw host strings (paf_encode_mismatches, impl/paf.c:739-784) ----------------
int pafgpu_encode_one(pafgpu_ctx *c, const uint64_t *ops, int64_t n_ops, int same_strand, int64_t query_start, int64_t query_end,
                      int64_t target_start, const char *query_seq, const char *target_seq, uint64_t **out_ops, int64_t *n_out,
                      int64_t *mismatches) {
    if (!c || !ops || n_ops < 0 || !out_ops || !n_out || !query_seq || !target_seq) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    // spans the record touches; zero-length M ops vanish (a1-edge ii)
    std::vector<uint64_t> kept;
    kept.reserve((size_t)n_ops);
    int64_t t_span = 0, q_span = 0;
    for (int64_t k = 0; k < n_ops; k++) {
        const int code = (int)(ops[k] >> 56);
        const int64_t len = (int64_t)(ops[k] & LEN_MASK);
        if (code == OP_M && len == 0) continue;
        kept.push_back(ops[k]);
        if (code != OP_I) t_span += len;
        if (code != OP_D) q_span += len;
    }
    *out_ops = nullptr;
    *n_out = 0;
    if (mismatches) *mismatches = 0;
    if (kept.empty()) { *out_ops = (uint64_t *)malloc(8); return 0; }
    const int64_t q_lo = same_strand ? query_start : query_end - q_span;
    if (q_lo < 0 || target_start < 0) return fail(c, PAFGPU_EBOUNDS, "record walks before the start of a sequence");
    GenomeStore &g = c->scratch;
    g.reset();
    const int64_t tid = genome_add_host(c, g, "t", target_seq + target_start, t_span);
    if (tid < 0) return (int)tid;
    const int64_t qid = genome_add_host(c, g, "q", query_seq + q_lo, q_span);
    if (qid < 0) return (int)qid;
    const int64_t op_begin[2] = {0, (int64_t)kept.size()};
    const int32_t qc = (int32_t)qid, tc = (int32_t)tid;
    const int64_t qs = 0, qe = q_span, ts = 0;
    const uint8_t strand = same_strand ? 1 : 0;
    pafgpu_batch b{1, (int64_t)kept.size(), op_begin, kept.data(), &qc, &tc, &qs, &qe, &ts, &strand};
    pafgpu_result res;
    if (int rc = encode_host(c, g, &b, &res)) return rc;
    uint64_t *o = (uint64_t *)malloc((size_t)std::max<int64_t>(res.n_out_ops, 1) * 8);
    int64_t mm = 0;
    if (int rc = fetch_ops(c, o, nullptr, &mm)) { free(o); return rc; }
    *out_ops = o;
    *n_out = res.n_out_ops;
    if (mismatches) *mismatches = mm;
    return 0;
}

}  // extern "C"

// ---- PAF text in, PAF text out, all of it on the device -------------------------------------------
namespace {

uint64_t host_name_hash(const std::string &s) {  // = name_hash (paftext.cuh)
    uint64_t h = 1469598103934665603ull;
    for (unsigned char ch : s) { h ^= ch; h *= 1099511628211ull; }
    return h ? h : 1;
}

// name -> contig id table of the genome on the device (a repeated name points at the newest contig, as the host map does)
int genome_sync_names(pafgpu_ctx *c, GenomeStore &g) {
    std::lock_guard<std::mutex> lk(g.mu);
    if (!g.names_dirty) return 0;
    uint32_t cap = 16;
    while (cap < 2 * g.index.size() + 2) cap *= 2;
    std::vector<NameEntry> tab(cap, NameEntry{0, 0, 0, -1, 0});
    std::string blob;
    for (const auto &kv : g.index) {
        const uint64_t h = host_name_hash(kv.first);
        uint32_t slot = (uint32_t)h & (cap - 1);
        while (tab[slot].hash) slot = (slot + 1) & (cap - 1);
        tab[slot] = NameEntry{h, (uint32_t)blob.size(), (uint32_t)kv.first.size(), (int32_t)kv.second, 0};
        blob += kv.first;
    }
    CU(g.d_names.ensure((size_t)cap * sizeof(NameEntry)));
    CU(g.d_name_blob.ensure(std::max<size_t>(blob.size(), 16)));
    CU(cudaMemcpyAsync(g.d_names.p, tab.data(), (size_t)cap * sizeof(NameEntry), cudaMemcpyHostToDevice, c->stream));
    if (!blob.empty()) CU(cudaMemcpyAsync(g.d_name_blob.p, blob.data(), blob.size(), cudaMemcpyHostToDevice, c->stream));
    CU(stream_sync(c));
    g.name_mask = cap - 1;
    g.names_dirty = false;
    return 0;
}

// carve typed arrays out of one device buffer
struct Carver {
    char *base;
    size_t at = 0;
    template <class T> T *take(size_t n) {
        at = (at + 15) & ~(size_t)15;
        T *p = reinterpret_cast<T *>(base + at);
        at += n * sizeof(T);
        return p;
    }
};

}  // namespace

extern "C" int pafgpu_add_mismatches_text(pafgpu_ctx *c, const char *text, int64_t n_bytes, pafgpu_text_result *res) {
    return pafgpu_add_mismatches_text_opts(c, text, n_bytes, nullptr, res);
}

// The text of the staged output ops (c->out, c->rec_out_begin; c->ser_prefix_begin holds the prefix offsets): one pass
// writes every op and line_begin (ser_write_kernel), `prefixes` then launches the kernel that renders the line heads,
// and the byte total comes back. The text buffer is sized from `guess` beforehand; the kernels write nothing beyond it,
// so a guess that was too small costs one more run at the exact size.
template <class F>
static int lay_out_and_write(pafgpu_ctx *c, int64_t R, int64_t K, SerialiseParams &S, int64_t guess, int64_t *total_out, F prefixes) {
    const int64_t nb = blocks_for(K, SER_OPS_PER_BLOCK);
    CU(c->ser_optoff.ensure((size_t)(nb + 2) * 8));
    CU(c->ser_oplen.ensure((size_t)(nb + 1) * 4));
    CU(c->ser_line_begin.ensure((size_t)(R + 1) * 8));
    if (K > 0) ser_block_records_kernel<<<blocks_for(R, 256), 256, 0, c->stream>>>(R, c->rec_out_begin.as<int64_t>(), c->ser_oplen.as<uint32_t>());
    S.block_rec = c->ser_oplen.as<uint32_t>();
    S.n_records = R; S.n_out = K;
    S.out = c->out.as<uint64_t>(); S.rec_out_begin = c->rec_out_begin.as<int64_t>();
    S.status = c->ser_optoff.as<unsigned long long>();
    S.prefix_begin = c->ser_prefix_begin.as<int64_t>();
    S.line_begin = c->ser_line_begin.as<int64_t>();
    S.keep_lo = c->trimmed ? c->trim_lo.as<int64_t>() : nullptr;
    S.keep_hi = c->trimmed ? c->trim_hi.as<int64_t>() : nullptr;
    int64_t total = 0;
    for (int attempt = 0;; attempt++) {
        CU(c->ser_text.ensure((size_t)std::max<int64_t>(guess, 8)));
        S.text = c->ser_text.as<char>(); S.text_cap = (int64_t)c->ser_text.cap;
        if (K > 0) {
            CU(cudaMemsetAsync(c->ser_optoff.p, 0, (size_t)(nb + 2) * 8, c->stream));
            ser_write_kernel<<<(unsigned)nb, SER_THREADS, 0, c->stream>>>(S);
        } else {
            line_begin_no_ops_kernel<<<blocks_for(R + 1, 256), 256, 0, c->stream>>>(R, S.prefix_begin, S.line_begin);
        }
        prefixes(S);
        CU(read_words(c, (uint64_t *)&total, c->ser_line_begin.as<uint64_t>() + R, 1));
        if (total <= S.text_cap) break;
        if (attempt) return fail(c, PAFGPU_ESTATE, "serialiser: the text buffer is still too small after growing it");
        guess = total;
    }
    *total_out = total;
    return 0;
}

extern "C" int pafgpu_add_mismatches_text_opts(pafgpu_ctx *c, const char *text, int64_t n_bytes, const pafgpu_text_opts *opts,
                                               pafgpu_text_result *res) {
    if (!c || !res || n_bytes < 0 || (n_bytes && !text)) return PAFGPU_EINVAL;
    const int mode = opts ? opts->mode : PAFGPU_MODE_ADD;
    if (mode != PAFGPU_MODE_ADD && mode != PAFGPU_MODE_ADD_TRIM && mode != PAFGPU_MODE_VIEW) return fail(c, PAFGPU_EINVAL, "text mode: add, add + trim, or view totals");
    if (n_bytes >= ((int64_t)1 << 31)) return fail(c, PAFGPU_EINVAL, "text chunk too large: cut it (32-bit byte offsets)");
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    memset(res, 0, sizeof *res);
    res->first_bad = -1;
    c->have_batch = false;
    c->ser_text_bytes = 0;
    if (n_bytes == 0) return 0;
    GenomeStore &g = c->G();
    if (int rc = genome_sync_names(c, g)) return rc;
    if (int rc = genome_sync_table(c, g)) return rc;
    // the text ends with a newline on the device (a last line without one still counts, sonLibFile.c:42-71)
    const bool add_nl = text[n_bytes - 1] != '\n';
    const int64_t n = n_bytes + (add_nl ? 1 : 0);
    const int64_t n_tiles_tx = (n + TX_TILE - 1) / TX_TILE;
    CU(c->tx_text.ensure((size_t)n + 64));
    CU(c->tx_tile_nl.ensure((size_t)n_tiles_tx * 4));
    CU(c->tx_tile_tab.ensure((size_t)n_tiles_tx * 4));
    CU(c->tx_nl_off.ensure((size_t)(n_tiles_tx + 1) * 8));
    CU(c->tx_tab_off.ensure((size_t)(n_tiles_tx + 1) * 8));
    CU(c->tx_tile_ops.ensure((size_t)n_tiles_tx * 4 + (size_t)n_tiles_tx * TX_THREADS));  // ops per tile, and behind them ops per thread
    CU(c->tx_op_off.ensure((size_t)(n_tiles_tx + 1) * 8));
    CU(c->tx_scal.ensure(32 * 8));
    CU(c->scalars.ensure(SC_COUNT * 8));
    char *d_text = c->tx_text.as<char>();
    uint64_t *scal = c->tx_scal.as<uint64_t>();  // [0] first bad record, [1] newlines, [2] tabs, [3] ops, [4..11] totals, [12] mismatches
    CU(cudaEventRecord(c->ev_tx[0], c->stream));
    CU(cudaMemcpyAsync(d_text, text, (size_t)n_bytes, cudaMemcpyHostToDevice, c->stream));
    if (add_nl) CU(cudaMemsetAsync(d_text + n_bytes, '\n', 1, c->stream));
    CU(cudaMemsetAsync(d_text + n, 0, 32, c->stream));
    CU(cudaMemsetAsync(scal, 0, 32 * 8, c->stream));
    CU(cudaMemsetAsync(scal, 0xff, 8, c->stream));
    tx_count_delims_kernel<<<(unsigned)n_tiles_tx, TX_THREADS, 0, c->stream>>>(d_text, n, c->tx_tile_nl.as<uint32_t>(), c->tx_tile_tab.as<uint32_t>());
    if (int rc = scan_u32(c, c->tx_tile_nl.as<uint32_t>(), n_tiles_tx, c->tx_nl_off.as<uint64_t>(), scal + 1)) return rc;
    if (int rc = scan_u32(c, c->tx_tile_tab.as<uint32_t>(), n_tiles_tx, c->tx_tab_off.as<uint64_t>(), scal + 2)) return rc;
    uint64_t counts[2];
    CU(read_words(c, counts, scal + 1, 2));
    int64_t R = (int64_t)counts[0];
    const int64_t n_tabs = (int64_t)counts[1];
    const int64_t R_all = R;
    if (R >= ((int64_t)1 << 31)) return fail(c, PAFGPU_EINVAL, "text chunk holds too many records: cut it");
    CU(c->tx_nl_pos.ensure((size_t)std::max<int64_t>(R, 1) * 4));
    CU(c->tx_tab_pos.ensure((size_t)(n_tabs + 1) * 4));
    CU(c->tx_line_tab0.ensure((size_t)(R + 1) * 4));
    // parsed columns
    const size_t Rz = (size_t)R + 1;
    CU(c->tx_cols.ensure(Rz * (7 * 4 + 13 * 8 + 3 + 2 * 4 + 8 + 4 + 8) + 64 * 16));
    Carver cv{c->tx_cols.as<char>()};
    TextRecords T{};
    T.line_begin = cv.take<uint32_t>(Rz); T.qn_off = cv.take<uint32_t>(Rz); T.qn_len = cv.take<uint32_t>(Rz);
    T.tn_off = cv.take<uint32_t>(Rz); T.tn_len = cv.take<uint32_t>(Rz); T.cg_begin = cv.take<uint32_t>(Rz); T.cg_end = cv.take<uint32_t>(Rz);
    T.qlen = cv.take<int64_t>(Rz); T.qs = cv.take<int64_t>(Rz); T.qe = cv.take<int64_t>(Rz); T.tlen = cv.take<int64_t>(Rz);
    T.ts = cv.take<int64_t>(Rz); T.te = cv.take<int64_t>(Rz); T.nmatch = cv.take<int64_t>(Rz); T.nbases = cv.take<int64_t>(Rz);
    T.mapq = cv.take<int64_t>(Rz); T.score = cv.take<int64_t>(Rz); T.tl = cv.take<int64_t>(Rz); T.cn = cv.take<int64_t>(Rz);
    T.s1 = cv.take<int64_t>(Rz);
    T.type = cv.take<uint8_t>(Rz); T.strand = cv.take<uint8_t>(Rz); T.has_cg = cv.take<uint8_t>(Rz);
    T.q_contig = cv.take<int32_t>(Rz); T.t_contig = cv.take<int32_t>(Rz);
    int64_t *op_begin = cv.take<int64_t>(Rz);
    uint32_t *prefix_len = cv.take<uint32_t>(Rz);
    uint32_t *nl_pos = c->tx_nl_pos.as<uint32_t>(), *tab_pos = c->tx_tab_pos.as<uint32_t>(), *line_tab0 = c->tx_line_tab0.as<uint32_t>();
    CU(cudaMemsetAsync(tab_pos + n_tabs, 0xff, 4, c->stream));  // sentinel behind the last tab
    CU(cudaMemsetAsync(line_tab0, 0, 4, c->stream));
    CU(cudaMemsetAsync(op_begin, 0, 8, c->stream));
    tx_scatter_delims_kernel<<<(unsigned)n_tiles_tx, TX_THREADS, 0, c->stream>>>(d_text, n, c->tx_nl_off.as<uint64_t>(), c->tx_tab_off.as<uint64_t>(),
                                                                                nl_pos, tab_pos, line_tab0);
    ParseParams PP{};
    PP.text = d_text; PP.n_records = R; PP.nl_pos = nl_pos; PP.tab_pos = tab_pos; PP.line_tab0 = line_tab0; PP.R = T;
    PP.names = g.d_names.as<NameEntry>(); PP.name_mask = g.name_mask; PP.name_blob = g.d_name_blob.as<char>();
    PP.first_bad = (unsigned long long *)scal;
    if (R > 0) tx_parse_records_kernel<<<blocks_for(R, 128), 128, 0, c->stream>>>(PP);
    OpsParams OP{};
    OP.text = d_text; OP.n = n; OP.n_records = R; OP.nl_off = c->tx_nl_off.as<uint64_t>(); OP.cg_begin = T.cg_begin; OP.cg_end = T.cg_end;
    OP.tile_ops = c->tx_tile_ops.as<uint32_t>(); OP.op_off = c->tx_op_off.as<uint64_t>(); OP.op_begin = op_begin;
    OP.first_bad = (unsigned long long *)scal;
    OP.thread_ops = reinterpret_cast<uint8_t *>(c->tx_tile_ops.as<uint32_t>() + n_tiles_tx);
    tx_ops_kernel<false><<<(unsigned)n_tiles_tx, TX_THREADS, 0, c->stream>>>(OP);
    if (int rc = scan_u32(c, c->tx_tile_ops.as<uint32_t>(), n_tiles_tx, c->tx_op_off.as<uint64_t>(), scal + 3)) return rc;
    uint64_t after_ops[4];
    CU(read_words(c, after_ops, scal, 4));
    int64_t N = (int64_t)after_ops[3];
    CU(c->tx_ops.ensure((size_t)std::max<int64_t>(N, 1) * 8));
    OP.ops = c->tx_ops.as<uint64_t>();
    tx_ops_kernel<true><<<(unsigned)n_tiles_tx, TX_THREADS, 0, c->stream>>>(OP);
    CU(cudaGetLastError());
    CU(cudaEventRecord(c->ev_tx[1], c->stream));
    int64_t first_bad = after_ops[0] == ~0ull ? -1 : (int64_t)after_ops[0];

    // ---- ingest + encode on the records before the first bad one; the ingest kernel may find an earlier one (a record
    // that walks outside its contig, or fails paf_check's op sums)
    pafgpu_result er{};
    for (int attempt = 0;; attempt++) {
        if (first_bad >= 0 && first_bad < R) {
            R = first_bad;
            CU(cudaMemcpyAsync(&N, op_begin + R, 8, cudaMemcpyDeviceToHost, c->stream));
            CU(stream_sync(c));
        }
        if (R == 0) break;
        pafgpu_batch b{R, N, op_begin, c->tx_ops.as<uint64_t>(), T.q_contig, T.t_contig, T.qs, T.qe, T.ts, T.strand};
        c->ingest_t_end = T.te;
        c->ingest_has_cg = T.has_cg;
        const int rc = encode_device(c, g, &b, &er);
        c->ingest_t_end = nullptr;
        c->ingest_has_cg = nullptr;
        if (rc == PAFGPU_EBOUNDS && c->bad_record >= 0 && c->bad_record < R && attempt < 4) {
            first_bad = c->bad_record;
            continue;
        }
        if (rc) return rc;
        break;
    }
    res->records = R;
    res->records_in = R_all;
    res->first_bad = first_bad >= 0 && first_bad < R_all ? first_bad : -1;
    if (res->first_bad >= 0) {
        uint32_t ends[2] = {0, 0};
        if (res->first_bad > 0) CU(cudaMemcpyAsync(&ends[0], nl_pos + res->first_bad - 1, 4, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaMemcpyAsync(&ends[1], nl_pos + res->first_bad, 4, cudaMemcpyDeviceToHost, c->stream));
        CU(stream_sync(c));
        res->bad_line_begin = res->first_bad > 0 ? (int64_t)ends[0] + 1 : 0;
        res->bad_line_end = std::min<int64_t>(ends[1], n_bytes);
    }
    if (R == 0) { c->have_batch = false; return 0; }
    res->n_ops = N; res->n_out_ops = c->n_out; res->aligned_bases = er.aligned_bases;
    res->ms_ingest = er.ms_ingest; res->ms_kernel = er.ms_kernel; res->ms_encode = er.ms_encode;
    cudaEventElapsedTime(&res->ms_parse, c->ev_tx[0], c->ev_tx[1]);

    // totals for the report: sums over the input ops by kind + the mismatches the kernel tallied
    tx_op_totals_kernel<<<(unsigned)std::min<int64_t>(blocks_for(N, 256 * 8), 1184), 256, 0, c->stream>>>(c->tx_ops.as<uint64_t>(), N,
                                                                                                     (unsigned long long *)(scal + 4));
    tx_sum_kernel<<<(unsigned)std::min<int64_t>(blocks_for(R, 256 * 8), 1184), 256, 0, c->stream>>>((const unsigned long long *)c->rec_mism.p, R,
                                                                                               (unsigned long long *)(scal + 12));
    auto totals = [&](int64_t records) -> int {
        uint64_t tot[9];
        CU(read_words(c, tot, scal + 4, 9));
        const int64_t mm = (int64_t)tot[8];
        res->stats[0] = (int64_t)tot[0] + (int64_t)tot[1] - mm;  // matches: M and = bases that are no mismatches
        res->stats[1] = (int64_t)tot[2] + mm;
        res->stats[2] = (int64_t)tot[3]; res->stats[3] = (int64_t)tot[5];
        res->stats[4] = (int64_t)tot[4]; res->stats[5] = (int64_t)tot[6];
        res->stats[6] = records;
        return 0;
    };
    if (mode == PAFGPU_MODE_VIEW) return totals(R);  // `view -s -t`: nothing is printed per record

    // ---- output text: prefix of every record measured, the lines laid out, prefixes and ops rendered in place
    int64_t K = c->n_out;
    CU(cudaEventRecord(c->ev_tx[2], c->stream));
    tx_empty_out_begin_kernel<<<blocks_for(R, 256), 256, 0, c->stream>>>(R, op_begin, c->op_rec.as<uint32_t>(), N,
                                                                         (const unsigned long long *)(c->scalars.as<uint64_t>() + SC_TOTAL),
                                                                         c->rec_out_begin.as<int64_t>());
    TextRecords TP = T;  // the columns the lines are printed from (trimmed coordinates in add + trim mode)
    const int64_t R_enc = R;
    if (mode == PAFGPU_MODE_ADD_TRIM) {
        // paf_trim_unreliable_tails / paf_trim_end_fraction on the encoded ops (trim.cuh), then the reference's paf_check
        // behind the trim; the first record that fails ends the output
        const bool fixed = opts->trim_score_fraction < 0;
        const size_t rb = (size_t)R * 8;
        CU(c->trim_coords.ensure(4 * rb));
        CU(c->trim_lo.ensure(rb));
        CU(c->trim_hi.ensure(rb));
        CU(c->trim_failed.ensure((size_t)R));
        int64_t *co = c->trim_coords.as<int64_t>();
        CU(cudaMemcpyAsync(co, T.qs, rb, cudaMemcpyDeviceToDevice, c->stream));
        CU(cudaMemcpyAsync(co + R, T.qe, rb, cudaMemcpyDeviceToDevice, c->stream));
        CU(cudaMemcpyAsync(co + 2 * R, T.ts, rb, cudaMemcpyDeviceToDevice, c->stream));
        CU(cudaMemcpyAsync(co + 3 * R, T.te, rb, cudaMemcpyDeviceToDevice, c->stream));
        TrimParams TR{};
        TR.n_records = R; TR.out = c->out.as<uint64_t>(); TR.rec_out_begin = c->rec_out_begin.as<int64_t>();
        TR.same_strand = T.strand;
        TR.qs = co; TR.qe = co + R; TR.ts = co + 2 * R; TR.te = co + 3 * R;
        TR.lo = c->trim_lo.as<int64_t>(); TR.hi = c->trim_hi.as<int64_t>(); TR.failed = c->trim_failed.as<uint8_t>();
        TR.score_fraction = opts->trim_score_fraction; TR.max_fraction = opts->trim_max_fraction;
        bool by_warp = K / R > 64;
        if (const char *e = getenv("PAFGPU_TRIM_WARP")) by_warp = atoi(e) != 0;
        if (fixed) trim_fixed_kernel<<<blocks_for(R, 128), 128, 0, c->stream>>>(TR, c->out.as<uint64_t>());
        else if (by_warp) trim_warp_kernel<<<blocks_for(R * 32, 128), 128, 0, c->stream>>>(TR);
        else trim_kernel<<<blocks_for(R, 128), 128, 0, c->stream>>>(TR);
        CU(cudaMemsetAsync(scal + 16, 0xff, 8, c->stream));
        tx_trim_check_kernel<<<blocks_for(R, 256), 256, 0, c->stream>>>(R, TR.qs, TR.qe, TR.ts, TR.te, T.qlen, T.tlen, op_begin, TR.failed,
                                                                        (unsigned long long *)(scal + 16));
        uint64_t bad2 = 0;
        CU(read_words(c, &bad2, scal + 16, 1));
        TP.qs = TR.qs; TP.qe = TR.qe; TP.ts = TR.ts; TP.te = TR.te;
        c->trimmed = true;
        if (bad2 != ~0ull && (int64_t)bad2 < R) {
            // the run ends at that record: its moved coordinates go back for the message
            R = (int64_t)bad2;
            uint8_t fl = 0;
            uint32_t ends[2] = {0, 0};
            for (int q = 0; q < 4; q++) CU(cudaMemcpyAsync(&res->bad_coords[q], co + q * R_enc + R, 8, cudaMemcpyDeviceToHost, c->stream));
            CU(cudaMemcpyAsync(&fl, TR.failed + R, 1, cudaMemcpyDeviceToHost, c->stream));
            CU(cudaMemcpyAsync(&K, c->rec_out_begin.as<int64_t>() + R, 8, cudaMemcpyDeviceToHost, c->stream));
            if (R > 0) CU(cudaMemcpyAsync(&ends[0], nl_pos + R - 1, 4, cudaMemcpyDeviceToHost, c->stream));
            CU(cudaMemcpyAsync(&ends[1], nl_pos + R, 4, cudaMemcpyDeviceToHost, c->stream));
            CU(stream_sync(c));
            res->first_bad = R;
            int64_t ob[2] = {0, 1};
            CU(cudaMemcpy(ob, op_begin + R, 16, cudaMemcpyDeviceToHost));
            res->bad_kind = (fl || ob[1] == ob[0]) ? 2 : 1;
            res->bad_line_begin = R > 0 ? (int64_t)ends[0] + 1 : 0;
            res->bad_line_end = std::min<int64_t>(ends[1], n_bytes);
            res->records = R;
            if (R == 0) { res->text_bytes = 0; c->ser_text_bytes = 0; return totals(R_enc); }
        }
    } else {
        c->trimmed = false;
    }
    PrefixParams PR{};
    PR.text = d_text; PR.n_records = R; PR.R = TP; PR.prefix_len = prefix_len;
    tx_prefix_kernel<false><<<blocks_for(R, TXP_THREADS), TXP_THREADS, 0, c->stream>>>(PR);
    CU(c->ser_prefix_begin.ensure((size_t)(R + 1) * 8));
    if (int rc = scan_u32(c, prefix_len, R, c->ser_prefix_begin.as<uint64_t>(), nullptr)) return rc;
    SerialiseParams S{};
    S.prefix = nullptr;
    int64_t total = 0;
    // first guess of the output size: 4 bytes per op (C2 averages 3.1) + line heads about as long as the input's
    const int64_t guess = 4 * K + 2 * n_bytes + R + 4096;
    if (int rc = lay_out_and_write(c, R, K, S, guess, &total, [&](const SerialiseParams &L) {
            PR.line_begin = L.line_begin; PR.out = L.text; PR.out_cap = L.text_cap;
            if (R > 0) tx_prefix_kernel<true><<<blocks_for(R, TXP_THREADS), TXP_THREADS, 0, c->stream>>>(PR);
        }))
        return rc;
    CU(cudaEventRecord(c->ev_tx[3], c->stream));
    if (int rc = totals(R_enc)) return rc;
    CU(cudaGetLastError());
    cudaEventElapsedTime(&res->ms_render, c->ev_tx[2], c->ev_tx[3]);
    res->text_bytes = total;
    c->ser_text_bytes = total;
    return 0;
}

extern "C" {

// One record against contigs that are already packed in the context's genome (the cache behind paf_encode_mismatches):
// a batch of one through the normal path. out_ops is malloc'd (caller frees).
int pafgpu_encode_one_resident(pafgpu_ctx *c, const uint64_t *ops, int64_t n_ops, int same_strand, int64_t query_start, int64_t query_end,
                               int64_t target_start, int64_t q_contig, int64_t t_contig, uint64_t **out_ops, int64_t *n_out, int64_t *mismatches) {
    if (!c || !ops || n_ops < 0 || !out_ops || !n_out) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    std::vector<uint64_t> kept;
    kept.reserve((size_t)n_ops);
    for (int64_t k = 0; k < n_ops; k++)
        if (!((int)(ops[k] >> 56) == OP_M && (ops[k] & LEN_MASK) == 0)) kept.push_back(ops[k]);  // zero-length M ops vanish (a1-edge ii)
    *out_ops = nullptr;
    *n_out = 0;
    if (mismatches) *mismatches = 0;
    if (kept.empty()) { *out_ops = (uint64_t *)malloc(8); return 0; }
    const int64_t op_begin[2] = {0, (int64_t)kept.size()};
    const int32_t qc = (int32_t)q_contig, tc = (int32_t)t_contig;
    const uint8_t strand = same_strand ? 1 : 0;
    pafgpu_batch b{1, (int64_t)kept.size(), op_begin, kept.data(), &qc, &tc, &query_start, &query_end, &target_start, &strand};
    pafgpu_result res;
    if (int rc = encode_host(c, c->G(), &b, &res)) return rc;
    uint64_t *o = (uint64_t *)malloc((size_t)std::max<int64_t>(res.n_out_ops, 1) * 8);
    int64_t mm = 0;
    if (int rc = fetch_ops(c, o, nullptr, &mm)) { free(o); return rc; }
    *out_ops = o;
    *n_out = res.n_out_ops;
    if (mismatches) *mismatches = mm;
    return 0;
}

// forget every contig (the arena keeps its capacity)
int pafgpu_genome_clear(pafgpu_ctx *c) {
    if (!c) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    CU(stream_sync(c));
    GenomeStore &g = c->G();
    g.reset();
    g.generation++;
    if (g.words.p) CU(cudaMemsetAsync(g.words.p, 0, g.words.cap, c->stream));
    if (g.flags.p) CU(cudaMemsetAsync(g.flags.p, 0, g.flags.cap, c->stream));
    if (g.rank.p) CU(cudaMemsetAsync(g.rank.p, 0, g.rank.cap, c->stream));
    CU(stream_sync(c));
    return 0;
}

int64_t pafgpu_genome_total_bases(const pafgpu_ctx *c) { return c ? c->G().next_off : PAFGPU_EINVAL; }

// ---- device trim (paf_trim_unreliable_tails, impl/paf.c:906-953) -------------------------------------
static int trim_batch(pafgpu_ctx *c, bool fixed, float score_fraction, float max_fraction, const int64_t *t_end, int64_t *coords,
                      int64_t *kept_ops, uint8_t *failed) {
    if (!c || !t_end || !coords) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->have_batch) return fail(c, PAFGPU_ESTATE, "no batch has been encoded");
    CU(cudaSetDevice(c->device));
    const int64_t R = c->n_records;
    const size_t rb = (size_t)std::max<int64_t>(R, 1) * 8;
    CU(c->trim_coords.ensure(4 * rb));
    CU(c->trim_lo.ensure(rb));
    CU(c->trim_hi.ensure(rb));
    CU(c->trim_failed.ensure((size_t)std::max<int64_t>(R, 1)));
    int64_t *co = c->trim_coords.as<int64_t>();
    // the batch arrays live in c->in_* (host batches) or in the caller's device memory (device batches)
    CU(cudaMemcpyAsync(co, c->last_qs, (size_t)R * 8, cudaMemcpyDeviceToDevice, c->stream));
    CU(cudaMemcpyAsync(co + R, c->last_qe, (size_t)R * 8, cudaMemcpyDeviceToDevice, c->stream));
    CU(cudaMemcpyAsync(co + 2 * R, c->last_ts, (size_t)R * 8, cudaMemcpyDeviceToDevice, c->stream));
    CU(cudaMemcpyAsync(co + 3 * R, t_end, (size_t)R * 8, cudaMemcpyHostToDevice, c->stream));
    TrimParams T{};
    T.n_records = R; T.out = c->out.as<uint64_t>(); T.rec_out_begin = c->rec_out_begin.as<int64_t>();
    T.same_strand = c->last_strand;
    T.qs = co; T.qe = co + R; T.ts = co + 2 * R; T.te = co + 3 * R;
    T.lo = c->trim_lo.as<int64_t>(); T.hi = c->trim_hi.as<int64_t>(); T.failed = c->trim_failed.as<uint8_t>();
    T.score_fraction = score_fraction; T.max_fraction = max_fraction;
    // long CIGARs (mean > 64 ops per record): one warp per record instead of one thread ($PAFGPU_TRIM_WARP = 0 / 1 forces)
    bool by_warp = R > 0 && c->n_out / R > 64;
    if (const char *e = getenv("PAFGPU_TRIM_WARP")) by_warp = atoi(e) != 0;
    if (R > 0 && !fixed && by_warp) trim_warp_kernel<<<blocks_for(R * 32, 128), 128, 0, c->stream>>>(T);
    else if (R > 0 && !fixed) trim_kernel<<<blocks_for(R, 128), 128, 0, c->stream>>>(T);
    if (R > 0 && fixed) trim_fixed_kernel<<<blocks_for(R, 128), 128, 0, c->stream>>>(T, c->out.as<uint64_t>());
    CU(cudaMemcpyAsync(coords, co, 4 * (size_t)R * 8, cudaMemcpyDeviceToHost, c->stream));
    if (failed) CU(cudaMemcpyAsync(failed, T.failed, (size_t)R, cudaMemcpyDeviceToHost, c->stream));
    std::vector<int64_t> lo, hi;
    if (kept_ops) {
        lo.resize((size_t)R); hi.resize((size_t)R);
        CU(cudaMemcpyAsync(lo.data(), T.lo, (size_t)R * 8, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaMemcpyAsync(hi.data(), T.hi, (size_t)R * 8, cudaMemcpyDeviceToHost, c->stream));
    }
    CU(stream_sync(c));
    CU(cudaGetLastError());
    if (kept_ops) for (int64_t i = 0; i < R; i++) kept_ops[i] = hi[(size_t)i] - lo[(size_t)i];
    c->trimmed = true;
    return 0;
}

int pafgpu_trim(pafgpu_ctx *c, float score_fraction, float max_fraction, const int64_t *t_end, int64_t *coords, int64_t *kept_ops,
                uint8_t *failed) {
    return trim_batch(c, false, score_fraction, max_fraction, t_end, coords, kept_ops, failed);
}

int pafgpu_trim_fixed(pafgpu_ctx *c, float end_fraction, const int64_t *t_end, int64_t *coords, int64_t *kept_ops) {
    if (!(end_fraction >= 0.0f && end_fraction <= 1.0f)) return c ? fail(c, PAFGPU_EINVAL, "end fraction outside [0, 1]") : PAFGPU_EINVAL;
    return trim_batch(c, true, 0.0f, end_fraction, t_end, coords, kept_ops, nullptr);
}

// ---- a batch that is already encoded (the input of `paffy trim`): ops go straight to the result buffers ----------
int pafgpu_stage_ops_host(pafgpu_ctx *c, const pafgpu_batch *b) {
    if (!c || !b || b->n_records < 0 || b->n_ops < 0) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    CU(cudaSetDevice(c->device));
    const int64_t R = b->n_records, N = b->n_ops;
    c->have_batch = false;
    c->trimmed = false;
    auto up = [&](DevBuf &d, const void *src, size_t bytes) -> int {
        CU(d.ensure(std::max<size_t>(bytes, 8)));
        if (bytes) CU(cudaMemcpyAsync(d.p, src, bytes, cudaMemcpyHostToDevice, c->stream));
        return 0;
    };
    int rc;
    if ((rc = up(c->rec_out_begin, b->op_begin, (size_t)(R + 1) * 8))) return rc;
    if ((rc = up(c->out, b->ops, (size_t)N * 8))) return rc;
    if ((rc = up(c->in_qs, b->q_start, (size_t)R * 8))) return rc;
    if ((rc = up(c->in_qe, b->q_end, (size_t)R * 8))) return rc;
    if ((rc = up(c->in_ts, b->t_start, (size_t)R * 8))) return rc;
    if ((rc = up(c->in_strand, b->same_strand, (size_t)R))) return rc;
    CU(c->rec_mism.ensure((size_t)std::max<int64_t>(R, 1) * 8));
    CU(cudaMemsetAsync(c->rec_mism.p, 0, (size_t)std::max<int64_t>(R, 1) * 8, c->stream));
    CU(stream_sync(c));
    c->last_strand = c->in_strand.as<uint8_t>();
    c->last_qs = c->in_qs.as<int64_t>(); c->last_qe = c->in_qe.as<int64_t>(); c->last_ts = c->in_ts.as<int64_t>();
    c->n_records = R; c->n_ops = N; c->n_out = N; c->n_tiles = 0; c->n_win = 0; c->aligned = 0; c->retried = 0;
    c->bad_record = -1;
    c->ser_text_bytes = 0;
    c->have_batch = true;
    c->staged_only = true;
    return 0;
}

// ---- device serialiser ---------------------------------------------------------------------------
int pafgpu_serialise(pafgpu_ctx *c, const char *prefix_bytes, const int64_t *prefix_begin, int64_t *text_bytes_out, float *ms_out) {
    if (!c || !prefix_bytes || !prefix_begin) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->have_batch) return fail(c, PAFGPU_ESTATE, "no batch has been encoded");
    CU(cudaSetDevice(c->device));
    const int64_t R = c->n_records, K = c->n_out;
    const int64_t pre_bytes = prefix_begin[R];
    CU(c->ser_prefix.ensure((size_t)std::max<int64_t>(pre_bytes, 8)));
    CU(c->ser_prefix_begin.ensure((size_t)(R + 1) * 8));
    CU(cudaMemcpyAsync(c->ser_prefix.p, prefix_bytes, (size_t)pre_bytes, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->ser_prefix_begin.p, prefix_begin, (size_t)(R + 1) * 8, cudaMemcpyHostToDevice, c->stream));
    CU(cudaEventRecord(c->ev[0], c->stream));
    SerialiseParams S{};
    S.prefix = c->ser_prefix.as<char>();
    int64_t total = 0;
    if (int rc = lay_out_and_write(c, R, K, S, 4 * K + pre_bytes + R + 4096, &total, [&](const SerialiseParams &L) {
            if (R > 0) prefix_copy_kernel<<<blocks_for(R * 32, 256), 256, 0, c->stream>>>(L);
        }))
        return rc;
    CU(cudaEventRecord(c->ev[1], c->stream));
    CU(stream_sync(c));
    CU(cudaGetLastError());
    c->ser_text_bytes = total;
    if (text_bytes_out) *text_bytes_out = total;
    if (ms_out) cudaEventElapsedTime(ms_out, c->ev[0], c->ev[1]);
    return 0;
}

int pafgpu_fetch_text(pafgpu_ctx *c, char *text, int64_t *line_begin) {
    if (!c) return PAFGPU_EINVAL;
    std::lock_guard<std::mutex> lk(c->mu);
    if (!c->have_batch) return fail(c, PAFGPU_ESTATE, "no batch has been serialised");
    CU(cudaSetDevice(c->device));
    if (text && c->ser_text_bytes)
        CU(cudaMemcpyAsync(text, c->ser_text.p, (size_t)c->ser_text_bytes, cudaMemcpyDeviceToHost, c->stream));
    if (line_begin)
        CU(cudaMemcpyAsync(line_begin, c->ser_line_begin.p, (size_t)(c->n_records + 1) * 8, cudaMemcpyDeviceToHost, c->stream));
    CU(stream_sync(c));
    return 0;
}

}  // extern "C"

int pafgpu_internal_fail(pafgpu_ctx *c, int code, const char *msg) { return fail(c, code, msg ? msg : ""); }

// ---- serialiser without allocations -------------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it). The one-pass serialiser (ser_write_kernel,
// serialise.cuh) needs scratch - one look-back word per block of SER_OPS_PER_BLOCK ops and the record
// of every block's first op - and here it lives in a file-scope device array of fixed size, which every GPU that loads
// the library holds until exit. All calls share it: it keeps the last call's contents (the part a call uses is cleared
// first), and calls that overlap on one GPU race.
namespace {
constexpr int64_t NOALLOC_MAX_OPS = 1ll << 26;                                  // 64 Mi output ops per call
constexpr int64_t NOALLOC_MAX_BLOCKS = NOALLOC_MAX_OPS / SER_OPS_PER_BLOCK;     // 32 768
constexpr size_t NOALLOC_SCRATCH_BYTES = (size_t)(NOALLOC_MAX_BLOCKS + 2) * 8 + (size_t)(NOALLOC_MAX_BLOCKS + 2) * 4;  // 393 240 B
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
void *g_noalloc_ptr[NOALLOC_MAX_DEVICES];  // device address of the array on each device, asked for once (that device current)

int noalloc_fail(int code, const std::string &msg) {
    g_create_error = msg;  // read with pafgpu_last_error(NULL)
    return code;
}
}  // namespace

extern "C" int pafgpu_serialise_no_alloc(int device, void *stream, const uint64_t *ops_dev, int64_t n_ops, const int64_t *op_begin_dev,
                                         int64_t n_records, const char *prefix_dev, const int64_t *prefix_begin_dev, char *text_dev,
                                         int64_t text_cap, int64_t *line_begin_dev, int64_t *total_bytes) {
    if (n_ops < 0 || n_records < 0 || text_cap < 0 || !op_begin_dev || !prefix_begin_dev || !line_begin_dev || (n_ops > 0 && !ops_dev) ||
        (text_cap > 0 && !text_dev) || (n_records > 0 && !prefix_dev))
        return noalloc_fail(PAFGPU_EINVAL, "pafgpu_serialise_no_alloc: bad argument");
    if (n_ops > 0 && n_records == 0) return noalloc_fail(PAFGPU_EINVAL, "pafgpu_serialise_no_alloc: ops without records");
    if (n_ops > NOALLOC_MAX_OPS)
        return noalloc_fail(PAFGPU_EINVAL, "pafgpu_serialise_no_alloc: " + std::to_string(n_ops) + " ops exceed the limit of " +
                                               std::to_string(NOALLOC_MAX_OPS) + " ops per call (fixed scratch array)");
    if (n_records >= (1ll << 31)) return noalloc_fail(PAFGPU_EINVAL, "pafgpu_serialise_no_alloc: at most 2^31 - 1 records per call");
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return noalloc_fail(PAFGPU_EINVAL, "pafgpu_serialise_no_alloc: bad device");
    auto cuda_fail = [](const char *what, cudaError_t e) { return noalloc_fail(PAFGPU_ENODEVICE, std::string(what) + ": " + cudaGetErrorString(e)); };
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return cuda_fail("cudaSetDevice", e);
    void *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_ptr[device]) {
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p) return cuda_fail("cudaGetSymbolAddress(g_noalloc_scratch)", e);
            g_noalloc_ptr[device] = p;
        }
        scratch = g_noalloc_ptr[device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t nb = blocks_for(n_ops, SER_OPS_PER_BLOCK);
    SerialiseParams S{};
    S.n_records = n_records; S.n_out = n_ops;
    S.out = ops_dev; S.rec_out_begin = op_begin_dev;
    S.status = static_cast<unsigned long long *>(scratch);
    uint32_t *block_rec = reinterpret_cast<uint32_t *>(S.status + NOALLOC_MAX_BLOCKS + 2);
    S.block_rec = block_rec;
    S.prefix = prefix_dev; S.prefix_begin = prefix_begin_dev;
    S.line_begin = line_begin_dev;
    S.text = text_dev; S.text_cap = text_cap;
    S.keep_lo = S.keep_hi = nullptr;
    if (n_ops > 0) {
        e = cudaMemsetAsync(S.status, 0, (size_t)(nb + 1) * 8, st);  // look-back words start at zero
        if (e != cudaSuccess) return cuda_fail("cudaMemsetAsync", e);
        ser_block_records_kernel<<<blocks_for(n_records, 256), 256, 0, st>>>(n_records, op_begin_dev, block_rec);
        ser_write_kernel<<<(unsigned)nb, SER_THREADS, 0, st>>>(S);
    } else {
        line_begin_no_ops_kernel<<<blocks_for(n_records + 1, 256), 256, 0, st>>>(n_records, prefix_begin_dev, line_begin_dev);
    }
    if (n_records > 0) prefix_copy_kernel<<<blocks_for(n_records * 32, 256), 256, 0, st>>>(S);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail("pafgpu_serialise_no_alloc: launch", e);
    if (total_bytes) {
        e = cudaMemcpyAsync(total_bytes, line_begin_dev + n_records, 8, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) return cuda_fail("pafgpu_serialise_no_alloc", e);
        if (*total_bytes > text_cap)
            return noalloc_fail(PAFGPU_EINVAL, "pafgpu_serialise_no_alloc: the text needs " + std::to_string(*total_bytes) + " bytes, the buffer holds " +
                                                   std::to_string(text_cap) + " (nothing was written beyond it)");
    }
    return 0;
}
