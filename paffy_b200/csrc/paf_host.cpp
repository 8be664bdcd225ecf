// paf_host.cpp - host side of the add_mismatches path above the batch C ABI: PAF text in, PAF text out.
//
// Mirrors, for this path only, what the reference's command mains do around paf_encode_mismatches
// (impl/paf_add_mismatches.c:106-140, impl/paf_view.c:148-192, impl/paf_trim.c:113-127): read records,
// look the contigs up, encode (on the GPU, as one batch), paf_check, write - with the reference's record
// order, tag normalisation, failure ordering (records before a bad one are still written) and exit codes.
// Parsing restates paf_parse / cigar_parse (impl/paf.c:70-111, 137-209) and printing restates
// paf_write_to_buffer (impl/paf.c:317-389); the trim pass (impl/paf.c:811-953) runs on the device (trim.cuh). Nothing is copied.
// There is no CPU implementation of the base comparison here: every M op goes through the CUDA kernels.
#include "../../include/paffy_b200.h"
#include "../../include/paf_b200_compat.h"

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <deque>
#include <map>
#include <mutex>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <strings.h>
#include <string>
#include <thread>
#include <vector>

#include "paf_text.h"

namespace {

using namespace paftext;


struct Parsed {
    std::vector<Rec> recs;
    std::vector<uint64_t> ops;
};

// parse the lines of text[0, n) (already free of '\r') on `threads` workers, keeping input order
void parse_all(const char *text, size_t n, int threads, Parsed &out) {
    threads = std::max(1, std::min<int>(threads, (int)(n / (1 << 16)) + 1));
    std::vector<size_t> cut(threads + 1, n);
    cut[0] = 0;
    for (int t = 1; t < threads; t++) {
        size_t p = n / threads * t;
        const void *nl = p < n ? memchr(text + p, '\n', n - p) : nullptr;
        cut[t] = nl ? (size_t)((const char *)nl - text) + 1 : n;
    }
    for (int t = 1; t <= threads; t++) cut[t] = std::max(cut[t], cut[t - 1]);
    std::vector<Parsed> part(threads);
    auto work = [&](int t) {
        Parsed &P = part[t];
        const char *p = text + cut[t], *e = text + cut[t + 1];
        while (p < e) {
            const char *nl = (const char *)memchr(p, '\n', (size_t)(e - p));
            const char *le = nl ? nl : e;
            if (le == p && !nl) break;
            Rec r;
            parse_line(p, le, r, P.ops);
            P.recs.push_back(r);
            p = nl ? nl + 1 : e;
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < threads; t++) th.emplace_back(work, t);
    work(0);
    for (auto &x : th) x.join();
    size_t nr = 0, no = 0;
    for (auto &P : part) { nr += P.recs.size(); no += P.ops.size(); }
    out.recs.reserve(nr);
    out.ops.reserve(no);
    for (auto &P : part) {
        const int64_t shift = (int64_t)out.ops.size();
        for (Rec &r : P.recs) { r.op_begin += shift; r.op_end += shift; out.recs.push_back(r); }
        out.ops.insert(out.ops.end(), P.ops.begin(), P.ops.end());
    }
}

int hw_threads() {
    if (const char *e = getenv("PAFFY_THREADS")) { const int t = atoi(e); if (t > 0) return t; }
    const unsigned n = std::thread::hardware_concurrency();
    return n ? (int)std::min(n, 32u) : 4;
}

template <class F>
void parallel_for(int64_t n, int threads, F f) {  // f(begin, end, thread)
    threads = (int)std::max<int64_t>(1, std::min<int64_t>(threads, n / 4096 + 1));
    std::vector<std::thread> th;
    for (int t = 1; t < threads; t++) th.emplace_back(f, n * t / threads, n * (t + 1) / threads, t);
    f(0, n / threads, 0);
    for (auto &x : th) x.join();
}

}  // namespace

#include "internal.h"

#include <chrono>
#include <cerrno>
#include <getopt.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace {

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// the whole line of a record as paf_write_to_buffer prints it
void render_line(const Rec &r, const uint64_t *ops, int64_t n, std::string &out) {
    render_prefix(r, out);
    if (r.has_cigar) render_ops(ops, n, out);
    out.push_back('\n');
}

std::string parse_error(const Rec &r) {
    char buf[128];
    switch (r.bad) {
        case 2: snprintf(buf, sizeof buf, "Got an unexpected strand character (%c) in a paf string", r.bad_char); break;
        case 3: snprintf(buf, sizeof buf, "Got an unexpected character paf cigar string: %c", r.bad_char ? r.bad_char : ' '); break;
        case 4: snprintf(buf, sizeof buf, "Got an unexpected tp value in a paf string"); break;
        default: snprintf(buf, sizeof buf, "Too few fields in a paf line"); break;
    }
    return buf;
}

struct Pinned {  // pinned host buffer that only grows (plain memory where no CUDA device exists: `add_mismatches -a`)
    void *p = nullptr;
    int64_t cap = 0;
    bool pageable = false;
    void release() {
        if (p) { if (pageable) free(p); else pafgpu_host_free(p); }
        p = nullptr;
        cap = 0;
    }
    bool ensure(int64_t bytes) {
        if (bytes <= cap) return true;
        release();
        cap = bytes + bytes / 4 + 4096;
        p = pafgpu_host_alloc(cap);
        pageable = false;
        if (!p && pafgpu_device_count() == 0) { p = malloc((size_t)cap); pageable = true; }
        if (!p) cap = 0;
        return p != nullptr;
    }
    bool ensure_exact(int64_t bytes) {  // no head room: the caller rounds to size classes itself
        if (bytes <= cap) return true;
        release();
        cap = bytes;
        p = pafgpu_host_alloc(cap);
        pageable = false;
        if (!p && pafgpu_device_count() == 0) { p = malloc((size_t)cap); pageable = true; }
        if (!p) cap = 0;
        return p != nullptr;
    }
    ~Pinned() { release(); }
    template <class T> T *as() { return static_cast<T *>(p); }
};

bool write_all(std::string *sink, const char *p, size_t n) {
    sink->append(p, n);
    return true;
}


}  // namespace

// ================================================================================================
// The file filter. Everything the reference does per record around paf_encode_mismatches, done per batch.
namespace {

struct Filter {
    pafgpu_ctx *ctx;
    pafgpu_file_opts o;
    pafgpu_file_report rep{};
    const pafgpu_filter *flt = nullptr;  // `| paffy filter` fused behind add_mismatches (impl/paf_filter.c:121-156)
    int threads = 1;
    bool device_text = false;  // plain add_mismatches: the PAF text is parsed and rendered on the device
    std::string *out = nullptr;  // output bytes of the chunk being processed, in input order (written by the ordered writer)
    Parsed in;
    std::string err;           // message of the record that ended the run
    // per-record
    std::vector<int32_t> qc, tc;
    // batch staging (pinned)
    Pinned p_ops, p_begin, p_qc, p_tc, p_qs, p_qe, p_ts, p_st, p_prefix, p_text, p_out_ops;
    std::vector<int64_t> prefix_begin, line_begin, out_begin, mism;
    std::vector<int64_t> map;  // batch record -> input record

    // records [a, b) -> one GPU batch. Returns 0, 1 (bad record: b is moved to it) or a negative library error.
    int run_batch(int64_t a, int64_t &b);
    int finish_add(int64_t a, int64_t b);
    int finish_trim(int64_t a, int64_t b);
    int finish_view(int64_t a, int64_t b);
};

int Filter::run_batch(int64_t a, int64_t &b) {
    std::vector<Rec> &R = in.recs;
    // `paffy trim` alone prints every op it was given; the encoder drops zero-length M ops (SURVEY a1-edge ii)
    const bool trim_only = o.mode == PAFGPU_MODE_TRIM;
    auto dropped = [&](uint64_t w) { return !trim_only && op_code(w) == OPM && op_len(w) == 0; };
    for (;;) {
        // SoA of the records that carry aligned work; the others are printed by the host as they are
        map.clear();
        int64_t n_ops = 0;
        for (int64_t i = a; i < b; i++) {
            if (!R[i].has_cigar) continue;
            int64_t kept = 0;
            for (int64_t k = R[i].op_begin; k < R[i].op_end; k++) kept += !dropped(in.ops[k]);
            if (!kept) continue;
            map.push_back(i);
            n_ops += kept;
        }
        const int64_t nb = (int64_t)map.size();
        if (nb == 0) return 0;
        if (!p_ops.ensure(n_ops * 8) || !p_begin.ensure((nb + 1) * 8) || !p_qc.ensure(nb * 4) || !p_tc.ensure(nb * 4) ||
            !p_qs.ensure(nb * 8) || !p_qe.ensure(nb * 8) || !p_ts.ensure(nb * 8) || !p_st.ensure(nb))
            return pafgpu_internal_fail(ctx, PAFGPU_ENODEVICE, "pinned host allocation failed");
        int64_t *begin = p_begin.as<int64_t>();
        begin[0] = 0;
        for (int64_t j = 0; j < nb; j++) {
            const Rec &r = R[map[j]];
            int64_t kept = 0;
            for (int64_t k = r.op_begin; k < r.op_end; k++) kept += !dropped(in.ops[k]);
            begin[j + 1] = begin[j] + kept;
        }
        uint64_t *ops = p_ops.as<uint64_t>();
        parallel_for(nb, threads, [&](int64_t j0, int64_t j1, int) {
            for (int64_t j = j0; j < j1; j++) {
                const Rec &r = R[map[j]];
                uint64_t *w = ops + begin[j];
                for (int64_t k = r.op_begin; k < r.op_end; k++)
                    if (!dropped(in.ops[k])) *w++ = in.ops[k];
                p_qc.as<int32_t>()[j] = qc[map[j]]; p_tc.as<int32_t>()[j] = tc[map[j]];
                p_qs.as<int64_t>()[j] = r.qs; p_qe.as<int64_t>()[j] = r.qe; p_ts.as<int64_t>()[j] = r.ts;
                p_st.as<uint8_t>()[j] = r.same_strand ? 1 : 0;
            }
        });
        pafgpu_batch B{nb, n_ops, begin, ops, p_qc.as<int32_t>(), p_tc.as<int32_t>(), p_qs.as<int64_t>(), p_qe.as<int64_t>(),
                       p_ts.as<int64_t>(), p_st.as<uint8_t>()};
        if (trim_only) return pafgpu_stage_ops_host(ctx, &B);
        pafgpu_result res;
        const int rc = pafgpu_encode_batch_host(ctx, &B, &res);
        if (rc == 0) {
            rep.aligned_bases += res.aligned_bases;
            return 0;
        }
        if (rc != PAFGPU_EBOUNDS) return rc;
        // the reference would read outside the sequence here (undefined behaviour); the run ends at that record
        const int64_t bad = pafgpu_last_bad_record(ctx);
        if (bad < 0 || bad >= nb) return rc;
        b = map[bad];
        err = std::string("ERROR: ") + pafgpu_last_error(ctx);
        if (b == a) return 1;
        const int again = run_batch(a, b);
        return again ? again : 1;
    }
}

// stats of one record as paf_stats_calc sees its encoded CIGAR (impl/paf.c:236-260): input ops + the mismatch count the
// kernel tallied for the record's M ops. s: matches, mismatches, query_inserts, query_deletes, insert bases, delete bases
static void rec_stats(const Filter &F, const Rec &r, int64_t mm, int64_t *s) {
    int64_t m = 0;
    for (int64_t k = r.op_begin; k < r.op_end; k++) {
        const int c = op_code(F.in.ops[k]);
        const int64_t n = op_len(F.in.ops[k]);
        if (c == OPM) m += n;
        else if (c == OPEQ) s[0] += n;
        else if (c == OPX) s[1] += n;
        else if (c == OPI) { s[2]++; s[4] += n; }
        else if (c == OPD) { s[3]++; s[5] += n; }
    }
    s[0] += m - mm;
    s[1] += mm;
}

// The test of paffy_filter_main (impl/paf_filter.c:123-134) on the record as add_mismatches prints it: float32 quotients
// widened to double, compared with the double thresholds (0/0 = NaN fails every >=, so a record without aligned bases
// never passes); true = the record is written.
static bool filter_writes(const pafgpu_filter &f, const Rec &r, const int64_t *s) {
    const double identity = (float)s[0] / (s[0] + s[1]);
    const double identity_with_gaps = (float)s[0] / (s[0] + s[1] + s[4] + s[5]);
    const bool pass = r.score >= f.min_alignment_score && r.chain_score >= f.min_chain_score &&
                      (f.max_tile_level == -1 || r.tile_level <= f.max_tile_level) && identity >= f.min_identity &&
                      identity_with_gaps >= f.min_identity_with_gaps;
    return f.invert ? !pass : pass;
}

// totals over records [a, b), as `paffy view -s` accumulates them (impl/paf_view.c:176-178)
static void add_stats(Filter &F, int64_t a, int64_t b) {
    int64_t *s = F.rep.stats;
    size_t j = 0;
    for (int64_t i = a; i < b; i++) {
        int64_t mm = 0;
        if (j < F.map.size() && F.map[j] == i) mm = F.mism[j++];
        rec_stats(F, F.in.recs[i], mm, s);
        s[6]++;
    }
}

// `paffy view` (impl/paf_view.c:152-181): the encoded CIGAR itself is never printed, so nothing but the per-record
// mismatch tallies of the kernel comes back from the device; the per-alignment line is paf_pretty_print's header
// (impl/paf.c:270-279: float32 quotients printed with %f).
int Filter::finish_view(int64_t a, int64_t b) {
    std::vector<Rec> &R = in.recs;
    const int64_t nb = (int64_t)map.size();
    double t0 = now_s();
    mism.assign((size_t)nb, 0);
    if (nb) {
        const int rc = pafgpu_fetch_ops(ctx, nullptr, nullptr, mism.data());
        if (rc) return rc;
    }
    rep.s_gpu += now_s() - t0;
    t0 = now_s();
    add_stats(*this, a, b);
    bool ok = true;
    if (o.mode & PAFGPU_VIEW_PER_ALIGNMENT) {
        std::vector<int64_t> mm_of((size_t)(b - a), 0);
        for (int64_t j = 0; j < nb; j++) mm_of[(size_t)(map[(size_t)j] - a)] = mism[(size_t)j];
        std::vector<std::string> part((size_t)threads);
        parallel_for(b - a, threads, [&](int64_t i0, int64_t i1, int t) {
            std::string &out_s = part[(size_t)t];
            std::vector<char> line;
            for (int64_t i = a + i0; i < a + i1; i++) {
                const Rec &r = R[(size_t)i];
                int64_t s[6] = {0, 0, 0, 0, 0, 0};
                rec_stats(*this, r, mm_of[(size_t)(i - a)], s);
                const float identity = (float)s[0] / (s[0] + s[1]);
                const float identity_with_gaps = (float)s[0] / (s[0] + s[1] + s[4] + s[5]);
                line.resize((size_t)r.qname_len + r.tname_len + 512);
                const int n = snprintf(line.data(), line.size(),
                                       "Query:%.*s\tQ-start:%lld\tQ-length:%lld\tTarget:%.*s\tT-start:%lld\tT-length:%lld\tSame-strand:%i\t"
                                       "Score:%lld\tIdentity:%f\tIdentity-with-gaps%f\tAligned-bases:%lld\tQuery-inserts:%lld\t"
                                       "Query-deletes:%lld\n",
                                       (int)r.qname_len, r.qname, (long long)r.qs, (long long)(r.qe - r.qs), (int)r.tname_len, r.tname,
                                       (long long)r.ts, (long long)(r.te - r.ts), (int)r.same_strand, (long long)r.score, identity,
                                       identity_with_gaps, (long long)(s[0] + s[1]), (long long)s[2], (long long)s[3]);
                out_s.append(line.data(), (size_t)n);
            }
        });
        rep.s_host += now_s() - t0;
        t0 = now_s();
        for (auto &p : part) ok = ok && write_all(out, p.data(), p.size());
    }
    rep.s_write += now_s() - t0;
    return ok ? 0 : pafgpu_internal_fail(ctx, PAFGPU_EIO, "write failed");
}

int Filter::finish_add(int64_t a, int64_t b) {
    std::vector<Rec> &R = in.recs;
    const int64_t nb = (int64_t)map.size();
    double t0 = now_s();
    mism.assign((size_t)nb, 0);
    line_begin.assign((size_t)nb + 1, 0);
    int64_t text_bytes = 0;
    if (nb) {
        // text before the CIGAR, rendered on the host threads; the device appends <len><op>... and '\n'
        std::vector<std::string> part((size_t)threads);
        std::vector<int64_t> plen((size_t)nb);
        parallel_for(nb, threads, [&](int64_t j0, int64_t j1, int t) {
            std::string &s = part[(size_t)t];
            for (int64_t j = j0; j < j1; j++) {
                const size_t before = s.size();
                render_prefix(R[map[j]], s);
                plen[(size_t)j] = (int64_t)(s.size() - before);
            }
        });
        prefix_begin.assign((size_t)nb + 1, 0);
        for (int64_t j = 0; j < nb; j++) prefix_begin[(size_t)j + 1] = prefix_begin[(size_t)j] + plen[(size_t)j];
        if (!p_prefix.ensure(prefix_begin[(size_t)nb] + 8)) return pafgpu_internal_fail(ctx, PAFGPU_ENODEVICE, "pinned host allocation failed");
        {
            char *w = p_prefix.as<char>();
            for (auto &s : part) { memcpy(w, s.data(), s.size()); w += s.size(); }
        }
        rep.s_host += now_s() - t0;
        t0 = now_s();
        int rc = pafgpu_serialise(ctx, p_prefix.as<char>(), prefix_begin.data(), &text_bytes, nullptr);
        if (rc) return rc;
        if (!p_text.ensure(text_bytes + 8)) return pafgpu_internal_fail(ctx, PAFGPU_ENODEVICE, "pinned host allocation failed");
        if ((rc = pafgpu_fetch_text(ctx, p_text.as<char>(), line_begin.data()))) return rc;
        if ((rc = pafgpu_fetch_ops(ctx, nullptr, nullptr, mism.data()))) return rc;
        rep.s_gpu += now_s() - t0;
        t0 = now_s();
    }
    add_stats(*this, a, b);
    // write in input order: device text for the batch records, host text for the records without aligned work
    const char *text = p_text.as<char>();
    bool ok = true;
    if (flt) {
        // fused `| paffy filter`: the counts are the input op sums + the kernel's tallies; lines that fail are skipped
        std::string plain;
        int64_t j = 0, run_from = -1;  // run_from: first device line of the pending run of written lines
        auto flush = [&](int64_t upto) {
            if (run_from >= 0) ok = ok && write_all(out, text + line_begin[(size_t)run_from], (size_t)(line_begin[(size_t)upto] - line_begin[(size_t)run_from]));
            run_from = -1;
        };
        for (int64_t i = a; i < b && ok; i++) {
            const bool dev = j < nb && map[(size_t)j] == i;
            int64_t st[6] = {0, 0, 0, 0, 0, 0};
            rec_stats(*this, R[(size_t)i], dev ? mism[(size_t)j] : 0, st);
            const bool keep = filter_writes(*flt, R[(size_t)i], st);
            if (dev) {
                if (keep && run_from < 0) run_from = j;
                if (!keep) flush(j);
                j++;
            } else {
                flush(j);
                if (keep) {
                    plain.clear();
                    render_line(R[(size_t)i], nullptr, 0, plain);
                    ok = ok && write_all(out, plain.data(), plain.size());
                }
            }
        }
        flush(nb);
    } else if (nb == b - a) {
        ok = write_all(out, text, (size_t)text_bytes);
    } else {
        std::string plain;
        int64_t j = 0, run_from = 0;
        for (int64_t i = a; i < b && ok; i++) {
            if (j < nb && map[(size_t)j] == i) { j++; continue; }
            ok = write_all(out, text + line_begin[(size_t)run_from], (size_t)(line_begin[(size_t)j] - line_begin[(size_t)run_from]));
            run_from = j;
            plain.clear();
            render_line(R[i], nullptr, 0, plain);
            ok = ok && write_all(out, plain.data(), plain.size());
        }
        ok = ok && write_all(out, text + line_begin[(size_t)run_from], (size_t)(line_begin[(size_t)nb] - line_begin[(size_t)run_from]));
    }
    rep.s_write += now_s() - t0;
    return ok ? 0 : pafgpu_internal_fail(ctx, PAFGPU_EIO, "write failed");
}

// add_mismatches | trim: paf_trim_unreliable_tails (impl/paf.c:811-953) runs on the device over the encoded ops
// (pafgpu_trim, one thread per record, the reference's float32 arithmetic); only the moved coordinates come back. The
// host re-renders the text before each CIGAR with them, the device serialiser prints the kept ops.
int Filter::finish_trim(int64_t a, int64_t b) {
    std::vector<Rec> &R = in.recs;
    const int64_t nb = (int64_t)map.size();
    double t0 = now_s();
    mism.assign((size_t)nb, 0);
    line_begin.assign((size_t)nb + 1, 0);
    int64_t text_bytes = 0;
    int rc = 0;
    int64_t stop = b;  // first record that fails after trimming
    std::vector<Rec> trimmed;
    if (nb) {
        std::vector<int64_t> t_end((size_t)nb), coords((size_t)nb * 4);
        std::vector<uint8_t> failed((size_t)nb);
        for (int64_t j = 0; j < nb; j++) t_end[(size_t)j] = R[(size_t)map[(size_t)j]].te;
        if (o.trim_score_fraction < 0)  // `paffy trim -f`: fixed end fraction (impl/paf_trim.c:119-121)
            rc = pafgpu_trim_fixed(ctx, o.trim_max_fraction, t_end.data(), coords.data(), nullptr);
        else
            rc = pafgpu_trim(ctx, o.trim_score_fraction, o.trim_max_fraction, t_end.data(), coords.data(), nullptr, failed.data());
        if (rc) return rc;
        if ((rc = pafgpu_fetch_ops(ctx, nullptr, nullptr, mism.data()))) return rc;
        rep.s_gpu += now_s() - t0;
        t0 = now_s();
        trimmed.resize((size_t)nb);
        for (int64_t j = 0; j < nb; j++) {
            Rec r = R[(size_t)map[(size_t)j]];
            r.qs = coords[(size_t)j]; r.qe = coords[(size_t)(nb + j)]; r.ts = coords[(size_t)(2 * nb + j)]; r.te = coords[(size_t)(3 * nb + j)];
            trimmed[(size_t)j] = r;
            // paf_check after the trim (impl/paf_trim.c:124): the coordinate tests; the op sums follow the coordinates
            // by construction of the trim. The reference's final assert (impl/paf.c:952) aborts as well.
            std::string why;
            Rec coords_only = r;
            coords_only.has_cigar = false;
            int bad = stop == b ? (failed[(size_t)j] || check_record(coords_only, nullptr, 0, &why)) : 0;
            if (stop == b && !bad && o.mode == PAFGPU_MODE_TRIM) {
                // paf_check's op sums after the trim (impl/paf.c:443-459): the trim moves the coordinates by exactly the
                // spans of the ops it drops, so kept sum - new span = input sum - input span
                const Rec &r0 = R[(size_t)map[(size_t)j]];
                int64_t q = 0, t = 0;
                for (int64_t k = r0.op_begin; k < r0.op_end; k++) {
                    if (op_code(in.ops[(size_t)k]) != OPD) q += op_len(in.ops[(size_t)k]);
                    if (op_code(in.ops[(size_t)k]) != OPI) t += op_len(in.ops[(size_t)k]);
                }
                const int64_t dq = q - (r0.qe - r0.qs), dt = t - (r0.te - r0.ts);
                char buf[160];
                if (dq) {
                    snprintf(buf, sizeof buf, "Paf cigar alignment does not match query length: %lld vs. %lld", (long long)(r.qe - r.qs + dq), (long long)(r.qe - r.qs));
                    why = buf; bad = 1;
                } else if (dt) {
                    snprintf(buf, sizeof buf, "Paf cigar alignment does not match target length: %lld vs. %lld", (long long)(r.te - r.ts + dt), (long long)(r.te - r.ts));
                    why = buf; bad = 1;
                }
            }
            if (bad) {
                stop = map[(size_t)j];
                if (failed[(size_t)j]) why = "trimmed alignment identity fell below the untrimmed identity (reference assert)";
                std::string line;
                render_prefix(r, line);
                err = "ERROR: " + why + " " + line;
            }
        }
        std::vector<std::string> part((size_t)threads);
        std::vector<int64_t> plen((size_t)nb);
        parallel_for(nb, threads, [&](int64_t j0, int64_t j1, int t) {
            std::string &s = part[(size_t)t];
            for (int64_t j = j0; j < j1; j++) {
                const size_t before = s.size();
                render_prefix(trimmed[(size_t)j], s);
                plen[(size_t)j] = (int64_t)(s.size() - before);
            }
        });
        prefix_begin.assign((size_t)nb + 1, 0);
        for (int64_t j = 0; j < nb; j++) prefix_begin[(size_t)j + 1] = prefix_begin[(size_t)j] + plen[(size_t)j];
        if (!p_prefix.ensure(prefix_begin[(size_t)nb] + 8)) return pafgpu_internal_fail(ctx, PAFGPU_ENODEVICE, "pinned host allocation failed");
        {
            char *w = p_prefix.as<char>();
            for (auto &s : part) { memcpy(w, s.data(), s.size()); w += s.size(); }
        }
        rep.s_host += now_s() - t0;
        t0 = now_s();
        if ((rc = pafgpu_serialise(ctx, p_prefix.as<char>(), prefix_begin.data(), &text_bytes, nullptr))) return rc;
        if (!p_text.ensure(text_bytes + 8)) return pafgpu_internal_fail(ctx, PAFGPU_ENODEVICE, "pinned host allocation failed");
        if ((rc = pafgpu_fetch_text(ctx, p_text.as<char>(), line_begin.data()))) return rc;
        rep.s_gpu += now_s() - t0;
        t0 = now_s();
    }
    add_stats(*this, a, b);
    // write in input order up to the first failing record
    const char *text = p_text.as<char>();
    bool ok = true;
    std::string plain;
    int64_t j = 0, run_from = 0;
    for (int64_t i = a; i < stop && ok; i++) {
        if (j < nb && map[(size_t)j] == i) { j++; continue; }
        ok = write_all(out, text + line_begin[(size_t)run_from], (size_t)(line_begin[(size_t)j] - line_begin[(size_t)run_from]));
        run_from = j;
        plain.clear();
        render_line(R[(size_t)i], nullptr, 0, plain);
        ok = ok && write_all(out, plain.data(), plain.size());
    }
    ok = ok && write_all(out, text + line_begin[(size_t)run_from], (size_t)(line_begin[(size_t)j] - line_begin[(size_t)run_from]));
    rep.s_write += now_s() - t0;
    if (!ok) return pafgpu_internal_fail(ctx, PAFGPU_EIO, "write failed");
    if (stop < b) {
        rep.failed_record = stop;
        return 1;
    }
    return 0;
}

}  // namespace

extern "C" {

int pafgpu_partition(const int64_t *weights, int64_t n, int parts, int64_t *cuts) {
    if (n < 0 || parts < 1 || !cuts || (n && !weights)) return PAFGPU_EINVAL;
    long double total = 0;
    for (int64_t i = 0; i < n; i++) total += (long double)std::max<int64_t>(weights[i], 0);
    cuts[0] = 0;
    long double acc = 0;
    int64_t i = 0;
    for (int p = 1; p < parts; p++) {
        // first record index at which the running weight reaches p / parts of the total
        const long double want = total * p / parts;
        while (i < n && acc + (long double)std::max<int64_t>(weights[i], 0) / 2 < want) acc += (long double)std::max<int64_t>(weights[i++], 0);
        cuts[p] = i;
    }
    cuts[parts] = n;
    return 0;
}

int pafgpu_add_mismatches_file(pafgpu_ctx *ctx, const char *in_path, const char *out_path, const pafgpu_file_opts *opts,
                               pafgpu_file_report *report) {
    return pafgpu_add_mismatches_file_filtered(ctx, in_path, out_path, opts, nullptr, report);
}

}  // extern "C"

// ================================================================================================
// The streaming pipeline: the reference reads, encodes and writes one record at a time
// (impl/paf_add_mismatches.c:106-140); here a reader thread cuts the input into chunks of whole lines, worker threads
// (one per context: `streams` sibling contexts on each of `gpus` devices, the packed genome shared on a device and
// cloned GPU to GPU between devices) run a chunk each through parse -> validate -> GPU batches -> text, and the calling
// thread writes the finished chunks in input order. Memory is bounded by the chunks in flight, the first output
// leaves before the input ends, and a record the reference would abort on ends the run with everything before it
// written. One job is thereby sharded over the GPUs of the box by contiguous record ranges with no
// collective: a range is a chunk, ranges are dealt to whichever GPU is free, outputs are concatenated in input order.
namespace {

struct ChunkIn {
    int64_t seq = 0;
    Pinned *buf = nullptr;   // leased from g_pinned (the device path copies straight out of it); whole lines, free of '\r'
    size_t n = 0;
    const char *data() const { return buf ? static_cast<const char *>(buf->p) : nullptr; }
};

// Pinned host buffers for finished output text, shared by all pipelines of the process and kept between calls (pinning
// memory costs more than filling it): a worker leases one for the text of a chunk, the writer hands it back.
struct PinnedPool {
    std::mutex mu;
    std::vector<Pinned *> idle;
    Pinned *get(int64_t bytes) {
        // sizes in classes (32 MiB steps below 256 MiB, 128 MiB steps above): a buffer leased for one chunk fits the next
        const int64_t step = bytes < ((int64_t)256 << 20) ? (int64_t)32 << 20 : (int64_t)128 << 20;
        bytes = (bytes + step - 1) / step * step;
        Pinned *p = nullptr;
        {
            std::lock_guard<std::mutex> lk(mu);
            // the smallest idle buffer that is large enough, else the largest one (it grows)
            size_t best = idle.size();
            for (size_t i = 0; i < idle.size(); i++) {
                if (best == idle.size()) { best = i; continue; }
                const bool fits_i = idle[i]->cap >= bytes, fits_b = idle[best]->cap >= bytes;
                if (fits_i != fits_b ? fits_i : (fits_i ? idle[i]->cap < idle[best]->cap : idle[i]->cap > idle[best]->cap)) best = i;
            }
            if (best < idle.size()) { p = idle[best]; idle.erase(idle.begin() + (ptrdiff_t)best); }
        }
        if (!p) p = new Pinned();
        if (!p->ensure_exact(bytes)) { delete p; return nullptr; }
        return p;
    }
    void put(Pinned *p) {
        if (!p) return;
        std::lock_guard<std::mutex> lk(mu);
        idle.push_back(p);
    }
};
PinnedPool g_pinned;

struct ChunkOut {
    int rc = 0;                // 0, 1 (a record ended the run; everything before it is in `bytes`), < 0 library error
    std::string err;
    std::string bytes;         // host-assembled output ...
    Pinned *lease = nullptr;   // ... or text that came off the device in one piece (written after `bytes`)
    int64_t lease_bytes = 0;
    int64_t n_records = 0;     // records of the chunk (parsed)
    pafgpu_file_report rep{};  // the chunk's share: records written, aligned bases, batches, stats, stage seconds;
                               // failed_record is chunk-relative
    int code = 0;              // library error code behind rc < 0
    bool ready = false;
};

// paf_check runs behind the encode (impl/paf_add_mismatches.c:130-133), so its message shows the record with its =/X
// CIGAR: the one record is encoded by itself (a batch of one). Where that is impossible - the record walks outside its
// contig - the input CIGAR is shown.
std::string record_as_the_reference_prints_it(pafgpu_ctx *ctx, const Rec &r, const std::vector<uint64_t> &ops) {
    std::vector<uint64_t> kept, enc;
    for (int64_t k = r.op_begin; k < r.op_end; k++)
        if (!(op_code(ops[(size_t)k]) == OPM && op_len(ops[(size_t)k]) == 0)) kept.push_back(ops[(size_t)k]);
    const uint64_t *shown = ops.data() + r.op_begin;
    int64_t n_shown = r.op_end - r.op_begin;
    if (ctx && r.has_cigar) {
        const std::string qn(r.qname, r.qname_len), tn(r.tname, r.tname_len);
        const int32_t qc = (int32_t)pafgpu_genome_find(ctx, qn.c_str()), tc = (int32_t)pafgpu_genome_find(ctx, tn.c_str());
        const int64_t begin[2] = {0, (int64_t)kept.size()};
        const uint8_t strand = r.same_strand ? 1 : 0;
        pafgpu_batch B{1, (int64_t)kept.size(), begin, kept.data(), &qc, &tc, &r.qs, &r.qe, &r.ts, &strand};
        pafgpu_result res;
        if (kept.empty()) {
            n_shown = 0;
        } else if (qc >= 0 && tc >= 0 && pafgpu_encode_batch_host(ctx, &B, &res) == 0) {
            enc.resize((size_t)std::max<int64_t>(res.n_out_ops, 1));
            if (pafgpu_fetch_ops(ctx, enc.data(), nullptr, nullptr) == 0) { shown = enc.data(); n_shown = res.n_out_ops; }
        }
    }
    std::string line;
    render_line(r, shown, n_shown, line);
    if (!line.empty() && line.back() == '\n') line.pop_back();
    return line;
}

// The message the reference would die with on the line [b, e) (parse error, unknown contig, paf_check), or the bounds
// message when the record passes all of those (it then walks outside its contig, which is undefined behaviour there).
std::string failing_line_message(pafgpu_ctx *ctx, const char *b, const char *e, int64_t index) {
    Rec r;
    std::vector<uint64_t> ops;
    parse_line(b, e, r, ops);
    if (r.bad) return "ERROR: " + parse_error(r);
    const std::string qn(r.qname, r.qname_len), tn(r.tname, r.tname_len);
    if (pafgpu_genome_find(ctx, qn.c_str()) < 0) return "No query sequence named: " + qn + " found";
    if (pafgpu_genome_find(ctx, tn.c_str()) < 0) return "No target sequence named: " + tn + " found";
    std::string why;
    if (check_record(r, ops.data() + r.op_begin, r.op_end - r.op_begin, &why)) return "ERROR: " + why + " " + record_as_the_reference_prints_it(ctx, r, ops);
    char buf[160];
    snprintf(buf, sizeof buf, "ERROR: record %lld: CIGAR/coordinates walk outside the contig", (long long)index);
    return buf;
}

// `paffy add_mismatches` for one chunk with the text handled on the device end to end (pafgpu_add_mismatches_text): the
// host only moves bytes
void process_chunk_device(Filter &F, const char *text, size_t n_text, ChunkOut &R) {
    pafgpu_text_result tr;
    double t0 = now_s();
    const pafgpu_text_opts to{F.o.mode >= PAFGPU_MODE_VIEW ? (int)PAFGPU_MODE_VIEW : F.o.mode, F.o.trim_score_fraction, F.o.trim_max_fraction};
    const int rc = pafgpu_add_mismatches_text_opts(F.ctx, text, (int64_t)n_text, &to, &tr);
    if (rc < 0) { R.rc = R.code = rc; R.err = pafgpu_last_error(F.ctx); return; }
    R.n_records = tr.records_in;
    R.rep = pafgpu_file_report{};
    R.rep.failed_record = -1;
    R.rep.records = tr.records;
    R.rep.aligned_bases = tr.aligned_bases;
    R.rep.batches = 1;
    for (int k = 0; k < 7; k++) R.rep.stats[k] = tr.stats[k];
    if (tr.text_bytes > 0) {
        R.lease = g_pinned.get(tr.text_bytes);
        if (!R.lease) { R.rc = R.code = PAFGPU_ENODEVICE; R.err = "pinned host allocation failed"; return; }
        const int frc = pafgpu_fetch_text(F.ctx, R.lease->as<char>(), nullptr);
        if (frc < 0) { R.rc = R.code = frc; R.err = pafgpu_last_error(F.ctx); return; }
        R.lease_bytes = tr.text_bytes;
    }
    R.rep.s_gpu = now_s() - t0;
    if (tr.first_bad >= 0) {
        R.rc = 1;
        R.rep.failed_record = tr.first_bad;
        if (tr.bad_kind == 0) {
            R.err = failing_line_message(F.ctx, text + tr.bad_line_begin, text + tr.bad_line_end, tr.first_bad);
        } else {
            // the record failed behind the trim (impl/paf_trim.c:124 / impl/paf.c:952): the message shows it with the moved coordinates
            Rec r;
            std::vector<uint64_t> ops;
            parse_line(text + tr.bad_line_begin, text + tr.bad_line_end, r, ops);
            r.qs = tr.bad_coords[0]; r.qe = tr.bad_coords[1]; r.ts = tr.bad_coords[2]; r.te = tr.bad_coords[3];
            std::string why = "trimmed alignment identity fell below the untrimmed identity (reference assert)";
            bool any_op = false;
            for (uint64_t w : ops) any_op = any_op || !(op_code(w) == OPM && op_len(w) == 0);
            if (!any_op) {
                R.err = "ERROR: paffy trim needs a cigar on every record (the reference aborts here)";
                return;
            }
            if (tr.bad_kind == 1) {
                Rec coords_only = r;
                coords_only.has_cigar = false;
                check_record(coords_only, nullptr, 0, &why);
            }
            std::string line;
            render_prefix(r, line);
            R.err = "ERROR: " + why + " " + line;
        }
    }
}

// everything the per-record loop of the reference does, for the records of one chunk
void process_chunk(Filter &F, const char *text, size_t n_text, ChunkOut &R) {
    pafgpu_ctx *ctx = F.ctx;
    if (F.device_text) { process_chunk_device(F, text, n_text, R); return; }
    const bool trim_only = F.o.mode == PAFGPU_MODE_TRIM;
    const bool view = F.o.mode >= PAFGPU_MODE_VIEW;  // low bits: PAFGPU_VIEW_PER_ALIGNMENT, PAFGPU_VIEW_AGGREGATE
    auto finish = [&](int64_t x, int64_t y) {
        return view ? F.finish_view(x, y) : F.o.mode == PAFGPU_MODE_ADD ? F.finish_add(x, y) : F.finish_trim(x, y);
    };
    const int64_t batch_bases = F.o.batch_bases > 0 ? F.o.batch_bases : (int64_t)1 << 34;
    const int64_t batch_ops = (int64_t)1 << 30;
    // `paffy trim` (alone or behind add_mismatches) on a record without ops - no cg tag, or only zero-length M ops, which the
    // encoder drops: the identity is 0 / 0 and the reference's closing assert fires (impl/paf.c:952); the run ends there
    auto no_ops_for_trim = [&](const Rec &r) {
        if (trim_only) return r.op_end == r.op_begin;
        if (F.o.mode != PAFGPU_MODE_ADD_TRIM) return false;
        for (int64_t k = r.op_begin; k < r.op_end; k++)
            if (!(op_code(F.in.ops[(size_t)k]) == OPM && op_len(F.in.ops[(size_t)k]) == 0)) return false;
        return true;
    };
    F.rep = pafgpu_file_report{};
    F.rep.failed_record = -1;
    F.err.clear();
    F.out = &R.bytes;
    F.in.recs.clear();
    F.in.ops.clear();

    double t0 = now_s();
    parse_all(text, n_text, F.threads, F.in);
    F.rep.s_parse = now_s() - t0;
    std::vector<Rec> &Rv = F.in.recs;
    const int64_t n = (int64_t)Rv.size();
    R.n_records = n;
    int rc = 0;
    // ---- the record that ends the run, found in input order before anything is computed: parse failures, unknown
    // contigs (impl/paf_add_mismatches.c:116-127) and paf_check (impl/paf.c:427-461; encoding keeps every sum it tests)
    t0 = now_s();
    int64_t stop = n;
    F.qc.assign((size_t)n, -1);
    F.tc.assign((size_t)n, -1);
    {
        // the lookups and checks of all records in parallel; the first failing record (input order) decides
        std::atomic<int64_t> first_bad{n};
        std::vector<uint8_t> bad((size_t)n, 0);
        parallel_for(n, F.threads, [&](int64_t i0, int64_t i1, int) {
            std::string name;
            for (int64_t i = i0; i < i1; i++) {
                Rec &r = Rv[(size_t)i];
                bool fails = r.bad != 0;
                if (!fails) {
                    if (no_ops_for_trim(r)) {
                        fails = true;
                    } else if (F.o.mode != PAFGPU_MODE_REMOVE && !trim_only) {
                        name.assign(r.qname, r.qname_len);
                        const int64_t q = pafgpu_genome_find(ctx, name.c_str());
                        name.assign(r.tname, r.tname_len);
                        const int64_t t = q >= 0 ? pafgpu_genome_find(ctx, name.c_str()) : -1;
                        F.qc[(size_t)i] = (int32_t)q;
                        F.tc[(size_t)i] = (int32_t)t;
                        fails = q < 0 || t < 0;
                    }
                    if (!fails) {
                        if (F.o.mode == PAFGPU_MODE_REMOVE && r.has_cigar) r.op_end = merge_matches(F.in.ops.data(), r.op_begin, r.op_end);
                        // (`view` prints without paf_check: impl/paf_view.c:152-181; `paffy trim` checks after the trim only,
                        // impl/paf_trim.c:124: finish_trim)
                        if (!view && !trim_only && check_record(r, F.in.ops.data() + r.op_begin, r.op_end - r.op_begin, nullptr)) fails = true;
                    }
                }
                if (fails) {
                    bad[(size_t)i] = 1;
                    int64_t cur = first_bad.load();
                    while (i < cur && !first_bad.compare_exchange_weak(cur, i)) {}
                    break;  // later records of this share cannot be the first
                }
            }
        });
        if (first_bad.load() < n) {
            const int64_t i = first_bad.load();
            Rec &r = Rv[(size_t)i];
            std::string msg;
            if (r.bad) {
                msg = "ERROR: " + parse_error(r);
            } else if (no_ops_for_trim(r)) {
                // no cg tag: the identity trim's closing assert sees 0/0 (impl/paf.c:952), the fixed trim reads a NULL
                // cigar (impl/paf.c:520): the reference dies on this record either way
                msg = "ERROR: paffy trim needs a cigar on every record (the reference aborts here)";
            } else if (F.o.mode != PAFGPU_MODE_REMOVE && !trim_only && (F.qc[(size_t)i] < 0 || F.tc[(size_t)i] < 0)) {
                if (F.qc[(size_t)i] < 0) msg = "No query sequence named: " + std::string(r.qname, r.qname_len) + " found";
                else msg = "No target sequence named: " + std::string(r.tname, r.tname_len) + " found";
            } else {
                std::string why;
                check_record(r, F.in.ops.data() + r.op_begin, r.op_end - r.op_begin, &why);
                // (add_mismatches checks behind the encode: the record is shown with its =/X CIGAR; `-a` shows the merged one)
                msg = "ERROR: " + why + " " + record_as_the_reference_prints_it(F.o.mode == PAFGPU_MODE_REMOVE ? nullptr : ctx, r, F.in.ops);
            }
            stop = i;
            F.err = msg;
            F.rep.failed_record = i;
            rc = 1;
        }
    }
    F.rep.s_host += now_s() - t0;

    // ---- batches
    int64_t a = 0;
    while (a < stop) {
        int64_t b = a, bases = 0, nops = 0;
        if (F.o.mode == PAFGPU_MODE_REMOVE) {
            b = std::min<int64_t>(stop, a + (1 << 20));
        } else {
            while (b < stop && (b == a || (bases < batch_bases && nops + (Rv[(size_t)b].op_end - Rv[(size_t)b].op_begin) < batch_ops))) {
                for (int64_t k = Rv[(size_t)b].op_begin; k < Rv[(size_t)b].op_end; k++)
                    if (op_code(F.in.ops[(size_t)k]) == OPM || (trim_only && op_code(F.in.ops[(size_t)k]) >= OPEQ)) bases += op_len(F.in.ops[(size_t)k]);
                nops += Rv[(size_t)b].op_end - Rv[(size_t)b].op_begin;
                b++;
            }
        }
        int brc = 0;
        if (F.o.mode == PAFGPU_MODE_REMOVE) {
            t0 = now_s();
            std::vector<std::string> part((size_t)F.threads);
            parallel_for(b - a, F.threads, [&](int64_t i0, int64_t i1, int t) {
                for (int64_t i = a + i0; i < a + i1; i++)
                    render_line(Rv[(size_t)i], F.in.ops.data() + Rv[(size_t)i].op_begin, Rv[(size_t)i].op_end - Rv[(size_t)i].op_begin, part[(size_t)t]);
            });
            F.rep.s_host += now_s() - t0;
            for (auto &sp : part) write_all(F.out, sp.data(), sp.size());
        } else {
            t0 = now_s();
            int64_t b2 = b;
            brc = F.run_batch(a, b2);
            F.rep.s_gpu += now_s() - t0;
            if (brc == 1) {  // a record walks outside its contig: the records before it are still written
                F.rep.failed_record = b2;  // (earlier than any record the checks above found: those lie at or behind `stop`)
                if (b2 > a) {
                    const int frc = finish(a, b2);
                    if (frc < 0) brc = frc;
                }
                F.rep.records += std::max<int64_t>(0, F.rep.failed_record - a);
            } else if (brc == 0) {
                brc = finish(a, b);
                if (brc == 1) F.rep.records += F.rep.failed_record - a;
            }
        }
        if (brc) { rc = brc; break; }
        F.rep.records += b - a;
        F.rep.batches++;
        a = b;
    }
    R.rc = rc;
    R.err = F.err;
    R.rep = F.rep;
    if (rc < 0) { R.code = rc; R.err = ctx ? pafgpu_last_error(ctx) : "library error"; }
}

// reads whole lines into chunks of about `chunk_bytes` (sonLibFile.c:42-71: '\r' dropped, a last line without '\n' counts).
// read(2) returns what a pipe holds right now: a chunk ends at the last newline once it is full OR the pipe has run
// dry, so a slow producer sees its first lines come out before it closes the stream. The bytes land in pinned memory.
struct ChunkReader {
    int fd;
    size_t chunk_bytes;
    bool seekable = false;    // a regular file: large reads are split over a few threads (pread)
    int64_t file_off = 0;
    std::vector<char> carry;  // the incomplete last line of the previous read (holds no newline)
    bool eof = false;
    int64_t seq = 0;
    // false at end of input (or read error / allocation failure: *failed)
    bool next(ChunkIn &c, bool *failed) {
        c.n = 0;
        if (eof && carry.empty()) return false;
        size_t cap = std::max(chunk_bytes, carry.size()) + (1 << 16);
        c.buf = g_pinned.get((int64_t)cap);
        if (!c.buf) { *failed = true; return false; }
        char *d = c.buf->as<char>();
        size_t have = carry.size();
        if (have) memcpy(d, carry.data(), have);
        carry.clear();
        size_t last_nl = 0;  // one past the last newline seen
        while (!eof) {
            if (have + 4096 > cap) {  // a line longer than the chunk: grow (keeping the bytes)
                Pinned *bigger = g_pinned.get((int64_t)(2 * cap));
                if (!bigger) { *failed = true; g_pinned.put(c.buf); c.buf = nullptr; return false; }
                memcpy(bigger->p, d, have);
                g_pinned.put(c.buf);
                c.buf = bigger;
                d = c.buf->as<char>();
                cap *= 2;
            }
            const size_t want = chunk_bytes > have ? std::min(chunk_bytes - have, cap - have) : std::min<size_t>(std::max<size_t>(chunk_bytes, 256), cap - have);
            ssize_t got;
            if (seekable && want >= ((size_t)8 << 20)) {
                // page-cache copies of one file run in parallel
                const int nt = 4;
                ssize_t part[nt];
                std::vector<std::thread> th;
                for (int t = 0; t < nt; t++)
                    th.emplace_back([&, t]() {
                        const size_t a0 = want * (size_t)t / nt, a1 = want * (size_t)(t + 1) / nt;
                        size_t at = a0;
                        while (at < a1) {
                            const ssize_t r = pread(fd, d + have + at, a1 - at, (off_t)(file_off + (int64_t)at));
                            if (r < 0) { if (errno == EINTR) continue; part[t] = -1; return; }
                            if (r == 0) break;
                            at += (size_t)r;
                        }
                        part[t] = (ssize_t)(at - a0);
                    });
                for (auto &x : th) x.join();
                got = 0;
                for (int t = 0; t < nt; t++) {
                    if (part[t] < 0) { got = -1; break; }
                    got += part[t];
                    if ((size_t)part[t] < want * (size_t)(t + 1) / nt - want * (size_t)t / nt) break;  // end of file inside this part
                }
                if (got > 0) file_off += got;
            } else {
                do got = ::read(fd, d + have, want); while (got < 0 && errno == EINTR);
                if (got > 0) file_off += got;
            }
            if (got < 0) { *failed = true; g_pinned.put(c.buf); c.buf = nullptr; return false; }
            if (got == 0) { eof = true; break; }
            for (size_t k = have + (size_t)got; k > have; k--)
                if (d[k - 1] == '\n') { last_nl = k; break; }
            have += (size_t)got;
            // full, or the source has nothing more for now: hand over the whole lines read so far
            if (last_nl && (have >= chunk_bytes || (size_t)got < want)) break;
        }
        if (!eof) {
            carry.assign(d + last_nl, d + have);
            have = last_nl;
        }
        if (have == 0) { g_pinned.put(c.buf); c.buf = nullptr; return false; }
        if (memchr(d, '\r', have)) have = (size_t)(std::remove(d, d + have, '\r') - d);  // the line reader drops '\r' (sonLibFile.c:49-51)
        c.n = have;
        c.seq = seq++;
        return true;
    }
};

std::vector<int> device_list(pafgpu_ctx *ctx, int gpus_opt) {
    std::vector<int> devs;
    if (!ctx) return devs;
    const int first = pafgpu_device_of(ctx), avail = pafgpu_device_count();
    devs.push_back(first);
    const char *e = getenv("PAFFY_GPUS");
    if (gpus_opt <= 0 && e && strchr(e, ',')) {  // an explicit list: the context's own device stays first
        for (const char *p = e; *p;) {
            const int d = atoi(p);
            if (d >= 0 && d < avail && std::find(devs.begin(), devs.end(), d) == devs.end()) devs.push_back(d);
            p = strchr(p, ',');
            if (!p) break;
            p++;
        }
        return devs;
    }
    int want = gpus_opt > 0 ? gpus_opt : (e ? (strcmp(e, "all") == 0 ? avail : atoi(e)) : 1);
    want = std::max(1, std::min(want, avail));
    for (int d = 0; (int)devs.size() < want && d < avail; d++)
        if (d != first) devs.push_back(d);
    return devs;
}

}  // namespace

extern "C" {

int pafgpu_add_mismatches_file_filtered(pafgpu_ctx *ctx, const char *in_path, const char *out_path, const pafgpu_file_opts *opts,
                                        const pafgpu_filter *filter, pafgpu_file_report *report) {
    if (!ctx && !(opts && opts->mode == PAFGPU_MODE_REMOVE)) return PAFGPU_EINVAL;  // -a is host-only text work
    if (filter && opts && opts->mode != PAFGPU_MODE_ADD) return pafgpu_internal_fail(ctx, PAFGPU_EINVAL, "the fused filter follows add_mismatches (mode 0) only");
    const pafgpu_file_opts o = opts ? *opts : pafgpu_file_opts{PAFGPU_MODE_ADD, 0.05f, 1.0f, 0, 0, 0, 0, 0};
    if (o.mode < 0 || o.mode > 7) return pafgpu_internal_fail(ctx, PAFGPU_EINVAL, "bad mode");
    const bool view = o.mode >= PAFGPU_MODE_VIEW;
    pafgpu_file_report rep{};
    rep.failed_record = -1;
    const double t_start = now_s();
    const bool trace = getenv("PAFFY_TRACE") != nullptr;  // a timeline of the pipeline on stderr
    std::mutex trace_mu;
    auto mark = [&](const char *what, int64_t seq, int64_t bytes) {
        if (!trace) return;
        std::lock_guard<std::mutex> lk(trace_mu);
        fprintf(stderr, "[paffy_b200 %8.3f ms] %-14s chunk %lld  %lld bytes\n", 1e3 * (now_s() - t_start), what, (long long)seq, (long long)bytes);
    };

    // ---- contexts: `streams` per device; device 0 of the list is the caller's, the others hold a clone of its genome.
    // The helpers belong to the caller's context and outlive this call (created once, reused by the next call).
    std::vector<pafgpu_ctx *> ctxs;
    const std::vector<int> devs = device_list(ctx, o.gpus);
    int streams = o.streams > 0 ? o.streams : 2;
    if (const char *e = getenv("PAFFY_STREAMS")) if (o.streams <= 0 && atoi(e) > 0) streams = atoi(e);
    streams = std::min(streams, 8);
    if (ctx) {
        for (size_t d = 0; d < devs.size(); d++)
            for (int k = 0; k < streams; k++) {
                pafgpu_ctx *h = ctx;
                if (d > 0 || k > 0)
                    if (int rc = pafgpu_helper(ctx, devs[d], k, &h)) return rc;
                ctxs.push_back(h);
            }
    } else {
        ctxs.assign(2, nullptr);  // `add_mismatches -a`: two host workers
    }
    const int W = (int)ctxs.size();
    mark("contexts ready", -1, 0);
    rep.gpus = (int64_t)devs.size();
    rep.workers = W;

    FILE *fin = in_path ? fopen(in_path, "rb") : stdin;
    if (!fin) { return pafgpu_internal_fail(ctx, PAFGPU_EIO, (std::string("cannot open ") + in_path).c_str()); }
    if (!out_path) fflush(stdout);
    const int out_fd = out_path ? open(out_path, O_WRONLY | O_CREAT | O_TRUNC, 0666) : fileno(stdout);
    if (out_fd < 0) {
        if (in_path) fclose(fin);
        return pafgpu_internal_fail(ctx, PAFGPU_EIO, (std::string("cannot open ") + out_path).c_str());
    }
    mark("output open", -1, 0);
    struct stat out_st;
    const bool out_regular = out_path && fstat(out_fd, &out_st) == 0 && S_ISREG(out_st.st_mode);
    int64_t out_off = 0;  // regular files: bytes written so far
    const bool no_mmap = getenv("PAFFY_NO_MMAP") != nullptr;
    auto write_out = [&](const char *p, size_t n) -> bool {
        if (n == 0) return true;
        bool ok = true;
        const size_t piece = (size_t)4 << 20;
        bool mapped = false;
        if (out_regular && n >= 2 * piece && !no_mmap) {
            // a large piece of a regular file: the file is extended and mapped, several threads copy into the mapping
            // (writes through write(2) to ONE file serialise on the inode lock; page faults on a mapping do not)
            const int64_t page = sysconf(_SC_PAGESIZE);
            const int64_t m0 = out_off / page * page;
            const size_t span = (size_t)(out_off - m0) + n;
            if (ftruncate(out_fd, (off_t)(out_off + (int64_t)n)) == 0) {
                void *map = mmap(nullptr, span, PROT_READ | PROT_WRITE, MAP_SHARED, out_fd, (off_t)m0);
                if (map != MAP_FAILED) {
                    char *dst = static_cast<char *>(map) + (out_off - m0);
                    const int nt = (int)std::min<size_t>((size_t)std::max(2, std::min(hw_threads() - 2, 16)), n / piece);
                    std::vector<std::thread> th;
                    for (int t = 0; t < nt; t++)
                        th.emplace_back([&, t]() {
                            const size_t a0 = n * (size_t)t / (size_t)nt, a1 = n * (size_t)(t + 1) / (size_t)nt;
                            memcpy(dst + a0, p + a0, a1 - a0);
                        });
                    for (auto &x : th) x.join();
                    munmap(map, span);
                    mapped = true;
                }
            }
        }
        if (mapped) {
        } else {
            size_t a0 = 0;
            while (a0 < n) {
                const ssize_t w = out_regular ? pwrite(out_fd, p + a0, n - a0, (off_t)(out_off + (int64_t)a0)) : ::write(out_fd, p + a0, n - a0);
                if (w < 0) { if (errno == EINTR) continue; ok = false; break; }
                a0 += (size_t)w;
            }
        }
        out_off += (int64_t)n;
        return ok;
    };

    // ---- queues
    // the text of a chunk is handled on the device end to end for add_mismatches, add_mismatches | trim and the totals of
    // view -s -t; per-alignment view lines, trim alone and the fused filter parse on host threads
    const bool device_text = ctx && !filter && !getenv("PAFFY_HOST_PARSE") &&
                             (o.mode == PAFGPU_MODE_ADD || o.mode == PAFGPU_MODE_ADD_TRIM ||
                              (o.mode >= PAFGPU_MODE_VIEW && !(o.mode & PAFGPU_VIEW_PER_ALIGNMENT)));
    const size_t chunk_bytes = (size_t)std::min<int64_t>(o.chunk_bytes > 0 ? o.chunk_bytes : (int64_t)32 << 20, (int64_t)1 << 30);
    std::mutex mu;
    std::condition_variable cv_job, cv_done, cv_room;
    std::deque<ChunkIn> jobs;
    std::map<int64_t, ChunkOut> done;
    bool reader_done = false, read_failed = false;
    std::atomic<bool> stop{false};
    int64_t in_flight = 0;  // chunks read and not yet written: bounds the memory
    const int64_t max_in_flight = W + 2;
    double s_read = 0;

    std::thread reader([&]() {
        ChunkReader rd{fileno(fin), chunk_bytes};
        {
            struct stat st;
            rd.seekable = in_path && fstat(rd.fd, &st) == 0 && S_ISREG(st.st_mode);
        }
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(mu);
                cv_room.wait(lk, [&] { return in_flight < max_in_flight || stop.load(); });
                if (stop.load()) break;
            }
            ChunkIn c;
            const double t0 = now_s();
            bool failed = false;
            const bool more = rd.next(c, &failed);
            s_read += now_s() - t0;
            if (more) mark("read", c.seq, (int64_t)c.n);
            std::lock_guard<std::mutex> lk(mu);
            if (failed) read_failed = true;
            if (!more) break;
            in_flight++;
            jobs.push_back(std::move(c));
            cv_job.notify_one();
        }
        std::lock_guard<std::mutex> lk(mu);
        reader_done = true;
        cv_job.notify_all();
        cv_done.notify_all();
    });

    const int threads_total = o.threads > 0 ? o.threads : hw_threads();
    std::vector<std::thread> workers;
    for (int w = 0; w < W; w++)
        workers.emplace_back([&, w]() {
            Filter F;
            F.ctx = ctxs[(size_t)w];
            F.flt = filter;
            F.o = o;
            F.device_text = device_text;
            F.threads = std::max(1, threads_total / W);
            for (;;) {
                ChunkIn c;
                {
                    std::unique_lock<std::mutex> lk(mu);
                    cv_job.wait(lk, [&] { return !jobs.empty() || reader_done || stop.load(); });
                    if (stop.load() || jobs.empty()) break;
                    c = std::move(jobs.front());
                    jobs.pop_front();
                }
                ChunkOut R;
                mark("start", c.seq, (int64_t)c.n);
                process_chunk(F, c.data(), c.n, R);
                mark("done", c.seq, R.lease_bytes + (int64_t)R.bytes.size());
                g_pinned.put(c.buf);
                c.buf = nullptr;
                R.ready = true;
                std::lock_guard<std::mutex> lk(mu);
                done.emplace(c.seq, std::move(R));
                cv_done.notify_all();
            }
        });

    // ---- ordered writer (this thread)
    int rc = 0;
    std::string err;
    int64_t next_seq = 0, records_before = 0;
    bool wrote_ok = true;
    for (;;) {
        ChunkOut R;
        {
            std::unique_lock<std::mutex> lk(mu);
            cv_done.wait(lk, [&] { return done.count(next_seq) || (reader_done && in_flight == 0); });
            auto it = done.find(next_seq);
            if (it == done.end()) break;  // input exhausted, everything written
            R = std::move(it->second);
            done.erase(it);
        }
        const double t0 = now_s();
        wrote_ok = write_out(R.bytes.data(), R.bytes.size()) && write_out(R.lease ? R.lease->as<char>() : nullptr, (size_t)R.lease_bytes);
        g_pinned.put(R.lease);
        R.lease = nullptr;
        rep.s_write += now_s() - t0;
        mark("written", next_seq, R.lease_bytes + (int64_t)R.bytes.size());
        rep.records += R.rep.records;
        rep.aligned_bases += R.rep.aligned_bases;
        rep.batches += R.rep.batches;
        for (int k = 0; k < 7; k++) rep.stats[k] += R.rep.stats[k];
        rep.s_parse += R.rep.s_parse; rep.s_gpu += R.rep.s_gpu; rep.s_host += R.rep.s_host; rep.s_write += R.rep.s_write;
        rep.chunks++;
        if (R.rc != 0 || !wrote_ok) {
            rc = !wrote_ok ? PAFGPU_EIO : R.rc;
            err = !wrote_ok ? "write failed" : R.err;
            if (R.rc == 1) rep.failed_record = records_before + R.rep.failed_record;
            break;
        }
        records_before += R.n_records;
        next_seq++;
        std::lock_guard<std::mutex> lk(mu);
        in_flight--;
        cv_room.notify_one();
    }
    {
        std::lock_guard<std::mutex> lk(mu);
        stop.store(true);
        cv_room.notify_all();
        cv_job.notify_all();
    }
    reader.join();
    for (auto &t : workers) t.join();
    for (auto &kv : done) g_pinned.put(kv.second.lease);  // chunks finished behind the one that ended the run
    for (auto &c : jobs) g_pinned.put(c.buf);
    if (rc == 0 && read_failed) { rc = PAFGPU_EIO; err = "read failed"; }
    rep.s_read = s_read;

    if (view && rc == 0 && (o.mode & PAFGPU_VIEW_AGGREGATE)) {  // impl/paf_view.c:183-192
        const int64_t *t = rep.stats;
        char line[512];
        const int len = snprintf(line, sizeof line, "Total-alignments:%lld\tAvg-Identity:%f\tAvg-Identity-with-gaps:%f\tAligned-bases:%lld\tAligned-bases-with-gaps:%lld\t"
                                 "Query-inserts:%lld\tQuery-deletes:%lld\n",
                                 (long long)t[6], (float)t[0] / (t[0] + t[1]), (float)t[0] / (t[0] + t[1] + t[4] + t[5]), (long long)(t[0] + t[1]),
                                 (long long)(t[0] + t[1] + t[4] + t[5]), (long long)t[2], (long long)t[3]);
        if (!write_out(line, (size_t)len) && rc == 0) { rc = PAFGPU_EIO; err = "write failed"; }
    }
    if (in_path) fclose(fin);
    if (out_path && close(out_fd) != 0 && rc == 0) { rc = PAFGPU_EIO; err = "write failed"; }
    mark("closed", -1, out_off);
    rep.s_total = now_s() - t_start;
    if (rc == 1) fprintf(stderr, "%s\n", err.c_str());
    if (rc != 0) pafgpu_internal_fail(ctx, rc, err.c_str());
    if (report) *report = rep;
    return rc;
}

}  // extern "C"


// ================================================================================================
// FASTA positionals of a subcommand -> packed genome in the context. With $PAFFY_B200_CACHE=<dir> the packed arena is
// persisted there, keyed by the paths, sizes and modification times of the FASTA files: the next process of a pipeline
// (tests/paf_pipeline_test.sh runs add_mismatches once per split_file shard, each re-reading both FASTA files) reads the
// packed file at copy speed instead of parsing the FASTA again.
namespace {
int load_genome(pafgpu_ctx *ctx, int argc, char *argv[], int first, bool verbose) {
    std::string key;
    for (int i = first; i < argc; i++) {
        struct stat st;
        char buf[96];
        if (stat(argv[i], &st) != 0) { key.clear(); break; }
        snprintf(buf, sizeof buf, "|%lld|%lld.%09ld;", (long long)st.st_size, (long long)st.st_mtim.tv_sec, (long)st.st_mtim.tv_nsec);
        key += argv[i];
        key += buf;
    }
    const char *dir = getenv("PAFFY_B200_CACHE");
    std::string cache;
    if (dir && *dir && !key.empty()) {
        uint64_t h = 1469598103934665603ull;
        for (unsigned char ch : key) { h ^= ch; h *= 1099511628211ull; }
        char name[40];
        snprintf(name, sizeof name, "/%016llx.pafb200", (unsigned long long)h);
        cache = std::string(dir) + name;
        if (pafgpu_genome_load_packed(ctx, cache.c_str(), key.c_str()) == 0) {
            if (verbose) fprintf(stderr, "Read %i packed sequences from %s\n", (int)pafgpu_genome_num_contigs(ctx), cache.c_str());
            return 0;
        }
    }
    for (int i = first; i < argc; i++) {
        if (verbose) fprintf(stderr, "Parsing sequence file : %s\n", argv[i]);
        if (pafgpu_genome_load_fasta(ctx, argv[i])) return 1;
    }
    if (verbose) fprintf(stderr, "Read %i sequences from sequence files\n", (int)pafgpu_genome_num_contigs(ctx));
    if (!cache.empty() && pafgpu_genome_save(ctx, cache.c_str(), key.c_str()) && verbose)
        fprintf(stderr, "could not write %s: %s\n", cache.c_str(), pafgpu_last_error(ctx));
    return 0;
}
}  // namespace

// ================================================================================================
// Drop-in subcommand entry point (impl/paf_add_mismatches.c:26-160).
namespace {
void add_mismatches_usage() {
    fprintf(stderr, "paffy add_mismatches [fasta_file]xN [options], version 0.1 (paffy_b200: CUDA sm_100a path)\n");
    fprintf(stderr, "Replace Ms with =/Xs in PAF cigar string\n");
    fprintf(stderr, "-i --inputFile : Input paf file. If not specified reads from stdin\n");
    fprintf(stderr, "-o --outputFile : Output paf file. If not specified outputs to stdout\n");
    fprintf(stderr, "-a : Remove mismatches, removing X and = encoding and replacing with M\n");
    fprintf(stderr, "-l --logLevel : Set the log level\n");
    fprintf(stderr, "-h --help : Print this help message\n");
}
}  // namespace

extern "C" int paffy_add_mismatches_main(int argc, char *argv[]) {
    const char *input = nullptr, *output = nullptr, *log_level = nullptr;
    bool remove = false;
    static struct option long_options[] = {{"logLevel", required_argument, 0, 'l'},   {"inputFile", required_argument, 0, 'i'},
                                           {"outputFile", required_argument, 0, 'o'}, {"removeMismatches", no_argument, 0, 'a'},
                                           {"help", no_argument, 0, 'h'},             {0, 0, 0, 0}};
    optind = 1;
    for (;;) {
        int idx = 0;
        const int key = getopt_long(argc, argv, "l:i:o:ha", long_options, &idx);
        if (key == -1) break;
        switch (key) {
            case 'l': log_level = optarg; break;
            case 'i': input = optarg; break;
            case 'o': output = optarg; break;
            case 'a': remove = true; break;
            case 'h': add_mismatches_usage(); return 0;
            default: add_mismatches_usage(); return 1;
        }
    }
    const bool verbose = log_level && (!strcasecmp(log_level, "INFO") || !strcasecmp(log_level, "DEBUG"));
    const double t0 = now_s();
    pafgpu_ctx *ctx = nullptr;
    const char *dev = getenv("PAFFY_B200_DEVICE");
    if (!remove && pafgpu_create(dev ? atoi(dev) : 0, &ctx)) {
        fprintf(stderr, "ERROR: %s\n", pafgpu_last_error(nullptr));
        exit(1);
    }
    if (ctx && load_genome(ctx, argc, argv, optind, verbose)) {
        fprintf(stderr, "ERROR: %s\n", pafgpu_last_error(ctx));
        exit(1);
    }
    pafgpu_file_opts o{remove ? PAFGPU_MODE_REMOVE : PAFGPU_MODE_ADD, 0.05f, 1.0f, 0, 0, 0, 0, 0};
    pafgpu_file_report rep;
    const int rc = pafgpu_add_mismatches_file(ctx, input, output, &o, &rep);
    if (rc < 0) fprintf(stderr, "ERROR: %s\n", pafgpu_last_error(ctx));
    pafgpu_destroy(ctx);
    if (rc) exit(1);  // st_errAbort / exit(1) in the reference
    if (verbose) fprintf(stderr, "Paffy add_mismatches is done!, %lld seconds have elapsed\n", (long long)(now_s() - t0));
    return 0;
}

// ================================================================================================
// Drop-in subcommand entry point `paffy view` (impl/paf_view.c:39-215).
namespace {
// FASTA file -> name -> sequence bytes as they stand in the file (fastaReadToFunction's rules, bioioC.c:155-209: the header
// is the whole line, '>' anywhere ends a record, newline / space / tab are dropped, any other byte must be a letter or
// '-', a repeated name replaces the earlier sequence). Only `view -a` needs the bytes on the host.
bool read_fasta_host(const char *path, std::map<std::string, std::string> &seqs) {
    FILE *fh = fopen(path, "rb");
    if (!fh) { fprintf(stderr, "ERROR: cannot open %s\n", path); return false; }
    std::string all;
    char tmp[1 << 16];
    size_t got;
    while ((got = fread(tmp, 1, sizeof tmp, fh)) > 0) all.append(tmp, got);
    fclose(fh);
    size_t p = all.find('>');
    while (p != std::string::npos) {
        const size_t nl = all.find('\n', p + 1);
        const std::string name = all.substr(p + 1, nl == std::string::npos ? std::string::npos : nl - p - 1);
        const size_t b = nl == std::string::npos ? all.size() : nl + 1;
        const size_t nx = all.find('>', b);
        const size_t e = nx == std::string::npos ? all.size() : nx;
        std::string seq;
        seq.reserve(e - b);
        for (size_t i = b; i < e; i++) {
            const unsigned char ch = (unsigned char)all[i];
            if (ch == '\n' || ch == ' ' || ch == '\t') continue;
            if (!((ch >= 'A' && ch <= 'Z') || (ch >= 'a' && ch <= 'z') || ch == '-')) {
                fprintf(stderr, "ERROR: !!Got an unexpected character in input fasta sequence: '%c' \n", ch);
                return false;
            }
            seq.push_back((char)ch);
        }
        seqs[name] = std::move(seq);
        p = nx;
    }
    return true;
}

// the record loop of paffy_view_main with -a (impl/paf_view.c:151-183) on the paf.h surface; totals as -s accumulates them
int view_with_alignment(int argc, char *argv[], int first_fasta, const char *input, const char *output, bool aggregate, bool verbose,
                        int64_t *totals) {
    std::map<std::string, std::string> seqs;
    for (int i = first_fasta; i < argc; i++) {
        if (verbose) fprintf(stderr, "Parsing sequence file : %s\n", argv[i]);
        if (!read_fasta_host(argv[i], seqs)) return 1;
    }
    FILE *fin = input ? fopen(input, "r") : stdin, *fout = output ? fopen(output, "w") : stdout;
    if (!fin || !fout) { fprintf(stderr, "ERROR: cannot open %s\n", !fin ? input : output); return 1; }
    int64_t cap = 100;
    char *line = (char *)malloc((size_t)cap);
    for (Paf *paf; (paf = paf_read_with_buffer(fin, true, &line, &cap)) != nullptr;) {
        auto q = seqs.find(paf->query_name), t = seqs.find(paf->target_name);
        if (q == seqs.end()) { fprintf(stderr, "No query sequence named: %s found\n", paf->query_name); exit(1); }
        if (t == seqs.end()) { fprintf(stderr, "No target sequence named: %s found\n", paf->target_name); exit(1); }
        paf_encode_mismatches(paf, &q->second[0], &t->second[0]);
        paf_pretty_print(paf, &q->second[0], &t->second[0], fout, true);
        if (aggregate) {
            paf_stats_calc(paf, &totals[0], &totals[1], &totals[2], &totals[3], &totals[4], &totals[5], false);
            totals[6]++;
        }
        paf_destruct(paf);
    }
    free(line);
    if (aggregate)
        fprintf(fout, "Total-alignments:%lld\tAvg-Identity:%f\tAvg-Identity-with-gaps:%f\tAligned-bases:%lld\tAligned-bases-with-gaps:%lld\t"
                      "Query-inserts:%lld\tQuery-deletes:%lld\n",
                (long long)totals[6], (float)totals[0] / (totals[0] + totals[1]), (float)totals[0] / (totals[0] + totals[1] + totals[4] + totals[5]),
                (long long)(totals[0] + totals[1]), (long long)(totals[0] + totals[1] + totals[4] + totals[5]), (long long)totals[2], (long long)totals[3]);
    if (input) fclose(fin);
    if (output) fclose(fout); else fflush(fout);
    return 0;
}

void view_usage() {
    fprintf(stderr, "paffy view [fasta_files]xN [options], version 0.1 (paffy_b200: CUDA sm_100a path)\n");
    fprintf(stderr, "Pretty print PAF alignments\n");
    fprintf(stderr, "-i --inputFile : Input paf file to invert. If not specified reads from stdin\n");
    fprintf(stderr, "-o --outputFile : Output paf file. If not specified outputs to stdout\n");
    fprintf(stderr, "-a --includeAlignment : Include base level alignment in output\n");
    fprintf(stderr, "-s --printAggregateStats : Print overall stats about the alignments at the end\n");
    fprintf(stderr, "-t --noPerAlignmentStats : Do not print stats about each paf\n");
    fprintf(stderr, "-u --errorIfIdentityLowerThanX : Float between 0 and 1. Assert identity is >= X.\n");
    fprintf(stderr, "-v --errorIfAlignedBasesLowerThanX : Integer >= 0. Assert total aligned based is >= X.\n");
    fprintf(stderr, "-l --logLevel : Set the log level\n");
    fprintf(stderr, "-h --help : Print this help message\n");
}
}  // namespace

extern "C" int paffy_view_main(int argc, char *argv[]) {
    const char *input = nullptr, *output = nullptr, *log_level = nullptr;
    bool include_alignment = false, aggregate = false, per_alignment = true;
    float min_identity = 0.0f;
    int64_t min_aligned = 0;
    static struct option long_options[] = {{"logLevel", required_argument, 0, 'l'},
                                           {"inputFile", required_argument, 0, 'i'},
                                           {"outputFile", required_argument, 0, 'o'},
                                           {"includeAlignment", no_argument, 0, 'a'},
                                           {"printAggregateStats", no_argument, 0, 's'},
                                           {"noPerAlignmentStats", no_argument, 0, 't'},
                                           {"errorIfIdentityLowerThanX", required_argument, 0, 'u'},
                                           {"errorIfAlignedBasesLowerThanX", required_argument, 0, 'v'},
                                           {"help", no_argument, 0, 'h'},
                                           {0, 0, 0, 0}};
    optind = 1;
    for (;;) {
        int idx = 0;
        const int key = getopt_long(argc, argv, "l:i:o:hastu:v:", long_options, &idx);
        if (key == -1) break;
        switch (key) {
            case 'l': log_level = optarg; break;
            case 'i': input = optarg; break;
            case 'o': output = optarg; break;
            case 'a': include_alignment = true; break;
            case 's': aggregate = true; break;
            case 't': per_alignment = false; break;
            case 'u': min_identity = (float)atof(optarg); break;
            case 'v': min_aligned = atoi(optarg); break;
            case 'h': view_usage(); return 0;
            default: view_usage(); return 1;
        }
    }
    if (optind >= argc) {
        fprintf(stderr, "Expected at least one sequence file\n");
        exit(1);
    }
    const bool verbose = log_level && (!strcasecmp(log_level, "INFO") || !strcasecmp(log_level, "DEBUG"));
    const double t0 = now_s();
    pafgpu_file_report rep{};
    if (include_alignment && per_alignment) {
        // -a prints the ORIGINAL bytes of both sequences (case kept) beside every record: the record loop of the reference
        // (impl/paf_view.c:151-183) over the paf.h surface - paf_encode_mismatches still compares the bases on the GPU,
        // against whole contigs that stay packed in HBM between records (paf_compat.cpp)
        if (view_with_alignment(argc, argv, optind, input, output, aggregate, verbose, rep.stats)) exit(1);
    } else {
        pafgpu_ctx *ctx = nullptr;
        const char *dev = getenv("PAFFY_B200_DEVICE");
        if (pafgpu_create(dev ? atoi(dev) : 0, &ctx)) {
            fprintf(stderr, "ERROR: %s\n", pafgpu_last_error(nullptr));
            exit(1);
        }
        if (load_genome(ctx, argc, argv, optind, verbose)) {
            fprintf(stderr, "ERROR: %s\n", pafgpu_last_error(ctx));
            exit(1);
        }
        pafgpu_file_opts o{PAFGPU_MODE_VIEW | (per_alignment ? PAFGPU_VIEW_PER_ALIGNMENT : 0) | (aggregate ? PAFGPU_VIEW_AGGREGATE : 0),
                           0.05f, 1.0f, 0, 0, 0, 0, 0};
        const int rc = pafgpu_add_mismatches_file(ctx, input, output, &o, &rep);
        if (rc < 0) fprintf(stderr, "ERROR: %s\n", pafgpu_last_error(ctx));
        pafgpu_destroy(ctx);
        if (rc) exit(1);
    }
    // the reference's closing asserts (impl/paf_view.c:195-196; asserts are compiled in, include.mk:32). The totals they
    // test are only accumulated under -s (:176-178): without it they see 0 / 0 = NaN, the first assert fires and the
    // reference aborts (exit 134) - here as well, after the output has been written.
    int64_t zero[7] = {0, 0, 0, 0, 0, 0, 0};
    const int64_t *t = aggregate ? rep.stats : zero;
    if (!((float)t[0] / (t[0] + t[1]) >= min_identity)) {
        fprintf(stderr, "paffy view: Assertion `(float)total_matches/(total_matches+total_mismatches) >= error_if_identity_lower_than_this' failed.\n");
        abort();
    }
    if (!(t[0] + t[1] >= min_aligned)) {
        fprintf(stderr, "paffy view: Assertion `total_matches+total_mismatches >= error_if_aligned_bases_lower_than_this' failed.\n");
        abort();
    }
    if (verbose) fprintf(stderr, "Paffy view is done!, %lld seconds have elapsed\n", (long long)(now_s() - t0));
    return 0;
}

// ================================================================================================
// Drop-in subcommand entry point `paffy trim` (impl/paf_trim.c:38-140): no FASTA; the parsed ops are staged on the
// device, trimmed there (identity trim by default, -f = fixed end fraction), paf_check'ed and serialised.
namespace {
void trim_usage() {
    fprintf(stderr, "paffy trim [options], version 0.1 (paffy_b200: CUDA sm_100a path)\n");
    fprintf(stderr, "Trims the ends of the PAF alignments\n");
    fprintf(stderr, "-i --inputFile : Input paf file. If not specified reads from stdin\n");
    fprintf(stderr, "-o --outputFile : Output paf file. If not specified outputs to stdout\n");
    fprintf(stderr, "-r --trimIdentity : Trim tails with alignment identity lower than this fraction of the overall alignment identity (default 0.05)\n");
    fprintf(stderr, "-t --trimFraction : Fraction (from 0 to 1) of aligned bases to trim from ends of the alignment, or the maximum fraction trimmed by identity (default 1.0)\n");
    fprintf(stderr, "-f --fixedTrim : Trim a constant fraction (-t) from each end instead of trimming by identity\n");
    fprintf(stderr, "-l --logLevel : Set the log level\n");
    fprintf(stderr, "-h --help : Print this help message\n");
}
}  // namespace

extern "C" int paffy_trim_main(int argc, char *argv[]) {
    const char *input = nullptr, *output = nullptr, *log_level = nullptr;
    float trim_end_fraction = 1.0f, trim_by_identity_fraction = 0.05f;  // impl/paf_trim.c:14-16
    bool trim_by_identity = true;
    static struct option long_options[] = {{"logLevel", required_argument, 0, 'l'},     {"inputFile", required_argument, 0, 'i'},
                                           {"outputFile", required_argument, 0, 'o'},   {"help", no_argument, 0, 'h'},
                                           {"trimFraction", required_argument, 0, 't'}, {"trimIdentity", required_argument, 0, 'r'},
                                           {"fixedTrim", no_argument, 0, 'f'},          {0, 0, 0, 0}};
    optind = 1;
    for (;;) {
        int idx = 0;
        const int key = getopt_long(argc, argv, "l:i:o:ht:r:f", long_options, &idx);
        if (key == -1) break;
        switch (key) {
            case 'l': log_level = optarg; break;
            case 'i': input = optarg; break;
            case 'o': output = optarg; break;
            case 't': trim_end_fraction = (float)atof(optarg); break;
            case 'r': trim_by_identity_fraction = (float)atof(optarg); break;
            case 'f': trim_by_identity = false; break;
            case 'h': trim_usage(); return 0;
            default: trim_usage(); return 1;
        }
    }
    const bool verbose = log_level && (!strcasecmp(log_level, "INFO") || !strcasecmp(log_level, "DEBUG"));
    const double t0 = now_s();
    if (!trim_by_identity && !(trim_end_fraction >= 0.0f && trim_end_fraction <= 1.0f)) {
        fprintf(stderr, "paffy trim: Assertion `percentage >= 0 && percentage <= 1.0' failed.\n");  // impl/paf.c:591
        abort();
    }
    pafgpu_ctx *ctx = nullptr;
    const char *dev = getenv("PAFFY_B200_DEVICE");
    if (pafgpu_create(dev ? atoi(dev) : 0, &ctx)) {
        fprintf(stderr, "ERROR: %s\n", pafgpu_last_error(nullptr));
        exit(1);
    }
    pafgpu_file_opts o{PAFGPU_MODE_TRIM, trim_by_identity ? trim_by_identity_fraction : -1.0f, trim_end_fraction, 0, 0, 0, 0, 0};
    if (trim_by_identity && trim_by_identity_fraction < 0) {  // the option's documented range is 0..1; < 0 is the -f switch in the ABI
        fprintf(stderr, "ERROR: paffy trim -r must be >= 0\n");
        exit(1);
    }
    pafgpu_file_report rep;
    const int rc = pafgpu_add_mismatches_file(ctx, input, output, &o, &rep);
    if (rc < 0) fprintf(stderr, "ERROR: %s\n", pafgpu_last_error(ctx));
    pafgpu_destroy(ctx);
    if (rc) exit(1);
    if (verbose) fprintf(stderr, "Paffy trim is done!, %lld seconds have elapsed\n", (long long)(now_s() - t0));
    return 0;
}
