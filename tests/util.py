"""Shared helpers for the parity tests: seeded random cases and the oracle comparison."""
import numpy as np

from oracle import pyoracle as O

OPS = "MID=X"
LEN_MASK = (1 << 56) - 1


def words(pairs):
    return np.array([(n & LEN_MASK) | (OPS.index(c) << 56) for n, c in pairs], dtype=np.uint64)


def pairs_of(ws):
    return [(int(w) & LEN_MASK, OPS[int(w) >> 56]) for w in ws]


def revcomp(seq: bytes) -> bytes:
    return bytes(O.lib().oracle_revcomp_char(bytes([b]))[0] for b in reversed(seq))


def random_seq(rng, n, main, p_main=0.9, extra=b"ACGT", p_weird=0.0, weird=b"NnRYKMacgt-"):
    """Low-entropy sequence so that long match runs and dense flips both occur."""
    r = rng.random(n)
    out = np.frombuffer(main * 1, dtype=np.uint8)[0] * np.ones(n, dtype=np.uint8)
    alt = np.frombuffer(extra, dtype=np.uint8)[rng.integers(0, len(extra), n)]
    out = np.where(r < p_main, out, alt)
    if p_weird > 0:
        w = np.frombuffer(weird, dtype=np.uint8)[rng.integers(0, len(weird), n)]
        # weird bytes come in short runs
        mask = rng.random(n) < p_weird
        mask = mask | np.roll(mask, 1) | np.roll(mask, 2)
        out = np.where(mask, w, out)
    return out.astype(np.uint8).tobytes()


def random_ops(rng, n_ops, max_len, p_m=0.6, allow_zero=True):
    pairs = []
    for _ in range(n_ops):
        c = "M" if rng.random() < p_m else "ID=X"[rng.integers(0, 4)]
        kind = rng.random()
        if kind < 0.25:
            n = int(rng.integers(0 if allow_zero else 1, 4))
        elif kind < 0.6:
            n = int(rng.integers(1, 70))
        else:
            n = int(rng.integers(1, max_len))
        if c != "M" and rng.random() < 0.7:
            n = min(n, 12)
        pairs.append((n, c))
    return pairs


def random_case(seed, n_records=40, contig_len=20000, n_ops=12, max_len=600, p_weird=0.0, p_main=0.9):
    """-> dict(contigs={name: bytes}, records=[dict(q,t,qs,qe,ts,strand,ops)])"""
    rng = np.random.default_rng(seed)
    contigs = {
        "tgt": random_seq(rng, contig_len, b"A", p_main, p_weird=p_weird),
        "qfwd": random_seq(rng, contig_len + 777, b"A", p_main, p_weird=p_weird),
        "qrev desc": random_seq(rng, contig_len + 31, b"T", p_main, p_weird=p_weird),
    }
    recs = []
    while len(recs) < n_records:
        ops = random_ops(rng, int(rng.integers(1, n_ops + 1)), max_len)
        t_span = sum(n for n, c in ops if c != "I")
        q_span = sum(n for n, c in ops if c != "D")
        strand = bool(rng.random() < 0.5)
        qn = "qfwd" if strand else "qrev desc"
        if t_span >= contig_len or q_span >= len(contigs[qn]):
            continue
        ts = int(rng.integers(0, contig_len - t_span + 1))
        qs = int(rng.integers(0, len(contigs[qn]) - q_span + 1))
        recs.append(dict(q=qn, t="tgt", qs=qs, qe=qs + q_span, ts=ts, strand=strand, ops=ops))
    return dict(contigs=contigs, records=recs)


def oracle_encode(case):
    """Per record: list of (len, op) pairs from the CPU oracle, and mismatch count inside M ops."""
    outs = []
    for r in case["records"]:
        q, t = case["contigs"][r["q"]], case["contigs"][r["t"]]
        enc = O.encode(r["ops"], r["strand"], r["qs"], r["qe"], r["ts"], q, t)
        base = sum(n for n, c in r["ops"] if c == "X")
        outs.append((enc, sum(n for n, c in enc if c == "X") - base))
    return outs


def flatten(case, ctx_find, drop_zero_m=True):
    """Records -> the SoA arrays of pafgpu_batch (host numpy)."""
    op_begin, ops, qc, tc, qs, qe, ts, st = [0], [], [], [], [], [], [], []
    for r in case["records"]:
        kept = [(n, c) for n, c in r["ops"] if not (drop_zero_m and c == "M" and n == 0)]
        ops.extend(kept)
        op_begin.append(len(ops))
        qc.append(ctx_find(r["q"])); tc.append(ctx_find(r["t"]))
        qs.append(r["qs"]); qe.append(r["qe"]); ts.append(r["ts"]); st.append(1 if r["strand"] else 0)
    return (np.array(op_begin, dtype=np.int64), words(ops), np.array(qc, dtype=np.int32), np.array(tc, dtype=np.int32),
            np.array(qs, dtype=np.int64), np.array(qe, dtype=np.int64), np.array(ts, dtype=np.int64),
            np.array(st, dtype=np.uint8))


def gpu_encode(ctx, case):
    """Run the CUDA path on a case; -> per record (pairs, mismatches)."""
    for name, seq in case["contigs"].items():
        ctx.add_contig(name, seq)
    arrs = flatten(case, ctx.find)
    res = ctx.encode_batch(*arrs)
    out, begin, mism = ctx.fetch_ops()
    assert begin[0] == 0 and begin[-1] == res.n_out_ops
    return [(pairs_of(out[begin[i]:begin[i + 1]]), int(mism[i])) for i in range(len(case["records"]))], res


def case_lines(case):
    """PAF text lines for a random case (coordinates valid for paf_check)."""
    lines = []
    for r in case["records"]:
        ql, tl = len(case["contigs"][r["q"]]), len(case["contigs"][r["t"]])
        t_span = sum(n for n, c in r["ops"] if c != "I")
        lines.append(f"{r['q']}\t{ql}\t{r['qs']}\t{r['qe']}\t{'+' if r['strand'] else '-'}\t{r['t']}\t{tl}\t{r['ts']}\t{r['ts'] + t_span}"
                     f"\t0\t0\t60\ttp:A:P\tcg:Z:{O.cigar_str(r['ops'])}")
    return lines


# ---- config C1: the reference's own fixture + a seeded synthetic chr10 -------------
C1_LEN = 133797422
C1_SEED = 20261104


def c1_genome():
    """The synthetic chr10 (query and target share the name, so one contig serves both): uint8 ASCII array."""
    rng = np.random.default_rng(C1_SEED)
    return np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, C1_LEN, dtype=np.uint8)]


def write_c1_fasta(path, width=100):
    g = c1_genome()
    full = (C1_LEN // width) * width
    body = np.empty((C1_LEN // width, width + 1), dtype=np.uint8)
    body[:, :width] = g[:full].reshape(-1, width)
    body[:, width] = 10
    with open(path, "wb") as fh:
        fh.write(b">chr10\n")
        fh.write(body.tobytes())
        if full < C1_LEN:
            fh.write(g[full:].tobytes() + b"\n")


def view_summary(stats):
    """The `paffy view -s` line (impl/paf_view.c:187-192) from the seven totals: float32 quotients printed with %f."""
    m, mm, qi, qd, qib, qdb, n = stats
    f = np.float32
    with np.errstate(all="ignore"):
        ident = f(m) / f(m + mm)
        ident_gaps = f(m) / f(m + mm + qib + qdb)
    return (f"Total-alignments:{n}\tAvg-Identity:{float(ident):f}\tAvg-Identity-with-gaps:{float(ident_gaps):f}\t"
            f"Aligned-bases:{m + mm}\tAligned-bases-with-gaps:{m + mm + qib + qdb}\tQuery-inserts:{qi}\tQuery-deletes:{qd}\n")
