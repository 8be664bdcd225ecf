"""GPU parity: the CUDA base-comparison path vs the CPU oracle, bit-exact (integer work)."""
import numpy as np
import pytest

from oracle import pyoracle as O
from tests import util

pytestmark = pytest.mark.gpu


def check_case(ctx, case):
    want = util.oracle_encode(case)
    got, res = util.gpu_encode(ctx, case)
    for i, (w, g) in enumerate(zip(want, got)):
        assert g[0] == w[0], f"record {i} {case['records'][i]}:\n gpu {O.cigar_str(g[0])[:400]}\n ora {O.cigar_str(w[0])[:400]}"
        assert g[1] == w[1], f"record {i}: mismatches gpu {g[1]} oracle {w[1]}"
    return res


def test_reference_unit_vectors(ctx):
    """tests/paf_unit_test.c:525-559: 5M all match -> 5=, all mismatch -> 5X, mixed -> 2=2X."""
    for q, t, m, want in [(b"ACGTA", b"ACGTA", 5, [(5, "=")]), (b"AAAAA", b"CCCCC", 5, [(5, "X")]),
                          (b"ACGT", b"ACTA", 4, [(2, "="), (2, "X")])]:
        out, mm = ctx.encode_one(util.words([(m, "M")]), True, 0, m, 0, q, t)
        assert util.pairs_of(out) == want
        assert mm == sum(n for n, c in want if c == "X")


def test_edge_pair(ctx):
    """per-op runs, vanishing 0M, case folding, IUPAC on both strands."""
    q = b"ACGTNnacgtMRWSYKVHDBXACGT"
    t = b"ACGTNNACGTKYWSRMBDHVXACGA"
    for ops, strand in [([(25, "M")], True), ([(25, "M")], False),
                        ([(3, "M"), (0, "M"), (2, "M"), (5, "M"), (4, "I"), (4, "D"), (11, "M")], False),
                        ([(2, "="), (2, "X"), (6, "M"), (15, "M")], True)]:
        want = O.encode(ops, strand, 0, 25, 0, q, t)
        out, _ = ctx.encode_one(util.words(ops), strand, 0, 25, 0, q, t)
        assert util.pairs_of(out) == want, (ops, strand)
    assert O.cigar_str(O.encode([(25, "M")], True, 0, 25, 0, q, t)) == "10=2X2=6X4=1X"


@pytest.mark.parametrize("seed", range(6))
def test_random_small(ctx, seed):
    check_case(ctx, util.random_case(seed, n_records=60, contig_len=30000, n_ops=14, max_len=700))


@pytest.mark.parametrize("seed", [11, 12])
def test_random_with_exceptions(ctx, seed):
    check_case(ctx, util.random_case(seed, n_records=60, contig_len=30000, n_ops=14, max_len=700, p_weird=0.002))


def test_long_ops_span_tiles(ctx):
    """M ops far longer than one warp tile (8192 bases), incl. perfect-match stretches crossing many tiles."""
    case = util.random_case(21, n_records=12, contig_len=400000, n_ops=6, max_len=120000, p_main=0.9995)
    res = check_case(ctx, case)
    assert res.n_tiles > 20


def test_huge_indels_escape_the_event_list(ctx):
    """I / D / = ops longer than 2^24 bases do not fit a 32-bit list entry: they are stored whole after the copy-out."""
    rng = np.random.default_rng(5)
    big = (1 << 24) + 12345
    tgt = util.random_seq(rng, 2 * big + 5000, b"A", 0.9)
    qry = util.random_seq(rng, 2 * big + 5000, b"A", 0.9)
    ops = [(300, "M"), (big, "D"), (700, "M"), (big + 7, "I"), (64, "M"), (big, "="), (1, "M"), (3, "I"), (900, "M")]
    q_span = sum(n for n, c in ops if c != "D")
    recs = [dict(q="q", t="t", qs=0, qe=q_span, ts=0, strand=True, ops=ops),
            dict(q="q", t="t", qs=17, qe=17 + q_span, ts=33, strand=True, ops=ops)]
    check_case(ctx, dict(contigs={"t": tgt, "q": qry}, records=recs))


def test_dense_flips_overflow_retry(ctx):
    """High-entropy genome: ~37% of bases start a run, exceeding the first capacity guess -> retry path."""
    case = util.random_case(31, n_records=30, contig_len=100000, n_ops=8, max_len=20000, p_main=0.25)
    check_case(ctx, case)


def test_many_short_ops(ctx):
    """More than 32 ops inside 32 windows (pre-encoded =/X input and 1-base M ops)."""
    rng = np.random.default_rng(5)
    case = util.random_case(41, n_records=1, contig_len=50000)
    ops = [(int(rng.integers(1, 3)), "M=XID"[int(rng.integers(0, 5))]) for _ in range(3000)]
    t_span = sum(n for n, c in ops if c != "I"); q_span = sum(n for n, c in ops if c != "D")
    case["records"] = [dict(q="qfwd", t="tgt", qs=5, qe=5 + q_span, ts=9, strand=True, ops=ops),
                       dict(q="qrev desc", t="tgt", qs=7, qe=7 + q_span, ts=1, strand=False, ops=ops)]
    check_case(ctx, case)


def test_bounds_error(ctx):
    import paffy_b200
    case = util.random_case(51, n_records=3, contig_len=5000)
    case["records"][1]["ts"] = 4990
    case["records"][1]["ops"] = [(100, "M")]
    case["records"][1]["qe"] = case["records"][1]["qs"] + 100
    with pytest.raises(paffy_b200.PafGpuError) as e:
        util.gpu_encode(ctx, case)
    assert e.value.code == -4 and "record 1" in str(e.value)


@pytest.mark.parametrize("seed", [61, 62])
def test_serialiser_matches_oracle_text(ctx, seed, tmp_path):
    """Device serialiser output == the oracle's add_mismatches file output, byte for byte."""
    case = util.random_case(seed, n_records=80, contig_len=30000, n_ops=14, max_len=2500, p_weird=0.001)
    fa = O.Fasta()
    for n, s in case["contigs"].items():
        fa.add(n, s)
    inp, outp = tmp_path / "in.paf", tmp_path / "out.paf"
    inp.write_text("\n".join(util.case_lines(case)) + "\n")
    rc, stats, aligned = fa.add_mismatches_file(str(inp), str(outp))
    assert rc == 0
    want = outp.read_bytes()
    got, res = util.gpu_encode(ctx, case)
    assert res.aligned_bases == aligned
    prefixes = [ln[: ln.index(b"\tcg:Z:") + 6] for ln in want.split(b"\n")[:-1]]
    begin = np.zeros(len(prefixes) + 1, dtype=np.int64)
    begin[1:] = np.cumsum([len(p) for p in prefixes])
    n, _ = ctx.serialise(b"".join(prefixes), begin)
    text, lines = ctx.fetch_text()
    assert n == len(want) and lines[-1] == n
    assert text.tobytes() == want
