"""Pins the CPU oracle (oracle/paf_oracle.c) to the reference: the reference's own unit-test answers
(tests/golden/reference_unit_vectors.json, restated from tests/paf_unit_test.c) and the committed outputs of
the unmodified reference CLI (tests/golden/*, written by tests/golden/make_golden.py). CPU only."""
import ctypes as C
import gzip
import hashlib
import json
import os

import pytest

from oracle import pyoracle as O
from tests import util

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
VEC = json.load(open(os.path.join(G, "reference_unit_vectors.json")))


def cig(text):
    """'5M3I' -> [(5,'M'),(3,'I')] through the oracle's own cigar_parse."""
    out = C.POINTER(C.c_uint64)()
    n = O.lib().oracle_cigar_parse(text.encode(), C.byref(out))
    assert n >= 0
    pairs = O.ops_to_pairs(out, n)
    if n:
        O.lib().free(out)
    return pairs


def make(rec):
    """(qname,qlen,qs,qe,strand,tname,tlen,ts,te,cigar) -> OPaf via the oracle's parser."""
    qn, ql, qs, qe, st, tn, tl, ts, te, cg = rec
    line = f"{qn}\t{ql}\t{qs}\t{qe}\t{st}\t{tn}\t{tl}\t{ts}\t{te}\t0\t0\t60" + (f"\tcg:Z:{cg}" if cg else "")
    rc, p = O.parse(line)
    assert rc == 0
    return p


def coords(p):
    return [p.query_start, p.query_end, p.target_start, p.target_end]


def cigar_of(p):
    return O.cigar_str(O.ops_to_pairs(p.ops, p.n_ops))


def test_revcomp_table():
    v = VEC["revcomp"]
    for a, b in zip(v["from"], v["to"]):
        assert O.lib().oracle_revcomp_char(a.encode()) == b.encode(), a


def test_cigar_parse_vectors():
    for v in VEC["cigar_parse"]:
        assert cig(v["text"]) == [tuple(x) for x in v["want"]], v["ref"]


def test_encode_vectors():
    for v in VEC["encode"]:
        got = O.encode(cig(v["cigar"]), v["strand"] == "+", v["qs"], v["qe"], v["ts"], v["q"].encode(), v["t"].encode())
        assert O.cigar_str(got) == v["want"], v["ref"]


def test_remove_stats_aligned_vectors():
    L = O.lib()
    for v in VEC["remove"]:
        pairs = cig(v["cigar"])
        arr = O.mk_ops(pairs)
        n = L.oracle_remove_mismatches(arr, len(pairs))
        assert O.cigar_str(O.ops_to_pairs(arr, n)) == v["want"], v["ref"]
    for v in VEC["stats"]:
        pairs = cig(v["cigar"])
        arr = O.mk_ops(pairs)
        c = (C.c_int64 * 6)()
        for _ in range(2 if v.get("twice") else 1):
            L.oracle_stats(arr, len(pairs), c)
        assert list(c) == v["want"], v["ref"]
    for v in VEC["aligned_bases"]:
        pairs = cig(v["cigar"])
        assert L.oracle_aligned_bases(O.mk_ops(pairs), len(pairs)) == v["want"], v["ref"]


def test_invert_vectors():
    for v in VEC["invert"]:
        p = make(v["rec"][:9] + [v["rec"][9]])
        p = make([*v["rec"][:8], v["rec"][8], v["rec"][9]])
        O.lib().oracle_paf_invert(C.byref(p))
        w = v["want"]
        assert [p.query_name.decode(), p.query_length, p.query_start, p.query_end, "+" if p.same_strand else "-",
                p.target_name.decode(), p.target_length, p.target_start, p.target_end, cigar_of(p)] == w, v["ref"]
        O.lib().oracle_paf_invert(C.byref(p))  # double inversion is the identity (tests/paf_unit_test.c:382-393)
        assert cigar_of(p) == v["rec"][9] and p.query_start == v["rec"][2]


def test_trim_vectors():
    L = O.lib()
    for v in VEC["trim_ends"]:
        p = make(v["rec"])
        L.oracle_trim_ends(C.byref(p), v["n"])
        assert coords(p) == v["want"][:4], v["ref"]
        if v["want"][4]:
            assert cigar_of(p) == v["want"][4], v["ref"]
    for v in VEC["trim_end_fraction"]:
        p = make(v["rec"])
        L.oracle_trim_end_fraction(C.byref(p), v["fraction"])
        assert coords(p) == v["want"][:4], v["ref"]
    for v in VEC["trim_unreliable_tails"]:
        p = make(v["rec"])
        L.oracle_trim_unreliable_tails(C.byref(p), v["score_fraction"], v["max_fraction"])
        assert coords(p) == v["want"][:4], v["ref"]
        assert cigar_of(p) == v["want"][4], v["ref"]


def test_check_valid_vectors():
    for v in VEC["check_valid"]:
        assert O.lib().oracle_paf_check(C.byref(make(v["rec"]))) == 0, v["rec"]


# ---- committed outputs of the unmodified reference CLI -------------------------------------------------------------
SETS = ["edge", "rand101", "rand102", "rand103"]


def run_oracle(stem, tmp_path, mode, src_suffix=".paf", **kw):
    fa = O.Fasta([os.path.join(G, stem + ".fa")])
    out = tmp_path / "out.paf"
    rc, stats, aligned = fa.add_mismatches_file(os.path.join(G, stem + src_suffix), str(out), mode, **kw)
    return rc, out.read_bytes(), stats, aligned


@pytest.mark.parametrize("stem", SETS)
def test_add_mismatches_matches_reference(stem, tmp_path):
    rc, got, _, _ = run_oracle(stem, tmp_path, 0)
    assert rc == 0
    assert got == open(os.path.join(G, stem + ".add.paf"), "rb").read()


@pytest.mark.parametrize("stem", SETS)
def test_remove_mismatches_matches_reference(stem, tmp_path):
    rc, got, _, _ = run_oracle(stem, tmp_path, 1, src_suffix=".add.paf")
    assert rc == 0
    assert got == open(os.path.join(G, stem + ".remove.paf"), "rb").read()


@pytest.mark.parametrize("stem", SETS[1:])
def test_add_then_trim_matches_reference(stem, tmp_path):
    rc, got, _, _ = run_oracle(stem, tmp_path, 2)
    assert rc == 0
    assert got == open(os.path.join(G, stem + ".trim.paf"), "rb").read()


@pytest.mark.parametrize("stem", SETS)
def test_view_summary_matches_reference(stem, tmp_path):
    rc, _, stats, _ = run_oracle(stem, tmp_path, 0)
    assert rc == 0
    assert util.view_summary(stats).encode() == open(os.path.join(G, stem + ".view.txt"), "rb").read()


def test_error_paths_match_reference(tmp_path):
    """A failing record ends the run with exit 1 after the earlier records were written (impl/paf_add_mismatches.c:117-133)."""
    errs = json.load(open(os.path.join(G, "errors.json")))
    fa = O.Fasta([os.path.join(G, "edge.fa")])
    for name, e in errs.items():
        inp, out = tmp_path / f"{name}.paf", tmp_path / f"{name}.out"
        inp.write_text("\n".join(e["input"]) + "\n")
        rc, _, _ = fa.add_mismatches_file(str(inp), str(out))
        assert e["exit"] == 1 and rc == 1, name
        assert out.read_text() == e["stdout"], name


def test_c1_fixture_matches_reference(tmp_path):
    """Config C1: the reference's own 207-record fixture over the seeded synthetic chr10 (124 949 228 aligned bp)."""
    info = json.load(open(os.path.join(G, "c1.json")))
    raw = gzip.open(os.path.join(G, "c1_human_chimp.paf.gz")).read()
    assert hashlib.md5(raw).hexdigest() == info["input_md5"]
    inp, out = tmp_path / "c1.paf", tmp_path / "c1.out"
    inp.write_bytes(raw)
    fa = O.Fasta()
    fa.add("chr10", util.c1_genome().tobytes())
    rc, stats, aligned = fa.add_mismatches_file(str(inp), str(out))
    got = out.read_bytes()
    assert rc == 0 and aligned == 124949228
    assert len(got) == info["add_bytes"] and hashlib.md5(got).hexdigest() == info["add_md5"]
    assert util.view_summary(stats) == info["view_summary"]
    rc, _, _ = fa.add_mismatches_file(str(inp), str(out), 2)
    got = out.read_bytes()
    assert rc == 0 and len(got) == info["trim_bytes"] and hashlib.md5(got).hexdigest() == info["trim_md5"]


@pytest.mark.skipif(not O.ref_available(), reason="compiled reference (oracle/_ref) not present")
@pytest.mark.parametrize("seed", [7, 8])
def test_live_against_compiled_reference(seed, tmp_path):
    """Where the reference is compiled: fresh random cases through both."""
    case = util.random_case(seed, n_records=40, contig_len=8000, n_ops=10, max_len=500, p_weird=0.004)
    case["records"] = [r for r in case["records"] if r["qe"] > r["qs"] and any(c != "I" and n for n, c in r["ops"])]
    fa_path, paf = tmp_path / "x.fa", tmp_path / "x.paf"
    fa_path.write_bytes(b"".join(b">" + n.encode() + b"\n" + s + b"\n" for n, s in case["contigs"].items()))
    paf.write_text("\n".join(util.case_lines(case)) + "\n")
    rc, want, err = O.ref_paffy(["add_mismatches", str(fa_path), "-i", str(paf)])
    assert rc == 0, err
    out = tmp_path / "o.paf"
    rc, _, _ = O.Fasta([str(fa_path)]).add_mismatches_file(str(paf), str(out))
    assert rc == 0 and out.read_bytes() == want
