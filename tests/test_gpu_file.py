"""GPU parity of the host layer above the C ABI: PAF text in -> CUDA kernels -> PAF text out, byte-compared with the
committed outputs of the unmodified reference CLI (tests/golden, written by tests/golden/make_golden.py)."""
import ctypes as C
import gzip
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

import paffy_b200
from tests import util

pytestmark = pytest.mark.gpu

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SETS = ["edge", "rand101", "rand102", "rand103"]


@pytest.fixture()
def fresh():
    c = paffy_b200.Context(0)
    yield c
    c.close()


def golden(name):
    return open(os.path.join(G, name), "rb").read()


@pytest.mark.parametrize("stem", SETS)
@pytest.mark.parametrize("batch_bases", [0, 3000])
def test_add_mismatches_file_matches_reference(fresh, stem, batch_bases, tmp_path):
    fresh.load_fasta(os.path.join(G, stem + ".fa"))
    out = tmp_path / "out.paf"
    # (plain add_mismatches parses the text on the device: one GPU batch per chunk of text)
    rc, rep = fresh.add_mismatches_file(os.path.join(G, stem + ".paf"), str(out), batch_bases=batch_bases, chunk_bytes=1500 if batch_bases else 0)
    assert rc == 0
    assert out.read_bytes() == golden(stem + ".add.paf")
    assert util.view_summary(list(rep.stats)).encode() == golden(stem + ".view.txt")
    if batch_bases and stem != "edge":
        assert rep.batches > 1


@pytest.mark.parametrize("stem", SETS[1:])
def test_add_then_trim_file_matches_reference(fresh, stem, tmp_path):
    fresh.load_fasta(os.path.join(G, stem + ".fa"))
    out = tmp_path / "out.paf"
    rc, rep = fresh.add_mismatches_file(os.path.join(G, stem + ".paf"), str(out), mode=2)
    assert rc == 0
    assert out.read_bytes() == golden(stem + ".trim.paf")


def test_error_paths_match_reference(fresh, tmp_path):
    """Exit code 1, the records before the failing one written, the reference's message (tests/golden/errors.json)."""
    errs = json.load(open(os.path.join(G, "errors.json")))
    fresh.load_fasta(os.path.join(G, "edge.fa"))
    for name, e in errs.items():
        inp, out = tmp_path / f"{name}.paf", tmp_path / f"{name}.out"
        inp.write_text("\n".join(e["input"]) + "\n")
        rc, rep = fresh.add_mismatches_file(str(inp), str(out))
        assert rc == 1 and e["exit"] == 1, name
        assert out.read_text() == e["stdout"], name
        assert rep.failed_record == 1
        msg = fresh.last_error()
        assert msg == e["stderr_message"], (name, msg)  # paf_check's message shows the record with its =/X CIGAR, as the reference's does
        os.environ["PAFFY_HOST_PARSE"] = "1"
        try:
            rc, rep = fresh.add_mismatches_file(str(inp), str(out))
        finally:
            del os.environ["PAFFY_HOST_PARSE"]
        assert rc == 1 and out.read_text() == e["stdout"] and fresh.last_error() == e["stderr_message"], name


def test_record_outside_contig_ends_run(fresh, tmp_path):
    """Declared lengths larger than the FASTA: the reference reads past the string (UB); here the run ends at that
    record with exit code 1 after the earlier records were written."""
    fresh.load_fasta(os.path.join(G, "edge.fa"))
    good = "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:25M"
    bad = "q\t100\t50\t75\t+\tt\t100\t50\t75\t0\t25\t60\tcg:Z:25M"
    inp, out = tmp_path / "in.paf", tmp_path / "out.paf"
    inp.write_text("\n".join([good, good, bad, good]) + "\n")
    rc, rep = fresh.add_mismatches_file(str(inp), str(out))
    assert rc == 1 and rep.failed_record == 2
    assert out.read_text().count("\n") == 2


def test_c1_fixture_matches_reference(fresh, tmp_path):
    """Config C1: the reference's own 207-record fixture over the seeded synthetic chr10 (124 949 228 aligned bp)."""
    info = json.load(open(os.path.join(G, "c1.json")))
    raw = gzip.open(os.path.join(G, "c1_human_chimp.paf.gz")).read()
    inp, out = tmp_path / "c1.paf", tmp_path / "c1.out"
    inp.write_bytes(raw)
    fresh.add_contig("chr10", util.c1_genome())
    rc, rep = fresh.add_mismatches_file(str(inp), str(out))
    got = out.read_bytes()
    assert rc == 0 and rep.aligned_bases == 124949228 and rep.records == 207
    assert len(got) == info["add_bytes"] and hashlib.md5(got).hexdigest() == info["add_md5"]
    assert util.view_summary(list(rep.stats)) == info["view_summary"]
    rc, rep = fresh.add_mismatches_file(str(inp), str(out), mode=2)
    got = out.read_bytes()
    assert rc == 0 and len(got) == info["trim_bytes"] and hashlib.md5(got).hexdigest() == info["trim_md5"]


@pytest.mark.parametrize("stem", SETS)
@pytest.mark.parametrize("batch_bases", [0, 3000])
def test_view_file_matches_reference(fresh, stem, batch_bases, tmp_path):
    """`paffy view -s X.fa -i X.paf` on the raw (M) input: per-alignment stats lines (paf_pretty_print header,
    impl/paf.c:270-279) + the aggregate line, from the kernel's per-record tallies; -s -t and neither flag too."""
    fresh.load_fasta(os.path.join(G, stem + ".fa"))
    want = golden(stem + ".viewlines.txt")
    out = tmp_path / "view.txt"
    rc, rep = fresh.add_mismatches_file(os.path.join(G, stem + ".paf"), str(out), mode=4 | 1 | 2, batch_bases=batch_bases)
    assert rc == 0 and out.read_bytes() == want
    lines = want.splitlines(keepends=True)
    rc, rep = fresh.add_mismatches_file(os.path.join(G, stem + ".paf"), str(out), mode=4 | 1, batch_bases=batch_bases)
    assert rc == 0 and out.read_bytes() == b"".join(lines[:-1])
    rc, rep = fresh.add_mismatches_file(os.path.join(G, stem + ".paf"), str(out), mode=4 | 2, batch_bases=batch_bases)
    assert rc == 0 and out.read_bytes() == lines[-1] == golden(stem + ".view.txt")


def test_view_c1_fixture_matches_reference(fresh, tmp_path):
    info = json.load(open(os.path.join(G, "c1.json")))
    inp, out = tmp_path / "c1.paf", tmp_path / "c1.view"
    inp.write_bytes(gzip.open(os.path.join(G, "c1_human_chimp.paf.gz")).read())
    fresh.add_contig("chr10", util.c1_genome())
    rc, rep = fresh.add_mismatches_file(str(inp), str(out), mode=4 | 1 | 2)
    got = out.read_bytes()
    assert rc == 0 and len(got) == info["viewlines_bytes"] and hashlib.md5(got).hexdigest() == info["viewlines_md5"]


def test_view_cli_drop_in(tmp_path):
    """`paffy_b200_cli view` takes the reference's arguments (impl/paf_view.c:39-113) and prints the reference's bytes."""
    exe = os.path.join(os.path.dirname(paffy_b200.lib.LIB_PATH), "paffy_b200_cli")
    fa, paf = os.path.join(G, "rand102.fa"), os.path.join(G, "rand102.paf")
    r = subprocess.run([exe, "view", "-s", fa, "-i", paf], capture_output=True)
    assert r.returncode == 0, r.stderr
    assert r.stdout == golden("rand102.viewlines.txt")
    r = subprocess.run([exe, "view", "--printAggregateStats", "--noPerAlignmentStats", fa], capture_output=True, input=golden("rand102.add.paf"))
    assert r.returncode == 0 and r.stdout == golden("rand102.view.txt")  # already encoded input: =/X pass through
    r = subprocess.run([exe, "view", "-i", paf], capture_output=True)
    assert r.returncode == 1 and b"Expected at least one sequence file" in r.stderr
    r = subprocess.run([exe, "view", "-t", "-u", "0.999999", fa, "-i", paf], capture_output=True)
    assert r.returncode == -6 and r.stdout == b""  # the reference's closing assert aborts (impl/paf_view.c:195)
    r = subprocess.run([exe, "view", "-s", "-t", "-u", "0.1", "-v", "10", fa, "-i", paf], capture_output=True)
    assert r.returncode == 0 and r.stdout == golden("rand102.view.txt")
    r = subprocess.run([exe, "view", "-t", "-u", "0.1", "-v", "10", fa, "-i", paf], capture_output=True)
    assert r.returncode == -6  # without -s the totals stay 0: NaN fails the assert, as in the reference
    r = subprocess.run([exe, "view", os.path.join(G, "edge.fa"), "-i", "/dev/stdin"], capture_output=True,
                       input=b"q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:25M\nq\t25\t0\t25\t+\tnope\t25\t0\t25\t0\t25\t60\tcg:Z:25M\n")
    assert r.returncode == 1 and b"No target sequence named: nope found" in r.stderr and r.stdout.count(b"\n") == 1


@pytest.mark.parametrize("stem", SETS[1:])
@pytest.mark.parametrize("batch_bases", [0, 3000])
def test_trim_alone_matches_reference(fresh, stem, batch_bases, tmp_path):
    """`paffy trim` by itself (impl/paf_trim.c:113-127; no FASTA, no encode): ops staged on the device, identity trim and
    the fixed trim `-f` (cigar_trim / cigar_trim_back, impl/paf.c:518-576: the one trim that splits ops), on encoded and
    on raw (M) input, vs the compiled reference's outputs."""
    out = tmp_path / "out.paf"
    enc, raw = os.path.join(G, stem + ".add.paf"), os.path.join(G, stem + ".paf")
    rc, rep = fresh.add_mismatches_file(enc, str(out), mode=3, batch_bases=batch_bases)
    assert rc == 0 and out.read_bytes() == golden(stem + ".trim.paf")
    rc, rep = fresh.add_mismatches_file(enc, str(out), mode=3, trim_r=-1.0, trim_t=0.3, batch_bases=batch_bases)
    assert rc == 0 and out.read_bytes() == golden(stem + ".trimf.paf")
    rc, rep = fresh.add_mismatches_file(raw, str(out), mode=3, trim_r=0.2, trim_t=0.5, batch_bases=batch_bases)
    assert rc == 0 and out.read_bytes() == golden(stem + ".trimraw.paf")
    rc, rep = fresh.add_mismatches_file(raw, str(out), mode=3, trim_r=-1.0, trim_t=0.5, batch_bases=batch_bases)
    assert rc == 0 and out.read_bytes() == golden(stem + ".trimfraw.paf")
    # add_mismatches | trim -f in one pass (mode 2 with the fixed trim)
    fresh.load_fasta(os.path.join(G, stem + ".fa"))
    rc, rep = fresh.add_mismatches_file(raw, str(out), mode=2, trim_r=-1.0, trim_t=0.3, batch_bases=batch_bases)
    assert rc == 0 and out.read_bytes() == golden(stem + ".trimf.paf")


def test_trim_c1_and_cli(fresh, tmp_path):
    """C1 (207 records, 20 k-op CIGARs): add_mismatches on the GPU, then `paffy_b200_cli trim` / `trim -f -t 0.3` on that
    file, md5 against the reference's `add_mismatches | trim`; error paths of the stand-alone trim."""
    info = json.load(open(os.path.join(G, "c1.json")))
    inp, enc, out = tmp_path / "c1.paf", tmp_path / "c1.add.paf", tmp_path / "c1.trim.paf"
    inp.write_bytes(gzip.open(os.path.join(G, "c1_human_chimp.paf.gz")).read())
    fresh.add_contig("chr10", util.c1_genome())
    rc, rep = fresh.add_mismatches_file(str(inp), str(enc))
    assert rc == 0
    exe = os.path.join(os.path.dirname(paffy_b200.lib.LIB_PATH), "paffy_b200_cli")
    r = subprocess.run([exe, "trim", "-i", str(enc), "-o", str(out)], capture_output=True)
    got = out.read_bytes()
    assert r.returncode == 0 and len(got) == info["trim_bytes"] and hashlib.md5(got).hexdigest() == info["trim_md5"], r.stderr
    r = subprocess.run([exe, "trim", "--fixedTrim", "--trimFraction", "0.3"], capture_output=True, input=enc.read_bytes())
    assert r.returncode == 0 and len(r.stdout) == info["trimf_bytes"] and hashlib.md5(r.stdout).hexdigest() == info["trimf_md5"]
    # paf_check after the trim: a CIGAR one base short of its coordinates ends the run after the records before it
    good = "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:20=1X4="
    bad = "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:20=1X3="
    r = subprocess.run([exe, "trim"], capture_output=True, input=("\n".join([good, good, bad, good]) + "\n").encode())
    assert r.returncode == 1 and r.stdout.count(b"\n") == 2
    assert b"ERROR: Paf cigar alignment does not match query length: 24 vs. 25" in r.stderr
    r = subprocess.run([exe, "trim", "-x"], capture_output=True)
    assert r.returncode == 1


@pytest.mark.parametrize("stem", SETS)
@pytest.mark.parametrize("batch_bases", [0, 3000])
def test_add_then_filter_matches_reference(fresh, stem, batch_bases, tmp_path):
    """`paffy add_mismatches | paffy filter` fused (impl/paf_filter.c:123-134 on the kernel's tallies): identity, identity
    with gaps inverted, alignment score / tile level; defaults drop only the records without aligned bases (0/0)."""
    fresh.load_fasta(os.path.join(G, stem + ".fa"))
    inp, out = os.path.join(G, stem + ".paf"), tmp_path / "out.paf"
    rc, rep = fresh.add_mismatches_filter_file(inp, str(out), min_identity=0.5, batch_bases=batch_bases)
    assert rc == 0 and out.read_bytes() == golden(stem + ".filter_u.paf")
    rc, rep = fresh.add_mismatches_filter_file(inp, str(out), min_identity_with_gaps=0.4, invert=True, batch_bases=batch_bases)
    assert rc == 0 and out.read_bytes() == golden(stem + ".filter_vx.paf")
    rc, rep = fresh.add_mismatches_filter_file(inp, str(out), min_alignment_score=1, max_tile_level=3, batch_bases=batch_bases)
    assert rc == 0 and out.read_bytes() == golden(stem + ".filter_t.paf")
    assert util.view_summary(list(rep.stats)).encode() == golden(stem + ".view.txt")  # the totals cover every record
    want = golden(stem + ".add.paf")
    gu, gvx = golden(stem + ".filter_u.paf"), golden(stem + ".filter_vx.paf")
    assert 0 < gu.count(b"\n") < want.count(b"\n") and 0 < gvx.count(b"\n") < want.count(b"\n")


def test_device_fasta_reader_matches_oracle(fresh, tmp_path):
    """pafgpu_genome_load_fasta: record boundaries on the host (memchr), separator removal + admission test + compaction on
    the device (fasta_count_kernel / fasta_compact_kernel), against the oracle's restatement of fastaReadToFunction
    (bioioC.c:155-209): text before the first '>', whole-line names, '>' in the middle of a line, spaces and tabs, ragged
    lines, tiles that end inside a line, an empty record, a repeated name, EOF right after a header, soft-masked / IUPAC /
    '-' bytes; then the inadmissible-byte error."""
    import numpy as np
    from oracle import pyoracle as O
    rng = np.random.default_rng(5)
    alpha = np.frombuffer(b"ACGTacgtNnRYKMSWBDHV-", dtype=np.uint8)
    def seq(n):
        return alpha[rng.integers(0, len(alpha) if n < 3000 else 8, n)].tobytes()
    def wrap(b, widths):
        out, i, k = [], 0, 0
        while i < len(b):
            w = widths[k % len(widths)]; k += 1
            out.append(b[i:i + w]); i += w
        return b"\n".join(out) + b"\n"
    big = seq(300000)
    fa = (b"junk before the first header\n\n>c1 with a description > and a bracket\n" + wrap(seq(777), [60]) +
          b">empty\n>c2\n" + wrap(seq(5000), [1, 7, 4096, 13]).replace(b"\n", b" \n\t", 3) + b"ACGT>c3 starts mid-line\n" +
          wrap(big, [100]) + b">c2\n" + wrap(seq(4096 * 3), [4095, 4097]) + b">last header at EOF")
    path = tmp_path / "quirks.fa"
    path.write_bytes(fa)
    want = O.Fasta([str(path)])
    fresh.load_fasta(str(path))
    assert len(want) == 5
    for name in ["c1 with a description > and a bracket", "empty", "c2", "c3 starts mid-line", "last header at EOF"]:
        w = want.get(name)
        cid = fresh.find(name)
        assert w is not None and cid >= 0, name
        assert fresh.contig_length(cid) == len(w), name
        if len(w):
            assert fresh.fetch_bases(cid, 0, len(w)) == w.upper(), name
    assert fresh.find("c1") < 0
    bad = tmp_path / "bad.fa"
    bad.write_bytes(b">ok\nACGT\n>broken\n" + wrap(seq(9000), [80]) + b"AC\rGT\nAC1GT\n")
    with pytest.raises(paffy_b200.lib.PafGpuError) as e:
        fresh.load_fasta(str(bad))
    assert e.value.code == -5 and "unexpected character in input fasta sequence: '\r'" in str(e.value)
    assert fresh.find("ok") >= 0 and fresh.find("broken") < 0


def test_cli_drop_in(tmp_path):
    """`paffy_b200_cli add_mismatches` takes the reference's arguments and prints the reference's bytes."""
    exe = os.path.join(os.path.dirname(paffy_b200.lib.LIB_PATH), "paffy_b200_cli")
    r = subprocess.run([exe, "add_mismatches", os.path.join(G, "rand101.fa"), "-i", os.path.join(G, "rand101.paf")], capture_output=True)
    assert r.returncode == 0, r.stderr
    assert r.stdout == golden("rand101.add.paf")
    out = tmp_path / "o.paf"
    r = subprocess.run([exe, "add_mismatches", "-a", "-i", os.path.join(G, "rand101.add.paf"), "-o", str(out)], capture_output=True)
    assert r.returncode == 0 and out.read_bytes() == golden("rand101.remove.paf")
    r = subprocess.run([exe, "add_mismatches", os.path.join(G, "edge.fa"), "-i", "/dev/stdin"], capture_output=True,
                       input=b"nope\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:25M\n")
    assert r.returncode == 1 and b"No query sequence named: nope found" in r.stderr


class Cigar(C.Structure):
    _fields_ = [("recs", C.POINTER(C.c_uint64)), ("length", C.c_int64), ("start", C.c_int64), ("capacity", C.c_int64)]


class Paf(C.Structure):
    _fields_ = [("query_name", C.c_char_p), ("query_length", C.c_int64), ("query_start", C.c_int64), ("query_end", C.c_int64),
                ("target_name", C.c_char_p), ("target_length", C.c_int64), ("target_start", C.c_int64), ("target_end", C.c_int64),
                ("cigar", C.POINTER(Cigar)), ("cigar_string", C.c_char_p), ("score", C.c_int64), ("mapping_quality", C.c_int64),
                ("num_matches", C.c_int64), ("num_bases", C.c_int64), ("tile_level", C.c_int64), ("chain_id", C.c_int64),
                ("chain_score", C.c_int64), ("same_strand", C.c_bool), ("type", C.c_char)]


def test_paf_encode_mismatches_symbol():
    """The reference-named entry point on the reference's struct layout (inc/paf.h:61-109), incl. cigar->start and
    ownership (recs replaced by a malloc'd array)."""
    assert C.sizeof(Paf) == 144 and C.sizeof(Cigar) == 32
    lib = paffy_b200.load()
    libc = C.CDLL(None)
    libc.malloc.restype = C.c_void_p
    libc.malloc.argtypes = [C.c_size_t]
    q, t = b"ACGTNnacgtMRWSYKVHDBXACGT", b"ACGTNNACGTKYWSRMBDHVXACGA"
    ops = util.words([(7, "I"), (25, "M")])  # start = 1 skips the first op, like a prefix-trimmed cigar
    mem = libc.malloc(16)
    C.memmove(mem, ops.ctypes.data, 16)
    cg = Cigar(C.cast(mem, C.POINTER(C.c_uint64)), 1, 1, 2)
    p = Paf(b"q", 25, 0, 25, b"t", 25, 0, 25, C.pointer(cg), None, 0, 60, 0, 0, -1, -1, -1, True, b"P")
    lib.paf_encode_mismatches.argtypes = [C.POINTER(Paf), C.c_char_p, C.c_char_p]
    lib.paf_encode_mismatches.restype = None
    lib.paf_encode_mismatches(C.byref(p), q, t)
    got = util.pairs_of([cg.recs[i] for i in range(cg.length)])
    from oracle import pyoracle as O
    assert cg.start == 0 and O.cigar_str(got) == "10=2X2=6X4=1X"
    p.same_strand = False
    C.memmove(mem := libc.malloc(8), util.words([(25, "M")]).ctypes.data, 8)
    cg2 = Cigar(C.cast(mem, C.POINTER(C.c_uint64)), 1, 0, 1)
    p.cigar = C.pointer(cg2)
    lib.pafgpu_paf_encode_mismatches.argtypes = [C.POINTER(Paf), C.c_char_p, C.c_char_p]
    lib.pafgpu_paf_encode_mismatches.restype = C.c_int64
    mm = lib.pafgpu_paf_encode_mismatches(C.byref(p), q, t)
    want = O.encode([(25, "M")], False, 0, 25, 0, q, t)
    assert util.pairs_of([cg2.recs[i] for i in range(cg2.length)]) == want
    assert mm == sum(n for n, c in want if c == "X")


@pytest.mark.parametrize("by_warp", ["0", "1"])
@pytest.mark.parametrize("seed", [71, 72, 73])
def test_device_trim_matches_oracle(ctx, seed, by_warp, monkeypatch):
    """pafgpu_trim (one thread per record, or one warp per record for long CIGARs: both forced here; float32 identity walk)
    vs the oracle's paf_trim_unreliable_tails: moved coordinates and kept op counts, both strands, several -r / -t."""
    import numpy as np
    from oracle import pyoracle as O
    monkeypatch.setenv("PAFGPU_TRIM_WARP", by_warp)
    case = util.random_case(seed, n_records=80, contig_len=30000, n_ops=14 if by_warp == "0" else 90, max_len=900 if by_warp == "0" else 200, p_main=0.97)
    case["records"] = [r for r in case["records"] if any(c == "M" and n for n, c in r["ops"])]
    got, res = util.gpu_encode(ctx, case)
    lines = util.case_lines(case)
    t_end = np.array([r["ts"] + sum(n for n, c in r["ops"] if c != "I") for r in case["records"]], dtype=np.int64)
    for sf, mf in [(0.05, 1.0), (0.3, 0.5), (0.01, 0.2)]:
        coords, kept, failed = ctx.trim(t_end, sf, mf)
        for i, (line, (enc, _)) in enumerate(zip(lines, got)):
            rc, p = O.parse(line.split("\tcg:Z:")[0] + "\tcg:Z:" + O.cigar_str(enc))
            assert rc == 0
            O.lib().oracle_trim_unreliable_tails(C.byref(p), sf, mf)
            want = [p.query_start, p.query_end, p.target_start, p.target_end]
            assert [int(coords[k][i]) for k in range(4)] == want, (i, sf, mf)
            assert int(kept[i]) == p.n_ops and not failed[i]


# ---- round 2: the streaming pipeline, one job over the GPUs of the box, the persisted packed genome --------------------
@pytest.mark.parametrize("stem", SETS[1:])
@pytest.mark.parametrize("mode,suffix", [(0, ".add.paf"), (2, ".trim.paf"), (4 | 1 | 2, ".viewlines.txt")])
def test_small_chunks_and_several_streams_give_the_same_bytes(fresh, stem, mode, suffix, tmp_path):
    """Chunks of a few lines dealt to 3 sibling contexts (shared genome, own streams) and written in input order."""
    fresh.load_fasta(os.path.join(G, stem + ".fa"))
    out = tmp_path / "out.paf"
    src = stem + ".paf"
    rc, rep = fresh.add_mismatches_file(os.path.join(G, src), str(out), mode=mode, streams=3, chunk_bytes=600)
    assert rc == 0
    assert out.read_bytes() == golden(stem + suffix)
    assert rep.chunks >= 3 and rep.workers == 3 and rep.gpus == 1


def test_one_job_sharded_over_the_gpus_of_the_box(fresh, tmp_path):
    """contiguous record ranges dealt to the GPUs (genome cloned GPU to GPU), outputs concatenated in input order,
    byte-equal to the reference's single-process output. With one visible GPU the same code runs with one device."""
    n_dev = paffy_b200.load().pafgpu_device_count()
    want_gpus = min(2, n_dev)
    for stem in SETS[1:]:
        c = paffy_b200.Context(0)
        try:
            c.load_fasta(os.path.join(G, stem + ".fa"))
            out = tmp_path / (stem + ".out")
            rc, rep = c.add_mismatches_file(os.path.join(G, stem + ".paf"), str(out), gpus=want_gpus, streams=2, chunk_bytes=900)
            assert rc == 0 and rep.gpus == want_gpus and rep.workers == 2 * want_gpus and rep.chunks >= 3
            assert out.read_bytes() == golden(stem + ".add.paf")
            assert util.view_summary(list(rep.stats)).encode() == golden(stem + ".view.txt")
        finally:
            c.close()


def test_genome_clone_between_contexts(tmp_path):
    """pafgpu_genome_clone: a context on (another) device encodes against the cloned genome exactly like the source."""
    L = paffy_b200.load()
    n_dev = L.pafgpu_device_count()
    a, b = paffy_b200.Context(0), paffy_b200.Context(n_dev - 1)
    try:
        a.load_fasta(os.path.join(G, "rand103.fa"))
        assert L.pafgpu_genome_clone(b.h, a.h) == 0
        assert b.num_contigs() == a.num_contigs()
        out = tmp_path / "o.paf"
        rc, rep = b.add_mismatches_file(os.path.join(G, "rand103.paf"), str(out))
        assert rc == 0 and out.read_bytes() == golden("rand103.add.paf")
    finally:
        a.close()
        b.close()


def test_packed_genome_save_and_load(tmp_path):
    """the packed arena written once and read back by a later process gives byte-identical output; a
    differing key or a damaged file is refused."""
    L = paffy_b200.load()
    L.pafgpu_genome_save.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
    L.pafgpu_genome_load_packed.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
    path = str(tmp_path / "g.pafb200").encode()
    for stem in ("edge", "rand102"):
        a = paffy_b200.Context(0)
        a.load_fasta(os.path.join(G, stem + ".fa"))
        assert L.pafgpu_genome_save(a.h, path, b"key-1") == 0
        n = a.num_contigs()
        a.close()
        b = paffy_b200.Context(0)
        try:
            assert L.pafgpu_genome_load_packed(b.h, path, b"another key") == -7  # PAFGPU_ESTATE
            assert L.pafgpu_genome_load_packed(b.h, path, b"key-1") == 0
            assert b.num_contigs() == n
            out = tmp_path / "o.paf"
            rc, rep = b.add_mismatches_file(os.path.join(G, stem + ".paf"), str(out))
            assert rc == 0 and out.read_bytes() == golden(stem + ".add.paf")
            blob = open(path, "rb").read()
            open(path, "wb").write(blob[: len(blob) // 2])
            assert L.pafgpu_genome_load_packed(b.h, path, b"key-1") == -6  # PAFGPU_EIO
        finally:
            b.close()


def test_failure_ordering_across_chunks(fresh, tmp_path):
    """A bad record in a late chunk: every record before it is written, nothing after it, whatever chunk finished first."""
    fresh.load_fasta(os.path.join(G, "edge.fa"))
    good = "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:25M"
    bad = "nosuch\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:25M"
    inp, out = tmp_path / "in.paf", tmp_path / "out.paf"
    inp.write_text("\n".join([good] * 211 + [bad] + [good] * 300) + "\n")
    rc, rep = fresh.add_mismatches_file(str(inp), str(out), streams=3, chunk_bytes=700)
    assert rc == 1 and rep.failed_record == 211 and rep.records == 211
    assert out.read_text().count("\n") == 211
    assert "No query sequence named: nosuch found" in fresh.last_error()


# ---- round 2: PAF text parsed and rendered on the device ------------------------------------------
def _decorated_lines(case):
    """Lines of a random case dressed up with everything paf_parse tolerates (impl/paf.c:137-209): extra and malformed tags,
    tags in any order, runs of tabs, no cg tag, an empty cg value, a repeated cg tag, 0M ops, negative numbers."""
    out = []
    for i, line in enumerate(util.case_lines(case)):
        f = line.split("\t")
        head, cg = f[:12], f[-1]
        k = i % 11
        if k == 1:
            f = head + ["NM:i:5", "zz:Z:abc", "tp:A:P", cg]
        elif k == 2:
            f = head + [cg, "tp:A:S", "tl:i:3"]
        elif k == 3:
            f = head[:3] + ["", ""] + head[3:] + ["", "tp:A:I", cg, ""]
        elif k == 4:
            f = head + ["tp:A:P"]
        elif k == 5:
            f = head + ["cg:Z:", "AS:i:77"]
        elif k == 6:
            f = head + ["cg:Z:5M", "tp:A:P", cg]
        elif k == 7:
            f = head + ["AS:i:-12", "cn:i:7", "s1:i:99", cg]
        elif k == 8:
            f = head + ["tp:A:P", cg.replace("M", "M0M", 1) + "0M"]
        elif k == 9:
            f = head + ["x", "ab:c", "tl:i:1", cg]
        elif k == 10:
            f = head + ["tl:i:2", cg]
        out.append("\t".join(f))
    return out


@pytest.mark.parametrize("seed", [5, 6])
def test_device_text_path_matches_the_reference_on_decorated_lines(fresh, seed, tmp_path):
    """The device parser / renderer against the compiled reference CLI (the oracle's file filter where the reference binary
    is absent), and against this library's own host-parse path: same bytes, same totals."""
    from oracle import pyoracle as O
    case = util.random_case(seed, n_records=150, contig_len=30000, n_ops=14, max_len=900, p_weird=0.002)
    fa, paf = tmp_path / "g.fa", tmp_path / "in.paf"
    with open(fa, "wb") as fh:
        for name, seq in case["contigs"].items():
            fh.write(b">" + name.encode() + b"\n" + seq + b"\n")
    lines = _decorated_lines(case)
    paf.write_text("\n".join(lines))  # no newline behind the last line
    want = tmp_path / "want.paf"
    if O.ref_available():
        r = subprocess.run([O.REF_PAFFY, "add_mismatches", str(fa), "-i", str(paf), "-o", str(want)], capture_output=True)
        assert r.returncode == 0, r.stderr[-500:]
    else:
        rc, _, _ = O.Fasta([str(fa)]).add_mismatches_file(str(paf), str(want))
        assert rc == 0
    fresh.load_fasta(str(fa))
    out = tmp_path / "out.paf"
    rc, rep = fresh.add_mismatches_file(str(paf), str(out), chunk_bytes=5000)
    assert rc == 0 and rep.records == len(lines) and rep.chunks > 3
    assert out.read_bytes() == want.read_bytes()
    os.environ["PAFFY_HOST_PARSE"] = "1"
    try:
        out2 = tmp_path / "out2.paf"
        rc, rep2 = fresh.add_mismatches_file(str(paf), str(out2), chunk_bytes=5000)
    finally:
        del os.environ["PAFFY_HOST_PARSE"]
    assert rc == 0 and out2.read_bytes() == want.read_bytes()
    assert list(rep2.stats) == list(rep.stats) and rep2.aligned_bases == rep.aligned_bases
    # add_mismatches | trim and the totals of view -s -t: device text path against the host-parse path and the reference.
    # `paffy trim` aborts on a record without ops (its identity is 0 / 0, impl/paf.c:952): lines 4 and 5 of every 11 have
    # no cigar, so the whole file ends at record 4 in both paths; the comparison with the reference runs without them.
    for host_parse in (False, True):
        if host_parse:
            os.environ["PAFFY_HOST_PARSE"] = "1"
        try:
            o = tmp_path / "cut.out"
            rc, r = fresh.add_mismatches_file(str(paf), str(o), mode=2, chunk_bytes=5000)
            assert rc == 1 and r.failed_record == 4 and o.read_bytes().count(b"\n") == 4
            assert "needs a cigar" in fresh.last_error()
        finally:
            os.environ.pop("PAFFY_HOST_PARSE", None)
    with_cigar = [ln for i, ln in enumerate(lines)
                  if i % 11 not in (4, 5) and any(not (c == "M" and n == 0) for n, c in case["records"][i]["ops"])]
    paf.write_text("\n".join(with_cigar) + "\n")
    r = subprocess.run([O.REF_PAFFY, "add_mismatches", str(fa), "-i", str(paf), "-o", str(want)], capture_output=True) if O.ref_available() else None
    for mode, ref_cmd in ((2, ["trim"]), (4 | 2, ["view", "-s", "-t", str(fa)])):
        got = {}
        for host_parse in (False, True):
            if host_parse:
                os.environ["PAFFY_HOST_PARSE"] = "1"
            try:
                o = tmp_path / f"m{mode}_{int(host_parse)}.out"
                rc, r = fresh.add_mismatches_file(str(paf), str(o), mode=mode, chunk_bytes=5000)
                assert rc == 0
                got[host_parse] = (o.read_bytes(), list(r.stats))
            finally:
                os.environ.pop("PAFFY_HOST_PARSE", None)
        assert got[False] == got[True], mode
        if O.ref_available():
            r = subprocess.run([O.REF_PAFFY] + ref_cmd + ["-i", str(want)], capture_output=True)
            assert r.returncode == 0 and r.stdout == got[False][0], mode


@pytest.mark.parametrize("seed", [8, 9])
def test_device_op_parser_on_numbers_of_any_width(fresh, seed, tmp_path):
    """cigar_parse (impl/paf.c:70-111) reads any run of digits, leading zeros included, and a bare letter as length 0. The
    device parser classifies 16 bytes per thread: numbers dressed with up to 40 leading zeros cross one or two of those
    windows, bare letters and 0-length ops of every kind sit at every offset."""
    import re
    from oracle import pyoracle as O
    rng = np.random.default_rng(seed)
    case = util.random_case(seed, n_records=400, contig_len=30000, n_ops=30, max_len=700, p_weird=0.002)
    fa, paf = tmp_path / "g.fa", tmp_path / "in.paf"
    with open(fa, "wb") as fh:
        for name, seq in case["contigs"].items():
            fh.write(b">" + name.encode() + b"\n" + seq + b"\n")

    def dress(m):
        n, c = m.group(1), m.group(2)
        if int(n) == 0 and rng.random() < 0.5:
            return c if rng.random() < 0.5 else "0" * int(rng.integers(1, 20)) + c  # length 0: bare letter, or zeros only
        return "0" * int(rng.integers(0, 41) if rng.random() < 0.3 else 0) + n + c

    lines = []
    for ln in util.case_lines(case):
        head, cg = ln.rsplit("\tcg:Z:", 1)
        lines.append(head + "\tcg:Z:" + re.sub(r"(\d+)([MID=X])", dress, cg))
    paf.write_text("\n".join(lines) + "\n")
    want = tmp_path / "want.paf"
    if O.ref_available():
        r = subprocess.run([O.REF_PAFFY, "add_mismatches", str(fa), "-i", str(paf), "-o", str(want)], capture_output=True)
        assert r.returncode == 0, r.stderr[-500:]
    else:
        rc, _, _ = O.Fasta([str(fa)]).add_mismatches_file(str(paf), str(want))
        assert rc == 0
    fresh.load_fasta(str(fa))
    for chunk_bytes in (0, 7001):
        out = tmp_path / f"out{chunk_bytes}.paf"
        rc, rep = fresh.add_mismatches_file(str(paf), str(out), chunk_bytes=chunk_bytes)
        assert rc == 0 and rep.records == len(lines)
        assert out.read_bytes() == want.read_bytes()


def test_device_text_path_locates_the_record_the_reference_aborts_on(fresh, tmp_path):
    """Every kind of failing record, each behind good ones and in a late chunk: same stdout, record index and message as the
    host-parse path (which tests/golden/errors.json pins to the reference)."""
    fresh.load_fasta(os.path.join(G, "edge.fa"))
    good = "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:25M"
    bads = ["q\t25\t0\t25\t*\tt\t25\t0\t25\t0\t25\t60\tcg:Z:25M",          # strand
            "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:20M5Q",        # cigar letter
            "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:20M5",         # digits without a letter
            "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\ttp:A:Z\tcg:Z:25M",  # tp value
            "q\t25\t0\t25\t+\tt\t25\t0\t25",                               # too few fields
            "nosuch\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:25M",     # unknown query
            "q\t25\t0\t25\t+\tnosuch\t25\t0\t25\t0\t25\t60\tcg:Z:25M",     # unknown target
            "q\t25\t0\t10\t+\tt\t25\t0\t9\t0\t0\t60\tcg:Z:9M",             # paf_check: query sum
            "q\t25\t0\t9\t+\tt\t25\t0\t10\t0\t0\t60\tcg:Z:9M",             # paf_check: target sum
            "q\t25\t30\t35\t+\tt\t25\t0\t5\t0\t0\t60\tcg:Z:5M",            # paf_check: coordinates
            "q\t100\t50\t75\t+\tt\t100\t50\t75\t0\t25\t60\tcg:Z:25M"]      # walks outside the contig
    for bad in bads:
        inp = tmp_path / "in.paf"
        inp.write_text("\n".join([good] * 57 + [bad] + [good] * 40) + "\n")
        got = {}
        for host_parse in (False, True):
            if host_parse:
                os.environ["PAFFY_HOST_PARSE"] = "1"
            try:
                out = tmp_path / f"out{int(host_parse)}.paf"
                rc, rep = fresh.add_mismatches_file(str(inp), str(out), chunk_bytes=900)
                got[host_parse] = (rc, rep.failed_record, rep.records, out.read_bytes(), fresh.last_error().replace("record 57", "record N").split("record")[0])
            finally:
                os.environ.pop("PAFFY_HOST_PARSE", None)
        assert got[False][0] == 1 and got[False][1] == 57 and got[False][2] == 57, bad
        assert got[False][3].count(b"\n") == 57
        assert got[False][:4] == got[True][:4], bad
        assert got[False][4] == got[True][4], (bad, got[False][4], got[True][4])


# ---- round 2: view -a, the closing asserts of view, the persisted genome behind the CLI ---------------------------------
CLI = os.path.join(os.path.dirname(paffy_b200.lib.LIB_PATH), "paffy_b200_cli")


@pytest.mark.parametrize("stem", SETS)
def test_view_with_alignment_matches_reference(stem, tmp_path):
    """`paffy view -a -s`: per-alignment lines, the three-row alignment blocks in the FASTA's original case, the aggregate
    line - byte-equal to the compiled reference's output."""
    fa, paf = os.path.join(G, stem + ".fa"), os.path.join(G, stem + ".paf")
    r = subprocess.run([CLI, "view", "-a", "-s", fa, "-i", paf], capture_output=True)
    assert r.returncode == 0, r.stderr[-500:]
    assert r.stdout == golden(stem + ".viewa.txt")
    out = tmp_path / "o.txt"
    r = subprocess.run([CLI, "view", "-a", "-s", fa, "-i", paf, "-o", str(out)], capture_output=True)
    assert r.returncode == 0 and out.read_bytes() == golden(stem + ".viewa.txt")


def test_view_without_aggregate_stats_aborts_like_the_reference():
    """impl/paf_view.c:195: the totals are only accumulated under -s; without it the identity assert sees NaN and the
    reference aborts. Same exit (SIGABRT) here."""
    fa, paf = os.path.join(G, "edge.fa"), os.path.join(G, "edge.add.paf")
    r = subprocess.run([CLI, "view", fa, "-i", paf], capture_output=True)
    assert r.returncode == -6 and b"Assertion" in r.stderr
    r = subprocess.run([CLI, "view", "-s", fa, "-i", paf], capture_output=True)
    assert r.returncode == 0


def test_cli_uses_the_packed_genome_cache(tmp_path):
    """$PAFFY_B200_CACHE: the first run parses the FASTA and leaves the packed arena, the second one loads it; a FASTA whose
    size / mtime changed is parsed again. Same bytes out every time."""
    import shutil
    cache = tmp_path / "cache"
    cache.mkdir()
    fa = tmp_path / "g.fa"
    shutil.copy(os.path.join(G, "rand101.fa"), fa)
    env = dict(os.environ, PAFFY_B200_CACHE=str(cache))
    args = [CLI, "add_mismatches", str(fa), "-i", os.path.join(G, "rand101.paf"), "-l", "INFO"]
    r1 = subprocess.run(args, capture_output=True, env=env)
    assert r1.returncode == 0 and r1.stdout == golden("rand101.add.paf") and b"Parsing sequence file" in r1.stderr
    assert len(list(cache.iterdir())) == 1
    r2 = subprocess.run(args, capture_output=True, env=env)
    assert r2.returncode == 0 and r2.stdout == golden("rand101.add.paf")
    assert b"packed sequences" in r2.stderr and b"Parsing sequence file" not in r2.stderr
    os.utime(fa, (1, 1))
    r3 = subprocess.run(args, capture_output=True, env=env)
    assert r3.returncode == 0 and r3.stdout == golden("rand101.add.paf") and b"Parsing sequence file" in r3.stderr


def test_paf_encode_mismatches_keeps_contigs_resident():
    """The library symbol called record after record with the same whole-contig strings (the loop of `paffy view` / Cactus):
    results equal the oracle's every time, on both strands, and a different string behind a reused pointer is noticed."""
    from oracle import pyoracle as O
    from tests.test_paf_api import Paf
    lib = paffy_b200.load()
    lib.paf_parse.restype = C.POINTER(Paf)
    lib.paf_parse.argtypes = [C.c_char_p, C.c_bool]
    lib.paf_encode_mismatches.argtypes = [C.POINTER(Paf), C.c_char_p, C.c_char_p]
    lib.paf_encode_mismatches.restype = None
    lib.paf_print.restype = C.c_void_p
    lib.paf_print.argtypes = [C.POINTER(Paf)]
    case = util.random_case(91, n_records=60, contig_len=20000, n_ops=10, max_len=700, p_weird=0.002)
    bufs = {name: C.create_string_buffer(seq, len(seq) + 1) for name, seq in case["contigs"].items()}
    want = util.oracle_encode(case)
    for rnd in range(2):
        for r, line, (enc, _) in zip(case["records"], util.case_lines(case), want):
            p = lib.paf_parse(C.create_string_buffer(line.encode()), True)
            lib.paf_encode_mismatches(p, bufs[r["q"]], bufs[r["t"]])
            s = lib.paf_print(p)
            got = C.string_at(s).decode()
            assert got.split("cg:Z:")[1] == O.cigar_str(enc), (rnd, line)
    # the same buffer now holds another sequence: the fingerprint of its head changes, the contig is packed again
    name = case["records"][0]["q"]
    other = bytes(reversed(case["contigs"][name]))
    C.memmove(bufs[name], other, len(other))
    case["contigs"][name] = other
    want2 = util.oracle_encode(case)
    for r, line, (enc, _) in zip(case["records"], util.case_lines(case), want2):
        p = lib.paf_parse(C.create_string_buffer(line.encode()), True)
        lib.paf_encode_mismatches(p, bufs[r["q"]], bufs[r["t"]])
        assert C.string_at(lib.paf_print(p)).decode().split("cg:Z:")[1] == O.cigar_str(enc)
