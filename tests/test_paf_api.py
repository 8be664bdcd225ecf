"""The reference's paf.h library surface as exported by libpaffy_b200.so (include/paf_b200_compat.h).

CPU: (1) the reference's own CuTest cases (tests/paf_unit_test.c + test_paf of tests/paf_test.c, compiled UNCHANGED by
`make -C oracle linkproof`) run against the library - every case that needs no base comparison; (2) the reference's
unchanged paffy_main.c linked against the library instead of impl/paf.c / paf_add_mismatches.c / paf_view.c / paf_trim.c
prints the same bytes as the reference binary for the subcommands that need no GPU; (3) the exported functions called
through ctypes on the reference's struct layout against the committed goldens of the compiled reference.
GPU (-m gpu): all 40 + 1 reference cases, and the linked binary's add_mismatches / view / trim against the goldens."""
import ctypes as C
import gzip
import os
import shutil
import subprocess

import pytest

import paffy_b200
from oracle import pyoracle as O

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

PAF_H_SYMBOLS = ["cigar_parse", "cigar_destruct", "paf_destruct", "paf_parse", "paf_read", "paf_read2", "paf_read_with_buffer",
                 "paf_print", "paf_write", "paf_write_with_buffer", "paf_stats_calc", "paf_pretty_print", "paf_check", "paf_invert",
                 "paf_get_number_of_aligned_bases", "paf_trim_ends", "paf_trim_end_fraction", "paf_remove_mismatches",
                 "paf_trim_unreliable_tails", "paf_trim_unreliable_prefix", "paf_shatter2", "cigar_number_of_records",
                 "paf_encode_mismatches"]


class Cigar(C.Structure):
    _fields_ = [("recs", C.POINTER(C.c_uint64)), ("length", C.c_int64), ("start", C.c_int64), ("capacity", C.c_int64)]


class Paf(C.Structure):  # inc/paf.h:88-109
    _fields_ = [("query_name", C.c_char_p), ("query_length", C.c_int64), ("query_start", C.c_int64), ("query_end", C.c_int64),
                ("target_name", C.c_char_p), ("target_length", C.c_int64), ("target_start", C.c_int64), ("target_end", C.c_int64),
                ("cigar", C.POINTER(Cigar)), ("cigar_string", C.c_char_p), ("score", C.c_int64), ("mapping_quality", C.c_int64),
                ("num_matches", C.c_int64), ("num_bases", C.c_int64), ("tile_level", C.c_int64), ("chain_id", C.c_int64),
                ("chain_score", C.c_int64), ("same_strand", C.c_bool), ("type", C.c_char)]


@pytest.fixture(scope="module")
def L():
    lib = paffy_b200.load()
    P = C.POINTER(Paf)
    lib.paf_parse.restype = P
    lib.paf_parse.argtypes = [C.c_char_p, C.c_bool]
    lib.paf_print.restype = C.c_void_p
    lib.paf_print.argtypes = [P]
    lib.paf_destruct.argtypes = [P]
    lib.paf_destruct.restype = None
    for name in ("paf_invert", "paf_remove_mismatches", "paf_check"):
        getattr(lib, name).argtypes = [P]
        getattr(lib, name).restype = None
    lib.paf_trim_unreliable_tails.argtypes = [P, C.c_float, C.c_float]
    lib.paf_trim_unreliable_tails.restype = None
    lib.paf_trim_end_fraction.argtypes = [P, C.c_float]
    lib.paf_trim_end_fraction.restype = None
    lib.paf_trim_ends.argtypes = [P, C.c_int64]
    lib.paf_trim_ends.restype = None
    lib.paf_get_number_of_aligned_bases.argtypes = [P]
    lib.paf_get_number_of_aligned_bases.restype = C.c_int64
    lib.paf_stats_calc.argtypes = [P] + [C.POINTER(C.c_int64)] * 6 + [C.c_bool]
    lib.paf_stats_calc.restype = None
    return lib


def parse(L, line, cigar=True):
    buf = C.create_string_buffer(line if isinstance(line, bytes) else line.encode())
    return L.paf_parse(buf, cigar)


def text(L, p):
    s = L.paf_print(p)
    out = C.string_at(s)
    C.CDLL(None).free(C.c_void_p(s))
    return out


def golden_lines(name):
    return open(os.path.join(G, name), "rb").read().splitlines()


def test_library_exports_the_paf_h_surface():
    lib = paffy_b200.load()
    for name in PAF_H_SYMBOLS:
        assert hasattr(lib, name), name
    hdr = open(os.path.join(os.path.dirname(G), "..", "include", "paf_b200_compat.h")).read()
    for name in PAF_H_SYMBOLS:
        assert name + "(" in hdr, f"{name} is not declared in include/paf_b200_compat.h"
    assert C.sizeof(Paf) == 144 and C.sizeof(Cigar) == 32
    # the sonLib-container functions ship as a static archive for host programs that bring their own sonLib
    ar = os.path.join(os.path.dirname(paffy_b200.lib.LIB_PATH), "libpaffy_b200_stpaf.a")
    syms = subprocess.run(["nm", "--defined-only", ar], capture_output=True, text=True).stdout
    for name in ("read_pafs", "write_pafs", "paf_shatter", "get_alignment_count_array", "increase_alignment_level_counts",
                 "sequenceCountArray_destruct", "decode_fasta_header", "cmp_intervals", "interval_destruct"):
        assert f" T {name}\n" in syms, name


@pytest.mark.parametrize("stem", ["edge", "rand101", "rand102", "rand103"])
def test_parse_print_round_trip_matches_reference_normalisation(L, stem):
    """paf_parse + paf_print = the tag normalisation the reference applies (AS:i:0, dropped tags, tag order): the raw input
    re-printed equals `paffy add_mismatches -a` on an =/X-free file ... which for the raw inputs is what the oracle's
    parse + print gives; and the reference's own outputs are fixed points."""
    for line in golden_lines(stem + ".paf"):
        rc, op = O.parse(line)
        if rc != 0:
            continue
        p = parse(L, line)
        assert text(L, p) == O.to_text(op).rstrip(b"\n")
        L.paf_destruct(p)
    for line in golden_lines(stem + ".add.paf"):
        p = parse(L, line)
        assert text(L, p) == line
        L.paf_destruct(p)
        q = parse(L, line, cigar=False)  # the cigar kept as text
        assert bool(q.contents.cigar) is False and text(L, q) == line
        L.paf_destruct(q)


@pytest.mark.parametrize("stem", ["rand101", "rand102", "rand103"])
def test_trims_and_remove_match_reference_cli_outputs(L, stem):
    add = golden_lines(stem + ".add.paf")
    for fn, args, out_name in ((L.paf_trim_unreliable_tails, (0.05, 1.0), ".trim.paf"), (L.paf_trim_end_fraction, (0.3,), ".trimf.paf"),
                               (L.paf_remove_mismatches, (), ".remove.paf")):
        want = golden_lines(stem + out_name)
        assert len(want) == len(add)
        for line, w in zip(add, want):
            p = parse(L, line)
            fn(p, *args)
            L.paf_check(p)
            assert text(L, p) == w
            L.paf_destruct(p)
    raw = [ln for ln in golden_lines(stem + ".paf") if O.parse(ln)[0] == 0]
    for fn, args, out_name in ((L.paf_trim_unreliable_tails, (0.2, 0.5), ".trimraw.paf"), (L.paf_trim_end_fraction, (0.5,), ".trimfraw.paf")):
        want = golden_lines(stem + out_name)
        for line, w in zip(raw, want):
            p = parse(L, line)
            fn(p, *args)
            assert text(L, p) == w
            L.paf_destruct(p)


def test_stats_invert_and_aligned_bases_match_the_oracle(L):
    for line in golden_lines("rand102.add.paf") + golden_lines("edge.add.paf"):
        p = parse(L, line)
        rc, op = O.parse(line)
        s = [C.c_int64(7) for _ in range(6)]
        L.paf_stats_calc(p, *[C.byref(x) for x in s], True)
        want = (C.c_int64 * 6)()
        O.lib().oracle_stats(op.ops, op.n_ops, want)
        assert [x.value for x in s] == list(want)
        L.paf_stats_calc(p, *[C.byref(x) for x in s], False)  # accumulates
        assert [x.value for x in s] == [2 * v for v in want]
        assert L.paf_get_number_of_aligned_bases(p) == O.lib().oracle_aligned_bases(op.ops, op.n_ops)
        L.paf_invert(p)
        O.lib().oracle_paf_invert(C.byref(op))
        assert text(L, p) == O.to_text(op).rstrip(b"\n")
        L.paf_invert(p)
        assert text(L, p) == line
        L.paf_destruct(p)


def test_fatal_paths_follow_st_errAbort():
    """paf_check failures and parse errors are fatal the reference's way: ERROR: <message> on stderr, exit code 1."""
    code = ("import ctypes as C, sys, paffy_b200\n"
            "L = paffy_b200.load(); L.paf_parse.restype = C.c_void_p; L.paf_check.argtypes = [C.c_void_p]\n"
            "p = L.paf_parse(C.create_string_buffer(sys.argv[1].encode()), True)\n"
            "L.paf_check(p); print('ok')\n")
    env = dict(os.environ, PYTHONPATH=os.path.dirname(os.path.dirname(G)))
    cases = [("q\t25\t0\t10\t+\tt\t25\t0\t9\t0\t0\t60\tcg:Z:9=", 1, b"ERROR: Paf cigar alignment does not match query length: 9 vs. 10 q\t25\t0\t10"),
             ("q\t25\t0\t10\t*\tt\t25\t0\t9\t0\t0\t60\tcg:Z:9=", 1, b"ERROR: Got an unexpected strand character (*) in a paf string"),
             ("q\t25\t0\t10\t+\tt\t25\t0\t10\t0\t0\t60\tcg:Z:9=1Q", 1, b"ERROR: Got an unexpected character paf cigar string: Q"),
             ("q\t25\t30\t10\t+\tt\t25\t0\t10\t0\t0\t60", 1, b"ERROR: Paf query start coordinates are invalid, q\t25\t30"),
             ("q\t25\t0\t10\t+\tt\t25\t0\t10\t0\t0\t60\tcg:Z:10M", 0, b"")]
    for line, rc, head in cases:
        r = subprocess.run(["python", "-c", code, line], capture_output=True, env=env)
        assert r.returncode == rc, (line, r.stderr)
        assert r.stderr.startswith(head), r.stderr
        assert (r.stdout == b"ok\n") == (rc == 0)


def _unit_dir(tmp_path):
    d = tmp_path / "tests"
    d.mkdir()
    with gzip.open(os.path.join(G, "c1_human_chimp.paf.gz"), "rb") as src, open(d / "human_chimp.paf", "wb") as dst:
        shutil.copyfileobj(src, dst)
    return tmp_path


def _run_unit_tests(tmp_path, skip):
    if not os.access(O.UNIT_TESTS, os.X_OK):
        pytest.skip("oracle/_ref/b200_unit_tests is not built (make -C oracle linkproof needs /root/reference)")
    r = subprocess.run([O.UNIT_TESTS] + skip, cwd=_unit_dir(tmp_path), capture_output=True, text=True)
    lines = r.stdout.splitlines()
    return r, [ln for ln in lines if ln.startswith("PASS ")], [ln for ln in lines if ln.startswith("FAIL ")]


def test_reference_unit_tests_pass_against_the_library(tmp_path):
    """SURVEY section 7 step 1 gate: tests/paf_unit_test.c and test_paf compiled unchanged, linked against the product."""
    r, passed, failed = _run_unit_tests(tmp_path, ["encode_mismatches"])
    assert r.returncode == 0 and not failed, r.stdout[-2000:] + r.stderr[-2000:]
    assert len(passed) == 38 and "PASS test_paf" in passed  # 40 unit cases - 3 that compare bases + test_paf


@pytest.mark.gpu
def test_reference_unit_tests_all_pass_on_the_gpu(tmp_path):
    r, passed, failed = _run_unit_tests(tmp_path, [])
    assert r.returncode == 0 and not failed, r.stdout[-2000:] + r.stderr[-2000:]
    assert len(passed) == 41
    assert {"PASS test_paf_encode_mismatches_all_match", "PASS test_paf_encode_mismatches_all_mismatch",
            "PASS test_paf_encode_mismatches_mixed"} <= set(passed)


CPU_SUBCOMMANDS = [["invert"], ["shatter"], ["filter", "-u", "0.9"], ["filter", "-t", "1", "-x"], ["chain"], ["tile"], ["dedupe"],
                   ["add_mismatches", "-a"], ["to_bed"], ["upconvert"]]


@pytest.mark.parametrize("cmd", CPU_SUBCOMMANDS, ids=lambda c: "_".join(c))
def test_linked_cli_prints_what_the_reference_cli_prints(cmd, tmp_path):
    """INTEGRATION.md section 1 executed: the reference's paffy_main.c + its other mains, with impl/paf.c replaced by the
    library. Same stdout and exit code as the reference binary (including the abort of `shatter` on =/X cigars)."""
    if not (os.access(O.LINKED_PAFFY, os.X_OK) and os.access(O.REF_PAFFY, os.X_OK)):
        pytest.skip("oracle/_ref/paffy_linked is not built")
    c1 = os.path.join(_unit_dir(tmp_path), "tests", "human_chimp.paf")
    for inp in (os.path.join(G, "rand101.add.paf"), os.path.join(G, "rand103.paf"), c1):
        a = subprocess.run([O.REF_PAFFY] + cmd + ["-i", inp], capture_output=True)
        b = subprocess.run([O.LINKED_PAFFY] + cmd + ["-i", inp], capture_output=True)
        assert a.returncode == b.returncode, (cmd, inp, b.stderr[-500:])
        assert a.stdout == b.stdout, (cmd, inp)


@pytest.mark.gpu
@pytest.mark.parametrize("stem", ["edge", "rand101", "rand102"])
def test_linked_cli_gpu_subcommands_match_goldens(stem, tmp_path):
    if not os.access(O.LINKED_PAFFY, os.X_OK):
        pytest.skip("oracle/_ref/paffy_linked is not built")
    fa, paf = os.path.join(G, stem + ".fa"), os.path.join(G, stem + ".paf")
    r = subprocess.run([O.LINKED_PAFFY, "add_mismatches", fa, "-i", paf], capture_output=True)
    assert r.returncode == 0 and r.stdout == open(os.path.join(G, stem + ".add.paf"), "rb").read(), r.stderr[-500:]
    r = subprocess.run([O.LINKED_PAFFY, "view", "-s", "-t", fa, "-i", os.path.join(G, stem + ".add.paf")], capture_output=True)
    assert r.returncode == 0 and r.stdout == open(os.path.join(G, stem + ".view.txt"), "rb").read(), r.stderr[-500:]
    if stem != "edge":
        r = subprocess.run([O.LINKED_PAFFY, "trim", "-i", os.path.join(G, stem + ".add.paf")], capture_output=True)
        assert r.returncode == 0 and r.stdout == open(os.path.join(G, stem + ".trim.paf"), "rb").read(), r.stderr[-500:]
