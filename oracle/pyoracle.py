"""ctypes bindings for the parity checkers in oracle/ (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module. Nothing under paffy_b200/ does.
"""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libpaf_oracle.so")
REF_DIR = os.path.join(HERE, "_ref")
REF_PAFFY = os.path.join(REF_DIR, "paffy")
REF_BENCH = os.path.join(REF_DIR, "ref_bench")
UNIT_TESTS = os.path.join(REF_DIR, "b200_unit_tests")   # the reference's unit tests linked against the product
LINKED_PAFFY = os.path.join(REF_DIR, "paffy_linked")   # the reference's paffy_main.c linked against the product


def build(ref=True):
    """Compile the C restatement (always) and, where /root/reference exists, oracle/_ref."""
    subprocess.run(["make", "-s", "-C", HERE, "oracle"], check=True)
    if ref and os.path.isdir("/root/reference/impl"):
        subprocess.run(["make", "-s", "-j8", "-C", HERE, "ref"], check=True)
        # the link proof of the drop-in boundary: the reference's unchanged unit tests and paffy_main.c against the
        # product library (which must have been built first)
        if os.path.exists(os.path.join(HERE, "..", "paffy_b200", "libpaffy_b200_stpaf.a")):
            subprocess.run(["make", "-s", "-j8", "-C", HERE, "linkproof"], check=True)


class OPaf(C.Structure):
    _fields_ = [("query_name", C.c_char_p), ("target_name", C.c_char_p),
                ("query_length", C.c_int64), ("query_start", C.c_int64), ("query_end", C.c_int64),
                ("target_length", C.c_int64), ("target_start", C.c_int64), ("target_end", C.c_int64),
                ("num_matches", C.c_int64), ("num_bases", C.c_int64), ("mapping_quality", C.c_int64),
                ("score", C.c_int64), ("tile_level", C.c_int64), ("chain_id", C.c_int64),
                ("chain_score", C.c_int64), ("same_strand", C.c_int), ("type", C.c_char),
                ("has_cigar", C.c_int), ("ops", C.POINTER(C.c_uint64)), ("n_ops", C.c_int64)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        L = C.CDLL(ORACLE_SO)
        u64p, i64p = C.POINTER(C.c_uint64), C.POINTER(C.c_int64)
        L.oracle_revcomp_char.restype = C.c_char
        L.oracle_revcomp_char.argtypes = [C.c_char]
        L.oracle_encode_mismatches.restype = C.c_int64
        L.oracle_encode_mismatches.argtypes = [u64p, C.c_int64, C.c_int, C.c_int64, C.c_int64, C.c_int64,
                                               C.c_char_p, C.c_char_p, C.POINTER(u64p)]
        L.oracle_remove_mismatches.restype = C.c_int64
        L.oracle_remove_mismatches.argtypes = [u64p, C.c_int64]
        L.oracle_stats.argtypes = [u64p, C.c_int64, i64p]
        L.oracle_aligned_bases.restype = C.c_int64
        L.oracle_aligned_bases.argtypes = [u64p, C.c_int64]
        L.oracle_cigar_parse.restype = C.c_int64
        L.oracle_cigar_parse.argtypes = [C.c_char_p, C.POINTER(u64p)]
        L.oracle_paf_parse.argtypes = [C.c_char_p, C.POINTER(OPaf)]
        L.oracle_paf_free.argtypes = [C.POINTER(OPaf)]
        L.oracle_paf_print.restype = C.c_int64
        L.oracle_paf_print.argtypes = [C.POINTER(OPaf), C.POINTER(C.c_void_p)]
        L.oracle_paf_check.argtypes = [C.POINTER(OPaf)]
        L.oracle_paf_invert.argtypes = [C.POINTER(OPaf)]
        L.oracle_trim_unreliable_tails.argtypes = [C.POINTER(OPaf), C.c_float, C.c_float]
        L.oracle_trim_ends.argtypes = [C.POINTER(OPaf), C.c_int64]
        L.oracle_trim_end_fraction.argtypes = [C.POINTER(OPaf), C.c_float]
        L.oracle_fasta_new.restype = C.c_void_p
        L.oracle_fasta_load.argtypes = [C.c_void_p, C.c_char_p]
        L.oracle_fasta_add.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int64]
        L.oracle_fasta_get.restype = C.c_void_p
        L.oracle_fasta_get.argtypes = [C.c_void_p, C.c_char_p, i64p]
        L.oracle_fasta_count.restype = C.c_int64
        L.oracle_fasta_count.argtypes = [C.c_void_p]
        L.oracle_fasta_free.argtypes = [C.c_void_p]
        L.oracle_add_mismatches_file.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int, C.c_float,
                                                 C.c_float, i64p, i64p]
        L.free = C.CDLL(None).free
        L.free.argtypes = [C.c_void_p]
        _lib = L
    return _lib


OPS = "MID=X"


def mk_ops(pairs):
    """[(length, 'M'), ...] -> ctypes uint64 array of CigarRecord words."""
    arr = (C.c_uint64 * max(1, len(pairs)))()
    for i, (n, c) in enumerate(pairs):
        arr[i] = (n & ((1 << 56) - 1)) | (OPS.index(c) << 56)
    return arr


def ops_to_pairs(arr, n):
    out = []
    for i in range(n):
        w = int(arr[i])
        ln = w & ((1 << 56) - 1)
        if ln >> 55:
            ln -= 1 << 56
        out.append((ln, OPS[w >> 56]))
    return out


def cigar_str(pairs):
    return "".join(f"{n}{c}" for n, c in pairs)


def encode(pairs, same_strand, qs, qe, ts, qseq, tseq):
    L = lib()
    arr = mk_ops(pairs)
    out = C.POINTER(C.c_uint64)()
    n = L.oracle_encode_mismatches(arr, len(pairs), int(same_strand), qs, qe, ts, qseq, tseq, C.byref(out))
    res = ops_to_pairs(out, n)
    L.free(out)
    return res


def parse(line):
    p = OPaf()
    rc = lib().oracle_paf_parse(line.encode() if isinstance(line, str) else line, C.byref(p))
    return rc, p


def to_text(p):
    out = C.c_void_p()
    n = lib().oracle_paf_print(C.byref(p), C.byref(out))
    s = C.string_at(out, n)
    lib().free(out)
    return s


class Fasta:
    def __init__(self, paths=()):
        self.h = lib().oracle_fasta_new()
        for p in paths:
            rc = lib().oracle_fasta_load(self.h, p.encode())
            if rc:
                raise ValueError(f"oracle_fasta_load({p}) -> {rc}")

    def add(self, name, seq):
        lib().oracle_fasta_add(self.h, name.encode(), seq, len(seq))

    def get(self, name):
        n = C.c_int64()
        p = lib().oracle_fasta_get(self.h, name.encode(), C.byref(n))
        return None if not p else C.string_at(p, n.value)

    def __len__(self):
        return lib().oracle_fasta_count(self.h)

    def add_mismatches_file(self, in_path, out_path, mode=0, trim_r=0.05, trim_t=1.0):
        stats = (C.c_int64 * 7)()
        aligned = C.c_int64()
        rc = lib().oracle_add_mismatches_file(self.h, in_path.encode(), out_path.encode(), mode, trim_r, trim_t,
                                              stats, C.byref(aligned))
        return rc, list(stats), aligned.value

    def __del__(self):
        try:
            lib().oracle_fasta_free(self.h)
        except Exception:
            pass


def ref_available():
    return os.access(REF_PAFFY, os.X_OK)


def ref_paffy(args, stdin=None):
    """Run the compiled, unmodified reference CLI. Returns (returncode, stdout bytes, stderr bytes)."""
    r = subprocess.run([REF_PAFFY] + list(args), input=stdin, capture_output=True)
    return r.returncode, r.stdout, r.stderr
