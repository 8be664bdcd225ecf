#!/usr/bin/env python
"""Per-record `paf_encode_mismatches(Paf *, query_seq, target_seq)` on config C1 (tests/human_chimp.paf, 207 records, one
synthetic chr10 for both names): the library symbol on the reference's struct layout, record after record with the same
whole-contig strings, as `paffy view` / Cactus call it - beside the reference's own function (oracle/_ref/libpaf_ref.so)
in the same loop. Prints one JSON object. usage: python tools/encode_one_bench.py [out.json]"""
import ctypes as C
import gzip
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import paffy_b200  # noqa: E402
from tests import util  # noqa: E402
from tests.test_paf_api import Paf  # noqa: E402


def run(lib, lines, genome, reps=1):
    lib.paf_parse.restype = C.POINTER(Paf)
    lib.paf_parse.argtypes = [C.c_char_p, C.c_bool]
    lib.paf_encode_mismatches.argtypes = [C.POINTER(Paf), C.c_char_p, C.c_char_p]
    lib.paf_encode_mismatches.restype = None
    lib.paf_print.restype = C.c_void_p
    lib.paf_print.argtypes = [C.POINTER(Paf)]
    lib.paf_destruct.argtypes = [C.POINTER(Paf)]
    times, outs = [], None
    for _ in range(reps):
        pafs = [lib.paf_parse(C.create_string_buffer(ln), True) for ln in lines]
        t0 = time.perf_counter()
        for p in pafs:
            lib.paf_encode_mismatches(p, genome, genome)
        times.append(time.perf_counter() - t0)
        if outs is None:
            import hashlib
            h = hashlib.md5()
            for p in pafs:
                s = lib.paf_print(p)
                h.update(C.string_at(s))
                C.CDLL(None).free(C.c_void_p(s))
            outs = h.hexdigest()
        for p in pafs:
            lib.paf_destruct(p)
    return times, outs


def main():
    lines = gzip.open(os.path.join(ROOT, "tests", "golden", "c1_human_chimp.paf.gz"), "rb").read().splitlines()
    aligned = 124949228
    g = util.c1_genome().tobytes()
    genome = C.create_string_buffer(g, len(g) + 1)
    out = {"config": "C1: tests/human_chimp.paf, 207 records, 124 949 228 aligned bp, one synthetic chr10 (133 797 422 bp) for query and target"}
    ours, md5_ours = run(paffy_b200.load(), lines, genome, reps=3)
    out["paffy_b200"] = {"first_pass_s": ours[0], "later_pass_s": min(ours[1:]), "records_per_s": len(lines) / min(ours[1:]),
                         "aligned_bp_per_s": aligned / min(ours[1:]),
                         "what": "first pass packs the contig into HBM (strlen + upload + pack, once per sequence pointer); later passes upload the record only"}
    ref_so = os.path.join(ROOT, "oracle", "_ref", "libpaf_ref.so")
    if os.path.exists(ref_so):
        ref, md5_ref = run(C.CDLL(ref_so), lines, genome, reps=1)
        out["reference"] = {"pass_s": ref[0], "records_per_s": len(lines) / ref[0], "aligned_bp_per_s": aligned / ref[0], "threads": 1}
        out["outputs_equal"] = md5_ours == md5_ref
    print(json.dumps(out))
    if len(sys.argv) > 1:
        json.dump(out, open(sys.argv[1], "w"), indent=1)


if __name__ == "__main__":
    main()
