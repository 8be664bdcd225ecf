#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference.

Run in the authoring container only (needs oracle/_ref/paffy, i.e. `make -C oracle ref`, which needs
/root/reference). Everything written here is an OUTPUT of the compiled reference CLI
(`paffy add_mismatches`, `paffy trim`, `paffy view -s -t`) on inputs this script also writes, so the
tests can run where /root/reference does not exist (the GPU box).

    python tests/golden/make_golden.py

Fixtures
  edge.{fa,paf}            SURVEY.md a1-edge cases (per-op runs, 0M, case folding, IUPAC, both strands,
                           pass-through =/X, record without cg, tag normalisation)
  rand<seed>.{fa,paf}      seeded random records over low-entropy contigs with soft-masked / N / IUPAC bytes
  *.add.paf                `paffy add_mismatches -i X.paf X.fa`
  *.remove.paf             `paffy add_mismatches -a` applied to *.add.paf
  *.trim.paf               `paffy add_mismatches ... | paffy trim`           (defaults -r 0.05 -t 1.0)
  *.trimf.paf              `paffy add_mismatches ... | paffy trim -f -t 0.3` (fixed end fraction)
  *.trimraw.paf            `paffy trim -r 0.2 -t 0.5` on the raw input (no add_mismatches before it)
  *.trimfraw.paf           `paffy trim -f -t 0.5` on the raw input
  *.filter_{u,vx,t}.paf    `paffy add_mismatches ... | paffy filter -u 0.5`, `... -v 0.4 -x`, `... -t 1 -s -1 -w 3`
  *.view.txt               `paffy add_mismatches ... | paffy view -s -t X.fa` summary line
  *.viewlines.txt          `paffy view -s X.fa -i X.paf` on the raw input: per-alignment stats lines + summary line
  *.viewa.txt              `paffy view -a -s X.fa -i X.paf`: the base-level alignment blocks (original-case bytes) as well
  errors.json              exit code / stdout / stderr head of the failing-record cases
  c1_human_chimp.paf.gz    the reference's own fixture tests/human_chimp.paf (data, gzip'd; config C1)
  c1.json                  size + md5 of the reference outputs for C1 with the seeded synthetic chr10
                           (tests/util.py c1_genome), and the view summary line
  reference_unit_vectors.json is hand-restated from tests/paf_unit_test.c and is NOT written here.
"""
import gzip
import hashlib
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import pyoracle as O  # noqa: E402
from tests import util  # noqa: E402

REF_FIXTURE = "/root/reference/tests/human_chimp.paf"


def run(args, stdin=None):
    r = subprocess.run([O.REF_PAFFY] + args, input=stdin, capture_output=True)
    return r.returncode, r.stdout, r.stderr


def must(args, stdin=None):
    rc, out, err = run(args, stdin)
    if rc != 0:
        raise SystemExit(f"reference failed: paffy {' '.join(args)} -> {rc}\n{err.decode()[-2000:]}")
    return out


def write(name, data):
    with open(os.path.join(HERE, name), "wb") as fh:
        fh.write(data)


def pipeline(stem, fa_path, paf_bytes):
    """Run the reference on one input set and store every output."""
    add = must(["add_mismatches", fa_path], paf_bytes)
    write(stem + ".add.paf", add)
    write(stem + ".remove.paf", must(["add_mismatches", "-a"], add))
    # the reference's own assert (impl/paf.c:952) can abort trim on adversarial records: no fixture then
    for suffix, args in ((".trim.paf", ["trim"]), (".trimf.paf", ["trim", "-f", "-t", "0.3"])):
        rc, out, err = run(args, add)
        path = os.path.join(HERE, stem + suffix)
        if rc == 0:
            write(stem + suffix, out)
        else:
            why = (err.decode(errors="replace").strip().splitlines() or ["no message"])[-1][:100]
            print(f"  {stem}: reference `paffy {' '.join(args)}` exits {rc} ({why}); no fixture")
            if os.path.exists(path):
                os.unlink(path)
    # trim straight on the raw (M) input, where M counts as a match (impl/paf.c:823): both trims
    for suffix, args in ((".trimraw.paf", ["trim", "-r", "0.2", "-t", "0.5"]), (".trimfraw.paf", ["trim", "-f", "-t", "0.5"])):
        rc, out, err = run(args, paf_bytes)
        if rc == 0:
            write(stem + suffix, out)
        elif os.path.exists(os.path.join(HERE, stem + suffix)):
            os.unlink(os.path.join(HERE, stem + suffix))
    write(stem + ".view.txt", must(["view", "-s", "-t", fa_path], add))
    # add_mismatches | filter (impl/paf_filter.c): identity, identity with gaps inverted, alignment score
    write(stem + ".filter_u.paf", must(["filter", "-u", "0.5"], add))
    write(stem + ".filter_vx.paf", must(["filter", "-v", "0.4", "-x"], add))
    write(stem + ".filter_t.paf", must(["filter", "-t", "1", "-s", "-1", "-w", "3"], add))
    # view on the raw (M) input: per-alignment lines + the summary line, i.e. paffy_view_main's own encode
    write(stem + ".viewlines.txt", must(["view", "-s", fa_path], paf_bytes))
    write(stem + ".viewa.txt", must(["view", "-a", "-s", fa_path], paf_bytes))


def fasta_bytes(contigs, width=60):
    out = []
    for name, seq in contigs.items():
        out.append(b">" + name.encode() + b"\n")
        for i in range(0, len(seq), width):
            out.append(seq[i:i + width] + b"\n")
    return b"".join(out)


def edge_set():
    q = b"ACGTNnacgtMRWSYKVHDBXACGT"
    t = b"ACGTNNACGTKYWSRMBDHVXACGA"
    contigs = {"q": q, "t": t, "q with a description": q[::-1], "t2": t + t}
    lines = [
        "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\ttp:A:P\tcg:Z:25M",
        "q\t25\t0\t25\t-\tt\t25\t0\t25\t0\t25\t60\ttp:A:P\tcg:Z:25M",
        # runs never join across ops; a zero-length M vanishes
        "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:3M0M2M5M4I4D11M",
        "q\t25\t0\t25\t-\tt\t25\t0\t25\t0\t25\t60\tcg:Z:3M0M2M5M4I4D11M",
        # existing =/X pass through
        "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tAS:i:77\tcg:Z:2=2X6M15M",
        # no cg tag: record passes through (tags normalised)
        "q\t25\t3\t9\t+\tt\t25\t5\t11\t6\t6\t60\ttp:A:S\txx:i:5\ts1:i:9\tcn:i:2\ttl:i:3",
        # dropped tags, tile level without tp, whole-line contig name
        "q with a description\t25\t1\t20\t-\tt2\t50\t20\t41\t0\t0\t0\ttl:i:2\tNM:i:4\tcg:Z:10M2D9M",
        "q\t25\t5\t6\t+\tt2\t50\t49\t50\t1\t1\t1\tcg:Z:1M",
        "q\t25\t0\t1\t-\tt2\t50\t0\t1\t1\t1\t1\tcg:Z:1M",
    ]
    return contigs, ("\n".join(lines) + "\n").encode()


def rand_set(seed):
    case = util.random_case(seed, n_records=70, contig_len=20000, n_ops=14, max_len=900, p_weird=0.003)
    recs = []
    for r in case["records"]:
        if r["qe"] == r["qs"] or sum(n for n, c in r["ops"] if c != "I") == 0:
            continue  # paf_check rejects empty spans at the contig end; keep the fixture valid
        recs.append(r)
    case["records"] = recs
    return case["contigs"], ("\n".join(util.case_lines(case)) + "\n").encode()


def error_cases(fa_path):
    good = "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:25M"
    out = {}
    cases = {
        "cigar_query_span": [good, "q\t25\t0\t10\t+\tt\t25\t0\t9\t0\t0\t60\tcg:Z:9M", good],
        "missing_query": [good, "nope\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:25M", good],
        "missing_target": [good, "q\t25\t0\t25\t+\tnope\t25\t0\t25\t0\t25\t60\tcg:Z:25M"],
    }
    for name, lines in cases.items():
        rc, so, se = run(["add_mismatches", fa_path], ("\n".join(lines) + "\n").encode())
        se = se.decode(errors="replace")
        # keep the message line only (backtraces are build specific)
        msg = [ln for ln in se.splitlines() if ln.startswith("ERROR") or ln.startswith("No ")]
        out[name] = {"input": lines, "exit": rc, "stdout": so.decode(), "stderr_message": msg[0] if msg else ""}
    return out


def c1_set(tmp):
    raw = open(REF_FIXTURE, "rb").read()
    with open(os.path.join(HERE, "c1_human_chimp.paf.gz"), "wb") as fh:
        with gzip.GzipFile(fileobj=fh, mode="wb", mtime=0, compresslevel=9) as gz:
            gz.write(raw)
    fa = os.path.join(tmp, "c1.fa")
    util.write_c1_fasta(fa)
    add = must(["add_mismatches", fa, "-i", REF_FIXTURE])
    trim = must(["trim"], add)
    view = must(["view", "-s", "-t", fa], add)
    remove = must(["add_mismatches", "-a"], add)
    viewlines = must(["view", "-s", fa, "-i", REF_FIXTURE])
    trimf = must(["trim", "-f", "-t", "0.3"], add)
    info = {
        "input_md5": hashlib.md5(raw).hexdigest(), "input_bytes": len(raw), "records": raw.count(b"\n"),
        "genome": "tests/util.py c1_genome(): numpy default_rng(20261104).integers(0,4,133797422) -> ACGT, one contig chr10",
        "add_bytes": len(add), "add_md5": hashlib.md5(add).hexdigest(),
        "trim_bytes": len(trim), "trim_md5": hashlib.md5(trim).hexdigest(),
        "remove_bytes": len(remove), "remove_md5": hashlib.md5(remove).hexdigest(),
        "view_summary": view.decode(),
        "trimf_bytes": len(trimf), "trimf_md5": hashlib.md5(trimf).hexdigest(),
        "viewlines_bytes": len(viewlines), "viewlines_md5": hashlib.md5(viewlines).hexdigest(),
        "add_line_md5": [hashlib.md5(ln).hexdigest()[:12] for ln in add.split(b"\n")[:-1]],
    }
    with open(os.path.join(HERE, "c1.json"), "w") as fh:
        json.dump(info, fh, indent=1)
    return info


def main():
    if not O.ref_available():
        raise SystemExit("oracle/_ref/paffy is missing: run `make -C oracle ref` where /root/reference exists")
    with tempfile.TemporaryDirectory() as tmp:
        for stem, (contigs, paf) in [("edge", edge_set())] + [(f"rand{s}", rand_set(s)) for s in (101, 102, 103)]:
            fa = fasta_bytes(contigs)
            write(stem + ".fa", fa)
            write(stem + ".paf", paf)
            pipeline(stem, os.path.join(HERE, stem + ".fa"), paf)
            print(stem, "ok")
        with open(os.path.join(HERE, "errors.json"), "w") as fh:
            json.dump(error_cases(os.path.join(HERE, "edge.fa")), fh, indent=1)
        info = c1_set(tmp)
        print("c1", info["add_bytes"], info["add_md5"], info["view_summary"].strip())


if __name__ == "__main__":
    main()
