#!/usr/bin/env python
"""bench.py - add_mismatches aligned bp/s on synthetic genomes + PAFs of BASELINE.json's shapes.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C2|C3|C4|small|tiny] [--scale f]
  python bench.py --impl reference ...      (the reference's own CPU code on the host cores)

One process per GPU (torchrun for N > 1; records are sharded, the packed genome is replicated, no
data-path collective). A step is one pass of the hot path (CIGAR ingest + base-comparison kernel +
stitch) over one batch of records. `value`: inputs resident in HBM. `e2e`: the C-ABI host call with
pinned host buffers - H2D of the batch, ingest + kernel + device serialiser, D2H of the PAF text.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "add_mismatches aligned bp/sec"
UNIT = "aligned bp/s"
# kernels of ours per batch: scalars_init, ingest, scan (one pass; also writes tile_first_op), words_out (scalars to the
# host), encode, stitch, words_out (number of output ops)
KERNELS_PER_STEP = 7
# per sub-batch of the e2e pipeline: the above + ser_block_records, ser_write, prefix_copy, words_out (text size)
KERNELS_PER_E2E_SUBBATCH = 11
# per chunk of the text path (counted in profiles/r02_launches_text_c2.csv): count_delims, 2 scans, words_out, scatter_delims,
# parse_records, ops (count), scan, words_out, ops (write), the 7 of a batch, op_totals, sum, words_out, empty_out_begin,
# prefix (measure), scan, ser_block_records, ser_write, prefix (render), words_out
KERNELS_PER_TEXT_CHUNK = 28


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class Clocks:
    """Sample nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        self.p.wait()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        sm = [float(r[0]) for r in rows if r[0].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in rows for i in range(4) if len(r) >= 6 and r[2 + i].strip() == "Active"})
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": float(rows[0][1]) if rows and rows[0][1].replace(".", "").isdigit() else None,
                "reasons": reasons, "samples": len(sm)}


def render_prefixes(rec):
    """Host-rendered text of every record up to and including '\\tcg:Z:' (what paf_write_to_buffer emits
    before the CIGAR, impl/paf.c:320-368) -> (uint8 array, int64 offsets)."""
    c = {k: rec[k].cpu().numpy() for k in ("contig", "ts", "te", "qs", "qe", "strand", "sum_m", "q_len")}
    ob = rec["op_begin"].cpu().numpy()
    ops = rec["ops"].cpu().numpy()
    tot = np.add.reduceat(ops & ((1 << 56) - 1), ob[:-1])
    clen = rec["clen"]
    parts = [f"q{ci + 1}\t{ql}\t{qs}\t{qe}\t{'+' if st else '-'}\tt{ci + 1}\t{clen}\t{ts}\t{te}\t{sm}\t{tt}\t60\ttp:A:P\tAS:i:{sm}\tcg:Z:"
             for ci, ql, qs, qe, st, ts, te, sm, tt in zip(c["contig"].tolist(), c["q_len"].tolist(), c["qs"].tolist(), c["qe"].tolist(),
                                                           c["strand"].tolist(), c["ts"].tolist(), c["te"].tolist(), c["sum_m"].tolist(),
                                                           tot.tolist())]
    lens = np.fromiter((len(p) for p in parts), dtype=np.int64, count=len(parts))
    begin = np.zeros(len(parts) + 1, dtype=np.int64)
    np.cumsum(lens, out=begin[1:])
    return np.frombuffer("".join(parts).encode(), dtype=np.uint8), begin


def cpu_sample(workload, scale, n_sample, device, n_pairs=4):
    """A bounded sample of the same workload (same generator, law and rates) on `n_pairs` contig pairs spread over the
    genome, both strands (even pairs are '+', odd ones '-'), written as FASTA + PAF for the compiled reference."""
    from paffy_b200 import synth
    n_contigs, clen, n_rec, law, s, g = synth.CONFIGS[workload]
    d = tempfile.mkdtemp(prefix="paffy_b200_cpu_")
    fa, paf = os.path.join(d, "sample.fa"), os.path.join(d, "sample.paf")
    total_m = 0
    with open(fa, "wb") as ffa, open(paf, "w") as fpaf:
        n_pairs = max(1, min(n_pairs, n_contigs))
        picks = sorted({(k * (n_contigs - 1)) // max(n_pairs - 1, 1) for k in range(n_pairs)}) if n_pairs > 1 else [0]
        if len(picks) > 1 and all(i % 2 == picks[0] % 2 for i in picks):
            picks[1] += 1  # both strands
        for i in picks:
            c = synth.make_contig(i, clen, s, g, synth.MASTER_SEED, device, keep_truth=False)
            p = synth.contig_records(c, max(1, n_sample // len(picks)), law, synth.MASTER_SEED + 4242, device)
            p["q_len"] = torch.full((p["n"],), c.Lq, dtype=torch.int64)
            p["clen"] = clen
            for name, asc in ((c.name_t, c.t_ascii), (c.name_q, c.q_ascii)):
                ffa.write(b">" + name.encode() + b"\n")
                ffa.write(asc.cpu().numpy().tobytes())
                ffa.write(b"\n")
            fpaf.write("\n".join(synth.paf_lines(p, range(p["n"]))) + "\n")
            total_m += int(p["sum_m"].sum().item())
            del c, p
    return d, fa, paf, total_m


def run_reference(fa, paf, threads, reps):
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_bench")
    if not os.access(exe, os.X_OK):
        return None
    env = dict(os.environ, REF_BENCH_REPS=str(reps))
    r = subprocess.run([exe, str(threads), paf, fa], capture_output=True, text=True, env=env)
    if r.returncode:
        sys.stderr.write(r.stderr)
        return None
    return json.loads(r.stdout)


def run_reference_cli(fa, paf, workdir):
    """The reference CLI end to end on the sample, as BASELINE.json asks: `paffy add_mismatches` single process, then
    one process per `paffy split_file` shard (one shard per target contig) run concurrently; wall = slowest."""
    exe = os.path.join(ROOT, "oracle", "_ref", "paffy")
    if not os.access(exe, os.X_OK):
        return None
    out = {}
    t0 = time.perf_counter()
    r = subprocess.run([exe, "add_mismatches", fa, "-i", paf, "-o", os.path.join(workdir, "single.out")], capture_output=True)
    if r.returncode:
        return None
    out["single_process_s"] = time.perf_counter() - t0
    prefix = os.path.join(workdir, "shard_")
    t0 = time.perf_counter()
    if subprocess.run([exe, "split_file", "-i", paf, "-p", prefix], capture_output=True).returncode:
        return out
    out["split_file_s"] = time.perf_counter() - t0
    shards = sorted(f for f in os.listdir(workdir) if f.startswith("shard_") and f.endswith(".paf"))
    t0 = time.perf_counter()
    procs = [subprocess.Popen([exe, "add_mismatches", fa, "-i", os.path.join(workdir, f), "-o", os.path.join(workdir, f + ".out")],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL) for f in shards]
    ok = all(p.wait() == 0 for p in procs)
    out["parallel_processes"] = len(shards)
    out["parallel_wall_s"] = time.perf_counter() - t0 if ok else None
    out["nproc"] = os.cpu_count()
    out["concurrency"] = len(shards)  # one process per split_file shard, all at once (each loads the whole FASTA of the sample)
    try:
        out["ram_gb"] = round(int(open("/proc/meminfo").readline().split()[1]) / 2 ** 20, 1)
    except Exception:
        out["ram_gb"] = None
    out["fasta_bytes"] = os.path.getsize(fa)
    return out


def file_parity(fa, paf, workdir, c):
    """The same sample file to file through the CUDA path (FASTA load, parse, kernels, device serialiser, write) and
    byte-compared with what the compiled reference CLI printed: `add_mismatches` (single.out of run_reference_cli),
    `add_mismatches | trim` and the `add_mismatches | view -s -t` summary line (SURVEY 8d, config C4's pipeline). The
    reference is the checker here, never the thing measured."""
    import hashlib
    import paffy_b200
    exe = os.path.join(ROOT, "oracle", "_ref", "paffy")
    ref_add = os.path.join(workdir, "single.out")
    if not (os.access(exe, os.X_OK) and os.path.exists(ref_add)):
        return None

    def md5(path):
        h = hashlib.md5()
        with open(path, "rb") as fh:
            for blk in iter(lambda: fh.read(1 << 24), b""):
                h.update(blk)
        return h.hexdigest()

    out = {}
    try:
        c2 = paffy_b200.Context(0)
        t0 = time.perf_counter()
        c2.load_fasta(fa)  # timed only: the comparison below runs on the caller's context (the whole genome)
        out["fasta_load_s"] = time.perf_counter() - t0
        c2.close()
        ours = os.path.join(workdir, "ours.add.paf")
        t0 = time.perf_counter()
        rc, rep = c.add_mismatches_file(paf, ours)
        out["add_file_to_file_s"] = time.perf_counter() - t0
        out["records"] = int(rep.records)
        out["add_bytes"] = os.path.getsize(ours)
        out["add_bytes_equal"] = rc == 0 and md5(ours) == md5(ref_add)
        ours_trim, ref_trim = os.path.join(workdir, "ours.trim.paf"), os.path.join(workdir, "ref.trim.paf")
        t0 = time.perf_counter()
        rc, rep = c.add_mismatches_file(paf, ours_trim, mode=2)
        out["add_trim_file_to_file_s"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        r = subprocess.run([exe, "trim", "-i", ref_add, "-o", ref_trim], capture_output=True)
        out["reference_trim_s"] = time.perf_counter() - t0
        out["trim_bytes_equal"] = (rc == 0 and md5(ours_trim) == md5(ref_trim)) if r.returncode == 0 else None
        ours_view = os.path.join(workdir, "ours.view.txt")
        t0 = time.perf_counter()
        rc, rep = c.add_mismatches_file(paf, ours_view, mode=4 | 2)
        out["view_file_to_file_s"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        r = subprocess.run([exe, "view", "-s", "-t", fa, "-i", ref_add], capture_output=True)
        out["reference_view_s"] = time.perf_counter() - t0
        out["view_summary_equal"] = (rc == 0 and open(ours_view, "rb").read() == r.stdout) if r.returncode == 0 else None
        out["view_summary"] = open(ours_view).read().strip()
    finally:
        pass
    out["what"] = ("CUDA path file to file on the cpu_baseline sample vs the compiled reference CLI: add_mismatches, "
                   "add_mismatches | trim (defaults), add_mismatches | view -s -t; md5 / bytes compared")
    return out


def reference_arm(args, rank):
    """The reference's own CPU implementation (oracle/_ref/ref_bench: unmodified paf_encode_mismatches on all
    host threads) on a bounded sample of the same workload. Rank 0 only."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    device = "cuda" if torch.cuda.is_available() else "cpu"
    n_sample = args.cpu_sample
    d, fa, paf, total_m = cpu_sample(args.workload, args.scale, n_sample, device)
    t0 = time.time()
    out = run_reference(fa, paf, cores, args.warmup + args.steps)
    if out is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_bench is not built (needs /root/reference at build time)"}))
        return
    reps = out["reps"][args.warmup:]
    secs = sum(r["encode_s"] for r in reps)
    bp = sum(r["aligned_bp"] for r in reps)
    value = bp / secs
    sample = (f"{reps[0]['records']} records of workload {args.workload} on 4 contig pairs (both strands), {reps[0]['aligned_bp']} aligned bp per step; "
              f"paf_encode_mismatches only (FASTA load {out['load_s']:.1f}s and PAF parse {reps[0]['parse_s']:.2f}s excluded)")
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * secs / len(reps), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic", "impl": "reference",
            "config": {"workload": f"{args.workload} (sampled)", "records_per_step": reps[0]["records"]},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.time() - t0}
    print(json.dumps(line))


def bind_near_gpu(index):
    """Pin this process (its host threads and the pinned buffers they first touch) to the CPUs of the GPU's NUMA node,
    so that staging copies do not cross sockets. No-op when the topology cannot be read."""
    try:
        bus = torch.cuda.get_device_properties(index).pci_bus_id
        dom = getattr(torch.cuda.get_device_properties(index), "pci_domain_id", 0)
        dev_id = torch.cuda.get_device_properties(index).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev_id:02x}.0"
        node = int(open(path + "/numa_node").read())
        if node < 0:
            return {"numa_node": node}
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return {"numa_node": node, "cpus": len(cpus)}
    except Exception as e:  # noqa: BLE001 - topology files differ between boxes
        return {"error": str(e)[:80]}


class DevArray:
    """A device buffer of the library as something torch.as_tensor can wrap without a copy."""
    def __init__(self, ptr, n, typestr="<i8"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


def make_input_text(ctx, rec, pre_b, pre_o):
    """The workload as PAF text (what `paffy add_mismatches` reads): the INPUT ops are staged on the device as they are and
    printed behind the host-rendered fields by the device serialiser. Not timed."""
    host = {k: rec[k].cpu().numpy() for k in ("op_begin", "qs", "qe", "ts", "strand")}
    ctx.stage_ops(host["op_begin"], rec["ops"].cpu().numpy().view(np.uint64), host["qs"], host["qe"], host["ts"], host["strand"])
    n, _ = ctx.serialise(pre_b, pre_o)
    text, _ = ctx.fetch_text(want_lines=False)
    return text[:n]


def md5_file(path):
    import hashlib
    h = hashlib.md5()
    with open(path, "rb") as fh:
        for blk in iter(lambda: fh.read(1 << 24), b""):
            h.update(blk)
    return h.hexdigest()


CPU_SAMPLE = {"C2": 100000, "C3": 500, "C4": 100000, "C5": 100000, "small": 20000, "tiny": 2000}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="paffy_b200")
    ap.add_argument("--workload", default=None, help="C2 (default at 1 GPU), C5 (default at N > 1: one job, strong scaling), C3, C4, small, tiny")
    ap.add_argument("--scale", type=float, default=1.0, help="fraction of the workload's record count")
    ap.add_argument("--scaling", default=None, choices=["strong", "weak"], help="N > 1: shard ONE record list (default) or one list per rank")
    ap.add_argument("--cpu-sample", type=int, default=0, help="records in the CPU-baseline sample (0 = per workload)")
    ap.add_argument("--cpu-contigs", type=int, default=4, help="contig pairs the CPU sample is drawn from (24 = all: 6 GB of FASTA per reference process)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-e2e-file", action="store_true")
    ap.add_argument("--e2e-subbatches", type=int, default=8)
    ap.add_argument("--e2e-contexts", type=int, default=8, help="sibling contexts (streams) the e2e chunks are dealt to; 8 hides the host round trips of a chunk behind the other chunks")
    ap.add_argument("--file-streams", type=int, default=3)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.workload is None:
        args.workload = "C2" if world == 1 else "C5"
    if not args.cpu_sample:
        args.cpu_sample = CPU_SAMPLE.get(args.workload, 100000)
    if args.impl == "reference":
        reference_arm(args, rank)
        return

    import paffy_b200
    from paffy_b200 import synth
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: paffy_b200 has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    affinity = bind_near_gpu(local) if os.environ.get("PAFGPU_BENCH_AFFINITY", "1") != "0" else None
    dist = None
    if world > 1:
        import torch.distributed as dist
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "NONE"  # keep stdout to the one JSON line (NCCL prints its version there)
        dist.init_process_group("nccl", device_id=dev)
    strong = world > 1 and args.scaling != "weak"

    ctx = paffy_b200.Context(local)
    t_gen = time.time()
    # strong scaling: every rank generates the SAME input-ordered record list (the job), pafgpu_partition cuts it into
    # `world` contiguous ranges of near-equal aligned bases, a rank keeps its own range; the packed genome is replicated
    rec = synth.build_workload(ctx, args.workload, dev, scale=args.scale, rank=0 if strong else rank)
    job_records = rec["n"]
    my_range = (0, rec["n"])
    if strong:
        cuts = paffy_b200.partition(rec["sum_m"].cpu().numpy(), world)
        my_range = (int(cuts[rank]), int(cuts[rank + 1]))
        rec = synth.shard(rec, *my_range)
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    t_gen = time.time() - t_gen
    R, N = rec["n"], rec["ops"].numel()
    dev_arrays = [rec["op_begin"], rec["ops"], rec["q_contig"], rec["t_contig"], rec["qs"], rec["qe"], rec["ts"], rec["strand"]]

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    # ---- device-resident steps ------------------------------------------------------------------
    res = None
    for _ in range(args.warmup):
        res = ctx.encode_batch(*dev_arrays)
    # size-independent parity properties at full size (generator truth)
    _, begin, mism = ctx.fetch_ops(want_ops=False)
    true_mism = rec["true_mism"].cpu().numpy()
    true_out = rec["true_out_ops"].cpu().numpy()
    parity = {"records": int(R), "mismatch_counts_equal": bool((mism == true_mism).all()),
              "output_op_counts_equal": bool((np.diff(begin) == true_out).all()), "n_out_ops": int(res.n_out_ops)}
    if not (parity["mismatch_counts_equal"] and parity["output_op_counts_equal"]) and not os.environ.get("PAFGPU_DEBUG"):
        raise SystemExit(f"parity check against generator truth failed: {parity}")

    clocks = Clocks(local) if rank == 0 else None
    barrier()
    t0 = time.perf_counter()
    ms_kernel = ms_stage = ms_ingest = ms_emit = 0.0
    for _ in range(args.steps):
        res = ctx.encode_batch(*dev_arrays)  # synchronous: returns when the step's kernels have finished
        ms_kernel += res.ms_kernel
        ms_stage += res.ms_encode
        ms_ingest += res.ms_ingest
        ms_emit += res.ms_emit
    torch.cuda.synchronize()
    own = time.perf_counter() - t0   # this rank's K steps; the closing barrier is not part of anybody's time
    barrier()
    elapsed = max_over_ranks(own)
    clk = clocks.stop() if clocks else None
    A, n_out = int(res.aligned_bases), int(res.n_out_ops)
    # outputs "gathered in input order": per rank (records, output ops, a checksum of the output words), rank = range order
    ops_ptr, _, _ = ctx.result_device_ptrs()
    out_sum = int(torch.as_tensor(DevArray(ops_ptr, n_out), device=dev).sum().item()) if n_out else 0
    mine = torch.tensor([float(A), float(R), float(n_out), float(out_sum % (1 << 52)), float(my_range[0]), float(my_range[1])], dtype=torch.float64, device=dev)
    if dist is not None:
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
    else:
        allr = [mine]
    shards = [{"rank": i, "records": [int(t[4].item()), int(t[5].item())], "aligned_bp": int(t[0].item()), "output_ops": int(t[2].item()),
               "output_checksum": int(t[3].item())} for i, t in enumerate(allr)]
    A_total = float(sum(sh["aligned_bp"] for sh in shards))
    value = A_total * args.steps / elapsed

    # ---- end to end through the host C ABI ----------------------------------------------------------
    # The batch lives in pinned host memory; every step copies it to the device, runs ingest + encode + device
    # serialiser and copies the finished PAF text back to pinned host memory. The records are cut into sub-batches
    # that are dealt to a few SIBLING contexts (own stream and buffers, the packed genome shared) so that the H2D copy
    # of one sub-batch, the kernels of another and the D2H copy of a third overlap.
    e2e = None
    pre_b = pre_o = None
    soa_md5 = None
    if not args.no_e2e:
        from concurrent.futures import ThreadPoolExecutor
        ctxs = [ctx] + [ctx.sibling() for _ in range(max(1, args.e2e_contexts) - 1)]
        host = [a.cpu() for a in dev_arrays]
        pre_b, pre_o = render_prefixes(rec)
        ob = host[0].numpy()
        n_sub = max(2, min(args.e2e_subbatches, R))
        cuts = [int(x) for x in paffy_b200.partition(rec["sum_m"].cpu().numpy(), n_sub)]
        subs = []
        h2d = 0
        for k in range(n_sub):
            r0, r1 = cuts[k], cuts[k + 1]
            if r1 <= r0:
                continue
            o0, o1 = int(ob[r0]), int(ob[r1])
            arrs = [torch.from_numpy(ob[r0:r1 + 1] - o0).pin_memory(), host[1][o0:o1].clone().pin_memory()]
            arrs += [h[r0:r1].clone().pin_memory() for h in host[2:]]
            pb = torch.from_numpy(pre_b[pre_o[r0]:pre_o[r1]].copy()).pin_memory()
            po = torch.from_numpy((pre_o[r0:r1 + 1] - pre_o[r0]).copy()).pin_memory()
            nps = [a.numpy() for a in arrs]
            nps[1] = nps[1].view(np.uint64)
            subs.append({"arrs": arrs, "np": nps, "pb": pb, "po": po, "out": None, "n": 0})
            h2d += sum(a.numel() * a.element_size() for a in arrs) + pb.numel() + po.numel() * 8

        stage_s = [0.0, 0.0, 0.0]

        def run_sub(c, sb):
            t_a = time.perf_counter()
            c.encode_batch(*sb["np"])
            t_b = time.perf_counter()
            nbytes, _ = c.serialise(sb["pb"].numpy(), sb["po"].numpy())
            t_c = time.perf_counter()
            stage_s[0] += t_b - t_a
            stage_s[1] += t_c - t_b
            if sb["out"] is None or sb["out"].numel() < nbytes:
                sb["out"] = torch.empty(int(nbytes * 1.05) + 64, dtype=torch.uint8).pin_memory()
            c.fetch_text(pinned=sb["out"].numpy(), want_lines=False)
            sb["n"] = nbytes
            stage_s[2] += time.perf_counter() - t_c
            return nbytes

        def worker(c, mine_, n_steps):
            return sum(run_sub(c, sb) for _ in range(n_steps) for sb in mine_)

        pool = ThreadPoolExecutor(len(ctxs))

        def e2e_steps(n_steps):
            # the K steps stream through the contexts back to back, as the batches of a file do: a context starts
            # its share of step k+1 as soon as it has delivered its share of step k (no drain between steps)
            futs = [pool.submit(worker, c, subs[k::len(ctxs)], n_steps) for k, c in enumerate(ctxs)]
            return sum(f.result() for f in futs) // n_steps

        d2h = e2e_steps(args.warmup)
        barrier()
        stage_s[:] = [0.0, 0.0, 0.0]
        t0 = time.perf_counter()
        d2h = e2e_steps(args.steps)
        own = time.perf_counter() - t0
        barrier()
        el = max_over_ranks(own)
        if world == 1:
            import hashlib
            h = hashlib.md5()
            for sb in subs:
                h.update(sb["out"].numpy()[:sb["n"]].tobytes())
            soa_md5 = h.hexdigest()
        e2e_soa = {"value": A_total * args.steps / el, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": 1e3 * el / args.steps,
               "what": "pafgpu_encode_batch_host + pafgpu_serialise + pafgpu_fetch_text on parsed SoA arrays and host-rendered line heads in pinned memory",
               "pipeline": f"{len(subs)} sub-batches per step dealt round-robin to {len(ctxs)} sibling contexts (one stream each, ONE packed genome) on the GPU; the K steps stream back to back, all of them inside the timed region",
               "host": {"cpus_visible": len(os.sched_getaffinity(0)), "affinity": affinity},
               "gpu_launches_per_step": KERNELS_PER_E2E_SUBBATCH * len(subs),
               "host_wall_ms_per_step_summed_over_threads": {"encode_batch_host": 1e3 * stage_s[0] / args.steps,
                                                              "serialise": 1e3 * stage_s[1] / args.steps,
                                                              "fetch_text": 1e3 * stage_s[2] / args.steps}}
        del subs

        # ---- `e2e`: PAF TEXT in pinned host memory -> PAF TEXT in pinned host memory (pafgpu_add_mismatches_text +
        # pafgpu_fetch_text): what the reference's interface carries, nothing pre-parsed and nothing pre-rendered - the
        # text is copied to the device, split into lines and fields, parsed, looked up, encoded, checked and rendered
        # there. Same sub-batch pipeline over the same sibling contexts.
        text = make_input_text(ctx, rec, pre_b, pre_o)
        nl_targets = [len(text) * k // n_sub for k in range(1, n_sub)]
        tcuts = [0]
        for t in nl_targets:
            w = text[t:t + (1 << 20)]
            tcuts.append(t + int(np.argmax(w == 10)) + 1)
        tcuts.append(len(text))
        tsubs = []
        for k in range(n_sub):
            a, b = tcuts[k], tcuts[k + 1]
            if b > a:
                tsubs.append({"in": torch.from_numpy(text[a:b].copy()).pin_memory(), "out": None, "n": 0})
        h2d_t = sum(sb["in"].numel() for sb in tsubs)

        tstage = {"add_mismatches_text": 0.0, "fetch_text": 0.0, "dev_parse": 0.0, "dev_ingest": 0.0, "dev_encode": 0.0, "dev_render": 0.0}

        def run_sub_text(c, sb):
            t_a = time.perf_counter()
            res_t = c.add_mismatches_text_resident(sb["in"].numpy())
            t_b = time.perf_counter()
            tstage["add_mismatches_text"] += t_b - t_a
            tstage["dev_parse"] += 1e-3 * res_t.ms_parse; tstage["dev_ingest"] += 1e-3 * res_t.ms_ingest
            tstage["dev_encode"] += 1e-3 * res_t.ms_encode; tstage["dev_render"] += 1e-3 * res_t.ms_render
            assert res_t.first_bad < 0
            nbytes = res_t.text_bytes
            if sb["out"] is None or sb["out"].numel() < nbytes:
                sb["out"] = torch.empty(int(nbytes * 1.05) + 64, dtype=torch.uint8).pin_memory()
            t_c = time.perf_counter()
            c.fetch_text(pinned=sb["out"].numpy(), want_lines=False)
            tstage["fetch_text"] += time.perf_counter() - t_c
            sb["n"] = nbytes
            return nbytes

        def text_steps(n_steps):
            futs = [pool.submit(lambda c, mine_: sum(run_sub_text(c, sb) for _ in range(n_steps) for sb in mine_), c, tsubs[k::len(ctxs)])
                    for k, c in enumerate(ctxs)]
            return sum(f.result() for f in futs) // n_steps

        d2h_t = text_steps(args.warmup)
        barrier()
        for k_ in tstage:
            tstage[k_] = 0.0
        t0 = time.perf_counter()
        d2h_t = text_steps(args.steps)
        own = time.perf_counter() - t0
        barrier()
        el = max_over_ranks(own)
        text_md5 = None
        if world == 1:
            import hashlib
            h = hashlib.md5()
            for sb in tsubs:
                h.update(sb["out"].numpy()[:sb["n"]].tobytes())
            text_md5 = h.hexdigest()
        e2e = {"value": A_total * args.steps / el, "unit": UNIT, "h2d_bytes_per_step": int(h2d_t), "d2h_bytes_per_step": int(d2h_t),
               "ms_per_step": 1e3 * el / args.steps,
               "what": "pafgpu_add_mismatches_text + pafgpu_fetch_text: PAF text in pinned host memory -> device (lines, fields, cg ops parsed there; contigs looked up; "
                       "encode; paf_check; whole lines rendered) -> PAF text in pinned host memory",
               "pipeline": f"{len(tsubs)} chunks of whole lines per step dealt round-robin to {len(ctxs)} sibling contexts (one stream each, ONE packed genome); the K steps stream back to back",
               "gpu_launches_per_step": KERNELS_PER_TEXT_CHUNK * len(tsubs),
               "wall_ms_per_step_summed_over_contexts": {k_: 1e3 * v_ / args.steps for k_, v_ in tstage.items()},
               "output_equals_soa_path_output": (text_md5 == soa_md5) if text_md5 else None,
               "soa": e2e_soa}
        del text, tsubs
        pool.shutdown()
        for c2 in ctxs[1:]:
            c2.close()

    # ---- end to end through the reference's own interface: PAF file in, PAF file out ---------------------------------
    # `paffy add_mismatches -i in.paf -o out.paf` with the FASTA already packed (its load is reported beside it): the text
    # is read in chunks, parsed, encoded and rendered on the device, written in input order (pafgpu_add_mismatches_file,
    # what paffy_add_mismatches_main runs). Files live in tmpfs. At N > 1 rank 0 drives all N GPUs of the box from one
    # process (the product's multi-GPU path) on the first million records of the job; the other ranks wait.
    e2e_file = None
    if not args.no_e2e_file:
        barrier()
        if rank == 0:
            try:
                base = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
                d = tempfile.mkdtemp(prefix="paffy_b200_bench_", dir=base)
                sub = synth.shard(rec, 0, min(R, 1_000_000)) if R > 1_000_000 else rec
                if pre_b is None or sub is not rec:
                    pre_b2, pre_o2 = render_prefixes(sub)
                else:
                    pre_b2, pre_o2 = pre_b, pre_o
                text = make_input_text(ctx, sub, pre_b2, pre_o2)
                fin, fout = os.path.join(d, "in.paf"), os.path.join(d, "out.paf")
                text.tofile(fin)
                A_file = int(sub["sum_m"].sum().item())
                del text
                rep = None
                for _ in range(3):
                    rc, rep = ctx.add_mismatches_file(fin, fout, gpus=world, streams=args.file_streams)
                    assert rc == 0, ctx.last_error()
                k_file = max(3, min(args.steps, 5))
                el = 0.0
                for _ in range(k_file):
                    if os.path.exists(fout):
                        os.unlink(fout)  # every pass writes a fresh file (freeing the old one's pages is not part of the pass)
                    t0 = time.perf_counter()
                    rc, rep = ctx.add_mismatches_file(fin, fout, gpus=world, streams=args.file_streams)
                    el += time.perf_counter() - t0
                el /= k_file
                out_md5 = md5_file(fout)
                # the same passes with the output thrown away (/dev/null): what the pipeline does when the page cache of the
                # output file is not the limit (creating a fresh 0.67 GB file costs ~0.15 s of kernel time on this box)
                t0 = time.perf_counter()
                for _ in range(k_file):
                    ctx.add_mismatches_file(fin, "/dev/null", gpus=world, streams=args.file_streams)
                el_null = (time.perf_counter() - t0) / k_file
                e2e_file = {"value": A_file / el, "unit": UNIT, "s_per_pass": el, "passes": k_file, "records": int(sub["n"]),
                            "aligned_bp": A_file, "input_bytes": os.path.getsize(fin), "output_bytes": os.path.getsize(fout),
                            "h2d_bytes_per_step": os.path.getsize(fin), "d2h_bytes_per_step": os.path.getsize(fout),
                            "gpus": int(rep.gpus), "workers": int(rep.workers), "chunks": int(rep.chunks),
                            "stage_s_summed_over_workers": {"read": rep.s_read, "gpu": rep.s_gpu, "write": rep.s_write},
                            "what": "pafgpu_add_mismatches_file (the record loop of paffy_add_mismatches_main): PAF text file in tmpfs -> PAF text file in tmpfs, "
                                    "text parsed and rendered on the device; genome already packed in HBM",
                            "output_to_dev_null": {"value": A_file / el_null, "s_per_pass": el_null},
                            "output_md5": out_md5,
                            "equals_soa_path_output": (out_md5 == soa_md5) if (soa_md5 and sub is rec) else None}
                for f in (fin, fout):
                    os.unlink(f)
                os.rmdir(d)
            except Exception as e:  # noqa: BLE001 - the extra figure must not take the bench line down
                e2e_file = {"error": repr(e)[:300]}
        barrier()

    # restore the whole batch as the library's last batch (the roofline section below re-reads its result)
    res = ctx.encode_batch(*dev_arrays)

    # ---- roofline of the dominant kernel ---------------------------------------------------------------
    peak, peak_src = peaks()
    b_alg = 0.5 * A + 8.0 * N + 8.0 * n_out  # 2 x A/4 sequence bytes + 8 B per input op + 8 B per output op
    k_ms = ms_kernel / args.steps
    s_ms = ms_stage / args.steps
    # One kernel (encode_kernel) reads the sequences and the op table and writes the output ops; the roofline is
    # all algorithmic bytes / its CUDA-event time (measured on the library's stream around that launch alone).
    achieved = b_alg / (k_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel": "encode_kernel (op table + compare + look-back + emit, single pass)", "kernel_ms": k_ms,
                "algorithmic_bytes_per_launch": b_alg, "bytes_per_aligned_bp": b_alg / A, "peak_source": peak_src,
                "kernel_share_of_step": ms_kernel / (elapsed * 1e3) if world == 1 else None,
                "encode_stage": {"ms": s_ms, "achieved": b_alg / (s_ms * 1e-3) / 1e9, "frac": b_alg / (s_ms * 1e-3) / 1e9 / peak,
                                 "what": "resets + encode_kernel + stitch_kernel + host read of the op count"}}
    # stage 3 + 4, the encode stage followed by the device serialiser on the resident result (line heads
    # rendered by the host and uploaded beforehand; the serialiser's own CUDA-event time), with its second byte figure:
    # the algorithmic bytes above + the output ops read again + the line heads read + every byte of the PAF text written
    if pre_b is not None and world == 1:
        try:
            ser_ms = []
            for _ in range(1 + min(args.steps, 5)):
                t_bytes, ms = ctx.serialise(pre_b, pre_o)
                ser_ms.append(ms)
            ser = sum(ser_ms[1:]) / len(ser_ms[1:])
            b34 = b_alg + 8.0 * n_out + float(len(pre_b)) + float(t_bytes)
            roofline["stage_3_plus_4"] = {"ms": s_ms + ser, "serialiser_ms": ser, "bytes": b34, "achieved": b34 / ((s_ms + ser) * 1e-3) / 1e9,
                                          "frac": b34 / ((s_ms + ser) * 1e-3) / 1e9 / peak, "text_bytes": int(t_bytes),
                                          "what": "encode stage + ser_block_records / ser_write / prefix_copy kernels (pafgpu_serialise), outputs left in HBM"}
        except Exception as e:  # noqa: BLE001
            roofline["stage_3_plus_4"] = {"error": repr(e)[:200]}
    try:
        mx, mean, sms = ctx.encode_balance()
        roofline["load_balance"] = {"max_sm_busy_ms": mx, "mean_sm_busy_ms": mean, "max_over_mean": mx / mean if mean else None, "sms": sms,
                                    "what": "per SM: the longest span from first to last tile among its persistent CTAs (globaltimer), last launch"}
    except Exception as e:  # noqa: BLE001
        roofline["load_balance"] = {"error": repr(e)[:120]}
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        try:
            t = json.load(open(tr)).get(args.workload)
            if t:
                roofline["traffic"] = t["bytes"] if isinstance(t, dict) else t
                roofline["traffic_source"] = t.get("source") if isinstance(t, dict) else "ncu --set full capture under profiles/ (see profiles/traffic.json)"
        except Exception:
            pass

    # ---- CPU baseline (rank 0, N = 1 only) ---------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        d, fa, paf, _ = cpu_sample(args.workload, args.scale, args.cpu_sample, dev, args.cpu_contigs)
        out = run_reference(fa, paf, cores, 2)
        if out is not None:
            r = out["reps"][-1]
            one = run_reference(fa, paf, 1, 1)["reps"][0] if cores > 1 else r
            cpu = {"value": r["aligned_bp"] / r["encode_s"], "unit": UNIT, "cores": cores, "kind": "reference",
                   "sample": f"{r['records']} records of {args.workload}'s generator on {args.cpu_contigs} contig pairs (both strands), {r['aligned_bp']} aligned bp; "
                             f"unmodified paf_encode_mismatches on {cores} pthreads, FASTA load ({out['load_s']:.1f} s) and PAF parse ({r['parse_s']:.2f} s) excluded",
                   "single_thread_value": one["aligned_bp"] / one["encode_s"], "fasta_load_s": out["load_s"], "parse_s": r["parse_s"]}
            cli = run_reference_cli(fa, paf, d)
            if cli:
                cli["aligned_bp"] = r["aligned_bp"]
                cli["single_process_bp_per_s_end_to_end"] = r["aligned_bp"] / cli["single_process_s"]
                if cli.get("parallel_wall_s"):
                    cli["parallel_bp_per_s_end_to_end"] = r["aligned_bp"] / cli["parallel_wall_s"]
                cli["what"] = ("reference CLI `paffy add_mismatches` file to file on the same sample (FASTA load, parse, encode, print): one process, "
                               "then one process per `paffy split_file` shard (per target contig) run concurrently, every process loading the whole FASTA")
                cpu["reference_cli"] = cli
                try:
                    # the sample's records live on contigs of the full genome already packed in `ctx`: the byte comparison
                    # runs against the 24-contig arena, not a genome of the sample's contigs alone
                    fp = file_parity(fa, paf, d, ctx)
                except Exception as e:  # the checker must not take the bench line down
                    fp = {"error": repr(e)}
                if fp:
                    parity["reference_cli"] = fp
                    for k in ("add_bytes_equal", "trim_bytes_equal", "view_summary_equal"):
                        cpu["parity_" + k] = fp.get(k)

    if rank == 0:
        wl = synth.CONFIGS[args.workload]
        config = {"workload": f"{args.workload}: {wl[0]}x{wl[1]} bp target+query genomes, {job_records} records in the job"
                              + (f", sharded over {world} GPUs by aligned bases (contiguous ranges, pafgpu_partition)" if strong else f", {R} records/GPU")
                              + f", scale {args.scale}",
                  "records_this_rank": int(R), "input_ops": int(N), "output_ops": n_out,
                  "aligned_bp_per_step_this_rank": A, "aligned_bp_per_step_all_ranks": int(A_total),
                  "sharding": ("ONE record list cut into contiguous ranges of near-equal aligned bases, one range per rank" if strong
                               else "one record list per rank") + "; packed genome replicated; no data-path collective (outputs concatenate in rank order)",
                  "shards": shards,
                  "l2": "inputs larger than L2 (packed genome 1.5 GB + op arrays); no flush needed",
                  "timing": "K synchronous steps; every rank times its own K steps between the opening barrier and its own last synchronize, the maximum over ranks is taken",
                  "parity_generator_truth": parity["mismatch_counts_equal"] and parity["output_op_counts_equal"]}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 * elapsed / args.steps, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None,
                "dtype": "u8", "data": "synthetic", "config": config,
                "device_ms": {"ingest": ms_ingest / args.steps, "encode_stage": ms_stage / args.steps, "encode_kernel": k_ms,
                              "stitch_kernel": ms_emit / args.steps, "retried": int(res.retried)},
                "gpu_launches": KERNELS_PER_STEP * args.steps, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "e2e_file": e2e_file,
                "clocks": clk, "parity": parity, "gen_s": t_gen}
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
