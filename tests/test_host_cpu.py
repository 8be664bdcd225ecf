"""CPU-only checks of the host logic that needs no base comparison: `add_mismatches -a` (text in, merged M ops out,
byte-compared with the reference CLI's committed output), record partitioning for multi-GPU sharding (single process
and world_size-2 gloo)."""
import os

import numpy as np
import pytest

import paffy_b200

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("stem", ["edge", "rand101", "rand102", "rand103"])
def test_remove_mismatches_file_matches_reference(stem, tmp_path):
    out = tmp_path / "o.paf"
    rc, rep = paffy_b200.remove_mismatches_file(os.path.join(G, stem + ".add.paf"), str(out))
    assert rc == 0
    assert out.read_bytes() == open(os.path.join(G, stem + ".remove.paf"), "rb").read()


def test_remove_mismatches_bad_record_ends_run(tmp_path):
    good = "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:10=2X13="
    bad = "q\t25\t0\t10\t+\tt\t25\t0\t9\t0\t0\t60\tcg:Z:9="
    inp, out = tmp_path / "i.paf", tmp_path / "o.paf"
    inp.write_text("\n".join([good, bad, good]) + "\n")
    rc, rep = paffy_b200.remove_mismatches_file(str(inp), str(out))
    assert rc == 1 and rep.failed_record == 1
    assert out.read_text() == "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tAS:i:0\tcg:Z:25M\n"


def test_partition_properties():
    rng = np.random.default_rng(3)
    for parts in (1, 2, 3, 8):
        w = rng.integers(1, 5_000_000, 5000)
        cuts = paffy_b200.partition(w, parts)
        assert cuts[0] == 0 and cuts[-1] == len(w) and (np.diff(cuts) >= 0).all()
        loads = np.add.reduceat(w, cuts[:-1])[: parts] if parts > 1 else np.array([w.sum()])
        assert abs(loads - w.sum() / parts).max() <= w.max()  # imbalance bounded by one record
    assert list(paffy_b200.partition([], 3)) == [0, 0, 0, 0]
    assert list(paffy_b200.partition([7], 4)) in ([0, 0, 0, 0, 1], [0, 0, 0, 1, 1], [0, 0, 1, 1, 1], [0, 1, 1, 1, 1])


def _rank_main(rank, world, port, q):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    from paffy_b200 import shard
    rng = np.random.default_rng(11)
    weights = rng.integers(100, 100000, 1000)
    lo, hi = shard.my_range(weights, rank, world)
    mine = np.arange(lo, hi, dtype=np.int64)
    gathered = shard.gather_in_order(mine.tobytes(), rank, world)
    if rank == 0:
        q.put((np.frombuffer(gathered, dtype=np.int64).tolist() == list(range(1000)), int(weights[lo:hi].sum()), int(weights.sum())))
    dist.destroy_process_group()


def test_sharding_world_size_2_gloo():
    """Two ranks take contiguous base-balanced ranges; rank 0 gathers the outputs in input order (no data-path
    collective: a gather of finished bytes only)."""
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    ok, mine, total = q.get(timeout=120)
    [p.join(60) for p in ps]
    assert ok and abs(mine - total / 2) < 100000


def test_cli_option_handling_without_a_gpu():
    """What the subcommand entry points decide before any device is touched, against the reference's mains: `view` without
    a FASTA exits 1 (impl/paf_view.c:110-113), -h prints the usage and returns 0, an unknown flag returns 1
    (impl/paf_trim.c:89-91), and nothing falls back to a CPU path: with no device the commands that compare bases fail."""
    import subprocess
    exe = os.path.join(os.path.dirname(paffy_b200.lib.LIB_PATH), "paffy_b200_cli")
    r = subprocess.run([exe, "view", "-i", os.path.join(G, "edge.paf")], capture_output=True)
    assert r.returncode == 1 and b"Expected at least one sequence file" in r.stderr and r.stdout == b""
    for cmd in ("add_mismatches", "view", "trim"):
        r = subprocess.run([exe, cmd, "-h"], capture_output=True)
        assert r.returncode == 0 and b"paffy " + cmd.encode() in r.stderr, cmd
        r = subprocess.run([exe, cmd, "--noSuchOption"], capture_output=True)
        assert r.returncode == 1, cmd
    if paffy_b200.load().pafgpu_device_count() == 0:
        for args in (["add_mismatches", os.path.join(G, "edge.fa"), "-i", os.path.join(G, "edge.paf")],
                     ["view", "-a", "-s", os.path.join(G, "edge.fa"), "-i", os.path.join(G, "edge.paf")],
                     ["view", os.path.join(G, "edge.fa"), "-i", os.path.join(G, "edge.paf")],
                     ["trim", "-i", os.path.join(G, "rand101.add.paf")]):
            r = subprocess.run([exe] + args, capture_output=True)
            assert r.returncode == 1 and r.stdout == b"" and b"ERROR" in r.stderr, args


# ---- the streaming pipeline (reader thread -> chunk workers -> ordered writer), on the mode that needs no GPU ----------
@pytest.mark.parametrize("chunk", [64, 1000, 1 << 20])
def test_chunked_pipeline_gives_the_same_bytes(chunk, tmp_path):
    """Chunks of whole lines processed by concurrent workers and written in input order: same bytes whatever the chunk size
    (64 bytes = every line its own chunk, lines longer than a chunk included)."""
    out = tmp_path / "o.paf"
    rc, rep = paffy_b200.remove_mismatches_file(os.path.join(G, "rand102.add.paf"), str(out), chunk_bytes=chunk)
    assert rc == 0
    want = open(os.path.join(G, "rand102.remove.paf"), "rb").read()
    assert out.read_bytes() == want
    assert rep.records == want.count(b"\n") and rep.workers == 2
    if chunk == 64:
        assert rep.chunks > rep.records // 4


def test_failure_in_a_later_chunk_keeps_input_order(tmp_path):
    good = "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:10=2X13="
    bad = "q\t25\t0\t10\t+\tt\t25\t0\t9\t0\t0\t60\tcg:Z:9="
    inp, out = tmp_path / "i.paf", tmp_path / "o.paf"
    inp.write_text("\n".join([good] * 137 + [bad] + [good] * 200) + "\n")
    rc, rep = paffy_b200.remove_mismatches_file(str(inp), str(out), chunk_bytes=300)
    assert rc == 1 and rep.failed_record == 137 and rep.records == 137
    assert out.read_text() == "q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tAS:i:0\tcg:Z:25M\n" * 137


def test_output_starts_before_the_input_ends():
    """The reference streams record by record (impl/paf_add_mismatches.c:106-140): in a pipe the first lines must come out
    while the producer still holds its end open."""
    import select
    import subprocess
    exe = os.path.join(os.path.dirname(paffy_b200.lib.LIB_PATH), "paffy_b200_cli")
    p = subprocess.Popen([exe, "add_mismatches", "-a"], stdin=subprocess.PIPE, stdout=subprocess.PIPE, bufsize=0)
    line = b"q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tcg:Z:10=2X13=\n"
    p.stdin.write(line * 3)
    ready, _, _ = select.select([p.stdout], [], [], 20)
    assert ready, "no output while stdin is still open"
    first = os.read(p.stdout.fileno(), 1 << 16)
    assert first.startswith(b"q\t25\t0\t25\t+\tt\t25\t0\t25\t0\t25\t60\tAS:i:0\tcg:Z:25M\n")
    p.stdin.write(line * 2)
    p.stdin.close()
    rest = p.stdout.read()
    assert p.wait() == 0 and (first + rest).count(b"\n") == 5


def _strong_rank(rank, world, port, q):
    import torch
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    from paffy_b200 import synth
    # what bench.py does at N > 1: every rank generates the SAME job, cuts it with pafgpu_partition, keeps its own range
    rec = synth.build_workload(None, "tiny", torch.device("cpu"), rank=0)
    cuts = paffy_b200.partition(rec["sum_m"].numpy(), world)
    lo, hi = int(cuts[rank]), int(cuts[rank + 1])
    mine = synth.shard(rec, lo, hi)
    assert mine["n"] == hi - lo and int(mine["op_begin"][0]) == 0 and int(mine["op_begin"][-1]) == mine["ops"].numel()
    payload = (lo, hi, int(mine["sum_m"].sum()), mine["ops"].numpy().tobytes(), mine["qs"].numpy().tobytes())
    objs = [None] * world if rank == 0 else None
    dist.gather_object(payload, objs, dst=0)
    if rank == 0:
        whole_ops = rec["ops"].numpy().tobytes()
        ok = objs[0][0] == 0 and objs[-1][1] == rec["n"] and all(objs[i][1] == objs[i + 1][0] for i in range(world - 1))
        ok = ok and b"".join(o[3] for o in objs) == whole_ops and b"".join(o[4] for o in objs) == rec["qs"].numpy().tobytes()
        q.put((ok, [o[2] for o in objs], int(rec["sum_m"].sum()), int(rec["sum_m"].max())))
    dist.destroy_process_group()


def test_one_job_strong_scaling_shards_world_size_2_gloo():
    """bench.py's N > 1 path: ONE input-ordered record list cut into contiguous ranges of near-equal aligned bases, the
    rank-ordered concatenation of the shards is the job again (no collective on the data path)."""
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_strong_rank, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    ok, loads, total, biggest = q.get(timeout=180)
    [p.join(60) for p in ps]
    assert ok and sum(loads) == total and abs(loads[0] - total / 2) <= biggest


def test_large_output_goes_through_the_mapped_writer(tmp_path):
    """Pieces of 8 MiB and more of a regular output file are written by several threads through a mapping of the file
    (extended chunk by chunk); the bytes must be the same as through write(2) ($PAFFY_NO_MMAP)."""
    lines = open(os.path.join(G, "rand101.add.paf"), "rb").read()
    want1 = open(os.path.join(G, "rand101.remove.paf"), "rb").read()
    reps = (20 << 20) // len(lines) + 1
    inp = tmp_path / "big.paf"
    inp.write_bytes(lines * reps)
    outs = []
    for env in ({}, {"PAFFY_NO_MMAP": "1"}):
        os.environ.update(env)
        try:
            out = tmp_path / f"o{len(outs)}.paf"
            rc, rep = paffy_b200.remove_mismatches_file(str(inp), str(out), chunk_bytes=12 << 20)
        finally:
            for k in env:
                del os.environ[k]
        assert rc == 0 and rep.chunks >= 2
        outs.append(out.read_bytes())
    assert outs[0] == want1 * reps and outs[1] == outs[0]
