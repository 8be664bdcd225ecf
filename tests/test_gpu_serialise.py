"""The one-pass device serialiser (serialise.cuh: ser_write_kernel) against a plain restatement of paf_write_to_buffer's
CIGAR loop (impl/paf.c:367-380: "<signed decimal length><op letter>", letters MID=X, 'N' for any other code) on op
lists the encode path never produces: lengths of every magnitude, negative lengths, unknown op codes, records without
ops (single ones and runs longer than a block's record table), records longer than several blocks. Driven through
pafgpu_serialise_no_alloc, which takes the op words as they are."""
import numpy as np
import pytest

from paffy_b200 import lib as L

pytestmark = pytest.mark.gpu

LETTERS = "MID=X"


def render(ops, op_begin, heads):
    """The text the reference would print for these records (heads = the bytes before each CIGAR)."""
    out, lines = [], [0]
    for r, head in enumerate(heads):
        parts = [head]
        for w in ops[op_begin[r]:op_begin[r + 1]]:
            w = int(w)
            n = w & ((1 << 56) - 1)
            if n >= 1 << 55:
                n -= 1 << 56
            code = w >> 56
            parts.append(f"{n}{LETTERS[code] if code <= 4 else 'N'}".encode())
        parts.append(b"\n")
        out.append(b"".join(parts))
        lines.append(lines[-1] + len(out[-1]))
    return b"".join(out), np.array(lines, dtype=np.int64)


def run_device(ops, op_begin, heads, slack=64):
    import torch
    dev = torch.device("cuda", 0)
    R = len(heads)
    hb = np.zeros(R + 1, dtype=np.int64)
    hb[1:] = np.cumsum([len(h) for h in heads])
    head_bytes = b"".join(heads) or b"\0"
    ops_d = torch.from_numpy(ops.view(np.int64)).to(dev) if len(ops) else torch.zeros(1, dtype=torch.int64, device=dev)
    begin_d = torch.from_numpy(op_begin).to(dev)
    heads_d = torch.frombuffer(bytearray(head_bytes), dtype=torch.uint8).to(dev)
    hb_d = torch.from_numpy(hb).to(dev)
    want, want_lines = render(ops, op_begin, heads)
    cap = len(want) + slack
    text_d = torch.full((cap + 256,), 0xEE, dtype=torch.uint8, device=dev)  # 256 guard bytes behind what the library is told about
    lines_d = torch.full((R + 1,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    try:
        total = L.serialise_no_alloc(0, ops_d.data_ptr(), len(ops), begin_d.data_ptr(), R, heads_d.data_ptr(), hb_d.data_ptr(), text_d.data_ptr(), cap,
                                     lines_d.data_ptr())
    finally:
        torch.cuda.synchronize()
        got = text_d.cpu().numpy()
        assert (got[cap:] == 0xEE).all()  # nothing behind the buffer was touched, whatever happened
    assert total == len(want)
    assert (got[len(want):] == 0xEE).all()
    assert (lines_d.cpu().numpy() == want_lines).all()
    assert got[:len(want)].tobytes() == want


def op_words(rng, n, kind):
    if kind == "short":
        lens = rng.integers(0, 10000, n)
    elif kind == "mixed":
        mags = rng.integers(0, 17, n)  # decimal magnitudes 1 .. 17 digits (below 2^55)
        lens = np.array([int(rng.integers(10 ** (m - 1) if m else 0, min(10 ** m, 1 << 55) if m else 1)) for m in mags], dtype=np.int64)
        neg = rng.random(n) < 0.15
        lens = np.where(neg, -lens, lens)
    else:  # edges
        pool = np.array([0, 1, 9, 10, 99, 100, 999, 1000, 9999, 10000, 99999, (1 << 32) - 1, 1 << 32, (1 << 32) + 1, 10 ** 10, (1 << 55) - 1,
                         -1, -9, -10, -9999, -10000, -(1 << 32), -(1 << 55)], dtype=np.int64)
        lens = pool[rng.integers(0, len(pool), n)]
    codes = rng.integers(0, 7 if kind != "short" else 5, n).astype(np.uint64)
    return (lens.astype(np.int64).view(np.uint64) & np.uint64((1 << 56) - 1)) | (codes << np.uint64(56))


def make_records(rng, counts):
    op_begin = np.zeros(len(counts) + 1, dtype=np.int64)
    op_begin[1:] = np.cumsum(counts)
    heads = [(b"q%d\t1\t0\t1\t+\tt\t1\t0\t1\t0\t0\t%d\tcg:Z:" % (i, int(rng.integers(0, 255)))) if c else (b"q%d\tno\tcigar" % i) for i, c in enumerate(counts)]
    return op_begin, heads


@pytest.mark.parametrize("kind", ["short", "mixed", "edges"])
def test_op_text_of_every_magnitude(kind):
    rng = np.random.default_rng(11)
    counts = rng.integers(1, 60, 700)
    op_begin, heads = make_records(rng, counts)
    run_device(op_words(rng, int(op_begin[-1]), kind), op_begin, heads)


def test_records_without_ops_and_records_longer_than_blocks():
    rng = np.random.default_rng(12)
    # a run of empty records first, single empties between, a run longer than a block's record table (2056) in the middle,
    # records of 1 .. 7000 ops (a block holds 2048), empties at the very end
    counts = [0] * 5 + [3, 0, 1, 0, 0, 7000, 1, 2048, 2047, 2049] + [0] * 2500 + [5, 4096] + [1] * 3000 + [0, 0, 900, 0, 0, 0]
    op_begin, heads = make_records(rng, counts)
    run_device(op_words(rng, int(op_begin[-1]), "mixed"), op_begin, heads)


def test_block_whose_text_does_not_fit_the_staging_buffer():
    # 2048 ops of 18 bytes each = 36 KB of text in one block: more than the 16 KB a block stages in shared memory
    rng = np.random.default_rng(13)
    n = 5000
    lens = np.full(n, -(10 ** 15), dtype=np.int64)
    ops = (lens.view(np.uint64) & np.uint64((1 << 56) - 1)) | (rng.integers(0, 5, n).astype(np.uint64) << np.uint64(56))
    op_begin, heads = make_records(rng, [1200, 0, 3000, 800])
    run_device(ops, op_begin, heads)


def test_no_ops_at_all_and_no_records():
    rng = np.random.default_rng(14)
    op_begin, heads = make_records(rng, [0, 0, 0])
    run_device(np.zeros(0, dtype=np.uint64), op_begin, heads)
    run_device(np.zeros(0, dtype=np.uint64), np.zeros(1, dtype=np.int64), [])


def test_text_buffer_too_small_is_reported_and_nothing_is_written_behind_it():
    import paffy_b200
    rng = np.random.default_rng(15)
    op_begin, heads = make_records(rng, [50, 60, 70])
    ops = op_words(rng, 180, "short")
    with pytest.raises(paffy_b200.PafGpuError) as e:
        run_device(ops, op_begin, heads, slack=-100)
    assert "the buffer holds" in str(e.value)
