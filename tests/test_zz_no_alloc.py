"""pafgpu_serialise_no_alloc (include/paffy_b200.h): the serialiser with its scratch in a fixed device array, for callers
that cannot allocate device memory. Collected last on purpose. Its output is compared
with the existing entry point's (pafgpu_serialise + pafgpu_fetch_text), byte for byte - the tolerance of every text test
here - on two calls in a row that share the scratch array."""
import numpy as np
import pytest

import paffy_b200
from paffy_b200 import lib as L
from tests import util

pytestmark = pytest.mark.gpu


def _prefixes(case):
    """The text before each record's CIGAR, rendered on the host as a caller of pafgpu_serialise does."""
    heads = [ln[: ln.index("\tcg:Z:") + 6].encode() for ln in util.case_lines(case)]
    begin = np.zeros(len(heads) + 1, dtype=np.int64)
    begin[1:] = np.cumsum([len(h) for h in heads])
    return b"".join(heads), begin


def _both_ways(ctx, case):
    """-> (text, line offsets) from the existing entry point, and from the no-alloc one on the same encoded batch."""
    import torch
    util.gpu_encode(ctx, case)
    heads, begin = _prefixes(case)
    n, _ = ctx.serialise(heads, begin)
    want_text, want_lines = ctx.fetch_text()
    want_text, want_lines = want_text.tobytes(), want_lines.copy()
    assert n == len(want_text)

    # the caller's side of the new entry point: inputs and outputs in device memory it owns (the test's own allocations)
    R = ctx.n_records
    ops_dev, begin_dev, _ = ctx.result_device_ptrs()
    _, out_begin, _ = ctx.fetch_ops(want_ops=False)
    n_ops = int(out_begin[-1])
    dev = torch.device("cuda", 0)
    heads_d = torch.frombuffer(bytearray(heads), dtype=torch.uint8).to(dev)
    hbegin_d = torch.from_numpy(begin).to(dev)
    text_d = torch.full((n + 64,), 0xEE, dtype=torch.uint8, device=dev)
    lines_d = torch.full((R + 1,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    total = L.serialise_no_alloc(0, ops_dev, n_ops, begin_dev, R, heads_d.data_ptr(), hbegin_d.data_ptr(), text_d.data_ptr(), n + 64,
                                 lines_d.data_ptr())
    got = text_d.cpu().numpy()
    assert total == n
    assert (got[n:] == 0xEE).all()  # nothing behind the text was touched
    return want_text, want_lines, got[:n].tobytes(), lines_d.cpu().numpy()


def test_no_alloc_serialiser_equals_the_existing_one_twice_in_a_row(ctx):
    # two calls in a row, the second on a smaller batch: it meets the first call's leftovers in the shared scratch array
    for seed, n_records, n_ops in ((71, 300, 40), (72, 37, 9)):
        case = util.random_case(seed, n_records=n_records, contig_len=30000, n_ops=n_ops, max_len=2500, p_weird=0.001)
        # (the SoA batch interface wants at least one op per record once zero-length M ops are dropped)
        case["records"] = [r for r in case["records"] if any(not (c == "M" and n == 0) for n, c in r["ops"])]
        want_text, want_lines, got_text, got_lines = _both_ways(ctx, case)
        assert got_text == want_text
        assert (got_lines == want_lines).all()


def test_no_alloc_serialiser_rejects_what_its_scratch_cannot_hold():
    with pytest.raises(paffy_b200.PafGpuError) as e:
        L.serialise_no_alloc(0, 1, L.NO_ALLOC_MAX_OPS + 1, 1, 1, 1, 1, 1, 1, 1)
    assert e.value.code == -2 and "exceed the limit" in str(e.value)
