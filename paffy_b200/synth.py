"""Seeded synthetic genomes + PAF records of the shapes BASELINE.json names.

Bench / test input generation only (torch ops, runs on CPU for small cases and on the GPU at full
size); nothing here is on the measured path. The generator knows the true per-record mismatch count
and =/X run count, which gives size-independent parity checks at full scale.

Model: query contig i = mutate(target contig i) with a contig-wide edit script (substitutions at rate
s to a different base; indel events at rate g, 50/50 I/D, length ~ Geometric(mean 3) capped at 50);
odd contigs are stored reverse-complemented so their records are '-' strand. A record is a window of
the target contig whose CIGAR is the edit script restricted to the window.
"""
import math

import numpy as np
import torch

MASTER_SEED = 20261104
ASCII = torch.tensor([65, 67, 71, 84], dtype=torch.uint8)  # A C G T


class Contig:
    pass


def _gen(seed, device):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return g


def make_contig(idx, length, s, g_rate, seed, device, keep_truth=True):
    """One target/query contig pair. Returns a Contig with ASCII tensors and the edit script."""
    gen = _gen(seed * 1000003 + idx, device)
    L = length
    t = torch.randint(0, 4, (L,), dtype=torch.uint8, device=device, generator=gen)
    sub = torch.rand(L, device=device, generator=gen) < s
    shift = torch.randint(1, 4, (L,), dtype=torch.uint8, device=device, generator=gen)
    qcode = torch.where(sub, (t + shift) & 3, t)
    del shift
    # indel events, spaced > 51 target bases apart so ops never touch
    cand = torch.nonzero(torch.rand(L, device=device, generator=gen) < g_rate).flatten()
    if cand.numel():
        ok = torch.ones_like(cand, dtype=torch.bool)
        ok[1:] = (cand[1:] - cand[:-1]) > 51
        ok &= (cand > 51) & (cand < L - 102)
        cand = cand[ok]
    E = cand.numel()
    u = torch.rand(E, device=device, generator=gen).clamp_(1e-12, 1.0)
    ev_len = (torch.log(u) / math.log(2.0 / 3.0)).floor().long().add_(1).clamp_(1, 50)  # Geometric(p=1/3), mean 3
    ev_del = torch.rand(E, device=device, generator=gen) < 0.5
    ev_pos = cand
    # deleted target bases
    delta = torch.zeros(L + 1, dtype=torch.int32, device=device)
    dpos, dlen = ev_pos[ev_del], ev_len[ev_del]
    delta.index_add_(0, dpos, torch.ones_like(dpos, dtype=torch.int32))
    delta.index_add_(0, dpos + dlen, -torch.ones_like(dpos, dtype=torch.int32))
    keep = torch.cumsum(delta[:L], 0) == 0
    del delta
    ins_at = torch.zeros(L, dtype=torch.int64, device=device)
    ipos, ilen = ev_pos[~ev_del], ev_len[~ev_del]
    ins_at.index_add_(0, ipos, ilen)
    qpos = torch.cumsum(keep.long(), 0) - keep.long() + torch.cumsum(ins_at, 0)
    del ins_at
    Lq = int(keep.sum().item() + ilen.sum().item())
    q = torch.randint(0, 4, (Lq,), dtype=torch.uint8, device=device, generator=gen)
    q[qpos[keep]] = qcode[keep]
    del qcode
    c = Contig()
    c.idx, c.L, c.Lq = idx, L, Lq
    c.rev = (idx % 2) == 1
    c.ev_pos, c.ev_len, c.ev_del = ev_pos, ev_len, ev_del
    # prefix sums over events for t -> q coordinate mapping
    c.cum_del = torch.cat([torch.zeros(1, dtype=torch.int64, device=device), torch.cumsum(torch.where(ev_del, ev_len, 0), 0)])
    c.cum_ins = torch.cat([torch.zeros(1, dtype=torch.int64, device=device), torch.cumsum(torch.where(ev_del, 0, ev_len), 0)])
    asc = ASCII.to(device)
    c.t_ascii = asc[t.long()]
    if c.rev:
        c.q_ascii = asc[(3 - q.flip(0)).long()]
    else:
        c.q_ascii = asc[q.long()]
    if keep_truth:
        subk = sub & keep
        c.cum_sub = torch.cat([torch.zeros(1, dtype=torch.int64, device=device), torch.cumsum(subk.long(), 0)])
        # flip[p]: aligned base p has a different match state than aligned base p-1 (both kept & adjacent)
        flip = torch.zeros(L, dtype=torch.bool, device=device)
        flip[1:] = (sub[1:] != sub[:-1]) & keep[1:] & keep[:-1]
        c.cum_flip = torch.cat([torch.zeros(1, dtype=torch.int64, device=device), torch.cumsum(flip.long(), 0)])
    c.name_t, c.name_q = f"t{idx + 1}", f"q{idx + 1}"
    return c


def contig_records(c, n, length_law, seed, device):
    """n records on contig c. Returns dict of per-record tensors + flat ops (uint64 words as int64)."""
    gen = _gen(seed * 7919 + 17 * c.idx + 5, device)
    L = c.L
    kind, a, b = length_law[:3]
    if kind == "normal":      # N(a, b^2) clipped
        ln = (torch.randn(n, device=device, generator=gen) * b + a).clamp_(length_law[3], length_law[4])
    elif kind == "uniform":
        ln = torch.rand(n, device=device, generator=gen) * (b - a) + a
    elif kind == "loguniform":
        ln = torch.exp(torch.rand(n, device=device, generator=gen) * (math.log(b) - math.log(a)) + math.log(a))
    else:
        raise ValueError(kind)
    ln = ln.long().clamp_(60, L - 200)
    ts = (torch.rand(n, device=device, generator=gen).double() * (L - 100 - ln).double()).long() + 50
    te = ts + ln
    E = c.ev_pos.numel()
    ev_end = c.ev_pos + torch.where(c.ev_del, c.ev_len, 0)  # first target base after the event
    # snap: ts / te-1 must not lie inside a deletion
    k = torch.searchsorted(c.ev_pos, ts, right=True) - 1  # last event with pos <= ts
    kk = k.clamp(min=0)
    inside = (k >= 0) & c.ev_del[kk] & (ts < ev_end[kk]) if E else torch.zeros_like(ts, dtype=torch.bool)
    ts = torch.where(inside, ev_end[kk], ts) if E else ts
    k = torch.searchsorted(c.ev_pos, te - 1, right=True) - 1
    kk = k.clamp(min=0)
    inside = (k >= 0) & c.ev_del[kk] & (te - 1 < ev_end[kk]) if E else torch.zeros_like(ts, dtype=torch.bool)
    te = torch.where(inside, c.ev_pos[kk], te) if E else te
    # events strictly inside the window: ts < pos < te
    k1 = torch.searchsorted(c.ev_pos, ts, right=True)
    k2 = torch.searchsorted(c.ev_pos, te, right=False)
    nev = (k2 - k1).clamp(min=0)
    nops = 2 * nev + 1
    op_begin = torch.cat([torch.zeros(1, dtype=torch.int64, device=device), torch.cumsum(nops, 0)])
    N = int(op_begin[-1].item())
    rec = torch.repeat_interleave(torch.arange(n, device=device), nops)
    local = torch.arange(N, device=device) - op_begin[rec]
    is_ev = (local & 1) == 1
    e_raw = k1[rec] + (local >> 1)
    e_idx = e_raw.clamp(max=max(E - 1, 0))   # event after this M (or the event itself)
    if E:
        prev_end = torch.where(local >> 1 > 0, ev_end[(e_raw - 1).clamp(min=0, max=E - 1)], ts[rec])
        nxt = torch.where((local >> 1) < nev[rec], c.ev_pos[e_idx], te[rec])
        m_len = nxt - prev_end
        length = torch.where(is_ev, c.ev_len[e_idx], m_len)
        code = torch.where(is_ev, torch.where(c.ev_del[e_idx], 2, 1), 0)
    else:
        length = te[rec] - ts[rec]
        code = torch.zeros_like(length)
    ops = length | (code << 56)
    # t -> q coordinates of kept bases
    def qmap(p):
        kd = torch.searchsorted(c.ev_pos, p, right=False)   # events with pos < p  (deletions fully before p)
        ki = torch.searchsorted(c.ev_pos, p, right=True)    # events with pos <= p (insertions before base p)
        return p - c.cum_del[kd] + c.cum_ins[ki]
    qs_ = qmap(ts)
    qe_ = qmap(te - 1) + 1
    if c.rev:
        qs, qe = c.Lq - qe_, c.Lq - qs_
    else:
        qs, qe = qs_, qe_
    out = dict(n=n, contig=torch.full((n,), c.idx, dtype=torch.int32, device=device), ts=ts, te=te, qs=qs, qe=qe,
               strand=torch.full((n,), 0 if c.rev else 1, dtype=torch.uint8, device=device), op_begin=op_begin, ops=ops,
               sum_m=None)
    is_m = ~is_ev
    sum_m = torch.zeros(n, dtype=torch.int64, device=device).index_add_(0, rec[is_m], length[is_m])
    out["sum_m"] = sum_m
    if hasattr(c, "cum_sub"):
        out["true_mism"] = c.cum_sub[te] - c.cum_sub[ts]
        # runs: per M op [a, b): 1 + flips in (a, b); flips never cross an event because keep breaks them
        m_a = torch.where(is_m, prev_end if E else ts[rec], 0)
        m_b = torch.where(is_m, nxt if E else te[rec], 0)
        runs = torch.where(is_m, 1 + c.cum_flip[m_b] - c.cum_flip[(m_a + 1).clamp(max=L)], 1)
        out["true_out_ops"] = torch.zeros(n, dtype=torch.int64, device=device).index_add_(0, rec, runs)
    return out


def concat_records(parts, shuffle_seed=None):
    """Concatenate per-contig record sets; optionally shuffle record order (re-flattening the ops)."""
    device = parts[0]["ops"].device
    keys = ["contig", "ts", "te", "qs", "qe", "strand", "sum_m"] + [k for k in ("true_mism", "true_out_ops") if k in parts[0]]
    rec = {k: torch.cat([p[k] for p in parts]) for k in keys}
    nops = torch.cat([p["op_begin"][1:] - p["op_begin"][:-1] for p in parts])
    ops = torch.cat([p["ops"] for p in parts])
    R = nops.numel()
    begin = torch.cat([torch.zeros(1, dtype=torch.int64, device=device), torch.cumsum(nops, 0)])
    if shuffle_seed is not None:
        perm = torch.randperm(R, device=device, generator=_gen(shuffle_seed, device))
        new_nops = nops[perm]
        new_begin = torch.cat([torch.zeros(1, dtype=torch.int64, device=device), torch.cumsum(new_nops, 0)])
        src_rec = torch.repeat_interleave(perm, new_nops)
        dst_rec = torch.repeat_interleave(torch.arange(R, device=device), new_nops)
        local = torch.arange(ops.numel(), device=device) - new_begin[dst_rec]
        ops = ops[begin[src_rec] + local]
        rec = {k: v[perm] for k, v in rec.items()}
        begin = new_begin
    rec["op_begin"], rec["ops"], rec["n"] = begin, ops, R
    return rec


CONFIGS = {
    # name: (n_contigs, contig_len, n_records, length law, s, g)
    "C2": (24, 125_000_000, 1_000_000, ("normal", 10_000, 2_000, 2_000, 20_000), 0.01, 0.001),
    "C3": (24, 125_000_000, 50_000, ("loguniform", 1e5, 5e6), 0.005, 0.0005),
    "C4": (24, 125_000_000, 10_000_000, ("uniform", 1_000, 5_000), 0.01, 0.001),
    # config 5 of BASELINE.json: 20 M records of C2's shape, ONE record list sharded over the GPUs of the box
    "C5": (24, 125_000_000, 20_000_000, ("normal", 10_000, 2_000, 2_000, 20_000), 0.01, 0.001),
    "tiny": (4, 200_000, 2_000, ("normal", 3_000, 1_000, 500, 8_000), 0.01, 0.001),
    "small": (4, 5_000_000, 40_000, ("normal", 10_000, 2_000, 2_000, 20_000), 0.01, 0.001),
}


def shard(rec, lo, hi):
    """Records [lo, hi) of a record set (the ops re-based), e.g. one rank's contiguous range of a job."""
    out = {}
    ob = rec["op_begin"]
    o0, o1 = int(ob[lo].item()), int(ob[hi].item())
    for k, v in rec.items():
        if k == "op_begin":
            out[k] = (ob[lo:hi + 1] - o0).contiguous()
        elif k == "ops":
            out[k] = v[o0:o1].contiguous()
        elif torch.is_tensor(v) and v.dim() == 1 and v.numel() == rec["n"]:
            out[k] = v[lo:hi].contiguous()
        else:
            out[k] = v
    out["n"] = hi - lo
    return out


def build_workload(ctx, name, device, seed=MASTER_SEED, scale=1.0, shuffle=True, keep_ascii=(), rank=0):
    """Generate config `name`, load the genome into ctx (device-resident ASCII -> packed on the GPU) and
    return the record set (torch tensors on `device`). keep_ascii: contig indices whose ASCII is kept
    (host bytes) for CPU-baseline FASTA files. `rank` offsets the record seed so ranks get disjoint shards
    (weak scaling); strong scaling generates the same list (rank 0's) everywhere and cuts it with shard()."""
    n_contigs, clen, n_rec, law, s, g = CONFIGS[name]
    n_rec = max(n_contigs, int(n_rec * scale))
    parts, kept = [], {}
    per = [n_rec // n_contigs + (1 if i < n_rec % n_contigs else 0) for i in range(n_contigs)]
    if ctx is not None:
        ctx.reserve(int(2.02 * n_contigs * clen) + 4096 * n_contigs)
    for i in range(n_contigs):
        c = make_contig(i, clen, s, g, seed, device)
        if ctx is not None:
            tid = ctx.add_contig(c.name_t, c.t_ascii)
            qid = ctx.add_contig(c.name_q, c.q_ascii)
        else:
            tid, qid = 2 * i, 2 * i + 1
        if i in keep_ascii:
            kept[i] = (c.name_t, c.t_ascii.cpu().numpy().tobytes(), c.name_q, c.q_ascii.cpu().numpy().tobytes(), c.L, c.Lq)
        p = contig_records(c, per[i], law, seed + 977 * rank, device)
        p["t_contig"] = torch.full((per[i],), tid, dtype=torch.int32, device=device)
        p["q_contig"] = torch.full((per[i],), qid, dtype=torch.int32, device=device)
        p["q_len"] = torch.full((per[i],), c.Lq, dtype=torch.int64, device=device)
        parts.append(p)
        del c
    extra = ["t_contig", "q_contig", "q_len"]
    rec = concat_records([{**p} for p in parts], None)
    for k in extra:
        rec[k] = torch.cat([p[k] for p in parts])
    if shuffle:
        # shuffle with the extra columns carried along
        device_ = rec["ops"].device
        R = rec["n"]
        perm = torch.randperm(R, device=device_, generator=_gen(seed + 31 + rank, device_))
        nops = rec["op_begin"][1:] - rec["op_begin"][:-1]
        new_nops = nops[perm]
        new_begin = torch.cat([torch.zeros(1, dtype=torch.int64, device=device_), torch.cumsum(new_nops, 0)])
        src_rec = torch.repeat_interleave(perm, new_nops)
        dst_rec = torch.repeat_interleave(torch.arange(R, device=device_), new_nops)
        local = torch.arange(rec["ops"].numel(), device=device_) - new_begin[dst_rec]
        rec["ops"] = rec["ops"][rec["op_begin"][src_rec] + local]
        for k in list(rec.keys()):
            if k not in ("ops", "op_begin", "n") and torch.is_tensor(rec[k]):
                rec[k] = rec[k][perm]
        rec["op_begin"] = new_begin
    rec["clen"] = clen
    rec["kept_ascii"] = kept
    return rec


def paf_lines(rec, idxs, names=None):
    """Render records `idxs` as PAF text lines (host loop; small samples only)."""
    ob = rec["op_begin"].cpu().numpy()
    ops = rec["ops"].cpu().numpy()
    cols = {k: rec[k].cpu().numpy() for k in ("contig", "ts", "te", "qs", "qe", "strand", "sum_m", "q_len")}
    out = []
    for i in idxs:
        ci = int(cols["contig"][i])
        w = ops[ob[i]:ob[i + 1]]
        cg = "".join(f"{int(x) & ((1 << 56) - 1)}{'MID=X'[int(x) >> 56]}" for x in w)
        tot = int(sum(int(x) & ((1 << 56) - 1) for x in w))
        out.append(f"q{ci + 1}\t{int(cols['q_len'][i])}\t{int(cols['qs'][i])}\t{int(cols['qe'][i])}\t"
                   f"{'+' if cols['strand'][i] else '-'}\tt{ci + 1}\t{rec['clen']}\t{int(cols['ts'][i])}\t{int(cols['te'][i])}\t"
                   f"{int(cols['sum_m'][i])}\t{tot}\t60\ttp:A:P\tAS:i:{int(cols['sum_m'][i])}\tcg:Z:{cg}")
    return out
