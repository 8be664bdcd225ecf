"""ctypes binding of libpaffy_b200.so (the C ABI declared in include/paffy_b200.h).

Plumbing only: every computation happens in the CUDA library. There is no CPU fallback; creating a
Context without a CUDA device raises PafGpuError.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PAFFY_B200_LIB") or os.path.join(HERE, "libpaffy_b200.so")  # (the override: kernel A/B runs with prebuilt variants)

# symbols include/paffy_b200.h declares (tests check that the library exports every one)
ABI_SYMBOLS = [
    "pafgpu_abi_version", "pafgpu_device_count", "pafgpu_create", "pafgpu_destroy", "pafgpu_last_error",
    "pafgpu_genome_reserve", "pafgpu_genome_add_contig", "pafgpu_genome_add_contig_device", "pafgpu_genome_load_fasta",
    "pafgpu_genome_find", "pafgpu_genome_contig_length", "pafgpu_genome_num_contigs", "pafgpu_genome_fetch",
    "pafgpu_encode_batch_host", "pafgpu_encode_batch_device", "pafgpu_fetch_ops", "pafgpu_result_device_ptrs",
    "pafgpu_serialise", "pafgpu_fetch_text", "pafgpu_encode_one", "pafgpu_time_encode",
    "pafgpu_last_bad_record", "pafgpu_host_alloc", "pafgpu_host_free", "pafgpu_add_mismatches_file", "pafgpu_partition", "pafgpu_trim",
    "pafgpu_trim_fixed", "pafgpu_stage_ops_host", "pafgpu_add_mismatches_file_filtered",
    "pafgpu_add_mismatches_text", "pafgpu_add_mismatches_text_opts", "pafgpu_create_sibling", "pafgpu_helper", "pafgpu_encode_balance", "pafgpu_encode_one_resident", "pafgpu_genome_clear", "pafgpu_genome_total_bases", "pafgpu_device_of", "pafgpu_genome_clone", "pafgpu_genome_save", "pafgpu_genome_load_packed",
    "pafgpu_serialise_no_alloc",
]
# reference-named entry points the library also exports (include/paf_b200_compat.h, include/paffy_b200.h)
REFERENCE_SYMBOLS = ["paffy_add_mismatches_main", "paffy_view_main", "paffy_trim_main", "paf_encode_mismatches", "pafgpu_paf_encode_mismatches"]


class PafGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"paffy_b200 error {code}: {msg}")
        self.code = code


class Batch(C.Structure):
    _fields_ = [("n_records", C.c_int64), ("n_ops", C.c_int64), ("op_begin", C.c_void_p), ("ops", C.c_void_p),
                ("q_contig", C.c_void_p), ("t_contig", C.c_void_p), ("q_start", C.c_void_p), ("q_end", C.c_void_p),
                ("t_start", C.c_void_p), ("same_strand", C.c_void_p)]


class FileOpts(C.Structure):
    _fields_ = [("mode", C.c_int), ("trim_score_fraction", C.c_float), ("trim_max_fraction", C.c_float), ("threads", C.c_int),
                ("batch_bases", C.c_int64), ("gpus", C.c_int), ("streams", C.c_int), ("chunk_bytes", C.c_int64)]


class FileFilter(C.Structure):
    """pafgpu_filter: the options of `paffy filter` (impl/paf_filter.c:26-33) with the reference's defaults."""
    _fields_ = [("min_chain_score", C.c_int64), ("min_alignment_score", C.c_int64), ("min_identity", C.c_double),
                ("min_identity_with_gaps", C.c_double), ("max_tile_level", C.c_int64), ("invert", C.c_int)]


class FileReport(C.Structure):
    _fields_ = [("records", C.c_int64), ("aligned_bases", C.c_int64), ("batches", C.c_int64), ("failed_record", C.c_int64),
                ("stats", C.c_int64 * 7), ("s_read", C.c_double), ("s_parse", C.c_double), ("s_gpu", C.c_double),
                ("s_host", C.c_double), ("s_write", C.c_double), ("s_total", C.c_double), ("chunks", C.c_int64), ("gpus", C.c_int64),
                ("workers", C.c_int64)]


class TextResult(C.Structure):
    _fields_ = [("records", C.c_int64), ("records_in", C.c_int64), ("n_ops", C.c_int64), ("n_out_ops", C.c_int64),
                ("aligned_bases", C.c_int64), ("text_bytes", C.c_int64), ("first_bad", C.c_int64), ("bad_line_begin", C.c_int64),
                ("bad_line_end", C.c_int64), ("bad_kind", C.c_int64), ("bad_coords", C.c_int64 * 4), ("stats", C.c_int64 * 7), ("ms_parse", C.c_float), ("ms_ingest", C.c_float),
                ("ms_kernel", C.c_float), ("ms_encode", C.c_float), ("ms_render", C.c_float)]


class Result(C.Structure):
    _fields_ = [("n_out_ops", C.c_int64), ("aligned_bases", C.c_int64), ("n_windows", C.c_int64), ("n_tiles", C.c_int64),
                ("retried", C.c_int64), ("ms_ingest", C.c_float), ("ms_encode", C.c_float), ("ms_kernel", C.c_float),
                ("ms_emit", C.c_float)]


_lib = None


def load():
    """Load (building if stale and nvcc is present) the CUDA library. Fails loudly when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.environ.get("PAFFY_B200_LIB") and (not os.path.exists(LIB_PATH) or (_build.stale() and os.path.exists("/usr/local/cuda/bin/nvcc"))):
        _build.build()
    if not os.path.exists(LIB_PATH):
        raise PafGpuError(-1, f"{LIB_PATH} is missing: run `python -m paffy_b200.build` (no CPU fallback exists)")
    L = C.CDLL(LIB_PATH)
    vp, i64, i64p = C.c_void_p, C.c_int64, C.POINTER(C.c_int64)
    L.pafgpu_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.pafgpu_destroy.argtypes = [vp]
    L.pafgpu_destroy.restype = None
    L.pafgpu_last_error.argtypes = [vp]
    L.pafgpu_last_error.restype = C.c_char_p
    L.pafgpu_genome_reserve.argtypes = [vp, i64]
    L.pafgpu_genome_add_contig.argtypes = [vp, C.c_char_p, vp, i64]
    L.pafgpu_genome_add_contig.restype = i64
    L.pafgpu_genome_add_contig_device.argtypes = [vp, C.c_char_p, vp, i64]
    L.pafgpu_genome_add_contig_device.restype = i64
    L.pafgpu_genome_load_fasta.argtypes = [vp, C.c_char_p]
    L.pafgpu_genome_find.argtypes = [vp, C.c_char_p]
    L.pafgpu_genome_find.restype = i64
    L.pafgpu_genome_contig_length.argtypes = [vp, i64]
    L.pafgpu_genome_contig_length.restype = i64
    L.pafgpu_genome_num_contigs.argtypes = [vp]
    L.pafgpu_genome_num_contigs.restype = i64
    L.pafgpu_genome_fetch.argtypes = [vp, i64, i64, i64, vp]
    L.pafgpu_encode_batch_host.argtypes = [vp, C.POINTER(Batch), C.POINTER(Result)]
    L.pafgpu_encode_batch_device.argtypes = [vp, C.POINTER(Batch), C.POINTER(Result)]
    L.pafgpu_fetch_ops.argtypes = [vp, vp, vp, vp]
    L.pafgpu_result_device_ptrs.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    L.pafgpu_serialise.argtypes = [vp, vp, vp, i64p, C.POINTER(C.c_float)]
    L.pafgpu_fetch_text.argtypes = [vp, vp, vp]
    L.pafgpu_encode_one.argtypes = [vp, vp, i64, C.c_int, i64, i64, i64, vp, vp, C.POINTER(vp), i64p, i64p]
    L.pafgpu_time_encode.argtypes = [vp, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L.pafgpu_last_bad_record.argtypes = [vp]
    L.pafgpu_last_bad_record.restype = i64
    L.pafgpu_host_alloc.argtypes = [i64]
    L.pafgpu_host_alloc.restype = vp
    L.pafgpu_host_free.argtypes = [vp]
    L.pafgpu_host_free.restype = None
    L.pafgpu_add_mismatches_file.argtypes = [vp, C.c_char_p, C.c_char_p, C.POINTER(FileOpts), C.POINTER(FileReport)]
    L.pafgpu_add_mismatches_file_filtered.argtypes = [vp, C.c_char_p, C.c_char_p, C.POINTER(FileOpts), C.POINTER(FileFilter), C.POINTER(FileReport)]
    L.pafgpu_partition.argtypes = [vp, i64, C.c_int, vp]
    L.pafgpu_create_sibling.argtypes = [vp, C.POINTER(vp)]
    L.pafgpu_add_mismatches_text.argtypes = [vp, vp, i64, C.POINTER(TextResult)]
    L.pafgpu_genome_clone.argtypes = [vp, vp]
    L.pafgpu_genome_save.argtypes = [vp, C.c_char_p, C.c_char_p]
    L.pafgpu_genome_load_packed.argtypes = [vp, C.c_char_p, C.c_char_p]
    L.pafgpu_stage_ops_host.argtypes = [vp, C.POINTER(Batch)]
    L.pafgpu_trim.argtypes = [vp, C.c_float, C.c_float, vp, vp, vp, vp]
    L.pafgpu_serialise_no_alloc.argtypes = [C.c_int, vp, vp, i64, vp, i64, vp, vp, vp, i64, vp, C.POINTER(i64)]
    _lib = L
    return L


i64_t = C.c_int64
NO_ALLOC_MAX_OPS = 1 << 26


def serialise_no_alloc(device, ops_dev, n_ops, op_begin_dev, n_records, prefix_dev, prefix_begin_dev, text_dev, text_cap, line_begin_dev,
                       stream=None, wait=True):
    """pafgpu_serialise_no_alloc: raw device pointers (ints) in, text written into the caller's device buffer; allocates nothing.
    -> bytes of the whole text (wait=True) or None (wait=False: only asynchronous work on `stream` was queued)."""
    L = load()
    total = i64_t(0)
    rc = L.pafgpu_serialise_no_alloc(device, stream, ops_dev, n_ops, op_begin_dev, n_records, prefix_dev, prefix_begin_dev, text_dev, text_cap,
                                     line_begin_dev, C.byref(total) if wait else None)
    if rc:
        raise PafGpuError(rc, (L.pafgpu_last_error(None) or b"").decode())
    return total.value if wait else None


def partition(weights, parts):
    """Contiguous near-equal-weight cuts of input-ordered records (multi-GPU sharding, no device needed)."""
    w = np.ascontiguousarray(weights, dtype=np.int64)
    cuts = np.zeros(parts + 1, dtype=np.int64)
    rc = load().pafgpu_partition(w.ctypes.data, w.size, parts, cuts.ctypes.data)
    if rc:
        raise PafGpuError(rc, "pafgpu_partition: bad arguments")
    return cuts


def remove_mismatches_file(in_path, out_path, threads=0, chunk_bytes=0):
    """`paffy add_mismatches -a` (host-only text work: no CUDA device needed). -> (rc, FileReport)"""
    o = FileOpts(1, 0.05, 1.0, threads, 0, 0, 0, chunk_bytes)
    rep = FileReport()
    rc = load().pafgpu_add_mismatches_file(None, in_path.encode() if in_path else None, out_path.encode() if out_path else None, C.byref(o), C.byref(rep))
    return rc, rep


def _ptr(a):
    """Host numpy array or torch tensor (any device) -> raw pointer."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()  # torch tensor


def _is_device(a):
    return not isinstance(a, np.ndarray) and getattr(a, "is_cuda", False)


class Context:
    """One GPU context: a packed genome store plus the batch encode / serialise pipeline."""

    def __init__(self, device=0):
        self.L = load()
        h = C.c_void_p()
        rc = self.L.pafgpu_create(device, C.byref(h))
        if rc:
            raise PafGpuError(rc, self.L.pafgpu_last_error(None).decode())
        self.h = h
        self.n_records = 0
        self.result = None

    def close(self):
        if getattr(self, "h", None):
            self.L.pafgpu_destroy(self.h)
            self.h = None

    def sibling(self):
        """A second context on the same GPU sharing this one's packed genome (own stream and buffers): copies and kernels
        of different batches overlap without a second genome copy."""
        s = Context.__new__(Context)
        s.L = self.L
        h = C.c_void_p()
        rc = self.L.pafgpu_create_sibling(self.h, C.byref(h))
        if rc:
            raise PafGpuError(rc, self.L.pafgpu_last_error(None).decode())
        s.h, s.n_records, s.result = h, 0, None
        return s

    def add_mismatches_text_resident(self, buf):
        """The same with the input in a caller-owned (pinned) uint8 numpy array and the output left on the device: fetch it
        with fetch_text(pinned=...). -> TextResult"""
        res = TextResult()
        self._check(self.L.pafgpu_add_mismatches_text(self.h, buf.ctypes.data, buf.size, C.byref(res)))
        self.text_bytes = res.text_bytes
        self.n_records = res.records
        return res

    def add_mismatches_text(self, text):
        """PAF text (bytes, free of \r) in, PAF text out, parsed / encoded / rendered on the device. -> (TextResult, bytes)"""
        res = TextResult()
        buf = np.frombuffer(text, dtype=np.uint8)
        self._check(self.L.pafgpu_add_mismatches_text(self.h, buf.ctypes.data, buf.size, C.byref(res)))
        out = np.empty(res.text_bytes, dtype=np.uint8)
        if res.text_bytes:
            self._check(self.L.pafgpu_fetch_text(self.h, out.ctypes.data, None))
        return res, out.tobytes()

    __del__ = close

    def _check(self, rc):
        if rc < 0:
            raise PafGpuError(rc, self.L.pafgpu_last_error(self.h).decode())
        return rc

    # ---- genome store
    def reserve(self, total_bases):
        self._check(self.L.pafgpu_genome_reserve(self.h, total_bases))

    def add_contig(self, name, seq):
        """seq: bytes / numpy uint8 (host) or a torch cuda uint8 tensor (device resident)."""
        if _is_device(seq):
            assert seq.is_contiguous()
            return self._check(self.L.pafgpu_genome_add_contig_device(self.h, name.encode(), seq.data_ptr(), seq.numel()))
        if isinstance(seq, (bytes, bytearray)):
            buf = (C.c_char * len(seq)).from_buffer_copy(seq) if len(seq) else None
            return self._check(self.L.pafgpu_genome_add_contig(self.h, name.encode(), C.cast(buf, C.c_void_p), len(seq)))
        seq = np.ascontiguousarray(seq, dtype=np.uint8)
        return self._check(self.L.pafgpu_genome_add_contig(self.h, name.encode(), seq.ctypes.data, seq.size))

    def load_fasta(self, path):
        self._check(self.L.pafgpu_genome_load_fasta(self.h, path.encode()))

    def find(self, name):
        return self.L.pafgpu_genome_find(self.h, name.encode())

    def contig_length(self, cid):
        return self.L.pafgpu_genome_contig_length(self.h, cid)

    def num_contigs(self):
        return self.L.pafgpu_genome_num_contigs(self.h)

    def fetch_bases(self, cid, start, length):
        out = np.empty(length, dtype=np.uint8)
        self._check(self.L.pafgpu_genome_fetch(self.h, cid, start, length, out.ctypes.data))
        return out.tobytes()

    # ---- batch encode
    def encode_batch(self, op_begin, ops, q_contig, t_contig, q_start, q_end, t_start, same_strand):
        """All eight arrays either host numpy (int64/uint64/int32/uint8) or torch cuda tensors."""
        arrs = [op_begin, ops, q_contig, t_contig, q_start, q_end, t_start, same_strand]
        dev = _is_device(ops)
        if not dev:
            want = [np.int64, np.uint64, np.int32, np.int32, np.int64, np.int64, np.int64, np.uint8]
            arrs = [a if not isinstance(a, np.ndarray) else np.ascontiguousarray(a, dtype=t) for a, t in zip(arrs, want)]
        n_records = int(arrs[0].shape[0]) - 1
        n_ops = int(arrs[1].shape[0])
        b = Batch(n_records, n_ops, *[_ptr(a) for a in arrs])
        res = Result()
        fn = self.L.pafgpu_encode_batch_device if dev else self.L.pafgpu_encode_batch_host
        self._check(fn(self.h, C.byref(b), C.byref(res)))
        self._keep = arrs  # device inputs are referenced by pafgpu_time_encode
        self.n_records = n_records
        self.result = res
        return res

    def stage_ops(self, op_begin, ops, q_start, q_end, t_start, same_strand):
        """Upload ops as if an encode had produced them (pafgpu_stage_ops_host): serialise / trim then work on them."""
        want = [np.int64, np.uint64, np.int64, np.int64, np.int64, np.uint8]
        a = [np.ascontiguousarray(x, dtype=t) for x, t in zip([op_begin, ops, q_start, q_end, t_start, same_strand], want)]
        n_records = int(a[0].shape[0]) - 1
        b = Batch(n_records, int(a[1].shape[0]), _ptr(a[0]), _ptr(a[1]), None, None, _ptr(a[2]), _ptr(a[3]), _ptr(a[4]), _ptr(a[5]))
        self._check(self.L.pafgpu_stage_ops_host(self.h, C.byref(b)))
        self.n_records = n_records

    def fetch_ops(self, want_ops=True, pinned=None):
        """-> (out_ops uint64[n_out], out_op_begin int64[R+1], rec_mismatch int64[R])"""
        n_out, R = self.result.n_out_ops, self.n_records
        ops = np.empty(n_out, dtype=np.uint64) if want_ops and pinned is None else pinned
        begin = np.empty(R + 1, dtype=np.int64)
        mism = np.empty(R, dtype=np.int64)
        self._check(self.L.pafgpu_fetch_ops(self.h, _ptr(ops) if want_ops else None, begin.ctypes.data, mism.ctypes.data))
        return ops, begin, mism

    def result_device_ptrs(self):
        """-> raw device pointers (output ops uint64[n_out], out_op_begin int64[R+1], rec_mismatch int64[R]) of the last batch."""
        a, b, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        self._check(self.L.pafgpu_result_device_ptrs(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def encode_balance(self):
        """-> (busiest SM's busy ms, mean busy ms over the SMs used, SMs used) of the last encode_kernel launch."""
        mx, mean, n = C.c_double(), C.c_double(), C.c_int()
        self._check(self.L.pafgpu_encode_balance(self.h, C.byref(mx), C.byref(mean), C.byref(n)))
        return mx.value, mean.value, n.value

    def time_encode(self, reps):
        stage, kernel = C.c_float(), C.c_float()
        self._check(self.L.pafgpu_time_encode(self.h, reps, C.byref(stage), C.byref(kernel)))
        return stage.value, kernel.value

    def serialise(self, prefix_bytes, prefix_begin):
        if isinstance(prefix_begin, np.ndarray):
            prefix_begin = np.ascontiguousarray(prefix_begin, dtype=np.int64)
        pb = np.frombuffer(prefix_bytes, dtype=np.uint8) if isinstance(prefix_bytes, (bytes, bytearray)) else prefix_bytes
        n, ms = C.c_int64(), C.c_float()
        self._check(self.L.pafgpu_serialise(self.h, _ptr(pb), _ptr(prefix_begin), C.byref(n), C.byref(ms)))
        self.text_bytes = n.value
        return n.value, ms.value

    def fetch_text(self, pinned=None, want_lines=True):
        """-> (text uint8[text_bytes], line offsets int64[R+1] or None). pinned: caller's (pinned) destination buffer."""
        buf = np.empty(self.text_bytes, dtype=np.uint8) if pinned is None else pinned
        lines = np.empty(self.n_records + 1, dtype=np.int64) if want_lines else None
        self._check(self.L.pafgpu_fetch_text(self.h, _ptr(buf), _ptr(lines)))
        return buf, lines

    def trim(self, t_end, score_fraction=0.05, max_fraction=1.0):
        """paf_trim_unreliable_tails on the encoded batch (device). -> (coords int64[4, R] = qs, qe, ts, te; kept ops; failed)"""
        R = self.n_records
        t_end = np.ascontiguousarray(t_end, dtype=np.int64)
        coords = np.empty((4, R), dtype=np.int64)
        kept = np.empty(R, dtype=np.int64)
        failed = np.empty(R, dtype=np.uint8)
        self._check(self.L.pafgpu_trim(self.h, score_fraction, max_fraction, t_end.ctypes.data, coords.ctypes.data, kept.ctypes.data,
                                       failed.ctypes.data))
        return coords, kept, failed

    def add_mismatches_file(self, in_path, out_path, mode=0, trim_r=0.05, trim_t=1.0, threads=0, batch_bases=0, gpus=0, streams=0, chunk_bytes=0):
        """The whole-file filter (paffy add_mismatches / -a / | trim; mode 4 | 1 | 2 = paffy view with per-alignment lines
        and / or the aggregate line). -> (rc, FileReport); rc 1 = a record ended the run the way the reference exits 1
        (message in self.last_error())."""
        o = FileOpts(mode, trim_r, trim_t, threads, batch_bases, gpus, streams, chunk_bytes)
        rep = FileReport()
        rc = self.L.pafgpu_add_mismatches_file(self.h, in_path.encode(), out_path.encode(), C.byref(o), C.byref(rep))
        if rc < 0:
            self._check(rc)
        return rc, rep

    def add_mismatches_filter_file(self, in_path, out_path, min_chain_score=-1, min_alignment_score=-1, min_identity=-1.0,
                                   min_identity_with_gaps=-1.0, max_tile_level=-1, invert=False, batch_bases=0, gpus=0, streams=0, chunk_bytes=0):
        """`paffy add_mismatches | paffy filter -s -t -u -v -w [-x]` in one pass -> (rc, FileReport)."""
        o = FileOpts(0, 0.05, 1.0, 0, batch_bases, gpus, streams, chunk_bytes)
        f = FileFilter(min_chain_score, min_alignment_score, min_identity, min_identity_with_gaps, max_tile_level, int(invert))
        rep = FileReport()
        rc = self.L.pafgpu_add_mismatches_file_filtered(self.h, in_path.encode(), out_path.encode(), C.byref(o), C.byref(f), C.byref(rep))
        if rc < 0:
            self._check(rc)
        return rc, rep

    def last_error(self):
        return self.L.pafgpu_last_error(self.h).decode()

    def encode_one(self, ops, same_strand, q_start, q_end, t_start, query_seq, target_seq):
        """paf_encode_mismatches on one record with whole-contig host strings -> (uint64 ops, mismatches)."""
        ops = np.ascontiguousarray(ops, dtype=np.uint64)
        out, n, mm = C.c_void_p(), C.c_int64(), C.c_int64()
        q = C.create_string_buffer(query_seq, len(query_seq) + 1)
        t = C.create_string_buffer(target_seq, len(target_seq) + 1)
        self._check(self.L.pafgpu_encode_one(self.h, ops.ctypes.data, ops.size, int(same_strand), q_start, q_end, t_start,
                                             C.cast(q, C.c_void_p), C.cast(t, C.c_void_p), C.byref(out), C.byref(n), C.byref(mm)))
        res = np.ctypeslib.as_array(C.cast(out, C.POINTER(C.c_uint64)), shape=(max(n.value, 1),))[: n.value].copy()
        C.CDLL(None).free(out)
        return res, mm.value
