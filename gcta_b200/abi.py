"""ctypes binding of include/gcta_b200.h (harness: tests/ and bench.py call the product through it).

There is no CPU fallback: if libgcta_b200.so is missing the import of ``lib()`` raises, and every
compute entry point returns GCTA_B200_ERR_NO_DEVICE without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgcta_b200.so")
FP16, BF16, FP16X2 = 0, 1, 2
ERRCAP = 512
TILE_M, TILE_N = 256, 256
IPC_HANDLE_BYTES = 64   # GCTA_B200_IPC_HANDLE_BYTES

EXPORTS = [
    "gcta_b200_make_grm", "gcta_b200_make_grm_multi", "gcta_b200_plan_set_chunk_hook", "gcta_b200_plan_own_bed", "gcta_b200_plan_create", "gcta_b200_plan_upload", "gcta_b200_plan_synth", "gcta_b200_plan_download_bed",
    "gcta_b200_plan_run", "gcta_b200_plan_run_streamed", "gcta_b200_plan_download", "gcta_b200_plan_device_results", "gcta_b200_plan_destroy",
    "gcta_b200_bed_decode", "gcta_b200_snp_stats", "gcta_b200_compact_rows", "gcta_b200_decode_emit_diag", "gcta_b200_diag_accum", "gcta_b200_diag_accum_static", "gcta_b200_nonmiss_planes", "gcta_b200_syrk_tiles", "gcta_b200_syrk_tiles_chained",
    "gcta_b200_paircount_tiles", "gcta_b200_finalize_tiles", "gcta_b200_version", "gcta_b200_device_count",
    "gcta_b200_partition", "gcta_b200_tile_list", "gcta_b200_allele_counts", "gcta_b200_nonmiss_matrix", "gcta_b200_count_tiles",
    "gcta_b200_plan_use_device_bed", "gcta_b200_plan_upload_rows", "gcta_b200_plan_run_to_host",
    "gcta_b200_plan_num_chunks", "gcta_b200_plan_chunk_rows", "gcta_b200_plan_upload_rows_async", "gcta_b200_plan_set_chunk_events",
    "gcta_b200_plan_use_own_bed", "gcta_b200_plan_geometry",
    "gcta_b200_exchange_slice", "gcta_b200_exchange_blob_bytes", "gcta_b200_exchange_create", "gcta_b200_exchange_connect", "gcta_b200_exchange_post",
    "gcta_b200_exchange_pull", "gcta_b200_exchange_finish", "gcta_b200_exchange_disconnect", "gcta_b200_exchange_destroy",
]


class Args(C.Structure):
    _fields_ = [
        ("bed", C.c_void_p), ("n_file", C.c_int64), ("m_file", C.c_int64),
        ("keep", C.c_void_p), ("n", C.c_int64), ("include", C.c_void_p), ("m", C.c_int64),
        ("mu", C.c_void_p), ("sex", C.c_void_p),
        ("xchr", C.c_int32), ("alg", C.c_int32), ("operand", C.c_int32), ("device", C.c_int32),
        ("part_rank", C.c_int32), ("part_world", C.c_int32), ("chunk_snps", C.c_int64),
    ]


class Out(C.Structure):
    _fields_ = [
        ("A_lower", C.c_void_p), ("N_lower", C.c_void_p), ("mu_out", C.c_void_p),
        ("row_begin", C.c_int64), ("row_end", C.c_int64), ("total_missing", C.c_int64),
        ("ms_h2d", C.c_float), ("ms_decode", C.c_float), ("ms_syrk", C.c_float), ("ms_paircount", C.c_float),
        ("ms_finalize", C.c_float), ("ms_d2h", C.c_float), ("ms_total", C.c_float), ("launches", C.c_int32),
    ]


class GctaB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is not built (python -m gcta_b200.build); gcta_b200 has no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.gcta_b200_version.restype = C.c_char_p
        L.gcta_b200_tile_list.restype = C.c_int64
        L.gcta_b200_tile_list.argtypes = [C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
        L.gcta_b200_partition.argtypes = [C.c_int64, C.c_int32, C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.gcta_b200_plan_destroy.restype = None
        L.gcta_b200_plan_destroy.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def _check(rc, err):
    if rc != 0:
        raise GctaB200Error(rc, err.value.decode(errors="replace"))


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return int(a)  # raw address (torch data_ptr())


def partition(n, rank, world):
    b, e = C.c_int64(), C.c_int64()
    rc = lib().gcta_b200_partition(n, rank, world, C.byref(b), C.byref(e))
    if rc:
        raise GctaB200Error(rc, "bad partition arguments")
    return b.value, e.value


def tile_list(n, row_begin=0, row_end=None) -> np.ndarray:
    row_end = n if row_end is None else row_end
    cnt = lib().gcta_b200_tile_list(n, row_begin, row_end, None)
    if cnt < 0:
        raise GctaB200Error(1, "bad tile_list arguments")
    t = np.zeros((cnt, 2), dtype=np.int32)
    lib().gcta_b200_tile_list(n, row_begin, row_end, t.ctypes.data)
    return t


def make_args(bed, n_file, m_file, keep=None, include=None, mu=None, sex=None, xchr=0, alg=0, operand=FP16,
              device=0, part_rank=0, part_world=1, chunk_snps=0):
    """Returns (Args, keepalive list).  bed may be a numpy array or a raw host address."""
    ka = []

    def arr(a, dt):
        if a is None:
            return None
        a = np.ascontiguousarray(a, dtype=dt)
        ka.append(a)
        return a.ctypes.data

    if isinstance(bed, np.ndarray):
        bed = np.ascontiguousarray(bed, dtype=np.uint8)
        ka.append(bed)
    a = Args()
    a.bed = _ptr(bed)
    a.n_file, a.m_file = n_file, m_file
    a.keep = arr(keep, np.int32)
    a.n = n_file if keep is None else len(keep)
    a.include = arr(include, np.int32)
    a.m = m_file if include is None else len(include)
    a.mu = arr(mu, np.float64)
    a.sex = arr(sex, np.int32)
    a.xchr, a.alg, a.operand, a.device = xchr, alg, operand, device
    a.part_rank, a.part_world, a.chunk_snps = part_rank, part_world, chunk_snps
    return a, ka


def make_grm(bed, n_file, m_file, want_N=True, **kw):
    """Whole path through gcta_b200_make_grm with host numpy buffers.
    -> (A_lower f32, N_lower i32 or None, mu f64, Out)"""
    a, ka = make_args(bed, n_file, m_file, **kw)
    n, m = a.n, a.m
    A = np.zeros(n * (n + 1) // 2, dtype=np.float32)
    N = np.zeros(n * (n + 1) // 2, dtype=np.int32) if want_N else None
    mu = np.zeros(m, dtype=np.float64)
    o = Out()
    o.A_lower, o.N_lower, o.mu_out = A.ctypes.data, _ptr(N), mu.ctypes.data
    err = C.create_string_buffer(ERRCAP)
    _check(lib().gcta_b200_make_grm(C.byref(a), C.byref(o), err, ERRCAP), err)
    return A, N, mu, o


def make_grm_multi(bed, n_file, m_file, n_gpus, want_N=True, **kw):
    """gcta_b200_make_grm_multi: the whole path on n_gpus devices of this process.
    -> (A_lower, N_lower or None, mu, Out, [Out per device])"""
    a, ka = make_args(bed, n_file, m_file, **kw)
    n, m = a.n, a.m
    A = np.zeros(n * (n + 1) // 2, dtype=np.float32)
    N = np.zeros(n * (n + 1) // 2, dtype=np.int32) if want_N else None
    mu = np.zeros(m, dtype=np.float64)
    o = Out()
    o.A_lower, o.N_lower, o.mu_out = A.ctypes.data, _ptr(N), mu.ctypes.data
    per = (Out * n_gpus)()
    err = C.create_string_buffer(ERRCAP)
    _check(lib().gcta_b200_make_grm_multi(C.byref(a), C.c_int32(n_gpus), C.byref(o), per, err, ERRCAP), err)
    return A, N, mu, o, list(per)


class Plan:
    """RAII wrapper of the plan API."""

    def __init__(self, n_file, m_file, **kw):
        self.args, self._ka = make_args(None, n_file, m_file, **kw)
        self.n, self.m = self.args.n, self.args.m
        self._h = C.c_void_p()
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_plan_create(C.byref(self.args), C.byref(self._h), err, ERRCAP), err)
        self.out = Out()

    def upload(self, bed_host):
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_plan_upload(self._h, C.c_void_p(_ptr(bed_host)), err, ERRCAP), err)

    def synth(self, seed, thr):
        thr = np.ascontiguousarray(thr, dtype=np.uint32)
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_plan_synth(self._h, C.c_uint64(seed), C.c_void_p(thr.ctypes.data), err, ERRCAP), err)

    def use_device_bed(self, d_bed=None) -> int:
        """make the plan read genotypes from caller-owned device memory; returns the row pitch in bytes"""
        pitch = C.c_int64()
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_plan_use_device_bed(self._h, C.c_void_p(_ptr(d_bed)), C.byref(pitch), err, ERRCAP), err)
        return pitch.value

    def upload_rows(self, bed_host, row_begin, row_end):
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_plan_upload_rows(self._h, C.c_void_p(_ptr(bed_host)), C.c_int64(row_begin), C.c_int64(row_end), err, ERRCAP), err)

    def chunk_rows(self):
        """[(row_begin, row_end)] of the file rows each SNP chunk reads"""
        L = lib()
        L.gcta_b200_plan_num_chunks.restype = C.c_int64
        out = []
        for c in range(L.gcta_b200_plan_num_chunks(self._h)):
            b, e = C.c_int64(), C.c_int64()
            if L.gcta_b200_plan_chunk_rows(self._h, C.c_int64(c), C.byref(b), C.byref(e)):
                raise GctaB200Error(1, "bad chunk index")
            out.append((b.value, e.value))
        return out

    def upload_rows_async(self, bed_host, row_begin, row_end, stream):
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_plan_upload_rows_async(self._h, C.c_void_p(_ptr(bed_host)), C.c_int64(row_begin), C.c_int64(row_end),
                                                      C.c_void_p(stream), err, ERRCAP), err)

    def set_chunk_events(self, events):
        """events: list of raw cudaEvent_t handles (torch.cuda.Event.cuda_event), one per chunk; [] clears"""
        arr = (C.c_void_p * max(1, len(events)))(*events)
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_plan_set_chunk_events(self._h, arr, C.c_int64(len(events)), err, ERRCAP), err)

    def download_bed(self, bed_host):
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_plan_download_bed(self._h, C.c_void_p(_ptr(bed_host)), err, ERRCAP), err)

    def run(self):
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_plan_run(self._h, C.byref(self.out), err, ERRCAP), err)
        return self.out

    def run_streamed(self, bed_host):
        """upload + run with the host->HBM copy of chunk c+1 hidden behind the compute of chunk c"""
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_plan_run_streamed(self._h, C.c_void_p(_ptr(bed_host)), C.byref(self.out), err, ERRCAP), err)
        return self.out

    def run_to_host(self, bed_host, A_lower, N_lower=None, mu=None):
        """upload + run + download, all pipelined (what gcta_b200_make_grm runs)"""
        self.out.A_lower, self.out.N_lower, self.out.mu_out = _ptr(A_lower), _ptr(N_lower), _ptr(mu)
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_plan_run_to_host(self._h, C.c_void_p(_ptr(bed_host)), C.byref(self.out), err, ERRCAP), err)
        return self.out

    def download(self, A_lower, N_lower=None, mu=None):
        self.out.A_lower, self.out.N_lower, self.out.mu_out = _ptr(A_lower), _ptr(N_lower), _ptr(mu)
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_plan_download(self._h, C.byref(self.out), err, ERRCAP), err)
        return self.out

    def device_results(self):
        dA, dN, cnt = C.c_void_p(), C.c_void_p(), C.c_int64()
        rc = lib().gcta_b200_plan_device_results(self._h, C.byref(dA), C.byref(dN), C.byref(cnt))
        if rc:
            raise GctaB200Error(rc, "plan has not run")
        return dA.value, dN.value, cnt.value

    def close(self):
        if self._h:
            lib().gcta_b200_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def exchange_slices(m_file: int, world: int, rank: int):
    """[(row_begin, row_end)] per segment: the file rows `rank` uploads in gcta_b200_exchange_post (no device needed)"""
    L = lib()
    L.gcta_b200_exchange_slice.restype = C.c_int64
    nseg = L.gcta_b200_exchange_slice(C.c_int64(m_file), C.c_int32(world), C.c_int32(rank), C.c_int64(0), None, None)
    out = []
    for s in range(nseg):
        b, e = C.c_int64(), C.c_int64()
        L.gcta_b200_exchange_slice(C.c_int64(m_file), C.c_int32(world), C.c_int32(rank), C.c_int64(s), C.byref(b), C.byref(e))
        out.append((b.value, e.value))
    return out


class Exchange:
    """gcta_b200_exchange_*: peer exchange of the packed rows between the ranks' plans (copy engines over NVLink).

    ex = Exchange(plan, rank, world); publish ex.blob to every rank; ex.connect(b"".join(blobs in rank order));
    per step: <host barrier> ex.post(bed_host) <host barrier> ex.pull(); plan.run_to_host(None, ...); ex.finish()."""

    def __init__(self, plan: "Plan", rank: int, world: int):
        L = lib()
        L.gcta_b200_exchange_blob_bytes.restype = C.c_int64
        self.plan = plan   # keeps the plan alive as long as the exchange
        self._h = C.c_void_p()
        nbytes = L.gcta_b200_exchange_blob_bytes(plan._h)
        blob = C.create_string_buffer(nbytes)
        err = C.create_string_buffer(ERRCAP)
        _check(L.gcta_b200_exchange_create(plan._h, C.c_int32(rank), C.c_int32(world), blob, C.byref(self._h), err, ERRCAP), err)
        self.blob = blob.raw
        self.world = world

    def connect(self, all_blobs: bytes):
        if len(all_blobs) != len(self.blob) * self.world:
            raise GctaB200Error(1, "exchange.connect: need one blob per rank")
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_exchange_connect(self._h, all_blobs, err, ERRCAP), err)

    def post(self, bed_host):
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_exchange_post(self._h, C.c_void_p(_ptr(bed_host)), err, ERRCAP), err)

    def pull(self):
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_exchange_pull(self._h, err, ERRCAP), err)

    def finish(self):
        err = C.create_string_buffer(ERRCAP)
        _check(lib().gcta_b200_exchange_finish(self._h, err, ERRCAP), err)

    def disconnect(self):
        """unmap the peers' buffers; every rank, then a host barrier, then close()"""
        if self._h:
            lib().gcta_b200_exchange_disconnect(self._h)

    def close(self):
        if self._h:
            lib().gcta_b200_exchange_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def call(name, *args):
    """Kernel-level entry point: appends the (err, errcap) pair and raises on failure."""
    err = C.create_string_buffer(ERRCAP)
    fn = getattr(lib(), name)
    rc = fn(*args, err, C.c_size_t(ERRCAP))
    _check(rc, err)
