"""ctypes binding of libfxgpu.so (include/fxgpu.h).  Plumbing only."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FXGPU_LIB") or os.path.join(HERE, "_build", "libfxgpu.so")   # FXGPU_LIB: kernel-variant experiments only

MODE_LITERAL, MODE_MULTI, MODE_EXACT = 0, 1, 2
TARGET_SEQ, TARGET_NAME, TARGET_COMMENT, TARGET_QUAL = 0, 1, 2, 3


class FxgError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("mode", C.c_int32), ("target", C.c_int32), ("invert", C.c_int32),
                ("count_only", C.c_int32), ("want_pattern_id", C.c_int32), ("batch_bytes", C.c_uint64),
                ("stream", C.c_void_p), ("route_runs", C.c_int32), ("reserved", C.c_int32)]


class Run(C.Structure):
    _fields_ = [("off", C.c_uint64), ("len", C.c_uint64)]


class Result(C.Structure):
    _fields_ = [("n_records", C.c_uint64), ("n_matched", C.c_uint64), ("n_selected", C.c_uint64),
                ("n_lines", C.c_uint64), ("consumed_bytes", C.c_uint64), ("canonical_ok", C.c_int32),
                ("violation", C.c_uint32), ("first_bad_offset", C.c_uint64), ("n_runs", C.c_uint64),
                ("runs", C.POINTER(Run)), ("flags", C.POINTER(C.c_uint8)), ("pattern_id", C.POINTER(C.c_int32)),
                ("run_pattern_id", C.POINTER(C.c_int32))]


NOALLOC_MAX_BYTES = 8 << 20


class NoAllocCounts(C.Structure):
    _fields_ = [("n_lines", C.c_uint64), ("n_records", C.c_uint64), ("n_matched", C.c_uint64),
                ("first_bad_offset", C.c_uint64), ("violation", C.c_uint32), ("canonical_ok", C.c_uint32)]


class SynthSpec(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("first_record", C.c_uint64), ("n_records", C.c_uint64),
                ("read_len", C.c_uint32), ("plant_mode", C.c_uint32), ("pattern_seed", C.c_uint64),
                ("n_patterns", C.c_uint32), ("pattern_len", C.c_uint32), ("plant_per_million", C.c_uint32),
                ("plant_revcomp", C.c_uint32)]


EXPORTS = ["fxg_device_count", "fxg_create", "fxg_destroy", "fxg_last_error", "fxg_set_patterns",
           "fxg_acquire_buffer", "fxg_submit_raw", "fxg_submit_raw_from", "fxg_submit_raw_interleaved_from",
           "fxg_submit_raw_pair_from", "fxg_scan_device_pairs_async", "fxg_submit_fields", "fxg_collect",
           "fxg_count", "fxg_find_record_cut", "fxg_scan_device_async", "fxg_scan_device_finish",
           "fxg_count_device_noalloc", "fxg_can_route", "fxg_last_launch_count", "fxg_last_d2h_bytes", "fxg_set_profiling", "fxg_last_kernel_ms", "fxg_synth_bytes", "fxg_synth_record_offset", "fxg_synth_host",
           "fxg_synth_device", "fxg_synth_pattern", "fxg_synth_name", "fxg_host_alloc", "fxg_host_free", "fxg_copy_to_host", "fxg_copy_to_device"]

_lib = None


def load_library():
    """Loads the in-tree CUDA library; raises loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FxgError("libfxgpu.so not built (%s): run `make` / __graft_entry__.build(); there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    lib.fxg_last_error.restype = C.c_char_p
    lib.fxg_last_error.argtypes = [C.c_void_p]
    lib.fxg_create.argtypes = [C.POINTER(Config), C.POINTER(C.c_void_p)]
    lib.fxg_destroy.argtypes = [C.c_void_p]
    lib.fxg_destroy.restype = None
    lib.fxg_set_patterns.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_uint32), C.c_uint32]
    lib.fxg_acquire_buffer.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
    lib.fxg_submit_raw.argtypes = [C.c_void_p, C.c_size_t, C.c_int]
    lib.fxg_submit_raw_from.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
    lib.fxg_submit_raw_interleaved_from.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
    lib.fxg_submit_raw_pair_from.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_int]
    lib.fxg_scan_device_pairs_async.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64]
    lib.fxg_submit_fields.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_uint32), C.c_uint32]
    lib.fxg_collect.argtypes = [C.c_void_p, C.POINTER(Result)]
    lib.fxg_count.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    lib.fxg_find_record_cut.argtypes = [C.c_void_p, C.c_size_t]
    lib.fxg_find_record_cut.restype = C.c_size_t
    lib.fxg_scan_device_async.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
    lib.fxg_scan_device_finish.argtypes = [C.c_void_p, C.POINTER(Result)]
    lib.fxg_count_device_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.fxg_can_route.argtypes = [C.c_void_p]
    lib.fxg_last_launch_count.argtypes = [C.c_void_p]
    lib.fxg_last_d2h_bytes.argtypes = [C.c_void_p]
    lib.fxg_last_d2h_bytes.restype = C.c_uint64
    lib.fxg_set_profiling.argtypes = [C.c_void_p, C.c_int]
    lib.fxg_last_kernel_ms.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.fxg_synth_bytes.argtypes = [C.POINTER(SynthSpec)]
    lib.fxg_synth_bytes.restype = C.c_uint64
    lib.fxg_synth_record_offset.argtypes = [C.POINTER(SynthSpec), C.c_uint64]
    lib.fxg_synth_record_offset.restype = C.c_uint64
    lib.fxg_synth_host.argtypes = [C.POINTER(SynthSpec), C.c_void_p]
    lib.fxg_synth_device.argtypes = [C.c_void_p, C.POINTER(SynthSpec), C.c_void_p]
    lib.fxg_synth_pattern.argtypes = [C.POINTER(SynthSpec), C.c_uint32, C.c_void_p]
    lib.fxg_synth_name.argtypes = [C.POINTER(SynthSpec), C.c_uint64, C.c_void_p, C.POINTER(C.c_uint32)]
    lib.fxg_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
    lib.fxg_host_free.argtypes = [C.c_void_p]
    lib.fxg_copy_to_host.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    lib.fxg_copy_to_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    lib.fxg_device_count.argtypes = [C.POINTER(C.c_int)]
    _lib = lib
    return lib


def find_record_cut(buf):
    lib = load_library()
    b = bytes(buf)
    return lib.fxg_find_record_cut(b, len(b))


def find_record_cut_ptr(ptr, n):
    """Same, over n bytes of host memory at address ptr (no copy)."""
    return load_library().fxg_find_record_cut(C.c_void_p(ptr), n)


def host_alloc(n):
    p = C.c_void_p()
    rc = load_library().fxg_host_alloc(C.byref(p), n)
    if rc != 0:
        raise FxgError("fxg_host_alloc(%d) failed: %s" % (n, load_library().fxg_last_error(None).decode()))
    return p.value


def host_free(ptr):
    load_library().fxg_host_free(C.c_void_p(ptr))


def pack_patterns(patterns):
    blob = b"".join(patterns)
    off = (C.c_uint32 * (len(patterns) + 1))()
    o = 0
    for i, p in enumerate(patterns):
        off[i] = o
        o += len(p)
    off[len(patterns)] = o
    return blob, off


class Context:
    """One fxg_ctx.  Thin: every method is one C-ABI call."""

    def __init__(self, mode, target=TARGET_SEQ, invert=False, count_only=False, want_pattern_id=False,
                 device=0, batch_bytes=0, stream=None, route_runs=False):
        self.lib = load_library()
        cfg = Config(device, mode, target, int(invert), int(count_only), int(want_pattern_id), batch_bytes,
                     C.c_void_p(stream) if stream else None, int(route_runs), 0)
        self.h = C.c_void_p()
        rc = self.lib.fxg_create(C.byref(cfg), C.byref(self.h))
        if rc != 0:
            raise FxgError("fxg_create failed (%d): %s" % (rc, self.lib.fxg_last_error(None).decode()))
        self.cfg = cfg

    def _check(self, rc):
        if rc != 0:
            raise FxgError("libfxgpu error %d: %s" % (rc, self.lib.fxg_last_error(self.h).decode()))

    def close(self):
        if self.h:
            self.lib.fxg_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_patterns(self, patterns):
        blob, off = pack_patterns(patterns)
        self._check(self.lib.fxg_set_patterns(self.h, blob, off, len(patterns)))

    def can_route(self):
        return bool(self.lib.fxg_can_route(self.h))

    # ---- host path
    def acquire(self):
        p = C.c_void_p()
        cap = C.c_size_t()
        self._check(self.lib.fxg_acquire_buffer(self.h, C.byref(p), C.byref(cap)))
        return p.value, cap.value

    def submit_raw_bytes(self, data, is_last=True):
        p, cap = self.acquire()
        assert len(data) <= cap
        C.memmove(p, data, len(data))
        self._check(self.lib.fxg_submit_raw(self.h, len(data), int(is_last)))

    def submit_raw_from(self, ptr, nbytes, is_last=False):
        self._check(self.lib.fxg_submit_raw_from(self.h, C.c_void_p(ptr), nbytes, int(is_last)))

    def submit_raw_interleaved_from(self, ptr, nbytes, is_last=False):
        self._check(self.lib.fxg_submit_raw_interleaved_from(self.h, C.c_void_p(ptr), nbytes, int(is_last)))

    def submit_raw_pair_from(self, ptr1, n1, ptr2, n2, is_last=False):
        self._check(self.lib.fxg_submit_raw_pair_from(self.h, C.c_void_p(ptr1), n1, C.c_void_p(ptr2), n2, int(is_last)))

    def submit_fields(self, fields):
        p, cap = self.acquire()
        blob = b"".join(fields)
        assert len(blob) <= cap
        C.memmove(p, blob, len(blob))
        _, off = pack_patterns(fields)
        self._check(self.lib.fxg_submit_fields(self.h, len(blob), off, len(fields)))

    def collect(self):
        r = Result()
        self._check(self.lib.fxg_collect(self.h, C.byref(r)))
        return r

    def count(self):
        t = C.c_uint64()
        self._check(self.lib.fxg_count(self.h, C.byref(t)))
        return t.value

    # ---- device-resident path
    def scan_device_async(self, dptr, nbytes):
        self._check(self.lib.fxg_scan_device_async(self.h, C.c_void_p(dptr), nbytes))

    def scan_device_pairs_async(self, dptr, nbytes, split):
        self._check(self.lib.fxg_scan_device_pairs_async(self.h, C.c_void_p(dptr), nbytes, split))

    def count_device_noalloc(self, dptr, nbytes, d_out):
        """Enqueues the allocation-free literal count; d_out = device address of one NoAllocCounts.  Returns the raw
        status (0 = ok) so that callers can look at refusals; fxg_last_error() has the text."""
        return self.lib.fxg_count_device_noalloc(self.h, C.c_void_p(dptr), nbytes, C.c_void_p(d_out))

    def last_error(self):
        return self.lib.fxg_last_error(self.h).decode()

    def scan_device_finish(self):
        r = Result()
        self._check(self.lib.fxg_scan_device_finish(self.h, C.byref(r)))
        return r

    def last_launch_count(self):
        return self.lib.fxg_last_launch_count(self.h)

    def last_d2h_bytes(self):
        return self.lib.fxg_last_d2h_bytes(self.h)

    def set_profiling(self, on=True):
        self._check(self.lib.fxg_set_profiling(self.h, int(on)))

    def last_kernel_ms(self):
        a, b = C.c_float(), C.c_float()
        self._check(self.lib.fxg_last_kernel_ms(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def copy_to_host(self, host_ptr, dptr, n):
        self._check(self.lib.fxg_copy_to_host(self.h, C.c_void_p(host_ptr), C.c_void_p(dptr), n))

    def copy_to_device(self, dptr, host_ptr, n):
        self._check(self.lib.fxg_copy_to_device(self.h, C.c_void_p(dptr), C.c_void_p(host_ptr), n))

    def synth_device(self, spec, dptr):
        self._check(self.lib.fxg_synth_device(self.h, C.byref(spec), C.c_void_p(dptr)))


def result_runs(r):
    return [(r.runs[i].off, r.runs[i].len) for i in range(r.n_runs)]


def synth_host(spec):
    lib = load_library()
    n = lib.fxg_synth_bytes(C.byref(spec))
    buf = C.create_string_buffer(n)
    rc = lib.fxg_synth_host(C.byref(spec), buf)
    if rc != 0:
        raise FxgError("fxg_synth_host failed")
    return buf.raw


def synth_pattern(spec, p):
    lib = load_library()
    buf = C.create_string_buffer(spec.pattern_len)
    lib.fxg_synth_pattern(C.byref(spec), p, buf)
    return buf.raw


def synth_name(spec, i):
    lib = load_library()
    buf = C.create_string_buffer(64)
    n = C.c_uint32()
    lib.fxg_synth_name(C.byref(spec), i, buf, C.byref(n))
    return buf.raw[:n.value]
