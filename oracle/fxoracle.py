"""ctypes binding of the oracle (oracle/_build/libfxoracle.so).  TEST INFRASTRUCTURE ONLY:
import it from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg, never from
the product package."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "libfxoracle.so")


class Opts(C.Structure):
    _fields_ = [(k, C.c_int) for k in ("H", "Q", "C", "X", "r", "I", "v", "c", "l", "S", "have_f")]


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB)
        L.fxo_job_new.restype = C.c_void_p
        L.fxo_job_new.argtypes = [C.POINTER(Opts)]
        L.fxo_job_free.argtypes = [C.c_void_p]
        L.fxo_job_set_pattern.argtypes = [C.c_void_p, C.c_char_p]
        L.fxo_job_load_patterns.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        L.fxo_job_run.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_size_t, C.c_char_p, C.c_char_p, C.c_size_t]
        L.fxo_job_stdout.restype = C.POINTER(C.c_uint8)
        L.fxo_job_stdout.argtypes = [C.c_void_p, C.POINTER(C.c_size_t)]
        L.fxo_job_matched.restype = C.c_long
        L.fxo_job_matched.argtypes = [C.c_void_p]
        L.fxo_revcomp.argtypes = [C.c_char_p, C.c_size_t]
        _lib = L
    return _lib


def run(data, flags="", pattern=None, pattern_file=None, data2=None):
    """Runs the restated search path over one in-memory (pair of) file(s).
    flags: string of option letters (e.g. "Hvc"); returns stdout bytes."""
    L = lib()
    o = Opts()
    for ch in flags:
        if ch in "HQCXrIvclS":
            setattr(o, ch, 1)
    o.have_f = 1 if pattern_file is not None else 0
    j = L.fxo_job_new(C.byref(o))
    try:
        if pattern_file is not None:
            L.fxo_job_load_patterns(j, pattern_file, len(pattern_file))
        else:
            L.fxo_job_set_pattern(j, pattern)
        L.fxo_job_run(j, b"file1", data, len(data), b"file2" if data2 is not None else None, data2,
                      len(data2) if data2 is not None else 0)
        n = C.c_size_t()
        p = L.fxo_job_stdout(j, C.byref(n))
        return C.string_at(p, n.value)
    finally:
        L.fxo_job_free(j)


def revcomp(s):
    b = C.create_string_buffer(s, len(s))
    lib().fxo_revcomp(b, len(s))
    return b.raw[:len(s)]


class AC:
    """fxo_ac_* (Aho-Corasick with the documented report order)."""

    def __init__(self, patterns):
        L = lib()
        L.fxo_ac_build.restype = C.c_void_p
        L.fxo_ac_build.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.c_int]
        L.fxo_ac_first.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.fxo_ac_free.argtypes = [C.c_void_p]
        n = len(patterns)
        self._keep = [C.create_string_buffer(p, len(p)) for p in patterns]
        arr = (C.c_char_p * n)(*[C.cast(b, C.c_char_p) for b in self._keep])
        lens = (C.c_size_t * n)(*[len(p) for p in patterns])
        self.h = L.fxo_ac_build(arr, lens, n)

    def first(self, text):
        end = C.c_size_t()
        return lib().fxo_ac_first(self.h, text, len(text), C.byref(end))

    def __del__(self):
        try:
            lib().fxo_ac_free(self.h)
        except Exception:
            pass
