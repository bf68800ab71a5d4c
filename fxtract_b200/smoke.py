"""smoke(): one small invocation of the hot path on cuda:0, checked against the oracle.
(The oracle is the checker here, never the thing measured or shipped.)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run():
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("smoke() needs a CUDA device: the search path has no CPU fallback")
    sys.path.insert(0, ROOT)
    import fxtract_b200 as fx
    from fxtract_b200 import lib
    from oracle import fxoracle
    spec = fx.SynthSpec(seed=1, first_record=0, n_records=20000, read_len=150)
    host = lib.synth_host(spec)
    n = len(host)
    one = fx.SynthSpec(seed=1, first_record=1234, n_records=1, read_len=150)
    pat = lib.synth_host(one).split(b"\n")[1][37:57]
    ctx = fx.Context(fx.MODE_LITERAL, count_only=False, batch_bytes=max(n, 1 << 20))
    ctx.set_patterns([pat, fxoracle.revcomp(pat)])
    d = torch.empty(n + 64, dtype=torch.uint8, device="cuda:0")
    ctx.synth_device(spec, d.data_ptr())
    ctx.scan_device_async(d.data_ptr(), n)
    r = ctx.scan_device_finish()
    out = b"".join(host[o:o + l] for o, l in lib.result_runs(r))
    want = fxoracle.run(host, "r", pattern=pat)
    assert r.canonical_ok == 1 and r.n_records == 20000, (r.canonical_ok, r.violation, r.n_records)
    assert out == want and r.n_selected >= 1, (len(out), len(want))
    ctx.close()
    print("smoke ok: %d records, %d selected, output identical to oracle" % (r.n_records, r.n_selected))
