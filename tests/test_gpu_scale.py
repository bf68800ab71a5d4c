"""GPU parity at BASELINE scale (round-1 verdict, "Next round" items 1 and 8).

* every BASELINE config (c2 .. c5) with its FULL-SIZE pattern set -- 10 000 20-mers + reverse complements, 1 M read
  ids, 1000 20-mers -- over >= 4 M synthetic records: the reference binary (oracle/_ref/fxtract, unmodified
  sources; one process per core over record-aligned shards) against the GPU path through the C ABI and through the
  drop-in `fxtract` CLI; sha256 of the output, counts for -c;
* the same file over 2+ real devices (CLI round-robin over FXTRACT_DEVICES, and bench.py-style contiguous rank
  shards through fxtract_b200.shard) equals the 1-GPU run and the reference;
* FXTRACT_FULL_SCALE=1: the 100 M-record c2 and c4 counts against the reference (about a minute of all host cores).

The checker legs are the ones bench.py uses before it prints a number (bench.verify_and_time_cpu)."""
import ctypes as C
import hashlib
import os
import shutil
import subprocess

import pytest

import helpers

pytestmark = pytest.mark.gpu

SCALE_RECORDS = int(os.environ.get("FXTRACT_SCALE_RECORDS", "4000000"))


@pytest.fixture(scope="module")
def env(built):
    import torch
    import bench
    import fxtract_b200 as fx
    from fxtract_b200 import lib
    if not os.path.exists(bench.REF_BIN):
        pytest.skip("oracle/_ref/fxtract not built")
    d = bench.scratch_dir()
    yield dict(bench=bench, fx=fx, lib=lib, torch=torch, dir=d)
    shutil.rmtree(d, ignore_errors=True)


def cli_digest(bench, w, pats, pattern_path, path, extra_env=None):
    e = dict(os.environ, FXTRACT_PATH="raw")
    e.update(extra_env or {})
    p = subprocess.run(bench.reference_argv(helpers.PRODUCT_CLI, w, pats, pattern_path, path), stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE, env=e, timeout=900)
    assert p.returncode == 0, p.stderr[-400:]
    if w["count_only"]:
        return {"count": int(p.stdout.split()[-1])}
    return {"sha256": hashlib.sha256(p.stdout).hexdigest(), "bytes": len(p.stdout)}


@pytest.mark.parametrize("name", ["c2", "c3", "c4", "c5v", "c5c"])
def test_baseline_config_at_scale(env, name):
    bench, fx, lib = env["bench"], env["fx"], env["lib"]
    w = bench.workload_def(name, SCALE_RECORDS)
    pattern_path, pats = bench.workload_pattern_file(w, 100_000_000, env["dir"])      # the full-size pattern set
    ok, ref, gpu, cpu, nbytes = bench.verify_and_time_cpu(fx, lib, w, pats, pattern_path, 0, SCALE_RECORDS, env["dir"], 0)
    assert ok, (name, ref, gpu)
    if name == "c3":
        assert len(pats) == 20000
    if name == "c4":
        assert len(pats) >= 1_000_000 and (ref.get("bytes", 0) > 0)
    # the drop-in binary on the same records as ONE file
    one = bench.write_shards(w, 0, SCALE_RECORDS, 1, env["dir"], "cli_" + name)[0]
    try:
        assert cli_digest(bench, w, pats, pattern_path, one) == ref, name
    finally:
        os.remove(one)


def test_multi_gpu_equals_single_gpu_and_reference(env):
    """SURVEY section 7 step 6 gate: 1 / 2 / ... GPU outputs sha256-equal to 1 GPU and to the reference."""
    bench, fx, lib, torch = env["bench"], env["fx"], env["lib"], env["torch"]
    ndev = torch.cuda.device_count()
    if ndev < 2:
        pytest.skip("needs >= 2 GPUs")
    from fxtract_b200 import shard
    devs = ",".join(str(i) for i in range(ndev))
    nrec = 2_000_000
    for name in ("c2", "c5v", "c5c", "c3", "c4"):
        w = bench.workload_def(name, nrec)
        pattern_path, pats = bench.workload_pattern_file(w, nrec, env["dir"])
        one = bench.write_shards(w, 0, nrec, 1, env["dir"], "mg_" + name)[0]
        try:
            dt, outs = bench.run_reference_shards(bench.REF_BIN, w, pats, pattern_path, [one], True)
            ref = bench.digest_reference(w, outs)
            os.remove(outs[0])
            single = cli_digest(bench, w, pats, pattern_path, one, {"FXTRACT_DEVICES": "0", "FXTRACT_BATCH_MIB": "32"})
            multi = cli_digest(bench, w, pats, pattern_path, one, {"FXTRACT_DEVICES": devs, "FXTRACT_BATCH_MIB": "32"})
            assert single == ref and multi == ref, (name, ref, single, multi)
            # bench.py-style sharding: rank r scans its own contiguous record range on device r
            data = open(one, "rb").read()
            parts, total = [], 0
            for r, (first, cnt) in enumerate(shard.record_ranges(nrec, ndev)):
                lo, hi = bench.record_offset(w, 0, first), bench.record_offset(w, 0, first + cnt)
                ctx = fx.Context(w["mode"], target=w["target"], invert=bool(w["invert"]), count_only=bool(w["count_only"]),
                                 batch_bytes=hi - lo + 64, device=r)
                ctx.set_patterns(pats)
                ctx.submit_raw_bytes(data[lo:hi], True)
                res = ctx.collect()
                assert res.canonical_ok and res.n_records == cnt
                total += shard.reduce_counts(res.n_selected)
                parts.append(b"".join(data[lo + o:lo + o + ln] for o, ln in lib.result_runs(res)))
                ctx.close()
            if w["count_only"]:
                assert {"count": total} == ref
            else:
                out = shard.gather_in_rank_order(b"".join(parts))
                assert {"sha256": hashlib.sha256(out).hexdigest(), "bytes": len(out)} == ref, name
        finally:
            os.remove(one)


@pytest.mark.skipif(os.environ.get("FXTRACT_FULL_SCALE") != "1", reason="set FXTRACT_FULL_SCALE=1 (about a minute of all host cores)")
@pytest.mark.parametrize("name", ["c2", "c4"])
def test_full_size_counts_against_reference(env, name):
    """100 M records: the GPU count over the whole shard (one scan, 1.4 M tiles, 64-bit line prefixes) against the
    reference binary run with -c over record-aligned shards, one per core, generated and deleted shard by shard."""
    bench, fx, lib, torch = env["bench"], env["fx"], env["lib"], env["torch"]
    nrec = 100_000_000
    w = bench.workload_def(name, nrec)
    w["cli"] = list(w["cli"]) + ["-c"]
    w["count_only"] = 1
    pattern_path, pats = bench.workload_pattern_file(w, nrec, env["dir"])
    cores = bench.host_cores()
    nsh = max(cores, 16)
    want = 0
    for g0 in range(0, nsh, cores):
        procs = []
        for k in range(g0, min(nsh, g0 + cores)):
            lo, hi = nrec * k // nsh, nrec * (k + 1) // nsh
            path = os.path.join(env["dir"], "full_%s_%d.fq" % (name, k))
            cmd = "%s fastq %s %s && %s; rc=$?; rm -f %s; exit $rc" % (
                bench.FXSYNTH, bench.spec_string(w, lo, hi - lo), path,
                " ".join("'%s'" % a for a in bench.reference_argv(bench.REF_BIN, w, pats, pattern_path, path)), path)
            procs.append(subprocess.Popen(["bash", "-c", cmd], stdout=subprocess.PIPE))
        for p in procs:
            out, _ = p.communicate()
            assert p.returncode == 0
            want += int(out.split()[-1])
    spec = fx.SynthSpec(seed=w["seed"], first_record=0, n_records=nrec, read_len=150)
    n = lib.load_library().fxg_synth_bytes(C.byref(spec))
    d = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
    ctx = fx.Context(w["mode"], target=w["target"], count_only=True, batch_bytes=1 << 20)
    ctx.set_patterns(pats)
    ctx.synth_device(spec, d.data_ptr())
    ctx.scan_device_async(d.data_ptr(), n)
    r = ctx.scan_device_finish()
    ctx.close()
    assert r.canonical_ok and r.n_records == nrec
    assert r.n_selected == want, (name, r.n_selected, want)
    with open(os.path.join(helpers.ROOT, "gpurun_out", "full_scale_%s.txt" % name), "w") as f:
        f.write("%s: 100 M records, GPU selected %d == reference -c total %d over %d shards\n" % (name, r.n_selected, want, nsh))
