#!/usr/bin/env python
"""bench.py -- FASTQ GB/s scanned by the B200 search path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--records R] [--workload c2] [--also c3,c4,c5v,c5c]
  python bench.py --impl reference ...      # the reference CPU binary on the host cores

A "step" is one pass of the hot path (fused raw-FASTQ scan: record index + canonical predicate + matcher, then
count / order-preserving compaction) over one rank's whole synthetic shard, resident in HBM.  Headline workload =
BASELINE.json configs[1] ("c2": one 20-mer over 100 M x 150 bp synthetic FASTQ, ~34.7 GB per GPU); with N GPUs
every rank holds its own contiguous 100 M-record range of the stream (weak scaling, no data-path collective; only
the match count is summed).  `value` = bytes scanned by all ranks / max-over-ranks device time.

The same process then measures the other BASELINE configs at the same size -- c3 (-r -f 10 000 20-mers), c4
(-H -X -f 1 M read ids), c5v / c5c (-f 1000 20-mers, -v / -c) -- into `workloads`, each with its own roofline
block; at N > 1 those are the sharded 1/2/4/8-GPU figures BASELINE config 5 asks for.

`e2e` = the headline work through the C ABI from HOST (pinned) buffers: record-aligned batches cut on the host,
cudaMemcpyAsync H2D (copy stream, two device buffers), kernels, results D2H, selected records written out;
`e2e.bound_gbs` is a pinned host->device copy of the same bytes measured in the same run (all ranks at once).
`e2e_cli` = the drop-in `fxtract` binary reading a file and writing to /dev/null.

Nothing is printed unverified: before timing, every workload is run on a bounded sample (the first --cpu-records
records of rank 0's shard) by the reference binary (oracle/_ref/fxtract, unmodified sources; all host cores, one
process per record-aligned shard) and by the GPU path; outputs must be sha256-equal (counts for -c).  That same
reference run is the `cpu_baseline` timing.  The full-size runs are then checked against what is known exactly
(c4: the number of ids; all: the count over the whole shard equals the sum over independently scanned slices).
Only these checker legs and `--impl reference` execute anything under oracle/; the measured GPU path never does.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import shutil
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "fastq_scan_throughput"
UNIT = "GB/s"
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "fxtract")
REF_BIN_O0 = os.path.join(ROOT, "oracle", "_ref", "fxtract_O0")
FXSYNTH = os.path.join(ROOT, "fxtract_b200", "_build", "fxsynth")
PRODUCT_CLI = os.path.join(ROOT, "fxtract_b200", "_build", "fxtract")
PATTERN_RECORD = 1234          # the single 20-mer of c2 is copied from this global record
COMP = bytes.maketrans(b"ACGT", b"TGCA")
KERNEL_NAMES = {
    "c2": "fxg::scan_tiles_kernel<Geom<24576,384,3,1>, 9 (literal fingerprint), false> (fxg_scan_kernel.cuh)",
    "c3": "fxg::scan_tiles_kernel<Geom<16384,512,1,2>, 19 (window matcher, 128 KiB filter, two groups), false> (fxg_scan_kernel.cuh)",
    "c4": "fxg::scan_tiles_kernel<Geom<24576,384,3,1>, 30 (exact: key hash), false> (fxg_scan_kernel.cuh)",
    "c5v": "fxg::scan_tiles_kernel<Geom<32768,512,2,1>, 19 (window matcher, 32 KiB filter), false> (fxg_scan_kernel.cuh)",
    "c5c": "fxg::scan_tiles_kernel<Geom<32768,512,2,1>, 19 (window matcher, 32 KiB filter), false> (fxg_scan_kernel.cuh)",
    "ms": "fxg::scan_tiles_kernel<Geom<32768,512,2,1>, 18..20 (window matcher, 32 KiB filter), false> (fxg_scan_kernel.cuh)",
    "hf": "fxg::scan_tiles_kernel<Geom<24576,384,3,1>, 0 (index pass), false> (fxg_scan_kernel.cuh); the matcher is fields_ac_kernel (fxg_fields.cu), inside pipeline_ms",
}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------- workloads
def workload_def(name, records):
    """dict(mode, target, invert, count_only, cli flags of the reference, synthetic-plant fields)."""
    w = dict(name=name, records=records, read_len=150, seed=1, plant_mode=0, pattern_seed=0, n_patterns=0,
             pattern_len=20, plant_per_million=0, plant_revcomp=0, target=0, invert=0, count_only=0, revcomp=False)
    if name == "c2":
        w.update(mode=0, desc="single 20-mer over synthetic %d x 150bp FASTQ" % records, cli=[])
    elif name == "c2c":
        w.update(mode=0, count_only=1, desc="single 20-mer, -c, over synthetic %d x 150bp FASTQ" % records, cli=["-c"])
    elif name == "c3":
        w.update(mode=1, plant_mode=1, pattern_seed=2, n_patterns=10000, plant_per_million=10000, plant_revcomp=1,
                 revcomp=True, desc="-r -f 10k 20-mers (+revcomp) over synthetic %d-read FASTQ" % records, cli=["-r"])
    elif name == "c4":
        w.update(mode=2, target=1, n_ids=1000000, desc="-HXf 1M read ids over synthetic %d-record FASTQ" % records,
                 cli=["-H", "-X"])
    elif name in ("c5v", "c5c"):
        w.update(mode=1, plant_mode=1, pattern_seed=4, n_patterns=1000, plant_per_million=10000, plant_revcomp=0,
                 invert=1 if name == "c5v" else 0, count_only=1 if name == "c5c" else 0,
                 desc="-f 1000 20-mers %s over synthetic %d-read FASTQ" % ("-v" if name == "c5v" else "-c", records),
                 cli=["-v"] if name == "c5v" else ["-c"])
    elif name == "ms":
        # the reference's own speed-test pattern set (speed_test/random_1000_patterns_minus_small.txt, committed under
        # tests/golden/patterns/): 1000 patterns of 16..60 bytes, not planted -- measures the filter, not the hits
        w.update(mode=1, pattern_source=os.path.join(ROOT, "tests", "golden", "patterns", "random_1000_patterns_minus_small.txt"),
                 desc="-f random_1000_patterns_minus_small.txt over synthetic %d-read FASTQ" % records, cli=[])
    elif name == "hf":
        # multi-pattern search on the read NAME (-H -f): 1000 name fragments ":x:y:" -- served by the index pass + the
        # Aho-Corasick field matcher (DFA in shared memory), not by the window matcher
        w.update(mode=1, target=1, name_fragments=1000,
                 desc="-H -f 1000 name fragments over synthetic %d-record FASTQ" % records, cli=["-H"])
    else:
        raise SystemExit("unknown workload %s" % name)
    return w


def spec_string(w, first, n):
    return "%d:%d:%d:%d:%d:%d:%d:%d:%d:%d" % (w["seed"], first, n, w["read_len"], w["plant_mode"], w["pattern_seed"],
                                              w["n_patterns"], w["pattern_len"], w["plant_per_million"], w["plant_revcomp"])


def fxsynth(*args):
    if not os.path.exists(FXSYNTH):
        raise SystemExit("bench.py: %s not built (make tools)" % FXSYNTH)
    return subprocess.run([FXSYNTH] + [str(a) for a in args], check=True, stdout=subprocess.PIPE).stdout


def record_offset(w, first, i):
    """byte offset of local record i of the range that starts at global record `first` (same as fxg_synth_record_offset)"""
    return int(fxsynth("bytes", spec_string(w, first, i)))


def workload_pattern_file(w, total_records, directory):
    """Writes the pattern file as the reference CLI takes it (-r expansion is the CLI's own job) and returns
    (path, pattern list exactly as the host CLI hands it to the library: expanded, de-duplicated, std::map order)."""
    path = os.path.join(directory, "patterns_%s.txt" % w["name"])
    if w["mode"] == 0:
        one = os.path.join(directory, "one.fq")
        fxsynth("fastq", spec_string(w, PATTERN_RECORD, 1), one)
        seq = open(one, "rb").read().split(b"\n")[1]
        os.remove(one)
        return None, [seq[37:57]]
    if w.get("pattern_source"):
        shutil.copyfile(w["pattern_source"], path)
        return path, sorted(set(l for l in open(path, "rb").read().split(b"\n") if l))
    if w.get("name_fragments"):
        n = min(w["name_fragments"], total_records)
        fxsynth("names", spec_string(w, 0, 1), max(1, total_records // n), n, path)
        frags = sorted(set(b":" + b":".join(l.split(b":")[4:6]) + b":" for l in open(path, "rb").read().split(b"\n") if l))
        with open(path, "wb") as f:
            f.write(b"\n".join(frags) + b"\n")
        return path, frags
    if w["mode"] == 1:
        fxsynth("patterns", spec_string(w, 0, 1), path)
        base = set(l for l in open(path, "rb").read().split(b"\n") if l)
        pats = set(base)
        if w["revcomp"]:
            pats |= set(p.translate(COMP)[::-1] for p in base)
        return path, sorted(pats)
    # exact ids: every (total/n_ids)-th read name of the whole stream + 1 % absent ids
    n_ids = min(w["n_ids"], total_records)
    stride = max(1, total_records // n_ids)
    fxsynth("names", spec_string(w, 0, 1), stride, n_ids, path)
    with open(path, "ab") as f:
        f.write(b"".join(b"SYN:1:FC:9:0000:00000:%d\n" % i for i in range(n_ids // 100)))
    ids = set(l for l in open(path, "rb").read().split(b"\n") if l)
    return path, sorted(ids)


# ---------------------------------------------------------------- clocks
class ClockSampler:
    FIELDS = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.samples = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) >= 8:
                try:
                    self.samples.append((time.time(), float(parts[1]), float(parts[2]), float(parts[3]), parts[4:8]))
                except ValueError:
                    pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self, windows):
        inside = [s for s in self.samples if any(a - 0.02 <= s[0] <= b + 0.06 for a, b in windows)]
        use = inside if len(inside) >= 2 else self.samples
        if not use:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted(set(n for s in use for n, v in zip(names, s[4]) if v.lower().startswith("active")))
        return {"sm_mhz": statistics.median(s[1] for s in use), "sm_max_mhz": max(s[2] for s in use),
                "power_w_max": max(s[3] for s in use), "reasons": reasons, "samples": len(use),
                "samples_in_timed_region": len(inside)}


# ---------------------------------------------------------------- reference runs (checker + CPU baseline)
def scratch_dir():
    for d in ("/dev/shm", tempfile.gettempdir()):
        try:
            if os.path.isdir(d) and shutil.disk_usage(d).free > (8 << 30):
                return tempfile.mkdtemp(prefix="fxbench_", dir=d)
        except Exception:
            pass
    return tempfile.mkdtemp(prefix="fxbench_")


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def write_shards(w, first, n_records, nshards, directory, tag):
    """The record range [first, first + n_records) as `nshards` record-aligned FASTQ files (fxsynth, no libfxgpu)."""
    paths, procs = [], []
    for k in range(nshards):
        lo, hi = n_records * k // nshards, n_records * (k + 1) // nshards
        p = os.path.join(directory, "%s_%03d.fq" % (tag, k))
        paths.append(p)
        if hi > lo:
            procs.append(subprocess.Popen([FXSYNTH, "fastq", spec_string(w, first + lo, hi - lo), p]))
        else:
            open(p, "wb").close()
    for pr in procs:
        if pr.wait() != 0:
            raise RuntimeError("fxsynth failed")
    return paths


def reference_argv(binary, w, pats, pattern_path, sample):
    argv = [binary] + list(w["cli"])
    if w["mode"] == 0:
        return argv + ["-S", pats[0].decode(), sample]
    return argv + ["-S", "-f", pattern_path, sample]


def run_reference_shards(binary, w, pats, pattern_path, shards, keep_output):
    """One reference process per shard, all at once (the reference has no threads).  Returns (seconds, per-shard
    stdout paths or None)."""
    outs = [s + ".out" if keep_output else os.devnull for s in shards]
    t0 = time.perf_counter()
    procs = []
    for s, o in zip(shards, outs):
        f = open(o, "wb")
        procs.append((subprocess.Popen(reference_argv(binary, w, pats, pattern_path, s), stdout=f, stderr=subprocess.DEVNULL), f))
    rcs = []
    for p, f in procs:
        rcs.append(p.wait())
        f.close()
    dt = time.perf_counter() - t0
    if any(rcs):
        raise RuntimeError("reference binary failed: %r" % rcs)
    return dt, (outs if keep_output else None)


def digest_reference(w, outs):
    """-c: sum of the printed counts; else sha256 + size of the shard outputs in input order."""
    if w["count_only"]:
        return {"count": sum(int(open(o, "rb").read().split()[-1]) for o in outs)}
    h, n = hashlib.sha256(), 0
    for o in outs:
        with open(o, "rb") as f:
            while True:
                b = f.read(1 << 24)
                if not b:
                    break
                h.update(b)
                n += len(b)
    return {"sha256": h.hexdigest(), "bytes": n}


def gpu_sample_digest(fx, lib, w, pats, shards, local):
    """The same sample through the C ABI from host memory (submit_raw_from, 64 MiB batches)."""
    total = sum(os.path.getsize(s) for s in shards)
    hptr = lib.host_alloc(total + 64)
    try:
        view = (C.c_uint8 * total).from_address(hptr)
        mem = memoryview(view).cast("B")
        off = 0
        for s in shards:
            with open(s, "rb") as f:
                off += f.readinto(mem[off:off + os.path.getsize(s)])
        ctx = fx.Context(w["mode"], target=w["target"], invert=bool(w["invert"]), count_only=bool(w["count_only"]),
                         batch_bytes=64 << 20, device=local)
        ctx.set_patterns(pats)
        h, nbytes, count, pend, pos = hashlib.sha256(), 0, 0, [], 0

        def drain():
            nonlocal nbytes, count
            base = pend.pop(0)
            r = ctx.collect()
            if not r.canonical_ok:
                raise RuntimeError("verification sample rejected: %#x" % r.violation)
            count += r.n_selected
            for i in range(r.n_runs):
                o, ln = r.runs[i].off, r.runs[i].len
                h.update(mem[base + o:base + o + ln])
                nbytes += ln
        while pos < total:
            n = min(64 << 20, total - pos)
            if pos + n < total:
                n = lib.find_record_cut_ptr(hptr + pos, n)
            if len(pend) == 2:
                drain()
            ctx.submit_raw_from(hptr + pos, n, pos + n >= total)
            pend.append(pos)
            pos += n
        while pend:
            drain()
        ctx.close()
        del mem, view
    finally:
        lib.host_free(hptr)
    return {"count": count} if w["count_only"] else {"sha256": h.hexdigest(), "bytes": nbytes}


def verify_and_time_cpu(fx, lib, w, pats, pattern_path, first, n_records, directory, local, with_o0=False):
    """Reference on all host cores over a bounded sample: the checker for the GPU path AND the cpu_baseline timing."""
    cores = host_cores()
    shards = write_shards(w, first, n_records, cores, directory, "v_" + w["name"])
    nbytes = sum(os.path.getsize(s) for s in shards)
    try:
        run_reference_shards(REF_BIN, w, pats, pattern_path, shards[:1], False)         # warm: binary + pattern file in page cache
        dt, outs = run_reference_shards(REF_BIN, w, pats, pattern_path, shards, True)
        ref = digest_reference(w, outs)
        for o in outs:
            os.remove(o)
        one_bytes = os.path.getsize(shards[0])
        one, _ = run_reference_shards(REF_BIN, w, pats, pattern_path, shards[:1], False)
        o0 = None
        if with_o0 and os.path.exists(REF_BIN_O0):
            o0, _ = run_reference_shards(REF_BIN_O0, w, pats, pattern_path, shards[:1], False)
        gpu = gpu_sample_digest(fx, lib, w, pats, shards, local)
    finally:
        for s in shards:
            if os.path.exists(s):
                os.remove(s)
    ok = ref == gpu
    cpu = {"value": nbytes / dt / 1e9, "unit": UNIT, "cores": cores, "kind": "reference",
           "sample": ("reference binary (oracle/_ref/fxtract, unmodified sources, g++ -O2) on the first %d records / %.0f MB of "
                      "the workload, one single-threaded process per core over %d record-aligned shards, page cache warm, "
                      "stdout to tmpfs; single_process_value = one process on one shard" % (n_records, nbytes / 1e6, cores)),
           "single_process_value": one_bytes / one / 1e9,
           "single_process_O0_value": (one_bytes / o0 / 1e9) if o0 else None}
    return ok, ref, gpu, cpu, nbytes


# ---------------------------------------------------------------- reference arm
def static_config(w, nbytes):
    return {"workload": w["desc"], "records_per_gpu": w["records"], "read_len": 150, "bytes_per_gpu": nbytes,
            "sharding": "contiguous record ranges, one per GPU; count summed, no data-path collective",
            "l2_policy": "input per step (%.1f GB) is far larger than the 126 MB L2" % (nbytes / 1e9)}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    w = workload_def(args.workload, args.records)
    nbytes_full = record_offset(w, 0, w["records"])
    cfg = static_config(w, nbytes_full)
    # the same keys as the GPU arm's config; what only a GPU run can fill stays None
    cfg.update({"verified_against": None, "verify_records": None, "expected_selected": None, "selected_records": None,
                "records": None, "runs_rank0": None})
    base = {"metric": METRIC, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic", "config": cfg}
    if not os.path.exists(REF_BIN) or not os.path.exists(FXSYNTH):
        base["unavailable"] = "oracle/_ref/fxtract or fxsynth not built (reference sources absent at build time)"
        print(json.dumps(base))
        return 0
    cores = host_cores()
    nrec = args.ref_records                      # per step: `cores` shards of the workload's first nrec records
    d = scratch_dir()
    try:
        pattern_path, pats = workload_pattern_file(w, w["records"] * args.gpus, d)
        shards = write_shards(w, 0, nrec, cores, d, "ref")
        nbytes = sum(os.path.getsize(s) for s in shards)
        for _ in range(args.warmup):
            run_reference_shards(REF_BIN, w, pats, pattern_path, shards, False)
        times = [run_reference_shards(REF_BIN, w, pats, pattern_path, shards, False)[0] for _ in range(args.steps)]
        one = run_reference_shards(REF_BIN, w, pats, pattern_path, shards[:1], False)[0]
        one_bytes = os.path.getsize(shards[0])
    finally:
        shutil.rmtree(d, ignore_errors=True)
    total = sum(times)
    value = nbytes * args.steps / total / 1e9
    smp = ("each step: %d concurrent single-threaded reference processes (oracle/_ref/fxtract, unmodified sources, g++ -O2), "
           "one per record-aligned shard of the workload's first %d records (%.0f MB in all), page cache warm, stdout to "
           "/dev/null; the full workload is %d records per GPU" % (cores, nrec, nbytes / 1e6, w["records"]))
    base.update({"value": value, "ms_per_step": 1e3 * total / args.steps,
                 "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": smp,
                                  "sample_records": nrec, "single_process_value": one_bytes / one / 1e9},
                 "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0})
    print(json.dumps(base))
    return 0


# ---------------------------------------------------------------- GPU arm
def bind_to_gpu_numa_node(torch, local):
    """Pin this rank to the CPUs of the NUMA node its GPU hangs off, before any pinned host buffer is allocated."""
    try:
        pr = torch.cuda.get_device_properties(local)
        if not hasattr(pr, "pci_bus_id"):
            return None
        bus = "%04x:%02x:%02x.0" % (getattr(pr, "pci_domain_id", 0), pr.pci_bus_id, getattr(pr, "pci_device_id", 0))
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read())
        if node < 0:
            return None
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except Exception:
        return None


def run_gpu_arm(args):
    import torch
    import fxtract_b200 as fx
    from fxtract_b200 import lib
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the search path has no CPU fallback)")
    claim_stdout()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    numa_node = None if args.no_numa_bind else bind_to_gpu_numa_node(torch, local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(vals_max, vals_sum):
        a = torch.tensor(vals_max, dtype=torch.float64, device="cuda")
        b = torch.tensor(vals_sum, dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(a, op=dist.ReduceOp.MAX)
            dist.all_reduce(b, op=dist.ReduceOp.SUM)
        return a.tolist(), b.tolist()

    L = lib.load_library()
    peak, peak_src = peaks()
    names = [args.workload] + [x for x in args.also.split(",") if x and x != args.workload]
    nrec = args.records
    first = rank * nrec
    stream = torch.cuda.Stream()
    sampler = ClockSampler(local) if rank == 0 else None
    windows = []
    scratch = scratch_dir()
    traffic = {}
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        pass
    results = {}
    d = None
    e2e = None
    verify_records = args.cpu_records if world == 1 else min(args.cpu_records, 1_000_000)

    try:
        with torch.cuda.stream(stream):
            for wi, name in enumerate(names):
                w = workload_def(name, nrec)
                spec = fx.SynthSpec(seed=w["seed"], first_record=first, n_records=nrec, read_len=w["read_len"],
                                    plant_mode=w["plant_mode"], pattern_seed=w["pattern_seed"], n_patterns=w["n_patterns"],
                                    pattern_len=w["pattern_len"], plant_per_million=w["plant_per_million"],
                                    plant_revcomp=w["plant_revcomp"])
                nbytes = L.fxg_synth_bytes(C.byref(spec))
                pattern_path, pats = workload_pattern_file(w, nrec * world, scratch)
                res = {"desc": w["desc"], "bytes_per_gpu": nbytes, "n_patterns": len(pats)}

                # ---- 1. checker + CPU baseline on a bounded sample (rank 0), before anything is timed
                ok = [1.0]
                if rank == 0 and not args.no_verify:
                    if not (os.path.exists(REF_BIN) and os.path.exists(FXSYNTH)):
                        res["verified_against"] = None
                        res["verify_note"] = "oracle/_ref/fxtract not built"
                    else:
                        good, ref, gpu, cpu, vbytes = verify_and_time_cpu(fx, lib, w, pats, pattern_path, first, min(verify_records, nrec),
                                                                          scratch, local, with_o0=(wi == 0 and world == 1))
                        res.update({"verified_against": "oracle/_ref", "verify_records": min(verify_records, nrec),
                                    "verify_bytes": vbytes, "verify_reference": ref, "verify_gpu": gpu, "verified": bool(good)})
                        if world == 1 and not args.no_cpu_baseline:
                            res["cpu_baseline"] = cpu
                        ok = [1.0 if good else 0.0]
                okv, _ = reduce([1.0 - ok[0]], [0.0])
                if okv[0] != 0.0:
                    raise SystemExit("bench.py: %s: GPU output differs from oracle/_ref on the verification sample: %r vs %r"
                                     % (name, res.get("verify_gpu"), res.get("verify_reference")))

                # ---- 2. the shard in HBM, device-timed steps
                if d is None or d.numel() < nbytes + 64:
                    d = None
                    d = torch.empty(nbytes + 64, dtype=torch.uint8, device="cuda")
                ctx = fx.Context(w["mode"], target=w["target"], invert=bool(w["invert"]), count_only=bool(w["count_only"]),
                                 batch_bytes=args.batch_mib << 20, stream=stream.cuda_stream, device=local)
                ctx.set_patterns(pats)
                ctx.synth_device(spec, d.data_ptr())
                ctx.set_profiling(True)

                def step(ptr=None, n=None, want=None):
                    ctx.scan_device_async(d.data_ptr() if ptr is None else ptr, nbytes if n is None else n)
                    r = ctx.scan_device_finish()
                    if not r.canonical_ok or (want is not None and r.n_records != want):
                        raise RuntimeError("scan rejected the synthetic shard: violation=%#x records=%d" % (r.violation, r.n_records))
                    return r

                steps = args.steps if wi == 0 else max(3, min(args.steps, args.also_steps))
                for _ in range(max(args.warmup, 3)):
                    r = step(want=nrec)
                launches = ctx.last_launch_count()
                ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                scan_ms, pipe_ms = [], []
                barrier()
                t_a = time.time()
                ev0.record(stream)
                for _ in range(steps):
                    r = step(want=nrec)
                    a, b = ctx.last_kernel_ms()
                    scan_ms.append(a)
                    pipe_ms.append(b)
                ev1.record(stream)
                barrier()
                windows.append((t_a, time.time()))
                dev_ms = ev0.elapsed_time(ev1)
                n_selected, n_runs = int(r.n_selected), int(r.n_runs)

                # ---- 3. full-size checks: what is known exactly, and whole == sum of independently scanned slices
                checks = {}
                if not args.no_verify:
                    cuts = [L.fxg_synth_record_offset(C.byref(spec), nrec * k // 3) for k in range(4)]
                    part = 0
                    scr = torch.empty(max(cuts[k + 1] - cuts[k] for k in range(3)) + 64, dtype=torch.uint8, device="cuda")
                    for k in range(3):
                        ln = cuts[k + 1] - cuts[k]
                        scr[:ln].copy_(d[cuts[k]:cuts[k + 1]])           # re-based: other tile alignment, other prefix spans
                        part += int(step(scr.data_ptr(), ln, nrec * (k + 1) // 3 - nrec * k // 3).n_selected)
                    del scr
                    checks["slices_sum"] = part
                    checks["slices_equal_whole"] = part == n_selected
                    if not checks["slices_equal_whole"]:
                        raise SystemExit("bench.py: %s: selected %d over the whole shard but %d over its three slices" % (name, n_selected, part))
                    if name == "c4":
                        stride = max(1, nrec * world // min(w["n_ids"], nrec * world))
                        expect = len([1 for i in range(min(w["n_ids"], nrec * world)) if first <= i * stride < first + nrec])
                        checks["expected_selected"] = expect
                        if expect != n_selected:
                            raise SystemExit("bench.py: c4: %d ids of this shard are in the list, %d records selected" % (expect, n_selected))
                    if name == "c2" and rank == (PATTERN_RECORD // nrec) and n_selected < 1:
                        raise SystemExit("bench.py: c2: the record the pattern was copied from was not selected")
                vmax, vsum = reduce([dev_ms], [float(nbytes), float(n_selected), float(r.n_records)])
                ms_per_step = vmax[0] / steps
                scan_avg = sum(scan_ms) / len(scan_ms)
                algo = nbytes + int(r.n_records)                 # N_in + 1 flag per record
                roof = {"bound": "hbm", "achieved": algo / (scan_avg * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": algo / (scan_avg * 1e-3) / 1e9 / peak, "traffic": None, "kernel": KERNEL_NAMES.get(name, "fxg::scan_tiles_kernel"),
                        "kernel_ms": scan_avg, "pipeline_ms": sum(pipe_ms) / len(pipe_ms), "algorithmic_bytes": algo, "peak_source": peak_src}
                per_byte = traffic.get(name, {}).get("dram_bytes_per_input_byte")
                if per_byte:
                    roof["traffic"] = per_byte * nbytes
                    roof["traffic_source"] = traffic[name].get("source")
                res.update({"value": vsum[0] / (ms_per_step * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": ms_per_step, "steps": steps,
                            "roofline": roof, "selected_records": int(vsum[1]), "records": int(vsum[2]), "runs_rank0": n_runs,
                            "launches_per_step": launches, "checks": checks})

                # ---- 4. end to end from host buffers through the C ABI (headline workload, and c5v: the one that
                #         selects nearly everything)
                if not args.no_e2e and (wi == 0 or name == "c5v"):
                    res["e2e"] = e2e_leg(args, torch, fx, lib, L, ctx, w, spec, d, nbytes, nrec, world, barrier, reduce, windows,
                                         numa_node, probe=(wi == 0))
                ctx.close()
                results[name] = res
    finally:
        shutil.rmtree(scratch, ignore_errors=True)

    # ---------------- the drop-in binary itself: file -> stdout
    cli = None
    if rank == 0 and world == 1 and not args.no_cli and os.path.exists(PRODUCT_CLI):
        try:
            cli = cli_leg(args, results)
        except Exception as e:        # the CLI leg is extra evidence; a failure must not lose the line
            cli = {"error": str(e)[:300]}
    if sampler:
        sampler.stop()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0

    head = results[names[0]]
    w0 = workload_def(names[0], nrec)
    cfg = static_config(w0, head["bytes_per_gpu"])
    cfg.update({"verified_against": head.get("verified_against"), "verify_records": head.get("verify_records"),
                "expected_selected": head["checks"].get("expected_selected"), "selected_records": head["selected_records"],
                "records": head["records"], "runs_rank0": head["runs_rank0"]})
    out = {"metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
           "warmup": max(args.warmup, 3), "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": cfg,
           "gpu_launches": head["launches_per_step"] * args.steps, "roofline": head["roofline"],
           "clocks": sampler.summary(windows) if sampler else None}
    if "e2e" in head:
        out["e2e"] = head["e2e"]
    if "cpu_baseline" in head:
        out["cpu_baseline"] = head["cpu_baseline"]
    if cli:
        out["e2e_cli"] = cli
    out["workloads"] = {k: {kk: vv for kk, vv in v.items() if kk not in ("verify_reference", "verify_gpu")} for k, v in results.items() if k != names[0]}
    out["verification"] = {k: {"verified": v.get("verified"), "records": v.get("verify_records"), "reference": v.get("verify_reference"),
                               "gpu": v.get("verify_gpu"), "checks": v.get("checks")} for k, v in results.items()}
    emit_line(out)
    return 0


def e2e_leg(args, torch, fx, lib, L, ctx, w, spec, d, nbytes, nrec, world, barrier, reduce, windows, numa_node, probe):
    avail = 0
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable"):
                avail = int(line.split()[1]) * 1024
    except Exception:
        pass
    e2e_bytes = nbytes
    limit = int(avail * 0.5 / max(world, 1)) if avail else (8 << 30)
    if args.e2e_gib > 0:
        limit = min(limit, int(args.e2e_gib * (1 << 30)))
    if e2e_bytes > limit:                 # host RAM too small for the whole shard: use a record-aligned prefix
        approx = int(limit // (nbytes / nrec))
        e2e_bytes = L.fxg_synth_record_offset(C.byref(spec), approx)
    hptr = lib.host_alloc(e2e_bytes + 64)
    ctx.copy_to_host(hptr, d.data_ptr(), e2e_bytes)
    batch = args.batch_mib << 20
    devnull = os.open(os.devnull, os.O_WRONLY)
    hview = (C.c_uint8 * e2e_bytes).from_address(hptr)
    hmem = memoryview(hview).cast("B")

    def e2e_pass():
        off, sel, d2h, wrote, pend = 0, 0, 0, 0, []

        def drain():
            nonlocal sel, d2h, wrote
            base = pend.pop(0)
            rr = ctx.collect()
            if not rr.canonical_ok:
                raise RuntimeError("e2e batch rejected: %#x" % rr.violation)
            sel += rr.n_selected
            d2h += ctx.last_d2h_bytes()                         # measured by the library: counters + the runs it copied
            for i in range(rr.n_runs):                          # selected records leave from the host's own copy
                o, ln = rr.runs[i].off, rr.runs[i].len
                wrote += os.write(devnull, hmem[base + o:base + o + ln])
        while off < e2e_bytes:
            n = min(batch, e2e_bytes - off)
            if off + n < e2e_bytes:
                n = lib.find_record_cut_ptr(hptr + off, n)
            if len(pend) == 2:
                drain()
            ctx.submit_raw_from(hptr + off, n, off + n >= e2e_bytes)
            pend.append(off)
            off += n
        while pend:
            drain()
        return sel, d2h, wrote

    e2e_pass()
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    barrier()
    t_c = time.time()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        sel, d2h, wrote = e2e_pass()
    barrier()
    t1 = time.perf_counter()
    windows.append((t_c, time.time()))
    secs = (t1 - t0) / e2e_steps
    out = {"unit": UNIT, "steps": e2e_steps, "batch_bytes": batch, "bytes_per_gpu": int(e2e_bytes), "selected_records_rank0": int(sel),
           "written_bytes_rank0": int(wrote), "numa_node_rank0": numa_node,
           "overlap": "H2D on a copy stream into one of two device buffers while the previous batch's kernels run"}
    # the bound: the same bytes, pinned host -> device, nothing else (all ranks at once: the box's aggregate ceiling)
    bound = None
    if probe:
        chunk = min(e2e_bytes, 4 << 30)
        ctx.copy_to_device(d.data_ptr(), hptr, chunk)
        barrier()
        t0 = time.perf_counter()
        done = 0
        while done < e2e_bytes:
            n = min(chunk, e2e_bytes - done)
            ctx.copy_to_device(d.data_ptr() + done, hptr + done, n)
            done += n
        barrier()
        bound = time.perf_counter() - t0
    os.close(devnull)
    del hmem, hview
    lib.host_free(hptr)
    vmax, vsum = reduce([secs, bound or 0.0], [float(e2e_bytes), float(d2h)])
    out.update({"value": vsum[0] / vmax[0] / 1e9, "h2d_bytes_per_step": int(vsum[0]), "d2h_bytes_per_step": int(vsum[1])})
    if bound:
        out["bound_gbs"] = vsum[0] / vmax[1] / 1e9
        out["bound"] = "pinned host->device cudaMemcpyAsync of the same bytes in the same run, all ranks concurrently (PCIe / host-ingest ceiling)"
        out["frac"] = out["value"] / out["bound_gbs"]
    return out


def cli_leg(args, results):
    """The drop-in binary on a file in tmpfs: wall clock of the whole process (CUDA context creation included) and
    the streaming rate it reports itself (FXTRACT_STATS).  Equality with the reference binary is checked on a
    smaller file of the same workload (the reference needs 3 s per GB); its rate on that file is reported beside."""
    d = scratch_dir()
    out = {}
    env = dict(os.environ, FXTRACT_STATS="1", FXTRACT_PATH="raw")

    def run_cli(argv, capture):
        t0 = time.perf_counter()
        p = subprocess.run(argv, stdout=subprocess.PIPE if capture else subprocess.DEVNULL, stderr=subprocess.PIPE, env=env)
        dt = time.perf_counter() - t0
        if p.returncode != 0:
            raise RuntimeError("fxtract failed: %s" % p.stderr[-300:])
        return dt, p

    try:
        for name in ("c2", "c5v"):
            if name not in results and name != "c2":
                continue
            # ---- equality: product CLI vs reference binary, same small file
            ws = workload_def(name, args.cli_check_records)
            pattern_path, pats = workload_pattern_file(ws, args.cli_records, d)
            fs = write_shards(ws, 0, args.cli_check_records, 1, d, "clis_" + name)[0]
            _, p = run_cli(reference_argv(PRODUCT_CLI, ws, pats, pattern_path, fs), True)
            sha = hashlib.sha256(p.stdout).hexdigest()
            t0 = time.perf_counter()
            r = subprocess.run(reference_argv(REF_BIN, ws, pats, pattern_path, fs), stdout=subprocess.PIPE, stderr=subprocess.DEVNULL)
            ref_dt = time.perf_counter() - t0
            same = hashlib.sha256(r.stdout).hexdigest() == sha
            ref_rate = os.path.getsize(fs) / ref_dt / 1e9
            os.remove(fs)
            if not same:
                raise SystemExit("bench.py: fxtract CLI output differs from oracle/_ref on %s" % name)
            # ---- timing: the big file
            w = workload_def(name, args.cli_records)
            f = write_shards(w, 0, args.cli_records, 1, d, "cli_" + name)[0]
            nbytes = os.path.getsize(f)
            argv = reference_argv(PRODUCT_CLI, w, pats, pattern_path, f)
            best, stats_rate, breakdown = None, None, None
            for rep in range(3):
                dt, p = run_cli(argv, False)
                if rep == 0:
                    continue                                    # first run: file and binary into the page cache
                if best is None or dt < best:
                    best = dt
                    for line in p.stderr.decode(errors="replace").splitlines():
                        if "raw path main thread" in line:
                            breakdown = line.split(":", 1)[1].strip()
                        elif "GB/s" in line:
                            try:
                                stats_rate = float(line.rsplit(",", 1)[1].split()[0])
                            except Exception:
                                pass
            os.remove(f)
            out[name] = {"value": nbytes / best / 1e9, "unit": UNIT, "bytes": nbytes, "records": args.cli_records,
                         "process_seconds": best, "streaming_gbs": stats_rate, "main_thread": breakdown,
                         "reference_process_gbs": ref_rate, "sha256_equal_to_reference": same, "check_records": args.cli_check_records,
                         "what": "fxtract %s <file in tmpfs> > /dev/null; value = file bytes / wall clock of the whole process "
                                 "(CUDA context creation included); streaming_gbs = the rate the binary reports for its read->GPU->write "
                                 "loop; sha256 of the output compared with the reference binary on the first check_records records"
                                 % " ".join(w["cli"] + ["-S"] + (["-f", "patterns"] if w["mode"] else ["<20-mer>"]))}
    finally:
        shutil.rmtree(d, ignore_errors=True)
    return out


_RESULT_FD = None


def claim_stdout():
    """Keep the process's stdout for the JSON line alone: native libraries write to fd 1 too (NCCL prints its
    version banner there), so fd 1 is pointed at stderr and the line goes to a private copy of the old fd 1."""
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def emit_line(obj):
    data = (json.dumps(obj) + "\n").encode()
    sys.stdout.flush()
    os.write(_RESULT_FD if _RESULT_FD is not None else 1, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", help="headline workload (value / roofline / e2e)")
    ap.add_argument("--also", default="c3,c4,c5v,c5c", help="further workloads measured into `workloads` ('' = none)")
    ap.add_argument("--also-steps", type=int, default=5)
    ap.add_argument("--records", type=int, default=100_000_000, help="records per GPU")
    ap.add_argument("--batch-mib", type=int, default=256, help="e2e staging batch size")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--e2e-gib", type=float, default=0, help="cap the e2e host shard (0 = whole shard if RAM allows)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cli", action="store_true")
    ap.add_argument("--no-verify", action="store_true", help="skip the oracle/_ref comparison and the slice check (experiments only)")
    ap.add_argument("--no-numa-bind", action="store_true", help="do not pin the rank to its GPU's NUMA node")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-records", type=int, default=4_000_000, help="records in the verification / cpu_baseline sample")
    ap.add_argument("--cli-records", type=int, default=48_000_000, help="records in the e2e_cli file (timing)")
    ap.add_argument("--cli-check-records", type=int, default=4_000_000, help="records in the file the CLI output is compared with the reference on")
    ap.add_argument("--ref-records", type=int, default=4_000_000, help="records per step of --impl reference")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
