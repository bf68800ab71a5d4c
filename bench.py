#!/usr/bin/env python
"""bench.py -- FASTQ GB/s scanned by the B200 search path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4|c5v|c5c] [--records R]
  python bench.py --impl reference ...      # the reference CPU binary on the host cores

A "step" is one pass of the hot path (fused raw-FASTQ scan: record index + canonical
predicate + matcher, then count / order-preserving compaction) over one rank's whole
synthetic shard, resident in HBM.  Default workload = BASELINE.json configs[1]: one 20-mer
over 100 M x 150 bp synthetic FASTQ (~34.6 GB) per GPU; with N GPUs every rank holds its own
contiguous 100 M-record range of the stream (weak scaling, no data-path collective; only the
match count is summed).  `value` = bytes scanned by all ranks / max-over-ranks device time.
`e2e` = the same work through the C ABI from HOST (pinned) buffers: record-aligned batches cut
on the host, cudaMemcpyAsync H2D, kernels, results D2H, selected records written out.

Only the `cpu_baseline` leg and `--impl reference` execute anything under oracle/ (the
reference binary built from the unmodified sources, oracle/_ref/fxtract); the measured GPU
path never does.
"""
import argparse
import ctypes as C
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
PATTERN_RECORD = 1234          # the single 20-mer of c2 is copied from this global record


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------- workloads
def workload_def(name, records):
    """Returns dict(mode, target, invert, count_only, flags (reference CLI), plant spec fields, npat)."""
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
                 desc="-f 1000 20-mers %s over synthetic FASTQ" % ("-v" if name == "c5v" else "-c"),
                 cli=["-v"] if name == "c5v" else ["-c"])
    else:
        raise SystemExit("unknown workload %s" % name)
    return w


def make_spec(fx, w, first, n):
    return fx.SynthSpec(seed=w["seed"], first_record=first, n_records=n, read_len=w["read_len"],
                        plant_mode=w["plant_mode"], pattern_seed=w["pattern_seed"], n_patterns=w["n_patterns"],
                        pattern_len=w["pattern_len"], plant_per_million=w["plant_per_million"],
                        plant_revcomp=w["plant_revcomp"])


COMP = bytes.maketrans(b"ACGT", b"TGCA")


def workload_patterns(fx, w, total_records):
    """Pattern list exactly as the host CLI would hand it over: -r expansion, de-dup, std::map order."""
    from fxtract_b200 import lib
    if w["mode"] == 0:
        one = make_spec(fx, w, PATTERN_RECORD, 1)
        rec = lib.synth_host(one)
        seq = rec.split(b"\n")[1]
        return [seq[37:57]]
    if w["mode"] == 1:
        spec = make_spec(fx, w, 0, 1)
        pats = set(lib.synth_pattern(spec, p) for p in range(w["n_patterns"]))
        if w["revcomp"]:
            pats |= set(p.translate(COMP)[::-1] for p in pats)
        return sorted(pats)
    # exact ids: every (total/n_ids)-th read name + 1 % absent ids
    spec = make_spec(fx, w, 0, 1)
    n_ids = min(w["n_ids"], total_records)
    stride = max(1, total_records // n_ids)
    ids = set(lib.synth_name(spec, i * stride) for i in range(n_ids))
    ids |= set(b"SYN:1:FC:9:0000:00000:%d" % i for i in range(n_ids // 100))
    return sorted(ids)


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


# ---------------------------------------------------------------- CPU reference timing
def _sample_dir():
    for d in ("/dev/shm", tempfile.gettempdir()):
        try:
            if os.path.isdir(d) and shutil.disk_usage(d).free > (4 << 30):
                return tempfile.mkdtemp(prefix="fxbench_", dir=d)
        except Exception:
            pass
    return tempfile.mkdtemp(prefix="fxbench_")


def write_sample(fx, w, n_records, directory):
    from fxtract_b200 import lib
    spec = make_spec(fx, w, 0, n_records)
    data = lib.synth_host(spec)
    path = os.path.join(directory, "sample.fq")
    with open(path, "wb") as f:
        f.write(data)
    return path, len(data)


def reference_argv(binary, w, pats, directory, sample, base_pats=None):
    argv = [binary] + list(w["cli"])
    if w["mode"] == 0:
        argv += ["-S", pats[0].decode(), sample]
    else:
        pf = os.path.join(directory, "patterns.txt")
        if not os.path.exists(pf):
            with open(pf, "wb") as f:     # under -r the CLI adds the reverse complements itself
                f.write(b"\n".join(base_pats if base_pats is not None else pats) + b"\n")
        argv += ["-S", "-f", pf, sample]
    return argv


def time_reference(argv, nproc):
    """nproc concurrent reference processes (the binary is single-threaded), stdout discarded."""
    t0 = time.perf_counter()
    procs = [subprocess.Popen(argv, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL) for _ in range(nproc)]
    rcs = [p.wait() for p in procs]
    dt = time.perf_counter() - t0
    if any(rcs):
        raise RuntimeError("reference binary failed: %r -> %r" % (argv, rcs))
    return dt


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import fxtract_b200 as fx
    w = workload_def(args.workload, args.records)
    base = {"metric": METRIC, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": w["desc"], "records_per_gpu": w["records"], "read_len": 150}}
    if not os.path.exists(REF_BIN):
        base["unavailable"] = "oracle/_ref/fxtract not built (reference sources absent at build time)"
        print(json.dumps(base))
        return 0
    cores = host_cores()
    nrec = args.ref_records
    d = _sample_dir()
    try:
        sample, nbytes = write_sample(fx, w, nrec, d)
        pats = workload_patterns(fx, w, nrec)
        argv = reference_argv(REF_BIN, w, pats, d, sample)
        for _ in range(args.warmup):
            time_reference(argv, cores)
        times = [time_reference(argv, cores) for _ in range(args.steps)]
        one = time_reference(argv, 1)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    total = sum(times)
    value = cores * nbytes * args.steps / total / 1e9
    smp = ("%d concurrent single-threaded reference processes (oracle/_ref/fxtract, unmodified sources, g++ -O2), "
           "each scanning the same %d-record / %.0f MB synthetic sample from page cache, stdout to /dev/null"
           % (cores, nrec, nbytes / 1e6))
    base.update({"value": value, "ms_per_step": 1e3 * total / args.steps,
                 "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": smp,
                                  "single_process_value": nbytes / one / 1e9},
                 "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0})
    print(json.dumps(base))
    return 0


def cpu_baseline_leg(fx, w, budget_records):
    """Bounded CPU timing of the reference binary on this box (rank 0, N=1 only)."""
    if not os.path.exists(REF_BIN):
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref/fxtract not built"}
    cores = host_cores()
    d = _sample_dir()
    try:
        sample, nbytes = write_sample(fx, w, budget_records, d)
        pats = workload_patterns(fx, w, budget_records)
        argv = reference_argv(REF_BIN, w, pats, d, sample)
        time_reference(argv, 1)                       # warm page cache
        one = min(time_reference(argv, 1) for _ in range(2))
        allc = min(time_reference(argv, cores) for _ in range(2))
        o0 = None
        if os.path.exists(REF_BIN_O0):
            argv0 = [REF_BIN_O0] + argv[1:]
            o0 = time_reference(argv0, 1)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    return {"value": cores * nbytes / allc / 1e9, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": ("reference binary (unmodified sources, -O2) on a %d-record / %.0f MB slice of the workload, page "
                       "cache warm, stdout to /dev/null; value = %d concurrent single-threaded processes (the reference "
                       "has no threads); single_process_value = one process" % (budget_records, nbytes / 1e6, cores)),
            "single_process_value": nbytes / one / 1e9,
            "single_process_O0_value": (nbytes / o0 / 1e9) if o0 else None}


# ---------------------------------------------------------------- GPU arm
def bind_to_gpu_numa_node(torch, local):
    """Pin this rank to the CPUs of the NUMA node its GPU hangs off, before any pinned host buffer is allocated:
    the e2e leg streams the whole shard from pinned memory, and first-touch placement on the far socket sends
    every H2D byte across the inter-socket link.  Returns the node, or None when the box does not say."""
    try:
        pr = torch.cuda.get_device_properties(local)
        bus = None
        if hasattr(pr, "pci_bus_id"):
            bus = "%04x:%02x:%02x.0" % (getattr(pr, "pci_domain_id", 0), pr.pci_bus_id, getattr(pr, "pci_device_id", 0))
        else:
            import pynvml
            pynvml.nvmlInit()
            for i in range(pynvml.nvmlDeviceGetCount()):
                h = pynvml.nvmlDeviceGetHandleByIndex(i)
                u = pynvml.nvmlDeviceGetUUID(h)
                u = u.decode() if isinstance(u, bytes) else u
                if u.replace("GPU-", "") == str(pr.uuid).replace("GPU-", ""):
                    b = pynvml.nvmlDeviceGetPciInfo(h).busId
                    b = b.decode() if isinstance(b, bytes) else b
                    bus = b.lower()[-12:]
        if bus is None:
            return None
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

    w = workload_def(args.workload, args.records)
    nrec = w["records"]
    first = rank * nrec
    spec = make_spec(fx, w, first, nrec)
    L = lib.load_library()
    nbytes = L.fxg_synth_bytes(C.byref(spec))
    stream = torch.cuda.Stream()
    pats = workload_patterns(fx, w, nrec * world)
    sampler = ClockSampler(local) if rank == 0 else None
    windows = []

    with torch.cuda.stream(stream):
        d = torch.empty(nbytes + 64, dtype=torch.uint8, device="cuda")
        ctx = fx.Context(w["mode"], target=w["target"], invert=bool(w["invert"]), count_only=bool(w["count_only"]),
                         batch_bytes=args.batch_mib << 20, stream=stream.cuda_stream, device=local)
        ctx.set_patterns(pats)
        ctx.synth_device(spec, d.data_ptr())
        ctx.set_profiling(True)

        def step():
            ctx.scan_device_async(d.data_ptr(), nbytes)
            r = ctx.scan_device_finish()
            if not r.canonical_ok or r.n_records != nrec:
                raise RuntimeError("scan rejected the synthetic shard: violation=%#x records=%d" % (r.violation, r.n_records))
            return r

        for _ in range(max(args.warmup, 3) if args.warmup >= 0 else 0):
            r = step()
        launches_per_step = ctx.last_launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        scan_ms, pipe_ms = [], []
        barrier()
        t_a = time.time()
        ev0.record(stream)
        for _ in range(args.steps):
            r = step()
            a, b = ctx.last_kernel_ms()
            scan_ms.append(a)
            pipe_ms.append(b)
        ev1.record(stream)
        barrier()
        t_b = time.time()
        windows.append((t_a, t_b))
        dev_ms = ev0.elapsed_time(ev1)
        n_selected, n_records = int(r.n_selected), int(r.n_records)
        n_runs = int(r.n_runs)

        # ---------------- end to end from host buffers through the C ABI
        e2e = None
        if not args.no_e2e:
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
                off, sel, d2h, pend = 0, 0, 0, []

                def drain():
                    nonlocal sel, d2h
                    base = pend.pop(0)
                    rr = ctx.collect()
                    if not rr.canonical_ok:
                        raise RuntimeError("e2e batch rejected: %#x" % rr.violation)
                    sel += rr.n_selected
                    d2h += 40 + 16 * (0 if w["count_only"] else 1) + (0 if w["count_only"] else 16 * 65536)
                    for i in range(rr.n_runs):                      # selected records leave from the host's own copy
                        o, ln = rr.runs[i].off, rr.runs[i].len
                        os.write(devnull, hmem[base + o:base + o + ln])
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
                return sel, d2h

            for _ in range(1):
                e2e_pass()
            e2e_steps = max(1, min(args.steps, args.e2e_steps))
            barrier()
            t_c = time.time()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                sel, d2h = e2e_pass()
            barrier()
            t1 = time.perf_counter()
            windows.append((t_c, time.time()))
            e2e_s = (t1 - t0) / e2e_steps
            os.close(devnull)
            lib.host_free(hptr)
            e2e = dict(seconds=e2e_s, bytes=e2e_bytes, d2h=d2h, selected=int(sel), steps=e2e_steps)
        ctx.close()

    # ---------------- reduce over ranks
    vals = torch.tensor([dev_ms, e2e["seconds"] if e2e else 0.0], dtype=torch.float64, device="cuda")
    sums = torch.tensor([float(nbytes), float(e2e["bytes"]) if e2e else 0.0, float(n_selected), float(n_records)],
                        dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    vals, sums = vals.tolist(), sums.tolist()
    if sampler:
        sampler.stop()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0

    peak, peak_src = peaks()
    ms_per_step = vals[0] / args.steps
    value = sums[0] / (ms_per_step * 1e-3) / 1e9
    scan_avg_ms = sum(scan_ms) / len(scan_ms)
    algo_bytes = nbytes + n_records                      # N_in + 1 flag per record
    achieved = algo_bytes / (scan_avg_ms * 1e-3) / 1e9
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
           "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "u8", "data": "synthetic",
           "config": {"workload": w["desc"], "records_per_gpu": nrec, "read_len": 150, "bytes_per_gpu": nbytes,
                      "sharding": "contiguous record ranges, one per GPU; count summed, no data-path collective",
                      "l2_policy": "input per step (%.1f GB) is far larger than the 126 MB L2" % (nbytes / 1e9),
                      "selected_records": int(sums[2]), "records": int(sums[3]), "runs_rank0": n_runs},
           "gpu_launches": launches_per_step * args.steps,
           "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                        "traffic": None, "kernel": "fxg::scan_tiles_kernel (fxg_scan.cu; <9,0> literal, <0,0> index only)", "kernel_ms": scan_avg_ms,
                        "pipeline_ms": sum(pipe_ms) / len(pipe_ms), "algorithmic_bytes": algo_bytes, "peak_source": peak_src},
           "clocks": sampler.summary(windows) if sampler else None}
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_file):
        try:
            t = json.load(open(traffic_file))
            per_byte = t.get(args.workload, {}).get("dram_bytes_per_input_byte")
            if per_byte:
                out["roofline"]["traffic"] = per_byte * nbytes
                out["roofline"]["traffic_source"] = t[args.workload].get("source")
        except Exception:
            pass
    if e2e:
        out["e2e"] = {"value": sums[1] / vals[1] / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(sums[1]),
                      "d2h_bytes_per_step": int(e2e["d2h"]) * world, "steps": e2e["steps"], "batch_bytes": args.batch_mib << 20,
                      "bytes_per_gpu": int(e2e["bytes"]), "selected_records_rank0": e2e["selected"],
                      "numa_node_rank0": numa_node}
    if world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline_leg(fx, w, args.cpu_records)
    emit_line(out)
    return 0


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
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--records", type=int, default=100_000_000, help="records per GPU")
    ap.add_argument("--batch-mib", type=int, default=256, help="e2e staging batch size")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--e2e-gib", type=float, default=0, help="cap the e2e host shard (0 = whole shard if RAM allows)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-numa-bind", action="store_true", help="do not pin the rank to its GPU's NUMA node")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-records", type=int, default=4_000_000, help="records in the cpu_baseline sample")
    ap.add_argument("--ref-records", type=int, default=1_000_000, help="records per step of --impl reference")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
