#!/usr/bin/env python
"""bench.py -- LiME classification hot path on B200: eBWT symbols/s.

    python bench.py --gpus N --steps K --warmup W            (ours; torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K ...  (the reference's CPU pipeline)

Workload (BASELINE.json configs[1]): single-end synthetic, 2 M reads x 100 bp + 100 genomes,
N = 402 M eBWT symbols per strand, two strands (F, RC), EBWT=1, alpha=16, beta=0.25, TaxRank=1.
One step = the whole job: per strand cluster detection (step 1), similarity accumulation and
clusterChoose (step 2), then Classify over both strands (step 3).  At N > 1 GPUs the position
range of every strand is sharded contiguously (weak scaling: every GPU holds 402 M symbols per
strand and the job has N x 2 M reads).

value  = symbols of all strands / step time, inputs resident in HBM (CUDA events).
e2e    = the same through the C ABI with HOST buffers: pinned host -> device copies of
         lcp/da/ebwt and the device -> host copy of the assignments inside the timed region.
The oracle / reference binaries are used only for cpu_baseline and --impl reference.
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOAD = dict(name="cfg2 single-end synthetic: 2M reads x 100bp + 100 genomes, 402M symbols/strand, "
                     "2 strands (F,RC), EBWT=1, alpha=16, beta=0.25, TaxRank=1",
                n=402_000_000, n_reads=2_000_000, n_refs=100, read_len=100, alpha=16, beta=0.25,
                ebwt=1, n_strands=2, tax_rank=1, seed=20261104 + 2)
CPU_SAMPLE_N = 1 << 28          # positions per strand of the CPU-timed prefix sample (~4.5 s per run on 16 cores)
METRIC = "ebwt_symbols_per_s"
UNIT = "symbols/s"


def _small(w, n, n_reads):
    w = dict(w)
    w.update(n=n, n_reads=n_reads, name=f"REDUCED test shape n={n} n_reads={n_reads} (not the benchmark)")
    return w


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0, period_ms=50):
        self.rows, self.proc, self.thr = [], None, None
        self.cmd = ["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                    "-lms", str(period_ms)]

    def start(self):
        try:
            self.proc = subprocess.Popen(self.cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        def pump():
            for line in self.proc.stdout:
                self.rows.append((time.time(), line.strip()))
        self.thr = threading.Thread(target=pump, daemon=True)
        self.thr.start()

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        if self.thr:
            self.thr.join(timeout=2)

    def summary(self, t0, t1):
        sm, mx, reasons, pw = [], [], set(), []
        for t, line in self.rows:
            if t < t0 or t > t1 + 0.05:
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw)}


def _extend_for_clocks(step, t0, min_seconds=0.7):
    """The timed region of this workload lasts tens of milliseconds, shorter than nvidia-smi's
    sampling period: keep running the same step (untimed) until the sampler has seen min_seconds
    of this load.  Returns the end of the sampled window."""
    while time.time() - t0 < min_seconds:
        step()
    return time.time()


# ------------------------------------------------------------------------------------------------
# the reference's CPU pipeline on a bounded sample (oracle/_ref binaries; the one place bench.py
# executes anything under oracle/)
# ------------------------------------------------------------------------------------------------
def _lineage_file(path, n_refs):
    from lime_b200.pipeline import synthetic_lineage
    rows, v, m = synthetic_lineage(n_refs)
    with open(path, "w") as f:
        f.write("Acc;Species_TaxID;Genus_TaxID;Family_TaxID;Order_TaxID;Class_TaxID;Phylum_TaxID\n")
        for r in rows:
            f.write(";".join(str(x) for x in r) + "\n")
    return v, m


def _scratch_dir():
    for base in ("/dev/shm", tempfile.gettempdir()):
        try:
            if os.path.isdir(base) and shutil.disk_usage(base).free > (8 << 30):
                return tempfile.mkdtemp(prefix="lime_bench_", dir=base)
        except OSError:
            pass
    return tempfile.mkdtemp(prefix="lime_bench_")


class CpuReference:
    """Runs the unmodified reference (ClusterLCP, ClusterBWT_DA, Classify from oracle/_ref) with
    the schedule of LiME_paired.sh: ClusterLCP of all strands concurrently (:44-53), ClusterBWT_DA
    strand after strand (:62-68), one Classify (:79); all with threads = host cores."""

    def __init__(self, w, sample_n):
        from oracle.oracle import Reference
        self.ref = Reference()
        self.kind = "reference" if self.ref.available() else "port"
        self.w, self.sample_n = w, min(sample_n, w["n"])
        self.cores = os.cpu_count() or 1
        self.dir = _scratch_dir()
        self.fastas = []

    def write_inputs(self, strands_host):
        """strands_host: list of (ebwt u8, lcp u32, da u32) numpy prefixes."""
        from lime_b200 import datagen
        for s, (e, l, d) in enumerate(strands_host):
            fasta = os.path.join(self.dir, f"strand{s}+Refs.fasta")
            datagen.write_strand(fasta, e, l, d)
            self.fastas.append(fasta)
        self.lineage = os.path.join(self.dir, "lineage.csv")
        self.v, self.m = _lineage_file(self.lineage, self.w["n_refs"])
        self.host = strands_host

    def step(self):
        """One pass of the reference pipeline over the sample; returns wall seconds."""
        w = self.w
        t0 = time.perf_counter()
        if self.kind == "reference":
            exe = lambda n: os.path.join(self.ref.dir, n)          # noqa: E731
            procs = [subprocess.Popen([exe("ClusterLCP"), f, str(w["n_reads"]), str(w["n_refs"]), str(w["alpha"]),
                                       str(self.cores)], cwd=self.dir, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                     for f in self.fastas]
            for p in procs:
                assert p.wait() == 0, "reference ClusterLCP failed"
            for f in self.fastas:
                r = subprocess.run([exe("ClusterBWT_DA_ebwt1" if w["ebwt"] else "ClusterBWT_DA_ebwt0"), f,
                                    str(w["read_len"]), str(w["beta"]), str(self.cores)], cwd=self.dir,
                                   stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True)
                assert r.returncode == 0, r.stderr
            out = os.path.join(self.dir, "out.csv")
            r = subprocess.run([exe("Classify_higher0"), str(len(self.fastas))] + [f + ".res" for f in self.fastas] +
                               [str(w["n_reads"]), str(w["n_refs"]), out, self.lineage, str(w["tax_rank"]),
                                str(self.cores)], cwd=self.dir, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True)
            assert r.returncode == 0, r.stderr
        else:                                                    # scalar C port (1 core)
            from oracle.oracle import Oracle
            from lime_b200.pipeline import norm_of
            orc = Oracle()
            bins, poss = [], []
            for (e, l, d) in self.host:
                cl, _ = orc.cluster_detect(l, d, w["n_reads"], w["alpha"])
                sim = orc.similarity(cl, d, e if w["ebwt"] else None, w["n_reads"], w["n_refs"])
                b, p = orc.choose(sim, norm_of(w["read_len"], w["alpha"]), np.float32(w["beta"]))
                bins.append(b); poss.append(p)
            orc.classify(bins, poss, w["n_refs"], self.v, None, w["tax_rank"], False)
        return time.perf_counter() - t0

    def describe(self):
        return (f"prefix of {self.sample_n} positions per strand x {len(self.fastas)} strands of the same workload "
                f"(all {self.w['n_reads']} reads, {self.w['n_refs']} genomes), files in {os.path.dirname(self.dir) or self.dir}, "
                f"page cache warm; ClusterLCP x{len(self.fastas)} concurrent, ClusterBWT_DA x{len(self.fastas)} sequential, "
                f"Classify {len(self.fastas)}; threads={self.cores if self.kind == 'reference' else 1}")

    def close(self):
        shutil.rmtree(self.dir, ignore_errors=True)


def _host_sample(w, sample_n, device):
    """The first sample_n positions of every strand, as numpy arrays in the file formats."""
    from lime_b200 import datagen
    out = []
    for s in range(w["n_strands"]):
        e, l, d = datagen.synth_arrays(w["n"], w["n_reads"], w["n_refs"], w["read_len"], seed=w["seed"] + 1000 * s,
                                       device=device, hi=sample_n)
        out.append((e.cpu().numpy(), l.cpu().numpy().view(np.uint32), d.cpu().numpy().view(np.uint32)))
    return out


def run_reference_arm(args, w):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    cpu = CpuReference(w, args.cpu_sample)
    try:
        cpu.write_inputs(_host_sample(w, cpu.sample_n, dev))
        for _ in range(args.warmup):
            cpu.step()
        times = [cpu.step() for _ in range(args.steps)]
        total = sum(times)
        symbols = cpu.sample_n * w["n_strands"]
        value = symbols * args.steps / total
        line = {
            "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": w["name"], "sample": cpu.describe()},
            "reads_per_s": w["n_reads"] * (cpu.sample_n / w["n"]) * args.steps / total,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cpu.cores if cpu.kind == "reference" else 1,
                             "kind": cpu.kind, "sample": cpu.describe()},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
        }
        print(json.dumps(line), flush=True)
    finally:
        cpu.close()


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def _alg_bytes(kname, st, w):
    """Algorithmic bytes of one launch of kernel family `kname` for a strand with stats st
    (DESIGN.md 'kernels')."""
    E = 1 if w["ebwt"] else 0
    n, C, P, nr, ng = st["n"], st["n_clusters"], st["cluster_positions"], st["n_reads_total"], w["n_refs"]
    if kname == "k1_detect":
        return 8 * n + 16 * C
    if kname in ("k2_warp", "k2_block"):
        return (4 + E) * P + 16 * C          # da (+ebwt) of the clustered positions + the records
    if kname == "k12_fused":
        return (8 + E) * n
    if kname == "k3_count":
        return nr * ng + 8 * nr
    if kname == "k3_emit":
        return nr * ng + 8 * st["n_records"]
    if kname == "k3_classify":
        return 8 * st["n_records"] + 8 * nr + 9 * nr
    return 0


def run_ours(args, w):
    import torch
    import torch.distributed as dist
    from lime_b200 import datagen
    from lime_b200.api import LimeContext, MEM_DEVICE, MEM_HOST
    from lime_b200.pipeline import run_job, synthetic_lineage

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        return run_ours_multi(args, w, world, rank, local)

    dev = torch.device("cuda", local)
    ctx = LimeContext(local)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    rows, v, m = synthetic_lineage(w["n_refs"])

    # ---- synthetic inputs, resident in HBM ----------------------------------------------------
    strands_dev, keep = [], []
    for s in range(w["n_strands"]):
        e, l, d = datagen.synth_arrays(w["n"], w["n_reads"], w["n_refs"], w["read_len"], seed=w["seed"] + 1000 * s,
                                       device=dev)
        keep.append((e, l, d))
        strands_dev.append({"ebwt": e.data_ptr(), "lcp": l.data_ptr(), "da": d.data_ptr(), "n": w["n"]})
    torch.cuda.synchronize()

    def step_resident():
        return run_job(ctx, strands_dev, w["n_reads"], w["n_refs"], w["read_len"], w["alpha"], w["beta"], w["ebwt"],
                       v, None, w["tax_rank"], False, mem=MEM_DEVICE, fused=not args.unfused)

    # ---- warm-up (also gathers the per-strand statistics the roofline needs) -------------------
    for _ in range(max(args.warmup, 1)):
        out, stats = step_resident()
    strand_stats = []
    for s in range(w["n_strands"]):
        # cluster statistics of strand s (outside the timed region)
        ctx.load(strands_dev[s]["lcp"], strands_dev[s]["da"], strands_dev[s]["ebwt"], w["n_reads"], w["n_refs"],
                 mem=MEM_DEVICE, n_total=w["n"])
        n_cl, max_len = ctx.cluster_detect(w["alpha"])
        cl = ctx.clusters()
        strand_stats.append({"n": w["n"], "n_clusters": int(n_cl), "max_len": int(max_len),
                             "cluster_positions": int(cl["len"].sum()), "n_records": int(stats[s][2]),
                             "n_reads_total": w["n_reads"]})
        del cl
    out, stats = step_resident()          # leave the slots in a consistent state
    out = tuple(x.copy() for x in out)    # classify() returns views of context-owned pinned buffers
    kinds = {chr(t): int((out[0] == t).sum()) for t in np.unique(out[0])}

    # ---- timed region: K steps, resident inputs -------------------------------------------------
    clocks = ClockSampler(local)
    clocks.start()
    time.sleep(0.3)
    ctx.prof_enable(True)
    ctx.prof_reset()
    l0 = ctx.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.time()
    ev0.record(stream)
    for _ in range(args.steps):
        step_resident()
    ev1.record(stream)
    torch.cuda.synchronize()
    t1 = time.time()
    ms_total = ev0.elapsed_time(ev1)
    launches = ctx.launch_count() - l0
    prof = ctx.prof()
    ctx.prof_enable(False)
    t1x = _extend_for_clocks(lambda: (step_resident(), torch.cuda.synchronize()), t0)
    clk = clocks.summary(t0, t1x)
    clk["window"] = "timed region + untimed repeats of the same step up to 0.7 s (nvidia-smi -lms 50)"

    ms_step = ms_total / args.steps
    symbols = w["n"] * w["n_strands"]
    value = symbols / (ms_step * 1e-3)

    # ---- roofline of the dominant kernel ----------------------------------------------------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_gbs, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback")
    kern = {k: x for k, x in prof.items() if x["launches"]}
    kernel_ms_total = sum(x["ms"] for x in kern.values())
    top = max(kern, key=lambda k: kern[k]["ms"])
    per_strand_launches = kern[top]["launches"] / (args.steps * (1 if top == "k3_classify" else w["n_strands"]))
    if top == "k3_classify":
        agg = dict(strand_stats[0]); agg["n_records"] = sum(s["n_records"] for s in strand_stats)
        alg = _alg_bytes(top, agg, w) * args.steps
    else:
        alg = sum(_alg_bytes(top, s, w) for s in strand_stats) * args.steps
    avg_launch_ms = kern[top]["ms"] / kern[top]["launches"]
    alg_per_launch = alg / kern[top]["launches"]
    achieved = alg_per_launch / (avg_launch_ms * 1e-3) / 1e9
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tr.get(top, {}).get("dram_bytes_per_launch")
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak_gbs, "unit": "GB/s",
                "frac": achieved / peak_gbs, "traffic": traffic, "peak_source": peak_src,
                "alg_bytes_per_launch": alg_per_launch, "avg_launch_ms": avg_launch_ms,
                "launches_per_strand": per_strand_launches,
                "kernel_share_of_step": kern[top]["ms"] / ms_total,
                "kernels_ms_per_step": {k: x["ms"] / args.steps for k, x in kern.items()},
                "kernel_time_fraction_of_step": kernel_ms_total / ms_total,
                "job_alg_bytes_per_step": sum((8 + w["ebwt"]) * s["n"] for s in strand_stats),
                "job_frac_of_peak": sum((8 + w["ebwt"]) * s["n"] for s in strand_stats) / (ms_step * 1e-3) / 1e9 / peak_gbs}

    # ---- e2e: host buffers through the C ABI ----------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        strands_host, pins = [], []
        for (e, l, d) in keep:
            pe = torch.empty(e.shape, dtype=e.dtype, pin_memory=True); pe.copy_(e)
            pl = torch.empty(l.shape, dtype=l.dtype, pin_memory=True); pl.copy_(l)
            pd = torch.empty(d.shape, dtype=d.dtype, pin_memory=True); pd.copy_(d)
            pins.append((pe, pl, pd))
            strands_host.append({"ebwt": pe.numpy(), "lcp": pl.numpy().view(np.uint32), "da": pd.numpy().view(np.uint32),
                                 "n": w["n"]})
        torch.cuda.synchronize()

        def step_host():
            return run_job(ctx, strands_host, w["n_reads"], w["n_refs"], w["read_len"], w["alpha"], w["beta"],
                           w["ebwt"], v, None, w["tax_rank"], False, mem=MEM_HOST, fused=not args.unfused)
        h2d = sum((8 + w["ebwt"]) * w["n"] for _ in range(w["n_strands"])) + 4 * w["n_refs"]
        d2h = 9 * w["n_reads"]
        e2e_steps = max(1, min(args.steps, 5))

        def time_host(note, h2d_bytes, check_counts=False):
            out_h, st_h = step_host()
            for a, b in zip(out, out_h):
                assert a.tobytes() == b.tobytes(), "host-buffer path disagrees with the resident path"
            if check_counts:
                assert [tuple(x) for x in st_h] == [tuple(x) for x in stats], "per-strand counts disagree with the resident path"
            torch.cuda.synchronize()
            ev0.record(stream)
            tw0 = time.perf_counter()
            for _ in range(e2e_steps):
                step_host()
            ev1.record(stream)
            torch.cuda.synchronize()
            tw1 = time.perf_counter()
            e_ms = max(ev0.elapsed_time(ev1), (tw1 - tw0) * 1e3) / e2e_steps
            return {"value": symbols / (e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e_ms, "steps": e2e_steps,
                    "reads_per_s": w["n_reads"] / (e_ms * 1e-3), "note": note}
        e2e = time_host("lime_load(LIME_MEM_HOST) from pinned host arrays per strand + assignments copied back; "
                        "PCIe-bound", h2d)

    # ---- CPU baseline beside it (rank 0, N = 1 only) ----------------------------------------------------
    cpu_baseline = None
    if not args.no_cpu:
        cpu = CpuReference(w, args.cpu_sample)
        try:
            sample = []
            for (e, l, d) in keep:
                k = cpu.sample_n
                sample.append((e[:k].cpu().numpy(), l[:k].cpu().numpy().view(np.uint32), d[:k].cpu().numpy().view(np.uint32)))
            cpu.write_inputs(sample)
            cpu.step()                                   # warms the page cache
            t = cpu.step()
            cpu_baseline = {"value": cpu.sample_n * w["n_strands"] / t, "unit": UNIT,
                            "cores": cpu.cores if cpu.kind == "reference" else 1, "kind": cpu.kind,
                            "sample": cpu.describe(), "seconds": t}
        finally:
            cpu.close()
    clocks.stop()

    # ---- e2e again with the narrowed lcp transfer (lime_set_alpha_limit): the same host arrays, lcp crosses
    # PCIe as min(lcp, 255) bytes (narrowed on host threads inside the timed region) and is widened on
    # the device.  Same call sequence and results (checked against the resident path first); the faster
    # verified mode is the e2e figure, both are reported.  Last GPU work of the run, and a failure
    # here only loses this leg.
    if e2e is not None and w["alpha"] <= 255 and not args.no_narrow:
        plain = dict(e2e)
        try:
            ctx.set_alpha_limit(w["alpha"])
            h2d_n = sum((5 + w["ebwt"]) * w["n"] for _ in range(w["n_strands"])) + 4 * w["n_refs"]
            nar = time_host("lime_set_alpha_limit(alpha) + lime_load(LIME_MEM_HOST): lcp narrowed to bytes by %d host "
                            "threads inside the timed region, widened on the device; pinned host arrays; assignments "
                            "copied back" % max(1, ctx.lib.lime_host_threads() - 1), h2d_n, check_counts=True)
            ctx.set_alpha_limit(0)
            if nar["value"] > plain["value"]:
                e2e = nar
            e2e["modes"] = {"plain": {k: plain[k] for k in ("value", "ms_per_step", "h2d_bytes_per_step")},
                            "narrow_lcp": {k: nar[k] for k in ("value", "ms_per_step", "h2d_bytes_per_step")}}
        except BaseException as ex:                      # noqa: BLE001 -- report, keep the plain figure
            try:
                ctx.set_alpha_limit(0)
            except Exception:
                pass
            e2e = plain
            e2e["modes"] = {"plain": {k: plain[k] for k in ("value", "ms_per_step", "h2d_bytes_per_step")},
                            "narrow_lcp": {"error": "%s: %s" % (type(ex).__name__, ex)}}
    if e2e is not None:
        del pins, strands_host

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 1),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": w["name"], "l2": "inputs (3.6 GB per strand) are larger than L2; no flush needed",
                   "parallelism": "1 GPU", "symbols_per_step": symbols, "reads_per_step": w["n_reads"],
                   "strand_stats": strand_stats, "assignments": kinds},
        "reads_per_s": w["n_reads"] / (ms_step * 1e-3),
        "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "gpu_launches": int(launches),
        "clocks": clk,
    }
    print(json.dumps(line), flush=True)
    ctx.close()


class _SyncTimer:
    """Per-phase wall timer with a device synchronise on both sides (diagnostic step only)."""

    def __init__(self, torch):
        self.torch, self.ms = torch, {}

    def __call__(self, name):
        outer = self

        class T:
            def __enter__(self_inner):
                outer.torch.cuda.synchronize()
                self_inner.t = time.perf_counter()

            def __exit__(self_inner, *a):
                outer.torch.cuda.synchronize()
                outer.ms[name] = outer.ms.get(name, 0.0) + (time.perf_counter() - self_inner.t) * 1e3
                return False
        return T()


class _HostTimer:
    """Per-phase host wall time WITHOUT device synchronisation (where the host thread spends a step)."""

    def __init__(self):
        self.ms = {}

    def __call__(self, name):
        outer = self

        class T:
            def __enter__(self_inner):
                self_inner.t = time.perf_counter()

            def __exit__(self_inner, *a):
                outer.ms[name] = outer.ms.get(name, 0.0) + (time.perf_counter() - self_inner.t) * 1e3
                return False
        return T()


def run_ours_multi(args, w, world, rank, local):
    """N > 1: weak scaling.  Every rank holds one 402 M-position shard of every strand of an
    N x 402 M-position job with N x 2 M reads; see lime_b200/multi.py for the exchange."""
    import torch
    import torch.distributed as dist
    from lime_b200 import datagen, multi
    from lime_b200.api import LimeContext
    from lime_b200.pipeline import synthetic_lineage

    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ctx = LimeContext(local)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    S, HALO = w["n"], 8192
    n_reads = w["n_reads"] * world
    read_frac = w["n_reads"] * (w["read_len"] + 1) / float(S)
    rows, v, m = synthetic_lineage(w["n_refs"])

    # shard k of strand s = second half of block k ++ first half of block k+1 (blocks are independent
    # draws of the runs model), so shard boundaries fall INSIDE segments and clusters straddle them
    shards = []
    for s in range(w["n_strands"]):
        bufs = [torch.zeros(S + HALO, dtype=torch.uint8, device=dev), torch.zeros(S + HALO, dtype=torch.int32, device=dev),
                torch.zeros(S + HALO, dtype=torch.int32, device=dev)]
        half = S // 2
        for blk, lo, hi, at in ((rank, half, S, 0), (rank + 1, 0, half, S - half)):
            e, l, d = datagen.synth_arrays(S, n_reads, w["n_refs"], w["read_len"], seed=w["seed"] + 1000 * s + 17 * blk,
                                           device=dev, lo=lo, hi=hi, read_frac=read_frac)
            bufs[0][at:at + hi - lo] = e; bufs[1][at:at + hi - lo] = l; bufs[2][at:at + hi - lo] = d
            del e, l, d
        torch.cuda.synchronize()
        h = multi.exchange_halo(dist, rank, world, [bufs[1], bufs[2], bufs[0]], S, HALO)
        shards.append(dict(ebwt=bufs[0], lcp=bufs[1], da=bufs[2], n_own=S, n_halo=h, pos_base=rank * S,
                           is_last=rank == world - 1))
    torch.cuda.synchronize()
    be = multi.CudaBackend(ctx, torch, dev)
    # LIME_MULTI=peer: updates go straight to the read owner's slab over NVLink (PeerJob); =exchange: partial
    # matrices + triples all-to-all (ShardedJob, also the exact fallback).  Default: whichever was measured
    # faster on this pool (profiles/README.md): peer adds on 2 GPUs, the exchange beyond (remote adds over
    # a footprint of several owners collapse to 0.3 G adds/s on 8 GPUs)
    mode = os.environ.get("LIME_MULTI", "auto")
    peer = mode == "peer" or (mode == "auto" and world <= 2)
    if peer:
        job = multi.PeerJob(dist, torch, be, rank, world, n_reads, w["n_refs"], w["n_strands"])
    else:
        job = multi.ShardedJob(dist, torch, be, rank, world, n_reads, w["n_refs"])

    def step(gather=True, src=None):
        job.stats = []
        if peer:
            job.begin()
        for s, sh in enumerate(src or shards):
            job.run_strand(s, sh, w["alpha"], w["beta"], w["ebwt"], w["read_len"])
        # every rank fetches the assignments of the reads it owns (9 B per read, as at N = 1)
        return job.finish(v, None, w["tax_rank"], False, gather=False)

    for _ in range(max(args.warmup, 1)):
        out = step()
    cnt = torch.tensor([int((out[0] == ord(c)).sum()) for c in "CUAH"], device=dev, dtype=torch.int64)
    dist.all_reduce(cnt)
    kinds = {c: int(x) for c, x in zip("CUAH", cnt.tolist()) if x}
    stats = list(job.stats)

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
        time.sleep(0.3)
    ctx.prof_enable(True); ctx.prof_reset()
    l0 = ctx.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.time()
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    ev1.record(stream)
    dist.barrier(); torch.cuda.synchronize()
    t1 = time.time()
    ms = torch.tensor([max(ev0.elapsed_time(ev1), 0.0)], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)                     # max over ranks
    ms_step = float(ms.item()) / args.steps
    launches = torch.tensor([ctx.launch_count() - l0], device=dev, dtype=torch.int64)
    dist.all_reduce(launches, op=dist.ReduceOp.SUM)
    prof = ctx.prof(); ctx.prof_enable(False)
    k12_all = torch.zeros(world, device=dev, dtype=torch.float64)
    k12_all[rank] = prof.get("k12_fused", {}).get("ms", 0.0) / max(prof.get("k12_fused", {}).get("launches", 1), 1)
    dist.all_reduce(k12_all)
    n_ext = torch.tensor([max(0, int(0.7 / max(ms_step * 1e-3, 1e-4)) - args.steps)], device=dev, dtype=torch.int64)
    dist.broadcast(n_ext, 0)                                       # same number of extra steps on every rank
    for _ in range(int(n_ext.item())):
        step()
    dist.barrier(); torch.cuda.synchronize()
    clk = clocks.summary(t0, time.time()) if rank == 0 else None
    if rank == 0:
        clk["window"] = "timed region + untimed repeats of the same step up to 0.7 s (nvidia-smi -lms 50)"
        clocks.stop()

    if os.environ.get("LIME_TRACE"):
        ht = _HostTimer()
        job.timer = ht
        dist.barrier(); torch.cuda.synchronize()
        t_h = time.perf_counter()
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        print(f"[trace] rank {rank}: 3 steps {1e3 * (time.perf_counter() - t_h):.2f} ms, host ms per phase (3 steps): "
              + ", ".join(f"{k}={v:.2f}" for k, v in ht.ms.items()), file=sys.stderr, flush=True)
    # one diagnostic step with synchronising per-phase timers (not part of the timed region)
    tm = _SyncTimer(torch)
    job.timer = tm
    step()
    job.timer = lambda name: multi._NoTimer()
    sent = sum(st.get("triples_sent", 0) for st in job.stats) * 8
    PH = ("zero", "compute", "export", "counts", "all_to_all", "import", "choose", "scalars", "classify", "gather", "check")
    phases = torch.tensor([tm.ms.get(k, 0.0) for k in PH], device=dev, dtype=torch.float64)
    dist.all_reduce(phases, op=dist.ReduceOp.MAX)
    sent_t = torch.tensor([sent], device=dev, dtype=torch.int64)
    dist.all_reduce(sent_t, op=dist.ReduceOp.MAX)

    # ---- e2e: the same step with every rank's shards in page-locked HOST memory (lime_load(LIME_MEM_HOST) per
    # strand inside the timed region) and each rank's assignments copied back to its host, max over ranks.
    # Every rank first proves, without collectives, that it can load its host shards; the leg runs only
    # if all ranks can (a rank dropping out of a step would leave the others in a collective).
    e2e = None
    if not args.no_e2e and os.environ.get("LIME_BENCH_MULTI_E2E", "1") != "0":
        ok, why, host_shards = 1, "", []
        try:
            for sh in shards:
                hs = dict(sh)
                for k in ("ebwt", "lcp", "da"):
                    t = torch.empty(sh[k].shape, dtype=sh[k].dtype, pin_memory=True)
                    t.copy_(sh[k])
                    hs[k] = t
                host_shards.append(hs)
            torch.cuda.synchronize()
            for hs in host_shards:
                be.load(hs["lcp"], hs["da"], hs["ebwt"] if w["ebwt"] else None, hs["n_own"], hs["n_halo"],
                        hs["pos_base"], hs["is_last"], n_reads, w["n_refs"])
            torch.cuda.synchronize()
        except BaseException as ex:                      # noqa: BLE001
            ok, why = 0, "%s: %s" % (type(ex).__name__, ex)
        flag = torch.tensor([ok], device=dev, dtype=torch.int64)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 1:
            out_d = tuple(np.array(x, copy=True) for x in step())
            out_h = tuple(np.array(x, copy=True) for x in step(src=host_shards))
            same = torch.tensor([int(all(a.tobytes() == b.tobytes() for a, b in zip(out_d, out_h)))], device=dev,
                                dtype=torch.int64)
            dist.all_reduce(same, op=dist.ReduceOp.MIN)
            e2e_steps = max(1, min(args.steps, 5))
            dist.barrier(); torch.cuda.synchronize()
            ev0.record(stream)
            tw0 = time.perf_counter()
            for _ in range(e2e_steps):
                step(src=host_shards)
            ev1.record(stream)
            torch.cuda.synchronize()
            tw1 = time.perf_counter()
            dist.barrier()
            e_ms = torch.tensor([max(ev0.elapsed_time(ev1), (tw1 - tw0) * 1e3)], device=dev, dtype=torch.float64)
            dist.all_reduce(e_ms, op=dist.ReduceOp.MAX)
            e_ms = float(e_ms.item()) / e2e_steps
            h2d = sum((8 + w["ebwt"]) * (hs["n_own"] + hs["n_halo"]) for hs in host_shards)
            h2d_t = torch.tensor([h2d], device=dev, dtype=torch.int64)
            dist.all_reduce(h2d_t)
            if int(same.item()) == 1:
                e2e = {"value": S * w["n_strands"] * world / (e_ms * 1e-3), "unit": UNIT,
                       "h2d_bytes_per_step": int(h2d_t.item()), "d2h_bytes_per_step": 9 * n_reads,
                       "ms_per_step": e_ms, "steps": e2e_steps, "reads_per_s": n_reads / (e_ms * 1e-3),
                       "note": "every rank: lime_load(LIME_MEM_HOST) of its shard of each strand from page-locked host "
                               "arrays + its assignments copied back; max over ranks; results equal to the resident step"}
            else:
                e2e = {"error": "host-buffer step disagrees with the resident step"}
        else:
            e2e = {"error": "host-shard pre-flight failed on some rank" + (": " + why if why else "")}
        del host_shards

    if rank == 0:
        symbols = S * w["n_strands"] * world
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak_gbs, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback")
        kern = {k: x for k, x in prof.items() if x["launches"]}
        top = max(kern, key=lambda k: kern[k]["ms"])
        alg_per_launch = (8 + w["ebwt"]) * (S + 0)
        avg_ms = kern[top]["ms"] / kern[top]["launches"]
        achieved = alg_per_launch / (avg_ms * 1e-3) / 1e9 if top == "k12_fused" else None
        ph = {k: float(x) for k, x in zip(PH, phases.tolist()) if x > 0}
        a2a_ms = ph.get("all_to_all", 0.0)
        par = (f"position-sharded x{world}, halo {HALO}; similarity rows live on the read owner's GPU and every update is a "
               "red.add over NVLink / NVSwitch peer memory issued by the fused kernel (no partial matrix, no all-to-all)" if peer else
               f"position-sharded x{world}, halo {HALO}, triples all-to-all by read owner")
        line = {
            "metric": METRIC, "value": symbols / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 1), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": w["name"] + f"; x{world} weak scaling: {world} position shards of {S} symbols per strand, {n_reads} reads",
                       "l2": "inputs (3.6 GB per strand and GPU) are larger than L2; no flush needed",
                       "parallelism": par + "; every rank copies the assignments of its own read range to its host",
                       "symbols_per_step": symbols, "reads_per_step": n_reads, "strand_stats": stats, "assignments": kinds},
            "reads_per_s": n_reads / (ms_step * 1e-3),
            "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak_gbs, "unit": "GB/s",
                         "frac": (achieved / peak_gbs) if achieved else None, "traffic": None, "peak_source": peak_src,
                         "alg_bytes_per_launch": alg_per_launch, "avg_launch_ms": avg_ms,
                         "kernels_ms_per_step_rank0": {k: x["ms"] / args.steps for k, x in kern.items()},
                         "k12_ms_per_launch_by_rank": [round(float(x), 3) for x in k12_all.tolist()],
                         "job_frac_of_aggregate_peak": (8 + w["ebwt"]) * symbols / (ms_step * 1e-3) / 1e9 / (peak_gbs * world)},
            "exchange": {"mode": "peer-memory adds" if peer else "triples all-to-all", "phases_ms_max_over_ranks": ph,
                         "all_to_all_bytes_sent_per_rank_max": int(sent_t.item()),
                         "all_to_all_GBps_per_rank": (int(sent_t.item()) / (a2a_ms * 1e-3) / 1e9) if a2a_ms > 0 else None,
                         "remote_updates_per_rank_and_step": (sum(st.get("updates", 0) for st in stats) // world * (world - 1) // world) if peer else None,
                         "nvlink_peak_GBps_per_direction": 900, "note": "phase times from one extra synchronised step"},
            "cpu_baseline": None, "e2e": e2e, "gpu_launches": int(launches.item()), "clocks": clk,
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    if peer:
        job.close()
    ctx.close()
    dist.destroy_process_group()


def main():
    if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "WARN"):
        os.environ["NCCL_DEBUG"] = "NONE"          # NCCL would print its version banner on stdout next to the JSON line
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=CPU_SAMPLE_N)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-narrow", action="store_true", help="skip the e2e leg with the narrowed lcp transfer")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--small", action="store_true", help="reduced shape for functional testing only")
    ap.add_argument("--unfused", action="store_true", help="cluster list between steps 1 and 2 (for comparison)")
    args = ap.parse_args()
    w = dict(WORKLOAD)
    if args.small:
        w = _small(w, 4_000_000, 20_000)
        args.cpu_sample = min(args.cpu_sample, 1 << 20)
    if args.impl == "reference":
        run_reference_arm(args, w)
    else:
        run_ours(args, w)


if __name__ == "__main__":
    main()
