#!/usr/bin/env python
"""bench.py -- LiME classification hot path on B200: eBWT symbols/s.

    python bench.py --gpus N --steps K --warmup W            (ours; torchrun for N > 1)
    python bench.py --impl reference --gpus N --steps K ...  (the reference's CPU pipeline)
    python bench.py --config cfg3|cfg4 ...                   (the north-star shapes, strong scaling)
    python bench.py --n 1000000000 --alpha 12 ...            (config 5 sweep points on the cfg2 model)

Default workload (BASELINE.json configs[1]): single-end synthetic, 2 M reads x 100 bp + 100 genomes,
N = 402 M eBWT symbols per strand, two strands (F, RC), EBWT=1, alpha=16, beta=0.25, TaxRank=1.
One step = the whole job: per strand cluster detection (step 1), similarity accumulation and
clusterChoose (step 2), then Classify over all strands (step 3).  At N > 1 GPUs the position
range of every strand is sharded contiguously (cfg2: weak scaling, every GPU holds 402 M symbols per
strand and the job has N x 2 M reads; cfg3 / cfg4: strong scaling of one fixed job).

value  = symbols of all strands / step time, inputs resident in HBM (CUDA events).
e2e    = the same through the C ABI with HOST buffers: pinned host -> device copies of
         lcp/da/ebwt and the device -> host copy of the assignments inside the timed region.
parity = BASELINE.md's gate: before a value is printed the GPU result is compared byte for byte with the
         reference's files (.out, .res.bin, .res.pos, CSV) of the SAME input (N = 1), resp. with the
         single-GPU job (N > 1).  A mismatch aborts the run.
The oracle / reference binaries are used only for cpu_baseline, the parity gate and --impl reference.
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

METRIC = "ebwt_symbols_per_s"
UNIT = "symbols/s"
HALO = 70_000          # positions a shard sees of the next one: more than sizeMaxBuf (65536, ClusterBWT_DA.cpp:558)
BLOCKS = 8             # strong-scaled configs: a strand is 8 independent draws of the runs model, whatever N

CONFIGS = {
    # BASELINE.json configs[1]; the line the driver records
    "cfg2": dict(name="cfg2 single-end synthetic: 2M reads x 100bp + 100 genomes, 402M symbols/strand, "
                      "2 strands (F,RC), EBWT=1, alpha=16, beta=0.25, TaxRank=1",
                 n=402_000_000, n_reads=2_000_000, n_refs=100, read_len=100, alpha=16, beta=0.25, ebwt=1, n_strands=2,
                 tax_rank=1, higher=False, trlcp=None, scaling="weak", seed=20261104 + 2),
    # configs[2]: setA2 shape, 4 strands x 1.5 G symbols, EBWT=0, 930 genomes (the row count of
    # Datasets/Reference_database.csv), strong scaling over 1-8 GPUs
    "cfg3": dict(name="cfg3 setA2-shape paired-end: 4 strands x 1.5G symbols, 10M reads x 100bp, 930 genomes, EBWT=0, "
                      "alpha=16, beta=0.25, TaxRank=1",
                 n=1_500_000_000, n_reads=10_000_000, n_refs=930, read_len=100, alpha=16, beta=0.25, ebwt=0, n_strands=4,
                 tax_rank=1, higher=False, trlcp=None, scaling="strong", seed=20261104 + 3),
    # configs[3]: the north-star run, 10 G symbols (4 strands x 2.5 G), HIGHER=1, lcp truncated at 32
    "cfg4": dict(name="cfg4 HIGHER=1 paired-end: 4 strands x 2.5G symbols (10G total), 10M reads x 150bp, 1000 genomes, "
                      "EBWT=1, lcp truncated at 32, alpha=16, beta=0.25, TaxRank=1",
                 n=2_500_000_000, n_reads=10_000_000, n_refs=1000, read_len=150, alpha=16, beta=0.25, ebwt=1, n_strands=4,
                 tax_rank=1, higher=True, trlcp=32, scaling="strong", seed=20261104 + 4),
}


def _small(w, n, n_reads):
    w = dict(w)
    w.update(n=n, n_reads=n_reads, name=f"REDUCED test shape n={n} n_reads={n_reads} of {w['name']} (not the benchmark)")
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


def _peak():
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        if "hbm_gbs" in peaks:
            return float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# inputs
# ------------------------------------------------------------------------------------------------
def _lineage(w):
    from lime_b200.pipeline import synthetic_lineage
    rows, v, m = synthetic_lineage(w["n_refs"])
    return rows, v, (m if w["higher"] else None)


def _lineage_file(path, w):
    rows, v, m = _lineage(w)
    with open(path, "w") as f:
        f.write("Acc;Species_TaxID;Genus_TaxID;Family_TaxID;Order_TaxID;Class_TaxID;Phylum_TaxID\n")
        for r in rows:
            f.write(";".join(str(x) for x in r) + "\n")
    return v, m


def _block_arrays(w, s, blk, n_blk, dev, lo=0, hi=None):
    """Positions [lo, hi) of block `blk` of strand s: an independent draw of the runs model with the job's read
    and genome ids (strong-scaled configs: the strand is the concatenation of BLOCKS blocks, whatever N)."""
    from lime_b200 import datagen
    read_frac = w["n_reads"] * (w["read_len"] + 1) / float(w["n"])
    return datagen.synth_arrays(n_blk, w["n_reads"], w["n_refs"], w["read_len"], seed=w["seed"] + 1000 * s + 17 * blk,
                                device=dev, lo=lo, hi=hi, read_frac=read_frac, trlcp=w["trlcp"])


def _strand_arrays(w, s, dev, hi=None):
    """A whole strand on one device (N = 1)."""
    import torch
    from lime_b200 import datagen
    if w["scaling"] == "weak":
        return datagen.synth_arrays(w["n"], w["n_reads"], w["n_refs"], w["read_len"], seed=w["seed"] + 1000 * s, device=dev,
                                    trlcp=w["trlcp"], hi=hi)
    n_blk = w["n"] // BLOCKS
    n = n_blk * BLOCKS if hi is None else min(hi, n_blk * BLOCKS)
    e = torch.empty(n, dtype=torch.uint8, device=dev)
    l = torch.empty(n, dtype=torch.int32, device=dev)
    d = torch.empty(n, dtype=torch.int32, device=dev)
    for b in range(BLOCKS):
        a0 = b * n_blk
        if a0 >= n:
            break
        k = min(n_blk, n - a0)
        eb, lb, db = _block_arrays(w, s, b, n_blk, dev, hi=k)
        e[a0:a0 + k] = eb; l[a0:a0 + k] = lb; d[a0:a0 + k] = db
        del eb, lb, db
    return e, l, d


def _scratch_dir():
    for base in ("/dev/shm", tempfile.gettempdir()):
        try:
            if os.path.isdir(base) and shutil.disk_usage(base).free > (12 << 30):
                return tempfile.mkdtemp(prefix="lime_bench_", dir=base)
        except OSError:
            pass
    return tempfile.mkdtemp(prefix="lime_bench_")


# ------------------------------------------------------------------------------------------------
# the reference's CPU pipeline (oracle/_ref binaries; the one place bench.py executes anything under oracle/)
# ------------------------------------------------------------------------------------------------
class CpuReference:
    """Runs the unmodified reference (ClusterLCP, ClusterBWT_DA, Classify from oracle/_ref) with
    the schedule of LiME_paired.sh: ClusterLCP of all strands concurrently (:44-53), ClusterBWT_DA
    strand after strand (:62-68), one Classify (:79); all with threads = host cores."""

    def __init__(self, w, sample_n):
        from oracle.oracle import Reference
        self.ref = Reference()
        self.kind = "reference" if self.ref.available() else "port"
        self.w, self.sample_n = w, min(sample_n, w["n"])
        self.same_config = self.sample_n == w["n"]
        self.cores = os.cpu_count() or 1
        self.dir = _scratch_dir()
        self.fastas = []

    def write_inputs(self, strands_host):
        """strands_host: list of (ebwt u8, lcp u32, da u32) numpy arrays."""
        from lime_b200 import datagen
        for s, (e, l, d) in enumerate(strands_host):
            fasta = os.path.join(self.dir, f"strand{s}+Refs.fasta")
            datagen.write_strand(fasta, e, l, d)
            self.fastas.append(fasta)
        self.lineage = os.path.join(self.dir, "lineage.csv")
        self.v, self.m = _lineage_file(self.lineage, self.w)
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
            r = subprocess.run([exe("Classify_higher1" if w["higher"] else "Classify_higher0"), str(len(self.fastas))] +
                               [f + ".res" for f in self.fastas] +
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
            orc.classify(bins, poss, w["n_refs"], self.v, self.m, w["tax_rank"], w["higher"])
        return time.perf_counter() - t0

    def files_of(self, s):
        """The reference's outputs of strand s after a step: (.out bytes, .res.bin bytes, .res.pos bytes)."""
        f = self.fastas[s]
        stem = f[:f.find(".fasta")]
        return open(stem + ".out", "rb").read(), open(f + ".res.bin", "rb").read(), open(f + ".res.pos", "rb").read()

    def csv(self):
        return open(os.path.join(self.dir, "out.csv"), "rb").read()

    def describe(self):
        what = ("the whole workload" if self.same_config else
                f"prefix of {self.sample_n} positions per strand (extrapolated per symbol)")
        return (f"{what}: {self.sample_n} positions x {len(self.fastas)} strands (all {self.w['n_reads']} reads, "
                f"{self.w['n_refs']} genomes), files in {os.path.dirname(self.dir) or self.dir}, "
                f"page cache warm; ClusterLCP x{len(self.fastas)} concurrent, ClusterBWT_DA x{len(self.fastas)} sequential, "
                f"Classify {len(self.fastas)}; threads={self.cores if self.kind == 'reference' else 1}")

    def close(self):
        shutil.rmtree(self.dir, ignore_errors=True)


def _host_sample(w, sample_n, device):
    """The first sample_n positions of every strand, as numpy arrays in the file formats."""
    out = []
    for s in range(w["n_strands"]):
        e, l, d = _strand_arrays(w, s, device, hi=sample_n)
        out.append((e.cpu().numpy(), l.cpu().numpy().view(np.uint32), d.cpu().numpy().view(np.uint32)))
        del e, l, d
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
            "higher_is_better": True, "scaling": w["scaling"], "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": w["name"], "sample": cpu.describe(), "same_config": cpu.same_config},
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
    n, nr, ng = st["n"], st["n_reads_total"], w["n_refs"]
    if kname == "k1_detect":
        return 8 * n + 16 * st["n_clusters"]
    if kname == "k12_fused":
        return (8 + E) * n
    if kname == "k3_choose":
        return nr * ng + 8 * st["n_records"] + 8 * nr
    if kname == "k3_classify":
        return 8 * st["n_records"] + 8 * nr + 9 * nr
    return 0


def _parity_gate(ctx, cpu, w, out, stats):
    """.out, .res.bin, .res.pos of every strand and the final CSV, byte for byte against
    the reference's files of the same input.  Returns the parity record; raises SystemExit on a mismatch."""
    from lime_b200 import api
    files = []
    for s in range(w["n_strands"]):
        out_ref, bin_ref, pos_ref = cpu.files_of(s)
        n_cl, max_len, n_rec = (int(x) for x in stats[s][:3])
        aux = (np.array([w["n_reads"], w["n_refs"], w["alpha"]], dtype="<u4").tobytes() +
               np.array([max_len, n_cl], dtype="<u8").tobytes())                       # ClusterLCP.cpp:304-308
        bin_g, pos_g = ctx.choose_result(s)
        for name, a, b in ((f"strand{s}.out", aux, out_ref), (f"strand{s}.res.bin", bin_g.tobytes(), bin_ref),
                           (f"strand{s}.res.pos", pos_g.tobytes(), pos_ref)):
            if a != b:
                raise SystemExit(f"bench.py: PARITY FAILED on {name} ({len(a)} vs {len(b)} bytes): no value is reported")
            files.append(name)
        del bin_g, pos_g
    path = os.path.join(cpu.dir, "gpu.csv")
    api.write_csv(path, out[0], out[1], out[2])
    if open(path, "rb").read() != cpu.csv():
        raise SystemExit("bench.py: PARITY FAILED on the classification CSV: no value is reported")
    files.append("out.csv")
    return {"checked": True, "against": "the unmodified reference (oracle/_ref) on the same input, byte for byte",
            "files": files, "n": cpu.sample_n, "strands": w["n_strands"]}


def run_ours(args, w):
    import torch
    from lime_b200.api import LimeContext, MEM_DEVICE, MEM_HOST
    from lime_b200.pipeline import run_job

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
    rows, v, m = _lineage(w)
    use_ebwt = bool(w["ebwt"])

    # ---- synthetic inputs, resident in HBM ----------------------------------------------------
    strands_dev, keep = [], []
    for s in range(w["n_strands"]):
        e, l, d = _strand_arrays(w, s, dev)
        if not use_ebwt:
            e = None                                     # the EBWT=0 build never opens .ebwt (ClusterBWT_DA.cpp:16-18)
        keep.append((e, l, d))
        strands_dev.append({"ebwt": e.data_ptr() if e is not None else None, "lcp": l.data_ptr(), "da": d.data_ptr(),
                            "n": int(l.shape[0])})
    n_pos = strands_dev[0]["n"]
    torch.cuda.synchronize()

    def step_resident():
        return run_job(ctx, strands_dev, w["n_reads"], w["n_refs"], w["read_len"], w["alpha"], w["beta"], w["ebwt"],
                       v, m, w["tax_rank"], w["higher"], mem=MEM_DEVICE, fused=not args.unfused)

    # ---- warm-up (also gathers the per-strand statistics the roofline needs) -------------------
    for _ in range(max(args.warmup, 3)):
        out, stats = step_resident()
    out = tuple(np.array(x, copy=True) for x in out)    # classify() returns views of context-owned pinned buffers
    stats = [tuple(int(x) for x in st) for st in stats]
    kinds = {chr(t): int((out[0] == t).sum()) for t in np.unique(out[0])}
    strand_stats = [{"n": n_pos, "n_clusters": st[0], "max_len": st[1], "n_records": st[2], "n_reads_total": w["n_reads"]}
                    for st in stats]

    # ---- CPU baseline beside it + the parity gate (rank 0, N = 1 only) -------------------------------------
    cpu_baseline, parity = None, {"checked": False, "why": "--no-cpu"}
    if not args.no_cpu:
        cpu = CpuReference(w, args.cpu_sample)
        try:
            k = cpu.sample_n
            sample = [((e[:k].cpu().numpy() if e is not None else np.zeros(k, np.uint8)), l[:k].cpu().numpy().view(np.uint32),
                       d[:k].cpu().numpy().view(np.uint32)) for (e, l, d) in keep]
            cpu.write_inputs(sample)
            del sample
            cpu.step()                                   # warms the page cache
            t = cpu.step()
            cpu_baseline = {"value": cpu.sample_n * w["n_strands"] / t, "unit": UNIT,
                            "cores": cpu.cores if cpu.kind == "reference" else 1, "kind": cpu.kind,
                            "sample": cpu.describe(), "seconds": t, "same_config": cpu.same_config}
            if cpu.kind == "reference":
                if cpu.same_config:
                    parity = _parity_gate(ctx, cpu, w, out, stats)
                else:
                    # the GPU path on the very prefix the reference ran on
                    pre = [{"ebwt": sd["ebwt"], "lcp": sd["lcp"], "da": sd["da"], "n": k} for sd in strands_dev]
                    out_p, stats_p = run_job(ctx, pre, w["n_reads"], w["n_refs"], w["read_len"], w["alpha"], w["beta"], w["ebwt"],
                                             v, m, w["tax_rank"], w["higher"], mem=MEM_DEVICE, fused=not args.unfused)
                    parity = _parity_gate(ctx, cpu, w, out_p, stats_p)
                    step_resident()                      # slots back to the full job
            else:
                parity = {"checked": False, "why": "oracle/_ref not built on this box"}
        finally:
            cpu.close()

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
    ms_total = ev0.elapsed_time(ev1)
    launches = ctx.launch_count() - l0
    prof = ctx.prof()
    ctx.prof_enable(False)
    # The reported K steps: no per-launch events, and two jobs in flight -- a second context on its own stream, the strands
    # of job k + 1 are queued (lime_load of borrowed device arrays + lime_strand_run never wait) before the host waits for
    # job k's assignments, so the result copy and the host's gaps of job k run beside the scoring of job k + 1.
    # LIME_BENCH_INFLIGHT=1: one job at a time.
    inflight = 1
    ctx2 = None
    free_now, total_mem = torch.cuda.mem_get_info()
    if (not args.unfused and os.environ.get("LIME_BENCH_INFLIGHT", "2") != "1"
            and free_now > 1.25 * (w["n_reads"] * ((w["n_refs"] + 15) // 16 * 16) + 20 * w["n_reads"]) + (4 << 30)):   # its matrix and slots
        ctx2 = LimeContext(local)
        stream2 = torch.cuda.Stream()
        ctx2.set_stream(stream2.cuda_stream)
        out2, stats2 = run_job(ctx2, strands_dev, w["n_reads"], w["n_refs"], w["read_len"], w["alpha"], w["beta"], w["ebwt"],
                               v, m, w["tax_rank"], w["higher"], mem=MEM_DEVICE, fused=True)
        assert all(a.tobytes() == np.asarray(b).tobytes() for a, b in zip(out, out2)), "second context disagrees"
        assert [tuple(int(y) for y in x) for x in stats2] == stats
        inflight = 2

    def steps_pair(k_steps, strands, mem, one_at_a_time):
        """k_steps jobs over `strands`; two in flight when the second context is there."""
        if inflight == 1:
            for _ in range(k_steps):
                one_at_a_time()
            return
        from lime_b200.pipeline import norm_of
        from lime_b200.api import LimeError
        pair, prev = (ctx, ctx2), None
        try:
            for k in range(k_steps):
                c = pair[k % 2]
                for s_, st in enumerate(strands):
                    c.load(st["lcp"], st["da"], st["ebwt"] if w["ebwt"] else None, w["n_reads"], w["n_refs"], mem=mem, n_total=st["n"])
                    c.strand_run(s_, w["alpha"], w["ebwt"], norm_of(w["read_len"], w["alpha"]), w["beta"])
                if prev is not None:
                    prev.classify(w["n_strands"], w["n_refs"], v, m, w["tax_rank"], w["higher"])
                prev = c
            prev.classify(w["n_strands"], w["n_refs"], v, m, w["tax_rank"], w["higher"])
        except LimeError:
            # a strand that must be redone exactly (not expected on this workload): one job at a time from here on
            torch.cuda.synchronize()
            for _ in range(k_steps):
                one_at_a_time()

    def steps_resident(k_steps):
        steps_pair(k_steps, strands_dev, MEM_DEVICE, step_resident)

    steps_resident(4)
    ev_b = torch.cuda.Event()
    torch.cuda.synchronize()
    ev0.record(stream)
    steps_resident(args.steps)
    if ctx2 is not None:
        ev_b.record(stream2)
        stream.wait_event(ev_b)
    ev1.record(stream)
    torch.cuda.synchronize()
    ms_plain = ev0.elapsed_time(ev1)
    t1x = _extend_for_clocks(lambda: (steps_resident(2), torch.cuda.synchronize()), t0)
    inflight_used = inflight
    clk = clocks.summary(t0, t1x)
    clk["window"] = "timed region + untimed repeats of the same step up to 0.7 s (nvidia-smi -lms 50)"

    ms_step = ms_plain / args.steps                   # the reported K steps; ms_total is the instrumented pass
    symbols = n_pos * w["n_strands"]
    value = symbols / (ms_step * 1e-3)

    # ---- roofline of the dominant kernel ----------------------------------------------------------
    peak_gbs, peak_src = _peak()
    kern = {k: x for k, x in prof.items() if x["launches"]}
    kernel_ms_total = sum(x["ms"] for x in kern.values())
    top = max(kern, key=lambda k: kern[k]["ms"])
    if top == "k3_classify":
        agg = dict(strand_stats[0]); agg["n_records"] = sum(s["n_records"] for s in strand_stats)
        alg = _alg_bytes(top, agg, w) * args.steps
    else:
        alg = sum(_alg_bytes(top, s, w) for s in strand_stats) * args.steps
    avg_launch_ms = kern[top]["ms"] / kern[top]["launches"]
    alg_per_launch = alg / kern[top]["launches"]
    achieved = alg_per_launch / (avg_launch_ms * 1e-3) / 1e9
    traffic, traffic_src = None, None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        ent = tr.get(top, {})
        if ent.get("workload", "cfg2") == args.config and not args.n:
            traffic = ent.get("dram_bytes_per_launch")
            traffic_src = "static: profiles/traffic.json (%s), one ncu --set full capture of this kernel on this workload" % ent.get("capture", "?")
    except Exception:
        pass
    job_alg = sum((8 + w["ebwt"]) * s["n"] for s in strand_stats)
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak_gbs, "unit": "GB/s",
                "frac": achieved / peak_gbs, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "alg_bytes_per_launch": alg_per_launch, "alg_bytes_per_symbol": 8 + w["ebwt"], "avg_launch_ms": avg_launch_ms,
                "launches_per_step": kern[top]["launches"] / args.steps,
                "kernel_share_of_step": kern[top]["ms"] / ms_total,
                "kernels_ms_per_step": {k: x["ms"] / args.steps for k, x in kern.items()},
                "kernel_time_fraction_of_step": kernel_ms_total / ms_total,
                "ms_per_step_with_kernel_events": ms_total / args.steps, "ms_per_step_plain": ms_plain / args.steps,
                "job_alg_bytes_per_step": job_alg,
                "job_frac_of_peak": job_alg / (ms_step * 1e-3) / 1e9 / peak_gbs}

    # ---- e2e: host buffers through the C ABI ----------------------------------------------------------
    e2e = None
    host_bytes = sum((8 + w["ebwt"]) * sd["n"] for sd in strands_dev)
    if not args.no_e2e and (host_bytes <= (24 << 30) or args.e2e):
        strands_host, pins = [], []
        for (e, l, d) in keep:
            pe = None
            if e is not None:
                pe = torch.empty(e.shape, dtype=e.dtype, pin_memory=True); pe.copy_(e)
            pl = torch.empty(l.shape, dtype=l.dtype, pin_memory=True); pl.copy_(l)
            pd = torch.empty(d.shape, dtype=d.dtype, pin_memory=True); pd.copy_(d)
            pins.append((pe, pl, pd))
            strands_host.append({"ebwt": pe.numpy() if pe is not None else None, "lcp": pl.numpy().view(np.uint32),
                                 "da": pd.numpy().view(np.uint32), "n": int(pl.shape[0])})
        torch.cuda.synchronize()

        def step_host():
            return run_job(ctx, strands_host, w["n_reads"], w["n_refs"], w["read_len"], w["alpha"], w["beta"],
                           w["ebwt"], v, m, w["tax_rank"], w["higher"], mem=MEM_HOST, fused=not args.unfused)
        d2h = 9 * w["n_reads"]
        e2e_steps = max(1, args.steps)

        def time_host(note, h2d_bytes):
            out_h, st_h = step_host()
            for a, b in zip(out, out_h):
                assert a.tobytes() == np.asarray(b).tobytes(), "host-buffer path disagrees with the resident path"
            assert [tuple(int(y) for y in x) for x in st_h] == stats, "per-strand counts disagree with the resident path"
            if ctx2 is not None:                        # the second context on the same host arrays, same answer
                steps_pair(2, strands_host, MEM_HOST, step_host)
                nr_ = w["n_reads"]                       # the page-locked result arrays of its last lime_classify
                out_2 = (ctx2.pinned("cls_type", (nr_,), np.uint8), ctx2.pinned("cls_tax", (nr_,), np.uint32),
                         ctx2.pinned("cls_sim", (nr_,), np.float32))
                for a, b in zip(out, out_2):
                    assert a.tobytes() == np.asarray(b).tobytes(), "host-buffer path (second context) disagrees with the resident path"
            torch.cuda.synchronize()
            ev0.record(stream)
            tw0 = time.perf_counter()
            steps_pair(e2e_steps, strands_host, MEM_HOST, step_host)       # two jobs in flight, as in the resident steps
            if ctx2 is not None:
                ev_b.record(stream2)
                stream.wait_event(ev_b)
            ev1.record(stream)
            torch.cuda.synchronize()
            tw1 = time.perf_counter()
            e_ms = max(ev0.elapsed_time(ev1), (tw1 - tw0) * 1e3) / e2e_steps
            return {"value": symbols / (e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e_ms, "steps": e2e_steps,
                    "reads_per_s": w["n_reads"] / (e_ms * 1e-3), "h2d_GBps": h2d_bytes / (e_ms * 1e-3) / 1e9,
                    "note": note + ("; two jobs in flight (the loads of job k + 1 start before job k's assignments are awaited)" if ctx2 is not None else "")}
        modes = {}
        both = [ctx] + ([ctx2] if ctx2 is not None else [])
        if w["alpha"] <= 255 and not args.no_narrow:
            # the drop-in executables know alpha from argv and name it (lime_set_alpha): lcp crosses PCIe as the bit
            # lcp < alpha, packed on host threads inside the timed region, expanded to bytes on the device
            workers = max(1, ctx.lib.lime_host_threads() - 1)
            da3 = w["n_reads"] + w["n_refs"] <= (1 << 24) and os.environ.get("LIME_PACK_DA", "1") != "0"
            for c_ in both:
                c_.set_alpha(w["alpha"])
            try:
                modes["lcp_bits"] = time_host(
                    "lime_set_alpha(alpha) + lime_load(LIME_MEM_HOST) per strand from pinned host arrays: lcp packed to one bit "
                    "per position (lcp < alpha) by %d host threads inside the timed region; assignments copied back"
                    % workers + ("; da as 3 bytes per position; bound by the packing threads (the bus runs below its rate)" if da3 else "; PCIe-bound"),
                    host_bytes - (5 if da3 else 4) * symbols + sum((sd["n"] + 7) // 8 for sd in strands_dev) + 4 * w["n_refs"])
            finally:
                for c_ in both:
                    c_.set_alpha(0)
            # the weaker promise alpha <= 255 (lime_set_alpha_limit): min(lcp, 255) as bytes
            for c_ in both:
                c_.set_alpha_limit(w["alpha"])
            try:
                modes["narrow_lcp"] = time_host(
                    "lime_set_alpha_limit(alpha) + lime_load(LIME_MEM_HOST) per strand from pinned host arrays: lcp narrowed to "
                    "bytes by %d host threads inside the timed region; assignments copied back; PCIe-bound"
                    % workers, host_bytes - 3 * symbols + 4 * w["n_refs"])
            finally:
                for c_ in both:
                    c_.set_alpha_limit(0)
        if args.e2e_plain or not modes:
            modes["plain"] = time_host("lime_load(LIME_MEM_HOST) per strand from pinned host arrays + assignments copied back; "
                                       "PCIe-bound", host_bytes + 4 * w["n_refs"])
        best = max(modes, key=lambda k: modes[k]["value"])
        e2e = dict(modes[best])
        e2e["mode"] = best
        e2e["modes"] = {k: {q: x[q] for q in ("value", "ms_per_step", "h2d_bytes_per_step", "h2d_GBps")} for k, x in modes.items()}
        del pins, strands_host
    elif not args.no_e2e:
        e2e = {"value": None, "note": "skipped: %.0f GB of page-locked host arrays per step; run with --e2e" % (host_bytes / 1e9)}
    clocks.stop()

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": w["scaling"], "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": w["name"], "config": args.config, "alpha": w["alpha"],
                   "l2": "inputs (%.1f GB per strand) are larger than L2; no flush needed" % ((8 + w["ebwt"]) * n_pos / 1e9),
                   "parallelism": "1 GPU, %d job(s) in flight" % inflight_used, "symbols_per_step": symbols, "reads_per_step": w["n_reads"],
                   "strand_stats": strand_stats, "assignments": kinds},
        "reads_per_s": w["n_reads"] / (ms_step * 1e-3),
        "parity": parity,
        "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "gpu_launches": int(launches),
        "clocks": clk,
    }
    print(json.dumps(line), flush=True)
    if ctx2 is not None:
        ctx2.close()
    ctx.close()


class _SyncTimer:
    """Per-phase wall timer with a device synchronise and a barrier on both sides (diagnostic step only)."""

    def __init__(self, torch, dist):
        self.torch, self.dist, self.ms = torch, dist, {}

    def __call__(self, name):
        outer = self

        class T:
            def __enter__(self_inner):
                outer.torch.cuda.synchronize()
                outer.dist.barrier()
                outer.torch.cuda.synchronize()
                self_inner.t = time.perf_counter()

            def __exit__(self_inner, *a):
                outer.torch.cuda.synchronize()
                outer.ms[name] = outer.ms.get(name, 0.0) + (time.perf_counter() - self_inner.t) * 1e3
                return False
        return T()


def _make_shards(w, world, rank, dev):
    """This rank's shard of every strand, halo in place.  weak: every rank holds one 402 M-position shard
    (second half of block k ++ first half of block k+1, so shard boundaries fall INSIDE segments and clusters
    straddle them), the job has world x n_reads reads.  strong: the strand is BLOCKS fixed blocks, rank k owns
    blocks [k * BLOCKS / world, (k+1) * BLOCKS / world)."""
    import torch
    from lime_b200 import datagen
    shards = []
    if w["scaling"] == "weak":
        S = w["n"]
        n_reads = w["n_reads"] * world
        read_frac = w["n_reads"] * (w["read_len"] + 1) / float(S)
        for s in range(w["n_strands"]):
            bufs = [torch.zeros(S + HALO, dtype=torch.uint8, device=dev), torch.zeros(S + HALO, dtype=torch.int32, device=dev),
                    torch.zeros(S + HALO, dtype=torch.int32, device=dev)]
            half = S // 2
            last = rank == world - 1
            for blk, lo, hi, at in ((rank, half, S, 0), (rank + 1, 0, half, S - half), (rank + 1, half, half + (0 if last else HALO), S)):
                if hi <= lo:
                    continue
                e, l, d = datagen.synth_arrays(S, n_reads, w["n_refs"], w["read_len"], seed=w["seed"] + 1000 * s + 17 * blk,
                                               device=dev, lo=lo, hi=hi, read_frac=read_frac, trlcp=w["trlcp"])
                bufs[0][at:at + hi - lo] = e; bufs[1][at:at + hi - lo] = l; bufs[2][at:at + hi - lo] = d
                del e, l, d
            shards.append(dict(ebwt=bufs[0] if w["ebwt"] else None, lcp=bufs[1], da=bufs[2], n_own=S, n_halo=0 if last else HALO,
                               pos_base=rank * S, is_last=last))
        return shards, n_reads, S * world
    assert BLOCKS % world == 0, "strong-scaled configs need a GPU count that divides %d" % BLOCKS
    n_blk = w["n"] // BLOCKS
    per = BLOCKS // world
    own = per * n_blk
    last = rank == world - 1
    for s in range(w["n_strands"]):
        bufs = [torch.zeros(own + HALO, dtype=torch.uint8, device=dev) if w["ebwt"] else None,
                torch.zeros(own + HALO, dtype=torch.int32, device=dev), torch.zeros(own + HALO, dtype=torch.int32, device=dev)]
        for j in range(per + (0 if last else 1)):
            blk = rank * per + j
            k = n_blk if j < per else HALO
            e, l, d = _block_arrays(w, s, blk, n_blk, dev, hi=k)
            at = j * n_blk
            if bufs[0] is not None:
                bufs[0][at:at + k] = e
            bufs[1][at:at + k] = l; bufs[2][at:at + k] = d
            del e, l, d
        shards.append(dict(ebwt=bufs[0], lcp=bufs[1], da=bufs[2], n_own=own, n_halo=0 if last else HALO,
                           pos_base=rank * own, is_last=last))
    return shards, w["n_reads"], n_blk * BLOCKS


def _multi_parity(args, w, world, rank, dev, dist, torch, ctx, be):
    """Before any timing: a reduced job of the same model (>= 50 M positions per strand) sharded over the ranks
    against the single-GPU job on the whole arrays, which every rank computes for itself."""
    from lime_b200 import datagen, multi
    from lime_b200.api import MEM_DEVICE
    from lime_b200.pipeline import run_job
    n = max(50_000_000, 8_000_000 * world) if not args.small else 4_000_000
    n_reads = max(1000, int(w["n_reads"] * (n / w["n"]) * (world if w["scaling"] == "weak" else 1)))
    n_reads = min(n_reads, n // 8)
    rows, v, m = _lineage(w)
    read_frac = min(0.9, n_reads * (w["read_len"] + 1) / float(n))
    whole, shards = [], []
    for s in range(w["n_strands"]):
        e, l, d = datagen.synth_arrays(n, n_reads, w["n_refs"], w["read_len"], seed=w["seed"] + 555 + s, device=dev,
                                       trlcp=w["trlcp"], read_frac=read_frac)
        whole.append((e, l, d))
        lo, hi = multi.shard_ranges(n, world)[rank]
        k = min(HALO, n - hi)
        shards.append(dict(ebwt=e[lo:hi + k].clone() if w["ebwt"] else None, lcp=l[lo:hi + k].clone(), da=d[lo:hi + k].clone(),
                           n_own=hi - lo, n_halo=k, pos_base=lo, is_last=rank == world - 1))
    torch.cuda.synchronize()
    strands = [{"ebwt": e.data_ptr() if w["ebwt"] else None, "lcp": l.data_ptr(), "da": d.data_ptr(), "n": n} for (e, l, d) in whole]
    exp, st1 = run_job(ctx, strands, n_reads, w["n_refs"], w["read_len"], w["alpha"], w["beta"], w["ebwt"], v, m,
                       w["tax_rank"], w["higher"], mem=MEM_DEVICE)
    exp = tuple(np.array(x, copy=True) for x in exp)
    flat_cap, inbox_cap = multi.xchg_capacities(max(sh["n_own"] + sh["n_halo"] for sh in shards), world)
    xj = multi.XchgJob(dist, torch, be, rank, world, n_reads, w["n_refs"], w["n_strands"], flat_cap, inbox_cap)
    got, st, route = multi.run_sharded(dist, torch, be, rank, world, n_reads, w["n_refs"], shards, w["alpha"], w["beta"], w["ebwt"],
                                       w["read_len"], v, m, w["tax_rank"], w["higher"], gather=True, xchg_job=xj)
    if not xj.closed:
        xj.close()
    same = all(np.asarray(a).tobytes() == b.tobytes() for a, b in zip(got, exp))
    same = same and all(js["n_clusters"] == int(s1[0]) and js["max_len"] == int(s1[1]) for js, s1 in zip(st, st1))
    flag = torch.tensor([int(same)], device=dev, dtype=torch.int64)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    del whole, shards
    torch.cuda.empty_cache()
    if int(flag.item()) != 1:
        raise SystemExit("bench.py: PARITY FAILED: the sharded job differs from the single-GPU job: no value is reported")
    return {"checked": True, "against": "the single-GPU job on the whole arrays, computed on every rank (assignments of all reads, "
            "nClusters, maxLen)", "n": n, "strands": w["n_strands"], "n_reads": n_reads, "route": route}


def run_ours_multi(args, w, world, rank, local):
    """N > 1: see lime_b200/multi.py for the exchange."""
    import torch
    import torch.distributed as dist
    from lime_b200 import multi
    from lime_b200.api import LimeContext

    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    if "LIME_PACK_DA" not in os.environ:
        # host loads of several ranks are bound by the host's cores and memory system, not by PCIe (2 ranks: 21.7 GB/s per
        # GPU, 8 ranks: 13.6): repacking da to 3 bytes costs host work and saves none where it counts (measured, G symbols/s:
        # N=2 13.5 -> 10.5, N=8 21.2 -> 11.7), so only the one-GPU e2e leg uses it
        os.environ["LIME_PACK_DA"] = "0"
    ctx = LimeContext(local)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    be = multi.CudaBackend(ctx, torch, dev)
    rows, v, m = _lineage(w)
    parity = _multi_parity(args, w, world, rank, dev, dist, torch, ctx, be) if not args.no_parity else {"checked": False, "why": "--no-parity"}

    shards, n_reads, n_strand_total = _make_shards(w, world, rank, dev)
    torch.cuda.synchronize()
    S_local = shards[0]["n_own"]
    # LIME_MULTI=exchange forces the exact triples route (partial matrices + all-to-all); default: the record exchange
    mode = os.environ.get("LIME_MULTI", "xchg")
    density = float(os.environ.get("LIME_XCHG_DENSITY", "0.25"))
    job = None

    def make_job(density):
        if mode == "exchange":
            return multi.ShardedJob(dist, torch, be, rank, world, n_reads, w["n_refs"])
        flat_cap, inbox_cap = multi.xchg_capacities(S_local + HALO, world, density)
        return multi.XchgJob(dist, torch, be, rank, world, n_reads, w["n_refs"], w["n_strands"], flat_cap, inbox_cap)
    free0 = torch.cuda.mem_get_info()[0]
    job = make_job(density)
    route = {"name": "xchg" if mode != "exchange" else "exchange"}
    ctx2 = job2 = None

    def step(src=None):
        nonlocal job
        for attempt in range(3):
            try:
                job.stats = []
                if hasattr(job, "begin"):
                    job.begin()
                for s, sh in enumerate(src or shards):
                    job.run_strand(s, sh, w["alpha"], w["beta"], w["ebwt"], w["read_len"])
                # every rank fetches the assignments of the reads it owns (9 B per read, as at N = 1)
                return job.finish(v, m, w["tax_rank"], w["higher"], gather=False)
            except multi.CarryDetected as e:
                job.close()
                if e.overflow and attempt == 0:
                    job = make_job(density * 4)            # lists too small for this input: once more, larger
                    route["regrown"] = True
                else:
                    job = multi.ShardedJob(dist, torch, be, rank, world, n_reads, w["n_refs"])
                    route["name"] = "exchange (fallback: %s)" % e
        raise SystemExit("bench.py: no route completed the job")

    def steps(k_steps, src=None):
        """k_steps whole jobs; with a second context two of them are in flight."""
        nonlocal job2
        out = None
        if job2 is None or job2.closed or getattr(job, "closed", True):
            for _ in range(k_steps):
                out = step(src)
            return out
        try:
            return multi.run_jobs_in_flight((job, job2), k_steps, src or shards, w["alpha"], w["beta"], w["ebwt"], w["read_len"],
                                            v, m, w["tax_rank"], w["higher"], gather=False)
        except multi.CarryDetected as e:
            # not expected on this workload: back to one job at a time (step() then takes the exact route)
            torch.cuda.synchronize()
            job2.close()
            route["inflight_fallback"] = str(e)
            for _ in range(k_steps):
                out = step(src)
            return out

    for _ in range(max(args.warmup, 3)):
        out = step()
    # Two jobs in flight (record exchange only): a second context with its own slabs, lists, inboxes and side stream, so
    # that the scoring of job k + 1 is queued before the host waits for job k -- the routing / applying / clusterChoose of
    # job k's last strand then run beside it.  Only when every rank has the memory for it (what the first job took, and a
    # quarter more); LIME_BENCH_INFLIGHT=1: one job at a time.
    if mode != "exchange" and not getattr(job, "closed", True) and os.environ.get("LIME_BENCH_INFLIGHT", "2") != "1":
        torch.cuda.synchronize()
        free1 = torch.cuda.mem_get_info()[0]
        fits = torch.tensor([int(free1 > 1.25 * max(free0 - free1, 0) + (2 << 30))], device=dev, dtype=torch.int64)
        dist.all_reduce(fits, op=dist.ReduceOp.MIN)
        if int(fits.item()) == 1:
            ctx2 = LimeContext(local)
            ctx2.set_stream(stream.cuda_stream)
            be2 = multi.CudaBackend(ctx2, torch, dev)
            job2 = multi.XchgJob(dist, torch, be2, rank, world, n_reads, w["n_refs"], w["n_strands"], job.flat_cap, job.inbox_cap)
        else:
            route["one_job_in_flight_because"] = "no memory for a second context on some rank"
    if job2 is not None:
        out_a = tuple(np.array(x, copy=True) for x in out)
        out_b = steps(2)                                     # the second job of the pair ran on the second context
        same2 = torch.tensor([int(all(a.tobytes() == np.asarray(b).tobytes() for a, b in zip(out_a, out_b)))], device=dev, dtype=torch.int64)
        dist.all_reduce(same2, op=dist.ReduceOp.MIN)
        if int(same2.item()) != 1:
            raise SystemExit("bench.py: the second context's job differs from the first one's")
        steps(2)
    route["jobs_in_flight"] = 2 if job2 is not None and not job2.closed else 1
    cnt = torch.tensor([int((np.asarray(out[0]) == ord(c)).sum()) for c in "CUAH"], device=dev, dtype=torch.int64)
    dist.all_reduce(cnt)
    kinds = {c: int(x) for c, x in zip("CUAH", cnt.tolist()) if x}
    stats = list(job.stats)

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
        time.sleep(0.3)
    ctxs = [ctx] + ([ctx2] if ctx2 is not None else [])
    for c in ctxs:
        c.prof_enable(True); c.prof_reset()
    l0 = sum(c.launch_count() for c in ctxs)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.time()
    ev0.record(stream)
    steps(args.steps)
    ev1.record(stream)
    dist.barrier(); torch.cuda.synchronize()
    ms_prof = ev0.elapsed_time(ev1)
    launches = torch.tensor([sum(c.launch_count() for c in ctxs) - l0], device=dev, dtype=torch.int64)
    dist.all_reduce(launches, op=dist.ReduceOp.SUM)
    prof = {}
    for c in ctxs:
        for k_, x in c.prof().items():
            y = prof.setdefault(k_, {"ms": 0.0, "launches": 0})
            y["ms"] += x["ms"]; y["launches"] += x["launches"]
        c.prof_enable(False)
    # the same K steps without the per-launch events of the kernel profile
    dist.barrier(); torch.cuda.synchronize()
    ev0.record(stream)
    steps(args.steps)
    ev1.record(stream)
    dist.barrier(); torch.cuda.synchronize()
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64)      # the reported K steps
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)                     # max over ranks
    ms_step = float(ms.item()) / args.steps
    k12_all = torch.zeros(world, device=dev, dtype=torch.float64)
    k12_all[rank] = prof.get("k12_fused", {}).get("ms", 0.0) / max(prof.get("k12_fused", {}).get("launches", 1), 1)
    dist.all_reduce(k12_all)
    n_ext = torch.tensor([max(0, int(0.7 / max(ms_step * 1e-3, 1e-4)) - 2 * args.steps)], device=dev, dtype=torch.int64)
    dist.broadcast(n_ext, 0)                                       # same number of extra steps on every rank
    steps(int(n_ext.item())) if int(n_ext.item()) else None
    dist.barrier(); torch.cuda.synchronize()
    clk = clocks.summary(t0, time.time()) if rank == 0 else None
    if rank == 0:
        clk["window"] = "timed region + untimed repeats of the same step up to 0.7 s (nvidia-smi -lms 50)"
        clocks.stop()

    # one diagnostic step with barrier + synchronise around every phase (not part of the timed region): where the
    # step goes on the slowest rank, and the NVLink rate of the routing kernel
    tm = _SyncTimer(torch, dist)
    job.timer = tm
    if hasattr(job, "_inner"):
        job._inner.timer = tm
    step()
    job.timer = lambda name: multi._NoTimer()
    if hasattr(job, "_inner"):
        job._inner.timer = job.timer
    PH = ("compute", "export", "counts", "all_to_all", "import", "apply", "choose", "scalars", "classify", "gather", "check")
    phases = torch.tensor([tm.ms.get(k, 0.0) for k in PH], device=dev, dtype=torch.float64)
    dist.all_reduce(phases, op=dist.ReduceOp.MAX)
    sent = sum(st.get("triples_sent", 0) for st in job.stats) * 8
    upd = sum(st.get("updates", 0) for st in job.stats)            # all ranks, all strands
    sent_t = torch.tensor([sent], device=dev, dtype=torch.int64)
    dist.all_reduce(sent_t, op=dist.ReduceOp.MAX)
    route_ms = prof.get("x_route", {}).get("ms", 0.0) / max(prof.get("x_route", {}).get("launches", 1), 1)

    # ---- e2e: the same step with every rank's shards in page-locked HOST memory (lime_load(LIME_MEM_HOST) per
    # strand inside the timed region, lcp packed to bits on host threads) and each rank's assignments copied
    # back to its host, max over ranks.
    e2e = None
    host_bytes = sum((8 + w["ebwt"]) * (sh["n_own"] + sh["n_halo"]) for sh in shards)
    if not args.no_e2e and os.environ.get("LIME_BENCH_MULTI_E2E", "1") != "0" and host_bytes <= (24 << 30):
        ok, why, host_shards = 1, "", []
        try:
            for sh in shards:
                hs = dict(sh)
                for k in ("ebwt", "lcp", "da"):
                    if sh[k] is None:
                        continue
                    t = torch.empty(sh[k].shape, dtype=sh[k].dtype, pin_memory=True)
                    t.copy_(sh[k])
                    hs[k] = t
                host_shards.append(hs)
            torch.cuda.synchronize()
        except BaseException as ex:                      # noqa: BLE001
            ok, why = 0, "%s: %s" % (type(ex).__name__, ex)
        flag = torch.tensor([ok], device=dev, dtype=torch.int64)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 1:
            if w["alpha"] <= 255 and not args.no_narrow:
                for c in ctxs:
                    c.set_alpha(w["alpha"])               # lcp crosses as the bit lcp < alpha
            try:
                out_d = tuple(np.array(x, copy=True) for x in step())
                out_h = tuple(np.array(x, copy=True) for x in step(src=host_shards))
                same = torch.tensor([int(all(a.tobytes() == b.tobytes() for a, b in zip(out_d, out_h)))], device=dev,
                                    dtype=torch.int64)
                dist.all_reduce(same, op=dist.ReduceOp.MIN)
                e2e_steps = max(1, args.steps)
                dist.barrier(); torch.cuda.synchronize()
                ev0.record(stream)
                tw0 = time.perf_counter()
                for _ in range(e2e_steps):                # one job at a time: the host threads are the bound here
                    step(src=host_shards)
                ev1.record(stream)
                torch.cuda.synchronize()
                tw1 = time.perf_counter()
                dist.barrier()
            finally:
                for c in ctxs:
                    c.set_alpha(0)
            e_ms = torch.tensor([max(ev0.elapsed_time(ev1), (tw1 - tw0) * 1e3)], device=dev, dtype=torch.float64)
            dist.all_reduce(e_ms, op=dist.ReduceOp.MAX)
            e_ms = float(e_ms.item()) / e2e_steps
            narrow = w["alpha"] <= 255 and not args.no_narrow
            da_b = 3 if narrow and n_reads + w["n_refs"] <= (1 << 24) and os.environ.get("LIME_PACK_DA", "1") != "0" else 4
            h2d = sum((da_b + w["ebwt"]) * (hs["n_own"] + hs["n_halo"]) + ((hs["n_own"] + hs["n_halo"] + 7) // 8 if narrow else 4 * (hs["n_own"] + hs["n_halo"]))
                      for hs in host_shards)
            h2d_t = torch.tensor([h2d], device=dev, dtype=torch.int64)
            dist.all_reduce(h2d_t)
            if int(same.item()) == 1:
                e2e = {"value": n_strand_total * w["n_strands"] / (e_ms * 1e-3), "unit": UNIT,
                       "h2d_bytes_per_step": int(h2d_t.item()), "d2h_bytes_per_step": 9 * n_reads,
                       "ms_per_step": e_ms, "steps": e2e_steps, "reads_per_s": n_reads / (e_ms * 1e-3),
                       "h2d_GBps_per_gpu": int(h2d_t.item()) / world / (e_ms * 1e-3) / 1e9,
                       "note": "every rank: lime_load(LIME_MEM_HOST) of its shard of each strand from page-locked host arrays"
                               + (" (lcp packed to the bit lcp < alpha%s by host threads inside the timed region)" % (", da to 3 bytes per position" if da_b == 3 else "") if narrow else "")
                               + " + its assignments copied back; max over ranks; results equal to the resident step"}
            else:
                e2e = {"error": "host-buffer step disagrees with the resident step"}
        else:
            e2e = {"error": "host-shard pre-flight failed on some rank" + (": " + why if why else "")}
        del host_shards
    elif not args.no_e2e:
        e2e = {"value": None, "note": "skipped (%.0f GB of page-locked host arrays per rank)" % (host_bytes / 1e9)}

    if rank == 0:
        symbols = n_strand_total * w["n_strands"]
        peak_gbs, peak_src = _peak()
        kern = {k: x for k, x in prof.items() if x["launches"]}
        top = max(kern, key=lambda k: kern[k]["ms"])
        alg_per_launch = (8 + w["ebwt"]) * S_local
        avg_ms = kern[top]["ms"] / kern[top]["launches"]
        achieved = alg_per_launch / (avg_ms * 1e-3) / 1e9 if top == "k12_fused" else None
        ph = {k: float(x) for k, x in zip(PH, phases.tolist()) if x > 0}
        par = (f"position-sharded x{world}, halo {HALO}; similarity rows live on the read owner's GPU only: the fused kernel appends "
               "update records, a routing kernel sorts them by owner and stores each run into the owner's inbox over NVLink / "
               "NVSwitch peer memory (CUDA IPC) followed by its counts and a flag, the owner polls the flags on the device and "
               "adds the records; no collective per strand, routing / applying / clusterChoose of strand s run on a side stream "
               "beside the scoring of strand s + 1; one all-gather of the scalars per job; %d job(s) in flight (a second context, so that "
               "the scoring of job k + 1 is queued before the host waits for job k)" % route.get("jobs_in_flight", 1) if route["name"] == "xchg" else
               f"position-sharded x{world}, halo {HALO}, partial matrices + triples all-to-all by read owner ({route['name']})")
        remote = upd * (world - 1) // world
        line = {
            "metric": METRIC, "value": symbols / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": w["scaling"],
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": w["name"] + (f"; x{world} weak scaling: {world} position shards of {S_local} symbols per strand, {n_reads} reads"
                                                if w["scaling"] == "weak" else f"; strong scaling over {world} GPUs: {S_local} symbols per strand and GPU"),
                       "config": args.config, "alpha": w["alpha"],
                       "l2": "inputs (%.1f GB per strand and GPU) are larger than L2; no flush needed" % ((8 + w["ebwt"]) * S_local / 1e9),
                       "parallelism": par + "; every rank copies the assignments of its own read range to its host",
                       "symbols_per_step": symbols, "reads_per_step": n_reads, "strand_stats": stats, "assignments": kinds},
            "reads_per_s": n_reads / (ms_step * 1e-3),
            "parity": parity,
            "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak_gbs, "unit": "GB/s",
                         "frac": (achieved / peak_gbs) if achieved else None, "traffic": None, "peak_source": peak_src,
                         "alg_bytes_per_launch": alg_per_launch, "avg_launch_ms": avg_ms,
                         "kernels_ms_per_step_rank0": {k: x["ms"] / args.steps for k, x in kern.items()},
                         "k12_ms_per_launch_by_rank": [round(float(x), 3) for x in k12_all.tolist()],
                         "job_frac_of_aggregate_peak": (8 + w["ebwt"]) * symbols / (ms_step * 1e-3) / 1e9 / (peak_gbs * world)},
            "exchange": {"route": route, "phases_ms_max_over_ranks": ph,
                         "note": "phase times from one extra step with a barrier and a synchronise around every phase (which "
                                 "serialises the side stream: 'apply' = routing wait + adds, no overlap with the next strand)",
                         "ms_per_step_of_the_profiled_run_rank0": ms_prof / args.steps,
                         "updates_per_step_all_ranks": upd,
                         "remote_record_bytes_per_rank_and_step": 8 * remote // world if route["name"] == "xchg" else None,
                         "route_kernel_ms_per_launch_rank0": route_ms if route["name"] == "xchg" else None,
                         "route_kernel_nvlink_GBps_rank0": (8 * remote / world / max(len(stats), 1) / (route_ms * 1e-3) / 1e9)
                         if route["name"] == "xchg" and route_ms > 0 else None,
                         "all_to_all_bytes_sent_per_rank_max": int(sent_t.item()) or None,
                         "nvlink_peak_GBps_per_direction": 900},
            "cpu_baseline": None, "e2e": e2e, "gpu_launches": int(launches.item()), "clocks": clk,
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    if job2 is not None and not job2.closed:
        job2.close()
    if hasattr(job, "close"):
        job.close()
    if ctx2 is not None:
        ctx2.close()
    ctx.close()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg2", choices=sorted(CONFIGS))
    ap.add_argument("--n", type=int, default=0, help="positions per strand instead of the config's (config 5 sweep)")
    ap.add_argument("--alpha", type=int, default=0, help="alpha instead of the config's (config 5 sweep)")
    ap.add_argument("--cpu-sample", type=int, default=0,
                    help="positions per strand the CPU reference runs on (default: the whole workload up to 2^29, a prefix beyond)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e", action="store_true", help="run the e2e leg even when it needs > 24 GB of page-locked memory")
    ap.add_argument("--e2e-plain", action="store_true", help="also time the e2e leg without the narrowed lcp transfer")
    ap.add_argument("--no-narrow", action="store_true", help="e2e without the narrowed lcp transfer")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the sharded-vs-single-GPU check before timing")
    ap.add_argument("--small", action="store_true", help="reduced shape for functional testing only")
    ap.add_argument("--unfused", action="store_true", help="cluster list between steps 1 and 2 (for comparison)")
    args = ap.parse_args()
    w = dict(CONFIGS[args.config])
    if args.n:
        w["n_reads"] = max(1000, int(w["n_reads"] * (args.n / w["n"])))
        w["n"] = args.n
        w["name"] += f" [sweep point: {args.n} symbols per strand, {w['n_reads']} reads]"
    if args.alpha:
        w["alpha"] = args.alpha
        w["name"] += f" [alpha {args.alpha}]"
    if args.small:
        w = _small(w, 4_000_000 if w["scaling"] == "weak" else 8_000_000, 20_000)
    if not args.cpu_sample:
        args.cpu_sample = w["n"] if w["n"] <= (1 << 29) else (1 << 28)
    args.cpu_sample = min(args.cpu_sample, w["n"])
    if args.impl == "reference":
        run_reference_arm(args, w)
    else:
        run_ours(args, w)


if __name__ == "__main__":
    main()
