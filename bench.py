#!/usr/bin/env python
"""bench.py — pair-HMM GCUPS of the B200 path on BASELINE.json's configs[1], with the reference CPU aligner beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

One "step" = one ``poa_realign`` sweep of the aligner over one synthetic chunk (100 kb contig,
60x, 20 kb ONT-like reads at ~10 % error, r94-g360 RLE model): banded forward + backward + posterior tuples of all
three lists for every read.  Metric: GCUPS = 1e-9 x credited cell-passes / s, where an alignment with C band
cells is credited 2*C cell-passes (forward + backward) however many the kernels execute.

* ``value``  inputs already resident in HBM, timed with CUDA events on the library's stream, L2 flushed between steps.
* ``e2e``    the same sweep through the C-ABI call a host program makes (``margin_b200_posteriors_batch``) with HOST
             buffers: H2D of symbols/anchors and D2H of the tuple lists inside the timed region.  ``e2e.value`` is the
             throughput with three host threads per GPU pushing chunk after chunk (how margin's OpenMP chunk loop
             drives the aligner); ``e2e.single_call`` is one thread issuing one call at a time.
* N > 1      one process per GPU (torchrun), every rank sweeps its own chunk (chunks share no state, so there is no
             data-path collective; weak scaling), time = max over ranks, value = sum of the ranks' cell-passes / time.
* ``--impl reference``  the UNMODIFIED reference aligner (oracle/_ref, built from /root/reference by oracle/Makefile)
             on the host cores, OpenMP over reads, on a bounded sample of the same chunk.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "pair-HMM GCUPS (banded forward+backward+posterior, poa_realign sweep)"
UNIT = "GCUPS"
DEFAULT_WORKLOAD = "c2_ont_100kb_60x_20kb"
MODEL_FILE = os.path.join(ROOT, "tests", "golden", "model_r94g360.bin")
FLOP_PER_CELL_PASS = 108.0
FLOP_PER_POSTERIOR_CELL = 30.0 + 14.0


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            d = json.load(fh)
        return dict(hbm_gbs=float(d.get("hbm_gbs", 6650.0)), sm_max_mhz=float(d.get("sm_max_mhz", 1965.0)),
                    source="measured")
    return dict(hbm_gbs=6650.0, sm_max_mhz=1965.0, source="fallback")


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            self._stop_evt.wait(0.02)

    def finish(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def build_items(workload, seed, scale=1.0):
    from margin_b200 import model, synth
    pm = model.load_flat_model(MODEL_FILE)
    items = synth.workload(workload, pm, seed=seed, scale=scale)
    return pm, items


def dist_setup(n_gpus):
    """torch.distributed only as plumbing for the barrier and the max-over-ranks (no data-path collective)."""
    world = env_int("WORLD_SIZE", 1)
    rank = env_int("RANK", 0)
    local = env_int("LOCAL_RANK", 0)
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_mod
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        kw = {"device_id": torch.device("cuda", local)} if backend == "nccl" else {}
        dist_mod.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
        dist = dist_mod
    return rank, local, world, dist


ALL_CORES = None


def unbind_cores():
    """Back to every core of the box (the `margin polish` run of rank 0 uses all of them, like the reference's)."""
    if ALL_CORES:
        try:
            os.sched_setaffinity(0, ALL_CORES)
        except OSError:
            pass


def bind_rank_to_cores(local, world):
    """One process per GPU: every rank keeps its host threads (and therefore the page-locked buffers they allocate) on a
    disjoint slice of the box's cores, so that eight ranks do not all queue on one NUMA node's memory (SCALE r01: end-to-end
    efficiency 0.58 at N=8 with unbound ranks).  Returns the cores this rank runs on."""
    global ALL_CORES
    try:
        avail = sorted(os.sched_getaffinity(0))
    except AttributeError:
        return None
    ALL_CORES = avail
    if world <= 1 or len(avail) < 2 * world or os.environ.get("MARGIN_B200_BENCH_NO_BIND"):
        return avail
    per = len(avail) // world
    mine = avail[local * per:(local + 1) * per]
    try:
        os.sched_setaffinity(0, mine)
    except OSError:
        return avail
    return mine


def reduce_scalar(dist, value, op, device):
    if dist is None:
        return value
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
    return float(t.item())


# ----------------------------------------------------------------------------------------------- reference arm

def time_reference(pm, items, steps, warmup, threads, sample_reads=None):
    """The reference's own CPU aligner (getAlignedPairsWithIndelsUsingAnchors per read, OpenMP over reads like the
    chunk loop of polish.c:476) from oracle/_ref; falls back to the C restatement (kind 'port') when the reference
    build did not travel.  Returns dict(value GCUPS, seconds per step, cells per step, kind, sample)."""
    from margin_b200 import abi
    from oracle import oracle_py
    n = len(items)
    k = min(n, sample_reads or max(32, 16 * threads))
    sample = items[:k]
    batch = abi.Batch(sample)
    orc = oracle_py.Oracle()
    # band cells of the sample: counted by the restatement's band builder (pure integer, same as the reference's)
    cells = 0
    for it in sample:
        kl, kr = orc.band(it[2], len(it[0]), len(it[1]), pm.p.dynamic_anchor_expansion, pm.p.diagonal_expansion)
        cells += int(((kr - kl) // 2 + 1).sum())
    if oracle_py.Reference.available():
        ref = oracle_py.Reference()
        sms = [ref.sm_from_model(pm.models[s]) for s in (0, 1)]
        run = lambda: ref.posteriors_batch(sms, pm.p, batch, 1, threads)
        kind = "reference"
    else:
        run = lambda: orc.posteriors_batch(pm.models, pm.p, batch, 1, threads, keep=False)
        kind = "port"
    for _ in range(warmup):
        run()
    t0 = time.perf_counter()
    for _ in range(steps):
        run()
    dt = (time.perf_counter() - t0) / max(steps, 1)
    return dict(value=2.0 * cells / dt * 1e-9, sec_per_step=dt, cells=cells, kind=kind, cores=threads,
                sample="%d of %d reads of the chunk (%d band cells), %d OpenMP threads" % (k, n, cells, threads))


def main_reference(args):
    rank, local, world, dist = dist_setup(args.gpus)
    if rank != 0:
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return 0
    from margin_b200 import model
    pm = model.load_flat_model(MODEL_FILE)
    items = build_chunk(pm, args.seed, 0)  # chunk 0 of the bench contig = BASELINE configs[1]
    threads = host_cores()
    r = time_reference(pm, items, args.steps, args.warmup, threads)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["sec_per_step"] * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "seed": args.seed, "params": "ont/r9.4 r94-g360 (RLE, repeat counts)",
                   "l2": "n/a (host)"},
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                         "sample": r["sample"]},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if args.polish_synth > 0:  # the reference's own CLI on the same synthetic contig as our arm's `polish` object
        line["polish"] = time_polish_synth(args.polish_synth, min(threads, args.polish_threads), args.seed, only_reference=True)
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


# ---------------------------------------------------------------------------------------------------- our arm

# Chunk lengths of the bench contig, cycled: chunks differ in size (and seed), so arenas, tile cuts and transfer sizes
# change from step to step the way they do in a real run (polish.c:476 hands every thread chunk after chunk).
CHUNK_LENGTH_FACTORS = (1.0, 0.8, 1.3, 0.9)
CHUNK_LEN = 100000


def chunk_length(k):
    return int(CHUNK_LEN * CHUNK_LENGTH_FACTORS[k % len(CHUNK_LENGTH_FACTORS)])


def build_chunk(pm, seed, k):
    """Chunk k of the synthetic contig (BASELINE configs[1] shape: 60x 20 kb ONT-like reads at ~10 % error; chunk 0 is
    exactly the 100 kb chunk of configs[1] with the bench seed)."""
    from margin_b200 import synth
    ch = synth.make_chunk(seed + 1000 * k, chunk_length(k), 60, 20000, synth.ONT, 0.01, int(pm.p.diagonal_expansion))
    return ch.items(pm)


def time_polish(threads):
    """End-to-end `margin polish` on the bundled ONT region (BASELINE configs[0]; 0.1 Mb of contig): the reference's
    CLI linked against this library (oracle/_ref/margin_b200cli) and the unmodified reference (oracle/_ref/margin_ref)
    with -t threads each.  Both binaries and the data are built/copied by oracle/Makefile where /root/reference
    exists and travel to the GPU box; returns None when they did not."""
    import hashlib
    import subprocess
    import tempfile
    ref = os.path.join(ROOT, "oracle", "_ref")
    data = os.path.join(ref, "data")
    bam = os.path.join(data, "HG002.r94g360.chr20_59M_100k.bam")
    if not (os.path.exists(os.path.join(ref, "margin_b200cli")) and os.path.exists(bam)):
        return None
    out = {"region": "HG002 chr20:59M 100 kb, ont r9.4 g360 params", "contig_mb": 0.1, "threads": threads}
    with tempfile.TemporaryDirectory() as tmp:
        for key, binary in (("reference", "margin_ref"), ("b200", "margin_b200cli")):
            if not os.path.exists(os.path.join(ref, binary)):
                continue
            prefix = os.path.join(tmp, key)
            cmd = [os.path.join(ref, binary), "polish", bam, os.path.join(data, "hg38.chr20_59M_100k.fa"),
                   os.path.join(data, "params", "polish", "ont", "r9.4", "allParams.np.human.r94-g360.json"),
                   "-t", str(threads), "-o", prefix]
            t0 = time.perf_counter()
            r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
            dt = time.perf_counter() - t0
            if r.returncode != 0:
                out[key] = {"error": r.stdout[-300:]}
                continue
            h = hashlib.md5()
            with open(prefix + ".fa", "rb") as fh:
                h.update(fh.read())
            md5 = h.hexdigest()
            out[key] = {"wall_s": dt, "mb_per_hour": 0.1 / dt * 3600.0, "fasta_md5": md5}
    if "reference" in out and "b200" in out and "fasta_md5" in out["reference"] and "fasta_md5" in out["b200"]:
        out["byte_identical"] = out["reference"]["fasta_md5"] == out["b200"]["fasta_md5"]
    return out


def time_polish_synth(contig_kb, threads, seed=42, skip_reference=False, diploid=False, depth=60, platform="ont", gpus=None,
                      precision=None, only_reference=False):
    """End-to-end `margin polish` on a SYNTHETIC contig of BASELINE configs[1]/[2]'s shape (60x ONT-like 20 kb reads at
    ~10 % error, r94-g360 params, 100 kb chunks; margin_b200/synth.py: write_polish_dataset + oracle/_ref/sam2bam):
    the unmodified reference CLI and the CLI linked against this library, -t `threads` each (the reference
    parallelises over chunks only, polish.c:476; the binding adds one of the visible GPUs per 16 threads of the team,
    compat/pairwiseAligner_b200.c: b200_device_of_thread).  Mb/hour counts
    the whole CLI run: BAM parsing, POA, bubble graph, stitching and FASTA output included."""
    import hashlib
    import subprocess
    import tempfile
    from margin_b200 import synth
    ref = os.path.join(ROOT, "oracle", "_ref")
    if not (os.path.exists(os.path.join(ref, "margin_b200cli")) and os.path.exists(os.path.join(ref, "sam2bam"))):
        return None
    params = os.path.join(ref, "data", "params", "polish", "ont", "r9.4", "allParams.np.human.r94-g360.json")
    read_len, err, draft_err, what = 20000, synth.ONT, 0.01, "20 kb ONT-like reads (~10 % error), ont r9.4 g360 params"
    if platform == "hifi":  # BASELINE configs[3]: 15 kb reads at ~1 % error, pb-hifi params (50 kb chunks, narrow bands)
        params = os.path.join(ref, "data", "params", "polish", "pacbio", "hifi", "allParams.hifi.json")
        read_len, err, draft_err, what = 15000, synth.HIFI, 0.001, "15 kb HiFi-like reads (~1 % error), pacbio hifi params"
    with tempfile.TemporaryDirectory() as tmp:
        t0 = time.perf_counter()
        fa, sam, n_reads, draft_len = synth.write_polish_dataset(tmp, seed, contig_kb * 1000, depth, read_len, err, draft_err,
                                                                 het=synth.HET if diploid else None)
        bam = os.path.join(tmp, "reads.bam")
        subprocess.run([os.path.join(ref, "sam2bam"), sam, bam], check=True, stderr=subprocess.DEVNULL)
        os.remove(sam)
        out = {"workload": "synthetic %d kb %scontig, %dx %s%s"
                           % (contig_kb, "diploid " if diploid else "", depth, what, ", margin polish --diploid" if diploid else ""),
               "contig_mb": draft_len / 1e6, "reads": n_reads, "threads": threads, "seed": seed,
               "dataset_s": time.perf_counter() - t0}
        arms = (("b200", "margin_b200cli"),) if skip_reference else (("reference", "margin_ref"), ("b200", "margin_b200cli"))
        if only_reference:
            arms = (("reference", "margin_ref"),)
        env = dict(os.environ)
        if gpus:  # the binding maps OpenMP thread t to device t % (visible devices)
            env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in range(gpus))
            out["gpus"] = gpus
            out["devices_used"] = min(gpus, max(1, -(-threads // int(os.environ.get("MARGIN_B200_THREADS_PER_GPU", "16")))))
        if precision:
            env["MARGIN_B200_PRECISION"] = precision
            out["precision"] = precision
        for key, binary in arms:
            prefix = os.path.join(tmp, key)
            cmd = [os.path.join(ref, binary), "polish", bam, fa, params, "-t", str(threads), "-o", prefix]
            if diploid:
                cmd.append("-2")  # phasing bubble graphs, then one polished FASTA per haplotype (<prefix>.fa.hap1/.hap2)
            t0 = time.perf_counter()
            r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=3000, env=env)
            dt = time.perf_counter() - t0
            if r.returncode != 0:
                out[key] = {"error": r.stdout[-300:]}
                continue
            h = hashlib.md5()  # diploid: both haplotype FASTAs in one digest
            for name in ([prefix + ".fa.hap1", prefix + ".fa.hap2"] if diploid else [prefix + ".fa"]):
                with open(name, "rb") as fh:
                    h.update(fh.read())
            md5 = h.hexdigest()
            out[key] = {"wall_s": dt, "mb_per_hour": draft_len / 1e6 / dt * 3600.0, "fasta_md5": md5}
    if "fasta_md5" in out.get("reference", {}) and "fasta_md5" in out.get("b200", {}):
        out["byte_identical"] = out["reference"]["fasta_md5"] == out["b200"]["fasta_md5"]
    return out


def abi_mod():
    from margin_b200 import abi
    return abi


def time_sustained(pm, batches, cells, device, precision, contexts=3, min_seconds=2.0):
    """`contexts` aligner contexts on one GPU, one host thread each; every thread pushes this rank's chunks, one after the
    other and round after round, through margin_b200_posteriors_batch (host buffers in, host lists out) until
    `min_seconds` have passed: how margin's OpenMP chunk loop drives the aligner (polish.c:476)."""
    from margin_b200 import abi
    from margin_b200.aligner import PairHmmAligner
    als = [PairHmmAligner(pm.models, pm.p, device=device) for _ in range(contexts)]
    for al in als:
        al.set_precision(precision)
        al.set_pair_format(abi.PAIRS_DELTA32)
    warm = threading.Barrier(contexts)

    def warm_up(al):  # every context sees every chunk once (arenas reach their steady size)
        for b in batches:
            al.posteriors_host(b, True)
        warm.wait()

    ws = [threading.Thread(target=warm_up, args=(al,)) for al in als]
    for t in ws:
        t.start()
    for t in ws:
        t.join()
    start = threading.Barrier(contexts + 1)
    errs, done = [], [0.0] * contexts
    calls = [0] * contexts
    stop = [False]

    def worker(i, al):
        try:
            start.wait()
            k = i
            while not stop[0]:
                al.posteriors_host(batches[k % len(batches)], True)
                done[i] += 2.0 * cells[k % len(batches)]
                calls[i] += 1
                k += 1
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    ths = [threading.Thread(target=worker, args=(i, al)) for i, al in enumerate(als)]
    for t in ths:
        t.start()
    start.wait()
    t0 = time.perf_counter()
    time.sleep(min_seconds)
    stop[0] = True
    for t in ths:
        t.join()
    dt = time.perf_counter() - t0
    for al in als:
        al.close()
    if errs:
        raise errs[0]
    n = sum(calls)
    return {"value": sum(done) / dt * 1e-9, "unit": UNIT, "contexts": contexts, "chunks": n,
            "ms_per_chunk": dt / max(n, 1) * 1e3, "wall_s": dt, "cell_passes": sum(done)}


def e2e_object(single_value, single_s, h2d, d2h, sustained):
    single = {"value": single_value, "unit": UNIT, "ms_per_step": single_s * 1e3, "host_threads": 1}
    buffers = "batch arrays in page-locked memory from margin_b200_pinned_alloc (as compat/pairwiseAligner_b200.c builds them)"
    if sustained is None:
        return dict(single, h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h, mode="single call", host_buffers=buffers)
    return {"value": sustained["value"], "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
            "ms_per_step": sustained["ms_per_chunk"],
            "mode": "%d host threads per GPU, every thread sweeps this rank's chunks in turn for %.1f s (%d calls)"
                    % (sustained["contexts"], sustained["wall_s"], sustained["chunks"]),
            "host_threads": sustained["contexts"], "host_buffers": buffers, "single_call": single}


def main_ours(args):
    import torch
    from margin_b200 import abi, model, scheduler
    from margin_b200.aligner import PairHmmAligner

    rank, local, world, dist = dist_setup(args.gpus)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    cores = bind_rank_to_cores(local, world)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    peaks = load_peaks()
    precision = abi.F32 if args.precision == "f32" else abi.F64

    # ---- the chunks: one synthetic contig of chunks_per_gpu x N chunks, sharded over the ranks the way the reference's
    # size-ordered dynamic schedule would (scheduler.shard_chunks: sizes are known from the chunk lengths alone, so every
    # rank derives the same assignment and generates only its own chunks; no data-path collective)
    pm = model.load_flat_model(MODEL_FILE)
    n_chunks = args.chunks_per_gpu * world
    mine = scheduler.shard_chunks([chunk_length(k) for k in range(n_chunks)], world)[rank]
    t_gen = time.perf_counter()
    chunk_items = {k: build_chunk(pm, args.seed, k) for k in mine}
    t_gen = time.perf_counter() - t_gen
    batches = [abi.Batch(chunk_items[k], compact=not args.plain_formats).to_pinned() for k in mine]
    al = PairHmmAligner(pm.models, pm.p, device=local)
    al.set_precision(precision)
    al.bench_workload = args.workload
    if not args.plain_formats:
        al.set_pair_format(abi.PAIRS_DELTA32)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def l2_flush():
        flush.zero_()
        torch.cuda.synchronize(dev)

    # ---- kernel-only leg: a chunk's inputs resident in HBM (upload outside the timed region), per-launch CUDA events
    al.set_kernel_timing(not args.no_kernel_events)
    for i in range(max(args.warmup, len(batches))):
        al.upload(batches[i % len(batches)])
        l2_flush()
        st = al.posteriors_resident(True)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize(dev)
    sampler = ClockSampler(local)
    sampler.start()
    dev_ms, launches, cells_done = 0.0, 0, 0
    ktimes, chunk_cells, stats_sum = {}, [0] * len(batches), {}
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        b = i % len(batches)
        al.upload(batches[b])
        l2_flush()
        st = al.posteriors_resident(True)
        dev_ms += st["gpu_ms"]
        launches += st["kernel_launches"]
        cells_done += st["band_cells"]
        chunk_cells[b] = st["band_cells"]
        for key in ("tiles", "wide_cells", "sequential_cells", "guard_seams", "guard_failures", "diagonals", "segments"):
            stats_sum[key] = stats_sum.get(key, 0) + int(st[key])
        for k, (ms, ln) in al.kernel_times().items():
            a = ktimes.setdefault(k, [0.0, 0])
            a[0] += ms
            a[1] += ln
    torch.cuda.synchronize(dev)
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.finish()
    if dist is not None:
        dist.barrier()
    for b in range(len(batches)):  # chunks the timed steps did not reach (steps < chunks)
        if not chunk_cells[b]:
            al.upload(batches[b])
            chunk_cells[b] = al.posteriors_resident(True)["band_cells"]
    ms_step = reduce_scalar(dist, dev_ms / args.steps, "MAX", dev)
    cell_passes_step = reduce_scalar(dist, 2.0 * cells_done / args.steps, "SUM", dev)
    value = cell_passes_step / (ms_step * 1e-3) * 1e-9

    al.set_kernel_timing(False)
    # ---- end-to-end leg: host buffers in, host tuple lists out, through the public C-ABI call
    for b in batches[:2]:
        res = al.posteriors_host(b, True)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize(dev)
    e2e_steps = max(len(batches), min(args.steps, 8))
    t0 = time.perf_counter()
    e2e_cells, d2h, h2d = 0, 0, 0
    for i in range(e2e_steps):
        b = i % len(batches)
        res = al.posteriors_host(batches[b], True)
        e2e_cells += chunk_cells[b]
        d2h += int(res["tuples"]) * int(res["tuple_bytes"]) + int(res["escapes"]) * 8 + 6 * (batches[b].n + 1) * 8
        h2d += batches[b].h2d_bytes
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    e2e_s = reduce_scalar(dist, e2e_s, "MAX", dev)
    e2e_value = reduce_scalar(dist, 2.0 * e2e_cells / e2e_steps, "SUM", dev) / e2e_s * 1e-9

    # ---- sustained end-to-end: three contexts on this GPU, each sweeping the rank's chunks from host buffers, so that
    # one chunk's copies overlap another's kernels (what `margin polish -t` does with several threads per GPU)
    sustained = None
    if not args.no_sustained:
        sustained = time_sustained(pm, batches, chunk_cells, local, precision, contexts=3, min_seconds=args.e2e_seconds)
        wall = reduce_scalar(dist, sustained["wall_s"], "MAX", dev)  # whole job: all ranks' chunks over the slowest rank
        sustained["value"] = reduce_scalar(dist, sustained.pop("cell_passes"), "SUM", dev) / wall * 1e-9

    # ---- roofline of the dominant kernel (CUDA events around each launch, same timed region)
    dom = max(ktimes.items(), key=lambda kv: kv[1][0])[0] if ktimes else "none"
    dom_ms = ktimes[dom][0] / max(ktimes[dom][1], 1) if ktimes else 0.0
    roof = roofline(dom, dom_ms, ktimes, cells_done / args.steps, args.steps, peaks, al)

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": al.arithmetic(), "data": "synthetic",
        "config": {"workload": args.workload, "seed": args.seed, "chunks_per_gpu": args.chunks_per_gpu,
                   "chunk_lengths_kb": [chunk_length(k) // 1000 for k in mine],
                   "reads_per_chunk": [b.n for b in batches], "band_cells_per_chunk": [int(c) for c in chunk_cells],
                   "sharding": "one contig of %d chunks, scheduler.shard_chunks (size-descending LPT), no collective" % n_chunks,
                   "params": "ont/r9.4 r94-g360 (RLE, repeat counts)", "l2": "flushed between steps (512 MiB write)",
                   "boundary_formats": "plain" if args.plain_formats else "symbols 1 B, anchors as 4 B runs, tuples 4 B step records + escapes",
                   "tiles_per_step": stats_sum.get("tiles", 0) / args.steps,
                   "wide_cell_share": stats_sum.get("wide_cells", 0) / max(cells_done, 1),
                   "sequential_cell_share": stats_sum.get("sequential_cells", 0) / max(cells_done, 1),
                   "guard_seams_per_step": stats_sum.get("guard_seams", 0) / args.steps,
                   "guard_failures": stats_sum.get("guard_failures", 0),
                   "wall_ms_per_step_incl_upload_and_flush": t_wall / args.steps * 1e3, "chunk_generation_s": t_gen,
                   "host_cores_of_rank0": len(cores) if cores else None},
        "clocks": clocks,
        # e2e.value is THROUGHPUT through margin_b200_posteriors_batch the way margin drives it (polish.c:476: several
        # host threads, one chunk each): 3 host threads per GPU sweeping distinct chunks for e2e_seconds, every chunk copied
        # in from page-locked host memory and its lists copied back.  `single_call` is one thread, one call at a time.
        "e2e": e2e_object(e2e_value, e2e_s, h2d // e2e_steps, d2h // e2e_steps, sustained),
        "gpu_launches": launches,
        "kernel_ms_per_step": {k: v[0] / args.steps for k, v in ktimes.items() if v[1]},
        "roofline": roof,
    }
    # ---- the FP32 fast mode beside the exact one (include/margin_b200.h: margin_b200_set_precision), same chunks, same
    # legs in short: the headline stays the reference's own arithmetic, this object says what the tolerance of
    # BASELINE.json's north_star (posteriors within 1e-4, threshold ties listed: tests/test_corpus.py) buys
    if precision == abi.F64 and not args.no_fast_mode:
        fast = PairHmmAligner(pm.models, pm.p, device=local)
        fast.set_precision(abi.F32)
        fast.set_pair_format(abi.PAIRS_DELTA32)
        fast.set_kernel_timing(True)
        for i in range(len(batches)):
            fast.upload(batches[i])
            fast.posteriors_resident(True)
        f_ms, f_cells, f_k = 0.0, 0, {}
        for i in range(args.steps):
            fast.upload(batches[i % len(batches)])
            l2_flush()
            st = fast.posteriors_resident(True)
            f_ms += st["gpu_ms"]
            f_cells += st["band_cells"]
            for k, (ms, ln) in fast.kernel_times().items():
                a = f_k.setdefault(k, [0.0, 0])
                a[0] += ms
                a[1] += ln
        f_ms_step = reduce_scalar(dist, f_ms / args.steps, "MAX", dev)
        f_value = reduce_scalar(dist, 2.0 * f_cells / args.steps, "SUM", dev) / (f_ms_step * 1e-3) * 1e-9
        f_dom = max(f_k.items(), key=lambda kv: kv[1][0])[0]
        f_roof = roofline(f_dom, f_k[f_dom][0] / max(f_k[f_dom][1], 1), f_k, f_cells / args.steps, args.steps, peaks, fast)
        fast.close()
        f_sus = None
        if not args.no_sustained:
            f_sus = time_sustained(pm, batches, chunk_cells, local, abi.F32, contexts=3, min_seconds=max(1.0, args.e2e_seconds / 2))
            wall = reduce_scalar(dist, f_sus["wall_s"], "MAX", dev)
            f_sus["value"] = reduce_scalar(dist, f_sus.pop("cell_passes"), "SUM", dev) / wall * 1e-9
        line["fast_mode"] = {"dtype": "f32", "value": f_value, "unit": UNIT, "ms_per_step": f_ms_step,
                             "kernel_ms_per_step": {k: v[0] / args.steps for k, v in f_k.items() if v[1]},
                             "e2e_value": f_sus["value"] if f_sus else None,
                             "roofline": {k: f_roof[k] for k in ("kernel", "bound", "achieved", "peak", "unit", "frac", "traffic")},
                             "parity": "posteriors within 1e-4 of the reference's except at logAdd piece boundaries "
                                       "(5 of 8.8 M tuples of a chunk up to 4.5e-4), tuple sets equal up to threshold ties "
                                       "(15 of 10 240 corpus cases), polished FASTA byte-identical (tests/test_corpus.py, "
                                       "tests/test_polish_e2e.py)"}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = time_reference(pm, chunk_items[mine[0]], 1, 1, host_cores())
        line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                                "sample": r["sample"]}
    al.close()
    for b in batches:
        b.release_pinned()
    # ---- the product-level number: polished Mb/hour of `margin polish` (reference CLI linked against this library) on a
    # synthetic contig, all N GPUs of the box visible to it; rank 0 runs it while the other ranks wait at the barrier
    if rank == 0 and args.polish_synth > 0:
        unbind_cores()
        line["polish"] = time_polish_synth(args.polish_synth, min(host_cores(), args.polish_threads), args.seed,
                                           skip_reference=True, gpus=world, precision=args.precision)
    if rank == 0 and world == 1 and args.polish:
        line["polish_bundled_region"] = time_polish(min(host_cores(), 8))
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture of this workload
NCU_TRAFFIC_SOURCE = ("profiles/r02_tiles_ncu_raw_selected.txt (ncu --set full, one launch of the 100 kb chunk of configs[1], "
                      "scripts/ncu_target.py)")
NCU_DRAM_BYTES_PER_LAUNCH = {  # keyed by (arithmetic, kernel class)
    ("f64", "forward"): 0.104837e9 + 1.132977e9,
    ("f64", "backward"): 1.305050e9 + 0.140474e9,
    ("f32", "forward"): 0.111883e9 + 0.779340e9,
    ("f32", "backward"): 0.953477e9 + 0.139526e9,
}


def al_workload(al):
    return getattr(al, "bench_workload", None)


def roofline(dom, dom_ms, ktimes, cells, steps, peaks, al):
    """Roofline of the dominant kernel.  The path is a log-space recurrence, so the pipe that bounds it is the
    non-tensor FP pipe of the arithmetic type the kernels compute in; the HBM view of the
    same launch is reported beside it."""
    import torch
    prop = torch.cuda.get_device_properties(torch.cuda.current_device())
    sms = prop.multi_processor_count
    dtype = al.arithmetic()
    lanes = 64 if dtype == "f64" else 128  # FP64 units per SM are half the FP32 lanes on sm_100
    peak_tflops = sms * lanes * 2 * peaks["sm_max_mhz"] * 1e6 * 1e-12
    steps = max(1, steps)
    # algorithmic work of one step attributed to the dominant kernel class.  HBM bytes per SURVEY.md §8 D-4: symbols
    # ~0.9 B, band descriptor ~2.7 B, output tuples ~2.6 B per band cell (forward state on chip: 0 B).  This design spills
    # the forward state once in each direction; those bytes are reported as `moved_bytes`, not as algorithmic bytes.
    f_spill = cells * al.f_bytes_per_cell()
    if dom == "forward":
        flop = cells * FLOP_PER_CELL_PASS
        bytes_alg = cells * (0.9 + 2.7)
        bytes_moved = bytes_alg + f_spill
    elif dom in ("backward", "fused"):
        flop = cells * ((2.0 if dom == "fused" else 1.0) * FLOP_PER_CELL_PASS + FLOP_PER_POSTERIOR_CELL)
        bytes_alg = cells * (0.9 + 2.7 + 2.6)
        bytes_moved = bytes_alg + (f_spill if dom == "backward" else 0.0)
    else:
        flop = 0.0
        bytes_alg = cells * 2.6
        bytes_moved = bytes_alg
    class_ms_step = ktimes[dom][0] / steps if ktimes else 0.0
    ach_tflops = flop / (class_ms_step * 1e-3) * 1e-12 if class_ms_step > 0 else 0.0
    ach_gbs = bytes_alg / (class_ms_step * 1e-3) * 1e-9 if class_ms_step > 0 else 0.0
    return {
        "kernel": dom, "bound": "fp64-pipe" if dtype == "f64" else "fp32-pipe", "achieved": ach_tflops,
        "peak": peak_tflops, "unit": "TFLOP/s", "frac": ach_tflops / peak_tflops if peak_tflops else None,
        "traffic": NCU_DRAM_BYTES_PER_LAUNCH.get((dtype, dom)), "traffic_source": NCU_TRAFFIC_SOURCE,
        "flop_per_cell": flop / max(cells, 1), "cells_per_launch": cells,
        "peak_source": "%d SMs x %d lanes x 2 x %.0f MHz (%s sm_max_mhz)" % (
            sms, lanes, peaks["sm_max_mhz"], peaks["source"]),
        "avg_launch_ms": dom_ms, "class_ms_per_step": class_ms_step,
        "hbm": {"achieved": ach_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": ach_gbs / peaks["hbm_gbs"], "algorithmic_bytes": bytes_alg, "moved_bytes": bytes_moved,
                "moved_gbs": bytes_moved / (class_ms_step * 1e-3) * 1e-9 if class_ms_step > 0 else 0.0,
                "f_spill_bytes_per_cell": al.f_bytes_per_cell(), "peak_source": peaks["source"]},
    }


def main_replay(args):
    """`--workload c1_replay`: every posterior batch `margin polish` sends to the library on the bundled ONT region
    (BASELINE configs[0]: round 1 with BAM-CIGAR anchors, round 2 with posterior anchors and per-anchor expansions, the
    stitcher's overlap alignment), dumped by the binding and replayed with inputs resident and through the host call.
    Reports GCUPS per batch, the share of band cells on the wide-tile and sequential paths, and the seam guard counts:
    what real data, unlike the synthetic chunk, does to the tiling."""
    import tempfile
    import torch
    from margin_b200 import abi, replay
    from margin_b200.aligner import PairHmmAligner
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    ref = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.exists(os.path.join(ref, "margin_b200cli")):
        print(json.dumps({"workload": "c1_replay", "unavailable": "oracle/_ref/margin_b200cli not built"}))
        return 0
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    rows, tot_cells, tot_ms, tot_e2e = [], 0, 0.0, 0.0
    launches = 0
    with tempfile.TemporaryDirectory() as tmp:
        t0 = time.perf_counter()
        dumps = replay.dump_bundled_region(os.path.join(tmp, "d"), ref)
        cli_s = time.perf_counter() - t0
        for path in dumps:
            items, models, params, want = replay.load_dump(path)
            batch = abi.Batch(items, compact=True)
            al = PairHmmAligner(models, params, device=0)
            al.set_pair_format(abi.PAIRS_DELTA32)
            al.upload(batch)
            for _ in range(args.warmup):
                st = al.posteriors_resident(bool(want))
            ms = 0.0
            for _ in range(args.steps):
                flush.zero_()
                torch.cuda.synchronize(dev)
                st = al.posteriors_resident(bool(want))
                ms += st["gpu_ms"]
                launches += st["kernel_launches"]
            ms /= args.steps
            al.posteriors_host(batch, bool(want))
            t0 = time.perf_counter()
            for _ in range(args.steps):
                al.posteriors_host(batch, bool(want))
            e2e_ms = (time.perf_counter() - t0) / args.steps * 1e3
            al.close()
            cells = int(st["band_cells"])
            lens = [len(it[1]) for it in items]
            rows.append({"batch": os.path.basename(path), "alignments": len(items), "want_indels": int(want),
                         "read_len_max": int(max(lens)) if lens else 0, "band_cells": cells, "tiles": int(st["tiles"]),
                         "wide_cell_share": st["wide_cells"] / max(cells, 1),
                         "sequential_cell_share": st["sequential_cells"] / max(cells, 1),
                         "guard_seams": int(st["guard_seams"]), "guard_failures": int(st["guard_failures"]),
                         "ms_resident": ms, "gcups_resident": 2.0 * cells / (ms * 1e-3) * 1e-9 if ms > 0 else 0.0,
                         "ms_host_call": e2e_ms, "gcups_host_call": 2.0 * cells / (e2e_ms * 1e-3) * 1e-9})
            tot_cells += cells
            tot_ms += ms
            tot_e2e += e2e_ms
    line = {"metric": METRIC, "value": 2.0 * tot_cells / (tot_ms * 1e-3) * 1e-9 if tot_ms else 0.0, "unit": UNIT, "n_gpus": 1,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "real (bundled HG002 chr20 100 kb region, batches dumped from margin polish)",
            "config": {"workload": "c1_replay", "batches": len(rows), "band_cells": tot_cells, "l2": "flushed between steps",
                       "margin_polish_wall_s": cli_s},
            "e2e": {"value": 2.0 * tot_cells / (tot_e2e * 1e-3) * 1e-9 if tot_e2e else 0.0, "unit": UNIT,
                    "mode": "one host thread, one call per batch, pageable numpy buffers"},
            "gpu_launches": launches, "replay": rows}
    print(json.dumps(line), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD)
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernel-events", action="store_true", help="no per-launch CUDA events in the resident leg")
    ap.add_argument("--plain-formats", action="store_true",
                    help="12-byte anchor triples in, 12-byte tuple records out instead of the compact boundary formats")
    ap.add_argument("--no-sustained", action="store_true", help="skip the two-context sustained end-to-end leg")
    ap.add_argument("--polish-synth", type=int, default=2000, metavar="KB",
                    help="polished Mb/hour of `margin polish` on a synthetic contig of KB kilobases (60x 20 kb ONT-like reads, "
                         "100 kb chunks; BASELINE configs[2] is 5000): the CLI linked against this library, or the unmodified "
                         "reference CLI with --impl reference; 0 = skip")
    ap.add_argument("--chunks-per-gpu", type=int, default=4, help="distinct chunks every rank sweeps (weak scaling)")
    ap.add_argument("--precision", default="f64", choices=["f64", "f32"],
                    help="arithmetic of the lane-group tile kernels (include/margin_b200.h: margin_b200_set_precision)")
    ap.add_argument("--e2e-seconds", type=float, default=2.0, help="timed window of the sustained end-to-end leg")
    ap.add_argument("--no-fast-mode", action="store_true", help="skip the FP32 fast-mode object of the line")
    ap.add_argument("--polish-threads", type=int, default=64, help="-t of the margin polish runs (capped at the host's cores)")
    ap.add_argument("--polish-skip-reference", action="store_true")
    ap.add_argument("--polish", action="store_true",
                    help="also time `margin polish` on the bundled region: reference CLI vs the CLI linked against "
                         "this library (oracle/_ref, built where /root/reference exists)")
    args = ap.parse_args()
    if args.impl == "reference":
        return main_reference(args)
    if args.workload == "c1_replay":
        return main_replay(args)
    return main_ours(args)


if __name__ == "__main__":
    sys.exit(main())
