#!/usr/bin/env python
"""bench.py — Seq simulated bases/s on B200 (BASELINE.json metric), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--molecules M]

A step = one pass of the Seq hot path (materialise -> Badread errors -> qscores -> FASTQ) over one
batch of synthetic molecules.  Workload at every N: BASELINE.json configs[1], the README bulk
pipeline shape Tsb->Trc->plA->Flp->Tag->Seq, 1 M molecules per GPU on a synthetic 3.1 Gbp
GRCh38-sized genome (weak scaling: every rank simulates its own 1 M-molecule shard of the stream,
Philox-keyed by global read ordinal; no collective on the data path).

value  : whole-job simulated bases/s with the molecule batch resident in HBM (CUDA events around
         the kernel sequence, max over ranks).
e2e    : same metric through tksm_seq_simulate_text(): MDF text in host memory -> parse -> H2D ->
         kernels -> FASTQ bytes back in host memory.  At N > 1 the product's sharding is what is timed: ONE
         MDF of N x M molecules, cut by tksm_seq_shard_plan into N contiguous shards of equal simulated bases,
         rank r streams shard r with the global ordinal of its first read (the rank outputs in rank order are the
         bytes of the one-GPU run; tests/test_sharding.py).
gates  : before the timed region, untimed: the error-free stage byte for byte against oracle/seq_oracle.py and the
         CUDA FASTQ byte for byte against the CPU twin on the first reads of the workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("TKSM_MODELS", os.path.join(ROOT, "tksm_b200", "models"))

GRCH38 = [248956422, 242193529, 198295559, 190214555, 181538259, 170805979, 159345973, 145138636, 138394717,
          133797422, 135086622, 133275309, 114364328, 107043718, 101991189, 90338345, 83257441, 80373285, 58617616,
          64444167, 46709983, 50818468, 156040895, 57227415, 16569]
NAMES = [f"chr{i}" for i in range(1, 23)] + ["chrX", "chrY", "chrM"]


# --workload: the default is the configuration BASELINE.json's metric is quoted on (configs[1]); the others time the
# shapes of the remaining configs on one GPU (extra data points for DESIGN.md, not the bench line the driver reads)
WORKLOADS = {
    "bulk": dict(shape=1, models="nanopore2020", identity=None,
                 desc="BASELINE configs[1]: README bulk Tsb->Trc->plA->Flp->Tag->Seq"),
    "scrna": dict(shape=3, models="nanopore2020", identity=None,
                  desc="BASELINE configs[2] shape: scRNA Tsb->SCB->PCR->Seq (barcode, UMI, adapter literals, PCR base mods)"),
    "pacbio": dict(shape=1, models="pacbio2016", identity="99,99.9,0.5",
                   desc="BASELINE configs[3] shape: high-identity pacbio2016 tables on the bulk molecules"),
    "skew": dict(shape=2, models="nanopore2020", identity=None,
                 desc="BASELINE configs[4] shape: single-interval molecules, lengths log-uniform 500 bp-100 kbp"),
}


# dram__bytes_read.sum + dram__bytes_write.sum over all launches of one step of the default workload (one ncu pass,
# profiles/r2_dram_per_launch.txt)
TRAFFIC_BYTES = 41.94e9
TRAFFIC_NOTE = "profiles/r2zz_dram_per_launch.txt: every launch of one step, ncu (serialised, cold caches); about 21 GB of it are the two largest align_kernel launches (7.6 GB of trace cells written and read back, plus the per-read byte streams), 8.7 GB the hand-off records between the error-injection and expansion kernels"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks and throttle reasons while the timed region runs."""

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu, self.rows, self.stop_ev = gpu, [], threading.Event()

    def run_nvml(self):
        """NVML directly (nvidia-ml-py): a query takes microseconds, so a 160 ms timed region gets tens of samples."""
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(self.gpu)
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        bits = [(0x8, 2), (0x40, 3), (0x20, 4), (0x4, 5)]   # hw_slowdown, hw_thermal, sw_thermal, sw_power_cap
        while not self.stop_ev.is_set():
            sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            r = int(get_reasons(h))
            row = [str(sm), str(mx), "", "", "", ""]
            for bit, col in bits:
                row[col] = "Active" if r & bit else "Not Active"
            self.rows.append(row)
            self.stop_ev.wait(0.005)

    def run(self):
        try:
            self.run_nvml()
            return
        except Exception:
            pass
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_ev.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop_ev.wait(0.2)

    def summary(self):
        self.stop_ev.set()
        self.join(timeout=6)
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if len(r) > 2 + i and r[2 + i] == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def cpu_port_run(seqs, ids, threads):
    """Times the reference's algorithm (oracle/badread_port.c: sequence_fragment + get_qscores, statement by statement
    in C) on the given molecules: error injection, alignment, qscores and FASTQ formatting.  (bases, seconds)."""
    from oracle import badread_port as bp
    model = bp.get_model("nanopore2020", "nanopore2020", (84.0, 99.0, 5.5))   # model load is not timed
    t0 = time.perf_counter()
    reads, quals, idn = bp.simulate(model, seqs, seed=1, threads=threads)
    fq, _ = bp.to_fastq(seqs, ids, reads, quals, idn)
    secs = time.perf_counter() - t0
    assert fq[:1] == b"@"
    return sum(len(r) for r in reads), secs


CPU_KIND = ("port: oracle/badread_port.c (the reference's sequence_fragment + get_qscores in C, banded NW for edlib) + "
            "FASTQ formatting; inputs from oracle/bulk_inputs.py (numpy), nothing of the library under test")


def cpu_model():
    try:
        return open("/proc/cpuinfo").read().split("model name")[1].split(":")[1].splitlines()[0].strip()
    except Exception:
        return "unknown"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--molecules", type=int, default=1_000_000)
    ap.add_argument("--genome-scale", type=float, default=1.0, help="fraction of the 3.1 Gbp genome (tests)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--identity", default="84,99,5.5", help="--badread-identity mean,max,stdev (experiments)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="molecules in the CPU baseline sample (0 = auto)")
    ap.add_argument("--workload", default="bulk", choices=sorted(WORKLOADS))
    ap.add_argument("--no-gates", action="store_true", help="skip the untimed correctness gates")
    ap.add_argument("--gate-reads", type=int, default=10000, help="reads of the workload the CPU twin is run on")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if wl["identity"] and args.identity == "84,99,5.5":
        args.identity = wl["identity"]

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    config = {"workload": f"{wl['desc']}, "
                          f"{args.molecules} molecules per GPU, synthetic {3.1 * args.genome_scale:.2f} Gbp genome, "
                          f"{wl['models']} error+qscore models, identity {args.identity}, FASTQ with qscores",
              "molecules_per_gpu": args.molecules, "sharding": f"independent molecule shards x{world}, no collective",
              "l2": "inputs+outputs (>3 GB per step) exceed the 126 MB L2; no explicit flush"}

    if args.impl == "reference":
        if rank != 0:
            return 0
        if args.workload != "bulk":
            raise SystemExit("--impl reference times the default workload only")
        return reference_arm(args, config)

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from tksm_b200.sequence import BADREAD, FASTQ, PERFECT, Sequencer

    # host parser threads: the ranks of one box share its cores
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    host_threads = max(2, min(32, (os.cpu_count() or 1) // max(local_world, 1)))
    s = Sequencer(identity=tuple(float(x) for x in args.identity.split(",")), seed=42, device=local, threads=host_threads)
    lengths = [max(1000, int(x * args.genome_scale)) for x in GRCH38]
    s.synth_reference(NAMES, lengths, n_frac=0.01, seed=42)
    have = os.path.exists(os.path.join(os.environ["TKSM_MODELS"], "badread", "nanopore2020.error.gz"))
    if not have:
        raise SystemExit("model files missing under tksm_b200/models/badread (run __graft_entry__.build())")
    s.load_models(wl["models"], wl["models"])
    if world == 1:
        mdf = s.synth_mdf(wl["shape"], args.molecules, seed=1000, min_len=500, max_len=100000)
        ordinal_base = 0
    else:
        # ONE stream of N x M molecules (every rank generates the same text), cut by the product's shard plan:
        # contiguous shards of about equal simulated bases, global ordinals
        full = s.synth_mdf(wl["shape"], args.molecules * world, seed=1000, min_len=500, max_len=100000)
        plan = s.shard_plan(full, world)
        lo, hi, ordinal_base, _ = plan[rank]
        mdf = full[lo:hi]
        del full
        config["sharding"] = (f"tksm_seq_shard_plan: one MDF of {args.molecules * world} molecules cut into {world} contiguous "
                              f"shards of equal simulated bases, rank r streams shard r, no collective")
    batch = s.batch(mdf, ordinal_base)
    gates = run_gates(s, mdf, ordinal_base, wl, args) if (rank == 0 and not args.no_gates) else "skipped"

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        st = batch.run(BADREAD, FASTQ)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    t0 = time.perf_counter()
    dev_ms, sim_ms, launches = 0.0, 0.0, 0
    for _ in range(args.steps):
        st = batch.run(BADREAD, FASTQ)
        dev_ms += st["kernel_ms"]
        sim_ms += st["simulate_ms"]
        launches += st["kernel_launches"]
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.summary()
    bases_step = st["read_bases"]
    reads_step = st["reads"]

    # end to end through the public host-buffer call (MDF text in host memory -> FASTQ bytes in host
    # memory): tksm_seq_simulate_text parses on the host threads, uploads, runs the kernels and
    # downloads the records, chunk by chunk.  Three untimed calls first (the page-locked staging and record buffers
    # are allocated on first use and the record buffer is sized from the first call's output; on a fresh box the
    # first large page-locked allocation alone can take a second), then e2e_steps timed calls.
    e2e_bases = 0
    e2e_stats = {"h2d_bytes": 0, "d2h_bytes": 0}
    e2e_wall = 0.0
    e2e_each = []
    if args.e2e_steps:
        for _ in range(3):
            with s.simulate_view(mdf, BADREAD, fastq=True, ordinal_base=ordinal_base) as rb:
                assert rb.nbytes > 0 and rb.array()[0] == ord("@")
        barrier()
        e2e_t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            t_call = time.perf_counter()
            with s.simulate_view(mdf, BADREAD, fastq=True, ordinal_base=ordinal_base) as rb:
                a = rb.array()
                assert a[0] == ord("@") and a[rb.nbytes - 1] == ord("\n")   # the records are in host memory
            e2e_each.append(round((time.perf_counter() - t_call) * 1e3, 2))
            e2e_bases += s.last_stats["read_bases"]
            e2e_stats = dict(s.last_stats)
        barrier()
        e2e_wall = time.perf_counter() - e2e_t0

    vals = torch.tensor([dev_ms, wall, e2e_wall, float(bases_step), float(e2e_bases), float(reads_step)],
                        dtype=torch.float64, device="cuda")
    if world > 1:
        mx = vals.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = vals.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    else:
        mx, sm = vals, vals
    dev_ms_max, wall_max, e2e_wall_max = mx[0].item(), mx[1].item(), max(mx[2].item(), 1e-9)
    total_bases_step, total_e2e_bases, total_reads_step = sm[3].item(), sm[4].item(), sm[5].item()

    if rank == 0:
        peak, peak_src = peaks()
        value = total_bases_step * args.steps / (dev_ms_max / 1e3)
        achieved = st["algorithmic_bytes"] / (st["kernel_ms"] / 1e3) / 1e9
        line = {
            "metric": "Seq simulated bases/s (errors+quals, FASTQ)", "value": value, "unit": "bases/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms_max / args.steps, "wall_ms_per_step": wall_max / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic", "config": config,
            "reads_per_s": total_reads_step * args.steps / (dev_ms_max / 1e3),
            "clocks": clocks,
            "e2e": {"value": total_e2e_bases / e2e_wall_max if args.e2e_steps else None, "unit": "bases/s",
                    "h2d_bytes_per_step": e2e_stats["h2d_bytes"],
                    "mdf_text_bytes_per_step": len(mdf),   # host memory -> library (parsed on the host), not a device copy
                    "d2h_bytes_per_step": e2e_stats["d2h_bytes"],
                    "ms_per_step": e2e_wall_max / args.e2e_steps * 1e3 if args.e2e_steps else None,
                    "host_threads": host_threads, "ms_each_call_rank0": e2e_each,
                    "note": "tksm_seq_simulate_text: MDF text in host memory -> parse (host threads) -> H2D -> kernels -> "
                            "D2H of the FASTQ bytes into a page-locked host buffer, chunks pipelined"},
            "gpu_launches": launches,
            "gates": gates,
            "status_or": int(st.get("status_or", 0)),   # OR of the per-read status words of the last step (8 = some read's qscores follow the edit script: alignment band left)
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         # dram__bytes_read.sum + dram__bytes_write.sum over the 33 launches of one step of this
                         # workload, one ncu capture (profiles/r2zz_dram_per_launch.txt); other sizes: not captured
                         "traffic": TRAFFIC_BYTES if (args.workload == "bulk" and args.molecules == 1_000_000) else None,
                         "traffic_unit": "bytes per step", "peak_source": peak_src,
                         "kernel": "work order + per length class: simulate_kernel<W,1> (error injection) | <W,2> (expansion) | "
                                   "align_kernel (read-to-fragment alignment, one thread per read) | <W,4> (qscores) + "
                                   "reclen/scan/format: whole kernel sequence of a step",
                         "traffic_note": TRAFFIC_NOTE,
                         "simulate_share": sim_ms / dev_ms if dev_ms else None,
                         "algorithmic_bytes_per_step": st["algorithmic_bytes"]},
        }
        # CPU baseline beside it, rank 0, N = 1 only, bounded sample
        if world == 1 and args.workload == "bulk":   # the CPU port is set up for the default models
            line["cpu_baseline"] = cpu_baseline(s, mdf, args)
        print(json.dumps(line), file=RESULT_OUT, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def cpu_baseline(s, mdf, args):
    from oracle import bulk_inputs as bi
    n = args.cpu_sample or 2000
    seqs, ids = bi.molecules(n, seed=1000)
    bases, secs = cpu_port_run(seqs, ids, 1)
    return {"value": bases / secs, "unit": "bases/s", "cores": 1, "kind": "port", "what": CPU_KIND, "cpu": cpu_model(),
            "sample": f"{len(seqs)} bulk-shape molecules ({sum(map(len, seqs))} raw bases), {secs:.1f} s"}


def _twin_chunk(kw, models, seqs, ids, lo):
    """One span of the gate's reads through the CPU twin (worker process)."""
    from oracle import seq_oracle as so
    from tksm_b200.sequence import Sequencer
    s = Sequencer(host_only=True, **kw)
    s.load_models(models, models)
    return so.twin_simulate(s.tables(), seqs, ids, ordinal_base=lo, badread=True, fastq=True)


def run_gates(s, mdf, ordinal_base, wl, args):
    """SURVEY.md 8(d), untimed: (1) error-free stage of the first reads byte for byte against oracle/seq_oracle.py;
    (2) CUDA FASTQ of the first --gate-reads reads byte for byte against the CPU twin.  The oracle needs reference
    text: the genomic intervals of those reads are fetched back from the device-generated genome one by one."""
    import multiprocessing as mp
    from oracle import seq_oracle as so
    from tksm_b200.sequence import BADREAD, PERFECT
    n = min(args.gate_reads, args.molecules)
    pos, cnt = 0, 0
    while cnt < n:
        nxt = mdf.find(b"\n+", pos + 1)
        if nxt < 0:
            pos = len(mdf)
            break
        pos = nxt + 1
        cnt += 1
    sub = mdf[:pos]
    mols = so.parse_mdf(sub)
    contigs = set(NAMES)
    refs, seqs = {}, []
    for _, ivs in mols:
        rebased = []
        for chrom, a, b, strand, mods in ivs:
            if chrom in contigs:
                key = f"{chrom}:{a}-{b}"
                if key not in refs:
                    refs[key] = s.fetch_reference(chrom, a, b - a).decode("latin-1")
                rebased.append((key, 0, b - a, strand, mods))
            else:
                rebased.append((chrom, a, b, strand, mods))
        seqs.append(so.molecule_sequence(refs, rebased))
    ids = [m for m, _ in mols]
    tables = s.tables()
    got_p = s.simulate(sub, PERFECT, fastq=False, ordinal_base=ordinal_base)
    want_p = so.twin_simulate(tables, seqs, ids, ordinal_base=ordinal_base, badread=False, fastq=False)
    if got_p != want_p:
        return "FAILED: error-free stage differs from oracle/seq_oracle.py"
    got = s.simulate(sub, BADREAD, fastq=True, ordinal_base=ordinal_base)
    kw = dict(identity=tuple(float(x) for x in args.identity.split(",")), seed=42,
              corr=(float(tables.calib_overshoot), float(tables.calib_kappa)))
    procs = min(os.cpu_count() or 1, 16)
    per = (len(seqs) + procs - 1) // procs
    jobs = [(kw, wl["models"], seqs[i:i + per], ids[i:i + per], ordinal_base + i) for i in range(0, len(seqs), per)]
    with mp.get_context("fork").Pool(len(jobs)) as pool:
        parts = pool.starmap(_twin_chunk, jobs)
    if got != b"".join(parts):
        return "FAILED: CUDA FASTQ differs from the CPU twin"
    return f"ok: error-free stage == oracle/seq_oracle.py and CUDA FASTQ == CPU twin, byte for byte, on the first {len(seqs)} reads"


def reference_arm(args, config):
    """The reference's CPU algorithm for this path on the host cores: oracle/badread_port.c with all threads, inputs
    from oracle/bulk_inputs.py (the Python reference itself needs /root/reference and python-edlib, neither of which
    is on the GPU box; BASELINE.md has its timing in the build container).  Nothing of tksm_b200/ is loaded here."""
    from oracle import bulk_inputs as bi
    n = args.cpu_sample or 4000
    seqs, ids = bi.molecules(n, seed=1000)
    threads = os.cpu_count() or 1
    vals = []
    for i in range(args.warmup + args.steps):
        bases, secs = cpu_port_run(seqs, ids, threads)
        if i >= args.warmup:
            vals.append((bases, secs))
    tb, ts = sum(b for b, _ in vals), sum(t for _, t in vals)
    v = tb / ts
    config = dict(config, sharding="host threads", molecules_per_step=len(seqs))
    line = {"impl": "reference", "metric": "Seq simulated bases/s (errors+quals, FASTQ)", "value": v, "unit": "bases/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ts / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": config,
            "cpu_baseline": {"value": v, "unit": "bases/s", "cores": threads, "kind": "port", "what": CPU_KIND, "cpu": cpu_model(),
                             "sample": f"{len(seqs)} bulk-shape molecules per step ({sum(map(len, seqs))} raw bases)"},
            "e2e": {"value": v, "unit": "bases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), file=RESULT_OUT, flush=True)
    return 0


# stdout carries the one JSON line and nothing else: whatever libraries print on fd 1 (NCCL's version banner when
# the box sets NCCL_DEBUG) goes to stderr
RESULT_OUT = sys.stdout

if __name__ == "__main__":
    sys.stdout.flush()
    RESULT_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.exit(main())
