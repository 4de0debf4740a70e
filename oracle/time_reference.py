#!/usr/bin/env python
"""oracle/time_reference.py — times the UNMODIFIED Python reference Seq module on BASELINE configs[0].  TEST INFRASTRUCTURE.

Runs only in the build container (needs /root/reference).  configs[0]: 10 k molecules from a synthetic 1 Mbp genome,
nanopore2020 error + qscore models (nanopore2023 is not shipped), FASTQ with qscores.  The one missing dependency,
python-edlib, is replaced by oracle/edlib_shim (oracle/nw.c), so alignment time is our stand-in's, not edlib's.
Prints start-up (model load) and steady-state throughput for -t 1 (on a subsample) and -t <cores>; the numbers go
into BASELINE.md by hand.
"""
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.make_goldens import REF, ref_env   # noqa: E402


def c1_inputs(n_mol=10000, seed=11):
    rng = np.random.default_rng(seed)
    contigs = {"chr1": 600000, "chr2": 400000}
    fasta, seqs = [], {}
    for name, n in contigs.items():
        g = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=n).tobytes().decode()
        seqs[name] = g
        fasta.append(f">{name}\n" + "\n".join(g[i:i + 60] for i in range(0, n, 60)) + "\n")
    lines = []
    for m in range(n_mol):   # Tsb shape: 1 + Geom exons of log-normal length on one strand, mean molecule ~1.5 kbp
        chrom = "chr1" if rng.random() < 0.6 else "chr2"
        n_ex = 1 + int(rng.geometric(1 / 8.0))
        strand = "+" if rng.random() < 0.5 else "-"
        pos = int(rng.integers(0, contigs[chrom] - 40000))
        lines.append(f"+m{m}\t1\t")
        for _ in range(n_ex):
            ln = max(20, int(rng.lognormal(np.log(150), 0.6)))
            lines.append(f"{chrom}\t{pos}\t{pos + ln}\t{strand}\t")
            pos += ln + max(50, int(rng.lognormal(np.log(1500), 0.8)))
            if pos > contigs[chrom] - 2000:
                break
    return "".join(fasta), "\n".join(lines) + "\n"


def run(td, mdf_name, threads):
    cmd = [sys.executable, os.path.join(REF, "py", "sequence.py"), "-i", mdf_name, "-r", "g.fa", "-o", f"out_t{threads}.fastq",
           "-t", str(threads), "--badread-error-model", "nanopore2020", "--badread-qscore-model", "nanopore2020"]
    t0 = time.time()
    subprocess.run(cmd, cwd=td, env=ref_env(), check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    wall = time.time() - t0
    bases = sum(len(l) - 1 for i, l in enumerate(open(os.path.join(td, f"out_t{threads}.fastq"))) if i % 4 == 1)
    return wall, bases


if __name__ == "__main__":
    fasta, mdf = c1_inputs()
    cores = os.cpu_count() or 1
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "g.fa"), "w").write(fasta)
        open(os.path.join(td, "full.mdf"), "w").write(mdf)
        recs = mdf.split("\n+")
        open(os.path.join(td, "empty.mdf"), "w").write(recs[0] + "\n")            # one molecule: start-up cost
        open(os.path.join(td, "sub.mdf"), "w").write("\n+".join(recs[:600]) + "\n")
        t_start, _ = run(td, "empty.mdf", 1)
        t1, b1 = run(td, "sub.mdf", 1)
        tn, bn = run(td, "full.mdf", cores)
    print(f"cpu: {open('/proc/cpuinfo').read().split('model name')[1].split(':')[1].splitlines()[0].strip()}, {cores} hardware threads")
    print(f"start-up (imports, model load, one molecule): {t_start:.1f} s")
    print(f"-t 1, 600 molecules: {b1} bases in {t1:.1f} s -> {b1 / max(t1 - t_start, 1e-9):.0f} bases/s steady state")
    print(f"-t {cores}, 10000 molecules: {bn} bases in {tn:.1f} s -> {bn / max(tn - t_start, 1e-9):.0f} bases/s steady state, "
          f"{bn / tn:.0f} bases/s including start-up")
