#!/usr/bin/env python
"""oracle/fit_calibration.py — measures the identity calibration of PB1.  TEST INFRASTRUCTURE.

PB1 stops a read when the identity of its own edit script, g-side, falls to an effective target; what the
reference stops on, and what the header prints, is the identity of an optimal alignment, a-side, which explains
part of the edits away:  a = g - kappa g^2  (a, g: 1 - identity).  kappa depends on the error model's mix of
substitutions, insertions and deletions.  The product measures it when the models are loaded, by simulating
random fragments on the device with and without the alignment stage (engine.cu: self_calibrate); this script
measures the same coefficient with the CPU twin, writes tests/golden/calibration.json (used by the CPU suite,
which has no device, and compared with the device's value by the GPU suite), and reports how the mean aligned
identity of the twin compares with the reference fixtures per length regime (py/tksm_badread.py:408-432: reads
of up to 1000 bases are re-aligned whole every 25 changes, longer ones in sampled windows), which is what the
model-independent overshoot constants in engine.cu were chosen from.

    python oracle/fit_calibration.py [model ...]
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("TKSM_MODELS", os.path.join(ROOT, "tksm_b200", "models"))
from oracle import seq_oracle as so   # noqa: E402

LEVELS = (0.78, 0.84, 0.90, 0.95)
N_MOL, LEN = 512, 1000


def calibration_molecules():
    """The fragments the device calibrates on (engine.cu: self_calibrate): splitmix64 bases."""
    x = np.uint64(0x9E3779B97F4A7C15)
    out = []
    state = np.uint64(12345)
    with np.errstate(over="ignore"):
        for _ in range(N_MOL):
            bases = np.empty(LEN, np.uint8)
            for i in range(0, LEN, 32):
                state = state + x
                z = state
                z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
                z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
                z = z ^ (z >> np.uint64(31))
                for t in range(min(32, LEN - i)):
                    bases[i + t] = (int(z) >> (2 * t)) & 3
            out.append(bytes(b"ACGT"[b] for b in bases))
    return out


def measure_kappa(em, qm):
    from tksm_b200.sequence import Sequencer
    seqs = calibration_molecules()
    ids = [f"c{i}" for i in range(len(seqs))]
    sx = sy = 0.0
    for level in LEVELS:
        pairs = []
        for qual in (False, True):
            s = Sequencer(identity=(100 * level, 100 * level, 0.0), seed=0x5EEDCA11B, host_only=True, corr=(0.0, 0.0),
                          skip_qual_compute=not qual)
            s.load_models(em, qm)
            _, _, ident = so.twin_simulate(s.tables(), seqs, ids, badread=True, fastq=True, want_lengths=True)
            pairs.append(1.0 - ident[:, 0].astype(float) / ident[:, 1])
        g, a = pairs
        sx += float((g * g * (g - a)).sum())
        sy += float((g ** 4).sum())
    return sx / sy


def _twin_hdr_mean(args):
    """Mean read_identity= of the twin on a slice of the calibration molecules at a given effective kappa."""
    em, qm, ident, lo, hi, kappa = args
    from oracle.make_goldens import dist_inputs
    from tksm_b200.sequence import Sequencer
    fasta, mdf = dist_inputs(seed=6, n_mol=4000)
    recs = mdf.split("\n+")
    recs = [recs[0]] + ["+" + r for r in recs[1:]]
    part = "\n".join(x.rstrip("\n") for x in recs[lo:hi]) + "\n"
    s = Sequencer(identity=ident, seed=4242, host_only=True, corr=(0.8, kappa))
    s.add_reference_text(fasta)
    s.load_models(em, qm)
    refs = so.load_fasta(fasta)
    mols = so.parse_mdf(part)
    seqs = [so.molecule_sequence(refs, iv) for _, iv in mols]
    fq = so.twin_simulate(s.tables(), seqs, [m for m, _ in mols], ordinal_base=lo, badread=True, fastq=True)
    vals = [float(ln.split(b"read_identity=")[1].split(b"%")[0]) / 100.0 for ln in fq.split(b"\n")[0::4] if ln]
    return sum(vals), len(vals)


def fit_affine(n_reads=2400, procs=8):
    """CALIB_OFFSET_TABLE / _RANDOM of engine.cu: the effective kappa at which the twin's mean header identity equals
    the reference's (tests/golden/ref_calib.json: molecules the distribution tests do not use) minus the kappa
    measured on random fragments, averaged over the shipped models of each family."""
    from multiprocessing import Pool
    cal = json.load(open(os.path.join(ROOT, "tests", "golden", "calibration.json")))
    ref = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_calib.json")))
    xs, ys = [], []
    per = (n_reads + procs - 1) // procs
    for name, info in ref["models"].items():
        c = info["config"]
        ident = tuple(float(x) for x in c["ident"].split(","))
        means = []
        for kappa in (0.38, 0.62):
            with Pool(procs) as p:
                parts = p.map(_twin_hdr_mean, [(c["em"], c["qm"], ident, i * per, min(n_reads, (i + 1) * per), kappa)
                                               for i in range(procs)])
            means.append(sum(a for a, _ in parts) / sum(b for _, b in parts))
        slope = (means[1] - means[0]) / (0.62 - 0.38)
        k0 = 0.38 + (info["hdr_identity_mean"] - means[0]) / slope
        print(f"{name}: twin header identity {means[0]:.5f} (kappa 0.38) / {means[1]:.5f} (0.62), reference "
              f"{info['hdr_identity_mean']:.5f} -> kappa {k0:.3f}; measured on random fragments {cal[name]:.3f}")
        xs.append(cal[name])
        ys.append(k0)
    # one offset for the models with a k-mer table, one for the table-less `random` model (three points carry no slope)
    names = list(ref["models"])
    tab = [y - x for nm, x, y in zip(names, xs, ys) if nm != "random"]
    rnd = [y - x for nm, x, y in zip(names, xs, ys) if nm == "random"]
    print(f"CALIB_OFFSET_TABLE = {np.mean(tab):.3f} (per model {[round(v, 3) for v in tab]}), "
          f"CALIB_OFFSET_RANDOM = {np.mean(rnd):.3f}")
    return float(np.mean(tab)), float(np.mean(rnd))


if __name__ == "__main__":
    if sys.argv[1:2] == ["fit"]:
        fit_affine()
        sys.exit(0)
    models = sys.argv[1:] or ["nanopore2020", "nanopore2018", "pacbio2016", "random"]
    path = os.path.join(ROOT, "tests", "golden", "calibration.json")
    cal = json.load(open(path)) if os.path.exists(path) else {}
    for m in models:
        cal[m] = round(measure_kappa(m, m if m != "random" else "random"), 4)
        print(m, "kappa", cal[m])
    json.dump(cal, open(path, "w"), indent=1, sort_keys=True)
