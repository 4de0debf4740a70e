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


if __name__ == "__main__":
    models = sys.argv[1:] or ["nanopore2020", "nanopore2018", "pacbio2016", "random"]
    path = os.path.join(ROOT, "tests", "golden", "calibration.json")
    cal = json.load(open(path)) if os.path.exists(path) else {}
    for m in models:
        cal[m] = round(measure_kappa(m, m if m != "random" else "random"), 4)
        print(m, "kappa", cal[m])
    json.dump(cal, open(path, "w"), indent=1, sort_keys=True)
