#!/usr/bin/env python
"""oracle/make_goldens.py — generates tests/golden/* by running the UNMODIFIED reference.

TEST INFRASTRUCTURE.  Runs only in the build container (needs /root/reference); the fixtures it
writes are committed so the GPU box and the CPU test-suite never read /root/reference.

    python oracle/make_goldens.py [deterministic] [align] [qscore] [identity] [dist]

What runs: /root/reference/py/sequence.py and /root/reference/py/tksm_badread.py exactly as they
are, with oracle/edlib_shim on PYTHONPATH because python-edlib (env.yaml:8, unpinned) is the one
dependency missing from this image.  src/sequence.cpp adds nothing but the
$TKSM_MODELS variable, so this is the reference Seq module.

Fixtures:
    deterministic.json   FASTA + MDF cases -> `--perfect` FASTA and FASTQ records (uuid masked)
    align_kmers.json     ERROR_MODEL_PY.align_kmers on sampled (k-mer, alternative) pairs
    qscore_resolve.json  QScoreModel.get_qscore: which model row a window cigar ends up sampling
    identity.json        IDENTITIES_PY.beta_parameters values
    ref_dist_<model>.json.gz  statistics of reads the reference simulated (unseeded) for the
                         distribution tests: per-read length ratio / identity, X/I/D rates,
                         homopolymer spectrum, qscore histograms
"""
import gzip
import importlib.util
import json
import os
import random
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
SHIM = os.path.join(HERE, "edlib_shim")
GOLD = os.path.join(ROOT, "tests", "golden")
MODELS = os.path.join(REF, "py", "tksm_models")
sys.path.insert(0, ROOT)


def ref_env():
    env = dict(os.environ)
    env["PYTHONPATH"] = SHIM + ":" + os.path.join(REF, "py")
    env["TKSM_MODELS"] = MODELS
    return env


def run_reference(args, cwd):
    cmd = [sys.executable, os.path.join(REF, "py", "sequence.py")] + args
    p = subprocess.run(cmd, cwd=cwd, env=ref_env(), capture_output=True, text=True)
    return p


def import_reference_badread():
    sys.path.insert(0, SHIM)
    spec = importlib.util.spec_from_file_location("tksm_badread", os.path.join(REF, "py", "tksm_badread.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def mask_uuid(text):
    out = []
    for ln in text.split("\n"):
        if ln[:1] in "@>" and " length=" in ln:
            ln = ln[0] + "UUID" + ln[ln.index(" "):]
        out.append(ln)
    return "\n".join(out)


# ---------------------------------------------------------------- deterministic stage

def deterministic_cases():
    rng = np.random.default_rng(20240611)

    def rand_seq(n, alphabet="ACGT"):
        return "".join(rng.choice(list(alphabet), size=n))

    cases = []
    # README.md:233-251 and 253-270 (with the trailing tab the MDF writer always emits)
    cases.append(dict(name="readme_mods", fasta=">1\nAGTCCCGTAA\n",
                      mdf="+m1\t1\t\n1\t0\t4\t+\t2C,3T\n1\t6\t9\t+\t1G\n"))
    cases.append(dict(name="readme_literal", fasta=">1\nAGTCCCGTAA\n",
                      mdf="+m2\t1\t\nTT\t0\t2\t+\t\n1\t0\t4\t+\t\n"))
    # SURVEY.md §4 item 4
    cases.append(dict(name="minus_depth_clamp_lower", fasta=">1 desc ignored\nAGTCCCGTAA\n",
                      mdf="+a\t1\tx=1;\n1\t0\t4\t-\t\n+b\t2\t\n1\t2\t7\t+\t\n+c\t1\t\n1\t8\t50\t+\t\n"
                          "+d\t1\t\nacgtn\t0\t5\t-\t0r\n+e\t0\t\n1\t0\t3\t+\t\n+f\t1\t\n1\t-4\t-1\t+\t-1G\n"))
    # soft-masked genome, N runs, IUPAC, CRLF-free multi-line records, several contigs, header with spaces
    g1 = rand_seq(3000)
    g1 = g1[:500] + g1[500:900].lower() + g1[900:1200] + "N" * 137 + g1[1337:2000] + "RYKMSWBDHVN" + g1[2011:]
    g2 = rand_seq(1500) + "nnnnnnnn" + rand_seq(400).lower()
    fasta = (">chrA first contig\n" + "\n".join(g1[i:i + 60] for i in range(0, len(g1), 60)) + "\n"
             ">chrB\n" + "\n".join(g2[i:i + 70] for i in range(0, len(g2), 70)) + "\n"
             ">chrA_dup x\nACGTACGTAC\n>chrA_dup y\nTTTTGGGGCC\n")
    mdf = []
    lens = {"chrA": len(g1), "chrB": len(g2)}
    for m in range(120):
        nseg = int(rng.integers(1, 7))
        lines = [f"+mol{m}\t{int(rng.integers(1, 4)) if m % 7 == 0 else 1}\ttid=T{m};CB=ACGT;"]
        for s in range(nseg):
            kind = rng.random()
            if kind < 0.2:
                lit = rand_seq(int(rng.integers(1, 40)), "ACGT" if rng.random() < 0.7 else "ACGTNacgtRY")
                a, b = 0, len(lit)
                chrom, clen = lit, len(lit)
            else:
                chrom = "chrA" if rng.random() < 0.6 else "chrB"
                clen = lens[chrom]
                a = int(rng.integers(0, clen - 1))
                b = min(clen, a + int(rng.integers(1, 400)))
                if rng.random() < 0.05:
                    b = clen + int(rng.integers(0, 100))     # clamped slice
            strand = "+" if rng.random() < 0.5 else "-"
            seglen = min(b, clen) - a
            mods = ""
            if seglen > 0 and rng.random() < 0.3:
                ps = sorted(set(int(x) for x in rng.integers(0, seglen, size=int(rng.integers(1, 5)))))
                mods = ",".join(f"{p}{rng.choice(list('ACGTacgtN'))}" for p in ps)
            lines.append(f"{chrom}\t{a}\t{b}\t{strand}\t{mods}")
        mdf.append("\n".join(lines))
    cases.append(dict(name="mixed_corpus", fasta=fasta, mdf="\n".join(mdf) + "\n"))
    # later FASTA records override earlier names; single-base and empty slices
    cases.append(dict(name="override_and_tiny", fasta=">c\nAAAAAAAAAA\n>c\nCCCCCGGGGG\n>d\nACGT\n",
                      mdf="+x\t1\t\nc\t3\t7\t+\t\n+y\t1\t\nd\t1\t2\t-\t\n+z\t1\t\nd\t2\t2\t+\t\nc\t0\t1\t+\t0T\n"))
    return cases


def make_deterministic():
    out = []
    for case in deterministic_cases():
        with tempfile.TemporaryDirectory() as td:
            open(os.path.join(td, "g.fa"), "w").write(case["fasta"])
            open(os.path.join(td, "in.mdf"), "w").write(case["mdf"])
            rec = dict(case)
            for ext in ("fasta", "fastq"):
                p = run_reference(["-i", "in.mdf", "-r", "g.fa", "--perfect", f"out.{ext}"], td)
                if p.returncode != 0:
                    raise SystemExit(f"reference failed on {case['name']}: {p.stderr[-2000:]}")
                rec["out_" + ext] = mask_uuid(open(os.path.join(td, f"out.{ext}")).read())
            out.append(rec)
            print(f"deterministic/{case['name']}: {rec['out_fasta'].count(chr(10)) // 2} records")
    json.dump(out, open(os.path.join(GOLD, "deterministic.json"), "w"), indent=0)


# ---------------------------------------------------------------- align_kmers

def make_align():
    bad = import_reference_badread()
    ak = bad.ERROR_MODEL_PY.align_kmers
    vec = [["ACGT", "ACGTT", ak("ACGT", "ACGTT")], ["ACGT", "ACT", ak("ACGT", "ACT")]]
    rng = random.Random(7)
    for model in ("nanopore2020", "nanopore2018", "pacbio2016"):
        path = os.path.join(MODELS, "badread", f"{model}.error.gz")
        lines = gzip.open(path, "rt").read().split("\n")
        lines = [l for l in lines if l]
        pick = [lines[0], lines[-1]] + rng.sample(lines, 150)
        for l in pick:
            items = [x.split(",") for x in l.strip().split(";") if x]
            kmer = items[0][0]
            for alt, _ in items:
                vec.append([kmer, alt, ak(kmer, alt)])
    json.dump(vec, open(os.path.join(GOLD, "align_kmers.json"), "w"), separators=(",", ":"))
    print(f"align_kmers: {len(vec)} vectors")


# ---------------------------------------------------------------- qscore row resolution

def make_qscore():
    bad = import_reference_badread()
    captured = {}

    def fake_choices(scores, weights=None):
        captured["scores"], captured["weights"] = list(scores), list(weights)
        return [scores[0]]

    out = {}
    rng = random.Random(11)
    for model in ("nanopore2020", "nanopore2018", "pacbio2016", "random", "ideal"):
        name = model if model in ("random", "ideal") else os.path.join(MODELS, "badread", f"{model}.qscore.gz")
        qm = bad.QSCOREMODEL_PY.QScoreModel(name, output=open(os.devnull, "w"))
        by_dist = {}
        for c in qm.scores:
            by_dist.setdefault((tuple(qm.scores[c]), tuple(qm.probabilities[c])), []).append(c)
        vec = []
        real_choices = bad.random.choices
        bad.random.choices = fake_choices
        try:
            for _ in range(1500):
                w = rng.choice([1, 3, 5, 7, 9]) if qm.kmer_size >= 9 else 1
                w = min(w, qm.kmer_size)
                ops = [rng.choices("=XI", weights=[0.8, 0.1, 0.1])[0] for _ in range(w)]
                cig = ops[0]
                for o in ops[1:]:
                    if rng.random() < 0.15:
                        cig += "D" * rng.choice([1, 1, 1, 2, 3, 5])
                    cig += o
                qm.get_qscore(cig)
                vec.append([cig, captured["scores"], captured["weights"]])
        finally:
            bad.random.choices = real_choices
        out[model] = {"kmer_size": qm.kmer_size, "n_cigars": len(qm.scores), "vectors": vec}
        print(f"qscore/{model}: k={qm.kmer_size}, {len(qm.scores)} cigars, {len(vec)} vectors")
    with gzip.open(os.path.join(GOLD, "qscore_resolve.json.gz"), "wt") as f:
        json.dump(out, f, separators=(",", ":"))


def make_identity():
    bad = import_reference_badread()
    rows = []
    for trip in [(84.0, 5.5, 99.0), (87.0, 4.0, 97.0), (99.0, 0.5, 99.9), (60.0, 8.0, 80.0), (95.0, 2.5, 100.0)]:
        a, b = bad.IDENTITIES_PY.beta_parameters(*trip)
        rows.append({"mean": trip[0], "stdev": trip[1], "max": trip[2], "a": a, "b": b})
    json.dump(rows, open(os.path.join(GOLD, "identity.json"), "w"), indent=0)
    print("identity:", rows[0])


# ---------------------------------------------------------------- tail noise

def tail_model_dict():
    """A small synthetic tail model in the layout KDE_noise_generator.save writes (no model is shipped)."""
    rng = np.random.default_rng(11)
    lx = [0, 3, 6, 12, 25, 50, 100, 200]
    ly = [200, 500, 1000, 2000]
    grid = (rng.random((len(ly), len(lx))) + 0.05) * np.array([[3, 2, 2, 1.5, 1, 1, 0.5, 0.25]])
    trans = rng.random((4, 4)) + 0.1
    trans[0, 0] += 2.0   # homopolymer-rich, like poly(A) read-through
    return {"lx": lx, "ly": ly, "grid": grid.tolist(), "bases": list("AGTC"), "ratio": 0.35,
            "trans": trans.tolist(), "begin": [0.25, 0.25, 0.25, 0.25]}


def make_tail():
    import random
    bad = import_reference_badread()
    dc = tail_model_dict()
    path = os.path.join(GOLD, "tail_model.json")
    gen = bad.TAIL_NOISE_MODEL_PY.KDE_noise_generator(dc["lx"], dc["ly"], dc["grid"], (dc["begin"], dc["trans"]),
                                                      dc["ratio"], dc["bases"])
    with open(path, "w") as f:
        gen.save(f)                                   # the reference's own writer
    gen = bad.TAIL_NOISE_MODEL_PY.KDE_noise_generator.load(path)   # and its own reader
    random.seed(20240611)
    out = {"n": 40000, "frag_lens": {}}
    for frag_len in (150, 340, 360, 500, 1400, 2000, 2001, 9000):
        lens, pairs, first = {}, np.zeros((4, 4), np.int64), np.zeros(4, np.int64)
        for _ in range(out["n"]):
            t = gen.noise_seq(frag_len)
            lens[len(t)] = lens.get(len(t), 0) + 1
            if t:
                idx = ["AGTC".index(c) for c in t]
                first[idx[0]] += 1
                for a, b in zip(idx, idx[1:]):
                    pairs[a, b] += 1
        out["frag_lens"][str(frag_len)] = {"lens": {str(k): v for k, v in sorted(lens.items())},
                                           "pairs": pairs.tolist(), "first": first.tolist()}
        print(f"tail/{frag_len}: empty {lens.get(0, 0) / out['n']:.3f}, lengths {sorted(lens)[:10]}")
    # end to end through the unmodified sequence.py: the tail is part of the read, not of error_free_length,
    # and the perfect file does not hold it (py/sequence.py:249-270)
    with tempfile.TemporaryDirectory() as td:
        rng = np.random.default_rng(3)
        genome = "".join("ACGT"[i] for i in rng.integers(0, 4, 60000))
        open(os.path.join(td, "g.fa"), "w").write(">c1\n" + genome + "\n")
        with open(os.path.join(td, "in.mdf"), "w") as f:
            for i in range(600):
                st = int(rng.integers(0, 59000))
                f.write(f"+m{i}\t1\t\n")
                f.write(f"c1\t{st}\t{st + 500}\t+\t\n")
        p = run_reference(["-i", "in.mdf", "-r", "g.fa", "-o", "out.fasta", "--badread-error-model", "random",
                           "--badread-qscore-model", "random", "--badread-identity", "99.5,99.9,0.1",
                           "--badread-tail-model", path], td)
        assert p.returncode == 0, p.stderr
        extra = []
        for line in open(os.path.join(td, "out.fasta")):
            if line.startswith(">"):
                kv = dict(t.split("=") for t in line.split()[1:])
                assert kv["error_free_length"] == "500"
                extra.append(int(kv["length"]) - 500)
        p = run_reference(["-i", "in.mdf", "-r", "g.fa", "--perfect", "perf.fasta", "--badread-tail-model", path], td)
        assert p.returncode == 0, p.stderr
        plen = {len(l.strip()) for i, l in enumerate(open(os.path.join(td, "perf.fasta"))) if i % 2 == 1}
        assert plen == {500}, plen
        out["e2e"] = {"frag_len": 500, "identity": "99.5,99.9,0.1", "extra": extra}
        print(f"tail/e2e: mean extra length {np.mean(extra):.2f}, with tail {np.mean(np.array(extra) > 1):.3f}")
    with gzip.open(os.path.join(GOLD, "ref_tail.json.gz"), "wt") as f:
        json.dump(out, f, separators=(",", ":"))


# ---------------------------------------------------------------- distributions

def dist_inputs(seed=5, n_mol=1500, lo=250, hi=4000):
    """Genome with planted homopolymer runs + single-interval molecules on both strands."""
    rng = np.random.default_rng(seed)
    g = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=400000)
    for p in rng.integers(0, len(g) - 20, size=9000):
        g[p:p + int(rng.integers(3, 11))] = rng.choice(np.frombuffer(b"ACGT", np.uint8))
    g = g.tobytes().decode()
    fasta = ">chr1\n" + "\n".join(g[i:i + 80] for i in range(0, len(g), 80)) + "\n"
    lines = []
    for m in range(n_mol):
        L = int(np.exp(rng.uniform(np.log(lo), np.log(hi))))
        a = int(rng.integers(0, len(g) - L))
        lines.append(f"+m{m}\t1\t\nchr1\t{a}\t{a + L}\t{'+' if rng.random() < 0.5 else '-'}\t")
    return fasta, "\n".join(lines) + "\n"


def reference_reads(fasta, mdf, c):
    """FASTQ of the unmodified reference for the MDF (independent single-threaded processes) + its --perfect FASTA."""
    # The reference draws read identities from numpy's global RandomState, which fork() copies
    # into every Pool worker unchanged (py/sequence.py:363-365, py/tksm_badread.py:741-745): with
    # `-t N` the N workers replay the same identity sequence.  Independent samples need `-t 1`,
    # so the MDF is cut into chunks that separate single-threaded reference processes simulate.
    recs = mdf.split("\n+")
    recs = [recs[0]] + ["+" + r for r in recs[1:]]
    nproc = max(1, min(os.cpu_count() or 1, 8))
    per = (len(recs) + nproc - 1) // nproc
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "g.fa"), "w").write(fasta)
        open(os.path.join(td, "in.mdf"), "w").write(mdf)
        procs = []
        for w in range(nproc):
            part = recs[w * per:(w + 1) * per]
            if not part:
                continue
            open(os.path.join(td, f"in{w}.mdf"), "w").write("\n".join(x.rstrip("\n") for x in part) + "\n")
            cmd = [sys.executable, os.path.join(REF, "py", "sequence.py"), "-i", f"in{w}.mdf", "-r", "g.fa",
                   "-o", f"out{w}.fastq", "-t", "1", "--badread-identity", c["ident"],
                   "--badread-error-model", c["em"], "--badread-qscore-model", c["qm"]]
            # progress text goes to a file: a pipe nobody drains would block the worker
            log = open(os.path.join(td, f"log{w}.txt"), "w")
            procs.append((w, subprocess.Popen(cmd, cwd=td, env=ref_env(), stdout=log, stderr=log)))
        fq = b""
        for w, p in procs:
            if p.wait() != 0:
                err = open(os.path.join(td, f"log{w}.txt")).read()
                raise SystemExit(f"reference failed on {c}: {err[-3000:]}")
            fq += open(os.path.join(td, f"out{w}.fastq"), "rb").read()
        p2 = run_reference(["-i", "in.mdf", "-r", "g.fa", "--perfect", "perf.fasta"], td)
        assert p2.returncode == 0
        perf = open(os.path.join(td, "perf.fasta"), "rb").read()
    return fq, perf


DIST_CONFIGS = {
    "nanopore2020": dict(em="nanopore2020", qm="nanopore2020", ident="84.0,99.0,5.5"),
    "nanopore2018": dict(em="nanopore2018", qm="nanopore2018", ident="84.0,99.0,5.5"),
    "pacbio2016": dict(em="pacbio2016", qm="pacbio2016", ident="84.0,99.0,5.5"),
    "pacbio2016_hifi": dict(em="pacbio2016", qm="pacbio2016", ident="99.0,99.9,0.5"),
    "random": dict(em="random", qm="random", ident="84.0,99.0,5.5"),
    "nanopore2020_const90": dict(em="nanopore2020", qm="ideal", ident="90.0,90.0,0.0"),
}


def make_dist(which=None):
    from oracle import stats as st
    n_mol = int(os.environ.get("DIST_N_MOL", "10000"))   # at least 10 k reads per model
    fasta, mdf = dist_inputs(n_mol=n_mol)
    for name, c in DIST_CONFIGS.items():
        if which and name not in which:
            continue
        fq, perf = reference_reads(fasta, mdf, c)
        summary = st.summarise(fq, perf)
        summary["config"] = c
        summary["inputs"] = {"seed": 5, "n_mol": n_mol, "lo": 250, "hi": 4000}
        for key in ("len_ratio", "identity", "hdr_identity", "sub_rate", "ins_rate", "del_rate", "mean_q", "hp_err",
                    "hp_del", "mean_q_eq", "mean_q_x", "mean_q_i"):
            summary[key] = [round(float(v), 5) for v in summary[key]]
        with gzip.open(os.path.join(GOLD, f"ref_dist_{name}.json.gz"), "wt") as f:
            json.dump(summary, f, separators=(",", ":"))
        print(f"dist/{name}: {summary['n_reads']} reads, rates {summary['rates']}, "
              f"mean identity {np.mean(summary['identity']):.4f}, mean len ratio {np.mean(summary['len_ratio']):.4f}")


def make_calib():
    """Reference runs the identity calibration is fitted on (oracle/fit_calibration.py): other molecules and another
    genome (dist_inputs seed 6) than the distribution tests use, so the gate does not check what the fit has seen."""
    fasta, mdf = dist_inputs(seed=6, n_mol=4000)
    out = {"inputs": {"seed": 6, "n_mol": 4000, "lo": 250, "hi": 4000}, "models": {}}
    for name in ("nanopore2020", "nanopore2018", "pacbio2016", "random"):
        c = DIST_CONFIGS[name]
        fq, _ = reference_reads(fasta, mdf, c)
        hdr = []
        for ln in fq.split(b"\n")[0::4]:
            if ln:
                tok = [t for t in ln.decode().split(" ") if t.startswith("read_identity=")][0]
                hdr.append(float(tok.split("=")[1].rstrip("%")) / 100.0)
        out["models"][name] = {"config": c, "n_reads": len(hdr), "hdr_identity_mean": float(np.mean(hdr)),
                               "hdr_identity_std": float(np.std(hdr))}
        print(f"calib/{name}: {len(hdr)} reads, mean header identity {np.mean(hdr):.5f}")
    json.dump(out, open(os.path.join(GOLD, "ref_calib.json"), "w"), indent=1)


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    subprocess.check_call(["make", "-C", HERE, "_build/libnw.so"])
    what = sys.argv[1:] or ["deterministic", "align", "qscore", "identity", "dist"]
    if "deterministic" in what:
        make_deterministic()
    if "align" in what:
        make_align()
    if "qscore" in what:
        make_qscore()
    if "identity" in what:
        make_identity()
    if "tail" in what:
        make_tail()
    if "calib" in what:
        make_calib()
    if "dist" in what:
        sel = [w[5:] for w in what if w.startswith("dist:")]
        make_dist(sel or None)
    for w in what:
        if w.startswith("dist:") and "dist" not in what:
            make_dist([w[5:]])
