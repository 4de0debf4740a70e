import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
os.environ.setdefault("TKSM_MODELS", "/root/repo/tksm_b200/models")
import numpy as np, multiprocessing as mp
import test_gpu_parity as t
from oracle import seq_oracle as so
from tksm_b200.sequence import BADREAD, Sequencer
rng = np.random.default_rng(31)
fasta = t.make_fasta(rng, [("chr1", 600000), ("chr2", 400000)])
s = Sequencer(seed=4321); s.add_reference_text(fasta); s.load_models("nanopore2020", "nanopore2020")
mdf = s.synth_mdf(0, 3000, seed=3, min_len=300, max_len=20000)
got = s.simulate(mdf, BADREAD, fastq=True)
print("status_or", s.last_stats["status_or"])
kw = dict(seed=4321, corr=(float(s.tables().calib_overshoot), float(s.tables().calib_kappa)))
text = mdf.decode(); recs = text.split("\n+"); recs = [recs[0]] + ["+" + r for r in recs[1:]]
depth = [int(r.split("\n", 1)[0].split("\t")[1]) for r in recs]
procs = 16; per = (len(recs) + procs - 1) // procs
jobs, lo = [], 0
for i in range(0, len(recs), per):
    jobs.append((kw, fasta, "\n".join(x.rstrip("\n") for x in recs[i:i + per]) + "\n", lo, True)); lo += sum(depth[i:i + per])
with mp.get_context("fork").Pool(len(jobs)) as pool:
    want = b"".join(pool.map(t._twin_span, jobs))
g = so.parse_records(got); w = so.parse_records(want)
bad = [(i, len(a[1])) for i, (a, b) in enumerate(zip(g, w)) if a != b]
lens = [len(a[1]) for a in g]
print("reads", len(g), "bad", len(bad), bad[:20], "long reads", sum(l > 2040 for l in lens), "of which >4800:", sum(l > 4800 for l in lens))
