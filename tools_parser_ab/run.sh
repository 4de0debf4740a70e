#!/bin/bash
# A/B timing of csrc/host.cpp's MdfParser::feed between two revisions, in one process, the two alternating rep by rep
# (the build container's clock is noisy: min and the median of the paired ratios are what to read), with a checksum
# of the parsed molecules (literals by content, not by pool offset) that must be the same on both sides.
#   tools_parser_ab/run.sh OLD_REV [REPS]       e.g. tools_parser_ab/run.sh 62c9621 100
# Inputs: a 23 Mbp random genome and the bulk / scRNA shapes of tksm_seq_synth_mdf (host_only handle, no GPU).
set -e
cd "$(dirname "$0")"
OLD=${1:?old revision}
REPS=${2:-30}
B=_build
mkdir -p $B
CS=../tksm_b200/csrc
git show "$OLD:tksm_b200/csrc/host.cpp" > $B/host_old.cpp
[ -f $B/g.fa ] || TKSM_MODELS=${TKSM_MODELS:-$PWD/../tksm_b200/models} python - <<'PY'
import sys
import numpy as np
sys.path.insert(0, "..")
from tksm_b200.sequence import Sequencer
rng = np.random.default_rng(3)
fasta = "".join(f">{n}\n{rng.choice(np.frombuffer(b'ACGT', np.uint8), size=1_000_000).tobytes().decode()}\n"
                for n in [f"chr{i}" for i in range(1, 23)] + ["chrX"])
open("_build/g.fa", "w").write(fasta)
s = Sequencer(host_only=True)
s.add_reference_text(fasta)
for shape, n in ((1, 300000), (3, 100000)):
    open(f"_build/s{shape}.mdf", "wb").write(s.synth_mdf(shape, n, seed=4))
PY
g++ -O2 -std=c++20 -I$CS -Dtk=tk_old -DRUN_NAME=run_old -DRUN_MAKE=run_make_old -c body.cpp -o $B/body_old.o
g++ -O2 -std=c++20 -I$CS -Dtk=tk_old -c $B/host_old.cpp -o $B/host_old.o
g++ -O2 -std=c++20 -I$CS -DRUN_NAME=run_new -DRUN_MAKE=run_make_new -c body.cpp -o $B/body_new.o
g++ -O2 -std=c++20 -I$CS -c $CS/host.cpp -o $B/host_new.o
g++ -O2 -std=c++20 -c ab.cpp -o $B/ab.o
g++ -o $B/ab $B/ab.o $B/body_old.o $B/host_old.o $B/body_new.o $B/host_new.o -lz
for f in s1 s3; do echo "== $f"; $B/ab $B/g.fa $B/$f.mdf "$REPS"; done
