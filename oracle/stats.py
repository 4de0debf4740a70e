"""oracle/stats.py — read-level statistics of simulated reads.  TEST INFRASTRUCTURE ONLY.

`summarise(fastq_bytes, perfect_fasta_bytes)` aligns every simulated read back to its error-free
molecule (matched by the molecule_id= header token) with oracle/nw.c and reduces the alignments
to the quantities north_star names for distributional parity: read length, per-read identity,
sub/ins/del rates, homopolymer error spectrum, qscore histograms.  The same function is applied to
reads of the unmodified reference (oracle/make_goldens.py -> tests/golden/ref_dist_*.json.gz) and
to reads of the CPU twin / CUDA path (tests/test_distribution.py), so alignment tie-breaking
cancels out of the comparison.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_nw = None


def nw_lib():
    global _nw
    if _nw is None:
        so = os.path.join(_HERE, "_build", "libnw.so")
        if not os.path.exists(so):
            subprocess.check_call(["make", "-C", _HERE, "_build/libnw.so"])
        _nw = ctypes.CDLL(so)
        _nw.nw_align.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_int, ctypes.c_char_p,
                                 ctypes.POINTER(ctypes.c_int)]
        _nw.nw_align.restype = ctypes.c_int
    return _nw


def align_ops(query, target):
    """numpy uint8 array of ops ('=', 'X', 'I', 'D'); I consumes the query (read)."""
    buf = ctypes.create_string_buffer(len(query) + len(target) + 1)
    n = nw_lib().nw_align(query, len(query), target, len(target), buf, None)
    if n < 0:
        raise MemoryError("nw_align")
    return np.frombuffer(buf.raw[:n], np.uint8)


def parse_fastx(data, fastq):
    lines = data.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    step = 4 if fastq else 2
    recs = []
    for i in range(0, len(lines), step):
        hdr = lines[i].decode()
        info = dict(tok.split("=", 1) for tok in hdr.split(" ")[1:] if "=" in tok)
        recs.append((info, lines[i + 1], lines[i + 3] if fastq else None))
    return recs


HP_MAX = 10      # homopolymer run lengths 1..HP_MAX (longer runs are pooled into HP_MAX)
HP_OBS = 16      # observed lengths 0..HP_OBS-1
POS_EDGE = 20    # qscore histograms: first / last POS_EDGE read bases vs the middle


_PAR = {}


def _summarise_part(span):
    lo, hi = span
    return _summarise_reads(_PAR["reads"][lo:hi], _PAR["perf"], lo)


def summarise(fastq_bytes, perfect_bytes, max_reads=None, procs=None):
    """Statistics of the reads; large inputs are cut into spans that worker processes (fork) reduce in parallel."""
    reads = parse_fastx(fastq_bytes, True)
    perf = {}
    for info, seq, _ in parse_fastx(perfect_bytes, False):
        perf.setdefault(info["molecule_id"], seq)
    if max_reads:
        reads = reads[:max_reads]
    procs = procs or min(os.cpu_count() or 1, 16)
    if len(reads) < 2000 or procs < 2:
        out = _summarise_reads(reads, perf, 0)
        out.pop("_tot")
        return out
    import multiprocessing as mp
    nw_lib()
    _PAR["reads"], _PAR["perf"] = reads, perf
    per = (len(reads) + procs - 1) // procs
    spans = [(i, min(len(reads), i + per)) for i in range(0, len(reads), per)]
    try:
        with mp.get_context("fork").Pool(len(spans)) as pool:
            parts = pool.map(_summarise_part, spans)
    finally:
        _PAR.clear()
    out = parts[0]
    for q in parts[1:]:
        for k, v in q.items():
            if k in ("rates", "n_reads", "_tot"):
                continue
            if isinstance(v, list) and v and isinstance(v[0], list):
                out[k] = (np.array(out[k]) + np.array(v)).tolist()
            elif k in ("ins_len", "del_len"):
                out[k] = (np.array(out[k]) + np.array(v)).tolist()
            else:
                out[k] = out[k] + v
    out["n_reads"] = len(reads)
    tot = np.array(out.pop("_tot"))
    for q in parts[1:]:
        tot = tot + np.array(q["_tot"])
    n_cols, n_raw = int(tot.sum()), int(tot[0] + tot[1] + tot[3])
    out["rates"] = {"X": tot[1] / n_raw, "I": tot[2] / n_raw, "D": tot[3] / n_raw, "identity": tot[0] / n_cols}
    return out


def _summarise_reads(reads, perf, first_index):
    E, X, I, D = ord("="), ord("X"), ord("I"), ord("D")
    len_ratio, identity, hdr_identity, sub, ins, dele, meanq, rawlen = [], [], [], [], [], [], [], []
    hp_err, hp_del, mq_cls = [], [], [[], [], []]   # per read: the unit the samples are independent in
    hp = np.zeros((HP_MAX + 1, HP_OBS), np.int64)
    hp1 = np.zeros((HP_MAX + 1, HP_OBS), np.int64)  # the same over ONE run per read and run length (independent units)
    qh = np.zeros((3, 3, 94), np.int64)           # [position class][op class =,X,I][q]
    qh1 = np.zeros((3, 3, 94), np.int64)          # the same over ONE sampled base per read, position and op class
    ins_len = np.zeros(12, np.int64)
    del_len = np.zeros(12, np.int64)
    tot = np.zeros(4, np.int64)
    for ridx, (info, seq, qual) in enumerate(reads, first_index):
        raw = perf[info["molecule_id"]]
        ops = align_ops(seq, raw)
        ne, nx, ni, nd = (int((ops == c).sum()) for c in (E, X, I, D))
        tot += (ne, nx, ni, nd)
        L = len(raw)
        rawlen.append(L)
        len_ratio.append(len(seq) / L)
        identity.append(ne / len(ops))
        hdr_identity.append(float(info["read_identity"].rstrip("%")) / 100.0)
        sub.append(nx / L)
        ins.append(ni / L)
        dele.append(nd / L)
        # run-length spectra of I and D
        for code, acc in ((I, ins_len), (D, del_len)):
            m = (ops == code).astype(np.int8)
            if m.any():
                d = np.diff(np.concatenate(([0], m, [0])))
                runs = np.flatnonzero(d == -1) - np.flatnonzero(d == 1)
                np.add.at(acc, np.minimum(runs, 11), 1)
        # qscores by op class of the read base and by position class
        q = np.frombuffer(qual, np.uint8).astype(np.int64) - 33
        meanq.append(float(q.mean()) if len(q) else 0.0)
        rops = ops[ops != D]                       # one op per read base
        cls = np.where(rops == E, 0, np.where(rops == X, 1, 2))
        pos = np.ones(len(q), np.int64)
        pos[:POS_EDGE] = 0
        pos[max(len(q) - POS_EDGE, 0):] = 2
        np.add.at(qh, (pos, cls, np.clip(q, 0, 93)), 1)
        # independent units for a chi-square test: bases of one read share its identity, so one base per
        # read is drawn (seeded by the read's index) for every (position class, op class) cell
        srng = np.random.default_rng(977 * ridx + 13)
        cell = pos * 3 + cls
        for ce in range(9):
            idx = np.flatnonzero(cell == ce)
            if len(idx):
                qh1[ce // 3, ce % 3, min(max(int(q[idx[srng.integers(len(idx))]]), 0), 93)] += 1
        for oc in range(3):
            sel = q[cls == oc]
            mq_cls[oc].append(float(sel.mean()) if len(sel) else -1.0)
        # homopolymer spectrum: per run of the molecule, count read bases of the run's base inside its columns
        rawa = np.frombuffer(raw, np.uint8)
        seqa = np.frombuffer(seq, np.uint8)
        tcol = np.flatnonzero(ops != I)            # alignment column of every molecule base
        qidx = np.cumsum(ops != D) - 1             # read index at every column (valid where op != D)
        bnd = np.flatnonzero(np.diff(rawa) != 0) + 1
        starts = np.concatenate(([0], bnd))
        ends = np.concatenate((bnd, [L]))
        colbase = np.where(ops != D, seqa[np.clip(qidx, 0, len(seqa) - 1)], 0)
        n_hp = n_hp_err = n_hp_del = 0
        # the run of each length class that goes into hp1: drawn with the read's own generator
        rls = np.minimum(ends - starts, HP_MAX)
        pick = {}
        for rl in range(1, HP_MAX + 1):
            idx = np.flatnonzero(rls == rl)
            if len(idx):
                pick[int(idx[srng.integers(len(idx))])] = rl
        for ri, (a, b) in enumerate(zip(starts, ends)):
            rl = b - a
            if rl < 3:
                obs = min(int((colbase[tcol[a]:tcol[b - 1] + 1] == rawa[a]).sum()), HP_OBS - 1)
                hp[min(rl, HP_MAX), obs] += 1
                if ri in pick:
                    hp1[min(rl, HP_MAX), obs] += 1
                continue
            c0, c1 = tcol[a], tcol[b - 1]
            # extend over insertion columns adjacent to the run
            while c0 > 0 and ops[c0 - 1] == I:
                c0 -= 1
            while c1 + 1 < len(ops) and ops[c1 + 1] == I:
                c1 += 1
            obs = int((colbase[c0:c1 + 1] == rawa[a]).sum())
            hp[min(rl, HP_MAX), min(obs, HP_OBS - 1)] += 1
            if ri in pick:
                hp1[min(rl, HP_MAX), min(obs, HP_OBS - 1)] += 1
            n_hp += 1
            n_hp_err += obs != rl
            n_hp_del += obs < rl
        hp_err.append(n_hp_err / n_hp if n_hp else -1.0)
        hp_del.append(n_hp_del / n_hp if n_hp else -1.0)
    n_cols = int(tot.sum())
    n_raw = int(tot[0] + tot[1] + tot[3])
    return {
        "_tot": tot.tolist(),
        "n_reads": len(reads),
        "raw_len": rawlen,
        "len_ratio": len_ratio,
        "identity": identity,
        "hdr_identity": hdr_identity,
        "sub_rate": sub, "ins_rate": ins, "del_rate": dele, "mean_q": meanq,
        "hp_err": hp_err, "hp_del": hp_del, "mean_q_eq": mq_cls[0], "mean_q_x": mq_cls[1], "mean_q_i": mq_cls[2],
        "rates": {"X": tot[1] / n_raw, "I": tot[2] / n_raw, "D": tot[3] / n_raw, "identity": tot[0] / n_cols},
        "hp": hp.tolist(), "hp1": hp1.tolist(), "qhist": qh.tolist(), "qhist1": qh1.tolist(), "ins_len": ins_len.tolist(), "del_len": del_len.tolist(),
    }
