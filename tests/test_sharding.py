"""Multi-GPU host logic on CPU (world_size 2, gloo): the shard plan cuts the MDF at record
boundaries, balances simulated bases, hands every rank the global ordinal of its first read, and
the rank outputs concatenate to the single-process bytes.  The CPU twin stands in for the device
(it is bit-identical to the CUDA path, tests/test_gpu_parity.py); the GPU suite runs the same plan
through the kernels."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import seq_oracle as so


def make_world():
    rng = np.random.default_rng(3)
    g = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=120_000).tobytes().decode()
    return ">chr1\n" + g[:60000] + "\n>chr2 second\n" + g[60000:] + "\n"


def make_mdf(s, n=240):
    mdf = s.synth_mdf(1, n, seed=9).decode()
    # depth > 1 records and an empty-depth record: ordinals must follow the expanded stream
    return "+d3\t3\tx;\nchr1\t100\t400\t+\t\n+d0\t0\t\nchr1\t5\t50\t-\t\n" + mdf + "+tail\t2\t\nchr2\t10\t900\t-\t\n"


def twin_bytes(s, fasta, mdf, ordinal_base):
    refs = so.load_fasta(fasta)
    mols = so.parse_mdf(mdf)
    seqs = [so.molecule_sequence(refs, iv) for _, iv in mols]
    ids = [m for m, _ in mols]
    return so.twin_simulate(s.tables(), seqs, ids, ordinal_base=ordinal_base, badread=True, fastq=True)


def _rank_main(rank, world, port, result_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from tksm_b200.sequence import Sequencer
    fasta = make_world()
    s = Sequencer(seed=5, host_only=True)
    s.add_reference_text(fasta)
    s.load_models("random", "random")
    mdf = make_mdf(s)
    plan = s.shard_plan(mdf, world)
    b0, b1, ob, nreads = plan[rank]
    shard = mdf.encode()[b0:b1].decode()
    out = twin_bytes(s, fasta, shard, ob) if b1 > b0 else b""
    # exchange sizes and read counts (the only collective of a sharded run: bookkeeping, not data)
    mine = torch.tensor([len(out), nreads, ob], dtype=torch.int64)
    allv = [torch.zeros(3, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(allv, mine)
    with open(os.path.join(result_dir, f"rank{rank}.fastq"), "wb") as f:
        f.write(out)
    dist.barrier()
    if rank == 0:
        whole = twin_bytes(s, fasta, mdf, 0)
        cat = b"".join(open(os.path.join(result_dir, f"rank{r}.fastq"), "rb").read() for r in range(world))
        assert cat == whole, "rank outputs do not concatenate to the single-process bytes"
        assert sum(int(v[0]) for v in allv) == len(whole)
        assert [int(v[2]) for v in allv] == [p[2] for p in plan]
        assert sum(int(v[1]) for v in allv) == len(so.parse_mdf(mdf)) 
    dist.barrier()
    dist.destroy_process_group()


def free_port():
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        return sk.getsockname()[1]


def test_two_ranks_concatenate_to_single_run(built, tmp_path):
    mp.spawn(_rank_main, args=(2, free_port(), str(tmp_path)), nprocs=2, join=True)


def test_shard_plan_properties(built):
    from tksm_b200.sequence import Sequencer
    s = Sequencer(seed=5, host_only=True)
    s.add_reference_text(make_world())
    mdf = make_mdf(s, 500).encode()
    total_reads = len(so.parse_mdf(mdf.decode()))
    for world in (1, 2, 3, 8, 1000):
        plan = s.shard_plan(mdf, world)
        assert plan[0][0] == 0 and plan[-1][1] == len(mdf)
        assert all(a[1] == b[0] for a, b in zip(plan, plan[1:]))              # contiguous
        assert all(b0 == len(mdf) or mdf[b0:b0 + 1] == b"+" for b0, _, _, _ in plan)  # record boundaries
        assert sum(p[3] for p in plan) == total_reads
        ob = 0
        for b0, b1, base, n in plan:
            assert base == ob and n == len(so.parse_mdf(mdf[b0:b1].decode()))
            ob += n
        if world in (2, 3, 8):
            # balanced by simulated bases within one record
            w = [sum(sum(e - st for _, st, e, _, _ in iv) for _, iv in so.parse_mdf(mdf[b0:b1].decode())) for b0, b1, _, _ in plan]
            assert max(w) - min(w) <= 0.15 * sum(w) / world + 6000, w
    assert s.shard_plan(b"", 2) == [(0, 0, 0, 0), (0, 0, 0, 0)]


def test_shard_plan_of_a_large_text_on_several_threads(built):
    """Above 8 MiB the light pass of the plan runs on host threads over ranges that start at header lines
    (csrc/host.cpp: plan_shards).  The plan must be the one the pass over the whole text gives: checked against a plain
    restatement of that pass (depth x sum of interval lengths per record, + 1 per read, cut where the running weight
    reaches i / world of the total), for LF and CRLF texts and with lines before the first header."""
    from tksm_b200.sequence import Sequencer
    s = Sequencer(seed=5, host_only=True)
    s.add_reference_text(make_world())
    base = s.synth_mdf(1, 45000, seed=8) + s.synth_mdf(3, 8000, seed=9) + s.synth_mdf(2, 300, seed=10, min_len=500, max_len=60000)
    assert len(base) > (8 << 20)

    def restated(text, world):
        recs, pos = [], 0
        for ln in text.split(b"\n"):
            line = ln[:-1] if ln.endswith(b"\r") else ln
            if line.startswith(b"+"):
                depth = int(line.split(b"\t")[1])
                recs.append([pos, max(depth, 0), 0])
            elif recs and line:
                f = line.split(b"\t")
                a, b = int(f[1]), int(f[2])
                if b > a >= 0:
                    recs[-1][2] += (b - a) * recs[-1][1]
            pos += len(ln) + 1
        total = sum(r[2] + r[1] for r in recs)
        begin, ordinal = [0], [0]
        acc = reads = 0
        shard = 1
        for i, (off, n, bases) in enumerate(recs):
            while shard < world and acc * world >= total * shard and i > 0:
                begin.append(off)
                ordinal.append(reads)
                shard += 1
            acc += bases + n
            reads += n
        while shard < world:
            begin.append(len(text))
            ordinal.append(reads)
            shard += 1
        begin.append(len(text))
        ordinal.append(reads)
        return [(begin[i], begin[i + 1], ordinal[i], ordinal[i + 1] - ordinal[i]) for i in range(world)]

    for text in (base, base.replace(b"\n", b"\r\n"), b"chr1\t5\t9\t+\t\n" + base):
        for world in (1, 3, 8):
            assert s.shard_plan(text, world) == restated(text, world), world
