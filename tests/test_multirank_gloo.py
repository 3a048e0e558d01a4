"""CPU, world_size 2 over gloo: the multi-GPU scheme of the path -- reads sharded by batch
(batch b -> rank b mod G), index replicated, ONE sum-reduce of counts / kmc / nmapread / counters at the end,
per-pair records kept sharded and merged by pair index -- gives bit-identical outputs to a single rank.
Each rank computes its shard with the oracle (no GPU here); the reduce and the merge are what is under test."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, mode, q):
    sys.path.insert(0, HERE)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch
    import torch.distributed as dist
    from oracle_abi import Oracle, abi, rpgg_io
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ix, p, seqs = _case(mode)
    npairs = len(seqs) // 2
    bsz = 128                                                       # pairs per batch; batch b belongs to rank b % world
    o = Oracle(ix)
    acc, recs = None, []
    for b in range((npairs + bsz - 1) // bsz):
        if b % world != rank:
            continue
        b0, b1 = b * bsz, min(npairs, (b + 1) * bsz)
        bases, off = rpgg_io.seqs_to_batch(seqs[2 * b0: 2 * b1])
        r, cnt, trkmc, kmc, nmap = o.run(p, bases, off, first_pair_index=b0, acc=acc)
        # keep the padded accumulators the oracle works on
        acc = (np.concatenate([trkmc, np.zeros(max(o.nk_tr, 1) - len(trkmc), np.uint64)]), np.concatenate([kmc, np.zeros(max(o.nloci, 1) - len(kmc), np.uint64)]),
               np.concatenate([nmap, np.zeros(max(o.nloci, 1) - len(nmap), np.uint32)]), cnt)
        recs.append(r.copy())
    trkmc, kmc, nmap, cnt = acc
    red = [torch.from_numpy(a.astype(np.int64)) for a in (trkmc, kmc, nmap, cnt)]
    for t in red:                                                    # the path's one exchange step: sum over ranks
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    mine = np.concatenate(recs) if recs else None
    gathered = [None] * world
    dist.all_gather_object(gathered, None if mine is None else mine.tobytes())
    if rank == 0:
        q.put(([t.numpy().copy() for t in red], gathered, mine.dtype.itemsize))
    dist.barrier()
    dist.destroy_process_group()


def _case(mode):
    sys.path.insert(0, HERE)
    from oracle_abi import abi
    if mode == "asgn":
        from common import synthetic_case
        ix, seqs, _ = synthetic_case(nloci=12, seed=5, npairs=700, read_seed=9, frac_locus=0.6, frac_chimeric=0.1, frac_border=0.1)
        p = abi.default_params(k=21, cth=30, flags=abi.F_KAM | abi.F_TRACE, max_batch_pairs=4096)
    else:
        import thread_common as tc
        _, ix, seqs, _ = tc.thread_case(write_text=False, frac_bg=0.2, nloci=6, seed=11, npairs=500)
        p = tc.thread_params(extra_flags=abi.F_TRACE, max_batch_pairs=4096)
    return ix, p, seqs


@pytest.mark.parametrize("mode", ["asgn", "thread"])
def test_two_ranks_sharded_reads_one_sum_reduce_equals_single_rank(mode):
    import torch.multiprocessing as mp
    from oracle_abi import Oracle, rpgg_io
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, mode, q)) for r in range(2)]
    for pr in procs:
        pr.start()
    red, gathered, itemsize = q.get(timeout=240)
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    ix, p, seqs = _case(mode)
    bases, off = rpgg_io.seqs_to_batch(seqs)
    recs, cnt, trkmc, kmc, nmap = Oracle(ix).run(p, bases, off)
    assert np.array_equal(red[0][: len(trkmc)].astype(np.uint64), trkmc)
    assert np.array_equal(red[1][: len(kmc)].astype(np.uint64), kmc)
    assert np.array_equal(red[2][: len(nmap)].astype(np.uint32), nmap)
    assert np.array_equal(red[3].astype(np.uint64), cnt)
    # records stay sharded; the writer merges them by pair index (first 8 bytes of every record)
    allrec = np.concatenate([np.frombuffer(g, np.uint8).reshape(-1, itemsize) for g in gathered if g])
    order = np.argsort(allrec[:, :8].copy().view("<u8").ravel(), kind="stable")
    assert allrec[order].tobytes() == recs.tobytes()
