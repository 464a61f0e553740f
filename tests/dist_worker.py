"""One rank of a multi-process AnchorFinder job (launched by test_dist_*.py).

usage: dist_worker.py MODE CONFIG SCALE K
  MODE standin: numpy stages (CPU, gloo);  cuda: the C ABI stages, gloo when the
  ranks share one GPU, NCCL otherwise.
Every rank holds an interleaved part of the job's sequences (a placement the
finder has to repair); rank 0 compares the gathered result with the oracle on
the whole job and exits non-zero on any difference."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, os.path.join(ROOT, "oracle"), HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np
import torch
import torch.distributed as dist

import oracle
from npge_b200 import synth
from npge_b200.dist import DistAnchorFinder


def placement(n, rank, world, how):
    if how == "interleaved":
        return [i for i in range(n) if i % world == rank]
    per = (n + world - 1) // world            # "blocks": caller order, contiguous
    return list(range(rank * per, min(n, (rank + 1) * per)))


def main():
    mode, config, scale, k = sys.argv[1], sys.argv[2], float(sys.argv[3]), int(sys.argv[4])
    how = sys.argv[5] if len(sys.argv) > 5 else "interleaved"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    device = 0
    if mode == "cuda":
        n_gpus = torch.cuda.device_count()
        device = rank % n_gpus
        torch.cuda.set_device(device)
        backend = "nccl" if n_gpus >= world else "gloo"
        stages = None
    else:
        from dist_standin import NumpyStages
        backend, stages = "gloo", NumpyStages()
    dist.init_process_group(backend, rank=rank, world_size=world)
    if config == "oneseq":  # one chromosome (+ a fragment): fewer sequences than ranks
        names, parts = synth.generate("tiny", scale=scale)
        seqs = [np.concatenate(parts), parts[1][50:400]]
    else:
        names, seqs = synth.generate(config, scale=scale)
    if config == "tiny":  # some short and some equal-sized sequences too
        seqs = seqs + [seqs[0][:k - 1], seqs[1][100:700], seqs[2][50:650], seqs[0][:0]]
    parts = [placement(len(seqs), r, world, how) for r in range(world)]
    mine = [seqs[i] for i in parts[rank]]
    # job-wide caller index: the ranks' sequences one after the other
    job = [seqs[i] for r in range(world) for i in parts[r]]
    f = DistAnchorFinder(device=device, stages=stages)
    fails = []
    cases = [(False, 2**62, None), (True, 2**62, None)]
    if mode == "cuda":
        cases += [(True, 1000, None), (False, 777, None), (True, 2**62, "used")]
    for similar, limit, used in cases:
        if used:
            first = oracle.find_ideal(job, k, similar=similar, max_frags=limit)
            used = first.block_hash[::3].copy()
        got_local = f.find(mine, k, similar=similar, max_frags=limit, used=used)
        if mode == "cuda":
            # the compact layout of the resident result: the same slice on every rank
            from npge_b200 import _abi
            f.stages.set_result_layout(_abi.LAYOUT_COMPACT)
            f.find_device(k, similar=similar, max_frags=limit, used=used)
            cl = f.fetch_compact()
            f.stages.set_result_layout(_abi.LAYOUT_FULL)
            if (cl.blocks() != got_local.blocks() or
                    list(cl.block_hash) != list(got_local.block_hash)):
                fails.append("rank %d similar=%s limit=%d: compact slice differs"
                             % (rank, similar, limit))
        got = f.gather(got_local)
        if rank != 0:
            continue
        want = oracle.find_ideal(job, k, similar=similar, max_frags=limit, used=used)
        tag = "similar=%s limit=%d used=%s" % (similar, limit, used is not None)
        if got.blocks() != want.blocks():
            fails.append(tag + ": blocks differ (%d vs %d)" % (got.n_blocks, want.n_blocks))
        if list(got.block_hash) != list(want.block_hash):
            fails.append(tag + ": block hashes differ")
        if list(got.new_used) != list(want.new_used):
            fails.append(tag + ": new_used differs")
        if mode == "cuda":
            if got.digest != want.digest:
                fails.append(tag + ": digest differs")
            if limit >= 2**62 and got_local.job["digest"] != want.digest:
                fails.append(tag + ": device digest differs")
            if got_local.job["n_fragments"] != want.n_fragments and limit >= 2**62:
                fails.append(tag + ": job counts differ")
        print("rank0", tag, "blocks", got.n_blocks, "frags", got.n_fragments,
              "cuts", got_local.job.get("bucket_cuts"), flush=True)
    if mode == "cuda" and f.stages.debug_check() != 0:
        fails.append("guard bytes behind a device buffer were overwritten")
    dist.barrier()
    dist.destroy_process_group()
    if fails:  # (a rank other than 0 fails on its own checks only)
        print("\n".join(fails), flush=True)
        sys.exit(1)


if __name__ == "__main__":
    main()
