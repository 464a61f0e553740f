#!/usr/bin/env python
"""bench.py -- AnchorFinder throughput on synthetic genomes (BASELINE.json).

A step = one pass of the hot path (AnchorFinder::run_impl,
src/algo/AnchorFinder.cpp:393-406) over the whole synthetic genome set:
  value   bases scanned per second with the packed genome resident in HBM and
          the anchors left in HBM (npge_anchor_find_device), CUDA events around
          the whole step (the same definition at every N), L2 flushed between
          steps
  e2e     the same metric through the drop-in call npge_anchor_find with HOST
          buffers: pinned ASCII bases in, Fragment arrays out, copies included
  parity  counts and 128-bit digest of every timed result against
          tests/golden/digests.json (written by the CPU oracle in hash-range
          slices, tests/golden/make_digests.py); the run FAILS (rc 1) on a mismatch
  c5      the north-star configuration (1 Gbp pangenome, k = 20) sharded over the
          N GPUs of the run (strong scaling), its own parity check, and at N = 1
          the k = 24 / 28 / 32 sweep
  --impl reference   the reference's CPU algorithm (oracle/ restatement; the
          reference itself cannot be built here) on the host cores,
          on the same full-size workload

Headline workload: BASELINE.json configs[1], 17 synthetic Brucella-shaped genomes
(~56 Mbp), anchor-size 20, parity mode (anchor-similar=false, no truncation);
the default options are timed beside it (`default_mode`).  At N>1 every rank
holds its own 17 strains (weak scaling), k-mers are repartitioned by hash range.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOAD = "brucella17"
C5 = "pangenome1g"
K = 20
BIG = 2 ** 62
METRIC = "Gbp scanned/sec (AnchorFinder)"
UNIT = "Gbp/s"


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def golden(workload, k, mode):
    """(n_fragments, n_blocks, digest) the CPU oracle holds for a workload, or None."""
    path = os.path.join(ROOT, "tests", "golden", "digests.json")
    if not os.path.exists(path):
        return None
    with open(path) as f:
        g = json.load(f).get("%s|k=%d|%s" % (workload, k, mode))
    if g is None:
        return None
    return g["n_fragments"], g["n_blocks"], tuple(int(x, 16) for x in g["digest"])


def check_parity(workload, k, mode, st):
    """Compares a job's counts and digest with the golden entry."""
    want = golden(workload, k, mode)
    got = (st["n_fragments_global"], st["n_blocks_global"], tuple(st["digest"]))
    out = {"workload": workload, "anchor_size": k, "mode": mode,
           "source": "tests/golden/digests.json", "fragments": got[0], "blocks": got[1],
           "digest": ["%016x" % d for d in got[2]]}
    if want is None:
        out["digest_equal"] = None
        out["note"] = "no golden entry"
    else:
        out["digest_equal"] = bool(got == want)
        if got != want:
            out["expected"] = {"fragments": want[0], "blocks": want[1],
                               "digest": ["%016x" % d for d in want[2]]}
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,"
              "clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.out = open(self.path, "w")
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu),
                 "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=self.out, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.out.close()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                 "sw_power_cap"]
        with open(self.path) as f:
            for line in f:
                parts = [x.strip() for x in line.split(",")]
                if len(parts) < 8:
                    continue
                try:
                    sm.append(float(parts[1]))
                    mx.append(float(parts[2]))
                    power.append(float(parts[3]))
                except ValueError:
                    continue
                for n, v in zip(names, parts[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)),
                "power_w_max": float(max(power)), "samples": len(sm),
                "reasons": sorted(reasons)}


def base_config(world, n_total, n_seqs):
    """The keys both arms share (the reference arm prints exactly these)."""
    return {
        "workload": WORKLOAD + (" x%d (one 17-strain cluster per GPU, independent "
                                "ancestors)" % world if world > 1 else ""),
        "anchor_size": K,
        "mode": "parity (anchor-similar=false, no truncation)",
        "bases_per_step": int(n_total), "sequences": int(n_seqs),
    }


# ---------------------------------------------------------------------------
def run_reference(args, rank, world):
    """The reference's CPU algorithm (two-pass Bloom AnchorFinder restated in
    oracle/anchor_oracle.cpp) on the host cores, all threads.  N = 1: the GPU
    arm's workload at full size (same config).  N > 1: the job is N clusters;
    every step runs ONE of them (rank 0's, 56 Mbp) so that the run stays within
    minutes -- per-base throughput of the CPU does not depend on which."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    from npge_b200 import synth
    cores = os.cpu_count() or 1
    if world == 1:
        names, seqs = synth.generate(WORKLOAD, scale=args.cpu_scale)
    else:
        names, seqs = synth.generate_shard(WORKLOAD, 0, world, scale=args.cpu_scale)
    n = synth.total_bases(seqs)
    gid = synth.genome_ids(names)
    times = []
    frags = 0
    for i in range(args.warmup + args.steps):
        t = time.perf_counter()
        r = oracle.find_faithful(seqs, K, similar=False, max_frags=BIG,
                                 workers=cores, genome_ids=gid, asserts=True)
        dt = time.perf_counter() - t
        frags = r.n_fragments
        if i >= args.warmup:
            times.append(dt)
    total = sum(times)
    value = n * len(times) / total / 1e9
    sample = ("%s%s at scale %.3g: %d sequences, %.2f Mbp, k=%d, parity mode; "
              "two-pass Bloom restatement (oracle find_faithful, range "
              "checks on), %d std::threads, one task per sequence"
              % (WORKLOAD, "" if world == 1 else " (one of the job's %d clusters per step)" % world,
                 args.cpu_scale, len(seqs), n / 1e6, K, cores))
    config = base_config(world, n * world if args.cpu_scale == 1.0 else n, len(seqs) * world)
    if world > 1 or args.cpu_scale != 1.0:
        config["cpu_sample_bases_per_step"] = n
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic", "config": config,
        "anchors_per_s": frags * len(times) / total,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores,
                         "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ---------------------------------------------------------------------------
def cpu_baseline(budget_s):
    """The CPU legs of SURVEY 8(d), on this box's host cores, beside the GPU run:
    the headline workload (brucella17, full size, all threads, parity mode: one
    run) and the designated CPU configuration C1 (brucella3) with workers = 1 and
    all cores, range checks on and off, parity mode and the default options."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    from npge_b200 import synth
    cores = os.cpu_count() or 1
    t_all = time.perf_counter()
    names, seqs = synth.generate("brucella3")
    n1 = synth.total_bases(seqs)
    gid = synth.genome_ids(names)
    legs = []
    for mode, similar, limit in (("parity", False, BIG), ("default", True, 100000)):
        for workers in (cores, 1):
            for asserts in (True, False):
                if time.perf_counter() - t_all > budget_s:
                    continue
                t = time.perf_counter()
                r = oracle.find_faithful(seqs, K, similar=similar, max_frags=limit,
                                         workers=workers, genome_ids=gid, asserts=asserts)
                dt = time.perf_counter() - t
                legs.append({"workload": "brucella3", "mode": mode, "workers": workers,
                             "range_checks": asserts, "seconds": dt,
                             "value": n1 / dt / 1e9, "fragments": r.n_fragments})
    names, seqs = synth.generate(WORKLOAD)
    n = synth.total_bases(seqs)
    gid = synth.genome_ids(names)
    t = time.perf_counter()
    r = oracle.find_faithful(seqs, K, similar=False, max_frags=BIG,
                             workers=cores, genome_ids=gid, asserts=True)
    dt = time.perf_counter() - t
    t1 = time.perf_counter()
    oracle.find_faithful(seqs, K, similar=True, max_frags=100000,
                         workers=cores, genome_ids=gid, asserts=True)
    dt_default = time.perf_counter() - t1
    return {
        "value": n / dt / 1e9, "unit": UNIT, "cores": cores, "kind": "port",
        "sample": ("%s at full size (%d sequences, %.2f Mbp), k=%d, parity "
                   "mode, oracle find_faithful (two-pass Bloom restatement of "
                   "AnchorFinder.cpp, range checks on), %d threads, one run of %.1f s"
                   % (WORKLOAD, len(seqs), n / 1e6, K, cores, dt)),
        "anchors_per_s": r.n_fragments / dt,
        "default_mode_value": n / dt_default / 1e9,
        "c1": {"workload": "brucella3", "bases": n1, "cores": cores, "unit": UNIT,
               "note": "SURVEY 8(d): the designated CPU configuration; the reference "
                       "runs one task per sequence (6 here), so more than 6 workers idle",
               "legs": legs},
    }


def timed_steps(runner, torch, flush, steps, k, similar, limit):
    """`steps` timed passes; returns (sum of whole-step ms, list of stats)."""
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    total, out = 0.0, []
    for _ in range(steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        e0.record()
        st = runner.step(k, similar, limit)
        e1.record()
        e1.synchronize()
        total += e0.elapsed_time(e1)
        out.append(st)
    return total, out


def max_over_ranks(x, world, dist, torch):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_c5(args, rank, world, runner_factory, torch, dist, flush, hbm_peak):
    """The 1 Gbp pangenome sharded over the N GPUs of this run (strong scaling)."""
    from npge_b200 import synth
    from npge_b200.engine import concat
    names, seqs = synth.generate(C5)
    n_seqs = len(seqs)
    n_total = synth.total_bases(seqs)
    per = (n_seqs + world - 1) // world  # caller order, contiguous: job index = index
    mine = seqs[rank * per:(rank + 1) * per]
    bases, offsets = concat(mine)
    del seqs, mine
    runner = runner_factory(bases, offsets)
    steps = max(3, min(args.steps, 5))
    out = {"workload": C5, "anchor_size": K, "n_gpus": world, "bases": n_total,
           "sequences": n_seqs, "scaling": "strong", "steps": steps,
           "mode": "parity (anchor-similar=false, no truncation)"}
    sweep = []
    for k in ((20, 24, 28, 32) if world == 1 else (20,)):
        runner.upload(k)
        for _ in range(2):
            runner.step(k, False, BIG)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms, sts = timed_steps(runner, torch, flush, steps, k, False, BIG)
        ms = max_over_ranks(ms, world, dist, torch) / steps
        st = runner.finish(sts[-1])
        par = check_parity(C5, k, "parity", st)
        row = {"anchor_size": k, "ms_per_step": ms, "gbp_s": n_total / (ms * 1e-3) / 1e9,
               "fragments": st["n_fragments_global"], "blocks": st["n_blocks_global"],
               "digest_equal": par["digest_equal"],
               "survey_model_frac": st["bytes_algorithmic_job"] / (ms * 1e-3) / 1e9 / hbm_peak / world,
               "compulsory_frac": st["bytes_moved_job"] / (ms * 1e-3) / 1e9 / hbm_peak / world}
        if k == K:
            out.update(row)
            out["parity"] = par
            out["kernel_ms_rank0"] = float(np.mean([x["ms_total"] for x in sts]))
            if world > 1:
                sent = float(st.get("nvlink_bytes_sent", 0))  # rank 0, the last timed step
                exposed = float(st.get("nvlink_ms", 0.0))
                window = float(st["ms_pass"][6])   # start of pass A .. last copy landed
                chunks = int(st.get("chunks", 1))
                pass_a = float(st["ms_pass"][0]) - exposed
                # the copies of chunk c start when pass A has finished it: they are in
                # flight from the end of the first chunk to the end of the window
                busy = max(window - pass_a / chunks, 1e-6)
                out["nvlink"] = {
                    "bytes_per_gpu": sent, "exposed_ms": exposed, "window_ms": window,
                    "pass_a_ms": pass_a, "chunks": chunks, "copy_ms": busy,
                    "gb_s": sent / (busy * 1e-3) / 1e9,
                    "frac_of_900": sent / (busy * 1e-3) / 900e9,
                    "survey_model_bytes_per_gpu": 12.0 * n_total / world * (world - 1) / world,
                    "exchange": st.get("exchange"),
                    "note": "rank 0: bytes it wrote into the other ranks' HBM over the time "
                            "its copies were in flight (window - pass A / chunks); exposed_ms "
                            "is what they add to the step after pass A"}
            # default options on the same resident genome
            for _ in range(2):
                runner.step(k, True, 100000)
            dms, dsts = timed_steps(runner, torch, flush, steps, k, True, 100000)
            dms = max_over_ranks(dms, world, dist, torch) / steps
            dst = runner.finish(dsts[-1])
            dpar = check_parity(C5, k, "default", dst)
            out["default_mode"] = {"ms_per_step": dms, "gbp_s": n_total / (dms * 1e-3) / 1e9,
                                   "digest_equal": dpar["digest_equal"],
                                   "fragments": dst["n_fragments_global"],
                                   "blocks": dst["n_blocks_global"]}
        else:
            sweep.append(row)
    if sweep:
        out["k_sweep"] = sweep
    runner.close()
    return out


def run_gpu(args, rank, world, local_rank):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the anchor finder has no "
                         "CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        guard_stdout()
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from npge_b200 import synth
    from npge_b200.engine import AnchorEngine, concat

    hbm_peak, peak_src = peaks()
    if world == 1:
        names, seqs = synth.generate(WORKLOAD)
        eng = AnchorEngine(local_rank)

        def runner_factory(b, o):
            return SingleRunner(eng, b, o)
    else:
        from npge_b200.dist import DistAnchorFinder
        names, seqs = synth.generate_shard(WORKLOAD, rank, world)
        finder = DistAnchorFinder(local_rank)

        def runner_factory(b, o):
            return DistRunner(finder, b, o)
    bases, offsets = concat(seqs)
    runner = runner_factory(bases, offsets)
    job_name = WORKLOAD if world == 1 else "%s_x%d" % (WORKLOAD, world)
    n_local = int(offsets[-1])
    n_total = n_local
    if world > 1:
        t = torch.tensor([n_local], dtype=torch.int64, device="cuda")
        dist.all_reduce(t)
        n_total = int(t.item())

    # L2 flush buffer: 512 MiB > 126 MB L2, rewritten between timed steps
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident: value ------------------------------------------
    runner.upload(K)
    for _ in range(args.warmup):
        runner.step(K, False, BIG)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    wall0 = time.perf_counter()
    step_ms, all_stats = timed_steps(runner, torch, flush, args.steps, K, False, BIG)
    barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.stop() if rank == 0 else None
    ev_ms = sum(st["ms_total"] for st in all_stats)
    hist_ms = sum(st["ms_hist"] for st in all_stats)
    pass_ms = [st["ms_pass"] for st in all_stats]
    kern_launches = sum(st["gpu_launches"] for st in all_stats)
    exch_ms, host_ms = {}, {}
    for st in all_stats:
        for nm, v in st.get("exchange_ms", {}).items():
            exch_ms[nm] = exch_ms.get(nm, 0.0) + v
        for nm, v in st.get("host_ms", {}).items():
            host_ms[nm] = host_ms.get(nm, 0.0) + v
    stats = all_stats[-1]
    # CUDA events around the whole step (kernels, exchanges, host tables and the
    # call's own synchronisation) at every N, max over ranks
    t_dev = max_over_ranks(step_ms, world, dist, torch) / 1e3
    parity_stats = runner.finish(stats)
    parity = check_parity(job_name, K, "parity", parity_stats)
    # the row next to the path: the reference's block_hash of
    # every anchor block, on the result that is still resident.  Reported beside
    # the headline, not part of it; never allowed to break the line.
    block_hash = None
    if world == 1 and not args.profile:
        try:
            best = None
            for _ in range(3):
                _, total = eng.anchor_block_hashes(names, want_hashes=False)
                ms = eng.block_hashes_ms()
                best = ms if best is None else min(best, ms)
            block_hash = {"kernel": "block_hash_kernel", "device_ms": best,
                          "blocks_per_s": parity_stats["n_blocks_global"] / (best * 1e-3),
                          "blockset_hash": "%016x" % total}
        except Exception as exc:  # noqa: BLE001
            block_hash = {"error": str(exc)[:200]}
    if args.profile:
        if rank == 0:
            emit({"profile_run": True, "steps": args.steps,
                  "ms_per_step": 1e3 * t_dev / args.steps,
                  "kernel_ms_per_step": ev_ms / args.steps,
                  "digest_equal": parity["digest_equal"],
                  "stats": {k: v for k, v in stats.items() if k != "digest"}})
        return

    # ---- default options (anchor-similar=true, 100000 fragments) -----------
    for _ in range(2):
        runner.step(K, True, 100000)
    dsteps = max(3, args.steps // 2)
    dflt_ms, dstats = timed_steps(runner, torch, flush, dsteps, K, True, 100000)
    dflt_ms = max_over_ranks(dflt_ms, world, dist, torch)
    default_stats = runner.finish(dstats[-1])
    parity_default = check_parity(job_name, K, "default", default_stats)

    # ---- end to end through the drop-in call, host buffers -------------------
    pinned = torch.from_numpy(bases).pin_memory()
    pinned_np = pinned.numpy()
    e2e_steps = max(3, min(args.steps, 5))
    runner.find_host(pinned_np, offsets, K, False, BIG)  # warm-up
    barrier()
    t0 = time.perf_counter()
    d2h = 0
    for _ in range(e2e_steps):
        d2h = runner.find_host(pinned_np, offsets, K, False, BIG)
    barrier()
    e2e_t = max_over_ranks(time.perf_counter() - t0, world, dist, torch)
    # the host-buffer call's own result against the golden digest
    h = runner.last_host
    e2e_phases = h.get("host_phases_ms")  # N > 1: rank 0's wall time by phase, last call
    e2e_parity = check_parity(job_name, K, "parity", {
        "n_fragments_global": h["n_fragments"], "n_blocks_global": h["n_blocks"],
        "digest": h["digest"]})["digest_equal"]
    h2d = int(bases.nbytes + offsets.nbytes)
    if world > 1:  # job totals of the copies and of the launches
        t = torch.tensor([h2d, int(d2h), kern_launches], dtype=torch.int64, device="cuda")
        dist.all_reduce(t)
        h2d, d2h, kern_launches = (int(x) for x in t.tolist())
    runner.find_host(pinned_np, offsets, K, True, 100000)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        d2h_default = runner.find_host(pinned_np, offsets, K, True, 100000)
    barrier()
    e2e_default_t = time.perf_counter() - t0
    e2e_full = None
    if world == 1:  # the full-layout call beside it (9 bytes per fragment over PCIe)
        runner.find_host(pinned_np, offsets, K, False, BIG, full=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            d2h_full = runner.find_host(pinned_np, offsets, K, False, BIG, full=True)
        torch.cuda.synchronize()
        e2e_full = {"value": n_total * e2e_steps / (time.perf_counter() - t0) / 1e9,
                    "d2h_bytes_per_step": int(d2h_full),
                    "call": "npge_anchor_find (fragment arrays seq / min_pos / ori out)"}
    del pinned, pinned_np

    # ---- the north-star configuration under the same clock ---------------------
    c5 = None
    if not args.no_c5:
        try:
            c5 = run_c5(args, rank, world, runner_factory, torch, dist, flush, hbm_peak)
        except Exception as exc:  # noqa: BLE001  (every rank fails alike or the job hangs)
            c5 = {"workload": C5, "error": str(exc)[:300]}

    if rank != 0:
        return
    value = n_total * args.steps / t_dev / 1e9
    nw = parity_stats.get("local_records", parity_stats["n_windows"])  # rank 0's
    # stages of one step (CUDA events inside the library):
    # count A + scan | pass A | count B + scan | pass B | bucket sort | scan + emit
    names_s = ["pass_a", "count_b_scan", "pass_b", "bucket_sort", "scan_emit"]
    stage = {"count_a_scan": hist_ms / args.steps}
    for i, nm in enumerate(names_s):
        stage[nm] = float(np.mean([p[i] for p in pass_ms]))
    stage["kernels"] = ev_ms / args.steps
    stage["step"] = step_ms / args.steps
    if world > 1:  # rank 0's view of the step
        for nm, v in exch_ms.items():
            stage["exchange_" + nm] = v / args.steps
        stage["host_wall_by_phase"] = {nm: v / args.steps for nm, v in host_ms.items()}
        stage["host_tables_and_syncs"] = (
            stage["step"] - stage["kernels"] - sum(exch_ms.values()) / args.steps)
    # dominant kernel: bucket_sort_kernel reads every record once and writes the
    # sorted unit back (8 B each way at k <= 21, 12 B above)
    rec_bytes = 12.0 if K > 21 else 8.0
    sort_ms = stage["bucket_sort"]
    sort_bytes = 2.0 * rec_bytes * nw
    achieved = sort_bytes / (sort_ms * 1e-3) / 1e9
    moved = parity_stats["bytes_moved"]
    moved_gbs = moved * args.steps / (ev_ms * 1e-3) / 1e9
    pipeline_bytes = parity_stats["bytes_algorithmic"]
    pipeline_gbs = pipeline_bytes * args.steps / (ev_ms * 1e-3) / 1e9
    config = base_config(world, n_total, int(len(offsets) - 1) * world)
    config.update({
        "l2": "512 MiB flush buffer rewritten between timed steps",
        "timed": "CUDA events around the whole step (kernels, exchanges, host tables, "
                 "the call's own synchronisation), max over ranks",
        "parallelism": ("1 GPU" if world == 1 else
                        "sequences sharded per GPU; k-mers repartitioned by hash "
                        "range x%d, exchange=%s" % (world, parity_stats.get("exchange", "?"))),
    })
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t_dev / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic", "config": config,
        "parity": parity,
        "anchors_per_s": parity_stats["n_fragments_global"] * args.steps / t_dev,
        "fragments": parity_stats["n_fragments_global"],
        "blocks": parity_stats["n_blocks_global"],
        "windows": parity_stats["n_windows_global"],
        "kernel_ms_per_step": ev_ms / args.steps,
        "wall_ms_per_step": 1e3 * wall / args.steps,
        "stage_ms": stage,
        "default_mode": {
            "value": n_total * dsteps / (dflt_ms * 1e-3) / 1e9, "unit": UNIT,
            "options": "anchor-similar=true, max-anchor-fragments=100000",
            "fragments": default_stats["n_fragments_global"],
            "blocks": default_stats["n_blocks_global"],
            "digest_equal": parity_default["digest_equal"]},
        "roofline": {
            "bound": "hbm", "kernel": runner.dominant_kernel,
            "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
            "frac": achieved / hbm_peak, "traffic": runner.ncu_traffic(nw),
            "peak_source": peak_src,
            "bytes_per_launch": sort_bytes, "ms_per_launch": sort_ms,
            "note": "the bucket sort is bound by shared-memory ranking, not by "
                    "HBM; see pipeline for the whole step",
            "pipeline": {"survey_model_bytes_per_step": pipeline_bytes,
                         "survey_model_gbs": pipeline_gbs,
                         "survey_model_frac": pipeline_gbs / hbm_peak,
                         "survey_model_bytes_per_base":
                             pipeline_bytes / float(n_local),
                         "compulsory_bytes_per_step": moved,
                         "compulsory_gbs": moved_gbs,
                         "compulsory_frac": moved_gbs / hbm_peak,
                         "over": "the library's own events, first to last kernel"}},
        "e2e": {"value": n_total * e2e_steps / e2e_t / 1e9, "unit": UNIT,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(d2h),
                "steps": e2e_steps,
                "call": runner.e2e_call, "digest_equal": e2e_parity,
                "phases_ms_rank0": e2e_phases,
                "full_layout": e2e_full,
                "default_mode": {
                    "value": n_total * e2e_steps / e2e_default_t / 1e9,
                    "d2h_bytes_per_step": int(d2h_default)}},
        "gpu_launches": int(kern_launches),
        "clocks": clocks,
    }
    if c5 is not None:
        line["c5"] = c5
    if block_hash is not None:
        line["block_hash"] = block_hash
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(args.cpu_budget)
    emit(line)
    bad = [p for p in (parity, parity_default, (c5 or {}).get("parity"))
           if p and p.get("digest_equal") is False]
    if e2e_parity is False:
        bad.append({"e2e": "digest of the host-buffer call differs"})
    if bad or (c5 and c5.get("default_mode", {}).get("digest_equal") is False):
        sys.stderr.write("bench.py: PARITY FAILURE against tests/golden/digests.json: %s\n"
                         % json.dumps(bad))
        sys.exit(1)


class SingleRunner:
    dominant_kernel = "bucket_sort_kernel"

    def __init__(self, eng, bases, offsets):
        self.eng, self.bases, self.offsets = eng, bases, offsets

    def upload(self, k):
        self.eng.upload((self.bases, self.offsets), k)

    def step(self, k, similar, max_frags):
        st = self.eng.find_device(k, similar=similar, max_frags=max_frags)
        st["n_fragments_global"] = st["n_fragments"]
        st["n_blocks_global"] = st["n_blocks"]
        st["n_windows_global"] = st["n_windows"]
        return st

    def finish(self, st):
        st["bytes_algorithmic_job"] = st["bytes_algorithmic"]
        st["bytes_moved_job"] = st["bytes_moved"]
        return st

    def close(self):
        pass

    e2e_call = ("npge_anchor_find_compact (pinned ASCII bases in; blocks + 4-byte "
                "(position, strand) fragments + layout table out, pinned)")

    def find_host(self, bases, offsets, k, similar, max_frags, full=False):
        if full:  # the 9-bytes-per-fragment arrays of npge_anchor_find
            r = self.eng.find_counts((bases, offsets), k, similar=similar,
                                     max_frags=max_frags)
        else:
            r = self.eng.find_compact((bases, offsets), k, similar=similar,
                                      max_frags=max_frags, counts=True)
        self.last_host = r  # counts + digest of the host-buffer call's own result
        return r["result_bytes"]

    def ncu_traffic(self, n_windows):
        """dram bytes (read + write) of one bucket_sort_kernel launch from the
        committed ncu --set full capture, scaled to this run's record count
        when the capture used the same workload; None when there is none."""
        path = os.path.join(ROOT, "profiles", "bucket_sort_traffic.json")
        if os.path.exists(path):
            with open(path) as f:
                t = json.load(f)
            if t.get("n_windows") == n_windows:
                return t["dram_bytes_per_launch"]
        return None


class DistRunner(SingleRunner):
    dominant_kernel = "bucket_sort_kernel"

    def __init__(self, finder, bases, offsets):
        self.f, self.bases, self.offsets = finder, bases, offsets

    def upload(self, k):
        self.f.upload((self.bases, self.offsets), k)

    @staticmethod
    def _job(st):
        st["n_fragments_global"] = st["n_fragments"]
        st["n_blocks_global"] = st["n_blocks"]
        st["n_windows_global"] = st["n_windows"]
        return st

    def step(self, k, similar, max_frags):
        # job-wide block count / digest: one more all-gather, after the timing
        return self.f.find_device(k, similar=similar, max_frags=max_frags,
                                  job_stats=False)

    def finish(self, st):
        return self._job(self.f.job_stats()) if st is self.f._last else self._job(st)

    e2e_call = ("DistAnchorFinder.find_compact_counts (pinned ASCII bases in, every rank's "
                "slice of the blocks + 4-byte (position, strand) fragments out, pinned)")

    def find_host(self, bases, offsets, k, similar, max_frags, full=False):
        fn = self.f.find_counts if full else self.f.find_compact_counts
        st = fn((bases, offsets), k, similar=similar, max_frags=max_frags)
        self.last_host = st
        return st["result_bytes"]

    def ncu_traffic(self, n_windows):
        return None


_RESULT_FD = None


def guard_stdout():
    """Libraries (NCCL's version banner) write to fd 1; the contract is ONE JSON
    line on stdout.  Everything but emit() goes to stderr from here on."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    text = json.dumps(line) + "\n"
    if _RESULT_FD is None:
        sys.stdout.write(text)
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_RESULT_FD, text.encode())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-scale", type=float, default=1.0,
                    help="chromosome length scale of the --impl reference workload "
                         "(1.0 = the GPU arm's workload)")
    ap.add_argument("--cpu-budget", type=float, default=60.0,
                    help="seconds the C1 legs of cpu_baseline may take")
    ap.add_argument("--no-c5", action="store_true",
                    help="skip the 1 Gbp pangenome block")
    ap.add_argument("--no-cpu", action="store_true",
                    help="skip the cpu_baseline leg")
    ap.add_argument("--profile", action="store_true",
                    help="headline steps only (for ncu): no default-mode, "
                         "e2e or cpu legs; prints the per-stage times")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world != args.gpus:
        if args.gpus > 1:
            raise SystemExit("launch --gpus %d with torch.distributed.run "
                             "--nproc-per-node %d" % (args.gpus, args.gpus))
    args.warmup = max(args.warmup, 3)
    run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
