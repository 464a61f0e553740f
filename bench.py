#!/usr/bin/env python
"""bench.py -- AnchorFinder throughput on synthetic genomes (BASELINE.json).

A step = one pass of the hot path (AnchorFinder::run_impl,
src/algo/AnchorFinder.cpp:393-406) over the whole synthetic genome set:
  value   bases scanned per second with the packed genome resident in HBM and
          the anchors left in HBM (npge_anchor_find_device), timed with CUDA
          events on the library's own stream, L2 flushed between steps
  e2e     the same metric through the drop-in call npge_anchor_find with HOST
          buffers: pinned ASCII bases in, Fragment arrays out, copies included
  --impl reference   the reference's CPU algorithm (oracle/ restatement; the
          reference itself cannot be built here) on the host cores

Workload: BASELINE.json configs[1], 17 synthetic Brucella-shaped genomes
(~56 Mbp), anchor-size 20, parity mode (anchor-similar=false, no truncation);
the default options are timed beside it (`default_mode`).  At N>1 every rank
holds its own 17 strains of the same ancestor (weak scaling), k-mers are
repartitioned by hash range with one NCCL all-to-all.
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


def make_workload(rank, world, scale=1.0):
    """Rank r's strains: the same ancestor for every rank (same seed), its own
    17 strains.  Returns (names, seqs, offset of this rank's first caller
    sequence index)."""
    from npge_b200 import synth
    if world == 1:
        names, seqs = synth.generate(WORKLOAD, scale=scale)
        return names, seqs
    return synth.generate_shard(WORKLOAD, rank, world, scale=scale)


# ---------------------------------------------------------------------------
def run_reference(args, rank, world):
    """The reference's CPU algorithm (two-pass Bloom AnchorFinder restated in
    oracle/anchor_oracle.cpp) on the host cores, all threads, on a bounded
    sample of the same workload."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    from npge_b200 import synth
    cores = os.cpu_count() or 1
    # bound the whole run to ~150 s of CPU work: probe the rate on a 1 % sample
    names, seqs = synth.generate(WORKLOAD, scale=0.01)
    t = time.perf_counter()
    oracle.find_faithful(seqs, K, similar=False, max_frags=BIG, workers=cores,
                         genome_ids=synth.genome_ids(names), asserts=True)
    rate = synth.total_bases(seqs) / (time.perf_counter() - t)
    full = 56.1e6
    scale = min(args.cpu_scale,
                rate * 150.0 / (args.steps + args.warmup) / full)
    scale = max(scale, 0.005)
    names, seqs = synth.generate(WORKLOAD, scale=scale)
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
    sample = ("%s at scale %.3g: %d sequences, %.2f Mbp, k=%d, parity mode; "
              "two-pass Bloom restatement (oracle find_faithful, range "
              "checks on), %d std::threads, one task per sequence"
              % (WORKLOAD, scale, len(seqs), n / 1e6, K, cores))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "anchor_size": K,
                   "mode": "parity (anchor-similar=false, no truncation)",
                   "sample_scale": scale, "bases_per_step": n},
        "anchors_per_s": frags * len(times) / total,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores,
                         "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ---------------------------------------------------------------------------
def cpu_baseline(scale):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle
    from npge_b200 import synth
    names, seqs = synth.generate(WORKLOAD, scale=scale)
    n = synth.total_bases(seqs)
    gid = synth.genome_ids(names)
    cores = os.cpu_count() or 1
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
        "sample": ("%s at scale %.3g (%d sequences, %.2f Mbp), k=%d, parity "
                   "mode, oracle find_faithful (two-pass Bloom restatement of "
                   "AnchorFinder.cpp, range checks on), %d threads, %.1f s"
                   % (WORKLOAD, scale, len(seqs), n / 1e6, K, cores, dt)),
        "anchors_per_s": r.n_fragments / dt,
        "default_mode_value": n / dt_default / 1e9,
    }


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
        bases, offsets = concat(seqs)
        eng = AnchorEngine(local_rank)
        runner = SingleRunner(eng, bases, offsets)
    else:
        from npge_b200.dist import DistAnchorFinder
        names, seqs = synth.generate_shard(WORKLOAD, rank, world)
        bases, offsets = concat(seqs)
        runner = DistRunner(DistAnchorFinder(local_rank), bases, offsets)
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
    stats = None
    for _ in range(args.warmup):
        stats = runner.step(K, False, BIG)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    ev_ms, pass_ms, kern_launches = 0.0, [], 0
    hist_ms = group_ms = step_ms = 0.0
    exch_ms, host_ms = {}, {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    wall0 = time.perf_counter()
    for _ in range(args.steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        e0.record()
        stats = runner.step(K, False, BIG)
        e1.record()
        e1.synchronize()
        step_ms += e0.elapsed_time(e1)
        for nm, v in stats.get("exchange_ms", {}).items():
            exch_ms[nm] = exch_ms.get(nm, 0.0) + v
        for nm, v in stats.get("host_ms", {}).items():
            host_ms[nm] = host_ms.get(nm, 0.0) + v
        ev_ms += stats["ms_total"]
        hist_ms += stats["ms_hist"]
        group_ms += stats["ms_group"]
        pass_ms.append(stats["ms_pass"])
        kern_launches += stats["gpu_launches"]
    barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.stop() if rank == 0 else None
    # one GPU: the library's own events on its stream (first kernel .. last
    # kernel of every step); several GPUs: events around the whole step, i.e.
    # kernels + collectives + the host tables between them
    t_dev = (ev_ms if world == 1 else step_ms) / 1e3
    if world > 1:
        t = torch.tensor([t_dev], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_dev = float(t.item())
    parity_stats = runner.finish(stats)
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
                  "stats": {k: v for k, v in stats.items() if k != "digest"}})
        return

    # ---- default options (anchor-similar=true, 100000 fragments) -----------
    for _ in range(2):
        runner.step(K, True, 100000)
    dflt_ms = 0.0
    dsteps = max(3, args.steps // 2)
    for _ in range(dsteps):
        flush.fill_(1)
        torch.cuda.synchronize()
        e0.record()
        st = runner.step(K, True, 100000)
        e1.record()
        e1.synchronize()
        dflt_ms += st["ms_total"] if world == 1 else e0.elapsed_time(e1)
        kern_launches += 0  # not part of the timed headline region
    if world > 1:
        t = torch.tensor([dflt_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dflt_ms = float(t.item())
    default_stats = runner.finish(st)

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
    e2e_t = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_t], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_t = float(t.item())
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

    if rank != 0:
        return
    value = n_total * args.steps / t_dev / 1e9
    nw = parity_stats.get("local_records", parity_stats["n_windows"])  # rank 0's
    passes = parity_stats["radix_passes"]
    # stages of one step (CUDA events inside npge_anchor_find_device):
    # count A + scan | pass A | count B + scan | pass B | bucket sort | scan + emit
    names = ["pass_a", "count_b_scan", "pass_b", "bucket_sort", "scan_emit"]
    stage = {"count_a_scan": hist_ms / args.steps}
    for i, nm in enumerate(names):
        stage[nm] = float(np.mean([p[i] for p in pass_ms]))
    if world > 1:  # rank 0's view of the step
        for nm, v in exch_ms.items():
            stage["exchange_" + nm] = v / args.steps
        stage["kernels"] = ev_ms / args.steps
        stage["step"] = step_ms / args.steps
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
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t_dev / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic",
        "config": {
            "workload": WORKLOAD + (" x%d (one 17-strain cluster per GPU, "
                                    "independent ancestors)" % world
                                    if world > 1 else ""),
            "anchor_size": K,
            "mode": "parity (anchor-similar=false, no truncation)",
            "bases_per_step": n_total, "sequences": int(len(offsets) - 1) * world,
            "l2": "512 MiB flush buffer rewritten between timed steps",
            "timed": ("library CUDA events on its stream, first to last kernel "
                      "of the step" if world == 1 else
                      "CUDA events around the whole step (kernels, NCCL "
                      "exchanges, host tables), max over ranks"),
            "parallelism": ("1 GPU" if world == 1 else
                            "sequences sharded per GPU; k-mers repartitioned by hash "
                            "range x%d, exchange=%s (peer: pass A scatters into the "
                            "owners' HBM over NVLink; bulk: copy kernel into peer "
                            "memory; collective: NCCL all-to-all)"
                            % (world, parity_stats.get("exchange", "?"))),
        },
        "anchors_per_s": parity_stats["n_fragments_global"] * args.steps / t_dev,
        "fragments": parity_stats["n_fragments_global"],
        "blocks": parity_stats["n_blocks_global"],
        "windows": parity_stats["n_windows_global"],
        "wall_ms_per_step": 1e3 * wall / args.steps,
        "stage_ms": stage,
        "default_mode": {
            "value": n_total * dsteps / (dflt_ms * 1e-3) / 1e9, "unit": UNIT,
            "options": "anchor-similar=true, max-anchor-fragments=100000",
            "fragments": default_stats["n_fragments_global"],
            "blocks": default_stats["n_blocks_global"]},
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
                         "compulsory_frac": moved_gbs / hbm_peak}},
        "e2e": {"value": n_total * e2e_steps / e2e_t / 1e9, "unit": UNIT,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(d2h),
                "steps": e2e_steps,
                "call": "npge_anchor_find (pinned ASCII bases in, fragment "
                        "arrays out, pinned)",
                "default_mode": {
                    "value": n_total * e2e_steps / e2e_default_t / 1e9,
                    "d2h_bytes_per_step": int(d2h_default)}},
        "gpu_launches": int(kern_launches),
        "clocks": clocks,
    }
    if block_hash is not None:
        line["block_hash"] = block_hash
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(args.cpu_scale)
    emit(line)


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
        return st

    def find_host(self, bases, offsets, k, similar, max_frags):
        r = self.eng.find_counts((bases, offsets), k, similar=similar,
                                 max_frags=max_frags)
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

    def find_host(self, bases, offsets, k, similar, max_frags):
        st = self.f.find_counts((bases, offsets), k, similar=similar,
                                max_frags=max_frags)
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
    ap.add_argument("--cpu-scale", type=float, default=0.2,
                    help="chromosome length scale of the bounded CPU sample")
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
