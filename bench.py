#!/usr/bin/env python3
"""bench.py -- headline benchmark of the fold hot path (BASELINE.json metric, config C2).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c5]

A step = one pass of the hot path (fold + backtrack + Hamming/reward) over one batch of synthetic
sequences: C2 = 65,536 random ACGU sequences of n=100 per GPU (weak scaling).  Prints ONE JSON line
on rank 0 (contract in the task statement): `value` = folds/s with inputs resident in HBM, timed
with CUDA events on the launching stream, max over ranks; `e2e` = the same through the C ABI with
HOST buffers (H2D + kernels + D2H inside the timed region); `roofline`, `cpu_baseline`, `clocks`.

`--impl reference` times the CPU implementation of the path (the oracle port of RNA.fold --
ViennaRNA itself is a third-party dependency absent from the reference tree and this image) on all
host threads, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from rnagent_b200 import synth  # noqa: E402

WORKLOADS = {
    # name: (count per GPU, n or None for mixed, seed, parameter set number, description)
    "c2": (65536, 100, 100, 1, "65536 random ACGU sequences, n=100, Turner 2004 (BASELINE.json configs[1])"),
    "c3": (8192, 400, 400, 1, "8192 random ACGU sequences, n=400, Turner 2004 (BASELINE.json configs[2])"),
    "c5": (16384, None, 1999, 2, "16384 random sequences, Eterna100 length mix 12-400, Turner 1999 (configs[4])"),
}
PARAM_NAMES = {1: "turner2004", 2: "turner1999"}


_JSON_OUT = None


def emit_json(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def make_workload(name, rank):
    count, n, seed, setno, desc = WORKLOADS[name]
    if n is None:
        seq, off = synth.random_sequences_mixed(synth.eterna_mixed_lengths(count, seed + rank), seed + 7919 * rank)
    else:
        seq, off = synth.random_sequences(count, n, seed + 7919 * rank)
    tgt = synth.random_targets(off, seed + 1 + 7919 * rank)
    return seq, off, tgt, setno, desc


def total_work(off):
    lens = np.diff(off)
    uniq, cnt = np.unique(lens, return_counts=True)
    return float(sum(synth.work_candidates(int(n)) * int(c) for n, c in zip(uniq, cnt)))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for k, nm in enumerate(names):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# CPU legs: the ONLY place (besides tests/ and smoke()) that executes oracle/.
# ------------------------------------------------------------------------------------------------
def cpu_fold_rate(seq, off, tgt, setno, budget_s, nthreads=0):
    """Oracle port of RNA.fold on the host cores over a bounded sample -> (folds/s, cores, sample B, seconds)."""
    import oracle
    P = oracle.Params(os.path.join(ROOT, "rnagent_b200", "data", f"{PARAM_NAMES[setno]}_37.par"))
    # explicit thread count: torchrun exports OMP_NUM_THREADS=1, which would silently serialise the OpenMP driver
    cores = nthreads or len(os.sched_getaffinity(0)) or 1
    nthreads = cores
    B = len(off) - 1
    probe = min(B, max(2 * cores, 16))
    t0 = time.perf_counter()
    P.fold_batch(seq[:off[probe]], off[:probe + 1], targets_u8=tgt[:off[probe]], nthreads=nthreads)
    dt = max(time.perf_counter() - t0, 1e-4)
    sample = int(min(B, max(probe, budget_s * probe / dt)))
    for _ in range(3):
        t0 = time.perf_counter()
        P.fold_batch(seq[:off[sample]], off[:sample + 1], targets_u8=tgt[:off[sample]], nthreads=nthreads)
        dt = time.perf_counter() - t0
        # the probe includes thread start-up and underestimates the rate: grow the sample until it fills the budget
        if dt >= 0.7 * budget_s or sample >= B:
            break
        sample = int(min(B, sample * min(4.0, budget_s / max(dt, 1e-3))))
    return sample / dt, cores, sample, dt


def cpu_executed_candidates(seq, off, setno, sample=48):
    """W_eff: min-plus candidates the oracle actually evaluates per fold (unpairable cells and inner
    pairs skipped), averaged over the first `sample` sequences -> (executed per fold, algorithmic W per fold)."""
    import oracle
    P = oracle.Params(os.path.join(ROOT, "rnagent_b200", "data", f"{PARAM_NAMES[setno]}_37.par"))
    k = min(sample, len(off) - 1)
    ex = alg = kern = 0
    canpair = np.zeros((256, 256), dtype=bool)
    for a_, b_ in ("AU", "UA", "CG", "GC", "GU", "UG"):
        canpair[ord(a_), ord(b_)] = True
    for b in range(k):
        raw = seq[off[b]:off[b + 1]]
        s_ = raw.tobytes().decode()
        n = len(s_)
        P.fold_dcal(s_)
        ex += P.last_candidates()
        alg += synth.work_candidates(n)
        # what the GPU kernels execute: interior-loop candidates of the cells whose OUTER ends can pair (unpairable inner
        # pairs hold a stored infinity and are still loaded), every multiloop split, hairpin / closing / extension terms
        for d in range(4, n):
            npair = int(canpair[raw[:n - d], raw[d:]].sum())
            kern += npair * synth.interior_candidates(d) + (n - d) * (max(0, d - 8) + 3)
        kern += sum(j - 4 for j in range(5, n + 1))
    k = max(k, 1)
    return ex / k, alg / k, kern / k


def cpu_env_rate(targets, budget_s, n_envs=32, seed=0):
    """C4 on the host: the reference's per-environment logic (RnaDesign.step / Solution.insert_nucleotide /
    find_next_unfilled / get_state, rlif/environments/RLIFenv.py:59-81, rlif/rna/solution.py:84-123 -- the host state
    machine of rnagent_b200.Solution mirrors it line for line) with the terminal evaluation served by the oracle fold.
    One Python process, like one worker of the reference's SubprocVecEnv.  -> (env steps/s, episodes/s, steps)"""
    import oracle
    from rnagent_b200 import DotBracket, Solution
    from rnagent_b200.utils import hamming_distance
    P = oracle.Params(os.path.join(ROOT, "rnagent_b200", "data", "turner2004_37.par"))
    cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True)
    dots = [DotBracket(t, None, k + 1) for k, t in enumerate(targets)]
    rng = np.random.default_rng(seed)
    sols = [Solution(target=dots[int(rng.integers(len(dots)))], config=cfg) for _ in range(n_envs)]
    steps = episodes = 0
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < budget_s:
        acts = rng.integers(0, 4, n_envs)
        for e in range(n_envs):
            sol = sols[e]
            sol.insert_nucleotide(int(acts[e]))
            if sol.index < sol.target.len - 1:
                sol.find_next_unfilled()
            else:
                folded, fe = P.fold(sol.string)
                hd, _ = hamming_distance(sol.target.seq, folded)
                sol.r = (1 - float(hd) / float(sol.target.len)) ** 9 + (.25 if hd == 0 else 0)
                episodes += 1
                sols[e] = sol = Solution(target=dots[int(rng.integers(len(dots)))], config=cfg)
            sol.get_state()
            steps += 1
    dt = time.perf_counter() - t0
    return steps / dt, episodes / dt, steps


def measured_traffic(kernel_key):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the named kernel from this build's
    `ncu --set full` capture, as summarised in profiles/r2_traffic.json (tools/ncu_summary.py) -> (bytes per fold, source)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))[kernel_key]
        return (t["dram_read_bytes"] + t["dram_write_bytes"]) / t["folds"], t.get("source")
    except (OSError, KeyError, ValueError):
        return None, None


def arm_config(desc, folds_per_gpu, world):
    return {"workload": desc, "folds_per_gpu_per_step": int(folds_per_gpu), "l2": "flushed between timed steps (256 MiB memset)",
            "parallelism": f"dp{world} (independent shards, no data-path collective)"}


def run_reference(args):
    """`--impl reference`: CPU path on all host threads, rank 0 only; same metric/unit/config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    seq, off, tgt, setno, desc = make_workload(args.workload, 0)
    n_label = WORKLOADS[args.workload][1] or "mixed"
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    rates, sample, cores = [], 0, 0
    for k in range(args.warmup + args.steps):
        r, cores, sample, dt = cpu_fold_rate(seq, off, tgt, setno, per_step)
        if k >= args.warmup:
            rates.append((sample, dt))
    folds = sum(s for s, _ in rates)
    secs = sum(d for _, d in rates)
    value = folds / secs
    line = {
        "impl": "reference", "metric": f"mfe_folds_per_sec_n{n_label}", "value": value, "unit": "folds/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        # the same `config` as the GPU arm (it names the workload); what this arm folded per step is stated beside it
        "config": arm_config(desc, len(off) - 1, args.gpus),
        "sample_per_step": sample, "us_per_fold_per_thread": 1e6 * cores / value,
        "cpu_baseline": {"value": value, "unit": "folds/s", "cores": cores, "kind": "port",
                         "sample": f"first {sample} sequences of the workload per step, OpenMP over {cores} host threads; "
                                   "oracle port of ViennaRNA 2.4.14 RNA.fold (ViennaRNA itself is not in the "
                                   "reference tree nor installable here)"},
        "e2e": {"value": value, "unit": "folds/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_json(line)
    return 0


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from rnagent_b200 import _lib, vienna

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the fold path has no CPU fallback (use --impl reference for the CPU leg)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib(local)
    seq, off, tgt, setno, desc = make_workload(args.workload, rank)
    vienna.set_vienna_params(setno)
    B, total = len(off) - 1, int(off[-1])
    W = total_work(off)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- resident leg: inputs and outputs live in HBM, kernels on a torch stream we hand to the C ABI
    st = torch.cuda.Stream(device=dev)
    d_seq, d_off, d_tgt = (torch.from_numpy(x).to(dev) for x in (seq, off, tgt))
    d_db = torch.empty_like(d_seq)
    d_mfe = torch.empty(B, dtype=torch.int32, device=dev)
    d_hd = torch.empty(B, dtype=torch.int32, device=dev)
    d_rw = torch.empty(B, dtype=torch.float64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step_resident():
        _lib.fold_hamming_raw(d_seq, d_off, B, target=d_tgt, db=d_db, mfe=d_mfe, hd=d_hd, reward=d_rw,
                              reward_exp=9.0, stream=st.cuda_stream)

    _lib.set_timing(True)
    with torch.cuda.stream(st):
        for _ in range(args.warmup):
            step_resident()
        barrier()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        step_ms, kern_ms, launches = [], [], 0
        for _ in range(args.steps):
            flush.zero_()                       # L2 flush between timed iterations (outside the events)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            step_resident()
            e1.record(st)
            e1.synchronize()
            step_ms.append(e0.elapsed_time(e1))
            kern_ms.append(_lib.last_kernel_ms())
            launches += _lib.last_launch_info()[0]
        barrier()
        clocks = sampler.stop() if rank == 0 else None
    resident_ms = max_over_ranks(sum(step_ms))
    checksum = int(d_mfe.sum().item())

    # ---- end-to-end leg: HOST buffers through the C ABI (pinned host memory, H2D, kernels, D2H in the timed region)
    def pinned(arr):
        t = torch.empty(arr.shape, dtype=torch.from_numpy(arr[:0].copy()).dtype, pin_memory=True)
        t.numpy()[...] = arr
        return t
    p_seq, p_tgt = pinned(seq), pinned(tgt)
    p_db = torch.empty(total, dtype=torch.uint8, pin_memory=True)
    p_mfe = torch.empty(B, dtype=torch.int32, pin_memory=True)
    p_hd = torch.empty(B, dtype=torch.int32, pin_memory=True)
    p_rw = torch.empty(B, dtype=torch.float64, pin_memory=True)
    h_mfe = p_mfe.numpy()

    def step_e2e():
        _lib.fold_hamming_raw(p_seq, off, B, target=p_tgt, db=p_db, mfe=p_mfe, hd=p_hd, reward=p_rw, reward_exp=9.0)

    for _ in range(max(1, min(args.warmup, 2))):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    assert int(h_mfe.astype(np.int64).sum()) == checksum, "host-buffer and device-buffer legs disagree"
    h2d = 2 * total + 4 * (2 * B + 1)            # sequences, targets, offsets + processing order
    d2h = total + 4 * B + 4 * B + 8 * B + 4      # structures, mfe, hd, reward, status

    # ---- the other two figures BASELINE.json's metric names: folds/s at n=400 and RL env steps/s
    extra = {}
    if args.workload == "c2" and not args.no_extra:
        extra = secondary_metrics(torch, dist, _lib, vienna, dev, rank, world, barrier, max_over_ranks,
                                  cpu_budget=0.0 if (world > 1 or args.no_cpu) else max(3.0, args.cpu_seconds / 2))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel (the fold kernel is the only kernel of the step)
    lds_peak, alu_peak = _lib.measure_onchip_peaks()
    k_ms = float(np.mean(kern_ms))
    cand_per_s = W / (k_ms * 1e-3)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_bytes = float(sum(synth.algorithmic_hbm_bytes(int(n)) for n in np.diff(off)))
    per_fold, traffic_src = measured_traffic({"c2": "wave_n100", "c3": "cluster_n400"}.get(args.workload, ""))
    traffic = per_fold * B if per_fold is not None else None
    w_exec = w_alg = w_kern = None
    if world == 1 and not args.no_cpu:
        w_exec, w_alg, w_kern = cpu_executed_candidates(seq, off, setno)
    roofline = {
        "bound": "smem-load/int32-alu (on-chip; neither hbm nor tensor binds this integer min-plus DP)",
        "achieved": cand_per_s / 1e9, "peak": lds_peak / 1e9, "unit": "G min-plus candidates/s (1 shared load each)",
        "frac": cand_per_s / lds_peak, "peak_source": "conflict-free ld.shared.b32 stream measured on this device by bench.py",
        "alu_peak": alu_peak / 1e9, "alu_frac": cand_per_s / alu_peak,
        "work_per_launch": W, "kernel_ms": k_ms, "traffic": traffic, "traffic_source": traffic_src,
        "frac_is": "ALGORITHMIC candidates (every cell, every interior shape) per second / measured shared-load peak; the "
                   "kernel skips unpairable cells, so executed shared loads are fewer -- see w_eff",
        "w_eff": None if w_exec is None else {
            "algorithmic_per_fold": w_alg, "oracle_executed_per_fold": w_exec, "oracle_ratio": w_exec / w_alg,
            "kernel_executed_per_fold": w_kern, "kernel_ratio": w_kern / w_alg,
            "kernel_executed_per_s": cand_per_s * w_kern / w_alg / 1e9, "kernel_executed_frac_of_peak": cand_per_s * w_kern / w_alg / lds_peak,
            "source": "first 48 sequences: the CPU oracle's candidate counter (outer AND inner pair pairable), and the count of "
                      "what the GPU kernels load (interior candidates of cells whose outer ends pair + every multiloop split)"},
        "hbm": {"bound": "hbm", "achieved": hbm_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                "frac": hbm_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak,
                "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6650 GB/s"},
    }

    cpu = None
    if world == 1 and not args.no_cpu:
        r, cores, sample, dt = cpu_fold_rate(seq, off, tgt, setno, args.cpu_seconds)
        cpu = {"value": r, "unit": "folds/s", "cores": cores, "kind": "port",
               "sample": f"first {sample} sequences of the same workload, {dt:.1f} s on {cores} host threads (OpenMP), "
                         f"oracle port of ViennaRNA RNA.fold, {1e6 * cores / r:.0f} us per fold per thread"}

    n_label = WORKLOADS[args.workload][1] or "mixed"
    line = {
        "metric": f"mfe_folds_per_sec_n{n_label}", "value": world * B * args.steps / (resident_ms * 1e-3),
        "unit": "folds/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": resident_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int16 tables / int32 arithmetic (int32 re-run of sequences that leave the int16 range)", "data": "synthetic",
        "config": arm_config(desc, B, world),
        "e2e": {"value": world * B * args.steps / e2e_s, "unit": "folds/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h},
        "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
        "checksum_mfe_dcal": checksum,
    }
    line.update(extra)
    emit_json(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def secondary_metrics(torch, dist, _lib, vienna, dev, rank, world, barrier, max_over_ranks, cpu_budget=0.0):
    """n=400 fold rate (config C3) and RL environment steps/s (config C4), every rank on its own shard."""
    import rnagent_b200
    out = {}
    lds_peak, _alu = _lib.measure_onchip_peaks()

    def roof(cand_per_s_one_gpu, note):
        return {"bound": "smem-load (on-chip)", "achieved": cand_per_s_one_gpu / 1e9, "peak": lds_peak / 1e9,
                "unit": "G min-plus candidates/s (1 shared load each), per GPU", "frac": cand_per_s_one_gpu / lds_peak, "note": note}
    # C3: 8192 x n=400 per GPU, device-resident, 2 timed passes
    seq, off, tgt, setno, desc = make_workload("c3", rank)
    vienna.set_vienna_params(setno)
    B = len(off) - 1
    st = torch.cuda.Stream(device=dev)
    d_seq, d_off, d_tgt = (torch.from_numpy(x).to(dev) for x in (seq, off, tgt))
    d_db = torch.empty_like(d_seq)
    d_mfe = torch.empty(B, dtype=torch.int32, device=dev)
    d_hd = torch.empty(B, dtype=torch.int32, device=dev)
    d_rw = torch.empty(B, dtype=torch.float64, device=dev)

    def run():
        _lib.fold_hamming_raw(d_seq, d_off, B, target=d_tgt, db=d_db, mfe=d_mfe, hd=d_hd, reward=d_rw, reward_exp=9.0,
                              stream=st.cuda_stream)
    with torch.cuda.stream(st):
        run()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        run()
        run()
        e1.record(st)
        e1.synchronize()
    ms = max_over_ranks(e0.elapsed_time(e1))
    W = total_work(off)
    k400 = _lib.last_kernel_ms()
    tr400, tr400_src = measured_traffic("cluster_n400")
    out["n400"] = {"metric": "mfe_folds_per_sec_n400", "value": world * 2 * B / (ms * 1e-3), "unit": "folds/s",
                   "workload": desc, "ms_per_pass": ms / 2, "candidates_per_s": world * 2 * W / (ms * 1e-3),
                   "routed": _lib.last_launch_info()[1], "kernel_ms": k400,
                   "roofline": dict(roof(W / (k400 * 1e-3), "fold_cluster_kernel (2-CTA cluster, DSMEM) + f5_backtrack_kernel + int32 re-run, CUDA events around the launches"),
                                    traffic=None if tr400 is None else tr400 * B, traffic_source=tr400_src)}
    if cpu_budget > 0 and rank == 0:
        r, cores, sample, dt = cpu_fold_rate(seq, off, tgt, setno, cpu_budget)
        out["n400"]["cpu_baseline"] = {"value": r, "unit": "folds/s", "cores": cores, "kind": "port",
                                       "sample": f"first {sample} sequences, {dt:.1f} s on {cores} host threads, {1e3 * cores / r:.1f} ms per fold per thread"}
    # C5: Turner 1999, Eterna100-shaped length mix 12..400 (wave, cluster and re-run kernels in one call)
    seq5, off5, tgt5, setno5, desc5 = make_workload("c5", rank)
    vienna.set_vienna_params(setno5)
    B5 = len(off5) - 1
    d5 = [torch.from_numpy(x).to(dev) for x in (seq5, off5, tgt5)]
    o5 = (torch.empty_like(d5[0]), torch.empty(B5, dtype=torch.int32, device=dev), torch.empty(B5, dtype=torch.int32, device=dev))

    def run5():
        _lib.fold_hamming_raw(d5[0], d5[1], B5, target=d5[2], db=o5[0], mfe=o5[1], hd=o5[2], reward_exp=9.0, stream=st.cuda_stream)
    with torch.cuda.stream(st):
        run5()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        run5()
        e1.record(st)
        e1.synchronize()
    ms5 = max_over_ranks(e0.elapsed_time(e1))
    k5 = _lib.last_kernel_ms()
    out["mixed"] = {"metric": "mfe_folds_per_sec_mixed_12_400", "value": world * B5 / (ms5 * 1e-3), "unit": "folds/s",
                    "workload": desc5, "ms_per_pass": ms5, "candidates_per_s": world * total_work(off5) / (ms5 * 1e-3),
                    "routed": _lib.last_launch_info()[1], "launches": _lib.last_launch_info()[0], "kernel_ms": k5,
                    "roofline": roof(total_work(off5) / (k5 * 1e-3), "all length classes of one call (wave + cluster kernels), CUDA events around the launches")}
    # A8 / N1: batched local search (Solution.permute) over 4096 near misses of n=100, the reference's Testing.yml settings
    vienna.set_vienna_params(1)
    rngp = np.random.default_rng(515 + rank)
    pseq, poff = synth.random_sequences(4096, 100, 5100 + rank)
    base = [pseq[poff[k]:poff[k + 1]].tobytes().decode() for k in range(4096)]
    ptargets, _ = rnagent_b200.fold_batch(base)
    near = []
    for s_ in base:
        s_ = list(s_)
        for _ in range(int(rngp.integers(1, 4))):
            s_[int(rngp.integers(0, 100))] = "ACGU"[int(rngp.integers(0, 4))]
        near.append("".join(s_))
    rnagent_b200.permute_batch(near, ptargets, rnagent_b200.permute.TESTING, seed=1)   # warm-up at full size (scratch growth)
    barrier()
    t0 = time.perf_counter()
    for rep in range(3):
        res = rnagent_b200.permute_batch(near, ptargets, rnagent_b200.permute.TESTING, seed=100 * rank + rep)
    psecs = max_over_ranks(time.perf_counter() - t0) / 3
    searched = int((res.rounds > 0).sum())
    out["permute"] = {"metric": "local_search_solutions_per_sec", "value": world * 4096 / psecs, "unit": "solutions/s",
                      "folds_per_s": world * res.folds / psecs, "folds": res.folds, "searched": searched,
                      "solved_by_search": int(((res.hd == 0) & (res.rounds > 0)).sum()),
                      "workload": "4096 near misses (1-3 point mutations of a solving sequence) of n=100 per GPU, "
                                  "permutation_budget 15, radius 5, threshold 15, GU allowed (rlif/config/Testing.yml:3-8); "
                                  "host strings in, host strings out",
                      "roofline": roof(res.folds * synth.work_candidates(100) / psecs, "one batched fold per search round; wall clock incl. host strings")}
    # C4: 4096 environments per GPU, n=100 targets (folds of random sequences), random actions, auto-reset
    vienna.set_vienna_params(1)
    sseq, soff = synth.random_sequences(256, 100, 4100 + rank)
    targets, _ = rnagent_b200.fold_batch([sseq[soff[k]:soff[k + 1]].tobytes().decode() for k in range(256)])
    cfg = dict(kernel_size=25, reward_exp=9, permute=False, encoding_type=2, use_nucleotides=True)
    # `settle` untimed steps first: environments that share a target finish in cohorts until their episode phases have drifted
    # apart (tools/env_terminal_hist.py: 15-22 % of the steps carry more than 148 terminal folds after 300 steps, none after
    # 1,500) -- the steady state of a long rollout is what is timed
    N, steps, settle = 4096, 400, 1500
    env = rnagent_b200.RnaDesignBatch(cfg, N, targets=targets, rank=rank, auto_reset=True, randomize=True, seed=rank)
    g = torch.Generator(device=dev)
    g.manual_seed(rank)
    acts = torch.randint(0, 4, (steps + settle, N), dtype=torch.int32, device=dev, generator=g)
    d_obs = torch.empty((N, 10, 50), dtype=torch.uint8, device=dev)
    d_r = torch.empty(N, dtype=torch.float64, device=dev)
    d_done = torch.empty(N, dtype=torch.uint8, device=dev)
    env.reset()
    for k in range(settle):
        env.step_device(acts[k], d_obs, d_r, d_done)
    env.stats(clear=True)
    barrier()
    t0 = time.perf_counter()
    for k in range(steps):
        env.step_device(acts[settle + k], d_obs, d_r, d_done)
    torch.cuda.synchronize()
    my_secs = time.perf_counter() - t0
    secs = max_over_ranks(my_secs)
    rank_us = [1e6 * my_secs / steps]
    if world > 1:                                        # per-rank step times: is the maximum one slow rank or all of them?
        g_ = [torch.zeros(1, dtype=torch.float64, device=dev) for _ in range(world)]
        dist.all_gather(g_, torch.tensor([rank_us[0]], dtype=torch.float64, device=dev))
        rank_us = [float(t.item()) for t in g_]
    stats = env.stats().all_reduce(device=dev)          # the path's only collective: rollout statistics over NCCL
    # the same steps as a ROLLOUT that keeps one reward row per step (what PPO stores): rewards deferred by up to 32 steps, the
    # terminal sequences of a window folded by one call (values identical, tests/test_env_gpu.py)
    d_rT = torch.empty((steps, N), dtype=torch.float64, device=dev)
    env.set_reward_delay(32)
    for k in range(64):
        env.step_device(acts[k], d_obs, d_rT[k], d_done)
    env.flush()
    env.stats(clear=True)
    barrier()
    t0 = time.perf_counter()
    for k in range(steps):
        env.step_device(acts[settle + k], d_obs, d_rT[k], d_done)
    env.flush()
    torch.cuda.synchronize()
    dsecs = max_over_ranks(time.perf_counter() - t0)
    dstats = env.stats().all_reduce(device=dev)
    env.set_reward_delay(1)
    env.close()
    out["env"] = {"metric": "rl_env_steps_per_sec", "value": stats.steps / secs, "unit": "env steps/s",
                  "per_gpu": stats.steps / secs / world, "us_per_vector_step": 1e6 * secs / steps,
                  "episodes_per_s": stats.episodes / secs, "envs_per_gpu": N, "steps_timed": steps,
                  "host_syncs_per_step": 0, "per_rank_us_per_vector_step": [round(v, 1) for v in rank_us],
                  "rewards": "step-synchronous: every step's rewards are complete when the step is (one fold latency per step); "
                             "rollout_deferred is the same loop as a rollout whose rewards may arrive up to 32 steps late",
                  "sum_of_rank_rates": sum(N * 1e6 / v for v in rank_us),
                  "rollout_deferred": {"value": dstats.steps / dsecs, "unit": "env steps/s", "per_gpu": dstats.steps / dsecs / world,
                                       "us_per_vector_step": 1e6 * dsecs / steps, "reward_delay_steps": 32,
                                       "episodes_per_s": dstats.episodes / dsecs,
                                       "note": "rnafold_env_config_delay(32): one reward row per step ([T, N] rollout buffer), the terminal folds of 32 steps "
                                               "in one call; rewards complete at the end of the timed region (20.4 / 82 / 119 / 127 M steps/s at "
                                               "delay 1 / 8 / 32 / 64, tools/env_speed.py)"},
                  "roofline": roof(stats.episodes / secs / world * synth.work_candidates(100),
                                   "terminal folds only (one per episode); the step is bound by ONE fold's latency, not by throughput"),
                  "workload": "4096 RnaDesign envs per GPU, n=100 targets (folds of random sequences), uniform random actions, "
                              "auto-reset, steady state (1,500 untimed steps first); observations, rewards and dones stay on the device",
                  "stats": stats.summary()}
    if cpu_budget > 0 and rank == 0:
        r_steps, r_eps, n_steps = cpu_env_rate(targets, cpu_budget)
        out["env"]["cpu_baseline"] = {"value": r_steps, "unit": "env steps/s", "cores": 1, "kind": "port", "episodes_per_s": r_eps,
                                      "sample": f"{n_steps} steps of 32 environments in one Python process (one SubprocVecEnv worker of the "
                                                "reference): host state machine of Solution + oracle fold at the terminal step"}
    # second C4 point: the C5 length mix as targets
    lens = synth.eterna_mixed_lengths(128, 77 + rank)
    mseq, moff = synth.random_sequences_mixed(lens, 4200 + rank)
    mtargets, _ = rnagent_b200.fold_batch([mseq[moff[k]:moff[k + 1]].tobytes().decode() for k in range(128)])
    env = rnagent_b200.RnaDesignBatch(cfg, N, targets=mtargets, rank=rank, auto_reset=True, randomize=True, seed=rank)
    msteps, msettle = 120, 400
    macts = torch.randint(0, 4, (msteps + msettle, N), dtype=torch.int32, device=dev, generator=g)
    env.reset()
    for k in range(msettle):
        env.step_device(macts[k], d_obs, d_r, d_done)
    env.stats(clear=True)
    barrier()
    t0 = time.perf_counter()
    for k in range(msteps):
        env.step_device(macts[msettle + k], d_obs, d_r, d_done)
    torch.cuda.synchronize()
    msecs = max_over_ranks(time.perf_counter() - t0)
    mstats = env.stats().all_reduce(device=dev)
    d_rT = torch.empty((msteps, N), dtype=torch.float64, device=dev)
    env.set_reward_delay(32)                             # capped by eight times the shortest episode of the data set
    env.stats(clear=True)
    barrier()
    t0 = time.perf_counter()
    for k in range(msteps):
        env.step_device(macts[msettle + k], d_obs, d_rT[k], d_done)
    env.flush()
    torch.cuda.synchronize()
    mdsecs = max_over_ranks(time.perf_counter() - t0)
    mdstats = env.stats().all_reduce(device=dev)
    env.close()
    out["env_mixed"] = {"metric": "rl_env_steps_per_sec_mixed_targets", "value": mstats.steps / msecs, "unit": "env steps/s",
                        "per_gpu": mstats.steps / msecs / world, "us_per_vector_step": 1e6 * msecs / msteps,
                        "episodes_per_s": mstats.episodes / msecs, "envs_per_gpu": N, "steps_timed": msteps,
                        "rollout_deferred": {"value": mdstats.steps / mdsecs, "unit": "env steps/s", "us_per_vector_step": 1e6 * mdsecs / msteps,
                                             "reward_delay_steps": "32 requested, capped by 8 x the shortest episode"},
                        "workload": "4096 RnaDesign envs per GPU, targets = folds of 128 random sequences with the Eterna100 length mix "
                                    "(12-400 nt), uniform random actions, auto-reset, 400 untimed steps first"}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="c2")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU-baseline sample budget")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the n=400 and env-steps measurements")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON: whatever libraries print there (NCCL's version banner under NCCL_DEBUG=VERSION,
    # for one) goes to stderr instead
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
