#!/usr/bin/env python
"""bench.py -- pod-node score evals/sec (and reconcile decisions/sec) on 1/2/4/8 B200.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload c3|c4|c5]

One "step" = one whole placement of the workload's synthetic snapshot (every round: filter,
score, admit) starting from tables already resident in HBM.  Workloads (BASELINE.json configs,
instantiated as SURVEY.md 8(d)):  c3 = 10k nodes x 100k pods (N=1 default),
c4 = 100k nodes x 1M pods sharded over the ranks (N>1 default), c2 = 1k pools x 100k VMs
(reported beside the headline as `reconcile`).  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

EMIT = None  # set in main(): prints the one JSON line on the real stdout
NODE_REC_BYTES = 32  # algorithmic bytes per pod-node eval: one packed node record
WORKLOADS = {
    "c3": dict(n=10_000, m=100_000, config_id=3, desc="config3: 10k nodes x 100k pods"),
    "c4": dict(n=100_000, m=1_000_000, config_id=4, desc="config4: 100k nodes x 1M pods"),
    "c5": dict(n=10_000, m=1_000_000, config_id=5, desc="config5: 10k nodes x 1M pods (sweep top)"),
    "tiny": dict(n=1_000, m=5_000, config_id=3, desc="tiny: 1k nodes x 5k pods (dev only)"),
}
C2 = dict(p=1_000, v=100_000)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """DRAM bytes per scoring-kernel launch from the committed ncu --set full capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "score_kernel_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return None


class ClockSampler:
    """Samples SM clocks / throttle reasons during the timed region (NVML, else nvidia-smi)."""

    def __init__(self, index=0, period=0.1):
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        self._nvml = None

    def start(self):
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nvml = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._nvml = None
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()
        return self

    def _run(self):
        names = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                 "hw_power_brake_slowdown": 0x80, "sw_power_cap": 0x4}
        while not self._stop.is_set():
            try:
                if self._nvml:
                    n = self._nvml
                    self.samples.append(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM))
                    r = n.nvmlDeviceGetCurrentClocksEventReasons(self._h) if hasattr(
                        n, "nvmlDeviceGetCurrentClocksEventReasons") else n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                    for k, bit in names.items():
                        if r & bit:
                            self.reasons.add(k)
                else:
                    out = subprocess.run(
                        ["nvidia-smi", f"--id={self.index}", "--query-gpu=clocks.sm,clocks.max.sm,"
                         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap",
                         "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout.strip()
                    f = [x.strip() for x in out.split(",")]
                    self.samples.append(int(f[0]))
                    self.max_mhz = int(f[1])
                    for k, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:]):
                        if v.lower().startswith("active"):
                            self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(self.period)

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def gen_workload(name):
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import synth

    w = WORKLOADS[name]
    return synth.gen_nodes(w["n"], w["config_id"]), synth.gen_pods(w["m"], w["config_id"])


# The CPU arm times a BOUNDED SAMPLE of the workload: every round of the first `frac` of the pods
# against the whole node table.  One rule for both the --impl reference arm and the in-line
# cpu_baseline (so that one config has one CPU number), sized for >= ~2 s of work on 32 host
# threads, so that handing the chunks to the oracle's thread pool is noise.
CPU_SAMPLE_FRAC = {"c3": 0.5, "c4": 0.02, "c5": 0.05, "tiny": 1.0}


def cpu_port(nodes, pods, frac, threads):
    """The CPU oracle (a port: the reference has no runnable code) on the first `frac` of the pods,
    every round, all host threads.  Returns (evals_per_s, evals, seconds, sample text)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as orc

    m = max(1, int(len(pods["req_gpu"]) * frac))
    sub = {k: np.ascontiguousarray(v[:m]) for k, v in pods.items()}
    t0 = time.perf_counter()
    r = orc.place(nodes, sub, nthreads=threads)
    dt = time.perf_counter() - t0
    return r["evals"] / dt, r["evals"], dt, f"all {r['rounds']} rounds of the first {m} pods x {len(nodes['flags'])} nodes"


def reference_arm(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path.  The reference is
    Markdown only (no code to run), so this times the oracle port on the host cores."""
    if rank != 0:
        return
    wl = args.workload or ("c3" if args.gpus == 1 else "c4")
    nodes, pods = gen_workload(wl)
    threads = os.cpu_count() or 1
    frac = CPU_SAMPLE_FRAC[wl]
    for _ in range(min(args.warmup, 2)):  # the thread pool exists and the tables are in cache after one
        cpu_port(nodes, pods, frac, threads)
    ev, secs, sample = 0, 0.0, ""
    for _ in range(args.steps):
        _, e, dt, sample = cpu_port(nodes, pods, frac, threads)
        ev += e
        secs += dt
    v = ev / secs
    line = {
        "impl": "reference", "metric": "pod_node_score_evals_per_sec", "value": v, "unit": "evals/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u32",
        "data": "synthetic",
        "config": {"workload": WORKLOADS[wl]["desc"], "nodes": WORKLOADS[wl]["n"], "pods": WORKLOADS[wl]["m"],
                   "note": "reference is Markdown only; CPU oracle port timed on host cores"},
        "cpu_baseline": {"value": v, "unit": "evals/s", "cores": threads, "kind": "port", "sample": sample,
                         "sample_frac": frac, "seconds_per_step": secs / args.steps},
        "e2e": {"value": v, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    EMIT(line)


_PINNED = []  # page-locked blocks stay alive until the process ends


def pinned_block(torch, n_elems):
    """uint32 array in page-locked host memory from the library's own allocator (k8sgpu_host_alloc)."""
    from k8sgpu_b200 import capi

    blk = capi.PinnedBlock(max(int(n_elems), 1))
    _PINNED.append(blk)
    return blk.array


def slab(torch, cols, order):
    """One pinned block with the columns back to back (the layout the device keeps): the library
    then moves a table with one copy instead of eight."""
    n = len(cols[order[0]])
    blk = pinned_block(torch, len(order) * n)
    out = {}
    for c, k in enumerate(order):
        out[k] = blk[c * n:(c + 1) * n]
        out[k][:] = cols[k]
    return out


def time_resident(ctx, steps, flush, barrier, dist, torch):
    """`steps` whole placements from tables resident in HBM; device time per step (CUDA events
    recorded by the library on the stream its kernels run on), max over ranks."""
    tot = dict(ms=0.0, evals=0, score_ms=0.0, score_evals=0, score_launches=0, launches=0)
    st = None
    for _ in range(steps):
        flush.zero_()  # evict the tables from L2 between timed steps (untimed)
        barrier()
        st = ctx.place_run()
        ms = st["device_ms"]
        if dist is not None:  # max over ranks, measured on the device
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        tot["ms"] += ms
        tot["evals"] += st["evals"]
        tot["score_ms"] += st["score_ms"]
        tot["score_evals"] += st["score_evals"]
        tot["score_launches"] += st["score_launches"]
        tot["launches"] += st["kernel_launches"]
    return tot, st


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-reconcile", action="store_true")
    ap.add_argument("--no-scaling-base", action="store_true")
    ap.add_argument("--exchange", default=None, choices=["allgather", "peer"],
                    help="N>1: force how a round's proposals travel (default: peer stores where peers can be mapped)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE line, the JSON: anything a library prints there meanwhile (NCCL's
    # version banner under NCCL_DEBUG) is sent to stderr instead
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)

    global EMIT
    EMIT = emit
    if args.impl == "reference":
        reference_arm(args, rank, world)
        return
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run "
                         f"--nproc-per-node {args.gpus}")
    args.warmup = max(args.warmup, 3)

    import torch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU path for --impl ours")
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import capi
    from k8sgpu_b200.digest import place_digest

    torch.cuda.set_device(local_rank)
    dist = None
    ctx = capi.Context(local_rank)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        # torch.distributed is the plumbing (id broadcast, barriers, max over ranks); the placement's
        # own collectives run on the library's NCCL communicator, inside k8sgpu_place_run.
        box = [capi.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        ctx.comm_init_rank(rank, world, box[0])
        if args.exchange:
            ctx.set_tuning(exchange=1 if args.exchange == "peer" else 0)
    wl = args.workload or ("c3" if world == 1 else "c4")
    nodes, pods = gen_workload(wl)
    N, M = len(nodes["flags"]), len(pods["req_gpu"])

    ctx.place_upload(nodes, pods)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        ctx.place_run()
    sampler = ClockSampler(local_rank, period=0.01).start() if rank == 0 else None
    barrier()
    wall0 = time.perf_counter()
    tot, st = time_resident(ctx, args.steps, flush, barrier, dist, torch)
    barrier()
    wall = time.perf_counter() - wall0
    clocks = sampler.stop() if sampler else None
    rounds, placed = st["rounds"], st["placed"]
    result = ctx.place_download()  # untimed: what was computed, as a digest (equal at every N)
    sha = place_digest(result)

    # end to end through the drop-in call: pinned host tables -> H2D -> rounds -> D2H of results,
    # one k8sgpu_place per step (collective across the ranks at N > 1)
    hn = slab(torch, nodes, ("free_gpu", "free_mcpu", "free_mib", "cap_gpu", "sku_id", "zone_id", "taint_mask", "flags"))
    hp = slab(torch, pods, ("req_gpu", "req_mcpu", "req_mib", "sku_mask", "zone_mask", "tol_mask", "pref_zone", "priority"))
    ho = slab(torch, {k: np.zeros(N, np.uint32) for k in ("free_gpu", "free_mcpu", "free_mib")},
              ("free_gpu", "free_mcpu", "free_mib"))
    ho["assign"] = pinned_block(torch, M).view(np.int32)[:M]
    ctx.place(hn, hp, out=ho)
    barrier()
    t0 = time.perf_counter()
    e_evals = 0
    for _ in range(args.steps):
        r = ctx.place(hn, hp, out=ho)
        e_evals += r["evals"]
    e_dt = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([e_dt], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e_dt = float(t.item())
    e2e = {"value": e_evals / e_dt, "unit": "evals/s", "h2d_bytes_per_step": int(r["h2d_bytes"]) * world,
           "d2h_bytes_per_step": int(r["d2h_bytes"]) * world, "ms_per_step": 1e3 * e_dt / args.steps,
           "result_sha256": place_digest(ho)}
    if world > 1:
        e2e["note"] = "every rank uploads the replicated tables and downloads the full result"
    launches_e2e = int(r["kernel_launches"])

    # N > 1 runs config 4; the like-for-like base of its scaling efficiency is config 4 on ONE GPU:
    # rank 0 times it on its own GPU after the sharded run (the other ranks wait at the barrier).
    same_base = None
    if world > 1 and not args.no_scaling_base:
        if rank == 0:
            with capi.Context(local_rank) as solo:
                solo.place_upload(nodes, pods)
                solo.place_run()
                b_tot, b_st = time_resident(solo, max(args.steps, 5), flush, torch.cuda.synchronize, None, torch)
                same_base = {"workload": WORKLOADS[wl]["desc"], "n_gpus": 1, "value": b_tot["evals"] / (b_tot["ms"] * 1e-3),
                             "unit": "evals/s", "ms_per_step": b_tot["ms"] / max(args.steps, 5), "steps": max(args.steps, 5),
                             "rounds": b_st["rounds"], "result_sha256": place_digest(solo.place_download())}
        barrier()

    if rank != 0:
        ctx.close()
        if dist is not None:
            dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    tot_ms, tot_evals, tot_score_ms, tot_score_evals = tot["ms"], tot["evals"], tot["score_ms"], tot["score_evals"]
    score_launches, launches = tot["score_launches"], tot["launches"]
    contract_gbs = tot_score_evals * NODE_REC_BYTES / (tot_score_ms * 1e-3) / 1e9 if tot_score_ms > 0 else 0.0
    traffic = ncu_traffic()
    ipe = float((traffic or {}).get("warp_instructions_per_warp_eval", 15.67))  # ncu: inst_executed / warp-evals
    if traffic and traffic.get("workload", "c3") != wl:
        traffic = None  # the committed ncu capture is of another workload's launch
    # What binds the scoring kernel is instruction issue (one warp instruction per clock per SM
    # sub-partition, 592 of them), not HBM: node records are re-used from shared memory by every
    # pod of a CTA.  Ceiling = issue rate / the instructions ncu counted per pod-node eval.
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965
    issue_peak = 148 * 4 * 32 * sm_mhz * 1e6 / ipe
    score_rate = tot_score_evals / (tot_score_ms * 1e-3) if tot_score_ms > 0 else 0.0
    line = {
        "metric": "pod_node_score_evals_per_sec", "value": tot_evals / (tot_ms * 1e-3), "unit": "evals/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_ms / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,  # total work per workload is fixed
        "dtype": "u32", "data": "synthetic",
        "config": {"workload": WORKLOADS[wl]["desc"], "nodes": N, "pods": M, "rounds": rounds, "placed": placed,
                   "evals_per_step": tot_evals // args.steps, "result_sha256": sha,
                   "parallelism": f"pods sharded x{world}, nodes replicated",
                   "exchange": "none" if world == 1 else {"peer": "peer stores + 1-word ncclAllReduce (forced)",
                                                          "allgather": "ncclAllGather of the proposals (forced)",
                                                          None: "in the library: peer stores from the score-decoding kernel + "
                                                                "1-word ncclAllReduce (ncclAllGather where peers cannot be mapped)"}[args.exchange],
                   "l2": "flushed between timed steps (256 MB write)", "wall_s_timed_region": wall,
                   "launches_per_e2e_step": launches_e2e},
        "roofline": {"bound": "instruction issue (INT32), not hbm: see roofline_contract", "achieved": score_rate,
                     "peak": issue_peak, "unit": "evals/s", "frac": score_rate / issue_peak,
                     "traffic": traffic["dram_bytes_per_launch"] if traffic else None,
                     "kernel": "score_kernel", "instructions_per_eval": ipe,
                     "peak_source": "592 SM sub-partitions x 32 lanes x SM clock under load / warp instructions per "
                                    "warp-eval measured by ncu (profiles/score_kernel_traffic.json)",
                     "evals_per_launch": tot_score_evals // max(score_launches, 1),
                     "avg_launch_ms": tot_score_ms / max(score_launches, 1),
                     "share_of_step": tot_score_ms / tot_ms if tot_ms else None},
        "roofline_contract": {"bound": "hbm", "achieved": contract_gbs, "peak": peak, "unit": "GB/s",
                              "ratio": contract_gbs / peak, "algorithmic_bytes_per_eval": NODE_REC_BYTES,
                              "peak_source": peak_src,
                              "note": "SURVEY 8(d) contract figure: 32 B (one node record) per eval against the HBM copy "
                                      "peak.  The records come from shared memory / L2, not DRAM (compare `traffic`), "
                                      "so this ratio exceeds 1 and is NOT a roofline fraction"},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
    }
    if world == 1 and not args.no_scaling_base:
        # Disclosure: the same workload with the scoring kernel made to evaluate the CPU and memory
        # differences of every pair (by default they are left out for
        # pods whose request the round's node table proves redundant -- every pod, with these
        # synthetic tables).  Same result bytes; not the headline.
        ctx.set_tuning(force_cls=18)
        ctx.place_run()
        a_tot, _ = time_resident(ctx, args.steps, flush, barrier, dist, torch)
        ctx.reset_tuning()
        line["all_resource_clauses"] = {"value": a_tot["evals"] / (a_tot["ms"] * 1e-3), "unit": "evals/s",
                                        "ms_per_step": a_tot["ms"] / args.steps,
                                        "result_sha256": place_digest(ctx.place_download()),
                                        "note": "force_cls=18: no CPU / memory difference left out of any eval"}
    if same_base is not None:
        eff = line["value"] / (world * same_base["value"])
        e_eff = e2e["value"] / (world * same_base["value"])
        line["scaling_base"] = dict(same_base, same_workload_efficiency=eff, same_workload_efficiency_e2e=e_eff,
                                    digest_equal=same_base["result_sha256"] == sha)
    if world == 1 and wl == "c3" and not args.no_scaling_base:
        # N>1 runs config 4 (BASELINE.json), so the like-for-like 1-GPU base of the scaling
        # series is config 4 on one GPU: measure it here too.
        n4, p4 = gen_workload("c4")
        ctx.place_upload(n4, p4)
        ctx.place_run()
        b_tot, st4 = time_resident(ctx, 5, flush, torch.cuda.synchronize, None, torch)
        line["scaling_base"] = {"workload": WORKLOADS["c4"]["desc"], "n_gpus": 1, "value": b_tot["evals"] / (b_tot["ms"] * 1e-3),
                                "unit": "evals/s", "ms_per_step": b_tot["ms"] / 5, "steps": 5, "rounds": st4["rounds"],
                                "result_sha256": place_digest(ctx.place_download()),
                                "note": "divide the --gpus N>1 values by N times this for same-workload scaling efficiency"}

    if not args.no_reconcile and world == 1:
        from k8sgpu_b200 import synth
        from k8sgpu_b200.digest import reconcile_digest

        t = synth.gen_reconcile(C2["p"], C2["v"])
        P, V = C2["p"], C2["v"]
        rargs = (t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"], t["vm_ordinal"])
        ctx.reconcile_upload(*rargs)
        for _ in range(args.warmup):
            ctx.reconcile_run()
        rms = 0.0
        for _ in range(args.steps):
            flush.zero_()
            torch.cuda.synchronize()
            rms += ctx.reconcile_run()["device_ms"]
        # end to end: pinned host tables in, host results out, one k8sgpu_reconcile_batch call per step.
        # Inputs in one pinned block in the order the library keeps them (one H2D copy), outputs in one
        # pinned block laid out as k8sgpu_reconcile_out_size says (adjacent copies merge).
        inb = pinned_block(torch, 3 * V + 2 * P)
        pin_in = []
        for i, a in enumerate((t["vm_owner_id"], t["vm_flags"], t["vm_ordinal"])):
            inb[i * V:(i + 1) * V] = a
            pin_in.append(inb[i * V:(i + 1) * V])
        inb[3 * V:3 * V + P] = t["desired"].view(np.uint32)
        inb[3 * V + P:3 * V + 2 * P] = t["pool_owner_id"]
        pargs = (inb[3 * V:3 * V + P].view(np.int32), inb[3 * V + P:3 * V + 2 * P], pin_in[0], pin_in[1],
                 t["n_owner_ids"], pin_in[2])
        cap = int(t["desired"].sum()) + 1
        lay = capi.reconcile_out_size(P, V, cap)
        outb = pinned_block(torch, lay["total"])
        sizes = {"pool_of": V, "current": P, "ready": P, "delta": P, "create_count": P, "victim_off": P + 1,
                 "member_off": P + 1, "create_off": P + 1, "victim_vm": V, "member_vm": V, "create_ordinal": cap}
        order = ["pool_of", "current", "delta", "ready", "create_count", "victim_off", "victim_vm", "member_off",
                 "member_vm", "create_off", "create_ordinal"]  # the order Context._rec_out asks for buffers in
        nxt = [0]

        def alloc(n, dt):
            k = order[nxt[0]]
            nxt[0] += 1
            return outb[lay[k]: lay[k] + sizes[k]].view(dt)

        def call():
            nxt[0] = 0
            return ctx.reconcile_batch(*pargs, alloc=alloc, create_capacity=cap - 1, trim=False)

        call()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            rr = call()
        rdt = time.perf_counter() - t0
        rsha = reconcile_digest(capi.Context._rec_trim({k: np.array(v) for k, v in rr.items() if isinstance(v, np.ndarray)}, P))
        alg = 12 * C2["v"] + 16 * C2["p"]
        line["reconcile"] = {
            "workload": f"config2: {C2['p']} pools x {C2['v']} VMs", "decisions_per_s": C2["p"] * args.steps / (rms * 1e-3),
            "vm_rows_per_s": C2["v"] * args.steps / (rms * 1e-3), "ms_per_step": rms / args.steps,
            "e2e_decisions_per_s": C2["p"] * args.steps / rdt, "e2e_ms_per_step": 1e3 * rdt / args.steps,
            "h2d_bytes_per_step": int(rr["h2d_bytes"]), "d2h_bytes_per_step": int(rr["d2h_bytes"]),
            "launches_per_step": int(rr["kernel_launches"]), "algorithmic_bytes": alg,
            "achieved_gbs": alg * args.steps / (rms * 1e-3) / 1e9, "result_sha256": rsha,
            "note": "1.2 MB of work: launch-latency bound at this size; see profiles/ for the scaled sweep",
        }
        if not args.no_cpu_baseline:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import oracle as orc

            threads = os.cpu_count() or 1
            best = 1e9
            for nt in (1, threads):
                orc.reconcile_batch(*rargs, nthreads=nt)
                t0 = time.perf_counter()
                for _ in range(5):
                    orc.reconcile_batch(*rargs, nthreads=nt)
                dt = (time.perf_counter() - t0) / 5
                if dt < best:
                    best, best_nt = dt, nt
            line["reconcile"]["cpu_baseline"] = {"value": C2["p"] / best, "unit": "decisions/s", "cores": best_nt,
                                                 "kind": "port", "sample": "whole config 2, best of 1 and all threads"}
            # BASELINE.json config 1 is CPU only (one pool against the fake ARM client: nothing to batch)
            line["reconcile"]["config1_cpu"] = {
                "ns_per_reconcile": orc.config1_ns_per_reconcile(), "cores": 1, "kind": "port",
                "sample": "replicas 3->5->2 with a settle pass each (K1-K5, 6 Reconciles) x 20000, fake in-memory ARM"}

    if not args.no_cpu_baseline and world == 1:
        threads = os.cpu_count() or 1
        frac = CPU_SAMPLE_FRAC[wl]
        cpu_port(nodes, pods, min(frac, 0.05), threads)  # thread pool up, tables in cache
        v, ev, dt, sample = cpu_port(nodes, pods, frac, threads)
        line["cpu_baseline"] = {"value": v, "unit": "evals/s", "cores": threads, "kind": "port", "sample": sample,
                                "sample_frac": frac, "seconds": dt}
    EMIT(line)
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
