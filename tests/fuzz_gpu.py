"""Randomised GPU-vs-oracle comparison (not collected by pytest: run it on the B200 box by hand):

    python tests/fuzz_gpu.py [seconds] [seed]

Every case draws random table shapes and contents -- including heavy skew (all pods alike, all
proposals on one node, a single giant pool), ragged sizes around the tile boundaries and every
kernel variant selectable by environment -- runs the CUDA path through the C-ABI and the CPU
oracle, and demands byte-identical results.  Prints the seed of any failing case and exits 1.
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import capi  # noqa: E402
import oracle as orc  # noqa: E402

NONE = 0xFFFFFFFF
NCPU = os.cpu_count() or 1
EDGE = [1, 2, 7, 8, 9, 63, 64, 65, 255, 256, 257, 511, 512, 513, 2047, 2048, 2049]


def pick_size(rng, hi):
    return int(rng.choice(EDGE)) if rng.random() < 0.35 else int(rng.integers(1, hi))


def random_nodes(rng, n):
    skew = rng.random()
    # (half the cases keep every node at <= 8 free GPUs: with pods that all ask for >= 1 GPU that is
    # the domain of the slot admission; the other half, and pods asking for 0 GPUs, veto it)
    caps = [1, 2, 4, 8, 16] if rng.random() < 0.5 else [1, 2, 4, 8]
    cap = rng.choice(caps, n).astype(np.uint32) if skew > 0.2 else np.full(n, 8, np.uint32)
    free = (rng.integers(0, 17, n) % (cap + 1)).astype(np.uint32) if skew > 0.1 else cap.copy()
    # (CPU / memory per GPU: the same everywhere, or short on some nodes -- whether enough GPUs implies
    # enough of both decides which evals the scoring kernel runs)
    per_cpu = rng.choice([4000, 12000]) if rng.random() < 0.6 else rng.choice([1500, 4000, 12000], n)
    per_mib = rng.choice([28672, 243200]) if rng.random() < 0.6 else rng.choice([9000, 28672, 243200], n)
    return dict(free_gpu=free, free_mcpu=(free * per_cpu).astype(np.uint32),
                free_mib=(free * per_mib).astype(np.uint32), cap_gpu=cap,
                sku_id=rng.integers(0, rng.choice([1, 2, 32]), n).astype(np.uint32),
                zone_id=rng.integers(0, rng.choice([1, 3, 32]), n).astype(np.uint32),
                taint_mask=np.where(rng.random(n) < rng.choice([0.0, 0.1, 0.5]), rng.integers(1, rng.choice([2, 8]), n), 0).astype(np.uint32),
                flags=(rng.random(n) > 0.03).astype(np.uint32))


def random_pods(rng, m):
    alike = rng.random() < 0.25
    sizes = [0, 1, 1, 1, 2, 4, 8, 40] if rng.random() < 0.4 else [1, 1, 1, 2, 4, 8]
    g = np.full(m, rng.choice([1, 2]), np.uint32) if alike else rng.choice(sizes, m).astype(np.uint32)
    anymask = lambda p: np.where(rng.random(m) < p, rng.integers(0, 2 ** 32, m, dtype=np.uint64), NONE).astype(np.uint32)
    return dict(req_gpu=g, req_mcpu=(g * 4000).astype(np.uint32), req_mib=(g * 28672).astype(np.uint32),
                sku_mask=np.full(m, NONE, np.uint32) if alike else anymask(0.3),
                zone_mask=np.full(m, NONE, np.uint32) if alike else anymask(0.2),
                tol_mask=np.zeros(m, np.uint32) if alike else
                (rng.integers(0, 8, m) if rng.random() < 0.5 else rng.choice([0, 0xFFFFFFFF], m)).astype(np.uint32),
                pref_zone=np.full(m, NONE, np.uint32) if alike else
                np.where(rng.random(m) < 0.5, rng.integers(0, 32, m), NONE).astype(np.uint32),
                priority=np.zeros(m, np.uint32))


def same(got, want, keys):
    return all(np.array_equal(got[k], want[k]) for k in keys)


def case_place(rng, ctx):
    n, m = pick_size(rng, 3000), pick_size(rng, 12000)
    nodes, pods = random_nodes(rng, n), random_pods(rng, m)
    flags = int(rng.random() < 0.2)
    env = dict(res=int(rng.choice([0, 1, 2, 3, 4, 4, 4, 4])), ppt=int(rng.choice([1, 2, 4])), coop=int(rng.choice([0, 1, 1])),
               order=int(rng.choice([0, 1, 1])),  # order: scoring order sorted by pod class (RES 4 + one-launch admission)
               slot_admit=int(rng.choice([0, 1, 1])))  # slot admission where its domain holds, else the sort
    if rng.random() < 0.3:  # 16: toleration clause always evaluated, 17: no clauses shared between a thread's pods
        env["force_cls"] = int(rng.choice([16, 17, 18]))  # 18: CPU / memory differences never left out
    if rng.random() < 0.5:
        env["splits"] = int(rng.integers(1, 9))
    if rng.random() < 0.5:  # few CTAs in the single-launch admission: every CTA loops over several tiles
        env["coop_grid"] = int(rng.choice([1, 2, 3, 5, 17, 64]))
    ctx.set_tuning(**env)
    try:
        if rng.random() < 0.3:
            gang = rng.integers(0, max(m // 5, 1), m).astype(np.uint32)
            gang[rng.random(m) < 0.3] = NONE
            mins = rng.integers(0, 7, max(m // 5, 1)).astype(np.uint32)
            got = ctx.place_gang(nodes, pods, gang, mins, flags=flags)
            want = orc.place_gang(nodes, pods, gang, mins, flags=flags, nthreads=NCPU)
            ok = same(got, want, ("assign", "free_gpu", "free_mcpu", "free_mib", "gang_ok")) and \
                got["gang_iterations"] == want["iterations"]
        else:
            got = ctx.place(nodes, pods, flags=flags)
            want = orc.place(nodes, pods, flags=flags, nthreads=NCPU)
            ok = same(got, want, ("assign", "free_gpu", "free_mcpu", "free_mib"))
        ok = ok and (got["rounds"], got["evals"], got["placed"]) == (want["rounds"], want["evals"], want["placed"])
    finally:
        ctx.reset_tuning()
    return ok, f"place n={n} m={m} flags={flags} tuning={env}"


def case_reconcile(rng, ctx):
    P, V = pick_size(rng, 4000), pick_size(rng, 150_000)
    # path 0 = one launch (few pools), 1 = pruned ranking (victims only), 2 = full stable sort
    members = bool(rng.random() < 0.6)
    env = dict(rec_path=int(rng.integers(0, 3)))
    if rng.random() < 0.5:
        env["rec_grid"] = int(rng.choice([1, 2, 3, 5, 17, 64]))
    if rng.random() < 0.5:
        env["rec_segments"] = int(rng.choice([1, 2, 3, 7, 40]))
    ctx.set_tuning(**env)
    try:
        n_ids = P + int(rng.integers(0, 50))
        powner = rng.permutation(n_ids)[:P].astype(np.uint32)
        if rng.random() < 0.2:  # one giant pool
            vowner = np.full(V, powner[0], np.uint32)
        else:
            vowner = rng.integers(0, n_ids + 20, V).astype(np.uint32)
        vowner[rng.random(V) < 0.05] = NONE
        vflags = (rng.random(V) < 0.9).astype(np.uint32)
        ords = rng.integers(0, 64, V).astype(np.uint32)
        ords[rng.random(V) < 0.1] = NONE
        desired = rng.integers(0, max(2 * V // max(P, 1), 3), P).astype(np.int32)
        args = (desired, powner, vowner, vflags, n_ids, ords if rng.random() < 0.8 else None)
        got, want = ctx.reconcile_batch(*args, members=members), orc.reconcile_batch(*args, nthreads=NCPU)
        ok = all(np.array_equal(got[k], want[k]) for k in want if members or not k.startswith("member_"))
    finally:
        ctx.reset_tuning()
    return ok, f"reconcile P={P} V={V} members={members} tuning={env}"


def main():
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
    seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    t0, n, bad = time.time(), 0, 0
    with capi.Context(0) as ctx:
        while time.time() - t0 < budget:
            seed = seed0 + n
            rng = np.random.default_rng(seed)
            ok, desc = (case_place if rng.random() < 0.7 else case_reconcile)(rng, ctx)
            n += 1
            if not ok:
                bad += 1
                print(f"MISMATCH seed={seed}: {desc}", flush=True)
    print(f"fuzz: {n} cases, {bad} mismatches, {time.time() - t0:.0f} s", flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
