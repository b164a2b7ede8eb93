"""The N>1 placement protocol (k8sgpu_b200.placement.ShardedPlacer) on CPU: world_size 2 and 3
over gloo, with the CPU oracle standing in for the GPU engine.  Checks that slicing the active
list across ranks + all-gather + replicated admission gives exactly the single-process answer
on every rank (G-invariance)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleEngine:
    """CPU stand-in with the CudaEngine interface (tests only: the product never imports the oracle)."""

    def __init__(self, orc):
        self.orc = orc

    def upload(self, nodes, pods):
        self.nodes, self.pods = nodes, pods
        self.n, self.m = len(nodes["flags"]), len(pods["req_gpu"])

    def begin(self, flags, max_rounds):
        self.flags = flags
        self.free = [self.nodes[c].copy() for c in ("free_gpu", "free_mcpu", "free_mib")]
        self.assign = np.full(self.m, -2, np.int32)
        self.active = np.arange(self.m, dtype=np.uint32)
        self.rounds, self.evals = 0, 0
        return self.m

    def score(self, lo, hi, slice_len):
        out = torch.full((max(slice_len, 1),), -7, dtype=torch.int32)  # padding is never read
        if hi > lo:
            best = self.orc.score_round(self.nodes, self.free, self.pods, self.active[lo:hi], flags=self.flags)
            out[: hi - lo] = torch.from_numpy(best)
        return out[:slice_len]

    def new_gather_buffer(self, numel):
        return torch.empty(max(numel, 1), dtype=torch.int32)

    def admit(self, all_best):
        na = len(self.active)
        best = all_best[:na].numpy().copy()
        self.evals += na * self.n
        self.rounds += 1
        self.active = self.orc.admit_round(self.n, self.free, self.pods, self.active, best, self.assign)
        return len(self.active)

    def end(self):
        return {"rounds": self.rounds, "evals": self.evals, "placed": int((self.assign >= 0).sum())}

    # gang steps, restated from k8sgpu.h "gang placement" (steps 2 and 3)
    def gang_begin(self, gang_id, min_available, flags, max_rounds):
        self.gang = np.asarray(gang_id, np.uint32)
        self.gmin = np.asarray(min_available, np.uint32)
        self.rejected = np.zeros(len(self.gmin), bool)
        self.iterations = 1
        return self.begin(flags, max_rounds)

    def gang_verdicts(self):
        has = self.gang != 0xFFFFFFFF
        gid = np.where(has, self.gang, 0).astype(np.int64)
        count = np.bincount(gid[has & (self.assign >= 0)], minlength=len(self.gmin))
        newly = ~self.rejected & (count < self.gmin)
        if not newly.any():
            return 0
        self.rejected |= newly
        hit = has & newly[gid]
        rel = hit & (self.assign >= 0)
        for col, req in zip(self.free, ("req_gpu", "req_mcpu", "req_mib")):
            np.add.at(col, self.assign[rel], self.pods[req][rel])
        self.assign[hit] = -3
        retry = (self.assign == -1) & (~has | ~self.rejected[gid])
        self.assign[retry] = -2
        self.active = np.nonzero(retry)[0].astype(np.uint32)
        if len(self.active):
            self.iterations += 1
        return len(self.active)

    def gang_end(self):
        st = self.end()
        st["gang_ok"] = (~self.rejected).astype(np.int32)
        st["gang_iterations"] = self.iterations
        return st

    def download(self):
        return {"assign": self.assign, "free_gpu": self.free[0], "free_mcpu": self.free[1], "free_mib": self.free[2]}


def _worker(rank, world, port, case, q):
    for p in (ROOT, os.path.join(ROOT, "oracle")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import placement, synth
    import oracle as orc

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        if case == "hetero":
            nodes, pods, flags = synth.gen_nodes(300, config_id=5), synth.gen_pods(1501, config_id=5), 0
        else:
            nodes, pods = synth.gen_identical(40, 400)
            flags = 1 if case == "ties_rot0" else 0
        pl = placement.ShardedPlacer(dist=dist, rank=rank, world=world, engine=OracleEngine(orc))
        if case == "gang":
            rng = np.random.default_rng(21)
            nodes, pods = synth.gen_nodes(60, config_id=5), synth.gen_pods(700, config_id=5)
            gang = rng.integers(0, 40, 700).astype(np.uint32)
            gang[rng.random(700) < 0.3] = 0xFFFFFFFF
            mins = rng.integers(0, 12, 40).astype(np.uint32)
            pl.upload(nodes, pods)
            st = pl.run_gang(gang, mins)
            got = pl.download()
            want = orc.place_gang(nodes, pods, gang, mins)
            ok = np.array_equal(st["gang_ok"], want["gang_ok"]) and st["gang_iterations"] == want["iterations"]
            ok = ok and want["iterations"] > 1  # the case must exercise a retry
        else:
            pl.upload(nodes, pods)
            st = pl.run(flags=flags)
            got = pl.download()
            want = orc.place(nodes, pods, flags=flags)
            ok = True
        ok = ok and all(np.array_equal(got[k], want[k]) for k in ("assign", "free_gpu", "free_mcpu", "free_mib"))
        ok = ok and st["rounds"] == want["rounds"] and st["evals"] == want["evals"] and st["placed"] == want["placed"]
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,case", [(2, "hetero"), (2, "ties_rot0"), (3, "hetero"), (2, "ties_rot"), (2, "gang")])
def test_sharded_placement_is_rank_count_invariant(orc, world, case):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(r, True) for r in range(world)]


def test_slices_tile_the_active_list():
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200.placement import ShardedPlacer

    for na in (0, 1, 7, 8, 9, 1000, 1001):
        for world in (1, 2, 3, 4, 8):
            cover, prev_hi = 0, 0
            for r in range(world):
                lo, hi, sl = ShardedPlacer.slice_of(na, r, world)
                assert lo == min(r * sl, na) and lo <= hi <= na and hi - lo <= sl
                assert lo == prev_hi or lo == na
                prev_hi = hi
                cover += hi - lo
            assert cover == na and sl * world >= na
