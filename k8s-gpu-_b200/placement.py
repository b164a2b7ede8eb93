"""Host-side driver of round-based placement across one process per GPU.

Every rank holds the same node and pod tables (replicated) and the same active list.  Per round
each rank scores only its contiguous slice of the active list on its own GPU, the proposals are
exchanged, and every rank runs the admission step on the full proposal list, so all replicas
stay bit-identical and the result cannot depend on the number of ranks (SURVEY.md 8(e),
single-collective form).

Two exchanges exist:
  * peer stores (default on GPUs, `p2p=True`): the kernel that decodes the scores writes each
    proposal straight into EVERY rank's exchange buffer over NVLink peer memory (CUDA IPC), so
    the all-gather is done by the producing kernel; the only collective left is a one-element
    all-reduce that acts as the barrier before admission;
  * NCCL / gloo all-gather of the slices (`p2p=False`, and the CPU tests).

The engine does the computing.  `CudaEngine` is the product: it calls libk8sgpu.so through the
C-ABI round entry points.  tests/ plug in a CPU stand-in to exercise this protocol with gloo.
"""
from __future__ import annotations

import os


class CudaEngine:
    """Round API of libk8sgpu.so on the caller's current torch CUDA stream."""

    def __init__(self, ctx):
        import torch

        self.torch = torch
        self.ctx = ctx
        self.device = torch.device("cuda", ctx.device)
        self._mine = None
        self.p2p_ready = False
        self._peer_local = 0
        self._peer_m = 0

    def upload(self, nodes, pods):
        self.ctx.place_upload(nodes, pods)
        self.m = len(pods["req_gpu"])

    def _stream(self):
        return self.torch.cuda.current_stream(self.device).cuda_stream

    def begin(self, flags, max_rounds):
        return self.ctx.round_begin(flags, max_rounds, self._stream())

    def score(self, lo, hi, slice_len):
        if self._mine is None or self._mine.numel() < slice_len:
            self._mine = self.torch.empty(max(slice_len, 1), dtype=self.torch.int32, device=self.device)
        out = self._mine[:slice_len]
        self.ctx.round_score(lo, hi, out.data_ptr(), self._stream())
        return out

    def new_gather_buffer(self, numel):
        return self.torch.empty(max(numel, 1), dtype=self.torch.int32, device=self.device)

    def admit(self, all_best):
        return self.ctx.round_admit(all_best.data_ptr(), self._stream())

    # ---- peer-store exchange
    def setup_p2p(self, dist, rank, world):
        """Allocate this rank's exchange buffer, swap IPC handles, map every peer's buffer.
        Done once; repeated only if a later upload has more pods than the buffer holds."""
        if self.p2p_ready and self.m <= self._peer_m:
            return
        if self.p2p_ready:  # a larger table: nobody may free a buffer that a peer still maps
            self.ctx.peer_close()
            dist.barrier()
        self._peer_m = max(self.m, 1)
        handle, self._peer_local = self.ctx.peer_export(self._peer_m)
        handles = [None] * world
        dist.all_gather_object(handles, handle)
        self.ctx.peer_import(rank, world, b"".join(handles))
        self._bar = self.torch.zeros(1, dtype=self.torch.int32, device=self.device)
        self.p2p_ready = True

    def score_p2p(self, lo, hi):
        self.ctx.round_score_p2p(lo, hi, self._stream())

    def admit_p2p(self, round_index):
        # even and odd rounds use different halves of the exchange buffer (see k8sgpu.h)
        return self.ctx.round_admit(self._peer_local + 4 * self._peer_m * (round_index & 1), self._stream())

    def end(self):
        return self.ctx.round_end(self._stream())

    # ---- gang placement, one step at a time (k8sgpu_place_gang_begin / _verdicts / _end)
    def gang_begin(self, gang_id, min_available, flags, max_rounds):
        return self.ctx.gang_begin(gang_id, min_available, flags, max_rounds, self._stream())

    def gang_verdicts(self):
        return self.ctx.gang_verdicts(self._stream())

    def gang_end(self):
        return self.ctx.gang_end(self._stream())

    def download(self):
        return self.ctx.place_download()


class ShardedPlacer:
    def __init__(self, ctx=None, dist=None, rank=0, world=1, engine=None, p2p=None):
        self.engine = engine if engine is not None else CudaEngine(ctx)
        self.ctx = ctx
        self.dist = dist if world > 1 else None
        self.rank, self.world = rank, world
        self._all = None
        if p2p is None:
            p2p = os.environ.get("K8SGPU_P2P", "1") != "0"
        self.p2p = bool(p2p) and self.dist is not None and isinstance(self.engine, CudaEngine)
        self.exchange = "none" if self.dist is None else ("peer-stores" if self.p2p else "all-gather")

    def upload(self, nodes, pods):
        self.engine.upload(nodes, pods)
        if self.p2p:
            try:
                self.engine.setup_p2p(self.dist, self.rank, self.world)
            except Exception as e:  # no peer access between these GPUs: use the collective
                ok = 0
                self._p2p_error = str(e)
            else:
                ok = 1
            t = self.engine.torch.tensor([ok], device=self.engine.device)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)  # all ranks take the same path
            if int(t.item()) != 1:
                self.p2p = False
                self.exchange = "all-gather (peer mapping unavailable)"

    @staticmethod
    def slice_of(n_active, rank, world):
        """Contiguous slice [lo,hi) of the active list that `rank` scores; equal-length slots."""
        slice_len = -(-n_active // world)
        lo = min(rank * slice_len, n_active)
        return lo, min(lo + slice_len, n_active), slice_len

    def _rounds(self, na, rounds, max_rounds):
        """Score / exchange / admit until no pod is active (or max_rounds); returns (active, rounds)."""
        eng = self.engine
        while na and (max_rounds == 0 or rounds < max_rounds):
            lo, hi, slice_len = self.slice_of(na, self.rank, self.world)
            if self.p2p:
                eng.score_p2p(lo, hi)                  # scores + peer stores into every rank's buffer
                self.dist.all_reduce(eng._bar)         # barrier: every rank's stores have landed
                na = eng.admit_p2p(rounds)
            else:
                mine = eng.score(lo, hi, slice_len)
                if self.dist is not None:
                    need = slice_len * self.world
                    if self._all is None or self._all.numel() < need:
                        self._all = eng.new_gather_buffer(need)
                    allb = self._all[:need]
                    self.dist.all_gather_into_tensor(allb, mine)
                else:
                    allb = mine
                na = eng.admit(allb)
            rounds += 1
        return na, rounds

    def _timed(self):
        eng = self.engine
        torch = getattr(eng, "torch", None)
        if torch is not None and isinstance(eng, CudaEngine):
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record()
            return ev
        return None

    @staticmethod
    def _stop(ev, st):
        if ev is not None:
            ev[1].record()
            ev[1].synchronize()
            st["step_ms"] = ev[0].elapsed_time(ev[1])
        return st

    def run(self, flags=0, max_rounds=0):
        """One whole placement from the uploaded tables.  Returns the library's stats dict plus
        step_ms (device time of the step on this rank)."""
        if self.world == 1 and isinstance(self.engine, CudaEngine):
            st = self.ctx.place_run(flags, max_rounds)  # the in-library loop: fewer host round trips
            st["step_ms"] = st["device_ms"]
            return st
        eng = self.engine
        ev = self._timed()
        if self.p2p:
            # no rank may start storing into a peer that is still reading the previous run's last round
            self.dist.all_reduce(eng._bar)
        na = eng.begin(flags, max_rounds)
        self._rounds(na, 0, max_rounds)
        return self._stop(ev, eng.end())

    def run_gang(self, gang_id, min_available, flags=0, max_rounds=0):
        """Gang placement (k8sgpu.h, "gang placement") with the rounds sharded like run(): rounds
        until no pod is active, then the verdict step on every replica (deterministic, so no
        collective), repeated while verdicts hands back a retry list.  The result has gang_ok."""
        eng = self.engine
        ev = self._timed()
        if self.p2p:
            self.dist.all_reduce(eng._bar)
        na = eng.gang_begin(gang_id, min_available, flags, max_rounds)
        rounds = 0
        while True:
            na, rounds = self._rounds(na, rounds, max_rounds)
            if na:  # max_rounds reached: no verdicts on half a placement
                break
            na = eng.gang_verdicts()
            if na == 0:
                break
        return self._stop(ev, eng.gang_end())

    def download(self):
        return self.engine.download()


def place(nodes: dict, pods: dict, flags: int = 0, max_rounds: int = 0, device: int = 0) -> dict:
    """Convenience: single-GPU placement of host tables (numpy uint32 columns)."""
    from . import capi

    with capi.Context(device) as ctx:
        return ctx.place(nodes, pods, flags, max_rounds)
