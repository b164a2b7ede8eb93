"""Multi-GPU check, launched with torchrun (one process per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/run_sharded_gpu.py

Every rank runs the sharded placement (its slice of the active list on its own GPU, NCCL
all-gather of the proposals, replicated admission) and checks that
  * its result is byte-identical to a plain single-GPU run of the same tables on its GPU,
  * all ranks hold the same result (hash all-gathered),
  * on the small case, the result equals the CPU oracle (G-invariance).
Not collected by pytest (needs >1 GPU); tests/test_sharded_gloo.py covers the protocol on CPU.
"""
import hashlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import capi, placement, synth  # noqa: E402


def digest(res):
    h = hashlib.sha256()
    for k in ("assign", "free_gpu", "free_mcpu", "free_mib"):
        h.update(np.ascontiguousarray(res[k]).tobytes())
    return h.digest()


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cases = [("small-hetero", synth.gen_nodes(3000, 5), synth.gen_pods(20_000, 5), 0, True),
             ("ties-rot0", *synth.gen_identical(500, 4100), 1, True),
             ("ties-rotated", *synth.gen_identical(500, 4100), 0, True),  # K12: G-invariant with rotation on
             ("c4-slice", synth.gen_nodes(100_000, 4), synth.gen_pods(60_000, 4), 0, False)]
    ok_all = True
    with capi.Context(local) as ctx:
      for p2p in (True, False):  # peer-store exchange, then the NCCL all-gather exchange
        pl = placement.ShardedPlacer(ctx, dist, rank, world, p2p=p2p)
        for name, nodes, pods, flags, check_oracle in cases:
            pl.upload(nodes, pods)
            name = f"{name} [{pl.exchange}]"
            st = pl.run(flags=flags)
            got = pl.download()
            single = ctx.place(nodes, pods, flags=flags)
            same_single = all(np.array_equal(got[k], single[k]) for k in ("assign", "free_gpu", "free_mcpu", "free_mib"))
            same_single = same_single and st["rounds"] == single["rounds"] and st["evals"] == single["evals"]
            d = torch.tensor(list(digest(got)), dtype=torch.uint8, device="cuda")
            alld = [torch.empty_like(d) for _ in range(world)]
            dist.all_gather(alld, d)
            same_ranks = all(torch.equal(alld[0], x) for x in alld)
            same_oracle = None
            if check_oracle and rank == 0:
                import oracle as orc

                want = orc.place(nodes, pods, flags=flags, nthreads=os.cpu_count() or 1)
                same_oracle = all(np.array_equal(got[k], want[k]) for k in ("assign", "free_gpu", "free_mcpu", "free_mib"))
            ok = same_single and same_ranks and same_oracle is not False
            ok_all = ok_all and ok
            if rank == 0:
                print(f"[sharded x{world}] {name}: rounds={st['rounds']} placed={st['placed']} evals={st['evals']} "
                      f"== single-GPU: {same_single}  all ranks equal: {same_ranks}  == oracle: {same_oracle}", flush=True)
        # gang placement with the rounds sharded (stepwise C-ABI) vs. the one-call form and the oracle
        rng = np.random.default_rng(33)
        nodes, pods = synth.gen_nodes(400, 5), synth.gen_pods(6000, 5)
        gang = rng.integers(0, 300, 6000).astype(np.uint32)
        gang[rng.random(6000) < 0.3] = 0xFFFFFFFF
        mins = rng.integers(0, 14, 300).astype(np.uint32)
        pl.upload(nodes, pods)
        st = pl.run_gang(gang, mins)
        got = pl.download()
        single = ctx.place_gang(nodes, pods, gang, mins)
        same = all(np.array_equal(got[k], single[k]) for k in ("assign", "free_gpu", "free_mcpu", "free_mib"))
        same = same and np.array_equal(st["gang_ok"], single["gang_ok"]) and st["rounds"] == single["rounds"]
        same = same and st["gang_iterations"] == single["gang_iterations"]
        same_oracle = None
        if rank == 0:
            import oracle as orc

            want = orc.place_gang(nodes, pods, gang, mins, nthreads=os.cpu_count() or 1)
            same_oracle = np.array_equal(got["assign"], want["assign"]) and np.array_equal(st["gang_ok"], want["gang_ok"])
            print(f"[sharded x{world}] gangs [{pl.exchange}]: rounds={st['rounds']} iterations={st['gang_iterations']} "
                  f"placed={st['placed']} == one-call single-GPU: {same}  == oracle: {same_oracle}", flush=True)
        ok_all = ok_all and same and same_oracle is not False
    flag = torch.tensor([1 if ok_all else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    if int(flag.item()) != 1:
        raise SystemExit("sharded placement mismatch")
    if rank == 0:
        print("sharded placement OK", flush=True)


if __name__ == "__main__":
    main()
