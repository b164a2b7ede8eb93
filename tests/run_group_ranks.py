"""One process per GPU through the C-ABI (k8sgpu_comm_init_rank), launched with torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29512 tests/run_group_ranks.py [--full]

Every rank creates a one-device context, rank 0 makes the ncclUniqueId, all join, and every
k8sgpu_place call is then collective.  Checks, for both exchanges (NCCL all-gather, peer stores):
  * every rank's result equals a plain single-GPU run on its own GPU, byte for byte,
  * the digest equals the committed oracle digest (tests/golden/digests.json) where one exists
    (--full adds config 4 at 100k nodes x 1M pods and the 1M-pod config-5 cases),
  * gang placement equals the single-GPU gang placement.
Not collected by pytest (needs > 1 GPU); torch.distributed only carries the id and the verdict.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import capi, synth  # noqa: E402
from k8sgpu_b200.digest import place_digest  # noqa: E402

sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_digests as md  # noqa: E402  (case table only; the oracle is not called here)

DOC = json.load(open(os.path.join(ROOT, "tests", "golden", "digests.json")))


def main():
    full = "--full" in sys.argv
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    box = [capi.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    names = ["c5_ties_rot0_2k_16k", "c3", "c5_1e5_h50_t5", "c5_ties_rotated_10k_100k"]
    if full:
        names += ["c4", "c5_1e6_h0_t0", "c5_1e6_h100_t50"]
    bad = 0
    with capi.Context(local) as grp, capi.Context(local) as solo:
        grp.comm_init_rank(rank, world, box[0])
        info = grp.group_info()
        assert info["world"] == world and info["rank"] == rank and info["nlocal"] == 1
        for exchange in (0, 1):
            grp.set_tuning(exchange=exchange)
            for name in names:
                c = md.PLACE_CASES[name]
                nodes, pods = md.place_inputs(c)
                got = grp.place(nodes, pods, flags=c["flags"])
                one = solo.place(nodes, pods, flags=c["flags"])
                same = all(np.array_equal(got[k], one[k]) for k in ("assign", "free_gpu", "free_mcpu", "free_mib"))
                same = same and (got["rounds"], got["placed"], got["evals"]) == (one["rounds"], one["placed"], one["evals"])
                pinned = place_digest(got) == DOC["place"][name]["sha256"]
                if not (same and pinned):
                    bad += 1
                if rank == 0:
                    print(f"[x{world} exchange={exchange}] {name}: rounds={got['rounds']} placed={got['placed']} "
                          f"== one GPU: {same}  == oracle digest: {pinned}  ({got['device_ms']:.2f} ms vs {one['device_ms']:.2f} ms)",
                          flush=True)
            # gangs
            rng = np.random.default_rng(5)
            nodes, pods = synth.gen_nodes(2000, 5), synth.gen_pods(30_000, 5)
            gang = rng.integers(0, 5000, 30_000).astype(np.uint32)
            gang[rng.random(30_000) < 0.3] = 0xFFFFFFFF
            mins = rng.integers(0, 7, 5000).astype(np.uint32)
            got, one = grp.place_gang(nodes, pods, gang, mins), solo.place_gang(nodes, pods, gang, mins)
            same = all(np.array_equal(got[k], one[k]) for k in ("assign", "free_gpu", "free_mcpu", "free_mib", "gang_ok"))
            same = same and got["gang_iterations"] == one["gang_iterations"]
            bad += 0 if same else 1
            if rank == 0:
                print(f"[x{world} exchange={exchange}] gangs: iterations={got['gang_iterations']} == one GPU: {same}", flush=True)
    t = torch.tensor([bad], device="cuda")
    dist.all_reduce(t)
    if rank == 0:
        print("GROUP RANKS:", "OK" if int(t.item()) == 0 else f"FAILED ({int(t.item())})", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(t.item()) == 0 else 1)


if __name__ == "__main__":
    main()
