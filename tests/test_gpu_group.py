"""Placement groups: several GPUs behind one k8sgpu_ctx (k8sgpu_create with ndev > 1), driven from
C++ (tests/cpp/test_group.cpp, the cgo caller's view) and from Python, against the one-GPU result
and the CPU oracle.  Needs >= 2 GPUs: skipped on a one-GPU box (run with `gpurun --gpus 2`)."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
NCPU = os.cpu_count() or 1


def gpu_count():
    import torch

    return torch.cuda.device_count()


def test_cpp_caller_drives_two_gpus():
    exe = os.path.join(HERE, "cpp", "test_group")
    if not os.path.exists(exe):
        subprocess.run(["make", "-s", "-C", os.path.join(HERE, "cpp")], check=True)
    r = subprocess.run([exe, "2"], capture_output=True, text=True, timeout=600)
    print(r.stdout, r.stderr)
    if "SKIP" in r.stdout:
        pytest.skip(r.stdout.strip())
    assert r.returncode == 0, r.stdout + r.stderr
    assert "OK:" in r.stdout and "0 mismatches" in r.stdout


@pytest.mark.parametrize("exchange", [0, 1])
def test_group_context_equals_oracle(orc, synth, exchange):
    if gpu_count() < 2:
        pytest.skip("needs 2 GPUs")
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import capi

    nodes, pods = synth.gen_nodes(3000, config_id=5), synth.gen_pods(40_000, config_id=5)
    want = orc.place(nodes, pods, nthreads=NCPU)
    with capi.Context(list(range(min(gpu_count(), 4)))) as ctx:
        info = ctx.group_info()
        assert info["world"] == info["nlocal"] == len(ctx.devices) and info["nccl_version"] > 0
        ctx.set_tuning(exchange=exchange)
        for _ in range(2):  # twice: exchange buffers are reused
            got = ctx.place(nodes, pods)
            for k in ("assign", "free_gpu", "free_mcpu", "free_mib"):
                assert np.array_equal(got[k], want[k]), k
            assert (got["rounds"], got["placed"], got["evals"]) == (want["rounds"], want["placed"], want["evals"])
        # resident form + a smaller table afterwards
        n2, p2 = synth.gen_identical(300, 2500)
        ctx.place_upload(n2, p2)
        ctx.place_run()
        got = ctx.place_download()
        assert np.array_equal(got["assign"], orc.place(n2, p2)["assign"])
        # batch reconcile on a group context runs on its first device
        t = synth.gen_reconcile(100, 20_000)
        args = (t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"], t["vm_ordinal"])
        g, w = ctx.reconcile_batch(*args), orc.reconcile_batch(*args)
        assert all(np.array_equal(g[k], w[k]) for k in w)
