"""Snapshot ingestion: manifests and ARM listings -> k8sgpu tables.
CPU tests check the parsers and the table builder against hand-derived answers; the gpu test
feeds the built tables to the CUDA path and to the oracle.

The manifests below are written for this test in the shape of the reference's samples
(AzureVmPool CR: README.md:267-286; Volcano Job: GPU调度平台搭建.md:642-673; job template :534).
"""
import json

import numpy as np
import pytest

NONE = 0xFFFFFFFF

POOL_YAML = """\
apiVersion: compute.my.domain/v1alpha1
kind: AzureVmPool
metadata:
  name: gpu-pool-prod
spec:
  replicas: 2
  resourceGroupName: "MyGpuResourceGroup"
  location: "eastus"
  vmSize: "Standard_NC4as_T4_v3"
  # the VNet must already exist
  vnetName: "MyVnet"
  subnetName: "default"
  imageReference:
    publisher: "Canonical"
    offer: "0001-com-ubuntu-server-jammy"
    sku: "22_04-lts-gen2"
    version: "latest"
  azureCredentialSecret: "azure-credentials"
---
apiVersion: compute.my.domain/v1alpha1
kind: AzureVmPool
metadata:
  name: h100-pool
  namespace: training
spec:
  replicas: 3
  resourceGroupName: MyGpuResourceGroup
  location: eastus
  vmSize: Standard_ND96isr_H100_v5
"""

JOB_YAML = """\
apiVersion: batch.volcano.sh/v1alpha1
kind: Job
metadata:
  name: fashion-mnist
spec:
  minAvailable: 1
  schedulerName: volcano
  queue: default
  tasks:
    - replicas: 2
      name: trainer
      template:
        spec:
          containers:
            - name: pytorch
              image: harbor.local/ai/train:latest
              command: ["python", "train.py"]
              resources:
                limits:
                  nvidia.com/gpu: 1
                  cpu: "4"
                  memory: 28Gi
          restartPolicy: OnFailure
    - replicas: 1
      name: big
      template:
        spec:
          nodeSelector:
            topology.kubernetes.io/zone: eastus-2
            node.kubernetes.io/instance-type: Standard_ND96isr_H100_v5
          tolerations:
          - key: gpu-taint-0
            operator: Exists
          containers:
          - name: main
            resources:
              limits:
                nvidia.com/gpu: 8
                cpu: 64000m
                memory: 1Ti
---
name: template-job
mode: single
singleInstanceType: gpu-1x-16c-32g-1gpu
"""


def vm(name, owner=None, managed="azurevmpool-operator", size="Standard_NC4as_T4_v3", zone="1", rg="MyGpuResourceGroup"):
    tags = {}
    if managed:
        tags["managed-by"] = managed
    if owner:
        tags["owner"] = owner
    return {"name": name, "resourceGroup": rg, "location": "eastus", "zones": [zone],
            "hardwareProfile": {"vmSize": size}, "tags": tags}


VMS = [
    vm("gpu-pool-prod-0", "default-gpu-pool-prod"),
    vm("h100-pool-1", "training-h100-pool", size="Standard_ND96isr_H100_v5", zone="2"),
    vm("gpu-pool-prod-3", "default-gpu-pool-prod", zone="2"),
    vm("stranger", "default-somebody-else"),
    vm("gpu-pool-prod-1", "default-gpu-pool-prod", managed="other-operator"),
    vm("gpu-pool-prod-7", "default-gpu-pool-prod"),
    vm("odd-size", None, managed=None, size="Standard_Unknown_v9", zone="3"),
    vm("elsewhere-0", "default-gpu-pool-prod", rg="OtherGroup"),
]


@pytest.fixture(scope="module")
def hm():
    import __graft_entry__ as g

    g.build()
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import hostmirror

    return hostmirror


@pytest.fixture()
def built(hm):
    s = hm.Snapshot()
    s.add_pools_yaml(POOL_YAML)
    s.add_vms_json(json.dumps(VMS))
    s.add_jobs_yaml(JOB_YAML)
    return s, s.build()


def test_pool_manifest_fields(built):
    s, b = built
    assert b["pool_keys"] == ["default/gpu-pool-prod", "training/h100-pool"]
    assert s.pool_field(0, "vmSize") == "Standard_NC4as_T4_v3" and s.pool_field(0, "vnetName") == "MyVnet"
    assert s.pool_field(0, "imageReference.offer") == "0001-com-ubuntu-server-jammy"
    assert s.pool_field(0, "azureCredentialSecret") == "azure-credentials"
    assert s.pool_field(1, "namespace") == "training" and s.pool_field(1, "resourceGroupName") == "MyGpuResourceGroup"
    assert b["reconcile"]["desired"].tolist() == [2, 3]


def test_reconcile_tables(built, orc):
    _, b = built
    r = b["reconcile"]
    assert r["pool_owner_id"].tolist() == [0, 1] and r["n_owner_ids"] == 2
    #              prod-0  h100-1  prod-3  stranger  wrong-mgr  prod-7  untagged  other-rg
    assert r["vm_owner_id"].tolist() == [0, 1, 0, 2, 0, 0, NONE, 2]
    assert r["vm_flags"].tolist() == [1, 1, 1, 1, 0, 1, 0, 1]
    assert r["vm_ordinal"].tolist() == [0, 1, 3, NONE, 1, 7, NONE, NONE]
    out = orc.reconcile_batch(r["desired"], r["pool_owner_id"], r["vm_owner_id"], r["vm_flags"], r["n_owner_ids"],
                              r["vm_ordinal"])
    assert out["current"].tolist() == [3, 1]             # the wrongly managed / foreign / other-group VMs do not count
    assert out["delta"].tolist() == [-1, 2]
    assert [b["vm_names"][v] for v in out["victim_vm"]] == ["gpu-pool-prod-0"]   # first in list order (README.md:214)
    assert out["create_ordinal"].tolist() == [0, 2]      # h100-pool-0 and h100-pool-2 are the lowest free names


def test_node_and_pod_tables(built):
    _, b = built
    n, p = b["nodes"], b["pods"]
    assert n["cap_gpu"].tolist() == [1, 8, 1, 1, 1, 1, 0, 1]
    assert n["free_mcpu"].tolist()[:2] == [4000, 96000] and n["free_mib"].tolist()[1] == 1945600
    assert n["sku_id"].tolist()[:2] == [0, 1]
    assert n["zone_id"].tolist() == [0, 1, 1, 0, 0, 0, 2, 0]        # eastus-1, eastus-2, eastus-3 in first-seen order
    assert n["flags"].tolist() == [1, 1, 1, 1, 1, 1, 0, 1]          # unknown vmSize: not schedulable
    assert b["pod_names"] == ["fashion-mnist-trainer-0", "fashion-mnist-trainer-1", "fashion-mnist-big-0", "template-job"]
    assert p["req_gpu"].tolist() == [1, 1, 8, 1]
    assert p["req_mcpu"].tolist() == [4000, 4000, 64000, 16000]
    assert p["req_mib"].tolist() == [28672, 28672, 1048576, 32768]
    assert p["sku_mask"].tolist() == [NONE, NONE, 2, NONE]
    assert p["zone_mask"].tolist() == [NONE, NONE, 2, NONE]
    assert p["tol_mask"].tolist() == [0, 0, 1, 0]


def test_quantity_parsers_and_errors(hm):
    s = hm.Snapshot()
    with pytest.raises(ValueError):
        s.add_pools_yaml("kind: AzureVmPool\nmetadata:\n  name: p\nspec:\n  replicas: -1\n")
    with pytest.raises(ValueError):
        s.add_vms_json('[{"name": "x", "tags": {"a": "b"}')      # truncated JSON
    with pytest.raises(ValueError):
        s.add_vms_json('{"name": "x"}')                            # not an array
    with pytest.raises(ValueError):
        s.add_jobs_yaml("kind: Job\nspec:\n  tasks:\n    - replicas: 1\n      template:\n        spec:\n"
                        "          containers:\n            - resources:\n                limits:\n"
                        "                  memory: lots\n")
    with pytest.raises(ValueError):
        s.add_pools_yaml("kind: AzureVmPool\n\tmetadata: {}\n")     # tab indentation
    s.add_vms_json("[]")
    s.add_jobs_yaml("singleInstanceType: gpu-2x-96c-1900g-8gpu\n")
    b = s.build()
    assert b["pods"]["req_gpu"].tolist() == [8] and b["pods"]["req_mib"].tolist() == [1900 * 1024]
    assert len(b["nodes"]["flags"]) == 0 and len(b["reconcile"]["desired"]) == 0


@pytest.mark.gpu
def test_ingested_snapshot_through_the_gpu(built, orc):
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import capi

    _, b = built
    r = b["reconcile"]
    args = (r["desired"], r["pool_owner_id"], r["vm_owner_id"], r["vm_flags"], r["n_owner_ids"], r["vm_ordinal"])
    with capi.Context(0) as ctx:
        got, want = ctx.reconcile_batch(*args), orc.reconcile_batch(*args)
        for k in want:
            assert np.array_equal(got[k], want[k]), k
        gp, wp = ctx.place(b["nodes"], b["pods"]), orc.place(b["nodes"], b["pods"])
        assert np.array_equal(gp["assign"], wp["assign"])
        # the 8-GPU pod can only land on the H100 VM in eastus-2; the 1-GPU pods take T4 nodes (best fit)
        assert b["node_names"][gp["assign"][2]] == "h100-pool-1"
        assert all(b["nodes"]["cap_gpu"][gp["assign"][i]] == 1 for i in (0, 1, 3))


def test_owner_tag_conflict_is_reported(hm):
    """team-a/pool and team/a-pool share the owner tag "team-a-pool" (README.md:238 builds it as
    "<namespace>-<name>"): the second is left out of the tables and listed in conflicts()."""
    mk = lambda ns, name: (f"apiVersion: compute.my.domain/v1alpha1\nkind: AzureVmPool\nmetadata:\n  name: {name}\n"
                           f"  namespace: {ns}\nspec:\n  replicas: 1\n  resourceGroupName: rg1\n  location: eastus\n"
                           f"  vmSize: Standard_NC4as_T4_v3\n")
    s = hm.Snapshot()
    s.add_pools_yaml(mk("team-a", "pool") + "---\n" + mk("team", "a-pool") + "---\n" + mk("team", "other"))
    s.add_vms_json("[]")
    b = s.build()
    assert b["pool_keys"] == ["team-a/pool", "team/other"]
    assert s.conflicts() == [("team/a-pool", "team-a/pool")]
