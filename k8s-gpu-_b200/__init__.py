"""k8sgpu_b200 -- B200-native batch reconcile + GPU pod placement (the one data-parallel
path of Andy-ckm/K8S-GPU-).  The directory is named `k8s-gpu-_b200`; the
root-level shim `k8sgpu_b200.py` registers it under the importable name `k8sgpu_b200`.
"""
__all__ = ["synth"]
