"""L1 norm per element — ``Blob::asum_data`` (src/utils/blob.cc:251-255) used by the debug dump."""
import torch

from ..device import ctx


def asum(t: torch.Tensor) -> float:
    if t is None or t.numel() == 0:
        return 0.0
    if t.is_cuda and ctx().is_native and t.dtype in (torch.float32, torch.bfloat16) and t.is_contiguous():
        from ..ops import kernels as K
        out = torch.zeros(1, dtype=torch.float32, device=t.device)
        K.asum(t, out)
        return float(out.item()) / t.numel()
    return float(t.float().abs().sum().item()) / t.numel()
