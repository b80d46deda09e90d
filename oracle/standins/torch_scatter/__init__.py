"""TEST INFRASTRUCTURE ONLY -- stand-in for the third-party `torch_scatter` package (conda `pytorch-scatter`,
unpinned in /root/reference/environment.yaml:10; absent from this image and from /root/reference).

The reference's Python path uses two of its functions: `scatter_sum` (droid_slam/geom/ba.py:8) and `scatter_mean`
(droid_slam/droid_net.py:18,67).  Their published definitions (pytorch_scatter README / docs):
    scatter_sum(src, index, dim, dim_size)[..., i, ...]  = sum of src[..., j, ...] over all j with index[j] == i
    scatter_mean = scatter_sum / clamp(scatter_sum(ones), min=1)      (floor division for integer tensors)
restated here over torch.Tensor.index_add_.  `index` is 1-d along `dim` (the only form the reference uses).
This file is written for this repo; nothing in it comes from the torch_scatter sources.
"""
import torch


def _norm_dim(src, dim):
    return dim + src.dim() if dim < 0 else dim


def scatter_sum(src, index, dim=-1, out=None, dim_size=None):
    dim = _norm_dim(src, dim)
    if index.dim() != 1 or index.numel() != src.shape[dim]:
        raise NotImplementedError("stand-in torch_scatter: index must be 1-d along `dim`")
    if dim_size is None:
        dim_size = int(index.max().item()) + 1 if index.numel() > 0 else 0
    shape = list(src.shape)
    shape[dim] = dim_size
    res = torch.zeros(shape, dtype=src.dtype, device=src.device) if out is None else out
    return res.index_add_(dim, index.to(device=src.device, dtype=torch.long), src)


scatter_add = scatter_sum


def scatter_mean(src, index, dim=-1, out=None, dim_size=None):
    dim = _norm_dim(src, dim)
    total = scatter_sum(src, index, dim, out, dim_size)
    n = total.shape[dim]
    count = torch.zeros(n, dtype=src.dtype if src.is_floating_point() else torch.long, device=src.device)
    count.index_add_(0, index.to(device=src.device, dtype=torch.long), torch.ones_like(index, dtype=count.dtype, device=src.device))
    count = count.clamp(min=1).view([n if d == dim else 1 for d in range(total.dim())])
    return total / count if total.is_floating_point() else total // count
