"""Densification on the flat parameter bucket: GaussianModel.densify_and_prune (scene/gaussian_model.py:645-659) as two
small kernels around the stream compaction of vcr_view_parallel.prune_and_grow.

The reference runs densify_and_clone (:628-643), densify_and_split_along_maxscaling (:579-626) and the final prune
(:651-657) one after the other, each rebuilding all six parameter tensors and both Adam moments with boolean-mask
indexing and torch.cat (~60 launches and as many allocations).  Every decision is a function of the state before the
round, so here one kernel decides per Gaussian (keep / append a clone / append two children), the kept rows of
parameters and moments are compacted once, and one kernel writes the new rows.  Same row order, same values.

View-parallel training: call it on every rank after ViewParallel.sync_densification_stats(); the inputs are then
identical on all ranks and so are the resulting buckets (the split is RNG-free).  CUDA only."""
import ctypes
import math
from typing import Optional

import torch

import vcr_native as _n
import vcr_view_parallel as VP

_RAW = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def densify_and_prune(bucket, optimizer, xyz_gradient_accum, denom, max_radii2D, max_grad: float, min_opacity: float, extent: float,
                      max_screen_size: Optional[float] = None, percent_dense: float = 0.01, grad_storage=None):
    """One densification round.  `bucket`: a GaussianBucket of the RAW parameters, groups named
    ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation") in that order (the reference's optimizer groups);
    `optimizer`: its BucketAdam or None; the three statistics tensors as the trainer keeps them ([P,1], [P,1], [P]).
    `max_radii2D` is accepted for signature parity only: the reference resets it to zero (densification_postfix) before
    its prune reads it, so it never decides anything.  Arguments after that as densify_and_prune(max_grad, min_opacity,
    extent, max_screen_size) + training_args.percent_dense.
    Returns (new_bucket, new_optimizer, info) with info = dict(n_kept, n_clone, n_split, P).  The statistics tensors have
    to be re-created with the new size (zeros) by the caller, as densification_postfix does.  One host read (three
    counters) to size the new bucket."""
    if tuple(bucket.names) != _RAW:
        raise ValueError(f"densify_and_prune needs a bucket of the raw parameter groups {_RAW}, got {tuple(bucket.names)}")
    if not bucket.flat.is_cuda:
        raise RuntimeError("densify_and_prune needs a CUDA bucket; there is no CPU fallback")
    P, dev = bucket.P, bucket.flat.device
    acc, den = xyz_gradient_accum.contiguous(), denom.contiguous()
    if acc.numel() != P or den.numel() != P or acc.dtype != torch.float32 or den.dtype != torch.float32:
        raise ValueError("xyz_gradient_accum and denom must be float32 with one entry per Gaussian")
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    flags = torch.empty(P, dtype=torch.uint8, device=dev)
    scratch = torch.empty(int(_n.vcr_densify_scratch_bytes(P)), dtype=torch.uint8, device=dev)
    totals = torch.empty(3, dtype=torch.int32, device=dev)
    p = bucket.params
    dense_limit = float(percent_dense * extent)
    world_limit = float(0.1 * extent) if max_screen_size else math.inf
    _n.check(_n.vcr_densify_decide(P, p["scaling"].data_ptr(), p["opacity"].data_ptr(), acc.data_ptr(), den.data_ptr(), float(max_grad),
                                   float(min_opacity), dense_limit, world_limit, flags.data_ptr(), scratch.data_ptr(),
                                   totals.data_ptr(), stream), "densify_decide")
    n_kept, n_clone, n_split = (int(x) for x in totals.tolist())           # the one host read of the round
    n_new = n_clone + 2 * n_split
    new, opt2, _ = VP.prune_and_grow(bucket, flags & 1, n_new=n_new, optimizer=optimizer, grad_storage=grad_storage, n_kept=n_kept)
    if n_new:
        arr_p, arr_i = ctypes.c_void_p * 6, ctypes.c_int * 6
        old_l = {name: (off, n) for name, off, n, _ in bucket.layout}
        new_l = {name: (off, n) for name, off, n, _ in new.layout}
        widths = [old_l[k][1] // P for k in _RAW]
        srcs = [bucket.flat.data_ptr() + 4 * old_l[k][0] for k in _RAW]
        dsts = [new.flat.data_ptr() + 4 * new_l[k][0] for k in _RAW]
        _n.check(_n.vcr_densify_fill(P, flags.data_ptr(), scratch.data_ptr(), n_kept, n_clone, n_split, arr_p(*srcs), arr_p(*dsts),
                                     arr_i(*widths), stream), "densify_fill")
    return new, opt2, dict(n_kept=n_kept, n_clone=n_clone, n_split=n_split, P=new.P)
