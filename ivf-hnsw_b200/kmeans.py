"""Coarse k-means for the centroid file.

The reference does not train its coarse centroids ("centroids file create by k-means tool, not here",
IndexIVF_HNSW.cpp:703; users download centroids_sift1b.fvecs, README.md:56).  Lloyd's iterations need exactly the
operation this library accelerates -- the nearest centroid of every training vector, K-A1 -- so the centroid file of
a 2^20-centroid index becomes minutes of GEMM instead of an offline artefact.  Device memory and the scatter-add of
the update step come from PyTorch (plumbing); the assignment runs through the C ABI (ivfhnsw_b200_coarse_dev).
"""
import numpy as np


def lloyd_centroids(index, learn_dev, nc, iters=2, seed=7, on_iteration=None):
    """Lloyd iterations driven by K-A1.  `index` is an IndexIVF_HNSW[_Grouping] handle of the right (d, nc); its
    quantizer is overwritten.  `learn_dev` is a float32 CUDA tensor [n][d].  Initial centroids: a seeded random sample
    of the learn set; an empty cluster keeps its previous centroid.  Returns the centroids as a CUDA tensor [nc][d]
    (also left in the index through build_quantizer)."""
    import torch
    dev = learn_dev.device
    n = learn_dev.shape[0]
    if n < nc:
        raise ValueError("need at least nc training vectors")
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    cent = learn_dev[torch.randperm(n, device=dev, generator=g)[:nc]].clone()
    lab = torch.empty(n, dtype=torch.int32, device=dev)
    dist = torch.empty(n, dtype=torch.float32, device=dev)
    ones = torch.ones(n, device=dev)
    for it in range(iters):
        index.build_quantizer(centroids=cent.cpu().numpy())
        index.coarse_dev(learn_dev, 1, dist, lab)
        index.sync()
        sums = torch.zeros_like(cent).index_add_(0, lab.long(), learn_dev)
        cnt = torch.zeros(nc, device=dev).index_add_(0, lab.long(), ones)
        cent = torch.where(cnt[:, None] > 0, sums / cnt.clamp(min=1)[:, None], cent)
        if on_iteration is not None:
            on_iteration(it, float(dist.mean().item()), int((cnt == 0).sum().item()))
    index.build_quantizer(centroids=np.ascontiguousarray(cent.cpu().numpy()))
    return cent
