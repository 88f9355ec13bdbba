"""Coarse k-means for the centroid file.

The reference does not train its coarse centroids ("centroids file create by k-means tool, not here",
IndexIVF_HNSW.cpp:703; users download centroids_sift1b.fvecs, README.md:56).  Lloyd's iterations need exactly the
operation this library accelerates -- the nearest centroid of every training vector, K-A1 -- so the centroid file of
a 2^20-centroid index becomes minutes of GEMM instead of an offline artefact.  The whole loop runs in the CUDA library
(ivfhnsw_b200_train_centroids, csrc/train.cu): faiss::Clustering::train with K-A1 as the assignment step, a
double-precision scatter-add update kernel and km_update_centroids' re-seeding of void clusters.
"""


def lloyd_centroids(index, learn_dev, nc, iters=2, seed=7, on_iteration=None):
    """`index` is an IndexIVF_HNSW[_Grouping] handle of the right (d, nc); its quantizer is overwritten.  `learn_dev`
    is a float32 CUDA tensor [n][d], n >= nc.  Returns the centroids as a CUDA tensor [nc][d] (also left in the index).
    on_iteration, when given, is called once with (iters - 1, mean squared distance of the last assignment, void
    clusters re-seeded)."""
    import torch
    if learn_dev.shape[0] < nc:
        raise ValueError("need at least nc training vectors")
    obj, nsplit = index.train_centroids(learn_dev.contiguous(), niter=iters, seed=seed)
    if on_iteration is not None:
        on_iteration(iters - 1, obj / learn_dev.shape[0], nsplit)
    return torch.from_numpy(index.get_centroids()).to(learn_dev.device)
