import sys; sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, torch, bench
dev = torch.device('cuda:0')
w = dict(bench.WORKLOADS["c2"])
P = bench.pkg()
means = bench.blob_means(w, torch, dev)
learn = bench.draw(w, torch, dev, means, w["nlearn"], 1234 + 2)
cent = learn[torch.randperm(w["nlearn"], device=dev, generator=torch.Generator(device=dev).manual_seed(7))[:w["nc"]]].clone()
idx = P.IndexIVF_HNSW_Grouping(w["d"], w["nc"], w["M"], 8, w["nsubc"])
for it in range(2):
    idx.build_quantizer(centroids=cent.cpu().numpy())
    res = {}
    for impl in (0, 1):
        idx.set_option("coarse_impl", impl)
        lab = torch.empty(learn.shape[0], dtype=torch.int32, device=dev); dt = torch.empty(learn.shape[0], device=dev)
        idx.coarse_dev(learn, 1, dt, lab); idx.sync()
        res[impl] = (lab.clone(), dt.clone())
    a, b = res[0][0].long(), res[1][0].long()
    bad = (a != b).nonzero().flatten()
    print("iter", it, "agree", (a == b).float().mean().item(), "nbad", bad.numel())
    if bad.numel():
        r = bad[:2000]
        xa = (learn[r].double() - cent[a[r]].double()).pow(2).sum(1); xb = (learn[r].double() - cent[b[r]].double()).pow(2).sum(1)
        print("   on disagreeing rows: impl0 better", (xa < xb).sum().item(), "impl1 better", (xb < xa).sum().item(), "equal", (xa == xb).sum().item())
        print("   example", r[:3].tolist(), a[r[:3]].tolist(), b[r[:3]].tolist(), xa[:3].tolist(), xb[:3].tolist(), res[0][1][r[:3]].tolist(), res[1][1][r[:3]].tolist())
    lab = res[1][0]
    sums = torch.zeros_like(cent).index_add_(0, lab.long(), learn)
    cnt = torch.zeros(w["nc"], device=dev).index_add_(0, lab.long(), torch.ones(w["nlearn"], device=dev))
    cent = torch.where(cnt[:, None] > 0, sums / cnt.clamp(min=1)[:, None], cent)
    print("   empty clusters", (cnt == 0).sum().item(), "dup centroids?", cent.shape[0] - torch.unique(cent, dim=0).shape[0])
