import torch, json, sys, os
sys.path.insert(0, "/root/repo")
from gcn_b200 import gemm
DEV = torch.device("cuda", 0)
g = torch.Generator(device=DEV).manual_seed(1)
for (K, M, N, mean) in [(400_000, 100, 256, 0.25), (400_000, 100, 256, 0.0), (2_449_029, 100, 256, 0.25), (2_449_029, 100, 256, 1.0)]:
    X = torch.randn(K, M, device=DEV, generator=g)
    G = torch.randn(K, N, device=DEV, generator=g) + mean
    dW, db = gemm.dw_colsum(X, G)
    ref = G.double().sum(0)
    e = float((db.double() - ref).abs().max() / ref.abs().max())
    et = float((G.sum(0).double() - ref).abs().max() / ref.abs().max())
    ew = float((dW.double() - X.double().t() @ G.double()).abs().max() / (X.double().t() @ G.double()).abs().max())
    print(json.dumps({"K": K, "mean": mean, "colsum_rel_err": e, "torch_sum_rel_err": et, "dW_rel_err": ew, "chain": os.environ.get("GCNB_XW_MAX_CHAIN", "64")}), flush=True)
