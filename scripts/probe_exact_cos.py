"""The exhaustive kernel in its cosine role (certificate re-runs): time per failing query, forced by switching the
bound filter of the fused path off is not possible, so near-duplicate queries are used to make the certificate fail."""
import importlib, json, sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
vs = importlib.import_module(P + ".vectorstore"); nat = importlib.import_module(P + "._native")
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
NDUP = int(sys.argv[1]) if len(sys.argv) > 1 else 30     # < 54: the screened pass settles; 64: it hands over to the full pass
for n in (1_250_000, 10_000_000):
    rows = torch.randn((n, 384), generator=g, device=dev).bfloat16()
    base = torch.randn((384,), generator=g, device=dev)
    pos = torch.arange(0, NDUP, device=dev) * (n // 70) + 123
    rows[pos] = (base[None, :] * (1 + 1e-3 * torch.randn((NDUP, 384), generator=g, device=dev))).bfloat16()
    ix = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
    for c0 in range(0, n, 1 << 20):
        ix.add(rows[c0:c0 + (1 << 20)].contiguous())
    for nf in (1, 8):
        q = torch.randn((128, 384), generator=g, device=dev).bfloat16()
        q[:nf] = (base[None, :] * (1 + 1e-3 * torch.randn((nf, 384), generator=g, device=dev))).bfloat16()
        q = q.contiguous()
        for _ in range(2): ix.search(q, 10)
        torch.cuda.synchronize()
        c0_, r0_ = ix.certificate_stats()
        nat.profile(True)
        for _ in range(5): ix.search(q, 10)
        torch.cuda.synchronize()
        pr = nat.profile_read(); nat.profile(False)
        c1_, r1_ = ix.certificate_stats()
        nfail = (r1_ - r0_) / 5
        ms = pr["exact_topk"]["ms"] / pr["exact_topk"]["launches"]
        full = pr.get("exact_full", {"ms": 0.0, "launches": 1})
        ms_full = full["ms"] / max(full["launches"], 1)
        print(json.dumps({"rows": n, "near_duplicates": NDUP, "failing_queries_per_search": nfail,
                          "screened_ms_per_search": round(ms, 4), "screened_ms_per_failing_query": round(ms / max(nfail, 1), 4),
                          "full_ms_per_search": round(ms_full, 4)}), flush=True)
    ix.close(); del rows
