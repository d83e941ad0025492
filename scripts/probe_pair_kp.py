"""CTA-pair similarity kernel: 16 vs 20 kept candidates per query, by shard size (library event timers)."""
import importlib, json, sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
vs = importlib.import_module(P + ".vectorstore"); nat = importlib.import_module(P + "._native")
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
for n in (1_250_000, 5_000_000, 10_000_000, 25_000_000, 50_000_000):
    ix = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
    for c0 in range(0, n, 1 << 20):
        ix.add(torch.randn((min(1 << 20, n - c0), 384), generator=g, device=dev).bfloat16().contiguous())
    for nq in (4096,):
        q = torch.randn((nq, 384), generator=g, device=dev).bfloat16().contiguous()
        for kp in (16, 20):
            with nat.options(sim_pair_kp=kp):
                for _ in range(3): ix.search(q, 10)
                torch.cuda.synchronize()
                c0_, r0_ = ix.certificate_stats()
                nat.profile(True)
                for _ in range(10): ix.search(q, 10)
                torch.cuda.synchronize()
                pr = nat.profile_read(); nat.profile(False)
                c1_, r1_ = ix.certificate_stats()
            print(json.dumps({"rows": n, "nq": nq, "kp": kp, "sim_ms": round(pr["sim_topk"]["ms"] / pr["sim_topk"]["launches"], 4),
                              "merge_ms": round(pr["merge_rescore"]["ms"] / pr["merge_rescore"]["launches"], 4),
                              "exact_ms": round(pr["exact_topk"]["ms"] / pr["exact_topk"]["launches"], 4),
                              "full_ms": round(pr["exact_full"]["ms"] / pr["exact_full"]["launches"], 4),
                              "rescored_per_search": (r1_ - r0_) / 10}), flush=True)
    ix.close()
