"""16 vs 20 candidates in the CTA-pair kernel on small shards, INCLUDING the certificate re-runs each choice causes:
mean device time per 4096-query search over many different query batches (library event timers)."""
import importlib, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
vs = importlib.import_module(P + ".vectorstore"); nat = importlib.import_module(P + "._native")
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
for n in (125_000, 400_000, 1_250_000):
    ix = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
    for c0 in range(0, n, 1 << 20):
        ix.add(torch.randn((min(1 << 20, n - c0), 384), generator=g, device=dev).bfloat16().contiguous())
    qs = [torch.randn((4096, 384), generator=g, device=dev).bfloat16().contiguous() for _ in range(24)]
    for kp in (16, 20):
        with nat.options(sim_pair_kp=kp):
            for q in qs[:3]: ix.search(q, 10)
            torch.cuda.synchronize()
            c0_, r0_ = ix.certificate_stats()
            nat.profile(True)
            for q in qs: ix.search(q, 10)
            torch.cuda.synchronize()
            pr = nat.profile_read(); nat.profile(False)
            c1_, r1_ = ix.certificate_stats()
        per = {k: v["ms"] / len(qs) for k, v in pr.items()}
        print(json.dumps({"rows": n, "kp": kp, "total_ms": round(sum(per.values()), 4), **{k: round(v, 4) for k, v in per.items()},
                          "reruns_per_search": (r1_ - r0_) / len(qs)}), flush=True)
    ix.close()
