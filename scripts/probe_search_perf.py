"""K8 tuning sweep (library event timers): cross-CTA bound modes x refresh period, for shard sizes and batch sizes."""
import importlib, json, sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
vs = importlib.import_module(P + ".vectorstore"); nat = importlib.import_module(P + "._native")
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
sizes = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "1250000,10000000").split(",")]
nqs = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "128").split(",")]
for n in sizes:
    ix = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
    for c0 in range(0, n, 1 << 20):
        ix.add(torch.randn((min(1 << 20, n - c0), 384), generator=g, device=dev).bfloat16().contiguous())
    for nq in nqs:
        q = torch.randn((nq, 384), generator=g, device=dev).bfloat16().contiguous()
        for bound in (3, 1, 2, 0):
            for refresh in (0, 1, 3, 7):
                with nat.options(sim_bound=bound, sim_refresh_mask=refresh):
                    for _ in range(3): ix.search(q, 10)
                    torch.cuda.synchronize()
                    nat.profile(True)
                    for _ in range(10): ix.search(q, 10)
                    torch.cuda.synchronize()
                    pr = nat.profile_read(); nat.profile(False)
                print(json.dumps({"rows": n, "nq": nq, "bound": bound, "refresh_mask": refresh,
                                  "sim_ms": round(pr["sim_topk"]["ms"] / pr["sim_topk"]["launches"], 4),
                                  "merge_ms": round(pr["merge_rescore"]["ms"] / pr["merge_rescore"]["launches"], 4)}), flush=True)
    ix.close()
