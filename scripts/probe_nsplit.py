"""CTA-pair similarity kernel at 4096 queries: time against the number of corpus slices (option sim_nsplit)."""
import importlib, json, sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
vs = importlib.import_module(P + ".vectorstore"); nat = importlib.import_module(P + "._native")
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_250_000
ix = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
for c0 in range(0, n, 1 << 20):
    ix.add(torch.randn((min(1 << 20, n - c0), 384), generator=g, device=dev).bfloat16().contiguous())
q = torch.randn((4096, 384), generator=g, device=dev).bfloat16().contiguous()
for ns in (0, 2, 4, 5, 9, 13, 14, 18, 23, 27, 37, 46, 74):
    for bound in (3, 0):
        with nat.options(sim_nsplit=ns, sim_bound=bound):
            for _ in range(3): ix.search(q, 10)
            torch.cuda.synchronize()
            nat.profile(True)
            for _ in range(8): ix.search(q, 10)
            torch.cuda.synchronize()
            pr = nat.profile_read(); nat.profile(False)
        print(json.dumps({"rows": n, "nsplit": ns, "bound": bound, "sim_ms": round(pr["sim_topk"]["ms"] / pr["sim_topk"]["launches"], 4),
                          "merge_ms": round(pr["merge_rescore"]["ms"] / pr["merge_rescore"]["launches"], 4)}), flush=True)
