"""CTA-pair similarity kernel: one drain group against two (option sim_drain), independent random queries."""
import importlib, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
vs = importlib.import_module(P + ".vectorstore"); nat = importlib.import_module(P + "._native")
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
for n, nqs in ((125_000, (4096,)), (1_000_000, (512, 4096)), (1_250_000, (4096,)), (2_400_000, (4096,)), (3_500_000, (4096,)), (5_000_000, (4096,))):
    ix = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
    for c0 in range(0, n, 1 << 20):
        ix.add(torch.randn((min(1 << 20, n - c0), 384), generator=g, device=dev).bfloat16().contiguous())
    for nq in nqs:
        q = torch.randn((nq, 384), generator=g, device=dev).bfloat16().contiguous()
        out = {"rows": n, "nq": nq}
        for dr in (1, 2, 1, 2):
            with nat.options(sim_drain=dr):
                for _ in range(3): ix.search(q, 10)
                torch.cuda.synchronize()
                nat.profile(True)
                for _ in range(8): ix.search(q, 10)
                torch.cuda.synchronize()
                pr = nat.profile_read(); nat.profile(False)
            out.setdefault(f"drain{dr}_sim", []).append(round(pr["sim_topk"]["ms"] / 8, 4))
        print(json.dumps(out), flush=True)
    ix.close()
