import importlib, json, sys, os
import torch
sys.path.insert(0, "/root/repo")
P = "copilot-for-consensus_b200"
vs = importlib.import_module(P + ".vectorstore")
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
for n in (1000, 100_000, 1_250_000):
    ix = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
    ix.add(torch.randn((n, 384), generator=g, device=dev).bfloat16().contiguous())
    for nq in (1, 8):
        q = torch.randn((nq, 384), generator=g, device=dev).contiguous()
        for _ in range(2): ix.search_l2(q, 10)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ix.search_l2(q, 10); e1.record(); torch.cuda.synchronize()
        print(json.dumps({"rows": n, "nq": nq, "ms_per_query": round(e0.elapsed_time(e1) / nq, 4)}), flush=True)
    ix.close()
