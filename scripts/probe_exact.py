"""Exhaustive kernel timing through the L2 entry point: queries x rows."""
import importlib, json, sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
vs = importlib.import_module(P + ".vectorstore")
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
for n, dt in ((1_250_000, "bfloat16"), (5_000_000, "bfloat16"), (1_250_000, "float32")):
    ix = vs.B200Index(384, device=0, master_dtype=dt, capacity=n)
    for c0 in range(0, n, 1 << 20):
        r = torch.randn((min(1 << 20, n - c0), 384), generator=g, device=dev)
        ix.add((r.bfloat16() if dt == "bfloat16" else r).contiguous())
    for nq in (1, 2, 8, 32):
        q = torch.randn((nq, 384), generator=g, device=dev).contiguous()
        for _ in range(2): ix.search_l2(q, 10)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ix.search_l2(q, 10); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        bytes_ = n * 384 * (2 if dt == "bfloat16" else 4)
        print(json.dumps({"rows": n, "master": dt, "nq": nq, "ms_total": round(ms, 4), "ms_per_query": round(ms / nq, 4),
                          "gbs": round(bytes_ * nq / (ms * 1e-3) / 1e9, 1)}), flush=True)
    ix.close()
