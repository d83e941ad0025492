"""Single-tile similarity kernel on small shards: time against live queries and rows (library event timers)."""
import importlib, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
vs = importlib.import_module(P + ".vectorstore"); nat = importlib.import_module(P + "._native")
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
def sim_ms(ix, q, reps=20):
    for _ in range(3): ix.search(q, 10)
    torch.cuda.synchronize()
    nat.profile(True)
    for _ in range(reps): ix.search(q, 10)
    torch.cuda.synchronize()
    pr = nat.profile_read(); nat.profile(False)
    return {k: round(v["ms"] / v["launches"], 4) for k, v in pr.items()}
for n in (312_500, 1_250_000, 5_000_000):
    ix = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
    for c0 in range(0, n, 1 << 20):
        ix.add(torch.randn((min(1 << 20, n - c0), 384), generator=g, device=dev).bfloat16().contiguous())
    for nq in (1, 8, 32, 64, 128):
        q = torch.randn((nq, 384), generator=g, device=dev).bfloat16().contiguous()
        out = {"rows": n, "nq": nq, "hbm_floor_ms": round(n * 768 / 6454.3e9 * 1e3, 4), "default": sim_ms(ix, q)}
        if nq == 128:
            for name, opts in (("no_bounds", {"sim_bound": 0}), ("bucket_only", {"sim_bound": 1})):
                with nat.options(**opts):
                    out[name] = sim_ms(ix, q)["sim_topk"]
        print(json.dumps(out), flush=True)
    ix.close()
