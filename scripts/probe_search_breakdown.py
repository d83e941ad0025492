"""Per-kernel breakdown of one search call (library event timers) for shard sizes 10 M / 5 M / 1.25 M rows."""
import importlib, json, sys, os
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = "copilot-for-consensus_b200"
vs = importlib.import_module(P + ".vectorstore"); nat = importlib.import_module(P + "._native")
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
n_max = 10_000_000
ix = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n_max)
for c0 in range(0, n_max, 1 << 20):
    ix.add(torch.randn((min(1 << 20, n_max - c0), 384), generator=g, device=dev).bfloat16().contiguous())
def run(ix, nq, k=10, iters=20):
    q = torch.randn((nq, 384), generator=g, device=dev).bfloat16().contiguous()
    for _ in range(3): ix.search(q, k)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): ix.search(q, k)
    e1.record(); torch.cuda.synchronize()
    wall = e0.elapsed_time(e1) / iters
    nat.profile(True)
    for _ in range(5): ix.search(q, k)
    torch.cuda.synchronize()
    pr = nat.profile_read(); nat.profile(False)
    return wall, {k_: round(v["ms"] / v["launches"], 4) for k_, v in pr.items()}
for n in (10_000_000, 5_000_000, 1_250_000):
    if n != n_max:
        ix.close()
        ix = vs.B200Index(384, device=0, master_dtype="bfloat16", capacity=n)
        for c0 in range(0, n, 1 << 20):
            ix.add(torch.randn((min(1 << 20, n - c0), 384), generator=g, device=dev).bfloat16().contiguous())
    for nq in (1, 128, 512, 4096):
        wall, br = run(ix, nq)
        print(json.dumps({"rows": n, "nq": nq, "ms_per_call": round(wall, 4), "hbm_floor_ms": round(n * 768 / 6454.3e9 * 1e3, 4), "kernels_ms": br}), flush=True)
# the exhaustive kernel (same code as the certificate's re-run), through the L2 entry point: ms per query
q8 = torch.randn((8, 384), generator=g, device=dev).contiguous()
for _ in range(2): ix.search_l2(q8, 10)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); ix.search_l2(q8, 10); e1.record(); torch.cuda.synchronize()
print(json.dumps({"exact_kernel_ms_per_query": round(e0.elapsed_time(e1) / 8, 4), "rows": len(ix), "master": "bf16",
                  "hbm_floor_ms": round(len(ix) * 768 / 6454.3e9 * 1e3, 4)}))
