"""BASELINE.json configs[2] / configs[4]: top-10 over a row-sharded bf16 corpus, queries replicated, NCCL
all-gather of the per-rank candidates + merge.  One process per GPU:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \\
      scripts/bench_search_sharded.py --rows-total 10000000 --queries 4096,128,1
  python scripts/bench_search_sharded.py --rows-total 50000000     # one GPU's 38.4 GB share of the 400 M corpus

Prints one JSON line per query-batch size from rank 0; times are device times (CUDA events), max over ranks."""
import argparse
import importlib
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
P = "copilot-for-consensus_b200"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows-total", type=int, default=10_000_000)
    ap.add_argument("--queries", default="4096,128,1")
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--exchange", default="nccl,peer", help="comma list: nccl (all-gather of hit records) and/or peer (stores into NVLink peer memory)")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    vs = importlib.import_module(P + ".vectorstore")
    sh = importlib.import_module(P + ".sharded")
    r0, r1 = sh.shard_range(a.rows_total, world, rank)
    n = r1 - r0
    g = torch.Generator(device="cuda").manual_seed(1 + rank)
    index = vs.B200Index(384, device=local, master_dtype="bfloat16", capacity=n)
    for c0 in range(0, n, 1 << 20):
        m = min(1 << 20, n - c0)
        index.add(torch.randn((m, 384), generator=g, device="cuda").bfloat16().contiguous())
    searchers = {x: sh.ShardedSearch.over_index(index, r0, exchange=x) for x in a.exchange.split(",") if world > 1 or x == "nccl"}
    hbm, tf = 6555.2e9, 1388.2e12                     # MEASURED_PEAKS.json (copy bandwidth, sustained bf16)
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            mp = json.load(f)
        hbm = float(mp.get("hbm_gbs", hbm / 1e9)) * 1e9
        tf = float(mp.get("bf16_tflops_sustained", tf / 1e12)) * 1e12
    except Exception:  # noqa: BLE001
        pass
    gq = torch.Generator(device="cuda").manual_seed(2)
    for nq, (xname, searcher) in [(int(x), it) for x in a.queries.split(",") for it in searchers.items()]:
        q = torch.randn((nq, 384), generator=gq, device="cuda").bfloat16().contiguous()
        for _ in range(3):
            searcher.search_replicated(q, a.k)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            ids, sc = searcher.search_replicated(q, a.k)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / a.iters], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        t_min = max(n * 768 / hbm, nq * n * 768.0 / tf) * 1e3      # per-rank roofline, ms
        if rank == 0:
            print(json.dumps({"metric": "top-10 queries/s", "exchange": xname, "rows_total": a.rows_total, "rows_per_gpu": n, "n_gpus": world,
                              "queries_per_batch": nq, "ms_per_batch": float(ms), "qps": nq / (float(ms) * 1e-3),
                              "roofline_ms_per_rank": t_min, "frac_of_roofline": t_min / float(ms),
                              "bound": "tensor" if nq > 212 else "hbm"}), flush=True)
    index.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
