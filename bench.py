#!/usr/bin/env python
"""bench.py -- the retrieval hot path on B200: embed a batch of chunks, then cosine top-10.

Workload (BASELINE.json configs[1]): MiniLM-L6-class encoder (384-d, 6 layers, random-init), one
batch of 512 chunks x 256 tokens per GPU per step, each resulting embedding searched top-10 over a
1 M x 384 bf16 in-memory corpus.  A "step" = encode the batch + search its embeddings.

  python bench.py [--gpus N] [--steps K] [--warmup W]          # this repo's CUDA path
  python bench.py --impl reference [...]                        # the reference's CPU path (port)

Multi-GPU (launched by torch.distributed.run, one rank per GPU): embedding is data-parallel with
no collective (every rank encodes its own 512 chunks); the 1 M-row corpus is row-sharded, queries
are all-gathered, each rank searches its shard, candidates are all-gathered and merged.  Per-GPU
work is constant in N => "scaling": "weak".  value = chunks embedded-and-searched per second over
all ranks, timed on the device with CUDA events, max over ranks.

Only the `cpu_baseline` leg and `--impl reference` touch oracle/ (as the thing being timed on the
host cores); the CUDA arm never does.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "copilot-for-consensus_b200"

METRIC = "chunks/sec embedded (MiniLM-L6 seq256) and searched top-10 (cosine, 384-d bf16 corpus)"
UNIT = "chunks/s"
FLOP_PER_CHUNK = 6.040e9          # 23 592 960 flop/token x 256 tokens


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=512, help="chunks per GPU per step")
    ap.add_argument("--seq", type=int, default=256)
    ap.add_argument("--corpus-rows", type=int, default=1_000_000, help="global corpus rows (sharded over GPUs)")
    ap.add_argument("--top-k", type=int, default=10)
    ap.add_argument("--no-extra", action="store_true", help="skip the 10M-row search-only measurement")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tf_burst": d["bf16_tflops"], "tf_sustained": d["bf16_tflops_sustained"],
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 20 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's call pattern on the host cores (oracle port; /root/reference is absent
# on the GPU box and is pure Python + third-party torch, so there is no compiled oracle/_ref)
# ------------------------------------------------------------------------------------------------
class CpuArm:
    """The reference's call pattern on the host cores, as a bounded sample of the workload:
    chunks go through the fp32 torch encoder ONE sequence per call (embedding/app/service.py:384-393
    calls embed() per chunk) and queries through the per-row Python loop of InMemoryVectorStore.query
    (inmemory.py:111-118) over a `slice_rows`-row slice of the corpus; the query time is scaled
    linearly to the full corpus (the loop is O(rows))."""

    def __init__(self, seq: int, corpus_rows: int, top_k: int, slice_rows: int):
        import torch

        from oracle import encode_oracle as eo
        from oracle.search_oracle import InMemoryPort   # pure Python/numpy; the C oracle library is not loaded here

        self.eo, self.seq, self.corpus_rows, self.top_k, self.slice_rows = eo, seq, corpus_rows, top_k, slice_rows
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        self.cfg = eo.MINILM_L6
        self.w = eo.make_weights(self.cfg, seed=0)
        self.ids, self.mask, _ = eo.synthetic_tokens(64, seq, self.cfg.vocab_size, seed=0)
        eo.encode_fast(self.cfg, self.w, self.ids[:1], self.mask[:1])     # warm-up
        rows = torch.from_numpy(np.random.default_rng(1).standard_normal((slice_rows, self.cfg.hidden), dtype=np.float32))
        rows = rows.bfloat16().float().numpy()
        self.port = InMemoryPort()
        for i0 in range(0, slice_rows, 32):                               # the service inserts <= 32 rows per call
            sl = rows[i0 : i0 + 32]
            self.port.add_embeddings([str(i0 + j) for j in range(len(sl))], [r.tolist() for r in sl], [{} for _ in sl])

    def sample(self, n_encode: int, n_query: int) -> dict:
        eo = self.eo
        t0 = time.perf_counter()
        embs = [eo.encode_fast(self.cfg, self.w, self.ids[i % 64 : i % 64 + 1], self.mask[i % 64 : i % 64 + 1])[0].numpy()
                for i in range(n_encode)]
        t_enc = (time.perf_counter() - t0) / n_encode
        t0 = time.perf_counter()
        for i in range(n_query):
            self.port.query(embs[i % n_encode].tolist(), top_k=self.top_k)
        t_q_slice = (time.perf_counter() - t0) / n_query
        scale = self.corpus_rows / self.slice_rows
        t_q = t_q_slice * scale
        return {
            "value": 1.0 / (t_enc + t_q), "unit": UNIT, "cores": self.cores, "kind": "port",
            "sample": (f"{n_encode} chunks x {self.seq} tokens encoded one per call (fp32 torch, {self.cores} threads): "
                       f"{t_enc*1e3:.1f} ms/chunk; {n_query} queries through the per-row numpy loop over a "
                       f"{self.slice_rows}-row slice: {t_q_slice*1e3:.1f} ms/query, scaled x{scale:.0f} to "
                       f"{self.corpus_rows} rows = {t_q*1e3:.0f} ms/query (single thread by construction)"),
            "encode_ms_per_chunk": t_enc * 1e3, "query_ms_per_chunk": t_q * 1e3,
        }


def run_reference(a) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm = CpuArm(a.seq, a.corpus_rows, a.top_k, slice_rows=20000)
    per_step, last = [], None
    t_begin = time.perf_counter()
    for i in range(a.warmup + a.steps):
        last = arm.sample(n_encode=4, n_query=1)
        if i >= a.warmup:
            per_step.append(last["value"])
        if time.perf_counter() - t_begin > 150 and len(per_step) >= 3:
            break
    v = float(np.mean(per_step))
    cb = dict(last); cb["value"] = v
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": len(per_step),
        "warmup": a.warmup, "ms_per_step": 1e3 / v, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"MiniLM-L6 embed seq {a.seq} one chunk per call + top-{a.top_k} over {a.corpus_rows}x384 "
                               "in-memory corpus, on CPU (reference path); each step is a bounded sample"},
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------
# CUDA arm
# ------------------------------------------------------------------------------------------------
def run_b200(a) -> None:
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the CUDA arm has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    nat = importlib.import_module(PKG + "._native")
    mio = importlib.import_module(PKG + ".embedding.model_io")
    prov_mod = importlib.import_module(PKG + ".embedding.provider")
    vs = importlib.import_module(PKG + ".vectorstore")
    sh = importlib.import_module(PKG + ".sharded")

    # --- model + corpus (synthetic, seeded) ---
    with tempfile.TemporaryDirectory() as td:
        mio.write_random_model(td, "all-MiniLM-L6-v2", seed=0)
        prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device=f"cuda:{local}")
    r0, r1 = sh.shard_range(a.corpus_rows, world, rank)
    g = torch.Generator(device=dev).manual_seed(1000 + rank)
    index = vs.B200Index(384, device=local, master_dtype="bfloat16", capacity=r1 - r0)
    for c0 in range(r0, r1, 1 << 20):
        n = min(1 << 20, r1 - c0)
        index.add(torch.randn((n, 384), generator=g, device=dev, dtype=torch.float32).bfloat16().contiguous())
    searcher = sh.ShardedSearch.over_index(index, r0)

    # per-step inputs: fresh token ids each step from a pool larger than L2 would need to be >126 MB; the
    # dominant streams (activations 100 MB+/layer, corpus 768 MB/shard at N=1) already exceed L2 every step
    tg = torch.Generator(device="cpu").manual_seed(rank)
    ids_h = torch.randint(1000, 30522, (a.batch, a.seq), generator=tg, dtype=torch.int32)
    ids_h[:, 0] = 101
    ids_h[:, -1] = 102
    lens_h = torch.full((a.batch,), a.seq, dtype=torch.int32)
    ids_pin, lens_pin = ids_h.pin_memory(), lens_h.pin_memory()
    d_ids, d_lens = ids_h.to(dev), lens_h.to(dev)

    def step_device():
        emb = prov.embed_tokens(d_ids, d_lens)                      # [batch, 384] fp32, stays in HBM
        return searcher.search_local_queries(emb, a.top_k)

    def step_e2e():
        # through the public API with HOST buffers: ids H2D, embeddings D2H, queries H2D, results D2H
        emb_h = prov.embed_tokens(ids_pin.numpy(), lens_pin.numpy())
        if world == 1:
            return index.search(emb_h, a.top_k, id_base=r0)
        q = torch.from_numpy(emb_h).to(dev)
        ids, sc = searcher.search_local_queries(q, a.top_k)
        return ids.cpu().numpy(), sc.cpu().numpy()

    def timed(fn, steps, warmup, sample_clocks=False):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        sampler = ClockSampler(local) if sample_clocks else None
        if sampler:
            sampler.start()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        clocks = sampler.stop() if sampler else None
        return float(ms.item()), clocks

    ms_total, clocks = timed(step_device, a.steps, a.warmup, sample_clocks=True)
    ms_step = ms_total / a.steps
    value = a.batch * world / (ms_step * 1e-3)
    ms_e2e_total, _ = timed(step_e2e, a.steps, max(3, a.warmup))
    e2e_value = a.batch * world / (ms_e2e_total / a.steps * 1e-3)

    # --- per-kernel timing for the roofline line (CUDA events around each launch, inside the library) ---
    pk = peaks()
    nat.profile(True)
    for _ in range(3):
        step_device()
    torch.cuda.synchronize()
    prof = nat.profile_read()
    nat.profile(False)
    kernels = {}
    tot_ms = sum(v["ms"] for v in prof.values()) or 1.0
    for name, v in prof.items():
        per = v["ms"] / max(v["launches"], 1)
        kernels[name] = {"launches_per_step": v["launches"] / 3, "ms_per_launch": per, "share_of_profiled_ms": v["ms"] / tot_ms,
                         "tflops": v["flops"] / (v["ms"] * 1e9) if v["ms"] else None,
                         "gbs": v["bytes"] / (v["ms"] * 1e6) if v["ms"] else None}
    gemm = [n for n in prof if n.startswith("gemm_") or n.startswith("ffn_")]
    g_ms = sum(prof[n]["ms"] for n in gemm)
    g_fl = sum(prof[n]["flops"] for n in gemm)
    g_n = sum(prof[n]["launches"] for n in gemm)
    roofline = None
    if g_ms > 0:
        ach = g_fl / (g_ms * 1e9)
        roofline = {"bound": "tensor", "kernel": "gemm_bf16_ast_kernel, gemm_ln_kernel, ffn_fused_kernel (the four linear maps of the encoder layer)",
                    "achieved": ach, "peak": pk["tf_sustained"], "unit": "TFLOP/s", "frac": ach / pk["tf_sustained"],
                    "peak_source": pk["source"] + " (sustained bf16 GEMM: kernel timed inside a long step)",
                    "flop_per_launch": g_fl / g_n, "ms_per_launch": g_ms / g_n, "traffic": gemm_traffic_per_launch()}
    sim = prof.get("sim_topk")
    search_roof = None
    if sim and sim["ms"] > 0:
        q_per_pass = a.batch * world
        tf = sim["flops"] / (sim["ms"] * 1e9)
        gbs = sim["bytes"] / (sim["ms"] * 1e6)
        search_roof = {"queries_per_pass": q_per_pass, "rows_per_shard": r1 - r0, "ms_per_launch": sim["ms"] / sim["launches"],
                       "tflops": tf, "frac_tensor": tf / pk["tf_sustained"], "gbs": gbs, "frac_hbm": gbs / pk["hbm_gbs"],
                       "bound": "tensor" if q_per_pass > 212 else "hbm"}

    # launches of OUR kernels per step: embed+LN, 6 x (QKV GEMM, attention, out-proj GEMM + LayerNorm, fused
    # feed-forward block + LayerNorm), pool, prep_queries, sim, merge
    launches_step = 1 + 6 * 4 + 1 + 3 + (1 if world > 1 else 0)
    gpu_launches = launches_step * a.steps

    # --- extra: BASELINE.json configs[2] shape at one GPU, search only (4096 queries x 10M rows bf16) ---
    extra = None
    if not a.no_extra and world == 1:
        try:
            del searcher
            index.close()
            n_big = 10_000_000
            big = vs.B200Index(384, device=local, master_dtype="bfloat16", capacity=n_big)
            for c0 in range(0, n_big, 1 << 20):
                n = min(1 << 20, n_big - c0)
                big.add(torch.randn((n, 384), generator=g, device=dev, dtype=torch.float32).bfloat16().contiguous())
            extra = {}
            for nq in (4096, 128, 1):
                q = torch.randn((nq, 384), generator=g, device=dev, dtype=torch.float32).bfloat16().contiguous()
                f = lambda: big.search(q, a.top_k)
                ms, _ = timed(f, 5, 3)
                ms /= 5
                t_min = max(n_big * 768 / (pk["hbm_gbs"] * 1e9), nq * n_big * 768.0 / (pk["tf_sustained"] * 1e12)) * 1e3
                extra[f"search_10Mx384_q{nq}"] = {"ms_per_batch": ms, "qps": nq / (ms * 1e-3), "roofline_ms": t_min,
                                                  "frac_of_roofline": t_min / ms,
                                                  "bound": "tensor" if nq > 212 else "hbm"}
            big.close()
        except Exception as exc:  # noqa: BLE001 - the extra block must never sink the headline line
            extra = {"error": str(exc)[:200]}
        # BASELINE.json configs[3] architecture at one GPU: 768-d, 12 layers, 12 heads x 64, seq 384
        try:
            with tempfile.TemporaryDirectory() as td:
                mio.write_random_model(td, "all-mpnet-base-v2", seed=0)
                big_prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device=f"cuda:{local}")
            bb, ss = 128, 384
            ids_b = torch.randint(1000, 30522, (bb, ss), generator=g, device=dev, dtype=torch.int32)
            lens_b = torch.full((bb,), ss, device=dev, dtype=torch.int32)
            f = lambda: big_prov.embed_tokens(ids_b, lens_b)
            ms, _ = timed(f, 5, 3)
            ms /= 5
            gflop_chunk = 70.67           # 768-d, 12 layers, seq 384
            extra = extra or {}
            extra["encode_mpnet_class_b128_s384"] = {"ms_per_batch": ms, "chunks_per_s": bb / (ms * 1e-3),
                                                     "tflops_algorithmic": bb * gflop_chunk / ms,
                                                     "frac_of_sustained_peak": bb * gflop_chunk / ms / pk["tf_sustained"]}
            big_prov.close()
        except Exception as exc:  # noqa: BLE001
            extra = extra or {}
            extra["encode_mpnet_class_b128_s384"] = {"error": str(exc)[:200]}

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        cpu = CpuArm(a.seq, a.corpus_rows, a.top_k, slice_rows=50000).sample(n_encode=32, n_query=3)

    if rank == 0:
        h2d = a.batch * a.seq * 4 + a.batch * 4 + a.batch * 384 * 4
        d2h = a.batch * 384 * 4 + a.batch * a.top_k * 12
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"MiniLM-L6 (384-d, 6 layers, random-init) embed {a.batch} chunks x seq {a.seq} per GPU "
                                   f"+ top-{a.top_k} over a {a.corpus_rows}x384 bf16 corpus (row-sharded over GPUs)",
                       "chunks_per_gpu_per_step": a.batch, "seq_len": a.seq, "corpus_rows_global": a.corpus_rows,
                       "parallelism": f"dp{world} embed, row-sharded search",
                       "l2_policy": "inputs larger than L2: per step the encoder streams >1 GB of activations and the "
                                    "search streams the 768 MB corpus shard; no explicit flush"},
            "encode_tflops_algorithmic": FLOP_PER_CHUNK * a.batch / (ms_step * 1e-3) / 1e12,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "provider.embed_tokens(host ids) -> host embeddings -> index.search(host queries) -> host ids/scores"},
            "gpu_launches": gpu_launches,
            "clocks": clocks,
            "roofline": roofline,
            "search_kernel": search_roof,
            "kernels": kernels,
            "cpu_baseline": cpu,
            "extra": extra,
            "peaks": pk,
        }))
    if world > 1:
        dist.destroy_process_group()


def gemm_traffic_per_launch():
    """dram__bytes_read.sum + dram__bytes_write.sum per GEMM launch (mean over the four linear shapes of a
    layer, weighted like the roofline's flops: one launch each), from the committed `ncu --set full`
    capture summarised by scripts/summarize_ncu_full.py; None when the capture is not there."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")) as f:
            t = json.load(f)["dram_bytes_per_launch"]
        g = [v for k, v in t.items() if k.startswith("gemm_") or k.startswith("ffn_")]
        return sum(g) / len(g) if g else None
    except Exception:  # noqa: BLE001
        return None


def main():
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
