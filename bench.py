#!/usr/bin/env python
"""bench.py -- the retrieval hot path on B200: embed a batch of chunks, then cosine top-10.

Workload (BASELINE.json configs[1]): MiniLM-L6-class encoder (384-d, 6 layers, random-init), one
batch of 512 chunks x 256 tokens per GPU per step, each resulting embedding searched top-10 over a
1 M x 384 bf16 in-memory corpus.  A "step" = encode the batch + search its embeddings.

  python bench.py [--gpus N] [--steps K] [--warmup W]          # this repo's CUDA path
  python bench.py --impl reference [...]                        # the reference's CPU path, same workload

Multi-GPU (launched by torch.distributed.run, one rank per GPU): embedding is data-parallel with
no collective (every rank encodes its own 512 chunks); the 1 M-row corpus is row-sharded, queries
are all-gathered, each rank searches its shard, hit records are all-gathered and merged.  Per-GPU
work is constant in N => "scaling": "weak".  value = chunks embedded-and-searched per second over
all ranks, timed on the device with CUDA events, max over ranks.

Besides the headline the line carries, under `roofline` (which the driver's record keeps):
  search_10M   BASELINE.json's second metric half and configs[2]: top-10 over 10 M x 384 bf16, 4096 / 128 / 1
               queries per batch, row-sharded over the N GPUs
  bert_base    configs[3]: 768-d, 12-layer, seq-384 encoder, data-parallel over the N GPUs
  sustained    a >= 2 s back-to-back run of the headline step with its own clock sample
  parity_check (N > 1) the sharded result compared, bit for bit, with one index holding every row

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
FLOP_PER_CHUNK_BERT_BASE = 70.67e9  # 768-d, 12 layers, seq 384


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
    ap.add_argument("--no-extra", action="store_true", help="skip the 10M-row search, bert-base and sustained legs")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def workload_config(a, world: int) -> dict:
    """The SAME dict in both arms (the driver compares them)."""
    return {
        "workload": f"MiniLM-L6 (384-d, 6 layers, random-init) embed {a.batch} chunks x seq {a.seq} per GPU "
                    f"+ top-{a.top_k} over a {a.corpus_rows}x384 bf16 corpus (row-sharded over GPUs)",
        "chunks_per_gpu_per_step": a.batch, "seq_len": a.seq, "corpus_rows_global": a.corpus_rows,
        "parallelism": f"dp{world} embed, row-sharded search",
        "l2_policy": "inputs larger than L2: per step the encoder streams >1 GB of activations and the "
                     "search streams the 768 MB corpus shard; no explicit flush",
    }


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
# CPU arm: the reference's own code path on the host cores.
#   store   the REAL InMemoryVectorStore, byte-compiled from /root/reference into oracle/_ref by
#           oracle/make_ref.py (kind "reference"); the behavioural port in oracle/search_oracle.py when
#           oracle/_ref was not built (kind "port")
#   encode  sentence-transformers is un-vendored and not installable, so the encoder
#           is the fp32 torch restatement (oracle/encode_oracle.py), called the way the service calls
#           embed(): one chunk per call (embedding/app/service.py:384-393)
# Nothing is extrapolated: every query runs over a store that really holds corpus_rows rows.
# ------------------------------------------------------------------------------------------------
def reference_store_class():
    from oracle import ref_loader

    if ref_loader.load():
        return sys.modules["copilot_vectorstore.inmemory"].InMemoryVectorStore, "reference"
    from oracle.search_oracle import InMemoryPort

    return InMemoryPort, "port"


class CpuArm:
    def __init__(self, seq: int, top_k: int):
        import torch

        from oracle import encode_oracle as eo

        self.eo, self.seq, self.top_k = eo, seq, top_k
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        self.cfg = eo.MINILM_L6
        self.w = eo.make_weights(self.cfg, seed=0)
        self.ids, self.mask, _ = eo.synthetic_tokens(64, seq, self.cfg.vocab_size, seed=0)
        eo.encode_fast(self.cfg, self.w, self.ids[:1], self.mask[:1])     # warm-up
        self.store_cls, self.kind = reference_store_class()
        self.store, self.rows = None, 0

    @property
    def store_name(self) -> str:
        return ("the reference's InMemoryVectorStore (oracle/_ref, byte-compiled from its source)" if self.kind == "reference"
                else "the InMemoryVectorStore port (oracle/search_oracle.py)")

    def fill(self, rows: int) -> float:
        """A fresh store of `rows` x 384 fp32 vectors (bf16-representable values), inserted the way the service
        does: add_embeddings with <= 32 rows per call (embedding/app/service.py:284-292).  Returns seconds."""
        self.store, self.rows = self.store_cls(), rows
        rng = np.random.default_rng(1)
        t0 = time.perf_counter()
        for c0 in range(0, rows, 8192):
            blk = rng.standard_normal((min(8192, rows - c0), self.cfg.hidden), dtype=np.float32)
            blk = (blk.view(np.uint32) & np.uint32(0xFFFF0000)).view(np.float32)      # bf16-representable
            for j0 in range(0, len(blk), 32):
                sl = blk[j0:j0 + 32]
                self.store.add_embeddings([str(c0 + j0 + j) for j in range(len(sl))], sl.tolist(), [{} for _ in sl])
        return time.perf_counter() - t0

    def encode_one_per_call(self, n: int):
        eo = self.eo
        t0 = time.perf_counter()
        embs = [eo.encode_fast(self.cfg, self.w, self.ids[i % 64:i % 64 + 1], self.mask[i % 64:i % 64 + 1])[0].numpy()
                for i in range(n)]
        return (time.perf_counter() - t0) / n, embs

    def encode_batched(self, n_batches: int, bs: int = 32) -> float:
        eo = self.eo
        t0 = time.perf_counter()
        for b in range(n_batches):
            lo = (b * bs) % 32
            eo.encode_fast(self.cfg, self.w, self.ids[lo:lo + bs], self.mask[lo:lo + bs])
        return (time.perf_counter() - t0) / (n_batches * bs)

    def query(self, vec) -> float:
        t0 = time.perf_counter()
        self.store.query(vec.tolist(), top_k=self.top_k)
        return time.perf_counter() - t0

    def step(self):
        """One unit of the workload: embed one chunk (one call), search its embedding over the whole store."""
        t_enc, embs = self.encode_one_per_call(1)
        return t_enc, self.query(embs[0])


def run_reference(a) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm = CpuArm(a.seq, a.top_k)
    fill_s = arm.fill(a.corpus_rows)
    enc, qry = [], []
    for i in range(a.warmup + a.steps):
        t_enc, t_q = arm.step()
        if i >= a.warmup:
            enc.append(t_enc); qry.append(t_q)
    step_s = float(np.mean(enc) + np.mean(qry))
    v = 1.0 / step_s
    cb = {"value": v, "unit": UNIT, "cores": arm.cores, "kind": arm.kind,
          "sample": (f"{len(enc)} steps, each = 1 chunk x {a.seq} tokens through the fp32 torch encoder in one call "
                     f"({arm.cores} threads; sentence-transformers itself is not installable) + 1 query(top_k={a.top_k}) of "
                     f"{arm.store_name} over a store that holds all {a.corpus_rows} rows (filled by add_embeddings, "
                     f"32 rows per call, in {fill_s:.0f} s, untimed); nothing scaled"),
          "encode_ms_per_chunk": float(np.mean(enc)) * 1e3, "query_ms_per_chunk": float(np.mean(qry)) * 1e3,
          "fill_s": fill_s}
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": len(enc),
        "warmup": a.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": workload_config(a, a.gpus),
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def cpu_baseline_block(a) -> dict:
    """`cpu_baseline` of the CUDA arm's line: BASELINE.md section 4's table on the host cores (encode one-per-call and
    batch-32; real add_embeddings / query over the 10 000-row corpus of configs[0]; the vectorised numpy restatement
    "for fairness"), then the headline workload itself on a full-size store -- about 30 s of CPU work."""
    from oracle import search_oracle as so

    arm = CpuArm(a.seq, a.top_k)
    t_one, embs = arm.encode_one_per_call(48)
    t_b32 = arm.encode_batched(2, 32)
    fill10k = arm.fill(10_000)
    q10k = float(np.mean([arm.query(embs[i % len(embs)]) for i in range(50)]))
    rows10k = np.asarray([arm.store.get(str(i)).vector for i in range(10_000)], dtype=np.float32)
    t0 = time.perf_counter()
    so.topk_numpy(np.stack(embs[:32]), rows10k, a.top_k)
    qnp10k = (time.perf_counter() - t0) / 32
    fill_s = arm.fill(a.corpus_rows)
    tq = float(np.mean([arm.query(embs[i]) for i in range(3)]))
    v = 1.0 / (t_one + tq)
    return {
        "value": v, "unit": UNIT, "cores": arm.cores, "kind": arm.kind,
        "sample": (f"48 chunks x {a.seq} tokens encoded one per call (fp32 torch restatement, {arm.cores} threads): {t_one*1e3:.1f} ms/chunk; "
                   f"3 x query(top_k={a.top_k}) of {arm.store_name} over a store holding all {a.corpus_rows} rows: {tq*1e3:.0f} ms/query "
                   "(single thread by construction); value = 1 / (encode + query) per chunk, nothing scaled"),
        "encode_ms_per_chunk": t_one * 1e3, "query_ms_per_chunk": tq * 1e3, "fill_s_full_store": fill_s,
        "config1": {"what": "BASELINE.md section 4 rows on BASELINE.json configs[0] (10 000-row corpus), bounded sample",
                    "encode_one_per_call_chunks_per_s": 1.0 / t_one, "encode_batch32_chunks_per_s": 1.0 / t_b32,
                    "ingest_rows_per_s": 10_000 / fill10k, "query_qps_reference_loop": 1.0 / q10k,
                    "query_qps_vectorised_numpy": 1.0 / qnp10k},
    }


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
    pk = peaks()

    def corpus_chunks(rank_: int, r0_: int, r1_: int, seed_base: int):
        """The shard of `rank_`, 1 Mi rows at a time (any rank can regenerate any shard: same seed, same generator)."""
        gg = torch.Generator(device=dev).manual_seed(seed_base + rank_)
        for c0 in range(r0_, r1_, 1 << 20):
            n = min(1 << 20, r1_ - c0)
            yield torch.randn((n, 384), generator=gg, device=dev, dtype=torch.float32).bfloat16().contiguous()

    def build_index(rows_global: int, seed_base: int, ranks=None):
        lo = sh.shard_range(rows_global, world, rank)[0] if ranks is None else 0
        todo = [rank] if ranks is None else ranks
        n_rows = sum(sh.shard_range(rows_global, world, r)[1] - sh.shard_range(rows_global, world, r)[0] for r in todo)
        ix = vs.B200Index(384, device=local, master_dtype="bfloat16", capacity=n_rows)
        for r in todo:
            s0, s1 = sh.shard_range(rows_global, world, r)
            for blk in corpus_chunks(r, s0, s1, seed_base):
                ix.add(blk)
        return ix, lo

    # --- model + corpus (synthetic, seeded) ---
    with tempfile.TemporaryDirectory() as td:
        mio.write_random_model(td, "all-MiniLM-L6-v2", seed=0)
        prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device=f"cuda:{local}")
    index, r0 = build_index(a.corpus_rows, 1000)
    searcher = sh.ShardedSearch.over_index(index, r0)      # N > 1: hit records through NVLink peer memory when available

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

    def timed(fn, steps, warmup, sample_clocks=False, min_seconds=0.0):
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
        t_host = time.perf_counter()
        e0.record()
        done = 0
        while done < steps or (min_seconds and world == 1 and time.perf_counter() - t_host < min_seconds):
            fn()
            done += 1
            if min_seconds and done % 16 == 0:
                torch.cuda.synchronize()         # keep the host clock meaningful for the >= min_seconds loop
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        clocks = sampler.stop() if sampler else None
        return float(ms.item()), clocks, done

    ms_total, clocks, _ = timed(step_device, a.steps, a.warmup, sample_clocks=True)
    ms_step = ms_total / a.steps
    value = a.batch * world / (ms_step * 1e-3)
    ms_e2e_total, _, _ = timed(step_e2e, a.steps, max(3, a.warmup))
    e2e_value = a.batch * world / (ms_e2e_total / a.steps * 1e-3)

    # --- per-kernel timing for the roofline line (CUDA events around each launch, inside the library) ---
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
    timed_region_s = ms_total * 1e-3
    roofline = {}
    if g_ms > 0:
        ach = g_fl / (g_ms * 1e9)
        # a timed region under a second runs at boost clocks: the burst figure is the honest denominator there
        use_burst = timed_region_s < 1.0
        peak = pk["tf_burst"] if use_burst else pk["tf_sustained"]
        roofline = {"bound": "tensor", "kernel": "gemm_bf16_ast_kernel, gemm_ln_kernel, ffn_fused_kernel (the four linear maps of the encoder layer)",
                    "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                    "frac_burst": ach / pk["tf_burst"], "frac_sustained": ach / pk["tf_sustained"],
                    "peak_source": pk["source"] + (" (burst bf16 GEMM: the timed region is %.2f s)" % timed_region_s if use_burst
                                                   else " (sustained bf16 GEMM: kernels timed inside a long run)"),
                    "flop_per_launch": g_fl / g_n, "ms_per_launch": g_ms / g_n, "traffic": traffic_per_launch("gemm_bf16_ast_kernel", "gemm_ln_kernel", "ffn_fused_kernel"),
                    "profiled_ms_per_step": tot_ms / 3, "timed_ms_per_step": ms_step}
    sim = prof.get("sim_topk")
    if sim and sim["ms"] > 0:
        q_per_pass = a.batch * world
        tf = sim["flops"] / (sim["ms"] * 1e9)
        gbs = sim["bytes"] / (sim["ms"] * 1e6)
        roofline["search_step"] = {"queries_per_pass": q_per_pass, "rows_per_shard": len(index), "ms_per_launch": sim["ms"] / sim["launches"],
                                   "tflops": tf, "frac_tensor_burst": tf / pk["tf_burst"], "frac_tensor_sustained": tf / pk["tf_sustained"],
                                   "gbs": gbs, "frac_hbm": gbs / pk["hbm_gbs"], "bound": "tensor" if q_per_pass > 212 else "hbm"}
        # The step searches the embeddings a RANDOM-INIT encoder produces, and those are nearly parallel to one another:
        # the 32 queries a drain warp serves then pass their thresholds at the same corpus columns, which is the cheapest
        # case for the fused select.  So the same launch is also timed with independent random unit queries (what a
        # trained model's embeddings of different texts look like to this kernel) -- both ways back to back, 5 searches
        # each, because a kernel timed inside the step (above) starts with cold caches and is not comparable.
        try:
            def sim_ms_back_to_back(qq):
                for _ in range(2):
                    index.search(qq, a.top_k, id_base=r0)
                torch.cuda.synchronize()
                nat.profile(True)
                for _ in range(5):
                    index.search(qq, a.top_k, id_base=r0)
                torch.cuda.synchronize()
                p2 = nat.profile_read()
                nat.profile(False)
                return p2["sim_topk"]["ms"] / p2["sim_topk"]["launches"]

            gq = torch.Generator(device=dev).manual_seed(4242 + rank)
            q_ind = torch.randn((q_per_pass, 384), generator=gq, device=dev, dtype=torch.float32)
            q_ind = (q_ind / q_ind.norm(dim=1, keepdim=True)).contiguous()
            q_own = prov.embed_tokens(d_ids, d_lens).repeat(world, 1).contiguous()
            roofline["search_step"]["back_to_back_ms_per_launch"] = {
                "own_queries": sim_ms_back_to_back(q_own), "independent_queries": sim_ms_back_to_back(q_ind),
                "note": "own = the step's embeddings (random-init encoder: nearly parallel vectors); independent = random unit "
                        "vectors, the harder case for the fused select (every lane of a drain warp inserts at different columns)"}
        except Exception as exc:  # noqa: BLE001
            roofline["search_step"]["back_to_back_ms_per_launch"] = {"error": str(exc)[:200]}
    roofline["whole_encode_tflops"] = FLOP_PER_CHUNK * a.batch / (ms_step * 1e-3) / 1e12
    chk, redo = index.certificate_stats()
    roofline["certificate"] = {"queries_checked": chk, "queries_rescored_exhaustively": redo}

    # launches of OUR kernels per step: embed+LN, 6 x (QKV GEMM, attention, out-proj GEMM + LayerNorm, fused
    # feed-forward block + LayerNorm), pool, prep_queries, sim, merge+rescore, the two (predicated) exhaustive passes, K10 merge
    launches_step = 1 + 6 * 4 + 1 + 5 + (1 if world > 1 else 0)
    gpu_launches = launches_step * a.steps

    # --- N > 1: the sharded result against ONE index holding every row, outside the timed region ---
    if world > 1:
        emb = prov.embed_tokens(d_ids, d_lens)
        ids_s, sc_s = searcher.search_local_queries(emb, a.top_k)
        q_all = torch.empty((world * a.batch, 384), dtype=emb.dtype, device=dev)
        dist.all_gather_into_tensor(q_all, emb.contiguous())
        ids_all = torch.empty((world * a.batch, a.top_k), dtype=torch.int64, device=dev)
        sc_all = torch.empty((world * a.batch, a.top_k), dtype=torch.float32, device=dev)
        dist.all_gather_into_tensor(ids_all, ids_s.contiguous())
        dist.all_gather_into_tensor(sc_all, sc_s.contiguous())
        if rank == 0:
            full, _ = build_index(a.corpus_rows, 1000, ranks=list(range(world)))
            ids_1, sc_1 = full.search(q_all, a.top_k)
            ok = bool(torch.equal(ids_1, ids_all) and torch.equal(sc_1.view(torch.int32), sc_all.view(torch.int32)))
            full.close()
            roofline["parity_check"] = {"sharded_equals_single": ok, "queries": world * a.batch, "rows": a.corpus_rows,
                                        "compared": "ids and fp32 score bits of every rank's search_local_queries result"}
        dist.barrier()

    extra_ok = not a.no_extra
    # --- sustained: the headline step back to back for >= 2 s, own clock sample (N = 1) ---
    if extra_ok and world == 1:
        ms_long, clk_long, n_long = timed(step_device, 50, 0, sample_clocks=True, min_seconds=2.0)
        ach_long = FLOP_PER_CHUNK * a.batch * n_long / (ms_long * 1e-3) / 1e12
        roofline["sustained"] = {"seconds": ms_long * 1e-3, "steps": n_long, "chunks_per_s": a.batch * n_long / (ms_long * 1e-3),
                                 "whole_encode_tflops": ach_long, "frac_sustained": ach_long / pk["tf_sustained"],
                                 "sm_mhz_median": clk_long["sm_mhz"], "reasons": clk_long["reasons"]}

    # --- BASELINE.json's metric, second half, and configs[2]: top-10 over 10 M x 384 bf16, row-sharded over N GPUs ---
    if extra_ok:
        try:
            del searcher
            index.close()
            n_big = 10_000_000
            big, b0 = build_index(n_big, 5000)
            bs = sh.ShardedSearch.over_index(big, b0)
            bs_nccl = sh.ShardedSearch.over_index(big, b0, exchange="nccl") if world > 1 and bs.exchange == "peer" else None
            rows_rank = len(big)
            s10 = {"rows_global": n_big, "rows_per_gpu": rows_rank, "algorithmic_bytes_per_pass_per_gpu": rows_rank * 768,
                   "traffic": {"single_tile_kernel_128_queries": traffic_per_launch("sim_topk_kernel<256, 20, 1>"),
                               "cta_pair_kernel_4096_queries": traffic_per_launch("sim_topk_kernel<256, 20, 2>"),
                               "note": "dram bytes per launch over ALL 10 M rows on one GPU (profiles/r02_ncu_search.md); algorithmic = 7.68e9"},
                   "exchange": bs.exchange,
                   "api": "ShardedSearch.search_replicated: b200_index_search_exchange (hit records stored into NVLink peer memory) -> "
                          "b200_merge_hits_wait" if bs.exchange == "peer" else
                          "ShardedSearch.search_replicated: b200_index_search_packed -> all-gather -> b200_merge_hits"}
            gq = torch.Generator(device=dev).manual_seed(77)
            for nq in (1, 128, 4096):      # ascending: the HBM-bound cases are not measured under the power cap the tensor-bound one leaves behind
                q = torch.randn((nq, 384), generator=gq, device=dev, dtype=torch.float32).bfloat16().contiguous()
                ms, _, n_it = timed(lambda: bs.search_replicated(q, a.top_k), 5, 3)
                ms /= n_it
                t_hbm = rows_rank * 768 / (pk["hbm_gbs"] * 1e9) * 1e3
                t_b = nq * rows_rank * 768.0 / (pk["tf_burst"] * 1e12) * 1e3
                t_s = nq * rows_rank * 768.0 / (pk["tf_sustained"] * 1e12) * 1e3
                bound = "tensor" if t_s > t_hbm else "hbm"
                s10[f"q{nq}"] = {"ms_per_batch": ms, "qps": nq / (ms * 1e-3), "bound": bound,
                                 "frac_burst": max(t_hbm, t_b) / ms, "frac_sustained": max(t_hbm, t_s) / ms,
                                 "gbs_per_gpu": rows_rank * 768 / (ms * 1e-3) / 1e9,
                                 "tflops_per_gpu": nq * rows_rank * 768.0 / (ms * 1e-3) / 1e12}
                if bs_nccl is not None:
                    ms2, _, n2 = timed(lambda: bs_nccl.search_replicated(q, a.top_k), 5, 3)
                    s10[f"q{nq}"]["ms_per_batch_nccl_all_gather"] = ms2 / n2
            chk, redo = big.certificate_stats()
            s10["certificate"] = {"queries_checked": chk, "queries_rescored_exhaustively": redo}
            roofline["search_10M"] = s10
            big.close()
        except Exception as exc:  # noqa: BLE001 - an extra leg must never sink the headline line
            roofline["search_10M"] = {"error": str(exc)[:300]}
        # configs[3]: 768-d, 12 layers, 12 heads x 64, seq 384, data-parallel (every rank its own batch of 128)
        try:
            with tempfile.TemporaryDirectory() as td:
                mio.write_random_model(td, "bert-base-class", seed=0)
                big_prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device=f"cuda:{local}")
            bb, ss = 128, 384
            gb = torch.Generator(device=dev).manual_seed(5 + rank)
            ids_b = torch.randint(1000, 30522, (bb, ss), generator=gb, device=dev, dtype=torch.int32)
            lens_b = torch.full((bb,), ss, device=dev, dtype=torch.int32)
            ms, _, n_it = timed(lambda: big_prov.embed_tokens(ids_b, lens_b), 5, 3)
            ms /= n_it
            tfl = bb * FLOP_PER_CHUNK_BERT_BASE / (ms * 1e-3) / 1e12
            roofline["bert_base"] = {"what": "768-d, 12 layers, 12 heads x 64, seq 384, batch 128 per GPU, data-parallel",
                                     "ms_per_batch": ms, "chunks_per_s": world * bb / (ms * 1e-3), "tflops_per_gpu": tfl,
                                     "frac_burst": tfl / pk["tf_burst"], "frac_sustained": tfl / pk["tf_sustained"]}
            big_prov.close()
        except Exception as exc:  # noqa: BLE001
            roofline["bert_base"] = {"error": str(exc)[:300]}

    # --- what an UNPATCHED service gets: embed(text) one text per call, query(list, 10) per vector, 10 000-row corpus ---
    unpatched = None
    if extra_ok and rank == 0 and world == 1:
        try:
            unpatched = unpatched_service_leg(prov, vs, local)
        except Exception as exc:  # noqa: BLE001
            unpatched = {"error": str(exc)[:300]}

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        cpu = cpu_baseline_block(a)

    if rank == 0:
        h2d = a.batch * a.seq * 4 + a.batch * 4 + a.batch * 384 * 4
        d2h = a.batch * 384 * 4 + a.batch * a.top_k * 12
        e2e = {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "api": "provider.embed_tokens(host ids) -> host embeddings -> index.search(host queries) -> host ids/scores"}
        if unpatched is not None:
            e2e["unpatched_service"] = unpatched
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": workload_config(a, world),
            "e2e": e2e,
            "gpu_launches": gpu_launches,
            "clocks": clocks,
            "roofline": roofline,
            "kernels": kernels,
            "cpu_baseline": cpu,
            "peaks": pk,
        }))
    if world > 1:
        dist.destroy_process_group()


def unpatched_service_leg(prov, vs, local: int) -> dict:
    """The drop-in's first-day behaviour, before any service is patched to batch: `provider.embed(text)` once per chunk
    (embedding/app/service.py:384-393) and `store.query(list, 10)` once per vector (reporting/app/service.py:821-828)
    over the 10 000-row corpus of BASELINE.json configs[0], through the reference-shaped API with Python lists."""
    import torch

    words = ["ba", "be", "bi", "bo", "bu", "ca", "ce", "ci", "co", "cu", "da", "de", "di", "do", "du"]
    rng = np.random.default_rng(3)
    texts = [" ".join(words[j] for j in rng.integers(0, len(words), 250)) for _ in range(64)]
    store = vs.B200VectorStore(device=f"cuda:{local}")
    rows = rng.standard_normal((10_000, 384), dtype=np.float32)
    store.add_embeddings_array([f"c{i}" for i in range(10_000)], rows)
    for t in texts[:4]:
        store.query(prov.embed(t), top_k=10)
    torch.cuda.synchronize()
    t_tok = t_emb = t_q = 0.0
    vecs = []
    for t in texts:
        t0 = time.perf_counter()
        prov.tokenizer.encode_arrays([t], prov.max_seq_length)
        t1 = time.perf_counter()
        v = prov.embed(t)
        t2 = time.perf_counter()
        vecs.append(v)
        t_tok += t1 - t0
        t_emb += t2 - t1
    for v in vecs:
        t0 = time.perf_counter()
        store.query(v, top_k=10)
        t_q += time.perf_counter() - t0
    n = len(texts)
    store.close()
    return {"corpus_rows": 10_000, "calls": n, "embed_calls_per_s": n / t_emb, "query_calls_per_s": n / t_q,
            "embed_ms_per_call": t_emb / n * 1e3, "of_which_tokenise_ms": t_tok / n * 1e3,
            "query_ms_per_call": t_q / n * 1e3,
            "note": "one ~256-word-piece text per embed() call; query() returns 10 SearchResult objects, each carrying its "
                    "384-float vector as a list"}


def traffic_per_launch(*prefixes: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch (mean over the kernels whose name starts with one of
    `prefixes`), from the committed `ncu --set full` capture summarised into profiles/traffic.json; None when absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)["dram_bytes_per_launch"]
        g = [v for k, v in t.items() if any(k.startswith(p) for p in prefixes)]
        return sum(g) / len(g) if g else None
    except Exception:  # noqa: BLE001
        return None


def main():
    a = parse()
    # The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner on fd 1 when
    # NCCL_DEBUG is set in the environment), so for the duration of the run fd 1 points at stderr and the line goes to
    # the saved descriptor at the end.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w", buffering=1)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
