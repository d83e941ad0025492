"""bert-base-class (K = 768 GEMMs): where the MMA issuer of CTA pair 0 spends its clocks, and the encode time with parts
of the epilogue switched off (library built with EXTRA=-DB200_GEMM_TIMING; results are wrong under flags 1..4)."""
import ctypes, importlib, os, sys, tempfile
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
P = "copilot-for-consensus_b200"
mio = importlib.import_module(P + ".embedding.model_io")
prov_mod = importlib.import_module(P + ".embedding.provider")
nat = importlib.import_module(P + "._native")
with tempfile.TemporaryDirectory() as td:
    mio.write_random_model(td, "bert-base-class", seed=0)
    prov = prov_mod.B200SentenceTransformerProvider(model_name=td, device="cuda:0")
B, S = 128, 384
ids = torch.randint(1000, 30522, (B, S), dtype=torch.int32, device="cuda")
lens = torch.full((B,), S, dtype=torch.int32, device="cuda")
L = nat.lib()
names = {0: "full epilogue", 1: "accumulators drained, nothing computed or stored", 2: "no TMA store", 3: "no fence, no store",
         4: "stores land in the first 256 rows (L2)"}
for defer in (1, 0):
    nat.set_option("defer_ln", defer)
    for flag in (0, 1, 2, 3, 4):
        buf = (ctypes.c_ulonglong * 8)()
        buf[7] = flag
        L.b200_debug_gemm_timing(buf)
        for _ in range(2): prov.embed_tokens(ids, lens)
        buf[7] = flag
        L.b200_debug_gemm_timing(buf)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): prov.embed_tokens(ids, lens)
        e1.record(); torch.cuda.synchronize()
        buf[7] = 0
        L.b200_debug_gemm_timing(buf)
        n = max(1, buf[3])
        print("defer_ln=%d flag %d (%s): encode %.3f ms | per tile: wait accumulator %.0f  wait operands %.0f  issue %.0f  total %.0f clk"
              " | first column tiles: %d of %d tiles, wait operands %.0f per tile"
              % (defer, flag, names[flag], e0.elapsed_time(e1) / 5, buf[0] / n, buf[1] / n, buf[2] / n, buf[4] / n,
                 buf[6], buf[3], buf[5] / max(1, buf[6])), flush=True)
