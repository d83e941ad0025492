"""Host-side logic that needs no GPU: factories' error contracts, config stand-ins, model-directory
I/O, and that constructing the CUDA-backed classes without a GPU fails loudly."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import pkg


def test_vector_store_factory_errors_need_no_gpu():
    vs, cfgm = pkg("vectorstore"), pkg("config")
    with pytest.raises(ValueError, match="vector_store config is required"):
        vs.create_vector_store(None)
    with pytest.raises(ValueError, match="Unknown vector_store driver: nope"):
        vs.create_vector_store(cfgm.AdapterConfig_VectorStore(vector_store_type="nope", driver=None))
    with pytest.raises(TypeError, match="Expected inmemory config, got object"):
        vs.create_vector_store(cfgm.AdapterConfig_VectorStore(vector_store_type="INMEMORY", driver=object()))


def test_embedding_factory_errors_need_no_gpu():
    emb, cfgm = pkg("embedding"), pkg("config")
    with pytest.raises(ValueError, match="embedding_backend config is required"):
        emb.create_embedding_provider(None)
    with pytest.raises(ValueError, match="Unknown embedding backend driver: nope"):
        emb.create_embedding_provider(cfgm.AdapterConfig_EmbeddingBackend(embedding_backend_type="nope", driver=None))
    with pytest.raises(TypeError, match="driver_config must be DriverConfig_EmbeddingBackend_Sentencetransformers"):
        emb.create_embedding_provider(cfgm.AdapterConfig_EmbeddingBackend(embedding_backend_type="sentencetransformers", driver=object()))


def test_config_standins_have_the_reference_defaults():
    cfgm = pkg("config")
    d = cfgm.DriverConfig_EmbeddingBackend_Sentencetransformers()
    assert (d.model_name, d.device, d.cache_dir) == ("all-MiniLM-L6-v2", "cpu", None)   # embedding_backend.py:64-73
    assert cfgm.DriverConfig_VectorStore_Inmemory() == cfgm.DriverConfig_VectorStore_Inmemory()


def test_model_dir_roundtrip(tmp_path):
    mio = pkg("embedding.model_io")
    cfg = mio.EncoderConfig(vocab_size=1500, hidden=128, layers=2, heads=4, intermediate=256, max_pos=64, max_seq_length=48)
    d = mio.write_random_model(str(tmp_path / "m"), cfg, seed=3)
    assert mio.resolve_model_dir(d, None) == d
    assert mio.resolve_model_dir("m", str(tmp_path)) == os.path.join(str(tmp_path), "m")
    assert mio.resolve_model_dir("missing-model", str(tmp_path)) is None
    cfg2, sd, vocab, lower = mio.load_model_dir(d)
    assert cfg2 == cfg and lower is True and len(vocab) == 1500
    want = mio.random_weights(cfg, seed=3)
    assert sd.keys() == want.keys() and all(torch.equal(sd[k], want[k]) for k in want)
    hf = json.load(open(os.path.join(d, "config.json")))
    assert hf["hidden_size"] == 128 and hf["hidden_act"] == "gelu"
    with pytest.raises(ValueError, match="hidden_act"):
        mio.config_from_hf({**hf, "hidden_act": "relu"})


def test_cuda_backed_classes_fail_loudly_without_gpu(tmp_path):
    if torch.cuda.is_available():
        pytest.skip("this check is for the GPU-less build container")
    mio, prov = pkg("embedding.model_io"), pkg("embedding.provider")
    d = mio.write_random_model(str(tmp_path / "m"), mio.EncoderConfig(vocab_size=1500, hidden=128, layers=1, heads=4,
                                                                     intermediate=256, max_pos=64, max_seq_length=48))
    with pytest.raises(RuntimeError, match="no CPU fallback|CUDA"):
        prov.B200SentenceTransformerProvider(model_name=d, device="cuda:0")
    with pytest.raises(ValueError, match="CUDA device"):
        prov.B200SentenceTransformerProvider(model_name=d, device="cpu")
    with pytest.raises(RuntimeError, match="not found locally"):
        prov.B200SentenceTransformerProvider(model_name="all-MiniLM-L6-v2", device="cuda:0", cache_dir=str(tmp_path))
    with pytest.raises(RuntimeError, match="no CPU fallback|CUDA"):
        pkg("vectorstore").B200Index(384, device="cuda:0")


def test_bench_reference_arm_prints_the_contract_line():
    import subprocess
    import sys

    from conftest import ROOT

    from oracle import ref_loader

    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1",
                          "--corpus-rows", "3000"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-500:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "chunks/s" and line["value"] > 0 and line["steps"] == 2
    # the real InMemoryVectorStore when oracle/_ref is built (python oracle/make_ref.py), the port otherwise; never scaled
    assert line["cpu_baseline"]["kind"] == ("reference" if ref_loader.available() else "port")
    assert "nothing scaled" in line["cpu_baseline"]["sample"] and line["e2e"]["h2d_bytes_per_step"] == 0
    assert abs(line["ms_per_step"] - 1e3 / line["value"]) < 1e-6
    assert line["config"]["corpus_rows_global"] == 3000 and line["config"]["parallelism"].startswith("dp1")


def test_model_dir_refuses_other_module_stacks_and_architectures(tmp_path):
    """SentenceTransformer.encode runs what modules.json lists; only Transformer -> mean Pooling -> Normalize on a
    BERT-architecture checkpoint is implemented, anything else must fail at load instead of embedding differently."""
    mio = pkg("embedding.model_io")
    cfg = mio.EncoderConfig(vocab_size=1500, hidden=128, layers=1, heads=4, intermediate=256, max_pos=64, max_seq_length=48)
    d = mio.write_random_model(str(tmp_path / "m"), cfg, seed=1)
    mio.load_model_dir(d)                                              # the stack write_random_model lays down is accepted
    pc = os.path.join(d, "1_Pooling", "config.json")
    pool = json.load(open(pc))
    pool.update(pooling_mode_cls_token=True, pooling_mode_mean_tokens=False)
    json.dump(pool, open(pc, "w"))
    with pytest.raises(ValueError, match="unsupported pooling"):
        mio.load_model_dir(d)
    pool.update(pooling_mode_cls_token=False, pooling_mode_mean_tokens=True)
    json.dump(pool, open(pc, "w"))
    mods = json.load(open(os.path.join(d, "modules.json")))
    json.dump(mods[:2], open(os.path.join(d, "modules.json"), "w"))    # no Normalize stage (paraphrase-MiniLM-* style)
    with pytest.raises(ValueError, match="module stack"):
        mio.load_model_dir(d)
    json.dump(mods, open(os.path.join(d, "modules.json"), "w"))
    hf = json.load(open(os.path.join(d, "config.json")))
    hf["model_type"] = "mpnet"
    json.dump(hf, open(os.path.join(d, "config.json"), "w"))
    with pytest.raises(ValueError, match="model_type 'mpnet'"):
        mio.load_model_dir(d)
    assert "all-mpnet-base-v2" not in mio.KNOWN_ARCHS and mio.KNOWN_ARCHS["bert-base-class"].hidden == 768


def test_factory_with_the_reference_packages_installed_behaves_like_the_reference_factory():
    """With copilot_config / copilot_vectorstore importable the drop-in factory validates the driver config through the
    reference's own schema validator (adapter_factory.py:61-67) and then fails / succeeds exactly where the reference's
    factory does; the store it builds IS a copilot_vectorstore.VectorStore.  Needs /root/reference (build container)."""
    import subprocess
    import sys

    from conftest import ROOT

    ref = "/root/reference/adapters"
    if not os.path.isdir(ref):
        pytest.skip("the reference tree is not on this machine")
    code = r"""
import importlib, sys
from copilot_config.generated.adapters.vector_store import (AdapterConfig_VectorStore, DriverConfig_VectorStore_Inmemory,
                                                            DriverConfig_VectorStore_Faiss)
import copilot_vectorstore.factory as rf
from copilot_vectorstore.interface import VectorStore
vs = importlib.import_module("copilot-for-consensus_b200.vectorstore")
def outcome(f, cfg):
    try:
        return ("ok", type(f(cfg)).__name__)
    except Exception as e:
        return (type(e).__name__, str(e))
for vtype, drv in [("inmemory", object()), ("inmemory", DriverConfig_VectorStore_Faiss(dimension=8)), ("nope", None)]:
    cfg = AdapterConfig_VectorStore(vector_store_type=vtype, driver=drv)
    a, b = outcome(rf.create_vector_store, cfg), outcome(vs.create_vector_store, cfg)
    assert a == b, (vtype, a, b)
st = vs.create_vector_store(AdapterConfig_VectorStore(vector_store_type="inmemory", driver=DriverConfig_VectorStore_Inmemory()))
assert isinstance(st, VectorStore) and type(st).__name__ == "B200VectorStore"
# the faiss slot is only taken over on request, and only for a flat index without persistence
fcfg = AdapterConfig_VectorStore(vector_store_type="faiss", driver=DriverConfig_VectorStore_Faiss(dimension=8, index_type="flat"))
vs.factory.configure(serve_faiss_flat=True)
st = vs.create_vector_store(fcfg)
assert type(st).__name__ == "B200VectorStore" and st.metric == "l2" and st.dimension == 8
vs.factory.configure(serve_faiss_flat=False)
print("factory parity ok")
"""
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([ref + "/copilot_config", ref + "/copilot_vectorstore", ROOT]))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, env=env, cwd="/tmp")
    assert r.returncode == 0 and "factory parity ok" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]


def test_host_out_is_a_fresh_writable_array_without_cuda():
    """`_native.host_out` hands the host layer its result arrays: page-locked when torch + CUDA can provide them, plain
    numpy otherwise (this box) -- either way a fresh, writable, C-contiguous array of the requested shape and dtype."""
    nat = pkg("_native")
    a = nat.host_out((3, 5), np.float32)
    b = nat.host_out((3, 5), np.float32)
    assert a.shape == (3, 5) and a.dtype == np.float32 and a.flags.c_contiguous and a.flags.writeable
    a[:] = 1.0
    b[:] = 2.0
    assert float(a.sum()) == 15.0 and float(b.sum()) == 30.0          # two calls never alias
    assert nat.host_out((0, 4), np.int64).shape == (0, 4)
    assert nat.host_out((2,), np.int64).dtype == np.int64
