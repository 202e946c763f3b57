"""CPU-only: the C-ABI library loads and exports every symbol include/llb200.h declares; the
class-swap boundary and module surface behave like the reference's (no compute calls)."""
import re
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent


def test_library_exports_every_declared_symbol():
    from lit_llama import _native

    header = (ROOT / "include" / "llb200.h").read_text()
    declared = set(re.findall(r"LL_API\s+[\w\s\*]+?\b(ll_\w+)\s*\(", header))
    assert len(declared) >= 18
    assert declared == set(_native.EXPORTS), declared ^ set(_native.EXPORTS)
    for name in declared:
        assert getattr(_native.lib, name) is not None
    assert _native._fn["ll_abi_version"]() == _native.ABI_VERSION


def test_quantization_ctx_swaps_and_restores():
    from lit_llama import LLaMA, LLaMAConfig
    from lit_llama.quantization import ColBlockQuantizedLinear
    from lit_llama.utils import EmptyInitOnDevice, quantization

    cfg = LLaMAConfig(block_size=64, vocab_size=100, n_layer=1, n_head=4, n_embd=128)
    orig = torch.nn.Linear
    with quantization("gptq.int4"):
        m = LLaMA(cfg)
    assert torch.nn.Linear is orig
    lin = m.transformer.h[0].mlp.c_proj
    assert isinstance(lin, ColBlockQuantizedLinear) and lin.bits == 4
    assert lin.quant_weight.shape == (128, 256) and lin.quant_weight.stride() == (1, 128)
    assert lin.quant_weight.dtype == torch.uint8 and lin.scales.shape == (128, 1) and lin.bias is None
    keys = set(m.state_dict())
    assert "lm_head.quant_weight" in keys and "transformer.h.0.attn.c_attn.zeros" in keys
    assert "transformer.wte.weight" in keys and "transformer.h.0.rms_1.scale" in keys
    assert not any(k.endswith(".bias") for k in keys)
    with quantization("gptq.int8"):
        assert LLaMA(cfg).lm_head.bits == 8
    with quantization(None):
        assert type(LLaMA(cfg).lm_head) is orig
    with pytest.raises(ValueError, match="Unknown quantization mode"):
        with quantization("nope"):
            pass
    with pytest.raises(RuntimeError, match="unknown quantization mode"):
        EmptyInitOnDevice(device="cpu", quantization_mode="nope")
    with pytest.raises(ValueError, match="only supported on the GPU"):
        EmptyInitOnDevice(device=torch.device("cpu"), quantization_mode="llm.int8")
    with pytest.raises(ZeroDivisionError):      # restored even when the body raises
        with quantization("gptq.int4"):
            1 / 0
    assert torch.nn.Linear is orig
    with EmptyInitOnDevice(device="cpu", dtype=torch.bfloat16, quantization_mode="gptq.int4"):
        m2 = LLaMA(cfg)
    assert torch.nn.Linear is orig
    assert m2.lm_head.scales.dtype == torch.bfloat16 and m2.transformer.wte.weight.dtype == torch.bfloat16


def test_grouped_checkpoint_reshapes_per_row_module(tmp_path):
    from lit_llama.quantization import ColBlockQuantizedLinear
    from lit_llama.utils import lazy_load, llama_model_lookup, find_multiple

    src = ColBlockQuantizedLinear(256, 64, False, bits=4, tile_cols=128)
    src.quant_weight.random_(0, 255); src.scales.normal_(); src.zeros.normal_()
    p = tmp_path / "ck.pth"
    torch.save({**{"l." + k: v for k, v in src.state_dict().items()}, "transformer.wte.weight": torch.zeros(8, 4096)}, p)
    with lazy_load(p) as ck:
        assert llama_model_lookup(ck) == "7B"
        assert ck["l.quant_weight"].stride() == (1, 64)       # torch.save keeps the column-major strides
        dst = ColBlockQuantizedLinear(256, 64, False, bits=4, tile_cols=-1)
        dst.load_state_dict({k[2:]: v for k, v in ck.items() if k.startswith("l.")})
    assert dst.tile_cols == 128 and dst.scales.shape == (64, 2)
    assert torch.equal(dst.quant_weight, src.quant_weight) and dst.quant_weight.stride() == (1, 64)
    assert find_multiple(100, 64) == 128 and find_multiple(128, 64) == 128


def test_cpu_compute_fails_loudly():
    from lit_llama import LLaMA, LLaMAConfig

    m = LLaMA(LLaMAConfig(block_size=16, vocab_size=16, n_layer=1, n_head=2, n_embd=8))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.zeros(1, 4, dtype=torch.long))


# ----------------------------------------------------------------------------- bench.py contract (reference arm, CPU)
def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` runs on the host cores only and prints ONE JSON line with the keys the driver reads."""
    import json
    import subprocess
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    import os

    env = dict(os.environ, LLB200_BENCH_MODEL="tiny", LLB200_BENCH_PROMPT="16", LLB200_BENCH_NEW="8", OMP_NUM_THREADS="1")
    out = subprocess.run([sys.executable, str(root / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600, cwd=root, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "decode_tokens_per_sec" and d["unit"] == "tokens/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0 and d["e2e"]["value"] == d["value"]
    assert "workload" in d["config"]
    # all the host threads, whatever OMP_NUM_THREADS torchrun exported; and none of this repo's libraries in that process
    assert d["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
