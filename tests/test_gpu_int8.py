"""GPU parity of the llm.int8 path (Linear8bitLt) against the oracle's restatement of LLM.int8().
bitsandbytes is un-vendored/unpinned in the reference and absent here: parity is "unpinned" (DESIGN.md section 1)."""
import numpy as np
import pytest
import torch

from oracle import quant as OQ

pytestmark = pytest.mark.gpu
DEV = "cuda"


def make(N, K, W, bias=None, dtype=torch.bfloat16):
    from lit_llama.quantization import Linear8bitLt

    lin = Linear8bitLt(K, N, bias=bias is not None, device=DEV, dtype=dtype)
    sd = {"weight": torch.from_numpy(W)}
    if bias is not None:
        sd["bias"] = torch.from_numpy(bias)
    lin.load_state_dict(sd)
    return lin


@pytest.mark.parametrize("N,K", [(64, 128), (4096, 4096), (1000, 384)])
def test_int8_weight_quantization_bit_exact(N, K):
    r = np.random.RandomState(N + K)
    W = (r.randn(N, K) * 0.03).astype(np.float32)
    W[3] = 0.0                                            # an all-zero row
    lin = make(N, K, W)
    cb, scb = OQ.int8_quantize_weight(W)
    assert lin.weight.dtype == torch.int8 and lin.weight.CB.data_ptr() == lin.weight.data_ptr()
    assert np.array_equal(lin.weight.CB.cpu().numpy(), cb)
    assert np.array_equal(lin.weight.SCB.cpu().numpy(), scb)


@pytest.mark.parametrize("M,N,K,outliers", [(1, 256, 512, 0), (4, 512, 1024, 3), (32, 4096, 4096, 5), (32, 1024, 11008, 11),
                                            (7, 384, 256, 256), (100, 320, 640, 2), (32, 1024, 4096, 700)])
@pytest.mark.parametrize("dt", ["bf16", "fp16"])
def test_int8_linear_vs_oracle(M, N, K, outliers, dt):
    """int8 part is exact (int32); the fp16 scaling / outlier part differs only by fp32 summation order and
    one fp16 rounding: <= 2 fp16 ulp of the result magnitude, then the cast to the activation dtype."""
    r = np.random.RandomState(M * 7 + N + K + outliers)
    W = (r.randn(N, K) * 0.03).astype(np.float32)
    bias = (r.randn(N) * 0.1).astype(np.float32)
    x = r.randn(M, K).astype(np.float32)
    cols = r.choice(K, size=outliers, replace=False)
    x[:, cols] *= 20.0                                    # features above the 6.0 threshold -> fp16 outlier path
    tdt = torch.bfloat16 if dt == "bf16" else torch.float16
    xr = OQ.bf16_round(x) if dt == "bf16" else OQ.fp16_round(x)
    lin = make(N, K, W, bias, tdt)
    y = lin(torch.from_numpy(xr).to(DEV, tdt)).float().cpu().numpy()
    cb, scb = OQ.int8_quantize_weight(W)
    b = OQ.bf16_round(bias) if dt == "bf16" else OQ.fp16_round(bias)
    ref = OQ.int8_linear_forward(xr, cb, scb, b, out_dtype=dt)
    ulp = 2.0**-7 if dt == "bf16" else 2.0**-10
    tol = np.abs(ref) * ulp + np.abs(ref).max() * 2.0**-10
    assert (np.abs(y - ref) <= tol).all(), float((np.abs(y - ref) / tol).max())
    # against the un-quantized fp32 linear: the algorithm's own error stays small
    full = xr @ W.T + bias
    assert np.abs(y - full).max() <= 0.05 * np.abs(full).max()
    # deterministic run to run (integer split-K atomics are order-independent)
    y2 = lin(torch.from_numpy(xr).to(DEV, tdt)).float().cpu().numpy()
    assert np.array_equal(y, y2)


def test_int8_mode_builds_and_runs_tiny_model(golden):
    """quantization('llm.int8') swaps nn.Linear for Linear8bitLt; an ordinary fp checkpoint loads and is
    re-quantized (ref quantization.py:50-65); logits stay close to the fp32 model's."""
    from lit_llama import LLaMA, LLaMAConfig
    from lit_llama.quantization import Linear8bitLt
    from lit_llama.utils import EmptyInitOnDevice
    from tests.util_tiny import TINY, tiny_weights

    with EmptyInitOnDevice(device=DEV, dtype=torch.bfloat16, quantization_mode="llm.int8"):
        m = LLaMA(LLaMAConfig(**TINY))
    assert isinstance(m.lm_head, Linear8bitLt) and isinstance(m.transformer.h[0].mlp.c_fc1, Linear8bitLt)
    m.load_state_dict({k: torch.from_numpy(v) for k, v in tiny_weights().items()})
    m.eval()
    g = golden["model_tiny"]
    idx = torch.from_numpy(g["idx"]).to(DEV)
    with torch.no_grad():
        logits = m(idx).float().cpu().numpy()
    ref = g["fp32.logits"]
    assert np.abs(logits - ref).max() <= 0.06 * np.abs(ref).max()
    with torch.no_grad():
        m.reset_cache()
        a = m(idx[:4, :10], 16, torch.arange(10, device=DEV))[:, -1].float().cpu().numpy()
    assert np.abs(a - ref[:4, 9]).max() <= 0.06 * np.abs(ref).max()


def test_int8_errors():
    from lit_llama.quantization import Linear8bitLt
    from lit_llama.utils import EmptyInitOnDevice

    with pytest.raises(RuntimeError, match="CUDA"):
        Linear8bitLt(64, 32)
    with pytest.raises(ValueError, match="only supported on the GPU"):
        EmptyInitOnDevice(device="cpu", quantization_mode="llm.int8")
