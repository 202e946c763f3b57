"""Lazy checkpoint streaming (ref lit_llama/utils.py:198-376).  CPU: the archive index and the placeholders
against ``torch.load`` on the same file.  GPU: the disk -> pinned -> HBM pipeline with the repack inside the load."""
import numpy as np
import pytest
import torch

from tests.util_tiny import TINY


def _checkpoint(tmp_path):
    """A state dict with everything the loader must cope with: fp32 / bf16 / uint8 / int64, a Parameter, a column-major
    buffer (the GPTQ quant_weight strides), a zero-size tensor, a scalar, two views of one storage."""
    g = torch.Generator().manual_seed(3)
    base = torch.randn(6, 10, generator=g)
    sd = {
        "a.weight": torch.nn.Parameter(torch.randn(5, 7, generator=g)),
        "a.bf16": torch.randn(33, 9, generator=g).bfloat16(),
        "q.quant_weight": torch.randint(0, 255, (12, 16), generator=g, dtype=torch.uint8).t().contiguous().t(),
        "q.colmajor": torch.randint(0, 255, (16, 12), generator=g, dtype=torch.uint8).t(),
        "idx": torch.arange(11, dtype=torch.int64),
        "empty": torch.zeros(0, 4),
        "scalar": torch.tensor(2.5),
        "view.lo": base[:3],
        "view.hi": base[3:, 2:7],
        "flag": torch.tensor([True, False, True]),
    }
    p = tmp_path / "ck.pth"
    torch.save(sd, p)
    return p, sd


def test_lazy_load_matches_torch_load(tmp_path):
    from lit_llama.utils import NotYetLoadedTensor, lazy_load

    p, sd = _checkpoint(tmp_path)
    want = torch.load(p, map_location="cpu", weights_only=True)
    with lazy_load(p) as ck:
        assert list(ck.keys()) == list(want.keys())
        for k, w in want.items():
            lt = ck[k]
            assert isinstance(lt, NotYetLoadedTensor), k
            assert lt.shape == w.shape and lt.dtype == w.dtype and lt.size() == w.size() and lt.dim() == w.dim(), k
            got = lt._load_tensor()
            assert got.stride() == w.stride() or w.numel() <= 1, k
            assert torch.equal(got.detach(), w.detach()), k
        assert isinstance(ck["a.weight"]._load_tensor(), torch.nn.Parameter)
        assert ck["q.colmajor"].stride() == (1, 12)                     # answered without reading
        assert torch.equal(ck["q.colmajor"].contiguous(), want["q.colmajor"].contiguous())
        # any torch function materialises its placeholder arguments (what load_state_dict's param.copy_(x) relies on)
        dst = torch.zeros(33, 9)
        dst.copy_(ck["a.bf16"])
        assert torch.equal(dst, want["a.bf16"].float())
        assert torch.equal(torch.add(ck["idx"], 1), want["idx"] + 1)
        assert torch.equal(ck["view.hi"].float(), want["view.hi"])


def test_lazy_load_refuses_foreign_globals(tmp_path):
    import pickle
    from lit_llama.utils import lazy_load

    p = tmp_path / "evil.pth"
    torch.save({"x": torch.zeros(2), "f": np.float64(1.0)}, p)         # a numpy scalar pickles a numpy global
    with pytest.raises(pickle.UnpicklingError):
        lazy_load(p)


def test_load_state_dict_from_placeholders(tmp_path):
    """The reference's usage (generate.py:124-136): ``with lazy_load(path) as ck: model.load_state_dict(ck)`` - here for
    a tiny LLaMA with quantized Linears (module construction and state-dict plumbing are host code)."""
    from lit_llama import LLaMA, LLaMAConfig
    from lit_llama.utils import lazy_load, quantization

    cfg = LLaMAConfig(**TINY)
    with quantization("gptq.int4", groupsize=32):
        src = LLaMA(cfg)
    g = torch.Generator().manual_seed(0)
    for t in src.state_dict().values():
        if t.dtype == torch.uint8:
            t.random_(0, 255, generator=g)
        elif t.is_floating_point():
            t.normal_(generator=g)
    p = tmp_path / "m.pth"
    torch.save(src.state_dict(), p)
    with quantization("gptq.int4"):                                    # per-row modules: the grouped checkpoint re-shapes them
        dst = LLaMA(cfg)
    with lazy_load(p) as ck:
        dst.load_state_dict(ck, strict=True)
    for (k, a), (_, b) in zip(src.state_dict().items(), dst.state_dict().items()):
        assert torch.equal(a, b), k
    assert dst.lm_head.quant_weight.stride() == (1, dst.lm_head.out_features)


@pytest.mark.gpu
def test_stream_load_state_dict_on_the_gpu(tmp_path):
    """disk -> pinned -> HBM with two 64 KB staging buffers (forces many turns and one grow), repack under the read: same
    bytes as the ordinary load, every Linear's GEMV stream built, and the model decodes to the same logits."""
    from lit_llama import LLaMA, LLaMAConfig
    from lit_llama.utils import EmptyInitOnDevice, quantization, stream_load_state_dict

    dev = torch.device("cuda")
    cfg = LLaMAConfig(**TINY)

    def fresh():
        with EmptyInitOnDevice(device=dev, dtype=torch.bfloat16), quantization("gptq.int4", groupsize=32):
            return LLaMA(cfg)

    src = fresh()
    g = torch.Generator(device=dev).manual_seed(0)
    with torch.no_grad():
        for k, t in src.state_dict().items():
            if t.dtype == torch.uint8:
                t.random_(0, 255, generator=g)
            elif k.endswith("scales"):
                t.copy_(torch.rand(t.shape, device=dev, generator=g) * 0.01 + 0.005)
            elif k.endswith("zeros"):
                t.copy_(torch.randint(5, 11, t.shape, device=dev, generator=g).to(t.dtype))
            elif k.endswith("scale"):
                t.fill_(1.0)
            elif t.is_floating_point():
                t.normal_(generator=g).mul_(0.05)
    p = tmp_path / "m.pth"
    torch.save(src.state_dict(), p)
    a, b = fresh(), fresh()
    a.load_state_dict(torch.load(p, map_location=dev, weights_only=True))
    info = stream_load_state_dict(b, p, buffer_bytes=64 << 10)
    assert info["repacked_layers"] == 1 + 5 * cfg.n_layer and not info["missing"] and not info["unexpected"]
    assert info["h2d_bytes"] >= sum(t.numel() * t.element_size() for t in src.state_dict().values())
    for (k, x), (_, y) in zip(a.state_dict().items(), b.state_dict().items()):
        assert torch.equal(x, y), k
    assert all(m.stream_for(1)[0] is not None for m in [b.lm_head, b.transformer.h[0].attn.c_attn])
    idx = torch.randint(0, cfg.vocab_size, (1, 6), device=dev)
    with torch.no_grad():
        la = a(idx, 16, torch.arange(6, device=dev))
        lb = b(idx, 16, torch.arange(6, device=dev))
    assert torch.equal(la, lb)


def test_stream_load_refuses_a_cpu_model(tmp_path):
    """The streaming loader is a CUDA pipeline (pinned staging, copy stream): a model on the CPU is refused loudly."""
    from lit_llama.utils import stream_load_state_dict

    m = torch.nn.Linear(4, 3)
    p = tmp_path / "l.pth"
    torch.save(m.state_dict(), p)
    with pytest.raises(RuntimeError, match="CUDA"):
        stream_load_state_dict(m, p)
