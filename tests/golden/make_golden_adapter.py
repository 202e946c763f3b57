"""Golden vectors for the LLaMA-Adapter callers, produced by running the UNMODIFIED reference
(/root/reference/lit_llama/adapter.py, adapter_v2.py) in the build container on the weights of
``tests/util_tiny.adapter_tiny_weights``.  Not runnable on the GPU box (the reference is absent there); the file it
writes, ``adapter_tiny.npz``, is committed beside it.

    python tests/golden/make_golden_adapter.py
"""
import importlib.util
import sys
from pathlib import Path

import numpy as np
import torch

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT))

from tests.util_tiny import ADAPTER_TINY, adapter_tiny_weights  # noqa: E402


def _reference():
    spec = importlib.util.spec_from_file_location("llb_make_golden", HERE / "make_golden.py")
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    mg._import_reference()                       # stubs `lightning`, puts /root/reference on sys.path
    from lit_llama import adapter, adapter_v2    # the reference's modules
    return adapter, adapter_v2


def build(adapter, adapter_v2, v2: bool):
    torch.manual_seed(0)
    model = adapter.LLaMA(adapter.LLaMAConfig(**ADAPTER_TINY))
    if v2:
        adapter_v2.add_adapter_v2_parameters_to_linear_layers(model)
    sd = {k: torch.from_numpy(v) for k, v in adapter_tiny_weights(v2=v2).items()}
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all("adapter" not in k and "gating" not in k for k in missing), missing
    assert not missing, missing
    return model.eval()


@torch.no_grad()
def run(model, out, tag):
    r = np.random.RandomState(7)
    idx = r.randint(0, ADAPTER_TINY["vocab_size"], size=(2, 12))
    out["idx"] = idx
    out[tag + "nocache"] = model(torch.from_numpy(idx)).float().numpy()
    for B in (1, 2):
        model.reset_cache()
        S = 16
        got = [model(torch.from_numpy(idx[:B, :8]), S, torch.arange(8)).float().numpy()[:, -1]]
        for t in range(8, 12):
            got.append(model(torch.from_numpy(idx[:B, t:t + 1]), S, torch.tensor([t])).float().numpy()[:, -1])
        out[f"{tag}cached_b{B}"] = np.stack(got, axis=1)          # [B, 5, V]
    model.reset_cache()


def main():
    adapter, adapter_v2 = _reference()
    out = {}
    run(build(adapter, adapter_v2, False), out, "v1.")
    run(build(adapter, adapter_v2, True), out, "v2.")
    # old checkpoints: ONE gating value for all heads (adapter.py:171-186)
    m = build(adapter, adapter_v2, False)
    sd = {"transformer.h.1.attn.gating_factor": torch.tensor([0.25])}
    m.load_state_dict(sd, strict=False)
    out["old_gate"] = m.transformer.h[1].attn.gating_factor.detach().numpy()
    np.savez_compressed(HERE / "adapter_tiny.npz", **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
