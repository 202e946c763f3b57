"""Per-row scales (tile_cols = -1) at the 7B layer shapes, 2 layers: this repo's fused chain, the numpy oracle and the
UNMODIFIED reference's Triton kernel path (ref quantization.py:185-331) on the same weights and tokens - who agrees with whom."""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "lit-llama-ja_b200")]
import tests.test_gpu_config2 as T  # noqa: E402
from baseline import ref_loader  # noqa: E402

m = T.build_cfg2_model(-1)
group = m.lm_head.in_features
o = T.oracle_of(m, -1)
NP = int(sys.argv[1]) if len(sys.argv) > 1 else 64
idx = T.tokens(NP + 6, seed=9)
ours = T.run_ours(m, idx, NP + 64, NP, fused=True)
want = T.run_oracle(o, idx, NP + 64, NP)
print("ours vs oracle:", T.metrics(ours, want))
ref = ref_loader.load()
rm = ref_loader.build_quantized_model(ref, ref.model.LLaMAConfig(**T.CFG2), tile_cols=-1, device=T.DEV, dtype=torch.bfloat16)
rm.load_state_dict(m.state_dict(), strict=True)
theirs = T.run_reference_gpu(rm, idx, NP + 64, NP)
print("reference Triton path vs oracle:", T.metrics(theirs, want))
print("ours vs reference Triton path:", T.metrics(ours, theirs))
# one Linear in isolation: the reference's module forward (Triton) against its own get_weight + F.linear
lin = rm.transformer.h[0].attn.c_attn
x = torch.randn((1, 1, 4096), device=T.DEV, dtype=torch.bfloat16)
with torch.no_grad():
    y_tri = lin(x).float()
    y_fb = torch.nn.functional.linear(x, lin.get_weight(torch.bfloat16)).float()
print("reference c_attn: Triton forward vs its own get_weight + F.linear: max |diff|", float((y_tri - y_fb).abs().max()), "max |y|", float(y_fb.abs().max()))
import triton
print("triton", triton.__version__)
