"""Perplexity of a LLaMA-Adapter v2 fine-tune (ref: /root/reference/evaluate/adapter_v2.py): as evaluate/adapter.py, with the
per-Linear ``adapter_bias`` / ``adapter_scale`` parameters added before the checkpoints are loaded."""
import importlib.util
from pathlib import Path
from typing import Optional

_spec = importlib.util.spec_from_file_location("llb_evaluate_adapter", Path(__file__).absolute().parent / "adapter.py")
_v1 = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(_v1)


def main(
    data_paths: str = "data/wikitext-2-raw/wiki.test.raw",
    *,
    accelerator: str = "auto",
    adapter_path: Path = Path("out/adapter_v2/alpaca/lit-llama-adapter-finetuned.pth"),
    checkpoint_path: Path = Path("checkpoints/lit-llama/7B/lit-llama.pth"),
    tokenizer_path: Path = Path("checkpoints/lit-llama/tokenizer.model"),
    dtype: str = "float32",
    quantize: Optional[str] = None,
) -> None:
    """ref evaluate/adapter_v2.py - one perplexity per text file."""
    _v1.run(data_paths, accelerator, adapter_path, checkpoint_path, tokenizer_path, dtype, quantize, v2=True)


if __name__ == "__main__":
    _v1.cli(main, "out/adapter_v2/alpaca/lit-llama-adapter-finetuned.pth")
