"""Drop-in ``lit_llama`` package for B200 (ref: lit_llama/__init__.py:1-2)."""
from lit_llama.model import LLaMAConfig, LLaMA, RMSNorm, build_rope_cache, apply_rope
from lit_llama.tokenizer import Tokenizer, HFTokenizer
