"""CPU oracle for the lit-llama-ja quantized-inference hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is imported by the product
package (``lit-llama-ja_b200/``); only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may use it, and there
only as the checker / the timed CPU baseline.

It is a plain-numpy restatement of the reference's algorithms (every function
cites the reference file:line it follows).  Parity status:

* GPTQ int4/int8 pack / unpack / dequant-matmul, RTN parameter search, RMSNorm,
  RoPE, KV-cache attention, MLP, LLaMA forward, generate loop: **pinned** against
  outputs of the unmodified reference run in the build container
  (``tests/golden/make_golden.py`` -> ``tests/golden/*.npz``); the reference's own
  test-suite holds no quantization vectors.
* LLaMA-Adapter v1 / v2 forward (``oracle/adapter.py``): **pinned** against outputs of the
  unmodified reference (``tests/golden/make_golden_adapter.py`` -> ``adapter_tiny.npz``).
* llm.int8 (bitsandbytes, un-vendored and unpinned in the reference's
  requirements.txt:7): **parity unpinned** - restated from the published
  LLM.int8() algorithm (Dettmers et al., arXiv:2208.07339).
"""
