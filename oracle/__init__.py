"""TEST INFRASTRUCTURE ONLY — the CPU oracle for the shard-execution hot path.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it, and only as the checker / the reported CPU baseline.  The product
package (``llm-sharding_b200/``) never imports it and fails loudly without its CUDA
library.

Contents
--------
``synth.py``         deterministic random-init shards in the reference's ``ModelSharder``
                     on-disk layout (``/root/reference/utils/model_sharder.py:64-94``) plus
                     a synthetic tokenizer; numpy PCG64, bf16-representable values.
``llama_oracle.py``  torch-CPU restatement of the arithmetic the reference runs through
                     Hugging Face ``transformers`` (third-party, pinned ``==4.53.2`` in
                     ``/root/reference/requirements.txt:8``, 5.5.0 installed here).
``ref_harness.py``   runs the REAL reference files from ``/root/reference`` (this container
                     only; the GPU box does not have them) behind a two-point compat shim,
                     and writes the golden fixtures under ``tests/golden/``.

Parity status: the reference has no tests / golden vectors of its own (SURVEY.md §4,
F7), so the restatement is pinned against fixtures produced by running the reference
itself here (``tests/golden/*.npz``, generator ``oracle/make_golden.py``).
"""
