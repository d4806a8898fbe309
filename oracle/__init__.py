"""CPU oracle for the pipeline hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import it;
the engine package never does (``tests/test_no_oracle_in_product.py`` enforces
that).  Every function cites the reference ``file:line`` it restates.

Parity pin status:
  * CIFAR parts  -- PINNED: restatement checked bit-for-bit against the
    reference's own classes imported from /root/reference (golden vectors under
    tests/golden/, generator oracle/make_golden.py).
  * GPT parts    -- stage structure pinned against the reference's
    partitions/gpt_model_parts.py; the block arithmetic lives in a third-party
    ``model`` module that is NOT in the reference tree (no version pinned), so
    that half is "parity unpinned": restated from the public nanoGPT semantics
    and cross-checked against transformers' GPT2LMHeadModel.
"""
