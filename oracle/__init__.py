"""CPU oracle for the ConvNeXt-block training hot path.  TEST INFRASTRUCTURE ONLY.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs
may import this package; nothing under `image_classification_b200/` does (tests/test_layout.py
greps for it).  It is the checker, never the thing measured or shipped.

Pinning status:
  * mix (MixUp/CutMix plan, pixels, soft targets), EMA and the warmup-cosine LR schedule are
    PINNED: `oracle/gen_golden.py` imports the reference's own `dataset.py` /
    `train_advanced_v4.py` from /root/reference (with stub modules for the absent `timm` and
    `albumentations`), runs them on seeded inputs and commits the outputs under tests/golden/;
    tests/test_oracle_golden.py checks this restatement against those vectors bit for bit.
  * the ConvNeXt block itself lives in `timm` (requirements.txt:3 `timm>=0.9.2`, unpinned, not
    vendored, not installable here) and the reference holds no test, golden vector or checkpoint
    for it: **parity unpinned** by the reference.  The restatement follows timm's published
    ConvNeXtBlock and is cross-checked against the architecture-identical
    torchvision `CNBlock` (tests/test_oracle_block.py), which is the strongest anchor available.
"""
