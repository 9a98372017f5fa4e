"""CPU oracle for the ViT-VS `dino` servoing hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it, and there only as the
checker or as the timed CPU baseline.  The product package (``vitvs_b200``)
never imports this package and fails loudly when its CUDA library is missing.

Parity status:
  * matching + IBVS stages: PINNED - the restatement in
    ``oracle/servo.py`` is checked against the reference's own functions
    executed from /root/reference (``tests/test_oracle_vs_reference.py`` here,
    committed fixtures ``tests/golden/*.npz`` everywhere).
  * descriptor stage (a7/a8: hooks, CLS drop, ``_log_bin``): PINNED the same way
    (the reference's ``ViTExtractor`` runs on the restated backbone).
  * backbone (a4-a6): the DINOv2 model code is a third-party, un-vendored,
    un-pinned torch.hub dependency (facebookresearch/dinov2, hub default
    branch) => "parity unpinned" by the reference itself.  The restatement in
    ``oracle/dinov2.py`` follows dino_patch/{attention,block}.py (vendored in
    the reference) and is cross-checked by weight copy against
    ``transformers.models.dinov2.Dinov2Model``.
"""
