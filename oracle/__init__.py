"""CPU oracle for the UFM dense-correspondence forward pass.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product path: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import it, and only as the
checker / reported CPU baseline.  The engine in ``ufm_b200/`` never imports this package and fails
loudly when its CUDA library is missing.

Parity status: the reference has NO golden vectors for UFM.  The oracle
(:mod:`oracle.ufm_oracle`, a plain PyTorch-CPU fp32 restatement) is pinned instead against outputs
of the reference itself: ``oracle/make_golden.py`` imports the unmodified reference classes from
``/root/reference`` in the build container (only ``torch.hub.load`` is substituted, see
:mod:`oracle.ref_harness`), runs them on seeded inputs and weights, and commits stage-boundary
tensors under ``tests/golden/``.  The DINOv2 encoder arithmetic is third-party
(``facebookresearch/dinov2`` via ``torch.hub``, un-vendored and unpinned by the reference), restated
from its published architecture: encoder parity is therefore "unpinned" against upstream DINOv2,
but pinned between oracle, reference wrapper and engine, which all consume the same restatement.
"""
