"""ORACLE — test infrastructure only.

CPU restatements of the reference hot path (gateway preprocessing + Xception-299 forward).
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product package
(``kubernetes-deep-learning_b200`` / ``kdl_b200``) never imports it and has no CPU fallback.

Parity status (also stated in DESIGN.md):
  * preprocessing: PINNED — against Pillow itself (the reference's real resize implementation,
    importable in this image) and against the media/image2.png known-answer hashes.
  * network: PARITY UNPINNED — TensorFlow 2.3.0 / Keras / the trained .h5 are not installable
    or reachable offline, and the reference ships no tests or golden logits that can be
    reproduced without them (guide.md:622-626 needs the .h5 and the URL image).
"""
