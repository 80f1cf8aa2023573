"""CPU oracle for the OE-VAE hot path -- TEST INFRASTRUCTURE ONLY.

This package is a CPU restatement (NumPy float64/float32 + a torch-CPU autograd
twin) of the reference algorithm in /root/reference/OE-VAE/{models,utils,plots}.py.
It is *not* part of the product: only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it, and only
as the checker or the timed CPU baseline.  The product path (``atlas_vae_b200``)
never imports it and fails loudly when ``liboevae.so`` is missing.

Parity status
-------------
* Pure-NumPy/sklearn pieces of the reference (``utils.loss_function``, ``get_idx``,
  ``jets_sorting``, ``get_4v``, ``filtering``, ``plots.loss_mapping``, ``get_rates``,
  ``best_threshold``, ``make_cut``, ``models.model_checkpoint``) are PINNED: the
  oracle is checked against outputs of the reference's own function bodies, executed
  from /root/reference by ``tests/golden/make_golden.py`` (AST extraction, because the
  modules import tensorflow/h5py/matplotlib at top level, none of which is installed).
* The TensorFlow-resident arithmetic (Dense, GradientTape, Adam, random.normal;
  ``models.py:7-132,181-186``) cannot run here (TF 2.1 absent, no network):
  for those pieces this oracle is **parity unpinned** -- two independent derivations
  (hand-written NumPy backward and torch autograd) are cross-checked instead.
"""
