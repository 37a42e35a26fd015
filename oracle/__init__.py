"""CPU oracle for the adversarial-SDF step.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and the CPU legs of ``bench.py`` (``cpu_baseline`` and
``--impl reference``) may import it, and there only as the checker or as the timed CPU
baseline -- never as the thing shipped.  The product path
(``deeplearninginassetpricing_paperreplication_b200``) fails loudly when its CUDA library
is missing; it never routes through this package.

Two independent restatements of the reference algorithm live here:

* ``torch_port``  -- op-for-op restatement of ``src/model.py`` on torch CPU tensors
  (the reference's arithmetic *is* torch, so this one is bit-identical to it and is what
  the goldens under ``tests/golden`` pin with ``torch.equal``).  Backward is autograd.
* ``closed_form`` -- numpy float64 forward AND hand-derived backward,
  independent of autograd.  It is the specification the CUDA kernels implement.

Parity status: PINNED.  The reference ships no tests of its own, so the pin is against
outputs of the reference itself run in the build container
(``tests/golden/make_golden.py`` imports ``/root/reference`` and writes the fixtures).
"""
