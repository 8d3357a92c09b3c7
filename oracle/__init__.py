"""CPU oracle for the Ape-X prioritized-replay / learner data path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and there only as the checker or as
the timed CPU baseline.  The product path (``ape-x-dqn_b200/``) never imports
this package and fails loudly when the CUDA library is missing.

Parity status ("pinned" vs "unpinned"), per function:

* n-step assembly (``nstep_oracle.LiteralExperienceBuffer``), initial priority
  (``nstep_oracle.initial_priority``) and the proportional sampling law
  (``replay_oracle.reference_law_sample``) are PINNED against outputs of the
  reference itself, generated in the build container by
  ``tests/golden/make_golden.py`` (imports /root/reference, writes
  ``tests/golden/*.json|npz``).
* The learner arithmetic (``td_oracle``) follows learner.py:43-52.  That function
  raises ``TypeError`` on this container's torch/numpy as written (SURVEY Q14:
  ndarray * Tensor); make_golden.py runs the unmodified function under an operator
  shim that restores torch 0.4's ndarray coercion and records its per-row |TD| and
  loss (``tests/golden/td_reference.npz``), so this part is PINNED too, with that one
  documented shim; it is also cross-checked against a torch fp64 evaluation.
* The batched actor policy (``policy_oracle``: duelling combine at batch 1, epsilon-greedy) is PINNED
  against the reference model run at batch 1 and ``Actor.epsilon_greedy_Q`` run on its own ``random``
  stream (``tests/golden/policy_reference.npz``).
* The lossless checkpoint packing (``pack_oracle``) has no counterpart in the reference (README.md:24 lists
  observation compression as a TODO): it restates this repository's own format, so there is nothing upstream to
  pin it against; it is checked by round trips and the kernels are checked byte for byte against it.
* The reference ships no tests, so there are no upstream golden vectors.
"""

from . import replay_oracle, td_oracle, nstep_oracle, policy_oracle  # noqa: F401
