"""CPU oracle for the DQN minibatch-update hot path of hbfs/dist-dqn.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product: only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import it, and there only as the checker or as the timed CPU
baseline -- never as a fallback for the CUDA path.

Parity status
-------------
* Data half (``replay_ref.py``): PINNED.  The restatement is checked, in
  ``tests/test_oracle_golden.py``, against ``tests/golden/replay_*.npz`` which
  ``oracle/gen_golden.py`` produced by importing and running the reference's own
  ``replay_memory.py`` / ``frame_buffer.py`` / ``utils.py`` from ``/root/reference/src``.
* Model half (``net_ref.py``, ``torch_ref.py``): PARITY UNPINNED.  The arithmetic lives in
  TensorFlow (un-vendored, un-pinned, API era r0.9-r0.11; call sites ``network.py:61-64,
  149-152, 169-202, 251-252, 304-311, 401-407, 417-424``), which is not installable on
  this image, and the reference ships no tests, golden vectors or fixtures for it.  The
  two restatements (numpy float64 with hand-written backward; torch-CPU autograd) are
  independent of each other and must agree to ~1e-10; a finite-difference check guards
  the hand-written backward.  The TF semantics restated are listed in ``net_ref.py``.
"""
