"""CPU oracle for the SevenNet E/F/S evaluation path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``sevennet_exp_b200/`` may import this
package.  Allowed users: ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``.

Parity status: PINNED for lmax<=2 (Wigner-3j goldens stored in the reference's
``tests/data/checkpoints/cp_0.pth``, the ``test_train.py:284-316`` metric
known-answers, parameter counts of ``test_model.py:164-175`` and edge counts of
``test_data.py:48``); lmax=3 spherical harmonics are pinned only through the
w3j recursion self-check (the reference's l3i5 checkpoint is absent), see
DESIGN.md.
"""
