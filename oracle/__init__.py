"""CPU oracle for the py_mc Metropolis-sweep / replica-exchange hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it, and there only as the
checker (or as the timed CPU baseline), never as the thing shipped.  The
product package ``py_mc_b200`` never imports this package and has no CPU
fallback.

What it restates (all citations relative to the reference tree):

* ``py_mc/mc.py:170-185``      one Metropolis trial (``MCstep``)
* ``examples/ising2D.py:15-52`` Ising energy / trialstep / newconfig
* ``examples/latgas.py:17-57``  lattice-gas flip moves
* ``examples/latgas_move.py:31-95`` lattice-gas particle moves
* ``py_mc/mc_mpi.py:102-140``  replica-exchange ``Xtrial``
* ``examples/2D_hardcore/model_setup.py:24-125`` hard discs

Parity pin: the reference ships no tests or golden vectors, so
the oracle is pinned against the reference *itself*, imported from
``/root/reference`` in the build container and driven with injected random
draws; the outputs are committed under ``tests/golden/`` together with the
script that made them (``tests/golden/make_golden.py``).  ``tests/test_oracle_*``
check the oracle against those fixtures on every run; ``tests/test_live_reference.py``
regenerates every fixture from the live reference (``/root/reference`` or the
git-ignored copy ``oracle/_ref`` made by ``oracle/make_ref.py``) wherever it is present
and compares array for array.  ``oracle/refarm.py`` times that same unmodified reference
for ``bench.py``'s CPU baseline.
"""
