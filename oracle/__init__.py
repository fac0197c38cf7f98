"""CPU oracle for the route -> probe -> pack -> infer -> unpack path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is product code: it may be
imported by ``tests/``, by ``__graft_entry__.smoke()`` and by ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs, and only as the checker or the
timed CPU baseline -- never by the ``die`` package.

Parity status: PINNED.  ``oracle/ref_port.py`` restates the reference's
pure-Python algorithms; ``tests/golden/make_golden.py`` executed the reference's
own classes (imported from /root/reference in the build container) and froze
their outputs as fixtures under ``tests/golden/``; ``tests/test_oracle_golden.py``
replays every fixture through this port.  The mock linear / MLP arithmetic has
no counterpart in the reference (its model is an echo, fake_model.py:52), so
for those two functions parity is "unpinned" and the tolerance is self-defined
(see ``oracle/models.py``).
"""
