"""CPU oracle for the BlackBird self-play hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``blackbird_b200/`` imports this
package; it is used by ``tests/``, by ``__graft_entry__.smoke()`` and by the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` as the checker /
baseline, never as the product path.

The oracle is a numpy restatement of the reference algorithm (every function
cites the reference file:line it follows).  It is pinned against golden vectors
produced by running the *unmodified* reference in the build container
(``tests/golden/make_golden.py``; fixtures in ``tests/golden/*.json``).  The
network forward (``oracle/network_np.py``) has no reference-side pin because
TensorFlow cannot be installed here: that part is "parity unpinned".
"""
