"""CPU oracle for the LMI query path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is product code.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it, and only as the checker / the timed CPU
baseline.  The product path (``learnedmetricindex_b200``) never imports it and
fails loudly when its CUDA library is missing.

Contents
--------
``lmi_oracle``   pandas-free numpy/torch-CPU restatement of
                 ``LearnedIndex.search`` (reference ``search/li/LearnedIndex.py``,
                 ``model.py``, ``utils.py``, ``PriorityQueue.py``), each function
                 citing the reference file:line it follows.
``faiss_shim``   stand-in for faiss-cpu 1.7.4 (not vendored in the reference,
                 not installable here): ``faiss.knn`` restated from its published
                 behaviour (exact f32 inner products, per-query top-k, sorted
                 descending, ``-1`` / ``-FLT_MAX`` padding).
``ref_harness``  imports the UNMODIFIED reference package ``li`` from
                 ``/root/reference/search`` under the shims (build container
                 only; ``/root/reference`` does not exist on the GPU box).
``make_golden``  runs the reference through ``ref_harness`` on small seeded
                 inputs and writes ``tests/golden/*.npz``.

Parity pinning: the reference ships no tests or golden vectors for this path
(SURVEY.md §4), so the oracle is pinned against the reference's own code
executed in this container (``make_golden`` → ``tests/golden``; and live in
``tests/test_oracle_vs_reference.py`` whenever ``/root/reference`` exists).
"""
