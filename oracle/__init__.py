"""CPU oracle for protein-clip's contrastive hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import it, and only as the checker or as the timed CPU
baseline.  The product path (protein-clip_b200/pclip -> libpclip.so) never imports this package and fails loudly
when the CUDA library is missing.

Parity status: PINNED.  The reference ships no golden vectors or tests of its own, so the
restatement in oracle/restate.py is pinned against outputs of the reference's own functions, imported from
/root/reference in the authoring container by oracle/make_golden.py and committed under tests/golden/*.npz
(tests/test_oracle_golden.py re-checks every vector on every run; tests/test_oracle_vs_reference.py additionally
calls the live reference whenever /root/reference is mounted).  The one exception is grad_cache
(luyug/GradCache @ 33695437, un-vendored and absent from /root/reference): its two decorators are restated from
the published algorithm and are "parity unpinned" at that boundary; they contain no loss arithmetic.
"""
