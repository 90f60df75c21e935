"""CPU oracle for the mused SWFD hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is product code.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and there only as the checker or the
timed CPU baseline.  The product package ``mused_b200`` never imports it and
fails loudly when its CUDA library is missing.

Parity status: the reference (kelaendi/mused @ a3612f3, swfd @ 3113ec3) ships
no golden vectors, KATs or asserting tests for this path,
so the restatement is pinned against OUTPUTS OF THE REFERENCE ITSELF, run in
the build container by ``oracle/gen_golden.py`` (imports /root/reference via
``oracle/ref_shim.py``), asserted bit-identical there and committed as
fixtures under ``tests/golden/``.

Third-party arithmetic the reference reaches (not vendored under
/root/reference, versions unpinned upstream; the versions used to make the
fixtures are recorded inside every fixture): scipy.linalg.svd(gesvd) --
LAPACK dgesvd; numpy matmul.
"""
