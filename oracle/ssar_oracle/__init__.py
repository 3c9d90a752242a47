"""CPU oracle for ssar_b200 — TEST INFRASTRUCTURE ONLY.

A CPU restatement (torch-CPU / numpy, fp32 like the reference) of the reference's algorithms for the hot path:
ssar/features/**, ssar/supervised/data.py:60-107, ssar/random/{latent,noise,patch}.py and
ssar/models/latent_n_noise2.py:134-178.  Each function cites the reference file:line it follows.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may import this
package, and only as the checker or the timed CPU baseline.  The product package `ssar_b200` never imports it
and has no CPU fallback.

Parity pinning: the reference ships no golden vectors or tests for this path, so the oracle is
pinned against outputs of the UNMODIFIED reference run in the build container (`tests/golden/make_golden.py`
imports /root/reference through `tests/golden/ref_loader.py`; the resulting fixtures are committed under
`tests/golden/*.npz`, and `tests/test_oracle_vs_golden.py` checks the oracle against them on CPU).
One boundary stays UNPINNED: `torchcubicspline` (un-vendored, un-pinned, not installable here) — the oracle and
the loader shim both follow the published upstream algorithm (natural cubic spline, coefficients
(t, a, b, two_c, three_d)).  It affects only chroma_cens' spline quantiser and spline_loop_latents.
"""
