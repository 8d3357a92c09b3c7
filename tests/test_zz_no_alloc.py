"""GPU test of apexb200_td_fwd_noalloc, the TD/loss entry point whose reduction scratch is a
static buffer inside the library instead of a caller-provided workspace.

It is meant for callers that cannot allocate device memory themselves (bindings without
the CUDA runtime, processes whose framework already owns all of HBM, loops that must not
hit cudaMalloc's device-wide synchronisation).  The test calls it through the C ABI and
compares every output with apexb200_td_fwd on the same inputs and, directly, with the CPU
oracle of learner.py:43-52 (oracle/td_oracle.py), on several calls in a row: each later
call is smaller than the first, so stale partial sums and the used ticket left in
the shared scratch by the first call would show up if the entry point did not handle them.
This file sorts last on purpose: with `pytest -x` every other test has run before it.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from oracle import td_oracle as to               # noqa: E402

pytestmark = pytest.mark.gpu

RTOL = 1e-5     # the tolerance of tests/test_gpu_parity.py


def _inputs(rng, B, nA):
    q = [torch.tensor(rng.standard_normal((B, nA)).astype(np.float32) * 2, device="cuda") for _ in range(3)]
    A = torch.tensor(rng.integers(0, nA, B).astype(np.int32), device="cuda")
    R = torch.tensor(np.clip(rng.standard_normal(B), -1, 1).astype(np.float32), device="cuda")
    G = torch.tensor(np.where(rng.random(B) < 0.05, 0, 0.970299).astype(np.float32), device="cuda")
    return q, A, R, G


def test_noalloc_matches_workspace_entry_point_on_consecutive_calls():
    import apex_b200
    from apex_b200 import _C, ops
    torch.cuda.set_device(0)
    rng = np.random.default_rng(2024)
    # (rows, actions, loss): big first, then smaller batches that reuse the same scratch bytes;
    # 700 rows takes the single-CTA kernel in the workspace entry point, so only tolerance-equal there.
    cases = [(70000, 4, _C.LOSS_HALF_SQ), (1500, 4, _C.LOSS_HALF_SQ), (5000, 18, _C.LOSS_HUBER), (700, 6, _C.LOSS_HALF_SQ)]
    for B, nA, kind in cases:
        q, A, R, G = _inputs(rng, B, nA)
        want = ops.td_forward(q[0], q[1], q[2], A, R, G, 0.6, 0.0, kind, 1.0)
        # outputs are allocated up front: the call under test itself must not need any allocation
        out = tuple(torch.full((n,), float("nan"), device="cuda") for n in (B, B, B, 1))
        torch.cuda.synchronize()
        got = ops.td_forward_noalloc(q[0], q[1], q[2], A, R, G, 0.6, 0.0, kind, 1.0, out=out)
        torch.cuda.synchronize()
        for name, g, w in zip(("delta", "prio", "loss", "loss_mean"), got, want):
            torch.testing.assert_close(g, w, rtol=RTOL, atol=1e-7, msg=lambda m: "%s B=%d: %s" % (name, B, m))
        if B > 1024:                                   # same kernel on both sides: identical bits
            assert all(torch.equal(g, w) for g, w in zip(got, want))
        # and against the oracle itself, not only against the sibling entry point
        o = to.td_forward(*(x.cpu().numpy() for x in q), A.cpu().numpy(), R.cpu().numpy(), G.cpu().numpy(), alpha=0.6,
                          loss_kind=kind, kappa=1.0)
        np.testing.assert_allclose(got[0].cpu().numpy(), o["delta"], rtol=RTOL, atol=1e-6)
        np.testing.assert_allclose(got[1].cpu().numpy(), o["prio"], rtol=RTOL, atol=1e-7)
        np.testing.assert_allclose(got[2].cpu().numpy(), o["loss"], rtol=RTOL, atol=1e-7)
        np.testing.assert_allclose(got[3].item(), o["loss_mean"], rtol=RTOL)
    # the documented limit is enforced with a clear error, before anything is launched
    B = _C.TD_NOALLOC_MAX_B + 1
    z = torch.zeros(B, 4, device="cuda")
    zi, zf = torch.zeros(B, dtype=torch.int32, device="cuda"), torch.zeros(B, device="cuda")
    with pytest.raises(_C.ApexError, match="exceeds the built-in scratch") as err:
        ops.td_forward_noalloc(z, z, z, zi, zf, zf, 0.6)
    assert err.value.code == _C.ERANGE
