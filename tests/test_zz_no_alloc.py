"""The workspace-free loss entry point (b200_ce_topk_fwd_static), through the C ABI.

It exists for callers that cannot obtain device memory at call time (for instance a validation hook
running after the framework has taken all of HBM, or a thread that must not trigger the device-wide
stall of cudaMalloc/cudaFree).  It must give exactly what the workspace-taking entry point gives - it
runs the same kernel on a scratch array built into the library - on repeated calls, including when the
batch size changes between calls (the array keeps stale contents from the previous call).

Each call is checked twice: bit for bit against the workspace-taking entry point, and against the
float64 oracle with the tolerances the other loss tests use.

Collected last on purpose (file name), so a run that stops at the first failure has already executed
every other test; it passes on its own in a fresh process as well.
"""
import numpy as np
import pytest
import torch

from oracle import cross_entropy as oce

import imagent_distributed_training_pytorch_with_slurm_b200._cabi as cabi

pytestmark = pytest.mark.gpu


def _case(n, c, seed, dtype):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(n, c, generator=g) * 3).to(dtype)
    t = torch.randint(0, c, (n,), generator=g)
    for r in range(0, n, 3):
        x[r, t[r]] += 6.0
    return x.cuda(), t.cuda()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_static_scratch_entry_matches_workspace_entry_on_consecutive_calls(dtype):
    # two calls in a row, the second smaller than the first so stale per-row slots are left behind
    for call, (n, c) in enumerate([(256, 1000), (97, 1000), (256, 1000)]):
        x, t = _case(n, c, 10 + call, dtype)
        want3, want_lse = cabi.ce_topk_fwd(x, t)
        out3 = torch.full((3,), float("nan"), device="cuda")
        lse = torch.full((n,), float("nan"), device="cuda")
        free_before = torch.cuda.memory_allocated()
        cabi.ce_topk_fwd_static(x, t, out3, lse)
        assert torch.cuda.memory_allocated() == free_before        # the wrapper created no tensor
        # same kernel, same inputs, fixed-order reduction: identical bits (tolerance of the other CE
        # tests is 2e-6 relative; this is stricter)
        assert torch.equal(out3, want3), (call, out3, want3)
        assert torch.equal(lse, want_lse)
        # and directly against the oracle (float64 accumulation), with the tolerances of test_ce_topk_matches_oracle:
        # loss 2e-6 relative, the two precision counts bit-exact, per-row log-sum-exp 2e-6
        loss, t1, t5, _, row = oce.ce_topk_oracle(x.float().cpu().numpy(), t.cpu().numpy())
        got = out3.cpu().numpy()
        assert abs(got[0] - loss) <= 2e-6 * abs(loss)
        assert got[1] == t1 and got[2] == t5
        assert np.allclose(lse.cpu().numpy(), row, rtol=2e-6, atol=2e-6)


def test_static_scratch_entry_rejects_batches_beyond_its_capacity():
    n = cabi.CE_STATIC_MAX_ROWS + 1
    x = torch.zeros(n, 8, device="cuda")
    t = torch.zeros(n, dtype=torch.int64, device="cuda")
    out3, lse = torch.zeros(3, device="cuda"), torch.zeros(n, device="cuda")
    with pytest.raises(cabi.B200Error, match="at most"):
        cabi.ce_topk_fwd_static(x, t, out3, lse)
    # the C entry point itself refuses too (cudaErrorInvalidValue = 1), without launching anything
    rc = cabi.load().b200_ce_topk_fwd_static(x.data_ptr(), t.data_ptr(), out3.data_ptr(), lse.data_ptr(), n, 8,
                                             cabi.F32, torch.cuda.current_stream().cuda_stream)
    assert rc == 1
    # the largest allowed batch works
    n = cabi.CE_STATIC_MAX_ROWS
    x, t = _case(n, 16, 3, torch.float32)
    out3, lse = torch.zeros(3, device="cuda"), torch.zeros(n, device="cuda")
    cabi.ce_topk_fwd_static(x, t, out3, lse)
    want3, _ = cabi.ce_topk_fwd(x, t)
    assert torch.equal(out3, want3)
