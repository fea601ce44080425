"""bqa_norm_noalloc (include/bqa_b200.h): the allocation-free variant of the norm reduction, through the C ABI,
against the existing bqa_norm entry point - on calls in a row that share the static device scratch.

Collected last on purpose; passes alone in a fresh process and after the
other GPU tests."""
import numpy as np
import pytest
from conftest import cuda_available

from bloqade_analog_b200 import _lib

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not cuda_available(), reason="needs a CUDA device")]

N_ATOMS = 16  # dim 65536: 256 CTAs per state
TOL = 1e-14   # the tolerance test_observables uses for bqa_norm


def _states(torch, n_states, dim, seed, scale):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return scale * torch.randn(n_states, dim, 2, dtype=torch.float64, device="cuda", generator=g)


def _check_against_bqa_norm(torch, plan, psi, out):
    n_states, dim = psi.shape[0], psi.shape[1]
    plan.norm_noalloc(psi.data_ptr(), n_states, out.data_ptr())
    torch.cuda.synchronize()
    got = out[:n_states].cpu().numpy()
    for s in range(n_states):
        ref = plan.norm(psi[s].data_ptr(), mem=_lib.MEM_DEVICE)
        assert abs(got[s] - ref) <= TOL * max(1.0, ref), (s, got[s], ref)
    host = np.sqrt((psi.cpu().numpy() ** 2).sum(axis=(1, 2)))
    np.testing.assert_allclose(got, host, rtol=1e-12)
    return got


def test_norm_noalloc_matches_bqa_norm_on_calls_in_a_row():
    import torch

    dim = 1 << N_ATOMS
    plan = _lib.Plan(N_ATOMS, 2, dim)
    out = torch.full((8,), -1.0, dtype=torch.float64, device="cuda")
    first = _states(torch, 5, dim, seed=11, scale=1.0)
    second = _states(torch, 2, dim, seed=12, scale=1e-3)  # fewer, much smaller states: stale partial sums would show
    a = _check_against_bqa_norm(torch, plan, first, out)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    b = _check_against_bqa_norm(torch, plan, second, out)
    assert float(out[2]) == a[2]  # entries beyond n_states are not touched
    c = _check_against_bqa_norm(torch, plan, first, out)
    assert np.array_equal(a, c)  # deterministic, independent of what the previous call left behind
    assert b.max() < 1e-2 * a.min()
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before  # nothing was allocated on the device
    plan.close()


def test_norm_noalloc_small_and_subspace_dims():
    import torch

    for n_atoms, dim in ((3, 8), (9, 512), (12, 4096)):
        plan = _lib.Plan(n_atoms, 2, dim)
        out = torch.zeros(4, dtype=torch.float64, device="cuda")
        _check_against_bqa_norm(torch, plan, _states(torch, 3, dim, seed=dim, scale=2.5), out)
        _check_against_bqa_norm(torch, plan, _states(torch, 4, dim, seed=dim + 1, scale=0.5), out)
        plan.close()


def test_norm_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch

    dim = 1 << N_ATOMS
    plan = _lib.Plan(N_ATOMS, 2, dim)
    psi = _states(torch, 1, dim, seed=1, scale=1.0)
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    max_states = 65536 // 256
    with pytest.raises(_lib.BackendError, match="exceed the static scratch"):
        plan.norm_noalloc(psi.data_ptr(), max_states + 1, out.data_ptr())
    with pytest.raises(_lib.BackendError, match="null"):
        plan.norm_noalloc(0, 1, out.data_ptr())
    with pytest.raises(_lib.BackendError, match="n_states"):
        plan.norm_noalloc(psi.data_ptr(), 0, out.data_ptr())
    # the error path leaves the entry point usable
    _check_against_bqa_norm(torch, plan, psi, out)
    plan.close()
