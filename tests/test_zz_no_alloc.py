"""lb_sigma_fwd_noalloc (include/loner_b200.h): the sigma-head inference entry point that allocates nothing -- its encode ->
MLP hand-off and ReLU masks live in a fixed scratch array inside the library.  No counterpart in the reference,
so the checker is the existing entry point lb_sigma_fwd with caller-owned buffers:
same kernels, the results must agree (tolerance of the other sigma tests; they are in fact bit-identical).

Collected last on purpose (file name): a failure here must not hide other tests.
Everything goes through the C-ABI (ctypes, loner_b200._lib.call).  Passes alone in a fresh process and after the rest
of the suite: it depends on no state but the library's own scratch, which every call fully rewrites before reading.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def rel_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def make_inputs(n, s, seed, plan):
    gen = torch.Generator().manual_seed(seed)
    params = ((torch.rand(plan.n_params, generator=gen) - 0.5) * 0.6).to(DEV)
    rays = torch.zeros(n, 13)
    rays[:, 0:3] = (torch.rand(n, 3, generator=gen) - 0.5) * 0.4
    d = torch.randn(n, 3, generator=gen)
    rays[:, 3:6] = d / d.norm(dim=1, keepdim=True)
    z = torch.rand(n, s, generator=gen).sort(dim=1).values * 0.7
    return params, rays.to(DEV), z.contiguous().to(DEV)


@pytest.mark.parametrize("multi_group", [True, False], ids=["scratch_path", "one_kernel_path"])
def test_noalloc_matches_sigma_fwd_on_two_calls_in_a_row(multi_group):
    """Two calls in a row (the second smaller, so most of the scratch still holds the first call's data) against
    lb_sigma_fwd.  multi_group: a tiny L2 budget forces the level-group path -- the one that needs the enc / act
    hand-off and therefore the built-in scratch; otherwise the one-kernel path runs and the scratch stays untouched."""
    from loner_b200 import ops
    plan = ops.Plan(log2_hashmap_size=15, l2_bytes=700 * 1024 if multi_group else None)
    assert plan.enc_optional == (not multi_group)
    for call, (n, s, seed) in enumerate([(300, 129, 1), (37, 64, 2)]):
        params, rays, z = make_inputs(n, s, seed, plan)
        want, _ = ops.sigma_fwd(plan, params, rays=rays, z=z, save_enc=multi_group)       # existing entry point
        got = torch.full((n, s), float("nan"), device=DEV)
        torch.cuda.synchronize()
        free0, alloc0 = torch.cuda.mem_get_info()[0], torch.cuda.memory_allocated()
        ops.sigma_fwd_noalloc(plan, params, got, rays=rays, z=z)
        torch.cuda.synchronize()
        if call == 1:        # after the first call's set-up: neither the driver's nor torch's allocations move
            assert torch.cuda.mem_get_info()[0] == free0 and torch.cuda.memory_allocated() == alloc0
        assert torch.isfinite(got).all()
        assert rel_err(got, want) < 1e-5, (call, rel_err(got, want))
        assert torch.equal(got, want), "same kernels, same inputs: the results should be bit-identical"
        # the x01 point source (tcnn.NetworkWithInputEncoding's argument) through the same entry point
        x01 = ((rays[:, None, 0:3] + rays[:, None, 3:6] * z[..., None]).reshape(-1, 3) + 1) / 2
        x01 = x01.contiguous()
        want_x, _ = ops.sigma_fwd(plan, params, x01=x01, save_enc=multi_group)
        got_x = torch.empty(n * s, device=DEV)
        ops.sigma_fwd_noalloc(plan, params, got_x, x01=x01)
        assert rel_err(got_x, want_x) < 1e-5


def test_noalloc_limit_and_empty_batch():
    from loner_b200 import ops
    from loner_b200._lib import LonerB200Error
    plan = ops.Plan(log2_hashmap_size=15, l2_bytes=700 * 1024)
    # exactly the maximum is accepted ...
    n, s = ops.NOALLOC_MAX_POINTS // 128, 128
    params, rays, z = make_inputs(n, s, 3, plan)
    got = torch.empty(n, s, device=DEV)
    ops.sigma_fwd_noalloc(plan, params, got, rays=rays, z=z)
    want, _ = ops.sigma_fwd(plan, params, rays=rays, z=z)
    assert rel_err(got, want) < 1e-5
    # ... one ray more is rejected with an error that names the limit, and nothing is launched or written
    params, rays, z = make_inputs(n + 1, s, 4, plan)
    out = torch.full((n + 1, s), -7.0, device=DEV)
    with pytest.raises(LonerB200Error, match="LB_NOALLOC_MAX_POINTS = 65536"):
        ops.sigma_fwd_noalloc(plan, params, out, rays=rays, z=z)
    torch.cuda.synchronize()
    assert (out == -7.0).all()
    # an empty batch is a no-op
    ops.sigma_fwd_noalloc(plan, params, torch.empty(0, s, device=DEV), rays=rays[:0], z=z[:0])
