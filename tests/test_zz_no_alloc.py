"""GPU: wd_adam_step_static -- the allocation-free variant of the fused optimiser step (its clipping scratch is a static
device array inside the library; include/wdmpnn.h) -- against the existing entry point wd_adam_step, through the C ABI,
on two calls in a row (the second call finds the first call's contents in the shared scratch).

This file is named so that pytest collects it LAST: a failure here must not hide
the tests before it.  It passes alone in a fresh process and after the other tests (it keeps no state of its own; the
static scratch is rewritten by every call before it is read).  Tolerance: 2e-6 absolute per step on parameters and
moments, the bound tests/test_gpu_optim.py holds wd_adam_step to (the two entry points run the same kernels in the same
order, so they are also expected to agree bit for bit)."""
import ctypes as C

import pytest
import torch

from polymer_chemprop_b200 import _capi

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
TOL = 2e-6


def _state(n, seed):
    g = torch.Generator(device=DEV).manual_seed(seed)
    p = torch.randn(n, device=DEV, generator=g)
    return p, torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)


@pytest.mark.parametrize("n,decoupled,max_norm", [(264300, 0, 1.0), (1001, 1, 0.5), (7, 0, 0.0)])
def test_static_scratch_adam_matches_wd_adam_step_on_two_calls_in_a_row(n, decoupled, max_norm):
    L, p = _capi.lib(), _capi.ptr
    stream = _capi.current_stream(DEV)
    ref_p, ref_m, ref_v = _state(n, 3)
    new_p, new_m, new_v = ref_p.clone(), ref_m.clone(), ref_v.clone()
    scratch = torch.zeros(int(L.wd_adam_scratch_floats()), device=DEV)
    norm_ref, norm_new = torch.zeros(1, device=DEV), torch.zeros(1, device=DEV)
    gen = torch.Generator(device=DEV).manual_seed(11)
    for step in (1, 2):                                       # two calls in a row
        grads = torch.randn(n, device=DEV, generator=gen) * (10.0 if step == 1 else 0.1)
        cfg = _capi.WdAdam(1e-3, 0.9, 0.999, 1e-8, 0.01, max_norm, step, decoupled)
        with torch.cuda.device(DEV):
            _capi.check(L.wd_adam_step(p(ref_p), p(grads), p(ref_m), p(ref_v), n, C.byref(cfg), p(scratch), p(norm_ref), stream))
            _capi.check(L.wd_adam_step_static(p(new_p), p(grads), p(new_m), p(new_v), n, C.byref(cfg), p(norm_new), stream))
        torch.cuda.synchronize()
        for a, b, name in ((new_p, ref_p, "params"), (new_m, ref_m, "exp_avg"), (new_v, ref_v, "exp_avg_sq")):
            assert torch.isfinite(a).all(), name
            assert float((a - b).abs().max()) <= TOL, (step, name)
            assert torch.equal(a, b), (step, name)            # same kernels, same order
        if max_norm > 0:
            assert float(norm_new) > 0 and abs(float(norm_new) - float(norm_ref)) <= TOL * max(1.0, float(norm_ref))
            assert abs(float(norm_new) - float(grads.double().norm())) <= 1e-4 * float(grads.double().norm())
    assert not torch.equal(new_p, _state(n, 3)[0])            # the steps did something


def test_static_scratch_adam_rejects_sizes_beyond_its_maximum():
    L, p = _capi.lib(), _capi.ptr
    limit = int(L.wd_adam_static_max_params())
    assert limit == 1 << 28
    buf = torch.zeros(16, device=DEV)
    before = buf.clone()
    cfg = _capi.WdAdam(1e-3, 0.9, 0.999, 1e-8, 0.0, 1.0, 1, 0)
    n0 = _capi.launch_count()
    with pytest.raises(ValueError, match="exceeds the maximum"):
        _capi.check(L.wd_adam_step_static(p(buf), p(buf), p(buf), p(buf), limit + 1, C.byref(cfg), None,
                                          _capi.current_stream(DEV)))
    torch.cuda.synchronize()
    assert _capi.launch_count() == n0 and torch.equal(buf, before)      # rejected before any launch
