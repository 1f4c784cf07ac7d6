"""lassim_eval_device_noalloc: the evaluation entry point for hosts that cannot
allocate device memory when they evaluate (their GPU memory belongs to another
allocator, they cannot take cudaMalloc's device-wide synchronisation, or they
run under a hard memory budget).  Its scratch is a fixed array inside the
library, so the checks here are: same costs as lassim_eval_device, twice in a
row (the array keeps the previous call's contents), and a clear refusal of
batches that do not fit.

Collected last on purpose (file name): a run that stops at the first failure
has then already run every other test.  Works alone in a fresh process too."""
import numpy as np
import pytest

from conftest import load_golden

pytestmark = pytest.mark.gpu


def test_noalloc_entry_point_matches_eval_device_twice():
    import torch
    from lassim_b200 import Engine
    p = load_golden("th2_problem")
    g = load_golden("th2_uniform")
    eng = Engine(0)
    eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        # two different batches, the second smaller than the first, so stale
        # counters or end states from call one would show up in call two
        for call, (lo, hi) in enumerate(((0, 96), (96, 128))):
            X = torch.from_numpy(np.ascontiguousarray(g["X"][lo:hi])).cuda()
            N = X.shape[0]
            want = torch.empty(N, dtype=torch.float64, device="cuda")
            got = torch.full((N,), float("nan"), dtype=torch.float64, device="cuda")
            eng.eval_device(N, X.data_ptr(), want.data_ptr(), stream=side.cuda_stream)
            side.synchronize()
            free_before = torch.cuda.mem_get_info()[0]          # cudaMemGetInfo
            eng.eval_device_noalloc(N, X.data_ptr(), got.data_ptr(), stream=side.cuda_stream)
            side.synchronize()
            if call == 1:
                # the promise itself: the device's free memory does not move (checked on the
                # second call; the first one may still load the library's kernel image lazily)
                assert torch.cuda.mem_get_info()[0] == free_before
            # same kernel, same inputs: the reduction tolerance of the parity tests
            np.testing.assert_allclose(got.cpu().numpy(), want.cpu().numpy(), rtol=1e-12)
            assert np.isfinite(got.cpu().numpy()).all()
    eng.close()


def test_noalloc_rejects_batches_beyond_the_built_in_scratch():
    from lassim_b200 import Engine, LassimError
    p = load_golden("th2_problem")
    eng = Engine(0)
    eng.set_problem(p["data"], p["sigma"], p["times"], p["y0"], p["mask"], p["pert"], 0.5)
    with pytest.raises(LassimError, match="scratch"):
        # 100,000 Th2 candidates would need ~220 MB; nothing is dereferenced
        eng.eval_device_noalloc(100000, 256, 256)
    with pytest.raises(LassimError):
        eng.eval_device_noalloc(8, None, None)
    eng.close()
