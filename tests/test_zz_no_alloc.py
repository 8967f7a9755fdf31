"""Allocation-free entry point ``s2lc_se_fc_fwd_noalloc`` (static device scratch, no cudaMalloc) against the existing
``s2lc_se_fc_fwd`` through the C ABI: two calls in a row with different shapes (the shared scratch keeps the first
call's contents), the size limit, and torch as the ground truth.  Collected last on purpose (file name)."""
import pytest
import torch
import torch.nn.functional as F

from conftest import rel_err

pytestmark = pytest.mark.gpu


def _case(lib, L, n, c, s, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)

    def mk(*shape, scale=1.0):
        return torch.randn(*shape, generator=g, device="cuda") * scale

    pooled = mk(n, c)
    w1, b1, w2, b2 = mk(s, c, scale=0.2), mk(s, scale=0.1), mk(c, s, scale=0.2), mk(c, scale=0.1)
    st = torch.cuda.current_stream().cuda_stream
    gate_old, gate_new = torch.empty(n, c, device="cuda"), torch.full((n, c), float("nan"), device="cuda")
    L.check(lib.s2lc_se_fc_fwd(pooled.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(),
                               gate_old.data_ptr(), n, c, s, st))
    L.check(lib.s2lc_se_fc_fwd_noalloc(pooled.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(),
                                       gate_new.data_ptr(), n, c, s, st))
    torch.cuda.synchronize()
    ref = torch.sigmoid(F.linear(F.silu(F.linear(pooled, w1, b1)), w2, b2))
    assert torch.equal(gate_new, gate_old)              # same kernels, same order of operations
    assert rel_err(gate_new, ref) < 1e-5                # the tolerance of test_se_fc in test_gpu_kernels.py


def test_se_fc_fwd_noalloc_matches_existing_entry_point():
    from s2lc_b200 import _lib as L
    lib = L.load()
    L.check(lib.s2lc_arch_check(0))
    _case(lib, L, 32, 3072, 128, 1)      # b5 deepest block at bs 32
    _case(lib, L, 5, 48, 12, 2)          # second call: smaller problem over the first call's leftovers
    _case(lib, L, 256, 640, 160, 3)      # N * S = 40960 of the 65536 floats
    # beyond the static array: refused with a clear error, nothing launched
    n, c, s = 300, 64, 256               # 76800 > 65536
    z = torch.zeros(n * c + s * c + s + c * s + c, device="cuda")
    rc = lib.s2lc_se_fc_fwd_noalloc(z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(), z.data_ptr(),
                                    z.data_ptr(), n, c, s, torch.cuda.current_stream().cuda_stream)
    assert rc == -1 and b"exceeds the static scratch" in lib.s2lc_last_error()
    torch.cuda.synchronize()
