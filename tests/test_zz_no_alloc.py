"""nrd_cross_entropy_static: the cross-entropy entry for callers that cannot allocate device memory.

Such callers (everything on the device already belongs to another allocator or a fixed arena, or the
hook they run in only receives input/output pointers) cannot pass the row scratch that
nrd_cross_entropy needs, so this entry keeps it inside the library.  The test calls it through the C
ABI twice in a row -- with different batch sizes, so the second call sees the first call's leftovers in
the shared array -- and compares loss and dlogits with the caller-scratch entry (same kernels: bit-exact,
asserted at the 1e-6 relative tolerance the other cross-entropy test uses).  It also checks that a
batch beyond the built-in limit is refused, and that no device memory is allocated by the calls.
This file sorts last on purpose: a run that stops at the first failure has then already run every
other test.
"""
import pytest
import torch

from util import *  # noqa: F401,F403
from ddp_cifar10_b200 import _lib

pytestmark = pytest.mark.gpu


def _inputs(B, seed):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(B, 10, generator=g) * 3).to(DEV), torch.randint(0, 10, (B,), generator=g).to(DEV)


def test_static_scratch_cross_entropy_matches_caller_scratch_entry_twice():
    torch.cuda.synchronize()
    for call, (B, scale) in enumerate([(300, 1.0), (37, 0.5)]):        # second call: smaller B over stale rows
        logits, labels = _inputs(B, 100 + call)
        loss_a = torch.empty((), device=DEV); dl_a = torch.empty(B, 10, device=DEV); rows = torch.empty(B, device=DEV)
        loss_b = torch.full((), -1.0, device=DEV); dl_b = torch.full((B, 10), -1.0, device=DEV)
        check(lib.nrd_cross_entropy(logits.data_ptr(), labels.data_ptr(), B, scale, loss_a.data_ptr(), dl_a.data_ptr(),
                                    rows.data_ptr(), st()), "nrd_cross_entropy")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
        check(lib.nrd_cross_entropy_static(logits.data_ptr(), labels.data_ptr(), B, scale, loss_b.data_ptr(), dl_b.data_ptr(), st()),
              "nrd_cross_entropy_static")
        torch.cuda.synchronize()
        after = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
        if call == 1:                       # after the first call's one-time set-up nothing may change
            assert after == before
        assert abs(float(loss_b) - float(loss_a)) <= 1e-6 * max(1.0, abs(float(loss_a)))
        assert relmax(dl_b, dl_a) <= 1e-6
        ref = torch.nn.functional.cross_entropy(logits.double().cpu(), labels.cpu())
        assert abs(float(loss_b) - float(ref)) < 1e-6 * max(1.0, abs(float(ref)))


def test_static_scratch_cross_entropy_rejects_oversize_batch():
    logits, labels = _inputs(4, 7)
    loss = torch.empty((), device=DEV)
    rc = lib.nrd_cross_entropy_static(logits.data_ptr(), labels.data_ptr(), 8193, 1.0, loss.data_ptr(), None, st())
    assert rc == -1 and "8192" in _lib.last_error()
    assert lib.nrd_cross_entropy_static(None, labels.data_ptr(), 4, 1.0, loss.data_ptr(), None, st()) == -1
