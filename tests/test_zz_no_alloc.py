"""The allocation-free local step, called straight through the C ABI.

Some hosts cannot let a library allocate: they have given every byte of device memory to their
own pool, or they call us from a place where cudaMalloc's device-wide synchronisation is not
acceptable.  For them libfedb200.so exports fedb_local_step_noalloc, which keeps its per-tile
gradient records in a fixed array inside the library instead of a caller buffer.  This test
drives that symbol with ctypes (no framework wrapper in between) and checks, on two consecutive
steps, that parameters, Adam moments, step counters and losses equal what the ordinary pair
fedb_fwd_bwd + fedb_reduce_opt produces from the same starting state.  The second step also
shows that whatever the first one left in the shared array does no harm.

The file name sorts last on purpose: if this fails under `pytest -x`, everything else has
already run.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

_P, _I32, _I64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64


def _engine(income):
    from fedb200 import FedEngine, FedProblem, make_shards
    probs = [FedProblem.custom(14, 2, make_shards("contiguous", 2500, 3), patience=0)]
    return FedEngine(probs, income.X_train[:2500], income.y_train[:2500], device="cuda", max_rounds=4)


def _bind():
    from fedb200.ops import _native as nat
    nat.lib(required=True)
    lib = ctypes.CDLL(nat.LIB_PATH)
    lib.fedb_local_step_noalloc.argtypes = [_P, _P, _I32, _P, _I32, _I32, _P, _P, _P, _P, _I32, _P, _P, _I32, _P, _P, _P,
                                            _P, _I64, _P]
    lib.fedb_local_step_noalloc.restype = _I32
    lib.fedb_noalloc_scratch_floats.restype = _I64
    lib.fedb_last_error.restype = ctypes.c_char_p
    return lib


def _call(lib, eng, groups_host):
    return lib.fedb_local_step_noalloc(
        groups_host.ctypes.data, eng.tb.d_groups.data_ptr(), eng.n_groups, eng.train_tiles.data_ptr(), len(eng.train_tiles_np),
        eng.tb.max_slab, eng.params.data_ptr(), eng.m.data_ptr(), eng.v.data_ptr(), eng.X.data_ptr(), eng.X.shape[1],
        eng.y.data_ptr(), eng.idx.data_ptr(), 0, eng.step.data_ptr(), eng.active.data_ptr(), eng.fed_state.data_ptr(),
        eng.loss.data_ptr(), eng.tb.smem_bytes, torch.cuda.current_stream().cuda_stream)


def test_noalloc_step_matches_regular_step_twice(income):
    lib = _bind()
    ref, sut = _engine(income), _engine(income)
    assert torch.equal(ref.params, sut.params)
    sut.scratch.fill_(float("nan"))  # the caller buffer must not be touched by the new entry point
    groups_host = np.ascontiguousarray(sut.tb.groups)
    for call in range(2):
        ref.enqueue_local_step(0)
        rc = _call(lib, sut, groups_host)
        assert rc == 0, lib.fedb_last_error().decode()
        torch.cuda.synchronize()
        for name in ("params", "m", "v", "loss"):   # same kernels, same order of operations: the bits agree
            assert torch.equal(getattr(sut, name), getattr(ref, name)), f"{name}, call {call}"
        assert torch.equal(sut.step, ref.step) and int(sut.step[0]) == call + 1
    assert bool(torch.isnan(sut.scratch).all())
    assert float((sut.params - _engine(income).params).abs().max()) > 0  # it did train


def test_noalloc_step_rejects_oversize(income):
    lib = _bind()
    eng = _engine(income)
    groups_host = np.ascontiguousarray(eng.tb.groups).copy()
    groups_host["max_tiles"][-1] = 1 << 20  # pretend the last client needs a million tiles
    before = eng.params.clone()
    rc = _call(lib, eng, groups_host)
    torch.cuda.synchronize()
    assert rc == -3 and b"built-in scratch" in lib.fedb_last_error()
    assert torch.equal(before, eng.params)
    assert lib.fedb_noalloc_scratch_floats() == 4 * 1024 * 1024
    # a table that reserves fewer records than one step writes is refused too (it would run past the group's range)
    groups_host = np.ascontiguousarray(eng.tb.groups).copy()
    groups_host["max_tiles"][0] = 1
    rc = _call(lib, eng, groups_host)
    torch.cuda.synchronize()
    assert rc == -2 and b"max_tiles" in lib.fedb_last_error()
    assert torch.equal(before, eng.params)
